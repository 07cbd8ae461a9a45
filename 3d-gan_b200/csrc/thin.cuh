// thin.cuh — convolutions whose big-side tensor has ONE channel
// (dcgan.py:161 Conv3d(1->ndf), dcgan.py:67 ConvTranspose3d(ngf->1), FID_ResNet.py:119
// stem, biggan.py:78 input conv).  These are HBM-bound, K = taps is tiny, and a
// 1-channel NDHWC tensor has no 16-byte channel vector for TMA to fetch, so they are
// lowered explicitly:
//   gather  (small <- big): col = im2col(big) [rows][Kp]  ->  tcgen05 GEMM with W[sc][Kp]
//   scatter (big <- small): col = tcgen05 GEMM small[rows][sc] x W[Kp][sc]  ->  col2im (+bias, act)
//   wgrad                 : col = im2col(big)  ->  tcgen05 split-K GEMM  dW[sc][Kp]
// rows = voxels of the small tensor, Kp = taps rounded up to a multiple of 8 (>= 16).
#pragma once
#include "common.cuh"

namespace b2c {
// 0: not a thin problem.  1: big side has one channel, *tg = g.  2: SMALL side has one channel and
// the stride is 1 (biggan.py:42): *tg = the mirrored problem (roles of the two tensors exchanged,
// filter reversed, padding (K-1)*dl - p); callers swap gather <-> scatter and small <-> big and pass
// flip = true to thin_pack / thin_wgrad.
int thin_mode(const b2c_conv_desc* c, const Geo& g, Geo* tg);
int thin_kp(const Geo& g);
size_t thin_packed_bytes(const Geo& g);
// torch filter [sc][1][T] (conv) or [sc][1][T] (convT: [x.c][y.c=1][T]) -> gather: [sc][Kp]; scatter: [Kp][sc]
int thin_pack(const Geo& g, bool gather, const float* w, void* wp, cudaStream_t st, bool flip);
size_t thin_workspace(const Geo& g, int kind /*0 gather, 1 scatter, 2 wgrad*/);
struct TcStats;
int thin_gather(const Geo& g, const void* big, const void* wp, void* out_small, const float* bias, int act,
                float slope, void* ws, size_t wsb, cudaStream_t st, const TcStats* stats = nullptr);
int thin_scatter(const Geo& g, const void* small, const void* wp, void* out_big, const float* bias, int act,
                 float slope, void* ws, size_t wsb, cudaStream_t st);
int thin_wgrad(const Geo& g, const void* small, const void* big, float* dw, void* ws, size_t wsb, cudaStream_t st,
               bool flip);
}  // namespace b2c
