// tc_conv.cuh — host interface of the tcgen05 kernels (tc_conv.cu).
#pragma once
#include "common.cuh"
#include "tc_plan.h"

namespace b2c {
// out = plan(in, w_packed); bf16 I/O, fp32 accumulate in TMEM.
// `stats` (optional): sums of the STORED output for the normalisation layer that follows, accumulated in the epilogue:
// kind 1 = per output channel, kind 2 = per sample; every CTA writes one row part[block][2][len] (nblocks rows are
// written, returned in `nblocks`), summed afterwards in a fixed order (stat_finalize, norm.cu).
struct TcStats { int kind, len; float* part; size_t part_bytes; mutable int nblocks; };
size_t tc_conv_stats_bytes(int len);
bool tc_conv_stats_supported(const ConvPlan& p, int kind, int bias_mod);
int tc_conv_launch(const ConvPlan& p, const void* in, const void* w_packed, void* out, const float* bias, int bias_mod,
                   int act, float slope, cudaStream_t st, const TcStats* stats = nullptr);
// dW[a][b][t] (torch layout, fp32) = sum small[v,a] * big[v*s+k*dl-p, b]; split-K through `ws`.
bool tc_wgrad_supported(const Geo& g);
size_t tc_wgrad_workspace(const Geo& g);
int tc_wgrad_launch(const Geo& g, const void* small, const void* big, float* dw, void* ws, size_t wsb,
                    cudaStream_t st);
// Fused gather-form kernel for one-channel big tensors with <= 64 taps and <= 128 output channels
// (no im2col matrix in HBM); wp = [sc][Kp] bf16 as packed by thin_pack(gather).
bool tc_thin_gather_supported(const Geo& g);
int tc_thin_gather_launch(const Geo& g, const void* big, const void* wp, int Kp, void* out_small, const float* bias,
                          int act, float slope, cudaStream_t st, const TcStats* stats = nullptr);   // stats: kind 2 only
// fused weight gradient of the same layer class: dw[sc][T] fp32 (taps reversed when flip != 0)
size_t tc_thin_wgrad_workspace(const Geo& g);
int tc_thin_wgrad_launch(const Geo& g, const void* small, const void* big, float* dw, int flip, void* ws, size_t wsb,
                         cudaStream_t st);
// col2im of the k4 / s2 / p1 one-channel layers (thin.cu), TMA-staged: col = [n*sd*sh*sw][64] bf16 in k4s2_perm order
int tc_col2im1_k4s2_launch(const Geo& g, const void* col, void* big, const float* bias, int act, float slope,
                           cudaStream_t st);
// Sticky device error word of the tensor-core kernels (synchronises the device).
unsigned tc_error_fetch_clear();
}  // namespace b2c
