// direct_conv.cuh — CUDA-core direct convolution (fp32 accumulate).  The exact
// fp32 path (north_star: fp32 max-rel <= 1e-4) and the catch-all for shapes the
// tcgen05 path does not take.
#pragma once
#include "common.cuh"

namespace b2c {
// small <- big gather : out[n,o,co]  = sum in[n,o*s+k*dl-p,ci] * W[t][co][ci]   (co in small.c, ci in big.c)
int direct_gather(const Geo& g, int dt, const void* in_big, const void* w, void* out_small, const float* bias,
                  int act, float slope, cudaStream_t st);
// big <- small scatter: out[n,i,co]  = sum [s|(i+p-k*dl)] in[n,(i+p-k*dl)/s,ci] * W[t][co][ci]  (co in big.c, ci in small.c)
int direct_scatter(const Geo& g, int dt, const void* in_small, const void* w, void* out_big, const float* bias,
                   int act, float slope, cudaStream_t st);
// dW[a][b][t] = sum small[n,o,a] * big[n,o*s+k*dl-p,b]   (torch layout, fp32, overwritten)
int direct_wgrad(const Geo& g, int dt, const void* small, const void* big, float* dw, cudaStream_t st);
// out[c] = sum_rows x[r][c]
// norm.cu: deterministic per-channel sum (dbias)
size_t channel_sum_workspace(int64_t rows, int64_t C);
int channel_sum(const void* x, int64_t rows, int C, int dt, float* out, void* ws, size_t wsb, cudaStream_t st);
// statistics handed from a conv epilogue to the following normalisation layer (norm.cu)
int stat_finalize(const float* part, int nblocks, int len, float* sum, float* sq, cudaStream_t st);
size_t sample_stats_workspace(int64_t N, int64_t L);
int sample_stats(const void* y, int64_t N, int64_t L, int dt, float* sum, float* sq, void* ws, size_t wsb, cudaStream_t st);
size_t channel_stats2_workspace(int64_t rows, int64_t C);
int channel_stats2(const void* y, int64_t rows, int64_t C, int dt, float* sum, float* sq, void* ws, size_t wsb, cudaStream_t st);
}  // namespace b2c
