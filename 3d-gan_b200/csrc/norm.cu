// norm.cu — HBM-bound normalisation / activation passes over NDHWC tensors:
// BatchNorm3d statistics + apply (+residual, +activation) and backward,
// whole-sample LayerNorm (+activation) and backward, pointwise activations.
// Every kernel moves 16-byte vectors, grids are sized in multiples of the SM
// count, and all reductions are deterministic two-stage trees (no atomics).
#include <mutex>

#include "common.cuh"

namespace b2c {

constexpr int kThreads = 256;
constexpr int kBlocksPerSM = 4;

// VEC consecutive fp32 values (offset a multiple of VEC) as 16-byte loads / stores.
template <int VEC>
__device__ __forceinline__ void ldvecf(const float* __restrict__ p, int64_t off, float* out, float dflt) {
  if (!p) {
#pragma unroll
    for (int k = 0; k < VEC; k++) out[k] = dflt;
  } else if (VEC % 4 == 0) {
#pragma unroll
    for (int q = 0; q < VEC / 4; q++) {
      const float4 v = __ldg(reinterpret_cast<const float4*>(p + off) + q);
      out[4 * q] = v.x; out[4 * q + 1] = v.y; out[4 * q + 2] = v.z; out[4 * q + 3] = v.w;
    }
  } else {
#pragma unroll
    for (int k = 0; k < VEC; k++) out[k] = p[off + k];
  }
}
template <int VEC>
__device__ __forceinline__ void stvecf(float* __restrict__ p, int64_t off, const float* v) {
  if (!p) return;
  if (VEC % 4 == 0) {
#pragma unroll
    for (int q = 0; q < VEC / 4; q++)
      reinterpret_cast<float4*>(p + off)[q] = make_float4(v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]);
  } else {
#pragma unroll
    for (int k = 0; k < VEC; k++) p[off + k] = v[k];
  }
}

template <typename T, int VEC>
__device__ __forceinline__ void ldvec(const T* __restrict__ p, int64_t off, float* out) {
  if (VEC > 1) { Vec<T> v; v.load(p + off); v.unpack(out); } else out[0] = ldf(p, off);
}
// a VEC-element load kept packed (16 bytes for bf16 x 8) until it is consumed
template <typename T, int VEC> struct Pk {
  Vec<T> v;
  __device__ __forceinline__ void load(const T* __restrict__ p, int64_t off) { v.load(p + off); }
  __device__ __forceinline__ void unpack(float* o) const { v.unpack(o); }
};
template <typename T> struct Pk<T, 1> {
  float v;
  __device__ __forceinline__ void load(const T* __restrict__ p, int64_t off) { v = ldf(p, off); }
  __device__ __forceinline__ void unpack(float* o) const { o[0] = v; }
};
template <typename T, int VEC>
__device__ __forceinline__ void stvec(T* __restrict__ p, int64_t off, float* v) {
  if (VEC > 1) { Vec<T> o; o.pack(v); o.store(p + off); } else stf(p, off, v[0]);
}

// ===========================================================================
// Per-channel two-value reduction over rows of an [rows][C] tensor.
//   out0[c] = sum_r f0(r,c),  out1[c] = sum_r f1(r,c)
// Stage 1: block b owns a contiguous slab of rows; a thread owns VEC adjacent
// channels (one 16 B load) and walks rows with stride (kThreads / lanes_per_row).
// Partial sums for the same channel group are combined through shared memory
// and written to ws[b][2][C].  Stage 2 sums ws over blocks in fixed order.
// ===========================================================================
template <typename T>
struct BnBwdF {            // g = dy*act'(z);  f0 = g, f1 = g*xhat
  const T* dy; const T* res;
  const float *mean, *invstd, *scale, *shift;
  int act; float slope;
};

template <typename T, int VEC, bool BWD>
__global__ void __launch_bounds__(kThreads, BWD ? 2 : 1)   // BWD: 140 registers left ONE resident CTA per SM (37 % of HBM)
chan_reduce_stage1(const T* __restrict__ x, int64_t rows, int C, BnBwdF<T> f, float* __restrict__ ws,
                   int64_t rows_per_block) {
  extern __shared__ float sm[];           // [kThreads][2*VEC]
  const int groups = (C + VEC - 1) / VEC;                 // channel groups per row
  const int lanes = groups < kThreads ? groups : kThreads; // threads across channels
  const int rstep = kThreads / lanes;                     // rows processed in parallel
  const int lane = threadIdx.x % lanes, rlane = threadIdx.x / lanes;
  const int64_t r0 = (int64_t)blockIdx.x * rows_per_block;
  const int64_t r1 = r0 + rows_per_block < rows ? r0 + rows_per_block : rows;
  float* wsb = ws + (int64_t)blockIdx.x * 2 * C;
  for (int g0 = 0; g0 < groups; g0 += lanes) {
    const int g = g0 + lane;
    const int c0 = g * VEC;
    float a[VEC], b[VEC];
#pragma unroll
    for (int i = 0; i < VEC; i++) a[i] = b[i] = 0.f;
    if (g < groups && rlane < rstep) {
      float mu[VEC], is[VEC], sc[VEC], sh[VEC];
      if (BWD) {
#pragma unroll
        for (int i = 0; i < VEC; i++) {
          int c = c0 + i < C ? c0 + i : C - 1;
          mu[i] = f.mean[c]; is[i] = f.invstd[c]; sc[i] = f.scale[c]; sh[i] = f.shift[c];
        }
      }
      // four rows in flight per thread: loads first, then the math
      for (int64_t rb = r0 + rlane; rb < r1; rb += 4 * (int64_t)rstep) {
        Pk<T, VEC> xr[4], dr[4], rr[4];
#pragma unroll
        for (int u = 0; u < 4; u++) {
          const int64_t r = rb + (int64_t)u * rstep;
          if (r < r1) {
            const int64_t off = r * C + c0;
            xr[u].load(x, off);
            if (BWD) { dr[u].load(f.dy, off); if (f.res) rr[u].load(f.res, off); }
          }
        }
#pragma unroll
        for (int u = 0; u < 4; u++) {
          if (rb + (int64_t)u * rstep >= r1) break;
          float xv[VEC], dv[VEC], rv[VEC];
          xr[u].unpack(xv);
          if (BWD) { dr[u].unpack(dv); if (f.res) rr[u].unpack(rv); }
#pragma unroll
          for (int i = 0; i < VEC; i++) {
            if (BWD) {
              float z = xv[i] * sc[i] + sh[i];
              if (f.res) z += rv[i];
              float gg = dv[i] * act_dz(z, f.act, f.slope);
              a[i] += gg; b[i] += gg * (xv[i] - mu[i]) * is[i];
            } else {
              a[i] += xv[i]; b[i] += xv[i] * xv[i];
            }
          }
        }
      }
    }
    // combine the rstep row-lanes of each channel group
    float* my = sm + threadIdx.x * 2 * VEC;
#pragma unroll
    for (int i = 0; i < VEC; i++) { my[i] = a[i]; my[VEC + i] = b[i]; }
    __syncthreads();
    if (rlane == 0 && g < groups) {
      for (int j = 1; j < rstep; j++) {
        const float* o = sm + (j * lanes + lane) * 2 * VEC;
#pragma unroll
        for (int i = 0; i < VEC; i++) { a[i] += o[i]; b[i] += o[VEC + i]; }
      }
#pragma unroll
      for (int i = 0; i < VEC; i++)
        if (c0 + i < C) { wsb[c0 + i] = a[i]; wsb[C + c0 + i] = b[i]; }
    }
    __syncthreads();
  }
}

// Second stage: out[c] = sum over the stage-1 CTAs of their partial sums.  32 channels x 16 slices per CTA: a
// slice sums every 16th partial row (128-byte coalesced rows, double accumulators), the slices are combined
// through shared memory in a fixed order (deterministic).  The one-thread-per-channel version walked all
// ~1200 partial rows serially: 50-90 us per call, 2-9 % of the FID / BigGAN step.
constexpr int kS2Chan = 32, kS2Slices = 16;
__global__ void __launch_bounds__(kS2Chan * kS2Slices)
chan_reduce_stage2(const float* __restrict__ ws, int nblocks, int C, float* __restrict__ out0, float* __restrict__ out1) {
  __shared__ double sa[kS2Slices][kS2Chan], sb[kS2Slices][kS2Chan];
  const int lc = threadIdx.x % kS2Chan, sl = threadIdx.x / kS2Chan;
  const int c = blockIdx.x * kS2Chan + lc;
  double a = 0.0, b = 0.0;
  if (c < C)
    for (int i = sl; i < nblocks; i += kS2Slices) { a += ws[(int64_t)i * 2 * C + c]; b += ws[(int64_t)i * 2 * C + C + c]; }
  sa[sl][lc] = a; sb[sl][lc] = b;
  __syncthreads();
  if (sl == 0 && c < C) {
    for (int j = 1; j < kS2Slices; j++) { a += sa[j][lc]; b += sb[j][lc]; }
    out0[c] = (float)a; out1[c] = (float)b;
  }
}

static int reduce_blocks(int64_t rows) {
  int64_t nb = (int64_t)num_sms() * kBlocksPerSM;
  if (nb > rows) nb = rows;
  return (int)(nb < 1 ? 1 : nb);
}

template <typename T, bool BWD>
static int chan_reduce(const T* x, int64_t rows, int64_t C, BnBwdF<T> f, float* o0, float* o1, void* ws,
                       size_t wsb, cudaStream_t st) {
  const int nb = reduce_blocks(rows);
  if (wsb < (size_t)nb * 2 * C * sizeof(float)) return B2C_ERR_WORKSPACE;
  const int64_t rpb = cdiv(rows, nb);
  constexpr int V = Vec<T>::N;
  float* w = static_cast<float*>(ws);
  bool vec_ok = (C % V == 0) && ((reinterpret_cast<uintptr_t>(x) & 15) == 0);
  if (BWD) vec_ok = vec_ok && ((reinterpret_cast<uintptr_t>(f.dy) & 15) == 0) &&
                    (!f.res || (reinterpret_cast<uintptr_t>(f.res) & 15) == 0);
  if (vec_ok)
    chan_reduce_stage1<T, V, BWD><<<nb, kThreads, kThreads * 2 * V * sizeof(float), st>>>(x, rows, (int)C, f, w, rpb);
  else
    chan_reduce_stage1<T, 1, BWD><<<nb, kThreads, kThreads * 2 * sizeof(float), st>>>(x, rows, (int)C, f, w, rpb);
  int rc = launched();
  if (rc) return rc;
  chan_reduce_stage2<<<(unsigned)cdiv(C, kS2Chan), kS2Chan * kS2Slices, 0, st>>>(w, nb, (int)C, o0, o1);
  return launched();
}

// ===========================================================================
// BatchNorm finalize: tiny [C] kernel.
// ===========================================================================
__global__ void bn_finalize_kernel(const float* sum, const float* sq, double inv_count, double unbias, int C,
                                   float eps, float mom, const float* gamma, const float* beta, float* mean,
                                   float* invstd, float* scale, float* shift, float* rmean, float* rvar) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  double m = (double)sum[c] * inv_count;
  double var = (double)sq[c] * inv_count - m * m;
  if (var < 0) var = 0;
  double is = rsqrt(var + (double)eps);
  double g = gamma ? (double)gamma[c] : 1.0, b = beta ? (double)beta[c] : 0.0;
  mean[c] = (float)m; invstd[c] = (float)is;
  scale[c] = (float)(g * is); shift[c] = (float)(b - m * g * is);
  if (rmean) rmean[c] = (float)((1.0 - mom) * rmean[c] + mom * m);
  if (rvar) rvar[c] = (float)((1.0 - mom) * rvar[c] + mom * var * unbias);
}

// ===========================================================================
// Elementwise BN apply (+residual, +act) and backward apply.
// ===========================================================================
// A thread's channel group is fixed whenever the grid stride is a multiple of C / VEC (always for
// power-of-two C): the per-channel parameters are then loaded once, as 16-byte vectors, and two
// vectors per thread are in flight per iteration.
template <typename T, int VEC>
__global__ void __launch_bounds__(kThreads)
bn_act_fwd_kernel(const T* __restrict__ x, T* __restrict__ y, const T* __restrict__ res, int64_t nvec, int C,
                  const float* __restrict__ scale, const float* __restrict__ shift, int act, float slope) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  float sc[VEC], sh[VEC];
  int cc = -1;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += 2 * stride) {
    const int64_t i2 = i + stride;
    const bool two = i2 < nvec;
    Pk<T, VEC> xa, xb, ra, rb;
    xa.load(x, i * VEC);
    if (res) ra.load(res, i * VEC);
    if (two) { xb.load(x, i2 * VEC); if (res) rb.load(res, i2 * VEC); }
#pragma unroll
    for (int u = 0; u < 2; u++) {
      if (u == 1 && !two) break;
      const int64_t off = (u ? i2 : i) * VEC;
      const int c0 = (int)(off % C);
      if (c0 != cc) {
        if (VEC > 1) { ldvecf<VEC>(scale, c0, sc, 1.f); ldvecf<VEC>(shift, c0, sh, 0.f); }
        else { sc[0] = scale[c0]; sh[0] = shift[c0]; }
        cc = c0;
      }
      float xv[VEC], rv[VEC];
      (u ? xb : xa).unpack(xv);
      if (res) (u ? rb : ra).unpack(rv);
#pragma unroll
      for (int k = 0; k < VEC; k++) {
        float z = xv[k] * sc[k] + sh[k];
        if (res) z += rv[k];
        xv[k] = act_f(z, act, slope);
      }
      stvec<T, VEC>(y, off, xv);
    }
  }
}

// Four vectors of x and dy per thread are in flight per iteration (128 bytes a thread; at ~128 registers two
// CTAs are resident, 64 KB per SM: two vectors left the pass at half of HBM).
template <typename T, int VEC>
__global__ void __launch_bounds__(kThreads, 2)
bn_act_bwd_kernel(const T* __restrict__ x, const T* __restrict__ dy, const T* __restrict__ res,
                  T* __restrict__ dx, T* __restrict__ dres, int64_t nvec, int C, float inv_rows,
                  const float* __restrict__ mean, const float* __restrict__ invstd,
                  const float* __restrict__ scale, const float* __restrict__ shift,
                  const float* __restrict__ s1, const float* __restrict__ s2, int act, float slope) {
  constexpr int U = 4;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  float sc[VEC], sh[VEC], ka[VEC], kb[VEC];   // dx = sc * (g - k1 - xhat * k2) = sc * g + ka + kb * x
  int cc = -1;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += U * stride) {
    Pk<T, VEC> xr[U], dr[U], rr[U];
#pragma unroll
    for (int u = 0; u < U; u++) {
      const int64_t iu = i + u * stride;
      if (iu < nvec) {
        xr[u].load(x, iu * VEC); dr[u].load(dy, iu * VEC);
        if (res) rr[u].load(res, iu * VEC);
      }
    }
#pragma unroll
    for (int u = 0; u < U; u++) {
      const int64_t iu = i + u * stride;
      if (iu >= nvec) break;
      const int64_t off = iu * VEC;
      const int c0 = (int)(off % C);
      if (c0 != cc) {
        float mu[VEC], is[VEC], k1[VEC], k2[VEC];
        if (VEC > 1) {
          ldvecf<VEC>(scale, c0, sc, 1.f); ldvecf<VEC>(shift, c0, sh, 0.f); ldvecf<VEC>(mean, c0, mu, 0.f);
          ldvecf<VEC>(invstd, c0, is, 1.f); ldvecf<VEC>(s1, c0, k1, 0.f); ldvecf<VEC>(s2, c0, k2, 0.f);
        } else { sc[0] = scale[c0]; sh[0] = shift[c0]; mu[0] = mean[c0]; is[0] = invstd[c0]; k1[0] = s1[c0]; k2[0] = s2[c0]; }
#pragma unroll
        for (int k = 0; k < VEC; k++) {
          kb[k] = -sc[k] * k2[k] * inv_rows * is[k];
          ka[k] = -sc[k] * k1[k] * inv_rows - kb[k] * mu[k];
        }
        cc = c0;
      }
      float xv[VEC], dv[VEC], rv[VEC], gv[VEC];
      xr[u].unpack(xv); dr[u].unpack(dv);
      if (res) rr[u].unpack(rv);
#pragma unroll
      for (int k = 0; k < VEC; k++) {
        float z = xv[k] * sc[k] + sh[k];
        if (res) z += rv[k];
        float g = dv[k] * act_dz(z, act, slope);
        gv[k] = g;
        xv[k] = fmaf(sc[k], g, fmaf(kb[k], xv[k], ka[k]));
      }
      stvec<T, VEC>(dx, off, xv);
      if (dres) stvec<T, VEC>(dres, off, gv);
    }
  }
}

static unsigned ew_grid(int64_t nvec) {
  int64_t nb = cdiv(nvec, kThreads), cap = (int64_t)num_sms() * 8;
  return (unsigned)(nb < cap ? (nb < 1 ? 1 : nb) : cap);
}
static bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

template <typename T>
static int bn_fwd_t(const void* x, void* y, const void* res, int64_t rows, int64_t C, const float* scale,
                    const float* shift, int act, float slope, cudaStream_t st) {
  constexpr int V = Vec<T>::N;
  const int64_t n = rows * C;
  if (n == 0) return B2C_OK;
  const T* xp = static_cast<const T*>(x); T* yp = static_cast<T*>(y); const T* rp = static_cast<const T*>(res);
  if (C % V == 0 && aligned16(x) && aligned16(y) && (!res || aligned16(res)))
    bn_act_fwd_kernel<T, V><<<ew_grid(n / V), kThreads, 0, st>>>(xp, yp, rp, n / V, (int)C, scale, shift, act, slope);
  else
    bn_act_fwd_kernel<T, 1><<<ew_grid(n), kThreads, 0, st>>>(xp, yp, rp, n, (int)C, scale, shift, act, slope);
  return launched();
}

template <typename T>
static int bn_bwd_t(const void* x, const void* dy, const void* res, void* dx, void* dres, int64_t rows, int64_t C,
                    const float* mean, const float* invstd, const float* scale, const float* shift,
                    const float* s1, const float* s2, int act, float slope, cudaStream_t st) {
  constexpr int V = Vec<T>::N;
  const int64_t n = rows * C;
  if (n == 0) return B2C_OK;
  const T* xp = static_cast<const T*>(x); const T* dp = static_cast<const T*>(dy);
  const T* rp = static_cast<const T*>(res); T* dxp = static_cast<T*>(dx); T* drp = static_cast<T*>(dres);
  const float inv = 1.0f / (float)rows;
  if (C % V == 0 && aligned16(x) && aligned16(dy) && aligned16(dx) && (!res || aligned16(res)) &&
      (!dres || aligned16(dres)))
    bn_act_bwd_kernel<T, V><<<ew_grid(n / V), kThreads, 0, st>>>(xp, dp, rp, dxp, drp, n / V, (int)C, inv, mean,
                                                                 invstd, scale, shift, s1, s2, act, slope);
  else
    bn_act_bwd_kernel<T, 1><<<ew_grid(n), kThreads, 0, st>>>(xp, dp, rp, dxp, drp, n, (int)C, inv, mean, invstd,
                                                             scale, shift, s1, s2, act, slope);
  return launched();
}

}  // namespace b2c

namespace b2c {

// ===========================================================================
// Whole-sample LayerNorm (+activation).  x [N][L]; gamma/beta [L] fp32.
// ===========================================================================
__device__ __forceinline__ void block_sum2(float& a, float& b) {
  __shared__ float sa[kThreads / 32], sb[kThreads / 32];
  a = warp_sum(a); b = warp_sum(b);
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  if (l == 0) { sa[w] = a; sb[w] = b; }
  __syncthreads();
  if (w == 0) {
    a = l < kThreads / 32 ? sa[l] : 0.f; b = l < kThreads / 32 ? sb[l] : 0.f;
    a = warp_sum(a); b = warp_sum(b);
  }
  __syncthreads();
}

// grid (chunks, N): per-chunk (sum x, sum x^2) of one sample.
template <typename T, int VEC>
__global__ void __launch_bounds__(kThreads)
ln_partial_kernel(const T* __restrict__ x, int64_t L, float* __restrict__ part) {
  const int n = blockIdx.y;
  const int64_t nvec = L / VEC;
  const int64_t per = (nvec + gridDim.x - 1) / gridDim.x;
  const int64_t v0 = (int64_t)blockIdx.x * per, v1 = v0 + per < nvec ? v0 + per : nvec;
  const T* xs = x + (int64_t)n * L;
  float a = 0.f, b = 0.f;
  for (int64_t i = v0 + threadIdx.x; i < v1; i += kThreads) {
    float xv[VEC];
    ldvec<T, VEC>(xs, i * VEC, xv);
#pragma unroll
    for (int k = 0; k < VEC; k++) { a += xv[k]; b += xv[k] * xv[k]; }
  }
  block_sum2(a, b);
  if (threadIdx.x == 0) {
    part[((int64_t)n * gridDim.x + blockIdx.x) * 2] = a;
    part[((int64_t)n * gridDim.x + blockIdx.x) * 2 + 1] = b;
  }
}

// Backward statistics (sum g, sum g*xhat), g = dy*act'(z)*gamma.  grid (chunks, ceil(N / kLnGroup)).
// gamma / beta are fp32 over the whole sample shape (twice the bytes of a bf16 activation), so a
// thread owns VEC adjacent positions, loads their gamma / beta ONCE and walks kLnGroup samples with
// per-sample accumulators: affine traffic drops from 8 to 8 / kLnGroup bytes per element.
constexpr int kLnGroup = 8;
template <typename T, int VEC>
__global__ void __launch_bounds__(kThreads, 2)
ln_partial_bwd_kernel(const T* __restrict__ x, const T* __restrict__ dy, int N, int64_t L,
                      const float* __restrict__ gamma, const float* __restrict__ beta,
                      const float* __restrict__ mean, const float* __restrict__ rstd, int act, float slope,
                      float* __restrict__ part) {
  const int n0 = blockIdx.y * kLnGroup;
  const int nn = N - n0 < kLnGroup ? N - n0 : kLnGroup;
  float a[kLnGroup], b[kLnGroup], mu[kLnGroup], rs[kLnGroup];
#pragma unroll
  for (int u = 0; u < kLnGroup; u++) {
    a[u] = b[u] = 0.f;
    mu[u] = u < nn ? mean[n0 + u] : 0.f; rs[u] = u < nn ? rstd[n0 + u] : 0.f;
  }
  const int64_t nvec = L / VEC;
  const int64_t stride = (int64_t)gridDim.x * kThreads;
  for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < nvec; i += stride) {
    const int64_t e = i * VEC;
    float ga[VEC], be[VEC];
    ldvecf<VEC>(gamma, e, ga, 1.f); ldvecf<VEC>(beta, e, be, 0.f);
#pragma unroll
    for (int h = 0; h < kLnGroup; h += 4) {
      Pk<T, VEC> xr[4], dr[4];
#pragma unroll
      for (int u = 0; u < 4; u++)
        if (h + u < nn) { xr[u].load(x, (int64_t)(n0 + h + u) * L + e); dr[u].load(dy, (int64_t)(n0 + h + u) * L + e); }
#pragma unroll
      for (int u = 0; u < 4; u++)
        if (h + u < nn) {
          float xv[VEC], dv[VEC];
          xr[u].unpack(xv); dr[u].unpack(dv);
#pragma unroll
          for (int k = 0; k < VEC; k++) {
            const float xh = (xv[k] - mu[h + u]) * rs[h + u];
            const float g = dv[k] * act_dz(xh * ga[k] + be[k], act, slope) * ga[k];
            a[h + u] += g; b[h + u] += g * xh;
          }
        }
    }
  }
#pragma unroll
  for (int u = 0; u < kLnGroup; u++) {
    block_sum2(a[u], b[u]);
    if (threadIdx.x == 0 && u < nn) {
      part[((int64_t)(n0 + u) * gridDim.x + blockIdx.x) * 2] = a[u];
      part[((int64_t)(n0 + u) * gridDim.x + blockIdx.x) * 2 + 1] = b[u];
    }
  }
}

// one thread per sample; chunks is a few hundred at most.
__global__ void ln_finalize_kernel(const float* part, int chunks, int N, double invL, float eps, int bwd,
                                   float* o0, float* o1) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= N) return;
  double a = 0, b = 0;
  for (int i = 0; i < chunks; i++) { a += part[((int64_t)n * chunks + i) * 2]; b += part[((int64_t)n * chunks + i) * 2 + 1]; }
  if (bwd) { o0[n] = (float)(a * invL); o1[n] = (float)(b * invL); }   // mean(g), mean(g*xhat)
  else {
    double m = a * invL, var = b * invL - m * m;
    if (var < 0) var = 0;
    o0[n] = (float)m; o1[n] = (float)rsqrt(var + (double)eps);
  }
}

// grid (position blocks, sample ranges).  A thread owns VEC adjacent positions, loads gamma / beta
// once and walks its sample range four samples at a time (loads issued before the math).
template <typename T, int VEC>
__global__ void __launch_bounds__(kThreads)
ln_apply_fwd_kernel(const T* __restrict__ x, T* __restrict__ y, int N, int64_t L,
                    const float* __restrict__ gamma, const float* __restrict__ beta,
                    const float* __restrict__ mean, const float* __restrict__ rstd, int act, float slope) {
  const int per = (N + gridDim.y - 1) / gridDim.y;
  const int nb = blockIdx.y * per, ne = nb + per < N ? nb + per : N;
  const int64_t nvec = L / VEC;
  const int64_t stride = (int64_t)gridDim.x * kThreads;
  for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < nvec; i += stride) {
    const int64_t e = i * VEC;
    float ga[VEC], be[VEC];
    ldvecf<VEC>(gamma, e, ga, 1.f); ldvecf<VEC>(beta, e, be, 0.f);
    for (int n0 = nb; n0 < ne; n0 += 4) {
      Pk<T, VEC> xr[4];
#pragma unroll
      for (int u = 0; u < 4; u++)
        if (n0 + u < ne) xr[u].load(x, (int64_t)(n0 + u) * L + e);
#pragma unroll
      for (int u = 0; u < 4; u++)
        if (n0 + u < ne) {
          const float mu = mean[n0 + u], rs = rstd[n0 + u];
          float xv[VEC];
          xr[u].unpack(xv);
#pragma unroll
          for (int k = 0; k < VEC; k++) xv[k] = act_f((xv[k] - mu) * rs * ga[k] + be[k], act, slope);
          stvec<T, VEC>(y, (int64_t)(n0 + u) * L + e, xv);
        }
    }
  }
}

// Thread owns VEC adjacent elements e and walks the batch: dx per sample and
// the full-shape affine gradients dgamma[e], dbeta[e] without atomics.
template <typename T, int VEC>
__global__ void __launch_bounds__(kThreads, 2)
ln_apply_bwd_kernel(const T* __restrict__ x, const T* __restrict__ dy, T* __restrict__ dx, int N, int64_t L,
                    const float* __restrict__ gamma, const float* __restrict__ beta,
                    const float* __restrict__ mean, const float* __restrict__ rstd,
                    const float* __restrict__ mg, const float* __restrict__ mgx, int act, float slope,
                    float* __restrict__ dgamma, float* __restrict__ dbeta) {
  const int64_t nvec = L / VEC;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += stride) {
    const int64_t e = i * VEC;
    float ga[VEC], be[VEC], dg[VEC], db[VEC];
    ldvecf<VEC>(gamma, e, ga, 1.f); ldvecf<VEC>(beta, e, be, 0.f);
#pragma unroll
    for (int k = 0; k < VEC; k++) dg[k] = db[k] = 0.f;
    for (int n0 = 0; n0 < N; n0 += 4) {
      Pk<T, VEC> xr[4], dr[4];
#pragma unroll
      for (int u = 0; u < 4; u++)
        if (n0 + u < N) { xr[u].load(x, (int64_t)(n0 + u) * L + e); dr[u].load(dy, (int64_t)(n0 + u) * L + e); }
#pragma unroll
      for (int u = 0; u < 4; u++)
        if (n0 + u < N) {
          const int n = n0 + u;
          const float mu = mean[n], rs = rstd[n], a = mg[n], b = mgx[n];
          float xv[VEC], dv[VEC];
          xr[u].unpack(xv); dr[u].unpack(dv);
#pragma unroll
          for (int k = 0; k < VEC; k++) {
            float xh = (xv[k] - mu) * rs;
            float g0 = dv[k] * act_dz(xh * ga[k] + be[k], act, slope);
            dg[k] += g0 * xh; db[k] += g0;
            xv[k] = rs * (g0 * ga[k] - a - xh * b);
          }
          if (dx) stvec<T, VEC>(dx, (int64_t)n * L + e, xv);
        }
    }
    stvecf<VEC>(dgamma, e, dg); stvecf<VEC>(dbeta, e, db);
  }
}

static int ln_chunks(int64_t N, int64_t L) {
  int64_t want = cdiv((int64_t)num_sms() * kBlocksPerSM, N);
  int64_t maxc = cdiv(L, (int64_t)kThreads * 8);
  if (want > maxc) want = maxc;
  return (int)(want < 1 ? 1 : want);
}
// backward statistics: position blocks per sample group (two resident blocks per SM)
static int ln_chunks_bwd(int64_t N, int64_t L) {
  int64_t want = cdiv((int64_t)num_sms() * 2, cdiv(N, (int64_t)kLnGroup));
  int64_t maxc = cdiv(L, (int64_t)kThreads * 8);
  if (want > maxc) want = maxc;
  return (int)(want < 1 ? 1 : want);
}
// (position blocks, sample ranges) for the batch-walking apply kernel
static dim3 ln_apply_grid(int64_t N, int64_t nvec) {
  const int64_t target = (int64_t)num_sms() * 8;
  int64_t bx = cdiv(nvec, kThreads);
  if (bx > target) bx = target;
  int64_t by = bx >= target / 2 ? 1 : cdiv(target / 2, bx);
  const int64_t maxy = cdiv(N, 4);
  if (by > maxy) by = maxy;
  return dim3((unsigned)bx, (unsigned)(by < 1 ? 1 : by));
}

template <typename T>
static int ln_fwd_t(const void* x, void* y, int64_t N, int64_t L, const float* gamma, const float* beta, float eps,
                    int act, float slope, float* mean, float* rstd, void* ws, cudaStream_t st) {
  constexpr int V = Vec<T>::N;
  const T* xp = static_cast<const T*>(x); T* yp = static_cast<T*>(y);
  const int chunks = ln_chunks(N, L);
  float* part = static_cast<float*>(ws);
  const bool vec = (L % V == 0) && aligned16(x) && aligned16(y);
  dim3 grid(chunks, (unsigned)N);
  if (vec) ln_partial_kernel<T, V><<<grid, kThreads, 0, st>>>(xp, L, part);
  else     ln_partial_kernel<T, 1><<<grid, kThreads, 0, st>>>(xp, L, part);
  int rc = launched(); if (rc) return rc;
  ln_finalize_kernel<<<(unsigned)cdiv(N, 128), 128, 0, st>>>(part, chunks, (int)N, 1.0 / (double)L, eps, 0, mean, rstd);
  rc = launched(); if (rc) return rc;
  if (vec) ln_apply_fwd_kernel<T, V><<<ln_apply_grid(N, L / V), kThreads, 0, st>>>(xp, yp, (int)N, L, gamma, beta, mean, rstd, act, slope);
  else     ln_apply_fwd_kernel<T, 1><<<ln_apply_grid(N, L), kThreads, 0, st>>>(xp, yp, (int)N, L, gamma, beta, mean, rstd, act, slope);
  return launched();
}

template <typename T>
static int ln_bwd_t(const void* x, const void* dy, void* dx, int64_t N, int64_t L, const float* gamma,
                    const float* beta, const float* mean, const float* rstd, int act, float slope, float* dgamma,
                    float* dbeta, void* ws, cudaStream_t st) {
  constexpr int V = Vec<T>::N;
  const T* xp = static_cast<const T*>(x); const T* dp = static_cast<const T*>(dy); T* dxp = static_cast<T*>(dx);
  const int chunks = ln_chunks_bwd(N, L);
  float* part = static_cast<float*>(ws);
  float* mg = part + (size_t)N * chunks * 2;
  float* mgx = mg + N;
  const bool vec = (L % V == 0) && aligned16(x) && aligned16(dy) && (!dx || aligned16(dx));
  dim3 grid(chunks, (unsigned)cdiv(N, (int64_t)kLnGroup));
  if (vec) ln_partial_bwd_kernel<T, V><<<grid, kThreads, 0, st>>>(xp, dp, (int)N, L, gamma, beta, mean, rstd, act, slope, part);
  else     ln_partial_bwd_kernel<T, 1><<<grid, kThreads, 0, st>>>(xp, dp, (int)N, L, gamma, beta, mean, rstd, act, slope, part);
  int rc = launched(); if (rc) return rc;
  ln_finalize_kernel<<<(unsigned)cdiv(N, 128), 128, 0, st>>>(part, chunks, (int)N, 1.0 / (double)L, 0.f, 1, mg, mgx);
  rc = launched(); if (rc) return rc;
  if (vec) ln_apply_bwd_kernel<T, V><<<ew_grid(L / V), kThreads, 0, st>>>(xp, dp, dxp, (int)N, L, gamma, beta, mean, rstd, mg, mgx, act, slope, dgamma, dbeta);
  else     ln_apply_bwd_kernel<T, 1><<<ew_grid(L), kThreads, 0, st>>>(xp, dp, dxp, (int)N, L, gamma, beta, mean, rstd, mg, mgx, act, slope, dgamma, dbeta);
  return launched();
}

// ===========================================================================
// Pointwise activation fwd / bwd (bwd from the output).
// ===========================================================================
template <typename T, int VEC, bool BWD>
__global__ void __launch_bounds__(kThreads)
act_kernel(const T* __restrict__ a, const T* __restrict__ dy, T* __restrict__ out, int64_t nvec, int act, float slope) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += stride) {
    const int64_t off = i * VEC;
    float av[VEC], dv[VEC];
    if (VEC > 1) { Vec<T> v; v.load(a + off); v.unpack(av); if (BWD) { v.load(dy + off); v.unpack(dv); } }
    else { av[0] = ldf(a, off); if (BWD) dv[0] = ldf(dy, off); }
#pragma unroll
    for (int k = 0; k < VEC; k++) av[k] = BWD ? dv[k] * act_dy(av[k], act, slope) : act_f(av[k], act, slope);
    if (VEC > 1) { Vec<T> v; v.pack(av); v.store(out + off); } else stf(out, off, av[0]);
  }
}

template <typename T, bool BWD>
static int act_t(const void* a, const void* dy, void* out, int64_t n, int act, float slope, cudaStream_t st) {
  constexpr int V = Vec<T>::N;
  if (n == 0) return B2C_OK;
  const T* ap = static_cast<const T*>(a); const T* dp = static_cast<const T*>(dy); T* op = static_cast<T*>(out);
  const int64_t nv = n / V, tail = n - nv * V;
  const bool vec = aligned16(a) && aligned16(out) && (!BWD || aligned16(dy));
  if (vec && nv > 0) {
    act_kernel<T, V, BWD><<<ew_grid(nv), kThreads, 0, st>>>(ap, dp, op, nv, act, slope);
    int rc = launched(); if (rc) return rc;
    if (tail) {
      act_kernel<T, 1, BWD><<<1, kThreads, 0, st>>>(ap + nv * V, BWD ? dp + nv * V : nullptr, op + nv * V, tail, act, slope);
      return launched();
    }
    return B2C_OK;
  }
  act_kernel<T, 1, BWD><<<ew_grid(n), kThreads, 0, st>>>(ap, dp, op, n, act, slope);
  return launched();
}

}  // namespace b2c

using namespace b2c;

// dbias of b2c_conv3d_wgrad: out[c] = sum_r x[r][c] through the deterministic two-stage tree
// (the sum-of-squares output lands in the scratch tail).  ws >= channel_sum_workspace(rows, C).
namespace b2c {
size_t channel_sum_workspace(int64_t rows, int64_t C) {
  if (rows <= 0 || C <= 0) return 0;
  return ((size_t)reduce_blocks(rows) * 2 + 1) * C * sizeof(float);
}
int channel_sum(const void* x, int64_t rows, int C, int dt, float* out, void* ws, size_t wsb, cudaStream_t st) {
  if (rows <= 0 || C <= 0) return B2C_OK;
  if (!ws || wsb < channel_sum_workspace(rows, C)) return B2C_ERR_WORKSPACE;
  float* sq = static_cast<float*>(ws);
  void* part = sq + C;
  const size_t pb = wsb - (size_t)C * sizeof(float);
  if (dt == B2C_BF16) {
    BnBwdF<__nv_bfloat16> f{};
    return chan_reduce<__nv_bfloat16, false>(static_cast<const __nv_bfloat16*>(x), rows, C, f, out, sq, part, pb, st);
  }
  BnBwdF<float> f{};
  return chan_reduce<float, false>(static_cast<const float*>(x), rows, C, f, out, sq, part, pb, st);
}
}  // namespace b2c

// ---- statistics handed from a conv epilogue to the normalisation layer that follows -------------------
namespace b2c {
// second stage of the fused conv statistics: out[i] = sum over the CTAs' partial rows part[b][2][len]
int stat_finalize(const float* part, int nblocks, int len, float* sum, float* sq, cudaStream_t st) {
  if (!part || !sum || !sq || nblocks <= 0 || len <= 0) return B2C_ERR_INVALID;
  chan_reduce_stage2<<<(unsigned)cdiv(len, kS2Chan), kS2Chan * kS2Slices, 0, st>>>(part, nblocks, len, sum, sq);
  return launched();
}

// [N][chunks][2] partial sums -> sum[N], sq[N]  (per-sample statistics computed by a separate pass)
__global__ void ln_sums_kernel(const float* part, int chunks, int N, float* sum, float* sq) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= N) return;
  double a = 0, b = 0;
  for (int i = 0; i < chunks; i++) { a += part[((int64_t)n * chunks + i) * 2]; b += part[((int64_t)n * chunks + i) * 2 + 1]; }
  sum[n] = (float)a; sq[n] = (float)b;
}
__global__ void ln_from_sums_kernel(const float* sum, const float* sq, int N, double invL, float eps, float* mean,
                                    float* rstd) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= N) return;
  const double m = (double)sum[n] * invL;
  double var = (double)sq[n] * invL - m * m;
  if (var < 0) var = 0;
  mean[n] = (float)m; rstd[n] = (float)rsqrt(var + (double)eps);
}

// Unfused fall-backs of the same contract (paths whose epilogue does not accumulate): one read of y.
size_t sample_stats_workspace(int64_t N, int64_t L) {
  if (N <= 0 || L <= 0) return 0;
  return (size_t)N * ln_chunks(N, L) * 2 * sizeof(float);
}
int sample_stats(const void* y, int64_t N, int64_t L, int dt, float* sum, float* sq, void* ws, size_t wsb,
                 cudaStream_t st) {
  if (!y || !sum || !sq || !ws || N <= 0 || L <= 0 || N > 65535) return B2C_ERR_INVALID;
  if (wsb < sample_stats_workspace(N, L)) return B2C_ERR_WORKSPACE;
  const int chunks = ln_chunks(N, L);
  float* part = static_cast<float*>(ws);
  dim3 grid(chunks, (unsigned)N);
  if (dt == B2C_BF16) {
    const __nv_bfloat16* yp = static_cast<const __nv_bfloat16*>(y);
    if (L % 8 == 0 && aligned16(y)) ln_partial_kernel<__nv_bfloat16, 8><<<grid, kThreads, 0, st>>>(yp, L, part);
    else ln_partial_kernel<__nv_bfloat16, 1><<<grid, kThreads, 0, st>>>(yp, L, part);
  } else {
    const float* yp = static_cast<const float*>(y);
    if (L % 4 == 0 && aligned16(y)) ln_partial_kernel<float, 4><<<grid, kThreads, 0, st>>>(yp, L, part);
    else ln_partial_kernel<float, 1><<<grid, kThreads, 0, st>>>(yp, L, part);
  }
  int rc = launched();
  if (rc) return rc;
  ln_sums_kernel<<<(unsigned)cdiv(N, 128), 128, 0, st>>>(part, chunks, (int)N, sum, sq);
  return launched();
}
size_t channel_stats2_workspace(int64_t rows, int64_t C) { return (size_t)reduce_blocks(rows) * 2 * C * sizeof(float); }
int channel_stats2(const void* y, int64_t rows, int64_t C, int dt, float* sum, float* sq, void* ws, size_t wsb,
                   cudaStream_t st) {
  if (dt == B2C_BF16) {
    BnBwdF<__nv_bfloat16> f{};
    return chan_reduce<__nv_bfloat16, false>(static_cast<const __nv_bfloat16*>(y), rows, C, f, sum, sq, ws, wsb, st);
  }
  BnBwdF<float> f{};
  return chan_reduce<float, false>(static_cast<const float*>(y), rows, C, f, sum, sq, ws, wsb, st);
}
}  // namespace b2c

// LayerNorm forward from statistics the producing convolution already accumulated (b2c_fusion.stat_*): no pass
// over x for mean / variance - finalize [N] values, then the apply pass.
extern "C" int b2c_ln_act_fwd_stats(const void* x, void* y, int64_t N, int64_t L, int dt, const float* gamma,
                                    const float* beta, float eps, int act, float slope, const float* sum,
                                    const float* sqsum, float* mean, float* rstd, b2c_stream s) {
  if (!x || !y || !sum || !sqsum || !mean || !rstd || !dt_ok(dt) || N <= 0 || L <= 0 || N > 65535) return B2C_ERR_INVALID;
  cudaStream_t st = static_cast<cudaStream_t>(s);
  ln_from_sums_kernel<<<(unsigned)cdiv(N, 128), 128, 0, st>>>(sum, sqsum, (int)N, 1.0 / (double)L, eps, mean, rstd);
  int rc = launched();
  if (rc) return rc;
  if (dt == B2C_BF16) {
    const __nv_bfloat16* xp = static_cast<const __nv_bfloat16*>(x); __nv_bfloat16* yp = static_cast<__nv_bfloat16*>(y);
    if (L % 8 == 0 && aligned16(x) && aligned16(y))
      ln_apply_fwd_kernel<__nv_bfloat16, 8><<<ln_apply_grid(N, L / 8), kThreads, 0, st>>>(xp, yp, (int)N, L, gamma, beta, mean, rstd, act, slope);
    else
      ln_apply_fwd_kernel<__nv_bfloat16, 1><<<ln_apply_grid(N, L), kThreads, 0, st>>>(xp, yp, (int)N, L, gamma, beta, mean, rstd, act, slope);
  } else {
    const float* xp = static_cast<const float*>(x); float* yp = static_cast<float*>(y);
    if (L % 4 == 0 && aligned16(x) && aligned16(y))
      ln_apply_fwd_kernel<float, 4><<<ln_apply_grid(N, L / 4), kThreads, 0, st>>>(xp, yp, (int)N, L, gamma, beta, mean, rstd, act, slope);
    else
      ln_apply_fwd_kernel<float, 1><<<ln_apply_grid(N, L), kThreads, 0, st>>>(xp, yp, (int)N, L, gamma, beta, mean, rstd, act, slope);
  }
  return launched();
}

extern "C" size_t b2c_stats_workspace(int64_t rows, int64_t C) {
  if (rows <= 0 || C <= 0) return 0;
  return (size_t)reduce_blocks(rows) * 2 * C * sizeof(float);
}

extern "C" int b2c_channel_stats(const void* x, int64_t rows, int64_t C, int dt, float* sum, float* sq, void* ws,
                                 size_t wsb, b2c_stream s) {
  if (!x || !sum || !sq || !ws || !dt_ok(dt) || rows <= 0 || C <= 0 || C > 0x7fffffff) return B2C_ERR_INVALID;
  cudaStream_t st = static_cast<cudaStream_t>(s);
  if (dt == B2C_BF16) {
    BnBwdF<__nv_bfloat16> f{};
    return chan_reduce<__nv_bfloat16, false>(static_cast<const __nv_bfloat16*>(x), rows, C, f, sum, sq, ws, wsb, st);
  }
  BnBwdF<float> f{};
  return chan_reduce<float, false>(static_cast<const float*>(x), rows, C, f, sum, sq, ws, wsb, st);
}

// ---- allocation-free variant of b2c_channel_stats -------------------------------------------
// The stage-1 partial sums live in a file-scope device array instead of a caller workspace, for
// callers that cannot allocate (CUDA graph capture, serving loops).  Every device that loads the
// library holds the array until exit.  In host code `g_stats_static` is only a stand-in for the
// symbol: its device address is looked up once per device (cudaGetSymbolAddress with that device
// current) and that pointer is what the kernels get.
namespace b2c {
static __device__ __align__(256) float g_stats_static[B2C_STATIC_SCRATCH_BYTES / sizeof(float)];

static int stats_static_ptr(float** out) {
  static std::mutex mu;
  static float* ptrs[64] = {};
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess) return B2C_ERR_CUDA;
  if (dev < 0 || dev >= 64) return B2C_ERR_UNSUPPORTED;
  std::lock_guard<std::mutex> lk(mu);
  if (!ptrs[dev]) {
    float* p = nullptr;
    if (cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_stats_static) != cudaSuccess || !p) {
      (void)cudaGetLastError();
      return B2C_ERR_CUDA;
    }
    ptrs[dev] = p;
  }
  *out = ptrs[dev];
  return B2C_OK;
}
}  // namespace b2c

extern "C" int b2c_channel_stats_static(const void* x, int64_t rows, int64_t C, int dt, float* sum, float* sq,
                                        b2c_stream s) {
  if (!x || !sum || !sq || !dt_ok(dt) || rows <= 0 || C <= 0 || C > 0x7fffffff) return B2C_ERR_INVALID;
  if (b2c_stats_workspace(rows, C) > (size_t)B2C_STATIC_SCRATCH_BYTES) return B2C_ERR_STATIC_LIMIT;
  float* scratch = nullptr;
  const int rc = stats_static_ptr(&scratch);
  if (rc) return rc;
  // Stage 1 overwrites every partial that stage 2 reads: nothing in the array has to start at zero.
  cudaStream_t st = static_cast<cudaStream_t>(s);
  if (dt == B2C_BF16) {
    BnBwdF<__nv_bfloat16> f{};
    return chan_reduce<__nv_bfloat16, false>(static_cast<const __nv_bfloat16*>(x), rows, C, f, sum, sq, scratch,
                                             B2C_STATIC_SCRATCH_BYTES, st);
  }
  BnBwdF<float> f{};
  return chan_reduce<float, false>(static_cast<const float*>(x), rows, C, f, sum, sq, scratch,
                                   B2C_STATIC_SCRATCH_BYTES, st);
}

extern "C" int b2c_bn_finalize(const float* sum, const float* sq, int64_t count, int64_t C, float eps, float momentum,
                               const float* gamma, const float* beta, float* mean, float* invstd, float* scale,
                               float* shift, float* rmean, float* rvar, b2c_stream s) {
  if (!sum || !sq || !mean || !invstd || !scale || !shift || count <= 0 || C <= 0) return B2C_ERR_INVALID;
  const double unbias = count > 1 ? (double)count / (double)(count - 1) : 1.0;
  bn_finalize_kernel<<<(unsigned)cdiv(C, 256), 256, 0, static_cast<cudaStream_t>(s)>>>(
      sum, sq, 1.0 / (double)count, unbias, (int)C, eps, momentum, gamma, beta, mean, invstd, scale, shift, rmean, rvar);
  return launched();
}

extern "C" int b2c_bn_act_fwd(const void* x, void* y, const void* res, int64_t rows, int64_t C, int dt,
                              const float* scale, const float* shift, int act, float slope, b2c_stream s) {
  if (!x || !y || !scale || !shift || !dt_ok(dt) || rows < 0 || C <= 0) return B2C_ERR_INVALID;
  cudaStream_t st = static_cast<cudaStream_t>(s);
  return dt == B2C_BF16 ? bn_fwd_t<__nv_bfloat16>(x, y, res, rows, C, scale, shift, act, slope, st)
                        : bn_fwd_t<float>(x, y, res, rows, C, scale, shift, act, slope, st);
}

extern "C" int b2c_bn_act_bwd_reduce(const void* x, const void* dy, const void* res, int64_t rows, int64_t C, int dt,
                                     const float* mean, const float* invstd, const float* scale, const float* shift,
                                     int act, float slope, float* s1, float* s2, void* ws, size_t wsb, b2c_stream s) {
  if (!x || !dy || !mean || !invstd || !scale || !shift || !s1 || !s2 || !ws || !dt_ok(dt) || rows <= 0 || C <= 0)
    return B2C_ERR_INVALID;
  cudaStream_t st = static_cast<cudaStream_t>(s);
  if (dt == B2C_BF16) {
    BnBwdF<__nv_bfloat16> f{static_cast<const __nv_bfloat16*>(dy), static_cast<const __nv_bfloat16*>(res),
                            mean, invstd, scale, shift, act, slope};
    return chan_reduce<__nv_bfloat16, true>(static_cast<const __nv_bfloat16*>(x), rows, C, f, s1, s2, ws, wsb, st);
  }
  BnBwdF<float> f{static_cast<const float*>(dy), static_cast<const float*>(res), mean, invstd, scale, shift, act, slope};
  return chan_reduce<float, true>(static_cast<const float*>(x), rows, C, f, s1, s2, ws, wsb, st);
}

extern "C" int b2c_bn_act_bwd_apply(const void* x, const void* dy, const void* res, void* dx, void* dres, int64_t rows,
                                    int64_t C, int dt, const float* mean, const float* invstd, const float* scale,
                                    const float* shift, const float* s1, const float* s2, int act, float slope,
                                    b2c_stream s) {
  if (!x || !dy || !dx || !mean || !invstd || !scale || !shift || !s1 || !s2 || !dt_ok(dt) || rows <= 0 || C <= 0)
    return B2C_ERR_INVALID;
  cudaStream_t st = static_cast<cudaStream_t>(s);
  return dt == B2C_BF16
             ? bn_bwd_t<__nv_bfloat16>(x, dy, res, dx, dres, rows, C, mean, invstd, scale, shift, s1, s2, act, slope, st)
             : bn_bwd_t<float>(x, dy, res, dx, dres, rows, C, mean, invstd, scale, shift, s1, s2, act, slope, st);
}

extern "C" size_t b2c_ln_workspace(int64_t N, int64_t L) {
  if (N <= 0 || L <= 0) return 0;
  const int c1 = ln_chunks(N, L), c2 = ln_chunks_bwd(N, L);
  return ((size_t)N * (c1 > c2 ? c1 : c2) * 2 + 2 * (size_t)N) * sizeof(float);
}

extern "C" int b2c_ln_act_fwd(const void* x, void* y, int64_t N, int64_t L, int dt, const float* gamma,
                              const float* beta, float eps, int act, float slope, float* mean, float* rstd, void* ws,
                              size_t wsb, b2c_stream s) {
  if (!x || !y || !mean || !rstd || !ws || !dt_ok(dt) || N <= 0 || L <= 0 || N > 65535) return B2C_ERR_INVALID;
  if (wsb < b2c_ln_workspace(N, L)) return B2C_ERR_WORKSPACE;
  cudaStream_t st = static_cast<cudaStream_t>(s);
  return dt == B2C_BF16 ? ln_fwd_t<__nv_bfloat16>(x, y, N, L, gamma, beta, eps, act, slope, mean, rstd, ws, st)
                        : ln_fwd_t<float>(x, y, N, L, gamma, beta, eps, act, slope, mean, rstd, ws, st);
}

extern "C" int b2c_ln_act_bwd(const void* x, const void* dy, void* dx, int64_t N, int64_t L, int dt,
                              const float* gamma, const float* beta, const float* mean, const float* rstd, int act,
                              float slope, float* dgamma, float* dbeta, void* ws, size_t wsb, b2c_stream s) {
  if (!x || !dy || !mean || !rstd || !ws || !dt_ok(dt) || N <= 0 || L <= 0 || N > 65535) return B2C_ERR_INVALID;
  if (wsb < b2c_ln_workspace(N, L)) return B2C_ERR_WORKSPACE;
  cudaStream_t st = static_cast<cudaStream_t>(s);
  return dt == B2C_BF16
             ? ln_bwd_t<__nv_bfloat16>(x, dy, dx, N, L, gamma, beta, mean, rstd, act, slope, dgamma, dbeta, ws, st)
             : ln_bwd_t<float>(x, dy, dx, N, L, gamma, beta, mean, rstd, act, slope, dgamma, dbeta, ws, st);
}

extern "C" int b2c_act_fwd(const void* x, void* y, int64_t n, int dt, int act, float slope, b2c_stream s) {
  if (!x || !y || !dt_ok(dt) || n < 0) return B2C_ERR_INVALID;
  cudaStream_t st = static_cast<cudaStream_t>(s);
  return dt == B2C_BF16 ? act_t<__nv_bfloat16, false>(x, nullptr, y, n, act, slope, st)
                        : act_t<float, false>(x, nullptr, y, n, act, slope, st);
}

extern "C" int b2c_act_bwd(const void* y, const void* dy, void* dx, int64_t n, int dt, int act, float slope,
                           b2c_stream s) {
  if (!y || !dy || !dx || !dt_ok(dt) || n < 0) return B2C_ERR_INVALID;
  cudaStream_t st = static_cast<cudaStream_t>(s);
  return dt == B2C_BF16 ? act_t<__nv_bfloat16, true>(y, dy, dx, n, act, slope, st)
                        : act_t<float, true>(y, dy, dx, n, act, slope, st);
}
