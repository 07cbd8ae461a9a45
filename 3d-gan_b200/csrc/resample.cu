// resample.cu — the x2 resampling steps of the BigGAN blocks on NDHWC tensors (HBM bound).
//   nearest x2 (+ channel drop)   utils.py:76-80  x[:, :Cout]; F.interpolate(scale_factor=2)
//   AvgPool3d(2)                  utils.py:113-116,128-129 / biggan.py:86,92
// ATen has no channels-last kernels for either in 3-D (it converts to NCDHW and back, three
// extra passes over the 8x larger tensor); here each pass reads and writes every element once
// with 16-byte vectors: a thread owns one channel vector of one LOW-resolution voxel and
// touches its 2x2x2 high-resolution block.
#include "common.cuh"

namespace b2c {
namespace {

constexpr int kThr = 256;
bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

struct RsGeo { int64_t n; int d, h, w, cx, c0, c; };      // low-resolution extents, channel slice

template <typename T, int VEC>
__device__ __forceinline__ void rs_ld(const T* p, float* o) {
  if (VEC > 1) { Vec<T> v; v.load(p); v.unpack(o); } else o[0] = ldf(p, 0);
}
template <typename T, int VEC>
__device__ __forceinline__ void rs_st(T* p, float* o) {
  if (VEC > 1) { Vec<T> v; v.pack(o); v.store(p); } else stf(p, 0, o[0]);
}

// y[n, 2d+a, 2h+b, 2w+e, c] = x[n, d, h, w, c0 + c]
template <typename T, int VEC>
__global__ void __launch_bounds__(kThr) up2_fwd_kernel(RsGeo g, const T* __restrict__ x, T* __restrict__ y, int64_t total) {
  const int64_t idx = (int64_t)blockIdx.x * kThr + threadIdx.x;
  if (idx >= total) return;
  const int cv = g.c / VEC;
  const int c = (int)(idx % cv) * VEC;
  int64_t v = idx / cv;
  const int w = (int)(v % g.w); v /= g.w;
  const int h = (int)(v % g.h); v /= g.h;
  const int d = (int)(v % g.d);
  const int64_t n = v / g.d;
  const T* src = x + (((n * g.d + d) * g.h + h) * g.w + w) * (int64_t)g.cx + g.c0 + c;
  T* dst = y + (((n * 2 * g.d + 2 * d) * 2 * g.h + 2 * h) * 2 * g.w + 2 * w) * (int64_t)g.c + c;
  const int64_t sw = g.c, sh = 2LL * g.w * g.c, sd = 4LL * g.h * g.w * g.c;
  if (VEC > 1) {
    Vec<T> val; val.load(src);
#pragma unroll
    for (int q = 0; q < 8; q++) val.store(dst + (q >> 2) * sd + ((q >> 1) & 1) * sh + (q & 1) * sw);
  } else {
    const T val = src[0];
#pragma unroll
    for (int q = 0; q < 8; q++) dst[(q >> 2) * sd + ((q >> 1) & 1) * sh + (q & 1) * sw] = val;
  }
}

// dx[n, d, h, w, k] = sum of the 2x2x2 block of dy for k in [c0, c0 + c), 0 for dropped channels
template <typename T, int VEC>
__global__ void __launch_bounds__(kThr) up2_bwd_kernel(RsGeo g, const T* __restrict__ dy, T* __restrict__ dx, int64_t total) {
  const int64_t idx = (int64_t)blockIdx.x * kThr + threadIdx.x;
  if (idx >= total) return;
  const int cv = g.cx / VEC;
  const int k = (int)(idx % cv) * VEC;
  int64_t v = idx / cv;
  T* dst = dx + v * (int64_t)g.cx + k;
  float acc[VEC];
#pragma unroll
  for (int i = 0; i < VEC; i++) acc[i] = 0.f;
  if (k >= g.c0 && k < g.c0 + g.c) {
    const int w = (int)(v % g.w); v /= g.w;
    const int h = (int)(v % g.h); v /= g.h;
    const int d = (int)(v % g.d);
    const int64_t n = v / g.d;
    const T* src = dy + (((n * 2 * g.d + 2 * d) * 2 * g.h + 2 * h) * 2 * g.w + 2 * w) * (int64_t)g.c + (k - g.c0);
    const int64_t sw = g.c, sh = 2LL * g.w * g.c, sd = 4LL * g.h * g.w * g.c;
#pragma unroll
    for (int q = 0; q < 8; q++) {
      float t[VEC];
      rs_ld<T, VEC>(src + (q >> 2) * sd + ((q >> 1) & 1) * sh + (q & 1) * sw, t);
#pragma unroll
      for (int i = 0; i < VEC; i++) acc[i] += t[i];
    }
  }
  rs_st<T, VEC>(dst, acc);
}

// g = OUTPUT (low-resolution) extents; input is [n, di, hi, wi, c] with di >= 2d etc. (floor)
template <typename T, int VEC>
__global__ void __launch_bounds__(kThr) pool2_fwd_kernel(RsGeo g, int hi, int wi, int di, const T* __restrict__ x,
                                                         T* __restrict__ y, int64_t total) {
  const int64_t idx = (int64_t)blockIdx.x * kThr + threadIdx.x;
  if (idx >= total) return;
  const int cv = g.c / VEC;
  const int c = (int)(idx % cv) * VEC;
  int64_t v = idx / cv;
  T* dst = y + v * (int64_t)g.c + c;
  const int w = (int)(v % g.w); v /= g.w;
  const int h = (int)(v % g.h); v /= g.h;
  const int d = (int)(v % g.d);
  const int64_t n = v / g.d;
  const T* src = x + (((n * di + 2 * d) * hi + 2 * h) * wi + 2 * w) * (int64_t)g.c + c;
  const int64_t sw = g.c, sh = (int64_t)wi * g.c, sd = (int64_t)hi * wi * g.c;
  float acc[VEC];
#pragma unroll
  for (int i = 0; i < VEC; i++) acc[i] = 0.f;
#pragma unroll
  for (int q = 0; q < 8; q++) {
    float t[VEC];
    rs_ld<T, VEC>(src + (q >> 2) * sd + ((q >> 1) & 1) * sh + (q & 1) * sw, t);
#pragma unroll
    for (int i = 0; i < VEC; i++) acc[i] += t[i];
  }
#pragma unroll
  for (int i = 0; i < VEC; i++) acc[i] *= 0.125f;
  rs_st<T, VEC>(dst, acc);
}

// one thread per INPUT-resolution channel vector: dx = dy[parent] / 8, 0 in the floor remainder
template <typename T, int VEC>
__global__ void __launch_bounds__(kThr) pool2_bwd_kernel(RsGeo g, int hi, int wi, int di, const T* __restrict__ dy,
                                                         T* __restrict__ dx, int64_t total) {
  const int64_t idx = (int64_t)blockIdx.x * kThr + threadIdx.x;
  if (idx >= total) return;
  const int cv = g.c / VEC;
  const int c = (int)(idx % cv) * VEC;
  int64_t v = idx / cv;
  T* dst = dx + v * (int64_t)g.c + c;
  const int w = (int)(v % wi); v /= wi;
  const int h = (int)(v % hi); v /= hi;
  const int d = (int)(v % di);
  const int64_t n = v / di;
  float t[VEC];
#pragma unroll
  for (int i = 0; i < VEC; i++) t[i] = 0.f;
  if ((d >> 1) < g.d && (h >> 1) < g.h && (w >> 1) < g.w) {
    rs_ld<T, VEC>(dy + (((n * g.d + (d >> 1)) * g.h + (h >> 1)) * g.w + (w >> 1)) * (int64_t)g.c + c, t);
#pragma unroll
    for (int i = 0; i < VEC; i++) t[i] *= 0.125f;
  }
  rs_st<T, VEC>(dst, t);
}

bool grid_ok(int64_t total, unsigned* blocks) {
  const int64_t b = cdiv(total, (int64_t)kThr);
  if (b > 0x7fffffffLL) return false;
  *blocks = (unsigned)b;
  return true;
}

template <typename T>
int up2(bool bwd, const void* a, void* b, const RsGeo& g, cudaStream_t st) {
  constexpr int V = Vec<T>::N;
  const bool vec = g.c % V == 0 && g.cx % V == 0 && g.c0 % V == 0 && aligned16(a) && aligned16(b);
  const int64_t vox = g.n * g.d * g.h * g.w;
  const int64_t total = vox * ((bwd ? g.cx : g.c) / (vec ? V : 1));
  if (total == 0) return B2C_OK;
  unsigned blocks;
  if (!grid_ok(total, &blocks)) return B2C_ERR_UNSUPPORTED;
  const T* ap = static_cast<const T*>(a); T* bp = static_cast<T*>(b);
  if (!bwd) { if (vec) up2_fwd_kernel<T, V><<<blocks, kThr, 0, st>>>(g, ap, bp, total); else up2_fwd_kernel<T, 1><<<blocks, kThr, 0, st>>>(g, ap, bp, total); }
  else      { if (vec) up2_bwd_kernel<T, V><<<blocks, kThr, 0, st>>>(g, ap, bp, total); else up2_bwd_kernel<T, 1><<<blocks, kThr, 0, st>>>(g, ap, bp, total); }
  return launched();
}

template <typename T>
int pool2(bool bwd, const void* a, void* b, const RsGeo& g, int di, int hi, int wi, cudaStream_t st) {
  constexpr int V = Vec<T>::N;
  const bool vec = g.c % V == 0 && aligned16(a) && aligned16(b);
  const int64_t vox = bwd ? g.n * di * hi * wi : g.n * g.d * g.h * g.w;
  const int64_t total = vox * (g.c / (vec ? V : 1));
  if (total == 0) return B2C_OK;
  unsigned blocks;
  if (!grid_ok(total, &blocks)) return B2C_ERR_UNSUPPORTED;
  const T* ap = static_cast<const T*>(a); T* bp = static_cast<T*>(b);
  if (!bwd) { if (vec) pool2_fwd_kernel<T, V><<<blocks, kThr, 0, st>>>(g, hi, wi, di, ap, bp, total); else pool2_fwd_kernel<T, 1><<<blocks, kThr, 0, st>>>(g, hi, wi, di, ap, bp, total); }
  else      { if (vec) pool2_bwd_kernel<T, V><<<blocks, kThr, 0, st>>>(g, hi, wi, di, ap, bp, total); else pool2_bwd_kernel<T, 1><<<blocks, kThr, 0, st>>>(g, hi, wi, di, ap, bp, total); }
  return launched();
}

// ---- MaxPool3d(k, stride, pad) on NDHWC (FID_ResNet.py:129 `nn.MaxPool3d(kernel_size=3, stride=2, padding=1)`) --
// ATen has no channels-last 3-D max-pool kernel: it converts to NCDHW and back (2.9 ms for the FID stem at batch
// 25); here one thread owns (output voxel, 16-byte channel vector) and reads the k^3 window once.  Padding counts
// as -inf (never materialised).  Ties go to the first element in (d, h, w) scan order, as in ATen.
struct MpGeo { int64_t n; int id, ih, iw, od, oh, ow, c, k, s, p; };

template <typename T, int VEC>
__global__ void __launch_bounds__(kThr) maxpool_fwd_kernel(MpGeo g, const T* __restrict__ x, T* __restrict__ y,
                                                           int64_t total) {
  const int64_t idx = (int64_t)blockIdx.x * kThr + threadIdx.x;
  if (idx >= total) return;
  const int cv = g.c / VEC;
  const int c = (int)(idx % cv) * VEC;
  int64_t v = idx / cv;
  T* dst = y + v * (int64_t)g.c + c;
  const int ow = (int)(v % g.ow); v /= g.ow;
  const int oh = (int)(v % g.oh); v /= g.oh;
  const int od = (int)(v % g.od);
  const int64_t n = v / g.od;
  float m[VEC];
#pragma unroll
  for (int i = 0; i < VEC; i++) m[i] = -INFINITY;
  for (int a = 0; a < g.k; a++) {
    const int jd = od * g.s + a - g.p;
    if (jd < 0 || jd >= g.id) continue;
    for (int b = 0; b < g.k; b++) {
      const int jh = oh * g.s + b - g.p;
      if (jh < 0 || jh >= g.ih) continue;
      const T* row = x + ((n * g.id + jd) * g.ih + jh) * (int64_t)g.iw * g.c + c;
      for (int e = 0; e < g.k; e++) {
        const int jw = ow * g.s + e - g.p;
        if (jw < 0 || jw >= g.iw) continue;
        float t[VEC];
        rs_ld<T, VEC>(row + (int64_t)jw * g.c, t);
#pragma unroll
        for (int i = 0; i < VEC; i++) m[i] = (t[i] > m[i] || t[i] != t[i]) ? t[i] : m[i];
      }
    }
  }
  rs_st<T, VEC>(dst, m);
}

// gather form of the backward (deterministic): an input voxel collects dy of every window that covers it and
// whose FIRST maximum (scan order) it is.  One thread per (input voxel, channel); not on a hot path (the
// reference only runs the FID network under no_grad, eval_utils.py:41-52).
template <typename T>
__global__ void __launch_bounds__(kThr) maxpool_bwd_kernel(MpGeo g, const T* __restrict__ x, const T* __restrict__ dy,
                                                           T* __restrict__ dx, int64_t total) {
  const int64_t idx = (int64_t)blockIdx.x * kThr + threadIdx.x;
  if (idx >= total) return;
  const int c = (int)(idx % g.c);
  int64_t v = idx / g.c;
  const int iw = (int)(v % g.iw); v /= g.iw;
  const int ih = (int)(v % g.ih); v /= g.ih;
  const int id = (int)(v % g.id);
  const int64_t n = v / g.id;
  const float self = ldf(x, idx);
  float acc = 0.f;
  for (int od = max(0, (id + g.p - g.k + g.s) / g.s); od < g.od && od * g.s - g.p <= id; od++)
    for (int oh = max(0, (ih + g.p - g.k + g.s) / g.s); oh < g.oh && oh * g.s - g.p <= ih; oh++)
      for (int ow = max(0, (iw + g.p - g.k + g.s) / g.s); ow < g.ow && ow * g.s - g.p <= iw; ow++) {
        // is (id, ih, iw) the first maximum of window (od, oh, ow)?
        bool win = true;
        for (int a = 0; a < g.k && win; a++) {
          const int jd = od * g.s + a - g.p;
          if (jd < 0 || jd >= g.id) continue;
          for (int b = 0; b < g.k && win; b++) {
            const int jh = oh * g.s + b - g.p;
            if (jh < 0 || jh >= g.ih) continue;
            for (int e = 0; e < g.k; e++) {
              const int jw = ow * g.s + e - g.p;
              if (jw < 0 || jw >= g.iw) continue;
              const float t = ldf(x, (((n * g.id + jd) * g.ih + jh) * (int64_t)g.iw + jw) * g.c + c);
              const bool before = (jd < id) || (jd == id && (jh < ih || (jh == ih && jw < iw)));
              if (t > self || (before && t == self)) { win = false; break; }
            }
          }
        }
        if (win) acc += ldf(dy, (((n * g.od + od) * g.oh + oh) * (int64_t)g.ow + ow) * g.c + c);
      }
  stf(dx, idx, acc);
}

template <typename T>
int maxpool_t(bool bwd, const void* a, const void* b, void* o, const MpGeo& g, cudaStream_t st) {
  constexpr int V = Vec<T>::N;
  unsigned blocks;
  if (!bwd) {
    const bool vec = g.c % V == 0 && aligned16(a) && aligned16(o);
    const int64_t total = g.n * g.od * g.oh * g.ow * (g.c / (vec ? V : 1));
    if (total == 0) return B2C_OK;
    if (!grid_ok(total, &blocks)) return B2C_ERR_UNSUPPORTED;
    if (vec) maxpool_fwd_kernel<T, V><<<blocks, kThr, 0, st>>>(g, static_cast<const T*>(a), static_cast<T*>(o), total);
    else maxpool_fwd_kernel<T, 1><<<blocks, kThr, 0, st>>>(g, static_cast<const T*>(a), static_cast<T*>(o), total);
    return launched();
  }
  const int64_t total = g.n * g.id * g.ih * g.iw * g.c;
  if (total == 0) return B2C_OK;
  if (!grid_ok(total, &blocks)) return B2C_ERR_UNSUPPORTED;
  maxpool_bwd_kernel<T><<<blocks, kThr, 0, st>>>(g, static_cast<const T*>(a), static_cast<const T*>(b), static_cast<T*>(o), total);
  return launched();
}

bool mp_geo(int64_t n, int d, int h, int w, int c, int k, int s, int p, MpGeo* g) {
  if (n < 0 || d <= 0 || h <= 0 || w <= 0 || c <= 0 || k < 1 || s < 1 || p < 0 || 2 * p > k) return false;
  const int od = (d + 2 * p - k) / s + 1, oh = (h + 2 * p - k) / s + 1, ow = (w + 2 * p - k) / s + 1;
  if (od <= 0 || oh <= 0 || ow <= 0) return false;
  *g = MpGeo{n, d, h, w, od, oh, ow, c, k, s, p};
  return true;
}

bool up_args_ok(const void* a, const void* b, int64_t n, int d, int h, int w, int cx, int c0, int c, int dt) {
  return a && b && n >= 0 && d > 0 && h > 0 && w > 0 && cx > 0 && c > 0 && c0 >= 0 && c0 + c <= cx && dt_ok(dt);
}

}  // namespace
}  // namespace b2c

using namespace b2c;

extern "C" int b2c_upsample2x_fwd(const void* x, void* y, int64_t n, int d, int h, int w, int cx, int c0, int c,
                                  int dt, b2c_stream s) {
  if (!up_args_ok(x, y, n, d, h, w, cx, c0, c, dt)) return B2C_ERR_INVALID;
  const RsGeo g{n, d, h, w, cx, c0, c};
  cudaStream_t st = static_cast<cudaStream_t>(s);
  return dt == B2C_BF16 ? up2<__nv_bfloat16>(false, x, y, g, st) : up2<float>(false, x, y, g, st);
}

extern "C" int b2c_upsample2x_bwd(const void* dy, void* dx, int64_t n, int d, int h, int w, int cx, int c0, int c,
                                  int dt, b2c_stream s) {
  if (!up_args_ok(dy, dx, n, d, h, w, cx, c0, c, dt)) return B2C_ERR_INVALID;
  const RsGeo g{n, d, h, w, cx, c0, c};
  cudaStream_t st = static_cast<cudaStream_t>(s);
  return dt == B2C_BF16 ? up2<__nv_bfloat16>(true, dy, dx, g, st) : up2<float>(true, dy, dx, g, st);
}

extern "C" int b2c_avgpool2_fwd(const void* x, void* y, int64_t n, int d, int h, int w, int c, int dt, b2c_stream s) {
  if (!x || !y || n < 0 || d < 2 || h < 2 || w < 2 || c <= 0 || !dt_ok(dt)) return B2C_ERR_INVALID;
  const RsGeo g{n, d / 2, h / 2, w / 2, c, 0, c};
  cudaStream_t st = static_cast<cudaStream_t>(s);
  return dt == B2C_BF16 ? pool2<__nv_bfloat16>(false, x, y, g, d, h, w, st) : pool2<float>(false, x, y, g, d, h, w, st);
}

extern "C" int b2c_avgpool2_bwd(const void* dy, void* dx, int64_t n, int d, int h, int w, int c, int dt, b2c_stream s) {
  if (!dy || !dx || n < 0 || d < 2 || h < 2 || w < 2 || c <= 0 || !dt_ok(dt)) return B2C_ERR_INVALID;
  const RsGeo g{n, d / 2, h / 2, w / 2, c, 0, c};
  cudaStream_t st = static_cast<cudaStream_t>(s);
  return dt == B2C_BF16 ? pool2<__nv_bfloat16>(true, dy, dx, g, d, h, w, st) : pool2<float>(true, dy, dx, g, d, h, w, st);
}

extern "C" int b2c_maxpool3d_fwd(const void* x, void* y, int64_t n, int d, int h, int w, int c, int k, int stride,
                                 int pad, int dt, b2c_stream s) {
  MpGeo g;
  if (!x || !y || !dt_ok(dt) || !mp_geo(n, d, h, w, c, k, stride, pad, &g)) return B2C_ERR_INVALID;
  cudaStream_t st = static_cast<cudaStream_t>(s);
  return dt == B2C_BF16 ? maxpool_t<__nv_bfloat16>(false, x, nullptr, y, g, st) : maxpool_t<float>(false, x, nullptr, y, g, st);
}

extern "C" int b2c_maxpool3d_bwd(const void* x, const void* dy, void* dx, int64_t n, int d, int h, int w, int c, int k,
                                 int stride, int pad, int dt, b2c_stream s) {
  MpGeo g;
  if (!x || !dy || !dx || !dt_ok(dt) || !mp_geo(n, d, h, w, c, k, stride, pad, &g)) return B2C_ERR_INVALID;
  cudaStream_t st = static_cast<cudaStream_t>(s);
  return dt == B2C_BF16 ? maxpool_t<__nv_bfloat16>(true, x, dy, dx, g, st) : maxpool_t<float>(true, x, dy, dx, g, st);
}

// ---- depth-to-space by 2 (and its adjoint) on NDHWC ---------------------------------------------------------
// z [n, jd, jh, jw, 8*c] (phase-major channels: r = rd*4 + rh*2 + rw) <-> y [n, od, oh, ow, c] with
// y[n, 2*j + r, :] = z[n, j, r, :], od <= 2*jd (cropped).  The second half of the phase-merged transposed
// convolution of b200conv3d/functional.py.  A thread moves one 16-byte channel vector; the z side is read /
// written fully contiguously, the y side in runs of 2*c elements (rw = 0, 1 are neighbours along w).
namespace b2c {
namespace {
struct D2sGeo { int64_t n; int jd, jh, jw, c, od, oh, ow; };

template <typename T, int VEC, bool INV>
__global__ void __launch_bounds__(kThr) d2s_kernel(D2sGeo g, const T* __restrict__ src, T* __restrict__ dst, int64_t total) {
  const int64_t idx = (int64_t)blockIdx.x * kThr + threadIdx.x;
  if (idx >= total) return;
  const int cv = g.c / VEC;
  const int ci = (int)(idx % cv) * VEC;
  int64_t v = idx / cv;
  const int r = (int)(v & 7); v >>= 3;
  const int jw = (int)(v % g.jw); v /= g.jw;
  const int jh = (int)(v % g.jh); v /= g.jh;
  const int jd = (int)(v % g.jd);
  const int64_t n = v / g.jd;
  const int od = 2 * jd + (r >> 2), oh = 2 * jh + ((r >> 1) & 1), ow = 2 * jw + (r & 1);
  const bool in = od < g.od && oh < g.oh && ow < g.ow;
  const int64_t zo = idx * VEC;
  const int64_t yo = (((n * g.od + od) * g.oh + oh) * (int64_t)g.ow + ow) * g.c + ci;
  float t[VEC];
  if (!INV) {                       // z -> y
    if (!in) return;
    rs_ld<T, VEC>(src + zo, t);
    rs_st<T, VEC>(dst + yo, t);
  } else {                          // y -> z (zero where the even grid overhangs y)
    if (in) rs_ld<T, VEC>(src + yo, t);
    else {
#pragma unroll
      for (int i = 0; i < VEC; i++) t[i] = 0.f;
    }
    rs_st<T, VEC>(dst + zo, t);
  }
}

template <typename T>
int d2s_t(bool inv, const void* src, void* dst, const D2sGeo& g, cudaStream_t st) {
  constexpr int V = Vec<T>::N;
  const bool vec = g.c % V == 0 && (reinterpret_cast<uintptr_t>(src) & 15) == 0 && (reinterpret_cast<uintptr_t>(dst) & 15) == 0;
  const int64_t total = g.n * g.jd * g.jh * g.jw * 8 * (g.c / (vec ? V : 1));
  if (total == 0) return B2C_OK;
  const int64_t b = cdiv(total, (int64_t)kThr);
  if (b > 0x7fffffffLL) return B2C_ERR_UNSUPPORTED;
  const unsigned blocks = (unsigned)b;
  const T* sp = static_cast<const T*>(src); T* dp = static_cast<T*>(dst);
  if (!inv) { if (vec) d2s_kernel<T, V, false><<<blocks, kThr, 0, st>>>(g, sp, dp, total); else d2s_kernel<T, 1, false><<<blocks, kThr, 0, st>>>(g, sp, dp, total); }
  else      { if (vec) d2s_kernel<T, V, true><<<blocks, kThr, 0, st>>>(g, sp, dp, total); else d2s_kernel<T, 1, true><<<blocks, kThr, 0, st>>>(g, sp, dp, total); }
  return launched();
}
}  // namespace
}  // namespace b2c

// ---- W-window gather (second generation of the W-folding of narrow stride-1 convolutions) ----------------
// x [rows, W, C] -> y [rows, W/F, (F + 2*pad) * C],  y[r, j, v*C + c] = x[r, F*j + v - pad, c]  (0 outside [0, W))
// turns a C -> Cout conv with 2*pad+1 taps along W into a (F+2*pad)*C -> F*Cout conv with ONE tap along W whose
// output [rows, W/F, F*Cout] IS the unfolded [rows, W, Cout] tensor; its block-Toeplitz filter has
// (2*pad+1) / (F + 2*pad) non-zero blocks (3/6 for F = 4 against 1/4 for the 3-tap folding).
// Adjoint (inverse != 0): x[r, w, c] = sum over the windows that contain w of y[r, j, (w - F*j + pad)*C + c].
namespace b2c {
namespace {
struct WinGeo { int64_t rows; int w, c, f, pad; };

template <typename T, int VEC, bool ADJ>
__global__ void __launch_bounds__(kThr) window_w_kernel(WinGeo g, const T* __restrict__ src, T* __restrict__ dst, int64_t total) {
  const int64_t idx = (int64_t)blockIdx.x * kThr + threadIdx.x;
  if (idx >= total) return;
  const int cv = g.c / VEC, win = g.f + 2 * g.pad, groups = g.w / g.f;
  const int ci = (int)(idx % cv) * VEC;
  int64_t t = idx / cv;
  float o[VEC];
  if (!ADJ) {                                  // one thread per 16-byte vector of y
    const int v = (int)(t % win); t /= win;
    const int j = (int)(t % groups);
    const int64_t r = t / groups;
    const int w = g.f * j + v - g.pad;
    if (w >= 0 && w < g.w) rs_ld<T, VEC>(src + (r * g.w + w) * (int64_t)g.c + ci, o);
    else {
#pragma unroll
      for (int i = 0; i < VEC; i++) o[i] = 0.f;
    }
    rs_st<T, VEC>(dst + idx * VEC, o);
  } else {                                     // one thread per 16-byte vector of x: gather form, no atomics
    const int w = (int)(t % g.w);
    const int64_t r = t / g.w;
#pragma unroll
    for (int i = 0; i < VEC; i++) o[i] = 0.f;
    // windows j with 0 <= w - F*j + pad < win
    int j0 = (w + g.pad - win + g.f) / g.f;    // ceil((w + pad - win + 1) / F)
    if (j0 < 0) j0 = 0;
    int j1 = (w + g.pad) / g.f;
    if (j1 > groups - 1) j1 = groups - 1;
    for (int j = j0; j <= j1; j++) {
      const int v = w - g.f * j + g.pad;
      float a[VEC];
      rs_ld<T, VEC>(src + ((r * groups + j) * (int64_t)win + v) * g.c + ci, a);
#pragma unroll
      for (int i = 0; i < VEC; i++) o[i] += a[i];
    }
    rs_st<T, VEC>(dst + idx * VEC, o);
  }
}

template <typename T>
int window_w_t(bool adj, const void* src, void* dst, const WinGeo& g, cudaStream_t st) {
  constexpr int V = Vec<T>::N;
  const bool vec = g.c % V == 0 && (reinterpret_cast<uintptr_t>(src) & 15) == 0 && (reinterpret_cast<uintptr_t>(dst) & 15) == 0;
  const int cv = g.c / (vec ? V : 1);
  const int64_t total = adj ? g.rows * g.w * cv : g.rows * (g.w / g.f) * (g.f + 2 * g.pad) * cv;
  if (total == 0) return B2C_OK;
  const int64_t b = cdiv(total, (int64_t)kThr);
  if (b > 0x7fffffffLL) return B2C_ERR_UNSUPPORTED;
  const unsigned blocks = (unsigned)b;
  const T* sp = static_cast<const T*>(src); T* dp = static_cast<T*>(dst);
  if (!adj) { if (vec) window_w_kernel<T, V, false><<<blocks, kThr, 0, st>>>(g, sp, dp, total); else window_w_kernel<T, 1, false><<<blocks, kThr, 0, st>>>(g, sp, dp, total); }
  else      { if (vec) window_w_kernel<T, V, true><<<blocks, kThr, 0, st>>>(g, sp, dp, total); else window_w_kernel<T, 1, true><<<blocks, kThr, 0, st>>>(g, sp, dp, total); }
  return launched();
}
}  // namespace
}  // namespace b2c

extern "C" int b2c_window_w(const void* src, void* dst, int64_t rows, int w, int c, int f, int pad, int dt, int inverse,
                            b2c_stream s) {
  using namespace b2c;
  if (!src || !dst || !dt_ok(dt) || rows < 0 || w <= 0 || c <= 0 || f <= 0 || pad < 0 || w % f != 0) return B2C_ERR_INVALID;
  WinGeo g{rows, w, c, f, pad};
  cudaStream_t st = static_cast<cudaStream_t>(s);
  return dt == B2C_BF16 ? window_w_t<__nv_bfloat16>(inverse != 0, src, dst, g, st) : window_w_t<float>(inverse != 0, src, dst, g, st);
}

extern "C" int b2c_depth_to_space2(const void* src, void* dst, int64_t n, int jd, int jh, int jw, int c, int od, int oh,
                                   int ow, int dt, int inverse, b2c_stream s) {
  using namespace b2c;
  if (!src || !dst || !dt_ok(dt) || n < 0 || jd <= 0 || jh <= 0 || jw <= 0 || c <= 0 || od <= 0 || oh <= 0 || ow <= 0 ||
      od > 2 * jd || oh > 2 * jh || ow > 2 * jw)
    return B2C_ERR_INVALID;
  D2sGeo g{n, jd, jh, jw, c, od, oh, ow};
  cudaStream_t st = static_cast<cudaStream_t>(s);
  return dt == B2C_BF16 ? d2s_t<__nv_bfloat16>(inverse != 0, src, dst, g, st) : d2s_t<float>(inverse != 0, src, dst, g, st);
}
