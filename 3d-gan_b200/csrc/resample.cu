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
