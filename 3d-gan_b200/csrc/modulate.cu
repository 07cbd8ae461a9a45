// modulate.cu — the style-modulation glue of the StyleGAN2 synthesis layers on NDHWC tensors
// (stylegan.py:396-444 `modulated_conv3d`, live branch, and the bias_act that follows it, :526-531):
//
//   prologue   xs[n,v,c]  = x[n,v,c] * s[n,c]                                         (stylegan.py:427)
//   epilogue   out[n,v,c] = act(y[n,v,c] * d[n,c] + strength * field[n,v] + b[c])     (:429-432, :531, :522-524)
//
// and their gradients.  All HBM bound: the reference (and stock ATen) runs them as 5 broadcasting
// elementwise passes forward and a dozen elementwise / reduction passes backward; here the
// prologue is one pass each way (the backward also produces ds[n,c] = sum_v gy * x) and the
// epilogue is one pass each way (the backward also produces dd[n,c] = sum_v g * y, the per-sample
// bias gradient sum_v g and the noise-strength gradient sum g * field).
//
// Mapping: one CTA per (voxel chunk, sample); thread = (voxel lane, channel vector); a thread's
// channel vector is fixed, so s / d / bias live in registers and the per-(n, c) sums are register
// accumulators, reduced over the voxel lanes in shared memory, then over the chunks by a second
// tiny kernel in a fixed order (deterministic, no atomics).
#include "common.cuh"

namespace b2c {
namespace {

constexpr int kThr = 256;

struct ModGrid { int cv, rows, chunks; int64_t vpc; };   // channel vectors, voxel lanes, chunks, voxels per chunk

template <typename T>
bool mod_grid(int64_t N, int64_t V, int64_t C, ModGrid* g) {
  constexpr int VEC = Vec<T>::N;
  if (C % VEC || C / VEC > kThr || N <= 0 || N > 65535 || V <= 0) return false;
  g->cv = (int)(C / VEC);
  g->rows = kThr / g->cv;
  int64_t want = cdiv(V, (int64_t)g->rows * 32);                 // >= 32 voxels per thread
  int64_t cap = cdiv((int64_t)num_sms() * 8, N);
  int64_t chunks = want < cap ? want : cap;
  if (chunks < 1) chunks = 1;
  g->vpc = cdiv(cdiv(V, chunks), g->rows) * g->rows;
  g->chunks = (int)cdiv(V, g->vpc);
  return true;
}

template <typename T>
__global__ void __launch_bounds__(kThr) modulate_fwd_kernel(const T* __restrict__ x, const float* __restrict__ s,
                                                             T* __restrict__ y, int64_t V, int C, ModGrid G) {
  constexpr int VEC = Vec<T>::N;
  const int lane = threadIdx.x / G.cv, cvec = threadIdx.x % G.cv;
  if (lane >= G.rows) return;
  const int64_t n = blockIdx.y;
  float sc[VEC];
#pragma unroll
  for (int i = 0; i < VEC; i++) sc[i] = s[n * C + cvec * VEC + i];
  const int64_t v0 = (int64_t)blockIdx.x * G.vpc, v1 = min(V, v0 + G.vpc);
  const T* xp = x + (n * V) * C + cvec * VEC;
  T* yp = y + (n * V) * C + cvec * VEC;
  for (int64_t v = v0 + lane; v < v1; v += 2 * G.rows) {
    const int64_t v2 = v + G.rows;
    Vec<T> a, b;
    float fa[VEC], fb[VEC];
    a.load(xp + v * C);
    if (v2 < v1) b.load(xp + v2 * C);
    a.unpack(fa);
#pragma unroll
    for (int i = 0; i < VEC; i++) fa[i] *= sc[i];
    a.pack(fa); a.store(yp + v * C);
    if (v2 < v1) {
      b.unpack(fb);
#pragma unroll
      for (int i = 0; i < VEC; i++) fb[i] *= sc[i];
      b.pack(fb); b.store(yp + v2 * C);
    }
  }
}

// reduce per-thread [VEC] accumulators over the voxel lanes of the CTA; threads with lane 0 hold the result
template <int VEC>
__device__ __forceinline__ void lanes_sum(float* acc, float* sm, int lane, int cvec, const ModGrid& G) {
  __syncthreads();
  if (lane < G.rows)
#pragma unroll
    for (int i = 0; i < VEC; i++) sm[(lane * G.cv + cvec) * VEC + i] = acc[i];
  __syncthreads();
  if (lane == 0) {
    for (int r = 1; r < G.rows; r++)
#pragma unroll
      for (int i = 0; i < VEC; i++) acc[i] += sm[(r * G.cv + cvec) * VEC + i];
  }
}

// dx = gy * s ; partial ds[n, chunk, c] = sum_v gy * x
template <typename T>
__global__ void __launch_bounds__(kThr) modulate_bwd_kernel(const T* __restrict__ x, const T* __restrict__ gy,
                                                             const float* __restrict__ s, T* __restrict__ dx,
                                                             float* __restrict__ part, int64_t V, int C, ModGrid G) {
  constexpr int VEC = Vec<T>::N;
  __shared__ float sm[kThr * VEC];
  const int lane = threadIdx.x / G.cv, cvec = threadIdx.x % G.cv;
  const int64_t n = blockIdx.y;
  float sc[VEC], acc[VEC];
#pragma unroll
  for (int i = 0; i < VEC; i++) { sc[i] = lane < G.rows ? s[n * C + cvec * VEC + i] : 0.f; acc[i] = 0.f; }
  if (lane < G.rows) {
    const int64_t v0 = (int64_t)blockIdx.x * G.vpc, v1 = min(V, v0 + G.vpc);
    const int64_t base = (n * V) * C + cvec * VEC;
    for (int64_t v = v0 + lane; v < v1; v += G.rows) {
      Vec<T> a, g;
      float fa[VEC], fg[VEC];
      a.load(x + base + v * C);
      g.load(gy + base + v * C);
      a.unpack(fa); g.unpack(fg);
#pragma unroll
      for (int i = 0; i < VEC; i++) { acc[i] = fmaf(fg[i], fa[i], acc[i]); fg[i] *= sc[i]; }
      g.pack(fg); g.store(dx + base + v * C);
    }
  }
  lanes_sum<VEC>(acc, sm, lane, cvec, G);
  if (lane == 0) {
    float* p = part + ((n * G.chunks + blockIdx.x) * (int64_t)C) + cvec * VEC;
#pragma unroll
    for (int i = 0; i < VEC; i++) p[i] = acc[i];
  }
}

// out = act(y * d + strength * field + b)
template <typename T>
__global__ void __launch_bounds__(kThr) mod_epilogue_fwd_kernel(const T* __restrict__ y, const float* __restrict__ d,
                                                                 const float* __restrict__ field, int64_t field_n,
                                                                 const float* __restrict__ strength,
                                                                 const float* __restrict__ bias, T* __restrict__ out,
                                                                 int64_t V, int C, int act, float slope, ModGrid G) {
  constexpr int VEC = Vec<T>::N;
  const int lane = threadIdx.x / G.cv, cvec = threadIdx.x % G.cv;
  if (lane >= G.rows) return;
  const int64_t n = blockIdx.y;
  float dc[VEC], bc[VEC];
#pragma unroll
  for (int i = 0; i < VEC; i++) {
    dc[i] = d ? d[n * C + cvec * VEC + i] : 1.f;
    bc[i] = bias ? bias[cvec * VEC + i] : 0.f;
  }
  const float st = field ? *strength : 0.f;
  const float* fp = field ? field + (field_n > 1 ? n * V : 0) : nullptr;
  const int64_t v0 = (int64_t)blockIdx.x * G.vpc, v1 = min(V, v0 + G.vpc);
  const int64_t base = (n * V) * C + cvec * VEC;
  for (int64_t v = v0 + lane; v < v1; v += G.rows) {
    Vec<T> a;
    float f[VEC];
    a.load(y + base + v * C);
    a.unpack(f);
    const float nz = fp ? st * fp[v] : 0.f;
#pragma unroll
    for (int i = 0; i < VEC; i++) f[i] = act_f(fmaf(f[i], dc[i], nz + bc[i]), act, slope);
    a.pack(f); a.store(out + base + v * C);
  }
}

// g = gout * act'(out); dy = g * d; partials: dd[n,chunk,c] = sum_v g * y, db[n,chunk,c] = sum_v g,
// dst[n,chunk] = sum_{v,c} g * field[v]
template <typename T>
__global__ void __launch_bounds__(kThr) mod_epilogue_bwd_kernel(const T* __restrict__ y, const T* __restrict__ out,
                                                                 const T* __restrict__ gout, const float* __restrict__ d,
                                                                 const float* __restrict__ field, int64_t field_n,
                                                                 T* __restrict__ dy, float* __restrict__ part_d,
                                                                 float* __restrict__ part_b, float* __restrict__ part_s,
                                                                 int64_t V, int C, int act, float slope, ModGrid G) {
  constexpr int VEC = Vec<T>::N;
  __shared__ float sm[kThr * VEC];
  __shared__ float sw[kThr / 32];
  const int lane = threadIdx.x / G.cv, cvec = threadIdx.x % G.cv;
  const int64_t n = blockIdx.y;
  float dc[VEC], ad[VEC], ab[VEC], as = 0.f;
#pragma unroll
  for (int i = 0; i < VEC; i++) { dc[i] = (d && lane < G.rows) ? d[n * C + cvec * VEC + i] : 1.f; ad[i] = 0.f; ab[i] = 0.f; }
  if (lane < G.rows) {
    const float* fp = field ? field + (field_n > 1 ? n * V : 0) : nullptr;
    const int64_t v0 = (int64_t)blockIdx.x * G.vpc, v1 = min(V, v0 + G.vpc);
    const int64_t base = (n * V) * C + cvec * VEC;
    for (int64_t v = v0 + lane; v < v1; v += G.rows) {
      Vec<T> a, o, g;
      float fy[VEC], fo[VEC], fg[VEC];
      g.load(gout + base + v * C);
      o.load(out + base + v * C);
      if (part_d) a.load(y + base + v * C);
      g.unpack(fg); o.unpack(fo);
      if (part_d) a.unpack(fy);
      float rs = 0.f;
#pragma unroll
      for (int i = 0; i < VEC; i++) {
        const float gi = fg[i] * act_dy(fo[i], act, slope);
        ab[i] += gi;
        if (part_d) ad[i] = fmaf(gi, fy[i], ad[i]);
        rs += gi;
        fg[i] = gi * dc[i];
      }
      if (fp) as = fmaf(rs, fp[v], as);
      g.pack(fg); g.store(dy + base + v * C);
    }
  }
  if (part_d) {
    lanes_sum<VEC>(ad, sm, lane, cvec, G);
    if (lane == 0) {
      float* p = part_d + ((n * G.chunks + blockIdx.x) * (int64_t)C) + cvec * VEC;
#pragma unroll
      for (int i = 0; i < VEC; i++) p[i] = ad[i];
    }
  }
  lanes_sum<VEC>(ab, sm, lane, cvec, G);
  if (lane == 0) {
    float* p = part_b + ((n * G.chunks + blockIdx.x) * (int64_t)C) + cvec * VEC;
#pragma unroll
    for (int i = 0; i < VEC; i++) p[i] = ab[i];
  }
  if (part_s) {
    as = warp_sum(as);
    if ((threadIdx.x & 31) == 0) sw[threadIdx.x >> 5] = as;
    __syncthreads();
    if (threadIdx.x == 0) {
      float t = 0.f;
      for (int w = 0; w < kThr / 32; w++) t += sw[w];
      part_s[n * G.chunks + blockIdx.x] = t;
    }
  }
}

// out[o, c] = sum_{j < inner} part[o, j, c]   (fixed order)
__global__ void sum_partials_kernel(const float* __restrict__ part, float* __restrict__ out, int64_t outer, int inner,
                                    int C) {
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= outer * C) return;
  const int64_t o = idx / C;
  const int c = (int)(idx % C);
  float t = 0.f;
  for (int j = 0; j < inner; j++) t += part[(o * inner + j) * C + c];
  out[idx] = t;
}

int sum_partials(const float* part, float* out, int64_t outer, int inner, int C, cudaStream_t st) {
  const int64_t tot = outer * C;
  sum_partials_kernel<<<(unsigned)cdiv(tot, 256), 256, 0, st>>>(part, out, outer, inner, C);
  return launched();
}

template <typename T>
int modulate_t(bool bwd, const void* x, const void* gy, const float* s, void* y, float* ds, int64_t N, int64_t V,
               int64_t C, void* ws, size_t wsb, cudaStream_t st) {
  ModGrid G;
  if (!mod_grid<T>(N, V, C, &G)) return B2C_ERR_UNSUPPORTED;
  dim3 grid(G.chunks, (unsigned)N);
  if (!bwd) {
    modulate_fwd_kernel<T><<<grid, kThr, 0, st>>>(static_cast<const T*>(x), s, static_cast<T*>(y), V, (int)C, G);
    return launched();
  }
  if (wsb < (size_t)N * G.chunks * C * sizeof(float)) return B2C_ERR_WORKSPACE;
  float* part = static_cast<float*>(ws);
  modulate_bwd_kernel<T><<<grid, kThr, 0, st>>>(static_cast<const T*>(x), static_cast<const T*>(gy), s,
                                                static_cast<T*>(y), part, V, (int)C, G);
  int rc = launched();
  if (rc) return rc;
  return sum_partials(part, ds, N, G.chunks, (int)C, st);
}

template <typename T>
int epilogue_bwd_t(const void* y, const void* out, const void* gout, const float* d, const float* field,
                   int64_t field_n, void* dy, float* dd, float* dbias, float* dstrength, int64_t N, int64_t V,
                   int64_t C, int act, float slope, void* ws, size_t wsb, cudaStream_t st) {
  ModGrid G;
  if (!mod_grid<T>(N, V, C, &G)) return B2C_ERR_UNSUPPORTED;
  const size_t np = (size_t)N * G.chunks * C;
  if (wsb < (2 * np + (size_t)N * G.chunks) * sizeof(float)) return B2C_ERR_WORKSPACE;
  float* pd = static_cast<float*>(ws);
  float* pb = pd + np;
  float* ps = pb + np;
  dim3 grid(G.chunks, (unsigned)N);
  mod_epilogue_bwd_kernel<T><<<grid, kThr, 0, st>>>(static_cast<const T*>(y), static_cast<const T*>(out),
                                                    static_cast<const T*>(gout), d, field, field_n, static_cast<T*>(dy),
                                                    dd ? pd : nullptr, pb, (field && dstrength) ? ps : nullptr, V, (int)C,
                                                    act, slope, G);
  int rc = launched();
  if (rc) return rc;
  if (dd && (rc = sum_partials(pd, dd, N, G.chunks, (int)C, st))) return rc;
  if (dbias && (rc = sum_partials(pb, dbias, 1, (int)(N * G.chunks), (int)C, st))) return rc;
  if (field && dstrength && (rc = sum_partials(ps, dstrength, 1, (int)(N * G.chunks), 1, st))) return rc;
  return B2C_OK;
}

}  // namespace
}  // namespace b2c

using namespace b2c;

extern "C" size_t b2c_mod_workspace(int64_t N, int64_t V, int64_t C) {
  if (N <= 0 || V <= 0 || C <= 0) return 0;
  // chunks <= ceil(8 * SMs / N) + 1 for either dtype
  const size_t chunks = (size_t)cdiv((int64_t)num_sms() * 8, N) + 1;
  return (2 * (size_t)N * chunks * C + (size_t)N * chunks) * sizeof(float);
}

extern "C" int b2c_modulate_fwd(const void* x, const float* s, void* y, int64_t N, int64_t V, int64_t C, int dt,
                                b2c_stream st) {
  if (!x || !s || !y || !dt_ok(dt)) return B2C_ERR_INVALID;
  cudaStream_t cs = static_cast<cudaStream_t>(st);
  return dt == B2C_BF16 ? modulate_t<__nv_bfloat16>(false, x, nullptr, s, y, nullptr, N, V, C, nullptr, 0, cs)
                        : modulate_t<float>(false, x, nullptr, s, y, nullptr, N, V, C, nullptr, 0, cs);
}

extern "C" int b2c_modulate_bwd(const void* x, const void* gy, const float* s, void* dx, float* ds, int64_t N,
                                int64_t V, int64_t C, int dt, void* ws, size_t wsb, b2c_stream st) {
  if (!x || !gy || !s || !dx || !ds || !ws || !dt_ok(dt)) return B2C_ERR_INVALID;
  cudaStream_t cs = static_cast<cudaStream_t>(st);
  return dt == B2C_BF16 ? modulate_t<__nv_bfloat16>(true, x, gy, s, dx, ds, N, V, C, ws, wsb, cs)
                        : modulate_t<float>(true, x, gy, s, dx, ds, N, V, C, ws, wsb, cs);
}

extern "C" int b2c_mod_epilogue_fwd(const void* y, const float* d, const float* field, int64_t field_n,
                                    const float* strength, const float* bias, void* out, int64_t N, int64_t V,
                                    int64_t C, int dt, int act, float slope, b2c_stream st) {
  if (!y || !out || !dt_ok(dt) || (field && (!strength || (field_n != 1 && field_n != N)))) return B2C_ERR_INVALID;
  cudaStream_t cs = static_cast<cudaStream_t>(st);
  ModGrid G;
  if (dt == B2C_BF16) {
    if (!mod_grid<__nv_bfloat16>(N, V, C, &G)) return B2C_ERR_UNSUPPORTED;
    mod_epilogue_fwd_kernel<__nv_bfloat16><<<dim3(G.chunks, (unsigned)N), kThr, 0, cs>>>(
        static_cast<const __nv_bfloat16*>(y), d, field, field_n, strength, bias, static_cast<__nv_bfloat16*>(out), V,
        (int)C, act, slope, G);
  } else {
    if (!mod_grid<float>(N, V, C, &G)) return B2C_ERR_UNSUPPORTED;
    mod_epilogue_fwd_kernel<float><<<dim3(G.chunks, (unsigned)N), kThr, 0, cs>>>(
        static_cast<const float*>(y), d, field, field_n, strength, bias, static_cast<float*>(out), V, (int)C, act,
        slope, G);
  }
  return launched();
}

extern "C" int b2c_mod_epilogue_bwd(const void* y, const void* out, const void* gout, const float* d,
                                    const float* field, int64_t field_n, void* dy, float* dd, float* dbias,
                                    float* dstrength, int64_t N, int64_t V, int64_t C, int dt, int act, float slope,
                                    void* ws, size_t wsb, b2c_stream st) {
  if (!out || !gout || !dy || !ws || !dt_ok(dt) || (dd && (!d || !y)) || (field && field_n != 1 && field_n != N))
    return B2C_ERR_INVALID;
  cudaStream_t cs = static_cast<cudaStream_t>(st);
  return dt == B2C_BF16 ? epilogue_bwd_t<__nv_bfloat16>(y, out, gout, d, field, field_n, dy, dd, dbias, dstrength, N, V,
                                                        C, act, slope, ws, wsb, cs)
                        : epilogue_bwd_t<float>(y, out, gout, d, field, field_n, dy, dd, dbias, dstrength, N, V, C, act,
                                                slope, ws, wsb, cs);
}
