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


// ---------------------------------------------------------------------------------------------
// AdaIN of the v1-like generator (stylegan.py:931-959) with the trilinear x2 that follows every
// second one (stylegan.py:1081 F.interpolate(scale_factor=2, mode='trilinear'), align_corners=False)
// folded in: tri() commutes with the per-(n,c) affine, so
//   y[n,o,c] = a[n,c] * tri(x)[n,o,c] + b[n,c],  a = s_std / sigma,  b = s_mean - mean * a
// is ONE pass that reads the low-resolution tensor and writes the high-resolution one.
// Per axis: out[2i] = .75 x[i] + .25 x[i-1], out[2i+1] = .75 x[i] + .25 x[i+1], indices clamped;
// the adjoint gathers positions 2i-1 .. 2i+2 with weights (.25, .75, .75, .25), indices clamped.

// partial sums over a chunk: p1 = sum_v x, p2 = sum_v x^2
template <typename T>
__global__ void __launch_bounds__(kThr) inst_stats_kernel(const T* __restrict__ x, float* __restrict__ p1,
                                                           float* __restrict__ p2, int64_t V, int C, ModGrid G) {
  constexpr int VEC = Vec<T>::N;
  __shared__ float sm[kThr * VEC];
  const int lane = threadIdx.x / G.cv, cvec = threadIdx.x % G.cv;
  const int64_t n = blockIdx.y;
  float a1[VEC], a2[VEC];
#pragma unroll
  for (int i = 0; i < VEC; i++) { a1[i] = 0.f; a2[i] = 0.f; }
  if (lane < G.rows) {
    const int64_t v0 = (int64_t)blockIdx.x * G.vpc, v1 = min(V, v0 + G.vpc);
    const T* xp = x + (n * V) * C + cvec * VEC;
    for (int64_t v = v0 + lane; v < v1; v += G.rows) {
      Vec<T> a;
      float f[VEC];
      a.load(xp + v * C);
      a.unpack(f);
#pragma unroll
      for (int i = 0; i < VEC; i++) { a1[i] += f[i]; a2[i] = fmaf(f[i], f[i], a2[i]); }
    }
  }
  lanes_sum<VEC>(a1, sm, lane, cvec, G);
  lanes_sum<VEC>(a2, sm, lane, cvec, G);
  if (lane == 0) {
    const int64_t o = ((n * G.chunks + blockIdx.x) * (int64_t)C) + cvec * VEC;
#pragma unroll
    for (int i = 0; i < VEC; i++) { p1[o + i] = a1[i]; p2[o + i] = a2[i]; }
  }
}

struct Vol { int d, h, w; };

// forward: UP ? trilinear x2 of x : x, times a[n,c], plus b[n,c].  V = OUTPUT voxels per sample.
template <typename T, bool UP>
__global__ void __launch_bounds__(kThr) adain_apply_kernel(const T* __restrict__ x, const float* __restrict__ a,
                                                            const float* __restrict__ b, T* __restrict__ y, Vol in,
                                                            int64_t V, int C, ModGrid G) {
  constexpr int VEC = Vec<T>::N;
  const int lane = threadIdx.x / G.cv, cvec = threadIdx.x % G.cv;
  if (lane >= G.rows) return;
  const int64_t n = blockIdx.y;
  float ac[VEC], bc[VEC];
#pragma unroll
  for (int i = 0; i < VEC; i++) { ac[i] = a[n * C + cvec * VEC + i]; bc[i] = b[n * C + cvec * VEC + i]; }
  const int64_t v0 = (int64_t)blockIdx.x * G.vpc, v1 = min(V, v0 + G.vpc);
  const int64_t vin = (int64_t)in.d * in.h * in.w;
  const T* xp = x + (n * vin) * C + cvec * VEC;
  T* yp = y + (n * V) * C + cvec * VEC;
  for (int64_t v = v0 + lane; v < v1; v += G.rows) {
    float f[VEC];
    if (!UP) {
      Vec<T> t; t.load(xp + v * C); t.unpack(f);
    } else {
      const int ow = (int)(v % (2 * in.w)), oh = (int)((v / (2 * in.w)) % (2 * in.h)), od = (int)(v / ((int64_t)4 * in.w * in.h));
      // neighbour on the side of the output's parity, clamped
      const int d0 = od >> 1, h0 = oh >> 1, w0 = ow >> 1;
      const int d1 = min(max(d0 + ((od & 1) ? 1 : -1), 0), in.d - 1);
      const int h1 = min(max(h0 + ((oh & 1) ? 1 : -1), 0), in.h - 1);
      const int w1 = min(max(w0 + ((ow & 1) ? 1 : -1), 0), in.w - 1);
#pragma unroll
      for (int i = 0; i < VEC; i++) f[i] = 0.f;
#pragma unroll
      for (int q = 0; q < 8; q++) {
        const int dd = (q & 4) ? d1 : d0, hh = (q & 2) ? h1 : h0, ww = (q & 1) ? w1 : w0;
        const float wt = ((q & 4) ? .25f : .75f) * ((q & 2) ? .25f : .75f) * ((q & 1) ? .25f : .75f);
        Vec<T> t;
        float g[VEC];
        t.load(xp + (((int64_t)dd * in.h + hh) * in.w + ww) * C);
        t.unpack(g);
#pragma unroll
        for (int i = 0; i < VEC; i++) f[i] = fmaf(wt, g[i], f[i]);
      }
    }
#pragma unroll
    for (int i = 0; i < VEC; i++) f[i] = fmaf(f[i], ac[i], bc[i]);
    Vec<T> o; o.pack(f); o.store(yp + v * C);
  }
}

// backward pass 1: t = UP ? tri^T(gy) : gy at every INPUT voxel; partials r1 = sum t, r2 = sum t * x;
// UP: t is written to `tbuf` (the dx buffer) for pass 2.   V = INPUT voxels per sample.
template <typename T, bool UP>
__global__ void __launch_bounds__(kThr) adain_bwd_reduce_kernel(const T* __restrict__ x, const T* __restrict__ gy,
                                                                 T* __restrict__ tbuf, float* __restrict__ p1,
                                                                 float* __restrict__ p2, Vol in, int64_t V, int C,
                                                                 ModGrid G) {
  constexpr int VEC = Vec<T>::N;
  __shared__ float sm[kThr * VEC];
  const int lane = threadIdx.x / G.cv, cvec = threadIdx.x % G.cv;
  const int64_t n = blockIdx.y;
  float a1[VEC], a2[VEC];
#pragma unroll
  for (int i = 0; i < VEC; i++) { a1[i] = 0.f; a2[i] = 0.f; }
  if (lane < G.rows) {
    const int64_t v0 = (int64_t)blockIdx.x * G.vpc, v1 = min(V, v0 + G.vpc);
    const int64_t base = (n * V) * C + cvec * VEC;
    const T* gp = gy + (UP ? 8 : 1) * (n * V) * C + cvec * VEC;
    const int OD = 2 * in.d, OH = 2 * in.h, OW = 2 * in.w;
    for (int64_t v = v0 + lane; v < v1; v += G.rows) {
      float t[VEC], fx[VEC];
      Vec<T> xv; xv.load(x + base + v * C); xv.unpack(fx);
      if (!UP) {
        Vec<T> g; g.load(gp + v * C); g.unpack(t);
      } else {
        const int iw = (int)(v % in.w), ih = (int)((v / in.w) % in.h), id = (int)(v / ((int64_t)in.w * in.h));
#pragma unroll
        for (int i = 0; i < VEC; i++) t[i] = 0.f;
        for (int a = 0; a < 4; a++) {
          const int od = min(max(2 * id - 1 + a, 0), OD - 1);
          const float wa = (a == 0 || a == 3) ? .25f : .75f;
          for (int b = 0; b < 4; b++) {
            const int oh = min(max(2 * ih - 1 + b, 0), OH - 1);
            const float wb = wa * ((b == 0 || b == 3) ? .25f : .75f);
            const T* row = gp + (((int64_t)od * OH + oh) * OW) * C;
#pragma unroll
            for (int e = 0; e < 4; e++) {
              const int ow = min(max(2 * iw - 1 + e, 0), OW - 1);
              const float we = wb * ((e == 0 || e == 3) ? .25f : .75f);
              Vec<T> g;
              float fg[VEC];
              g.load(row + (int64_t)ow * C);
              g.unpack(fg);
#pragma unroll
              for (int i = 0; i < VEC; i++) t[i] = fmaf(we, fg[i], t[i]);
            }
          }
        }
        Vec<T> o; o.pack(t); o.store(tbuf + base + v * C);
      }
#pragma unroll
      for (int i = 0; i < VEC; i++) { a1[i] += t[i]; a2[i] = fmaf(t[i], fx[i], a2[i]); }
    }
  }
  lanes_sum<VEC>(a1, sm, lane, cvec, G);
  lanes_sum<VEC>(a2, sm, lane, cvec, G);
  if (lane == 0) {
    const int64_t o = ((n * G.chunks + blockIdx.x) * (int64_t)C) + cvec * VEC;
#pragma unroll
    for (int i = 0; i < VEC; i++) { p1[o + i] = a1[i]; p2[o + i] = a2[i]; }
  }
}

// backward pass 2: dx = ct[n,c] * t + cx[n,c] * x + c0[n,c]   (t may alias dx)
template <typename T>
__global__ void __launch_bounds__(kThr) adain_bwd_apply_kernel(const T* __restrict__ x, const T* t,
                                                                const float* __restrict__ ct, const float* __restrict__ cx,
                                                                const float* __restrict__ c0, T* dx, int64_t V, int C,
                                                                ModGrid G) {
  constexpr int VEC = Vec<T>::N;
  const int lane = threadIdx.x / G.cv, cvec = threadIdx.x % G.cv;
  if (lane >= G.rows) return;
  const int64_t n = blockIdx.y;
  float k1[VEC], k2[VEC], k3[VEC];
#pragma unroll
  for (int i = 0; i < VEC; i++) {
    const int64_t o = n * C + cvec * VEC + i;
    k1[i] = ct[o]; k2[i] = cx[o]; k3[i] = c0[o];
  }
  const int64_t v0 = (int64_t)blockIdx.x * G.vpc, v1 = min(V, v0 + G.vpc);
  const int64_t base = (n * V) * C + cvec * VEC;
  for (int64_t v = v0 + lane; v < v1; v += G.rows) {
    Vec<T> a, b;
    float fx[VEC], ft[VEC];
    a.load(x + base + v * C); b.load(t + base + v * C);
    a.unpack(fx); b.unpack(ft);
#pragma unroll
    for (int i = 0; i < VEC; i++) ft[i] = fmaf(k1[i], ft[i], fmaf(k2[i], fx[i], k3[i]));
    b.pack(ft); b.store(dx + base + v * C);
  }
}

template <typename T>
int inst_stats_t(const void* x, float* s1, float* s2, int64_t N, int64_t V, int64_t C, void* ws, size_t wsb,
                 cudaStream_t st) {
  ModGrid G;
  if (!mod_grid<T>(N, V, C, &G)) return B2C_ERR_UNSUPPORTED;
  const size_t np = (size_t)N * G.chunks * C;
  if (wsb < 2 * np * sizeof(float)) return B2C_ERR_WORKSPACE;
  float* p1 = static_cast<float*>(ws);
  float* p2 = p1 + np;
  inst_stats_kernel<T><<<dim3(G.chunks, (unsigned)N), kThr, 0, st>>>(static_cast<const T*>(x), p1, p2, V, (int)C, G);
  int rc = launched();
  if (rc) return rc;
  if ((rc = sum_partials(p1, s1, N, G.chunks, (int)C, st))) return rc;
  return sum_partials(p2, s2, N, G.chunks, (int)C, st);
}

template <typename T>
int adain_fwd_t(const void* x, const float* a, const float* b, void* y, int64_t N, Vol in, int64_t C, int up,
                cudaStream_t st) {
  ModGrid G;
  const int64_t V = (int64_t)in.d * in.h * in.w * (up ? 8 : 1);
  if (!mod_grid<T>(N, V, C, &G)) return B2C_ERR_UNSUPPORTED;
  dim3 grid(G.chunks, (unsigned)N);
  if (up) adain_apply_kernel<T, true><<<grid, kThr, 0, st>>>(static_cast<const T*>(x), a, b, static_cast<T*>(y), in, V, (int)C, G);
  else adain_apply_kernel<T, false><<<grid, kThr, 0, st>>>(static_cast<const T*>(x), a, b, static_cast<T*>(y), in, V, (int)C, G);
  return launched();
}

template <typename T>
int adain_bwd_reduce_t(const void* x, const void* gy, void* tbuf, float* r1, float* r2, int64_t N, Vol in, int64_t C,
                       int up, void* ws, size_t wsb, cudaStream_t st) {
  ModGrid G;
  const int64_t V = (int64_t)in.d * in.h * in.w;
  if (!mod_grid<T>(N, V, C, &G)) return B2C_ERR_UNSUPPORTED;
  const size_t np = (size_t)N * G.chunks * C;
  if (wsb < 2 * np * sizeof(float)) return B2C_ERR_WORKSPACE;
  float* p1 = static_cast<float*>(ws);
  float* p2 = p1 + np;
  dim3 grid(G.chunks, (unsigned)N);
  if (up) adain_bwd_reduce_kernel<T, true><<<grid, kThr, 0, st>>>(static_cast<const T*>(x), static_cast<const T*>(gy),
                                                                  static_cast<T*>(tbuf), p1, p2, in, V, (int)C, G);
  else adain_bwd_reduce_kernel<T, false><<<grid, kThr, 0, st>>>(static_cast<const T*>(x), static_cast<const T*>(gy),
                                                                nullptr, p1, p2, in, V, (int)C, G);
  int rc = launched();
  if (rc) return rc;
  if ((rc = sum_partials(p1, r1, N, G.chunks, (int)C, st))) return rc;
  return sum_partials(p2, r2, N, G.chunks, (int)C, st);
}

template <typename T>
int adain_bwd_apply_t(const void* x, const void* t, const float* ct, const float* cx, const float* c0, void* dx,
                      int64_t N, int64_t V, int64_t C, cudaStream_t st) {
  ModGrid G;
  if (!mod_grid<T>(N, V, C, &G)) return B2C_ERR_UNSUPPORTED;
  adain_bwd_apply_kernel<T><<<dim3(G.chunks, (unsigned)N), kThr, 0, st>>>(static_cast<const T*>(x), static_cast<const T*>(t),
                                                                          ct, cx, c0, static_cast<T*>(dx), V, (int)C, G);
  return launched();
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

extern "C" int b2c_instance_stats(const void* x, float* sum, float* sqsum, int64_t N, int64_t V, int64_t C, int dt,
                                  void* ws, size_t wsb, b2c_stream st) {
  if (!x || !sum || !sqsum || !ws || !dt_ok(dt)) return B2C_ERR_INVALID;
  cudaStream_t cs = static_cast<cudaStream_t>(st);
  return dt == B2C_BF16 ? inst_stats_t<__nv_bfloat16>(x, sum, sqsum, N, V, C, ws, wsb, cs)
                        : inst_stats_t<float>(x, sum, sqsum, N, V, C, ws, wsb, cs);
}

extern "C" int b2c_adain_fwd(const void* x, const float* a, const float* b, void* y, int64_t N, int d, int h, int w,
                             int64_t C, int dt, int up, b2c_stream st) {
  if (!x || !a || !b || !y || !dt_ok(dt) || d <= 0 || h <= 0 || w <= 0) return B2C_ERR_INVALID;
  cudaStream_t cs = static_cast<cudaStream_t>(st);
  Vol in{d, h, w};
  return dt == B2C_BF16 ? adain_fwd_t<__nv_bfloat16>(x, a, b, y, N, in, C, up, cs)
                        : adain_fwd_t<float>(x, a, b, y, N, in, C, up, cs);
}

extern "C" int b2c_adain_bwd_reduce(const void* x, const void* gy, void* t, float* r1, float* r2, int64_t N, int d,
                                    int h, int w, int64_t C, int dt, int up, void* ws, size_t wsb, b2c_stream st) {
  if (!x || !gy || !r1 || !r2 || !ws || (up && !t) || !dt_ok(dt) || d <= 0 || h <= 0 || w <= 0) return B2C_ERR_INVALID;
  cudaStream_t cs = static_cast<cudaStream_t>(st);
  Vol in{d, h, w};
  return dt == B2C_BF16 ? adain_bwd_reduce_t<__nv_bfloat16>(x, gy, t, r1, r2, N, in, C, up, ws, wsb, cs)
                        : adain_bwd_reduce_t<float>(x, gy, t, r1, r2, N, in, C, up, ws, wsb, cs);
}

extern "C" int b2c_adain_bwd_apply(const void* x, const void* t, const float* ct, const float* cx, const float* c0,
                                   void* dx, int64_t N, int64_t V, int64_t C, int dt, b2c_stream st) {
  if (!x || !t || !ct || !cx || !c0 || !dx || !dt_ok(dt)) return B2C_ERR_INVALID;
  cudaStream_t cs = static_cast<cudaStream_t>(st);
  return dt == B2C_BF16 ? adain_bwd_apply_t<__nv_bfloat16>(x, t, ct, cx, c0, dx, N, V, C, cs)
                        : adain_bwd_apply_t<float>(x, t, ct, cx, c0, dx, N, V, C, cs);
}
