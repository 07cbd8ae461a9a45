// thin.cu — see thin.cuh.
#include "thin.cuh"

#include <cstdlib>

#include "tc_conv.cuh"

namespace b2c {

typedef __nv_bfloat16 bf16;

// ---------------------------------------------------------------------------
// im2col of a 1-channel tensor: col[v][t] = big[n, o*s + k*dl - p] (0 outside),
// t = (kd*k1 + kh)*k2 + kw, zero for t in [T, Kp).  One thread per (v, kd): it
// produces the k1*k2 taps of one kd plane, contiguous in the col row, so a warp
// writes whole 128-byte lines.  KHW > 0: compile-time plane size with 16-byte
// stores (k1*k2 % 8 == 0); KHW == 0: any size, scalar stores.
template <int KHW>
__global__ void __launch_bounds__(256)
im2col1_kernel(Geo g, const bf16* __restrict__ big, bf16* __restrict__ col, int Kp, int64_t total) {
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= total) return;
  const int kd = (int)(idx % g.k[0]);
  int64_t v = idx / g.k[0];
  const int64_t row = v;
  const int ow = (int)(v % g.sw); v /= g.sw;
  const int oh = (int)(v % g.sh); v /= g.sh;
  const int od = (int)(v % g.sd);
  const int n = (int)(v / g.sd);
  const int khw = g.k[1] * g.k[2];
  bf16* dst = col + row * Kp + (int64_t)kd * khw;
  const int id = od * g.s[0] + kd * g.dl[0] - g.p[0];
  const bool dok = id >= 0 && id < g.bd;
  const bf16* plane = big + ((int64_t)n * g.bd + (dok ? id : 0)) * g.bh * g.bw;
  const int ih0 = oh * g.s[1] - g.p[1], iw0 = ow * g.s[2] - g.p[2];
  const bf16 zero = __float2bfloat16_rn(0.f);
  if (KHW > 0) {
    __align__(16) bf16 buf[KHW > 0 ? KHW : 8];
    int j = 0;
    for (int kh = 0; kh < g.k[1]; kh++) {
      const int ih = ih0 + kh * g.dl[1];
      const bool hok = dok && ih >= 0 && ih < g.bh;
      const bf16* rowp = plane + (int64_t)(hok ? ih : 0) * g.bw;
#pragma unroll 4
      for (int kw = 0; kw < g.k[2]; kw++, j++) {
        const int iw = iw0 + kw * g.dl[2];
        buf[j < KHW ? j : 0] = (hok && iw >= 0 && iw < g.bw) ? rowp[iw] : zero;
      }
    }
#pragma unroll
    for (int q = 0; q < KHW / 8; q++)
      reinterpret_cast<uint4*>(dst)[q] = reinterpret_cast<const uint4*>(buf)[q];
  } else {
    int j = 0;
    for (int kh = 0; kh < g.k[1]; kh++) {
      const int ih = ih0 + kh * g.dl[1];
      const bool hok = dok && ih >= 0 && ih < g.bh;
      const bf16* rowp = plane + (int64_t)(hok ? ih : 0) * g.bw;
      for (int kw = 0; kw < g.k[2]; kw++, j++) {
        const int iw = iw0 + kw * g.dl[2];
        dst[j] = (hok && iw >= 0 && iw < g.bw) ? rowp[iw] : zero;
      }
    }
  }
  if (kd == 0) {
    bf16* r = col + row * Kp;
    for (int t = g.k[0] * khw; t < Kp; t++) r[t] = zero;
  }
}

// ---------------------------------------------------------------------------
// col2im onto a 1-channel tensor (a gather over the taps that divide):
//   big[i] = act(bias + sum_t [s | (i + p - k*dl)] col[(i + p - k*dl)/s][t])
__global__ void __launch_bounds__(256)
col2im1_kernel(Geo g, const bf16* __restrict__ col, bf16* __restrict__ big, int Kp, const float* __restrict__ bias,
               int act, float slope, int64_t total) {
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= total) return;
  int64_t v = idx;
  const int iw = (int)(v % g.bw); v /= g.bw;
  const int ih = (int)(v % g.bh); v /= g.bh;
  const int id = (int)(v % g.bd);
  const int n = (int)(v / g.bd);
  float acc = bias ? bias[0] : 0.f;
  for (int kd = 0; kd < g.k[0]; kd++) {
    const int nd = id + g.p[0] - kd * g.dl[0];
    if (nd < 0 || nd % g.s[0]) continue;
    const int od = nd / g.s[0];
    if (od >= g.sd) continue;
    for (int kh = 0; kh < g.k[1]; kh++) {
      const int nh = ih + g.p[1] - kh * g.dl[1];
      if (nh < 0 || nh % g.s[1]) continue;
      const int oh = nh / g.s[1];
      if (oh >= g.sh) continue;
      const bf16* base = col + ((((int64_t)n * g.sd + od) * g.sh + oh) * g.sw) * Kp + (kd * g.k[1] + kh) * g.k[2];
      for (int kw = 0; kw < g.k[2]; kw++) {
        const int nw = iw + g.p[2] - kw * g.dl[2];
        if (nw < 0 || nw % g.s[2]) continue;
        const int ow = nw / g.s[2];
        if (ow >= g.sw) continue;
        acc += __bfloat162float(base[(int64_t)ow * Kp + kw]);
      }
    }
  }
  big[idx] = __float2bfloat16_rn(act_f(acc, act, slope));
}

// ---------------------------------------------------------------------------
// Fast path for k = 4, stride 2, pad 1, dilation 1 (dcgan.py:67 and the data gradient of
// dcgan.py:161).  Output voxel i = 2j + r (per axis) receives exactly two taps per axis:
//   a = 0: small voxel j,            tap k = 1 + r
//   a = 1: small voxel j - 1 + 2r,   tap k = 3 (r = 0) or 0 (r = 1)
// The GEMM writes its 64 columns in the order t' = a*8 + r (a, r = 3-bit (d,h,w) codes), so
// that one thread per small voxel j gathers the 2x2x2 output block it owns from the 27
// neighbouring col rows (L1-resident reuse) and writes coalesced pairs along w.
__host__ __device__ inline int k4s2_tap(int r, int a) { return r == 0 ? 1 + 2 * a : 2 - 2 * a; }
__host__ __device__ inline int k4s2_perm(int tp) {          // t' -> torch tap index (kd*4+kh)*4+kw
  const int a = tp >> 3, r = tp & 7;
  const int kd = k4s2_tap((r >> 2) & 1, (a >> 2) & 1), kh = k4s2_tap((r >> 1) & 1, (a >> 1) & 1),
            kw = k4s2_tap(r & 1, a & 1);
  return (kd * 4 + kh) * 4 + kw;
}

__device__ __forceinline__ float bf_lo(uint32_t v) { return __uint_as_float(v << 16); }
__device__ __forceinline__ float bf_hi(uint32_t v) { return __uint_as_float(v & 0xffff0000u); }

// One thread per small voxel j, a warp = 32 consecutive jw.  Per (dd, dh) neighbour line each
// lane reads ONE 32-byte sector of its own col row (groups a0 = (ad,ah,0) and a0+1 = (ad,ah,1));
// the w-neighbours' contributions (group a0+1 of rows jw-1 / jw+1) come through warp shuffles,
// only lanes 0 / 31 fetch them from memory.
__global__ void __launch_bounds__(256)
col2im1_k4s2_kernel(const bf16* __restrict__ col, bf16* __restrict__ big, int n, int sd, int sh, int sw,
                    const float* __restrict__ bias, int act, float slope) {
  const int lane = threadIdx.x & 31;
  const int jw = blockIdx.x * 64 + (threadIdx.x & 63);
  const int jh = blockIdx.y * 4 + (threadIdx.x >> 6);
  const int jd = blockIdx.z % sd, in = blockIdx.z / sd;
  const bool live = jw < sw && jh < sh;          // jh is warp-uniform
  float acc[8];
  const float b0 = bias ? bias[0] : 0.f;
#pragma unroll
  for (int r = 0; r < 8; r++) acc[r] = b0;
  const uint4 zero4 = make_uint4(0u, 0u, 0u, 0u);
#pragma unroll
  for (int dd = -1; dd <= 1; dd++) {
    const int od = jd + dd;
    if (od < 0 || od >= sd) continue;              // block-uniform
#pragma unroll
    for (int dh = -1; dh <= 1; dh++) {
      const int oh = jh + dh;
      if (oh < 0 || oh >= sh || jh >= sh) continue;  // warp-uniform
      const int a0 = ((dd != 0) << 2) | ((dh != 0) << 1);
      const bf16* line = col + (((int64_t)in * sd + od) * sh + oh) * (int64_t)sw * 64 + a0 * 8;
      uint4 A0 = zero4, A1 = zero4;
      if (live) {
        const uint4* p = reinterpret_cast<const uint4*>(line + (int64_t)jw * 64);
        A0 = __ldg(p); A1 = __ldg(p + 1);
      }
      uint4 L, R;                                   // group a0+1 of rows jw-1 and jw+1
      L.x = __shfl_up_sync(0xffffffffu, A1.x, 1); L.y = __shfl_up_sync(0xffffffffu, A1.y, 1);
      L.z = __shfl_up_sync(0xffffffffu, A1.z, 1); L.w = __shfl_up_sync(0xffffffffu, A1.w, 1);
      R.x = __shfl_down_sync(0xffffffffu, A1.x, 1); R.y = __shfl_down_sync(0xffffffffu, A1.y, 1);
      R.z = __shfl_down_sync(0xffffffffu, A1.z, 1); R.w = __shfl_down_sync(0xffffffffu, A1.w, 1);
      if (lane == 0) L = (live && jw > 0) ? __ldg(reinterpret_cast<const uint4*>(line + (int64_t)(jw - 1) * 64) + 1) : zero4;
      if (lane == 31) R = (live && jw + 1 < sw) ? __ldg(reinterpret_cast<const uint4*>(line + (int64_t)(jw + 1) * 64) + 1) : zero4;
      if (jw + 1 >= sw) R = zero4;
      const uint32_t a0v[4] = {A0.x, A0.y, A0.z, A0.w}, lv[4] = {L.x, L.y, L.z, L.w}, rv[4] = {R.x, R.y, R.z, R.w};
#pragma unroll
      for (int m = 0; m < 4; m++) {                 // m = rd*2 + rh; register m holds (rw = 0 | rw = 1)
        const bool ok = (dd == 0 || ((m >> 1) & 1) == (dd > 0)) && (dh == 0 || (m & 1) == (dh > 0));
        if (ok) {
          acc[2 * m] += bf_lo(a0v[m]) + bf_lo(lv[m]);           // rw = 0: own row (k=1) + row jw-1 (k=3)
          acc[2 * m + 1] += bf_hi(a0v[m]) + bf_hi(rv[m]);       // rw = 1: own row (k=2) + row jw+1 (k=0)
        }
      }
    }
  }
  if (!live) return;
  const int bh = 2 * sh, bw = 2 * sw;
#pragma unroll
  for (int rd = 0; rd < 2; rd++)
#pragma unroll
    for (int rh = 0; rh < 2; rh++) {
      const float v0 = act_f(acc[rd * 4 + rh * 2], act, slope), v1 = act_f(acc[rd * 4 + rh * 2 + 1], act, slope);
      __nv_bfloat162 o = __floats2bfloat162_rn(v0, v1);
      bf16* dst = big + (((int64_t)in * 2 * sd + 2 * jd + rd) * bh + 2 * jh + rh) * bw + 2 * jw;
      *reinterpret_cast<__nv_bfloat162*>(dst) = o;
    }
}

static bool is_k4s2(const Geo& g) {
  for (int i = 0; i < 3; i++) if (g.k[i] != 4 || g.s[i] != 2 || g.p[i] != 1 || g.dl[i] != 1) return false;
  return g.bd == 2 * g.sd && g.bh == 2 * g.sh && g.bw == 2 * g.sw && (int64_t)g.n * g.sd <= 65535;
}

// packed filters --------------------------------------------------------------
// gather : wp[a][t]  (Kp columns, zero padded)   = w[a][t]
// scatter: wp[t][a]  (Kp rows,    zero padded)   = w[a][t]
__global__ void __launch_bounds__(256)
thin_pack_kernel(const float* __restrict__ w, bf16* __restrict__ wp, int A, int T, int Kp, int gather, int k4s2,
                 int flip) {
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (int64_t)A * Kp) return;
  int a, t;
  if (gather) { a = (int)(idx / Kp); t = (int)(idx % Kp); }
  else        { t = (int)(idx / A);  a = (int)(idx % A); if (k4s2) t = k4s2_perm(t); }
  // flip: the mirrored problem (thin_mode == 2) sees the filter with all three axes reversed
  wp[idx] = __float2bfloat16_rn(t < T ? w[(int64_t)a * T + (flip ? T - 1 - t : t)] : 0.f);
}

// dw[a][t] = tmp[a][t] for t < T  (tmp has Kp columns)
__global__ void __launch_bounds__(256)
thin_unpad_kernel(const float* __restrict__ tmp, float* __restrict__ dw, int A, int T, int Kp, int flip) {
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (int64_t)A * T) return;
  const int t = (int)(idx % T);
  dw[idx] = tmp[(idx / T) * Kp + (flip ? T - 1 - t : t)];
}

// ---------------------------------------------------------------------------
static bool thin_shape_ok(const Geo& g) {
  if (g.bc != 1 || g.sc % 8 != 0 || g.sc < 16) return false;
  const int64_t rows = (int64_t)g.n * g.sd * g.sh * g.sw;
  return g.taps() <= 1024 && rows < 0x7fffffffLL && g.k[1] * g.k[2] <= 64;
}

int thin_mode(const b2c_conv_desc* c, const Geo& g, Geo* tg) {
  if (c->io_dtype != B2C_BF16 || c->algo == B2C_ALGO_DIRECT) return 0;
  if (thin_shape_ok(g)) { *tg = g; return 1; }
  // one channel on the SMALL side with unit stride (biggan.py:42 output conv f -> 1,
  // stylegan.py:1038): i = o + k*dl - p  <=>  o = i + (K-1-k)*dl - ((K-1)*dl - p), i.e. the same
  // problem with the two tensors' roles exchanged, the filter reversed and the padding mirrored.
  if (g.sc != 1) return 0;
  Geo m = g;
  m.sc = g.bc; m.sd = g.bd; m.sh = g.bh; m.sw = g.bw;
  m.bc = 1;    m.bd = g.sd; m.bh = g.sh; m.bw = g.sw;
  for (int i = 0; i < 3; i++) {
    if (g.s[i] != 1) return 0;
    m.p[i] = (g.k[i] - 1) * g.dl[i] - g.p[i];
    if (m.p[i] < 0) return 0;
  }
  if (!thin_shape_ok(m)) return 0;
  *tg = m;
  return 2;
}
int thin_kp(const Geo& g) { int kp = (g.taps() + 7) / 8 * 8; return kp < 16 ? 16 : kp; }
size_t thin_packed_bytes(const Geo& g) { return (size_t)thin_kp(g) * g.sc * sizeof(bf16); }

int thin_pack(const Geo& g, bool gather, const float* w, void* wp, cudaStream_t st, bool flip) {
  const int Kp = thin_kp(g);
  const int64_t n = (int64_t)g.sc * Kp;
  thin_pack_kernel<<<(unsigned)cdiv(n, 256), 256, 0, st>>>(w, static_cast<bf16*>(wp), g.sc, g.taps(), Kp,
                                                          gather ? 1 : 0, is_k4s2(g) ? 1 : 0, flip ? 1 : 0);
  return launched();
}

static int64_t rows_of(const Geo& g) { return (int64_t)g.n * g.sd * g.sh * g.sw; }
static size_t align256(size_t v) { return (v + 255) & ~(size_t)255; }
static size_t col_bytes(const Geo& g) { return align256((size_t)rows_of(g) * thin_kp(g) * sizeof(bf16)); }

// the GEMM [rows][kin] x [nout][kin]^T as a 1x1x1 convolution over a 1 x 1 x 1 x rows lattice
static Geo gemm_geo(int64_t rows, int nout, int kin) {
  Geo p;
  p.n = 1;
  p.sc = nout; p.sd = 1; p.sh = 1; p.sw = (int)rows;
  p.bc = kin;  p.bd = 1; p.bh = 1; p.bw = (int)rows;
  for (int i = 0; i < 3; i++) { p.k[i] = 1; p.s[i] = 1; p.p[i] = 0; p.dl[i] = 1; }
  return p;
}

static bool is_pointwise(const Geo& g);
static int pw_wgrad_blocks(int64_t rows);

size_t thin_workspace(const Geo& g, int kind) {
  size_t b = col_bytes(g);
  if (kind == 2 && is_pointwise(g)) {          // partial column sums of the pointwise weight gradient
    const size_t pw = align256((size_t)pw_wgrad_blocks(rows_of(g)) * g.sc * sizeof(float));
    if (pw > b) b = pw;
  }
  if (kind == 2) {
    b += align256((size_t)g.sc * thin_kp(g) * sizeof(float));
    b += tc_wgrad_workspace(gemm_geo(rows_of(g), g.sc, thin_kp(g)));
    const size_t fused = tc_thin_wgrad_workspace(g);
    if (fused > b) b = fused;
  }
  return b;
}

static int im2col_launch(const Geo& g, const void* big, void* col, cudaStream_t st) {
  const int Kp = thin_kp(g);
  const int64_t total = rows_of(g) * g.k[0];
  if (total == 0) return B2C_OK;
  const int64_t blocks = cdiv(total, 256);
  if (blocks > 0x7fffffffLL) return B2C_ERR_UNSUPPORTED;
  const bf16* b = static_cast<const bf16*>(big);
  bf16* c = static_cast<bf16*>(col);
  const int khw = g.k[1] * g.k[2];
  if (khw == 16) im2col1_kernel<16><<<(unsigned)blocks, 256, 0, st>>>(g, b, c, Kp, total);
  else if (khw == 8) im2col1_kernel<8><<<(unsigned)blocks, 256, 0, st>>>(g, b, c, Kp, total);
  else im2col1_kernel<0><<<(unsigned)blocks, 256, 0, st>>>(g, b, c, Kp, total);
  return launched();
}

// ---------------------------------------------------------------------------
// 1x1x1 filters on a one-channel tensor (fromrgb / torgb, stylegan.py:542-546,758; their gradients):
// no reduction over taps is left, the three passes are an outer product, a row dot product and a
// weighted column sum - pure HBM streams.  Through the K = 16 GEMM lowering they ran at 1 TB/s.
static bool is_pointwise(const Geo& g) {
  for (int i = 0; i < 3; i++) if (g.k[i] != 1 || g.s[i] != 1 || g.p[i] != 0) return false;
  return g.sd == g.bd && g.sh == g.bh && g.sw == g.bw && g.sc % 8 == 0 && g.sc <= 512;
}

// out[v][a] = act(big[v] * w[a] + bias[a]); w[a] = wp[a * wstride].  A thread owns one 8-channel group (its
// weights and bias stay in registers) and walks the rows of its block's slab, 256 / (sc/8) rows per pass.
__global__ void __launch_bounds__(256)
pw_outer_kernel(const bf16* __restrict__ big, const bf16* __restrict__ wp, int wstride, const float* __restrict__ bias,
                bf16* __restrict__ out, int sc, int act, float slope, int64_t rows, int64_t rows_per_block) {
  const int cv = sc >> 3;
  const int lanes = cv < 256 ? cv : 256, rstep = 256 / lanes;
  const int lane = threadIdx.x % lanes, rl = threadIdx.x / lanes;
  if (rl >= rstep) return;
  const int64_t r0 = (int64_t)blockIdx.x * rows_per_block;
  const int64_t r1 = r0 + rows_per_block < rows ? r0 + rows_per_block : rows;
  for (int c0 = lane * 8; c0 < sc; c0 += lanes * 8) {
    float w[8], bb[8];
#pragma unroll
    for (int i = 0; i < 8; i++) {
      w[i] = __bfloat162float(wp[(int64_t)(c0 + i) * wstride]);
      bb[i] = bias ? bias[c0 + i] : 0.f;
    }
    for (int64_t rb = r0 + rl; rb < r1; rb += 4 * (int64_t)rstep) {
      float x[4];
#pragma unroll
      for (int u = 0; u < 4; u++) {
        const int64_t r = rb + (int64_t)u * rstep;
        x[u] = r < r1 ? __bfloat162float(big[r]) : 0.f;
      }
#pragma unroll
      for (int u = 0; u < 4; u++) {
        const int64_t r = rb + (int64_t)u * rstep;
        if (r >= r1) break;
        float o[8];
#pragma unroll
        for (int i = 0; i < 8; i++) o[i] = act_f(fmaf(x[u], w[i], bb[i]), act, slope);
        Vec<bf16> pk; pk.pack(o); pk.store(out + r * sc + c0);
      }
    }
  }
}

// out[v] = act(bias[0] + sum_a small[v][a] * w[a]); LPR lanes share a row (power of two <= 32, sc <= 512: at most
// two 8-channel groups per lane, weights in registers); a thread has four rows in flight.
__global__ void __launch_bounds__(256)
pw_rowdot_kernel(const bf16* __restrict__ small, const bf16* __restrict__ wp, int wstride, const float* __restrict__ bias,
                 bf16* __restrict__ out, int sc, int lpr, int act, float slope, int64_t rows, int64_t rows_per_block) {
  const int l = threadIdx.x % lpr, rl = threadIdx.x / lpr, rstep = 256 / lpr;
  const int64_t r0 = (int64_t)blockIdx.x * rows_per_block;
  const int64_t r1 = r0 + rows_per_block < rows ? r0 + rows_per_block : rows;
  const int ca = l * 8, cb = ca + lpr * 8;
  const bool ha = ca < sc, hb = cb < sc;
  float wa[8], wb[8];
#pragma unroll
  for (int i = 0; i < 8; i++) {
    wa[i] = ha ? __bfloat162float(wp[(int64_t)(ca + i) * wstride]) : 0.f;
    wb[i] = hb ? __bfloat162float(wp[(int64_t)(cb + i) * wstride]) : 0.f;
  }
  const float b0 = bias ? bias[0] : 0.f;
  // block-uniform trip count (the row offset rl is added inside): every lane reaches the full-mask shuffles
  for (int64_t base = r0; base < r1; base += 4 * (int64_t)rstep) {
    const int64_t rb = base + rl;
    Vec<bf16> pa[4], pb[4];
#pragma unroll
    for (int u = 0; u < 4; u++) {
      const int64_t r = rb + (int64_t)u * rstep;
      pa[u].zero(); pb[u].zero();
      if (r < r1) {
        if (ha) pa[u].load(small + r * sc + ca);
        if (hb) pb[u].load(small + r * sc + cb);
      }
    }
#pragma unroll
    for (int u = 0; u < 4; u++) {
      const int64_t r = rb + (int64_t)u * rstep;
      float xa[8], xb[8];
      pa[u].unpack(xa); pb[u].unpack(xb);
      float acc = 0.f;
#pragma unroll
      for (int i = 0; i < 8; i++) acc = fmaf(xa[i], wa[i], fmaf(xb[i], wb[i], acc));
      for (int o = lpr >> 1; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
      if (r < r1 && l == 0) out[r] = __float2bfloat16_rn(act_f(acc + b0, act, slope));
    }
  }
}

// part[b][a] = sum over the rows of slab b of small[v][a] * big[v]   (deterministic two-stage column sum)
__global__ void __launch_bounds__(256)
pw_wgrad_stage1(const bf16* __restrict__ small, const bf16* __restrict__ big, float* __restrict__ part, int sc,
                int64_t rows, int64_t rows_per_block) {
  extern __shared__ float pw_sm[];           // [256][8]
  const int cv = sc >> 3;
  const int lanes = cv < 256 ? cv : 256, rstep = 256 / lanes;
  const int lane = threadIdx.x % lanes, rl = threadIdx.x / lanes;
  const int64_t r0 = (int64_t)blockIdx.x * rows_per_block;
  const int64_t r1 = r0 + rows_per_block < rows ? r0 + rows_per_block : rows;
  for (int g0 = 0; g0 < cv; g0 += lanes) {
    const int c0 = (g0 + lane) * 8;
    float a[8];
#pragma unroll
    for (int i = 0; i < 8; i++) a[i] = 0.f;
    if (g0 + lane < cv && rl < rstep)
      for (int64_t rb = r0 + rl; rb < r1; rb += 8 * (int64_t)rstep) {
        Vec<bf16> pk[8]; float xs[8];
#pragma unroll
        for (int u = 0; u < 8; u++) {
          const int64_t r = rb + (int64_t)u * rstep;
          if (r < r1) { pk[u].load(small + r * sc + c0); xs[u] = __bfloat162float(big[r]); }
        }
#pragma unroll
        for (int u = 0; u < 8; u++) {
          if (rb + (int64_t)u * rstep >= r1) break;
          float y[8]; pk[u].unpack(y);
#pragma unroll
          for (int i = 0; i < 8; i++) a[i] = fmaf(y[i], xs[u], a[i]);
        }
      }
    float* my = pw_sm + threadIdx.x * 8;
#pragma unroll
    for (int i = 0; i < 8; i++) my[i] = a[i];
    __syncthreads();
    if (rl == 0 && g0 + lane < cv) {
      for (int j = 1; j < rstep; j++) {
        const float* o = pw_sm + (j * lanes + lane) * 8;
#pragma unroll
        for (int i = 0; i < 8; i++) a[i] += o[i];
      }
#pragma unroll
      for (int i = 0; i < 8; i++) part[(int64_t)blockIdx.x * sc + c0 + i] = a[i];
    }
    __syncthreads();
  }
}
__global__ void __launch_bounds__(256)
pw_wgrad_stage2(const float* __restrict__ part, int nblocks, int sc, float* __restrict__ dw) {
  const int a = blockIdx.x * blockDim.x + threadIdx.x;
  if (a >= sc) return;
  double acc = 0.0;
  for (int b = 0; b < nblocks; b++) acc += part[(int64_t)b * sc + a];
  dw[a] = (float)acc;
}
static int pw_wgrad_blocks(int64_t rows) {
  int64_t nb = (int64_t)num_sms() * 4;
  if (nb > rows) nb = rows;
  return (int)(nb < 1 ? 1 : nb);
}

int thin_gather(const Geo& g, const void* big, const void* wp, void* out_small, const float* bias, int act,
                float slope, void* ws, size_t wsb, cudaStream_t st, const TcStats* stats) {
  if (is_pointwise(g)) {
    const int64_t rows = rows_of(g);
    if (rows == 0) return B2C_OK;
    const int nb = pw_wgrad_blocks(rows) * 2;
    pw_outer_kernel<<<nb, 256, 0, st>>>(static_cast<const bf16*>(big), static_cast<const bf16*>(wp), thin_kp(g), bias,
                                        static_cast<bf16*>(out_small), g.sc, act, slope, rows, cdiv(rows, (int64_t)nb));
    return launched();
  }
  if (tc_thin_gather_supported(g) && !knobs().thin_unfused) {
    const int rc = tc_thin_gather_launch(g, big, wp, thin_kp(g), out_small, bias, act, slope, st, stats);
    if (rc != B2C_ERR_UNSUPPORTED) return rc;
  }
  if (!ws || wsb < thin_workspace(g, 0)) return B2C_ERR_WORKSPACE;
  int rc = im2col_launch(g, big, ws, st);
  if (rc) return rc;
  ConvPlan plan;
  rc = plan_gather(gemm_geo(rows_of(g), g.sc, thin_kp(g)), &plan);
  if (rc) return rc;
  return tc_conv_launch(plan, ws, wp, out_small, bias, 0, act, slope, st);
}

int thin_scatter(const Geo& g, const void* small, const void* wp, void* out_big, const float* bias, int act,
                 float slope, void* ws, size_t wsb, cudaStream_t st) {
  if (is_pointwise(g)) {
    const int64_t rows = rows_of(g);
    if (rows == 0) return B2C_OK;
    int lpr = 1;
    while (2 * lpr < g.sc / 8 && lpr < 32) lpr <<= 1;        // one or two 8-channel groups per lane (sc <= 512)
    const int nb = pw_wgrad_blocks(rows) * 2;
    // scatter packing: wp[t][a], one tap -> w[a] = wp[a]
    pw_rowdot_kernel<<<nb, 256, 0, st>>>(static_cast<const bf16*>(small), static_cast<const bf16*>(wp), 1, bias,
                                         static_cast<bf16*>(out_big), g.sc, lpr, act, slope, rows, cdiv(rows, (int64_t)nb));
    return launched();
  }
  if (!ws || wsb < thin_workspace(g, 1)) return B2C_ERR_WORKSPACE;
  const int Kp = thin_kp(g);
  ConvPlan plan;
  int rc = plan_gather(gemm_geo(rows_of(g), Kp, g.sc), &plan);
  if (rc) return rc;
  rc = tc_conv_launch(plan, small, wp, ws, nullptr, 0, B2C_ACT_NONE, 0.f, st);
  if (rc) return rc;
  const int64_t total = (int64_t)g.n * g.bd * g.bh * g.bw;
  if (total == 0) return B2C_OK;
  if (is_k4s2(g) && !knobs().col2im_ldg)
    return tc_col2im1_k4s2_launch(g, ws, out_big, bias, act, slope, st);      // TMA-staged (tc_thin.cuh)
  if (is_k4s2(g)) {
    dim3 grid((unsigned)cdiv(g.sw, 64), (unsigned)cdiv(g.sh, 4), (unsigned)(g.n * g.sd));
    col2im1_k4s2_kernel<<<grid, 256, 0, st>>>(static_cast<const bf16*>(ws), static_cast<bf16*>(out_big), g.n, g.sd,
                                              g.sh, g.sw, bias, act, slope);
    return launched();
  }
  col2im1_kernel<<<(unsigned)cdiv(total, 256), 256, 0, st>>>(g, static_cast<const bf16*>(ws),
                                                             static_cast<bf16*>(out_big), Kp, bias, act, slope, total);
  return launched();
}

int thin_wgrad(const Geo& g, const void* small, const void* big, float* dw, void* ws, size_t wsb, cudaStream_t st,
               bool flip) {
  if (!ws || wsb < thin_workspace(g, 2)) return B2C_ERR_WORKSPACE;
  if (is_pointwise(g)) {
    const int64_t rows = rows_of(g);
    const int nb = pw_wgrad_blocks(rows);
    if ((size_t)nb * g.sc * sizeof(float) > wsb) return B2C_ERR_WORKSPACE;
    float* part = static_cast<float*>(ws);
    if (rows > 0)
      pw_wgrad_stage1<<<nb, 256, 256 * 8 * sizeof(float), st>>>(static_cast<const bf16*>(small),
                                                               static_cast<const bf16*>(big), part, g.sc, rows,
                                                               cdiv(rows, (int64_t)nb));
    int rc = launched();
    if (rc) return rc;
    pw_wgrad_stage2<<<(unsigned)cdiv(g.sc, 256), 256, 0, st>>>(part, rows > 0 ? nb : 0, g.sc, dw);
    return launched();
  }
  const int Kp = thin_kp(g), T = g.taps();
  if (tc_thin_gather_supported(g) && !knobs().thin_unfused) {
    const int frc = tc_thin_wgrad_launch(g, small, big, dw, flip ? 1 : 0, ws, wsb, st);
    if (frc != B2C_ERR_UNSUPPORTED) return frc;
  }
  int rc = im2col_launch(g, big, ws, st);
  if (rc) return rc;
  uint8_t* p = static_cast<uint8_t*>(ws) + col_bytes(g);
  float* tmp = reinterpret_cast<float*>(p);
  p += align256((size_t)g.sc * Kp * sizeof(float));
  const Geo gg = gemm_geo(rows_of(g), g.sc, Kp);
  const size_t tcws = tc_wgrad_workspace(gg);
  float* target = (Kp == T && !flip) ? dw : tmp;
  rc = tc_wgrad_launch(gg, small, ws, target, p, tcws, st);
  if (rc) return rc;
  if (target == tmp) {
    const int64_t n = (int64_t)g.sc * T;
    thin_unpad_kernel<<<(unsigned)cdiv(n, 256), 256, 0, st>>>(tmp, dw, g.sc, T, Kp, flip ? 1 : 0);
    return launched();
  }
  return B2C_OK;
}

}  // namespace b2c
