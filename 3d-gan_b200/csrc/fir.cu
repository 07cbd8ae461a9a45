// fir.cu — depthwise FIR resampling on NDHWC tensors (stylegan.py:172-200 `_upfirdn3d_ref`):
// zero-insert upsampling, pad / crop, one [kd][kh][kw] filter shared by all channels
// (F.conv3d(groups=C) over a repeated filter in the reference), decimation.  HBM bound.
//
//   y[n, o, c] = sum_k f[k] * xup[n, o*down + k - pad, c],   xup[j] = x[j / up] if up | j, else 0
//
// `f` arrives as CORRELATION taps (the caller has applied the reference's flip and gain).
// Padding is never materialised (bounds checks; negative pad = crop), decimation and zero
// insertion are index arithmetic: one read of x, one write of y.  The backward pass w.r.t. x
// is the adjoint gather (the filter is a constant buffer in the reference, it has no gradient).
#include "common.cuh"
#include "ptx.cuh"

namespace b2c {
namespace {

constexpr int kThr = 128;
constexpr int kMaxTaps = 64;

struct FirP {
  int64_t n;
  int id, ih, iw, od, oh, ow, c;
  int kd, kh, kw, up, down;
  int pad[3];
  float f[kMaxTaps];
};

bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

template <typename T, int VEC>
__device__ __forceinline__ void fir_ld(const T* p, float* o) {
  if (VEC > 1) { Vec<T> v; v.load(p); v.unpack(o); } else o[0] = ldf(p, 0);
}
template <typename T, int VEC>
__device__ __forceinline__ void fir_st(T* p, float* o) {
  if (VEC > 1) { Vec<T> v; v.pack(o); v.store(p); } else stf(p, 0, o[0]);
}

// smallest k >= 0 with k == v (mod m)
__device__ __forceinline__ int first_tap(int v, int m) {
  const int r = v % m;
  return r < 0 ? r + m : r;
}

// generic forward: one thread per (output voxel, channel vector)
// UP / DOWN: compile-time rates (1 or 2) so the index arithmetic is shifts; 0 = take them from P.
template <typename T, int VEC, int UP, int DOWN>
__global__ void __launch_bounds__(kThr) fir_fwd_kernel(const __grid_constant__ FirP Pin, const T* __restrict__ x,
                                                        T* __restrict__ y, int64_t total) {
  const FirP& P0 = Pin;
  struct { int64_t n; int id, ih, iw, od, oh, ow, c, kd, kh, kw, up, down; int pad[3]; const float* f; } P =
      {P0.n, P0.id, P0.ih, P0.iw, P0.od, P0.oh, P0.ow, P0.c, P0.kd, P0.kh, P0.kw, UP ? UP : P0.up, DOWN ? DOWN : P0.down,
       {P0.pad[0], P0.pad[1], P0.pad[2]}, P0.f};
  const int64_t idx = (int64_t)blockIdx.x * kThr + threadIdx.x;
  if (idx >= total) return;
  const int cv = P.c / VEC;
  const int c = (int)(idx % cv) * VEC;
  int64_t v = idx / cv;
  T* dst = y + v * (int64_t)P.c + c;
  const int ow = (int)(v % P.ow); v /= P.ow;
  const int oh = (int)(v % P.oh); v /= P.oh;
  const int od = (int)(v % P.od);
  const int64_t n = v / P.od;
  float acc[VEC];
#pragma unroll
  for (int i = 0; i < VEC; i++) acc[i] = 0.f;
  // only taps with (o*down + k - pad) % up == 0 touch a real sample: start at that residue, step by up
  const int a0 = first_tap(P.pad[0] - od * P.down, P.up), b0 = first_tap(P.pad[1] - oh * P.down, P.up),
            e0 = first_tap(P.pad[2] - ow * P.down, P.up);
  for (int a = a0; a < P.kd; a += P.up) {
    const int jd = od * P.down + a - P.pad[0];
    if (jd < 0 || jd >= P.id * P.up) continue;
    for (int b = b0; b < P.kh; b += P.up) {
      const int jh = oh * P.down + b - P.pad[1];
      if (jh < 0 || jh >= P.ih * P.up) continue;
      const T* row = x + ((n * P.id + jd / P.up) * P.ih + jh / P.up) * (int64_t)P.iw * P.c + c;
      for (int e = e0; e < P.kw; e += P.up) {
        const int jw = ow * P.down + e - P.pad[2];
        if (jw < 0 || jw >= P.iw * P.up) continue;
        float t[VEC];
        fir_ld<T, VEC>(row + (int64_t)(jw / P.up) * P.c, t);
        const float fk = P.f[(a * P.kh + b) * P.kw + e];
#pragma unroll
        for (int i = 0; i < VEC; i++) acc[i] += fk * t[i];
      }
    }
  }
  fir_st<T, VEC>(dst, acc);
}

// generic adjoint: dx[n, i, c] = sum over (o, k) with o*down + k - pad == i*up of f[k] * dy[n, o, c]
template <typename T, int VEC, int UP, int DOWN>
__global__ void __launch_bounds__(kThr) fir_bwd_kernel(const __grid_constant__ FirP Pin, const T* __restrict__ dy,
                                                        T* __restrict__ dx, int64_t total) {
  const FirP& P0 = Pin;
  struct { int64_t n; int id, ih, iw, od, oh, ow, c, kd, kh, kw, up, down; int pad[3]; const float* f; } P =
      {P0.n, P0.id, P0.ih, P0.iw, P0.od, P0.oh, P0.ow, P0.c, P0.kd, P0.kh, P0.kw, UP ? UP : P0.up, DOWN ? DOWN : P0.down,
       {P0.pad[0], P0.pad[1], P0.pad[2]}, P0.f};
  const int64_t idx = (int64_t)blockIdx.x * kThr + threadIdx.x;
  if (idx >= total) return;
  const int cv = P.c / VEC;
  const int c = (int)(idx % cv) * VEC;
  int64_t v = idx / cv;
  T* dst = dx + v * (int64_t)P.c + c;
  const int iw = (int)(v % P.iw); v /= P.iw;
  const int ih = (int)(v % P.ih); v /= P.ih;
  const int id = (int)(v % P.id);
  const int64_t n = v / P.id;
  float acc[VEC];
#pragma unroll
  for (int i = 0; i < VEC; i++) acc[i] = 0.f;
  // only taps with (i*up + pad - k) % down == 0 reach an output: start at that residue, step by down
  const int a0 = first_tap(id * P.up + P.pad[0], P.down), b0 = first_tap(ih * P.up + P.pad[1], P.down),
            e0 = first_tap(iw * P.up + P.pad[2], P.down);
  for (int a = a0; a < P.kd; a += P.down) {
    const int td = id * P.up + P.pad[0] - a;
    if (td < 0 || td / P.down >= P.od) continue;
    for (int b = b0; b < P.kh; b += P.down) {
      const int th = ih * P.up + P.pad[1] - b;
      if (th < 0 || th / P.down >= P.oh) continue;
      const T* row = dy + ((n * P.od + td / P.down) * P.oh + th / P.down) * (int64_t)P.ow * P.c + c;
      for (int e = e0; e < P.kw; e += P.down) {
        const int tw = iw * P.up + P.pad[2] - e;
        if (tw < 0 || tw / P.down >= P.ow) continue;
        float t[VEC];
        fir_ld<T, VEC>(row + (int64_t)(tw / P.down) * P.c, t);
        const float fk = P.f[(a * P.kh + b) * P.kw + e];
#pragma unroll
        for (int i = 0; i < VEC; i++) acc[i] += fk * t[i];
      }
    }
  }
  fir_st<T, VEC>(dst, acc);
}

// up = down = 1, kw <= 4: a thread owns WB = 4 adjacent outputs along w of one channel vector and
// loads kw + 3 input vectors per (kd, kh) line instead of 4 * kw.
template <typename T, int VEC>
__global__ void __launch_bounds__(kThr) fir_fwd_w4_kernel(const __grid_constant__ FirP P, const T* __restrict__ x,
                                                           T* __restrict__ y, int64_t total) {
  constexpr int WB = 4, KW = 4;
  const int64_t idx = (int64_t)blockIdx.x * kThr + threadIdx.x;
  if (idx >= total) return;
  const int cv = P.c / VEC;
  const int c = (int)(idx % cv) * VEC;
  int64_t v = idx / cv;
  const int wblocks = (P.ow + WB - 1) / WB;
  const int ow0 = (int)(v % wblocks) * WB; v /= wblocks;
  const int oh = (int)(v % P.oh); v /= P.oh;
  const int od = (int)(v % P.od);
  const int64_t n = v / P.od;
  float acc[WB][VEC];
#pragma unroll
  for (int q = 0; q < WB; q++)
#pragma unroll
    for (int i = 0; i < VEC; i++) acc[q][i] = 0.f;
  for (int a = 0; a < P.kd; a++) {
    const int jd = od + a - P.pad[0];
    if (jd < 0 || jd >= P.id) continue;
    for (int b = 0; b < P.kh; b++) {
      const int jh = oh + b - P.pad[1];
      if (jh < 0 || jh >= P.ih) continue;
      const T* row = x + ((n * P.id + jd) * P.ih + jh) * (int64_t)P.iw * P.c + c;
      const float* fk = P.f + (a * P.kh + b) * P.kw;
      float in[WB + KW - 1][VEC];
#pragma unroll
      for (int j = 0; j < WB + KW - 1; j++) {
        const int jw = ow0 + j - P.pad[2];
        if (j < WB + P.kw - 1 && jw >= 0 && jw < P.iw) fir_ld<T, VEC>(row + (int64_t)jw * P.c, in[j]);
        else {
#pragma unroll
          for (int i = 0; i < VEC; i++) in[j][i] = 0.f;
        }
      }
#pragma unroll
      for (int e = 0; e < KW; e++) {
        if (e < P.kw) {
          const float w = fk[e];
#pragma unroll
          for (int q = 0; q < WB; q++)
#pragma unroll
            for (int i = 0; i < VEC; i++) acc[q][i] += w * in[q + e][i];
        }
      }
    }
  }
  T* dst = y + (((n * P.od + od) * P.oh + oh) * (int64_t)P.ow + ow0) * P.c + c;
#pragma unroll
  for (int q = 0; q < WB; q++)
    if (ow0 + q < P.ow) fir_st<T, VEC>(dst + (int64_t)q * P.c, acc[q]);
}


// up = down = 1, 4x4x4 rank-1 filter f[a][b][e] = fd[a] * fh[b] * fw[e] (the reference's [1,3,3,1]-derived
// filter is one: constant along d, stylegan.py:462-465).  A thread owns WB = 2 adjacent outputs along w of
// one channel vector and marches along d over DCH outputs: every input plane is filtered along w and h in
// registers once (4 lines x 5 loads, 4 + 4 taps) and scattered into a rolling window of 4 output
// accumulators along d: 12 multiply-adds per output instead of 64, 10 loads per output instead of 28.
struct FirSep { float fd[4], fh[4], fw[4]; int dch; };

// DOWN = 2 (decimating FIR of the discriminator's skip path, stylegan.py:244-247): the same march with input
// rows / columns / planes at stride 2; a plane feeds K / DOWN = 2 outputs along d.
template <typename T, int VEC, int DOWN>
__global__ void __launch_bounds__(kThr) fir_sep_kernel(const __grid_constant__ FirP P, const __grid_constant__ FirSep S,
                                                        const T* __restrict__ x, T* __restrict__ y, int64_t total) {
  constexpr int WB = 2, K = 4, NACC = K / DOWN, NCOL = DOWN * (WB - 1) + K;
  const int64_t idx = (int64_t)blockIdx.x * kThr + threadIdx.x;
  if (idx >= total) return;
  const int cv = P.c / VEC;
  const int c = (int)(idx % cv) * VEC;
  int64_t v = idx / cv;
  const int wblocks = (P.ow + WB - 1) / WB;
  const int ow0 = (int)(v % wblocks) * WB; v /= wblocks;
  const int oh = (int)(v % P.oh); v /= P.oh;
  const int dchunks = (P.od + S.dch - 1) / S.dch;
  const int od0 = (int)(v % dchunks) * S.dch;
  const int64_t n = v / dchunks;
  const int nout = min(S.dch, P.od - od0);
  float acc[NACC][WB][VEC];
#pragma unroll
  for (int s = 0; s < NACC; s++)
#pragma unroll
    for (int q = 0; q < WB; q++)
#pragma unroll
      for (int i = 0; i < VEC; i++) acc[s][q][i] = 0.f;
  T* dst = y + (((n * P.od + od0) * P.oh + oh) * (int64_t)P.ow + ow0) * P.c + c;
  const int64_t dst_plane = (int64_t)P.oh * P.ow * P.c;
  const int nplanes = (nout - 1) * DOWN + K;
  for (int base = 0; base < nplanes; base += K) {
#pragma unroll
    for (int s = 0; s < K; s++) {
      const int r = base + s;                       // input plane index relative to the chunk
      const int jd = od0 * DOWN + r - P.pad[0];
      if (r < nplanes && jd >= 0 && jd < P.id) {
        // all 4 x 5 row loads of the plane are issued before the first use (predicated, no branches):
        // the kernel runs at 2-3 CTAs per SM, so memory-level parallelism has to come from each thread
        Vec<T> raw[K][NCOL];
        const T* plane = x + (n * P.id + jd) * (int64_t)P.ih * P.iw * P.c + c;
#pragma unroll
        for (int b = 0; b < K; b++) {
          const int jh = oh * DOWN + b - P.pad[1];
          const bool hok = jh >= 0 && jh < P.ih;
#pragma unroll
          for (int j = 0; j < NCOL; j++) {
            const int jw = ow0 * DOWN + j - P.pad[2];
            const bool ok = hok && jw >= 0 && jw < P.iw;
            raw[b][j].load(plane + (ok ? ((int64_t)jh * P.iw + jw) * P.c : 0));     // plane[0] is always valid
            if (!ok) raw[b][j].zero();
          }
        }
        float p[WB][VEC];
#pragma unroll
        for (int q = 0; q < WB; q++)
#pragma unroll
          for (int i = 0; i < VEC; i++) p[q][i] = 0.f;
#pragma unroll
        for (int b = 0; b < K; b++) {
          float in[NCOL][VEC];
#pragma unroll
          for (int j = 0; j < NCOL; j++) raw[b][j].unpack(in[j]);
          const float hb = S.fh[b];
#pragma unroll
          for (int q = 0; q < WB; q++)
#pragma unroll
            for (int i = 0; i < VEC; i++) {
              float t = S.fw[0] * in[q * DOWN][i];
              t = fmaf(S.fw[1], in[q * DOWN + 1][i], t);
              t = fmaf(S.fw[2], in[q * DOWN + 2][i], t);
              t = fmaf(S.fw[3], in[q * DOWN + 3][i], t);
              p[q][i] = fmaf(hb, t, p[q][i]);
            }
        }
        // plane r feeds outputs o = (r - a) / DOWN for the taps a = r % DOWN + DOWN * j; output o lives in slot o % NACC
#pragma unroll
        for (int j = 0; j < NACC; j++) {
          const float da = S.fd[(s % DOWN) + DOWN * j];
#pragma unroll
          for (int q = 0; q < WB; q++)
#pragma unroll
            for (int i = 0; i < VEC; i++)
              acc[(s / DOWN - j + NACC) % NACC][q][i] = fmaf(da, p[q][i], acc[(s / DOWN - j + NACC) % NACC][q][i]);
        }
      }
      // output o = (r + 1 - K) / DOWN is complete after plane r when DOWN | r + 1: its slot is ((s + 1) / DOWN) % NACC
      if ((s + 1) % DOWN == 0) {
        const int o = (r + 1 - K) / DOWN;
        if (r + 1 >= K && o < nout) {
#pragma unroll
          for (int q = 0; q < WB; q++)
            if (ow0 + q < P.ow) fir_st<T, VEC>(dst + o * dst_plane + (int64_t)q * P.c, acc[((s + 1) / DOWN) % NACC][q]);
        }
#pragma unroll
        for (int q = 0; q < WB; q++)
#pragma unroll
          for (int i = 0; i < VEC; i++) acc[((s + 1) / DOWN) % NACC][q][i] = 0.f;
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------------------------
// Unit-rate 4x4x4 rank-1 FIR, shared-memory tiled (the 128^3 / 64^3 / 32^3 sites of stylegan.py:427-432 and their
// adjoints: 33 of the 47 FIR milliseconds of a BASELINE config 4 step ran in fir_sep_kernel<.., 1>, whose threads
// each pull a 4 x 5 window per plane through L1: 10 loads per output, 20 % of the copy peak).
//
// A CTA of 256 threads owns a 16 x 16 (h, w) tile of CG = 4 channel vectors (32 bf16 channels) and marches along d
// over `dch` outputs.  Every input plane of the tile (19 x 19 voxels x 64 B = 23 KB) is copied ONCE into shared memory
// with cp.async (three buffers: two planes in flight while one is filtered; out-of-range voxels are stored as zeros,
// padding is never materialised), then a thread = (4 outputs along h, one w, one channel vector) reads 7 rows x 4
// columns as conflict-free LDS.128 (a warp reads 512 contiguous bytes), filters along w, then h, in registers
// (15 multiply-adds per output element instead of 64) and scatters the plane into the rolling window of 4 output
// accumulators along d.  HBM sees each input byte 1.4 times (tile halo) and each output byte once.
constexpr int kFtT = 16, kFtR = kFtT + 3, kFtCG = 4, kFtThreads = 256, kFtBufs = 3;
constexpr int kFtVecs = kFtR * kFtR * kFtCG;                       // 16-byte vectors per plane buffer

__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

template <typename T>
__global__ void __launch_bounds__(kFtThreads, 1) fir_tile_kernel(const __grid_constant__ FirP P, const __grid_constant__ FirSep S,
                                                                  const T* __restrict__ x, T* __restrict__ y,
                                                                  int htiles, int wtiles, int cgroups, int dchunks) {
  constexpr int VEC = Vec<T>::N, K = 4;
  extern __shared__ uint4 ft_smem[];
  int b = blockIdx.x;
  const int cgi = b % cgroups; b /= cgroups;
  const int wt = b % wtiles; b /= wtiles;
  const int ht = b % htiles; b /= htiles;
  const int dc = b % dchunks;
  const int64_t n = b / dchunks;
  const int tid = threadIdx.x;
  const int cv = tid % kFtCG, wl = (tid / kFtCG) % kFtT, hb = tid / (kFtCG * kFtT);
  const int oh0 = ht * kFtT, ow0 = wt * kFtT, od0 = dc * S.dch, c0 = cgi * kFtCG * VEC;
  const int nout = min(S.dch, P.od - od0);
  const int nplanes = nout - 1 + K;
  const uint32_t sbase = ptx::smem_u32(ft_smem);

  auto load_plane = [&](int r) {
    if (r < nplanes) {
      const int jd = od0 + r - P.pad[0];
      const bool dok = jd >= 0 && jd < P.id;
      const uint32_t dst0 = sbase + (uint32_t)(r % kFtBufs) * (kFtVecs * 16u);
      const T* plane = x + (n * P.id + (dok ? jd : 0)) * (int64_t)P.ih * P.iw * P.c + c0;
      for (int i = tid; i < kFtVecs; i += kFtThreads) {
        const int cvv = i % kFtCG, col = (i / kFtCG) % kFtR, row = i / (kFtCG * kFtR);
        const int jh = oh0 + row - P.pad[1], jw = ow0 + col - P.pad[2];
        if (dok && jh >= 0 && jh < P.ih && jw >= 0 && jw < P.iw)
          cp_async16(dst0 + (uint32_t)i * 16u, plane + ((int64_t)jh * P.iw + jw) * P.c + cvv * VEC);
        else
          ft_smem[(r % kFtBufs) * kFtVecs + i] = make_uint4(0u, 0u, 0u, 0u);
      }
    }
    cp_async_commit();                                  // one group per call, empty or not: wait counts stay uniform
  };

  float acc[K][4][VEC];
#pragma unroll
  for (int s = 0; s < K; s++)
#pragma unroll
    for (int q = 0; q < 4; q++)
#pragma unroll
      for (int i = 0; i < VEC; i++) acc[s][q][i] = 0.f;
  const int oh = oh0 + hb * 4, ow = ow0 + wl;
  T* dst = y + (((n * P.od + od0) * P.oh + oh) * (int64_t)P.ow + ow) * P.c + c0 + cv * VEC;
  const int64_t dst_plane = (int64_t)P.oh * P.ow * P.c, dst_row = (int64_t)P.ow * P.c;
  const bool wok = ow < P.ow;

  load_plane(0);
  load_plane(1);
  for (int base = 0; base < nplanes; base += K) {
#pragma unroll
    for (int s = 0; s < K; s++) {
      const int r = base + s;
      load_plane(r + 2);                                // buffer (r + 2) % 3 was released by the barrier that ended plane r - 1
      cp_async_wait<2>();                               // this thread's copies of plane r have landed
      __syncthreads();                                  // ... and everybody else's (and the zero stores)
      if (r < nplanes) {
        const uint4* pl = ft_smem + (r % kFtBufs) * kFtVecs;
        float p[4][VEC];
#pragma unroll
        for (int q = 0; q < 4; q++)
#pragma unroll
          for (int i = 0; i < VEC; i++) p[q][i] = 0.f;
#pragma unroll
        for (int bb = 0; bb < 4 + K - 1; bb++) {
          const uint4* rowp = pl + ((hb * 4 + bb) * kFtR + wl) * kFtCG + cv;
          float t[VEC];
#pragma unroll
          for (int e = 0; e < K; e++) {
            Vec<T> raw;
            *reinterpret_cast<uint4*>(&raw.raw) = rowp[e * kFtCG];
            float in[VEC];
            raw.unpack(in);
#pragma unroll
            for (int i = 0; i < VEC; i++) t[i] = e == 0 ? S.fw[0] * in[i] : fmaf(S.fw[e], in[i], t[i]);
          }
#pragma unroll
          for (int q = 0; q < 4; q++)
            if (bb - q >= 0 && bb - q < K) {
#pragma unroll
              for (int i = 0; i < VEC; i++) p[q][i] = fmaf(S.fh[bb - q], t[i], p[q][i]);
            }
        }
        // plane r feeds output o = r - j through tap j; output o lives in slot o % K
#pragma unroll
        for (int j = 0; j < K; j++)
#pragma unroll
          for (int q = 0; q < 4; q++)
#pragma unroll
            for (int i = 0; i < VEC; i++)
              acc[(s - j + K) % K][q][i] = fmaf(S.fd[j], p[q][i], acc[(s - j + K) % K][q][i]);
      }
      // output o = r + 1 - K is complete after plane r: slot (s + 1) % K
      {
        const int o = r + 1 - K;
        if (o >= 0 && o < nout && wok) {
#pragma unroll
          for (int q = 0; q < 4; q++)
            if (oh + q < P.oh) fir_st<T, VEC>(dst + o * dst_plane + q * dst_row, acc[(s + 1) % K][q]);
        }
#pragma unroll
        for (int q = 0; q < 4; q++)
#pragma unroll
          for (int i = 0; i < VEC; i++) acc[(s + 1) % K][q][i] = 0.f;
      }
      __syncthreads();                                  // plane r's buffer may be overwritten (by load_plane(r + 3))
    }
  }
  cp_async_wait<0>();
}

// ---------------------------------------------------------------------------------------------------------------
// Adjoint of the decimating (down = 2) 4x4x4 rank-1 FIR: the gradient of the discriminator's skip path
// (stylegan.py:244-247).  dx[i] = sum over (o, k) with 2 o + k - pad == i of f[k] dy[o]: per dimension a fine sample
// sees the two taps of its parity.  A thread owns a 2 x 2 x 2 block of fine outputs of one channel vector: the 8
// outputs read the same 3 x 3 x 3 coarse neighbourhood (27 loads instead of 8 x 8), filtered separably in registers
// with the per-dimension coefficient table cf[dim][fine parity][coarse offset] built on the host (one of the three
// entries is zero).  Writes (8/9 of the bytes) are 16-byte vectors, two adjacent voxels per thread.
struct FirUp2 { float cf[3][2][3]; int ob[3]; };       // ob: coarse index of offset 0 for fine index 0, per dim

template <typename T, int VEC>
__global__ void __launch_bounds__(kThr) fir_adj2_kernel(const __grid_constant__ FirP P, const __grid_constant__ FirUp2 U,
                                                         const T* __restrict__ dy, T* __restrict__ dx, int64_t total) {
  const int64_t idx = (int64_t)blockIdx.x * kThr + threadIdx.x;
  if (idx >= total) return;
  const int cv = P.c / VEC;
  const int c = (int)(idx % cv) * VEC;
  int64_t v = idx / cv;
  const int mw_n = (P.iw + 1) / 2, mh_n = (P.ih + 1) / 2, md_n = (P.id + 1) / 2;
  const int mw = (int)(v % mw_n); v /= mw_n;
  const int mh = (int)(v % mh_n); v /= mh_n;
  const int md = (int)(v % md_n);
  const int64_t n = v / md_n;
  float out[2][2][2][VEC];
#pragma unroll
  for (int a = 0; a < 2; a++)
#pragma unroll
    for (int b = 0; b < 2; b++)
#pragma unroll
      for (int e = 0; e < 2; e++)
#pragma unroll
        for (int i = 0; i < VEC; i++) out[a][b][e][i] = 0.f;
  const int od0 = md + U.ob[0], oh0 = mh + U.ob[1], ow0 = mw + U.ob[2];
#pragma unroll
  for (int jd = 0; jd < 3; jd++) {
    const int od = od0 + jd - 1;
    if (od < 0 || od >= P.od) continue;
    float ph[2][2][VEC];
#pragma unroll
    for (int b = 0; b < 2; b++)
#pragma unroll
      for (int e = 0; e < 2; e++)
#pragma unroll
        for (int i = 0; i < VEC; i++) ph[b][e][i] = 0.f;
#pragma unroll
    for (int jh = 0; jh < 3; jh++) {
      const int oh = oh0 + jh - 1;
      if (oh < 0 || oh >= P.oh) continue;
      const T* row = dy + ((n * P.od + od) * P.oh + oh) * (int64_t)P.ow * P.c + c;
      float tw[2][VEC];
#pragma unroll
      for (int e = 0; e < 2; e++)
#pragma unroll
        for (int i = 0; i < VEC; i++) tw[e][i] = 0.f;
#pragma unroll
      for (int jw = 0; jw < 3; jw++) {
        const int ow = ow0 + jw - 1;
        if (ow < 0 || ow >= P.ow) continue;
        float in[VEC];
        fir_ld<T, VEC>(row + (int64_t)ow * P.c, in);
#pragma unroll
        for (int e = 0; e < 2; e++)
#pragma unroll
          for (int i = 0; i < VEC; i++) tw[e][i] = fmaf(U.cf[2][e][jw], in[i], tw[e][i]);
      }
#pragma unroll
      for (int b = 0; b < 2; b++)
#pragma unroll
        for (int e = 0; e < 2; e++)
#pragma unroll
          for (int i = 0; i < VEC; i++) ph[b][e][i] = fmaf(U.cf[1][b][jh], tw[e][i], ph[b][e][i]);
    }
#pragma unroll
    for (int a = 0; a < 2; a++)
#pragma unroll
      for (int b = 0; b < 2; b++)
#pragma unroll
        for (int e = 0; e < 2; e++)
#pragma unroll
          for (int i = 0; i < VEC; i++) out[a][b][e][i] = fmaf(U.cf[0][a][jd], ph[b][e][i], out[a][b][e][i]);
  }
#pragma unroll
  for (int a = 0; a < 2; a++) {
    const int id = 2 * md + a;
    if (id >= P.id) continue;
#pragma unroll
    for (int b = 0; b < 2; b++) {
      const int ih = 2 * mh + b;
      if (ih >= P.ih) continue;
#pragma unroll
      for (int e = 0; e < 2; e++) {
        const int iw = 2 * mw + e;
        if (iw < P.iw) fir_st<T, VEC>(dx + (((n * P.id + id) * P.ih + ih) * (int64_t)P.iw + iw) * P.c + c, out[a][b][e]);
      }
    }
  }
}

// coefficient table of fir_adj2_kernel for one dimension: fine index i = 2 m + a, q = i + pad, tap parity par = q & 1,
// taps k = par (coarse (q - par) / 2) and k = par + 2 (one coarse sample lower)
static void adj2_dim(const float* f, int pad, float cf[2][3], int* ob) {
  auto fdiv2 = [](int v) { return v >= 0 ? v / 2 : -((-v + 1) / 2); };
  const int par0 = ((pad % 2) + 2) % 2;
  const int base = fdiv2(pad - par0);               // coarse index of the high tap for i = 0; offset 0 = base
  *ob = base;
  for (int a = 0; a < 2; a++) {
    for (int j = 0; j < 3; j++) cf[a][j] = 0.f;
    const int q = a + pad, par = ((q % 2) + 2) % 2, hi = fdiv2(q - par);      // relative to m = 0
    const int off = hi - base;                      // 0 or 1
    cf[a][off + 1] += f[par];                       // table index j = offset + 1
    cf[a][off] += f[par + 2];
  }
}

// f[a][b][e] == fd[a] * fh[b] * fw[e]?  (host; 64 taps)
bool factor_rank1(const float* f, FirSep* S) {
  int pa = 0, pb = 0, pe = 0;
  float best = 0.f;
  for (int t = 0; t < 64; t++)
    if (fabsf(f[t]) > best) { best = fabsf(f[t]); pa = t / 16; pb = (t / 4) % 4; pe = t % 4; }
  if (best == 0.f) return false;
  const float piv = f[(pa * 4 + pb) * 4 + pe];
  for (int k = 0; k < 4; k++) {
    S->fd[k] = f[(k * 4 + pb) * 4 + pe] / piv;
    S->fh[k] = f[(pa * 4 + k) * 4 + pe] / piv;
    S->fw[k] = f[(pa * 4 + pb) * 4 + k];
  }
  for (int a = 0; a < 4; a++)
    for (int b = 0; b < 4; b++)
      for (int e = 0; e < 4; e++)
        if (fabsf(S->fd[a] * S->fh[b] * S->fw[e] - f[(a * 4 + b) * 4 + e]) > 1e-6f * best) return false;
  return true;
}

template <typename T>
int fir_run(bool bwd, const FirP& P, const void* a, void* b, cudaStream_t st) {
  constexpr int V = Vec<T>::N;
  const bool vec = P.c % V == 0 && aligned16(a) && aligned16(b);
  const T* ap = static_cast<const T*>(a); T* bp = static_cast<T*>(b);
  const int cvn = P.c / (vec ? V : 1);
  if (P.up == 1 && P.down == 2 && !bwd && vec && P.kd == 4 && P.kh == 4 && P.kw == 4) {
    FirSep S;
    if (factor_rank1(P.f, &S)) {
      S.dch = 16;
      const int64_t tot = P.n * ((P.od + S.dch - 1) / S.dch) * P.oh * ((P.ow + 1) / 2) * cvn;
      if (tot == 0) return B2C_OK;
      const int64_t blk = cdiv(tot, (int64_t)kThr);
      if (blk > 0x7fffffffLL) return B2C_ERR_UNSUPPORTED;
      fir_sep_kernel<T, V, 2><<<(unsigned)blk, kThr, 0, st>>>(P, S, ap, bp, tot);
      return launched();
    }
  }
  if (P.up == 1 && P.down == 1 && P.kw <= 4 && vec) {
    // forward as is; the adjoint of a unit-rate FIR is the FIR with the filter reversed and pad' = K-1-pad
    FirP Q = P;
    if (bwd) {
      Q.id = P.od; Q.ih = P.oh; Q.iw = P.ow; Q.od = P.id; Q.oh = P.ih; Q.ow = P.iw;
      Q.pad[0] = P.kd - 1 - P.pad[0]; Q.pad[1] = P.kh - 1 - P.pad[1]; Q.pad[2] = P.kw - 1 - P.pad[2];
      const int T_ = P.kd * P.kh * P.kw;
      for (int t = 0; t < T_; t++) Q.f[t] = P.f[T_ - 1 - t];
    }
    FirSep S;
    if (Q.kd == 4 && Q.kh == 4 && Q.kw == 4 && factor_rank1(Q.f, &S)) {
      if (Q.c % (kFtCG * V) == 0 && Q.oh >= kFtT && Q.ow >= kFtT && Q.od >= 8) {
        // shared-memory tiled march (see fir_tile_kernel); small volumes keep the register-window kernel below
        S.dch = Q.od >= 64 ? 32 : Q.od;
        const int ht = (Q.oh + kFtT - 1) / kFtT, wt = (Q.ow + kFtT - 1) / kFtT, cg = Q.c / (kFtCG * V),
                  dc = (Q.od + S.dch - 1) / S.dch;
        const int64_t blk = Q.n * dc * ht * wt * cg;
        if (blk == 0) return B2C_OK;
        if (blk > 0x7fffffffLL) return B2C_ERR_UNSUPPORTED;
        constexpr int smem = kFtBufs * kFtVecs * 16;
        if (cudaFuncSetAttribute(fir_tile_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess)
          return B2C_ERR_CUDA;
        fir_tile_kernel<T><<<(unsigned)blk, kFtThreads, smem, st>>>(Q, S, ap, bp, ht, wt, cg, dc);
        return launched();
      }
      S.dch = Q.od >= 64 ? 32 : 16;
      const int64_t tot = Q.n * ((Q.od + S.dch - 1) / S.dch) * Q.oh * ((Q.ow + 1) / 2) * cvn;
      if (tot == 0) return B2C_OK;
      const int64_t blk = cdiv(tot, (int64_t)kThr);
      if (blk > 0x7fffffffLL) return B2C_ERR_UNSUPPORTED;
      fir_sep_kernel<T, V, 1><<<(unsigned)blk, kThr, 0, st>>>(Q, S, ap, bp, tot);
      return launched();
    }
    const int64_t total = Q.n * Q.od * Q.oh * ((Q.ow + 3) / 4) * cvn;
    if (total == 0) return B2C_OK;
    const int64_t blocks = cdiv(total, (int64_t)kThr);
    if (blocks > 0x7fffffffLL) return B2C_ERR_UNSUPPORTED;
    fir_fwd_w4_kernel<T, V><<<(unsigned)blocks, kThr, 0, st>>>(Q, ap, bp, total);
    return launched();
  }
  if (bwd && P.up == 1 && P.down == 2 && vec && P.kd == 4 && P.kh == 4 && P.kw == 4) {
    FirSep S;
    if (factor_rank1(P.f, &S)) {
      FirUp2 U;
      adj2_dim(S.fd, P.pad[0], U.cf[0], &U.ob[0]);
      adj2_dim(S.fh, P.pad[1], U.cf[1], &U.ob[1]);
      adj2_dim(S.fw, P.pad[2], U.cf[2], &U.ob[2]);
      const int64_t tot = P.n * ((P.id + 1) / 2) * ((P.ih + 1) / 2) * ((P.iw + 1) / 2) * cvn;
      if (tot == 0) return B2C_OK;
      const int64_t blk = cdiv(tot, (int64_t)kThr);
      if (blk > 0x7fffffffLL) return B2C_ERR_UNSUPPORTED;
      fir_adj2_kernel<T, V><<<(unsigned)blk, kThr, 0, st>>>(P, U, ap, bp, tot);
      return launched();
    }
  }
  const int64_t total = (bwd ? P.n * P.id * P.ih * P.iw : P.n * P.od * P.oh * P.ow) * cvn;
  if (total == 0) return B2C_OK;
  const int64_t blocks = cdiv(total, (int64_t)kThr);
  if (blocks > 0x7fffffffLL) return B2C_ERR_UNSUPPORTED;
  const unsigned g = (unsigned)blocks;
  const int rate = P.up == 1 && P.down == 2 ? 12 : (P.up == 2 && P.down == 1 ? 21 : 0);
#define B2C_FIR_LAUNCH(KERN, U, D)                                                 \
  do {                                                                             \
    if (vec) KERN<T, V, U, D><<<g, kThr, 0, st>>>(P, ap, bp, total);               \
    else KERN<T, 1, U, D><<<g, kThr, 0, st>>>(P, ap, bp, total);                   \
  } while (0)
  if (!bwd) {
    if (rate == 12) B2C_FIR_LAUNCH(fir_fwd_kernel, 1, 2);
    else if (rate == 21) B2C_FIR_LAUNCH(fir_fwd_kernel, 2, 1);
    else B2C_FIR_LAUNCH(fir_fwd_kernel, 0, 0);
  } else {
    if (rate == 12) B2C_FIR_LAUNCH(fir_bwd_kernel, 1, 2);
    else if (rate == 21) B2C_FIR_LAUNCH(fir_bwd_kernel, 2, 1);
    else B2C_FIR_LAUNCH(fir_bwd_kernel, 0, 0);
  }
#undef B2C_FIR_LAUNCH
  return launched();
}

int fir_entry(bool bwd, const b2c_fir_desc* d, const float* taps, const void* a, void* b, b2c_stream s) {
  if (!d || !taps || !a || !b || !dt_ok(d->dtype)) return B2C_ERR_INVALID;
  const int T_ = d->k[0] * d->k[1] * d->k[2];
  if (d->n < 0 || d->c <= 0 || d->up < 1 || d->down < 1 || T_ < 1 || T_ > kMaxTaps) return B2C_ERR_INVALID;
  for (int i = 0; i < 3; i++) if (d->in[i] <= 0 || d->out[i] <= 0 || d->k[i] < 1) return B2C_ERR_INVALID;
  FirP P;
  P.n = d->n; P.id = d->in[0]; P.ih = d->in[1]; P.iw = d->in[2];
  P.od = d->out[0]; P.oh = d->out[1]; P.ow = d->out[2]; P.c = d->c;
  P.kd = d->k[0]; P.kh = d->k[1]; P.kw = d->k[2]; P.up = d->up; P.down = d->down;
  for (int i = 0; i < 3; i++) P.pad[i] = d->pad[i];
  for (int t = 0; t < kMaxTaps; t++) P.f[t] = t < T_ ? taps[t] : 0.f;      // taps: HOST pointer
  cudaStream_t st = static_cast<cudaStream_t>(s);
  return d->dtype == B2C_BF16 ? fir_run<__nv_bfloat16>(bwd, P, a, b, st) : fir_run<float>(bwd, P, a, b, st);
}

}  // namespace
}  // namespace b2c

extern "C" int b2c_fir3d_fwd(const b2c_fir_desc* d, const float* taps, const void* x, void* y, b2c_stream s) {
  return b2c::fir_entry(false, d, taps, x, y, s);
}
extern "C" int b2c_fir3d_bwd(const b2c_fir_desc* d, const float* taps, const void* dy, void* dx, b2c_stream s) {
  return b2c::fir_entry(true, d, taps, dy, dx, s);
}
