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
