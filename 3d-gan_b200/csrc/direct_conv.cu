// direct_conv.cu — CUDA-core direct convolution kernels (see direct_conv.cuh).
#include "direct_conv.cuh"

namespace b2c {

__device__ __forceinline__ void decode_vox(int64_t v, int D, int H, int W, int& n, int& d, int& h, int& w) {
  w = (int)(v % W); v /= W;
  h = (int)(v % H); v /= H;
  d = (int)(v % D); n = (int)(v / D);
}

// ---------------------------------------------------------------------------
// gather: out(small)[v,co] = sum_{t,ci} in(big)[n, o*s+k*dl-p, ci] * W[t][co][ci]
// WARP=false: one thread per (v,co), co fastest (coalesced stores; used when
//             big.c is small, e.g. the 1-channel stems).
// WARP=true : one warp per (v,co); lanes stride ci so both operand streams are
//             coalesced; shuffle tree at the end.
template <typename T, bool WARP>
__global__ void __launch_bounds__(256)
gather_kernel(Geo g, const T* __restrict__ in, const T* __restrict__ w, T* __restrict__ out,
              const float* __restrict__ bias, int act, float slope, int64_t total) {
  const int lane = threadIdx.x & 31;
  int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (WARP) idx >>= 5;
  if (idx >= total) return;
  const int co = (int)(idx % g.sc);
  int n, od, oh, ow;
  decode_vox(idx / g.sc, g.sd, g.sh, g.sw, n, od, oh, ow);
  float acc = 0.f;
  for (int kd = 0; kd < g.k[0]; kd++) {
    const int id = od * g.s[0] + kd * g.dl[0] - g.p[0];
    if (id < 0 || id >= g.bd) continue;
    for (int kh = 0; kh < g.k[1]; kh++) {
      const int ih = oh * g.s[1] + kh * g.dl[1] - g.p[1];
      if (ih < 0 || ih >= g.bh) continue;
      for (int kw = 0; kw < g.k[2]; kw++) {
        const int iw = ow * g.s[2] + kw * g.dl[2] - g.p[2];
        if (iw < 0 || iw >= g.bw) continue;
        const int t = (kd * g.k[1] + kh) * g.k[2] + kw;
        const T* ip = in + ((((int64_t)n * g.bd + id) * g.bh + ih) * g.bw + iw) * g.bc;
        const T* wp = w + ((int64_t)t * g.sc + co) * g.bc;
        if (WARP) { for (int ci = lane; ci < g.bc; ci += 32) acc += ldf(ip, ci) * ldf(wp, ci); }
        else      { for (int ci = 0; ci < g.bc; ci++) acc += ldf(ip, ci) * ldf(wp, ci); }
      }
    }
  }
  if (WARP) { acc = warp_sum(acc); if (lane) return; }
  if (bias) acc += bias[co];
  stf(out, idx, act_f(acc, act, slope));
}

// ---------------------------------------------------------------------------
// scatter form as a gather over the taps that divide:
// out(big)[v,co] = sum_{t,ci} [s | (i+p-k*dl)] in(small)[n,(i+p-k*dl)/s,ci] * W[t][co][ci]
template <typename T, bool WARP>
__global__ void __launch_bounds__(256)
scatter_kernel(Geo g, const T* __restrict__ in, const T* __restrict__ w, T* __restrict__ out,
               const float* __restrict__ bias, int act, float slope, int64_t total) {
  const int lane = threadIdx.x & 31;
  int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (WARP) idx >>= 5;
  if (idx >= total) return;
  const int co = (int)(idx % g.bc);
  int n, id, ih, iw;
  decode_vox(idx / g.bc, g.bd, g.bh, g.bw, n, id, ih, iw);
  float acc = 0.f;
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
      for (int kw = 0; kw < g.k[2]; kw++) {
        const int nw = iw + g.p[2] - kw * g.dl[2];
        if (nw < 0 || nw % g.s[2]) continue;
        const int ow = nw / g.s[2];
        if (ow >= g.sw) continue;
        const int t = (kd * g.k[1] + kh) * g.k[2] + kw;
        const T* ip = in + ((((int64_t)n * g.sd + od) * g.sh + oh) * g.sw + ow) * g.sc;
        const T* wp = w + ((int64_t)t * g.bc + co) * g.sc;
        if (WARP) { for (int ci = lane; ci < g.sc; ci += 32) acc += ldf(ip, ci) * ldf(wp, ci); }
        else      { for (int ci = 0; ci < g.sc; ci++) acc += ldf(ip, ci) * ldf(wp, ci); }
      }
    }
  }
  if (WARP) { acc = warp_sum(acc); if (lane) return; }
  if (bias) acc += bias[co];
  stf(out, idx, act_f(acc, act, slope));
}

// BLOCK variant of the gather: one 256-thread block per output element, threads
// stride the flattened (tap, ci) reduction (ci fastest -> coalesced).  For heads
// with few outputs and a very long K (dcgan.py:181: Conv3d(2048->1, k4) on 4^3).
template <typename T>
__global__ void __launch_bounds__(256)
gather_block_kernel(Geo g, const T* __restrict__ in, const T* __restrict__ w, T* __restrict__ out,
                    const float* __restrict__ bias, int act, float slope) {
  __shared__ float red[8];
  const int64_t idx = blockIdx.x;
  const int co = (int)(idx % g.sc);
  int n, od, oh, ow;
  decode_vox(idx / g.sc, g.sd, g.sh, g.sw, n, od, oh, ow);
  const int T_ = g.taps();
  const int64_t K = (int64_t)T_ * g.bc;
  float acc = 0.f;
  for (int64_t j = threadIdx.x; j < K; j += 256) {
    const int t = (int)(j / g.bc), ci = (int)(j - (int64_t)t * g.bc);
    const int kw = t % g.k[2], kh = (t / g.k[2]) % g.k[1], kd = t / (g.k[2] * g.k[1]);
    const int id = od * g.s[0] + kd * g.dl[0] - g.p[0];
    const int ih = oh * g.s[1] + kh * g.dl[1] - g.p[1];
    const int iw = ow * g.s[2] + kw * g.dl[2] - g.p[2];
    if (id < 0 || id >= g.bd || ih < 0 || ih >= g.bh || iw < 0 || iw >= g.bw) continue;
    acc += ldf(in, ((((int64_t)n * g.bd + id) * g.bh + ih) * g.bw + iw) * g.bc + ci) *
           ldf(w, ((int64_t)t * g.sc + co) * g.bc + ci);
  }
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 8; i++) s += red[i];
    if (bias) s += bias[co];
    stf(out, idx, act_f(s, act, slope));
  }
}

template <typename T>
static int launch_fwdlike(bool scatter, const Geo& g, const void* in, const void* w, void* out, const float* bias,
                          int act, float slope, cudaStream_t st) {
  const int64_t vox = (int64_t)g.n * (scatter ? (int64_t)g.bd * g.bh * g.bw : (int64_t)g.sd * g.sh * g.sw);
  const int oc = scatter ? g.bc : g.sc, ic = scatter ? g.sc : g.bc;
  const int64_t total = vox * oc;
  if (total == 0) return B2C_OK;
  const bool warp = ic >= 32;
  const int64_t threads = warp ? total * 32 : total;
  const int64_t blocks = cdiv(threads, 256);
  if (blocks > 0x7fffffffLL) return B2C_ERR_UNSUPPORTED;
  const T* ip = static_cast<const T*>(in); const T* wp = static_cast<const T*>(w); T* op = static_cast<T*>(out);
  if (!scatter && total <= 4096 && (int64_t)g.taps() * ic >= 8192) {
    gather_block_kernel<T><<<(unsigned)total, 256, 0, st>>>(g, ip, wp, op, bias, act, slope);
    return launched();
  }
  if (scatter) {
    if (warp) scatter_kernel<T, true><<<(unsigned)blocks, 256, 0, st>>>(g, ip, wp, op, bias, act, slope, total);
    else      scatter_kernel<T, false><<<(unsigned)blocks, 256, 0, st>>>(g, ip, wp, op, bias, act, slope, total);
  } else {
    if (warp) gather_kernel<T, true><<<(unsigned)blocks, 256, 0, st>>>(g, ip, wp, op, bias, act, slope, total);
    else      gather_kernel<T, false><<<(unsigned)blocks, 256, 0, st>>>(g, ip, wp, op, bias, act, slope, total);
  }
  return launched();
}

int direct_gather(const Geo& g, int dt, const void* in, const void* w, void* out, const float* bias, int act,
                  float slope, cudaStream_t st) {
  return dt == B2C_BF16 ? launch_fwdlike<__nv_bfloat16>(false, g, in, w, out, bias, act, slope, st)
                        : launch_fwdlike<float>(false, g, in, w, out, bias, act, slope, st);
}
int direct_scatter(const Geo& g, int dt, const void* in, const void* w, void* out, const float* bias, int act,
                   float slope, cudaStream_t st) {
  return dt == B2C_BF16 ? launch_fwdlike<__nv_bfloat16>(true, g, in, w, out, bias, act, slope, st)
                        : launch_fwdlike<float>(true, g, in, w, out, bias, act, slope, st);
}

// ---------------------------------------------------------------------------
// wgrad: dW[a][b][t] = sum_{n,o} small[n,o,a] * big[n,o*s+k*dl-p,b]
// Threads run over the wider channel dimension (coalesced), each keeps ACC
// accumulators for a chunk of (tap, other-channel) pairs; blocks split the
// voxel range and merge with fp32 atomics into the zeroed output.
constexpr int kWgAcc = 32;
template <typename T, bool OVER_A>
__global__ void __launch_bounds__(128)
wgrad_kernel(Geo g, const T* __restrict__ sm, const T* __restrict__ bg, float* __restrict__ dw, int64_t vox,
             int64_t vox_per_block) {
  const int T_ = g.taps();
  const int wide = OVER_A ? g.sc : g.bc, other = OVER_A ? g.bc : g.sc;
  const int ch = blockIdx.y * 128 + threadIdx.x;          // index in the wide dim
  const int pair0 = blockIdx.z * kWgAcc;                  // (t, other) pairs, other fastest
  const int npairs = T_ * other;
  const int64_t v0 = (int64_t)blockIdx.x * vox_per_block;
  const int64_t v1 = v0 + vox_per_block < vox ? v0 + vox_per_block : vox;
  float acc[kWgAcc];
#pragma unroll
  for (int i = 0; i < kWgAcc; i++) acc[i] = 0.f;
  const bool live = ch < wide;
  for (int64_t v = v0; v < v1; v++) {
    int n, od, oh, ow;
    decode_vox(v, g.sd, g.sh, g.sw, n, od, oh, ow);
    const T* sp = sm + v * g.sc;
    float sv = 0.f;
    if (OVER_A && live) sv = ldf(sp, ch);
#pragma unroll
    for (int i = 0; i < kWgAcc; i++) {
      const int pr = pair0 + i;
      if (pr >= npairs) break;
      const int t = pr / other, oc = pr - t * other;
      const int kw = t % g.k[2], kh = (t / g.k[2]) % g.k[1], kd = t / (g.k[2] * g.k[1]);
      const int id = od * g.s[0] + kd * g.dl[0] - g.p[0];
      const int ih = oh * g.s[1] + kh * g.dl[1] - g.p[1];
      const int iw = ow * g.s[2] + kw * g.dl[2] - g.p[2];
      if (id < 0 || id >= g.bd || ih < 0 || ih >= g.bh || iw < 0 || iw >= g.bw) continue;
      const T* bp = bg + ((((int64_t)n * g.bd + id) * g.bh + ih) * g.bw + iw) * g.bc;
      if (live) acc[i] += OVER_A ? sv * ldf(bp, oc) : ldf(sp, oc) * ldf(bp, ch);
    }
  }
  if (!live) return;
#pragma unroll
  for (int i = 0; i < kWgAcc; i++) {
    const int pr = pair0 + i;
    if (pr >= npairs) break;
    const int t = pr / other, oc = pr - t * other;
    const int a = OVER_A ? ch : oc, b = OVER_A ? oc : ch;
    atomicAdd(dw + ((int64_t)a * g.bc + b) * T_ + t, acc[i]);
  }
}

template <typename T>
static int wgrad_t(const Geo& g, const void* sm, const void* bg, float* dw, cudaStream_t st) {
  const int T_ = g.taps();
  const int64_t nel = (int64_t)g.sc * g.bc * T_;
  int rc = cuda_rc(cudaMemsetAsync(dw, 0, nel * sizeof(float), st));
  if (rc) return rc;
  const int64_t vox = (int64_t)g.n * g.sd * g.sh * g.sw;
  const bool over_a = g.sc >= g.bc;
  const int wide = over_a ? g.sc : g.bc, other = over_a ? g.bc : g.sc;
  const int64_t gy = cdiv(wide, 128), gz = cdiv((int64_t)T_ * other, kWgAcc);
  if (gy > 65535 || gz > 65535) return B2C_ERR_UNSUPPORTED;
  int64_t gx = cdiv((int64_t)num_sms() * 8, gy * gz);
  if (gx < 1) gx = 1;
  if (gx > vox) gx = vox;
  const int64_t vpb = cdiv(vox, gx);
  gx = cdiv(vox, vpb);
  dim3 grid((unsigned)gx, (unsigned)gy, (unsigned)gz);
  const T* sp = static_cast<const T*>(sm); const T* bp = static_cast<const T*>(bg);
  if (over_a) wgrad_kernel<T, true><<<grid, 128, 0, st>>>(g, sp, bp, dw, vox, vpb);
  else        wgrad_kernel<T, false><<<grid, 128, 0, st>>>(g, sp, bp, dw, vox, vpb);
  return launched();
}

int direct_wgrad(const Geo& g, int dt, const void* sm, const void* bg, float* dw, cudaStream_t st) {
  return dt == B2C_BF16 ? wgrad_t<__nv_bfloat16>(g, sm, bg, dw, st) : wgrad_t<float>(g, sm, bg, dw, st);
}

}  // namespace b2c
