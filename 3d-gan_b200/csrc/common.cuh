// common.cuh — shared device/host helpers for libb2c (sm_100a only).
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>

#include "../../include/b2c.h"

namespace b2c {

extern std::atomic<uint64_t> g_launches;   // b2c_launch_count()
extern thread_local int t_last_algo;       // b2c_last_algo()

inline int cuda_rc(cudaError_t e) { return e == cudaSuccess ? B2C_OK : B2C_ERR_CUDA; }
// Call after every kernel launch: counts it and converts launch errors.
inline int launched() {
  g_launches.fetch_add(1, std::memory_order_relaxed);
  return cuda_rc(cudaGetLastError());
}
inline bool dt_ok(int dt) { return dt == B2C_BF16 || dt == B2C_F32; }
inline int esize(int dt) { return dt == B2C_BF16 ? 2 : 4; }
inline int64_t cdiv(int64_t a, int64_t b) { return (a + b - 1) / b; }
int num_sms();

// Experiment switches (include/b2c_debug.h lists them).  Read from the environment ONCE, when the library first needs
// them - never on the launch path: getenv is not safe against a concurrent setenv from another Python thread.
struct Knobs {
  int swap_min;        // B2C_SWAP_MIN   : smallest Cout that takes the swapped (channels on M) epilogue
  bool mt1;            // B2C_MT1        : never use 256 x 256 tiles
  bool no_pair;        // B2C_NO_PAIR    : no CTA pairs (cta_group::2)
  int dbg;             // B2C_DBG        : timing experiments (8: no MMAs, 16: no loads)
  int stages;          // B2C_STAGES     : cap of the smem pipeline depth (0 = no cap)
  bool wg_mt1;         // B2C_WG_MT1     : weight-gradient kernel with one M sub-tile
  bool thin_unfused;   // B2C_THIN_UNFUSED: one-channel layers through im2col + GEMM + col2im
  bool col2im_ldg;     // B2C_COL2IM_LDG : k4 / s2 col2im with global 16-byte loads instead of the TMA-staged kernel
};
const Knobs& knobs();

// ---- element access -------------------------------------------------------
__device__ __forceinline__ float ldf(const __nv_bfloat16* p, int64_t i) { return __bfloat162float(p[i]); }
__device__ __forceinline__ float ldf(const float* p, int64_t i) { return p[i]; }
__device__ __forceinline__ void stf(__nv_bfloat16* p, int64_t i, float v) { p[i] = __float2bfloat16_rn(v); }
__device__ __forceinline__ void stf(float* p, int64_t i, float v) { p[i] = v; }

// 16-byte vectors: 8 bf16 or 4 fp32.
template <typename T> struct Vec;
template <> struct Vec<__nv_bfloat16> {
  static constexpr int N = 8;
  uint4 raw;
  __device__ __forceinline__ void load(const __nv_bfloat16* p) { raw = *reinterpret_cast<const uint4*>(p); }
  __device__ __forceinline__ void store(__nv_bfloat16* p) const { *reinterpret_cast<uint4*>(p) = raw; }
  __device__ __forceinline__ void zero() { raw = make_uint4(0u, 0u, 0u, 0u); }
  __device__ __forceinline__ void unpack(float* f) const {
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&raw);
#pragma unroll
    for (int i = 0; i < 4; i++) { float2 t = __bfloat1622float2(h[i]); f[2 * i] = t.x; f[2 * i + 1] = t.y; }
  }
  __device__ __forceinline__ void pack(const float* f) {
    __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&raw);
#pragma unroll
    for (int i = 0; i < 4; i++) h[i] = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
  }
};
template <> struct Vec<float> {
  static constexpr int N = 4;
  float4 raw;
  __device__ __forceinline__ void load(const float* p) { raw = *reinterpret_cast<const float4*>(p); }
  __device__ __forceinline__ void store(float* p) const { *reinterpret_cast<float4*>(p) = raw; }
  __device__ __forceinline__ void zero() { raw = make_float4(0.f, 0.f, 0.f, 0.f); }
  __device__ __forceinline__ void unpack(float* f) const { f[0] = raw.x; f[1] = raw.y; f[2] = raw.z; f[3] = raw.w; }
  __device__ __forceinline__ void pack(const float* f) { raw = make_float4(f[0], f[1], f[2], f[3]); }
};

// ---- activations ------------------------------------------------------------
// tanhf is ~60 instructions; when the activation switch is if-converted the compiler
// evaluates it for EVERY element even for relu/none (measured: 5x slower epilogues).
// A noinline call keeps it behind a real (warp-uniform) branch.
static __device__ __noinline__ float tanh_exact(float z) { return tanhf(z); }

__device__ __forceinline__ float act_f(float z, int act, float slope) {
  if (act == B2C_ACT_TANH) return tanh_exact(z);
  if (act == B2C_ACT_RELU) return fmaxf(z, 0.f);
  if (act == B2C_ACT_LRELU) return z > 0.f ? z : slope * z;
  return z;
}
// d act / d z from the pre-activation
__device__ __forceinline__ float act_dz(float z, int act, float slope) {
  if (act == B2C_ACT_TANH) { float t = tanh_exact(z); return 1.f - t * t; }
  if (act == B2C_ACT_RELU) return z > 0.f ? 1.f : 0.f;
  if (act == B2C_ACT_LRELU) return z > 0.f ? 1.f : slope;
  return 1.f;
}
// d act / d z from the output y = act(z)
__device__ __forceinline__ float act_dy(float y, int act, float slope) {
  switch (act) {
    case B2C_ACT_RELU:  return y > 0.f ? 1.f : 0.f;
    case B2C_ACT_LRELU: return y > 0.f ? 1.f : slope;
    case B2C_ACT_TANH:  return 1.f - y * y;
    default:            return 1.f;
  }
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// ---- conv geometry shared by direct and tensor-core paths --------------------
// Every pass is expressed on a (small, big) tensor pair: small index o maps to
// big index o*stride + k*dil - pad.  conv: small=y, big=x; convT: small=x, big=y.
struct Geo {
  int n;
  int sc, sd, sh, sw;      // small tensor: channels, D,H,W
  int bc, bd, bh, bw;      // big tensor
  int k[3], s[3], p[3], dl[3];
  __host__ __device__ int taps() const { return k[0] * k[1] * k[2]; }
};

int check_desc(const b2c_conv_desc* c);
Geo make_geo(const b2c_conv_desc* c);

}  // namespace b2c
