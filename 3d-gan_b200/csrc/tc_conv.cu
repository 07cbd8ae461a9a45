// tc_conv.cu — tcgen05/TMEM implicit-GEMM convolution kernels for sm_100a.
//
//   tc_conv_kernel  : out[voxel, co] = sum_{tap, ci} in[voxel + off(tap), ci] * W[tap][co][ci]
//                     (conv fwd, conv dgrad, convT fwd, convT dgrad via tc_plan.h)
//   tc_wgrad_kernel : dW[tap][a][b]  = sum_voxel small[voxel, a] * big[voxel + off(tap), b]
//
// Structure of both: warp 0 = TMA producer (one elected lane), warp 1 = TMEM
// allocator + single-thread tcgen05.mma issuer, warps 2..5 = epilogue (one warp
// per 32-lane TMEM sub-partition).  Operands are staged by TMA into 128-byte
// swizzled shared memory through an mbarrier ring; the accumulator lives in
// TMEM; im2col never exists: each filter tap is a TMA box of the NDHWC input at
// a shifted coordinate, with the zero padding supplied by TMA's OOB fill.
#include "tc_conv.cuh"

#include <mutex>

#include "ptx.cuh"

namespace b2c {

using namespace ptx;

constexpr int kThreadsTc = 192;
constexpr uint32_t kABytes = kBlockM * kBlockK * 2;   // 16 KB: 128 rows x 128 B

struct alignas(64) TcConvParams {
  CUtensorMap in_maps[kMaxMaps];
  CUtensorMap w_map;
  SubDesc subs[kMaxSubs];
  TapDesc taps[kMaxTaps];
  int32_t nsubs;
  int32_t box[4];
  int32_t nout, block_n, tiles_n, kchunks;
  int32_t stages;
  uint32_t idesc, tmem_cols;
  __nv_bfloat16* out;
  const float* bias;
  int32_t bias_mod;      // bias index = column % bias_mod (point-convT: columns are (tap, co))
  int32_t act;
  float slope;
};

__device__ __forceinline__ uint32_t pack_bf16x2(float a, float b) {
  __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&h);
}

__global__ void __launch_bounds__(kThreadsTc, 1) tc_conv_kernel(const __grid_constant__ TcConvParams P) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t sbase = (raw + 1023u) & ~1023u;               // SW128 needs 1024-B alignment
  uint8_t* sgen = smem_raw + (sbase - raw);
  const int stages = P.stages;
  const uint32_t b_bytes = (uint32_t)P.block_n * 128u;
  const uint32_t stage_bytes = kABytes + b_bytes;
  const uint32_t bar0 = sbase + stages * stage_bytes;           // full[stages], empty[stages], acc
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(sgen + stages * stage_bytes + (2 * stages + 1) * 8);
  auto full_bar = [&](int s) { return bar0 + 8u * s; };
  auto empty_bar = [&](int s) { return bar0 + 8u * (stages + s); };
  const uint32_t acc_bar = bar0 + 8u * (2 * stages);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  // ---- tile decode ---------------------------------------------------------
  const int nt = blockIdx.x % P.tiles_n;
  const int mt = blockIdx.x / P.tiles_n;
  int si = 0;
  for (int s = 1; s < P.nsubs; s++) if (mt >= P.subs[s].tile0) si = s;
  const SubDesc& sub = P.subs[si];
  int local = mt - sub.tile0;
  const int tw = local % sub.tiles[3]; local /= sub.tiles[3];
  const int th = local % sub.tiles[2]; local /= sub.tiles[2];
  const int td = local % sub.tiles[1]; local /= sub.tiles[1];
  const int tn = local;
  const int q0n = tn * P.box[0], q0d = td * P.box[1], q0h = th * P.box[2], q0w = tw * P.box[3];
  const int iters = sub.ntaps * P.kchunks;

  // ---- setup ---------------------------------------------------------------
  if (warp == 0 && lane == 0) {
    for (int s = 0; s < stages; s++) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), 1); }
    mbar_init(acc_bar, 1);
    fence_barrier_init();
    prefetch_tmap(&P.w_map);
    prefetch_tmap(&P.in_maps[0]);
  }
  if (warp == 1) {
    tmem_alloc(smem_u32(tmem_slot), P.tmem_cols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(tmem_slot);

  if (warp == 0) {
    // ===== TMA producer =====
    if (lane == 0) {
      int it = 0;
      for (int t = 0; t < sub.ntaps; t++) {
        const TapDesc tp = P.taps[sub.tap0 + t];
        const CUtensorMap* im = &P.in_maps[tp.map];
        for (int kc = 0; kc < P.kchunks; kc++, it++) {
          const int s = it % stages;
          const uint32_t ph = (uint32_t)(it / stages) & 1u;
          if (!mbar_wait(empty_bar(s), ph ^ 1u)) break;
          mbar_arrive_expect_tx(full_bar(s), stage_bytes);
          const uint32_t a_dst = sbase + s * stage_bytes;
          tma_load_5d(a_dst, im, full_bar(s), kc * kBlockK, q0w + tp.off_w, q0h + tp.off_h, q0d + tp.off_d, q0n);
          tma_load_3d(a_dst + kABytes, &P.w_map, full_bar(s), kc * kBlockK, nt * P.block_n, tp.wtap);
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer (single thread) =====
    if (lane == 0) {
      constexpr uint64_t hi = smem_desc_hi(16, 1024, 2);         // K-major, SWIZZLE_128B
      bool ok = true;
      for (int it = 0; it < iters && ok; it++) {
        const int s = it % stages;
        const uint32_t ph = (uint32_t)(it / stages) & 1u;
        ok = mbar_wait(full_bar(s), ph);
        tc_fence_after();
        const uint32_t a_addr = sbase + s * stage_bytes, b_addr = a_addr + kABytes;
#pragma unroll
        for (int k = 0; k < kBlockK / 16; k++)
          umma_f16(tmem_base, smem_desc(a_addr + k * 32, hi), smem_desc(b_addr + k * 32, hi), P.idesc,
                   (it | k) != 0 ? 1u : 0u);
        umma_commit(empty_bar(s));       // frees the smem slot when these MMAs retire
      }
      if (iters > 0) umma_commit(acc_bar);   // accumulator complete
      else mbar_arrive(acc_bar);
    }
  } else {
    // ===== epilogue: TMEM -> registers -> bias/activation -> bf16 -> global =====
    const int grp = warp & 3;                    // TMEM sub-partition of this warp
    const int row = grp * 32 + lane;
    int r = row;
    const int rw = r % P.box[3]; r /= P.box[3];
    const int rh = r % P.box[2]; r /= P.box[2];
    const int rd = r % P.box[1]; r /= P.box[1];
    const int qn = q0n + r, qd = q0d + rd, qh = q0h + rh, qw = q0w + rw;
    const bool valid = qn < sub.ext[0] && qd < sub.ext[1] && qh < sub.ext[2] && qw < sub.ext[3];
    __nv_bfloat16* orow = P.out + sub.obase + qn * sub.ostr[0] + qd * sub.ostr[1] + qh * sub.ostr[2] +
                          qw * sub.ostr[3];
    mbar_wait(acc_bar, 0);
    tc_fence_after();
    const uint32_t trow = tmem_base + ((uint32_t)(grp * 32) << 16);
    for (int c0 = 0; c0 < P.block_n; c0 += 16) {
      uint32_t v[16];
      if (iters > 0) { tmem_ld16(trow + c0, v); tmem_ld_wait(); }
      else {
#pragma unroll
        for (int j = 0; j < 16; j++) v[j] = 0u;
      }
      const int col = nt * P.block_n + c0;
      if (valid) {
#pragma unroll
        for (int j = 0; j < 16; j += 8) {
          if (col + j < P.nout) {                 // nout % 8 == 0: whole 8-group in range
            float f[8];
#pragma unroll
            for (int i = 0; i < 8; i++) {
              float x = __uint_as_float(v[j + i]);
              if (P.bias) x += __ldg(P.bias + (col + j + i) % P.bias_mod);
              f[i] = act_f(x, P.act, P.slope);
            }
            uint4 o;
            o.x = pack_bf16x2(f[0], f[1]); o.y = pack_bf16x2(f[2], f[3]);
            o.z = pack_bf16x2(f[4], f[5]); o.w = pack_bf16x2(f[6], f[7]);
            *reinterpret_cast<uint4*>(orow + col + j) = o;
          }
        }
      }
    }
  }

  // ---- teardown --------------------------------------------------------------
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, P.tmem_cols);
}


// =============================================================================
// Weight gradient.  GEMM with the voxel index as K: both operands are read in
// their natural NDHWC form, i.e. MN-major (channels contiguous), one TMA box of
// 64 voxels x 64 channels per 8 KB slab.
//   A (M = 128 small-tensor channels) : 2 slabs,  B (N = block_n big channels).
// grid = taps * a_tiles * b_tiles * splits; partial tiles go to
// ws[split][tap][a][b] (fp32) and are reduced + transposed to torch layout by
// wgrad_reduce_kernel (deterministic: fixed split order, no atomics).
// =============================================================================
constexpr int kWgVox = 64;                                   // voxels per stage (K)
constexpr uint32_t kSlabBytes = kWgVox * 128;                // 8 KB

struct alignas(64) TcWgradParams {
  CUtensorMap s_map;
  CUtensorMap b_maps[kMaxMaps];
  TapDesc taps[kMaxTaps];
  int32_t ntaps;
  int32_t box[4], tiles[4];          // voxel tile box (product 64) and counts over {n,d,h,w}
  int32_t ktiles, ktiles_per_split, splits;
  int32_t a_ch, b_ch, a_tiles, b_tiles, block_n;
  int32_t stages;
  uint32_t idesc, tmem_cols;
  float* ws;
};

__global__ void __launch_bounds__(kThreadsTc, 1) tc_wgrad_kernel(const __grid_constant__ TcWgradParams P) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t sbase = (raw + 1023u) & ~1023u;
  uint8_t* sgen = smem_raw + (sbase - raw);
  const int stages = P.stages;
  const int b_slabs = P.block_n / 64;
  const uint32_t a_bytes = 2 * kSlabBytes, b_bytes = b_slabs * kSlabBytes;
  const uint32_t stage_bytes = a_bytes + b_bytes;
  const uint32_t bar0 = sbase + stages * stage_bytes;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(sgen + stages * stage_bytes + (2 * stages + 1) * 8);
  auto full_bar = [&](int s) { return bar0 + 8u * s; };
  auto empty_bar = [&](int s) { return bar0 + 8u * (stages + s); };
  const uint32_t acc_bar = bar0 + 8u * (2 * stages);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  int id = blockIdx.x;
  const int bt = id % P.b_tiles; id /= P.b_tiles;
  const int at = id % P.a_tiles; id /= P.a_tiles;
  const int ti = id % P.ntaps;
  const int sp = id / P.ntaps;
  const TapDesc tp = P.taps[ti];
  const int kt0 = sp * P.ktiles_per_split;
  const int kt1 = kt0 + P.ktiles_per_split < P.ktiles ? kt0 + P.ktiles_per_split : P.ktiles;
  const int iters = kt1 - kt0;

  if (warp == 0 && lane == 0) {
    for (int s = 0; s < stages; s++) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), 1); }
    mbar_init(acc_bar, 1);
    fence_barrier_init();
    prefetch_tmap(&P.s_map);
    prefetch_tmap(&P.b_maps[tp.map]);
  }
  if (warp == 1) { tmem_alloc(smem_u32(tmem_slot), P.tmem_cols); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(tmem_slot);

  if (warp == 0) {
    if (lane == 0) {
      const CUtensorMap* bm = &P.b_maps[tp.map];
      for (int it = 0; it < iters; it++) {
        int kt = kt0 + it;
        const int tw = kt % P.tiles[3]; kt /= P.tiles[3];
        const int th = kt % P.tiles[2]; kt /= P.tiles[2];
        const int td = kt % P.tiles[1]; kt /= P.tiles[1];
        const int q0n = kt * P.box[0], q0d = td * P.box[1], q0h = th * P.box[2], q0w = tw * P.box[3];
        const int s = it % stages;
        const uint32_t ph = (uint32_t)(it / stages) & 1u;
        if (!mbar_wait(empty_bar(s), ph ^ 1u)) break;
        mbar_arrive_expect_tx(full_bar(s), stage_bytes);
        const uint32_t a_dst = sbase + s * stage_bytes;
        for (int j = 0; j < 2; j++)
          tma_load_5d(a_dst + j * kSlabBytes, &P.s_map, full_bar(s), at * 128 + j * 64, q0w, q0h, q0d, q0n);
        for (int j = 0; j < b_slabs; j++)
          tma_load_5d(a_dst + a_bytes + j * kSlabBytes, bm, full_bar(s), bt * P.block_n + j * 64, q0w + tp.off_w,
                      q0h + tp.off_h, q0d + tp.off_d, q0n);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      // MN-major SW128: LBO = slab pitch (next 64 channels), SBO = 1024 (next 8 voxel rows)
      constexpr uint64_t hi = smem_desc_hi(kSlabBytes, 1024, 2);
      bool ok = true;
      for (int it = 0; it < iters && ok; it++) {
        const int s = it % stages;
        const uint32_t ph = (uint32_t)(it / stages) & 1u;
        ok = mbar_wait(full_bar(s), ph);
        tc_fence_after();
        const uint32_t a_addr = sbase + s * stage_bytes, b_addr = a_addr + a_bytes;
#pragma unroll
        for (int k = 0; k < kWgVox / 16; k++)     // 16 voxel rows of 128 B per MMA
          umma_f16(tmem_base, smem_desc(a_addr + k * 2048, hi), smem_desc(b_addr + k * 2048, hi), P.idesc,
                   (it | k) != 0 ? 1u : 0u);
        umma_commit(empty_bar(s));
      }
      if (iters > 0) umma_commit(acc_bar);
      else mbar_arrive(acc_bar);
    }
  } else {
    const int grp = warp & 3;
    const int a = at * 128 + grp * 32 + lane;
    const bool valid = a < P.a_ch;
    float* orow = P.ws + (((int64_t)sp * P.ntaps + ti) * P.a_ch + a) * P.b_ch;
    mbar_wait(acc_bar, 0);
    tc_fence_after();
    const uint32_t trow = tmem_base + ((uint32_t)(grp * 32) << 16);
    for (int c0 = 0; c0 < P.block_n; c0 += 16) {
      uint32_t v[16];
      if (iters > 0) { tmem_ld16(trow + c0, v); tmem_ld_wait(); }
      else {
#pragma unroll
        for (int j = 0; j < 16; j++) v[j] = 0u;
      }
      const int col = bt * P.block_n + c0;
      if (valid) {
#pragma unroll
        for (int j = 0; j < 16; j += 4)
          if (col + j < P.b_ch)                 // b_ch % 8 == 0
            *reinterpret_cast<uint4*>(orow + col + j) = make_uint4(v[j], v[j + 1], v[j + 2], v[j + 3]);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, P.tmem_cols);
}

// dw[(a*B + b)*T + t] = sum_split ws[split][tapidx(t)][a][b]   (0 when the tap only sees padding)
struct TapIndex { int16_t idx[kMaxTaps]; };
__global__ void __launch_bounds__(256)
wgrad_reduce_kernel(const float* __restrict__ ws, float* __restrict__ dw, int A, int B, int T, int ntaps, int splits,
                    TapIndex ti) {
  const int64_t ab = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (ab >= (int64_t)A * B) return;
  const int64_t plane = (int64_t)A * B;
  for (int t = 0; t < T; t++) {
    float acc = 0.f;
    const int idx = ti.idx[t];
    if (idx >= 0)
      for (int s = 0; s < splits; s++) acc += ws[((int64_t)s * ntaps + idx) * plane + ab];
    dw[ab * T + t] = acc;
  }
}


// =============================================================================
// Host side
// =============================================================================
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// cuTensorMapEncodeTiled is resolved through the runtime so that libb2c.so has
// no link-time dependency on libcuda.so (it must load in a GPU-less container).
static EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  });
  return fn;
}

// bf16 view `m` of `base`, box = {64 channels, box[1..rank-1]}, 128-B swizzle, zero OOB fill.
static int encode_map(CUtensorMap* out, const void* base, const MapDesc& m, int rank, const int* box) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) return B2C_ERR_NO_DEVICE;
  cuuint64_t dims[5], strides[4];
  cuuint32_t bx[5], es[5];
  for (int i = 0; i < rank; i++) {
    dims[i] = (cuuint64_t)m.dims[i];
    bx[i] = (cuuint32_t)box[i];
    es[i] = 1;
    if (i > 0) strides[i - 1] = (cuuint64_t)m.strides[i] * 2;
  }
  void* addr = const_cast<void*>(static_cast<const void*>(static_cast<const __nv_bfloat16*>(base) + m.base));
  CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, addr, dims, strides, bx, es,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? B2C_OK : B2C_ERR_CUDA;
}

static uint32_t tmem_cols_for(int n) { uint32_t c = 32; while ((int)c < n) c <<= 1; return c; }
constexpr int kSmemBudget = 227 * 1024;

int tc_conv_launch(const ConvPlan& p, const void* in, const void* w, void* out, const float* bias, int bias_mod,
                   int act, float slope, cudaStream_t st) {
  TcConvParams P;
  memset(&P, 0, sizeof(P));
  const int abox[5] = {kBlockK, p.box[3], p.box[2], p.box[1], p.box[0]};
  for (int i = 0; i < p.nmaps; i++) {
    int rc = encode_map(&P.in_maps[i], in, p.maps[i], 5, abox);
    if (rc) return rc;
  }
  MapDesc wm;
  wm.base = 0;
  const int T = [&] { int mx = 0; for (int i = 0; i < p.ntaps_total; i++) if (p.taps[i].wtap > mx) mx = p.taps[i].wtap; return mx + 1; }();
  wm.dims[0] = p.kin; wm.dims[1] = p.nout; wm.dims[2] = T;
  wm.strides[0] = 1; wm.strides[1] = p.kin; wm.strides[2] = (int64_t)p.kin * p.nout;
  const int wbox[3] = {kBlockK, p.block_n, 1};
  int rc = encode_map(&P.w_map, w, wm, 3, wbox);
  if (rc) return rc;
  memcpy(P.subs, p.subs, sizeof(P.subs));
  memcpy(P.taps, p.taps, sizeof(P.taps));
  P.nsubs = p.nsubs;
  for (int i = 0; i < 4; i++) P.box[i] = p.box[i];
  P.nout = p.nout; P.block_n = p.block_n; P.tiles_n = p.tiles_n; P.kchunks = p.kchunks;
  const uint32_t stage_bytes = kABytes + (uint32_t)p.block_n * 128u;
  int stages = (kSmemBudget - 1024 - 256) / (int)stage_bytes;
  if (stages > 8) stages = 8;
  if (stages < 2) return B2C_ERR_UNSUPPORTED;
  P.stages = stages;
  P.idesc = idesc_bf16(kBlockM, p.block_n, 0, 0);
  P.tmem_cols = tmem_cols_for(p.block_n);
  P.out = static_cast<__nv_bfloat16*>(out);
  P.bias = bias; P.bias_mod = bias_mod > 0 ? bias_mod : p.nout; P.act = act; P.slope = slope;
  const size_t smem = (size_t)stages * stage_bytes + 1024 + 256;
  // the attribute is per device; set it on every call (cheap) rather than caching per device
  if (cudaFuncSetAttribute(tc_conv_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBudget) != cudaSuccess)
    return B2C_ERR_CUDA;
  const int64_t grid = (int64_t)p.m_tiles * p.tiles_n;
  if (grid <= 0) return B2C_OK;
  tc_conv_kernel<<<(unsigned)grid, kThreadsTc, smem, st>>>(P);
  return launched();
}

// ---- wgrad ---------------------------------------------------------------------
struct WgradCfg { ConvPlan plan; int box[4], tiles[4]; int ktiles, splits, kps, a_tiles, b_tiles, block_n; };

static int wgrad_cfg(const Geo& g, WgradCfg* c) {
  if (g.sc % 8 != 0 || g.bc % 8 != 0 || g.sc < 16 || g.bc < 16) return B2C_ERR_UNSUPPORTED;
  // phase maps + taps are those of the gather form (lattice = small tensor, reads from big)
  int rc = plan_gather(g, &c->plan);
  if (rc) return rc;
  const int ext[4] = {g.n, g.sd, g.sh, g.sw};
  int rem = kWgVox;
  auto p2 = [](int v) { int p = 1; while (p < v) p <<= 1; return p; };
  c->box[3] = p2(ext[3]) < rem ? p2(ext[3]) : rem; rem /= c->box[3];
  c->box[2] = p2(ext[2]) < rem ? p2(ext[2]) : rem; rem /= c->box[2];
  c->box[1] = p2(ext[1]) < rem ? p2(ext[1]) : rem; rem /= c->box[1];
  c->box[0] = rem;
  int64_t kt = 1;
  for (int i = 0; i < 4; i++) { c->tiles[i] = (ext[i] + c->box[i] - 1) / c->box[i]; kt *= c->tiles[i]; }
  if (kt > 0x7fffffff) return B2C_ERR_UNSUPPORTED;
  c->ktiles = (int)kt;
  int bn = ((g.bc + 63) / 64) * 64;
  if (bn > 256) bn = 256;
  c->block_n = bn;
  c->a_tiles = (g.sc + 127) / 128;
  c->b_tiles = (g.bc + bn - 1) / bn;
  const int64_t units = (int64_t)c->plan.ntaps_total * c->a_tiles * c->b_tiles;
  if (units <= 0) return B2C_ERR_UNSUPPORTED;
  int64_t splits = cdiv((int64_t)num_sms() * 3, units);
  if (splits > c->ktiles) splits = c->ktiles;
  // bound the fp32 partial buffer to 1 GiB
  const int64_t plane = (int64_t)c->plan.ntaps_total * g.sc * g.bc * 4;
  while (splits > 1 && splits * plane > (1LL << 30)) splits--;
  if (splits < 1) splits = 1;
  c->kps = (int)cdiv(c->ktiles, splits);
  c->splits = (int)cdiv(c->ktiles, c->kps);
  if (units * c->splits > 0x7fffffffLL) return B2C_ERR_UNSUPPORTED;
  return B2C_OK;
}

bool tc_wgrad_supported(const Geo& g) { WgradCfg c; return wgrad_cfg(g, &c) == B2C_OK; }

size_t tc_wgrad_workspace(const Geo& g) {
  WgradCfg c;
  if (wgrad_cfg(g, &c)) return 0;
  return (size_t)c.splits * c.plan.ntaps_total * g.sc * g.bc * sizeof(float);
}

int tc_wgrad_launch(const Geo& g, const void* small, const void* big, float* dw, void* ws, size_t wsb,
                    cudaStream_t st) {
  WgradCfg c;
  int rc = wgrad_cfg(g, &c);
  if (rc) return rc;
  const size_t need = (size_t)c.splits * c.plan.ntaps_total * g.sc * g.bc * sizeof(float);
  if (!ws || wsb < need) return B2C_ERR_WORKSPACE;
  TcWgradParams P;
  memset(&P, 0, sizeof(P));
  const int box5[5] = {64, c.box[3], c.box[2], c.box[1], c.box[0]};
  MapDesc sm;
  sm.base = 0;
  sm.dims[0] = g.sc; sm.dims[1] = g.sw; sm.dims[2] = g.sh; sm.dims[3] = g.sd; sm.dims[4] = g.n;
  sm.strides[0] = 1; sm.strides[1] = g.sc; sm.strides[2] = (int64_t)g.sw * g.sc;
  sm.strides[3] = (int64_t)g.sh * g.sw * g.sc; sm.strides[4] = (int64_t)g.sd * g.sh * g.sw * g.sc;
  rc = encode_map(&P.s_map, small, sm, 5, box5);
  if (rc) return rc;
  for (int i = 0; i < c.plan.nmaps; i++) {
    rc = encode_map(&P.b_maps[i], big, c.plan.maps[i], 5, box5);
    if (rc) return rc;
  }
  memcpy(P.taps, c.plan.taps, sizeof(P.taps));
  P.ntaps = c.plan.ntaps_total;
  for (int i = 0; i < 4; i++) { P.box[i] = c.box[i]; P.tiles[i] = c.tiles[i]; }
  P.ktiles = c.ktiles; P.ktiles_per_split = c.kps; P.splits = c.splits;
  P.a_ch = g.sc; P.b_ch = g.bc; P.a_tiles = c.a_tiles; P.b_tiles = c.b_tiles; P.block_n = c.block_n;
  const uint32_t stage_bytes = 2 * kSlabBytes + (c.block_n / 64) * kSlabBytes;
  int stages = (kSmemBudget - 1024 - 256) / (int)stage_bytes;
  if (stages > 8) stages = 8;
  P.stages = stages;
  P.idesc = idesc_bf16(128, c.block_n, 1, 1);
  P.tmem_cols = tmem_cols_for(c.block_n);
  P.ws = static_cast<float*>(ws);
  if (cudaFuncSetAttribute(tc_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBudget) != cudaSuccess)
    return B2C_ERR_CUDA;
  const size_t smem = (size_t)stages * stage_bytes + 1024 + 256;
  const int64_t grid = (int64_t)P.ntaps * c.a_tiles * c.b_tiles * c.splits;
  tc_wgrad_kernel<<<(unsigned)grid, kThreadsTc, smem, st>>>(P);
  rc = launched();
  if (rc) return rc;
  TapIndex ti;
  const int T = g.taps();
  for (int t = 0; t < kMaxTaps; t++) ti.idx[t] = -1;
  for (int i = 0; i < c.plan.ntaps_total; i++) ti.idx[c.plan.taps[i].wtap] = (int16_t)i;
  const int64_t ab = (int64_t)g.sc * g.bc;
  wgrad_reduce_kernel<<<(unsigned)cdiv(ab, 256), 256, 0, st>>>(P.ws, dw, g.sc, g.bc, T, P.ntaps, c.splits, ti);
  return launched();
}

unsigned tc_error_fetch_clear() {
  unsigned v = 0, z = 0;
  if (cudaDeviceSynchronize() != cudaSuccess) return 0x80000000u;
  if (cudaMemcpyFromSymbol(&v, g_tc_error, sizeof(v)) != cudaSuccess) return 0x80000000u;
  cudaMemcpyToSymbol(g_tc_error, &z, sizeof(z));
  return v;
}

}  // namespace b2c
