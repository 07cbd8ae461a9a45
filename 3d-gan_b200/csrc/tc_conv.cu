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
#include <stdlib.h>

#include "ptx.cuh"

namespace b2c {

using namespace ptx;

constexpr int kThreadsTc = 192;
constexpr uint32_t kABytes = kBlockM * kBlockK * 2;   // 16 KB: 128 rows x 128 B

struct alignas(64) TcConvParams {
  CUtensorMap in_maps[kMaxMaps];
  CUtensorMap out_maps[kMaxSubs];
  CUtensorMap w_map;
  SubDesc subs[kMaxSubs];
  TapDesc taps[kMaxTaps];
  int32_t nsubs;
  int32_t box[4];
  int32_t mt, split_dim, sub_ext;
  int32_t nout, block_n, tiles_n, kchunks;
  int32_t num_tiles;
  int32_t interleave;    // all sub-problems have identical tile grids: order tiles spatial-major, phase-minor
  int32_t stages;
  uint32_t idesc, tmem_cols;
  const float* bias;
  int32_t bias_mod;      // bias index = column % bias_mod (point-convT: columns are (tap, co))
  int32_t kb;            // channels per stage: 64 (128-B rows, SWIZZLE_128B) or 32 (64-B rows, SWIZZLE_64B)
  int32_t act;
  float slope;
  int32_t swap;          // 1: channels on UMMA M (filter tile = A), voxels on N: D^T = W * X^T (narrow Cout)
  int32_t w_rows;        // filter rows staged per stage
  int32_t nacc;          // TMEM accumulator sets: 2 (epilogue of tile i overlaps tile i+1) or 1 (256 x 256 tiles)
  int32_t dbg;           // B2C_DBG env (diagnostics only): 1 no TMA store, 2 no tcgen05.ld, 4 no st.shared
  // statistics of the STORED output (after bias / activation / bf16 rounding) for the normalisation layer that
  // follows (dcgan.py:47-65 ConvT -> BatchNorm3d, dcgan.py:161-179 Conv -> LayerNorm): every CTA accumulates its
  // tiles in shared memory in a fixed order and writes ONE partial row; a second stage sums the rows (deterministic)
  int32_t stat_kind;     // 0 none, 1 per output channel (sum over voxels), 2 per sample (sum over voxels x channels)
  int32_t stat_len;      // channels (kind 1; = bias_mod for the point GEMM) or samples (kind 2)
  int32_t sb_log[3];     // log2 of the sub-tile box extents {w, h, d} (powers of two)
  float* stat_part;      // [gridDim.x][2][stat_len]
};

// hardware tanh (abs error ~5e-4 of full scale): far inside bf16 output resolution near +-1
__device__ __forceinline__ float tanh_approx(float x) {
  float y;
  asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

__device__ __forceinline__ uint32_t pack_bf16x2(float a, float b) {
  __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&h);
}

// activation over 64 accumulator values: one uniform branch, never a per-element switch
__device__ __forceinline__ void apply_act64(uint32_t* v, int act, float slope) {
  if (act == B2C_ACT_RELU) {
#pragma unroll
    for (int i = 0; i < 64; i++) v[i] = __float_as_uint(fmaxf(__uint_as_float(v[i]), 0.f));
  } else if (act == B2C_ACT_LRELU) {
#pragma unroll
    for (int i = 0; i < 64; i++) {
      const float x = __uint_as_float(v[i]);
      v[i] = __float_as_uint(x > 0.f ? x : x * slope);
    }
  } else if (act == B2C_ACT_TANH) {
#pragma unroll
    for (int i = 0; i < 64; i++) v[i] = __float_as_uint(tanh_approx(__uint_as_float(v[i])));
  }
}
__device__ __forceinline__ void st_shared_b16(uint32_t addr, uint16_t v) {
  asm volatile("st.shared.b16 [%0], %1;" ::"r"(addr), "h"(v) : "memory");
}

__device__ __forceinline__ float bf16_round(float x) { return __bfloat162float(__float2bfloat16_rn(x)); }
__device__ __forceinline__ uint4 ld_shared_v4(uint32_t addr) {
  uint4 v;
  asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
  return v;
}

// Which rows of a 128-row sub-tile lie inside the output lattice (the TMA store clips the others).
struct RowClip {
  int lim[4];            // valid extent of the sub-tile box per lattice dim {n, d, h, w}
  int l3, l2, l1;        // log2 of the box extents {w, h, d}
  bool full;
  __device__ __forceinline__ bool ok(int r) const {
    if (full) return true;
    const int iw = r & ((1 << l3) - 1), ih = (r >> l3) & ((1 << l2) - 1), id = (r >> (l3 + l2)) & ((1 << l1) - 1);
    const int in = r >> (l3 + l2 + l1);
    return iw < lim[3] && ih < lim[2] && id < lim[1] && in < lim[0];
  }
};

struct TileCoord { int si, nt, q[4], iters; };

// NCTA == 2: `tile` indexes PAIR tiles; the two CTAs of a pair take M tiles 2j and 2j+1 of
// the same sub-problem and the same N tile (an odd last one is fully out of range: TMA
// zero-fills its loads and clips its stores).
template <int NCTA>
__device__ __forceinline__ TileCoord decode_tile(const TcConvParams& P, int tile, int rank) {
  TileCoord c;
  c.nt = tile % P.tiles_n;
  const int mtile = tile / P.tiles_n;
  int si = 0, local;
  if (NCTA == 2) {
    si = P.nsubs > 1 ? mtile % P.nsubs : 0;
    local = 2 * (P.nsubs > 1 ? mtile / P.nsubs : mtile) + rank;
  } else if (P.interleave) {
    // the stride^3 sub-pixel phases of one spatial tile read the same input voxels:
    // run them back to back so the input is fetched from HBM once, not once per phase
    si = mtile % P.nsubs;
    local = mtile / P.nsubs;
  } else {
    for (int s = 1; s < P.nsubs; s++) if (mtile >= P.subs[s].tile0) si = s;
    local = mtile - P.subs[si].tile0;
  }
  c.si = si;
  const SubDesc& sub = P.subs[si];
  const int tw = local % sub.tiles[3]; local /= sub.tiles[3];
  const int th = local % sub.tiles[2]; local /= sub.tiles[2];
  const int td = local % sub.tiles[1]; local /= sub.tiles[1];
  c.q[0] = local * P.box[0]; c.q[1] = td * P.box[1]; c.q[2] = th * P.box[2]; c.q[3] = tw * P.box[3];
  c.iters = sub.ntaps * P.kchunks;
  return c;
}

// validity of the rows of the 128-row sub-tile whose origin is q (lattice coordinates {n, d, h, w})
__device__ __forceinline__ RowClip make_clip(const TcConvParams& P, const TileCoord& c, const int* q) {
  RowClip k;
  const SubDesc& sub = P.subs[c.si];
  k.l3 = P.sb_log[0]; k.l2 = P.sb_log[1]; k.l1 = P.sb_log[2];
  const int sb[4] = {kBlockM >> (k.l3 + k.l2 + k.l1), 1 << k.l1, 1 << k.l2, 1 << k.l3};
  k.full = true;
#pragma unroll
  for (int i = 0; i < 4; i++) {
    k.lim[i] = sub.ext[i] - q[i];
    if (k.lim[i] < sb[i]) k.full = false;
  }
  return k;
}

constexpr uint32_t kStagingBytes = 2 * kABytes;      // two 128-row x 64-column bf16 slabs (ring)
constexpr int kEpiThreads = 128;

// Persistent, warp-specialised: CTA b processes tiles b, b + grid, ...
//   warp 0      TMA producer: runs ahead across tiles through the smem ring
//   warp 1      tcgen05.mma issuer; two TMEM accumulator sets so that the epilogue of
//               tile i overlaps the main loop of tile i+1
//   warps 2..5  epilogue: TMEM -> registers -> bias/activation -> bf16 -> swizzled smem
//               slab -> TMA store (coalesced, clipped at the tensor edge)
// STATS: the epilogue also accumulates the statistics of P.stat_kind (a separate instantiation, so that launches
// without them keep the register allocation of the plain kernel).
template <int NCTA, bool STATS>
__global__ void __launch_bounds__(kThreadsTc, 1) tc_conv_kernel(const __grid_constant__ TcConvParams P) {
  extern __shared__ uint8_t smem_raw[];
  const int rank = NCTA == 2 ? (int)cluster_ctarank() : 0;     // rank 0 = leader: issues every MMA of the pair
  const int first_tile = NCTA == 2 ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;
  const int tile_step = NCTA == 2 ? (int)(gridDim.x >> 1) : (int)gridDim.x;
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t sbase = (raw + 1023u) & ~1023u;               // SW128 needs 1024-B alignment
  uint8_t* sgen = smem_raw + (sbase - raw);
  const int stages = P.stages;
  const uint32_t row_bytes = (uint32_t)P.kb * 2u;              // 128 or 64
  const uint32_t a_sub = (uint32_t)kBlockM * row_bytes;        // one 128-row sub-tile
  const uint32_t a_bytes = (uint32_t)P.mt * a_sub;
  const uint32_t b_rows = (uint32_t)P.w_rows / NCTA;           // pair: each CTA stages half of the filter tile
  const uint32_t b_bytes = b_rows * row_bytes;
  const uint32_t stage_bytes = a_bytes + b_bytes;
  const uint32_t staging = sbase + stages * stage_bytes;
  const uint32_t bar0 = staging + kStagingBytes;                // full[stages], empty[stages], tfull[2], tempty[2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(sgen + stages * stage_bytes + kStagingBytes + (2 * stages + 4) * 8);
  auto full_bar = [&](int s) { return bar0 + 8u * s; };
  auto empty_bar = [&](int s) { return bar0 + 8u * (stages + s); };
  auto tfull_bar = [&](int a) { return bar0 + 8u * (2 * stages + a); };
  auto tempty_bar = [&](int a) { return bar0 + 8u * (2 * stages + 2 + a); };

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    for (int s = 0; s < stages; s++) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), 1); }
    for (int a = 0; a < 2; a++) { mbar_init(tfull_bar(a), 1); mbar_init(tempty_bar(a), NCTA * kEpiThreads / 32); }
    fence_barrier_init();
    prefetch_tmap(&P.w_map);
    prefetch_tmap(&P.in_maps[0]);
    prefetch_tmap(&P.out_maps[0]);
  }
  if (warp == 1) {
    if (NCTA == 2) { tmem_alloc_pair(smem_u32(tmem_slot), P.tmem_cols); tmem_relinquish_pair(); }
    else { tmem_alloc(smem_u32(tmem_slot), P.tmem_cols); tmem_relinquish(); }
  }
  tc_fence_before();
  if (NCTA == 2) cluster_sync_all();      // peer barriers are initialised before anyone signals them
  else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(tmem_slot);
  const uint32_t acc_cols = (uint32_t)P.mt * (uint32_t)P.block_n;

  if (warp == 0) {
    // ===== TMA producer =====
    if (lane == 0) {
      uint32_t it = 0;
      bool ok = true;
      for (int tile = first_tile; tile < P.num_tiles && ok; tile += tile_step) {
        const TileCoord c = decode_tile<NCTA>(P, tile, rank);
        const SubDesc& sub = P.subs[c.si];
        for (int t = 0; t < sub.ntaps && ok; t++) {
          const TapDesc tp = P.taps[sub.tap0 + t];
          const CUtensorMap* im = &P.in_maps[tp.map];
          for (int kc = 0; kc < P.kchunks; kc++, it++) {
            const uint32_t s = it % stages;
            const uint32_t ph = (it / stages) & 1u;
            if (!mbar_wait(empty_bar(s), ph ^ 1u)) { ok = false; break; }
            const uint32_t a_dst = sbase + s * stage_bytes;
            if (NCTA == 2) {
              // both CTAs of the pair complete the LEADER's full barrier; only the leader arms it
              const uint32_t fb = mapa_u32(full_bar(s), 0);
              if (rank == 0) mbar_arrive_expect_tx(full_bar(s), 2 * stage_bytes);
              tma_load_5d_pair(a_dst, im, fb, kc * P.kb, c.q[3] + tp.off_w, c.q[2] + tp.off_h,
                               c.q[1] + tp.off_d, c.q[0]);
              tma_load_3d_pair(a_dst + a_bytes, &P.w_map, fb, kc * P.kb, c.nt * P.block_n + rank * (int)b_rows,
                               tp.wtap);
            } else if (P.dbg & 16) {
              mbar_arrive(full_bar(s));
            } else {
              mbar_arrive_expect_tx(full_bar(s), stage_bytes);
              tma_load_5d(a_dst, im, full_bar(s), kc * P.kb, c.q[3] + tp.off_w, c.q[2] + tp.off_h,
                          c.q[1] + tp.off_d, c.q[0]);
              tma_load_3d(a_dst + a_bytes, &P.w_map, full_bar(s), kc * P.kb, c.nt * P.block_n, tp.wtap);
            }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer (single thread; in a pair only the leader CTA issues) =====
    if (lane == 0 && rank == 0) {
      // K-major: SWIZZLE_128B (8-row atoms of 1024 B) or, for 32-channel stages, SWIZZLE_64B (512 B)
      const uint64_t hi = P.kb == 32 ? smem_desc_hi(16, 512, 4) : smem_desc_hi(16, 1024, 2);
      const int nk = P.kb / 16;
      uint32_t it = 0, tcount = 0;
      bool ok = true;
      for (int tile = first_tile; tile < P.num_tiles && ok; tile += tile_step, tcount++) {
        const TileCoord c = decode_tile<NCTA>(P, tile, 0);
        const uint32_t acc = P.nacc == 2 ? (tcount & 1u) : 0u;
        const uint32_t aph = (P.nacc == 2 ? (tcount >> 1) : tcount) & 1u;
        // the epilogues (of both CTAs) have drained this accumulator set
        ok = NCTA == 2 ? mbar_wait_cluster(tempty_bar(acc), aph ^ 1u) : mbar_wait(tempty_bar(acc), aph ^ 1u);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * acc_cols;
        for (int i = 0; i < c.iters && ok; i++, it++) {
          const uint32_t s = it % stages;
          const uint32_t ph = (it / stages) & 1u;
          ok = NCTA == 2 ? mbar_wait_cluster(full_bar(s), ph) : mbar_wait(full_bar(s), ph);
          tc_fence_after();
          const uint32_t a_addr = sbase + s * stage_bytes, b_addr = a_addr + a_bytes;
          if (P.swap) {
            // narrow Cout: the filter tile is the M = 128 operand, all 128*mt voxel rows are N
#pragma unroll
            for (int k = 0; k < kBlockK / 16; k++)
              if (k < nk)
                umma_f16(d_tmem, smem_desc(b_addr + k * 32, hi), smem_desc(a_addr + k * 32, hi), P.idesc,
                         (i | k) != 0 ? 1u : 0u);
          } else
          for (int m = 0; m < P.mt; m++) {
#pragma unroll
            for (int k = 0; k < kBlockK / 16; k++) {
              if (k >= nk) break;
              const uint64_t ad = smem_desc(a_addr + m * a_sub + k * 32, hi), bd = smem_desc(b_addr + k * 32, hi);
              if (NCTA == 2) umma_f16_pair(d_tmem + m * P.block_n, ad, bd, P.idesc, (i | k) != 0 ? 1u : 0u);
              else if (!(P.dbg & 8) || (i | k) == 0) umma_f16(d_tmem + m * P.block_n, ad, bd, P.idesc, (i | k) != 0 ? 1u : 0u);
            }
          }
          // frees the smem slot (in both CTAs of a pair) when these MMAs retire
          if (NCTA == 2) umma_commit_pair(empty_bar(s)); else umma_commit(empty_bar(s));
        }
        if (c.iters > 0) {                   // accumulator complete
          if (NCTA == 2) umma_commit_pair(tfull_bar(acc)); else umma_commit(tfull_bar(acc));
        } else {
          mbar_arrive(tfull_bar(acc));
          if (NCTA == 2) mbar_arrive_cluster(mapa_u32(tfull_bar(acc), 1));
        }
      }
    }
  } else {
    // ===== epilogue =====
    const int grp = warp & 3;                    // TMEM sub-partition of this warp
    const int row = grp * 32 + lane;             // row of the 128-row sub-tile
    const int et = threadIdx.x - 64;             // 0..127
    const uint32_t lane_off = (uint32_t)(grp * 32) << 16;
    uint32_t tcount = 0, slab_count = 0;
    // fused statistics, in shared memory behind the barriers.  kind 2 (per sample): the CTA's sums [2][stat_len] and a
    // [128][2] scratch.  kind 1 (per channel): the sums of the CTA's CURRENT N tile [2][256] and a [4 warps][64][2]
    // scratch (the 256 x 256 tiles leave < 6 KB of shared memory); they are added to the CTA's row of the partial
    // buffer in global memory whenever the N tile changes (never, when the grid is a multiple of tiles_n).
    float* sacc = reinterpret_cast<float*>(sgen + stages * stage_bytes + kStagingBytes + 256);
    float* sscr = sacc + (P.stat_kind == 1 ? 512 : ((2 * P.stat_len + 3) & ~3));
    float* gpart = P.stat_part + (size_t)blockIdx.x * 2 * P.stat_len;
    float ks = 0.f, kq = 0.f;                    // swap mode: this thread's channel, over every tile of the CTA
    int cur_nt = -1;
    auto flush_nt = [&]() {                      // all 128 epilogue threads; entry e is always handled by thread e % 128
      named_bar_sync(2, kEpiThreads);
      if (cur_nt >= 0)
        for (int e = et; e < 2 * P.block_n; e += kEpiThreads) {
          const int which = e >= P.block_n ? 1 : 0, col = cur_nt * P.block_n + (e - which * P.block_n);
          if (col < P.nout) gpart[which * P.stat_len + (P.bias_mod < P.nout ? col % P.bias_mod : col)] += sacc[e];
        }
      for (int e = et; e < 512; e += kEpiThreads) sacc[e] = 0.f;
      named_bar_sync(2, kEpiThreads);
    };
    if (STATS && !P.swap) {
      if (P.stat_kind == 1) {
        for (int i = et; i < 2 * P.stat_len; i += kEpiThreads) gpart[i] = 0.f;
      } else {
        for (int i = et; i < 2 * P.stat_len; i += kEpiThreads) sacc[i] = 0.f;
      }
      named_bar_sync(1, kEpiThreads);
    }
    // A failed (timed-out) wait only raises g_tc_error: the loop is never left early, so the
    // named barriers of the 128 epilogue threads stay matched and the kernel drains.
    for (int tile = first_tile; tile < P.num_tiles; tile += tile_step, tcount++) {
      const TileCoord c = decode_tile<NCTA>(P, tile, rank);
      const uint32_t acc = P.nacc == 2 ? (tcount & 1u) : 0u;
      const uint32_t aph = (P.nacc == 2 ? (tcount >> 1) : tcount) & 1u;
      if (NCTA == 2) mbar_wait_cluster(tfull_bar(acc), aph); else mbar_wait(tfull_bar(acc), aph);
      tc_fence_after();
      const CUtensorMap* om = &P.out_maps[c.si];
      const int nslab = (P.block_n + 63) >> 6;
      if (P.swap) {
        // TMEM lane = output channel, column = voxel: transpose through the two staging slabs
        // (slab s = channels [64s, 64s+64) of 128 voxel rows, 128-B swizzled like every TMA tile)
        const int cch = grp * 32 + lane;
        const int chan = c.nt * 128 + cch;
        const int nch = P.nout - c.nt * 128 < 128 ? P.nout - c.nt * 128 : 128;
        const int ns = (nch + 63) >> 6;
        const bool warp_live = grp * 32 < ns * 64;
        float bv = 0.f;
        if (P.bias && chan < P.nout) bv = __ldg(P.bias + chan % P.bias_mod);
        const uint32_t sdst = staging + (uint32_t)(cch >> 6) * kABytes + (uint32_t)(cch & 7) * 2u;
        const uint32_t chunk = (uint32_t)(cch & 63) >> 3;
        for (int m = 0; m < P.mt; m++) {
          int q[4] = {c.q[0], c.q[1], c.q[2], c.q[3]};
          q[P.split_dim] += m * P.sub_ext;
          const RowClip clip = make_clip(P, c, q);
          if (et == 0) tma_store_wait_read<0>();
          named_bar_sync(1, kEpiThreads);
          if (warp_live) {
            float ts = 0.f, tq = 0.f;              // two-level sum: 128 voxels here, then the CTA's running sums
            for (int h = 0; h < 2; h++) {
              uint32_t v[64];
              const uint32_t tcol = tmem_base + lane_off + acc * acc_cols + m * 128 + h * 64;
              if (c.iters > 0) {
#pragma unroll
                for (int g = 0; g < 4; g++) tmem_ld16(tcol + g * 16, v + g * 16);
                tmem_ld_wait();
              } else {
#pragma unroll
                for (int i = 0; i < 64; i++) v[i] = 0u;
              }
#pragma unroll
              for (int i = 0; i < 64; i++) v[i] = __float_as_uint(__uint_as_float(v[i]) + bv);
              apply_act64(v, P.act, P.slope);
#pragma unroll
              for (int j = 0; j < 64; j++) {
                const __nv_bfloat16 hv = __float2bfloat16_rn(__uint_as_float(v[j]));
                st_shared_b16(sdst + (uint32_t)(h * 64 + j) * 128u + ((chunk ^ (uint32_t)(j & 7)) << 4),
                              *reinterpret_cast<const uint16_t*>(&hv));
                if (STATS && P.stat_kind == 1 && clip.ok(h * 64 + j)) {   // column = voxel h*64 + j of sub-tile m
                  const float r = __bfloat162float(hv);
                  ts += r; tq += r * r;
                }
              }
            }
            ks += ts; kq += tq;
          }
          fence_proxy_async();
          named_bar_sync(1, kEpiThreads);
          if (et == 0) {
            for (int sl = 0; sl < ns; sl++)
              tma_store_5d(om, staging + sl * kABytes, c.nt * 128 + sl * 64, q[3], q[2], q[1], q[0]);
            tma_store_commit();
          }
        }
      } else {
      if (STATS && P.stat_kind == 1 && c.nt != cur_nt) { flush_nt(); cur_nt = c.nt; }
      for (int m = 0; m < P.mt; m++) {
        int q[4] = {c.q[0], c.q[1], c.q[2], c.q[3]};
        q[P.split_dim] += m * P.sub_ext;
        const RowClip clip = make_clip(P, c, q);
        float rs = 0.f, rq = 0.f;                // kind 2: this thread's row (voxel), over the channels of the tile
        for (int j = 0; j < nslab; j++, slab_count++) {
          const uint32_t sbuf = staging + (slab_count & 1u) * kABytes;
          // the TMA store that last read this slab buffer (2 slabs ago) must be done with it
          if (et == 0) tma_store_wait_read<1>();
          named_bar_sync(1, kEpiThreads);
          const uint32_t tcol = tmem_base + lane_off + acc * acc_cols + m * P.block_n + j * 64;
          uint32_t v[64];
          const int ncol = P.block_n - j * 64 < 64 ? P.block_n - j * 64 : 64;   // multiple of 16
          if (c.iters > 0) {
#pragma unroll
            for (int g = 0; g < 4; g++)
              if (g * 16 < ncol) tmem_ld16(tcol + g * 16, v + g * 16);
            tmem_ld_wait();
          } else {
#pragma unroll
            for (int i = 0; i < 64; i++) v[i] = 0u;
          }
          const int col0 = c.nt * P.block_n + j * 64;
          if (P.bias) {
            if (P.bias_mod >= P.nout) {            // plain per-channel bias: no index arithmetic
              const float* bp = P.bias + col0;
              const int nb = P.nout - col0;          // valid columns of this slab (warp-uniform)
              if (nb >= 64 && (reinterpret_cast<uintptr_t>(bp) & 15) == 0) {
#pragma unroll
                for (int q = 0; q < 16; q++) {
                  const float4 b4 = __ldg(reinterpret_cast<const float4*>(bp) + q);
                  v[4 * q] = __float_as_uint(__uint_as_float(v[4 * q]) + b4.x);
                  v[4 * q + 1] = __float_as_uint(__uint_as_float(v[4 * q + 1]) + b4.y);
                  v[4 * q + 2] = __float_as_uint(__uint_as_float(v[4 * q + 2]) + b4.z);
                  v[4 * q + 3] = __float_as_uint(__uint_as_float(v[4 * q + 3]) + b4.w);
                }
              } else {
#pragma unroll
                for (int i = 0; i < 64; i++)
                  if (i < nb && i < ncol) v[i] = __float_as_uint(__uint_as_float(v[i]) + __ldg(bp + i));
              }
            } else {                               // point convT: columns are (tap, co), bias_mod = Cout >= 8
              const int b0 = col0 % P.bias_mod;    // one modulo per slab
#pragma unroll
              for (int i = 0; i < 64; i++) {
                if (i < ncol) {
                  int bi = b0 + i;
                  if (P.bias_mod >= 64) { if (bi >= P.bias_mod) bi -= P.bias_mod; } else bi %= P.bias_mod;
                  v[i] = __float_as_uint(__uint_as_float(v[i]) + __ldg(P.bias + bi));
                }
              }
            }
          }
          apply_act64(v, P.act, P.slope);
#pragma unroll
          for (int ch = 0; ch < 8; ch++) {
            if (ch * 8 < ncol) {
              const uint32_t p0 = pack_bf16x2(__uint_as_float(v[ch * 8 + 0]), __uint_as_float(v[ch * 8 + 1])),
                             p1 = pack_bf16x2(__uint_as_float(v[ch * 8 + 2]), __uint_as_float(v[ch * 8 + 3])),
                             p2 = pack_bf16x2(__uint_as_float(v[ch * 8 + 4]), __uint_as_float(v[ch * 8 + 5])),
                             p3 = pack_bf16x2(__uint_as_float(v[ch * 8 + 6]), __uint_as_float(v[ch * 8 + 7]));
              st_shared_v4(sbuf + row * 128 + ((ch ^ (row & 7)) << 4), p0, p1, p2, p3);
              if (STATS && P.stat_kind == 2 && col0 + ch * 8 < P.nout) {      // the stored (rounded) values; nout % 8 == 0
                const uint32_t pk[4] = {p0, p1, p2, p3};
#pragma unroll
                for (int i = 0; i < 4; i++) {
                  const float lo = __uint_as_float(pk[i] << 16), hi = __uint_as_float(pk[i] & 0xffff0000u);
                  rs += lo + hi; rq += lo * lo + hi * hi;
                }
              }
            }
          }
          fence_proxy_async();                   // make the generic-proxy smem writes visible to TMA
          named_bar_sync(1, kEpiThreads);
          if (et == 0) {
            tma_store_5d(om, sbuf, col0, q[3], q[2], q[1], q[0]);
            tma_store_commit();
          }
          if (STATS && P.stat_kind == 1) {
            // per-channel sums of the slab just staged (its bf16 values, as stored).  A warp owns 32 rows: lane =
            // (8 rows, 8 channels); the four row groups of a warp are combined with shuffles, the four warps through
            // the scratch in a fixed order, then ONE thread adds to the sums of this N tile.
            const int ch8 = lane & 7, rg = (et >> 5) * 4 + (lane >> 3);
            float s8[8], q8[8];
#pragma unroll
            for (int k = 0; k < 8; k++) s8[k] = q8[k] = 0.f;
            if (ch8 * 8 < ncol) {
#pragma unroll
              for (int i = 0; i < 8; i++) {
                const int r = rg * 8 + i;
                if (clip.ok(r)) {
                  const uint4 raw = ld_shared_v4(sbuf + r * 128 + ((ch8 ^ (r & 7)) << 4));
                  const uint32_t pk[4] = {raw.x, raw.y, raw.z, raw.w};
#pragma unroll
                  for (int k = 0; k < 4; k++) {
                    const float lo = __uint_as_float(pk[k] << 16), hi = __uint_as_float(pk[k] & 0xffff0000u);
                    s8[2 * k] += lo; q8[2 * k] += lo * lo; s8[2 * k + 1] += hi; q8[2 * k + 1] += hi * hi;
                  }
                }
              }
            }
#pragma unroll
            for (int k = 0; k < 8; k++) {
              s8[k] += __shfl_xor_sync(0xffffffffu, s8[k], 8);  q8[k] += __shfl_xor_sync(0xffffffffu, q8[k], 8);
              s8[k] += __shfl_xor_sync(0xffffffffu, s8[k], 16); q8[k] += __shfl_xor_sync(0xffffffffu, q8[k], 16);
            }
            if (lane < 8) {
              float* sc = sscr + ((et >> 5) * 64 + ch8 * 8) * 2;
#pragma unroll
              for (int k = 0; k < 8; k++) { sc[2 * k] = s8[k]; sc[2 * k + 1] = q8[k]; }
            }
            named_bar_sync(2, kEpiThreads);
            const int cc = et & 63, which = et >> 6;
            const float a = sscr[cc * 2 + which] + sscr[(64 + cc) * 2 + which] + sscr[(128 + cc) * 2 + which] +
                            sscr[(192 + cc) * 2 + which];
            if (cc < ncol) sacc[which * P.block_n + j * 64 + cc] += a;
          }
        }
        if (STATS && P.stat_kind == 2) {
          // per-sample sums: rows of one sample are contiguous in the box (n is its slowest dim).  With >= 32 rows per
          // sample a warp's rows share the sample: shuffle reduction, then thread 0 adds the four warp sums in order;
          // smaller samples (4^3 outputs ...) go through the per-row scratch.
          const bool rv = clip.ok(row);
          const int rl = P.sb_log[0] + P.sb_log[1] + P.sb_log[2];       // log2(rows per sample in the sub-tile)
          float ws_ = rv ? rs : 0.f, wq_ = rv ? rq : 0.f;
          if (rl >= 5) {
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) { ws_ += __shfl_xor_sync(0xffffffffu, ws_, o); wq_ += __shfl_xor_sync(0xffffffffu, wq_, o); }
            if (lane == 0) { sscr[grp * 2] = ws_; sscr[grp * 2 + 1] = wq_; }
            named_bar_sync(2, kEpiThreads);
            if (et == 0) {
#pragma unroll
              for (int g2 = 0; g2 < 4; g2++) {
                const int samp = (g2 * 32) >> rl;
                if (samp < clip.lim[0]) {
                  sacc[q[0] + samp] += sscr[g2 * 2];
                  sacc[P.stat_len + q[0] + samp] += sscr[g2 * 2 + 1];
                }
              }
            }
          } else {
            sscr[row * 2] = ws_;
            sscr[row * 2 + 1] = wq_;
            named_bar_sync(2, kEpiThreads);
            const int samp = et >> 1, which = et & 1;
            if (samp < (kBlockM >> rl) && samp < clip.lim[0]) {
              float a = 0.f;
              for (int r = samp << rl; r < (samp + 1) << rl; r++) a += sscr[r * 2 + which];
              sacc[which * P.stat_len + q[0] + samp] += a;
            }
          }
        }
      }
      }
      // all TMEM reads of this accumulator set are complete (tmem_ld_wait above)
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if (NCTA == 2) mbar_arrive_cluster(mapa_u32(tempty_bar(acc), 0));   // the leader's barrier counts both CTAs
        else mbar_arrive(tempty_bar(acc));
      }
    }
    if (STATS) {
      if (P.swap) {
        const int cch = grp * 32 + lane;          // swap mode has one N tile: channel = TMEM lane
        if (cch < P.nout) { gpart[cch] = ks; gpart[P.stat_len + cch] = kq; }
      } else if (P.stat_kind == 1) {
        flush_nt();
      } else {
        named_bar_sync(1, kEpiThreads);
        for (int i = et; i < 2 * P.stat_len; i += kEpiThreads) gpart[i] = sacc[i];
      }
    }
    if (et == 0) tma_store_wait<0>();
  }

  // ---- teardown --------------------------------------------------------------
  tc_fence_before();
  if (NCTA == 2) cluster_sync_all();      // the peer may still read this CTA's smem / signal its barriers
  else __syncthreads();
  if (warp == 1) {
    if (NCTA == 2) tmem_dealloc_pair(tmem_base, P.tmem_cols);
    else tmem_dealloc(tmem_base, P.tmem_cols);
  }
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
  int32_t a_ch, b_ch, a_tiles, b_tiles, block_n;   // M side (128-row tiles) / N side (block_n columns)
  int32_t m_big;                     // 1: M side = big tensor (tap-shifted), N side = small tensor
  int32_t tpack;                     // > 1: narrow big tensor (<= 64 ch): N = tpack taps x one 64-column slab each
  int32_t ngroups;                   // tap groups = ceil(ntaps / tpack), or M-unit groups = ceil(ntaps * a_tiles / mt)
  int32_t mt;                        // M tiles (tap, a_tile) per CTA sharing one N-side tile: 1 or 2 accumulators
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
  const uint32_t a_tile_bytes = 2 * kSlabBytes;
  const uint32_t a_bytes = P.mt * a_tile_bytes, b_bytes = b_slabs * kSlabBytes;
  const uint32_t stage_bytes = a_bytes + b_bytes;
  const uint32_t bar0 = sbase + stages * stage_bytes;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(sgen + stages * stage_bytes + (2 * stages + 1) * 8);
  auto full_bar = [&](int s) { return bar0 + 8u * s; };
  auto empty_bar = [&](int s) { return bar0 + 8u * (stages + s); };
  const uint32_t acc_bar = bar0 + 8u * (2 * stages);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  int id = blockIdx.x;
  const int bt = id % P.b_tiles; id /= P.b_tiles;
  int at = 0, ti, nmt = 1;                   // first M unit of this CTA and how many it owns
  int at2 = 0, ti2 = 0;                      // second M unit (mt == 2)
  if (P.mt > 1) {
    const int gi = id % P.ngroups, u = gi * 2, total = P.ntaps * P.a_tiles;
    ti = u / P.a_tiles; at = u - ti * P.a_tiles;
    if (u + 1 < total) { nmt = 2; ti2 = (u + 1) / P.a_tiles; at2 = (u + 1) - ti2 * P.a_tiles; }
  } else {
    at = id % P.a_tiles; id /= P.a_tiles;
    ti = id % P.ngroups;                     // tap (tpack == 1) or tap group
  }
  const int sp = id / P.ngroups;
  const TapDesc tp = P.taps[ti * P.tpack];
  const TapDesc tp2 = P.taps[ti2];
  const uint32_t tx_bytes = stage_bytes - (uint32_t)(P.mt - nmt) * a_tile_bytes;
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
        mbar_arrive_expect_tx(full_bar(s), tx_bytes);
        const uint32_t a_dst = sbase + s * stage_bytes;
        // the big tensor is read at the tap-shifted voxel box, the small one at the box itself
        const CUtensorMap* mm = P.m_big ? bm : &P.s_map;
        const CUtensorMap* nm = P.m_big ? &P.s_map : bm;
        const int mw = P.m_big ? tp.off_w : 0, mh = P.m_big ? tp.off_h : 0, md = P.m_big ? tp.off_d : 0;
        const int nw = tp.off_w - mw, nh = tp.off_h - mh, nd = tp.off_d - md;
        for (int j = 0; j < 2; j++)
          tma_load_5d(a_dst + j * kSlabBytes, mm, full_bar(s), at * 128 + j * 64, q0w + mw, q0h + mh, q0d + md, q0n);
        if (nmt > 1) {
          // second M tile: its own tap shift when the big tensor sits on M, else the next 128 small channels
          const CUtensorMap* mm2 = P.m_big ? &P.b_maps[tp2.map] : &P.s_map;
          const int w2 = P.m_big ? tp2.off_w : 0, h2 = P.m_big ? tp2.off_h : 0, d2 = P.m_big ? tp2.off_d : 0;
          for (int j = 0; j < 2; j++)
            tma_load_5d(a_dst + a_tile_bytes + j * kSlabBytes, mm2, full_bar(s), at2 * 128 + j * 64, q0w + w2, q0h + h2,
                        q0d + d2, q0n);
        }
        if (P.tpack > 1) {
          // slab j = tap (group * tpack + j) of the <= 64-channel big tensor (a short last group
          // re-reads its last tap; the epilogue drops those columns)
          for (int j = 0; j < b_slabs; j++) {
            const int tj = ti * P.tpack + j < P.ntaps ? ti * P.tpack + j : P.ntaps - 1;
            const TapDesc tq = P.taps[tj];
            tma_load_5d(a_dst + a_bytes + j * kSlabBytes, &P.b_maps[tq.map], full_bar(s), 0, q0w + tq.off_w,
                        q0h + tq.off_h, q0d + tq.off_d, q0n);
          }
        } else
        for (int j = 0; j < b_slabs; j++)
          tma_load_5d(a_dst + a_bytes + j * kSlabBytes, nm, full_bar(s), bt * P.block_n + j * 64, q0w + nw, q0h + nh,
                      q0d + nd, q0n);
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
        if (nmt > 1) {
#pragma unroll
          for (int k = 0; k < kWgVox / 16; k++)
            umma_f16(tmem_base + (uint32_t)P.block_n, smem_desc(a_addr + a_tile_bytes + k * 2048, hi),
                     smem_desc(b_addr + k * 2048, hi), P.idesc, (it | k) != 0 ? 1u : 0u);
        }
        umma_commit(empty_bar(s));
      }
      if (iters > 0) umma_commit(acc_bar);
      else mbar_arrive(acc_bar);
    }
  } else {
    const int grp = warp & 3;
    mbar_wait(acc_bar, 0);
    tc_fence_after();
    for (int m = 0; m < nmt; m++) {
    const int a = (m ? at2 : at) * 128 + grp * 32 + lane;
    const bool valid = a < P.a_ch;
    float* orow = P.ws + (((int64_t)sp * P.ntaps + (m ? ti2 : ti)) * P.a_ch + a) * P.b_ch;
    const uint32_t trow = tmem_base + ((uint32_t)(grp * 32) << 16) + (uint32_t)(m * P.block_n);
    for (int c0 = 0; c0 < P.block_n; c0 += 16) {
      uint32_t v[16];
      if (iters > 0) { tmem_ld16(trow + c0, v); tmem_ld_wait(); }
      else {
#pragma unroll
        for (int j = 0; j < 16; j++) v[j] = 0u;
      }
      if (P.tpack > 1) {
        const int tap = ti * P.tpack + (c0 >> 6), col = c0 & 63;
        if (valid && tap < P.ntaps) {
          float* prow = P.ws + (((int64_t)sp * P.ntaps + tap) * P.a_ch + a) * P.b_ch;
#pragma unroll
          for (int j = 0; j < 16; j += 4)
            if (col + j < P.b_ch)
              *reinterpret_cast<uint4*>(prow + col + j) = make_uint4(v[j], v[j + 1], v[j + 2], v[j + 3]);
        }
        continue;
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
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, P.tmem_cols);
}

// dw[(a*B + b)*T + t] = sum_split ws[split][tapidx(t)][m][n]   (0 when the tap only sees padding)
// where (m, n) = (a, b), or (b, a) when the GEMM ran with the big tensor on the M side.
// A block owns AB consecutive (m,n) pairs (plane % 64 == 0: both channel counts are multiples of 8):
// the partials are read as float4 along (m,n) - a warp reads 512 contiguous bytes per tap, every load of
// the thread is issued before the first use when there is one split - staged in shared memory as
// [t][AB + 1] (conflict-free float4 rows in, odd-pitch columns out) and written back as runs of T
// contiguous floats (t fastest).  The first version moved 4-byte loads, 32 pairs per block, and ran the
// 1 GB reduce of the 1024 <-> 2048 layers at 1.4 TB/s.
struct TapIndex { int16_t idx[kMaxTaps]; };
template <int AB>
__global__ void __launch_bounds__(256)
wgrad_reduce_kernel(const float* __restrict__ ws, float* __restrict__ dw, int64_t plane, int T, int ntaps, int splits,
                    int m_big, int n_ch, int big_ch, TapIndex ti) {
  extern __shared__ float red_sm[];            // [T][AB + 1]
  constexpr int Q = AB / 4;
  const int64_t mn0 = (int64_t)blockIdx.x * AB;
  const int nmn = plane - mn0 < AB ? (int)(plane - mn0) : AB;
  const int total = Q * T;
  for (int i0 = threadIdx.x; i0 < total; i0 += 8 * 256) {
    float4 acc[8];
#pragma unroll
    for (int u = 0; u < 8; u++) {
      const int i = i0 + u * 256;
      acc[u] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (i < total) {
        const int t = i / Q, a = (i - t * Q) * 4;
        const int idx = ti.idx[t];
        if (idx >= 0 && a < nmn)
          acc[u] = __ldg(reinterpret_cast<const float4*>(ws + (int64_t)idx * plane + mn0 + a));
      }
    }
    for (int s = 1; s < splits; s++) {
#pragma unroll
      for (int u = 0; u < 8; u++) {
        const int i = i0 + u * 256;
        if (i < total) {
          const int t = i / Q, a = (i - t * Q) * 4;
          const int idx = ti.idx[t];
          if (idx >= 0 && a < nmn) {
            const float4 v = __ldg(reinterpret_cast<const float4*>(ws + ((int64_t)s * ntaps + idx) * plane + mn0 + a));
            acc[u].x += v.x; acc[u].y += v.y; acc[u].z += v.z; acc[u].w += v.w;
          }
        }
      }
    }
#pragma unroll
    for (int u = 0; u < 8; u++) {
      const int i = i0 + u * 256;
      if (i < total) {
        const int t = i / Q, a = (i - t * Q) * 4;
        float* d = red_sm + t * (AB + 1) + a;
        d[0] = acc[u].x; d[1] = acc[u].y; d[2] = acc[u].z; d[3] = acc[u].w;
      }
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < nmn * T; i += blockDim.x) {
    const int q = i / T, t = i - q * T;
    int64_t ab = mn0 + q;                      // (a*B + b) when M = small
    if (m_big) { const int64_t m = ab / n_ch, n = ab - m * n_ch; ab = n * big_ch + m; }   // m = b, n = a
    dw[ab * T + t] = red_sm[t * (AB + 1) + q];
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

// bf16 view `m` of `base`, box = {64 (or 32) channels, box[1..rank-1]}, swizzle span = the box row
// (128 B, or 64 B for 32-channel boxes), zero OOB fill.
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
                  CU_TENSOR_MAP_INTERLEAVE_NONE, box[0] == 32 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_128B,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? B2C_OK : B2C_ERR_CUDA;
}

static uint32_t tmem_cols_for(int n) { uint32_t c = 32; while ((int)c < n) c <<= 1; return c; }
constexpr int kSmemBudget = 227 * 1024;

size_t tc_conv_stats_bytes(int len) { return (size_t)num_sms() * 2 * (size_t)len * sizeof(float); }

bool tc_conv_stats_supported(const ConvPlan& p, int kind, int bias_mod) {
  // point GEMM (columns = (tap, channel)): a channel must map to ONE accumulator slot of the N tile
  if (kind == 1) return bias_mod == 0 || (!p.swap && bias_mod >= p.block_n && bias_mod % p.block_n == 0);
  if (kind == 2) return !p.swap && bias_mod == 0 && p.subs[0].ext[0] <= 1024;
  return false;
}

int tc_conv_launch(const ConvPlan& p, const void* in, const void* w, void* out, const float* bias, int bias_mod,
                   int act, float slope, cudaStream_t st, const TcStats* stats) {
  TcConvParams P;
  memset(&P, 0, sizeof(P));
  size_t stat_smem = 0;
  if (stats && stats->kind) {
    if (!tc_conv_stats_supported(p, stats->kind, bias_mod) || !stats->part || stats->len <= 0 ||
        stats->part_bytes < tc_conv_stats_bytes(stats->len))
      return B2C_ERR_INVALID;
    P.stat_kind = stats->kind; P.stat_len = stats->len; P.stat_part = stats->part;
    stat_smem = p.swap ? 0 : stats->kind == 1 ? 4096 : (((size_t)2 * stats->len + 3) & ~(size_t)3) * sizeof(float) + 1024;
  }
  const int abox[5] = {p.kb, p.box[3], p.box[2], p.box[1], p.box[0]};
  for (int i = 0; i < p.nmaps; i++) {
    int rc = encode_map(&P.in_maps[i], in, p.maps[i], 5, abox);
    if (rc) return rc;
  }
  // output lattices (one per sub-problem): box = one 128-row sub-tile x 64 channels
  int sbox[4] = {p.box[0], p.box[1], p.box[2], p.box[3]};
  sbox[p.split_dim] = p.sub_ext;
  const int obox[5] = {kBlockK, sbox[3], sbox[2], sbox[1], sbox[0]};
  for (int s = 0; s < p.nsubs; s++) {
    const SubDesc& sd = p.subs[s];
    MapDesc om;
    om.base = sd.obase;
    om.dims[0] = p.nout; om.dims[1] = sd.ext[3]; om.dims[2] = sd.ext[2]; om.dims[3] = sd.ext[1]; om.dims[4] = sd.ext[0];
    om.strides[0] = 1; om.strides[1] = sd.ostr[3]; om.strides[2] = sd.ostr[2]; om.strides[3] = sd.ostr[1];
    om.strides[4] = sd.ostr[0];
    int rc = encode_map(&P.out_maps[s], out, om, 5, obox);
    if (rc) return rc;
  }
  // CTA pairs (cta_group::2): two M tiles of one sub-problem share the filter tile, each CTA
  // staging half of it -> 1.5x fewer L2->SM bytes per FLOP and deeper pipelines.
  int uniform = 1;
  for (int s2 = 1; s2 < p.nsubs; s2++)
    for (int i = 0; i < 4; i++) if (p.subs[s2].tiles[i] != p.subs[0].tiles[i]) uniform = 0;
  const int64_t sub_tiles = (int64_t)p.subs[0].tiles[0] * p.subs[0].tiles[1] * p.subs[0].tiles[2] * p.subs[0].tiles[3];
  // measured (profiles/r02_exp_pair_tiles.log, isolated launches): D.L2 fwd 1375 -> 1485 TFLOP/s, D.L4 fwd / dgrad
  // 1311 / 1411 -> 1457 / 1536, G.L5 dgrad 1375 -> 1465, G.L2 fwd 1246 -> 1395: on by default (B2C_NO_PAIR=1: off)
  const bool no_pair = knobs().no_pair;
  const bool pair = !no_pair && !p.swap && uniform && sub_tiles >= 2 && p.block_n % 32 == 0;
  const int ncta = pair ? 2 : 1;
  MapDesc wm;
  wm.base = 0;
  const int T = [&] { int mx = 0; for (int i = 0; i < p.ntaps_total; i++) if (p.taps[i].wtap > mx) mx = p.taps[i].wtap; return mx + 1; }();
  wm.dims[0] = p.kin; wm.dims[1] = p.nout; wm.dims[2] = T;
  wm.strides[0] = 1; wm.strides[1] = p.kin; wm.strides[2] = (int64_t)p.kin * p.nout;
  const int w_rows = p.swap ? 128 : p.block_n;
  const int wbox[3] = {p.kb, w_rows / ncta, 1};
  int rc = encode_map(&P.w_map, w, wm, 3, wbox);
  if (rc) return rc;
  memcpy(P.subs, p.subs, sizeof(P.subs));
  memcpy(P.taps, p.taps, sizeof(P.taps));
  P.nsubs = p.nsubs;
  for (int i = 0; i < 4; i++) P.box[i] = p.box[i];
  P.mt = p.mt; P.split_dim = p.split_dim; P.sub_ext = p.sub_ext;
  P.nout = p.nout; P.block_n = p.block_n; P.tiles_n = p.tiles_n; P.kchunks = p.kchunks;
  const uint32_t stage_bytes = ((uint32_t)p.mt * kBlockM + (uint32_t)(w_rows / ncta)) * (uint32_t)p.kb * 2u;
  P.swap = p.swap; P.w_rows = w_rows; P.kb = p.kb;
  const int fixed = 1024 /*alignment slack*/ + (int)kStagingBytes + 256 /*barriers + tmem slot*/ + (int)stat_smem;
  {
    int sbox2[4] = {p.box[0], p.box[1], p.box[2], p.box[3]};
    sbox2[p.split_dim] = p.sub_ext;
    auto lg = [](int v) { int l = 0; while ((1 << l) < v) l++; return l; };
    P.sb_log[0] = lg(sbox2[3]); P.sb_log[1] = lg(sbox2[2]); P.sb_log[2] = lg(sbox2[1]);
  }
  int stages = (kSmemBudget - fixed) / (int)stage_bytes;
  if (stages > 8) stages = 8;
  P.dbg = knobs().dbg;                                         // timing experiments only (8: no MMAs, 16: no loads)
  if (knobs().stages >= 2 && knobs().stages < stages) stages = knobs().stages;                               // experiments
  if (stages < 2) return B2C_ERR_UNSUPPORTED;
  P.stages = stages;
  P.idesc = p.swap ? idesc_bf16(128, 128 * p.mt, 0, 0) : idesc_bf16(kBlockM * ncta, p.block_n, 0, 0);
  P.nacc = p.mt * p.block_n > 256 ? 1 : 2;
  P.tmem_cols = tmem_cols_for(P.nacc * p.mt * p.block_n);
  if (P.tmem_cols > 512) return B2C_ERR_UNSUPPORTED;
  P.bias = bias; P.bias_mod = bias_mod > 0 ? bias_mod : p.nout; P.act = act; P.slope = slope;
  const size_t smem = (size_t)stages * stage_bytes + fixed;
  const int64_t tiles = pair ? (int64_t)p.nsubs * ((sub_tiles + 1) / 2) * p.tiles_n : (int64_t)p.m_tiles * p.tiles_n;
  if (tiles <= 0) return B2C_OK;
  if (tiles > 0x7fffffffLL) return B2C_ERR_UNSUPPORTED;
  P.num_tiles = (int32_t)tiles;
  P.interleave = p.nsubs > 1 && uniform;
  const int sms = num_sms();
  // the attribute is per device; set it on every call (cheap) rather than caching per device
  if (pair) {
    auto kern2 = P.stat_kind ? tc_conv_kernel<2, true> : tc_conv_kernel<2, false>;
    if (cudaFuncSetAttribute(kern2, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBudget) != cudaSuccess)
      return B2C_ERR_CUDA;
    const int64_t clusters = tiles < sms / 2 ? tiles : sms / 2;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(2 * clusters));
    cfg.blockDim = dim3(kThreadsTc);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    if (cudaLaunchKernelEx(&cfg, kern2, P) != cudaSuccess) { cudaGetLastError(); return B2C_ERR_CUDA; }
    if (stats) stats->nblocks = (int)(2 * clusters);
    return launched();
  }
  auto kern1 = P.stat_kind ? tc_conv_kernel<1, true> : tc_conv_kernel<1, false>;
  if (cudaFuncSetAttribute(kern1, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBudget) != cudaSuccess)
    return B2C_ERR_CUDA;
  const unsigned grid = (unsigned)(tiles < sms ? tiles : sms);
  kern1<<<grid, kThreadsTc, smem, st>>>(P);
  if (stats) stats->nblocks = (int)grid;
  return launched();
}

// ---- wgrad ---------------------------------------------------------------------
struct WgradCfg { ConvPlan plan; int box[4], tiles[4]; int ktiles, splits, kps, a_tiles, b_tiles, block_n, m_big, m_ch, n_ch, tpack, ngroups, mt; };

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
  // Which tensor's channels sit on N (<= 256 columns per tile)?  The wider N tile wins: a
  // 128 x 256 tile moves 2/3 of the L2 bytes per FLOP of a 128 x 128 one.
  auto ru64 = [](int v) { int r = (v + 63) / 64 * 64; return r > 256 ? 256 : r; };
  c->m_big = ru64(g.sc) > ru64(g.bc) ? 1 : 0;
  c->m_ch = c->m_big ? g.bc : g.sc;
  c->n_ch = c->m_big ? g.sc : g.bc;
  int bn = ru64(c->n_ch);
  c->tpack = 1;
  c->ngroups = c->plan.ntaps_total;
  if (g.bc <= 64 && c->plan.ntaps_total >= 2) {
    // narrow big tensor (utils.py hidden 32 / 64-channel 3^3 layers): one tap would fill a single
    // 64-column slab, so up to four taps share a CTA - the small tensor's tile is read once for
    // all of them and N = 256 keeps the tensor pipe on its efficient shape
    c->m_big = 0; c->m_ch = g.sc; c->n_ch = g.bc;
    c->tpack = c->plan.ntaps_total >= 4 ? 4 : c->plan.ntaps_total;
    c->ngroups = (c->plan.ntaps_total + c->tpack - 1) / c->tpack;
    bn = 64 * c->tpack;
  }
  c->block_n = bn;
  c->a_tiles = (c->m_ch + 127) / 128;
  c->b_tiles = c->tpack > 1 ? 1 : (c->n_ch + bn - 1) / bn;
  // Two M tiles (two taps, or two 128-channel tiles of one tap) per CTA share the N-side tile: the
  // kernel is bound by L2 -> SM bytes (measured: loads alone take as long as loads + MMAs), and
  // (2 x 16 + 32) KB per two 128 x 256 x 64 stages is 2/3 of the bytes of (16 + 32) KB per one.
  // With the big tensor on M any two units pair up (the small N tile does not depend on the tap);
  // with the small tensor on M the pair must be two channel tiles of one tap.
  c->mt = 1;
  if (c->tpack == 1 && 2 * bn <= 512 && c->plan.ntaps_total * c->a_tiles >= 2 && (c->m_big || c->a_tiles % 2 == 0) &&
      !knobs().wg_mt1) {
    c->mt = 2;
    c->ngroups = (c->plan.ntaps_total * c->a_tiles + 1) / 2;
  }
  const int64_t units = (int64_t)c->ngroups * (c->mt > 1 ? 1 : c->a_tiles) * c->b_tiles;
  if (units <= 0) return B2C_ERR_UNSUPPORTED;
  // split-K count: 2.5 .. 10 waves of the SMs, at least 64 K-iterations per CTA (16 when the
  // problem has too few tiles to fill the machine otherwise: 1x1x1 filters, 32-channel layers);
  // pick the count whose last wave is fullest (ties -> fewer splits = less partial-sum traffic)
  const int sms = num_sms();
  int64_t splits = 1;
  {
    double best = -1.0;
    int64_t lo = cdiv((int64_t)sms * 5 / 2, units), hi = cdiv((int64_t)sms * 10, units);
    int64_t cap = c->ktiles / 64;
    if (cap < lo) cap = c->ktiles / 16 < lo ? c->ktiles / 16 : lo;
    if (cap < 1) cap = 1;
    if (lo > cap) lo = cap;
    if (hi > cap) hi = cap;
    for (int64_t sp = lo < 1 ? 1 : lo; sp <= hi; sp++) {
      const double waves = (double)(units * sp) / sms;
      const double eff = waves / (double)cdiv(units * sp, sms);
      if (eff > best + 0.02) { best = eff; splits = sp; }
    }
  }
  if (splits > c->ktiles) splits = c->ktiles;
  // bound the fp32 partial buffer to 1 GiB
  const int64_t plane = (int64_t)c->plan.ntaps_total * g.sc * g.bc * 4;
  while (splits > 1 && splits * plane > (1LL << 30)) splits--;
  if (splits < 1) splits = 1;
  c->kps = (int)cdiv(c->ktiles, splits);
  c->splits = (int)cdiv(c->ktiles, c->kps);
  if (units * c->splits > 0x7fffffffLL) return B2C_ERR_UNSUPPORTED;   // grid = units * splits
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
  P.a_ch = c.m_ch; P.b_ch = c.n_ch; P.a_tiles = c.a_tiles; P.b_tiles = c.b_tiles; P.block_n = c.block_n;
  P.m_big = c.m_big; P.tpack = c.tpack; P.ngroups = c.ngroups; P.mt = c.mt;
  const uint32_t stage_bytes = c.mt * 2 * kSlabBytes + (c.block_n / 64) * kSlabBytes;
  int stages = (kSmemBudget - 1024 - 256) / (int)stage_bytes;
  if (stages > 8) stages = 8;
  P.stages = stages;
  P.idesc = idesc_bf16(128, c.block_n, 1, 1);
  P.tmem_cols = tmem_cols_for(c.mt * c.block_n);
  P.ws = static_cast<float*>(ws);
  if (cudaFuncSetAttribute(tc_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBudget) != cudaSuccess)
    return B2C_ERR_CUDA;
  const size_t smem = (size_t)stages * stage_bytes + 1024 + 256;
  const int64_t grid = (int64_t)c.ngroups * (c.mt > 1 ? 1 : c.a_tiles) * c.b_tiles * c.splits;
  tc_wgrad_kernel<<<(unsigned)grid, kThreadsTc, smem, st>>>(P);
  rc = launched();
  if (rc) return rc;
  TapIndex ti;
  const int T = g.taps();
  for (int t = 0; t < kMaxTaps; t++) ti.idx[t] = -1;
  for (int i = 0; i < c.plan.ntaps_total; i++) ti.idx[c.plan.taps[i].wtap] = (int16_t)i;
  const int64_t ab = (int64_t)g.sc * g.bc;
  if (ab % 4 != 0) return B2C_ERR_UNSUPPORTED;      // wgrad_cfg: both channel counts are multiples of 8
  if (T <= 64)
    wgrad_reduce_kernel<128><<<(unsigned)cdiv(ab, 128), 256, (size_t)T * 129 * sizeof(float), st>>>(
        P.ws, dw, ab, T, P.ntaps, c.splits, c.m_big, c.n_ch, g.bc, ti);
  else
    wgrad_reduce_kernel<64><<<(unsigned)cdiv(ab, 64), 256, (size_t)T * 65 * sizeof(float), st>>>(
        P.ws, dw, ab, T, P.ntaps, c.splits, c.m_big, c.n_ch, g.bc, ti);
  return launched();
}

#include "tc_thin.cuh"

bool tc_thin_gather_supported(const Geo& g) {
  if (g.bc != 1 || g.sc % 8 != 0 || g.sc < 16 || g.sc > 128 || g.taps() > 64) return false;
  const int64_t rows = (int64_t)g.n * g.sd * g.sh * g.sw;
  if ((int64_t)g.n * g.bd * g.bh * g.bw >= 0x7fffffffLL) return false;                  // 32-bit element offsets
  for (int i = 0; i < 3; i++)
    if ((g.k[i] - 1) * g.dl[i] > 127) return false;
  return rows > 0 && rows < 0x7fffffffLL - kBlockM;
}

// how the producer warps cut im2col rows out of the one-channel tensor
static int thin_src_fill(const Geo& g, const void* big, ThinSrc* S) {
  if (!tc_thin_gather_supported(g)) return B2C_ERR_UNSUPPORTED;
  if ((reinterpret_cast<uintptr_t>(big) & 3) != 0) return B2C_ERR_UNSUPPORTED;
  S->g = g;
  S->big = static_cast<const __nv_bfloat16*>(big);
  S->rows = (int64_t)g.n * g.sd * g.sh * g.sw;
  S->T = g.taps();
  bool k4s2 = g.bw % 2 == 0;
  for (int i = 0; i < 3; i++) k4s2 = k4s2 && g.k[i] == 4 && g.s[i] == 2 && g.p[i] == 1 && g.dl[i] == 1;
  k4s2 = k4s2 && g.bd == 2 * g.sd && g.bh == 2 * g.sh && g.bw == 2 * g.sw;
  S->fast_k4s2 = (k4s2 && g.sw % 32 == 0) ? 2 : 0;
  bool k3s1 = g.bd == g.sd && g.bh == g.sh && g.bw == g.sw && g.sw % 32 == 0;
  for (int i = 0; i < 3; i++) k3s1 = k3s1 && g.k[i] == 3 && g.s[i] == 1 && g.p[i] == 1 && g.dl[i] == 1;
  if (k3s1) S->fast_k4s2 = 3;           // thin_build_row: 3 = the k3 / s1 / p1 line loader
  for (int kd = 0, t = 0; kd < g.k[0]; kd++)
    for (int kh = 0; kh < g.k[1]; kh++)
      for (int kw = 0; kw < g.k[2]; kw++, t++) {
        const int dd = kd * g.dl[0], dh = kh * g.dl[1], dw = kw * g.dl[2];
        S->tap[t].d = (int8_t)dd; S->tap[t].h = (int8_t)dh; S->tap[t].w = (int8_t)dw;
        S->tap[t].off = (dd * g.bh + dh) * g.bw + dw;
      }
  return B2C_OK;
}

int tc_thin_gather_launch(const Geo& g, const void* big, const void* wp, int Kp, void* out_small, const float* bias,
                          int act, float slope, cudaStream_t st, const TcStats* stats) {
  if (!tc_thin_gather_supported(g) || Kp > 64 || Kp % 8 != 0) return B2C_ERR_UNSUPPORTED;
  if (stats && stats->kind && (stats->kind != 2 || stats->len != g.n || !stats->part ||
                               stats->part_bytes < 2 * tc_conv_stats_bytes(stats->len)))
    return B2C_ERR_INVALID;
  if ((reinterpret_cast<uintptr_t>(wp) & 15) != 0) return B2C_ERR_UNSUPPORTED;
  ThinGatherParams P;
  memset(&P, 0, sizeof(P));
  int rc = thin_src_fill(g, big, &P.src);
  if (rc) return rc;
  const int64_t rows = P.src.rows;
  MapDesc om;
  memset(&om, 0, sizeof(om));
  om.base = 0;
  om.dims[0] = g.sc; om.dims[1] = (int32_t)rows;
  om.strides[0] = 1; om.strides[1] = g.sc;
  const int obox[2] = {64, kBlockM};
  rc = encode_map(&P.out_map, out_small, om, 2, obox);
  if (rc) return rc;
  P.wp = static_cast<const __nv_bfloat16*>(wp);
  P.bias = bias; P.Kp = Kp; P.sc = g.sc;
  P.block_n = (g.sc + 15) / 16 * 16;
  P.num_tiles = (int32_t)cdiv(rows, (int64_t)kBlockM);
  P.act = act; P.slope = slope;
  P.idesc = idesc_bf16(kBlockM, P.block_n, 0, 0);
  P.tmem_cols = tmem_cols_for(2 * P.block_n);
  size_t smem = 1024 + (size_t)(1 + kThinStages + 2) * kThinTile + 256;
  if (stats && stats->kind) {
    P.stat_kind = 2; P.stat_len = stats->len; P.stat_part = stats->part;
    P.vps = g.sd * g.sh * g.sw;
    smem += (((size_t)2 * stats->len + 3) & ~(size_t)3) * sizeof(float) + 2048;
  }
  auto kern = P.stat_kind ? thin_gather_kernel<true> : thin_gather_kernel<false>;
  if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess)
    return B2C_ERR_CUDA;
  const int sms = num_sms();
  const int grid = P.num_tiles < 2 * sms ? P.num_tiles : 2 * sms;        // two resident CTAs per SM
  kern<<<grid, kThinThreads, smem, st>>>(P);
  if (stats) stats->nblocks = grid;
  return launched();
}

int tc_col2im1_k4s2_launch(const Geo& g, const void* col, void* big, const float* bias, int act, float slope,
                           cudaStream_t st) {
  Col2imK4S2Params P;
  memset(&P, 0, sizeof(P));
  MapDesc m;
  m.base = 0;
  m.dims[0] = 64; m.dims[1] = g.sw; m.dims[2] = g.sh; m.dims[3] = g.sd; m.dims[4] = g.n;
  m.strides[0] = 1; m.strides[1] = 64; m.strides[2] = (int64_t)g.sw * 64; m.strides[3] = (int64_t)g.sh * g.sw * 64;
  m.strides[4] = (int64_t)g.sd * g.sh * g.sw * 64;
  const int box[5] = {64, kC2RW, kC2RH, 1, 1};
  int rc = encode_map(&P.col_map, col, m, 5, box);
  if (rc) return rc;
  P.big = static_cast<__nv_bfloat16*>(big); P.bias = bias; P.act = act; P.slope = slope;
  P.n = g.n; P.sd = g.sd; P.sh = g.sh; P.sw = g.sw;
  P.htiles = (g.sh + kC2H - 1) / kC2H; P.wtiles = (g.sw + kC2W - 1) / kC2W;
  // d chunks: enough CTAs for ~4 waves of one-CTA-per-SM blocks, at least 8 output planes per chunk (2 halo planes each)
  const int64_t tiles = (int64_t)g.n * P.htiles * P.wtiles;
  int chunks = (int)cdiv((int64_t)4 * num_sms(), tiles > 0 ? tiles : 1);
  if (chunks > (g.sd + 7) / 8) chunks = (g.sd + 7) / 8;
  if (chunks < 1) chunks = 1;
  P.dch = (g.sd + chunks - 1) / chunks;
  P.dchunks = (g.sd + P.dch - 1) / P.dch;
  const int64_t grid = tiles * P.dchunks;
  if (grid == 0) return B2C_OK;
  if (grid > 0x7fffffffLL) return B2C_ERR_UNSUPPORTED;
  const size_t smem = 1024 + (size_t)kC2Ring * kC2PlaneStride + 64;
  if (cudaFuncSetAttribute(col2im1_k4s2_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess)
    return B2C_ERR_CUDA;
  col2im1_k4s2_tma_kernel<<<(unsigned)grid, kC2Threads, smem, st>>>(P);
  return launched();
}

// split-K over 128-voxel tiles: one CTA (= one fp32 partial) per contiguous tile range, two CTAs per SM
static void thin_wgrad_split(const Geo& g, int* tiles, int* per_cta, int* splits) {
  const int64_t rows = (int64_t)g.n * g.sd * g.sh * g.sw;
  *tiles = (int)cdiv(rows, (int64_t)kBlockM);
  const int want = 2 * num_sms();
  *per_cta = (int)cdiv((int64_t)*tiles, (int64_t)want);
  if (*per_cta < 1) *per_cta = 1;
  *splits = (int)cdiv((int64_t)*tiles, (int64_t)*per_cta);
}

size_t tc_thin_wgrad_workspace(const Geo& g) {
  if (!tc_thin_gather_supported(g)) return 0;
  int tiles, per, splits;
  thin_wgrad_split(g, &tiles, &per, &splits);
  return (size_t)splits * g.sc * 64 * sizeof(float);
}

int tc_thin_wgrad_launch(const Geo& g, const void* small, const void* big, float* dw, int flip, void* ws, size_t wsb,
                         cudaStream_t st) {
  if (!tc_thin_gather_supported(g)) return B2C_ERR_UNSUPPORTED;
  if (!ws || wsb < tc_thin_wgrad_workspace(g)) return B2C_ERR_WORKSPACE;
  ThinWgradParams P;
  memset(&P, 0, sizeof(P));
  int rc = thin_src_fill(g, big, &P.src);
  if (rc) return rc;
  MapDesc dm;
  memset(&dm, 0, sizeof(dm));
  dm.base = 0;
  dm.dims[0] = g.sc; dm.dims[1] = (int32_t)P.src.rows;
  dm.strides[0] = 1; dm.strides[1] = g.sc;
  const int dbox[2] = {64, kWgVox};
  rc = encode_map(&P.dy_map, small, dm, 2, dbox);
  if (rc) return rc;
  int splits;
  thin_wgrad_split(g, &P.num_tiles, &P.tiles_per_cta, &splits);
  P.ws = static_cast<float*>(ws);
  P.sc = g.sc;
  P.ach = g.sc > 64 ? 2 : 1;
  P.stages = P.ach == 2 ? 2 : 3;                 // <= 96 KB of tiles: two CTAs per SM
  P.idesc = idesc_bf16(128, 64, 1, 1);
  P.tmem_cols = tmem_cols_for(64);
  const size_t smem = 1024 + (size_t)P.stages * (P.ach * 2 * kSlabBytes + kThinTile) + 256;
  if (cudaFuncSetAttribute(thin_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess)
    return B2C_ERR_CUDA;
  thin_wgrad_kernel<<<splits, kThinWgThreads, smem, st>>>(P);
  rc = launched();
  if (rc) return rc;
  const int T = g.taps();
  thin_wgrad_reduce_kernel<<<(unsigned)cdiv((int64_t)g.sc * T, 256), 256, 0, st>>>(P.ws, dw, g.sc, T, splits, flip);
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
