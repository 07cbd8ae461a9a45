// tc_thin.cuh — part of the tc_conv.cu translation unit (shares its PTX helpers, tensor-map encoder
// and sticky error word).  Fused gather-form kernel for convolutions whose big-side tensor has ONE
// channel and at most 64 taps (dcgan.py:161 Conv3d(1 -> ndf, k4, s2, p1); biggan.py:78 input conv):
//
//   out[v, c] = act(bias[c] + sum_t A[v, t] * W[c, t]),   A[v, t] = big[n, o*s + k*dl - p]  (0 outside)
//
// The im2col matrix A never exists in HBM: four producer warps gather one 128-voxel x 64-tap tile
// straight into 128-byte-swizzled shared memory (the K-major operand layout tcgen05 reads), one warp
// issues tcgen05.mma into two TMEM accumulator sets, four epilogue warps do bias / activation /
// bf16 and hand 128 x 64 slabs to TMA stores.  HBM traffic = |big| + |out| (the roofline of this
// layer class); the input tensor (134 MB at BASELINE config 2) stays L2 resident.
//
// Roles (288 threads): warps 0-3 producers, warp 4 MMA issuer, warps 5-8 epilogue.

constexpr int kThinThreads = 288;
constexpr int kThinStages = 3;            // 6 x 16 KB tiles per CTA: two CTAs fit one SM
constexpr uint32_t kThinTile = kBlockM * 128;          // 16 KB: 128 rows x 128 B (64 taps)

struct ThinTap { int8_t d, h, w, _pad; int32_t off; };      // tap offsets {kd*dl, kh*dl, kw*dl} and their element offset

// the one-channel source tensor and how im2col rows are cut from it (shared by the forward and the
// weight-gradient kernels)
struct ThinSrc {
  Geo g;
  const __nv_bfloat16* big;
  int64_t rows;
  int32_t T, fast_k4s2;
  ThinTap tap[64];
};

struct alignas(64) ThinGatherParams {
  CUtensorMap out_map;                // [rows][sc] bf16, box {64 ch, 128 rows}
  ThinSrc src;
  const __nv_bfloat16* wp;            // [sc][Kp]
  const float* bias;
  int32_t Kp, sc, block_n, num_tiles, act;
  float slope;
  uint32_t idesc, tmem_cols;
  // per-sample sums of the stored output for the LayerNorm that follows (dcgan.py:161-162), as in tc_conv_kernel
  int32_t stat_kind, stat_len, vps;   // 0 / 2; samples; voxels (rows) per sample
  float* stat_part;                   // [gridDim.x][2][stat_len]
};

// element t (0..63) of row r lives at r*128 + ((t/8) ^ (r & 7))*16 + (t % 8)*2  (SWIZZLE_128B)
__device__ __forceinline__ uint32_t thin_chunk(uint32_t tile, int r, int c) {
  return tile + (uint32_t)r * 128u + ((uint32_t)(c ^ (r & 7)) << 4);
}

// One im2col row through the tap table, branch free: every load is unconditional (an out-of-range tap
// reads element 0 and is masked afterwards), so the compiler issues all of them before the first use.
template <int TMAX>
__device__ __forceinline__ void thin_gather_row(const ThinSrc& P, const uint16_t* __restrict__ xb, int base,
                                                int id0, int ih0, int iw0, uint32_t* pk) {
  const Geo& g = P.g;
  uint32_t e[TMAX];
#pragma unroll
  for (int t = 0; t < TMAX; t++) {
    const ThinTap tp = P.tap[t];
    const int id = id0 + tp.d, ih = ih0 + tp.h, iw = iw0 + tp.w;
    const bool ok = t < P.T && id >= 0 && id < g.bd && ih >= 0 && ih < g.bh && iw >= 0 && iw < g.bw;
    const uint32_t val = (uint32_t)__ldg(xb + (ok ? base + tp.off : 0));
    e[t] = ok ? val : 0u;
  }
#pragma unroll
  for (int i = 0; i < 32; i++) pk[i] = 2 * i + 1 < TMAX ? (e[2 * i] | (e[2 * i + 1] << 16)) : 0u;
}

// Row r of the 128-voxel x 64-tap im2col tile `tile`, written to the 128-byte-swizzled tile at `dst`.
__device__ __forceinline__ void thin_build_row(const ThinSrc& S, const uint16_t* __restrict__ xb, uint32_t dst, int tile,
                                               int r, int warp, int lane) {
  const Geo& g = S.g;
  const int64_t v = (int64_t)tile * kBlockM + r;
  if (S.fast_k4s2 == 2) {
    // k = 4, stride 2, pad 1, sw % 32 == 0: the 32 lanes of a warp are 32 consecutive ow of ONE (n, od, oh)
    // line (rows % 32 == 0, so a warp is all valid or all tail).  Taps (kd, kh, 0..3) = input elements
    // 2*ow-1 .. 2*ow+2: each lane loads only its aligned middle pair (2*ow, 2*ow+1) per line - 16
    // independent loads issued back to back - and takes the outer two elements from its neighbours'
    // pairs through warp shuffles.  All index arithmetic is 32-bit and warp-uniform where possible.
    const uint32_t v0 = (uint32_t)tile * kBlockM + (uint32_t)(warp * 32);        // < 2^31 (host check)
    if (v0 >= (uint32_t)S.rows) {
#pragma unroll
      for (int c = 0; c < 8; c++) st_shared_v4(thin_chunk(dst, r, c), 0u, 0u, 0u, 0u);
    } else {
      uint32_t t = v0 / (uint32_t)g.sw;
      const int ow = (int)(v0 - t * (uint32_t)g.sw) + lane;
      const uint32_t t2 = t / (uint32_t)g.sh;
      const int oh = (int)(t - t2 * (uint32_t)g.sh);
      const uint32_t n = t2 / (uint32_t)g.sd;
      const int od = (int)(t2 - n * (uint32_t)g.sd);
      const int id0 = 2 * od - 1, ih0 = 2 * oh - 1;
      const int sH = g.bw, sD = g.bh * g.bw;
      // element offset of (n, id0, ih0, 2*ow); lines with id / ih out of range are never dereferenced
      const int base = ((int)n * g.bd + id0) * sD + ih0 * sH + 2 * ow;
      // all loads of the tile row are issued before anything consumes them: the middle pairs, and on
      // the two edge lanes the one element their missing neighbour would have supplied
      uint32_t mid[16], edge[16];
      const bool left = lane == 0 && ow > 0, right = lane == 31 && 2 * ow + 2 < g.bw;
#pragma unroll
      for (int l = 0; l < 16; l++) {
        const int id = id0 + (l >> 2), ih = ih0 + (l & 3);
        const bool ok = id >= 0 && id < g.bd && ih >= 0 && ih < g.bh;   // warp-uniform
        const uint16_t* px = xb + base + (l >> 2) * sD + (l & 3) * sH;
        mid[l] = ok ? __ldg(reinterpret_cast<const uint32_t*>(px)) : 0u;
        edge[l] = 0u;
        if (ok && left) edge[l] = (uint32_t)__ldg(px - 1);
        if (ok && right) edge[l] = (uint32_t)__ldg(px + 2);
      }
      uint32_t w4[4];
#pragma unroll
      for (int l = 0; l < 16; l++) {
        uint32_t e0 = __shfl_up_sync(0xffffffffu, mid[l], 1) >> 16;                 // x[2*ow - 1]
        uint32_t e3 = __shfl_down_sync(0xffffffffu, mid[l], 1) & 0xffffu;           // x[2*ow + 2]
        if (lane == 0) e0 = edge[l];
        if (lane == 31) e3 = edge[l];
        w4[(l & 1) * 2] = e0 | (mid[l] << 16);
        w4[(l & 1) * 2 + 1] = (mid[l] >> 16) | (e3 << 16);
        if (l & 1) st_shared_v4(thin_chunk(dst, r, l >> 1), w4[0], w4[1], w4[2], w4[3]);
      }
    }
  } else if (S.fast_k4s2 == 3) {
    // k = 3, stride 1, pad 1, sw % 32 == 0 (biggan.py:78 input conv, the data gradient of biggan.py:42): the 32 lanes
    // of a warp are 32 consecutive ow of ONE (n, od, oh) line.  Taps (kd, kh, 0..2) = input elements ow-1 .. ow+1:
    // each lane loads its own element of the nine (kd, kh) lines - nine coalesced 64-byte warp loads instead of 27
    // predicated scalar ones - and takes the outer two from its neighbours through warp shuffles.
    const uint32_t v0 = (uint32_t)tile * kBlockM + (uint32_t)(warp * 32);
    if (v0 >= (uint32_t)S.rows) {
#pragma unroll
      for (int c = 0; c < 8; c++) st_shared_v4(thin_chunk(dst, r, c), 0u, 0u, 0u, 0u);
    } else {
      uint32_t t = v0 / (uint32_t)g.sw;
      const int ow = (int)(v0 - t * (uint32_t)g.sw) + lane;
      const uint32_t t2 = t / (uint32_t)g.sh;
      const int oh = (int)(t - t2 * (uint32_t)g.sh);
      const uint32_t n = t2 / (uint32_t)g.sd;
      const int od = (int)(t2 - n * (uint32_t)g.sd);
      const int sH = g.bw, sD = g.bh * g.bw;
      const int base = ((int)n * g.bd + od - 1) * sD + (oh - 1) * sH + ow;
      const bool left = lane == 0 && ow > 0, right = lane == 31 && ow + 1 < g.bw;
      uint32_t mid[9], edge[9];
#pragma unroll
      for (int l = 0; l < 9; l++) {
        const int id = od - 1 + l / 3, ih = oh - 1 + l % 3;
        const bool ok = id >= 0 && id < g.bd && ih >= 0 && ih < g.bh;     // warp-uniform
        const uint16_t* px = xb + base + (l / 3) * sD + (l % 3) * sH;
        mid[l] = ok ? (uint32_t)__ldg(px) : 0u;
        edge[l] = 0u;
        if (ok && left) edge[l] = (uint32_t)__ldg(px - 1);
        if (ok && right) edge[l] = (uint32_t)__ldg(px + 1);
      }
      uint32_t e[32];
#pragma unroll
      for (int l = 0; l < 9; l++) {
        uint32_t e0 = __shfl_up_sync(0xffffffffu, mid[l], 1), e2 = __shfl_down_sync(0xffffffffu, mid[l], 1);
        if (lane == 0) e0 = edge[l];
        if (lane == 31) e2 = edge[l];
        e[3 * l] = e0; e[3 * l + 1] = mid[l]; e[3 * l + 2] = e2;
      }
#pragma unroll
      for (int i = 27; i < 32; i++) e[i] = 0u;
#pragma unroll
      for (int c = 0; c < 4; c++)
        st_shared_v4(thin_chunk(dst, r, c), e[8 * c] | (e[8 * c + 1] << 16), e[8 * c + 2] | (e[8 * c + 3] << 16),
                     e[8 * c + 4] | (e[8 * c + 5] << 16), e[8 * c + 6] | (e[8 * c + 7] << 16));
#pragma unroll
      for (int c = 4; c < 8; c++) st_shared_v4(thin_chunk(dst, r, c), 0u, 0u, 0u, 0u);    // the weight gradient reads 64 taps
    }
  } else
  if (v >= S.rows) {
#pragma unroll
    for (int c = 0; c < 8; c++) st_shared_v4(thin_chunk(dst, r, c), 0u, 0u, 0u, 0u);
  } else {
    // any geometry with <= 64 taps (host-built tap table, 32-bit indices): all predicated loads of the
    // row are issued before the first store consumes one
    uint32_t q = (uint32_t)v;
    uint32_t t1 = q / (uint32_t)g.sw;
    const int ow = (int)(q - t1 * (uint32_t)g.sw);
    uint32_t t2 = t1 / (uint32_t)g.sh;
    const int oh = (int)(t1 - t2 * (uint32_t)g.sh);
    const uint32_t n = t2 / (uint32_t)g.sd;
    const int od = (int)(t2 - n * (uint32_t)g.sd);
    const int id0 = od * g.s[0] - g.p[0], ih0 = oh * g.s[1] - g.p[1], iw0 = ow * g.s[2] - g.p[2];
    const int base = (((int)n * g.bd + id0) * g.bh + ih0) * g.bw + iw0;
    uint32_t pk[32];
    if (S.T <= 32) thin_gather_row<32>(S, xb, base, id0, ih0, iw0, pk);
    else thin_gather_row<64>(S, xb, base, id0, ih0, iw0, pk);
#pragma unroll
    for (int c = 0; c < 8; c++) st_shared_v4(thin_chunk(dst, r, c), pk[4 * c], pk[4 * c + 1], pk[4 * c + 2], pk[4 * c + 3]);
  }
}

template <bool STATS>
__global__ void __launch_bounds__(kThinThreads, 2) thin_gather_kernel(const __grid_constant__ ThinGatherParams P) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t sbase = (raw + 1023u) & ~1023u;
  uint8_t* sgen = smem_raw + (sbase - raw);
  const uint32_t w_tile = sbase;                                   // 128 x 64 filter tile (zero padded)
  const uint32_t a_ring = sbase + kThinTile;                       // kThinStages A tiles
  const uint32_t staging = a_ring + kThinStages * kThinTile;       // two output slabs
  const uint32_t bar0 = staging + 2 * kThinTile;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(sgen + (bar0 - sbase) + (2 * kThinStages + 4) * 8);
  auto afull = [&](int s) { return bar0 + 8u * s; };
  auto aempty = [&](int s) { return bar0 + 8u * (kThinStages + s); };
  auto tfull = [&](int a) { return bar0 + 8u * (2 * kThinStages + a); };
  auto tempty = [&](int a) { return bar0 + 8u * (2 * kThinStages + 2 + a); };
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (int s = 0; s < kThinStages; s++) { mbar_init(afull(s), 4); mbar_init(aempty(s), 1); }
    for (int a = 0; a < 2; a++) { mbar_init(tfull(a), 1); mbar_init(tempty(a), 4); }
    fence_barrier_init();
    prefetch_tmap(&P.out_map);
  }
  if (warp == 4) { tmem_alloc(smem_u32(tmem_slot), P.tmem_cols); tmem_relinquish(); }
  if (warp < 4) {
    // filter tile: row = channel (zero rows beyond sc), 8 chunks of 8 taps (zero beyond Kp)
    const int r = threadIdx.x;
    for (int c = 0; c < 8; c++) {
      uint4 v = make_uint4(0u, 0u, 0u, 0u);
      if (r < P.sc && c * 8 < P.Kp) v = __ldg(reinterpret_cast<const uint4*>(P.wp + (int64_t)r * P.Kp + c * 8));
      st_shared_v4(thin_chunk(w_tile, r, c), v.x, v.y, v.z, v.w);
    }
    fence_proxy_async();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(tmem_slot);

  if (warp < 4) {
    // ===== producers: thread r builds row r of every A tile =====
    const int r = threadIdx.x;
    const uint16_t* xb = reinterpret_cast<const uint16_t*>(P.src.big);
    uint32_t it = 0;
    for (int tile = blockIdx.x; tile < P.num_tiles; tile += gridDim.x, it++) {
      const uint32_t s = it % kThinStages, ph = (it / kThinStages) & 1u;
      mbar_wait(aempty(s), ph ^ 1u);
      const uint32_t dst = a_ring + s * kThinTile;
      thin_build_row(P.src, xb, dst, tile, r, warp, lane);
      fence_proxy_async();                      // generic-proxy stores -> visible to the tensor core's async proxy
      __syncwarp();
      if (lane == 0) mbar_arrive(afull(s));
    }
  } else if (warp == 4) {
    // ===== MMA issuer =====
    if (lane == 0) {
      constexpr uint64_t hi = smem_desc_hi(16, 1024, 2);
      uint32_t it = 0;
      bool ok = true;
      for (int tile = blockIdx.x; tile < P.num_tiles && ok; tile += gridDim.x, it++) {
        const uint32_t s = it % kThinStages, ph = (it / kThinStages) & 1u;
        const uint32_t acc = it & 1u, aph = (it >> 1) & 1u;
        ok = mbar_wait(tempty(acc), aph ^ 1u);
        ok = ok && mbar_wait(afull(s), ph);
        tc_fence_after();
        const uint32_t a_addr = a_ring + s * kThinTile;
        const uint32_t d_tmem = tmem_base + acc * (uint32_t)P.block_n;
        const int nk = (P.Kp + 15) / 16;
#pragma unroll
        for (int k = 0; k < 4; k++)
          if (k < nk) umma_f16(d_tmem, smem_desc(a_addr + k * 32, hi), smem_desc(w_tile + k * 32, hi), P.idesc, k ? 1u : 0u);
        umma_commit(aempty(s));
        umma_commit(tfull(acc));
      }
    }
  } else {
    // ===== epilogue =====
    const int grp = warp & 3;
    const int row = grp * 32 + lane;
    const int et = threadIdx.x - 160;                    // 0..127
    const uint32_t lane_off = (uint32_t)(grp * 32) << 16;
    const int nslab = (P.block_n + 63) >> 6;
    uint32_t it = 0, slab_count = 0;
    float* sacc = reinterpret_cast<float*>(sgen + (bar0 - sbase) + 256);
    float* sscr = sacc + ((2 * P.stat_len + 3) & ~3);
    if (STATS) {
      for (int i = et; i < 2 * P.stat_len; i += 128) sacc[i] = 0.f;
      named_bar_sync(2, 128);
    }
    for (int tile = blockIdx.x; tile < P.num_tiles; tile += gridDim.x, it++) {
      const uint32_t acc = it & 1u, aph = (it >> 1) & 1u;
      mbar_wait(tfull(acc), aph);
      tc_fence_after();
      float rs = 0.f, rq = 0.f;
      for (int j = 0; j < nslab; j++, slab_count++) {
        const uint32_t sbuf = staging + (slab_count & 1u) * kThinTile;
        if (et == 0) tma_store_wait_read<1>();
        named_bar_sync(2, 128);
        uint32_t v[64];
        const int ncol = P.block_n - j * 64 < 64 ? P.block_n - j * 64 : 64;
        const uint32_t tcol = tmem_base + lane_off + acc * (uint32_t)P.block_n + j * 64;
#pragma unroll
        for (int q = 0; q < 4; q++)
          if (q * 16 < ncol) tmem_ld16(tcol + q * 16, v + q * 16);
        tmem_ld_wait();
        if (P.bias) {
          const float* bp = P.bias + j * 64;
          const int nb = P.sc - j * 64;
#pragma unroll
          for (int i = 0; i < 64; i++)
            if (i < nb && i < ncol) v[i] = __float_as_uint(__uint_as_float(v[i]) + __ldg(bp + i));
        }
        apply_act64(v, P.act, P.slope);
#pragma unroll
        for (int ch = 0; ch < 8; ch++)
          if (ch * 8 < ncol) {
            const uint32_t p0 = pack_bf16x2(__uint_as_float(v[ch * 8 + 0]), __uint_as_float(v[ch * 8 + 1])),
                           p1 = pack_bf16x2(__uint_as_float(v[ch * 8 + 2]), __uint_as_float(v[ch * 8 + 3])),
                           p2 = pack_bf16x2(__uint_as_float(v[ch * 8 + 4]), __uint_as_float(v[ch * 8 + 5])),
                           p3 = pack_bf16x2(__uint_as_float(v[ch * 8 + 6]), __uint_as_float(v[ch * 8 + 7]));
            st_shared_v4(sbuf + row * 128 + ((ch ^ (row & 7)) << 4), p0, p1, p2, p3);
            if (STATS && j * 64 + ch * 8 < P.sc) {            // the stored (rounded) values; sc % 8 == 0
              const uint32_t pk[4] = {p0, p1, p2, p3};
#pragma unroll
              for (int i = 0; i < 4; i++) {
                const float lo = __uint_as_float(pk[i] << 16), hi = __uint_as_float(pk[i] & 0xffff0000u);
                rs += lo + hi; rq += lo * lo + hi * hi;
              }
            }
          }
        fence_proxy_async();
        named_bar_sync(2, 128);
        if (et == 0) {
          tma_store_2d(&P.out_map, sbuf, j * 64, tile * kBlockM);
          tma_store_commit();
        }
      }
      if (STATS) {
        // rows of the tile are consecutive voxels of the flattened [n][d][h][w] lattice: sample = row / vps.  A warp
        // whose 32 rows lie in ONE sample (always, when vps % 32 == 0) is reduced with shuffles; thread 0 then adds
        // the four warp sums in a fixed order.  Warps that straddle a sample boundary leave their rows in the scratch.
        const int64_t r0 = (int64_t)tile * kBlockM;
        const bool rv = r0 + row < P.src.rows;
        const int nrow = rv ? (int)((r0 + row) / P.vps) : -1;
        const int nfirst = __shfl_sync(0xffffffffu, nrow, 0);
        const bool uni = __all_sync(0xffffffffu, nrow == nfirst || nrow < 0) && nfirst >= 0;
        float ws_ = rv ? rs : 0.f, wq_ = rv ? rq : 0.f;
        if (uni) {
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) { ws_ += __shfl_xor_sync(0xffffffffu, ws_, o); wq_ += __shfl_xor_sync(0xffffffffu, wq_, o); }
          if (lane == 0) { sscr[grp * 4] = ws_; sscr[grp * 4 + 1] = wq_; sscr[grp * 4 + 2] = __int_as_float(nfirst); }
        } else {
          if (lane == 0) sscr[grp * 4 + 2] = __int_as_float(-1);
          sscr[16 + row * 3] = ws_; sscr[16 + row * 3 + 1] = wq_; sscr[16 + row * 3 + 2] = __int_as_float(nrow);
        }
        named_bar_sync(2, 128);
        if (et == 0) {
          for (int g2 = 0; g2 < 4; g2++) {
            const int n = __float_as_int(sscr[g2 * 4 + 2]);
            if (n >= 0) {
              if (n < P.stat_len) { sacc[n] += sscr[g2 * 4]; sacc[P.stat_len + n] += sscr[g2 * 4 + 1]; }
            } else {
              for (int r = g2 * 32; r < g2 * 32 + 32; r++) {
                const int nr = __float_as_int(sscr[16 + r * 3 + 2]);
                if (nr >= 0 && nr < P.stat_len) { sacc[nr] += sscr[16 + r * 3]; sacc[P.stat_len + nr] += sscr[16 + r * 3 + 1]; }
              }
            }
          }
        }
        // the scratch is rewritten only after the next tile's slab barriers: thread 0 passes them after its reads
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(tempty(acc));
    }
    if (STATS) {
      named_bar_sync(2, 128);
      float* part = P.stat_part + (size_t)blockIdx.x * 2 * P.stat_len;
      for (int i = et; i < 2 * P.stat_len; i += 128) part[i] = sacc[i];
    }
    if (et == 0) tma_store_wait<0>();
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 4) tmem_dealloc(tmem_base, P.tmem_cols);
}

// =============================================================================
// Weight gradient of the same layer class, fused the same way:
//
//   dW[c, t] = sum_v dY[v, c] * A[v, t]
//
// a GEMM with the voxel index as K.  Both operands are MN-major: dY is read in its natural [voxel][channel]
// form by TMA (64 x 64 slabs), and the im2col tile the producers build ([voxel][tap], 128-byte swizzled
// rows) is byte for byte the tile of the forward kernel - there it is a K-major A operand, here an
// MN-major B operand with N = 64 taps.  M = 128 channel rows (rows >= sc hold zeros or are dropped).
// Each CTA owns a contiguous range of 128-voxel tiles and writes one fp32 partial [sc][64] to the
// workspace; thin_wgrad_reduce_kernel sums the partials in a fixed order (deterministic) and writes
// torch's [sc][T] layout (tap-reversed for the mirrored Cout = 1 convolutions).
//
// Roles (192 threads): warps 0-3 producers then epilogue, warp 4 TMA (dY), warp 5 MMA issuer.
// =============================================================================
constexpr int kThinWgThreads = 192;

struct alignas(64) ThinWgradParams {
  CUtensorMap dy_map;                 // [rows][sc] bf16, box {64 ch, 64 rows}
  ThinSrc src;
  float* ws;                          // [splits][sc][64] fp32
  int32_t sc, ach, num_tiles, tiles_per_cta, stages;
  uint32_t idesc, tmem_cols;
};

__global__ void __launch_bounds__(kThinWgThreads, 2) thin_wgrad_kernel(const __grid_constant__ ThinWgradParams P) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t sbase = (raw + 1023u) & ~1023u;
  uint8_t* sgen = smem_raw + (sbase - raw);
  const int stages = P.stages;
  const uint32_t a_bytes = (uint32_t)P.ach * 2u * kSlabBytes;            // 2 voxel halves x ach channel slabs
  const uint32_t stage_bytes = a_bytes + kThinTile;
  const uint32_t bar0 = sbase + stages * stage_bytes;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(sgen + stages * stage_bytes + (2 * stages + 1) * 8);
  auto full_bar = [&](int s) { return bar0 + 8u * s; };
  auto empty_bar = [&](int s) { return bar0 + 8u * (stages + s); };
  const uint32_t acc_bar = bar0 + 8u * (2 * stages);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  const int t0 = blockIdx.x * P.tiles_per_cta;
  const int t1 = t0 + P.tiles_per_cta < P.num_tiles ? t0 + P.tiles_per_cta : P.num_tiles;
  const int iters = t1 > t0 ? t1 - t0 : 0;

  if (threadIdx.x == 0) {
    for (int s = 0; s < stages; s++) { mbar_init(full_bar(s), 5); mbar_init(empty_bar(s), 1); }
    mbar_init(acc_bar, 1);
    fence_barrier_init();
    prefetch_tmap(&P.dy_map);
  }
  if (warp == 5) { tmem_alloc(smem_u32(tmem_slot), P.tmem_cols); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(tmem_slot);

  if (warp < 4) {
    // ===== producers: thread r builds row r (one voxel, 64 taps) of every B tile =====
    const int r = threadIdx.x;
    const uint16_t* xb = reinterpret_cast<const uint16_t*>(P.src.big);
    for (int it = 0; it < iters; it++) {
      const int s = it % stages;
      const uint32_t ph = (uint32_t)(it / stages) & 1u;
      mbar_wait(empty_bar(s), ph ^ 1u);
      thin_build_row(P.src, xb, sbase + s * stage_bytes + a_bytes, t0 + it, r, warp, lane);
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) mbar_arrive(full_bar(s));
    }
    // ===== epilogue: thread r owns channel r of the partial =====
    mbar_wait(acc_bar, 0);
    tc_fence_after();
    const uint32_t trow = tmem_base + ((uint32_t)(warp * 32) << 16);
    float* orow = P.ws + ((int64_t)blockIdx.x * P.sc + r) * 64;
#pragma unroll
    for (int c0 = 0; c0 < 64; c0 += 16) {
      uint32_t v[16];
      if (iters > 0) { tmem_ld16(trow + c0, v); tmem_ld_wait(); }
      else {
#pragma unroll
        for (int j = 0; j < 16; j++) v[j] = 0u;
      }
      if (r < P.sc) {
#pragma unroll
        for (int j = 0; j < 16; j += 4) *reinterpret_cast<uint4*>(orow + c0 + j) = make_uint4(v[j], v[j + 1], v[j + 2], v[j + 3]);
      }
    }
  } else if (warp == 4) {
    if (lane == 0) {
      for (int it = 0; it < iters; it++) {
        const int s = it % stages;
        const uint32_t ph = (uint32_t)(it / stages) & 1u;
        if (!mbar_wait(empty_bar(s), ph ^ 1u)) break;
        mbar_arrive_expect_tx(full_bar(s), a_bytes);
        const uint32_t a_dst = sbase + s * stage_bytes;
        const int v0 = (t0 + it) * kBlockM;
        for (int h = 0; h < 2; h++)
          for (int j = 0; j < P.ach; j++)
            tma_load_2d(a_dst + (uint32_t)(h * P.ach + j) * kSlabBytes, &P.dy_map, full_bar(s), j * 64, v0 + h * 64);
      }
    }
  } else {
    if (lane == 0) {
      // MN-major SW128: LBO = next 64 channels (one slab on), SBO = 1024 (next 8 voxel rows).  With one
      // channel slab (sc <= 64) accumulator rows 64..127 read whatever follows - they are never stored.
      constexpr uint64_t hi = smem_desc_hi(kSlabBytes, 1024, 2);
      bool ok = true;
      for (int it = 0; it < iters && ok; it++) {
        const int s = it % stages;
        const uint32_t ph = (uint32_t)(it / stages) & 1u;
        ok = mbar_wait(full_bar(s), ph);
        tc_fence_after();
        const uint32_t a_addr = sbase + s * stage_bytes, b_addr = a_addr + a_bytes;
#pragma unroll
        for (int k = 0; k < kBlockM / 16; k++)      // 16 voxel rows per MMA
          umma_f16(tmem_base, smem_desc(a_addr + (uint32_t)((k >> 2) * P.ach) * kSlabBytes + (k & 3) * 2048, hi),
                   smem_desc(b_addr + k * 2048, hi), P.idesc, (it | k) != 0 ? 1u : 0u);
        umma_commit(empty_bar(s));
      }
      if (iters > 0) umma_commit(acc_bar);
      else mbar_arrive(acc_bar);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 5) tmem_dealloc(tmem_base, P.tmem_cols);
}

// dw[c][t] = sum_split ws[split][c][flip ? T-1-t : t]
__global__ void __launch_bounds__(256)
thin_wgrad_reduce_kernel(const float* __restrict__ ws, float* __restrict__ dw, int sc, int T, int splits, int flip) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= sc * T) return;
  const int c = idx / T, t = idx - c * T;
  const float* p = ws + (int64_t)c * 64 + (flip ? T - 1 - t : t);
  const int64_t pitch = (int64_t)sc * 64;
  float acc = 0.f;
  int s = 0;
  for (; s + 8 <= splits; s += 8) {
    float v[8];
#pragma unroll
    for (int j = 0; j < 8; j++) v[j] = p[(s + j) * pitch];
#pragma unroll
    for (int j = 0; j < 8; j++) acc += v[j];
  }
  for (; s < splits; s++) acc += p[s * pitch];
  dw[idx] = acc;
}


// =============================================================================
// col2im onto a ONE-channel tensor for k = 4, stride 2, pad 1 (dcgan.py:67 ConvTranspose3d(ngf, 1) and the data
// gradient of dcgan.py:161), TMA-staged.  The GEMM in front leaves col[small voxel][64] (128 B rows, columns in the
// order t' = a*8 + r of thin.cu:k4s2_perm); output voxel block 2j + r gathers, per (dd, dh) neighbour line, group
// a0 = (dd != 0, dh != 0, 0) of row jw and group a0 + 1 of rows jw - 1 / jw + 1.
//
// The first version read those 16-byte groups straight from global memory, a warp = 32 consecutive jw: every load
// instruction touched 32 different 128-byte lines (one L1 tag lookup each) and the kernel ran at 1.6 TB/s of
// algorithmic bytes.  Here a CTA owns a 4 (h) x 64 (w) tile of small voxels and marches along d (4 x 32 tiles with two
// CTAs per SM measured 2 % slower): each col plane of the tile plus its halo (6 x 66 rows x 128 B) arrives by ONE TMA
// box copy (out-of-range rows / lines / planes are
// zero-filled by TMA: the padding logic disappears) into a ring of four 128-byte-swizzled planes - three in use, one
// in flight - and the 27 groups a thread needs per output plane are conflict-free LDS.128 (8 consecutive rows of one
// chunk column hit 8 different bank groups under the swizzle).  HBM sees each col row 1.55x (h / w halo) instead of
// nine sector requests per row through L1.
// =============================================================================
constexpr int kC2H = 4, kC2W = 64, kC2RH = kC2H + 2, kC2RW = kC2W + 2, kC2Ring = 4, kC2Threads = kC2H * kC2W;
constexpr uint32_t kC2PlaneBytes = kC2RH * kC2RW * 128;                 // 50688: what one TMA box delivers
constexpr uint32_t kC2PlaneStride = (kC2PlaneBytes + 1023u) & ~1023u;   // ring slots stay 1024-byte aligned (swizzle atom)

struct alignas(64) Col2imK4S2Params {
  CUtensorMap col_map;                // [n][sd][sh][sw][64] bf16, box {64, kC2RW, kC2RH, 1, 1}
  __nv_bfloat16* big;
  const float* bias;
  int32_t n, sd, sh, sw, dch, htiles, wtiles, dchunks, act;
  float slope;
};

__global__ void __launch_bounds__(kC2Threads, 1) col2im1_k4s2_tma_kernel(const __grid_constant__ Col2imK4S2Params P) {
  extern __shared__ uint8_t c2_smem_raw[];
  const uint32_t base = (smem_u32(c2_smem_raw) + 1023u) & ~1023u;
  const uint32_t bars = base + kC2Ring * kC2PlaneStride;
  auto full_bar = [&](int s) { return bars + (uint32_t)s * 8u; };
  int b = blockIdx.x;
  const int wt = b % P.wtiles; b /= P.wtiles;
  const int ht = b % P.htiles; b /= P.htiles;
  const int dc = b % P.dchunks;
  const int in = b / P.dchunks;
  const int d0 = dc * P.dch;
  const int nout = min(P.dch, P.sd - d0);
  const int tid = threadIdx.x, jwl = tid % kC2W, jhl = tid / kC2W;
  const int jw = wt * kC2W + jwl, jh = ht * kC2H + jhl;
  const bool live = jw < P.sw && jh < P.sh;
  if (tid == 0) {
    prefetch_tmap(&P.col_map);
    for (int s = 0; s < kC2Ring; s++) mbar_init(full_bar(s), 1);
    fence_barrier_init();
  }
  __syncthreads();
  auto load_plane = [&](int q) {                       // q: plane index relative to d0 - 1 (thread 0 only)
    const int slot = q & (kC2Ring - 1);
    mbar_arrive_expect_tx(full_bar(slot), kC2PlaneBytes);
    tma_load_5d(base + (uint32_t)slot * kC2PlaneStride, &P.col_map, full_bar(slot), 0, wt * kC2W - 1, ht * kC2H - 1,
                d0 - 1 + q, in);
  };
  if (tid == 0) { load_plane(0); load_plane(1); load_plane(2); }
  const float b0 = P.bias ? P.bias[0] : 0.f;
  const int bh = 2 * P.sh, bw = 2 * P.sw;
  for (int o = 0; o < nout; o++) {
    if (tid == 0 && o + 3 < nout + 2) load_plane(o + 3);     // its slot was released by the barrier that ended o - 1
    if (o == 0) { mbar_wait(full_bar(0), 0); mbar_wait(full_bar(1), 0); }
    mbar_wait(full_bar((o + 2) & (kC2Ring - 1)), ((o + 2) >> 2) & 1);
    float acc[8];
#pragma unroll
    for (int r = 0; r < 8; r++) acc[r] = b0;
#pragma unroll
    for (int dd = -1; dd <= 1; dd++) {
      const uint32_t plane = base + (uint32_t)((o + 1 + dd) & (kC2Ring - 1)) * kC2PlaneStride;
#pragma unroll
      for (int dh = -1; dh <= 1; dh++) {
        const int a0 = ((dd != 0) << 2) | ((dh != 0) << 1);
        const int lin = (jhl + 1 + dh) * kC2RW + jwl + 1;        // own row in the box; neighbours lin -+ 1
        const uint4 A0 = ld_shared_v4(plane + (uint32_t)lin * 128u + (uint32_t)((a0 ^ (lin & 7)) << 4));
        const uint4 L = ld_shared_v4(plane + (uint32_t)(lin - 1) * 128u + (uint32_t)(((a0 + 1) ^ ((lin - 1) & 7)) << 4));
        const uint4 R = ld_shared_v4(plane + (uint32_t)(lin + 1) * 128u + (uint32_t)(((a0 + 1) ^ ((lin + 1) & 7)) << 4));
        const uint32_t a0v[4] = {A0.x, A0.y, A0.z, A0.w}, lv[4] = {L.x, L.y, L.z, L.w}, rv[4] = {R.x, R.y, R.z, R.w};
#pragma unroll
        for (int m = 0; m < 4; m++) {                 // m = rd*2 + rh; register m holds (rw = 0 | rw = 1)
          const bool ok = (dd == 0 || ((m >> 1) & 1) == (dd > 0)) && (dh == 0 || (m & 1) == (dh > 0));
          if (ok) {
            acc[2 * m] += __uint_as_float(a0v[m] << 16) + __uint_as_float(lv[m] << 16);                  // rw = 0
            acc[2 * m + 1] += __uint_as_float(a0v[m] & 0xffff0000u) + __uint_as_float(rv[m] & 0xffff0000u);   // rw = 1
          }
        }
      }
    }
    if (live) {
      const int jd = d0 + o;
#pragma unroll
      for (int rd = 0; rd < 2; rd++)
#pragma unroll
        for (int rh = 0; rh < 2; rh++) {
          const float v0 = act_f(acc[rd * 4 + rh * 2], P.act, P.slope), v1 = act_f(acc[rd * 4 + rh * 2 + 1], P.act, P.slope);
          __nv_bfloat16* dst = P.big + (((int64_t)in * 2 * P.sd + 2 * jd + rd) * bh + 2 * jh + rh) * bw + 2 * jw;
          *reinterpret_cast<__nv_bfloat162*>(dst) = __floats2bfloat162_rn(v0, v1);
        }
    }
    __syncthreads();                                   // plane o (slot o & 3) is free for plane o + 4
  }
}
