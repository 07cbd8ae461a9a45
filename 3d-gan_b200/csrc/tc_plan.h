// tc_plan.h — host-side plan of the tcgen05 implicit-GEMM convolution.
//
// Every conv / conv-transpose pass is lowered to "sub-problems": each owns an
// output lattice (a strided view of the output tensor) and a list of taps; a
// tap names the input view it reads (a TMA tensor map over a stride-phase of
// the input), a constant voxel offset and the filter slab it multiplies with.
//   gather form  (conv fwd, convT dgrad): 1 sub-problem, stride^3 input phase
//                maps — a strided conv reads dense boxes of each phase.
//   scatter form (convT fwd, conv dgrad): stride^3 sub-problems (sub-pixel
//                phases of the output), each a dense stride-1 conv over the
//                taps that land on that phase.  No zero insertion anywhere.
// Padding is never materialised: out-of-range box coordinates are zero-filled
// by TMA.  The plan is pure integer logic and is unit-tested on the CPU.
#pragma once
#include <stdint.h>

#include "common.cuh"

namespace b2c {

constexpr int kMaxMaps = 8;
constexpr int kMaxSubs = 8;
constexpr int kMaxTaps = 128;
constexpr int kBlockM = 128;     // output voxels per tile (UMMA M)
constexpr int kBlockK = 64;      // channels per pipeline stage (128 B of bf16)

struct MapDesc {                 // a 5-D strided view, element units, dim0 = channels
  int64_t base;
  int32_t dims[5];               // {C, W', H', D', N}
  int64_t strides[5];            // strides[0] == 1
};
struct TapDesc { int16_t map, wtap, off_d, off_h, off_w, _pad; };
struct SubDesc {
  int32_t ext[4];                // lattice extents {n, d, h, w}
  int32_t tiles[4];              // tile counts per lattice dim
  int64_t obase;                 // element offset of lattice origin in the output
  int64_t ostr[4];               // element strides of lattice steps {n, d, h, w}
  int32_t tap0, ntaps;
  int32_t tile0;                 // first global (M) tile index of this sub-problem
  int32_t _pad;
};
struct ConvPlan {
  int32_t nmaps, nsubs, ntaps_total;
  int32_t box[4];                // CTA tile box {bn, bd, bh, bw}, product == kBlockM * mt
  int32_t mt;                    // 128-row UMMA sub-tiles per CTA tile (1, 2 or 4); mt * block_n <= 256
  int32_t swap;                  // 1: nout <= 128 -> channels ride UMMA M (block_n = 128), voxels UMMA N = 128*mt
  int32_t split_dim, sub_ext;    // sub-tile m covers box[split_dim] range [m*sub_ext, (m+1)*sub_ext)
  int32_t kin, nout;             // contracted / produced channels
  int32_t block_n, tiles_n;      // N tile and count
  int32_t kchunks;               // ceil(kin / kb)
  int32_t kb;                    // channels per pipeline stage: kBlockK, or 32 when kin <= 32 (64-byte rows)
  int32_t m_tiles;               // total M tiles over all sub-problems
  MapDesc maps[kMaxMaps];
  SubDesc subs[kMaxSubs];
  TapDesc taps[kMaxTaps];
};

// Returns B2C_OK, or B2C_ERR_UNSUPPORTED when the shape does not fit the
// tensor-core path (caller then uses the direct kernels).
int plan_gather(const Geo& g, ConvPlan* p);    // out = small (sc), in = big (bc)
int plan_scatter(const Geo& g, ConvPlan* p);   // out = big (bc),   in = small (sc)

// Reference executor of a plan on host fp32 arrays (tests only).
void plan_run_host(const ConvPlan& p, const float* in, const float* w /*[tap][nout][kin]*/, float* out);

}  // namespace b2c
