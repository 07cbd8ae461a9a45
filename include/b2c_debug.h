/*
 * b2c_debug.h — test / diagnostics entry points of libb2c.so.  Not part of the
 * drop-in boundary (include/b2c.h); nothing on the product path calls these.
 */
#ifndef B2C_DEBUG_H_
#define B2C_DEBUG_H_
#include "b2c.h"
#ifdef __cplusplus
extern "C" {
#endif

/* Synchronises the device, returns and clears the sticky error word of the
 * tensor-core kernels (bit0: an mbarrier wait ran out of its cycle budget;
 * bit31: the CUDA context reported an error).                               */
unsigned b2c_debug_tc_error(void);

/* Executes, ON THE HOST in fp32, the tensor-core plan the library would build
 * for (desc, pass in {FWD, DGRAD}): sub-problems, phase maps, taps, offsets.
 * `in`, `w_packed` ([tap][n_out][k_in]) and `out` are host fp32 NDHWC arrays.
 * Lets the CPU test-suite verify the integer lowering (stride phases,
 * sub-pixel decomposition, OOB padding) without a GPU.
 * info[0..7] = {nsubs, nmaps, ntaps, box_n, box_d, box_h, box_w, m_tiles}.   */
int b2c_debug_plan_run(const b2c_conv_desc* desc, int pass, const float* in,
                       const float* w_packed, float* out, int32_t* info);

/* Experiment switches of the tensor-core kernels, read from the ENVIRONMENT once, when the library first plans a
 * launch (csrc/common.cuh `Knobs`; changing them later in the process has no effect).  They change tile plans, hence
 * summation order and speed, never the arithmetic contract; the product path sets none of them.
 *   B2C_SWAP_MIN=n      smallest Cout that takes the swapped (channels on UMMA M) epilogue (default 0: all <= 128)
 *   B2C_MT1=1           never use the 256 x 256 single-accumulator tiles
 *   B2C_NO_PAIR=1       no CTA pairs (tcgen05 cta_group::2)
 *   B2C_STAGES=n        cap of the shared-memory pipeline depth
 *   B2C_WG_MT1=1        weight-gradient kernel with one M sub-tile per CTA
 *   B2C_THIN_UNFUSED=1  one-channel layers through im2col + GEMM + col2im instead of the fused gather kernels
 *   B2C_COL2IM_LDG=1    k4 / s2 one-channel col2im with global loads (round-1 kernel) instead of the TMA-staged one
 *   B2C_DBG=bits        timing experiments that BREAK the result (1 no TMA store, 2 no tcgen05.ld, 4 no st.shared,
 *                       8 no MMAs, 16 no loads)                                                                  */

#ifdef __cplusplus
}
#endif
#endif
