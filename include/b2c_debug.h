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

#ifdef __cplusplus
}
#endif
#endif
