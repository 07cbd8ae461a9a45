// debug_api.cu — see include/b2c_debug.h.
#include "../../include/b2c_debug.h"
#include "tc_conv.cuh"

using namespace b2c;

extern "C" unsigned b2c_debug_tc_error(void) { return tc_error_fetch_clear(); }

extern "C" int b2c_debug_plan_run(const b2c_conv_desc* c, int pass, const float* in, const float* w, float* out,
                                  int32_t* info) {
  int rc = check_desc(c);
  if (rc) return rc;
  if (pass != B2C_PASS_FWD && pass != B2C_PASS_DGRAD) return B2C_ERR_INVALID;
  const Geo g = make_geo(c);
  const bool scatter = (pass == B2C_PASS_FWD) == (c->transposed != 0);
  ConvPlan plan;
  rc = scatter ? plan_scatter(g, &plan) : plan_gather(g, &plan);
  if (rc) return rc;
  if (info) {
    info[0] = plan.nsubs; info[1] = plan.nmaps; info[2] = plan.ntaps_total;
    for (int i = 0; i < 4; i++) info[3 + i] = plan.box[i];
    info[7] = plan.m_tiles;
  }
  if (in && w && out) plan_run_host(plan, in, w, out);
  return B2C_OK;
}
