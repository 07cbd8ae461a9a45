// api.cu — library-level entry points, descriptor validation and geometry.
#include <cstdlib>
#include <mutex>

#include "common.cuh"

namespace b2c {

std::atomic<uint64_t> g_launches{0};
thread_local int t_last_algo = B2C_ALGO_AUTO;

int num_sms() {
  static int sms[64];
  static std::once_flag once;
  std::call_once(once, [] { for (int& s : sms) s = 0; });
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (sms[dev] == 0) {
    int v = 0;
    if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || v <= 0) v = 148;
    sms[dev] = v;
  }
  return sms[dev];
}

const Knobs& knobs() {
  static Knobs k;
  static std::once_flag once;
  std::call_once(once, [] {
    auto num = [](const char* n) { const char* e = getenv(n); return e ? atoi(e) : 0; };
    auto set = [](const char* n) { return getenv(n) != nullptr; };
    k.swap_min = num("B2C_SWAP_MIN"); k.mt1 = set("B2C_MT1"); k.no_pair = set("B2C_NO_PAIR"); k.dbg = num("B2C_DBG");
    k.stages = num("B2C_STAGES"); k.wg_mt1 = set("B2C_WG_MT1"); k.thin_unfused = set("B2C_THIN_UNFUSED");
    k.col2im_ldg = set("B2C_COL2IM_LDG");
  });
  return k;
}

static int out_extent(int in, int k, int s, int p, int d) { return (in + 2 * p - d * (k - 1) - 1) / s + 1; }

int check_desc(const b2c_conv_desc* c) {
  if (!c || !dt_ok(c->io_dtype)) return B2C_ERR_INVALID;
  const int xs[3] = {c->x.d, c->x.h, c->x.w}, ys[3] = {c->y.d, c->y.h, c->y.w};
  if (c->x.n != c->y.n || c->x.n <= 0 || c->x.c <= 0 || c->y.c <= 0) return B2C_ERR_INVALID;
  if (c->algo != B2C_ALGO_AUTO && c->algo != B2C_ALGO_DIRECT && c->algo != B2C_ALGO_TC) return B2C_ERR_INVALID;
  for (int i = 0; i < 3; i++) {
    if (c->k[i] <= 0 || c->stride[i] <= 0 || c->dil[i] <= 0 || c->pad[i] < 0) return B2C_ERR_INVALID;
    if (xs[i] <= 0 || ys[i] <= 0) return B2C_ERR_INVALID;
    if (!c->transposed) {
      if (out_extent(xs[i], c->k[i], c->stride[i], c->pad[i], c->dil[i]) != ys[i]) return B2C_ERR_INVALID;
    } else {
      const int base = (xs[i] - 1) * c->stride[i] - 2 * c->pad[i] + c->dil[i] * (c->k[i] - 1) + 1;
      const int op = ys[i] - base;   // output_padding
      if (op < 0 || (op >= c->stride[i] && op >= c->dil[i])) return B2C_ERR_INVALID;
    }
  }
  return B2C_OK;
}

Geo make_geo(const b2c_conv_desc* c) {
  Geo g;
  const b2c_shape& sm = c->transposed ? c->x : c->y;
  const b2c_shape& bg = c->transposed ? c->y : c->x;
  g.n = c->x.n;
  g.sc = sm.c; g.sd = sm.d; g.sh = sm.h; g.sw = sm.w;
  g.bc = bg.c; g.bd = bg.d; g.bh = bg.h; g.bw = bg.w;
  for (int i = 0; i < 3; i++) { g.k[i] = c->k[i]; g.s[i] = c->stride[i]; g.p[i] = c->pad[i]; g.dl[i] = c->dil[i]; }
  return g;
}

}  // namespace b2c

extern "C" int b2c_version(void) { return 1001; }
extern "C" int b2c_last_algo(void) { return b2c::t_last_algo; }
extern "C" uint64_t b2c_launch_count(void) { return b2c::g_launches.load(); }
extern "C" const char* b2c_strerror(int c) {
  switch (c) {
    case B2C_OK: return "ok";
    case B2C_ERR_INVALID: return "invalid argument";
    case B2C_ERR_UNSUPPORTED: return "unsupported configuration";
    case B2C_ERR_WORKSPACE: return "workspace too small";
    case B2C_ERR_CUDA: return "CUDA error";
    case B2C_ERR_NO_DEVICE: return "no sm_100 device";
    case B2C_ERR_STATIC_LIMIT: return "problem exceeds the static scratch limit (B2C_STATIC_SCRATCH_BYTES)";
    default: return "unknown error";
  }
}
