// conv_api.cu — b2c_conv3d_{workspace,fwd,dgrad,wgrad}: validation and dispatch
// between the tcgen05 implicit-GEMM path (bf16) and the direct CUDA-core path
// (fp32, or shapes the tensor-core path does not take).
#include "direct_conv.cuh"
#include "tc_conv.cuh"
#include "thin.cuh"

using namespace b2c;

namespace {

bool tc_allowed(const b2c_conv_desc* c) { return c->io_dtype == B2C_BF16 && c->algo != B2C_ALGO_DIRECT; }

// ConvTranspose3d on a 1x1x1 input with stride-independent geometry (dcgan.py:47,
// z -> 4^3): every output voxel sees exactly one tap, so the layer is the plain
// GEMM  y[n, (tap,co)] = x[n,:] . Wp[tap][co][:]  — a 1x1x1 conv with taps*Cout
// output channels over the SAME packed filter memory.
bool is_point_convT(const b2c_conv_desc* c) {
  if (!c->transposed) return false;
  if (c->x.d != 1 || c->x.h != 1 || c->x.w != 1) return false;
  for (int i = 0; i < 3; i++) if (c->pad[i] != 0) return false;
  return c->y.d == c->dil[0] * (c->k[0] - 1) + 1 && c->y.h == c->dil[1] * (c->k[1] - 1) + 1 &&
         c->y.w == c->dil[2] * (c->k[2] - 1) + 1 && c->dil[0] == 1 && c->dil[1] == 1 && c->dil[2] == 1;
}

// `stats` (forward only): filled by the kernel that can accumulate them in its epilogue (nblocks > 0 afterwards);
// left untouched (nblocks == 0) by the others - the caller then falls back to one pass over the output.
int fwd_like(const b2c_conv_desc* c, bool scatter, const void* in, const void* w, void* out, const float* bias,
             int act, float slope, void* ws, size_t wsb, cudaStream_t st, const TcStats* stats = nullptr) {
  const Geo g = make_geo(c);
  Geo tg;
  if (const int tm = thin_mode(c, g, &tg)) {
    t_last_algo = B2C_ALGO_TC;
    // the fused one-channel gather kernel accumulates per-sample sums (dcgan.py:161-162 Conv(1 -> ndf) -> LayerNorm)
    const bool st_ok = stats && stats->kind == 2 && tm == 1 && !scatter && tg.n == c->y.n &&
                       (int64_t)tg.sd * tg.sh * tg.sw >= 4;
    return scatter != (tm == 2) ? thin_scatter(tg, in, w, out, bias, act, slope, ws, wsb, st)
                                : thin_gather(tg, in, w, out, bias, act, slope, ws, wsb, st, st_ok ? stats : nullptr);
  }
  if (tc_allowed(c)) {
    ConvPlan plan;
    int rc;
    int bias_mod = 0;
    if (scatter && c->transposed && is_point_convT(c)) {
      Geo pg = g;                       // small = y viewed as 1^3 with taps*Cout channels, big = x
      const int T = g.taps();
      pg.sc = T * c->y.c; pg.sd = pg.sh = pg.sw = 1;
      pg.bc = c->x.c; pg.bd = pg.bh = pg.bw = 1;
      for (int i = 0; i < 3; i++) { pg.k[i] = 1; pg.s[i] = 1; pg.p[i] = 0; pg.dl[i] = 1; }
      rc = plan_gather(pg, &plan);
      bias_mod = c->y.c;
    } else {
      rc = scatter ? plan_scatter(g, &plan) : plan_gather(g, &plan);
    }
    if (rc == B2C_OK) {
      t_last_algo = B2C_ALGO_TC;
      const bool st_ok = stats && stats->kind && tc_conv_stats_supported(plan, stats->kind, bias_mod);
      return tc_conv_launch(plan, in, w, out, bias, bias_mod, act, slope, st, st_ok ? stats : nullptr);
    }
    if (c->algo == B2C_ALGO_TC) return B2C_ERR_UNSUPPORTED;
  } else if (c->algo == B2C_ALGO_TC) {
    return B2C_ERR_UNSUPPORTED;
  }
  t_last_algo = B2C_ALGO_DIRECT;
  return scatter ? direct_scatter(g, c->io_dtype, in, w, out, bias, act, slope, st)
                 : direct_gather(g, c->io_dtype, in, w, out, bias, act, slope, st);
}

int stat_len_of(const b2c_conv_desc* c, int kind) { return kind == B2C_STAT_CHANNEL ? c->y.c : c->y.n; }

}  // namespace

extern "C" size_t b2c_conv3d_stats_workspace(const b2c_conv_desc* c, int kind) {
  if (check_desc(c) || (kind != B2C_STAT_CHANNEL && kind != B2C_STAT_SAMPLE)) return 0;
  const int64_t rows = (int64_t)c->y.n * c->y.d * c->y.h * c->y.w;
  size_t fused = 2 * tc_conv_stats_bytes(stat_len_of(c, kind));        // one partial row per CTA (<= 2 per SM)
  size_t unfused = kind == B2C_STAT_CHANNEL ? channel_stats2_workspace(rows, c->y.c)
                                            : sample_stats_workspace(c->y.n, rows / c->y.n * c->y.c);
  return fused > unfused ? fused : unfused;
}

namespace {
size_t align256(size_t v) { return (v + 255) & ~(size_t)255; }
// filter-gradient part of the wgrad workspace; the dbias scratch follows it (256-byte aligned)
size_t wgrad_core_ws(const b2c_conv_desc* c, const Geo& g) {
  Geo tg;
  if (thin_mode(c, g, &tg)) return thin_workspace(tg, 2);
  return tc_allowed(c) ? tc_wgrad_workspace(g) : 0;
}
int64_t y_rows(const b2c_conv_desc* c) { return (int64_t)c->y.n * c->y.d * c->y.h * c->y.w; }
}  // namespace

extern "C" size_t b2c_conv3d_workspace(const b2c_conv_desc* c, int pass) {
  if (check_desc(c)) return 0;
  const Geo g = make_geo(c);
  if (pass == B2C_PASS_WGRAD) return align256(wgrad_core_ws(c, g)) + channel_sum_workspace(y_rows(c), c->y.c);
  Geo tg;
  if (const int tm = thin_mode(c, g, &tg)) {
    const bool scatter = (pass == B2C_PASS_FWD) == (c->transposed != 0);
    return thin_workspace(tg, scatter != (tm == 2) ? 1 : 0);
  }
  return 0;
}

extern "C" int b2c_conv3d_fwd(const b2c_conv_desc* c, const void* x, const void* wp, void* y, const b2c_fusion* f,
                              void* ws, size_t wsb, b2c_stream s) {
  int rc = check_desc(c);
  if (rc) return rc;
  if (!x || !wp || !y) return B2C_ERR_INVALID;
  const float* bias = f ? f->bias : nullptr;
  const int act = f ? f->act : B2C_ACT_NONE;
  const float slope = f ? f->slope : 0.f;
  if (act < B2C_ACT_NONE || act > B2C_ACT_TANH) return B2C_ERR_INVALID;
  cudaStream_t st = static_cast<cudaStream_t>(s);
  const int kind = f ? f->stat_kind : B2C_STAT_NONE;
  if (kind == B2C_STAT_NONE) return fwd_like(c, c->transposed != 0, x, wp, y, bias, act, slope, ws, wsb, st);
  // statistics of the stored output for the BatchNorm3d / LayerNorm that follows: the tail of the workspace holds
  // the per-CTA partial rows (or the scratch of the one-pass fall-back)
  if ((kind != B2C_STAT_CHANNEL && kind != B2C_STAT_SAMPLE) || !f->stat_sum || !f->stat_sqsum) return B2C_ERR_INVALID;
  if (kind == B2C_STAT_SAMPLE && c->y.n > 65535) return B2C_ERR_UNSUPPORTED;
  const size_t core = align256(b2c_conv3d_workspace(c, B2C_PASS_FWD)), extra = b2c_conv3d_stats_workspace(c, kind);
  if (!ws || wsb < core + extra) return B2C_ERR_WORKSPACE;
  float* part = reinterpret_cast<float*>(static_cast<uint8_t*>(ws) + core);
  const int len = stat_len_of(c, kind);
  TcStats ts{kind, len, part, extra, 0};
  rc = fwd_like(c, c->transposed != 0, x, wp, y, bias, act, slope, ws, core, st, &ts);
  if (rc) return rc;
  if (ts.nblocks > 0) return stat_finalize(part, ts.nblocks, len, f->stat_sum, f->stat_sqsum, st);
  const int64_t rows = (int64_t)c->y.n * c->y.d * c->y.h * c->y.w;
  return kind == B2C_STAT_CHANNEL
             ? channel_stats2(y, rows, c->y.c, c->io_dtype, f->stat_sum, f->stat_sqsum, part, extra, st)
             : sample_stats(y, c->y.n, rows / c->y.n * c->y.c, c->io_dtype, f->stat_sum, f->stat_sqsum, part, extra, st);
}

extern "C" int b2c_conv3d_dgrad(const b2c_conv_desc* c, const void* dy, const void* wp, void* dx, void* ws,
                                size_t wsb, b2c_stream s) {
  int rc = check_desc(c);
  if (rc) return rc;
  if (!dy || !wp || !dx) return B2C_ERR_INVALID;
  // dgrad of a conv is the scatter form; dgrad of a convT is the gather form.
  const Geo g = make_geo(c);
  const bool scatter = !c->transposed;
  cudaStream_t st = static_cast<cudaStream_t>(s);
  Geo tg;
  if (const int tm = thin_mode(c, g, &tg)) {
    t_last_algo = B2C_ALGO_TC;
    return scatter != (tm == 2) ? thin_scatter(tg, dy, wp, dx, nullptr, B2C_ACT_NONE, 0.f, ws, wsb, st)
                                : thin_gather(tg, dy, wp, dx, nullptr, B2C_ACT_NONE, 0.f, ws, wsb, st);
  }
  if (tc_allowed(c)) {
    ConvPlan plan;
    rc = scatter ? plan_scatter(g, &plan) : plan_gather(g, &plan);
    if (rc == B2C_OK) {
      t_last_algo = B2C_ALGO_TC;
      return tc_conv_launch(plan, dy, wp, dx, nullptr, 0, B2C_ACT_NONE, 0.f, st);
    }
    if (c->algo == B2C_ALGO_TC) return B2C_ERR_UNSUPPORTED;
  } else if (c->algo == B2C_ALGO_TC) {
    return B2C_ERR_UNSUPPORTED;
  }
  t_last_algo = B2C_ALGO_DIRECT;
  return scatter ? direct_scatter(g, c->io_dtype, dy, wp, dx, nullptr, B2C_ACT_NONE, 0.f, st)
                 : direct_gather(g, c->io_dtype, dy, wp, dx, nullptr, B2C_ACT_NONE, 0.f, st);
}

extern "C" int b2c_conv3d_wgrad(const b2c_conv_desc* c, const void* x, const void* dy, float* dw, float* dbias,
                                void* ws, size_t wsb, b2c_stream s) {
  int rc = check_desc(c);
  if (rc) return rc;
  if (!x || !dy) return B2C_ERR_INVALID;
  cudaStream_t st = static_cast<cudaStream_t>(s);
  const Geo g = make_geo(c);
  const void* small = c->transposed ? x : dy;
  const void* big = c->transposed ? dy : x;
  if (dw) {
    bool done = false;
    Geo tg;
    if (const int tm = thin_mode(c, g, &tg)) {
      t_last_algo = B2C_ALGO_TC;
      rc = tm == 2 ? thin_wgrad(tg, big, small, dw, ws, wsb, st, true)
                   : thin_wgrad(tg, small, big, dw, ws, wsb, st, false);
      if (rc) return rc;
      done = true;
    } else if (tc_allowed(c) && tc_wgrad_supported(g)) {
      t_last_algo = B2C_ALGO_TC;
      rc = tc_wgrad_launch(g, small, big, dw, ws, wsb, st);
      if (rc) return rc;
      done = true;
    } else if (c->algo == B2C_ALGO_TC) {
      return B2C_ERR_UNSUPPORTED;
    }
    if (!done) {
      t_last_algo = B2C_ALGO_DIRECT;
      rc = direct_wgrad(g, c->io_dtype, small, big, dw, st);
      if (rc) return rc;
    }
  }
  if (dbias) {
    const size_t core = align256(wgrad_core_ws(c, g));
    if (!ws || wsb < core) return B2C_ERR_WORKSPACE;
    rc = channel_sum(dy, y_rows(c), c->y.c, c->io_dtype, dbias, static_cast<uint8_t*>(ws) + core, wsb - core, st);
    if (rc) return rc;
  }
  return B2C_OK;
}
