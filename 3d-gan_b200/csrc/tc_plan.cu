// tc_plan.cpp — see tc_plan.h.
#include "tc_plan.h"

#include <cstdlib>

#include <string.h>

namespace b2c {

static int floordiv(int a, int b) { int q = a / b; return (a % b != 0 && ((a < 0) != (b < 0))) ? q - 1 : q; }
static int posmod(int a, int b) { int m = a % b; return m < 0 ? m + b : m; }
static int pow2ceil(int v) { int p = 1; while (p < v) p <<= 1; return p; }
static int imin(int a, int b) { return a < b ? a : b; }

static void choose_box(const int ext[4], int box[4], int rows) {
  int rem = rows;
  box[3] = imin(rem, pow2ceil(ext[3])); rem /= box[3];
  box[2] = imin(rem, pow2ceil(ext[2])); rem /= box[2];
  box[1] = imin(rem, pow2ceil(ext[1])); rem /= box[1];
  box[0] = rem;
}

static int finish(ConvPlan* p, int kin, int nout) {
  if (kin % 8 != 0 || nout % 8 != 0 || kin < 16) return B2C_ERR_UNSUPPORTED;  // 16-B TMA strides / vector stores
  p->kin = kin; p->nout = nout;
  p->kb = kin <= 32 ? 32 : kBlockK;       // 32-channel layers (utils.py hidden convs): no zero-padded half rows
  p->kchunks = (kin + p->kb - 1) / p->kb;
  int bn = ((nout + 15) / 16) * 16;
  if (bn > 256) bn = 256;
  // A UMMA with M = 128 costs about the same for N = 32 .. 128 as for N = 256, so narrow
  // output-channel counts put the CHANNELS on M (one zero-padded 128-row filter tile) and up
  // to 256 voxels on N; the epilogue transposes back to NDHWC.
  // (Short reductions - the K = 64 / 128 GEMMs of the one-channel layers - are epilogue bound and
  // keep the direct epilogue.)
  int min_taps = p->nsubs > 0 ? p->subs[0].ntaps : 0;
  for (int s = 1; s < p->nsubs; s++) if (p->subs[s].ntaps < min_taps) min_taps = p->subs[s].ntaps;
  const int swap_min = knobs().swap_min;
  p->swap = (bn <= 128 && nout >= swap_min && (int64_t)min_taps * p->kchunks >= 8) ? 1 : 0;
  if (p->swap) bn = 128;
  p->block_n = bn;
  p->tiles_n = (nout + bn - 1) / bn;
  int ext[4] = {0, 0, 0, 0};
  for (int s = 0; s < p->nsubs; s++)
    for (int i = 0; i < 4; i++) if (p->subs[s].ext[i] > ext[i]) ext[i] = p->subs[s].ext[i];
  // Narrow N tiles are L2-bandwidth bound at 128 rows: give the CTA mt sub-tiles that share
  // every B (filter) stage, keeping mt * block_n fp32 accumulator columns <= 256 so that
  // two accumulator sets (epilogue overlap) fit the 512 TMEM columns.
  int mt = bn <= 128 ? 2 : 1;
  int64_t vox = 1;
  for (int i = 0; i < 4; i++) vox *= pow2ceil(ext[i] > 0 ? ext[i] : 1);
  while (mt > 1 && vox < (int64_t)kBlockM * mt) mt >>= 1;
  // box / sub-tile split / tile counts for `m` sub-tiles per CTA; false if the box cannot be split
  auto layout = [&](int m) -> bool {
    choose_box(ext, p->box, kBlockM * m);
    int sd = 0;
    while (sd < 3 && p->box[sd] == 1) sd++;
    if (p->box[sd] < m || p->box[sd] % m != 0) return false;
    p->split_dim = sd; p->sub_ext = p->box[sd] / m; p->mt = m;
    int64_t t0 = 0;
    for (int s = 0; s < p->nsubs; s++) {
      SubDesc& sub = p->subs[s];
      int64_t nt = 1;
      for (int i = 0; i < 4; i++) { sub.tiles[i] = (sub.ext[i] + p->box[i] - 1) / p->box[i]; nt *= sub.tiles[i]; }
      sub.tile0 = (int32_t)t0;
      t0 += nt;
    }
    p->m_tiles = (int32_t)t0;
    return t0 * p->tiles_n <= 0x7fffffffLL;
  };
  for (;; mt >>= 1) {
    if (layout(mt)) break;
    if (mt == 1) return B2C_ERR_UNSUPPORTED;
  }
  // Wide N (256 columns): a 256 x 256 tile on ONE accumulator set moves 2/3 of the L2 -> SM bytes per
  // MAC of a 128 x 256 tile (the kernel is bound by those bytes and by the power they cost), at the
  // price of an exposed epilogue and half as many tiles: taken when the reduction is long and the
  // last wave of the persistent grid does not get much emptier (measured gain ~1.2x per tile).
  if (bn == 256 && mt == 1 && !p->swap && (int64_t)min_taps * p->kchunks >= 32 && vox >= 2 * kBlockM &&
      !knobs().mt1) {
    const int sms = num_sms();
    auto eff = [&](int64_t t) { return (double)t / (double)(((t + sms - 1) / sms) * sms); };
    const int64_t tiles1 = (int64_t)p->m_tiles * p->tiles_n;
    if (layout(2)) {
      const int64_t tiles2 = (int64_t)p->m_tiles * p->tiles_n;
      if (1.2 * eff(tiles2) < eff(tiles1)) layout(1);
    } else if (!layout(1)) {
      return B2C_ERR_UNSUPPORTED;
    }
  }
  for (int i = 0; i < 4; i++) if (p->box[i] > 256) return B2C_ERR_UNSUPPORTED;
  return B2C_OK;
}

int plan_gather(const Geo& g, ConvPlan* p) {
  memset(p, 0, sizeof(*p));
  if (g.s[0] * g.s[1] * g.s[2] > kMaxMaps || g.taps() > kMaxTaps) return B2C_ERR_UNSUPPORTED;
  int map_of_phase[kMaxMaps];
  for (int i = 0; i < kMaxMaps; i++) map_of_phase[i] = -1;
  SubDesc& sd = p->subs[0];
  p->nsubs = 1;
  sd.ext[0] = g.n; sd.ext[1] = g.sd; sd.ext[2] = g.sh; sd.ext[3] = g.sw;
  sd.obase = 0;
  sd.ostr[3] = g.sc; sd.ostr[2] = (int64_t)g.sw * g.sc; sd.ostr[1] = (int64_t)g.sh * g.sw * g.sc;
  sd.ostr[0] = (int64_t)g.sd * g.sh * g.sw * g.sc;
  sd.tap0 = 0;
  int nt = 0;
  for (int kd = 0; kd < g.k[0]; kd++)
    for (int kh = 0; kh < g.k[1]; kh++)
      for (int kw = 0; kw < g.k[2]; kw++) {
        const int od = kd * g.dl[0] - g.p[0], oh = kh * g.dl[1] - g.p[1], ow = kw * g.dl[2] - g.p[2];
        const int pd = posmod(od, g.s[0]), ph = posmod(oh, g.s[1]), pw = posmod(ow, g.s[2]);
        const int ed = (g.bd - pd + g.s[0] - 1) / g.s[0], eh = (g.bh - ph + g.s[1] - 1) / g.s[1],
                  ew = (g.bw - pw + g.s[2] - 1) / g.s[2];
        if (ed <= 0 || eh <= 0 || ew <= 0) continue;          // phase is empty: tap only sees padding
        const int phase = (pd * g.s[1] + ph) * g.s[2] + pw;
        if (map_of_phase[phase] < 0) {
          MapDesc& m = p->maps[p->nmaps];
          m.base = (((int64_t)pd * g.bh + ph) * g.bw + pw) * g.bc;
          m.dims[0] = g.bc; m.dims[1] = ew; m.dims[2] = eh; m.dims[3] = ed; m.dims[4] = g.n;
          m.strides[0] = 1; m.strides[1] = (int64_t)g.s[2] * g.bc; m.strides[2] = (int64_t)g.s[1] * g.bw * g.bc;
          m.strides[3] = (int64_t)g.s[0] * g.bh * g.bw * g.bc; m.strides[4] = (int64_t)g.bd * g.bh * g.bw * g.bc;
          map_of_phase[phase] = p->nmaps++;
        }
        TapDesc& t = p->taps[nt++];
        t.map = (int16_t)map_of_phase[phase];
        t.wtap = (int16_t)((kd * g.k[1] + kh) * g.k[2] + kw);
        t.off_d = (int16_t)floordiv(od, g.s[0]); t.off_h = (int16_t)floordiv(oh, g.s[1]);
        t.off_w = (int16_t)floordiv(ow, g.s[2]);
      }
  sd.ntaps = nt;
  p->ntaps_total = nt;
  return finish(p, g.bc, g.sc);
}

int plan_scatter(const Geo& g, ConvPlan* p) {
  memset(p, 0, sizeof(*p));
  if (g.s[0] * g.s[1] * g.s[2] > kMaxSubs) return B2C_ERR_UNSUPPORTED;
  MapDesc& m = p->maps[0];
  p->nmaps = 1;
  m.base = 0;
  m.dims[0] = g.sc; m.dims[1] = g.sw; m.dims[2] = g.sh; m.dims[3] = g.sd; m.dims[4] = g.n;
  m.strides[0] = 1; m.strides[1] = g.sc; m.strides[2] = (int64_t)g.sw * g.sc;
  m.strides[3] = (int64_t)g.sh * g.sw * g.sc; m.strides[4] = (int64_t)g.sd * g.sh * g.sw * g.sc;
  int nt = 0;
  for (int rd = 0; rd < g.s[0]; rd++)
    for (int rh = 0; rh < g.s[1]; rh++)
      for (int rw = 0; rw < g.s[2]; rw++) {
        const int ed = (g.bd - rd + g.s[0] - 1) / g.s[0], eh = (g.bh - rh + g.s[1] - 1) / g.s[1],
                  ew = (g.bw - rw + g.s[2] - 1) / g.s[2];
        if (ed <= 0 || eh <= 0 || ew <= 0) continue;
        SubDesc& sd = p->subs[p->nsubs++];
        sd.ext[0] = g.n; sd.ext[1] = ed; sd.ext[2] = eh; sd.ext[3] = ew;
        sd.obase = (((int64_t)rd * g.bh + rh) * g.bw + rw) * g.bc;
        sd.ostr[0] = (int64_t)g.bd * g.bh * g.bw * g.bc; sd.ostr[1] = (int64_t)g.s[0] * g.bh * g.bw * g.bc;
        sd.ostr[2] = (int64_t)g.s[1] * g.bw * g.bc; sd.ostr[3] = (int64_t)g.s[2] * g.bc;
        sd.tap0 = nt;
        for (int kd = 0; kd < g.k[0]; kd++) {
          const int nd = rd + g.p[0] - kd * g.dl[0];
          if (posmod(nd, g.s[0])) continue;
          for (int kh = 0; kh < g.k[1]; kh++) {
            const int nh = rh + g.p[1] - kh * g.dl[1];
            if (posmod(nh, g.s[1])) continue;
            for (int kw = 0; kw < g.k[2]; kw++) {
              const int nw = rw + g.p[2] - kw * g.dl[2];
              if (posmod(nw, g.s[2])) continue;
              if (nt >= kMaxTaps) return B2C_ERR_UNSUPPORTED;
              TapDesc& t = p->taps[nt++];
              t.map = 0;
              t.wtap = (int16_t)((kd * g.k[1] + kh) * g.k[2] + kw);
              t.off_d = (int16_t)floordiv(nd, g.s[0]); t.off_h = (int16_t)floordiv(nh, g.s[1]);
              t.off_w = (int16_t)floordiv(nw, g.s[2]);
            }
          }
        }
        sd.ntaps = nt - sd.tap0;
      }
  p->ntaps_total = nt;
  return finish(p, g.sc, g.bc);
}

void plan_run_host(const ConvPlan& p, const float* in, const float* w, float* out) {
  for (int s = 0; s < p.nsubs; s++) {
    const SubDesc& sd = p.subs[s];
    for (int n = 0; n < sd.ext[0]; n++)
      for (int d = 0; d < sd.ext[1]; d++)
        for (int h = 0; h < sd.ext[2]; h++)
          for (int x = 0; x < sd.ext[3]; x++) {
            float* o = out + sd.obase + n * sd.ostr[0] + d * sd.ostr[1] + h * sd.ostr[2] + x * sd.ostr[3];
            for (int co = 0; co < p.nout; co++) {
              double acc = 0.0;
              for (int t = sd.tap0; t < sd.tap0 + sd.ntaps; t++) {
                const TapDesc& tp = p.taps[t];
                const MapDesc& m = p.maps[tp.map];
                const int cw = x + tp.off_w, chh = h + tp.off_h, cd = d + tp.off_d;
                if (cw < 0 || cw >= m.dims[1] || chh < 0 || chh >= m.dims[2] || cd < 0 || cd >= m.dims[3]) continue;
                const float* ip = in + m.base + cw * m.strides[1] + chh * m.strides[2] + cd * m.strides[3] +
                                  n * m.strides[4];
                const float* wp = w + ((int64_t)tp.wtap * p.nout + co) * p.kin;
                for (int ci = 0; ci < p.kin; ci++) acc += (double)ip[ci] * wp[ci];
              }
              o[co] = (float)acc;
            }
          }
  }
}

}  // namespace b2c
