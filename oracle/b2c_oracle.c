/*
 * b2c_oracle.c — TEST INFRASTRUCTURE ONLY.  CPU restatement (plain C) of the
 * arithmetic on the reference's conv3d hot path, exported under the SAME C ABI
 * as the product library (include/b2c.h) but operating on HOST pointers.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load this library; the product package
 * (3d-gan_b200/b200conv3d) never does.
 *
 * Where the algorithm lives: the reference (julschoen/3D-GAN) has no arithmetic
 * of its own on this path; every FLOP is a PyTorch layer call:
 *   nn.Conv3d / F.conv3d              dcgan.py:161-181, utils.py:8-14,
 *                                     FID_ResNet.py:9-16,70-75,119-125,
 *                                     stylegan.py:168-170
 *   nn.ConvTranspose3d                dcgan.py:47-67, stylegan.py:170
 *   nn.BatchNorm3d (training mode)    dcgan.py:48-64, utils.py:62-65,
 *                                     FID_ResNet.py:71-76,127
 *   nn.LayerNorm([C,D,H,W])           dcgan.py:162-178
 *   ReLU / LeakyReLU / Tanh           dcgan.py:49,68,163, utils.py:68
 * The dependency is torch (unpinned by the reference; torch 2.11.0+cu128 in
 * this image).  This file restates torch's documented definitions of those
 * ops.  Parity pinning: the reference ships NO tests or golden vectors
 * (SURVEY.md §4), so this oracle is pinned against outputs of the reference's
 * own modules imported in the build container (tests/golden/ npz files, generated
 * by oracle/make_golden.py) and against torch CPU fp32 ops
 * (tests/test_oracle.py).
 *
 * Accumulation is in double so the oracle is order-independent.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include "../include/b2c.h"

typedef uint16_t bf16_t;

static inline float bf16_to_f(bf16_t v) {
  uint32_t u = ((uint32_t)v) << 16; float f; memcpy(&f, &u, 4); return f;
}
static inline bf16_t f_to_bf16(float f) {          /* round-to-nearest-even */
  uint32_t u; memcpy(&u, &f, 4);
  if ((u & 0x7fffffffu) > 0x7f800000u) return (bf16_t)((u >> 16) | 0x40);
  u += 0x7fffu + ((u >> 16) & 1u);
  return (bf16_t)(u >> 16);
}
static inline double ld(const void* p, int64_t i, int dt) {
  return dt == B2C_BF16 ? (double)bf16_to_f(((const bf16_t*)p)[i])
                        : (double)((const float*)p)[i];
}
static inline void st(void* p, int64_t i, int dt, double v) {
  if (dt == B2C_BF16) ((bf16_t*)p)[i] = f_to_bf16((float)v);
  else ((float*)p)[i] = (float)v;
}
static inline int esize(int dt) { return dt == B2C_BF16 ? 2 : 4; }
static inline int dt_ok(int dt) { return dt == B2C_BF16 || dt == B2C_F32; }

static inline double act_f(double z, int act, double slope) {
  switch (act) {
    case B2C_ACT_RELU:  return z > 0 ? z : 0.0;
    case B2C_ACT_LRELU: return z > 0 ? z : slope * z;
    case B2C_ACT_TANH:  return tanh(z);
    default:            return z;
  }
}
/* derivative w.r.t. the pre-activation z */
static inline double act_dz(double z, int act, double slope) {
  switch (act) {
    case B2C_ACT_RELU:  return z > 0 ? 1.0 : 0.0;
    case B2C_ACT_LRELU: return z > 0 ? 1.0 : slope;
    case B2C_ACT_TANH:  { double t = tanh(z); return 1.0 - t * t; }
    default:            return 1.0;
  }
}
/* derivative expressed from the output y = act(z) */
static inline double act_dy(double y, int act, double slope) {
  switch (act) {
    case B2C_ACT_RELU:  return y > 0 ? 1.0 : 0.0;
    case B2C_ACT_LRELU: return y > 0 ? 1.0 : slope;
    case B2C_ACT_TANH:  return 1.0 - y * y;
    default:            return 1.0;
  }
}

int b2c_version(void) { return 100; }
int b2c_last_algo(void) { return B2C_ALGO_DIRECT; }
uint64_t b2c_launch_count(void) { return 0; }
const char* b2c_strerror(int c) {
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

/* ---- descriptor validation (same rules as the product library) ------- */
static int out_extent(int in, int k, int s, int p, int d) {
  return (in + 2 * p - d * (k - 1) - 1) / s + 1;
}
static int check_desc(const b2c_conv_desc* c) {
  if (!c || !dt_ok(c->io_dtype)) return B2C_ERR_INVALID;
  const int xs[3] = {c->x.d, c->x.h, c->x.w}, ys[3] = {c->y.d, c->y.h, c->y.w};
  if (c->x.n != c->y.n || c->x.n <= 0 || c->x.c <= 0 || c->y.c <= 0) return B2C_ERR_INVALID;
  for (int i = 0; i < 3; i++) {
    if (c->k[i] <= 0 || c->stride[i] <= 0 || c->dil[i] <= 0 || c->pad[i] < 0) return B2C_ERR_INVALID;
    if (xs[i] <= 0 || ys[i] <= 0) return B2C_ERR_INVALID;
    if (!c->transposed) {
      if (out_extent(xs[i], c->k[i], c->stride[i], c->pad[i], c->dil[i]) != ys[i]) return B2C_ERR_INVALID;
    } else {
      int base = (xs[i] - 1) * c->stride[i] - 2 * c->pad[i] + c->dil[i] * (c->k[i] - 1) + 1;
      int op = ys[i] - base;                 /* output_padding */
      if (op < 0 || (op >= c->stride[i] && op >= c->dil[i])) return B2C_ERR_INVALID;
    }
  }
  return B2C_OK;
}
static inline int ntaps(const b2c_conv_desc* c) { return c->k[0] * c->k[1] * c->k[2]; }

/* ---- filter packing -------------------------------------------------- */
size_t b2c_filter_packed_bytes(const b2c_conv_desc* c, int pass) {
  (void)pass;
  if (check_desc(c)) return 0;
  return (size_t)ntaps(c) * c->x.c * c->y.c * esize(c->io_dtype);
}
/* torch filter is [A][B][taps]; packed is [tap][n_out][k_in].
 * n_out is the channel count of the tensor the pass produces.            */
int b2c_filter_pack(const b2c_conv_desc* c, int pass, const float* w, void* wp, b2c_stream s) {
  (void)s;
  int rc = check_desc(c); if (rc) return rc;
  if (pass != B2C_PASS_FWD && pass != B2C_PASS_DGRAD) return B2C_ERR_INVALID;
  const int T = ntaps(c);
  const int A = c->transposed ? c->x.c : c->y.c;   /* torch dim 0 */
  const int B = c->transposed ? c->y.c : c->x.c;   /* torch dim 1 */
  const int n_out = pass == B2C_PASS_FWD ? c->y.c : c->x.c;
  /* conv fwd / convT dgrad produce torch-dim-0 channels; the other two dim 1 */
  const int out_is_A = (pass == B2C_PASS_FWD) != (c->transposed != 0);
  const int k_in = out_is_A ? B : A;
  for (int t = 0; t < T; t++)
    for (int o = 0; o < n_out; o++)
      for (int i = 0; i < k_in; i++) {
        int a = out_is_A ? o : i, b = out_is_A ? i : o;
        st(wp, ((int64_t)t * n_out + o) * k_in + i, c->io_dtype,
           w[((int64_t)a * B + b) * T + t]);
      }
  return B2C_OK;
}

/* ---- convolution cores ----------------------------------------------- *
 * "gather" form  (conv fwd, convT dgrad):
 *    out[n,o,co] = sum_{k,ci} in[n, o*s + k*dil - p, ci] * Wp[k][co][ci]
 * "scatter" form (convT fwd, conv dgrad), written as a gather over taps:
 *    out[n,i,co] = sum_{k,ci} [s | (i+p-k*dil)] in[n,(i+p-k*dil)/s,ci] * Wp[k][co][ci]
 * `small` is the tensor on the strided-down side of the layer, `big` the
 * other one; the geometry (k,s,p,dil) always maps small index o to big
 * index o*s + k*dil - p.                                                   */
typedef struct { int n, c, d, h, w; } shp;
static shp mk(b2c_shape s) { shp r = {s.n, s.c, s.d, s.h, s.w}; return r; }

static void conv_gather(const b2c_conv_desc* c, shp big, shp small, const void* in /*big*/,
                        const void* wp, void* out /*small*/, const b2c_fusion* f) {
  const int dt = c->io_dtype, KD = c->k[0], KH = c->k[1], KW = c->k[2];
  const int64_t vox = (int64_t)small.n * small.d * small.h * small.w;
#pragma omp parallel for schedule(static)
  for (int64_t v = 0; v < vox; v++) {
    int ow = (int)(v % small.w); int64_t r = v / small.w;
    int oh = (int)(r % small.h); r /= small.h;
    int od = (int)(r % small.d); int n = (int)(r / small.d);
    for (int co = 0; co < small.c; co++) {
      double acc = 0.0;
      for (int kd = 0; kd < KD; kd++) {
        int id = od * c->stride[0] + kd * c->dil[0] - c->pad[0];
        if (id < 0 || id >= big.d) continue;
        for (int kh = 0; kh < KH; kh++) {
          int ih = oh * c->stride[1] + kh * c->dil[1] - c->pad[1];
          if (ih < 0 || ih >= big.h) continue;
          for (int kw = 0; kw < KW; kw++) {
            int iw = ow * c->stride[2] + kw * c->dil[2] - c->pad[2];
            if (iw < 0 || iw >= big.w) continue;
            int t = (kd * KH + kh) * KW + kw;
            int64_t ib = ((((int64_t)n * big.d + id) * big.h + ih) * big.w + iw) * big.c;
            int64_t wb = ((int64_t)t * small.c + co) * big.c;
            for (int ci = 0; ci < big.c; ci++) acc += ld(in, ib + ci, dt) * ld(wp, wb + ci, dt);
          }
        }
      }
      if (f) {
        if (f->bias) acc += (double)f->bias[co];
        acc = act_f(acc, f->act, (double)f->slope);
      }
      st(out, v * small.c + co, dt, acc);
    }
  }
}

static void conv_scatter(const b2c_conv_desc* c, shp big, shp small, const void* in /*small*/,
                         const void* wp, void* out /*big*/, const b2c_fusion* f) {
  const int dt = c->io_dtype, KD = c->k[0], KH = c->k[1], KW = c->k[2];
  const int64_t vox = (int64_t)big.n * big.d * big.h * big.w;
#pragma omp parallel for schedule(static)
  for (int64_t v = 0; v < vox; v++) {
    int iw = (int)(v % big.w); int64_t r = v / big.w;
    int ih = (int)(r % big.h); r /= big.h;
    int id = (int)(r % big.d); int n = (int)(r / big.d);
    for (int co = 0; co < big.c; co++) {
      double acc = 0.0;
      for (int kd = 0; kd < KD; kd++) {
        int nd = id + c->pad[0] - kd * c->dil[0];
        if (nd < 0 || nd % c->stride[0]) continue;
        int od = nd / c->stride[0]; if (od >= small.d) continue;
        for (int kh = 0; kh < KH; kh++) {
          int nh = ih + c->pad[1] - kh * c->dil[1];
          if (nh < 0 || nh % c->stride[1]) continue;
          int oh = nh / c->stride[1]; if (oh >= small.h) continue;
          for (int kw = 0; kw < KW; kw++) {
            int nw = iw + c->pad[2] - kw * c->dil[2];
            if (nw < 0 || nw % c->stride[2]) continue;
            int ow = nw / c->stride[2]; if (ow >= small.w) continue;
            int t = (kd * KH + kh) * KW + kw;
            int64_t ib = ((((int64_t)n * small.d + od) * small.h + oh) * small.w + ow) * small.c;
            int64_t wb = ((int64_t)t * big.c + co) * small.c;
            for (int ci = 0; ci < small.c; ci++) acc += ld(in, ib + ci, dt) * ld(wp, wb + ci, dt);
          }
        }
      }
      if (f) {
        if (f->bias) acc += (double)f->bias[co];
        acc = act_f(acc, f->act, (double)f->slope);
      }
      st(out, v * big.c + co, dt, acc);
    }
  }
}

size_t b2c_conv3d_workspace(const b2c_conv_desc* c, int pass) { (void)c; (void)pass; return 0; }

int b2c_conv3d_fwd(const b2c_conv_desc* c, const void* x, const void* wp, void* y,
                   const b2c_fusion* f, void* ws, size_t wsb, b2c_stream s) {
  (void)ws; (void)wsb; (void)s;
  int rc = check_desc(c); if (rc) return rc;
  if (!x || !wp || !y) return B2C_ERR_INVALID;
  if (!c->transposed) conv_gather(c, mk(c->x), mk(c->y), x, wp, y, f);
  else                conv_scatter(c, mk(c->y), mk(c->x), x, wp, y, f);
  /* b2c_fusion.stat_*: sums of the STORED output for the normalisation layer that follows - what
   * F.batch_norm / F.layer_norm compute first (torch.nn.functional.batch_norm, layer_norm) */
  if (f && f->stat_kind != B2C_STAT_NONE) {
    if ((f->stat_kind != B2C_STAT_CHANNEL && f->stat_kind != B2C_STAT_SAMPLE) || !f->stat_sum || !f->stat_sqsum)
      return B2C_ERR_INVALID;
    const int64_t vox = (int64_t)c->y.d * c->y.h * c->y.w, C = c->y.c, N = c->y.n;
    if (f->stat_kind == B2C_STAT_CHANNEL) {
#pragma omp parallel for schedule(static)
      for (int64_t ch = 0; ch < C; ch++) {
        double a = 0, b = 0;
        for (int64_t r = 0; r < N * vox; r++) { double v = ld(y, r * C + ch, c->io_dtype); a += v; b += v * v; }
        f->stat_sum[ch] = (float)a; f->stat_sqsum[ch] = (float)b;
      }
    } else {
#pragma omp parallel for schedule(static)
      for (int64_t n = 0; n < N; n++) {
        double a = 0, b = 0;
        for (int64_t e = 0; e < vox * C; e++) { double v = ld(y, n * vox * C + e, c->io_dtype); a += v; b += v * v; }
        f->stat_sum[n] = (float)a; f->stat_sqsum[n] = (float)b;
      }
    }
  }
  return B2C_OK;
}
size_t b2c_conv3d_stats_workspace(const b2c_conv_desc* c, int kind) { (void)c; (void)kind; return 0; }

int b2c_conv3d_dgrad(const b2c_conv_desc* c, const void* dy, const void* wp, void* dx,
                     void* ws, size_t wsb, b2c_stream s) {
  (void)ws; (void)wsb; (void)s;
  int rc = check_desc(c); if (rc) return rc;
  if (!dy || !wp || !dx) return B2C_ERR_INVALID;
  if (!c->transposed) conv_scatter(c, mk(c->x), mk(c->y), dy, wp, dx, NULL);
  else                conv_gather(c, mk(c->y), mk(c->x), dy, wp, dx, NULL);
  return B2C_OK;
}

/* dW[a][b][k] = sum_{n,o} small[n,o,a] * big[n, o*s + k*dil - p, b]
 * conv : small = dy (a = y.c), big = x  (b = x.c)  -> dW [y.c][x.c][k]
 * convT: small = x  (a = x.c), big = dy (b = y.c)  -> dW [x.c][y.c][k]     */
int b2c_conv3d_wgrad(const b2c_conv_desc* c, const void* x, const void* dy, float* dw,
                     float* dbias, void* ws, size_t wsb, b2c_stream s) {
  (void)ws; (void)wsb; (void)s;
  int rc = check_desc(c); if (rc) return rc;
  if (!x || !dy) return B2C_ERR_INVALID;
  const int dt = c->io_dtype, KD = c->k[0], KH = c->k[1], KW = c->k[2], T = KD * KH * KW;
  shp small = c->transposed ? mk(c->x) : mk(c->y), big = c->transposed ? mk(c->y) : mk(c->x);
  const void* S = c->transposed ? x : dy; const void* Bg = c->transposed ? dy : x;
  if (dw) {
#pragma omp parallel for collapse(2) schedule(static)
    for (int a = 0; a < small.c; a++)
      for (int t = 0; t < T; t++) {
        int kd = t / (KH * KW), kh = (t / KW) % KH, kw = t % KW;
        double* acc = (double*)calloc(big.c, sizeof(double));
        for (int n = 0; n < small.n; n++)
          for (int od = 0; od < small.d; od++) {
            int id = od * c->stride[0] + kd * c->dil[0] - c->pad[0];
            if (id < 0 || id >= big.d) continue;
            for (int oh = 0; oh < small.h; oh++) {
              int ih = oh * c->stride[1] + kh * c->dil[1] - c->pad[1];
              if (ih < 0 || ih >= big.h) continue;
              for (int ow = 0; ow < small.w; ow++) {
                int iw = ow * c->stride[2] + kw * c->dil[2] - c->pad[2];
                if (iw < 0 || iw >= big.w) continue;
                double sv = ld(S, ((((int64_t)n * small.d + od) * small.h + oh) * small.w + ow) * small.c + a, dt);
                int64_t bb = ((((int64_t)n * big.d + id) * big.h + ih) * big.w + iw) * big.c;
                for (int b = 0; b < big.c; b++) acc[b] += sv * ld(Bg, bb + b, dt);
              }
            }
          }
        for (int b = 0; b < big.c; b++) dw[((int64_t)a * big.c + b) * T + t] = (float)acc[b];
        free(acc);
      }
  }
  if (dbias) {
    const int64_t vox = (int64_t)c->y.n * c->y.d * c->y.h * c->y.w;
    for (int co = 0; co < c->y.c; co++) {
      double a = 0.0;
      for (int64_t v = 0; v < vox; v++) a += ld(dy, v * c->y.c + co, dt);
      dbias[co] = (float)a;
    }
  }
  return B2C_OK;
}

/* ---- layout ----------------------------------------------------------- */
int b2c_transpose(const void* src, void* dst, int64_t batch, int64_t rows, int64_t cols,
                  int sdt, int ddt, b2c_stream s) {
  (void)s;
  if (!src || !dst || !dt_ok(sdt) || !dt_ok(ddt) || batch < 0 || rows < 0 || cols < 0) return B2C_ERR_INVALID;
#pragma omp parallel for schedule(static)
  for (int64_t b = 0; b < batch; b++)
    for (int64_t i = 0; i < rows; i++)
      for (int64_t j = 0; j < cols; j++)
        st(dst, (b * cols + j) * rows + i, ddt, ld(src, (b * rows + i) * cols + j, sdt));
  return B2C_OK;
}
int b2c_cast(const void* src, void* dst, int64_t n, int sdt, int ddt, b2c_stream s) {
  (void)s;
  if (!src || !dst || !dt_ok(sdt) || !dt_ok(ddt) || n < 0) return B2C_ERR_INVALID;
  for (int64_t i = 0; i < n; i++) st(dst, i, ddt, ld(src, i, sdt));
  return B2C_OK;
}

/* ---- BatchNorm3d (training mode) -------------------------------------- */
size_t b2c_stats_workspace(int64_t rows, int64_t C) { (void)rows; (void)C; return 0; }
int b2c_channel_stats(const void* x, int64_t rows, int64_t C, int dt, float* sum, float* sq,
                      void* ws, size_t wsb, b2c_stream s) {
  (void)ws; (void)wsb; (void)s;
  if (!x || !sum || !sq || !dt_ok(dt) || rows <= 0 || C <= 0) return B2C_ERR_INVALID;
#pragma omp parallel for schedule(static)
  for (int64_t c = 0; c < C; c++) {
    double a = 0, b = 0;
    for (int64_t r = 0; r < rows; r++) { double v = ld(x, r * C + c, dt); a += v; b += v * v; }
    sum[c] = (float)a; sq[c] = (float)b;
  }
  return B2C_OK;
}
/* Host stand-in of the allocation-free variant (no reference counterpart; the product
 * version keeps its partials in a static device array - the oracle needs no scratch). */
int b2c_channel_stats_static(const void* x, int64_t rows, int64_t C, int dt, float* sum, float* sq,
                             b2c_stream s) {
  return b2c_channel_stats(x, rows, C, dt, sum, sq, NULL, 0, s);
}
int b2c_bn_finalize(const float* sum, const float* sq, int64_t count, int64_t C, float eps,
                    float momentum, const float* gamma, const float* beta, float* mean,
                    float* invstd, float* scale, float* shift, float* rmean, float* rvar,
                    b2c_stream s) {
  (void)s;
  if (!sum || !sq || !mean || !invstd || !scale || !shift || count <= 0 || C <= 0) return B2C_ERR_INVALID;
  for (int64_t c = 0; c < C; c++) {
    double m = (double)sum[c] / count;
    double var = (double)sq[c] / count - m * m; if (var < 0) var = 0;
    double is = 1.0 / sqrt(var + (double)eps);
    double g = gamma ? gamma[c] : 1.0, b = beta ? beta[c] : 0.0;
    mean[c] = (float)m; invstd[c] = (float)is;
    scale[c] = (float)(g * is); shift[c] = (float)(b - m * g * is);
    if (rmean) rmean[c] = (float)((1.0 - momentum) * rmean[c] + momentum * m);
    if (rvar) {
      double unb = count > 1 ? var * count / (double)(count - 1) : var;
      rvar[c] = (float)((1.0 - momentum) * rvar[c] + momentum * unb);
    }
  }
  return B2C_OK;
}
int b2c_bn_act_fwd(const void* x, void* y, const void* res, int64_t rows, int64_t C, int dt,
                   const float* scale, const float* shift, int act, float slope, b2c_stream s) {
  (void)s;
  if (!x || !y || !scale || !shift || !dt_ok(dt)) return B2C_ERR_INVALID;
#pragma omp parallel for schedule(static)
  for (int64_t r = 0; r < rows; r++)
    for (int64_t c = 0; c < C; c++) {
      double z = ld(x, r * C + c, dt) * scale[c] + shift[c];
      if (res) z += ld(res, r * C + c, dt);
      st(y, r * C + c, dt, act_f(z, act, slope));
    }
  return B2C_OK;
}
int b2c_bn_act_bwd_reduce(const void* x, const void* dy, const void* res, int64_t rows, int64_t C,
                          int dt, const float* mean, const float* invstd, const float* scale,
                          const float* shift, int act, float slope, float* s1, float* s2,
                          void* ws, size_t wsb, b2c_stream s) {
  (void)ws; (void)wsb; (void)s;
  if (!x || !dy || !mean || !invstd || !scale || !shift || !s1 || !s2 || !dt_ok(dt)) return B2C_ERR_INVALID;
#pragma omp parallel for schedule(static)
  for (int64_t c = 0; c < C; c++) {
    double a = 0, b = 0;
    for (int64_t r = 0; r < rows; r++) {
      double xv = ld(x, r * C + c, dt);
      double z = xv * scale[c] + shift[c]; if (res) z += ld(res, r * C + c, dt);
      double g = ld(dy, r * C + c, dt) * act_dz(z, act, slope);
      a += g; b += g * (xv - mean[c]) * invstd[c];
    }
    s1[c] = (float)a; s2[c] = (float)b;
  }
  return B2C_OK;
}
int b2c_bn_act_bwd_apply(const void* x, const void* dy, const void* res, void* dx, void* dres,
                         int64_t rows, int64_t C, int dt, const float* mean, const float* invstd,
                         const float* scale, const float* shift, const float* s1, const float* s2,
                         int act, float slope, b2c_stream s) {
  (void)s;
  if (!x || !dy || !dx || !mean || !invstd || !scale || !shift || !s1 || !s2 || !dt_ok(dt)) return B2C_ERR_INVALID;
#pragma omp parallel for schedule(static)
  for (int64_t r = 0; r < rows; r++)
    for (int64_t c = 0; c < C; c++) {
      double xv = ld(x, r * C + c, dt);
      double z = xv * scale[c] + shift[c]; if (res) z += ld(res, r * C + c, dt);
      double g = ld(dy, r * C + c, dt) * act_dz(z, act, slope);
      double xh = (xv - mean[c]) * invstd[c];
      st(dx, r * C + c, dt, scale[c] * (g - s1[c] / (double)rows - xh * s2[c] / (double)rows));
      if (dres) st(dres, r * C + c, dt, g);
    }
  return B2C_OK;
}

/* ---- LayerNorm over the whole sample + activation ---------------------- */
size_t b2c_ln_workspace(int64_t N, int64_t L) { (void)N; (void)L; return 0; }
/* LayerNorm forward from per-sample sums delivered by the producing convolution (same formula) */
int b2c_ln_act_fwd_stats(const void* x, void* y, int64_t N, int64_t L, int dt, const float* gamma,
                         const float* beta, float eps, int act, float slope, const float* sum,
                         const float* sqsum, float* mean, float* rstd, b2c_stream s) {
  (void)s;
  if (!x || !y || !sum || !sqsum || !mean || !rstd || !dt_ok(dt) || N <= 0 || L <= 0) return B2C_ERR_INVALID;
#pragma omp parallel for schedule(static)
  for (int64_t n = 0; n < N; n++) {
    double m = (double)sum[n] / L, var = (double)sqsum[n] / L - m * m; if (var < 0) var = 0;
    double rs = 1.0 / sqrt(var + (double)eps);
    mean[n] = (float)m; rstd[n] = (float)rs;
    for (int64_t e = 0; e < L; e++) {
      double z = (ld(x, n * L + e, dt) - m) * rs;
      z = z * (gamma ? gamma[e] : 1.0) + (beta ? beta[e] : 0.0);
      st(y, n * L + e, dt, act_f(z, act, slope));
    }
  }
  return B2C_OK;
}
int b2c_ln_act_fwd(const void* x, void* y, int64_t N, int64_t L, int dt, const float* gamma,
                   const float* beta, float eps, int act, float slope, float* mean, float* rstd,
                   void* ws, size_t wsb, b2c_stream s) {
  (void)ws; (void)wsb; (void)s;
  if (!x || !y || !mean || !rstd || !dt_ok(dt) || N <= 0 || L <= 0) return B2C_ERR_INVALID;
#pragma omp parallel for schedule(static)
  for (int64_t n = 0; n < N; n++) {
    double a = 0, b = 0;
    for (int64_t e = 0; e < L; e++) { double v = ld(x, n * L + e, dt); a += v; b += v * v; }
    double m = a / L, var = b / L - m * m; if (var < 0) var = 0;
    double rs = 1.0 / sqrt(var + (double)eps);
    mean[n] = (float)m; rstd[n] = (float)rs;
    for (int64_t e = 0; e < L; e++) {
      double z = (ld(x, n * L + e, dt) - m) * rs;
      z = z * (gamma ? gamma[e] : 1.0) + (beta ? beta[e] : 0.0);
      st(y, n * L + e, dt, act_f(z, act, slope));
    }
  }
  return B2C_OK;
}
int b2c_ln_act_bwd(const void* x, const void* dy, void* dx, int64_t N, int64_t L, int dt,
                   const float* gamma, const float* beta, const float* mean, const float* rstd,
                   int act, float slope, float* dgamma, float* dbeta, void* ws, size_t wsb,
                   b2c_stream s) {
  (void)ws; (void)wsb; (void)s;
  if (!x || !dy || !mean || !rstd || !dt_ok(dt) || N <= 0 || L <= 0) return B2C_ERR_INVALID;
  double* dg = dgamma ? (double*)calloc(L, sizeof(double)) : NULL;
  double* db = dbeta ? (double*)calloc(L, sizeof(double)) : NULL;
  for (int64_t n = 0; n < N; n++) {
    double m = mean[n], rs = rstd[n], a = 0, b = 0;
    for (int64_t e = 0; e < L; e++) {
      double xh = (ld(x, n * L + e, dt) - m) * rs, ga = gamma ? gamma[e] : 1.0;
      double z = xh * ga + (beta ? beta[e] : 0.0);
      double g0 = ld(dy, n * L + e, dt) * act_dz(z, act, slope);
      if (dg) dg[e] += g0 * xh;
      if (db) db[e] += g0;
      double g = g0 * ga; a += g; b += g * xh;
    }
    if (dx)
      for (int64_t e = 0; e < L; e++) {
        double xh = (ld(x, n * L + e, dt) - m) * rs, ga = gamma ? gamma[e] : 1.0;
        double z = xh * ga + (beta ? beta[e] : 0.0);
        double g = ld(dy, n * L + e, dt) * act_dz(z, act, slope) * ga;
        st(dx, n * L + e, dt, rs * (g - a / L - xh * b / L));
      }
  }
  if (dg) { for (int64_t e = 0; e < L; e++) dgamma[e] = (float)dg[e]; free(dg); }
  if (db) { for (int64_t e = 0; e < L; e++) dbeta[e] = (float)db[e]; free(db); }
  return B2C_OK;
}

/* ---- pointwise ---------------------------------------------------------- */
int b2c_act_fwd(const void* x, void* y, int64_t n, int dt, int act, float slope, b2c_stream s) {
  (void)s;
  if (!x || !y || !dt_ok(dt) || n < 0) return B2C_ERR_INVALID;
  for (int64_t i = 0; i < n; i++) st(y, i, dt, act_f(ld(x, i, dt), act, slope));
  return B2C_OK;
}
int b2c_act_bwd(const void* y, const void* dy, void* dx, int64_t n, int dt, int act, float slope,
                b2c_stream s) {
  (void)s;
  if (!y || !dy || !dx || !dt_ok(dt) || n < 0) return B2C_ERR_INVALID;
  for (int64_t i = 0; i < n; i++) st(dx, i, dt, ld(dy, i, dt) * act_dy(ld(y, i, dt), act, slope));
  return B2C_OK;
}

/* ---- x2 resampling (utils.py:76-80 nearest x2 + channel drop; nn.AvgPool3d(2)) ----------- */
int b2c_upsample2x_fwd(const void* x, void* y, int64_t n, int d, int h, int w, int cx, int c0, int c, int dt,
                       b2c_stream s) {
  (void)s;
  if (!x || !y || !dt_ok(dt) || c0 < 0 || c0 + c > cx) return B2C_ERR_INVALID;
  for (int64_t in = 0; in < n; in++)
    for (int od = 0; od < 2 * d; od++)
      for (int oh = 0; oh < 2 * h; oh++)
        for (int ow = 0; ow < 2 * w; ow++)
          for (int k = 0; k < c; k++) {
            int64_t src = ((((int64_t)in * d + od / 2) * h + oh / 2) * w + ow / 2) * cx + c0 + k;
            int64_t dst = ((((int64_t)in * 2 * d + od) * 2 * h + oh) * 2 * w + ow) * c + k;
            st(y, dst, dt, ld(x, src, dt));
          }
  return B2C_OK;
}
int b2c_upsample2x_bwd(const void* dy, void* dx, int64_t n, int d, int h, int w, int cx, int c0, int c, int dt,
                       b2c_stream s) {
  (void)s;
  if (!dy || !dx || !dt_ok(dt) || c0 < 0 || c0 + c > cx) return B2C_ERR_INVALID;
  for (int64_t in = 0; in < n; in++)
    for (int id = 0; id < d; id++)
      for (int ih = 0; ih < h; ih++)
        for (int iw = 0; iw < w; iw++)
          for (int k = 0; k < cx; k++) {
            double acc = 0.0;
            if (k >= c0 && k < c0 + c)
              for (int q = 0; q < 8; q++) {
                int64_t src = ((((int64_t)in * 2 * d + 2 * id + (q >> 2)) * 2 * h + 2 * ih + ((q >> 1) & 1)) * 2 * w +
                               2 * iw + (q & 1)) * c + (k - c0);
                acc += ld(dy, src, dt);
              }
            st(dx, ((((int64_t)in * d + id) * h + ih) * w + iw) * cx + k, dt, acc);
          }
  return B2C_OK;
}
int b2c_avgpool2_fwd(const void* x, void* y, int64_t n, int d, int h, int w, int c, int dt, b2c_stream s) {
  (void)s;
  if (!x || !y || !dt_ok(dt) || d < 2 || h < 2 || w < 2) return B2C_ERR_INVALID;
  int od = d / 2, oh = h / 2, ow = w / 2;
  for (int64_t in = 0; in < n; in++)
    for (int a = 0; a < od; a++)
      for (int b = 0; b < oh; b++)
        for (int e = 0; e < ow; e++)
          for (int k = 0; k < c; k++) {
            double acc = 0.0;
            for (int q = 0; q < 8; q++)
              acc += ld(x, ((((int64_t)in * d + 2 * a + (q >> 2)) * h + 2 * b + ((q >> 1) & 1)) * w + 2 * e + (q & 1)) * c + k, dt);
            st(y, ((((int64_t)in * od + a) * oh + b) * ow + e) * c + k, dt, acc * 0.125);
          }
  return B2C_OK;
}
int b2c_avgpool2_bwd(const void* dy, void* dx, int64_t n, int d, int h, int w, int c, int dt, b2c_stream s) {
  (void)s;
  if (!dy || !dx || !dt_ok(dt) || d < 2 || h < 2 || w < 2) return B2C_ERR_INVALID;
  int od = d / 2, oh = h / 2, ow = w / 2;
  for (int64_t in = 0; in < n; in++)
    for (int a = 0; a < d; a++)
      for (int b = 0; b < h; b++)
        for (int e = 0; e < w; e++)
          for (int k = 0; k < c; k++) {
            double v = 0.0;
            if (a / 2 < od && b / 2 < oh && e / 2 < ow)
              v = 0.125 * ld(dy, ((((int64_t)in * od + a / 2) * oh + b / 2) * ow + e / 2) * c + k, dt);
            st(dx, ((((int64_t)in * d + a) * h + b) * w + e) * c + k, dt, v);
          }
  return B2C_OK;
}

/* ---- depthwise FIR resampling (stylegan.py:172-200) ---------------------------------------- */
static int fir_ok(const b2c_fir_desc* d, const float* taps, const void* a, const void* b) {
  if (!d || !taps || !a || !b || !dt_ok(d->dtype)) return 0;
  int T = d->k[0] * d->k[1] * d->k[2];
  return d->n >= 0 && d->c > 0 && d->up >= 1 && d->down >= 1 && T >= 1 && T <= 64;
}
int b2c_fir3d_fwd(const b2c_fir_desc* d, const float* taps, const void* x, void* y, b2c_stream s) {
  (void)s;
  if (!fir_ok(d, taps, x, y)) return B2C_ERR_INVALID;
  for (int64_t n = 0; n < d->n; n++)
    for (int od = 0; od < d->out[0]; od++)
      for (int oh = 0; oh < d->out[1]; oh++)
        for (int ow = 0; ow < d->out[2]; ow++)
          for (int c = 0; c < d->c; c++) {
            double acc = 0.0;
            for (int a = 0; a < d->k[0]; a++)
              for (int b = 0; b < d->k[1]; b++)
                for (int e = 0; e < d->k[2]; e++) {
                  int jd = od * d->down + a - d->pad[0], jh = oh * d->down + b - d->pad[1], jw = ow * d->down + e - d->pad[2];
                  if (jd < 0 || jh < 0 || jw < 0 || jd % d->up || jh % d->up || jw % d->up) continue;
                  jd /= d->up; jh /= d->up; jw /= d->up;
                  if (jd >= d->in[0] || jh >= d->in[1] || jw >= d->in[2]) continue;
                  acc += (double)taps[(a * d->k[1] + b) * d->k[2] + e] *
                         ld(x, (((n * d->in[0] + jd) * d->in[1] + jh) * d->in[2] + jw) * d->c + c, d->dtype);
                }
            st(y, (((n * d->out[0] + od) * d->out[1] + oh) * d->out[2] + ow) * d->c + c, d->dtype, acc);
          }
  return B2C_OK;
}
int b2c_fir3d_bwd(const b2c_fir_desc* d, const float* taps, const void* dy, void* dx, b2c_stream s) {
  (void)s;
  if (!fir_ok(d, taps, dy, dx)) return B2C_ERR_INVALID;
  for (int64_t n = 0; n < d->n; n++)
    for (int id = 0; id < d->in[0]; id++)
      for (int ih = 0; ih < d->in[1]; ih++)
        for (int iw = 0; iw < d->in[2]; iw++)
          for (int c = 0; c < d->c; c++) {
            double acc = 0.0;
            for (int a = 0; a < d->k[0]; a++)
              for (int b = 0; b < d->k[1]; b++)
                for (int e = 0; e < d->k[2]; e++) {
                  int td = id * d->up + d->pad[0] - a, th = ih * d->up + d->pad[1] - b, tw = iw * d->up + d->pad[2] - e;
                  if (td < 0 || th < 0 || tw < 0 || td % d->down || th % d->down || tw % d->down) continue;
                  td /= d->down; th /= d->down; tw /= d->down;
                  if (td >= d->out[0] || th >= d->out[1] || tw >= d->out[2]) continue;
                  acc += (double)taps[(a * d->k[1] + b) * d->k[2] + e] *
                         ld(dy, (((n * d->out[0] + td) * d->out[1] + th) * d->out[2] + tw) * d->c + c, d->dtype);
                }
            st(dx, (((n * d->in[0] + id) * d->in[1] + ih) * d->in[2] + iw) * d->c + c, d->dtype, acc);
          }
  return B2C_OK;
}

/* ---- style modulation glue (stylegan.py:396-444 live branch + bias_act, :522-531) ---------- */
size_t b2c_mod_workspace(int64_t N, int64_t V, int64_t C) { (void)N; (void)V; (void)C; return 16; }

int b2c_modulate_fwd(const void* x, const float* sc, void* y, int64_t N, int64_t V, int64_t C, int dt, b2c_stream s) {
  (void)s;
  if (!x || !sc || !y || !dt_ok(dt)) return B2C_ERR_INVALID;
  for (int64_t n = 0; n < N; n++)
    for (int64_t v = 0; v < V; v++)
      for (int64_t c = 0; c < C; c++) {
        int64_t i = (n * V + v) * C + c;
        st(y, i, dt, ld(x, i, dt) * (double)sc[n * C + c]);          /* stylegan.py:427 */
      }
  return B2C_OK;
}

int b2c_modulate_bwd(const void* x, const void* gy, const float* sc, void* dx, float* ds, int64_t N, int64_t V,
                     int64_t C, int dt, void* ws, size_t wsb, b2c_stream s) {
  (void)s; (void)ws; (void)wsb;
  if (!x || !gy || !sc || !dx || !ds || !dt_ok(dt)) return B2C_ERR_INVALID;
  for (int64_t n = 0; n < N; n++)
    for (int64_t c = 0; c < C; c++) {
      double acc = 0.0;
      for (int64_t v = 0; v < V; v++) {
        int64_t i = (n * V + v) * C + c;
        double g = ld(gy, i, dt);
        acc += g * ld(x, i, dt);
        st(dx, i, dt, g * (double)sc[n * C + c]);
      }
      ds[n * C + c] = (float)acc;
    }
  return B2C_OK;
}

int b2c_mod_epilogue_fwd(const void* y, const float* d, const float* field, int64_t field_n, const float* strength,
                         const float* bias, void* out, int64_t N, int64_t V, int64_t C, int dt, int act, float slope,
                         b2c_stream s) {
  (void)s;
  if (!y || !out || !dt_ok(dt) || (field && (!strength || (field_n != 1 && field_n != N)))) return B2C_ERR_INVALID;
  for (int64_t n = 0; n < N; n++)
    for (int64_t v = 0; v < V; v++) {
      double nz = field ? (double)*strength * (double)field[(field_n > 1 ? n * V : 0) + v] : 0.0;   /* :522-524 */
      for (int64_t c = 0; c < C; c++) {
        int64_t i = (n * V + v) * C + c;
        double z = ld(y, i, dt) * (d ? (double)d[n * C + c] : 1.0) + nz + (bias ? (double)bias[c] : 0.0);   /* :429-432, :531 */
        st(out, i, dt, act_f(z, act, slope));
      }
    }
  return B2C_OK;
}

int b2c_mod_epilogue_bwd(const void* y, const void* out, const void* gout, const float* d, const float* field,
                         int64_t field_n, void* dy, float* dd, float* dbias, float* dstrength, int64_t N, int64_t V,
                         int64_t C, int dt, int act, float slope, void* ws, size_t wsb, b2c_stream s) {
  (void)s; (void)ws; (void)wsb;
  if (!out || !gout || !dy || !dt_ok(dt) || (dd && (!d || !y)) || (field && field_n != 1 && field_n != N))
    return B2C_ERR_INVALID;
  double* accb = (double*)calloc((size_t)C, sizeof(double));
  double accs = 0.0;
  if (!accb) return B2C_ERR_INVALID;
  for (int64_t n = 0; n < N; n++)
    for (int64_t c = 0; c < C; c++) {
      double accd = 0.0;
      for (int64_t v = 0; v < V; v++) {
        int64_t i = (n * V + v) * C + c;
        double o = ld(out, i, dt), dz;
        switch (act) {                       /* derivative from the output, as the product kernels take it */
          case B2C_ACT_RELU:  dz = o > 0 ? 1.0 : 0.0; break;
          case B2C_ACT_LRELU: dz = o > 0 ? 1.0 : (double)slope; break;
          case B2C_ACT_TANH:  dz = 1.0 - o * o; break;
          default:            dz = 1.0;
        }
        double g = ld(gout, i, dt) * dz;
        accb[c] += g;
        if (dd) accd += g * ld(y, i, dt);
        if (field) accs += g * (double)field[(field_n > 1 ? n * V : 0) + v];
        st(dy, i, dt, g * (d ? (double)d[n * C + c] : 1.0));
      }
      if (dd) dd[n * C + c] = (float)accd;
    }
  if (dbias) for (int64_t c = 0; c < C; c++) dbias[c] = (float)accb[c];
  if (field && dstrength) *dstrength = (float)accs;
  free(accb);
  return B2C_OK;
}

/* ---- AdaIN (+ trilinear x2) of the v1-like generator (stylegan.py:931-959, 1081) ------------ */
int b2c_instance_stats(const void* x, float* sum, float* sqsum, int64_t N, int64_t V, int64_t C, int dt, void* ws,
                       size_t wsb, b2c_stream s) {
  (void)s; (void)ws; (void)wsb;
  if (!x || !sum || !sqsum || !dt_ok(dt)) return B2C_ERR_INVALID;
  for (int64_t n = 0; n < N; n++)
    for (int64_t c = 0; c < C; c++) {
      double a = 0.0, b = 0.0;
      for (int64_t v = 0; v < V; v++) { double t = ld(x, (n * V + v) * C + c, dt); a += t; b += t * t; }
      sum[n * C + c] = (float)a; sqsum[n * C + c] = (float)b;
    }
  return B2C_OK;
}

static inline int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

/* torch upsample_trilinear3d, align_corners=False, scale 2: src = (dst + .5) / 2 - .5 clamped at 0 */
static void tri_src(int o, int n, int* i0, int* i1, double* w1) {
  double src = (o + 0.5) / 2.0 - 0.5;
  if (src < 0) src = 0;
  *i0 = (int)src;
  *i1 = *i0 + (*i0 < n - 1 ? 1 : 0);
  *w1 = src - *i0;
}

int b2c_adain_fwd(const void* x, const float* a, const float* b, void* y, int64_t N, int d, int h, int w, int64_t C,
                  int dt, int up, b2c_stream s) {
  (void)s;
  if (!x || !a || !b || !y || !dt_ok(dt)) return B2C_ERR_INVALID;
  int f = up ? 2 : 1, od = d * f, oh = h * f, ow = w * f;
  for (int64_t n = 0; n < N; n++)
    for (int zd = 0; zd < od; zd++)
      for (int zh = 0; zh < oh; zh++)
        for (int zw = 0; zw < ow; zw++)
          for (int64_t c = 0; c < C; c++) {
            double v;
            if (!up) v = ld(x, (((n * d + zd) * h + zh) * (int64_t)w + zw) * C + c, dt);
            else {
              int d0, d1, h0, h1, w0, w1; double fd, fh, fw;
              tri_src(zd, d, &d0, &d1, &fd); tri_src(zh, h, &h0, &h1, &fh); tri_src(zw, w, &w0, &w1, &fw);
              v = 0.0;
              for (int q = 0; q < 8; q++) {
                int dd = (q & 4) ? d1 : d0, hh = (q & 2) ? h1 : h0, ww = (q & 1) ? w1 : w0;
                double wt = ((q & 4) ? fd : 1 - fd) * ((q & 2) ? fh : 1 - fh) * ((q & 1) ? fw : 1 - fw);
                v += wt * ld(x, (((n * d + dd) * h + hh) * (int64_t)w + ww) * C + c, dt);
              }
            }
            st(y, (((n * od + zd) * oh + zh) * (int64_t)ow + zw) * C + c, dt, v * (double)a[n * C + c] + (double)b[n * C + c]);
          }
  return B2C_OK;
}

int b2c_adain_bwd_reduce(const void* x, const void* gy, void* t, float* r1, float* r2, int64_t N, int d, int h, int w,
                         int64_t C, int dt, int up, void* ws, size_t wsb, b2c_stream s) {
  (void)s; (void)ws; (void)wsb;
  if (!x || !gy || !r1 || !r2 || (up && !t) || !dt_ok(dt)) return B2C_ERR_INVALID;
  int64_t V = (int64_t)d * h * w;
  double* tt = (double*)calloc((size_t)(V * C), sizeof(double));
  if (!tt) return B2C_ERR_INVALID;
  for (int64_t n = 0; n < N; n++) {
    memset(tt, 0, (size_t)(V * C) * sizeof(double));
    if (!up) {
      for (int64_t i = 0; i < V * C; i++) tt[i] = ld(gy, n * V * C + i, dt);
    } else {                                   /* scatter form of the adjoint */
      int od = 2 * d, oh = 2 * h, ow = 2 * w;
      for (int zd = 0; zd < od; zd++)
        for (int zh = 0; zh < oh; zh++)
          for (int zw = 0; zw < ow; zw++) {
            int d0, d1, h0, h1, w0, w1; double fd, fh, fw;
            tri_src(zd, d, &d0, &d1, &fd); tri_src(zh, h, &h0, &h1, &fh); tri_src(zw, w, &w0, &w1, &fw);
            for (int q = 0; q < 8; q++) {
              int dd = (q & 4) ? d1 : d0, hh = (q & 2) ? h1 : h0, ww = (q & 1) ? w1 : w0;
              double wt = ((q & 4) ? fd : 1 - fd) * ((q & 2) ? fh : 1 - fh) * ((q & 1) ? fw : 1 - fw);
              for (int64_t c = 0; c < C; c++)
                tt[(((int64_t)dd * h + hh) * w + ww) * C + c] +=
                    wt * ld(gy, (((n * od + zd) * oh + zh) * (int64_t)ow + zw) * C + c, dt);
            }
          }
    }
    for (int64_t c = 0; c < C; c++) {
      double a1 = 0.0, a2 = 0.0;
      for (int64_t v = 0; v < V; v++) {
        double tv = tt[v * C + c];
        if (up) { st(t, (n * V + v) * C + c, dt, tv); }
        a1 += tv; a2 += tv * ld(x, (n * V + v) * C + c, dt);
      }
      r1[n * C + c] = (float)a1; r2[n * C + c] = (float)a2;
    }
  }
  free(tt);
  return B2C_OK;
}

int b2c_adain_bwd_apply(const void* x, const void* t, const float* ct, const float* cx, const float* c0, void* dx,
                        int64_t N, int64_t V, int64_t C, int dt, b2c_stream s) {
  (void)s;
  if (!x || !t || !ct || !cx || !c0 || !dx || !dt_ok(dt)) return B2C_ERR_INVALID;
  for (int64_t n = 0; n < N; n++)
    for (int64_t v = 0; v < V; v++)
      for (int64_t c = 0; c < C; c++) {
        int64_t i = (n * V + v) * C + c, k = n * C + c;
        st(dx, i, dt, (double)ct[k] * ld(t, i, dt) + (double)cx[k] * ld(x, i, dt) + (double)c0[k]);
      }
  return B2C_OK;
}

/* ---- nn.MaxPool3d(k, stride, pad) on NDHWC (FID_ResNet.py:129) ------------------------------- */
static int mp_out(int i, int k, int s, int p) { return (i + 2 * p - k) / s + 1; }
int b2c_maxpool3d_fwd(const void* x, void* y, int64_t n, int d, int h, int w, int c, int k, int stride, int pad, int dt,
                      b2c_stream s) {
  (void)s;
  if (!x || !y || !dt_ok(dt) || k < 1 || stride < 1 || pad < 0 || 2 * pad > k) return B2C_ERR_INVALID;
  int od = mp_out(d, k, stride, pad), oh = mp_out(h, k, stride, pad), ow = mp_out(w, k, stride, pad);
  if (od <= 0 || oh <= 0 || ow <= 0) return B2C_ERR_INVALID;
  for (int64_t in = 0; in < n; in++)
    for (int a = 0; a < od; a++)
      for (int b = 0; b < oh; b++)
        for (int e = 0; e < ow; e++)
          for (int ch = 0; ch < c; ch++) {
            double m = -INFINITY;
            for (int kd = 0; kd < k; kd++)
              for (int kh = 0; kh < k; kh++)
                for (int kw = 0; kw < k; kw++) {
                  int jd = a * stride + kd - pad, jh = b * stride + kh - pad, jw = e * stride + kw - pad;
                  if (jd < 0 || jd >= d || jh < 0 || jh >= h || jw < 0 || jw >= w) continue;
                  double t = ld(x, ((((int64_t)in * d + jd) * h + jh) * w + jw) * c + ch, dt);
                  if (t > m) m = t;
                }
            st(y, ((((int64_t)in * od + a) * oh + b) * ow + e) * c + ch, dt, m);
          }
  return B2C_OK;
}
int b2c_maxpool3d_bwd(const void* x, const void* dy, void* dx, int64_t n, int d, int h, int w, int c, int k, int stride,
                      int pad, int dt, b2c_stream s) {
  (void)s;
  if (!x || !dy || !dx || !dt_ok(dt) || k < 1 || stride < 1 || pad < 0 || 2 * pad > k) return B2C_ERR_INVALID;
  int od = mp_out(d, k, stride, pad), oh = mp_out(h, k, stride, pad), ow = mp_out(w, k, stride, pad);
  if (od <= 0 || oh <= 0 || ow <= 0) return B2C_ERR_INVALID;
  int64_t total = n * d * h * w * c;
  double* acc = (double*)calloc((size_t)total, sizeof(double));
  if (!acc) return B2C_ERR_INVALID;
  for (int64_t in = 0; in < n; in++)            /* scatter form: dy goes to the first maximum of its window */
    for (int a = 0; a < od; a++)
      for (int b = 0; b < oh; b++)
        for (int e = 0; e < ow; e++)
          for (int ch = 0; ch < c; ch++) {
            double m = -INFINITY; int64_t arg = -1;
            for (int kd = 0; kd < k; kd++)
              for (int kh = 0; kh < k; kh++)
                for (int kw = 0; kw < k; kw++) {
                  int jd = a * stride + kd - pad, jh = b * stride + kh - pad, jw = e * stride + kw - pad;
                  if (jd < 0 || jd >= d || jh < 0 || jh >= h || jw < 0 || jw >= w) continue;
                  int64_t i = ((((int64_t)in * d + jd) * h + jh) * w + jw) * c + ch;
                  double t = ld(x, i, dt);
                  if (t > m) { m = t; arg = i; }
                }
            if (arg >= 0) acc[arg] += ld(dy, ((((int64_t)in * od + a) * oh + b) * ow + e) * c + ch, dt);
          }
  for (int64_t i = 0; i < total; i++) st(dx, i, dt, acc[i]);
  free(acc);
  return B2C_OK;
}

/* ---- W-window gather (include/b2c.h b2c_window_w): the im2col of F.conv3d restricted to the W axis and to
 * F-voxel groups - y[r, j, v*C + c] = x[r, F*j + v - pad, c]; inverse: the adjoint overlap-add ---- */
int b2c_window_w(const void* src, void* dst, int64_t rows, int w, int c, int f, int pad, int dt, int inverse,
                 b2c_stream s) {
  (void)s;
  if (!src || !dst || !dt_ok(dt) || rows < 0 || w <= 0 || c <= 0 || f <= 0 || pad < 0 || w % f != 0) return B2C_ERR_INVALID;
  const int win = f + 2 * pad, groups = w / f;
  if (!inverse) {
#pragma omp parallel for schedule(static)
    for (int64_t r = 0; r < rows; r++)
      for (int j = 0; j < groups; j++)
        for (int v = 0; v < win; v++) {
          const int wi = f * j + v - pad;
          for (int ch = 0; ch < c; ch++)
            st(dst, ((r * groups + j) * (int64_t)win + v) * c + ch, dt,
               (wi >= 0 && wi < w) ? ld(src, (r * w + wi) * (int64_t)c + ch, dt) : 0.0);
        }
  } else {
#pragma omp parallel for schedule(static)
    for (int64_t r = 0; r < rows; r++)
      for (int wi = 0; wi < w; wi++)
        for (int ch = 0; ch < c; ch++) {
          double a = 0;
          for (int j = 0; j < groups; j++) {
            const int v = wi - f * j + pad;
            if (v >= 0 && v < win) a += ld(src, ((r * groups + j) * (int64_t)win + v) * c + ch, dt);
          }
          st(dst, (r * w + wi) * (int64_t)c + ch, dt, a);
        }
  }
  return B2C_OK;
}

/* ---- depth-to-space by 2 on NDHWC (include/b2c.h; phase-merged form of F.conv_transpose3d, stylegan.py:170) ---- */
int b2c_depth_to_space2(const void* src, void* dst, int64_t n, int jd, int jh, int jw, int c, int od, int oh, int ow,
                        int dt, int inverse, b2c_stream s) {
  (void)s;
  if (!src || !dst || !dt_ok(dt) || n < 0 || jd <= 0 || jh <= 0 || jw <= 0 || c <= 0 || od <= 0 || oh <= 0 || ow <= 0 ||
      od > 2 * jd || oh > 2 * jh || ow > 2 * jw)
    return B2C_ERR_INVALID;
  for (int64_t in = 0; in < n; in++)
    for (int a = 0; a < jd; a++)
      for (int b = 0; b < jh; b++)
        for (int e = 0; e < jw; e++)
          for (int r = 0; r < 8; r++) {
            int o0 = 2 * a + (r >> 2), o1 = 2 * b + ((r >> 1) & 1), o2 = 2 * e + (r & 1);
            int inside = o0 < od && o1 < oh && o2 < ow;
            for (int ch = 0; ch < c; ch++) {
              int64_t zi = (((((int64_t)in * jd + a) * jh + b) * jw + e) * 8 + r) * c + ch;
              int64_t yi = ((((int64_t)in * od + o0) * oh + o1) * ow + o2) * c + ch;
              if (!inverse) { if (inside) st(dst, yi, dt, ld(src, zi, dt)); }
              else st(dst, zi, dt, inside ? ld(src, yi, dt) : 0.0);
            }
          }
  return B2C_OK;
}

/* ---- gradient averaging over peer memory (include/b2c.h b2c_peer_*): host stand-in.  The "ranks" are buffers of ONE
 * process; a call does rank `rank`'s share of the two-shot scheme - slice `rank` of every buffer is summed in rank
 * order (fp32, like the kernel), scaled by 1 / world and written back into every buffer - so calling it once per rank
 * (any order: the slices are disjoint) averages the buckets.  No flags are needed on the host. ---- */
int b2c_peer_alloc(size_t bytes, void** ptr, unsigned char* handle) {
  if (!ptr || !handle || bytes == 0) return B2C_ERR_INVALID;
  void* p = calloc(1, bytes);
  if (!p) return B2C_ERR_CUDA;
  memset(handle, 0, B2C_PEER_HANDLE_BYTES);
  memcpy(handle, &p, sizeof(p));              /* same-process "handle": the address itself */
  *ptr = p;
  return B2C_OK;
}
int b2c_peer_open(const unsigned char* handle, void** ptr) {
  if (!ptr || !handle) return B2C_ERR_INVALID;
  memcpy(ptr, handle, sizeof(*ptr));
  return *ptr ? B2C_OK : B2C_ERR_INVALID;
}
int b2c_peer_close(void* ptr) { return ptr ? B2C_OK : B2C_ERR_INVALID; }
int b2c_peer_free(void* ptr) {
  if (!ptr) return B2C_ERR_INVALID;
  free(ptr);
  return B2C_OK;
}
int b2c_peer_allreduce_avg(void* const* bufs, void* const* flags, int rank, int world, int64_t count, uint32_t seq,
                           int ctas, b2c_stream s) {
  (void)s;
  if (!bufs || !flags || world < 1 || world > B2C_PEER_MAX_WORLD || rank < 0 || rank >= world || count < 0 || seq == 0 ||
      seq > 0x3fffffffu || ctas < 1 || ctas > B2C_PEER_MAX_CTAS)
    return B2C_ERR_INVALID;
  for (int i = 0; i < world; i++)
    if (!bufs[i] || !flags[i]) return B2C_ERR_INVALID;
  if (world == 1 || count == 0) return B2C_OK;
  const int64_t per = ((count + world - 1) / world + 3) & ~(int64_t)3;
  const int64_t lo = rank * per, hi = lo + per < count ? lo + per : count;
  const float scale = 1.f / (float)world;
  for (int64_t i = lo; i < hi; i++) {
    float a = ((const float*)bufs[0])[i];
    for (int r = 1; r < world; r++) a += ((const float*)bufs[r])[i];
    a *= scale;
    for (int r = 0; r < world; r++) ((float*)bufs[r])[i] = a;
  }
  return B2C_OK;
}
unsigned b2c_peer_error(void) { return 0; }
