/*
 * b2c.h — C ABI of the B200-native conv3d hot path (libb2c.so).
 *
 * This is the drop-in boundary ("Face 2" of SURVEY.md §8(b)).  The reference
 * (julschoen/3D-GAN) has no FFI of its own: every op on its hot path is a
 * PyTorch layer call.  Each entry point below therefore cites the reference
 * call site whose PyTorch op it replaces.  The Python host side
 * (3d-gan_b200/b200conv3d) binds these with ctypes and mirrors the torch
 * nn.Module / functional API ("Face 1").
 *
 * Rules of the ABI
 *   - extern "C", plain pointers and sizes, no torch types.
 *   - All activation tensors are physically NDHWC (torch channels_last_3d),
 *     dense.  Filters arrive in torch layout (fp32) and are re-packed by
 *     b2c_filter_pack into the layout the kernels stream.
 *   - The caller owns every buffer (device memory); the library allocates no
 *     device memory, launches on the given stream and never synchronises.
 *   - Return value: 0 = OK, negative = B2C_ERR_*.  b2c_strerror() maps a code
 *     to a static string.  No C++ exception crosses the ABI.
 *   - Re-entrant: the only global state is a mutex-guarded cache of TMA
 *     descriptors encoder entry point and device attributes.
 */
#ifndef B2C_H_
#define B2C_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* cudaStream_t without pulling in cuda_runtime.h for C callers. */
typedef void* b2c_stream;

/* ---- error codes ---------------------------------------------------- */
#define B2C_OK                 0
#define B2C_ERR_INVALID       -1   /* bad argument / inconsistent shapes     */
#define B2C_ERR_UNSUPPORTED   -2   /* legal torch config we do not implement */
#define B2C_ERR_WORKSPACE     -3   /* workspace too small                    */
#define B2C_ERR_CUDA          -4   /* a CUDA runtime / driver call failed    */
#define B2C_ERR_NO_DEVICE     -5   /* no sm_100 device / driver available    */
#define B2C_ERR_STATIC_LIMIT  -6   /* problem larger than the static scratch */

/* ---- dtypes --------------------------------------------------------- */
#define B2C_BF16 0
#define B2C_F32  2

/* ---- activations (epilogues the reference models use) ---------------
 * dcgan.py:49-68 ReLU/Tanh, dcgan.py:163 LeakyReLU(0.2), dcgan.py:121
 * LeakyReLU(0.1), utils.py:68 ReLU, stylegan.py:103-113 lrelu.           */
#define B2C_ACT_NONE  0
#define B2C_ACT_RELU  1
#define B2C_ACT_LRELU 2
#define B2C_ACT_TANH  3

/* ---- passes --------------------------------------------------------- */
#define B2C_PASS_FWD   0
#define B2C_PASS_DGRAD 1
#define B2C_PASS_WGRAD 2

/* ---- kernel selection (for tests / profiling; AUTO in production) --- */
#define B2C_ALGO_AUTO   0
#define B2C_ALGO_DIRECT 1   /* CUDA-core direct conv, fp32 accumulate        */
#define B2C_ALGO_TC     2   /* tcgen05 implicit GEMM (bf16 only)             */

typedef struct { int32_t n, c, d, h, w; } b2c_shape;   /* logical NCDHW extents */

/* Describes one nn.Conv3d (transposed=0) or nn.ConvTranspose3d
 * (transposed=1) layer call.  x is the layer's forward input, y its forward
 * output, whichever pass is being computed.
 *   conv : y.d = (x.d + 2*pad - dil*(k-1) - 1)/stride + 1
 *   convT: y.d = (x.d-1)*stride - 2*pad + dil*(k-1) + output_padding + 1
 * Filter (torch layout, fp32): conv [y.c, x.c, kD,kH,kW];
 *                              convT [x.c, y.c, kD,kH,kW].   groups must be 1
 * (depthwise FIR of stylegan.py:197 has its own entry point).              */
typedef struct {
  b2c_shape x, y;
  int32_t   k[3], stride[3], pad[3], dil[3];
  int32_t   transposed;
  int32_t   io_dtype;      /* B2C_BF16 or B2C_F32; accumulation is fp32     */
  int32_t   algo;          /* B2C_ALGO_*                                    */
} b2c_conv_desc;

/* Epilogue fused into the forward kernels: bias, activation, and - for the
 * BatchNorm3d / LayerNorm that follows the convolution in the reference's
 * layer tables (dcgan.py:47-65 ConvT -> BatchNorm3d, dcgan.py:161-179
 * Conv -> LayerNorm, utils.py:62-75, FID_ResNet.py:70-76) - the statistics
 * of the STORED output y (after bias / activation, rounded to io_dtype):
 *   B2C_STAT_CHANNEL: stat_sum[c] = sum over (n,d,h,w) of y, stat_sqsum = of y^2   [y.c]
 *   B2C_STAT_SAMPLE : stat_sum[n] = sum over (c,d,h,w) of y, stat_sqsum = of y^2   [y.n]
 * accumulated in the GEMM epilogue (warp-shuffle-free shared-memory trees, one
 * partial row per CTA, fixed-order second stage: deterministic, no atomics),
 * so the normalisation layer needs no statistics pass over y.  Kernels that
 * cannot accumulate (fp32 / CUDA-core paths) compute the same sums with one
 * pass over y inside the same call.  The workspace must then hold
 * align256(b2c_conv3d_workspace(desc, FWD)) + b2c_conv3d_stats_workspace(desc, kind). */
#define B2C_STAT_NONE    0
#define B2C_STAT_CHANNEL 1
#define B2C_STAT_SAMPLE  2
typedef struct {
  const float* bias;       /* [y.c] fp32 or NULL                            */
  int32_t      act;        /* B2C_ACT_*                                     */
  float        slope;      /* LeakyReLU negative slope                      */
  int32_t      stat_kind;  /* B2C_STAT_*                                    */
  float*       stat_sum;   /* [y.c] or [y.n] fp32, written; or NULL         */
  float*       stat_sqsum; /* same shape                                    */
} b2c_fusion;

/* ---- library ------------------------------------------------------- */
int         b2c_version(void);
const char* b2c_strerror(int code);
/* Last kernel family chosen by b2c_conv3d_* on this thread (B2C_ALGO_*). */
int         b2c_last_algo(void);
/* Number of kernel launches issued by this library since process start.  */
uint64_t    b2c_launch_count(void);

/* ---- filters ------------------------------------------------------- *
 * Re-pack a torch-layout fp32 filter into the streaming layout of `pass`:
 *   [tap][n_out][k_in] with k_in contiguous, dtype = desc->io_dtype, where
 *   (n_out,k_in) = (y.c,x.c) for the pass that produces y and (x.c,y.c) for
 *   the pass that produces dx.  Replaces the implicit weight handling of
 *   nn.Conv3d / nn.ConvTranspose3d (dcgan.py:47-67,161-181, utils.py:8-14,
 *   FID_ResNet.py:9-16,70-75,119-125, stylegan.py:164-170).               */
size_t b2c_filter_packed_bytes(const b2c_conv_desc*, int pass);
int    b2c_filter_pack(const b2c_conv_desc*, int pass, const float* w_torch,
                       void* w_packed, b2c_stream);

/* ---- convolution --------------------------------------------------- *
 * fwd  : y  = act(conv(x, w) + bias)     F.conv3d / F.conv_transpose3d
 * dgrad: dx = conv_backward_input(dy, w)  (autograd of the same calls)
 * wgrad: dw (torch layout, fp32, overwritten) and optional dbias [y.c].
 * Workspace: b2c_conv3d_workspace(desc, pass) bytes (may be 0).
 * Call sites replaced: dcgan.py:45-70,158-182; utils.py:8-14,47-132;
 * biggan.py:42,78; FID_ResNet.py:65-102,119-125; stylegan.py:164-170.     */
size_t b2c_conv3d_workspace(const b2c_conv_desc*, int pass);
size_t b2c_conv3d_stats_workspace(const b2c_conv_desc*, int stat_kind);
int b2c_conv3d_fwd  (const b2c_conv_desc*, const void* x, const void* w_packed,
                     void* y, const b2c_fusion*, void* ws, size_t ws_bytes,
                     b2c_stream);
int b2c_conv3d_dgrad(const b2c_conv_desc*, const void* dy, const void* w_packed,
                     void* dx, void* ws, size_t ws_bytes, b2c_stream);
int b2c_conv3d_wgrad(const b2c_conv_desc*, const void* x, const void* dy,
                     float* dw_torch, float* dbias, void* ws, size_t ws_bytes,
                     b2c_stream);

/* ---- layout -------------------------------------------------------- *
 * dst[b][j][i] = (dst_dtype) src[b][i][j]   (src is [batch][rows][cols]).
 * NCDHW -> NDHWC is rows=C, cols=D*H*W; the inverse swaps them.  Also used
 * to permute LayerNorm's [C,D,H,W] affine (dcgan.py:162-178).             */
int b2c_transpose(const void* src, void* dst, int64_t batch, int64_t rows,
                  int64_t cols, int src_dtype, int dst_dtype, b2c_stream);
/* dst[i] = (dst_dtype) src[i] */
int b2c_cast(const void* src, void* dst, int64_t n, int src_dtype,
             int dst_dtype, b2c_stream);

/* ---- statistics ---------------------------------------------------- *
 * x is [rows][C] (NDHWC flattened).  per-channel: sum[c], sqsum[c] over
 * rows (BatchNorm3d, dcgan.py:48-64, utils.py:62-65, FID_ResNet.py:71-76).
 * Deterministic two-stage reduction; ws from b2c_stats_workspace.         */
size_t b2c_stats_workspace(int64_t rows, int64_t C);
int b2c_channel_stats(const void* x, int64_t rows, int64_t C, int dtype,
                      float* sum, float* sqsum, void* ws, size_t ws_bytes,
                      b2c_stream);
/* Allocation-free variant of b2c_channel_stats.  NO COUNTERPART IN THE REFERENCE:
 * added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  Same inputs, outputs and arithmetic (bit-identical results) as
 * b2c_channel_stats, but the stage-1 partial sums go to a file-scope
 * `static __device__` array of B2C_STATIC_SCRATCH_BYTES (4 MiB) inside the
 * library instead of a caller workspace; every GPU that loads the library holds
 * that array until exit.  Calls no cudaMalloc / cudaFree / allocating routine.
 * Limit: b2c_stats_workspace(rows, C) <= B2C_STATIC_SCRATCH_BYTES (C <= 885 at
 * 148 SMs), otherwise B2C_ERR_STATIC_LIMIT.  All calls on a device share the
 * array (it keeps the last call's partials; nothing needs clearing because
 * stage 1 overwrites every slot stage 2 reads): calls that overlap on the GPU
 * (different streams, or concurrent graph launches) RACE - serialise them.
 * The first call on each device looks the array's address up
 * (cudaGetSymbolAddress); make that call once before starting a capture.      */
#define B2C_STATIC_SCRATCH_BYTES (4u << 20)
int b2c_channel_stats_static(const void* x, int64_t rows, int64_t C, int dtype,
                             float* sum, float* sqsum, b2c_stream);
/* BatchNorm3d training-mode finalize (momentum update exactly like torch:
 * running = (1-m)*running + m*batch, unbiased var for running, biased for
 * normalisation).  Writes mean[C], invstd[C], scale[C]=gamma*invstd,
 * shift[C]=beta-mean*scale.  running_* may be NULL.                        */
int b2c_bn_finalize(const float* sum, const float* sqsum, int64_t count,
                    int64_t C, float eps, float momentum, const float* gamma,
                    const float* beta, float* mean, float* invstd,
                    float* scale, float* shift, float* running_mean,
                    float* running_var, b2c_stream);
/* y = act(x*scale[c] + shift[c] (+ residual))  — BN apply + ReLU (+ the
 * Bottleneck residual add, FID_ResNet.py:97-100) in one HBM pass.          */
int b2c_bn_act_fwd(const void* x, void* y, const void* residual, int64_t rows,
                   int64_t C, int dtype, const float* scale,
                   const float* shift, int act, float slope, b2c_stream);
/* Backward of the above. Step 1 reduces per channel
 *   s1[c] = sum g, s2[c] = sum g*xhat   with g = dy*act'(z), z the
 *   pre-activation recomputed from x (and residual), xhat=(x-mean)*invstd.
 * Step 2 writes dx = scale*(g - s1/M - xhat*s2/M) and, when dres != NULL,
 * dres = g (gradient of the residual branch).  dgamma = s2, dbeta = s1.     */
int b2c_bn_act_bwd_reduce(const void* x, const void* dy, const void* residual,
                          int64_t rows, int64_t C, int dtype,
                          const float* mean, const float* invstd,
                          const float* scale, const float* shift, int act,
                          float slope, float* s1, float* s2, void* ws,
                          size_t ws_bytes, b2c_stream);
int b2c_bn_act_bwd_apply(const void* x, const void* dy, const void* residual,
                         void* dx, void* dres, int64_t rows, int64_t C,
                         int dtype, const float* mean, const float* invstd,
                         const float* scale, const float* shift,
                         const float* s1, const float* s2, int act,
                         float slope, b2c_stream);

/* ---- LayerNorm([C,D,H,W]) + LeakyReLU  (dcgan.py:162-179) ---------- *
 * x is [N][L] with L = D*H*W*C (NDHWC); gamma/beta are [L] fp32 already
 * permuted to NDHWC order (b2c_transpose).  Per-sample mean/rstd [N].      */
size_t b2c_ln_workspace(int64_t N, int64_t L);
int b2c_ln_act_fwd(const void* x, void* y, int64_t N, int64_t L, int dtype,
                   const float* gamma, const float* beta, float eps, int act,
                   float slope, float* mean, float* rstd, void* ws,
                   size_t ws_bytes, b2c_stream);
/* The same forward when the producing convolution already delivered the
 * per-sample sums (b2c_fusion.stat_kind = B2C_STAT_SAMPLE): finalizes mean /
 * rstd from sum / sqsum [N] and runs only the apply pass (no workspace).     */
int b2c_ln_act_fwd_stats(const void* x, void* y, int64_t N, int64_t L, int dtype,
                         const float* gamma, const float* beta, float eps,
                         int act, float slope, const float* sum,
                         const float* sqsum, float* mean, float* rstd,
                         b2c_stream);
/* dgamma/dbeta [L] fp32 in NDHWC order (may be NULL when the affine is
 * frozen, trainer.py:278-279); dx may be NULL when x needs no grad.         */
int b2c_ln_act_bwd(const void* x, const void* dy, void* dx, int64_t N,
                   int64_t L, int dtype, const float* gamma, const float* beta,
                   const float* mean, const float* rstd, int act, float slope,
                   float* dgamma, float* dbeta, void* ws, size_t ws_bytes,
                   b2c_stream);

/* ---- pointwise ------------------------------------------------------ *
 * y = act(x); dx = dy * act'(.) evaluated from the OUTPUT y (valid for
 * relu/lrelu with slope>0/tanh — all in-place-safe, dcgan.py:49,163,68).   */
int b2c_act_fwd(const void* x, void* y, int64_t n, int dtype, int act,
                float slope, b2c_stream);
int b2c_act_bwd(const void* y, const void* dy, void* dx, int64_t n, int dtype,
                int act, float slope, b2c_stream);

/* ---- x2 resampling on NDHWC tensors ------------- *
 * nearest x2 of the channel slice [c0, c0 + c) of x [n,d,h,w,cx] into
 * y [n,2d,2h,2w,c]: replaces `x[:, :Cout]` + F.interpolate(scale_factor=2)
 * (utils.py:76-80, biggan.py:26,31).  bwd: dx [n,d,h,w,cx] = 2x2x2 block
 * sums of dy inside the slice, 0 outside.                                  */
int b2c_upsample2x_fwd(const void* x, void* y, int64_t n, int d, int h, int w,
                       int cx, int c0, int c, int dtype, b2c_stream);
int b2c_upsample2x_bwd(const void* dy, void* dx, int64_t n, int d, int h,
                       int w, int cx, int c0, int c, int dtype, b2c_stream);
/* nn.AvgPool3d(2) (utils.py:113-116,128-129; biggan.py:86,92): x [n,d,h,w,c]
 * -> y [n,d/2,h/2,w/2,c] (floor); bwd writes every element of dx.          */
int b2c_avgpool2_fwd(const void* x, void* y, int64_t n, int d, int h, int w,
                     int c, int dtype, b2c_stream);
int b2c_avgpool2_bwd(const void* dy, void* dx, int64_t n, int d, int h, int w,
                     int c, int dtype, b2c_stream);

/* ---- depthwise FIR resampling (stylegan.py:172-200 `_upfirdn3d_ref`) -- *
 * y[n,o,c] = sum_k taps[k] * xup[n, o*down + k - pad, c] with xup the
 * zero-inserted (x up) input; x [n,in,c], y [n,out,c] dense NDHWC.  One
 * filter for all channels (the reference repeats f over groups = C,
 * stylegan.py:195-197).  `taps` is a HOST array of k[0]*k[1]*k[2] <= 64
 * correlation taps: the caller applies the reference's flip and gain
 * (stylegan.py:190-192).  pad may be negative (crop, stylegan.py:187).
 * bwd = gradient w.r.t. x (the filter is a constant buffer).               */
typedef struct {
  int64_t n;
  int32_t c, in[3], out[3], k[3], pad[3], up, down, dtype;
} b2c_fir_desc;
int b2c_fir3d_fwd(const b2c_fir_desc*, const float* taps, const void* x,
                  void* y, b2c_stream);
int b2c_fir3d_bwd(const b2c_fir_desc*, const float* taps, const void* dy,
                  void* dx, b2c_stream);

/* ---- style modulation glue (stylegan.py:396-444 `modulated_conv3d`, live --- *
 * branch, + the bias_act of the calling layer, stylegan.py:522-531) --------- *
 * x, y, out [N, V, C] dense NDHWC (V = D*H*W voxels), bf16 or fp32.            *
 *   modulate:  y   = x * s[n,c]                         (stylegan.py:427)      *
 *   epilogue:  out = act(y * d[n,c] + *strength * field[n|0, v] + bias[c])     *
 *              (demodulation :429-432, noise :522-524, bias_act :531)          *
 * s, d [N,C] fp32; d may be NULL (demodulate=False); field [field_n, V] fp32   *
 * with field_n = N or 1 (broadcast over the batch) or NULL; strength a DEVICE  *
 * scalar; bias [C] fp32 or NULL.  The backward passes also produce the small   *
 * reductions: ds[n,c] = sum_v gy*x;  dd[n,c] = sum_v g*y, dbias[c] = sum g,    *
 * dstrength = sum g*field  with g = gout * act'(out)  (deterministic order).   *
 * C must be a multiple of the 16-byte vector (8 bf16 / 4 fp32) and <= 2048 bf16*
 * / 1024 fp32, else B2C_ERR_UNSUPPORTED.  Workspace: b2c_mod_workspace bytes.  */
size_t b2c_mod_workspace(int64_t N, int64_t V, int64_t C);
int b2c_modulate_fwd(const void* x, const float* s, void* y, int64_t N,
                     int64_t V, int64_t C, int dtype, b2c_stream);
int b2c_modulate_bwd(const void* x, const void* gy, const float* s, void* dx,
                     float* ds, int64_t N, int64_t V, int64_t C, int dtype,
                     void* ws, size_t ws_bytes, b2c_stream);
int b2c_mod_epilogue_fwd(const void* y, const float* d, const float* field,
                         int64_t field_n, const float* strength,
                         const float* bias, void* out, int64_t N, int64_t V,
                         int64_t C, int dtype, int act, float slope, b2c_stream);
int b2c_mod_epilogue_bwd(const void* y, const void* out, const void* gout,
                         const float* d, const float* field, int64_t field_n,
                         void* dy, float* dd, float* dbias, float* dstrength,
                         int64_t N, int64_t V, int64_t C, int dtype, int act,
                         float slope, void* ws, size_t ws_bytes, b2c_stream);

/* ---- AdaIN of the v1-like generator (stylegan.py:931-959), optionally fused *
 * with the trilinear x2 that follows it (stylegan.py:1081, align_corners =     *
 * False): y = a[n,c] * (up ? tri2x(x) : x) + b[n,c], x [N, d*h*w, C] NDHWC.   *
 * The caller derives a = s_std / sigma, b = s_mean - mean * a from            *
 * b2c_instance_stats (sum and sum of squares per (n, c), fp32) and combines   *
 * the backward reductions r1 = sum t, r2 = sum t*x (t = up ? tri2x^T(gy) : gy *
 * at the INPUT resolution; for up the kernel stores t in `t`) into            *
 * dx = ct[n,c]*t + cx[n,c]*x + c0[n,c] (b2c_adain_bwd_apply; t may alias dx). *
 * Workspace: b2c_mod_workspace(N, V, C) bytes; same channel restrictions.     */
int b2c_instance_stats(const void* x, float* sum, float* sqsum, int64_t N,
                       int64_t V, int64_t C, int dtype, void* ws,
                       size_t ws_bytes, b2c_stream);
int b2c_adain_fwd(const void* x, const float* a, const float* b, void* y,
                  int64_t N, int d, int h, int w, int64_t C, int dtype, int up,
                  b2c_stream);
int b2c_adain_bwd_reduce(const void* x, const void* gy, void* t, float* r1,
                         float* r2, int64_t N, int d, int h, int w, int64_t C,
                         int dtype, int up, void* ws, size_t ws_bytes,
                         b2c_stream);
int b2c_adain_bwd_apply(const void* x, const void* t, const float* ct,
                        const float* cx, const float* c0, void* dx, int64_t N,
                        int64_t V, int64_t C, int dtype, b2c_stream);

/* ---- nn.MaxPool3d(k, stride, pad) on NDHWC (FID_ResNet.py:129) -------------- *
 * x [n,d,h,w,c] -> y [n,od,oh,ow,c], o = (i + 2*pad - k) / stride + 1 (floor);   *
 * padding counts as -inf; ties go to the first element in scan order (ATen).    *
 * bwd: dx from (x, dy), deterministic gather form (recomputes the arg-max).     */
int b2c_maxpool3d_fwd(const void* x, void* y, int64_t n, int d, int h, int w,
                      int c, int k, int stride, int pad, int dtype, b2c_stream);
int b2c_maxpool3d_bwd(const void* x, const void* dy, void* dx, int64_t n, int d,
                      int h, int w, int c, int k, int stride, int pad, int dtype,
                      b2c_stream);

/* ---- depth-to-space by 2 on NDHWC (second half of the phase-merged stride-2   *
 * transposed convolution; replaces the sub-pixel scatter of F.conv_transpose3d  *
 * at stylegan.py:170 for the narrow up-sampling layers of stylegan.py:262-282)  *
 * z [n,jd,jh,jw,8*c], channel = r*c + ch, r = rd*4 + rh*2 + rw                  *
 * y [n,od,oh,ow,c],  y[n, 2*j + r] = z[n, j, r];  od <= 2*jd (cropped).         *
 * inverse == 0: z -> y (elements of y beyond 2*j are not touched: none exist);  *
 * inverse != 0: y -> z, zero where the even grid overhangs y (the adjoint).     */
int b2c_depth_to_space2(const void* src, void* dst, int64_t n, int jd, int jh,
                        int jw, int c, int od, int oh, int ow, int dtype,
                        int inverse, b2c_stream);

/* ---- W-window gather on NDHWC (x [rows, W, C], rows = N*D*H) ------------------ *
 * y [rows, W/F, (F+2*pad)*C],  y[r, j, v*C + c] = x[r, F*j + v - pad, c] (0 outside).    *
 * Lowers the narrow stride-1 convolutions of utils.py:57-60,105-109 (32- / 64-channel  *
 * 3^3 convs at 128^3 / 64^3) and stylegan.py:285-287 to a (F+2*pad)*C -> F*Cout conv   *
 * with one tap along W and a block-Toeplitz filter: its output [rows, W/F, F*Cout] is  *
 * the result [rows, W, Cout] itself, M = F*Cout = 128 fills the UMMA tile and half of   *
 * the issued MACs are useful (a quarter with the three-tap folding).  inverse != 0:    *
 * the adjoint (overlap-add, gather form), y -> x; used by autograd.                     */
int b2c_window_w(const void* src, void* dst, int64_t rows, int w, int c, int f,
                 int pad, int dtype, int inverse, b2c_stream);

/* ---- gradient averaging over NVLink peer memory ----------------------- *
 * Replaces the gradient reduction of nn.DataParallel (trainer.py:70-72: the replicas'      *
 * gradients are summed onto GPU 0; with batch-mean losses, trainer.py:228-241,272, that is *
 * the average over the global batch).  One process per GPU; every rank keeps its gradient  *
 * buckets in memory its peers can address (CUDA IPC over NVLink / NVSwitch):               *
 *   b2c_peer_alloc  cudaMalloc + zero fill + export handle (64 bytes, to be sent to peers) *
 *   b2c_peer_open   map a peer's allocation from its handle (enables peer access)          *
 *   b2c_peer_close / b2c_peer_free   undo the above                                        *
 * b2c_peer_allreduce_avg averages `count` fp32 values IN PLACE over `world` ranks with ONE *
 * kernel per rank (two-shot over peer pointers): rank r waits until every peer's kernel    *
 * has started (their gradients are then complete: stream order), reads slice r of every    *
 * rank's bucket over NVLink, adds the `world` values in rank order 0, 1, ... (every rank   *
 * therefore ends with bit-identical values, and the same values run after run), scales by  *
 * 1 / world and writes the result into slice r of EVERY rank's bucket; a second flag       *
 * exchange tells each rank that all of its slices have landed.  The kernel is small on     *
 * purpose - `ctas` blocks (<= B2C_PEER_MAX_CTAS) of 192 threads, <= 80 registers, no      *
 * shared memory - so that its blocks fit on an SM NEXT TO a resident block of the          *
 * persistent convolution kernels of the backward pass it overlaps (a collective that       *
 * takes whole SMs makes the statically scheduled 148-block conv grids wait for it).        *
 *   bufs[i]   (host array, `world` entries) device address of the bucket on rank i, as     *
 *             mapped in THIS process (bufs[rank] = the local allocation)                   *
 *   flags[i]  likewise: rank i's flag block, B2C_PEER_FLAG_BYTES, zero at start, never     *
 *             reset: `seq` must be 1, 2, 3, ... over the calls that share the flag blocks, *
 *             the same sequence (and the same `ctas`) on every rank                        *
 * Calls that share flag blocks must be ordered on one stream per rank.  A peer that does   *
 * not arrive within ~10 s makes the kernel give up (result undefined) and                  *
 * b2c_peer_error() return nonzero from then on.                                            */
#define B2C_PEER_MAX_WORLD 8
#define B2C_PEER_MAX_CTAS 160
#define B2C_PEER_FLAG_BYTES (B2C_PEER_MAX_CTAS * 16 * 4)
#define B2C_PEER_HANDLE_BYTES 64
int b2c_peer_alloc(size_t bytes, void** ptr, unsigned char* handle /*[B2C_PEER_HANDLE_BYTES]*/);
int b2c_peer_open(const unsigned char* handle, void** ptr);
int b2c_peer_close(void* ptr);
int b2c_peer_free(void* ptr);
int b2c_peer_allreduce_avg(void* const* bufs, void* const* flags, int rank, int world,
                           int64_t count, uint32_t seq, int ctas, b2c_stream);
unsigned b2c_peer_error(void);

#ifdef __cplusplus
}
#endif
#endif /* B2C_H_ */
