// pack.cu — filter re-packing, layout transposition and dtype casts.
// All HBM-bound: coalesced reads AND writes through a padded shared tile.
#include "common.cuh"
#include "thin.cuh"

namespace b2c {

// ---------------------------------------------------------------------------
// Batched 2-D transpose with dtype conversion:
//   dst[b][j][i] = (D) src[b][i][j],  src [batch][rows][cols].
// 32x32 tile, 32x8 threads, +1 padding -> conflict-free, both sides coalesced.
template <typename S, typename D>
__global__ void transpose_kernel(const S* __restrict__ src, D* __restrict__ dst, int64_t rows,
                                 int64_t cols, int64_t tiles_c, int64_t tiles_r) {
  __shared__ float tile[32][33];
  int64_t t = blockIdx.x;
  const int64_t tc = t % tiles_c; t /= tiles_c;
  const int64_t tr = t % tiles_r;
  const int64_t b = t / tiles_r;
  const S* s = src + b * rows * cols;
  D* d = dst + b * rows * cols;
  const int tx = threadIdx.x, ty = threadIdx.y;
#pragma unroll
  for (int k = 0; k < 32; k += 8) {
    int64_t r = tr * 32 + ty + k, c = tc * 32 + tx;
    if (r < rows && c < cols) tile[ty + k][tx] = ldf(s, r * cols + c);
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < 32; k += 8) {
    int64_t c = tc * 32 + ty + k, r = tr * 32 + tx;
    if (r < rows && c < cols) stf(d, c * rows + r, tile[tx][ty + k]);
  }
}

template <typename S, typename D>
static int transpose_launch(const void* src, void* dst, int64_t batch, int64_t rows, int64_t cols,
                            cudaStream_t st) {
  const int64_t tr = cdiv(rows, 32), tc = cdiv(cols, 32);
  const int64_t blocks = batch * tr * tc;
  if (blocks == 0) return B2C_OK;
  if (blocks > 0x7fffffffLL) return B2C_ERR_UNSUPPORTED;
  transpose_kernel<S, D><<<(unsigned)blocks, dim3(32, 8), 0, st>>>(
      static_cast<const S*>(src), static_cast<D*>(dst), rows, cols, tc, tr);
  return launched();
}

// ---------------------------------------------------------------------------
// Elementwise cast, 4 elements per thread, grid-stride.
template <typename S, typename D>
__global__ void cast_kernel(const S* __restrict__ src, D* __restrict__ dst, int64_t n) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    stf(dst, i, ldf(src, i));
}

template <typename S, typename D>
static int cast_launch(const void* src, void* dst, int64_t n, cudaStream_t st) {
  if (n == 0) return B2C_OK;
  const int threads = 256;
  const int64_t blocks = cdiv(n, threads) < (int64_t)num_sms() * 16 ? cdiv(n, threads) : (int64_t)num_sms() * 16;
  cast_kernel<S, D><<<(unsigned)blocks, threads, 0, st>>>(static_cast<const S*>(src), static_cast<D*>(dst), n);
  return launched();
}

// ---------------------------------------------------------------------------
// Filter pack: torch [A][B][T] fp32 -> [T][n_out][k_in] of type D.
//   out_is_A: packed[t][a][b] = w[a][b][t]     (k_in = B)
//   else    : packed[t][b][a] = w[a][b][t]     (k_in = A)
// For fixed a the source is a [B][T] matrix; a 32x32 smem tile over (b,t) makes
// the source read coalesced along t and the packed write coalesced along b
// (out_is_A).  For !out_is_A the write is coalesced along a instead, so the
// tile runs over (a, t) at fixed b.  One kernel, parametrised by strides.
template <typename D>
__global__ void pack_kernel(const float* __restrict__ w, D* __restrict__ out, int outer, int inner,
                            int T, int64_t src_outer_stride, int64_t src_inner_stride,
                            int64_t dst_outer_stride /* in elements, per outer index */,
                            int64_t dst_tap_stride, int tiles_i, int tiles_t) {
  // outer: the channel index that stays fixed per block;
  // inner: the channel index that becomes contiguous in the packed layout.
  __shared__ float tile[32][33];
  int64_t bid = blockIdx.x;
  const int tt = (int)(bid % tiles_t); bid /= tiles_t;
  const int ti = (int)(bid % tiles_i);
  const int o = (int)(bid / tiles_i);
  const int tx = threadIdx.x, ty = threadIdx.y;
#pragma unroll
  for (int k = 0; k < 32; k += 8) {
    int i = ti * 32 + ty + k, t = tt * 32 + tx;
    if (i < inner && t < T) tile[ty + k][tx] = w[o * src_outer_stride + i * src_inner_stride + t];
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < 32; k += 8) {
    int t = tt * 32 + ty + k, i = ti * 32 + tx;
    if (i < inner && t < T) stf(out, t * dst_tap_stride + o * dst_outer_stride + i, tile[tx][ty + k]);
  }
  (void)outer;
}

template <typename D>
static int pack_launch(const float* w, void* out, int A, int B, int T, bool out_is_A, cudaStream_t st) {
  // out_is_A : packed[t][a][b]; fixed (outer) = a, contiguous (inner) = b.
  // !out_is_A: packed[t][b][a]; fixed (outer) = b, contiguous (inner) = a.
  const int outer = out_is_A ? A : B, inner = out_is_A ? B : A;
  const int64_t so = out_is_A ? (int64_t)B * T : (int64_t)T;
  const int64_t si = out_is_A ? (int64_t)T : (int64_t)B * T;
  const int tiles_i = (int)cdiv(inner, 32), tiles_t = (int)cdiv(T, 32);
  const int64_t blocks = (int64_t)outer * tiles_i * tiles_t;
  if (blocks > 0x7fffffffLL) return B2C_ERR_UNSUPPORTED;
  pack_kernel<D><<<(unsigned)blocks, dim3(32, 8), 0, st>>>(
      w, static_cast<D*>(out), outer, inner, T, so, si, (int64_t)inner, (int64_t)A * B, tiles_i, tiles_t);
  return launched();
}

}  // namespace b2c

using namespace b2c;

extern "C" int b2c_transpose(const void* src, void* dst, int64_t batch, int64_t rows, int64_t cols,
                             int sdt, int ddt, b2c_stream s) {
  if (!src || !dst || !dt_ok(sdt) || !dt_ok(ddt) || batch < 0 || rows < 0 || cols < 0) return B2C_ERR_INVALID;
  cudaStream_t st = static_cast<cudaStream_t>(s);
  if (sdt == B2C_F32 && ddt == B2C_F32) return transpose_launch<float, float>(src, dst, batch, rows, cols, st);
  if (sdt == B2C_F32 && ddt == B2C_BF16) return transpose_launch<float, __nv_bfloat16>(src, dst, batch, rows, cols, st);
  if (sdt == B2C_BF16 && ddt == B2C_F32) return transpose_launch<__nv_bfloat16, float>(src, dst, batch, rows, cols, st);
  return transpose_launch<__nv_bfloat16, __nv_bfloat16>(src, dst, batch, rows, cols, st);
}

extern "C" int b2c_cast(const void* src, void* dst, int64_t n, int sdt, int ddt, b2c_stream s) {
  if (!src || !dst || !dt_ok(sdt) || !dt_ok(ddt) || n < 0) return B2C_ERR_INVALID;
  cudaStream_t st = static_cast<cudaStream_t>(s);
  if (sdt == B2C_F32 && ddt == B2C_F32) return cast_launch<float, float>(src, dst, n, st);
  if (sdt == B2C_F32 && ddt == B2C_BF16) return cast_launch<float, __nv_bfloat16>(src, dst, n, st);
  if (sdt == B2C_BF16 && ddt == B2C_F32) return cast_launch<__nv_bfloat16, float>(src, dst, n, st);
  return cast_launch<__nv_bfloat16, __nv_bfloat16>(src, dst, n, st);
}

extern "C" size_t b2c_filter_packed_bytes(const b2c_conv_desc* c, int pass) {
  (void)pass;
  if (check_desc(c)) return 0;
  const Geo g = make_geo(c);
  Geo tg;
  if (thin_mode(c, g, &tg)) return thin_packed_bytes(tg);
  return (size_t)c->k[0] * c->k[1] * c->k[2] * c->x.c * c->y.c * esize(c->io_dtype);
}

extern "C" int b2c_filter_pack(const b2c_conv_desc* c, int pass, const float* w, void* wp, b2c_stream s) {
  int rc = check_desc(c);
  if (rc) return rc;
  if (!w || !wp || (pass != B2C_PASS_FWD && pass != B2C_PASS_DGRAD)) return B2C_ERR_INVALID;
  const int T = c->k[0] * c->k[1] * c->k[2];
  const int A = c->transposed ? c->x.c : c->y.c;   // torch dim 0
  const int B = c->transposed ? c->y.c : c->x.c;   // torch dim 1
  // conv fwd / convT dgrad produce torch-dim-0 channels; the other two dim 1
  const bool out_is_A = (pass == B2C_PASS_FWD) != (c->transposed != 0);
  cudaStream_t st = static_cast<cudaStream_t>(s);
  const Geo g = make_geo(c);
  Geo tg;
  if (const int tm = thin_mode(c, g, &tg)) return thin_pack(tg, /*gather=*/out_is_A != (tm == 2), w, wp, st, tm == 2);
  if (c->io_dtype == B2C_BF16) return pack_launch<__nv_bfloat16>(w, wp, A, B, T, out_is_A, st);
  return pack_launch<float>(w, wp, A, B, T, out_is_A, st);
}
