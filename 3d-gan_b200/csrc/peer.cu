// peer.cu — gradient averaging over NVLink peer memory (include/b2c.h, "gradient averaging over NVLink peer
// memory"; the reference's counterpart is nn.DataParallel's gradient reduce, trainer.py:70-72).
//
// One kernel per rank, two-shot over peer pointers.  Design constraints, in the order they shaped it:
//  * the kernel overlaps the backward pass, whose convolution kernels are PERSISTENT (148 blocks, ~225 KB of shared
//    memory and ~47.6 K registers each, static tile schedule).  A collective whose blocks need an SM to themselves
//    makes the displaced conv blocks start when the others finish: the conv kernel takes twice as long.  So this
//    kernel must fit BESIDE a conv block: 192 threads x <= 80 registers (15.4 K of the 17.9 K registers left), no
//    shared memory at all (the conv kernels leave < 2 KB);
//  * no grid-wide barrier (the blocks of this kernel are not guaranteed to be resident together): block c of rank r
//    only ever synchronises with block c of the other ranks, through its own row of the flag block;
//  * NVLink loads are latency bound (~2 us): every thread keeps 6-8 sixteen-byte peer loads in flight;
//  * determinism: slice r is summed by rank r alone, in rank order, and broadcast - all replicas stay bit-identical.
#include <mutex>

#include "common.cuh"

namespace b2c {

__device__ unsigned g_peer_error = 0;

struct PeerParams {
  float* buf[B2C_PEER_MAX_WORLD];
  uint32_t* flag[B2C_PEER_MAX_WORLD];
  int64_t count;
  int rank;
  uint32_t seq;
};

constexpr int kPeerThreads = 192;

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float4 ld_peer(const float* p) {     // never from a stale cache line: the data is a peer's
  float4 v;
  asm volatile("ld.volatile.global.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
  return v;
}
__device__ __forceinline__ uint64_t global_ns() {
  uint64_t t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

// Flag block of a rank: [B2C_PEER_MAX_CTAS][16] words; word [c][s] is written by block c of rank s only.
// Values only grow (2 * seq + phase), so a peer that is already one call ahead still satisfies the wait.
template <int W>
__device__ __forceinline__ void exchange(const PeerParams& P, int c, uint32_t val) {
  __syncthreads();                                   // every thread's stores are fenced (see callers)
  if (threadIdx.x < W) {
    st_release_sys(P.flag[threadIdx.x] + c * 16 + P.rank, val);
    const uint32_t* mine = P.flag[P.rank] + c * 16 + threadIdx.x;
    const uint64_t t0 = global_ns();
    uint32_t spins = 0;
    while ((int32_t)(ld_acquire_sys(mine) - val) < 0) {
      // the peer may be milliseconds behind (rank skew): poll gently, the SM is shared with a convolution block
      __nanosleep(spins < 64 ? 100 : 1000);
      if ((++spins & 1023u) == 0 && global_ns() - t0 > 10000000000ull) { atomicExch(&g_peer_error, 1u); break; }
    }
  }
  __syncthreads();
}

// U = float4 positions per thread per iteration
template <int W, int U>
__global__ void __launch_bounds__(kPeerThreads, 4) peer_allreduce_kernel(const PeerParams P) {
  const int c = blockIdx.x;
  exchange<W>(P, c, 2u * P.seq + 1u);                // every rank's gradients are complete
  const int64_t per = ((P.count + W - 1) / W + 3) & ~(int64_t)3;          // slice length, whole float4s
  const int64_t lo = (int64_t)P.rank * per;
  const int64_t hi = lo + per < P.count ? lo + per : P.count;
  const float scale = 1.f / (float)W;
  constexpr int64_t kBlock = (int64_t)kPeerThreads * U * 4;              // floats per block iteration
  for (int64_t base = lo + (int64_t)c * kBlock; base < hi; base += (int64_t)gridDim.x * kBlock) {
    float4 v[U][W];
#pragma unroll
    for (int u = 0; u < U; u++) {
      const int64_t i = base + ((int64_t)u * kPeerThreads + threadIdx.x) * 4;
      if (i + 4 <= hi) {                              // count need not be a multiple of 4: the tail is handled below
#pragma unroll
        for (int r = 0; r < W; r++) v[u][r] = ld_peer(P.buf[r] + i);
      }
    }
#pragma unroll
    for (int u = 0; u < U; u++) {
      const int64_t i = base + ((int64_t)u * kPeerThreads + threadIdx.x) * 4;
      if (i + 4 <= hi) {
        float4 a = v[u][0];
#pragma unroll
        for (int r = 1; r < W; r++) { a.x += v[u][r].x; a.y += v[u][r].y; a.z += v[u][r].z; a.w += v[u][r].w; }
        a.x *= scale; a.y *= scale; a.z *= scale; a.w *= scale;
#pragma unroll
        for (int r = 0; r < W; r++) *reinterpret_cast<float4*>(P.buf[r] + i) = a;
      } else if (i < hi) {                            // last 1..3 values of the bucket
        for (int64_t j = i; j < hi; j++) {
          float a = *reinterpret_cast<volatile float*>(P.buf[0] + j);
          for (int r = 1; r < W; r++) a += *reinterpret_cast<volatile float*>(P.buf[r] + j);
          a *= scale;
          for (int r = 0; r < W; r++) P.buf[r][j] = a;
        }
      }
    }
  }
  __threadfence_system();                            // my stores to the peers are visible before my flag is
  exchange<W>(P, c, 2u * P.seq + 2u);                // every slice of MY bucket has landed; nobody still reads it
}

template <int W, int U>
static int peer_launch(const PeerParams& P, int ctas, cudaStream_t st) {
  peer_allreduce_kernel<W, U><<<ctas, kPeerThreads, 0, st>>>(P);
  return launched();
}

}  // namespace b2c

using namespace b2c;

extern "C" int b2c_peer_alloc(size_t bytes, void** ptr, unsigned char* handle) {
  if (!ptr || !handle || bytes == 0) return B2C_ERR_INVALID;
  static_assert(sizeof(cudaIpcMemHandle_t) == B2C_PEER_HANDLE_BYTES, "handle size");
  void* p = nullptr;
  if (cudaMalloc(&p, bytes) != cudaSuccess) { cudaGetLastError(); return B2C_ERR_CUDA; }
  cudaIpcMemHandle_t h;
  if (cudaMemset(p, 0, bytes) != cudaSuccess || cudaDeviceSynchronize() != cudaSuccess ||
      cudaIpcGetMemHandle(&h, p) != cudaSuccess) {
    cudaGetLastError();
    cudaFree(p);
    return B2C_ERR_CUDA;
  }
  memcpy(handle, &h, sizeof(h));
  *ptr = p;
  return B2C_OK;
}

extern "C" int b2c_peer_open(const unsigned char* handle, void** ptr) {
  if (!ptr || !handle) return B2C_ERR_INVALID;
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, sizeof(h));
  void* p = nullptr;
  if (cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { cudaGetLastError(); return B2C_ERR_CUDA; }
  *ptr = p;
  return B2C_OK;
}

extern "C" int b2c_peer_close(void* ptr) {
  if (!ptr) return B2C_ERR_INVALID;
  return cuda_rc(cudaIpcCloseMemHandle(ptr));
}

extern "C" int b2c_peer_free(void* ptr) {
  if (!ptr) return B2C_ERR_INVALID;
  return cuda_rc(cudaFree(ptr));
}

extern "C" int b2c_peer_allreduce_avg(void* const* bufs, void* const* flags, int rank, int world, int64_t count,
                                      uint32_t seq, int ctas, b2c_stream stream) {
  if (!bufs || !flags || world < 1 || world > B2C_PEER_MAX_WORLD || rank < 0 || rank >= world || count < 0 ||
      seq == 0 || seq > 0x3fffffffu || ctas < 1 || ctas > B2C_PEER_MAX_CTAS)
    return B2C_ERR_INVALID;
  PeerParams P;
  for (int i = 0; i < B2C_PEER_MAX_WORLD; i++) { P.buf[i] = nullptr; P.flag[i] = nullptr; }
  for (int i = 0; i < world; i++) {
    if (!bufs[i] || !flags[i] || (reinterpret_cast<uintptr_t>(bufs[i]) & 15)) return B2C_ERR_INVALID;
    P.buf[i] = static_cast<float*>(bufs[i]);
    P.flag[i] = static_cast<uint32_t*>(flags[i]);
  }
  P.count = count; P.rank = rank; P.seq = seq;
  if (world == 1 || count == 0) return B2C_OK;          // nothing to average; no flags are touched on any rank
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  switch (world) {
    case 2: return peer_launch<2, 6>(P, ctas, st);
    case 3: return peer_launch<3, 4>(P, ctas, st);
    case 4: return peer_launch<4, 2>(P, ctas, st);
    case 5: return peer_launch<5, 2>(P, ctas, st);
    case 6: return peer_launch<6, 1>(P, ctas, st);
    case 7: return peer_launch<7, 1>(P, ctas, st);
    default: return peer_launch<8, 1>(P, ctas, st);
  }
}

extern "C" unsigned b2c_peer_error(void) {
  unsigned v = 0;
  if (cudaMemcpyFromSymbol(&v, g_peer_error, sizeof(v)) != cudaSuccess) { cudaGetLastError(); return 0xffffffffu; }
  return v;
}
