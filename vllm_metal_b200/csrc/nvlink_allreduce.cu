// One-shot sum all-reduce over NVLink peer memory for the tensor-parallel decode step.
//
// The reference has no tensor parallelism (vllm_metal/platform.py:524-529 pins world size 1);
// SURVEY.md 8(e) shards this path over heads / MLP columns, which leaves one [M, hidden] sum after
// o_proj and one after down_proj per layer -- 512 KiB at the headline shape, far below the size
// where a ring pays, and latency-bound in NCCL (~20 us each, 64 per step).  This kernel does the
// exchange itself through CUDA-IPC mapped peer buffers (NVSwitch gives every pair full bandwidth):
//
//   CTA b owns slice b of the vector on EVERY rank.
//   1. copy my slice of x into my half of the symmetric staging buffer   (local HBM)
//   2. release-store the call's epoch into flags[b][rank] of every peer  (NVLink, 4 B each)
//   3. acquire-spin until every peer's epoch shows up in my flags[b][*]
//   4. read slice b from every peer's staging buffer (16-byte NVLink loads), add in RANK ORDER in
//      fp32 (bit-identical result on every rank), write x in place.
//
// Two staging halves alternate by epoch parity, so no trailing barrier is needed: a peer can only
// overwrite half h two calls later, and it cannot pass the barrier of the call in between before
// I have signalled it, i.e. before I finished reading.  The epoch lives in device memory (one
// counter per CTA), so the kernel replays inside a CUDA graph unchanged.
#include <cstring>

#include "common.cuh"
#include "../../include/b200_decoder_ops.h"

namespace b200 {
namespace ar {

constexpr int kMaxWorld = 8;
constexpr int kThreads = 512;

struct Params {
  void* x;                    // [numel] in place
  char* symm[kMaxWorld];      // staging buffers of every rank (peer-mapped), 2 * half_bytes each
  uint32_t* flags[kMaxWorld];  // [grid][world] epoch slots of every rank (peer-mapped)
  uint32_t* epoch;            // local [grid]
  int64_t nvec;               // 16-byte vectors in x
  int64_t half_bytes;
  int rank, world;
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
// peer data: a COHERENT 128-bit load (relaxed, system scope).  The staging halves are rewritten by
// the peers every second call while kernels of this rank may still be resident (PDL), so the
// non-coherent .nc path -- which presumes the data is read-only for the kernel's lifetime -- is not
// allowed here; the acquire + __syncthreads below order these loads after the peer's release.
__device__ __forceinline__ uint4 ld_peer_u4(const void* p) {
  uint4 r;
  asm volatile("ld.relaxed.sys.global.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p)
               : "memory");
  return r;
}

template <typename T>
__device__ __forceinline__ void accumulate(float (&acc)[16 / sizeof(T)], const uint4& v) {
  const T* e = reinterpret_cast<const T*>(&v);
#pragma unroll
  for (int i = 0; i < (int)(16 / sizeof(T)); ++i) acc[i] += to_float(e[i]);
}

template <typename T, int WORLD>
__global__ void __launch_bounds__(kThreads) allreduce_oneshot_kernel(const Params p) {
  pdl_trigger();
  pdl_wait();
  constexpr int VE = 16 / (int)sizeof(T);
  const int b = blockIdx.x;
  const uint32_t e = p.epoch[b] + 1u;
  // slices are whole 128-byte lines (8 vectors): no two CTAs ever share a line of a peer's staging buffer
  const int64_t per = ((p.nvec + gridDim.x - 1) / gridDim.x + 7) / 8 * 8;
  const int64_t lo = b * per;
  const int64_t hi = lo + per < p.nvec ? lo + per : p.nvec;  // (lo may exceed nvec: empty slice, flags still exchanged)
  const int64_t half_off = (int64_t)(e & 1u) * p.half_bytes;
  uint4* xv = reinterpret_cast<uint4*>(p.x);
  uint4* mine = reinterpret_cast<uint4*>(p.symm[p.rank] + half_off);

  // 1. publish my slice.  No per-thread fence: bar.sync orders every thread's stores before the
  // signalling threads' st.release.sys, which is cumulative over them (PTX memory model).
  for (int64_t v = lo + threadIdx.x; v < hi; v += kThreads) mine[v] = xv[v];
  __syncthreads();
  // 2. + 3. per-CTA flag exchange with every peer
  if (threadIdx.x < WORLD && (int)threadIdx.x != p.rank) {
    const int peer = threadIdx.x;
    st_release_sys(p.flags[peer] + (size_t)b * WORLD + p.rank, e);
    const uint32_t* slot = p.flags[p.rank] + (size_t)b * WORLD + peer;
    uint32_t spins = 0;
    while ((int32_t)(ld_acquire_sys(slot) - e) < 0) {
      if (++spins > (1u << 27)) __trap();  // a dead peer must fault this rank, not hang the box
    }
  }
  __syncthreads();
  // 4. gather + sum in rank order
  for (int64_t v = lo + threadIdx.x; v < hi; v += kThreads) {
    uint4 part[WORLD];
#pragma unroll
    for (int r = 0; r < WORLD; ++r)
      part[r] = (r == p.rank) ? xv[v] : ld_peer_u4(reinterpret_cast<const uint4*>(p.symm[r] + half_off) + v);
    float acc[VE];
#pragma unroll
    for (int i = 0; i < VE; ++i) acc[i] = 0.f;
#pragma unroll
    for (int r = 0; r < WORLD; ++r) accumulate<T>(acc, part[r]);
    uint4 o;
    T* oe = reinterpret_cast<T*>(&o);
#pragma unroll
    for (int i = 0; i < VE; ++i) oe[i] = from_float<T>(acc[i]);
    xv[v] = o;
  }
  __syncthreads();
  if (threadIdx.x == 0) p.epoch[b] = e;
}

template <typename T>
static int launch(const Params& p, int grid, cudaStream_t st) {
  switch (p.world) {
    case 2: B200_CUDA_OK(launch_pdl(allreduce_oneshot_kernel<T, 2>, dim3(grid), dim3(kThreads), 0, st, p)); break;
    case 4: B200_CUDA_OK(launch_pdl(allreduce_oneshot_kernel<T, 4>, dim3(grid), dim3(kThreads), 0, st, p)); break;
    case 8: B200_CUDA_OK(launch_pdl(allreduce_oneshot_kernel<T, 8>, dim3(grid), dim3(kThreads), 0, st, p)); break;
    default: return set_error(B200_ERR_INVALID_ARGUMENT, "allreduce: world size must be 2, 4 or 8, got %d", p.world);
  }
  return check_launch("allreduce_oneshot_kernel");
}

}  // namespace ar
}  // namespace b200

extern "C" {

int b200_ipc_alloc(int64_t bytes, void** ptr, unsigned char* handle64) {
  using namespace b200;
  B200_CHECK_ARG(bytes > 0 && ptr && handle64, "ipc_alloc: bad arguments");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle is 64 bytes");
  void* d = nullptr;
  B200_CUDA_OK(cudaMalloc(&d, (size_t)bytes));
  B200_CUDA_OK(cudaMemset(d, 0, (size_t)bytes));
  B200_CUDA_OK(cudaDeviceSynchronize());
  cudaIpcMemHandle_t h;
  B200_CUDA_OK(cudaIpcGetMemHandle(&h, d));
  memcpy(handle64, &h, 64);
  *ptr = d;
  return B200_OK;
}

int b200_ipc_open(const unsigned char* handle64, void** ptr) {
  using namespace b200;
  B200_CHECK_ARG(ptr && handle64, "ipc_open: bad arguments");
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, 64);
  B200_CUDA_OK(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return B200_OK;
}

int b200_ipc_close(void* ptr) {
  using namespace b200;
  B200_CUDA_OK(cudaIpcCloseMemHandle(ptr));
  return B200_OK;
}

int b200_ipc_free(void* ptr) {
  using namespace b200;
  B200_CUDA_OK(cudaFree(ptr));
  return B200_OK;
}

int b200_allreduce_sum(void* x, int64_t numel, int dtype, void* const* symm_ptrs, void* const* flag_ptrs,
                       void* epoch, int rank, int world, int64_t half_bytes, int grid, void* stream) {
  using namespace b200;
  if (numel == 0) return B200_OK;
  B200_CHECK_ARG(x && symm_ptrs && flag_ptrs && epoch && numel > 0, "allreduce: bad arguments");
  B200_CHECK_ARG(rank >= 0 && rank < world && world <= ar::kMaxWorld, "allreduce: rank %d of world %d", rank, world);
  B200_CHECK_ARG(grid > 0 && grid <= 1024, "allreduce: grid %d out of range", grid);
  const int64_t bytes = numel * (int64_t)dtype_size(dtype);
  B200_CHECK_ARG(bytes % 16 == 0 && ((uintptr_t)x & 15) == 0, "allreduce: x must be 16-byte aligned and a multiple of 16 bytes");
  B200_CHECK_ARG(bytes <= half_bytes, "allreduce: %lld bytes exceed the staging half (%lld)", (long long)bytes,
                 (long long)half_bytes);
  ar::Params p;
  p.x = x;
  for (int r = 0; r < world; ++r) {
    p.symm[r] = reinterpret_cast<char*>(symm_ptrs[r]);
    p.flags[r] = reinterpret_cast<uint32_t*>(flag_ptrs[r]);
  }
  p.epoch = reinterpret_cast<uint32_t*>(epoch);
  p.nvec = bytes / 16;
  p.half_bytes = half_bytes;
  p.rank = rank;
  p.world = world;
  cudaStream_t st = (cudaStream_t)stream;
  return B200_DISPATCH_DTYPE(dtype, T, [&]() -> int { return ar::launch<T>(p, grid, st); });
}

}  // extern "C"
