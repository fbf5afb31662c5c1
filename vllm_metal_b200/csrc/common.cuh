// Shared helpers for the sm_100a kernels behind include/b200_paged_ops.h.
#pragma once

#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <cstdarg>
#include <cstdio>

#include "../../include/b200_paged_ops.h"

namespace b200 {

// ---- thread-local error string (c_abi.cu owns the storage) -----------------
int set_error(int code, const char* fmt, ...);
void clear_error();

#define B200_CHECK_ARG(cond, ...)                                   \
  do {                                                              \
    if (!(cond)) return ::b200::set_error(B200_ERR_INVALID_ARGUMENT, __VA_ARGS__); \
  } while (0)

#define B200_CUDA_OK(expr)                                                          \
  do {                                                                              \
    cudaError_t e__ = (expr);                                                       \
    if (e__ != cudaSuccess)                                                         \
      return ::b200::set_error(B200_ERR_CUDA, "%s failed: %s (%s:%d)", #expr,       \
                               cudaGetErrorString(e__), __FILE__, __LINE__);        \
  } while (0)

inline int check_launch(const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess)
    return set_error(B200_ERR_CUDA, "launch of %s failed: %s", what, cudaGetErrorString(e));
  return B200_OK;
}

int sm_count();  // cached per process (current device at first call)

constexpr float kLog2e = 1.4426950408889634f;

// ---- dtype helpers ---------------------------------------------------------
template <typename T>
struct TypeTag;
template <>
struct TypeTag<float> {
  static constexpr int value = B200_F32;
};
template <>
struct TypeTag<__half> {
  static constexpr int value = B200_F16;
};
template <>
struct TypeTag<__nv_bfloat16> {
  static constexpr int value = B200_BF16;
};

inline size_t dtype_size(int dtype) { return dtype == B200_F32 ? 4 : 2; }

__device__ __forceinline__ float to_float(float v) { return v; }
__device__ __forceinline__ float to_float(__half v) { return __half2float(v); }
__device__ __forceinline__ float to_float(__nv_bfloat16 v) { return __bfloat162float(v); }

template <typename T>
__device__ __forceinline__ T from_float(float v);
template <>
__device__ __forceinline__ float from_float<float>(float v) {
  return v;
}
template <>
__device__ __forceinline__ __half from_float<__half>(float v) {
  return __float2half_rn(v);
}
template <>
__device__ __forceinline__ __nv_bfloat16 from_float<__nv_bfloat16>(float v) {
  return __float2bfloat16_rn(v);
}

// Dispatch a generic lambda on a b200_dtype.  `fn` receives a value of the
// element type as a tag.
#define B200_DISPATCH_DTYPE(dtype, T, ...)                                          \
  [&]() -> int {                                                                    \
    switch (dtype) {                                                                \
      case B200_F32: {                                                              \
        using T = float;                                                            \
        return __VA_ARGS__();                                                       \
      }                                                                             \
      case B200_F16: {                                                              \
        using T = __half;                                                           \
        return __VA_ARGS__();                                                       \
      }                                                                             \
      case B200_BF16: {                                                             \
        using T = __nv_bfloat16;                                                    \
        return __VA_ARGS__();                                                       \
      }                                                                             \
      default:                                                                      \
        return ::b200::set_error(B200_ERR_UNSUPPORTED, "unsupported dtype code %d", \
                                 (int)(dtype));                                     \
    }                                                                               \
  }()

// ---- warp helpers ----------------------------------------------------------
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// 128-bit streaming load that does not allocate in L1 (data read once).
__device__ __forceinline__ uint4 ld_stream_u4(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}

// ---- programmatic dependent launch (PDL) -----------------------------------
// Every kernel of the decode step is launched with the programmatic-stream-serialization
// attribute: its CTAs may become resident (barrier init, TMEM alloc, weight prefetch) while the
// previous kernel in the stream drains, and block in pdl_wait() until that kernel has completed
// and its writes are visible.  RULE: a kernel launched through launch_pdl() must execute
// pdl_wait() in every thread before it reads anything a previous kernel wrote and before it
// writes anything at all.  pdl_trigger() lets the *next* kernel start its prologue.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

bool pdl_enabled();  // c_abi.cu (b200_set_pdl); default on

template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                              Args... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
}

}  // namespace b200
