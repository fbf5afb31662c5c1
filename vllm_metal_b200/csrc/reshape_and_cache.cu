// reshape_and_cache / copy_blocks for sm_100a.
//
// Replaces vllm_metal/metal/kernels_v2/reshape_and_cache.metal:36-96 (one
// threadgroup per token, strided element copy).  Here: grid (token, K|V); each
// thread moves 16-byte vectors of the contiguous [H_kv, D] row, so a 2 KiB row
// (8 heads x 128 x bf16) is 128 coalesced 128-bit stores.  HBM/launch bound:
// 2*(2*H_kv*D*esize) + 8 bytes per token.
#include "common.cuh"

namespace b200 {

template <int VEC_BYTES>
__global__ void __launch_bounds__(128)
reshape_and_cache_kernel(const uint8_t* __restrict__ key, const uint8_t* __restrict__ value,
                         uint8_t* __restrict__ key_cache, uint8_t* __restrict__ value_cache,
                         const int64_t* __restrict__ slot_mapping, int row_bytes, int block_size,
                         int64_t num_slots, int64_t block_stride_bytes,
                         int64_t token_stride_bytes) {
  pdl_trigger();
  pdl_wait();
  const int t = blockIdx.x;
  const int64_t slot = slot_mapping[t];
  if (slot < 0 || slot >= num_slots) return;  // padding (reshape_and_cache.metal:61-65)
  const int64_t blk = slot / block_size;
  const int64_t off = slot - blk * block_size;
  const uint8_t* src = (blockIdx.y == 0 ? key : value) + (int64_t)t * row_bytes;
  uint8_t* dst = (blockIdx.y == 0 ? key_cache : value_cache) + blk * block_stride_bytes +
                 off * token_stride_bytes;
  if (VEC_BYTES == 16) {
    const uint4* s = reinterpret_cast<const uint4*>(src);
    uint4* d = reinterpret_cast<uint4*>(dst);
    const int n = row_bytes >> 4;
    for (int i = threadIdx.x; i < n; i += blockDim.x) d[i] = s[i];
  } else {
    const uint16_t* s = reinterpret_cast<const uint16_t*>(src);
    uint16_t* d = reinterpret_cast<uint16_t*>(dst);
    const int n = row_bytes >> 1;
    for (int i = threadIdx.x; i < n; i += blockDim.x) d[i] = s[i];
  }
}

__global__ void __launch_bounds__(256)
copy_blocks_kernel(uint8_t* __restrict__ key_cache, uint8_t* __restrict__ value_cache,
                   const int32_t* __restrict__ src, const int32_t* __restrict__ dst,
                   int64_t block_bytes, int64_t num_blocks) {
  pdl_trigger();
  pdl_wait();
  const int pair = blockIdx.x;
  const int64_t s = src[pair], d = dst[pair];
  if (s < 0 || d < 0 || s >= num_blocks || d >= num_blocks || s == d) return;
  uint8_t* base = blockIdx.y == 0 ? key_cache : value_cache;
  const uint4* sp = reinterpret_cast<const uint4*>(base + s * block_bytes);
  uint4* dp = reinterpret_cast<uint4*>(base + d * block_bytes);
  const int64_t n = block_bytes >> 4;
  for (int64_t i = blockIdx.z * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.z * blockDim.x)
    dp[i] = sp[i];
}

}  // namespace b200

extern "C" int b200_reshape_and_cache(const void* key, const void* value, void* key_cache,
                                      void* value_cache, const int64_t* slot_mapping,
                                      int num_tokens, int num_kv_heads, int head_dim,
                                      int block_size, int64_t num_blocks,
                                      int64_t kv_block_stride, int64_t kv_token_stride,
                                      int dtype, void* stream) {
  using namespace b200;
  B200_CHECK_ARG(num_tokens >= 0, "reshape_and_cache: num_tokens must be >= 0, got %d", num_tokens);
  if (num_tokens == 0) return B200_OK;
  B200_CHECK_ARG(key && value && key_cache && value_cache && slot_mapping,
                 "reshape_and_cache: null pointer argument");
  B200_CHECK_ARG(num_kv_heads > 0 && head_dim > 0 && block_size > 0 && num_blocks > 0,
                 "reshape_and_cache: non-positive dimension (kv_heads=%d head_dim=%d block=%d)",
                 num_kv_heads, head_dim, block_size);
  B200_CHECK_ARG(dtype == B200_F32 || dtype == B200_F16 || dtype == B200_BF16,
                 "reshape_and_cache: unsupported dtype code %d", dtype);
  const int64_t esize = (int64_t)dtype_size(dtype);
  const int64_t row_elems = (int64_t)num_kv_heads * head_dim;
  B200_CHECK_ARG(kv_token_stride >= row_elems && kv_block_stride >= kv_token_stride * block_size,
                 "reshape_and_cache: cache strides overlap (block=%lld token=%lld row=%lld)",
                 (long long)kv_block_stride, (long long)kv_token_stride, (long long)row_elems);
  const int row_bytes = (int)(row_elems * esize);
  const int64_t bsb = kv_block_stride * esize, tsb = kv_token_stride * esize;
  const bool vec16 = (row_bytes % 16 == 0) && (bsb % 16 == 0) && (tsb % 16 == 0) &&
                     ((uintptr_t)key % 16 == 0) && ((uintptr_t)value % 16 == 0) &&
                     ((uintptr_t)key_cache % 16 == 0) && ((uintptr_t)value_cache % 16 == 0);
  dim3 grid(num_tokens, 2);
  cudaStream_t st = (cudaStream_t)stream;
  auto* k8 = (const uint8_t*)key;
  auto* v8 = (const uint8_t*)value;
  auto* kc8 = (uint8_t*)key_cache;
  auto* vc8 = (uint8_t*)value_cache;
  const int64_t num_slots = num_blocks * block_size;
  if (vec16)
    launch_pdl(reshape_and_cache_kernel<16>, dim3(grid), dim3(128), 0, st, k8, v8, kc8, vc8, slot_mapping, row_bytes,
                                                       block_size, num_slots, bsb, tsb);
  else
    launch_pdl(reshape_and_cache_kernel<2>, dim3(grid), dim3(128), 0, st, k8, v8, kc8, vc8, slot_mapping, row_bytes,
                                                      block_size, num_slots, bsb, tsb);
  return check_launch("reshape_and_cache_kernel");
}

extern "C" int b200_copy_blocks(void* key_cache, void* value_cache, const int32_t* src,
                                const int32_t* dst, int num_pairs, int64_t block_bytes,
                                int64_t num_blocks, void* stream) {
  using namespace b200;
  if (num_pairs == 0) return B200_OK;
  B200_CHECK_ARG(num_pairs > 0 && key_cache && value_cache && src && dst,
                 "copy_blocks: bad arguments");
  B200_CHECK_ARG(block_bytes > 0 && block_bytes % 16 == 0,
                 "copy_blocks: block_bytes must be a positive multiple of 16, got %lld",
                 (long long)block_bytes);
  int z = (int)((block_bytes / 16 + 255) / 256);
  if (z > 64) z = 64;
  dim3 grid(num_pairs, 2, z);
  launch_pdl(copy_blocks_kernel, dim3(grid), dim3(256), 0, (cudaStream_t)stream, 
      (uint8_t*)key_cache, (uint8_t*)value_cache, src, dst, block_bytes, num_blocks);
  return check_launch("copy_blocks_kernel");
}
