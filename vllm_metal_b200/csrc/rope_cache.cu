// Fused per-head RMSNorm -> RoPE -> KV-cache write.
//
// One launch does what the reference's sdpa_forward does in five steps between the q/k/v
// projections and the attention kernel (vllm_metal/attention/impls/sdpa.py:348-368 q/k norm +
// RoPE, :542-544 cast, :638 reshape_and_cache -> kernels_v2/reshape_and_cache.metal:36-96):
//
//   q[t,h,:]  <- rope(norm(q[t,h,:]))                       in place (strided rows of the qkv buffer)
//   k[t,h,:]  <- rope(norm(k[t,h,:]))                       in place AND to key_cache[slot[t]]
//   value_cache[slot[t]] <- v[t,:,:]                        bit-exact copy
//
// The arithmetic is the composition of the standalone ops bit for bit: the norm result is
// rounded to T before the rotation reads it, angles come from the same fp32 formula as
// rope_kernel, slot < 0 (padding) skips the cache write only.  One CTA per token, one warp per
// head; a head row is staged in shared memory as fp32 so any head_dim / rot_dim / pair layout
// works.  HBM bytes per token: (Hq + 2*Hkv)*D*esize read, (Hq + Hkv)*D*esize in-place write,
// 2*Hkv*D*esize cache write.
#include "common.cuh"
#include "../../include/b200_decoder_ops.h"

namespace b200 {

constexpr int RC_MAX_D = 256;
constexpr int RC_WARPS = 8;

template <typename T>
__global__ void __launch_bounds__(RC_WARPS * 32)
norm_rope_cache_kernel(T* __restrict__ q, T* __restrict__ k, const T* __restrict__ v,
                       const T* __restrict__ q_norm_w, const T* __restrict__ k_norm_w,
                       const int32_t* __restrict__ positions, const int64_t* __restrict__ slot_mapping,
                       T* __restrict__ key_cache, T* __restrict__ value_cache, int num_q_heads,
                       int num_kv_heads, int head_dim, int rot_dim, float log2_base, int traditional,
                       float eps, int64_t q_stride, int64_t k_stride, int64_t v_stride, int block_size,
                       int64_t num_slots, int64_t kv_block_stride, int64_t kv_token_stride) {
  pdl_trigger();
  pdl_wait();
  __shared__ float rows[RC_WARPS][RC_MAX_D];
  const int t = blockIdx.x;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const float pos = positions ? (float)positions[t] : 0.f;
  const int64_t slot = slot_mapping ? slot_mapping[t] : -1;
  const bool write_cache = slot >= 0 && slot < num_slots;
  int64_t cache_off = 0;
  if (write_cache) {
    const int64_t blk = slot / block_size;
    cache_off = blk * kv_block_stride + (slot - blk * block_size) * kv_token_stride;
  }
  const int half = rot_dim >> 1;
  float* row = rows[warp];
  const int total = num_q_heads + 2 * num_kv_heads;
  // heads are spread over gridDim.y CTAs per token so a 64-token decode step fills the machine
  for (int h = blockIdx.y * RC_WARPS + warp; h < total; h += RC_WARPS * gridDim.y) {
    if (h >= num_q_heads + num_kv_heads) {
      // ---- V head: bit-exact copy into the cache ----
      if (!write_cache) continue;
      const int hv = h - num_q_heads - num_kv_heads;
      const T* src = v + (int64_t)t * v_stride + (int64_t)hv * head_dim;
      T* dst = value_cache + cache_off + (int64_t)hv * head_dim;
      if (sizeof(T) == 2 && head_dim % 8 == 0 && ((uintptr_t)src % 16 == 0) && ((uintptr_t)dst % 16 == 0)) {
        for (int i = lane; i < head_dim / 8; i += 32)
          reinterpret_cast<uint4*>(dst)[i] = reinterpret_cast<const uint4*>(src)[i];
      } else {
        for (int i = lane; i < head_dim; i += 32) dst[i] = src[i];
      }
      continue;
    }
    const bool is_q = h < num_q_heads;
    T* base = is_q ? q + (int64_t)t * q_stride + (int64_t)h * head_dim
                   : k + (int64_t)t * k_stride + (int64_t)(h - num_q_heads) * head_dim;
    const T* nw = is_q ? q_norm_w : k_norm_w;
    const bool do_norm = is_q ? (q_norm_w != nullptr) : (k_norm_w != nullptr);
    // ---- load (+ per-head RMSNorm, rounded to T like the standalone op) ----
    float ss = 0.f;
    for (int i = lane; i < head_dim; i += 32) {
      const float x = to_float(base[i]);
      row[i] = x;
      ss += x * x;
    }
    if (do_norm) {
      ss = warp_sum(ss);
      const float inv = rsqrtf(ss / (float)head_dim + eps);
      for (int i = lane; i < head_dim; i += 32)
        row[i] = to_float(from_float<T>(row[i] * inv * to_float(nw[i])));
    }
    __syncwarp();
    T* cdst = (!is_q && write_cache) ? key_cache + cache_off + (int64_t)(h - num_q_heads) * head_dim : nullptr;
    // ---- rotate the first rot_dim dims ----
    for (int i = lane; i < half; i += 32) {
      const float inv_freq = exp2f(-log2_base * (2.f * (float)i / (float)rot_dim));
      float sn, cs;
      sincosf(pos * inv_freq, &sn, &cs);
      const int i0 = traditional ? 2 * i : i;
      const int i1 = traditional ? 2 * i + 1 : i + half;
      const float a = row[i0], b = row[i1];
      const T o0 = from_float<T>(a * cs - b * sn), o1 = from_float<T>(b * cs + a * sn);
      base[i0] = o0;
      base[i1] = o1;
      if (cdst) {
        cdst[i0] = o0;
        cdst[i1] = o1;
      }
    }
    // ---- pass-through dims (rot_dim < head_dim, or no RoPE at all) ----
    for (int i = rot_dim + lane; i < head_dim; i += 32) {
      const T o = from_float<T>(row[i]);
      if (do_norm) base[i] = o;
      if (cdst) cdst[i] = o;
    }
    __syncwarp();
  }
}

}  // namespace b200

extern "C" int b200_norm_rope_cache(void* q, void* k, const void* v, const void* q_norm_weight,
                                    const void* k_norm_weight, const int32_t* positions,
                                    const int64_t* slot_mapping, void* key_cache, void* value_cache,
                                    int num_tokens, int num_q_heads, int num_kv_heads, int head_dim,
                                    int rot_dim, float base, int traditional, float eps,
                                    int64_t q_stride, int64_t k_stride, int64_t v_stride,
                                    int block_size, int64_t num_blocks, int64_t kv_block_stride,
                                    int64_t kv_token_stride, int dtype, void* stream) {
  using namespace b200;
  B200_CHECK_ARG(num_tokens >= 0, "norm_rope_cache: num_tokens must be >= 0, got %d", num_tokens);
  if (num_tokens == 0) return B200_OK;
  B200_CHECK_ARG(q && k && v && key_cache && value_cache && slot_mapping,
                 "norm_rope_cache: null pointer argument");
  B200_CHECK_ARG(num_q_heads > 0 && num_kv_heads > 0 && head_dim > 0 && head_dim <= RC_MAX_D,
                 "norm_rope_cache: head_dim must be in [1, %d], got %d (heads %d/%d)", RC_MAX_D, head_dim,
                 num_q_heads, num_kv_heads);
  B200_CHECK_ARG(rot_dim >= 0 && rot_dim % 2 == 0 && rot_dim <= head_dim,
                 "norm_rope_cache: rot_dim (%d) must be even and <= head_dim (%d)", rot_dim, head_dim);
  B200_CHECK_ARG(rot_dim == 0 || (positions != nullptr && base > 0.f),
                 "norm_rope_cache: RoPE needs positions and a positive base");
  B200_CHECK_ARG(block_size > 0 && num_blocks > 0, "norm_rope_cache: bad cache geometry");
  const int64_t row_elems = (int64_t)num_kv_heads * head_dim;
  B200_CHECK_ARG(kv_token_stride >= row_elems && kv_block_stride >= kv_token_stride * block_size,
                 "norm_rope_cache: cache strides overlap (block=%lld token=%lld row=%lld)",
                 (long long)kv_block_stride, (long long)kv_token_stride, (long long)row_elems);
  cudaStream_t st = (cudaStream_t)stream;
  const float l2b = rot_dim > 0 ? log2f(base) : 0.f;
  return B200_DISPATCH_DTYPE(dtype, T, [&]() -> int {
    const int total_heads = num_q_heads + 2 * num_kv_heads;
    const int ysplit = num_tokens >= 1024 ? 1 : (total_heads + RC_WARPS - 1) / RC_WARPS;
    launch_pdl(norm_rope_cache_kernel<T>, dim3(num_tokens, ysplit), dim3(RC_WARPS * 32), 0, st, 
        (T*)q, (T*)k, (const T*)v, (const T*)q_norm_weight, (const T*)k_norm_weight, positions,
        slot_mapping, (T*)key_cache, (T*)value_cache, num_q_heads, num_kv_heads, head_dim, rot_dim, l2b,
        traditional, eps, q_stride, k_stride, v_stride, block_size, num_blocks * block_size,
        kv_block_stride, kv_token_stride);
    return check_launch("norm_rope_cache_kernel");
  });
}
