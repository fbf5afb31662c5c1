// Elementwise / reduction ops either side of the hot path (include/b200_decoder_ops.h):
// RMSNorm (+ fused residual add), RoPE, SiLU*mul, greedy argmax, residual add.
// All HBM/launch bound; one CTA per row, fp32 math, one rounding to the I/O type.
#include "common.cuh"
#include "../../include/b200_decoder_ops.h"

namespace b200 {

__device__ __forceinline__ float block_sum(float v, float* red) {
  v = warp_sum(v);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) red[warp] = v;
  __syncthreads();
  const int nw = (blockDim.x + 31) >> 5;
  float t = lane < nw ? red[lane] : 0.f;
  t = warp_sum(t);
  __syncthreads();
  return t;
}

template <typename T>
__global__ void __launch_bounds__(256)
rms_norm_kernel(const T* __restrict__ x, T* __restrict__ residual, const T* __restrict__ weight,
                T* __restrict__ out, int dim, int64_t x_stride, int64_t res_stride,
                int64_t out_stride, float eps) {
  pdl_trigger();
  pdl_wait();
  __shared__ float red[32];
  const int r = blockIdx.x;
  const T* xr = x + (int64_t)r * x_stride;
  T* rr = residual ? residual + (int64_t)r * res_stride : nullptr;
  float ss = 0.f;
  for (int d = threadIdx.x; d < dim; d += blockDim.x) {
    float v = to_float(xr[d]);
    if (rr) {
      v = to_float(from_float<T>(v + to_float(rr[d])));  // residual stream kept in T
      rr[d] = from_float<T>(v);
    }
    ss += v * v;
  }
  ss = block_sum(ss, red);
  const float inv = rsqrtf(ss / (float)dim + eps);
  T* orow = out + (int64_t)r * out_stride;
  for (int d = threadIdx.x; d < dim; d += blockDim.x) {
    const float v = rr ? to_float(rr[d]) : to_float(xr[d]);
    const float w = weight ? to_float(weight[d]) : 1.f;
    orow[d] = from_float<T>(v * inv * w);
  }
}

// Vectorised row kernel for 16-bit types: one 16-byte load per thread per pass, the row
// stays in registers between the sum-of-squares and the scale (one HBM read, one write).
// Requires dim % 8 == 0, 16-byte aligned rows, dim <= blockDim.x * 8 * NV.
template <typename T, int NV>
__global__ void __launch_bounds__(1024)
rms_norm_vec_kernel(const T* __restrict__ x, T* __restrict__ residual, const T* __restrict__ weight,
                    T* __restrict__ out, int dim, int64_t x_stride, int64_t res_stride,
                    int64_t out_stride, float eps, int out_f16) {
  pdl_trigger();
  pdl_wait();
  __shared__ float red[32];
  const int r = blockIdx.x;
  const int nvec = dim >> 3;
  const uint4* xr = reinterpret_cast<const uint4*>(x + (int64_t)r * x_stride);
  uint4* rr = residual ? reinterpret_cast<uint4*>(residual + (int64_t)r * res_stride) : nullptr;
  float v[NV][8];
  float ss = 0.f;
#pragma unroll
  for (int j = 0; j < NV; ++j) {
    const int i = threadIdx.x + j * blockDim.x;
    if (i < nvec) {
      uint4 a = xr[i];
      const T* ae = reinterpret_cast<const T*>(&a);
      if (rr) {
        uint4 b = rr[i];
        const T* be = reinterpret_cast<const T*>(&b);
        uint4 c;
        T* ce = reinterpret_cast<T*>(&c);
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          ce[e] = from_float<T>(to_float(ae[e]) + to_float(be[e]));  // residual stream kept in T
          v[j][e] = to_float(ce[e]);
        }
        rr[i] = c;
      } else {
#pragma unroll
        for (int e = 0; e < 8; ++e) v[j][e] = to_float(ae[e]);
      }
#pragma unroll
      for (int e = 0; e < 8; ++e) ss += v[j][e] * v[j][e];
    }
  }
  ss = block_sum(ss, red);
  const float inv = rsqrtf(ss / (float)dim + eps);
  uint4* orow = reinterpret_cast<uint4*>(out + (int64_t)r * out_stride);
  const uint4* wv = reinterpret_cast<const uint4*>(weight);
#pragma unroll
  for (int j = 0; j < NV; ++j) {
    const int i = threadIdx.x + j * blockDim.x;
    if (i < nvec) {
      uint4 c;
      T* ce = reinterpret_cast<T*>(&c);
      if (weight) {
        uint4 w = __ldg(wv + i);
        const T* we = reinterpret_cast<const T*>(&w);
#pragma unroll
        for (int e = 0; e < 8; ++e) ce[e] = from_float<T>(v[j][e] * inv * to_float(we[e]));
      } else {
#pragma unroll
        for (int e = 0; e < 8; ++e) ce[e] = from_float<T>(v[j][e] * inv);
      }
      if (out_f16) {  // the value rounded to T, stored as fp16 (exact below 65504, saturating): the stream-K GEMM's x
        __half* he = reinterpret_cast<__half*>(&c);
#pragma unroll
        for (int e = 0; e < 8; ++e) he[e] = __float2half_rn(fminf(fmaxf(to_float(ce[e]), -65504.f), 65504.f));
      }
      orow[i] = c;
    }
  }
}

// One CTA per token; threads over (head, pair).
template <typename T>
__global__ void __launch_bounds__(256)
rope_kernel(T* __restrict__ q, T* __restrict__ k, const int32_t* __restrict__ positions,
            int num_q_heads, int num_kv_heads, int head_dim, int rot_dim, float log2_base,
            int traditional, int64_t q_stride, int64_t k_stride) {
  pdl_trigger();
  pdl_wait();
  const int t = blockIdx.x;
  const float pos = (float)positions[t];
  const int half = rot_dim >> 1;
  const int total = (num_q_heads + (k ? num_kv_heads : 0)) * half;
  for (int idx = threadIdx.x; idx < total; idx += blockDim.x) {
    const int h = idx / half, i = idx - h * half;
    T* base = h < num_q_heads ? q + (int64_t)t * q_stride + (int64_t)h * head_dim
                              : k + (int64_t)t * k_stride + (int64_t)(h - num_q_heads) * head_dim;
    const float inv_freq = exp2f(-log2_base * (2.f * (float)i / (float)rot_dim));
    float sn, cs;
    sincosf(pos * inv_freq, &sn, &cs);
    const int i0 = traditional ? 2 * i : i;
    const int i1 = traditional ? 2 * i + 1 : i + half;
    const float a = to_float(base[i0]), b = to_float(base[i1]);
    base[i0] = from_float<T>(a * cs - b * sn);
    base[i1] = from_float<T>(b * cs + a * sn);
  }
}

template <typename T>
__global__ void __launch_bounds__(256)
silu_mul_kernel(const T* __restrict__ gate_up, T* __restrict__ out, int inter, int64_t in_stride,
                int64_t out_stride) {
  pdl_trigger();
  pdl_wait();
  const int r = blockIdx.y;
  const T* g = gate_up + (int64_t)r * in_stride;
  const T* u = g + inter;
  T* o = out + (int64_t)r * out_stride;
  constexpr int VE = 16 / (int)sizeof(T);
  // 16-byte path when every row segment is aligned (the decode layout: inter % 8 == 0)
  const bool vec = (inter % VE == 0) && ((reinterpret_cast<uintptr_t>(g) | reinterpret_cast<uintptr_t>(u) |
                                          reinterpret_cast<uintptr_t>(o)) % 16 == 0);
  if (vec) {
    const int nvec = inter / VE;
    for (int v = blockIdx.x * blockDim.x + threadIdx.x; v < nvec; v += gridDim.x * blockDim.x) {
      const uint4 gv = ld_stream_u4(reinterpret_cast<const uint4*>(g) + v);
      const uint4 uv = ld_stream_u4(reinterpret_cast<const uint4*>(u) + v);
      const T* ge = reinterpret_cast<const T*>(&gv);
      const T* ue = reinterpret_cast<const T*>(&uv);
      uint4 ov;
      T* oe = reinterpret_cast<T*>(&ov);
#pragma unroll
      for (int i = 0; i < VE; ++i) {
        const float a = to_float(ge[i]);
        const float s = a / (1.f + __expf(-a));
        oe[i] = from_float<T>(to_float(from_float<T>(s)) * to_float(ue[i]));
      }
      reinterpret_cast<uint4*>(o)[v] = ov;
    }
    return;
  }
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < inter; i += gridDim.x * blockDim.x) {
    const float a = to_float(g[i]);
    const float s = a / (1.f + __expf(-a));
    o[i] = from_float<T>(to_float(from_float<T>(s)) * to_float(u[i]));
  }
}

// Greedy top-1 per row.  A CLUSTER of kArgmaxSplit CTAs shares one row: each CTA scans its slice
// with 16-byte loads, the per-CTA winners meet in rank 0's shared memory through DSMEM.  Ties go
// to the lowest index (mx.argmax / numpy semantics, v1/sampling_batch.py:342-348).
constexpr int kArgmaxSplit = 8;
constexpr int kArgmaxThreads = 256;

__device__ __forceinline__ void argmax_take(float& best, int& bi, float v, int i) {
  if (v > best || (v == best && i < bi)) {
    best = v;
    bi = i;
  }
}

template <typename T>
__global__ void __cluster_dims__(kArgmaxSplit, 1, 1) __launch_bounds__(kArgmaxThreads)
argmax_kernel(const T* __restrict__ logits, int32_t* __restrict__ out, int vocab, int64_t stride) {
  pdl_trigger();
  pdl_wait();
  __shared__ float sv[kArgmaxThreads / 32];
  __shared__ int si[kArgmaxThreads / 32];
  __shared__ float cv[kArgmaxSplit];
  __shared__ int ci[kArgmaxSplit];
  const int r = blockIdx.y;
  const int rank = blockIdx.x;  // == cluster rank (cluster spans x)
  const T* row = logits + (int64_t)r * stride;
  float best = -INFINITY;
  int bi = 0x7fffffff;
  constexpr int VE = 16 / (int)sizeof(T);
  // slice boundaries in units of VE elements so vector loads stay aligned
  const int nvec = vocab / VE;
  const int v_lo = (int)((int64_t)nvec * rank / kArgmaxSplit), v_hi = (int)((int64_t)nvec * (rank + 1) / kArgmaxSplit);
  if ((reinterpret_cast<uintptr_t>(row) & 15) == 0) {
    for (int v = v_lo + threadIdx.x; v < v_hi; v += kArgmaxThreads) {
      const uint4 raw = ld_stream_u4(reinterpret_cast<const uint4*>(row) + v);
      const T* e = reinterpret_cast<const T*>(&raw);
#pragma unroll
      for (int j = 0; j < VE; ++j) argmax_take(best, bi, to_float(e[j]), v * VE + j);
    }
  } else {
    for (int i = v_lo * VE + threadIdx.x; i < v_hi * VE; i += kArgmaxThreads) argmax_take(best, bi, to_float(row[i]), i);
  }
  if (rank == kArgmaxSplit - 1)  // tail that does not fill a vector
    for (int i = nvec * VE + threadIdx.x; i < vocab; i += kArgmaxThreads) argmax_take(best, bi, to_float(row[i]), i);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float ov = __shfl_xor_sync(0xffffffffu, best, o);
    const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
    argmax_take(best, bi, ov, oi);
  }
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) {
    sv[warp] = best;
    si[warp] = bi;
  }
  __syncthreads();
  if (warp == 0) {
    constexpr int nw = kArgmaxThreads / 32;
    best = lane < nw ? sv[lane] : -INFINITY;
    bi = lane < nw ? si[lane] : 0x7fffffff;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ov = __shfl_xor_sync(0xffffffffu, best, o);
      const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
      argmax_take(best, bi, ov, oi);
    }
    if (lane == 0) {
      // publish into rank 0's cv/ci[rank] through distributed shared memory
      uint32_t av = (uint32_t)__cvta_generic_to_shared(&cv[rank]), ai = (uint32_t)__cvta_generic_to_shared(&ci[rank]);
      uint32_t rv, ri;
      asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(rv) : "r"(av), "r"(0));
      asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(ri) : "r"(ai), "r"(0));
      asm volatile("st.shared::cluster.f32 [%0], %1;" ::"r"(rv), "f"(best) : "memory");
      asm volatile("st.shared::cluster.s32 [%0], %1;" ::"r"(ri), "r"(bi) : "memory");
    }
  }
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
  if (rank == 0 && threadIdx.x == 0) {
    best = cv[0];
    bi = ci[0];
#pragma unroll
    for (int j = 1; j < kArgmaxSplit; ++j) argmax_take(best, bi, cv[j], ci[j]);
    out[r] = bi == 0x7fffffff ? 0 : bi;
  }
}

template <typename T>
__global__ void __launch_bounds__(256)
add_inplace_kernel(T* __restrict__ y, const T* __restrict__ x, int dim, int64_t y_stride,
                   int64_t x_stride) {
  pdl_trigger();
  pdl_wait();
  const int r = blockIdx.y;
  T* yr = y + (int64_t)r * y_stride;
  const T* xr = x + (int64_t)r * x_stride;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < dim; i += gridDim.x * blockDim.x)
    yr[i] = from_float<T>(to_float(yr[i]) + to_float(xr[i]));
}

__global__ void advance_decode_kernel(int32_t* __restrict__ input_ids, int32_t* __restrict__ positions,
                                      int32_t* __restrict__ seq_lens, int64_t* __restrict__ slot_mapping,
                                      const int32_t* __restrict__ block_tables,
                                      const int32_t* __restrict__ sampled, int batch, int max_blocks,
                                      int block_size) {
  pdl_trigger();
  pdl_wait();
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= batch) return;
  const int pos = positions[r] + 1;
  positions[r] = pos;
  seq_lens[r] += 1;
  input_ids[r] = sampled[r];
  int page = pos / block_size;
  if (page >= max_blocks) page = max_blocks - 1;
  slot_mapping[r] = (int64_t)block_tables[(int64_t)r * max_blocks + page] * block_size + pos % block_size;
}

}  // namespace b200

extern "C" int b200_advance_decode(int32_t* input_ids, int32_t* positions, int32_t* seq_lens,
                                   int64_t* slot_mapping, const int32_t* block_tables,
                                   const int32_t* sampled, int batch, int max_blocks, int block_size,
                                   void* stream) {
  using namespace b200;
  if (batch == 0) return B200_OK;
  B200_CHECK_ARG(batch > 0 && max_blocks > 0 && block_size > 0 && input_ids && positions && seq_lens &&
                     slot_mapping && block_tables && sampled,
                 "advance_decode: bad arguments");
  launch_pdl(advance_decode_kernel, dim3((batch + 127) / 128), dim3(128), 0, (cudaStream_t)stream, 
      input_ids, positions, seq_lens, slot_mapping, block_tables, sampled, batch, max_blocks, block_size);
  return check_launch("advance_decode_kernel");
}

extern "C" int b200_rms_norm(const void* x, void* residual, const void* weight, void* out,
                             int rows, int dim, int64_t x_stride, int64_t res_stride,
                             int64_t out_stride, float eps, int dtype, int out_dtype, void* stream) {
  using namespace b200;
  if (rows == 0) return B200_OK;
  B200_CHECK_ARG(rows > 0 && dim > 0 && x && out, "rms_norm: bad arguments");
  const int out_f16 = (out_dtype == B200_F16 && dtype == B200_BF16) ? 1 : 0;
  B200_CHECK_ARG(out_dtype == dtype || out_f16, "rms_norm: out dtype must be the input dtype, or fp16 for bf16 input");
  cudaStream_t st = (cudaStream_t)stream;
  const int threads = dim >= 1024 ? 256 : (dim >= 256 ? 128 : 32);
  const bool aligned = dim % 8 == 0 && dtype != B200_F32 && x_stride % 8 == 0 && res_stride % 8 == 0 &&
                       out_stride % 8 == 0 && (uintptr_t)x % 16 == 0 && (uintptr_t)out % 16 == 0 &&
                       (uintptr_t)residual % 16 == 0 && (uintptr_t)weight % 16 == 0;
  if (aligned && dim >= 512 && dim <= 1024 * 8 * 2) {
    // threads: one 16-byte vector each (two for dim > 8192), rounded up to whole warps
    const int nvec = dim / 8;
    const int nv = nvec > 1024 ? 2 : 1;
    const int vt = ((nvec + nv - 1) / nv + 31) / 32 * 32;
    if (dtype == B200_BF16) {
      using T = __nv_bfloat16;
      if (nv == 1)
        launch_pdl(rms_norm_vec_kernel<T, 1>, dim3(rows), dim3(vt), 0, st, (const T*)x, (T*)residual, (const T*)weight, (T*)out,
                                                        dim, x_stride, res_stride, out_stride, eps, out_f16);
      else
        launch_pdl(rms_norm_vec_kernel<T, 2>, dim3(rows), dim3(vt), 0, st, (const T*)x, (T*)residual, (const T*)weight, (T*)out,
                                                        dim, x_stride, res_stride, out_stride, eps, out_f16);
    } else {
      using T = __half;
      if (nv == 1)
        launch_pdl(rms_norm_vec_kernel<T, 1>, dim3(rows), dim3(vt), 0, st, (const T*)x, (T*)residual, (const T*)weight, (T*)out,
                                                        dim, x_stride, res_stride, out_stride, eps, out_f16);
      else
        launch_pdl(rms_norm_vec_kernel<T, 2>, dim3(rows), dim3(vt), 0, st, (const T*)x, (T*)residual, (const T*)weight, (T*)out,
                                                        dim, x_stride, res_stride, out_stride, eps, out_f16);
    }
    return check_launch("rms_norm_vec_kernel");
  }
  B200_CHECK_ARG(!out_f16, "rms_norm: fp16 output needs the vectorised path (dim %% 8 == 0, 512 <= dim <= 16384, aligned rows)");
  return B200_DISPATCH_DTYPE(dtype, T, [&]() -> int {
    launch_pdl(rms_norm_kernel<T>, dim3(rows), dim3(threads), 0, st, (const T*)x, (T*)residual, (const T*)weight,
                                                 (T*)out, dim, x_stride, res_stride, out_stride,
                                                 eps);
    return check_launch("rms_norm_kernel");
  });
}

extern "C" int b200_rope(void* q, void* k, const int32_t* positions, int num_tokens,
                         int num_q_heads, int num_kv_heads, int head_dim, int rot_dim, float base,
                         int traditional, int64_t q_stride, int64_t k_stride, int dtype,
                         void* stream) {
  using namespace b200;
  if (num_tokens == 0) return B200_OK;
  B200_CHECK_ARG(num_tokens > 0 && q && positions, "rope: bad arguments");
  B200_CHECK_ARG(rot_dim > 0 && rot_dim % 2 == 0 && rot_dim <= head_dim,
                 "rope: rot_dim (%d) must be even and <= head_dim (%d)", rot_dim, head_dim);
  B200_CHECK_ARG(base > 0.f, "rope: base must be positive");
  cudaStream_t st = (cudaStream_t)stream;
  return B200_DISPATCH_DTYPE(dtype, T, [&]() -> int {
    launch_pdl(rope_kernel<T>, dim3(num_tokens), dim3(256), 0, st, (T*)q, (T*)k, positions, num_q_heads, num_kv_heads,
                                               head_dim, rot_dim, log2f(base), traditional,
                                               q_stride, k_stride);
    return check_launch("rope_kernel");
  });
}

extern "C" int b200_silu_mul(const void* gate_up, void* out, int rows, int inter,
                             int64_t in_stride, int64_t out_stride, int dtype, void* stream) {
  using namespace b200;
  if (rows == 0) return B200_OK;
  B200_CHECK_ARG(rows > 0 && inter > 0 && gate_up && out, "silu_mul: bad arguments");
  cudaStream_t st = (cudaStream_t)stream;
  dim3 grid((inter + 2047) / 2048, rows);  // 256 threads x 8 elements per CTA pass
  return B200_DISPATCH_DTYPE(dtype, T, [&]() -> int {
    launch_pdl(silu_mul_kernel<T>, dim3(grid), dim3(256), 0, st, (const T*)gate_up, (T*)out, inter, in_stride,
                                             out_stride);
    return check_launch("silu_mul_kernel");
  });
}

extern "C" int b200_argmax(const void* logits, int32_t* out, int rows, int vocab, int64_t stride,
                           int dtype, void* stream) {
  using namespace b200;
  if (rows == 0) return B200_OK;
  B200_CHECK_ARG(rows > 0 && vocab > 0 && logits && out, "argmax: bad arguments");
  cudaStream_t st = (cudaStream_t)stream;
  return B200_DISPATCH_DTYPE(dtype, T, [&]() -> int {
    launch_pdl(argmax_kernel<T>, dim3(kArgmaxSplit, rows), dim3(kArgmaxThreads), 0, st, (const T*)logits, out, vocab, stride);
    return check_launch("argmax_kernel");
  });
}

extern "C" int b200_add_inplace(void* y, const void* x, int rows, int dim, int64_t y_stride,
                                int64_t x_stride, int dtype, void* stream) {
  using namespace b200;
  if (rows == 0) return B200_OK;
  B200_CHECK_ARG(rows > 0 && dim > 0 && y && x, "add_inplace: bad arguments");
  cudaStream_t st = (cudaStream_t)stream;
  dim3 grid((dim + 1023) / 1024, rows);
  return B200_DISPATCH_DTYPE(dtype, T, [&]() -> int {
    launch_pdl(add_inplace_kernel<T>, dim3(grid), dim3(256), 0, st, (T*)y, (const T*)x, dim, y_stride, x_stride);
    return check_launch("add_inplace_kernel");
  });
}
