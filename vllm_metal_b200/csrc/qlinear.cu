// Affine group-quantized linear (MLX layout) for small and medium batch:
//   y[m,n] = sum_k x[m,k] * (scales[n,k/g] * q[n,k] + biases[n,k/g]) (+ bias[n])
// Replaces mx.quantized_matmul(..., transpose=True) as the reference calls it
// (vllm_metal/gguf/mlx_native.py:212-222; nn.QuantizedLinear from
// attention/impls/sdpa.py:297-343,739).
//
// Weight-streaming kernel: a CTA owns NT output rows and walks K in 64-element
// steps through a 4-stage cp.async ring holding the packed weight tile (NT x 32
// or 64 bytes) and the activation tile (MT x 128 bytes).  Each step the packed
// words are dequantised (s*q+b in fp32, one rounding to the activation type)
// into a bf16/fp16 tile, and 4 warps run m16n8k16 MMAs with the weight rows as
// the MMA M dimension and the batch rows as N, fp32 accumulation.  Algorithmic
// bytes: N*K*(bits/8 + 2*sb_bytes/g) + (M*K + M*N)*2.
#include "common.cuh"

namespace b200 {

// qgemm_tc.cu: tcgen05 / TMEM / TMA path
bool qgemm_tc_supported(const void* x, const uint32_t* w, int M, int N, int K, int group_size, int bits, int x_dtype,
                        int dtype, int sb_dtype, int64_t x_stride);
int qgemm_tc_launch(const void* x, const uint32_t* w, const void* scales, const void* biases,
                    const void* bias, void* y, int M, int N, int K, int group_size, int bits, int dtype,
                    int64_t x_stride, int64_t y_stride, bool prepacked, cudaStream_t st);
int qgemm_tc_set_variant(int v);
void qgemm_tc_set_trace(void* dev_ptr);

int qgemm_tc_set_tuning(int ksplit, int dequant_warps);

static int g_qlinear_kernel = 0;  // 0 auto, 1 mma.sync kernel, 2 tcgen05 kernel

namespace ql {

constexpr int KT = 64;      // k elements per pipeline step
constexpr int STAGES = 4;
constexpr int PITCH = KT + 8;  // smem row pitch in elements (144 B: conflict-free ldmatrix)

__device__ __forceinline__ void cp_async_16(uint32_t dst, const void* src, int src_bytes) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src),
               "r"(src_bytes)
               : "memory");
}
__device__ __forceinline__ void cp_async_commit() {
  asm volatile("cp.async.commit_group;" ::: "memory");
}
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void ldmatrix_x4(uint32_t& r0, uint32_t& r1, uint32_t& r2,
                                            uint32_t& r3, uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3)
               : "r"(addr));
}

template <typename T>
struct MmaOp;
template <>
struct MmaOp<__nv_bfloat16> {
  __device__ static __forceinline__ void run(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2,
                                             uint32_t a3, uint32_t b0, uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, "
        "{%8,%9}, {%0,%1,%2,%3};"
        : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
        : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
  }
  __device__ static __forceinline__ uint32_t pack(float lo, float hi) {
    __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&v);
  }
};
template <>
struct MmaOp<__half> {
  __device__ static __forceinline__ void run(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2,
                                             uint32_t a3, uint32_t b0, uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, "
        "{%8,%9}, {%0,%1,%2,%3};"
        : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
        : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
  }
  __device__ static __forceinline__ uint32_t pack(float lo, float hi) {
    __half2 v = __floats2half2_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&v);
  }
};

__device__ __forceinline__ float sb_load(const void* p, int64_t idx, int sb_dtype) {
  if (sb_dtype == B200_BF16) return __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(p)[idx]);
  if (sb_dtype == B200_F16) return __half2float(reinterpret_cast<const __half*>(p)[idx]);
  return reinterpret_cast<const float*>(p)[idx];
}

struct Params {
  const void* x;
  const uint32_t* w;
  const void* scales;
  const void* biases;
  const void* bias;
  void* y;
  int M, N, K, group_size, sb_dtype;
  int64_t x_stride, y_stride;
};

// T: activation type (bf16/fp16); BITS: 4|8; MT: batch rows per CTA (16|32|64);
// RB: 16-row blocks per warp (1|2) -> NT = 64*RB output rows per CTA.
template <typename T, int BITS, int MT, int RB>
__global__ void __launch_bounds__(128)
qgemm_mma_kernel(const Params p) {
  pdl_trigger();
  pdl_wait();
  constexpr int NT = 64 * RB;
  constexpr int U4_PER_ROW = KT * BITS / 128;   // 16-byte vectors of packed words per row-step
  constexpr int ELEMS_PER_U4 = 128 / BITS;      // 32 (int4) | 16 (int8)
  constexpr int W_U4 = NT * U4_PER_ROW;         // packed vectors per stage
  constexpr int W_PER_THREAD = W_U4 / 128;
  constexpr int X_CHUNKS = MT * (KT / 8);       // 16-byte activation chunks per stage
  constexpr int X_PER_THREAD = X_CHUNKS / 128;
  constexpr int STAGE_BYTES = W_U4 * 16 + MT * PITCH * 2;
  constexpr int MT8 = MT / 8;
  static_assert(W_PER_THREAD >= 1 && X_PER_THREAD >= 1, "tile too small for 128 threads");

  extern __shared__ __align__(128) uint8_t smem[];
  // [STAGES][packed W | x tile]  then the dequantised W tile [NT][PITCH]
  T* wtile = reinterpret_cast<T*>(smem + STAGES * STAGE_BYTES);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int n0 = blockIdx.x * NT;
  const int m0 = blockIdx.y * MT;
  const int nk = p.K / KT;
  const int words_per_row = p.K * BITS / 32;
  const uint32_t smem_base = (uint32_t)__cvta_generic_to_shared(smem);

  auto issue = [&](int kt, int stage) {
    if (kt < nk) {
      const uint32_t sb = smem_base + stage * STAGE_BYTES;
#pragma unroll
      for (int i = 0; i < W_PER_THREAD; ++i) {
        const int u = tid + i * 128;
        const int row = u / U4_PER_ROW, part = u % U4_PER_ROW;
        const int n = n0 + row;
        const bool ok = n < p.N;
        const uint32_t* src = p.w + (int64_t)(ok ? n : 0) * words_per_row +
                              (int64_t)kt * (KT * BITS / 32) + part * 4;
        cp_async_16(sb + u * 16, src, ok ? 16 : 0);
      }
#pragma unroll
      for (int i = 0; i < X_PER_THREAD; ++i) {
        const int c = tid + i * 128;
        const int row = c / (KT / 8), ch = c % (KT / 8);
        const int m = m0 + row;
        const bool ok = m < p.M;
        const T* src = reinterpret_cast<const T*>(p.x) + (int64_t)(ok ? m : 0) * p.x_stride +
                       (int64_t)kt * KT + ch * 8;
        cp_async_16(sb + W_U4 * 16 + row * (PITCH * 2) + ch * 16, src, ok ? 16 : 0);
      }
    }
    cp_async_commit();
  };

  float acc[RB][MT8][4];
#pragma unroll
  for (int r = 0; r < RB; ++r)
#pragma unroll
    for (int j = 0; j < MT8; ++j) acc[r][j][0] = acc[r][j][1] = acc[r][j][2] = acc[r][j][3] = 0.f;

#pragma unroll
  for (int s = 0; s < STAGES - 1; ++s) issue(s, s);

  const int groups_per_row = p.K / p.group_size;
  int stage = 0;
  for (int kt = 0; kt < nk; ++kt) {
    // scale / bias of this thread's packed vectors for this step (global, L2 resident)
    float sc[W_PER_THREAD], bs[W_PER_THREAD];
#pragma unroll
    for (int i = 0; i < W_PER_THREAD; ++i) {
      const int u = tid + i * 128;
      const int row = u / U4_PER_ROW, part = u % U4_PER_ROW;
      const int n = min(n0 + row, p.N - 1);
      const int k = kt * KT + part * ELEMS_PER_U4;
      const int64_t gi = (int64_t)n * groups_per_row + k / p.group_size;
      sc[i] = sb_load(p.scales, gi, p.sb_dtype);
      bs[i] = sb_load(p.biases, gi, p.sb_dtype);
    }
    cp_async_wait<STAGES - 2>();
    __syncthreads();  // stage kt landed for everyone; compute(kt-1) finished
    {
      int nst = stage + STAGES - 1;
      if (nst >= STAGES) nst -= STAGES;
      issue(kt + STAGES - 1, nst);
    }
    // ---- dequantise packed[stage] -> wtile ---------------------------------------
    const uint8_t* sp = smem + stage * STAGE_BYTES;
#pragma unroll
    for (int i = 0; i < W_PER_THREAD; ++i) {
      const int u = tid + i * 128;
      const int row = u / U4_PER_ROW, part = u % U4_PER_ROW;
      const uint4 pk = *reinterpret_cast<const uint4*>(sp + u * 16);
      const uint32_t words[4] = {pk.x, pk.y, pk.z, pk.w};
      T* dst = wtile + row * PITCH + part * ELEMS_PER_U4;
      if (BITS == 4) {
#pragma unroll
        for (int wi = 0; wi < 4; ++wi) {
          const uint32_t wd = words[wi];
          uint4 o;
          o.x = MmaOp<T>::pack(fmaf(sc[i], (float)(wd & 0xF), bs[i]),
                               fmaf(sc[i], (float)((wd >> 4) & 0xF), bs[i]));
          o.y = MmaOp<T>::pack(fmaf(sc[i], (float)((wd >> 8) & 0xF), bs[i]),
                               fmaf(sc[i], (float)((wd >> 12) & 0xF), bs[i]));
          o.z = MmaOp<T>::pack(fmaf(sc[i], (float)((wd >> 16) & 0xF), bs[i]),
                               fmaf(sc[i], (float)((wd >> 20) & 0xF), bs[i]));
          o.w = MmaOp<T>::pack(fmaf(sc[i], (float)((wd >> 24) & 0xF), bs[i]),
                               fmaf(sc[i], (float)(wd >> 28), bs[i]));
          *reinterpret_cast<uint4*>(dst + wi * 8) = o;
        }
      } else {
#pragma unroll
        for (int wi = 0; wi < 4; wi += 2) {
          const uint32_t a = words[wi], b = words[wi + 1];
          uint4 o;
          o.x = MmaOp<T>::pack(fmaf(sc[i], (float)(a & 0xFF), bs[i]),
                               fmaf(sc[i], (float)((a >> 8) & 0xFF), bs[i]));
          o.y = MmaOp<T>::pack(fmaf(sc[i], (float)((a >> 16) & 0xFF), bs[i]),
                               fmaf(sc[i], (float)(a >> 24), bs[i]));
          o.z = MmaOp<T>::pack(fmaf(sc[i], (float)(b & 0xFF), bs[i]),
                               fmaf(sc[i], (float)((b >> 8) & 0xFF), bs[i]));
          o.w = MmaOp<T>::pack(fmaf(sc[i], (float)((b >> 16) & 0xFF), bs[i]),
                               fmaf(sc[i], (float)(b >> 24), bs[i]));
          *reinterpret_cast<uint4*>(dst + (wi / 2) * 8) = o;
        }
      }
    }
    __syncthreads();
    // ---- MMA: warp rows [warp*16*RB, +16*RB) x MT batch rows ------------------------
    const uint32_t wt = smem_base + STAGES * STAGE_BYTES;
    const uint32_t xt = smem_base + stage * STAGE_BYTES + W_U4 * 16;
#pragma unroll
    for (int ks = 0; ks < KT / 16; ++ks) {
      uint32_t a[RB][4];
#pragma unroll
      for (int r = 0; r < RB; ++r) {
        const int row = (warp * RB + r) * 16 + (lane & 15);
        ldmatrix_x4(a[r][0], a[r][1], a[r][2], a[r][3],
                    wt + row * (PITCH * 2) + ks * 32 + (lane >> 4) * 16);
      }
#pragma unroll
      for (int j2 = 0; j2 < MT / 16; ++j2) {
        const int mj = lane >> 3, rr = lane & 7;
        const int xrow = j2 * 16 + (mj >> 1) * 8 + rr;
        uint32_t b0, b1, b2, b3;
        ldmatrix_x4(b0, b1, b2, b3, xt + xrow * (PITCH * 2) + ks * 32 + (mj & 1) * 16);
#pragma unroll
        for (int r = 0; r < RB; ++r) {
          MmaOp<T>::run(acc[r][2 * j2], a[r][0], a[r][1], a[r][2], a[r][3], b0, b1);
          MmaOp<T>::run(acc[r][2 * j2 + 1], a[r][0], a[r][1], a[r][2], a[r][3], b2, b3);
        }
      }
    }
    if (++stage == STAGES) stage = 0;
  }
  cp_async_wait<0>();

  // ---- epilogue: c0,c1 -> (n = g, m = 2t,2t+1); c2,c3 -> (n = g+8, ...) --------------
  const int g = lane >> 2, tq = lane & 3;
  T* yp = reinterpret_cast<T*>(p.y);
  const T* bp = reinterpret_cast<const T*>(p.bias);
#pragma unroll
  for (int r = 0; r < RB; ++r) {
#pragma unroll
    for (int half = 0; half < 2; ++half) {
      const int n = n0 + (warp * RB + r) * 16 + g + half * 8;
      if (n >= p.N) continue;
      const float bv = bp ? to_float(bp[n]) : 0.f;
#pragma unroll
      for (int j = 0; j < MT8; ++j) {
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          const int m = m0 + j * 8 + 2 * tq + e;
          if (m < p.M) yp[(int64_t)m * p.y_stride + n] = from_float<T>(acc[r][j][half * 2 + e] + bv);
        }
      }
    }
  }
}

template <typename T, int BITS, int MT, int RB>
static int launch(const Params& p, cudaStream_t st) {
  constexpr int NT = 64 * RB;
  constexpr int stage_bytes = NT * (KT * BITS / 128) * 16 + MT * PITCH * 2;
  constexpr int smem = STAGES * stage_bytes + NT * PITCH * 2;
  static bool configured_dev[16] = {false};  // cudaFuncSetAttribute is per device
  int dev_ = 0;
  B200_CUDA_OK(cudaGetDevice(&dev_));
  bool& configured = configured_dev[dev_ & 15];
  if (!configured) {
    B200_CUDA_OK(cudaFuncSetAttribute(qgemm_mma_kernel<T, BITS, MT, RB>,
                                      cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    configured = true;
  }
  dim3 grid((p.N + NT - 1) / NT, (p.M + MT - 1) / MT);
  launch_pdl(qgemm_mma_kernel<T, BITS, MT, RB>, dim3(grid), dim3(128), smem, st, p);
  return check_launch("qgemm_mma_kernel");
}

template <typename T, int BITS>
static int dispatch_tile(const Params& p, cudaStream_t st) {
  const int sms = sm_count() > 0 ? sm_count() : 148;
  // 128-row CTAs only when they still give every SM a few CTAs.
  const bool big = (p.N / 128) >= 3 * sms;
  if (p.M <= 16) return big ? launch<T, BITS, 16, 2>(p, st) : launch<T, BITS, 16, 1>(p, st);
  if (p.M <= 32) return big ? launch<T, BITS, 32, 2>(p, st) : launch<T, BITS, 32, 1>(p, st);
  return big ? launch<T, BITS, 64, 2>(p, st) : launch<T, BITS, 64, 1>(p, st);
}

// ---- row dequantisation (mx.dequantize / quantized embedding gather) ---------------
template <typename T, int BITS>
__global__ void __launch_bounds__(256)
dequantize_rows_kernel(const uint32_t* __restrict__ w, const void* __restrict__ scales,
                       const void* __restrict__ biases, const int32_t* __restrict__ ids,
                       T* __restrict__ out, int N, int K, int group_size, int sb_dtype) {
  pdl_trigger();
  pdl_wait();
  constexpr int PER_WORD = 32 / BITS;
  const int r = blockIdx.x;
  int64_t row = ids ? ids[r] : r;
  if (row < 0 || row >= N) row = 0;
  const int words = K / PER_WORD;
  const int groups = K / group_size;
  for (int wi = threadIdx.x; wi < words; wi += blockDim.x) {
    const uint32_t wd = w[row * words + wi];
    const int k0 = wi * PER_WORD;
    const int64_t gi = row * groups + k0 / group_size;  // a word never straddles a group
    const float s = sb_load(scales, gi, sb_dtype), b = sb_load(biases, gi, sb_dtype);
#pragma unroll
    for (int e = 0; e < PER_WORD; ++e) {
      const uint32_t q = (wd >> (e * BITS)) & ((1u << BITS) - 1u);
      out[(int64_t)r * K + k0 + e] = from_float<T>(fmaf(s, (float)q, b));
    }
  }
}

// ---- fp32 activations: MLX's own factorisation, y = sum_g (s_g * sum_k x_k q_k + b_g * sum_k x_k), all fp32.
// The reference exercises this only in tests (tests/test_gguf_mlx_native.py:70-86, atol = max|ref| * 4e-6);
// one warp per output element, lanes stride the row's packed words (coalesced), no tensor cores.
template <int BITS>
__global__ void __launch_bounds__(256)
qgemv_f32_kernel(const float* __restrict__ x, const uint32_t* __restrict__ w, const void* __restrict__ scales,
                 const void* __restrict__ biases, const float* __restrict__ bias, float* __restrict__ y, int M,
                 int N, int K, int group_size, int sb_dtype, int64_t x_stride, int64_t y_stride) {
  pdl_trigger();
  pdl_wait();
  constexpr int PER_WORD = 32 / BITS;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n = blockIdx.x * (blockDim.x >> 5) + warp;
  const int m = blockIdx.y;
  if (n >= N || m >= M) return;
  const int words = K / PER_WORD, groups = K / group_size;
  const float* xr = x + (int64_t)m * x_stride;
  const uint32_t* wr = w + (int64_t)n * words;
  float acc = 0.f;
  for (int wi = lane; wi < words; wi += 32) {
    const uint32_t wd = wr[wi];
    const int k0 = wi * PER_WORD;
    const int64_t gi = (int64_t)n * groups + k0 / group_size;
    const float s = sb_load(scales, gi, sb_dtype), b = sb_load(biases, gi, sb_dtype);
    float xq = 0.f, xs = 0.f;
#pragma unroll
    for (int e = 0; e < PER_WORD; ++e) {
      const float xv = xr[k0 + e];
      xq = fmaf(xv, (float)((wd >> (e * BITS)) & ((1u << BITS) - 1u)), xq);
      xs += xv;
    }
    acc += s * xq + b * xs;
  }
  acc = warp_sum(acc);
  if (lane == 0) y[(int64_t)m * y_stride + n] = acc + (bias ? bias[n] : 0.f);
}

}  // namespace ql
}  // namespace b200

static int check_quant_args(const char* what, int N, int K, int group_size, int bits,
                            int sb_dtype) {
  using namespace b200;
  B200_CHECK_ARG(bits == 4 || bits == 8, "%s: bits must be 4 or 8, got %d", what, bits);
  B200_CHECK_ARG(group_size == 32 || group_size == 64 || group_size == 128,
                 "%s: group_size must be 32, 64 or 128, got %d", what, group_size);
  B200_CHECK_ARG(N > 0 && K > 0 && K % group_size == 0,
                 "%s: K (%d) must be a positive multiple of group_size (%d); N=%d", what, K,
                 group_size, N);
  B200_CHECK_ARG(sb_dtype == B200_F32 || sb_dtype == B200_F16 || sb_dtype == B200_BF16,
                 "%s: unsupported scales/biases dtype code %d", what, sb_dtype);
  return B200_OK;
}

extern "C" int b200_qlinear(const void* x, const uint32_t* w, const void* scales,
                            const void* biases, const void* bias, void* y, int M, int N, int K,
                            int group_size, int bits, int x_dtype, int dtype, int sb_dtype, int64_t x_stride,
                            int64_t y_stride, void* stream) {
  using namespace b200;
  int rc = check_quant_args("qlinear", N, K, group_size, bits, sb_dtype);
  if (rc != B200_OK) return rc;
  B200_CHECK_ARG(M >= 0, "qlinear: M must be >= 0, got %d", M);
  if (M == 0) return B200_OK;
  B200_CHECK_ARG(x && w && scales && biases && y, "qlinear: null pointer argument");
  B200_CHECK_ARG(x_dtype == dtype || (x_dtype == B200_F16 && dtype == B200_BF16),
                 "qlinear: x dtype %d with result dtype %d (x must match, or be fp16 for bf16 results)", x_dtype, dtype);
  if (dtype == B200_F32) {
    B200_CHECK_ARG(y_stride >= N && x_stride >= K, "qlinear: bad fp32 strides");
    cudaStream_t stf = (cudaStream_t)stream;
    const dim3 grid((N + 7) / 8, M), block(256);
    const float* xf = reinterpret_cast<const float*>(x);
    const float* bf = reinterpret_cast<const float*>(bias);
    float* yf = reinterpret_cast<float*>(y);
    if (bits == 4)
      launch_pdl(ql::qgemv_f32_kernel<4>, grid, block, 0, stf, xf, w, scales, biases, bf, yf, M, N, K, group_size,
                 sb_dtype, x_stride, y_stride);
    else
      launch_pdl(ql::qgemv_f32_kernel<8>, grid, block, 0, stf, xf, w, scales, biases, bf, yf, M, N, K, group_size,
                 sb_dtype, x_stride, y_stride);
    return check_launch("qgemv_f32_kernel");
  }
  B200_CHECK_ARG(dtype == B200_BF16 || dtype == B200_F16,
                 "qlinear: activations must be fp32, bf16 or fp16 (dtype code %d)", dtype);
  B200_CHECK_ARG(K % ql::KT == 0, "qlinear: K (%d) must be a multiple of %d", K, ql::KT);
  cudaStream_t st0 = (cudaStream_t)stream;
  if (g_qlinear_kernel != 1) {
    const bool ok = y_stride >= N &&
                    qgemm_tc_supported(x, w, M, N, K, group_size, bits, x_dtype, dtype, sb_dtype, x_stride);
    if (ok) return qgemm_tc_launch(x, w, scales, biases, bias, y, M, N, K, group_size, bits, dtype, x_stride,
                                   y_stride, false, st0);
    if (g_qlinear_kernel == 2 || x_dtype != dtype)
      return set_error(B200_ERR_UNSUPPORTED,
                       "qlinear: tcgen05 kernel does not take x_dtype=%d dtype=%d sb_dtype=%d K=%d x_stride=%lld "
                       "(it reads fp16 x -- b200_cast_bf16_to_f16 -- and writes bf16)",
                       x_dtype, dtype, sb_dtype, K, (long long)x_stride);
  }
  B200_CHECK_ARG(x_stride >= K && x_stride % 8 == 0 && ((uintptr_t)x % 16) == 0,
                 "qlinear: x rows must be 16-byte aligned (stride %lld)", (long long)x_stride);
  B200_CHECK_ARG(((uintptr_t)w % 16) == 0, "qlinear: packed weights must be 16-byte aligned");
  B200_CHECK_ARG(y_stride >= N, "qlinear: y_stride (%lld) < N (%d)", (long long)y_stride, N);
  ql::Params p{x, w, scales, biases, bias, y, M, N, K, group_size, sb_dtype, x_stride, y_stride};
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == B200_BF16)
    return bits == 4 ? ql::dispatch_tile<__nv_bfloat16, 4>(p, st)
                     : ql::dispatch_tile<__nv_bfloat16, 8>(p, st);
  return bits == 4 ? ql::dispatch_tile<__half, 4>(p, st) : ql::dispatch_tile<__half, 8>(p, st);
}

// Streaming ("prepacked") weight layout: same shape as the MLX layout, nibbles permuted inside
// every 32-bit word so that nibbles (i, i+4) hold elements (2i, 2i+1) of the word's 8 elements.
// The tcgen05 kernel's (x >> 4i) & 0x000F000F extraction then yields adjacent-k pairs directly.
namespace b200 {
namespace ql {
__global__ void prepack_int4_kernel(const uint32_t* __restrict__ in, uint32_t* __restrict__ out, int64_t n,
                                    int inverse) {
  // no early pdl_trigger(): the quantized GEMM prefetches weights BEFORE its own pdl_wait (they are
  // constants on the decode path), so a dependent launch must not start until this grid is done
  pdl_wait();
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const uint32_t w = in[i];
    uint32_t o = 0;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      if (!inverse) {
        o |= ((w >> (8 * j)) & 0xFu) << (4 * j);             // element 2j   -> nibble j
        o |= ((w >> (8 * j + 4)) & 0xFu) << (4 * j + 16);    // element 2j+1 -> nibble j+4
      } else {
        o |= ((w >> (4 * j)) & 0xFu) << (8 * j);
        o |= ((w >> (4 * j + 16)) & 0xFu) << (8 * j + 4);
      }
    }
    out[i] = o;
  }
}
}  // namespace ql
}  // namespace b200

extern "C" int b200_qlinear_prepack(const uint32_t* w, uint32_t* out, int N, int K, int bits, int inverse,
                                    void* stream) {
  using namespace b200;
  B200_CHECK_ARG(w && out, "qlinear_prepack: null pointer argument");
  B200_CHECK_ARG(bits == 4 || bits == 8, "qlinear_prepack: bits must be 4 or 8, got %d", bits);
  B200_CHECK_ARG(N > 0 && K > 0 && (K * bits) % 32 == 0, "qlinear_prepack: bad shape N=%d K=%d", N, K);
  const int64_t words = (int64_t)N * ((int64_t)K * bits / 32);
  cudaStream_t st = (cudaStream_t)stream;
  if (bits == 8) {  // int8 streams in the MLX order already
    if (w != out) {
      B200_CUDA_OK(cudaMemcpyAsync(out, w, words * 4, cudaMemcpyDeviceToDevice, st));
      B200_CUDA_OK(cudaStreamSynchronize(st));
    }
    return B200_OK;
  }
  const int threads = 256;
  const int64_t want = (words + threads - 1) / threads;
  const int blocks = (int)(want < 148 * 16 ? want : 148 * 16);
  launch_pdl(ql::prepack_int4_kernel, dim3(blocks), dim3(threads), 0, st, w, out, words, inverse);
  B200_CUDA_OK(cudaGetLastError());
  // load-time call: make the streaming layout stream-complete before any qlinear can prefetch it
  cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
  B200_CUDA_OK(cudaStreamIsCapturing(st, &cap));
  if (cap == cudaStreamCaptureStatusNone) B200_CUDA_OK(cudaStreamSynchronize(st));
  return B200_OK;
}

extern "C" int b200_qlinear_prepacked(const void* x, const uint32_t* wp, const void* scales,
                                      const void* biases, const void* bias, void* y, int M, int N, int K,
                                      int group_size, int bits, int x_dtype, int dtype, int sb_dtype, int64_t x_stride,
                                      int64_t y_stride, void* stream) {
  using namespace b200;
  int rc = check_quant_args("qlinear_prepacked", N, K, group_size, bits, sb_dtype);
  if (rc != B200_OK) return rc;
  B200_CHECK_ARG(M >= 0, "qlinear_prepacked: M must be >= 0, got %d", M);
  if (M == 0) return B200_OK;
  B200_CHECK_ARG(x && wp && scales && biases && y, "qlinear_prepacked: null pointer argument");
  B200_CHECK_ARG(y_stride >= N, "qlinear_prepacked: y_stride (%lld) < N (%d)", (long long)y_stride, N);
  if (!qgemm_tc_supported(x, wp, M, N, K, group_size, bits, x_dtype, dtype, sb_dtype, x_stride))
    return set_error(B200_ERR_UNSUPPORTED,
                     "qlinear_prepacked: the streaming layout is only read by the tcgen05 kernel, which does not "
                     "take x_dtype=%d (fp16 needed) dtype=%d sb_dtype=%d K=%d x_stride=%lld",
                     x_dtype, dtype, sb_dtype, K, (long long)x_stride);
  return qgemm_tc_launch(x, wp, scales, biases, bias, y, M, N, K, group_size, bits, dtype, x_stride, y_stride,
                         true, (cudaStream_t)stream);
}

extern "C" int b200_set_qlinear_trace(void* dev_ptr) {
  b200::qgemm_tc_set_trace(dev_ptr);
  return B200_OK;
}

extern "C" int b200_set_qlinear_variant(int variant) { return b200::qgemm_tc_set_variant(variant); }

extern "C" int b200_set_qlinear_kernel(int kernel) {
  if (kernel < 0 || kernel > 2)
    return b200::set_error(B200_ERR_INVALID_ARGUMENT, "set_qlinear_kernel: expected 0, 1 or 2, got %d", kernel);
  b200::g_qlinear_kernel = kernel;
  return B200_OK;
}

extern "C" int b200_set_qlinear_tuning(int ksplit, int dequant_warps) {
  return b200::qgemm_tc_set_tuning(ksplit, dequant_warps);
}

extern "C" int b200_dequantize_rows(const uint32_t* w, const void* scales, const void* biases,
                                    const int32_t* ids, void* out, int num_rows, int N, int K,
                                    int group_size, int bits, int sb_dtype, int out_dtype,
                                    void* stream) {
  using namespace b200;
  int rc = check_quant_args("dequantize_rows", N, K, group_size, bits, sb_dtype);
  if (rc != B200_OK) return rc;
  if (num_rows == 0) return B200_OK;
  B200_CHECK_ARG(num_rows > 0 && w && scales && biases && out, "dequantize_rows: bad arguments");
  cudaStream_t st = (cudaStream_t)stream;
  return B200_DISPATCH_DTYPE(out_dtype, T, [&]() -> int {
    if (bits == 4)
      launch_pdl(ql::dequantize_rows_kernel<T, 4>, dim3(num_rows), dim3(256), 0, st, w, scales, biases, ids, (T*)out, N,
                                                                K, group_size, sb_dtype);
    else
      launch_pdl(ql::dequantize_rows_kernel<T, 8>, dim3(num_rows), dim3(256), 0, st, w, scales, biases, ids, (T*)out, N,
                                                                K, group_size, sb_dtype);
    return check_launch("dequantize_rows_kernel");
  });
}


// Stacked-expert form (mx.gather_qmm over rows already grouped by expert; mlx-lm QuantizedSwitchLinear, named in
// vllm_metal/quant/awq_loader.py:171-173): expert e multiplies the `counts[e]` consecutive rows of x that follow the
// rows of experts 0..e-1.  One b200_qlinear launch per expert that has rows, issued from this one call -- the host
// cost per expert is a C loop iteration, not a trip through the Python shim.
extern "C" int b200_qlinear_experts(const void* x, const uint32_t* w, const void* scales, const void* biases,
                                    const void* bias, void* y, const int32_t* counts, int num_experts, int N, int K,
                                    int group_size, int bits, int x_dtype, int dtype, int sb_dtype, int64_t x_stride,
                                    int64_t y_stride, void* stream) {
  using namespace b200;
  B200_CHECK_ARG(counts && num_experts > 0, "qlinear_experts: bad arguments");
  int rc = check_quant_args("qlinear_experts", N, K, group_size, bits, sb_dtype);
  if (rc != B200_OK) return rc;
  const size_t w_words = (size_t)N * ((size_t)K * bits / 32);
  const size_t sb_bytes = (size_t)N * (K / group_size) * dtype_size(sb_dtype);
  const size_t xe = dtype_size(x_dtype), ye = dtype_size(dtype);
  int64_t at = 0;
  for (int e = 0; e < num_experts; ++e) {
    const int c = counts[e];
    B200_CHECK_ARG(c >= 0, "qlinear_experts: negative row count for expert %d", e);
    if (c == 0) continue;
    rc = b200_qlinear(reinterpret_cast<const char*>(x) + (size_t)at * x_stride * xe, w + e * w_words,
                      reinterpret_cast<const char*>(scales) + e * sb_bytes,
                      reinterpret_cast<const char*>(biases) + e * sb_bytes,
                      bias ? reinterpret_cast<const char*>(bias) + (size_t)e * N * ye : nullptr,
                      reinterpret_cast<char*>(y) + (size_t)at * y_stride * ye, c, N, K, group_size, bits, x_dtype, dtype,
                      sb_dtype, x_stride, y_stride, stream);
    if (rc != B200_OK) return rc;
    at += c;
  }
  return B200_OK;
}
