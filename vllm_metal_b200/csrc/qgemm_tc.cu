// Affine group-quantized GEMM on the 5th-gen tensor cores (tcgen05 + TMEM + TMA).
//
//   y[m,n] = sum_k x[m,k] * (scales[n,k/g] * q[n,k] + biases[n,k/g]) (+ bias[n])
//
// Replaces mx.quantized_matmul(x, w, scales, biases, transpose=True, ...) as the reference
// calls it (vllm_metal/gguf/mlx_native.py:212-222; nn.QuantizedLinear from
// attention/impls/sdpa.py:297-343,739) for every batch size: decode (M <= 64) is
// weight-streaming / HBM-bound, prefill (M >= 256) is tensor-bound.
//
// One CTA owns 128 output features (UMMA M = 128 = weight rows) and a batch tile of NB rows
// of x (UMMA N = NB in {16..256}); the fp32 accumulator [128 x NB] lives in TMEM.
//
//   warps 0..DQW-1     dequant: DQW/4 threads per weight row, each unpacks its share of the 64-wide
//                      k block (int4/int8 -> s*q+b, bf16x2 HFMA2, single rounding) -> A ring in the
//                      canonical K-major SWIZZLE_128B layout
//   warp DQW   (1 thr)  TMA: packed uint32 weight tiles [128 rows x 128 B] -> smem ring (PS deep)
//   warp DQW+1 (1 thr)  TMA: x tiles [NB x 64] bf16, 128B-swizzled K-major  -> B ring (NS deep)
//   warp DQW+2 (1 thr)  tcgen05.mma.cta_group::1.kind::f16 128xNBx16, 4 per 64-wide k block;
//                      tcgen05.commit frees the A/B stage and finally signals the epilogue
//   warps 0..DQW-1     epilogue: tcgen05.ld -> (+bias) -> y (warp w reads TMEM lanes 32*(w%4)..)
//
// Matrices with few row tiles (o_proj, down_proj: N = hidden) are split along K over a thread
// block CLUSTER (1 x ksplit x 1); partial accumulators are exchanged through distributed
// shared memory and reduced in-cluster, so the op stays a single launch with no workspace.
//
// Algorithmic bytes: N*K*(bits/8 + 2*sizeof(sb)/g) + (M*K + M*N)*2.
#include "tcgen05.cuh"

namespace b200 {
namespace qtc {
using namespace tc;

constexpr int ROWS = 128;  // weight rows per CTA (UMMA M)
constexpr int KB = 64;     // k elements per A/B stage (128 bytes of 16-bit)
constexpr int A_STAGE_BYTES = ROWS * 128;
constexpr int P_STAGE_BYTES = ROWS * 128;

// ------------------------------------------------------------- dequant traits ----
template <typename T>
struct Cvt;
template <>
struct Cvt<__nv_bfloat16> {
  using T2 = __nv_bfloat162;
  static constexpr int kFmt = 1;
  static constexpr uint32_t kMagic = 0x43004300u;  // bf16 128.0: 128 + q exact for q < 128
  __device__ static __forceinline__ T2 magic_off() { return __float2bfloat162_rn(-128.f); }
  __device__ static __forceinline__ T2 bcast(uint16_t raw) {
    __nv_bfloat16_raw r;
    r.x = raw;
    return __bfloat162bfloat162(__nv_bfloat16(r));
  }
  __device__ static __forceinline__ uint32_t f2(float lo, float hi) {
    T2 v = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&v);
  }
  __device__ static __forceinline__ float to_f(uint16_t raw) {
    return __uint_as_float((uint32_t)raw << 16);
  }
};
template <>
struct Cvt<__half> {
  using T2 = __half2;
  static constexpr int kFmt = 0;
  static constexpr uint32_t kMagic = 0x64006400u;  // fp16 1024.0: 1024 + q exact
  __device__ static __forceinline__ T2 magic_off() { return __float2half2_rn(-1024.f); }
  __device__ static __forceinline__ T2 bcast(uint16_t raw) {
    __half_raw r;
    r.x = raw;
    return __half2half2(__half(r));
  }
  __device__ static __forceinline__ uint32_t f2(float lo, float hi) {
    T2 v = __floats2half2_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&v);
  }
  __device__ static __forceinline__ float to_f(uint16_t raw) {
    __half_raw r;
    r.x = raw;
    return __half2float(__half(r));
  }
};

// 8 nibbles of one packed word -> 8 values s*q+b in memory order (k0..k7), as 4 packed pairs.
// PRE: the word was nibble-permuted at load time (b200_qlinear_prepack) so that nibbles
// (i, i+4) hold (k_{2i}, k_{2i+1}); the extraction then yields the pairs in memory order
// and the four PRMTs disappear.
template <typename T, bool PRE>
__device__ __forceinline__ uint4 dequant_word4(uint32_t w, typename Cvt<T>::T2 s2,
                                               typename Cvt<T>::T2 b2) {
  using C = Cvt<T>;
  using T2 = typename C::T2;
  uint32_t a[4];
  const uint32_t mask = 0x000F000Fu, magic = C::kMagic;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    // (nibble i, nibble i+4) as the low/high halves: one LOP3 ((x & mask) | magic) on the
    // shifted word (both constants live in registers so ptxas keeps it a single instruction)
    uint32_t v;
    asm("lop3.b32 %0, %1, %2, %3, 0xEA;" : "=r"(v) : "r"(w >> (4 * i)), "r"(mask), "r"(magic));
    T2 q = __hadd2(*reinterpret_cast<T2*>(&v), C::magic_off());  // exact: q in [0, 15]
    T2 r = __hfma2(s2, q, b2);                                    // s*q+b, one rounding
    a[i] = *reinterpret_cast<uint32_t*>(&r);
  }
  uint4 o;
  if constexpr (PRE) {
    o.x = a[0]; o.y = a[1]; o.z = a[2]; o.w = a[3];
  } else {
    o.x = __byte_perm(a[0], a[1], 0x5410);  // (k0, k1)
    o.y = __byte_perm(a[2], a[3], 0x5410);  // (k2, k3)
    o.z = __byte_perm(a[0], a[1], 0x7632);  // (k4, k5)
    o.w = __byte_perm(a[2], a[3], 0x7632);  // (k6, k7)
  }
  return o;
}
// 4 bytes of one packed word -> 4 fp16 values s*q+b: a byte under the exponent byte 0x64 reads 1024 + q exactly
// (q < 1024), the subtraction is exact and the FMA rounds once -- the same single rounding as an fp32 fma followed
// by a conversion, in 2 PRMT + 2 HADD2 + 2 HFMA2 instead of 4 I2F + 4 FFMA + 2 F2F per word (the int8 GEMM at
// M >= 256 was dequant-bound on the conversions: 0.28-0.47 of the bf16 peak, profiles/r02_gemm_bigm.txt).
template <typename T>
__device__ __forceinline__ uint2 dequant_word8(uint32_t w, float s, float b) {
  const __half2 s2 = __float2half2_rn(s), b2 = __float2half2_rn(b), off = __float2half2_rn(-1024.f);
  const uint32_t lo = __byte_perm(w, 0x64646464u, 0x5140), hi = __byte_perm(w, 0x64646464u, 0x7362);
  const __half2 q0 = __hadd2(*reinterpret_cast<const __half2*>(&lo), off);
  const __half2 q1 = __hadd2(*reinterpret_cast<const __half2*>(&hi), off);
  const __half2 r0 = __hfma2(s2, q0, b2), r1 = __hfma2(s2, q1, b2);
  uint2 o;
  o.x = *reinterpret_cast<const uint32_t*>(&r0);
  o.y = *reinterpret_cast<const uint32_t*>(&r1);
  return o;
}

struct Params {
  const uint16_t* scales;
  const uint16_t* biases;
  const void* bias;  // dense [N] of T or null
  void* y;
  int M, N, K, group_size;
  int64_t y_stride;
  int num_kb;        // K / 64
  int kb_per_split;  // k blocks per cluster rank
  unsigned long long* trace;  // debug: 8 globaltimer stamps per CTA (b200_set_qlinear_trace), else null
};

static unsigned long long* g_trace = nullptr;  // host side: b200_set_qlinear_trace

// In-loop cycle counters perturb the dequant chain: compiled in only with -DQGEMM_LOOP_TRACE.
#ifdef QGEMM_LOOP_TRACE
__device__ __forceinline__ long long lt_clock() { return clock64(); }
#else
__device__ __forceinline__ long long lt_clock() { return 0; }
#endif

__device__ __forceinline__ void trace_put(const Params& p, int slot, unsigned long long v) {
  if (p.trace != nullptr)
    p.trace[((size_t)(blockIdx.x * gridDim.y + blockIdx.y) * gridDim.z + blockIdx.z) * 64 + slot] = v;
}

__device__ __forceinline__ void trace_stamp(const Params& p, int slot) {
  if (p.trace != nullptr) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    p.trace[((size_t)(blockIdx.x * gridDim.y + blockIdx.y) * gridDim.z + blockIdx.z) * 64 + slot] = t;
  }
}

constexpr int SB_STAGE_BYTES = 2 * ROWS * 16;  // 8 groups of scales + 8 of biases per row
constexpr int SBS = 2;                         // scale/bias ring depth

// The dynamic shared window is declared __align__(1024) (SWIZZLE_128B stage bases); no slack.
template <int NB, int PS, int NS, bool SBT, bool TA>
constexpr int smem_bytes() {
  return PS * P_STAGE_BYTES + NS * ((TA ? 0 : A_STAGE_BYTES) + NB * 128) + (SBT ? SBS * SB_STAGE_BYTES : 0) + 256;
}

__device__ __forceinline__ float4 ld_dsmem_v4(uint32_t local_addr, uint32_t peer) {
  uint32_t remote;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(local_addr), "r"(peer));
  float4 v;
  asm volatile("ld.shared::cluster.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "r"(remote)
               : "memory");
  return v;
}

// T: activation / scale type; BITS: 4|8; G: quantisation group; NB: batch tile (UMMA N);
// PS/NS: ring depths; SBT: scales/biases staged by TMA (8 groups per stage) instead of per-thread
// global loads; PRE: packed words are in the prepacked nibble order (see dequant_word4).
// TA: the dequantised A tile goes to TENSOR MEMORY (tcgen05.st, one lane per weight row) and the MMA
// reads it from there -- no shared-memory A ring, no swizzle, no generic->async proxy fence.
template <typename T, int BITS, int G, int NB, int PS, int NS, int DQW, bool SBT, bool PRE, bool TA>
__global__ void __launch_bounds__((DQW + 3) * 32, NB <= 64 ? 2 : 1)
qgemm_tc_kernel(const __grid_constant__ CUtensorMap w_map, const __grid_constant__ CUtensorMap x_map,
                const __grid_constant__ CUtensorMap s_map, const __grid_constant__ CUtensorMap b_map,
                const Params p) {
  // T is the type of scales / biases / bias / y; the tensor-core OPERANDS are always fp16 (x arrives as fp16, the
  // weights are dequantised to fp16): tcgen05 kind::f16 faults on mixed fp16/bf16 operands, and bf16's 8
  // significant bits on s*q+b break the 2e-2/1e-2 bound at K >= 25600 (tests/test_gpu_layer_shapes.py).
  using C = Cvt<__half>;
  using CI = Cvt<T>;
  using T2 = typename C::T2;
  constexpr int KPP = 1024 / BITS;        // k per packed stage (128 B per row): 256 | 128
  constexpr int KB_PER_P = KPP / KB;      // 4 | 2
  constexpr int CH_PER_KB = KB * BITS / 128;  // 16-byte packed chunks per k block: 2 | 4
  constexpr int B_STAGE_BYTES = NB * 128;
  constexpr int ACC_COLS = NB < 32 ? 32 : NB;
  constexpr int A_COLS = KB / 2;  // 32-bit TMEM columns of one 64-wide k block of 16-bit values
  constexpr int TMEM_NEED = ACC_COLS + (TA ? NS * A_COLS : 0);
  constexpr int TMEM_COLS = TMEM_NEED <= 32 ? 32 : TMEM_NEED <= 64 ? 64 : TMEM_NEED <= 128 ? 128
                            : TMEM_NEED <= 256 ? 256 : 512;
  static_assert(TMEM_NEED <= 512, "accumulator + A ring exceed tensor memory");
  static_assert(!TA || (DQW == 8 && (NB > 64 || TMEM_COLS <= 256)), "TA: 8 dequant warps, two CTAs must fit TMEM");
  constexpr int GPP = KPP / G;  // groups per packed stage (>= 1)
  constexpr int TPR = DQW / 4;                // dequant threads per weight row: 1 | 2
  constexpr int CH_T = CH_PER_KB / TPR;       // packed chunks per thread per k block
  constexpr int W_TMA = DQW, W_X = DQW + 1, W_MMA = DQW + 2;
  constexpr int SB_PER = SBT ? (8 * G) / KPP : 1;  // packed stages per scale/bias stage
  static_assert(DQW == 4 || DQW == 8, "4 or 8 dequant warps");
  static_assert(CH_PER_KB % TPR == 0, "chunks must split evenly over the row's threads");
  static_assert(!SBT || (8 * G) % KPP == 0, "a scale/bias stage must cover whole packed stages");

  // PDL: the next kernel may start its prologue now; OUR weight/scale TMA traffic (constant data)
  // starts before the previous kernel has finished, only the x producer and the epilogue wait.
  pdl_trigger();
  if (threadIdx.x == 0) trace_stamp(p, 0);  // CTA start
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t base = smem_u32(smem_raw);
  if ((base & 1023u) != 0) __trap();  // SWIZZLE_128B stages need 1 KiB alignment: fail loudly
  const uint32_t sm_p = base;
  const uint32_t sm_a = sm_p + PS * P_STAGE_BYTES;
  const uint32_t sm_b = sm_a + (TA ? 0 : NS * A_STAGE_BYTES);
  const uint32_t sm_sb = sm_b + NS * B_STAGE_BYTES;
  const uint32_t sm_bar = sm_sb + (SBT ? SBS * SB_STAGE_BYTES : 0);
  // barriers (8 B each): full_p[PS], empty_p[PS], full_ab[NS], empty_ab[NS], tmem_full,
  // full_sb[SBS], empty_sb[SBS]; then the tmem pointer
  const uint32_t bar_full_p = sm_bar, bar_empty_p = sm_bar + 8 * PS;
  const uint32_t bar_full_ab = sm_bar + 16 * PS, bar_empty_ab = bar_full_ab + 8 * NS;
  const uint32_t bar_tmem_full = bar_empty_ab + 8 * NS;
  const uint32_t bar_full_sb = bar_tmem_full + 8, bar_empty_sb = bar_full_sb + 8 * SBS;
  const uint32_t tmem_ptr_addr = bar_empty_sb + 8 * SBS;
  static_assert(8 * (2 * PS + 2 * NS + 1 + 2 * SBS) + 4 <= 256, "barrier block overflows its 256 bytes");
  uint8_t* gen_base = smem_raw;
  volatile uint32_t* tmem_ptr_gen = reinterpret_cast<volatile uint32_t*>(gen_base + (tmem_ptr_addr - base));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n0 = blockIdx.x * ROWS;
  const int m0 = blockIdx.z * NB;
  const int ksplit = gridDim.y;
  const int krank = ksplit > 1 ? (int)cluster_ctarank() : 0;
  const int kb_begin = krank * p.kb_per_split;
  int my_kb = p.num_kb - kb_begin;
  if (my_kb > p.kb_per_split) my_kb = p.kb_per_split;
  if (my_kb < 0) my_kb = 0;
  const int my_ps = (my_kb + KB_PER_P - 1) / KB_PER_P;
  const int gps0 = kb_begin / KB_PER_P;  // global index of this rank's first packed stage

  if (threadIdx.x == 0) {
    for (int i = 0; i < PS; ++i) {
      mbar_init(bar_full_p + 8 * i, 1);
      mbar_init(bar_empty_p + 8 * i, DQW);
    }
    for (int i = 0; i < NS; ++i) {
      mbar_init(bar_full_ab + 8 * i, DQW + 1);  // dequant warps + the x producer's expect_tx arrive
      mbar_init(bar_empty_ab + 8 * i, 1);
    }
    mbar_init(bar_tmem_full, 1);
    for (int i = 0; i < SBS; ++i) {
      mbar_init(bar_full_sb + 8 * i, 1);
      mbar_init(bar_empty_sb + 8 * i, DQW);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == W_MMA) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_ptr_addr),
                 "n"(TMEM_COLS)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_gen;
  if (threadIdx.x == 0) trace_stamp(p, 1);  // setup done (barriers, TMEM)

  if (warp == W_TMA) {
    // ===== packed-weight TMA producer (whole-warp loop, issue under elect_one(): see tcgen05.cuh) =====
    for (int ps = 0; ps < my_ps; ++ps) {
      const int st = ps % PS;
      if constexpr (SBT) {
        // scale/bias stages sit on the GLOBAL grid of 8 groups (TMA needs the inner start
        // coordinate 16-byte aligned); a cluster rank may start in the middle of one
        const int gps = gps0 + ps;
        if (ps == 0 || gps % SB_PER == 0) {
          const int sbi = gps / SB_PER - gps0 / SB_PER, sst = sbi % SBS;
          mbar_wait(bar_empty_sb + 8 * sst, ((sbi / SBS) & 1) ^ 1);
          const int g0 = (gps / SB_PER) * 8;
          if (elect_one()) {
            mbar_expect_tx(bar_full_sb + 8 * sst, SB_STAGE_BYTES);
            tma_load_2d(sm_sb + sst * SB_STAGE_BYTES, &s_map, bar_full_sb + 8 * sst, g0, n0);
            tma_load_2d(sm_sb + sst * SB_STAGE_BYTES + ROWS * 16, &b_map, bar_full_sb + 8 * sst, g0, n0);
          }
          __syncwarp();
        }
      }
      mbar_wait(bar_empty_p + 8 * st, ((ps / PS) & 1) ^ 1);
#ifdef QGEMM_LOOP_TRACE
      if (lane == 0 && ps < 16) trace_stamp(p, 32 + ps);  // issue time of packed stage ps
#endif
      if (elect_one()) {
        mbar_expect_tx(bar_full_p + 8 * st, P_STAGE_BYTES);
        tma_load_2d(sm_p + st * P_STAGE_BYTES, &w_map, bar_full_p + 8 * st,
                    (kb_begin + ps * KB_PER_P) * (KB * BITS / 32), n0);
      }
      __syncwarp();
    }
  } else if (warp == W_X) {
    // ===== x tile TMA producer =====
    pdl_wait();  // x is the previous kernel's output
    if (lane == 0) trace_stamp(p, 2);  // previous kernel complete
    for (int kb = 0; kb < my_kb; ++kb) {
      const int st = kb % NS;
      mbar_wait(bar_empty_ab + 8 * st, ((kb / NS) & 1) ^ 1);
      if (elect_one()) {
        mbar_expect_tx(bar_full_ab + 8 * st, B_STAGE_BYTES);
        tma_load_2d(sm_b + st * B_STAGE_BYTES, &x_map, bar_full_ab + 8 * st, (kb_begin + kb) * KB, m0);
      }
      __syncwarp();
    }
  } else if (warp == W_MMA) {
    // ===== MMA issuer: whole-warp loop with warp-uniform operands, issue under elect_one().  Under `if (lane == 0)`
    // ptxas wrapped every tcgen05.mma in an ELECT + R2UR waterfall (~80 cycles per MMA): at NB = 64 (32 cycles of
    // tensor pipe per MMA) that WAS the "0.3 us per k block whatever the math" of round 1's ablation. =====
    constexpr uint32_t idesc = make_idesc(C::kFmt, ROWS, NB);
    const uint32_t tb = __shfl_sync(0xffffffffu, tmem_base, 0);
    for (int kb = 0; kb < my_kb; ++kb) {
      const int st = kb % NS;
      mbar_wait(bar_full_ab + 8 * st, (kb / NS) & 1);
      tc_fence_after();
      if (elect_one()) {
        const uint64_t db = make_smem_desc(sm_b + st * B_STAGE_BYTES);
        if constexpr (TA) {
          const uint32_t ta = tb + ACC_COLS + st * A_COLS;
#pragma unroll
          for (int k = 0; k < KB / 16; ++k)  // 16 elements = 8 TMEM columns of A, 32 B of the B row
            tc_mma_f16_ts(tb, ta + 8 * k, db + 2 * k, idesc, (kb | k) != 0);
        } else {
          const uint64_t da = make_smem_desc(sm_a + st * A_STAGE_BYTES);
#pragma unroll
          for (int k = 0; k < KB / 16; ++k)  // +32 B (16 elements) inside the 128 B swizzle row
            tc_mma_f16(tb, da + 2 * k, db + 2 * k, idesc, (kb | k) != 0);
        }
        tc_commit(bar_empty_ab + 8 * st);
      }
      __syncwarp();
    }
    if (elect_one()) tc_commit(bar_tmem_full);
    __syncwarp();
  } else {
    // ===== dequant warps: row n0 + r is shared by TPR threads (hsel picks the chunk share) =====
    const int r = threadIdx.x & (ROWS - 1);
    const int hsel = threadIdx.x >> 7;  // 0 .. TPR-1
    const int n = min(n0 + r, p.N - 1);
    const int groups_per_row = p.K / G;
    const uint16_t* srow = p.scales + (int64_t)n * groups_per_row;
    const uint16_t* brow = p.biases + (int64_t)n * groups_per_row;
    const uint32_t xr = (uint32_t)(r & 7);
    uint16_t s_nxt[GPP], b_nxt[GPP];
    // rows of scales/biases are 2*K/G bytes: vector loads need every stage start aligned
    const bool sb_vec = (GPP == 1) || ((groups_per_row % GPP) == 0);
    auto load_sb = [&](int ps) {
      const int g0 = (kb_begin * KB + ps * KPP) / G;  // stage starts are group aligned
      if (sb_vec && g0 + GPP <= groups_per_row) {
        if constexpr (GPP == 8) {
          const uint4 sv = __ldg(reinterpret_cast<const uint4*>(srow + g0));
          const uint4 bv = __ldg(reinterpret_cast<const uint4*>(brow + g0));
          const uint32_t sw[4] = {sv.x, sv.y, sv.z, sv.w}, bw[4] = {bv.x, bv.y, bv.z, bv.w};
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            s_nxt[2 * i] = (uint16_t)(sw[i] & 0xFFFFu);
            s_nxt[2 * i + 1] = (uint16_t)(sw[i] >> 16);
            b_nxt[2 * i] = (uint16_t)(bw[i] & 0xFFFFu);
            b_nxt[2 * i + 1] = (uint16_t)(bw[i] >> 16);
          }
        } else if constexpr (GPP == 4) {
          const uint2 sv = __ldg(reinterpret_cast<const uint2*>(srow + g0));
          const uint2 bv = __ldg(reinterpret_cast<const uint2*>(brow + g0));
          s_nxt[0] = (uint16_t)(sv.x & 0xFFFFu); s_nxt[1] = (uint16_t)(sv.x >> 16);
          s_nxt[2] = (uint16_t)(sv.y & 0xFFFFu); s_nxt[3] = (uint16_t)(sv.y >> 16);
          b_nxt[0] = (uint16_t)(bv.x & 0xFFFFu); b_nxt[1] = (uint16_t)(bv.x >> 16);
          b_nxt[2] = (uint16_t)(bv.y & 0xFFFFu); b_nxt[3] = (uint16_t)(bv.y >> 16);
        } else if constexpr (GPP == 2) {
          const uint32_t sv = __ldg(reinterpret_cast<const uint32_t*>(srow + g0));
          const uint32_t bv = __ldg(reinterpret_cast<const uint32_t*>(brow + g0));
          s_nxt[0] = (uint16_t)(sv & 0xFFFFu); s_nxt[1] = (uint16_t)(sv >> 16);
          b_nxt[0] = (uint16_t)(bv & 0xFFFFu); b_nxt[1] = (uint16_t)(bv >> 16);
        } else {
          s_nxt[0] = __ldg(srow + g0);
          b_nxt[0] = __ldg(brow + g0);
        }
        return;
      }
#pragma unroll
      for (int i = 0; i < GPP; ++i) {
        const bool ok = g0 + i < groups_per_row;
        s_nxt[i] = ok ? __ldg(srow + g0 + i) : (uint16_t)0;
        b_nxt[i] = ok ? __ldg(brow + g0 + i) : (uint16_t)0;
      }
    };
    // SBT: this packed stage's GPP scales and biases sit in the TMA-staged ring, 16 B per row
    auto load_sb_smem = [&](int ps) {
      const int gps = gps0 + ps;
      const int sbi = gps / SB_PER - gps0 / SB_PER, sst = sbi % SBS;
      if (ps == 0 || gps % SB_PER == 0) mbar_wait(bar_full_sb + 8 * sst, (sbi / SBS) & 1);
      const uint32_t sa = sm_sb + sst * SB_STAGE_BYTES + r * 16 + (gps % SB_PER) * (GPP * 2);
      const uint32_t ba = sa + ROWS * 16;
      uint32_t sw[4] = {0, 0, 0, 0}, bw[4] = {0, 0, 0, 0};
      if constexpr (GPP == 8) {
        asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(sw[0]), "=r"(sw[1]), "=r"(sw[2]), "=r"(sw[3]) : "r"(sa));
        asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(bw[0]), "=r"(bw[1]), "=r"(bw[2]), "=r"(bw[3]) : "r"(ba));
      } else if constexpr (GPP == 4) {
        asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(sw[0]), "=r"(sw[1]) : "r"(sa));
        asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(bw[0]), "=r"(bw[1]) : "r"(ba));
      } else if constexpr (GPP == 2) {
        asm volatile("ld.shared.u32 %0, [%1];" : "=r"(sw[0]) : "r"(sa));
        asm volatile("ld.shared.u32 %0, [%1];" : "=r"(bw[0]) : "r"(ba));
      } else {
        asm volatile("ld.shared.u16 %0, [%1];" : "=r"(sw[0]) : "r"(sa));
        asm volatile("ld.shared.u16 %0, [%1];" : "=r"(bw[0]) : "r"(ba));
      }
#pragma unroll
      for (int i = 0; i < GPP; ++i) {
        s_nxt[i] = (uint16_t)((sw[i >> 1] >> (16 * (i & 1))) & 0xFFFFu);
        b_nxt[i] = (uint16_t)((bw[i >> 1] >> (16 * (i & 1))) & 0xFFFFu);
      }
    };
    if constexpr (!SBT) {
      if (my_ps > 0) load_sb(0);
    }
    long long w_full = 0, w_empty = 0, w_sb = 0;
    const long long c_loop = lt_clock();
    for (int ps = 0; ps < my_ps; ++ps) {
      const int pst = ps % PS;
      {
        const long long c_s = lt_clock();
        if constexpr (SBT) load_sb_smem(ps);
        w_sb += lt_clock() - c_s;
      }
      // this stage's scales/biases, broadcast once; then prefetch the next stage's
      // (G == 32 with two threads per row: a thread only ever touches group 2j + hsel of k block j)
      constexpr bool SEL = (G < KB) && (TPR == 2);
      constexpr int NG = SEL ? GPP / 2 : GPP;
      T2 s2[NG], b2[NG];
      float sf[NG], bf[NG];
#pragma unroll
      for (int i = 0; i < NG; ++i) {
        uint16_t sr, br;
        if constexpr (SEL) {
          sr = hsel ? s_nxt[2 * i + 1] : s_nxt[2 * i];
          br = hsel ? b_nxt[2 * i + 1] : b_nxt[2 * i];
        } else {
          sr = s_nxt[i];
          br = b_nxt[i];
        }
        if constexpr (BITS == 4) {
          s2[i] = __float2half2_rn(CI::to_f(sr));
          b2[i] = __float2half2_rn(CI::to_f(br));
        } else {
          sf[i] = CI::to_f(sr);
          bf[i] = CI::to_f(br);
        }
      }
      (void)s2; (void)b2; (void)sf; (void)bf;
      if constexpr (SBT) {
        // values are in registers now: hand the scale/bias stage back after its last packed stage
        const int gps = gps0 + ps;
        if (gps % SB_PER == SB_PER - 1 || ps == my_ps - 1) {
          __syncwarp();
          if (lane == 0) mbar_arrive(bar_empty_sb + 8 * ((gps / SB_PER - gps0 / SB_PER) % SBS));
        }
      } else {
        if (ps + 1 < my_ps) load_sb(ps + 1);
      }
      const long long c_a = lt_clock();
      mbar_wait(bar_full_p + 8 * pst, (ps / PS) & 1);
      w_full += lt_clock() - c_a;
      if (ps == 0 && threadIdx.x == 0) trace_stamp(p, 3);  // first packed stage landed
#ifdef QGEMM_LOOP_TRACE
      if (ps < 16 && threadIdx.x == 0) trace_stamp(p, 16 + ps);  // dequant sees packed stage ps
#endif
      const uint32_t prow = sm_p + pst * P_STAGE_BYTES + r * 128;
#pragma unroll
      for (int j = 0; j < KB_PER_P; ++j) {
        const int kb = ps * KB_PER_P + j;
        if (kb < my_kb) {
          const int st = kb % NS;
          // all of this thread's packed chunks first (independent loads), then the ring slot
          uint4 pk[CH_T];
#pragma unroll
          for (int cc = 0; cc < CH_T; ++cc) {
            const uint32_t c = (uint32_t)(hsel * CH_T + cc);
            const uint32_t src = prow + ((((uint32_t)(j * CH_PER_KB) + c) ^ xr) << 4);
            asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];"
                         : "=r"(pk[cc].x), "=r"(pk[cc].y), "=r"(pk[cc].z), "=r"(pk[cc].w)
                         : "r"(src));
          }
          if constexpr (!TA) mbar_wait(bar_empty_ab + 8 * st, ((kb / NS) & 1) ^ 1);
          const uint32_t arow = sm_a + st * A_STAGE_BYTES + r * 128;
          uint32_t av[16];  // TA: this thread's 32 k of the block, in k order
          (void)av;
#pragma unroll
          for (int cc = 0; cc < CH_T; ++cc) {
            const int c = hsel * CH_T + cc;
            if constexpr (BITS == 4) {  // chunk = 4 words x 8 nibbles = 32 k
              const uint32_t words[4] = {pk[cc].x, pk[cc].y, pk[cc].z, pk[cc].w};
              // group of this 32-k chunk: one per k block for G >= 64, one per chunk for G == 32
              T2 sg, bg;
              if constexpr (G >= KB) {
                sg = s2[(j * KB) / G];
                bg = b2[(j * KB) / G];
              } else if constexpr (SEL) {
                sg = s2[j];
                bg = b2[j];
              } else {  // one thread per row: cc is a compile-time chunk index
                sg = s2[j * 2 + (cc & 1)];
                bg = b2[j * 2 + (cc & 1)];
              }
              uint4 o[4];
#pragma unroll
              for (int wi = 0; wi < 4; ++wi) o[wi] = dequant_word4<__half, PRE>(words[wi], sg, bg);
              if constexpr (TA) {
#pragma unroll
                for (int wi = 0; wi < 4; ++wi) {
                  av[wi * 4 + 0] = o[wi].x; av[wi * 4 + 1] = o[wi].y;
                  av[wi * 4 + 2] = o[wi].z; av[wi * 4 + 3] = o[wi].w;
                }
              } else {
#pragma unroll
                for (int wi = 0; wi < 4; ++wi) {
                  const uint32_t chunk = (uint32_t)(c * 4 + wi);  // 8 elements = 16 B of the A row
                  asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(arow + ((chunk ^ xr) << 4)),
                               "r"(o[wi].x), "r"(o[wi].y), "r"(o[wi].z), "r"(o[wi].w));
                }
              }
            } else {  // chunk = 4 words x 4 bytes = 16 k
              float sc, bi;
              if constexpr (G >= KB) {
                sc = sf[(j * KB) / G];
                bi = bf[(j * KB) / G];
              } else if constexpr (SEL) {  // G == 32: chunks 0,1 -> first group of the k block, 2,3 -> second
                sc = sf[j];
                bi = bf[j];
              } else {
                sc = sf[j * 2 + (cc >> 1)];
                bi = bf[j * 2 + (cc >> 1)];
              }
              const uint2 lo0 = dequant_word8<__half>(pk[cc].x, sc, bi), lo1 = dequant_word8<__half>(pk[cc].y, sc, bi);
              const uint2 hi0 = dequant_word8<__half>(pk[cc].z, sc, bi), hi1 = dequant_word8<__half>(pk[cc].w, sc, bi);
              if constexpr (TA) {
                av[cc * 8 + 0] = lo0.x; av[cc * 8 + 1] = lo0.y; av[cc * 8 + 2] = lo1.x; av[cc * 8 + 3] = lo1.y;
                av[cc * 8 + 4] = hi0.x; av[cc * 8 + 5] = hi0.y; av[cc * 8 + 6] = hi1.x; av[cc * 8 + 7] = hi1.y;
              } else {
                const uint32_t chunk = (uint32_t)(c * 2);
                asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(arow + ((chunk ^ xr) << 4)),
                             "r"(lo0.x), "r"(lo0.y), "r"(lo1.x), "r"(lo1.y));
                asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(arow + (((chunk + 1) ^ xr) << 4)),
                             "r"(hi0.x), "r"(hi0.y), "r"(hi1.x), "r"(hi1.y));
              }
            }
          }
          if constexpr (TA) {
            // the math above ran ahead of the ring: only the store waits for the slot
            const long long c_e = lt_clock();
            mbar_wait(bar_empty_ab + 8 * st, ((kb / NS) & 1) ^ 1);
            w_empty += lt_clock() - c_e;
            tc_fence_after();
            tmem_st16(tmem_base + ((uint32_t)((warp & 3) * 32) << 16) + ACC_COLS + st * A_COLS + hsel * 16, av);
            tc_fence_before();
          } else {
            fence_proxy_async();  // generic-proxy stores -> visible to the tensor core (async proxy)
          }
          __syncwarp();
          if (lane == 0) mbar_arrive(bar_full_ab + 8 * st);
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_empty_p + 8 * pst);
    }

    if (threadIdx.x == 0) {
      trace_stamp(p, 4);  // dequant done
      trace_put(p, 8, (unsigned long long)(lt_clock() - c_loop));
      trace_put(p, 9, (unsigned long long)w_full);
      trace_put(p, 10, (unsigned long long)w_empty);
      trace_put(p, 11, (unsigned long long)w_sb);
    }
    // ===== epilogue (dequant warps: thread <-> TMEM lane r <-> output feature n0 + r;
    //       with 8 warps the batch columns are split in two halves) =====
    if (my_kb > 0) {
      mbar_wait(bar_tmem_full, 0);
      tc_fence_after();
    }
    if (threadIdx.x == 0) trace_stamp(p, 5);  // accumulator complete
    pdl_wait();  // y may still be read by the previous kernel when this rank had no k blocks
    const uint32_t taddr = tmem_base + ((uint32_t)((warp & 3) * 32) << 16);
    const int nn = n0 + r;
    constexpr int CW = NB < 32 ? 16 : 32;  // columns per tcgen05.ld
    constexpr int ESPLIT = (NB / TPR >= CW) ? TPR : 1;  // column shares among the row's threads
    constexpr int ECOLS = NB / ESPLIT;
    const bool epi_active = hsel < ESPLIT;
    const int ec0 = hsel * ECOLS;
    if (ksplit == 1) {
      T* yp = reinterpret_cast<T*>(p.y);
      const float bias_v =
          (p.bias != nullptr && nn < p.N) ? to_float(reinterpret_cast<const T*>(p.bias)[nn]) : 0.f;
      if (epi_active) {
#pragma unroll 1
        for (int c0 = ec0; c0 < ec0 + ECOLS; c0 += CW) {
          uint32_t v[CW];
          if (my_kb > 0) {
            tmem_ld<CW>(taddr + c0, v);
          } else {
#pragma unroll
            for (int i = 0; i < CW; ++i) v[i] = 0u;
          }
          if (nn < p.N) {
#pragma unroll
            for (int i = 0; i < CW; ++i) {
              const int m = m0 + c0 + i;
              if (m < p.M) yp[(int64_t)m * p.y_stride + nn] = from_float<T>(__uint_as_float(v[i]) + bias_v);
            }
          }
        }
      }
    } else if (epi_active) {
      // split-K: stage the fp32 partial [NB][128] in our own smem (the operand rings are idle now)
      const uint32_t red = TA ? sm_p : sm_a;
#pragma unroll 1
      for (int c0 = ec0; c0 < ec0 + ECOLS; c0 += CW) {
        uint32_t v[CW];
        if (my_kb > 0) {
          tmem_ld<CW>(taddr + c0, v);
        } else {
#pragma unroll
          for (int i = 0; i < CW; ++i) v[i] = 0u;
        }
#pragma unroll
        for (int i = 0; i < CW; ++i)
          asm volatile("st.shared.u32 [%0], %1;" ::"r"(red + (uint32_t)(((c0 + i) * ROWS + r) * 4)), "r"(v[i])
                       : "memory");
      }
    }
  }

  if (ksplit > 1) {
    // every thread of every CTA of the cluster: partials are published ...
    tc_fence_before();
    cluster_sync_all();
    if constexpr (NB <= 64) {
      if (warp < DQW) {
        // Rank j owns the batch columns c with c % ksplit == j.  A warp sums, for one such column,
        // 4 consecutive output features per lane (float4) over all peers' partials through
        // DSMEM, in peer order (deterministic), and writes 8 contiguous bytes per lane.
        const int r4 = lane * 4;
        const uint32_t red = TA ? sm_p : sm_a;
        const int cols_owned = NB / ksplit;
        T* yp = reinterpret_cast<T*>(p.y);
        const T* bp = reinterpret_cast<const T*>(p.bias);
        const int nn = n0 + r4;
        for (int ci = warp; ci < cols_owned; ci += DQW) {
          const int c = krank + ci * ksplit;
          float4 v[8];
#pragma unroll
          for (int peer = 0; peer < 8; ++peer)
            if (peer < ksplit) v[peer] = ld_dsmem_v4(red + (uint32_t)((c * ROWS + r4) * 4), (uint32_t)peer);
          float4 acc = v[0];
#pragma unroll
          for (int peer = 1; peer < 8; ++peer)
            if (peer < ksplit) {
              acc.x += v[peer].x;
              acc.y += v[peer].y;
              acc.z += v[peer].z;
              acc.w += v[peer].w;
            }
          const int m = m0 + c;
          if (m >= p.M) continue;
          const float o[4] = {acc.x, acc.y, acc.z, acc.w};
          if (nn + 3 < p.N && ((p.y_stride & 3) == 0) && (((uintptr_t)yp & 7) == 0)) {
            const float b0 = bp ? to_float(bp[nn]) : 0.f, b1 = bp ? to_float(bp[nn + 1]) : 0.f;
            const float b2v = bp ? to_float(bp[nn + 2]) : 0.f, b3 = bp ? to_float(bp[nn + 3]) : 0.f;
            uint2 pkd;
            pkd.x = CI::f2(o[0] + b0, o[1] + b1);
            pkd.y = CI::f2(o[2] + b2v, o[3] + b3);
            *reinterpret_cast<uint2*>(yp + (int64_t)m * p.y_stride + nn) = pkd;
          } else {
#pragma unroll
            for (int e = 0; e < 4; ++e)
              if (nn + e < p.N)
                yp[(int64_t)m * p.y_stride + nn + e] =
                    from_float<T>(o[e] + (bp ? to_float(bp[nn + e]) : 0.f));
          }
        }
      }
    }
    // ... and nobody leaves (freeing its smem) before every peer has finished reading
    cluster_sync_all();
  }

  tc_fence_before();
  __syncthreads();
  if (threadIdx.x == 0) trace_stamp(p, 6);  // epilogue (and cluster reduce) done
  if (warp == W_MMA) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(TMEM_COLS)
                 : "memory");
  }
}

struct HostArgs {
  const void* x;
  const uint32_t* w;
  const void* scales;
  const void* biases;
  const void* bias;
  void* y;
  int M, N, K, group_size, bits, dtype;
  int64_t x_stride, y_stride;
  bool prepacked;
};

// Scales/biases can be staged by TMA when their rows are 16-byte multiples and 16-byte aligned.
static bool sb_tma_ok(const HostArgs& a) {
  const int64_t row_bytes = (int64_t)(a.K / a.group_size) * 2;
  return row_bytes % 16 == 0 && (uintptr_t)a.scales % 16 == 0 && (uintptr_t)a.biases % 16 == 0;
}

template <typename T, int BITS, int G, int NB, int PS, int NS, int DQW, bool SBT = false, bool PRE = false,
          bool TA = false>
static int launch_cfg(const HostArgs& a, int ksplit, cudaStream_t st) {
  constexpr int smem = smem_bytes<NB, PS, NS, SBT, TA>();
  auto kern = qgemm_tc_kernel<T, BITS, G, NB, PS, NS, DQW, SBT, PRE, TA>;
  constexpr int KB_PER_P = (1024 / BITS) / KB;
  static bool configured_dev[16] = {false};  // cudaFuncSetAttribute is per device
  int dev_ = 0;
  B200_CUDA_OK(cudaGetDevice(&dev_));
  bool& configured = configured_dev[dev_ & 15];
  if (!configured) {
    B200_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    configured = true;
  }
  CUtensorMap w_map, x_map, s_map, b_map;
  const uint64_t words = (uint64_t)a.K * BITS / 32;
  int rc = make_map_2d(&w_map, CU_TENSOR_MAP_DATA_TYPE_UINT32, a.w, words, (uint64_t)a.N, words * 4, 32, ROWS,
                       CU_TENSOR_MAP_L2_PROMOTION_L2_256B);
  if (rc != B200_OK) return rc;
  rc = make_map_2d(&x_map, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, a.x, (uint64_t)a.K, (uint64_t)a.M, (uint64_t)a.x_stride * 2, KB, NB,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_128B);
  if (rc != B200_OK) return rc;
  if constexpr (SBT) {
    const uint64_t groups = (uint64_t)(a.K / a.group_size);
    rc = make_map_2d(&s_map, CU_TENSOR_MAP_DATA_TYPE_UINT16, a.scales, groups, (uint64_t)a.N, groups * 2, 8, ROWS,
                     CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_SWIZZLE_NONE);
    if (rc != B200_OK) return rc;
    rc = make_map_2d(&b_map, CU_TENSOR_MAP_DATA_TYPE_UINT16, a.biases, groups, (uint64_t)a.N, groups * 2, 8, ROWS,
                     CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_SWIZZLE_NONE);
    if (rc != B200_OK) return rc;
  } else {
    s_map = w_map;
    b_map = w_map;
  }
  Params p;
  p.scales = reinterpret_cast<const uint16_t*>(a.scales);
  p.biases = reinterpret_cast<const uint16_t*>(a.biases);
  p.bias = a.bias;
  p.y = a.y;
  p.M = a.M;
  p.N = a.N;
  p.K = a.K;
  p.group_size = a.group_size;
  p.y_stride = a.y_stride;
  p.num_kb = a.K / KB;
  // ranks start on packed-stage (hence group) boundaries
  p.trace = g_trace;
  p.kb_per_split = ((p.num_kb + ksplit - 1) / ksplit + KB_PER_P - 1) / KB_PER_P * KB_PER_P;
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3((a.N + ROWS - 1) / ROWS, ksplit, (a.M + NB - 1) / NB);
  cfg.blockDim = dim3((DQW + 3) * 32);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 1;
  attr[0].val.clusterDim.y = ksplit;
  attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
  cfg.attrs = attr;
  cfg.numAttrs = 2;
  B200_CUDA_OK(cudaLaunchKernelEx(&cfg, kern, w_map, x_map, s_map, b_map, p));
  return B200_OK;
}

// Tuning knobs (b200_set_qlinear_tuning): 0 = heuristic.
static int g_force_ksplit = 0;
static int g_force_dqw = 0;
static int g_variant = 0;  // b200_set_qlinear_variant: 0 = A in TMEM, 1 = legacy smem A + LDG scales, 2 = smem A + TMA scales

// M <= 64 weight-streaming configurations: NB batch tile, (PS, NS) ring depths.
template <typename T, int BITS, int G, int NB, bool PRE>
static int launch_stream(const HostArgs& a, int ksplit, cudaStream_t st) {
  constexpr bool kSbtShape = (8 * G) % (1024 / BITS) == 0 && (8 * G) / (1024 / BITS) <= 4;
  const bool dq8 = g_force_dqw ? g_force_dqw == 8 : true;
  if constexpr (kSbtShape) {
    if (g_variant != 1 && sb_tma_ok(a)) {
      if (g_variant == 2) return launch_cfg<T, BITS, G, NB, 2, 3, 8, true, PRE, false>(a, ksplit, st);
      return launch_cfg<T, BITS, G, NB, 4, 4, 8, true, PRE, true>(a, ksplit, st);
    }
  }
  return dq8 ? launch_cfg<T, BITS, G, NB, 3, 2, 8, false, PRE>(a, ksplit, st)
             : launch_cfg<T, BITS, G, NB, 3, 2, 4, false, PRE>(a, ksplit, st);
}

template <typename T, int BITS, int G>
static int dispatch_g(const HostArgs& a, cudaStream_t st) {
  const int tiles = (a.N + ROWS - 1) / ROWS;
  const int num_kb = a.K / KB;
  constexpr int KB_PER_P = (1024 / BITS) / KB;
  if (a.M <= 64) {
    // Weight-streaming regime.  Two CTAs fit per SM; matrices with few row tiles are split along
    // K over a cluster until every SM has CTAs streaming, keeping >= 2 packed stages per rank.
    const int sms = sm_count() > 0 ? sm_count() : 148;
    int ksplit = 1;
    // (tiles >= sms: every SM already streams; measured best unsplit -- profiles/r01_kernels_v3_sweep.txt)
    // ... and >= 12 k blocks per rank unless the split still leaves SMs idle (TP shards): below that the per-CTA fill + cluster reduce (~6 us) outweighs
    // the extra parallelism, and tiles * ksplit stays within one wave of 2 CTAs/SM
    // (qkv 6144x4096: ksplit 8 = 384 CTAs, two waves, 19.1 us; ksplit 4 = 13.8 us -- r01_kernels_v4 sweep)
    while (tiles < sms && ksplit < 8 && tiles * ksplit < 2 * sms && num_kb / (ksplit * 2) >= 2 * KB_PER_P &&
           tiles * ksplit * 2 <= 2 * sms && (num_kb / (ksplit * 2) >= 12 || tiles * ksplit * 2 <= sms))
      ksplit *= 2;
    if (g_force_ksplit > 0) ksplit = g_force_ksplit;
    while (ksplit > 1 && num_kb / ksplit < KB_PER_P) ksplit /= 2;
    if (a.M <= 16)
      return a.prepacked ? launch_stream<T, BITS, G, 16, BITS == 4>(a, ksplit, st)
                         : launch_stream<T, BITS, G, 16, false>(a, ksplit, st);
    return a.prepacked ? launch_stream<T, BITS, G, 64, BITS == 4>(a, ksplit, st)
                       : launch_stream<T, BITS, G, 64, false>(a, ksplit, st);
  }
  if (a.prepacked && BITS == 4) {
    if (a.M <= 128) return launch_cfg<T, BITS, G, 128, 3, 3, 8, false, BITS == 4>(a, 1, st);
    return launch_cfg<T, BITS, G, 256, 2, 3, 8, false, BITS == 4>(a, 1, st);
  }
  if (a.M <= 128) return launch_cfg<T, BITS, G, 128, 3, 3, 8>(a, 1, st);
  return launch_cfg<T, BITS, G, 256, 2, 3, 8>(a, 1, st);
}

template <typename T, int BITS>
static int dispatch(const HostArgs& a, cudaStream_t st) {
  switch (a.group_size) {
    case 32: return dispatch_g<T, BITS, 32>(a, st);
    case 64: return dispatch_g<T, BITS, 64>(a, st);
    case 128: return dispatch_g<T, BITS, 128>(a, st);
    default: return set_error(B200_ERR_INVALID_ARGUMENT, "qgemm_tc: bad group size %d", a.group_size);
  }
}

}  // namespace qtc

// Shapes the tensor-core path takes; everything else runs on the mma.sync kernel in qlinear.cu.
bool qgemm_tc_supported(const void* x, const uint32_t* w, int M, int N, int K, int group_size, int bits,
                        int x_dtype, int dtype, int sb_dtype, int64_t x_stride) {
  if (x_dtype != B200_F16) return false;  // operands are fp16 (see the kernel head): a bf16 x is cast first
  if (dtype != B200_BF16) return false;   // fp16 results stay on the mma.sync kernel
  if (sb_dtype != dtype) return false;
  if (K % 64 != 0 || K % group_size != 0) return false;
  if ((uintptr_t)x % 16 || (uintptr_t)w % 16 || (x_stride * 2) % 16) return false;
  if (((int64_t)K * bits / 8) % 16) return false;
  (void)M;
  (void)N;
  return tc::encode_fn() != nullptr;
}

int qgemm_tc_set_tuning(int ksplit, int dequant_warps) {
  if (!(ksplit == 0 || ksplit == 1 || ksplit == 2 || ksplit == 4 || ksplit == 8))
    return set_error(B200_ERR_INVALID_ARGUMENT, "qlinear tuning: ksplit must be 0, 1, 2, 4 or 8, got %d", ksplit);
  if (!(dequant_warps == 0 || dequant_warps == 4 || dequant_warps == 8))
    return set_error(B200_ERR_INVALID_ARGUMENT, "qlinear tuning: dequant_warps must be 0, 4 or 8, got %d",
                     dequant_warps);
  qtc::g_force_ksplit = ksplit;
  qtc::g_force_dqw = dequant_warps;
  return B200_OK;
}

void qgemm_tc_set_trace(void* dev_ptr) { qtc::g_trace = reinterpret_cast<unsigned long long*>(dev_ptr); }

int qgemm_tc_set_variant(int v) {
  if (v < 0 || v > 2) return set_error(B200_ERR_INVALID_ARGUMENT, "qlinear variant must be 0, 1 or 2, got %d", v);
  qtc::g_variant = v;
  return B200_OK;
}

int qgemm_tc_launch(const void* x, const uint32_t* w, const void* scales, const void* biases,
                    const void* bias, void* y, int M, int N, int K, int group_size, int bits, int dtype,
                    int64_t x_stride, int64_t y_stride, bool prepacked, cudaStream_t st) {
  qtc::HostArgs a{x, w, scales, biases, bias, y, M, N, K, group_size, bits, dtype, x_stride, y_stride, prepacked};
  return bits == 4 ? qtc::dispatch<__nv_bfloat16, 4>(a, st) : qtc::dispatch<__nv_bfloat16, 8>(a, st);
}

}  // namespace b200
