// Persistent stream-K affine-quantized GEMM for the decode regime (and, tiled over M, for any M):
//
//   y[m,n] = sum_k x[m,k] * (s[n,k/g] * q[n,k] + b[n,k/g]) (+ bias[n])          (EPI_PLAIN)
//   y[m,i] = silu(gate[m,i]) * up[m,i],  gate|up = rows [0,I) | [I,2I) of W      (EPI_SWIGLU)
//
// Replaces mx.quantized_matmul(x, w, scales, biases, transpose=True, ...) as the reference's decode
// step calls it (vllm_metal/gguf/mlx_native.py:212-222; nn.QuantizedLinear from
// attention/impls/sdpa.py:297-343,739; the SwiGLU the reference fuses in compiled_mlp.py:1-22).
//
// Why a second kernel next to qgemm_tc.cu: at M <= 64 the op is a weight stream (HBM-bound), and
// round 1's one-CTA-per-row-tile kernel spent it on per-CTA prologue/epilogue, a 0.76-wave grid and
// one mbarrier hand-off per 64 k.  Here:
//
//   * ONE persistent CTA per SM.  The (row super-tile x k-slab) units of the matrix are cut into
//     gridDim.x equal contiguous ranges (stream-K), so every SM streams the same number of bytes
//     whatever N and K are; barrier init, TMEM allocation and the tensor-map fetch happen once.
//   * The weights live in a load-time "streaming" layout (b200_qstream_prepack): per unit one
//     contiguous blob [2 tiles][2 k-halves][2 chunks][128 rows][16 B] + fp16 (scale, bias) pairs,
//     fetched by ONE cp.async.bulk per unit (no tensor map), read back by the dequant threads with
//     conflict-free 16-byte shared loads.
//   * A super-tile is 256 weight rows = two UMMA M=128 tiles that share every x tile (B operand):
//     x is re-read from L2 by every CTA, so sharing halves the L2->SM traffic per weight byte.
//   * 16 dequant warps (one thread per row and k-half) expand int4/int8 -> fp16 with the magic-number
//     trick (int4: 1 SHF + 4 LOP3 + 4 HADD2 + 4 HFMA2 per 8 weights, no PRMT; int8: 2 PRMT + 2 HADD2
//     + 2 HFMA2 per 4) and store the A tile straight into TENSOR MEMORY (tcgen05.st); the MMA is the
//     TS form, fp16 A x bf16/fp16 B -> fp32 in TMEM.  fp16 keeps 11 significant bits of s*q+b (bf16: 8),
//     which is what holds the 2e-2/1e-2 parity bound at K = 25600 / 28672.
//   * Accumulators are double-buffered in TMEM; 4 epilogue warps drain tile i (bias, SwiGLU, bf16,
//     smem transpose, 16-byte coalesced stores) while the main loop runs tile i+1.
//   * A super-tile cut between CTAs is finished by whichever CTA arrives last: every partial goes to
//     an fp32 workspace slot, an atomic counter elects the finisher, and it adds the partials in
//     k ORDER -- the result is deterministic and independent of arrival order.
//
// Algorithmic bytes per launch: N*K*(bits/8 + 4/g) + (M*K + M*N_out)*2.
#include "tcgen05.cuh"

namespace b200 {
namespace qs {
using namespace tc;

constexpr int TILE_ROWS = 128;          // UMMA M
constexpr int SUPER_ROWS = 2 * TILE_ROWS;
constexpr int DQ_WARPS = 16;            // (tile, k-half, lane quarter)
constexpr int W_WPROD = DQ_WARPS, W_XPROD = DQ_WARPS + 1, W_MMA = DQ_WARPS + 2, W_EPI = DQ_WARPS + 3;
constexpr int EPI_WARPS = 4;
constexpr int THREADS = (DQ_WARPS + 3 + EPI_WARPS) * 32;  // 736
constexpr int W_PACKED_BYTES = 2 * 2 * 2 * TILE_ROWS * 16;  // 16 KiB of packed words per unit
constexpr int EPI_PLAIN = 0, EPI_SWIGLU = 1;
constexpr int STAGE_M = 64;             // output rows staged per epilogue pass
constexpr int64_t X16_BYTES = 8ll << 20; // workspace region holding the fp16 copy of a bf16 x (M * K <= 4 Mi elements)

template <int BITS>
struct Geo {
  static constexpr int KT = 256 / BITS;   // weights per dequant thread and unit (32 B packed): 64 | 32
  static constexpr int KS = 2 * KT;       // k per unit ("slab"): 128 | 64
  static constexpr int A_COLS = KS / 2;   // TMEM columns of one tile's A slab: 64 | 32
  static constexpr int MMAS = KS / 16;    // k steps per tile and slab: 8 | 4
};
// (scale, bias) entries per row and slab in the blob
__host__ __device__ constexpr int groups_per_slab(int bits, int g) {
  return (2 * (256 / bits)) / g > 0 ? (2 * (256 / bits)) / g : 1;
}
__host__ __device__ constexpr int unit_bytes(int bits, int g) {
  return W_PACKED_BYTES + 2 * groups_per_slab(bits, g) * TILE_ROWS * 4;
}

struct Params {
  const uint8_t* blob;
  const void* bias;        // dense [N] in the x dtype, or null (EPI_PLAIN only)
  void* y;
  float* workspace;        // [2 * grid][2][128][NB] fp32 partial accumulators
  int* counters;           // arrivals [m_tiles * super_tiles] | finished [..] at +2^17 ints; zero between launches
  int M, N, K;             // N = weight rows (2*I for SwiGLU)
  int n_out;               // output columns: N, or I for SwiGLU
  int super_tiles, slabs;  // T2, S
  int m_tiles;
  int64_t y_stride;
  int y_fp16;              // output / bias / SwiGLU rounding dtype: 1 = fp16, 0 = bf16 (x, the B operand, is always fp16)
  int y_store_f16;         // with y_fp16 = 0: store the bf16-rounded result as fp16 (saturating) for a following qstream
  unsigned long long* trace;
};

__device__ __forceinline__ void bulk_load(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(dst), "l"(src), "r"(bytes), "r"(bar)
               : "memory");
}
__device__ __forceinline__ void named_bar_sync(int id, int threads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(threads) : "memory");
}
__device__ __forceinline__ void tmem_st32(uint32_t taddr, const uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,"
      "%29,%30,%31,%32};"
      ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]),
        "r"(v[8]), "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]), "r"(v[16]),
        "r"(v[17]), "r"(v[18]), "r"(v[19]), "r"(v[20]), "r"(v[21]), "r"(v[22]), "r"(v[23]), "r"(v[24]),
        "r"(v[25]), "r"(v[26]), "r"(v[27]), "r"(v[28]), "r"(v[29]), "r"(v[30]), "r"(v[31])
      : "memory");
  asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
      "{%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};"
      ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]),
        "r"(v[8]), "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15])
      : "memory");
  asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t lop3_and_or(uint32_t a, uint32_t mask, uint32_t orv) {
  uint32_t v;
  asm("lop3.b32 %0, %1, %2, %3, 0xEA;" : "=r"(v) : "r"(a), "r"(mask), "r"(orv));  // (a & mask) | orv
  return v;
}
__device__ __forceinline__ uint32_t h2_as_u32(__half2 h) { return *reinterpret_cast<uint32_t*>(&h); }
__device__ __forceinline__ __half2 u32_as_h2(uint32_t u) { return *reinterpret_cast<__half2*>(&u); }

// One streaming-layout int4 word (nibbles at bit 0,16 | 4,20 | 8,24 | 12,28 = k0,k1 | k2,k3 | k4,k5 | k6,k7)
// -> 8 fp16 values s*q+b in k order.  0x6400 = 1024.0 (ulp 1), 0x5400 = 64.0 (ulp 1/16): a nibble OR-ed
// into mantissa bits 0..3 / 4..7 reads 1024+q / 64+q; the subtraction is exact, the FMA rounds once.
__device__ __forceinline__ void dequant_word4(uint32_t w, __half2 s2, __half2 b2, uint32_t* out) {
  const uint32_t m_lo = 0x000F000Fu, m_hi = 0x00F000F0u, k_lo = 0x64006400u, k_hi = 0x54005400u;
  const __half2 o_lo = __float2half2_rn(-1024.f), o_hi = __float2half2_rn(-64.f);
  const uint32_t w8 = w >> 8;
  const __half2 q0 = __hadd2(u32_as_h2(lop3_and_or(w, m_lo, k_lo)), o_lo);
  const __half2 q1 = __hadd2(u32_as_h2(lop3_and_or(w, m_hi, k_hi)), o_hi);
  const __half2 q2 = __hadd2(u32_as_h2(lop3_and_or(w8, m_lo, k_lo)), o_lo);
  const __half2 q3 = __hadd2(u32_as_h2(lop3_and_or(w8, m_hi, k_hi)), o_hi);
  out[0] = h2_as_u32(__hfma2(s2, q0, b2));
  out[1] = h2_as_u32(__hfma2(s2, q1, b2));
  out[2] = h2_as_u32(__hfma2(s2, q2, b2));
  out[3] = h2_as_u32(__hfma2(s2, q3, b2));
}
// One int8 word (bytes k0..k3) -> 4 fp16 values: byte OR-ed under 0x64 reads 1024+q (q < 1024: exact).
__device__ __forceinline__ void dequant_word8(uint32_t w, __half2 s2, __half2 b2, uint32_t* out) {
  const __half2 off = __float2half2_rn(-1024.f);
  const __half2 q0 = __hadd2(u32_as_h2(__byte_perm(w, 0x64646464u, 0x5140)), off);
  const __half2 q1 = __hadd2(u32_as_h2(__byte_perm(w, 0x64646464u, 0x7362)), off);
  out[0] = h2_as_u32(__hfma2(s2, q0, b2));
  out[1] = h2_as_u32(__hfma2(s2, q1, b2));
}

__device__ __forceinline__ void trace_stamp(const Params& p, int slot) {
  if (p.trace != nullptr) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    p.trace[(size_t)blockIdx.x * 16 + slot] = t;
  }
}

// fine-grained pipeline trace of CTA 0 (SM clock): event e of unit i at trace[148 * 16 + i * 8 + e], i < 64
__device__ __forceinline__ void unit_stamp(const Params& p, int i, int e) {
  if (p.trace != nullptr && blockIdx.x == 0 && i < 64) {
    long long t;
    asm volatile("mov.u64 %0, %%clock64;" : "=l"(t));
    p.trace[148 * 16 + i * 8 + e] = (unsigned long long)t;
  }
}


__device__ __forceinline__ int ld_acquire_gpu(const int* ptr) {
  int v;
  asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(ptr) : "memory");
  return v;
}
// fp16: 0 = bf16, 1 = fp16, 2 = rounded to bf16 then stored as fp16 (exact below 65504, saturating)
__device__ __forceinline__ uint16_t to_out16(float v, int fp16) {
  if (fp16 == 1) {
    const __half h = __float2half_rn(v);
    return *reinterpret_cast<const uint16_t*>(&h);
  }
  const __nv_bfloat16 h = __float2bfloat16_rn(v);
  if (fp16 == 2) {
    const __half g = __float2half_rn(fminf(fmaxf(__bfloat162float(h), -65504.f), 65504.f));
    return *reinterpret_cast<const uint16_t*>(&g);
  }
  return *reinterpret_cast<const uint16_t*>(&h);
}
__device__ __forceinline__ float round_out(float v, int fp16) {
  return fp16 == 1 ? __half2float(__float2half_rn(v)) : __bfloat162float(__float2bfloat16_rn(v));
}
// gate and up are rounded to the activation dtype first, as the unfused sequence stores them
// (oracle/decoder.py silu_mul: silu rounded, then the product rounded)
__device__ __forceinline__ uint16_t swiglu_out(float gate, float up, int fp16) {
  const float g = round_out(gate, fp16), u = round_out(up, fp16);
  const float sl = round_out(g / (1.f + __expf(-g)), fp16);
  return to_out16(sl * u, fp16);
}
__device__ __forceinline__ float load_bias(const Params& p, int n) {
  if (p.bias == nullptr || n >= p.N) return 0.f;
  return p.y_fp16 == 1 ? __half2float(reinterpret_cast<const __half*>(p.bias)[n])
                  : __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(p.bias)[n]);
}
// four consecutive outputs n..n+3 of one y row
__device__ __forceinline__ void store_out4(uint16_t* yrow, int n, int n_out, const uint16_t (&o)[4], bool vec_ok) {
  if (n >= n_out) return;
  if (vec_ok && n + 4 <= n_out) {
    *reinterpret_cast<uint2*>(yrow + n) = make_uint2((uint32_t)o[0] | ((uint32_t)o[1] << 16), (uint32_t)o[2] | ((uint32_t)o[3] << 16));
  } else {
#pragma unroll
    for (int q = 0; q < 4; ++q)
      if (n + q < n_out) yrow[n + q] = o[q];
  }
}

template <int NB>
struct Cfg {
  static constexpr int ACC_COLS = NB < 32 ? 32 : NB;
  static constexpr int ACC_BUFS = NB <= 64 ? 2 : 1;
};

template <int BITS, int G, int NB>
struct Smem {
  static constexpr int WB = unit_bytes(BITS, G);
  static constexpr int XB = NB * Geo<BITS>::KS * 2;
  static constexpr int STAGE = STAGE_M * SUPER_ROWS * 2;  // 32 KiB bf16 output staging
  static constexpr int XST = NB <= 32 ? 4 : 3;
  static constexpr int BUDGET = 227 * 1024 - 1024 /*barriers*/ - 1024 /*align slack*/;
  static constexpr int WST_RAW = (BUDGET - STAGE - XST * XB) / WB;
  static constexpr int WST = WST_RAW > 12 ? 12 : WST_RAW;
  static constexpr int A_SLOT_COLS = 2 * Geo<BITS>::A_COLS;
  static constexpr int A_SLOTS_RAW = (512 - 2 * Cfg<NB>::ACC_BUFS * Cfg<NB>::ACC_COLS) / A_SLOT_COLS;
  static constexpr int A_SLOTS = A_SLOTS_RAW > 4 ? 4 : A_SLOTS_RAW;
  static constexpr int TOTAL = WST * WB + XST * XB + STAGE + 1024 + 1024;
  static_assert(WST >= 3, "weight ring too shallow");
  static_assert(A_SLOTS >= 2, "A ring too shallow");
};

// v in [0, V): which CTA's range holds it (ranges are [floor(c*V/P), floor((c+1)*V/P)))
__device__ __forceinline__ int cta_of(long long v, long long V, int P) {
  return (int)(((v + 1) * P + V - 1) / V) - 1;
}

template <int BITS, int G, int NB, int EPI>
__global__ void __launch_bounds__(THREADS, 1)
qgemm_stream_kernel(const __grid_constant__ CUtensorMap x_map, const Params p) {
  using GE = Geo<BITS>;
  using SM = Smem<BITS, G, NB>;
  constexpr int KT = GE::KT, KS = GE::KS, A_COLS = GE::A_COLS, MMAS = GE::MMAS;
  constexpr int GS = groups_per_slab(BITS, G);
  constexpr int NG = GS >= 2 ? GS / 2 : 1;  // (scale, bias) entries one dequant thread uses per slab
  constexpr int WB = SM::WB, XB = SM::XB, WST = SM::WST, XST = SM::XST, AS = SM::A_SLOTS;
  constexpr int ACC_COLS = Cfg<NB>::ACC_COLS, ACC_BUFS = Cfg<NB>::ACC_BUFS;
  constexpr int TM_ACC = 0, TM_A = 2 * ACC_BUFS * ACC_COLS;
  constexpr int XBOXES = KS / 64;  // 64-k (128 B) swizzle atoms per x slab
  static_assert(TM_A + AS * SM::A_SLOT_COLS <= 512, "tensor memory overflow");

  pdl_trigger();
  if (threadIdx.x == 0) trace_stamp(p, 0);
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t sm_x = base;                    // XST x tiles (1 KiB aligned: SWIZZLE_128B)
  const uint32_t sm_stage = sm_x + XST * XB;     // output staging
  const uint32_t sm_w = sm_stage + SM::STAGE;    // WST weight blobs
  const uint32_t sm_bar = sm_w + WST * WB;
  const uint32_t bar_w_full = sm_bar, bar_w_empty = bar_w_full + 8 * WST;
  const uint32_t bar_x_full = bar_w_empty + 8 * WST, bar_x_empty = bar_x_full + 8 * XST;
  const uint32_t bar_a_full = bar_x_empty + 8 * XST, bar_a_empty = bar_a_full + 8 * AS;
  const uint32_t bar_acc_full = bar_a_empty + 8 * AS, bar_acc_empty = bar_acc_full + 8 * 2;
  const uint32_t tmem_ptr_addr = bar_acc_empty + 8 * 2;
  const uint32_t bar_fix = tmem_ptr_addr + 8;  // 2 single-use barriers of the split finish
  static_assert(8 * (2 * WST + 2 * XST + 2 * AS + 4) + 8 + 16 <= 1024, "barrier block overflow");
  uint8_t* gen = smem_raw + (base - smem_u32(smem_raw));
  volatile uint32_t* tmem_ptr_gen = reinterpret_cast<volatile uint32_t*>(gen + (tmem_ptr_addr - base));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int P = gridDim.x;
  const long long U = (long long)p.super_tiles * p.slabs;  // units per m tile
  const long long V = U * p.m_tiles;
  const long long v_begin = (long long)blockIdx.x * V / P, v_end = (long long)(blockIdx.x + 1) * V / P;

  if (threadIdx.x == 0) {
    for (int i = 0; i < WST; ++i) {
      mbar_init(bar_w_full + 8 * i, 1);
      mbar_init(bar_w_empty + 8 * i, DQ_WARPS);
    }
    for (int i = 0; i < XST; ++i) {
      mbar_init(bar_x_full + 8 * i, 1);
      mbar_init(bar_x_empty + 8 * i, 1);
    }
    for (int i = 0; i < AS; ++i) {
      mbar_init(bar_a_full + 8 * i, DQ_WARPS);
      mbar_init(bar_a_empty + 8 * i, 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(bar_acc_full + 8 * i, 1);
      mbar_init(bar_acc_empty + 8 * i, EPI_WARPS);
      mbar_init(bar_fix + 8 * i, 1);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == W_MMA) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_ptr_addr), "n"(512)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_gen;
  if (threadIdx.x == 0) trace_stamp(p, 1);

  // The three single-issuer roles run their loops with the WHOLE warp (uniform control flow, uniform operands)
  // and issue under elect_one(): see tcgen05.cuh.
  const int vb = (int)v_begin, ve = (int)v_end, Ui = (int)U;
  if (warp == W_WPROD) {
    // ===== weight blobs: one bulk copy per unit; constants, so no pdl_wait (see b200_qstream header note) =====
    int st = 0;
    uint32_t ph = 1;
    int u = vb % Ui;
    for (int v = vb; v < ve; ++v) {
      mbar_wait_parked(bar_w_empty + 8 * st, ph);
#ifdef QS_ABL_NO_W
      if (v - vb >= WST) {
        if (elect_one()) mbar_arrive(bar_w_full + 8 * st);
      } else
#endif
      if (elect_one()) {
        mbar_expect_tx(bar_w_full + 8 * st, WB);
        bulk_load(sm_w + st * WB, p.blob + (size_t)u * WB, WB, bar_w_full + 8 * st);
      }
      __syncwarp();
      if (++u == Ui) u = 0;
      if (++st == WST) { st = 0; ph ^= 1; }
    }
  } else if (warp == W_XPROD) {
    // ===== x tiles (B operand): [NB rows x KS k], 128B-swizzled K-major, through the tensor map =====
    pdl_wait();  // x is the previous kernel's output
    if (lane == 0) trace_stamp(p, 2);
    int st = 0;
    uint32_t ph = 1;
    int mt = vb / Ui;
    int j = (vb % Ui) % p.slabs, T = (vb % Ui) / p.slabs;
    for (int v = vb; v < ve; ++v) {
      mbar_wait_parked(bar_x_empty + 8 * st, ph);
#ifdef QS_ABL_NO_X
      if (v - vb >= XST) {
        if (elect_one()) mbar_arrive(bar_x_full + 8 * st);
      } else
#endif
      if (elect_one()) {
        mbar_expect_tx(bar_x_full + 8 * st, XB);
#pragma unroll
        for (int b = 0; b < XBOXES; ++b)
          tma_load_2d(sm_x + st * XB + b * (NB * 128), &x_map, bar_x_full + 8 * st, j * KS + b * 64, mt * NB);
      }
      __syncwarp();
      if (++j == p.slabs) {
        j = 0;
        if (++T == p.super_tiles) { T = 0; ++mt; }
      }
      if (++st == XST) { st = 0; ph ^= 1; }
    }
  } else if (warp == W_MMA) {
    // ===== MMA issuer: D[128 x NB] (+)= A[128 x 16] (TMEM, fp16) * B[16 x NB] (smem), two tiles per slab =====
    // fp16 A (TMEM) x fp16 B (smem) -> fp32: kind::f16 faults on mixed fp16/bf16 operands (measured)
    constexpr uint32_t idesc = make_idesc(0, TILE_ROWS, NB);
    const uint32_t tb = __shfl_sync(0xffffffffu, tmem_base, 0);  // warp-uniform for ptxas
    int xs = 0, as = 0, seg = 0;
    uint32_t xph = 0, aph = 0;
    int v = vb;
    while (v < ve) {
      const int tile_id = v / p.slabs;
      const int seg_end = min((tile_id + 1) * p.slabs, ve);
      const int buf = seg % ACC_BUFS;
      mbar_wait(bar_acc_empty + 8 * buf, ((seg / ACC_BUFS) & 1) ^ 1);
      uint32_t acc = 0;  // the first k step of a segment overwrites the accumulator (both tiles)
      for (; v < seg_end; ++v) {
        mbar_wait(bar_x_full + 8 * xs, xph);
        mbar_wait(bar_a_full + 8 * as, aph);
#ifdef QS_UNIT_TRACE
        if (lane == 0) unit_stamp(p, v - vb, 3);
#endif
        tc_fence_after();
        if (elect_one()) {
          const uint64_t db0 = make_smem_desc(sm_x + xs * XB);
#pragma unroll
          for (int t = 0; t < 2; ++t) {
            const uint32_t d = tb + TM_ACC + (buf * 2 + t) * ACC_COLS;
            const uint32_t a = tb + TM_A + (as * 2 + t) * A_COLS;
#pragma unroll
            for (int k = 0; k < MMAS; ++k) {
              // k-step inside a 128B swizzle atom: +32 B (>>4 = 2); next atom (64 k): + NB * 128 B
              const uint64_t db = db0 + (uint64_t)((k >> 2) * ((NB * 128) >> 4) + 2 * (k & 3));
#ifndef QS_ABL_NO_MMA
              tc_mma_f16_ts(d, a + 8 * k, db, idesc, k == 0 ? acc : 1u);
#else
              (void)d; (void)a; (void)db;
#endif
            }
          }
          tc_commit(bar_x_empty + 8 * xs);
          tc_commit(bar_a_empty + 8 * as);
        }
        __syncwarp();
        acc = 1u;
#ifdef QS_UNIT_TRACE
        if (lane == 0) unit_stamp(p, v - vb, 4);
#endif
        if (++xs == XST) { xs = 0; xph ^= 1; }
        if (++as == AS) { as = 0; aph ^= 1; }
      }
      if (elect_one()) tc_commit(bar_acc_full + 8 * buf);
      __syncwarp();
      ++seg;
    }
    if (lane == 0) trace_stamp(p, 6);
  } else if (warp < DQ_WARPS) {
    // ===== dequant: thread = (tile t, k-half h, row r); 32 B packed -> KT fp16 -> TMEM =====
    const int th = warp >> 2;                 // t * 2 + h
    const int t = th >> 1, h = th & 1;
    const int r = (warp & 3) * 32 + lane;
    const uint32_t w_off0 = (uint32_t)(((th * 2 + 0) * TILE_ROWS + r) * 16);
    const uint32_t w_off1 = (uint32_t)(((th * 2 + 1) * TILE_ROWS + r) * 16);
    const uint32_t sb_off = (uint32_t)(W_PACKED_BYTES + ((t * GS + (GS >= 2 ? h * NG : 0)) * TILE_ROWS + r) * 4);
    const uint32_t lane_addr = ((uint32_t)((warp & 3) * 32) << 16);
    int i = 0;
    for (long long v = v_begin; v < v_end; ++v, ++i) {
      const int ws = i % WST, as = i % AS;
      mbar_wait(bar_w_full + 8 * ws, (i / WST) & 1);
      if (i == 0 && threadIdx.x == 0) trace_stamp(p, 5);
#ifdef QS_UNIT_TRACE
      if (threadIdx.x == 0) unit_stamp(p, i, 0);
#endif
      const uint32_t wbase = sm_w + ws * WB;
      uint4 pk0, pk1;
      asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(pk0.x), "=r"(pk0.y), "=r"(pk0.z), "=r"(pk0.w) : "r"(wbase + w_off0));
      asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(pk1.x), "=r"(pk1.y), "=r"(pk1.z), "=r"(pk1.w) : "r"(wbase + w_off1));
      uint32_t sb[NG];
#pragma unroll
      for (int g = 0; g < NG; ++g)
        asm volatile("ld.shared.u32 %0, [%1];" : "=r"(sb[g]) : "r"(wbase + sb_off + g * (TILE_ROWS * 4)));
      // the words are in registers: hand the weight stage back to the producer
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_w_empty + 8 * ws);
      const uint32_t words[8] = {pk0.x, pk0.y, pk0.z, pk0.w, pk1.x, pk1.y, pk1.z, pk1.w};
      __half2 s2[NG], b2[NG];
#pragma unroll
      for (int g = 0; g < NG; ++g) {
        s2[g] = u32_as_h2(__byte_perm(sb[g], 0, 0x1010));
        b2[g] = u32_as_h2(__byte_perm(sb[g], 0, 0x3232));
      }
      uint32_t av[KT / 2];
#ifdef QS_ABL_NO_MATH
#pragma unroll
      for (int q = 0; q < KT / 2; ++q) av[q] = words[q & 7] ^ sb[0];
      if (false)
#endif
      if constexpr (BITS == 4) {
#pragma unroll
        for (int wi = 0; wi < 8; ++wi) dequant_word4(words[wi], s2[NG == 2 ? (wi >> 2) : 0], b2[NG == 2 ? (wi >> 2) : 0], av + 4 * wi);
      } else {
#pragma unroll
        for (int wi = 0; wi < 8; ++wi) dequant_word8(words[wi], s2[NG == 2 ? (wi >> 2) : 0], b2[NG == 2 ? (wi >> 2) : 0], av + 2 * wi);
      }
      // the math ran ahead of the ring: only the TMEM store waits for the slot (MMA of slab i - AS done)
#ifdef QS_UNIT_TRACE
      if (threadIdx.x == 0) unit_stamp(p, i, 1);
#endif
      mbar_wait(bar_a_empty + 8 * as, ((i / AS) & 1) ^ 1);
#ifdef QS_UNIT_TRACE
      if (threadIdx.x == 0) unit_stamp(p, i, 2);
#endif
      tc_fence_after();
      const uint32_t taddr = tmem_base + lane_addr + TM_A + (as * 2 + t) * A_COLS + h * (KT / 2);
#ifndef QS_ABL_NO_STTM
      if constexpr (BITS == 4) tmem_st32(taddr, av); else tmem_st16(taddr, av);
#else
      if (av[0] == 0x12345678u && av[KT / 2 - 1] == 0x9abcdef0u) tmem_st16(taddr, reinterpret_cast<const uint32_t(&)[16]>(av));
#endif
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_a_full + 8 * as);
#ifdef QS_UNIT_TRACE
      if (threadIdx.x == 0) unit_stamp(p, i, 5);
#endif
    }
    if (threadIdx.x == 0) trace_stamp(p, 3);
  } else {
    // ===== epilogue warps: TMEM lane quarter = warp % 4; thread <-> weight row r of both tiles =====
    const int r = (warp & 3) * 32 + lane;
    const int ew = warp & 3;
    const int et = threadIdx.x - W_EPI * 32;   // 0..127
    const uint32_t lane_addr = ((uint32_t)((warp & 3) * 32) << 16);
    constexpr int OUT_COLS = EPI == EPI_SWIGLU ? TILE_ROWS : SUPER_ROWS;
    uint16_t* yp = reinterpret_cast<uint16_t*>(p.y);
    pdl_wait();  // y, the workspace and the counters may still be in use by the previous kernel
    if (et == 0) trace_stamp(p, 9);
    int seg = 0;
    int v = (int)v_begin;
    const int ve = (int)v_end;
    while (v < ve) {
      const int tile_id = v / p.slabs;  // mt * super_tiles + T
      const int j0 = v - tile_id * p.slabs;
      const int seg_end = min((tile_id + 1) * p.slabs, ve);
      const int j1 = seg_end - tile_id * p.slabs;
      const int mt = tile_id / p.super_tiles, T = tile_id % p.super_tiles;
      const int buf = seg % ACC_BUFS;
      const bool whole = (j0 == 0 && j1 == p.slabs);
      mbar_wait_parked(bar_acc_full + 8 * buf, (seg / ACC_BUFS) & 1);
      if (seg == 0 && et == 0) trace_stamp(p, 7);
      tc_fence_after();
      const uint32_t acc0 = tmem_base + lane_addr + TM_ACC + (buf * 2) * ACC_COLS;
      if (!whole) {
        // park the fp32 partial as [tile t][column c][row r] (coalesced both here and in the finish), publish it
        float* slot = p.workspace + ((size_t)(2 * blockIdx.x + (j0 == 0 ? 1 : 0)) * 2 * TILE_ROWS) * NB;
#pragma unroll
        for (int t = 0; t < 2; ++t) {
#pragma unroll 1
          for (int c0 = 0; c0 < NB; c0 += 16) {
            uint32_t vv[16];
            tmem_ld<16>(acc0 + t * ACC_COLS + c0, vv);
#pragma unroll
            for (int q = 0; q < 16; ++q) __stcg(slot + ((size_t)(t * NB + c0 + q)) * TILE_ROWS + r, __uint_as_float(vv[q]));
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(bar_acc_empty + 8 * buf);
        __threadfence();
        named_bar_sync(1, EPI_WARPS * 32);
        if (et == 0) atomicAdd(p.counters + tile_id, 1);
      } else {
        const int n_base = EPI == EPI_SWIGLU ? T * TILE_ROWS : T * SUPER_ROWS;  // first output column of this tile
        for (int mh = 0; mh < NB; mh += STAGE_M) {
          constexpr int MH = NB < STAGE_M ? NB : STAGE_M;
#pragma unroll 1
          for (int c0 = mh; c0 < mh + MH; c0 += 16) {
            float a0[16], a1[16];
            {
              uint32_t vv[16];
              tmem_ld<16>(acc0 + c0, vv);
#pragma unroll
              for (int q = 0; q < 16; ++q) a0[q] = __uint_as_float(vv[q]);
              tmem_ld<16>(acc0 + ACC_COLS + c0, vv);
#pragma unroll
              for (int q = 0; q < 16; ++q) a1[q] = __uint_as_float(vv[q]);
            }
            if constexpr (EPI == EPI_SWIGLU) {
#pragma unroll
              for (int q = 0; q < 16; ++q) {
                const uint16_t o = swiglu_out(a0[q], a1[q], p.y_fp16);
                asm volatile("st.shared.u16 [%0], %1;" ::"r"(sm_stage + (uint32_t)(((c0 - mh + q) * OUT_COLS + r) * 2)), "h"(o) : "memory");
              }
            } else {
              const float bias0 = load_bias(p, n_base + r), bias1 = load_bias(p, n_base + TILE_ROWS + r);
#pragma unroll
              for (int q = 0; q < 16; ++q) {
                const uint32_t row = sm_stage + (uint32_t)((c0 - mh + q) * OUT_COLS * 2);
                asm volatile("st.shared.u16 [%0], %1;" ::"r"(row + (uint32_t)(r * 2)), "h"(to_out16(a0[q] + bias0, p.y_fp16)) : "memory");
                asm volatile("st.shared.u16 [%0], %1;" ::"r"(row + (uint32_t)((TILE_ROWS + r) * 2)), "h"(to_out16(a1[q] + bias1, p.y_fp16)) : "memory");
              }
            }
          }
          if (mh + STAGE_M >= NB) {  // last TMEM read of this buffer: give it back to the MMA warp
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(bar_acc_empty + 8 * buf);
          }
          named_bar_sync(1, EPI_WARPS * 32);
          // staged [MH rows][OUT_COLS] -> y, one 16-byte vector per lane; a warp covers 512 B of a row
          {
            constexpr int VEC_PER_ROW = OUT_COLS / 8;
            const bool vec_ok = (p.y_stride % 8 == 0) && ((reinterpret_cast<uintptr_t>(p.y) & 15) == 0);
            for (int idx = et; idx < MH * VEC_PER_ROW; idx += EPI_WARPS * 32) {
              const int mr = idx / VEC_PER_ROW, vc = idx % VEC_PER_ROW;
              const int m = mt * NB + mh + mr, n = n_base + vc * 8;
              if (m >= p.M || n >= p.n_out) continue;
              uint4 val;
              asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];"
                           : "=r"(val.x), "=r"(val.y), "=r"(val.z), "=r"(val.w)
                           : "r"(sm_stage + (uint32_t)((mr * OUT_COLS + vc * 8) * 2)));
              uint16_t* dst = yp + (int64_t)m * p.y_stride + n;
              if (vec_ok && n + 8 <= p.n_out) {
                *reinterpret_cast<uint4*>(dst) = val;
              } else {
                const uint16_t* e = reinterpret_cast<const uint16_t*>(&val);
                for (int q = 0; q < 8 && n + q < p.n_out; ++q) dst[q] = e[q];
              }
            }
          }
          named_bar_sync(1, EPI_WARPS * 32);  // staging free for the next pass
        }
      }
      v = seg_end;
      ++seg;
    }
    if (et == 0) trace_stamp(p, 8);

  }

  // ----- finish the cut super-tiles (dequant + epilogue warps, 640 threads).  The CTAs that hold pieces of one
  // split its (valid) columns between them; each waits until every piece is published, pulls ITS columns of
  // every piece into the (now idle) weight ring with bulk copies -- all in flight at once -- and adds them in
  // k ORDER: deterministic whatever the timing.  Lane l owns rows 4l..4l+3, each warp takes every 20th column. -----
  if (warp < DQ_WARPS || warp >= W_EPI) {
    constexpr int FIN_WARPS = DQ_WARPS + EPI_WARPS, FIN_THREADS = FIN_WARPS * 32;
    const int wj = warp < DQ_WARPS ? warp : DQ_WARPS + (warp - W_EPI);
    const bool leader = threadIdx.x == W_EPI * 32;  // the thread that published this CTA's pieces
    uint16_t* yp = reinterpret_cast<uint16_t*>(p.y);
    const int vb = (int)v_begin, ve = (int)v_end;
    int pend_tile[2];
    int npend = 0;
    if (ve > vb) {
      const int t_first = vb / p.slabs, t_last = (ve - 1) / p.slabs;
      const int end_first = min((t_first + 1) * p.slabs, ve);
      if (!(vb == t_first * p.slabs && end_first == (t_first + 1) * p.slabs)) pend_tile[npend++] = t_first;
      if (t_last != t_first && ve != (t_last + 1) * p.slabs) pend_tile[npend++] = t_last;
    }
    int f_cfirst[2], f_nseg[2], f_ca[2], f_cb[2];
    uint32_t f_off[2];
#pragma unroll
    for (int pi = 0; pi < 2; ++pi) {
      f_cfirst[pi] = f_nseg[pi] = f_ca[pi] = f_cb[pi] = 0;
      if (pi < npend) {
        const int tile_id = pend_tile[pi];
        const int mt = tile_id / p.super_tiles;
        const long long v_first = (long long)tile_id * p.slabs;
        f_cfirst[pi] = cta_of(v_first, V, P);
        f_nseg[pi] = cta_of(v_first + p.slabs - 1, V, P) - f_cfirst[pi] + 1;
        const int s_me = blockIdx.x - f_cfirst[pi];
        const int mvalid = min(NB, p.M - mt * NB);
        f_ca[pi] = s_me * mvalid / f_nseg[pi];
        f_cb[pi] = (s_me + 1) * mvalid / f_nseg[pi];
      }
    }
    f_off[0] = 0;
    f_off[1] = (uint32_t)(f_nseg[0] * 2 * (f_cb[0] - f_ca[0]) * TILE_ROWS * 4);
    // both fetches in flight together when they fit the ring, else one after the other
    const bool together = npend == 2 && f_off[1] + (uint32_t)(f_nseg[1] * 2 * (f_cb[1] - f_ca[1]) * TILE_ROWS * 4) <= (uint32_t)(WST * WB);
    if (!together) f_off[1] = 0;
    auto fetch = [&](int pi) {  // leader only
      const int ncol = f_cb[pi] - f_ca[pi];
      if (ncol <= 0) return;
      uint32_t spins = 0;
      while (ld_acquire_gpu(p.counters + pend_tile[pi]) < f_nseg[pi]) {
        if (++spins > (1u << 26)) __trap();
      }
      if (pi == 0) trace_stamp(p, 11);
      asm volatile("fence.proxy.async;" ::: "memory");  // peers' generic-proxy stores -> this thread's bulk reads
      const uint32_t bytes = (uint32_t)(ncol * TILE_ROWS * 4);
      mbar_expect_tx(bar_fix + 8 * pi, bytes * 2 * f_nseg[pi]);
      for (int sgi = 0; sgi < f_nseg[pi]; ++sgi) {
        const float* slot = p.workspace + ((size_t)(2 * (f_cfirst[pi] + sgi) + (sgi == 0 ? 1 : 0)) * 2 * TILE_ROWS) * NB;
#pragma unroll
        for (int t = 0; t < 2; ++t)
          bulk_load(sm_w + f_off[pi] + (uint32_t)(sgi * 2 + t) * bytes, slot + (size_t)(t * NB + f_ca[pi]) * TILE_ROWS, bytes,
                    bar_fix + 8 * pi);
      }
    };
    if (leader) {
      if (npend > 0) fetch(0);
      if (together) fetch(1);
    }
    for (int pi = 0; pi < npend; ++pi) {
      const int tile_id = pend_tile[pi];
      const int mt = tile_id / p.super_tiles, T = tile_id % p.super_tiles;
      const int ca = f_ca[pi], cb = f_cb[pi], nseg = f_nseg[pi];
      if (pi == 1 && !together) {
        named_bar_sync(2, FIN_THREADS);  // ring reads of piece 0 are over
        if (leader) fetch(1);
      }
      if (cb > ca) {
        mbar_wait_parked(bar_fix + 8 * pi, 0);
        if (leader) trace_stamp(p, 12 + pi);
        const int n_base = EPI == EPI_SWIGLU ? T * TILE_ROWS : T * SUPER_ROWS;
        const bool vec_ok = (p.y_stride % 4 == 0) && ((reinterpret_cast<uintptr_t>(p.y) & 7) == 0);
        const uint32_t col_bytes = TILE_ROWS * 4, piece_bytes = (uint32_t)(cb - ca) * col_bytes;
        for (int c = ca + wj; c < cb; c += FIN_WARPS) {
          float a[2][4];
#pragma unroll
          for (int t = 0; t < 2; ++t)
#pragma unroll
            for (int q = 0; q < 4; ++q) a[t][q] = 0.f;
          const uint32_t src0 = sm_w + f_off[pi] + (uint32_t)(c - ca) * col_bytes + lane * 16;
          for (int sgi = 0; sgi < nseg; ++sgi) {
#pragma unroll
            for (int t = 0; t < 2; ++t) {
              float4 pv;
              asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];"
                           : "=f"(pv.x), "=f"(pv.y), "=f"(pv.z), "=f"(pv.w)
                           : "r"(src0 + (uint32_t)(sgi * 2 + t) * piece_bytes));
              a[t][0] += pv.x; a[t][1] += pv.y; a[t][2] += pv.z; a[t][3] += pv.w;
            }
          }
          const int m = mt * NB + c;
          uint16_t* yrow = yp + (int64_t)m * p.y_stride;
          if constexpr (EPI == EPI_SWIGLU) {
            const int n = n_base + 4 * lane;
            uint16_t o[4];
#pragma unroll
            for (int q = 0; q < 4; ++q) o[q] = swiglu_out(a[0][q], a[1][q], p.y_fp16);
            store_out4(yrow, n, p.n_out, o, vec_ok);
          } else {
#pragma unroll
            for (int t = 0; t < 2; ++t) {
              const int n = n_base + t * TILE_ROWS + 4 * lane;
              uint16_t o[4];
#pragma unroll
              for (int q = 0; q < 4; ++q) o[q] = to_out16(a[t][q] + load_bias(p, n + q), p.y_fp16);
              store_out4(yrow, n, p.n_out, o, vec_ok);
            }
          }
        }
      }
    }
    if (npend > 0) {
      named_bar_sync(2, FIN_THREADS);  // every reader of the pieces is done
      if (leader) {
        int* done = p.counters + (1 << 17);
        for (int pi = 0; pi < npend; ++pi) {
          const int old = atomicAdd(done + pend_tile[pi], 1);
          if (old == f_nseg[pi] - 1) {  // every piece holder is past its reads: re-arm for the next launch (graph replay)
            p.counters[pend_tile[pi]] = 0;
            done[pend_tile[pi]] = 0;
          }
        }
        trace_stamp(p, 10);
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (threadIdx.x == 0) trace_stamp(p, 4);
  if (warp == W_MMA) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(512) : "memory");
  }
}

// ------------------------------------------------------------- load-time layout transform ----
struct PackParams {
  const uint32_t* w;       // MLX layout [N, K*bits/32]
  const void* scales;      // [N, K/g]
  const void* biases;
  uint8_t* blob;
  int N, K, bits, group, sb_dtype;
  int super_tiles, slabs;
  int swiglu_i;            // 0, or I: rows [0,I) are gate, [I,2I) up; super-tile T = gate rows 128T.. | up rows 128T..
  int* overflow;           // set to 1 when a scale/bias does not fit fp16
};

__device__ __forceinline__ int src_row(const PackParams& p, int T, int t, int r) {
  if (p.swiglu_i > 0) {
    const int i = T * TILE_ROWS + r;
    return i < p.swiglu_i ? t * p.swiglu_i + i : -1;
  }
  const int n = T * SUPER_ROWS + t * TILE_ROWS + r;
  return n < p.N ? n : -1;
}

__device__ __forceinline__ float sb_value(const void* ptr, int64_t idx, int dt) {
  if (dt == B200_BF16) return __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(ptr)[idx]);
  if (dt == B200_F16) return __half2float(reinterpret_cast<const __half*>(ptr)[idx]);
  return reinterpret_cast<const float*>(ptr)[idx];
}

// int4: MLX word (k0..k7 at bits 4i) -> streaming word (k0,k2,k4,k6 at bits 0,4,8,12; k1,k3,k5,k7 at 16,20,24,28)
__device__ __forceinline__ uint32_t perm4(uint32_t w) {
  uint32_t o = 0;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    o |= ((w >> (8 * i)) & 0xFu) << (4 * i);
    o |= ((w >> (8 * i + 4)) & 0xFu) << (4 * i + 16);
  }
  return o;
}
__device__ __forceinline__ uint32_t unperm4(uint32_t w) {
  uint32_t o = 0;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    o |= ((w >> (4 * i)) & 0xFu) << (8 * i);
    o |= ((w >> (4 * i + 16)) & 0xFu) << (8 * i + 4);
  }
  return o;
}

__global__ void qstream_pack_kernel(const PackParams p) {
  const int kt = 256 / p.bits, ks = 2 * kt;
  const int gs = groups_per_slab(p.bits, p.group);
  const int ub = unit_bytes(p.bits, p.group);
  const int words_per_row = p.K * p.bits / 32;
  const int64_t units = (int64_t)p.super_tiles * p.slabs;
  const int chunks = W_PACKED_BYTES / 16;           // 1024 packed chunks per unit
  const int sbn = 2 * gs * TILE_ROWS;               // (scale, bias) entries per unit
  const int64_t total = units * (chunks + sbn);
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t u = i / (chunks + sbn);
    const int e = (int)(i % (chunks + sbn));
    const int T = (int)(u / p.slabs), j = (int)(u % p.slabs);
    uint8_t* ublob = p.blob + u * ub;
    if (e < chunks) {
      const int r = e % TILE_ROWS, cc = (e / TILE_ROWS) & 1, th = e / (2 * TILE_ROWS);
      const int t = th >> 1, h = th & 1;
      const int n = src_row(p, T, t, r);
      const int k0 = j * ks + h * kt + cc * (kt / 2);       // first k of this 16-byte chunk
      const int w0 = k0 * p.bits / 32;
      uint4 o = make_uint4(0, 0, 0, 0);
      uint32_t* ov = reinterpret_cast<uint32_t*>(&o);
      if (n >= 0) {
        for (int q = 0; q < 4; ++q) {
          if (w0 + q < words_per_row) {
            const uint32_t wd = p.w[(int64_t)n * words_per_row + w0 + q];
            ov[q] = p.bits == 4 ? perm4(wd) : wd;
          }
        }
      }
      reinterpret_cast<uint4*>(ublob)[e] = o;
    } else {
      const int s = e - chunks;
      const int r = s % TILE_ROWS, gidx = (s / TILE_ROWS) % gs, t = s / (TILE_ROWS * gs);
      const int n = src_row(p, T, t, r);
      const int k = j * ks + gidx * (ks / gs);
      uint32_t o = 0;
      if (n >= 0 && k < p.K) {
        const int64_t gi = (int64_t)n * (p.K / p.group) + k / p.group;
        const float sv = sb_value(p.scales, gi, p.sb_dtype), bv = sb_value(p.biases, gi, p.sb_dtype);
        const __half sh = __float2half_rn(sv), bh = __float2half_rn(bv);
        if ((isinf(__half2float(sh)) && !isinf(sv)) || (isinf(__half2float(bh)) && !isinf(bv))) *p.overflow = 1;
        o = (uint32_t)(*reinterpret_cast<const unsigned short*>(&sh)) |
            ((uint32_t)(*reinterpret_cast<const unsigned short*>(&bh)) << 16);
      }
      reinterpret_cast<uint32_t*>(ublob + W_PACKED_BYTES)[s] = o;
    }
  }
}

// blob -> MLX-layout packed words (bit-exact inverse of the integer part)
__global__ void qstream_unpack_kernel(const PackParams p, uint32_t* out) {
  const int kt = 256 / p.bits, ks = 2 * kt;
  const int ub = unit_bytes(p.bits, p.group);
  const int words_per_row = p.K * p.bits / 32;
  const int64_t total = (int64_t)p.N * words_per_row;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int n = (int)(i / words_per_row), wk = (int)(i % words_per_row);
    int T, t, r;
    if (p.swiglu_i > 0) {
      t = n >= p.swiglu_i ? 1 : 0;
      const int ii = n - t * p.swiglu_i;
      T = ii / TILE_ROWS;
      r = ii % TILE_ROWS;
    } else {
      T = n / SUPER_ROWS;
      t = (n % SUPER_ROWS) / TILE_ROWS;
      r = n % TILE_ROWS;
    }
    const int k0 = wk * 32 / p.bits;
    const int j = k0 / ks, kin = k0 % ks;
    const int h = kin / kt, cc = (kin % kt) / (kt / 2);
    const int q = ((kin % (kt / 2)) * p.bits) / 32;
    const uint8_t* ublob = p.blob + ((int64_t)T * p.slabs + j) * ub;
    const int e = ((t * 2 + h) * 2 + cc) * TILE_ROWS + r;
    const uint32_t wd = reinterpret_cast<const uint32_t*>(ublob)[e * 4 + q];
    out[i] = p.bits == 4 ? unperm4(wd) : wd;
  }
}

// bf16 -> fp16 for the B operand.  Exact for every bf16 whose magnitude is a normal fp16 (8 significant bits fit
// in 11); magnitudes above 65504 saturate (bf16 activations of that size do not occur after a norm / SwiGLU; the
// reference would carry them, this path clamps -- stated in DESIGN.md), below 2^-24 round to zero.
__global__ void cast_bf16_f16_sat_kernel(const __nv_bfloat16* __restrict__ x, __half* __restrict__ y, int rows, int cols,
                                         int64_t x_stride) {
  pdl_trigger();
  pdl_wait();
  const int vec_per_row = cols >> 3;
  const int64_t total = (int64_t)rows * vec_per_row;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int r = (int)(i / vec_per_row), c = (int)(i % vec_per_row) * 8;
    const uint4 v = *reinterpret_cast<const uint4*>(x + (int64_t)r * x_stride + c);
    const __nv_bfloat162* b2 = reinterpret_cast<const __nv_bfloat162*>(&v);
    uint4 o;
    __half2* h2 = reinterpret_cast<__half2*>(&o);
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      float2 f = __bfloat1622float2(b2[q]);
      f.x = fminf(fmaxf(f.x, -65504.f), 65504.f);
      f.y = fminf(fmaxf(f.y, -65504.f), 65504.f);
      h2[q] = __floats2half2_rn(f.x, f.y);
    }
    *reinterpret_cast<uint4*>(y + (int64_t)r * cols + c) = o;
  }
}

// ----------------------------------------------------------------- host side ----
struct HostArgs {
  const void* x;
  const uint8_t* blob;
  const void* bias;
  void* y;
  float* workspace;
  size_t workspace_bytes;
  int* counters;
  int M, N, K, group, bits, dtype, epi, store_f16;
  int64_t x_stride, y_stride;
};

static int g_force_grid = 0;
static unsigned long long* g_trace = nullptr;

template <int BITS, int G, int NB, int EPI>
static int launch_cfg(const HostArgs& a, cudaStream_t st) {
  using SM = Smem<BITS, G, NB>;
  auto kern = qgemm_stream_kernel<BITS, G, NB, EPI>;
  static int configured_dev[16] = {0};
  int dev = 0;
  B200_CUDA_OK(cudaGetDevice(&dev));
  if (dev < 16 && !configured_dev[dev]) {
    B200_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SM::TOTAL));
    configured_dev[dev] = 1;
  }
  CUtensorMap x_map;
  int rc = make_map_2d(&x_map, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, a.x, (uint64_t)a.K, (uint64_t)a.M,
                       (uint64_t)a.x_stride * 2, 64, NB,
                       CU_TENSOR_MAP_L2_PROMOTION_L2_128B);
  if (rc != B200_OK) return rc;
  Params p;
  p.blob = a.blob;
  p.bias = a.bias;
  p.y = a.y;
  p.workspace = a.workspace;
  p.counters = a.counters;
  p.M = a.M;
  p.N = a.N;
  p.K = a.K;
  p.n_out = EPI == EPI_SWIGLU ? a.N / 2 : a.N;
  p.super_tiles = EPI == EPI_SWIGLU ? (a.N / 2 + TILE_ROWS - 1) / TILE_ROWS : (a.N + SUPER_ROWS - 1) / SUPER_ROWS;
  p.slabs = (a.K + Geo<BITS>::KS - 1) / Geo<BITS>::KS;
  p.m_tiles = (a.M + NB - 1) / NB;
  p.y_stride = a.y_stride;
  p.y_fp16 = a.dtype == B200_F16 ? 1 : (a.store_f16 ? 2 : 0);
  p.y_store_f16 = a.store_f16;
  p.trace = g_trace;
  const long long V = (long long)p.super_tiles * p.slabs * p.m_tiles;
  const int sms = sm_count() > 0 ? sm_count() : 148;
  // Persistent grid.  Few super-tiles (T < SMs): cut every tile into the same number s of equal k pieces
  // (grid = T * s): each CTA then holds ONE piece, and the finish is one exchange.  Otherwise (or when that
  // would idle > 15% of the SMs) plain stream-K over all SMs.  Every CTA keeps >= 2 slabs.
  const long long tiles = (long long)p.super_tiles * p.m_tiles;
  long long grid = sms;
  if (tiles < sms) {
    long long s = sms / tiles;
    if (s > p.slabs / 2) s = p.slabs / 2 > 0 ? p.slabs / 2 : 1;
    if (tiles * s * 100 >= (long long)sms * 85 || V / 2 < sms) grid = tiles * s;
  }
  if (grid > V / 2) grid = V / 2;
  if (grid < 1) grid = 1;
  if (g_force_grid > 0) grid = g_force_grid < V ? g_force_grid : V;
  B200_CHECK_ARG((size_t)grid * 2 * 2 * TILE_ROWS * NB * sizeof(float) <= a.workspace_bytes,
                 "qstream: workspace too small (%zu bytes) for grid %lld", a.workspace_bytes, grid);
  B200_CHECK_ARG(V < (1ll << 31), "qstream: problem too large");
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3((unsigned)grid);
  cfg.blockDim = dim3(THREADS);
  cfg.dynamicSmemBytes = SM::TOTAL;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  B200_CUDA_OK(cudaLaunchKernelEx(&cfg, kern, x_map, p));
  return B200_OK;
}

template <int BITS, int G>
static int dispatch_nb(const HostArgs& a, cudaStream_t st) {
  if (a.epi == EPI_SWIGLU) {
    if (a.M <= 16) return launch_cfg<BITS, G, 16, EPI_SWIGLU>(a, st);
    if (a.M <= 32) return launch_cfg<BITS, G, 32, EPI_SWIGLU>(a, st);
    return launch_cfg<BITS, G, 64, EPI_SWIGLU>(a, st);
  }
  if (a.M <= 16) return launch_cfg<BITS, G, 16, EPI_PLAIN>(a, st);
  if (a.M <= 32) return launch_cfg<BITS, G, 32, EPI_PLAIN>(a, st);
  return launch_cfg<BITS, G, 64, EPI_PLAIN>(a, st);
}

template <int BITS>
static int dispatch_g(const HostArgs& a, cudaStream_t st) {
  switch (a.group) {
    case 32: return dispatch_nb<BITS, 32>(a, st);
    case 64: return dispatch_nb<BITS, 64>(a, st);
    case 128: return dispatch_nb<BITS, 128>(a, st);
    default: return set_error(B200_ERR_INVALID_ARGUMENT, "qstream: bad group size %d", a.group);
  }
}

static int fill_pack(PackParams& pp, int N, int K, int bits, int group, int swiglu) {
  B200_CHECK_ARG(bits == 4 || bits == 8, "qstream: bits must be 4 or 8, got %d", bits);
  B200_CHECK_ARG(group == 32 || group == 64 || group == 128, "qstream: group_size must be 32, 64 or 128, got %d", group);
  B200_CHECK_ARG(N > 0 && K > 0 && K % group == 0 && (K * bits) % 32 == 0, "qstream: bad shape N=%d K=%d", N, K);
  B200_CHECK_ARG(!swiglu || N % 2 == 0, "qstream: SwiGLU needs an even row count, got %d", N);
  pp.N = N;
  pp.K = K;
  pp.bits = bits;
  pp.group = group;
  pp.swiglu_i = swiglu ? N / 2 : 0;
  pp.super_tiles = swiglu ? (N / 2 + TILE_ROWS - 1) / TILE_ROWS : (N + SUPER_ROWS - 1) / SUPER_ROWS;
  const int ks = 2 * (256 / bits);
  pp.slabs = (K + ks - 1) / ks;
  return B200_OK;
}

}  // namespace qs
}  // namespace b200

extern "C" {

int64_t b200_qstream_blob_bytes(int N, int K, int bits, int group_size, int swiglu) {
  using namespace b200;
  qs::PackParams pp{};
  if (qs::fill_pack(pp, N, K, bits, group_size, swiglu) != B200_OK) return -1;
  return (int64_t)pp.super_tiles * pp.slabs * qs::unit_bytes(bits, group_size);
}

int64_t b200_qstream_workspace_bytes(void) {
  const int sms = b200::sm_count() > 0 ? b200::sm_count() : 148;
  // [2 slots per CTA][2 tiles][128 rows][NB <= 64] fp32 | fp16 copy of a bf16 x | 1 MiB of tile counters at the end
  return (int64_t)sms * 2 * 2 * b200::qs::TILE_ROWS * 64 * 4 + b200::qs::X16_BYTES + (1 << 20);
}

int b200_cast_bf16_to_f16(const void* x, void* y, int rows, int cols, int64_t x_stride, void* stream) {
  using namespace b200;
  B200_CHECK_ARG(rows >= 0 && cols >= 0 && cols % 8 == 0, "cast_bf16_to_f16: cols (%d) must be a multiple of 8", cols);
  if (rows == 0 || cols == 0) return B200_OK;
  B200_CHECK_ARG(x && y, "cast_bf16_to_f16: null pointer argument");
  B200_CHECK_ARG(((uintptr_t)x % 16) == 0 && ((uintptr_t)y % 16) == 0 && (x_stride * 2) % 16 == 0 && x_stride >= cols,
                 "cast_bf16_to_f16: rows must be 16-byte aligned");
  const int64_t vecs = (int64_t)rows * (cols / 8);
  int grid = (int)((vecs + 255) / 256);
  if (grid > 148 * 8) grid = 148 * 8;
  B200_CUDA_OK(launch_pdl(qs::cast_bf16_f16_sat_kernel, dim3(grid), dim3(256), 0, (cudaStream_t)stream,
                          reinterpret_cast<const __nv_bfloat16*>(x), reinterpret_cast<__half*>(y), rows, cols, x_stride));
  return B200_OK;
}

int b200_qstream_prepack(const uint32_t* w, const void* scales, const void* biases, void* blob, int N, int K,
                         int bits, int group_size, int sb_dtype, int swiglu, void* stream) {
  using namespace b200;
  B200_CHECK_ARG(w && scales && biases && blob, "qstream_prepack: null pointer argument");
  B200_CHECK_ARG(sb_dtype == B200_BF16 || sb_dtype == B200_F16 || sb_dtype == B200_F32,
                 "qstream_prepack: unsupported scales/biases dtype code %d", sb_dtype);
  qs::PackParams pp{};
  int rc = qs::fill_pack(pp, N, K, bits, group_size, swiglu);
  if (rc != B200_OK) return rc;
  pp.w = w;
  pp.scales = scales;
  pp.biases = biases;
  pp.blob = reinterpret_cast<uint8_t*>(blob);
  pp.sb_dtype = sb_dtype;
  cudaStream_t st = (cudaStream_t)stream;
  int* flag = nullptr;
  B200_CUDA_OK(cudaMalloc(&flag, sizeof(int)));
  B200_CUDA_OK(cudaMemsetAsync(flag, 0, sizeof(int), st));
  pp.overflow = flag;
  qs::qstream_pack_kernel<<<148 * 8, 256, 0, st>>>(pp);
  cudaError_t e = cudaGetLastError();
  int h = 0;
  if (e == cudaSuccess) e = cudaMemcpyAsync(&h, flag, sizeof(int), cudaMemcpyDeviceToHost, st);
  // load-time call: the blob is stream-complete before any qstream launch may prefetch it (the kernel reads
  // weights BEFORE its programmatic-dependency wait)
  if (e == cudaSuccess) e = cudaStreamSynchronize(st);
  cudaFree(flag);
  if (e != cudaSuccess) return set_error(B200_ERR_CUDA, "qstream_prepack failed: %s", cudaGetErrorString(e));
  if (h != 0)
    return set_error(B200_ERR_UNSUPPORTED, "qstream_prepack: a scale or bias does not fit fp16; keep the MLX layout");
  return B200_OK;
}

int b200_qstream_unpack(const void* blob, uint32_t* w_out, int N, int K, int bits, int group_size, int swiglu,
                        void* stream) {
  using namespace b200;
  B200_CHECK_ARG(blob && w_out, "qstream_unpack: null pointer argument");
  qs::PackParams pp{};
  int rc = qs::fill_pack(pp, N, K, bits, group_size, swiglu);
  if (rc != B200_OK) return rc;
  pp.blob = reinterpret_cast<uint8_t*>(const_cast<void*>(blob));
  qs::qstream_unpack_kernel<<<148 * 8, 256, 0, (cudaStream_t)stream>>>(pp, w_out);
  return check_launch("qstream_unpack_kernel");
}

int b200_qstream_linear(const void* x, const void* blob, const void* bias, void* y, int M, int N, int K,
                        int group_size, int bits, int x_dtype, int dtype, int epilogue, int64_t x_stride,
                        int64_t y_stride, void* workspace, int64_t workspace_bytes, void* stream) {
  using namespace b200;
  B200_CHECK_ARG(M >= 0, "qstream_linear: M must be >= 0, got %d", M);
  if (M == 0) return B200_OK;
  B200_CHECK_ARG(x && blob && y && workspace, "qstream_linear: null pointer argument");
  B200_CHECK_ARG(dtype == B200_BF16 || dtype == B200_F16, "qstream_linear: y must be bf16 or fp16");
  B200_CHECK_ARG(x_dtype == B200_BF16 || x_dtype == B200_F16, "qstream_linear: x must be bf16 or fp16");
  B200_CHECK_ARG(epilogue >= 0 && epilogue <= 3, "qstream_linear: bad epilogue %d", epilogue);
  const int store_f16 = (epilogue & 2) ? 1 : 0;  // bf16-rounded values stored as fp16 (feeds the next qstream_linear)
  epilogue &= 1;
  B200_CHECK_ARG(!(store_f16 && dtype != B200_BF16), "qstream_linear: the fp16-store flag applies to bf16 results");
  B200_CHECK_ARG(!(epilogue == qs::EPI_SWIGLU && bias), "qstream_linear: SwiGLU epilogue takes no dense bias");
  qs::PackParams pp{};
  int rc = qs::fill_pack(pp, N, K, bits, group_size, epilogue == qs::EPI_SWIGLU);
  if (rc != B200_OK) return rc;
  const int n_out = epilogue == qs::EPI_SWIGLU ? N / 2 : N;
  B200_CHECK_ARG(y_stride >= n_out, "qstream_linear: y_stride (%lld) < output columns (%d)", (long long)y_stride, n_out);
  B200_CHECK_ARG(((uintptr_t)x % 16) == 0 && (x_stride * 2) % 16 == 0 && x_stride >= K,
                 "qstream_linear: x rows must be 16-byte aligned (stride %lld)", (long long)x_stride);
  B200_CHECK_ARG(((uintptr_t)blob % 16) == 0 && ((uintptr_t)workspace % 16) == 0, "qstream_linear: blob/workspace alignment");
  B200_CHECK_ARG(tc::encode_fn() != nullptr, "qstream_linear: cuTensorMapEncodeTiled unavailable");
  const int64_t counter_bytes = 1 << 20;
  B200_CHECK_ARG(workspace_bytes > counter_bytes + qs::X16_BYTES, "qstream_linear: workspace too small");
  cudaStream_t st = (cudaStream_t)stream;
  uint8_t* x16 = reinterpret_cast<uint8_t*>(workspace) + (workspace_bytes - counter_bytes - qs::X16_BYTES);
  if (x_dtype == B200_BF16) {
    // the tensor cores take fp16 x fp16 only (A is dequantised to fp16): stage an fp16 copy of x
    B200_CHECK_ARG((int64_t)M * K * 2 <= qs::X16_BYTES,
                   "qstream_linear: bf16 x of %d x %d exceeds the staging region; pass fp16 (b200_cast_bf16_to_f16)", M, K);
    rc = b200_cast_bf16_to_f16(x, x16, M, K, x_stride, stream);
    if (rc != B200_OK) return rc;
    x = x16;
    x_stride = K;
  }
  const int nb = M <= 16 ? 16 : M <= 32 ? 32 : 64;  // M > 64: 64-row m tiles, weights re-streamed per tile
  B200_CHECK_ARG((int64_t)pp.super_tiles * ((M + nb - 1) / nb) <= (1 << 17),
                 "qstream_linear: too many (m tile, super-tile) pairs for the counters");
  qs::HostArgs a;
  a.x = x;
  a.blob = reinterpret_cast<const uint8_t*>(blob);
  a.bias = bias;
  a.y = y;
  a.workspace = reinterpret_cast<float*>(workspace);
  a.workspace_bytes = (size_t)(workspace_bytes - counter_bytes - qs::X16_BYTES);
  a.counters = reinterpret_cast<int*>(reinterpret_cast<uint8_t*>(workspace) + (workspace_bytes - counter_bytes));
  a.M = M;
  a.N = N;
  a.K = K;
  a.group = group_size;
  a.bits = bits;
  a.dtype = dtype;
  a.epi = epilogue;
  a.store_f16 = store_f16;
  a.x_stride = x_stride;
  a.y_stride = y_stride;
  return bits == 4 ? qs::dispatch_g<4>(a, st) : qs::dispatch_g<8>(a, st);
}

int b200_set_qstream_grid(int grid) {
  b200::qs::g_force_grid = grid > 0 ? grid : 0;
  return B200_OK;
}

int b200_set_qstream_trace(void* dev_ptr) {
  b200::qs::g_trace = reinterpret_cast<unsigned long long*>(dev_ptr);
  return B200_OK;
}

}  // extern "C"
