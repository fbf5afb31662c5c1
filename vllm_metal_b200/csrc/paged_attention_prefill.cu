// Tensor-core varlen prefill / mixed-batch paged attention for sm_100a: tcgen05.mma + TMEM + TMA.
//
// Replaces the reference's FlashAttention-2 tile kernel and its tensor-unit variant
// (vllm_metal/metal/kernels_v2/pagedattention_tiled.metal:114-525, host paged_ops.cpp:375-434;
// pagedattention_nax.metal:406-449, host paged_ops.cpp:328-373).  Same contract: per (q head, Q block of one
// sequence) gather K/V through the block table, S = Q K^T, scale*log2e, mask (causal, kv >= seq_len, window
// kv < q_abs+1-W), online softmax, O += P V, sink folded once per row, out = O / (l + 1e-6); the Q-block ->
// sequence lookup is the reference's `cu_seqlens_q[i]/BQ + i` bound (:162-171).  Like the NAX variant, P is
// rounded to the operand type before the PV product (relaxed precision); everything else is fp32.
//
// One CTA = two 128-row Q tiles (256 consecutive query rows) of one (sequence, q head), ping-ponged FlashAttention-4
// style so that one tile's softmax overlaps the other tile's tensor-core work and K/V are fetched once for both:
//   warp 0      TMA producer: Q tile once; per 128-token KV block the K and V pages of the block table
//               (one box [64 d x block_size tokens] per page and 64-wide half of the head) into two 2-deep rings
//   warp 1      MMA issuer: S_i(j) = Q_i K_j^T (SS form, both K-major, 128B swizzle), O_i += P_i(j) V_j (TS form: P
//               from TMEM, V MN-major straight from the token-major cache layout), tiles i = 0, 1 interleaved
//   warps 2..5 / 6..9  softmax of Q tile 0 / 1: thread = query row; S row TMEM -> registers, mask, running max with
//               lazy rescale (O and l are rescaled only when the max grows by more than 2^8, FlashAttention-4's
//               rule), P -> TMEM over S's own columns, final sink fold, normalise, store
// TMEM: S0/P0 [0,128) | S1/P1 [128,256) | O0 [256,256+D) | O1 [256+D,256+2D).
//
// Flops per launch: 4 * D * sum_i q_i * (ctx_i + (q_i + 1) / 2) per q head (causal).
#include "paged_attention.cuh"
#include "tcgen05.cuh"

namespace b200 {
namespace pf {
using namespace tc;

constexpr int BQ = 128, BK = 128, BQ2 = 2 * BQ;
constexpr int THREADS = 10 * 32;          // TMA, MMA, 2 x 4 softmax warps
constexpr int TM_SP = 0, TM_O = 2 * BK;   // S_i / P_i at TM_SP + i * 128, O_i at TM_O + i * D
constexpr float kRescaleThreshold = 8.f;  // log2 units

struct Params {
  void* out;
  const int32_t* block_tables;
  const int32_t* seq_lens;
  const int32_t* cu;
  const float* sinks;
  int num_seqs, num_q_tokens, num_kv_heads, group;
  int block_size, pages_per_tile, max_blocks_per_seq, sliding_window, min_qlen;
  int64_t out_stride;
  float scale, softcap;
  int is_fp16;
  int out_f16;  // bf16 call storing its (bf16-rounded) output as fp16
};

// MN-major SWIZZLE_128B operand (cute::UMMA canonical layout ((8,n),(8,k)):((1,LBO),(8,SBO)) in 16-byte units):
// 64 elements of N contiguous in a 128-byte row, rows = k, 8-row groups SBO apart, 64-wide N chunks LBO apart.
__device__ __forceinline__ uint64_t make_smem_desc_mn(uint32_t addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = (uint64_t)((addr & 0x3FFFFu) >> 4);
  d |= (uint64_t)(lbo_bytes >> 4) << 16;
  d |= (uint64_t)(sbo_bytes >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}
__device__ __forceinline__ float ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ uint32_t pack2(float lo, float hi, int fp16) {
  if (fp16) {
    const __half2 v = __floats2half2_rn(lo, hi);
    return *reinterpret_cast<const uint32_t*>(&v);
  }
  const __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<const uint32_t*>(&v);
}
__device__ __forceinline__ void tmem_st32(uint32_t taddr, const uint32_t* v) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,"
      "%29,%30,%31,%32};"
      ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]),
        "r"(v[8]), "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]), "r"(v[16]),
        "r"(v[17]), "r"(v[18]), "r"(v[19]), "r"(v[20]), "r"(v[21]), "r"(v[22]), "r"(v[23]), "r"(v[24]),
        "r"(v[25]), "r"(v[26]), "r"(v[27]), "r"(v[28]), "r"(v[29]), "r"(v[30]), "r"(v[31])
      : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

template <int D>
struct Smem {
  static constexpr int HALVES = D / 64;
  static constexpr int Q_BYTES = BQ * D * 2;
  static constexpr int KV_BYTES = BK * D * 2;   // one K (or V) tile: HALVES x [128 tokens x 128 B]
  static constexpr int HALF_BYTES = BK * 128;   // 16 KiB
  static constexpr int STAGES = 2;
  static constexpr int TOTAL = 2 * Q_BYTES + 2 * STAGES * KV_BYTES + 1024 /*barriers*/ + 1024 /*align*/;
};

__device__ __forceinline__ float tanh_approx(float x) {
  float y;
  asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// SOFTCAP is a template flag so that the (rare) tanh path never sits in the hot loop's instruction stream: with it
// inline the unrolled softmax body was ~150 KB of SASS and the four softmax warps ran out of the instruction cache.
template <int D, bool SOFTCAP>
__global__ void __launch_bounds__(THREADS, 1)
prefill_tc_kernel(const __grid_constant__ CUtensorMap q_map, const __grid_constant__ CUtensorMap k_map,
                  const __grid_constant__ CUtensorMap v_map, const Params p) {
  using SM = Smem<D>;
  constexpr int HALVES = SM::HALVES, STAGES = SM::STAGES;
  pdl_trigger();
  pdl_wait();  // the metadata (seq_lens advance on the device), q, the caches come from earlier kernels of the step
  // ---- which (sequence, local 256-row Q super-tile) is this CTA: largest i with cu[i]/BQ2 + i <= tile (:162-171)
  // CTAs are dispatched in blockIdx order: hand out the LAST tiles first -- under a causal mask they walk the most
  // KV blocks, so the long ones start early and the short ones fill the tail of the last wave
  const int tile = gridDim.x - 1 - blockIdx.x, h = blockIdx.y;
  int lo_i = 0, hi_i = p.num_seqs;
  while (hi_i - lo_i > 1) {
    const int mid = (lo_i + hi_i) >> 1;
    if (p.cu[mid] / BQ2 + mid <= tile) lo_i = mid; else hi_i = mid;
  }
  const int seq = lo_i;
  const int q_start = p.cu[seq], q_len = p.cu[seq + 1] - q_start;
  const int local = tile - (q_start / BQ2 + seq);
  if (local * BQ2 >= q_len || q_len < p.min_qlen) return;  // padding tile, or a sequence the row kernels own
  const int row0 = local * BQ2;                      // first query row of the super-tile inside the sequence
  const int nrows_all = min(BQ2, q_len - row0);
  const bool two = nrows_all > BQ;                    // the second 128-row Q tile exists
  const int seq_len = p.seq_lens[seq];
  const int qabs0 = seq_len - q_len + row0;           // absolute position of row 0
  // KV positions any row of the super-tile can see: [kv_lo, kv_hi); both Q tiles walk this same block range (the
  // tile that cannot see a block masks it: at most one block at either end)
  const int kv_hi = min(seq_len, qabs0 + nrows_all);
  if (kv_hi <= 0) return;
  const int kv_lo = p.sliding_window >= 0 ? max(0, qabs0 + 1 - p.sliding_window) : 0;
  const int j0 = kv_lo / BK, j1 = (kv_hi - 1) / BK;   // KV blocks j0..j1
  const int nblk = j1 - j0 + 1;

  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t sm_q = base;                         // 2 Q tiles
  const uint32_t sm_k = sm_q + 2 * SM::Q_BYTES;
  const uint32_t sm_v = sm_k + STAGES * SM::KV_BYTES;
  const uint32_t sm_bar = sm_v + STAGES * SM::KV_BYTES;
  const uint32_t bar_q = sm_bar;
  const uint32_t bar_k_full = bar_q + 8, bar_k_empty = bar_k_full + 8 * STAGES;
  const uint32_t bar_v_full = bar_k_empty + 8 * STAGES, bar_v_empty = bar_v_full + 8 * STAGES;
  const uint32_t bar_s_full = bar_v_empty + 8 * STAGES;   // [2 tiles]
  const uint32_t bar_p_full = bar_s_full + 16;            // [2 tiles]
  const uint32_t bar_o_full = bar_p_full + 16;            // [2 tiles]
  const uint32_t tmem_ptr_addr = bar_o_full + 16;
  uint8_t* gen = smem_raw + (base - smem_u32(smem_raw));
  volatile uint32_t* tmem_ptr_gen = reinterpret_cast<volatile uint32_t*>(gen + (tmem_ptr_addr - base));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    mbar_init(bar_q, 1);
    for (int i = 0; i < STAGES; ++i) {
      mbar_init(bar_k_full + 8 * i, 1);
      mbar_init(bar_k_empty + 8 * i, 1);
      mbar_init(bar_v_full + 8 * i, 1);
      mbar_init(bar_v_empty + 8 * i, 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(bar_s_full + 8 * i, 1);
      mbar_init(bar_p_full + 8 * i, 4);
      mbar_init(bar_o_full + 8 * i, 1);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_ptr_addr), "n"(512)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_gen;

  const int kvh = h / p.group;
  if (warp == 0) {
    // ===== TMA producer =====
    if (elect_one()) {
      mbar_expect_tx(bar_q, (two ? 2 : 1) * SM::Q_BYTES);
      for (int i = 0; i < (two ? 2 : 1); ++i)
#pragma unroll
        for (int hf = 0; hf < HALVES; ++hf)
          tma_load_2d(sm_q + i * SM::Q_BYTES + hf * SM::HALF_BYTES, &q_map, bar_q, h * D + hf * 64, q_start + row0 + i * BQ);
    }
    __syncwarp();
    const int32_t* table = p.block_tables + (int64_t)seq * p.max_blocks_per_seq;
    const int page_bytes = p.block_size * 128;
    int st = 0;
    uint32_t ph = 1;
    for (int it = 0; it < nblk; ++it) {
      const int tok0 = (j0 + it) * BK;
      // block ids of this tile's pages: lane l holds page l (pages wholly past seq_len read block 0 and are masked)
      int bid_l = 0;
      if (lane < p.pages_per_tile) {
        const int pg = tok0 / p.block_size + lane;
        if (pg < p.max_blocks_per_seq && pg * p.block_size < seq_len) bid_l = table[pg];
      }
      mbar_wait(bar_k_empty + 8 * st, ph);
      if (elect_one()) mbar_expect_tx(bar_k_full + 8 * st, SM::KV_BYTES);
      __syncwarp();
      for (int pg = 0; pg < p.pages_per_tile; ++pg) {
        const int bid = __shfl_sync(0xffffffffu, bid_l, pg);
        if (elect_one()) {
#pragma unroll
          for (int hf = 0; hf < HALVES; ++hf)
            tma_load_2d(sm_k + st * SM::KV_BYTES + hf * SM::HALF_BYTES + pg * page_bytes, &k_map, bar_k_full + 8 * st,
                        kvh * D + hf * 64, bid * p.block_size);
        }
        __syncwarp();
      }
      mbar_wait(bar_v_empty + 8 * st, ph);
      if (elect_one()) mbar_expect_tx(bar_v_full + 8 * st, SM::KV_BYTES);
      __syncwarp();
      for (int pg = 0; pg < p.pages_per_tile; ++pg) {
        const int bid = __shfl_sync(0xffffffffu, bid_l, pg);
        if (elect_one()) {
#pragma unroll
          for (int hf = 0; hf < HALVES; ++hf)
            tma_load_2d(sm_v + st * SM::KV_BYTES + hf * SM::HALF_BYTES + pg * page_bytes, &v_map, bar_v_full + 8 * st,
                        kvh * D + hf * 64, bid * p.block_size);
        }
        __syncwarp();
      }
      if (++st == STAGES) { st = 0; ph ^= 1; }
    }
  } else if (warp == 1) {
    // ===== MMA issuer.  The two Q tiles ping-pong: while one tile's softmax warps work on S_i(j), the tensor pipe
    // runs the other tile's S / PV.  S_i and P_i share tensor memory (P overwrites the first half of S's columns):
    // tcgen05.mma executes in issue order, so S_i(j+1), issued after PV_i(j), cannot overtake it. =====
    const uint32_t fmt = p.is_fp16 ? 0u : 1u;
    const uint32_t idesc_s = (1u << 4) | (fmt << 7) | (fmt << 10) | ((uint32_t)(BK >> 3) << 17) | ((uint32_t)(BQ >> 4) << 24);
    // PV: B (V) is MN-major (bit 16)
    const uint32_t idesc_o = (1u << 4) | (fmt << 7) | (fmt << 10) | (1u << 16) | ((uint32_t)(D >> 3) << 17) |
                             ((uint32_t)(BQ >> 4) << 24);
    const uint32_t tb = __shfl_sync(0xffffffffu, tmem_base, 0);
    const int ntile = two ? 2 : 1;
    mbar_wait(bar_q, 0);
    auto issue_s = [&](int it, int i, bool last_tile) {  // S_i(it) = Q_i K_it^T into SP_i
      const int st = it % STAGES;
      if (i == 0) mbar_wait(bar_k_full + 8 * st, (it / STAGES) & 1);
      tc_fence_after();
      if (elect_one()) {
#pragma unroll
        for (int k = 0; k < D / 16; ++k) {
          const uint32_t off = (uint32_t)((k >> 2) * SM::HALF_BYTES + (k & 3) * 32);
          tc_mma_f16(tb + TM_SP + i * BK, make_smem_desc(sm_q + i * SM::Q_BYTES + off),
                     make_smem_desc(sm_k + st * SM::KV_BYTES + off), idesc_s, k == 0 ? 0u : 1u);
        }
        if (last_tile) tc_commit(bar_k_empty + 8 * st);
        tc_commit(bar_s_full + 8 * i);
      }
      __syncwarp();
    };
    for (int i = 0; i < ntile; ++i) issue_s(0, i, i == ntile - 1);
    for (int it = 0; it < nblk; ++it) {
      const int st = it % STAGES;
      for (int i = 0; i < ntile; ++i) {
        mbar_wait(bar_p_full + 8 * i, it & 1);
        if (i == 0) {
          mbar_wait(bar_v_full + 8 * st, (it / STAGES) & 1);
          // V rows at positions >= seq_len are whatever the pool holds (another sequence's page, never-written
          // memory): their P is 0, but 0 * NaN would poison O, so those rows are zeroed before the product
          const int nvalid = seq_len - (j0 + it) * BK;
          if (nvalid < BK) {
            for (int idx = lane; idx < (BK - nvalid) * 8 * HALVES; idx += 32) {
              const int row = nvalid + idx / (8 * HALVES), rem = idx % (8 * HALVES);
              asm volatile("st.shared.v4.u32 [%0], {%1,%1,%1,%1};" ::"r"(sm_v + st * SM::KV_BYTES + (rem >> 3) * SM::HALF_BYTES +
                                                                         row * 128 + (rem & 7) * 16), "r"(0u) : "memory");
            }
            fence_proxy_async();
            __syncwarp();
          }
        }
        tc_fence_after();
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < BK / 16; ++k) {
            const uint64_t dv = make_smem_desc_mn(sm_v + st * SM::KV_BYTES + k * 16 * 128, SM::HALF_BYTES, 1024);
            tc_mma_f16_ts(tb + TM_O + i * D, tb + TM_SP + i * BK + 8 * k, dv, idesc_o, (it | k) != 0 ? 1u : 0u);
          }
          if (i == ntile - 1) tc_commit(bar_v_empty + 8 * st);
          tc_commit(bar_o_full + 8 * i);
        }
        __syncwarp();
        if (it + 1 < nblk) issue_s(it + 1, i, i == ntile - 1);  // overwrites P_i(it): ordered after PV_i(it) by issue order
      }
    }
  } else {
    // ===== softmax / correction / epilogue: warps 2..5 own Q tile 0, warps 6..9 Q tile 1; thread <-> query row =====
    const int ti = (warp - 2) >> 2;
    if (ti == 0 || two) {
      const int r = (warp & 3) * 32 + lane;
      const uint32_t lane_addr = ((uint32_t)((warp & 3) * 32) << 16);
      const int nrows = min(BQ, nrows_all - ti * BQ);
      const int qabs = qabs0 + ti * BQ + r;
      // this row sees KV positions [row_lo, row_hi); rows past the tile's valid rows see nothing
      const int row_hi = r < nrows ? min(seq_len, qabs + 1) : 0;
      const int row_lo = p.sliding_window >= 0 ? max(0, qabs + 1 - p.sliding_window) : 0;
      const float sc = p.scale * kLog2e;
      const float cap_in = SOFTCAP ? p.scale / p.softcap : 0.f, cap_out = SOFTCAP ? p.softcap * kLog2e : 0.f;  // tanh(s*scale/c)*c
      const float inv_sc = SOFTCAP ? 1.f / sc : 0.f;
      (void)cap_in; (void)cap_out; (void)inv_sc;
      const uint32_t t_sp = tmem_base + lane_addr + TM_SP + ti * BK, t_o = tmem_base + lane_addr + TM_O + ti * D;
      float m_used = -INFINITY, m_true = -INFINITY, l = 0.f;
      for (int it = 0; it < nblk; ++it) {
        mbar_wait(bar_s_full + 8 * ti, it & 1);
        tc_fence_after();
        float s[BK];
        {
          // all four 32-column loads in flight, ONE wait: a wait per load exposed the TMEM latency four times
          uint32_t v0[32], v1[32], v2[32], v3[32];
          tmem_ld32_nowait(t_sp, v0);
          tmem_ld32_nowait(t_sp + 32, v1);
          tmem_ld32_nowait(t_sp + 64, v2);
          tmem_ld32_nowait(t_sp + 96, v3);
          tmem_ld_wait();
#pragma unroll
          for (int q = 0; q < 32; ++q) {
            s[q] = __uint_as_float(v0[q]);
            s[32 + q] = __uint_as_float(v1[q]);
            s[64 + q] = __uint_as_float(v2[q]);
            s[96 + q] = __uint_as_float(v3[q]);
          }
        }
        const int kv0 = (j0 + it) * BK;
        // The softmax threads are instruction-issue bound (a polynomial ex2 on the FMA pipe made the kernel 20% SLOWER),
        // so the row is touched as little as possible: the max is taken over the RAW scores (scale > 0 commutes with
        // max), the scale and the max subtraction are one packed FFMA2 per column pair, the row sum one packed FADD2.
        float m_raw = -INFINITY;
        // blocks wholly inside every row's [row_lo, row_hi) of this warp need no mask (the common case away from
        // the causal diagonal and the window edge)
        const bool need_mask = __any_sync(0xffffffffu, kv0 < row_lo || kv0 + BK > row_hi);
        if (SOFTCAP || need_mask) {
#pragma unroll
          for (int c = 0; c < BK; ++c) {
            float x = SOFTCAP ? tanh_approx(s[c] * cap_in) * cap_out * inv_sc : s[c];  // back in raw-score units
            const int kv = kv0 + c;
            x = (kv >= row_lo && kv < row_hi) ? x : -INFINITY;
            s[c] = x;
            m_raw = fmaxf(m_raw, x);
          }
        } else {
#pragma unroll
          for (int c = 0; c < BK; ++c) m_raw = fmaxf(m_raw, s[c]);
        }
        const float m_blk = m_raw * sc;  // log2 domain (-inf stays -inf: sc > 0)
        m_true = fmaxf(m_true, m_blk);
        const float m_new = (m_blk > m_used + kRescaleThreshold) ? m_blk : m_used;
        const float alpha = (m_used == -INFINITY) ? 0.f : ex2(m_used - m_new);
        const float m_eff = (m_new == -INFINITY) ? 0.f : m_new;
        const float2 sc2 = make_float2(sc, sc), nm2 = make_float2(-m_eff, -m_eff);
        float2 sum2 = make_float2(0.f, 0.f);
        uint32_t pk[BK / 2];
#pragma unroll
        for (int c = 0; c < BK; c += 2) {
          const float2 x2 = __ffma2_rn(make_float2(s[c], s[c + 1]), sc2, nm2);  // s * scale*log2e - m
          const float2 p2 = make_float2(ex2(x2.x), ex2(x2.y));
          sum2 = __fadd2_rn(sum2, p2);
          pk[c >> 1] = pack2(p2.x, p2.y, p.is_fp16);
        }
        const float sum = sum2.x + sum2.y;
        l = l * alpha + sum;
        // O_i belongs to PV_i of the previous block until that has completed (it had: S_i(it) was issued after it)
        if (it > 0) {
          mbar_wait(bar_o_full + 8 * ti, (it - 1) & 1);
          tc_fence_after();
          if (__any_sync(0xffffffffu, m_new != m_used)) {
#pragma unroll
            for (int c0 = 0; c0 < D; c0 += 32) {
              uint32_t v[32];
              tmem_ld<32>(t_o + c0, v);
#pragma unroll
              for (int q = 0; q < 32; ++q) v[q] = __float_as_uint(__uint_as_float(v[q]) * alpha);
              tmem_st32(t_o + c0, v);
            }
          }
        }
        m_used = m_new;
        tmem_st32(t_sp, pk);        // P_i over the first half of S_i's columns (every thread has read its S row)
        tmem_st32(t_sp + 32, pk + 32);
        tmem_st_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(bar_p_full + 8 * ti);
      }
      // ---- epilogue: sink fold (pagedattention_tiled.metal sinks, pagedattention.metal:1318-1328), normalise, store
      mbar_wait(bar_o_full + 8 * ti, (nblk - 1) & 1);
      tc_fence_after();
      // everything is relative to m_used; the reference's l is relative to m_true (or to max(m_true, sink))
      float m_fin = m_true;
      float sink_term = 0.f;
      if (p.sinks != nullptr) {
        const float sk = p.sinks[h] * kLog2e;
        m_fin = fmaxf(m_fin, sk);
        sink_term = ex2(sk - m_fin);
      }
      const float f = (m_used == -INFINITY) ? 0.f : ex2(m_used - m_fin);
      const float inv = 1.f / (l * f + sink_term + 1e-6f);
      const float oscale = f * inv;
      uint16_t* orow = reinterpret_cast<uint16_t*>(p.out) + (int64_t)(q_start + row0 + ti * BQ + r) * p.out_stride + (int64_t)h * D;
      const bool vec_ok = (p.out_stride % 8 == 0) && ((reinterpret_cast<uintptr_t>(p.out) & 15) == 0);
#pragma unroll
      for (int c0 = 0; c0 < D; c0 += 32) {
        uint32_t v[32];
        tmem_ld<32>(t_o + c0, v);
        if (r < nrows && row_hi > 0) {
          uint32_t o[16];
#pragma unroll
          for (int q = 0; q < 16; ++q) {
            o[q] = pack2(__uint_as_float(v[2 * q]) * oscale, __uint_as_float(v[2 * q + 1]) * oscale, p.is_fp16);
            if (p.out_f16) {  // the bf16 pair, re-stored as fp16 (exact below 65504)
              const float2 f = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&o[q]));
              o[q] = pack2(fminf(fmaxf(f.x, -65504.f), 65504.f), fminf(fmaxf(f.y, -65504.f), 65504.f), 1);
            }
          }
          if (vec_ok) {
#pragma unroll
            for (int q = 0; q < 4; ++q)
              *reinterpret_cast<uint4*>(orow + c0 + 8 * q) = make_uint4(o[4 * q], o[4 * q + 1], o[4 * q + 2], o[4 * q + 3]);
          } else {
#pragma unroll
            for (int q = 0; q < 16; ++q) *reinterpret_cast<uint32_t*>(orow + c0 + 2 * q) = o[q];
          }
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(512) : "memory");
  }
}

template <int D, bool SOFTCAP>
static int launch(const b200_paged_attn_args& a, int min_qlen, cudaStream_t st) {
  using SM = Smem<D>;
  auto kern = prefill_tc_kernel<D, SOFTCAP>;
  static int configured_dev[16] = {0};
  int dev = 0;
  B200_CUDA_OK(cudaGetDevice(&dev));
  if (dev < 16 && !configured_dev[dev]) {
    B200_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SM::TOTAL));
    configured_dev[dev] = 1;
  }
  const CUtensorMapDataType dt = a.dtype == B200_BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16;
  CUtensorMap q_map, k_map, v_map;
  int rc = make_map_2d(&q_map, dt, a.query, (uint64_t)a.num_q_heads * D, (uint64_t)a.num_q_tokens, (uint64_t)a.q_stride * 2,
                       64, BQ, CU_TENSOR_MAP_L2_PROMOTION_L2_128B);
  if (rc != B200_OK) return rc;
  const uint64_t tokens = (uint64_t)a.num_blocks * a.block_size;
  rc = make_map_2d(&k_map, dt, a.key_cache, (uint64_t)a.num_kv_heads * D, tokens, (uint64_t)a.kv_token_stride * 2, 64,
                   (uint32_t)a.block_size, CU_TENSOR_MAP_L2_PROMOTION_L2_128B);
  if (rc != B200_OK) return rc;
  rc = make_map_2d(&v_map, dt, a.value_cache, (uint64_t)a.num_kv_heads * D, tokens, (uint64_t)a.kv_token_stride * 2, 64,
                   (uint32_t)a.block_size, CU_TENSOR_MAP_L2_PROMOTION_L2_128B);
  if (rc != B200_OK) return rc;
  Params p;
  p.out = a.out;
  p.block_tables = a.block_tables;
  p.seq_lens = a.seq_lens;
  p.cu = a.cu_seqlens_q;
  p.sinks = a.sinks;
  p.num_seqs = a.num_seqs;
  p.num_q_tokens = a.num_q_tokens;
  p.num_kv_heads = a.num_kv_heads;
  p.group = a.num_q_heads / a.num_kv_heads;
  p.block_size = a.block_size;
  p.pages_per_tile = BK / a.block_size;
  p.max_blocks_per_seq = a.max_blocks_per_seq;
  p.sliding_window = a.sliding_window;
  p.min_qlen = min_qlen;
  p.out_stride = a.out_stride;
  p.scale = a.scale;
  p.softcap = a.softcap;
  p.is_fp16 = a.dtype == B200_F16 ? 1 : 0;
  p.out_f16 = out_is_f16(a);
  // upper bound on Q tiles: every sequence adds at most one partial tile (reference grid T/BQ + S)
  dim3 grid((unsigned)(a.num_q_tokens / BQ2 + a.num_seqs), (unsigned)a.num_q_heads);
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid;
  cfg.blockDim = dim3(THREADS);
  cfg.dynamicSmemBytes = SM::TOTAL;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  B200_CUDA_OK(cudaLaunchKernelEx(&cfg, kern, q_map, k_map, v_map, p));
  return B200_OK;
}

}  // namespace pf

bool prefill_tc_supported(const b200_paged_attn_args& a) {
  if (a.dtype != B200_BF16 && a.dtype != B200_F16) return false;
  if (a.head_dim != 64 && a.head_dim != 128) return false;
  if (a.block_size < 8 || a.block_size > 128 || (128 % a.block_size) != 0) return false;
  // the TMA maps view each cache as one [num_blocks * block_size tokens, H_kv * D] matrix
  if (a.kv_head_stride != a.head_dim) return false;
  if (a.kv_token_stride != (int64_t)a.num_kv_heads * a.head_dim) return false;
  if (a.kv_block_stride != (int64_t)a.block_size * a.kv_token_stride) return false;
  if (((uintptr_t)a.query % 16) || ((uintptr_t)a.key_cache % 16) || ((uintptr_t)a.value_cache % 16)) return false;
  if ((a.q_stride * 2) % 16) return false;
  if ((int64_t)a.num_blocks * a.block_size >= (1ll << 31)) return false;
  return tc::encode_fn() != nullptr;
}

int launch_prefill_tc(const b200_paged_attn_args& a, int min_qlen, cudaStream_t st) {
  if (a.softcap > 0.f) return a.head_dim == 128 ? pf::launch<128, true>(a, min_qlen, st) : pf::launch<64, true>(a, min_qlen, st);
  return a.head_dim == 128 ? pf::launch<128, false>(a, min_qlen, st) : pf::launch<64, false>(a, min_qlen, st);
}

}  // namespace b200
