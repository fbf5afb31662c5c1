// Bandwidth path of paged attention: GQA-shared, split-KV, tensor-core decode.
//
// One CTA per (kv head, query row, KV partition).  The CTA streams its slice of
// the paged K/V ONCE for the whole GQA group (the reference re-streams it per
// q head, pagedattention.metal:957-960): the H_q/H_kv query heads of the group
// are the M rows of an m16n8k16 MMA, 16-token KV tiles are the N (QK^T) / K
// (PV) dimension.  Each of the 4 warps owns every 4th tile of the partition and
// runs its own cp.async ring (16-byte coalesced row segments, zero-filled
// outside the row's [lo, hi) range, XOR-swizzled for conflict-free ldmatrix),
// keeps an online-softmax state in the log2 domain, and the warps are merged
// through shared memory at the end.  Partials go to the split-KV workspace
// (fp32) or, single pass, straight to `out` with the sink fold and the
// 1/(l + 1e-6) normalisation of pagedattention.metal:1318-1328,1383.
//
// HBM-bound: 2 * L * D * esize bytes per (row, kv head); the MMA work
// (4*L*16*D flop) is <2% of the tensor pipe at roofline.
//
// Window mode (WIN; reference pagedattention.metal:1396-1914, host paged_ops.cpp:468-476,508-515): with a
// `window_seqlen_q` hint (spec-decode verification, every segment a K+1-row window) one CTA owns
// floor(16 / G) consecutive query rows of ONE sequence -- they fill the MMA rows the GQA group leaves empty -- and
// streams the KV once for all of them.  Each MMA row keeps its own [lo, hi) range (causal staircase, window start,
// partition clip), tiles are dealt to warps by ABSOLUTE tile index and the split-KV plan is keyed on the row count,
// so every row's arithmetic is the one the expanded (one row per CTA) dispatch does: the two are bitwise equal
// (reference tests/test_spec_window_parity.py).  Tiles on which the rows' masks differ do P.V one query row at a
// time against a V fragment zeroed outside that row's range, so a non-finite V at a position only later rows may see
// cannot reach earlier rows through 0 * inf (test_windowed_skips_masked_inf_v).
#include "paged_attention.cuh"

namespace b200 {

constexpr int kDecodeWarps = 4;
constexpr int kTileTokens = 16;

template <typename T>
struct Mma;
template <>
struct Mma<__nv_bfloat16> {
  __device__ static __forceinline__ void run(float (&c)[4], const uint32_t (&a)[4], uint32_t b0,
                                             uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, "
        "{%8,%9}, {%0,%1,%2,%3};"
        : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
  }
  __device__ static __forceinline__ uint32_t pack(float lo, float hi) {
    __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&v);
  }
};
template <>
struct Mma<__half> {
  __device__ static __forceinline__ void run(float (&c)[4], const uint32_t (&a)[4], uint32_t b0,
                                             uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, "
        "{%8,%9}, {%0,%1,%2,%3};"
        : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
  }
  __device__ static __forceinline__ uint32_t pack(float lo, float hi) {
    __half2 v = __floats2half2_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&v);
  }
};

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
__device__ __forceinline__ void ldmatrix_x4_trans(uint32_t& r0, uint32_t& r1, uint32_t& r2,
                                                  uint32_t& r3, uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3)
               : "r"(addr));
}

struct DecodeParams {
  void* out;
  const void* query;
  const uint8_t* key_cache;
  const uint8_t* value_cache;
  const int32_t* block_tables;
  const int32_t* seq_lens;
  const int32_t* cu_seqlens_q;
  const float* sinks;
  int num_seqs, num_q_heads, num_kv_heads, group;
  int block_shift, max_blocks_per_seq, sliding_window;
  int64_t q_stride, out_stride;
  uint32_t block_stride16, token_stride16;  // cache strides in 16-byte units
  int64_t head_stride_bytes;
  int64_t num_blocks;
  float scale, softcap;
  SplitWorkspace ws;
  int partitioned;
  int win_rows, win_subs;  // window mode: query rows per CTA, CTAs per sequence
};

template <typename T, int D, int STAGES, bool WIN>
__global__ void __launch_bounds__(kDecodeWarps * 32, (D >= 128 ? 2 : 3))
paged_decode_mma_kernel(const DecodeParams p) {
  pdl_trigger();
  pdl_wait();
  constexpr int ROW_BYTES = D * 2;
  constexpr int CHUNKS = D / 8;               // 16-byte chunks per row
  constexpr int ROWS_PER_INSTR = 32 / CHUNKS;  // rows one warp-wide cp.async covers
  constexpr int INSTRS = kTileTokens / ROWS_PER_INSTR;
  constexpr int TILE_BYTES = kTileTokens * ROW_BYTES;
  constexpr int STAGE_BYTES = 2 * TILE_BYTES;  // K tile then V tile
  constexpr int KSTEPS = D / 16;
  static_assert(STAGE_BYTES >= 16 * D * 4, "merge buffer must fit in one stage");

  extern __shared__ __align__(128) uint8_t smem_raw[];
  __shared__ float sm_m[kDecodeWarps][16];
  __shared__ float sm_l[kDecodeWarps][16];

  const int kvh = blockIdx.x, part = blockIdx.z;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = lane >> 2, tq = lane & 3;
  const int G = p.group;

  // ---- the CTA's query rows: one (t0), or in window mode nrows consecutive rows of sequence s ----
  int s, t0, nrows, eff0;  // eff0 = KV positions row t0 may see (pagedattention.metal:913); row t0+i sees eff0+i
  if constexpr (WIN) {
    s = blockIdx.y / p.win_subs;
    const int r0 = (blockIdx.y - s * p.win_subs) * p.win_rows;
    const int q_start = p.cu_seqlens_q[s];
    const int q_len = p.cu_seqlens_q[s + 1] - q_start;
    if (r0 >= q_len) return;  // shorter segment than the hint: nothing in this sub-window
    nrows = min(p.win_rows, q_len - r0);
    t0 = q_start + r0;
    eff0 = p.seq_lens[s] - q_len + r0 + 1;
  } else {
    t0 = blockIdx.y;
    const RowInfo ri = row_info(p.cu_seqlens_q, p.seq_lens, p.num_seqs, t0, p.sliding_window, p.ws.skip_qlen);
    s = ri.seq;
    nrows = 1;
    eff0 = ri.eff;
    if (eff0 <= 0) return;  // :917-920
  }
  // KV range [lo, hi) of row t0+qi inside this CTA's partition.  0: the row does not exist (never written),
  // 1: its partition is empty (:922-926), 2: live.
  auto row_range = [&](int qi, int& lo, int& hi) -> int {
    lo = hi = 0;
    const int eff = eff0 + qi;
    if (qi >= nrows || eff <= 0) return 0;
    int beg = 0, base = 0, end = eff;
    if (p.sliding_window >= 0) {  // window: :1129-1133; partitions start at the window's first tile (RowInfo)
      beg = max(0, eff - p.sliding_window);
      base = beg & ~15;
    }
    if (p.partitioned) {
      const int pstart = base + part * p.ws.partition_size;
      if (pstart >= eff) return 1;
      beg = max(beg, pstart);
      end = min(end, pstart + p.ws.partition_size);
    }
    lo = beg;
    hi = max(end, beg);
    return 2;
  };
  // union of the rows' ranges (what is streamed) and their intersection (where no row masks anything)
  int beg, end, ulo = 0, uhi = 0x7fffffff;
  {
    int umin = 0x7fffffff, umax = 0, live = 0;
    for (int qi = 0; qi < nrows; ++qi) {
      int lo, hi;
      const int st = row_range(qi, lo, hi);
      if (st == 0) continue;
      live |= st == 2;
      if (hi > lo) {
        umin = min(umin, lo);
        umax = max(umax, hi);
      }
      ulo = max(ulo, lo);
      uhi = min(uhi, hi);
    }
    if (!live) return;
    beg = umax > umin ? umin : 0;
    end = umax > umin ? umax : 0;
  }
  const int first_tile = beg >> 4;
  const int n_tiles = end > beg ? ((end - 1) >> 4) - first_tile + 1 : 0;

  // ---- this thread's two MMA rows (g and g+8): query row, head within the group, KV range ----
  int qi0 = 0, qi1 = 0, hl0 = g, hl1 = g + 8;
  if constexpr (WIN) {
    qi0 = g / G;
    hl0 = g - qi0 * G;
    qi1 = (g + 8) / G;
    hl1 = g + 8 - qi1 * G;
  }
  const bool rv0 = WIN ? qi0 < nrows : g < G;
  const bool rv1 = WIN ? qi1 < nrows : g + 8 < G;
  int lo0 = beg, hi0 = end, lo1 = beg, hi1 = end;
  if constexpr (WIN) {
    row_range(rv0 ? qi0 : nrows, lo0, hi0);
    row_range(rv1 ? qi1 : nrows, lo1, hi1);
  }

  // ---- Q fragments (rows = heads of the GQA group [x query rows], zero beyond) ------------
  uint32_t qf[KSTEPS][4];
  {
    const T* q0 = reinterpret_cast<const T*>(p.query) + (int64_t)(t0 + qi0) * p.q_stride +
                  ((int64_t)kvh * G + hl0) * D;
    const T* q1 = reinterpret_cast<const T*>(p.query) + (int64_t)(t0 + qi1) * p.q_stride +
                  ((int64_t)kvh * G + hl1) * D;
#pragma unroll
    for (int ks = 0; ks < KSTEPS; ++ks) {
      const int c = ks * 16 + 2 * tq;
      qf[ks][0] = rv0 ? *reinterpret_cast<const uint32_t*>(q0 + c) : 0u;
      qf[ks][1] = rv1 ? *reinterpret_cast<const uint32_t*>(q1 + c) : 0u;
      qf[ks][2] = rv0 ? *reinterpret_cast<const uint32_t*>(q0 + c + 8) : 0u;
      qf[ks][3] = rv1 ? *reinterpret_cast<const uint32_t*>(q1 + c + 8) : 0u;
    }
  }

  const uint32_t warp_smem =
      (uint32_t)__cvta_generic_to_shared(smem_raw) + warp * (STAGES * STAGE_BYTES);
  const int32_t* table = p.block_tables + (int64_t)s * p.max_blocks_per_seq;
  const int block_mask = (1 << p.block_shift) - 1;
  const uint8_t* kbase = p.key_cache + (int64_t)kvh * p.head_stride_bytes;
  const uint8_t* vbase = p.value_cache + (int64_t)kvh * p.head_stride_bytes;

  auto issue_tile = [&](int j, int stage) {
    if (j < n_tiles) {
      const int tok = (first_tile + j) * kTileTokens + (lane & 15);
      int page = tok >> p.block_shift;
      page = min(page, p.max_blocks_per_seq - 1);
      int64_t blk = table[page];
      if (blk < 0 || blk >= p.num_blocks) blk = 0;
      uint32_t rowoff = (uint32_t)blk * p.block_stride16 +
                        (uint32_t)(tok & block_mask) * p.token_stride16;
      if (tok < beg || tok >= end) rowoff = 0xFFFFFFFFu;  // zero-fill rows outside the range
      const uint32_t sbase = warp_smem + stage * STAGE_BYTES;
      const int c = lane % CHUNKS;
#pragma unroll
      for (int i = 0; i < INSTRS; ++i) {
        const int r = i * ROWS_PER_INSTR + lane / CHUNKS;
        const uint32_t ro = __shfl_sync(0xffffffffu, rowoff, r);
        const bool ok = ro != 0xFFFFFFFFu;
        const int64_t goff = (int64_t)(ok ? ro : 0u) * 16 + c * 16;
        const uint32_t dst = sbase + r * ROW_BYTES + ((c ^ (r & 7)) << 4);
        cp_async_16(dst, kbase + goff, ok ? 16 : 0);
        cp_async_16(dst + TILE_BYTES, vbase + goff, ok ? 16 : 0);
      }
    }
    cp_async_commit();
  };

  float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f;
  float o[D / 8][4];
#pragma unroll
  for (int i = 0; i < D / 8; ++i) o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f;

  const float sc_mul = p.softcap > 0.f ? p.scale : p.scale * kLog2e;

  // tiles are dealt to warps by absolute index (tile & 3 == warp), so a row's tile -> warp map does not depend on
  // where the CTA's range starts
  const int j0 = (warp - first_tile) & (kDecodeWarps - 1);
#pragma unroll
  for (int st = 0; st < STAGES - 1; ++st) issue_tile(j0 + st * kDecodeWarps, st);

  int stage = 0;
  for (int j = j0; j < n_tiles; j += kDecodeWarps) {
    cp_async_wait<STAGES - 2>();
    __syncwarp();
    {
      int nst = stage + STAGES - 1;
      if (nst >= STAGES) nst -= STAGES;
      issue_tile(j + (STAGES - 1) * kDecodeWarps, nst);
    }
    const uint32_t kt = warp_smem + stage * STAGE_BYTES;
    const uint32_t vt = kt + TILE_BYTES;

    // ---- S = Q K^T : 16 (heads) x 16 (tokens) ---------------------------------
    float sacc[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
    {
      const int mj = lane >> 3, r = lane & 7;
      const int trow = (mj >> 1) * 8 + r;
#pragma unroll
      for (int ks = 0; ks < KSTEPS; ++ks) {
        const int chunk = ks * 2 + (mj & 1);
        uint32_t b0, b1, b2, b3;
        ldmatrix_x4(b0, b1, b2, b3, kt + trow * ROW_BYTES + ((chunk ^ r) << 4));
        Mma<T>::run(sacc[0], qf[ks], b0, b1);
        Mma<T>::run(sacc[1], qf[ks], b2, b3);
      }
    }
    // ---- scale, softcap, log2 domain, range mask -------------------------------
    const int tok0 = (first_tile + j) * kTileTokens + 2 * tq;
#pragma unroll
    for (int nt = 0; nt < 2; ++nt) {
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        float v = sacc[nt][e] * sc_mul;
        if (p.softcap > 0.f) v = tanhf(v / p.softcap) * p.softcap * kLog2e;  // :1120-1122,1138
        const int tok = tok0 + nt * 8 + (e & 1);
        const bool in = e < 2 ? (tok >= lo0 && tok < hi0) : (tok >= lo1 && tok < hi1);
        sacc[nt][e] = in ? v : -INFINITY;
      }
    }
    float mx0 = fmaxf(fmaxf(sacc[0][0], sacc[0][1]), fmaxf(sacc[1][0], sacc[1][1]));
    float mx1 = fmaxf(fmaxf(sacc[0][2], sacc[0][3]), fmaxf(sacc[1][2], sacc[1][3]));
    mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1));
    mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
    mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1));
    mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
    const float mn0 = fmaxf(m0, mx0), mn1 = fmaxf(m1, mx1);
    const float mu0 = mn0 == -INFINITY ? 0.f : mn0;  // all masked so far: clamp (:1169-1176)
    const float mu1 = mn1 == -INFINITY ? 0.f : mn1;
    const float c0 = exp2f(m0 - mu0), c1 = exp2f(m1 - mu1);
    m0 = mn0;
    m1 = mn1;
    float ps0 = 0.f, ps1 = 0.f;
#pragma unroll
    for (int nt = 0; nt < 2; ++nt) {
      sacc[nt][0] = exp2f(sacc[nt][0] - mu0);
      sacc[nt][1] = exp2f(sacc[nt][1] - mu0);
      sacc[nt][2] = exp2f(sacc[nt][2] - mu1);
      sacc[nt][3] = exp2f(sacc[nt][3] - mu1);
      ps0 += sacc[nt][0] + sacc[nt][1];
      ps1 += sacc[nt][2] + sacc[nt][3];
    }
    l0 = l0 * c0 + ps0;
    l1 = l1 * c1 + ps1;
#pragma unroll
    for (int i = 0; i < D / 8; ++i) {
      o[i][0] *= c0;
      o[i][1] *= c0;
      o[i][2] *= c1;
      o[i][3] *= c1;
    }
    // ---- O += P V ---------------------------------------------------------------
    uint32_t pa[4];
    pa[0] = Mma<T>::pack(sacc[0][0], sacc[0][1]);
    pa[1] = Mma<T>::pack(sacc[0][2], sacc[0][3]);
    pa[2] = Mma<T>::pack(sacc[1][0], sacc[1][1]);
    pa[3] = Mma<T>::pack(sacc[1][2], sacc[1][3]);
    const int tile_beg = (first_tile + j) * kTileTokens;
    // every streamed token of this tile is inside every row's range?  (tokens outside [beg, end) are zero-filled)
    const bool uniform = !WIN || (max(tile_beg, beg) >= ulo && min(tile_beg + kTileTokens, end) <= uhi);
    if (uniform) {
      const int mj = lane >> 3, r = lane & 7;
      const int trow = (mj & 1) * 8 + r;
#pragma unroll
      for (int n2 = 0; n2 < KSTEPS; ++n2) {
        const int chunk = n2 * 2 + (mj >> 1);
        uint32_t v0, v1, v2, v3;
        ldmatrix_x4_trans(v0, v1, v2, v3, vt + trow * ROW_BYTES + ((chunk ^ r) << 4));
        Mma<T>::run(o[2 * n2], pa, v0, v1);
        Mma<T>::run(o[2 * n2 + 1], pa, v2, v3);
      }
    } else {
      // the rows' masks differ on this tile: one query row at a time, V zeroed outside that row's range -- what the
      // expanded dispatch's zero-filled load gives that row; the other rows keep their accumulators
      const int mj = lane >> 3, r = lane & 7;
      const int trow = (mj & 1) * 8 + r;
      const int tk = tile_beg + 2 * tq;  // this lane's B-fragment tokens: tk, tk+1 (b0) and tk+8, tk+9 (b1)
      for (int qi = 0; qi < nrows; ++qi) {
        int lo, hi;
        row_range(qi, lo, hi);
        const uint32_t m01 = ((tk >= lo && tk < hi) ? 0x0000FFFFu : 0u) | ((tk + 1 >= lo && tk + 1 < hi) ? 0xFFFF0000u : 0u);
        const uint32_t m89 = ((tk + 8 >= lo && tk + 8 < hi) ? 0x0000FFFFu : 0u) | ((tk + 9 >= lo && tk + 9 < hi) ? 0xFFFF0000u : 0u);
        // only query row qi's MMA rows take this pass's result: a zero P row times a non-finite V is NaN, not 0
        const bool t0_ = rv0 && qi0 == qi, t1_ = rv1 && qi1 == qi;
        uint32_t pq[4];
        pq[0] = t0_ ? pa[0] : 0u;
        pq[1] = t1_ ? pa[1] : 0u;
        pq[2] = t0_ ? pa[2] : 0u;
        pq[3] = t1_ ? pa[3] : 0u;
#pragma unroll
        for (int n2 = 0; n2 < KSTEPS; ++n2) {
          const int chunk = n2 * 2 + (mj >> 1);
          uint32_t v0, v1, v2, v3;
          ldmatrix_x4_trans(v0, v1, v2, v3, vt + trow * ROW_BYTES + ((chunk ^ r) << 4));
#pragma unroll
          for (int h = 0; h < 2; ++h) {
            float (&oo)[4] = o[2 * n2 + h];
            float d[4] = {oo[0], oo[1], oo[2], oo[3]};
            Mma<T>::run(d, pq, (h ? v2 : v0) & m01, (h ? v3 : v1) & m89);
            if (t0_) {
              oo[0] = d[0];
              oo[1] = d[1];
            }
            if (t1_) {
              oo[2] = d[2];
              oo[3] = d[3];
            }
          }
        }
      }
    }
    if (++stage == STAGES) stage = 0;
  }
  cp_async_wait<0>();
  __syncwarp();

  // ---- per-warp state -> shared memory, then merge the 4 warps -------------------
  l0 += __shfl_xor_sync(0xffffffffu, l0, 1);
  l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
  l1 += __shfl_xor_sync(0xffffffffu, l1, 1);
  l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
  float* so = reinterpret_cast<float*>(smem_raw + warp * (STAGES * STAGE_BYTES));
  if (tq == 0) {
    sm_m[warp][g] = m0;
    sm_l[warp][g] = l0;
    sm_m[warp][g + 8] = m1;
    sm_l[warp][g + 8] = l1;
  }
#pragma unroll
  for (int i = 0; i < D / 8; ++i) {
    if (rv0) *reinterpret_cast<float2*>(so + g * D + i * 8 + 2 * tq) = make_float2(o[i][0], o[i][1]);
    if (rv1)
      *reinterpret_cast<float2*>(so + (g + 8) * D + i * 8 + 2 * tq) = make_float2(o[i][2], o[i][3]);
  }
  __syncthreads();

  for (int idx = threadIdx.x; idx < nrows * G * D; idx += blockDim.x) {
    const int row = idx / D, d = idx - row * D;  // MMA row = query row * G + head within the group
    const int qi = WIN ? row / G : 0;
    const int t = t0 + qi;
    if (WIN && eff0 + qi <= 0) continue;  // :917-920
    float M = -INFINITY;
#pragma unroll
    for (int w = 0; w < kDecodeWarps; ++w) M = fmaxf(M, sm_m[w][row]);
    float L = 0.f, acc = 0.f;
#pragma unroll
    for (int w = 0; w < kDecodeWarps; ++w) {
      const float mw = sm_m[w][row];
      if (mw != -INFINITY) {  // skip warps that saw nothing (:1288-1289)
        const float wgt = exp2f(mw - M);
        L += sm_l[w][row] * wgt;
        const float* sw =
            reinterpret_cast<const float*>(smem_raw + w * (STAGES * STAGE_BYTES));
        acc += sw[row * D + d] * wgt;
      }
    }
    const int h = kvh * G + (row - qi * G);
    if (p.partitioned) {
      const int64_t prow = ((int64_t)t * p.num_q_heads + h) * p.ws.num_partitions + part;
      p.ws.tmp_out[prow * D + d] = acc;
      if (d == 0) {
        p.ws.max_l2[prow] = M;
        p.ws.exp_sum[prow] = L;
      }
    } else {
      if (M == -INFINITY) M = 0.f;
      float c = 1.f;
      if (p.sinks != nullptr) {  // :1318-1328
        const float sk = p.sinks[h] * kLog2e;
        const float M2 = fmaxf(M, sk);
        c = exp2f(M - M2);
        L = L * c + exp2f(sk - M2);
      }
      T* op = reinterpret_cast<T*>(p.out) + (int64_t)t * p.out_stride + (int64_t)h * D;
      store_out<T>(op, d, acc * c / (L + 1e-6f), p.ws.out_f16);  // :1383
    }
  }
}

static int log2_exact_i(int v) {
  int s = 0;
  while ((1 << s) < v) ++s;
  return (1 << s) == v ? s : -1;
}

bool decode_mma_supported(const b200_paged_attn_args& a) {
  if (a.dtype != B200_BF16 && a.dtype != B200_F16) return false;
  if (a.head_dim != 64 && a.head_dim != 128) return false;
  const int G = a.num_q_heads / a.num_kv_heads;
  if (G < 1 || G > 16) return false;
  if (log2_exact_i(a.block_size) < 0) return false;
  if (a.kv_block_stride % 8 || a.kv_token_stride % 8 || a.kv_head_stride % 8) return false;
  if (a.q_stride % 2) return false;
  if (((uintptr_t)a.key_cache | (uintptr_t)a.value_cache) % 16) return false;
  if ((uintptr_t)a.query % 4) return false;
  // row offsets are carried as 32-bit counts of 16-byte units
  if ((a.num_blocks * a.kv_block_stride) / 8 >= 0xFFFFFFF0ll) return false;
  return true;
}

template <typename T, int D, int STAGES, bool WIN>
static int launch_one_mode(const DecodeParams& p, dim3 grid, cudaStream_t st) {
  constexpr int smem = kDecodeWarps * STAGES * 2 * kTileTokens * D * 2;
  static bool configured_dev[16] = {false};  // cudaFuncSetAttribute is per device
  int dev_ = 0;
  B200_CUDA_OK(cudaGetDevice(&dev_));
  bool& configured = configured_dev[dev_ & 15];
  if (!configured) {
    B200_CUDA_OK(cudaFuncSetAttribute(paged_decode_mma_kernel<T, D, STAGES, WIN>,
                                      cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    configured = true;
  }
  launch_pdl(paged_decode_mma_kernel<T, D, STAGES, WIN>, dim3(grid), dim3(kDecodeWarps * 32), smem, st, p);
  return check_launch("paged_decode_mma_kernel");
}

template <typename T, int D, int STAGES>
static int launch_one(const DecodeParams& p, dim3 grid, cudaStream_t st) {
  if (p.win_rows > 1) return launch_one_mode<T, D, STAGES, true>(p, grid, st);
  return launch_one_mode<T, D, STAGES, false>(p, grid, st);
}

// Window mode (a `window_seqlen_q` hint > 1): query rows of one sequence that share a CTA -- the MMA rows the GQA
// group leaves free.  1 = per-row dispatch.
int decode_window_rows(const b200_paged_attn_args& a) {
  if (a.window_seqlen_q <= 1) return 1;
  const int G = a.num_q_heads / a.num_kv_heads;
  const int r = G > 0 ? 16 / G : 1;
  return r < 1 ? 1 : (r > a.window_seqlen_q ? a.window_seqlen_q : r);
}

int launch_decode_mma(const b200_paged_attn_args& a, const SplitWorkspace* ws, cudaStream_t st, int skip_qlen) {
  DecodeParams p{};
  p.out = a.out;
  p.query = a.query;
  p.key_cache = (const uint8_t*)a.key_cache;
  p.value_cache = (const uint8_t*)a.value_cache;
  p.block_tables = a.block_tables;
  p.seq_lens = a.seq_lens;
  p.cu_seqlens_q = a.cu_seqlens_q;
  p.sinks = a.sinks;
  p.num_seqs = a.num_seqs;
  p.num_q_heads = a.num_q_heads;
  p.num_kv_heads = a.num_kv_heads;
  p.group = a.num_q_heads / a.num_kv_heads;
  p.block_shift = log2_exact_i(a.block_size);
  p.max_blocks_per_seq = a.max_blocks_per_seq;
  p.sliding_window = a.sliding_window;
  p.q_stride = a.q_stride;
  p.out_stride = a.out_stride;
  p.block_stride16 = (uint32_t)(a.kv_block_stride / 8);
  p.token_stride16 = (uint32_t)(a.kv_token_stride / 8);
  p.head_stride_bytes = a.kv_head_stride * 2;
  p.num_blocks = a.num_blocks;
  p.scale = a.scale;
  p.softcap = a.softcap;
  p.partitioned = ws != nullptr;
  p.ws = SplitWorkspace{};
  if (ws) p.ws = *ws;
  p.ws.skip_qlen = skip_qlen;
  p.ws.out_f16 = out_is_f16(a);
  dim3 grid(a.num_kv_heads, a.num_q_tokens, ws ? ws->num_partitions : 1);
  p.win_rows = skip_qlen > 0 ? 1 : decode_window_rows(a);
  p.win_subs = 1;
  if (p.win_rows > 1) {  // one CTA per (sequence, sub-window of win_rows rows); paged_ops.cpp:508-515
    p.win_subs = (a.window_seqlen_q + p.win_rows - 1) / p.win_rows;
    grid.y = (unsigned)(a.num_seqs * p.win_subs);
  }
  if (a.dtype == B200_BF16) {
    if (a.head_dim == 128) return launch_one<__nv_bfloat16, 128, 3>(p, grid, st);
    return launch_one<__nv_bfloat16, 64, 4>(p, grid, st);
  }
  if (a.head_dim == 128) return launch_one<__half, 128, 3>(p, grid, st);
  return launch_one<__half, 64, 4>(p, grid, st);
}

}  // namespace b200
