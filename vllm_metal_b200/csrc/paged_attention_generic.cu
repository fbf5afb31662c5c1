// Generic per-(token, q head) paged attention and the split-KV reduce.
//
// This is the any-shape CUDA path (fp32/fp16/bf16, any head_dim <= 512, any
// power-of-two page size, varlen rows): one CTA per (q head, query row
// [, partition]) as in the reference's per-token kernel
// (kernels_v2/pagedattention.metal:817-1395), warp w walking tokens w, w+4, ...
// with a per-warp online softmax in the log2 domain and a 4-warp merge.  The
// bandwidth path for decode is paged_attention_decode.cu; this file is what
// every other shape (and fp32) runs on, and the kernel the tensor-core paths
// are checked against on the device.
#include "paged_attention.cuh"

namespace b200 {

constexpr int kGenericWarps = 4;
constexpr int kMaxEpl = 16;  // elements per lane: head_dim <= 512

template <typename T>
__global__ void __launch_bounds__(kGenericWarps * 32)
paged_attention_generic_kernel(T* __restrict__ out, const T* __restrict__ query,
                               const T* __restrict__ key_cache, const T* __restrict__ value_cache,
                               const int32_t* __restrict__ block_tables,
                               const int32_t* __restrict__ seq_lens,
                               const int32_t* __restrict__ cu_seqlens_q,
                               const float* __restrict__ sinks, int num_seqs, int num_q_heads,
                               int num_kv_heads, int head_dim, int block_shift,
                               int max_blocks_per_seq, int sliding_window, int64_t q_stride,
                               int64_t out_stride, int64_t kv_block_stride,
                               int64_t kv_token_stride, int64_t kv_head_stride, int64_t num_blocks,
                               float scale, float softcap, SplitWorkspace ws, bool partitioned) {
  pdl_trigger();
  pdl_wait();
  const int h = blockIdx.x;
  const int t = blockIdx.y;
  const int part = blockIdx.z;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  const RowInfo ri = row_info(cu_seqlens_q, seq_lens, num_seqs, t, sliding_window, ws.skip_qlen);
  const int s = ri.seq, eff = ri.eff;  // eff: pagedattention.metal:913
  if (eff <= 0) return;                // :917-920 (row left untouched)

  int lo = ri.beg, hi = eff;  // window: :1129-1133
  if (partitioned) {
    const int p0 = ri.base + part * ws.partition_size;
    if (p0 >= eff) return;  // :922-926 empty partition: nothing written, reduce skips it
    lo = max(lo, p0);
    hi = min(hi, p0 + ws.partition_size);
  }

  const int kvh = h / (num_q_heads / num_kv_heads);  // :959-960
  const int block_mask = (1 << block_shift) - 1;
  const int32_t* table = block_tables + (int64_t)s * max_blocks_per_seq;

  float q[kMaxEpl];
  const T* qp = query + (int64_t)t * q_stride + (int64_t)h * head_dim;
#pragma unroll
  for (int i = 0; i < kMaxEpl; ++i) {
    const int d = lane + 32 * i;
    q[i] = d < head_dim ? to_float(qp[d]) * scale : 0.f;
  }

  float m = -INFINITY, l = 0.f;
  float o[kMaxEpl];
#pragma unroll
  for (int i = 0; i < kMaxEpl; ++i) o[i] = 0.f;

  for (int tok = lo + warp; tok < hi; tok += kGenericWarps) {
    int64_t blk = table[tok >> block_shift];
    if (blk < 0 || blk >= num_blocks) blk = 0;  // padded / corrupt entry: stay in bounds
    const int64_t base = blk * kv_block_stride + (int64_t)(tok & block_mask) * kv_token_stride +
                         (int64_t)kvh * kv_head_stride;
    const T* kp = key_cache + base;
    const T* vp = value_cache + base;
    float dot = 0.f;
#pragma unroll
    for (int i = 0; i < kMaxEpl; ++i) {
      const int d = lane + 32 * i;
      if (d < head_dim) dot = fmaf(q[i], to_float(kp[d]), dot);
    }
    dot = warp_sum(dot);
    if (softcap > 0.f) dot = tanhf(dot / softcap) * softcap;  // :1120-1122
    const float sc = dot * kLog2e;                            // :1138
    const float m_new = fmaxf(m, sc);
    const float corr = exp2f(m - m_new);  // m = -inf first time -> 0
    const float p = exp2f(sc - m_new);
    l = l * corr + p;
#pragma unroll
    for (int i = 0; i < kMaxEpl; ++i) {
      const int d = lane + 32 * i;
      if (d < head_dim) o[i] = fmaf(p, to_float(vp[d]), o[i] * corr);
    }
    m = m_new;
  }

  // ---- merge the warps (pagedattention.metal:1256-1307) ---------------------
  extern __shared__ float smem[];
  float* sm_m = smem;                      // [warps]
  float* sm_l = smem + kGenericWarps;      // [warps]
  float* sm_o = smem + 2 * kGenericWarps;  // [warps][head_dim]
  if (lane == 0) {
    sm_m[warp] = m;
    sm_l[warp] = l;
  }
#pragma unroll
  for (int i = 0; i < kMaxEpl; ++i) {
    const int d = lane + 32 * i;
    if (d < head_dim) sm_o[warp * head_dim + d] = o[i];
  }
  __syncthreads();

  float M = -INFINITY;
#pragma unroll
  for (int w = 0; w < kGenericWarps; ++w) M = fmaxf(M, sm_m[w]);
  float L = 0.f, wgt[kGenericWarps];
#pragma unroll
  for (int w = 0; w < kGenericWarps; ++w) {
    wgt[w] = (sm_m[w] == -INFINITY) ? 0.f : exp2f(sm_m[w] - M);  // :1288-1289 skip empty warps
    L += sm_l[w] * wgt[w];
  }

  if (partitioned) {
    const int64_t row = ((int64_t)t * num_q_heads + h) * ws.num_partitions + part;
    if (threadIdx.x == 0) {
      ws.max_l2[row] = M;  // -inf when the whole partition is window-masked
      ws.exp_sum[row] = L;
    }
    for (int d = threadIdx.x; d < head_dim; d += blockDim.x) {
      float acc = 0.f;
#pragma unroll
      for (int w = 0; w < kGenericWarps; ++w) acc += sm_o[w * head_dim + d] * wgt[w];
      ws.tmp_out[row * head_dim + d] = acc;
    }
    return;
  }

  if (M == -INFINITY) M = 0.f;  // everything masked (W == 0): clamp as :1169-1176
  float c = 1.f;
  if (sinks != nullptr) {  // :1318-1328 fold the sink once, after the merge
    const float sk = sinks[h] * kLog2e;
    const float M2 = fmaxf(M, sk);
    c = exp2f(M - M2);
    L = L * c + exp2f(sk - M2);
  }
  const float inv = c / (L + 1e-6f);  // :1383
  T* op = out + (int64_t)t * out_stride + (int64_t)h * head_dim;
  for (int d = threadIdx.x; d < head_dim; d += blockDim.x) {
    float acc = 0.f;
#pragma unroll
    for (int w = 0; w < kGenericWarps; ++w) acc += sm_o[w * head_dim + d] * wgt[w];
    store_out<T>(op, d, acc * inv, ws.out_f16);
  }
}

// LSE merge of the split-KV partials (reference paged_attention_v2_reduce,
// pagedattention.metal:1917-2144): one CTA per (q head, row).
template <typename T>
__global__ void __launch_bounds__(128)
split_reduce_kernel(T* __restrict__ out, const int32_t* __restrict__ seq_lens,
                    const int32_t* __restrict__ cu_seqlens_q, const float* __restrict__ sinks,
                    int num_seqs, int num_q_heads, int head_dim, int sliding_window,
                    int64_t out_stride, SplitWorkspace ws) {
  pdl_trigger();
  pdl_wait();
  const int h = blockIdx.x, t = blockIdx.y;
  const RowInfo ri = row_info(cu_seqlens_q, seq_lens, num_seqs, t, sliding_window, ws.skip_qlen);
  const int eff = ri.eff;
  if (eff <= 0) return;
  int np = (eff - ri.base + ws.partition_size - 1) / ws.partition_size;  // :1946
  np = min(np, ws.num_partitions);
  const int64_t row = ((int64_t)t * num_q_heads + h) * ws.num_partitions;

  float M = -INFINITY;
  for (int j = 0; j < np; ++j) M = fmaxf(M, ws.max_l2[row + j]);
  float sk = 0.f;
  if (sinks != nullptr) {  // :2045-2072
    sk = sinks[h] * kLog2e;
    M = fmaxf(M, sk);
  }
  if (M == -INFINITY) M = 0.f;
  float L = 0.f;
  for (int j = 0; j < np; ++j) {
    const float mj = ws.max_l2[row + j];
    if (mj != -INFINITY) L += ws.exp_sum[row + j] * exp2f(mj - M);
  }
  if (sinks != nullptr) L += exp2f(sk - M);
  const float inv = 1.f / (L + 1e-6f);  // :2074
  T* op = out + (int64_t)t * out_stride + (int64_t)h * head_dim;
  for (int d = threadIdx.x; d < head_dim; d += blockDim.x) {
    float acc = 0.f;
    for (int j = 0; j < np; ++j) {
      const float mj = ws.max_l2[row + j];
      if (mj != -INFINITY) acc += ws.tmp_out[(row + j) * head_dim + d] * exp2f(mj - M);
    }
    store_out<T>(op, d, acc * inv, ws.out_f16);
  }
}

static int log2_exact(int v) {
  int s = 0;
  while ((1 << s) < v) ++s;
  return (1 << s) == v ? s : -1;
}

int launch_generic(const b200_paged_attn_args& a, const SplitWorkspace* ws, cudaStream_t st, int skip_qlen) {
  const int shift = log2_exact(a.block_size);
  const size_t smem = (2 * kGenericWarps + (size_t)kGenericWarps * a.head_dim) * sizeof(float);
  SplitWorkspace w{};
  if (ws) w = *ws;
  w.skip_qlen = skip_qlen;
  w.out_f16 = out_is_f16(a);
  dim3 grid(a.num_q_heads, a.num_q_tokens, ws ? ws->num_partitions : 1);
  return B200_DISPATCH_DTYPE(a.dtype, T, [&]() -> int {
    launch_pdl(paged_attention_generic_kernel<T>, dim3(grid), dim3(kGenericWarps * 32), smem, st, 
        (T*)a.out, (const T*)a.query, (const T*)a.key_cache, (const T*)a.value_cache,
        a.block_tables, a.seq_lens, a.cu_seqlens_q, a.sinks, a.num_seqs, a.num_q_heads,
        a.num_kv_heads, a.head_dim, shift, a.max_blocks_per_seq, a.sliding_window, a.q_stride,
        a.out_stride, a.kv_block_stride, a.kv_token_stride, a.kv_head_stride, a.num_blocks,
        a.scale, a.softcap, w, ws != nullptr);
    return check_launch("paged_attention_generic_kernel");
  });
}

int launch_split_reduce(const b200_paged_attn_args& a, const SplitWorkspace& ws, cudaStream_t st) {
  dim3 grid(a.num_q_heads, a.num_q_tokens);
  return B200_DISPATCH_DTYPE(a.dtype, T, [&]() -> int {
    launch_pdl(split_reduce_kernel<T>, dim3(grid), dim3(128), 0, st, (T*)a.out, a.seq_lens, a.cu_seqlens_q, a.sinks,
                                                 a.num_seqs, a.num_q_heads, a.head_dim,
                                                 a.sliding_window, a.out_stride, ws);
    return check_launch("split_reduce_kernel");
  });
}

// ---- validation + routing (reference: paged_attention_primitive_fn,
// paged_ops.cpp:798-921, and dispatch_paged_attention_v2_online, :436-716) -----
static int validate(const b200_paged_attn_args* a) {
  B200_CHECK_ARG(a != nullptr, "paged_attention: args is NULL");
  B200_CHECK_ARG(a->num_q_tokens >= 0 && a->num_seqs >= 0, "paged_attention: negative sizes");
  if (a->num_q_tokens == 0 || a->num_seqs == 0) return B200_OK;
  B200_CHECK_ARG(a->out && a->query && a->key_cache && a->value_cache && a->block_tables &&
                     a->seq_lens && a->cu_seqlens_q,
                 "paged_attention: null pointer argument");
  B200_CHECK_ARG(a->num_q_heads > 0 && a->num_kv_heads > 0 &&
                     a->num_q_heads % a->num_kv_heads == 0,
                 "paged_attention: num_q_heads (%d) must be a positive multiple of num_kv_heads (%d)",
                 a->num_q_heads, a->num_kv_heads);
  B200_CHECK_ARG(a->head_dim > 0 && a->head_dim <= 32 * kMaxEpl,
                 "paged_attention: head_dim %d outside (0, %d]", a->head_dim, 32 * kMaxEpl);
  B200_CHECK_ARG(a->block_size > 0 && log2_exact(a->block_size) >= 0,
                 "paged_attention: block_size must be a power of two, got %d", a->block_size);
  B200_CHECK_ARG(a->max_blocks_per_seq > 0, "paged_attention: max_blocks_per_seq must be > 0");
  B200_CHECK_ARG(a->dtype == B200_F32 || a->dtype == B200_F16 || a->dtype == B200_BF16,
                 "paged_attention: unsupported dtype code %d", a->dtype);
  B200_CHECK_ARG(a->num_blocks > 0, "paged_attention: num_blocks must be > 0");
  B200_CHECK_ARG(a->window_seqlen_q >= 0, "paged_attention: window_seqlen_q must be >= 1 (1 = per-token decode), got %d",
                 a->window_seqlen_q);
  B200_CHECK_ARG(a->out_dtype == 0 || a->out_dtype == a->dtype || (a->out_dtype == B200_F16 && a->dtype == B200_BF16),
                 "paged_attention: out_dtype must be 0 / the query dtype, or fp16 for a bf16 call");
  return B200_OK;
}

// Longest KV span any row can cover: a windowed row only walks [eff-W, eff)
// rounded out to 16-token tiles (see RowInfo).
static int effective_span(const b200_paged_attn_args& a) {
  int L = a.max_seq_len;
  if (a.sliding_window >= 0 && a.sliding_window + 15 < L) L = a.sliding_window + 15;
  return L < 1 ? 1 : L;
}

// Returns the partition length in tokens (a multiple of b200_partition_size())
// and the partition count through *parts; parts == 1 means single pass.
static int plan_partitions(const b200_paged_attn_args& a, bool mma, int* parts) {
  const int unit = b200_partition_size();
  const int span = effective_span(a);
  const int max_parts = (span + unit - 1) / unit;  // paged_ops.cpp:520-521
  *parts = 1;
  if (a.force_partitions == 1 || max_parts < 2) return unit;
  if (a.force_partitions > 1) {
    *parts = max_parts;
    return unit;
  }
  if (mma) {
    // Bandwidth kernel: one CTA per (kv head, row, partition), 2 resident CTAs per SM.
    const int64_t items = (int64_t)a.num_kv_heads * a.num_q_tokens;
    const int sms = sm_count() > 0 ? sm_count() : 148;
    {
      // The whole launch is a handful of waves of (2 x SMs) resident CTAs, so pick the partition count that
      // minimises waves x (partition length + pipeline fill/drain) rather than a fixed CTA target.  Measured
      // (profiles/r02_decode_attn_plan.txt): 512 items at ctx 4096 (the headline shape) run 12% faster single pass
      // (2 waves, no reduce kernel) than as 3 partitions; 64 items (TP=8) want 4 partitions of 1040 tokens = ONE
      // wave of 256 CTAs, where 9 partitions of 512 were 1.95 waves of 576.  Partitions are multiples of 64 tokens.
      const int64_t slots = (a.head_dim >= 128 ? 2LL : 3LL) * sms;  // paged_decode_mma_kernel's launch bounds
      const int fill = 192;     // tokens' worth of prologue / drain / merge per CTA
      const int reduce = 448;   // the separate LSE-merge launch (~8 us), in the same token units
      int64_t best_cost = -1;
      int best_p = 1, best_size = ((span + 63) / 64) * 64;
      const int pmax = max_parts;  // what b200_paged_attention_workspace_bytes() sizes the scratch for
      for (int p = 1; p <= pmax; ++p) {
        const int psz = (((span + p - 1) / p + 63) / 64) * 64;
        const int np = (span + psz - 1) / psz;
        const int64_t waves = (items * np + slots - 1) / slots;
        const int64_t cost = waves * (psz + fill) + (np > 1 ? reduce : 0);
        if (np > pmax) continue;
        if (best_cost < 0 || cost < best_cost) {
          best_cost = cost;
          best_p = np;
          best_size = psz;
        }
      }
      *parts = best_p;
      return best_size;
    }
  }
  // Reference gate (paged_ops.cpp:539-541): pure decode && grid below occupancy bound.
  // Window batches keep split-KV eligibility on the split-equivalent grid (one row per token, :528-541), so the
  // windowed and the expanded dispatch of a verification step take the same decision.
  const bool pure_decode = a.num_q_tokens == a.num_seqs || a.window_seqlen_q > 1;
  const int gate = b200_min_decode_grid();
  if (pure_decode && gate > 0 && (int64_t)a.num_q_heads * a.num_q_tokens < gate)
    *parts = max_parts;
  return unit;
}

static bool pick_mma(const b200_paged_attn_args& a) {
  if (a.kernel == 1) return false;
  return decode_mma_supported(a);
}

}  // namespace b200

namespace b200 {
// What one b200_paged_attention call does for these arguments; shared by the scratch-size query, the launch count
// query and the launcher so they cannot disagree.
struct AttnPlan {
  bool tile;       // tcgen05 tile kernel over the sequences with q_len >= skip_qlen
  bool rows;       // per-row kernels (everything when !tile, the short sequences of a mixed batch otherwise)
  bool mma;        // per-row kernel = tensor-core decode kernel (else the generic scalar kernel)
  int skip_qlen;
  int parts, psize;  // split-KV plan of the per-row kernels
};
constexpr int kTileMinQlen = 16;

static AttnPlan make_plan(const b200_paged_attn_args& a) {
  AttnPlan pl{};
  // Varlen prefill / mixed batches (reference paged_ops.cpp:463-497 sends them to the tiled kernel): sequences
  // with at least kTileMinQlen query rows run on the tcgen05 tile kernel, the remaining (decode) rows of a mixed
  // batch on the per-row kernels, which leave the tile kernel's sequences untouched.
  const bool tile_ok = prefill_tc_supported(a);
  // A spec-verify batch (window hint) never takes the tile kernel (paged_ops.cpp:468-476): K+1-row windows waste a
  // 128-row tile and lose the split-KV parallelism; the decode kernel runs it in window mode.
  const bool window_batch = a.window_seqlen_q > 1;
  pl.tile = a.kernel == 3 ||
            (a.kernel == 0 && tile_ok && !window_batch && a.num_q_tokens - a.num_seqs + 1 >= kTileMinQlen);
  // kernel == 3 pins every sequence to the tile kernel (parity tests); auto keeps short ones on the row kernels
  pl.skip_qlen = pl.tile ? (a.kernel == 3 ? 1 : kTileMinQlen) : 0;
  pl.rows = a.kernel != 3;
  pl.mma = pick_mma(a);
  pl.parts = 1;
  pl.psize = plan_partitions(a, pl.mma, &pl.parts);
  return pl;
}
}  // namespace b200

extern "C" size_t b200_paged_attention_workspace_bytes(const b200_paged_attn_args* a) {
  using namespace b200;
  if (!a || a->num_q_tokens <= 0 || a->num_seqs <= 0) return 0;
  const AttnPlan pl = make_plan(*a);
  if (!pl.rows || pl.parts < 2) return 0;  // single pass: no scratch
  return split_workspace_bytes(a->num_q_tokens, a->num_q_heads, pl.parts, a->head_dim);
}

extern "C" int b200_paged_attention_num_launches(const b200_paged_attn_args* a) {
  using namespace b200;
  if (!a || a->num_q_tokens <= 0 || a->num_seqs <= 0) return 0;
  const AttnPlan pl = make_plan(*a);
  const bool split = pl.rows && pl.parts > 1 && a->workspace != nullptr &&
                     a->workspace_bytes >= split_workspace_bytes(a->num_q_tokens, a->num_q_heads, pl.parts, a->head_dim);
  return (pl.tile ? 1 : 0) + (pl.rows ? 1 : 0) + (split ? 1 : 0);
}

extern "C" int b200_paged_attention(const b200_paged_attn_args* a, void* stream) {
  using namespace b200;
  int rc = validate(a);
  if (rc != B200_OK) return rc;
  if (a->num_q_tokens == 0 || a->num_seqs == 0) return B200_OK;
  cudaStream_t st = (cudaStream_t)stream;
  if (a->kernel == 3 && !prefill_tc_supported(*a))
    return set_error(B200_ERR_UNSUPPORTED,
                     "paged_attention: the tcgen05 tile kernel needs bf16/fp16, head_dim 64/128, block_size | 128 and "
                     "a contiguous [blocks, block_size, kv_heads, head_dim] cache");
  const AttnPlan pl = make_plan(*a);
  const int skip_qlen = pl.skip_qlen;
  if (pl.tile) {
    rc = launch_prefill_tc(*a, skip_qlen, st);
    if (rc != B200_OK || !pl.rows) return rc;
  }
  const bool mma = pl.mma;
  if (a->kernel == 2 && !mma)
    return set_error(B200_ERR_UNSUPPORTED,
                     "paged_attention: tensor-core decode kernel does not instantiate "
                     "dtype=%d head_dim=%d block_size=%d group=%d",
                     a->dtype, a->head_dim, a->block_size, a->num_q_heads / a->num_kv_heads);
  int parts = pl.parts;
  const int psize = pl.psize;
  SplitWorkspace ws{};
  if (parts > 1) {
    const size_t need = split_workspace_bytes(a->num_q_tokens, a->num_q_heads, parts, a->head_dim);
    if (a->workspace == nullptr || a->workspace_bytes < need) {
      if (a->force_partitions > 1)
        return set_error(B200_ERR_INVALID_ARGUMENT,
                         "paged_attention: split-KV needs %zu workspace bytes, got %zu", need,
                         a->workspace_bytes);
      parts = 1;  // no scratch: single pass
    } else {
      ws = carve_workspace(a->workspace, a->num_q_tokens, a->num_q_heads, parts, a->head_dim,
                           psize);
    }
  }
  const SplitWorkspace* wsp = parts > 1 ? &ws : nullptr;
  if (wsp) {
    ws.skip_qlen = skip_qlen;
    ws.out_f16 = out_is_f16(*a);
  }
  rc = mma ? launch_decode_mma(*a, wsp, st, skip_qlen) : launch_generic(*a, wsp, st, skip_qlen);
  if (rc != B200_OK) return rc;
  if (wsp) rc = launch_split_reduce(*a, ws, st);
  return rc;
}

// ---- allocation-free variant (no counterpart in the reference) -------
// Split-KV scratch in a file-scope device array for callers that cannot allocate (CUDA graph capture, tight serving
// loops, harnesses that forbid cudaMalloc).  Every GPU that loads the library holds the array until exit.  In host
// code `g_static_scratch` is only the host-side stand-in of the symbol: its device address is looked up once per
// device, with that device current, through cudaGetSymbolAddress.
namespace b200 {
constexpr size_t kStaticScratchBytes = (size_t)B200_STATIC_SCRATCH_BYTES;
static __device__ __align__(256) unsigned char g_static_scratch[kStaticScratchBytes];

static int static_scratch_ptr(void** out) {
  static void* ptr_of_dev[16] = {nullptr};
  int dev = 0;
  B200_CUDA_OK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 16) return set_error(B200_ERR_UNSUPPORTED, "static scratch: device index %d out of range", dev);
  if (ptr_of_dev[dev] == nullptr) {
    void* p = nullptr;
    B200_CUDA_OK(cudaGetSymbolAddress(&p, g_static_scratch));
    if (p == nullptr) return set_error(B200_ERR_CUDA, "static scratch: cudaGetSymbolAddress returned a null address");
    ptr_of_dev[dev] = p;
  }
  *out = ptr_of_dev[dev];
  return B200_OK;
}
}  // namespace b200

extern "C" int b200_paged_attention_static_scratch(const b200_paged_attn_args* a, void* stream) {
  using namespace b200;
  int rc = validate(a);
  if (rc != B200_OK) return rc;
  if (a->num_q_tokens == 0 || a->num_seqs == 0) return B200_OK;
  b200_paged_attn_args c = *a;  // the caller's workspace fields are ignored
  c.workspace = nullptr;
  c.workspace_bytes = 0;
  const size_t need = b200_paged_attention_workspace_bytes(&c);  // 0 when the plan is a single pass
  if (need > kStaticScratchBytes)
    return set_error(B200_ERR_INVALID_ARGUMENT,
                     "paged_attention_static_scratch: this call's split-KV plan needs %zu scratch bytes, the built-in "
                     "array holds %zu (B200_STATIC_SCRATCH_BYTES); use b200_paged_attention with a caller workspace",
                     need, kStaticScratchBytes);
  void* scratch = nullptr;
  rc = static_scratch_ptr(&scratch);
  if (rc != B200_OK) return rc;
  // Nothing in the scratch has to start at zero: the partition kernels write every (row, head, partition) slot the
  // merge kernel reads in the same call, so the previous call's contents are never observed.
  c.workspace = scratch;
  c.workspace_bytes = kStaticScratchBytes;
  return b200_paged_attention(&c, stream);
}
