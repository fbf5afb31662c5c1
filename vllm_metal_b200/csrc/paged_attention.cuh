// Internal interface between the paged-attention router and its kernels.
#pragma once
#include "common.cuh"

namespace b200 {

// Split-KV scratch layout (fp32 partials; SURVEY.md appendix A.4 -- the
// reference keeps tmp_out in the query dtype, paged_ops.cpp:647-653):
//   tmp_out  float [T, H_q, P, D]   un-normalised partial numerators
//   max_l2   float [T, H_q, P]      running max in the log2 domain
//   exp_sum  float [T, H_q, P]      sum of 2^(s - max)
struct SplitWorkspace {
  float* tmp_out;
  float* max_l2;
  float* exp_sum;
  int num_partitions;  // P
  int partition_size;  // tokens
  int skip_qlen;       // > 0: rows of sequences with q_len >= skip_qlen belong to the tile kernel (left untouched here)
  int out_f16;         // bf16 call whose `out` buffer is fp16: store the bf16-rounded value as fp16 (args.out_dtype)
};

// One output element: rounded to the call's dtype T; with out_f16 (T = bf16) that value is stored as fp16 -- exact
// below 65504 -- so the o_proj that follows (fp16 tensor-core operands) needs no cast launch.
template <typename T>
__device__ __forceinline__ void store_out(T* row, int d, float v, int out_f16) {
  const T t = from_float<T>(v);
  if (out_f16) reinterpret_cast<__half*>(row)[d] = __float2half_rn(fminf(fmaxf(to_float(t), -65504.f), 65504.f));
  else row[d] = t;
}

inline size_t split_workspace_bytes(int64_t T, int64_t Hq, int64_t P, int64_t D) {
  return (size_t)(T * Hq * P * (D + 2)) * sizeof(float);
}

inline SplitWorkspace carve_workspace(void* ws, int64_t T, int64_t Hq, int64_t P, int64_t D,
                                      int partition_size) {
  SplitWorkspace w;
  w.tmp_out = reinterpret_cast<float*>(ws);
  w.max_l2 = w.tmp_out + T * Hq * P * D;
  w.exp_sum = w.max_l2 + T * Hq * P;
  w.num_partitions = (int)P;
  w.partition_size = partition_size;
  w.skip_qlen = 0;
  w.out_f16 = 0;
  return w;
}

// Row (token) -> sequence lookup; reference find_seq_idx,
// pagedattention.metal:772-784 (binary search over cu_seqlens_q).
__device__ __forceinline__ int find_seq_idx(const int32_t* __restrict__ cu, int num_seqs, int t) {
  int lo = 0, hi = num_seqs;  // invariant: cu[lo] <= t < cu[hi]
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if (cu[mid] <= t) lo = mid; else hi = mid;
  }
  return lo;
}

// Per-row KV range.  `eff` = number of KV positions row t may see
// (pagedattention.metal:913); with a sliding window W only [eff-W, eff) is kept
// (:1129-1133).  Split-KV partitions are laid out from `base` (the window start
// rounded down to a 16-token tile), so windowed rows need ceil((W+15)/psize)
// partitions instead of ceil(max_seq_len/psize): pages wholly outside the window
// are never touched (the reference iterates and masks them).
struct RowInfo {
  int seq, eff, beg, base;
};
__device__ __forceinline__ RowInfo row_info(const int32_t* __restrict__ cu,
                                            const int32_t* __restrict__ seq_lens, int num_seqs,
                                            int t, int sliding_window, int skip_qlen = 0) {
  RowInfo r;
  r.seq = find_seq_idx(cu, num_seqs, t);
  const int q_start = cu[r.seq];
  const int q_len = cu[r.seq + 1] - q_start;
  r.eff = seq_lens[r.seq] - q_len + (t - q_start) + 1;
  if (skip_qlen > 0 && q_len >= skip_qlen) r.eff = 0;  // mixed batch: the tile kernel owns this sequence
  r.beg = 0;
  r.base = 0;
  if (sliding_window >= 0) {
    r.beg = max(0, r.eff - sliding_window);
    r.base = r.beg & ~15;
  }
  return r;
}

// Launchers implemented in paged_attention_generic.cu / paged_attention_decode.cu.
// `ws == nullptr` => single pass.
int launch_generic(const b200_paged_attn_args& a, const SplitWorkspace* ws, cudaStream_t st, int skip_qlen = 0);
inline int out_is_f16(const b200_paged_attn_args& a) { return a.out_dtype == B200_F16 && a.dtype == B200_BF16 ? 1 : 0; }
bool decode_mma_supported(const b200_paged_attn_args& a);
int launch_decode_mma(const b200_paged_attn_args& a, const SplitWorkspace* ws, cudaStream_t st, int skip_qlen = 0);
int decode_window_rows(const b200_paged_attn_args& a);  // > 1: window mode (query rows of one sequence per CTA)
int launch_split_reduce(const b200_paged_attn_args& a, const SplitWorkspace& ws, cudaStream_t st);
// Tile kernel (paged_attention_prefill.cu): sequences with q_len >= min_qlen, one CTA per (128-row Q tile, q head).
bool prefill_tc_supported(const b200_paged_attn_args& a);
int launch_prefill_tc(const b200_paged_attn_args& a, int min_qlen, cudaStream_t st);

}  // namespace b200
