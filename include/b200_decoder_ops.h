/*
 * b200_decoder_ops.h -- C-ABI of the ops either side of the hot path
 * (SURVEY.md section 8(f) rows N1-N3): RMSNorm, RoPE, SiLU*mul, greedy argmax.
 * In the reference these are MLX built-ins called by the mlx_lm layer code
 * around `sdpa_forward` (attention/impls/sdpa.py:348-368 q/k norm + RoPE,
 * attention/impls/varlen_rope_compat.py:70-82 batched single-token RoPE,
 * v1/sampling_batch.py:342-348 mlx_greedy_tokens).  Same conventions as
 * b200_paged_ops.h: device pointers, `stream` is a cudaStream_t, 0 on success.
 */
#ifndef B200_DECODER_OPS_H_
#define B200_DECODER_OPS_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* out[r,:] = x[r,:] * rsqrt(mean(x[r,:]^2) + eps) * weight   (mx.fast.rms_norm).
 * If `residual` != NULL: residual[r,:] += x[r,:] first (in place, rounded to
 * `dtype`), and the norm is taken of the updated residual (fused
 * add + norm used between decoder sub-blocks).  `weight` may be NULL (=1).
 * out_dtype = dtype, or B200_F16 with dtype = B200_BF16: the bf16-rounded result stored as fp16
 * (exact below 65504, saturating) -- the operand type of b200_qstream_linear, saving its staging cast. */
int b200_rms_norm(const void* x, void* residual, const void* weight, void* out,
                  int rows, int dim, int64_t x_stride, int64_t res_stride,
                  int64_t out_stride, float eps, int dtype, int out_dtype, void* stream);

/* In-place rotary embedding of q [T, Hq, D] and k [T, Hk, D] (k may be NULL)
 * at per-token positions (int32 [T]); the first rot_dim dims of each head are
 * rotated.  traditional = 0: half-split pairs (i, i + rot_dim/2) (mlx
 * nn.RoPE(traditional=False), the Llama/Qwen3 default); 1: interleaved pairs
 * (2i, 2i+1).  inv_freq[i] = base^(-2i/rot_dim) computed in fp32. */
int b200_rope(void* q, void* k, const int32_t* positions, int num_tokens,
              int num_q_heads, int num_kv_heads, int head_dim, int rot_dim,
              float base, int traditional, int64_t q_stride, int64_t k_stride,
              int dtype, void* stream);

/* Fused per-head RMSNorm -> RoPE -> KV-cache write: the steps between the q/k/v projections
 * and the attention kernel in the reference's sdpa_forward (attention/impls/sdpa.py:348-368
 * q/k norm + RoPE, :542-544, :638 reshape_and_cache) as ONE launch.  Bit-identical to
 * b200_rms_norm (per head, weight q_norm_weight / k_norm_weight [head_dim], NULL = no norm)
 * followed by b200_rope followed by b200_reshape_and_cache:
 *   q [T, Hq, D] (row stride q_stride) and k [T, Hk, D] (k_stride) are updated in place,
 *   the updated k row and the v row [T, Hk, D] (v_stride) are written to
 *   key_cache / value_cache [num_blocks, block_size, Hk, D] at slot_mapping[t]
 *   (int64; negative or out-of-range = padding: q/k are still updated, nothing is cached).
 * rot_dim = 0 disables RoPE (positions may then be NULL). */
int b200_norm_rope_cache(void* q, void* k, const void* v, const void* q_norm_weight,
                         const void* k_norm_weight, const int32_t* positions,
                         const int64_t* slot_mapping, void* key_cache, void* value_cache,
                         int num_tokens, int num_q_heads, int num_kv_heads, int head_dim,
                         int rot_dim, float base, int traditional, float eps,
                         int64_t q_stride, int64_t k_stride, int64_t v_stride,
                         int block_size, int64_t num_blocks, int64_t kv_block_stride,
                         int64_t kv_token_stride, int dtype, void* stream);

/* out[r, i] = silu(gate_up[r, i]) * gate_up[r, inter + i], i in [0, inter). */
int b200_silu_mul(const void* gate_up, void* out, int rows, int inter,
                  int64_t in_stride, int64_t out_stride, int dtype, void* stream);

/* out[r] = index of the first maximum of logits[r, :vocab] (mx.argmax). */
int b200_argmax(const void* logits, int32_t* out, int rows, int vocab,
                int64_t stride, int dtype, void* stream);

/* y[r,:] += x[r,:] (residual add, rounded to dtype). */
int b200_add_inplace(void* y, const void* x, int rows, int dim, int64_t y_stride,
                     int64_t x_stride, int dtype, void* stream);

/* Device-side step bookkeeping for back-to-back decode steps (the host-side
 * equivalent is the reference's per-step packing, v1/model_runner.py:1200-1240 +
 * attention/context.py:219-223): for every request r in [0, batch):
 *   input_ids[r] = sampled[r]; positions[r] += 1; seq_lens[r] += 1;
 *   slot_mapping[r] = block_tables[r, pos / block_size] * block_size + pos % block_size
 * with pos the new position.  block_tables is int32 [batch, max_blocks]. */
int b200_advance_decode(int32_t* input_ids, int32_t* positions, int32_t* seq_lens,
                        int64_t* slot_mapping, const int32_t* block_tables,
                        const int32_t* sampled, int batch, int max_blocks, int block_size,
                        void* stream);

/* ---- tensor-parallel sum over NVLink peer memory ----------------------
 * The reference runs single-device (vllm_metal/platform.py:524-529); this is the one exchange
 * step the head/column sharding of the path adds: y = sum over ranks of the o_proj / down_proj
 * partial outputs.  One process per GPU; buffers are shared with CUDA IPC.
 *
 * b200_ipc_alloc: cudaMalloc + zero `bytes` and export a 64-byte cudaIpcMemHandle_t.
 * b200_ipc_open : map a peer's handle (lazy peer access); b200_ipc_close / b200_ipc_free undo them.
 * b200_allreduce_sum: in-place sum of x [numel] of `dtype` over `world` in {2,4,8} ranks.
 *   symm_ptrs[r] : rank r's staging buffer (2 * half_bytes; own pointer at [rank], peers mapped)
 *   flag_ptrs[r] : rank r's flag array, uint32 [grid * world], zero-initialised
 *   epoch        : local uint32 [grid], zero-initialised
 *   grid         : CTAs; must be identical on every rank and every call using these buffers
 * Every rank must call it in the same order with the same numel.  Result is bit-identical on
 * all ranks (fixed rank-order fp32 accumulation, one rounding to `dtype`). */
int b200_ipc_alloc(int64_t bytes, void** ptr, unsigned char* handle64);
int b200_ipc_open(const unsigned char* handle64, void** ptr);
int b200_ipc_close(void* ptr);
int b200_ipc_free(void* ptr);
int b200_allreduce_sum(void* x, int64_t numel, int dtype, void* const* symm_ptrs,
                       void* const* flag_ptrs, void* epoch, int rank, int world,
                       int64_t half_bytes, int grid, void* stream);
/* Fused tensor-parallel epilogue of a decoder sub-block: s = all-reduce(x) (as above, x [rows, hidden] contiguous,
 * left untouched), residual += s (in place, rounded to dtype), out = rms_norm(residual) * weight -- bit-identical
 * to b200_allreduce_sum followed by b200_rms_norm(x, residual, ...), in one launch.  Shares buffers, flags and
 * epochs with b200_allreduce_sum (same `grid`); hidden % 8 == 0, hidden <= 8192; out_dtype as in b200_rms_norm.
 * twoshot != 0: row r is reduced + normed on rank r % world only and the results are gathered (less NVLink traffic,
 * one more flag exchange; same bits): the staging buffers must then be 6 * half_bytes (2 halves of partials + 2 x 2
 * halves of results) and the flag arrays 2 * grid * world. */
int b200_allreduce_add_rmsnorm(void* x, void* residual, const void* weight, void* out, int rows, int hidden,
                               float eps, int dtype, int out_dtype, void* const* symm_ptrs,
                               void* const* flag_ptrs, void* epoch, int rank, int world,
                               int64_t half_bytes, int grid, int twoshot, void* stream);

/* Push form of the fused tensor-parallel epilogue with a ROW-SHARDED residual stream: row r belongs to rank
 * r % world.  Non-owners store their partial of r into the owner's packet buffer; the owner adds the partials in rank
 * order (the arithmetic of b200_allreduce_add_rmsnorm, bit for bit), adds its residual row in place, norms, and
 * stores the normed row into every peer's buffer.  On return (stream order) `out` holds every normed row on every
 * rank; `residual` is updated -- and must have been valid -- on the rows this rank owns ONLY (the other rows are
 * neither read nor written: a caller keeps using this entry point, or re-materialises the stream, consistently
 * within a forward pass).  Packets are 8-byte words {4 bytes of data, call epoch}: no flag arrays, no fences, two
 * one-way NVLink trips per call.
 *   box_ptrs[r] : rank r's packet buffer of box_bytes, zero-initialised (own pointer at [rank], peers mapped);
 *                 box_bytes >= 4 * ceil(rows / world) * world * hidden * 4 (2 parities x inbox + outbox)
 *   epoch       : local uint32 [grid], zero-initialised, used by this entry point only; grid % world == 0 (clusters of `world` CTAs)
 * Every rank must call it in the same order with the same rows / hidden. */
int b200_allreduce_add_rmsnorm_push(const void* x, void* residual, const void* weight, void* out, int rows,
                                    int hidden, float eps, int dtype, int out_dtype, void* const* box_ptrs,
                                    int64_t box_bytes, void* epoch, int rank, int world, int grid, void* stream);

/* Greedy token of a vocab-parallel head (lm_head column-parallel over the vocabulary +
 * a tiny (max, index) exchange instead of gathering logits).  `logits` [rows, shard_vocab] is this rank's slice
 * (global token id = vocab_offset + column); out[r] = index of the first maximum over ALL ranks' slices -- what
 * b200_argmax / mx.argmax (v1/sampling_batch.py:342-348) gives on the gathered row -- identical on every rank.
 *   slot_ptrs[r] : rank r's slot table, 2 * world * max_rows * 16 bytes, zero-initialised (own pointer at [rank],
 *                  peers mapped through b200_ipc_open); epoch: local uint32 [max_rows], zero-initialised.
 * Two launches (arg-max of the slice, then the exchange: every rank stores its winner into all peers' tables and
 * polls its own).  Every rank must call it in the same order with the same `rows`. */
int b200_argmax_vocab_parallel(const void* logits, int32_t* out, int rows, int shard_vocab, int64_t stride,
                               int dtype, int vocab_offset, void* const* slot_ptrs, void* epoch, int rank,
                               int world, int max_rows, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200_DECODER_OPS_H_ */
