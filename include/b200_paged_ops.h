/*
 * b200_paged_ops.h -- C-ABI of the B200-native decode hot path.
 *
 * This is the drop-in boundary "B3" of SURVEY.md section 8(b): every entry
 * point below replaces one binding of the reference's native module
 * `_paged_ops` (vllm_metal/metal/paged_ops.cpp:1618-1900) or, for the
 * quantized linears, one MLX call the reference makes
 * (vllm_metal/gguf/mlx_native.py:212-239).  Plain pointers and sizes only; no
 * torch / MLX types.  All pointers are DEVICE pointers unless stated
 * otherwise; `stream` is a `cudaStream_t` passed as `void*` (NULL = legacy
 * default stream).  Every function returns 0 on success and a negative
 * `b200_status` otherwise; `b200_last_error()` then returns a thread-local,
 * NUL-terminated message.  Nothing here allocates device memory, synchronises
 * the device or falls back to the host: work is enqueued on `stream` and the
 * caller owns every buffer (same ownership rule as the reference: caches are
 * updated in place, paged_ops.cpp:1743-1749).
 */
#ifndef B200_PAGED_OPS_H_
#define B200_PAGED_OPS_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200_ABI_VERSION 3

/* Element types of activations / caches (reference: dtype_to_metal,
 * paged_ops.cpp:185-196 -- float, half, bfloat16 are the supported set). */
typedef enum b200_dtype {
  B200_F32 = 0,
  B200_F16 = 1,
  B200_BF16 = 2
} b200_dtype;

typedef enum b200_status {
  B200_OK = 0,
  B200_ERR_INVALID_ARGUMENT = -1, /* reference: std::invalid_argument -> ValueError */
  B200_ERR_RUNTIME = -2,          /* reference: std::runtime_error   -> RuntimeError */
  B200_ERR_UNSUPPORTED = -3,      /* shape / dtype the kernels do not instantiate */
  B200_ERR_CUDA = -4              /* a CUDA runtime call failed (message has the code) */
} b200_status;

/* Thread-local message of the last failing call on this thread ("" if none). */
const char* b200_last_error(void);
int b200_abi_version(void);

/* Programmatic dependent launch (sm_90+): when enabled (default) every kernel of this library is
 * launched with cudaLaunchAttributeProgrammaticStreamSerialization, so its prologue overlaps the
 * previous kernel's tail; each kernel executes griddepcontrol.wait before touching data a
 * previous kernel produced.  No reference analogue (MLX batches a whole step into one Metal
 * command buffer, paged_ops.cpp:597-631); 0 restores classic stream order for debugging. */
int b200_set_pdl(int enabled);

/* Split-KV partition length in tokens.  Reference: PARTITION_SIZE = 512
 * (vllm_metal/metal/constants.py:4, exported at paged_ops.cpp:1619-1622). */
int b200_partition_size(void);

/* Occupancy threshold of the split-KV gate: split-KV decode is engaged when
 * `num_q_heads * total_q_tokens < b200_min_decode_grid()`.  Reference:
 * min_decode_grid() = 8 * gpu_core_count (paged_ops.cpp:75-115); here it is a
 * multiple of the SM count of the current device.  Returns <0 on error. */
int b200_min_decode_grid(void);

/* Number of SMs of the current device (<0 on error). */
int b200_sm_count(void);

/* ---------------------------------------------------------------------------
 * reshape_and_cache -- replaces `_paged_ops.reshape_and_cache`
 * (binding paged_ops.cpp:1718-1749, primitive :1117-1204, kernel
 * kernels_v2/reshape_and_cache.metal:36-96).
 *
 * For every token t in [0, num_tokens): slot = slot_mapping[t]; slot < 0 is
 * padding and is skipped (reshape_and_cache.metal:61-65); otherwise rows
 * key[t], value[t] ([num_kv_heads, head_dim], contiguous) are copied bit for
 * bit to cache[slot / block_size][slot % block_size].  Slots >= num_blocks *
 * block_size are skipped as well (memory safety; the reference would write out
 * of bounds).  Caches are [num_blocks, block_size, num_kv_heads, head_dim]
 * with element strides `kv_block_stride` (dim 0) and `kv_token_stride`
 * (dim 1); the [num_kv_heads, head_dim] row is contiguous.
 * --------------------------------------------------------------------------- */
int b200_reshape_and_cache(const void* key, const void* value, void* key_cache,
                           void* value_cache, const int64_t* slot_mapping,
                           int num_tokens, int num_kv_heads, int head_dim,
                           int block_size, int64_t num_blocks,
                           int64_t kv_block_stride, int64_t kv_token_stride,
                           int dtype, void* stream);

/* Copy-on-write block copies, `cache[dst[i]] = cache[src[i]]` for K and V.
 * Reference: MetalPagedKVCache.copy_blocks (attention/caches/kv_cache.py:327-354).
 * `block_bytes` = bytes of one block of one cache; src/dst are device int32. */
int b200_copy_blocks(void* key_cache, void* value_cache, const int32_t* src,
                     const int32_t* dst, int num_pairs, int64_t block_bytes,
                     int64_t num_blocks, void* stream);

/* ---------------------------------------------------------------------------
 * paged_attention -- replaces `_paged_ops.paged_attention_primitive`
 * (binding paged_ops.cpp:1790-1854; argument set = bind_paged_attn_buffers,
 * paged_ops.cpp:244-283; router :436-716; kernels
 * kernels_v2/pagedattention.metal:817-1395 (per token), :1917-2144 (reduce),
 * kernels_v2/pagedattention_tiled.metal:114-525 (varlen prefill / mixed)).
 *
 *   query, out    [num_q_tokens, num_q_heads, head_dim], row stride q_stride /
 *                 out_stride elements, heads contiguous
 *   key/value     [num_blocks, block_size, num_kv_heads, head_dim] with element
 *                 strides kv_block_stride, kv_token_stride, kv_head_stride
 *   block_tables  int32 [num_seqs, max_blocks_per_seq], rows padded with 0
 *   seq_lens      int32 [num_seqs]  KV length INCLUDING this step's tokens
 *   cu_seqlens_q  int32 [num_seqs + 1], starts at 0, ends at num_q_tokens
 *   sinks         float32 [num_q_heads] or NULL
 *
 * Row r = (sequence s, position q_pos of q_len) attends KV positions
 * [max(0, eff - W), eff) with eff = seq_len - q_len + q_pos + 1
 * (pagedattention.metal:913, :1129-1133); W = sliding_window, < 0 disables.
 * Softmax in the log2 domain with fp32 accumulation, optional softcap
 * tanh(s/c)*c, sink folded once per row, out = O / (l + 1e-6)
 * (pagedattention.metal:1138, :1318-1328, :1383).  Rows with eff <= 0 are left
 * untouched (:917-920).
 *
 * `workspace` is scratch for split-KV partials; size it with
 * b200_paged_attention_workspace_bytes() for the same args: it returns exactly what the call's own plan needs
 * (0 when that plan is single pass).  Passing less is legal: the call then runs single pass.  `force_partitions`: 0 = heuristic gate,
 * 1 = never split, >1 = split into partitions of b200_partition_size() tokens
 * whenever max_seq_len allows (used by the parity tests to pin both routes).
 * `kernel`: 0 = auto, 1 = generic scalar kernel, 2 = tensor-core decode kernel, 3 = tcgen05 tile kernel for
 * every sequence (2 and 3 fail with B200_ERR_UNSUPPORTED if the shape is not instantiated).  Auto sends
 * sequences with >= 16 query rows to the tile kernel and the rest to the per-row kernels.
 * --------------------------------------------------------------------------- */
typedef struct b200_paged_attn_args {
  void* out;
  const void* query;
  const void* key_cache;
  const void* value_cache;
  const int32_t* block_tables;
  const int32_t* seq_lens;
  const int32_t* cu_seqlens_q;
  const float* sinks;
  void* workspace;
  size_t workspace_bytes;
  int32_t num_q_tokens;
  int32_t num_seqs;
  int32_t num_q_heads;
  int32_t num_kv_heads;
  int32_t head_dim;
  int32_t block_size;
  int32_t max_blocks_per_seq;
  int32_t max_seq_len;
  int32_t sliding_window;
  int32_t dtype;
  int32_t force_partitions;
  int32_t kernel;
  int64_t q_stride;
  int64_t out_stride;
  int64_t kv_block_stride;
  int64_t kv_token_stride;
  int64_t kv_head_stride;
  int64_t num_blocks;
  float scale;
  float softcap;
  int32_t out_dtype; /* 0 or `dtype`: out has the query dtype; B200_F16 with dtype = B200_BF16: the bf16-rounded
                        result is stored as fp16 (feeds the fp16-operand o_proj without a cast launch) */
  int32_t window_seqlen_q; /* 0 / 1: per-row dispatch.  > 1: spec-verify window hint of the reference binding
                              (paged_ops.cpp:841-899): must equal the longest cu_seqlens_q segment (the caller checks,
                              as the reference does host-side); the batch runs the decode kernel in window mode --
                              floor(16 / (H_q/H_kv)) query rows of one sequence per CTA share every KV read
                              (pagedattention.metal:1396-1914) -- and never the tile kernel (paged_ops.cpp:468-476).
                              Results are bitwise those of the expanded one-row-per-segment dispatch. */
} b200_paged_attn_args;

size_t b200_paged_attention_workspace_bytes(const b200_paged_attn_args* args);
/* Kernels one b200_paged_attention call launches for these args (tile kernel + per-row kernel + split-KV merge,
 * from the same plan the call itself uses; `workspace` / `workspace_bytes` as they will be passed). */
int b200_paged_attention_num_launches(const b200_paged_attn_args* args);
int b200_paged_attention(const b200_paged_attn_args* args, void* stream);

/* Allocation-free variant of b200_paged_attention.  It has NO counterpart in the reference: it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  Same arguments, routing, kernels and results as b200_paged_attention, except that
 * args->workspace / workspace_bytes are ignored: the split-KV partials live in a file-scope device array of
 * B200_STATIC_SCRATCH_BYTES that every GPU loading the library holds until exit.  A call whose split-KV plan needs
 * more (b200_paged_attention_workspace_bytes(args) > B200_STATIC_SCRATCH_BYTES) is rejected with
 * B200_ERR_INVALID_ARGUMENT; single-pass plans need none.  The call allocates nothing (no cudaMalloc / cudaFree, no
 * allocating library routine).  All calls on a device share the one array, which keeps the last call's contents
 * (nothing in it has to start at zero: each call writes every slot it reads); calls that overlap on the GPU --
 * different streams, or graph branches without a dependency -- RACE on it and must be serialised by the caller. */
#define B200_STATIC_SCRATCH_BYTES (32u << 20)
int b200_paged_attention_static_scratch(const b200_paged_attn_args* args, void* stream);

/* ---------------------------------------------------------------------------
 * Affine group-quantized linears -- replace the MLX calls the reference makes:
 * `mx.quantized_matmul(x, w, scales, biases, transpose=True, group_size, bits)`
 * (vllm_metal/gguf/mlx_native.py:212-222; reached through
 * nn.QuantizedLinear from attention/impls/sdpa.py:297,306,340,343,739) and
 * `mx.dequantize(w[ids], scales[ids], biases[ids])` (mlx_native.py:232-239).
 *
 *   w       uint32 [N, K*bits/32]; element k of a row in word k/(32/bits) at
 *           bit offset (k % (32/bits))*bits, lowest bits first
 *   scales  [N, K/group_size], biases same; dtype sb_dtype
 *   x       [M, K] row stride x_stride elements; y [M, N] row stride y_stride
 *   bias    dense [N] of dtype `dtype` or NULL (gguf/wrappers.py:71-75)
 *
 *   y[m,n] = sum_k x[m,k] * (scales[n,k/g] * q[n,k] + biases[n,k/g]) (+ bias[n])
 * fp32 accumulation, one rounding to `dtype`.  bits in {4, 8}; group_size in
 * {32, 64, 128}; K % group_size == 0.
 * x_dtype = dtype, or B200_F16 with dtype = B200_BF16: the tcgen05 kernels multiply fp16 x by weights
 * dequantised to fp16 (bf16 would cost 3 significant bits of s*q+b and kind::f16 does not mix the two),
 * so bf16 callers stage x through b200_cast_bf16_to_f16 (or have the producer emit fp16); with a bf16 x
 * b200_qlinear runs the mma.sync kernel.
 * --------------------------------------------------------------------------- */
int b200_qlinear(const void* x, const uint32_t* w, const void* scales,
                 const void* biases, const void* bias, void* y, int M, int N,
                 int K, int group_size, int bits, int x_dtype, int dtype, int sb_dtype,
                 int64_t x_stride, int64_t y_stride, void* stream);

/* Stacked experts (mixture-of-experts MLP): `mx.gather_qmm(x, w, scales, biases, rhs_indices, transpose=True, ...)` of
 * mlx-lm's QuantizedSwitchLinear (named in vllm_metal/quant/awq_loader.py:171-173) over rows ALREADY GROUPED by
 * expert.  w [E, N, K*bits/32], scales / biases [E, N, K/group_size], bias [E, N] or NULL, all contiguous;
 * counts: HOST int32 [E]; expert e multiplies the counts[e] rows of x that follow the rows of experts 0..e-1, results
 * in the same rows of y.  One b200_qlinear launch per expert with rows, same kernels and bits as b200_qlinear. */
int b200_qlinear_experts(const void* x, const uint32_t* w, const void* scales, const void* biases,
                         const void* bias, void* y, const int32_t* counts, int num_experts, int N, int K,
                         int group_size, int bits, int x_dtype, int dtype, int sb_dtype, int64_t x_stride,
                         int64_t y_stride, void* stream);

/* Test / tuning knob: which kernel b200_qlinear runs.  0 = auto (tcgen05 + TMEM + TMA kernel
 * whenever the shape qualifies, else the mma.sync kernel), 1 = always the mma.sync kernel,
 * 2 = tcgen05 kernel or B200_ERR_UNSUPPORTED.  Process-wide; returns 0 on success. */
int b200_set_qlinear_kernel(int kernel);

/* Tuning knob of the tcgen05 kernel's weight-streaming (M <= 64) regime, used by the sweep in
 * profiles/bench_kernels.py: `ksplit` = K split over a thread-block cluster (0 = heuristic, else
 * 1, 2, 4 or 8), `dequant_warps` = 0 (default), 4 or 8.  Process-wide; returns 0 on success. */
int b200_set_qlinear_tuning(int ksplit, int dequant_warps);

/* Kernel-variant knob of the M <= 64 regime for the sweep.  0 = default: dequantised A tile in
 * TENSOR MEMORY (tcgen05.st / TS-form MMA), TMA-staged scales/biases, packed ring 4 / operand
 * ring 4; 1 = A tile in shared memory, per-thread scale loads, rings 3/2; 2 = A tile in shared
 * memory, TMA-staged scales, rings 2/3.  Process-wide. */
int b200_set_qlinear_variant(int variant);

/* Debug/profiling: when `dev_ptr` is non-null every tcgen05 qlinear CTA writes up to 64 x uint64
 * stamps (%globaltimer: start, setup, dependency met, first weights, dequant done, accumulator
 * done, epilogue done; cycle counters; per-stage times) at dev_ptr[cta * 64 ...]; null (default) disables it. */
int b200_set_qlinear_trace(void* dev_ptr);

/* Load-time repack into the streaming layout read by b200_qlinear_prepacked: same shape as
 * `w` ([N, K*bits/32] uint32); for bits == 4 the 8 nibbles of every word are permuted so that
 * nibbles (i, i+4) hold elements (2i, 2i+1); bits == 8 is a copy.  `inverse` != 0 undoes it
 * (bit-exact round trip).  `out` may alias `w`.  Reference counterpart: none -- MLX streams its
 * own layout; this is the B200 analogue of a loader-side weight transform such as
 * quant/awq_loader.py:49-224. */
int b200_qlinear_prepack(const uint32_t* w, uint32_t* out, int N, int K, int bits,
                         int inverse, void* stream);

/* b200_qlinear over weights in the streaming layout (scales / biases unchanged).  Only the
 * tcgen05 kernel reads this layout: shapes it does not take fail with B200_ERR_UNSUPPORTED. */
int b200_qlinear_prepacked(const void* x, const uint32_t* wp, const void* scales,
                           const void* biases, const void* bias, void* y, int M, int N,
                           int K, int group_size, int bits, int x_dtype, int dtype, int sb_dtype,
                           int64_t x_stride, int64_t y_stride, void* stream);

/* out[r, :] = dequantize(row ids[r]) for r in [0, num_rows); ids == NULL means
 * rows 0..num_rows-1 (plain mx.dequantize).  ids are int32 (token ids). */
int b200_dequantize_rows(const uint32_t* w, const void* scales,
                         const void* biases, const int32_t* ids, void* out,
                         int num_rows, int N, int K, int group_size, int bits,
                         int sb_dtype, int out_dtype, void* stream);

/* ---- streaming-layout quantized linear: persistent stream-K kernel (csrc/qgemm_stream.cu) ----------------
 * The decode-regime (M <= 64) replacement for mx.quantized_matmul as nn.QuantizedLinear calls it
 * (vllm_metal/gguf/mlx_native.py:212-222, attention/impls/sdpa.py:297-343,739) and, with
 * epilogue = 1, for the SwiGLU the reference fuses in vllm_metal/compiled_mlp.py:1-22
 * (y[m,i] = silu(gate[m,i]) * up[m,i], gate|up = rows [0,N/2) | [N/2,N) of the weight).
 *
 * b200_qstream_prepack turns an MLX-affine triple (w uint32 [N, K*bits/32], scales/biases [N, K/g]) into
 * ONE device blob of b200_qstream_blob_bytes(...) bytes: per (256-row super-tile, k-slab) unit the
 * packed words in the order the kernel's dequant threads read them, followed by fp16 (scale, bias)
 * pairs.  It is a load-time call and synchronises the stream: the kernel prefetches weights BEFORE its
 * programmatic-dependency wait, so a blob must be stream-complete before the first b200_qstream_linear
 * that reads it.  Fails with B200_ERR_UNSUPPORTED when a scale/bias does not fit fp16 (keep the MLX
 * layout and b200_qlinear then).  b200_qstream_unpack is the bit-exact inverse of the integer part.
 *
 * b200_qstream_linear: x [M, K] of x_dtype fp16 or bf16 (row stride x_stride elements), y [M, N] (or
 * [M, N/2] with epilogue = 1) of `dtype` (bf16/fp16), optional dense bias [N] in `dtype` (epilogue 0 only).
 * The tensor-core operands are fp16 x fp16 (tcgen05 kind::f16 faults on mixed fp16/bf16 operands, and the
 * dequantised weights need fp16's 11 significant bits): a bf16 x is first copied to fp16 inside the
 * workspace (exact below 65504 in magnitude, saturating above; M * K <= 4 Mi elements), so producers that
 * can emit fp16 directly save that launch.  `workspace` is
 * b200_qstream_workspace_bytes() of device memory, ZERO-initialised once by the caller and shared by
 * all calls on one stream (fp32 partials of super-tiles cut between CTAs + their arrival counters). */
int64_t b200_qstream_blob_bytes(int N, int K, int bits, int group_size, int swiglu);
int64_t b200_qstream_workspace_bytes(void);
int b200_qstream_prepack(const uint32_t* w, const void* scales, const void* biases, void* blob,
                         int N, int K, int bits, int group_size, int sb_dtype, int swiglu,
                         void* stream);
int b200_qstream_unpack(const void* blob, uint32_t* w_out, int N, int K, int bits,
                        int group_size, int swiglu, void* stream);
int b200_qstream_linear(const void* x, const void* blob, const void* bias, void* y, int M, int N,
                        int K, int group_size, int bits, int x_dtype, int dtype, int epilogue,
                        int64_t x_stride, int64_t y_stride, void* workspace,
                        int64_t workspace_bytes, void* stream);
/* bf16 [rows, cols] (row stride x_stride elements) -> contiguous fp16, saturating at +-65504; cols % 8 == 0. */
int b200_cast_bf16_to_f16(const void* x, void* y, int rows, int cols, int64_t x_stride, void* stream);
/* tuning / tracing knobs (0 / NULL = default) */
int b200_set_qstream_grid(int grid);
int b200_set_qstream_trace(void* dev_ptr);

#ifdef __cplusplus
}
#endif
#endif /* B200_PAGED_OPS_H_ */
