"""Kernel constants, single-sourced (reference: vllm_metal/metal/constants.py).

``PARTITION_SIZE`` is also compiled into the library (``b200_partition_size``);
``get_ops()`` asserts the two agree at load time.
"""

# Split-KV partition length in tokens (vllm_metal/metal/constants.py:4).
PARTITION_SIZE = 512

# Kernel page sizes the attention kernels accept directly; other cache block
# sizes are translated (vllm_metal/attention/impls/sdpa.py:49,115-163).
KERNEL_BLOCK_SIZES = (32, 16, 8)

# C-ABI dtype codes (include/b200_paged_ops.h: b200_dtype).
DTYPE_F32, DTYPE_F16, DTYPE_BF16 = 0, 1, 2

# Spec-verify window mode (vllm_metal/metal/constants.py:8-19).  The reference shares each KV read between
# PA_WINDOW_ROWS = 2 query rows of a threadgroup; here the decode kernel's 16-row MMA tile holds
# floor(16 / (H_q / H_kv)) query rows of one sequence (4 for Llama-3's 4-head groups), so the row count per CTA
# follows the model, not a constant.  The head-size bound of the hint is the reference's.
PA_WINDOW_MAX_HEAD_SIZE = 256
