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
