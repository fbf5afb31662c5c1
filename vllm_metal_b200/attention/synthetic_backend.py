"""Synthetic attention backend that only advertises the kernel block alignment.

Mirrors ``vllm_metal/attention/synthetic_backend.py:13-47``: vLLM's block-size selection
(``Platform.update_block_size_for_backend``) reads ``get_supported_kernel_block_sizes`` from a
backend class; the real paged attention of this plugin is ``attention/impls/sdpa.py`` calling
the C-ABI kernels, never a vLLM attention backend.  Kept in its own module so importing the
platform during plugin discovery does not pull ``vllm.v1.attention.backend`` in early.
"""

from vllm.v1.attention.backend import AttentionBackend, MultipleOf


class B200Backend(AttentionBackend):
    """Advertises ``MultipleOf(16)``: the decode kernels take pages of 8/16/32 tokens and larger
    cache blocks are translated (``_pick_kernel_block_size``), so every multiple of 16 works and
    16 is the default.  The dispatch methods raise on purpose (reference behaviour)."""

    @staticmethod
    def get_name() -> str:
        return "B200_PAGED_ATTN"

    @staticmethod
    def get_supported_kernel_block_sizes() -> list:
        return [MultipleOf(16)]

    @staticmethod
    def get_impl_cls():
        raise NotImplementedError("B200Backend is a block-size marker, not a dispatchable backend")

    @staticmethod
    def get_builder_cls():
        raise NotImplementedError("B200Backend is a block-size marker, not a dispatchable backend")

    @staticmethod
    def get_kv_cache_shape(*args, **kwargs):
        raise NotImplementedError("B200Backend is a block-size marker, not a dispatchable backend")
