"""Paged-attention host side: per-step context, KV cache, SDPA forward."""
