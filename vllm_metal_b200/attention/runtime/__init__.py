"""Paged-attention runtimes (reference: vllm_metal/attention/runtime/)."""
