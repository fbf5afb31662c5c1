"""Decoder definitions that stand in for the third-party ``mlx_lm`` models the
reference imports (Llama / Qwen3 / gpt-oss attention blocks)."""
