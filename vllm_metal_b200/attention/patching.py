"""Find a model's attention modules and install paged-attention wrappers over them (reference:
vllm_metal/attention/patching.py:14-101).  One loop shared by every runtime; a layer whose wrapper cannot be
built must raise -- a silently skipped layer would keep its unpaged attention."""

from __future__ import annotations

from collections.abc import Callable
from typing import Any

# attribute names probed on each layer, in priority order (reference _ATTN_ATTR_NAMES)
ATTN_ATTR_NAMES = ("self_attn", "linear_attn", "attention")


def find_layers(model: Any) -> Any:
    """``model.model.layers`` / ``model.language_model.layers`` / ``model.layers`` (reference find_layers)."""
    root = model
    for name in ("language_model", "model"):
        if hasattr(root, name) and not hasattr(root, "layers"):
            root = getattr(root, name)
    if hasattr(root, "layers"):
        return root.layers
    raise ValueError(f"Cannot find transformer layers in model of type {type(model)}")


def find_attn_attr(layer: Any) -> str | None:
    for name in ATTN_ATTR_NAMES:
        if hasattr(layer, name):
            return name
    return None


def walk_and_wrap(model: Any, wrap_layer: Callable[[int, Any], Any], *, only_layers: list[int] | None = None) -> int:
    """Call ``wrap_layer(layer_idx, attn)`` for every layer that has an attention submodule and install what it
    returns (the same object = rebound in place).  Returns the number of patched layers."""
    only = set(only_layers) if only_layers is not None else None
    patched = 0
    for idx, layer in enumerate(find_layers(model)):
        if only is not None and idx not in only:
            continue
        attr = find_attn_attr(layer)
        if attr is None:
            continue
        attn = getattr(layer, attr)
        wrapper = wrap_layer(idx, attn)
        if wrapper is None:
            raise RuntimeError(f"walk_and_wrap: wrap_layer returned None for layer {idx} ({type(attn).__name__}); "
                               f"it must return a wrapper or raise.")
        if wrapper is not attn:
            setattr(layer, attr, wrapper)
        patched += 1
    return patched
