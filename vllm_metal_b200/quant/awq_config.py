"""Preflight for AWQ checkpoints: canonicalise the HF ``quantization_config`` or refuse it.

Mirror of the reference's ``vllm_metal/quant/awq_config.py:62-131``
(``normalize_quant_config`` + ``UnsupportedQuantizationConfigError``): the one AWQ
flavour admitted is 4-bit, group 128, asymmetric (``zero_point``), GEMM packing;
AutoAWQ's alias keys ``w_bit`` / ``q_group_size`` are accepted.  Everything else is
rejected *before* any weight is touched, naming the offending field, so a config
this build was not validated on never reaches ``transform_awq_weights``.
"""

from __future__ import annotations

from collections.abc import Mapping
from typing import Any

SUPPORTED = {"bits": 4, "group_size": 128, "zero_point": True, "version": "gemm"}


class UnsupportedQuantizationConfigError(ValueError):
    """An AWQ ``quantization_config`` outside the supported subset (raised pre-load)."""


def _first_of(raw: Mapping[str, Any], *names: str) -> Any:
    return next((raw[n] for n in names if n in raw), None)


def normalize_quant_config(raw: Mapping[str, Any]) -> dict[str, Any]:
    """Return ``{"quant_method": "awq", "bits": 4, "group_size": 128, "zero_point": True,
    "version": "gemm"}`` or raise :class:`UnsupportedQuantizationConfigError`."""
    method = raw.get("quant_method")
    if method != "awq":
        raise UnsupportedQuantizationConfigError(
            f"quant_method={method!r} is not handled by the AWQ preflight; expected 'awq'.")
    bits = _first_of(raw, "bits", "w_bit")
    if bits is None:
        raise UnsupportedQuantizationConfigError("missing 'bits' (or alias 'w_bit') in quantization_config")
    if bits != SUPPORTED["bits"]:
        raise UnsupportedQuantizationConfigError(f"bits={bits!r} is not supported; only bits=4 is")
    group = _first_of(raw, "group_size", "q_group_size")
    if group is None:
        raise UnsupportedQuantizationConfigError(
            "missing 'group_size' (or alias 'q_group_size') in quantization_config")
    if group != SUPPORTED["group_size"]:
        raise UnsupportedQuantizationConfigError(f"group_size={group!r} is not supported; only group_size=128 is")
    zero_point = raw.get("zero_point", True)
    if zero_point is not True:
        raise UnsupportedQuantizationConfigError(
            f"symmetric quantization is not supported; zero_point=true is required, got zero_point={zero_point!r}.")
    version = raw.get("version", "gemm")
    if not isinstance(version, str):
        raise UnsupportedQuantizationConfigError(f"version={version!r} must be a string")
    if version.lower() == "gemv":
        raise UnsupportedQuantizationConfigError(
            "version='gemv' is not supported (different intra-int32 packing); only 'gemm' is")
    if version.lower() != "gemm":
        raise UnsupportedQuantizationConfigError(
            f"version={version!r} is not supported; only 'gemm' (case-insensitive) is")
    return {"quant_method": "awq", **SUPPORTED}
