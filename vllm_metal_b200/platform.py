"""vLLM platform plugin for the B200 decode path (boundary B1).

Mirrors the hooks ``vllm_metal/platform.py:25-1022`` overrides -- device queries, request
validation, config gates, worker class, block-size hook, synthetic attention backend -- with the
device being ``cuda:<local_rank>`` of compute capability 10.x instead of the single Apple GPU.
Only host-side glue: no kernels are launched from here.  What the reference's platform does for
Ray / MLX wired memory / hybrid-model TurboQuant realignment is outside the hot path and is not
reproduced (DESIGN.md "out of scope").
"""

from __future__ import annotations

import logging
from typing import TYPE_CHECKING

import torch
from vllm.platforms.interface import Platform, PlatformEnum

if TYPE_CHECKING:
    from vllm.config import VllmConfig

logger = logging.getLogger(__name__)

WORKER_CLS = "vllm_metal_b200.v1.worker.B200Worker"


class B200Platform(Platform):
    """Out-of-tree platform (reference: ``MetalPlatform``, platform.py:25-56)."""

    _enum = PlatformEnum.OOT
    device_name: str = "cuda"
    device_type: str = "cuda"
    dispatch_key: str = "CUDA"
    ray_device_key: str = "GPU"
    device_control_env_var: str = "CUDA_VISIBLE_DEVICES"
    # what the quant-method hook accepts: MLX-style affine group quantisation (+ AWQ repacked to it)
    supported_quantization: list[str] = ["mlx_affine", "awq"]

    # ---- device queries (platform.py:58-200) ---------------------------------
    @classmethod
    def is_available(cls) -> bool:
        try:
            return torch.cuda.is_available() and torch.cuda.get_device_capability(0)[0] == 10
        except Exception:  # noqa: BLE001
            return False

    @classmethod
    def get_device_name(cls, device_id: int = 0) -> str:
        return torch.cuda.get_device_name(device_id) if torch.cuda.is_available() else "B200 (no CUDA device)"

    @classmethod
    def get_device_total_memory(cls, device_id: int = 0) -> int:
        return int(torch.cuda.get_device_properties(device_id).total_memory)

    @classmethod
    def get_device_available_memory(cls, device_id: int = 0) -> int:
        free, _ = torch.cuda.mem_get_info(device_id)
        return int(free)

    @classmethod
    def get_device_count(cls) -> int:
        return torch.cuda.device_count()

    @classmethod
    def set_device(cls, device: torch.device | int | None = 0) -> None:
        torch.cuda.set_device(0 if device is None else device)

    @classmethod
    def current_device(cls) -> int:
        return torch.cuda.current_device()

    @classmethod
    def synchronize(cls, device_id: int = 0) -> None:
        torch.cuda.synchronize(device_id)

    @classmethod
    def manual_seed_all(cls, seed: int) -> None:
        torch.cuda.manual_seed_all(seed)

    @classmethod
    def get_torch_device(cls, device_id: int = 0) -> torch.device:
        return torch.device("cuda", device_id)

    @classmethod
    def is_pin_memory_available(cls) -> bool:
        return True  # the runner stages every step's metadata through pinned buffers

    @classmethod
    def get_device_communicator_cls(cls) -> str:
        return "vllm.distributed.device_communicators.cuda_communicator.CudaCommunicator"

    # ---- request validation (platform.py:209-235) ----------------------------
    @classmethod
    def validate_request(cls, processed_inputs: object, params: object) -> None:
        """Reject sampling controls the greedy/top-1 device sampler cannot apply (same three the
        reference rejects: they need vLLM logits processors)."""
        from vllm.sampling_params import SamplingParams

        if not isinstance(params, SamplingParams):
            return
        bad = [name for name, on in (("min_p", params.min_p > 0.0), ("logit_bias", bool(params.logit_bias)),
                                     ("min_tokens", params.min_tokens > 0)) if on]
        if bad:
            try:
                from vllm.exceptions import VLLMValidationError as _Err

                raise _Err("vllm-metal-b200 does not support sampling controls backed by vLLM logits "
                           f"processors ({', '.join(bad)}).", parameter=bad[0])
            except ImportError:
                raise ValueError("vllm-metal-b200 does not support sampling controls backed by vLLM "
                                 f"logits processors ({', '.join(bad)}).") from None

    # ---- config gates (platform.py:342-797, the parts the hot path needs) -----
    @classmethod
    def check_and_update_config(cls, vllm_config: "VllmConfig") -> None:
        parallel = vllm_config.parallel_config
        if parallel.worker_cls in ("auto", None):
            parallel.worker_cls = WORKER_CLS
        if getattr(parallel, "pipeline_parallel_size", 1) != 1:
            raise ValueError("vllm-metal-b200 shards by tensor parallelism only (pipeline_parallel_size must be 1)")
        tp = getattr(parallel, "tensor_parallel_size", 1)
        if tp not in (1, 2, 4, 8):
            raise ValueError(f"tensor_parallel_size must be 1, 2, 4 or 8 on one 8xB200 box, got {tp}")
        # the plugin's own all-reduce runs after o_proj / down_proj; vLLM's custom one is unused
        parallel.disable_custom_all_reduce = True
        cache = vllm_config.cache_config
        if cache is not None and getattr(cache, "block_size", None) is None:
            cache.block_size = 16
        comp = getattr(vllm_config, "compilation_config", None)
        if comp is not None and hasattr(comp, "custom_ops") and not comp.custom_ops:
            comp.custom_ops = ["none"]  # no torch.compile / inductor on this path (CUDA graphs are ours)

    @classmethod
    def support_hybrid_kv_cache(cls) -> bool:
        return True

    # ---- block size / attention backend (platform.py:803-1006) ----------------
    @classmethod
    def _find_non_ssm_backend(cls, vllm_config: "VllmConfig"):
        from .attention.synthetic_backend import B200Backend

        return B200Backend

    @classmethod
    def update_block_size_for_backend(cls, vllm_config: "VllmConfig") -> None:
        """Default to 16-token pages and keep user sizes that are multiples of 16; other sizes
        are served through block-size translation (attention/impls/sdpa.py)."""
        cache = vllm_config.cache_config
        if cache is None:
            return
        if not getattr(cache, "user_specified_block_size", False) or cache.block_size is None:
            cache.block_size = 16
        elif cache.block_size % 8 != 0:
            raise ValueError(f"block_size {cache.block_size} has no kernel block size in (32, 16, 8) dividing it")

    @classmethod
    def get_attn_backend_cls(cls, selected_backend, attn_selector_config, num_heads: int | None = None) -> str:
        if getattr(attn_selector_config, "use_sparse", False):
            raise NotImplementedError("Sparse attention is not supported by vllm-metal-b200.")
        if selected_backend:
            logger.info("Ignoring %s: attention is handled by the plugin's model runner.", selected_backend)
        return "vllm_metal_b200.attention.synthetic_backend.B200Backend"
