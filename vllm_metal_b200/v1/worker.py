"""vLLM worker for the B200 decode path (boundary B2).

Mirrors the method set of ``MetalWorker(WorkerBase)`` (vllm_metal/v1/worker.py:65-404): device
init, model load, KV-cache sizing/spec/initialisation, ``execute_model`` -> ``sample_tokens``.
The engine-side scheduler drives it; everything on the device goes through
``v1/model_runner.py::PagedModelRunner`` -> ``sdpa_forward`` -> the C-ABI kernels.  Only the
greedy decode + chunked-prefill contract of the hot path is implemented: LoRA, multimodal,
speculative drafts, sleep mode and pipeline parallelism raise or no-op exactly where the
reference's worker has its own gates (DESIGN.md "out of scope").
"""

from __future__ import annotations

import logging
import time
from typing import Any

import torch

from vllm.v1.kv_cache_interface import FullAttentionSpec, KVCacheConfig, KVCacheSpec, SlidingWindowSpec
from vllm.v1.worker.worker_base import WorkerBase

from ..attention.caches.kv_cache import B200PagedKVCache
from ..lm.decoder import PRESETS, Model, ModelArgs
from .model_runner import PagedModelRunner

logger = logging.getLogger(__name__)


def model_args_from_hf_config(hf_config: Any, quant: dict | None = None) -> ModelArgs:
    """Map a HF/mlx-lm ``config.json`` onto the decoder's ``ModelArgs`` (weights themselves are
    loaded or random-initialised by ``Model``).  ``quant`` is the ``quantization`` dict."""
    import dataclasses

    q = quant or getattr(hf_config, "quantization", None) or {}
    n_heads = int(hf_config.num_attention_heads)
    hidden = int(hf_config.hidden_size)
    fields = dict(
        num_layers=int(hf_config.num_hidden_layers), hidden=hidden, n_heads=n_heads,
        n_kv_heads=int(getattr(hf_config, "num_key_value_heads", n_heads)),
        head_dim=int(getattr(hf_config, "head_dim", None) or hidden // n_heads),
        intermediate=int(hf_config.intermediate_size), vocab=int(hf_config.vocab_size),
        rms_eps=float(getattr(hf_config, "rms_norm_eps", 1e-6)),
        rope_base=float(getattr(hf_config, "rope_theta", 10000.0)),
        tie_embeddings=bool(getattr(hf_config, "tie_word_embeddings", False)),
        qk_norm=str(getattr(hf_config, "model_type", "")).startswith("qwen3"),
        sliding_window=int(getattr(hf_config, "sliding_window", None) or -1),
        layer_types=list(getattr(hf_config, "layer_types", None) or []),
        name=str(getattr(hf_config, "model_type", "hf")),
        bits=int(q.get("bits", 4)), group_size=int(q.get("group_size", 64)),
    )
    known = {f.name for f in dataclasses.fields(ModelArgs)}
    return dataclasses.replace(PRESETS["tiny"], **{k: v for k, v in fields.items() if k in known})


class B200Worker(WorkerBase):
    """One process per GPU; ``rank`` is the tensor-parallel rank."""

    def __init__(self, vllm_config, local_rank: int, rank: int, distributed_init_method: str,
                 is_driver_worker: bool = False, **kwargs: Any) -> None:
        super().__init__(vllm_config=vllm_config, local_rank=local_rank, rank=rank,
                         distributed_init_method=distributed_init_method, is_driver_worker=is_driver_worker)
        self.parallel_config.disable_custom_all_reduce = True  # the plugin's own NVLink kernel is used
        self.model: Model | None = None
        self.kv_cache: B200PagedKVCache | None = None
        self.model_runner: PagedModelRunner | None = None
        self._allreduce = None
        self._pending: list[int] | None = None

    # ---- device / distributed (worker.py:117-193) -----------------------------
    def init_device(self) -> None:
        import torch.distributed as dist

        torch.cuda.set_device(self.local_rank)
        self.device = torch.device("cuda", self.local_rank)
        world = self.parallel_config.tensor_parallel_size
        if world > 1:
            if not dist.is_initialized():
                dist.init_process_group("nccl", init_method=self.distributed_init_method, rank=self.rank,
                                        world_size=world, device_id=self.device)
            from ..distributed.nvlink_allreduce import NVLinkAllReduce

            hidden = int(self.model_config.hf_config.hidden_size)
            max_tokens = int(self.scheduler_config.max_num_batched_tokens)
            self._allreduce = NVLinkAllReduce(device=self.device, max_bytes=min(max_tokens, 256) * hidden * 2)

    def load_model(self) -> None:
        args = model_args_from_hf_config(self.model_config.hf_config)
        world = self.parallel_config.tensor_parallel_size
        self.model = Model(args, torch.bfloat16, self.device, seed=int(self.model_config.seed or 0),
                           rank=self.rank, world=world, allreduce=self._allreduce)

    def get_model(self) -> Any:
        return self.model

    # ---- KV cache (worker.py:199-250; cache_policy.py:383-411) ----------------
    def get_cache_block_size_bytes(self) -> int:
        a = self.model.args
        world = self.parallel_config.tensor_parallel_size
        return 2 * a.num_layers * self.cache_config.block_size * (a.n_kv_heads // world) * a.head_dim * 2

    def determine_available_memory(self) -> int:
        free, total = torch.cuda.mem_get_info(self.device)
        budget = int(total * self.cache_config.gpu_memory_utilization) - (total - free)
        return max(budget, 0)

    def get_kv_cache_spec(self) -> dict[str, KVCacheSpec]:
        a = self.model.args
        world = self.parallel_config.tensor_parallel_size
        windows = a.sliding_window_per_layer()
        specs: dict[str, KVCacheSpec] = {}
        for i in range(a.num_layers):  # layer names as the reference: layers.{i}.self_attn
            common = dict(block_size=self.cache_config.block_size, num_kv_heads=a.n_kv_heads // world,
                          head_size=a.head_dim, dtype=torch.bfloat16)
            w = windows[i] if windows else -1
            specs[f"layers.{i}.self_attn"] = (SlidingWindowSpec(sliding_window=w, **common) if w and w > 0
                                               else FullAttentionSpec(**common))
        return specs

    def initialize_cache(self, num_gpu_blocks: int, num_cpu_blocks: int) -> None:
        self.cache_config.num_gpu_blocks = num_gpu_blocks
        self.cache_config.num_cpu_blocks = num_cpu_blocks

    def initialize_from_config(self, kv_cache_config: KVCacheConfig) -> None:
        a = self.model.args
        world = self.parallel_config.tensor_parallel_size
        self.kv_cache = B200PagedKVCache(a.num_layers, a.n_kv_heads // world, a.head_dim,
                                         int(kv_cache_config.num_blocks), self.cache_config.block_size,
                                         torch.bfloat16, self.device,
                                         sliding_window_per_layer=a.sliding_window_per_layer())
        self.model_runner = PagedModelRunner(self.model, self.kv_cache,
                                             max_batch=int(self.scheduler_config.max_num_seqs),
                                             max_model_len=int(self.model_config.max_model_len),
                                             use_cuda_graph=not self.model_config.enforce_eager)

    def compile_or_warm_up_model(self) -> float:
        start = time.perf_counter()
        torch.cuda.synchronize(self.device)
        return time.perf_counter() - start

    # ---- step (worker.py:267-285; model_runner.py:1138-1410) -------------------
    def execute_model(self, scheduler_output: Any) -> None:
        """Pack the scheduled requests (decode rows first, then prefill chunks -- the reference's
        order, model_runner.py batch packing) and run the step; tokens are handed out by
        ``sample_tokens`` (deferred sampling contract)."""
        tokens, decodes, prefills = self._pack(scheduler_output)
        self._pending = self.model_runner.execute(tokens, decodes, prefills) if tokens else []

    def sample_tokens(self, grammar_output: Any = None) -> list[int]:
        out, self._pending = self._pending or [], None
        return out

    def _pack(self, so: Any):
        """``scheduler_output`` duck type used here: ``.decodes`` = [(token, block_ids, seq_len)],
        ``.prefills`` = [(tokens, block_ids, start_pos)] -- the per-step facts the reference's
        runner extracts from vLLM's SchedulerOutput before ``prepare_grouped``."""
        tokens: list[int] = []
        decodes, prefills = [], []
        for tok, blocks, seq_len in getattr(so, "decodes", []):
            tokens.append(int(tok))
            decodes.append((list(blocks), int(seq_len)))
        for toks, blocks, start in getattr(so, "prefills", []):
            tokens.extend(int(t) for t in toks)
            prefills.append((list(blocks), len(toks), int(start)))
        return tokens, decodes, prefills

    # ---- gates the reference worker also has (worker.py:259-356) ---------------
    def reset_mm_cache(self) -> None:
        return None

    def add_lora(self, lora_request: Any) -> bool:
        raise NotImplementedError("LoRA is outside the B200 decode hot path")

    def remove_lora(self, lora_id: int) -> bool:
        raise NotImplementedError("LoRA is outside the B200 decode hot path")

    def pin_lora(self, lora_id: int) -> bool:
        raise NotImplementedError("LoRA is outside the B200 decode hot path")

    def list_loras(self) -> set[int]:
        return set()

    def check_health(self) -> None:
        torch.cuda.synchronize(self.device)

    def shutdown(self) -> None:
        if self._allreduce is not None:
            self._allreduce.close()
            self._allreduce = None
