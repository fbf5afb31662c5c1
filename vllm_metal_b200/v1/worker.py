"""vLLM worker for the B200 decode path (boundary B2).

Mirrors the method set of ``MetalWorker(WorkerBase)`` (vllm_metal/v1/worker.py:65-404): device
init, model load, KV-cache sizing/spec/initialisation, ``execute_model`` -> ``sample_tokens``.
The engine-side scheduler drives it; everything on the device goes through
``v1/model_runner.py::PagedModelRunner`` -> ``sdpa_forward`` -> the C-ABI kernels.  Only the
greedy decode + chunked-prefill contract of the hot path is implemented, plus greedy verification of
scheduled draft tokens in the decode kernel's window mode with a prompt-lookup drafter
(``v1/spec_decode.py``): LoRA, multimodal, model-based drafters, sleep mode and pipeline parallelism
raise or no-op exactly where the reference's worker has its own gates (DESIGN.md "out of scope").
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


class _RequestState:
    """What the runner keeps per live request (reference RequestState, model_runner.py:1150-1200): every known
    token id (prompt + generated), one block-id list per KV-cache group and how many tokens are already cached."""

    __slots__ = ("tokens", "block_ids", "num_computed")

    def __init__(self, tokens: list[int], block_ids: list[int], num_computed: int) -> None:
        self.tokens, self.block_ids, self.num_computed = tokens, block_ids, num_computed


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
        self._pending = None
        self._requests: dict[str, _RequestState] = {}
        self._drafter = self._build_drafter(getattr(vllm_config, "speculative_config", None))
        self._drafts: dict[str, list[int]] = {}
        # one-step-ahead decode pipelining (reference v1/decode_pipeline.py): on with vLLM's async scheduling, or forced
        # by B200_DECODE_PIPELINE=1
        import os

        from .decode_pipeline import DecodePipeline

        want = os.environ.get("B200_DECODE_PIPELINE")
        enabled = (want == "1") if want is not None else bool(getattr(self.scheduler_config, "async_scheduling", False))
        self._pipeline = DecodePipeline() if enabled else None

    @staticmethod
    def _build_drafter(spec: Any):
        """Reference model_runner.py:905-934: Gemma4 MTP / draft_model / ngram.  The n-gram (prompt-lookup) drafter is
        host-only; model-based drafters are outside the decode hot path and raise."""
        if spec is None:
            return None
        from .spec_decode import NgramDrafter

        if getattr(spec, "method", None) != "ngram":
            raise NotImplementedError(
                f"speculative method {getattr(spec, 'method', None)!r} is outside the B200 decode hot path (supported: ngram)")
        return NgramDrafter(int(spec.num_speculative_tokens), int(getattr(spec, "prompt_lookup_min", None) or 2),
                            int(getattr(spec, "prompt_lookup_max", None) or 5))

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
        """Builds the decoder for the HF config with RANDOM-INITIALISED affine-quantized weights (``seed``):
        the BASELINE configs are all random-init and no checkpoint reader is wired (the AWQ repack path is
        ``quant/awq_loader.py``; safetensors loading is out of scope).  ``load_format`` other than
        "dummy"/"auto" therefore raises instead of silently ignoring a checkpoint."""
        lf = str(getattr(getattr(self.vllm_config, "load_config", None), "load_format", "dummy"))
        if lf not in ("dummy", "auto", "LoadFormat.DUMMY", "LoadFormat.AUTO"):
            raise NotImplementedError(f"load_format={lf}: only random-init ('dummy') weights are wired on this path")
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
        """Bytes the KV cache may take: the utilisation budget minus everything resident NOW -- called after
        ``load_model`` (vLLM's order), so the weights are inside ``total - free``; the stream-K workspace and
        graph pools allocated later are covered by a fixed 1 GiB reserve."""
        if self.model is None:
            raise RuntimeError("determine_available_memory must follow load_model (weights are not resident yet)")
        torch.cuda.synchronize(self.device)
        free, total = torch.cuda.mem_get_info(self.device)
        budget = int(total * self.cache_config.gpu_memory_utilization) - (total - free) - (1 << 30)
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
        groups = getattr(kv_cache_config, "kv_cache_groups", None) or []
        self._num_kv_groups = max(1, len(groups))
        group_sizes = None
        if len(groups) > 1:
            # vLLM grouped the layers (full + sliding-window specs): adopt its layout as is -- per-layer page size,
            # window and the scheduler group whose block ids the layer reads (reference adopt_layout, runtime/mha.py:74)
            from ..attention.caches.mha_layout import MHAKVCacheLayout

            names = tuple(f"layers.{i}.self_attn" for i in range(a.num_layers))
            layout = MHAKVCacheLayout.from_config(kv_cache_config, names)
            self.kv_cache = B200PagedKVCache.from_layout(layout, torch.bfloat16, self.device)
            group_sizes = layout.group_block_sizes
        else:
            self.kv_cache = B200PagedKVCache(a.num_layers, a.n_kv_heads // world, a.head_dim,
                                             int(kv_cache_config.num_blocks), self.cache_config.block_size,
                                             torch.bfloat16, self.device,
                                             sliding_window_per_layer=a.sliding_window_per_layer())
        self.model_runner = PagedModelRunner(self.model, self.kv_cache,
                                             max_batch=int(self.scheduler_config.max_num_seqs),
                                             max_model_len=int(self.model_config.max_model_len),
                                             use_cuda_graph=not self.model_config.enforce_eager and group_sizes is None)
        self.model_runner.group_block_sizes = group_sizes

    def compile_or_warm_up_model(self) -> float:
        """Capture the decode CUDA graphs the first real steps would otherwise pay for (reference
        ``compile_or_warm_up_model`` returns CompilationTimes, worker.py:252-257; vLLM 0.22 takes the
        elapsed seconds): one dummy decode step per batch bucket on the padding block."""
        start = time.perf_counter()
        r = self.model_runner
        if r is not None and r.use_cuda_graph:
            for b in sorted({1, min(8, r.max_batch), r.max_batch}):
                r.decode_step([0] * b, [[0]] * b, [1] * b)  # block 0 is the reserved padding block
        torch.cuda.synchronize(self.device)
        return time.perf_counter() - start

    # ---- step (worker.py:267-285; model_runner.py:1138-1410, 2733-2783) ----------
    def execute_model(self, scheduler_output: Any) -> None:
        """Consume vLLM's ``SchedulerOutput``: update the per-request state (reference
        ``_start_paged_forward``, model_runner.py:1138-1240), pack decode rows first, then prefill
        chunks (the reference's packing order, attention/context.py:198-283), and run the step.
        Returns ``None``: tokens are handed out by ``sample_tokens`` (the deferred-sampling contract of
        the reference, model_runner.py:2588-2596)."""
        so = scheduler_output
        for field in ("scheduled_new_reqs", "scheduled_cached_reqs", "num_scheduled_tokens", "finished_req_ids"):
            if not hasattr(so, field):
                raise NotImplementedError(
                    f"execute_model expects vllm.v1.core.sched.output.SchedulerOutput; {type(so).__name__} has no "
                    f"'{field}'")
        spec_tokens = dict(so.scheduled_spec_decode_tokens or {})  # req_id -> scheduled draft token ids
        for rid in list(so.finished_req_ids) + list(so.preempted_req_ids or ()):
            self._requests.pop(rid, None)
        for nr in so.scheduled_new_reqs:
            if nr.prompt_token_ids is None:
                raise NotImplementedError("prompt embeddings are outside the B200 decode hot path")
            if len(nr.block_ids) != getattr(self, "_num_kv_groups", 1):
                raise RuntimeError(f"request {nr.req_id}: {len(nr.block_ids)} block-id lists for "
                                   f"{getattr(self, '_num_kv_groups', 1)} KV-cache groups")
            self._requests[nr.req_id] = _RequestState(list(nr.prompt_token_ids), [list(b) for b in nr.block_ids],
                                                      int(nr.num_computed_tokens))
        cached = so.scheduled_cached_reqs
        for i, rid in enumerate(cached.req_ids):
            st = self._requests[rid]
            st.num_computed = int(cached.num_computed_tokens[i])
            new_blocks = cached.new_block_ids[i]
            if rid in cached.resumed_req_ids:
                st.block_ids = [list(b) for b in new_blocks] if new_blocks is not None else [[] for _ in st.block_ids]
                if rid in cached.all_token_ids:
                    st.tokens = list(cached.all_token_ids[rid])
            elif new_blocks is not None:
                for have, more in zip(st.block_ids, new_blocks, strict=True):
                    have.extend(more)
            if cached.new_token_ids and cached.new_token_ids[i]:  # pipeline-parallel peers only; we sample here
                st.tokens = st.tokens[:st.num_computed] + list(cached.new_token_ids[i])
        decodes, prefills = [], []  # (req_id, state, n)
        windows = False
        for rid, n in so.num_scheduled_tokens.items():
            st = self._requests[rid]
            drafts = spec_tokens.get(rid) or []
            if drafts:
                # a verification window: the last committed token + the scheduled drafts (model_runner.py:1200-1209)
                if st.num_computed + 1 != len(st.tokens) or n != 1 + len(drafts):
                    raise RuntimeError(f"request {rid}: {len(drafts)} draft tokens scheduled as {n} rows from "
                                       f"{st.num_computed} of {len(st.tokens)} known tokens")
                windows = True
                decodes.append((rid, st, int(n)))
                continue
            if st.num_computed + n > len(st.tokens):
                raise RuntimeError(f"request {rid}: scheduled {n} tokens from {st.num_computed} but only "
                                   f"{len(st.tokens)} are known")
            (decodes if n == 1 and st.num_computed > 0 else prefills).append((rid, st, int(n)))
        tokens: list[int] = []
        dec_rows, pre_rows = [], []
        grouped = getattr(self.model_runner, "group_block_sizes", None) is not None
        for rid, st, n in decodes:
            tokens.append(st.tokens[st.num_computed])
            tokens.extend(spec_tokens.get(rid) or [])
            blocks = st.block_ids if grouped else st.block_ids[0]
            dec_rows.append((blocks, st.num_computed, n) if windows else (blocks, st.num_computed))
        for _, st, n in prefills:
            tokens.extend(st.tokens[st.num_computed:st.num_computed + n])
            pre_rows.append((st.block_ids if grouped else st.block_ids[0], n, st.num_computed))
        order = decodes + prefills
        r = self.model_runner
        pipe = getattr(self, "_pipeline", None)
        if pipe is not None:
            reason = None if not tokens else pipe.block_reason(
                so, has_prefill_rows=bool(pre_rows), has_spec_tokens=windows, runner=r, batch=len(dec_rows),
                drafter=getattr(self, "_drafter", None))
            if tokens and reason is None:
                # deferred step: input tokens the host has not seen yet are gathered from the pending step's argmax on
                # the device; nothing here waits for the GPU
                rows = [pipe.pending_row(rid) if st.tokens[st.num_computed] < 0 else -1 for rid, st, _ in decodes]
                if any(row < 0 and st.tokens[st.num_computed] < 0 for row, (_, st, _) in zip(rows, decodes, strict=True)):
                    raise RuntimeError("decode pipeline: a placeholder token has no pending row")
                handle = r.submit_decode(tokens, [b for b, _ in dec_rows], [n for _, n in dec_rows], rows)
                self._pending = (order, handle, "deferred")
                return
            pipe.begin_step()  # synchronous step: the pending tokens land before any state is read
            tokens = self._retokenize(decodes, prefills, spec_tokens)
        if not tokens:
            sampled = []
        elif windows and not pre_rows:
            # pure verification step: window-mode layout, the rows of a window share every KV read
            sampled = r.verify(tokens, dec_rows)
        elif windows:
            # any prefill segment keeps the batch on the expanded layout (attention/context.py:276-280): one segment,
            # hence one greedy token, per window row
            flat = r.execute(tokens, dec_rows, pre_rows)
            sampled, at = [], 0
            for _, _, n in decodes:
                sampled.append(flat[at:at + n])
                at += n
            sampled.extend(flat[at:])
        elif not pre_rows and r.use_cuda_graph and len(dec_rows) <= r.max_batch:
            # pure decode: the CUDA-graph path (pinned staging -> one H2D -> graph replay -> D2H token ids)
            sampled = r.decode_step(tokens, [b for b, _ in dec_rows], [n for _, n in dec_rows])
        else:
            sampled = r.execute(tokens, dec_rows, pre_rows)
        self._pending = (order, sampled, spec_tokens)

    def sample_tokens(self, grammar_output: Any = None):
        """Greedy tokens of the step as a ``ModelRunnerOutput`` (reference ``_build_output``,
        model_runner.py:2450-2457).  A prefill chunk that does not reach the end of the known tokens
        samples nothing (``[]``), exactly as vLLM expects for partial prefills."""
        from vllm.v1.outputs import ModelRunnerOutput

        if grammar_output is not None:
            raise NotImplementedError("structured output (grammar bitmask) is outside the B200 decode hot path")
        if self._pending is None:
            return None
        (order, sampled, spec_tokens), self._pending = self._pending, None
        if spec_tokens == "deferred":
            # value-independent bookkeeping now, token values when the engine asks (decode_pipeline.py invariants)
            from .decode_pipeline import PENDING_TOKEN_PLACEHOLDER

            entries = []
            for row, (rid, st, n) in enumerate(order):
                st.tokens.append(PENDING_TOKEN_PLACEHOLDER)
                st.num_computed += n
                entries.append((rid, st, len(st.tokens) - 1, row))
            self._drafts = {}
            return self._pipeline.submit(sampled, entries)
        from .spec_decode import verify_greedy

        req_ids, out = [], []
        self._drafts = {}
        for (rid, st, n), tok in zip(order, sampled, strict=True):
            req_ids.append(rid)
            drafts = spec_tokens.get(rid) or []
            if drafts:
                # accepted drafts + the target's own token at the first mismatch (or the bonus token); the KV rows of
                # rejected drafts are overwritten when real tokens reach those positions
                emitted = verify_greedy(tok, drafts)
                st.tokens.extend(emitted)
                st.num_computed += len(emitted)
                out.append(emitted)
            elif st.num_computed + n >= len(st.tokens):  # reached the last known token: this step samples
                t = int(tok[-1]) if isinstance(tok, (list, tuple)) else int(tok)
                st.tokens.append(t)
                out.append([t])
                st.num_computed += n
            else:
                out.append([])
                st.num_computed += n
            drafter = getattr(self, "_drafter", None)
            if drafter is not None and out[-1]:
                max_len = int(getattr(getattr(self, "model_config", None), "max_model_len", 1 << 30))
                self._drafts[rid] = drafter.propose(st.tokens, max_len - len(st.tokens) - 1)
        return ModelRunnerOutput(req_ids=req_ids, req_id_to_index={r: i for i, r in enumerate(req_ids)},
                                 sampled_token_ids=out, logprobs=None, prompt_logprobs_dict={}, pooler_output=[])

    def _retokenize(self, decodes, prefills, spec_tokens) -> list[int]:
        """The packed input ids of a step, rebuilt after ``begin_step`` replaced placeholders by real tokens."""
        tokens: list[int] = []
        for rid, st, _n in decodes:
            tokens.append(st.tokens[st.num_computed])
            tokens.extend(spec_tokens.get(rid) or [])
        for _, st, n in prefills:
            tokens.extend(st.tokens[st.num_computed:st.num_computed + n])
        return tokens

    def take_draft_token_ids(self):
        """Drafts for the requests that sampled last step (reference worker.py:286-288 -> model_runner.py:730); ``None``
        without a drafter.  vLLM schedules them as ``scheduled_spec_decode_tokens`` of the next step."""
        if getattr(self, "_drafter", None) is None or not getattr(self, "_drafts", None):
            return None
        from vllm.v1.outputs import DraftTokenIds

        drafts, self._drafts = self._drafts, {}
        return DraftTokenIds(req_ids=list(drafts), draft_token_ids=[drafts[r] for r in drafts])

    def get_supported_tasks(self) -> tuple[str, ...]:
        return ("generate",)

    def profile(self, is_start: bool = True, profile_prefix: str | None = None) -> None:
        """``cudaProfilerStart/Stop`` + an NVTX range, the CUDA analogue of the reference's Metal capture
        hook (worker.py:358-390): `ncu/nsys --capture-range=cudaProfilerApi` then record exactly the
        steps between the two calls."""
        if is_start:
            torch.cuda.profiler.start()
            torch.cuda.nvtx.range_push(profile_prefix or "b200_worker")
            self._profiling = True
        elif getattr(self, "_profiling", False):
            torch.cuda.nvtx.range_pop()
            torch.cuda.synchronize(self.device)
            torch.cuda.profiler.stop()
            self._profiling = False

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
