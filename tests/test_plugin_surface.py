"""Host-side plugin surface (B1 platform, B2 worker, quant-method hook) on CPU.

These mirror what the reference's own tests check about its plugin layer without a device
(tests/test_platform.py, tests/test_block_size_translation.py): entry points resolve, the
synthetic backend advertises MultipleOf(16), the marker quant config registers, the worker
produces the reference's layer names and per-layer spec types.  No kernels are called.
"""

from __future__ import annotations

import types

import pytest
import torch

vllm = pytest.importorskip("vllm")


def test_register_entry_points_resolve():
    import vllm_metal_b200 as pkg

    assert pkg.PLATFORM_CLS == "vllm_metal_b200.platform.B200Platform"
    # no sm_100 device in the CPU container -> vLLM must fall through to its other platforms
    assert pkg.register() in (None, pkg.PLATFORM_CLS)
    from vllm_metal_b200.platform import WORKER_CLS, B200Platform

    mod, _, cls = WORKER_CLS.rpartition(".")
    assert hasattr(__import__(mod, fromlist=[cls]), cls)
    assert B200Platform.device_type == "cuda" and B200Platform.dispatch_key == "CUDA"
    assert B200Platform.get_attn_backend_cls(None, types.SimpleNamespace(use_sparse=False, use_mla=False)) \
        == "vllm_metal_b200.attention.synthetic_backend.B200Backend"
    with pytest.raises(NotImplementedError):
        B200Platform.get_attn_backend_cls(None, types.SimpleNamespace(use_sparse=True))


def test_pyproject_declares_the_reference_entry_point_groups(monkeypatch):
    """vLLM finds the plugin through the two groups the reference declares (pyproject.toml:104-112); each
    target must resolve to a callable, and register() stays inert without the operator's opt-in."""
    import importlib
    import os
    import tomllib

    import vllm_metal_b200 as pkg

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    with open(os.path.join(root, "pyproject.toml"), "rb") as f:
        eps = tomllib.load(f)["project"]["entry-points"]
    assert set(eps) == {"vllm.platform_plugins", "vllm.general_plugins"}
    for group in eps.values():
        for target in group.values():
            mod, _, fn = target.partition(":")
            assert callable(getattr(importlib.import_module(mod), fn))
    monkeypatch.delenv("VLLM_B200_PLUGIN", raising=False)
    assert pkg.register() is None
    monkeypatch.setenv("VLLM_B200_PLUGIN", "1")
    monkeypatch.setattr(pkg, "_cuda_sm100_available", lambda: True)
    assert pkg.register() == pkg.PLATFORM_CLS


def test_synthetic_backend_advertises_multiple_of_16_and_refuses_dispatch():
    from vllm.v1.attention.backend import MultipleOf

    from vllm_metal_b200.attention.synthetic_backend import B200Backend

    sizes = B200Backend.get_supported_kernel_block_sizes()
    assert len(sizes) == 1 and isinstance(sizes[0], MultipleOf) and sizes[0].base == 16
    for fn in (B200Backend.get_impl_cls, B200Backend.get_builder_cls, B200Backend.get_kv_cache_shape):
        with pytest.raises(NotImplementedError):
            fn()


def test_platform_config_gates():
    from vllm_metal_b200.platform import WORKER_CLS, B200Platform

    def cfg(tp=1, pp=1, block=None, user=False):
        return types.SimpleNamespace(
            parallel_config=types.SimpleNamespace(worker_cls="auto", tensor_parallel_size=tp,
                                                  pipeline_parallel_size=pp, disable_custom_all_reduce=False),
            cache_config=types.SimpleNamespace(block_size=block, user_specified_block_size=user),
            compilation_config=None)

    c = cfg(tp=2)
    B200Platform.check_and_update_config(c)
    assert c.parallel_config.worker_cls == WORKER_CLS and c.parallel_config.disable_custom_all_reduce
    assert c.cache_config.block_size == 16
    with pytest.raises(ValueError):
        B200Platform.check_and_update_config(cfg(pp=2))
    with pytest.raises(ValueError):
        B200Platform.check_and_update_config(cfg(tp=3))
    c = cfg(block=544, user=True)  # the reference's hybrid example: translated 544 -> 32 x 17
    B200Platform.update_block_size_for_backend(c)
    assert c.cache_config.block_size == 544
    with pytest.raises(ValueError):
        B200Platform.update_block_size_for_backend(cfg(block=7, user=True))


def test_validate_request_rejects_logits_processor_controls():
    from vllm.sampling_params import SamplingParams

    from vllm_metal_b200.platform import B200Platform

    B200Platform.validate_request(None, SamplingParams(temperature=0.0))
    for kw in (dict(min_p=0.1), dict(logit_bias={1: 2.0}), dict(min_tokens=3, max_tokens=8)):
        with pytest.raises(Exception, match="logits"):
            B200Platform.validate_request(None, SamplingParams(**kw))


def test_marker_quant_config_registers_idempotently():
    import vllm_metal_b200 as pkg
    from vllm.model_executor.layers.quantization import get_quantization_config

    pkg.register_quant_config()
    pkg.register_quant_config()
    cls = get_quantization_config("mlx_affine")
    c = cls.from_config({"quantization": {"bits": 8, "group_size": 32}})
    assert (c.bits, c.group_size) == (8, 32) and cls.get_name() == "mlx_affine"
    with pytest.raises(ValueError):
        cls(bits=3)
    with pytest.raises(ValueError):
        cls(group_size=48)


def test_worker_kv_cache_spec_names_and_window_types():
    from vllm.v1.kv_cache_interface import FullAttentionSpec, SlidingWindowSpec

    from vllm_metal_b200.v1.worker import B200Worker, model_args_from_hf_config

    hf = types.SimpleNamespace(num_attention_heads=8, hidden_size=512, num_key_value_heads=4, head_dim=64,
                               intermediate_size=1024, vocab_size=1000, num_hidden_layers=4, rms_norm_eps=1e-5,
                               rope_theta=150000.0, tie_word_embeddings=False, model_type="gpt_oss",
                               sliding_window=128,
                               layer_types=["sliding_attention", "full_attention"] * 2,
                               quantization={"bits": 4, "group_size": 32})
    args = model_args_from_hf_config(hf)
    assert (args.num_layers, args.n_kv_heads, args.head_dim, args.group_size) == (4, 4, 64, 32)
    assert args.sliding_window_per_layer() == [128, -1, 128, -1]
    w = object.__new__(B200Worker)
    w.model = types.SimpleNamespace(args=args)
    w.parallel_config = types.SimpleNamespace(tensor_parallel_size=2)
    w.cache_config = types.SimpleNamespace(block_size=16)
    specs = w.get_kv_cache_spec()
    assert list(specs) == [f"layers.{i}.self_attn" for i in range(4)]  # cache_policy.py:383,406-411
    assert isinstance(specs["layers.0.self_attn"], SlidingWindowSpec)
    assert isinstance(specs["layers.1.self_attn"], FullAttentionSpec)
    assert specs["layers.1.self_attn"].num_kv_heads == 2  # sharded over TP=2
    assert w.get_cache_block_size_bytes() == 2 * 4 * 16 * 2 * 64 * 2


class _SpyRunner:
    """Stands where PagedModelRunner does: records what the worker packs, returns one token per sequence."""

    use_cuda_graph = False

    def __init__(self):
        self.calls = []
        self.next_tok = 1000

    def execute(self, tokens, decodes, prefills):
        self.calls.append((list(tokens), [(list(b), n) for b, n in decodes], [(list(b), q, o) for b, q, o in prefills]))
        out = list(range(self.next_tok, self.next_tok + len(decodes) + len(prefills)))
        self.next_tok += len(out)
        return out


def _sched(new=(), cached=None, num=None, finished=()):
    from vllm.v1.core.sched.output import CachedRequestData, SchedulerOutput

    cached = cached or CachedRequestData(req_ids=[], resumed_req_ids=set(), new_token_ids=[], all_token_ids={},
                                         new_block_ids=[], num_computed_tokens=[], num_output_tokens=[])
    num = num or {}
    return SchedulerOutput(scheduled_new_reqs=list(new), scheduled_cached_reqs=cached, num_scheduled_tokens=num,
                           total_num_scheduled_tokens=sum(num.values()), scheduled_spec_decode_tokens={},
                           scheduled_encoder_inputs={}, num_common_prefix_blocks=[0], finished_req_ids=set(finished),
                           free_encoder_mm_hashes=[])


def test_worker_drives_real_scheduler_outputs_through_chunked_prefill_decode_finish():
    """B2 with vLLM's own types (reference MetalWorker.execute_model(SchedulerOutput) -> sample_tokens ->
    ModelRunnerOutput, v1/worker.py:267-285): new request -> chunked prefill (no token for the partial chunk)
    -> mixed step with a second request's decode row packed FIRST -> finish frees the state."""
    from vllm.v1.core.sched.output import CachedRequestData, NewRequestData
    from vllm.v1.outputs import ModelRunnerOutput

    from vllm_metal_b200.v1.worker import B200Worker

    w = object.__new__(B200Worker)
    w._requests, w._pending = {}, None
    spy = w.model_runner = _SpyRunner()

    def new_req(rid, prompt, blocks):
        return NewRequestData(req_id=rid, prompt_token_ids=prompt, mm_features=[], sampling_params=None,
                              pooling_params=None, block_ids=(blocks,), num_computed_tokens=0, lora_request=None)

    # step 1: request A arrives, 5-token prompt, only 3 scheduled (chunked prefill); B arrives whole (2 tokens)
    so = _sched(new=[new_req("A", [10, 11, 12, 13, 14], [7]), new_req("B", [20, 21], [9])], num={"A": 3, "B": 2})
    assert w.execute_model(so) is None  # deferred-sampling contract
    out = w.sample_tokens(None)
    assert isinstance(out, ModelRunnerOutput)
    assert spy.calls[-1] == ([10, 11, 12, 20, 21], [], [([7], 3, 0), ([9], 2, 0)])
    assert out.req_ids == ["A", "B"] and out.sampled_token_ids == [[], [1001]]  # A's chunk is partial: no token
    assert out.req_id_to_index == {"A": 0, "B": 1}

    # step 2: A finishes its prompt (2 more tokens, a new block), B decodes: decode rows are packed first
    cached = CachedRequestData(req_ids=["A", "B"], resumed_req_ids=set(), new_token_ids=[[], []], all_token_ids={},
                               new_block_ids=[([8],), None], num_computed_tokens=[3, 2], num_output_tokens=[0, 1])
    w.execute_model(_sched(cached=cached, num={"A": 2, "B": 1}))
    out = w.sample_tokens(None)
    assert spy.calls[-1] == ([1001, 13, 14], [([9], 2)], [([7, 8], 2, 3)])
    assert out.req_ids == ["B", "A"] and out.sampled_token_ids == [[1002], [1003]]

    # step 3: both decode; B finished last step -> its state is dropped, A continues alone
    cached = CachedRequestData(req_ids=["A"], resumed_req_ids=set(), new_token_ids=[[]], all_token_ids={},
                               new_block_ids=[None], num_computed_tokens=[5], num_output_tokens=[1])
    w.execute_model(_sched(cached=cached, num={"A": 1}, finished=["B"]))
    out = w.sample_tokens(None)
    assert "B" not in w._requests
    assert spy.calls[-1] == ([1003], [([7, 8], 5)], [])
    assert out.sampled_token_ids == [[1004]] and w._requests["A"].tokens[-3:] == [14, 1003, 1004]
    assert w.sample_tokens(None) is None  # nothing pending


def test_worker_rejects_what_it_does_not_implement():
    from vllm_metal_b200.v1.worker import B200Worker

    w = object.__new__(B200Worker)
    w._requests, w._pending, w.model_runner = {}, None, _SpyRunner()
    with pytest.raises(NotImplementedError, match="SchedulerOutput"):
        w.execute_model(types.SimpleNamespace(decodes=[], prefills=[]))  # the old private duck type
    with pytest.raises(NotImplementedError, match="supported: ngram"):
        B200Worker._build_drafter(types.SimpleNamespace(method="eagle", num_speculative_tokens=3))
    w.execute_model(_sched())
    with pytest.raises(NotImplementedError, match="structured"):
        w.sample_tokens(object())
    assert w.get_supported_tasks() == ("generate",) and w.take_draft_token_ids() is None


def test_worker_verifies_scheduled_drafts_and_proposes_ngram_drafts():
    """N4: vLLM schedules the worker's own drafts back as ``scheduled_spec_decode_tokens``; a pure verification step
    goes through ``runner.verify`` in the window-mode layout (one (blocks, seq_len, K+1) request per window), greedy
    acceptance follows the reference (v1/spec_decode.py:272-292), a step with a prefill keeps the expanded layout."""
    from vllm.v1.core.sched.output import CachedRequestData, NewRequestData
    from vllm.v1.outputs import DraftTokenIds

    from vllm_metal_b200.v1.spec_decode import NgramDrafter, verify_greedy
    from vllm_metal_b200.v1.worker import B200Worker

    assert verify_greedy([5, 6, 7], [5, 6]) == [5, 6, 7]       # all drafts hold: + bonus token
    assert verify_greedy([5, 9, 7], [5, 6]) == [5, 9]          # second draft rejected: target's token instead
    assert verify_greedy([4, 6, 7], [5, 6]) == [4]
    d = NgramDrafter(3, min_n=2, max_n=3)
    assert d.propose([1, 2, 3, 4, 9, 1, 2, 3]) == [4, 9, 1]    # suffix (1,2,3) seen at 0 -> what followed it
    assert d.propose([1, 2, 3, 4, 9, 1, 2, 3], room=1) == [4]
    assert d.propose([1, 2, 3, 4]) == [] and d.propose([7, 7]) == []

    class _VerifyRunner(_SpyRunner):
        def __init__(self, rows):
            super().__init__()
            self.rows, self.verified = rows, []

        def verify(self, tokens, reqs):
            self.verified.append((list(tokens), [(list(b), s, n) for b, s, n in reqs]))
            return [self.rows.pop(0) for _ in reqs]

    w = object.__new__(B200Worker)
    w._requests, w._pending, w._drafts = {}, None, {}
    w._drafter = B200Worker._build_drafter(types.SimpleNamespace(method="ngram", num_speculative_tokens=2,
                                                                 prompt_lookup_min=2, prompt_lookup_max=3))
    spy = w.model_runner = _VerifyRunner(rows=[[31, 32, 50], [60, 61, 62]])
    spy.next_tok = 30
    prompt = [30, 31, 32, 33, 7, 30, 31]
    w.execute_model(_sched(new=[NewRequestData(req_id="A", prompt_token_ids=prompt, mm_features=[], sampling_params=None,
                                               pooling_params=None, block_ids=([3, 4],), num_computed_tokens=0,
                                               lora_request=None)], num={"A": 7}))
    assert w.sample_tokens(None).sampled_token_ids == [[30]]
    drafts = w.take_draft_token_ids()  # history ...30 31 | 30 -> suffix (31, 30)? no; (30, 31, 30) no; n=2: (31, 30) no
    assert drafts is None or isinstance(drafts, DraftTokenIds)
    st = w._requests["A"]
    assert st.tokens == prompt + [30] and st.num_computed == 7

    def cached(n_computed):
        return CachedRequestData(req_ids=["A"], resumed_req_ids=set(), new_token_ids=[[]], all_token_ids={},
                                 new_block_ids=[None], num_computed_tokens=[n_computed], num_output_tokens=[1])

    # verification step 1: drafts [31, 32]; target rows say 31, 32, then 50 -> both accepted + bonus
    so = _sched(cached=cached(7), num={"A": 3})
    so.scheduled_spec_decode_tokens = {"A": [31, 32]}
    w.execute_model(so)
    out = w.sample_tokens(None)
    assert spy.verified[-1] == ([30, 31, 32], [([3, 4], 7, 3)])
    assert out.sampled_token_ids == [[31, 32, 50]]
    assert st.tokens == prompt + [30, 31, 32, 50] and st.num_computed == 10
    drafts = w.take_draft_token_ids()
    assert drafts is None or drafts.req_ids == ["A"]
    assert w.take_draft_token_ids() is None  # handed out once

    # verification step 2: first draft rejected -> only the target's token, KV position count follows
    so = _sched(cached=cached(10), num={"A": 3})
    so.scheduled_spec_decode_tokens = {"A": [1, 2]}
    w.execute_model(so)
    assert w.sample_tokens(None).sampled_token_ids == [[60]]
    assert st.tokens[-2:] == [50, 60] and st.num_computed == 11

    # a window scheduled together with a prefill: expanded layout through execute (3-tuple decode request)
    spy.next_tok = 70
    so = _sched(new=[NewRequestData(req_id="B", prompt_token_ids=[1, 2], mm_features=[], sampling_params=None,
                                    pooling_params=None, block_ids=([9],), num_computed_tokens=0, lora_request=None)],
                cached=cached(11), num={"A": 2, "B": 2})
    so.scheduled_spec_decode_tokens = {"A": [70]}

    def execute3(tokens, decodes, prefills):
        spy.calls.append((list(tokens), [tuple(x) for x in decodes], [tuple(x) for x in prefills]))
        return [70, 71, 99]  # two window rows (expanded), then B's prefill

    spy.execute = execute3
    w.execute_model(so)
    out = w.sample_tokens(None)
    assert spy.calls[-1] == ([60, 70, 1, 2], [([3, 4], 11, 2)], [([9], 2, 0)])
    assert out.req_ids == ["A", "B"] and out.sampled_token_ids == [[70, 71], [99]]

    # malformed window: the row count must be 1 + drafts
    so = _sched(cached=cached(13), num={"A": 2})
    so.scheduled_spec_decode_tokens = {"A": [1, 2]}
    with pytest.raises(RuntimeError, match="draft tokens scheduled"):
        w.execute_model(so)


def test_worker_defers_sampling_one_step_ahead():
    """N4 deferred sampling (reference v1/decode_pipeline.py invariants): eligible pure-decode steps return an
    ``AsyncModelRunnerOutput``; the next step is submitted with placeholder inputs that name the pending step's rows;
    bookkeeping happens before submit, values are backfilled on resolve; an ineligible step resolves first."""
    from vllm.v1.core.sched.output import CachedRequestData, NewRequestData
    from vllm.v1.outputs import AsyncModelRunnerOutput, ModelRunnerOutput

    from vllm_metal_b200.v1.decode_pipeline import PENDING_TOKEN_PLACEHOLDER, DecodePipeline
    from vllm_metal_b200.v1.worker import B200Worker

    class _Handle:
        def __init__(self, toks):
            self.toks, self.reads = toks, 0

        def tokens(self):
            self.reads += 1
            return self.toks

    class _PipeRunner(_SpyRunner):
        use_cuda_graph, max_batch = True, 4

        def __init__(self):
            super().__init__()
            self.submits, self.handles = [], []

        def submit_decode(self, tokens, blocks, seq_lens, pending_rows):
            self.submits.append((list(tokens), [list(b) for b in blocks], list(seq_lens), list(pending_rows)))
            h = _Handle(list(range(self.next_tok, self.next_tok + len(tokens))))
            self.next_tok += len(tokens)
            self.handles.append(h)
            return h

        def decode_step(self, tokens, blocks, seq_lens):
            raise AssertionError("the synchronous graph path must not run on an eligible step")

    w = object.__new__(B200Worker)
    w._requests, w._pending, w._drafts, w._drafter = {}, None, {}, None
    w._pipeline = DecodePipeline()
    spy = w.model_runner = _PipeRunner()

    def new_req(rid, prompt, blocks):
        return NewRequestData(req_id=rid, prompt_token_ids=prompt, mm_features=[], sampling_params=None,
                              pooling_params=None, block_ids=(blocks,), num_computed_tokens=0, lora_request=None)

    def cached(ids, computed):
        return CachedRequestData(req_ids=ids, resumed_req_ids=set(), new_token_ids=[[] for _ in ids], all_token_ids={},
                                 new_block_ids=[None for _ in ids], num_computed_tokens=computed,
                                 num_output_tokens=[1 for _ in ids])

    # prefill: new requests -> synchronous path (gate: "new requests scheduled")
    w.execute_model(_sched(new=[new_req("A", [1, 2, 3], [7]), new_req("B", [4, 5], [9])], num={"A": 3, "B": 2}))
    out = w.sample_tokens(None)
    assert isinstance(out, ModelRunnerOutput) and out.sampled_token_ids == [[1000], [1001]]
    assert w._pipeline.blocked == {"new requests scheduled": 1}

    # step 1: pure decode -> deferred; real tokens go in, nothing is read back
    w.execute_model(_sched(cached=cached(["A", "B"], [3, 2]), num={"A": 1, "B": 1}))
    out1 = w.sample_tokens(None)
    assert isinstance(out1, AsyncModelRunnerOutput)
    assert spy.submits[-1] == ([1000, 1001], [[7], [9]], [3, 2], [-1, -1])
    a, b = w._requests["A"], w._requests["B"]
    assert a.tokens[-1] == PENDING_TOKEN_PLACEHOLDER and a.num_computed == 4 and b.num_computed == 3
    assert spy.handles[0].reads == 0

    # step 2 is submitted BEFORE step 1's output is taken: placeholders name step 1's rows; B comes first this time
    w.execute_model(_sched(cached=cached(["B", "A"], [3, 4]), num={"B": 1, "A": 1}))
    assert spy.submits[-1][2:] == ([3, 4], [1, 0]) and spy.handles[0].reads == 0
    out2 = w.sample_tokens(None)          # submitting step 2 resolves step 1 on the host
    assert spy.handles[0].reads == 1
    assert a.tokens[-2:] == [1002, PENDING_TOKEN_PLACEHOLDER] and b.tokens[-2:] == [1003, PENDING_TOKEN_PLACEHOLDER]
    r1 = out1.get_output()                # the engine takes step 1 late: cached, no second read
    assert r1.req_ids == ["A", "B"] and r1.sampled_token_ids == [[1002], [1003]] and spy.handles[0].reads == 1

    # step 3: A finishes, a new request arrives -> ineligible: the pending step 2 is resolved first, then the
    # synchronous path runs with real tokens only
    w.execute_model(_sched(new=[new_req("C", [8, 9], [11])], cached=cached(["B"], [4]), num={"B": 1, "C": 2},
                           finished=["A"]))
    assert spy.handles[1].reads == 1 and b.tokens[-1] == 1004 and PENDING_TOKEN_PLACEHOLDER not in b.tokens
    assert a.tokens[-1] == 1005           # backfilled through the direct reference although A was dropped
    assert "A" not in w._requests
    assert spy.calls[-1] == ([1004, 8, 9], [([9], 4)], [([11], 2, 0)])
    out3 = w.sample_tokens(None)
    assert isinstance(out3, ModelRunnerOutput) and out3.req_ids == ["B", "C"]
    r2 = out2.get_output()
    assert r2.req_ids == ["B", "A"] and r2.sampled_token_ids == [[1004], [1005]]
    assert out1.get_output() is r1        # the last two resolved steps are kept for a late reader
    assert w._pipeline.submitted == 2
