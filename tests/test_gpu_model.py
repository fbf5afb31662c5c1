"""GPU parity at the model level: a packed prefill + greedy decode loop through the
product runner (sdpa_forward -> C-ABI kernels) vs the CPU oracle decoder holding the
same quantised weights.  North-star bar: greedy token ids identical; logits within
2e-2 max-abs / 1e-2 relative.  A mismatch is accepted only as a documented near-tie
(the reference does the same, tests/test_paged_deterministic.py:14-17,64-74): the
oracle's own logits of the two candidate tokens differ by less than the GEMM tolerance.
"""

from __future__ import annotations

import dataclasses

import numpy as np
import pytest
import torch

from .gpu_helpers import oracle_from_model, to_np

pytestmark = pytest.mark.gpu


def _run(preset_overrides, steps, prompt_lens, block_size=16, seed=0, cuda_graph=False):
    from vllm_metal_b200.attention.caches.kv_cache import B200PagedKVCache
    from vllm_metal_b200.lm.decoder import PRESETS, Model
    from vllm_metal_b200.v1.model_runner import PagedModelRunner

    args = dataclasses.replace(PRESETS["tiny"], **preset_overrides)
    model = Model(args, torch.bfloat16, "cuda", seed=seed)
    max_len = max(prompt_lens) + steps + 1
    blocks_per_seq = (max_len + block_size - 1) // block_size
    nb = 1 + blocks_per_seq * len(prompt_lens)
    cache = B200PagedKVCache(args.num_layers, args.n_kv_heads, args.head_dim, nb, block_size, torch.bfloat16,
                             "cuda", sliding_window_per_layer=args.sliding_window_per_layer())
    runner = PagedModelRunner(model, cache, max_batch=len(prompt_lens), max_model_len=max_len + block_size,
                              use_cuda_graph=cuda_graph)
    oracle = oracle_from_model(model, nb, block_size)
    rng = np.random.default_rng(seed + 1)
    perm = rng.permutation(nb - 1) + 1
    tables = [perm[i * blocks_per_seq:(i + 1) * blocks_per_seq].tolist() for i in range(len(prompt_lens))]
    prompts = [rng.integers(0, args.vocab, n).tolist() for n in prompt_lens]

    # packed prefill of every prompt in one step
    flat = [t for p in prompts for t in p]
    pre = [(tables[i], len(p), 0) for i, p in enumerate(prompts)]
    got = runner.execute(flat, [], pre)
    ologits = oracle.forward(flat, [], pre)
    want = oracle.greedy(ologits).tolist()
    exact, near = 0, 0

    def compare(got, want, ologits):
        nonlocal exact, near
        for i, (g, w) in enumerate(zip(got, want, strict=True)):
            if g == w:
                exact += 1
                continue
            gap = float(ologits[i, w] - ologits[i, g])
            assert gap <= 2e-2 + 1e-2 * abs(float(ologits[i, w])), (
                f"token mismatch seq {i}: got {g} want {w}, oracle logit gap {gap}")
            near += 1

    compare(got, want, ologits)
    seq_lens = list(prompt_lens)
    cur = want  # teacher forcing with the oracle's tokens keeps both sides on one trajectory
    for _ in range(steps):
        dec = [(tables[i], seq_lens[i]) for i in range(len(prompts))]
        if cuda_graph:
            got = runner.decode_step(cur, tables, seq_lens)
        else:
            got = runner.execute(cur, dec, [])
        ologits = oracle.forward(cur, dec, [])
        want = oracle.greedy(ologits).tolist()
        compare(got, want, ologits)
        cur = want
        seq_lens = [n + 1 for n in seq_lens]
    return exact, near


def test_tiny_qwen3_style_greedy_matches_oracle():
    exact, near = _run({}, steps=24, prompt_lens=[37, 5, 64])
    assert near <= 0.05 * (exact + near), f"{near} near-tie mismatches of {exact + near}"


def test_tiny_llama_style_int8_g32():
    exact, near = _run(dict(qk_norm=False, tie_embeddings=False, bits=8, group_size=32), steps=8,
                       prompt_lens=[20, 33])
    assert near <= 0.05 * (exact + near)


def test_tiny_gpt_oss_style_sinks_and_window():
    ov = dict(qk_norm=False, sinks=True, sliding_window=16, layer_types=["sliding_attention", "full_attention"],
              mlp=False, tie_embeddings=False)
    exact, near = _run(ov, steps=24, prompt_lens=[50, 17])
    assert near <= 0.05 * (exact + near)


def test_tiny_gpt_oss_style_moe_model():
    """N2 at the model level: sinks + alternating sliding window + a 4-expert top-2 MoE MLP with gpt-oss's clipped
    SwiGLU, router / expert biases -- packed prefill + greedy decode vs the oracle decoder (oracle/moe.py)."""
    ov = dict(qk_norm=False, sinks=True, sliding_window=16, layer_types=["sliding_attention", "full_attention"],
              tie_embeddings=False, num_experts=4, experts_per_tok=2, moe_activation="gpt_oss", moe_bias=True,
              intermediate=256)
    exact, near = _run(ov, steps=16, prompt_lens=[33, 9])
    assert near <= 0.05 * (exact + near)
    exact, near = _run(dict(num_experts=8, experts_per_tok=2, moe_routing="softmax_topk", intermediate=128), steps=8,
                       prompt_lens=[12, 20, 5])
    assert near <= 0.05 * (exact + near)


def test_tiny_cuda_graph_decode_matches_oracle():
    exact, near = _run({}, steps=40, prompt_lens=[30, 31, 500], cuda_graph=True)
    assert near <= 0.05 * (exact + near)


def _bf16_ulp(x: float) -> float:
    x = abs(float(x))
    return 2.0 ** -133 if x == 0.0 else 2.0 ** (int(np.floor(np.log2(x))) - 7)


def test_cfg_a_qwen3_0p6b_128_free_running_greedy_steps():
    """BASELINE configs[0] at its stated size: Qwen3-0.6B shapes (28 layers, hidden 1024, 16q/8kv x 128,
    vocab 151936, q/k norm, tied int4 g64 head), bs 1, 512-token prompt, block 16, 128 greedy
    decode steps.  The product runs FREE (CUDA graph, its own tokens fed back, as in
    /root/reference/tests/test_paged_deterministic.py:50-74); the oracle is stepped along the
    product's trajectory and must name the same token at every step.  A differing token is only
    tolerated -- and printed -- when the oracle's own logits of the two candidates are closer than the
    north-star parity bound on those logits themselves (2e-2 max-abs + 1e-2 relative: they are bf16
    GEMM outputs over 28 layers of bf16 activations, the product accumulates in other orders --
    tensor-core tiles, cluster split-K, stream-K pieces -- than the oracle's fp32 loop, so a candidate's
    logit may legitimately land a couple of bf16 ulps (1 ulp = 0.0156 at |logit| in [1, 2)) away).  With
    random-init weights and a 151936-way argmax the top two logits are routinely 0-2 ulp apart, so such
    ties do occur; anything wider fails, and at most 5% of the steps may be ties at all (an actual
    numerical bug flips far more: the logits' spread is ~40 ulp)."""
    from vllm_metal_b200.attention.caches.kv_cache import B200PagedKVCache
    from vllm_metal_b200.lm.decoder import PRESETS, Model
    from vllm_metal_b200.v1.model_runner import PagedModelRunner

    args = PRESETS["qwen3-0.6b"]
    steps, prompt_len, bs = 128, 512, 16
    model = Model(args, torch.bfloat16, "cuda", seed=0)
    blocks = (prompt_len + steps + bs) // bs
    nb = 1 + blocks
    cache = B200PagedKVCache(args.num_layers, args.n_kv_heads, args.head_dim, nb, bs, torch.bfloat16, "cuda")
    runner = PagedModelRunner(model, cache, max_batch=1, max_model_len=blocks * bs, use_cuda_graph=True)
    oracle = oracle_from_model(model, nb, bs)
    rng = np.random.default_rng(1)  # token ids seed 1, page permutation seed 2
    prompt = rng.integers(0, args.vocab, prompt_len).tolist()
    table = (np.random.default_rng(2).permutation(nb - 1) + 1).tolist()

    got = runner.execute(prompt, [], [(table, prompt_len, 0)])
    ologits = oracle.forward(prompt, [], [(table, prompt_len, 0)])
    ties = []

    def check(step, got_tok, ologits):
        want = int(oracle.greedy(ologits)[0])
        if got_tok == want:
            return
        gap = float(ologits[0, want] - ologits[0, got_tok])
        ulp = _bf16_ulp(ologits[0, want])
        print(f"step {step}: product {got_tok} vs oracle {want}, oracle logit gap {gap:.6f} (1 bf16 ulp = {ulp:.6f})")
        bound = 2e-2 + 1e-2 * abs(float(ologits[0, want]))
        assert gap <= bound, f"step {step}: token {got_tok} != oracle {want}; oracle top-2 gap {gap} > parity bound {bound}"
        ties.append(step)

    check(-1, got[0], ologits)
    seq_len, cur, tokens = prompt_len, got[0], []
    for step in range(steps):
        nxt = runner.decode_step([cur], [table], [seq_len])[0]
        ologits = oracle.forward([cur], [(table, seq_len)], [])
        check(step, nxt, ologits)
        tokens.append(nxt)
        cur, seq_len = nxt, seq_len + 1
    assert len(tokens) == steps
    print(f"cfg a: {steps} free-running greedy steps, {len(ties)} near-ties at steps {ties}")
    assert len(ties) <= 0.05 * steps, f"{len(ties)} near-tie steps of {steps}"


def test_worker_with_vllm_scheduler_outputs_matches_oracle():
    """B2 end to end on the device: ``B200Worker.execute_model(SchedulerOutput)`` / ``sample_tokens`` driven the way
    vLLM's engine core does (new request -> chunked prefill in two chunks -> 16 free-running decode steps, a second
    request joining mid-way), real runner + kernels underneath, every sampled token checked against the oracle
    stepped along the same trajectory (<= 1 bf16 ulp tie rule of the cfg-a test)."""
    pytest.importorskip("vllm")
    from vllm.v1.core.sched.output import CachedRequestData, NewRequestData, SchedulerOutput

    from vllm_metal_b200.attention.caches.kv_cache import B200PagedKVCache
    from vllm_metal_b200.lm.decoder import PRESETS, Model
    from vllm_metal_b200.v1.model_runner import PagedModelRunner
    from vllm_metal_b200.v1.worker import B200Worker

    args, bs = PRESETS["tiny"], 16
    model = Model(args, torch.bfloat16, "cuda", seed=3)
    nb = 24
    cache = B200PagedKVCache(args.num_layers, args.n_kv_heads, args.head_dim, nb, bs, torch.bfloat16, "cuda")
    w = object.__new__(B200Worker)
    w._requests, w._pending = {}, None
    w.model, w.kv_cache = model, cache
    w.model_runner = PagedModelRunner(model, cache, max_batch=4, max_model_len=8 * bs, use_cuda_graph=True)
    oracle = oracle_from_model(model, nb, bs)
    rng = np.random.default_rng(11)
    prompts = {"A": rng.integers(0, args.vocab, 40).tolist(), "B": rng.integers(0, args.vocab, 21).tolist()}
    blocks = {"A": [5, 9, 2, 17, 11], "B": [3, 20, 7]}
    known = {k: list(v) for k, v in prompts.items()}  # what the engine knows: prompt + sampled tokens
    computed = {"A": 0, "B": 0}
    live: list[str] = []

    def sched(new, num):
        new_reqs = [NewRequestData(req_id=r, prompt_token_ids=prompts[r], mm_features=[], sampling_params=None,
                                   pooling_params=None, block_ids=(blocks[r][:(computed[r] + num[r] + bs - 1) // bs],),
                                   num_computed_tokens=0, lora_request=None) for r in new]
        old = [r for r in live if r not in new and r in num]
        nbk = []
        for r in old:  # blocks the scheduler hands out as the sequence grows
            have = (computed[r] + bs - 1) // bs
            need = (computed[r] + num[r] + bs - 1) // bs
            nbk.append((blocks[r][have:need],) if need > have else None)
        cached = CachedRequestData(req_ids=old, resumed_req_ids=set(), new_token_ids=[[] for _ in old], all_token_ids={},
                                   new_block_ids=nbk, num_computed_tokens=[computed[r] for r in old],
                                   num_output_tokens=[len(known[r]) - len(prompts[r]) for r in old])
        return SchedulerOutput(scheduled_new_reqs=new_reqs, scheduled_cached_reqs=cached, num_scheduled_tokens=num,
                               total_num_scheduled_tokens=sum(num.values()), scheduled_spec_decode_tokens={},
                               scheduled_encoder_inputs={}, num_common_prefix_blocks=[0], finished_req_ids=set(),
                               free_encoder_mm_hashes=[])

    def step(new, num):
        for r in new:
            live.append(r)
        assert w.execute_model(sched(new, num)) is None
        out = w.sample_tokens(None)
        # the oracle runs the same rows in the worker's packing order (decode rows first)
        dec = [r for r in out.req_ids if num[r] == 1 and computed[r] > 0]
        pre = [r for r in out.req_ids if r not in dec]
        assert out.req_ids == dec + pre
        ids = [known[r][computed[r]] for r in dec] + [t for r in pre for t in known[r][computed[r]:computed[r] + num[r]]]
        ologits = oracle.forward(ids, [(blocks[r], computed[r]) for r in dec],
                                 [(blocks[r], num[r], computed[r]) for r in pre])
        want = oracle.greedy(ologits).tolist()
        for i, r in enumerate(out.req_ids):
            toks = out.sampled_token_ids[i]
            if computed[r] + num[r] < len(known[r]):
                assert toks == []  # partial prefill chunk samples nothing
            else:
                assert len(toks) == 1
                if toks[0] != want[i]:
                    gap = float(ologits[i, want[i]] - ologits[i, toks[0]])
                    assert gap <= 2.0 ** (int(np.floor(np.log2(abs(float(ologits[i, want[i]])) + 1e-30))) - 7), (r, toks, want[i], gap)
                known[r].append(toks[0])
            computed[r] += num[r]

    step(["A"], {"A": 24})            # chunk 1 of A's 40-token prompt
    step([], {"A": 16})               # chunk 2 finishes the prompt -> first token
    for i in range(16):
        if i == 4:
            step(["B"], {"A": 1, "B": 21})  # B joins: A's decode row is packed before B's prefill
        else:
            step([], {r: 1 for r in live})
    assert len(known["A"]) == 40 + 17 and len(known["B"]) == 21 + 12


def test_spec_verify_windows_match_expanded_and_oracle():
    """A K+1-row verification step through ``PagedModelRunner.verify``: the window-mode layout (one segment per
    request, KV shared between the rows of a window) gives the logits of the expanded layout bit for bit
    (reference tests/test_spec_window_parity.py at the model level), and every row's greedy token is the oracle's."""
    from vllm_metal_b200.attention.caches.kv_cache import B200PagedKVCache
    from vllm_metal_b200.lm.decoder import PRESETS, Model
    from vllm_metal_b200.v1.model_runner import PagedModelRunner

    args = dataclasses.replace(PRESETS["tiny"], sliding_window=24, layer_types=["sliding_attention", "full_attention"])
    model = Model(args, torch.bfloat16, "cuda", seed=3)
    prompt_lens, windows, bs = [45, 130, 7], [4, 6, 1], 16
    blocks_per_seq = (max(prompt_lens) + max(windows) + bs) // bs
    nb = 1 + blocks_per_seq * len(prompt_lens)
    rng = np.random.default_rng(5)
    perm = rng.permutation(nb - 1) + 1
    tables = [perm[i * blocks_per_seq:(i + 1) * blocks_per_seq].tolist() for i in range(len(prompt_lens))]
    prompts = [rng.integers(0, args.vocab, n).tolist() for n in prompt_lens]
    flat = [t for p in prompts for t in p]
    pre = [(tables[i], len(p), 0) for i, p in enumerate(prompts)]
    drafts = [rng.integers(0, args.vocab, w).tolist() for w in windows]
    win_ids = [t for d in drafts for t in d]
    reqs = [(tables[i], prompt_lens[i], windows[i]) for i in range(len(prompts))]

    logits = {}
    for merged in (False, True):
        cache = B200PagedKVCache(args.num_layers, args.n_kv_heads, args.head_dim, nb, bs, torch.bfloat16, "cuda",
                                 sliding_window_per_layer=args.sliding_window_per_layer())
        runner = PagedModelRunner(model, cache, max_batch=8, max_model_len=256, use_cuda_graph=False)
        runner.merge_verify_windows = merged
        runner.execute(flat, [], pre)
        toks = runner.verify(win_ids, reqs)
        assert [len(t) for t in toks] == windows
        logits[merged] = (runner.last_verify_logits.clone(), toks)
    assert torch.equal(logits[False][0], logits[True][0])
    assert logits[False][1] == logits[True][1]

    oracle = oracle_from_model(model, nb, bs)
    oracle.forward(flat, [], pre)
    ologits = oracle.forward(win_ids, reqs, [])  # expanded layout: one segment (= one logits row) per window row
    assert ologits.shape[0] == sum(windows)
    got = to_np(logits[True][0])
    np.testing.assert_allclose(got, ologits, atol=2e-2 + 1e-2 * np.abs(ologits).max(), rtol=0)
    want = oracle.greedy(ologits).tolist()
    flat_got = [t for row in logits[True][1] for t in row]
    for i, (g, w) in enumerate(zip(flat_got, want, strict=True)):
        if g != w:  # documented near-tie rule
            assert float(ologits[i, w] - ologits[i, g]) <= 2e-2 + 1e-2 * abs(float(ologits[i, w]))


def test_worker_speculative_steps_emit_the_plain_greedy_stream():
    """N4 on the device: the worker verifies scheduled drafts in window mode (``runner.verify``) and emits exactly the
    tokens plain greedy decoding emits -- right drafts are accepted with a bonus token, wrong ones cost nothing but
    the step (reference SpeculativeDecodeController.verify_greedy, v1/spec_decode.py:243-294)."""
    pytest.importorskip("vllm")
    from vllm.v1.core.sched.output import CachedRequestData, NewRequestData, SchedulerOutput

    from vllm_metal_b200.attention.caches.kv_cache import B200PagedKVCache
    from vllm_metal_b200.lm.decoder import PRESETS, Model
    from vllm_metal_b200.v1.model_runner import PagedModelRunner
    from vllm_metal_b200.v1.worker import B200Worker

    args, bs, nb = PRESETS["tiny"], 16, 24
    model = Model(args, torch.bfloat16, "cuda", seed=4)
    rng = np.random.default_rng(12)
    prompts = {"A": rng.integers(0, args.vocab, 37).tolist(), "B": rng.integers(0, args.vocab, 18).tolist()}
    blocks = {"A": [5, 9, 2, 17, 11], "B": [3, 20, 7, 1]}

    def make_worker():
        cache = B200PagedKVCache(args.num_layers, args.n_kv_heads, args.head_dim, nb, bs, torch.bfloat16, "cuda")
        w = object.__new__(B200Worker)
        w._requests, w._pending, w._drafts, w._drafter = {}, None, {}, None
        w.model, w.kv_cache = model, cache
        w.model_runner = PagedModelRunner(model, cache, max_batch=4, max_model_len=8 * bs, use_cuda_graph=False)
        return w

    def sched(new, num, computed, spec):
        new_reqs = [NewRequestData(req_id=r, prompt_token_ids=prompts[r], mm_features=[], sampling_params=None,
                                   pooling_params=None, block_ids=(blocks[r],), num_computed_tokens=0,
                                   lora_request=None) for r in new]
        old = [r for r in num if r not in new]
        cached = CachedRequestData(req_ids=old, resumed_req_ids=set(), new_token_ids=[[] for _ in old], all_token_ids={},
                                   new_block_ids=[None for _ in old], num_computed_tokens=[computed[r] for r in old],
                                   num_output_tokens=[1 for _ in old])
        return SchedulerOutput(scheduled_new_reqs=new_reqs, scheduled_cached_reqs=cached, num_scheduled_tokens=num,
                               total_num_scheduled_tokens=sum(num.values()), scheduled_spec_decode_tokens=spec,
                               scheduled_encoder_inputs={}, num_common_prefix_blocks=[0], finished_req_ids=set(),
                               free_encoder_mm_hashes=[])

    def run(w, num, computed, new=(), spec=None):
        w.execute_model(sched(list(new), num, computed, spec or {}))
        out = w.sample_tokens(None)
        return dict(zip(out.req_ids, out.sampled_token_ids, strict=True))

    # plain greedy: prefill both, then 14 one-token steps
    w = make_worker()
    plain = {r: [] for r in prompts}
    computed = {r: 0 for r in prompts}
    got = run(w, {r: len(prompts[r]) for r in prompts}, computed, new=prompts)
    for r in prompts:
        plain[r] += got[r]
        computed[r] = len(prompts[r])
    for _ in range(14):
        got = run(w, {r: 1 for r in prompts}, computed)
        for r in prompts:
            plain[r] += got[r]
            computed[r] += 1

    # speculative: the same requests; drafts alternate between the true continuation (accepted + bonus), a wrong
    # first draft (rejected at once) and a right-then-wrong pair
    w = make_worker()
    spec_out = {r: [] for r in prompts}
    computed = {r: 0 for r in prompts}
    got = run(w, {r: len(prompts[r]) for r in prompts}, computed, new=prompts)
    for r in prompts:
        spec_out[r] += got[r]
        computed[r] = len(prompts[r])
    step = 0
    while min(len(v) for v in spec_out.values()) < 12:
        spec, num = {}, {}
        for r in prompts:
            done = len(spec_out[r])
            truth = plain[r][done:done + 3]
            if len(truth) < 3:
                drafts = []
            elif step % 3 == 0:
                drafts = truth                                      # all three accepted, bonus token on top
            elif step % 3 == 1:
                drafts = [(truth[0] + 1) % args.vocab, truth[1]]    # wrong at once
            else:
                drafts = [truth[0], (truth[1] + 1) % args.vocab]    # one accepted, then the correction
            if drafts:
                spec[r] = drafts
            num[r] = 1 + len(drafts)
        got = run(w, num, computed, spec=spec)
        for r in prompts:
            emitted = got[r]
            want_n = {0: 4, 1: 1, 2: 2}[step % 3] if r in spec else 1
            assert len(emitted) == want_n, (r, step, emitted, spec.get(r))
            spec_out[r] += emitted
            computed[r] += len(emitted)
        step += 1
    for r in prompts:
        n = min(len(spec_out[r]), len(plain[r]))
        assert spec_out[r][:n] == plain[r][:n], r


def test_worker_deferred_sampling_emits_the_synchronous_stream():
    """N4 deferred sampling on the device: with the one-step-ahead pipeline every decode step is submitted before the
    previous step's tokens are read (inputs gathered from the previous argmax on the GPU); the emitted tokens are the
    synchronous worker's, token for token, across a block boundary and a request joining / leaving mid-way."""
    pytest.importorskip("vllm")
    from vllm.v1.core.sched.output import CachedRequestData, NewRequestData, SchedulerOutput
    from vllm.v1.outputs import AsyncModelRunnerOutput

    from vllm_metal_b200.attention.caches.kv_cache import B200PagedKVCache
    from vllm_metal_b200.lm.decoder import PRESETS, Model
    from vllm_metal_b200.v1.decode_pipeline import DecodePipeline
    from vllm_metal_b200.v1.model_runner import PagedModelRunner
    from vllm_metal_b200.v1.worker import B200Worker

    args, bs, nb = PRESETS["tiny"], 16, 24
    model = Model(args, torch.bfloat16, "cuda", seed=6)
    rng = np.random.default_rng(13)
    prompts = {"A": rng.integers(0, args.vocab, 29).tolist(), "B": rng.integers(0, args.vocab, 14).tolist(),
               "C": rng.integers(0, args.vocab, 9).tolist()}
    blocks = {"A": [5, 9, 2, 17], "B": [3, 20, 7], "C": [1, 12, 4]}

    def run(pipelined: bool) -> dict[str, list[int]]:
        cache = B200PagedKVCache(args.num_layers, args.n_kv_heads, args.head_dim, nb, bs, torch.bfloat16, "cuda")
        w = object.__new__(B200Worker)
        w._requests, w._pending, w._drafts, w._drafter = {}, None, {}, None
        w._pipeline = DecodePipeline() if pipelined else None
        w.model, w.kv_cache = model, cache
        w.model_runner = PagedModelRunner(model, cache, max_batch=4, max_model_len=4 * bs, use_cuda_graph=True)
        emitted = {r: [] for r in prompts}
        computed = {r: 0 for r in prompts}
        outstanding = []  # async outputs the "engine" has not taken yet

        def take(out):
            got = out.get_output() if isinstance(out, AsyncModelRunnerOutput) else out
            for rid, toks in zip(got.req_ids, got.sampled_token_ids, strict=True):
                emitted[rid] += toks

        def step(new, num, finished=()):
            new_reqs = [NewRequestData(req_id=r, prompt_token_ids=prompts[r], mm_features=[], sampling_params=None,
                                       pooling_params=None, block_ids=(blocks[r],), num_computed_tokens=0,
                                       lora_request=None) for r in new]
            old = [r for r in num if r not in new]
            cached = CachedRequestData(req_ids=old, resumed_req_ids=set(), new_token_ids=[[] for _ in old],
                                       all_token_ids={}, new_block_ids=[None for _ in old],
                                       num_computed_tokens=[computed[r] for r in old], num_output_tokens=[1 for _ in old])
            so = SchedulerOutput(scheduled_new_reqs=new_reqs, scheduled_cached_reqs=cached, num_scheduled_tokens=num,
                                 total_num_scheduled_tokens=sum(num.values()), scheduled_spec_decode_tokens={},
                                 scheduled_encoder_inputs={}, num_common_prefix_blocks=[0],
                                 finished_req_ids=set(finished), free_encoder_mm_hashes=[])
            w.execute_model(so)
            out = w.sample_tokens(None)
            for r, n in num.items():
                computed[r] += n
            # the engine takes output k only after step k+1 was submitted (async scheduling)
            outstanding.append(out)
            while len(outstanding) > 1:
                take(outstanding.pop(0))

        step(["A", "B"], {"A": 29, "B": 14})
        for i in range(22):
            if i == 6:
                step(["C"], {"A": 1, "B": 1, "C": 9})       # C joins: a synchronous step in the middle
            elif i == 15:
                step([], {"A": 1, "C": 1}, finished=["B"])   # B leaves: the batch shrinks, rows move
            elif i > 15:
                step([], {"A": 1, "C": 1})
            elif i > 6:
                step([], {"A": 1, "B": 1, "C": 1})
            else:
                step([], {"A": 1, "B": 1})
        while outstanding:
            take(outstanding.pop(0))
        if pipelined:
            assert w._pipeline.submitted >= 18, w._pipeline.blocked
        return emitted

    sync, piped = run(False), run(True)
    assert {r: len(v) for r, v in piped.items()} == {r: len(v) for r, v in sync.items()}
    assert piped == sync
