"""One-step-ahead decode pipelining for the B200 worker ("deferred sampling").

Reference: vllm_metal/v1/decode_pipeline.py:1-26, 182-347.  A fully serial decode step launches the step, then blocks
on the sampled tokens while the GPU idles through the next step's host work.  Here step ``k`` is submitted without
waiting (``PagedModelRunner.submit_decode``: graph replay + an async read-back behind an event) and handed to the
engine as an ``AsyncModelRunnerOutput``; step ``k+1`` takes its input tokens from step ``k``'s argmax ON THE DEVICE
(an index gather in the stream, no host round trip) while the host still has only placeholders, and step ``k``'s
tokens are read when the engine asks (or when the next step is submitted), by which time the GPU is already busy
with ``k+1``.

Invariants (the reference's, pinned by tests/test_plugin_surface.py):
- ``submit()`` only on gate-eligible steps: pure one-token decode rows, CUDA-graph path, no new / resumed /
  preempted requests, no drafts, no prefill rows;
- all value-independent bookkeeping (placeholder append, computed-token advance) happens BEFORE ``submit()``;
  ``resolve()`` only backfills token values through direct request-state references (a request dropped between
  submit and resolve is never re-inserted);
- an ineligible step calls ``begin_step()`` first, so the synchronous path never sees a placeholder.
"""

from __future__ import annotations

from typing import Any

PENDING_TOKEN_PLACEHOLDER = -1  # legal token ids are non-negative: a leaked placeholder is detectable


class DecodePipeline:
    """Holds at most one submitted-but-unresolved decode step."""

    KEEP = 2  # resolved outputs kept for a late ``get_output()``

    def __init__(self) -> None:
        self._pending: tuple[int, Any, list] | None = None  # (serial, runner handle, entries)
        self._serial = 0
        self._resolved: dict[int, Any] = {}
        self.submitted = 0
        self.blocked: dict[str, int] = {}

    # -- gate -------------------------------------------------------------------------------------------------
    def block_reason(self, so: Any, *, has_prefill_rows: bool, has_spec_tokens: bool, runner: Any, batch: int,
                     drafter: Any) -> str | None:
        """First fact that rules the pipeline out for this step, or ``None`` (decode_pipeline.py:88-143)."""
        cached = so.scheduled_cached_reqs
        blockers = (
            (bool(so.scheduled_new_reqs), "new requests scheduled"),
            (has_prefill_rows, "prefill-phase requests"),
            (bool(getattr(cached, "resumed_req_ids", None)), "resumed requests"),
            (bool(getattr(so, "preempted_req_ids", None)), "preempted requests"),
            (bool(getattr(so, "scheduled_encoder_inputs", None)), "encoder inputs scheduled"),
            (has_spec_tokens or drafter is not None, "speculative decode"),
            (bool(getattr(so, "has_structured_output_requests", False)), "structured output"),
            (not getattr(runner, "use_cuda_graph", False), "no CUDA-graph decode path"),
            (batch < 1 or batch > getattr(runner, "max_batch", 0), "batch outside the graph buckets"),
        )
        for blocked, reason in blockers:
            if blocked:
                self.blocked[reason] = self.blocked.get(reason, 0) + 1
                return reason
        return None

    # -- pending rows -----------------------------------------------------------------------------------------
    def pending_row(self, req_id: str) -> int:
        """Row of ``req_id`` in the pending step (its token is still on the device), or -1."""
        if self._pending is None:
            return -1
        for rid, _st, _idx, row in self._pending[2]:
            if rid == req_id:
                return row
        return -1

    def begin_step(self) -> None:
        """Resolve the pending step before a synchronous step mutates request state."""
        if self._pending is not None:
            self._resolve_pending()

    def submit(self, handle: Any, entries: list) -> B200AsyncModelRunnerOutput:
        """``entries``: ``(req_id, state, token_index, row)`` per request -- where each deferred token lands."""
        if self._pending is not None:  # its tokens were this step's inputs on the device; the host catches up now
            self._resolve_pending()
        self._serial += 1
        self._pending = (self._serial, handle, list(entries))
        self.submitted += 1
        return B200AsyncModelRunnerOutput(self, self._serial)

    def _resolve_pending(self) -> None:
        from vllm.v1.outputs import ModelRunnerOutput

        serial, handle, entries = self._pending
        self._pending = None
        toks = handle.tokens()
        req_ids, out = [], []
        for rid, st, idx, row in entries:
            tok = int(toks[row])
            if st.tokens[idx] != PENDING_TOKEN_PLACEHOLDER:
                raise RuntimeError(f"request {rid}: token {idx} was already resolved")
            st.tokens[idx] = tok
            req_ids.append(rid)
            out.append([tok])
        self._resolved[serial] = ModelRunnerOutput(
            req_ids=req_ids, req_id_to_index={r: i for i, r in enumerate(req_ids)}, sampled_token_ids=out,
            logprobs=None, prompt_logprobs_dict={}, pooler_output=[])
        for old in [k for k in self._resolved if k <= serial - self.KEEP]:  # the engine pops outputs in order
            del self._resolved[old]

    def resolve(self, serial: int):
        if self._pending is not None and self._pending[0] == serial:
            self._resolve_pending()
        if serial not in self._resolved:
            raise RuntimeError(f"decode pipeline: step {serial} is neither pending nor among the last resolved steps")
        return self._resolved[serial]


try:
    from vllm.v1.outputs import AsyncModelRunnerOutput as _AsyncBase
except Exception:  # pragma: no cover - vLLM is present in this image; keep the module importable without it
    _AsyncBase = object


class B200AsyncModelRunnerOutput(_AsyncBase):
    """Deferred step output; ``get_output()`` blocks until the tokens are read back."""

    def __init__(self, pipeline: DecodePipeline, serial: int) -> None:
        self._pipeline, self._serial = pipeline, serial

    def get_output(self):
        return self._pipeline.resolve(self._serial)
