"""Speculative decoding around the paged decode path: greedy verification of scheduled draft
tokens and a prompt-lookup (n-gram) drafter.

The reference verifies a K+1-row window per request with the target model's greedy tokens
(``SpeculativeDecodeController.verify_greedy``, vllm_metal/v1/spec_decode.py:243-294) and drafts, for
``method="ngram"``, by matching the longest suffix n-gram of the committed history against an earlier occurrence
(vllm_metal/v1/ngram_proposer.py:1-33, which wraps vLLM's numba kernel).  Here the verification rows run through
``PagedModelRunner.verify`` -- the decode kernel's window mode shares every KV read between a window's rows -- and
the drafter is a plain suffix match over the host-side token list (histories are a few thousand ints).
"""

from __future__ import annotations

from collections.abc import Sequence


def verify_greedy(row_tokens: Sequence[int], draft_token_ids: Sequence[int]) -> list[int]:
    """Tokens a request emits this step.  ``row_tokens[j]`` = the target's greedy token after window row j (row 0 is
    the last committed token, row j > 0 is draft j-1).  Drafts are accepted while they equal the target's token; the
    first mismatch emits the target's token instead and stops; if every draft holds, the last row's token is the
    bonus (spec_decode.py:272-292)."""
    if len(row_tokens) != len(draft_token_ids) + 1:
        raise ValueError(f"{len(draft_token_ids)} drafts need {len(draft_token_ids) + 1} target rows, got "
                         f"{len(row_tokens)}")
    out: list[int] = []
    for j, draft in enumerate(draft_token_ids):
        target = int(row_tokens[j])
        out.append(target)
        if target != int(draft):
            return out
    out.append(int(row_tokens[len(draft_token_ids)]))
    return out


class NgramDrafter:
    """Prompt-lookup drafts: find the longest n-gram (``max_n`` down to ``min_n`` tokens) that ends the history and
    also occurs earlier, and propose up to ``k`` tokens that followed its EARLIEST earlier occurrence (vLLM
    ``method="ngram"`` semantics: longest match first, leftmost occurrence)."""

    def __init__(self, num_speculative_tokens: int, min_n: int = 2, max_n: int = 5) -> None:
        if num_speculative_tokens < 1 or min_n < 1 or max_n < min_n:
            raise ValueError("NgramDrafter: need num_speculative_tokens >= 1 and 1 <= min_n <= max_n")
        self.k, self.min_n, self.max_n = int(num_speculative_tokens), int(min_n), int(max_n)

    def propose(self, tokens: Sequence[int], room: int | None = None) -> list[int]:
        """``room``: tokens the request may still take (model length / block budget); drafts never exceed it."""
        k = self.k if room is None else min(self.k, int(room))
        n_tok = len(tokens)
        if k <= 0:
            return []
        for n in range(min(self.max_n, n_tok - 1), self.min_n - 1, -1):
            suffix = list(tokens[n_tok - n:])
            for start in range(0, n_tok - n):  # earlier occurrences only; the one at n_tok - n is the suffix itself
                if list(tokens[start:start + n]) == suffix:
                    follow = list(tokens[start + n:start + n + k])
                    if follow:
                        return [int(t) for t in follow]
                    break
        return []
