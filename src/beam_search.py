"""Client-side beam search over the served pipeline.

Upstream Petals keeps the hypotheses of a beam in one batched cache and tells the servers, with ``hypo_ids``, which
row each hypothesis continues (petals/server/backend.py:154-158).  Here a hypothesis is a session: every server of the
route (and this client's own blocks) holds one cache per hypothesis, the last stage returns the ``k`` most likely next
tokens with their log-probabilities (``return_logprobs`` in the request), and between steps
:meth:`RpcTransport.reorder_hypotheses` makes session *i* continue the hypothesis the search kept in slot *i*
(``rpc_reorder`` -> ``KVCacheManager.reorder``).  The reference itself only samples (src/main.py:62-227).
"""
from __future__ import annotations

from typing import List, Optional, Sequence, Tuple
from uuid import uuid4

import torch


def beam_search(tx, stage0, kv_manager, prompt_ids: Sequence[int], num_beams: int, max_new_tokens: int,
                eos_token_id: Optional[int] = None, length_penalty: float = 1.0) -> Tuple[List[int], float, List[List[int]]]:
    """Returns (best continuation, its score, all final hypotheses).

    ``tx``: an :class:`RpcTransport`; ``stage0``: the client's :class:`Stage0` (embedding + first blocks);
    ``kv_manager``: ``stage0.new_kv_manager()``.  Scores are sums of token log-probabilities, divided by
    ``length ** length_penalty`` when hypotheses are compared at the end (HF's convention)."""
    B = int(num_beams)
    L = len(prompt_ids)
    max_length = L + max_new_tokens
    base = str(uuid4())
    sids = [f"{base}-h{i}" for i in range(B)]
    kvs = [kv_manager.allocate(s, stage0.kv_capacity) for s in sids]
    saved_t, tx.temperature = tx.temperature, 0.0
    tx.extra_metadata = dict(tx.extra_metadata, return_logprobs=2 * B)
    ids = torch.tensor([list(prompt_ids)], dtype=torch.long)
    try:
        # every hypothesis starts from the same prompt: one prefill per session (the caches must exist everywhere)
        reply = None
        for sid, kv in zip(sids, kvs):
            hidden, _ = stage0(ids, None, None, kv)
            tx.send_prefill(L, hidden, session_id=sid, max_length=max_length)
            tx.recv_token()
            reply = reply or dict(tx.last_reply)
        scores = [0.0] * B
        seqs: List[List[int]] = [[] for _ in range(B)]
        cand = list(zip(reply["topk_logprobs"], reply["topk_ids"]))[:B]          # first step: B best tokens of ONE row
        beams = [(lp, 0, tok) for lp, tok in cand]
        finished: List[Tuple[float, List[int]]] = []
        cur_len = L
        for step in range(max_new_tokens):
            parents = [p for _, p, _ in beams]
            if parents != list(range(B)):
                tx.reorder_hypotheses(sids, parents)
                kv_manager.reorder(sids, parents)
            seqs = [seqs[p] + [tok] for _, p, tok in beams]
            scores = [sc for sc, _, _ in beams]
            cur_len += 1
            if step == max_new_tokens - 1:
                break
            pool = []
            for i, (sid, kv) in enumerate(zip(sids, kvs)):
                if eos_token_id is not None and seqs[i][-1] == eos_token_id:
                    finished.append((scores[i] / (len(seqs[i]) ** length_penalty), seqs[i]))
                    scores[i] = float("-inf")           # a finished hypothesis keeps its slot but cannot be extended
                hidden, _ = stage0(torch.tensor([[seqs[i][-1]]], dtype=torch.long), None, None, kv)
                tx.send_decode_step(cur_len, hidden, session_id=sid, max_length=max_length, generated_tokens=seqs[i])
                tx.recv_token()
                r = tx.last_reply
                pool += [(scores[i] + lp, i, tok) for lp, tok in zip(r["topk_logprobs"], r["topk_ids"])]
            pool.sort(key=lambda c: (-c[0], c[1], c[2]))
            beams = pool[:B]
        final = finished + [(scores[i] / (len(seqs[i]) ** length_penalty), seqs[i]) for i in range(B)
                            if scores[i] != float("-inf")]
        final.sort(key=lambda c: -c[0])
        return final[0][1], final[0][0], [s for _, s in final]
    finally:
        tx.temperature = saved_t
        tx.extra_metadata = {k: v for k, v in tx.extra_metadata.items() if k != "return_logprobs"}
        for sid in sids:
            try:
                tx.close_session(sid)
            except Exception:  # noqa: BLE001
                pass
            kv_manager.close(sid)
