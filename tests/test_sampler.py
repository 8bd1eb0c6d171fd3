"""Sampler semantics vs the reference's ``_sample_token`` (src/rpc_handler.py:327-403)."""
import pytest
import torch

from petals_b200.runtime.sampler import SamplingParams, apply_repetition_penalty, filtered_probs, sample_token
from tests.refstub import load_reference_module


def test_greedy_and_defaults():
    logits = torch.tensor([0.1, 3.0, -1.0, 2.9])
    assert sample_token(logits, SamplingParams(temperature=0.0)) == 1
    p = SamplingParams.from_metadata({})
    assert (p.temperature, p.top_p, p.top_k, p.repetition_penalty) == (0.8, 0.9, 0, 1.5)


def test_repetition_penalty_counts_and_triple():
    logits = torch.tensor([2.0, -2.0, 1.0, 4.0])
    out = apply_repetition_penalty(logits, [0, 0, 1], 2.0)
    assert out.tolist() == [2.0 / 4, -2.0 * 2, 1.0, 4.0]
    out = apply_repetition_penalty(logits, [3, 3, 3], 2.0)      # 3 repeats: 2**3 then an extra 2**3
    assert out[3].item() == pytest.approx(4.0 / 8 / 8)


def test_topk_topp_distribution():
    logits = torch.tensor([5.0, 4.0, 3.0, 2.0, 1.0, 0.0])
    pr = filtered_probs(logits, SamplingParams(1.0, 1.0, 2, 1.0))
    assert (pr > 0).sum() == 2 and pr.sum().item() == pytest.approx(1.0)
    pr = filtered_probs(logits, SamplingParams(1.0, 0.7, 0, 1.0))
    full = torch.softmax(logits, -1)
    keep = int((torch.cumsum(full, -1) <= 0.7).sum().clamp(min=1))
    assert (pr > 0).sum() == keep
    pr = filtered_probs(logits, SamplingParams(1.0, 1e-6, 0, 1.0))
    assert (pr > 0).sum() == 1                                     # the first entry always survives


@pytest.mark.parametrize("seed", range(2))
def test_distribution_matches_reference(seed):
    ref = load_reference_module("rpc_handler")
    if ref is None:
        pytest.skip("reference sources not available")
    h = ref.StageConnectionHandler.__new__(ref.StageConnectionHandler)
    g = torch.Generator().manual_seed(seed)
    V = 200
    logits = torch.randn(1, V, generator=g) * 3
    gen = torch.randint(0, V, (30,), generator=g).tolist() + [7, 7, 7]
    for (t, p, k, rp) in [(0.8, 0.9, 0, 1.5), (1.0, 0.92, 50, 1.5), (0.7, 1.0, 10, 1.0), (1.3, 0.5, 0, 1.2)]:
        ours = filtered_probs(logits[0].clone(), SamplingParams(t, p, k, rp), gen)
        # empirical check of the reference sampler against our closed-form distribution
        torch.manual_seed(seed)
        n = 1500
        counts = torch.zeros(V)
        for _ in range(n):
            counts[h._sample_token(logits.clone(), t, p, k, repetition_penalty=rp, generated_tokens=list(gen))] += 1
        emp = counts / n
        assert set(torch.nonzero(emp).flatten().tolist()) <= set(torch.nonzero(ours).flatten().tolist())
        assert (emp - ours).abs().max().item() < 0.05
    assert h._sample_token(logits.clone(), 0.0, 0.9, 0) == int(logits.argmax())
