"""Device sampling kernel vs the torch restatement of the reference's sampler."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("V", [1000, 32000, 50257])
@pytest.mark.parametrize("cfg", [(0.8, 0.9, 0, 1.5), (1.0, 0.92, 50, 1.5), (0.7, 1.0, 10, 1.0), (1.3, 0.5, 0, 1.2),
                                 (1.0, 1.0, 0, 1.0), (0.5, 1e-6, 0, 1.0)])
def test_final_distribution_matches_reference_rules(V, cfg):
    from petals_b200.ops.sampling import DeviceSampler
    from petals_b200.runtime.sampler import SamplingParams, filtered_probs

    t, p, k, rp = cfg
    g = torch.Generator().manual_seed(V + int(100 * t))
    T = 3
    logits = (torch.randn(T, V, generator=g) * 3).cuda()
    gen = [torch.randint(0, V, (30,), generator=g).tolist() + [7, 7, 7], [], [5, 9, 5, 5]]
    params = SamplingParams(t, p, k, rp)
    ds = DeviceSampler("cuda", T, V)
    u = torch.tensor([0.0, 0.5, 0.999], device="cuda")
    ids = ds.sample(logits, params, gen, uniforms=u)
    torch.cuda.synchronize()
    for r in range(T):
        ref = filtered_probs(logits[r].clone(), params, gen[r]).double().cpu()
        got = ds.probs[r].double().cpu()
        got = got / got.sum()
        assert set(torch.nonzero(got).flatten().tolist()) == set(torch.nonzero(ref).flatten().tolist()), "kept set differs"
        assert (got - ref).abs().max().item() < 1e-5
        cdf = torch.cumsum(got, 0)
        want = float(u[r]) * float(cdf[-1])
        exp_idx = int(torch.nonzero((cdf > want) & (got > 0)).flatten()[0])
        assert int(ids[r]) == exp_idx


def test_sampling_statistics():
    from petals_b200.ops.sampling import DeviceSampler
    from petals_b200.runtime.sampler import SamplingParams, filtered_probs

    V, N = 500, 4000
    g = torch.Generator().manual_seed(1)
    row = (torch.randn(V, generator=g) * 2).cuda()
    params = SamplingParams(0.9, 0.9, 40, 1.0)
    ds = DeviceSampler("cuda", 32, V)
    counts = torch.zeros(V)
    gen = torch.Generator(device="cuda").manual_seed(3)
    for _ in range(N // 32):
        ids = ds.sample(row.expand(32, V).contiguous(), params, [[]] * 32, generator=gen)
        counts += torch.bincount(ids.cpu(), minlength=V).float()
    ref = filtered_probs(row.clone(), params, []).cpu()
    assert (counts / counts.sum() - ref).abs().max().item() < 0.03


def test_stateful_sampler_matches_stateless_with_philox_uniforms():
    """b200_sample_stateful keeps parameters, recent tokens and the Philox stream on the device; fed the same
    uniform numbers (host restatement of Philox4x32-10) the stateless kernel must pick the same tokens, step by step."""
    from petals_b200.ops.sampling import DeviceSampler, SamplerPool, philox_uniform
    from petals_b200.runtime.sampler import SamplingParams

    dev = torch.device("cuda")
    V = 4096
    g = torch.Generator(device="cuda").manual_seed(7)
    params = [SamplingParams(0.9, 0.92, 40, 1.3), SamplingParams(1.2, 0.8, 0, 1.5)]
    seeds = [1234, 99]
    pool = SamplerPool(dev, 8, V, max_batch=4)
    rows = [pool.open(p, s, recent=[5, 5, 5] if i == 0 else []) for i, (p, s) in enumerate(zip(params, seeds))]
    idx = torch.tensor(rows, dtype=torch.int32, device=dev)
    ref = DeviceSampler(dev, 2, V)
    hist = [[5, 5, 5], []]
    out = torch.zeros(2, dtype=torch.int64, device=dev)
    for step in range(12):
        logits = torch.randn(2, V, device=dev, generator=g) * 3.0
        pool.sample(logits, idx, out)
        u = torch.tensor([philox_uniform(seeds[i], step) for i in range(2)], dtype=torch.float32, device=dev)
        want = ref.sample(logits, params, hist, uniforms=u).clone()
        assert torch.equal(out, want), f"step {step}: {out.tolist()} vs {want.tolist()}"
        for i in range(2):
            hist[i].append(int(want[i]))
    assert pool.states[rows[0], 1].item() == 12 and pool.states[rows[1], 1].item() == 12   # draws made
