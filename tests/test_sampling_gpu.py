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
