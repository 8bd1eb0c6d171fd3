"""Training-side RPCs (upstream rpc_forward without a session, rpc_backward): whole sequences through a served span,
nothing cached; the gradient w.r.t. the span's input equals autograd through the fp32 oracle."""
import torch

from tests.test_cpu_pipeline import _args


def _oracle(model, start, end, seed):
    from petals_b200.models.config import resolve_config
    from petals_b200.models.torch_ref import TorchSpan
    from petals_b200.weights.loader import load_span

    cfg = resolve_config(model)
    end = cfg.n_layers if end is None else end
    w = load_span(model, cfg, start, end, need_embed=False, need_head=False, dtype=torch.float32, device="cpu", seed=seed)
    return cfg, TorchSpan(cfg, w, torch.device("cpu"), torch.float32)


def _check(model, splits, stage, start, end):
    from petals_b200.net.rpc import RpcClient
    from src.main import parse_splits, run_stage_server_fixed

    h = run_stage_server_fixed(_args(model, splits, stage, "", seed=9), torch.device("cpu"), parse_splits(splits), block=False)
    try:
        cfg, ref = _oracle(model, start, end, 9)
        torch.manual_seed(0)
        x = torch.randn(2, 7, cfg.hidden)
        gy = torch.randn(2, 7, cfg.hidden)
        xr = x.clone().requires_grad_(True)
        y_ref = ref.forward_stateless(xr)
        (gx_ref,) = torch.autograd.grad(y_ref, xr, gy)
        c = RpcClient()
        addr = h.rpc.address
        (y,), md = c.call(addr, "StageConnectionHandler.rpc_forward", "k", [x], {})
        assert md.get("stateless") and torch.allclose(y, y_ref.detach(), atol=1e-5, rtol=1e-5)
        (gx,), _ = c.call(addr, "StageConnectionHandler.rpc_backward", "k", [x, gy], {})
        assert gx.shape == x.shape and torch.allclose(gx, gx_ref, atol=1e-5, rtol=1e-4)
        # the stateless pass sees what a session's prefill sees
        sess = {"session_id": "s", "seq_len": 7, "cur_len": 7, "is_prefill": True, "max_length": 32}
        out, md2 = c.call(addr, "StageConnectionHandler.rpc_forward", "k", [x[:1]], sess)
        if "token_id" not in md2:
            assert torch.allclose(out[0].reshape(7, -1), y[0], atol=1e-5, rtol=1e-5)
        assert len(h.handler.kv.sessions) == 1, "a stateless call opens no session"
        try:
            c.call(addr, "StageConnectionHandler.rpc_backward", "k", [x, gy[:, :3]], {})
            raise AssertionError("mismatched shapes must be refused")
        except Exception as e:  # noqa: BLE001
            assert "differ in shape" in str(e)
        c.close()
    finally:
        h.stop()


def test_stateless_forward_and_backward_llama_segment():
    _check("tiny-llama-gqa", "1,3", 1, 1, 3)


def test_stateless_forward_and_backward_gpt2_last_stage():
    from petals_b200.models.config import resolve_config

    n = resolve_config("tiny-gpt2").n_layers
    _check("tiny-gpt2", "1", 1, 1, n)


import pytest  # noqa: E402


@pytest.mark.gpu
@pytest.mark.parametrize("model", ["tiny-llama-gqa", "tiny-gpt2"])
def test_native_stage_untiles_blocks_for_stateless_passes(model):
    """A native (sm_100a) stage keeps only packed tiles; for the training-side passes it un-tiles one block at a time.
    Forward must match the stage's own kernels (a session prefill), backward the fp32 autograd oracle."""
    from petals_b200.models.config import resolve_config
    from petals_b200.models.torch_ref import TorchSpan
    from petals_b200.runtime.stages import StageSegment, StageWeights
    from petals_b200.weights.loader import load_span

    cfg = resolve_config(model)
    dev = torch.device("cuda", 0)
    dt = torch.bfloat16
    start, end = 1, cfg.n_layers - 1 if cfg.n_layers > 2 else cfg.n_layers
    w = load_span(model, cfg, start, end, need_embed=False, need_head=False, dtype=dt, device=dev, seed=2)
    w32 = {"layers": [{k: v.float() for k, v in L.items()} for L in w["layers"]]}
    ref = TorchSpan(cfg, w32, dev, torch.float32)
    stage = StageSegment(StageWeights(cfg, w, "segment", start, end, dev, dt), start, end, kv_capacity=64, max_batch=2)
    assert stage.backend == "native"
    torch.manual_seed(1)
    x = (torch.randn(2, 9, cfg.hidden, device=dev) * 0.5).to(dt)
    gy = (torch.randn(2, 9, cfg.hidden, device=dev) * 0.5).to(dt)
    y = stage.forward_stateless(x)
    kv = stage.new_kv_manager().allocate("s", 64)
    y_kernel, _ = stage(x[:1], None, None, kv)
    scale = y.float().abs().max().item()
    assert (y[0].float() - y_kernel.reshape(9, -1).float()).abs().max().item() < 0.03 * scale
    xr = x.float().requires_grad_(True)
    (g_ref,) = torch.autograd.grad(ref.forward_stateless(xr), xr, gy.float())
    g = stage.backward(x, gy)
    assert g.shape == x.shape
    assert (g.float() - g_ref).abs().max().item() < 0.05 * g_ref.abs().max().item()


def test_client_chains_stateless_passes_over_two_servers():
    """RpcTransport.forward_stateless / backward_stateless: the route's spans one after the other, gradients flow back
    in reverse; equal to autograd through the oracle over all served blocks."""
    from src.main import parse_splits, run_stage_server_fixed
    from src.rpc_transport import RpcTransport

    model, splits = "tiny-llama-gqa", "1,3"
    cuts = parse_splits(splits)
    handles = []
    try:
        h1 = run_stage_server_fixed(_args(model, splits, 1, "", seed=9), torch.device("cpu"), cuts, block=False)
        handles.append(h1)
        peers = h1.dht.get_visible_maddrs()[0]
        handles.append(run_stage_server_fixed(_args(model, splits, 2, peers, seed=9), torch.device("cpu"), cuts, block=False))
        cfg, ref = _oracle(model, 1, None, 9)
        tx = RpcTransport(device=torch.device("cpu"), dht_initial_peers=[peers],
                          stage_keys=["mini_petals:stage1", "mini_petals:stage2"], request_timeout=30.0)
        torch.manual_seed(3)
        x, gy = torch.randn(1, 6, cfg.hidden), torch.randn(1, 6, cfg.hidden)
        y, ctx = tx.forward_stateless(x)
        gx = tx.backward_stateless(ctx, gy)
        tx.shutdown()
        xr = x.clone().requires_grad_(True)
        y_ref = ref.forward_stateless(xr)
        (gx_ref,) = torch.autograd.grad(y_ref, xr, gy)
        assert torch.allclose(y, y_ref.detach(), atol=1e-5, rtol=1e-5)
        assert torch.allclose(gx, gx_ref, atol=1e-5, rtol=1e-4)
    finally:
        for h in handles:
            h.stop()


def test_deep_prompts_forward_and_gradients():
    """Deep prompt tuning over the wire: per-block prefixes are added to the block inputs; rpc_backward returns the
    gradients w.r.t. the input AND the prompts - equal to autograd through the oracle."""
    from petals_b200.net.rpc import RpcClient
    from src.main import parse_splits, run_stage_server_fixed

    model, splits = "tiny-llama-gqa", "1,3"
    h = run_stage_server_fixed(_args(model, splits, 1, "", seed=9), torch.device("cpu"), parse_splits(splits), block=False)
    try:
        cfg, ref = _oracle(model, 1, 3, 9)
        torch.manual_seed(2)
        x, gy = torch.randn(2, 6, cfg.hidden), torch.randn(2, 6, cfg.hidden)
        prompts = torch.randn(2, 2, 3, cfg.hidden) * 0.1                 # [blocks, batch, prefix, H]
        xr, pr = x.clone().requires_grad_(True), prompts.clone().requires_grad_(True)
        hcur = xr
        for li, L in enumerate(ref.layers):
            hcur = ref.block_stateless(torch.cat([hcur[:, :3] + pr[li], hcur[:, 3:]], 1), L)
        gx_ref, gp_ref = torch.autograd.grad(hcur, (xr, pr), gy)
        c = RpcClient()
        addr = h.rpc.address
        (y,), _ = c.call(addr, "StageConnectionHandler.rpc_forward", "k", [x, prompts], {})
        assert torch.allclose(y, hcur.detach(), atol=1e-5, rtol=1e-5)
        (gx, gp), _ = c.call(addr, "StageConnectionHandler.rpc_backward", "k", [x, gy, prompts], {})
        assert torch.allclose(gx, gx_ref, atol=1e-5, rtol=1e-4) and torch.allclose(gp, gp_ref, atol=1e-5, rtol=1e-4)
        try:
            c.call(addr, "StageConnectionHandler.rpc_forward", "k", [x, prompts[:1]], {})
            raise AssertionError("a prompt tensor for the wrong number of blocks must be refused")
        except Exception as e:  # noqa: BLE001
            assert "prompts must be" in str(e)
        c.close()
    finally:
        h.stop()
