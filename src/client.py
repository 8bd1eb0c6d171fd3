"""Programmatic client: what ``python -m src.main --stage 0`` does, as a call.

    from src.client import generate
    tokens, stats = generate("llama-2-7b", "8,16,24", "Hello, how are you?", initial_peers="/ip4/10.0.0.1/tcp/8000",
                             max_new_tokens=64, temperature=0.8, top_p=0.9)

The reference only has the CLI (src/main.py:62-227); this wraps the same loop (`run_rank0`): stage 0 of the model runs
here (embedding + the first blocks), every further span is found through the registry, failures are recovered by
journal replay.  Keyword arguments are the client-side flags of the CLI (``push``, ``stream``, ``device_handoff``,
``use_load_balancing``, ``total_blocks``, ``repetition_penalty``, ``request_timeout``, ``seed``, ``kv_capacity`` ...).
"""
from __future__ import annotations

from typing import List, Optional, Tuple

import torch


def generate(model: str, splits: str, prompt: str, initial_peers: str, *, max_new_tokens: int = 64,
             temperature: float = 1.0, top_p: float = 0.92, top_k: int = 50, dtype: str = "fp16",
             device: Optional[torch.device] = None, dht=None, on_token=None, **flags) -> Tuple[List[int], dict]:
    """Returns (generated token ids, timing / recovery statistics)."""
    from .main import build_parser, parse_splits, run_rank0

    a = build_parser().parse_args(["--model", model, "--splits", splits, "--stage", "0", "--dtype", dtype,
                                   "--dht_initial_peers", initial_peers or "", "--prompt", prompt,
                                   "--max_new_tokens", str(max_new_tokens), "--temperature", str(temperature),
                                   "--top_p", str(top_p), "--top_k", str(top_k), "--dht_port", "0", "--rpc_port", "0"])
    known = set(vars(a))
    for k, v in flags.items():
        if k not in known:
            raise TypeError(f"generate() got an unknown client flag {k!r}")
        setattr(a, k, v)
    if device is None:
        device = torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else torch.device("cpu")
    device = torch.device(device)
    a.dtype = {"fp16": torch.float16, "bf16": torch.bfloat16, "fp32": torch.float32}[dtype]
    if device.type == "cpu":
        a.dtype = torch.float32
    if on_token is not None:
        a.on_token = on_token
    return run_rank0(a, device, parse_splits(splits), dht=dht)
