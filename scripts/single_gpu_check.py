#!/usr/bin/env python
"""Single-process oracle + micro-benchmark (reference: scripts/single_gpu_check.py:16-319).

Runs the whole model in one process two ways — HF transformers (when the architecture can be
instantiated from the config) and this framework's engine on the same weights — and prints the
top-5 prefill logits, the greedy first token, TTFT, decode tokens/s and throughput for both.

    python scripts/single_gpu_check.py --model tiny-llama-gqa --dtype bf16 --max_new_tokens 32
"""
import argparse
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="tiny-llama-gqa")
    ap.add_argument("--dtype", default="bf16", choices=["fp16", "bf16", "fp32"])
    ap.add_argument("--prompt", default="Hello, how are you?")
    ap.add_argument("--max_new_tokens", type=int, default=32)
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--no_hf", action="store_true")
    a = ap.parse_args()
    import torch

    from petals_b200.models.config import resolve_config
    from petals_b200.parallel.pipeline import PipelineEngine
    from petals_b200.utils.tokenizer import load_tokenizer
    from petals_b200.weights.loader import full_hf_state_dict

    cfg = resolve_config(a.model)
    dev = torch.device("cuda", 0) if torch.cuda.is_available() else torch.device("cpu")
    dtype = {"fp16": torch.float16, "bf16": torch.bfloat16, "fp32": torch.float32}[a.dtype]
    if dev.type == "cpu":
        dtype = torch.float32
    tok = load_tokenizer(a.model, cfg)
    ids = torch.tensor(tok.encode(a.prompt), dtype=torch.long)
    sync = (lambda: torch.cuda.synchronize()) if dev.type == "cuda" else (lambda: None)

    eng = PipelineEngine(a.model, [], dtype, dev, kv_capacity=max(256, len(ids) + a.max_new_tokens + 8), seed=a.seed)
    sync()
    t0 = time.perf_counter()
    outs, stats = eng.generate([ids], a.max_new_tokens)
    sync()
    print(f"[framework/{eng.backend}] tokens: {outs[0][:16]}{'...' if len(outs[0]) > 16 else ''}")
    print(f"[framework/{eng.backend}] TTFT {stats['ttft_s'] * 1e3:.2f} ms | decode {stats['decode_tok_per_s']:.1f} tok/s | "
          f"total {stats['total_s']:.3f} s | throughput {a.max_new_tokens / stats['total_s']:.1f} tok/s")

    if not a.no_hf and cfg.layer_param_count() * cfg.n_layers < 2e9:
        import transformers

        hf = transformers.AutoModelForCausalLM.from_config(transformers.AutoConfig.for_model(**cfg.to_hf_dict()))
        sd = full_hf_state_dict(cfg, a.seed)
        if not cfg.is_llama:
            sd = {("transformer." + k): v for k, v in sd.items()}
            sd["lm_head.weight"] = sd["transformer.wte.weight"]
        hf.load_state_dict(sd, strict=False)
        hf = hf.to(dev, dtype).eval()
        x = ids[None].to(dev)
        with torch.no_grad():
            sync(); t0 = time.perf_counter()
            out = hf(x, use_cache=True)
            sync(); ttft = time.perf_counter() - t0
            top = out.logits[0, -1].float().topk(5)
            print("[hf] prefill top-5:", list(zip(top.indices.tolist(), [round(v, 3) for v in top.values.tolist()])))
            gen = [int(out.logits[0, -1].argmax())]
            past = out.past_key_values
            t1 = time.perf_counter()
            for _ in range(a.max_new_tokens - 1):
                out = hf(torch.tensor([[gen[-1]]], device=dev), past_key_values=past, use_cache=True)
                past = out.past_key_values
                gen.append(int(out.logits[0, -1].argmax()))
            sync(); dec = time.perf_counter() - t1
        print(f"[hf] tokens: {gen[:16]}{'...' if len(gen) > 16 else ''}")
        print(f"[hf] TTFT {ttft * 1e3:.2f} ms | decode {(a.max_new_tokens - 1) / dec:.1f} tok/s")
        agree = sum(int(x == y) for x, y in zip(gen, outs[0]))
        print(f"greedy agreement with HF: {agree}/{len(gen)} tokens"
              + ("" if dtype == torch.float32 else "  (16-bit paths may diverge after a near-tie)"))


if __name__ == "__main__":
    main()
