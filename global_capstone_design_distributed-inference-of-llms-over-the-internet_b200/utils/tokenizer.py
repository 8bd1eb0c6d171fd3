"""Tokenizer loading with an offline fallback.

The reference calls ``AutoTokenizer.from_pretrained(model)`` (src/main.py:98-100), which needs
the hub.  Here a local checkpoint directory with tokenizer files is used when present;
otherwise a byte-level tokenizer (UTF-8 bytes folded into the model's vocabulary) keeps the CLI
usable with random-init weights on a machine without network access.
"""
from __future__ import annotations

import os
from typing import List, Optional

from ..models.config import ModelConfig


class ByteTokenizer:
    def __init__(self, cfg: ModelConfig):
        self.vocab = cfg.vocab
        self.eos_token_id: Optional[int] = None   # random weights: never stop on a made-up EOS
        self.pad_token_id: Optional[int] = 0
        self.bos_token_id = cfg.bos_token_id

    def encode(self, text: str) -> List[int]:
        ids = [3 + b for b in text.encode("utf-8")] or [3]
        return [i % self.vocab for i in ids]

    def decode(self, ids, skip_special_tokens: bool = True) -> str:
        out = bytearray()
        for i in ids:
            i = int(i) - 3
            out.append(i if 0 <= i < 256 else ord("?"))
        return out.decode("utf-8", errors="replace")


class HFTokenizer:
    def __init__(self, tok):
        self.tok = tok
        if tok.pad_token is None:
            tok.pad_token = tok.eos_token
        self.eos_token_id = tok.eos_token_id if tok.eos_token_id is not None else tok.pad_token_id
        self.pad_token_id = tok.pad_token_id
        self.bos_token_id = tok.bos_token_id

    def encode(self, text: str) -> List[int]:
        return list(self.tok(text, return_tensors=None)["input_ids"])

    def decode(self, ids, skip_special_tokens: bool = True) -> str:
        return self.tok.decode(list(ids), skip_special_tokens=skip_special_tokens)


def load_tokenizer(model: str, cfg: ModelConfig):
    if os.path.isdir(model) and any(os.path.exists(os.path.join(model, f))
                                    for f in ("tokenizer.json", "tokenizer.model", "vocab.json")):
        try:
            from transformers import AutoTokenizer

            return HFTokenizer(AutoTokenizer.from_pretrained(model, local_files_only=True))
        except Exception:  # noqa: BLE001 - fall back to bytes
            pass
    return ByteTokenizer(cfg)
