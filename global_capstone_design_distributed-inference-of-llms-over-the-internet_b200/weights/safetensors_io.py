"""Minimal safetensors reader/writer (format: u64 LE header length, JSON header, raw data).

Only the tensors a stage asks for are touched: the file is memory-mapped and each
tensor is a zero-copy view until it is moved to the device.  This is what makes
span-selective loading possible — the reference materialises the whole model on
the CPU first (src/llama_partition.py:495-499); the dead upstream loader filters
shards by block prefix (petals/server/from_pretrained.py:94-128, 216-224).
"""
from __future__ import annotations

import json
import mmap
import os
import struct
from typing import Dict, Iterable, List, Optional

import warnings

import torch

_DTYPES = {
    "F32": torch.float32, "F16": torch.float16, "BF16": torch.bfloat16, "F64": torch.float64,
    "I64": torch.int64, "I32": torch.int32, "I16": torch.int16, "I8": torch.int8, "U8": torch.uint8,
    "BOOL": torch.bool, "F8_E4M3": getattr(torch, "float8_e4m3fn", torch.uint8),
}
_NAMES = {v: k for k, v in _DTYPES.items()}


class SafetensorsFile:
    def __init__(self, path: str):
        self.path = path
        self._f = open(path, "rb")
        (hlen,) = struct.unpack("<Q", self._f.read(8))
        if hlen > 100 << 20:
            raise ValueError(f"{path}: implausible safetensors header length {hlen}")
        self.header: Dict[str, dict] = json.loads(self._f.read(hlen))
        self.metadata = self.header.pop("__metadata__", {})
        self._base = 8 + hlen
        self._mm = mmap.mmap(self._f.fileno(), 0, access=mmap.ACCESS_READ)

    def keys(self) -> List[str]:
        return list(self.header.keys())

    def __contains__(self, name: str) -> bool:
        return name in self.header

    def get(self, name: str) -> torch.Tensor:
        info = self.header[name]
        b, e = info["data_offsets"]
        dt = _DTYPES[info["dtype"]]
        buf = memoryview(self._mm)[self._base + b: self._base + e]
        if e == b:
            return torch.empty(info["shape"], dtype=dt)
        with warnings.catch_warnings():
            warnings.simplefilter("ignore", UserWarning)   # read-only mmap view; nothing writes through it
            t = torch.frombuffer(buf, dtype=dt)
        return t.view(info["shape"])

    def close(self):
        try:
            self._mm.close()
        except BufferError:
            pass  # live tensor views keep the map alive; the OS reclaims it with them
        self._f.close()


def save_file(tensors: Dict[str, torch.Tensor], path: str, metadata: Optional[Dict[str, str]] = None) -> None:
    header: Dict[str, dict] = {}
    off = 0
    order = sorted(tensors.keys())
    for k in order:
        t = tensors[k]
        n = t.numel() * t.element_size()
        header[k] = {"dtype": _NAMES[t.dtype], "shape": list(t.shape), "data_offsets": [off, off + n]}
        off += n
    if metadata:
        header["__metadata__"] = metadata
    hj = json.dumps(header, separators=(",", ":")).encode()
    hj += b" " * ((8 - len(hj) % 8) % 8)
    with open(path, "wb") as f:
        f.write(struct.pack("<Q", len(hj)))
        f.write(hj)
        for k in order:
            t = tensors[k].detach().contiguous().cpu()
            f.write(t.view(torch.uint8).numpy().tobytes() if t.numel() else b"")


class CheckpointIndex:
    """Name -> tensor access over an HF checkpoint directory, opening only needed shards.

    Search order matches upstream Petals (from_pretrained.py:131): sharded safetensors,
    single safetensors, sharded .bin, single .bin.
    """

    def __init__(self, directory: str):
        self.dir = directory
        self._files: Dict[str, object] = {}
        self.weight_map: Dict[str, str] = {}
        p = lambda f: os.path.join(directory, f)  # noqa: E731
        if os.path.exists(p("model.safetensors.index.json")):
            self.weight_map = json.load(open(p("model.safetensors.index.json")))["weight_map"]
        elif os.path.exists(p("model.safetensors")):
            sf = SafetensorsFile(p("model.safetensors"))
            self._files["model.safetensors"] = sf
            self.weight_map = {k: "model.safetensors" for k in sf.keys()}
        elif os.path.exists(p("pytorch_model.bin.index.json")):
            self.weight_map = json.load(open(p("pytorch_model.bin.index.json")))["weight_map"]
        elif os.path.exists(p("pytorch_model.bin")):
            sd = torch.load(p("pytorch_model.bin"), map_location="cpu", weights_only=True)
            self._files["pytorch_model.bin"] = sd
            self.weight_map = {k: "pytorch_model.bin" for k in sd.keys()}
        else:
            raise FileNotFoundError(f"no model.safetensors / pytorch_model.bin (or index) under {directory}")
        self.shards_opened: List[str] = list(self._files)

    def names(self) -> Iterable[str]:
        return self.weight_map.keys()

    def has(self, name: str) -> bool:
        return name in self.weight_map

    def get(self, name: str) -> torch.Tensor:
        shard = self.weight_map[name]
        f = self._files.get(shard)
        if f is None:
            path = os.path.join(self.dir, shard)
            f = SafetensorsFile(path) if shard.endswith(".safetensors") else torch.load(
                path, map_location="cpu", weights_only=True)
            self._files[shard] = f
            self.shards_opened.append(shard)
        return f.get(name) if isinstance(f, SafetensorsFile) else f[name]
