"""Device-to-device hand-off of activations between server PROCESSES that share one NVSwitch box.

The reference moves every hidden state through the client over TCP (client -> stage 1 -> client -> stage 2 ...,
src/rpc_transport.py:399-465, tensors serialised by hivemind); between two GPUs of one box that is a device->host
copy, two socket writes and a host->device copy per hop.  Here a stage that knows its consumer lives in the same box
answers with a *descriptor* instead of bytes: it copies its output into a ring slot of an arena it exported once with
CUDA IPC, and the RPC reply (or the pushed request) carries only ``{handle, offset, shape, dtype, seq}``.  The consumer
opens the handle once (cached), and pulls the slot with one peer copy over NVLink straight into its own memory.  DHT,
routing, journaling, recovery, batching and sampling stay exactly as they are - the TCP messages just stop carrying
the payload.

Every slot is framed ``[seq | payload | seq]``: the ring is reused, so a consumer that comes very late (a client that
re-routes after a failure) could find its slot overwritten.  The pulled copy is checked against the descriptor's
``seq`` and a mismatch raises :class:`StaleDeviceTensor`, which the client treats like any failed hop (it replays from
its journal, which it filled from its own pull at the time the descriptor arrived).
"""
from __future__ import annotations

import os
import threading
from typing import Dict, Optional, Tuple

import torch

_DTYPES = {"float16": torch.float16, "bfloat16": torch.bfloat16, "float32": torch.float32, "int64": torch.int64}
_HDR = 16            # bytes of the seq word in front of / behind the payload (keeps payloads 16-byte aligned)


class StaleDeviceTensor(RuntimeError):
    """The producer's ring slot no longer holds the tensor the descriptor names."""


def box_id() -> str:
    """Identifies 'the same machine': CUDA IPC handles only mean something inside one OS instance."""
    try:
        with open("/proc/sys/kernel/random/boot_id") as f:
            return f.read().strip()
    except OSError:
        import socket
        return socket.gethostname()


def _rt():
    from cuda.bindings import driver, runtime
    return driver, runtime


def _check(err, what: str):
    if int(err) != 0:
        raise RuntimeError(f"{what} failed with CUDA error {int(err)}")


class DeviceExporter:
    """Producer side: a ring arena on this GPU whose slots other processes of the box can pull."""

    def __init__(self, device: torch.device, arena_bytes: int = 64 << 20):
        self.device = torch.device(device)
        driver, runtime = _rt()
        with torch.cuda.device(self.device):
            # a private allocation (not a slice of the caching allocator's pool), so that the exported range holds
            # nothing but hand-off slots
            err, ptr = runtime.cudaMalloc(arena_bytes)
            _check(err, "cudaMalloc(arena)")
            err, handle = runtime.cudaIpcGetMemHandle(ptr)
            _check(err, "cudaIpcGetMemHandle")
        self.base = int(ptr)
        self.bytes = arena_bytes
        self.handle = bytes(handle.reserved)
        self.arena = _wrap(self.base, arena_bytes, self.device)          # uint8 view over the arena
        self.head = 0
        self.seq = 0
        self.lock = threading.Lock()
        self.box = box_id()

    def close(self) -> None:
        """Free the arena (importers that still hold it open keep their mapping until they close it)."""
        if self.base:
            _, runtime = _rt()
            self.arena = None
            with torch.cuda.device(self.device):
                torch.cuda.synchronize(self.device)
                runtime.cudaFree(self.base)
            self.base = 0

    def export(self, t: torch.Tensor) -> dict:
        """Copy ``t`` into the next slot and describe it.  Returns after the copy has completed on the device."""
        if not t.is_cuda:
            raise ValueError("only device tensors are handed off")
        t = t.detach().contiguous()
        n = t.numel() * t.element_size()
        need = _HDR + ((n + 15) // 16) * 16 + _HDR
        if need > self.bytes:
            raise ValueError(f"tensor of {n} bytes does not fit the {self.bytes}-byte hand-off arena")
        with self.lock:
            if self.head + need > self.bytes:
                self.head = 0
            off = self.head
            self.head += need
            self.seq += 1
            seq = self.seq
            slot = self.arena[off:off + need]
            tail = _HDR + ((n + 15) // 16) * 16
            slot[:_HDR].view(torch.int64).fill_(seq)
            slot[_HDR:_HDR + n].copy_(t.reshape(-1).view(torch.uint8))
            slot[tail:tail + _HDR].view(torch.int64).fill_(seq)
            torch.cuda.current_stream(self.device).synchronize()
        return {"box": self.box, "pid": os.getpid(), "gpu": _physical_index(self.device), "handle": self.handle,
                "arena_bytes": self.bytes, "offset": off, "nbytes": n, "shape": list(t.shape),
                "dtype": str(t.dtype).replace("torch.", ""), "seq": seq}


class DeviceImporter:
    """Consumer side: opens exporters' arenas (once each) and pulls slots into local tensors."""

    def __init__(self, device: torch.device):
        self.device = torch.device(device)
        self.box = box_id()
        self._open: Dict[Tuple[int, bytes], Tuple[int, int]] = {}
        self.lock = threading.Lock()
        self.pulled_bytes = 0
        self.pulls = 0

    def close(self) -> None:
        """Unmap every arena this process opened."""
        _, runtime = _rt()
        with self.lock:
            for ptr, src in self._open.values():
                with torch.cuda.device(src):
                    runtime.cudaIpcCloseMemHandle(ptr)
            self._open.clear()

    def accepts(self, desc: Optional[dict]) -> bool:
        return bool(desc) and desc.get("box") == self.box and self.device.type == "cuda"

    def _arena(self, desc: dict) -> Tuple[int, int]:
        """(base address of the exporter's arena in THIS process, its device index here); opened once per exporter."""
        key = (int(desc["pid"]), bytes(desc["handle"]))
        with self.lock:
            a = self._open.get(key)
            if a is None:
                if int(desc["pid"]) == os.getpid():
                    raise RuntimeError("a process does not import its own arena")
                driver, runtime = _rt()
                h = runtime.cudaIpcMemHandle_t()
                h.reserved = bytes(desc["handle"])
                # CUDA IPC maps the allocation into the context of the device it physically lives on: open it with THAT
                # device current (as torch's own CUDA IPC does), then copy across with cudaMemcpyPeerAsync.  (Opened
                # under the consumer's device instead, every access to the mapping is an illegal address.)
                src = local_index(int(desc["gpu"]))
                with torch.cuda.device(src):
                    err, ptr = runtime.cudaIpcOpenMemHandle(h, runtime.cudaIpcMemLazyEnablePeerAccess)
                    _check(err, "cudaIpcOpenMemHandle")
                a = self._open[key] = (int(ptr), src)
                if len(self._open) > 64:            # exporters come and go (client processes): unmap the oldest
                    old_ptr, old_src = self._open.pop(next(iter(self._open)))
                    with torch.cuda.device(old_src):
                        runtime.cudaIpcCloseMemHandle(old_ptr)
            return a

    def pull(self, desc: dict, local: Optional[torch.Tensor] = None) -> torch.Tensor:
        """One peer copy of the slot (frame included) into this GPU's memory, then the frame check."""
        if desc.get("box") != self.box:
            raise RuntimeError("device descriptor from another machine")
        n, off = int(desc["nbytes"]), int(desc["offset"])
        tail = _HDR + ((n + 15) // 16) * 16
        total = tail + _HDR
        if off < 0 or off + total > int(desc["arena_bytes"]):
            raise ValueError("hand-off descriptor points outside its arena")
        buf = torch.empty(total, dtype=torch.uint8, device=self.device)
        if local is not None and int(desc["pid"]) == os.getpid():
            buf.copy_(local[off:off + total])                       # producer and consumer in one process (tests)
        else:
            base, src = self._arena(desc)
            _, runtime = _rt()
            dst = self.device.index if self.device.index is not None else torch.cuda.current_device()
            with torch.cuda.device(self.device):
                err, = runtime.cudaMemcpyPeerAsync(buf.data_ptr(), dst, base + off, src, total,
                                                   torch.cuda.current_stream(self.device).cuda_stream)
                _check(err, "cudaMemcpyPeerAsync(hand-off slot)")
        seqs = torch.stack([buf[:8].view(torch.int64)[0], buf[tail:tail + 8].view(torch.int64)[0]]).tolist()
        if seqs[0] != int(desc["seq"]) or seqs[1] != int(desc["seq"]):
            raise StaleDeviceTensor(f"hand-off slot holds seq {seqs}, descriptor says {desc['seq']}")
        self.pulls += 1
        self.pulled_bytes += n
        return buf[_HDR:_HDR + n].view(_DTYPES[desc["dtype"]]).reshape(desc["shape"])


def visible_gpus() -> list:
    """Physical indices of the GPUs this process can address, in its own numbering order."""
    vis = os.environ.get("CUDA_VISIBLE_DEVICES")
    if vis:
        try:
            return [int(v) for v in vis.split(",") if v.strip() != ""]
        except ValueError:
            pass
    return list(range(torch.cuda.device_count()))


def local_index(physical: int) -> int:
    """This process's device index of the GPU with the given physical index (raises when it is not visible here)."""
    vis = visible_gpus()
    if physical not in vis:
        raise RuntimeError(f"GPU {physical} of the exporting process is not visible here (CUDA_VISIBLE_DEVICES)")
    return vis.index(physical)


def _physical_index(device: torch.device) -> int:
    idx = device.index if device.index is not None else torch.cuda.current_device()
    vis = visible_gpus()
    return vis[idx] if idx < len(vis) else idx


class _RawCuda:
    """Minimal ``__cuda_array_interface__`` carrier so torch can view a raw device pointer without copying."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 3,
                                         "strides": None}


def _wrap(ptr: int, nbytes: int, device: Optional[torch.device]) -> torch.Tensor:
    t = torch.as_tensor(_RawCuda(ptr, nbytes), device=device) if device is not None else torch.as_tensor(_RawCuda(ptr, nbytes))
    return t
