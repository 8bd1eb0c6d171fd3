"""Imports reference modules (read-only oracle) with stub ``hivemind`` packages."""
import importlib
import logging
import os
import sys
import types

REF_CANDIDATES = [os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "baseline", "_ref"),
                  "/root/reference"]


def reference_root():
    for c in REF_CANDIDATES:
        if os.path.exists(os.path.join(c, "src", "load_balancing.py")):
            return c
    return None


def install_hivemind_stub():
    if "hivemind" in sys.modules and getattr(sys.modules["hivemind"], "_is_stub", False):
        return

    class PeerID(str):
        @classmethod
        def from_base58(cls, s):
            return cls(s)

        def to_base58(self):
            return str(self)

    def mk(name, **attrs):
        m = types.ModuleType(name)
        m.__dict__.update(attrs)
        m._is_stub = True
        sys.modules[name] = m
        return m

    class _Dummy:
        def __init__(self, *a, **k):
            pass

    mk("hivemind", DHT=_Dummy, PeerID=PeerID, get_logger=logging.getLogger, get_dht_time=lambda: 0.0,
       P2PContext=_Dummy, deserialize_torch_tensor=None, serialize_torch_tensor=None, MSGPackSerializer=_Dummy)
    mk("hivemind.p2p", P2P=_Dummy, P2PContext=_Dummy, PeerID=PeerID)
    mk("hivemind.utils", get_logger=logging.getLogger)
    mk("hivemind.utils.logging", get_logger=logging.getLogger)
    mk("hivemind.compression", serialize_torch_tensor=None, deserialize_torch_tensor=None)
    mk("hivemind.compression.serialization", serialize_torch_tensor=None, deserialize_torch_tensor=None,
       deserialize_tensor_stream=None)
    mk("hivemind.moe", )
    mk("hivemind.moe.server", )
    mk("hivemind.moe.server.connection_handler", ConnectionHandler=_Dummy)
    mk("hivemind.proto", runtime_pb2=types.SimpleNamespace(ExpertRequest=_Dummy, ExpertResponse=_Dummy))
    mk("hivemind.utils.serializer", MSGPackSerializer=_Dummy)
    mk("hivemind.utils.asyncio", aiter_with_timeout=None, iter_as_aiter=None)
    mk("hivemind.utils.streaming", split_for_streaming=None)
    mk("async_timeout", timeout=_Dummy)


def load_reference_module(name: str):
    """e.g. 'load_balancing' -> the reference's src/load_balancing.py as an isolated module."""
    root = reference_root()
    if root is None:
        return None
    install_hivemind_stub()
    path = os.path.join(root, "src", name + ".py")
    modname = f"_refsrc_{name}"
    if modname in sys.modules:
        return sys.modules[modname]
    pkg = types.ModuleType("_refsrc")
    pkg.__path__ = [os.path.join(root, "src")]
    sys.modules.setdefault("_refsrc", pkg)
    spec = importlib.util.spec_from_file_location(f"_refsrc.{name}", path, submodule_search_locations=None)
    mod = importlib.util.module_from_spec(spec)
    mod.__package__ = "_refsrc"
    sys.modules[f"_refsrc.{name}"] = mod
    try:
        spec.loader.exec_module(mod)
    except Exception:  # noqa: BLE001
        sys.modules.pop(f"_refsrc.{name}", None)
        return None
    sys.modules[modname] = mod
    return mod
