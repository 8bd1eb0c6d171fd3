"""Public modules of the framework (same names as the reference's ``src`` package).

``python -m src.main --model ... --splits ... --stage N`` is the command-line surface;
the modules re-implement the reference's client/server/control-plane roles on top of
``petals_b200`` (kernels, engines, DHT/RPC)."""
__version__ = "0.1.0"
