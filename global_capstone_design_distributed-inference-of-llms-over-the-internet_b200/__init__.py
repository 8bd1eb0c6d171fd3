"""Blackwell-native pipeline-parallel LLM inference (Mini-Petals capabilities).

Sub-packages
  models/    architecture configs + plain-torch restatements (CPU path and oracle)
  ops/       ctypes bindings of the sm_100a kernel library (csrc/), weight packing
  runtime/   session KV cache, stage engines (torch + CUDA-graph native), scheduler,
             memory budget, task pools
  parallel/  NCCL p2p pipeline chain, tensor-parallel spans over peer memory
  weights/   span-selective HF checkpoint reader, random init, fp8 quantiser
  net/       in-process/TCP DHT and framed-TCP RPC (cross-host fallback)
  utils/     logging, timing, metrics

Import it as ``petals_b200`` (see ``petals_b200/__init__.py``).
"""
__version__ = "0.1.0"
