"""aero-optical-flow_b200: B200 (sm_100a) calcFlow hot path of aero-optical-flow.

The product is csrc/ (CUDA kernels + C ABI, built into libaof_b200.so, declared in
include/aof.h) and cpp/ (header-compatible C++ shims of the reference classes).
`capi` is the ctypes view of the same ABI used by tests and bench.py; `synth`
makes deterministic test frames.  The directory name is not a Python identifier:
load it with __graft_entry__.load_package() (module name `aof_b200`).
"""
from . import capi, sharding, synth  # noqa: F401
from .capi import AofError, Gftt, LkFlow, LkFlowMulti, Px4Flow, Px4FlowMulti, SadDense, lib  # noqa: F401
