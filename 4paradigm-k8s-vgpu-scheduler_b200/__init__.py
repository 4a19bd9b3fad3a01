"""B200-native Filter scoring for the 4paradigm k8s vGPU scheduler extender.

Only the hot path lives here: csrc/ holds the sm_100a kernels and
the C ABI (libvgpuscore.so); the Python modules are the ctypes harness that stands
in for the Go cgo shim on hosts without Go.  The directory name is not a legal
Python identifier; import it through ``vgpu_b200`` (repo root, ``vgpu_b200.py``).
"""
from .model import (ANNO_NOUSE, ANNO_NUMA, ANNO_USE, MEMPCT_UNSET, BatchResult, PodBatch, Snapshot)  # noqa: F401
from . import capi, hostlib, sharding, workload  # noqa: F401
from .capi import Context, Group, VgsError, load_library  # noqa: F401
