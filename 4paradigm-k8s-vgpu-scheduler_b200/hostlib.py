"""ctypes driver of libvgpuhost.so: the C++ mirror of the extender's Filter path
(host/scheduler.hpp).  Test harness only -- production hosts are the Go extender (cgo)
or C++ programs linking scheduler.hpp directly."""
from __future__ import annotations

import ctypes as C
import os
from typing import Dict, List, Optional, Sequence

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "host", "libvgpuhost.so")
ABSENT = -(1 << 63)
_lib = None


def load():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} is missing: run __graft_entry__.build()")
        from .capi import load_library
        load_library()  # libvgpuscore.so first (rpath also finds it)
        lib = C.CDLL(LIB_PATH)
        lib.vgh_new.restype = C.c_void_p
        lib.vgh_new.argtypes = [C.c_int] * 5
        lib.vgh_free.argtypes = [C.c_void_p]
        lib.vgh_add_node.argtypes = [C.c_void_p, C.c_char_p, C.c_int, C.POINTER(C.c_char_p), C.c_void_p, C.c_void_p,
                                     C.c_void_p, C.POINTER(C.c_char_p), C.c_void_p]
        lib.vgh_del_pod.argtypes = [C.c_void_p, C.c_char_p]
        lib.vgh_register_node.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p]
        lib.vgh_rm_node_devices.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p]
        lib.vgh_on_add_pod.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p, C.c_char_p, C.c_char_p, C.c_char_p, C.c_int]
        lib.vgh_filter.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p, C.c_char_p, C.c_int, C.POINTER(C.c_char_p),
                                   C.POINTER(C.c_char_p), C.c_int, C.c_void_p, C.c_int, C.POINTER(C.c_char_p),
                                   C.c_char_p, C.c_int]
        lib.vgh_last_timing.argtypes = [C.c_void_p, C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(C.c_int)]
        _lib = lib
    return _lib


def _strs(xs: Sequence[str]):
    return (C.c_char_p * max(len(xs), 1))(*[x.encode() for x in xs])


class Scheduler:
    """vgpu::Scheduler (host/scheduler.hpp)"""

    def __init__(self, max_nodes: int, device: int = 0, incremental: bool = False, default_mem: int = 0,
                 default_cores: int = 0):
        self.lib = load()
        self.h = self.lib.vgh_new(device, max_nodes, int(incremental), default_mem, default_cores)

    def close(self):
        if self.h:
            self.lib.vgh_free(self.h)
            self.h = None

    def add_node(self, node_id: str, devices: List[dict]):
        """devices: dicts with ID, Count, Devmem, Devcore, Type, Numa (nodes.go DeviceInfo)"""
        n = len(devices)
        arr = lambda k: np.array([d[k] for d in devices], dtype=np.int32)
        cnt, mem, core, numa = arr("Count"), arr("Devmem"), arr("Devcore"), arr("Numa")
        self.lib.vgh_add_node(self.h, node_id.encode(), n, _strs([d["ID"] for d in devices]), cnt.ctypes.data,
                              mem.ctypes.data, core.ctypes.data, _strs([d["Type"] for d in devices]), numa.ctypes.data)

    def last_timing(self):
        """(total microseconds of the last Filter, microseconds inside libvgpuscore calls, took the resident path)"""
        t, g, f = C.c_double(), C.c_double(), C.c_int()
        self.lib.vgh_last_timing(self.h, C.byref(t), C.byref(g), C.byref(f))
        return t.value, g.value, bool(f.value)

    def del_pod(self, uid: str):
        self.lib.vgh_del_pod(self.h, uid.encode())

    def register_node(self, node_id: str, register_anno: str) -> int:
        """one vendor's node register annotation (scheduler.go:152-222, data half) -> devices that were new"""
        return int(self.lib.vgh_register_node(self.h, node_id.encode(), register_anno.encode()))

    def rm_node_devices(self, node_id: str, register_anno: str):
        self.lib.vgh_rm_node_devices(self.h, node_id.encode(), register_anno.encode())

    def on_add_pod(self, uid: str, node_anno: Optional[str], ids_anno: Optional[str], terminated: bool = False):
        """pod informer add/update event (scheduler.go:72-96)"""
        enc = lambda x: None if x is None else x.encode()
        self.lib.vgh_on_add_pod(self.h, b"default", uid.encode(), uid.encode(), enc(node_anno), enc(ids_anno),
                                int(terminated))

    def filter(self, uid: str, containers: List[Dict[str, int]], annos: Dict[str, str], node_names: List[str]):
        """containers[i]: {"gpu":, "gpumem":, "gpumem-percentage":, "gpucores":, "mlunum":, "mlumem":, "dcunum":,
        "dcumem":, "dcucores":} (absent keys allowed).
        Returns dict(node, ids, failed{}, passed[]) ; raises RuntimeError for the reference's (nil, err)."""
        res = np.full((max(len(containers), 1), 9), ABSENT, dtype=np.int64)
        for i, c in enumerate(containers):
            for k, key in enumerate(("gpu", "gpumem", "gpumem-percentage", "gpucores", "mlunum", "mlumem", "dcunum",
                                     "dcumem", "dcucores")):
                if key in c:
                    res[i, k] = c[key]
        cap = 1 << 16
        out = C.create_string_buffer(cap)
        rc = self.lib.vgh_filter(self.h, b"default", uid.encode(), uid.encode(), len(annos), _strs(list(annos.keys())),
                                 _strs(list(annos.values())), len(containers), res.ctypes.data, len(node_names),
                                 _strs(node_names), out, cap)
        if rc != 0:
            raise RuntimeError(out.value.decode())
        r = {"node": None, "ids": None, "failed": {}, "passed": []}
        for line in out.value.decode().split("\n"):
            if line.startswith("node="):
                r["node"] = line[5:]
            elif line.startswith("ids="):
                r["ids"] = line[4:]
            elif line.startswith("failed="):
                for item in filter(None, line[7:].split(";")):
                    k, v = item.split(":", 1)
                    r["failed"][k] = v
            elif line.startswith("pass="):
                r["passed"] = [x for x in line[5:].split(",") if x]
        return r
