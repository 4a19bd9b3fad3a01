"""ctypes wrapper of oracle/libvgpuoracle.so (TEST INFRASTRUCTURE ONLY).

Takes the same packed Snapshot / PodBatch objects the product harness uses (duck
typed: .planes [8,N,D], .ndev, .type_names / .ctr_begin, .req_begin, .nums, ...),
so the oracle and the GPU read identical bytes.  PARITY UNPINNED: see score_ref.h.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from types import SimpleNamespace
from typing import Optional

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libvgpuoracle.so")
_lib = None

ANNO_USE = "nvidia.com/use-gputype"
ANNO_NOUSE = "nvidia.com/nouse-gputype"
ANNO_NUMA = "nvidia.com/numa-bind"

RESULT_DTYPE = np.dtype([("best_row", "<i4"), ("score", "<f4"), ("fit_count", "<i4"), ("n_picks", "<i4")])
PICK_DTYPE = np.dtype([("slot", "<i4"), ("usedmem", "<i4"), ("usedcores", "<i4")])


class _Annos(C.Structure):
    _fields_ = [("use", C.c_char_p), ("nouse", C.c_char_p), ("numa", C.c_char_p), ("dcu_use", C.c_char_p),
                ("dcu_nouse", C.c_char_p), ("mlu_use", C.c_char_p), ("mlu_nouse", C.c_char_p)]


class _Snap(C.Structure):
    _fields_ = [("n_nodes", C.c_int32), ("dmax", C.c_int32)] + \
               [(k, C.c_void_p) for k in ("used", "count", "usedmem", "totalmem", "usedcores", "totalcore",
                                          "numa", "type_id", "ndev")] + \
               [("n_types", C.c_int32), ("type_names", C.POINTER(C.c_char_p))]


class _Pods(C.Structure):
    _fields_ = [("n_pods", C.c_int32)] + \
               [(k, C.c_void_p) for k in ("ctr_begin", "req_begin", "nums", "memreq", "mempct", "coresreq")] + \
               [("req_type", C.POINTER(C.c_char_p)), ("annos", C.POINTER(_Annos))]


class _PodsPacked(C.Structure):
    _fields_ = [("n_pods", C.c_int32)] + \
               [(k, C.c_void_p) for k in ("ctr_begin", "req_begin", "nums", "memreq", "mempct", "coresreq",
                                          "typemask", "numa_bind")]


def build(force: bool = False) -> str:
    """compile the C restatement (building the checker is not using it)"""
    src = [os.path.join(_HERE, f) for f in ("score_ref.c", "score_ref.h", "upper_table.inc")]
    if force or not os.path.exists(_SO) or any(os.path.getmtime(s) > os.path.getmtime(_SO) for s in src):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return _SO


def _load():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        lib = C.CDLL(_SO)
        lib.ref_score_batch.argtypes = [C.POINTER(_Snap), C.POINTER(_Pods), C.c_void_p, C.c_void_p, C.c_void_p,
                                        C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_double),
                                        C.POINTER(C.c_double)]
        lib.ref_score_batch_packed.argtypes = [C.POINTER(_Snap), C.POINTER(_PodsPacked), C.c_void_p, C.c_void_p,
                                               C.c_void_p, C.c_void_p, C.c_int, C.POINTER(C.c_double)]
        lib.ref_typemask.argtypes = [C.POINTER(_Annos), C.c_char_p, C.c_int32, C.POINTER(C.c_char_p), C.c_int,
                                     C.POINTER(C.c_int), C.POINTER(C.c_int)]
        lib.ref_typemask.restype = C.c_uint64
        lib.ref_encode_pod_devices.argtypes = [C.c_int, C.c_void_p, C.POINTER(C.c_char_p), C.POINTER(C.c_char_p),
                                               C.c_void_p, C.c_void_p, C.c_char_p, C.c_int]
        _lib = lib
    return _lib


def _b(s):
    return None if s is None else s.encode("utf-8")


def _snap_struct(snap):
    names = (C.c_char_p * max(len(snap.type_names), 1))(*[t.encode() for t in snap.type_names])
    planes = [np.ascontiguousarray(snap.planes[f], dtype=np.int32) for f in range(8)]
    ndev = np.ascontiguousarray(snap.ndev, dtype=np.int32)
    s = _Snap(snap.planes.shape[1], snap.planes.shape[2], *[p.ctypes.data for p in planes], ndev.ctypes.data,
              len(snap.type_names), names)
    return s, (names, planes, ndev)


def _annos_struct(an):
    return _Annos(_b(an.get(ANNO_USE)), _b(an.get(ANNO_NOUSE)), _b(an.get(ANNO_NUMA)),
                  _b(an.get("hygon.com/use-dcutype")), _b(an.get("hygon.com/nouse-dcutype")),
                  _b(an.get("cambricon.com/use-mlutype")), _b(an.get("cambricon.com/nouse-mlutype")))


def to_upper(s: bytes) -> bytes:
    """strings.ToUpper as the restatement computes it (any NUL-free bytes)"""
    lib = _load()
    lib.ref_strings_to_upper.argtypes = [C.c_char_p, C.c_size_t, C.c_char_p, C.c_size_t]
    lib.ref_strings_to_upper.restype = C.c_size_t
    n = lib.ref_strings_to_upper(s, len(s), None, 0)
    buf = C.create_string_buffer(n + 1)
    lib.ref_strings_to_upper(s, len(s), buf, n + 1)
    return buf.raw[:n]


def typemask(annos: dict, req_type: str, type_names, memreq: int = 0) -> tuple:
    """(mask, flags): flags bit 0 = numaAssert, bit 1 = in-use "370" cards fail (Cambricon, Memreq == 0)"""
    lib = _load()
    names = (C.c_char_p * max(len(type_names), 1))(*[t.encode() for t in type_names])
    a = _annos_struct(annos)
    numa, idle = C.c_int(0), C.c_int(0)
    m = lib.ref_typemask(C.byref(a), req_type.encode(), int(memreq), names, len(type_names), C.byref(numa),
                         C.byref(idle))
    return int(m), int(numa.value) | (int(idle.value) << 1)


def _outputs(batch, n_nodes, full):
    P = len(batch.ctr_begin) - 1
    results = np.zeros(P, dtype=RESULT_DTYPE)
    picks = np.zeros(int(np.maximum(batch.nums, 0).sum()), dtype=PICK_DTYPE)
    fit = np.zeros((P, n_nodes), dtype=np.uint8) if full else None
    scores = np.zeros((P, n_nodes), dtype=np.float32) if full else None
    return results, picks, fit, scores


def _wrap(results, picks, fit, scores, **extra):
    return SimpleNamespace(best_row=results["best_row"].copy(), score=results["score"].copy(),
                           fit_count=results["fit_count"].copy(), n_picks=results["n_picks"].copy(),
                           picks=np.stack([picks["slot"], picks["usedmem"], picks["usedcores"]], axis=1)
                           if len(picks) else np.zeros((0, 3), np.int32),
                           fit=None if fit is None else fit.astype(bool), scores=scores, **extra)


def score_faithful(snap, batch, full: bool = False, log_io: bool = False):
    """R1: pointer-of-structs + string checkType per device, single thread."""
    lib = _load()
    s, keep = _snap_struct(snap)
    P = len(batch.ctr_begin) - 1
    rt = (C.c_char_p * max(len(batch.nums), 1))(*[t.encode() for t in batch.req_type])
    an = (_Annos * max(P, 1))(*[_annos_struct(a) for a in batch.annos])
    arrs = [np.ascontiguousarray(getattr(batch, k), dtype=np.int32)
            for k in ("ctr_begin", "req_begin", "nums", "memreq", "mempct", "coresreq")]
    pods = _Pods(P, *[a.ctypes.data for a in arrs], rt, an)
    results, picks, fit, scores = _outputs(batch, snap.planes.shape[1], full)
    ts, tb = C.c_double(0), C.c_double(0)
    fd = os.open(os.devnull, os.O_WRONLY) if log_io else -1
    try:
        rc = lib.ref_score_batch(C.byref(s), C.byref(pods), results.ctypes.data, picks.ctypes.data,
                                 None if fit is None else fit.ctypes.data,
                                 None if scores is None else scores.ctypes.data, 1 if log_io else 0, fd,
                                 C.byref(ts), C.byref(tb))
    finally:
        if fd >= 0:
            os.close(fd)
    if rc != 0:
        raise RuntimeError("oracle: unsupported input (dmax > 12 or > 8 requests per container)")
    return _wrap(results, picks, fit, scores, seconds_scoring=ts.value, seconds_rebuild=tb.value)


def score_packed(snap, batch, full: bool = False, threads: Optional[int] = None):
    """R2: packed planes + oracle-built type masks, host threads."""
    lib = _load()
    s, keep = _snap_struct(snap)
    P = len(batch.ctr_begin) - 1
    R = len(batch.nums)
    masks = np.zeros(R, dtype=np.uint64)
    numa = np.zeros(R, dtype=np.uint8)
    cache = {}
    ctr_counts = np.diff(batch.ctr_begin)
    pod_of_ctr = np.repeat(np.arange(P), ctr_counts)
    ctr_of_req = np.repeat(np.arange(len(batch.req_begin) - 1), np.diff(batch.req_begin))
    for r in range(R):
        a = batch.annos[int(pod_of_ctr[ctr_of_req[r]])]
        key = (batch.req_type[r], a.get(ANNO_USE), a.get(ANNO_NOUSE), a.get(ANNO_NUMA),
               a.get("hygon.com/use-dcutype"), a.get("hygon.com/nouse-dcutype"),
               a.get("cambricon.com/use-mlutype"), a.get("cambricon.com/nouse-mlutype"),
               int(batch.memreq[r]) != 0 if "MLU" in batch.req_type[r] else None)
        if key not in cache:
            cache[key] = typemask(a, batch.req_type[r], snap.type_names, int(batch.memreq[r]))
        masks[r], numa[r] = cache[key]
    arrs = [np.ascontiguousarray(getattr(batch, k), dtype=np.int32)
            for k in ("ctr_begin", "req_begin", "nums", "memreq", "mempct", "coresreq")]
    pods = _PodsPacked(P, *[a.ctypes.data for a in arrs], masks.ctypes.data, numa.ctypes.data)
    results, picks, fit, scores = _outputs(batch, snap.planes.shape[1], full)
    sec = C.c_double(0)
    nt = threads or os.cpu_count() or 1
    rc = lib.ref_score_batch_packed(C.byref(s), C.byref(pods), results.ctypes.data, picks.ctypes.data,
                                    None if fit is None else fit.ctypes.data,
                                    None if scores is None else scores.ctypes.data, nt, C.byref(sec))
    if rc != 0:
        raise RuntimeError("oracle: unsupported input")
    return _wrap(results, picks, fit, scores, seconds=sec.value, threads=nt)


def encode_pod_devices(ctrs) -> str:
    """util.EncodePodDevices restated: ctrs[c] = [(uuid, type, mem, cores), ...]"""
    lib = _load()
    flat = [d for c in ctrs for d in c]
    per = np.array([len(c) for c in ctrs], dtype=np.int32)
    mem = np.array([d[2] for d in flat], dtype=np.int32)
    cor = np.array([d[3] for d in flat], dtype=np.int32)
    uu = (C.c_char_p * max(len(flat), 1))(*[d[0].encode() for d in flat])
    tt = (C.c_char_p * max(len(flat), 1))(*[d[1].encode() for d in flat])
    cap = 64 + sum(len(d[0]) + len(d[1]) + 32 for d in flat) + len(ctrs)
    buf = C.create_string_buffer(cap)
    n = lib.ref_encode_pod_devices(len(ctrs), per.ctypes.data, uu, tt, mem.ctypes.data, cor.ctypes.data, buf, cap)
    if n < 0:
        raise RuntimeError("oracle encode: buffer too small")
    return buf.value.decode()
