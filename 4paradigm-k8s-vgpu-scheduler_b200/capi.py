"""ctypes binding of libvgpuscore.so -- the same calls, in the same order, that the
Go shim makes through cgo.  There is no fallback: if the shared
library is missing or no GPU is usable, this module raises.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Dict, List, Optional, Tuple

import numpy as np

from .model import ANNO_DCU_NOUSE, ANNO_DCU_USE, ANNO_MLU_NOUSE, ANNO_MLU_USE, ANNO_NOUSE, ANNO_NUMA, ANNO_USE, BatchResult, PodBatch, Snapshot

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "csrc", "libvgpuscore.so")

VGS_OK = 0
VGS_EINVAL, VGS_ECUDA, VGS_ETOOBIG, VGS_EDMAX, VGS_ERANGE, VGS_ESTATE, VGS_EINTERNAL = range(-1, -8, -1)
VGS_ELIMIT = -8
VGS_REQ_NUMA_BIND = 1
VGS_REQ_IDLE_SHARE = 2
VGS_TYPE_IDLE_SHARE = 1
NOALLOC_MAX_NODES, NOALLOC_MAX_DEVICES, NOALLOC_MAX_PODS, NOALLOC_MAX_REQS = 16384, 8, 256, 1024

# every symbol include/vgpuscore.h declares
EXPORTS = (
    "vgs_abi_version", "vgs_strerror", "vgs_init", "vgs_destroy", "vgs_last_error",
    "vgs_snapshot_upload", "vgs_snapshot_set_device", "vgs_snapshot_refresh", "vgs_snapshot_delta", "vgs_snapshot_download",
    "vgs_snapshot_type_flags", "vgs_host_typemask_mlu", "vgs_host_type_flags", "vgs_host_decode_pod_devices",
    "vgs_exchange_block", "vgs_resolve_local", "vgs_exchange_merge", "vgs_snapshot_upload_async",
    "vgs_score_batch", "vgs_drain_batch", "vgs_pods_upload", "vgs_score_resident", "vgs_resolve_resident",
    "vgs_results_download", "vgs_sync", "vgs_keys_device", "vgs_keys_merge", "vgs_results_device",
    "vgs_verify_download", "vgs_last_timing", "vgs_host_typemask", "vgs_host_numa_bind",
    "vgs_host_encode_pod_devices", "vgs_host_decode_node_devices", "vgs_score_batch_noalloc",
    "vgs_tuning_defaults", "vgs_last_plan", "vgs_snapshot_upload_narrow", "vgs_host_to_upper",
    "vgs_group_init", "vgs_group_destroy", "vgs_group_last_error", "vgs_group_snapshot_upload",
    "vgs_group_snapshot_type_flags", "vgs_group_snapshot_delta", "vgs_group_score_batch", "vgs_group_last_timing",
)


class VgsTuning(C.Structure):
    _fields_ = [("slice_min", C.c_int32), ("slice_small", C.c_int32), ("slice_small_pairs", C.c_int64),
                ("node_words", C.c_int32), ("tables_in_smem", C.c_int32), ("tables_smem_kb", C.c_int32),
                ("scout_stride", C.c_int32), ("scout_split", C.c_int32), ("class_order", C.c_int32),
                ("zero_copy_upload", C.c_int32), ("general_legacy", C.c_int32), ("max_ctas", C.c_int32),
                ("filter_one_launch", C.c_int32), ("drain_chunk", C.c_int32), ("alloc_pad_kb", C.c_int32),
                ("scouts_node_sliced", C.c_int32), ("slice_min_nodes", C.c_int32)]


class VgsConfig(C.Structure):
    _fields_ = [("device", C.c_int32), ("max_nodes", C.c_int32), ("max_devices", C.c_int32),
                ("max_pods", C.c_int32), ("max_reqs", C.c_int32), ("max_picks", C.c_int32),
                ("row_base", C.c_int32), ("verify", C.c_int32), ("stream", C.c_void_p),
                ("tuning", C.POINTER(VgsTuning))]


class VgsClassPlan(C.Structure):
    _fields_ = [("pods", C.c_int32), ("kernel", C.c_int32), ("words_pod_sliced", C.c_int32),
                ("words_node_sliced", C.c_int32), ("tables_in_smem", C.c_int32), ("scout_stride", C.c_int32),
                ("scout_split", C.c_int32), ("tiles_per_cta_scout", C.c_int32), ("tiles_per_cta_main", C.c_int32),
                ("launches", C.c_int32)]


class VgsPlan(C.Structure):
    _fields_ = [("cls", VgsClassPlan * 4), ("dict_full", C.c_int32), ("drain_chunks", C.c_int32),
                ("drain_rescans", C.c_int32), ("drain_chain_pods", C.c_int32)]


KERNEL_NONE, KERNEL_PAIR, KERNEL_SLICED, KERNEL_GENERAL = range(4)
VGS_ENCCL = -9
EXCHANGE_NCCL, EXCHANGE_PEER = 0, 1


class VgsGroupConfig(C.Structure):
    _fields_ = [("n_devices", C.c_int32), ("devices", C.c_int32 * 8), ("max_nodes", C.c_int32),
                ("max_devices", C.c_int32), ("max_pods", C.c_int32), ("max_reqs", C.c_int32), ("max_picks", C.c_int32),
                ("exchange", C.c_int32), ("tuning", C.POINTER(VgsTuning))]


class VgsTiming(C.Structure):
    _fields_ = [("snapshot_h2d_ms", C.c_float), ("derive_ms", C.c_float), ("pods_h2d_ms", C.c_float),
                ("score_ms", C.c_float), ("resolve_ms", C.c_float), ("d2h_ms", C.c_float),
                ("kernel_launches", C.c_int32), ("reserved", C.c_int32)]


REQ_DTYPE = np.dtype([("ctr", "<i4"), ("nums", "<i4"), ("memreq", "<i4"), ("mempct", "<i4"),
                      ("coresreq", "<i4"), ("flags", "<u4"), ("typemask", "<u8")], align=True)
POD_DTYPE = np.dtype([("req_begin", "<i4"), ("n_reqs", "<i4")])
RESULT_DTYPE = np.dtype([("best_row", "<i4"), ("score", "<f4"), ("fit_count", "<i4"), ("n_picks", "<i4")])
PICK_DTYPE = np.dtype([("slot", "<i4"), ("usedmem", "<i4"), ("usedcores", "<i4")])
DELTA_DTYPE = np.dtype([("row", "<i4"), ("slot", "<i4"), ("d_used", "<i4"), ("d_usedmem", "<i4"),
                        ("d_usedcores", "<i4")])
assert REQ_DTYPE.itemsize == 32 and RESULT_DTYPE.itemsize == 16 and PICK_DTYPE.itemsize == 12

_lib = None


class VgsError(RuntimeError):
    def __init__(self, code: int, text: str):
        super().__init__(f"libvgpuscore: {text} (status {code})")
        self.code = code


def load_library() -> C.CDLL:
    """dlopen the in-tree library; raise loudly when it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                           "(there is no CPU fallback for the scoring path)")
    lib = C.CDLL(LIB_PATH)
    vp, i32, u64p, i32p = C.c_void_p, C.c_int32, C.POINTER(C.c_uint64), C.POINTER(C.c_int32)
    lib.vgs_abi_version.restype = C.c_int
    lib.vgs_strerror.restype = C.c_char_p
    lib.vgs_strerror.argtypes = [C.c_int]
    lib.vgs_init.argtypes = [C.POINTER(VgsConfig), C.POINTER(vp)]
    lib.vgs_destroy.argtypes = [vp]
    lib.vgs_destroy.restype = None
    lib.vgs_last_error.argtypes = [vp]
    lib.vgs_last_error.restype = C.c_char_p
    lib.vgs_snapshot_upload.argtypes = [vp, i32, i32, C.POINTER(vp), vp, u64p]
    lib.vgs_snapshot_set_device.argtypes = [vp, i32, i32, C.POINTER(vp), vp, u64p]
    lib.vgs_snapshot_upload_async.argtypes = [vp, i32, i32, C.POINTER(vp), vp, u64p]
    lib.vgs_snapshot_upload_narrow.argtypes = [vp, i32, i32, C.POINTER(vp), vp, vp, i32, u64p]
    lib.vgs_snapshot_refresh.argtypes = [vp]
    lib.vgs_snapshot_type_flags.argtypes = [vp, i32, vp]
    lib.vgs_host_typemask_mlu.argtypes = [i32, C.c_char_p, C.c_char_p, C.POINTER(C.c_char_p), i32,
                                          C.POINTER(C.c_uint32)]
    lib.vgs_host_typemask_mlu.restype = C.c_uint64
    lib.vgs_host_to_upper.argtypes = [C.c_char_p, C.c_char_p, i32]
    lib.vgs_host_to_upper.restype = i32
    lib.vgs_host_type_flags.argtypes = [C.c_char_p]
    lib.vgs_host_type_flags.restype = C.c_uint32
    lib.vgs_snapshot_delta.argtypes = [vp, i32, vp]
    lib.vgs_snapshot_download.argtypes = [vp, C.POINTER(vp), vp]
    lib.vgs_score_batch.argtypes = [vp, i32, vp, i32, vp, vp, vp]
    lib.vgs_drain_batch.argtypes = [vp, i32, vp, i32, vp, vp, vp]
    lib.vgs_pods_upload.argtypes = [vp, i32, vp, i32, vp]
    lib.vgs_score_resident.argtypes = [vp]
    lib.vgs_resolve_resident.argtypes = [vp]
    lib.vgs_results_download.argtypes = [vp, vp, vp]
    lib.vgs_sync.argtypes = [vp]
    lib.vgs_keys_device.argtypes = [vp, C.POINTER(vp), C.POINTER(vp)]
    lib.vgs_keys_merge.argtypes = [vp, vp, vp, i32, i32, i32]
    lib.vgs_results_device.argtypes = [vp, C.POINTER(vp), C.POINTER(vp)]
    lib.vgs_exchange_block.argtypes = [vp, C.POINTER(vp), u64p]
    lib.vgs_resolve_local.argtypes = [vp]
    lib.vgs_exchange_merge.argtypes = [vp, vp, i32, C.c_uint64]
    lib.vgs_verify_download.argtypes = [vp, vp, vp]
    lib.vgs_last_timing.argtypes = [vp, C.POINTER(VgsTiming)]
    lib.vgs_last_plan.argtypes = [vp, C.POINTER(VgsPlan)]
    lib.vgs_group_init.argtypes = [C.POINTER(VgsGroupConfig), C.POINTER(vp)]
    lib.vgs_group_destroy.argtypes = [vp]
    lib.vgs_group_destroy.restype = None
    lib.vgs_group_last_error.argtypes = [vp]
    lib.vgs_group_last_error.restype = C.c_char_p
    lib.vgs_group_snapshot_upload.argtypes = [vp, i32, i32, C.POINTER(vp), vp]
    lib.vgs_group_snapshot_type_flags.argtypes = [vp, i32, vp]
    lib.vgs_group_snapshot_delta.argtypes = [vp, i32, vp]
    lib.vgs_group_score_batch.argtypes = [vp, i32, vp, i32, vp, vp, vp]
    lib.vgs_group_last_timing.argtypes = [vp, C.POINTER(C.c_double), C.POINTER(C.c_double)]
    lib.vgs_tuning_defaults.argtypes = [C.POINTER(VgsTuning)]
    lib.vgs_tuning_defaults.restype = None
    lib.vgs_score_batch_noalloc.argtypes = [i32, vp, i32, i32, i32, C.POINTER(vp), vp, i32, vp, i32, vp, vp, vp]
    lib.vgs_host_typemask.argtypes = [C.c_char_p, C.c_char_p, C.c_char_p, C.POINTER(C.c_char_p), i32]
    lib.vgs_host_typemask.restype = C.c_uint64
    lib.vgs_host_numa_bind.argtypes = [C.c_char_p]
    lib.vgs_host_numa_bind.restype = i32
    lib.vgs_host_encode_pod_devices.argtypes = [i32, vp, C.POINTER(C.c_char_p), C.POINTER(C.c_char_p), vp,
                                                C.c_char_p, i32]
    lib.vgs_host_encode_pod_devices.restype = i32
    _lib = lib
    return lib


def _b(s: Optional[str]) -> Optional[bytes]:
    return None if s is None else s.encode("utf-8")


def host_typemask(req_type: str, annos: Dict[str, str], type_names: List[str]) -> int:
    lib = load_library()
    arr = (C.c_char_p * len(type_names))(*[t.encode("utf-8") for t in type_names])
    use, nouse = (ANNO_DCU_USE, ANNO_DCU_NOUSE) if req_type == "DCU" else (ANNO_USE, ANNO_NOUSE)
    return int(lib.vgs_host_typemask(_b(req_type), _b(annos.get(use)), _b(annos.get(nouse)), arr, len(type_names)))


def host_typemask_mlu(memreq: int, annos: Dict[str, str], type_names: List[str]) -> Tuple[int, int]:
    """Cambricon request -> (typemask, request flags)"""
    lib = load_library()
    arr = (C.c_char_p * len(type_names))(*[t.encode("utf-8") for t in type_names])
    fl = C.c_uint32()
    m = lib.vgs_host_typemask_mlu(int(memreq), _b(annos.get(ANNO_MLU_USE)), _b(annos.get(ANNO_MLU_NOUSE)), arr,
                                  len(type_names), C.byref(fl))
    return int(m), int(fl.value)


def host_to_upper(s: bytes) -> bytes:
    """vgs_host_to_upper: Go 1.19 strings.ToUpper of a byte string (any bytes, NUL-free)"""
    lib = load_library()
    n = lib.vgs_host_to_upper(s, None, 0)
    buf = C.create_string_buffer(n + 1)
    lib.vgs_host_to_upper(s, buf, n + 1)
    return buf.raw[:n]


def host_type_flags(type_name: str) -> int:
    return int(load_library().vgs_host_type_flags(type_name.encode("utf-8")))


def host_numa_bind(annos: Dict[str, str]) -> int:
    return int(load_library().vgs_host_numa_bind(_b(annos.get(ANNO_NUMA))))


def decode_node_devices(anno: str, max_dev: int = 64):
    """util.DecodeNodeDevices -> list of dicts (ID, Count, Devmem, Devcore, Type, Numa, Health); raises on
    what the reference rejects"""
    lib = load_library()
    raw = anno.encode("utf-8")
    cols = [np.zeros(max_dev, dtype=np.int32) for _ in range(9)]
    lib.vgs_host_decode_node_devices.restype = C.c_int32
    n = lib.vgs_host_decode_node_devices(C.c_char_p(raw), max_dev, *[C.c_void_p(c.ctypes.data) for c in cols])
    if n < 0:
        raise VgsError(n, "node annotations not decode successfully")
    cnt, mem, core, numa, health, io, il, to, tl = cols
    return [dict(ID=raw[io[i]:io[i] + il[i]].decode(), Count=int(cnt[i]), Devmem=int(mem[i]), Devcore=int(core[i]),
                 Type=raw[to[i]:to[i] + tl[i]].decode(), Numa=int(numa[i]), Health=bool(health[i])) for i in range(n)]


def decode_pod_devices(anno: str, max_ctrs: int = 64, max_dev: int = 1024) -> List[List[Tuple[str, str, int, int]]]:
    """util.DecodePodDevices: the vgpu-ids-new annotation -> [[(uuid, type, usedmem, usedcores), ...], ...]
    (a malformed entry gives [] like the reference, which drops the error)"""
    lib = load_library()
    raw = anno.encode("utf-8")
    ctr_n = np.zeros(max(max_ctrs, 1), dtype=np.int32)
    cols = [np.zeros(max(max_dev, 1), dtype=np.int32) for _ in range(6)]
    lib.vgs_host_decode_pod_devices.restype = C.c_int32
    nc = lib.vgs_host_decode_pod_devices(C.c_char_p(raw), max_ctrs, max_dev, C.c_void_p(ctr_n.ctypes.data),
                                         *[C.c_void_p(c.ctypes.data) for c in cols])
    if nc < 0:
        raise VgsError(nc, "decode_pod_devices")
    uo, ul, to, tl, mem, cores = cols
    out, d = [], 0
    for c in range(nc):
        devs = []
        for _ in range(int(ctr_n[c])):
            devs.append((raw[uo[d]:uo[d] + ul[d]].decode(), raw[to[d]:to[d] + tl[d]].decode(), int(mem[d]), int(cores[d])))
            d += 1
        out.append(devs)
    return out


def encode_pod_devices(ctrs: List[List[Tuple[str, str, int, int]]]) -> str:
    """ctrs[c] = [(uuid, type, usedmem, usedcores), ...] -> the vgpu-ids-new annotation value"""
    lib = load_library()
    flat = [d for c in ctrs for d in c]
    per = np.array([len(c) for c in ctrs], dtype=np.int32)
    picks = np.zeros(len(flat), dtype=PICK_DTYPE)
    for i, d in enumerate(flat):
        picks[i] = (0, d[2], d[3])
    uu = (C.c_char_p * max(len(flat), 1))(*[d[0].encode() for d in flat])
    tt = (C.c_char_p * max(len(flat), 1))(*[d[1].encode() for d in flat])
    cap = 64 + sum(len(d[0]) + len(d[1]) + 32 for d in flat) + len(ctrs)
    buf = C.create_string_buffer(cap)
    n = lib.vgs_host_encode_pod_devices(len(ctrs), per.ctypes.data, uu, tt, picks.ctypes.data, buf, cap)
    if n < 0:
        raise VgsError(n, "encode_pod_devices")
    return buf.value.decode()


def pack_batch(batch: PodBatch, type_names: List[str]):
    """PodBatch -> (vgs_pod[], vgs_req[]) exactly as the Go shim would fill them."""
    P = batch.n_pods
    pods = np.zeros(P, dtype=POD_DTYPE)
    reqs = np.zeros(batch.n_reqs, dtype=REQ_DTYPE)
    reqs["nums"], reqs["memreq"], reqs["mempct"], reqs["coresreq"] = (batch.nums, batch.memreq,
                                                                      batch.mempct, batch.coresreq)
    cache: Dict[tuple, int] = {}
    numa_cache: Dict[Optional[str], int] = {}
    ctr_of_req = np.zeros(batch.n_reqs, dtype=np.int32)
    n_ctrs = len(batch.req_begin) - 1
    if n_ctrs:
        counts = np.diff(batch.req_begin)
        ctr_of_req = np.repeat(np.arange(n_ctrs, dtype=np.int32), counts)
    pod_of_ctr = np.repeat(np.arange(P, dtype=np.int32), np.diff(batch.ctr_begin)) if P else np.zeros(0, np.int32)
    first_ctr = batch.ctr_begin[:-1]
    for p in range(P):
        r0, r1 = batch.pod_req_range(p)
        pods[p] = (r0, r1 - r0)
    if batch.n_reqs:
        pod_of_req = pod_of_ctr[ctr_of_req]
        reqs["ctr"] = ctr_of_req - first_ctr[pod_of_req]
        for r in range(batch.n_reqs):
            an = batch.annos[int(pod_of_req[r])]
            if batch.req_type[r] == "MLU":  # the mask depends on Memreq (cambricon/device.go:94-99)
                key = ("MLU", int(batch.memreq[r]) != 0, an.get(ANNO_MLU_USE), an.get(ANNO_MLU_NOUSE))
                mf = cache.get(key)
                if mf is None:
                    mf = host_typemask_mlu(int(batch.memreq[r]), an, type_names)
                    cache[key] = mf
                reqs["typemask"][r], reqs["flags"][r] = mf
                continue
            key = (batch.req_type[r], an.get(ANNO_USE), an.get(ANNO_NOUSE), an.get(ANNO_DCU_USE), an.get(ANNO_DCU_NOUSE))
            m = cache.get(key)
            if m is None:
                m = host_typemask(batch.req_type[r], an, type_names)
                cache[key] = m
            reqs["typemask"][r] = m
            nb = an.get(ANNO_NUMA)
            f = numa_cache.get(nb)
            if f is None:
                # numaAssert comes from the NVIDIA vendor only (nvidia/device.go:107-112)
                f = host_numa_bind(an)
                numa_cache[nb] = f
            reqs["flags"][r] = VGS_REQ_NUMA_BIND if (f and batch.req_type[r] == "NVIDIA") else 0
    return pods, reqs


def score_batch_noalloc(device: int, stream: int, n_nodes: int, dmax: int, row_base: int, plane_ptrs: List[int],
                        ndev_ptr: int, pods: np.ndarray, reqs: np.ndarray, results_ptr: int, picks_ptr: int) -> int:
    """vgs_score_batch_noalloc: planes / outputs are addresses of device memory the caller
    owns; returns the status code (0 = ok) instead of raising so callers can test limits."""
    lib = load_library()
    ptrs = (C.c_void_p * 8)(*plane_ptrs)
    return int(lib.vgs_score_batch_noalloc(device, C.c_void_p(stream) if stream else None, n_nodes, dmax, row_base,
                                           ptrs, C.c_void_p(ndev_ptr), len(pods),
                                           pods.ctypes.data if len(pods) else None, len(reqs),
                                           reqs.ctypes.data if len(reqs) else None, C.c_void_p(results_ptr),
                                           C.c_void_p(picks_ptr)))


def narrow_planes(snap: Snapshot) -> List[np.ndarray]:
    """[8 planes, ndev] of a snapshot, each in the narrowest of int8 / int16 / int32 that holds its values -- what a
    host packer filling the planes from the DeviceUsage structs would write for vgs_snapshot_upload_narrow"""
    out = []
    for a in [snap.planes[f] for f in range(8)] + [snap.ndev]:
        lo, hi = (int(a.min()), int(a.max())) if a.size else (0, 0)
        dt = np.int8 if -128 <= lo and hi <= 127 else (np.int16 if -32768 <= lo and hi <= 32767 else np.int32)
        out.append(np.ascontiguousarray(a, dtype=dt))
    return out


class Context:
    """One GPU, one shard of the node set.  Thin, 1:1 over the C ABI."""

    def __init__(self, max_nodes: int, max_devices: int = 8, max_pods: int = 4096, max_reqs: Optional[int] = None,
                 max_picks: Optional[int] = None, device: int = 0, row_base: int = 0, verify: bool = False,
                 stream: int = 0, tuning: Optional[Dict[str, int]] = None):
        """tuning: vgs_tuning fields to change from their defaults, e.g. {"slice_min": 1, "node_words": 0}"""
        self.lib = load_library()
        max_reqs = max_reqs or max_pods * 4
        max_picks = max_picks or max_reqs * 4
        tune = VgsTuning()
        self.lib.vgs_tuning_defaults(C.byref(tune))
        for k, val in (tuning or {}).items():
            if k not in dict((f[0], 1) for f in VgsTuning._fields_):
                raise KeyError(f"vgs_tuning has no field {k!r}")
            setattr(tune, k, int(val))
        self.tuning = tune
        cfg = VgsConfig(device, max(max_nodes, 1), max_devices, max_pods, max_reqs, max_picks, row_base,
                        1 if verify else 0, C.c_void_p(stream) if stream else None, C.pointer(tune))
        self.cfg = cfg
        self.h = C.c_void_p()
        rc = self.lib.vgs_init(C.byref(cfg), C.byref(self.h))
        if rc != VGS_OK:
            raise VgsError(rc, self.lib.vgs_strerror(rc).decode())
        self.verify = verify
        self.n_nodes = 0
        self._batch: Optional[PodBatch] = None
        self._n_picks = 0

    def close(self):
        if getattr(self, "h", None) and self.h.value:
            self.lib.vgs_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _check(self, rc: int):
        if rc != VGS_OK:
            detail = self.lib.vgs_last_error(self.h).decode()
            raise VgsError(rc, f"{self.lib.vgs_strerror(rc).decode()}: {detail}")

    # -- snapshot -------------------------------------------------------------------
    def upload(self, snap: Snapshot) -> int:
        self.set_type_flags([host_type_flags(t) for t in snap.type_names])
        ptrs = (C.c_void_p * 8)(*[snap.planes[f].ctypes.data for f in range(8)])
        ver = C.c_uint64()
        self._check(self.lib.vgs_snapshot_upload(self.h, snap.n_nodes, snap.dmax, ptrs, snap.ndev.ctypes.data,
                                                 C.byref(ver)))
        self.n_nodes, self.dmax, self.type_names = snap.n_nodes, snap.dmax, snap.type_names
        return ver.value

    def upload_ptrs(self, n_nodes: int, dmax: int, plane_ptrs: List[int], ndev_ptr: int, device: bool,
                    asynchronous: bool = False) -> int:
        """planes given as raw addresses (pinned host memory or this GPU's memory); asynchronous: queue the
        upload and return (vgs_snapshot_upload_async) -- the planes must stay put until the next call"""
        ptrs = (C.c_void_p * 8)(*plane_ptrs)
        ver = C.c_uint64()
        fn = self.lib.vgs_snapshot_set_device if device else (
            self.lib.vgs_snapshot_upload_async if asynchronous else self.lib.vgs_snapshot_upload)
        self._check(fn(self.h, n_nodes, dmax, ptrs, C.c_void_p(ndev_ptr), C.byref(ver)))
        self.n_nodes, self.dmax = n_nodes, dmax
        return ver.value

    def upload_narrow_ptrs(self, n_nodes: int, dmax: int, plane_ptrs: List[int], ndev_ptr: int, widths,
                           asynchronous: bool = False) -> int:
        """vgs_snapshot_upload_narrow: planes of 1-, 2- or 4-byte elements (widths[0..7] for the planes, widths[8] for
        ndev), host memory given as raw addresses; None leaves a plane as it is on the GPU (same shape resident)"""
        ptrs = (C.c_void_p * 8)(*plane_ptrs)
        w = np.ascontiguousarray(widths, dtype=np.uint8)
        assert len(w) == 9
        ver = C.c_uint64()
        self._check(self.lib.vgs_snapshot_upload_narrow(self.h, n_nodes, dmax, ptrs, C.c_void_p(ndev_ptr or None), w.ctypes.data,
                                                        1 if asynchronous else 0, C.byref(ver)))
        self.n_nodes, self.dmax = n_nodes, dmax
        return ver.value

    def upload_narrow(self, snap: Snapshot) -> int:
        """the snapshot through the narrow wire format, every plane in the narrowest type its values fit"""
        self.set_type_flags([host_type_flags(t) for t in snap.type_names])
        arrs = narrow_planes(snap)
        ver = self.upload_narrow_ptrs(snap.n_nodes, snap.dmax, [a.ctypes.data for a in arrs[:8]], arrs[8].ctypes.data,
                                      [a.dtype.itemsize for a in arrs])
        self.type_names = snap.type_names
        return ver

    def refresh(self):
        self._check(self.lib.vgs_snapshot_refresh(self.h))

    def set_type_flags(self, flags: List[int]):
        """vgs_snapshot_type_flags: VGS_TYPE_* per type id"""
        arr = np.ascontiguousarray(flags, dtype=np.uint32)
        self._check(self.lib.vgs_snapshot_type_flags(self.h, len(arr), arr.ctypes.data if len(arr) else None))

    def delta(self, deltas: np.ndarray):
        deltas = np.ascontiguousarray(deltas, dtype=DELTA_DTYPE)
        self._check(self.lib.vgs_snapshot_delta(self.h, len(deltas), deltas.ctypes.data))

    def download_snapshot(self) -> Tuple[np.ndarray, np.ndarray]:
        planes = np.zeros((8, self.n_nodes, self.dmax), dtype=np.int32)
        ndev = np.zeros(self.n_nodes, dtype=np.int32)
        ptrs = (C.c_void_p * 8)(*[planes[f].ctypes.data for f in range(8)])
        self._check(self.lib.vgs_snapshot_download(self.h, ptrs, ndev.ctypes.data))
        return planes, ndev

    # -- scoring --------------------------------------------------------------------
    def score_packed(self, pods: np.ndarray, reqs: np.ndarray) -> Tuple[np.ndarray, np.ndarray]:
        """vgs_score_batch on already packed records -> (results, picks)"""
        results = np.zeros(len(pods), dtype=RESULT_DTYPE)
        picks = np.zeros(int(reqs["nums"].sum()) if len(reqs) else 0, dtype=PICK_DTYPE)
        self._check(self.lib.vgs_score_batch(self.h, len(pods), pods.ctypes.data, len(reqs), reqs.ctypes.data,
                                             results.ctypes.data, picks.ctypes.data))
        self._np, self._nr, self._n_picks = len(pods), len(reqs), len(picks)
        return results, picks

    def drain(self, batch: PodBatch, type_names: Optional[List[str]] = None) -> BatchResult:
        """vgs_drain_batch: pods placed one after another, each seeing the previous winners' usage"""
        pods, reqs = pack_batch(batch, type_names if type_names is not None else self.type_names)
        results = np.zeros(len(pods), dtype=RESULT_DTYPE)
        picks = np.zeros(int(reqs["nums"].sum()) if len(reqs) else 0, dtype=PICK_DTYPE)
        self._check(self.lib.vgs_drain_batch(self.h, len(pods), pods.ctypes.data, len(reqs), reqs.ctypes.data,
                                             results.ctypes.data, picks.ctypes.data))
        return self._to_result(results, picks)

    def score(self, batch: PodBatch, type_names: Optional[List[str]] = None) -> BatchResult:
        pods, reqs = pack_batch(batch, type_names if type_names is not None else self.type_names)
        results, picks = self.score_packed(pods, reqs)
        return self._to_result(results, picks, want_verify=self.verify)

    def _to_result(self, results, picks, want_verify=False) -> BatchResult:
        out = BatchResult(results["best_row"].copy(), results["score"].copy(), results["fit_count"].copy(),
                          results["n_picks"].copy(),
                          np.stack([picks["slot"], picks["usedmem"], picks["usedcores"]], axis=1)
                          if len(picks) else np.zeros((0, 3), np.int32))
        if want_verify:
            P, N = len(results), self.n_nodes
            words = (N + 31) // 32
            bm = np.zeros((P, words), dtype=np.uint32)
            sc = np.zeros((P, N), dtype=np.float32)
            self._check(self.lib.vgs_verify_download(self.h, bm.ctypes.data, sc.ctypes.data))
            bits = np.unpackbits(bm.view(np.uint8), axis=1, bitorder="little")[:, :N]
            out.fit, out.scores = bits.astype(bool), sc
        return out

    # split phases
    def pods_upload(self, pods: np.ndarray, reqs: np.ndarray):
        self._np, self._nr = len(pods), len(reqs)
        self._n_picks = int(reqs["nums"].sum()) if len(reqs) else 0
        self._check(self.lib.vgs_pods_upload(self.h, len(pods), pods.ctypes.data, len(reqs), reqs.ctypes.data))

    def score_resident(self):
        self._check(self.lib.vgs_score_resident(self.h))

    def resolve_resident(self):
        self._check(self.lib.vgs_resolve_resident(self.h))

    def results_download(self) -> Tuple[np.ndarray, np.ndarray]:
        results = np.zeros(self._np, dtype=RESULT_DTYPE)
        picks = np.zeros(self._n_picks, dtype=PICK_DTYPE)
        self._check(self.lib.vgs_results_download(self.h, results.ctypes.data, picks.ctypes.data))
        return results, picks

    def exchange_block(self) -> Tuple[int, int]:
        """(device address, bytes) of the [keys | counts | picks] block a shard ships in one collective"""
        ptr, n = C.c_void_p(), C.c_uint64()
        self._check(self.lib.vgs_exchange_block(self.h, C.byref(ptr), C.byref(n)))
        return int(ptr.value), int(n.value)

    def resolve_local(self):
        self._check(self.lib.vgs_resolve_local(self.h))

    def exchange_merge(self, gathered_ptr: int, n_shards: int, stride_bytes: int):
        self._check(self.lib.vgs_exchange_merge(self.h, C.c_void_p(gathered_ptr), n_shards, stride_bytes))

    def sync(self):
        self._check(self.lib.vgs_sync(self.h))

    def keys_device(self) -> Tuple[int, int]:
        k, c = C.c_void_p(), C.c_void_p()
        self._check(self.lib.vgs_keys_device(self.h, C.byref(k), C.byref(c)))
        return k.value, c.value

    def results_device(self) -> Tuple[int, int]:
        r, p = C.c_void_p(), C.c_void_p()
        self._check(self.lib.vgs_results_device(self.h, C.byref(r), C.byref(p)))
        return r.value, p.value

    def keys_merge(self, keys_ptr: int, counts_ptr: int, n_shards: int, stride: int, count_stride: int = 0):
        self._check(self.lib.vgs_keys_merge(self.h, C.c_void_p(keys_ptr), C.c_void_p(counts_ptr), n_shards, stride,
                                            count_stride))

    def plan(self) -> VgsPlan:
        """vgs_last_plan: which kernels the last score pass used, and their launch geometry"""
        pl = VgsPlan()
        self._check(self.lib.vgs_last_plan(self.h, C.byref(pl)))
        return pl

    def timing(self) -> VgsTiming:
        t = VgsTiming()
        self._check(self.lib.vgs_last_timing(self.h, C.byref(t)))
        return t


class Group:
    """vgs_group_*: the node set sharded over several GPUs of one box inside the library (one process, one host
    thread), exchange by NCCL all-gather or by peer-mapped stores.  Same calls as a Context."""

    def __init__(self, devices: List[int], max_nodes: int, max_devices: int = 8, max_pods: int = 4096,
                 max_reqs: Optional[int] = None, max_picks: Optional[int] = None, exchange: int = EXCHANGE_NCCL,
                 tuning: Optional[Dict[str, int]] = None):
        self.lib = load_library()
        max_reqs = max_reqs or max_pods * 4
        max_picks = max_picks or max_reqs * 4
        tune = VgsTuning()
        self.lib.vgs_tuning_defaults(C.byref(tune))
        for k, val in (tuning or {}).items():
            setattr(tune, k, int(val))
        self._tune = tune
        cfg = VgsGroupConfig(len(devices), (C.c_int32 * 8)(*devices), max(max_nodes, 1), max_devices, max_pods, max_reqs,
                             max_picks, exchange, C.pointer(tune))
        self.h = C.c_void_p()
        rc = self.lib.vgs_group_init(C.byref(cfg), C.byref(self.h))
        if rc != VGS_OK:
            raise VgsError(rc, self.lib.vgs_strerror(rc).decode())

    def close(self):
        if getattr(self, "h", None) and self.h.value:
            self.lib.vgs_group_destroy(self.h)
            self.h = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _check(self, rc: int):
        if rc != VGS_OK:
            raise VgsError(rc, f"{self.lib.vgs_strerror(rc).decode()}: {self.lib.vgs_group_last_error(self.h).decode()}")

    def upload(self, snap: Snapshot):
        flags = np.ascontiguousarray([host_type_flags(t) for t in snap.type_names], dtype=np.uint32)
        self._check(self.lib.vgs_group_snapshot_type_flags(self.h, len(flags), flags.ctypes.data if len(flags) else None))
        ptrs = (C.c_void_p * 8)(*[snap.planes[f].ctypes.data for f in range(8)])
        self._check(self.lib.vgs_group_snapshot_upload(self.h, snap.n_nodes, snap.dmax, ptrs, snap.ndev.ctypes.data))
        self.type_names = snap.type_names

    def delta(self, deltas: np.ndarray):
        deltas = np.ascontiguousarray(deltas, dtype=DELTA_DTYPE)
        self._check(self.lib.vgs_group_snapshot_delta(self.h, len(deltas), deltas.ctypes.data))

    def score_packed(self, pods: np.ndarray, reqs: np.ndarray):
        results = np.zeros(len(pods), dtype=RESULT_DTYPE)
        picks = np.zeros(int(reqs["nums"].sum()) if len(reqs) else 0, dtype=PICK_DTYPE)
        self._check(self.lib.vgs_group_score_batch(self.h, len(pods), pods.ctypes.data, len(reqs), reqs.ctypes.data,
                                                   results.ctypes.data, picks.ctypes.data))
        return results, picks

    def score(self, batch: PodBatch) -> BatchResult:
        pods, reqs = pack_batch(batch, self.type_names)
        r, pk = self.score_packed(pods, reqs)
        return BatchResult(r["best_row"].copy(), r["score"].copy(), r["fit_count"].copy(), r["n_picks"].copy(),
                           np.stack([pk["slot"], pk["usedmem"], pk["usedcores"]], axis=1) if len(pk)
                           else np.zeros((0, 3), np.int32))

    def timing(self):
        t, x = C.c_double(), C.c_double()
        self._check(self.lib.vgs_group_last_timing(self.h, C.byref(t), C.byref(x)))
        return t.value, x.value
