"""Packed host-side data model of the Filter scoring pass.

The Go extender keeps its own ``NodeUsage`` / ``DeviceUsage`` structs
(pkg/scheduler/nodes.go:44-48, pkg/util/types.go:105-117) and packs them into the
int32 planes below right before the cgo call; this module is the same packed form
for hosts without Go (tests, bench).  Nothing here computes a fit or a score.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence

import numpy as np

# plane order == include/vgpuscore.h VGS_PLANE_*
PLANES = ("used", "count", "usedmem", "totalmem", "usedcores", "totalcore", "numa", "type_id")
P_USED, P_COUNT, P_USEDMEM, P_TOTALMEM, P_USEDCORES, P_TOTALCORE, P_NUMA, P_TYPEID = range(8)

MEMPCT_UNSET = 101  # pkg/device/nvidia/device.go:136
ANNO_USE = "nvidia.com/use-gputype"  # nvidia/device.go:20
ANNO_NOUSE = "nvidia.com/nouse-gputype"  # :21
ANNO_NUMA = "nvidia.com/numa-bind"  # :22
ANNO_DCU_USE = "hygon.com/use-dcutype"  # pkg/device/hygon/device.go:20
ANNO_DCU_NOUSE = "hygon.com/nouse-dcutype"  # :21
ANNO_MLU_USE = "cambricon.com/use-mlutype"  # pkg/device/cambricon/device.go:19
ANNO_MLU_NOUSE = "cambricon.com/nouse-mlutype"  # :20


@dataclass
class Snapshot:
    """Usage snapshot: what getNodesUsage (scheduler.go:249-310) hands to calcScore."""

    planes: np.ndarray  # int32 [8, N, D], C-contiguous; planes[f, row, slot]
    ndev: np.ndarray  # int32 [N] = len(node.Devices)
    type_names: List[str]  # distinct DeviceUsage.Type strings; planes[P_TYPEID] indexes it

    def __post_init__(self):
        self.planes = np.ascontiguousarray(self.planes, dtype=np.int32)
        self.ndev = np.ascontiguousarray(self.ndev, dtype=np.int32)
        assert self.planes.ndim == 3 and self.planes.shape[0] == 8
        assert self.ndev.shape == (self.planes.shape[1],)

    @property
    def n_nodes(self) -> int:
        return int(self.planes.shape[1])

    @property
    def dmax(self) -> int:
        return int(self.planes.shape[2])

    def rows(self, lo: int, hi: int) -> "Snapshot":
        """Contiguous node-range shard [lo, hi)."""
        return Snapshot(self.planes[:, lo:hi, :].copy(), self.ndev[lo:hi].copy(), self.type_names)

    @staticmethod
    def from_nodes(nodes: Sequence[Sequence[dict]], dmax: Optional[int] = None) -> "Snapshot":
        """nodes[row] = list of device dicts with the DeviceUsage field names
        (Used, Count, Usedmem, Totalmem, Usedcores, Totalcore, Numa, Type)."""
        n = len(nodes)
        d = dmax or max([len(x) for x in nodes] + [1])
        planes = np.zeros((8, n, d), dtype=np.int32)
        ndev = np.zeros(n, dtype=np.int32)
        names: List[str] = []
        idx: Dict[str, int] = {}
        for r, devs in enumerate(nodes):
            ndev[r] = len(devs)
            for s, dv in enumerate(devs):
                t = dv.get("Type", "NVIDIA-Tesla V100-PCIE-32GB")
                if t not in idx:
                    idx[t] = len(names)
                    names.append(t)
                planes[P_USED, r, s] = dv.get("Used", 0)
                planes[P_COUNT, r, s] = dv.get("Count", 10)
                planes[P_USEDMEM, r, s] = dv.get("Usedmem", 0)
                planes[P_TOTALMEM, r, s] = dv.get("Totalmem", 32768)
                planes[P_USEDCORES, r, s] = dv.get("Usedcores", 0)
                planes[P_TOTALCORE, r, s] = dv.get("Totalcore", 100)
                planes[P_NUMA, r, s] = dv.get("Numa", 0)
                planes[P_TYPEID, r, s] = idx[t]
        return Snapshot(planes, ndev, names)


@dataclass
class PodBatch:
    """P pending pods = P times the [][]ContainerDeviceRequest of
    k8sutil.Resourcereqs (pkg/k8sutil/pod.go:26-40) plus each pod's annotations."""

    ctr_begin: np.ndarray  # int32 [P+1]  containers of pod p (empty containers included)
    req_begin: np.ndarray  # int32 [C+1]  requests of container c
    nums: np.ndarray  # int32 [R]
    memreq: np.ndarray
    mempct: np.ndarray
    coresreq: np.ndarray
    req_type: List[str] = field(default_factory=list)  # [R] ContainerDeviceRequest.Type
    annos: List[Dict[str, str]] = field(default_factory=list)  # [P] pod annotations

    def __post_init__(self):
        for k in ("ctr_begin", "req_begin", "nums", "memreq", "mempct", "coresreq"):
            setattr(self, k, np.ascontiguousarray(getattr(self, k), dtype=np.int32))
        if not self.req_type:
            self.req_type = ["NVIDIA"] * len(self.nums)
        if not self.annos:
            self.annos = [{} for _ in range(self.n_pods)]

    @property
    def n_pods(self) -> int:
        return len(self.ctr_begin) - 1

    @property
    def n_reqs(self) -> int:
        return len(self.nums)

    def pod_req_range(self, p: int):
        c0, c1 = int(self.ctr_begin[p]), int(self.ctr_begin[p + 1])
        return int(self.req_begin[c0]), int(self.req_begin[c1])

    @property
    def pick_begin(self) -> np.ndarray:
        """exclusive prefix sum of nums: request r owns picks[pb[r]:pb[r]+nums[r]]"""
        pb = np.zeros(self.n_reqs + 1, dtype=np.int64)
        np.cumsum(np.maximum(self.nums, 0), out=pb[1:])
        return pb

    @staticmethod
    def from_pods(pods: Sequence[dict]) -> "PodBatch":
        """pods[p] = {"containers": [[(nums, memreq, mempct, coresreq[, type]), ...], ...],
        "annos": {...}}; a container is the list of its per-vendor requests."""
        ctr_begin, req_begin = [0], [0]
        nums, memreq, mempct, cores, rtype, annos = [], [], [], [], [], []
        for pod in pods:
            for ctr in pod["containers"]:
                for rq in ctr:
                    nums.append(rq[0])
                    memreq.append(rq[1])
                    mempct.append(rq[2])
                    cores.append(rq[3])
                    rtype.append(rq[4] if len(rq) > 4 else "NVIDIA")
                req_begin.append(len(nums))
            ctr_begin.append(len(req_begin) - 1)
            annos.append(dict(pod.get("annos", {})))
        return PodBatch(np.array(ctr_begin), np.array(req_begin), np.array(nums, dtype=np.int32),
                        np.array(memreq, dtype=np.int32), np.array(mempct, dtype=np.int32),
                        np.array(cores, dtype=np.int32), rtype, annos)

    def select(self, idx: Sequence[int]) -> "PodBatch":
        """sub-batch of the given pods (order kept)"""
        pods = []
        for p in idx:
            c0, c1 = int(self.ctr_begin[p]), int(self.ctr_begin[p + 1])
            ctrs = []
            for c in range(c0, c1):
                r0, r1 = int(self.req_begin[c]), int(self.req_begin[c + 1])
                ctrs.append([(int(self.nums[r]), int(self.memreq[r]), int(self.mempct[r]),
                              int(self.coresreq[r]), self.req_type[r]) for r in range(r0, r1)])
            pods.append({"containers": ctrs, "annos": self.annos[p]})
        return PodBatch.from_pods(pods)


@dataclass
class BatchResult:
    """Per pod: winner NodeScore (score.go:29-33) under the canonical tie rule."""

    best_row: np.ndarray  # int32 [P], -1 = no node fits
    score: np.ndarray  # float32 [P]
    fit_count: np.ndarray  # int32 [P]
    n_picks: np.ndarray  # int32 [P]
    picks: np.ndarray  # int32 [sum nums, 3] (slot, usedmem, usedcores), request-major
    fit: Optional[np.ndarray] = None  # verify: bool [P, N]
    scores: Optional[np.ndarray] = None  # verify: float32 [P, N]

    def pod_devices(self, batch: PodBatch, p: int):
        """util.PodDevices of pod p's winner: one list of (slot, mem, cores) per container."""
        if self.best_row[p] < 0:
            return None
        pb = batch.pick_begin
        out = []
        for c in range(int(batch.ctr_begin[p]), int(batch.ctr_begin[p + 1])):
            devs = []
            for r in range(int(batch.req_begin[c]), int(batch.req_begin[c + 1])):
                for j in range(int(batch.nums[r])):
                    s, m, k = self.picks[int(pb[r]) + j]
                    devs.append((int(s), int(m), int(k)))
            out.append(devs)
        return out
