"""Synthetic snapshots and pod batches of BASELINE.json's configs.

Deterministic everywhere: all randomness comes from a splitmix64 counter stream
implemented here in integer numpy (no dependence on numpy's bit generators).
Request shapes follow the reference's examples (examples/nvidia/*.yaml): vGPU
count, gpumem MB or gpumem-percentage, gpucores, use/nouse-gputype, numa-bind.
"""
from __future__ import annotations

from typing import Dict, List, Tuple

import numpy as np

from .model import ANNO_NOUSE, ANNO_NUMA, ANNO_USE, MEMPCT_UNSET, PodBatch, Snapshot

SEED = 0xB200

# (DeviceUsage.Type as registered by the device plugin "NVIDIA-"+model, Totalmem MiB)
GPU_TYPES: List[Tuple[str, int]] = [
    ("NVIDIA-Tesla V100-PCIE-32GB", 32510),
    ("NVIDIA-Tesla V100-SXM2-16GB", 16160),
    ("NVIDIA-NVIDIA A100-SXM4-40GB", 40960),
    ("NVIDIA-NVIDIA A100-SXM4-80GB", 81920),
    ("NVIDIA-NVIDIA H100 80GB HBM3", 81559),
    ("NVIDIA-Tesla T4", 15360),
    ("NVIDIA-NVIDIA GeForce RTX 2080 Ti", 11264),
    ("NVIDIA-NVIDIA B200", 183359),
]
TYPE_TOKENS = ["V100", "a100", "H100", "t4", "2080", "b200", "Tesla", "sxm", "A100-SXM4-80GB", "GeForce"]


class SplitMix:
    """Vectorised splitmix64 stream."""

    def __init__(self, seed: int):
        self.state = np.uint64(seed & 0xFFFFFFFFFFFFFFFF)

    def u64(self, n: int) -> np.ndarray:
        with np.errstate(over="ignore"):
            g = np.uint64(0x9E3779B97F4A7C15)
            idx = np.arange(1, n + 1, dtype=np.uint64)
            z = self.state + g * idx
            self.state = self.state + g * np.uint64(n)
            z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
            z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
            return z ^ (z >> np.uint64(31))

    def below(self, n: int, bound) -> np.ndarray:
        """uniform ints in [0, bound) (bound scalar or array); tiny modulo bias is irrelevant"""
        return (self.u64(n) >> np.uint64(11)) % np.asarray(bound, dtype=np.uint64)

    def choice(self, n: int, values) -> np.ndarray:
        v = np.asarray(values)
        return v[self.below(n, len(v)).astype(np.int64)]

    def unit(self, n: int) -> np.ndarray:
        return (self.u64(n) >> np.uint64(11)).astype(np.float64) / float(1 << 53)


def make_snapshot(n_nodes: int, seed: int = SEED, dmax: int = 8, count: int = 10) -> Snapshot:
    rng = SplitMix(seed)
    D = dmax
    planes = np.zeros((8, n_nodes, D), dtype=np.int32)
    ntype = rng.below(n_nodes, len(GPU_TYPES)).astype(np.int32)
    totalmem = np.array([m for _, m in GPU_TYPES], dtype=np.int64)[ntype]
    kind = rng.unit(n_nodes)  # <0.10 idle, >=0.95 saturated
    CH = 65536
    cores_choice = np.array([0, 0, 10, 25, 30, 50], dtype=np.int64)
    for lo in range(0, n_nodes, CH):
        hi = min(lo + CH, n_nodes)
        n = hi - lo
        used = rng.below(n * D, count + 1).astype(np.int64).reshape(n, D)
        tm = np.repeat(totalmem[lo:hi, None], D, axis=1)
        draws_m = 500 + rng.below(n * D * count, np.repeat(np.maximum(tm // 4 - 500 + 1, 1).reshape(-1), count))
        draws_m = draws_m.astype(np.int64).reshape(n, D, count)
        draws_c = cores_choice[rng.below(n * D * count, len(cores_choice)).astype(np.int64)].reshape(n, D, count)
        live = np.arange(count)[None, None, :] < used[:, :, None]
        usedmem = np.minimum(tm, (draws_m * live).sum(axis=2))
        usedcores = np.minimum(100, (draws_c * live).sum(axis=2))
        k = kind[lo:hi]
        idle, sat = k < 0.10, k >= 0.95
        used[idle], usedmem[idle], usedcores[idle] = 0, 0, 0
        used[sat], usedcores[sat] = count, 100
        usedmem[sat] = tm[sat]
        planes[0, lo:hi] = used
        planes[1, lo:hi] = count
        planes[2, lo:hi] = usedmem
        planes[3, lo:hi] = tm
        planes[4, lo:hi] = usedcores
        planes[5, lo:hi] = 100
        planes[6, lo:hi] = (np.arange(D) >= (D + 1) // 2).astype(np.int32)[None, :]
        planes[7, lo:hi] = ntype[lo:hi, None]
    ndev = np.full(n_nodes, D, dtype=np.int32)
    return Snapshot(planes, ndev, [t for t, _ in GPU_TYPES])


def _type_annos(rng: SplitMix, n: int) -> List[Dict[str, str]]:
    sel = rng.unit(n)
    ntok = 1 + rng.below(n, 2).astype(int)
    t0 = rng.below(n, len(TYPE_TOKENS)).astype(int)
    t1 = rng.below(n, len(TYPE_TOKENS)).astype(int)
    out = []
    for i in range(n):
        toks = TYPE_TOKENS[t0[i]] if ntok[i] == 1 else TYPE_TOKENS[t0[i]] + "," + TYPE_TOKENS[t1[i]]
        if sel[i] < 0.25:
            out.append({ANNO_USE: toks})
        elif sel[i] < 0.50:
            out.append({ANNO_NOUSE: toks})
        else:
            out.append({})
    return out


def pods_single(n_pods: int, seed: int) -> PodBatch:
    """config 2: one container, Nums=1, Memreq U{1000..16000 step 500}, Cores in {0,10,20,30,50}"""
    rng = SplitMix(seed)
    mem = 1000 + 500 * rng.below(n_pods, 31).astype(np.int32)
    cores = rng.choice(n_pods, [0, 10, 20, 30, 50]).astype(np.int32)
    ar = np.arange(n_pods + 1, dtype=np.int32)
    return PodBatch(ar, ar, np.ones(n_pods, np.int32), mem, np.full(n_pods, MEMPCT_UNSET, np.int32), cores)


def pods_mixed(n_pods: int, seed: int) -> PodBatch:
    """config 3/5: 50% Memreq / 50% MemPct in {10,25,50,100}; Nums in {1,1,1,2};
    25% use-gputype, 25% nouse-gputype, 5% exclusive (0, 100, 100)"""
    rng = SplitMix(seed)
    nums = rng.choice(n_pods, [1, 1, 1, 2]).astype(np.int32)
    by_pct = rng.unit(n_pods) < 0.5
    mem = (1000 + 500 * rng.below(n_pods, 31)).astype(np.int32)
    pct = rng.choice(n_pods, [10, 25, 50, 100]).astype(np.int32)
    cores = rng.choice(n_pods, [0, 10, 20, 30, 50]).astype(np.int32)
    excl = rng.unit(n_pods) < 0.05
    memreq = np.where(by_pct, 0, mem).astype(np.int32)
    mempct = np.where(by_pct, pct, MEMPCT_UNSET).astype(np.int32)
    memreq[excl], mempct[excl], cores[excl] = 0, 100, 100
    annos = _type_annos(rng, n_pods)
    ar = np.arange(n_pods + 1, dtype=np.int32)
    return PodBatch(ar, ar, nums, memreq, mempct, cores, ["NVIDIA"] * n_pods, annos)


def pods_multi(n_pods: int, seed: int) -> PodBatch:
    """config 4: 2-4 containers totalling 2-8 vGPUs (30% one container with Nums 2..8),
    10% numa-bind, memory by MB or percentage"""
    rng = SplitMix(seed)
    single = rng.unit(n_pods) < 0.30
    nctr = 2 + rng.below(n_pods, 3).astype(int)
    numa = rng.unit(n_pods) < 0.10
    big = 2 + rng.below(n_pods, 7).astype(int)
    pods = []
    draws = rng.below(n_pods * 4 * 4, 1 << 30).astype(np.int64).reshape(n_pods, 4, 4)
    for p in range(n_pods):
        ctrs = []
        if single[p]:
            shape = [int(big[p])]
        else:
            shape = [1] * int(nctr[p])
            extra = int(draws[p, 0, 0] % 5)  # spread up to 4 more devices, total <= 8
            for e in range(extra):
                shape[int(draws[p, 1, e % 4] % len(shape))] += 1
        for c, n in enumerate(shape):
            d = draws[p, c]
            if d[1] % 2:
                mem, pct = 0, [10, 25, 50][int(d[2] % 3)]
            else:
                mem, pct = 1000 + 500 * int(d[2] % 21), MEMPCT_UNSET
            cores = [0, 0, 10, 20, 30][int(d[3] % 5)]
            ctrs.append([(n, mem, pct, cores)])
        pods.append({"containers": ctrs, "annos": ({ANNO_NUMA: "true"} if numa[p] else {})})
    return PodBatch.from_pods(pods)


def make_stress_snapshot(n_nodes: int, seed: int = SEED + 77, dmax: int = 8) -> Snapshot:
    """A snapshot built against the assumptions the synthetic configs are kind to (robustness number, not a
    BASELINE config): Count differs per device (1..32), nodes are ragged (0..8 devices), 48 device types mixed
    INSIDE nodes, two numa nodes in random order, and the fill level falls with the row index -- the kernels walk
    rows downwards, so candidate scores keep RISING along the walk and there are hundreds of distinct score values
    (the sliced kernels track the 32 largest exactly)."""
    rng = SplitMix(seed)
    D = dmax
    names, mems = [], []
    for t, m in GPU_TYPES:
        for k in range(6):
            names.append(f"{t}-rev{k}")
            mems.append(m - 256 * k)
    mems = np.array(mems, dtype=np.int64)
    planes = np.zeros((8, n_nodes, D), dtype=np.int32)
    count = 1 + rng.below(n_nodes * D, 32).astype(np.int64).reshape(n_nodes, D)
    # free fraction grows with the row: row 0 is nearly full (high scores), the last row nearly idle
    frac = (np.arange(n_nodes, dtype=np.float64)[:, None] + 1.0) / n_nodes
    noise = rng.unit(n_nodes * D).reshape(n_nodes, D) * 0.15
    free = np.clip(np.round(count * np.clip(frac + noise - 0.05, 0.0, 1.0)), 0, count).astype(np.int64)
    used = count - free
    tid = rng.below(n_nodes * D, len(names)).astype(np.int64).reshape(n_nodes, D)
    tm = mems[tid]
    usedmem = np.minimum(tm, (used * tm) // (count * 2) + rng.below(n_nodes * D, 2000).astype(np.int64).reshape(n_nodes, D) * (used > 0))
    usedcores = np.minimum(100, (used * 60) // count + 5 * rng.below(n_nodes * D, 5).astype(np.int64).reshape(n_nodes, D) * (used > 0))
    planes[0], planes[1], planes[2], planes[3] = used, count, usedmem, tm
    planes[4], planes[5] = usedcores, 100
    planes[6] = rng.below(n_nodes * D, 2).astype(np.int32).reshape(n_nodes, D)
    planes[7] = tid
    ndev = rng.below(n_nodes, D + 1).astype(np.int32)
    ndev[rng.unit(n_nodes) < 0.6] = D  # most nodes are full-width, the rest ragged (incl. 0 devices)
    return Snapshot(planes, ndev, names)


CONFIGS = {
    # name: (n_nodes, n_pods, pod generator)
    "cfg1": (64, 1, None),
    "cfg2": (10_000, 256, pods_single),
    "cfg3": (100_000, 4096, pods_mixed),
    "cfg4": (100_000, 4096, pods_multi),
    "cfg5": (1_000_000, 16384, pods_mixed),
}


def make_config(name: str, n_nodes: int = None, n_pods: int = None):
    """(Snapshot, PodBatch) of a BASELINE.json config, optionally resized."""
    if name == "stress":
        return make_stress_snapshot(n_nodes or 100_000), pods_mixed(n_pods or 4096, SEED + 177)
    N, P, gen = CONFIGS[name]
    N = n_nodes or N
    P = n_pods or P
    idx = int(name[3:])
    snap = make_snapshot(N, SEED + idx)
    if gen is None:  # config 1: examples/nvidia/default_use.yaml, 1 vGPU 3000M 30%
        batch = PodBatch.from_pods([{"containers": [[(1, 3000, MEMPCT_UNSET, 30)]]}] * P)
    else:
        batch = gen(P, SEED + 100 + idx)
    return snap, batch
