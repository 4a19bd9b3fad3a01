#!/usr/bin/env python
"""bench.py -- pod x node fit-evals/s of the Filter scoring pass (BASELINE.json metric).

  python bench.py --gpus N --steps K --warmup W          (under torchrun for N > 1)
  python bench.py --impl reference ...                   CPU arm: the oracle restatement

Step = one pass of the hot path over one pod batch against the resident snapshot:
K0 prepare (raw planes -> walk-ordered tiles) + K1 score + K3 resolve; for N > 1 every shard
resolves its own best node, the [keys | counts | picks] blocks are all-gathered once (NCCL) and
merged (K2x).  Headline workload = BASELINE config 3 (100k nodes x 8 devices x 4096 mixed pods);
for N > 1 every rank holds its own 100k-node shard (weak scaling, N x 100k nodes) and the same pods.

`value`   : device-timed (CUDA events per step on the launching stream, max over ranks),
            inputs resident in HBM, L2 flushed between steps.
`e2e`     : the same metric through the host-buffer C ABI calls the Go shim makes
            (snapshot upload, queued + pods upload / score / resolve / download): pinned
            host planes and request records in, results out, every step.  The planes travel in the narrow
            wire format (vgs_snapshot_upload_narrow: int8 where the values fit, widened on the GPU);
            `e2e_int32_planes` is the same step with int32 planes, the figure of the earlier rounds.
`e2e_delta`: the production shape of the same calls: snapshot resident, each step ships the usage deltas of
            the previous step's placements (vgs_snapshot_delta) + the pod batch.
`roofline`: the score kernels (K1) against the measured HBM peak with the algorithmic bytes of SURVEY 8(d),
            and `issue_fraction` = K1 warp instructions (ncu launch list under profiles/) / K1 time of this
            run / issue slots.  The batch regime is integer-issue bound; see DESIGN.md.
`configs` : the other BASELINE configs as device-timed legs of their own, each with a sampled oracle check:
            N = 1: cfg2, cfg4, one cfg5 shard (125k nodes x 16384 pods); N > 1: cfg5 (1M nodes) and cfg4,
            both strong-scaled over the N shards.
`check`   : a sample of the headline's pods re-scored by the CPU oracle over all nodes (not timed).
`roofline_hbm_regime` (N = 1): one pod against 1M resident nodes, where the pass IS HBM-bound.
`filter_latency` (N = 1): wall clock of the single-pod vgs_score_batch call, snapshot resident.
`cpu_baseline` (N = 1): oracle R2 (packed, all host cores) + R1 / R1-io (the reference's own shape, one thread).
The reference (Go) cannot be built or run here; its CPU arm is the oracle's packed
multi-threaded restatement ("port"), timed on this box's host cores.
"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "pod_x_node_fit_evals_per_sec"
UNIT = "fit-evals/s"
WORKLOAD = "cfg3: 100k nodes x 8 devices x 4096 mixed pods (mem MB / mem %, use/nouse-gputype, exclusive)"
NODES_PER_GPU = 100_000
N_PODS = 4096
L2_FLUSH_BYTES = 256 << 20


def measured_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def algorithmic_bytes(n_nodes, dmax, n_pods, n_reqs):
    # snapshot once + request records + results
    r = n_reqs / max(n_pods, 1)
    return n_nodes * (dmax * 32 + 4) + n_pods * (16 + r * 32) + n_pods * 80


class ClockSampler:
    """NVML sampling of SM clock / throttle reasons while the timed regions run."""

    def __init__(self, index):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._t = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _run(self):
        nv = self.nv
        names = {
            getattr(nv, "nvmlClocksThrottleReasonHwSlowdown", 0x8): "hw_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonSwPowerCap", 0x4): "sw_power_cap",
        }
        while not self._stop.is_set():
            try:
                util = nv.nvmlDeviceGetUtilizationRates(self.h).gpu
                mhz = nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                self.samples.append((mhz, util))
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop.wait(0.02)

    def start(self):
        if self.nv:
            self._t = threading.Thread(target=self._run, daemon=True)
            self._t.start()

    def stop(self):
        self._stop.set()
        if self._t:
            self._t.join()
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": [], "samples": 0}
        busy = [m for m, u in self.samples if u > 0] or [m for m, _ in self.samples]
        busy.sort()
        return {"sm_mhz": busy[len(busy) // 2], "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(self.samples)}


def cpu_arm(snap, batch, n_pods_sample, threads=None):
    """oracle packed restatement on a bounded sample: first n_pods_sample pods x all nodes"""
    import oracle
    sub = batch.select(list(range(min(n_pods_sample, batch.n_pods))))
    t0 = time.perf_counter()
    r = oracle.score_packed(snap, sub, threads=threads)
    dt = time.perf_counter() - t0
    return sub.n_pods * snap.n_nodes / r.seconds, r.threads, dt, sub.n_pods


def run_reference(args):
    """--impl reference: the reference's own CPU algorithm (restated: no Go toolchain on this
    image) on the host cores, rank 0 only."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import vgpu_b200 as v
    snap, batch = v.workload.make_config("cfg3", NODES_PER_GPU, N_PODS)
    threads = os.cpu_count() or 1
    sample_pods = 128
    for _ in range(max(args.warmup, 1)):
        cpu_arm(snap, batch, 16, threads)
    secs, evals = 0.0, 0
    for _ in range(args.steps):
        val, th, dt, sp = cpu_arm(snap, batch, sample_pods, threads)
        evals += sp * snap.n_nodes
        secs += sp * snap.n_nodes / val
    value = evals / secs
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * secs / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "sample": f"{sample_pods} pods x {snap.n_nodes} nodes per step"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port",
                         "sample": f"{sample_pods} of {N_PODS} pods x all {snap.n_nodes} nodes per step, "
                                   "oracle packed restatement (reference Go not runnable here)"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def run_sweep(args):
    """--sweep: the CPU/GPU crossover: N x P grid, single-container mixed pods
    (config 3 mix).  GPU times are wall clock around the C ABI calls the Go shim would make
    (median of reps): `resident` = vgs_score_batch with the snapshot already on the GPU,
    `with_upload` = vgs_snapshot_upload + vgs_score_batch.  CPU arms on bounded samples:
    R1 = faithful single-thread restatement, R1io = R1 + the reference's log lines to /dev/null,
    R2 = packed multi-thread restatement.  One JSON line per cell."""
    import statistics

    import numpy as np

    import oracle
    import vgpu_b200 as v

    Ns = [int(x) for x in args.sweep_nodes.split(",")]
    Ps = [int(x) for x in args.sweep_pods.split(",")]
    threads = os.cpu_count() or 1
    out = open(args.sweep_out, "w") if args.sweep_out else None
    for N in Ns:
        snap = v.workload.make_snapshot(N, v.workload.SEED + 3)
        ctx = v.Context(max_nodes=N, max_pods=max(Ps), max_reqs=max(Ps), max_picks=4 * max(Ps))
        t0 = time.perf_counter()
        ctx.upload(snap)
        for P in Ps:
            _, batch = v.workload.make_config("cfg3", 8, P)
            pods, reqs = v.capi.pack_batch(batch, snap.type_names)
            reps = 30 if N * P <= 1e9 else 8
            res_t, up_t, k1_t = [], [], []
            for r in range(reps + 3):
                t0 = time.perf_counter()
                ctx.score_packed(pods, reqs)
                t1 = time.perf_counter()
                if r >= 3:
                    res_t.append(t1 - t0)
                    k1_t.append(ctx.timing().score_ms * 1e-3)
            for r in range(max(3, reps // 4) + 1):
                t0 = time.perf_counter()
                ctx.upload(snap)
                ctx.score_packed(pods, reqs)
                t1 = time.perf_counter()
                if r >= 1:
                    up_t.append(t1 - t0)
            # CPU arms, bounded: at most ~2e7 pairs for R1, ~2e8 for R2
            def sample(limit):
                sp = max(1, min(P, int(limit // max(N, 1))))
                return batch.select(list(range(sp))), sp
            b1, sp1 = sample(2e7)
            r1 = oracle.score_faithful(snap, b1)
            bio, spio = sample(2e6)
            rio = oracle.score_faithful(snap, bio, log_io=True)
            b2, sp2 = sample(3e8)
            r2 = oracle.score_packed(snap, b2, threads=threads)
            scale = lambda secs, sp: secs * P / sp
            cell = {
                "nodes": N, "pods": P, "pairs": N * P,
                "gpu_resident_s": statistics.median(res_t), "gpu_with_upload_s": statistics.median(up_t),
                "gpu_k1_s": statistics.median(k1_t),
                "cpu_r1_s": scale(r1.seconds_scoring, sp1), "cpu_r1_rebuild_s": scale(r1.seconds_rebuild, sp1),
                "cpu_r1io_s": scale(rio.seconds_scoring, spio), "cpu_r2_s": scale(r2.seconds, sp2),
                "cpu_threads": threads, "cpu_sample_pods": {"r1": sp1, "r1io": spio, "r2": sp2},
            }
            line = json.dumps(cell)
            print(line, flush=True)
            if out:
                out.write(line + "\n")
        ctx.close()
    if out:
        out.close()


def latest_k1_profile():
    """profiles/rNN_k1_profile.json of the newest round (tools/make_k1_profile.py writes it from the ncu launch
    list of this command): warp instructions and DRAM bytes of the K1 phase of one cfg3 step, with the tree it
    was captured on."""
    import glob
    files = sorted(glob.glob(os.path.join(ROOT, "profiles", "r*_k1_profile.json")))
    if not files:
        return None
    try:
        with open(files[-1]) as f:
            d = json.load(f)
        d["file"] = os.path.relpath(files[-1], ROOT)
        return d
    except Exception:
        return None


def pin_to_gpu_numa(torch, local):
    """Run this rank on the cores next to its GPU, so that the pinned planes it allocates (first touch) and the packing
    of its batches sit on that GPU's NUMA node: with N ranks uploading 26 MB each per e2e step, planes on a remote node
    go over the inter-socket link.  Best effort: returns the node, or None when the topology cannot be read."""
    try:
        pr = torch.cuda.get_device_properties(local)
        bus = "%04x:%02x:%02x.0" % (pr.pci_domain_id, pr.pci_bus_id, pr.pci_device_id)
        base = "/sys/bus/pci/devices/" + bus
        cpus = set()
        for part in open(base + "/local_cpulist").read().strip().split(","):
            a, _, b = part.partition("-")
            cpus.update(range(int(a), int(b or a) + 1))
        cpus &= os.sched_getaffinity(0)
        if not cpus:
            return None
        os.sched_setaffinity(0, cpus)
        node = int(open(base + "/numa_node").read().strip())
        return node if node >= 0 else None  # (-1: a single-node host, e.g. the VMs this was measured on -- nothing to gain)
    except Exception:
        return None


def sampled_check(v, np, snap_whole, batch, results, picks, n_sample=48):
    """not timed: a sample of pods re-scored by the CPU oracle over ALL nodes; bit-exact or not"""
    import oracle
    P = batch.n_pods
    idx = list(range(0, P, max(1, P // n_sample)))
    want = oracle.score_packed(snap_whole, batch.select(idx))
    pb = batch.pick_begin
    ok = (np.array_equal(results["best_row"][idx], want.best_row)
          and np.array_equal(results["score"][idx].view(np.uint32), want.score.view(np.uint32))
          and np.array_equal(results["fit_count"][idx], want.fit_count))
    flat = np.stack([picks["slot"], picks["usedmem"], picks["usedcores"]], axis=1)
    got = [flat[int(pb[batch.pod_req_range(p)[0]]):int(pb[batch.pod_req_range(p)[1]])] for p in idx]
    got = np.concatenate(got) if got else np.zeros((0, 3), np.int32)
    ok = ok and np.array_equal(got, want.picks)
    return {"pods_checked": len(idx), "nodes": snap_whole.n_nodes, "bit_exact_vs_oracle": bool(ok)}


class Leg:
    """one resident workload on this rank's shard: context, pinned planes, device-timed steps"""

    tuning = None  # vgs_tuning overrides (--tuning), A/B runs only

    def __init__(self, v, torch, dist, local, rank, world, snap, batch, n_total, row_lo=None):
        self.v, self.torch, self.dist, self.world, self.rank = v, torch, dist, world, rank
        self.snap, self.batch = snap, batch
        self.pods, self.reqs = v.capi.pack_batch(batch, snap.type_names)
        self.n_picks = int(self.reqs["nums"].sum())
        P, R = len(self.pods), len(self.reqs)
        self.sc = v.sharding.ShardedScorer(n_total, snap.dmax, max_pods=P, max_reqs=max(R, P),
                                           max_picks=max(self.n_picks, R), device=local, rank=rank, world=world,
                                           tuning=Leg.tuning)
        assert self.sc.hi - self.sc.lo == snap.n_nodes, (self.sc.lo, self.sc.hi, snap.n_nodes)
        self.ctx = self.sc.ctx
        self.pin_planes = torch.from_numpy(snap.planes).pin_memory()
        self.pin_ndev = torch.from_numpy(snap.ndev).pin_memory()
        self.plane_ptrs = [self.pin_planes[f].data_ptr() for f in range(8)]
        self.ctx.upload_ptrs(snap.n_nodes, snap.dmax, self.plane_ptrs, self.pin_ndev.data_ptr(), device=False)
        self.ctx.type_names = snap.type_names
        self.ctx.pods_upload(self.pods, self.reqs)
        self.dev = torch.device("cuda", local)

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def step(self):
        self.ctx.refresh()                      # K0
        self.sc.step_resident(self.n_picks)     # K1 (+ local K3 + all-gather + K2x), else K3

    def timed(self, steps, warmup, flush):
        """W untimed steps, then K steps bracketed by CUDA events on the library's stream, L2 flushed between
        steps; returns (ms per step = max over ranks of the summed step times / K, K1 ms, launches)"""
        torch = self.torch
        torch.cuda.set_stream(self.sc.stream)
        for _ in range(warmup):
            flush.zero_()
            self.step()
        self.barrier()
        l0 = self.ctx.timing().kernel_launches
        starts = [torch.cuda.Event(enable_timing=True) for _ in range(steps)]
        ends = [torch.cuda.Event(enable_timing=True) for _ in range(steps)]
        score_ms = []
        self.barrier()
        for i in range(steps):
            flush.zero_()          # evict the snapshot from L2 between steps (not timed)
            starts[i].record()
            self.step()
            ends[i].record()
            if i % 8 == 7 or i == steps - 1:
                score_ms.append(self.ctx.timing().score_ms)  # syncs; last step's K1 time
        self.barrier()
        launches = self.ctx.timing().kernel_launches - l0
        step_ms = [s.elapsed_time(e) for s, e in zip(starts, ends)]
        total_ms = torch.tensor([sum(step_ms)], dtype=torch.float64, device=self.dev)
        k1_ms = torch.tensor([sum(score_ms) / len(score_ms)], dtype=torch.float64, device=self.dev)
        if self.world > 1:
            self.dist.all_reduce(total_ms, op=self.dist.ReduceOp.MAX)
            self.dist.all_reduce(k1_ms, op=self.dist.ReduceOp.MAX)
        return float(total_ms.item()) / steps, float(k1_ms.item()), int(launches)

    def close(self):
        self.sc.close()
        del self.pin_planes, self.pin_ndev


def whole_snapshot(v, np, make_shard, world):
    shards = [make_shard(r) for r in range(world)]
    if world == 1:
        return shards[0]
    return v.Snapshot(np.concatenate([s_.planes for s_ in shards], axis=1), np.concatenate([s_.ndev for s_ in shards]),
                      shards[0].type_names)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--nodes-per-gpu", type=int, default=NODES_PER_GPU)
    ap.add_argument("--pods", type=int, default=N_PODS)
    ap.add_argument("--config", default="cfg3")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-hbm-leg", action="store_true", help="skip the single-pod 1M-node (HBM-bound regime) leg")
    ap.add_argument("--no-latency", action="store_true", help="skip the single-pod Filter latency leg")
    ap.add_argument("--no-configs", action="store_true", help="skip the other BASELINE configs (cfg2 / cfg4 / cfg5)")
    ap.add_argument("--check", action="store_true", help="(kept for old command lines: the oracle check always runs)")
    ap.add_argument("--tuning", default="", help="vgs_tuning overrides for A/B runs, e.g. early_scouts=0,scout_split=4")
    ap.add_argument("--sweep", action="store_true", help="CPU/GPU crossover grid instead of the bench line")
    ap.add_argument("--sweep-nodes", default="64,512,4096,32768,100000,1000000")
    ap.add_argument("--sweep-pods", default="1,16,256,4096")
    ap.add_argument("--sweep-out", default="")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    if args.sweep:
        return run_sweep(args)
    args.warmup = max(args.warmup, 3)
    if args.tuning:
        Leg.tuning = {k: int(x) for k, x in (kv.split("=") for kv in args.tuning.split(","))}

    import numpy as np
    import torch
    import torch.distributed as dist

    import vgpu_b200 as v

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA GPU: the scoring path has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    numa = pin_to_gpu_numa(torch, local) if world > 1 else None  # host buffers next to this rank's GPU (e2e legs)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    v.load_library()
    cfg_idx = 77 if args.config == "stress" else int(args.config[3:])

    # ---- headline workload: every rank builds its own shard (rows [rank*npg, (rank+1)*npg)) -------
    npg = args.nodes_per_gpu
    n_total = npg * world
    make_shard = (lambda r: v.workload.make_stress_snapshot(npg, v.workload.SEED + 77 + 1000 * r)) if args.config == "stress" \
        else (lambda r: v.workload.make_snapshot(npg, v.workload.SEED + cfg_idx + 1000 * r))
    snap = make_shard(rank)
    _, batch = v.workload.make_config(args.config, 8, args.pods)
    leg = Leg(v, torch, dist, local, rank, world, snap, batch, n_total)
    ctx, sc, pods, reqs, n_picks = leg.ctx, leg.sc, leg.pods, leg.reqs, leg.n_picks
    P, R, D = len(pods), len(reqs), snap.dmax
    flush = torch.empty(L2_FLUSH_BYTES, dtype=torch.uint8, device=dev)
    barrier = leg.barrier

    sampler = ClockSampler(local)
    sampler.start()
    ms_per_step, k1_ms, launches = leg.timed(args.steps, args.warmup, flush)
    value = P * n_total / (ms_per_step * 1e-3)
    results, picks = ctx.results_download()
    placed = int((results["best_row"] >= 0).sum())

    # ---- end-to-end leg: host buffers through the C ABI, every step -----------------------
    e2e = e2e_delta = None
    if not args.no_e2e:
        h2d = snap.planes.nbytes + snap.ndev.nbytes + pods.nbytes + reqs.nbytes
        d2h = results.nbytes + picks.nbytes

        def e2e_step():
            # pinned H2D + K0, queued: the request packing below overlaps the copy
            ctx.upload_ptrs(npg, D, leg.plane_ptrs, leg.pin_ndev.data_ptr(), device=False, asynchronous=True)
            ctx.pods_upload(pods, reqs)
            sc.step_resident(n_picks)
            return ctx.results_download()

        def run_e2e(fn, n):
            for _ in range(3):
                fn()
            barrier()
            t0 = time.perf_counter()
            for _ in range(n):
                fn()
            barrier()
            dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
            if world > 1:
                dist.all_reduce(dt, op=dist.ReduceOp.MAX)
            return float(dt.item()) / n

        n_e2e = max(5, min(args.steps, 30))
        sec = run_e2e(e2e_step, n_e2e)
        e2e_int32 = {"value": P * n_total / sec, "unit": UNIT, "h2d_bytes_per_step": int(h2d),
                     "d2h_bytes_per_step": int(d2h), "ms_per_step": 1e3 * sec, "steps": n_e2e,
                     "api": "vgs_snapshot_upload_async + vgs_pods_upload/score/resolve/results_download (host buffers, "
                            "int32 planes: the e2e figure of the earlier rounds)"}
        # the same step with the snapshot in the narrow wire format (vgs_snapshot_upload_narrow): every plane in the
        # narrowest integer type its values fit, widened on the GPU by the kernel that pulls it over PCIe
        narrow = v.capi.narrow_planes(snap)
        widths = [a.dtype.itemsize for a in narrow]
        pin_narrow = [torch.from_numpy(a.reshape(-1)).pin_memory() for a in narrow]
        nptrs = [t.data_ptr() for t in pin_narrow]

        def e2e_step_narrow():
            ctx.upload_narrow_ptrs(npg, D, nptrs[:8], nptrs[8], widths, asynchronous=True)
            ctx.pods_upload(pods, reqs)
            sc.step_resident(n_picks)
            return ctx.results_download()

        ra, rb = e2e_step(), e2e_step_narrow()
        if not all(np.array_equal(x, y) for x, y in zip(ra, rb)):
            print("bench: narrow-plane upload changed the results", file=sys.stderr)
            sys.exit(3)
        sec = run_e2e(e2e_step_narrow, n_e2e)
        h2d_n = sum(a.nbytes for a in narrow) + pods.nbytes + reqs.nbytes
        e2e = {"value": P * n_total / sec, "unit": UNIT, "h2d_bytes_per_step": int(h2d_n),
               "d2h_bytes_per_step": int(d2h), "ms_per_step": 1e3 * sec, "steps": n_e2e,
               "api": "vgs_snapshot_upload_narrow (queued) + vgs_pods_upload/score/resolve/results_download (host buffers)",
               "plane_bytes_per_element": widths, "results_equal_int32_upload": True}
        if numa is not None:
            e2e["host_numa_node"] = numa  # rank 0's; every rank runs on the cores next to its own GPU
        # ... and with the registration half of the snapshot (Count, Totalmem, Totalcore, Numa, type, len(Devices)) left
        # resident: the upload of a Filter whose node registry has not changed -- Used / Usedmem / Usedcores only
        uptrs = [nptrs[f] if f in (0, 2, 4) else None for f in range(8)]

        def e2e_step_usage():
            ctx.upload_narrow_ptrs(npg, D, uptrs, None, widths, asynchronous=True)
            ctx.pods_upload(pods, reqs)
            sc.step_resident(n_picks)
            return ctx.results_download()

        rc_ = e2e_step_usage()
        if not all(np.array_equal(x, y) for x, y in zip(ra, rc_)):
            print("bench: usage-plane upload changed the results", file=sys.stderr)
            sys.exit(3)
        sec = run_e2e(e2e_step_usage, n_e2e)
        e2e_usage = {"value": P * n_total / sec, "unit": UNIT, "ms_per_step": 1e3 * sec, "steps": n_e2e,
                     "h2d_bytes_per_step": int(sum(narrow[f].nbytes for f in (0, 2, 4)) + pods.nbytes + reqs.nbytes),
                     "d2h_bytes_per_step": int(d2h), "results_equal_int32_upload": True,
                     "api": "vgs_snapshot_upload_narrow with the usage planes only (Used, Usedmem, Usedcores; the "
                            "registration planes stay resident) + pods upload / score / resolve / download"}

        # the production path (scheduler.go:280-297 folded into deltas): the snapshot stays resident, each step
        # ships the usage changes of the previous step's placements + the pod batch, and reads the results back
        pod_of_req = np.repeat(np.arange(P), pods["n_reqs"])
        pod_of_pick = pod_of_req[np.repeat(np.arange(R), reqs["nums"])]
        state = {"sign": 1, "deltas": None, "bytes": 0}

        def deltas_of(res, pk):
            rows = res["best_row"][pod_of_pick]
            ok = rows >= 0
            d = np.zeros(int(ok.sum()), dtype=v.capi.DELTA_DTYPE)
            d["row"], d["slot"] = rows[ok], pk["slot"][ok]
            d["d_used"], d["d_usedmem"], d["d_usedcores"] = 1, pk["usedmem"][ok], pk["usedcores"][ok]
            return d

        def e2e_delta_step():
            if state["sign"] > 0:      # commit the previous winners ...
                state["deltas"] = deltas_of(*state["last"])
                d = state["deltas"]
            else:                      # ... and take them away again next step: the snapshot stays bounded
                d = state["deltas"].copy()
                for k in ("d_used", "d_usedmem", "d_usedcores"):
                    d[k] = -d[k]
            state["sign"] = -state["sign"]
            state["bytes"] = d.nbytes
            ctx.delta(d)
            ctx.pods_upload(pods, reqs)
            sc.step_resident(n_picks)
            state["last"] = ctx.results_download()
            return state["last"]

        ctx.upload_ptrs(npg, D, leg.plane_ptrs, leg.pin_ndev.data_ptr(), device=False)
        ctx.pods_upload(pods, reqs)
        sc.step_resident(n_picks)
        state["last"] = ctx.results_download()
        sec = run_e2e(e2e_delta_step, n_e2e + (n_e2e & 1))
        if state["sign"] < 0:  # leave the snapshot as uploaded
            e2e_delta_step()
        e2e_delta = {"value": P * n_total / sec, "unit": UNIT, "ms_per_step": 1e3 * sec,
                     "h2d_bytes_per_step": int(state["bytes"] + pods.nbytes + reqs.nbytes), "d2h_bytes_per_step": int(d2h),
                     "api": "vgs_snapshot_delta (previous step's placements) + vgs_pods_upload/score/resolve/"
                            "results_download; snapshot resident"}
        ctx.upload_ptrs(npg, D, leg.plane_ptrs, leg.pin_ndev.data_ptr(), device=False)
        ctx.pods_upload(pods, reqs)
    clocks = sampler.stop()

    # ---- metric (ii): single-pod Filter latency through the C ABI, snapshot resident -------------
    latency = None
    if world == 1 and not args.no_latency:
        one = batch.select([0])
        p1, r1 = v.capi.pack_batch(one, snap.type_names)
        ts, k1s = [], []
        # the call as a cgo shim makes it: caller-owned result buffers, no allocation per call
        res1 = np.zeros(1, v.capi.RESULT_DTYPE)
        pk1 = np.zeros(max(int(r1["nums"].sum()), 1), v.capi.PICK_DTYPE)
        call = (ctx.h, 1, p1.ctypes.data, len(r1), r1.ctypes.data, res1.ctypes.data, pk1.ctypes.data)
        for i in range(200):
            t0 = time.perf_counter()
            rc = ctx.lib.vgs_score_batch(*call)
            ts.append(time.perf_counter() - t0)
            assert rc == 0
            if i % 4 == 0:
                k1s.append(ctx.timing().score_ms)
        ts, k1s = sorted(ts[20:]), sorted(k1s[5:])
        k1 = k1s[len(k1s) // 2] * 1e-3
        b1 = algorithmic_bytes(npg, D, 1, 1)
        latency = {"p50_us": 1e6 * ts[len(ts) // 2], "p99_us": 1e6 * ts[int(len(ts) * 0.99)], "nodes": npg, "pods": 1,
                   "api": "vgs_score_batch, snapshot resident: one launch (pod as launch parameters, last CTA resolves into "
                          "pinned buffers) + one stream synchronisation; k1_us = that kernel",
                   "k1_us": 1e6 * k1, "k1_algorithmic_GBps": b1 / k1 / 1e9}
        ctx.pods_upload(pods, reqs)  # restore the batch for anything that follows

    prof = latest_k1_profile()
    peak, peak_src = measured_peak()
    # ---- the regime where the pass IS HBM-bound (P <= ~6): one pod, 1M nodes, snapshot
    # (256 MB) larger than L2, so every call streams it from HBM ----------------------------------
    hbm_regime = None
    if world == 1 and not args.no_hbm_leg and args.config == "cfg3":
        reps = max(1, 1_000_000 // npg)
        big = v.Snapshot(np.tile(snap.planes, (1, reps, 1)), np.tile(snap.ndev, reps), snap.type_names)
        with v.Context(max_nodes=big.n_nodes, max_devices=D, max_pods=1, device=local) as c1:
            c1.upload(big)
            one = batch.select([0])
            p1, r1 = v.capi.pack_batch(one, snap.type_names)
            ks = []
            for i in range(40):
                c1.score_packed(p1, r1)
                ks.append(c1.timing().score_ms)
            k = sorted(ks[10:])[len(ks[10:]) // 2] * 1e-3
            b = algorithmic_bytes(big.n_nodes, D, 1, 1)
            tr = (prof or {}).get("p1_dram_bytes_per_launch")
            hbm_regime = {"workload": f"{big.n_nodes} nodes x {D} devices x 1 pod, snapshot resident (256 MB > L2)",
                          "bound": "hbm", "kernel": "k1_score_single<8,true,false,true>: the whole call in one launch (score, last CTA "
                                                    "resolves the winner)",
                          # the fraction that follows from what the kernel moves (ncu DRAM bytes per launch), on this
                          # run's CUDA-event time; the algorithmic-bytes figure (all 8 prepared words per device,
                          # of which the kernel reads 4-6) is kept beside it under its own name
                          "achieved": (tr / k / 1e9) if tr else None, "peak": peak, "unit": "GB/s",
                          "frac": (tr / k / 1e9 / peak) if tr else None, "traffic": tr,
                          "algorithmic_bytes": int(b), "algorithmic_GBps": b / k / 1e9,
                          "kernel_us": 1e6 * k, "peak_source": peak_src,
                          "traffic_source": (prof or {}).get("file")}
        del big

    # ---- the other BASELINE configs, each a resident device-timed leg with its own oracle check ------------------
    configs = {}
    steps_c, warm_c = max(3, min(args.steps, 10)), 3
    if not args.no_configs and args.config == "cfg3":
        leg.close()
        plan = []
        if world == 1:
            plan = [("cfg2", "cfg2", None, None, "10k nodes x 8 devices x 256 single-container pods (Nums 1, MB, cores)"),
                    ("cfg4", "cfg4", None, None, "100k nodes x 4096 multi-container pods, 2-8 vGPUs, 10% numa-bind"),
                    ("cfg5_shard", "cfg5", 125_000, 16384, "one of cfg5's eight shards: 125k nodes x 16384 mixed pods"),
                    ("stress", "stress", None, None, "robustness, not a BASELINE config: 100k ragged nodes, Count 1..32 per device, 48 "
                     "types mixed inside nodes, scores rising along the walk; the cfg3 pod mix")]
        else:
            plan = [("cfg5", "cfg5", 1_000_000, 16384, f"1M nodes x 16384 mixed pods, node-sharded x{world} (strong)"),
                    ("cfg4_strong", "cfg4", None, None, f"100k nodes x 4096 multi-container pods, node-sharded x{world} (strong)")]
        for name, cfg, n_nodes, n_pods, what in plan:
            whole, b = v.workload.make_config(cfg, n_nodes, n_pods)
            lo, hi = v.sharding.shard_range(whole.n_nodes, world, rank)
            lg = Leg(v, torch, dist, local, rank, world, whole.rows(lo, hi), b, whole.n_nodes)
            ms, k1c, ln = lg.timed(steps_c, warm_c, flush)
            rs, pk = lg.ctx.results_download()
            pl = lg.ctx.plan()
            rec = {"workload": what, "nodes_total": whole.n_nodes, "pods": b.n_pods, "requests": b.n_reqs,
                   "ms_per_step": ms, "value": b.n_pods * whole.n_nodes / (ms * 1e-3), "unit": UNIT, "k1_ms": k1c,
                   "steps": steps_c, "gpu_launches": ln, "scaling": "strong" if world > 1 else "single GPU",
                   "kernels": [["none", "pair", "sliced", "general"][pl.cls[i].kernel] for i in range(4)],
                   "class_pods": [pl.cls[i].pods for i in range(4)], "score_set_overflow": bool(pl.dict_full)}
            if rank == 0:
                rec["check"] = sampled_check(v, np, whole, b, rs, pk)
            configs[name] = rec
            lg.close()
            del whole, lg

    # ---- sequential placement of the whole cfg3 batch: vgs_drain_batch, snapshot resident, the C
    # call with host buffers (request records up, results + picks down); chunked path against the per-pod chain,
    # pod for pod, and the first pods against P sequential oracle Filters ------------------------------------------
    drain = None
    if world == 1 and not args.no_configs and args.config == "cfg3":
        if not configs:
            leg.close()
        import oracle
        whole_d, bd = v.workload.make_config("cfg3")
        pd_, rd_ = v.capi.pack_batch(bd, whole_d.type_names)
        res_d = {k: np.zeros(len(pd_), dtype=v.capi.RESULT_DTYPE) for k in (0, -1)}
        pk_d = {k: np.zeros(int(rd_["nums"].sum()), dtype=v.capi.PICK_DTYPE) for k in (0, -1)}
        t_d, plan_d = {}, None
        for chunk in (0, -1):
            with v.Context(max_nodes=whole_d.n_nodes, max_devices=D, max_pods=len(pd_), max_reqs=len(rd_),
                           max_picks=int(rd_["nums"].sum()), device=local, tuning={"drain_chunk": chunk}) as cd:
                ts = []
                for _ in range(4 if chunk == 0 else 2):
                    cd.upload(whole_d)
                    t0 = time.perf_counter()
                    rc = cd.lib.vgs_drain_batch(cd.h, len(pd_), pd_.ctypes.data, len(rd_), rd_.ctypes.data,
                                                res_d[chunk].ctypes.data, pk_d[chunk].ctypes.data)
                    ts.append(time.perf_counter() - t0)
                    cd._check(rc)
                t_d[chunk] = min(ts[1:])
                if chunk == 0:
                    plan_d = cd.plan()
        n_chk = 96
        planes_o = whole_d.planes.copy()
        ok_o = True
        pb = bd.pick_begin
        for q in range(n_chk):
            ro = oracle.score_packed(v.Snapshot(planes_o, whole_d.ndev, whole_d.type_names), bd.select([q]))
            r0, r1 = bd.pod_req_range(q)
            mine = pk_d[0][int(pb[r0]):int(pb[r1])]
            ok_o = ok_o and int(ro.best_row[0]) == int(res_d[0]["best_row"][q]) and int(ro.fit_count[0]) == int(res_d[0]["fit_count"][q]) \
                and ro.score[:1].view(np.uint32)[0] == res_d[0]["score"][q:q + 1].view(np.uint32)[0] \
                and np.array_equal(ro.picks, np.stack([mine["slot"], mine["usedmem"], mine["usedcores"]], axis=1))
            if ro.best_row[0] >= 0:
                for sl, m_, c_ in ro.picks:
                    planes_o[0, ro.best_row[0], sl] += 1
                    planes_o[2, ro.best_row[0], sl] += m_
                    planes_o[4, ro.best_row[0], sl] += c_
        drain = {"workload": f"the cfg3 batch placed pod after pod: {len(pd_)} pods x {whole_d.n_nodes} nodes, each pod scored "
                             "against the snapshot as the winners before it left it (P sequential Filters + addPod)",
                 "api": "vgs_drain_batch, snapshot resident, host buffers; best wall clock of 3 calls",
                 "ms": 1e3 * t_d[0], "us_per_pod": 1e6 * t_d[0] / len(pd_), "chunks": plan_d.drain_chunks,
                 "rescans": plan_d.drain_rescans, "per_pod_chain_ms": 1e3 * t_d[-1],
                 "equals_per_pod_chain": bool(np.array_equal(res_d[0], res_d[-1]) and np.array_equal(pk_d[0], pk_d[-1])),
                 "check": {"pods_checked": n_chk, "bit_exact_vs_sequential_oracle": bool(ok_o)},
                 "placed": int((res_d[0]["best_row"] >= 0).sum())}
        del whole_d

    # ---- the same sharded pass through vgs_group_* : ONE process (rank 0) driving all N GPUs, exchange by NCCL
    # all-gather and by peer-mapped stores.  The other ranks wait on the rendezvous store (a CPU wait: an NCCL
    # barrier would keep their GPUs spinning).  Wall clock per vgs_group_score_batch call, snapshot resident.
    group = None
    whole = None
    if world > 1 and args.config == "cfg3" and not args.no_configs:
        if not configs:
            leg.close()
        torch.cuda.synchronize()
        store = dist.distributed_c10d._get_default_store()
        if rank == 0:
            whole = whole_snapshot(v, np, make_shard, world)
            group = {}
            for name, ex in (("nccl", v.capi.EXCHANGE_NCCL), ("peer", v.capi.EXCHANGE_PEER)):
                try:
                    with v.Group(list(range(world)), max_nodes=whole.n_nodes, max_devices=D, max_pods=P, max_reqs=max(R, P),
                                 max_picks=max(n_picks, R), exchange=ex) as g:
                        g.upload(whole)
                        for _ in range(3):
                            rg, pg = g.score_packed(pods, reqs)
                        ts, xs = [], []
                        for _ in range(20):
                            t0 = time.perf_counter()
                            rg, pg = g.score_packed(pods, reqs)
                            ts.append(time.perf_counter() - t0)
                            xs.append(g.timing()[1])
                        ts.sort(), xs.sort()
                        group[name] = {"ms_per_call": 1e3 * ts[len(ts) // 2], "value": P * whole.n_nodes / ts[len(ts) // 2],
                                       "unit": UNIT, "exchange_us": xs[len(xs) // 2],
                                       "equals_torchrun_result": bool(np.array_equal(rg, results) and np.array_equal(pg, picks))}
                except Exception as e:  # reported, not fatal: the torchrun path above is the measured headline
                    group[name] = {"error": str(e)[:300]}
            group["api"] = ("vgs_group_score_batch: request records H2D to every GPU + score + local resolve + exchange + "
                            "merge + results D2H, one caller thread + one helper thread per further GPU, snapshot resident; median wall clock of 20 calls")
            store.set("vgs_group_leg_done", "1")
        else:
            store.wait(["vgs_group_leg_done"])

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    if whole is None:
        whole = whole_snapshot(v, np, make_shard, world)
    check = sampled_check(v, np, whole, batch, results, picks)
    if not check["bit_exact_vs_oracle"]:
        print(json.dumps({"error": "result differs from the oracle", "check": check}), flush=True)
        raise SystemExit(3)
    for name, rec in configs.items():
        if not rec["check"]["bit_exact_vs_oracle"]:
            print(json.dumps({"error": f"{name}: result differs from the oracle", "check": rec["check"]}), flush=True)
            raise SystemExit(3)

    abytes = algorithmic_bytes(npg, D, P, R)
    achieved = abytes / (k1_ms * 1e-3) / 1e9
    props = torch.cuda.get_device_properties(local)
    sm_hz = 1e6 * float(clocks.get("sm_mhz") or clocks.get("sm_max_mhz") or 1965.0)
    issue_peak = props.multi_processor_count * 4 * sm_hz
    winst = (prof or {}).get("k1_warp_instructions_per_step") if (args.config == "cfg3" and npg == NODES_PER_GPU and P == N_PODS) else None
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "int32", "data": "synthetic",
        "config": {"workload": WORKLOAD if args.config == "cfg3" else f"{args.config}: {npg} nodes per GPU x {P} pods",
                   "nodes_per_gpu": npg,
                   "nodes_total": n_total, "devices_per_node": D, "pods": P, "requests": R,
                   "parallelism": f"node-shard x{world}", "l2": "256 MiB memset between steps (untimed)",
                   "placed_pods": placed},
        "e2e": e2e,
        "gpu_launches": int(launches),
        "clocks": clocks,
        "roofline": {"bound": "hbm", "kernel": "K1 phase of one step (every score launch between the K0 and K3 events)",
                     "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": (prof or {}).get("k1_dram_bytes_per_step") if winst else None,
                     "peak_source": peak_src, "algorithmic_bytes": int(abytes),
                     "kernel_ms": k1_ms,
                     # SURVEY 8(d) "always": warp instructions of the K1 phase (ncu launch list of this command,
                     # file named below) / K1 time of THIS run / (SMs x 4 schedulers x SM clock sampled in this run)
                     "issue_fraction": (winst / (k1_ms * 1e-3) / issue_peak) if winst else None,
                     "warp_instructions": winst, "issue_peak_per_s": issue_peak,
                     "profile": (prof or {}).get("file"), "profile_tree": (prof or {}).get("tree"),
                     "note": "batch regime is instruction-issue bound, not HBM bound (SURVEY F6, DESIGN.md 4)"},
        "device_evals_per_s": P * n_total * D / (ms_per_step * 1e-3),
        "check": check,
    }
    if e2e:
        line["e2e_int32_planes"] = e2e_int32
        line["e2e_usage_planes"] = e2e_usage
    if e2e_delta:
        line["e2e_delta"] = e2e_delta
    if configs:
        line["configs"] = configs
    if drain:
        line["drain"] = drain
        if not (drain["equals_per_pod_chain"] and drain["check"]["bit_exact_vs_sequential_oracle"]):
            print(json.dumps({"error": "drain: result differs", "drain": drain}), flush=True)
            raise SystemExit(3)
    if group:
        line["group"] = group
    if latency:
        line["filter_latency"] = latency
    if hbm_regime:
        line["roofline_hbm_regime"] = hbm_regime
    if not args.no_cpu and world == 1:  # reported baselines, measured once (N = 1), bounded samples
        import oracle
        val, th, dt, sp = cpu_arm(snap, batch, 64)   # calibrate
        sample = max(64, min(P, int(64 * 12.0 / max(dt, 1e-3))))
        val, th, dt, sp = cpu_arm(snap, batch, sample)
        line["cpu_baseline"] = {"value": val, "unit": UNIT, "cores": th, "kind": "port",
                                "sample": f"first {sp} of {P} pods x all {npg} nodes ({dt:.1f} s), oracle packed "
                                          "restatement (R2); reference Go not runnable here"}
        # R1: the reference's own shape (pointers to structs, string checkType per device, one thread), and
        # R1-io: the same with the reference's per-device log lines formatted and written to /dev/null
        rows = {}
        for key, io, budget in (("r1_faithful_single_thread", False, 4.0), ("r1_io_with_log_lines", True, 4.0)):
            r = oracle.score_faithful(snap, batch.select([0]), log_io=io)
            n = max(1, min(32, int(budget / max(r.seconds_scoring, 1e-4))))
            r = oracle.score_faithful(snap, batch.select(list(range(n))), log_io=io)
            rows[key] = {"value": n * npg / r.seconds_scoring, "unit": UNIT, "cores": 1,
                         "sample": f"first {n} pods x all {npg} nodes ({r.seconds_scoring:.1f} s)",
                         "snapshot_rebuild_s_per_filter": r.seconds_rebuild / n}
        line["cpu_baseline"]["other_rows"] = rows
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
