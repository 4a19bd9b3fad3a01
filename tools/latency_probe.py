"""single-pod Filter latency through the C ABI: where do the microseconds go?"""
import ctypes as C, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import vgpu_b200 as v
N = int(sys.argv[1]) if len(sys.argv) > 1 else 100_000
snap, batch = v.workload.make_config("cfg3", N, 64)
for max_pods in (1, 4096):
    with v.Context(max_nodes=N, max_pods=max_pods) as c:
        c.upload(snap)
        p1, r1 = v.capi.pack_batch(batch.select([0]), snap.type_names)
        res = np.zeros(1, v.capi.RESULT_DTYPE); pk = np.zeros(8, v.capi.PICK_DTYPE)
        lib, h = c.lib, c.h
        args = (h, 1, p1.ctypes.data, 1, r1.ctypes.data, res.ctypes.data, pk.ctypes.data)
        ts = []
        for i in range(300):
            t0 = time.perf_counter(); rc = lib.vgs_score_batch(*args); ts.append(time.perf_counter() - t0)
            assert rc == 0
        ts = np.array(ts[50:]) * 1e6
        tm = c.timing()
        ta = []
        for i in range(200):
            t0 = time.perf_counter(); c.score_packed(p1, r1); ta.append(time.perf_counter() - t0)
        ta = np.array(ta[50:]) * 1e6
        print(f"max_pods={max_pods}: raw ctypes p50 {np.median(ts):.1f} us p99 {np.percentile(ts, 99):.1f}; "
              f"Context.score_packed p50 {np.median(ta):.1f}; gpu phases (us): pods_h2d {1e3*tm.pods_h2d_ms:.1f} "
              f"score {1e3*tm.score_ms:.1f} resolve {1e3*tm.resolve_ms:.1f} d2h {1e3*tm.d2h_ms:.1f}")
