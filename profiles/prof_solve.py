"""Short profiling driver: one cfg2-shaped solve capped at --iters pivots."""
import argparse, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mvolps_b200 import capi, instances
from mvolps_b200.lp import EngineLP

ap = argparse.ArgumentParser()
ap.add_argument("--m", type=int, default=1024)
ap.add_argument("--n", type=int, default=2048)
ap.add_argument("--iters", type=int, default=1500)
ap.add_argument("--G", type=int, default=0)
ap.add_argument("--reps", type=int, default=2)
a = ap.parse_args()
L = capi.load_product()
assert L.mvb_device_init(0) == 0, L.mvb_last_error()
t = EngineLP.from_instance(L, instances.dense_lp(a.m, a.n, 1234))
w = t.clone()
for _ in range(a.reps):
    w.assign(t)
    r = w.solve(max_iters=a.iters, ctas_per_lp=a.G)
    ms = L.mvb_last_kernel_ms()
    print(f"pivots {r['pivots']} kernel {ms:.3f} ms -> {ms * 1e3 / max(r['pivots'], 1):.2f} us/pivot, "
          f"{16 * (a.m + 1) * (a.n + 1) * r['pivots'] / ms / 1e6:.0f} GB/s algorithmic")
