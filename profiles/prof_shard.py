"""Short profiling driver: configs[4]-shaped column-sharded solve on ONE GPU (world = 1), few pivots."""
import argparse, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mvolps_b200 import capi, instances
from mvolps_b200.dist import ShardGroup

ap = argparse.ArgumentParser()
ap.add_argument("--m", type=int, default=16384)
ap.add_argument("--n", type=int, default=32768)
ap.add_argument("--pivots", type=int, default=64)
a = ap.parse_args()
L = capi.load_product()
assert L.mvb_device_init(0) == 0, L.mvb_last_error()
g = ShardGroup(L, instances.dense_lp(a.m, a.n, 5), 1)
r = g.solve(max_iters=a.pivots)
ms = g.last_ms()
print(f"pivots {r['pivots']} {ms:.2f} ms -> {1e3 * ms / r['pivots']:.1f} us/pivot, "
      f"{16 * (a.m + 1) * (a.n + 1) * r['pivots'] / ms / 1e6:.0f} GB/s algorithmic")
