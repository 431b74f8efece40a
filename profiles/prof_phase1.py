"""The general kernel (simplex_kernel: composite phase 1 + phase 2, eager tile_update, one group
barrier per pivot) on LPs with equality / >= rows -- the path every LP that needs phase 1 takes.
ncu target of profiles/r2_simplex_kernel_summary.txt."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mvolps_b200 import capi, instances
from mvolps_b200.lp import EngineLP
L = capi.load_product()
assert L.mvb_device_init(0) == 0
m, n = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (512, 1024)
inst = instances.general_lp(m, n, 31)
tmpl = EngineLP.from_instance(L, inst)
work = tmpl.clone()
for k in range(3):
    work.assign(tmpl)
    r = work.solve()
    ms = L.mvb_last_kernel_ms()
    print(f"general LP {m}x{n}: status {r['status']}, {r['iters']} iterations ({r['iters_primal1']} phase 1, "
          f"{r['iters_primal2']} phase 2, {r['pivots']} pivots), last kernel {ms:.2f} ms = {1e3 * ms / max(r['iters'], 1):.2f} us "
          f"per iteration; algorithmic {16.0 * (m + 1) * (n + 1) * r['pivots'] / (ms * 1e-3) / 1e9:.0f} GB/s")
