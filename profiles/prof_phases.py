"""Phase clocks of the private-state loop (library built with `make PROF=1`)."""
import ctypes as C, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mvolps_b200 import capi, instances
from mvolps_b200.lp import EngineLP
L = capi.load_product()
L.mvb_device_init(0)
m, n = int(sys.argv[1]) if len(sys.argv) > 1 else 1024, int(sys.argv[2]) if len(sys.argv) > 2 else 2048
G = int(sys.argv[3]) if len(sys.argv) > 3 else 0
lp = EngineLP.from_instance(L, instances.dense_lp(m, n, 1234))
t = lp.clone()
t.solve(max_iters=200, ctas_per_lp=G)
buf = np.zeros(256 * 16, np.int64)
L.mvb_debug_prof.argtypes = [C.c_void_p, C.c_int]
L.mvb_debug_prof(None, 1)
t.assign(lp)
r = t.solve(max_iters=3000, ctas_per_lp=G)
L.mvb_debug_prof(buf.ctypes.data, 0)
ms = L.mvb_last_kernel_ms()
p = buf.reshape(256, 16)
names = ["price:argmin tail", "iter:col gather+ratio loop", "iter:step+beta+swap", "iter:row derive+d", "-", "apply(flip)", "flush+arrive", "iter:tma wait"]
print(f"m={m} n={n} pivots={r['pivots']} kernel {ms:.3f} ms -> {1e3*ms/r['iters']:.2f} us/iter")
used = p[:, 0] > 0
print("ctas", used.sum())
names += ["price:loop", "price:warp1", "price:sync", "price:warp2", "iter:group barrier wait", "-", "iter:ratio argmin", "-"]
for i, nm in enumerate(names):
    c = p[used, i] / r["iters"]
    print(f"{nm:12s} mean {c.mean():8.0f} clk  min {c.min():8.0f}  max {c.max():8.0f}   ({c.mean()/1965:.2f} us)")
print("sum", p[used, :16].sum(1).mean() / r["iters"])
