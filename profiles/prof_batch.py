"""Phase clocks of the private-state loop on an HBM-streaming batch (library built with -DMVB_PROF,
e.g. tools/build_variant.sh prof -DMVB_PROF; MVOLPS_LIB=build/variants/prof.so)."""
import ctypes as C, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mvolps_b200 import capi, instances
from mvolps_b200.lp import EngineLP, solve_batch
L = capi.load_product()
L.mvb_device_init(0)
nlp = int(sys.argv[1]) if len(sys.argv) > 1 else 592
m, n = (int(sys.argv[2]), int(sys.argv[3])) if len(sys.argv) > 3 else (256, 512)
tmpl = [EngineLP.from_instance(L, instances.dense_lp(m, n, sd)) for sd in range(16)]
work = [tmpl[k % 16].clone() for k in range(nlp)]
solve_batch(L, work)
for w, k in zip(work, range(nlp)):
    w.assign(tmpl[k % 16])
have_prof = hasattr(L, "mvb_debug_prof")  # only in -DMVB_PROF builds; without it: timing only
if have_prof:
    L.mvb_debug_prof.argtypes = [C.c_void_p, C.c_int]
    L.mvb_debug_prof(None, 1)
res = solve_batch(L, work)
buf = np.zeros(256 * 16, np.int64)
if have_prof:
    L.mvb_debug_prof(buf.ctypes.data, 0)
ms = L.mvb_last_kernel_ms()
iters = sum(r["iters"] for r in res)
G, groups, nt = C.c_int(), C.c_int(), C.c_int()
L.mvb_last_launch(C.byref(G), C.byref(groups), C.byref(nt))
hsum = sum(r["pivot_hash"] for r in res) & ((1 << 64) - 1)  # same LPs, same pivots: equal at every depth / build
print(f"{nlp} LPs m={m} n={n}: kernel {ms:.2f} ms, {iters} iterations -> {iters / ms / 1e3:.2f} M LP-pivots/s; "
      f"G={G.value} groups={groups.value} threads={nt.value} hash_sum={hsum}")
if not have_prof:
    sys.exit(0)
p = buf.reshape(256, 16)
names = ["price:argmin tail", "iter:col gather+ratio loop", "iter:step+beta+swap", "iter:row derive+d+price", "iter:final argmin",
         "apply(flip)", "flush+arrive", "iter:tma wait", "price:loop", "price:warp1", "price:sync", "price:warp2",
         "iter:group barrier wait", "-", "iter:ratio argmin", "-"]
# g_prof is indexed by blockIdx.x (first 256 CTAs); every CTA solved nlp / groups LPs
tot = p[: min(groups.value, 256)].sum(0)
per_iter = tot / (iters * min(groups.value, 256) / groups.value)
for nm, c in zip(names, per_iter):
    if c > 0:
        print(f"{nm:28s} {c:8.0f} clk  ({c / 1965:.2f} us)")
print(f"sum {per_iter.sum():.0f} clk = {per_iter.sum() / 1965:.2f} us per iteration per CTA")
