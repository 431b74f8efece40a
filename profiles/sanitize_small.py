"""Small end-to-end exercise for compute-sanitizer (memcheck / racecheck): primal, phase 1,
dual warm start with a cut row, a batch, a sharded group, the lazy loop with several CTAs per LP."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mvolps_b200 import capi, instances
from mvolps_b200.dist import ShardGroup
from mvolps_b200.lp import EngineLP, GlpProb, solve_batch

L = capi.load_product()
assert L.mvb_device_init(0) == 0
L.glp_term_out(0)
for G in (1, 3):
    lp = EngineLP.from_instance(L, instances.dense_lp(33, 70, 3))
    print("primal", G, lp.solve(ctas_per_lp=G)["iters"])
    lp = EngineLP.from_instance(L, instances.general_lp(20, 30, 12))
    print("phase1", G, lp.solve(ctas_per_lp=G)["iters"])
    lp = EngineLP.from_instance(L, instances.mdkp(40, 10, 21), row_cap=12)
    lp.solve(ctas_per_lp=G)
    x = lp.primal()[:40]
    j = [k for k in range(40) if x[k] != np.floor(x[k])][0]
    lp.set_bounds(j, -np.inf, np.floor(x[j]))
    row = np.zeros(40); row[:8] = 1.0
    lp.append_row(row, -np.inf, np.floor(row @ x) - 1.0)
    print("dual", G, lp.solve(ctas_per_lp=G)["iters_dual"])
lps = [EngineLP.from_instance(L, instances.dense_lp(24 + 3 * k, 40 + 5 * k, 100 + k)) for k in range(5)]
print("batch", [r["iters"] for r in solve_batch(L, lps)])
g = ShardGroup(L, instances.dense_lp(33, 70, 3), 3)
print("shard", g.solve()["iters"])
g.free()
os.environ["MVOLPS_SHARD_EXCHANGE"] = "push"   # the fused exchange: pushes into the peers' mailboxes + arrival counters
g = ShardGroup(L, instances.dense_lp(33, 70, 3), 3)
print("shard push", g.solve()["iters"], g.last_exchange())
g.free()
del os.environ["MVOLPS_SHARD_EXCHANGE"]
# more LPs than SMs: the narrow-CTA launch shape (128 / 256 threads, several CTAs per SM)
many = [EngineLP.from_instance(L, instances.dense_lp(12 + k % 5, 20 + k % 7, 300 + k % 11)) for k in range(160)]
print("wide batch", sum(r["iters"] for r in solve_batch(L, many)))
wide = [EngineLP.from_instance(L, instances.dense_lp(40, 300, 400 + k % 3)) for k in range(150)]
print("wide batch 256", sum(r["iters"] for r in solve_batch(L, wide)))
# B&B rounds with queued clones, queued cut rows and the batched row gather
inst = instances.mdkp(20, 5, 3)
print("bnb cuts", GlpProb.from_instance(L, inst).bnb(cut_mode=2, batch_width=16, max_iters=120)["launches_aux"])
p = GlpProb.from_file(L, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "examples", "small3.lp"))
print("bnb", p.bnb(cut_mode=2)["iterations"])
