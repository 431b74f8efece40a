"""Column-sharded tableau (include/mvolps_dist.h) == single-GPU engine == oracle,
bit for bit.  The ranks are emulated as a single-process group of shards on one
GPU (device-to-device all-gather); the NCCL path shares every kernel."""
import numpy as np
import pytest

from mvolps_b200 import instances
from mvolps_b200.dist import ShardGroup
from mvolps_b200.lp import EngineLP

pytestmark = pytest.mark.gpu

DUAL_KEYS = ("status", "iters", "pivots", "flips", "pivot_hash", "bland_iters", "primal_feasible",
             "iters_primal1", "iters_primal2", "iters_dual")


@pytest.mark.parametrize("m,n,seed", [(8, 12, 2), (33, 70, 3), (64, 128, 4), (100, 37, 6), (128, 256, 5)])
@pytest.mark.parametrize("world", [1, 2, 3, 8])
def test_sharded_equals_single_and_oracle(oracle, product, m, n, seed, world):
    inst = instances.dense_lp(m, n, seed)
    a = EngineLP.from_instance(oracle, inst)
    ra = a.solve(log_pivots=1)
    b = EngineLP.from_instance(product, inst)
    rb = b.solve(log_pivots=1)
    g = ShardGroup(product, inst, world)
    rg = g.solve(log_pivots=1)
    for k in ("status", "iters", "pivots", "flips", "pivot_hash", "bland_iters", "primal_feasible"):
        assert ra[k] == rb[k] == rg[k], (k, ra[k], rb[k], rg[k])
    assert ra["obj"] == rg["obj"]  # bit-exact fp64
    la, _ = a.pivot_log()
    for k in range(world):
        lg, _ = g.pivot_log(k)
        assert np.array_equal(la, lg)
        assert np.array_equal(a.primal(), g.primal(k))
    g.free()


@pytest.mark.parametrize("m,n,seed,world", [(33, 70, 3, 2), (64, 128, 4, 3), (128, 256, 5, 8), (200, 520, 8, 8)])
def test_pushed_exchange_equals_oracle(oracle, product, monkeypatch, m, n, seed, world):
    """The fused exchange of the one-process-per-GPU path (price kernel pushes its record into every
    peer's mailbox and raises the arrival counters, decide kernel waits on them; 4-slot mailbox)
    between the shards of a single-process group: same kernels, peers = the other shards' mailboxes."""
    monkeypatch.setenv("MVOLPS_SHARD_EXCHANGE", "push")
    inst = instances.dense_lp(m, n, seed)
    a = EngineLP.from_instance(oracle, inst)
    ra = a.solve(log_pivots=1)
    g = ShardGroup(product, inst, world)
    rg = g.solve(log_pivots=1)
    assert "pushed" in g.last_exchange()
    for k in ("status", "iters", "pivots", "flips", "pivot_hash", "bland_iters", "primal_feasible"):
        assert ra[k] == rg[k], (k, ra[k], rg[k])
    assert ra["obj"] == rg["obj"]
    la, _ = a.pivot_log()
    for k in range(world):
        lg, _ = g.pivot_log(k)
        assert np.array_equal(la, lg)
        assert np.array_equal(a.primal(), g.primal(k))
    g.free()
    monkeypatch.delenv("MVOLPS_SHARD_EXCHANGE")
    g = ShardGroup(product, inst, world)
    assert g.solve()["pivot_hash"] == ra["pivot_hash"] and "host copies" in g.last_exchange()
    g.free()


def test_sharded_iteration_limit_and_resume_free(product):
    inst = instances.dense_lp(64, 128, 4)
    g = ShardGroup(product, inst, 4)
    r = g.solve(max_iters=7)
    assert r["status"] == 3 and r["iters"] == 7
    g.free()


@pytest.mark.parametrize("m,n,seed", [(20, 30, 12), (12, 10, 7), (33, 70, 3), (64, 128, 4), (100, 37, 6), (128, 256, 5)])
@pytest.mark.parametrize("world", [1, 2, 3, 8])
def test_sharded_phase1_equals_single_and_oracle(oracle, product, m, n, seed, world):
    """Mixed <=, >=, = rows: the slack basis is neither primal nor dual feasible; the sharded path runs
    the composite phase 1 (infeasibility costs of the local columns summed in the specification's
    chunk order, no exchange of its own) and goes on in phase 2 when the basis turns feasible."""
    inst = instances.general_lp(m, n, seed)
    a = EngineLP.from_instance(oracle, inst)
    ra = a.solve(log_pivots=1)
    assert ra["iters_primal1"] > 0
    b = EngineLP.from_instance(product, inst)
    rb = b.solve(log_pivots=1)
    g = ShardGroup(product, inst, world)
    rg = g.solve(log_pivots=1)
    for k in DUAL_KEYS:
        assert ra[k] == rb[k] == rg[k], (k, ra[k], rb[k], rg[k])
    assert ra["obj"] == rg["obj"]
    la, _ = a.pivot_log()
    for k in range(world):
        lg, _ = g.pivot_log(k)
        assert np.array_equal(la, lg)
        assert np.array_equal(a.primal(), g.primal(k))
    g.free()


def test_sharded_phase1_pushed_infeasible_and_limit(oracle, product, monkeypatch):
    inst = instances.general_lp(64, 128, 4)
    a = EngineLP.from_instance(oracle, inst)
    ra = a.solve(max_iters=11)
    g = ShardGroup(product, inst, 3)
    rg = g.solve(max_iters=11)
    for k in DUAL_KEYS:
        assert ra[k] == rg[k], (k, ra[k], rg[k])
    g.free()
    # infeasible: two rows that contradict each other (phase 1 ends with no eligible column)
    inst = instances.general_lp(20, 30, 12)
    inst["A"][1] = inst["A"][0]
    inst["row_lb"][0], inst["row_ub"][0] = 5.0, float("inf")
    inst["row_lb"][1], inst["row_ub"][1] = -float("inf"), 3.0
    a = EngineLP.from_instance(oracle, inst)
    ra = a.solve(log_pivots=1)
    assert ra["status"] == 1
    monkeypatch.setenv("MVOLPS_SHARD_EXCHANGE", "push")
    g = ShardGroup(product, inst, 4)
    rg = g.solve(log_pivots=1)
    assert "pushed" in g.last_exchange()
    for k in DUAL_KEYS:
        assert ra[k] == rg[k], (k, ra[k], rg[k])
    la, _ = a.pivot_log()
    lg, _ = g.pivot_log(2)
    assert np.array_equal(la, lg)
    g.free()


def test_sharded_rejects_unsupported(product):
    inst = instances.dense_lp(8, 12, 2)
    inst["col_lb"] = np.full(12, 1.0)  # non-basic values would not be 0
    with pytest.raises(RuntimeError):
        ShardGroup(product, inst, 2)


@pytest.mark.parametrize("m,n,seed", [(8, 12, 2), (33, 70, 3), (64, 128, 4), (100, 37, 6), (128, 256, 5)])
@pytest.mark.parametrize("world", [1, 2, 3, 8])
def test_sharded_dual_simplex_equals_single_and_oracle(oracle, product, m, n, seed, world):
    """Covering rows: the slack basis is dual feasible, not primal feasible; the column-sharded path
    runs the dual simplex (K5a on the replicated basic values, K5b over the local columns of the
    leaving row, one record exchange per pivot) -- same pivots, same bits as the oracle."""
    inst = instances.covering_lp(m, n, seed)
    a = EngineLP.from_instance(oracle, inst)
    ra = a.solve(log_pivots=1)
    assert ra["status"] == 0 and ra["iters_dual"] == ra["iters"] > 0
    b = EngineLP.from_instance(product, inst)
    rb = b.solve(log_pivots=1)
    g = ShardGroup(product, inst, world)
    rg = g.solve(log_pivots=1)
    for k in DUAL_KEYS:
        assert ra[k] == rb[k] == rg[k], (k, ra[k], rb[k], rg[k])
    assert ra["obj"] == rg["obj"]
    la, _ = a.pivot_log()
    for k in range(world):
        lg, _ = g.pivot_log(k)
        assert np.array_equal(la, lg)
        assert np.array_equal(a.primal(), g.primal(k))
    g.free()


@pytest.mark.parametrize("m,n,seed,world", [(33, 70, 3, 2), (128, 256, 5, 8), (200, 520, 8, 8)])
def test_sharded_dual_pushed_exchange_equals_oracle(oracle, product, monkeypatch, m, n, seed, world):
    monkeypatch.setenv("MVOLPS_SHARD_EXCHANGE", "push")
    inst = instances.covering_lp(m, n, seed)
    a = EngineLP.from_instance(oracle, inst)
    ra = a.solve(log_pivots=1)
    g = ShardGroup(product, inst, world)
    rg = g.solve(log_pivots=1)
    assert "pushed" in g.last_exchange()
    for k in DUAL_KEYS:
        assert ra[k] == rg[k], (k, ra[k], rg[k])
    assert ra["obj"] == rg["obj"]
    la, _ = a.pivot_log()
    lg, _ = g.pivot_log(world - 1)
    assert np.array_equal(la, lg)
    assert np.array_equal(a.primal(), g.primal(0))
    g.free()


def test_sharded_dual_infeasible_and_iteration_limit(oracle, product):
    inst = instances.covering_lp(33, 70, 3)
    a = EngineLP.from_instance(oracle, inst)
    ra = a.solve(max_iters=9)
    g = ShardGroup(product, inst, 3)
    rg = g.solve(max_iters=9)
    assert ra["status"] == rg["status"] == 3 and rg["iters"] == 9 and ra["pivot_hash"] == rg["pivot_hash"]
    g.free()
    # a covering row nobody can satisfy: every column boxed, demand above what the boxes allow
    inst = instances.covering_lp(20, 30, 7)
    inst["col_ub"] = np.full(30, 1.0)
    inst["row_lb"][4] = float(inst["A"][4].sum()) + 1.0
    inst["row_ub"][4] = float("inf")
    a = EngineLP.from_instance(oracle, inst)
    ra = a.solve(log_pivots=1)
    g = ShardGroup(product, inst, 4)
    rg = g.solve(log_pivots=1)
    assert ra["status"] == rg["status"] != 0
    for k in DUAL_KEYS:
        assert ra[k] == rg[k], (k, ra[k], rg[k])
    g.free()


def test_glp_simplex_routes_to_the_shard_group(oracle, product):
    """mvb_glp_use_shards: the column-sharded engine behind the reference's own call (bs.cpp:117).
    A single-process group of 3 shards on this GPU; glp_simplex on a fresh slack-feasible LP is then
    solved column-sharded: status, objective bits, primal values and statuses of the plain solve."""
    from mvolps_b200 import capi
    from mvolps_b200.lp import GlpProb
    inst = instances.dense_lp(64, 128, 4)
    a = GlpProb.from_instance(oracle, inst)
    b = GlpProb.from_instance(product, inst)
    c = GlpProb.from_instance(product, inst)
    assert a.simplex() == 0 and b.simplex() == 0
    l0 = product.mvb_kernel_launches()
    assert product.mvb_glp_use_shards(None, 3, 1) == 0
    try:
        assert c.simplex() == 0
        launched = product.mvb_kernel_launches() - l0
        # an LP that needs the composite phase 1 is routed as well ...
        g = GlpProb.from_instance(product, instances.general_lp(20, 30, 12))
        go = GlpProb.from_instance(oracle, instances.general_lp(20, 30, 12))
        l2 = product.mvb_kernel_launches()
        assert g.simplex() == 0 and go.simplex() == 0 and g.obj_val() == go.obj_val()
        assert product.mvb_kernel_launches() - l2 >= 3 * 3 * oracle.glp_get_it_cnt(go.h)
        assert product.glp_get_it_cnt(g.h) == oracle.glp_get_it_cnt(go.h)
        # ... one whose non-basic variables do not start at 0 stays on the single-GPU path
        nz = instances.general_lp(20, 30, 12)
        nz["col_lb"] = np.full(30, 1.0)
        nz["col_ub"] = nz["col_ub"] + 1.0
        g1, g1o = GlpProb.from_instance(product, nz), GlpProb.from_instance(oracle, nz)
        l3 = product.mvb_kernel_launches()
        r1, r1o = g1.simplex(), g1o.simplex()
        assert r1 == r1o and g1.status() == g1o.status()
        assert product.mvb_kernel_launches() - l3 < 10
    finally:
        product.mvb_glp_use_shards(None, 0, 0)
    its = product.glp_get_it_cnt(b.h)
    assert launched >= 3 * 3 * its  # price / decide / update per shard and pivot: the sharded kernels ran
    # a dual-feasible start (covering rows) is routed too and runs the sharded dual simplex
    cov = instances.covering_lp(33, 70, 3)
    ca, cb = GlpProb.from_instance(oracle, cov), GlpProb.from_instance(product, cov)
    assert product.mvb_glp_use_shards(None, 3, 1) == 0
    try:
        l1 = product.mvb_kernel_launches()
        assert ca.simplex() == 0 and cb.simplex() == 0
        assert product.mvb_kernel_launches() - l1 >= 3 * 3 * oracle.glp_get_it_cnt(ca.h)
    finally:
        product.mvb_glp_use_shards(None, 0, 0)
    assert ca.status() == cb.status() == capi.GLP_OPT and ca.obj_val() == cb.obj_val()
    assert np.array_equal(ca.col_prim(), cb.col_prim())
    assert product.glp_get_it_cnt(cb.h) == oracle.glp_get_it_cnt(ca.h)
    assert product.glp_get_it_cnt(c.h) == its == oracle.glp_get_it_cnt(a.h)
    assert a.status() == b.status() == c.status() == capi.GLP_OPT
    assert a.obj_val() == b.obj_val() == c.obj_val()
    assert np.array_equal(a.col_prim(), c.col_prim())
    for j in range(1, 129):
        assert product.glp_get_col_stat(c.h, j) == oracle.glp_get_col_stat(a.h, j)
    for i in range(1, 65):
        assert product.glp_get_row_stat(c.h, i) == oracle.glp_get_row_stat(a.h, i)
        assert product.glp_get_row_prim(c.h, i) == oracle.glp_get_row_prim(a.h, i)
