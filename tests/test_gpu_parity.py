"""GPU <-> oracle bit parity through the engine C ABI (the parity tests proper)."""
import numpy as np
import pytest

from mvolps_b200 import instances
from mvolps_b200.lp import EngineLP, solve_batch

pytestmark = pytest.mark.gpu

KEYS = ["status", "iters", "pivots", "flips", "iters_primal1", "iters_primal2", "iters_dual",
        "bland_iters", "pivot_hash", "primal_feasible"]


def _pair(oracle, product, inst, **opts):
    a = EngineLP.from_instance(oracle, inst)
    b = EngineLP.from_instance(product, inst)
    ra = a.solve(log_pivots=1, **{k: v for k, v in opts.items() if k != "ctas_per_lp"})
    rb = b.solve(log_pivots=1, **opts)
    return a, b, ra, rb


def _assert_same(a, b, ra, rb):
    for k in KEYS:
        assert ra[k] == rb[k], (k, ra[k], rb[k])
    la, _ = a.pivot_log()
    lb, _ = b.pivot_log()
    assert np.array_equal(la, lb)
    assert ra["obj"] == rb["obj"]  # bit-exact fp64
    Ta, ba = a.tableau()
    Tb, bb = b.tableau()
    assert np.array_equal(Ta, Tb)
    assert np.array_equal(ba, bb)
    assert np.array_equal(a.primal(), b.primal())
    for x, y in zip(a.basis(), b.basis()):
        assert np.array_equal(x, y)


@pytest.mark.parametrize("m,n,seed", [(4, 6, 1), (8, 12, 2), (33, 70, 3), (64, 128, 4), (128, 256, 5),
                                       (100, 37, 6), (257, 300, 7)])
@pytest.mark.parametrize("G", [1, 4, 0])
def test_dense_lp_primal_parity(oracle, product, m, n, seed, G):
    a, b, ra, rb = _pair(oracle, product, instances.dense_lp(m, n, seed), ctas_per_lp=G)
    assert ra["status"] == 0 and ra["pivots"] > 0
    _assert_same(a, b, ra, rb)


@pytest.mark.parametrize("m,n,seed", [(6, 8, 11), (20, 30, 12), (50, 80, 13), (100, 150, 14), (40, 200, 15)])
@pytest.mark.parametrize("G", [1, 3, 0])
def test_general_lp_phase1_parity(oracle, product, m, n, seed, G):
    a, b, ra, rb = _pair(oracle, product, instances.general_lp(m, n, seed), ctas_per_lp=G)
    assert ra["iters_primal1"] > 0
    _assert_same(a, b, ra, rb)


@pytest.mark.parametrize("seed", [21, 22, 23])
@pytest.mark.parametrize("G", [1, 2])
def test_warm_start_dual_parity(oracle, product, seed, G):
    """Branching-style bound changes and an appended cut row re-solved by the dual simplex."""
    inst = instances.mdkp(40, 10, seed)
    a = EngineLP.from_instance(oracle, inst, row_cap=12)
    b = EngineLP.from_instance(product, inst, row_cap=12)
    ra, rb = a.solve(log_pivots=1), b.solve(log_pivots=1, ctas_per_lp=G)
    _assert_same(a, b, ra, rb)
    x = a.primal()[:40]
    frac = [j for j in range(40) if x[j] != np.floor(x[j])]
    assert frac
    j = frac[0]
    for lp in (a, b):
        lp.set_bounds(j, -np.inf, np.floor(x[j]))  # bs.cpp:274 drops the lower bound
    ra, rb = a.solve(log_pivots=1), b.solve(log_pivots=1, ctas_per_lp=G)
    assert ra["iters_dual"] > 0
    _assert_same(a, b, ra, rb)
    # cut row: sum of the first 8 columns <= floor(current activity) (append grows past row_cap)
    for t in range(4):
        xa = a.primal()[:40]
        row = np.zeros(40)
        row[t:t + 8] = 1.0
        rhs = np.floor(row @ xa) - 1.0
        for lp in (a, b):
            lp.append_row(row, -np.inf, rhs)
        ra, rb = a.solve(log_pivots=1), b.solve(log_pivots=1, ctas_per_lp=G)
        _assert_same(a, b, ra, rb)
    ca, cb = a.clone(), b.clone()
    for lp in (ca, cb):
        lp.set_bounds(j, np.floor(x[j]) + 1.0, np.inf)  # bs.cpp:282
    ra, rb = ca.solve(log_pivots=1), cb.solve(log_pivots=1, ctas_per_lp=G)
    _assert_same(ca, cb, ra, rb)


def test_batch_parity(oracle, product):
    insts = [instances.dense_lp(24 + 3 * k, 40 + 5 * k, 100 + k) for k in range(37)]
    A = [EngineLP.from_instance(oracle, i) for i in insts]
    B = [EngineLP.from_instance(product, i) for i in insts]
    ra = solve_batch(oracle, A, log_pivots=1)
    rb = solve_batch(product, B, log_pivots=1)
    for a, b, x, y in zip(A, B, ra, rb):
        _assert_same(a, b, x, y)


@pytest.mark.parametrize("kind", ["infeasible", "unbounded"])
def test_status_parity(oracle, product, kind):
    inst = instances.dense_lp(10, 15, 5)
    if kind == "infeasible":
        inst["row_lb"][0] = inst["row_ub"][0] + 100.0  # row 0: lb > max activity ... force conflict
        inst["row_ub"][0] = np.inf
        inst["col_ub"][:] = 0.5
    else:
        inst["A"][:, 3] = -np.abs(inst["A"][:, 3])
    a, b, ra, rb = _pair(oracle, product, inst)
    assert ra["status"] == (1 if kind == "infeasible" else 2)
    _assert_same(a, b, ra, rb)
