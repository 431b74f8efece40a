"""The resident kernel (simplex_res_kernel, engine_res.cuh: one CTA per LP, tableau in shared
memory, eager updates) against the oracle, bit for bit: primal phase 2, dual warm starts after
bound changes and appended cut rows, hand-over from phase 1, wide batches (grid-stride over more
LPs than resident CTAs), infeasible / unbounded / iteration-limited solves."""
import numpy as np
import pytest

from mvolps_b200 import instances
from mvolps_b200.lp import EngineLP, solve_batch

pytestmark = pytest.mark.gpu

KEYS = ["status", "iters", "pivots", "flips", "iters_primal1", "iters_primal2", "iters_dual",
        "bland_iters", "pivot_hash", "primal_feasible"]


def _same(a, b, ra, rb):
    for k in KEYS:
        assert ra[k] == rb[k], (k, ra[k], rb[k])
    la, _ = a.pivot_log()
    lb, _ = b.pivot_log()
    assert np.array_equal(la, lb)
    assert ra["obj"] == rb["obj"]
    Ta, ba = a.tableau()
    Tb, bb = b.tableau()
    assert np.array_equal(Ta, Tb) and np.array_equal(ba, bb)
    assert np.array_equal(a.primal(), b.primal())
    for x, y in zip(a.basis(), b.basis()):
        assert np.array_equal(x, y)


@pytest.mark.parametrize("m,n,seed", [(4, 6, 1), (8, 12, 2), (33, 70, 3), (64, 128, 4), (50, 200, 9), (100, 37, 6),
                                       (90, 250, 17), (20, 1000, 18)])
def test_resident_primal_parity(oracle, product, m, n, seed):
    inst = instances.dense_lp(m, n, seed)
    a, b = EngineLP.from_instance(oracle, inst), EngineLP.from_instance(product, inst)
    ra, rb = a.solve(log_pivots=1), b.solve(log_pivots=1)
    assert product.mvb_last_kernel_mode() == 2
    assert ra["status"] == 0 and ra["pivots"] > 0
    _same(a, b, ra, rb)


def test_resident_and_lazy_kernels_agree(product):
    """ctas_per_lp > 0 asks for the lazy private-state loop: same bits from both kernels."""
    inst = instances.dense_lp(64, 128, 4)
    a, b = EngineLP.from_instance(product, inst), EngineLP.from_instance(product, inst)
    ra = a.solve(log_pivots=1)
    assert product.mvb_last_kernel_mode() == 2
    rb = b.solve(log_pivots=1, ctas_per_lp=1)
    assert product.mvb_last_kernel_mode() == 1
    _same(a, b, ra, rb)


@pytest.mark.parametrize("seed", [21, 22, 23, 24])
def test_resident_dual_warm_start_parity(oracle, product, seed):
    """B&B-style: bound change on a fractional column (bs.cpp:274 drops the lower bound), then cut
    rows appended to the resident-solved tableau, each re-solved by the dual simplex."""
    inst = instances.mdkp(40, 10, seed)
    a = EngineLP.from_instance(oracle, inst, row_cap=12)
    b = EngineLP.from_instance(product, inst, row_cap=12)
    ra, rb = a.solve(log_pivots=1), b.solve(log_pivots=1)
    _same(a, b, ra, rb)
    x = a.primal()[:40]
    frac = [j for j in range(40) if x[j] != np.floor(x[j])]
    j = frac[0]
    for lp in (a, b):
        lp.set_bounds(j, -np.inf, np.floor(x[j]))
    ra, rb = a.solve(log_pivots=1), b.solve(log_pivots=1)
    assert product.mvb_last_kernel_mode() == 2 and ra["iters_dual"] > 0
    _same(a, b, ra, rb)
    for t in range(5):
        xa = a.primal()[:40]
        row = np.zeros(40)
        row[t:t + 8] = 1.0
        rhs = np.floor(row @ xa) - 1.0
        for lp in (a, b):
            lp.append_row(row, -np.inf, rhs)
        ra, rb = a.solve(log_pivots=1), b.solve(log_pivots=1)
        _same(a, b, ra, rb)
    # clone of a resident-solved LP, different branch on each copy
    a2, b2 = a.clone(), b.clone()
    xa = a.primal()[:40]
    frac = [j for j in range(40) if xa[j] != np.floor(xa[j])]
    if frac:
        j = frac[0]
        for lp in (a, b):
            lp.set_bounds(j, -np.inf, np.floor(xa[j]))
        for lp in (a2, b2):
            lp.set_bounds(j, np.ceil(xa[j]), np.inf)
        ra, ra2 = solve_batch(oracle, [a, a2], log_pivots=1)
        rb, rb2 = solve_batch(product, [b, b2], log_pivots=1)
        _same(a, b, ra, rb)
        _same(a2, b2, ra2, rb2)


@pytest.mark.parametrize("m,n,seed", [(6, 8, 11), (20, 30, 12), (50, 80, 13), (40, 200, 15)])
def test_resident_takes_over_after_phase1(oracle, product, m, n, seed):
    """An LP with equality / >= rows: phase 1 in the general kernel, which hands the LP back once a
    primal feasible basis is reached; phase 2 then runs resident, tallies carried on."""
    inst = instances.general_lp(m, n, seed)
    a, b = EngineLP.from_instance(oracle, inst), EngineLP.from_instance(product, inst)
    ra, rb = a.solve(log_pivots=1), b.solve(log_pivots=1)
    assert ra["iters_primal1"] > 0
    if ra["iters_primal2"] > 0:
        assert product.mvb_last_kernel_mode() == 2
    _same(a, b, ra, rb)


def test_resident_wide_batch(oracle, product):
    """More LPs than co-resident CTAs (grid-stride), mixed shapes in one launch."""
    shapes = [(12 + k % 7, 20 + 3 * (k % 5), 500 + k % 13) for k in range(700)]
    la = [EngineLP.from_instance(oracle, instances.dense_lp(*s)) for s in shapes[:40]]
    lb = [EngineLP.from_instance(product, instances.dense_lp(*s)) for s in shapes]
    ra = solve_batch(oracle, la, log_pivots=1)
    rb = solve_batch(product, lb, log_pivots=1)
    assert product.mvb_last_kernel_mode() == 2
    for k in range(40):
        _same(la[k], lb[k], ra[k], rb[k])
    by_shape = {}
    for s, r in zip(shapes, rb):
        assert by_shape.setdefault(s, r["pivot_hash"]) == r["pivot_hash"]  # equal LPs, equal runs


def test_resident_statuses(oracle, product):
    # unbounded: maximise x with a free upper bound
    inf = float("inf")
    unb = dict(A=np.array([[1.0, -1.0]]), c=np.array([1.0, 1.0]), sense="max", row_lb=np.array([-inf]),
               row_ub=np.array([4.0]), col_lb=np.zeros(2), col_ub=np.full(2, inf), integer=np.zeros(2, bool))
    a, b = EngineLP.from_instance(oracle, unb), EngineLP.from_instance(product, unb)
    ra, rb = a.solve(log_pivots=1), b.solve(log_pivots=1)
    assert ra["status"] == 2
    _same(a, b, ra, rb)
    # dual infeasible child: bounds that cannot be met
    inst = instances.mdkp(12, 4, 1)
    a, b = EngineLP.from_instance(oracle, inst), EngineLP.from_instance(product, inst)
    a.solve(), b.solve()
    for lp in (a, b):
        for j in range(12):
            lp.set_bounds(j, 3.0, 3.0)
    ra, rb = a.solve(log_pivots=1), b.solve(log_pivots=1)
    assert ra["status"] == 1
    _same(a, b, ra, rb)
    # iteration limit
    inst = instances.dense_lp(33, 70, 3)
    a, b = EngineLP.from_instance(oracle, inst), EngineLP.from_instance(product, inst)
    ra, rb = a.solve(log_pivots=1, max_iters=7), b.solve(log_pivots=1, max_iters=7)
    assert ra["status"] == 3 and ra["iters"] == 7
    _same(a, b, ra, rb)
    ra, rb = a.solve(log_pivots=1), b.solve(log_pivots=1)  # resume to optimality
    assert ra["status"] == 0
    _same(a, b, ra, rb)


# ---------------------------------------------------------------- cluster-resident kernel
@pytest.fixture()
def cres(product):
    """The cluster-resident kernel is opt-in (engine option "cres")."""
    old = product.mvb_engine_option(b"cres", 1)
    yield product
    product.mvb_engine_option(b"cres", old)


@pytest.mark.parametrize("m,n,seed,csize", [(128, 256, 5, 2), (100, 400, 31, 2), (257, 300, 7, 4), (200, 520, 8, 8),
                                             (256, 512, 0, 8), (300, 100, 33, 2)])
def test_cluster_resident_primal_parity(oracle, product, cres, m, n, seed, csize):
    """simplex_cres_kernel: the tableau in the distributed shared memory of a cluster of 2 / 4 / 8
    CTAs (row blocks), candidates exchanged through st.shared::cluster + the cluster barrier, the
    scaled pivot row sent by cp.async.bulk shared -> shared::cluster: oracle bits."""
    import ctypes as C
    inst = instances.dense_lp(m, n, seed)
    a, b = EngineLP.from_instance(oracle, inst), EngineLP.from_instance(product, inst)
    ra, rb = a.solve(log_pivots=1), b.solve(log_pivots=1)
    assert product.mvb_last_kernel_mode() == 3
    G, groups, nt = C.c_int(), C.c_int(), C.c_int()
    product.mvb_last_launch(C.byref(G), C.byref(groups), C.byref(nt))
    assert G.value == csize
    assert ra["status"] == 0 and ra["pivots"] > 0
    _same(a, b, ra, rb)


@pytest.mark.parametrize("nvars,nrows,seed", [(300, 100, 41), (500, 60, 42), (512, 256, 43)])
def test_cluster_resident_dual_warm_start_parity(oracle, product, cres, nvars, nrows, seed):
    """B&B-style on a tableau that needs a cluster: bound change, cut rows, clones."""
    inst = instances.mdkp(nvars, nrows, seed)
    a = EngineLP.from_instance(oracle, inst, row_cap=nrows + 2)
    b = EngineLP.from_instance(product, inst, row_cap=nrows + 2)
    ra, rb = a.solve(log_pivots=1), b.solve(log_pivots=1)
    assert product.mvb_last_kernel_mode() == 3
    _same(a, b, ra, rb)
    x = a.primal()[:nvars]
    frac = [j for j in range(nvars) if x[j] != np.floor(x[j])]
    j = frac[0]
    a2, b2 = a.clone(), b.clone()
    for lp in (a, b):
        lp.set_bounds(j, -np.inf, np.floor(x[j]))
    for lp in (a2, b2):
        lp.set_bounds(j, np.ceil(x[j]), np.inf)
    ra, ra2 = solve_batch(oracle, [a, a2], log_pivots=1)
    rb, rb2 = solve_batch(product, [b, b2], log_pivots=1)
    assert product.mvb_last_kernel_mode() == 3 and ra["iters_dual"] > 0
    _same(a, b, ra, rb)
    _same(a2, b2, ra2, rb2)
    for t in range(3):
        xa = a.primal()[:nvars]
        row = np.zeros(nvars)
        row[t:t + 8] = 1.0
        rhs = np.floor(row @ xa) - 1.0
        for lp in (a, b):
            lp.append_row(row, -np.inf, rhs)
        ra, rb = a.solve(log_pivots=1), b.solve(log_pivots=1)
        _same(a, b, ra, rb)


def test_cluster_resident_phase1_handover_batch_and_limits(oracle, product, cres):
    inst = instances.general_lp(100, 300, 14)
    a, b = EngineLP.from_instance(oracle, inst), EngineLP.from_instance(product, inst)
    ra, rb = a.solve(log_pivots=1), b.solve(log_pivots=1)
    assert ra["iters_primal1"] > 0 and ra["iters_primal2"] > 0 and product.mvb_last_kernel_mode() == 3
    _same(a, b, ra, rb)
    # more LPs than clusters in one launch, two cluster sizes in one batch
    shapes = [(128, 256, 60 + k % 5) if k % 2 else (200, 520, 70 + k % 3) for k in range(90)]
    la = [EngineLP.from_instance(oracle, instances.dense_lp(*s)) for s in shapes[:6]]
    lb = [EngineLP.from_instance(product, instances.dense_lp(*s)) for s in shapes]
    ra = solve_batch(oracle, la, log_pivots=1)
    rb = solve_batch(product, lb, log_pivots=1)
    for k in range(6):
        _same(la[k], lb[k], ra[k], rb[k])
    by_shape = {}
    for s, r in zip(shapes, rb):
        assert by_shape.setdefault(s, r["pivot_hash"]) == r["pivot_hash"]
    # iteration limit and resume
    inst = instances.dense_lp(128, 256, 5)
    a, b = EngineLP.from_instance(oracle, inst), EngineLP.from_instance(product, inst)
    ra, rb = a.solve(log_pivots=1, max_iters=9), b.solve(log_pivots=1, max_iters=9)
    assert ra["status"] == 3
    _same(a, b, ra, rb)
    ra, rb = a.solve(log_pivots=1), b.solve(log_pivots=1)
    _same(a, b, ra, rb)
    # the lazy loop on the same LP (cluster-resident switched off) gives the same bits
    product.mvb_engine_option(b"cres", 0)
    try:
        a2, c = EngineLP.from_instance(oracle, inst), EngineLP.from_instance(product, inst)
        ra2, rc = a2.solve(log_pivots=1), c.solve(log_pivots=1)
        assert product.mvb_last_kernel_mode() == 1
        _same(a2, c, ra2, rc)
        d = EngineLP.from_instance(product, inst)
        product.mvb_engine_option(b"cres", 1)
        rd = d.solve(log_pivots=1)
        assert product.mvb_last_kernel_mode() == 3
        _same(a2, d, ra2, rd)
    finally:
        product.mvb_engine_option(b"cres", 1)


def test_free_many_on_the_device(product):
    """mvb_lp_free_many: the slabs go back to the engine's free lists, new LPs reuse them and solve as before."""
    import ctypes as C
    inst = instances.mdkp(40, 10, 21)
    ref = EngineLP.from_instance(product, inst).solve()
    lps = [EngineLP.from_instance(product, inst) for _ in range(40)]
    kids = [lp.clone() for lp in lps[:10]]          # queued clones keep their sources' slabs alive until the flush
    arr = (C.c_void_p * len(lps))(*[lp.h for lp in lps])
    product.mvb_lp_free_many(arr, len(lps))
    for lp in lps:
        lp.h = None
    again = [EngineLP.from_instance(product, inst) for _ in range(40)]
    for r in solve_batch(product, again + kids):
        assert r["pivot_hash"] == ref["pivot_hash"] and r["obj"] == ref["obj"]
