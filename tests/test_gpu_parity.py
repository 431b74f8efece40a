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


def test_golden_fixtures_full_sizes(product):
    """tests/golden/lp_golden.json (cut from the oracle, HiGHS-anchored) incl. the full BASELINE
    sizes: configs[1] m=1024 n=2048 (13403 pivots) and configs[3] m=256 n=512."""
    import json
    import os
    with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "lp_golden.json")) as f:
        gold = json.load(f)["cases"]
    lps = [EngineLP.from_instance(product, getattr(instances, c["gen"])(*c["args"])) for c in gold]
    res = solve_batch(product, lps)  # one launch for all of them
    for c, r in zip(gold, res):
        assert r["status"] == c["status"] and r["iters"] == c["iters"], c
        assert r["pivot_hash"] == int(c["pivot_hash"]), c
        assert np.float64(r["obj"]).tobytes().hex() == c["obj_hex"], c
        if c["highs_obj"] is not None:
            assert abs(r["obj"] - c["highs_obj"]) <= 1e-9 * max(1.0, abs(c["highs_obj"])), c


def test_full_size_dense_lp_sharded_and_properties(product):
    """configs[1] at full size: sharded over 8 emulated ranks == golden hash; and
    size-independent properties of the final tableau (dictionary identity, optimality)."""
    import json
    import os
    from mvolps_b200.dist import ShardGroup
    with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "lp_golden.json")) as f:
        c = [x for x in json.load(f)["cases"] if x["args"] == [1024, 2048, 1234]][0]
    inst = instances.dense_lp(1024, 2048, 1234)
    g = ShardGroup(product, inst, 8)
    r = g.solve()
    assert r["pivot_hash"] == int(c["pivot_hash"]) and np.float64(r["obj"]).tobytes().hex() == c["obj_hex"]
    lp = EngineLP.from_instance(product, inst)
    lp.solve()
    assert np.array_equal(lp.primal(), g.primal())
    g.free()
    T, beta = lp.tableau()
    head, nb, at = lp.basis()
    x = lp.primal()
    assert np.allclose(T[1:] @ x[nb], x[head], rtol=1e-9, atol=1e-9)      # x_B = T x_N
    assert np.allclose(x[2048:], inst["A"] @ x[:2048], rtol=1e-9, atol=1e-7)  # auxiliaries = A x
    d = T[0]
    assert np.all(d[at == 0] >= -1e-9) and np.all(d[at == 1] <= 1e-9)     # dual feasible = optimal
    assert np.all(x[:2048] >= -1e-9) and np.all(x[2048:] <= inst["row_ub"] * (1 + 1e-9))


def test_queued_clones_chains_and_parent_release(oracle, product):
    """glp_copy_prob's device copy is queued and carried out by one clone_kernel in front of the next
    device operation (util.cpp:34, bs.cpp:116): chains of clones, a parent freed before the flush
    (its slab is parked), an assign over a queued source, and reads of a queued destination must all
    see the bytes a synchronous copy would have produced."""
    inst = instances.mdkp(40, 10, 31)
    a = EngineLP.from_instance(oracle, inst)
    b = EngineLP.from_instance(product, inst)
    ra, rb = a.solve(log_pivots=1), b.solve(log_pivots=1)
    _assert_same(a, b, ra, rb)
    x = a.primal()[:40]
    j = [k for k in range(40) if x[k] != np.floor(x[k])][0]
    Ta, _ = a.tableau()

    def children(lp):
        c1 = lp.clone()
        c2 = c1.clone()            # clone of a queued destination: flushes the queue first
        c3 = lp.clone()
        c1.set_bounds(j, -np.inf, np.floor(x[j]))
        c2.set_bounds(j, np.floor(x[j]) + 1.0, np.inf)
        return c1, c2, c3

    a1, a2, a3 = children(a)
    b1, b2, b3 = children(b)
    Tb3, _ = b3.tableau()          # read of a queued destination
    assert np.array_equal(Tb3, Ta)
    b.free()                       # parent released while b3's... (already flushed) and again below
    b4 = b3.clone()
    b3.free()                      # source of a queued clone released before the flush
    Tb4, _ = b4.tableau()
    assert np.array_equal(Tb4, Ta)
    b5 = b4.clone()
    b4.assign(b1)                  # overwriting the source of a queued clone: flushes first
    Tb5, _ = b5.tableau()
    assert np.array_equal(Tb5, Ta)
    for lp in (a3, b5):
        lp.set_bounds(j, -np.inf, np.floor(x[j]))
    res_a = solve_batch(oracle, [a1, a2, a3], log_pivots=1)
    res_b = solve_batch(product, [b1, b2, b5], log_pivots=1)
    for p, q, r, s in zip([a1, a2, a3], [b1, b2, b5], res_a, res_b):
        _assert_same(p, q, r, s)
    # b4 now equals b1 before its solve: same result as a1
    rb4 = b4.solve(log_pivots=1)
    assert rb4["pivot_hash"] == res_a[0]["pivot_hash"] and rb4["obj"] == res_a[0]["obj"]


def test_many_small_lps_use_narrow_ctas_and_match(oracle, product):
    """More LPs than SMs in one launch: the engine runs 128/256-thread CTAs, several per SM
    (launch_solve); results are those of the oracle, LP by LP."""
    import ctypes as C
    insts = [instances.mdkp(30 + (k % 7), 8 + (k % 5), 500 + k) for k in range(200)]
    A = [EngineLP.from_instance(oracle, i) for i in insts]
    B = [EngineLP.from_instance(product, i) for i in insts]
    ra = solve_batch(oracle, A, log_pivots=1)
    old = product.mvb_engine_option(b"no_res", 1)  # the lazy private-state loop, not the resident kernel
    try:
        rb = solve_batch(product, B, log_pivots=1)
    finally:
        product.mvb_engine_option(b"no_res", old)
    G, groups, nt = C.c_int(), C.c_int(), C.c_int()
    product.mvb_last_launch(C.byref(G), C.byref(groups), C.byref(nt))
    assert product.mvb_last_kernel_mode() == 1
    assert G.value == 1 and nt.value == 128 and groups.value > 148
    for a, b, x, y in zip(A, B, ra, rb):
        _assert_same(a, b, x, y)
    # dual warm starts through the same launch shape
    for a, b in zip(A, B):
        xa = a.primal()
        n = a.n
        frac = [k for k in range(n) if xa[k] != np.floor(xa[k])]
        if frac:
            for lp in (a, b):
                lp.set_bounds(frac[0], -np.inf, np.floor(xa[frac[0]]))
    ra = solve_batch(oracle, A, log_pivots=1)
    rb = solve_batch(product, B, log_pivots=1)
    for a, b, x, y in zip(A, B, ra, rb):
        _assert_same(a, b, x, y)


def test_hbm_streaming_batch_matches_oracle(oracle, product):
    """configs[3] shape: more 256x512 LPs than SMs in one launch, tableaux (340 MB) far beyond L2:
    256-thread CTAs, two per SM, pending depth 12 (simplex_ps_stream_kernel).  Every copy must reproduce the oracle's pivot
    sequence of its template (this regime once exposed a race on the TMA slot index)."""
    import ctypes as C
    insts = [instances.dense_lp(256, 512, 900 + k) for k in range(16)]
    ref = []
    for inst in insts:
        a = EngineLP.from_instance(oracle, inst)
        ref.append((a.solve(), a.primal()))
    tmpl = [EngineLP.from_instance(product, inst) for inst in insts]
    work = [tmpl[k % 16].clone() for k in range(160)]
    res = solve_batch(product, work)
    G, groups, nt = C.c_int(), C.c_int(), C.c_int()
    product.mvb_last_launch(C.byref(G), C.byref(groups), C.byref(nt))
    assert G.value == 1 and nt.value == 256 and groups.value == 160
    # a streaming batch runs simplex_ps_stream_kernel: 12 pending pivots per flush pass here (the
    # cp.async ring does not fit next to them with two CTAs per SM)
    opt = product.mvb_engine_option
    assert opt(b"last_stream", -1) == 1 and opt(b"last_lz_k", -1) == 12 and opt(b"last_ring", -1) == 0
    for k, (w, r) in enumerate(zip(work, res)):
        ra, xa = ref[k % 16]
        for key in ("status", "iters", "pivots", "pivot_hash", "obj"):
            assert r[key] == ra[key], (k, key)
        assert np.array_equal(w.primal(), xa)
    # A/B: depth 8 with the flush reads staged through the per-thread cp.async ring, then the same
    # depth through registers: the same bits every time (tableau bytes included)
    T12 = work[3].tableau()
    try:
        opt(b"lz_k", 8)
        work2 = [tmpl[k % 16].clone() for k in range(160)]
        res2 = solve_batch(product, work2)
        assert opt(b"last_stream", -1) == 1 and opt(b"last_lz_k", -1) == 8 and opt(b"last_ring", -1) == 8
        assert opt(b"lz_ring", 0) == 8
        work3 = [tmpl[k % 16].clone() for k in range(160)]
        res3 = solve_batch(product, work3)
        assert opt(b"last_ring", -1) == 0 and opt(b"last_lz_k", -1) == 8
    finally:
        opt(b"lz_ring", 1)
        opt(b"lz_k", 0)
    for r, r2, r3 in zip(res, res2, res3):
        assert r == r2 == r3
    for w in (work2[3], work3[3]):
        Tw = w.tableau()
        assert np.array_equal(T12[0], Tw[0]) and np.array_equal(T12[1], Tw[1])


@pytest.mark.parametrize("m,n,nlp", [(100, 300, 900), (256, 130, 700), (64, 1000, 600)])
def test_hbm_streaming_batch_ragged_shapes(oracle, product, m, n, nlp):
    """Streaming launches whose rows do not divide into ring blocks of 4 and whose CTAs hold several
    rows per pass (narrow tableaux): tails of the cp.async ring flush."""
    insts = [instances.dense_lp(m, n, 700 + k) for k in range(4)]
    ref = []
    for inst in insts:
        a = EngineLP.from_instance(oracle, inst)
        ref.append((a.solve(), a.primal(), a.tableau()))
    tmpl = [EngineLP.from_instance(product, inst) for inst in insts]
    work = [tmpl[k % 4].clone() for k in range(nlp)]
    product.mvb_engine_option(b"no_res", 1)   # these fit the resident kernel: keep them on the lazy loop
    try:
        res = solve_batch(product, work)
    finally:
        product.mvb_engine_option(b"no_res", 0)
    assert product.mvb_last_kernel_mode() == 1
    for k in (0, 1, 2, 3, nlp - 1):
        ra, xa, (Ta, ba) = ref[k % 4]
        for key in ("status", "iters", "pivots", "pivot_hash", "obj"):
            assert res[k][key] == ra[key], (k, key)
        assert np.array_equal(work[k].primal(), xa)
        Tb, bb = work[k].tableau()
        assert np.array_equal(Ta, Tb) and np.array_equal(ba, bb)


def _kat_instances():
    """The hand-checked LPs of tests/test_oracle_kats.py (Klee-Minty cube, a transportation problem
    with equality rows, a degenerate vertex, Beale's cycling example)."""
    INF = float("inf")

    def inst(A, c, lo, hi, sense="max"):
        A = np.array(A, dtype=np.float64)
        n = A.shape[1]
        return dict(A=A, c=np.array(c, float), sense=sense, row_lb=np.array(lo, float), row_ub=np.array(hi, float),
                    col_lb=np.zeros(n), col_ub=np.full(n, INF), integer=np.zeros(n, bool))

    n = 8
    km = np.zeros((n, n))
    for i in range(n):
        for j in range(i):
            km[i, j] = 2.0 * 2.0 ** (i - j)
        km[i, i] = 1.0
    yield "klee_minty_8", inst(km, [2.0 ** (n - 1 - j) for j in range(n)], [-INF] * n,
                               [5.0 ** (i + 1) for i in range(n)]), -(5.0 ** n)
    cost = np.array([[8, 6, 10, 9], [9, 12, 13, 7], [14, 9, 16, 5]], float)
    A = np.zeros((7, 12))
    for i in range(3):
        A[i, 4 * i: 4 * i + 4] = 1
    for j in range(4):
        A[3 + j, j::4] = 1
    rhs = [20, 30, 25, 10, 25, 15, 25]
    yield "transportation", inst(A, cost.ravel(), rhs, rhs, sense="min"), 585.0
    yield "degenerate_vertex", inst([[1, 0], [0, 1], [1, 1], [1, 2]], [1, 1], [-INF] * 4, [2, 2, 4, 6]), -4.0
    yield "beale", inst([[0.25, -8, -1, 9], [0.5, -12, -0.5, 3], [0, 0, 1, 0]], [0.75, -20, 0.5, -6],
                        [-INF] * 3, [0, 0, 1]), -1.25


@pytest.mark.parametrize("name,inst,obj", list(_kat_instances()), ids=[k[0] for k in _kat_instances()])
def test_known_answer_lps_on_the_gpu(oracle, product, name, inst, obj):
    """Known optima reached by the CUDA engine with the oracle's pivots, bit for bit (exponentially
    many Dantzig pivots, phase 1 on equality rows, degenerate ties, the Bland switch)."""
    a, b, ra, rb = _pair(oracle, product, inst)
    _assert_same(a, b, ra, rb)
    assert rb["status"] == 0 and abs(rb["obj"] - obj) <= 1e-9 * max(1.0, abs(obj))
