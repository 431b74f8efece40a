"""CPU tests of the host side: the glpk-subset shim (include/glpk.h), the LP/MPS
readers, and the batched B&B driver (include/mvolps_bnb.h) against (i) the
reference's own unmodified driver, recorded in tests/golden/bnb_golden.json,
(ii) that driver run live where oracle/_ref exists, (iii) SciPy/HiGHS MILP."""
import hashlib
import json
import os
import re
import subprocess

import numpy as np
import pytest

from mvolps_b200 import capi, instances
from mvolps_b200.lp import GlpProb

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")
SMALL3 = os.path.join(ROOT, "examples", "small3.lp")
FLAGS = {"-bs": "node_strategy", "-vs": "var_strategy", "-cm": "cut_mode"}


def _gold():
    with open(os.path.join(GOLDEN, "bnb_golden.json")) as f:
        return json.load(f)["cases"]


def _kwargs(flags):
    return {FLAGS[flags[i]]: int(flags[i + 1]) for i in range(0, len(flags), 2)}


def _prob(lib, case, tmp_path):
    if case["gen"] is None:
        return GlpProb.from_file(lib, SMALL3)
    inst = getattr(instances, case["gen"])(*case["args"])
    path = str(tmp_path / (case["name"] + ".lp"))
    instances.write_lp(inst, path)
    return GlpProb.from_file(lib, path)


# ---------------------------------------------------------------- shim
def test_read_lp_small3(oracle):
    p = GlpProb.from_file(oracle, SMALL3)
    L, h = oracle, p.h
    assert (p.nrows, p.ncols, L.glp_get_num_int(h)) == (4, 3, 3)
    assert L.glp_get_obj_dir(h) == capi.GLP_MAX
    assert [L.glp_get_obj_coef(h, j) for j in range(4)] == [0.0, 1.0, 1.0, 1.0]
    assert [L.glp_get_row_type(h, i) for i in range(1, 5)] == [capi.GLP_UP] * 4
    assert [L.glp_get_row_ub(h, i) for i in range(1, 5)] == [4877.0, 6750.0, 4587.0, 6911.0]
    assert [L.glp_get_col_type(h, j) for j in range(1, 4)] == [capi.GLP_DB] * 3
    assert [L.glp_get_col_kind(h, j) for j in range(1, 4)] == [capi.GLP_IV] * 3
    ind = np.zeros(4, np.int32)
    val = np.zeros(4)
    assert L.glp_get_mat_row(h, 3, capi.iptr(ind), capi.dptr(val)) == 3
    assert dict(zip(ind[1:].tolist(), val[1:].tolist())) == {1: 2.0, 2: 9.0, 3: 4.0}


def test_simplex_status_and_accessors(oracle):
    p = GlpProb.from_file(oracle, SMALL3)
    assert p.status() == capi.GLP_UNDEF
    assert p.simplex() == 0 and p.status() == capi.GLP_OPT
    x = p.col_prim()
    assert abs(p.obj_val() - x.sum()) < 1e-9 and p.obj_val() > 1202.0
    L, h = oracle, p.h
    stats = [L.glp_get_col_stat(h, j) for j in range(1, 4)] + [L.glp_get_row_stat(h, i) for i in range(1, 5)]
    assert stats.count(capi.GLP_BS) == 4  # m basic variables
    # glp_eval_tab_row: x_B = sum alpha_k xN_k reproduces the basic variable's value
    for j in range(1, 4):
        if L.glp_get_col_stat(h, j) != capi.GLP_BS:
            continue
        ind, val = p.eval_tab_row(4 + j)
        tot = 0.0
        for k, a in zip(ind, val):
            assert L.glp_get_row_stat(h, k) != capi.GLP_BS if k <= 4 else L.glp_get_col_stat(h, k - 4) != capi.GLP_BS
            tot += a * (L.glp_get_row_prim(h, k) if k <= 4 else L.glp_get_col_prim(h, k - 4))
        assert abs(tot - x[j - 1]) <= 1e-9 * max(1.0, abs(x[j - 1]))
        assert list(ind) == sorted(ind)  # documented ordering: ascending variable index


def test_copy_prob_carries_basis_and_bound_replacement(oracle):
    """bs.cpp:114-117 (copy, re-solve: 0 iterations) and bs.cpp:274/282 (type replaced)."""
    L = oracle
    p = GlpProb.from_file(L, SMALL3)
    p.simplex()
    q = p.copy()
    assert q.simplex() == 0 and L.glp_get_it_cnt(q.h) == 0 and q.obj_val() == p.obj_val()
    x = p.col_prim()
    j = int(np.argmax(x != np.floor(x))) + 1
    L.glp_set_col_bnds(q.h, j, capi.GLP_UP, 0.0, float(np.floor(x[j - 1])))
    assert L.glp_get_col_type(q.h, j) == capi.GLP_UP
    assert L.glp_get_col_lb(q.h, j) == -np.finfo(np.float64).max  # lower bound dropped
    q.simplex()
    assert q.status() == capi.GLP_OPT and q.obj_val() <= p.obj_val()
    assert L.glp_get_col_prim(q.h, j) <= np.floor(x[j - 1]) + 1e-9
    L.glp_erase_prob(q.h)
    assert q.nrows == 0 and q.ncols == 0


def test_add_cut_row_like_cutpool(oracle):
    """cut.cpp:23,40-43: glp_add_rows(1) + glp_set_row_bnds(GLP_LO) + glp_set_mat_row with ind[0] junk."""
    L = oracle
    p = GlpProb.from_file(L, SMALL3)
    p.simplex()
    z0 = p.obj_val()
    r = L.glp_add_rows(p.h, 1)
    assert r == 5
    ind = np.array([0, 1, 2, 3], np.int32)
    val = np.array([123.0, -1.0, -1.0, -1.0])
    L.glp_set_mat_row(p.h, r, 3, capi.iptr(ind), capi.dptr(val))
    L.glp_set_row_bnds(p.h, r, capi.GLP_LO, -1200.0, 0.0)  # x1+x2+x3 <= 1200
    p.simplex()
    assert p.status() == capi.GLP_OPT and abs(p.obj_val() - 1200.0) < 1e-9 < z0 - 1200.0


def test_read_mps_roundtrip(oracle, tmp_path):
    mps = tmp_path / "t.mps"
    mps.write_text(
        "NAME          T\nROWS\n N  COST\n L  C1\n G  C2\n E  C3\nCOLUMNS\n"
        "    MARKER    'MARKER'  'INTORG'\n"
        "    X  COST  1.0  C1  1.0\n    X  C2  1.0\n"
        "    MARKER    'MARKER'  'INTEND'\n"
        "    Y  COST  2.0  C1  1.0\n    Y  C3  1.0\n"
        "RHS\n    RHS  C1  4.0  C2  1.0\n    RHS  C3  1.5\n"
        "BOUNDS\n UP BND  X  3.0\nENDATA\n")
    p = GlpProb.from_file(oracle, str(mps))
    L, h = oracle, p.h
    assert (p.nrows, p.ncols) == (3, 2)
    assert [L.glp_get_row_type(h, i) for i in (1, 2, 3)] == [capi.GLP_UP, capi.GLP_LO, capi.GLP_FX]
    assert L.glp_get_col_kind(h, 1) == capi.GLP_IV and L.glp_get_col_kind(h, 2) == capi.GLP_CV
    assert L.glp_get_obj_dir(h) == capi.GLP_MIN
    p.simplex()
    assert p.status() == capi.GLP_OPT and abs(p.obj_val() - 4.0) < 1e-12  # x=1, y=1.5


def test_read_errors(oracle, tmp_path):
    bad = tmp_path / "bad.lp"
    bad.write_text("Maximize\n obj: x +\nSubject To\n c1: x <= \nEnd\n")
    with pytest.raises(ValueError):
        GlpProb.from_file(oracle, str(bad))
    with pytest.raises(ValueError):
        GlpProb.from_file(oracle, str(tmp_path / "missing.lp"))


def test_bv_kind_reported_for_01_integer(oracle, tmp_path):
    lp = tmp_path / "b.lp"
    lp.write_text("Maximize\n obj: x + y\nSubject To\n c: x + y <= 1.5\nBinary\n x\nGeneral\n y\nEnd\n")
    p = GlpProb.from_file(oracle, str(lp))
    assert oracle.glp_get_col_kind(p.h, 1) == capi.GLP_BV and oracle.glp_get_col_kind(p.h, 2) == capi.GLP_IV


# ---------------------------------------------------------------- B&B driver
@pytest.mark.parametrize("case", _gold(), ids=lambda c: c["name"])
def test_bnb_matches_recorded_reference_driver(oracle, case, tmp_path):
    """Batched driver == the reference's bs.cpp run (recorded): tree, solution string, count."""
    out = _prob(oracle, case, tmp_path).bnb(**_kwargs(case["flags"]))
    tree = [ln for ln in out["tree"].splitlines() if ln.strip()]
    assert len(tree) == case["tree_lines"]
    assert tree[: len(case["tree"])] == case["tree"]
    assert hashlib.sha256("\n".join(tree).encode()).hexdigest() == case["tree_sha256"]
    assert out["solution"].strip() == case["solution"]
    assert out["iterations"] == case["iterations"]


@pytest.mark.parametrize("width", [1, 2, 7, 64])
def test_bnb_batch_width_invariance(oracle, width, tmp_path):
    """FIFO batches of any width give the serial run's oids/prunes/incumbents."""
    case = [c for c in _gold() if c["name"] == "mdkp_12x4_s1"][0]
    base = _prob(oracle, case, tmp_path).bnb(batch_width=1)
    out = _prob(oracle, case, tmp_path).bnb(batch_width=width)
    for k in ("tree", "solution", "events", "event_hash", "iterations", "oids", "incumbent"):
        assert out[k] == base[k], k
    if width > 1:
        assert out["launches"] < base["launches"]


@pytest.mark.parametrize("args,vs,width", [((12, 4, 1), 0, 64), ((15, 5, 2), 2, 8), ((20, 5, 3), 1, 2)])
def test_best_first_speculation_is_sequence_identical(oracle, args, vs, width):
    """-bs 1 (util.cpp:170-187 pops one node per iteration): the children of the next `width` nodes in
    pop order are solved ahead of their turn in one launch; the serial order adopts them, so tree,
    events, ids, incumbent AND the solve statistics are those of the one-node-per-launch run."""
    inst = instances.mdkp(*args)
    base = GlpProb.from_instance(oracle, inst).bnb(node_strategy=1, var_strategy=vs, no_speculation=1)
    out = GlpProb.from_instance(oracle, inst).bnb(node_strategy=1, var_strategy=vs, batch_width=width)
    for k in ("tree", "solution", "events", "event_hash", "iterations", "oids", "incumbent", "incumbent_oid",
              "lp_solves", "cached_solves", "simplex_iters"):
        assert out[k] == base[k], k
    assert base["spec_nodes"] == 0 and out["spec_nodes"] >= out["spec_adopted"] > 0
    assert out["launches"] < base["launches"]
    # with cuts the children depend on the cut pool at branching time: no speculation
    cut = GlpProb.from_instance(oracle, inst).bnb(node_strategy=1, cut_mode=2, batch_width=width, max_iters=40)
    assert cut["spec_nodes"] == 0


def test_node_level_vfp_strategy(oracle, tmp_path):
    """-vs 3 (not in the reference): util.cpp:197-213's rule on the node's own LP solution instead of
    the never-solved root object.  Same optimum on small3 in fewer nodes; -vs 1 stays the reference's
    behaviour (== -vs 0); speculation under best-first stays sequence-identical."""
    case = [c for c in _gold() if c["name"] == "small3"][0]
    vo = _prob(oracle, case, tmp_path).bnb(var_strategy=0)
    vfp_ref = _prob(oracle, case, tmp_path).bnb(var_strategy=1)
    vfp = _prob(oracle, case, tmp_path).bnb(var_strategy=3)
    assert vfp_ref["tree"] == vo["tree"]
    assert vfp["incumbent"] == vo["incumbent"] and abs(vfp["incumbent"] - case["highs_ilp"]) <= 1e-9 * abs(case["highs_ilp"])
    assert vfp["iterations"] < vo["iterations"]
    a = _prob(oracle, case, tmp_path).bnb(var_strategy=3, node_strategy=1, no_speculation=1)
    b = _prob(oracle, case, tmp_path).bnb(var_strategy=3, node_strategy=1, batch_width=8)
    assert a["tree"] == b["tree"] and a["events"] == b["events"] and a["simplex_iters"] == b["simplex_iters"]


@pytest.mark.parametrize("name", ["small3", "mdkp_12x4_s1", "mdkp_12x4_s4_cm1"])
def test_bnb_optimum_vs_highs(oracle, name, tmp_path):
    case = [c for c in _gold() if c["name"] == name][0]
    out = _prob(oracle, case, tmp_path).bnb(**_kwargs(case["flags"]))
    assert out["has_incumbent"]
    assert abs(out["incumbent"] - case["highs_ilp"]) <= 1e-9 * abs(case["highs_ilp"])
    assert np.array_equal(out["x"], np.round(out["x"]))


def test_bnb_textbook_gmi_is_valid(oracle, tmp_path):
    """cut_mode 2 (textbook GMI off the tableau row) never cuts off the ILP optimum."""
    for name in ("small3", "mdkp_12x4_s1", "mdkp_12x4_s4_cm1"):
        case = [c for c in _gold() if c["name"] == name][0]
        out = _prob(oracle, case, tmp_path).bnb(cut_mode=2)
        assert out["has_incumbent"], name
        assert abs(out["incumbent"] - case["highs_ilp"]) <= 1e-9 * abs(case["highs_ilp"]), name


def test_bnb_node_budget_and_events(oracle, tmp_path):
    case = [c for c in _gold() if c["name"] == "mdkp_15x5_s2"][0]
    out = _prob(oracle, case, tmp_path).bnb(max_iters=50)
    assert out["hit_limit"] == 1 and out["iterations"] == 50
    full = _prob(oracle, case, tmp_path).bnb()
    ev, evf = out["events"].splitlines(), full["events"].splitlines()
    assert ev and evf[: len(ev)] == ev  # the budgeted run is a prefix of the full run
    assert re.match(r"^\w+ \d+ -?\d+ ", ev[0])


def test_timed_event_stream(oracle, tmp_path):
    """timed_events: the reference's field order (message.h:141-226: timeSpan first), same events."""
    case = [c for c in _gold() if c["name"] == "small3"][0]
    plain = _prob(oracle, case, tmp_path).bnb()
    timed = _prob(oracle, case, tmp_path).bnb(timed_events=1)
    a, b = plain["events"].splitlines(), timed["events"].splitlines()
    assert len(a) == len(b) > 0 and plain["event_hash"] == timed["event_hash"]
    ts = [float(l.split(" ", 1)[0]) for l in b]
    assert [l.split(" ", 1)[1] for l in b] == a and ts == sorted(ts) and ts[0] >= 0.0


@pytest.mark.parametrize("kw", [dict(), dict(cut_mode=2, max_iters=400), dict(node_strategy=1, var_strategy=2, batch_width=8),
                                dict(root_heuristic=1, max_iters=300)])
def test_event_hash_without_the_event_text(oracle, tmp_path, kw):
    """no_event_text (what the CLI does without -ev and bench.py inside its timed region): no line is
    formatted or kept, the event hash -- (event, oid, pid) of every event in order, then (oid, prune
    tag) of every node -- and everything else are those of the run that keeps the text."""
    case = [c for c in _gold() if c["name"] == "mdkp_15x5_s2"][0]
    full = _prob(oracle, case, tmp_path).bnb(**kw)
    bare = _prob(oracle, case, tmp_path).bnb(no_event_text=1, **kw)
    assert full["events"] and bare["events"] == ""
    for key in ("event_hash", "tree", "solution", "iterations", "oids", "incumbent", "lp_solves", "simplex_iters"):
        assert full[key] == bare[key], key


def test_bnb_live_reference_driver(oracle, ref_bin, tmp_path):
    """Where /root/reference was compiled (oracle/_ref): run it now and compare stdout."""
    inst = instances.mdkp(11, 4, 9)
    path = str(tmp_path / "live.lp")
    instances.write_lp(inst, path)
    p = subprocess.run([ref_bin, "-f", path], capture_output=True, text=True, timeout=300)
    assert p.returncode == 0
    tree = [ln for ln in p.stdout.splitlines() if re.fullmatch(r"\s*-\d+( [IFB])?\s*", ln)]
    sol = [ln.strip() for ln in p.stdout.splitlines() if "Solution is:" in ln]
    out = GlpProb.from_file(oracle, path).bnb()
    assert [ln for ln in out["tree"].splitlines() if ln.strip()] == tree
    assert out["solution"].strip() == (sol[-1] if sol else "")


def test_bound_setters_are_forwarded_and_unchanged_lps_are_answered_from_the_held_solution(oracle):
    """The shim forwards only the bounds its setters logged (glp_prob::dirty) and answers a solve of
    an LP that is unchanged since its optimal solve without touching the engine (bs.cpp:117 re-solves
    such copies).  Both must be invisible: same values as a cold solve of the same problem."""
    L = oracle
    inst = instances.mdkp(30, 8, 77)
    p = GlpProb.from_instance(L, inst)
    assert p.simplex() == 0 and p.status() == capi.GLP_OPT
    x0, z0, it0 = p.col_prim(), p.obj_val(), L.glp_get_it_cnt(p.h)
    assert it0 > 0
    # unchanged: zero iterations, same values (also on a copy)
    assert p.simplex() == 0 and L.glp_get_it_cnt(p.h) == 0
    assert np.array_equal(p.col_prim(), x0) and p.obj_val() == z0
    q = p.copy()
    assert q.simplex() == 0 and L.glp_get_it_cnt(q.h) == 0 and np.array_equal(q.col_prim(), x0)
    # setting a bound to the value it already has changes nothing
    j = int(np.argmax(x0 != np.floor(x0))) + 1
    L.glp_set_col_bnds(q.h, j, L.glp_get_col_type(q.h, j), L.glp_get_col_lb(q.h, j), L.glp_get_col_ub(q.h, j))
    assert q.simplex() == 0 and L.glp_get_it_cnt(q.h) == 0 and q.obj_val() == z0
    # a real change is forwarded: warm dual re-solve == cold solve of the changed problem
    for h in (q.h,):
        L.glp_set_col_bnds(h, j, capi.GLP_DB, 0.0, float(np.floor(x0[j - 1])))
        L.glp_set_row_bnds(h, 1, capi.GLP_UP, 0.0, float(np.floor(0.9 * L.glp_get_row_ub(h, 1))))
    assert q.simplex() == 0 and q.status() == capi.GLP_OPT and L.glp_get_it_cnt(q.h) > 0
    inst2 = instances.mdkp(30, 8, 77)
    cold = GlpProb.from_instance(L, inst2)
    L.glp_set_col_bnds(cold.h, j, capi.GLP_DB, 0.0, float(np.floor(x0[j - 1])))
    L.glp_set_row_bnds(cold.h, 1, capi.GLP_UP, 0.0, float(np.floor(0.9 * L.glp_get_row_ub(cold.h, 1))))
    assert cold.simplex() == 0 and cold.status() == capi.GLP_OPT
    assert abs(cold.obj_val() - q.obj_val()) <= 1e-9 * max(1.0, abs(cold.obj_val()))
    assert q.obj_val() < z0
    # the original is untouched by what happened to its copy
    assert p.simplex() == 0 and p.obj_val() == z0


def test_bnb_counts_cached_resolves(oracle, tmp_path):
    """Every popped non-root node was solved to optimality as a child: its re-solve (bs.cpp:117) is
    answered from the held solution and counted, results are those of the recorded reference run."""
    case = _gold()[0]
    p = _prob(oracle, case, tmp_path)
    out = p.bnb(**_kwargs(case["flags"]))
    assert out["lp_solves"] >= out["cached_solves"] > 0
    assert out["iterations"] == case["iterations"]


# ---------------------------------------------------------------- round-2 driver options
def test_cut_frequency_and_selection(oracle, tmp_path):
    """-cf (2test.cpp:135-149 parses it, util.cpp:259-261 stores it, nothing reads it): here the
    share of branched nodes that run the cut step, decided by a hash of the oid -> reproducible;
    cut_select 1 = the rule cut.cpp:16-19 leaves as a TODO (most violated cut of the node's own).
    Textbook GMI cuts never cut off the optimum whatever the frequency / selection."""
    for name in ("small3", "mdkp_12x4_s1", "mdkp_12x4_s4_cm1"):
        case = [c for c in _gold() if c["name"] == name][0]
        base = _prob(oracle, case, tmp_path).bnb(cut_mode=0)
        full = _prob(oracle, case, tmp_path).bnb(cut_mode=2)
        assert _prob(oracle, case, tmp_path).bnb(cut_mode=2, cut_freq=1.0)["event_hash"] == full["event_hash"]
        none = _prob(oracle, case, tmp_path).bnb(cut_mode=2, cut_freq=0.0)
        assert none["tree"] == base["tree"] and none["launches_aux"] == 0  # chance 0: the cut step never runs
        half = _prob(oracle, case, tmp_path).bnb(cut_mode=2, cut_freq=0.5)
        half2 = _prob(oracle, case, tmp_path).bnb(cut_mode=2, cut_freq=0.5)
        assert half["events"] == half2["events"]  # reproducible
        sel = _prob(oracle, case, tmp_path).bnb(cut_mode=2, cut_select=1)
        for out in (full, half, sel):
            assert out["has_incumbent"] and abs(out["incumbent"] - base["incumbent"]) <= 1e-9 * abs(base["incumbent"]), name


def test_gmi_rows_come_from_the_node_itself(oracle, tmp_path):
    """A node whose columns produce no textbook cut gets no row (it used to inherit the last cut of
    another subtree, valid only under that subtree's bounds).  With every row valid in the subtree
    it is added to, a tree that closes never loses the true ILP optimum (HiGHS); it may end above it,
    at a point with a negative entry, because a down child loses its lower bound (bs.cpp:274)."""
    from scipy.optimize import Bounds, LinearConstraint, milp
    closed = 0
    for args in ((12, 4, 1), (12, 4, 4), (11, 4, 9), (14, 5, 3), (16, 4, 6), (10, 3, 5)):
        inst = instances.mdkp(*args)
        cut = GlpProb.from_instance(oracle, inst).bnb(cut_mode=2, max_iters=20000)
        if cut["hit_limit"]:
            continue
        closed += 1
        ref = milp(-inst["c"], constraints=LinearConstraint(inst["A"], -np.inf, inst["row_ub"]),
                   integrality=np.ones(args[0]), bounds=Bounds(0, 3))
        assert cut["has_incumbent"] and cut["incumbent"] >= -ref.fun - 1e-9, args
        if np.all(cut["x"] >= 0):
            assert abs(cut["incumbent"] + ref.fun) <= 1e-9 * abs(ref.fun), args
    assert closed >= 3


def test_root_heuristic_gives_first_incumbent(oracle):
    inst = instances.mdkp(40, 10, 7)
    plain = GlpProb.from_instance(oracle, inst).bnb(max_iters=300)
    heur = GlpProb.from_instance(oracle, inst).bnb(max_iters=300, root_heuristic=1)
    assert heur["has_incumbent"] and heur["events"].splitlines()[1].startswith("heuristic 1 ")
    x = heur["x"] if not plain["has_incumbent"] or heur["incumbent"] >= plain["incumbent"] else None
    if x is not None:
        assert np.all(inst["A"] @ x <= inst["row_ub"] + 1e-9) and np.array_equal(x, np.round(x))
        assert abs(inst["c"] @ x - heur["incumbent"]) < 1e-9
    # with an incumbent from the start the bound test (bs.cpp:210) can only prune more
    assert heur["oids"] <= plain["oids"]


def test_batched_row_gather_equals_single_calls(oracle):
    """mvb_lp_get_tab_rows_batch (all cut candidates of a round in one engine call) == one
    mvb_lp_get_tab_rows per LP."""
    import ctypes as C
    from mvolps_b200.lp import EngineLP
    lps = [EngineLP.from_instance(oracle, instances.dense_lp(12 + k, 20, k)) for k in range(4)]
    for lp in lps:
        lp.solve()
    rows = [[0, 3, 5], [], [12, 1], [7]]
    flat = np.array(sum(rows, []), np.int32)
    cnt = np.array([len(r) for r in rows], np.int32)
    out = np.zeros((len(flat), 20))
    arr = (C.c_void_p * 4)(*[lp.h for lp in lps])
    assert oracle.mvb_lp_get_tab_rows_batch(arr, 4, capi.iptr(flat), capi.iptr(cnt), capi.dptr(out)) == 0
    k = 0
    for lp, r in zip(lps, rows):
        if r:
            assert np.array_equal(lp.tab_rows(r), out[k:k + len(r)])
        k += len(r)


def test_integer_snap_and_what_happens_without_it(oracle, tmp_path):
    """The shim reports values within 1e-9 of an integer as that integer, in front of the
    reference's exact test trunc(x) != x (util.cpp:435-451).  max x, 0.1 x <= 0.3, x integer: the
    vertex is 0.3/0.1 = 2.9999999999999996 in fp64.  With the snap the root's violated list is
    empty (integral root: bs.cpp:134-149 stops there); without it x is on the violated list, the
    root branches on it and the tree has three nodes."""
    lp = tmp_path / "snap.lp"
    lp.write_text("Maximize\n obj: x\nSubject To\n c1: 0.1 x <= 0.3\nBounds\n 0 <= x <= 10\nGeneral\n x\nEnd\n")
    assert 0.3 / 0.1 != 3.0
    old = oracle.mvb_glp_snap_tol(-1.0)
    assert old == 1e-9
    try:
        p = GlpProb.from_file(oracle, str(lp))
        p.simplex()
        assert p.col_prim()[0] == 3.0
        a = GlpProb.from_file(oracle, str(lp)).bnb()
        assert a["oids"] == 1 and a["tree"].strip() == "-1 I"
        oracle.mvb_glp_snap_tol(0.0)
        p = GlpProb.from_file(oracle, str(lp))
        p.simplex()
        assert p.col_prim()[0] == 0.3 / 0.1  # the raw tableau value
        b = GlpProb.from_file(oracle, str(lp)).bnb()
        assert b["oids"] == 3 and "branched 1 " in b["events"]  # x was on the violated list
        assert [ln.strip() for ln in b["tree"].splitlines() if ln.strip()] == ["-1", "-2 I", "-3 I"]
    finally:
        oracle.mvb_glp_snap_tol(old)


def test_rowless_problem_and_objective_constant(oracle):
    """GLPK solves bound-only problems (no rows); a change of the constant term moves the held
    objective value without a re-solve."""
    L = oracle
    p = GlpProb(L)
    L.glp_set_obj_dir(p.h, capi.GLP_MAX)
    L.glp_add_cols(p.h, 3)
    for j, (t, lo, hi, c) in enumerate([(capi.GLP_DB, 0.0, 4.0, 2.0), (capi.GLP_DB, -1.0, 5.0, -3.0),
                                        (capi.GLP_FX, 7.0, 7.0, 1.0)], 1):
        L.glp_set_col_bnds(p.h, j, t, lo, hi)
        L.glp_set_obj_coef(p.h, j, c)
    assert p.simplex() == 0 and p.status() == capi.GLP_OPT
    assert list(p.col_prim()) == [4.0, -1.0, 7.0] and p.obj_val() == 8.0 + 3.0 + 7.0
    L.glp_set_obj_coef(p.h, 0, 100.0)
    assert p.obj_val() == 118.0
    L.glp_set_col_bnds(p.h, 1, capi.GLP_LO, 0.0, 0.0)
    assert p.simplex() == 0 and p.status() == capi.GLP_UNBND
    q = GlpProb.from_file(L, SMALL3)
    q.simplex()
    z = q.obj_val()
    L.glp_set_obj_coef(q.h, 0, 5.0)
    assert abs(q.obj_val() - (z + 5.0)) < 1e-9
    assert q.simplex() == 0 and L.glp_get_it_cnt(q.h) == 0 and abs(q.obj_val() - (z + 5.0)) < 1e-9


def _event_gold():
    with open(os.path.join(GOLDEN, "event_golden.json")) as f:
        return json.load(f)["cases"]


@pytest.mark.parametrize("case", _event_gold(), ids=lambda c: c["name"])
def test_event_stream_equals_the_references_own(oracle, case, tmp_path):
    """mvb_bnb_event_text against what the reference's IPCDispatch::write (message.cpp:32-181) sends
    when it runs with --server: same records (message.h:141-226), same field order and formatting,
    event by event (the reference's timeSpan field aside).  Fixture: tests/golden/event_golden.json,
    captured from the unmodified reference by the zmqpp stub."""
    out = _prob(oracle, case, tmp_path).bnb(**_kwargs(case["flags"]))
    ev = out["events"].splitlines()
    assert ev[:30] == case["head"]
    assert len(ev) == case["events"]
    assert hashlib.sha256("\n".join(ev).encode()).hexdigest() == case["events_sha256"]
