"""The reference-facing boundary on the GPU: the glpk-subset shim and the batched B&B driver of the
PRODUCT library (sm_100a kernels) against the same host code on the CPU oracle and against the
reference's own driver run recorded in tests/golden/bnb_golden.json."""
import hashlib
import json
import os

import numpy as np
import pytest

from mvolps_b200 import capi, instances
from mvolps_b200.lp import GlpProb

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SMALL3 = os.path.join(ROOT, "examples", "small3.lp")
FLAGS = {"-bs": "node_strategy", "-vs": "var_strategy", "-cm": "cut_mode"}
SAME = ("tree", "solution", "events", "event_hash", "iterations", "oids", "incumbent", "has_incumbent",
        "incumbent_oid", "lp_solves", "cached_solves", "simplex_iters", "launches", "hit_limit")


def _gold():
    with open(os.path.join(ROOT, "tests", "golden", "bnb_golden.json")) as f:
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


@pytest.mark.parametrize("case", _gold(), ids=lambda c: c["name"])
def test_bnb_product_equals_oracle_and_recorded_reference(oracle, product, case, tmp_path):
    """Tree text, solution string, event stream, counters: identical on the GPU engine, the CPU oracle
    and (tree / solution / count) the reference's bs.cpp run.  Covers dual warm starts after bound
    changes, cut rows appended to the device tableau (-cm 1), GMI rows read off it, queued clones."""
    kw = _kwargs(case["flags"])
    a = _prob(oracle, case, tmp_path).bnb(**kw)
    b = _prob(product, case, tmp_path).bnb(**kw)
    for k in SAME:
        assert a[k] == b[k], k
    tree = [ln for ln in b["tree"].splitlines() if ln.strip()]
    assert hashlib.sha256("\n".join(tree).encode()).hexdigest() == case["tree_sha256"]
    assert b["solution"].strip() == case["solution"]
    assert b["iterations"] == case["iterations"]


@pytest.mark.parametrize("cut_mode", [0, 2])
def test_bnb_larger_tree_product_equals_oracle(oracle, product, cut_mode):
    """A few hundred nodes of the configs[2] family (batches wider than the SM count reach the
    narrow-CTA launch shape) with and without textbook GMI cuts."""
    inst = instances.mdkp(60, 15, 2024)
    a = GlpProb.from_instance(oracle, inst).bnb(cut_mode=cut_mode, batch_width=256, max_iters=600)
    b = GlpProb.from_instance(product, inst).bnb(cut_mode=cut_mode, batch_width=256, max_iters=600)
    for k in SAME:
        assert a[k] == b[k], k
    assert b["iterations"] == 600 or b["hit_limit"] == 0


def test_best_first_speculation_product_equals_unspeculated_oracle(oracle, product):
    """Best-first on the GPU with the children of 32 nodes per launch == the oracle run that solves
    one node's children per launch (two clones of one parent tableau per speculated node)."""
    inst = instances.mdkp(40, 10, 7)
    a = GlpProb.from_instance(oracle, inst).bnb(node_strategy=1, no_speculation=1, max_iters=800)
    b = GlpProb.from_instance(product, inst).bnb(node_strategy=1, batch_width=32, max_iters=800)
    for k in SAME:
        if k != "launches":
            assert a[k] == b[k], k
    assert b["spec_adopted"] > 0 and b["launches"] < a["launches"] / 3


def test_shim_accessors_and_tab_rows_on_the_device(oracle, product):
    """glp_simplex / glp_get_* / glp_eval_tab_row (gmi.cpp:36) through the product shim: bit-equal to
    the oracle shim, rows read off the device tableau."""
    pa, pb = GlpProb.from_file(oracle, SMALL3), GlpProb.from_file(product, SMALL3)
    assert pa.simplex() == 0 and pb.simplex() == 0
    assert pa.status() == pb.status() == capi.GLP_OPT
    assert pa.obj_val() == pb.obj_val()
    assert np.array_equal(pa.col_prim(), pb.col_prim())
    m, n = pa.nrows, pa.ncols
    for j in range(1, n + 1):
        assert oracle.glp_get_col_stat(pa.h, j) == product.glp_get_col_stat(pb.h, j)
        if product.glp_get_col_stat(pb.h, j) == capi.GLP_BS:
            ia, va = pa.eval_tab_row(m + j)
            ib, vb = pb.eval_tab_row(m + j)
            assert list(ia) == list(ib) and np.array_equal(np.asarray(va), np.asarray(vb))
    for i in range(1, m + 1):
        assert oracle.glp_get_row_stat(pa.h, i) == product.glp_get_row_stat(pb.h, i)
    # branch like bs.cpp:274 on a copy, re-solve (dual simplex from the cloned device tableau)
    x = pa.col_prim()
    j = int(np.argmax(x != np.floor(x))) + 1
    qa, qb = pa.copy(), pb.copy()
    for L, q in ((oracle, qa), (product, qb)):
        L.glp_set_col_bnds(q.h, j, capi.GLP_UP, 0.0, float(np.floor(x[j - 1])))
        assert q.simplex() == 0
    assert qa.status() == qb.status() and qa.obj_val() == qb.obj_val()
    assert np.array_equal(qa.col_prim(), qb.col_prim())
    assert oracle.glp_get_it_cnt(qa.h) == product.glp_get_it_cnt(qb.h) > 0


REF_B200 = os.path.join(ROOT, "oracle", "_ref", "mvolps_ref_b200")


@pytest.mark.parametrize("case", _gold(), ids=lambda c: c["name"])
def test_reference_objects_on_the_product_library(product, case, tmp_path):
    """The strongest drop-in proof: the reference's UNMODIFIED 2test/bs/util/cut/gmi/message objects
    (compiled where they lie under /root/reference by oracle/Makefile `ref_b200`, linked against
    libglpk_b200.so exactly as INTEGRATION.md section 1 says) run on the GPU; tree text, solution
    string and iteration count of bs.cpp:329-345 equal the recorded run of the same objects on the
    CPU oracle (tests/golden/bnb_golden.json).  Every glp_simplex / glp_eval_tab_row /
    glp_copy_prob of the reference's serial loop goes through the sm_100a engine here."""
    import re
    import subprocess
    if not os.path.exists(REF_B200):
        pytest.fail("oracle/_ref/mvolps_ref_b200 is missing: __graft_entry__.build() makes it where "
                    "/root/reference exists, and it travels to the GPU box with the snapshot")
    if case["gen"] is None:
        path = SMALL3
    else:
        path = str(tmp_path / (case["name"] + ".lp"))
        instances.write_lp(getattr(instances, case["gen"])(*case["args"]), path)
    p = subprocess.run([REF_B200, "-f", path] + case["flags"], capture_output=True, text=True, timeout=900)
    assert p.returncode == case["returncode"], p.stderr[-2000:]
    lines = p.stdout.splitlines()
    tree = [ln for ln in lines if re.fullmatch(r"\s*-\d+( [IFB])?\s*", ln)]
    sol = [ln for ln in lines if "Solution is:" in ln]
    iters = [ln for ln in lines if "Solution found after" in ln]
    assert hashlib.sha256("\n".join(tree).encode()).hexdigest() == case["tree_sha256"]
    assert (sol[-1].strip() if sol else "") == case["solution"]
    assert (int(iters[-1].split("after")[1].split()[0]) if iters else None) == case["iterations"]


def _event_gold():
    with open(os.path.join(ROOT, "tests", "golden", "event_golden.json")) as f:
        return json.load(f)["cases"]


@pytest.mark.parametrize("case", _event_gold(), ids=lambda c: c["name"])
def test_event_stream_on_the_gpu_equals_the_references_own(product, case, tmp_path):
    """The product's event text == the messages the reference's IPCDispatch::write sends with
    --server (message.cpp:32-181; fixture captured from the unmodified reference)."""
    ev = _prob(product, case, tmp_path).bnb(**_kwargs(case["flags"]))["events"].splitlines()
    assert ev[:30] == case["head"] and len(ev) == case["events"]
    assert hashlib.sha256("\n".join(ev).encode()).hexdigest() == case["events_sha256"]


def test_node_level_vfp_and_round2_options_on_the_gpu(oracle, product):
    """-vs 3 (most fractional on the node's own LP solution), -cf, the cut selection rule and the root
    heuristic through the product: identical to the same driver on the oracle engine."""
    inst = instances.mdkp(30, 8, 11)
    for kw in (dict(var_strategy=3, max_iters=500), dict(var_strategy=3, node_strategy=1, batch_width=16, max_iters=400),
               dict(cut_mode=2, cut_freq=0.5, max_iters=300), dict(cut_mode=2, cut_select=1, max_iters=300),
               dict(root_heuristic=1, max_iters=500)):
        a = GlpProb.from_instance(oracle, inst).bnb(**kw)
        b = GlpProb.from_instance(product, inst).bnb(**kw)
        for k in SAME:
            assert a[k] == b[k], (kw, k)
        assert a["launches_aux"] == b["launches_aux"]


@pytest.mark.parametrize("kw", [dict(batch_width=64, max_iters=6000), dict(batch_width=7, max_iters=900),
                                dict(batch_width=256, root_heuristic=1, max_iters=4000), dict(batch_width=32)])
def test_pipelined_rounds_equal_the_unpipelined_loop(product, monkeypatch, kw):
    """FIFO without cuts keeps one launch in flight (mvb_solve_batch_begin / _end): the children of a
    round are solved while the host consumes the next round.  Everything the run reports -- tree, ids,
    events, counters, incumbent -- equals the loop that waits for every launch (MVOLPS_BNB_NO_PIPELINE)."""
    inst = instances.mdkp(30, 8, 4) if "max_iters" not in kw else instances.mdkp(200, 50, 2024)
    a = GlpProb.from_instance(product, inst).bnb(**kw)
    monkeypatch.setenv("MVOLPS_BNB_NO_PIPELINE", "1")
    b = GlpProb.from_instance(product, inst).bnb(**kw)
    monkeypatch.delenv("MVOLPS_BNB_NO_PIPELINE")
    for k in SAME:
        assert a[k] == b[k], k
    if a["x"] is not None:
        assert np.array_equal(a["x"], b["x"])


def test_solve_batch_in_two_halves(oracle, product):
    """mvb_solve_batch_begin / _end through the C ABI: other LPs are created, cloned, changed and
    freed while the launch is in flight; results equal the one-call solve and the oracle."""
    import ctypes as C
    from mvolps_b200 import capi
    from mvolps_b200.lp import EngineLP, solve_batch
    # (a batch of resident-kernel LPs: one launch, left in flight; with an LP that needs phase 1 the
    # batch takes several launches and is solved inside begin)
    mixed = [instances.mdkp(40, 10, 30 + k) for k in range(24)]
    for extra in ([], [instances.general_lp(20, 30, 12)]):
        _two_halves(oracle, product, mixed + extra)


def _two_halves(oracle, product, insts):
    import ctypes as C
    from mvolps_b200.lp import EngineLP, solve_batch
    ref = [EngineLP.from_instance(oracle, i).solve() for i in insts]
    lps = [EngineLP.from_instance(product, i) for i in insts]
    n = len(lps)
    arr = (C.c_void_p * n)(*[lp.h for lp in lps])
    opts = capi.Opts(0, 0, 0, 0)
    assert product.mvb_solve_batch_begin(arr, n, C.byref(opts)) == 0
    assert product.mvb_solve_batch_begin(arr, n, C.byref(opts)) != 0  # one batch at a time
    other = EngineLP.from_instance(product, instances.mdkp(40, 10, 77))
    twin = other.clone()
    twin.set_bounds(3, 0.0, 1.0)
    other.free()
    if len(insts) % 2 == 0:   # a one-call solve in between: the batch in flight comes home first, end hands it out
        rt = twin.solve()
    res = (capi.Result * n)()
    assert product.mvb_solve_batch_end(res) == 0
    assert product.mvb_solve_batch_end(res) != 0
    if len(insts) % 2:
        rt = twin.solve()
    for r, want, lp in zip(res, ref, lps):
        for f in ("status", "iters", "pivots", "pivot_hash", "obj", "iters_primal1", "iters_dual"):
            assert getattr(r, f) == want[f], f
    o = EngineLP.from_instance(oracle, instances.mdkp(40, 10, 77))
    o.set_bounds(3, 0.0, 1.0)
    ro = o.solve()
    assert rt["pivot_hash"] == ro["pivot_hash"] and rt["obj"] == ro["obj"]
    # and the batch again in one call: the same results
    again = [EngineLP.from_instance(product, i) for i in insts]
    for r, r2 in zip(res, solve_batch(product, again)):
        assert r.pivot_hash == r2["pivot_hash"] and r.obj == r2["obj"]
