#!/usr/bin/env python
"""Regenerates the fixtures in this directory.  Run from the repo root:
    python tests/golden/make_golden.py

lp_golden.json  : for seeded instances, the oracle's status / iteration count /
                  pivot hash / objective bits and the HiGHS objective (SciPy) as
                  a solver-independent anchor.
bnb_golden.json : for small ILPs, the printed tree / solution string / event
                  hash of the REFERENCE's unmodified driver (bs.cpp, util.cpp,
                  gmi.cpp, cut.cpp compiled from /root/reference by
                  oracle/Makefile `ref`) running on the oracle shim, and the
                  SciPy/HiGHS MILP optimum.  /root/reference is needed only
                  here, not when the tests run.
"""
import hashlib
import json
import os
import re
import subprocess
import sys
import tempfile

import numpy as np
from scipy.optimize import Bounds, LinearConstraint, linprog, milp

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from mvolps_b200 import capi, instances  # noqa: E402
from mvolps_b200.lp import EngineLP  # noqa: E402

ORACLE = os.path.join(ROOT, "oracle", "_build", "libglpk_oracle.so")
REF_BIN = os.path.join(ROOT, "oracle", "_ref", "mvolps_ref_oracle")

LP_CASES = [
    ("dense_lp", [4, 6, 1]), ("dense_lp", [8, 12, 2]), ("dense_lp", [33, 70, 3]),
    ("dense_lp", [64, 128, 4]), ("dense_lp", [128, 256, 5]), ("dense_lp", [100, 37, 6]),
    ("dense_lp", [257, 300, 7]), ("general_lp", [6, 8, 11]), ("general_lp", [20, 30, 12]),
    ("general_lp", [50, 80, 13]), ("general_lp", [100, 150, 14]), ("general_lp", [40, 200, 15]),
    ("mdkp", [40, 10, 21]), ("mdkp", [40, 10, 22]), ("mdkp", [200, 50, 2024]),
    # full BASELINE sizes: configs[3] node LP and configs[1] (about 15 s on one core)
    ("dense_lp", [256, 512, 0]), ("dense_lp", [256, 512, 1]), ("dense_lp", [1024, 2048, 1234]),
]

BNB_CASES = [  # (name, generator, args, reference CLI flags)
    ("small3", None, None, []),
    ("small3_cm1", None, None, ["-cm", "1"]),
    ("small3_best", None, None, ["-bs", "1"]),
    ("small3_vgo", None, None, ["-vs", "2"]),
    ("mdkp_12x4_s1", "mdkp", [12, 4, 1], []),
    ("mdkp_15x5_s2", "mdkp", [15, 5, 2], []),
    ("mdkp_10x3_s3_best", "mdkp", [10, 3, 3], ["-bs", "1"]),
    ("mdkp_12x4_s4_cm1", "mdkp", [12, 4, 4], ["-cm", "1"]),
    ("mdkp_15x5_s2_best", "mdkp", [15, 5, 2], ["-bs", "1"]),
    ("mdkp_12x4_s1_vfp", "mdkp", [12, 4, 1], ["-vs", "1"]),
    ("mdkp_20x5_s3_best_vgo", "mdkp", [20, 5, 3], ["-bs", "1", "-vs", "2"]),
]


def highs_lp(inst):
    A = inst["A"]
    m, n = A.shape
    c = -inst["c"] if inst["sense"] == "max" else inst["c"]
    cons = LinearConstraint(A, inst["row_lb"], inst["row_ub"])
    r = milp(c, constraints=cons, bounds=Bounds(inst["col_lb"], inst["col_ub"]),
             integrality=np.zeros(n))
    return float(r.fun) if r.status == 0 else None


def highs_ilp(inst):
    A = inst["A"]
    c = -inst["c"] if inst["sense"] == "max" else inst["c"]
    cons = LinearConstraint(A, inst["row_lb"], inst["row_ub"])
    r = milp(c, constraints=cons, bounds=Bounds(inst["col_lb"], inst["col_ub"]),
             integrality=inst["integer"].astype(int))
    return (float(-r.fun) if inst["sense"] == "max" else float(r.fun)) if r.status == 0 else None


def small3_inst():
    A = np.array([[3, 5, 7], [7, 2, 5], [2, 9, 4], [5, 4, 11]], float)
    return dict(A=A, c=np.ones(3), sense="max", row_lb=np.full(4, -np.inf),
                row_ub=np.array([4877, 6750, 4587, 6911], float), col_lb=np.zeros(3),
                col_ub=np.full(3, 1000.0), integer=np.ones(3, bool))


def reference_events(path, flags):
    """What the reference's IPCDispatch::write (message.cpp:32-181) sends to its visualiser when
    started with --server: the zmqpp stub (oracle/stubs/zmqpp/zmqpp.hpp) appends every message to
    $MVOLPS_ZMQ_DUMP.  Returned without the leading timeSpan field and the trailing blank."""
    with tempfile.NamedTemporaryFile(suffix=".ev", delete=False) as t:
        dump = t.name
    env = dict(os.environ, MVOLPS_ZMQ_DUMP=dump)
    subprocess.run([REF_BIN, "--server", "5555", "-f", path] + flags, capture_output=True, text=True, timeout=600,
                   env=env, check=True)
    raw = [ln for ln in open(dump).read().splitlines() if ln.strip()]
    os.unlink(dump)
    assert raw and raw[-1] == "END"  # ~IPCDispatch (message.h:232-246) closes the stream with this token
    return [" ".join(ln.split()[1:]) for ln in raw[:-1]]


def event_golden():
    """event_golden.json: the reference's own event stream per B&B case (message.h:141-226 records)."""
    out = []
    for name, gen, args, flags in BNB_CASES:
        if gen is None:
            path, tmp = os.path.join(ROOT, "examples", "small3.lp"), None
        else:
            tmp = tempfile.NamedTemporaryFile(suffix=".lp", delete=False)
            tmp.close()
            path = tmp.name
            instances.write_lp(getattr(instances, gen)(*args), path)
        ev = reference_events(path, flags)
        out.append(dict(name=name, gen=gen, args=args, flags=flags, events=len(ev),
                        events_sha256=hashlib.sha256("\n".join(ev).encode()).hexdigest(), head=ev[:30]))
        if tmp:
            os.unlink(path)
    with open(os.path.join(HERE, "event_golden.json"), "w") as f:
        json.dump({"generator": "tests/golden/make_golden.py --events-only",
                   "source": "reference IPCDispatch::write (message.cpp:32-181) run with --server, messages captured by "
                             "the zmqpp stub; the timeSpan field is dropped",
                   "cases": out}, f, indent=1)
    print(f"wrote {len(out)} event streams")


def main():
    subprocess.run(["make", "-C", os.path.join(ROOT, "oracle"), "all", "ref"], check=True)
    if "--events-only" in sys.argv:
        return event_golden()
    orc = capi.load_library(ORACLE)
    orc.glp_term_out(0)
    cases = []
    for gen, args in LP_CASES:
        inst = getattr(instances, gen)(*args)
        lp = EngineLP.from_instance(orc, inst)
        r = lp.solve()
        cases.append(dict(gen=gen, args=args, status=r["status"], iters=r["iters"], pivots=r["pivots"],
                          pivot_hash=str(r["pivot_hash"]), obj_hex=np.float64(r["obj"]).tobytes().hex(),
                          obj=r["obj"], highs_obj=highs_lp(inst)))
    with open(os.path.join(HERE, "lp_golden.json"), "w") as f:
        json.dump({"generator": "tests/golden/make_golden.py", "cases": cases}, f, indent=1)

    out = []
    for name, gen, args, flags in BNB_CASES:
        if gen is None:
            path = os.path.join(ROOT, "examples", "small3.lp")
            inst = small3_inst()
            tmp = None
        else:
            inst = getattr(instances, gen)(*args)
            tmp = tempfile.NamedTemporaryFile(suffix=".lp", delete=False)
            tmp.close()
            path = tmp.name
            instances.write_lp(inst, path)
        p = subprocess.run([REF_BIN, "-f", path] + flags, capture_output=True, text=True, timeout=600)
        lines = p.stdout.splitlines()
        # the tree is the block of "-<oid>" lines; the solution string follows it (bs.cpp:329-345)
        tree = [ln for ln in lines if re.fullmatch(r"\s*-\d+( [IFB])?\s*", ln)]
        sol = [ln for ln in lines if "Solution is:" in ln]
        iters = [ln for ln in lines if "Solution found after" in ln]
        digest = hashlib.sha256("\n".join(tree).encode()).hexdigest()
        out.append(dict(name=name, gen=gen, args=args, flags=flags, returncode=p.returncode,
                        tree_sha256=digest, tree_lines=len(tree), tree=tree if len(tree) <= 100 else tree[:100], solution=sol[-1].strip() if sol else "",
                        iterations=int(iters[-1].split("after")[1].split()[0]) if iters else None,
                        highs_ilp=highs_ilp(inst)))
        if tmp:
            os.unlink(path)
    with open(os.path.join(HERE, "bnb_golden.json"), "w") as f:
        json.dump({"generator": "tests/golden/make_golden.py",
                   "source": "reference bs.cpp/util.cpp/gmi.cpp/cut.cpp (unmodified) on the oracle shim",
                   "cases": out}, f, indent=1)
    print(f"wrote {len(cases)} LP cases, {len(out)} B&B cases")
    event_golden()


if __name__ == "__main__":
    main()
