"""Golden hashes bench.py asserts in-process (tests/golden/bench_golden.json), all produced by the
CPU ORACLE (oracle/_build/libglpk_oracle.so; test infrastructure) in this container:

  dense     configs[1], m=1024 n=2048 seed 1234: pivot hash of the full solve
  shard     configs[4], m=16384 n=32768 seed 5: pivot hash after 100 / 200 / 300 pivots (the
            column-sharded GPU path must reproduce them at every N; about one minute of 8 threads)
  bnb       configs[2], MDKP 200x50 seed 2024, FIFO: event hash / node count of the runs bench.py
            makes (key = instance, node strategy, cut mode, root heuristic, node budget)
  bnb_close MDKP 30x8 seed 4 run to the proven optimum

Usage: python tests/golden/make_bench_golden.py [--skip-shard]   (re-uses the shard entry then)
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from mvolps_b200 import capi, instances  # noqa: E402
from mvolps_b200.lp import EngineLP, GlpProb  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden", "bench_golden.json")
M64 = (1 << 64) - 1


def prefix_hashes(log, at):
    h, out = 1469598103934665603, {}
    for k, (row, col) in enumerate(log.tolist(), 1):
        for v in (1, row, col):  # mode tag MVB_MODE_PRIMAL2, row, column
            h = ((h ^ v) * 1099511628211) & M64
        if k in at:
            out[str(k)] = str(h)
    return out


def main():
    orc = capi.load_library(os.path.join(ROOT, "oracle", "_build", "libglpk_oracle.so"))
    orc.glp_term_out(capi.GLP_OFF)
    old = json.load(open(OUT)) if os.path.exists(OUT) else {}
    g = {"generator": "tests/golden/make_bench_golden.py (CPU oracle)"}

    only_c3 = "--only-c3" in sys.argv
    if only_c3:
        g["dense"] = old["dense"]
    else:
        lp = EngineLP.from_instance(orc, instances.dense_lp(1024, 2048, 1234))
        r = lp.solve(threads=8)
        g["dense"] = {"1024x2048_s1234": {"pivot_hash": str(r["pivot_hash"]), "pivots": r["pivots"], "obj_hex": float(r["obj"]).hex()}}
    print("dense", g["dense"], flush=True)

    if (only_c3 or "--skip-shard" in sys.argv) and "shard" in old:
        g["shard"] = old["shard"]
    else:
        lp = EngineLP.from_instance(orc, instances.dense_lp(16384, 32768, 5))
        r = lp.solve(max_iters=300, threads=8, log_pivots=1)
        log, _ = lp.pivot_log()
        hs = prefix_hashes(log, (100, 200, 300))
        assert hs["300"] == str(r["pivot_hash"])
        g["shard"] = {"16384x32768": hs}
        lp.free()
    print("shard", g["shard"], flush=True)

    g["bnb"] = {}
    # (vars, rows, batch width, cut mode, root heuristic, node budget); the 512 x 256 line is the
    # BASELINE configs[3] node-LP shape (frontier batches of 4096 node LPs)
    runs = ((200, 50, 256, 0, 0, 80000), (200, 50, 256, 2, 0, 20000), (200, 50, 256, 0, 1, 80000),
            (200, 50, 256, 0, 0, 10000), (200, 50, 256, 2, 0, 10000), (512, 256, 4096, 0, 0, 8192))
    if "--only-c3" in sys.argv:
        g = old
        runs = runs[-1:]
    for nv, nr, bw, cm, rh, nodes in runs:
        inst = instances.mdkp(nv, nr, 2024)
        out = GlpProb.from_instance(orc, inst).bnb(cut_mode=cm, root_heuristic=rh, max_iters=nodes, batch_width=bw,
                                                   total_budget=1)
        key = f"mdkp_{nv}x{nr}_s2024_bs0_cm{cm}_rh{rh}_n{nodes}"
        g["bnb"][key] = {"event_hash": str(out["event_hash"]), "iterations": out["iterations"],
                         "incumbent": out["incumbent"], "has_incumbent": out["has_incumbent"], "oids": out["oids"]}
        print(key, g["bnb"][key], flush=True)

    if not only_c3:
        out = GlpProb.from_instance(orc, instances.mdkp(30, 8, 4)).bnb()
        g["bnb_close"] = {"mdkp_30x8_s4": {"incumbent": out["incumbent"], "iterations": out["iterations"],
                                          "event_hash": str(out["event_hash"])}}
    print(g["bnb_close"], flush=True)
    with open(OUT, "w") as f:
        json.dump(g, f, indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
