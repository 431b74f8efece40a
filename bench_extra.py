"""bench.py --workload batch|shard: the multi-LP and the column-sharded configs
(BASELINE.json configs[3] and configs[4]).  Same timing rules as bench.py: CUDA
events on the engine stream, L2 flushed between steps, max over ranks."""
import ctypes as C
import json
import time

import numpy as np

from mvolps_b200 import capi, instances
from mvolps_b200.dist import ShardGroup, nccl_comm
from mvolps_b200.lp import EngineLP, solve_batch


def _reduce(dist, vals_sum, vals_max):
    if dist is None:
        return vals_sum, vals_max
    import torch
    a = torch.tensor(vals_sum, dtype=torch.float64, device="cuda")
    b = torch.tensor(vals_max, dtype=torch.float64, device="cuda")
    dist.all_reduce(a)
    dist.all_reduce(b, op=dist.ReduceOp.MAX)
    return a.tolist(), b.tolist()


def run_batch(args, L, rank, world, local, dist, ClockSampler, peaks):
    """configs[3]: `--lps` node LPs (m=256 n=512, cfg-2 distribution) dealt round-robin over the
    ranks, each rank one mvb_solve_batch launch per step; no data-path collective."""
    m, n = args.m or 256, args.n or 512
    mine = [k for k in range(args.lps) if k % world == rank]
    seeds = sorted({k % args.distinct for k in mine})
    tmpl = {sd: EngineLP.from_instance(L, instances.dense_lp(m, n, sd)) for sd in seeds}
    work = [tmpl[k % args.distinct].clone() for k in mine]

    def step():
        for w, k in zip(work, mine):
            w.assign(tmpl[k % args.distinct])
        L.mvb_flush_l2()
        L.mvb_timer_record(0)
        res = solve_batch(L, work, ctas_per_lp=args.ctas_per_lp)
        L.mvb_timer_record(1)
        return res, L.mvb_timer_elapsed_ms(0, 1)

    for _ in range(max(args.warmup, 3)):
        step()
    if dist is not None:
        dist.barrier()
    l0 = L.mvb_kernel_launches()
    piv, ms = 0, 0.0
    with ClockSampler(local) as clk:
        for _ in range(args.steps):
            res, t = step()
            assert all(r["status"] == 0 for r in res)
            piv += sum(r["pivots"] for r in res)
            ms += t
    launches = int(L.mvb_kernel_launches() - l0)
    (tot_piv,), (max_ms,) = _reduce(dist, [piv], [ms])
    if rank != 0:
        return
    G, groups, nt = C.c_int(), C.c_int(), C.c_int()
    L.mvb_last_launch(C.byref(G), C.byref(groups), C.byref(nt))
    peak, src = peaks()
    alg = 16.0 * (m + 1) * (n + 1) * tot_piv
    ach = alg / (max_ms * 1e-3) / 1e9
    print(json.dumps({
        "metric": "lp_pivots_per_sec", "value": tot_piv / (max_ms * 1e-3), "unit": "pivots/s", "n_gpus": world,
        "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": max_ms / args.steps,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"{args.lps} node LPs m={m} n={n} ({args.distinct} distinct seeds, cloned on the device), "
                               "full solves, dealt round-robin over the ranks (BASELINE configs[3])",
                   "lps_per_sec": args.lps * args.steps / (max_ms * 1e-3), "ctas_per_lp": G.value,
                   "groups": groups.value, "l2": "flushed between steps"},
        "roofline": {"bound": "hbm", "achieved": ach, "peak": peak * world, "unit": "GB/s", "frac": ach / (peak * world),
                     "traffic": None, "peak_source": src, "kernel": "simplex_ps_kernel (batched, G CTAs per LP)"},
        "gpu_launches": launches, "clocks": clk.summary(),
    }))


def run_shard(args, L, rank, world, local, dist, ClockSampler, peaks):
    """configs[4]: one dense LP, tableau column-sharded over the ranks, `--pivots` iterations per step."""
    m, n = args.m or 16384, args.n or 32768
    inst = instances.dense_lp(m, n, 5)
    comm = nccl_comm(L, rank, world, dist) if dist is not None else None

    def fresh():
        return ShardGroup(L, inst, world, comm=comm, rank=rank)

    tot_ms, tot_piv, r = 0.0, 0, None
    l0 = None
    with ClockSampler(local) as clk:
        for k in range(max(args.warmup, 1) + args.steps):
            g = fresh()  # reload: a step starts from the slack basis
            L.mvb_flush_l2()
            if dist is not None:
                dist.barrier()
            if k == max(args.warmup, 1):
                l0 = L.mvb_kernel_launches()
            r = g.solve(max_iters=args.pivots)
            how = g.last_exchange()
            if k >= max(args.warmup, 1):
                tot_ms += g.last_ms()
                tot_piv += r["pivots"]
            g.free()
    launches = int(L.mvb_kernel_launches() - l0)
    _, (max_ms,) = _reduce(dist, [0.0], [tot_ms])
    if rank != 0:
        return
    peak, src = peaks()
    ach = 16.0 * (m + 1) * (n + 1) * tot_piv / (max_ms * 1e-3) / 1e9
    print(json.dumps({
        "metric": "simplex_pivots_per_sec", "value": tot_piv / (max_ms * 1e-3), "unit": "pivots/s", "n_gpus": world,
        "steps": args.steps, "warmup": max(args.warmup, 1), "ms_per_step": max_ms / args.steps,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"dense LP m={m} n={n} seed=5, tableau column-sharded over {world} GPU(s), first "
                               f"{args.pivots} pivots from the slack basis per step (BASELINE configs[4])",
                   "exchange": "one exchange per pivot: 64 B header + 8*(m+1) B column per rank; " + how,
                   "pivot_hash": str(r["pivot_hash"]), "l2": "tableau shard larger than L2"},
        "roofline": {"bound": "hbm", "achieved": ach, "peak": peak * world, "unit": "GB/s", "frac": ach / (peak * world),
                     "traffic": 8551384000 if world == 1 else None, "peak_source": src,
                     "traffic_source": "dram__bytes_read.sum + dram__bytes_write.sum of one shard_update_kernel launch at world 1, "
                                       "ncu --set full, profiles/r1e_shard_update_summary.txt" if world == 1 else None,
                     "kernel": "shard_update_kernel (eager rank-1 update of the local columns; the price / exchange / "
                               "decide chain of the next iteration overlaps it on a second stream)"},
        "gpu_launches": launches, "clocks": clk.summary(),
    }))


def run_bnb(args, L, rank, world, local, dist, ClockSampler, peaks):
    """configs[2]: MDKP 200 vars x 50 rows (seed 2024), FIFO B&B with a node budget through
    mvb_bnb_run (batched node LPs on the device); the same driver on the CPU oracle is the baseline.
    N > 1: ONE tree, node-sharded (mvb_bnb_run_dist): lock-step ramp-up, then every rank works on
    its share of the open nodes with the incumbent exchanged by NCCL max-reduction every round;
    the node budget is per rank."""
    import os
    from mvolps_b200.lp import GlpProb
    inst = instances.mdkp(args.n or 200, args.m or 50, args.inst_seed)
    prob = GlpProb.from_instance(L, inst)
    comm = nccl_comm(L, rank, world, dist) if dist is not None else None

    def step(lib, p, c=None):
        t0 = time.perf_counter()
        out = p.bnb(cut_mode=args.cut_mode, batch_width=args.batch_width, max_iters=args.nodes, comm=c,
                    no_steal=int(args.no_steal), node_strategy=args.node_strategy,
                    no_speculation=int(args.no_speculation))
        return out, time.perf_counter() - t0

    for _ in range(max(args.warmup, 1)):
        step(L, prob, comm)
    if dist is not None:
        dist.barrier()
    l0 = L.mvb_kernel_launches()
    nodes, secs, out = 0, 0.0, None
    with ClockSampler(local) as clk:
        for _ in range(args.steps):
            out, dt = step(L, prob, comm)
            nodes += out["total_nodes"]
            secs += dt
    launches = int(L.mvb_kernel_launches() - l0)
    _, (max_s,) = _reduce(dist, [0.0], [secs])
    if rank != 0:
        return
    line = {
        "metric": "bnb_nodes_per_sec", "value": nodes / max_s, "unit": "nodes/s", "n_gpus": world,
        "steps": args.steps, "warmup": max(args.warmup, 1), "ms_per_step": 1e3 * max_s / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"MDKP {args.n or 200} vars x {args.m or 50} rows seed {args.inst_seed}, {'best-first' if args.node_strategy else 'FIFO'} B&B, node budget "
                               f"{args.nodes} per rank, cut mode {args.cut_mode}, batch width {args.batch_width} "
                               "(BASELINE configs[2]); host wall clock around mvb_bnb_run[_dist] (host logic + H2D/D2H "
                               "included)" + ("; one tree node-sharded over the ranks, NCCL incumbent max-reduction "
                                              "per round" if world > 1 else ""),
                   "nodes_total": out["total_nodes"], "nodes_rank0": out["iterations"], "split_at": out["split_at"],
                   "lp_solves_rank0": out["lp_solves"], "engine_launches_rank0": out["launches"],
                   "event_hash_rank0": str(out["event_hash"]), "incumbent": out["incumbent"],
                   "hit_limit_rank0": out["hit_limit"], "nodes_sent_rank0": out["nodes_sent"],
                   "nodes_received_rank0": out["nodes_received"], "steal_rounds": out["steal_rounds"], "spec_nodes": out["spec_nodes"],
                   "spec_adopted": out["spec_adopted"]},
        "gpu_launches": launches, "clocks": clk.summary(),
    }
    if not args.no_cpu_baseline and world == 1:
        orc = capi.load_library(os.path.join(os.path.dirname(os.path.abspath(__file__)), "oracle", "_build",
                                             "libglpk_oracle.so"))
        orc.glp_term_out(capi.GLP_OFF)
        po = GlpProb.from_instance(orc, inst)
        oo, dt = step(orc, po)
        line["cpu_baseline"] = {"value": oo["iterations"] / dt, "unit": "nodes/s", "cores": os.cpu_count(),
                                "kind": "port", "sample": f"same driver on the oracle engine, {oo['iterations']} nodes, "
                                f"{dt:.1f} s; event hash {'equal' if oo['event_hash'] == out['event_hash'] else 'DIFFERENT'}"}
        ref = reference_driver_nodes_per_sec(inst, args)
        if ref is not None:
            line["cpu_baseline_reference_driver"] = ref
    print(json.dumps(line))


def reference_driver_nodes_per_sec(inst, args, seconds=10.0):
    """the reference's UNMODIFIED driver (2test.cpp bs.cpp util.cpp cut.cpp gmi.cpp
    message.cpp compiled where they lie under /root/reference by oracle/Makefile -> oracle/_ref/
    mvolps_ref_oracle) on the CPU oracle shim, single-threaded like the reference, default logging,
    on the same instance.  It has no node budget (bs.cpp:320 stops at 200 000 iterations), so it
    runs for `seconds` and the nodes it got through are read off its own log: printInfo
    (util.cpp:414-473) logs one solution line per popped feasible node (a slight under-count:
    infeasible nodes log none)."""
    import os
    import re
    import subprocess
    import tempfile
    exe = os.path.join(os.path.dirname(os.path.abspath(__file__)), "oracle", "_ref", "mvolps_ref_oracle")
    if not os.path.exists(exe):
        return None
    with tempfile.TemporaryDirectory() as td:
        lp = os.path.join(td, "inst.lp")
        instances.write_lp(inst, lp)
        cmd = [exe, "-f", lp, "-cm", str(min(args.cut_mode, 1)), "-bs", str(args.node_strategy)]
        t0 = time.perf_counter()
        try:
            out = subprocess.run(cmd, capture_output=True, text=True, timeout=seconds).stdout
            done = True
        except subprocess.TimeoutExpired as e:
            out = e.stdout.decode(errors="replace") if isinstance(e.stdout, bytes) else (e.stdout or "")
            done = False
        dt = time.perf_counter() - t0
    popped = sum(1 for ln in out.splitlines() if "*(x[" in ln)
    m = re.search(r"Solution found after (\d+) iterations", out)
    nodes = int(m.group(1)) if (done and m) else popped
    if nodes == 0:
        return None
    return {"value": nodes / dt, "unit": "nodes/s", "cores": 1, "kind": "reference",
            "sample": f"unmodified reference driver (oracle/_ref/mvolps_ref_oracle) on the oracle shim, default logging, "
                      f"{'ran to the end' if done else 'stopped after'} {dt:.1f} s, {nodes} nodes by its own log"}
