"""The legs of bench.py beyond the configs[1] headline: the multi-LP batch
(configs[3]), the column-sharded tableau (configs[4]) and the B&B tree
(configs[2]).  Every function returns the dict rank 0 prints (None on the
other ranks).  Same timing rules as bench.py: CUDA events on the engine
stream, L2 flushed between steps, max over ranks; golden hashes
(tests/golden/bench_golden.json, written by tests/golden/make_bench_golden.py
from the CPU oracle) are ASSERTED in-process."""
import ctypes as C
import json
import os
import time

import numpy as np

from mvolps_b200 import capi, instances
from mvolps_b200.dist import ShardGroup, nccl_comm
from mvolps_b200.lp import EngineLP, GlpProb, solve_batch

ROOT = os.path.dirname(os.path.abspath(__file__))
ORACLE_LIB = os.path.join(ROOT, "oracle", "_build", "libglpk_oracle.so")


_INST = {}  # generated instances, kept between the legs of one invocation


def golden():
    with open(os.path.join(ROOT, "tests", "golden", "bench_golden.json")) as f:
        return json.load(f)


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
    """configs[3] shape: `--lps` node LPs (m=256 n=512, cfg-2 distribution) dealt round-robin over
    the ranks, each rank one mvb_solve_batch launch per step; no data-path collective."""
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
        return res, L.mvb_timer_elapsed_ms(0, 1), L.mvb_last_kernel_ms()

    for _ in range(max(args.warmup, 3)):
        step()
    if dist is not None:
        dist.barrier()
    l0 = L.mvb_kernel_launches()
    piv, ms, kms = 0, 0.0, 0.0
    with ClockSampler(local) as clk:
        for _ in range(args.steps):
            res, t, k = step()
            assert all(r["status"] == 0 for r in res)
            piv += sum(r["pivots"] for r in res)
            ms += t
            kms += k
    launches = int(L.mvb_kernel_launches() - l0)
    (tot_piv,), (max_ms,) = _reduce(dist, [piv], [ms])
    if rank != 0:
        return None
    G, groups, nt = C.c_int(), C.c_int(), C.c_int()
    L.mvb_last_launch(C.byref(G), C.byref(groups), C.byref(nt))
    mode = L.mvb_last_kernel_mode() if hasattr(L, "mvb_last_kernel_mode") else -1
    peak, src = peaks()
    alg = 16.0 * (m + 1) * (n + 1) * piv      # this rank, per timed region
    # DRAM bytes the launch really needs: the resident kernel loads and stores every tableau once;
    # the streaming kernel moves ~16*R*ld/K per pivot (K pending updates per flush pass)
    R, ld = m + 1, (n + 15) // 16 * 16
    per_lp_once = 16.0 * R * ld
    opt = getattr(L, "mvb_engine_option", None)
    lz_k = opt(b"last_lz_k", -1) if opt and mode == 1 else 0
    stream = bool(opt(b"last_stream", -1)) if opt and mode == 1 else False
    # streaming kernel: a flush pass (read + write of the tableau) every lz_k pivots, plus the pivot
    # column (one 32-byte sector per row) and the pivot row of every pivot
    est = piv * (per_lp_once / lz_k + 32.0 * R + 8.0 * ld) if lz_k > 0 else None
    return {
        "metric": "lp_pivots_per_sec", "value": tot_piv / (max_ms * 1e-3), "unit": "pivots/s", "n_gpus": world,
        "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": max_ms / args.steps,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"{args.lps} node LPs m={m} n={n} ({args.distinct} distinct seeds, cloned on the device), "
                               "full solves, dealt round-robin over the ranks (BASELINE configs[3] shape)",
                   "lps_per_sec": args.lps * args.steps / (max_ms * 1e-3), "ctas_per_lp": G.value,
                   "groups": groups.value, "threads": nt.value, "kernel_mode": mode, "l2": "flushed between steps"},
        "roofline": {"bound": "hbm" if mode != 2 else "shared-memory / issue (tableau resident on chip)",
                     "algorithmic_GBs_rank0": alg / (kms * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                     "min_dram_bytes_per_step_rank0": per_lp_once * len(mine),
                     "pending_depth": lz_k, "streaming_kernel": stream,
                     "dram_bytes_estimate_rank0": est,
                     "dram_frac_estimate": None if est is None else est / (kms * 1e-3) / 1e9 / peak,
                     "dram_frac_ncu": "0.57 (3.70 TB/s) on 592 LPs of 256x512, profiles/r2_stream_kernel_summary.txt"
                                      if stream else None,
                     "note": "frac is reported from DRAM bytes (ncu capture under profiles/), never from the "
                             "algorithmic 16*R*C*pivots: lazy / on-chip updates move far fewer bytes",
                     "peak_source": src, "kernel_ms_per_launch_rank0": kms / args.steps},
        "gpu_launches": launches, "clocks": clk.summary(),
    }


def run_shard(args, L, rank, world, local, dist, ClockSampler, peaks, steps=None, warmup=None, comm=None):
    """configs[4]: one dense LP m=16384 n=32768, tableau column-sharded over the ranks, `--pivots`
    iterations per step from the slack basis; the pivot hash is asserted against the oracle's."""
    m, n = args.m or 16384, args.n or 32768
    steps = steps or args.steps
    warmup = max(1, args.warmup if warmup is None else warmup)
    inst = _INST.get((m, n, 5))
    if inst is None:
        inst = _INST[(m, n, 5)] = instances.dense_lp(m, n, 5)
    own_comm = comm is None and dist is not None
    if own_comm:
        comm = nccl_comm(L, rank, world, dist)
    gold = golden()["shard"].get(f"{m}x{n}", {}).get(str(args.pivots))

    from mvolps_b200.dist import slice_instance
    sl = {rank: slice_instance(inst, world, rank)}  # this rank's columns of A (host, pageable) and the bounds
    h2d = sl[rank]["A"].nbytes + sl[rank]["cost"].nbytes * 2 + sl[rank]["lb"].nbytes * 2
    d2h = 8 * (m + n) + 4 * (m + 2 * n) + 12 * (m + 1)
    tot_ms, tot_piv, r, how = 0.0, 0, None, ""
    e2e_s = 0.0
    l0 = None
    with ClockSampler(local) as clk:
        for k in range(warmup + steps):
            if dist is not None:
                dist.barrier()
            t0 = time.perf_counter()
            g = ShardGroup(L, inst, world, comm=comm, rank=rank, slices=sl)  # reload: a step starts from the slack basis
            t_load = time.perf_counter() - t0
            L.mvb_flush_l2()
            if dist is not None:
                dist.barrier()
            if k == warmup:
                l0 = L.mvb_kernel_launches()
            t1 = time.perf_counter()
            r = g.solve(max_iters=args.pivots)
            x = g.primal()
            t_solve = time.perf_counter() - t1
            how = g.last_exchange()
            if gold is not None and str(r["pivot_hash"]) != gold:
                raise AssertionError(f"column-sharded pivot hash {r['pivot_hash']} != oracle golden {gold} "
                                     f"({args.pivots} pivots, rank {rank} of {world})")
            if k >= warmup:
                tot_ms += g.last_ms()
                tot_piv += r["pivots"]
                e2e_s += t_load + t_solve
            g.free()
    launches = int(L.mvb_kernel_launches() - l0)
    _, (max_ms, max_e2e) = _reduce(dist, [0.0], [tot_ms, e2e_s])
    if own_comm:
        L.mvb_comm_free(comm)
    if rank != 0:
        return None
    peak, src = peaks()
    us_per_pivot = 1e3 * max_ms / tot_piv
    c0, nloc = C.c_int(), C.c_int()
    L.mvb_shard_partition(n, world, 0, C.byref(c0), C.byref(nloc))
    per_gpu_bytes = 16.0 * (m + 1) * nloc.value          # K4 on this GPU's columns, read + write once
    ach = per_gpu_bytes / (us_per_pivot * 1e-6) / 1e9
    return {
        "metric": "simplex_pivots_per_sec", "value": tot_piv / (max_ms * 1e-3), "unit": "pivots/s", "n_gpus": world,
        "steps": steps, "warmup": warmup, "ms_per_step": max_ms / steps, "us_per_pivot": us_per_pivot,
        "higher_is_better": True, "scaling": "strong", "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"dense LP m={m} n={n} seed=5, tableau column-sharded over {world} GPU(s), first "
                               f"{args.pivots} pivots from the slack basis per step (BASELINE configs[4])",
                   "exchange": "one exchange per pivot: 64 B header + 8*(m+1) B column per rank; " + how,
                   "pivot_hash": str(r["pivot_hash"]),
                   "pivot_hash_ok": None if gold is None else True,
                   "pivot_hash_golden": "tests/golden/bench_golden.json (CPU oracle, tests/golden/make_bench_golden.py); "
                                        "asserted on every rank and step",
                   "l2": "tableau shard larger than L2 (flushed between steps as well)"},
        "roofline": {"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                     "per": "GPU (rank 0's columns; every rank streams its own shard concurrently)",
                     "traffic": per_gpu_bytes,
                     "traffic_source": "computed: 16*(m+1)*local columns per shard_update_kernel launch (ncu at world 1: "
                                       "8.55 GB per launch = 0.995 of it, profiles/r1e_shard_update_summary.txt)",
                     "achieved_is": "algorithmic = real traffic here (eager update, every element read and written "
                                    "once per pivot) / WHOLE per-pivot time incl. the exposed part of the price / "
                                    "exchange / decide chain",
                     "peak_source": src,
                     "kernel": "shard_update_kernel (eager rank-1 update of the local columns; the price / exchange / "
                               "decide chain of the next iteration overlaps it on a second stream)"},
        "e2e": {"value": tot_piv / max_e2e, "unit": "pivots/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                "what": f"host wall clock, max over ranks: mvb_shard_create + mvb_shard_load of this rank's columns from "
                        f"pageable host memory + mvb_shard_solve ({args.pivots} pivots) + mvb_shard_get_primal; the load is "
                        "paid once per LP, so a full solve (tens of thousands of pivots) amortises it"},
        "gpu_launches": launches, "clocks": clk.summary(),
    }


def _bnb_once(L, prob, comm, **kw):
    t0 = time.perf_counter()
    out = prob.bnb(comm=comm, **kw)
    return out, time.perf_counter() - t0


def run_bnb(args, L, rank, world, local, dist, ClockSampler, peaks, steps=None, warmup=None, comm=None,
            cpu_baseline=True):
    """configs[2]: MDKP 200 vars x 50 rows (seed 2024), FIFO B&B through mvb_bnb_run[_dist] (batched
    node LPs on the device).  `--nodes` is the node budget of the WHOLE job at every N (strong
    scaling); N > 1: one tree, node-sharded (NCCL incumbent max-reduction per round, node work
    stealing) or, with --dist-mode 1, round-synchronous and sequence-identical.  At N = 1 (and in
    the sequence-identical mode at every N) the event hash is asserted against the oracle's."""
    steps = steps or args.steps
    warmup = max(1, args.warmup if warmup is None else warmup)
    inst = instances.mdkp(args.n or 200, args.m or 50, args.inst_seed)
    prob = GlpProb.from_instance(L, inst)
    own_comm = comm is None and dist is not None
    if own_comm:
        comm = nccl_comm(L, rank, world, dist)
    kw = dict(cut_mode=args.cut_mode, batch_width=args.batch_width, max_iters=args.nodes,
              no_steal=int(args.no_steal), node_strategy=args.node_strategy,
              no_speculation=int(args.no_speculation), total_budget=1, root_heuristic=int(args.root_heuristic),
              dist_mode=args.dist_mode, cut_freq=args.cut_freq, no_event_text=1)
    key = f"mdkp_{args.n or 200}x{args.m or 50}_s{args.inst_seed}_bs{args.node_strategy}_cm{args.cut_mode}_rh{int(args.root_heuristic)}_n{args.nodes}"
    gold = golden()["bnb"].get(key) if args.cut_freq == 1.0 else None
    for _ in range(warmup):
        _bnb_once(L, prob, comm, **kw)
    if dist is not None:
        dist.barrier()
    l0 = L.mvb_kernel_launches()
    nodes, secs, out = 0, 0.0, None
    with ClockSampler(local) as clk:
        for _ in range(steps):
            if dist is not None:
                dist.barrier()
            out, dt = _bnb_once(L, prob, comm, **kw)
            nodes += out["total_nodes"]
            secs += dt
            if gold is not None and (world == 1 or args.dist_mode == 1):
                if str(out["event_hash"]) != gold["event_hash"] or out["iterations"] != gold["iterations"]:
                    raise AssertionError(f"B&B event hash {out['event_hash']} / {out['iterations']} nodes != oracle "
                                         f"golden {gold} ({key}, rank {rank} of {world})")
    launches = int(L.mvb_kernel_launches() - l0)
    (sent, recv), (max_s, steal_rounds) = _reduce(dist, [out["nodes_sent"], out["nodes_received"]],
                                                  [secs, out["steal_rounds"]])
    if own_comm:
        L.mvb_comm_free(comm)
    if rank != 0:
        return None
    line = {
        "metric": "bnb_nodes_per_sec", "value": nodes / max_s, "unit": "nodes/s", "n_gpus": world,
        "steps": steps, "warmup": warmup, "ms_per_step": 1e3 * max_s / steps,
        "higher_is_better": True, "scaling": "strong", "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"MDKP {args.n or 200} vars x {args.m or 50} rows seed {args.inst_seed}, "
                               f"{'best-first' if args.node_strategy else 'FIFO'} B&B, node budget {args.nodes} for the "
                               f"whole job, cut mode {args.cut_mode}"
                               + (f" at frequency {args.cut_freq}" if args.cut_freq != 1.0 else "")
                               + f", batch width {args.batch_width}"
                               + (", root rounding heuristic" if args.root_heuristic else "")
                               + (" (BASELINE configs[3] node-LP shape)" if (args.n, args.m) == (512, 256) else
                                  " (BASELINE configs[2])") + "; host wall clock around mvb_bnb_run[_dist] (host logic + "
                               "H2D/D2H included); event lines not formatted (the reference formats them only with "
                               "--server), event hash computed and asserted"
                               + ("" if world == 1 else
                                  ("; one tree, round-synchronous sequence-identical mode" if args.dist_mode == 1 else
                                   "; one tree node-sharded over the ranks, NCCL incumbent max-reduction per round, "
                                   "node work stealing" + (" off" if args.no_steal else ""))),
                   "nodes_total": out["total_nodes"], "nodes_rank0": out["iterations"], "split_at": out["split_at"],
                   "lp_solves_rank0": out["lp_solves"], "engine_launches_rank0": out["launches"],
                   "row_gathers_rank0": out["launches_aux"],
                   "event_hash_rank0": str(out["event_hash"]),
                   "event_hash_ok": None if gold is None or not (world == 1 or args.dist_mode == 1) else True,
                   "has_incumbent": out["has_incumbent"], "incumbent": out["incumbent"],
                   "hit_limit_rank0": out["hit_limit"], "nodes_sent": int(sent), "nodes_received": int(recv),
                   "steal_rounds": int(steal_rounds), "spec_nodes": out["spec_nodes"],
                   "spec_adopted": out["spec_adopted"]},
        "gpu_launches": launches, "clocks": clk.summary(),
    }
    if cpu_baseline and world == 1 and not args.no_cpu_baseline:
        orc = capi.load_library(ORACLE_LIB)
        orc.glp_term_out(capi.GLP_OFF)
        po = GlpProb.from_instance(orc, inst)
        kwo = dict(kw)
        kwo["max_iters"] = min(args.nodes, 20000)
        oo, dt = _bnb_once(orc, po, None, **kwo)
        line["cpu_baseline"] = {"value": oo["iterations"] / dt, "unit": "nodes/s", "cores": 1,
                                "kind": "port", "sample": f"same driver on the oracle engine, first {oo['iterations']} "
                                f"nodes of the same tree, {dt:.1f} s, 1 thread; host has {os.cpu_count()} cores"}
        ref = reference_driver_nodes_per_sec(inst, args)
        if ref is not None:
            line["cpu_baseline_reference_driver"] = ref
    return line


def run_bnb_close(args, L, rank, world, local, dist, comm=None):
    """Time to the proven optimum on an MDKP that closes (30 vars x 8 rows, seed 4; 161 k nodes in
    the serial FIFO order): same optimum asserted at every N."""
    inst = instances.mdkp(30, 8, 4)
    gold = golden()["bnb_close"]["mdkp_30x8_s4"]
    prob = GlpProb.from_instance(L, inst)
    own_comm = comm is None and dist is not None
    if own_comm:
        comm = nccl_comm(L, rank, world, dist)
    kw = dict(batch_width=args.batch_width, no_steal=int(args.no_steal), dist_mode=args.dist_mode, no_event_text=1)
    _bnb_once(L, prob, comm, **kw)
    if dist is not None:
        dist.barrier()
    out, dt = _bnb_once(L, prob, comm, **kw)
    if out["incumbent"] != gold["incumbent"]:
        raise AssertionError(f"closing instance: optimum {out['incumbent']} != golden {gold['incumbent']}")
    if world == 1 and out["iterations"] != gold["iterations"]:
        raise AssertionError(f"closing instance: {out['iterations']} nodes != golden {gold['iterations']}")
    (sent,), (max_s, steal_rounds) = _reduce(dist, [out["nodes_sent"]], [dt, out["steal_rounds"]])
    if own_comm:
        L.mvb_comm_free(comm)
    if rank != 0:
        return None
    return {"workload": "MDKP 30 vars x 8 rows seed 4, FIFO, no cuts, run to the proven optimum",
            "ms_to_optimum": 1e3 * max_s, "optimum": out["incumbent"], "optimum_ok": True,
            "nodes_total": out["total_nodes"], "nodes_serial": gold["iterations"], "nodes_sent": int(sent),
            "steal_rounds": int(steal_rounds), "split_at": out["split_at"]}


def reference_driver_nodes_per_sec(inst, args, seconds=10.0):
    """the reference's UNMODIFIED driver (2test.cpp bs.cpp util.cpp cut.cpp gmi.cpp
    message.cpp compiled where they lie under /root/reference by oracle/Makefile -> oracle/_ref/
    mvolps_ref_oracle) on the CPU oracle shim, single-threaded like the reference, default logging,
    on the same instance.  It has no node budget (bs.cpp:320 stops at 200 000 iterations), so it
    runs for `seconds` and the nodes it got through are read off its own log: printInfo
    (util.cpp:414-473) logs one solution line per popped feasible node (a slight under-count:
    infeasible nodes log none)."""
    import re
    import subprocess
    import tempfile
    exe = os.path.join(ROOT, "oracle", "_ref", "mvolps_ref_oracle")
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
