#!/usr/bin/env python
"""bench.py -- simplex pivots/s of the LP-relaxation hot path on B200.

Headline (metric / value / e2e / roofline / cpu_baseline), at every N:
BASELINE.json configs[1] -- one dense feasible LP relaxation m=1024 n=2048 fp64
per GPU (the SAME seed-1234 instance on every rank, so the N-GPU curve measures
the hardware and not instance differences), solved from the slack basis to
optimality by the persistent cooperative kernel.  One step = one full solve.
`value`: inputs resident in HBM (a pristine device tableau is cloned
device-to-device inside the timed region), CUDA events on the engine's launch
stream, L2 flushed between steps.  `e2e`: the same solve through the
reference-facing call glp_simplex() on a host glp_prob (dense build, H2D, solve,
D2H inside the timed region).  This headline has no data-path collective
(independent LPs: "open nodes sharded across GPUs").

The same invocation then runs, at the same N, the two sharded paths of the
north_star and attaches them to the one JSON line:
  "sharded": BASELINE configs[4], one tableau m=16384 n=32768 column-sharded over
             the N GPUs, one exchange per pivot, 300 pivots per step; the pivot
             hash is asserted against the CPU oracle's (tests/golden/bench_golden.json)
  "bnb":     BASELINE configs[2], MDKP 200x50 B&B with a fixed node budget for the
             WHOLE job (strong scaling), without and with GMI cuts, one tree
             node-sharded over the N GPUs (NCCL incumbent max-reduction + work
             stealing); plus the time to the proven optimum on an instance that closes
(--headline-only skips them; --workload X runs one leg alone.)

--impl reference: the CPU restatement (oracle) on all host threads, bounded
sample (the reference's own LP engine, GLPK, is absent from this image; see
DESIGN.md).
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from mvolps_b200 import capi, instances  # noqa: E402
from mvolps_b200.lp import EngineLP, GlpProb  # noqa: E402

M, N_COLS, SEED = 1024, 2048, 1234
ORACLE_LIB = os.path.join(ROOT, "oracle", "_build", "libglpk_oracle.so")
METRIC = "simplex_pivots_per_sec"
REF_SAMPLE_PIVOTS = 4000
# one config for both arms (the reference arm prints exactly this dict)
CONFIG = {
    "workload": f"dense LP m={M} n={N_COLS} seed={SEED} (BASELINE configs[1]), one LP per GPU (same instance on every "
                "rank), solved from the slack basis",
    "l2": "flushed between steps (256 MiB memset); within a step the 2x16.8 MB ping-pong tableau is L2-resident",
    "reference_arm_sample": f"the CPU arm times the first {REF_SAMPLE_PIVOTS} pivots of the same solve per step (the "
                            "per-pivot cost of a dense tableau is constant: 16*R*C bytes whatever the pivot), the GPU arm "
                            "the whole solve",
}
# DRAM traffic and L2 throughput of the timed launch (one full configs[1] solve), from the ncu capture
# of this very command committed under profiles/ (see NCU_SOURCE)
NCU_DRAM_BYTES_PER_LAUNCH = 17550848 + 3976704
NCU_LTS_THROUGHPUT_PCT = 11.2
NCU_SOURCE = "profiles/r1d_bench_ps_summary.txt"


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled during the timed region."""

    def __init__(self, index):
        self.rows, self.stop = [], False
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")
        self.cmd = ["nvidia-smi", f"--id={index}", f"--query-gpu={q}", "--format=csv,noheader,nounits"]
        self.t = threading.Thread(target=self.run, daemon=True)

    def run(self):
        while not self.stop:
            try:
                out = subprocess.run(self.cmd, capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            time.sleep(0.1)

    def __enter__(self):
        self.t.start()
        return self

    def __exit__(self, *a):
        self.stop = True
        self.t.join(timeout=6)

    def summary(self):
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unsampled"]}
        sm = sorted(float(r[0]) for r in self.rows)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for k, n in enumerate(names) if any(r[2 + k].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(self.rows[0][1]), "reasons": reasons,
                "samples": len(self.rows)}


def cpu_sample(threads, max_iters):
    """Oracle on the same workload, first `max_iters` pivots of the solve."""
    orc = capi.load_library(ORACLE_LIB)
    lp = EngineLP.from_instance(orc, instances.dense_lp(M, N_COLS, SEED))
    t0 = time.perf_counter()
    r = lp.solve(max_iters=max_iters, threads=threads)
    dt = time.perf_counter() - t0
    return r["pivots"] / dt, r["pivots"], dt


def run_reference(args, rank, world):
    """The reference arm: the path's CPU implementation on the box's host cores.  GLPK (what the
    reference links) is absent, so this is the oracle port (kind "port") with OpenMP over the
    tableau rows on every host thread; the reference itself is single-threaded, so the 1-thread
    rate is printed beside it."""
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    for _ in range(args.warmup):
        cpu_sample(cores, 50)
    tot_p, tot_t = 0, 0.0
    for _ in range(args.steps):
        _, p, dt = cpu_sample(cores, REF_SAMPLE_PIVOTS)
        tot_p += p
        tot_t += dt
    v = tot_p / tot_t
    v1, p1, dt1 = cpu_sample(1, 1500)
    sample = (f"first {REF_SAMPLE_PIVOTS} pivots of the m={M} n={N_COLS} seed={SEED} solve per step, OpenMP over tableau "
              f"rows on {cores} threads")
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": "pivots/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * tot_t / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic", "config": CONFIG,
        "cpu_baseline": {"value": v, "unit": "pivots/s", "cores": cores, "kind": "port", "sample": sample,
                         "single_thread": {"value": v1, "unit": "pivots/s", "cores": 1,
                                           "sample": f"first {p1} pivots, {dt1:.1f} s (the reference is single-threaded)"}},
        "e2e": {"value": v, "unit": "pivots/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def run_dense(args, L, rank, world, local, dist):
    """configs[1], the headline: returns rank 0's line (None elsewhere)."""
    inst = instances.dense_lp(M, N_COLS, SEED)
    template = EngineLP.from_instance(L, inst)      # pristine tableau, resident in HBM
    work = template.clone()
    host_prob = GlpProb.from_instance(L, inst)      # what a GLPK client holds on the host

    def barrier():
        if dist is not None:
            dist.barrier()

    def device_step():
        L.mvb_timer_record(0)
        work.assign(template)                       # device-to-device, part of the step
        r = work.solve(ctas_per_lp=args.ctas_per_lp)
        L.mvb_timer_record(1)
        ms = L.mvb_timer_elapsed_ms(0, 1)
        return r, ms, L.mvb_last_kernel_ms()

    def e2e_step():
        t0 = time.perf_counter()
        p = host_prob.copy()                        # host copy: no engine state yet
        p.simplex()                                 # dense build + H2D + solve + D2H
        obj = p.obj_val()
        its = L.glp_get_it_cnt(p.h)
        dt = time.perf_counter() - t0
        p.free()
        return its, obj, dt

    warm = max(args.warmup, 3)
    for _ in range(warm):
        device_step()
        L.mvb_flush_l2()
    e2e_step()

    barrier()
    launches0 = L.mvb_kernel_launches()
    piv = 0
    step_ms = kern_ms = 0.0
    gold = None
    try:
        from bench_extra import golden
        gold = golden()["dense"].get(f"{M}x{N_COLS}_s{SEED}")
    except Exception:
        gold = None
    with ClockSampler(local) as clk:
        for _ in range(args.steps):
            L.mvb_flush_l2()
            r, ms, kms = device_step()
            assert r["status"] == 0, r
            if gold is not None and str(r["pivot_hash"]) != gold["pivot_hash"]:
                raise AssertionError(f"configs[1] pivot hash {r['pivot_hash']} != oracle golden {gold['pivot_hash']}")
            piv += r["pivots"]
            step_ms += ms
            kern_ms += kms
        launches = int(L.mvb_kernel_launches() - launches0)
        barrier()
        e_piv, e_t = 0, 0.0
        for _ in range(args.steps):
            its, obj, dt = e2e_step()
            e_piv += its
            e_t += dt
    barrier()
    tot_piv, max_ms, e_tot_piv, e_max_t = piv, step_ms, e_piv, e_t
    if dist is not None:
        import torch
        t = torch.tensor([piv, e_piv], dtype=torch.float64, device="cuda")
        dist.all_reduce(t)
        tm = torch.tensor([step_ms, e_t], dtype=torch.float64, device="cuda")
        dist.all_reduce(tm, op=dist.ReduceOp.MAX)
        tot_piv, e_tot_piv = t[0].item(), t[1].item()
        max_ms, e_max_t = tm[0].item(), tm[1].item()
    if rank != 0:
        return None
    G, groups, nt = C.c_int(), C.c_int(), C.c_int()
    L.mvb_last_launch(C.byref(G), C.byref(groups), C.byref(nt))
    peak, peak_src = peaks()
    pivots_per_step = piv / args.steps
    alg_bytes = 16.0 * (M + 1) * (N_COLS + 1) * pivots_per_step   # per launch (= per solve)
    k_s = kern_ms / args.steps * 1e-3
    achieved = alg_bytes / k_s / 1e9
    m, n = M, N_COLS
    h2d = 8 * (m * n + 2 * n + 2 * (n + m)) + 4 * (m + 2 * n)
    d2h = 8 * (n + m) + 4 * (m + 2 * n) + 96
    return {
        "metric": METRIC, "value": tot_piv / (max_ms * 1e-3), "unit": "pivots/s", "n_gpus": world,
        "steps": args.steps, "warmup": warm, "ms_per_step": max_ms / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic",
        "config": CONFIG,
        "launch": {"pivots_per_step": pivots_per_step, "ctas_per_lp": G.value, "threads_per_cta": nt.value,
                   "pivot_hash_ok": None if gold is None else True},
        "roofline": {
            # the contract's keys, read this way: `achieved` is the ALGORITHMIC rate of the rank-1
            # updates (16*R*C bytes per pivot / kernel time); the launch itself is NOT an HBM
            # measurement -- the tableau stays in L2 and K pending updates are applied per pass
            "bound": "latency/issue (L2-resident tableau, lazy rank-1 updates); the HBM-bound kernel of this path is "
                     "the column-sharded update, see sharded.roofline",
            "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
            "frac_is": "algorithmic bytes / time / measured HBM copy peak (comparable with SURVEY 8d's 16*R*C rule); "
                       "DRAM really moves `traffic` bytes per launch (dram_frac)",
            "traffic": NCU_DRAM_BYTES_PER_LAUNCH, "dram_frac": NCU_DRAM_BYTES_PER_LAUNCH / k_s / 1e9 / peak,
            "lts_throughput_pct": NCU_LTS_THROUGHPUT_PCT,
            "us_per_pivot": 1e6 * k_s / pivots_per_step,
            "floor_us_per_pivot": {"eager_L2_pass": 3.05, "source": "tools/micro_update.cu: one 16.8 MB A->B rank-1 pass "
                                   "out of L2 (an eager kernel cannot go below it; the lazy loop pays it once per K pivots)"},
            "peak_source": peak_src,
            "traffic_source": "dram__bytes_read.sum + dram__bytes_write.sum and lts__throughput of this launch, ncu --set full, "
                              + NCU_SOURCE,
            "kernel": "simplex_ps_kernel (one cooperative launch = whole solve; lazy rank-1 updates, K pivots per tableau pass)",
            "algorithmic_bytes_per_pivot": 16 * (M + 1) * (N_COLS + 1),
            "kernel_ms_per_launch": kern_ms / args.steps,
            "frac_of_8TBs": achieved / 8000.0},
        "e2e": {"value": e_tot_piv / e_max_t, "unit": "pivots/s", "h2d_bytes_per_step": h2d,
                "d2h_bytes_per_step": d2h, "api": "glp_simplex on a host glp_prob (dense build + H2D + solve + D2H)"},
        "gpu_launches": launches,
        "clocks": clk.summary(),
    }


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--ctas-per-lp", type=int, default=0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--headline-only", action="store_true", help="configs[1] only: skip the sharded / bnb legs")
    ap.add_argument("--workload", default="all", choices=["all", "dense", "batch", "shard", "bnb"],
                    help="all (default, the driver's line): configs[1] headline + the sharded and bnb legs; "
                         "dense: configs[1] alone; batch: configs[3] shape, node LPs m=256 n=512 dealt over the ranks; "
                         "shard: configs[4] alone; bnb: configs[2] alone")
    ap.add_argument("--lps", type=int, default=4096, help="batch: LPs in total")
    ap.add_argument("--distinct", type=int, default=128, help="batch: distinct seeds (the rest are clones)")
    # --rows/--cols: spellings torchrun's own parser does not mistake for its options
    ap.add_argument("--m", "--rows", dest="m", type=int, default=0)
    ap.add_argument("--n", "--cols", dest="n", type=int, default=0)
    ap.add_argument("--pivots", type=int, default=300, help="shard: pivot budget per step")
    ap.add_argument("--nodes", type=int, default=80000, help="bnb: node budget of the whole job per step")
    ap.add_argument("--cut-mode", type=int, default=0, help="bnb: 0 none, 1 reference gmi.cpp, 2 textbook GMI")
    ap.add_argument("--cut-freq", type=float, default=1.0, help="bnb: -cf, share of branched nodes that generate cuts")
    ap.add_argument("--batch-width", type=int, default=256, help="bnb: FIFO nodes per engine launch")
    ap.add_argument("--inst-seed", type=int, default=2024, help="bnb: instance seed (2024 = BASELINE configs[2])")
    ap.add_argument("--node-strategy", type=int, default=0, help="bnb: 0 FIFO (reference default), 1 best-first")
    ap.add_argument("--root-heuristic", action="store_true", help="bnb: root LP rounded down as first incumbent")
    ap.add_argument("--dist-mode", type=int, default=0, help="bnb, N > 1: 0 node-sharded with stealing, 1 round-synchronous")
    ap.add_argument("--no-speculation", action="store_true", help="bnb, best-first: one node per launch")
    ap.add_argument("--no-steal", action="store_true", help="bnb, N > 1: no node hand-over between the ranks")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        return run_reference(args, rank, world)

    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    L = capi.load_product()  # no fallback: raises when the CUDA library is missing
    if L.mvb_device_init(local) != 0:
        raise RuntimeError("CUDA engine unavailable: " + L.mvb_last_error().decode())
    L.glp_term_out(capi.GLP_OFF)
    import bench_extra as X
    if args.workload in ("batch", "shard", "bnb"):
        fn = {"batch": X.run_batch, "shard": X.run_shard, "bnb": X.run_bnb}[args.workload]
        line = fn(args, L, rank, world, local, dist, ClockSampler, peaks)
        if args.workload == "bnb" and args.cut_mode == 0 and not args.node_strategy and (args.n or 200) == 200:
            close = X.run_bnb_close(args, L, rank, world, local, dist)
            if line is not None:
                line["time_to_optimum"] = close
        if rank == 0:
            print(json.dumps(line))
        return

    line = run_dense(args, L, rank, world, local, dist)
    if args.workload == "all" and not args.headline_only:
        t_legs = time.perf_counter()
        comm = None
        if dist is not None:
            from mvolps_b200.dist import nccl_comm
            comm = nccl_comm(L, rank, world, dist)   # one NCCL communicator of the engine for every leg
        sh = X.run_shard(args, L, rank, world, local, dist, ClockSampler, peaks, steps=2, warmup=1, comm=comm)
        shn = None
        if world > 1:  # the same tableau with ncclAllGather as the per-pivot exchange (A/B, hash asserted)
            os.environ["MVOLPS_SHARD_EXCHANGE"] = "nccl"
            nargs = argparse.Namespace(**vars(args))
            nargs.pivots = 100
            shn = X.run_shard(nargs, L, rank, world, local, dist, ClockSampler, peaks, steps=1, warmup=1, comm=comm)
            os.environ.pop("MVOLPS_SHARD_EXCHANGE", None)
        X._INST.clear()
        bn = X.run_bnb(args, L, rank, world, local, dist, ClockSampler, peaks, steps=2, warmup=1, comm=comm)
        cargs = argparse.Namespace(**vars(args))
        cargs.cut_mode, cargs.nodes = 2, 20000
        bc = X.run_bnb(cargs, L, rank, world, local, dist, ClockSampler, peaks, steps=1, warmup=1, comm=comm,
                       cpu_baseline=False)
        hargs = argparse.Namespace(**vars(args))
        hargs.root_heuristic = True
        bh = X.run_bnb(hargs, L, rank, world, local, dist, ClockSampler, peaks, steps=1, warmup=1, comm=comm,
                       cpu_baseline=False)
        c3 = argparse.Namespace(**vars(args))  # configs[3] node-LP shape: frontier batches of 4096 LPs of 256 x 512
        c3.m, c3.n, c3.batch_width, c3.nodes = 256, 512, 4096, 8192
        b3 = X.run_bnb(c3, L, rank, world, local, dist, ClockSampler, peaks, steps=1, warmup=1, comm=comm,
                       cpu_baseline=False)
        close = X.run_bnb_close(args, L, rank, world, local, dist, comm=comm)
        if comm is not None:
            L.mvb_comm_free(comm)
        if line is not None:
            line["sharded"] = sh
            if shn is not None:
                line["sharded"]["nccl_allgather_check"] = {
                    "value": shn["value"], "unit": "pivots/s", "us_per_pivot": shn["us_per_pivot"], "pivots": 100,
                    "exchange": shn["config"]["exchange"], "pivot_hash_ok": shn["config"]["pivot_hash_ok"]}
            line["bnb"] = bn
            line["bnb"]["with_gmi_cuts"] = {k: bc[k] for k in ("value", "unit", "ms_per_step", "config", "gpu_launches")}
            line["bnb"]["with_root_heuristic"] = {k: bh[k] for k in ("value", "unit", "ms_per_step", "config")}
            line["bnb"]["configs3_shape"] = {k: b3[k] for k in ("value", "unit", "ms_per_step", "config", "gpu_launches")}
            line["bnb"]["time_to_optimum"] = close
            line["legs_wall_s"] = time.perf_counter() - t_legs
            line["gpu_launches_legs"] = sh["gpu_launches"] + bn["gpu_launches"] + bc["gpu_launches"] + b3["gpu_launches"]
    if rank != 0:
        return
    if not args.no_cpu_baseline and world == 1:
        v, p, dt = cpu_sample(1, 1 << 30)  # the whole solve: ~12 s on one core
        line["cpu_baseline"] = {"value": v, "unit": "pivots/s", "cores": 1, "kind": "port",
                                "sample": f"oracle (oracle/simplex_oracle.c), full solve of the same LP ({p} pivots), "
                                          f"{dt:.1f} s, 1 thread; host has {os.cpu_count()} cores"}
    print(json.dumps(line))


if __name__ == "__main__":
    try:
        main()
    finally:
        try:
            import torch.distributed as _d
            if _d.is_available() and _d.is_initialized():
                _d.destroy_process_group()
        except Exception:
            pass
