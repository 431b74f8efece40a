"""Host logic of the multi-GPU layer (include/mvolps_dist.h) on CPU: world size 2
over gloo for the exchange steps, plus the partition / slicing rules.  The
device kernels of the sharded path are covered by tests/test_gpu_shard.py."""
import ctypes as C
import os
import socket

import numpy as np
import pytest
import torch.distributed as dist
import torch.multiprocessing as mp

from mvolps_b200 import capi, instances
from mvolps_b200.dist import partition, slice_instance, torch_comm


@pytest.fixture(scope="module")
def lib():
    return capi.load_library(capi.PRODUCT_LIB)


def test_partition_matches_c(lib):
    for n in (1, 15, 16, 17, 100, 513, 2048, 32768, 32769):
        for world in (1, 2, 3, 4, 8):
            cover = []
            for r in range(world):
                c0, nc = C.c_int(), C.c_int()
                lib.mvb_shard_partition(n, world, r, C.byref(c0), C.byref(nc))
                assert (c0.value, nc.value) == partition(n, world, r)
                assert nc.value >= 0 and (nc.value == 0 or c0.value % 16 == 0)
                cover += list(range(c0.value, c0.value + nc.value))
            assert cover == list(range(n))  # contiguous, disjoint, complete


def test_slices_reassemble_instance():
    inst = instances.dense_lp(20, 70, 3)
    parts = [slice_instance(inst, 3, r) for r in range(3)]
    assert np.array_equal(np.hstack([p["A"] for p in parts]), inst["A"])
    assert np.array_equal(np.concatenate([p["cost"] for p in parts]), -inst["c"])  # max -> internal min
    for p in parts:
        assert p["lb"].shape == (70 + 20,) and np.array_equal(p["ub"][70:], inst["row_ub"])


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        lib = capi.load_library(capi.PRODUCT_LIB)
        comm = torch_comm(dist)
        out = []
        # rounds of a node-sharded tree: each rank reports the best integral node of its batch
        rounds = [(-np.inf, 1196.0), (1201.0, 1199.0), (1202.0, 1202.0), (1150.0, -np.inf)]
        best, best_owner = -np.inf, -1
        for mine in rounds:
            inc, owner = C.c_double(mine[rank]), C.c_int(-1)
            assert lib.mvb_incumbent_exchange(C.byref(comm), C.byref(inc), C.byref(owner)) == 0
            if inc.value > best:  # strict, first found kept (bs.cpp:172)
                best, best_owner = inc.value, owner.value
            out.append((inc.value, owner.value, best, best_owner))
        # the NCCL-id plumbing (dist.nccl_comm) is a 128-byte broadcast from rank 0: same call over gloo
        import torch
        t = torch.arange(128, dtype=torch.uint8) if rank == 0 else torch.zeros(128, dtype=torch.uint8)
        dist.broadcast(t, 0)
        out.append(t.tolist() == list(range(128)))
        # dealing of a FIFO batch over the ranks: round-robin, disjoint, complete
        batch = list(range(37))
        mine_lps = [k for k in batch if k % world == rank]
        counts = [None] * world
        dist.all_gather_object(counts, mine_lps)
        out.append(sorted(sum(counts, [])) == batch)
        q.put((rank, out))
    finally:
        dist.destroy_process_group()


def _bnb_worker(rank, world, port, q, oracle_lib):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from mvolps_b200.lp import GlpProb
        orc = capi.load_library(oracle_lib)  # the driver is engine-agnostic: CPU engine here
        orc.glp_term_out(capi.GLP_OFF)
        comm = torch_comm(dist)
        res = {}
        for name, args in (("mdkp_12x4_s1", (12, 4, 1)), ("mdkp_15x5_s2", (15, 5, 2))):
            p = GlpProb.from_instance(orc, instances.mdkp(*args))
            out = p.bnb(batch_width=16, comm=comm)
            res[name] = dict(incumbent=out["incumbent"], x=None if out["x"] is None else out["x"].tolist(),
                             iterations=out["iterations"], total=out["total_nodes"], split_at=out["split_at"],
                             inc_rank=out["incumbent_rank"], has=out["has_incumbent"])
        q.put((rank, res))
    finally:
        dist.destroy_process_group()


def test_node_sharded_bnb_world2_gloo():
    """mvb_bnb_run_dist on 2 ranks (gloo, CPU engine): exact search -> the serial optimum, the
    same x on both ranks, work split between the ranks after the lock-step ramp-up."""
    import json
    from mvolps_b200.lp import GlpProb
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    oracle_lib = os.path.join(root, "oracle", "_build", "libglpk_oracle.so")
    gold = {c["name"]: c for c in json.load(open(os.path.join(root, "tests", "golden", "bnb_golden.json")))["cases"]}
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_bnb_worker, args=(r, 2, port, q, oracle_lib)) for r in range(2)]
    for p in procs:
        p.start()
    res = dict(q.get(timeout=300) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    orc = capi.load_library(oracle_lib)
    orc.glp_term_out(capi.GLP_OFF)
    for name, args in (("mdkp_12x4_s1", (12, 4, 1)), ("mdkp_15x5_s2", (15, 5, 2))):
        serial = GlpProb.from_instance(orc, instances.mdkp(*args)).bnb()
        a, b = res[0][name], res[1][name]
        assert a["has"] and b["has"]
        assert a["incumbent"] == b["incumbent"] == serial["incumbent"]
        assert a["x"] == b["x"] and a["inc_rank"] == b["inc_rank"] and a["total"] == b["total"]
        assert abs(np.dot(instances.mdkp(*args)["c"], a["x"]) - a["incumbent"]) < 1e-9
        assert a["split_at"] == b["split_at"] and a["split_at"] >= 0
        # the lock-step ramp-up is replicated work and counts once
        assert a["total"] == a["split_at"] + (a["iterations"] - a["split_at"]) + (b["iterations"] - b["split_at"])
        assert 0 < a["iterations"] < a["total"] and 0 < b["iterations"] < b["total"]
        if name == "mdkp_12x4_s1":
            assert abs(a["incumbent"] - gold[name]["highs_ilp"]) < 1e-9


def _steal_worker(rank, world, port, q, oracle_lib):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from mvolps_b200.lp import GlpProb
        orc = capi.load_library(oracle_lib)
        orc.glp_term_out(capi.GLP_OFF)
        comm = torch_comm(dist)
        res = {}
        for args, cm in (((15, 5, 2), 0), ((20, 5, 3), 2)):
            for ns in (1, 0):
                out = GlpProb.from_instance(orc, instances.mdkp(*args)).bnb(batch_width=8, comm=comm, no_steal=ns,
                                                                            cut_mode=cm)
                res[(args, cm, ns)] = {k: out[k] for k in ("incumbent", "iterations", "total_nodes", "nodes_sent", "split_at",
                                                          "nodes_received", "steal_rounds", "has_incumbent")}
                res[(args, cm, ns)]["x"] = out["x"].tolist()
                res[(args, cm, ns)]["stolen_in_tree"] = out["tree"].count("\n")
        q.put((rank, res))
    finally:
        dist.destroy_process_group()


def test_work_stealing_world3_gloo():
    """Node hand-over (mvb_comm::sendrecv_host) on 3 ranks over gloo with the CPU
    engine: every node sent is received by exactly one rank, the search stays exact (serial
    optimum, with and without cut rows travelling along), and the ranks end up balanced."""
    from mvolps_b200.lp import GlpProb
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    oracle_lib = os.path.join(root, "oracle", "_build", "libglpk_oracle.so")
    world = 3
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_steal_worker, args=(r, world, port, q, oracle_lib)) for r in range(world)]
    for p in procs:
        p.start()
    res = dict(q.get(timeout=600) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    orc = capi.load_library(oracle_lib)
    orc.glp_term_out(capi.GLP_OFF)
    for args, cm in (((15, 5, 2), 0), ((20, 5, 3), 2)):
        serial = GlpProb.from_instance(orc, instances.mdkp(*args)).bnb(cut_mode=cm)
        off = [res[r][(args, cm, 1)] for r in range(world)]
        on = [res[r][(args, cm, 0)] for r in range(world)]
        for rr in off + on:
            assert rr["has_incumbent"] and rr["incumbent"] == serial["incumbent"]
        assert all(rr["x"] == on[0]["x"] for rr in on)
        assert all(rr["nodes_sent"] == rr["nodes_received"] == rr["steal_rounds"] == 0 for rr in off)
        sent, got = sum(rr["nodes_sent"] for rr in on), sum(rr["nodes_received"] for rr in on)
        assert sent == got > 0                                  # nothing lost, nothing duplicated
        assert len({rr["steal_rounds"] for rr in on}) == 1      # same matching on every rank
        its_on, its_off = [rr["iterations"] for rr in on], [rr["iterations"] for rr in off]
        assert max(its_on) - min(its_on) < max(its_off) - min(its_off)  # balanced
        res_split = on[0]["split_at"]
        assert sum(its_on) - (world - 1) * res_split == on[0]["total_nodes"]


def _plan(lib, cnt):
    cnt = np.asarray(cnt, dtype=np.float64)
    d, t, k = (np.zeros(8, np.int32) for _ in range(3))
    n = lib.mvb_bnb_steal_plan(capi.dptr(cnt), len(cnt), capi.iptr(d), capi.iptr(t), capi.iptr(k), 8)
    return [(int(d[i]), int(t[i]), int(k[i])) for i in range(n)]


def test_steal_matching_rule(lib):
    """The matching is a pure function of the open-node counts (identical on every rank): the
    poorest rank below MVB_STEAL_LOW = 4 takes half of the difference from the richest, one pair
    per rank and round, ties to the lowest rank, at most MVB_STEAL_MAX = 256 nodes, retired ranks
    (count < 0) neither give nor take."""
    assert _plan(lib, [10, 10, 10]) == []                       # nobody is short
    assert _plan(lib, [0, 100]) == [(1, 0, 50)]
    assert _plan(lib, [100, 0]) == [(0, 1, 50)]
    assert _plan(lib, [3, 4]) == []                             # difference < 2
    assert _plan(lib, [0, 1]) == []
    assert _plan(lib, [0, 2]) == [(1, 0, 1)]
    assert _plan(lib, [0, 0, 40, 20]) == [(2, 0, 20), (3, 1, 10)]  # two pairs, poorest/lowest rank first
    assert _plan(lib, [5000, 0]) == [(0, 1, 256)]               # capped
    assert _plan(lib, [-1, 0, 9]) == [(2, 1, 4)]                # rank 0 retired
    assert _plan(lib, [-1, -1, 7]) == []
    assert _plan(lib, [2, 3, 50, 1]) == [(2, 3, 24)]            # ranks 0 and 1 are short too, but 3 - 2 < 2
    # conservation and monotonicity on random counts
    rng = np.random.default_rng(0)
    for _ in range(200):
        cnt = rng.integers(-1, 600, size=int(rng.integers(2, 9))).astype(float)
        used = set()
        for d, t, k in _plan(lib, cnt):
            assert d not in used and t not in used and d != t
            used |= {d, t}
            assert cnt[t] < 4 and cnt[d] - cnt[t] >= 2 and 1 <= k <= min(256, (cnt[d] - cnt[t]) // 2)
            assert cnt[d] - k >= cnt[t] + k - 1                 # the donor stays at least as rich


def test_one_rank_world_equals_serial():
    # a 1-rank world: no partner -> no hand-over, the run equals the serial one
    from mvolps_b200.lp import GlpProb
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    orc = capi.load_library(os.path.join(root, "oracle", "_build", "libglpk_oracle.so"))
    orc.glp_term_out(capi.GLP_OFF)
    comm = capi.Comm(None, 0, 1, capi.ALLGATHER_FN(), capi.ALLREDUCE_FN(), capi.DESTROY_FN(), capi.SENDRECV_FN())
    a = GlpProb.from_instance(orc, instances.mdkp(12, 4, 1)).bnb(comm=comm)
    b = GlpProb.from_instance(orc, instances.mdkp(12, 4, 1)).bnb()
    assert a["tree"] == b["tree"] and a["nodes_sent"] == 0 and a["event_hash"] == b["event_hash"]


def test_incumbent_exchange_world2_gloo(lib):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = dict(q.get(timeout=120) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert res[0] == res[1]  # replicated decisions
    rounds = res[0][:4]
    assert rounds[0][:2] == (1196.0, 1)
    assert rounds[1][:2] == (1201.0, 0)
    assert rounds[2][:2] == (1202.0, 0)      # tie: lowest rank owns it
    assert rounds[3] == (1150.0, 0, 1202.0, 0)  # a worse round does not replace the incumbent
    assert res[0][4] is True and res[0][5] is True


def _sync_worker(rank, world, port, q, oracle_lib, cases):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from mvolps_b200.lp import GlpProb
        orc = capi.load_library(oracle_lib)
        orc.glp_term_out(capi.GLP_OFF)
        comm = torch_comm(dist)
        res = []
        for args, kw in cases:
            out = GlpProb.from_instance(orc, instances.mdkp(*args)).bnb(comm=comm, dist_mode=1, **kw)
            res.append({k: out[k] for k in ("tree", "events", "event_hash", "solution", "iterations", "oids", "incumbent",
                                            "lp_solves", "cached_solves", "simplex_iters", "solved_here", "split_at",
                                            "total_nodes", "launches")})
        q.put((rank, res))
    finally:
        dist.destroy_process_group()


SYNC_CASES = [((12, 4, 1), dict(batch_width=16)), ((15, 5, 2), dict(batch_width=64, max_iters=1500)),
              ((14, 4, 4), dict(batch_width=8, cut_mode=2, max_iters=600)),
              ((12, 4, 4), dict(batch_width=8, cut_mode=1)),
              ((20, 5, 3), dict(node_strategy=1, var_strategy=2, batch_width=8)),
              ((16, 4, 6), dict(batch_width=32, cut_mode=2, cut_freq=0.5, cut_select=1, max_iters=800))]


@pytest.mark.parametrize("world", [2, 3])
def test_round_synchronous_tree_is_the_serial_tree(world):
    """MVB_DIST_SYNC (SURVEY 8e's default mode) over gloo with the CPU engine: the LPs of a round
    are solved by their owner ranks, the solutions (and the tableau rows of the cut step) travel by
    all-gather, and every rank consumes them in the serial order -> tree text, event stream,
    solution string, oids and all counters equal the 1-process run, on every rank; the device work
    is split (solved_here)."""
    from mvolps_b200.lp import GlpProb
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    oracle_lib = os.path.join(root, "oracle", "_build", "libglpk_oracle.so")
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_sync_worker, args=(r, world, port, q, oracle_lib, SYNC_CASES)) for r in range(world)]
    for p in procs:
        p.start()
    res = dict(q.get(timeout=600) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    orc = capi.load_library(oracle_lib)
    orc.glp_term_out(capi.GLP_OFF)
    for k, (args, kw) in enumerate(SYNC_CASES):
        serial = GlpProb.from_instance(orc, instances.mdkp(*args)).bnb(**kw)
        for r in range(world):
            out = res[r][k]
            for key in ("tree", "events", "event_hash", "solution", "iterations", "oids", "incumbent", "lp_solves",
                        "cached_solves", "simplex_iters", "launches"):
                assert out[key] == serial[key], (args, kw, r, key)
            assert out["total_nodes"] == serial["iterations"] and out["split_at"] >= 0
        here = [res[r][k]["solved_here"] for r in range(world)]
        assert all(h > 0 for h in here)
        # after the ramp-up every LP is solved exactly once, by its owner
        assert max(here) < serial["lp_solves"] - serial["cached_solves"]
