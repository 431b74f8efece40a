"""Multi-process parity of the real NCCL / CUDA-IPC paths (not collected by pytest: it needs one
process per GPU).  Run on a box with N >= 2 GPUs:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
        --master-port 29511 tests/dist_parity_nccl.py

Checks, bit for bit against the CPU oracle (test infrastructure) on every rank:
  * column-sharded tableau, one process per GPU: the pushed exchange (price kernel writes into the
    peers' cudaIpc mailboxes, atomicAdd_system / ld.acquire.sys hand-shake) AND ncclAllGather
    (MVOLPS_SHARD_EXCHANGE=nccl): pivot log, hash, objective bits, primal vector; primal
    phase 2 and the dual simplex
  * node-sharded B&B (NCCL incumbent max-reduction, ncclSend/Recv work stealing): serial optimum and x
  * round-synchronous B&B (MVB_DIST_SYNC, results by ncclAllGather): tree text, event stream, counters
    of the serial run, without and with GMI cut rows
Prints one JSON line on rank 0 and exits non-zero on any mismatch.
"""
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mvolps_b200 import capi, instances  # noqa: E402
from mvolps_b200.dist import ShardGroup, nccl_comm  # noqa: E402
from mvolps_b200.lp import EngineLP, GlpProb  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    L = capi.load_product()
    assert L.mvb_device_init(local) == 0, L.mvb_last_error()
    L.glp_term_out(capi.GLP_OFF)
    orc = capi.load_library(os.path.join(ROOT, "oracle", "_build", "libglpk_oracle.so"))
    orc.glp_term_out(capi.GLP_OFF)
    comm = nccl_comm(L, rank, world, dist)
    report = {"world": world, "shard": [], "bnb_steal": [], "bnb_sync": []}

    # primal phase 2 (dense_lp) and the dual simplex (covering_lp: dual feasible slack basis)
    for gen, m, n, seed in (("dense_lp", 33, 70, 3), ("dense_lp", 64, 128, 4), ("dense_lp", 128, 256, 5),
                            ("dense_lp", 200, 520, 8), ("dense_lp", 256, 512, 0),
                            ("covering_lp", 33, 70, 3), ("covering_lp", 128, 256, 5), ("covering_lp", 200, 520, 8)):
        make = getattr(instances, gen)
        a = EngineLP.from_instance(orc, make(m, n, seed))
        ra = a.solve(log_pivots=1)
        la, _ = a.pivot_log()
        assert (ra["iters_dual"] == ra["iters"]) == (gen == "covering_lp")
        for mode in ("push", "nccl"):
            if mode == "nccl":
                os.environ["MVOLPS_SHARD_EXCHANGE"] = "nccl"
            else:
                os.environ.pop("MVOLPS_SHARD_EXCHANGE", None)
            g = ShardGroup(L, make(m, n, seed), world, comm=comm, rank=rank)
            rg = g.solve(log_pivots=1)
            how = g.last_exchange()
            for k in ("status", "iters", "pivots", "flips", "pivot_hash", "bland_iters", "primal_feasible", "iters_dual"):
                assert ra[k] == rg[k], (gen, m, n, seed, mode, k, ra[k], rg[k])
            assert ra["obj"] == rg["obj"]
            lg, _ = g.pivot_log()
            assert np.array_equal(la, lg) and np.array_equal(a.primal(), g.primal())
            assert ("ncclAllGather" in how) == (mode == "nccl"), how
            report["shard"].append({"lp": gen, "m": m, "n": n, "iters": rg["iters"], "dual": rg["iters_dual"],
                                    "exchange": how.split(" (")[0]})
            g.free()
    os.environ.pop("MVOLPS_SHARD_EXCHANGE", None)

    for args, kw in (((15, 5, 2), {}), ((20, 5, 3), dict(cut_mode=2)), ((30, 8, 4), dict(batch_width=64, max_iters=0))):
        inst = instances.mdkp(*args)
        serial = GlpProb.from_instance(orc, inst).bnb(**kw)
        out = GlpProb.from_instance(L, inst).bnb(comm=comm, **kw)
        assert out["has_incumbent"] and out["incumbent"] == serial["incumbent"], (args, out["incumbent"], serial["incumbent"])
        x = torch.tensor(out["x"], device="cuda")
        xs = [torch.empty_like(x) for _ in range(world)]
        dist.all_gather(xs, x)
        assert all(torch.equal(xs[0], t) for t in xs)
        report["bnb_steal"].append({"inst": list(args), "incumbent": out["incumbent"], "total_nodes": out["total_nodes"],
                                    "serial_nodes": serial["iterations"], "nodes_sent_rank0": out["nodes_sent"],
                                    "steal_rounds": out["steal_rounds"]})

    for args, kw in (((12, 4, 1), dict(batch_width=16)), ((15, 5, 2), dict(batch_width=64, max_iters=1500)),
                     ((14, 4, 4), dict(batch_width=8, cut_mode=2, max_iters=600)),
                     ((20, 5, 3), dict(node_strategy=1, var_strategy=2, batch_width=8)),
                     ((60, 15, 2024), dict(batch_width=256, max_iters=2000))):
        inst = instances.mdkp(*args)
        serial = GlpProb.from_instance(orc, inst).bnb(**kw)
        out = GlpProb.from_instance(L, inst).bnb(comm=comm, dist_mode=1, **kw)
        for key in ("tree", "events", "event_hash", "solution", "iterations", "oids", "incumbent", "lp_solves",
                    "cached_solves", "simplex_iters", "launches"):
            assert out[key] == serial[key], (args, kw, key)
        report["bnb_sync"].append({"inst": list(args), "nodes": out["iterations"], "lp_solves": out["lp_solves"],
                                   "solved_on_rank0": out["solved_here"], "event_hash": str(out["event_hash"])})
    L.mvb_comm_free(comm)
    dist.barrier()
    if rank == 0:
        report["ok"] = True
        print(json.dumps(report))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
