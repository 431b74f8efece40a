"""Column-sharded tableau (include/mvolps_dist.h) == single-GPU engine == oracle,
bit for bit.  The ranks are emulated as a single-process group of shards on one
GPU (device-to-device all-gather); the NCCL path shares every kernel."""
import numpy as np
import pytest

from mvolps_b200 import instances
from mvolps_b200.dist import ShardGroup
from mvolps_b200.lp import EngineLP

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("m,n,seed", [(8, 12, 2), (33, 70, 3), (64, 128, 4), (100, 37, 6), (128, 256, 5)])
@pytest.mark.parametrize("world", [1, 2, 3, 8])
def test_sharded_equals_single_and_oracle(oracle, product, m, n, seed, world):
    inst = instances.dense_lp(m, n, seed)
    a = EngineLP.from_instance(oracle, inst)
    ra = a.solve(log_pivots=1)
    b = EngineLP.from_instance(product, inst)
    rb = b.solve(log_pivots=1)
    g = ShardGroup(product, inst, world)
    rg = g.solve(log_pivots=1)
    for k in ("status", "iters", "pivots", "flips", "pivot_hash", "bland_iters", "primal_feasible"):
        assert ra[k] == rb[k] == rg[k], (k, ra[k], rb[k], rg[k])
    assert ra["obj"] == rg["obj"]  # bit-exact fp64
    la, _ = a.pivot_log()
    for k in range(world):
        lg, _ = g.pivot_log(k)
        assert np.array_equal(la, lg)
        assert np.array_equal(a.primal(), g.primal(k))
    g.free()


@pytest.mark.parametrize("m,n,seed,world", [(33, 70, 3, 2), (64, 128, 4, 3), (128, 256, 5, 8), (200, 520, 8, 8)])
def test_pushed_exchange_equals_oracle(oracle, product, monkeypatch, m, n, seed, world):
    """The fused exchange of the one-process-per-GPU path (price kernel pushes its record into every
    peer's mailbox and raises the arrival counters, decide kernel waits on them; 4-slot mailbox)
    between the shards of a single-process group: same kernels, peers = the other shards' mailboxes."""
    monkeypatch.setenv("MVOLPS_SHARD_EXCHANGE", "push")
    inst = instances.dense_lp(m, n, seed)
    a = EngineLP.from_instance(oracle, inst)
    ra = a.solve(log_pivots=1)
    g = ShardGroup(product, inst, world)
    rg = g.solve(log_pivots=1)
    assert "pushed" in g.last_exchange()
    for k in ("status", "iters", "pivots", "flips", "pivot_hash", "bland_iters", "primal_feasible"):
        assert ra[k] == rg[k], (k, ra[k], rg[k])
    assert ra["obj"] == rg["obj"]
    la, _ = a.pivot_log()
    for k in range(world):
        lg, _ = g.pivot_log(k)
        assert np.array_equal(la, lg)
        assert np.array_equal(a.primal(), g.primal(k))
    g.free()
    monkeypatch.delenv("MVOLPS_SHARD_EXCHANGE")
    g = ShardGroup(product, inst, world)
    assert g.solve()["pivot_hash"] == ra["pivot_hash"] and "host copies" in g.last_exchange()
    g.free()


def test_sharded_iteration_limit_and_resume_free(product):
    inst = instances.dense_lp(64, 128, 4)
    g = ShardGroup(product, inst, 4)
    r = g.solve(max_iters=7)
    assert r["status"] == 3 and r["iters"] == 7
    g.free()


def test_sharded_rejects_unsupported(product):
    inst = instances.general_lp(20, 30, 12)  # needs phase 1
    with pytest.raises(RuntimeError):
        ShardGroup(product, inst, 2)
