"""Harness over include/mvolps_dist.h: column-sharded LPs and communicators.

`partition(n, world, rank)` is a pure-Python restatement of
mvb_shard_partition (used by the CPU tests and to slice instances);
ShardGroup drives either a single-process group of shards on one GPU or one
NCCL rank per process (the unique id travels through torch.distributed).
"""
import ctypes as C

import numpy as np

from . import capi

LD_ALIGN = 16


def partition(n, world, rank):
    per = (n + world - 1) // world
    per = (per + LD_ALIGN - 1) // LD_ALIGN * LD_ALIGN
    a, b = min(n, rank * per), min(n, (rank + 1) * per)
    return a, b - a


def slice_instance(inst, world, rank):
    """The data rank `rank` loads: its columns of A and c, all bounds."""
    A = np.ascontiguousarray(inst["A"], dtype=np.float64)
    m, n = A.shape
    c0, nc = partition(n, world, rank)
    cost = -np.asarray(inst["c"], np.float64) if inst["sense"] == "max" else np.asarray(inst["c"], np.float64)
    lb = np.ascontiguousarray(np.concatenate([inst["col_lb"], inst["row_lb"]]), dtype=np.float64)
    ub = np.ascontiguousarray(np.concatenate([inst["col_ub"], inst["row_ub"]]), dtype=np.float64)
    return dict(m=m, n=n, col0=c0, ncols=nc, A=np.ascontiguousarray(A[:, c0:c0 + nc]),
                cost=np.ascontiguousarray(cost[c0:c0 + nc]), lb=lb, ub=ub)


def nccl_comm(lib, rank, world, dist):
    """NCCL communicator of the engine; `dist` = an initialised torch.distributed."""
    import torch
    buf = (C.c_ubyte * 128)()
    if rank == 0 and lib.mvb_nccl_unique_id(buf):
        raise RuntimeError(lib.mvb_last_error().decode())
    t = torch.tensor(list(buf), dtype=torch.uint8)
    if dist.get_backend() == "nccl":
        t = t.cuda()
    dist.broadcast(t, 0)
    ident = (C.c_ubyte * 128)(*t.cpu().tolist())
    comm = lib.mvb_comm_nccl(ident, rank, world)
    if not comm:
        raise RuntimeError(lib.mvb_last_error().decode())
    return comm


class ShardGroup:
    """The shards this process holds: all of them (local group) or one (NCCL)."""

    def __init__(self, lib, inst, world, comm=None, rank=0, slices=None):
        """`slices`: {rank: slice_instance(inst, world, rank)} prepared by the caller (bench.py times the
        load without the host-side column slicing of the harness)."""
        self.lib, self.world = lib, world
        self.shards, self.comms = [], []
        ranks = range(world) if comm is None else [rank]
        for r in ranks:
            sl = slices[r] if slices is not None else slice_instance(inst, world, r)
            c = lib.mvb_comm_local(r, world) if comm is None else comm
            if comm is None:
                self.comms.append(c)
            h = lib.mvb_shard_create(sl["m"], sl["n"], sl["col0"], sl["ncols"], c)
            if not h:
                raise RuntimeError(lib.mvb_last_error().decode())
            if lib.mvb_shard_load(h, capi.dptr(sl["A"]), capi.dptr(sl["cost"]), capi.dptr(sl["lb"]),
                                  capi.dptr(sl["ub"])):
                raise RuntimeError(lib.mvb_last_error().decode())
            self.shards.append(h)
        self.m, self.n = sl["m"], sl["n"]

    def solve(self, max_iters=0, log_pivots=0):
        k = len(self.shards)
        arr = (C.c_void_p * k)(*self.shards)
        res = capi.Result()
        opts = capi.Opts(max_iters, log_pivots, 0, 0)
        if self.lib.mvb_shard_solve(arr, k, C.byref(opts), C.byref(res)):
            raise RuntimeError(self.lib.mvb_last_error().decode())
        return {f: getattr(res, f) for f, _ in capi.Result._fields_ if f != "reserved"}

    def primal(self, k=0):
        x = np.zeros(self.n + self.m)
        self.lib.mvb_shard_get_primal(self.shards[k], capi.dptr(x))
        return x

    def pivot_log(self, k=0, cap=1 << 20):
        buf = np.zeros((cap, 2), np.int32)
        total = self.lib.mvb_shard_get_pivot_log(self.shards[k], capi.iptr(buf), cap)
        return buf[: min(total, cap)].copy(), total

    def last_ms(self):
        return self.lib.mvb_shard_last_ms(self.shards[0])

    def last_exchange(self):
        return ["host copies (single-process group)", "ncclAllGather",
                "pushed into the peers' mailboxes by the price kernel (CUDA IPC over NVLink)"][
            self.lib.mvb_shard_last_exchange(self.shards[0])]

    def free(self):
        for h in self.shards:
            self.lib.mvb_shard_free(h)
        for c in self.comms:
            self.lib.mvb_comm_free(c)
        self.shards, self.comms = [], []


def torch_comm(dist):
    """mvb_comm whose host reductions run over an initialised torch.distributed group
    (gloo on CPU in the tests).  Keep the returned object alive while C code uses it."""
    import torch

    def allreduce_max(ctx, vals, count):
        t = torch.tensor([vals[i] for i in range(count)], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        for i in range(count):
            vals[i] = float(t[i])
        return 0

    def sendrecv_host(ctx, send, sbytes, dst, recv, rbytes, src):
        if dst >= 0:
            t = torch.frombuffer((C.c_char * sbytes).from_address(send), dtype=torch.uint8)
            dist.send(t.clone(), dst)
        if src >= 0:
            t = torch.empty(rbytes, dtype=torch.uint8)
            dist.recv(t, src)
            C.memmove(recv, t.numpy().ctypes.data, rbytes)
        return 0

    def allgather_host(ctx, send, recv, nbytes):
        t = torch.frombuffer((C.c_char * nbytes).from_address(send), dtype=torch.uint8).clone()
        outs = [torch.empty(nbytes, dtype=torch.uint8) for _ in range(dist.get_world_size())]
        dist.all_gather(outs, t)
        for r, o in enumerate(outs):
            C.memmove(recv + r * nbytes, o.numpy().ctypes.data, nbytes)
        return 0

    cb = capi.ALLREDUCE_FN(allreduce_max)
    sr = capi.SENDRECV_FN(sendrecv_host)
    ag = capi.ALLGATHER_HOST_FN(allgather_host)
    comm = capi.Comm(None, dist.get_rank(), dist.get_world_size(), capi.ALLGATHER_FN(), cb, capi.DESTROY_FN(), sr, ag)
    comm._keep = (cb, sr, ag)
    return comm
