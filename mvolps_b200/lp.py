"""Thin object wrappers over the C ABIs for the tests and the bench.

EngineLP  : one `mvb_lp` (include/mvolps_engine.h) -- tableau-level access.
GlpProb   : one `glp_prob` (include/glpk.h) -- what the reference driver holds
            (util.cpp:25-42), built from an instance dict or an .lp/.mps file.
"""
import ctypes as C

import numpy as np

from . import capi


def _inst_bounds(inst):
    A = np.ascontiguousarray(inst["A"], dtype=np.float64)
    m, n = A.shape
    lb = np.concatenate([inst["col_lb"], inst["row_lb"]]).astype(np.float64)
    ub = np.concatenate([inst["col_ub"], inst["row_ub"]]).astype(np.float64)
    cost = np.asarray(inst["c"], dtype=np.float64)
    cost = -cost if inst["sense"] == "max" else cost.copy()
    return A, m, n, np.ascontiguousarray(cost), np.ascontiguousarray(lb), np.ascontiguousarray(ub)


class EngineLP:
    def __init__(self, lib, handle, m, n):
        self.lib, self.h, self.n = lib, handle, n
        self._m0 = m

    @classmethod
    def from_instance(cls, lib, inst, row_cap=0):
        A, m, n, cost, lb, ub = _inst_bounds(inst)
        h = lib.mvb_lp_create(m, n, max(row_cap, m))
        if not h:
            raise RuntimeError(lib.mvb_last_error().decode())
        self = cls(lib, h, m, n)
        self.load(A, cost, lb, ub)
        return self

    def load(self, A, cost, lb, ub):
        if self.lib.mvb_lp_load(self.h, capi.dptr(A), capi.dptr(cost), capi.dptr(lb), capi.dptr(ub)):
            raise RuntimeError(self.lib.mvb_last_error().decode())

    @property
    def m(self):
        m = C.c_int()
        n = C.c_int()
        self.lib.mvb_lp_dims(self.h, C.byref(m), C.byref(n))
        return m.value

    def clone(self):
        h = self.lib.mvb_lp_clone(self.h)
        if not h:
            raise RuntimeError(self.lib.mvb_last_error().decode())
        return EngineLP(self.lib, h, self.m, self.n)

    def assign(self, src):
        if self.lib.mvb_lp_assign(self.h, src.h):
            raise RuntimeError(self.lib.mvb_last_error().decode())

    def set_bounds(self, k, lb, ub):
        if self.lib.mvb_lp_set_bounds(self.h, k, lb, ub):
            raise RuntimeError(self.lib.mvb_last_error().decode())

    def append_row(self, a, lb, ub):
        a = np.ascontiguousarray(a, dtype=np.float64)
        if self.lib.mvb_lp_append_row(self.h, capi.dptr(a), lb, ub):
            raise RuntimeError(self.lib.mvb_last_error().decode())

    def solve(self, **opts):
        return solve_batch(self.lib, [self], **opts)[0]

    def primal(self):
        x = np.zeros(self.n + self.m)
        self.lib.mvb_lp_get_primal(self.h, capi.dptr(x))
        return x

    def basis(self):
        m = self.m
        head = np.zeros(m, np.int32)
        nb = np.zeros(self.n, np.int32)
        at = np.zeros(self.n, np.int32)
        self.lib.mvb_lp_get_basis(self.h, capi.iptr(head), capi.iptr(nb), capi.iptr(at))
        return head, nb, at

    def tableau(self):
        m = self.m
        T = np.zeros((m + 1, self.n))
        beta = np.zeros(m + 1)
        self.lib.mvb_lp_get_tableau(self.h, capi.dptr(T), capi.dptr(beta))
        return T, beta

    def tab_rows(self, rows):
        rows = np.ascontiguousarray(rows, dtype=np.int32)
        out = np.zeros((len(rows), self.n))
        if self.lib.mvb_lp_get_tab_rows(self.h, capi.iptr(rows), len(rows), capi.dptr(out)):
            raise RuntimeError(self.lib.mvb_last_error().decode())
        return out

    def pivot_log(self, cap=1 << 20):
        buf = np.zeros((cap, 2), np.int32)
        total = self.lib.mvb_lp_get_pivot_log(self.h, capi.iptr(buf), cap)
        return buf[: min(total, cap)].copy(), total

    def free(self):
        if self.h:
            self.lib.mvb_lp_free(self.h)
            self.h = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


def solve_batch(lib, lps, max_iters=0, log_pivots=0, ctas_per_lp=0, threads=0):
    n = len(lps)
    arr = (C.c_void_p * n)(*[lp.h for lp in lps])
    res = (capi.Result * n)()
    opts = capi.Opts(max_iters, log_pivots, ctas_per_lp, threads)
    if lib.mvb_solve_batch(arr, n, C.byref(opts), res):
        raise RuntimeError(lib.mvb_last_error().decode())
    return [
        {f: getattr(r, f) for f, _ in capi.Result._fields_ if f != "reserved"} for r in res
    ]


class GlpProb:
    """A glp_prob built through the same calls a GLPK client would make."""

    def __init__(self, lib, handle=None):
        self.lib = lib
        self.h = handle or lib.glp_create_prob()

    @classmethod
    def from_file(cls, lib, path):
        self = cls(lib)
        p = path.encode()
        rc = lib.glp_read_mps(self.h, capi.GLP_MPS_FILE, None, p) if path.endswith(".mps") \
            else lib.glp_read_lp(self.h, None, p)
        if rc:
            raise ValueError(f"cannot read {path}")
        return self

    @classmethod
    def from_instance(cls, lib, inst):
        self = cls(lib)
        L, h = lib, self.h
        A = np.asarray(inst["A"], dtype=np.float64)
        m, n = A.shape
        L.glp_set_obj_dir(h, capi.GLP_MAX if inst["sense"] == "max" else capi.GLP_MIN)
        L.glp_add_cols(h, n)
        L.glp_add_rows(h, m)

        def typed(lo, hi):
            if lo == hi:
                return capi.GLP_FX
            if np.isfinite(lo) and np.isfinite(hi):
                return capi.GLP_DB
            if np.isfinite(lo):
                return capi.GLP_LO
            if np.isfinite(hi):
                return capi.GLP_UP
            return capi.GLP_FR

        for j in range(n):
            lo, hi = float(inst["col_lb"][j]), float(inst["col_ub"][j])
            L.glp_set_col_bnds(h, j + 1, typed(lo, hi), lo if np.isfinite(lo) else 0.0,
                               hi if np.isfinite(hi) else 0.0)
            L.glp_set_obj_coef(h, j + 1, float(inst["c"][j]))
            if inst["integer"][j]:
                L.glp_set_col_kind(h, j + 1, capi.GLP_IV)
        ind = np.arange(0, n + 1, dtype=np.int32)
        for i in range(m):
            lo, hi = float(inst["row_lb"][i]), float(inst["row_ub"][i])
            L.glp_set_row_bnds(h, i + 1, typed(lo, hi), lo if np.isfinite(lo) else 0.0,
                               hi if np.isfinite(hi) else 0.0)
            val = np.concatenate([[0.0], A[i]])
            L.glp_set_mat_row(h, i + 1, n, capi.iptr(ind), capi.dptr(val))
        return self

    def simplex(self):
        return self.lib.glp_simplex(self.h, None)

    @property
    def nrows(self):
        return self.lib.glp_get_num_rows(self.h)

    @property
    def ncols(self):
        return self.lib.glp_get_num_cols(self.h)

    def status(self):
        return self.lib.glp_get_status(self.h)

    def obj_val(self):
        return self.lib.glp_get_obj_val(self.h)

    def col_prim(self):
        return np.array([self.lib.glp_get_col_prim(self.h, j + 1) for j in range(self.ncols)])

    def eval_tab_row(self, k):
        n = self.ncols
        ind = np.zeros(n + 1, np.int32)
        val = np.zeros(n + 1)
        ln = self.lib.glp_eval_tab_row(self.h, k, capi.iptr(ind), capi.dptr(val))
        return ind[1: ln + 1].copy(), val[1: ln + 1].copy()

    def copy(self):
        c = GlpProb(self.lib)
        self.lib.glp_copy_prob(c.h, self.h, capi.GLP_ON)
        return c

    def bnb(self, node_strategy=0, var_strategy=0, cut_mode=0, batch_width=0, max_iters=0, comm=None,
            no_steal=0, no_speculation=0, timed_events=0, total_budget=0, cut_select=0, root_heuristic=0,
            dist_mode=0, cut_freq=1.0, no_event_text=0):
        """comm: None (serial driver) or a communicator (capi.Comm / pointer) for the node-sharded run."""
        L = self.lib
        p = capi.BnbParams(node_strategy, var_strategy, cut_mode, batch_width, max_iters, 1, no_steal, no_speculation,
                           timed_events, total_budget, cut_select, root_heuristic, dist_mode, 1.0 - cut_freq,
                           no_event_text)
        if comm is None:
            hb = L.mvb_bnb_run(self.h, C.byref(p))
        else:
            hb = L.mvb_bnb_run_dist(self.h, C.byref(p), C.byref(comm) if isinstance(comm, capi.Comm) else comm)
        if not hb:
            raise RuntimeError("mvb_bnb_run failed: " + L.mvb_last_error().decode())
        st = L.mvb_bnb_get_stats(hb).contents
        out = {f: getattr(st, f) for f, _ in capi.BnbStats._fields_}
        out["tree"] = L.mvb_bnb_tree_text(hb).decode()
        out["solution"] = L.mvb_bnb_solution_text(hb).decode()
        out["events"] = L.mvb_bnb_event_text(hb).decode()
        x = np.zeros(self.ncols + 1)
        got = L.mvb_bnb_get_incumbent(hb, capi.dptr(x), len(x))
        out["x"] = x[1:].copy() if got else None
        L.mvb_bnb_free(hb)
        return out

    def free(self):
        if self.h:
            self.lib.glp_delete_prob(self.h)
            self.h = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass
