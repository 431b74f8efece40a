"""ctypes bindings of the C ABIs in include/{glpk.h,mvolps_engine.h,mvolps_bnb.h}.

`load_product()` opens mvolps_b200/lib/libglpk_b200.so (CUDA engine) and fails
loudly when it is missing -- there is no CPU fallback in the product.
`load_library(path)` binds the same symbols of any library exporting them; the
tests use it to open the test-only CPU oracle build.
"""
import ctypes as C
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PRODUCT_LIB = os.path.join(ROOT, "mvolps_b200", "lib", "libglpk_b200.so")

# ---- constants (include/glpk.h, include/mvolps_engine.h, include/mvolps_bnb.h)
GLP_MIN, GLP_MAX = 1, 2
GLP_CV, GLP_IV, GLP_BV = 1, 2, 3
GLP_FR, GLP_LO, GLP_UP, GLP_DB, GLP_FX = 1, 2, 3, 4, 5
GLP_BS, GLP_NL, GLP_NU, GLP_NF, GLP_NS = 1, 2, 3, 4, 5
GLP_UNDEF, GLP_FEAS, GLP_INFEAS, GLP_NOFEAS, GLP_OPT, GLP_UNBND = 1, 2, 3, 4, 5, 6
GLP_ON, GLP_OFF = 1, 0
GLP_MPS_FILE = 2
MVB_OPT, MVB_NOFEAS, MVB_UNBND, MVB_ITLIM, MVB_UNSOLVED = 0, 1, 2, 3, 4
MVB_AT_LB, MVB_AT_UB, MVB_AT_FR, MVB_AT_FX = 0, 1, 2, 3

GLPK_SYMBOLS = [  # the 34 symbols the reference leaves undefined
    "glp_create_prob", "glp_delete_prob", "glp_erase_prob", "glp_copy_prob",
    "glp_read_lp", "glp_read_mps", "glp_version", "glp_term_out",
    "glp_get_num_rows", "glp_get_num_cols", "glp_get_num_int",
    "glp_get_obj_dir", "glp_set_obj_dir", "glp_get_obj_coef", "glp_set_obj_coef",
    "glp_get_obj_val", "glp_add_rows", "glp_get_row_type", "glp_get_row_lb",
    "glp_get_row_ub", "glp_set_row_bnds", "glp_get_mat_row", "glp_set_mat_row",
    "glp_get_row_stat", "glp_get_col_type", "glp_get_col_lb", "glp_get_col_ub",
    "glp_set_col_bnds", "glp_get_col_kind", "glp_get_col_prim", "glp_get_col_stat",
    "glp_simplex", "glp_get_status", "glp_eval_tab_row",
]
GLPK_EXTRA_SYMBOLS = ["glp_add_cols", "glp_set_col_kind", "glp_get_row_prim", "glp_get_it_cnt", "mvb_glp_snap_tol"]
ENGINE_SYMBOLS = [
    "mvb_backend", "mvb_last_error", "mvb_device_init", "mvb_kernel_launches",
    "mvb_timer_record", "mvb_timer_elapsed_ms", "mvb_flush_l2", "mvb_last_launch", "mvb_last_kernel_ms", "mvb_last_kernel_mode", "mvb_engine_option",
    "mvb_host_staging", "mvb_lp_create", "mvb_lp_free", "mvb_lp_free_many", "mvb_lp_load", "mvb_lp_clone", "mvb_lp_assign",
    "mvb_lp_dims", "mvb_lp_set_bounds", "mvb_lp_append_row", "mvb_solve_batch", "mvb_solve_batch_begin", "mvb_solve_batch_end",
    "mvb_lp_get_primal", "mvb_lp_get_basis", "mvb_lp_get_tab_rows", "mvb_lp_get_tab_rows_batch", "mvb_lp_get_tab_rows_noalloc",
    "mvb_lp_get_tableau", "mvb_lp_get_pivot_log",
]
BNB_SYMBOLS = [
    "mvb_bnb_run", "mvb_bnb_run_dist", "mvb_bnb_steal_plan", "mvb_bnb_free", "mvb_bnb_get_stats", "mvb_bnb_tree_text",
    "mvb_bnb_solution_text", "mvb_bnb_event_text", "mvb_bnb_get_incumbent",
]
DIST_SYMBOLS = [  # include/mvolps_dist.h: exported by the product only
    "mvb_nccl_unique_id", "mvb_comm_nccl", "mvb_comm_local", "mvb_comm_free", "mvb_shard_partition",
    "mvb_incumbent_exchange",
    "mvb_shard_create", "mvb_shard_free", "mvb_shard_load", "mvb_shard_solve",
    "mvb_shard_get_primal", "mvb_shard_get_basis", "mvb_shard_get_pivot_log", "mvb_shard_last_ms", "mvb_shard_last_exchange",
    "mvb_glp_use_shards", "mvb_shard_start_mode",
]
ALL_SYMBOLS = GLPK_SYMBOLS + GLPK_EXTRA_SYMBOLS + ENGINE_SYMBOLS + BNB_SYMBOLS


class Result(C.Structure):
    _fields_ = [
        ("status", C.c_int), ("iters", C.c_int), ("pivots", C.c_int), ("flips", C.c_int),
        ("iters_primal1", C.c_int), ("iters_primal2", C.c_int), ("iters_dual", C.c_int),
        ("bland_iters", C.c_int), ("pivot_hash", C.c_ulonglong), ("obj", C.c_double),
        ("primal_feasible", C.c_int), ("reserved", C.c_int),
    ]


class Opts(C.Structure):
    _fields_ = [("max_iters", C.c_int), ("log_pivots", C.c_int),
                ("ctas_per_lp", C.c_int), ("threads", C.c_int)]


class BnbParams(C.Structure):
    _fields_ = [("node_strategy", C.c_int), ("var_strategy", C.c_int), ("cut_mode", C.c_int),
                ("batch_width", C.c_int), ("max_iters", C.c_int), ("quiet", C.c_int), ("no_steal", C.c_int),
                ("no_speculation", C.c_int), ("timed_events", C.c_int), ("total_budget", C.c_int),
                ("cut_select", C.c_int), ("root_heuristic", C.c_int), ("dist_mode", C.c_int),
                ("cut_skip", C.c_double), ("no_event_text", C.c_int)]


class BnbStats(C.Structure):
    _fields_ = [
        ("iterations", C.c_int), ("oids", C.c_int), ("lp_solves", C.c_int), ("launches", C.c_int),
        ("simplex_iters", C.c_longlong), ("has_incumbent", C.c_int), ("incumbent_oid", C.c_int),
        ("incumbent", C.c_double), ("hit_limit", C.c_int), ("event_hash", C.c_ulonglong),
        ("split_at", C.c_int), ("incumbent_rank", C.c_int), ("total_nodes", C.c_longlong),
        ("cached_solves", C.c_int), ("nodes_sent", C.c_int), ("nodes_received", C.c_int),
        ("steal_rounds", C.c_int), ("spec_nodes", C.c_int), ("spec_adopted", C.c_int),
        ("launches_aux", C.c_int), ("solved_here", C.c_int),
    ]


_P = C.c_void_p
_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int)


def _bind(lib):
    sig = {
        "glp_create_prob": (_P, []),
        "glp_delete_prob": (None, [_P]),
        "glp_erase_prob": (None, [_P]),
        "glp_copy_prob": (None, [_P, _P, C.c_int]),
        "glp_read_lp": (C.c_int, [_P, _P, C.c_char_p]),
        "glp_read_mps": (C.c_int, [_P, C.c_int, _P, C.c_char_p]),
        "glp_version": (C.c_char_p, []),
        "glp_term_out": (C.c_int, [C.c_int]),
        "glp_get_num_rows": (C.c_int, [_P]),
        "glp_get_num_cols": (C.c_int, [_P]),
        "glp_get_num_int": (C.c_int, [_P]),
        "glp_get_obj_dir": (C.c_int, [_P]),
        "glp_set_obj_dir": (None, [_P, C.c_int]),
        "glp_get_obj_coef": (C.c_double, [_P, C.c_int]),
        "glp_set_obj_coef": (None, [_P, C.c_int, C.c_double]),
        "glp_get_obj_val": (C.c_double, [_P]),
        "glp_add_rows": (C.c_int, [_P, C.c_int]),
        "glp_add_cols": (C.c_int, [_P, C.c_int]),
        "glp_get_row_type": (C.c_int, [_P, C.c_int]),
        "glp_get_row_lb": (C.c_double, [_P, C.c_int]),
        "glp_get_row_ub": (C.c_double, [_P, C.c_int]),
        "glp_set_row_bnds": (None, [_P, C.c_int, C.c_int, C.c_double, C.c_double]),
        "glp_get_mat_row": (C.c_int, [_P, C.c_int, _ip, _dp]),
        "glp_set_mat_row": (None, [_P, C.c_int, C.c_int, _ip, _dp]),
        "glp_get_row_stat": (C.c_int, [_P, C.c_int]),
        "glp_get_row_prim": (C.c_double, [_P, C.c_int]),
        "glp_get_col_type": (C.c_int, [_P, C.c_int]),
        "glp_get_col_lb": (C.c_double, [_P, C.c_int]),
        "glp_get_col_ub": (C.c_double, [_P, C.c_int]),
        "glp_set_col_bnds": (None, [_P, C.c_int, C.c_int, C.c_double, C.c_double]),
        "glp_get_col_kind": (C.c_int, [_P, C.c_int]),
        "glp_set_col_kind": (None, [_P, C.c_int, C.c_int]),
        "glp_get_col_prim": (C.c_double, [_P, C.c_int]),
        "glp_get_col_stat": (C.c_int, [_P, C.c_int]),
        "glp_simplex": (C.c_int, [_P, _P]),
        "glp_get_status": (C.c_int, [_P]),
        "glp_get_it_cnt": (C.c_int, [_P]),
        "mvb_glp_snap_tol": (C.c_double, [C.c_double]),
        "glp_eval_tab_row": (C.c_int, [_P, C.c_int, _ip, _dp]),
        "mvb_backend": (C.c_char_p, []),
        "mvb_last_error": (C.c_char_p, []),
        "mvb_device_init": (C.c_int, [C.c_int]),
        "mvb_kernel_launches": (C.c_ulonglong, []),
        "mvb_timer_record": (C.c_int, [C.c_int]),
        "mvb_timer_elapsed_ms": (C.c_double, [C.c_int, C.c_int]),
        "mvb_flush_l2": (C.c_int, []),
        "mvb_last_kernel_ms": (C.c_double, []),
        "mvb_last_kernel_mode": (C.c_int, []),
        "mvb_engine_option": (C.c_int, [C.c_char_p, C.c_int]),
        "mvb_last_launch": (C.c_int, [_ip, _ip, _ip]),
        "mvb_host_staging": (_P, [C.c_size_t]),
        "mvb_lp_free_many": (None, [C.POINTER(_P), C.c_int]),
        "mvb_lp_create": (_P, [C.c_int, C.c_int, C.c_int]),
        "mvb_lp_free": (None, [_P]),
        "mvb_lp_load": (C.c_int, [_P, _dp, _dp, _dp, _dp]),
        "mvb_lp_clone": (_P, [_P]),
        "mvb_lp_assign": (C.c_int, [_P, _P]),
        "mvb_lp_dims": (C.c_int, [_P, _ip, _ip]),
        "mvb_lp_set_bounds": (C.c_int, [_P, C.c_int, C.c_double, C.c_double]),
        "mvb_lp_append_row": (C.c_int, [_P, _dp, C.c_double, C.c_double]),
        "mvb_solve_batch": (C.c_int, [C.POINTER(_P), C.c_int, C.POINTER(Opts), C.POINTER(Result)]),
        "mvb_solve_batch_begin": (C.c_int, [C.POINTER(_P), C.c_int, C.POINTER(Opts)]),
        "mvb_solve_batch_end": (C.c_int, [C.POINTER(Result)]),
        "mvb_lp_get_primal": (C.c_int, [_P, _dp]),
        "mvb_lp_get_basis": (C.c_int, [_P, _ip, _ip, _ip]),
        "mvb_lp_get_tab_rows": (C.c_int, [_P, _ip, C.c_int, _dp]),
        "mvb_lp_get_tab_rows_batch": (C.c_int, [C.POINTER(_P), C.c_int, _ip, _ip, _dp]),
        "mvb_lp_get_tab_rows_noalloc": (C.c_int, [_P, _ip, C.c_int, _dp]),
        "mvb_lp_get_tableau": (C.c_int, [_P, _dp, _dp]),
        "mvb_lp_get_pivot_log": (C.c_int, [_P, _ip, C.c_int]),
        "mvb_bnb_run": (_P, [_P, C.POINTER(BnbParams)]),
        "mvb_bnb_run_dist": (_P, [_P, C.POINTER(BnbParams), _P]),
        "mvb_bnb_steal_plan": (C.c_int, [_dp, C.c_int, _ip, _ip, _ip, C.c_int]),
        "mvb_bnb_free": (None, [_P]),
        "mvb_bnb_get_stats": (C.POINTER(BnbStats), [_P]),
        "mvb_bnb_tree_text": (C.c_char_p, [_P]),
        "mvb_bnb_solution_text": (C.c_char_p, [_P]),
        "mvb_bnb_event_text": (C.c_char_p, [_P]),
        "mvb_bnb_get_incumbent": (C.c_int, [_P, _dp, C.c_int]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(lib, name)  # AttributeError = missing export
        fn.restype = res
        fn.argtypes = args
    dist = {
        "mvb_nccl_unique_id": (C.c_int, [C.c_void_p]),
        "mvb_comm_nccl": (_P, [C.c_void_p, C.c_int, C.c_int]),
        "mvb_comm_local": (_P, [C.c_int, C.c_int]),
        "mvb_comm_free": (None, [_P]),
        "mvb_incumbent_exchange": (C.c_int, [_P, _dp, _ip]),
        "mvb_shard_partition": (None, [C.c_int, C.c_int, C.c_int, _ip, _ip]),
        "mvb_shard_create": (_P, [C.c_int, C.c_int, C.c_int, C.c_int, _P]),
        "mvb_shard_free": (None, [_P]),
        "mvb_shard_load": (C.c_int, [_P, _dp, _dp, _dp, _dp]),
        "mvb_shard_solve": (C.c_int, [C.POINTER(_P), C.c_int, C.POINTER(Opts), C.POINTER(Result)]),
        "mvb_shard_get_primal": (C.c_int, [_P, _dp]),
        "mvb_shard_get_pivot_log": (C.c_int, [_P, _ip, C.c_int]),
        "mvb_shard_last_ms": (C.c_double, [_P]),
        "mvb_shard_last_exchange": (C.c_int, [_P]),
        "mvb_shard_get_basis": (C.c_int, [_P, _ip, _ip, _ip]),
        "mvb_glp_use_shards": (C.c_int, [_P, C.c_int, C.c_longlong]),
        "mvb_shard_start_mode": (C.c_int, [C.c_int, C.c_int, _dp, _dp, _dp]),
    }
    if hasattr(lib, "mvb_shard_create"):  # the oracle build has no multi-GPU layer
        for name, (res, args) in dist.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
    return lib


ALLGATHER_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p)
ALLREDUCE_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, _dp, C.c_int)
DESTROY_FN = C.CFUNCTYPE(None, C.c_void_p)
SENDRECV_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_void_p, C.c_size_t, C.c_int)
ALLGATHER_HOST_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t)


class Comm(C.Structure):
    """struct mvb_comm (include/mvolps_dist.h): a caller-supplied communicator."""
    _fields_ = [("ctx", C.c_void_p), ("rank", C.c_int), ("world", C.c_int),
                ("allgather_dev", ALLGATHER_FN), ("allreduce_max", ALLREDUCE_FN), ("destroy", DESTROY_FN),
                ("sendrecv_host", SENDRECV_FN), ("allgather_host", ALLGATHER_HOST_FN)]


def load_library(path):
    if not os.path.exists(path):
        raise FileNotFoundError(path)
    return _bind(C.CDLL(path, mode=os.RTLD_LOCAL | os.RTLD_NOW))


def load_product():
    """The CUDA engine.  No fallback: a missing extension is an error."""
    alt = os.environ.get("MVOLPS_LIB")  # A/B timing of another BUILD of the same CUDA library
    if alt:
        return load_library(alt)
    if not os.path.exists(PRODUCT_LIB):
        raise RuntimeError(
            f"{PRODUCT_LIB} is missing: build it with `python __graft_entry__.py` "
            "(there is no CPU fallback in the product path)")
    return load_library(PRODUCT_LIB)


def dptr(a):
    return a.ctypes.data_as(_dp)


def iptr(a):
    return a.ctypes.data_as(_ip)
