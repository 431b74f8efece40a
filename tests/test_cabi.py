"""The drop-in boundary: the product library loads without a GPU, exports every
symbol include/*.h declares (and the 34 glp_* symbols the reference's objects
leave undefined), never links the oracle, and fails loudly --
no CPU fallback -- when asked to compute without a device."""
import ctypes as C
import os
import re
import subprocess

import pytest

from mvolps_b200 import capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
INCLUDE = os.path.join(ROOT, "include")


def _declared(header):
    src = open(os.path.join(INCLUDE, header)).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b((?:glp|mvb)_[a-z0-9_]+)\s*\(", src)))


def _have_product():
    return os.path.exists(capi.PRODUCT_LIB)


@pytest.fixture(scope="module")
def product_nogpu():
    if not _have_product():
        subprocess.run(["make", "-C", ROOT, "-j8", "all"], check=True, capture_output=True)
    return capi.load_library(capi.PRODUCT_LIB)


@pytest.mark.parametrize("header", ["glpk.h", "mvolps_engine.h", "mvolps_bnb.h", "mvolps_dist.h"])
def test_every_declared_symbol_is_exported(product_nogpu, header):
    if not os.path.exists(os.path.join(INCLUDE, header)):
        pytest.skip(header + " not present")
    names = _declared(header)
    assert names
    raw = C.CDLL(capi.PRODUCT_LIB)
    missing = [n for n in names if not hasattr(raw, n)]
    assert not missing, missing


def test_reference_undefined_symbols_are_covered(product_nogpu):
    assert len(capi.GLPK_SYMBOLS) == 34
    decl = set(_declared("glpk.h"))
    assert set(capi.GLPK_SYMBOLS) <= decl
    raw = C.CDLL(capi.PRODUCT_LIB)
    assert all(hasattr(raw, n) for n in capi.GLPK_SYMBOLS)


def test_product_is_cuda_backend_and_does_not_link_oracle(product_nogpu):
    assert product_nogpu.mvb_backend() == b"cuda-sm100a"
    out = subprocess.run(["ldd", capi.PRODUCT_LIB], capture_output=True, text=True).stdout
    assert "oracle" not in out
    nm = subprocess.run(["nm", "-D", "--defined-only", capi.PRODUCT_LIB], capture_output=True, text=True).stdout
    assert "simplex_kernel" in nm or "mvb_solve_batch" in nm
    # sm_100a SASS is embedded
    co = subprocess.run(["cuobjdump", "-lelf", capi.PRODUCT_LIB], capture_output=True, text=True).stdout
    assert "sm_100a" in co


def test_no_cpu_fallback_without_device(product_nogpu):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    L = product_nogpu
    assert L.mvb_device_init(0) != 0
    assert b"no CPU fallback" in L.mvb_last_error() or b"CUDA" in L.mvb_last_error()
    assert not L.mvb_lp_create(2, 2, 2)
    # the reference-facing call reports failure instead of computing on the host
    p = L.glp_create_prob()
    assert L.glp_read_lp(p, None, os.path.join(ROOT, "examples", "small3.lp").encode()) == 0
    assert L.glp_simplex(p, None) != 0
    assert L.glp_get_status(p) == capi.GLP_UNDEF
    L.glp_delete_prob(p)


def test_product_sources_do_not_reference_oracle():
    pkg = os.path.join(ROOT, "mvolps_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cpp", ".cu", ".cuh", ".hpp", ".h")):
                txt = open(os.path.join(dp, f)).read()
                assert "libglpk_oracle" not in txt and "oracle/_" not in txt, f
                assert not re.search(r'#include\s*[<"][^>"]*oracle', txt), f
                assert not re.search(r"^\s*(from|import)\s+\S*oracle", txt, flags=re.M), f


def test_shard_start_mode_is_host_logic(product_nogpu):
    """mvb_shard_start_mode (include/mvolps_dist.h): which simplex the column-sharded path would start
    with from the slack basis -- pure host logic, so it is checked here against the mode the oracle's
    own solve takes (iters_primal2 / iters_dual / iters_primal1 of solve_one)."""
    import numpy as np
    from mvolps_b200 import instances
    from mvolps_b200.dist import slice_instance
    from mvolps_b200.lp import EngineLP
    from conftest import ORACLE_LIB
    if not os.path.exists(ORACLE_LIB):
        subprocess.run(["make", "-C", os.path.join(ROOT, "oracle"), "all"], check=True, capture_output=True)
    orc = capi.load_library(ORACLE_LIB)
    L = product_nogpu
    cases = [(instances.dense_lp(8, 12, 2), 1), (instances.covering_lp(33, 70, 3), 2),
             (instances.general_lp(20, 30, 12), 3), (instances.mdkp(12, 4, 1), 1)]
    for inst, want in cases:
        sl = slice_instance(inst, 1, 0)
        got = L.mvb_shard_start_mode(sl["m"], sl["n"], capi.dptr(sl["cost"]), capi.dptr(sl["lb"]), capi.dptr(sl["ub"]))
        assert got == want
        r = EngineLP.from_instance(orc, inst).solve()
        took = 3 if r["iters_primal1"] > 0 else 1 if r["iters_primal2"] == r["iters"] else 2
        assert took == want, (r, want)
    inst = instances.dense_lp(8, 12, 2)
    inst["col_lb"] = np.full(12, 1.0)   # non-basic values would not be 0: one GPU
    sl = slice_instance(inst, 1, 0)
    assert L.mvb_shard_start_mode(sl["m"], sl["n"], capi.dptr(sl["cost"]), capi.dptr(sl["lb"]), capi.dptr(sl["ub"])) == 0
