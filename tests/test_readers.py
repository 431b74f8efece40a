"""CPLEX-LP and MPS readers (glp_read_lp / glp_read_mps, util.cpp:283-293) beyond the subset the
B&B tests use: ranges, free / negative / infinite bounds, objective constant,
double-sided rows, free-format MPS with OBJSENSE, RANGES and every bound type.  Each file is
solved through the shim on the CPU oracle and compared with SciPy/HiGHS on the same matrices."""
import numpy as np
import pytest
from scipy.optimize import linprog

from mvolps_b200 import capi
from mvolps_b200.lp import GlpProb

INF = float("inf")


def _solve(lib, path):
    p = GlpProb.from_file(lib, str(path))
    assert p.simplex() == 0
    return p


def test_lp_format_bounds_ranges_constant(oracle, tmp_path):
    f = tmp_path / "g.lp"
    f.write_text(
        "\\ comment line\n"
        "Minimize\n obj: 2 x1 - 3 x2 + x3 + 0.5 x4 + 10\n"
        "Subject To\n"
        " r1: x1 + x2 + x3 >= -4\n"
        " r2: - 2 <= x1 - x2 <= 6\n"
        " r3: x2 + x4 = 3\n"
        " r4: x1 + 2 x3 - x4 <= 8\n"
        "Bounds\n"
        " -5 <= x1 <= 5\n"
        " x2 <= 4\n"
        " x3 free\n"
        " x4 >= -1\n"
        " -inf <= x3 <= 7\n"
        "End\n")
    p = _solve(oracle, f)
    L, h = oracle, p.h
    assert (p.nrows, p.ncols) == (4, 4)
    assert [L.glp_get_row_type(h, i) for i in (1, 2, 3, 4)] == [capi.GLP_LO, capi.GLP_DB, capi.GLP_FX, capi.GLP_UP]
    assert (L.glp_get_row_lb(h, 2), L.glp_get_row_ub(h, 2)) == (-2.0, 6.0)
    assert [L.glp_get_col_type(h, j) for j in (1, 2, 3, 4)] == [capi.GLP_DB, capi.GLP_DB, capi.GLP_UP, capi.GLP_LO]
    assert (L.glp_get_col_lb(h, 2), L.glp_get_col_ub(h, 2)) == (0.0, 4.0)  # default lower bound 0 kept
    assert L.glp_get_obj_coef(h, 0) == 10.0
    A = np.array([[1, 1, 1, 0], [1, -1, 0, 0], [0, 1, 0, 1], [1, 0, 2, -1.0]])
    c = np.array([2, -3, 1, 0.5])
    r = linprog(c, A_ub=np.vstack([-A[0], A[1], -A[1], A[3]]), b_ub=[4, 6, 2, 8], A_eq=A[2:3], b_eq=[3],
                bounds=[(-5, 5), (0, 4), (None, 7), (-1, None)], method="highs")
    assert p.status() == capi.GLP_OPT
    assert abs(p.obj_val() - (r.fun + 10.0)) <= 1e-9 * max(1.0, abs(r.fun))


def test_lp_format_maximize_general_and_negative_upper(oracle, tmp_path):
    f = tmp_path / "h.lp"
    f.write_text(
        "Maximize\n z: 3 a + 2 b - c\n"
        "Subject To\n c1: a + b + c <= 10\n c2: a - b >= -2\n"
        "Bounds\n c <= -1\n c >= -6\n 0 <= a <= 4\n"
        "General\n a b\nEnd\n")
    p = _solve(oracle, f)
    L, h = oracle, p.h
    assert L.glp_get_obj_dir(h) == capi.GLP_MAX and L.glp_get_num_int(h) == 2
    assert (L.glp_get_col_lb(h, 3), L.glp_get_col_ub(h, 3)) == (-6.0, -1.0)
    r = linprog([-3, -2, 1], A_ub=[[1, 1, 1], [-1, 1, 0]], b_ub=[10, 2], bounds=[(0, 4), (0, None), (-6, -1)],
                method="highs")
    assert abs(p.obj_val() + r.fun) <= 1e-9 * abs(r.fun)


MPS_FREE = """NAME prob FREE
OBJSENSE
    MAX
ROWS
 N obj
 L lim1
 G lim2
 E eq1
 L rng1
 G rng2
COLUMNS
 x obj 1 lim1 1
 x lim2 1 rng1 1
 MARKER 'MARKER' 'INTORG'
 y obj 2 lim1 1
 y eq1 -1 rng2 1
 MARKER 'MARKER' 'INTEND'
 z obj -1 eq1 1
 z rng1 1 rng2 1
 w obj 0.5 lim2 1
RHS
 rhs obj -7 lim1 6
 rhs lim2 1 eq1 0.5
 rhs rng1 9 rng2 -3
RANGES
 rng rng1 4 rng2 5
BOUNDS
 UP bnd x 4
 LO bnd y -1
 UP bnd y 3
 MI bnd z
 UP bnd z 6
 FR bnd w
 UP bnd w 2
ENDATA
"""


def test_mps_free_format_objsense_ranges_bounds(oracle, tmp_path):
    f = tmp_path / "r.mps"
    f.write_text(MPS_FREE)
    p = _solve(oracle, f)
    L, h = oracle, p.h
    assert (p.nrows, p.ncols) == (5, 4)
    assert L.glp_get_obj_dir(h) == capi.GLP_MAX
    assert L.glp_get_obj_coef(h, 0) == 7.0  # RHS on the objective row = minus the constant
    # RANGES: L row -> [rhs - |R|, rhs], G row -> [rhs, rhs + |R|]
    assert (L.glp_get_row_type(h, 4), L.glp_get_row_lb(h, 4), L.glp_get_row_ub(h, 4)) == (capi.GLP_DB, 5.0, 9.0)
    assert (L.glp_get_row_type(h, 5), L.glp_get_row_lb(h, 5), L.glp_get_row_ub(h, 5)) == (capi.GLP_DB, -3.0, 2.0)
    assert L.glp_get_col_kind(h, 2) == capi.GLP_IV and L.glp_get_col_kind(h, 1) == capi.GLP_CV
    assert (L.glp_get_col_lb(h, 2), L.glp_get_col_ub(h, 2)) == (-1.0, 3.0)
    assert L.glp_get_col_type(h, 3) == capi.GLP_UP and L.glp_get_col_ub(h, 3) == 6.0
    assert L.glp_get_col_type(h, 4) == capi.GLP_UP and L.glp_get_col_ub(h, 4) == 2.0
    c = np.array([1, 2, -1, 0.5])
    A = np.array([[1, 1, 0, 0], [1, 0, 0, 1], [0, -1, 1, 0], [1, 0, 1, 0], [0, 1, 1, 0.0]])
    r = linprog(-c, A_ub=np.vstack([A[0], -A[1], A[3], -A[3], A[4], -A[4]]), b_ub=[6, -1, 9, -5, 2, 3],
                A_eq=A[2:3], b_eq=[0.5], bounds=[(0, 4), (-1, 3), (None, 6), (None, 2)], method="highs")
    assert r.status == 0 and p.status() == capi.GLP_OPT
    assert abs(p.obj_val() - (-r.fun + 7.0)) <= 1e-9 * max(1.0, abs(r.fun))


def test_mps_fixed_format_bv_li_ui_fx_pl(oracle, tmp_path):
    f = tmp_path / "b.mps"
    f.write_text(
        "NAME          BND\nROWS\n N  COST\n L  C1\n L  C2\nCOLUMNS\n"
        "    A         COST      -1.0   C1         1.0\n"
        "    B         COST      -2.0   C1         1.0\n    B         C2         1.0\n"
        "    C         COST      -1.5   C2         1.0\n"
        "    D         COST       1.0   C1        -1.0\n"
        "    E         COST      -0.5   C2         1.0\n"
        "RHS\n    RHS       C1         5.5   C2         7.0\n"
        "BOUNDS\n BV BND       A\n LI BND       B         1\n UI BND       B         4\n"
        " FX BND       C         2.5\n MI BND       D\n PL BND       D\n UP BND       D         3.0\n"
        " UP BND       E         -0.0\nENDATA\n")
    p = _solve(oracle, f)
    L, h = oracle, p.h
    assert L.glp_get_col_kind(h, 1) == capi.GLP_BV
    assert L.glp_get_col_kind(h, 2) == capi.GLP_IV and (L.glp_get_col_lb(h, 2), L.glp_get_col_ub(h, 2)) == (1.0, 4.0)
    assert L.glp_get_col_type(h, 3) == capi.GLP_FX and L.glp_get_col_lb(h, 3) == 2.5
    assert L.glp_get_col_type(h, 4) == capi.GLP_UP and L.glp_get_col_ub(h, 4) == 3.0
    r = linprog([-1, -2, -1.5, 1, -0.5], A_ub=[[1, 1, 0, -1, 0], [0, 1, 1, 0, 1]], b_ub=[5.5, 7.0],
                bounds=[(0, 1), (1, 4), (2.5, 2.5), (None, 3), (0, 0)], method="highs")
    assert p.status() == capi.GLP_OPT
    assert abs(p.obj_val() - r.fun) <= 1e-9 * abs(r.fun)


@pytest.mark.parametrize("text", [
    "Maximize\n obj: x\nSubject To\n c1: x + <= 3\nEnd\n",            # dangling operator
    "Maximize\n obj: x\nSubject To\n c1: x <= 3\nBounds\n x >= \nEnd\n",  # bound without a value
    "NAME T\nROWS\n N C\n L R1\nCOLUMNS\n X C 1 R9 1\nRHS\nENDATA\n",     # unknown row (as .mps)
])
def test_reader_rejects_malformed(oracle, tmp_path, text):
    f = tmp_path / ("bad.mps" if text.startswith("NAME") else "bad.lp")
    f.write_text(text)
    with pytest.raises(ValueError):
        GlpProb.from_file(oracle, str(f))
