"""Synthetic instances of the BASELINE.md configs (seeds recorded there)."""
import numpy as np

INF = float("inf")


def dense_lp(m, n, seed):
    """cfg2/cfg4/cfg5 distribution: max c.x, A x <= b, x >= 0; slack basis feasible.
    A~U[0,1), b = 0.25*rowsum*(1+U[0,0.1)), c~U[0.5,1.5)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    A = rng.random((m, n))
    b = 0.25 * A.sum(axis=1) * (1.0 + 0.1 * rng.random(m))
    c = 0.5 + rng.random(n)
    return dict(A=A, c=c, sense="max", row_lb=np.full(m, -INF), row_ub=b,
                col_lb=np.zeros(n), col_ub=np.full(n, INF), integer=np.zeros(n, bool))


def mdkp(n, m, seed, U=3):
    """cfg3: multi-dimensional knapsack, integer weights/profits in 1..99,
    0 <= x <= U general integer, capacity floor(0.25*U*rowsum)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    W = rng.integers(1, 100, size=(m, n)).astype(np.float64)
    p = rng.integers(1, 100, size=n).astype(np.float64)
    cap = np.floor(0.25 * U * W.sum(axis=1))
    return dict(A=W, c=p, sense="max", row_lb=np.full(m, -INF), row_ub=cap,
                col_lb=np.zeros(n), col_ub=np.full(n, float(U)), integer=np.ones(n, bool))


def covering_lp(m, n, seed, frac_le=0.25, boxed=0.3):
    """min c.x, covering rows a.x >= b mixed with some packing rows a.x <= b', 0 <= x (<= u), c > 0:
    the slack basis is dual feasible but not primal feasible, so the solve is the dual simplex from
    the first iteration (the start a B&B child has after a bound change).  Feasible by construction
    around a known point x0."""
    rng = np.random.Generator(np.random.PCG64(seed))
    A = np.round(rng.random((m, n)) * 8) / 8
    x0 = rng.integers(0, 3, size=n).astype(np.float64)
    act = A @ x0
    row_lb = np.full(m, -INF)
    row_ub = np.full(m, INF)
    kind = rng.random(m)
    for i in range(m):
        if kind[i] < frac_le:
            row_ub[i] = act[i] + float(rng.integers(0, 4))
        else:
            row_lb[i] = max(act[i] - float(rng.integers(0, 3)), 0.125)
    col_ub = np.where(rng.random(n) < boxed, x0 + rng.integers(1, 3, size=n), INF)
    c = 0.5 + rng.random(n)
    return dict(A=A, c=c, sense="min", row_lb=row_lb, row_ub=row_ub,
                col_lb=np.zeros(n), col_ub=col_ub, integer=np.zeros(n, bool))


def general_lp(m, n, seed, frac_ge=0.3, frac_eq=0.1, boxed=0.5):
    """Mixed-sense feasible LP exercising phase 1 / dual / bound flips: rows of
    type <=, >=, = built around a known interior point so the LP is feasible."""
    rng = np.random.Generator(np.random.PCG64(seed))
    A = np.round(rng.normal(size=(m, n)) * 4) / 4
    x0 = rng.integers(0, 4, size=n).astype(np.float64)
    act = A @ x0
    row_lb = np.full(m, -INF)
    row_ub = np.full(m, INF)
    kind = rng.random(m)
    for i in range(m):
        slack = float(rng.integers(0, 5))
        if kind[i] < frac_eq:
            row_lb[i] = row_ub[i] = act[i]
        elif kind[i] < frac_eq + frac_ge:
            row_lb[i] = act[i] - slack
        else:
            row_ub[i] = act[i] + slack
    col_lb = np.zeros(n)
    col_ub = np.where(rng.random(n) < boxed, x0 + rng.integers(1, 4, size=n), INF)
    c = np.round(rng.normal(size=n) * 8) / 8
    # keep it bounded: box everything that could run away
    col_ub = np.where(np.isinf(col_ub), x0 + 10.0, col_ub)
    return dict(A=A, c=c, sense="max", row_lb=row_lb, row_ub=row_ub,
                col_lb=col_lb, col_ub=col_ub, integer=np.zeros(n, bool))


def write_lp(inst, path):
    """CPLEX-LP text of an instance (used to feed the CLI / reference driver)."""
    A, c = inst["A"], inst["c"]
    m, n = A.shape

    def expr(row):
        t = []
        for j, v in enumerate(row):
            if v != 0:
                t.append(f"{'+' if v >= 0 else '-'} {float(abs(v))!r} x{j + 1}")
        return " ".join(t) if t else "0 x1"

    with open(path, "w") as f:
        f.write("Maximize\n" if inst["sense"] == "max" else "Minimize\n")
        f.write(" obj: " + expr(c) + "\nSubject To\n")
        for i in range(m):
            lo, hi = float(inst["row_lb"][i]), float(inst["row_ub"][i])
            if lo == hi:
                f.write(f" c{i + 1}: {expr(A[i])} = {lo!r}\n")
            elif np.isfinite(hi):
                f.write(f" c{i + 1}: {expr(A[i])} <= {hi!r}\n")
                if np.isfinite(lo):
                    f.write(f" c{i + 1}lo: {expr(A[i])} >= {lo!r}\n")
            elif np.isfinite(lo):
                f.write(f" c{i + 1}: {expr(A[i])} >= {lo!r}\n")
        f.write("Bounds\n")
        for j in range(n):
            lo, hi = inst["col_lb"][j], inst["col_ub"][j]
            los = "-inf" if np.isneginf(lo) else repr(float(lo))
            his = "+inf" if np.isposinf(hi) else repr(float(hi))
            f.write(f" {los} <= x{j + 1} <= {his}\n")
        ints = [f"x{j + 1}" for j in range(n) if inst["integer"][j]]
        if ints:
            f.write("General\n " + " ".join(ints) + "\n")
        f.write("End\n")
