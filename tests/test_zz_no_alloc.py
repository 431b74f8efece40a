"""mvb_lp_get_tab_rows_noalloc (include/mvolps_engine.h): the row gather behind glp_eval_tab_row
with its staging in a static device array instead of engine-grown buffers, for callers that may
not allocate.  Checked through the C ABI against mvb_lp_get_tab_rows, bit for bit (a gather moves
doubles, it computes nothing), on calls in a row that share the scratch.  Named test_zz_* so that
pytest collects it last."""
import numpy as np
import pytest

from mvolps_b200 import capi, instances
from mvolps_b200.lp import EngineLP

pytestmark = pytest.mark.gpu

SCRATCH_BYTES = 1 << 20  # MVB_NOALLOC_SCRATCH_BYTES


def _rows_noalloc(lib, lp, rows):
    rows = np.ascontiguousarray(rows, dtype=np.int32)
    out = np.full((len(rows), lp.n), np.nan)
    rc = lib.mvb_lp_get_tab_rows_noalloc(lp.h, capi.iptr(rows), len(rows), capi.dptr(out))
    return rc, out


def test_no_alloc_row_gather_matches_the_allocating_entry_point(product):
    a = EngineLP.from_instance(product, instances.dense_lp(64, 128, 4))
    b = EngineLP.from_instance(product, instances.mdkp(40, 10, 21))
    assert a.solve()["status"] == 0 and b.solve()["status"] == 0
    # first call: every row of the larger LP, in a scrambled order
    rows_a = np.random.default_rng(7).permutation(a.m + 1)
    rc, got = _rows_noalloc(product, a, rows_a)
    assert rc == 0, product.mvb_last_error().decode()
    assert np.array_equal(got, a.tab_rows(rows_a))
    # second call straight after it: another LP, fewer and repeated rows -- whatever the first call
    # left in the shared scratch must not show
    rows_b = np.array([0, b.m, 3, 3, 1], np.int32)
    rc, got = _rows_noalloc(product, b, rows_b)
    assert rc == 0, product.mvb_last_error().decode()
    assert np.array_equal(got, b.tab_rows(rows_b))
    # and the first LP again
    rc, got = _rows_noalloc(product, a, rows_a[:5])
    assert rc == 0, product.mvb_last_error().decode()
    assert np.array_equal(got, a.tab_rows(rows_a[:5]))
    T, _ = a.tableau()
    assert np.array_equal(got, T[rows_a[:5]])


def test_no_alloc_row_gather_rejects_what_does_not_fit(product):
    lp = EngineLP.from_instance(product, instances.dense_lp(300, 512, 3))
    # unsolved LP: the slack-basis tableau [c; A] is on the device, rows are readable
    per_row = 8 * lp.n
    fits = (SCRATCH_BYTES - 1024) // per_row          # 1024 = align256(4 * nrows) for nrows <= 256
    assert 0 < fits < lp.m + 1
    rows = np.arange(fits, dtype=np.int32)
    rc, got = _rows_noalloc(product, lp, rows)
    assert rc == 0, product.mvb_last_error().decode()
    assert np.array_equal(got, lp.tab_rows(rows))
    rc, _ = _rows_noalloc(product, lp, np.arange(fits + 1, dtype=np.int32))   # one row too many
    assert rc != 0
    msg = product.mvb_last_error().decode()
    assert "MVB_NOALLOC_SCRATCH_BYTES" in msg and str(SCRATCH_BYTES) in msg
    rc, _ = _rows_noalloc(product, lp, np.array([lp.m + 1], np.int32))        # row out of range
    assert rc != 0 and "out of range" in product.mvb_last_error().decode()
    # an empty request is a no-op
    assert product.mvb_lp_get_tab_rows_noalloc(lp.h, None, 0, None) == 0


def test_no_alloc_row_gather_refuses_while_jobs_are_queued(product):
    """A queued clone would have to be flushed first and the flush may allocate: clear error, and
    the call works again once the queue has been carried out by an allocating entry point."""
    lp = EngineLP.from_instance(product, instances.mdkp(40, 10, 22))
    assert lp.solve()["status"] == 0
    h = product.mvb_lp_clone(lp.h)
    assert h
    try:
        rows = np.array([1, 2], np.int32)
        rc, _ = _rows_noalloc(product, lp, rows)
        if rc != 0:  # the clone is still queued
            assert "queued" in product.mvb_last_error().decode()
        want = lp.tab_rows(rows)   # flushes the queue
        rc, got = _rows_noalloc(product, lp, rows)
        assert rc == 0, product.mvb_last_error().decode()
        assert np.array_equal(got, want)
    finally:
        product.mvb_lp_free(h)
