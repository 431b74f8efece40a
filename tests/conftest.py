import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from mvolps_b200 import capi  # noqa: E402

ORACLE_LIB = os.path.join(ROOT, "oracle", "_build", "libglpk_oracle.so")
REF_BIN = os.path.join(ROOT, "oracle", "_ref", "mvolps_ref_oracle")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def oracle():
    """The CPU oracle (test infrastructure).  Built on demand where gcc is present."""
    if not os.path.exists(ORACLE_LIB):
        subprocess.run(["make", "-C", os.path.join(ROOT, "oracle"), "all"], check=True,
                       capture_output=True)
    lib = capi.load_library(ORACLE_LIB)
    lib.glp_term_out(capi.GLP_OFF)
    return lib


@pytest.fixture(scope="session")
def product():
    """The CUDA engine; a missing library or device is a failure, not a skip."""
    lib = capi.load_product()
    if lib.mvb_device_init(0) != 0:
        pytest.fail("CUDA engine unavailable: " + lib.mvb_last_error().decode())
    lib.glp_term_out(capi.GLP_OFF)
    return lib


@pytest.fixture(scope="session")
def ref_bin():
    if not os.path.exists(REF_BIN):
        if not os.path.isdir("/root/reference"):
            pytest.skip("reference driver binary not built and /root/reference absent")
        subprocess.run(["make", "-C", os.path.join(ROOT, "oracle"), "ref"], check=True,
                       capture_output=True)
    return REF_BIN
