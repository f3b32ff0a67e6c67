import importlib
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")
    # make sure the product library and the checkers exist (built in-tree; no-op when fresh)
    lib = os.path.join(ROOT, "ithaca-fv_b200", "lib", "libithaca_b200.so")
    fvc = os.path.join(ROOT, "oracle", "libfvc_oracle.so")
    ref = os.path.join(ROOT, "oracle", "_ref", "libpod_ref.so")
    need_ref = (not os.path.exists(ref)) and os.path.isdir("/root/reference/src/thirdparty/Eigen")
    if not os.path.exists(lib) or need_ref or (os.path.exists(os.path.join(ROOT, "oracle", "fvc_oracle.c")) and not os.path.exists(fvc)):
        sys.path.insert(0, ROOT)
        import __graft_entry__ as g
        g.build()


@pytest.fixture(scope="session")
def pkg():
    return importlib.import_module("ithaca-fv_b200")


@pytest.fixture(scope="session")
def pod(pkg):
    p = pkg.Pod(0)
    yield p
    p.close()
