"""Build check for the TMA ring kernels (gram.cu, lift.cu, conv.cu): a consumer warp that read a stage with ordinary shared-memory
loads hands it back one iteration late, behind the wait for the next stage (mbar_release_previous, common.cuh), and in the
SASS of the build under test that release must sit directly behind the wait, ahead of the loop body's first LDS and DMMA. The
builds of round 2 up to r2p released at the end of the stage's own iteration and ptxas put the arrive right behind the last LDS
issue, ahead of the last DMMAs: it could overtake loads still in flight and the producer's next TMA box landed under them (one
warp's fragment of one stage wrong in 8 of 40 passes of scripts/exp/lift_stress.py). scripts/sass_ring_check.py walks the SASS."""
import importlib.util
import os
import shutil

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.join(HERE, "..")
CSRC = os.path.join(ROOT, "ithaca-fv_b200", "csrc")


def _checker():
    spec = importlib.util.spec_from_file_location("sass_ring_check", os.path.join(ROOT, "scripts", "sass_ring_check.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


@pytest.mark.parametrize("obj,kernel,min_bodies", [
    ("lift.o", "lift_kernel", 5),            # one loop body per column-fragment count NB = 2, 3, 4, 7, 13
    ("conv.o", "conv_contract_kernel", 2),   # KT = 16 and KT = 32
    ("gram.o", "gram_kernel", 30),           # one per (rows, columns) fragment shape of gram_run
])
def test_consumer_release_sits_directly_behind_the_next_wait(obj, kernel, min_bodies):
    if not (shutil.which("cuobjdump") or os.path.exists("/usr/local/cuda/bin/cuobjdump")):
        pytest.skip("cuobjdump not available")
    path = os.path.join(CSRC, obj)
    if not os.path.exists(path):
        pytest.skip(obj + " not built (run __graft_entry__.build())")
    seen, bad = _checker().check(path, kernel)
    assert seen >= min_bodies, seen
    assert not bad, bad
