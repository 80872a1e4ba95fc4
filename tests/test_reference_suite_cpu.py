"""The reference's own unittest suite (python/tests/test_*.py, 16 tests: dynamics incl. the
LIP_matrices pickle, body, goal, restrictions, tools) run against this package with
``mpc_interface`` aliased to it -- see _run_reference_suite.py.  The files are read where
they lie (the read-only checkout, or the staged oracle/_ref).  The reference's example
formulation (biped_formulation.py) is built the same way and checked against the goldens,
receding-horizon ticks included.  Here (no GPU) the host bookkeeping is what is being
tested, on the CPU test doubles; the same suite runs on the CUDA backend in
test_gpu_reference_live.py."""
import os
import subprocess
import sys

import pytest

from oracle import make_ref, reference

HERE = os.path.dirname(os.path.abspath(__file__))


def test_staged_reference_is_unmodified():
    """oracle/_ref is a byte-for-byte copy of the reference's python/ tree: every file has
    the sha256 recorded (hashes only) in oracle/REF_MANIFEST.json."""
    path, kind = reference.find()
    if path is None:
        pytest.skip("no reference tree here")
    assert make_ref.verify(path), kind


@pytest.mark.skipif(not reference.available(), reason="no reference tree here")
def test_reference_unittests_pass_against_this_package():
    tests = os.path.join(reference.find()[0], "tests")
    done = subprocess.run([sys.executable, os.path.join(HERE, "_run_reference_suite.py"), tests],
                          capture_output=True, text=True, timeout=600)
    summary = done.stdout.strip().splitlines()[-1] if done.stdout.strip() else ""
    assert done.returncode == 0 and summary.split()[1:] == ["0", "0"], done.stderr[-3000:]
    assert int(summary.split()[0]) >= 16
    # and the reference's example script (biped_formulation.py) built on this package
    assert "example ok" in done.stdout
