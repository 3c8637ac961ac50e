"""The repo's own memcheck (-m gpu). compute-sanitizer is closed on the GPU pool ("runs under it have left GPUs needing a
reset"), so the small configurations of every kernel family (profiles/sanitize.py: both BVH builders, wide-node emission,
staged and fused renders on two streams, tail kernel, ray re-sorting, the deep adversarial scene, trace batches, film
resolve, scatter entry) run through libprt_b200_checked.so - the same sources with -DPRT_CHECKED=1, where every
data-derived index is tested before the access - and no check may fire; then the stack limit is lowered on purpose and
the check MUST fire (the checker is alive)."""
import os
import re
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = pytest.mark.gpu


def test_checked_library_runs_clean_and_catches_a_provoked_overflow():
    env = dict(os.environ, PRT_CHECKED="1")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "profiles", "sanitize.py")], capture_output=True, text=True, env=env, timeout=900)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    m = re.search(r"check_flags 0x([0-9a-f]+) provoked 0x([0-9a-f]+)", r.stdout)
    assert m, r.stdout[-1000:]
    flags, provoked = int(m.group(1), 16), int(m.group(2), 16)
    assert flags == 0x40000000, "a bounds check fired: 0x%x" % flags          # bit 30 = the checked library really ran
    assert provoked == 0x40000001, "PRT_CHK_STACK did not fire with a 28-entry stack: 0x%x" % provoked


def test_non_finite_vertices_are_harmless():
    """NaN / +-inf vertices (a broken exporter, an uninitialised buffer): the build and every walk terminate, no index leaves its
    array, such triangles are never hit and the rest of the scene traces exactly as the list loop does."""
    env = dict(os.environ, PRT_CHECKED="1", PYTHONPATH=os.pathsep.join([ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "oracle")]))
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tests", "nonfinite_check.py")], capture_output=True, text=True, env=env, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "nonfinite ok 1 check_flags 0x40000000" in r.stdout, r.stdout[-1000:]


def test_regular_library_has_no_checks():
    if os.environ.get("PRT_CHECKED") == "1":
        pytest.skip("this run IS the checked library (PRT_CHECKED=1)")
    from parallelraytracing_b200.capi import Context
    with Context(0) as c:
        assert c.check_flags() == 0
