"""Run-to-run identity of every kernel family (scripts/all_lanes_workload.py): builder variants, sort, renderer lanes,
rt_trace, feature scenes.  Two fresh processes print the same checksums line by line; so does a third one whose builder
climbs without the in-warp levels' compaction and without the look before the exchange (other interleavings of the same
protocol).  The reference's builder is not deterministic in this sense (unfenced refit)."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(tmp_path, extra_env=None):
    env = dict(os.environ, TMPDIR=str(tmp_path))
    env.update(extra_env or {})
    r = subprocess.run([sys.executable, os.path.join(ROOT, "scripts", "all_lanes_workload.py")], capture_output=True, text=True, timeout=600, env=env)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if " crc " in l]
    assert len(lines) > 100 and lines[-1].startswith("workload done")
    return lines


def test_two_processes_print_the_same_checksums(tmp_path):
    a = _run(tmp_path)
    b = _run(tmp_path)
    c = _run(tmp_path, {"RT_BU_NO_COMPACT": "1", "RT_BU_NO_PEEK": "1"})
    for x, y in zip(a, b):
        assert x == y, "run-to-run difference:\n%s\n%s" % (x, y)
    for x, y in zip(a, c):
        assert x == y, "builder switches change an output:\n%s\n%s" % (x, y)
    assert len(a) == len(b) == len(c)
