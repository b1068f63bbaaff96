"""-m gpu: every tile path of the library under BVC_GUARD=1 (4 KB guard bands around each device allocation, verified at every
wait / shard call / destroy; csrc/bvc_api.cu).  compute-sanitizer is closed on this pool, so this is the out-of-bounds-WRITE
check the suite can run itself; tools/sanitize_run.py also compares every result with the oracle."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(code_or_script, extra_env):
    env = dict(os.environ, BVC_GUARD="1", **extra_env)
    return subprocess.run([sys.executable] + code_or_script, cwd=ROOT, env=env, capture_output=True, text=True, timeout=900)


def test_all_tile_paths_leave_the_guard_bands_intact():
    r = _run([os.path.join(ROOT, "tools", "sanitize_run.py")], {})
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "all tiles agree with the oracle" in r.stdout and "BVC_GUARD" not in r.stderr


def test_guard_bands_catch_a_write_behind_an_allocation():
    code = ("import basevarc_b200 as B\n"
            "t = B.synth_tile(50, 100, 1.0, 'lowpass', seed=3)\n"
            "c = B.BaseTypeCaller(50, max_sites=t.n_sites, max_entries=t.n_entries)\n"
            "try:\n    c.call_sites(t)\nexcept B.BvcError as e:\n    print('CAUGHT', e)\n    import os; os._exit(0)\n"
            "print('MISSED')\n")
    r = _run(["-c", code], {"BVC_GUARD_SELFTEST": "1"})
    assert "CAUGHT" in r.stdout and "BVC_GUARD" in r.stdout and "back band, byte 0" in r.stdout, r.stdout + r.stderr
