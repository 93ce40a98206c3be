"""Race hunting for the warp-specialised tensor-core head (csrc/head_tc.cuh): with one role slowed down per
loop iteration, every probability must stay bit-identical to the undisturbed run. Runs tools/stress_roles.py in
a subprocess against the DEBUG build of the library (a deadlock traps after ~2 s and kills that process only;
the device watchdog says which warp waited on which barrier)."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.gpu
def test_role_delays_do_not_change_results_or_deadlock():
    env = dict(os.environ, STRESS_N="60000")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "stress_roles.py")], env=env, capture_output=True,
                       text=True, timeout=600)
    assert r.returncode == 0 and "STRESS OK" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]
