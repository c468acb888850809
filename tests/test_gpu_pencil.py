"""The pencil kernel (csrc/stiffness_pencil.cuh) against the oracle, for both cells-per-CTA settings.

The kernel choice is latched per process (BD_STIFFNESS_VARIANT), so the cases run in a worker process."""
import json
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.mark.parametrize("cpb", ["0", "1"])
def test_pencil_kernel_matches_oracle(cpb):
    env = dict(os.environ, BD_STIFFNESS_VARIANT="pencil", BD_PENCIL_CPB=cpb)
    r = subprocess.run([sys.executable, os.path.join(HERE, "pencil_worker.py")], env=env, capture_output=True,
                       text=True, timeout=900)
    lines = [json.loads(l) for l in r.stdout.splitlines() if l.startswith("{")]
    assert r.returncode == 0, (r.stdout[-3000:], r.stderr[-3000:])
    assert len(lines) == 16 and all(l["ok"] for l in lines)
