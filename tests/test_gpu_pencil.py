"""The pencil kernel (csrc/stiffness_pencil.cuh) against the oracle, for the three G-ring settings and the FFMA2 variant.

The settings are latched per process (BD_PENCIL_RING, BD_PENCIL_PAIRED), so the cases run in a worker process."""
import json
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.mark.parametrize("degree", [5, 6, 7])
@pytest.mark.parametrize("ring,paired", [("0", "0"), ("1", "0"), ("2", "0"), ("0", "1"), ("0", "2")])
def test_pencil_kernel_matches_oracle(ring, paired, degree):
    env = dict(os.environ, BD_PENCIL_RING=ring, BD_PENCIL_PAIRED=paired, CUDA_LAUNCH_BLOCKING="1")
    env.pop("BD_STIFFNESS_VARIANT", None)
    r = subprocess.run([sys.executable, os.path.join(HERE, "pencil_worker.py"), str(degree)], env=env,
                       capture_output=True, text=True, timeout=900)
    lines = [json.loads(l) for l in r.stdout.splitlines() if l.startswith("{")]
    assert r.returncode == 0, (r.stdout[-2000:], r.stderr[-1500:])
    assert len(lines) == 4 and all(l["ok"] for l in lines)
