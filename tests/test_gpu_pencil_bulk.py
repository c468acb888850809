"""The experimental bulk-copy pencil kernel (BD_PENCIL_BULK=1, csrc/stiffness_pencil_bulk.cuh) against the oracle.

The kernel was written after the round's GPU budget was spent and has not run on a GPU yet, so this test is
opt-in: it only runs with BD_EXPERIMENTAL=1 (first thing to do with the next GPU call, under a timeout: the kernel
waits on an mbarrier)."""
import json
import os
import subprocess
import sys

import pytest

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(os.environ.get("BD_EXPERIMENTAL") != "1",
                                 reason="experimental kernel, not yet validated on a GPU (set BD_EXPERIMENTAL=1)")]

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.mark.parametrize("paired", ["0", "1"])
def test_bulk_pencil_kernel_matches_oracle(paired):
    env = dict(os.environ, BD_PENCIL_BULK="1", BD_PENCIL_PAIRED=paired, CUDA_LAUNCH_BLOCKING="1")
    env.pop("BD_STIFFNESS_VARIANT", None)
    r = subprocess.run([sys.executable, os.path.join(HERE, "pencil_bulk_worker.py")], env=env, capture_output=True,
                       text=True, timeout=300)
    lines = [json.loads(l) for l in r.stdout.splitlines() if l.startswith("{")]
    assert r.returncode == 0, (r.stdout[-2000:], r.stderr[-1500:])
    assert len(lines) == 3 and all(l["ok"] for l in lines)
