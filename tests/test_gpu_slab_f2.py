"""The fp32 slab kernel with FFMA2 phases (BD_SLAB_F2=1, csrc/stiffness_slab_f2*.cuh) against the oracle."""
import json
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))


def test_ffma2_slab_kernel_matches_oracle():
    env = dict(os.environ, BD_SLAB_F2="1", CUDA_LAUNCH_BLOCKING="1")
    env.pop("BD_STIFFNESS_VARIANT", None)
    r = subprocess.run([sys.executable, os.path.join(HERE, "slab_f2_worker.py")], env=env, capture_output=True,
                       text=True, timeout=600)
    lines = [json.loads(l) for l in r.stdout.splitlines() if l.startswith("{")]
    assert r.returncode == 0, (r.stdout[-2000:], r.stderr[-1500:])
    assert len(lines) == 6 and all(l["ok"] for l in lines)
