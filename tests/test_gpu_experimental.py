"""Kernel variants that were written after the round's GPU budget was spent and have not run on a GPU yet.
Opt-in: BD_EXPERIMENTAL=1 (run them first, under a timeout, with the next GPU call)."""
import json
import os
import subprocess
import sys

import pytest

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(os.environ.get("BD_EXPERIMENTAL") != "1",
                                 reason="experimental kernels, not yet validated on a GPU (set BD_EXPERIMENTAL=1)")]

HERE = os.path.dirname(os.path.abspath(__file__))


def _run(env_extra, cases):
    env = dict(os.environ, CUDA_LAUNCH_BLOCKING="1", **env_extra)
    env.pop("BD_STIFFNESS_VARIANT", None)
    r = subprocess.run([sys.executable, os.path.join(HERE, "variant_worker.py")] + cases, env=env,
                       capture_output=True, text=True, timeout=600)
    lines = [json.loads(l) for l in r.stdout.splitlines() if l.startswith("{")]
    assert r.returncode == 0, (r.stdout[-2000:], r.stderr[-1500:])
    assert len(lines) == len(cases) and all(l["ok"] for l in lines)
    return lines


def test_dmma_kernel_with_value_prefetch():
    """BD_MMA8_PREFETCH=2 (csrc/stiffness_mma8.cuh): the next cell's input values are prefetched too."""
    lines = _run({"BD_MMA8_PREFETCH": "2"},
                 ["6:1:f64:4000", "6:1:f64:400000", "7:0:f64:300000", "5:1:f64:300000", "6:0:f64:300000",
                  "7:0:f32:300000", "6:0:f32:300000"])
    assert all("mma8" in l["kernel"] for l in lines)


def test_fp64_pencil_kernel_with_five_row_ring():
    """BD_PENCIL_RING=3 (csrc/stiffness_pencil.cuh): ring of 5 G rows restarted per cell, fp64, nq = 9."""
    lines = _run({"BD_PENCIL_RING": "3"}, ["7:1:f64:4000", "7:1:f64:450000"])
    assert all("pencil" in l["kernel"] for l in lines)
