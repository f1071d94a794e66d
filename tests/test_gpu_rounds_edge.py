"""GPU parity: the batched-affine round kernels on their special slots.  The halving rounds only run by default on MSMs of
>= 4 M (point, window) entries, where random inputs never put P and P, or P and -P, or the identity into one slot; the
edge-case MSMs of test_gpu_msm.py are far smaller and take the XYZZ-only path.  Here the same edge cases are pushed through
the round kernels by forcing the number of rounds (FB_MSM_ROUNDS), with the first round's slot map written by the scatter or
by the separate map kernel (FB_MSM_DIRECT), its pass 2 gathering again or reading the points parked by pass 1
(FB_MSM_REGATHER), and with and without the x-only table copy (FB_MSM_XTAB).  Checked against known discrete logs."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.mark.parametrize("env", [
    {"FB_MSM_ROUNDS": "1"},
    {"FB_MSM_ROUNDS": "3"},
    {"FB_MSM_ROUNDS": "8"},
    {"FB_MSM_ROUNDS": "3", "FB_MSM_DIRECT": "0"},
    {"FB_MSM_ROUNDS": "3", "FB_MSM_REGATHER": "0"},
    {"FB_MSM_ROUNDS": "2", "FB_MSM_XTAB": "0"},
    {"FB_MSM_ROUNDS": "0"},
], ids=lambda e: ",".join("%s=%s" % (k[7:], v) for k, v in e.items()))
def test_round_kernels_on_special_slots(gpu, env):
    e = dict(os.environ)
    e.update(env)
    r = subprocess.run([sys.executable, os.path.join(HERE, "rounds_edge_check.py")], env=e, capture_output=True, text=True,
                       timeout=600, cwd=os.path.dirname(HERE))
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    assert r.stdout.strip().splitlines()[-1].startswith("ok ")


@pytest.mark.parametrize("waves", [1, 2, 3, 5, 6])
def test_round_kernels_chains_per_thread(gpu, waves):
    """Pass 2 of the rounds on `waves` times the grid of pass 1 (FB_MSM_WAVES forced): pass 1 walks each thread's slots as
    that many chains and derives every chain's inverse from the CTA's one inversion.  2^20 points (smaller jobs clamp to one
    chain), random and skewed scalars, known discrete logs."""
    e = dict(os.environ)
    e["FB_MSM_WAVES"] = str(waves)
    r = subprocess.run([sys.executable, os.path.join(HERE, "rounds_edge_check.py"), "large"], env=e, capture_output=True,
                       text=True, timeout=900, cwd=os.path.dirname(HERE))
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    assert r.stdout.strip().splitlines()[-1].startswith("ok ")
