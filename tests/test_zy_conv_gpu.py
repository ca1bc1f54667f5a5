"""GPU checks for the kernels added after the last GPU session of round 1 - csrc/conv.cu: conv + bias + ReLU fwd / dgrad / wgrad, 2x2 max-pool, log-softmax + NLL, and an MNIST step with native_ops=True against the cuDNN / ATen model.
Collected after the other kernel tiers (file name) and run in child processes (see tests/_zy_checks.py).  They are marked
xfail(strict=False) until their first hardware run: a first-run fault is reported as XFAIL instead of stopping a `-x` run
before the tests collected after this file; a pass is reported as XPASS."""
import os
import subprocess
import sys

import pytest

pytestmark = [pytest.mark.gpu, pytest.mark.xfail(reason="kernel not yet run on hardware (added after the round's last GPU session)", strict=False)]
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(name):
    r = subprocess.run([sys.executable, "-m", "tests._zy_checks", name], cwd=ROOT, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0 and ("ok " + name) in r.stdout, (r.stdout[-2000:] + r.stderr[-4000:])


@pytest.mark.parametrize("name", ["conv_kernels", "pool_and_log_softmax", "mnist_step", "batchnorm_kernels"])
def test_conv_tier(name):
    _run(name)
