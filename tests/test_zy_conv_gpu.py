"""GPU checks for csrc/conv.cu: conv + bias + ReLU fwd / dgrad / wgrad, 2x2 max-pool, log-softmax + NLL, an MNIST step with native_ops=True against the cuDNN / ATen model, and csrc/batchnorm.cu against torch.nn.BatchNorm2d.
Each check runs in a CHILD process (tests/_zy_checks.py): a fault in one kernel tier cannot poison the CUDA context of the pytest
process and with it every test collected later."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(name):
    r = subprocess.run([sys.executable, "-m", "tests._zy_checks", name], cwd=ROOT, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0 and ("ok " + name) in r.stdout, (r.stdout[-2000:] + r.stderr[-4000:])


@pytest.mark.parametrize("name", ["conv_kernels", "pool_and_log_softmax", "mnist_step", "batchnorm_kernels"])
def test_conv_tier(name):
    _run(name)
