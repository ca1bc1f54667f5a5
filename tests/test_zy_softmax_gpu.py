"""GPU checks for csrc/softmax.cu against the fp32 restatement, and the unfused CoreAttention path of the modular GPT layer.
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


@pytest.mark.parametrize("name", ["softmax_kernels", "unfused_attention_layer"])
def test_softmax_tier(name):
    _run(name)
