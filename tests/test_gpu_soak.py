"""Short soak of the CUDA path (tools/soak.py): random shapes and lengths, every forward run twice and compared bit
for bit, a side stream of unrelated kernels perturbing CTA placement.  The long version (tens of thousands of
iterations) is run by hand; profiles/README.md records the last one."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("noise", ["", "1"])
def test_short_soak(noise):
    env = dict(os.environ)
    if noise:
        env["SOAK_NOISE"] = noise
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "soak.py"), "150", "7"], env=env, cwd=ROOT,
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "soak ok" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]
