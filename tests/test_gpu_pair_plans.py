"""The persistent GEMM picks its plan per shape (single CTAs or cta_group::2 CTA pairs, gemm2.cu) and the FFN block
is one fused launch by default with a two-GEMM alternative (TTS_FFN_FUSE=0, c_api.cu).  The library reads both
switches once, so the non-default settings are exercised in child processes: the operator parity tests and the
golden-vector forward must pass with pairs forced on for every shape that allows them (2), with pairs off (0), with
the two-GEMM FFN, and with the chained GEMM launches (TTS_GEMM_CHAIN=1, in the mode "2" run)."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.mark.parametrize("mode", ["0", "2"])
def test_gemm_plans_under_forced_pair_mode(mode):
    env = dict(os.environ, TTS_GEMM_PAIR=mode, TTS_FFN_FUSE="0" if mode == "2" else "1",
               TTS_GEMM_CHAIN="1" if mode == "2" else "0")
    cmd = [sys.executable, "-m", "pytest", "-x", "-q", "-m", "gpu", "-p", "no:cacheprovider",
           os.path.join(HERE, "test_gpu_ops.py"), os.path.join(HERE, "test_gpu_forward.py"),
           "-k", "linear or conv1d or ffn or forward"]
    r = subprocess.run(cmd, env=env, cwd=os.path.dirname(HERE), capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-2000:]
