"""north_star: "matching loss curves over 100 steps".  tools/loss_curves.py trains the DCGAN
(trainer.py:208-303 schedule) with the product layers in bf16 and with stock torch layers in
fp32 from the same weights / data / noise; the full 100-step record is committed as
profiles/r01_loss_curves_100steps.json.  Here: a 40-step run on the GPU box."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_loss_curves_track_fp32_reference():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "loss_curves.py"), "--steps", "40"],
                       capture_output=True, text=True, timeout=600)
    print(r.stdout[-2000:], r.stderr[-2000:])
    assert r.returncode == 0 and "PASS" in r.stdout
