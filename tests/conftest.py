import glob
import json
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def golden_cases(prefix):
    return sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN, prefix + "*.npz"))
                  if not os.path.basename(p).startswith("feed_"))   # feed_*: fixtures of the data feed (tests/feed_cases.py)


class Golden:
    """One fixture minted by tests/golden/make_golden.py from the reference."""

    def __init__(self, case):
        self.case = case
        z = np.load(os.path.join(GOLDEN, case + ".npz"))
        self.z = z
        self.meta = json.loads(bytes(z["meta"]).decode())
        self.inputs = {k[3:]: torch.from_numpy(z[k]) for k in z.files if k.startswith("in.")}
        self.params = {k[2:]: torch.from_numpy(z[k]) for k in z.files if k.startswith("p.")}
        self.R = torch.from_numpy(z["R"])
        self.out = torch.from_numpy(z["out64"]).double()
        self.grads = {k[4:]: torch.from_numpy(z[k]).double() for k in z.files if k.startswith("g64.")}
        self.gin = {k[6:]: torch.from_numpy(z[k]).double() for k in z.files if k.startswith("gin64.")}

    def noise(self, kind, name=None):
        """rel-L2 distance of the reference's own fp32 run from its fp64 run."""
        key = "err32.out" if kind == "out" else f"err32.{kind}.{name}"
        return float(self.z[key])


def rel_l2(a, b):
    a = a.detach().double().cpu()
    b = b.detach().double().cpu()
    den = float(b.norm())
    return float((a - b).norm()) / (den if den > 0 else 1.0)


@pytest.fixture
def golden():
    return Golden


def pytest_terminal_summary(terminalreporter):
    """Worst error / noise and error / tolerance per tensor class over the session (tests/parity.py)."""
    try:
        import parity
    except Exception:
        return
    if not parity.RATIOS:
        return
    terminalreporter.write_line("parity margins (worst over the session): class  n  err/noise  err/tol")
    for cls, r in sorted(parity.RATIOS.items()):
        terminalreporter.write_line(f"  {cls:20s} {r['n']:5d}  {r['err_over_noise']:9.2f}  {r['err_over_tol']:7.3f}")
