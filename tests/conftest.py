import glob
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
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def load_golden(name):
    z = np.load(os.path.join(GOLDEN, name + ".npz"), allow_pickle=False)
    return {k: z[k] for k in z.files}


def golden_names(prefix):
    return sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN, prefix + "*.npz")))


def t(a):
    return torch.from_numpy(np.ascontiguousarray(a))


def digest_close(tensor, g, prefix, rtol):
    """Compare a tensor with a stored digest (sum / abs-sum / strided samples).
    rtol is relative to the digest's abs-max (gradients) -- SURVEY §4.3."""
    f = tensor.detach().double().flatten().cpu()
    step = int(g[prefix + ".step"])
    samples = torch.from_numpy(g[prefix + ".samples"]).double()
    scale = float(g[prefix + ".absmax"]) or 1.0
    got = f[::step]
    assert got.numel() == samples.numel()
    err = (got - samples).abs().max().item() / scale
    assert err <= rtol, f"{prefix}: sample err {err:.3e} > {rtol}"
    abssum = float(g[prefix + ".abssum"])
    assert abs(f.abs().sum().item() - abssum) <= max(10 * rtol, 1e-12) * max(abssum, 1e-30), f"{prefix}: abssum"  # fp64 sum order


@pytest.fixture(scope="session")
def cuda_lib():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from yolo_scratch_b200 import _lib
    return _lib.lib()
