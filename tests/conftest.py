import contextlib
import os
import sys

import pytest

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (REPO, os.path.join(REPO, "tests", "golden")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def pytest_collection_modifyitems(config, items):
    import torch
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


@contextlib.contextmanager
def backbone_precision(precision):
    """Class-wide ``ResNetFPN_8_2.conv_precision`` for the duration: "f16" = own tcgen05 convolutions (the default, what
    bench.py times), "fp32" = exact cuDNN executor (with cuDNN's TF32 switched off, so that differences against the
    fp32 golden vectors come from the matcher kernels only)."""
    import torch
    from onepose_st_b200.nn_common import ResNetFPN_8_2
    old_p, old_tf32 = ResNetFPN_8_2.conv_precision, torch.backends.cudnn.allow_tf32
    ResNetFPN_8_2.conv_precision = precision
    if precision == "fp32":
        torch.backends.cudnn.allow_tf32 = False
    try:
        yield
    finally:
        ResNetFPN_8_2.conv_precision, torch.backends.cudnn.allow_tf32 = old_p, old_tf32


@pytest.fixture(scope="session")
def golden():
    import torch
    gdir = os.path.join(REPO, "tests", "golden")

    def load(name):
        return torch.load(os.path.join(gdir, name + ".pt"), map_location="cpu")
    return load
