"""Import shims that let the UNMODIFIED reference (/root/reference) be imported in the
build container, where pytorch_lightning / timm / yacs / kornia / ray / h5py are absent.

Test-side only (used by make_golden.py and by the optional live-reference tests, which
skip when /root/reference does not exist, e.g. on the GPU box).  SURVEY.md Appendix A.
"""
import os
import sys
import types

REFERENCE_ROOT = "/root/reference"


def reference_available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "src", "models", "OnePosePlus"))


def _mod(name, **attrs):
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    sys.modules[name] = m
    return m


def install():
    """Install the shims and return (OnePosePlus_model, LoFTR_for_OnePose_Plus, loftr default_cfg)."""
    if not reference_available():
        raise RuntimeError("reference tree not present")
    import torch
    # the kornia 0.4.1 restatement lives in the oracle (single source of truth for that boundary)
    repo = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    if repo not in sys.path:
        sys.path.insert(0, repo)
    from oracle.ref_ops import create_meshgrid as _cm, spatial_expectation2d as _se

    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    lo = os.path.join(REFERENCE_ROOT, "submodules", "LoFTR", "src")
    if lo not in sys.path:
        sys.path.insert(0, lo)

    class BaseProfiler:  # src/utils/profiler.py subclasses this
        def __init__(self, *a, **k):
            pass

    if "pytorch_lightning" not in sys.modules:
        _mod("pytorch_lightning", _logger=None)
        _mod("pytorch_lightning.profiler", BaseProfiler=BaseProfiler)
        _mod("pytorch_lightning.utilities", rank_zero_only=lambda f: f)
        _mod("pytorch_lightning.utilities.cloud_io", get_filesystem=None)
        _mod("pytorch_lightning.utilities.distributed", rank_zero_warn=None)
        _mod("pytorch_lightning.utilities.exceptions", MisconfigurationException=Exception)
    if "timm" not in sys.modules:
        _mod("timm")
        _mod("timm.models")
        _mod("timm.models.registry", register_model=lambda f: f)
    if "yacs" not in sys.modules:
        class CfgNode(dict):
            __getattr__ = dict.__getitem__
            __setattr__ = dict.__setitem__
        _mod("yacs")
        _mod("yacs.config", CfgNode=CfgNode)
    if "kornia" not in sys.modules:
        def create_meshgrid(h, w, normalized_coordinates=True, device="cpu"):
            return _cm(h, w, normalized_coordinates, device)
        dsnt = _mod("kornia.geometry.subpix.dsnt", spatial_expectation2d=_se)
        _mod("kornia")
        _mod("kornia.geometry")
        _mod("kornia.geometry.subpix", dsnt=dsnt)
        _mod("kornia.utils")
        _mod("kornia.utils.grid", create_meshgrid=create_meshgrid)
    if "src.KeypointFreeSfM" not in sys.modules:
        import importlib
        importlib.import_module("src")
        pkg = _mod("src.KeypointFreeSfM")
        pkg.__path__ = [os.path.join(REFERENCE_ROOT, "src", "KeypointFreeSfM")]

    from src.models.OnePosePlus.OnePosePlusModel import OnePosePlus_model
    from src.KeypointFreeSfM.loftr_for_sfm import LoFTR_for_OnePose_Plus, default_cfg
    return OnePosePlus_model, LoFTR_for_OnePose_Plus, default_cfg
