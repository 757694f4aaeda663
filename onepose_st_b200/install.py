"""Make the reference's import paths resolve to the sm_100a drop-in classes, without editing the reference.

    import onepose_st_b200.install as opb; opb.install()        # before the reference imports its models
    from src.models.OnePosePlus.OnePosePlusModel import OnePosePlus_model      # -> onepose_st_b200 class
    from src.KeypointFreeSfM.loftr_for_sfm import LoFTR_for_OnePose_Plus, default_cfg

Call sites this serves (reference file:line): inference.py:20,117-120,177-178;
src/inference/inference_OnePosePlus.py:30-40; src/KeypointFreeSfM/coarse_match/coarse_match_worker.py:18-29,35;
src/KeypointFreeSfM/post_optimization/matcher_model/fine_match_worker.py:11-30;
src/local_feature_object_detector/local_feature_2D_detector.py:23-37.
Only the model modules are replaced; every other ``src.*`` module keeps coming from the reference tree.
"""
import importlib
import sys
import types


def _module(name: str, **attrs) -> types.ModuleType:
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    m.__opb200__ = True
    sys.modules[name] = m
    return m


def install() -> None:
    """Register replacement modules under the reference's dotted names.  Idempotent."""
    from . import loftr, onepose_plus

    # 2D-3D matcher: `from src.models.OnePosePlus.OnePosePlusModel import OnePosePlus_model`
    _module("src.models.OnePosePlus.OnePosePlusModel", OnePosePlus_model=onepose_plus.OnePosePlus_model)
    # 2D-2D matcher: `from ..loftr_for_sfm import LoFTR_for_OnePose_Plus, default_cfg`
    pkg = _module("src.KeypointFreeSfM.loftr_for_sfm", LoFTR_for_OnePose_Plus=loftr.LoFTR_for_OnePose_Plus,
                  default_cfg=loftr.default_cfg)
    pkg.__path__ = []  # a package: `from .loftr_for_sfm.loftr import ...` style imports resolve below
    _module("src.KeypointFreeSfM.loftr_for_sfm.loftr", LoFTR_for_OnePose_Plus=loftr.LoFTR_for_OnePose_Plus)
    _module("src.KeypointFreeSfM.loftr_for_sfm.utils", __path__=[])
    _module("src.KeypointFreeSfM.loftr_for_sfm.utils.loftr_for_onepose_plus_cfg", get_cfg_defaults=lambda: loftr.default_cfg,
            lower_config=loftr.lower_config)
    # attach to already-imported parents so attribute access works too
    for parent, child in (("src.models.OnePosePlus", "OnePosePlusModel"), ("src.KeypointFreeSfM", "loftr_for_sfm")):
        if parent in sys.modules:
            setattr(sys.modules[parent], child, sys.modules[f"{parent}.{child}"])


def installed() -> bool:
    m = sys.modules.get("src.models.OnePosePlus.OnePosePlusModel")
    return bool(m is not None and getattr(m, "__opb200__", False))


def uninstall() -> None:
    for name in [n for n, m in sys.modules.items() if getattr(m, "__opb200__", False)]:
        del sys.modules[name]
    importlib.invalidate_caches()
