"""CPU checks of the boundary: the C-ABI library loads and exports every declared symbol, the drop-in classes
keep the reference's state_dict names, and install() redirects the reference's import paths."""
import ctypes
import os
import sys

import pytest
import torch

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    from onepose_st_b200 import _lib
    if not os.path.exists(_lib.LIB_PATH):
        _lib.build()
    return _lib


def test_library_exports_every_declared_symbol(lib):
    decl = lib.declared_symbols()
    assert len(decl) >= 16 and "opb200_coarse_match" in decl and "opb200_transformer" in decl
    assert lib.missing_symbols() == []
    l = lib.lib()
    assert l.opb200_version() == 100
    assert lib.launch_count() >= 0
    l.opb200_packed_layer_bytes.restype = ctypes.c_size_t
    assert l.opb200_packed_layer_bytes(256) == 10 * 256 * 256 * 4 + 4 * 256 * 4  # hi + lo planes + LN params


def test_ops_refuse_cpu_tensors(lib):
    from onepose_st_b200 import ops
    with pytest.raises(lib.Opb200Error):
        ops.fine_match(torch.zeros(2, 128), torch.zeros(2, 25, 128))
    with pytest.raises(lib.Opb200Error):
        ops.coarse_match(torch.zeros(1, 4, 256), torch.zeros(1, 4, 256), scale=1.0, thr=0.1, border_rm=0, border_mode=0,
                         grid_a=None, grid_b=(2, 2))


def test_state_dict_names_match_the_oracle_init():
    """The oracle's init_* state dicts carry the reference's own parameter names (they drive the reference modules
    in make_golden.py): strict loading proves the drop-in classes expose the same names and shapes."""
    from oracle import ref_ops as R
    from onepose_st_b200.loftr import LoFTR_for_OnePose_Plus, default_cfg
    from onepose_st_b200.onepose_plus import OnePosePlus_model
    from onepose_st_b200.synthetic import onepose_plus_cfg
    m = OnePosePlus_model(onepose_plus_cfg())
    m.load_state_dict(R.init_onepose_plus_state_dict(seed=1), strict=True)
    assert "dense_pos_encoding.pe" not in m.state_dict()  # non-persistent buffer, as in the reference
    l = LoFTR_for_OnePose_Plus(default_cfg)
    l.load_state_dict(R.init_loftr_state_dict(seed=1), strict=True)


def test_models_are_picklable():
    import pickle
    from onepose_st_b200.onepose_plus import OnePosePlus_model
    from onepose_st_b200.synthetic import onepose_plus_cfg
    m = OnePosePlus_model(onepose_plus_cfg()).eval()
    m2 = pickle.loads(pickle.dumps(m))  # Ray ships the model object into tasks
    assert set(m2.state_dict()) == set(m.state_dict())


def test_install_redirects_reference_imports():
    import onepose_st_b200.install as opb
    from onepose_st_b200 import loftr, onepose_plus
    saved = {k: v for k, v in sys.modules.items() if k.startswith("src")}
    try:
        opb.install()
        assert opb.installed()
        from src.models.OnePosePlus.OnePosePlusModel import OnePosePlus_model
        from src.KeypointFreeSfM.loftr_for_sfm import LoFTR_for_OnePose_Plus, default_cfg
        assert OnePosePlus_model is onepose_plus.OnePosePlus_model
        assert LoFTR_for_OnePose_Plus is loftr.LoFTR_for_OnePose_Plus and default_cfg["fine_window_size"] == 9
    finally:
        opb.uninstall()
        for k in [k for k in sys.modules if k.startswith("src")]:
            del sys.modules[k]
        sys.modules.update(saved)
    assert not opb.installed()


def test_unsupported_configs_raise():
    from onepose_st_b200.onepose_plus import CoarseMatching, LocalFeatureTransformer
    from onepose_st_b200.synthetic import onepose_plus_cfg
    cfg = onepose_plus_cfg()
    bad = dict(cfg["loftr_coarse"], attention="full")
    with pytest.raises(NotImplementedError):
        LocalFeatureTransformer(bad)
    bad = dict(cfg["coarse_matching"], type="sinkhorn")
    with pytest.raises(NotImplementedError):
        CoarseMatching(bad)


def test_synthetic_sequence_on_cpu_has_planted_structure():
    from onepose_st_b200 import synthetic
    from onepose_st_b200.onepose_plus import OnePosePlus_model
    torch.manual_seed(0)
    m = OnePosePlus_model(synthetic.onepose_plus_cfg()).eval()
    m.load_state_dict(synthetic.calibrate_onepose_state_dict(m.state_dict()))
    seq = synthetic.make_onepose_sequence(m, 300, 2, 64, 64, n_plant=20, device=torch.device("cpu"))
    assert seq["frames"].shape == (2, 1, 64, 64) and seq["descriptors3d_coarse_db"].shape == (1, 256, 300)
    assert not torch.equal(seq["frames"][0], seq["frames"][1])
    assert seq["planted_cells"].min() >= 2 * 8 + 2  # interior cells only (border_rm = 2 on an 8x8 grid)
