"""GPU parity of the drop-in models (reference operator API on the sm_100a kernels) against the golden
vectors of the UNMODIFIED reference models, plus the standalone operator entry points."""
import copy

import pytest
import torch

import cases
from oracle import ref_ops as R

pytestmark = pytest.mark.gpu

DEV = "cuda"


def to_dev(d):
    return {k: (v.to(DEV) if isinstance(v, torch.Tensor) else v) for k, v in d.items()}


def build_onepose(c, materialize=True):
    from onepose_st_b200.onepose_plus import OnePosePlus_model
    cfg = copy.deepcopy(c["cfg"])
    cfg["coarse_matching"]["materialize_conf_matrix"] = materialize
    m = OnePosePlus_model(cfg)
    m.load_state_dict(c["sd"], strict=True)
    return m.to(DEV).eval()


@pytest.fixture(autouse=True)
def _fp32_convs():
    # exact-fp32 (cuDNN) backbone here, so that differences against the fp32 golden vectors come from the matcher
    # kernels only; the default fp16 tcgen05 backbone is compared in test_gpu_bench_config.py / test_gpu_pose.py
    from conftest import backbone_precision
    with backbone_precision("fp32"):
        yield


@pytest.mark.parametrize("name,kw", [("onepose_full", {}), ("onepose_full_mask", {"with_mask": True})])
def test_onepose_model_vs_golden(golden, name, kw):
    c = cases.case_onepose_full(**kw)
    model = build_onepose(c)
    data = to_dev(c["data"])
    model(data)
    g = golden(name)
    for k in ("bs", "q_hw_i", "q_hw_c", "q_hw_f", "b_ids", "i_ids", "j_ids", "gt_mask", "m_bids", "mkpts_3d_db",
              "mkpts_query_c", "mconf", "W", "expec_f", "mkpts_query_f", "conf_matrix"):
        assert k in data, f"missing data key {k}"
    err = (data["conf_matrix"].cpu() - g["conf_matrix"]).abs().max().item()
    print(f"{name}: conf max abs err {err:.2e}; M={len(data['mconf'])} (ref {len(g['mconf'])})")
    assert err < 1e-3  # north_star: confidence values within max-abs 1e-3
    got = set(zip(data["i_ids"].tolist(), data["j_ids"].tolist()))
    ref = set(zip(g["i_ids"].tolist(), g["j_ids"].tolist()))
    # sets may differ only where the reference confidence sits within 1e-3 of the threshold
    for (i, j) in got ^ ref:
        assert abs(g["conf_matrix"][0, i, j].item() - 0.1) < 2e-3, (i, j, g["conf_matrix"][0, i, j].item())
    assert len(got ^ ref) <= 2
    if got == ref:
        assert torch.equal(data["i_ids"].cpu(), g["i_ids"]) and torch.equal(data["j_ids"].cpu(), g["j_ids"])
        assert torch.equal(data["mkpts_query_c"].cpu(), g["mkpts_query_c"])
        assert torch.equal(data["mkpts_3d_db"].cpu(), g["mkpts_3d_db"])
        assert (data["mconf"].cpu() - g["mconf"]).abs().max().item() < 1e-3
        e_exp = (data["expec_f"].cpu() - g["expec_f"]).abs().max().item()
        e_pts = (data["mkpts_query_f"].cpu() - g["mkpts_query_f"]).abs().max().item()
        print(f"{name}: expec_f err {e_exp:.2e}, mkpts_query_f err {e_pts:.2e} px")
        assert e_exp < 2e-3 and e_pts < 2e-2
    # same-confidence-matrix contract on the model's own matrix
    keep = R._border_keep_onepose(1, data["conf_matrix"].shape[1], 16, 16, 2, "cpu")
    rb, ri, rj, rc = R.coarse_select(data["conf_matrix"].cpu(), 0.1, keep)
    assert torch.equal(ri, data["i_ids"].cpu()) and torch.equal(rj, data["j_ids"].cpu()) and torch.equal(rc, data["mconf"].cpu())


def test_onepose_model_default_omits_conf_matrix_and_matches():
    c = cases.case_onepose_full()
    a, b = build_onepose(c, True), build_onepose(c, False)
    da, db = to_dev(c["data"]), to_dev(c["data"])
    a(da)
    b(db)
    assert "conf_matrix" not in db
    # the default (single-pass) coarse matcher sweeps the similarities with one fp16 product and re-evaluates the
    # captured elements exactly: same index sets, confidences inside the 1e-3 contract (measured ~1e-5 here)
    for k in ("i_ids", "j_ids"):
        assert torch.equal(da[k], db[k])
    print("default vs materialising path: mconf max abs diff", (da["mconf"] - db["mconf"]).abs().max().item())
    for k, tol in (("mconf", 1e-3), ("mkpts_query_f", 1e-3), ("expec_f", 1e-4)):
        assert (da[k] - db[k]).abs().max().item() < tol, k


def test_onepose_model_no_matches():
    c = cases.case_onepose_full()
    model = build_onepose(c, False)
    data = to_dev(c["data"])
    data["descriptors3d_coarse_db"] = torch.randn_like(data["descriptors3d_coarse_db"])  # nothing planted
    model(data)
    assert data["mconf"].shape == (0,) and data["expec_f"].shape == (0, 3)
    assert data["mkpts_query_f"].shape == (0, 2) and data["mkpts_3d_db"].shape == (0, 3)
    assert data["b_ids"].dtype == torch.int64


@pytest.mark.parametrize("materialize", [True, False])
@pytest.mark.parametrize("mode", ["full", "coarse", "fine_only", "extract"])
def test_loftr_model_vs_golden(golden, mode, materialize):
    """Both routes of CoarseMatching against the reference's golden outputs: ``materialize`` = two-pass path that also
    writes conf_matrix; default = the single-pass screening sweep (what LoFTR callers and tools/stage_bench.py run)."""
    from onepose_st_b200.loftr import LoFTR_for_OnePose_Plus
    c = cases.case_loftr_full(mode=mode)
    cfg = copy.deepcopy(c["cfg"])
    cfg["match_coarse"]["materialize_conf_matrix"] = materialize
    m = LoFTR_for_OnePose_Plus(cfg, enable_fine_matching=(mode != "coarse"))
    m.load_state_dict(c["sd"], strict=True)
    m = m.to(DEV).eval()
    data = to_dev(c["data"])
    kw = dict(extract_coarse_feature=True, extract_fine_feature=True) if mode == "extract" else {}
    m(data, **kw)
    g = golden(f"loftr_{mode}")
    if "conf_matrix" in g:
        assert ("conf_matrix" in data) == materialize
        assert len(g["mconf"]) > 30 and g["mconf"].min() < 0.3 and g["mconf"].max() > 0.9  # a spread, not saturation
        if materialize:
            err = (data["conf_matrix"].cpu() - g["conf_matrix"]).abs().max().item()
            print(f"loftr {mode}: conf_matrix max abs err {err:.2e}")
            assert err < 1e-3
        got = set(zip(data["i_ids"].tolist(), data["j_ids"].tolist()))
        ref = set(zip(g["i_ids"].tolist(), g["j_ids"].tolist()))
        for (i, j) in got ^ ref:  # only where the reference sits within 1e-3 of the threshold
            assert abs(g["conf_matrix"][0, i, j].item() - 0.2) < 1e-3, (i, j, g["conf_matrix"][0, i, j].item())
        assert len(got ^ ref) <= 1
        if got != ref:
            pytest.skip("a threshold-straddling pair differs; the remaining keys cannot be compared row by row")
        assert torch.equal(data["i_ids"].cpu(), g["i_ids"]) and torch.equal(data["j_ids"].cpu(), g["j_ids"])
        err = (data["mconf"].cpu() - g["mconf"]).abs().max().item()
        print(f"loftr {mode} ({'two-pass' if materialize else 'default single-pass'}): mconf max abs err {err:.2e}")
        assert err < 1e-3  # north_star: confidence values within max-abs 1e-3
    for k, v in g.items():
        if k == "conf_matrix" and not materialize:
            continue
        assert k in data, k
        if k in ("conf_matrix", "mconf"):
            continue
        if v.dtype in (torch.int64, torch.bool):
            assert torch.equal(data[k].cpu(), v), k
        else:
            tol = 2e-3 if k == "expec_f" else (2e-2 if k.startswith("mkpts") else 1e-3)
            err = (data[k].cpu().float() - v.float()).abs().max().item()
            assert err < tol, f"{k}: {err}"
    print(f"loftr {mode}: M={len(data['mconf'])}")


@pytest.mark.parametrize("masked", [False, True])
def test_linear_attention_operator(golden, masked):
    from onepose_st_b200.nn_common import LinearAttention
    c = cases.case_linattn(masked)
    qm = None if c["q_mask"] is None else c["q_mask"].to(DEV)
    km = None if c["kv_mask"] is None else c["kv_mask"].to(DEV)
    out = LinearAttention(32)(c["q"].to(DEV), c["k"].to(DEV), c["v"].to(DEV), qm, km)
    err = (out.cpu() - golden(f"linattn_{int(masked)}")["out"]).abs().max().item()
    print("linear attention operator err", err)
    assert err < 1e-4


@pytest.mark.parametrize("d_model", [256, 128])
@pytest.mark.parametrize("masked", [False, True])
def test_encoder_layer_operator(golden, d_model, masked):
    from onepose_st_b200.nn_common import LoFTREncoderLayer
    c = cases.case_encoder_layer(d_model, masked)
    lay = LoFTREncoderLayer(d_model, 8)
    lay.load_state_dict({k[len("t.layers.0."):]: v for k, v in c["sd"].items()}, strict=True)
    lay = lay.to(DEV).eval()
    xm = None if c["x_mask"] is None else c["x_mask"].to(DEV)
    sm = None if c["source_mask"] is None else c["source_mask"].to(DEV)
    out = lay(c["x"].to(DEV), c["source"].to(DEV), xm, sm)
    err = (out.cpu() - golden(f"enc_layer_{d_model}_{int(masked)}")["out"]).abs().max().item()
    print(f"encoder layer d={d_model} masked={masked}: err {err:.2e}")
    assert err < 2e-4


def test_ops_fail_loudly_without_cuda_tensors():
    from onepose_st_b200 import ops, _lib
    with pytest.raises(_lib.Opb200Error):
        ops.fine_match(torch.randn(3, 128), torch.randn(3, 25, 128))


def test_keypoint_encoding_kernel_vs_golden(golden):
    """normalize_3d_keypoints + KeypointEncoding_linear on the fused kernel (csrc/kptenc.cu) against the
    reference's golden output; both entry points -- the module's forward (keypoints already normalised) and the model's
    fused route (normalisation inside the kernel, broadcast 3D model included)."""
    from onepose_st_b200.onepose_plus import KeypointEncoding_linear, normalize_3d_keypoints
    c = cases.case_kpt_encoding()
    g = golden("kpt_encoding")
    enc = KeypointEncoding_linear(3, 256, [32, 64, 128], norm_method="instancenorm")
    enc.load_state_dict({k[len("kpt_3d_pos_encoding."):]: v for k, v in c["sd"].items()}, strict=True)
    enc = enc.to(DEV).eval()
    kp, desc = c["kpts"].to(DEV), c["desc"].to(DEV)
    with torch.no_grad():
        assert enc._kernel_ok(kp, desc)
        a = enc(normalize_3d_keypoints(kp), desc)                  # forward(kpts_normalised, descriptors)
        b = enc.encode(kp, desc, True)                             # fused normalisation
        eager = desc + enc.mlp(normalize_3d_keypoints(kp)).transpose(2, 1)
        one = enc.encode(kp[:1].expand(3, -1, -1), desc[:1].expand(3, -1, -1), True)   # stride-0 batch views
    for name, t in (("forward", a), ("fused", b), ("eager", eager)):
        err = (t.cpu() - g["out"]).abs().max().item()
        print(f"keypoint encoding {name}: max abs err vs golden {err:.2e}")
        assert err < 2e-5
    assert (one[2] - b[0]).abs().max().item() < 1e-6 and one.shape == (3, 256, 77)
