"""Pose parity (BASELINE.json north_star: "the resulting PnP poses must agree within 1 cm / 1 degree").

A geometrically consistent synthetic scene: 3D points are the back-projections of planted coarse cells under a
known pose, so the 2D-3D matches of the matcher determine that pose.  Matches of the CUDA path and of the CPU
oracle go through the reference's OpenCV RANSAC-PnP settings (src/utils/metric_utils.py:179-209: EPnP, 5 px
reprojection error, 10000 iterations) and the two poses are compared with each other and with the ground truth."""
import copy

import numpy as np
import pytest
import torch

import cases
from oracle import ref_ops as R

pytestmark = pytest.mark.gpu


def _rodrigues(rvec):
    import cv2
    return cv2.Rodrigues(np.asarray(rvec, dtype=np.float64))[0]


def _scene(H=256, W=256, L=600, n_plant=220, seed=3):
    cfg = copy.deepcopy(R.ONEPOSE_PLUS_CFG)
    sd = cases.calibrated_onepose_sd(seed)
    img = cases.rand(1000 + seed, 1, 1, H, W)
    with torch.no_grad():
        fc, ff = R.resnet_fpn_8_2(sd, "backbone.", img)
    hc, wc = fc.shape[2:]
    fcp = (fc + R.position_encoding_sine(256, hc, wc)).flatten(2)[0].t()
    g = torch.Generator().manual_seed(70 + seed)
    Kmat = np.array([[300.0, 0, W / 2], [0, 300.0, H / 2], [0, 0, 1]])
    Rgt = _rodrigues([0.2, -0.35, 0.1])
    tgt = np.array([0.02, -0.01, 0.6])
    interior = torch.tensor([r * wc + c for r in range(2, hc - 1) for c in range(2, wc - 1)])
    cells = interior[torch.randperm(interior.numel(), generator=g)[:n_plant]]
    pts = torch.randperm(L, generator=g)[:n_plant]
    kp = (torch.rand(1, L, 3, generator=g) - 0.5) * 0.3
    # planted 3D points: back-project the cell's pixel (8c, 8r) at a random depth, camera -> object frame
    yy, xx = (cells // wc).numpy(), (cells % wc).numpy()
    depth = 0.5 + 0.2 * torch.rand(n_plant, generator=g).numpy()
    pix = np.stack([xx * 8.0, yy * 8.0, np.ones(n_plant)], 0)
    xcam = (np.linalg.inv(Kmat) @ pix) * depth
    xobj = Rgt.T @ (xcam - tgt[:, None])
    kp[0, pts] = torch.from_numpy(xobj.T).float()
    dc = torch.randn(1, 256, L, generator=g)
    df = torch.randn(1, 128, L, generator=g)
    target_sim = 10 + 15 * torch.rand(n_plant, generator=g)
    dc[0, :, pts] = cases.whitened_plant(fcp, cells, target_sim * (256 * 0.0801) * (hc * wc) / 256)
    df[0, :, pts] = ff[0][:, torch.from_numpy(yy) * 4, torch.from_numpy(xx) * 4] * 0.05
    data = dict(query_image=img, keypoints3d=kp, descriptors3d_db=df, descriptors3d_coarse_db=dc)
    return cfg, sd, data, Kmat, Rgt, tgt


def _pnp(pts3d, pts2d, Kmat):
    import cv2
    cv2.setRNGSeed(0)
    ok, rvec, tvec, inl = cv2.solvePnPRansac(np.ascontiguousarray(pts3d, dtype=np.float64), np.ascontiguousarray(pts2d, dtype=np.float64),
                                             Kmat, np.zeros((8, 1)), reprojectionError=5, iterationsCount=10000,
                                             flags=cv2.SOLVEPNP_EPNP)
    assert ok
    return _rodrigues(rvec), tvec.reshape(3), 0 if inl is None else len(inl)


def _pose_diff(Ra, ta, Rb, tb):
    cos = np.clip((np.trace(Ra.T @ Rb) - 1) / 2, -1, 1)
    return np.degrees(np.arccos(cos)), np.linalg.norm(ta - tb) * 100.0  # degrees, centimetres


def test_pnp_pose_parity_with_oracle():
    pytest.importorskip("cv2")
    from onepose_st_b200.onepose_plus import OnePosePlus_model
    cfg, sd, data, Kmat, Rgt, tgt = _scene()
    ref = {k: v.clone() for k, v in data.items()}
    with torch.no_grad():
        R.onepose_plus_forward(sd, cfg, ref)
    from conftest import backbone_precision
    with backbone_precision("fp32"):
        model = OnePosePlus_model(copy.deepcopy(cfg))
        model.load_state_dict(sd, strict=True)
        model = model.cuda().eval()
        ours = {k: v.cuda() for k, v in data.items()}
        model(ours)
    m_ref, m_ours = len(ref["mconf"]), len(ours["mconf"])
    print(f"matches: ours {m_ours}, oracle {m_ref}")
    assert m_ref >= 100 and abs(m_ours - m_ref) <= 2
    Ro, to, no = _pnp(ref["mkpts_3d_db"].numpy(), ref["mkpts_query_f"].numpy(), Kmat)
    Rg, tg, ng = _pnp(ours["mkpts_3d_db"].cpu().numpy(), ours["mkpts_query_f"].cpu().numpy(), Kmat)
    d_deg, d_cm = _pose_diff(Rg, tg, Ro, to)
    g_deg, g_cm = _pose_diff(Rg, tg, Rgt, tgt)
    print(f"pose ours vs oracle: {d_deg:.4f} deg, {d_cm:.4f} cm (inliers {ng} vs {no}); ours vs ground truth: {g_deg:.3f} deg, {g_cm:.3f} cm")
    assert d_deg < 1.0 and d_cm < 1.0   # north_star: 1 cm / 1 degree
    assert g_deg < 1.0 and g_cm < 1.0


def test_pnp_pose_parity_tf32_grade_backbone():
    """Same scene with the default conv_precision="f16", i.e. the fp16-operand tcgen05 backbone (the configuration
    bench.py times): the pose must still agree with the fp32 oracle within north_star's 1 cm / 1 degree, and the
    match set must be essentially the oracle's."""
    pytest.importorskip("cv2")
    from onepose_st_b200.onepose_plus import OnePosePlus_model
    from onepose_st_b200.nn_common import FusedBackboneF16
    cfg, sd, data, Kmat, Rgt, tgt = _scene()
    ref = {k: v.clone() for k, v in data.items()}
    with torch.no_grad():
        R.onepose_plus_forward(sd, cfg, ref)
    from conftest import backbone_precision
    with backbone_precision("f16"):
        model = OnePosePlus_model(copy.deepcopy(cfg))
        model.load_state_dict(sd, strict=True)
        model = model.cuda().eval()
        ours = {k: v.cuda() for k, v in data.items()}
        model(ours)
        assert isinstance(model.backbone.fused(), FusedBackboneF16)
    ref_pairs = set(zip(ref["i_ids"].tolist(), ref["j_ids"].tolist()))
    our_pairs = set(zip(ours["i_ids"].cpu().tolist(), ours["j_ids"].cpu().tolist()))
    common = len(ref_pairs & our_pairs)
    print(f"matches: ours {len(our_pairs)}, oracle {len(ref_pairs)}, common {common}")
    assert common >= 0.9 * len(ref_pairs) and len(our_pairs) <= 1.1 * len(ref_pairs)
    Ro, to, no = _pnp(ref["mkpts_3d_db"].numpy(), ref["mkpts_query_f"].numpy(), Kmat)
    Rg, tg, ng = _pnp(ours["mkpts_3d_db"].cpu().numpy(), ours["mkpts_query_f"].cpu().numpy(), Kmat)
    d_deg, d_cm = _pose_diff(Rg, tg, Ro, to)
    g_deg, g_cm = _pose_diff(Rg, tg, Rgt, tgt)
    print(f"[tf32-grade backbone] pose ours vs oracle: {d_deg:.4f} deg, {d_cm:.4f} cm (inliers {ng} vs {no}); vs ground truth: {g_deg:.3f} deg, {g_cm:.3f} cm")
    assert d_deg < 1.0 and d_cm < 1.0
    assert g_deg < 1.0 and g_cm < 1.0
