"""Deterministic inputs for the golden-vector cases.

Shared by ``make_golden.py`` (which feeds them to the UNMODIFIED reference modules in the
build container) and by the tests (which feed them to the oracle on CPU and to the CUDA
product on the GPU box).  Everything is drawn from seeded ``torch.Generator``s on CPU so
the inputs can be rebuilt anywhere the same torch build is present; weights come from the
oracle's seeded ``init_*_state_dict`` (parameter names = the reference's own).
"""
import copy
import math

import torch

from oracle import ref_ops as R


def _g(seed):
    return torch.Generator().manual_seed(seed)


def randn(seed, *shape):
    return torch.randn(*shape, generator=_g(seed))


def rand(seed, *shape):
    return torch.rand(*shape, generator=_g(seed))


def layer_weights(d_model, n_layers, seed, prefix="t."):
    sd = {}
    R._init_transformer(sd, prefix, _g(seed), d_model, n_layers)
    return sd


# ------------------------------------------------------------------ operator-level cases
def case_linattn(masked):
    q, k, v = randn(1, 2, 37, 8, 32), randn(2, 2, 53, 8, 32), randn(3, 2, 53, 8, 32)
    qm = km = None
    if masked:
        qm = rand(4, 2, 37) > 0.2
        km = rand(5, 2, 53) > 0.2
    return dict(q=q, k=k, v=v, q_mask=qm, kv_mask=km)


def case_encoder_layer(d_model, masked):
    sd = layer_weights(d_model, 1, 10 + d_model)
    x, src = randn(11, 2, 50, d_model), randn(12, 2, 70, d_model)
    xm = sm = None
    if masked:
        xm, sm = rand(13, 2, 50) > 0.2, rand(14, 2, 70) > 0.2
    return dict(sd=sd, x=x, source=src, x_mask=xm, source_mask=sm)


def case_tr_onepose_coarse(masked):
    sd = layer_weights(256, 6, 20)
    d3 = randn(21, 1, 256, 90)
    d2 = randn(22, 1, 64, 256)
    qm = (rand(23, 1, 64) > 0.15) if masked else None
    return dict(sd=sd, desc3d=d3, desc2d=d2, query_mask=qm, layer_names=["self", "cross"] * 3, nhead=8)


def case_tr_loftr_coarse(masked):
    sd = layer_weights(256, 8, 30)
    f0, f1 = randn(31, 1, 48, 256), randn(32, 1, 64, 256)
    m0 = m1 = None
    if masked:
        m0, m1 = rand(33, 1, 48) > 0.15, rand(34, 1, 64) > 0.15
    return dict(sd=sd, feat0=f0, feat1=f1, mask0=m0, mask1=m1, layer_names=["self", "cross"] * 4, nhead=8)


def case_tr_onepose_fine():
    sd = layer_weights(128, 2, 40)
    return dict(sd=sd, desc3d=randn(41, 7, 128, 1), desc2d=randn(42, 7, 25, 128), query_mask=None,
                layer_names=["self", "cross"], nhead=8)


def case_tr_loftr_fine():
    sd = layer_weights(128, 2, 50)
    return dict(sd=sd, feat0=randn(51, 7, 81, 128), feat1=randn(52, 7, 81, 128), mask0=None, mask1=None,
                layer_names=["self", "cross"], nhead=8)


def planted_pair(seed, n, L, S, C, n_plant, hc, wc, border, noise=0.1, gain=1.0):
    """Stage-level CoarseMatching input (SURVEY.md section 8d config 3 style): f2d ~ N(0,1),
    f3d[i] = gain_i * f2d[pi(i)] + noise for planted i, N(0,1) otherwise.  ``gain`` in (0,1]
    draws per-point gains from U[gain,1] so confidences straddle the threshold."""
    g = _g(seed)
    f2d = torch.randn(n, S, C, generator=g)
    f3d = torch.randn(n, L, C, generator=g)
    for b in range(n):
        cells = [r * wc + c for r in range(hc) for c in range(wc)]
        perm = torch.randperm(len(cells), generator=g)[:n_plant]
        pts = torch.randperm(L, generator=g)[:n_plant]
        gains = gain + (1 - gain) * torch.rand(n_plant, generator=g)
        f3d[b, pts] = gains[:, None] * f2d[b, torch.tensor(cells)[perm]] + noise * torch.randn(n_plant, C, generator=g)
    return f3d, f2d


def case_cm_onepose(variant):
    hc, wc, L = 16, 16, 200
    n = 2 if variant == "batch2" else 1
    f3d, f2d = planted_pair(60, n, L, hc * wc, 256, 120, hc, wc, 2, gain=0.55)
    data = {"q_hw_i": torch.Size([hc * 8, wc * 8]), "q_hw_c": torch.Size([hc, wc]),
            "keypoints3d": randn(61, n, L, 3)}
    mask = None
    if variant == "mask":
        mask = torch.ones(n, hc, wc, dtype=torch.bool)
        mask[:, :, 11:] = False
        mask = mask.flatten(-2)
    if variant == "scale":
        data["query_image_scale"] = torch.tensor([[1.5, 0.75]]).repeat(n, 1)
    return dict(feat3d=f3d, feat2d=f2d, data=data, mask_query=mask, thr=0.1, border_rm=2, temperature=0.08)


def case_cm_loftr(variant):
    h0, w0, h1, w1 = 16, 14, 14, 18
    g = _g(70)
    S0, S1 = h0 * w0, h1 * w1
    f1 = torch.randn(1, S1, 256, generator=g)
    f0 = torch.randn(1, S0, 256, generator=g)
    perm = torch.randperm(S1, generator=g)[:160]
    rows = torch.randperm(S0, generator=g)[:160]
    gains = 0.6 + 0.4 * torch.rand(160, generator=g)
    f0[0, rows] = gains[:, None] * f1[0, perm] + 0.1 * torch.randn(160, 256, generator=g)
    data = {"hw0_i": torch.Size([h0 * 8, w0 * 8]), "hw1_i": torch.Size([h1 * 8, w1 * 8]),
            "hw0_c": torch.Size([h0, w0]), "hw1_c": torch.Size([h1, w1])}
    m0 = m1 = None
    if variant == "mask":
        mk0 = torch.ones(1, h0, w0, dtype=torch.bool)
        mk0[:, 13:] = False
        mk1 = torch.ones(1, h1, w1, dtype=torch.bool)
        mk1[:, :, 15:] = False
        data["mask0"], data["mask1"] = mk0, mk1
        m0, m1 = mk0.flatten(-2), mk1.flatten(-2)
    if variant == "scale":
        data["scale0"] = torch.tensor([[1.25, 0.5]])
        data["scale1"] = torch.tensor([[2.0, 1.5]])
    return dict(feat0=f0, feat1=f1, data=data, mask0=m0, mask1=m1, thr=0.2, border_rm=2, temperature=0.1)


def case_fine_preprocess(kind):
    """kind: 'onepose' (W=5) or 'loftr' (W=9).  Two batch elements, windows on the border too."""
    hc = wc = 8
    hf = wf = 32
    g = _g(80)
    M = 23 if kind == "onepose" else 9
    b_ids = torch.randint(0, 2, (M,), generator=g).sort()[0]
    j_ids = torch.randint(0, hc * wc, (M,), generator=g)
    j_ids[:4] = torch.tensor([0, wc - 1, hc * wc - wc, hc * wc - 1])  # four corners
    if kind == "onepose":
        L = 40
        i_ids = torch.randint(0, L, (M,), generator=g)
        data = {"b_ids": b_ids, "i_ids": i_ids, "j_ids": j_ids,
                "q_hw_f": torch.Size([hf, wf]), "q_hw_c": torch.Size([hc, wc])}
        return dict(data=data, desc3d_fine=randn(81, 2, 128, L), feat_f=randn(82, 2, 128, hf, wf), W=5)
    i_ids = torch.randint(0, hc * wc, (M,), generator=g)
    data = {"b_ids": b_ids, "i_ids": i_ids, "j_ids": j_ids,
            "hw0_f": torch.Size([hf, wf]), "hw0_c": torch.Size([hc, wc])}
    return dict(data=data, feat_f0=randn(83, 2, 128, hf, wf), feat_f1=randn(84, 2, 128, hf, wf), W=9)


def case_fine_matching(kind):
    M = 29
    g = _g(90)
    if kind == "onepose":
        WW = 25
        win = torch.randn(M, WW, 128, generator=g)
        r = torch.randint(0, WW, (M,), generator=g)
        q = win[torch.arange(M), r] * 0.35 + 0.05 * torch.randn(M, 128, generator=g)
        data = {"q_hw_i": torch.Size([64, 64]), "q_hw_f": torch.Size([32, 32]),
                "mkpts_query_c": torch.rand(M, 2, generator=g) * 64, "mkpts_3d_db": torch.randn(M, 3, generator=g),
                "b_ids": torch.zeros(M, dtype=torch.long)}
        return dict(feat3d=q[:, None, :], win=win, data=data)
    WW = 81
    f1 = torch.randn(M, WW, 128, generator=g)
    f0 = torch.randn(M, WW, 128, generator=g)
    r = torch.randint(0, WW, (M,), generator=g)
    f0[:, WW // 2] = f1[torch.arange(M), r] * 0.35 + 0.05 * torch.randn(M, 128, generator=g)
    data = {"hw0_i": torch.Size([64, 64]), "hw0_f": torch.Size([32, 32]),
            "mkpts0_c": torch.rand(M, 2, generator=g) * 64, "mkpts1_c": torch.rand(M, 2, generator=g) * 64,
            "mconf": torch.rand(M, generator=g), "b_ids": torch.zeros(M, dtype=torch.long),
            "scale0": torch.tensor([[1.25, 0.5]]), "scale1": torch.tensor([[2.0, 1.5]])}
    return dict(feat_f0=f0, feat_f1=f1, data=data)


def case_kpt_encoding():
    sd = R.init_onepose_plus_state_dict(seed=3)
    sd = {k: v for k, v in sd.items() if k.startswith("kpt_3d_pos_encoding.")}
    kpts = (rand(100, 2, 77, 3) - 0.5) * 0.3
    return dict(sd=sd, kpts=kpts, desc=randn(101, 2, 256, 77))


def case_backbone():
    sd = R.init_onepose_plus_state_dict(seed=4)
    sd = {k: v for k, v in sd.items() if k.startswith("backbone.")}
    return dict(sd=sd, image=rand(110, 2, 1, 64, 80))


# ------------------------------------------------------------------ model-level cases
def calibrated_onepose_sd(seed=0):
    """Random-init weights rescaled so that planted matches survive the (random) transformer:
    every encoder layer's norm2 (weight, bias) x0.05 and the keypoint-MLP output layer x0.05.
    Same recipe as ``onepose_st_b200.synthetic.calibrate_onepose_state_dict``."""
    sd = R.init_onepose_plus_state_dict(seed=seed)
    for k in sd:
        if ".norm2." in k:
            sd[k] = sd[k] * 0.05
    sd["kpt_3d_pos_encoding.encoder.9.weight"] = sd["kpt_3d_pos_encoding.encoder.9.weight"] * 0.05
    return sd


def whitened_plant(feat_c_pe, cells, temps):
    """Descriptors d_i = T_i * (X^T X + lam I)^-1 x_{c(i)} so that <d_i, x_j> peaks at j = c(i)
    even though random-conv features share a large common component.  feat_c_pe: [S,C]."""
    X = feat_c_pe.double()
    sg = X.t() @ X
    lam = 1e-3 * sg.diag().mean()
    wm = torch.linalg.inv(sg + lam * torch.eye(X.shape[1], dtype=torch.double))
    return (temps.double()[None, :] * (wm @ X[cells].t())).float()  # [C, n_plant]


def case_onepose_full(H=128, W=128, L=400, n_plant=150, seed=0, with_mask=False):
    cfg = copy.deepcopy(R.ONEPOSE_PLUS_CFG)
    sd = calibrated_onepose_sd(seed)
    img = rand(1000 + seed, 1, 1, H, W)
    with torch.no_grad():
        fc, ff = R.resnet_fpn_8_2(sd, "backbone.", img)
    hc, wc = fc.shape[2:]
    fcp = (fc + R.position_encoding_sine(256, hc, wc)).flatten(2)[0].t()  # [S,C]
    g = _g(7 + seed)
    kp = (torch.rand(1, L, 3, generator=g) - 0.5) * 0.3
    dc = torch.randn(1, 256, L, generator=g)
    df = torch.randn(1, 128, L, generator=g)
    cells = torch.randperm(hc * wc, generator=g)[:n_plant]
    pts = torch.randperm(L, generator=g)[:n_plant]
    # leverage of a cell under the whitening is ~C/S, so T = target_sim * (C*(T0+1e-4)) * S/C
    target_sim = 6 + 19 * torch.rand(n_plant, generator=g)
    dc[0, :, pts] = whitened_plant(fcp, cells, target_sim * (256 * 0.0801) * (hc * wc) / 256)
    yy, xx = cells // wc, cells % wc
    df[0, :, pts] = ff[0][:, yy * 4, xx * 4] * 0.05
    data = dict(query_image=img, keypoints3d=kp, descriptors3d_db=df, descriptors3d_coarse_db=dc)
    if with_mask:
        m = torch.ones(1, hc, wc, dtype=torch.bool)
        m[:, :, wc - 3:] = False
        data["query_image_mask"] = m
    return dict(cfg=cfg, sd=sd, data=data)


def calibrated_loftr_sd(seed=0, H=96, W=128, temp=400.0, lam_rel=1e-2):
    """Random-init LoFTR weights calibrated so that an image matched against itself gives mutual matches with
    confidences spread over (thr, 1) while the features stay O(1..10) (well conditioned in fp32):
    the 1x1 ``layer3_outconv`` is composed with the whitening map sqrt(temp) (X^T X + lam I)^-1/2 of the coarse
    features X of a seeded calibration image (random-conv features share a large common component; whitened,
    <f_i, f_j> is temp x the leverage matrix, which peaks on the diagonal), and norm2 x0.05 in both stacks keeps the
    random transformer from washing that structure out.  Same recipe as ``synthetic.calibrate_loftr_state_dict``."""
    sd = R.init_loftr_state_dict(seed=seed)
    for k in sd:
        if ".norm2." in k and (k.startswith("loftr_coarse.") or k.startswith("loftr_fine.")):
            sd[k] = sd[k] * 0.05
    sd["backbone.layer1_outconv2.3.weight"] = sd["backbone.layer1_outconv2.3.weight"] * 0.25  # fine logits O(10)
    with torch.no_grad():
        fc, _ = R.resnet_fpn_8_2(sd, "backbone.", rand(999, 1, 1, H, W))
    x = fc.flatten(2)[0].t().double()  # [S, C]
    sg = x.t() @ x
    ev, u = torch.linalg.eigh(sg + lam_rel * sg.diag().mean() * torch.eye(x.shape[1], dtype=torch.double))
    wm = (u * ev.rsqrt()[None, :]) @ u.t()
    w = sd["backbone.layer3_outconv.weight"][:, :, 0, 0].double()
    sd["backbone.layer3_outconv.weight"] = ((temp ** 0.5) * (wm @ w)).float()[:, :, None, None].contiguous()
    return sd


def case_loftr_full(H=96, W=128, seed=0, mode="full"):
    """mode: 'full' (coarse+fine), 'coarse' (enable_fine_matching=False), 'fine_only'
    (caller supplies mkpts0_c/mkpts1_c), 'extract' (fine_only + feature extraction)."""
    cfg = copy.deepcopy(R.LOFTR_CFG)
    sd = calibrated_loftr_sd(seed, H, W)
    img0 = rand(2000 + 2 * seed, 1, 1, H, W)
    img1 = img0.clone() if mode in ("full", "coarse") else rand(2001 + 2 * seed, 1, 1, H, W)
    data = dict(image0=img0, image1=img1, scale0=torch.tensor([[1.0, 1.0]]), scale1=torch.tensor([[1.0, 1.0]]))
    if mode in ("fine_only", "extract"):
        g = _g(2100 + seed)
        M = 31
        # below 0 exercises the reference's clip; the upper end stays < dim-4 because the reference
        # itself indexes out of bounds for points that round up to the last+1 coarse cell (loftr.py:87-99)
        k0 = torch.stack([torch.rand(M, generator=g) * (W - 1.5) - 3, torch.rand(M, generator=g) * (H - 1.5) - 3], 1)
        k1 = torch.stack([torch.rand(M, generator=g) * (W - 1.5) - 3, torch.rand(M, generator=g) * (H - 1.5) - 3], 1)
        data["mkpts0_c"], data["mkpts1_c"] = k0, k1
        data["scale0"] = torch.tensor([[1.0, 1.0]])
        data["scale1"] = torch.tensor([[1.0, 1.0]])
    return dict(cfg=cfg, sd=sd, data=data, mode=mode)
