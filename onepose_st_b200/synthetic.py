"""Seeded synthetic workloads of the shapes BASELINE.json names.

Random-init weights find zero matches on random inputs (max confidence ~1e-4), which would leave the
fine stage and the match gather idle.  The generators here therefore *plant* correspondences: the 3D
model's coarse descriptors are whitened copies of the query's own coarse features at chosen cells, its
fine descriptors are the query's fine features at the same places, and a few weights are rescaled so
the (random) transformer keeps the planted structure.  The same recipe generates the golden fixtures
(tests/golden/cases.py) on CPU; here it runs on the device through the product's own backbone.
"""
import copy
from typing import Dict, Optional

import torch

# configs/experiment/inference_demo.yaml:11-95 of the reference, with loftr_backbone.pretrained=None
ONEPOSE_PLUS_CFG = {
    "loftr_backbone": {"type": "ResNetFPN", "resolution": [8, 2],
                       "resnetfpn": {"block_type": "BasicBlock", "initial_dim": 128, "block_dims": [128, 196, 256],
                                     "output_layers": [3, 1]},
                       "pretrained": None, "pretrained_fix": False},
    "interpol_type": "bilinear",
    "keypoints_encoding": {"enable": True, "type": "mlp_linear", "descriptor_dim": 256,
                           "keypoints_encoder": [32, 64, 128], "norm_method": "instancenorm"},
    "positional_encoding": {"enable": True, "pos_emb_shape": [256, 256]},
    "loftr_coarse": {"type": "LoFTR", "d_model": 256, "d_ffm": 128, "nhead": 8, "layer_names": ["self", "cross"],
                     "layer_iter_n": 3, "dropout": 0.0, "attention": "linear", "norm_method": "layernorm",
                     "kernel_fn": "elu + 1", "d_kernel": 16, "redraw_interval": 2, "rezero": None, "final_proj": False},
    "coarse_matching": {"type": "dual-softmax", "thr": 0.1, "feat_norm_method": "sqrt_feat_dim", "border_rm": 2,
                        "dual_softmax": {"temperature": 0.08},
                        "train": {"train_padding": True, "train_coarse_percent": 0.3, "train_pad_num_gt_min": 200}},
    "loftr_fine": {"enable": True, "window_size": 5, "coarse_layer_norm": False, "type": "LoFTR", "d_model": 128,
                   "nhead": 8, "layer_names": ["self", "cross"], "layer_iter_n": 1, "dropout": 0.0, "attention": "linear",
                   "norm_method": "layernorm", "kernel_fn": "elu + 1", "d_kernel": 16, "redraw_interval": 2,
                   "rezero": None, "final_proj": False},
    "fine_matching": {"enable": True, "type": "s2d", "s2d": {"type": "heatmap"}},
}


def onepose_plus_cfg() -> dict:
    return copy.deepcopy(ONEPOSE_PLUS_CFG)


def calibrate_onepose_state_dict(sd: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
    """Every encoder layer's norm2 (weight, bias) x0.05 and the keypoint-MLP output layer x0.05, so that
    planted matches survive the random-init transformer (same recipe as tests/golden/cases.py)."""
    out = {}
    for k, v in sd.items():
        v = v.clone()
        if ".norm2." in k:
            v = v * 0.05
        if k == "kpt_3d_pos_encoding.encoder.9.weight":
            v = v * 0.05
        out[k] = v
    return out


@torch.no_grad()
def calibrate_loftr_state_dict(sd: Dict[str, torch.Tensor], height: int = 512, width: int = 512,
                               gain: float = 3.5, lam_rel: float = 1e-2) -> Dict[str, torch.Tensor]:
    """Random-init LoFTR weights calibrated so that an image matched against a noisy copy of itself gives mutual matches
    with confidences spread over (thr, 1) while the features stay O(1..10) (same recipe as tests/golden/cases.py):
    ``layer3_outconv`` (1x1) is composed with the whitening map sqrt(T) (X^T X + lam I)^-1/2 of the coarse features X of
    a seeded calibration image of the given size, T = gain x the number of coarse cells (random-conv features share a
    large common component; whitened, <f_i, f_j> is T x the leverage matrix, which peaks on the diagonal); norm2 x0.05
    in both stacks keeps the random transformer from washing that structure out; layer1_outconv2.3 x0.25 keeps the fine
    logits O(10).  Runs the backbone once on the CPU (plain PyTorch modules)."""
    from .nn_common import ResNetFPN_8_2
    out = {}
    for k, v in sd.items():
        v = v.detach().clone().cpu()
        if ".norm2." in k and (k.startswith("loftr_coarse.") or k.startswith("loftr_fine.")):
            v = v * 0.05
        if k == "backbone.layer1_outconv2.3.weight":
            v = v * 0.25
        out[k] = v
    bb = ResNetFPN_8_2({"block_type": "BasicBlock", "initial_dim": 128, "block_dims": [128, 196, 256]})
    bb.load_state_dict({k[len("backbone."):]: v for k, v in out.items() if k.startswith("backbone.")}, strict=True)
    g = torch.Generator().manual_seed(999)
    fc, _ = bb.eval()(torch.rand(1, 1, height, width, generator=g))
    x = fc.flatten(2)[0].t().double()  # [S, C]
    sg = x.t() @ x
    ev, u = torch.linalg.eigh(sg + lam_rel * sg.diag().mean() * torch.eye(x.shape[1], dtype=torch.double))
    wm = (u * ev.rsqrt()[None, :]) @ u.t()
    w = out["backbone.layer3_outconv.weight"][:, :, 0, 0].double()
    temp = gain * x.shape[0]
    out["backbone.layer3_outconv.weight"] = ((temp ** 0.5) * (wm @ w)).float()[:, :, None, None].contiguous()
    dev = next(iter(sd.values())).device
    return {k: v.to(dev) for k, v in out.items()}


def _whitened(feat_sc: torch.Tensor, cells: torch.Tensor, gains: torch.Tensor) -> torch.Tensor:
    """d_i = g_i (X^T X + lam I)^-1 x_{c(i)}: <d_i, x_j> peaks at j = c(i) although random-conv features
    share a large common component.  feat_sc [S, C] -> [C, n]."""
    x = feat_sc.double()
    sg = x.t() @ x
    lam = 1e-3 * sg.diag().mean()
    wm = torch.linalg.inv(sg + lam * torch.eye(x.shape[1], dtype=torch.double, device=x.device))
    return (gains.double()[None, :] * (wm @ x[cells].t())).float()


@torch.no_grad()
def make_onepose_sequence(model, n_points: int, n_frames: int, height: int = 512, width: int = 512,
                          n_plant: int = 1000, seed: int = 0, frame_noise: float = 0.02,
                          device: Optional[torch.device] = None) -> dict:
    """One 3D model (``n_points`` points) and ``n_frames`` query crops of it.

    Frame f = clamp(base + frame_noise * N(0,1); seed 1000+f) -- a static object seen with sensor noise,
    so the planted 2D-3D correspondences hold in every frame while no two frames are bit-identical.
    Returns {'keypoints3d' [1,L,3], 'descriptors3d_db' [1,128,L], 'descriptors3d_coarse_db' [1,256,L],
    'frames' [F,1,H,W] (on ``device``), 'planted_points' [n], 'planted_cells' [n]}."""
    dev = device or next(model.parameters()).device
    g = torch.Generator().manual_seed(1000 + seed)
    base = torch.rand(1, 1, height, width, generator=g).to(dev)
    fc, ff = model.backbone(base)
    hc, wc = fc.shape[2:]
    pe = model.dense_pos_encoding.token_table(hc, wc) if model.dense_pos_encoding is not None else 0.0
    feat_sc = fc.flatten(2)[0].t() + pe  # [S, C]
    g = torch.Generator().manual_seed(7 + seed)
    kp = ((torch.rand(1, n_points, 3, generator=g) - 0.5) * 0.3).to(dev)
    dc = torch.randn(1, 256, n_points, generator=g).to(dev)
    df = torch.randn(1, 128, n_points, generator=g).to(dev)
    interior = torch.tensor([r * wc + c for r in range(2, hc) for c in range(2, wc)])
    n_plant = min(n_plant, n_points, interior.numel())
    cells = interior[torch.randperm(interior.numel(), generator=g)[:n_plant]].to(dev)
    pts = torch.randperm(n_points, generator=g)[:n_plant].to(dev)
    target_sim = (6 + 19 * torch.rand(n_plant, generator=g)).to(dev)
    temp = model.config["coarse_matching"]["dual_softmax"]["temperature"] + 1e-4
    dc[0][:, pts] = _whitened(feat_sc, cells, target_sim * (256 * temp) * (hc * wc) / 256)
    yy, xx = cells // wc, cells % wc
    stride = ff.shape[2] // hc
    df[0][:, pts] = ff[0][:, yy * stride, xx * stride] * 0.05
    frames = torch.empty(n_frames, 1, height, width, device=dev)
    for f in range(n_frames):
        gf = torch.Generator().manual_seed(1000 + seed + 1 + f)
        noise = torch.randn(1, height, width, generator=gf).to(dev)
        frames[f] = (base[0] + frame_noise * noise).clamp_(0, 1)
    return {"keypoints3d": kp, "descriptors3d_db": df, "descriptors3d_coarse_db": dc, "frames": frames,
            "planted_points": pts, "planted_cells": cells}


def planted_coarse_pair(seed: int, n_batch: int, L: int, S: int, C: int, n_plant: int, hc: int, wc: int,
                        noise: float = 0.1, device="cpu"):
    """Stage-level CoarseMatching input: f2d ~ N(0,1); f3d[i] = f2d[pi(i)] +
    noise*N(0,1) for planted i (pi a random injection into interior cells), N(0,1) otherwise."""
    g = torch.Generator().manual_seed(seed)
    f2d = torch.randn(n_batch, S, C, generator=g)
    f3d = torch.randn(n_batch, L, C, generator=g)
    interior = torch.tensor([r * wc + c for r in range(2, hc) for c in range(2, wc)])
    n_plant = min(n_plant, L, interior.numel())
    for b in range(n_batch):
        cells = interior[torch.randperm(interior.numel(), generator=g)[:n_plant]]
        pts = torch.randperm(L, generator=g)[:n_plant]
        f3d[b, pts] = f2d[b, cells] + noise * torch.randn(n_plant, C, generator=g)
    return f3d.to(device), f2d.to(device)


def planted_fine_windows(seed: int, M: int, W: int, Cf: int = 128, device="cpu"):
    """Stage-level FineMatching input: win ~ N(0,1) [M,WW,Cf]; the query vector is a
    noisy copy of a random window row."""
    g = torch.Generator().manual_seed(seed)
    ww = W * W
    win = torch.randn(M, ww, Cf, generator=g)
    r = torch.randint(0, ww, (M,), generator=g)
    q = win[torch.arange(M), r] + 0.05 * torch.randn(M, Cf, generator=g)
    return q.to(device), win.to(device), r.to(device)
