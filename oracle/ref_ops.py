"""Functional torch-CPU restatement of the reference hot path (TEST INFRASTRUCTURE).

See ``oracle/__init__.py`` for the rules about who may import this.  All
citations are relative to /root/reference (mizeller/OnePose_ST @ 320dac4);
"OPP" = src/models/OnePosePlus, "LoFTR" = submodules/LoFTR/src/loftr.

Every function works on plain tensors and a flat ``sd`` (state dict with the
reference's parameter names).  Nothing here keeps state.
"""
from __future__ import annotations

import math
from typing import Dict, Optional, Sequence, Tuple

import torch
import torch.nn.functional as F

Tensor = torch.Tensor

__all__ = [
    "create_meshgrid", "spatial_expectation2d", "position_encoding_sine",
    "normalize_3d_keypoints", "keypoint_encoding_linear", "linear_attention",
    "encoder_layer", "transformer_onepose", "transformer_loftr", "resnet_fpn_8_2",
    "dual_softmax_conf", "coarse_select", "coarse_matching_onepose", "coarse_matching_loftr",
    "unfold_windows", "fine_preprocess_onepose", "fine_preprocess_loftr",
    "fine_heatmap_expectation", "fine_matching_onepose", "fine_matching_loftr",
    "sample_feature_from_featuremap", "onepose_plus_forward", "loftr_forward",
    "ONEPOSE_PLUS_CFG", "LOFTR_CFG", "init_onepose_plus_state_dict", "init_loftr_state_dict",
]


# --------------------------------------------------------------------------------------
# kornia==0.4.1 restatement (dependency absent from the image; requirements.txt pins 0.4.1).
# Call sites: OPP/utils/fine_matching.py:86-87, LoFTR/utils/fine_matching.py:49-50.
# --------------------------------------------------------------------------------------
def create_meshgrid(height: int, width: int, normalized_coordinates: bool = True,
                    device="cpu", dtype=torch.float32) -> Tensor:
    """kornia.utils.grid.create_meshgrid (0.4.1): returns [1,H,W,2], last dim (x, y).

    Normalised coordinates are ``(linspace(0, n-1, n) / (n-1) - 0.5) * 2``.
    """
    xs = torch.linspace(0, width - 1, width, device=device, dtype=dtype)
    ys = torch.linspace(0, height - 1, height, device=device, dtype=dtype)
    if normalized_coordinates:
        xs = (xs / (width - 1) - 0.5) * 2
        ys = (ys / (height - 1) - 0.5) * 2
    gx = xs.view(1, width).expand(height, width)
    gy = ys.view(height, 1).expand(height, width)
    return torch.stack([gx, gy], dim=-1).unsqueeze(0)


def spatial_expectation2d(heat: Tensor, normalized_coordinates: bool = True) -> Tensor:
    """kornia.geometry.subpix.dsnt.spatial_expectation2d (0.4.1): [B,N,H,W] -> [B,N,2] (x,y)."""
    b, n, h, w = heat.shape
    grid = create_meshgrid(h, w, normalized_coordinates, heat.device, heat.dtype)
    px = grid[..., 0].reshape(-1)
    py = grid[..., 1].reshape(-1)
    flat = heat.reshape(b, n, -1)
    ex = (flat * px).sum(-1, keepdim=True)
    ey = (flat * py).sum(-1, keepdim=True)
    return torch.cat([ex, ey], dim=-1)


# --------------------------------------------------------------------------------------
# Position encodings
# --------------------------------------------------------------------------------------
def position_encoding_sine(d_model: int, height: int, width: int, temp_bug_fix: bool = False,
                           dtype=torch.float32) -> Tensor:
    """PE slice ``pe[:, :, :height, :width]`` -> [1,C,H,W].

    OPP/utils/position_encoding.py:13-42 (always the floor-div form) and
    LoFTR/utils/position_encoding.py:11-42 with ``temp_bug_fix=False`` (cfg line 28).
    ``-log(1e4) / d_model // 2`` parses as ``(-log(1e4)/d_model) // 2 == -1.0``.
    Positions are 1-based (cumsum of ones).
    """
    ypos = torch.arange(1, height + 1, dtype=dtype).view(1, height, 1).expand(1, height, width)
    xpos = torch.arange(1, width + 1, dtype=dtype).view(1, 1, width).expand(1, height, width)
    k = torch.arange(0, d_model // 2, 2, dtype=dtype)
    if temp_bug_fix:
        factor = -math.log(10000.0) / (d_model // 2)
    else:
        factor = -math.log(10000.0) / d_model // 2
    div = torch.exp(k * factor).view(-1, 1, 1)
    pe = torch.zeros(d_model, height, width, dtype=dtype)
    pe[0::4] = torch.sin(xpos * div)
    pe[1::4] = torch.cos(xpos * div)
    pe[2::4] = torch.sin(ypos * div)
    pe[3::4] = torch.cos(ypos * div)
    return pe.unsqueeze(0)


def normalize_3d_keypoints(kpts: Tensor) -> Tensor:
    """OPP/utils/normalize.py:17-28.  Extent comes from batch element 0 only."""
    extent = kpts[0].max(dim=0).values - kpts[0].min(dim=0).values  # [3]
    centre = kpts.mean(dim=-2)  # [B,3]
    scaling = extent.max() * 0.6
    return (kpts - centre[:, None, :]) / scaling


def keypoint_encoding_linear(sd: Dict[str, Tensor], prefix: str, kpts: Tensor, desc: Tensor,
                             norm_method: str = "instancenorm") -> Tensor:
    """OPP/utils/position_encoding.py:46-79: ``desc + MLP(kpts)^T``; desc is [B,C,L].

    With ``norm_method="instancenorm"`` the reference applies ``nn.InstanceNorm1d(c)``
    to a [B,L,c] tensor, i.e. it normalises over the channel dim per point
    (biased variance, eps 1e-5, no affine).
    """
    idx = sorted({int(k[len(prefix):].split(".")[0]) for k in sd if k.startswith(prefix)})
    x = kpts
    for n, li in enumerate(idx):
        x = F.linear(x, sd[f"{prefix}{li}.weight"], sd[f"{prefix}{li}.bias"])
        if n < len(idx) - 1:
            if norm_method == "instancenorm":
                # InstanceNorm1d on [B,L,c]: dim 1 (the L points) plays "channel", the
                # statistics run over the last dim (biased variance, eps 1e-5, no affine)
                x = F.instance_norm(x, eps=1e-5)
            elif norm_method == "layernorm":
                x = F.layer_norm(x, x.shape[-1:], sd[f"{prefix}{li + 1}.weight"], sd[f"{prefix}{li + 1}.bias"])
            else:
                raise NotImplementedError(norm_method)
            x = F.relu(x)
    return desc + x.transpose(2, 1)


# --------------------------------------------------------------------------------------
# Linear attention + encoder layer + the two transformer schedules
# --------------------------------------------------------------------------------------
def linear_attention(q: Tensor, k: Tensor, v: Tensor, q_mask: Optional[Tensor] = None,
                     kv_mask: Optional[Tensor] = None, eps: float = 1e-6) -> Tensor:
    """OPP/loftr_module/linear_attention.py:29-61 (LoFTR copy :20-47 is the same math).

    q [N,L,H,D]; k,v [N,S,H,D] -> [N,L,H,D].
    """
    Q = F.elu(q) + 1
    K = F.elu(k) + 1
    if q_mask is not None:
        Q = Q * q_mask[:, :, None, None]
    if kv_mask is not None:
        K = K * kv_mask[:, :, None, None]
        v = v * kv_mask[:, :, None, None]
    s_len = v.size(1)
    v = v / s_len
    KV = torch.einsum("nshd,nshv->nhdv", K, v)
    Z = 1 / (torch.einsum("nlhd,nhd->nlh", Q, K.sum(dim=1)) + eps)
    out = torch.einsum("nlhd,nhdv,nlh->nlhv", Q, KV, Z) * s_len
    return out.contiguous()


def encoder_layer(sd: Dict[str, Tensor], prefix: str, x: Tensor, source: Tensor, nhead: int,
                  x_mask: Optional[Tensor] = None, source_mask: Optional[Tensor] = None) -> Tensor:
    """OPP/loftr_module/transformer.py:65-94; LoFTR/loftr_module/transformer.py:35-58."""
    bs, _, c = x.shape
    d = c // nhead
    q = F.linear(x, sd[prefix + "q_proj.weight"]).view(bs, -1, nhead, d)
    k = F.linear(source, sd[prefix + "k_proj.weight"]).view(bs, -1, nhead, d)
    v = F.linear(source, sd[prefix + "v_proj.weight"]).view(bs, -1, nhead, d)
    msg = linear_attention(q, k, v, x_mask, source_mask)
    msg = F.linear(msg.view(bs, -1, c), sd[prefix + "merge.weight"])
    msg = F.layer_norm(msg, (c,), sd[prefix + "norm1.weight"], sd[prefix + "norm1.bias"], 1e-5)
    msg = F.linear(torch.cat([x, msg], dim=2), sd[prefix + "mlp.0.weight"])
    msg = F.linear(F.relu(msg), sd[prefix + "mlp.2.weight"])
    msg = F.layer_norm(msg, (c,), sd[prefix + "norm2.weight"], sd[prefix + "norm2.bias"], 1e-5)
    return x + msg


def transformer_onepose(sd: Dict[str, Tensor], prefix: str, desc3d_cf: Tensor, desc2d: Tensor,
                        layer_names: Sequence[str], nhead: int,
                        query_mask: Optional[Tensor] = None) -> Tuple[Tensor, Tensor]:
    """OPP/loftr_module/transformer.py:133-171.

    ``desc3d_cf`` is channels-first [N,C,L]; returns (desc3d [N,L,C], desc2d [N,P,C]).
    One layer object serves both streams and both streams read the pre-update tensors.
    """
    d3 = desc3d_cf.permute(0, 2, 1)
    d2 = desc2d
    for i, name in enumerate(layer_names):
        p = f"{prefix}layers.{i}."
        if name == "self":
            n2 = encoder_layer(sd, p, d2, d2, nhead, query_mask, query_mask)
            n3 = encoder_layer(sd, p, d3, d3, nhead)
        elif name == "cross":
            n2 = encoder_layer(sd, p, d2, d3, nhead, x_mask=query_mask)
            n3 = encoder_layer(sd, p, d3, d2, nhead, source_mask=query_mask)
        else:
            raise NotImplementedError(name)
        d2, d3 = n2, n3
    return d3, d2


def transformer_loftr(sd: Dict[str, Tensor], prefix: str, feat0: Tensor, feat1: Tensor,
                      layer_names: Sequence[str], nhead: int, mask0: Optional[Tensor] = None,
                      mask1: Optional[Tensor] = None) -> Tuple[Tensor, Tensor]:
    """LoFTR/loftr_module/transformer.py:80-101: the cross step is sequential."""
    for i, name in enumerate(layer_names):
        p = f"{prefix}layers.{i}."
        if name == "self":
            feat0 = encoder_layer(sd, p, feat0, feat0, nhead, mask0, mask0)
            feat1 = encoder_layer(sd, p, feat1, feat1, nhead, mask1, mask1)
        elif name == "cross":
            feat0 = encoder_layer(sd, p, feat0, feat1, nhead, mask0, mask1)
            feat1 = encoder_layer(sd, p, feat1, feat0, nhead, mask1, mask0)
        else:
            raise KeyError(name)
    return feat0, feat1


# --------------------------------------------------------------------------------------
# ResNetFPN_8_2 backbone (eval mode)
# --------------------------------------------------------------------------------------
def _bn(sd, p, x):
    return F.batch_norm(x, sd[p + "running_mean"], sd[p + "running_var"], sd[p + "weight"],
                        sd[p + "bias"], False, 0.0, 1e-5)


def _basic_block(sd, p, x, stride):
    y = F.relu(_bn(sd, p + "bn1.", F.conv2d(x, sd[p + "conv1.weight"], None, stride, 1)))
    y = _bn(sd, p + "bn2.", F.conv2d(y, sd[p + "conv2.weight"], None, 1, 1))
    if stride != 1:
        x = _bn(sd, p + "downsample.1.", F.conv2d(x, sd[p + "downsample.0.weight"], None, stride, 0))
    return F.relu(x + y)


def _fpn_head(sd, p, x):
    y = F.conv2d(x, sd[p + "0.weight"], None, 1, 1)
    y = F.leaky_relu(_bn(sd, p + "1.", y), 0.01)
    return F.conv2d(y, sd[p + "3.weight"], None, 1, 1)


def resnet_fpn_8_2(sd: Dict[str, Tensor], prefix: str, x: Tensor) -> Tuple[Tensor, Tensor]:
    """OPP/backbone/resnet.py:141-164 and LoFTR/backbone/resnet_fpn.py:100-118.

    Returns (coarse 1/8 [N,256,H/8,W/8], fine 1/2 [N,128,H/2,W/2]).
    """
    p = prefix
    x0 = F.relu(_bn(sd, p + "bn1.", F.conv2d(x, sd[p + "conv1.weight"], None, 2, 3)))
    x1 = _basic_block(sd, p + "layer1.1.", _basic_block(sd, p + "layer1.0.", x0, 1), 1)
    x2 = _basic_block(sd, p + "layer2.1.", _basic_block(sd, p + "layer2.0.", x1, 2), 1)
    x3 = _basic_block(sd, p + "layer3.1.", _basic_block(sd, p + "layer3.0.", x2, 2), 1)
    x3_out = F.conv2d(x3, sd[p + "layer3_outconv.weight"])
    up3 = F.interpolate(x3_out, scale_factor=2.0, mode="bilinear", align_corners=True)
    x2_out = _fpn_head(sd, p + "layer2_outconv2.", F.conv2d(x2, sd[p + "layer2_outconv.weight"]) + up3)
    up2 = F.interpolate(x2_out, scale_factor=2.0, mode="bilinear", align_corners=True)
    x1_out = _fpn_head(sd, p + "layer1_outconv2.", F.conv2d(x1, sd[p + "layer1_outconv.weight"]) + up2)
    return x3_out, x1_out


# --------------------------------------------------------------------------------------
# Coarse matching
# --------------------------------------------------------------------------------------
def dual_softmax_conf(feat_a: Tensor, feat_b: Tensor, temperature: float, temp_eps: float,
                      mask_a: Optional[Tensor] = None, mask_b: Optional[Tensor] = None,
                      additive_mask: bool = False) -> Tensor:
    """conf = softmax(sim, dim=1) * softmax(sim, dim=2), sim = (a/sqrt C)(b/sqrt C)^T / (T + eps).

    OPP/utils/coarse_matching.py:102-115 (``temp_eps=1e-4``; mask added as -1e9, only the
    query side is maskable) and LoFTR/utils/coarse_matching.py:109-119 (``temp_eps=0``;
    ``masked_fill_(-1e9)`` with both masks).
    """
    c = feat_a.shape[-1]
    a = feat_a / c ** 0.5
    b = feat_b / c ** 0.5
    sim = torch.einsum("nlc,nsc->nls", a, b) / (temperature + temp_eps)
    if mask_b is not None or mask_a is not None:
        n, l, s = sim.shape
        ma = mask_a if mask_a is not None else torch.ones(n, l, dtype=torch.bool, device=sim.device)
        mb = mask_b if mask_b is not None else torch.ones(n, s, dtype=torch.bool, device=sim.device)
        valid = (ma[..., None] * mb[:, None]).bool()
        if additive_mask:
            neg = torch.zeros_like(sim)
            neg[~valid] = -1e9
            sim = sim + neg
        else:
            sim = sim.masked_fill(~valid, -1e9)
    return F.softmax(sim, 1) * F.softmax(sim, 2)


def coarse_select(conf: Tensor, thr: float, border_keep: Tensor) -> Tuple[Tensor, Tensor, Tensor, Tensor]:
    """Threshold + border + mutual nearest neighbour on a materialised conf matrix.

    OPP/utils/coarse_matching.py:145-172; LoFTR/utils/coarse_matching.py:172-195.
    ``border_keep`` is a bool tensor broadcastable to conf ([N,L,S]) that is False where the
    reference's ``mask_border`` writes False.  Ties inside a row resolve to the first j
    (``mask.max(dim=2)``); output is in ascending (b, i) order (``torch.where``).
    """
    mask = (conf > thr) & border_keep
    mask = mask & (conf == conf.max(dim=2, keepdim=True)[0]) & (conf == conf.max(dim=1, keepdim=True)[0])
    mask_v, all_j = mask.max(dim=2)
    b_ids, i_ids = torch.where(mask_v)
    j_ids = all_j[b_ids, i_ids]
    mconf = conf[b_ids, i_ids, j_ids]
    return b_ids, i_ids, j_ids, mconf


def _border_keep_onepose(n, l, hc, wc, b, device):
    """OPP/utils/coarse_matching.py:10-20: ``m[:, :, -b:0]`` is an empty slice, so only the
    first ``b`` rows / columns of the query grid are removed."""
    keep = torch.ones(1, 1, hc, wc, dtype=torch.bool, device=device)
    if b > 0:
        keep[:, :, :b] = False
        keep[:, :, :, :b] = False
    return keep.reshape(1, 1, hc * wc)


def _border_keep_loftr(h0, w0, h1, w1, b, device, mask0=None, mask1=None, n=1):
    """LoFTR/utils/coarse_matching.py:8-43: four-sided border on both grids; with padding
    masks the far border is measured from the valid extent of each image."""
    keep = torch.ones(n, h0, w0, h1, w1, dtype=torch.bool, device=device)
    if b > 0:
        keep[:, :b] = False
        keep[:, :, :b] = False
        keep[:, :, :, :b] = False
        keep[:, :, :, :, :b] = False
        if mask0 is None:
            keep[:, -b:] = False
            keep[:, :, -b:] = False
            keep[:, :, :, -b:] = False
            keep[:, :, :, :, -b:] = False
        else:
            h0s, w0s = mask0.sum(1).max(-1)[0].int(), mask0.sum(-1).max(-1)[0].int()
            h1s, w1s = mask1.sum(1).max(-1)[0].int(), mask1.sum(-1).max(-1)[0].int()
            for bi in range(n):
                keep[bi, int(h0s[bi]) - b:] = False
                keep[bi, :, int(w0s[bi]) - b:] = False
                keep[bi, :, :, int(h1s[bi]) - b:] = False
                keep[bi, :, :, :, int(w1s[bi]) - b:] = False
    return keep.reshape(n, h0 * w0, h1 * w1)


def coarse_matching_onepose(feat3d: Tensor, feat2d: Tensor, data: dict, thr: float = 0.1,
                            border_rm: int = 2, temperature: float = 0.08,
                            mask_query: Optional[Tensor] = None) -> None:
    """OPP/utils/coarse_matching.py:76-242 in eval mode (training-time GT padding omitted)."""
    n, l, _ = feat3d.shape
    conf = dual_softmax_conf(feat3d, feat2d, temperature, 1e-4, None, mask_query, additive_mask=True)
    data["conf_matrix"] = conf
    hc, wc = int(data["q_hw_c"][0]), int(data["q_hw_c"][1])
    keep = _border_keep_onepose(n, l, hc, wc, border_rm, conf.device)
    b_ids, i_ids, j_ids, mconf = coarse_select(conf, thr, keep)
    scale = data["q_hw_i"][0] / data["q_hw_c"][0]
    if "query_image_scale" in data:
        scale = scale * data["query_image_scale"][b_ids][:, [1, 0]]
    mk_q = torch.stack([j_ids % wc, j_ids // wc], dim=1) * scale
    mk_3d = data["keypoints3d"][b_ids, i_ids]
    nz = mconf != 0
    data.update(b_ids=b_ids, i_ids=i_ids, j_ids=j_ids, gt_mask=mconf == 0, m_bids=b_ids[nz],
                mkpts_3d_db=mk_3d[nz], mkpts_query_c=mk_q[nz], mconf=mconf[nz])


def coarse_matching_loftr(feat0: Tensor, feat1: Tensor, data: dict, thr: float = 0.2,
                          border_rm: int = 2, temperature: float = 0.1,
                          mask0: Optional[Tensor] = None, mask1: Optional[Tensor] = None) -> None:
    """LoFTR/utils/coarse_matching.py:87-261 (dual_softmax, eval mode)."""
    n = feat0.shape[0]
    conf = dual_softmax_conf(feat0, feat1, temperature, 0.0, mask0, mask1, additive_mask=False)
    data["conf_matrix"] = conf
    h0, w0 = int(data["hw0_c"][0]), int(data["hw0_c"][1])
    h1, w1 = int(data["hw1_c"][0]), int(data["hw1_c"][1])
    keep = _border_keep_loftr(h0, w0, h1, w1, border_rm, conf.device,
                              data.get("mask0") if "mask0" in data else None,
                              data.get("mask1") if "mask0" in data else None, n)
    b_ids, i_ids, j_ids, mconf = coarse_select(conf, thr, keep)
    scale = data["hw0_i"][0] / data["hw0_c"][0]
    scale0 = scale * data["scale0"][b_ids] if "scale0" in data else scale
    scale1 = scale * data["scale1"][b_ids] if "scale1" in data else scale
    mk0 = torch.stack([i_ids % w0, i_ids // w0], dim=1) * scale0
    mk1 = torch.stack([j_ids % w1, j_ids // w1], dim=1) * scale1
    nz = mconf != 0
    data.update(b_ids=b_ids, i_ids=i_ids, j_ids=j_ids, gt_mask=mconf == 0, m_bids=b_ids[nz],
                mkpts0_c=mk0[nz], mkpts1_c=mk1[nz], mconf=mconf[nz])


# --------------------------------------------------------------------------------------
# Fine level
# --------------------------------------------------------------------------------------
def unfold_windows(feat_f: Tensor, W: int, stride: int) -> Tensor:
    """``F.unfold(feat, (W,W), stride, padding=W//2)`` rearranged 'n (c ww) l -> n l ww c'.

    OPP/loftr_module/fine_preprocess.py:47-48; LoFTR/loftr_module/fine_preprocess.py:40-43.
    """
    n, c = feat_f.shape[:2]
    u = F.unfold(feat_f, kernel_size=(W, W), stride=stride, padding=W // 2)  # [N, C*WW, L]
    return u.view(n, c, W * W, -1).permute(0, 3, 2, 1)


def fine_preprocess_onepose(data: dict, desc3d_fine: Tensor, feat_f: Tensor, W: int,
                            d_model: int) -> Tuple[Tensor, Tensor]:
    """OPP/loftr_module/fine_preprocess.py:32-55 -> ([M,Cf,1], [M,WW,Cf])."""
    data["W"] = W
    if data["b_ids"].shape[0] == 0:
        return (torch.empty(0, d_model, 1, device=desc3d_fine.device),
                torch.empty(0, W * W, d_model, device=desc3d_fine.device))
    stride = data["q_hw_f"][0] // data["q_hw_c"][0]
    win = unfold_windows(feat_f, W, stride)
    f3d = desc3d_fine.permute(0, 2, 1)[data["b_ids"], data["i_ids"], :].unsqueeze(-1)
    return f3d, win[data["b_ids"], data["j_ids"]]


def fine_preprocess_loftr(data: dict, feat_f0: Tensor, feat_f1: Tensor, W: int,
                          d_model: int) -> Tuple[Tensor, Tensor]:
    """LoFTR/loftr_module/fine_preprocess.py:29-59 with ``fine_concat_coarse_feat=False``."""
    data["W"] = W
    if data["b_ids"].shape[0] == 0:
        e = torch.empty(0, W * W, d_model, device=feat_f0.device)
        return e, e.clone()
    stride = data["hw0_f"][0] // data["hw0_c"][0]
    w0 = unfold_windows(feat_f0, W, stride)[data["b_ids"], data["i_ids"]]
    w1 = unfold_windows(feat_f1, W, stride)[data["b_ids"], data["j_ids"]]
    return w0, w1


def fine_heatmap_expectation(query_vec: Tensor, win: Tensor) -> Tensor:
    """Correlation -> softmax -> spatial expectation + std.  Returns expec_f [M,3] = (x, y, std).

    OPP/utils/fine_matching.py:78-94; LoFTR/utils/fine_matching.py:43-57.
    """
    m, ww, c = win.shape
    w = int(math.sqrt(ww))
    sim = torch.einsum("mc,mrc->mr", query_vec, win)
    heat = torch.softmax(sim / c ** 0.5, dim=1).view(-1, w, w)
    coords = spatial_expectation2d(heat[None], True)[0]
    grid = create_meshgrid(w, w, True, heat.device).reshape(1, -1, 2)
    var = torch.sum(grid ** 2 * heat.view(-1, ww, 1), dim=1) - coords ** 2
    std = torch.sum(torch.sqrt(torch.clamp(var, min=1e-10)), -1)
    return torch.cat([coords, std.unsqueeze(1)], -1)


def fine_matching_onepose(feat3d: Tensor, win: Tensor, data: dict) -> None:
    """OPP/utils/fine_matching.py:28-110.  feat3d [M,1,Cf], win [M,WW,Cf]."""
    m, ww, c = win.shape
    w = int(math.sqrt(ww))
    scale = data["q_hw_i"][0] / data["q_hw_f"][0]
    if m == 0:
        data.update(expec_f=torch.empty(0, 3, device=feat3d.device),
                    mkpts_3d_db=data["mkpts_3d_db"], mkpts_query_f=data["mkpts_query_c"])
        return
    assert feat3d.shape[1] % 2 == 1
    expec = fine_heatmap_expectation(feat3d[:, feat3d.shape[1] // 2, :], win)
    data["expec_f"] = expec
    coords = expec[:, :2]
    qs = scale * data["query_image_scale"][data["b_ids"]][:, [1, 0]] if "query_image_scale" in data else scale
    data["mkpts_query_f"] = data["mkpts_query_c"] + (coords * (w // 2) * qs)[: len(data["mkpts_query_c"])]


def fine_matching_loftr(feat_f0: Tensor, feat_f1: Tensor, data: dict) -> None:
    """LoFTR/utils/fine_matching.py:15-74 (incl. the ``'scale0' in data`` test for scale1)."""
    m, ww, c = feat_f0.shape
    w = int(math.sqrt(ww))
    scale = data["hw0_i"][0] / data["hw0_f"][0]
    if m == 0:
        data.update(expec_f=torch.empty(0, 3, device=feat_f0.device),
                    mkpts0_f=data["mkpts0_c"], mkpts1_f=data["mkpts1_c"])
        return
    expec = fine_heatmap_expectation(feat_f0[:, ww // 2, :], feat_f1)
    data["expec_f"] = expec
    coords = expec[:, :2]
    scale1 = scale * data["scale1"][data["b_ids"]] if "scale0" in data else scale
    data["mkpts0_f"] = data["mkpts0_c"]
    data["mkpts1_f"] = data["mkpts1_c"] + (coords * (w // 2) * scale1)[: len(data["mconf"])]


def sample_feature_from_featuremap(feature_map: Tensor, kpts: Tensor, imghw: Tensor,
                                   sample_mode: str = "bilinear") -> Tensor:
    """src/KeypointFreeSfM/loftr_for_sfm/utils/sample_feature_from_featuremap.py:28-81
    (point mode: ``patch_feature_size=None``, ``norm_feature=False``) -> [L, C]."""
    assert feature_map.dim() == 4 and feature_map.shape[0] == 1
    h, w = imghw[0], imghw[1]
    grid = kpts[None, :, None, :] - 1 / 2 + 0.5
    rescale = torch.stack([(w - 1) * 1, (h - 1) * 1]).to(grid)
    grid = grid / rescale[None, None, None] * 2 - 1
    out = F.grid_sample(feature_map, grid.float(), mode=sample_mode, align_corners=True)  # 1,C,L,1
    return out[0, :, :, 0].t()


# --------------------------------------------------------------------------------------
# Model-level forwards
# --------------------------------------------------------------------------------------
ONEPOSE_PLUS_CFG = {
    # configs/experiment/inference_demo.yaml:11-95, with loftr_backbone.pretrained=None
    "loftr_backbone": {"type": "ResNetFPN", "resolution": [8, 2],
                       "resnetfpn": {"block_type": "BasicBlock", "initial_dim": 128,
                                     "block_dims": [128, 196, 256], "output_layers": [3, 1]},
                       "pretrained": None, "pretrained_fix": False},
    "interpol_type": "bilinear",
    "keypoints_encoding": {"enable": True, "type": "mlp_linear", "descriptor_dim": 256,
                           "keypoints_encoder": [32, 64, 128], "norm_method": "instancenorm"},
    "positional_encoding": {"enable": True, "pos_emb_shape": [256, 256]},
    "loftr_coarse": {"type": "LoFTR", "d_model": 256, "d_ffm": 128, "nhead": 8,
                     "layer_names": ["self", "cross"], "layer_iter_n": 3, "dropout": 0.0,
                     "attention": "linear", "norm_method": "layernorm", "kernel_fn": "elu + 1",
                     "d_kernel": 16, "redraw_interval": 2, "rezero": None, "final_proj": False},
    "coarse_matching": {"type": "dual-softmax", "thr": 0.1, "feat_norm_method": "sqrt_feat_dim",
                        "border_rm": 2, "dual_softmax": {"temperature": 0.08},
                        "train": {"train_padding": True, "train_coarse_percent": 0.3,
                                  "train_pad_num_gt_min": 200}},
    "loftr_fine": {"enable": True, "window_size": 5, "coarse_layer_norm": False, "type": "LoFTR",
                   "d_model": 128, "nhead": 8, "layer_names": ["self", "cross"], "layer_iter_n": 1,
                   "dropout": 0.0, "attention": "linear", "norm_method": "layernorm",
                   "kernel_fn": "elu + 1", "d_kernel": 16, "redraw_interval": 2, "rezero": None,
                   "final_proj": False},
    "fine_matching": {"enable": True, "type": "s2d", "s2d": {"type": "heatmap"}},
}

LOFTR_CFG = {
    # src/KeypointFreeSfM/loftr_for_sfm/utils/loftr_for_onepose_plus_cfg.py:10-50 after lower_config
    "backbone_type": "ResNetFPN", "resolution": (8, 2), "fine_window_size": 9,
    "fine_concat_coarse_feat": False,
    "resnetfpn": {"initial_dim": 128, "block_dims": [128, 196, 256]},
    "coarse": {"d_model": 256, "d_ffn": 256, "nhead": 8, "layer_names": ["self", "cross"] * 4,
               "attention": "linear", "temp_bug_fix": False},
    "match_coarse": {"thr": 0.2, "border_rm": 2, "match_type": "dual_softmax",
                     "dsmax_temperature": 0.1, "skh_iters": 3, "skh_init_bin_score": 1.0,
                     "skh_prefilter": True, "train_coarse_percent": 0.4, "train_pad_num_gt_min": 200},
    "fine": {"d_model": 128, "d_ffn": 128, "nhead": 8, "layer_names": ["self", "cross"],
             "attention": "linear"},
}


def onepose_plus_forward(sd: Dict[str, Tensor], cfg: dict, data: dict) -> None:
    """OPP/OnePosePlusModel.py:95-203 (eval mode, fine matching enabled)."""
    img = data["query_image"]
    data.update(bs=img.size(0), q_hw_i=img.shape[2:])
    feat_c, feat_f = resnet_fpn_8_2(sd, "backbone.", img)
    data.update(q_hw_c=feat_c.shape[2:], q_hw_f=feat_f.shape[2:])
    cc = cfg["loftr_coarse"]
    if cfg["positional_encoding"]["enable"]:
        feat_c = feat_c + position_encoding_sine(cc["d_model"], feat_c.shape[2], feat_c.shape[3]).to(feat_c)
    q2d = feat_c.flatten(2).permute(0, 2, 1)
    desc_c = data["descriptors3d_coarse_db"] if "descriptors3d_coarse_db" in data else data["descriptors3d_db"]
    if cfg["keypoints_encoding"]["enable"]:
        kn = normalize_3d_keypoints(data["keypoints3d"])
        desc_c = keypoint_encoding_linear(sd, "kpt_3d_pos_encoding.encoder.", kn, desc_c,
                                          cfg["keypoints_encoding"]["norm_method"])
    qmask = data["query_image_mask"].flatten(-2) if "query_image_mask" in data else None
    names = list(cc["layer_names"]) * cc["layer_iter_n"]
    d3, q2d = transformer_onepose(sd, "loftr_coarse.", desc_c, q2d, names, cc["nhead"], qmask)
    cm = cfg["coarse_matching"]
    coarse_matching_onepose(d3, q2d, data, cm["thr"], cm["border_rm"], cm["dual_softmax"]["temperature"], qmask)
    if not cfg["fine_matching"]["enable"]:
        data["mkpts_query_f"] = data["mkpts_query_c"]
        return
    cf = cfg["loftr_fine"]
    f3d, win = fine_preprocess_onepose(data, data["descriptors3d_db"], feat_f, cf["window_size"], cf["d_model"])
    if win.size(0) != 0 and cf["enable"]:
        fnames = list(cf["layer_names"]) * cf["layer_iter_n"]
        f3d, win = transformer_onepose(sd, "loftr_fine.", f3d, win, fnames, cf["nhead"])
    else:
        f3d = f3d.permute(0, 2, 1)
    fine_matching_onepose(f3d, win, data)


def loftr_forward(sd: Dict[str, Tensor], cfg: dict, data: dict, enable_fine_matching: bool = True,
                  extract_coarse_feature: bool = False, extract_fine_feature: bool = False) -> None:
    """src/KeypointFreeSfM/loftr_for_sfm/loftr.py:34-166."""
    i0, i1 = data["image0"], data["image1"]
    data.update(bs=i0.size(0), hw0_i=i0.shape[2:], hw1_i=i1.shape[2:])
    if data["hw0_i"] == data["hw1_i"]:
        fc, ff = resnet_fpn_8_2(sd, "backbone.", torch.cat([i0, i1], dim=0))
        (fc0, fc1), (ff0, ff1) = fc.split(data["bs"]), ff.split(data["bs"])
    else:
        (fc0, ff0), (fc1, ff1) = resnet_fpn_8_2(sd, "backbone.", i0), resnet_fpn_8_2(sd, "backbone.", i1)
    data.update(hw0_c=fc0.shape[2:], hw1_c=fc1.shape[2:], hw0_f=ff0.shape[2:], hw1_f=ff1.shape[2:])
    fc0_b, fc1_b, ff0_b, ff1_b = fc0, fc1, ff0, ff1
    cc, mc = cfg["coarse"], cfg["match_coarse"]
    if "mkpts0_c" not in data:
        c = cc["d_model"]
        s0 = (fc0 + position_encoding_sine(c, fc0.shape[2], fc0.shape[3], cc["temp_bug_fix"]).to(fc0)).flatten(2).permute(0, 2, 1)
        s1 = (fc1 + position_encoding_sine(c, fc1.shape[2], fc1.shape[3], cc["temp_bug_fix"]).to(fc1)).flatten(2).permute(0, 2, 1)
        m0 = m1 = None
        if "mask0" in data:
            m0, m1 = data["mask0"].flatten(-2), data["mask1"].flatten(-2)
        s0, s1 = transformer_loftr(sd, "loftr_coarse.", s0, s1, cc["layer_names"], cc["nhead"], m0, m1)
        coarse_matching_loftr(s0, s1, data, mc["thr"], mc["border_rm"], mc["dsmax_temperature"], m0, m1)
    else:
        # fine-only entry, loftr.py:79-115 (clips the caller's tensors in place)
        k0, k1 = data["mkpts0_c"], data["mkpts1_c"]
        b_ids = torch.zeros((k0.shape[0],), device=k0.device).long()
        k0[:, 0] = torch.clip(k0[:, 0], min=0, max=data["hw0_i"][1] - 2)
        k0[:, 1] = torch.clip(k0[:, 1], min=0, max=data["hw0_i"][0] - 2)
        k1[:, 0] = torch.clip(k1[:, 0], min=0, max=data["hw1_i"][1] - 2)
        k1[:, 1] = torch.clip(k1[:, 1], min=0, max=data["hw1_i"][0] - 2)
        scale = data["hw0_i"][0] / data["hw0_c"][0]
        sc0 = scale * data["scale0"][b_ids][:, [1, 0]] if "scale0" in data else scale
        sc1 = scale * data["scale1"][b_ids][:, [1, 0]] if "scale1" in data else scale
        r0, r1 = torch.round(k0 / sc0), torch.round(k1 / sc1)
        i_ids = (r0[:, 1] * data["hw0_c"][1] + r0[:, 0]).long()
        j_ids = (r1[:, 1] * data["hw1_c"][1] + r1[:, 0]).long()
        data.update(m_bids=b_ids, b_ids=b_ids, i_ids=i_ids, j_ids=j_ids, mconf=torch.ones_like(b_ids))
    if enable_fine_matching:
        fcfg = cfg["fine"]
        w0, w1 = fine_preprocess_loftr(data, ff0, ff1, cfg["fine_window_size"], fcfg["d_model"])
        if w0.size(0) != 0:
            w0, w1 = transformer_loftr(sd, "loftr_fine.", w0, w1, fcfg["layer_names"], fcfg["nhead"])
        fine_matching_loftr(w0, w1, data)
    else:
        data.update(mkpts0_f=data["mkpts0_c"], mkpts1_f=data["mkpts1_c"])
    if extract_coarse_feature:
        hw0 = data["scale0"].squeeze(0) * torch.tensor(data["hw0_i"]).to(data["scale0"])
        hw1 = data["scale1"].squeeze(0) * torch.tensor(data["hw1_i"]).to(data["scale1"])
        data["feat_coarse_b_0"] = sample_feature_from_featuremap(fc0_b, data["mkpts0_f"], hw0, "nearest")
        data["feat_coarse_b_1"] = sample_feature_from_featuremap(fc1_b, data["mkpts1_f"], hw1, "nearest")
    if extract_fine_feature:
        hw0 = data["scale0"].squeeze(0) * torch.tensor(data["hw0_i"]).to(data["scale0"])
        hw1 = data["scale1"].squeeze(0) * torch.tensor(data["hw1_i"]).to(data["scale1"])
        data["feat_ext0"] = sample_feature_from_featuremap(ff0_b, data["mkpts0_f"], hw0)
        data["feat_ext1"] = sample_feature_from_featuremap(ff1_b, data["mkpts1_f"], hw1)


# --------------------------------------------------------------------------------------
# Seeded random-init weights (checkpoints are not in the reference tree, README.md:44).
# Shapes / names are the reference's; values are the oracle's own deterministic draw so
# the same weights can be rebuilt on a machine that has no /root/reference.
# --------------------------------------------------------------------------------------
def _xavier(g, out_f, in_f):
    a = math.sqrt(6.0 / (in_f + out_f))
    return (torch.rand(out_f, in_f, generator=g) * 2 - 1) * a


def _kaiming_conv(g, out_c, in_c, k):
    std = math.sqrt(2.0 / (out_c * k * k))
    return torch.randn(out_c, in_c, k, k, generator=g) * std


def _init_bn(sd, p, c, g):
    sd[p + "weight"] = 1 + 0.1 * torch.randn(c, generator=g)
    sd[p + "bias"] = 0.1 * torch.randn(c, generator=g)
    sd[p + "running_mean"] = 0.1 * torch.randn(c, generator=g)
    sd[p + "running_var"] = 1 + 0.1 * torch.rand(c, generator=g)
    sd[p + "num_batches_tracked"] = torch.tensor(0, dtype=torch.long)


def _init_backbone(sd, p, g, initial_dim=128, block_dims=(128, 196, 256)):
    sd[p + "conv1.weight"] = _kaiming_conv(g, initial_dim, 1, 7)
    _init_bn(sd, p + "bn1.", initial_dim, g)
    in_c = initial_dim
    for li, (dim, stride) in enumerate(zip(block_dims, (1, 2, 2)), start=1):
        for bi in range(2):
            q = f"{p}layer{li}.{bi}."
            st = stride if bi == 0 else 1
            sd[q + "conv1.weight"] = _kaiming_conv(g, dim, in_c, 3)
            sd[q + "conv2.weight"] = _kaiming_conv(g, dim, dim, 3)
            _init_bn(sd, q + "bn1.", dim, g)
            _init_bn(sd, q + "bn2.", dim, g)
            if st != 1:
                sd[q + "downsample.0.weight"] = _kaiming_conv(g, dim, in_c, 1)
                _init_bn(sd, q + "downsample.1.", dim, g)
            in_c = dim
    d1, d2, d3 = block_dims
    sd[p + "layer3_outconv.weight"] = _kaiming_conv(g, d3, d3, 1)
    sd[p + "layer2_outconv.weight"] = _kaiming_conv(g, d3, d2, 1)
    sd[p + "layer2_outconv2.0.weight"] = _kaiming_conv(g, d3, d3, 3)
    _init_bn(sd, p + "layer2_outconv2.1.", d3, g)
    sd[p + "layer2_outconv2.3.weight"] = _kaiming_conv(g, d2, d3, 3)
    sd[p + "layer1_outconv.weight"] = _kaiming_conv(g, d2, d1, 1)
    sd[p + "layer1_outconv2.0.weight"] = _kaiming_conv(g, d2, d2, 3)
    _init_bn(sd, p + "layer1_outconv2.1.", d2, g)
    sd[p + "layer1_outconv2.3.weight"] = _kaiming_conv(g, d1, d2, 3)


def _init_transformer(sd, p, g, d_model, n_layers):
    for i in range(n_layers):
        q = f"{p}layers.{i}."
        for nm in ("q_proj", "k_proj", "v_proj", "merge"):
            sd[q + nm + ".weight"] = _xavier(g, d_model, d_model)
        sd[q + "mlp.0.weight"] = _xavier(g, 2 * d_model, 2 * d_model)
        sd[q + "mlp.2.weight"] = _xavier(g, d_model, 2 * d_model)
        for nm in ("norm1", "norm2"):
            sd[q + nm + ".weight"] = 1 + 0.1 * torch.randn(d_model, generator=g)
            sd[q + nm + ".bias"] = 0.1 * torch.randn(d_model, generator=g)


def init_onepose_plus_state_dict(cfg: dict = None, seed: int = 0) -> Dict[str, Tensor]:
    """State dict with the parameter names of ``OnePosePlus_model`` (OPP/OnePosePlusModel.py:25-93)."""
    cfg = cfg or ONEPOSE_PLUS_CFG
    g = torch.Generator().manual_seed(seed)
    sd: Dict[str, Tensor] = {}
    rf = cfg["loftr_backbone"]["resnetfpn"]
    _init_backbone(sd, "backbone.", g, rf["initial_dim"], tuple(rf["block_dims"]))
    ke = cfg["keypoints_encoding"]
    chans = [3] + list(ke["keypoints_encoder"]) + [ke["descriptor_dim"]]
    for n in range(1, len(chans)):
        li = (n - 1) * 3  # Linear, InstanceNorm1d, ReLU triplets
        bound = 1 / math.sqrt(chans[n - 1])
        sd[f"kpt_3d_pos_encoding.encoder.{li}.weight"] = (torch.rand(chans[n], chans[n - 1], generator=g) * 2 - 1) * bound
        sd[f"kpt_3d_pos_encoding.encoder.{li}.bias"] = (torch.rand(chans[n], generator=g) * 2 - 1) * bound
    sd[f"kpt_3d_pos_encoding.encoder.{(len(chans) - 2) * 3}.bias"].zero_()
    cc, cf = cfg["loftr_coarse"], cfg["loftr_fine"]
    _init_transformer(sd, "loftr_coarse.", g, cc["d_model"], len(cc["layer_names"]) * cc["layer_iter_n"])
    _init_transformer(sd, "loftr_fine.", g, cf["d_model"], len(cf["layer_names"]) * cf["layer_iter_n"])
    return sd


def init_loftr_state_dict(cfg: dict = None, seed: int = 0) -> Dict[str, Tensor]:
    """State dict with the parameter names of ``LoFTR_for_OnePose_Plus`` (loftr.py:17-32)."""
    cfg = cfg or LOFTR_CFG
    g = torch.Generator().manual_seed(seed)
    sd: Dict[str, Tensor] = {}
    rf = cfg["resnetfpn"]
    _init_backbone(sd, "backbone.", g, rf["initial_dim"], tuple(rf["block_dims"]))
    _init_transformer(sd, "loftr_coarse.", g, cfg["coarse"]["d_model"], len(cfg["coarse"]["layer_names"]))
    _init_transformer(sd, "loftr_fine.", g, cfg["fine"]["d_model"], len(cfg["fine"]["layer_names"]))
    return sd
