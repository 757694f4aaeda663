"""Drop-in 2D-2D matcher used by Keypoint-Free SfM: the API of
``src/KeypointFreeSfM/loftr_for_sfm/loftr.py`` + ``submodules/LoFTR/src/loftr`` on the sm_100a kernels.
Inference only."""
import math

import torch
import torch.nn as nn
import torch.nn.functional as F

from . import ops
from .nn_common import (HeadGraphs, LinearAttention, LoFTREncoderLayer, PositionEncodingSine as _PE, ResNetFPN_8_2, _TransformerBase,
                        fine_expectation)

__all__ = ["LoFTR_for_OnePose_Plus", "LocalFeatureTransformer", "CoarseMatching", "FinePreprocess", "FineMatching",
           "PositionEncodingSine", "default_cfg", "lower_config", "sample_feature_from_featuremap", "build_backbone"]

# src/KeypointFreeSfM/loftr_for_sfm/utils/loftr_for_onepose_plus_cfg.py:10-50 after lower_config()
default_cfg = {
    "backbone_type": "ResNetFPN", "resolution": (8, 2), "fine_window_size": 9, "fine_concat_coarse_feat": False,
    "resnetfpn": {"initial_dim": 128, "block_dims": [128, 196, 256]},
    "coarse": {"d_model": 256, "d_ffn": 256, "nhead": 8, "layer_names": ["self", "cross"] * 4, "attention": "linear",
               "temp_bug_fix": False},
    "match_coarse": {"thr": 0.2, "border_rm": 2, "match_type": "dual_softmax", "dsmax_temperature": 0.1, "skh_iters": 3,
                     "skh_init_bin_score": 1.0, "skh_prefilter": True, "train_coarse_percent": 0.4,
                     "train_pad_num_gt_min": 200},
    "fine": {"d_model": 128, "d_ffn": 128, "nhead": 8, "layer_names": ["self", "cross"], "attention": "linear"},
}


def lower_config(cfg):
    """Accepts a yacs CfgNode or a plain dict (loftr_for_onepose_plus_cfg.py:4-7)."""
    if not hasattr(cfg, "items"):
        return cfg
    return {str(k).lower(): lower_config(v) for k, v in cfg.items()}


def build_backbone(config):
    """submodules/LoFTR/src/loftr/backbone/__init__.py:4-11."""
    if config["backbone_type"] == "ResNetFPN":
        if tuple(config["resolution"]) == (8, 2):
            return ResNetFPN_8_2(config["resnetfpn"])
        raise NotImplementedError("only the (8, 2) ResNetFPN resolution is supported")
    raise ValueError(f"LOFTR.BACKBONE_TYPE {config['backbone_type']} not supported.")


class PositionEncodingSine(_PE):
    """submodules/LoFTR/src/loftr/utils/position_encoding.py:6-42 (default temp_bug_fix=True there)."""

    def __init__(self, d_model, max_shape=(256, 256), temp_bug_fix=True):
        super().__init__(d_model, max_shape, temp_bug_fix)


class LocalFeatureTransformer(_TransformerBase):
    """submodules/LoFTR/src/loftr/loftr_module/transformer.py:61-101 (cross step is sequential)."""
    schedule = 1

    def __init__(self, config):
        super().__init__()
        self.config = config
        self._build(config["d_model"], config["nhead"], config["layer_names"], attention=config["attention"])

    def forward(self, feat0, feat1, mask0=None, mask1=None, _add0=None, _add1=None, _planes_only=False):
        assert self.d_model == feat0.size(2), "the feature number of src and transformer must be equal"
        return self._run(feat0, feat1, mask0, mask1, _add0, _add1, planes_only=_planes_only)


class CoarseMatching(nn.Module):
    """submodules/LoFTR/src/loftr/utils/coarse_matching.py:59-261 (dual_softmax, eval mode)."""

    def __init__(self, config):
        super().__init__()
        self.config = config
        self.thr = config["thr"]
        self.border_rm = config["border_rm"]
        self.train_coarse_percent = config["train_coarse_percent"]
        self.train_pad_num_gt_min = config["train_pad_num_gt_min"]
        self.match_type = config["match_type"]
        if self.match_type == "dual_softmax":
            self.temperature = config["dsmax_temperature"]
        elif self.match_type == "sinkhorn":
            raise NotImplementedError("match_type 'sinkhorn' is never configured by the reference and is not implemented")
        else:
            raise NotImplementedError()
        self.materialize_conf_matrix = bool(config.get("materialize_conf_matrix", False))
        self.precision = ops.PREC_BF16X3

    def forward(self, feat_c0, feat_c1, data, mask_c0=None, mask_c1=None):
        self.finish(self.launch(feat_c0, feat_c1, data, mask_c0, mask_c1), data)

    def launch(self, feat_c0, feat_c1, data, mask_c0=None, mask_c1=None, _status=None) -> "ops.CoarseLaunch":
        """Enqueue the matching kernels (no host sync; needs data['hw0_c'] / data['hw1_c'])."""
        if self.training:
            raise NotImplementedError("training-time coarse matching is out of scope of the inference path")
        planes = feat_c0 if isinstance(feat_c0, ops.FeaturePlanes) else None  # internal fused hand-over
        C = planes.c if planes is not None else feat_c0.shape[2]
        h0, w0 = int(data["hw0_c"][0]), int(data["hw0_c"][1])
        h1, w1 = int(data["hw1_c"][0]), int(data["hw1_c"][1])
        vhw = None
        if "mask0" in data:  # mask_border_with_padding: valid extents of both images
            k0, k1 = data["mask0"], data["mask1"]
            vhw = torch.stack([k0.sum(1).max(-1)[0], k0.sum(-1).max(-1)[0], k1.sum(1).max(-1)[0], k1.sum(-1).max(-1)[0]], 1).int()
        if (mask_c0 is None) != (mask_c1 is None):
            raise ValueError("mask_c0 and mask_c1 must be given together")
        return ops.coarse_match_launch(
            feat_c0, feat_c1, scale=1.0 / (C * self.temperature), thr=self.thr, border_rm=self.border_rm, border_mode=1,
            grid_a=(h0, w0), grid_b=(h1, w1), mask_a=mask_c0, mask_b=mask_c1, mask_fill=1, valid_hw=vhw,
            materialize_conf=self.materialize_conf_matrix, precision=self.precision, planes=planes, status=_status)

    def finish(self, handle: "ops.CoarseLaunch", data, copy: bool = False):
        """Read the match count (the forward's host sync) and fill the reference's coarse-level ``data`` keys."""
        w0, w1 = int(data["hw0_c"][1]), int(data["hw1_c"][1])
        b_ids, i_ids, j_ids, mconf, conf = ops.coarse_match_finish(handle, copy=copy)
        if conf is not None:
            data.update({"conf_matrix": conf})
        scale = data["hw0_i"][0] / data["hw0_c"][0]
        scale0 = scale * data["scale0"][b_ids] if "scale0" in data else scale
        scale1 = scale * data["scale1"][b_ids] if "scale1" in data else scale
        mkpts0_c = torch.stack([i_ids % w0, i_ids // w0], dim=1) * scale0
        mkpts1_c = torch.stack([j_ids % w1, j_ids // w1], dim=1) * scale1
        data.update({"b_ids": b_ids, "i_ids": i_ids, "j_ids": j_ids, "gt_mask": mconf == 0, "m_bids": b_ids,
                     "mkpts0_c": mkpts0_c, "mkpts1_c": mkpts1_c, "mconf": mconf})


class FinePreprocess(nn.Module):
    """submodules/LoFTR/src/loftr/loftr_module/fine_preprocess.py:7-59 with fine_concat_coarse_feat=False."""

    def __init__(self, config):
        super().__init__()
        self.config = config
        self.cat_c_feat = config["fine_concat_coarse_feat"]
        if self.cat_c_feat:
            raise NotImplementedError("fine_concat_coarse_feat=True is not used by the reference configuration")
        self.W = config["fine_window_size"]
        self.d_model_f = config["fine"]["d_model"]

    def forward(self, feat_f0, feat_f1, feat_c0, feat_c1, data):
        W = self.W
        data.update({"W": W})
        if data["b_ids"].shape[0] == 0:
            e = torch.empty(0, W ** 2, self.d_model_f, device=feat_f0.device)
            return e, torch.empty_like(e)
        stride = data["hw0_f"][0] // data["hw0_c"][0]
        flag = data.setdefault("_opb200_err_flag", torch.zeros(1, dtype=torch.int32, device=feat_f0.device))
        w0 = ops.fine_gather(feat_f0, data["b_ids"], data["i_ids"], W, stride, int(data["hw0_c"][0]), int(data["hw0_c"][1]), flag)
        w1 = ops.fine_gather(feat_f1, data["b_ids"], data["j_ids"], W, stride, int(data["hw1_c"][0]), int(data["hw1_c"][1]), flag)
        return w0, w1


class FineMatching(nn.Module):
    """submodules/LoFTR/src/loftr/utils/fine_matching.py:9-74."""

    def forward(self, feat_f0, feat_f1, data):
        M, WW, C = feat_f0.shape
        W = int(math.sqrt(WW))
        scale = data["hw0_i"][0] / data["hw0_f"][0]
        self.M, self.W, self.WW, self.C, self.scale = M, W, WW, C, scale
        if M == 0:
            assert self.training is False, "M is always >0, when training, see coarse_matching.py"
            data.update({"expec_f": torch.empty(0, 3, device=feat_f0.device), "mkpts0_f": data["mkpts0_c"],
                         "mkpts1_f": data["mkpts1_c"]})
            return
        expec = fine_expectation(feat_f0[:, WW // 2, :], feat_f1)
        data.update({"expec_f": expec})
        coords = expec[:, :2]
        # the reference tests 'scale0' (not 'scale1') to decide whether scale1 applies (fine_matching.py:68)
        scale1 = scale * data["scale1"][data["b_ids"]] if "scale0" in data else scale
        mkpts1_f = data["mkpts1_c"] + (coords * (W // 2) * scale1)[: len(data["mconf"])]
        data.update({"mkpts0_f": data["mkpts0_c"], "mkpts1_f": mkpts1_f})


def sample_feature_from_featuremap(feature_map, kpts, imghw, norm_feature=False, patch_feature_size=None,
                                   sample_mode="bilinear"):
    """src/KeypointFreeSfM/loftr_for_sfm/utils/sample_feature_from_featuremap.py:28-81, point mode
    (what loftr.py:130-166 calls): M-point sampling of the pre-transformer backbone maps on the own kernel
    (csrc/fine.cu sample_features_kernel: the reference's fp32 coordinate arithmetic + grid_sample align_corners=True)."""
    if patch_feature_size is not None:
        raise NotImplementedError("patch sampling is not used by the matcher")
    if feature_map.dim() == 4:
        assert feature_map.shape[0] == 1, "batch should to be 1"
    else:
        feature_map = feature_map.unsqueeze(0)
    if feature_map.is_cuda and sample_mode in ("bilinear", "nearest"):
        return ops.sample_features(feature_map[0], kpts, float(imghw[0]), float(imghw[1]), sample_mode == "bilinear", norm_feature)
    h, w = imghw[0], imghw[1]
    grid = kpts[None, :, None, :] - 1 / 2 + 0.5
    rescale = torch.stack([(w - 1) * 1, (h - 1) * 1]).to(grid)
    grid = grid / rescale[None, None, None] * 2 - 1
    out = F.grid_sample(feature_map, grid.float(), mode=sample_mode, align_corners=True)[0, :, :, 0].t()
    return F.normalize(out, p=2, dim=-1) if norm_feature else out


class LoFTR_for_OnePose_Plus(nn.Module):
    """src/KeypointFreeSfM/loftr_for_sfm/loftr.py:16-166."""

    def __init__(self, config, enable_fine_matching=True):
        super().__init__()
        config = lower_config(config)
        self.config = config
        self.enable_fine_matching = enable_fine_matching
        self.backbone = build_backbone(config)
        self.pos_encoding = PositionEncodingSine(config["coarse"]["d_model"], temp_bug_fix=config["coarse"]["temp_bug_fix"])
        self.loftr_coarse = LocalFeatureTransformer(config["coarse"])
        self.coarse_matching = CoarseMatching(config["match_coarse"])
        self.fine_preprocess = FinePreprocess(config)
        self.loftr_fine = LocalFeatureTransformer(config["fine"])
        self.fine_matching = FineMatching()
        self._graphs = HeadGraphs()

    # CUDA graph of the head (nn_common.HeadGraphs): opt-in, no padding masks, weights not modified in place
    cuda_graph = False

    def __getstate__(self):
        st = self.__dict__.copy()
        st["_graphs"] = HeadGraphs()
        return st

    def _head(self, data, image0, image1, coarse, conv_precision=None):
        """Backbone of both images (+ coarse LoFTR stack and coarse matching kernels when ``coarse``): everything before
        the forward's host sync.  Returns ((feat_c0, feat_c1, feat_f0, feat_f1), launch handle | status words, hw dict) and
        sets the hw*_c / hw*_f keys of ``data``."""
        # status words of the forward's one device -> host copy: [0..2] coarse matching, [3] fp16 range flag of the backbone
        status = torch.zeros(4, dtype=torch.int32, device=image0.device) if image0.is_cuda else None
        bkw = dict(_precision=conv_precision, _range_flag=None if status is None else status[3:])
        if image0.shape[2:] == image1.shape[2:]:
            feats_c, feats_f = self.backbone(torch.cat([image0, image1], dim=0), **bkw)
            (feat_c0, feat_c1), (feat_f0, feat_f1) = feats_c.split(image0.size(0)), feats_f.split(image0.size(0))
        else:
            (feat_c0, feat_f0), (feat_c1, feat_f1) = self.backbone(image0, **bkw), self.backbone(image1, **bkw)
        hw = {"hw0_c": tuple(feat_c0.shape[2:]), "hw1_c": tuple(feat_c1.shape[2:]), "hw0_f": tuple(feat_f0.shape[2:]),
              "hw1_f": tuple(feat_f1.shape[2:])}
        data.update({"hw0_c": feat_c0.shape[2:], "hw1_c": feat_c1.shape[2:], "hw0_f": feat_f0.shape[2:],
                     "hw1_f": feat_f1.shape[2:]})
        handle = status  # fine-only entry: no coarse launch, the caller reads the range flag from the status words
        if coarse:
            s0 = feat_c0.flatten(2).permute(0, 2, 1)
            s1 = feat_c1.flatten(2).permute(0, 2, 1)
            add0 = self.pos_encoding.token_table(*feat_c0.shape[2:])
            add1 = add0 if feat_c1.shape[2:] == feat_c0.shape[2:] else self.pos_encoding.token_table(*feat_c1.shape[2:]).clone()
            mask_c0 = mask_c1 = None
            if "mask0" in data:
                mask_c0, mask_c1 = data["mask0"].flatten(-2), data["mask1"].flatten(-2)
            planes = self.loftr_coarse(s0, s1, mask_c0, mask_c1, _add0=add0, _add1=add1, _planes_only=True)
            handle = self.coarse_matching.launch(planes, None, data, mask_c0=mask_c0, mask_c1=mask_c1, _status=status)
        return (feat_c0, feat_c1, feat_f0, feat_f1), handle, hw

    @torch.no_grad()
    def forward(self, data, **kwargs):
        if self.training:
            raise NotImplementedError("onepose_st_b200 implements the inference forward only; call .eval()")
        data.update({"bs": data["image0"].size(0), "hw0_i": data["image0"].shape[2:], "hw1_i": data["image1"].shape[2:]})
        coarse = "mkpts0_c" not in data
        if self.cuda_graph and coarse and "mask0" not in data and data["image0"].is_cuda:
            key = (self.coarse_matching.precision, getattr(self.loftr_coarse, "precision", ops.PREC_BF16X3),
                   self.backbone.conv_precision, self.coarse_matching.materialize_conf_matrix)
            proto = {k: data[k] for k in ("bs", "hw0_i", "hw1_i")}
            feats, handle, hw = self._graphs.run(key, (data["image0"], data["image1"]),
                                                 lambda i0, i1: self._head(dict(proto), i0, i1, True))
            feat_c0, feat_c1, feat_f0, feat_f1 = feats
            data.update({k: torch.Size(v) for k, v in hw.items()})
            self.coarse_matching.finish(handle, data, copy=True)
        else:
            (feat_c0, feat_c1, feat_f0, feat_f1), handle, _ = self._head(data, data["image0"], data["image1"], coarse)
            if coarse:
                self.coarse_matching.finish(handle, data)
        out_of_range = (handle.extra and handle.extra[0]) if coarse else (handle is not None and int(handle[3].item()))
        if out_of_range:
            # the fp16 backbone stored a value beyond the fp16 range: repeat the head with exact-fp32 convolutions
            if not self.__dict__.get("_warned_fp16_range"):
                import logging
                logging.getLogger("onepose_st_b200").warning(
                    "backbone activations left the fp16 range; repeating this forward with conv_precision='fp32' "
                    "(set model.backbone.conv_precision = 'fp32' to avoid the repeated work)")
                self.__dict__["_warned_fp16_range"] = True
            (feat_c0, feat_c1, feat_f0, feat_f1), handle, _ = self._head(data, data["image0"], data["image1"], coarse, "fp32")
            if coarse:
                self.coarse_matching.finish(handle, data)
        if not coarse:
            # fine-only entry (loftr.py:79-115): clips the caller's tensors in place, batch size 1
            k0, k1 = data["mkpts0_c"], data["mkpts1_c"]
            b_ids = torch.zeros((k0.shape[0],), device=k0.device).long()
            k0[:, 0] = torch.clip(k0[:, 0], min=0, max=data["hw0_i"][1] - 2)
            k0[:, 1] = torch.clip(k0[:, 1], min=0, max=data["hw0_i"][0] - 2)
            k1[:, 0] = torch.clip(k1[:, 0], min=0, max=data["hw1_i"][1] - 2)
            k1[:, 1] = torch.clip(k1[:, 1], min=0, max=data["hw1_i"][0] - 2)
            scale = data["hw0_i"][0] / data["hw0_c"][0]
            scale0 = scale * data["scale0"][b_ids][:, [1, 0]] if "scale0" in data else scale
            scale1 = scale * data["scale1"][b_ids][:, [1, 0]] if "scale1" in data else scale
            r0, r1 = torch.round(k0 / scale0), torch.round(k1 / scale1)
            i_ids = (r0[:, 1] * data["hw0_c"][1] + r0[:, 0]).long()
            j_ids = (r1[:, 1] * data["hw1_c"][1] + r1[:, 0]).long()
            feat_c0_t = feat_c1_t = None
            data.update({"m_bids": b_ids, "b_ids": b_ids, "i_ids": i_ids, "j_ids": j_ids, "mconf": torch.ones_like(b_ids)})
        if self.enable_fine_matching:
            w0, w1 = self.fine_preprocess(feat_f0, feat_f1, None, None, data)
            if w0.size(0) != 0:
                w0, w1 = self.loftr_fine(w0, w1)
            self.fine_matching(w0, w1, data)
            flag = data.pop("_opb200_err_flag", None)
            if flag is not None and "mkpts0_c" in data and int(flag.item()):
                raise IndexError("coarse keypoints fall outside the coarse grid (the reference raises IndexError too)")
        else:
            data.update({"mkpts0_f": data["mkpts0_c"], "mkpts1_f": data["mkpts1_c"]})
        if kwargs.get("extract_coarse_feature"):
            hw0 = data["scale0"].squeeze(0) * torch.tensor(data["hw0_i"]).to(data["scale0"])
            hw1 = data["scale1"].squeeze(0) * torch.tensor(data["hw1_i"]).to(data["scale1"])
            data.update({"feat_coarse_b_0": sample_feature_from_featuremap(feat_c0, data["mkpts0_f"], hw0, sample_mode="nearest"),
                         "feat_coarse_b_1": sample_feature_from_featuremap(feat_c1, data["mkpts1_f"], hw1, sample_mode="nearest")})
        if kwargs.get("extract_fine_feature"):
            hw0 = data["scale0"].squeeze(0) * torch.tensor(data["hw0_i"]).to(data["scale0"])
            hw1 = data["scale1"].squeeze(0) * torch.tensor(data["hw1_i"]).to(data["scale1"])
            data.update({"feat_ext0": sample_feature_from_featuremap(feat_f0, data["mkpts0_f"], hw0),
                         "feat_ext1": sample_feature_from_featuremap(feat_f1, data["mkpts1_f"], hw1)})
