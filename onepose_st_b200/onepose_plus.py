"""Drop-in 2D-3D matcher: the reference's ``src/models/OnePosePlus`` operator API on the sm_100a kernels.

Class names, constructor / forward signatures, config keys, ``state_dict`` names and the ``data`` dict
contract follow the reference (file:line in each docstring).  Inference only (``eval`` mode): training
branches of the reference (GT padding in CoarseMatching, losses) are out of scope and raise.
"""
import logging
import math
from math import log

import threading

import torch
import torch.nn as nn

from . import _lib, ops
from .nn_common import (HeadGraphs, LinearAttention, LoFTREncoderLayer, PositionEncodingSine, ResNetFPN_8_2,
                        _TransformerBase, _cache_lock, fine_expectation, tensor_key)

logger = logging.getLogger("onepose_st_b200")

__all__ = ["OnePosePlus_model", "LocalFeatureTransformer", "CoarseMatching", "FinePreprocess", "FineMatching",
           "PositionEncodingSine", "KeypointEncoding_linear", "LoFTREncoderLayer", "LinearAttention",
           "normalize_3d_keypoints", "build_backbone"]


class _NullProfiler:
    """Stands in for the reference's PassThroughProfiler (src/utils/profiler.py): record_function is a no-op."""

    class _Ctx:
        def __enter__(self):
            return self

        def __exit__(self, *a):
            return False

    def record_function(self, name):
        return self._Ctx()

    def profile(self, name):
        return self._Ctx()


def build_backbone(config):
    """src/models/OnePosePlus/backbone/__init__.py:7-14."""
    if config["type"] == "ResNetFPN":
        if list(config["resolution"]) == [8, 2]:
            return ResNetFPN_8_2(config["resnetfpn"])
        raise NotImplementedError("only the (8, 2) ResNetFPN resolution is supported")
    raise ValueError("loftr_backbone.type / resolution are not supported")


def normalize_3d_keypoints(kpts):
    """src/models/OnePosePlus/utils/normalize.py:17-28 (extent of batch element 0 scales every element)."""
    extent = kpts[0].max(dim=0).values - kpts[0].min(dim=0).values
    return (kpts - kpts.mean(dim=-2)[:, None, :]) / (extent.max() * 0.6)


class KeypointEncoding_linear(nn.Module):
    """src/models/OnePosePlus/utils/position_encoding.py:46-79.  Small MLP on L points; depends only on the
    3D model, so OnePosePlus_model caches its result across the frames of a sequence."""

    def __init__(self, inp_dim, feature_dim, layers, norm_method="batchnorm"):
        super().__init__()
        chans = [inp_dim] + list(layers) + [feature_dim]
        mods = []
        for i in range(1, len(chans)):
            mods.append(nn.Linear(chans[i - 1], chans[i], bias=True))
            if i < len(chans) - 1:
                if norm_method == "instancenorm":
                    mods.append(nn.InstanceNorm1d(chans[i]))
                elif norm_method == "layernorm":
                    mods.append(nn.LayerNorm(chans[i]))
                elif norm_method == "batchnorm":
                    mods.append(nn.BatchNorm1d(chans[i]))
                else:
                    raise NotImplementedError(norm_method)
                mods.append(nn.ReLU())
        self.encoder = nn.Sequential(*mods)
        self.norm_method = norm_method
        nn.init.constant_(self.encoder[-1].bias, 0.0)

    def mlp(self, kpts):
        x = kpts
        for m in self.encoder:
            if isinstance(m, nn.InstanceNorm1d):
                # the reference feeds [B, L, c] to InstanceNorm1d(c): statistics run over the channel dim of
                # each point (biased variance, eps 1e-5, no affine); spelled out to avoid torch's shape warning
                x = torch.nn.functional.instance_norm(x, eps=m.eps)
            else:
                x = m(x)
        return x

    def _kernel_ok(self, kpts, descriptors):
        lin = [m for m in self.encoder if isinstance(m, nn.Linear)]
        return (kpts.is_cuda and descriptors.is_cuda and not self.training and not torch.is_grad_enabled()
                and self.norm_method == "instancenorm" and [tuple(m.weight.shape) for m in lin] == [(32, 3), (64, 32), (128, 64), (256, 128)]
                and kpts.dim() == 3 and kpts.shape[-1] == 3 and descriptors.dim() == 3 and descriptors.shape[1] == 256
                and descriptors.shape[2] == kpts.shape[1] and descriptors.shape[0] == kpts.shape[0])

    def _packed(self):
        lin = [m for m in self.encoder if isinstance(m, nn.Linear)]
        ps = [t for m in lin for t in (m.weight, m.bias)]
        key = tuple((p.data_ptr(), p._version, str(p.device)) for p in ps)
        hit = self.__dict__.get("_packed_cache")
        if hit is None or hit[0] != key:
            with _cache_lock:
                flat = torch.cat([t for m in lin for t in (m.weight.detach().float().t().contiguous().flatten(),
                                                           m.bias.detach().float().flatten())])
                torch.cuda.current_stream(flat.device).synchronize()
                hit = (key, flat)
                self.__dict__["_packed_cache"] = hit
        return hit[1]

    def encode(self, kpts, descriptors, normalize: bool):
        """descriptors + MLP(kpts)^T on the fused kernel (csrc/kptenc.cu); ``normalize`` folds normalize_3d_keypoints in."""
        return ops.keypoint_encode(kpts, descriptors, self._packed(), normalize)

    def forward(self, kpts, descriptors):
        if self._kernel_ok(kpts, descriptors):
            return self.encode(kpts, descriptors, False)
        return descriptors + self.mlp(kpts).transpose(2, 1)

    def __getstate__(self):
        st = self.__dict__.copy()
        st.pop("_packed_cache", None)
        return st


class LocalFeatureTransformer(_TransformerBase):
    """src/models/OnePosePlus/loftr_module/transformer.py:97-171."""
    schedule = 0

    def __init__(self, config):
        super().__init__()
        self.config = config
        if config.get("type", "LoFTR") != "LoFTR":
            raise ValueError("unsupported encoder layer type")
        if config.get("final_proj"):
            raise NotImplementedError("final_proj=True is not supported")
        ri = config.get("redraw_interval")
        if ri is not None:
            assert ri % 2 == 0, "redraw_interval must be divisible by 2 since each attetnion layer is repeatedly called twice."
        self.norm_method = config["norm_method"]
        self._build(config["d_model"], config["nhead"], list(config["layer_names"]) * config["layer_iter_n"],
                    dropout=config.get("dropout", 0.0), attention=config["attention"], kernel_fn=config["kernel_fn"],
                    redraw_interval=ri, d_kernel=config.get("d_kernel"), rezero=config.get("rezero"),
                    norm_method=config["norm_method"])

    def forward(self, desc3d_db, desc2d_query, query_mask=None, return_middle_layer_features=False, _add2d=None,
                _planes_only=False):
        """desc3d_db [N, C, L] (channels first), desc2d_query [N, P, C] -> (desc3d [N,L,C], desc2d [N,P,C]).
        ``_planes_only`` (internal): return an ``ops.FeaturePlanes`` handle for CoarseMatching instead."""
        x3 = desc3d_db.permute(0, 2, 1)  # view; the stage-in kernel reads it strided
        if _planes_only and not return_middle_layer_features:
            return self._run(x3, desc2d_query, mask1=query_mask, add1=_add2d, planes_only=True)
        if not return_middle_layer_features:
            d3, d2 = self._run(x3, desc2d_query, mask1=query_mask, add1=_add2d)
            return d3, d2
        n = len(self.layers)
        m3, m2 = self._run(x3, desc2d_query, mask1=query_mask, add1=_add2d, upto=n - 1)
        tail = ops.transformer([self.layers[-1].packed()], [0 if self.layer_names[-1] == "self" else 1], 0, self.d_model,
                               self.nhead, m3, m2, mask1=query_mask, big_groups=(self.d_model == 256))
        return tail[0], tail[1], m3, m2


class CoarseMatching(nn.Module):
    """src/models/OnePosePlus/utils/coarse_matching.py:56-251 (eval mode)."""

    def __init__(self, config, profiler=None):
        super().__init__()
        self.config = config
        if config["feat_norm_method"] != "sqrt_feat_dim":
            raise NotImplementedError("only feat_norm_method='sqrt_feat_dim' is implemented")
        self.type = config["type"]
        if self.type != "dual-softmax":
            raise NotImplementedError()
        self.temperature = config["dual_softmax"]["temperature"]
        self.thr = config["thr"]
        self.border_rm = config["border_rm"]
        self.train_coarse_percent = config["train"]["train_coarse_percent"]
        self.train_pad_num_gt_min = config["train"]["train_pad_num_gt_min"]
        self.profiler = profiler or _NullProfiler()
        # conf_matrix [N, L, S] is read only by the training losses; inference callers never touch it, so it is
        # not written to HBM unless asked for (SURVEY.md 8b "data dict -- outputs")
        self.materialize_conf_matrix = bool(config.get("materialize_conf_matrix", False))
        self.precision = ops.PREC_BF16X3

    def forward(self, feat_db_3d, feat_query, data, mask_query=None):
        self.finish(self.launch(feat_db_3d, feat_query, data, mask_query), data)

    def launch(self, feat_db_3d, feat_query, data, mask_query=None, _status=None) -> "ops.CoarseLaunch":
        """Enqueue the matching kernels (no host sync; needs data['q_hw_c'])."""
        if self.training:
            raise NotImplementedError("training-time coarse matching (GT padding) is out of scope of the inference path")
        if "mask0" in data:
            raise NotImplementedError
        planes = feat_db_3d if isinstance(feat_db_3d, ops.FeaturePlanes) else None  # internal fused hand-over
        C = planes.c if planes is not None else feat_db_3d.shape[2]
        hc, wc = int(data["q_hw_c"][0]), int(data["q_hw_c"][1])
        scale = 1.0 / (C * (self.temperature + 1e-4))  # (a/sqrt C).(b/sqrt C) / (T + 1e-4)
        with self.profiler.record_function("LoFTR/coarse-matching/get_coarse_match"):
            return ops.coarse_match_launch(
                feat_db_3d, feat_query, scale=scale, thr=self.thr, border_rm=self.border_rm, border_mode=0, grid_a=None,
                grid_b=(hc, wc), mask_b=mask_query, mask_fill=0, materialize_conf=self.materialize_conf_matrix,
                precision=self.precision, planes=planes, status=_status)

    def finish(self, handle: "ops.CoarseLaunch", data, copy: bool = False):
        """Read the match count (the forward's host sync) and fill the reference's coarse-level ``data`` keys."""
        wc = int(data["q_hw_c"][1])
        b_ids, i_ids, j_ids, mconf, conf = ops.coarse_match_finish(handle, copy=copy)
        if conf is not None:
            data.update({"conf_matrix": conf})
        scale_c = data["q_hw_i"][0] / data["q_hw_c"][0]
        scale_total = scale_c * data["query_image_scale"][b_ids][:, [1, 0]] if "query_image_scale" in data else scale_c
        mkpts_query = torch.stack([j_ids % wc, j_ids // wc], dim=1) * scale_total
        mkpts_3d_db = data["keypoints3d"][b_ids, i_ids]
        # mconf > thr > 0 always: the reference's `mconf != 0` filter (GT padding only) keeps everything
        data.update({"b_ids": b_ids, "i_ids": i_ids, "j_ids": j_ids, "gt_mask": mconf == 0, "m_bids": b_ids,
                     "mkpts_3d_db": mkpts_3d_db, "mkpts_query_c": mkpts_query, "mconf": mconf})


class FinePreprocess(nn.Module):
    """src/models/OnePosePlus/loftr_module/fine_preprocess.py:8-55: windows are gathered directly for the M
    matches instead of unfolding all H_c*W_c of them."""

    def __init__(self, config, cf_res=None, feat_ids=None, feat_dims=None):
        super().__init__()
        self.config = config
        self.W = config["window_size"]
        self.coarse_id, self.fine_id = [int(log(r, 2)) for r in cf_res]
        self.feat_ids, self.feat_dims = feat_ids, feat_dims

    def forward(self, data, feat_3D, feat_query_f):
        data.update({"W": self.W})
        dev = feat_3D.device
        if data["b_ids"].shape[0] == 0:
            w = getattr(self, "W_MAX", self.W)
            return (torch.empty(0, self.config["d_model"], 1, device=dev),
                    torch.empty(0, w ** 2, self.config["d_model"], device=dev))
        stride = data["q_hw_f"][0] // data["q_hw_c"][0]
        hc, wc = int(data["q_hw_c"][0]), int(data["q_hw_c"][1])
        flag = data.setdefault("_opb200_err_flag", torch.zeros(1, dtype=torch.int32, device=dev))
        f3d = ops.gather_rows(feat_3D, data["b_ids"], data["i_ids"], flag).unsqueeze(-1)
        win = ops.fine_gather(feat_query_f, data["b_ids"], data["j_ids"], self.W, stride, hc, wc, flag)
        return f3d, win


class FineMatching(nn.Module):
    """src/models/OnePosePlus/utils/fine_matching.py:10-110."""

    def __init__(self, config):
        super().__init__()
        self.config = config
        self._type = config["s2d"]["type"]
        if self._type != "heatmap":
            raise NotImplementedError()

    def forward(self, feat_db_3d, feat_query_unfolded, data):
        M, WW, C = feat_query_unfolded.shape
        W = int(math.sqrt(WW))
        scale = data["q_hw_i"][0] / data["q_hw_f"][0]
        self.M, self.W, self.WW, self.C, self.scale = M, W, WW, C, scale
        if M == 0:
            assert self.training is False, "M is always >0, when training, see coarse_matching.py"
            logger.warning("No matches found in coarse-level.")
            data.update({"expec_f": torch.empty(0, 3, device=feat_db_3d.device), "mkpts_3d_db": data["mkpts_3d_db"],
                         "mkpts_query_f": data["mkpts_query_c"]})
            return
        L = feat_db_3d.shape[1]
        assert L % 2 == 1
        expec = fine_expectation(feat_db_3d[:, L // 2, :], feat_query_unfolded)
        data.update({"expec_f": expec})
        coords = expec[:, :2]
        qs = scale * data["query_image_scale"][data["b_ids"]][:, [1, 0]] if "query_image_scale" in data else scale
        mk = data["mkpts_query_c"] + (coords * (W // 2) * qs)[: len(data["mkpts_query_c"])]
        data.update({"mkpts_3d_db": data["mkpts_3d_db"], "mkpts_query_f": mk})


def _get_feat_dims(config):
    dims = [1, *config["resnetfpn"]["block_dims"]]
    return [dims[i] for i in config["resnetfpn"]["output_layers"]]


class OnePosePlus_model(nn.Module):
    """src/models/OnePosePlus/OnePosePlusModel.py:24-203.  ``forward(data)`` mutates ``data`` in place."""

    def __init__(self, config, profiler=None, debug=False):
        super().__init__()
        self.config = config
        self.profiler = profiler or _NullProfiler()
        self.debug = debug
        bb = config["loftr_backbone"]
        if list(bb["resnetfpn"].get("output_layers", [3, 1])) != [3, 1]:
            raise NotImplementedError("backbone output_layers must be [3, 1]")
        self.backbone = build_backbone(bb)
        if config["positional_encoding"]["enable"]:
            self.dense_pos_encoding = PositionEncodingSine(config["loftr_coarse"]["d_model"],
                                                           max_shape=config["positional_encoding"]["pos_emb_shape"])
        else:
            self.dense_pos_encoding = None
        if config["keypoints_encoding"]["enable"]:
            if config["keypoints_encoding"]["type"] != "mlp_linear":
                raise NotImplementedError
            ke = config["keypoints_encoding"]
            self.kpt_3d_pos_encoding = KeypointEncoding_linear(inp_dim=3, feature_dim=ke["descriptor_dim"],
                                                               layers=ke["keypoints_encoder"], norm_method=ke["norm_method"])
        else:
            self.kpt_3d_pos_encoding = None
        self.loftr_coarse = LocalFeatureTransformer(config["loftr_coarse"])
        self.coarse_matching = CoarseMatching(config["coarse_matching"], profiler=self.profiler)
        self.fine_preprocess = FinePreprocess(config["loftr_fine"], cf_res=bb["resolution"],
                                              feat_ids=bb["resnetfpn"].get("output_layers", [3, 1]), feat_dims=_get_feat_dims(bb))
        self.loftr_fine = LocalFeatureTransformer(config["loftr_fine"])
        self.fine_matching = FineMatching(config["fine_matching"])
        self.loftr_backbone_pretrained = bb["pretrained"]
        if self.loftr_backbone_pretrained is not None:
            ckpt = torch.load(self.loftr_backbone_pretrained, "cpu")["state_dict"]
            for k in list(ckpt.keys()):
                if "backbone" in k:
                    ckpt[k[k.find("backbone") + len("backbone") + 1:]] = ckpt[k]
                ckpt.pop(k)
            self.backbone.load_state_dict(ckpt)
            if bb["pretrained_fix"]:
                for p in self.backbone.parameters():
                    p.requires_grad = False
        self._kpt_cache = None  # (key, encoded descriptors): the 3D model is fixed across a sequence
        self._graphs = HeadGraphs()  # CUDA graphs of the head, see _graphed_head

    def _encode_3d(self, data):
        desc = data["descriptors3d_coarse_db"] if "descriptors3d_coarse_db" in data else data["descriptors3d_db"]
        if self.kpt_3d_pos_encoding is None:
            return desc
        kp = data["keypoints3d"]

        def encode():
            enc = self.kpt_3d_pos_encoding
            fused = enc._kernel_ok(kp, desc)  # normalisation + MLP + add in one kernel (else: eager PyTorch on the GPU)
            if kp.shape[0] > 1 and kp.stride(0) == 0 and desc.stride(0) == 0:
                # one 3D model broadcast over the batch: encode it once, hand out a broadcast view (stride 0), which is
                # also what lets the coarse transformer reuse its first self layer on this stream
                one = enc.encode(kp[:1], desc[:1], True) if fused else enc(normalize_3d_keypoints(kp[:1]), desc[:1])
                return one.expand(kp.shape[0], -1, -1)
            return enc.encode(kp, desc, True) if fused else enc(normalize_3d_keypoints(kp), desc)

        if tensor_key(kp) is None or tensor_key(desc) is None:  # inference tensors track no version: never cached
            return encode()
        key = (tensor_key(kp), tensor_key(desc),
               tuple((p.data_ptr(), p._version) for p in self.kpt_3d_pos_encoding.parameters()))
        hit = self._kpt_cache
        if hit is None or hit[0] != key:
            with _cache_lock:
                hit = self._kpt_cache
                if hit is None or hit[0] != key:
                    enc = encode()
                    if enc.is_cuda:
                        torch.cuda.current_stream(enc.device).synchronize()  # other streams read it too
                    # the entry holds kp and desc: while it is current their addresses cannot be handed to other tensors
                    # (a freed-and-reallocated 3D model of the same shape must not hit), and a serial for the CUDA graphs
                    hit = (key, enc, kp, desc, (hit[4] + 1) if hit is not None else 1)
                    self._kpt_cache = hit
        return hit[1]

    def __getstate__(self):
        st = self.__dict__.copy()
        st["_kpt_cache"] = None
        st["_graphs"] = HeadGraphs()
        return st

    def _head(self, data, img, qmask, conv_precision=None):
        """Backbone -> coarse LoFTR stack -> coarse matching kernels: everything before the forward's host sync.
        Returns (coarse launch handle, fine feature map, (hc, wc)); sets data['q_hw_c']."""
        # status words of the forward's one device -> host copy: [0..2] coarse matching, [3] fp16 range flag of the backbone
        status = torch.zeros(4, dtype=torch.int32, device=img.device) if img.is_cuda else None
        feat_c, feat_f = self.backbone(img, _precision=conv_precision, _range_flag=None if status is None else status[3:])
        data.update({"q_hw_c": feat_c.shape[2:]})
        n, c, hc, wc = feat_c.shape
        # 'n c h w -> n (h w) c' (+ PE) happens inside the transformer's stage-in kernel
        q2d = feat_c.flatten(2).permute(0, 2, 1)
        add2d = self.dense_pos_encoding.token_table(hc, wc) if self.dense_pos_encoding is not None else None
        desc3d = self._encode_3d(data)
        # the coarse features are consumed by CoarseMatching only: hand them over as bf16 planes
        planes = self.loftr_coarse(desc3d, q2d, query_mask=qmask, _add2d=add2d, _planes_only=True)
        return self.coarse_matching.launch(planes, None, data, mask_query=qmask, _status=status), feat_f, feat_c.shape[2:]

    # ---- CUDA graph of the head (nn_common.HeadGraphs) ----------------------------------------------------------
    # Opt-in: a replay reuses static buffers and assumes the weights are not modified in place between forwards.
    cuda_graph = False

    def _graphed_head(self, data, img):
        kp = data["keypoints3d"]
        desc = data["descriptors3d_coarse_db"] if "descriptors3d_coarse_db" in data else data["descriptors3d_db"]
        # the per-sequence cache is consulted on every forward (a replay never reaches _head): a new 3D model rebuilds it
        # here, which changes the serial in cache_ids() below and drops the graphs captured for the previous model
        self._encode_3d(data)
        key = (tuple(kp.shape), tuple(kp.stride()), tuple(desc.shape), tuple(desc.stride()), self.coarse_matching.precision,
               getattr(self.loftr_coarse, "precision", ops.PREC_BF16X3), self.backbone.conv_precision,
               self.coarse_matching.materialize_conf_matrix)

        def cache_ids():  # tensors the captured kernels read by address besides the weights: (address, rebuild serial)
            c0, k0 = self.loftr_coarse.__dict__.get("_stream0_cache"), self._kpt_cache
            return (None if c0 is None else (c0[1].data_ptr(), c0[4]), None if k0 is None else (k0[1].data_ptr(), k0[4]))

        def head(static_img):
            handle, feat_f, hw_c = self._head(dict(data), static_img, None)
            return handle, feat_f, tuple(hw_c)

        handle, feat_f, hw_c = self._graphs.run(key, (img,), head, guard=cache_ids)
        return handle, feat_f, torch.Size(hw_c)

    @torch.no_grad()
    def forward(self, data):
        if self.training:
            raise NotImplementedError("onepose_st_b200 implements the inference forward only; call .eval()")
        img = data["query_image"]
        data.update({"bs": img.size(0), "q_hw_i": img.shape[2:]})
        qmask = data["query_image_mask"].flatten(-2) if "query_image_mask" in data else None
        cacheable = tensor_key(data["keypoints3d"]) is not None and tensor_key(data["descriptors3d_db"]) is not None \
            and ("descriptors3d_coarse_db" not in data or tensor_key(data["descriptors3d_coarse_db"]) is not None)
        if self.cuda_graph and qmask is None and img.is_cuda and cacheable:
            handle, feat_f, hw_c = self._graphed_head(data, img)
            data.update({"q_hw_c": hw_c, "q_hw_f": feat_f.shape[2:]})
            self.coarse_matching.finish(handle, data, copy=True)
        else:
            handle, feat_f, hw_c = self._head(data, img, qmask)
            data.update({"q_hw_f": feat_f.shape[2:]})
            self.coarse_matching.finish(handle, data)
        if handle.extra and handle.extra[0]:
            # the fp16 backbone stored a value beyond the fp16 range (unnormalised images, very large folded BatchNorm
            # gains): its maps cannot be trusted for this input -- the head is repeated with exact-fp32 convolutions
            if not self.__dict__.get("_warned_fp16_range"):
                logger.warning("backbone activations left the fp16 range; repeating this forward with conv_precision='fp32' "
                               "(set model.backbone.conv_precision = 'fp32' to avoid the repeated work)")
                self.__dict__["_warned_fp16_range"] = True
            handle, feat_f, hw_c = self._head(data, img, qmask, conv_precision="fp32")
            data.update({"q_hw_f": feat_f.shape[2:]})
            self.coarse_matching.finish(handle, data)
        if not self.config["fine_matching"]["enable"]:
            data.update({"mkpts_3d_db": data["mkpts_3d_db"], "mkpts_query_f": data["mkpts_query_c"]})
            return
        f3d, win = self.fine_preprocess(data, data["descriptors3d_db"], feat_f)
        if win.size(0) != 0 and self.config["loftr_fine"]["enable"]:
            f3d, win = self.loftr_fine(f3d, win)
        else:
            f3d = f3d.permute(0, 2, 1)
        self.fine_matching(f3d, win, data)
        flag = data.pop("_opb200_err_flag", None)
        if flag is not None and self.debug and int(flag.item()):
            raise IndexError("coarse match ids out of range in fine_preprocess")
