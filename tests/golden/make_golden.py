"""Generate the golden vectors by running the UNMODIFIED reference modules.

Run in the build container only (needs /root/reference):

    python tests/golden/make_golden.py

Imports the reference through ``ref_shims`` (no reference source is copied), feeds it the
seeded inputs of ``cases.py`` with weights loaded through ``load_state_dict(strict=True)``
(which also proves the parameter names line up), and stores the OUTPUTS under
``tests/golden/<case>.pt``.  Inputs are not stored: they are rebuilt from ``cases.py``.
"""
import copy
import os
import sys
import warnings

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, REPO)
sys.path.insert(0, HERE)
warnings.filterwarnings("ignore")

import ref_shims  # noqa: E402
import cases  # noqa: E402

OPP_model, LoFTR_model, loftr_default_cfg = ref_shims.install()
cwd = os.getcwd()

from src.models.OnePosePlus.loftr_module.linear_attention import LinearAttention as OPP_LinAttn  # noqa: E402
from src.models.OnePosePlus.loftr_module.transformer import (  # noqa: E402
    LoFTREncoderLayer as OPP_Layer, LocalFeatureTransformer as OPP_Transformer)
from src.models.OnePosePlus.loftr_module.fine_preprocess import FinePreprocess as OPP_FinePre  # noqa: E402
from src.models.OnePosePlus.utils.coarse_matching import CoarseMatching as OPP_Coarse  # noqa: E402
from src.models.OnePosePlus.utils.fine_matching import FineMatching as OPP_Fine  # noqa: E402
from src.utils.profiler import PassThroughProfiler  # noqa: E402
from src.models.OnePosePlus.utils.position_encoding import (  # noqa: E402
    PositionEncodingSine as OPP_PE, KeypointEncoding_linear as OPP_KptEnc)
from src.models.OnePosePlus.utils.normalize import normalize_3d_keypoints as opp_norm3d  # noqa: E402
from src.models.OnePosePlus.backbone import build_backbone as opp_build_backbone  # noqa: E402
from loftr.loftr_module.transformer import LocalFeatureTransformer as L_Transformer  # noqa: E402
from loftr.loftr_module.fine_preprocess import FinePreprocess as L_FinePre  # noqa: E402
from loftr.utils.coarse_matching import CoarseMatching as L_Coarse  # noqa: E402
from loftr.utils.fine_matching import FineMatching as L_Fine  # noqa: E402
from loftr.utils.position_encoding import PositionEncodingSine as L_PE  # noqa: E402
from oracle import ref_ops as R  # noqa: E402  (config dicts only)

OPP_CFG = R.ONEPOSE_PLUS_CFG
OUT = {}


def tensors_only(d):
    return {k: v.clone() for k, v in d.items() if isinstance(v, torch.Tensor)}


def strip(sd, prefix):
    return {k[len(prefix):]: v for k, v in sd.items() if k.startswith(prefix)}


@torch.no_grad()
def main():
    # ---- linear attention
    for masked in (False, True):
        c = cases.case_linattn(masked)
        att = OPP_LinAttn(32, kernel_fn="elu + 1")
        OUT[f"linattn_{int(masked)}"] = {"out": att(c["q"], c["k"], c["v"], c["q_mask"], c["kv_mask"])}
    # ---- encoder layer
    for d_model in (256, 128):
        for masked in (False, True):
            c = cases.case_encoder_layer(d_model, masked)
            lay = OPP_Layer(d_model, 8, 0.0, "linear", "elu + 1", 2, 16, rezero=None, norm_method="layernorm")
            lay.load_state_dict(strip(c["sd"], "t.layers.0."), strict=True)
            lay.eval()
            OUT[f"enc_layer_{d_model}_{int(masked)}"] = {"out": lay(c["x"], c["source"], c["x_mask"], c["source_mask"])}
    # ---- transformers
    for masked in (False, True):
        c = cases.case_tr_onepose_coarse(masked)
        tr = OPP_Transformer(OPP_CFG["loftr_coarse"])
        tr.load_state_dict(strip(c["sd"], "t."), strict=True)
        tr.eval()
        d3, d2 = tr(c["desc3d"], c["desc2d"], query_mask=c["query_mask"])
        OUT[f"tr_onepose_coarse_{int(masked)}"] = {"desc3d": d3, "desc2d": d2}
        c = cases.case_tr_loftr_coarse(masked)
        tr = L_Transformer(loftr_default_cfg["coarse"])
        tr.load_state_dict(strip(c["sd"], "t."), strict=True)
        tr.eval()
        f0, f1 = tr(c["feat0"], c["feat1"], c["mask0"], c["mask1"])
        OUT[f"tr_loftr_coarse_{int(masked)}"] = {"feat0": f0, "feat1": f1}
    c = cases.case_tr_onepose_fine()
    tr = OPP_Transformer(OPP_CFG["loftr_fine"])
    tr.load_state_dict(strip(c["sd"], "t."), strict=True)
    d3, d2 = tr.eval()(c["desc3d"], c["desc2d"])
    OUT["tr_onepose_fine"] = {"desc3d": d3, "desc2d": d2}
    c = cases.case_tr_loftr_fine()
    tr = L_Transformer(loftr_default_cfg["fine"])
    tr.load_state_dict(strip(c["sd"], "t."), strict=True)
    f0, f1 = tr.eval()(c["feat0"], c["feat1"])
    OUT["tr_loftr_fine"] = {"feat0": f0, "feat1": f1}
    # ---- coarse matching
    for variant in ("plain", "mask", "scale", "batch2"):
        c = cases.case_cm_onepose(variant)
        cm = OPP_Coarse(OPP_CFG["coarse_matching"], profiler=PassThroughProfiler()).eval()
        data = dict(c["data"])
        cm(c["feat3d"], c["feat2d"], data, mask_query=c["mask_query"])
        OUT[f"cm_onepose_{variant}"] = {k: data[k] for k in ("conf_matrix", "b_ids", "i_ids", "j_ids", "gt_mask", "m_bids",
                                                              "mkpts_3d_db", "mkpts_query_c", "mconf")}
    for variant in ("plain", "mask", "scale"):
        c = cases.case_cm_loftr(variant)
        cm = L_Coarse(loftr_default_cfg["match_coarse"]).eval()
        data = dict(c["data"])
        cm(c["feat0"], c["feat1"], data, mask_c0=c["mask0"], mask_c1=c["mask1"])
        OUT[f"cm_loftr_{variant}"] = {k: data[k] for k in ("conf_matrix", "b_ids", "i_ids", "j_ids", "gt_mask", "m_bids",
                                                            "mkpts0_c", "mkpts1_c", "mconf")}
    # ---- fine preprocess
    c = cases.case_fine_preprocess("onepose")
    fp = OPP_FinePre(OPP_CFG["loftr_fine"], cf_res=[8, 2], feat_ids=[3, 1], feat_dims=[256, 128])
    data = dict(c["data"])
    a, b = fp(data, c["desc3d_fine"], c["feat_f"])
    OUT["fine_preprocess_onepose"] = {"feat3d": a, "win": b}
    c = cases.case_fine_preprocess("loftr")
    fp = L_FinePre(loftr_default_cfg)
    data = dict(c["data"])
    a, b = fp(c["feat_f0"], c["feat_f1"], None, None, data)
    OUT["fine_preprocess_loftr"] = {"win0": a, "win1": b}
    # ---- fine matching
    c = cases.case_fine_matching("onepose")
    data = dict(c["data"])
    OPP_Fine(OPP_CFG["fine_matching"]).eval()(c["feat3d"], c["win"], data)
    OUT["fine_matching_onepose"] = {k: data[k] for k in ("expec_f", "mkpts_query_f", "mkpts_3d_db")}
    c = cases.case_fine_matching("loftr")
    data = dict(c["data"])
    L_Fine().eval()(c["feat_f0"], c["feat_f1"], data)
    OUT["fine_matching_loftr"] = {k: data[k] for k in ("expec_f", "mkpts0_f", "mkpts1_f")}
    # ---- encodings / backbone
    OUT["pe_onepose"] = {"pe": OPP_PE(256, max_shape=(256, 256)).pe[:, :, :16, :24].clone()}
    OUT["pe_loftr"] = {"pe": L_PE(256, temp_bug_fix=False).pe[:, :, :12, :20].clone()}
    c = cases.case_kpt_encoding()
    ke = OPP_KptEnc(3, 256, [32, 64, 128], "instancenorm")
    ke.load_state_dict(strip(c["sd"], "kpt_3d_pos_encoding."), strict=True)
    kn = opp_norm3d(c["kpts"])
    OUT["kpt_encoding"] = {"kpts_normalized": kn, "out": ke.eval()(kn, c["desc"])}
    c = cases.case_backbone()
    bb = opp_build_backbone(OPP_CFG["loftr_backbone"])
    bb.load_state_dict(strip(c["sd"], "backbone."), strict=True)
    fc, ff = bb.eval()(c["image"])
    OUT["backbone"] = {"feat_c": fc, "feat_f_sub": ff[:, ::8, ::4, ::4].clone()}
    # ---- full models
    for nm, kw in (("onepose_full", {}), ("onepose_full_mask", {"with_mask": True})):
        c = cases.case_onepose_full(**kw)
        m = OPP_model(copy.deepcopy(c["cfg"]))
        m.load_state_dict(c["sd"], strict=True)
        data = dict(c["data"])
        m.eval()(data)
        OUT[nm] = {k: data[k] for k in ("conf_matrix", "b_ids", "i_ids", "j_ids", "m_bids", "mkpts_3d_db", "mkpts_query_c",
                                        "mconf", "expec_f", "mkpts_query_f")}
        print(nm, "M =", len(data["mconf"]))
    for mode in ("full", "coarse", "fine_only", "extract"):
        c = cases.case_loftr_full(mode=mode)
        m = LoFTR_model(loftr_default_cfg, enable_fine_matching=(mode != "coarse"))
        m.load_state_dict(c["sd"], strict=True)
        data = {k: (v.clone() if isinstance(v, torch.Tensor) else v) for k, v in c["data"].items()}
        kw = dict(extract_coarse_feature=True, extract_fine_feature=True) if mode == "extract" else {}
        m.eval()(data, **kw)
        keys = [k for k in ("conf_matrix", "b_ids", "i_ids", "j_ids", "m_bids", "mkpts0_c", "mkpts1_c", "mconf", "expec_f",
                            "mkpts0_f", "mkpts1_f", "feat_coarse_b_0", "feat_coarse_b_1", "feat_ext0", "feat_ext1") if k in data]
        OUT[f"loftr_{mode}"] = {k: data[k] for k in keys}
        print("loftr", mode, "M =", len(data["mconf"]))

    os.chdir(cwd)
    total = 0
    for name, d in OUT.items():
        d = {k: (v.detach().clone().contiguous() if isinstance(v, torch.Tensor) else v) for k, v in d.items()}
        path = os.path.join(HERE, name + ".pt")
        torch.save(d, path)
        total += os.path.getsize(path)
    print(f"wrote {len(OUT)} golden files, {total / 1e6:.2f} MB")


if __name__ == "__main__":
    main()
