"""GPU parity: fine-level window gather and correlation/softmax-expectation kernels vs golden/oracle."""
import pytest
import torch

import cases
from oracle import ref_ops as R

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("channels_last", [False, True])
def test_fine_gather_vs_golden(golden, channels_last):
    from onepose_st_b200 import ops
    dev = "cuda"
    c = cases.case_fine_preprocess("onepose")
    d = c["data"]
    f = c["feat_f"].to(dev)
    if channels_last:
        f = f.contiguous(memory_format=torch.channels_last)
    win = ops.fine_gather(f, d["b_ids"].to(dev), d["j_ids"].to(dev), 5, 4, 8, 8)
    f3d = ops.gather_rows(c["desc3d_fine"].to(dev), d["b_ids"].to(dev), d["i_ids"].to(dev))
    g = golden("fine_preprocess_onepose")
    assert torch.equal(win.cpu(), g["win"])          # pure data movement: bit exact
    assert torch.equal(f3d.cpu()[:, :, None], g["feat3d"])
    c = cases.case_fine_preprocess("loftr")
    d = c["data"]
    w0 = ops.fine_gather(c["feat_f0"].to(dev), d["b_ids"].to(dev), d["i_ids"].to(dev), 9, 4, 8, 8)
    w1 = ops.fine_gather(c["feat_f1"].to(dev), d["b_ids"].to(dev), d["j_ids"].to(dev), 9, 4, 8, 8)
    g = golden("fine_preprocess_loftr")
    assert torch.equal(w0.cpu(), g["win0"]) and torch.equal(w1.cpu(), g["win1"])


def test_fine_gather_out_of_range_sets_flag():
    from onepose_st_b200 import ops
    dev = "cuda"
    f = torch.randn(1, 128, 32, 32, device=dev)
    flag = torch.zeros(1, dtype=torch.int32, device=dev)
    ops.fine_gather(f, torch.zeros(2, dtype=torch.long, device=dev), torch.tensor([3, 64], device=dev), 5, 4, 8, 8, flag)
    assert flag.item() == 1


def test_fine_match_vs_golden(golden):
    from onepose_st_b200 import ops
    dev = "cuda"
    c = cases.case_fine_matching("onepose")
    e = ops.fine_match(c["feat3d"].to(dev)[:, 0, :], c["win"].to(dev))
    g = golden("fine_matching_onepose")
    err = (e.cpu() - g["expec_f"]).abs().max().item()
    print("fine match onepose expec_f err", err)
    assert err < 2e-5
    c = cases.case_fine_matching("loftr")
    f0 = c["feat_f0"].to(dev)
    e = ops.fine_match(f0[:, 81 // 2, :], c["feat_f1"].to(dev))
    g = golden("fine_matching_loftr")
    err = (e.cpu() - g["expec_f"]).abs().max().item()
    print("fine match loftr expec_f err", err)
    assert err < 2e-5


@pytest.mark.parametrize("W,M", [(5, 2000), (9, 20000)])
def test_fine_match_sweep_vs_oracle(W, M):
    """BASELINE config 5 (fine sweep) shapes, SURVEY 8d planted windows."""
    from onepose_st_b200 import ops
    dev = "cuda"
    WW = W * W
    g = torch.Generator().manual_seed(31)
    win = torch.randn(M, WW, 128, generator=g)
    r = torch.randint(0, WW, (M,), generator=torch.Generator().manual_seed(32))
    q = win[torch.arange(M), r] * 0.3 + 0.05 * torch.randn(M, 128, generator=g)
    ref = R.fine_heatmap_expectation(q, win)
    e = ops.fine_match(q.to(dev), win.to(dev))
    err = (e.cpu() - ref).abs().max().item()
    print(f"fine sweep W={W} M={M}: expec_f max abs err {err:.2e}")
    assert err < 5e-5
    # size-independent property: the expectation is pulled towards the planted cell on both axes
    lin = torch.linspace(-1, 1, W)
    ec = e.cpu()
    assert torch.corrcoef(torch.stack([ec[:, 0], lin[r % W]]))[0, 1].item() > 0.9
    assert torch.corrcoef(torch.stack([ec[:, 1], lin[r // W]]))[0, 1].item() > 0.9
    assert (ec[:, :2].abs() <= 1.0).all() and (ec[:, 2] >= 0).all()


def test_backbone_glue_ops_match_torch():
    """FPN glue kernels (channels-last fp32) against the PyTorch ops they fuse."""
    import torch.nn.functional as F
    from onepose_st_b200 import ops
    g = torch.Generator().manual_seed(5)
    src = torch.randn(2, 12, 5, 7, generator=g).cuda().contiguous(memory_format=torch.channels_last)
    dst = torch.randn(2, 12, 10, 14, generator=g).cuda().contiguous(memory_format=torch.channels_last)
    want = dst + F.interpolate(src, scale_factor=2.0, mode="bilinear", align_corners=True)
    got = ops.upsample2x_add_(dst.clone(memory_format=torch.channels_last), src)
    assert (got - want).abs().max().item() < 1e-5
    x = torch.randn(3, 8, 6, 4, generator=g).cuda().contiguous(memory_format=torch.channels_last)
    b = torch.randn(8, generator=g).cuda()
    want = F.leaky_relu(x + b[None, :, None, None], 0.01)
    got = ops.bias_leaky_relu_(x.clone(memory_format=torch.channels_last), b, 0.01)
    assert (got - want).abs().max().item() < 1e-6
    with pytest.raises(Exception):
        ops.upsample2x_add_(dst.contiguous(), src)  # NCHW-contiguous input is refused, not silently converted


def test_backbone_fused_vs_plain_module(golden):
    """Fused inference backbone (folded BN, cuDNN fused blocks, glue kernels) vs the reference's golden maps."""
    import cases
    from onepose_st_b200.nn_common import ResNetFPN_8_2
    c = cases.case_backbone()
    m = ResNetFPN_8_2({"block_type": "BasicBlock", "initial_dim": 128, "block_dims": [128, 196, 256]})
    m.load_state_dict({k[len("backbone."):]: v for k, v in c["sd"].items()}, strict=True)
    m = m.cuda().eval()
    from conftest import backbone_precision
    with backbone_precision("fp32"), torch.no_grad():
        fc, ff = m(c["image"].cuda())
    g = golden("backbone")
    gc, gf = g["feat_c"], g["feat_f_sub"]
    ec = (fc.cpu() - gc).abs().max().item() / gc.abs().max().item()
    ef = (ff.cpu()[:, ::8, ::4, ::4] - gf).abs().max().item() / gf.abs().max().item()
    print(f"backbone rel err coarse {ec:.2e} fine {ef:.2e}")
    assert ec < 1e-4 and ef < 1e-4


def test_backbone_f16_tensor_core_path_is_tf32_grade(golden):
    """The tcgen05 fp16 backbone (conv_precision="f16", the default) against the
    reference's golden fp32 maps: its error must be of the size of the error the reference's own cuDNN TF32
    convolutions make on the same input (both have 11-bit operand significands)."""
    import cases
    from onepose_st_b200.nn_common import ResNetFPN_8_2, FusedBackbone, FusedBackboneF16
    c = cases.case_backbone()
    m = ResNetFPN_8_2({"block_type": "BasicBlock", "initial_dim": 128, "block_dims": [128, 196, 256]})
    m.load_state_dict({k[len("backbone."):]: v for k, v in c["sd"].items()}, strict=True)
    m = m.cuda().eval()
    assert FusedBackboneF16.supports(m)
    g = golden("backbone")
    gc, gf = g["feat_c"], g["feat_f_sub"]

    def errs(fc, ff):
        return ((fc.cpu() - gc).abs().max().item() / gc.abs().max().item(),
                (ff.cpu()[:, ::8, ::4, ::4] - gf).abs().max().item() / gf.abs().max().item())
    old = torch.backends.cudnn.allow_tf32
    try:
        torch.backends.cudnn.allow_tf32 = True
        with torch.no_grad():
            assert m.conv_precision == "f16"                        # the default, whatever the global TF32 switch says
            ex = m.fused()
            assert isinstance(ex, FusedBackboneF16)
            flag = torch.zeros(1, dtype=torch.int32, device="cuda")
            e16 = errs(*m(c["image"].cuda(), _range_flag=flag))
            assert int(flag.item()) == 0                            # nothing left the fp16 range
            etf = errs(*FusedBackbone(m)(c["image"].cuda()))       # cuDNN executor with TF32 convolutions
        torch.backends.cudnn.allow_tf32 = False
        with torch.no_grad():
            assert isinstance(m.fused(), FusedBackboneF16)          # the global switch does not pick the executor
            assert isinstance(m.fused("fp32"), FusedBackbone)
            m.conv_precision = "fp32"
            assert isinstance(m.fused(), FusedBackbone)
    finally:
        torch.backends.cudnn.allow_tf32 = old
    print(f"backbone rel err vs fp32 golden: fp16 tcgen05 coarse {e16[0]:.2e} fine {e16[1]:.2e}; cuDNN TF32 coarse {etf[0]:.2e} fine {etf[1]:.2e}")
    assert e16[0] < 5e-3 and e16[1] < 5e-3
    assert e16[0] < 3 * etf[0] + 1e-4 and e16[1] < 3 * etf[1] + 1e-4


@pytest.mark.parametrize("hw,n", [((480, 640), 2), ((136, 104), 3), ((64, 64), 1)])
def test_backbone_f16_vs_fp32_executor_other_sizes(hw, n):
    """tcgen05 backbone against the exact-fp32 cuDNN executor on non-square / small / ragged-tile image sizes."""
    from onepose_st_b200.nn_common import ResNetFPN_8_2, FusedBackbone, FusedBackboneF16
    torch.manual_seed(3)
    m = ResNetFPN_8_2({"block_type": "BasicBlock", "initial_dim": 128, "block_dims": [128, 196, 256]})
    for mod in m.modules():  # non-trivial BatchNorm statistics so that the folding is exercised
        if isinstance(mod, torch.nn.BatchNorm2d):
            mod.running_mean.normal_(0, 0.1)
            mod.running_var.uniform_(0.5, 1.5)
            mod.weight.data.uniform_(0.5, 1.5)
            mod.bias.data.normal_(0, 0.1)
    m = m.cuda().eval()
    img = torch.rand(n, 1, hw[0], hw[1], device="cuda")
    old = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = False
    try:
        with torch.no_grad():
            rc, rf = FusedBackbone(m)(img)
            gc, gf = FusedBackboneF16(m)(img)
    finally:
        torch.backends.cudnn.allow_tf32 = old
    assert gc.shape == rc.shape and gf.shape == rf.shape
    ec = (gc - rc).abs().max().item() / rc.abs().max().item()
    ef = (gf - rf).abs().max().item() / rf.abs().max().item()
    print(f"backbone {hw} x{n}: rel err coarse {ec:.2e} fine {ef:.2e}")
    assert ec < 5e-3 and ef < 5e-3


def test_backbone_f16_range_flag_and_fp32_repeat():
    """fp16 activations have TF32's significand but not its exponent range: an input that drives an activation beyond
    65504 must raise the kernels' range flag, and the matcher must then repeat the forward on the exact-fp32 executor
    (same result as a model configured with conv_precision='fp32')."""
    import copy
    import cases
    from onepose_st_b200.nn_common import ResNetFPN_8_2
    from onepose_st_b200.onepose_plus import OnePosePlus_model
    from conftest import backbone_precision
    torch.manual_seed(5)
    m = ResNetFPN_8_2({"block_type": "BasicBlock", "initial_dim": 128, "block_dims": [128, 196, 256]}).cuda().eval()
    img = torch.rand(1, 1, 64, 64, device="cuda")
    flag = torch.zeros(1, dtype=torch.int32, device="cuda")
    with torch.no_grad():
        m(img, _range_flag=flag)
        assert int(flag.item()) == 0
        m(img * 3.0e5, _range_flag=flag)          # stem output ~ 1e5 x the normal activations: beyond fp16
        assert int(flag.item()) == 1
        flag.zero_()
        bad = img.clone()
        bad[0, 0, 5, 5] = float("inf")
        m(bad, _range_flag=flag)
        assert int(flag.item()) == 1
    # model level: the flagged forward equals the fp32-configured forward
    c = cases.case_onepose_full()
    model = OnePosePlus_model(copy.deepcopy(c["cfg"]))
    model.load_state_dict(c["sd"], strict=True)
    model = model.cuda().eval()
    data = {k: (v.cuda() if isinstance(v, torch.Tensor) else v) for k, v in c["data"].items()}
    data["query_image"] = data["query_image"] * 3.0e5
    ref = dict(data)
    with backbone_precision("f16"):
        model(data)
        assert model.__dict__.get("_warned_fp16_range")
    with backbone_precision("fp32"):
        m2 = OnePosePlus_model(copy.deepcopy(c["cfg"]))
        m2.load_state_dict(c["sd"], strict=True)
        m2 = m2.cuda().eval()
        m2(ref)
    assert torch.equal(data["i_ids"], ref["i_ids"]) and torch.equal(data["j_ids"], ref["j_ids"])
    assert torch.equal(data["mconf"], ref["mconf"])


@pytest.mark.parametrize("W", [5, 9])
def test_fine_bf16_arm_tolerance(W):
    """BASELINE.json configs[4] bf16 arm: bf16 windows / query vectors (half the bytes), fp32 accumulation, against the fp32
    kernel and an fp64 evaluation of the same planted inputs.  Stated tolerance: expec_f (x, y) < 2e-2 of the window
    half-width (bf16 keeps 8 significant bits of every feature; the logits of a 128-channel correlation move by ~1e-2)."""
    from onepose_st_b200 import ops, synthetic
    q, win, r = synthetic.planted_fine_windows(31, 3000, W, 128, device="cuda")
    e32 = ops.fine_match(q, win)
    e16 = ops.fine_match(q.to(torch.bfloat16), win.to(torch.bfloat16))
    sim = torch.einsum("mc,mrc->mr", q.double(), win.double()) / 128 ** 0.5
    p = torch.softmax(sim, 1)
    g = torch.linspace(-1, 1, W, dtype=torch.double, device="cuda")
    gx, gy = g.repeat(W), g.repeat_interleave(W)
    ref = torch.stack([(p * gx).sum(1), (p * gy).sum(1)], 1)
    err32 = (e32[:, :2].double() - ref).abs().max().item()
    err16 = (e16[:, :2].double() - ref).abs().max().item()
    print(f"W={W}: expec_f (x, y) max abs err vs fp64: fp32 kernel {err32:.2e}, bf16 arm {err16:.2e}")
    assert err32 < 1e-4 and err16 < 2e-2
    # the gather's bf16 output is the rounded fp32 output
    feat = torch.randn(2, 128, 32, 32, device="cuda")
    b = torch.randint(0, 2, (50,), device="cuda")
    j = torch.randint(0, 64, (50,), device="cuda")
    w32 = ops.fine_gather(feat, b, j, W, 4, 8, 8)
    w16 = ops.fine_gather(feat, b, j, W, 4, 8, 8, dtype=torch.bfloat16)
    assert w16.dtype == torch.bfloat16 and torch.equal(w32.to(torch.bfloat16), w16)


@pytest.mark.parametrize("mode", ["nearest", "bilinear"])
def test_sample_features_vs_grid_sample(mode):
    """the M-point sampling of sample_feature_from_featuremap on the own kernel against the reference's
    formulation (coord_normalization + F.grid_sample, align_corners=True), channels-first and channels-last maps, points on
    and beyond the border, optional L2 normalisation."""
    import torch.nn.functional as F
    from onepose_st_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(4)
    C, H, W = (256, 12, 16) if mode == "nearest" else (128, 48, 64)
    fm = torch.randn(1, C, H, W, device="cuda", generator=g)
    img_h, img_w = 96.0, 128.0
    k = torch.rand(500, 2, device="cuda", generator=g) * torch.tensor([img_w + 6, img_h + 6], device="cuda") - 3
    k[:4] = torch.tensor([[0.0, 0.0], [img_w - 1, img_h - 1], [img_w - 1, 0.0], [63.5, 47.5]], device="cuda")
    grid = k[None, :, None, :] - 1 / 2 + 0.5
    rescale = torch.tensor([(img_w - 1) * 1, (img_h - 1) * 1], device="cuda")
    grid = grid / rescale[None, None, None] * 2 - 1
    want = F.grid_sample(fm, grid.float(), mode=mode, align_corners=True)[0, :, :, 0].t()
    for feat in (fm[0], fm.contiguous(memory_format=torch.channels_last)[0]):
        got = ops.sample_features(feat, k, img_h, img_w, mode == "bilinear")
        err = (got - want).abs().max().item()
        print(f"sample_features {mode}: max abs err vs F.grid_sample {err:.2e}")
        assert err < (1e-6 if mode == "nearest" else 2e-5)
    gn = ops.sample_features(fm[0], k, img_h, img_w, mode == "bilinear", normalize=True)
    assert (gn - F.normalize(want, p=2, dim=-1)).abs().max().item() < 1e-5


def test_track_feature_mean_is_numpy_mean():
    """per-track np.mean(stack(features), axis=0) of feature_aggregation_and_update, bit for bit (fp32, row order)."""
    import numpy as np
    from onepose_st_b200 import ops
    rng = np.random.default_rng(2)
    counts = rng.integers(1, 9, 300)
    counts[5] = 0
    start = np.concatenate([[0], np.cumsum(counts)])
    rows = rng.normal(size=(start[-1], 128)).astype(np.float32)
    got = ops.track_feature_mean(torch.from_numpy(rows).cuda(), torch.from_numpy(start).cuda()).cpu().numpy()
    for t in range(len(counts)):
        if counts[t]:
            want = np.mean(np.stack([rows[r] for r in range(start[t], start[t + 1])], axis=0), axis=0, keepdims=False)
            assert np.array_equal(got[t], want), t
    assert not got[5].any()


def test_detector_batched_matching_equals_per_view_loop():
    """match_query_to_references: all reference views in one LoFTR forward give the matches of the reference's per-view loop."""
    from onepose_st_b200 import synthetic
    from onepose_st_b200.detector import match_query_to_references
    from onepose_st_b200.loftr import LoFTR_for_OnePose_Plus, default_cfg
    torch.manual_seed(0)
    m = LoFTR_for_OnePose_Plus(default_cfg, enable_fine_matching=True).eval()
    m.load_state_dict(synthetic.calibrate_loftr_state_dict(m.state_dict(), 128, 160))
    m = m.cuda()
    g = torch.Generator().manual_seed(21)
    query = torch.rand(1, 1, 128, 160, generator=g).cuda()
    refs = [(query + 0.02 * torch.randn(query.shape, generator=g).cuda()).clamp(0, 1) for _ in range(5)]
    refs.append(torch.rand(1, 1, 96, 128, generator=g).cuda())          # a view of another size: its own forward
    batched = match_query_to_references(m, query, refs)
    for k, r in enumerate(refs):
        d = {"image0": r, "image1": query}
        m(d)
        assert torch.equal(batched[k]["mkpts0_f"], d["mkpts0_f"]) and torch.equal(batched[k]["mconf"], d["mconf"]), k
        assert (batched[k]["mkpts1_f"] - d["mkpts1_f"]).abs().max().item() < 1e-4 if len(d["mconf"]) else True
    assert sum(len(b["mconf"]) for b in batched[:5]) > 50
