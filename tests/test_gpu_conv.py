"""GPU parity: tcgen05 implicit-GEMM convolution (NHWC fp16, fp32 accumulation) vs torch's fp32 convolution evaluated on
the SAME fp16-rounded operands (so the only difference is fp32 summation order), plus the fp16 stem / upsample glue."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _ref(x16, w, b, res16, k, stride, act, slope):
    """fp32 reference on the fp16-rounded operands: x16 [N,H,W,Cs] fp16, w [Cout,Cin,k,k] fp32."""
    cin = w.shape[1]
    x = x16[..., :cin].float().permute(0, 3, 1, 2)
    y = F.conv2d(x, w.half().float(), b, stride, k // 2)
    if res16 is not None:
        y = y + res16[..., : w.shape[0]].float().permute(0, 3, 1, 2)
    if act == 1:
        y = F.relu(y)
    elif act == 2:
        y = F.leaky_relu(y, slope)
    return y.permute(0, 2, 3, 1)


CASES = [
    # n, H, W, cin, cout, k, stride, bias, res, act
    (1, 64, 48, 128, 128, 3, 1, True, False, 1),
    (2, 37, 53, 128, 128, 3, 1, True, True, 1),      # ragged tiles
    (1, 40, 40, 196, 196, 3, 1, True, True, 1),      # 208-channel storage, short last K chunk
    (1, 64, 64, 128, 196, 3, 2, True, False, 1),     # stride 2
    (1, 33, 47, 196, 256, 3, 2, True, False, 1),     # stride 2, odd sizes
    (1, 64, 64, 128, 196, 1, 2, True, False, 0),     # 1x1 stride-2 shortcut
    (2, 32, 32, 256, 256, 1, 1, False, False, 0),    # 1x1 lateral
    (1, 48, 48, 256, 256, 3, 1, True, False, 2),     # head conv + LeakyReLU
    (1, 48, 48, 256, 196, 3, 1, False, False, 0),
    (8, 128, 128, 128, 128, 3, 1, True, True, 1),    # enough tiles for the two-sub-tile (256-pixel) variant
    (8, 128, 128, 196, 196, 3, 1, True, True, 1),
    (8, 128, 128, 256, 256, 3, 1, True, False, 2),
]


@pytest.mark.parametrize("n,H,W,cin,cout,k,stride,use_bias,use_res,act", CASES)
def test_conv_f16_vs_torch(n, H, W, cin, cout, k, stride, use_bias, use_res, act):
    from onepose_st_b200 import ops
    from onepose_st_b200.nn_common import pack_conv_f16, _cs
    old = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = False
    try:
        g = torch.Generator(device="cuda").manual_seed(1234 + cin + cout + k + stride)
        cin_s, cout_s = _cs(cin), _cs(cout)
        x16 = torch.zeros(n, H, W, cin_s, dtype=torch.float16, device="cuda")
        x16[..., :cin] = torch.randn(n, H, W, cin, device="cuda", generator=g).half()
        w = torch.randn(cout, cin, k, k, device="cuda", generator=g) / (cin * k * k) ** 0.5
        b = torch.randn(cout, device="cuda", generator=g) if use_bias else None
        wp, bp = pack_conv_f16(w, b)
        ho, wo = (H + 2 * (k // 2) - k) // stride + 1, (W + 2 * (k // 2) - k) // stride + 1
        res16 = None
        if use_res:
            res16 = torch.zeros(n, ho, wo, cout_s, dtype=torch.float16, device="cuda")
            res16[..., :cout] = torch.randn(n, ho, wo, cout, device="cuda", generator=g).half()
        slope = 0.01
        y16, y32 = ops.conv2d_nhwc_f16(x16, wp, k, stride, bp, res16, act, slope, out_f16=True, out_f32=True)
        want = _ref(x16, w, b, res16, k, stride, act, slope)
        assert tuple(y32.shape) == (n, ho, wo, cout_s)
        err = (y32[..., :cout] - want).abs().max().item()
        assert err < 2e-4, err                                   # fp32 accumulation-order noise only
        assert torch.equal(y32.half(), y16)                      # fp16 output = rounded fp32 output
        if cout_s > cout:
            assert (y32[..., cout:] == 0).all()                  # pad channels stay zero
    finally:
        torch.backends.cudnn.allow_tf32 = old


@pytest.mark.parametrize("hw", [(96, 80), (104, 72), (512, 512)])
def test_stem_and_upsample_f16(hw):
    from onepose_st_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(7)
    img = torch.rand(2, 1, hw[0], hw[1], device="cuda", generator=g)
    w = torch.randn(128, 1, 7, 7, device="cuda", generator=g) * 0.1
    b = torch.randn(128, device="cuda", generator=g) * 0.1
    tap_major = w.reshape(128, 49).t().contiguous()
    a32 = ops.stem_conv_relu(img, tap_major, b)                  # fp32 CUDA-core kernel, channels-last [N,C,H,W]
    a16 = ops.stem_conv_relu_f16(img, tap_major, b)              # tensor-core kernel (fp16 operands), [N,H,W,C] fp16
    old = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = False
    try:
        want32 = F.relu(F.conv2d(img, w, b, 2, 3)).permute(0, 2, 3, 1)
        want16 = F.relu(F.conv2d(img.half().float(), w.half().float(), b, 2, 3)).permute(0, 2, 3, 1)
    finally:
        torch.backends.cudnn.allow_tf32 = old
    assert (a32.permute(0, 2, 3, 1) - want32).abs().max().item() < 1e-5
    assert (a16.float() - want16).abs().max().item() < 2e-3     # same fp16 operands: fp16 rounding of the O(1) output only
    assert (a16.float() - want32).abs().max().item() < 5e-3     # TF32-grade w.r.t. the fp32 stem
    src = torch.randn(2, 12, 10, 208, device="cuda", generator=g).half()
    dst = torch.randn(2, 24, 20, 208, device="cuda", generator=g).half()
    want = dst.float() + F.interpolate(src.float().permute(0, 3, 1, 2), scale_factor=2.0, mode="bilinear",
                                       align_corners=True).permute(0, 2, 3, 1)
    got = ops.upsample2x_add_f16_(dst.clone(), src)
    assert (got.float() - want).abs().max().item() < 4e-3        # one fp16 rounding of an O(4) value


@pytest.mark.parametrize("n,H,W,cin,cout", [(2, 24, 40, 128, 196), (1, 64, 64, 196, 256), (8, 128, 128, 128, 196)])
def test_lateral_conv_with_fused_upsample_add(n, H, W, cin, cout):
    """1x1 lateral convolution + bilinear x2 (align_corners=True) of the coarser FPN level in one epilogue."""
    from onepose_st_b200 import ops
    from onepose_st_b200.nn_common import pack_conv_f16, _cs
    old = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = False
    try:
        g = torch.Generator(device="cuda").manual_seed(99 + H)
        cin_s, cout_s = _cs(cin), _cs(cout)
        x16 = torch.zeros(n, H, W, cin_s, dtype=torch.float16, device="cuda")
        x16[..., :cin] = torch.randn(n, H, W, cin, device="cuda", generator=g).half()
        up16 = torch.zeros(n, H // 2, W // 2, cout_s, dtype=torch.float16, device="cuda")
        up16[..., :cout] = torch.randn(n, H // 2, W // 2, cout, device="cuda", generator=g).half()
        w = torch.randn(cout, cin, 1, 1, device="cuda", generator=g) / cin ** 0.5
        wp, _ = pack_conv_f16(w, None)
        got = ops.conv2d_nhwc_f16(x16, wp, 1, 1, None, None, 0, up=up16, out_f16=False, out_f32=True)
        want = _ref(x16, w, None, None, 1, 1, 0, 0.0) + F.interpolate(
            up16[..., :cout].float().permute(0, 3, 1, 2), scale_factor=2.0, mode="bilinear", align_corners=True).permute(0, 2, 3, 1)
        err = (got[..., :cout] - want).abs().max().item()
        assert err < 2e-4, err
        if cout_s > cout:
            assert (got[..., cout:] == 0).all()
    finally:
        torch.backends.cudnn.allow_tf32 = old


def test_conv_cluster_multicast_option_matches_the_default():
    """Clusters of two CTAs multicasting the weight slabs (TMA multicast, one box per tap) against the default kernel (one halo
    box per kernel column for 3x3 stride 1: the taps are added in another order, so the fp32 sums may round differently
    before the fp16 store -- at most an fp16 ulp or two apart)."""
    from onepose_st_b200 import _lib, ops
    from onepose_st_b200.nn_common import pack_conv_f16, _cs
    g = torch.Generator(device="cuda").manual_seed(5)
    lib = _lib.lib()
    for cin, cout, hw in [(128, 128, 128), (196, 196, 128), (256, 256, 128), (196, 196, 120)]:
        x16 = torch.zeros(8, hw, hw, _cs(cin), dtype=torch.float16, device="cuda")
        x16[..., :cin] = torch.randn(8, hw, hw, cin, device="cuda", generator=g).half()
        wp, bp = pack_conv_f16(torch.randn(cout, cin, 3, 3, device="cuda", generator=g) / (cin * 9) ** 0.5,
                               torch.randn(cout, device="cuda", generator=g))
        prev = lib.opb200_conv_cluster(0)
        try:
            plain = ops.conv2d_nhwc_f16(x16, wp, 3, 1, bp, None, 1)
            lib.opb200_conv_cluster(1)
            clustered = ops.conv2d_nhwc_f16(x16, wp, 3, 1, bp, None, 1)
        finally:
            lib.opb200_conv_cluster(prev)
        assert torch.allclose(plain.float(), clustered.float(), rtol=2e-3, atol=2e-3)
        assert (plain != clustered).float().mean().item() < 0.05


@pytest.mark.parametrize("n,H,W,cin,cout,k,stride,use_res", [(2, 32, 48, 128, 128, 3, 1, True), (1, 40, 24, 128, 196, 3, 2, False),
                                                            (1, 24, 24, 196, 256, 1, 2, False), (2, 16, 16, 256, 256, 3, 1, True)])
def test_conv_split_precision_is_fp32_grade(n, H, W, cin, cout, k, stride, use_res):
    """Split-precision convolution (opb200_conv2d_nhwc_f16x3: [hi | lo] fp16 planes, three products) against torch fp32 on the
    SAME fp32 inputs -- no pre-rounding of the reference's operands: the error must be fp32-grade (~1e-6 relative), three
    orders below the single-product fp16 kernel's."""
    from onepose_st_b200 import ops
    from onepose_st_b200.nn_common import pack_conv_f16x3, pack_conv_f16, _cs
    old = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = False
    try:
        g = torch.Generator(device="cuda").manual_seed(77 + cin + cout)
        cin_s, cout_s = _cs(cin), _cs(cout)
        x = torch.randn(n, H, W, cin, device="cuda", generator=g)
        w = torch.randn(cout, cin, k, k, device="cuda", generator=g) / (cin * k * k) ** 0.5
        b = torch.randn(cout, device="cuda", generator=g)
        xs = torch.zeros(n, H, W, cin_s, device="cuda")
        xs[..., :cin] = x
        x2 = ops.split_f16_planes(xs)
        assert (x2[..., :cin_s].float() + x2[..., cin_s:].float() - xs).abs().max().item() < 1e-6
        wp, bp, s = pack_conv_f16x3(w, b)
        Ho, Wo = (H + 2 * (k // 2) - k) // stride + 1, (W + 2 * (k // 2) - k) // stride + 1
        res2 = None
        want = F.conv2d(x.permute(0, 3, 1, 2), w, b, stride, k // 2).permute(0, 2, 3, 1)
        if use_res:
            r = torch.zeros(n, Ho, Wo, cout_s, device="cuda")
            r[..., :cout] = torch.randn(n, Ho, Wo, cout, device="cuda", generator=g)
            res2 = ops.split_f16_planes(r)
            want = want + r[..., :cout]
        want = F.relu(want)
        y16, y32 = ops.conv2d_nhwc_f16(x2, wp, k, stride, bp, res2, 1, out_f16=True, out_f32=True, split_scale=s)
        err = (y32[..., :cout] - want).abs().max().item() / want.abs().max().item()
        rec = (y16[..., :cout_s].float() + y16[..., cout_s:].float() - y32).abs().max().item()
        # the single-product kernel on the same inputs, for scale
        wp1, bp1 = pack_conv_f16(w, b)
        y1 = ops.conv2d_nhwc_f16(xs.half(), wp1, k, stride, bp1, None if res2 is None else r.half(), 1, out_f16=False, out_f32=True)
        err1 = (y1[..., :cout] - want).abs().max().item() / want.abs().max().item()
        print(f"conv {cin}->{cout} k{k} s{stride}: split rel err {err:.2e} (single fp16 product {err1:.2e}), hi+lo reconstruction {rec:.2e}")
        assert err < 1e-5 and err < 0.05 * err1 and rec < 2e-6 * max(1.0, want.abs().max().item())
        if cout_s > cout:
            assert (y32[..., cout:] == 0).all() and (y16[..., cout:cout_s] == 0).all()
    finally:
        torch.backends.cudnn.allow_tf32 = old


def test_backbone_split_precision_vs_golden(golden):
    """conv_precision='f16x3': the whole ResNetFPN_8_2 on the own split-precision convolutions against the reference's fp32
    golden maps -- fp32-grade (the same bound the cuDNN fp32 executor is held to)."""
    import cases
    from onepose_st_b200.nn_common import ResNetFPN_8_2, FusedBackboneF16x3
    c = cases.case_backbone()
    m = ResNetFPN_8_2({"block_type": "BasicBlock", "initial_dim": 128, "block_dims": [128, 196, 256]})
    m.load_state_dict({k[len("backbone."):]: v for k, v in c["sd"].items()}, strict=True)
    m = m.cuda().eval()
    m.conv_precision = "f16x3"
    flag = torch.zeros(1, dtype=torch.int32, device="cuda")
    with torch.no_grad():
        assert isinstance(m.fused(), FusedBackboneF16x3)
        fc, ff = m(c["image"].cuda(), _range_flag=flag)
    g = golden("backbone")
    gc, gf = g["feat_c"], g["feat_f_sub"]
    ec = (fc.cpu() - gc).abs().max().item() / gc.abs().max().item()
    ef = (ff.cpu()[:, ::8, ::4, ::4] - gf).abs().max().item() / gf.abs().max().item()
    print(f"split-precision backbone rel err coarse {ec:.2e} fine {ef:.2e}")
    assert int(flag.item()) == 0 and ec < 1e-4 and ef < 1e-4
