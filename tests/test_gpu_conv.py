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


def test_conv_cluster_multicast_option_is_bit_identical():
    """Clusters of two CTAs multicasting the weight slabs (TMA multicast) must give the same bits as the plain kernel."""
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
        assert torch.equal(plain, clustered)
