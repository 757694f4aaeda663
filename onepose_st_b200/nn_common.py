"""Building blocks shared by the OnePosePlus (2D-3D) and LoFTR (2D-2D) drop-in matchers.

Everything that is a hot-path operator calls the sm_100a C-ABI through ``ops``; the modules here only
hold parameters under the reference's names (so reference checkpoints load with ``strict=True``),
cache packed weights, and move pointers around.  The ResNet-FPN backbone (SURVEY.md 8a row a2 / first of the
"next" rows in 8f) runs on own tcgen05 convolutions at TF32-grade precision, or on cuDNN when exact fp32 is asked.
"""
import math
import threading
from typing import Optional, Sequence

import torch
import torch.nn as nn
import torch.nn.functional as F

from . import _lib, ops

_cache_lock = threading.RLock()  # check-then-set of the per-sequence caches (host threads of pipeline.StreamPipeline)


def tensor_key(t: torch.Tensor):
    """Identity of a tensor's current contents for the per-sequence caches: (address, version counter, shape, strides).
    Only meaningful while the cache entry holds a reference to ``t`` (the address cannot be recycled then).  ``None``
    for inference tensors, which track no version: callers must not cache on them."""
    if t.is_inference():
        return None
    return (t.data_ptr(), t._version, tuple(t.shape), tuple(t.stride()))


# --------------------------------------------------------------------------------------------------
# Backbone: ResNetFPN_8_2 (reference: src/models/OnePosePlus/backbone/resnet.py:85-164 and
# submodules/LoFTR/src/loftr/backbone/resnet_fpn.py:43-118).  Same parameter names, PyTorch/cuDNN math.
# --------------------------------------------------------------------------------------------------
class BasicBlock(nn.Module):
    def __init__(self, cin: int, cout: int, stride: int = 1):
        super().__init__()
        self.conv1 = nn.Conv2d(cin, cout, 3, stride, 1, bias=False)
        self.conv2 = nn.Conv2d(cout, cout, 3, 1, 1, bias=False)
        self.bn1 = nn.BatchNorm2d(cout)
        self.bn2 = nn.BatchNorm2d(cout)
        self.relu = nn.ReLU(inplace=True)
        self.downsample = None
        if stride != 1:
            self.downsample = nn.Sequential(nn.Conv2d(cin, cout, 1, stride, 0, bias=False), nn.BatchNorm2d(cout))

    def forward(self, x):
        y = self.bn2(self.conv2(self.relu(self.bn1(self.conv1(x)))))
        if self.downsample is not None:
            x = self.downsample(x)
        return self.relu(x + y)


class ResNetFPN_8_2(nn.Module):
    """ResNet + FPN with 1/8 (coarse) and 1/2 (fine) outputs.  ``forward`` returns (coarse, fine)."""

    def __init__(self, config):
        super().__init__()
        if config.get("block_type", "BasicBlock") != "BasicBlock":
            raise NotImplementedError("only BasicBlock backbones are supported")
        d0 = config["initial_dim"]
        d1, d2, d3 = config["block_dims"]
        self.conv1 = nn.Conv2d(1, d0, 7, 2, 3, bias=False)
        self.bn1 = nn.BatchNorm2d(d0)
        self.relu = nn.ReLU(inplace=True)
        self.layer1 = nn.Sequential(BasicBlock(d0, d1, 1), BasicBlock(d1, d1, 1))
        self.layer2 = nn.Sequential(BasicBlock(d1, d2, 2), BasicBlock(d2, d2, 1))
        self.layer3 = nn.Sequential(BasicBlock(d2, d3, 2), BasicBlock(d3, d3, 1))
        self.layer3_outconv = nn.Conv2d(d3, d3, 1, bias=False)
        self.layer2_outconv = nn.Conv2d(d2, d3, 1, bias=False)
        self.layer2_outconv2 = self._head(d3, d3, d2)
        self.layer1_outconv = nn.Conv2d(d1, d2, 1, bias=False)
        self.layer1_outconv2 = self._head(d2, d2, d1)
        for m in self.modules():
            if isinstance(m, nn.Conv2d):
                nn.init.kaiming_normal_(m.weight, mode="fan_out", nonlinearity="relu")

    @staticmethod
    def _head(cin, cmid, cout):
        return nn.Sequential(nn.Conv2d(cin, cmid, 3, 1, 1, bias=False), nn.BatchNorm2d(cmid), nn.LeakyReLU(),
                             nn.Conv2d(cmid, cout, 3, 1, 1, bias=False))

    def forward(self, x, _precision: Optional[str] = None, _range_flag: Optional[torch.Tensor] = None):
        """``_precision`` overrides ``conv_precision`` for this call; ``_range_flag`` (int32 [1] on the device, optional)
        receives bit 0 when the fp16 executor had to store a value outside the fp16 range (see ``conv_precision``)."""
        if x.is_cuda and not self.training and not torch.is_grad_enabled():
            fb = self.fused(_precision)
            return fb(x, _range_flag) if isinstance(fb, (FusedBackboneF16, FusedBackboneF16x3)) else fb(x)
        x1 = self.layer1(self.relu(self.bn1(self.conv1(x))))
        x2 = self.layer2(x1)
        x3 = self.layer3(x2)
        c3 = self.layer3_outconv(x3)
        up = lambda t: F.interpolate(t, scale_factor=2.0, mode="bilinear", align_corners=True)
        c2 = self.layer2_outconv2(self.layer2_outconv(x2) + up(c3))
        c1 = self.layer1_outconv2(self.layer1_outconv(x1) + up(c2))
        return c3, c1

    # Precision of the inference executor, an explicit per-module (or per-class) setting -- no global switch decides it:
    #   "f16"  (default) own tcgen05 implicit-GEMM convolutions on fp16 operands with fp32 accumulation: the 11-bit
    #          significand of the TF32 operands the reference's cuDNN convolutions use on a GPU, but fp16's exponent
    #          range (|x| <= 65504).  The kernels flag any stored value that leaves that range and the matchers then
    #          repeat the backbone with "fp32" for that call (OnePosePlus_model / LoFTR_for_OnePose_Plus.forward).
    #   "f16x3" the same kernels on split operands ([hi | lo] fp16 planes, three products per k-block): fp32-grade feature
    #          maps on own kernels at ~3x the convolution time -- the configuration in which the whole forward holds the
    #          fp32 contract without library code (bench.py --backbone f16x3).
    #   "fp32" exact-fp32 cuDNN executor (library code; what the fp32-grade parity tests select).
    conv_precision = "f16"

    def fused(self, precision: Optional[str] = None):
        """Inference executor (BN folded into the convolutions, NHWC): ``FusedBackboneF16`` (own tcgen05 implicit-GEMM
        convolutions) or ``FusedBackbone`` (cuDNN fp32); rebuilt whenever a parameter or buffer changes."""
        precision = precision or self.conv_precision
        if precision not in ("f16", "f16x3", "fp32"):
            raise ValueError(f"conv_precision must be 'f16', 'f16x3' or 'fp32', got {precision!r}")
        f16 = precision if (precision != "fp32" and FusedBackboneF16.supports(self)) else False
        key = tuple((t.data_ptr(), t._version) for t in list(self.parameters()) + list(self.buffers()))
        cache = self.__dict__.setdefault("_fused_cache", {})
        fb = cache.get(f16)
        if fb is None or fb[0] != key:
            with _cache_lock:
                fb = cache.get(f16)
                if fb is None or fb[0] != key:
                    fb = (key, FusedBackboneF16x3(self) if f16 == "f16x3" else (FusedBackboneF16(self) if f16 else FusedBackbone(self)))
                    dev = self.conv1.weight.device
                    if dev.type == "cuda":
                        torch.cuda.current_stream(dev).synchronize()  # folded weights are read from other streams too
                    cache[f16] = fb
        return fb[1]

    def __getstate__(self):
        st = self.__dict__.copy()
        st.pop("_fused_cache", None)
        return st


def _fold_bn(conv: nn.Conv2d, bn: Optional[nn.BatchNorm2d]):
    """Eval-mode BatchNorm folded into the preceding bias-free convolution: (w', b')."""
    w = conv.weight.detach().float()
    if bn is None:
        b = conv.bias.detach().float() if conv.bias is not None else torch.zeros(w.shape[0], device=w.device)
        return w.contiguous(memory_format=torch.channels_last), b
    g = bn.weight.detach().float() / torch.sqrt(bn.running_var.float() + bn.eps)
    w = w * g[:, None, None, None]
    b = bn.bias.detach().float() - bn.running_mean.float() * g
    if conv.bias is not None:
        b = b + conv.bias.detach().float() * g
    return w.contiguous(memory_format=torch.channels_last), b.contiguous()


def _cs(c: int) -> int:
    """Channel count padded to the multiple of 16 the fp16 tensor-core convolution works in."""
    return (c + 15) // 16 * 16


def pack_conv_f16(w: torch.Tensor, b: Optional[torch.Tensor]):
    """[Cout, Cin, k, k] fp32 (BatchNorm already folded) -> ([k*k*Cout_s, Cin_s] fp16 tap-major, [Cout_s] fp32 bias or
    None), zero in the padded rows / columns."""
    cout, cin, k, _ = w.shape
    cin_s, cout_s = _cs(cin), _cs(cout)
    wp = torch.zeros(k * k, cout_s, cin_s, dtype=torch.float32, device=w.device)
    wp[:, :cout, :cin] = w.detach().float().permute(2, 3, 0, 1).reshape(k * k, cout, cin)
    bp = None
    if b is not None:
        bp = torch.zeros(cout_s, dtype=torch.float32, device=w.device)
        bp[:cout] = b.detach().float()
    return wp.reshape(k * k * cout_s, cin_s).half().contiguous(), bp


def pack_conv_f16x3(w: torch.Tensor, b: Optional[torch.Tensor]):
    """Split-precision packing: [Cout, Cin, k, k] fp32 -> ([k*k*Cout_s, 2*Cin_s] fp16 = [hi | lo] planes of s*w, bias,
    s) with s the largest power of two keeping max|s*w| <= 1024 (so that the lo plane, ~2^-11 of the hi plane, stays in
    fp16's normal range for all but negligible weights; the kernel multiplies the accumulator by 1/s)."""
    cout, cin, k, _ = w.shape
    cin_s, cout_s = _cs(cin), _cs(cout)
    wf = w.detach().float()
    wmax = float(wf.abs().max().item())
    s = 2.0 ** math.floor(math.log2(1024.0 / wmax)) if wmax > 0 else 1.0
    s = min(max(s, 2.0 ** -10), 2.0 ** 20)
    ws = (wf * s).permute(2, 3, 0, 1).reshape(k * k, cout, cin)
    hi = ws.half()
    lo = (ws - hi.float()).half()
    wp = torch.zeros(k * k, cout_s, 2 * cin_s, dtype=torch.float16, device=w.device)
    wp[:, :cout, :cin] = hi
    wp[:, :cout, cin_s: cin_s + cin] = lo
    bp = None
    if b is not None:
        bp = torch.zeros(cout_s, dtype=torch.float32, device=w.device)
        bp[:cout] = b.detach().float()
    return wp.reshape(k * k * cout_s, 2 * cin_s).contiguous(), bp, s


class FusedBackbone:
    """Inference-only executor of a ``ResNetFPN_8_2``: same math with BatchNorm folded into the conv
    weights, NHWC (channels-last) activations end to end, and cuDNN's fused conv+bias+ReLU /
    conv+bias+residual+ReLU kernels for the BasicBlocks.  Outputs are channels-last, which is also the
    layout the transformer stage-in and the fine window gather read coalesced.  (PyTorch/cuDNN library
    code: the backbone is not one of north_star's hand-written kernels)"""

    def __init__(self, m: "ResNetFPN_8_2"):
        self.stem = _fold_bn(m.conv1, m.bn1)
        # own stem kernel (Conv2d(1, 128, 7, 2, 3) + bias + ReLU, channels-last output): weights tap-major [49][Cout]
        self.stem_tap_major = None
        if tuple(m.conv1.weight.shape[1:]) == (1, 7, 7) and m.conv1.weight.shape[0] == 128 and m.conv1.stride == (2, 2) \
                and m.conv1.padding == (3, 3):
            self.stem_tap_major = self.stem[0].reshape(128, 49).t().contiguous()
        self.blocks = []
        for layer in (m.layer1, m.layer2, m.layer3):
            for blk in layer:
                ds = _fold_bn(blk.downsample[0], blk.downsample[1]) if blk.downsample is not None else None
                self.blocks.append((_fold_bn(blk.conv1, blk.bn1), _fold_bn(blk.conv2, blk.bn2), ds, blk.conv1.stride))
        self.l3_out = _fold_bn(m.layer3_outconv, None)
        self.l2_out = _fold_bn(m.layer2_outconv, None)
        self.l1_out = _fold_bn(m.layer1_outconv, None)
        self.head2 = (_fold_bn(m.layer2_outconv2[0], m.layer2_outconv2[1]), _fold_bn(m.layer2_outconv2[3], None),
                      m.layer2_outconv2[2].negative_slope)
        self.head1 = (_fold_bn(m.layer1_outconv2[0], m.layer1_outconv2[1]), _fold_bn(m.layer1_outconv2[3], None),
                      m.layer1_outconv2[2].negative_slope)

    @staticmethod
    def _conv_relu(x, wb, stride, padding):
        return torch.cudnn_convolution_relu(x, wb[0], wb[1], stride, padding, (1, 1), 1)

    def _block(self, x, blk):
        c1, c2, ds, stride = blk
        y = self._conv_relu(x, c1, stride, (1, 1))
        if ds is not None:
            x = F.conv2d(x, ds[0], ds[1], stride)
        return torch.cudnn_convolution_add_relu(y, c2[0], x, 1.0, c2[1], (1, 1), (1, 1), (1, 1), 1)

    @staticmethod
    def _head(x, head):
        (w0, b0), (w1, b1), slope = head
        y = F.conv2d(x, w0, None, 1, 1).contiguous(memory_format=torch.channels_last)
        return F.conv2d(ops.bias_leaky_relu_(y, b0, slope), w1, None, 1, 1)

    def __call__(self, x):
        if self.stem_tap_major is not None:
            x = ops.stem_conv_relu(x, self.stem_tap_major, self.stem[1])
        else:  # non-default initial_dim: cuDNN stem (its NCHW result is converted once)
            x = self._conv_relu(x.float(), self.stem, (2, 2), (3, 3)).contiguous(memory_format=torch.channels_last)
        feats = []
        for i, blk in enumerate(self.blocks):
            x = self._block(x, blk)
            if i % 2 == 1:
                feats.append(x)
        x1, x2, x3 = feats
        c3 = F.conv2d(x3, self.l3_out[0])
        if c3.shape[1] % 4 == 0 and x2.shape[2] == 2 * x3.shape[2] and x1.shape[2] == 2 * x2.shape[2] \
                and x2.shape[3] == 2 * x3.shape[3] and x1.shape[3] == 2 * x2.shape[3]:
            cl = lambda t: t.contiguous(memory_format=torch.channels_last)
            c2 = self._head(ops.upsample2x_add_(cl(F.conv2d(x2, self.l2_out[0])), cl(c3)), self.head2)
            c1 = self._head(ops.upsample2x_add_(cl(F.conv2d(x1, self.l1_out[0])), cl(c2)), self.head1)
            return c3, c1
        raise NotImplementedError("ResNetFPN_8_2 on CUDA needs image sides that are multiples of 8")


class FusedBackboneF16:
    """Inference-only executor of a ``ResNetFPN_8_2`` on the tcgen05 convolution kernel (csrc/conv.cu): BatchNorm
    folded, NHWC fp16 activations (channels padded to a multiple of 16), fp32 accumulation, bias / residual / ReLU /
    LeakyReLU fused into the convolution epilogues; only the 7x7 stem and the bilinear x2 + add of the FPN are
    separate (own) kernels.  The two outputs are written in fp32 by the last epilogues and returned as
    channels-last [N,C,H,W] views, like ``FusedBackbone``."""

    @staticmethod
    def supports(m: "ResNetFPN_8_2") -> bool:
        d0 = m.conv1.weight.shape[0]
        dims = [m.layer1[0].conv1.weight.shape[0], m.layer2[0].conv1.weight.shape[0], m.layer3[0].conv1.weight.shape[0]]
        return (tuple(m.conv1.weight.shape) == (128, 1, 7, 7) and d0 == 128 and all(_cs(d) in (128, 208, 256) for d in dims)
                and _cs(dims[2]) == dims[2] and _cs(dims[0]) == dims[0])

    def __init__(self, m: "ResNetFPN_8_2"):
        w, b = _fold_bn(m.conv1, m.bn1)
        self.stem = (w.reshape(128, 49).t().contiguous(), b)
        pk = lambda conv, bn: pack_conv_f16(*_fold_bn(conv, bn)) if (bn is not None or conv.bias is not None) \
            else (pack_conv_f16(conv.weight, None)[0], None)
        self.blocks = []
        for layer in (m.layer1, m.layer2, m.layer3):
            for blk in layer:
                ds = pk(blk.downsample[0], blk.downsample[1]) if blk.downsample is not None else None
                self.blocks.append((pk(blk.conv1, blk.bn1), pk(blk.conv2, blk.bn2), ds, blk.conv1.stride[0]))
        self.l3_out, self.l2_out, self.l1_out = pk(m.layer3_outconv, None), pk(m.layer2_outconv, None), pk(m.layer1_outconv, None)
        self.head2 = (pk(m.layer2_outconv2[0], m.layer2_outconv2[1]), pk(m.layer2_outconv2[3], None),
                      m.layer2_outconv2[2].negative_slope)
        self.head1 = (pk(m.layer1_outconv2[0], m.layer1_outconv2[1]), pk(m.layer1_outconv2[3], None),
                      m.layer1_outconv2[2].negative_slope)

    @staticmethod
    def _block(x, blk, rf):
        (w1, b1), (w2, b2), ds, stride = blk
        y = ops.conv2d_nhwc_f16(x, w1, 3, stride, b1, None, 1, range_flag=rf)
        if ds is not None:
            x = ops.conv2d_nhwc_f16(x, ds[0], 1, stride, ds[1], None, 0, range_flag=rf)
        return ops.conv2d_nhwc_f16(y, w2, 3, 1, b2, x, 1, range_flag=rf)

    @staticmethod
    def _head(x, head, out_f32, rf):
        (w0, b0), (w1, b1), slope = head
        y = ops.conv2d_nhwc_f16(x, w0, 3, 1, b0, None, 2, slope, range_flag=rf)
        return ops.conv2d_nhwc_f16(y, w1, 3, 1, b1, None, 0, out_f16=not out_f32, out_f32=out_f32, range_flag=rf)

    def __call__(self, x, range_flag: Optional[torch.Tensor] = None):
        """``range_flag``: int32 [1] device tensor that collects the kernels' "value left the fp16 range" bit."""
        if x.shape[2] % 8 or x.shape[3] % 8:
            raise NotImplementedError("ResNetFPN_8_2 on CUDA needs image sides that are multiples of 8")
        rf = range_flag
        x = ops.stem_conv_relu_f16(x, self.stem[0], self.stem[1], rf)
        feats = []
        for i, blk in enumerate(self.blocks):
            x = self._block(x, blk, rf)
            if i % 2 == 1:
                feats.append(x)
        x1, x2, x3 = feats
        c3_16, c3 = ops.conv2d_nhwc_f16(x3, self.l3_out[0], 1, 1, self.l3_out[1], None, 0, out_f16=True, out_f32=True,
                                        range_flag=rf)
        # lateral 1x1 convolution + bilinear x2 of the coarser level, one kernel per FPN level
        l2 = ops.conv2d_nhwc_f16(x2, self.l2_out[0], 1, 1, self.l2_out[1], None, 0, up=c3_16, range_flag=rf)
        c2 = self._head(l2, self.head2, False, rf)
        l1 = ops.conv2d_nhwc_f16(x1, self.l1_out[0], 1, 1, self.l1_out[1], None, 0, up=c2, range_flag=rf)
        c1 = self._head(l1, self.head1, True, rf)
        return c3.permute(0, 3, 1, 2), c1.permute(0, 3, 1, 2)


class FusedBackboneF16x3:
    """Inference-only executor of a ``ResNetFPN_8_2`` on the SPLIT-PRECISION tcgen05 convolutions (csrc/conv.cu, terms = 3):
    every activation and weight travels as two fp16 planes [hi | lo] (22 significant bits) and each k-block is issued for
    hi*hi + lo*hi + hi*lo with fp32 accumulation -- fp32-grade feature maps on own kernels (conv_precision = "f16x3"), at
    three times the tensor work and twice the activation traffic of "f16".  The 7x7 stem runs on the exact fp32 CUDA-core
    kernel and is split once; the rest mirrors ``FusedBackboneF16``."""

    supports = staticmethod(lambda m: FusedBackboneF16.supports(m))

    def __init__(self, m: "ResNetFPN_8_2"):
        w, b = _fold_bn(m.conv1, m.bn1)
        self.stem = (w.reshape(128, 49).t().contiguous(), b)
        pk = lambda conv, bn: pack_conv_f16x3(*_fold_bn(conv, bn)) if (bn is not None or conv.bias is not None) \
            else pack_conv_f16x3(conv.weight, None)
        self.blocks = []
        for layer in (m.layer1, m.layer2, m.layer3):
            for blk in layer:
                ds = pk(blk.downsample[0], blk.downsample[1]) if blk.downsample is not None else None
                self.blocks.append((pk(blk.conv1, blk.bn1), pk(blk.conv2, blk.bn2), ds, blk.conv1.stride[0]))
        self.l3_out, self.l2_out, self.l1_out = pk(m.layer3_outconv, None), pk(m.layer2_outconv, None), pk(m.layer1_outconv, None)
        self.head2 = (pk(m.layer2_outconv2[0], m.layer2_outconv2[1]), pk(m.layer2_outconv2[3], None),
                      m.layer2_outconv2[2].negative_slope)
        self.head1 = (pk(m.layer1_outconv2[0], m.layer1_outconv2[1]), pk(m.layer1_outconv2[3], None),
                      m.layer1_outconv2[2].negative_slope)

    @staticmethod
    def _conv(x, p, k, stride, res, act, rf, slope=0.0, up=None, out_f16=True, out_f32=False):
        w, b, s = p
        return ops.conv2d_nhwc_f16(x, w, k, stride, b, res, act, slope, out_f16=out_f16, out_f32=out_f32, up=up, range_flag=rf,
                                   split_scale=s)

    def _block(self, x, blk, rf):
        c1, c2, ds, stride = blk
        y = self._conv(x, c1, 3, stride, None, 1, rf)
        if ds is not None:
            x = self._conv(x, ds, 1, stride, None, 0, rf)
        return self._conv(y, c2, 3, 1, x, 1, rf)

    def _head(self, x, head, out_f32, rf):
        c0, c1, slope = head
        y = self._conv(x, c0, 3, 1, None, 2, rf, slope)
        return self._conv(y, c1, 3, 1, None, 0, rf, out_f16=not out_f32, out_f32=out_f32)

    def __call__(self, x, range_flag: Optional[torch.Tensor] = None):
        if x.shape[2] % 8 or x.shape[3] % 8:
            raise NotImplementedError("ResNetFPN_8_2 on CUDA needs image sides that are multiples of 8")
        rf = range_flag
        s32 = ops.stem_conv_relu(x, self.stem[0], self.stem[1])                       # [N,128,Ho,Wo] channels-last fp32
        x = ops.split_f16_planes(s32.permute(0, 2, 3, 1), rf)                          # NHWC view -> [hi | lo] fp16
        feats = []
        for i, blk in enumerate(self.blocks):
            x = self._block(x, blk, rf)
            if i % 2 == 1:
                feats.append(x)
        x1, x2, x3 = feats
        c3_16, c3 = self._conv(x3, self.l3_out, 1, 1, None, 0, rf, out_f16=True, out_f32=True)
        l2 = self._conv(x2, self.l2_out, 1, 1, None, 0, rf, up=c3_16)
        c2 = self._head(l2, self.head2, False, rf)
        l1 = self._conv(x1, self.l1_out, 1, 1, None, 0, rf, up=c2)
        c1 = self._head(l1, self.head1, True, rf)
        return c3.permute(0, 3, 1, 2), c1.permute(0, 3, 1, 2)


# --------------------------------------------------------------------------------------------------
# CUDA graphs of the shape-static head of a matcher forward (backbone -> coarse stack -> coarse matching kernels).
# --------------------------------------------------------------------------------------------------
class HeadGraphs:
    """The head is ~70-110 kernel launches of fixed shape for given input sizes: captured once per (CUDA stream, key)
    and replayed as ONE launch, which keeps the host ahead of the GPU when several forwards are in flight
    (pipeline.StreamPipeline).  Each graph owns its workspaces (``ops.workspaces.scope``) and static input buffers;
    whatever ``fn`` returns stays alive with the graph and is overwritten by the next replay on that stream."""

    _lock = threading.Lock()
    _serial = 0

    def __init__(self, max_entries: int = 16):
        self.entries = {}  # insertion / use ordered: the first key is the least recently used
        self.max_entries = max_entries

    def __len__(self):
        return len(self.entries)

    def _drop(self, key):
        """Forget one graph and free the workspaces it owned (they live in the graph's private memory pool)."""
        ent = self.entries.pop(key, None)
        if ent is not None:
            ops.workspaces.drop_scope(ent[5])

    def clear(self):
        for key in list(self.entries):
            self._drop(key)

    def __del__(self):
        try:
            self.clear()
        except Exception:  # interpreter shutdown
            pass

    def run(self, key, inputs, fn, guard=None):
        """``fn(*static_inputs)`` replayed on the current stream with ``inputs`` copied into the static buffers.
        ``guard()`` (optional) is compared with its value at capture time: a mismatch (a per-sequence cache was rebuilt
        at another address) drops the graph."""
        dev = inputs[0].device
        key = (torch.cuda.current_stream(dev).cuda_stream, dev.index) + tuple(key) + tuple(tuple(t.shape) for t in inputs)
        ent = self.entries.get(key)
        if ent is not None and guard is not None and ent[4] != guard():
            self._drop(key)
            ent = None
        if ent is None:
            static = [torch.empty_like(t, memory_format=torch.contiguous_format) for t in inputs]
            for st, t in zip(static, inputs):
                st.copy_(t)
            fn(*static)  # eager run: builds every cache / workspace / function attribute the capture relies on
            torch.cuda.current_stream(dev).synchronize()
            with HeadGraphs._lock:
                HeadGraphs._serial += 1
                scope = ("graph", HeadGraphs._serial)
                g = torch.cuda.CUDAGraph()
                l0 = _lib.launch_count()
                with ops.workspaces.scope(scope), torch.cuda.graph(g, capture_error_mode="thread_local"):
                    out = fn(*static)
                n_launch = _lib.launch_count() - l0  # this library's kernels inside one replay
            ent = (g, static, out, n_launch, None if guard is None else guard(), scope)
            while len(self.entries) >= self.max_entries:
                self._drop(next(iter(self.entries)))  # least recently used
            self.entries[key] = ent
        else:
            self.entries[key] = self.entries.pop(key)  # mark as most recently used
        g, static, out, n_launch = ent[:4]
        for st, t in zip(static, inputs):
            st.copy_(t, non_blocking=True)
        g.replay()
        _lib.launch_count_add(n_launch)
        return out


# --------------------------------------------------------------------------------------------------
# Sine position encoding (reference: OnePosePlus utils/position_encoding.py:13-42; LoFTR :11-42).
# ``pe`` stays a non-persistent buffer of the reference's shape; a token-major [H*W, C] table of the
# current feature-map size is cached and added inside the transformer's stage-in kernel.
# --------------------------------------------------------------------------------------------------
class PositionEncodingSine(nn.Module):
    def __init__(self, d_model, max_shape=(256, 256), temp_bug_fix=False):
        super().__init__()
        h, w = tuple(max_shape)
        ypos = torch.arange(1, h + 1, dtype=torch.float32).view(1, h, 1).expand(1, h, w)
        xpos = torch.arange(1, w + 1, dtype=torch.float32).view(1, 1, w).expand(1, h, w)
        # NB: without temp_bug_fix the reference evaluates -log(1e4)/d_model//2 == -1.0 (floor division)
        coef = (-math.log(10000.0) / (d_model // 2)) if temp_bug_fix else (-math.log(10000.0) / d_model // 2)
        div = torch.exp(torch.arange(0, d_model // 2, 2).float() * coef)[:, None, None]
        pe = torch.zeros(d_model, h, w)
        pe[0::4], pe[1::4] = torch.sin(xpos * div), torch.cos(xpos * div)
        pe[2::4], pe[3::4] = torch.sin(ypos * div), torch.cos(ypos * div)
        self.register_buffer("pe", pe.unsqueeze(0), persistent=False)
        self._tables = {}

    def forward(self, x):
        return x + self.pe[:, :, : x.size(2), : x.size(3)]

    def token_table(self, h: int, w: int) -> torch.Tensor:
        """[h*w, C] fp32 table equal to ``pe[0, :, :h, :w]`` rearranged 'c h w -> (h w) c'."""
        key = (h, w, self.pe.device)
        t = self._tables.get(key)
        if t is None:
            with _cache_lock:
                t = self._tables.get(key)
                if t is None:
                    t = self.pe[0, :, :h, :w].reshape(self.pe.shape[1], h * w).t().contiguous()
                    if t.is_cuda:
                        torch.cuda.current_stream(t.device).synchronize()  # other streams may read it right away
                    if len(self._tables) >= 8:  # tables stay alive (CUDA graphs of the matcher head read them by address)
                        self._tables = {}
                    self._tables[key] = t
        return t


# --------------------------------------------------------------------------------------------------
# Encoder layer parameter holder + transformer base
# --------------------------------------------------------------------------------------------------
class LinearAttention(nn.Module):
    """elu+1 linear attention (reference: loftr_module/linear_attention.py:22-61).  Inside the encoder
    stack the attention is fused into the tcgen05 kernels; this standalone module serves callers that
    use the operator on its own."""

    def __init__(self, d_head=None, kernel_fn="elu + 1", eps=1e-6, redraw_interval=2, d_kernel=None):
        super().__init__()
        if kernel_fn not in (None, "elu + 1"):
            raise ValueError(f"unsupported kernel_fn {kernel_fn!r}")
        self.eps = eps

    def forward(self, queries, keys, values, q_mask=None, kv_mask=None):
        return ops.linear_attention(queries, keys, values, q_mask, kv_mask, self.eps)


class LoFTREncoderLayer(nn.Module):
    """Holds q/k/v/merge/mlp/norm parameters under the reference's names
    (transformer.py:10-63) and packs them for the kernels on demand."""

    def __init__(self, d_model, nhead, dropout=0.0, attention="linear", kernel_fn="elu + 1", redraw_interval=1,
                 d_kernel=None, rezero=None, norm_method="layernorm"):
        super().__init__()
        if attention != "linear":
            raise NotImplementedError("only attention='linear' is implemented (FullAttention is never configured)")
        if kernel_fn != "elu + 1":
            raise ValueError(f"unsupported kernel_fn {kernel_fn!r}")
        if norm_method != "layernorm":
            raise NotImplementedError("encoder layers support norm_method='layernorm' only")
        if rezero is not None:
            raise NotImplementedError("rezero is not supported")
        if dropout not in (0, 0.0, None):
            raise NotImplementedError("inference path: dropout must be 0")
        self.dim, self.nhead = d_model // nhead, nhead
        self.q_proj = nn.Linear(d_model, d_model, bias=False)
        self.k_proj = nn.Linear(d_model, d_model, bias=False)
        self.v_proj = nn.Linear(d_model, d_model, bias=False)
        self.attention = LinearAttention(self.dim)
        self.merge = nn.Linear(d_model, d_model, bias=False)
        self.mlp = nn.Sequential(nn.Linear(2 * d_model, 2 * d_model, bias=False), nn.ReLU(True),
                                 nn.Linear(2 * d_model, d_model, bias=False))
        self.norm1 = nn.LayerNorm(d_model)
        self.norm2 = nn.LayerNorm(d_model)
        self._packed = None
        self._packed_key = None

    def _params(self):
        return (self.q_proj.weight, self.k_proj.weight, self.v_proj.weight, self.merge.weight, self.mlp[0].weight,
                self.mlp[2].weight, self.norm1.weight, self.norm1.bias, self.norm2.weight, self.norm2.bias)

    def packed(self) -> torch.Tensor:
        ps = self._params()
        key = tuple((p.data_ptr(), p._version, str(p.device)) for p in ps)
        if self._packed is None or key != self._packed_key:
            with _cache_lock:
                if self._packed is None or key != self._packed_key:
                    packed = ops.pack_layer_weights(*ps)
                    torch.cuda.current_stream(packed.device).synchronize()  # other streams may read it right away
                    self._packed, self._packed_key = packed, key
        return self._packed

    def __getstate__(self):  # the packed buffer is a cache: rebuild after unpickling (Ray ships models by pickle)
        st = self.__dict__.copy()
        st["_packed"], st["_packed_key"] = None, None
        return st

    def forward(self, x, source, x_mask=None, source_mask=None):
        """Single layer call x <- layer(x, source): a one-step sequential cross schedule whose second
        half (the update of ``source``) is discarded."""
        d_model = x.shape[-1]
        out, _ = ops.transformer([self.packed()], [1], 1, d_model, self.nhead, x, source, mask0=x_mask, mask1=source_mask,
                                 big_groups=(d_model == 256))
        return out


class _TransformerBase(nn.Module):
    schedule = 0  # 0: OnePosePlus (simultaneous), 1: LoFTR (sequential cross)
    reuse_static_stream0 = True

    def __getstate__(self):
        st = self.__dict__.copy()
        st.pop("_stream0_cache", None)
        return st

    def _build(self, d_model: int, nhead: int, layer_names: Sequence[str], **layer_kw):
        self.d_model, self.nhead = d_model, nhead
        self.layer_names = list(layer_names)
        for n in self.layer_names:
            if n not in ("self", "cross"):
                raise NotImplementedError(f"unknown layer name {n!r}")
        self.layers = nn.ModuleList([LoFTREncoderLayer(d_model, nhead, **layer_kw) for _ in self.layer_names])
        for p in self.parameters():
            if p.dim() > 1:
                nn.init.xavier_uniform_(p)

    def _run(self, x0, x1, mask0=None, mask1=None, add0=None, add1=None, upto: Optional[int] = None,
             planes_only: bool = False):
        layers = list(self.layers)[:upto]
        kinds = [0 if n == "self" else 1 for n in self.layer_names[:upto]]
        prec = getattr(self, "precision", ops.PREC_BF16X3)
        # A first *self* layer sees stream 0 alone.  When stream 0 is one tensor broadcast over the batch (stride 0: the
        # fixed 3D model of a sequence matched against many frames) its output is computed once and reused for every
        # forward until the tensor, the layer's weights or the precision change.
        if (self.reuse_static_stream0 and self.schedule == 0 and kinds and kinds[0] == 0 and mask0 is None
                and x0.dim() == 3 and x0.shape[0] > 1 and x0.stride(0) == 0 and self.d_model == 256
                and tensor_key(x0) is not None and (add0 is None or tensor_key(add0) is not None)):
            key = (tensor_key(x0), prec, None if add0 is None else tensor_key(add0),
                   tuple((p.data_ptr(), p._version) for p in layers[0].parameters()))
            with _cache_lock:
                hit = self.__dict__.get("_stream0_cache")
                if hit is None or hit[0] != key:
                    y0, _ = ops.transformer([layers[0].packed()], [3], 0, self.d_model, self.nhead, x0[:1],
                                            x0.new_zeros(1, 128, self.d_model), add0=add0, big_groups=True, precision=prec)
                    torch.cuda.current_stream(x0.device).synchronize()  # other streams may consume the cached tensor
                    # the entry keeps x0 / add0 alive: while it is current no other tensor can take their addresses, so an
                    # equal (address, version) key means the same contents; the serial tells CUDA graphs of a rebuild
                    hit = (key, y0, x0, add0, (hit[4] + 1) if hit is not None else 1)
                    self.__dict__["_stream0_cache"] = hit
            x0, add0, kinds = hit[1].expand(x0.shape[0], -1, -1), None, [2] + kinds[1:]
        return ops.transformer([l.packed() for l in layers], kinds, self.schedule, self.d_model, self.nhead, x0, x1,
                               add0=add0, add1=add1, mask0=mask0, mask1=mask1, big_groups=(self.d_model == 256),
                               precision=prec, planes_only=planes_only)


# --------------------------------------------------------------------------------------------------
# Fine matching core shared by both variants
# --------------------------------------------------------------------------------------------------
def fine_expectation(query_vec: torch.Tensor, win: torch.Tensor) -> torch.Tensor:
    """query_vec [M, Cf] (strided view allowed), win [M, WW, Cf] -> expec_f [M,3] = (x, y, std)."""
    return ops.fine_match(query_vec, win)
