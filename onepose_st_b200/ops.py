"""Thin Python wrappers over the C-ABI in include/opb200.h: they only extract pointers, sizes and
strides from torch tensors, own the workspaces, and raise on any non-zero status.  No math here."""
import ctypes
import threading
from typing import List, Optional, Sequence

import torch

from . import _lib

PREC_BF16 = 1
PREC_BF16X3 = 3

_c_i64 = ctypes.c_int64
_c_sz = ctypes.c_size_t


def _need_cuda(*ts):
    for t in ts:
        if t is not None and not t.is_cuda:
            raise _lib.Opb200Error("onepose_st_b200 ops need CUDA tensors (there is no CPU path)")


def _f32(t: torch.Tensor) -> torch.Tensor:
    return t if t.dtype == torch.float32 else t.float()


class _Workspaces:
    """One growable byte buffer per (device, stream, tag); reused across calls so addresses stay stable
    (needed for CUDA-graph capture) and no allocation happens in steady state."""

    def __init__(self):
        self._bufs = {}
        self._tl = threading.local()

    def scope(self, name):
        """Context manager: workspaces requested inside belong to ``name`` instead of to the current stream.  A CUDA
        graph capture uses it so that every graph owns its scratch (torch captures all graphs on one shared side
        stream, and graphs replayed concurrently on different streams must not share workspaces)."""
        ws = self

        class _Scope:
            def __enter__(self_):
                self_.prev = getattr(ws._tl, "scope", None)
                ws._tl.scope = name

            def __exit__(self_, *exc):
                ws._tl.scope = self_.prev
                return False
        return _Scope()

    def drop_scope(self, name) -> None:
        for k in [k for k in self._bufs if k[1] == ("scope", name)]:
            del self._bufs[k]

    def get(self, device, tag: str, nbytes: int) -> torch.Tensor:
        # one buffer per (device, stream | scope, tag): forwards in flight on different streams never share scratch
        sc = getattr(self._tl, "scope", None)
        owner = ("scope", sc) if sc is not None else torch.cuda.current_stream(device).cuda_stream
        key = (device.index if device.index is not None else torch.cuda.current_device(), owner, tag)
        buf = self._bufs.get(key)
        if buf is None or buf.numel() < nbytes:
            buf = torch.empty(int(nbytes), dtype=torch.uint8, device=device)
            self._bufs[key] = buf
        return buf


workspaces = _Workspaces()


class _Stage:
    """Optional CUDA-event bracket around one C-ABI call.  ``stage_events`` is None in normal use (zero
    cost); bench.py sets it to a dict and reads per-op device durations after a synchronize.  Events are
    recorded on the stream the kernels are launched on (torch's current stream)."""

    def __init__(self, name):
        self.name = name

    def __enter__(self):
        if stage_events is not None:
            self.e0 = torch.cuda.Event(enable_timing=True)
            self.e1 = torch.cuda.Event(enable_timing=True)
            self.e0.record()
        return self

    def __exit__(self, *a):
        if stage_events is not None:
            self.e1.record()
            stage_events.setdefault(self.name, []).append((self.e0, self.e1))
        return False


stage_events = None


def split_planes(x: torch.Tensor):
    """fp32 [rows, cols] (any strides) -> (hi, lo) bf16 planes."""
    _need_cuda(x)
    x = _f32(x)
    rows, cols = x.shape
    hi = torch.empty(rows, cols, dtype=torch.bfloat16, device=x.device)
    lo = torch.empty_like(hi)
    st = _lib.lib().opb200_split_planes(_lib.ptr(x), _c_i64(x.stride(0)), _c_i64(x.stride(1)), rows, cols,
                                        _lib.ptr(hi), _lib.ptr(lo), None, _lib.stream_ptr())
    _lib.check(st, "opb200_split_planes")
    return hi, lo


def packed_layer_bytes(d_model: int) -> int:
    l = _lib.lib()
    l.opb200_packed_layer_bytes.restype = _c_sz
    return int(l.opb200_packed_layer_bytes(d_model))


def pack_layer_weights(wq, wk, wv, wm, w1, w2, ln1_w, ln1_b, ln2_w, ln2_b) -> torch.Tensor:
    """Pack one LoFTREncoderLayer's parameters into the bf16 hi/lo plane layout the kernels read."""
    ts = [_f32(t.detach()).contiguous() for t in (wq, wk, wv, wm, w1, w2, ln1_w, ln1_b, ln2_w, ln2_b)]
    _need_cuda(*ts)
    d_model = ts[0].shape[0]
    out = torch.empty(packed_layer_bytes(d_model), dtype=torch.uint8, device=ts[0].device)
    st = _lib.lib().opb200_pack_layer_weights(*[_lib.ptr(t) for t in ts], d_model, _lib.ptr(out), _lib.stream_ptr())
    _lib.check(st, "opb200_pack_layer_weights")
    return out


class FeaturePlanes:
    """Final token features of an encoder stack left in its workspace as bf16 hi/lo planes (x = hi + lo), the
    operand format of ``coarse_match``: passing them on skips the fp32 copy-out and the re-splitting."""

    def __init__(self, ptrs, padded, n, len0, len1, c, ws):
        self.ptrs, self.padded, self.n, self.len0, self.len1, self.c = ptrs, padded, n, len0, len1, c
        self.ws = ws  # keeps the workspace alive; valid until the next transformer call of the same width


def transformer(packed_layers: Sequence[torch.Tensor], layer_kinds: Sequence[int], schedule: int, d_model: int,
                nhead: int, x0: torch.Tensor, x1: torch.Tensor, add0: Optional[torch.Tensor] = None,
                add1: Optional[torch.Tensor] = None, mask0: Optional[torch.Tensor] = None,
                mask1: Optional[torch.Tensor] = None, big_groups: bool = True, precision: int = PREC_BF16X3,
                planes_only: bool = False):
    """x0 [N, len0, C], x1 [N, len1, C] as (possibly permuted) fp32 views -> (out0, out1) contiguous, or with
    ``planes_only`` a ``FeaturePlanes`` handle (no fp32 outputs are written)."""
    _need_cuda(x0, x1)
    x0, x1 = _f32(x0), _f32(x1)
    n, len0, c = x0.shape
    n1, len1, c1 = x1.shape
    if n != n1 or c != d_model or c1 != d_model:
        raise _lib.Opb200Error(f"transformer: inconsistent shapes {tuple(x0.shape)} / {tuple(x1.shape)} for d_model {d_model}")
    l = _lib.lib()
    l.opb200_transformer_workspace_bytes.restype = _c_sz
    nbytes = int(l.opb200_transformer_workspace_bytes(d_model, n, len0, len1, int(big_groups)))
    ws = workspaces.get(x0.device, "transformer%d" % d_model, nbytes)
    out0 = None if planes_only else torch.empty(n, len0, c, dtype=torch.float32, device=x0.device)
    out1 = None if planes_only else torch.empty(n, len1, c, dtype=torch.float32, device=x0.device)
    nl = len(packed_layers)
    w_arr = (ctypes.c_void_p * nl)(*[t.data_ptr() for t in packed_layers])
    k_arr = (ctypes.c_int * nl)(*[int(k) for k in layer_kinds])

    def m8(m, length):
        if m is None:
            return None
        m = m.reshape(n, length)
        return (m if m.dtype in (torch.bool, torch.uint8) else (m != 0)).contiguous()

    mk0, mk1 = m8(mask0, len0), m8(mask1, len1)
    a0 = None if add0 is None else _f32(add0).contiguous()
    a1 = None if add1 is None else _f32(add1).contiguous()
    with _Stage("transformer%d" % d_model):
        st = l.opb200_transformer(w_arr, k_arr, nl, int(schedule), d_model, nhead, n, len0, len1, int(big_groups),
                              _lib.ptr(x0), _c_i64(x0.stride(0)), _c_i64(x0.stride(1)), _c_i64(x0.stride(2)), _lib.ptr(a0),
                              _lib.ptr(x1), _c_i64(x1.stride(0)), _c_i64(x1.stride(1)), _c_i64(x1.stride(2)), _lib.ptr(a1),
                              _lib.ptr(mk0), _lib.ptr(mk1), _lib.ptr(out0), _lib.ptr(out1), _lib.ptr(ws), _c_sz(ws.numel()),
                              int(precision), _lib.stream_ptr())
    _lib.check(st, "opb200_transformer")
    if planes_only:
        ptrs = (ctypes.c_void_p * 4)()
        padded = (ctypes.c_int * 2)()
        st = l.opb200_transformer_planes(_lib.ptr(ws), d_model, n, len0, len1, int(big_groups), ptrs, padded)
        _lib.check(st, "opb200_transformer_planes")
        return FeaturePlanes([int(p or 0) for p in ptrs], (int(padded[0]), int(padded[1])), n, len0, len1, c, ws)
    return out0, out1


class CoarseLaunch:
    """Everything ``coarse_match_launch`` enqueued and allocated; ``coarse_match_finish`` turns it into results."""

    def __init__(self, outs, cnt, conf, cap, retry):
        self.outs, self.cnt, self.conf, self.cap, self.retry = outs, cnt, conf, cap, retry
        self.extra = []  # status words beyond the first three, filled by coarse_match_finish


def coarse_match_launch(feat_a: torch.Tensor, feat_b: torch.Tensor, *, scale: float, thr: float, border_rm: int,
                        border_mode: int, grid_a, grid_b, mask_a: Optional[torch.Tensor] = None,
                        mask_b: Optional[torch.Tensor] = None, mask_fill: int = 0,
                        valid_hw: Optional[torch.Tensor] = None, materialize_conf: bool = False,
                        precision: int = PREC_BF16X3, algo: int = 0, planes: Optional[FeaturePlanes] = None,
                        status: Optional[torch.Tensor] = None) -> CoarseLaunch:
    """Enqueues the dual-softmax + mutual-NN kernels on the current stream; no host synchronisation (capturable in
    a CUDA graph).  With ``planes`` (from ``transformer(..., planes_only=True)``) feat_a / feat_b are ignored.
    ``status`` (optional, zeroed int32 [>= 3]): where the call leaves (match count, overflow, rerun code); words beyond the
    third belong to the caller (the matchers keep the backbone's fp16 range flag there) and reach the host with the same
    copy, as ``CoarseLaunch.extra`` after ``coarse_match_finish``."""
    if planes is not None:
        n, L, S, C = planes.n, planes.len0, planes.len1, planes.c
        if planes.padded != ((L + 127) // 128 * 128, (S + 127) // 128 * 128):
            raise _lib.Opb200Error("coarse_match: feature planes are not padded to 128-row groups (coarse-level stack only)")
        dev = planes.ws.device
    else:
        _need_cuda(feat_a, feat_b)
        feat_a, feat_b = _f32(feat_a).contiguous(), _f32(feat_b).contiguous()
        n, L, C = feat_a.shape
        S = feat_b.shape[1]
        dev = feat_a.device
    if not (0.0 < thr < 1.0):
        raise NotImplementedError(f"coarse matching threshold must be in (0,1), got {thr}")
    l = _lib.lib()
    l.opb200_coarse_match_workspace_bytes.restype = _c_sz
    nbytes = int(l.opb200_coarse_match_workspace_bytes(n, L, S, C, ctypes.c_float(thr)))
    ws = workspaces.get(dev, "coarse", nbytes)
    cap = n * min(L, S)
    b_ids = torch.empty(cap, dtype=torch.int64, device=dev)
    i_ids = torch.empty(cap, dtype=torch.int64, device=dev)
    j_ids = torch.empty(cap, dtype=torch.int64, device=dev)
    mconf = torch.empty(cap, dtype=torch.float32, device=dev)
    cnt = status if status is not None else torch.zeros(3, dtype=torch.int32, device=dev)
    conf = torch.empty(n, L, S, dtype=torch.float32, device=dev) if materialize_conf else None

    def m8(m, length):
        if m is None:
            return None
        m = m.reshape(n, length)
        return (m if m.dtype in (torch.bool, torch.uint8) else (m != 0)).contiguous()

    ma, mb = m8(mask_a, L), m8(mask_b, S)
    vhw = None if valid_hw is None else valid_hw.to(device=dev, dtype=torch.int32).contiguous()
    h0, w0 = (int(grid_a[0]), int(grid_a[1])) if grid_a is not None else (1, L)
    h1, w1 = int(grid_b[0]), int(grid_b[1])
    with _Stage("coarse_match"):
        if planes is not None:
            head = [ctypes.c_void_p(p) for p in planes.ptrs]
            fn = l.opb200_coarse_match_planes
        else:
            head = [_lib.ptr(feat_a), _lib.ptr(feat_b)]
            fn = l.opb200_coarse_match
        st = fn(*head, _lib.ptr(ma), _lib.ptr(mb), int(mask_fill), n, L, S, C,
                               ctypes.c_float(scale), ctypes.c_float(thr), int(border_rm), int(border_mode), h0, w0, h1, w1,
                               _lib.ptr(vhw), _lib.ptr(b_ids), _lib.ptr(i_ids), _lib.ptr(j_ids), _lib.ptr(mconf), cap,
                               _lib.ptr(cnt), _lib.ptr(conf), _lib.ptr(ws), _c_sz(ws.numel()), int(precision), int(algo),
                               _lib.stream_ptr())
    _lib.check(st, "opb200_coarse_match")
    retry = dict(feat_a=feat_a, feat_b=feat_b, scale=scale, thr=thr, border_rm=border_rm, border_mode=border_mode,
                 grid_a=grid_a, grid_b=grid_b, mask_a=mask_a, mask_b=mask_b, mask_fill=mask_fill, valid_hw=valid_hw,
                 materialize_conf=materialize_conf, precision=precision, planes=planes)
    return CoarseLaunch((b_ids, i_ids, j_ids, mconf), cnt, conf, cap, retry)


def coarse_match_finish(h: CoarseLaunch, copy: bool = False):
    """The one host sync of the matcher (reading the match count; the reference syncs at the same point through
    torch.where).  Returns (b_ids, i_ids, j_ids, mconf, conf_matrix|None); ``copy`` detaches the results from the
    launch's buffers (needed when those are the static buffers of a CUDA graph)."""
    words = h.cnt.tolist()
    (m, overflow, rerun), h.extra = words[:3], words[3:]
    h.rerun = rerun  # 0: the launched algorithm's result stands; 1 / 2: it asked for the two-pass / split-product repeat
    if rerun:  # 1: the single-pass statistics hit their dynamic-range guard -> two-pass path;
        kw = dict(h.retry)  # 2: features outside the fp16 screening range -> single pass on the three split products
        a, b = kw.pop("feat_a"), kw.pop("feat_b")
        return coarse_match(a, b, algo=1 if rerun == 1 else 2, **kw)
    b_ids, i_ids, j_ids, mconf = h.outs
    if overflow:
        import os
        if os.environ.get("OPB200_DEBUG_OVERFLOW"):
            print("DEBUG overflow: count", m, "cap", h.cap)
            m = min(m, h.cap)
            return b_ids[:m], i_ids[:m], j_ids[:m], mconf[:m], h.conf
        raise _lib.Opb200Error("coarse_match: candidate/match buffer overflow (internal capacity bound violated)")
    if copy:
        return b_ids[:m].clone(), i_ids[:m].clone(), j_ids[:m].clone(), mconf[:m].clone(), h.conf
    return b_ids[:m], i_ids[:m], j_ids[:m], mconf[:m], h.conf


def coarse_match(feat_a: torch.Tensor, feat_b: torch.Tensor, _info: Optional[dict] = None, **kw):
    """Dual-softmax + mutual-NN selection: ``coarse_match_launch`` + ``coarse_match_finish``.
    Returns (b_ids, i_ids, j_ids, mconf, conf_matrix|None).  ``_info`` (tests): receives {'rerun': code}."""
    h = coarse_match_launch(feat_a, feat_b, **kw)
    out = coarse_match_finish(h)
    if _info is not None:
        _info["rerun"] = h.rerun
    return out


def _err_flag(dev) -> torch.Tensor:
    return torch.zeros(1, dtype=torch.int32, device=dev)


def fine_gather(feat: torch.Tensor, b_ids: torch.Tensor, cell_ids: torch.Tensor, W: int, stride: int, hc: int, wc: int,
                err_flag: Optional[torch.Tensor] = None, dtype: torch.dtype = torch.float32) -> torch.Tensor:
    """feat [N,Cf,Hf,Wf] (any memory format) -> windows [M, W*W, Cf] of the coarse cells cell_ids, as fp32 or (``dtype``
    = torch.bfloat16) bf16 windows (half the bytes; BASELINE.json configs[4])."""
    _need_cuda(feat, b_ids, cell_ids)
    if dtype not in (torch.float32, torch.bfloat16):
        raise _lib.Opb200Error("fine_gather: dtype must be torch.float32 or torch.bfloat16")
    feat = _f32(feat)
    n, cf, hf, wf = feat.shape
    m = int(b_ids.shape[0])
    out = torch.empty(m, W * W, cf, dtype=dtype, device=feat.device)
    if m == 0:
        return out
    flag = err_flag if err_flag is not None else _err_flag(feat.device)
    b_ids, cell_ids = b_ids.long().contiguous(), cell_ids.long().contiguous()
    with _Stage("fine_gather"):
        fn = _lib.lib().opb200_fine_gather if dtype == torch.float32 else _lib.lib().opb200_fine_gather_bf16
        st = fn(_lib.ptr(feat), _c_i64(feat.stride(0)), _c_i64(feat.stride(1)), _c_i64(feat.stride(2)),
                _c_i64(feat.stride(3)), n, cf, hf, wf, _lib.ptr(b_ids), _lib.ptr(cell_ids), m, int(W),
                int(stride), int(hc), int(wc), _lib.ptr(out), _lib.ptr(flag), _lib.stream_ptr())
    _lib.check(st, "opb200_fine_gather")
    return out


def gather_rows(src_ncl: torch.Tensor, b_ids: torch.Tensor, i_ids: torch.Tensor,
                err_flag: Optional[torch.Tensor] = None) -> torch.Tensor:
    """src [N, C, L] -> [M, C] rows src[b, :, i]."""
    _need_cuda(src_ncl, b_ids, i_ids)
    src = _f32(src_ncl)
    n, c, L = src.shape
    m = int(b_ids.shape[0])
    out = torch.empty(m, c, dtype=torch.float32, device=src.device)
    if m == 0:
        return out
    flag = err_flag if err_flag is not None else _err_flag(src.device)
    b_ids, i_ids = b_ids.long().contiguous(), i_ids.long().contiguous()
    st = _lib.lib().opb200_gather_rows(_lib.ptr(src), _c_i64(src.stride(0)), _c_i64(src.stride(1)), _c_i64(src.stride(2)), n, c, L,
                                       _lib.ptr(b_ids), _lib.ptr(i_ids), m, _lib.ptr(out), _lib.ptr(flag), _lib.stream_ptr())
    _lib.check(st, "opb200_gather_rows")
    return out


def fine_match(q: torch.Tensor, win: torch.Tensor) -> torch.Tensor:
    """q [M, Cf] (row stride free), win [M, WW, Cf] contiguous -> expec_f [M, 3].  bf16 windows (``fine_gather(...,
    dtype=torch.bfloat16)``) take the bf16 arm: q is rounded to bf16 too, products and sums stay fp32."""
    _need_cuda(q, win)
    if win.dtype == torch.bfloat16:
        q, win = q.to(torch.bfloat16), win.contiguous()
        m, ww, cf = win.shape
        w = int(round(ww ** 0.5))
        if q.stride(1) != 1:
            q = q.contiguous()
        out = torch.empty(m, 3, dtype=torch.float32, device=win.device)
        if m == 0:
            return out
        with _Stage("fine_match"):
            st = _lib.lib().opb200_fine_match_bf16(_lib.ptr(q), _c_i64(q.stride(0)), _lib.ptr(win), m, w, cf, _lib.ptr(out),
                                                   _lib.stream_ptr())
        _lib.check(st, "opb200_fine_match_bf16")
        return out
    q, win = _f32(q), _f32(win).contiguous()
    m, ww, cf = win.shape
    w = int(round(ww ** 0.5))
    if q.stride(1) != 1:
        q = q.contiguous()
    out = torch.empty(m, 3, dtype=torch.float32, device=win.device)
    if m == 0:
        return out
    with _Stage("fine_match"):
        st = _lib.lib().opb200_fine_match(_lib.ptr(q), _c_i64(q.stride(0)), _lib.ptr(win), m, w, cf, _lib.ptr(out), _lib.stream_ptr())
    _lib.check(st, "opb200_fine_match")
    return out


def upsample2x_add_(dst: torch.Tensor, src: torch.Tensor) -> torch.Tensor:
    """dst [N,C,2H,2W] += bilinear x2 (align_corners=True) of src [N,C,H,W]; both channels-last fp32; in place."""
    _need_cuda(dst, src)
    n, c, hs, ws = src.shape
    if tuple(dst.shape) != (n, c, 2 * hs, 2 * ws):
        raise _lib.Opb200Error(f"upsample2x_add_: shapes {tuple(src.shape)} -> {tuple(dst.shape)}")
    if not (src.is_contiguous(memory_format=torch.channels_last) and dst.is_contiguous(memory_format=torch.channels_last)
            and src.dtype == torch.float32 and dst.dtype == torch.float32):
        raise _lib.Opb200Error("upsample2x_add_: channels-last fp32 tensors required")
    with _Stage("backbone_glue"):
        st = _lib.lib().opb200_upsample2x_add_nhwc(_lib.ptr(src), n, hs, ws, c, _lib.ptr(dst), _lib.stream_ptr())
    _lib.check(st, "opb200_upsample2x_add_nhwc")
    return dst


def stem_conv_relu(image: torch.Tensor, weight_tap_major: torch.Tensor, bias: torch.Tensor) -> torch.Tensor:
    """image [N,1,H,W] fp32 -> relu(conv7x7/2(image) + bias) as a channels-last [N,Cout,Ho,Wo] tensor."""
    _need_cuda(image, weight_tap_major, bias)
    n, cin, h, w = image.shape
    if cin != 1:
        raise _lib.Opb200Error("stem_conv_relu: single-channel images only")
    image = _f32(image).contiguous()
    cout = weight_tap_major.shape[1]
    ho, wo = (h - 1) // 2 + 1, (w - 1) // 2 + 1
    out = torch.empty(n, cout, ho, wo, dtype=torch.float32, device=image.device, memory_format=torch.channels_last)
    with _Stage("backbone_glue"):
        st = _lib.lib().opb200_stem_conv7x7_s2_relu_nhwc(_lib.ptr(image), _lib.ptr(weight_tap_major), _lib.ptr(bias), n, h, w, cout,
                                                         _lib.ptr(out), _lib.stream_ptr())
    _lib.check(st, "opb200_stem_conv7x7_s2_relu_nhwc")
    return out


def bias_leaky_relu_(x: torch.Tensor, bias: torch.Tensor, slope: float) -> torch.Tensor:
    """x [N,C,H,W] channels-last fp32 <- leaky_relu(x + bias[c]); in place."""
    _need_cuda(x, bias)
    n, c, h, w = x.shape
    if not (x.is_contiguous(memory_format=torch.channels_last) and x.dtype == torch.float32):
        raise _lib.Opb200Error("bias_leaky_relu_: channels-last fp32 tensor required")
    bias = _f32(bias).contiguous()
    with _Stage("backbone_glue"):
        st = _lib.lib().opb200_bias_leaky_relu_nhwc(_lib.ptr(x), _lib.ptr(bias), ctypes.c_longlong(n * h * w), c,
                                                    ctypes.c_float(slope), _lib.stream_ptr())
    _lib.check(st, "opb200_bias_leaky_relu_nhwc")
    return x


def linear_attention(q, k, v, q_mask=None, kv_mask=None, eps: float = 1e-6) -> torch.Tensor:
    """Standalone elu+1 linear attention: q [N,L,H,D], k/v [N,S,H,D] -> [N,L,H,D]."""
    _need_cuda(q, k, v)
    q, k, v = _f32(q).contiguous(), _f32(k).contiguous(), _f32(v).contiguous()
    n, L, H, D = q.shape
    S = k.shape[1]
    l = _lib.lib()
    l.opb200_linear_attention_workspace_bytes.restype = _c_sz
    ws = workspaces.get(q.device, "linattn", int(l.opb200_linear_attention_workspace_bytes(n, S, H, D)))
    out = torch.empty_like(q)

    def m8(m, length):
        if m is None:
            return None
        m = m.reshape(n, length)
        return (m if m.dtype in (torch.bool, torch.uint8) else (m != 0)).contiguous()

    qm, km = m8(q_mask, L), m8(kv_mask, S)
    st = l.opb200_linear_attention(_lib.ptr(q), _lib.ptr(k), _lib.ptr(v), _lib.ptr(qm), _lib.ptr(km), n, L, S, H, D,
                                   ctypes.c_float(eps), _lib.ptr(out), _lib.ptr(ws), _c_sz(ws.numel()), _lib.stream_ptr())
    _lib.check(st, "opb200_linear_attention")
    return out


# ------------------------------------------------------------------------------------------------
# fp16 NHWC backbone (tcgen05 implicit-GEMM convolutions).  Tensors here are plain contiguous
# [N, H, W, Cs] fp16 with Cs = channels padded to a multiple of 16 (pad channels are zero).
# ------------------------------------------------------------------------------------------------
def conv2d_nhwc_f16(x: torch.Tensor, w: torch.Tensor, ksize: int, stride: int = 1, bias: Optional[torch.Tensor] = None,
                    res: Optional[torch.Tensor] = None, act: int = 0, slope: float = 0.0, out_f16: bool = True,
                    out_f32: bool = False, stage: str = "backbone_conv", up: Optional[torch.Tensor] = None,
                    range_flag: Optional[torch.Tensor] = None, split_scale: Optional[float] = None):
    """y = act(conv_k(x) + bias + res): x [N,Hi,Wi,Cin_s] fp16, w [k*k*Cout_s, Cin_s] fp16 tap-major, bias [Cout_s] fp32,
    res [N,Ho,Wo,Cout_s] fp16, up [N,Ho/2,Wo/2,Cout_s] fp16 (bilinear x2, align_corners=True, added before act).  Returns the
    fp16 output, the fp32 output, or (fp16, fp32) when both are asked.  ``range_flag`` (int32 [1], optional): bit 0 is set
    by the kernel when a value leaves the fp16 range on its way into the fp16 output.
    ``split_scale`` (float, optional) selects the split-precision kernel (opb200_conv2d_nhwc_f16x3): every fp16 tensor then
    carries [hi | lo] planes along its last axis (x [N,Hi,Wi,2*Cin_s], w [k*k*Cout_s, 2*Cin_s], res / up / fp16 output
    [.., 2*Cout_s]) and ``split_scale`` is the power of two the packed weights were multiplied by."""
    _need_cuda(x, w, bias, res)
    if x.dtype != torch.float16 or w.dtype != torch.float16 or not x.is_contiguous() or not w.is_contiguous() or x.dim() != 4:
        raise _lib.Opb200Error("conv2d_nhwc_f16: contiguous fp16 [N,H,W,C] input and fp16 weights required")
    n, hi, wi, cin_s = x.shape
    planes = 1 if split_scale is None else 2
    if cin_s % planes:
        raise _lib.Opb200Error("conv2d_nhwc_f16: split input needs an even channel count ([hi | lo])")
    cin_s //= planes
    if w.dim() != 2 or w.shape[1] != planes * cin_s or w.shape[0] % (ksize * ksize):
        raise _lib.Opb200Error(f"conv2d_nhwc_f16: weight shape {tuple(w.shape)} does not fit k={ksize}, Cin_s={cin_s}")
    cout_s = w.shape[0] // (ksize * ksize)
    pad = ksize // 2
    ho, wo = (hi + 2 * pad - ksize) // stride + 1, (wi + 2 * pad - ksize) // stride + 1
    if res is not None and (res.dtype != torch.float16 or tuple(res.shape) != (n, ho, wo, planes * cout_s) or not res.is_contiguous()):
        raise _lib.Opb200Error("conv2d_nhwc_f16: residual must be contiguous fp16 [N,Ho,Wo,Cout_s]")
    if up is not None and (up.dtype != torch.float16 or not up.is_contiguous() or ho % 2 or wo % 2
                           or tuple(up.shape) != (n, ho // 2, wo // 2, planes * cout_s)):
        raise _lib.Opb200Error("conv2d_nhwc_f16: up must be contiguous fp16 [N,Ho/2,Wo/2,Cout_s]")
    if bias is not None and (bias.dtype != torch.float32 or bias.numel() != cout_s):
        raise _lib.Opb200Error("conv2d_nhwc_f16: bias must be fp32 [Cout_s]")
    y16 = torch.empty(n, ho, wo, planes * cout_s, dtype=torch.float16, device=x.device) if out_f16 else None
    y32 = torch.empty(n, ho, wo, cout_s, dtype=torch.float32, device=x.device) if out_f32 else None
    with _Stage(stage):
        if split_scale is None:
            st = _lib.lib().opb200_conv2d_nhwc_f16(_lib.ptr(x), n, hi, wi, cin_s, _lib.ptr(w), ksize, stride, cout_s, _lib.ptr(bias),
                                                   _lib.ptr(res), _lib.ptr(up), _lib.ptr(y16), _lib.ptr(y32), act,
                                                   ctypes.c_float(slope), _lib.ptr(range_flag), _lib.stream_ptr())
        else:
            st = _lib.lib().opb200_conv2d_nhwc_f16x3(_lib.ptr(x), n, hi, wi, cin_s, _lib.ptr(w), ksize, stride, cout_s,
                                                     _lib.ptr(bias), _lib.ptr(res), _lib.ptr(up), _lib.ptr(y16), _lib.ptr(y32), act,
                                                     ctypes.c_float(slope), ctypes.c_float(1.0 / split_scale),
                                                     _lib.ptr(range_flag), _lib.stream_ptr())
    _lib.check(st, "opb200_conv2d_nhwc_f16")
    if out_f16 and out_f32:
        return y16, y32
    return y16 if out_f16 else y32


def stem_conv_relu_f16(image: torch.Tensor, weight_tap_major: torch.Tensor, bias: torch.Tensor,
                       range_flag: Optional[torch.Tensor] = None) -> torch.Tensor:
    """image [N,1,H,W] fp32 -> relu(conv7x7/2(image) + bias) as [N,Ho,Wo,Cout] fp16."""
    _need_cuda(image, weight_tap_major, bias)
    n, cin, h, w = image.shape
    if cin != 1:
        raise _lib.Opb200Error("stem_conv_relu_f16: single-channel images only")
    image = _f32(image).contiguous()
    cout = weight_tap_major.shape[1]
    ho, wo = (h - 1) // 2 + 1, (w - 1) // 2 + 1
    out = torch.empty(n, ho, wo, cout, dtype=torch.float16, device=image.device)
    with _Stage("backbone_glue"):
        st = _lib.lib().opb200_stem_conv7x7_s2_relu_nhwc_f16(_lib.ptr(image), _lib.ptr(weight_tap_major), _lib.ptr(bias), n, h, w,
                                                             cout, _lib.ptr(out), _lib.ptr(range_flag), _lib.stream_ptr())
    _lib.check(st, "opb200_stem_conv7x7_s2_relu_nhwc_f16")
    return out


def upsample2x_add_f16_(dst: torch.Tensor, src: torch.Tensor) -> torch.Tensor:
    """dst [N,2H,2W,C] fp16 += bilinear x2 (align_corners=True) of src [N,H,W,C] fp16; in place."""
    _need_cuda(dst, src)
    n, hs, ws, c = src.shape
    if tuple(dst.shape) != (n, 2 * hs, 2 * ws, c) or dst.dtype != torch.float16 or src.dtype != torch.float16 \
            or not dst.is_contiguous() or not src.is_contiguous():
        raise _lib.Opb200Error(f"upsample2x_add_f16_: contiguous fp16 NHWC {tuple(src.shape)} -> {tuple(dst.shape)} required")
    with _Stage("backbone_glue"):
        st = _lib.lib().opb200_upsample2x_add_nhwc_f16(_lib.ptr(src), n, hs, ws, c, _lib.ptr(dst), _lib.stream_ptr())
    _lib.check(st, "opb200_upsample2x_add_nhwc_f16")
    return dst


def coarse_screen_max_delta(max_delta: float = -1.0) -> float:
    """Set (>= 0) or query (< 0) the bound on the fp16 similarity error up to which the default CoarseMatching sweeps ONE
    fp16 product (include/opb200.h: opb200_coarse_screen_max_delta).  Returns the previous value."""
    l = _lib.lib()
    l.opb200_coarse_screen_max_delta.restype = ctypes.c_float
    return float(l.opb200_coarse_screen_max_delta(ctypes.c_float(max_delta)))


def coarse_diagnostics(device, n: int, L: int, S: int, C: int, thr: float):
    """(largest row norm of side a, of side b, candidates captured, largest measured |exact - screening| similarity in
    octaves) of the last default-path coarse_match call on the current stream with these sizes (synchronises)."""
    l = _lib.lib()
    l.opb200_coarse_match_workspace_bytes.restype = _c_sz
    nbytes = int(l.opb200_coarse_match_workspace_bytes(n, L, S, C, ctypes.c_float(thr)))
    ws = workspaces.get(device, "coarse", nbytes)
    na, nb, nc, ob = ctypes.c_float(0), ctypes.c_float(0), ctypes.c_int(0), ctypes.c_float(0)
    st = l.opb200_coarse_norms(_lib.ptr(ws), n, L, S, C, ctypes.c_float(thr), ctypes.byref(na), ctypes.byref(nb),
                               ctypes.byref(nc), ctypes.byref(ob), _lib.stream_ptr())
    _lib.check(st, "opb200_coarse_norms")
    return float(na.value), float(nb.value), int(nc.value), float(ob.value)


def keypoint_encode(kpts: torch.Tensor, desc: torch.Tensor, packed_w: torch.Tensor, normalize: bool) -> torch.Tensor:
    """kpts [N, L, 3], desc [N, 256, L] (views allowed), packed_w from KeypointEncoding_linear._packed() -> [N, 256, L]
    = desc + MLP(normalised kpts)^T (include/opb200.h: opb200_keypoint_encode)."""
    _need_cuda(kpts, desc, packed_w)
    kpts, desc = _f32(kpts), _f32(desc)
    n, L, _ = kpts.shape
    l = _lib.lib()
    l.opb200_keypoint_encode_packed_floats.restype = _c_sz
    if packed_w.numel() != int(l.opb200_keypoint_encode_packed_floats()) or packed_w.dtype != torch.float32:
        raise _lib.Opb200Error("keypoint_encode: packed weights have the wrong size")
    out = torch.empty(n, 256, L, dtype=torch.float32, device=kpts.device)
    stats = torch.empty(4 * n, dtype=torch.float32, device=kpts.device)
    with _Stage("keypoint_encode"):
        st = l.opb200_keypoint_encode(_lib.ptr(kpts), _c_i64(kpts.stride(0)), _c_i64(kpts.stride(1)), _c_i64(kpts.stride(2)),
                                      _lib.ptr(desc), _c_i64(desc.stride(0)), _c_i64(desc.stride(1)), _c_i64(desc.stride(2)),
                                      _lib.ptr(packed_w), n, L, int(bool(normalize)), _lib.ptr(out), _lib.ptr(stats),
                                      _lib.stream_ptr())
    _lib.check(st, "opb200_keypoint_encode")
    return out


def sample_features(feature_map: torch.Tensor, kpts: torch.Tensor, img_h: float, img_w: float, bilinear: bool,
                    normalize: bool = False) -> torch.Tensor:
    """feature_map [C, H, W] (any strides) sampled at kpts [M, 2] (x, y) given in an image of size (img_h, img_w):
    F.grid_sample(align_corners=True, zero padding) of the reference's sample_feature_from_featuremap -> [M, C]."""
    _need_cuda(feature_map, kpts)
    fm = _f32(feature_map)
    k = _f32(kpts).contiguous()
    c, h, w = fm.shape
    m = int(k.shape[0])
    out = torch.empty(m, c, dtype=torch.float32, device=fm.device)
    if m == 0:
        return out
    with _Stage("sample_features"):
        st = _lib.lib().opb200_sample_features(_lib.ptr(fm), _c_i64(fm.stride(0)), _c_i64(fm.stride(1)), _c_i64(fm.stride(2)), c, h, w,
                                               _lib.ptr(k), m, ctypes.c_float(float(img_h)), ctypes.c_float(float(img_w)),
                                               int(bool(bilinear)), int(bool(normalize)), _lib.ptr(out), _lib.stream_ptr())
    _lib.check(st, "opb200_sample_features")
    return out


def track_feature_mean(rows: torch.Tensor, track_start: torch.Tensor) -> torch.Tensor:
    """rows [R, C] fp32 grouped by track, track_start [T + 1] int64 -> [T, C]: per-track mean with the arithmetic of
    np.mean(axis=0) on float32 rows (sequential fp32 sum in row order, then one division)."""
    _need_cuda(rows, track_start)
    rows = _f32(rows).contiguous()
    ts = track_start.long().contiguous()
    t = int(ts.numel()) - 1
    out = torch.empty(max(t, 0), rows.shape[1], dtype=torch.float32, device=rows.device)
    if t <= 0:
        return out
    st = _lib.lib().opb200_track_feature_mean(_lib.ptr(rows), _lib.ptr(ts), t, int(rows.shape[1]), _lib.ptr(out), _lib.stream_ptr())
    _lib.check(st, "opb200_track_feature_mean")
    return out


def split_f16_planes(x_nhwc: torch.Tensor, range_flag: Optional[torch.Tensor] = None) -> torch.Tensor:
    """fp32 [N,H,W,C] contiguous -> fp16 [N,H,W,2C] = [hi | lo] planes (input format of the split-precision convolutions)."""
    _need_cuda(x_nhwc)
    if x_nhwc.dtype != torch.float32 or not x_nhwc.is_contiguous() or x_nhwc.dim() != 4 or x_nhwc.shape[3] % 8:
        raise _lib.Opb200Error("split_f16_planes: contiguous fp32 [N,H,W,C] with C % 8 == 0 required")
    n, h, w, c = x_nhwc.shape
    out = torch.empty(n, h, w, 2 * c, dtype=torch.float16, device=x_nhwc.device)
    with _Stage("backbone_glue"):
        st = _lib.lib().opb200_split_f16_planes(_lib.ptr(x_nhwc), ctypes.c_longlong(n * h * w), c, _lib.ptr(out), _lib.ptr(range_flag),
                                                _lib.stream_ptr())
    _lib.check(st, "opb200_split_f16_planes")
    return out
