#!/usr/bin/env python
"""Per-tile timeline of the encoder's tcgen05 GEMM launches (debug): arms opb200_debug_trace, runs ONE coarse
LoFTR stack on bench-sized input and prints, per launch, the mean tile period per CTA and how it splits into
mainloop (MMA start -> accumulator full), epilogue (accumulator full -> done) and their overlap."""
import argparse
import ctypes
import os
import sys

import torch

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
from onepose_st_b200 import _lib, synthetic  # noqa: E402
from onepose_st_b200.onepose_plus import LocalFeatureTransformer  # noqa: E402


def report(t, n, names, launches, mhz=1965.0):
    for i in range(min(launches, n)):
        x = t[i]
        valid = x[:, :, 3] > 0
        if not valid.any():
            break
        tiles = valid.sum(1)
        main_ = (x[:, :, 2] - x[:, :, 1])[valid]
        epi = (x[:, :, 3] - x[:, :, 2])[valid]
        per = []
        for c in range(148):
            k = int(tiles[c])
            if k >= 3:
                per.append((x[c, k - 1, 3] - x[c, 0, 3]) / (k - 1))
        us = lambda c: float(c) / mhz
        print(f"launch {i:2d} {names[i % len(names)]:26s} tiles/CTA {int(tiles.min())}-{int(tiles.max())}  "
              f"period {us(torch.tensor(per).mean()) if per else float('nan'):6.2f} us  mainloop {us(main_.mean()):6.2f} "
              f"(min {us(main_.min()):5.2f})  epilogue {us(epi.mean()):6.2f} (min {us(epi.min()):5.2f})")


def trace_conv(a):
    from onepose_st_b200 import ops
    from onepose_st_b200.nn_common import pack_conv_f16, _cs
    n, s = a.batch, 512
    shapes = [("3x3 128->128 @256", s // 2, 128, 128, 3, 1), ("3x3 196->196 @128", s // 4, 196, 196, 3, 1),
              ("3x3 256->256 @64", s // 8, 256, 256, 3, 1), ("3x3 256->256 @128", s // 4, 256, 256, 3, 1),
              ("3x3 196->196 @256", s // 2, 196, 196, 3, 1), ("3x3 196->128 @256", s // 2, 196, 128, 3, 1),
              ("1x1 128->196 @256", s // 2, 128, 196, 1, 1), ("3x3/2 128->196 @256", s // 2, 128, 196, 3, 2)]
    args = []
    for name, hw, cin, cout, k, stride in shapes:
        x16 = torch.zeros(n, hw, hw, _cs(cin), dtype=torch.float16, device="cuda")
        x16[..., :cin] = torch.randn(n, hw, hw, cin, device="cuda").half()
        wp, bp = pack_conv_f16(torch.randn(cout, cin, k, k, device="cuda") / (cin * k * k) ** 0.5, torch.randn(cout, device="cuda"))
        args.append((x16, wp, k, stride, bp))
    for x16, wp, k, stride, bp in args:
        ops.conv2d_nhwc_f16(x16, wp, k, stride, bp, None, 1)
    torch.cuda.synchronize()
    slot = 148 * 32 * 4
    buf = torch.zeros(16 * slot, dtype=torch.int64, device="cuda")
    lib = _lib.lib()
    cnt = lib.opb200_debug_trace(ctypes.c_void_p(buf.data_ptr()), ctypes.c_longlong(buf.numel() * 8))
    for x16, wp, k, stride, bp in args:
        ops.conv2d_nhwc_f16(x16, wp, k, stride, bp, None, 1)
    torch.cuda.synchronize()
    lib.opb200_debug_trace(None, ctypes.c_longlong(0))
    report(buf.view(-1, 148, 32, 4).cpu().double(), cnt, [s_[0] for s_ in shapes], len(shapes))


def trace_fine(a):
    tr = LocalFeatureTransformer(synthetic.onepose_plus_cfg()["loftr_fine"]).cuda().eval()
    M = a.matches
    f3 = torch.randn(M, 128, 1, device="cuda")
    win = torch.randn(M, 25, 128, device="cuda")
    with torch.no_grad():
        for _ in range(2):
            tr(f3, win)
        torch.cuda.synchronize()
        slot = 148 * 32 * 4
        buf = torch.zeros(16 * slot, dtype=torch.int64, device="cuda")
        lib = _lib.lib()
        n = lib.opb200_debug_trace(ctypes.c_void_p(buf.data_ptr()), ctypes.c_longlong(buf.numel() * 8))
        tr(f3, win)
        torch.cuda.synchronize()
        lib.opb200_debug_trace(None, ctypes.c_longlong(0))
    report(buf.view(-1, 148, 32, 4).cpu().double(), n, ["QKV128", "MERGE+LN128", "MLP1+ReLU128", "MLP2+LN+res128"], 8)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--conv", action="store_true")
    ap.add_argument("--fine", action="store_true")
    ap.add_argument("--matches", type=int, default=5300)
    ap.add_argument("--batch", type=int, default=8)
    ap.add_argument("--points", type=int, default=8000)
    ap.add_argument("--grid", type=int, default=64)
    ap.add_argument("--launches", type=int, default=10)
    a = ap.parse_args()
    torch.manual_seed(0)
    if a.conv:
        return trace_conv(a)
    if a.fine:
        return trace_fine(a)
    tr = LocalFeatureTransformer(synthetic.onepose_plus_cfg()["loftr_coarse"]).cuda().eval()
    d3 = torch.randn(a.batch, 256, a.points, device="cuda")
    d2 = torch.randn(a.batch, a.grid * a.grid, 256, device="cuda")
    with torch.no_grad():
        for _ in range(2):
            tr(d3, d2)
        torch.cuda.synchronize()
        slot = 148 * 32 * 4
        buf = torch.zeros(64 * slot, dtype=torch.int64, device="cuda")
        lib = _lib.lib()
        n = lib.opb200_debug_trace(ctypes.c_void_p(buf.data_ptr()), ctypes.c_longlong(buf.numel() * 8))
        tr(d3, d2)
        torch.cuda.synchronize()
        lib.opb200_debug_trace(None, ctypes.c_longlong(0))
    mhz = 1965.0
    t = buf.view(-1, 148, 32, 4).cpu().double()
    names = ["QKV", "ATTN", "MERGE+LN", "MLP1+ReLU", "MLP2+LN+res"]
    for i in range(min(a.launches, n)):
        x = t[i]
        valid = x[:, :, 3] > 0
        if not valid.any():
            break
        tiles = valid.sum(1)
        t0 = torch.where(valid, x[:, :, 0], torch.full_like(x[:, :, 0], float("inf"))).min()
        end = x[:, :, 3].max()
        main_ = (x[:, :, 2] - x[:, :, 1])[valid]
        epi = (x[:, :, 3] - x[:, :, 2])[valid]
        # steady-state period: (last done - first accumulator full) / (tiles - 1) on CTAs with >= 3 tiles
        per = []
        for c in range(148):
            k = int(tiles[c])
            if k >= 3:
                per.append((x[c, k - 1, 3] - x[c, 0, 3]) / (k - 1))
        first_full = (x[:, 0, 2] - x[:, 0, 0])[valid[:, 0]]
        # gap between one tile's accumulator-full and the next tile's MMA start (negative = MMA ran ahead of the epilogue)
        us = lambda c: float(c) / mhz
        print(f"launch {i:2d} {names[i % 5]:12s} tiles/CTA {int(tiles.min())}-{int(tiles.max())}  span {us(end - t0):7.1f} us  "
              f"period {us(torch.tensor(per).mean()) if per else float('nan'):6.2f} us  mainloop {us(main_.mean()):6.2f} "
              f"(min {us(main_.min()):5.2f})  epilogue {us(epi.mean()):6.2f} (min {us(epi.min()):5.2f})  "
              f"first-tile fill+mainloop {us(first_full.mean()):5.2f} us")


if __name__ == "__main__":
    main()
