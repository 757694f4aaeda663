#!/usr/bin/env python
"""Stage-level timing of the hot-path operators on one GPU (CUDA events, L2 flushed between iterations).

Covers the BASELINE.json configs that are stage-level by nature:
  coarse       CoarseMatching on planted features (config 3: 30k points x 64x64; also 5k / 8k)
  transformer  the coarse / fine LoFTR encoder stacks
  fine         FineMatching sweep W in {5, 9}, M in {2k..20k} (config 5), fp32 error vs an fp64 evaluation
Each line is JSON with the op's algorithmic work and the fraction of the measured roofline.
Also the thing to put under ncu: ``ncu -k regex:gemm_kernel ... python tools/stage_bench.py coarse --iters 1``.
"""
import argparse
import json
import math
import os
import sys

import torch

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
from onepose_st_b200 import ops, synthetic  # noqa: E402


def peaks():
    p = os.path.join(REPO, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return d["hbm_gbs"], d["bf16_tflops"], "measured"
    return 6650.0, 1590.0, "fallback"


def time_op(fn, iters, warmup=3):
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    ts = []
    for i in range(iters):
        flush.fill_(i & 0xFF)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2], ts[0]


def run_coarse(a):
    hbm, tf, src = peaks()
    hc = wc = a.grid
    S, C = hc * wc, 256
    for L in a.points:
        f3d, f2d = synthetic.planted_coarse_pair(21, 1, L, S, C, min(3000, L // 2), hc, wc, device="cuda")
        scale = 1.0 / (C * 0.0801)
        for mat in ([False, True] if a.materialize else [False]):
            def fn():
                return ops.coarse_match(f3d, f2d, scale=scale, thr=0.1, border_rm=2, border_mode=0, grid_a=None,
                                        grid_b=(hc, wc), materialize_conf=mat, precision=a.precision, algo=a.algo)
            ops.stage_events = {}
            med, best = time_op(fn, a.iters, warmup=0 if a.iters == 1 else 3)
            evs = ops.stage_events["coarse_match"][-a.iters:]
            ops.stage_events = None
            dev_ms = sorted(x.elapsed_time(y) for x, y in evs)[len(evs) // 2]
            m = fn()[0].numel()
            flops = 2.0 * L * S * C
            line = {"op": "coarse_match", "L": L, "S": S, "C": C, "materialize_conf": mat, "precision": a.precision, "algo": a.algo,
                    "matches": m, "ms_call_incl_host_sync": med, "ms_device": dev_ms,
                    "algorithmic_gflop": flops / 1e9, "tflops": flops / dev_ms / 1e9,
                    "frac_tensor": flops / dev_ms / 1e9 / tf, "peak": f"{tf} TFLOP/s bf16 burst ({src})"}
            if mat:
                by = 4.0 * L * S + 4.0 * C * (L + S)
                line.update({"algorithmic_mb": by / 1e6, "gbs": by / dev_ms / 1e6, "frac_hbm": by / dev_ms / 1e6 / hbm})
            print(json.dumps(line), flush=True)


def run_transformer(a):
    hbm, tf, src = peaks()
    from onepose_st_b200.onepose_plus import LocalFeatureTransformer
    cfgs = synthetic.onepose_plus_cfg()
    torch.manual_seed(0)
    S = a.grid * a.grid
    tr = LocalFeatureTransformer(cfgs["loftr_coarse"]).cuda().eval()
    tr.precision = a.precision
    for L in a.points:
        d3 = torch.randn(1, 256, L, device="cuda")
        d2 = torch.randn(1, S, 256, device="cuda")
        ops.stage_events = {}
        med, best = time_op(lambda: tr(d3, d2), a.iters, warmup=0 if a.iters == 1 else 3)
        evs = ops.stage_events["transformer256"][-a.iters:]
        ops.stage_events = None
        dev_ms = sorted(x.elapsed_time(y) for x, y in evs)[len(evs) // 2]
        flops = 120.0 * (L + S) * 256 * 256
        print(json.dumps({"op": "loftr_coarse (6 layers x 2 streams)", "L": L, "S": S, "precision": a.precision,
                          "ms_device": dev_ms, "algorithmic_gflop": flops / 1e9, "tflops": flops / dev_ms / 1e9,
                          "frac_tensor": flops / dev_ms / 1e9 / tf, "peak": f"{tf} TFLOP/s bf16 burst ({src})"}), flush=True)
    trf = LocalFeatureTransformer(cfgs["loftr_fine"]).cuda().eval()
    trf.precision = a.precision
    for M in a.matches:
        f3 = torch.randn(M, 128, 1, device="cuda")
        win = torch.randn(M, 25, 128, device="cuda")
        ops.stage_events = {}
        time_op(lambda: trf(f3, win), a.iters, warmup=0 if a.iters == 1 else 3)
        evs = ops.stage_events["transformer128"][-a.iters:]
        ops.stage_events = None
        dev_ms = sorted(x.elapsed_time(y) for x, y in evs)[len(evs) // 2]
        flops = 17.0e6 * M
        print(json.dumps({"op": "loftr_fine (W=5)", "M": M, "ms_device": dev_ms, "algorithmic_gflop": flops / 1e9,
                          "tflops": flops / dev_ms / 1e9}), flush=True)


def run_fine(a):
    """BASELINE config 5: fine-level sweep W in {5, 9} x M, fp32 and bf16 arms, error of each against an fp64 evaluation and
    of the bf16 arm against the fp32 one (the "bf16 vs fp32 tolerance report")."""
    hbm, tf, src = peaks()
    for W in (5, 9):
        for M in a.matches:
            q, win, r = synthetic.planted_fine_windows(31, M, W, device="cuda")
            # fp64 evaluation of the same formula (torch, on the device) for the error report
            sim = torch.einsum("mc,mrc->mr", q.double(), win.double()) / math.sqrt(128.0)
            heat = torch.softmax(sim, dim=1)
            lin = (torch.linspace(0, W - 1, W, device="cuda", dtype=torch.float64) / (W - 1) - 0.5) * 2
            gx, gy = lin.repeat(W), lin.repeat_interleave(W)
            ex, ey = heat @ gx, heat @ gy
            vx = (heat @ (gx * gx) - ex * ex).clamp_min(1e-10)
            vy = (heat @ (gy * gy) - ey * ey).clamp_min(1e-10)
            ref = torch.stack([ex, ey, vx.sqrt() + vy.sqrt()], 1)
            del sim, heat
            got32 = None
            for dt, nm, bpe in ((torch.float32, "fp32", 4.0), (torch.bfloat16, "bf16", 2.0)):
                qq, ww = q.to(dt), win.to(dt)
                ops.stage_events = {}
                time_op(lambda: ops.fine_match(qq, ww), a.iters, warmup=0 if a.iters == 1 else 3)
                evs = ops.stage_events["fine_match"][-a.iters:]
                ops.stage_events = None
                dev_ms = sorted(x.elapsed_time(y) for x, y in evs)[len(evs) // 2]
                got = ops.fine_match(qq, ww).double()
                if got32 is None:
                    got32 = got
                by = M * (bpe * 128 * (W * W + 1) + 20)
                print(json.dumps({"op": "fine_match", "dtype": nm, "W": W, "M": M, "ms_device": dev_ms, "algorithmic_mb": by / 1e6,
                                  "gbs": by / dev_ms / 1e6, "frac_hbm": by / dev_ms / 1e6 / hbm,
                                  "max_abs_err_vs_fp64": (got - ref).abs().max().item(),
                                  "max_abs_err_xy_vs_fp64": (got[:, :2] - ref[:, :2]).abs().max().item(),
                                  "max_abs_err_vs_fp32_kernel": (got - got32).abs().max().item(),
                                  "mkpts_err_px_at_scale2": (got[:, :2] - ref[:, :2]).abs().max().item() * (W // 2) * 2,
                                  "peak": f"{hbm} GB/s ({src})"}), flush=True)


def run_loftr(a):
    """BASELINE config 4: LoFTR wsize9 pair matching on 512x512 pairs (identical images + calibrated weights so that
    matches exist), coarse-only and coarse+fine, pairs/s through LoFTR_for_OnePose_Plus.forward."""
    from onepose_st_b200.loftr import LoFTR_for_OnePose_Plus, default_cfg
    torch.manual_seed(0)
    for fine in (False, True):
        m = LoFTR_for_OnePose_Plus(default_cfg, enable_fine_matching=fine).eval()
        m.load_state_dict(synthetic.calibrate_loftr_state_dict(m.state_dict(), a.size, a.size))
        m = m.cuda()
        for nb in a.pair_batch:
            g = torch.Generator().manual_seed(2000)
            img0 = torch.rand(nb, 1, a.size, a.size, generator=g).cuda()
            img1 = img0.clone()
            one = torch.ones(nb, 2, device="cuda")
            out = {}

            def fn():
                d = {"image0": img0, "image1": img1, "scale0": one, "scale1": one}
                m(d)
                out["m"] = int(d["mconf"].shape[0])
            m.cuda_graph = False
            med, best = time_op(fn, a.iters, warmup=0 if a.iters == 1 else 3)
            line = {"op": "LoFTR_for_OnePose_Plus.forward", "fine": fine, "pairs_per_forward": nb, "image": a.size,
                    "matches_per_pair": out["m"] / nb, "ms_per_forward": med, "pairs_per_s": nb / med * 1e3}
            # throughput mode: CUDA-graph head + three host threads x CUDA streams (pipeline.StreamPipeline)
            from onepose_st_b200 import pipeline
            m.cuda_graph = True
            pipe = pipeline.StreamPipeline(3)
            pipe.warm(fn)
            n_fw = 24
            pipe.map(lambda _: fn(), range(6))
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            pipe.map(lambda _: fn(), range(n_fw))
            e1.record()
            torch.cuda.synchronize()
            line["pairs_per_s_pipelined_graph"] = nb * n_fw / e0.elapsed_time(e1) * 1e3
            m.cuda_graph = False
            print(json.dumps(line), flush=True)


def run_conv(a):
    """Every convolution shape of ResNetFPN_8_2 at 512x512 (BASELINE configs 1-4), N images per launch: own tcgen05
    fp16 implicit GEMM vs cuDNN TF32 channels-last (what the reference runs on a GPU)."""
    import torch.nn.functional as F
    from onepose_st_b200.nn_common import pack_conv_f16, _cs
    hbm, tf, src = peaks()
    n = a.pair_batch[-1] if a.pair_batch != [1, 4] else 8
    h2, h4, h8 = a.size // 2, a.size // 4, a.size // 8
    shapes = [("layer1 3x3 128->128", h2, 128, 128, 3, 1, 4), ("layer2.0 3x3/2 128->196", h2, 128, 196, 3, 2, 1),
              ("layer2 3x3 196->196", h4, 196, 196, 3, 1, 3), ("layer3.0 3x3/2 196->256", h4, 196, 256, 3, 2, 1),
              ("layer3 3x3 256->256", h8, 256, 256, 3, 1, 3), ("head2 3x3 256->256", h4, 256, 256, 3, 1, 1),
              ("head2 3x3 256->196", h4, 256, 196, 3, 1, 1), ("head1 3x3 196->196", h2, 196, 196, 3, 1, 1),
              ("head1 3x3 196->128", h2, 196, 128, 3, 1, 1), ("lateral1 1x1 128->196", h2, 128, 196, 1, 1, 1),
              ("lateral2 1x1 196->256", h4, 196, 256, 1, 1, 1), ("lateral3 1x1 256->256", h8, 256, 256, 1, 1, 1)]
    tot_own = tot_cudnn = tot_flop = 0.0
    for name, hw, cin, cout, k, stride, count in shapes:
        x16 = torch.zeros(n, hw, hw, _cs(cin), dtype=torch.float16, device="cuda")
        x16[..., :cin] = torch.randn(n, hw, hw, cin, device="cuda").half()
        w = torch.randn(cout, cin, k, k, device="cuda") / (cin * k * k) ** 0.5
        b = torch.randn(cout, device="cuda")
        wp, bp = pack_conv_f16(w, b)
        ops.stage_events = {}
        time_op(lambda: ops.conv2d_nhwc_f16(x16, wp, k, stride, bp, None, 1), a.iters, warmup=0 if a.iters == 1 else 3)
        evs = ops.stage_events["backbone_conv"][-a.iters:]
        ops.stage_events = None
        own_ms = sorted(x.elapsed_time(y) for x, y in evs)[len(evs) // 2]
        xc = x16[..., :cin].float().permute(0, 3, 1, 2).contiguous(memory_format=torch.channels_last)
        wc = w.contiguous(memory_format=torch.channels_last)
        cu_ms, _ = time_op(lambda: torch.cudnn_convolution_relu(xc, wc, b, (stride, stride), (k // 2, k // 2), (1, 1), 1)
                           if k == 3 else F.conv2d(xc, wc, b, stride), a.iters, warmup=0 if a.iters == 1 else 3)
        ho = (hw + 2 * (k // 2) - k) // stride + 1
        flop = 2.0 * n * ho * ho * cin * cout * k * k
        tot_own += own_ms * count
        tot_cudnn += cu_ms * count
        tot_flop += flop * count
        print(json.dumps({"op": "conv " + name, "images": n, "out_hw": ho, "ms_own_fp16_tcgen05": own_ms,
                          "tflops_own": flop / own_ms / 1e9, "frac_tensor": flop / own_ms / 1e9 / tf,
                          "ms_cudnn_tf32": cu_ms, "tflops_cudnn": flop / cu_ms / 1e9, "per_backbone": count}), flush=True)
    print(json.dumps({"op": "backbone convolutions total", "images": n, "ms_own": tot_own, "ms_cudnn_tf32": tot_cudnn,
                      "algorithmic_gflop": tot_flop / 1e9, "tflops_own": tot_flop / tot_own / 1e9,
                      "frac_tensor": tot_flop / tot_own / 1e9 / tf, "peak": f"{tf} TFLOP/s bf16/fp16 burst ({src})"}), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("what", choices=["coarse", "transformer", "fine", "loftr", "conv"])
    ap.add_argument("--size", type=int, default=512)
    ap.add_argument("--pair-batch", type=int, nargs="+", default=[1, 4])
    ap.add_argument("--points", type=int, nargs="+", default=[5000, 8000, 30000])
    ap.add_argument("--matches", type=int, nargs="+", default=[2000, 5000, 10000, 20000])
    ap.add_argument("--grid", type=int, default=64)
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--precision", type=int, default=3)
    ap.add_argument("--algo", type=int, default=0, help="coarse: 0 = fp16 screening sweep, 1 = two-pass, 2 = single pass x3")
    ap.add_argument("--materialize", action="store_true")
    a = ap.parse_args()
    {"coarse": run_coarse, "transformer": run_transformer, "fine": run_fine, "loftr": run_loftr, "conv": run_conv}[a.what](a)


if __name__ == "__main__":
    main()
