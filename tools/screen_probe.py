"""Probe of the screening sweep on the bench workload's real coarse features: operand norms, the rigorous bound, and what
the single fp16 product actually costs in confidence error against the two-pass path and an fp64 dual softmax.
Run on the GPU box:  python tools/screen_probe.py"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from onepose_st_b200 import ops, synthetic  # noqa: E402
from onepose_st_b200.onepose_plus import OnePosePlus_model, normalize_3d_keypoints  # noqa: E402


def main():
    dev = torch.device("cuda")
    torch.manual_seed(0)
    cfg = synthetic.onepose_plus_cfg()
    model = OnePosePlus_model(cfg).eval()
    model.load_state_dict(synthetic.calibrate_onepose_state_dict(model.state_dict()), strict=True)
    model = model.to(dev)
    N, L, S = 8, 8000, 4096
    seq = synthetic.make_onepose_sequence(model, L, N, 512, 512, 1000, device=dev)
    with torch.no_grad():
        fc, ff = model.backbone(seq["frames"][:N])
        q2d = fc.flatten(2).permute(0, 2, 1)
        add2d = model.dense_pos_encoding.token_table(64, 64)
        kp = seq["keypoints3d"].expand(N, -1, -1)
        dc = seq["descriptors3d_coarse_db"].expand(N, -1, -1)
        d3 = model.kpt_3d_pos_encoding(normalize_3d_keypoints(kp[:1]), dc[:1]).expand(N, -1, -1)
        f3d, f2d = model.loftr_coarse(d3, q2d, _add2d=add2d)
    na, nb = f3d.flatten(0, 1).norm(dim=1), f2d.flatten(0, 1).norm(dim=1)
    scale, thr = 1.0 / (256 * 0.0801), 0.1
    c = scale * 1.4426950408889634
    print(json.dumps({"norm_a": [na.min().item(), na.median().item(), na.max().item()],
                      "norm_b": [nb.min().item(), nb.median().item(), nb.max().item()],
                      "abs_a_max": f3d.abs().max().item(), "abs_b_max": f2d.abs().max().item(),
                      "mean_b_norm": f2d[0].mean(0).norm().item(), "centered_b_norm_max": (f2d[0] - f2d[0].mean(0)).norm(dim=1).max().item(),
                      "mean_a_norm": f3d[0].mean(0).norm().item(), "centered_a_norm_max": (f3d[0] - f3d[0].mean(0)).norm(dim=1).max().item(),
                      "delta_wc_octaves": c * 2 ** -10 * na.max().item() * nb.max().item()}))
    kw = dict(scale=scale, thr=thr, border_rm=2, border_mode=0, grid_a=None, grid_b=(64, 64))
    ref = ops.coarse_match(f3d, f2d, algo=1, **kw)
    ref_set = set(zip(ref[0].tolist(), ref[1].tolist(), ref[2].tolist()))
    ref_conf = dict(zip(zip(ref[0].tolist(), ref[1].tolist(), ref[2].tolist()), ref[3].tolist()))
    # fp64 dual softmax of frame 0
    sim = (f3d[0].double() @ f2d[0].double().t()) * scale
    conf64 = (torch.softmax(sim, 0) * torch.softmax(sim, 1))
    # per-element error of the fp16 product, frame 0, a sample of rows
    y16 = (f3d[0].half().float().double() @ f2d[0].half().float().double().t()) * c
    y = sim * 1.4426950408889634
    err = (y16 - y).abs()
    big = conf64 > 0.01
    print(json.dumps({"fp16_product_err_octaves": {"max_all": err.max().item(), "rms_all": err.pow(2).mean().sqrt().item(),
                                                   "max_where_conf_gt_0.01": err[big].max().item(), "n_conf_gt_0.01": int(big.sum())}}))
    for md in (0.25, 0.6, 2.0):
        ops.coarse_screen_max_delta(md)
        info = {}
        out = ops.coarse_match(f3d, f2d, _info=info, **kw)
        diag = ops.coarse_diagnostics(dev, N, L, S, 256, thr)
        got = set(zip(out[0].tolist(), out[1].tolist(), out[2].tolist()))
        gc = dict(zip(zip(out[0].tolist(), out[1].tolist(), out[2].tolist()), out[3].tolist()))
        e2 = max(abs(gc[k] - ref_conf[k]) for k in got & ref_set)
        e64 = max(abs(gc[k] - conf64[k[1], k[2]].item()) for k in got if k[0] == 0)
        diff = got ^ ref_set
        marg = max((abs(ref_conf.get(k, gc.get(k)) - thr) for k in diff), default=0.0)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            h = ops.coarse_match_launch(f3d, f2d, **kw)
        e1.record()
        torch.cuda.synchronize()
        print(json.dumps({"max_delta": md, "rerun": info["rerun"], "M": len(got), "set_diff_vs_two_pass": len(diff),
                          "set_diff_max_dist_to_thr": marg, "mconf_err_vs_two_pass": e2, "mconf_err_vs_fp64_frame0": e64,
                          "norms_candidates": diag, "ms_per_call_launch_only": e0.elapsed_time(e1) / 10}))
    ops.coarse_screen_max_delta(0.6)
    # the sweep kernel by itself (event bracket inside opb200_coarse_match), default kernel vs the register-buffered variant
    from onepose_st_b200 import _lib
    for algo in (0, 3, 2):
        _lib.coarse_kernel_timing(True)
        for _ in range(3):
            ops.coarse_match_launch(f3d, f2d, algo=algo, **kw)
        torch.cuda.synchronize()
        _lib.coarse_kernel_ms()
        for _ in range(20):
            h = ops.coarse_match_launch(f3d, f2d, algo=algo, **kw)
        torch.cuda.synchronize()
        ms, n = _lib.coarse_kernel_ms()
        _lib.coarse_kernel_timing(False)
        out = ops.coarse_match_finish(h)
        diag = ops.coarse_diagnostics(dev, N, L, S, 256, thr)
        fl = 2.0 * N * L * S * 256
        print(json.dumps({"algo": algo, "sweep_kernel_us": 1e3 * ms / max(n, 1), "launches": n, "tflops": fl / (ms / max(n, 1) * 1e-3) / 1e12,
                          "M": int(out[0].shape[0]), "candidates": diag[2], "observed_err": diag[3]}))


if __name__ == "__main__":
    main()
