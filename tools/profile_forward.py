#!/usr/bin/env python
"""Kernel-level table of ONE OnePosePlus_model.forward (bench.py's workload, B frames per forward) from the torch
profiler: device time per kernel name, which torch op launched it, and how much of the wall time the GPU sat idle.
Diagnostic only (never a bench number)."""
import argparse
import collections
import os
import sys

import torch
from torch.profiler import ProfilerActivity, profile

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
from onepose_st_b200 import synthetic  # noqa: E402
from onepose_st_b200.onepose_plus import OnePosePlus_model  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=8)
    ap.add_argument("--points", type=int, default=8000)
    ap.add_argument("--size", type=int, default=512)
    ap.add_argument("--top", type=int, default=45)
    a = ap.parse_args()
    dev = torch.device("cuda")
    torch.manual_seed(0)
    model = OnePosePlus_model(synthetic.onepose_plus_cfg()).eval()
    model.load_state_dict(synthetic.calibrate_onepose_state_dict(model.state_dict()), strict=True)
    model = model.to(dev)
    seq = synthetic.make_onepose_sequence(model, a.points, a.batch, a.size, a.size, 1000, device=dev)
    B = a.batch
    kp = seq["keypoints3d"].expand(B, -1, -1)  # one 3D model for every frame of the batch: broadcast views
    d_f = seq["descriptors3d_db"].expand(B, -1, -1)
    d_c = seq["descriptors3d_coarse_db"].expand(B, -1, -1)

    def forward():
        data = {"keypoints3d": kp, "descriptors3d_db": d_f, "descriptors3d_coarse_db": d_c, "query_image": seq["frames"][:B]}
        model(data)
        return data

    for _ in range(3):
        forward()
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
        forward()
        torch.cuda.synchronize()
    by_kernel = collections.defaultdict(lambda: [0.0, 0, ""])
    spans = []
    launched_by = {}
    for e in prof.events():
        if e.device_type == torch.autograd.DeviceType.CPU and e.kernels:
            for k in e.kernels:
                launched_by[k.name[:90]] = e.name
    for e in prof.events():
        if e.device_type == torch.autograd.DeviceType.CUDA:
            dur = e.time_range.end - e.time_range.start
            spans.append((e.time_range.start, e.time_range.end))
            r = by_kernel[e.name[:90]]
            r[0] += dur
            r[1] += 1
            r[2] = launched_by.get(e.name[:90], "C ABI")
    busy = sum(r[0] for r in by_kernel.values())
    if spans:
        t0, t1 = min(s for s, _ in spans), max(t for _, t in spans)
        print(f"device span {(t1 - t0):.0f} us, kernel time {busy:.0f} us ({100 * busy / max(1, t1 - t0):.0f}% busy), "
              f"{sum(r[1] for r in by_kernel.values())} launches, {B} frames")
    for name, (t, n, op) in sorted(by_kernel.items(), key=lambda kv: -kv[1][0])[: a.top]:
        print(f"{t:9.1f} us {n:4d} x  {name:90s} <- {op}")


if __name__ == "__main__":
    main()
