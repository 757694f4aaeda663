#!/usr/bin/env python
"""Times opb200_coarse_match's kernels at the bench shape (8 frames x 8000 points x 4096 cells) on random features,
with the OPB200_SIM_DEBUG experiment modes of the sweep kernel; CUDA events around the whole call."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from onepose_st_b200 import ops, synthetic
N, L, S, C = 8, 8000, 4096, 256
f3d, f2d = synthetic.planted_coarse_pair(21, N, L, S, C, 1000, 64, 64, device="cuda")
kw = dict(scale=1.0 / (C * 0.0801), thr=0.1, border_rm=2, border_mode=0, grid_a=None, grid_b=(64, 64))
for _ in range(3):
    ops.coarse_match(f3d, f2d, **kw)
torch.cuda.synchronize()
ts = []
for _ in range(10):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); ops.coarse_match(f3d, f2d, **kw); e1.record(); torch.cuda.synchronize()
    ts.append(e0.elapsed_time(e1))
print("debug", os.environ.get("OPB200_SIM_DEBUG", "0"), "coarse_match call ms (min, median):", min(ts), sorted(ts)[5])

if os.environ.get("OPB200_SIM_TRACE"):
    import ctypes
    from onepose_st_b200 import _lib
    slot = 148 * 32 * 4
    buf = torch.zeros(slot, dtype=torch.int64, device="cuda")
    lib = _lib.lib()
    lib.opb200_debug_trace(ctypes.c_void_p(buf.data_ptr()), ctypes.c_longlong(buf.numel() * 8))
    ops.coarse_match(f3d, f2d, **kw)
    torch.cuda.synchronize()
    lib.opb200_debug_trace(ctypes.c_void_p(0), ctypes.c_longlong(0))
    x = buf.view(148, 32, 4).cpu()
    for cta in (0, 77):
        t0 = int(x[cta, 0, 0])
        print(f"CTA {cta}: every 4th tile: MMA issue start, MMA issued, epilogue sees full, epilogue releases (SM clocks from the first stamp)")
        for it in range(30):
            print(f"  {it:2d}: " + " ".join(f"{int(v) - t0:8d}" for v in x[cta, it]))
