#!/usr/bin/env python
"""Times opb200_coarse_match at the bench shape (8 frames x 8000 points x 4096 cells) on planted random features.

OPB200_SIM_DEBUG=1|2|3|4 selects an experiment mode of the sweep kernel (1: epilogue warps only hand the accumulators
back = MMA + TMA alone; 2: no MMAs are issued = epilogue alone; 3: TMEM reads only; 4: arithmetic without TMEM reads --
results are garbage in these modes, run under `ncu -k regex:sim_fast` for the kernel time).  OPB200_SIM_TRACE=1 prints
the per-tile SM-clock stamps of two CTAs (opb200_debug_trace).  OPB200_SIM_DEBUG_COUNT=1 prints the number of captured
candidates per call."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from onepose_st_b200 import _lib as _lib0
if os.environ.get("OPB200_LIB_ALT"):  # A/B on one box: another build of the library
    _lib0.LIB_PATH = os.path.abspath(os.environ["OPB200_LIB_ALT"])
from onepose_st_b200 import ops, synthetic
N, L, S, C = int(os.environ.get("OPB200_N", "8")), 8000, 4096, 256
f3d, f2d = synthetic.planted_coarse_pair(21, N, L, S, C, 1000, 64, 64, device="cuda")
kw = dict(scale=1.0 / (C * 0.0801), thr=0.1, border_rm=2, border_mode=0, grid_a=None, grid_b=(64, 64),
          algo=int(os.environ.get("OPB200_ALGO", "0")))  # 3 = the register-buffered variant (sim_sweep_kernel)
for _ in range(3):
    ops.coarse_match(f3d, f2d, **kw)
torch.cuda.synchronize()
ts = []
for _ in range(10):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); ops.coarse_match(f3d, f2d, **kw); e1.record(); torch.cuda.synchronize()
    ts.append(e0.elapsed_time(e1))
print("debug", os.environ.get("OPB200_SIM_DEBUG", "0"), "coarse_match call ms (min, median):", min(ts), sorted(ts)[5])
from onepose_st_b200 import _lib
_lib.coarse_kernel_timing(True)  # event bracket around the sweep kernel alone (inside opb200_coarse_match)
for _ in range(20):
    ops.coarse_match(f3d, f2d, **kw)
torch.cuda.synchronize()
_lib.coarse_kernel_timing(False)
k_ms, k_n = _lib.coarse_kernel_ms()
print("debug", os.environ.get("OPB200_SIM_DEBUG", "0"), "algo", kw["algo"], "sweep kernel us (mean of %d back-to-back calls):" % k_n, 1e3 * k_ms / max(k_n, 1),
      "-> TFLOP/s", 2.0 * N * L * S * C / (k_ms / max(k_n, 1)) / 1e9)

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
        print(f"CTA {cta}: every 4th tile: MMA issue start, MMA issued, an epilogue warp sees full, last release of the stage (SM clocks from the first stamp)")
        for it in range(32):
            if os.environ.get("OPB200_SIM_DEBUG") == "8":  # one warp's profile: first TMEM read landed, chunk loop done, tail done, end (clocks since entry)
                print(f"  {it:2d}: " + " ".join(f"{int(v):6d}" for v in x[cta, it]))
            elif os.environ.get("OPB200_SIM_DEBUG") == "5":  # per quarter: epilogue clocks of that warp, flags (1 = second sweep, 2 = re-swept)
                print(f"  {it:2d}: " + " ".join(f"{int(v) & ((1 << 48) - 1):6d}/{int(v) >> 48}" for v in x[cta, it]))
            else:
                print(f"  {it:2d}: " + " ".join(f"{int(v) - t0:8d}" for v in x[cta, it]))
