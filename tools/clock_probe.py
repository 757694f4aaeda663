#!/usr/bin/env python
"""SM clock and board power while ONE kernel configuration runs back to back (NVML samples every 20 ms): shows whether a
kernel is paced by the power cap (clock well under the maximum with sw_power_cap set).
usage: clock_probe.py [debug_mode]   -- runs opb200_coarse_match at the bench shape for ~3 s"""
import os, sys, threading, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if len(sys.argv) > 1:
    os.environ["OPB200_SIM_DEBUG"] = sys.argv[1]
import torch, pynvml
from onepose_st_b200 import ops, synthetic
N, L, S, C = 8, 8000, 4096, 256
f3d, f2d = synthetic.planted_coarse_pair(21, N, L, S, C, 1000, 64, 64, device="cuda")
kw = dict(scale=1.0 / (C * 0.0801), thr=0.1, border_rm=2, border_mode=0, grid_a=None, grid_b=(64, 64))
pynvml.nvmlInit()
h = pynvml.nvmlDeviceGetHandleByIndex(0)
samples, stop = [], False
def sampler():
    while not stop:
        samples.append((pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM), pynvml.nvmlDeviceGetPowerUsage(h) / 1e3,
                        pynvml.nvmlDeviceGetCurrentClocksThrottleReasons(h)))
        time.sleep(0.02)
for _ in range(5):
    ops.coarse_match(f3d, f2d, **kw)
torch.cuda.synchronize()
th = threading.Thread(target=sampler); th.start()
t0 = time.time(); n = 0
while time.time() - t0 < 3.0:
    for _ in range(50):
        ops.coarse_match(f3d, f2d, **kw)
    torch.cuda.synchronize(); n += 50
dt = time.time() - t0
stop = True; th.join()
mhz = sorted(s[0] for s in samples[5:]); pw = sorted(s[1] for s in samples[5:])
reasons = 0
for s in samples[5:]: reasons |= s[2]
print(f"debug {os.environ.get('OPB200_SIM_DEBUG', '0')}: {1e3 * dt / n:.3f} ms per call; SM clock median {mhz[len(mhz) // 2]} MHz (min {mhz[0]}, max {mhz[-1]}), "
      f"power median {pw[len(pw) // 2]:.0f} W (max {pw[-1]:.0f}), throttle reasons 0x{reasons:x} (0x4 = sw_power_cap)")
