"""Probe: does cuDNN's fused conv+bias(+add)+relu path work on this GPU, and what does BN folding +
channels_last buy for ResNetFPN_8_2?  Prints timings and max-abs differences vs the plain module."""
import sys, os, time
import torch
import torch.nn.functional as F
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from onepose_st_b200.nn_common import ResNetFPN_8_2

torch.manual_seed(0)
cfg = {"block_type": "BasicBlock", "initial_dim": 128, "block_dims": [128, 196, 256]}
m = ResNetFPN_8_2(cfg).cuda().eval()
for mod in m.modules():
    if isinstance(mod, torch.nn.BatchNorm2d):
        mod.running_mean.normal_(0, 0.1); mod.running_var.uniform_(0.5, 1.5); mod.weight.data.uniform_(0.5, 1.5); mod.bias.data.normal_(0, 0.1)
x = torch.rand(1, 1, 512, 512, device="cuda")

def timeit(fn, n=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n

with torch.no_grad():
    for tf32 in (False, True):
        torch.backends.cudnn.allow_tf32 = tf32
        ref_c, ref_f = m(x)
        print(f"tf32={tf32} plain NCHW          : {timeit(lambda: m(x)):.3f} ms")
        xl = x.contiguous(memory_format=torch.channels_last)
        ml = m.to(memory_format=torch.channels_last)
        c, f = ml(xl)
        print(f"tf32={tf32} plain channels_last : {timeit(lambda: ml(xl)):.3f} ms  diff {(c-ref_c).abs().max().item():.2e} {(f-ref_f).abs().max().item():.2e}")
        m = m.to(memory_format=torch.contiguous_format)
        try:
            from onepose_st_b200.nn_common import FusedBackbone
            fb = FusedBackbone(m)
            c, f = fb(x)
            print(f"tf32={tf32} fused folded-BN     : {timeit(lambda: fb(x)):.3f} ms  diff {(c-ref_c).abs().max().item():.2e} {(f-ref_f).abs().max().item():.2e}  strides {c.stride()} {f.stride()}")
            g = torch.cuda.CUDAGraph()
            xs = x.clone()
            s = torch.cuda.Stream()
            with torch.cuda.stream(s):
                fb(xs)
                torch.cuda.synchronize()
                with torch.cuda.graph(g):
                    out = fb(xs)
            torch.cuda.synchronize()
            print(f"tf32={tf32} fused + CUDA graph  : {timeit(lambda: g.replay()):.3f} ms  diff {(out[0]-ref_c).abs().max().item():.2e}")
        except Exception as e:
            import traceback; traceback.print_exc()
