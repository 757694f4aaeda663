"""Mainloop throughput probe: opb200_debug_gemm with the store-free epilogue (precision | 16)."""
import ctypes, os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from onepose_st_b200 import _lib
l = _lib.lib()
def run(M, N, K, bn, prec, nostore=True, iters=20):
    a = torch.randn(M, K, device="cuda").bfloat16(); b = torch.randn(N, K, device="cuda").bfloat16()
    a2 = torch.randn(M, K, device="cuda").bfloat16(); b2 = torch.randn(N, K, device="cuda").bfloat16()
    c = torch.empty(M, N, device="cuda")
    def go():
        st = l.opb200_debug_gemm(_lib.ptr(a), _lib.ptr(a2), _lib.ptr(b), _lib.ptr(b2), M, N, K, bn, prec | (16 if nostore else 0), _lib.ptr(c), _lib.stream_ptr())
        _lib.check(st, "gemm")
    for _ in range(3): go()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters): go()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    fl = 2.0 * M * N * K * (3 if prec == 3 else 1)
    print(f"M={M} N={N} K={K} bn={bn} prec={prec} nostore={nostore}: {ms*1e3:.1f} us  executed {fl/ms/1e9:.0f} TFLOP/s", flush=True)
for M in (12288, 49152, 148 * 128 * 4):
    for (N, K) in ((256, 256), (256, 512), (768, 256), (512, 512)):
        for bn in (256, 128):
            run(M, N, K, bn, 3)
    run(M, 256, 512, 256, 1)
    run(M, 256, 512, 256, 3, nostore=False)
