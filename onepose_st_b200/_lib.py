"""ctypes loader for libopb200.so (the sm_100a C-ABI library).

The library is built in-tree by ``onepose_st_b200/csrc/Makefile`` (``__graft_entry__.build()``).  There
is no fallback: if the library is missing or the device is not sm_100, every op raises.
"""
import ctypes
import os
import subprocess
import threading

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libopb200.so")
_lock = threading.Lock()
_lib = None


class Opb200Error(RuntimeError):
    pass


def build(verbose: bool = False) -> str:
    """Compile libopb200.so for sm_100a with nvcc (cross-compiles without a GPU)."""
    cmd = ["make", "-C", os.path.join(_HERE, "csrc"), "-j", str(min(8, os.cpu_count() or 1))]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise Opb200Error("building libopb200.so failed:\n" + res.stdout[-4000:] + res.stderr[-4000:])
    if verbose:
        print(res.stdout[-2000:])
    return LIB_PATH


def lib() -> ctypes.CDLL:
    """Return the loaded library; raises (loudly) when it has not been built."""
    global _lib
    if _lib is None:
        with _lock:
            if _lib is None:
                if not os.path.exists(LIB_PATH):
                    raise Opb200Error(
                        f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                        "(onepose_st_b200 has no CPU or PyTorch fallback path)")
                l = ctypes.CDLL(LIB_PATH)
                l.opb200_last_error.restype = ctypes.c_char_p
                l.opb200_version.restype = ctypes.c_int
                _lib = l
    return _lib


def declared_symbols():
    """Every entry point include/opb200.h declares (parsed from the header)."""
    import re
    hdr = os.path.join(os.path.dirname(_HERE), "include", "opb200.h")
    with open(hdr) as f:
        text = f.read()
    return sorted(set(re.findall(r"\b(opb200_[a-z0-9_]+)\s*\(", text)))


def missing_symbols():
    """Declared-but-not-exported symbols of the built library (loads it; launches nothing)."""
    l = lib()
    return [s for s in declared_symbols() if not hasattr(l, s)]


def launch_count() -> int:
    l = lib()
    l.opb200_launch_count.restype = ctypes.c_longlong
    return int(l.opb200_launch_count())


def coarse_kernel_timing(enable: bool) -> None:
    """Arm / disarm the CUDA-event bracket around CoarseMatching's sweep kernel (bench.py roofline)."""
    check(lib().opb200_coarse_kernel_timing(ctypes.c_int(1 if enable else 0)), "opb200_coarse_kernel_timing")


def coarse_kernel_ms():
    """(summed device time in ms, number of launches) of the sweep kernels recorded since the last call."""
    tot, n = ctypes.c_double(0.0), ctypes.c_longlong(0)
    check(lib().opb200_coarse_kernel_ms(ctypes.byref(tot), ctypes.byref(n)), "opb200_coarse_kernel_ms")
    return float(tot.value), int(n.value)


def check(status: int, what: str) -> None:
    if status != 0:
        msg = lib().opb200_last_error()
        raise Opb200Error(f"{what} failed with status {status}: {msg.decode() if msg else ''}")


def ptr(t):
    """Device pointer of a tensor (or NULL)."""
    return ctypes.c_void_p(0 if t is None else t.data_ptr())


def stream_ptr():
    import torch
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def launch_count_add(n: int) -> None:
    """Account for ``n`` kernels of this library launched by a CUDA-graph replay."""
    lib().opb200_launch_count_add(ctypes.c_longlong(int(n)))
