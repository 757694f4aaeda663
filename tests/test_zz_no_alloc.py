"""opb200_keypoint_encode_static: the entry point that allocates nothing (scratch in a file-scope __device__ array of the
library, csrc/kptenc.cu) against opb200_keypoint_encode on the same inputs, through the C-ABI, on two calls in a row.

Collected last on purpose (file name).  Self-contained: passes alone in a fresh
process and after the other tests (it shares nothing with them but the loaded library).
"""
import ctypes

import pytest
import torch

from onepose_st_b200 import _lib

pytestmark = pytest.mark.gpu

TOL = 2e-5   # the tolerance of test_keypoint_encoding_kernel_vs_golden; the two entry points run the same kernels,
             # so the results are expected to be identical, which is asserted as well


def _i64(v):
    return ctypes.c_int64(int(v))


def _encode(entry, kpts, desc, packed, out, normalize, stats=None):
    l = _lib.lib()
    args = [_lib.ptr(kpts), _i64(kpts.stride(0)), _i64(kpts.stride(1)), _i64(kpts.stride(2)),
            _lib.ptr(desc), _i64(desc.stride(0)), _i64(desc.stride(1)), _i64(desc.stride(2)),
            _lib.ptr(packed), ctypes.c_int(kpts.shape[0]), ctypes.c_int(kpts.shape[1]), ctypes.c_int(int(normalize)), _lib.ptr(out)]
    if stats is not None:
        args.append(_lib.ptr(stats))
    args.append(_lib.stream_ptr())
    return getattr(l, entry)(*args)


def test_keypoint_encode_static_matches_the_workspace_entry_point_twice_in_a_row():
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    l = _lib.lib()
    l.opb200_keypoint_encode_packed_floats.restype = ctypes.c_size_t
    n_packed = int(l.opb200_keypoint_encode_packed_floats())
    max_batch = int(l.opb200_keypoint_encode_static_max_batch())
    assert max_batch == 1024
    g = torch.Generator().manual_seed(20261105)
    packed = (torch.randn(n_packed, generator=g) * 0.2).to(dev)
    # two calls in a row on different shapes: the second (smaller batch, ragged L) runs over the first one's leftovers
    for call, (n, L, normalize) in enumerate([(5, 1000, True), (2, 77, True), (3, 130, False)]):
        kpts = (torch.randn(n, L, 3, generator=g) * torch.tensor([0.3, 0.1, 0.2]) + torch.tensor([0.5, -1.0, 2.0])).to(dev)
        desc = torch.randn(n, 256, L, generator=g).to(dev)
        ref = torch.full((n, 256, L), float("nan"), device=dev)
        got = torch.full((n, 256, L), float("nan"), device=dev)
        stats = torch.empty(4 * n, device=dev)
        torch.cuda.synchronize()
        _lib.check(_encode("opb200_keypoint_encode", kpts, desc, packed, ref, normalize, stats), "opb200_keypoint_encode")
        allocated = torch.cuda.memory_allocated(dev)
        _lib.check(_encode("opb200_keypoint_encode_static", kpts, desc, packed, got, normalize), "opb200_keypoint_encode_static")
        torch.cuda.synchronize()
        assert torch.cuda.memory_allocated(dev) == allocated
        assert torch.isfinite(ref).all() and torch.isfinite(got).all()
        err = (got - ref).abs().max().item()
        print(f"call {call}: n={n} L={L} normalize={normalize}: max abs diff static vs workspace entry point {err:.2e}")
        assert err <= TOL
        assert torch.equal(got, ref)
        # and the result is not trivially the input: the MLP term is there
        assert (got - desc).abs().max().item() > 1e-3


def test_keypoint_encode_static_rejects_batches_beyond_the_static_scratch():
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    l = _lib.lib()
    l.opb200_keypoint_encode_packed_floats.restype = ctypes.c_size_t
    packed = torch.zeros(int(l.opb200_keypoint_encode_packed_floats()), device=dev)
    n = int(l.opb200_keypoint_encode_static_max_batch()) + 1
    kpts = torch.zeros(1, 4, 3, device=dev).expand(n, -1, -1)      # stride-0 batch views: nothing of size n is allocated
    desc = torch.zeros(1, 256, 4, device=dev).expand(n, -1, -1)
    out = torch.zeros(1, device=dev)                                 # never written: the call must fail before any launch
    before = _lib.launch_count()
    st = _encode("opb200_keypoint_encode_static", kpts, desc, packed, out, True)
    assert st == -1
    assert _lib.launch_count() == before
    msg = l.opb200_last_error().decode()
    assert "exceeds the static scratch" in msg and "1024" in msg
    with pytest.raises(_lib.Opb200Error):
        _lib.check(st, "opb200_keypoint_encode_static")
