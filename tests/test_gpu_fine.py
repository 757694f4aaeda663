"""GPU parity: fine-level window gather and correlation/softmax-expectation kernels vs golden/oracle."""
import pytest
import torch

import cases
from oracle import ref_ops as R

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("channels_last", [False, True])
def test_fine_gather_vs_golden(golden, channels_last):
    from onepose_st_b200 import ops
    dev = "cuda"
    c = cases.case_fine_preprocess("onepose")
    d = c["data"]
    f = c["feat_f"].to(dev)
    if channels_last:
        f = f.contiguous(memory_format=torch.channels_last)
    win = ops.fine_gather(f, d["b_ids"].to(dev), d["j_ids"].to(dev), 5, 4, 8, 8)
    f3d = ops.gather_rows(c["desc3d_fine"].to(dev), d["b_ids"].to(dev), d["i_ids"].to(dev))
    g = golden("fine_preprocess_onepose")
    assert torch.equal(win.cpu(), g["win"])          # pure data movement: bit exact
    assert torch.equal(f3d.cpu()[:, :, None], g["feat3d"])
    c = cases.case_fine_preprocess("loftr")
    d = c["data"]
    w0 = ops.fine_gather(c["feat_f0"].to(dev), d["b_ids"].to(dev), d["i_ids"].to(dev), 9, 4, 8, 8)
    w1 = ops.fine_gather(c["feat_f1"].to(dev), d["b_ids"].to(dev), d["j_ids"].to(dev), 9, 4, 8, 8)
    g = golden("fine_preprocess_loftr")
    assert torch.equal(w0.cpu(), g["win0"]) and torch.equal(w1.cpu(), g["win1"])


def test_fine_gather_out_of_range_sets_flag():
    from onepose_st_b200 import ops
    dev = "cuda"
    f = torch.randn(1, 128, 32, 32, device=dev)
    flag = torch.zeros(1, dtype=torch.int32, device=dev)
    ops.fine_gather(f, torch.zeros(2, dtype=torch.long, device=dev), torch.tensor([3, 64], device=dev), 5, 4, 8, 8, flag)
    assert flag.item() == 1


def test_fine_match_vs_golden(golden):
    from onepose_st_b200 import ops
    dev = "cuda"
    c = cases.case_fine_matching("onepose")
    e = ops.fine_match(c["feat3d"].to(dev)[:, 0, :], c["win"].to(dev))
    g = golden("fine_matching_onepose")
    err = (e.cpu() - g["expec_f"]).abs().max().item()
    print("fine match onepose expec_f err", err)
    assert err < 2e-5
    c = cases.case_fine_matching("loftr")
    f0 = c["feat_f0"].to(dev)
    e = ops.fine_match(f0[:, 81 // 2, :], c["feat_f1"].to(dev))
    g = golden("fine_matching_loftr")
    err = (e.cpu() - g["expec_f"]).abs().max().item()
    print("fine match loftr expec_f err", err)
    assert err < 2e-5


@pytest.mark.parametrize("W,M", [(5, 2000), (9, 20000)])
def test_fine_match_sweep_vs_oracle(W, M):
    """BASELINE config 5 (fine sweep) shapes, SURVEY 8d planted windows."""
    from onepose_st_b200 import ops
    dev = "cuda"
    WW = W * W
    g = torch.Generator().manual_seed(31)
    win = torch.randn(M, WW, 128, generator=g)
    r = torch.randint(0, WW, (M,), generator=torch.Generator().manual_seed(32))
    q = win[torch.arange(M), r] * 0.3 + 0.05 * torch.randn(M, 128, generator=g)
    ref = R.fine_heatmap_expectation(q, win)
    e = ops.fine_match(q.to(dev), win.to(dev))
    err = (e.cpu() - ref).abs().max().item()
    print(f"fine sweep W={W} M={M}: expec_f max abs err {err:.2e}")
    assert err < 5e-5
    # size-independent property: the expectation is pulled towards the planted cell on both axes
    lin = torch.linspace(-1, 1, W)
    ec = e.cpu()
    assert torch.corrcoef(torch.stack([ec[:, 0], lin[r % W]]))[0, 1].item() > 0.9
    assert torch.corrcoef(torch.stack([ec[:, 1], lin[r // W]]))[0, 1].item() > 0.9
    assert (ec[:, :2].abs() <= 1.0).all() and (ec[:, 2] >= 0).all()
