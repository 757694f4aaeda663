"""GPU parity: the tcgen05 encoder stack (C-ABI opb200_transformer) against the golden vectors of the
reference modules and against the oracle at larger sizes."""
import pytest
import torch

import cases
from oracle import ref_ops as R

pytestmark = pytest.mark.gpu

# fp32 contract of BASELINE.json north_star: max-abs 1e-3.  The bf16x3 split path carries the residual stream as
# hi + lo bf16 planes (2^-17 relative per layer); measured worst case 2.1e-4 after 16 layers on O(10) features.
TOL_X3 = 5e-4
# single-pass bf16 products ("fast" mode): 8 mantissa bits per operand, 6-8 stacked layers.
TOL_X1 = 1.5e-1


def pack(sd, prefix, n_layers, dev):
    from onepose_st_b200 import ops
    out = []
    for i in range(n_layers):
        p = f"{prefix}layers.{i}."
        g = lambda k: sd[p + k].to(dev)
        out.append(ops.pack_layer_weights(g("q_proj.weight"), g("k_proj.weight"), g("v_proj.weight"), g("merge.weight"),
                                          g("mlp.0.weight"), g("mlp.2.weight"), g("norm1.weight"), g("norm1.bias"),
                                          g("norm2.weight"), g("norm2.bias")))
    return out


def kinds(names):
    return [0 if n == "self" else 1 for n in names]


def maxerr(a, b):
    return (a.float().cpu() - b.float()).abs().max().item()


@pytest.mark.parametrize("masked", [False, True])
@pytest.mark.parametrize("prec", [3, 1])
def test_onepose_coarse_vs_golden(golden, masked, prec):
    from onepose_st_b200 import ops
    c = cases.case_tr_onepose_coarse(masked)
    dev = "cuda"
    pk = pack(c["sd"], "t.", 6, dev)
    d3 = c["desc3d"].to(dev).permute(0, 2, 1)  # channels-first input, as the reference passes it
    qm = None if c["query_mask"] is None else c["query_mask"].to(dev)
    o3, o2 = ops.transformer(pk, kinds(c["layer_names"]), 0, 256, 8, d3, c["desc2d"].to(dev), mask1=qm, precision=prec)
    g = golden(f"tr_onepose_coarse_{int(masked)}")
    e3, e2 = maxerr(o3, g["desc3d"]), maxerr(o2, g["desc2d"])
    print(f"onepose coarse masked={masked} prec={prec}: max abs err 3d {e3:.2e} 2d {e2:.2e}")
    tol = TOL_X3 if prec == 3 else TOL_X1
    assert e3 < tol and e2 < tol


@pytest.mark.parametrize("masked", [False, True])
def test_loftr_coarse_vs_golden(golden, masked):
    from onepose_st_b200 import ops
    c = cases.case_tr_loftr_coarse(masked)
    dev = "cuda"
    pk = pack(c["sd"], "t.", 8, dev)
    m0 = None if c["mask0"] is None else c["mask0"].to(dev)
    m1 = None if c["mask1"] is None else c["mask1"].to(dev)
    o0, o1 = ops.transformer(pk, kinds(c["layer_names"]), 1, 256, 8, c["feat0"].to(dev), c["feat1"].to(dev), mask0=m0, mask1=m1)
    g = golden(f"tr_loftr_coarse_{int(masked)}")
    e0, e1 = maxerr(o0, g["feat0"]), maxerr(o1, g["feat1"])
    print(f"loftr coarse masked={masked}: max abs err {e0:.2e} {e1:.2e}")
    assert e0 < TOL_X3 and e1 < TOL_X3


def test_fine_vs_golden(golden):
    from onepose_st_b200 import ops
    dev = "cuda"
    c = cases.case_tr_onepose_fine()
    pk = pack(c["sd"], "t.", 2, dev)
    o3, o2 = ops.transformer(pk, kinds(c["layer_names"]), 0, 128, 8, c["desc3d"].to(dev).permute(0, 2, 1), c["desc2d"].to(dev),
                             big_groups=False)
    g = golden("tr_onepose_fine")
    e3, e2 = maxerr(o3, g["desc3d"]), maxerr(o2, g["desc2d"])
    print(f"onepose fine: {e3:.2e} {e2:.2e}")
    assert e3 < TOL_X3 and e2 < TOL_X3
    c = cases.case_tr_loftr_fine()
    pk = pack(c["sd"], "t.", 2, dev)
    o0, o1 = ops.transformer(pk, kinds(c["layer_names"]), 1, 128, 8, c["feat0"].to(dev), c["feat1"].to(dev), big_groups=False)
    g = golden("tr_loftr_fine")
    e0, e1 = maxerr(o0, g["feat0"]), maxerr(o1, g["feat1"])
    print(f"loftr fine: {e0:.2e} {e1:.2e}")
    assert e0 < TOL_X3 and e1 < TOL_X3


@pytest.mark.parametrize("n,L,S", [(1, 5000, 4096), (2, 1000, 1024)])
def test_onepose_coarse_vs_oracle_large(n, L, S):
    """Full-size sequence lengths (BASELINE configs[0] shape): oracle on CPU vs CUDA."""
    from onepose_st_b200 import ops
    dev = "cuda"
    sd = cases.layer_weights(256, 6, 77)
    d3 = cases.randn(78, n, 256, L)
    d2 = cases.randn(79, n, S, 256)
    qm = cases.rand(80, n, S) > 0.1
    names = ["self", "cross"] * 3
    with torch.no_grad():
        r3, r2 = R.transformer_onepose(sd, "t.", d3, d2, names, 8, qm)
    pk = pack(sd, "t.", 6, dev)
    o3, o2 = ops.transformer(pk, kinds(names), 0, 256, 8, d3.to(dev).permute(0, 2, 1), d2.to(dev), mask1=qm.to(dev))
    e3, e2 = maxerr(o3, r3), maxerr(o2, r2)
    print(f"onepose coarse n={n} L={L} S={S}: {e3:.2e} {e2:.2e}")
    assert e3 < TOL_X3 and e2 < TOL_X3


def test_fine_many_matches_vs_oracle():
    from onepose_st_b200 import ops
    dev = "cuda"
    M = 1500
    sd = cases.layer_weights(128, 2, 91)
    f0, f1 = cases.randn(92, M, 81, 128), cases.randn(93, M, 81, 128)
    with torch.no_grad():
        r0, r1 = R.transformer_loftr(sd, "t.", f0, f1, ["self", "cross"], 8)
    pk = pack(sd, "t.", 2, dev)
    o0, o1 = ops.transformer(pk, [0, 1], 1, 128, 8, f0.to(dev), f1.to(dev), big_groups=False)
    e0, e1 = maxerr(o0, r0), maxerr(o1, r1)
    print(f"loftr fine M={M}: {e0:.2e} {e1:.2e}")
    assert e0 < TOL_X3 and e1 < TOL_X3
