"""GPU: forwards issued from two host threads on two CUDA streams (pipeline.StreamPipeline) give bit-identical
matches to the same forwards run one after the other -- per-stream workspaces, no shared scratch."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def test_stream_pipeline_matches_sequential():
    from onepose_st_b200 import pipeline, sharding, synthetic
    from onepose_st_b200.onepose_plus import OnePosePlus_model
    dev = torch.device("cuda")
    torch.manual_seed(0)
    model = OnePosePlus_model(synthetic.onepose_plus_cfg()).eval()
    model.load_state_dict(synthetic.calibrate_onepose_state_dict(model.state_dict()), strict=True)
    model = model.to(dev)
    B, n_batches = 2, 6
    seq = synthetic.make_onepose_sequence(model, 1500, B * n_batches, 256, 256, 300, device=dev)
    kp = seq["keypoints3d"].expand(B, -1, -1).contiguous()
    d_f = seq["descriptors3d_db"].expand(B, -1, -1).contiguous()
    d_c = seq["descriptors3d_coarse_db"].expand(B, -1, -1).contiguous()

    def run(k):
        data = {"keypoints3d": kp, "descriptors3d_db": d_f, "descriptors3d_coarse_db": d_c,
                "query_image": seq["frames"][k * B: (k + 1) * B]}
        model(data)
        return sharding.rows_onepose(data).cpu(), data["b_ids"].cpu()

    with torch.no_grad():
        want = [run(k) for k in range(n_batches)]
        torch.cuda.synchronize()
        for n_streams in (1, 2, 3):
            got = pipeline.StreamPipeline(n_streams).map(run, range(n_batches))
            torch.cuda.synchronize()
            for (r0, b0), (r1, b1) in zip(want, got):
                assert r0.shape[0] > 20
                assert torch.equal(b0, b1) and torch.equal(r0, r1)


def test_stream_pipeline_propagates_errors_and_orders_results():
    from onepose_st_b200 import pipeline
    pipe = pipeline.StreamPipeline(2)
    out = pipe.map(lambda i: (torch.full((4,), float(i), device="cuda") * 2).sum().item(), range(9))
    assert out == [8.0 * i for i in range(9)]
    with pytest.raises(ZeroDivisionError):
        pipe.map(lambda i: 1 // (i - 3), range(6))
    assert pipe.map(lambda i: i, []) == []


def test_broadcast_3d_model_reuses_first_self_layer():
    """A 3D model passed as a broadcast (stride-0) batch is encoded once and its first self-attention layer is
    computed once per sequence; the matches must equal those of the plain batched evaluation."""
    from onepose_st_b200 import sharding, synthetic
    from onepose_st_b200.onepose_plus import OnePosePlus_model
    dev = torch.device("cuda")
    torch.manual_seed(0)
    model = OnePosePlus_model(synthetic.onepose_plus_cfg()).eval()
    model.load_state_dict(synthetic.calibrate_onepose_state_dict(model.state_dict()), strict=True)
    model = model.to(dev)
    B = 3
    seq = synthetic.make_onepose_sequence(model, 1500, B, 256, 256, 300, device=dev)

    def run(broadcast):
        f = (lambda t: t.expand(B, -1, -1)) if broadcast else (lambda t: t.expand(B, -1, -1).contiguous())
        data = {"keypoints3d": f(seq["keypoints3d"]), "descriptors3d_db": f(seq["descriptors3d_db"]),
                "descriptors3d_coarse_db": f(seq["descriptors3d_coarse_db"]), "query_image": seq["frames"][:B]}
        model(data)
        return data

    with torch.no_grad():
        plain = run(False)
        assert "_stream0_cache" not in model.loftr_coarse.__dict__
        cached = run(True)
        assert "_stream0_cache" in model.loftr_coarse.__dict__
        again = run(True)                                   # second forward: served from the cache
    for d in (cached, again):
        assert torch.equal(plain["b_ids"], d["b_ids"]) and torch.equal(plain["i_ids"], d["i_ids"])
        assert torch.equal(plain["j_ids"], d["j_ids"]) and plain["mconf"].numel() > 50
        assert (plain["mconf"] - d["mconf"]).abs().max().item() < 1e-4   # fp32 round trip of the cached stream
        assert (sharding.rows_onepose(plain) - sharding.rows_onepose(d)).abs().max().item() < 1e-3


def test_cuda_graph_head_matches_eager():
    """The forward with its head replayed as a CUDA graph (model.cuda_graph = True) against the eager forward: same
    matches for several frames through the same graph, and after switching to another 3D model and back."""
    from onepose_st_b200 import synthetic
    from onepose_st_b200.onepose_plus import OnePosePlus_model
    dev = torch.device("cuda")
    torch.manual_seed(0)
    model = OnePosePlus_model(synthetic.onepose_plus_cfg()).eval()
    model.load_state_dict(synthetic.calibrate_onepose_state_dict(model.state_dict()), strict=True)
    model = model.to(dev)
    B = 2
    seq_a = synthetic.make_onepose_sequence(model, 1500, 3 * B, 256, 256, 300, device=dev)
    seq_b = synthetic.make_onepose_sequence(model, 1500, B, 256, 256, 300, seed=77, device=dev)

    def run(seq, k):
        f = lambda t: t.expand(B, -1, -1)
        data = {"keypoints3d": f(seq["keypoints3d"]), "descriptors3d_db": f(seq["descriptors3d_db"]),
                "descriptors3d_coarse_db": f(seq["descriptors3d_coarse_db"]), "query_image": seq["frames"][k * B: (k + 1) * B]}
        model(data)
        return {k_: data[k_].clone() for k_ in ("b_ids", "i_ids", "j_ids", "mconf", "mkpts_query_f", "mkpts_3d_db")}

    order = [(seq_a, 0), (seq_a, 1), (seq_b, 0), (seq_a, 2), (seq_a, 0)]
    with torch.no_grad():
        model.cuda_graph = False
        want = [run(s, k) for s, k in order]
        model.cuda_graph = True
        got = [run(s, k) for s, k in order]
        assert len(model._graphs) >= 1   # (graphs of a replaced 3D model are dropped with their workspaces)
    for w, g in zip(want, got):
        assert w["mconf"].numel() > 50
        for k_ in ("b_ids", "i_ids", "j_ids"):
            assert torch.equal(w[k_], g[k_]), k_
        assert (w["mconf"] - g["mconf"]).abs().max().item() < 1e-5
        assert (w["mkpts_query_f"] - g["mkpts_query_f"]).abs().max().item() < 1e-3


def test_loftr_cuda_graph_head_matches_eager():
    """LoFTR_for_OnePose_Plus with the head replayed as a CUDA graph: same matches as the eager forward."""
    from onepose_st_b200 import synthetic
    from onepose_st_b200.loftr import LoFTR_for_OnePose_Plus, default_cfg
    torch.manual_seed(0)
    m = LoFTR_for_OnePose_Plus(default_cfg, enable_fine_matching=True).eval()
    m.load_state_dict(synthetic.calibrate_loftr_state_dict(m.state_dict(), 256, 256))
    m = m.cuda()
    g = torch.Generator().manual_seed(11)
    pairs = []
    for _ in range(3):
        img = torch.rand(2, 1, 256, 256, generator=g).cuda()
        pairs.append((img, (img + 0.01 * torch.randn(img.shape, generator=g).cuda()).clamp(0, 1)))
    one = torch.ones(2, 2, device="cuda")

    def run(i0, i1):
        d = {"image0": i0, "image1": i1, "scale0": one, "scale1": one}
        m(d)
        return {k: d[k].clone() for k in ("b_ids", "i_ids", "j_ids", "mconf", "mkpts0_f", "mkpts1_f")}

    with torch.no_grad():
        m.cuda_graph = False
        want = [run(*p) for p in pairs]
        m.cuda_graph = True
        got = [run(*p) for p in pairs + pairs[:1]]
    assert len(m._graphs) == 1
    for w, g_ in zip(want + want[:1], got):
        assert w["mconf"].numel() > 20
        for k in ("b_ids", "i_ids", "j_ids"):
            assert torch.equal(w[k], g_[k]), k
        assert (w["mconf"] - g_["mconf"]).abs().max().item() < 1e-5
        assert (w["mkpts1_f"] - g_["mkpts1_f"]).abs().max().item() < 1e-3


@pytest.mark.parametrize("graph", [False, True])
def test_new_3d_model_at_a_recycled_address_is_not_served_from_the_caches(graph):
    """The usual evaluation loop `for batch in loader: batch = to_cuda(batch); model(batch)` frees one object's tensors
    and allocates the next object's, of the same shape, at the same addresses with version counter 0.  The per-sequence
    caches (encoded 3D descriptors, first self layer of the 3D stream, CUDA-graph guard) must not mistake the new
    object for the old one: they hold references to the tensors they were built from."""
    from onepose_st_b200 import synthetic
    from onepose_st_b200.onepose_plus import OnePosePlus_model
    dev = torch.device("cuda")
    torch.manual_seed(0)
    model = OnePosePlus_model(synthetic.onepose_plus_cfg()).eval()
    model.load_state_dict(synthetic.calibrate_onepose_state_dict(model.state_dict()), strict=True)
    model = model.to(dev)
    B = 2
    seqs = [synthetic.make_onepose_sequence(model, 1500, B, 256, 256, 300, seed=s, device=dev) for s in (0, 77)]
    host = [{k: seq[k].cpu() for k in ("keypoints3d", "descriptors3d_db", "descriptors3d_coarse_db", "frames")} for seq in seqs]

    def forward(h, cached):
        # fresh device tensors every time, as a data loader produces them
        kp, d_f, d_c = h["keypoints3d"].to(dev), h["descriptors3d_db"].to(dev), h["descriptors3d_coarse_db"].to(dev)
        f = (lambda t: t.expand(B, -1, -1)) if cached else (lambda t: t.expand(B, -1, -1).contiguous())
        data = {"keypoints3d": f(kp), "descriptors3d_db": f(d_f), "descriptors3d_coarse_db": f(d_c),
                "query_image": h["frames"].to(dev)}
        model(data)
        return {k: data[k].clone() for k in ("b_ids", "i_ids", "j_ids", "mconf")}

    with torch.no_grad():
        model.cuda_graph = False
        model.loftr_coarse.reuse_static_stream0 = False
        want = [forward(h, False) for h in host]           # no broadcast, no graph: nothing is cached
        assert not torch.equal(want[0]["i_ids"], want[1]["i_ids"])
        model.loftr_coarse.reuse_static_stream0 = True
        model.cuda_graph = graph
        for rnd in range(3):                                # objects alternate; every forward allocates anew
            for h, w in zip(host, want):
                got = forward(h, True)
                torch.cuda.synchronize()
                assert torch.equal(got["i_ids"], w["i_ids"]) and torch.equal(got["j_ids"], w["j_ids"]), rnd
                assert (got["mconf"] - w["mconf"]).abs().max().item() < 1e-4


def test_inference_mode_inputs_bypass_the_caches():
    """Tensors created under torch.inference_mode() track no version counter: the forward must neither raise nor cache."""
    from onepose_st_b200 import synthetic
    from onepose_st_b200.onepose_plus import OnePosePlus_model
    dev = torch.device("cuda")
    torch.manual_seed(0)
    model = OnePosePlus_model(synthetic.onepose_plus_cfg()).eval()
    model.load_state_dict(synthetic.calibrate_onepose_state_dict(model.state_dict()), strict=True)
    model = model.to(dev)
    B = 2
    seq = synthetic.make_onepose_sequence(model, 1500, B, 256, 256, 300, device=dev)
    keys = ("keypoints3d", "descriptors3d_db", "descriptors3d_coarse_db")
    with torch.no_grad():
        plain = {k: seq[k].expand(B, -1, -1) for k in keys}
        plain["query_image"] = seq["frames"]
        model(plain)
    for graph in (False, True):
        model.cuda_graph = graph
        model._kpt_cache = None
        with torch.inference_mode():
            data = {k: (seq[k] + 0).expand(B, -1, -1) for k in keys}   # (+ 0: new tensors born in inference mode)
            assert data["keypoints3d"].is_inference()
            data["query_image"] = seq["frames"] + 0
            model(data)
            assert model._kpt_cache is None
            assert torch.equal(data["i_ids"], plain["i_ids"]) and torch.equal(data["j_ids"], plain["j_ids"])


def test_head_graphs_release_their_workspaces():
    """Dropping a graph (guard mismatch, LRU eviction, clear) frees the scratch buffers it owned."""
    from onepose_st_b200 import ops
    from onepose_st_b200.nn_common import HeadGraphs
    hg = HeadGraphs(max_entries=2)
    x = torch.ones(4, device="cuda")

    def fn(t):
        ws = ops.workspaces.get(t.device, "probe", 1 << 20)
        return t * 2 + ws[0]

    def scopes():
        return {k[1] for k in ops.workspaces._bufs if isinstance(k[1], tuple) and k[1][0] == "scope"}

    before = scopes()
    for i in range(4):
        hg.run((i,), (x,), fn)
    assert len(hg) == 2 and len(scopes() - before) == 2       # least recently used graphs were dropped with their scratch
    state = {"v": 0}
    hg.run((99,), (x,), fn, guard=lambda: state["v"])
    state["v"] = 1
    hg.run((99,), (x,), fn, guard=lambda: state["v"])        # guard mismatch: recaptured, old scratch released
    assert len(hg) == 2 and len(scopes() - before) == 2
    hg.clear()
    assert len(hg) == 0 and scopes() == before
