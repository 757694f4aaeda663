"""Host-side sharding logic and the final match gather, on CPU (gloo, world_size 2 and 3)."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from onepose_st_b200 import sharding


def test_shard_units_partition():
    for n, w in [(0, 1), (1, 4), (7, 2), (500, 8), (2000, 8), (9, 3)]:
        seen = []
        for r in range(w):
            u = sharding.shard_units(n, r, w)
            assert u == sorted(u)
            assert all(sharding.owner_of(k, w) == (r, i) for i, k in enumerate(u))
            seen += u
        assert sorted(seen) == list(range(n))
    with pytest.raises(ValueError):
        sharding.shard_units(4, 2, 2)


def _unit_rows(k, width):
    """Deterministic ragged payload of unit k (count varies, some units empty)."""
    m = (k * 7 + 3) % 5 if k % 4 else 0
    return torch.arange(m * width, dtype=torch.float32).reshape(m, width) + 1000.0 * k


def test_gather_single_process():
    rr = sharding.RaggedRows(6, "cpu", capacity_rows=2)  # tiny capacity: exercises growth
    for k in range(5):
        rr.append(_unit_rows(k, 6))
    rows, counts = sharding.gather_matches(rr.rows(), rr.counts, 5)
    assert counts.tolist() == [_unit_rows(k, 6).shape[0] for k in range(5)]
    assert torch.equal(rows, torch.cat([_unit_rows(k, 6) for k in range(5)]))
    with pytest.raises(ValueError):
        sharding.gather_matches(rr.rows(), rr.counts[:-1], 5)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n_units, width, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        rr = sharding.RaggedRows(width, "cpu")
        for k in sharding.shard_units(n_units, rank, world):
            rr.append(_unit_rows(k, width))
        rows, counts = sharding.gather_matches(rr.rows(), rr.counts, n_units)
        # plain lists: a tensor in an mp.Queue is shared through an fd that dies with the worker
        q.put((rank, rows.tolist(), counts.tolist()))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,n_units,width", [(2, 9, 6), (2, 1, 5), (3, 10, 5), (2, 0, 6)])
def test_gather_gloo(world, n_units, width):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n_units, width, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    want_rows = torch.cat([_unit_rows(k, width) for k in range(n_units)]) if n_units else torch.empty(0, width)
    want_counts = [_unit_rows(k, width).shape[0] for k in range(n_units)]
    for rank, rows, counts in got:
        assert counts == want_counts, rank
        assert rows == want_rows.tolist(), rank


def test_rows_layout():
    d = {"mkpts_3d_db": torch.rand(4, 3), "mkpts_query_f": torch.rand(4, 2), "mconf": torch.rand(4)}
    r = sharding.rows_onepose(d)
    assert r.shape == (4, 6) and torch.equal(r[:, 3:5], d["mkpts_query_f"]) and torch.equal(r[:, 5], d["mconf"])
    d = {"mkpts0_f": torch.rand(3, 2), "mkpts1_f": torch.rand(3, 2), "mconf": torch.rand(3)}
    r = sharding.rows_loftr(d)
    assert r.shape == (3, 5) and torch.equal(r[:, 2:4], d["mkpts1_f"])
