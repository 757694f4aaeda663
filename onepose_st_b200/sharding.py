"""Data-parallel sharding of independent units (query frames of a sequence / image pairs of the SfM
pair list) over the GPUs of one box, and the final gather of matches.

The reference shards the same way with Ray tasks over contiguous chunks
(src/inference/inference_OnePosePlus.py:81-100; src/KeypointFreeSfM/coarse_match/coarse_match.py:74-89) and
returns python objects through the Ray object store.  Here: one process per GPU (torchrun), unit k runs on
rank ``k mod G`` (round-robin balances the data-dependent match counts better than contiguous chunks), every
rank keeps its matches on the device in one ragged row buffer, and ONE collective round at the end of the
run (sizes, then rows padded to the largest rank -- the pad-to-largest pattern of src/utils/comm.py:113-138)
brings them together.  There is no exchange step during compute, so there is no other collective.

Row formats: 2D-3D ``[M, 6] = (x3d, y3d, z3d, u, v, conf)``; 2D-2D ``[M, 5] = (u0, v0, u1, v1, conf)`` -- the
reference's own ``(N, 5)`` match rows (coarse_match_worker.py:73-75).
"""
from typing import List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist


def shard_units(n_units: int, rank: int, world_size: int) -> List[int]:
    """Unit ids owned by ``rank``: k with k mod world_size == rank, ascending."""
    if not (0 <= rank < world_size):
        raise ValueError(f"rank {rank} outside world of {world_size}")
    return list(range(rank, n_units, world_size))


def owner_of(unit: int, world_size: int) -> Tuple[int, int]:
    """(rank, local position) of a unit under the round-robin shard."""
    return unit % world_size, unit // world_size


def rows_onepose(data: dict) -> torch.Tensor:
    """[M, 6] fp32 rows (x3d, y3d, z3d, u, v, conf) from an OnePosePlus ``data`` dict after forward."""
    return torch.cat([data["mkpts_3d_db"].float(), data["mkpts_query_f"].float(), data["mconf"].float()[:, None]], dim=1)


def rows_loftr(data: dict) -> torch.Tensor:
    """[M, 5] fp32 rows (u0, v0, u1, v1, conf) from a LoFTR ``data`` dict after forward."""
    return torch.cat([data["mkpts0_f"].float(), data["mkpts1_f"].float(), data["mconf"].float()[:, None]], dim=1)


class RaggedRows:
    """Per-rank device-resident accumulation of match rows, one segment per local unit (in unit order)."""

    def __init__(self, width: int, device, capacity_rows: int = 1 << 16):
        self.width = width
        self.buf = torch.empty(capacity_rows, width, dtype=torch.float32, device=device)
        self.used = 0
        self.counts: List[int] = []

    def append(self, rows: torch.Tensor) -> None:
        m = int(rows.shape[0])
        if rows.dim() != 2 or rows.shape[1] != self.width:
            raise ValueError(f"expected [M, {self.width}] rows, got {tuple(rows.shape)}")
        if self.used + m > self.buf.shape[0]:
            grown = torch.empty(max(2 * self.buf.shape[0], self.used + m), self.width, dtype=torch.float32,
                                device=self.buf.device)
            grown[: self.used] = self.buf[: self.used]
            self.buf = grown
        self.buf[self.used: self.used + m] = rows
        self.used += m
        self.counts.append(m)

    def rows(self) -> torch.Tensor:
        return self.buf[: self.used]

    def reset(self) -> None:
        self.used = 0
        self.counts = []


def _all_gather(out: torch.Tensor, inp: torch.Tensor, group) -> None:
    """out [world, ...] <- every rank's ``inp``; NCCL takes the flat tensor, gloo a list of views."""
    if out.device.type == "cuda":
        dist.all_gather_into_tensor(out, inp.contiguous(), group=group)
    else:
        dist.all_gather(list(out.unbind(0)), inp.contiguous(), group=group)


def gather_matches(rows: torch.Tensor, counts: Sequence[int], n_units: int, group: Optional[dist.ProcessGroup] = None
                   ) -> Tuple[torch.Tensor, torch.Tensor]:
    """All ranks contribute their ragged rows; every rank receives ``(rows_all [sum M, width],
    counts_all [n_units] int64)`` in ascending UNIT order (unit k's rows are
    ``rows_all[offsets[k]: offsets[k] + counts_all[k]]`` with ``offsets = cumsum(counts_all) - counts_all``).

    ``rows`` holds this rank's segments back to back in local unit order (``shard_units`` order) and
    ``counts`` their lengths.  Two collectives: an all_gather of the per-unit counts (padded to the largest
    shard) and an all_gather of the rows padded to the largest rank.  Without an initialised process group
    (single process) it degenerates to a local reorder."""
    dev = rows.device
    width = rows.shape[1]
    if not (dist.is_available() and dist.is_initialized()):
        world, rank = 1, 0
    else:
        world, rank = dist.get_world_size(group), dist.get_rank(group)
    my_units = shard_units(n_units, rank, world)
    if len(counts) != len(my_units):
        raise ValueError(f"rank {rank} owns {len(my_units)} units but reported {len(counts)} counts")
    if int(sum(counts)) != rows.shape[0]:
        raise ValueError("counts do not add up to the number of rows")
    max_units = (n_units + world - 1) // world
    cnt = torch.zeros(max_units, dtype=torch.int64, device=dev)
    if len(counts):
        cnt[: len(counts)] = torch.as_tensor(list(counts), dtype=torch.int64, device=dev)
    if world == 1:
        return rows.clone(), cnt[:n_units]
    cnt_all = torch.empty(world, max_units, dtype=torch.int64, device=dev)
    _all_gather(cnt_all, cnt, group)
    per_rank = cnt_all.sum(dim=1)
    max_rows = int(per_rank.max().item())  # host sync 1: how far to pad
    padded = torch.zeros(max_rows, width, dtype=torch.float32, device=dev)
    padded[: rows.shape[0]] = rows
    rows_all = torch.empty(world, max_rows, width, dtype=torch.float32, device=dev)
    _all_gather(rows_all, padded, group)
    # unit id of every real row, rank-major (the order the rows sit in rows_all once the padding is dropped):
    # unit of (rank r, local position k) is k * world + r
    unit_ids = (torch.arange(max_units, device=dev)[None, :] * world + torch.arange(world, device=dev)[:, None]).reshape(-1)
    unit_of_row = torch.repeat_interleave(unit_ids, cnt_all.reshape(-1))  # host sync 2 (output size)
    real = torch.arange(max_rows, device=dev)[None, :] < per_rank[:, None]
    rows_rank_major = rows_all[real]  # [sum M, width]; same length as unit_of_row
    order = torch.sort(unit_of_row, stable=True).indices
    counts_units = cnt_all.t().reshape(-1)[:n_units]  # [max_units, world] flattened = unit order
    return rows_rank_major[order], counts_units
