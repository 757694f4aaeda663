"""Parity report of a product forward against the oracle's forward on the same inputs (TEST INFRASTRUCTURE).

Used as the checker by ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s parity block; never on a product
path.  The quantities are the ones BASELINE.json's north_star names: confidence error (max-abs), mutual-NN index sets
(with every differing pair's distance to the oracle's nearest decision boundary), fine keypoint error.
"""
from __future__ import annotations

from typing import Dict, Optional

import torch


def boundary_margin(conf: torch.Tensor, i: int, j: int, thr: float) -> float:
    """Distance of the oracle confidence conf[i, j] to the nearest decision of the selection rule
    (OPP/utils/coarse_matching.py:145-166): the threshold, the runner-up of its row, the runner-up of its column."""
    c = conf[i, j].item()
    row, col = conf[i].clone(), conf[:, j].clone()
    row[j], col[i] = -1.0, -1.0
    return min(abs(c - thr), abs(c - row.max().item()), abs(c - col.max().item()))


def onepose_frame_report(ref: Dict[str, torch.Tensor], got: Dict[str, torch.Tensor], thr: float,
                         batch_index: Optional[int] = None) -> dict:
    """``ref``: the oracle's ``data`` of ONE frame (N = 1, holds conf_matrix); ``got``: the product's ``data`` (CPU tensors),
    optionally restricted to batch element ``batch_index`` of a batched forward."""
    conf = ref["conf_matrix"][0]
    if batch_index is None:
        sel = torch.ones_like(got["b_ids"], dtype=torch.bool)
    else:
        sel = got["b_ids"] == batch_index
    gi, gj = got["i_ids"][sel].tolist(), got["j_ids"][sel].tolist()
    gconf, gf = got["mconf"][sel], got["mkpts_query_f"][sel]
    ri, rj = ref["i_ids"].tolist(), ref["j_ids"].tolist()
    gmap = {p: k for k, p in enumerate(zip(gi, gj))}
    rmap = {p: k for k, p in enumerate(zip(ri, rj))}
    common = sorted(set(gmap) & set(rmap))
    diff = sorted(set(gmap) ^ set(rmap))
    margins = [boundary_margin(conf, i, j, thr) for (i, j) in diff]
    gk = torch.tensor([gmap[p] for p in common], dtype=torch.long)
    rk = torch.tensor([rmap[p] for p in common], dtype=torch.long)
    # confidence error over every pair either side selected (the product's value where it has one, else it did not
    # select the pair: then the oracle's margin to the decision boundary is what bounds the disagreement)
    conf_err = (gconf[gk] - ref["mconf"][rk]).abs().max().item() if common else 0.0
    got_vs_matrix = max((abs(gconf[gmap[p]].item() - conf[p[0], p[1]].item()) for p in gmap), default=0.0)
    kp_err = (gf[gk] - ref["mkpts_query_f"][rk]).abs().max().item() if common else 0.0
    return {"matches_ref": len(rmap), "matches_got": len(gmap), "common": len(common), "set_diff": len(diff),
            "set_diff_max_margin": max(margins, default=0.0), "mconf_max_abs_err": conf_err,
            "mconf_vs_oracle_matrix_max_abs_err": got_vs_matrix, "mkpts_query_f_max_abs_err_px": kp_err}


def merge_reports(reports) -> dict:
    out = {"frames": len(reports)}
    for k in ("matches_ref", "matches_got", "common", "set_diff"):
        out[k] = sum(r[k] for r in reports)
    for k in ("set_diff_max_margin", "mconf_max_abs_err", "mconf_vs_oracle_matrix_max_abs_err", "mkpts_query_f_max_abs_err_px"):
        out[k] = max((r[k] for r in reports), default=0.0)
    return out
