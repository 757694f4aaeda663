"""CPU restatement (numpy) of the Keypoint-Free SfM match post-processing (TEST INFRASTRUCTURE).

Reference (all under src/KeypointFreeSfM/coarse_match): ``Match2Pts2D`` + ``agg_groupby_2d`` (utils.py:5-60),
``points2D_worker`` / ``update_matches`` / ``transform_points2D`` (coarse_match_worker.py:87-184), driven by
``detector_free_coarse_matching`` (coarse_match.py:155-186).  Input: per image pair the LoFTR match rows
``[M, 5] = (u0, v0, u1, v1, conf)``; output: per image the merged integer keypoints ranked by their summed confidence,
and per pair the matches rewritten as indices into those keypoint lists.

Pinned against the reference's own functions by tests/golden/make_golden_sfm.py -> tests/golden/sfm_merge.npz.
"""
from __future__ import annotations

from typing import Dict, List, Sequence, Tuple

import numpy as np


def merge_pair_matches(matches: Sequence[np.ndarray], pairs: Sequence[Tuple[int, int]], n_images: int):
    """``matches[p]`` [M_p, 5] float, ``pairs[p] = (img0, img1)`` (pair order = insertion order of the reference's dict).
    Returns (keypoints: list of [K_i, 2] float32, scores: list of [K_i] float32, match_ids: list of [M_p, 2] int64)."""
    # Match2Pts2D.__getitem__ (utils.py:37-43): for every image, the (x, y, conf) of each pair it takes part in, in pair order
    per_image: List[List[np.ndarray]] = [[] for _ in range(n_images)]
    for m, (a, b) in zip(matches, pairs):
        per_image[a].append(m[:, [0, 1, 4]])
        per_image[b].append(m[:, [2, 3, 4]])
    kpt_ids: List[Dict[Tuple[int, int], int]] = []
    keypoints, scores = [], []
    for i in range(n_images):
        if not per_image[i]:
            kpt_ids.append({})
            keypoints.append(np.empty((0, 2), np.float32))
            scores.append(np.empty((0,), np.float32))
            continue
        kpts = np.concatenate(per_image[i], 0)
        # points2D_worker (coarse_match_worker.py:100-110) -> agg_groupby_2d(keys.astype(int), conf, 'sum') (utils.py:15-18)
        keys = kpts[:, :2].astype(int)
        uniq, group = np.unique(keys, axis=0, return_inverse=True)
        sums = np.bincount(group.reshape(-1), weights=kpts[:, -1])
        # sorted(..., key=score, reverse=True): stable, ties keep np.unique's lexicographic (x, y) order
        order = sorted(range(len(uniq)), key=lambda k: sums[k], reverse=True)
        kpt_ids.append({tuple(uniq[k]): r for r, k in enumerate(order)})
        # transform_points2D (coarse_match_worker.py:167-181)
        keypoints.append(uniq[order].astype(np.float32))
        scores.append(sums[order].astype(np.float32))
    # update_matches (coarse_match_worker.py:122-158)
    match_ids = []
    for m, (a, b) in zip(matches, pairs):
        k0, k1 = m[:, :2].astype(int), m[:, 2:4].astype(int)
        ids = np.array([[kpt_ids[a][tuple(p0)], kpt_ids[b][tuple(p1)]] for p0, p1 in zip(k0, k1)], dtype=np.int64).reshape(-1, 2)
        match_ids.append(ids)
    return keypoints, scores, match_ids
