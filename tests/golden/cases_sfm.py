"""Seeded input of the SfM match post-processing case (shared by make_golden_sfm.py, the oracle test and the GPU test)."""
import numpy as np


def case_sfm_merge(n_images=7, seed=11):
    """LoFTR-like match rows for a small pair list: coordinates on the 8-pixel coarse grid plus a sub-pixel fine offset,
    many keypoints shared between pairs (so that scores accumulate and ranks tie), every image in at least one pair."""
    rng = np.random.default_rng(seed)
    pairs = [(a, b) for a in range(n_images) for b in range(a + 1, n_images) if (a + b) % 3 != 0]
    pairs += [(4, 1), (6, 2)]  # an image on the second side too, out of order
    pool = {i: rng.integers(2, 60, size=(150, 2)) * 8 for i in range(n_images)}  # recurring coarse cells per image
    matches = []
    for (a, b) in pairs:
        m = int(rng.integers(20, 200))
        ia, ib = rng.integers(0, 150, m), rng.integers(0, 150, m)
        p0 = pool[a][ia] + rng.uniform(-0.4, 0.4, (m, 2)) * (rng.random((m, 1)) < 0.3)
        p1 = pool[b][ib] + rng.uniform(-3.9, 3.9, (m, 2))
        conf = np.round(rng.uniform(0.2, 1.0, m), 2)  # coarse confidences repeat: exact score ties occur
        matches.append(np.concatenate([p0, p1, conf[:, None]], 1).astype(np.float32))
    return matches, pairs, n_images
