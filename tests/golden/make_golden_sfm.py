"""Golden vectors of the SfM match post-processing, produced by the UNMODIFIED reference functions
(src/KeypointFreeSfM/coarse_match/utils.py, coarse_match_worker.py).  Build container only:

    python tests/golden/make_golden_sfm.py      ->  tests/golden/sfm_merge.npz
"""
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(0, HERE)
import ref_shims  # noqa: E402
import cases_sfm  # noqa: E402

ref_shims.install()
if "ray" not in sys.modules:  # coarse_match_worker.py decorates its wrappers with ray.remote at import time
    ray = types.ModuleType("ray")
    ray.remote = lambda *a, **k: (a[0] if a and callable(a[0]) else (lambda f: f))
    sys.modules["ray"] = ray
pl = sys.modules["pytorch_lightning"]
if not hasattr(pl, "seed_everything"):
    pl.seed_everything = lambda *a, **k: None

from src.KeypointFreeSfM.coarse_match.utils import Match2Pts2D  # noqa: E402
from src.KeypointFreeSfM.coarse_match.coarse_match_worker import points2D_worker, update_matches, transform_points2D  # noqa: E402


def main():
    matches, pairs, n_images = cases_sfm.case_sfm_merge()
    names = [f"img{i:03d}.png" for i in range(n_images)]
    split = " "
    md = {split.join((names[a], names[b])): m for m, (a, b) in zip(matches, pairs)}
    all_kpts = Match2Pts2D(md, names, name_split=split)
    keypoints = points2D_worker(all_kpts[0:n_images], verbose=False)
    upd = update_matches(md, keypoints, verbose=False, pair_name_split=split)
    kp, sc = transform_points2D({k: v for k, v in keypoints.items() if isinstance(v, dict)}, verbose=False)
    out = {}
    for i, n in enumerate(names):
        out[f"kpts_{i}"] = kp[n]
        out[f"scores_{i}"] = sc[n]
    for p, (a, b) in enumerate(pairs):
        out[f"ids_{p}"] = upd[split.join((names[a], names[b]))].astype(np.int64)
    np.savez_compressed(os.path.join(HERE, "sfm_merge.npz"), **out)
    print("wrote sfm_merge.npz:", sum(len(kp[n]) for n in names), "keypoints,", sum(len(m) for m in matches), "matches")


if __name__ == "__main__":
    main()
