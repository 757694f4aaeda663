"""First-frame object detection by matching, batched.

The reference's ``LocalFeatureObjectDetector.match_worker`` (src/local_feature_object_detector/local_feature_2D_detector.py:
89-144) runs one LoFTR forward per reference view (~15, sequentially, batch 1) and reads the matches back after each.  The
drop-in matcher is written for N >= 1, so here all reference views that share the query's size go through ONE forward
(image0 = the stacked reference views, image1 = the query repeated) and the matches are split by ``m_bids``; the per-view
results are what the reference's loop produces (same kernels, per-element independence of the batch)."""
from typing import Dict, List, Sequence

import torch


@torch.no_grad()
def match_query_to_references(matcher, query: torch.Tensor, references: Sequence[torch.Tensor]) -> List[Dict[str, torch.Tensor]]:
    """``query`` [1,1,H,W], ``references`` list of [1,1,Hr,Wr] on the matcher's device.  Returns, per reference view, the dict
    {'mkpts0_f' [M,2] (reference view), 'mkpts1_f' [M,2] (query), 'mconf' [M]} of that pair.  Views of equal size are matched
    in one batched forward."""
    out: List[Dict[str, torch.Tensor]] = [None] * len(references)
    by_shape: Dict[tuple, List[int]] = {}
    for k, r in enumerate(references):
        by_shape.setdefault(tuple(r.shape[-2:]), []).append(k)
    for shape, idx in by_shape.items():
        refs = torch.cat([references[k] for k in idx], dim=0)
        data = {"image0": refs, "image1": query.expand(len(idx), -1, -1, -1).contiguous()}
        matcher(data)
        b = data["m_bids"]
        for j, k in enumerate(idx):
            sel = b == j
            out[k] = {"mkpts0_f": data["mkpts0_f"][sel], "mkpts1_f": data["mkpts1_f"][sel], "mconf": data["mconf"][sel]}
    return out
