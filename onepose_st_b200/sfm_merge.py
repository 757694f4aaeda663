"""Keypoint-Free SfM match post-processing on the GPU.

Drop-in for the chain ``Match2Pts2D`` -> ``points2D_worker`` -> ``update_matches`` -> ``transform_points2D`` of
src/KeypointFreeSfM/coarse_match (utils.py:5-60, coarse_match_worker.py:87-184; driver coarse_match.py:155-186): given the
LoFTR match rows of all image pairs it returns, per image, the merged integer keypoints ranked by summed confidence and,
per pair, the matches as indices into those lists -- bit-identical to the reference (integer keys; scores added in fp64 in
the reference's order).  It consumes the gathered ``[sum M, 5]`` buffer of ``sharding.gather_matches`` directly.

Own kernels (csrc/sfm_merge.cu) build the keys and add the scores; the sorts and prefix sums between them are library
primitives (torch.sort / cumsum), like cuBLAS for a plain GEMM.  No CPU path."""
import ctypes
from typing import Dict, List, Sequence, Tuple

import torch

from . import _lib

_c_ll = ctypes.c_longlong


def merge_pair_matches(rows: torch.Tensor, pair_counts: Sequence[int], pair_images, n_images: int
                       ) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor, torch.Tensor]:
    """rows [sum M, 5] fp32 on the device (pairs back to back, in pair order); ``pair_counts`` [P] rows per pair;
    ``pair_images`` [P, 2] image ids.  Returns, all on the device:
      keypoints [K, 2] float32  -- all images back to back, image i owns image_offsets[i] .. image_offsets[i + 1] - 1,
                                   inside an image in rank order (summed confidence descending; ties: (x, y) ascending)
      scores    [K]    float32
      image_offsets [n_images + 1] int64
      match_ids [sum M, 2] int64 -- (rank of the match's keypoint in image0's list, in image1's list)"""
    if not rows.is_cuda:
        raise _lib.Opb200Error("merge_pair_matches needs CUDA tensors (there is no CPU path)")
    dev = rows.device
    rows = rows.float().contiguous()
    n_rows = int(rows.shape[0])
    if rows.dim() != 2 or rows.shape[1] != 5:
        raise _lib.Opb200Error(f"rows must be [M, 5], got {tuple(rows.shape)}")
    cnt = torch.as_tensor(pair_counts, dtype=torch.int64, device=dev).flatten()
    pim = torch.as_tensor(pair_images, dtype=torch.int32, device=dev).reshape(-1, 2).contiguous()
    if int(cnt.numel()) != int(pim.shape[0]):
        raise _lib.Opb200Error("pair_counts and pair_images disagree on the number of pairs")
    offs = torch.zeros(n_images + 1, dtype=torch.int64, device=dev)
    if n_rows == 0:
        return (torch.empty(0, 2, device=dev), torch.empty(0, device=dev), offs, torch.empty(0, 2, dtype=torch.int64, device=dev))
    pair_of_row = torch.repeat_interleave(torch.arange(cnt.numel(), device=dev, dtype=torch.int32), cnt)
    if int(pair_of_row.numel()) != n_rows:
        raise _lib.Opb200Error("pair_counts do not add up to the number of rows")
    l = _lib.lib()
    keys = torch.empty(2 * n_rows, dtype=torch.int64, device=dev)
    bad = torch.zeros(1, dtype=torch.int32, device=dev)
    st = l.opb200_sfm_keys(_lib.ptr(rows), _c_ll(n_rows), _lib.ptr(pair_of_row), _lib.ptr(pim), _lib.ptr(keys), _lib.ptr(bad),
                           _lib.stream_ptr())
    _lib.check(st, "opb200_sfm_keys")
    skeys, order = torch.sort(keys, stable=True)                 # equal keys keep their (pair, row) order
    head = torch.ones_like(skeys, dtype=torch.bool)
    head[1:] = skeys[1:] != skeys[:-1]
    group_of_sorted = torch.cumsum(head, 0) - 1                  # group index of every sorted entry
    start = torch.nonzero(head).flatten()
    n_groups = int(start.numel())                                # host sync (output size), as np.unique has
    if int(bad.item()):
        raise ValueError("match coordinates must be finite, non-negative and below 2^20")
    start = torch.cat([start, torch.tensor([2 * n_rows], device=dev)])
    sums = torch.empty(n_groups, dtype=torch.float64, device=dev)
    st = l.opb200_sfm_group_sums(_lib.ptr(rows), _lib.ptr(order), _lib.ptr(start), _c_ll(n_groups), _lib.ptr(sums),
                                 _lib.stream_ptr())
    _lib.check(st, "opb200_sfm_group_sums")
    gkeys = skeys[start[:-1]]                                    # ascending: (image, x, y) lexicographic
    gimg = gkeys >> 40
    # rank inside an image: score descending, ties in key order -- two stable sorts (score, then image)
    o1 = torch.sort(sums, descending=True, stable=True).indices
    o2 = torch.sort(gimg[o1], stable=True).indices
    ranked = o1[o2]                                              # group ids in output order
    per_image = torch.bincount(gimg, minlength=n_images)
    offs[1:] = torch.cumsum(per_image, 0)
    rk = gkeys[ranked]
    keypoints = torch.stack([(rk >> 20) & 0xFFFFF, rk & 0xFFFFF], dim=1).float()
    scores = sums[ranked].float()
    pos = torch.empty(n_groups, dtype=torch.int64, device=dev)
    pos[ranked] = torch.arange(n_groups, device=dev)
    rank_in_image = pos - offs[gimg]
    ids = torch.empty(2 * n_rows, dtype=torch.int64, device=dev)
    ids[order] = rank_in_image[group_of_sorted]
    return keypoints, scores, offs, ids.view(n_rows, 2)


def to_reference_dicts(keypoints, scores, image_offsets, match_ids, pair_counts, pair_names: List[str], image_names: List[str]
                       ) -> Tuple[Dict, Dict, Dict]:
    """The reference's containers (coarse_match.py:155-186): (final_keypoints {image: [K,2] float32}, final_scores
    {image: [K] float32}, updated_matches {pair: [M,2] int}) as numpy arrays on the host."""
    kp, sc, off, ids = keypoints.cpu().numpy(), scores.cpu().numpy(), image_offsets.cpu().tolist(), match_ids.cpu().numpy()
    fk = {n: kp[off[i]: off[i + 1]] for i, n in enumerate(image_names) if off[i + 1] > off[i]}
    fs = {n: sc[off[i]: off[i + 1]] for i, n in enumerate(image_names) if off[i + 1] > off[i]}
    um, r0 = {}, 0
    for name, c in zip(pair_names, list(pair_counts)):
        um[name] = ids[r0: r0 + int(c)].astype(int)
        r0 += int(c)
    return fk, fs, um
