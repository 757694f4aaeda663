"""RANSAC-PnP on the GPU behind the reference's ``ransac_PnP`` signature (src/utils/metric_utils.py:121-209).

``ransac_PnP`` is the drop-in for one frame (numpy in, numpy out, like the reference); ``ransac_pnp_batch`` is what the
matcher pipeline uses: it consumes the device-resident ``[sum M, 6]`` match rows (x3d, y3d, z3d, u, v, conf) of a whole
batch of frames (``sharding.rows_onepose``) and returns one pose per frame without the matches leaving HBM.
No CPU fallback: the kernels live in libopb200.so (csrc/pnp.cu)."""
import ctypes
from typing import Optional, Sequence, Tuple

import numpy as np
import torch

from . import _lib, ops

_c_sz = ctypes.c_size_t


def ransac_pnp_batch(rows: torch.Tensor, counts, intrinsics, reproj_error_px: float = 5.0, iterations: int = 10000,
                     seed: int = 0) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    """rows [sum M, >= 5] fp32 on the device with (x3d, y3d, z3d, u, v) in the first five columns; ``counts`` per-frame row
    counts (sequence or int tensor, frames back to back); ``intrinsics`` [n_frames, 4] = (fx, fy, cx, cy) or one [3, 3] / [4]
    for all frames.  Returns (pose [n_frames, 3, 4] float64, inlier mask [sum M] bool, counts [n_frames, 2] int32 =
    (inliers of the refined pose, best RANSAC score)), all on the device."""
    if not rows.is_cuda:
        raise _lib.Opb200Error("ransac_pnp_batch needs CUDA tensors (there is no CPU path)")
    rows = rows.float()
    if rows.dim() != 2 or rows.shape[1] < 5 or rows.stride(1) != 1:
        rows = rows.contiguous()
        if rows.dim() != 2 or rows.shape[1] < 5:
            raise _lib.Opb200Error(f"rows must be [M, >= 5], got {tuple(rows.shape)}")
    dev = rows.device
    cnt = torch.as_tensor(counts, dtype=torch.int64, device=dev).flatten()
    n_frames = int(cnt.numel())
    if n_frames < 1:
        raise _lib.Opb200Error("ransac_pnp_batch: no frames")
    offsets = torch.zeros(n_frames + 1, dtype=torch.int32, device=dev)
    offsets[1:] = torch.cumsum(cnt, 0).to(torch.int32)
    k = torch.as_tensor(intrinsics, dtype=torch.float64, device=dev)
    if k.shape[-2:] == (3, 3):
        k = torch.stack([k[..., 0, 0], k[..., 1, 1], k[..., 0, 2], k[..., 1, 2]], dim=-1)
    k = k.reshape(-1, 4)
    if k.shape[0] == 1:
        k = k.expand(n_frames, 4)
    if k.shape[0] != n_frames:
        raise _lib.Opb200Error(f"intrinsics: {k.shape[0]} cameras for {n_frames} frames")
    k = k.contiguous()
    l = _lib.lib()
    l.opb200_ransac_pnp_workspace_bytes.restype = _c_sz
    ws = ops.workspaces.get(dev, "pnp", int(l.opb200_ransac_pnp_workspace_bytes(n_frames, int(iterations))))
    pose = torch.empty(n_frames, 3, 4, dtype=torch.float64, device=dev)
    mask = torch.zeros(max(1, rows.shape[0]), dtype=torch.uint8, device=dev)
    out_counts = torch.zeros(n_frames, 2, dtype=torch.int32, device=dev)
    stride = int(rows.stride(0)) if rows.shape[0] else int(rows.shape[1])
    base = rows.data_ptr() if rows.shape[0] else mask.data_ptr()
    with ops._Stage("ransac_pnp"):
        st = l.opb200_ransac_pnp(ctypes.c_void_p(base), stride, ctypes.c_void_p(base + 12), stride, _lib.ptr(offsets), n_frames,
                                 _lib.ptr(k), ctypes.c_float(reproj_error_px), int(iterations), ctypes.c_uint(seed & 0xFFFFFFFF),
                                 _lib.ptr(pose), _lib.ptr(mask), _lib.ptr(out_counts), _lib.ptr(ws), _c_sz(ws.numel()),
                                 _lib.stream_ptr())
    _lib.check(st, "opb200_ransac_pnp")
    return pose, mask[: rows.shape[0]].bool(), out_counts


def ransac_PnP(K, pts_2d, pts_3d, scale=1, pnp_reprojection_error=5, img_hw=None, use_pycolmap_ransac=False,
               iterations: int = 10000, seed: int = 0, device: Optional[torch.device] = None):
    """Drop-in for the reference's ``ransac_PnP``: returns (pose [3,4], pose_homo [4,4], inlier indices) as numpy float64 /
    int arrays.  ``use_pycolmap_ransac`` and ``img_hw`` are accepted for signature compatibility (both reference back ends
    are replaced by the same device solver); a failed solve returns the identity pose and no inliers, like the reference."""
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    p2 = torch.as_tensor(np.ascontiguousarray(pts_2d), dtype=torch.float32) if not isinstance(pts_2d, torch.Tensor) else pts_2d.float()
    p3 = torch.as_tensor(np.ascontiguousarray(pts_3d), dtype=torch.float32) if not isinstance(pts_3d, torch.Tensor) else pts_3d.float()
    p2, p3 = p2.to(dev).reshape(-1, 2), p3.to(dev).reshape(-1, 3) * float(scale)
    rows = torch.cat([p3, p2], dim=1)
    kk = torch.as_tensor(np.asarray(K, dtype=np.float64)) if not isinstance(K, torch.Tensor) else K.double()
    pose, mask, counts = ransac_pnp_batch(rows, [rows.shape[0]], kk, float(pnp_reprojection_error), iterations, seed)
    pose = pose[0].cpu().numpy()
    n_in = int(counts[0, 0].item())
    if n_in == 0:
        return np.eye(4)[:3], np.eye(4), np.array([]).astype(bool)
    pose[:, 3] /= float(scale)
    pose_homo = np.concatenate([pose, np.array([[0.0, 0.0, 0.0, 1.0]])], axis=0)
    return pose, pose_homo, torch.nonzero(mask).flatten().cpu().numpy()
