"""CPU restatement (numpy, float64) of the RANSAC-PnP the CUDA kernels in csrc/pnp.cu run (TEST INFRASTRUCTURE).

The reference's pose step is ``ransac_PnP`` (src/utils/metric_utils.py:121-209): OpenCV ``solvePnPRansac`` (EPnP model,
5 px reprojection error, 10000 iterations) or pycolmap's ``absolute_pose_estimation`` (P3P RANSAC + non-linear refinement;
not installable here).  Both are third-party code absent from /root/reference; their published behaviour is: draw minimal
samples, score hypotheses by the number of points whose reprojection error is below the threshold, re-estimate on the
inliers of the best hypothesis.  This file restates that scheme the way the GPU runs it, function by function, so that
the kernels can be checked step by step (P3P roots, hypothesis scores, refined pose) on small inputs; the end-to-end
oracle of the pose is OpenCV itself (tests/test_gpu_pnp.py, tests/test_gpu_pose.py).

    sample  : 4 distinct indices from a counter-based hash (hash32, below) -- identical on CPU and GPU
    model   : Grunert's P3P on the first three (Haralick et al., "Review and analysis of solutions of the three point
              perspective pose estimation problem", IJCV 1994, eqs. for the quartic in v = s3/s1), up to four poses,
              the fourth point's reprojection error picks one
    score   : number of points with squared reprojection error < thr^2 (ties: smaller hypothesis index)
    refine  : Gauss-Newton on the inliers (left-multiplied so(3) x R^3 increments), inliers re-evaluated, twice
"""
from __future__ import annotations

import numpy as np


def hash32(x: np.ndarray) -> np.ndarray:
    """lowbias32 (Chris Wellons): the counter-based generator both sides use for the minimal samples."""
    x = np.asarray(x, dtype=np.uint64) & 0xFFFFFFFF
    x ^= x >> 16
    x = (x * 0x7FEB352D) & 0xFFFFFFFF
    x ^= x >> 15
    x = (x * 0x846CA68B) & 0xFFFFFFFF
    x ^= x >> 16
    return x


def sample4(h: int, m: int, seed: int) -> np.ndarray:
    """Four distinct indices in [0, m) for hypothesis ``h``: draw k is hash32(seed + 4*h*8 + attempt), rejected while it
    repeats an earlier index (at most 8 attempts per draw, then the next free index)."""
    out = []
    ctr = (seed + h * 32) & 0xFFFFFFFF
    for k in range(4):
        idx = None
        for a in range(8):
            c = int(hash32(np.uint64((ctr + k * 8 + a) & 0xFFFFFFFF))) % m
            if c not in out:
                idx = c
                break
        if idx is None:
            idx = next(i for i in range(m) if i not in out)
        out.append(idx)
    return np.array(out)


def solve_quartic(a4, a3, a2, a1, a0):
    """Real roots of a4 x^4 + ... + a0 (Ferrari via the resolvent cubic, each root polished by two Newton steps)."""
    if abs(a4) < 1e-14 * max(abs(a3), abs(a2), abs(a1), abs(a0), 1e-300):
        return []
    b, c, d, e = a3 / a4, a2 / a4, a1 / a4, a0 / a4
    # depressed quartic y^4 + p y^2 + q y + r, x = y - b/4
    p = c - 3 * b * b / 8
    q = d - b * c / 2 + b ** 3 / 8
    r = e - b * d / 4 + b * b * c / 16 - 3 * b ** 4 / 256
    roots = []
    if abs(q) < 1e-14:
        disc = p * p - 4 * r
        if disc >= 0:
            for z in ((-p + np.sqrt(disc)) / 2, (-p - np.sqrt(disc)) / 2):
                if z >= 0:
                    roots += [np.sqrt(z), -np.sqrt(z)]
    else:
        # resolvent cubic 8 m^3 + 8 p m^2 + (2 p^2 - 8 r) m - q^2 = 0 : any real root m > 0
        A, B, C = p, p * p / 4 - r, -q * q / 8
        # m^3 + A m^2 + B m + C = 0 -> t^3 + P t + Q, m = t - A/3
        P = B - A * A / 3
        Q = 2 * A ** 3 / 27 - A * B / 3 + C
        D = Q * Q / 4 + P ** 3 / 27
        if D >= 0:
            sD = np.sqrt(D)
            t = np.cbrt(-Q / 2 + sD) + np.cbrt(-Q / 2 - sD)
        else:
            rho = np.sqrt(-P / 3)
            th = np.arccos(np.clip(-Q / (2 * rho ** 3), -1, 1))
            t = 2 * rho * np.cos(th / 3)  # the largest of the three real roots
        m = t - A / 3
        if m <= 0:
            return []
        s = np.sqrt(2 * m)
        for sg in (1.0, -1.0):
            disc = -(2 * p + 2 * m + sg * 2 * q / s)
            if disc >= 0:
                roots += [(sg * s + np.sqrt(disc)) / 2, (sg * s - np.sqrt(disc)) / 2]
    out = []
    for y in roots:
        x = y - b / 4
        for _ in range(2):
            f = (((a4 * x + a3) * x + a2) * x + a1) * x + a0
            df = ((4 * a4 * x + 3 * a3) * x + 2 * a2) * x + a1
            if df != 0:
                x = x - f / df
        out.append(x)
    return out


def p3p_grunert(P, f):
    """P [3,3] world points (rows), f [3,3] unit bearing vectors (rows) -> list of (R [3,3], t [3]) with R P_i + t = s_i f_i."""
    a2 = np.sum((P[1] - P[2]) ** 2)
    b2 = np.sum((P[0] - P[2]) ** 2)
    c2 = np.sum((P[0] - P[1]) ** 2)
    if min(a2, b2, c2) < 1e-18:
        return []
    ca, cb, cg = f[1] @ f[2], f[0] @ f[2], f[0] @ f[1]
    k1 = (a2 - c2) / b2
    k2 = (a2 + c2) / b2
    A4 = (k1 - 1) ** 2 - 4 * c2 / b2 * ca * ca
    A3 = 4 * (k1 * (1 - k1) * cb - (1 - k2) * ca * cg + 2 * c2 / b2 * ca * ca * cb)
    A2 = 2 * (k1 * k1 - 1 + 2 * k1 * k1 * cb * cb + 2 * (b2 - c2) / b2 * ca * ca - 4 * k2 * ca * cb * cg + 2 * (b2 - a2) / b2 * cg * cg)
    A1 = 4 * (-k1 * (1 + k1) * cb + 2 * a2 / b2 * cg * cg * cb - (1 - k2) * ca * cg)
    A0 = (1 + k1) ** 2 - 4 * a2 / b2 * cg * cg
    sols = []
    for v in solve_quartic(A4, A3, A2, A1, A0):
        if not np.isfinite(v) or v <= 0:
            continue
        den = 2 * (cg - v * ca)
        if abs(den) < 1e-12:
            continue
        u = ((k1 - 1) * v * v - 2 * k1 * cb * v + 1 + k1) / den
        if u <= 0:
            continue
        d = 1 + v * v - 2 * v * cb
        if d <= 0:
            continue
        s1 = np.sqrt(b2 / d)
        s = np.array([s1, u * s1, v * s1])
        Q = s[:, None] * f
        # the three side lengths must be reproduced (spurious roots of the squared equations are dropped here)
        if abs(np.sum((Q[1] - Q[2]) ** 2) - a2) > 1e-6 * a2 or abs(np.sum((Q[0] - Q[1]) ** 2) - c2) > 1e-6 * c2:
            continue
        R = _frame(Q) @ _frame(P).T
        t = Q[0] - R @ P[0]
        sols.append((R, t))
    return sols


def _frame(X):
    e1 = X[1] - X[0]
    e1 = e1 / np.linalg.norm(e1)
    e3 = np.cross(e1, X[2] - X[0])
    e3 = e3 / np.linalg.norm(e3)
    e2 = np.cross(e3, e1)
    return np.stack([e1, e2, e3], 1)


def reproj_sq(R, t, K4, p3, p2):
    fx, fy, cx, cy = K4
    X = p3 @ R.T + t
    z = X[:, 2]
    ok = z > 1e-9
    zs = np.where(ok, z, 1.0)
    du = fx * X[:, 0] / zs + cx - p2[:, 0]
    dv = fy * X[:, 1] / zs + cy - p2[:, 1]
    return np.where(ok, du * du + dv * dv, np.inf)


def hypothesis(h, K4, p3, p2, seed):
    """Pose of hypothesis ``h`` or None."""
    fx, fy, cx, cy = K4
    idx = sample4(h, len(p3), seed)
    f = np.stack([(p2[idx[:3], 0] - cx) / fx, (p2[idx[:3], 1] - cy) / fy, np.ones(3)], 1)
    f = f / np.linalg.norm(f, axis=1, keepdims=True)
    best, best_e = None, np.inf
    for R, t in p3p_grunert(p3[idx[:3]], f):
        e = reproj_sq(R, t, K4, p3[idx[3:4]], p2[idx[3:4]])[0]
        if e < best_e:
            best, best_e = (R, t), e
    return best


def _exp_so3(w):
    th = np.linalg.norm(w)
    Kx = np.array([[0, -w[2], w[1]], [w[2], 0, -w[0]], [-w[1], w[0], 0]])
    if th < 1e-12:
        return np.eye(3) + Kx
    return np.eye(3) + np.sin(th) / th * Kx + (1 - np.cos(th)) / (th * th) * (Kx @ Kx)


def refine(R, t, K4, p3, p2, inl, iters=10):
    """Gauss-Newton on the reprojection error of the inliers; X' = Exp(w) (R P + t) + dt."""
    fx, fy, cx, cy = K4
    for _ in range(iters):
        X = p3[inl] @ R.T + t
        z = X[:, 2]
        ru = p2[inl, 0] - (fx * X[:, 0] / z + cx)
        rv = p2[inl, 1] - (fy * X[:, 1] / z + cy)
        n = len(z)
        J = np.zeros((2 * n, 6))
        ju = np.stack([fx / z, np.zeros(n), -fx * X[:, 0] / z ** 2], 1)
        jv = np.stack([np.zeros(n), fy / z, -fy * X[:, 1] / z ** 2], 1)
        for row, jp in ((0, ju), (1, jv)):
            # d X / d w = -[X]x : (w x X) ; columns of -[X]x are e_k x X ... J_w = jp @ (-[X]x)
            Jw = np.stack([jp[:, 2] * X[:, 1] - jp[:, 1] * X[:, 2],
                           jp[:, 0] * X[:, 2] - jp[:, 2] * X[:, 0],
                           jp[:, 1] * X[:, 0] - jp[:, 0] * X[:, 1]], 1)
            J[row::2, :3] = Jw
            J[row::2, 3:] = jp
        r = np.empty(2 * n)
        r[0::2], r[1::2] = ru, rv
        H = J.T @ J + 1e-9 * np.eye(6)
        d = np.linalg.solve(H, J.T @ r)
        E = _exp_so3(d[:3])
        R, t = E @ R, E @ t + d[3:]
    return R, t


def ransac_pnp(K4, p3, p2, thr=5.0, iterations=1024, seed=0):
    """(R, t, inlier mask) or None."""
    p3, p2 = np.asarray(p3, np.float64), np.asarray(p2, np.float64)
    m = len(p3)
    if m < 4:
        return None
    best, best_n = None, -1
    for h in range(iterations):
        hyp = hypothesis(h, K4, p3, p2, seed)
        if hyp is None:
            continue
        n = int((reproj_sq(hyp[0], hyp[1], K4, p3, p2) < thr * thr).sum())
        if n > best_n:
            best, best_n = hyp, n
    if best is None or best_n < 4:
        return None
    R, t = best
    inl = reproj_sq(R, t, K4, p3, p2) < thr * thr
    for _ in range(2):
        R, t = refine(R, t, K4, p3, p2, inl)
        inl2 = reproj_sq(R, t, K4, p3, p2) < thr * thr
        if inl2.sum() < 4:
            break
        inl = inl2
    return R, t, inl
