"""CPU oracle of the batched two-view triangulation + gating.  TEST INFRASTRUCTURE ONLY.

Restates `triangulate_points` (`src/modules/frontend.py:115-148`): linear (DLT) triangulation as done by
`cv2.triangulatePoints` (OpenCV 4.x `cvTriangulatePoints`: per point the 4x4 system
    [u1 P1[2] - P1[0];  v1 P1[2] - P1[1];  u2 P2[2] - P2[0];  v2 P2[2] - P2[1]] X = 0,
solved by the right singular vector of the smallest singular value), de-homogenisation, depth in the SECOND camera,
reprojection error in the second camera, gates `depth > min_depth` and `err < max_reproj_err` (`config.py:31-32`).
PINNED to the reference: `tests/golden/reference_triangulation.npz` holds the outputs of the reference's own
function (`tests/golden/make_reference_fixtures.py`), `tests/test_triangulate.py` checks this file against them.
The SVD here is a one-sided Jacobi (Hestenes) iteration, the same the CUDA kernel runs.
"""
from __future__ import annotations

import numpy as np


def _null_vector(A, sweeps=12):
    """A (N,4,4) -> (N,4) right singular vector of the smallest singular value, by one-sided Jacobi."""
    A = A.copy()
    n = A.shape[0]
    V = np.tile(np.eye(4), (n, 1, 1))
    for _ in range(sweeps):
        for p in range(3):
            for q in range(p + 1, 4):
                al = (A[:, :, p] ** 2).sum(1); be = (A[:, :, q] ** 2).sum(1); ga = (A[:, :, p] * A[:, :, q]).sum(1)
                rot = np.abs(ga) > 1e-300 + 1e-17 * np.sqrt(al * be)
                g = np.where(rot, ga, 1.0)
                zeta = (be - al) / (2.0 * g)
                t = np.where(zeta >= 0, 1.0, -1.0) / (np.abs(zeta) + np.sqrt(1.0 + zeta * zeta))
                c = 1.0 / np.sqrt(1.0 + t * t)
                s = c * t
                c = np.where(rot, c, 1.0); s = np.where(rot, s, 0.0)
                for M in (A, V):
                    mp, mq = M[:, :, p].copy(), M[:, :, q].copy()
                    M[:, :, p] = c[:, None] * mp - s[:, None] * mq
                    M[:, :, q] = s[:, None] * mp + c[:, None] * mq
    k = (A ** 2).sum(1).argmin(1)
    return V[np.arange(n), :, k]


def triangulate(P1, P2, uv1, uv2, min_depth, max_reproj_err):
    """P1, P2 (3,4) = K [R|t] of the two views; uv1, uv2 (N,2) -> (X (N,3) f64, keep (N,) bool, depth2, err2)."""
    P1, P2 = np.asarray(P1, np.float64), np.asarray(P2, np.float64)
    u1 = np.asarray(uv1, np.float64); u2 = np.asarray(uv2, np.float64)
    A = np.stack([u1[:, 0, None] * P1[2] - P1[0], u1[:, 1, None] * P1[2] - P1[1],
                  u2[:, 0, None] * P2[2] - P2[0], u2[:, 1, None] * P2[2] - P2[1]], axis=1)
    v = _null_vector(A)
    X = v[:, :3] / v[:, 3:4]
    h = X @ P2[:, :3].T + P2[:, 3]
    depth2 = h[:, 2]
    with np.errstate(divide="ignore", invalid="ignore"):
        err = np.linalg.norm(h[:, :2] / h[:, 2:3] - u2, axis=1)
    keep = (depth2 > min_depth) & (err < max_reproj_err)
    return X, keep, depth2, err
