"""ORACLE -- TEST INFRASTRUCTURE ONLY.  The same BA problem handed to an INDEPENDENT solver that is in the image:
scipy.optimize.least_squares (trust-region reflective, sparse finite-difference Jacobian).  SURVEY.md §8(c) item 3 and
§8(d) "CPU baseline beside it (3)": converged cost for cross-solver sanity, wall time as a second CPU comparator for
C1 / C3.  scipy is declared by the reference (pyproject.toml:14) and never called by it.

scipy's robust losses act per SCALAR residual, Spec a0's Huber per 2-vector, so only loss="linear" is comparable."""
from __future__ import annotations

import time

import numpy as np
import scipy.sparse as sp
from scipy.optimize import least_squares

from oracle.ba_numpy import exp_so3


def _unpack(p, x, free):
    poses = p.poses.copy()
    for i, c in enumerate(free):
        d = x[6 * i:6 * i + 6]
        E = exp_so3(d[:3])
        poses[c, :9] = (E @ p.poses[c, :9].reshape(3, 3)).reshape(9)
        poses[c, 9:] = E @ p.poses[c, 9:] + d[3:]
    return poses, p.points + x[6 * len(free):].reshape(-1, 3)


def residuals(p, poses, pts):
    R = poses[p.obs_cam, :9].reshape(-1, 3, 3)
    xc = np.einsum("nij,nj->ni", R, pts[p.obs_pt]) + poses[p.obs_cam, 9:]
    return np.stack([p.intr[0] * xc[:, 0] / xc[:, 2] + p.intr[2] - p.obs_uv[:, 0],
                     p.intr[1] * xc[:, 1] / xc[:, 2] + p.intr[3] - p.obs_uv[:, 1]], 1).reshape(-1)


def solve(p, max_nfev=100, ftol=1e-12, xtol=1e-12, gtol=1e-12):
    """Returns dict(poses, points, cost, nfev, njev, seconds); cost = 1/2 sum r^2 like Spec a0's (trivial loss)."""
    free = [c for c in range(p.nc) if not p.fixed_cam[c]]
    fi = {c: i for i, c in enumerate(free)}
    nx = 6 * len(free) + 3 * p.np_
    rows, cols = [], []
    for k in range(2):
        r = 2 * np.arange(p.no) + k
        f = np.array([fi.get(int(c), -1) for c in p.obs_cam])
        for j in range(6):
            rows.append(r[f >= 0]); cols.append(6 * f[f >= 0] + j)
        for j in range(3):
            rows.append(r); cols.append(6 * len(free) + 3 * p.obs_pt.astype(np.int64) + j)
    rows, cols = np.concatenate(rows), np.concatenate(cols)
    spars = sp.csr_matrix((np.ones(len(rows), np.int8), (rows, cols)), shape=(2 * p.no, nx))
    t0 = time.perf_counter()
    res = least_squares(lambda x: residuals(p, *_unpack(p, x, free)), np.zeros(nx), jac_sparsity=spars, method="trf",
                        x_scale="jac", ftol=ftol, xtol=xtol, gtol=gtol, max_nfev=max_nfev)
    dt = time.perf_counter() - t0
    poses, pts = _unpack(p, res.x, free)
    return dict(poses=poses, points=pts, cost=float(res.cost), nfev=int(res.nfev), njev=int(res.njev), seconds=dt,
                status=int(res.status))
