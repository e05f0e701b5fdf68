"""CPU oracle of motion-only bundle adjustment (pose refinement against fixed landmarks), SURVEY §8 row f-2.

TEST INFRASTRUCTURE ONLY (imported by tests/ only; never by visualodometry_b200/).  PARITY UNPINNED for the LM rule:
the reference has no pose refinement of its own -- it calls `cv2.solvePnPRansac` (`src/modules/vo.py:131`) and takes
what OpenCV returns.  What IS pinned is the boundary: the projection / residual convention
(`src/modules/frontend.py:128-140`, `vo.py:135-149`: `x_c = R_cw X + t_cw`, `uv = K x_c / z`, zero distortion,
residual in pixels) and the minimiser itself (`tests/test_pose.py` checks the converged pose against
`cv2.solvePnP(..., SOLVEPNP_ITERATIVE)` started from the same pose, the function OpenCV's RANSAC ends with).

The algorithm is Spec a0 restricted to one camera block:
    r = (fx x/z + cx - u, fy y/z + cy - v);  z < z_min => weight 0, cost 0;  Huber(a): w = a/|r| beyond a
    J_c = dr/dx [-[x]x | I]  (left perturbation [dw, dt]);  H = sum w J^T J,  g = sum w J^T r
    D = clamp(diag H, 1e-6, 1e32);  (H + D/mu) d = -g;  R <- Exp(dw) R,  t <- Exp(dw) t + dt
    rho = (cost - cost_trial) / (1/2 (d^T D d / mu - g^T d));  accept if rho > 1e-3:
    mu <- min(1e16, mu / max(1/3, 1 - (2 rho - 1)^3)), nu <- 2;  else mu <- mu / nu, nu <- 2 nu.   mu0 = 1e4
"""
from __future__ import annotations

from dataclasses import dataclass, field

import numpy as np


@dataclass
class PoseSummary:
    status: int                 # 1 max iterations, 2 converged, 3 numerical failure
    iterations: int
    accepted: int
    rejected: int
    n_invalid: int
    initial_cost: float
    final_cost: float
    final_mu: float
    cost_hist: list = field(default_factory=list)
    mu_hist: list = field(default_factory=list)
    accept_hist: list = field(default_factory=list)


def _exp_so3(w):
    t2 = float(w @ w)
    if t2 < 1e-16:
        a, b = 1.0 - t2 / 6.0, 0.5 - t2 / 24.0
    else:
        t = np.sqrt(t2)
        a, b = np.sin(t) / t, (1.0 - np.cos(t)) / t2
    W = np.array([[0, -w[2], w[1]], [w[2], 0, -w[0]], [-w[1], w[0], 0.0]])
    return np.eye(3) + a * W + b * (W @ W)


def _terms(R, t, X, uv, intr, z_min, loss, a):
    fx, fy, cx, cy = intr
    xc = X @ R.T + t
    z = xc[:, 2]
    valid = z >= z_min
    iz = np.where(valid, 1.0 / np.where(valid, z, 1.0), 0.0)
    r = np.stack([fx * xc[:, 0] * iz + cx - uv[:, 0].astype(np.float64),
                  fy * xc[:, 1] * iz + cy - uv[:, 1].astype(np.float64)], axis=1)
    s = (r * r).sum(1)
    w = np.ones_like(s)
    rho = s.copy()
    if loss == "huber":
        big = s > a * a
        rs = np.sqrt(np.where(big, s, 1.0))
        rho = np.where(big, 2.0 * a * rs - a * a, s)
        w = np.where(big, a / rs, 1.0)
    r[~valid] = 0.0
    w[~valid] = 0.0
    rho[~valid] = 0.0
    return xc, iz, r, w, rho, valid


def _jac_cam(xc, iz, intr):
    fx, fy = intr[0], intr[1]
    x, y, z = xc[:, 0], xc[:, 1], xc[:, 2]
    a0, a2 = fx * iz, -fx * x * iz * iz
    b1, b2 = fy * iz, -fy * y * iz * iz
    zero = np.zeros_like(x)
    j0 = np.stack([a2 * y, a0 * z - a2 * x, -a0 * y, a0, zero, a2], axis=1)
    j1 = np.stack([-b1 * z + b2 * y, -b2 * x, b1 * x, zero, b1, b2], axis=1)
    return j0, j1


def refine_pose(T_cw, X, uv, intr, max_iters=10, loss="trivial", loss_scale=1.0, z_min=1e-3, init_mu=1e4,
                ftol=0.0, gtol=0.0, ptol=0.0):
    """T_cw (4,4), X (N,3), uv (N,2), intr (fx, fy, cx, cy) -> (T_cw refined, PoseSummary)."""
    T = np.asarray(T_cw, np.float64)
    R, t = T[:3, :3].copy(), T[:3, 3].copy()
    X = np.asarray(X, np.float64).reshape(-1, 3)
    uv = np.asarray(uv, np.float32).reshape(-1, 2)
    mu, nu = init_mu, 2.0
    sm = PoseSummary(1, 0, 0, 0, 0, 0.0, 0.0, mu)
    have_lin = False
    cost = 0.0
    it = 0
    while it < max_iters:
        fresh = not have_lin
        if not have_lin:
            xc, iz, r, w, rho, valid = _terms(R, t, X, uv, intr, z_min, loss, loss_scale)
            j0, j1 = _jac_cam(xc, iz, intr)
            H = np.einsum("n,ni,nj->ij", w, j0, j0) + np.einsum("n,ni,nj->ij", w, j1, j1)
            g = (w * r[:, 0]) @ j0 + (w * r[:, 1]) @ j1
            cost = 0.5 * float(rho.sum())
            ninv = int((~valid).sum())
            if it == 0:
                sm.initial_cost = cost
            have_lin = True
        D = np.clip(np.diag(H), 1e-6, 1e32)
        step_ok = True
        try:
            L = np.linalg.cholesky(H + np.diag(D / mu))
            d = -np.linalg.solve(L.T, np.linalg.solve(L, g))
            step_ok = bool(np.isfinite(d).all())
        except np.linalg.LinAlgError:
            step_ok = False
        accept, stop, rho_q, tc = False, 0, 0.0, 0.0
        if gtol > 0.0 and fresh and np.abs(g).max() <= gtol:
            sm.status = 2
            break
        if step_ok:
            E = _exp_so3(d[:3])
            Rt, tt = E @ R, E @ t + d[3:]
            _, _, _, _, rho_t, valid_t = _terms(Rt, tt, X, uv, intr, z_min, loss, loss_scale)
            tc = 0.5 * float(rho_t.sum())
            model = 0.5 * (float(d @ (D * d)) / mu - float(g @ d))
            if model > 0.0:
                rho_q = (cost - tc) / model
                # a step that pushes MORE observations behind the cheirality gate (they cost 0) is rejected
                accept = bool(np.isfinite(tc) and rho_q > 1e-3 and int((~valid_t).sum()) <= ninv)
        if accept:
            if ftol > 0.0 and abs(cost - tc) <= ftol * cost:
                stop = 2
            xn = np.sqrt(float((R * R).sum() + t @ t))
            if ptol > 0.0 and np.linalg.norm(d) <= ptol * (xn + ptol):
                stop = 2
        mu_used = mu
        if accept:
            tq = 2.0 * rho_q - 1.0
            mu = min(1e16, mu / max(1.0 / 3.0, 1.0 - tq ** 3))
            nu = 2.0
            cost, R, t, ninv = tc, Rt, tt, int((~valid_t).sum())
            have_lin = False
            sm.accepted += 1
        else:
            mu, nu = mu / nu, 2.0 * nu
            sm.rejected += 1
        sm.cost_hist.append(cost); sm.mu_hist.append(mu_used); sm.accept_hist.append(int(accept))
        if not stop and not (mu > 1e-32):
            stop = 3
        it += 1
        if stop:
            sm.status = stop
            break
    sm.iterations, sm.final_cost, sm.final_mu, sm.n_invalid = it, cost, mu, ninv
    out = np.eye(4)
    out[:3, :3], out[:3, 3] = R, t
    return out, sm


def track_pose(T_cw_prior, X, uv, intr, reproj_err=1.0, rounds=6, iters_per_round=10, z_min=1e-3, init_mu=1e4):
    """CPU restatement of `b2ba_track_pose` (include/b2ba.h): the stand-in for the reference's
    `cv2.solvePnPRansac(X, uv, K, None, reprojectionError=pnp_reproj_err)` (`src/modules/vo.py:135-141`) when a pose
    prior exists.  `rounds` x [`refine_pose` on the current inliers (Huber a = thr_r; last of several rounds:
    trivial loss) -> re-classify EVERY observation: inlier = z >= z_min and |reprojection error| < thr_r], with the
    graduated threshold thr_r = reproj_err * 2^(rounds - 2 - r) >= reproj_err (6 rounds: 16, 8, 4, 2, 1, 1).
    Returns (ok, T_cw, inlier indices, PoseSummary of the last round).  PINNED to the reference's call in
    tests/test_pose.py::test_oracle_tracker_matches_solvepnpransac (same inlier set, pose to the RANSAC refinement's
    own tolerance)."""
    T = np.asarray(T_cw_prior, np.float64).copy()
    X = np.asarray(X, np.float64).reshape(-1, 3)
    uv = np.asarray(uv, np.float32).reshape(-1, 2)
    mask = np.ones(len(X), bool)
    sm = None
    for r in range(rounds):
        loss = "trivial" if (rounds > 1 and r == rounds - 1) else "huber"
        thr = reproj_err * float(1 << max(0, rounds - 2 - r))
        T, sm = refine_pose(T, X[mask], uv[mask], intr, max_iters=iters_per_round, loss=loss, loss_scale=thr,
                            z_min=z_min, init_mu=init_mu)
        _, _, res, _, rho, valid = _terms(T[:3, :3], T[:3, 3], X, uv, intr, z_min, "trivial", 1.0)
        mask = valid & (rho < thr * thr)
        if mask.sum() < 6:
            break
    ok = bool(mask.sum() >= 6 and np.isfinite(T).all())
    return ok, T, np.flatnonzero(mask), sm


def _splitmix64(x):
    """x: uint64 array -> uint64 array (wrapping arithmetic), the hash `b2ba_ransac_pose` samples with."""
    x = np.asarray(x, np.uint64)
    with np.errstate(over="ignore"):
        x = x + np.uint64(0x9E3779B97F4A7C15)
        x = (x ^ (x >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        x = (x ^ (x >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        return x ^ (x >> np.uint64(31))


def ransac_samples(n, n_hyp, seed, k=4):
    """the (n_hyp, k) observation indices `b2ba_ransac_pose` draws: hash(seed + j*k + q) mod n, collisions moved on."""
    with np.errstate(over="ignore"):
        base = np.uint64(seed) + np.arange(n_hyp * k, dtype=np.uint64)
    idx = (_splitmix64(base) % np.uint64(n)).astype(np.int64).reshape(n_hyp, k)
    for j in range(n_hyp):
        for q in range(k):
            i = int(idx[j, q])
            again = True
            while again:
                again = False
                for p in range(q):
                    if idx[j, p] == i:
                        i = (i + 1) % n
                        again = True
            idx[j, q] = i
    return idx


def ransac_pose(T_cw_prior, X, uv, intr, reproj_err=1.0, n_hyp=1024, hyp_iters=8, seed=0, refine_iters=10, z_min=1e-3):
    """CPU restatement of `b2ba_ransac_pose` (include/b2ba.h), the GPU counterpart of the reference's
    `cv2.solvePnPRansac` call (`src/modules/vo.py:135-141`): hypotheses = LM fits of 4 sampled matches from the prior,
    consensus = z >= z_min and |reprojection error| < reproj_err over all matches, best consensus (lowest index on ties)
    refined by two rounds of LM on the consensus set (Huber, then trivial) with re-classification.
    Returns (ok, T_cw, inlier indices, PoseSummary, winning consensus)."""
    T0 = np.asarray(T_cw_prior, np.float64)
    X = np.asarray(X, np.float64).reshape(-1, 3)
    uv = np.asarray(uv, np.float32).reshape(-1, 2)
    n = len(X)
    if n < 6:
        return False, T0.copy(), np.zeros(0, np.int64), None, 0
    idx = ransac_samples(n, n_hyp, seed)
    best, best_T = -1, T0
    for j in range(n_hyp):
        T, sm = refine_pose(T0, X[idx[j]], uv[idx[j]], intr, max_iters=hyp_iters, z_min=z_min)
        c = 0
        if sm.status != 3 and np.isfinite(T).all():
            _, _, _, _, rho, valid = _terms(T[:3, :3], T[:3, 3], X, uv, intr, z_min, "trivial", 1.0)
            c = int((valid & (rho < reproj_err * reproj_err)).sum())
        if c > best:
            best, best_T = c, T
    T = best_T
    _, _, _, _, rho, valid = _terms(T[:3, :3], T[:3, 3], X, uv, intr, z_min, "trivial", 1.0)
    mask = valid & (rho < reproj_err * reproj_err)
    sm = None
    for r in range(2):
        T, sm = refine_pose(T, X[mask], uv[mask], intr, max_iters=refine_iters, loss="huber" if r == 0 else "trivial",
                            loss_scale=reproj_err, z_min=z_min)
        _, _, _, _, rho, valid = _terms(T[:3, :3], T[:3, 3], X, uv, intr, z_min, "trivial", 1.0)
        mask = valid & (rho < reproj_err * reproj_err)
        if mask.sum() < 6:
            break
    ok = bool(mask.sum() >= 6 and np.isfinite(T).all())
    return ok, T, np.flatnonzero(mask), sm, best
