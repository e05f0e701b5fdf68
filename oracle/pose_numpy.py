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
                accept = bool(np.isfinite(tc) and rho_q > 1e-3)
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
