"""ORACLE -- TEST INFRASTRUCTURE ONLY (never imported by visualodometry_b200/).

Readable numpy restatement of "Spec a0" (SURVEY.md §8): Ceres-style Levenberg-Marquardt with Schur
elimination of the landmarks, every intermediate of an iteration returned by name.  It is the SECOND,
independently written implementation of the spec next to oracle/ba_cpu.cpp (C++): ba_cpu.cpp walks
point-major segments block by block with a closed-form adjugate; this file builds the whole Jacobian
as one scipy.sparse matrix, forms H = J^T W J and does the Schur complement with sparse block algebra
and numpy's LAPACK Cholesky.  tests/test_oracle_numpy.py holds the two against each other stage by
stage and trajectory by trajectory; the GPU engine is checked against both.

PARITY UNPINNED: the reference (cteufel13/VisualOdometry) ships no bundle adjustment.
What is pinned to the reference is the observation model this file shares with ba_cpu.cpp:
  x_c = R_cw X + t_cw, uv = K x_c / z, zero distortion   (src/modules/frontend.py:128-139, vo.py:139)
  residual = projected - observed, pixels                (frontend.py:140)
  cheirality gate z >= min_depth = 1e-3                  (frontend.py:135, src/config/config.py:31)

Spec a0 in one paragraph.  Unknowns: T_cw,c = (R_c, t_c) of the non-fixed cameras and X_p.  Left
perturbation [dw, dt]: R <- Exp(dw) R, t <- Exp(dw) t + dt, X <- X + dX.  Loss: trivial, or Huber(a)
as IRLS weight w = rho'(s), s = |r|^2.  H = sum w J^T J, g = sum w J^T r, D = clamp(diag H, 1e-6,
1e32); solve (H + D / mu) delta = -g; model decrease = (delta^T D delta / mu - g^T delta) / 2; gain
ratio rho = (cost - cost_trial) / model decrease; accept if rho > 1e-3 AND the trial state does not
put more observations behind the cheirality gate than the current one: mu <- min(1e16, mu /
max(1/3, 1 - (2 rho - 1)^3)), nu <- 2; else mu <- mu / nu, nu <- 2 nu.  mu0 = 1e4.  The linearisation
is reused after a rejected step.
"""
from __future__ import annotations

from dataclasses import dataclass, field

import numpy as np
import scipy.sparse as sp

DIAG_MIN, DIAG_MAX, MIN_RATIO, MU_MAX, MU_MIN = 1e-6, 1e32, 1e-3, 1e16, 1e-32


def exp_so3(w):
    """Rodrigues with the spec's series branch (theta^2 < 1e-16)."""
    w = np.asarray(w, np.float64)
    t2 = float(w @ w)
    if t2 < 1e-16:
        a, b = 1.0 - t2 / 6.0, 0.5 - t2 / 24.0
    else:
        t = np.sqrt(t2)
        a, b = np.sin(t) / t, (1.0 - np.cos(t)) / t2
    W = np.array([[0.0, -w[2], w[1]], [w[2], 0.0, -w[0]], [-w[1], w[0], 0.0]])
    return np.eye(3) + a * W + b * (W @ W)


def observe(poses, pts, obs_cam, obs_pt, obs_uv, intr, z_min=1e-3, loss="trivial", loss_scale=1.0, jac=True):
    """a1: residual, robust weight / loss value and the analytic 2x6 / 2x3 Jacobians of every observation.

    Observations behind the gate get r = 0, w = 0, rho = 0 and zero Jacobians (they cost nothing)."""
    fx, fy, cx, cy = intr
    R = poses[obs_cam, :9].reshape(-1, 3, 3)
    xc = np.einsum("nij,nj->ni", R, pts[obs_pt]) + poses[obs_cam, 9:]
    x, y, z = xc[:, 0], xc[:, 1], xc[:, 2]
    valid = z >= z_min
    iz = np.where(valid, 1.0 / np.where(valid, z, 1.0), 0.0)
    r = np.stack([fx * x * iz + cx - obs_uv[:, 0].astype(np.float64),
                  fy * y * iz + cy - obs_uv[:, 1].astype(np.float64)], 1)
    r[~valid] = 0.0
    s = (r * r).sum(1)
    w, rho = np.ones_like(s), s.copy()
    if loss == "huber":
        out = s > loss_scale * loss_scale
        rs = np.sqrt(np.where(out, s, 1.0))
        rho = np.where(out, 2.0 * loss_scale * rs - loss_scale * loss_scale, s)
        w = np.where(out, loss_scale / rs, 1.0)
    w[~valid], rho[~valid] = 0.0, 0.0
    res = dict(r=r, w=w, rho=rho, valid=valid)
    if jac:
        a0, a2 = fx * iz, -fx * x * iz * iz
        b1, b2 = fy * iz, -fy * y * iz * iz
        n, zero = len(x), np.zeros_like(x)
        drdx = np.zeros((n, 2, 3))
        drdx[:, 0, 0], drdx[:, 0, 2], drdx[:, 1, 1], drdx[:, 1, 2] = a0, a2, b1, b2
        # d x_c / d[dw, dt] = [ -[x_c]x | I ]
        mskew = np.stack([np.stack([zero, z, -y], 1), np.stack([-z, zero, x], 1), np.stack([y, -x, zero], 1)], 1)
        dxdc = np.concatenate([mskew, np.broadcast_to(np.eye(3), (n, 3, 3))], 2)
        Jc = drdx @ dxdc
        Jp = drdx @ R
        Jc[~valid], Jp[~valid] = 0.0, 0.0
        res.update(Jc=Jc, Jp=Jp)
    return res


def _free_index(fixed_cam):
    fidx = np.full(len(fixed_cam), -1, np.int64)
    free = np.flatnonzero(np.asarray(fixed_cam) == 0)
    fidx[free] = np.arange(len(free))
    return fidx, len(free)


def linearize(p, poses, pts, opts):
    """a1 + a2: per-observation terms and the block sums B_c (6x6), g_c, C_p (3x3), g_p."""
    o = observe(poses, pts, p.obs_cam, p.obs_pt, p.obs_uv, p.intr, opts["z_min"], opts["loss"], opts["loss_scale"])
    w = o["w"]
    B = np.zeros((p.nc, 6, 6)); gc = np.zeros((p.nc, 6)); C = np.zeros((p.np_, 3, 3)); gp = np.zeros((p.np_, 3))
    np.add.at(B, p.obs_cam, np.einsum("n,nki,nkj->nij", w, o["Jc"], o["Jc"]))
    np.add.at(gc, p.obs_cam, np.einsum("n,nki,nk->ni", w, o["Jc"], o["r"]))
    np.add.at(C, p.obs_pt, np.einsum("n,nki,nkj->nij", w, o["Jp"], o["Jp"]))
    np.add.at(gp, p.obs_pt, np.einsum("n,nki,nk->ni", w, o["Jp"], o["r"]))
    o.update(B=B, gc=gc, C=C, gp=gp, cost=0.5 * float(o["rho"].sum()), n_invalid=int((~o["valid"]).sum()))
    return o


def compute_step(p, lin, mu, solver="chol", pcg_max_iters=200, pcg_tol=1e-10):
    """a3-a7: damping, landmark elimination, reduced camera solve, back-substitution, model decrease."""
    fidx, nfree = _free_index(p.fixed_cam)
    n = 6 * nfree
    Dc = np.clip(np.einsum("cii->ci", lin["B"]), DIAG_MIN, DIAG_MAX)
    Dp = np.clip(np.einsum("pii->pi", lin["C"]), DIAG_MIN, DIAG_MAX)
    Cd = lin["C"] + np.einsum("pi,ij->pij", Dp / mu, np.eye(3))
    out = dict(Dc=Dc, Dp=Dp, ok=False, pcg_iters=0)
    try:
        np.linalg.cholesky(Cd)                       # every damped landmark block must be SPD
        Cinv = np.linalg.inv(Cd)
    except np.linalg.LinAlgError:
        return out
    # W = sum_k E_k placed at (camera block, landmark block): (n) x (3 Np) sparse; E_k = w J_c^T J_p
    f = fidx[p.obs_cam]
    use = f >= 0
    E = np.einsum("n,nki,nkj->nij", lin["w"], lin["Jc"], lin["Jp"])[use]
    rows = (6 * f[use])[:, None, None] + np.arange(6)[None, :, None] + np.zeros((1, 1, 3), np.int64)
    cols = (3 * p.obs_pt[use].astype(np.int64))[:, None, None] + np.arange(3)[None, None, :] + np.zeros((1, 6, 1), np.int64)
    W = sp.csr_matrix((E.ravel(), (rows.ravel(), cols.ravel())), shape=(n, 3 * p.np_))
    Cinv_bd = sp.bsr_matrix((Cinv, np.arange(p.np_), np.arange(p.np_ + 1)), shape=(3 * p.np_, 3 * p.np_)).tocsr()
    WC = W @ Cinv_bd
    free = np.flatnonzero(fidx >= 0)
    Bdamp = np.stack([lin["B"][c] + np.diag(Dc[c] / mu) for c in free]) if nfree else np.zeros((0, 6, 6))
    Bd = sp.bsr_matrix((Bdamp, np.arange(nfree), np.arange(nfree + 1)), shape=(n, n)).tocsr()
    S = np.asarray((Bd - WC @ W.T).todense()) if n else np.zeros((0, 0))
    gred = lin["gc"][free].reshape(-1) - WC @ lin["gp"].reshape(-1) if n else np.zeros(0)
    out.update(S=S, gred=gred, Cinv=Cinv)
    if n:
        if solver == "pcg":
            dc, it = _pcg(S, -gred, nfree, pcg_max_iters, pcg_tol)
            out["pcg_iters"] = it
            if dc is None:
                return out
        else:
            try:
                L = np.linalg.cholesky(S)
            except np.linalg.LinAlgError:
                return out
            dc = np.linalg.solve(L.T, np.linalg.solve(L, -gred))
    else:
        dc = np.zeros(0)
    if not np.isfinite(dc).all():
        return out
    # delta_p = -C'^-1 (g_p + W^T delta_c)
    t = lin["gp"] + (W.T @ dc).reshape(-1, 3)
    dp = -np.einsum("pij,pj->pi", Cinv, t)
    dDd = float((dc * dc * Dc[free].reshape(-1)).sum() + (dp * dp * Dp).sum())
    gd = float(dc @ lin["gc"][free].reshape(-1) + (dp * lin["gp"]).sum())
    out.update(dc=dc, dp=dp, model_decrease=0.5 * (dDd / mu - gd), ok=True)
    return out


def _pcg(S, b, nfree, max_it, tol):
    """a6: block-Jacobi (6x6) preconditioned CG, x0 = 0, stop at |r| <= tol |b| (checked BEFORE each iteration)."""
    n = 6 * nfree
    try:
        Minv = np.stack([np.linalg.inv(S[6 * f:6 * f + 6, 6 * f:6 * f + 6]) for f in range(nfree)])
        for f in range(nfree):
            np.linalg.cholesky(S[6 * f:6 * f + 6, 6 * f:6 * f + 6])
    except np.linalg.LinAlgError:
        return None, 0
    pre = lambda v: np.einsum("fij,fj->fi", Minv, v.reshape(nfree, 6)).reshape(n)
    x, r = np.zeros(n), b.copy()
    bn = np.linalg.norm(b)
    if bn == 0.0:
        return x, 0
    z = pre(r); q = z.copy(); rz = r @ z
    done = 0
    for _ in range(max_it):
        if np.linalg.norm(r) <= tol * bn:
            break
        Sq = S @ q
        pq = q @ Sq
        if not (pq > 0.0) or not np.isfinite(pq):
            return None, done
        a = rz / pq
        x += a * q; r -= a * Sq
        z = pre(r); rz2 = r @ z
        q = z + (rz2 / rz) * q
        rz = rz2
        done += 1
    return x, done


def apply_step(p, poses, pts, st):
    fidx, _ = _free_index(p.fixed_cam)
    out = poses.copy()
    for c in range(p.nc):
        f = fidx[c]
        if f < 0:
            continue
        d = st["dc"][6 * f:6 * f + 6]
        E = exp_so3(d[:3])
        out[c, :9] = (E @ poses[c, :9].reshape(3, 3)).reshape(9)
        out[c, 9:] = E @ poses[c, 9:] + d[3:]
    return out, pts + st["dp"]


@dataclass
class Summary:
    status: int = 1
    iterations: int = 0
    accepted: int = 0
    rejected: int = 0
    n_invalid: int = 0
    pcg_iters_total: int = 0
    initial_cost: float = 0.0
    final_cost: float = 0.0
    final_mu: float = 0.0
    cost_hist: list = field(default_factory=list)
    mu_hist: list = field(default_factory=list)
    accept_hist: list = field(default_factory=list)


def _opts(loss, loss_scale, z_min):
    return dict(loss=loss, loss_scale=loss_scale, z_min=z_min)


def solve(p, max_iters=10, loss="trivial", loss_scale=1.0, z_min=1e-3, init_mu=1e4, ftol=0.0, gtol=0.0, ptol=0.0,
          solver="chol", pcg_max_iters=200, pcg_tol=1e-10):
    """Full LM (Spec a0).  Returns (poses, points, Summary) like oracle.ba_cpu.solve."""
    opts = _opts(loss, loss_scale, z_min)
    poses, pts = np.array(p.poses, np.float64), np.array(p.points, np.float64)
    sm, mu, nu, lin, cost, ninv = Summary(), init_mu, 2.0, None, 0.0, 0
    it = 0
    while it < max_iters:
        if lin is None:
            lin = linearize(p, poses, pts, opts)
            cost, ninv = lin["cost"], lin["n_invalid"]
            if it == 0:
                sm.initial_cost = cost
            if gtol > 0.0:
                fidx, _ = _free_index(p.fixed_cam)
                gmax = max(np.abs(lin["gc"][fidx >= 0]).max(initial=0.0), np.abs(lin["gp"]).max(initial=0.0))
                if gmax <= gtol:
                    sm.status = 2
                    break
        st = compute_step(p, lin, mu, solver, pcg_max_iters, pcg_tol)
        sm.pcg_iters_total += st["pcg_iters"]
        accept, tcost, rho, converged = False, cost, 0.0, False
        if st["ok"] and st["model_decrease"] > 0.0:
            tposes, tpts = apply_step(p, poses, pts, st)
            t = observe(tposes, tpts, p.obs_cam, p.obs_pt, p.obs_uv, p.intr, z_min, loss, loss_scale, jac=False)
            tcost, ninv_t = 0.5 * float(t["rho"].sum()), int((~t["valid"]).sum())
            rho = (cost - tcost) / st["model_decrease"]
            accept = bool(np.isfinite(tcost) and rho > MIN_RATIO and ninv_t <= ninv)
        sm.mu_hist.append(mu)
        if accept:
            if ftol > 0.0 and abs(cost - tcost) <= ftol * cost:
                converged = True
            if ptol > 0.0:
                dn = np.sqrt(float(st["dc"] @ st["dc"] + (st["dp"] ** 2).sum()))
                if dn <= ptol * (np.sqrt(float((pts ** 2).sum())) + ptol):
                    converged = True
            poses, pts, cost, ninv, lin = tposes, tpts, tcost, ninv_t, None
            tq = 2.0 * rho - 1.0
            mu, nu = min(MU_MAX, mu / max(1.0 / 3.0, 1.0 - tq ** 3)), 2.0
            sm.accepted += 1
        else:
            mu, nu = mu / nu, 2.0 * nu
            sm.rejected += 1
        sm.cost_hist.append(cost); sm.accept_hist.append(int(accept))
        it += 1
        if converged:
            sm.status = 2
            break
        if not mu > MU_MIN:
            sm.status = 3
            break
    sm.iterations, sm.final_cost, sm.final_mu = it, cost, mu
    t = observe(poses, pts, p.obs_cam, p.obs_pt, p.obs_uv, p.intr, z_min, loss, loss_scale, jac=False)
    sm.n_invalid = int((~t["valid"]).sum())
    return poses, pts, sm


def stage(p, mu=1e4, loss="trivial", loss_scale=1.0, z_min=1e-3, solver="chol"):
    """Every intermediate of ONE LM iteration at radius mu, named like oracle.ba_cpu.stage (C as upper (Np,6))."""
    poses, pts = np.asarray(p.poses, np.float64), np.asarray(p.points, np.float64)
    lin = linearize(p, poses, pts, _opts(loss, loss_scale, z_min))
    st = compute_step(p, lin, mu, solver)
    C6 = lin["C"][:, [0, 0, 0, 1, 1, 2], [0, 1, 2, 1, 2, 2]]
    out = dict(r=lin["r"], w=lin["w"], Jc=lin["Jc"], Jp=lin["Jp"], B=lin["B"], gc=lin["gc"], C=C6, gp=lin["gp"],
               cost=lin["cost"], ok=st["ok"], n_invalid=lin["n_invalid"])
    if st["ok"]:
        tposes, tpts = apply_step(p, poses, pts, st)
        t = observe(tposes, tpts, p.obs_cam, p.obs_pt, p.obs_uv, p.intr, z_min, loss, loss_scale, jac=False)
        out.update(S=st["S"], gred=st["gred"], dc=st["dc"], dp=st["dp"], model_decrease=st["model_decrease"],
                   trial_cost=0.5 * float(t["rho"].sum()), tposes=tposes, tpoints=tpts)
    return out
