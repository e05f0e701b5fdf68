"""Oracle (oracle/ba_cpu.cpp) self-consistency on CPU: finite differences, dense normal equations,
scipy cross-solver, convergence to ground truth, index canonical order, Huber, gauge, edge cases.
(SURVEY §7.4 test pyramid; the reference brings no tests of its own.)"""
import numpy as np
import pytest

from visualodometry_b200 import synth


def _resid(p, poses, pts):
    R = poses[p.obs_cam, :9].reshape(-1, 3, 3)
    xc = np.einsum("nij,nj->ni", R, pts[p.obs_pt]) + poses[p.obs_cam, 9:]
    return np.stack([p.intr[0] * xc[:, 0] / xc[:, 2] + p.intr[2] - p.obs_uv[:, 0],
                     p.intr[1] * xc[:, 1] / xc[:, 2] + p.intr[3] - p.obs_uv[:, 1]], 1)


@pytest.fixture(scope="module")
def small():
    return synth.make_window(3, n_cams=4, n_pts=50, n_obs=150)


def test_jacobians_finite_differences(oracle, small):
    p, st, h = small, oracle.stage(small), 1e-6
    for k in range(6):
        d = np.zeros(6); d[k] = h

        def pert(sgn):
            P = p.poses.copy()
            for c in range(p.nc):
                E = synth.so3_exp(sgn * d[:3])
                P[c, :9] = (E @ p.poses[c, :9].reshape(3, 3)).reshape(9)
                P[c, 9:] = E @ p.poses[c, 9:] + sgn * d[3:]
            return _resid(p, P, p.points)
        fd = (pert(1) - pert(-1)) / (2 * h)
        assert np.abs(fd - st["Jc"][:, :, k]).max() <= 1e-7 * np.abs(st["Jc"]).max()
    for k in range(3):
        X1, X2 = p.points.copy(), p.points.copy()
        X1[:, k] += h; X2[:, k] -= h
        fd = (_resid(p, p.poses, X1) - _resid(p, p.poses, X2)) / (2 * h)
        assert np.abs(fd - st["Jp"][:, :, k]).max() <= 1e-7 * np.abs(st["Jp"]).max()


def test_schur_step_equals_dense_normal_equations(oracle, small):
    p, mu = small, 1e4
    st = oracle.stage(p, mu=mu)
    free = [c for c in range(p.nc) if not p.fixed_cam[c]]
    fi = {c: i for i, c in enumerate(free)}
    nc6 = 6 * len(free)
    J = np.zeros((2 * p.no, nc6 + 3 * p.np_))
    for i in range(p.no):
        c, q = p.obs_cam[i], p.obs_pt[i]
        if c in fi:
            J[2 * i:2 * i + 2, 6 * fi[c]:6 * fi[c] + 6] = st["Jc"][i]
        J[2 * i:2 * i + 2, nc6 + 3 * q:nc6 + 3 * q + 3] = st["Jp"][i]
    r = st["r"].reshape(-1)
    H, g = J.T @ J, J.T @ r
    D = np.clip(np.diag(H), 1e-6, 1e32)
    delta = np.linalg.solve(H + np.diag(D) / mu, -g)
    assert np.abs(delta[:nc6] - st["dc"]).max() <= 1e-9 * np.abs(st["dc"]).max()
    assert np.abs(delta[nc6:].reshape(-1, 3) - st["dp"]).max() <= 1e-9 * np.abs(st["dp"]).max()
    md = 0.5 * (delta @ (D * delta) / mu - g @ delta)
    assert abs(md - st["model_decrease"]) <= 1e-10 * abs(md)
    assert abs(0.5 * r @ r - st["cost"]) <= 1e-13 * st["cost"]
    # S is the Schur complement of the damped H
    Hd = H + np.diag(D) / mu
    S = Hd[:nc6, :nc6] - Hd[:nc6, nc6:] @ np.linalg.solve(Hd[nc6:, nc6:], Hd[nc6:, :nc6])
    assert np.abs(S - st["S"]).max() <= 1e-10 * np.abs(S).max()


def test_converges_and_matches_scipy(oracle):
    from scipy.optimize import least_squares
    p = synth.make_window(6, n_cams=4, n_pts=60, n_obs=220)
    poses, pts, sm = oracle.solve(p, max_iters=50, ftol=1e-14)
    assert sm.final_cost < sm.initial_cost * 1e-2 and sm.status == 2
    assert sm.n_invalid == 0      # no observation escaped through the cheirality gate
    free = [c for c in range(p.nc) if not p.fixed_cam[c]]

    def unpack(x, base_poses, base_pts):
        P = base_poses.copy()
        for i, c in enumerate(free):
            d = x[6 * i:6 * i + 6]
            E = synth.so3_exp(d[:3])
            P[c, :9] = (E @ base_poses[c, :9].reshape(3, 3)).reshape(9)
            P[c, 9:] = E @ base_poses[c, 9:] + d[3:]
        return P, base_pts + x[6 * len(free):].reshape(-1, 3)

    nx = 6 * len(free) + 3 * p.np_
    # (1) an independent solver started from the same initial guess must not find a lower minimum
    res = least_squares(lambda x: _resid(p, *unpack(x, p.poses, p.points)).reshape(-1), np.zeros(nx),
                        method="trf", x_scale="jac", xtol=1e-15, ftol=1e-15, gtol=1e-12, max_nfev=300)
    assert sm.final_cost <= res.cost * (1 + 1e-6)
    # (2) started AT the oracle's solution it must not be able to improve it: it is a minimum
    res2 = least_squares(lambda x: _resid(p, *unpack(x, poses, pts)).reshape(-1), np.zeros(nx),
                         method="trf", x_scale="jac", xtol=1e-15, ftol=1e-15, gtol=1e-12, max_nfev=100)
    assert abs(0.5 * (_resid(p, poses, pts) ** 2).sum() - sm.final_cost) <= 1e-12 * sm.final_cost
    assert res2.cost >= sm.final_cost * (1 - 1e-6)


def test_noise_free_recovers_zero_cost(oracle):
    p = synth.make_window(6, n_cams=5, n_pts=80, n_obs=300, noise_px=0.0)
    poses, pts, sm = oracle.solve(p, max_iters=40)
    # observations are f32-rounded pixels: the floor is ~ (eps_f32 * 1e3 px)^2 per residual
    assert sm.final_cost / p.no < 1e-6
    assert synth.rmse_px(p, poses, pts) < 2e-3


def test_huber_weights_and_cost(oracle):
    p = synth.make_window(2, n_cams=5, n_pts=100, n_obs=400, outlier_frac=0.2)
    st = oracle.stage(p, loss="huber", loss_scale=1.0)
    s = (st["r"] ** 2).sum(1)
    w = np.where(s <= 1.0, 1.0, 1.0 / np.sqrt(np.maximum(s, 1e-300)))
    assert np.abs(st["w"] - w).max() < 1e-14
    rho = np.where(s <= 1.0, s, 2 * np.sqrt(s) - 1)
    assert abs(0.5 * rho.sum() - st["cost"]) <= 1e-12 * st["cost"]


def test_index_canonical_order(oracle):
    p = synth.make_large(1, n_cams=30, n_pts=400, k=4, mode="random")
    ix = oracle.index(p)
    key = p.obs_pt[ix.pm_perm].astype(np.int64) * p.nc + p.obs_cam[ix.pm_perm]
    assert (np.diff(key) > 0).all()                     # strictly increasing: (pt, cam) unique
    key = p.obs_cam[ix.cm_perm].astype(np.int64) * p.np_ + p.obs_pt[ix.cm_perm]
    assert (np.diff(key) > 0).all()
    assert np.array_equal(ix.pt_ptr, np.concatenate([[0], np.cumsum(np.bincount(p.obs_pt, minlength=p.np_))]))
    assert np.array_equal(ix.cam_ptr, np.concatenate([[0], np.cumsum(np.bincount(p.obs_cam, minlength=p.nc))]))
    assert ix.fidx[0] == -1 and np.array_equal(ix.fidx[1:], np.arange(p.nc - 1))
    # block structure: every co-visible free pair present, diagonals present, sorted
    pairs = set()
    for q in range(p.np_):
        cams = sorted(ix.fidx[p.obs_cam[p.obs_pt == q]])
        cams = [c for c in cams if c >= 0]
        pairs.update((a, b) for i, a in enumerate(cams) for b in cams[i:])
    pairs.update((f, f) for f in range(p.nc - 1))
    got = [(f, int(c)) for f in range(p.nc - 1) for c in ix.blk_col[ix.blk_ptr[f]:ix.blk_ptr[f + 1]]]
    assert got == sorted(pairs)


def test_rejects_bad_input(oracle):
    p = synth.make_window(3, n_cams=4, n_pts=50, n_obs=150)
    bad = synth.BAProblem(p.poses, p.points, p.obs_cam.copy(), p.obs_pt.copy(), p.obs_uv, p.fixed_cam, p.intr)
    bad.obs_cam[0] = 99
    with pytest.raises(ValueError):
        oracle.solve(bad)
    dup = synth.BAProblem(p.poses, p.points, np.append(p.obs_cam, p.obs_cam[0]), np.append(p.obs_pt, p.obs_pt[0]),
                          np.vstack([p.obs_uv, p.obs_uv[:1]]), p.fixed_cam, p.intr)
    with pytest.raises(ValueError):
        oracle.solve(dup)


def test_empty_and_ragged(oracle):
    p = synth.make_window(3, n_cams=4, n_pts=50, n_obs=150)
    # a landmark with no observation and a camera with none: both must stay put, solve must run
    pts = np.vstack([p.points, [[1.0, 2.0, 30.0]]])
    poses = np.vstack([p.poses, p.poses[-1:]])
    q = synth.BAProblem(poses, pts, p.obs_cam, p.obs_pt, p.obs_uv, np.append(p.fixed_cam, 0).astype(np.uint8), p.intr)
    po, xo, sm = oracle.solve(q, max_iters=5)
    assert np.array_equal(xo[-1], pts[-1]) and np.allclose(po[-1], poses[-1], atol=1e-12)
    assert sm.final_cost < sm.initial_cost


def test_pcg_matches_cholesky(oracle):
    # random visibility: S is a well-connected graph and block-Jacobi PCG converges to 1e-10
    p = synth.make_large(2, n_cams=40, n_pts=2000, k=5, mode="random", path_len=40.0)
    a = oracle.solve(p, max_iters=8, solver="chol")
    b = oracle.solve(p, max_iters=8, solver="pcg")
    assert np.abs(a[0] - b[0]).max() < 1e-7 and abs(a[2].final_cost - b[2].final_cost) <= 1e-10 * a[2].final_cost
    assert b[2].pcg_iters_total > 0


def test_batch_equals_single(oracle):
    b = synth.make_batch(3, n_cams=4, n_pts=60, n_obs=200)
    po, xo, sms = oracle.solve_batch(b, max_iters=5, threads=2)
    for i in range(b.nb):
        w = b.window(i)
        p1, x1, s1 = oracle.solve(w, max_iters=5)
        assert np.array_equal(po[b.cam_off[i]:b.cam_off[i + 1]], p1)
        assert np.array_equal(xo[b.pt_off[i]:b.pt_off[i + 1]], x1)
        assert sms[i].final_cost == s1.final_cost
