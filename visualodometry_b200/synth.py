"""Deterministic synthetic bundle-adjustment problems (SURVEY.md §8 row d).

Shapes follow BASELINE.json `configs`; intrinsics are the reference's own
(`src/modules/dataset_loader.py:52-54` KITTI, `:85-87` Malaga).  Pose convention is the
reference's projection convention `x_c = R_cw X + t_cw` (`src/modules/frontend.py:128-139`).
Everything is numpy + `default_rng(seed)` so the same problem is rebuilt bit-identically on
the GPU box (no file travels).
"""
from __future__ import annotations

from dataclasses import dataclass, field

import numpy as np

# fx, fy, cx, cy, width, height
KITTI = (718.856, 718.856, 607.1928, 185.2157, 1241, 376)      # dataset_loader.py:52-54
MALAGA = (621.18428, 621.18428, 404.0076, 309.05989, 800, 600)  # dataset_loader.py:85-87, :96
PARKING = (331.37, 369.568, 320.0, 240.0, 640, 480)             # K.txt not in the reference: synthetic choice


@dataclass
class BAProblem:
    """One BA problem in the engine's array layout (SURVEY §8 row b)."""

    poses: np.ndarray        # (Nc,12) f64 row-major [R_cw(9) | t_cw(3)], perturbed initial guess
    points: np.ndarray       # (Np,3) f64, perturbed initial guess
    obs_cam: np.ndarray      # (No,) i32
    obs_pt: np.ndarray       # (No,) i32
    obs_uv: np.ndarray       # (No,2) f32
    fixed_cam: np.ndarray    # (Nc,) u8
    intr: np.ndarray         # (4,) f64 fx,fy,cx,cy
    gt_poses: np.ndarray = field(default=None, repr=False)
    gt_points: np.ndarray = field(default=None, repr=False)

    @property
    def nc(self): return self.poses.shape[0]
    @property
    def np_(self): return self.points.shape[0]
    @property
    def no(self): return self.obs_cam.shape[0]


def rot_y(a):
    c, s = np.cos(a), np.sin(a)
    return np.array([[c, 0, s], [0, 1, 0], [-s, 0, c]])


def so3_exp(w):
    """Rodrigues; same branch rule as Spec a0 (theta^2 < 1e-16 -> series)."""
    w = np.asarray(w, dtype=np.float64)
    t2 = float(w @ w)
    if t2 < 1e-16:
        a, b = 1.0 - t2 / 6.0, 0.5 - t2 / 24.0
    else:
        t = np.sqrt(t2)
        a, b = np.sin(t) / t, (1.0 - np.cos(t)) / t2
    W = np.array([[0, -w[2], w[1]], [w[2], 0, -w[0]], [-w[1], w[0], 0]])
    return np.eye(3) + a * W + b * (W @ W)


def _pose_cw(R_wc, C):
    R = R_wc.T
    return np.concatenate([R.reshape(9), -R @ C])


def _project(poses, X, intr):
    """poses (Nc,12), X (Np,3) -> u,v,z each (Nc,Np)."""
    R = poses[:, :9].reshape(-1, 3, 3)
    t = poses[:, 9:]
    xc = np.einsum("cij,pj->cpi", R, X) + t[:, None, :]
    z = xc[..., 2]
    with np.errstate(divide="ignore", invalid="ignore"):
        u = intr[0] * xc[..., 0] / z + intr[2]
        v = intr[1] * xc[..., 1] / z + intr[3]
    return u, v, z


def _perturb(rng, gt_poses, gt_points, ref_pos, sig_rot=0.01, sig_t=0.05, sig_pt=0.02):
    poses = gt_poses.copy()
    for c in range(1, poses.shape[0]):
        dR = so3_exp(rng.normal(0, sig_rot, 3))
        R = dR @ poses[c, :9].reshape(3, 3)
        t = dR @ poses[c, 9:] + rng.normal(0, sig_t, 3)
        poses[c, :9], poses[c, 9:] = R.reshape(9), t
    depth = np.linalg.norm(gt_points - ref_pos[None, :], axis=1)
    points = gt_points + rng.normal(0, 1, gt_points.shape) * (sig_pt * depth)[:, None]
    return poses, points


def make_window(seed=0, n_cams=10, n_pts=2000, n_obs=15000, cam=KITTI, noise_px=0.5,
                outlier_frac=0.0, outlier_px=20.0, spacing=1.0):
    """C1 / C3 / one C4 unit: forward-moving window, KITTI-05-shaped by default."""
    rng = np.random.default_rng(seed)
    fx, fy, cx, cy, W, H = cam
    intr = np.array([fx, fy, cx, cy])
    gt_poses = np.empty((n_cams, 12))
    for c in range(n_cams):
        yaw = 0.0 if c == 0 else rng.normal(0, 0.02)
        gt_poses[c] = _pose_cw(rot_y(yaw), np.array([0.0, 0.0, c * spacing]))
    z_mid = (n_cams // 2) * spacing

    def draw(n):
        return np.stack([rng.uniform(-15, 15, n), rng.uniform(-3, 2, n),
                         z_mid + rng.uniform(5, 60, n)], axis=1)

    X = draw(n_pts)
    for _ in range(200):
        u, v, z = _project(gt_poses, X, intr)
        vis = (z > 1.0) & (u >= 0) & (u < W) & (v >= 0) & (v < H)
        bad = vis.sum(0) < 2
        if not bad.any():
            break
        X[bad] = draw(int(bad.sum()))
    else:
        raise RuntimeError("could not place all points in >=2 views")
    # protect the first two views of each point, thin the rest down to ~n_obs
    order = np.cumsum(vis, axis=0)
    protect = vis & (order <= 2)
    extra = vis & ~protect
    want = n_obs - int(protect.sum())
    n_extra = int(extra.sum())
    if 0 <= want < n_extra:
        idx = np.flatnonzero(extra.ravel())
        drop = rng.permutation(n_extra)[: n_extra - want]
        keep = extra.ravel().copy()
        keep[idx[drop]] = False
        extra = keep.reshape(extra.shape)
    keep = protect | extra
    cam_i, pt_i = np.nonzero(keep)          # camera-major generation order
    uv = np.stack([u[cam_i, pt_i], v[cam_i, pt_i]], axis=1) + rng.normal(0, noise_px, (cam_i.size, 2))
    if outlier_frac > 0:
        out = rng.random(cam_i.size) < outlier_frac
        uv[out] += rng.normal(0, outlier_px, (int(out.sum()), 2))
    poses, points = _perturb(rng, gt_poses, X, np.array([0.0, 0.0, z_mid]))
    fixed = np.zeros(n_cams, np.uint8)
    fixed[0] = 1
    return BAProblem(poses, points, cam_i.astype(np.int32), pt_i.astype(np.int32),
                     uv.astype(np.float32), fixed, intr, gt_poses, X)


@dataclass
class BABatch:
    """B independent problems packed with CSR offsets (C4; SURVEY §8 row b batched form).
    obs_cam / obs_pt are window-LOCAL indices."""

    poses: np.ndarray      # (sum Nc,12)
    points: np.ndarray     # (sum Np,3)
    obs_cam: np.ndarray    # (sum No,) i32 local
    obs_pt: np.ndarray     # (sum No,) i32 local
    obs_uv: np.ndarray     # (sum No,2) f32
    fixed_cam: np.ndarray  # (sum Nc,) u8
    intr: np.ndarray       # (4,)
    cam_off: np.ndarray    # (B+1,) i32
    pt_off: np.ndarray     # (B+1,) i32
    obs_off: np.ndarray    # (B+1,) i64

    @property
    def nb(self): return self.cam_off.shape[0] - 1
    @property
    def no(self): return int(self.obs_off[-1])

    def window(self, b) -> BAProblem:
        c0, c1 = self.cam_off[b], self.cam_off[b + 1]
        p0, p1 = self.pt_off[b], self.pt_off[b + 1]
        o0, o1 = self.obs_off[b], self.obs_off[b + 1]
        return BAProblem(self.poses[c0:c1], self.points[p0:p1], self.obs_cam[o0:o1],
                         self.obs_pt[o0:o1], self.obs_uv[o0:o1], self.fixed_cam[c0:c1], self.intr)


def pack_batch(problems) -> BABatch:
    cam_off = np.zeros(len(problems) + 1, np.int32)
    pt_off = np.zeros(len(problems) + 1, np.int32)
    obs_off = np.zeros(len(problems) + 1, np.int64)
    for i, p in enumerate(problems):
        cam_off[i + 1] = cam_off[i] + p.nc
        pt_off[i + 1] = pt_off[i] + p.np_
        obs_off[i + 1] = obs_off[i] + p.no
    cat = lambda k: np.ascontiguousarray(np.concatenate([getattr(p, k) for p in problems]))
    return BABatch(cat("poses"), cat("points"), cat("obs_cam"), cat("obs_pt"), cat("obs_uv"),
                   cat("fixed_cam"), problems[0].intr.copy(), cam_off, pt_off, obs_off)


def make_batch(n_windows=4096, seed0=0, distinct=None, **kw) -> BABatch:
    """C4: `n_windows` independent C1-shaped windows, seeds seed0..seed0+n-1.
    `distinct` (optional) caps the number of distinct geometries that are generated; the
    remaining windows reuse a geometry with FRESH pixel noise and a fresh initial perturbation
    (so no two windows are identical) -- used only to bound host set-up time in benches."""
    if distinct is None or distinct >= n_windows:
        return pack_batch([make_window(seed=seed0 + i, **kw) for i in range(n_windows)])
    base = [make_window(seed=seed0 + i, **kw) for i in range(distinct)]
    out = []
    for i in range(n_windows):
        b = base[i % distinct]
        if i < distinct:
            out.append(b)
            continue
        rng = np.random.default_rng(seed0 + i)
        uv = (b.obs_uv.astype(np.float64) + rng.normal(0, 0.25, b.obs_uv.shape)).astype(np.float32)
        poses, points = _perturb(rng, b.gt_poses, b.gt_points, b.gt_points.mean(0))
        out.append(BAProblem(poses, points, b.obs_cam, b.obs_pt, uv, b.fixed_cam, b.intr,
                             b.gt_poses, b.gt_points))
    return pack_batch(out)


def make_large(seed=0, n_cams=1000, n_pts=200_000, k=5, mode="local", cam=KITTI, noise_px=0.5,
               path_len=None):
    """C5: BAL-style problem, exactly k observations per point, observation order shuffled.
    mode 'local' : cameras on a straight path along +x (1 m spacing unless path_len is given) looking
                   sideways (+z); each point is seen by the k cameras nearest to it along the path
                   (+- jitter) -> block-banded reduced camera system S.
    mode 'random': cameras on a ring of radius 40 m looking at the centre, points in a 16 m cube at
                   the centre, each point seen by ANY k cameras -> dense S (BAL-style stress)."""
    rng = np.random.default_rng(seed)
    fx, fy, cx, cy, W, H = cam
    intr = np.array([fx, fy, cx, cy])
    gt_poses = np.empty((n_cams, 12))
    if mode == "local":
        path_len = float(n_cams) if path_len is None else float(path_len)
        step = path_len / n_cams
        for c in range(n_cams):
            yaw = 0.0 if c == 0 else rng.normal(0, 0.02)
            gt_poses[c] = _pose_cw(rot_y(yaw), np.array([c * step, 0.0, 0.0]))
        X = np.stack([rng.uniform(0, path_len, n_pts), rng.uniform(-3, 2, n_pts),
                      rng.uniform(5, 60, n_pts)], axis=1)
        centre = np.rint(X[:, 0] / step).astype(np.int64) + rng.integers(-2, 3, n_pts)
        centre = np.clip(centre, k // 2, n_cams - 1 - (k - k // 2 - 1))
        cams = centre[:, None] + (np.arange(k) - k // 2)[None, :]
        depth_ref = X[:, 2]
    elif mode == "random":
        radius = 40.0
        for c in range(n_cams):
            th = 2 * np.pi * c / n_cams
            C = np.array([radius * np.sin(th), 0.0, -radius * np.cos(th)])
            # camera z axis points at the origin: R_wc = rot_y(-th) has z_cam = (-sin th, 0, cos th)
            gt_poses[c] = _pose_cw(rot_y(-th) @ rot_y(0.0 if c == 0 else rng.normal(0, 0.02)), C)
        X = rng.uniform(-8, 8, (n_pts, 3)) * np.array([1.0, 0.25, 1.0])
        # k distinct cameras per point: random keys, take the k smallest
        cams = np.argsort(rng.random((n_pts, n_cams), dtype=np.float32), axis=1)[:, :k]
        cams.sort(axis=1)
        depth_ref = np.full(n_pts, radius)
    else:
        raise ValueError(mode)
    pt_i = np.repeat(np.arange(n_pts), k)
    cam_i = cams.reshape(-1)
    R = gt_poses[cam_i, :9].reshape(-1, 3, 3)
    xc = np.einsum("nij,nj->ni", R, X[pt_i]) + gt_poses[cam_i, 9:]
    assert (xc[:, 2] > 1.0).all()
    uv = np.stack([fx * xc[:, 0] / xc[:, 2] + cx, fy * xc[:, 1] / xc[:, 2] + cy], axis=1)
    uv += rng.normal(0, noise_px, uv.shape)
    # shuffle observation order: the engine's index builder must restore point-major order
    perm = rng.permutation(cam_i.size)
    poses, _ = _perturb(rng, gt_poses, X, np.zeros(3), sig_pt=0.0)
    points = X + rng.normal(0, 1, X.shape) * (0.02 * depth_ref)[:, None]
    fixed = np.zeros(n_cams, np.uint8)
    fixed[0] = 1
    return BAProblem(poses, points, cam_i[perm].astype(np.int32), pt_i[perm].astype(np.int32),
                     uv[perm].astype(np.float32), fixed, intr, gt_poses, X)


def rmse_px(p: BAProblem, poses=None, points=None):
    """Reprojection RMSE in pixels over all observations (frontend.py:140 metric)."""
    poses = p.poses if poses is None else poses
    points = p.points if points is None else points
    R = poses[p.obs_cam, :9].reshape(-1, 3, 3)
    xc = np.einsum("nij,nj->ni", R, points[p.obs_pt]) + poses[p.obs_cam, 9:]
    du = p.intr[0] * xc[:, 0] / xc[:, 2] + p.intr[2] - p.obs_uv[:, 0].astype(np.float64)
    dv = p.intr[1] * xc[:, 1] / xc[:, 2] + p.intr[3] - p.obs_uv[:, 1].astype(np.float64)
    return float(np.sqrt(np.mean(du * du + dv * dv)))
