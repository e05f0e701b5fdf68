"""Synthetic KITTI-05-shaped monocular sequence + id-matched synthetic frontend (BASELINE config 2).

The learned / SIFT frontend of the reference (`src/modules/frontend.py:9-112`) is out of scope;
this module stands in for it with the same two calls the VO loop makes:

* ``process_image(img) -> feats`` with ``feats["keypoints"]`` a ``(1,N,2)`` float32 torch tensor
  (`frontend.py:66-70`); the "image" is a 1-element array carrying the frame index,
* ``match_frames(feats0, feats1) -> (M,2) int`` index pairs (`frontend.py:75-112`).

World: random 3D points in a corridor around a planar closed loop (KITTI camera axes: x right, y down,
z forward), 0.8 m per frame, <= 2048 visible points per frame with persistent ids, pixel noise N(0, 0.5 px),
5 % wrong matches.  Everything is seeded and numpy-only.
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np

try:
    import torch
except Exception:  # pragma: no cover
    torch = None

KITTI_K = np.array([[718.856, 0.0, 607.1928], [0.0, 718.856, 185.2157], [0.0, 0.0, 1.0]])   # dataset_loader.py:52-54
KITTI_SIZE = (1241, 376)


def _loop_path(n_frames: int, step: float, side: float = 400.0, radius: float = 40.0):
    """Planar rounded-rectangle loop sampled every `step` metres: positions (n,3) and yaw (n,)."""
    segs = []                      # (kind, length, curvature)
    straight = side - 2 * radius
    for _ in range(4):
        segs.append((straight, 0.0))
        segs.append((0.5 * np.pi * radius, 1.0 / radius))
    per = sum(s[0] for s in segs)
    pos = np.zeros((n_frames, 3))
    yaw = np.zeros(n_frames)
    x = z = th = 0.0
    si, left = 0, segs[0][0]
    for i in range(n_frames):
        pos[i] = (x, 0.0, z)
        yaw[i] = th
        d = step
        while d > 0:
            adv = min(d, left)
            k = segs[si][1]
            if k == 0.0:
                x += adv * np.sin(th); z += adv * np.cos(th)
            else:
                th2 = th + adv * k
                x += (np.cos(th) - np.cos(th2)) / k
                z += (np.sin(th2) - np.sin(th)) / k
                th = th2
            d -= adv; left -= adv
            if left <= 1e-12:
                si = (si + 1) % len(segs); left = segs[si][0]
    return pos, yaw, per


def _R_wc(yaw):
    """camera->world rotation for heading `yaw` about the y (down) axis."""
    c, s = np.cos(yaw), np.sin(yaw)
    return np.array([[c, 0.0, s], [0.0, 1.0, 0.0], [-s, 0.0, c]])


@dataclass
class Frame:
    uv: np.ndarray      # (N,2) float32 noisy pixels
    wid: np.ndarray     # (N,) int64 world point ids (sorted)


class SynthSequence:
    """Ground-truth trajectory, world points and per-frame observations."""

    def __init__(self, n_frames=2700, step=0.8, seed=0, max_kp=2048, noise_px=0.5, pts_per_frame=60,
                 K=KITTI_K, size=KITTI_SIZE):
        self.K, self.size, self.n_frames = np.asarray(K, np.float64), size, n_frames
        rng = np.random.default_rng(seed)
        pos, yaw, _ = _loop_path(n_frames, step)
        self.T_wc = np.tile(np.eye(4), (n_frames, 1, 1))
        for i in range(n_frames):
            self.T_wc[i, :3, :3] = _R_wc(yaw[i])
            self.T_wc[i, :3, 3] = pos[i]
        # world points: born ahead of each frame (camera frame: x lateral, y height, z depth), ids ordered by birth
        loc = np.stack([rng.uniform(-25, 25, (n_frames, pts_per_frame)), rng.uniform(-3, 2, (n_frames, pts_per_frame)),
                        rng.uniform(25, 65, (n_frames, pts_per_frame))], axis=2)
        self.points = np.einsum("nij,nkj->nki", self.T_wc[:, :3, :3], loc) + self.T_wc[:, None, :3, 3]
        self.points = self.points.reshape(-1, 3)
        self.birth = np.repeat(np.arange(n_frames), pts_per_frame)
        self.ppf = pts_per_frame
        self.frames = []
        fx, fy, cx, cy = self.K[0, 0], self.K[1, 1], self.K[0, 2], self.K[1, 2]
        back = int(np.ceil(70.0 / step))
        for i in range(n_frames):
            lo, hi = max(0, i - back) * pts_per_frame, (i + 1) * pts_per_frame
            Xw = self.points[lo:hi]
            R, t = self.T_wc[i, :3, :3], self.T_wc[i, :3, 3]
            Xc = (Xw - t) @ R                       # R_cw (Xw - t), R_cw = R^T
            z = Xc[:, 2]
            ok = z > 2.0
            u = fx * Xc[:, 0] / np.where(ok, z, 1.0) + cx
            v = fy * Xc[:, 1] / np.where(ok, z, 1.0) + cy
            ok &= (u >= 0) & (u < size[0]) & (v >= 0) & (v < size[1])
            ids = lo + np.flatnonzero(ok)
            if ids.size > max_kp:
                ids = ids[-max_kp:]                  # keep the newest (farthest ahead) ones
            uv = np.stack([u[ids - lo], v[ids - lo]], axis=1) + rng.normal(0.0, noise_px, (ids.size, 2))
            self.frames.append(Frame(uv.astype(np.float32), ids.astype(np.int64)))

    def image(self, i: int) -> np.ndarray:
        """The stand-in "image" handed to process_frame: carries the frame index only."""
        return np.array([[i]], dtype=np.int32)


class SynthFrontend:
    """Drop-in for FeatureFrontend (`frontend.py:9`): same two calls, id-matched keypoints."""

    def __init__(self, seq: SynthSequence, outlier_frac=0.05, seed=1):
        self.seq, self.outlier_frac, self.seed = seq, outlier_frac, seed

    def process_image(self, img) -> dict:
        i = int(np.asarray(img).reshape(-1)[0])
        fr = self.seq.frames[i]
        kp = torch.from_numpy(fr.uv).unsqueeze(0) if torch is not None else fr.uv[None]
        return {"keypoints": kp, "wid": fr.wid, "frame": i}

    def match_frames(self, feats0: dict, feats1: dict) -> np.ndarray:
        _, i0, i1 = np.intersect1d(feats0["wid"], feats1["wid"], assume_unique=True, return_indices=True)
        m = np.stack([i0, i1], axis=1).astype(int)
        if self.outlier_frac > 0 and len(m) > 0:
            rng = np.random.default_rng((self.seed, feats0["frame"], feats1["frame"]))
            bad = rng.random(len(m)) < self.outlier_frac
            m[bad, 1] = rng.integers(0, len(feats1["wid"]), int(bad.sum()))
        return m


def ate_sim3(est: np.ndarray, gt: np.ndarray) -> float:
    """RMS absolute trajectory error after a similarity (Umeyama) alignment of est (n,3) to gt (n,3)."""
    mu_e, mu_g = est.mean(0), gt.mean(0)
    E, G = est - mu_e, gt - mu_g
    U, D, Vt = np.linalg.svd(G.T @ E / len(est))
    S = np.eye(3)
    if np.linalg.det(U) * np.linalg.det(Vt) < 0:
        S[2, 2] = -1
    R = U @ S @ Vt
    var_e = (E ** 2).sum() / len(est)
    s = np.trace(np.diag(D) @ S) / var_e if var_e > 0 else 1.0
    al = (s * (R @ E.T)).T + mu_g
    return float(np.sqrt(((al - gt) ** 2).sum(1).mean()))
