"""Generates tests/golden/reference_triangulation.npz by RUNNING THE REFERENCE's own code
(/root/reference/src/modules/frontend.py: triangulate_points, which calls cv2.triangulatePoints and
cv2.projectPoints) on seeded synthetic two-view scenes.  Run in the build container only; the
GPU box never reads /root/reference -- the .npz is committed.

The reference has no BA, so these vectors pin what it DOES define on the hot path's boundary:
projection convention x_c = R_cw X + t_cw, pixel residual, cheirality gate, reprojection gate.
"""
import os
import sys
import types

import numpy as np

REF = "/root/reference/src"
sys.path.insert(0, REF)
# the learned frontend is not installed; triangulate_points does not use it
lg = types.ModuleType("lightglue")
lg.LightGlue = lg.SuperPoint = object
lgu = types.ModuleType("lightglue.utils")
lgu.rbd = lambda x: x
sys.modules.setdefault("lightglue", lg)
sys.modules.setdefault("lightglue.utils", lgu)

import cv2  # noqa: E402
from config.config import get_config  # noqa: E402
from modules.frontend import triangulate_points  # noqa: E402

sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", ".."))
from visualodometry_b200 import synth  # noqa: E402


def scene(seed, cam, identity_second):
    rng = np.random.default_rng(seed)
    fx, fy, cx, cy, W, H = cam
    K = np.array([[fx, 0, cx], [0, fy, cy], [0, 0, 1.0]])
    T1, T2 = np.eye(4), np.eye(4)
    R1 = synth.so3_exp(rng.normal(0, 0.03, 3))
    T1[:3, :3], T1[:3, 3] = R1, rng.normal(0, 0.3, 3) + np.array([0.8, 0, 0])
    if not identity_second:
        T2[:3, :3], T2[:3, 3] = synth.so3_exp(rng.normal(0, 0.03, 3)), rng.normal(0, 0.3, 3) - np.array([0.8, 0, 0.5])
    n = 400
    X = np.stack([rng.uniform(-8, 8, n), rng.uniform(-3, 3, n), rng.uniform(-2, 40, n)], 1)  # some behind the cameras

    def proj(T):
        xc = X @ T[:3, :3].T + T[:3, 3]
        return (xc[:, :2] / xc[:, 2:3]) * [fx, fy] + [cx, cy]

    p1 = (proj(T1) + rng.normal(0, 1.5, (n, 2))).astype(np.float32)
    p2 = (proj(T2) + rng.normal(0, 1.5, (n, 2))).astype(np.float32)
    return K, T1, T2, p1, p2


out = {}
cases = [("kitti", synth.KITTI, False), ("malaga", synth.MALAGA, False), ("parking", synth.PARKING, True)]
for i, (name, cam, ident) in enumerate(cases):
    K, T1, T2, p1, p2 = scene(100 + i, cam, ident)
    cfg = get_config(name)
    pts3d, mask = triangulate_points(T1, T2, p1, p2, K, cfg)          # the reference call
    # the intermediate quantities triangulate_points computes internally (frontend.py:128-143)
    P1, P2 = K @ T1[:3, :], K @ T2[:3, :]
    pts4d = cv2.triangulatePoints(P1, P2, p1.T, p2.T)
    all3d = (pts4d[:3] / pts4d[3]).T
    depth2 = (T2[:3, :3] @ all3d.T + T2[:3, 3:4]).T[:, 2]
    proj2, jac = cv2.projectPoints(all3d, T2[:3, :3], T2[:3, 3], K, None)
    assert np.array_equal(all3d[mask], pts3d)
    out.update({f"{name}_K": K, f"{name}_T1": T1, f"{name}_T2": T2, f"{name}_p1": p1, f"{name}_p2": p2,
                f"{name}_pts3d": pts3d, f"{name}_mask": mask, f"{name}_all3d": all3d, f"{name}_depth2": depth2,
                f"{name}_proj2": proj2.reshape(-1, 2), f"{name}_jac": jac,
                f"{name}_min_depth": cfg.min_depth, f"{name}_max_reproj_err": cfg.max_reproj_err})
np.savez_compressed(os.path.join(os.path.dirname(__file__), "reference_triangulation.npz"), **out)
print("wrote", len(out), "arrays; masks:", [int(out[f"{n}_mask"].sum()) for n, _, _ in cases])
