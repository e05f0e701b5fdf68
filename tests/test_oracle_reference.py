"""Pins the oracle's a1 stage (projection, residual, cheirality, Jacobians) to vectors produced by the
REFERENCE's own triangulate_points / cv2.projectPoints calls (src/modules/frontend.py:115-148),
committed in tests/golden/reference_triangulation.npz (generator: tests/golden/make_reference_fixtures.py).
The LM/Schur stages above a1 have no reference counterpart (parity unpinned)."""
import numpy as np
import pytest

from visualodometry_b200 import synth
from visualodometry_b200.ba import poses_from_T

CASES = ["kitti", "malaga", "parking"]


def _problem(g, name):
    """one camera (T2), every triangulated point observed once at p2 (frontend.py:138-140)."""
    K, T2, X, uv = g[f"{name}_K"], g[f"{name}_T2"], g[f"{name}_all3d"], g[f"{name}_p2"]
    n = X.shape[0]
    return synth.BAProblem(poses_from_T(T2[None]), np.ascontiguousarray(X), np.zeros(n, np.int32),
                           np.arange(n, dtype=np.int32), uv, np.zeros(1, np.uint8),
                           np.array([K[0, 0], K[1, 1], K[0, 2], K[1, 2]]))


@pytest.mark.parametrize("name", CASES)
def test_residual_matches_reference_projection(oracle, golden, name):
    p = _problem(golden, name)
    st = oracle.stage(p, z_min=float(golden[f"{name}_min_depth"]))
    front = golden[f"{name}_depth2"] > float(golden[f"{name}_min_depth"])        # frontend.py:135
    ref_r = golden[f"{name}_proj2"] - golden[f"{name}_p2"].astype(np.float64)    # frontend.py:140
    scale = np.maximum(1.0, np.abs(golden[f"{name}_proj2"]).max(axis=1))
    # the reference pipeline is float32 here (cv2.triangulatePoints returns f32, so projectPoints
    # computes in f32): agreement to a few f32 ulps of the pixel coordinate is exact agreement
    assert golden[f"{name}_proj2"].dtype == np.float32
    assert (np.abs(st["r"][front] - ref_r[front]).max(axis=1) <= 8 * np.finfo(np.float32).eps * scale[front]).all()
    # cheirality: behind-camera observations carry zero weight and zero residual
    assert (st["w"][~front] == 0).all() and (st["r"][~front] == 0).all()
    assert (st["w"][front] == 1).all()


@pytest.mark.parametrize("name", CASES)
def test_gates_reproduce_reference_mask(oracle, golden, name):
    p = _problem(golden, name)
    st = oracle.stage(p, z_min=float(golden[f"{name}_min_depth"]))
    err = np.linalg.norm(st["r"], axis=1)
    mask = (st["w"] > 0) & (err < float(golden[f"{name}_max_reproj_err"]))       # frontend.py:135,143,146
    assert np.array_equal(mask, golden[f"{name}_mask"])


@pytest.mark.parametrize("name", CASES)
def test_translation_jacobian_matches_cv2(oracle, golden, name):
    p = _problem(golden, name)
    st = oracle.stage(p, z_min=float(golden[f"{name}_min_depth"]))
    front = st["w"] > 0
    jac = golden[f"{name}_jac"].reshape(-1, 2, golden[f"{name}_jac"].shape[1])   # cols 0-2 rvec, 3-5 tvec
    ref = jac[:, :, 3:6]
    assert np.abs(st["Jc"][front][:, :, 3:6] - ref[front]).max() <= 1e-5 * np.abs(ref[front]).max()


def test_rotation_jacobian_matches_cv2_at_identity(oracle, golden):
    """With T_cw2 = I (the 'parking' fixture) the left perturbation equals cv2's d/d rvec at rvec=0."""
    name = "parking"
    assert np.allclose(golden[f"{name}_T2"], np.eye(4))
    p = _problem(golden, name)
    st = oracle.stage(p, z_min=float(golden[f"{name}_min_depth"]))
    front = st["w"] > 0
    jac = golden[f"{name}_jac"].reshape(-1, 2, golden[f"{name}_jac"].shape[1])
    ref = jac[:, :, 0:3]
    assert np.abs(st["Jc"][front][:, :, 0:3] - ref[front]).max() <= 1e-5 * np.abs(ref[front]).max()
