"""Motion-only BA: oracle/pose_numpy.py against ground truth and OpenCV's own PnP refinement
(CPU), and b2ba_refine_poses against the oracle (GPU)."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import pose_numpy  # noqa: E402

KITTI = np.array([718.856, 718.856, 607.1928, 185.2157])      # dataset_loader.py:52-54


def make_pose_problem(seed, n=300, noise=0.5, outliers=0.0, behind=0, rot=0.02, trans=0.3):
    """A camera looking down +z at n points; returns (T_gt, T_init, X, uv)."""
    rng = np.random.default_rng(seed)
    w = rng.normal(0, 0.2, 3)
    T_gt = np.eye(4)
    T_gt[:3, :3] = pose_numpy._exp_so3(w)
    T_gt[:3, 3] = rng.normal(0, 1.0, 3)
    xc = np.stack([rng.uniform(-15, 15, n), rng.uniform(-3, 2, n), rng.uniform(5, 60, n)], axis=1)
    X = (xc - T_gt[:3, 3]) @ T_gt[:3, :3]                    # R^T (x_c - t)
    uv = np.stack([KITTI[0] * xc[:, 0] / xc[:, 2] + KITTI[2], KITTI[1] * xc[:, 1] / xc[:, 2] + KITTI[3]], axis=1)
    uv += rng.normal(0, noise, uv.shape)
    bad = rng.random(n) < outliers
    uv[bad] += rng.normal(0, 30.0, (int(bad.sum()), 2))
    if behind:                                                # landmarks behind the camera: cheirality gate
        X[:behind] = (np.stack([rng.uniform(-5, 5, behind), rng.uniform(-2, 2, behind), rng.uniform(-30, -5, behind)], axis=1)
                      - T_gt[:3, 3]) @ T_gt[:3, :3]
    T0 = np.eye(4)
    T0[:3, :3] = pose_numpy._exp_so3(rng.normal(0, rot, 3)) @ T_gt[:3, :3]
    T0[:3, 3] = T_gt[:3, 3] + rng.normal(0, trans, 3)
    return T_gt, T0, X, uv.astype(np.float32)


def pose_err(Ta, Tb):
    return max(np.abs(Ta[:3, :3] - Tb[:3, :3]).max(), np.abs(Ta[:3, 3] - Tb[:3, 3]).max())


# ------------------------------------------------------------------------------------------------ CPU: the oracle
def test_oracle_recovers_ground_truth_noise_free():
    T_gt, T0, X, uv = make_pose_problem(0, noise=0.0)
    T, sm = pose_numpy.refine_pose(T0, X, uv, KITTI, max_iters=20)
    assert sm.final_cost < 1e-6 * sm.initial_cost          # uv are float32: ~1e-5 px residuals remain
    assert pose_err(T, T_gt) < 1e-6


def test_oracle_minimiser_agrees_with_opencv_pnp_refinement():
    """cv2.solvePnP(ITERATIVE, useExtrinsicGuess) is the refinement cv2.solvePnPRansac ends with (vo.py:131): the
    converged pose of the same reprojection problem must agree."""
    import cv2
    T_gt, T0, X, uv = make_pose_problem(1, noise=0.5)
    T, sm = pose_numpy.refine_pose(T0, X, uv, KITTI, max_iters=30, ftol=1e-14)
    K = np.array([[KITTI[0], 0, KITTI[2]], [0, KITTI[1], KITTI[3]], [0, 0, 1.0]])
    rvec, _ = cv2.Rodrigues(T0[:3, :3])
    ok, rvec, tvec = cv2.solvePnP(X, uv.astype(np.float64), K, None, rvec.copy(), T0[:3, 3].reshape(3, 1).copy(),
                                  useExtrinsicGuess=True, flags=cv2.SOLVEPNP_ITERATIVE)
    assert ok
    Tcv = np.eye(4)
    Tcv[:3, :3], Tcv[:3, 3] = cv2.Rodrigues(rvec)[0], tvec.ravel()
    assert pose_err(T, Tcv) < 1e-4
    # and our minimum is at least as low as OpenCV's
    _, _, _, _, rho_cv, _ = pose_numpy._terms(Tcv[:3, :3], Tcv[:3, 3], X, uv, KITTI, 1e-3, "trivial", 1.0)
    assert sm.final_cost <= 0.5 * rho_cv.sum() * (1 + 1e-9)


def test_oracle_huber_resists_outliers_and_gates_cheirality():
    T_gt, T0, X, uv = make_pose_problem(2, noise=0.5, outliers=0.15, behind=7)
    T_l2, _ = pose_numpy.refine_pose(T0, X, uv, KITTI, max_iters=20)
    T_h, sm = pose_numpy.refine_pose(T0, X, uv, KITTI, max_iters=20, loss="huber", loss_scale=2.0)
    assert sm.n_invalid == 7
    assert pose_err(T_h, T_gt) < 0.5 * pose_err(T_l2, T_gt)


# ------------------------------------------------------------------------------------------------ GPU: the engine
def _engine(**kw):
    import torch
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    from visualodometry_b200 import BAConfig, BundleAdjuster
    K = np.array([[KITTI[0], 0, KITTI[2]], [0, KITTI[1], KITTI[3]], [0, 0, 1.0]])
    return BundleAdjuster(K, BAConfig(**kw))


@pytest.mark.gpu
@pytest.mark.parametrize("loss", ["trivial", "huber"])
def test_refine_poses_matches_oracle_ragged_batch(loss):
    """ragged batch: different sizes, outliers, landmarks behind the camera, a problem without observations"""
    specs = [dict(n=300), dict(n=37, noise=1.0), dict(n=2048, outliers=0.1), dict(n=500, behind=11), dict(n=0), dict(n=5)]
    probs = [make_pose_problem(10 + i, **s) for i, s in enumerate(specs)]
    off = np.concatenate([[0], np.cumsum([p[2].shape[0] for p in probs])]).astype(np.int64)
    T0 = np.stack([p[1] for p in probs])
    X = np.concatenate([p[2] for p in probs]); uv = np.concatenate([p[3] for p in probs])
    kw = dict(max_iters=12, loss=loss, loss_scale=2.0)
    ba = _engine(record_history=True, **kw)
    T, sms = ba.refine_poses(T0, X, uv, off)
    hc = ba.dump("POSE_HIST_COST", (len(probs), 64)); hm = ba.dump("POSE_HIST_MU", (len(probs), 64))
    ha = ba.dump("POSE_HIST_ACCEPT", (len(probs), 64), np.int32)
    for b, p in enumerate(probs):
        To, so = pose_numpy.refine_pose(p[1], p[2], p[3], KITTI, **kw)
        s = sms[b]
        assert (s.status, s.iterations, s.n_invalid) == (so.status, so.iterations, so.n_invalid), b
        ch = np.array([so.initial_cost] + so.cost_hist)
        # the accept sequence / mu are compared while the PREVIOUS step still moved the cost by > 1e-4 relative: LM
        # converges quadratically here, the step after that is already rho = noise / noise
        moving = np.concatenate([[True], (ch[:-1] - ch[1:])[:-1] > 1e-4 * np.maximum(ch[:-2], 1e-30)])
        assert np.array_equal(ha[b, :so.iterations][moving], np.array(so.accept_hist)[moving]), b
        np.testing.assert_allclose(hc[b, :so.iterations], so.cost_hist, rtol=1e-9, atol=1e-9)
        np.testing.assert_allclose(hm[b, :so.iterations][moving], np.array(so.mu_hist)[moving], rtol=1e-6)
        assert abs(s.initial_cost - so.initial_cost) <= 1e-12 * max(1.0, so.initial_cost)
        assert abs(s.final_cost - so.final_cost) <= 1e-9 * max(1.0, so.final_cost)
        assert pose_err(T[b], To) <= 1e-8, b
    ba.close()


@pytest.mark.gpu
def test_refine_pose_device_arrays_and_stopping_tests():
    import torch
    T_gt, T0, X, uv = make_pose_problem(3, noise=0.3)
    ba = _engine(max_iters=50, ftol=1e-12)
    T_h, s_h = ba.refine_pose(T0, X, uv)
    T_d, s_d = ba.refine_pose(T0, torch.as_tensor(X).cuda(), torch.as_tensor(uv).cuda())
    To, so = pose_numpy.refine_pose(T0, X, uv, KITTI, max_iters=50, ftol=1e-12)
    assert s_h.status == 2 and s_h.iterations == so.iterations < 50
    assert np.array_equal(T_h, T_d) and s_h.final_cost == s_d.final_cost          # same kernel, same inputs
    assert pose_err(T_h, To) <= 1e-8 and pose_err(T_h, T_gt) < 1e-2
    ba.close()


@pytest.mark.gpu
def test_vo_loop_with_pose_refiner_tracks_at_least_as_well():
    """the optional hook after solvePnPRansac (vo_loop.VOLoop.pose_refiner): the loop runs, every refinement
    lowers the reprojection cost of the PnP inliers, and the trajectory stays as close to ground truth."""
    import cv2
    from visualodometry_b200 import vo_loop, vo_synth
    seq = vo_synth.SynthSequence(n_frames=150, seed=0)
    ba = _engine(max_iters=5, loss="huber", loss_scale=2.0)

    def run(refine):
        cv2.setRNGSeed(0)
        vo = vo_loop.VOLoop(seq.K, vo_loop.LoopConfig(), vo_synth.SynthFrontend(seq))
        log = []
        if refine:
            def hook(T, X, uv):
                T2, sm = ba.refine_pose(T, X, uv)
                log.append((sm.initial_cost, sm.final_cost, sm.status))
                return T2 if sm.ok else T
            vo.pose_refiner = hook
        for i in range(seq.n_frames):
            vo.process_frame(seq.image(i))
        tr = np.array(vo.trajectory)
        return vo_synth.ate_sim3(tr, seq.T_wc[-len(tr):, :3, 3]), len(tr), log
    ate0, n0, _ = run(False)
    ate1, n1, log = run(True)
    assert n1 == n0 == seq.n_frames and len(log) > 100
    assert all(f <= i * (1 + 1e-12) and st in (1, 2) for i, f, st in log)
    assert ate1 <= 1.5 * ate0 + 0.05
    ba.close()
