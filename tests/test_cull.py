"""Outlier culling after BA on the device (`b2ba_cull`): pinned bit for bit to the host rule of
adapter.cull_window -- which is the reference's landmark gate (`frontend.py:139-142`, `config.py:31-32`) applied to the
engine's reprojection errors, themselves pinned to cv2.projectPoints (tests/test_adapter.py)."""
import copy

import numpy as np
import pytest

from visualodometry_b200 import synth

pytestmark = pytest.mark.gpu


def K_of(p):
    return np.array([[p.intr[0], 0, p.intr[2]], [0, p.intr[1], p.intr[3]], [0, 0, 1]])


def _host_rule(err, obs_pt, npts, max_err):
    bad = (err < 0) | (err > max_err)           # adapter.cull_window
    keep = ~bad
    return keep, np.bincount(obs_pt[keep], minlength=npts).astype(np.int32), np.flatnonzero(keep).astype(np.int32)


@pytest.mark.parametrize("shape", ["c1", "c3_huber", "tiny", "behind_camera"])
def test_cull_equals_host_rule(shape):
    from visualodometry_b200 import BAConfig, BundleAdjuster
    if shape == "c1":
        p, loss = synth.make_window(3, n_pts=800, n_obs=6000, outlier_frac=0.08), "trivial"
    elif shape == "c3_huber":
        p, loss = synth.make_window(4, n_cams=20, n_pts=500, n_obs=7000, cam=synth.MALAGA, outlier_frac=0.1), "huber"
    elif shape == "tiny":
        p, loss = synth.make_window(5, n_cams=3, n_pts=20, n_obs=45, outlier_frac=0.2), "trivial"
    else:
        p, loss = synth.make_window(6, n_pts=300, n_obs=2000), "trivial"
        p.points[::17, 2] = -5.0                 # landmarks behind every camera: error -1 -> culled
    ba = BundleAdjuster(K_of(p), BAConfig(max_iters=5, loss=loss, loss_scale=2.0))
    ba.set_problem(p.poses, p.points, p.obs_cam, p.obs_pt, p.obs_uv, p.fixed_cam)
    if shape != "behind_camera":
        ba.solve()
    err = ba.reproj_errors()
    for max_err in (0.0, 1.0, 2.5, 1e9):
        keep, live, idx = ba.cull(max_err)
        k0, l0, i0 = _host_rule(err, p.obs_pt, p.np_, max_err)
        assert np.array_equal(keep, k0) and np.array_equal(live, l0) and np.array_equal(idx, i0), (shape, max_err)
    if shape == "behind_camera":
        assert (err < 0).sum() > 0 and not keep[err < 0].any()
    ba.close()


def test_cull_batch_and_more_than_one_tile():
    """A batch of ragged windows (window-local indices, several 2048-observation tiles per launch)."""
    from visualodometry_b200 import BAConfig, BundleAdjuster
    b = synth.pack_batch([synth.make_window(seed=11 + i, n_cams=6 + i, n_pts=300 + 150 * i, n_obs=1500 + 900 * i, outlier_frac=0.05)
                          for i in range(5)])
    ba = BundleAdjuster(K_of(b), BAConfig(max_iters=3))
    ba.set_problem(b.poses, b.points, b.obs_cam, b.obs_pt, b.obs_uv, b.fixed_cam, b.cam_off, b.pt_off, b.obs_off)
    ba.solve()
    err = ba.reproj_errors()
    gpt = b.obs_pt + np.repeat(b.pt_off[:-1], np.diff(b.obs_off)).astype(np.int32)
    keep, live, idx = ba.cull(1.2)
    k0, l0, i0 = _host_rule(err, gpt, int(b.pt_off[-1]), 1.2)
    assert 0 < k0.sum() < k0.size
    assert np.array_equal(keep, k0) and np.array_equal(live, l0) and np.array_equal(idx, i0)
    ba.close()


def test_cull_apply_equals_a_fresh_problem_of_the_kept_observations():
    """optimise -> cull(apply) -> optimise on the device == set_problem of the kept observations at the refined state."""
    from visualodometry_b200 import BAConfig, BundleAdjuster
    p = synth.make_window(8, n_pts=600, n_obs=4500, outlier_frac=0.1, outlier_px=30.0)
    cfg = BAConfig(max_iters=8, loss="huber", loss_scale=2.0)
    ba = BundleAdjuster(K_of(p), cfg)
    ba.set_problem(p.poses, p.points, p.obs_cam, p.obs_pt, p.obs_uv, p.fixed_cam)
    ba.solve()
    poses1, pts1 = ba.get()
    keep, live, idx = ba.cull(3.0, apply=True)
    assert 0.02 < (~keep).mean() < 0.8, (~keep).mean()
    s2 = ba.solve()[0]
    poses2, pts2 = ba.get()
    ref = BundleAdjuster(K_of(p), cfg)
    ref.set_problem(poses1, pts1, p.obs_cam[keep], p.obs_pt[keep], p.obs_uv[keep], p.fixed_cam)
    s3 = ref.solve()[0]
    poses3, pts3 = ref.get()
    assert (s2.status, s2.iterations, s2.accepted) == (s3.status, s3.iterations, s3.accepted)
    assert s2.initial_cost == s3.initial_cost and s2.final_cost == s3.final_cost
    assert np.array_equal(poses2, poses3) and np.array_equal(pts2, pts3)
    # the index of the culled problem is the canonical one of the kept observations
    assert np.array_equal(ba.dump("PT_PTR", (p.np_ + 1,), np.int32), ref.dump("PT_PTR", (p.np_ + 1,), np.int32))
    # every surviving observation was within the gate when the second solve started
    assert s2.initial_cost <= 0.5 * keep.sum() * 3.0 ** 2
    ba.close(); ref.close()


def _vo_window_from(p):
    """keyframe dicts + landmark dict (the reference's structures, vo.py:17,57) holding the observations of problem p"""
    from visualodometry_b200.ba import T_from_poses

    class VO:
        pass
    vo = VO()
    vo.K = K_of(p)
    vo.map_points = {100 + i: p.points[i].copy() for i in range(p.np_)}
    vo.T_wc, vo.last_pos = np.eye(4), np.zeros(3)
    T_cw = T_from_poses(p.poses)
    window = []
    for j in range(p.nc):
        m = p.obs_cam == j
        window.append({"feats": {"keypoints": p.obs_uv[m][None].astype(np.float32)}, "ids": (100 + p.obs_pt[m]).astype(int),
                       "T_wc": np.linalg.inv(T_cw[j])})
    return vo, window


def test_adapter_device_cull_equals_host_cull():
    """adapter.cull_window_device edits the reference's structures exactly like adapter.cull_window."""
    from visualodometry_b200 import BAConfig, BundleAdjuster, adapter
    p = synth.make_window(2, n_cams=6, n_pts=300, n_obs=1500, outlier_frac=0.1, outlier_px=25.0)
    vo, window = _vo_window_from(p)
    vo2, window2 = copy.deepcopy(vo), copy.deepcopy(window)
    ba = BundleAdjuster(vo.K, BAConfig(max_iters=5, loss="huber", loss_scale=2.0))
    for v, w, device in ((vo, window, True), (vo2, window2, False)):
        packed = adapter.pack_window(w, v.map_points)
        res = ba.optimize(packed.T_cw, packed.points, packed.obs_cam, packed.obs_pt, packed.obs_uv, packed.fixed_cam)
        adapter.write_back(v, w, packed, res.T_cw, res.points)
        n = adapter.cull_window_device(v, w, packed, ba, 2.0) if device else adapter.cull_window(v, w, packed, ba.reproj_errors(), 2.0)
        assert n > 0
    assert sorted(vo.map_points) == sorted(vo2.map_points)
    assert any((a["ids"] == -1).any() for a in window)
    for a, b_ in zip(window, window2):
        assert np.array_equal(a["ids"], b_["ids"])
    ba.close()
