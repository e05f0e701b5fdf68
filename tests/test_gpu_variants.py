"""GPU parity tests (-m gpu) that close the round-1 gaps (VERDICT r1 "What's weak" 1-2):
every window-kernel instantiation is forced through the C ABI and held against the oracle; the engine is held
against the SECOND oracle (oracle/ba_numpy.py) and against an independent solver (scipy least_squares); the
cheirality accept rule; two handles on two devices in one process; the error exit of the pipelined batch call."""
import time

import numpy as np
import pytest

from visualodometry_b200 import synth

pytestmark = pytest.mark.gpu


def K_of(p):
    return np.array([[p.intr[0], 0, p.intr[2]], [0, p.intr[1], p.intr[3]], [0, 0, 1]])


def rel(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-300)) if a.size else 0.0


def pts_err(x, xo):
    return float((np.linalg.norm(x - xo, axis=1) / np.maximum(1.0, np.linalg.norm(xo, axis=1))).max())


def _c1(seed, **kw):
    return synth.make_window(seed, n_cams=10, n_pts=600, n_obs=4500, **kw)


def _c3(seed):
    return synth.make_window(seed, n_cams=20, n_pts=500, n_obs=7000, cam=synth.MALAGA if seed % 2 else synth.PARKING,
                             outlier_frac=0.1)


def _check_batch(ba, b, oracle, loss, iters=8):
    ba.set_problem(b.poses, b.points, b.obs_cam, b.obs_pt, b.obs_uv, b.fixed_cam, b.cam_off, b.pt_off, b.obs_off)
    sms = ba.solve()
    poses, pts = ba.get()
    po, xo, sos = oracle.solve_batch(b, max_iters=iters, loss=loss, threads=0)
    assert [(s.status, s.iterations, s.n_invalid) for s in sms] == [(s.status, s.iterations, s.n_invalid) for s in sos]
    ha = ba.dump("HIST_ACCEPT", (b.nb, 64), np.int32)[:, :iters]
    for w, so in enumerate(sos):
        ch = np.array([so.initial_cost] + list(so.cost_hist[:iters]))
        moving = np.concatenate([[True], (ch[:-1] - ch[1:])[:-1] > 1e-7 * ch[:-2]])
        assert np.array_equal(ha[w][moving], np.array(so.accept_hist[:iters])[moving]), w
    assert rel([s.final_cost for s in sms], [s.final_cost for s in sos]) <= 1e-9
    assert np.abs(poses - po).max() <= 1e-5 and pts_err(pts, xo) <= 1e-5


# (windows, shape, loss, BAConfig overrides) -> the instantiation of window_lm_kernel that runs (csrc/ba_window.cu)
VARIANTS = {
    "CfgBig_1cta_24_windows_of_20_keyframes": (24, _c3, "huber", {}),
    "CfgSmall_256_threads": (24, _c1, "trivial", {"variant": "small256"}),
    "CfgSmall4_128_threads": (24, _c1, "trivial", {}),
    "CfgSmallC_8cta_cluster_12_windows": (12, _c1, "trivial", {}),
    "CfgBigC_8cta_cluster_12_windows": (12, _c3, "huber", {}),
    "CfgSmallC_8cta_cluster_forced_3_windows": (3, _c1, "trivial", {"variant": "cluster8"}),
    "CfgBigC_8cta_cluster_forced_2_windows": (2, _c3, "huber", {"variant": "cluster8"}),
    "CfgSmallC16_3_windows": (3, _c1, "trivial", {}),
    "CfgBigC16_2_windows": (2, _c3, "huber", {}),
    "CfgSmall4_single_window_no_cluster": (1, _c1, "trivial", {"cluster": False}),
    "CfgBig_single_window_no_cluster": (1, _c3, "huber", {"cluster": False}),
}


@pytest.mark.parametrize("name", list(VARIANTS))
def test_every_window_kernel_instantiation(oracle, name):
    from visualodometry_b200 import BAConfig, BundleAdjuster
    nwin, make, loss, kw = VARIANTS[name]
    b = synth.pack_batch([make(300 + i) for i in range(nwin)])
    ba = BundleAdjuster(K_of(b), BAConfig(max_iters=8, loss=loss, record_history=True, **kw))
    _check_batch(ba, b, oracle, loss)
    ba.close()


def test_engine_stage_against_the_numpy_oracle():
    """the engine against the SECOND, independently written oracle (scipy.sparse algebra), one LM iteration"""
    from oracle import ba_numpy
    from visualodometry_b200 import BAConfig, BundleAdjuster
    for p, loss in ((synth.make_window(0), "trivial"), (_c3(1), "huber")):
        ba = BundleAdjuster(K_of(p), BAConfig(loss=loss))
        ba.set_problem(p.poses, p.points, p.obs_cam, p.obs_pt, p.obs_uv, p.fixed_cam)
        st = ba_numpy.stage(p, mu=1e4, loss=loss)
        n = st["S"].shape[0]
        free = np.flatnonzero(p.fixed_cam == 0)
        for k, key, shp, tol in (("RESIDUALS", "r", (p.no, 2), 1e-12), ("WEIGHTS", "w", (p.no,), 1e-12),
                                 ("JC", "Jc", (p.no, 2, 6), 1e-11), ("JP", "Jp", (p.no, 2, 3), 1e-11),
                                 ("B", "B", (p.nc, 6, 6), 1e-11), ("GC", "gc", (p.nc, 6), 1e-10), ("C", "C", (p.np_, 6), 1e-11),
                                 ("GP", "gp", (p.np_, 3), 1e-10), ("S", "S", (n, n), 1e-10), ("GRED", "gred", (n,), 1e-9),
                                 ("DC", "dc", (n,), 1e-8), ("DP", "dp", (p.np_, 3), 1e-8)):
            got, ref = ba.dump(k, shp), st[key]
            if k in ("B", "GC"):
                got, ref = got[free], ref[free]
            assert rel(got, ref) <= tol, (k, rel(got, ref))
        ba.close()


@pytest.mark.parametrize("shape", ["c1", "c3"])
def test_engine_converged_cost_against_scipy(shape):
    """cross-solver sanity for the ENGINE (not only the oracle): scipy's trust-region solver, started from the same
    initial guess, must not find a lower minimum, and must agree on the converged cost (trivial loss; scipy's robust
    losses act per scalar residual and are not comparable)."""
    from oracle import scipy_ba
    from visualodometry_b200 import BAConfig, BundleAdjuster
    p = synth.make_window(7, n_cams=10, n_pts=400, n_obs=3000) if shape == "c1" else \
        synth.make_window(8, n_cams=20, n_pts=300, n_obs=4500, cam=synth.MALAGA)
    ba = BundleAdjuster(K_of(p), BAConfig(max_iters=60, ftol=1e-13))
    ba.set_problem(p.poses, p.points, p.obs_cam, p.obs_pt, p.obs_uv, p.fixed_cam)
    sm = ba.solve()[0]
    poses, pts = ba.get()
    ba.close()
    sc = scipy_ba.solve(p, max_nfev=200)
    assert sm.n_invalid == 0
    assert abs(0.5 * (scipy_ba.residuals(p, poses, pts) ** 2).sum() - sm.final_cost) <= 1e-10 * sm.final_cost
    assert sm.final_cost <= sc["cost"] * (1 + 1e-6)
    assert abs(sm.final_cost - sc["cost"]) <= 1e-4 * sc["cost"]


def _near_gate_problem():
    import os
    import sys
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    from test_oracle_numpy import _near_gate_problem as mk
    return mk()


def test_steps_that_discard_observations_are_rejected(oracle):
    """an accepted step never increases the number of observations behind the cheirality gate; same
    accept sequence as the oracle on a problem where the old rule let 9 observations escape."""
    from oracle import ba_numpy
    from visualodometry_b200 import BAConfig, BundleAdjuster
    p = _near_gate_problem()
    n0 = int((~ba_numpy.observe(p.poses, p.points, p.obs_cam, p.obs_pt, p.obs_uv, p.intr, jac=False)["valid"]).sum())
    for kw in ({}, {"cluster": False}):
        ba = BundleAdjuster(K_of(p), BAConfig(max_iters=15, record_history=True, **kw))
        ba.set_problem(p.poses, p.points, p.obs_cam, p.obs_pt, p.obs_uv, p.fixed_cam)
        sm = ba.solve()[0]
        po, xo, so = oracle.solve(p, max_iters=15)
        assert sm.n_invalid <= n0 and sm.n_invalid == so.n_invalid
        assert (sm.accepted, sm.rejected) == (so.accepted, so.rejected) and so.rejected > 0
        assert abs(sm.final_cost - so.final_cost) <= 1e-9 * so.final_cost
        ba.close()


def test_two_handles_on_two_devices_in_one_process(oracle):
    """the >48 KB shared-memory opt-in of the window kernel is per device; tensors live on config.device;
    the engine gives torch's current device back."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    from visualodometry_b200 import BAConfig, BundleAdjuster
    p = synth.make_window(5, n_cams=6, n_pts=300, n_obs=1500)
    po, xo, so = oracle.solve(p, max_iters=6)
    torch.cuda.set_device(0)
    for dev in (0, 1):
        for cluster in (True, False):
            ba = BundleAdjuster(K_of(p), BAConfig(max_iters=6, device=dev, cluster=cluster))
            t = lambda a: torch.as_tensor(a).to(f"cuda:{dev}")
            ba.set_problem(t(p.poses), t(p.points), t(p.obs_cam), t(p.obs_pt), t(p.obs_uv), t(p.fixed_cam))
            sm = ba.solve()[0]
            poses, pts = ba.get(device=True)
            assert poses.device.index == dev and torch.cuda.current_device() == 0
            assert abs(sm.final_cost - so.final_cost) <= 1e-9 * so.final_cost
            assert np.abs(poses.cpu().numpy() - po).max() <= 1e-5
            ba.close()


def test_batch_call_error_exit_leaves_no_work_in_flight():
    """ADVICE r1 / VERDICT r1 item 12: a bad index in the LAST chunk makes b2ba_optimize_batch fail; the call must
    not return while earlier chunks still write into the caller's output buffers."""
    import torch
    from types import SimpleNamespace
    from visualodometry_b200 import BAConfig, BundleAdjuster
    b = synth.make_batch(40, seed0=40, distinct=8)
    bad_cam = b.obs_cam.copy()
    bad_cam[-3] = 99                                   # last window -> last chunk
    pin = lambda a: torch.from_numpy(np.ascontiguousarray(a)).pin_memory().numpy()
    hb = SimpleNamespace(poses=pin(b.poses), points=pin(b.points), obs_cam=pin(bad_cam), obs_pt=pin(b.obs_pt),
                         obs_uv=pin(b.obs_uv), fixed_cam=pin(b.fixed_cam), cam_off=b.cam_off, pt_off=b.pt_off,
                         obs_off=b.obs_off)
    out = (pin(np.zeros_like(b.poses)), pin(np.zeros_like(b.points)))
    ba = BundleAdjuster(K_of(b), BAConfig(max_iters=10))
    with pytest.raises(RuntimeError, match="out of range"):
        ba.optimize_batch(hb, out=out, chunks=5)
    snap = (out[0].copy(), out[1].copy())
    time.sleep(0.3)
    torch.cuda.synchronize()
    assert np.array_equal(out[0], snap[0]) and np.array_equal(out[1], snap[1])
    # offsets must start at 0 (same check as set_problem_batch)
    hb.obs_off = b.obs_off + 1
    with pytest.raises(RuntimeError, match="start at 0"):
        ba.optimize_batch(hb, out=out)
    # and the handle is still usable
    hb.obs_off, hb.obs_cam = b.obs_off, pin(b.obs_cam)
    res = ba.optimize_batch(hb, out=out)
    assert all(s.status in (1, 2) for s in res.summaries)
    ba.close()


_BUILDER_SCRIPT = r"""
import sys, numpy as np
sys.path.insert(0, sys.argv[1])
from visualodometry_b200 import BAConfig, BundleAdjuster, synth
ps = [synth.make_window(s, n_cams=4 + 4 * s, n_pts=120 + 90 * s, n_obs=700 + 900 * s) for s in range(5)]   # 4 ... 20 cameras
e = ps[2]; e.obs_cam, e.obs_pt, e.obs_uv = e.obs_cam[:0], e.obs_pt[:0], e.obs_uv[:0]   # a window without observations
keep = ps[3].obs_pt != 5                                    # and an unobserved landmark
ps[3].obs_cam, ps[3].obs_pt, ps[3].obs_uv = ps[3].obs_cam[keep], ps[3].obs_pt[keep], ps[3].obs_uv[keep]
b = synth.pack_batch(ps)
K = np.array([[b.intr[0], 0, b.intr[2]], [0, b.intr[1], b.intr[3]], [0, 0, 1.0]])
ba = BundleAdjuster(K, BAConfig(max_iters=4, loss="huber", loss_scale=2.0))
ba.set_problem(b.poses, b.points, b.obs_cam, b.obs_pt, b.obs_uv, b.fixed_cam, b.cam_off, b.pt_off, b.obs_off)
out = {k: ba.dump(k, shp, np.int32) for k, shp in (("PM_PERM", (b.no,)), ("CM_PERM", (b.no,)), ("PT_PTR", (b.points.shape[0] + 1,)),
                                                    ("CAM_PTR", (b.poses.shape[0] + 1,)), ("SCHED", (b.points.shape[0],)))}
sm = ba.solve()
out["poses"], out["pts"] = ba.get()
out["cost"] = np.array([s.final_cost for s in sm])
np.savez(sys.argv[2], **out)
"""


def test_window_index_kernel_equals_the_radix_sort_builder(tmp_path):
    """The one-CTA-per-window shared-memory index builder (csrc/ba_index.cu window_index_kernel) and the device-wide
    radix-sort builder it replaced on the window path (forced with B2BA_INDEX_SORT=1) produce the same streams, offsets
    and schedule -- hence bit-identical solves -- on a ragged batch with 4 ... 20 cameras per window, a window without
    observations and an unobserved landmark."""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    outs = []
    for sort in ("0", "1"):
        f = str(tmp_path / f"idx{sort}.npz")
        env = dict(os.environ, B2BA_INDEX_SORT=sort)
        r = subprocess.run([sys.executable, "-c", _BUILDER_SCRIPT, root, f], env=env, capture_output=True, text=True, timeout=300)
        assert r.returncode == 0, r.stderr[-2000:]
        outs.append(np.load(f))
    for k in outs[0].files:
        assert np.array_equal(outs[0][k], outs[1][k]), k
