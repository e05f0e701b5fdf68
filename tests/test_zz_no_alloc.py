"""GPU test (-m gpu) of b2ba_solve_window_noalloc, the entry point that allocates nothing (include/b2ba.h,
no counterpart in the reference).  Collected LAST on purpose (file name):
a failure here must not hide other tests.  Called through the C ABI (ctypes binding) and held
against the existing entry points (b2ba_set_problem_batch + b2ba_solve + b2ba_get) on the same inputs, on two calls in a
row -- the static scratch array keeps the previous call's contents, the second call must not see them."""
import ctypes as C

import numpy as np
import pytest

from visualodometry_b200 import synth

pytestmark = pytest.mark.gpu


def K_of(p):
    return np.array([[p.intr[0], 0, p.intr[2]], [0, p.intr[1], p.intr[3]], [0, 0, 1]])


def _summaries(buf, nb):
    from visualodometry_b200 import _lib
    raw = buf.cpu().numpy().tobytes()
    return [_lib.Summary.from_buffer_copy(raw[48 * i:48 * (i + 1)]) for i in range(nb)]


@pytest.mark.parametrize("nwin,kw,loss", [(3, {}, "trivial"),                    # latency path: one 16-CTA cluster per window
                                          (24, {}, "trivial"),                   # batched path: one CTA per window
                                          (2, {"cluster": False}, "huber")])     # 20-keyframe windows, 512-thread CTAs
def test_noalloc_entry_point_equals_solve(nwin, kw, loss):
    import torch
    from visualodometry_b200 import BAConfig, BundleAdjuster, _lib
    assert C.sizeof(_lib.Summary) == 48
    if loss == "huber":
        ws = [synth.make_window(70 + i, n_cams=20, n_pts=500, n_obs=7000, cam=synth.MALAGA, outlier_frac=0.1) for i in range(nwin)]
    else:
        ws = [synth.make_window(50 + i, n_cams=10, n_pts=400 + 50 * (i % 3), n_obs=3000) for i in range(nwin)]
    b = synth.pack_batch(ws)
    ba = BundleAdjuster(K_of(b), BAConfig(max_iters=8, loss=loss, **kw))
    ba.set_problem(b.poses, b.points, b.obs_cam, b.obs_pt, b.obs_uv, b.fixed_cam, b.cam_off, b.pt_off, b.obs_off)
    ref_sm = ba.solve()
    ref_poses, ref_pts = ba.get()
    assert all(s.final_cost < s.initial_cost for s in ref_sm)
    dev = torch.device("cuda", 0)
    sums = torch.zeros(48 * b.nb, dtype=torch.uint8, device=dev)
    for call in range(2):
        if call == 1:
            # leave OTHER contents in the shared scratch between the two calls: a different problem's solve through the
            # same entry point (its landmarks / scratch rows / trial landmarks overwrite the array)
            o = synth.pack_batch([synth.make_window(90, n_cams=8, n_pts=900, n_obs=5000)])
            other = BundleAdjuster(K_of(o), BAConfig(max_iters=3))
            other.set_problem(o.poses, o.points, o.obs_cam, o.obs_pt, o.obs_uv, o.fixed_cam, o.cam_off, o.pt_off, o.obs_off)
            op, ox = torch.as_tensor(o.poses).to(dev), torch.as_tensor(o.points).to(dev)
            other.solve_noalloc(op, ox, torch.zeros(48, dtype=torch.uint8, device=dev))
            torch.cuda.synchronize()
            other.close()
        poses = torch.as_tensor(np.ascontiguousarray(b.poses)).to(dev)
        pts = torch.as_tensor(np.ascontiguousarray(b.points)).to(dev)
        sums.zero_()
        torch.cuda.synchronize()
        free_before = torch.cuda.mem_get_info(0)[0]
        ba.solve_noalloc(poses, pts, sums)           # -> b2ba_solve_window_noalloc through ctypes
        torch.cuda.synchronize()
        if call == 1:                                # (the first call on a device may load the module lazily)
            assert torch.cuda.mem_get_info(0)[0] >= free_before, "device memory was allocated by the no-alloc entry point"
        got_p, got_x = poses.cpu().numpy(), pts.cpu().numpy()
        # same kernel on the same data: the tolerances of tests/test_gpu_parity.py (poses 1e-5, landmarks 1e-5 relative,
        # cost 1e-9) are met with a wide margin -- in fact bit for bit
        assert np.abs(got_p - ref_poses).max() <= 1e-5
        assert (np.linalg.norm(got_x - ref_pts, axis=1) / np.maximum(1.0, np.linalg.norm(ref_pts, axis=1))).max() <= 1e-5
        assert np.array_equal(got_p, ref_poses) and np.array_equal(got_x, ref_pts), call
        for s, r in zip(_summaries(sums, b.nb), ref_sm):
            assert (s.status, s.iterations, s.accepted, s.rejected, s.n_invalid) == (r.status, r.iterations, r.accepted, r.rejected, r.n_invalid)
            assert abs(s.final_cost - r.final_cost) <= 1e-9 * r.final_cost and s.initial_cost == r.initial_cost
    # the handle's own state was not touched: its result is still the one of the reference solve
    p2, x2 = ba.get()
    assert np.array_equal(p2, ref_poses) and np.array_equal(x2, ref_pts)
    ba.close()


def test_noalloc_entry_point_rejects_what_it_cannot_hold():
    import torch
    from visualodometry_b200 import BAConfig, BundleAdjuster, _lib
    dev = torch.device("cuda", 0)
    ws = [synth.make_window(10 + (i % 4), n_cams=5, n_pts=2100, n_obs=9000) for i in range(16)]      # 33600 landmarks
    b = synth.pack_batch(ws)
    assert b.points.shape[0] > _lib.NOALLOC_MAX_POINTS
    ba = BundleAdjuster(K_of(b), BAConfig(max_iters=2))
    ba.set_problem(b.poses, b.points, b.obs_cam, b.obs_pt, b.obs_uv, b.fixed_cam, b.cam_off, b.pt_off, b.obs_off)
    poses, pts = torch.as_tensor(b.poses).to(dev), torch.as_tensor(b.points).to(dev)
    with pytest.raises(RuntimeError, match="exceed the static scratch"):
        ba.solve_noalloc(poses, pts, torch.zeros(48 * b.nb, dtype=torch.uint8, device=dev))
    assert torch.equal(poses.cpu(), torch.as_tensor(b.poses))                  # nothing ran
    ba.close()
    # a large-path problem (more than 24 cameras) has no window kernel to run
    p = synth.make_large(1, n_cams=40, n_pts=800, mode="local")
    ba = BundleAdjuster(K_of(p), BAConfig(max_iters=2))
    ba.set_problem(p.poses, p.points, p.obs_cam, p.obs_pt, p.obs_uv, p.fixed_cam)
    with pytest.raises(RuntimeError, match="window problem"):
        ba.solve_noalloc(torch.as_tensor(p.poses).to(dev), torch.as_tensor(p.points).to(dev),
                         torch.zeros(48, dtype=torch.uint8, device=dev))
    ba.close()
