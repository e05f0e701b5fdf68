"""The restated VO frame loop (visualodometry_b200/vo_loop.py) against the REFERENCE's own loop.

Golden: tests/golden/reference_vo_traj.npz = poses of the unmodified /root/reference/src/modules/vo.py
VisualOdometry on a seeded synthetic sequence (tests/golden/make_reference_vo_fixture.py).  The restatement makes
the same OpenCV calls in the same order, so the bar is exact equality; 1e-9 is allowed for a different
OpenCV build on another machine."""
import os

import numpy as np
import pytest

cv2 = pytest.importorskip("cv2")

from visualodometry_b200 import vo_loop, vo_synth  # noqa: E402

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def run_loop(vo, seq):
    poses, kf = [], []
    for i in range(seq.n_frames):
        before = vo.keyframe
        vo.process_frame(seq.image(i))
        if vo.keyframe is not before:
            kf.append(i)
        poses.append(vo.T_wc.copy())
    return np.array(poses), np.array(kf)


@pytest.fixture(scope="module")
def gold():
    return np.load(os.path.join(ROOT, "tests", "golden", "reference_vo_traj.npz"))


def test_restated_loop_matches_reference_fixture(gold):
    seq = vo_synth.SynthSequence(n_frames=int(gold["n_frames"]), seed=int(gold["seed"]))
    cv2.setRNGSeed(0)
    vo = vo_loop.VOLoop(seq.K, vo_loop.LoopConfig(), vo_synth.SynthFrontend(seq))
    poses, kf = run_loop(vo, seq)
    same_cv = str(gold["cv2_version"]) == cv2.__version__
    assert np.array_equal(kf, gold["keyframe_frames"])
    assert np.abs(poses - gold["T_wc"]).max() <= (0.0 if same_cv else 1e-9)
    assert vo.next_pt_id == int(gold["next_pt_id"])
    assert np.array_equal(np.array(sorted(vo.map_points)), gold["map_ids"])
    assert np.abs(np.array([vo.map_points[p] for p in gold["map_ids"]]) - gold["map_xyz"]).max() <= (0.0 if same_cv else 1e-9)
    assert np.array_equal(vo.keyframe["ids"], gold["last_ids"])


@pytest.mark.skipif(not os.path.isdir("/root/reference/src"), reason="reference tree only exists in the build container")
def test_restated_loop_matches_live_reference():
    """Same check against the reference imported live, on a different seed than the fixture."""
    import sys
    sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
    import make_reference_vo_fixture as mk
    seq = vo_synth.SynthSequence(n_frames=120, seed=11)
    cv2.setRNGSeed(0)
    ref_poses, ref_kf = mk.run(mk.reference_vo(seq, vo_synth.SynthFrontend(seq)), seq)
    cv2.setRNGSeed(0)
    poses, kf = run_loop(vo_loop.VOLoop(seq.K, vo_loop.LoopConfig(), vo_synth.SynthFrontend(seq)), seq)
    assert np.array_equal(kf, ref_kf) and np.array_equal(poses, ref_poses)


def test_loop_config_equals_reference_kitti_config():
    if not os.path.isdir("/root/reference/src"):
        pytest.skip("reference tree only exists in the build container")
    import sys
    sys.path.insert(0, "/root/reference/src")
    from config.config import get_config
    ref = get_config("kitti")
    mine = vo_loop.LoopConfig()
    for k, v in vars(mine).items():
        assert getattr(ref, k) == v, k


def test_synthetic_sequence_shape():
    seq = vo_synth.SynthSequence(n_frames=120, seed=0)
    n = [len(f.wid) for f in seq.frames]
    assert max(n) <= 2048 and min(n[80:]) > 800
    assert all(f.uv.dtype == np.float32 for f in seq.frames)
    fe = vo_synth.SynthFrontend(seq)
    f0, f1 = fe.process_image(seq.image(100)), fe.process_image(seq.image(102))
    m = fe.match_frames(f0, f1)
    assert m.shape[1] == 2 and len(m) > 500
    good = f0["wid"][m[:, 0]] == f1["wid"][m[:, 1]]
    assert 0.90 < good.mean() < 0.99          # 5 % wrong matches
