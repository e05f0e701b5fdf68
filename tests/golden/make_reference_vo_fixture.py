"""Generates tests/golden/reference_vo_traj.npz by RUNNING THE REFERENCE's own frame loop
(/root/reference/src/modules/vo.py: VisualOdometry, unmodified) on a seeded synthetic sequence with the
synthetic id-matched frontend (visualodometry_b200/vo_synth.py).  `lightglue` and `rerun` are not installed
(learned frontend / viewer: out of scope, the viewer is unavailable on this box) and are replaced by inert
stub modules; `_visualize` is a no-op.  Run in the build container only; the GPU box never reads
/root/reference -- the .npz is committed and pins visualodometry_b200/vo_loop.py (tests/test_vo_loop.py).
"""
import contextlib
import io
import os
import sys
import types

import numpy as np

N_FRAMES, SEED = 260, 3


def stub_modules():
    lg = types.ModuleType("lightglue")
    lg.LightGlue = lg.SuperPoint = object
    sys.modules.setdefault("lightglue", lg)

    class _Any:
        def __getattr__(self, k):
            return _Any()

        def __call__(self, *a, **k):
            return _Any()

    rr = types.ModuleType("rerun")

    def _ga(k):
        if k.startswith("__"):
            raise AttributeError(k)
        return _Any()
    rr.__getattr__ = _ga
    sys.modules.setdefault("rerun", rr)


def reference_vo(seq, frontend):
    """The unmodified reference class, synthetic frontend swapped in after construction."""
    stub_modules()
    if "/root/reference/src" not in sys.path:
        sys.path.insert(0, "/root/reference/src")
    from config.config import get_config
    from modules.vo import VisualOdometry
    cfg = get_config("kitti")
    cfg.extractor_type = "none"            # neither SuperPoint nor SIFT is constructed (frontend.py:17-34)
    vo = VisualOdometry(seq.K, cfg)
    vo.frontend = frontend
    vo._visualize = lambda *a, **k: None
    return vo


def run(vo, seq):
    poses, kf_frames = [], []
    with contextlib.redirect_stdout(io.StringIO()):
        for i in range(seq.n_frames):
            before = vo.keyframe
            vo.process_frame(seq.image(i))
            if vo.keyframe is not before:
                kf_frames.append(i)
            poses.append(vo.T_wc.copy())
    return np.array(poses), np.array(kf_frames)


if __name__ == "__main__":
    ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..")
    sys.path.insert(0, ROOT)
    import cv2
    from visualodometry_b200 import vo_synth
    seq = vo_synth.SynthSequence(n_frames=N_FRAMES, seed=SEED)
    cv2.setRNGSeed(0)
    vo = reference_vo(seq, vo_synth.SynthFrontend(seq))
    poses, kf = run(vo, seq)
    pids = np.array(sorted(vo.map_points))
    out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "reference_vo_traj.npz")
    np.savez_compressed(out, n_frames=N_FRAMES, seed=SEED, T_wc=poses, keyframe_frames=kf, next_pt_id=vo.next_pt_id,
                        map_ids=pids, map_xyz=np.array([vo.map_points[p] for p in pids]),
                        last_ids=vo.keyframe["ids"], cv2_version=cv2.__version__)
    print(out, poses.shape, len(kf), "keyframes", vo.next_pt_id, "landmarks created")
