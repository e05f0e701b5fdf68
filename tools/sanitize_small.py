"""Developer diagnostic (GPU box): small problems through every kernel family, for compute-sanitizer runs."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from visualodometry_b200 import BAConfig, BundleAdjuster, synth  # noqa: E402


def K_of(p):
    return np.array([[p.intr[0], 0, p.intr[2]], [0, p.intr[1], p.intr[3]], [0, 0, 1]])


def run(p, offs=None, **kw):
    ba = BundleAdjuster(K_of(p), BAConfig(max_iters=3, **kw))
    if offs is None:
        ba.set_problem(p.poses, p.points, p.obs_cam, p.obs_pt, p.obs_uv, p.fixed_cam)
    else:
        ba.set_problem(p.poses, p.points, p.obs_cam, p.obs_pt, p.obs_uv, p.fixed_cam, *offs)
    sm = ba.solve()
    ba.get()
    ba.close()
    return [(s.status, s.accepted, round(s.final_cost, 3)) for s in sm][:2]


if __name__ == "__main__":
    w = synth.make_window(3, n_cams=6, n_pts=150, n_obs=600)
    print("cluster small", run(w))
    print("1-CTA small  ", run(w, cluster=False))
    w2 = synth.make_window(1, n_cams=16, cam=synth.MALAGA, outlier_frac=0.1, n_pts=300, n_obs=2500)
    print("cluster big  ", run(w2, loss="huber"))
    print("1-CTA big    ", run(w2, loss="huber", cluster=False))
    b = synth.make_batch(20, seed0=7, n_cams=5, n_pts=120, n_obs=500)
    print("batch        ", run(b, (b.cam_off, b.pt_off, b.obs_off)))
    lg = synth.make_large(2, n_cams=30, n_pts=600, mode="random")
    print("large pcg    ", run(lg, solver="pcg"))
