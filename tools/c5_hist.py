"""Per-LM-iteration history of the C5 problem (cost, mu, accept, CG iterations): A/B tool for solver changes.
usage: python tools/c5_hist.py [local|random] [n_cams] [n_pts]   (PYTHONPATH picks the package under test)"""
import sys
import numpy as np
from visualodometry_b200 import BAConfig, BundleAdjuster, synth

mode = sys.argv[1] if len(sys.argv) > 1 else "random"
nc = int(sys.argv[2]) if len(sys.argv) > 2 else 1000
npts = int(sys.argv[3]) if len(sys.argv) > 3 else 200000
p = synth.make_large(0, n_cams=nc, n_pts=npts, mode=mode)
K = np.array([[p.intr[0], 0, p.intr[2]], [0, p.intr[1], p.intr[3]], [0, 0, 1.0]])
ba = BundleAdjuster(K, BAConfig(max_iters=20, record_history=True, solver="auto"))
ba.set_problem(p.poses, p.points, p.obs_cam, p.obs_pt, p.obs_uv, p.fixed_cam)
sm = ba.solve()[0]
print("summary", sm)
print("cost", ba.dump("HIST_COST", (1, 64))[0, :20])
print("mu", ba.dump("HIST_MU", (1, 64))[0, :20])
print("accept", ba.dump("HIST_ACCEPT", (1, 64), np.int32)[0, :20])
