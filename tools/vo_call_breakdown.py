"""Per-call cost of BundleAdjuster.optimize on VO-like windows of varying size (the adapter's call pattern)."""
import time
import numpy as np
import torch
from visualodometry_b200 import BAConfig, BundleAdjuster, synth
from visualodometry_b200.ba import poses_from_T, T_from_poses

rng = np.random.default_rng(0)
ps = [synth.make_window(s, n_pts=int(rng.integers(700, 1400)), n_obs=int(rng.integers(5000, 9000))) for s in range(24)]
p0 = ps[0]
K = np.array([[p0.intr[0], 0, p0.intr[2]], [0, p0.intr[1], p0.intr[3]], [0, 0, 1.0]])
ba = BundleAdjuster(K, BAConfig(window=10, max_iters=10, loss="huber", loss_scale=5.0))
for rep in range(3):
    t = dict(set=0.0, solve=0.0, get=0.0)
    for p in ps:
        fx = p.fixed_cam.copy(); fx[:2] = 1
        torch.cuda.synchronize(); a = time.perf_counter()
        ba.set_problem(p.poses, p.points, p.obs_cam, p.obs_pt, p.obs_uv, fx)
        torch.cuda.synchronize(); b = time.perf_counter()
        ba.solve()
        torch.cuda.synchronize(); c = time.perf_counter()
        ba.get()
        d = time.perf_counter()
        t["set"] += b - a; t["solve"] += c - b; t["get"] += d - c
    print(f"rep {rep}: per call set_problem {1e3*t['set']/len(ps):.3f} ms  solve {1e3*t['solve']/len(ps):.3f} ms  get {1e3*t['get']/len(ps):.3f} ms")
