import sys, numpy as np
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
from test_pose import make_pose_problem, KITTI
from visualodometry_b200 import BAConfig, BundleAdjuster
K = np.array([[KITTI[0], 0, KITTI[2]], [0, KITTI[1], KITTI[3]], [0, 0, 1.0]])
ba = BundleAdjuster(K, BAConfig())
T_gt, T0, X, uv = make_pose_problem(1, n=1100, noise=0.3, outliers=0.5, rot=0.03, trans=0.8)
import time
for i in range(3):
    t=time.perf_counter(); r = ba.ransac_pose(T0, X, uv, 1.0, n_hyp=1024, seed=i); print(r[0], len(r[2]), r[4], (time.perf_counter()-t)*1e3)
