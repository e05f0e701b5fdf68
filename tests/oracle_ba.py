"""Test helper: the CPU oracle behind the same `optimize(...)` call as BundleAdjuster, so host-side logic
(adapter, VO loop) can be exercised on the CPU.  Test infrastructure only (imports oracle/)."""
import numpy as np

from oracle import ba_cpu
from visualodometry_b200 import synth
from visualodometry_b200.ba import BAResult, BASummary, T_from_poses, poses_from_T


class OracleBA:
    def __init__(self, K, config):
        self.config = config
        K = np.asarray(K, np.float64)
        self.intr = np.array([K[0, 0], K[1, 1], K[0, 2], K[1, 2]])
        self.calls = 0

    def optimize(self, T_cw, points, obs_cam, obs_pt, obs_uv, fixed_cam) -> BAResult:
        self.calls += 1
        p = synth.BAProblem(poses_from_T(T_cw), np.ascontiguousarray(points, np.float64),
                            np.asarray(obs_cam, np.int32), np.asarray(obs_pt, np.int32),
                            np.asarray(obs_uv, np.float32), np.asarray(fixed_cam).astype(np.uint8), self.intr)
        c = self.config
        solver = "chol" if p.nc <= 24 else "pcg"
        po, xo, so = ba_cpu.solve(p, max_iters=c.max_iters, loss=c.loss, loss_scale=c.loss_scale, z_min=c.z_min,
                                  init_mu=c.init_radius, solver=solver)
        sm = BASummary(so.status, so.iterations, so.accepted, so.rejected, so.n_invalid, so.pcg_iters_total,
                       so.initial_cost, so.final_cost, so.final_mu)
        return BAResult(T_from_poses(po), xo, sm, [sm])
