"""Host-side restatement of the reference's VO frame loop, the CALLER of the BA hot path.

`/root/reference` does not exist on the GPU box, so BASELINE config 2 (end-to-end VO with the BA step on a
B200) cannot import `src/modules/vo.py` there.  `VOLoop` follows `VisualOdometry` step for step — same state
attributes (`K`, `cfg`, `map_points`, `next_pt_id`, `T_wc`, `last_pos`, `baseline_speed`, `keyframe`,
`initialized`, `frame_id`), same OpenCV calls in the same order with the same arguments, same seam
`_create_keyframe(curr_feats, curr_ids, ref_indices, curr_indices)` — so that `adapter.install()` hooks it exactly
as it hooks the reference class.  It is pinned against the reference itself: `tests/golden/reference_vo_traj.npz`
holds the trajectory of the unmodified `VisualOdometry` on a seeded synthetic sequence
(`tests/golden/make_reference_vo_fixture.py`), and `tests/test_vo_loop.py` requires the same poses from this file.

Not restated: visualisation (`vo.py:301-368`, rerun: unavailable on this box) and the console prints.
"""
from __future__ import annotations

from dataclasses import dataclass

import cv2
import numpy as np


@dataclass
class LoopConfig:
    """The fields of the reference's VOConfig (`src/config/config.py:4-46`) the frame loop reads, with the
    values `get_config("kitti")` produces (`config.py:49-61`)."""
    global_scale: float = 20.0
    min_median_flow: float = 40.0
    min_inliers: int = 10
    init_ransac_prob: float = 0.999
    init_ransac_thresh: float = 1.0
    min_depth: float = 0.001
    max_reproj_err: float = 5.0
    pnp_reproj_err: float = 1.0
    kf_min_tracked: int = 80
    turn_thresh: float = 0.01
    move_thresh: float = 0.01
    turn_smoothing: float = 0.2
    trans_smoothing: float = 0.4
    baseline_lr: float = 0.002
    scale_clamp_min: float = 0.5
    scale_clamp_max: float = 3.0


def _kp(feats) -> np.ndarray:
    k = feats["keypoints"][0]
    return k.cpu().numpy() if hasattr(k, "cpu") else np.asarray(k)


def triangulate(T_cw1, T_cw2, pts1, pts2, K, cfg):
    """Two-view linear triangulation with the cheirality and reprojection gates of
    `triangulate_points` (`frontend.py:115-148`): returns (accepted points, accept mask)."""
    if len(pts1) == 0:
        return np.empty((0, 3)), np.zeros(0, dtype=bool)
    X4 = cv2.triangulatePoints(K @ T_cw1[:3, :], K @ T_cw2[:3, :], pts1.T, pts2.T)
    X = (X4[:3] / X4[3]).T
    depth2 = (T_cw2[:3, :3] @ X.T + T_cw2[:3, 3:4]).T[:, 2]
    proj, _ = cv2.projectPoints(X, T_cw2[:3, :3], T_cw2[:3, 3], K, None)
    err = np.linalg.norm(proj.reshape(-1, 2) - pts2, axis=1)
    keep = (depth2 > cfg.min_depth) & (err < cfg.max_reproj_err)
    return X[keep], keep


class VOLoop:
    MAX_MAP_POINTS = 20000          # vo.py:38

    def __init__(self, K: np.ndarray, config, frontend):
        self.K, self.cfg, self.frontend = K, config, frontend
        self.frame_id = 0
        self.map_points = {}
        self.next_pt_id = 0
        self.T_wc = np.eye(4)                 # camera->world (vo.py:19; see SURVEY §8 a-R2 on the naming)
        self.last_pos = np.zeros(3)
        self.baseline_speed = 1.0
        self.is_turning = False
        self.keyframe = None
        self.trajectory = []
        self.initialized = False
        # NOT in the reference: optional motion-only BA after the PnP, a callable
        # (T_cw, X, uv) -> T_cw, e.g. `lambda T, X, uv: ba.refine_pose(T, X, uv)[0]`.  None = the reference's loop.
        self.pose_refiner = None
        # NOT in the reference either: replacement of `triangulate`, same signature and return
        # values, e.g. the CUDA kernel behind `BundleAdjuster.triangulate`.  None = cv2, as the reference does.
        self.triangulator = None
        # NOT in the reference: replacement of the `cv2.solvePnPRansac` call (`vo.py:135-141`) by a tracker that starts
        # from the previous pose, a callable (X, uv, T_cw_prior) -> (ok, T_cw, inlier indices), e.g.
        # `lambda X, uv, T: ba.track_pose(T, X, uv, cfg.pnp_reproj_err)[:3]` (robust motion-only BA on the GPU, no
        # sampling).  When it reports no consensus the loop falls through to cv2.solvePnPRansac, so it never resets
        # where the reference would not.  None = cv2.solvePnPRansac only, as the reference does (the pinned loop).
        self.pnp_solver = None
        self.pnp_fallbacks = 0
        # NOT in the reference: `dense_map = True` keeps `map_points` in a LandmarkTable (a dict with a dense id -> slot
        # mirror, landmark_table.py) and answers the matched-id look-ups of vo.py:121-130 and the pruning of
        # vo.py:41-47 with array operations; results are identical to the dict scans (tests/test_landmark_table.py).
        self.dense_map = False

    def _new_map(self):
        if self.dense_map:
            from .landmark_table import LandmarkTable
            return LandmarkTable()
        return {}

    # ------------------------------------------------------------------ vo.py:49-250
    def process_frame(self, img) -> None:
        if self.dense_map and not hasattr(self.map_points, "lookup"):
            from .landmark_table import LandmarkTable
            self.map_points = LandmarkTable(self.map_points)
        feats = self.frontend.process_image(img)
        ids = np.full(feats["keypoints"].shape[1], -1, dtype=int)
        if self.keyframe is None:                                          # vo.py:56-61
            self.keyframe = {"feats": feats, "ids": ids, "T_wc": np.eye(4)}
            self._end_frame()
            return
        matches = self.frontend.match_frames(self.keyframe["feats"], feats)
        ref_idx, cur_idx = matches[:, 0], matches[:, 1]
        if not self.initialized:
            self._initialize(feats, ids, ref_idx, cur_idx)
        else:
            self._track(feats, ids, ref_idx, cur_idx)
        self._end_frame()

    def _end_frame(self):
        self.trajectory.append(self.T_wc[:3, 3].copy())                     # vo.py:323 (inside _visualize)
        self.frame_id += 1

    def _flow(self, feats, ref_idx, cur_idx):
        a, b = _kp(self.keyframe["feats"])[ref_idx], _kp(feats)[cur_idx]
        return a, b, (np.median(np.linalg.norm(a - b, axis=1)) if len(a) > 0 else 0.0)

    # ------------------------------------------------------------------ vo.py:71-117
    def _initialize(self, feats, ids, ref_idx, cur_idx):
        uv_ref, uv_cur, flow = self._flow(feats, ref_idx, cur_idx)
        if flow < self.cfg.min_median_flow:
            return
        E, _ = cv2.findEssentialMat(uv_ref, uv_cur, self.K, method=cv2.RANSAC, prob=self.cfg.init_ransac_prob,
                                    threshold=self.cfg.init_ransac_thresh)
        if E is None:
            return
        _, R, t, _ = cv2.recoverPose(E, uv_ref, uv_cur, self.K)
        T_cw = np.eye(4)
        T_cw[:3, :3], T_cw[:3, 3] = R, t.flatten() * self.cfg.global_scale
        self.T_wc = np.linalg.inv(T_cw)
        self.last_pos = np.zeros(3)
        self.baseline_speed = np.linalg.norm(self.T_wc[:3, 3] - self.last_pos)
        self.initialized = True
        self._create_keyframe(feats, ids, ref_idx, cur_idx)

    # ------------------------------------------------------------------ vo.py:119-245
    def _track(self, feats, ids, ref_idx, cur_idx):
        cfg, mp = self.cfg, self.map_points
        kf_ids = self.keyframe["ids"][ref_idx]
        if self.dense_map and hasattr(mp, "lookup"):
            valid, X = mp.lookup(kf_ids)
        else:
            valid = np.array([(p != -1 and p in mp) for p in kf_ids])
        if not np.sum(valid) > cfg.min_inliers:
            self._reset_system()
            return
        if not (self.dense_map and hasattr(mp, "lookup")):
            X = np.array([mp[p] for p in kf_ids[valid]])
        uv = _kp(feats)[cur_idx[valid]]
        ok = False
        if self.pnp_solver is not None:
            ok, T_cw, inliers = self.pnp_solver(X, uv, np.linalg.inv(self.T_wc))
            if ok:
                R = T_cw[:3, :3]
                inliers = np.asarray(inliers).reshape(-1, 1)
            else:
                self.pnp_fallbacks += 1                # no consensus around the prior: the reference's own call decides
        if not ok:
            ok, rvec, tvec, inliers = cv2.solvePnPRansac(X, uv, self.K, None, reprojectionError=cfg.pnp_reproj_err)
            if not ok:
                self._reset_system()
                return
            R, _ = cv2.Rodrigues(rvec)
            T_cw = np.eye(4)
            T_cw[:3, :3], T_cw[:3, 3] = R, tvec.flatten()
        if self.pose_refiner is not None and inliers is not None and len(inliers) >= 6:
            sel = inliers.flatten()
            T_cw = self.pose_refiner(T_cw, X[sel], uv[sel])
            R = T_cw[:3, :3]
        T_raw = np.linalg.inv(T_cw)
        self._speed_control(R, T_raw)
        self.last_pos = self.T_wc[:3, 3].copy()
        if inliers is not None:
            sel = inliers.flatten()
            ids[cur_idx[valid][sel]] = kf_ids[valid][sel]
        tracked = np.sum(ids != -1)
        _, _, flow = self._flow(feats, ref_idx, cur_idx)
        if flow > cfg.min_median_flow or tracked < cfg.kf_min_tracked:      # vo.py:219-228
            self._create_keyframe(feats, ids, ref_idx, cur_idx)

    # ------------------------------------------------------------------ vo.py:152-202 (scale-drift heuristic)
    def _speed_control(self, R_cw, T_raw):
        cfg = self.cfg
        delta = T_raw[:3, 3] - self.last_pos
        speed = np.linalg.norm(delta)
        rot, _ = cv2.Rodrigues(R_cw @ self.T_wc[:3, :3])
        turning = np.linalg.norm(rot) > cfg.turn_thresh
        if not speed > cfg.move_thresh:
            self.T_wc = T_raw
            return
        self.is_turning = bool(turning)
        mix = cfg.turn_smoothing if turning else cfg.trans_smoothing
        scale = (mix * self.baseline_speed + (1 - mix) * speed) / speed
        if not turning:
            self.baseline_speed = (1 - cfg.baseline_lr) * self.baseline_speed + cfg.baseline_lr * speed
        scale = np.clip(scale, cfg.scale_clamp_min, cfg.scale_clamp_max)
        self.T_wc[:3, 3] = self.last_pos + delta * scale
        self.T_wc[:3, :3] = T_raw[:3, :3]

    # ------------------------------------------------------------------ vo.py:252-288 -- the BA seam
    def _create_keyframe(self, curr_feats, curr_ids, ref_indices, curr_indices) -> None:
        T_ref, T_cur = np.linalg.inv(self.keyframe["T_wc"]), np.linalg.inv(self.T_wc)
        fresh = curr_ids[curr_indices] == -1
        if np.sum(fresh) > 0:
            X, keep = (self.triangulator or triangulate)(T_ref, T_cur, _kp(self.keyframe["feats"])[ref_indices[fresh]],
                                                         _kp(curr_feats)[curr_indices[fresh]], self.K, self.cfg)
            slots = curr_indices[fresh][keep]
            for j in range(len(X)):
                # a keypoint matched twice gets the LAST id, and every id is stored (vo.py:277-284)
                self.map_points[self.next_pt_id] = X[j]
                curr_ids[slots[j]] = self.next_pt_id
                self.next_pt_id += 1
        self.keyframe = {"feats": curr_feats, "ids": curr_ids, "T_wc": self.T_wc.copy()}
        floor = self.next_pt_id - self.MAX_MAP_POINTS                         # vo.py:35-47
        if floor > 0:
            if self.dense_map and hasattr(self.map_points, "prune_below"):
                self.map_points.prune_below(floor)
            else:
                for pid in [p for p in self.map_points if p < floor]:
                    del self.map_points[pid]

    # ------------------------------------------------------------------ vo.py:290-299
    def _reset_system(self):
        self.initialized = False
        self.map_points = self._new_map()
        self.keyframe = None
        self.trajectory = []
        self.last_pos = np.zeros(3)
        self.baseline_speed = 1.0
