"""VO-side keyframe-window adapter (SURVEY §8 row b / row f-1).

The reference keeps ONE keyframe dict ``{"feats", "ids", "T_wc"}`` (`src/modules/vo.py:57,286`), a
landmark dict ``map_points: id -> xyz`` (`vo.py:17,281`) and the current pose ``T_wc`` (camera->world,
`vo.py:19,101,149`).  Bundle adjustment needs the last W keyframes, so this adapter

* wraps ``VisualOdometry._create_keyframe`` (`vo.py:252-288`) without touching the reference's source
  (``install(vo, ...)`` replaces the bound method on the instance),
* keeps the ``deque(maxlen=W)`` of keyframe dicts the reference itself creates (the very same dict
  objects, so a write-back into ``kf["T_wc"]`` is seen by the reference's next
  ``np.linalg.inv(self.keyframe["T_wc"])`` at `vo.py:260`),
* back-fills the ids of newly triangulated landmarks into the PREVIOUS keyframe's ``ids`` (the
  reference only tags the current frame, `vo.py:278-283`, although both frames observe the landmark),
* packs the window into the engine's arrays (poses ``inv(T_wc)`` as ``[R|t]``, observations
  ``(cam, point, uv)`` from ``kf["feats"]["keypoints"][0]`` / ``kf["ids"]``, landmarks from
  ``map_points``), runs ``BundleAdjuster`` and writes the result back to ``kf["T_wc"]``, ``vo.T_wc``,
  ``vo.last_pos`` (`vo.py:204`) and ``vo.map_points[pid]``.

Error behaviour follows the reference's style (no exceptions in the frame loop, `vo.py:241-245`): a
numerical failure (status 3) or a window that is too small only skips the write-back; it never calls
``_reset_system``.  Invalid arguments / CUDA errors still raise (they are programming errors).

All arithmetic is in libb2ba.so; this file only moves arrays.
"""
from __future__ import annotations

from collections import deque
from dataclasses import dataclass

import numpy as np

from .ba import BAConfig, BundleAdjuster, T_from_poses, poses_from_T


def _keypoints_numpy(feats) -> np.ndarray:
    """kf["feats"]["keypoints"] is a (1,N,2) float32 torch tensor on config.device
    (`frontend.py:47-49,66-70`); numpy arrays of the same shape are accepted too."""
    kp = feats["keypoints"]
    kp = kp[0]
    if hasattr(kp, "detach"):
        kp = kp.detach().cpu().numpy()
    return np.asarray(kp, np.float32).reshape(-1, 2)


@dataclass
class PackedWindow:
    """Engine-side arrays of one window plus what is needed to write the result back."""
    T_cw: np.ndarray        # (Nc,4,4) f64 world->camera
    points: np.ndarray      # (Np,3) f64
    obs_cam: np.ndarray     # (No,) i32
    obs_pt: np.ndarray      # (No,) i32
    obs_uv: np.ndarray      # (No,2) f32
    fixed_cam: np.ndarray   # (Nc,) bool
    pids: np.ndarray        # (Np,) int64 landmark ids (slot -> id)
    obs_kp: np.ndarray = None   # (No,) keypoint index of every observation in its keyframe (for culling)

    @property
    def poses(self):
        return poses_from_T(self.T_cw)


def pack_window(window, map_points, n_fixed: int = 2, min_obs: int = 2) -> PackedWindow | None:
    """Keyframe dicts + landmark dict -> BA arrays.

    Landmarks seen by fewer than ``min_obs`` window keyframes are left out (they do not constrain
    the cameras); duplicate (keyframe, landmark) pairs keep the first keypoint (the engine rejects
    duplicates, include/b2ba.h B2BA_E_DUPLICATE).  Slots are assigned in increasing id order so the
    packing is deterministic.  The `n_fixed` oldest keyframes are gauge-fixed (default 2: a monocular window
    has 7 gauge freedoms, two fixed poses pin all of them incl. the scale; at least one camera stays free).
    Returns None when nothing can be optimised.
    """
    kfs = list(window)
    nc = len(kfs)
    if nc < 2:
        return None
    cams, pids, uvs, kps = [], [], [], []
    for j, kf in enumerate(kfs):
        ids = np.asarray(kf["ids"]).astype(np.int64)
        kp = _keypoints_numpy(kf["feats"])
        sel = np.flatnonzero(ids >= 0)
        if sel.size == 0:
            continue
        # one observation per (keyframe, landmark): keep the first keypoint carrying the id
        _, first = np.unique(ids[sel], return_index=True)
        sel = sel[np.sort(first)]
        cams.append(np.full(sel.size, j, np.int32))
        pids.append(ids[sel])
        uvs.append(kp[sel])
        kps.append(sel)
    if not cams:
        return None
    cam = np.concatenate(cams)
    pid = np.concatenate(pids)
    uv = np.concatenate(uvs).astype(np.float32)
    kpi = np.concatenate(kps)
    upid, inv, cnt = np.unique(pid, return_inverse=True, return_counts=True)
    # landmarks that still exist in the map (`_prune_map`, vo.py:35-47, may have dropped old ids): looked up once per
    # landmark, not once per observation
    keep_pt = cnt >= min_obs
    cand = np.flatnonzero(keep_pt)                  # only landmarks that can enter the problem are looked up
    if cand.size == 0:
        return None
    dense = hasattr(map_points, "lookup")           # LandmarkTable: array gathers instead of per-id dict look-ups
    if dense:
        keep_pt[cand], _ = map_points.lookup(upid[cand])
    else:
        keep_pt[cand] = np.fromiter((p in map_points for p in upid[cand].tolist()), bool, cand.size)
    if not keep_pt.any():
        return None
    slot = np.cumsum(keep_pt) - 1
    keep_obs = keep_pt[inv]
    obs_pt = slot[inv[keep_obs]].astype(np.int32)
    obs_cam = cam[keep_obs]
    obs_uv = np.ascontiguousarray(uv[keep_obs])
    upid = upid[keep_pt]
    if dense:
        points = np.ascontiguousarray(map_points.lookup(upid)[1])
    else:
        points = np.array([map_points[p] for p in upid.tolist()], np.float64).reshape(-1, 3)
    T_cw = np.stack([np.linalg.inv(np.asarray(kf["T_wc"], np.float64)) for kf in kfs])
    fixed = np.zeros(nc, bool)
    fixed[:max(1, min(n_fixed, nc - 1))] = True
    return PackedWindow(T_cw, points, obs_cam, obs_pt, obs_uv, fixed, upid, kpi[keep_obs])


def write_back(vo, window, packed: PackedWindow, T_cw: np.ndarray, points: np.ndarray) -> None:
    """Refined poses / landmarks -> the reference's state (kf["T_wc"], vo.T_wc, vo.last_pos, vo.map_points)."""
    kfs = list(window)
    for j, kf in enumerate(kfs):
        if not packed.fixed_cam[j]:
            kf["T_wc"] = np.linalg.inv(T_cw[j])
    newest = kfs[-1]
    vo.T_wc = newest["T_wc"].copy()
    vo.last_pos = vo.T_wc[:3, 3].copy()          # vo.py:204
    mp = vo.map_points
    if hasattr(mp, "update_rows"):
        mp.update_rows(packed.pids, points)
        return
    rows = np.array(points, np.float64)             # one copy; every landmark gets its own row (an ndarray(3,) view)
    mp.update((p, x) for p, x in zip(packed.pids.tolist(), rows) if p in mp)


def cull_window(vo, window, packed: PackedWindow, err: np.ndarray, max_err: float) -> int:
    """Outlier culling after BA with the reference's own landmark gate (`frontend.py:139-142`:
    reprojection error < `max_reproj_err`, `config.py:32`; depth gate `config.py:31` = err < 0 here): an
    observation that fails it loses its id in its keyframe (`kf["ids"][kp] = -1`, the reference's "untracked"
    marker, `vo.py:53`), a landmark whose window observations ALL fail is dropped from `vo.map_points`.
    Returns the number of observations culled."""
    bad = (err < 0) | (err > max_err)
    if not bad.any():
        return 0
    kfs = list(window)
    for j, kf in enumerate(kfs):
        m = bad & (packed.obs_cam == j)
        if m.any():
            kf["ids"][packed.obs_kp[m]] = -1
    left = np.bincount(packed.obs_pt[~bad], minlength=len(packed.pids))
    for p in packed.pids[left == 0].tolist():
        vo.map_points.pop(int(p), None)
    return int(bad.sum())


def cull_window_device(vo, window, packed: PackedWindow, ba, max_err: float) -> int:
    """`cull_window` with the decision made on the device (`b2ba_cull`: error, gate, per-landmark survivor counts in
    one kernel); the host only edits the reference's own structures (keyframe id arrays, the landmark dict).  Same
    result as `cull_window(vo, window, packed, ba.reproj_errors(), max_err)`, bit for bit (tests/test_cull.py)."""
    keep, live, _ = ba.cull(max_err)
    bad = ~keep
    if not bad.any():
        return 0
    kfs = list(window)
    for j in np.unique(packed.obs_cam[bad]).tolist():
        m = bad & (packed.obs_cam == j)
        kfs[j]["ids"][packed.obs_kp[m]] = -1
    for p in packed.pids[live == 0].tolist():
        vo.map_points.pop(int(p), None)
    return int(bad.sum())


def refine_window(vo, window, ba: BundleAdjuster, n_fixed: int = 2, cull_px: float | None = None):
    """Pack -> BundleAdjuster.optimize -> write back (-> cull).  Returns the BASummary (or None if skipped)."""
    packed = pack_window(window, vo.map_points, n_fixed=n_fixed)
    if packed is None:
        return None
    res = ba.optimize(packed.T_cw, packed.points, packed.obs_cam, packed.obs_pt, packed.obs_uv, packed.fixed_cam)
    if not res.summary.ok or not np.isfinite(res.T_cw).all() or not np.isfinite(res.points).all():
        return res.summary                        # numerical failure: keep the reference's state
    write_back(vo, window, packed, res.T_cw, res.points)
    if cull_px is not None and hasattr(ba, "cull"):
        cull_window_device(vo, window, packed, ba, cull_px)
    elif cull_px is not None and hasattr(ba, "reproj_errors"):
        cull_window(vo, window, packed, ba.reproj_errors(), cull_px)
    return res.summary


class WindowAdapter:
    """Keyframe window + BA call, hooked onto one VisualOdometry instance."""

    def __init__(self, vo, config: BAConfig | None = None, ba=None, n_fixed: int = 2, cull_px: float | None = None):
        self.vo = vo
        self.cull_px = cull_px                    # e.g. config.max_reproj_err (config.py:32); None = no culling
        self.config = config or BAConfig()
        self.ba = ba if ba is not None else BundleAdjuster(vo.K, self.config)
        self.n_fixed = n_fixed
        self.window = deque(maxlen=self.config.window)
        self.summaries = []
        self._orig = None

    # -- the wrapped method: same signature as VisualOdometry._create_keyframe (vo.py:252-258)
    def _create_keyframe(self, curr_feats, curr_ids, ref_indices, curr_indices) -> None:
        vo = self.vo
        prev = vo.keyframe
        self._orig(curr_feats, curr_ids, ref_indices, curr_indices)
        new = vo.keyframe
        if new is None or new is prev:
            return
        if prev is None or not self.window or self.window[-1] is not prev:
            self.window.clear()                   # first keyframe pair, or the VO was reset (vo.py:290-299)
            if prev is not None:
                self.window.append(prev)
        if prev is not None:
            backfill_ids(prev, new, ref_indices, curr_indices)
        self.window.append(new)
        self.summaries.append(refine_window(vo, self.window, self.ba, self.n_fixed, self.cull_px))

    def install(self):
        self._orig = self.vo._create_keyframe
        self.vo._create_keyframe = self._create_keyframe
        return self

    def uninstall(self):
        if self._orig is not None:
            self.vo._create_keyframe = self._orig
            self._orig = None


def backfill_ids(prev_kf, new_kf, ref_indices, curr_indices) -> int:
    """Give the previous keyframe the ids of the landmarks just triangulated from the (prev, new)
    matches: the reference tags only the new frame (`vo.py:278-283`).  Returns how many were set."""
    pid_new = np.asarray(new_kf["ids"])[np.asarray(curr_indices, np.int64)]
    ref = np.asarray(ref_indices, np.int64)
    pid_prev = prev_kf["ids"][ref]
    m = (pid_prev < 0) & (pid_new >= 0)
    prev_kf["ids"][ref[m]] = pid_new[m]
    return int(m.sum())


def install(vo, config: BAConfig | None = None, ba=None, n_fixed: int = 2, cull_px: float | None = None) -> WindowAdapter:
    """The one call a maintainer adds after constructing VisualOdometry."""
    return WindowAdapter(vo, config, ba, n_fixed, cull_px).install()
