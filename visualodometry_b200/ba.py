"""Host-side mirror of the optimiser interface (SURVEY §8 row b).

The reference has no optimiser API; modules are plain classes taking ``(K, config)``
(`src/modules/vo.py:10`, `src/modules/frontend.py:10`) and functions take / return numpy f64
arrays plus a config dataclass (`frontend.py:115-122`, `src/config/config.py:4-46`).
``BundleAdjuster(K, BAConfig)`` follows that style.  All arithmetic happens in libb2ba.so
(CUDA, sm_100a) through the C ABI in include/b2ba.h; this file only packs arrays.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field

import numpy as np

from . import _lib

try:  # torch is plumbing (device memory / streams); numpy inputs work without it
    import torch
except Exception:  # pragma: no cover
    torch = None


@dataclass
class BAConfig:
    """BA options, same dataclass style as the reference's VOConfig (config.py:4-46)."""

    window: int = 10               # keyframes kept by the adapter's deque
    max_iters: int = 10
    loss: str = "trivial"          # "trivial" | "huber"
    loss_scale: float = 1.0        # Huber a, pixels (cf. config.py:32,35 pixel thresholds)
    z_min: float = 1e-3            # = VOConfig.min_depth (config.py:31)
    init_radius: float = 1e4
    ftol: float = 0.0
    gtol: float = 0.0
    ptol: float = 0.0
    solver: str = "auto"           # "auto" | "chol" (window kernel) | "pcg" | "dense" (large path: PCG / dense DMMA Cholesky)
    pcg_max_iters: int = 200
    pcg_tol: float = 1e-10
    device: int = 0
    record_history: bool = False
    cluster: bool = True           # few windows: one 8-CTA cluster per window (latency path); False = one CTA per window
    variant: str = "auto"          # "auto" | "small256" | "cluster8": force a window-kernel instantiation (tests)
    time_phases: bool = False      # CUDA-event timing of set-up / solve / LM phases (BundleAdjuster.timings())

    def to_options(self) -> _lib.Options:
        return _lib.Options(self.max_iters, {"trivial": 0, "huber": 1}[self.loss], self.loss_scale,
                            self.z_min, self.init_radius, self.ftol, self.gtol, self.ptol,
                            {"auto": 0, "chol": 1, "pcg": 2, "dense": 3}[self.solver], self.pcg_max_iters,
                            self.pcg_tol, (1 if self.record_history else 0) | (0 if self.cluster else 2)
                            | {"auto": 0, "small256": 4, "cluster8": 8}[self.variant] | (16 if self.time_phases else 0), 0)


@dataclass
class BASummary:
    status: int
    iterations: int
    accepted: int
    rejected: int
    n_invalid: int
    pcg_iters: int
    initial_cost: float
    final_cost: float
    final_mu: float

    @property
    def ok(self):  # numerical failure -> the VO adapter skips the write-back
        return self.status in (_lib.MAX_ITERS, _lib.CONVERGED)


class BAResult:
    """T_cw (Nc,4,4) f64, points (Np,3) f64, summary (the worst of the batch), summaries (per window).  The 4 x 4
    matrices are built from the engine's [R|t] rows on first access (a 4096-window batch has 40960 of them)."""

    def __init__(self, T_cw, points, summary, summaries=None, poses=None):
        self._T, self._poses = T_cw, poses
        self.points, self.summary = points, summary
        self.summaries = summaries if summaries is not None else []

    @property
    def T_cw(self) -> np.ndarray:
        if self._T is None:
            self._T = T_from_poses(self._poses)
        return self._T

    @property
    def poses(self) -> np.ndarray:
        return self._poses if self._poses is not None else poses_from_T(self._T)


class SummaryList:
    """Per-window summaries of a batch call: a sequence of BASummary over the C array the call filled, converted on
    access (4096 Python objects per call are not free); `.array` is the structured numpy view of the same memory."""
    DTYPE = np.dtype([("status", "<i4"), ("iterations", "<i4"), ("accepted", "<i4"), ("rejected", "<i4"), ("n_invalid", "<i4"),
                      ("pcg_iters", "<i4"), ("initial_cost", "<f8"), ("final_cost", "<f8"), ("final_mu", "<f8")])

    def __init__(self, carray):
        self._c = carray
        self.array = np.frombuffer(carray, dtype=self.DTYPE) if len(carray) else np.empty(0, self.DTYPE)

    def __len__(self):
        return len(self._c)

    def __getitem__(self, i):
        if isinstance(i, slice):
            return [self[j] for j in range(*i.indices(len(self)))]
        return BundleAdjuster._summary(self._c[i])

    def __iter__(self):
        return (BundleAdjuster._summary(x) for x in self._c)

    def worst(self):
        return self[int(np.argmax(self.array["status"]))]


def poses_from_T(T_cw) -> np.ndarray:
    """(Nc,4,4) world->camera matrices -> (Nc,12) [R row-major | t]."""
    T = np.asarray(T_cw, np.float64).reshape(-1, 4, 4)
    return np.ascontiguousarray(np.concatenate([T[:, :3, :3].reshape(-1, 9), T[:, :3, 3]], axis=1))


def T_from_poses(poses) -> np.ndarray:
    P = np.asarray(poses, np.float64).reshape(-1, 12)
    T = np.tile(np.eye(4), (P.shape[0], 1, 1))
    T[:, :3, :3] = P[:, :9].reshape(-1, 3, 3)
    T[:, :3, 3] = P[:, 9:]
    return T


def _is_cuda_tensor(a):
    return torch is not None and isinstance(a, torch.Tensor) and a.is_cuda


class _Arg:
    """Pointer + keep-alive for one array argument, host (numpy) or device (torch CUDA)."""

    def __init__(self, a, np_dtype, torch_dtype, device_mode, device=0):
        if device_mode:   # on the ADJUSTER's device (config.device), not torch's current one
            t = a if isinstance(a, torch.Tensor) else torch.as_tensor(np.asarray(a))
            t = t.to(device=torch.device("cuda", device), dtype=getattr(torch, torch_dtype)).contiguous()
            self.keep, self.ptr = t, C.c_void_p(t.data_ptr())
        else:
            if torch is not None and isinstance(a, torch.Tensor):
                a = a.detach().cpu().numpy()
            n = np.ascontiguousarray(a, dtype=np_dtype)
            self.keep, self.ptr = n, C.c_void_p(n.ctypes.data)


class BundleAdjuster:
    """Windowed / batched / large-scale BA on one B200.  Drop-in for the (missing) optimiser
    stage at the tail of VisualOdometry._create_keyframe (vo.py:286-288)."""

    def __init__(self, K: np.ndarray, config: BAConfig | None = None):
        self.config = config or BAConfig()
        K = np.asarray(K, np.float64)
        self.intr = np.array([K[0, 0], K[1, 1], K[0, 2], K[1, 2]], np.float64)   # zero skew/distortion (vo.py:139)
        self._lib = _lib.load()
        self._h = C.c_void_p()
        opts = self.config.to_options()
        rc = self._lib.b2ba_create(C.byref(self._h), self.config.device, C.byref(opts))
        if rc != _lib.OK:
            raise RuntimeError(f"b2ba_create failed ({rc}): a CUDA device is required (no CPU fallback)")
        self._nb = 0
        self._shape = None
        self.rank, self.nranks = 0, 1

    def init_distributed(self, group=None):
        """Observation-sharded solves (C5) across the ranks of a torch.distributed process group:
        rank 0 creates an ncclUniqueId, it is broadcast through the group, and every rank builds the
        engine's own NCCL communicator (NVLink / NVSwitch) from it."""
        import torch.distributed as dist
        from . import shard  # noqa: F401
        self.rank, self.nranks = dist.get_rank(group), dist.get_world_size(group)
        if self.nranks == 1:
            return self
        uid = np.zeros(128, np.uint8)
        if self.rank == 0:
            self._check(self._lib.b2ba_nccl_unique_id(uid.ctypes.data), "b2ba_nccl_unique_id")
        dev = "cuda" if dist.get_backend(group) == "nccl" else "cpu"
        t = torch.from_numpy(uid).to(dev)
        dist.broadcast(t, src=0, group=group)
        uid = t.cpu().numpy().copy()
        self._check(self._lib.b2ba_create_comm(self._h, uid.ctypes.data, self.rank, self.nranks), "b2ba_create_comm")
        return self

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            self._lib.b2ba_destroy(self._h)
            self._h = C.c_void_p()

    __del__ = close

    # ------------------------------------------------------------------ low level
    def _check(self, rc, what):
        if rc < 0:
            msg = self._lib.b2ba_last_error(self._h)
            raise RuntimeError(f"{what} failed ({rc}): {msg.decode() if msg else ''}")
        return rc

    def set_problem(self, poses, points, obs_cam, obs_pt, obs_uv, fixed_cam, cam_off=None, pt_off=None,
                    obs_off=None):
        """poses (Nc,12) [R|t] world->camera; batched when the CSR offsets are given."""
        dev = any(_is_cuda_tensor(a) for a in (poses, points, obs_cam, obs_pt, obs_uv))
        a = [_Arg(poses, np.float64, "float64", dev, self.config.device), _Arg(points, np.float64, "float64", dev, self.config.device),
             _Arg(obs_cam, np.int32, "int32", dev, self.config.device), _Arg(obs_pt, np.int32, "int32", dev, self.config.device),
             _Arg(obs_uv, np.float32, "float32", dev, self.config.device), _Arg(fixed_cam, np.uint8, "uint8", dev, self.config.device)]
        nc = int(np.prod(a[0].keep.shape)) // 12
        npts = int(np.prod(a[1].keep.shape)) // 3
        no = int(np.prod(a[2].keep.shape))
        intr = self.intr.ctypes.data_as(C.POINTER(C.c_double))
        mem = _lib.MEM_DEVICE if dev else _lib.MEM_HOST
        if cam_off is None:
            if self.nranks > 1:
                from .shard import point_ranges
                op = obs_pt.detach().cpu().numpy() if _is_cuda_tensor(obs_pt) else np.asarray(obs_pt)
                lo, hi = point_ranges(op, npts, self.nranks)[self.rank]
                self._check(self._lib.b2ba_set_point_range(self._h, lo, hi), "b2ba_set_point_range")
            rc = self._lib.b2ba_set_problem(self._h, nc, npts, no, *[x.ptr for x in a], intr, mem)
            self._nb = 1
        else:
            co = np.ascontiguousarray(cam_off, np.int32)
            po = np.ascontiguousarray(pt_off, np.int32)
            oo = np.ascontiguousarray(obs_off, np.int64)
            rc = self._lib.b2ba_set_problem_batch(self._h, len(co) - 1, co.ctypes.data, po.ctypes.data,
                                                  oo.ctypes.data, *[x.ptr for x in a], intr, mem)
            self._nb = len(co) - 1
        self._check(rc, "b2ba_set_problem")
        self._shape = (nc, npts, no)
        return self

    def reset(self):
        self._check(self._lib.b2ba_reset(self._h), "b2ba_reset")

    def solve(self, max_iters=0, stream=None, wait=True):
        """Run LM. wait=False only enqueues (summaries via .summaries())."""
        sp = C.c_void_p(stream) if stream else None
        if not wait:
            self._check(self._lib.b2ba_solve(self._h, max_iters, None, sp), "b2ba_solve")
            return None
        sm = (_lib.Summary * self._nb)()
        self._check(self._lib.b2ba_solve(self._h, max_iters, sm, sp), "b2ba_solve")
        return [self._summary(s) for s in sm]

    def summaries(self):
        sm = (_lib.Summary * self._nb)()
        self._check(self._lib.b2ba_get_summaries(self._h, sm), "b2ba_get_summaries")
        return [self._summary(s) for s in sm]

    def solve_noalloc(self, poses, points, summaries, max_iters=0, stream=None):
        """`b2ba_solve_window_noalloc` (no counterpart in the reference; added for callers that cannot call cudaMalloc):
        the window solve of the problem set with set_problem, scratch in the library's static device array, nothing
        allocated, nothing synchronised.  poses (Nc,12) / points (Np,3): CUDA float64 tensors, refined in place;
        summaries: CUDA uint8 tensor of >= 48 * B bytes (b2ba_summary per window).  At most
        `_lib.NOALLOC_MAX_POINTS` landmarks; overlapping calls on one GPU race (one shared scratch array)."""
        for t, n in ((poses, self._shape[0] * 12), (points, self._shape[1] * 3)):
            if not _is_cuda_tensor(t) or t.dtype != torch.float64 or t.numel() != n or not t.is_contiguous():
                raise ValueError("solve_noalloc: contiguous CUDA float64 tensors of the problem's size required")
        if not _is_cuda_tensor(summaries) or summaries.numel() * summaries.element_size() < 48 * self._nb:
            raise ValueError("solve_noalloc: summaries must be a CUDA buffer of 48 bytes per window")
        rc = self._lib.b2ba_solve_window_noalloc(self._h, poses.data_ptr(), points.data_ptr(), summaries.data_ptr(),
                                                 max_iters, C.c_void_p(stream) if stream else None)
        self._check(rc, "b2ba_solve_window_noalloc")

    def timings(self) -> dict:
        """Device times of the last set_problem / solve (config.time_phases): {"setup_ms", "solve_ms", "phases": {...}}."""
        t = _lib.Timings()
        self._check(self._lib.b2ba_get_timings(self._h, C.byref(t)), "b2ba_get_timings")
        return {"setup_ms": t.setup_ms, "solve_ms": t.solve_ms, "iterations": t.iterations_enqueued,
                "phases": {k: t.phase_ms[i] for i, k in enumerate(_lib.PHASES)}}

    @staticmethod
    def _summary(s):
        return BASummary(s.status, s.iterations, s.accepted, s.rejected, s.n_invalid, s.pcg_iters,
                         s.initial_cost, s.final_cost, s.final_mu)

    def get(self, device=False, out=None):
        """Current poses (Nc,12) / landmarks (Np,3).  `out=(poses, points)`: caller-owned C-contiguous f64
        numpy arrays (e.g. pinned host memory) to fill instead of allocating new ones."""
        nc, npts, _ = self._shape
        if out is not None:
            poses, pts = out
            for a, shp in ((poses, nc * 12), (pts, npts * 3)):
                if a.dtype != np.float64 or a.size != shp or not a.flags.c_contiguous:
                    raise ValueError("get(out=...): C-contiguous float64 arrays of the problem's size required")
            self._check(self._lib.b2ba_get(self._h, poses.ctypes.data, pts.ctypes.data, _lib.MEM_HOST), "b2ba_get")
            return poses, pts
        if device:
            poses = torch.empty((nc, 12), dtype=torch.float64, device=torch.device("cuda", self.config.device))
            pts = torch.empty((npts, 3), dtype=torch.float64, device=torch.device("cuda", self.config.device))
            self._check(self._lib.b2ba_get(self._h, poses.data_ptr(), pts.data_ptr(), _lib.MEM_DEVICE), "b2ba_get")
        else:
            poses, pts = np.empty((nc, 12)), np.empty((npts, 3))
            self._check(self._lib.b2ba_get(self._h, poses.ctypes.data, pts.ctypes.data, _lib.MEM_HOST), "b2ba_get")
        return poses, pts

    def triangulate(self, T_cw1, T_cw2, pts1, pts2, min_depth=1e-3, max_reproj_err=5.0):
        """`triangulate_points(T_cw1, T_cw2, pts1, pts2, K, config)` of the reference (`frontend.py:115-148`) on the
        GPU: returns (X[keep] (M,3) f64, keep (N,) bool) like the reference; `K` is the adjuster's."""
        n = len(pts1)
        if n == 0:
            return np.empty((0, 3)), np.zeros(0, dtype=bool)
        K = np.array([[self.intr[0], 0, self.intr[2]], [0, self.intr[1], self.intr[3]], [0, 0, 1.0]])
        P1 = np.ascontiguousarray(K @ np.asarray(T_cw1, np.float64)[:3, :])
        P2 = np.ascontiguousarray(K @ np.asarray(T_cw2, np.float64)[:3, :])
        dev = _is_cuda_tensor(pts1) or _is_cuda_tensor(pts2)
        a, b = _Arg(pts1, np.float32, "float32", dev, self.config.device), _Arg(pts2, np.float32, "float32", dev, self.config.device)
        dp = C.POINTER(C.c_double)
        if dev:
            X = torch.empty((n, 3), dtype=torch.float64, device=torch.device("cuda", self.config.device)); keep = torch.empty(n, dtype=torch.uint8, device=torch.device("cuda", self.config.device))
            xp, kp = C.c_void_p(X.data_ptr()), C.c_void_p(keep.data_ptr())
        else:
            X = np.empty((n, 3), np.float64); keep = np.empty(n, np.uint8)
            xp, kp = C.c_void_p(X.ctypes.data), C.c_void_p(keep.ctypes.data)
        rc = self._lib.b2ba_triangulate(self._h, n, P1.ctypes.data_as(dp), P2.ctypes.data_as(dp), a.ptr, b.ptr,
                                        float(min_depth), float(max_reproj_err), xp, kp,
                                        _lib.MEM_DEVICE if dev else _lib.MEM_HOST)
        self._check(rc, "b2ba_triangulate")
        if dev:
            X, keep = X.cpu().numpy(), keep.cpu().numpy()
        keep = keep.astype(bool)
        return X[keep], keep

    def reproj_errors(self) -> np.ndarray:
        """(No,) reprojection error in pixels of every observation at the current (after solve: refined) state, in
        the caller's observation order; -1 behind the cheirality gate.  The reference's landmark gate quantity
        (`frontend.py:139-142`), used for culling after BA (`adapter.cull_window`)."""
        if self._shape is None:
            raise RuntimeError("reproj_errors: no problem set")
        out = np.empty(self._shape[2], np.float64)
        self._check(self._lib.b2ba_reproj_errors(self._h, out.ctypes.data, _lib.MEM_HOST), "b2ba_reproj_errors")
        return out

    def cull(self, max_err: float, apply: bool = False):
        """Outlier culling after BA on the device (`b2ba_cull`): the reference's landmark gate
        (`frontend.py:139-142`, `config.py:31-32`) applied to every observation at the current (refined) state.
        Returns (keep (No,) bool, live (Np,) int32 = surviving observations per landmark, kept_index (n_kept,) int32).
        `apply=True` makes the culled problem the handle's problem (compacted on the device, index rebuilt), so the
        next `solve()` runs on it without passing the observation arrays through the host again."""
        if self._shape is None:
            raise RuntimeError("cull: no problem set")
        nc, npts, no = self._shape
        keep = np.empty(no, np.uint8)
        live = np.empty(npts, np.int32)
        idx = np.empty(no, np.int32)
        n = C.c_int64(0)
        self._check(self._lib.b2ba_cull(self._h, float(max_err), keep.ctypes.data, live.ctypes.data, idx.ctypes.data,
                                        C.byref(n), 1 if apply else 0, _lib.MEM_HOST), "b2ba_cull")
        if apply:
            self._shape = (nc, npts, int(n.value))
        return keep.astype(bool), live, idx[: int(n.value)].copy()

    def dump(self, what: str, shape, dtype=np.float64):
        out = np.zeros(shape, dtype)
        n = self._lib.b2ba_dump(self._h, _lib.D[what], out.ctypes.data, out.nbytes)
        self._check(n, f"b2ba_dump({what})")
        if n != out.nbytes:
            raise RuntimeError(f"b2ba_dump({what}): wrote {n} bytes, expected {out.nbytes}")
        return out

    # ------------------------------------------------------------------ the drop-in call
    def optimize(self, T_cw, points, obs_cam, obs_pt, obs_uv, fixed_cam) -> BAResult:
        """T_cw (Nc,4,4) f64, points (Np,3) f64, obs_cam/obs_pt (No,) i32, obs_uv (No,2) f32,
        fixed_cam (Nc,) bool -> refined T_cw / points + summary.  numpy (host) or torch CUDA."""
        if _is_cuda_tensor(T_cw):
            poses = torch.cat([T_cw[:, :3, :3].reshape(-1, 9), T_cw[:, :3, 3]], dim=1).contiguous()
        else:
            poses = poses_from_T(T_cw)
        fixed = fixed_cam if _is_cuda_tensor(fixed_cam) else np.asarray(fixed_cam).astype(np.uint8)
        self.set_problem(poses, points, obs_cam, obs_pt, obs_uv, fixed)
        sm = self.solve()
        poses_o, pts_o = self.get()
        return BAResult(T_from_poses(poses_o), pts_o, sm[0], sm)

    # ------------------------------------------------------------------ motion-only BA
    def refine_poses(self, T_cw, points, obs_uv, obs_off=None, max_iters=0):
        """Refine B camera poses against FIXED landmarks: the step the reference leaves to cv2.solvePnPRansac
        (`vo.py:131-150`: `X` = map points of the matched ids, `uv` = current keypoints).  T_cw (B,4,4) or (4,4);
        points (No,3) = the landmark of every observation; obs_uv (No,2); obs_off (B+1,) CSR offsets (one problem
        when omitted).  numpy (host) or torch CUDA arrays.  Returns (T_cw (B,4,4), [BASummary])."""
        dev = any(_is_cuda_tensor(a) for a in (points, obs_uv))
        if _is_cuda_tensor(T_cw):
            T_cw = T_cw.detach().cpu().numpy()
        poses = poses_from_T(T_cw)
        nb = poses.shape[0]
        a_pts, a_uv = _Arg(points, np.float64, "float64", dev, self.config.device), _Arg(obs_uv, np.float32, "float32", dev, self.config.device)
        no = int(np.prod(a_pts.keep.shape)) // 3
        off = np.array([0, no], np.int64) if obs_off is None else np.ascontiguousarray(obs_off, np.int64)
        if off.shape != (nb + 1,) or int(off[-1]) != no or int(np.prod(a_uv.keep.shape)) != 2 * no:
            raise ValueError("refine_poses: obs_off must have B+1 entries ending at the observation count")
        a_pose = _Arg(poses, np.float64, "float64", dev, self.config.device)
        sms = (_lib.Summary * max(nb, 1))()
        rc = self._lib.b2ba_refine_poses(self._h, nb, off.ctypes.data, a_pose.ptr, a_pts.ptr, a_uv.ptr,
                                         self.intr.ctypes.data_as(C.POINTER(C.c_double)), int(max_iters), sms,
                                         _lib.MEM_DEVICE if dev else _lib.MEM_HOST)
        self._check(rc, "b2ba_refine_poses")
        out = a_pose.keep.cpu().numpy() if dev else a_pose.keep
        return T_from_poses(out), [self._summary(sms[i]) for i in range(nb)]

    def refine_pose(self, T_cw, points, obs_uv, max_iters=0):
        """One pose: (4,4) in, (4,4) out + its summary."""
        T, sm = self.refine_poses(np.asarray(T_cw, np.float64).reshape(1, 4, 4), points, obs_uv, None, max_iters)
        return T[0], sm[0]

    def track_pose(self, T_cw_prior, points, obs_uv, reproj_err=1.0, rounds=6, iters_per_round=10):
        """Robust pose tracking from a prior: the GPU stand-in for the reference's
        `cv2.solvePnPRansac(X, uv, K, None, reprojectionError=pnp_reproj_err)` (`vo.py:135-141`, `config.py:35`) once the
        tracker has a previous pose.  `rounds` x [motion-only LM on the current inliers (Huber a = thr_r; last
        round: trivial loss) -> re-classify every observation by |reprojection error| < thr_r], thr_r = reproj_err x
        2^(rounds - 2 - r) >= reproj_err (graduated: 16, 8, 4, 2, 1, 1 for 6 rounds).  Deterministic
        (no sampling).  Returns (ok, T_cw (4,4), inliers (n,) int -- indices like solvePnPRansac's, BASummary);
        ok is False when fewer than 6 inliers remain or the pose is not finite (the caller resets, `vo.py:142-143`)."""
        dev = any(_is_cuda_tensor(a) for a in (points, obs_uv))
        pose = poses_from_T(np.asarray(T_cw_prior, np.float64).reshape(1, 4, 4))
        a_pts, a_uv = _Arg(points, np.float64, "float64", dev, self.config.device), _Arg(obs_uv, np.float32, "float32", dev, self.config.device)
        n = int(np.prod(a_pts.keep.shape)) // 3
        if int(np.prod(a_uv.keep.shape)) != 2 * n:
            raise ValueError("track_pose: points (N,3) and obs_uv (N,2) required")
        a_pose = _Arg(pose, np.float64, "float64", dev, self.config.device)
        a_inl = _Arg(np.zeros(max(n, 1), np.uint8), np.uint8, "uint8", dev, self.config.device)
        cnt, sm = C.c_int32(0), _lib.Summary()
        rc = self._lib.b2ba_track_pose(self._h, n, a_pose.ptr, a_pts.ptr, a_uv.ptr,
                                       self.intr.ctypes.data_as(C.POINTER(C.c_double)), float(reproj_err), int(rounds),
                                       int(iters_per_round), a_inl.ptr, C.byref(cnt), C.byref(sm),
                                       _lib.MEM_DEVICE if dev else _lib.MEM_HOST)
        self._check(rc, "b2ba_track_pose")
        out = a_pose.keep.cpu().numpy() if dev else a_pose.keep
        mask = (a_inl.keep.cpu().numpy() if dev else a_inl.keep)[:n]
        T = T_from_poses(out)[0]
        inl = np.flatnonzero(mask)
        ok = bool(cnt.value >= 6 and np.isfinite(T).all())
        return ok, T, inl, self._summary(sm)

    def ransac_pose(self, T_cw_prior, points, obs_uv, reproj_err=1.0, n_hyp=1024, hyp_iters=8, seed=0, refine_iters=10):
        """Sampled-consensus pose on the GPU: the counterpart of the reference's
        `cv2.solvePnPRansac(X, uv, K, None, reprojectionError=pnp_reproj_err)` (`vo.py:135-141`) for a tracker with a
        previous pose.  `n_hyp` hypotheses at once (each the LM fit of 4 hashed-sampled matches from the prior), scored
        by consensus over all matches, the best one refined on its consensus set (include/b2ba.h b2ba_ransac_pose).
        Returns (ok, T_cw (4,4), inlier indices, BASummary, consensus of the winning hypothesis)."""
        dev = any(_is_cuda_tensor(a) for a in (points, obs_uv))
        pose = poses_from_T(np.asarray(T_cw_prior, np.float64).reshape(1, 4, 4))
        a_pts, a_uv = _Arg(points, np.float64, "float64", dev, self.config.device), _Arg(obs_uv, np.float32, "float32", dev, self.config.device)
        n = int(np.prod(a_pts.keep.shape)) // 3
        if int(np.prod(a_uv.keep.shape)) != 2 * n:
            raise ValueError("ransac_pose: points (N,3) and obs_uv (N,2) required")
        a_pose = _Arg(pose, np.float64, "float64", dev, self.config.device)
        a_inl = _Arg(np.zeros(max(n, 1), np.uint8), np.uint8, "uint8", dev, self.config.device)
        cnt, best, sm = C.c_int32(0), C.c_int32(0), _lib.Summary()
        rc = self._lib.b2ba_ransac_pose(self._h, n, a_pose.ptr, a_pts.ptr, a_uv.ptr,
                                        self.intr.ctypes.data_as(C.POINTER(C.c_double)), float(reproj_err), int(n_hyp),
                                        int(hyp_iters), int(seed), int(refine_iters), a_inl.ptr, C.byref(cnt),
                                        C.byref(best), C.byref(sm), _lib.MEM_DEVICE if dev else _lib.MEM_HOST)
        self._check(rc, "b2ba_ransac_pose")
        out = a_pose.keep.cpu().numpy() if dev else a_pose.keep
        mask = (a_inl.keep.cpu().numpy() if dev else a_inl.keep)[:n]
        T = T_from_poses(out)[0]
        ok = bool(cnt.value >= 6 and np.isfinite(T).all())
        return ok, T, np.flatnonzero(mask), self._summary(sm), int(best.value)

    def optimize_batch(self, batch, out=None, chunks=0) -> BAResult:
        """C4: a synth.BABatch-like object with CSR offsets; poses in (Nc,12) layout.  Host arrays go through the
        pipelined C-ABI call (copies of one chunk of windows overlap the solve of another); `out=(poses, points)`
        are optional caller-owned (e.g. pinned) result arrays."""
        arrs = (batch.poses, batch.points, batch.obs_cam, batch.obs_pt, batch.obs_uv, batch.fixed_cam)
        if self.config.record_history or any(_is_cuda_tensor(a) for a in arrs):    # history dumps need handle state
            self.set_problem(*arrs, batch.cam_off, batch.pt_off, batch.obs_off)
            sm = self.solve()
            poses_o, pts_o = self.get()
        else:
            a = [_Arg(batch.poses, np.float64, "float64", False, self.config.device), _Arg(batch.points, np.float64, "float64", False, self.config.device),
                 _Arg(batch.obs_cam, np.int32, "int32", False, self.config.device), _Arg(batch.obs_pt, np.int32, "int32", False, self.config.device),
                 _Arg(batch.obs_uv, np.float32, "float32", False, self.config.device), _Arg(batch.fixed_cam, np.uint8, "uint8", False, self.config.device)]
            co = np.ascontiguousarray(batch.cam_off, np.int32)
            po = np.ascontiguousarray(batch.pt_off, np.int32)
            oo = np.ascontiguousarray(batch.obs_off, np.int64)
            nb, nc, npts = len(co) - 1, int(co[-1]), int(po[-1])
            poses_o, pts_o = out if out is not None else (np.empty((nc, 12)), np.empty((npts, 3)))
            for arr, size in ((poses_o, nc * 12), (pts_o, npts * 3)):
                if arr.dtype != np.float64 or arr.size != size or not arr.flags.c_contiguous:
                    raise ValueError("optimize_batch(out=...): C-contiguous float64 arrays of the problem's size required")
            sms = (_lib.Summary * nb)()
            rc = self._lib.b2ba_optimize_batch(self._h, nb, co.ctypes.data, po.ctypes.data, oo.ctypes.data,
                                               *[x.ptr for x in a], self.intr.ctypes.data_as(C.POINTER(C.c_double)), 0,
                                               poses_o.ctypes.data, pts_o.ctypes.data, sms, chunks)
            self._check(rc, "b2ba_optimize_batch")
            assert C.sizeof(_lib.Summary) == SummaryList.DTYPE.itemsize
            sm = SummaryList(sms)
            self._nb, self._shape = 0, None
            return BAResult(None, pts_o, sm.worst(), sm, poses=poses_o)
        worst = max(sm, key=lambda s: s.status)
        return BAResult(None, pts_o, worst, sm, poses=poses_o)
