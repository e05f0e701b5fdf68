"""ctypes binding of the CPU oracle (oracle/ba_cpu.cpp).  TEST INFRASTRUCTURE ONLY: imported by
tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs; never by
visualodometry_b200/.  PARITY UNPINNED (the reference ships no BA; see ba_cpu.cpp header)."""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from dataclasses import dataclass

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


class Opts(C.Structure):
    _fields_ = [("max_iters", C.c_int32), ("loss", C.c_int32), ("loss_scale", C.c_double),
                ("z_min", C.c_double), ("init_mu", C.c_double), ("ftol", C.c_double),
                ("gtol", C.c_double), ("ptol", C.c_double), ("solver", C.c_int32),
                ("pcg_max_iters", C.c_int32), ("pcg_tol", C.c_double), ("threads", C.c_int32),
                ("flags", C.c_int32)]


class Summary(C.Structure):
    _fields_ = [("status", C.c_int32), ("iterations", C.c_int32), ("accepted", C.c_int32),
                ("rejected", C.c_int32), ("n_invalid", C.c_int32), ("pcg_iters_total", C.c_int32),
                ("initial_cost", C.c_double), ("final_cost", C.c_double), ("final_mu", C.c_double),
                ("cost_hist", C.c_double * 64), ("mu_hist", C.c_double * 64),
                ("accept_hist", C.c_int32 * 64)]


def make_opts(max_iters=10, loss="trivial", loss_scale=1.0, z_min=1e-3, init_mu=1e4, ftol=0.0,
              gtol=0.0, ptol=0.0, solver="chol", pcg_max_iters=200, pcg_tol=1e-10, threads=1, lin_once=False,
              fast=False):
    """lin_once: store r / w / J per observation at linearisation and re-use them in the Schur and back-substitution
    passes (bit-identical results, fewer flops).  fast: run the -O3 -march=native -ffp-contract=fast build (timing only;
    contraction changes the rounding, so the parity tests keep the strict build)."""
    del fast
    return Opts(max_iters, {"trivial": 0, "huber": 1}[loss], loss_scale, z_min, init_mu, ftol, gtol,
                ptol, {"chol": 0, "pcg": 1}[solver], pcg_max_iters, pcg_tol, threads, 1 if lin_once else 0)


def build(force=False):
    so = os.path.join(_HERE, "libba_cpu.so")
    src = os.path.join(_HERE, "ba_cpu.cpp")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return so


FAST_FLAGS = ["-O3", "-march=native", "-ffp-contract=fast", "-fopenmp", "-fPIC", "-std=c++17"]
_FAST = None


def build_fast(force=False):
    """Timing-only build of the same source: -O3 -march=native, FMA contraction on.  Compiled ON THE MACHINE THAT
    RUNS IT (the GPU box's host CPU differs from the build container's, so a -march=native binary must not travel):
    the output is keyed by the CPU model and lives under oracle/_fast/ (git-ignored)."""
    import hashlib
    tag = hashlib.sha1(cpu_model().encode()).hexdigest()[:10]
    d = os.path.join(_HERE, "_fast")
    os.makedirs(d, exist_ok=True)
    so, src = os.path.join(d, f"libba_cpu_fast_{tag}.so"), os.path.join(_HERE, "ba_cpu.cpp")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["/usr/bin/g++", *FAST_FLAGS, "-shared", "-o", so, src])
    return so


def cpu_model():
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("model name"):
                return line.split(":", 1)[1].strip()
    except OSError:
        pass
    return "unknown"


def lib(fast=False):
    global _LIB, _FAST
    if fast:
        if _FAST is None:
            _FAST = C.CDLL(build_fast())
            _FAST.ba_cpu_index.restype = C.c_int64
        return _FAST
    if _LIB is None:
        so = os.path.join(_HERE, "libba_cpu.so")
        if not os.path.exists(so):
            build()
        _LIB = C.CDLL(so)
        _LIB.ba_cpu_index.restype = C.c_int64
    return _LIB


def _p(a, t):
    return None if a is None else a.ctypes.data_as(C.POINTER(t))


def _prep(p):
    return (np.ascontiguousarray(p.obs_cam, np.int32), np.ascontiguousarray(p.obs_pt, np.int32),
            np.ascontiguousarray(p.obs_uv, np.float32), np.ascontiguousarray(p.fixed_cam, np.uint8),
            np.ascontiguousarray(p.intr, np.float64))


def solve(p, **kw):
    """Full LM on one problem. Returns (poses, points, Summary)."""
    oc, op, uv, fx, intr = _prep(p)
    poses = np.array(p.poses, np.float64, order="C", copy=True)
    pts = np.array(p.points, np.float64, order="C", copy=True)
    o, sm = make_opts(**kw), Summary()
    rc = lib(kw.get("fast", False)).ba_cpu_solve(C.c_int(p.nc), C.c_int(p.np_), C.c_int64(p.no), _p(poses, C.c_double),
                            _p(pts, C.c_double), _p(oc, C.c_int32), _p(op, C.c_int32), _p(uv, C.c_float),
                            _p(fx, C.c_uint8), _p(intr, C.c_double), C.byref(o), C.byref(sm))
    if rc < 0:
        raise ValueError(f"ba_cpu_solve: invalid problem ({rc})")
    return poses, pts, sm


def solve_batch(b, **kw):
    """C4: all windows of a BABatch. Returns (poses, points, [Summary])."""
    poses = np.array(b.poses, np.float64, order="C", copy=True)
    pts = np.array(b.points, np.float64, order="C", copy=True)
    oc, op, uv, fx, intr = _prep(b)
    o = make_opts(**kw)
    sms = (Summary * b.nb)()
    co = np.ascontiguousarray(b.cam_off, np.int32)
    po = np.ascontiguousarray(b.pt_off, np.int32)
    oo = np.ascontiguousarray(b.obs_off, np.int64)
    rc = lib(kw.get("fast", False)).ba_cpu_solve_batch(C.c_int(b.nb), _p(co, C.c_int32), _p(po, C.c_int32), _p(oo, C.c_int64),
                                  _p(poses, C.c_double), _p(pts, C.c_double), _p(oc, C.c_int32),
                                  _p(op, C.c_int32), _p(uv, C.c_float), _p(fx, C.c_uint8),
                                  _p(intr, C.c_double), C.byref(o), sms)
    if rc < 0:
        raise ValueError("ba_cpu_solve_batch: invalid problem")
    return poses, pts, list(sms)


@dataclass
class IndexArrays:
    pm_perm: np.ndarray
    pt_ptr: np.ndarray
    cm_perm: np.ndarray
    cam_ptr: np.ndarray
    fidx: np.ndarray
    blk_ptr: np.ndarray
    blk_col: np.ndarray


def index(p) -> IndexArrays:
    oc, op, _, fx, _ = _prep(p)
    nfree = int((fx == 0).sum())
    a = IndexArrays(np.empty(p.no, np.int32), np.empty(p.np_ + 1, np.int32), np.empty(p.no, np.int32),
                    np.empty(p.nc + 1, np.int32), np.empty(p.nc, np.int32), np.empty(nfree + 1, np.int32),
                    np.empty(min(nfree * (nfree + 1) // 2, nfree + p.no * 64), np.int32))
    nb = lib().ba_cpu_index(C.c_int(p.nc), C.c_int(p.np_), C.c_int64(p.no), _p(oc, C.c_int32), _p(op, C.c_int32),
                            _p(fx, C.c_uint8), _p(a.pm_perm, C.c_int32), _p(a.pt_ptr, C.c_int32),
                            _p(a.cm_perm, C.c_int32), _p(a.cam_ptr, C.c_int32), _p(a.fidx, C.c_int32),
                            _p(a.blk_ptr, C.c_int32), _p(a.blk_col, C.c_int32))
    if nb < 0:
        raise ValueError(f"ba_cpu_index: invalid problem ({nb})")
    a.blk_col = a.blk_col[:nb].copy()
    return a


def stage(p, mu=1e4, **kw):
    """All intermediates of one LM iteration at radius mu (dict of arrays)."""
    oc, op, uv, fx, intr = _prep(p)
    poses = np.ascontiguousarray(p.poses, np.float64)
    pts = np.ascontiguousarray(p.points, np.float64)
    n = 6 * int((fx == 0).sum())
    o = make_opts(**kw)
    out = dict(r=np.zeros((p.no, 2)), w=np.zeros(p.no), Jc=np.zeros((p.no, 2, 6)), Jp=np.zeros((p.no, 2, 3)),
               B=np.zeros((p.nc, 6, 6)), gc=np.zeros((p.nc, 6)), C=np.zeros((p.np_, 6)), gp=np.zeros((p.np_, 3)),
               S=np.zeros((n, n)), gred=np.zeros(n), dc=np.zeros(n), dp=np.zeros((p.np_, 3)),
               scalars=np.zeros(4), tposes=np.zeros((p.nc, 12)), tpoints=np.zeros((p.np_, 3)))
    order = ["r", "w", "Jc", "Jp", "B", "gc", "C", "gp", "S", "gred", "dc", "dp", "scalars", "tposes", "tpoints"]
    rc = lib().ba_cpu_stage(C.c_int(p.nc), C.c_int(p.np_), C.c_int64(p.no), _p(poses, C.c_double),
                            _p(pts, C.c_double), _p(oc, C.c_int32), _p(op, C.c_int32), _p(uv, C.c_float),
                            _p(fx, C.c_uint8), _p(intr, C.c_double), C.byref(o), C.c_double(mu),
                            *[_p(out[k], C.c_double) for k in order])
    if rc < 0:
        raise ValueError(f"ba_cpu_stage: invalid problem ({rc})")
    out.update(cost=out["scalars"][0], model_decrease=out["scalars"][1], trial_cost=out["scalars"][2],
               ok=bool(out["scalars"][3]))
    return out


def num_procs():
    return int(lib().ba_cpu_num_procs())
