"""CPU-side checks of the boundary: the C-ABI library loads and exports every symbol include/b2ba.h
declares, refuses to run without a GPU (no CPU fallback), and the host-side mirror packs arrays the
way the reference holds them (T_wc 4x4 f64 vo.py:19, K 3x3 vo.py:11)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    from visualodometry_b200 import _build, _lib
    _build.build()
    return _lib.load()


def test_exports_every_declared_symbol(lib):
    from visualodometry_b200 import _lib
    hdr = open(os.path.join(ROOT, "include", "b2ba.h")).read()
    declared = sorted(set(re.findall(r"\b(b2ba_[a-z_]+)\s*\(", hdr)))
    assert declared == sorted(_lib.SYMBOLS)
    for s in declared:
        assert getattr(lib, s) is not None
    assert lib.b2ba_version() >= 100


def test_struct_layouts_match_header(lib):
    from visualodometry_b200 import _lib
    assert C.sizeof(_lib.Options) == 80 and C.sizeof(_lib.Summary) == 48
    o = _lib.Options()
    lib.b2ba_default_options(C.byref(o))
    assert (o.max_iters, o.loss, o.loss_scale, o.z_min, o.init_mu) == (10, 0, 1.0, 1e-3, 1e4)
    assert (o.solver, o.pcg_max_iters, o.pcg_tol) == (0, 200, 1e-10)


def test_no_cpu_fallback(lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from visualodometry_b200 import BundleAdjuster, _lib
    h = C.c_void_p()
    assert lib.b2ba_create(C.byref(h), 0, None) == _lib.E_CUDA and not h.value
    with pytest.raises(RuntimeError, match="CUDA device is required"):
        BundleAdjuster(np.eye(3))


def test_product_never_imports_oracle():
    """The oracle is test infrastructure: nothing under visualodometry_b200/ may import, link or run it."""
    pkg = os.path.join(ROOT, "visualodometry_b200")
    bad = re.compile(r"(from\s+oracle|import\s+oracle|libba_cpu|ba_cpu\.|#include[^\n]*oracle)")
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                assert not bad.search(open(os.path.join(dp, f)).read()), f


def test_pose_layout_roundtrip():
    from visualodometry_b200 import T_from_poses, poses_from_T, synth
    rng = np.random.default_rng(0)
    T = np.tile(np.eye(4), (5, 1, 1))
    for i in range(5):
        T[i, :3, :3] = synth.so3_exp(rng.normal(0, 0.5, 3))
        T[i, :3, 3] = rng.normal(0, 3, 3)
    P = poses_from_T(T)
    assert P.shape == (5, 12) and P.flags.c_contiguous
    assert np.array_equal(P[:, :9].reshape(5, 3, 3), T[:, :3, :3]) and np.array_equal(P[:, 9:], T[:, :3, 3])
    assert np.array_equal(T_from_poses(P), T)


def test_config_maps_to_options():
    from visualodometry_b200 import BAConfig
    o = BAConfig(max_iters=7, loss="huber", loss_scale=2.5, solver="pcg", record_history=True).to_options()
    assert (o.max_iters, o.loss, o.loss_scale, o.solver, o.flags) == (7, 1, 2.5, 2, 1)
    assert BAConfig().z_min == 1e-3      # = VOConfig.min_depth, config.py:31


def test_synthetic_shapes():
    from visualodometry_b200 import synth
    p = synth.make_window(0)
    assert (p.nc, p.np_, p.no) == (10, 2000, 15000) and np.bincount(p.obs_pt, minlength=2000).min() >= 2
    assert p.obs_uv.dtype == np.float32 and p.obs_cam.dtype == np.int32 and p.fixed_cam[0] == 1
    q = synth.make_window(0)
    assert np.array_equal(p.obs_uv, q.obs_uv) and np.array_equal(p.points, q.points)      # deterministic
    b = synth.make_batch(3, n_pts=100, n_obs=500)
    assert b.nb == 3 and b.obs_off[-1] == b.obs_cam.shape[0] and b.window(2).np_ == 100
    c5 = synth.make_large(0, n_cams=50, n_pts=1000, k=5)
    assert c5.no == 5000 and (np.bincount(c5.obs_pt) == 5).all()
