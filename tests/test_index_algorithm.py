"""The ALGORITHM of the one-CTA-per-window index kernel (csrc/ba_index.cu `window_index_kernel`), restated step by step
in numpy and held against the oracle's std::stable_sort index on the CPU: a camera bit mask per landmark, positions by
population counts, the landmark schedule as a stable counting sort over 32 buckets, per-camera bit vectors over the slots
with a prefix count per 32 slots.  (The kernel itself is held against the oracle and against the radix-sort builder on the
GPU: tests/test_gpu_parity.py::test_index_bit_exact, tests/test_gpu_variants.py::test_window_index_kernel_equals_...)"""
import numpy as np
import pytest

from visualodometry_b200 import synth


def popc(x):
    return np.array([bin(int(v)).count("1") for v in np.atleast_1d(x)], np.int64)


def window_index(nc, npts, cam, pt):
    """-> sched, pt_ptr (schedule order), pm (original index per position), cm, cam_ptr; mirrors the kernel's six steps"""
    no = cam.size
    mask = np.zeros(npts, np.uint32)
    for c, p in zip(cam.tolist(), pt.tolist()):                       # 1. atomicOr; a bit set twice = duplicate
        assert not (mask[p] >> c) & 1, "duplicate observation"
        mask[p] |= np.uint32(1 << c)
    cnt = popc(mask)
    bucket = 31 - np.minimum(cnt, 31)                                 # 2. stable counting sort over 32 buckets
    start = np.concatenate([[0], np.cumsum(np.bincount(bucket, minlength=32))])[:32]
    sched = np.empty(npts, np.int64)
    pos_of = np.empty(npts, np.int64)
    run = start.copy()
    for p in range(npts):                                             # landmark id order inside a bucket
        sched[run[bucket[p]]] = p
        pos_of[p] = run[bucket[p]]
        run[bucket[p]] += 1
    ptr = np.concatenate([[0], np.cumsum(cnt[sched])])                # 3. exclusive scan in schedule order
    nchunk = (npts + 31) // 32                                        # 4. per-camera bit vectors + prefix per 32 slots
    bv = np.zeros((nc, nchunk), np.uint32)
    for q in range(npts):
        for c in range(nc):
            if (mask[sched[q]] >> c) & 1:
                bv[c, q >> 5] |= np.uint32(1 << (q & 31))
    per_chunk = np.array([[bin(int(v)).count("1") for v in row] for row in bv]).reshape(nc, nchunk)
    cp = np.concatenate([np.zeros((nc, 1), np.int64), np.cumsum(per_chunk, axis=1)], axis=1)[:, :nchunk]
    cam_ptr = np.concatenate([[0], np.cumsum(per_chunk.sum(axis=1))])
    pm = np.full(no, -1, np.int64)                                    # 6. both streams
    cm = np.full(no, -1, np.int64)
    cm_slot = np.empty(no, np.int64)
    for i, (c, p) in enumerate(zip(cam.tolist(), pt.tolist())):
        s = pos_of[p]
        a_pm = ptr[s] + bin(int(mask[p]) & ((1 << c) - 1)).count("1")
        a_cm = cam_ptr[c] + cp[c, s >> 5] + bin(int(bv[c, s >> 5]) & ((1 << (s & 31)) - 1)).count("1")
        assert pm[a_pm] == -1 and cm[a_cm] == -1
        pm[a_pm], cm[a_cm], cm_slot[a_cm] = i, i, s
    return sched, ptr, pm, cm, cm_slot, cam_ptr


@pytest.mark.parametrize("shape", [dict(seed=1, n_cams=4, n_pts=40, n_obs=110), dict(seed=2, n_cams=10, n_pts=300, n_obs=2100),
                                   dict(seed=3, n_cams=20, n_pts=150, n_obs=1900), dict(seed=4, n_cams=7, n_pts=97, n_obs=400)])
def test_mask_and_popcount_index_equals_the_oracles_stable_sort(oracle, shape):
    p = synth.make_window(**shape)
    rng = np.random.default_rng(shape["seed"])
    perm = rng.permutation(p.no)                                      # any input order gives the same index
    p.obs_cam, p.obs_pt, p.obs_uv = p.obs_cam[perm], p.obs_pt[perm], p.obs_uv[perm]
    keep = p.obs_pt != 3                                              # an unobserved landmark
    p.obs_cam, p.obs_pt, p.obs_uv = p.obs_cam[keep], p.obs_pt[keep], p.obs_uv[keep]
    sched, ptr, pm, cm, cm_slot, cam_ptr = window_index(p.nc, p.np_, p.obs_cam, p.obs_pt)
    cnt = np.bincount(p.obs_pt, minlength=p.np_)
    assert np.array_equal(sched, np.argsort(-np.minimum(cnt, 31), kind="stable"))     # the schedule the tests document
    ix = oracle.index(p)
    # canonical point-major order = the landmarks' runs taken in id order
    pos_of = np.argsort(sched)
    pm_canon = np.concatenate([pm[ptr[pos_of[q]]:ptr[pos_of[q] + 1]] for q in range(p.np_)]) if p.no else pm
    assert np.array_equal(pm_canon, ix.pm_perm)
    assert np.array_equal(np.concatenate([[0], np.cumsum(cnt)]), ix.pt_ptr)
    # canonical camera-major order = every camera's run re-ordered from slot order to landmark-id order
    assert np.array_equal(cam_ptr, ix.cam_ptr)
    cm_canon = np.concatenate([cm[a:b][np.argsort(sched[cm_slot[a:b]], kind="stable")] for a, b in zip(cam_ptr[:-1], cam_ptr[1:])])
    assert np.array_equal(cm_canon, ix.cm_perm)
    # inside a camera's run the slots ascend (the order the window kernel walks them in)
    for a, b in zip(cam_ptr[:-1], cam_ptr[1:]):
        assert (np.diff(cm_slot[a:b]) > 0).all()
