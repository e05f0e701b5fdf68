"""Multi-GPU partitioning of the BA hot path (SURVEY §8 row e).

* C4 (independent windows): contiguous blocks of windows per rank, no collective.
* C5 (one large problem): landmarks are cut into `nranks` contiguous ranges balanced by
  OBSERVATION count, so every landmark's segment, C_p, C'_p^-1, delta_p and trial update stay
  local; cameras are replicated and the reduced camera system is all-reduced.
Pure numpy: also exercised on CPU with gloo (tests/test_shard_gloo.py)."""
from __future__ import annotations

import numpy as np


def window_ranges(n_windows: int, nranks: int):
    """[(w_lo, w_hi)] per rank, sizes differ by at most one."""
    base, rem = divmod(int(n_windows), int(nranks))
    out, lo = [], 0
    for r in range(nranks):
        hi = lo + base + (1 if r < rem else 0)
        out.append((lo, hi))
        lo = hi
    return out


def point_ranges(obs_pt, n_points: int, nranks: int):
    """[(p_lo, p_hi)] per rank: contiguous landmark ranges that tile [0, n_points) with
    (nearly) equal observation counts.  Deterministic; identical on every rank."""
    cnt = np.bincount(np.asarray(obs_pt, np.int64), minlength=int(n_points))
    pre = np.concatenate([[0], np.cumsum(cnt)])           # pre[p] = observations of landmarks < p
    total = int(pre[-1])
    cuts = [0]
    for r in range(1, nranks):
        target = (total * r) // nranks
        p = int(np.searchsorted(pre, target, side="left"))
        cuts.append(min(max(p, cuts[-1]), n_points))
    cuts.append(int(n_points))
    return [(cuts[r], cuts[r + 1]) for r in range(nranks)]


def shard_problem(p, p_lo: int, p_hi: int):
    """The sub-problem a rank owns: all cameras, landmarks [p_lo, p_hi) renumbered from 0 and
    their observations (original relative order kept).  Used by the CPU/gloo tests and the
    CPU baseline; the CUDA engine takes the full problem plus the range instead."""
    from .synth import BAProblem
    m = (p.obs_pt >= p_lo) & (p.obs_pt < p_hi)
    return BAProblem(p.poses, p.points[p_lo:p_hi], p.obs_cam[m], (p.obs_pt[m] - p_lo).astype(np.int32),
                     p.obs_uv[m], p.fixed_cam, p.intr)
