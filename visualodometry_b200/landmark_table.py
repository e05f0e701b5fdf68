"""Dense id -> slot landmark table behind the reference's `map_points` dict.

The reference keeps its map as a Python dict `id -> xyz` (`src/modules/vo.py:17`) and scans it one id at a time
in the frame loop: `valid = [(p != -1 and p in self.map_points) for p in kf_ids]`, `X = [self.map_points[p] for p in
kf_ids[valid]]` (`vo.py:121-130`), and `_prune_map` (`vo.py:35-47`).  Landmark ids are handed out consecutively
(`vo.py:281-283`) and old ones are pruned below a moving floor, so the live ids fill a window of the integers: a dense
array indexed by `id - base` answers the same questions with two gathers.

`LandmarkTable` IS a dict (the reference's code keeps working on it unchanged: `in`, `[]`, `del`, iteration, `update`)
that mirrors every write into `(capacity, 3)` float64 + a validity mask, and adds the vectorised forms:

    valid, X = table.lookup(ids)        # == the two list comprehensions of vo.py:121-130, bit for bit
    table.update_rows(ids, rows)        # == dict.update(zip(ids, rows)) for ids that are present
    table.prune_below(floor)            # == the loop of vo.py:41-47

Host-side only (numpy); nothing here touches the GPU -- it removes Python-level dict scans from the caller of the path.
"""
from __future__ import annotations

import numpy as np


class LandmarkTable(dict):
    def __init__(self, *args, capacity: int = 1 << 15, **kw):
        super().__init__()
        self._base = 0
        self._xyz = np.zeros((capacity, 3), np.float64)
        self._ok = np.zeros(capacity, bool)
        if args or kw:
            self.update(*args, **kw)

    # ------------------------------------------------------------------ mirror maintenance
    def _slot(self, k: int) -> int:
        i = k - self._base
        if i < 0:
            self._rebase(k)
            i = k - self._base
        if i >= self._ok.size:
            self._grow(k)
            i = k - self._base
        return i

    def _rebase(self, k_low: int) -> None:
        """make room below: shift the window so that k_low is addressable"""
        shift = self._base - k_low
        n = self._ok.size
        live = np.flatnonzero(self._ok)
        need = (live.max() + 1 if live.size else 0) + shift
        cap = n
        while cap < need:
            cap *= 2
        xyz, ok = np.zeros((cap, 3), np.float64), np.zeros(cap, bool)
        xyz[live + shift], ok[live + shift] = self._xyz[live], True
        self._xyz, self._ok, self._base = xyz, ok, k_low

    def _grow(self, k_high: int) -> None:
        """make room above: first drop the dead prefix (ids pruned away), then double if that is not enough"""
        live = np.flatnonzero(self._ok)
        lo = int(live.min()) if live.size else self._ok.size
        n = self._ok.size
        need = k_high - (self._base + lo) + 1
        cap = n
        while cap < need:
            cap *= 2
        xyz, ok = np.zeros((cap, 3), np.float64), np.zeros(cap, bool)
        if live.size:
            xyz[live - lo], ok[live - lo] = self._xyz[live], True
        self._xyz, self._ok, self._base = xyz, ok, self._base + lo

    # ------------------------------------------------------------------ dict protocol (every write is mirrored)
    def __setitem__(self, k, v):
        super().__setitem__(k, v)
        i = self._slot(int(k))
        self._xyz[i] = v
        self._ok[i] = True

    def __delitem__(self, k):
        super().__delitem__(k)
        self._ok[int(k) - self._base] = False

    def pop(self, k, *default):
        if k in self:
            v = super().pop(k)
            self._ok[int(k) - self._base] = False
            return v
        if default:
            return default[0]
        raise KeyError(k)

    def popitem(self):
        k, v = super().popitem()
        self._ok[int(k) - self._base] = False
        return k, v

    def clear(self):
        super().clear()
        self._ok[:] = False

    def setdefault(self, k, default=None):
        if k not in self:
            self[k] = default
        return self[k]

    def update(self, *args, **kw):
        for src in args:
            for k, v in (src.items() if hasattr(src, "items") else src):
                self[k] = v
        for k, v in kw.items():
            self[k] = v

    # ------------------------------------------------------------------ vectorised forms
    def lookup(self, ids):
        """`valid[i] = ids[i] != -1 and ids[i] in table`, `X = stack(table[p] for p in ids[valid])` (vo.py:121-130)."""
        ids = np.asarray(ids, dtype=np.int64)
        i = ids - self._base
        inb = (ids != -1) & (i >= 0) & (i < self._ok.size)
        valid = inb.copy()
        valid[inb] = self._ok[i[inb]]
        return valid, self._xyz[i[valid]]

    def update_rows(self, ids, rows) -> int:
        """`dict.update((p, x) for p, x in zip(ids, rows) if p in table)`; rows (N,3).  Returns how many were present."""
        ids = np.asarray(ids, dtype=np.int64)
        valid, _ = self.lookup(ids)
        rows = np.array(np.asarray(rows, np.float64)[valid])           # one copy; the dict holds views of its rows
        sel = ids[valid]
        self._xyz[sel - self._base] = rows
        super().update(zip(sel.tolist(), rows))
        return int(valid.sum())

    def prune_below(self, floor: int) -> int:
        """delete every id < floor (the reference's `_prune_map` loop, vo.py:41-47); returns how many"""
        hi = min(max(int(floor) - self._base, 0), self._ok.size)
        dead = np.flatnonzero(self._ok[:hi])
        for k in (dead + self._base).tolist():
            super().__delitem__(k)
        self._ok[:hi] = False
        return int(dead.size)
