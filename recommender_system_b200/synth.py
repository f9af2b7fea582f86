"""Synthetic inputs of the shapes whose upstream files are missing.

The reference checkout lists instML1M.in and inst60000-2000-200-10-20.in in
.MISSING_LARGE_BLOBS, so the two headline shapes are regenerated here with the
statistics SURVEY.md section 8(d) records (dimensions, nonzero count, row/column
length distributions, ratings 1..5).  The generator uses its own PRNG, never
libc random(), so the reference's initialisation sequence is untouched.
All outputs are (row, col)-sorted and duplicate free like every upstream sample.
"""
from __future__ import annotations

import numpy as np

from .io import ENTRY_DTYPE, Instance


def _entries(rows, cols, vals) -> np.ndarray:
    e = np.zeros(len(rows), dtype=ENTRY_DTYPE)
    e["row"], e["col"], e["value"] = rows, cols, vals
    return e


def _pick_items(rng, counts, items, log_w=None, chunk=4096):
    """For every user pick counts[u] distinct items (ascending); weighted by
    exp(log_w) through Gumbel top-k when log_w is given, uniform otherwise."""
    rows, cols = [], []
    kmax = int(counts.max()) if len(counts) else 0
    for lo in range(0, len(counts), chunk):
        c = counts[lo:lo + chunk]
        keys = rng.gumbel(size=(len(c), items)).astype(np.float32)
        if log_w is not None:
            keys += log_w[None, :].astype(np.float32)
        kk = min(kmax, items)
        top = np.argpartition(-keys, kk - 1, axis=1)[:, :kk] if kk < items else \
            np.tile(np.arange(items), (len(c), 1))
        # order the kk candidates of each user by key so that a prefix is a weighted sample
        order = np.argsort(-np.take_along_axis(keys, top, axis=1), axis=1)
        top = np.take_along_axis(top, order, axis=1)
        for u in range(len(c)):
            sel = np.sort(top[u, :c[u]])
            rows.append(np.full(len(sel), lo + u, dtype=np.int32))
            cols.append(sel.astype(np.int32))
    return np.concatenate(rows), np.concatenate(cols)


def inst60000_like(users=60000, items=2000, features=200, iters=100, alpha=0.001,
                   lo=10, hi=20, seed=12346) -> Instance:
    """Shape of inst60000-2000-200-10-20: every user rates lo..hi uniform items."""
    rng = np.random.default_rng(seed)
    counts = rng.integers(lo, hi + 1, size=users)
    counts = np.minimum(counts, items)
    rows, cols = _pick_items(rng, counts, items)
    vals = rng.integers(1, 6, size=len(rows)).astype(np.float64)
    return Instance(iters, alpha, features, users, items, _entries(rows, cols, vals))


def ml_like(users=6040, items=3952, nnz=1_000_209, features=100, iters=100, alpha=0.0001,
            min_row=20, max_row=2314, seed=12345) -> Instance:
    """MovieLens-1M shaped: log-normal user activity, Zipf-like item popularity."""
    rng = np.random.default_rng(seed)
    max_row = min(max_row, items)
    min_row = min(min_row, max_row)
    raw = rng.lognormal(mean=0.0, sigma=1.0, size=users)
    # scale so that the clipped counts add up to nnz
    a, b = 1e-3, 1e6
    for _ in range(80):
        mid = (a * b) ** 0.5
        tot = np.clip(np.rint(raw * mid), min_row, max_row).sum()
        if tot < nnz:
            a = mid
        else:
            b = mid
    counts = np.clip(np.rint(raw * b), min_row, max_row).astype(np.int64)
    diff = int(nnz - counts.sum())
    idx = rng.permutation(users)
    step = 1 if diff > 0 else -1
    p = 0
    while diff != 0 and p < 50 * users:
        u = idx[p % users]
        if min_row <= counts[u] + step <= max_row:
            counts[u] += step
            diff -= step
        p += 1
    log_w = -0.5 * np.log(np.arange(1, items + 1, dtype=np.float64))
    log_w = log_w[rng.permutation(items)]
    rows, cols = _pick_items(rng, counts, items, log_w=log_w, chunk=1024)
    vals = rng.integers(1, 6, size=len(rows)).astype(np.float64)
    return Instance(iters, alpha, features, users, items, _entries(rows, cols, vals))


def random_instance(users, items, features, density, iters, alpha, seed=0, values="int") -> Instance:
    """Small random problems for the parity tests (rows may be empty; a row may be full)."""
    rng = np.random.default_rng(seed)
    mask = rng.random((users, items)) < density
    rows, cols = np.nonzero(mask)
    if values == "int":
        vals = rng.integers(1, 6, size=len(rows)).astype(np.float64)
    else:
        vals = rng.uniform(-3.0, 7.0, size=len(rows))
    return Instance(iters, alpha, features, users, items,
                    _entries(rows.astype(np.int32), cols.astype(np.int32), vals))
