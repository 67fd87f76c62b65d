"""Host-side symbolic step of the reduced solve (no GPU): elimination order and fill of the tile pattern.

Replaces the ordering + symbolic factorisation inside Eigen::SimplicialLDLT (reference src/BundleAdjuster.cpp:753-761).
The library works on sorted column lists and merges a row's columns into its elimination-tree parent; the check is
the textbook dense symbolic Cholesky (every pair of columns of an eliminated row becomes coupled) on the same
permuted pattern."""
import ctypes as C
import numpy as np
import pytest

from ba_b200 import capi


def _fill(nT, edges):
    lib = capi.lib()
    ea = np.ascontiguousarray([e[0] for e in edges], np.uint32); eb = np.ascontiguousarray([e[1] for e in edges], np.uint32)
    perm = np.zeros(nT, np.uint32); row_ptr = np.zeros(nT + 1, np.uint32)
    cap = nT * (nT - 1) // 2 + 1
    row_col = np.zeros(cap, np.uint32)
    p = lambda a: a.ctypes.data_as(C.c_void_p)
    rc = lib.bagpu_symbolic_tile_fill(C.c_uint32(nT), C.c_uint32(len(edges)), p(ea), p(eb), p(perm), p(row_ptr), p(row_col), C.c_uint32(cap))
    assert rc == 0
    return perm.astype(int), row_ptr.astype(int), row_col[:row_ptr[-1]].astype(int)


def _dense_fill(nT, edges, perm):
    A = np.zeros((nT, nT), bool)
    for a, b in edges:
        x, y = perm[a], perm[b]
        if x != y:
            A[min(x, y), max(x, y)] = True
    for k in range(nT):
        cols = np.flatnonzero(A[k, k + 1:]) + k + 1
        for i, c in enumerate(cols):
            A[c, cols[i + 1:]] = True
    return A


def _graphs():
    rng = np.random.default_rng(7)
    out = []
    # banded chain (sequential video): bandwidth 3 tiles
    n = 97
    out.append((n, [(i, j) for i in range(n) for j in range(i + 1, min(n, i + 4))]))
    # pose graph: chain + loop closures
    n = 160
    e = [(i, i + 1) for i in range(n - 1)] + [(int(i), int(max(0, i - rng.integers(5, 60)))) for i in range(20, n, 7)]
    out.append((n, e))
    # random sparse, given in both orientations and with duplicates and self edges
    n = 61
    e = [(int(a), int(b)) for a, b in rng.integers(0, n, (150, 2))]
    out.append((n, e + [(b, a) for a, b in e[:40]]))
    # tiny (below the dissection threshold), empty, and dense
    out.append((5, [(0, 4), (1, 2)]))
    out.append((12, []))
    n = 20
    out.append((n, [(i, j) for i in range(n) for j in range(i + 1, n)]))
    return out


@pytest.mark.parametrize("case", range(6))
def test_fill_matches_dense_symbolic_cholesky(case):
    nT, edges = _graphs()[case]
    perm, row_ptr, row_col = _fill(nT, edges)
    assert sorted(perm.tolist()) == list(range(nT))                     # a permutation
    want = _dense_fill(nT, edges, perm)
    got = np.zeros((nT, nT), bool)
    for k in range(nT):
        cols = row_col[row_ptr[k]:row_ptr[k + 1]]
        assert np.all(np.diff(cols) > 0) and np.all(cols > k)            # strictly ascending, strict upper triangle
        got[k, cols] = True
    assert np.array_equal(got, want)


def test_dissection_shortens_the_dependency_chain_of_a_banded_matrix():
    # natural order of a block-tridiagonal matrix is a chain of nT dependent rows; the dissected order must give an
    # elimination tree whose height is far below nT (that is what the dataflow solver runs concurrently)
    nT = 256
    edges = [(i, i + 1) for i in range(nT - 1)]
    perm, row_ptr, row_col = _fill(nT, edges)
    level = np.zeros(nT, int)
    for k in range(nT):                                                 # parent = first column of the row
        if row_ptr[k + 1] > row_ptr[k]:
            par = row_col[row_ptr[k]]
            level[par] = max(level[par], level[k] + 1)
    assert level.max() < nT // 4


def _pack(starts, lo, hi):
    lib = capi.lib()
    starts = np.ascontiguousarray(starts, np.uint32)
    n = len(starts) - 1
    first = np.zeros(n + 1, np.uint32); count = np.zeros(n + 1, np.uint32); longs = np.zeros(n + 1, np.uint32)
    nt = C.c_uint32(n + 1); nl = C.c_uint32(n + 1)
    p = lambda a: a.ctypes.data_as(C.c_void_p)
    rc = lib.bagpu_pack_warp_tasks(C.c_uint32(n), p(starts), C.c_uint32(lo), C.c_uint32(hi), p(first), p(count), C.byref(nt), p(longs), C.byref(nl))
    assert rc == 0
    return first[:nt.value].astype(int), count[:nt.value].astype(int), longs[:nl.value].astype(int)


@pytest.mark.parametrize("seed", [0, 1, 2])
def test_warp_tasks_partition_the_short_tracks(seed):
    # every observation of a landmark with <= 32 observations lies in exactly one task, tasks hold whole landmarks
    # and at most 32 observations, landmarks with longer tracks are listed apart, empty landmarks cost nothing
    rng = np.random.default_rng(seed)
    n = 5000
    track = rng.geometric(0.2, n)                       # BAL-like: mean 5, long tail
    track[rng.integers(0, n, 40)] = rng.integers(33, 200, 40)
    track[rng.integers(0, n, 200)] = 0
    track[rng.integers(0, n, 20)] = 32
    starts = np.concatenate([[0], np.cumsum(track)])
    for lo, hi in [(0, n), (1234, 4321), (0, 0), (n - 1, n)]:   # the whole stream and a shard of it
        first, count, longs = _pack(starts, lo, hi)
        base = starts[lo]
        assert np.array_equal(longs, [l for l in range(lo, hi) if track[l] > 32])
        covered = np.zeros(starts[hi] - base, int)
        for f, c in zip(first, count):
            assert 0 < c <= 32
            assert (f + base) in starts and (f + c + base) in starts         # whole landmarks
            covered[f:f + c] += 1
        want = np.zeros_like(covered)
        for l in range(lo, hi):
            if track[l] <= 32:
                want[starts[l] - base:starts[l + 1] - base] = 1
        assert np.array_equal(covered, want)
        assert np.all(np.diff(first) > 0)                                      # in stream order
        if len(count) > 50:
            assert count.mean() > 24                                           # greedy packing keeps the lanes busy
