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


def _plan(T, row_ptr, row_col):
    lib = capi.lib()
    nT = len(row_ptr) - 1
    rp = np.ascontiguousarray(row_ptr, np.uint32); rc = np.ascontiguousarray(row_col, np.uint32)
    n_tiles = nT + len(rc)
    cap_t = 3 * n_tiles + 8; cap_m = n_tiles * max(1, nT) + 8
    tasks = np.zeros((cap_t, 8), np.int32); terms = np.zeros((cap_m, 3), np.int32)
    nt = C.c_uint32(cap_t); nm = C.c_uint32(cap_m)
    p = lambda a: a.ctypes.data_as(C.c_void_p)
    ret = lib.bagpu_chol_task_plan(C.c_int(T), C.c_uint32(nT), p(rp), p(rc), p(tasks), C.byref(nt), p(terms), C.byref(nm))
    assert ret == 0
    return tasks[:nt.value].astype(int), terms[:nm.value].astype(int)


def _replay(tasks, terms, row_ptr, row_col, near, workers):
    """The kernel's scheduling rule on the host: `workers` CTAs claim tasks in list order from one counter; a CTA holds its
    task until the flags it waits on are set by other tasks.  Returns the number of tasks completed (all of them, unless the
    order can deadlock with this many CTAs)."""
    nT = len(row_ptr) - 1
    tile_done = set(); diag_done = set(); x_done = set(); row_count = np.zeros(nT, int)

    def run(t):
        kind, i, j, slot, lb, le = t[:6]
        if kind in (0, 1):
            for sa, sb, _k in terms[lb:le]:
                while not (sa in tile_done and sb in tile_done):
                    yield
            if kind == 0:
                diag_done.add(i)
            else:
                while i not in diag_done:
                    yield
                tile_done.add(slot)
        elif kind == 2:
            while slot not in tile_done:
                yield
            while j not in x_done:
                yield
            row_count[i] += 1
        else:
            hi, lo = i, j
            while hi not in diag_done:
                yield
            for r in range(hi, lo - 1, -1):
                cols = row_col[row_ptr[r]:row_ptr[r + 1]]
                n_in = sum(1 for q, c in enumerate(cols) if q < near and c <= hi and c <= r + near)
                for c in cols[:n_in]:
                    assert c in x_done                                   # x of the near columns is the chain's own, already solved
                while row_count[r] < len(cols) - n_in:
                    yield
                x_done.add(r)

    nxt = 0; held = [None] * workers; finished = 0
    while True:
        progress = False
        for w in range(workers):
            if held[w] is None and nxt < len(tasks):
                held[w] = run(tasks[nxt]); nxt += 1; progress = True
            if held[w] is not None:
                try:
                    next(held[w])
                except StopIteration:
                    held[w] = None; finished += 1; progress = True
        if not progress:
            return finished, len(x_done)


@pytest.mark.parametrize("case", range(6))
@pytest.mark.parametrize("T", [48, 120])
def test_task_plan_covers_every_tile_once_and_cannot_deadlock(case, T):
    nT, edges = _graphs()[case]
    _perm, row_ptr, row_col = _fill(nT, edges)
    tasks, terms = _plan(T, row_ptr, row_col)
    near = 3 if T <= 72 else 1
    slot = {}
    for k in range(nT):
        slot[(k, k)] = k + row_ptr[k]
        for q, c in enumerate(row_col[row_ptr[k]:row_ptr[k + 1]]):
            slot[(k, int(c))] = k + row_ptr[k] + 1 + q
    have = set(slot)
    # factorisation: one task per stored tile, with its slot, and exactly the update terms of the left-looking rule
    fact = [t for t in tasks if t[0] in (0, 1)]
    assert sorted((t[1], t[2]) for t in fact) == sorted(have)
    for kind, i, j, sl, lb, le, _a, _b in fact:
        assert (kind == 0) == (i == j) and sl == slot[(i, j)]
        want = sorted(k for k in range(i) if (k, i) in have and (k, j) in have)
        assert sorted(terms[lb:le, 2].tolist()) == want
        for sa, sb, k in terms[lb:le]:
            assert sa == slot[(k, i)] and sb == slot[(k, j)]
    # backward substitution: chains partition the rows, consecutive rows of a chain are coupled, every off-diagonal tile is
    # either a tile task or one of the chain's near columns
    chains = [t for t in tasks if t[0] == 4]
    rows = []
    for _k, hi, lo, *_ in chains:
        assert 0 <= lo <= hi < nT and hi - lo + 1 <= 16
        for r in range(lo, hi):
            assert (r, r + 1) in have
        rows += range(lo, hi + 1)
    assert sorted(rows) == list(range(nT))
    chain_hi = {}
    for _k, hi, lo, *_ in chains:
        for r in range(lo, hi + 1):
            chain_hi[r] = hi
    tile_tasks = sorted((t[1], t[2]) for t in tasks if t[0] == 2)
    want = sorted((i, j) for (i, j) in have if i != j and not (j <= chain_hi[i] and j <= i + near
                                                              and list(row_col[row_ptr[i]:row_ptr[i + 1]]).index(j) < near))
    assert tile_tasks == want
    assert len(tasks) == len(fact) + len(chains) + len(tile_tasks)
    # the order: with as few as two CTAs every task completes (a task only ever waits for tasks that were claimed before it,
    # or — the far tiles of a chain — for x the chain publishes while it descends)
    for workers in (2, 3, 148):
        done, solved = _replay(tasks, terms, row_ptr, row_col, near, workers)
        assert done == len(tasks) and solved == nT, (workers, done, len(tasks))


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


def _work_items(n_ap, L, rec_start, lmrec_start, keys, counts):
    lib = capi.lib()
    rec_start = np.ascontiguousarray(rec_start, np.uint32); lmrec_start = np.ascontiguousarray(lmrec_start, np.uint32)
    keys = np.ascontiguousarray(keys, np.uint64); counts = np.ascontiguousarray(counts, np.uint32)
    cap = int(len(keys) + counts.sum() // 512 + rec_start[-1] // 512 + lmrec_start[-1] // 512 + 4 * n_ap + 64)
    items = np.zeros((cap, 10), np.uint32); sym = np.zeros((cap, 10), np.uint32)
    ni = C.c_uint32(cap); ns = C.c_uint32(cap); npart = C.c_uint32(0); ngrp = C.c_uint32(0)
    p = lambda a: a.ctypes.data_as(C.c_void_p)
    rc = lib.bagpu_build_work_items(C.c_uint32(n_ap), C.c_int(L), p(rec_start), p(lmrec_start), C.c_uint32(len(keys)), p(keys), p(counts),
                                    p(items), C.byref(ni), p(sym), C.byref(ns), C.byref(npart), C.byref(ngrp))
    assert rc == 0
    return items[:ni.value].astype(np.int64), sym[:ns.value].astype(np.int64), npart.value, ngrp.value


@pytest.mark.parametrize("L", [1, 3])
def test_work_items_cover_every_block_exactly_once(L):
    # columns of an item: begin, end, row, col, kind, mfirst, mcount, mself, mgroup, pad
    rng = np.random.default_rng(11 + L)
    n_ap = 37
    rec = np.concatenate([[0], np.cumsum(rng.integers(0, 1500, n_ap))])           # some poses need several chunks
    lmrec = np.concatenate([[0], np.cumsum(rng.integers(0, 700, n_ap))]) if L == 1 else np.zeros(n_ap + 1, int)
    keys, counts = [], []
    for r in range(n_ap):
        for c in range(r, min(n_ap, r + 6)):
            if c == r and rng.random() < 0.7:
                continue                                                            # same-pose pairs are rare
            if rng.random() < 0.2:
                continue
            keys.append((r * n_ap + c) * 2); counts.append(int(rng.choice([1, 7, 130, 512, 513, 1600])))
            if L == 1 and c != r and rng.random() < 0.5:                           # measurement/reference run of the same block
                keys.append((r * n_ap + c) * 2 + 1); counts.append(int(rng.choice([3, 40, 900])))
    items, sym, n_partial, n_groups = _work_items(n_ap, L, rec, lmrec, keys, counts)
    allit = np.concatenate([items, sym])
    assert np.all(allit[:, 1] > allit[:, 0]) and np.all(allit[:, 1] - allit[:, 0] <= 512)
    # diagonal items tile the record ranges of every pose
    for kind, starts in ((0, rec), (1, lmrec)):
        d = items[items[:, 4] == kind]
        for r in range(n_ap):
            mine = d[d[:, 2] == r]
            assert np.all(mine[:, 3] == r)
            got = sorted((int(b), int(e)) for b, e in mine[:, :2])
            want = [(b, min(b + 512, int(starts[r + 1]))) for b in range(int(starts[r]), int(starts[r + 1]), 512)]
            assert got == want
    # pair-list items tile the runs in list order, with the run's block and kind
    off = np.concatenate([[0], np.cumsum(counts)])
    pl = allit[allit[:, 4] >= 2]
    pl = pl[np.argsort(pl[:, 0])]
    k = 0
    for bi, key in enumerate(keys):
        r, c = divmod(key >> 1, n_ap)
        kind = 4 if key & 1 else (3 if r == c else 2)
        pos = off[bi]
        while pos < off[bi + 1]:
            it = pl[k]; k += 1
            assert (it[0], it[1]) == (pos, min(pos + 512, off[bi + 1])) and (it[2], it[3], it[4]) == (r, c, kind)
            pos = it[1]
    assert k == len(pl)
    assert np.all(sym[:, 4] == 3) and not np.any(items[:, 4] == 3)
    # chunk groups: the items of one block (all its runs) share mcount consecutive partial slots and one counter
    blocks = {}
    for it in allit:
        blocks.setdefault((it[2], it[3], it[4] < 2, it[4] == 3), []).append(it)
    used = []
    groups = set()
    for its in blocks.values():
        mc = len(its)
        assert all(it[6] == mc for it in its)
        if mc > 1:
            assert len({it[5] for it in its}) == 1 and len({it[8] for it in its}) == 1
            assert sorted(it[7] for it in its) == list(range(its[0][5], its[0][5] + mc))
            used += [it[7] for it in its]; groups.add(its[0][8])
    assert sorted(used) == list(range(n_partial)) and sorted(groups) == list(range(n_groups))
