"""world_size-2 (gloo, CPU) test of the multi-GPU host logic: landmark sharding.

Each rank keeps the projection residuals of a contiguous landmark range, builds its partial reduced
system (with the CPU oracle standing in for the kernels), the partial S / rhs are all-reduced, and the
sum must equal the unsharded system — the property the NCCL path of ba_b200/csrc/capi.cu relies on
(Schur contributions are independent per landmark).
"""
import os
import sys
import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from ba_b200 import synth
from ba_b200.dist import landmark_range, lm_obs_start_of, shard_arrays
from ba_b200.problem import Problem


def _problem():
    return synth.make_c1(n_poses=8, n_landmarks=60, obs_per_lm=5, seed=31)


def _build(arrays, lm_dim, pose_dim):
    from oracle.oracle_py import Oracle
    p = Problem(lm_dim, pose_dim)
    p.set_bulk(**arrays)
    s = Oracle(lm_dim, pose_dim)
    o = s.default_options(); o.use_robust_norm_for_proj_residuals = 0
    s.set_options(o)
    s.load(p); s.build()
    return s.get_system()


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    p = _problem()
    a = p.arrays()
    shard, (lo, hi) = shard_arrays(a, rank, world)
    sysm = _build(shard, 1, 6)
    S = torch.from_numpy(np.triu(sysm["S"]).copy()); rhs = torch.from_numpy(sysm["rhs_sc"].copy()); bp = torch.from_numpy(sysm["rhs_p"].copy())
    # rhs_sc of a shard already contains its own b_p; both are sums over the shard's residuals
    for t in (S, rhs, bp):
        dist.all_reduce(t)
    n_local = torch.tensor([len(shard["proj_z"])])
    dist.all_reduce(n_local)
    if rank == 0:
        q.put((S.numpy(), rhs.numpy(), bp.numpy(), int(n_local.item()), (lo, hi)))
    dist.destroy_process_group()


def test_landmark_ranges_cover_and_balance():
    a = _problem().arrays()
    st = lm_obs_start_of(a)
    for world in (1, 2, 4, 8):
        rs = [landmark_range(st, r, world) for r in range(world)]
        assert rs[0][0] == 0 and rs[-1][1] == len(st) - 1
        for (l0, h0), (l1, h1) in zip(rs[:-1], rs[1:]):
            assert h0 == l1
        counts = [st[h] - st[l] for l, h in rs]
        assert sum(counts) == st[-1] and max(counts) - min(counts) <= 2 * np.diff(st).max()


def test_partial_systems_sum_to_the_full_system():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for pr in procs:
        pr.start()
    S, rhs, bp, n_total, _ = q.get(timeout=120)
    for pr in procs:
        pr.join(60)
        assert pr.exitcode == 0
    a = _problem().arrays()
    full = _build(a, 1, 6)
    assert n_total == len(a["proj_z"])
    assert np.allclose(S, np.triu(full["S"]), rtol=1e-12, atol=1e-9)
    assert np.allclose(rhs, full["rhs_sc"], rtol=1e-12, atol=1e-9)
    assert np.allclose(bp, full["rhs_p"], rtol=1e-12, atol=1e-9)
