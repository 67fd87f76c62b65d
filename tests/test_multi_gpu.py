"""2-GPU parity (skipped with fewer than two GPUs): a landmark-sharded Solve() — NCCL all-reduce of S, small
reductions over NVLink peer memory (ba_b200/csrc/peer_reduce.cu) — against the single-GPU Solve() on the same
problem, and the peer-memory reductions against the NCCL-only path.

The sharded sums associate differently from the single-GPU ones, so agreement is to rounding, amplified by the
conditioning of the reduced system: tolerances as in test_parity_gpu.py (1e-9 on costs)."""
import os
import subprocess
import sys

import numpy as np
import pytest

from ba_b200 import synth

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _n_gpus():
    import torch
    return torch.cuda.device_count()


def _run_sharded(tmp_path, tag, name, scale, iters, p2p, port):
    out = str(tmp_path / ("%s.npz" % tag))
    env = dict(os.environ, BAGPU_P2P="1" if p2p else "0")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", str(port), os.path.join(ROOT, "tests", "_multi_gpu_worker.py"), out, name, str(scale), str(iters)]
    r = subprocess.run(cmd, env=env, cwd=ROOT, capture_output=True, text=True, timeout=240)
    assert r.returncode == 0, r.stderr[-2000:]
    return np.load(out)


def _single(name, scale, iters):
    from ba_b200.capi import BundleAdjusterGPU
    p = synth.make(name, scale)
    s = BundleAdjusterGPU(p.lm_dim, p.pose_dim)
    o = s.default_options(); o.error_change_threshold = 0; o.param_change_threshold = 0; o.record_deltas = 1
    if hasattr(p, "gravity"):
        o.gravity[0], o.gravity[1], o.gravity[2] = [float(x) for x in p.gravity]
    s.set_options(o); s.load(p)
    sm = s.solve(iters)
    tr = s.get_trace()
    res = dict(iterations=sm.iterations, result=sm.result, post=np.array([t["post_solve_norm"] for t in tr]),
               median=np.array([t["median_maha"] for t in tr]), poses=s.get_poses()["t_wp"],
               lms=s.get_landmarks()["x_w"] if p.lm_dim else np.zeros(0))
    s.close()
    return res


@pytest.mark.parametrize("name,scale,iters", [("c1", 1.0, 6), ("c4", 0.02, 5)])
def test_two_gpu_solve_matches_single_gpu(tmp_path, name, scale, iters):
    if _n_gpus() < 2:
        pytest.skip("needs two GPUs")
    one = _single(name, scale, iters)
    two = _run_sharded(tmp_path, "p2p", name, scale, iters, True, 29611)
    nccl = _run_sharded(tmp_path, "nccl", name, scale, iters, False, 29612)
    for got in (two, nccl):
        assert int(got["iterations"]) == one["iterations"] and int(got["result"]) == one["result"]
        np.testing.assert_allclose(got["post"], one["post"], rtol=1e-9)
        # the exact median is an order statistic of the same multiset: identical up to the rounding of its inputs
        np.testing.assert_allclose(got["median"], one["median"], rtol=1e-9)
        np.testing.assert_allclose(got["poses"], one["poses"], rtol=0, atol=1e-7)
        np.testing.assert_allclose(got["lms"], one["lms"], rtol=0, atol=1e-6)
    # both reduction paths sum the ranks' contributions; the peer path does it in rank order on every rank
    np.testing.assert_allclose(two["post"], nccl["post"], rtol=1e-10)
