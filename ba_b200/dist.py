"""Multi-GPU plumbing: landmark sharding and the NCCL communicator hand-off.

The path shards by landmark: every projection residual touches exactly one
landmark, so V, b_l, W and the Schur contribution W V^-1 W^T are independent per landmark
(reference src/BundleAdjuster.cpp:414-443,452-470).  Each rank uploads the whole problem, keeps
the observations of a contiguous landmark range balanced by observation count, builds a partial
S / b, and the partial systems are all-reduced (NCCL, inside the C++ library); the reduced camera
solve is replicated.  `landmark_range` is the same rule the C++ side applies
(ba_b200/csrc/capi.cu, setup()), kept here so the CPU tests can check it with gloo.
"""
import os
import numpy as np


def landmark_range(lm_obs_start, rank, world):
    """Contiguous landmark range [lo, hi) of `rank`; lm_obs_start = CSR offsets (n_lms + 1)."""
    n_lms = len(lm_obs_start) - 1
    if world <= 1:
        return 0, n_lms
    n_obs = float(lm_obs_start[-1])
    per = n_obs / world
    lo, hi = 0, n_lms
    for l in range(n_lms + 1):
        if float(lm_obs_start[l]) <= per * rank:
            lo = l
        if float(lm_obs_start[l]) <= per * (rank + 1):
            hi = l
    if rank == world - 1:
        hi = n_lms
    if rank == 0:
        lo = 0
    return lo, hi


def lm_obs_start_of(problem_arrays):
    lm = problem_arrays["proj_lm"]
    n_lms = len(problem_arrays["lm_x_w"])
    counts = np.bincount(lm, minlength=n_lms)
    return np.concatenate([[0], np.cumsum(counts)]).astype(np.int64)


def shard_arrays(a, rank, world):
    """Arrays of the shard `rank` sees: all poses / landmarks, only its landmarks' residuals."""
    lo, hi = landmark_range(lm_obs_start_of(a), rank, world)
    keep = (a["proj_lm"] >= lo) & (a["proj_lm"] < hi)
    out = dict(a)
    for k in ("proj_z", "proj_meas_pose", "proj_lm", "proj_cam", "proj_weight"):
        out[k] = np.ascontiguousarray(a[k][keep])
    return out, (lo, hi)


def init_nccl_for(solver):
    """One process per GPU: rank 0 creates the ncclUniqueId, torch.distributed broadcasts it."""
    import torch
    import torch.distributed as dist
    from .capi import BundleAdjusterGPU
    rank, world = dist.get_rank(), dist.get_world_size()
    if world == 1:
        return
    # prefer the NCCL that torch already loaded
    if "BAGPU_NCCL_LIB" not in os.environ:
        try:
            import nvidia.nccl
            cand = os.path.join(os.path.dirname(nvidia.nccl.__file__), "lib", "libnccl.so.2")
            if os.path.exists(cand):
                os.environ["BAGPU_NCCL_LIB"] = cand
        except Exception:
            pass
    uid = torch.zeros(128, dtype=torch.uint8)
    if rank == 0:
        uid = torch.tensor(list(BundleAdjusterGPU.nccl_unique_id()), dtype=torch.uint8)
    dev = torch.device("cuda", torch.cuda.current_device()) if dist.get_backend() == "nccl" else torch.device("cpu")
    uid = uid.to(dev)
    dist.broadcast(uid, src=0)
    solver.comm_init(rank, world, bytes(uid.cpu().numpy().tobytes()))
