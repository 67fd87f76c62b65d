"""Worker of tests/test_multi_gpu.py: one rank of a landmark-sharded Solve() (launched by torch.distributed.run).

Rank 0 writes the trace, the refined poses and landmarks to the file named on the command line.  Run once with
BAGPU_P2P=1 (small reductions over NVLink peer memory) and once with BAGPU_P2P=0 (NCCL only)."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import numpy as np
import torch
import torch.distributed as dist
from _mg_problems import make, options
from ba_b200.capi import BundleAdjusterGPU

out, name, scale, iters = sys.argv[1], sys.argv[2], float(sys.argv[3]), int(sys.argv[4])
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("gloo", rank=rank, world_size=world)
p = make(name, scale)
s = BundleAdjusterGPU(p.lm_dim, p.pose_dim)
o = s.default_options(); o.error_change_threshold = 0; o.param_change_threshold = 0; o.record_deltas = 1
for k_, v_ in options(name).items():
    setattr(o, k_, v_)
if hasattr(p, "gravity"):
    o.gravity[0], o.gravity[1], o.gravity[2] = [float(x) for x in p.gravity]
s.set_options(o)
ids = [BundleAdjusterGPU.nccl_unique_id() if rank == 0 else None]
dist.broadcast_object_list(ids, src=0)
s.comm_init(rank, world, ids[0])
if os.environ.get("MG_RELOAD") == "1":
    # a smaller problem first in the SAME handle: the peer-mapped system window must be re-allocated for the larger one
    s.load(make("c1", 0.2)); s.solve(2)
s.load(p)
sm = s.solve(iters)
tr = s.get_trace()
poses = s.get_poses(); lms = s.get_landmarks() if p.lm_dim else None
# every rank must have taken the same accept / reject branches: identical traces, bit for bit
mine = json.dumps([[t[k] for k in ("pre_solve_norm", "post_solve_norm", "median_maha", "delta_norm", "inner_iterations", "trust_region_after")] for t in tr])
all_tr = [None] * world
dist.all_gather_object(all_tr, mine)
if rank == 0 and not all(t == all_tr[0] for t in all_tr):
    for r_, t_ in enumerate(all_tr):
        print("bagpu-test rank %d trace %s" % (r_, t_), file=sys.stderr)
assert all(t == all_tr[0] for t in all_tr), "ranks disagree on the trace"
other = s.get_other_residuals()     # collective: every rank calls it
if rank == 0:
    np.savez(out, un_maha=other["un_maha"], bin_maha=other["bin_maha"], imu_maha=other["imu_maha"], un_cov_inv=other["un_cov_inv"], iterations=sm.iterations, result=sm.result,
             post=[t["post_solve_norm"] for t in tr], pre=[t["pre_solve_norm"] for t in tr],
             median=[t["median_maha"] for t in tr], delta=[t["delta_norm"] for t in tr],
             poses=np.asarray(poses["t_wp"] if isinstance(poses, dict) else poses, float),
             lms=np.asarray(lms["x_w"] if isinstance(lms, dict) else lms, float) if lms is not None else np.zeros(0),
             reliable=np.asarray(lms["is_reliable"]) if lms is not None else np.zeros(0),
             outliers=np.asarray(lms["num_outlier_residuals"]) if lms is not None else np.zeros(0))
s.close()
dist.barrier()
dist.destroy_process_group()
