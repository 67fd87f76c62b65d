"""Development aid: wall-clock split of one end-to-end Solve(K) (host buffers in, refined state out)."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ba_b200 import synth
from ba_b200.capi import BundleAdjusterGPU
name = sys.argv[1]; K = int(sys.argv[2]) if len(sys.argv) > 2 else 10
p = synth.make(name, 1.0)
a = p.arrays()
if len(sys.argv) > 3 and sys.argv[3] == "pinned":
    from ba_b200.capi import pinned_like
    from ba_b200.problem import Problem
    pp = Problem(p.lm_dim, p.pose_dim); pp.set_bulk(**{k: pinned_like(v) for k, v in a.items()}); p = pp
    import numpy as np
    n_l = len(a["lm_x_w"])
    lm_out = dict(x_w=pinned_like(np.zeros((n_l, 4))), x_s=pinned_like(np.zeros((n_l, 4))),
                  is_reliable=pinned_like(np.zeros(n_l, np.uint8)), num_outlier_residuals=pinned_like(np.zeros(n_l, np.uint32)))
for rep in range(3):
    t = [time.time()]
    s = BundleAdjusterGPU(p.lm_dim, p.pose_dim)
    o = s.default_options(); o.error_change_threshold = 0; o.param_change_threshold = 0
    s.set_options(o); t.append(time.time())
    s.load(p); t.append(time.time())
    sm = s.solve(K); t.append(time.time())
    s.get_poses(); s.get_landmarks(out=lm_out if 'lm_out' in globals() else None); t.append(time.time())
    s.close(); t.append(time.time())
    print("rep", rep, "create %.1f load %.1f solve %.1f get %.1f close %.1f total %.1f ms" % tuple(
        [(t[i + 1] - t[i]) * 1e3 for i in range(5)] + [(t[5] - t[0]) * 1e3]), "iters", sm.iterations, flush=True)
