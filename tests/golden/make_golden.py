"""Generates tests/golden/*.npz: small seeded problems plus the CPU oracle's results on them.

The reference itself cannot run here (Eigen3/Sophus/Calibu/TBB absent), so these are oracle
outputs, not reference outputs ("parity unpinned", oracle/minimath.h).  They freeze the oracle:
the CPU suite checks that the oracle still reproduces them, the GPU suite checks the CUDA path
against them without needing the oracle's numbers recomputed.
Run from the repo root:  python tests/golden/make_golden.py
"""
import os
import sys
import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", ".."))
from ba_b200 import synth  # noqa: E402
from oracle.oracle_py import Oracle  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))

CASES = {
    "c1_tiny": (lambda: synth.make_c1(n_poses=8, n_landmarks=60, obs_per_lm=5, seed=21, outlier_frac=0.05), 6),
    "c2_tiny": (lambda: synth.make_c2(n_cams=16, n_points=120, seed=22, window=8), 5),
    "c4_tiny": (lambda: synth.make_c4(n_cams=20, n_points=150, seed=24), 5),
    "c5_tiny": (lambda: synth.make_c5(n_poses=40, seed=25), 5),
    "c3_tiny": (lambda: synth.make_c3(n_keyframes=16, n_landmarks=60, obs_per_lm=5, seed=23), 4),
}


def run_case(make, iters, solver_cls=Oracle):
    p = make()
    s = solver_cls(p.lm_dim, p.pose_dim)
    o = s.default_options()
    o.error_change_threshold = 0; o.param_change_threshold = 0; o.record_deltas = 1
    if hasattr(p, "gravity"):
        o.gravity[0], o.gravity[1], o.gravity[2] = [float(x) for x in p.gravity]
    s.set_options(o)
    s.load(p)
    summ = s.solve(iters)
    tr = s.get_trace()
    out = {"in_" + k: v for k, v in p.arrays().items()}
    out["meta"] = np.array([p.lm_dim, p.pose_dim, iters])
    out["gravity"] = np.array(getattr(p, "gravity", [0, 0, 0]), float)
    keys = sorted(tr[0].keys())
    out["trace_keys"] = np.array(keys)
    out["trace"] = np.array([[float(t[k]) for k in keys] for t in tr])
    for i in range(len(tr)):
        dp, dl = s.get_deltas(i)
        out["delta_p_%d" % i] = dp; out["delta_l_%d" % i] = dl
    ps = s.get_poses()
    out["out_t_wp"] = ps["t_wp"]; out["out_v_w"] = ps["v_w"]; out["out_b"] = ps["b"]
    if p.lm_dim:
        lm = s.get_landmarks()
        out["out_x_w"] = lm["x_w"]; out["out_outliers"] = lm["num_outlier_residuals"]; out["out_reliable"] = lm["is_reliable"]
        pr = s.get_proj_residuals()
        out["out_residual"] = pr["residual"]; out["out_weight"] = pr["weight"]
    out["summary"] = np.array([summ.post_solve_norm, summ.pre_solve_norm, summ.delta_norm, summ.trust_region_size, summ.result, summ.iterations], float)
    return out


if __name__ == "__main__":
    for name, (make, iters) in CASES.items():
        out = run_case(make, iters)
        np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
        print(name, "post", out["summary"][0], "bytes", os.path.getsize(os.path.join(HERE, name + ".npz")))
