"""Developer tool: run the CUDA path and the CPU oracle on the same synthetic problem and print
where they differ (run under gpurun).  Not part of the product."""
import sys, time, json
import numpy as np
sys.path.insert(0, ".")
from ba_b200 import synth
from ba_b200.capi import BundleAdjusterGPU
from ba_b200.abi import PATTERN_U, PATTERN_W, PATTERN_S
from oracle.oracle_py import Oracle


def setup(cls, p, iters_opts):
    s = cls(p.lm_dim, p.pose_dim)
    o = s.default_options()
    o.error_change_threshold = 0; o.param_change_threshold = 0; o.record_deltas = 1
    if hasattr(p, "gravity"):
        o.gravity[0], o.gravity[1], o.gravity[2] = [float(x) for x in p.gravity]
    for k, v in iters_opts.items():
        setattr(o, k, v)
    s.set_options(o)
    s.load(p)
    return s


def rel(a, b):
    a = np.asarray(a, float); b = np.asarray(b, float)
    d = np.linalg.norm(a - b); n = max(np.linalg.norm(b), 1e-300)
    return d / n


def compare(name, p, iters=5, **opts):
    print("=====", name, p.counts)
    g = setup(BundleAdjusterGPU, p, opts); o = setup(Oracle, p, opts)
    g.build(); o.build()
    sg, so = g.get_system(), o.get_system()
    for k in ("S", "rhs_sc", "rhs_p", "rhs_l", "v_inv"):
        if sg[k] is None or np.size(sg[k]) == 0:
            continue
        print("  build %-7s rel diff %.3e   (|oracle| %.3e)" % (k, rel(np.triu(sg[k]) if k == "S" else sg[k], np.triu(so[k]) if k == "S" else so[k]), np.linalg.norm(so[k])))
    for w, nm in ((PATTERN_U, "U"), (PATTERN_W, "W"), (PATTERN_S, "S")):
        if w == PATTERN_W and p.lm_dim == 0:
            continue
        cg, rg = g.get_pattern(w); co, ro = o.get_pattern(w)
        print("  pattern %s equal: %s (nnzb %d)" % (nm, bool(np.array_equal(cg, co) and np.array_equal(rg, ro)), len(ro)))
    g = setup(BundleAdjusterGPU, p, opts); o = setup(Oracle, p, opts)
    t = time.time(); sg = g.solve(iters); tg = time.time() - t
    t = time.time(); so = o.solve(iters); to = time.time() - t
    print("  solve: gpu %.3fs oracle %.3fs   result gpu %d oracle %d iters %d/%d" % (tg, to, sg.result, so.result, sg.iterations, so.iterations))
    trg, tro = g.get_trace(), o.get_trace()
    for i, (a, b) in enumerate(zip(trg, tro)):
        keys = ["proj_error_build", "binary_error_build", "unary_error_build", "inertial_error_build", "median_maha", "sd_numerator", "sd_denominator", "sd_norm", "gn_norm",
                "pre_solve_norm", "post_solve_norm", "delta_norm", "trust_region_after"]
        worst = max((abs(a[k] - b[k]) / max(abs(b[k]), 1e-300), k) for k in keys if b[k] != 0 or a[k] != 0)
        dpg, dlg = g.get_deltas(i); dpo, dlo = o.get_deltas(i)
        print("  it %d inner %d/%d post %.9g/%.9g worst-scalar %.2e (%s) delta_p rel %.2e delta_l rel %.2e" % (
            i, a["inner_iterations"], b["inner_iterations"], a["post_solve_norm"], b["post_solve_norm"], worst[0], worst[1],
            rel(dpg, dpo), rel(dlg, dlo) if len(dlo) else 0.0))
    pg, po = g.get_poses(), o.get_poses()
    print("  final poses rel %.2e  cost rel %.2e" % (rel(pg["t_wp"], po["t_wp"]), abs(sg.post_solve_norm - so.post_solve_norm) / max(so.post_solve_norm, 1e-300)))
    if p.lm_dim:
        lg, lo = g.get_landmarks(), o.get_landmarks()
        print("  final landmarks x_w rel %.2e outliers equal %s reliable equal %s" % (rel(lg["x_w"], lo["x_w"]), bool(np.array_equal(lg["num_outlier_residuals"], lo["num_outlier_residuals"])), bool(np.array_equal(lg["is_reliable"], lo["is_reliable"]))))
        rg, ro = g.get_proj_residuals(), o.get_proj_residuals()
        print("  residual rel %.2e weight rel %.2e" % (rel(rg["residual"], ro["residual"]), rel(rg["weight"], ro["weight"])))
    ph, launches = g.get_phase_ms()
    print("  gpu phases ms:", {k: round(v, 3) for k, v in ph.items()}, "launches", launches)


if __name__ == "__main__":
    which = sys.argv[1:] or ["c1", "c2", "c4", "c5"]
    if "c1" in which:
        compare("c1-mini", synth.make_c1(n_poses=12, n_landmarks=200, obs_per_lm=8))
        compare("c1-mini-tvs", synth.make_c1(n_poses=10, n_landmarks=150, obs_per_lm=6, t_vs=[0.01, -0.02, 0.03, 0.9993, 0.1, -0.05, 0.02]))
    if "c2" in which:
        compare("c2-mini", synth.make_c2(n_cams=100, n_points=3000))
    if "c4" in which:
        compare("c4-mini", synth.make_c4(n_cams=60, n_points=2000))
    if "c5" in which:
        compare("c5-mini", synth.make_c5(n_poses=200))
    if "c3" in which:
        compare("c3-mini", synth.make_c3(n_keyframes=40, n_landmarks=300, obs_per_lm=10))
        compare("c3-mini-d9-nolm", synth.make_c3(n_keyframes=30, n_landmarks=0, pose_dim=9, lm_dim=0))
    if "c2mid" in which:
        compare("c2-mid", synth.make("c2", 0.1), iters=3)
    if "c4mid" in which:
        compare("c4-mid", synth.make("c4", 0.02), iters=3)
    if "c1full" in which:
        compare("c1-full", synth.make_c1(), iters=10)
