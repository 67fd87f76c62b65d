"""Diagnostic: run a workload at given scales and print result / trace (development aid)."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ba_b200 import synth
from ba_b200.capi import BundleAdjusterGPU
name = sys.argv[1]; iters = int(sys.argv[2])
for sc in [float(x) for x in sys.argv[3:]]:
    p = synth.make(name, sc)
    s = BundleAdjusterGPU(p.lm_dim, p.pose_dim)
    o = s.default_options(); o.error_change_threshold = 0; o.param_change_threshold = 0
    if hasattr(p, "gravity"):
        o.gravity[0], o.gravity[1], o.gravity[2] = [float(x) for x in p.gravity]
    s.set_options(o); s.load(p)
    t0 = time.time(); sm = s.solve(iters); dt = time.time() - t0
    ph, n = s.get_phase_ms()
    print("scale", sc, "result", sm.result, "iters", sm.iterations, "wall %.3f" % dt, "launches", n)
    print("  phases", {k: round(v, 3) for k, v in ph.items()})
    for t in s.get_trace():
        print("  pre %.9g post %.9g gn %.4g sd %.4g inner %d res %d tr %.4g" % (t["pre_solve_norm"], t["post_solve_norm"], t["gn_norm"], t["sd_norm"], t["inner_iterations"], t["result"], t["trust_region_after"]))
    s.close()
