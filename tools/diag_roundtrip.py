"""Development aid: trace of the noise-free round-trip problem (tests/test_parity_gpu.py)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from ba_b200 import synth
from ba_b200.capi import BundleAdjusterGPU
for anchor in (0, 50):
    p = synth.make_c4(n_cams=400, n_points=60000, lm_dim=3, anchor_every=anchor)
    a = p.arrays()
    z, _ = synth.project(p.truth["T"][a["proj_meas_pose"]], p.truth["X"][a["proj_lm"]])
    p.set_bulk(proj_z=z)
    s = BundleAdjusterGPU(3, 6)
    o = s.default_options(); o.error_change_threshold = 0; o.param_change_threshold = 0
    s.set_options(o); s.load(p)
    sm = s.solve(25)
    print("anchor", anchor, "result", sm.result, "iters", sm.iterations)
    for t in s.get_trace():
        print("  pre %.6g post %.6g gn %.4g sd %.4g inner %d tr %.4g" % (t["pre_solve_norm"], t["post_solve_norm"], t["gn_norm"], t["sd_norm"], t["inner_iterations"], t["trust_region_after"]))
    s.close()
