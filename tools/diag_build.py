"""Development aid: time only the Schur build (bagpu_build) of a workload, n times."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ba_b200 import synth
from ba_b200.capi import BundleAdjusterGPU
name = sys.argv[1]; n = int(sys.argv[2]); sc = float(sys.argv[3]) if len(sys.argv) > 3 else 1.0
p = synth.make(name, sc)
s = BundleAdjusterGPU(p.lm_dim, p.pose_dim)
o = s.default_options(); o.error_change_threshold = 0; o.param_change_threshold = 0
s.set_options(o); s.load(p)
for i in range(n):
    s.build()
ph, nl = s.get_phase_ms()
print(name, "builds", n, {k: round(v / n, 4) for k, v in ph.items() if k != "total"})
s.close()
