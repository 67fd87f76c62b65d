"""ba_b200 — B200-native (sm_100a CUDA) optimisation inner loop of arpg/ba's ba::BundleAdjuster.

  ba_b200.problem.Problem       host-side bookkeeping with the reference's Add* API
  ba_b200.capi.BundleAdjusterGPU  the solver behind the C ABI of include/ba_b200.h
  ba_b200.synth                 seeded synthetic problems of the BASELINE.json shapes
"""
from .problem import Problem  # noqa: F401
