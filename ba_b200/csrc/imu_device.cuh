// imu_device.cuh — inertial residual device code (placeholder until the RK4 integrator lands).
#pragma once
#include "ba_gpu.h"
#include "launch.h"
namespace bagpu {
template <int D> __device__ void imu_build_one(const DeviceProblem&, const OtherResiduals&, int) {}
template <int D> __device__ void imu_sd_one(const DeviceProblem&, const OtherResiduals&, int) {}
template <int D> __device__ void imu_eval_one(const DeviceProblem&, const OtherResiduals&, int) {}
}
