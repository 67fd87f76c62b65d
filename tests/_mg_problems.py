"""Problems of the multi-GPU tests (shared by tests/test_multi_gpu.py and its torchrun worker)."""
from ba_b200 import synth


def make(name, scale):
    if name == "c3mini":     # visual-inertial <double,1,15,0>: 40 keyframes, 300 landmarks, 39 IMU residuals
        return synth.make_c3(n_keyframes=40, n_landmarks=300, obs_per_lm=10)
    if name == "c5mini":     # pose graph <double,0,6,0>: binary + unary residuals, no landmarks
        return synth.make_c5(n_poses=200)
    return synth.make(name, scale)


def options(name):
    """Option overrides: the visual-inertial mini runs with IMU noise comparable to the visual information
    (cond(S) = 2.2e9 instead of 1e13 at the reference's default sigmas, see tests/test_parity_gpu.py::test_c3_visual_inertial)."""
    if name == "c3mini":
        return dict(gyro_sigma=2e-2, accel_sigma=2e-1, gyro_bias_sigma=1e-2, accel_bias_sigma=1e-1)
    return {}
