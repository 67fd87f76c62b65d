"""The ODO / UTM / IMU text-log reader and replay of the reference's integration application
(reference applications/unary_binary_imu_test/main.cpp:56-287) — include/ba/LogReplay.h, SURVEY.md 8f rank 4.

The application's data file (log.dat) is not in the reference tree (.MISSING_LARGE_BLOBS), so the test writes a synthetic
log of the same format and compares the C++ replay (driven by tests/cpp/log_replay_driver.cpp) with a literal Python
restatement of the application's bookkeeping: one pose per UTM fix dead-reckoned along the gyro path, one unary constraint
per fix, one inertial residual per interval with the measurements InterpolationBufferT::GetRange returns."""
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from ba_b200 import synth  # noqa: E402
from oracle.interp_buffer_py import InterpolationBufferRef  # noqa: E402

SRC = os.path.join(ROOT, "tests", "cpp", "log_replay_driver.cpp")
EXE = os.path.join(ROOT, "tests", "cpp", "build", "log_replay_driver")


def build_driver():
    os.makedirs(os.path.dirname(EXE), exist_ok=True)
    deps = [SRC] + [os.path.join(ROOT, "include", "ba", f) for f in ("LogReplay.h", "BundleAdjuster.h", "InterpolationBuffer.h", "MiniTypes.h")]
    if not (os.path.exists(EXE) and all(os.path.getmtime(d) <= os.path.getmtime(EXE) for d in deps)):
        subprocess.check_call(["g++", "-std=c++17", "-O2", "-Wall", "-I" + os.path.join(ROOT, "include"), SRC, "-o", EXE,
                               "-L" + os.path.join(ROOT, "ba_b200"), "-lba_b200", "-Wl,-rpath," + os.path.join(ROOT, "ba_b200")])
    return EXE


def write_log(path, seconds=8.0, seed=3):
    """A vehicle driving a gentle left curve: wheel odometry at 20 Hz, IMU at 100 Hz, a GPS fix every second; records
    interleaved by time like a real log."""
    rng = np.random.default_rng(seed)
    rec = []
    for t in np.arange(0.0, seconds, 0.05):
        v = 2.0 + 0.2 * np.sin(0.3 * t)
        rec.append((t + 0.001, "ODO %.6f %.6f %.6f" % (t + 0.001, v + 0.05, v - 0.05)))
    for t in np.arange(0.0, seconds, 0.01):
        w = np.array([0.0, 0.0, 0.07]) + rng.normal(0, 1e-3, 3)
        a = np.array([0.1 * np.cos(0.3 * t), 0.14, 9.8]) + rng.normal(0, 1e-2, 3)
        rec.append((t + 0.002, "IMU %.6f %.9f %.9f %.9f %.9f %.9f %.9f" % (t + 0.002, w[0], w[1], w[2], a[0], a[1], a[2])))
    for k, t in enumerate(np.arange(0.5, seconds, 1.0)):
        rec.append((t, "UTM %.6f %.4f %.4f %.4f" % (t, 500000.0 + 2.0 * t, 4100000.0 + 0.07 * t * t, 35.0 + 0.01 * k)))
    rec.sort(key=lambda r: r[0])
    with open(path, "w") as f:
        f.write("\n".join(r[1] for r in rec) + "\nXYZ 1 2\n")      # an unknown symbol: reported, its tokens skipped one by one
    return [r[1] for r in rec]


def replay_py(lines):
    """main.cpp:56-209 restated: returns (times, initial poses [q xyzw, t], IMU measurement counts)."""
    buf = InterpolationBufferRef()
    speed, gyro_last = 0.0, 0.0
    odo_first, inc_t = True, 0.0
    gyro_update = np.array([0, 0, 0, 1.0, 0, 0, 0])
    nodes_t, poses, counts = [], [], []
    last_gps = 0.0

    def axis(i, ang):
        w = np.zeros((1, 3)); w[0, i] = ang
        return synth.so3_exp(w)[0]

    for ln in lines:
        tok = ln.split()
        if tok[0] == "ODO":
            t, rr, rl = map(float, tok[1:4])
            if odo_first:
                inc_t = t; odo_first = False
                continue
            speed = 0.5 * (rr + rl)
            inc_t = t
        elif tok[0] == "IMU":
            t = float(tok[1]); w = np.array(list(map(float, tok[2:5]))); a = np.array(list(map(float, tok[5:8])))
            if gyro_last != 0:
                dt = t - gyro_last
                q = synth.quat_mul(synth.quat_mul(axis(2, w[2] * dt), axis(1, w[1] * dt)), axis(0, w[0] * dt))
                upd = np.concatenate([q, [0.0, speed * dt, 0.0]])
                gyro_update = synth.se3_mul(gyro_update[None], upd[None])[0]
            gyro_last = t
            buf.AddElement(np.concatenate([w, a, [t]]))
        elif tok[0] == "UTM":
            t = float(tok[1])
            if not poses:
                poses.append(np.array([0, 0, 0, 1.0, 0, 0, 0]))
            else:
                poses.append(synth.se3_mul(poses[-1][None], gyro_update[None])[0])
            nodes_t.append(t)
            if len(poses) >= 2:
                counts.append(len(buf.GetRange(last_gps, t)))
            gyro_update = np.array([0, 0, 0, 1.0, 0, 0, 0])
            last_gps = t
    return np.array(nodes_t), np.array(poses), np.array(counts)


def run_driver(tmp_path, solve):
    exe = build_driver()
    lines = write_log(tmp_path / "log.dat")
    r = subprocess.run([exe, str(tmp_path / "log.dat"), str(tmp_path / "out.bin"), "1" if solve else "0"], capture_output=True, text=True)
    return r, lines


def test_replay_bookkeeping_matches_the_application(tmp_path):
    r, lines = run_driver(tmp_path, False)
    assert r.returncode == 0, r.stderr
    assert "Unknown symbol <XYZ>" in r.stderr
    d = np.fromfile(tmp_path / "out.bin")
    n, ni, nrec = int(d[0]), int(d[1]), int(d[2])
    times, poses, counts = replay_py(lines)
    assert n == len(times) == 8 and ni == n - 1 and nrec == len(lines)
    body = d[3:3 + 8 * n].reshape(n, 8)
    assert np.array_equal(body[:, 0], times)
    assert np.allclose(body[:, 1:], poses, rtol=0, atol=1e-12)
    assert np.array_equal(d[3 + 8 * n:3 + 8 * n + ni].astype(int), counts)
    assert counts.min() > 90           # ~100 IMU samples + the two interpolated ends per one-second interval


def oracle_problem(lines):
    """The problem the replay builds, as arrays for the oracle: poses at the dead-reckoned estimates, a GPS prior per pose
    with covariance diag(1000, 1000, 30000, DBL_MAX, DBL_MAX, DBL_MAX), an inertial residual per interval."""
    from ba_b200.problem import Problem
    times, poses, counts = replay_py(lines)
    buf = InterpolationBufferRef()
    utm = []
    for ln in lines:
        tok = ln.split()
        if tok[0] == "IMU":
            v = list(map(float, tok[1:8]))
            buf.AddElement(np.array(v[1:] + v[:1]))
        elif tok[0] == "UTM":
            utm.append(list(map(float, tok[2:5])))
    utm = np.array(utm); utm -= utm[0]
    n = len(times)
    p = Problem(0, 9)
    cov_inv = np.diag(1.0 / np.array([1000, 1000, 30000, sys.float_info.max, sys.float_info.max, sys.float_info.max]))
    meas, starts = [], [0]
    for k in range(1, n):
        m = np.array(buf.GetRange(times[k - 1], times[k]))
        meas.append(m); starts.append(starts[-1] + len(m))
    p.set_bulk(pose_t_wp=poses, pose_v_w=np.zeros((n, 3)), pose_b=np.zeros((n, 6)), pose_active=np.ones(n, np.uint8),
               pose_mask_used=np.zeros(n, np.uint8), pose_mask=np.ones((n, 9), np.uint8),
               un_pose=np.arange(n), un_t_wp=np.concatenate([np.tile([0, 0, 0, 1.0], (n, 1)), utm], axis=1),
               un_cov_inv=np.tile(cov_inv.reshape(1, 36), (n, 1)), un_use_rot=np.ones(n, np.uint8),
               imu_p1=np.arange(n - 1), imu_p2=np.arange(1, n), imu_meas_start=np.array(starts), imu_meas=np.concatenate(meas),
               imu_weight=np.ones(n - 1))
    p.gravity = np.array([0, 0, 9.8])
    return p


@pytest.mark.gpu
def test_replay_solves_on_the_gpu(tmp_path):
    """Solve(25, 0.2) of the replayed problem through the CUDA path reports what the reference's algorithm yields for its own
    application: without projection residuals the inertial residuals never enter the system (quirk Q3), the
    velocity block of S is exactly zero, the factorisation fails.  The CPU oracle, fed the same problem, says the same."""
    from oracle.oracle_py import Oracle
    r, lines = run_driver(tmp_path, True)
    d = np.fromfile(tmp_path / "out.bin")
    n, ni = int(d[0]), int(d[1])
    o = 3 + 8 * n + ni
    result, iters, pre, post, unary, inertial = d[o:o + 6]
    p = oracle_problem(lines)
    orc = Oracle(0, 9)
    opt = orc.default_options(); opt.trust_region_size = 100000
    opt.gravity[0], opt.gravity[1], opt.gravity[2] = 0, 0, 9.8
    orc.set_options(opt); orc.load(p)
    so = orc.solve(25, 0.2)
    assert int(result) == so.result and int(iters) == so.iterations, (result, so.result, r.stderr)
    assert r.returncode == (3 if so.result in (4, 5) else 0)
