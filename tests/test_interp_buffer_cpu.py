"""include/ba/InterpolationBuffer.h (binary search) against the literal restatement of the reference's
ba::InterpolationBufferT (oracle/interp_buffer_py.py, reference include/ba/InterpolationBuffer.h:38-210):
bit-exact elements, indices and range lengths, including queries that hit a stored time stamp exactly and
ranges clipped to the buffer.  Host-only (the caller-side step before AddImuResidual)."""
import os
import struct
import subprocess
import numpy as np
import pytest

from oracle.interp_buffer_py import InterpolationBufferRef

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "tests", "cpp", "interp_buffer_driver.cpp")
EXE = os.path.join(ROOT, "tests", "cpp", "build", "interp_buffer_driver")


def build_driver():
    os.makedirs(os.path.dirname(EXE), exist_ok=True)
    deps = [SRC] + [os.path.join(ROOT, "include", "ba", h) for h in ("InterpolationBuffer.h", "BundleAdjuster.h", "MiniTypes.h")]
    if not (os.path.exists(EXE) and all(os.path.getmtime(d) <= os.path.getmtime(EXE) for d in deps)):
        subprocess.check_call(["g++", "-std=c++17", "-O2", "-Wall", "-ffp-contract=off", "-I" + os.path.join(ROOT, "include"), SRC, "-o", EXE])
    return EXE


def run_cpp(tmp_path, rows, queries):
    fin, fout = str(tmp_path / "in.bin"), str(tmp_path / "out.bin")
    with open(fin, "wb") as f:
        f.write(struct.pack("<I", len(rows))); f.write(np.ascontiguousarray(rows, dtype=np.float64).tobytes())
        f.write(struct.pack("<I", len(queries))); f.write(np.ascontiguousarray(queries, dtype=np.float64).tobytes())
    subprocess.check_call([build_driver(), fin, fout])
    return np.fromfile(fout, dtype=np.float64)


def make_rows(rng, n, kind):
    if kind == "regular":
        dt = np.full(n, 0.005)
    elif kind == "jitter":
        dt = 0.005 * rng.uniform(0.5, 1.5, n)
    else:   # long gaps and bursts: the average_dt guess lands far from the bracket, on either side
        dt = np.where(rng.uniform(size=n) < 0.05, rng.uniform(0.5, 2.0, n), rng.uniform(1e-4, 1e-2, n))
    t = 100.0 + np.cumsum(dt)
    return np.concatenate([rng.normal(size=(n, 6)), t[:, None]], axis=1)


@pytest.mark.parametrize("kind", ["regular", "jitter", "gaps"])
@pytest.mark.parametrize("n", [1, 2, 3, 400])
def test_matches_the_reference_walk(tmp_path, kind, n):
    rng = np.random.default_rng(100 + n)
    rows = make_rows(rng, n, kind)
    t = rows[:, 6]
    ref = InterpolationBufferRef()
    for r in rows:
        ref.AddElement(r)
    q = []
    inside = rng.uniform(t[0], t[-1], 60) if n > 1 else np.array([t[0]])
    stamps = t[rng.integers(0, n, 40)]                                   # exact hits on stored time stamps
    for x in np.concatenate([inside, stamps, [t[0], t[-1], t[0] - 1.0]]):
        q.append((0, x, 0))
    for op in (1, 2):
        for _ in range(30):
            a, b = np.sort(rng.uniform(t[0] - 0.1, t[-1] + 0.1, 2))
            q.append((op, a, b))
        for _ in range(20):                                              # pose times that coincide with IMU samples
            i, j = np.sort(rng.integers(0, n, 2))
            q.append((op, t[i], t[j]))
        q += [(op, t[0] - 5, t[-1] + 5), (op, t[-1] + 1, t[-1] + 2), (op, t[0] - 2, t[0] - 1), (op, t[0], t[0])]
    out = run_cpp(tmp_path, rows, np.array(q))
    assert out[0] == ref.start_time and out[1] == ref.end_time and out[2] == ref.average_dt
    o = 3
    for op, a, b in q:
        if op == 0:
            if n == 1:
                e, idx = rows[0], 0                                      # the source divides by average_dt = 0 here
            else:
                e, idx = ref.GetElement(a)
            assert np.array_equal(out[o:o + 7], e), (op, a)
            assert out[o + 7] == idx, (a, out[o + 7], idx)
            o += 8
        else:
            if n == 1:
                # one element: the clipped start is inside the buffer iff a <= t[0]; both ends return that element
                r = np.array([rows[0], rows[0]]) if a <= t[0] else np.zeros((0, 7))
            else:
                r = ref.GetRange(a, b)
            cnt = int(out[o]); o += 1
            assert cnt == len(r), (op, a, b, cnt, len(r))
            assert np.array_equal(out[o:o + 7 * cnt].reshape(-1, 7), r), (op, a, b)
            o += 7 * cnt
    assert out[o] == -1 and out[o + 1] == 0 and o + 2 == len(out)       # Clear()


def test_range_is_what_add_imu_residual_integrates(tmp_path):
    """First and last rows sit exactly on the pose times, stored samples in between are untouched, times increase."""
    rng = np.random.default_rng(5)
    rows = make_rows(rng, 300, "jitter")
    t = rows[:, 6]
    a, b = t[17] + 0.001, t[250] - 0.002
    out = run_cpp(tmp_path, rows, np.array([(2, a, b)]))
    cnt = int(out[3]); r = out[4:4 + 7 * cnt].reshape(-1, 7)
    assert r[0, 6] == a and r[-1, 6] == b and np.all(np.diff(r[:, 6]) > 0)
    assert np.array_equal(r[1:-1], rows[18:250])
    s = (a - t[17]) / (t[18] - t[17])
    assert np.array_equal(r[0, :6], rows[17, :6] * (1 - s) + rows[18, :6] * s)
