"""The on-disk dump format of the path (reference src/BundleAdjuster.cpp:600-620, Utils.h:66) and its
binary twin (ba_b200/dumpio.py): format, round trips, and the reference's own cross-check of the dumps
(reference matlab/marginalization.py:47-52: S and rhs rebuilt from the dumped Jacobians).  CPU only: the
files come from the oracle; the GPU's s.txt / rhs.txt are compared with them in test_zz_dump_gpu.py."""
import os
import numpy as np
import pytest

from ba_b200 import dumpio, synth
from ba_b200.abi import PATTERN_S
from oracle.oracle_py import Oracle


def make_solver(p, **opts):
    s = Oracle(p.lm_dim, p.pose_dim)
    o = s.default_options()
    o.error_change_threshold = 0
    o.param_change_threshold = 0
    for k, v in opts.items():
        setattr(o, k, v)
    s.set_options(o)
    s.load(p)
    return s


def test_csv_is_eigens_long_csv_format(tmp_path):
    a = np.array([[1.0, -2.5, 1 / 3], [1e-300, 0.0, 123456789.123456789]])
    p = str(tmp_path / "m.txt")
    dumpio.write_csv(p, a)
    text = open(p).read()
    assert text.count("\n") == 1 and not text.endswith("\n")           # "\n" between rows only
    assert text.split("\n")[0] == "1, -2.5, 0.33333333333333331"        # ", " between coefficients
    assert np.array_equal(dumpio.read_csv(p), a)                        # 17 digits: bit-exact round trip
    assert np.array_equal(np.loadtxt(p, delimiter=","), a)              # what marginalization.py does
    v = np.array([0.1, -7.0, 2e10])
    dumpio.write_csv(p, v)
    assert open(p).read() == "0.10000000000000001\n-7\n20000000000"
    assert np.array_equal(dumpio.read_csv(p), v)
    open(p, "w").close()
    assert dumpio.read_csv(p).shape == (0,)
    # Eigen's kLongFmt variant (";\n" row separator, Utils.h:65) is accepted too
    open(p, "w").write("1, 2;\n3, 4")
    assert np.array_equal(dumpio.read_csv(p), [[1, 2], [3, 4]])


def test_binary_twin_round_trips(tmp_path):
    rng = np.random.default_rng(3)
    p = str(tmp_path / "x.bad")
    a = rng.normal(size=(7, 5))
    dumpio.write_bin(p, a)
    assert np.array_equal(dumpio.read_bin(p), a)
    v = rng.normal(size=11)
    dumpio.write_bin(p, v)
    assert np.array_equal(dumpio.read_bin(p), v)
    # block-CSC with an explicit zero block kept by the pattern and an empty column
    d = np.zeros((6, 9))
    d[0:2, 0:3] = rng.normal(size=(2, 3)); d[4:6, 0:3] = rng.normal(size=(2, 3)); d[2:4, 6:9] = rng.normal(size=(2, 3))
    cp, ri, val = dumpio.dense_to_block_csc(d, 2, 3)
    assert cp.tolist() == [0, 2, 2, 3] and ri.tolist() == [0, 2, 1]
    assert np.array_equal(dumpio.block_csc_to_dense(3, cp, ri, val), d)
    pat = (np.array([0, 3, 3, 4], np.uint32), np.array([0, 1, 2, 1], np.uint32))
    cp2, ri2, val2 = dumpio.dense_to_block_csc(d, 2, 3, pat)
    assert len(val2) == 4 and not val2[1].any()
    dumpio.write_bin(p, block_csc=(cp2, ri2, val2), shape_blocks=(3, 3))
    assert np.array_equal(dumpio.read_bin(p), d)
    cp3, ri3, val3, shape = dumpio.read_bin(p, dense=False)
    assert shape == (3, 3) and np.array_equal(cp3, cp2) and np.array_equal(ri3, ri2) and np.array_equal(val3, val2)
    open(p, "wb").write(b"nonsense")
    with pytest.raises(ValueError):
        dumpio.read_bin(p)


@pytest.mark.parametrize("lm_dim", [1, 3])
def test_oracle_dump_is_consistent_with_its_system(tmp_path, lm_dim):
    """The five files hold what get_system returns, and S / rhs follow from the dumped Jacobians the way the
    reference's marginalization.py rebuilds them."""
    if lm_dim == 1:
        p = synth.make_c1(n_poses=6, n_landmarks=40, obs_per_lm=4, seed=11)
    else:
        p = synth.make_c2(n_cams=10, n_points=50, seed=11, window=6)
    s = make_solver(p)
    s.build()
    s.dump(str(tmp_path))
    d = dumpio.read_dump(str(tmp_path))
    assert sorted(d) == ["j_l", "j_pr", "r_pr", "rhs", "s"]
    sysm = s.get_system()
    n = s.n_active_poses * 6
    assert d["s"].shape == (n, n) and np.array_equal(d["s"], np.triu(sysm["S"]))   # bit-exact through the text
    assert np.array_equal(d["rhs"], sysm["rhs_sc"])
    n_res = len(p.arrays()["proj_z"])
    assert d["j_pr"].shape == (2 * n_res, n) and d["r_pr"].shape == (2 * n_res,)
    assert d["j_l"].shape == (2 * n_res, s.n_active_landmarks * lm_dim)
    # r_pr = sqrt(w) r of the stored residuals (cpp:1386-1387)
    pr = s.get_proj_residuals()
    assert np.allclose(d["r_pr"], (pr["residual"] * np.sqrt(pr["weight"])[:, None]).reshape(-1), rtol=1e-14, atol=0)
    jp, jl, r = d["j_pr"], d["j_l"], d["r_pr"]
    assert np.allclose(jp.T @ r, sysm["rhs_p"], rtol=1e-11, atol=1e-11 * np.abs(sysm["rhs_p"]).max())
    assert np.allclose(jl.T @ r, sysm["rhs_l"], rtol=1e-11, atol=1e-11 * np.abs(sysm["rhs_l"]).max())
    V = jl.T @ jl
    L = lm_dim
    Vinv = np.zeros_like(V)
    for l in range(s.n_active_landmarks):
        Vinv[L * l:L * l + L, L * l:L * l + L] = np.linalg.inv(V[L * l:L * l + L, L * l:L * l + L])
    assert np.allclose(V - V * (Vinv != 0), 0)                                     # V is block diagonal
    W = jp.T @ jl
    S = jp.T @ jp - W @ Vinv @ W.T
    rhs = jp.T @ r - W @ Vinv @ (jl.T @ r)
    scale = np.abs(S).max()
    assert np.allclose(np.triu(S), d["s"], rtol=0, atol=1e-11 * scale)
    assert np.allclose(rhs, d["rhs"], rtol=0, atol=1e-11 * np.abs(rhs).max())
    # CSV -> binary twin with the symbolic S pattern (explicit zero blocks kept) -> same matrices
    out = dumpio.csv_dump_to_bin(str(tmp_path), pose_dim=6, lm_dim=lm_dim, s_pattern=s.get_pattern(PATTERN_S))
    for name, path in out.items():
        assert np.array_equal(dumpio.read_bin(path), d[name]), name
    cp, ri, val, shape = dumpio.read_bin(out["s"], dense=False)
    cps, ris = s.get_pattern(PATTERN_S)
    assert np.array_equal(cp, cps) and np.array_equal(ri, ris) and shape == (s.n_active_poses, s.n_active_poses)
    assert os.path.getsize(out["j_pr"]) < os.path.getsize(os.path.join(str(tmp_path), "j_pr.txt")) / 2


def test_pose_graph_dump_has_no_projection_files(tmp_path):
    p = synth.make_c5(n_poses=30, seed=5)
    s = make_solver(p)
    s.build()
    s.dump(str(tmp_path))
    d = dumpio.read_dump(str(tmp_path))
    assert sorted(d) == ["rhs", "s"]
    sysm = s.get_system()
    assert np.array_equal(d["s"], np.triu(sysm["S"])) and np.array_equal(d["rhs"], sysm["rhs_sc"])


def _dogleg_from_dump(directory, trust_region):
    """The step the reference's dogleg prototype takes (reference matlab/dogleg.m:12-41) from the dumped J and r,
    with the beta of src/BundleAdjuster.cpp:1007-1008 (same expression in dogleg.m:36: -b^2, not -b)."""
    d = dumpio.read_dump(directory)
    J = np.hstack([d["j_pr"], d["j_l"]]); r = d["r_pr"]
    n_p = d["j_pr"].shape[1]
    g = J.T @ r
    top, bottom = g @ g, np.linalg.norm(J @ g) ** 2
    sd = top / bottom * g
    gn = np.linalg.solve(J.T @ J, g)
    out = dict(top=top, bottom=bottom, sd=sd, gn=gn, n_p=n_p)
    if np.linalg.norm(sd) >= trust_region:
        out["step"], out["branch"] = trust_region / np.linalg.norm(sd) * sd, "sd"
    elif np.linalg.norm(gn) <= trust_region:
        out["step"], out["branch"] = gn, "gn"
    else:
        diff = gn - sd
        a, b, c = diff @ diff, 2 * (diff @ sd), sd @ sd - trust_region ** 2
        beta = (-b * b + np.sqrt(b * b - 4 * a * c)) / (2 * a)
        out["step"], out["branch"], out["beta"] = sd + beta * diff, "blend", beta
        out["textbook"] = sd + (-b + np.sqrt(b * b - 4 * a * c)) / (2 * a) * diff
    return out


@pytest.mark.parametrize("trust_region,branch", [(1e-3, "sd"), (0.1, "blend"), (10.0, "gn")])
def test_dogleg_step_follows_the_prototype(tmp_path, monkeypatch, trust_region, branch):
    p = synth.make_c1(n_poses=6, n_landmarks=40, obs_per_lm=4, seed=11)
    s = make_solver(p, record_deltas=1, write_reduced_camera_matrix=1, trust_region_size=trust_region)
    monkeypatch.chdir(tmp_path)
    s.solve(1)
    tr = s.get_trace()[0]
    ref = _dogleg_from_dump(str(tmp_path), trust_region)
    assert ref["branch"] == branch and tr["inner_iterations"] == 1
    assert abs(tr["sd_numerator"] - ref["top"]) <= 1e-12 * ref["top"]
    assert abs(tr["sd_denominator"] - ref["bottom"]) <= 1e-12 * ref["bottom"]
    assert abs(tr["sd_norm"] - np.linalg.norm(ref["sd"])) <= 1e-12 * tr["sd_norm"]
    if branch != "sd":                                                 # the GN step is only computed when needed
        assert abs(tr["gn_norm"] - np.linalg.norm(ref["gn"])) <= 1e-9 * tr["gn_norm"]
    dp, dl = s.get_deltas(0)
    step = np.concatenate([dp, dl])
    assert np.linalg.norm(step - ref["step"]) <= 1e-9 * np.linalg.norm(step)
    # summary_.delta_norm adds the two norms (cpp:26)
    assert abs(tr["delta_norm"] - (np.linalg.norm(dp) + np.linalg.norm(dl))) <= 1e-12 * tr["delta_norm"]
    if branch == "blend":
        # quirk Q4: the step is not on the trust-region boundary, the textbook root would be
        assert abs(np.linalg.norm(ref["textbook"]) - trust_region) < 1e-12
        assert abs(np.linalg.norm(step) - trust_region) > 1e-5
