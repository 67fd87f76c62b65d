"""s.txt / rhs.txt written by the CUDA path (bagpu_dump, options.write_reduced_camera_matrix; reference
src/BundleAdjuster.cpp:600-606) against the same files from the CPU oracle, read through ba_b200.dumpio."""
import os
import numpy as np
import pytest

from ba_b200 import dumpio, synth
from ba_b200.abi import PATTERN_S

pytestmark = pytest.mark.gpu


def _mk(cls, p, **opts):
    s = cls(p.lm_dim, p.pose_dim)
    o = s.default_options()
    o.error_change_threshold = 0
    o.param_change_threshold = 0
    for k, v in opts.items():
        setattr(o, k, v)
    s.set_options(o)
    s.load(p)
    return s


def _pair(p, **opts):
    from ba_b200.capi import BundleAdjusterGPU
    from oracle.oracle_py import Oracle
    return _mk(BundleAdjusterGPU, p, **opts), _mk(Oracle, p, **opts)


def rel(a, b):
    return np.linalg.norm(np.asarray(a) - np.asarray(b)) / max(np.linalg.norm(b), 1e-300)


@pytest.mark.parametrize("make", [lambda: synth.make_c1(n_poses=8, n_landmarks=120, obs_per_lm=5, seed=21),
                                  lambda: synth.make_c2(n_cams=14, n_points=150, seed=21, window=8),
                                  lambda: synth.make_c5(n_poses=40, seed=21)], ids=["lm1", "lm3", "pose_graph"])
def test_dump_files_match_the_oracles(tmp_path, make):
    p = make()
    g, o = _pair(p)
    g.build(); o.build()
    dg, do = str(tmp_path / "gpu"), str(tmp_path / "oracle")
    os.mkdir(dg); os.mkdir(do)
    g.dump(dg); o.dump(do)
    a, b = dumpio.read_dump(dg), dumpio.read_dump(do)
    # j_pr / r_pr / j_l: the build never forms J; the dump evaluates the rows for the files (k_jacobian_rows)
    assert sorted(a) == sorted(b)
    for k in ("j_pr", "r_pr", "j_l"):
        if k in b:
            assert a[k].shape == b[k].shape, k
            assert rel(a[k], b[k]) < 1e-11, (k, rel(a[k], b[k]))
            assert np.array_equal(a[k] != 0, b[k] != 0), k                # same block pattern, element for element
    text = open(os.path.join(dg, "s.txt")).read()
    assert not text.endswith("\n") and ", " in text                    # Eigen kLongCsvFmt (Utils.h:66)
    n = g.n_active_poses * p.pose_dim
    assert a["s"].shape == b["s"].shape == (n, n) and not np.tril(a["s"], -1).any()
    assert rel(a["s"], b["s"]) < 1e-12 and rel(a["rhs"], b["rhs"]) < 1e-11
    assert np.array_equal(a["s"], np.triu(g.get_system()["S"]))       # the text round-trips bit-exactly
    # binary twin on the symbolic pattern: same matrix, same block structure as the oracle's
    out = dumpio.csv_dump_to_bin(dg, pose_dim=p.pose_dim, lm_dim=p.lm_dim, s_pattern=g.get_pattern(PATTERN_S))
    cp, ri, val, shape = dumpio.read_bin(out["s"], dense=False)
    cpo, rio = o.get_pattern(PATTERN_S)
    assert np.array_equal(cp, cpo) and np.array_equal(ri, rio)
    assert np.array_equal(dumpio.read_bin(out["s"]), a["s"])


def test_write_reduced_camera_matrix_option_dumps_every_iteration(tmp_path, monkeypatch):
    p = synth.make_c1(n_poses=8, n_landmarks=120, obs_per_lm=5, seed=22)
    dg, do = str(tmp_path / "gpu"), str(tmp_path / "oracle")
    os.mkdir(dg); os.mkdir(do)
    g, o = _pair(p, write_reduced_camera_matrix=1)
    monkeypatch.chdir(dg)
    sg = g.solve(3)
    monkeypatch.chdir(do)
    so = o.solve(3)
    assert sg.iterations == so.iterations == 3
    a, b = dumpio.read_dump(dg), dumpio.read_dump(do)
    # the files left behind are those of the last iteration's system
    assert rel(a["s"], b["s"]) < 1e-6 and rel(a["rhs"], b["rhs"]) < 1e-6
    g1, _ = _pair(p)
    g1.build()
    first = g1.get_system()
    assert rel(a["rhs"], first["rhs_sc"]) > 1e-3                       # not the first iteration's
    assert abs(sg.post_solve_norm - so.post_solve_norm) <= 1e-9 * so.post_solve_norm


@pytest.mark.parametrize("trust_region,branch", [(1e-3, "sd"), (0.1, "blend"), (10.0, "gn")])
def test_dogleg_step_follows_the_prototype_on_the_gpu(tmp_path, monkeypatch, trust_region, branch):
    """The three branches of the dogleg rule (reference matlab/dogleg.m:12-41, beta of cpp:1007-1008) rebuilt in NumPy
    from the oracle's dumped J and r, against the step the CUDA path takes."""
    from tests.test_dump_cpu import _dogleg_from_dump
    p = synth.make_c1(n_poses=6, n_landmarks=40, obs_per_lm=4, seed=11)
    g, o = _pair(p, record_deltas=1, trust_region_size=trust_region)
    o.options.write_reduced_camera_matrix = 1
    o.set_options(o.options)
    monkeypatch.chdir(tmp_path)
    o.solve(1); g.solve(1)
    ref = _dogleg_from_dump(str(tmp_path), trust_region)
    assert ref["branch"] == branch
    tg = g.get_trace()[0]
    assert tg["inner_iterations"] == 1
    assert abs(tg["sd_numerator"] - ref["top"]) <= 1e-11 * ref["top"]
    assert abs(tg["sd_denominator"] - ref["bottom"]) <= 1e-11 * ref["bottom"]
    dp, dl = g.get_deltas(0)
    step = np.concatenate([dp, dl])
    assert np.linalg.norm(step - ref["step"]) <= 1e-9 * np.linalg.norm(step)
    if branch == "blend":
        assert abs(np.linalg.norm(step) - trust_region) > 1e-5           # quirk Q4 reproduced
