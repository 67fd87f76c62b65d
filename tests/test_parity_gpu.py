"""GPU parity tests: the CUDA path (through the C ABI of include/ba_b200.h) against the CPU oracle on
identical seeded inputs, against the committed golden fixtures, and — at sizes the oracle cannot
reach in seconds — through size-independent properties.

Bars (BASELINE.json north_star): double precision, final cost and per-iteration deltas within 1e-9
relative; block sparsity and indexing bit-exact.  Where the reduced system is ill conditioned the
achievable agreement of two correct solvers is cond(S) * 2^-53; those cases state their tolerance.
"""
import os
import numpy as np
import pytest

from ba_b200 import synth
from ba_b200.abi import PATTERN_S, PATTERN_U, PATTERN_W
from ba_b200.problem import Problem

pytestmark = pytest.mark.gpu

HERE = os.path.join(os.path.dirname(__file__), "golden")


def _mk(cls, p, **opts):
    s = cls(p.lm_dim, p.pose_dim)
    o = s.default_options()
    o.error_change_threshold = 0
    o.param_change_threshold = 0
    o.record_deltas = 1
    if hasattr(p, "gravity"):
        o.gravity[0], o.gravity[1], o.gravity[2] = [float(x) for x in p.gravity]
    for k, v in opts.items():
        setattr(o, k, v)
    s.set_options(o)
    s.load(p)
    return s


def gpu(p, **opts):
    from ba_b200.capi import BundleAdjusterGPU
    return _mk(BundleAdjusterGPU, p, **opts)


def oracle(p, **opts):
    from oracle.oracle_py import Oracle
    return _mk(Oracle, p, **opts)


def rel(a, b):
    a = np.asarray(a, float); b = np.asarray(b, float)
    return np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300)


TRACE_KEYS = ["proj_error_build", "binary_error_build", "unary_error_build", "inertial_error_build", "median_maha",
              "sd_numerator", "sd_norm", "gn_norm", "pre_solve_norm", "post_solve_norm", "delta_norm", "trust_region_after"]


def check_solve(p, iters, tol, cost_tol=1e-9, **opts):
    g, o = gpu(p, **opts), oracle(p, **opts)
    sg, so = g.solve(iters), o.solve(iters)
    assert sg.result == so.result and sg.iterations == so.iterations
    tg, to = g.get_trace(), o.get_trace()
    assert len(tg) == len(to)
    for i, (a, b) in enumerate(zip(tg, to)):
        assert a["inner_iterations"] == b["inner_iterations"] and a["result"] == b["result"]
        for k in TRACE_KEYS:
            assert abs(a[k] - b[k]) <= tol * max(abs(b[k]), 1e-300) + 1e-300, (i, k, a[k], b[k])
        dpg, dlg = g.get_deltas(i); dpo, dlo = o.get_deltas(i)
        assert rel(dpg, dpo) < tol, (i, rel(dpg, dpo))
        if len(dlo):
            assert rel(dlg, dlo) < tol, (i, rel(dlg, dlo))
    assert abs(sg.post_solve_norm - so.post_solve_norm) <= cost_tol * so.post_solve_norm
    pg, po = g.get_poses(), o.get_poses()
    assert rel(pg["t_wp"], po["t_wp"]) < max(1e-9, cost_tol)
    assert np.array_equal(pg["mask_used"], po["mask_used"]) and np.array_equal(pg["mask"], po["mask"])
    if p.lm_dim:
        lg, lo = g.get_landmarks(), o.get_landmarks()
        assert rel(lg["x_w"], lo["x_w"]) < max(1e-9, cost_tol)
        assert np.array_equal(lg["num_outlier_residuals"], lo["num_outlier_residuals"])
        assert np.array_equal(lg["is_reliable"], lo["is_reliable"])
        rg, ro = g.get_proj_residuals(), o.get_proj_residuals()
        # stored residuals are those of the final state: they agree as closely as the states do
        assert rel(rg["residual"], ro["residual"]) < max(1e-8, tol) and rel(rg["weight"], ro["weight"]) < max(1e-9, cost_tol)
    return g, o


def check_build(p, **opts):
    g, o = gpu(p, **opts), oracle(p, **opts)
    g.build(); o.build()
    a, b = g.get_system(), o.get_system()
    assert rel(np.triu(a["S"]), np.triu(b["S"])) < 1e-12
    for k in ("rhs_sc", "rhs_p"):
        assert rel(a[k], b[k]) < 1e-11, k
    if p.lm_dim:
        assert rel(a["rhs_l"], b["rhs_l"]) < 1e-11 and rel(a["v_inv"], b["v_inv"]) < 1e-9
    # block structure: bit-exact
    for w in (PATTERN_U, PATTERN_S) + ((PATTERN_W,) if p.lm_dim else ()):
        cg, rg = g.get_pattern(w); co, ro = o.get_pattern(w)
        assert np.array_equal(cg, co) and np.array_equal(rg, ro)


# ---- the reference's CPU-runnable configuration (BASELINE configs[0]) -----------------------------
def test_c1_monocular_full_size():
    p = synth.make_c1()          # 50 poses, 5k inverse-depth landmarks, 100k projection residuals, 10 iterations
    check_build(p)
    check_solve(p, 10, 1e-7)     # the last iterations' deltas are ~1e-6 of the first; 6e-9 measured


@pytest.mark.parametrize("name,make,iters,tol", [
    ("c1", lambda: synth.make_c1(n_poses=12, n_landmarks=200, obs_per_lm=8), 6, 1e-9),
    ("c1_tvs", lambda: synth.make_c1(n_poses=10, n_landmarks=150, obs_per_lm=6, t_vs=[0.01, -0.02, 0.03, 0.9993, 0.1, -0.05, 0.02]), 6, 1e-9),
    ("c1_outliers", lambda: synth.make_c1(n_poses=10, n_landmarks=150, obs_per_lm=6, outlier_frac=0.1), 6, 1e-9),
    ("c2", lambda: synth.make_c2(n_cams=100, n_points=3000), 5, 1e-8),    # cond(S) = 7.4e8 (numpy): cond * 2^-53 = 8e-8
    ("c4", lambda: synth.make_c4(n_cams=60, n_points=2000), 5, 1e-8),     # cond(S) ~ 6e8
    ("c4_l3", lambda: synth.make_c4(n_cams=60, n_points=2000, lm_dim=3), 5, 1e-8),
    ("c5", lambda: synth.make_c5(n_poses=200), 3, 1e-8),
])
def test_small_configs(name, make, iters, tol):
    p = make()
    check_build(p)
    check_solve(p, iters, tol)


def test_c3_visual_inertial():
    # default IMU sigmas make cond(S) ~ 1e13: the build must agree to rounding, the step only to cond * eps
    p = synth.make_c3(n_keyframes=40, n_landmarks=300, obs_per_lm=10)
    check_build(p)
    # with sensor noise comparable to the visual information cond(S) = 2.2e9 (numpy, on the oracle's S), so two
    # correct solvers agree to cond * 2^-53 = 2.5e-7 on the first step (1.5e-7 measured on the Gauss-Newton norm) ...
    sig = dict(gyro_sigma=2e-2, accel_sigma=2e-1, gyro_bias_sigma=1e-2, accel_bias_sigma=1e-1)
    check_solve(p, 1, 1e-6, cost_tol=1e-7, **sig)     # a 1e-6 step difference moves the cost by ~3e-8
    # ... and later steps only to cond * (state difference): the cost, which is insensitive to the weak directions,
    # and the refined state still agree closely after 4 iterations
    g, o = gpu(p, **sig), oracle(p, **sig)
    sg, so = g.solve(4), o.solve(4)
    assert sg.iterations == so.iterations == 4
    for a, b in zip(g.get_trace(), o.get_trace()):
        assert a["inner_iterations"] == b["inner_iterations"]
        assert abs(a["post_solve_norm"] - b["post_solve_norm"]) <= 1e-6 * b["post_solve_norm"]
    assert rel(g.get_poses()["t_wp"], o.get_poses()["t_wp"]) < 1e-6
    g, o = gpu(p), oracle(p)
    sg, so = g.solve(5), o.solve(5)
    assert abs(sg.post_solve_norm - so.post_solve_norm) < 1e-3 * so.post_solve_norm


def test_gauss_newton_mode_and_error_increase():
    p = synth.make_c1(n_poses=10, n_landmarks=150, obs_per_lm=6)
    check_solve(p, 4, 1e-9, use_dogleg=0)
    g, o = gpu(p, use_dogleg=0), oracle(p, use_dogleg=0)
    sg, so = g.solve(3, 25.0), o.solve(3, 25.0)     # an absurd damping makes the error increase -> rollback
    assert sg.result == so.result == 1              # ErrorIncreased
    assert rel(g.get_poses()["t_wp"], o.get_poses()["t_wp"]) < 1e-12


def test_default_exit_thresholds():
    p = synth.make_c1(n_poses=10, n_landmarks=150, obs_per_lm=6)
    from ba_b200.capi import BundleAdjusterGPU
    from oracle.oracle_py import Oracle
    g = BundleAdjusterGPU(1, 6); g.load(p)
    o = Oracle(1, 6); o.load(p)
    sg, so = g.solve(50), o.solve(50)
    assert sg.result == so.result and sg.iterations == so.iterations and sg.result in (2, 3)


def test_edge_cases_inactive_and_empty():
    # inactive landmarks, a landmark without residuals, a user-regularised pose, all poses active (auto gauge)
    p = synth.make_c1(n_poses=10, n_landmarks=120, obs_per_lm=5, n_inactive_poses=0)
    a = p.arrays()
    act = a["lm_active"].copy(); act[::7] = 0
    xw = np.concatenate([a["lm_x_w"], [[0.3, 0.2, 0.1, 1.0]]]); rp = np.concatenate([a["lm_ref_pose"], [3]])
    p.set_bulk(lm_x_w=xw, lm_ref_pose=rp, lm_ref_cam=np.zeros(len(xw)), lm_active=np.concatenate([act, [1]]))
    # RegularizePose(1, translation) expressed on the bulk arrays (reference BundleAdjuster.h:608-631)
    mu = np.zeros(10, np.uint8); mu[1] = 1
    mk = np.ones((10, 6), np.uint8); mk[1, :3] = 0
    p.set_bulk(pose_mask_used=mu, pose_mask=mk)
    check_build(p)
    check_solve(p, 4, 1e-8)
    # no residuals at all: Solve returns immediately (cpp:282-285)
    e = Problem(1, 6); e.AddCamera(synth.CAM); e.AddPose([0, 0, 0, 1, 0, 0, 0])
    assert gpu(e).solve(3).iterations == 0


def test_second_solve_continues_from_the_first():
    p = synth.make_c1(n_poses=10, n_landmarks=150, obs_per_lm=6)
    g, o = gpu(p), oracle(p)
    g.solve(2); o.solve(2)
    sg, so = g.solve(2), o.solve(2)              # trust region and state persist (quirk Q5)
    assert abs(sg.post_solve_norm - so.post_solve_norm) <= 1e-9 * so.post_solve_norm
    assert rel(g.get_poses()["t_wp"], o.get_poses()["t_wp"]) < 1e-9


@pytest.mark.parametrize("name", ["c1_tiny", "c2_tiny", "c4_tiny", "c5_tiny"])
def test_against_golden_fixtures(name):
    from tests.golden.make_golden import CASES, run_case
    from ba_b200.capi import BundleAdjusterGPU
    g = np.load(os.path.join(HERE, name + ".npz"))
    make, iters = CASES[name]
    out = run_case(make, iters, BundleAdjusterGPU)
    keys = list(g["trace_keys"])
    for k in ("post_solve_norm", "pre_solve_norm", "gn_norm", "median_maha", "delta_norm"):
        c = keys.index(k)
        assert np.allclose(out["trace"][:, c], g["trace"][:, c], rtol=1e-8), k
    n0 = np.linalg.norm(g["delta_p_0"])
    for i in range(len(g["trace"])):   # late steps are ~1e-9 of the first: compare on the scale of the first step
        d = np.linalg.norm(out["delta_p_%d" % i] - g["delta_p_%d" % i])
        assert d < 1e-7 * max(np.linalg.norm(g["delta_p_%d" % i]), 1e-4 * n0), (i, d)
    assert rel(out["out_t_wp"], g["out_t_wp"]) < 1e-9
    assert abs(out["summary"][0] - g["summary"][0]) <= 1e-9 * g["summary"][0]


# ---- size-independent properties at sizes the oracle cannot reach quickly --------------------------
def test_noise_free_round_trip_recovers_ground_truth():
    # "encode -> perturb -> solve" round trip: exact observations, perturbed state, the solver must
    # return to the generating state and to zero cost
    p = synth.make_c4(n_cams=400, n_points=60000, lm_dim=3)
    a = p.arrays()
    z, _ = synth.project(p.truth["T"][a["proj_meas_pose"]], p.truth["X"][a["proj_lm"]])
    p.set_bulk(proj_z=z)
    g = gpu(p)
    s = g.solve(15)
    assert s.post_solve_norm < 1e-10
    T = g.get_poses()["t_wp"]
    assert np.abs(T[:, 4:] - p.truth["T"][:, 4:]).max() < 1e-6
    X = g.get_landmarks()["x_w"]
    assert np.abs(X[:, :3] - p.truth["X"]).max() < 1e-5


def test_gn_step_solves_the_assembled_system_mid_size():
    # linear consistency at 0.5M residuals: S delta_p = rhs_sc for the recorded Gauss-Newton step
    p = synth.make("c2", 0.1)
    g = gpu(p, use_dogleg=0)
    g.build()
    sysm = g.get_system()
    S = np.triu(sysm["S"]); S = S + np.triu(S, 1).T
    g2 = gpu(p, use_dogleg=0)
    g2.solve(1)
    dp, dl = g2.get_deltas(0)
    r = S @ dp - sysm["rhs_sc"]
    assert np.linalg.norm(r) < 1e-9 * np.linalg.norm(sysm["rhs_sc"])


def test_gn_step_solves_the_assembled_system_pose_graph():
    # 1000-pose graph with loop closures: 125 tile rows, nested-dissection order and symbolic fill of the tile
    # structure are exercised (sorted column lists merged into the elimination-tree parent); the recorded
    # Gauss-Newton step must solve the assembled reduced system
    p = synth.make("c5", 0.05)
    g = gpu(p, use_dogleg=0)
    g.build()
    sysm = g.get_system()
    S = np.triu(sysm["S"]); S = S + np.triu(S, 1).T
    g2 = gpu(p, use_dogleg=0)
    g2.solve(1)
    dp, _ = g2.get_deltas(0)
    r = S @ dp - sysm["rhs_sc"]
    assert np.linalg.norm(r) < 1e-8 * np.linalg.norm(sysm["rhs_sc"])


def test_cost_decreases_monotonically_c2_scaled():
    p = synth.make("c2", 0.2)
    g = gpu(p)
    g.solve(5)
    tr = g.get_trace()
    for t in tr:
        assert t["post_solve_norm"] <= t["pre_solve_norm"]
    assert tr[-1]["post_solve_norm"] < 0.5 * tr[0]["pre_solve_norm"]


def test_conditioning_summary_matches_oracle():
    # inactive reference poses with active measurement poses: conditioning residuals (reference BundleAdjuster.h:503-510),
    # counted at set-up and their maha / weight summed on the device (src/BundleAdjuster.cpp:680-704)
    p = synth.make_c1(n_poses=12, n_landmarks=200, obs_per_lm=6, n_inactive_poses=3)
    g, o = gpu(p), oracle(p)
    sg, so = g.solve(3), o.solve(3)
    assert so.num_cond_proj_residuals > 0
    assert sg.num_cond_proj_residuals == so.num_cond_proj_residuals
    assert sg.num_proj_residuals == so.num_proj_residuals
    assert abs(sg.cond_proj_error - so.cond_proj_error) <= 1e-9 * so.cond_proj_error


def test_page_locked_host_buffers_give_identical_results():
    # the same arrays handed over from bagpu_host_alloc buffers (direct DMA) and from ordinary memory (staged copy)
    from ba_b200.capi import pinned_like, release_pinned
    p = synth.make_c4(n_cams=40, n_points=3000, seed=7)
    a = p.arrays()
    pp = Problem(p.lm_dim, p.pose_dim)
    pp.set_bulk(**{k: pinned_like(v) for k, v in a.items()})
    g1, g2 = gpu(p), gpu(pp)
    s1, s2 = g1.solve(3), g2.solve(3)
    # same bytes in, same kernels: only the order of the atomic error sums differs from run to run
    assert abs(s1.post_solve_norm - s2.post_solve_norm) <= 1e-12 * s1.post_solve_norm
    assert rel(g1.get_poses()["t_wp"], g2.get_poses()["t_wp"]) < 1e-12
    assert rel(g1.get_landmarks()["x_w"], g2.get_landmarks()["x_w"]) < 1e-12
    g2.close()
    release_pinned()


def test_patterns_after_device_upload_and_shuffled_residual_order():
    # residual ids in random order: the device's stable landmark-major sort must reproduce the reference's per-landmark
    # ascending residual lists, and the host pattern query works from ids fetched back from the device
    p = synth.make_c1(n_poses=9, n_landmarks=80, obs_per_lm=5)
    a = p.arrays()
    rng = np.random.default_rng(5)
    perm = rng.permutation(len(a["proj_z"]))
    p.set_bulk(proj_z=a["proj_z"][perm], proj_meas_pose=a["proj_meas_pose"][perm], proj_lm=a["proj_lm"][perm],
               proj_cam=a["proj_cam"][perm], proj_weight=a["proj_weight"][perm])
    check_build(p)
    check_solve(p, 3, 1e-8)


def test_invalid_ids_are_rejected_at_upload():
    from ba_b200.capi import BundleAdjusterGPU
    p = synth.make_c1(n_poses=6, n_landmarks=30, obs_per_lm=4)
    a = p.arrays()
    bad = a["proj_lm"].copy(); bad[3] = 10 ** 6
    p.set_bulk(proj_lm=bad)
    s = BundleAdjusterGPU(p.lm_dim, p.pose_dim)
    with pytest.raises(RuntimeError, match="out of range"):
        s.load(p)


def test_landmarks_stay_on_device_until_asked_for():
    """Solve() leaves the refined landmarks on the device; get_landmarks reads them straight into the caller's
    (optionally preallocated) buffers, sync_state refreshes the host copies, and a problem that is set up again
    without new landmarks (only the residuals are re-sent) continues from the refined ones."""
    import ctypes as C
    from ba_b200.abi import _p
    p = synth.make("c1", 0.3)
    a = p.arrays()
    g = gpu(p)
    g.solve(3)
    lm1 = g.get_landmarks()
    n = len(a["lm_x_w"])
    out = dict(x_w=np.full((n, 4), np.nan), x_s=np.full((n, 4), np.nan), is_reliable=np.full(n, 7, np.uint8),
               num_outlier_residuals=np.full(n, 9, np.uint32))
    lm2 = g.get_landmarks(out=out)
    assert lm2["x_w"] is out["x_w"]
    for k in lm1:
        assert np.array_equal(lm1[k], lm2[k]), k
    assert not np.array_equal(lm1["x_w"], a["lm_x_w"])          # the solve moved them
    g.sync_state()
    lm3 = g.get_landmarks()
    for k in lm1:
        assert np.array_equal(lm1[k], lm3[k]), k
    # reference run: 3 + 2 iterations in two calls on one handle without touching the problem in between
    r = gpu(p); r.solve(3); r.solve(2)
    want = r.get_landmarks()
    # same, but the projection residuals are sent again between the calls: full set-up from the host vectors, whose
    # landmark part has to be fetched from the device first
    h = gpu(p); h.solve(3)
    h._check(h._f("set_proj_residuals")(h.h, C.c_uint32(len(a["proj_z"])), _p(a["proj_z"], np.float64), _p(a["proj_meas_pose"], np.uint32),
                                        _p(a["proj_lm"], np.uint32), _p(a["proj_cam"], np.uint32), _p(a["proj_weight"], np.float64)))
    h.solve(2)
    got = h.get_landmarks()
    np.testing.assert_allclose(got["x_w"], want["x_w"], rtol=0, atol=1e-9)
    np.testing.assert_allclose(h.get_poses()["t_wp"], r.get_poses()["t_wp"], rtol=0, atol=1e-9)


# ---- block-Jacobi PCG (options.reduced_solver = 1; north_star: "... and with a block-Jacobi PCG otherwise") -------------
@pytest.mark.parametrize("name,mk", [
    ("c1", lambda: synth.make_c1(n_poses=12, n_landmarks=300, obs_per_lm=8)),
    ("c2", lambda: synth.make("c2", 0.02)),
    ("c5", lambda: synth.make_c5(n_poses=120)),
])
def test_pcg_matches_tile_cholesky(name, mk):
    """Same system, two solvers: the Gauss-Newton step of the PCG (relative residual 1e-14) against the LDL^T's, and the
    residual of both in the conditioning-free form |S dp - rhs| / |rhs| from the dumped system."""
    p = mk()
    a, b = gpu(p, reduced_solver=0), gpu(p, reduced_solver=1, pcg_tolerance=1e-14, pcg_max_iterations=4000)
    sa, sb = a.solve(1), b.solve(1)
    assert sa.result == sb.result and sa.iterations == sb.iterations == 1
    ta, tb = a.get_trace()[0], b.get_trace()[0]
    assert abs(ta["gn_norm"] - tb["gn_norm"]) <= 1e-9 * ta["gn_norm"]
    assert abs(ta["post_solve_norm"] - tb["post_solve_norm"]) <= 1e-9 * ta["post_solve_norm"]
    dpa, _ = a.get_deltas(0); dpb, _ = b.get_deltas(0)
    assert rel(dpb, dpa) < 1e-8, rel(dpb, dpa)


def test_pcg_full_solve_matches_oracle():
    p = synth.make_c1(n_poses=12, n_landmarks=300, obs_per_lm=8)
    check_solve(p, 4, 1e-8, cost_tol=1e-9, reduced_solver=1, pcg_tolerance=1e-14, pcg_max_iterations=4000)
    # reduced_solver = 2: PCG below 65 poses, the tile LDL^T above — both within the same bar of the oracle
    check_solve(p, 3, 1e-8, cost_tol=1e-9, reduced_solver=2, pcg_tolerance=1e-14, pcg_max_iterations=4000)


def test_pcg_out_of_iterations_is_a_solver_error():
    p = synth.make("c2", 0.02)
    g = gpu(p, reduced_solver=1, pcg_tolerance=1e-14, pcg_max_iterations=3)
    s = g.solve(2)
    assert s.result == 5 and s.iterations == 1      # SolverError, Solve() stops like the reference (cpp:639-644)


# ---- bit reproducibility (the reference is a serial, deterministic program) -----------------------------------------------
@pytest.mark.parametrize("name,scale", [("c2", 0.2), ("c4", 0.05), ("c1", 1.0)])
def test_two_runs_agree_bit_for_bit(name, scale):
    p = synth.make(name, scale)
    out = []
    for _ in range(2):
        g = gpu(p)
        g.solve(4)
        out.append((g.get_trace(), g.get_poses()["t_wp"].copy(), g.get_landmarks()["x_w"].copy()))
        g.close()
    (t0, p0, l0), (t1, p1, l1) = out
    for a, b in zip(t0, t1):
        for k in TRACE_KEYS:
            assert a[k] == b[k], (k, a[k], b[k])
    assert np.array_equal(p0, p1) and np.array_equal(l0, l1)


# ---- conditioning-free comparison of the Gauss-Newton step -----------------------------------------------------------------
@pytest.mark.parametrize("name,make", [
    ("c2", lambda: synth.make_c2(n_cams=100, n_points=3000)),
    ("c4", lambda: synth.make_c4(n_cams=60, n_points=2000)),
    ("c5", lambda: synth.make_c5(n_poses=200)),
])
def test_gauss_newton_step_in_the_residual_norm(name, make):
    """|S (dp_gpu - dp_oracle)| / |rhs| <= 1e-12 on the configurations whose step parity bar is 1e-8: measured in the norm the
    solve controls, the two steps are the same to rounding; their 1e-9 .. 1e-8 distance in the Euclidean norm is the
    conditioning of S (tests/test_sensitivity_cpu.py quantifies it on the oracle's own system)."""
    p = make()
    g, o = gpu(p, use_dogleg=0), oracle(p, use_dogleg=0)
    g.solve(1); o.solve(1)
    dpg, _ = g.get_deltas(0); dpo, _ = o.get_deltas(0)
    o2 = oracle(p, use_dogleg=0); o2.build()
    sys_ = o2.get_system()
    U = np.triu(sys_["S"]); S = U + np.triu(U, 1).T
    assert np.linalg.norm(S @ (dpg - dpo)) <= 1e-12 * np.linalg.norm(sys_["rhs_sc"])
    assert np.linalg.norm(S @ dpg - sys_["rhs_sc"]) <= 1e-11 * np.linalg.norm(sys_["rhs_sc"])


def test_against_golden_c3_tiny():
    """The visual-inertial fixture at the reference's default IMU sigmas: cond(S) ~ 1e13, so only what does not pass through
    the reduced solve is held to 1e-9 (the first build: cost, median, inertial error); the costs after a step loosely
    (test_c3_visual_inertial has the well-conditioned variant at 1e-6)."""
    from tests.golden.make_golden import CASES, run_case
    from ba_b200.capi import BundleAdjusterGPU
    g = np.load(os.path.join(HERE, "c3_tiny.npz"))
    make, iters = CASES["c3_tiny"]
    out = run_case(make, iters, BundleAdjusterGPU)
    keys = list(g["trace_keys"])
    for k in ("pre_solve_norm", "median_maha", "proj_error_build", "inertial_error_build", "sd_norm"):
        c = keys.index(k)
        assert abs(out["trace"][0, c] - g["trace"][0, c]) <= 1e-9 * abs(g["trace"][0, c]), k
    # cond(S) = 8.9e12 on this fixture: two float64 solves of the ORACLE's own first system differ by 1.6e-5 from each other
    # and by 1e-4 from an extended-precision solve (tests/test_sensitivity_cpu.py::test_c3_tiny...), which moves the cost
    # after one step by ~1e-3 and compounds over the four iterations
    c = keys.index("post_solve_norm")
    assert abs(out["trace"][0, c] - g["trace"][0, c]) <= 5e-3 * g["trace"][0, c]
    assert np.allclose(out["trace"][:, c], g["trace"][:, c], rtol=0.1)
