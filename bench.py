#!/usr/bin/env python
"""bench.py — headline benchmark of the hot path (BASELINE.json: "LM iterations/sec and Schur-build
ms/iter at 1/2/4/8 B200 vs host CPU").

A step is one outer iteration of ba::BundleAdjuster::Solve (reference src/BundleAdjuster.cpp:298-663):
BuildProblem + Schur complement + reduced camera solve + dogleg step with its residual evaluations.
Headline workload (N = 1 and every N of the scaling run): BASELINE configs[3], the 10M-observation problem
north_star quotes its roofline and scaling targets on (10k cameras, 2M landmarks, <double,1,6,0>); BASELINE
configs[1] (BAL-shaped <double,3,6,0>, 1,778 cameras, 993k points, 5.0M observations) is measured in the same
run and reported as the keyed block `also.c2` of the same JSON line (`--workload` / `--also` choose).  With
N > 1 (torchrun, one process per GPU) the landmarks are sharded across ranks and the partial reduced systems
are all-reduced: strong scaling.

  value        whole-job LM iterations/s, inputs resident in HBM, device timeline (CUDA events on the
               solver's stream), max over ranks
  e2e          the same metric through the public API with host buffers: upload of the whole problem,
               K iterations, download of the refined poses/landmarks, all inside the timed region
  roofline     dominant kernel of the Schur build vs the measured HBM copy bandwidth
  cpu_baseline the CPU oracle (port of the reference loop, single-threaded like the reference's
               projection path): 2 iterations of the FULL workload, N = 1 only
  parity       first two iterations of the end-to-end GPU trace against that oracle run (same problem, same start)
`--impl reference` times that CPU implementation alone on the full workload (the reference itself cannot be
built here).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    "c1": dict(desc="C1 monocular <double,1,6,0>: 50 poses, 5k inverse-depth landmarks, 100k residuals"),
    "c2": dict(desc="C2 BAL-shaped <double,3,6,0>: 1778 cameras, 993k points, 5.0M observations"),
    "c3": dict(desc="C3 visual-inertial <double,1,15,0>: 500 keyframes, 50k landmarks, 2M projections + 499 IMU"),
    "c4": dict(desc="C4 large <double,1,6,0>: 10k cameras (every 64th fixed), 2M landmarks, 10M observations"),
    "c5": dict(desc="C5 pose graph <double,0,6,0>: 20k poses, binary + unary residuals"),
}


def algorithmic_bytes(a, L, d, nnzb):
    """compulsory traffic of one fused, matrix-free build (J never stored), split over the
    kernels that carry it.  W blocks (6 x L doubles per pose-landmark pair) are written once by k_linearize and
    read once by k_reduce_blocks; the records the kernels really move are larger (12 / 16 doubles per observation),
    which shows up as roofline.traffic > algorithmic bytes, not in `achieved`."""
    n_o = len(a["proj_z"]); n_l = len(a["lm_x_w"]); P = int(a["pose_active"].sum())
    n_pl = n_o if L == 3 else n_o + n_l
    per_kernel = {
        "k_residual": n_o * (40 + 8) + n_l * 40,
        "select": n_o * 8,
        "k_linearize": n_o * (40 + 8) + n_pl * 48 * L + n_l * (40 + 8 * (L * L + 2 * L)),
        "k_reduce_blocks": n_pl * 48 * L + nnzb * 8 * d * d + P * (56 + 16 * d),
    }
    total = n_o * 96 + n_l * (40 + 8 * (L * L + 2 * L)) + n_pl * 96 * L + nnzb * 8 * d * d + P * (56 + 16 * d)
    return total, per_kernel


def make_workload(name, scale):
    from ba_b200 import synth
    return synth.make(name, scale)


def options_for(solver, p, keep_on_device):
    o = solver.default_options()
    o.error_change_threshold = 0          # run exactly K iterations (quirk Q9: the defaults stop early)
    o.param_change_threshold = 0
    # `value` times the iterations alone: no end-of-Solve() hand-over of the state.  (The timed Solve(K) then starts
    # from the warm-up's refined poses and the original landmark positions — a perturbed state like the initial one;
    # every iteration does the full work.  The e2e leg runs without this flag.)
    o.keep_on_device = 1 if keep_on_device else 0
    if hasattr(p, "gravity"):
        o.gravity[0], o.gravity[1], o.gravity[2] = [float(x) for x in p.gravity]
    return o


class ClockSampler:
    Q = "index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, gpu_index=0):
        self.samples, self.stop, self.idx = [], False, gpu_index
        self.t = threading.Thread(target=self.run, daemon=True)

    def run(self):
        while not self.stop:
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.idx), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.samples.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            time.sleep(0.1)

    def __enter__(self):
        self.t.start(); return self

    def __exit__(self, *a):
        self.stop = True; self.t.join(2)

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        sm = [float(s[1]) for s in self.samples]
        reasons = set()
        for s in self.samples:
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), s[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(self.samples[0][2]), "reasons": sorted(reasons),
                "power_w_max": max(float(s[3]) for s in self.samples), "samples": len(self.samples)}


def measured_peak():
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def cpu_reference_run(name, steps, warmup, budget_s=420.0, scale=1.0):
    """The CPU implementation of the path (oracle port; projection, sparse algebra and solve on 1 thread like the
    reference, cpp:1345-1347) on the FULL workload (`scale` 1.0).  Only the number of iterations is bounded: when
    `steps` + `warmup` iterations would not fit `budget_s`, fewer are run and the sample says so."""
    from oracle.oracle_py import Oracle
    per_iter_full = {"c1": 0.09, "c2": 7.0, "c3": 5.0, "c4": 19.0, "c5": 0.3}[name] * scale   # rough, only to bound the iterations
    warm = min(warmup, 1)
    steps_run = max(1, min(steps, int(budget_s / per_iter_full) - warm))
    p = make_workload(name, scale)
    s = Oracle(p.lm_dim, p.pose_dim)
    s.set_options(options_for(s, p, False))
    s.load(p)
    trace = []
    if warm:
        s.solve(warm); trace += s.get_trace()
    t0 = time.time()
    s.solve(steps_run)
    dt = time.time() - t0
    trace += s.get_trace()
    ph, _ = s.get_phase_ms()
    n_res = max(1, p.counts["proj"] + p.counts["binary"] + p.counts["unary"] + p.counts["imu"])
    build_ms = (ph["residual"] + ph["median"] + ph["accumulate"] + ph["other_residuals"] + ph["schur"]) / steps_run
    return dict(value=steps_run / dt, build_ms=build_ms, ms_per_step=dt / steps_run * 1e3, scale=scale, trace=trace,
                sample="%d timed iteration(s) (%d asked) after %d warm-up on %.4g of the workload (%d residuals): full size, nothing extrapolated"
                       % (steps_run, steps, warm, scale, n_res),
                # projection / sparse algebra / solve are serial as in the reference (cpp:1345-1347); only the IMU functor is
                # spread over threads, like its tbb::parallel_reduce (cpp:1494) — oracle InertialResiduals, same split rule
                cores=max(1, min(os.cpu_count() or 1, p.counts["imu"] // 8)))


def trace_parity(gpu_trace, cpu_trace, n=2):
    """Largest relative difference between the first `n` outer iterations of the GPU trace (end-to-end Solve() from the
    initial state) and the CPU oracle's on the same full-size problem."""
    keys = ("pre_solve_norm", "post_solve_norm", "median_maha", "proj_error_build", "delta_norm", "gn_norm", "sd_norm")
    per_key = {}
    m = min(n, len(gpu_trace), len(cpu_trace))
    for i in range(m):
        for k in keys:
            a, b = gpu_trace[i][k], cpu_trace[i][k]
            per_key[k] = max(per_key.get(k, 0.0), abs(a - b) / max(abs(b), 1e-300))
    cost_keys = ("pre_solve_norm", "post_solve_norm", "median_maha", "proj_error_build")
    return {"iterations_compared": m, "max_rel_diff_costs": max([per_key[k] for k in cost_keys] or [0.0]),
            "max_rel_diff_steps": max([per_key[k] for k in keys if k not in cost_keys] or [0.0]), "per_key": per_key,
            "note": "costs, median and build errors are conditioning-free; the step norms inherit cond(S) * 2^-53 from the reduced solve (a long camera chain with two fixed cameras is gauge-weak)",
            "against": "CPU oracle (port of the reference loop, parity unpinned) on the same full-size problem, same initial state"}


def l2_note(name, scale):
    """Timing rule: inputs larger than L2, or say that they are not.  Per iteration the build streams the landmark-sorted
    observations (40 B each), writes and re-reads a 96-byte linearisation per observation (G and h records; 128 bytes with
    the gv record of the inverse-depth configurations) and one 128-byte record per landmark."""
    n = make_counts(name) * scale
    if name == "c5":
        n = 0
    mb = n * (40 + 8 + 96 + (32 if name in ("c1", "c3", "c4") else 0)) / 1e6
    if mb > 126:
        return "inputs larger than L2: every iteration streams %.0f MB of sorted observations and linearisation records (L2 = 126 MB)" % mb
    return ("working set of the build (%.0f MB) fits in L2, no flush between iterations: this configuration is "
            "launch/latency-bound (SURVEY 8d), its number is not an HBM figure" % mb)


def make_counts(name):
    return {"c1": 100000, "c2": 5000000, "c3": 2000499, "c4": 10000000, "c5": 21200}[name]


def e2e_cpp_run(p, K, tmp_dir):
    """The API north_star names: the C++ drop-in class (include/ba/BundleAdjuster.h) driven through AddPose / AddLandmark /
    AddProjectionResidual / ... / Solve(K) / getters by tests/cpp/host_api_driver.cpp.  The problem is handed to the driver as
    a file (read before the clock starts); timed: the Add* calls (host-side SoA marshalling into page-locked buffers),
    Solve() (upload + set-up + K iterations + read-back of poses and landmarks), the getters of the refined state."""
    from ba_b200.capi import build_cpp_driver
    from ba_b200.problem import write_driver_problem
    exe = build_cpp_driver()
    prob, outp = os.path.join(tmp_dir, "bench_problem.bin"), os.path.join(tmp_dir, "bench_out.bin")
    write_driver_problem(prob, p)
    best = None
    for _rep in range(2):          # the first run pays the CUDA context and module load
        env = dict(os.environ, BA_DRIVER_NO_RESIDUALS="1")
        r = subprocess.run([exe, prob, outp, str(K)], capture_output=True, text=True, env=env)
        if r.returncode != 0:
            return {"error": "driver exit %d: %s" % (r.returncode, r.stderr[-300:])}
        t = None
        for line in r.stderr.splitlines():
            if line.startswith('{"driver_timing"'):
                t = json.loads(line)["driver_timing"]
        if t is None:
            return {"error": "no timing line"}
        if best is None or t["solve_s"] < best["solve_s"]:
            best = t
    for f in (prob, outp):
        try:
            os.remove(f)
        except OSError:
            pass
    its = best["iterations"]
    return {"value": its / (best["add_s"] + best["solve_s"] + best["get_state_s"]), "unit": "iter/s", "iterations": its,
            "add_calls_s": best["add_s"], "solve_s": best["solve_s"], "getters_s": best["get_state_s"],
            "note": "one process, best of 2 (the first pays the CUDA context): Add* calls + Solve(K) + GetPose / GetLandmark over the whole state, wall clock inside the C++ driver"}


def config_for(name, scale, gpus):
    return {"workload": WORKLOADS[name]["desc"] + (" (scale %.3g)" % scale if scale != 1.0 else ""),
            "solver": "dogleg, nested-dissection tile LDL^T (dataflow kernel)", "sharding": "landmarks over %d GPU(s)" % gpus,
            "l2": l2_note(name, scale)}


def run_gpu_workload(name, args, K, W, rank, world, local_rank, with_e2e, with_cpu):
    """One workload through the CUDA path: device-resident iterations (value), end-to-end Solve() (e2e), roofline of the
    build, and (N = 1) the CPU oracle on the same full-size problem with a trace comparison."""
    import torch
    import torch.distributed as dist
    from ba_b200.capi import BundleAdjusterGPU
    from ba_b200.dist import init_nccl_for
    p = make_workload(name, args.scale)
    a = p.arrays()
    L, d = p.lm_dim, p.pose_dim

    def fresh(keep):
        s = BundleAdjusterGPU(L, d)
        s.set_options(options_for(s, p, keep))
        if world > 1:
            init_nccl_for(s)
        return s

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    # ---- value: K iterations with everything resident in HBM -------------------------------------
    s = fresh(True)
    s.load(p)
    if W:
        s.solve(W)
    barrier()
    with ClockSampler(local_rank) as clk:
        sm = s.solve(K)
        barrier()
    if sm.iterations != K or sm.result in (4, 5):
        raise SystemExit("bench.py: the timed Solve() of %s stopped after %d of %d iterations (result %d): no number" % (name, sm.iterations, K, sm.result))
    ph, launches = s.get_phase_ms()
    t_ms = torch.tensor([ph["total"]], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t_ms, op=dist.ReduceOp.MAX)
    total_ms = float(t_ms.item())
    build_ms = (ph["residual"] + ph["median"] + ph["accumulate"] + ph["other_residuals"] + ph["schur"] + ph["allreduce"]) / K
    trace = s.get_trace()
    peak, peak_src = measured_peak()
    roofline = None
    if L:
        from ba_b200.abi import PATTERN_S
        nnzb = len(s.get_pattern(PATTERN_S)[1])
        tot_b, per_kernel = algorithmic_bytes(a, L, d, nnzb)
        # per-kernel device time: CUDA events recorded by the library on its launch stream around each
        # kernel of the build, averaged over the K timed iterations (one launch of each per iteration)
        kt = {"k_residual": ph["residual"] / K, "select": ph["median"] / K, "k_linearize": ph["accumulate"] / K, "k_reduce_blocks": ph["schur"] / K}
        dom = max(("k_linearize", "k_reduce_blocks", "k_residual"), key=lambda k: kt[k])
        ach = per_kernel[dom] / world / (kt[dom] * 1e-3) / 1e9
        t_build = sum(kt.values())
        traffic = None
        try:
            tj = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))
            if world == 1 and args.scale == 1.0:
                traffic = tj.get(name, {}).get(dom)
        except Exception:
            pass
        roofline = {"bound": "hbm", "kernel": dom, "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak, "traffic": traffic,
                    "algorithmic_bytes_per_launch": per_kernel[dom] / world, "launch_ms": kt[dom], "peak_source": peak_src,
                    "note": "achieved = SURVEY 8d algorithmic bytes of this kernel / its CUDA-event time; the kernel gathers per list entry the G records (64 B) of a pose pair's two observations and 80 B of the landmark record from L2 (LmSize 1, off-diagonal: two 32-byte gv records and 32 B of the landmark record); traffic is the ncu figure of the capture named in profiles/ncu_traffic.json (DESIGN.md 4)",
                    "build": {"algorithmic_bytes": tot_b / world, "ms": t_build, "achieved": tot_b / world / (t_build * 1e-3) / 1e9,
                              "frac": tot_b / world / (t_build * 1e-3) / 1e9 / peak, "kernel_ms": kt, "kernel_bytes": per_kernel, "nnzb_S": nnzb}}
    s.close()

    e2e_its, e2e_times, h2d, d2h, e2e_trace = None, [], 0, 0, None
    if with_e2e:
        # ---- e2e: host buffers in, refined state out, every call ------------------------------------------
        # The caller's arrays live in page-locked host memory (bagpu_host_alloc, what include/ba/BundleAdjuster.h keeps
        # its observation buffers in); everything from handing them to the ABI to reading the refined state is timed.
        from ba_b200.capi import pinned_like, release_pinned
        from ba_b200.problem import Problem
        pp = Problem(L, d)
        pp.set_bulk(**{k: pinned_like(v) for k, v in a.items()})
        if hasattr(p, "gravity"):
            pp.gravity = p.gravity
        s2 = fresh(False)
        s2.load(pp); s2.solve(1); s2.close()                     # warm the allocator / module load
        h2d = sum(v.nbytes for k, v in a.items())
        d2h = a["pose_t_wp"].nbytes + a["pose_v_w"].nbytes + a["pose_b"].nbytes + 2 * a["lm_x_w"].nbytes + len(a["lm_x_w"]) * 5
        n_l = len(a["lm_x_w"])
        lm_out = dict(x_w=pinned_like(np.zeros((n_l, 4))), x_s=pinned_like(np.zeros((n_l, 4))),
                      is_reliable=pinned_like(np.zeros(n_l, np.uint8)), num_outlier_residuals=pinned_like(np.zeros(n_l, np.uint32)))
        e2e_times = []
        for _rep in range(3):                                    # median of three complete end-to-end calls
            s3 = fresh(False)                                    # handle + NCCL communicator: set-up, like the process group
            barrier()
            t0 = time.time()
            s3.load(pp)
            sm3 = s3.solve(K)
            if sm3.iterations != K:
                raise SystemExit("bench.py: the e2e Solve() stopped after %d of %d iterations" % (sm3.iterations, K))
            res = s3.get_poses(); lm = s3.get_landmarks(out=lm_out) if L else None
            e2e_trace = s3.get_trace()
            barrier()
            t_e2e = torch.tensor([time.time() - t0], dtype=torch.float64, device="cuda")
            if world > 1:
                dist.all_reduce(t_e2e, op=dist.ReduceOp.MAX)
            e2e_times.append(float(t_e2e.item()))
            s3.close()
        t_e2e = torch.tensor([sorted(e2e_times)[1]], dtype=torch.float64)
        e2e_its = K / float(t_e2e.item())
        del pp
        release_pinned()

    out = None
    if rank == 0:
        out = {"value": K / (total_ms * 1e-3), "ms_per_step": total_ms / K, "schur_build_ms_per_iter": build_ms,
               "phases_ms_per_iter": {k: v / K for k, v in ph.items() if k != "total"},
               "gpu_launches": int(launches), "clocks": clk.summary(),
               "final_cost": trace[-1]["post_solve_norm"] if trace else None,
               "first_cost": trace[0]["pre_solve_norm"] if trace else None}
        if with_e2e:
            out["e2e"] = {"value": e2e_its, "unit": "iter/s", "h2d_bytes_per_step": h2d / K, "d2h_bytes_per_step": d2h / K,
                          "seconds": e2e_times, "note": "median of 3; one Solve(K) call from page-locked host arrays: upload of the whole problem + symbolic set-up on the device + K iterations + download of the refined poses and landmarks"}
        if roofline:
            out["roofline"] = roofline
        if with_e2e and world == 1 and not args.no_e2e_cpp:
            import tempfile
            with tempfile.TemporaryDirectory() as td:
                out["e2e_cpp"] = e2e_cpp_run(p, K, td)
        if with_cpu:
            # the CPU oracle on the SAME full-size problem: 2 iterations from the same initial state, timed, and its
            # trace compared with the end-to-end GPU run's (full-size parity, not only the minis of tests/)
            r = cpu_reference_run(name, 2, 0, scale=args.scale)
            out["cpu_baseline"] = {"value": r["value"], "unit": "iter/s", "cores": r["cores"], "kind": "port", "sample": r["sample"],
                                   "schur_build_ms_per_iter": r["build_ms"]}
            if e2e_trace:
                out["parity"] = trace_parity(e2e_trace, r["trace"], 2)
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ba_b200")
    ap.add_argument("--workload", default="c4", choices=sorted(WORKLOADS),
                    help="headline workload; default c4 = the 10M-observation problem north_star quotes its targets on")
    ap.add_argument("--also", default="c2", help="second workload reported as a keyed block of the same line ('' = none)")
    ap.add_argument("--scale", type=float, default=1.0, help="shrink the workload (development only; invalidates the number)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e-cpp", action="store_true", help="skip the leg that drives the C++ drop-in class")
    ap.add_argument("--no-e2e", action="store_true", help="development / profiler runs only: skip the end-to-end leg (the line then has no e2e)")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    K, W = max(1, args.steps), max(0, args.warmup)
    name = args.workload
    base = {"metric": "lm_iterations_per_sec", "unit": "iter/s", "n_gpus": args.gpus, "steps": K, "warmup": W,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": config_for(name, args.scale, args.gpus)}

    if args.impl == "reference":
        if rank != 0:
            return
        r = cpu_reference_run(name, K, W, scale=args.scale)
        out = dict(base)
        out.update({"impl": "reference", "value": r["value"], "ms_per_step": r["ms_per_step"], "n_gpus": args.gpus,
                    "schur_build_ms_per_iter": r["build_ms"], "gpu_launches": 0, "scaling": "strong",
                    "cpu_baseline": {"value": r["value"], "unit": "iter/s", "cores": r["cores"], "kind": "port", "sample": r["sample"],
                                     "note": "the reference is unbuildable here (Eigen3/Sophus/Calibu/TBB absent); its projection path is single-threaded (src/BundleAdjuster.cpp:1345-1347); only the IMU functor is threaded, as under its tbb::parallel_reduce (cpp:1494): `cores` = those threads, 1 when there are no IMU residuals"},
                    "e2e": {"value": r["value"], "unit": "iter/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}})
        print(json.dumps(out))
        return

    # stdout carries exactly one line (the JSON): NCCL / the process group write a version banner there, so file
    # descriptor 1 points at stderr until the line is printed
    if os.environ.get("NCCL_DEBUG", "VERSION").upper() == "VERSION":
        os.environ["NCCL_DEBUG"] = "WARN"
    sys.stdout.flush()
    saved_stdout = os.dup(1)
    os.dup2(2, 1)
    import torch
    import torch.distributed as dist
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    with_cpu = world == 1 and not args.no_cpu_baseline
    head = run_gpu_workload(name, args, K, W, rank, world, local_rank, not args.no_e2e, with_cpu)
    second = None
    if args.also and args.also != name:
        second = run_gpu_workload(args.also, args, K, W, rank, world, local_rank, not args.no_e2e, with_cpu)
    if rank == 0:
        out = dict(base)
        out.update(head)
        if second is not None:
            second["config"] = config_for(args.also, args.scale, args.gpus)
            out["also"] = {args.also: second}
        sys.stdout.flush()
        os.dup2(saved_stdout, 1)
        print(json.dumps(out), flush=True)
        os.dup2(2, 1)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
