#!/usr/bin/env python
"""bench.py -- BA observations*iterations / second (BASELINE.json metric) on N B200s.

Default workload (N=1): C4 = 4096 KITTI-05-shaped windows (10 keyframes, 2000 landmarks, ~15000
observations each) x 10 LM iterations per step, one window per CTA, no data-path collective
(independent windows; weak scaling: 4096 windows per GPU).  `--workload c5` runs the sharded
BAL-style problem.  One JSON line on stdout (rank 0).  See DESIGN.md "Measurement".
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

FP64_PEAK_FALLBACK_TFLOPS = 37.1   # round-1 measurement (profiles/r1_fp64_peak.log); only used if the in-bench
                                   # measurement (b2ba_peak_flops) fails.  MEASURED_PEAKS.json carries no fp64 figure.


def fp64_peak(device):
    """FP64 peak of THIS device, measured now (libb2ba.so: DFMA chains and DMMA m8n8k4, CUDA events).  DFMA and DMMA
    share one pipe on B200 (tools/fp64_pipes.cu, profiles/r2_fp64_pipes.log), so the peak is the higher of the two."""
    import ctypes as C
    from visualodometry_b200 import _lib
    a, b = C.c_double(0), C.c_double(0)
    try:
        rc = _lib.load().b2ba_peak_flops(int(device), C.byref(a), C.byref(b))
    except Exception:
        rc = -1
    if rc != 0 or not (a.value > 0 and b.value > 0):
        return {"tflops": FP64_PEAK_FALLBACK_TFLOPS, "kind": "fallback (round-1 measurement)", "dfma": None, "dmma": None}
    return {"tflops": max(a.value, b.value), "kind": "measured in this run (b2ba_peak_flops)", "dfma": a.value, "dmma": b.value}


def algorithmic_flops_per_iter(pt_ptr_counts):
    """a1+a2+a7 = 570 flop/obs, Schur a4 = 108 k + 216 k(k+1)/2 per landmark."""
    k = np.asarray(pt_ptr_counts, np.float64)
    return float(570.0 * k.sum() + (108.0 * k + 108.0 * k * (k + 1)).sum())


def algorithmic_bytes_per_iter(no, npts, nc, nblk):
    """2*16*No + 72*Np + 2*96*Nc + 288*nnzblk(S)."""
    return 2.0 * 16 * no + 72.0 * npts + 2.0 * 96 * nc + 288.0 * nblk


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return json.load(open(p)), "measured"
    return {"hbm_gbs": 6650.0}, "fallback"


class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, dev):
        self.p, self.dev = None, dev

    def start(self):
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(self.dev), f"--query-gpu={self.Q}",
                                       "--format=csv,noheader,nounits", "-lms", "100"],
                                      stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.p = None

    def stop(self):
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.p.terminate()
        out = self.p.communicate()[0]
        sm, mx, reasons = [], [], set()
        for line in out.splitlines():
            f = [x.strip() for x in line.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        busy = [s for s in sm if s > 0.5 * max(sm)] or sm
        return {"sm_mhz": float(np.median(busy)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons),
                "samples": len(sm)}


def make_workload(args, rank):
    from visualodometry_b200 import synth
    if args.workload == "c4":
        b = synth.make_batch(args.windows, seed0=rank * args.windows, distinct=args.distinct)
        name = (f"C4: {args.windows} independent KITTI-05-shaped windows per GPU (10 keyframes, 2000 landmarks, "
                f"~15000 observations), {args.iters} LM iterations per step")
    elif args.workload == "c1":
        b = synth.pack_batch([synth.make_window(rank)])
        name = f"C1: one KITTI-05-shaped window, {args.iters} LM iterations per step"
    elif args.workload == "c3":
        b = synth.pack_batch([synth.make_window(rank, n_cams=20, cam=synth.MALAGA, outlier_frac=0.1, n_obs=30000)])
        name = f"C3: one Malaga-shaped 20-keyframe window, Huber, {args.iters} LM iterations per step"
    elif args.workload == "c5":
        b = synth.pack_batch([synth.make_large(0, mode=args.c5_mode)])     # same problem on every rank
        name = (f"C5: BAL-style, 1000 cameras, 200k landmarks, 1M observations, visibility '{args.c5_mode}', "
                f"{args.iters} LM iterations per step, landmarks sharded across ranks")
    else:
        raise SystemExit(f"unknown workload {args.workload}")
    return b, name


def K_of(intr):
    return np.array([[intr[0], 0, intr[2]], [0, intr[1], intr[3]], [0, 0, 1]])


CPU_VARIANTS = (
    ("strict", dict(fast=False, lin_once=False), "-O2 -ffp-contract=off, every pass re-evaluates its observations (the parity oracle's build)"),
    ("fast", dict(fast=True, lin_once=True), "-O3 -march=native -ffp-contract=fast, linearise once per accepted state (r, w, J stored per observation)"),
)


def cpu_baseline(batch, iters, loss, budget_windows, threads=0, with_scipy=False):
    """Oracle (C++ fp64 port of Spec a0; the reference has no BA) on a bounded sample of the same workload, on the
    GPU box's host cores.  Two builds of the same source are timed (CPU_VARIANTS); `value` is the FASTER one.  A single
    window (C1 / C3) is timed on one core and with OpenMP inside the window on all cores; scipy.optimize.least_squares
    (SURVEY 8(d) comparator 3) is timed beside them when asked."""
    from oracle import ba_cpu
    from visualodometry_b200 import synth
    nw = min(batch.nb, budget_windows)
    sample = synth.pack_batch([batch.window(i) for i in range(nw)])
    cores = ba_cpu.num_procs() if threads <= 0 else threads
    large = sample.nb == 1 and sample.cam_off[-1] > 24       # C5: one large problem, PCG, OpenMP inside the solve
    single = sample.nb == 1 and not large
    ba_cpu.lib(fast=False); ba_cpu.lib(fast=True)            # both builds exist before anything is timed
    res, t_all = {}, time.perf_counter()
    for name, kw, flags in CPU_VARIANTS:
        # best of 3 full passes: the first ones pay for the thread pool, page faults and glibc's adaptive mmap threshold
        # (measured: 937 / 399 / 199 ms for the same call), which are not the solver's cost.  C5 (seconds per pass): one.
        dt, done = float("inf"), iters
        for _ in range(1 if large else 3):
            t0 = time.perf_counter()
            if large:
                _, _, so = ba_cpu.solve(sample.window(0), max_iters=iters, loss=loss, solver="pcg", threads=cores, **kw)
                done = so.iterations
            elif single:
                ba_cpu.solve(sample.window(0), max_iters=iters, loss=loss, threads=cores, **kw)
            else:
                ba_cpu.solve_batch(sample, max_iters=iters, loss=loss, threads=cores, **kw)
            dt = min(dt, time.perf_counter() - t0)
        res[name] = {"value": sample.no * done / dt, "seconds": dt, "cores": cores, "build": flags}
        if single:
            d1 = float("inf")
            for _ in range(3):
                t0 = time.perf_counter()
                ba_cpu.solve(sample.window(0), max_iters=iters, loss=loss, threads=1, **kw)
                d1 = min(d1, time.perf_counter() - t0)
            res[name + "_one_core"] = {"value": sample.no * iters / d1, "seconds": d1, "cores": 1, "build": flags}
    if with_scipy and single and loss == "trivial":
        from oracle import scipy_ba
        sc = scipy_ba.solve(sample.window(0), max_nfev=iters)
        res["scipy_least_squares"] = {"value": sample.no * sc["nfev"] / sc["seconds"], "seconds": sc["seconds"], "cores": "scipy/BLAS default",
                                      "build": f"scipy.optimize.least_squares(trf, sparse 2-point Jacobian), {sc['nfev']} evaluations, cost {sc['cost']:.6g}"}
    best = max((k for k in res if k != "scipy_least_squares"), key=lambda k: res[k]["value"])
    what = ("the whole C5 problem, OpenMP inside the solve" if large else
            "one window, OpenMP inside the window" if single else
            f"first {nw} windows of the workload ({sample.no} observations), OpenMP across windows")
    return {"value": res[best]["value"], "unit": "obs*iter/s", "cores": res[best]["cores"], "kind": "port",
            "sample": f"{what}, {iters} LM iterations, oracle/ba_cpu.cpp [{best}: {res[best]['build']}], {res[best]['seconds']:.2f} s",
            "variants": res, "cpu_model": ba_cpu.cpu_model(), "host_cores": ba_cpu.num_procs(),
            "seconds_total": time.perf_counter() - t_all}


def run_c2(args, impl):
    """BASELINE config 2: end-to-end VO on the synthetic KITTI-05-shaped sequence, frames/sec.  The frame loop is
    visualodometry_b200/vo_loop.py (the reference's loop restated and pinned to it, tests/test_vo_loop.py), the
    frontend is synthetic (id-matched keypoints), BA runs once per keyframe through adapter.install():
    impl "b200" -> BundleAdjuster on the GPU (host buffers in/out every call), "reference" -> the CPU oracle."""
    import cv2
    from visualodometry_b200 import BAConfig, adapter, vo_loop, vo_synth
    seq = vo_synth.SynthSequence(n_frames=args.frames, seed=0)
    cfg = BAConfig(window=10, max_iters=args.iters, loss="huber", loss_scale=2.0)

    def one_run(kind, gpu_tracker=False, gpu_tri=False):
        cv2.setRNGSeed(0)
        lcfg = vo_loop.LoopConfig()
        vo = vo_loop.VOLoop(seq.K, lcfg, vo_synth.SynthFrontend(seq))
        # the GPU arms keep the map in the dense id -> slot table (landmark_table.py); results are
        # bit-identical to the dict scans of the reference (tests/test_landmark_table.py)
        vo.dense_map = kind == "gpu"
        ad, t_ba, n_obs, nbytes = None, [0.0], [0], [0, 0]
        if gpu_tracker:
            # SURVEY 8 row f-2 widened: the tracker's cv2.solvePnPRansac call (vo.py:135-141) replaced by the
            # sampling-free GPU tracker (b2ba_track_pose) started from the previous pose
            from visualodometry_b200 import BundleAdjuster
            trk = BundleAdjuster(seq.K, BAConfig(ftol=1e-10))   # the refinement rounds stop when the cost stops moving
            if gpu_tracker == "ransac":      # b2ba_ransac_pose: 1024 hypotheses per frame, sampled consensus
                vo.pnp_solver = lambda X, uv, T: trk.ransac_pose(T, X, uv, lcfg.pnp_reproj_err, n_hyp=1024, seed=vo.frame_id)[:3]
            else:                            # b2ba_track_pose: graduated robust refinement from the previous pose
                vo.pnp_solver = lambda X, uv, T: trk.track_pose(T, X, uv, lcfg.pnp_reproj_err)[:3]
            if gpu_tri:                      # triangulate_points (frontend.py:115-148) on the GPU
                vo.triangulator = lambda T1, T2, p1, p2, K, c: trk.triangulate(T1, T2, p1, p2, c.min_depth, c.max_reproj_err)
        if kind != "none":
            if kind == "gpu":
                from visualodometry_b200 import BundleAdjuster
                solver = BundleAdjuster(seq.K, cfg)
            else:
                sys.path.insert(0, os.path.join(ROOT, "tests"))
                from oracle_ba import OracleBA
                solver = OracleBA(seq.K, cfg)
            inner = solver.optimize

            def timed(*a):
                t = time.perf_counter()
                r = inner(*a)
                t_ba[0] += time.perf_counter() - t
                n_obs[0] += len(a[2]) * r.summary.iterations
                # bytes the call moves across PCIe: poses (12 f64 per keyframe), landmarks, observation arrays in;
                # refined poses and landmarks + the summary out
                nbytes[0] += 96 * len(a[0]) + 24 * len(a[1]) + 16 * len(a[2]) + len(a[5])
                nbytes[1] += 96 * len(a[0]) + 24 * len(a[1]) + 48
                return r
            solver.optimize = timed
            ad = adapter.install(vo, cfg, ba=solver)
        t0 = time.perf_counter()
        for i in range(seq.n_frames):
            vo.process_frame(seq.image(i))
        dt = time.perf_counter() - t0
        nkf = len([s_ for s_ in ad.summaries if s_ is not None]) if ad else 0
        ate = vo_synth.ate_sim3(np.array(vo.trajectory), seq.T_wc[-len(vo.trajectory):, :3, 3]) if len(vo.trajectory) > 10 else None
        return {"fps": seq.n_frames / dt, "s": dt, "ba_s": t_ba[0], "ba_calls": nkf, "ba_obs_iters": n_obs[0],
                "ba_h2d_bytes": nbytes[0], "ba_d2h_bytes": nbytes[1],
                "pnp_fallbacks_to_ransac": vo.pnp_fallbacks if gpu_tracker else None,
                "frames_since_last_reset": len(vo.trajectory), "ate_sim3_m": ate}

    plain = one_run("none")
    if impl == "reference":
        r = one_run("cpu")
    else:
        one_run("gpu")                       # warm-up (module load, allocations)
        r = one_run("gpu")
        r_trk = one_run("gpu", gpu_tracker="track")
        r_rsc = one_run("gpu", gpu_tracker="ransac")
        r_all = one_run("gpu", gpu_tracker="ransac", gpu_tri=True)
    name = (f"C2: end-to-end VO, synthetic KITTI-05-shaped sequence ({args.frames} frames, 1241x376 K, <=2048 keypoints "
            f"per frame, 5% wrong matches), 10-keyframe window BA ({args.iters} LM iterations, Huber) once per keyframe")
    out = {"metric": "VO frames/sec", "value": r["fps"], "unit": "frames/s", "n_gpus": 1, "steps": 1, "warmup": 1,
           "ms_per_step": 1e3 * r["s"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
           "data": "synthetic", "config": {"workload": name},
           "vo": {"with_ba": r, "without_ba": plain,
                  "ba_ms_per_keyframe": 1e3 * r["ba_s"] / max(1, r["ba_calls"]),
                  "ba_obs_iters_per_s": r["ba_obs_iters"] / r["ba_s"] if r["ba_s"] > 0 else None},
           # a step = the whole sequence; every BA call copies its window in and its result out (host arrays)
           "e2e": {"value": r["fps"], "unit": "frames/s", "h2d_bytes_per_step": int(r["ba_h2d_bytes"]),
                   "d2h_bytes_per_step": int(r["ba_d2h_bytes"])}}
    if impl != "reference":
        # `value` is the SAME-RESULT configuration: the reference's loop (OpenCV RANSAC and triangulation untouched,
        # pinned bit for bit by tests/test_vo_loop.py) with the window BA on the GPU.  The arms below replace further
        # stages of the loop by GPU kernels that are NOT result-identical to the OpenCV calls they stand in for (a
        # sampled consensus with its own random stream): different algorithm, different trajectory -- reported beside
        # the headline, never as it.
        out["vo"]["different_algorithm_arms"] = {
            "note": "PnP consensus / triangulation on the GPU: behaviourally pinned to cv2.solvePnPRansac / triangulate_points "
                    "(tests/test_pose.py, tests/test_triangulate.py), not result-identical; trajectories differ from the pinned loop",
            "gpu_tracker_sampling_free": r_trk, "gpu_ransac": r_rsc, "gpu_ransac_and_gpu_triangulation": r_all}
        out["config"]["gpu_stages_of_the_headline"] = ("window BA (b2ba_optimize) once per keyframe; landmark look-ups through the dense "
                                                       "id -> slot table (same results as the dict scans); everything else as in the reference")
    if impl == "reference":
        from oracle import ba_cpu
        out["impl"] = "reference"
        out["cpu_baseline"] = {"value": r["fps"], "unit": "frames/s", "cores": 1, "kind": "port",
                               "sample": f"whole sequence; BA = oracle/ba_cpu.cpp single thread per window ({ba_cpu.num_procs()} cores on the box)"}
    else:
        out["gpu_launches"] = r["ba_calls"]
    print(json.dumps(out))


def run_pnp(args):
    """SURVEY §8 row f-2 (motion-only BA): `--windows` pose problems x 2048 fixed landmarks (one KITTI frame's
    keypoints, frontend.py max_keypoints), `--iters` LM iterations; value = observations*iterations/s with the
    arrays resident on the device, e2e = the same call with host arrays; CPU = oracle/pose_numpy.py (numpy)."""
    import torch
    from visualodometry_b200 import BAConfig, BundleAdjuster
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from test_pose import KITTI as intr, make_pose_problem
    B, n = min(args.windows, 1024), 2048
    probs = [make_pose_problem(s_, n=n) for s_ in range(32)]
    T0 = np.stack([probs[i % 32][1] for i in range(B)])
    X = np.concatenate([probs[i % 32][2] for i in range(B)]); uv = np.concatenate([probs[i % 32][3] for i in range(B)])
    off = (np.arange(B + 1) * n).astype(np.int64)
    K = np.array([[intr[0], 0, intr[2]], [0, intr[1], intr[3]], [0, 0, 1.0]])
    ba = BundleAdjuster(K, BAConfig(max_iters=args.iters))
    Xd, uvd = torch.as_tensor(X).cuda(), torch.as_tensor(uv).cuda()

    def timed(fn, reps):
        for _ in range(args.warmup):
            fn()
        torch.cuda.synchronize(); t0 = time.perf_counter()
        for _ in range(reps):
            fn()
        torch.cuda.synchronize()
        return (time.perf_counter() - t0) / reps
    t_dev = timed(lambda: ba.refine_poses(T0, Xd, uvd, off), args.steps)
    t_e2e = timed(lambda: ba.refine_poses(T0, X, uv, off), args.steps)
    _, sms = ba.refine_poses(T0, X, uv, off)
    work = float(sum(s_.iterations for s_ in sms)) * n
    out = {"metric": "BA observations*iterations/sec", "value": work / t_dev, "unit": "obs*iter/s", "n_gpus": 1,
           "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t_dev, "higher_is_better": True,
           "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
           "config": {"workload": f"f-2 motion-only BA: {B} poses x {n} fixed landmarks, {args.iters} LM iterations "
                                  "(pose_lm_kernel, one CTA per pose; timed with host wall clock around the synchronous call)"},
           "e2e": {"value": work / t_e2e, "unit": "obs*iter/s", "ms_per_step": 1e3 * t_e2e,
                   "h2d_bytes_per_step": int(X.nbytes + uv.nbytes + T0.shape[0] * 96), "d2h_bytes_per_step": int(T0.shape[0] * 96)},
           "gpu_launches": args.steps}
    # algorithmic flops per observation*iteration: projection + J_c (~60), 27 weighted sums (108), trial (~40)
    tf = 208.0 * work / t_dev / 1e12
    pk64 = fp64_peak(0)
    out["roofline"] = {"bound": "tensor", "kernel": "pose_lm_kernel", "achieved": tf, "peak": pk64["tflops"],
                       "unit": "TFLOP/s", "frac": tf / pk64["tflops"], "traffic": None, "peak_kind": pk64["kind"],
                       "note": "FP64 pipe (DFMA); 208 algorithmic flop per observation*iteration; the 64 KB of one pose "
                               "problem stay in L1/L2 across iterations"}
    if not args.no_cpu:
        from oracle import pose_numpy
        m = min(B, 16)
        t0 = time.perf_counter()
        its = 0
        for i in range(m):
            _, so = pose_numpy.refine_pose(T0[i], X[i * n:(i + 1) * n], uv[i * n:(i + 1) * n], intr, max_iters=args.iters)
            its += so.iterations
        dt = time.perf_counter() - t0
        out["cpu_baseline"] = {"value": its * n / dt, "unit": "obs*iter/s", "cores": 1, "kind": "port",
                               "sample": f"first {m} poses, oracle/pose_numpy.py (numpy, one core), {dt:.2f} s"}
    print(json.dumps(out))


def c5_measure(mode, solver, iters, steps, warmup, rank, world, local, time_phases=True):
    """BASELINE config 5 (1000 cameras / 200k landmarks / 1M observations) on `world` GPUs, STRONG scaling: one
    problem, S sharded by block row, small state replicated, exchanges through peer memory.  Device-timed
    (CUDA events on the launching stream, L2 flushed between steps), max over ranks.  Returns the record of rank 0."""
    import torch
    import torch.distributed as dist
    from visualodometry_b200 import BAConfig, BundleAdjuster, synth
    p = synth.make_large(0, mode=mode)                       # the same problem on every rank
    out = {}
    cfgs = [(False, steps)] + ([(True, 1)] if time_phases else [])
    for timed, nsteps in cfgs:
        ba = BundleAdjuster(K_of(p.intr), BAConfig(max_iters=iters, device=local, solver=solver, time_phases=timed))
        if world > 1:
            ba.init_distributed()
        ba.set_problem(p.poses, p.points, p.obs_cam, p.obs_pt, p.obs_uv, p.fixed_cam)
        stream = torch.cuda.current_stream().cuda_stream
        flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")

        def sync():
            torch.cuda.synchronize()
            if world > 1:
                dist.barrier()
            torch.cuda.synchronize()
        for _ in range(warmup):
            ba.reset(); ba.solve(stream=stream, wait=False)
        sync()
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(nsteps)]
        for i in range(nsteps):
            flush.fill_(i)
            ev[i][0].record()
            ba.reset(); ba.solve(stream=stream, wait=False)
            ev[i][1].record()
        sync()
        ms = float(np.mean([a.elapsed_time(b) for a, b in ev]))
        sm = ba.summaries()[0]
        if timed:
            t = ba.timings()
            out["phases_ms_per_lm_iteration"] = {k: v / max(1, t["iterations"]) for k, v in t["phases"].items()}
            out["setup_ms"] = t["setup_ms"]
        else:
            tt = torch.tensor([ms], dtype=torch.float64, device="cuda")
            if world > 1:
                dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            ms = float(tt.item())
            out.update({"obs_iter_per_s": p.no * sm.iterations / (ms * 1e-3), "ms_per_step": ms,
                        "ms_per_lm_iteration": ms / max(1, sm.iterations), "lm_iterations": sm.iterations,
                        "pcg_iterations": sm.pcg_iters, "final_cost": sm.final_cost,
                        "solver": "banded Cholesky (replicated)" if sm.pcg_iters == 0 and solver == "auto" else
                                  "dense DMMA Cholesky (replicated)" if solver == "dense" else
                                  "block-Jacobi PCG, S sharded by block row" if world > 1 else "block-Jacobi PCG"})
        ba.close()
        del flush
    return out


def c5_subrecord(args, rank, world, local):
    """The strong-scaling half of the metric (BASELINE config 5), attached to the default line on EVERY N so that the
    driver's 1/2/4/8-GPU runs show it: both visibilities, obs*iter/s, ms per LM iteration and its split."""
    rec = {"workload": "C5: BAL-style, 1000 cameras, 200k landmarks, 1M observations, 20 LM iterations per step; ONE problem "
                       f"on {world} GPU(s) (strong scaling)",
           "partition": ("S sharded by block row (a rank forms and multiplies the rows it owns), back-substitution sharded by "
                         "landmark range, poses / landmarks / LM control replicated; exchanges = stores into peer memory "
                         "(CUDA IPC over NVLink) + epoch flags inside the engine's kernels, no collective call in the LM loop")
                        if world > 1 else "single GPU",
           "scaling": "strong", "n_gpus": world}
    for mode in ("local", "random"):
        rec[mode] = c5_measure(mode, "auto", 20, max(2, min(args.steps, 3)), 2, rank, world, local)
    return rec


def run_reference(args, rank, world):
    """--impl reference: the reference has no BA; the CPU arm is the oracle port on
    the host cores, all threads, on a bounded sample of the same workload."""
    if rank != 0:
        return
    if args.workload == "c2":
        return run_c2(args, "reference")
    if args.workload == "pnp":
        raise SystemExit("bench.py: --workload pnp reports its CPU baseline (oracle/pose_numpy.py) in the b200 line")
    b, name = make_workload(argparse.Namespace(**{**vars(args), "windows": min(args.windows, args.cpu_windows),
                                                  "distinct": None}), 0)
    loss = "huber" if args.workload == "c3" else "trivial"
    vals = []
    for i in range(args.warmup + args.steps):
        cb = cpu_baseline(b, args.iters, loss, args.cpu_windows)
        if i >= args.warmup:
            vals.append(cb)
    v = float(np.mean([c["value"] for c in vals]))
    ms = 1e3 * b.no * args.iters / v
    cb = dict(vals[-1], value=v)
    print(json.dumps({"impl": "reference", "metric": "BA observations*iterations/sec", "value": v,
                      "unit": "obs*iter/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
                      "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                      "dtype": "f64", "data": "synthetic", "config": {"workload": name, "sample": cb["sample"]},
                      "cpu_baseline": cb,
                      "e2e": {"value": v, "unit": "obs*iter/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


def bind_to_gpu_numa_node(local):
    """Multi-rank runs: keep a rank's threads -- and with them the pinned host buffers it allocates next (first touch) --
    on the NUMA node its GPU hangs off, so that the ranks' H2D / D2H streams do not all cross one socket's memory
    controller (the end-to-end number decayed from 1 to 8 ranks).  Best effort: returns what it did."""
    try:
        import torch
        bus = torch.cuda.get_device_properties(local).pci_bus_id
        dom = torch.cuda.get_device_properties(local).pci_domain_id
        dev = torch.cuda.get_device_properties(local).pci_device_id
        path = f"/sys/bus/pci/devices/{dom:04x}:{bus:02x}:{dev:02x}.0/numa_node"
        node = int(open(path).read().strip())
        if node < 0:
            return {"node": node, "bound": False}
        cpus = []
        for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
            lo, _, hi = part.partition("-")
            cpus += list(range(int(lo), int(hi or lo) + 1))
        allowed = sorted(set(cpus) & os.sched_getaffinity(0))
        if not allowed:
            return {"node": node, "bound": False}
        os.sched_setaffinity(0, allowed)
        return {"node": node, "bound": True, "cpus": len(allowed)}
    except Exception as e:   # no sysfs entry, no permission: run unbound
        return {"bound": False, "why": type(e).__name__}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c4", choices=["c1", "c2", "c3", "c4", "c5", "pnp"])
    ap.add_argument("--frames", type=int, default=2700, help="C2: frames of the synthetic sequence")
    ap.add_argument("--windows", type=int, default=4096)
    ap.add_argument("--distinct", type=int, default=256, help="distinct geometries generated (set-up time bound)")
    ap.add_argument("--iters", type=int, default=None, help="LM iterations per step (10 windows, 20 C5)")
    ap.add_argument("--c5-mode", default="local", choices=["local", "random"])
    ap.add_argument("--solver", default="auto", choices=["auto", "pcg", "dense"],
                    help="C5: auto = banded Cholesky if S is block-banded else PCG; dense = blocked DMMA Cholesky")
    ap.add_argument("--cpu-windows", type=int, default=128)
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-c5", action="store_true", help="default workload: skip the C5 strong-scaling sub-record")
    args = ap.parse_args()
    if args.iters is None:
        args.iters = 20 if args.workload == "c5" else 10
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        return run_reference(args, rank, world)

    import torch
    import torch.distributed as dist
    from visualodometry_b200 import BAConfig, BundleAdjuster

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the b2ba engine has no CPU fallback")
    if args.workload == "c2":
        return run_c2(args, "b200") if rank == 0 else None
    if args.workload == "pnp":
        return run_pnp(args) if rank == 0 else None
    torch.cuda.set_device(local)
    numa = bind_to_gpu_numa_node(local) if world > 1 else {"bound": False, "why": "single rank"}
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    b, name = make_workload(args, rank)
    loss = "huber" if args.workload == "c3" else "trivial"
    c5 = args.workload == "c5"
    ba = BundleAdjuster(K_of(b.intr), BAConfig(max_iters=args.iters, loss=loss, device=local,
                                               solver=args.solver if c5 else "auto"))
    if c5 and world > 1:
        ba.init_distributed()
    offs = (None, None, None) if c5 else (b.cam_off, b.pt_off, b.obs_off)
    ba.set_problem(b.poses, b.points, b.obs_cam, b.obs_pt, b.obs_uv, b.fixed_cam, *offs)
    stream = torch.cuda.current_stream().cuda_stream
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda") if b.no * 32 < (256 << 20) else None

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step():
        ba.reset()
        ba.solve(stream=stream, wait=False)

    for _ in range(args.warmup):
        step()
    # ---- timed region: device-resident inputs, CUDA events on the launching stream
    sampler = ClockSampler(local)
    barrier()
    sampler.start()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    t_wall = time.perf_counter()
    for i in range(args.steps):
        if flush is not None:
            flush.fill_(i)          # L2 flush between timed iterations (small workloads)
        ev[i][0].record()
        step()
        ev[i][1].record()
    barrier()
    t_wall = time.perf_counter() - t_wall
    clocks = sampler.stop()
    ms_steps = [a.elapsed_time(c) for a, c in ev]
    ms_total = float(sum(ms_steps))
    sms = ba.summaries()
    iters_done = sum(s.iterations for s in sms)

    # ---- e2e: host buffers in, host results out, through the public API, copies inside the timing
    pin = lambda a: torch.from_numpy(np.ascontiguousarray(a)).pin_memory().numpy()
    hb = dict(poses=pin(b.poses), points=pin(b.points), obs_cam=pin(b.obs_cam), obs_pt=pin(b.obs_pt),
              obs_uv=pin(b.obs_uv), fixed=pin(b.fixed_cam))
    h2d = sum(v.nbytes for v in hb.values()) + b.cam_off.nbytes + b.pt_off.nbytes + b.obs_off.nbytes
    d2h = hb["poses"].nbytes + hb["points"].nbytes + 48 * b.nb
    out_host = (pin(np.empty_like(b.poses, dtype=np.float64)), pin(np.empty_like(b.points, dtype=np.float64)))
    e2e_steps = max(2, min(args.steps, 3))
    from types import SimpleNamespace
    hbatch = SimpleNamespace(poses=hb["poses"], points=hb["points"], obs_cam=hb["obs_cam"], obs_pt=hb["obs_pt"],
                             obs_uv=hb["obs_uv"], fixed_cam=hb["fixed"], cam_off=b.cam_off, pt_off=b.pt_off,
                             obs_off=b.obs_off)
    if not c5:
        ba.optimize_batch(hbatch, out=out_host)     # warm-up of the pipelined path (child engines, streams)
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        if c5:
            ba.set_problem(hb["poses"], hb["points"], hb["obs_cam"], hb["obs_pt"], hb["obs_uv"], hb["fixed"], *offs)
            ba.solve(stream=stream)
            ba.get(out=out_host)            # refined poses / landmarks into pinned host buffers
        else:
            # the public batch call: chunked H2D copies / index build + solve / D2H read-back overlap on three streams
            ba.optimize_batch(hbatch, out=out_host)
    barrier()
    t_e2e = (time.perf_counter() - t0) / e2e_steps

    units = float(b.no) * args.iters          # obs*iter per step per rank
    if c5:
        units = units / world                  # ONE problem shared by all ranks (strong scaling)
    tt = torch.tensor([ms_total, t_e2e * 1e3], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        un = torch.tensor([units], dtype=torch.float64, device="cuda")
        dist.all_reduce(un, op=dist.ReduceOp.SUM)
        units_all = float(un.item())
    else:
        units_all = units
    ms_total, ms_e2e = float(tt[0].item()), float(tt[1].item())
    c5_rec = None
    l2_note = "inputs larger than L2" if flush is None else "L2 flushed between timed steps"
    if args.workload == "c4" and not args.no_c5:
        flush = None
        ba.close()
        torch.cuda.empty_cache()
        try:
            c5_rec = c5_subrecord(args, rank, world, local)     # every rank takes part
        except Exception as e:   # the C4 line of this run must not depend on the sub-record
            c5_rec = {"error": f"{type(e).__name__}: {e}"[:300]}
    if rank == 0:
        pk, pk_kind = peaks()
        pk64 = fp64_peak(local)
        counts = np.diff(np.concatenate([[0], np.cumsum(np.bincount(
            (b.obs_pt.astype(np.int64) + np.repeat(b.pt_off[:-1].astype(np.int64), np.diff(b.obs_off))),
            minlength=int(b.pt_off[-1])))]))
        flops_iter = algorithmic_flops_per_iter(counts)
        nblk = sum(int(n * (n + 1) // 2) for n in (np.diff(b.cam_off) - 1))
        bytes_iter = algorithmic_bytes_per_iter(b.no, int(b.pt_off[-1]), int(b.cam_off[-1]), nblk)
        ms_kernel = ms_total / args.steps       # one window_lm_kernel launch per step
        tf = flops_iter * args.iters / (ms_kernel * 1e-3) / 1e12
        gbs = bytes_iter * args.iters / (ms_kernel * 1e-3) / 1e9
        prof = os.path.join(ROOT, "profiles", "traffic.json")
        traffic = json.load(open(prof)).get(args.workload) if os.path.exists(prof) else None
        out = {"metric": "BA observations*iterations/sec", "value": units_all * args.steps / (ms_total * 1e-3),
               "unit": "obs*iter/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
               "ms_per_step": ms_total / args.steps, "higher_is_better": True,
               "scaling": "strong" if c5 else "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
               "config": {"workload": name, "windows_per_gpu": b.nb, "observations_per_gpu": int(b.no),
                          "lm_iterations": args.iters, "lm_iterations_executed": int(iters_done),
                          "distinct_geometries": args.distinct if args.workload == "c4" else 1,
                          "partition": (f"ONE problem on {world} rank(s): S sharded by block row, back-substitution by landmark range, "
                                        f"small state replicated, peer-memory exchanges inside the kernels (DESIGN 5)") if c5 else
                                       f"{world} x independent window shards, no collective",
                          "l2": l2_note},
               "roofline": {"bound": "tensor", "kernel": "window_lm_kernel", "achieved": tf, "peak": pk64["tflops"],
                            "unit": "TFLOP/s", "frac": tf / pk64["tflops"], "traffic": traffic,
                            "peak_kind": pk64["kind"], "peak_dfma_tflops": pk64["dfma"], "peak_dmma_tflops": pk64["dmma"],
                            "traffic_source": "profiles/traffic.json: dram bytes of the committed ncu --set full capture of this "
                                              "kernel, scaled per window to this launch (not a measurement of this run)",
                            "note": "fp64 pipe bound (DMMA and DFMA share ONE pipe on B200: profiles/r2_fp64_pipes.log); "
                                    "peak = max(DFMA, DMMA) measured in this run; algorithmic flops per SURVEY 8(d)",
                            "hbm_view": {"achieved": gbs, "peak": pk["hbm_gbs"], "unit": "GB/s",
                                         "frac": gbs / pk["hbm_gbs"], "peak_kind": pk_kind}},
               "clocks": clocks, "gpu_launches": args.steps,
               "e2e": {"value": units_all / (ms_e2e * 1e-3), "unit": "obs*iter/s", "h2d_bytes_per_step": int(h2d),
                       "d2h_bytes_per_step": int(d2h), "ms_per_step": ms_e2e},
               "wall_s_timed_region": t_wall, "numa_binding_rank0": numa}
        if c5:
            # dominant kernel: the cooperative PCG (block-sparse mat-vec, HBM/L2 bound).  Algorithmic bytes of
            # the step: 46.4 B per observation*iteration + 288 B per STORED (upper) S block per CG iteration
            # (the mat-vec reads every upper block once) + 7 vectors of 6*Nfree doubles.
            pcg_its = sum(s_.pcg_iters for s_ in sms)
            ncam = int(b.cam_off[-1])
            pairs = set()
            # nnzblk(S) from the visibility: only needed here, on the host, for the byte count
            order = np.lexsort((b.obs_cam, b.obs_pt))
            oc, op = b.obs_cam[order], b.obs_pt[order]
            k = np.bincount(op)
            if (k == k[0]).all():
                cams = oc.reshape(-1, int(k[0]))
                for i in range(cams.shape[1]):
                    for j in range(i, cams.shape[1]):
                        pairs.update((cams[:, i].astype(np.int64) * ncam + cams[:, j]).tolist())
            nupper = len(pairs)
            byt = (46.4 * b.no * iters_done + pcg_its * (288.0 * nupper + 7 * 48.0 * ncam))
            gbs5 = byt / (ms_total / args.steps * 1e-3) / 1e9
            banded = pcg_its == 0    # block-banded S: substructured Cholesky (a chain of dependent 6x6 pivots)
            out["roofline"] = {"bound": "hbm",
                               "kernel": ("large_sblock_kernel / band_forward_kernel (+ z, backsub passes)" if banded
                                          else "large_pcg_kernel (+ z, sblock, backsub passes)"),
                               "achieved": gbs5,
                               "peak": pk["hbm_gbs"], "unit": "GB/s", "frac": gbs5 / pk["hbm_gbs"], "traffic": None,
                               "note": f"whole-step algorithmic bytes / step time; {pcg_its} CG iterations per step; "
                                       + ("the banded solve is latency-bound (DESIGN 4.2), so the HBM fraction only "
                                          "says how far the step is from a pure streaming pass; " if banded else "")
                                       + f"peak {pk_kind}"}
            out["gpu_launches"] = args.steps * (int(iters_done) * (22 if banded else 15) + 2)
            out["config"].pop("windows_per_gpu"); out["config"].pop("distinct_geometries")
            out["config"]["pcg_iterations_per_step"] = int(pcg_its)
        if not args.no_cpu and world == 1:
            if c5:
                out["cpu_baseline"] = cpu_baseline(b, args.iters, loss, 1)
            else:
                out["cpu_baseline"] = cpu_baseline(b, args.iters, loss, args.cpu_windows, with_scipy=args.workload in ("c1", "c3"))
        if c5_rec is not None:
            out["c5"] = c5_rec
        print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
