#!/usr/bin/env python
"""bench.py -- headline benchmark of the B200-native bundle-adjustment path.

A "step" is ONE Levenberg-Marquardt iteration of rig_calibrator's solve (1 fused residual+Jacobian+
Schur build, 1 PCG solve of the reduced camera system, 1 back-substitution, 1 candidate-cost pass,
accept/reject) on the synthetic config named in config.workload (BASELINE.json config 5 by default:
2000 ref frames x 1M landmarks x 8 obs/track, 3-sensor rig).

  value : LM iterations/s, parameters resident in HBM (CUDA events around the LM loop, max over ranks)
  e2e   : LM iterations/s through the C ABI call mvba_solve() with HOST parameter arrays
          (host->device upload and device->host write-back inside the timed region)
  roofline / cpu_baseline : see DESIGN.md "Measurement"

`--impl reference` times the reference's CPU algorithm (the oracle restatement of the functors +
Ceres-1.14-style numeric-diff LM; Ceres itself is not available on this box) on a bounded sample.
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", type=int, default=5)
    ap.add_argument("--scale", type=float, default=1.0, help="shrink the workload (debug only; invalidates the number)")
    ap.add_argument("--cpu-sample", type=float, default=None, help="fraction of the workload timed on the CPU for the in-line cpu_baseline (default per config)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--preconditioner", type=int, default=0)
    return ap.parse_args()


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc, self.nvml = index, [], None, None

    def _poll_nvml(self):
        """NVML in this process (what nvidia-smi reads, without a second process taking the driver's locks every poll: one
        nvidia-smi poll inside a 70 ms timed region was seen to double the host-side wall time)."""
        n, h = self.nvml
        reasons = (("hw_slowdown", n.nvmlClocksThrottleReasonHwSlowdown), ("hw_thermal_slowdown", n.nvmlClocksThrottleReasonHwThermalSlowdown),
                   ("sw_thermal_slowdown", n.nvmlClocksThrottleReasonSwThermalSlowdown), ("sw_power_cap", n.nvmlClocksThrottleReasonSwPowerCap))
        mx = n.nvmlDeviceGetMaxClockInfo(h, n.NVML_CLOCK_SM)
        while not self.done:
            try:
                sm = n.nvmlDeviceGetClockInfo(h, n.NVML_CLOCK_SM)
                mask = n.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                self.rows.append([str(self.index), str(sm), str(mx), ""] + ["Active" if mask & bit else "Not Active" for _, bit in reasons])
            except Exception:
                pass
            time.sleep(0.01)

    def start(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nvml = (pynvml, pynvml.nvmlDeviceGetHandleByIndex(self.index))
            self.done = False
            threading.Thread(target=self._poll_nvml, daemon=True).start()
            return
        except Exception:
            self.nvml = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def mark(self):
        """Samples taken before this call (nvidia-smi start-up, warm-up) are not part of the timed region."""
        self.first = len(self.rows)

    def stop(self):
        if self.nvml is not None:
            time.sleep(0.03)
            self.done = True
        elif self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        else:
            time.sleep(0.15)
            self.proc.terminate()
        sm, mx, reasons = [], [], set()
        for r in self.rows[max(0, getattr(self, "first", 0) - 1):]:
            try:
                sm.append(float(r[1])); mx.append(float(r[2]))
            except Exception:
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[4:8]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": sorted(reasons),
                "samples": len(sm), "source": "nvml" if self.nvml is not None else "nvidia-smi"}


def algorithmic_bytes(pr):
    """bytes the unfused residual/Jacobian + J'J/Schur kernels must move for one build."""
    T = len(pr.sensors)
    free_pose = 7 if pr.float_poses is not None and pr.float_poses[0] else 0
    sens = pr.cam_sensor[pr.cid]
    direct = pr.cam_beg[pr.cid] == pr.cam_end[pr.cid]
    pf = np.full(pr.num_obs, 3.0)  # xyz
    pf += free_pose * np.where(direct, 1, 2)
    for t in range(T):
        m = sens == t
        extra = 0
        if pr.float_rig is not None and pr.float_rig[t] and t != 0: extra += 7
        if pr.float_focal is not None and pr.float_focal[t]: extra += 1
        if pr.float_center is not None and pr.float_center[t]: extra += 2
        if pr.float_dist is not None and pr.float_dist[t]: extra += pr.sensors[t].nd
        pf[m] += extra
    N, P = pr.num_obs, pr.num_landmarks
    k = N / max(P, 1)
    b_rj = float(np.sum(32 + 16 * pf + 24 / k))
    b_schur = float(np.sum(16 * pf + 16) + 72 * P)
    return b_rj, b_schur, float(pf.mean())


def host_cores():
    """Host threads this process may use.  Not OMP_NUM_THREADS: torch.distributed.run exports OMP_NUM_THREADS=1."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except Exception:
        return max(1, os.cpu_count() or 1)


CPU_SAMPLE = {1: 1.0, 2: 1.0, 3: 1.0, 4: 0.25, 5: 0.25}   # default fraction of each config timed for the in-line cpu_baseline


def _oracle_solve(pr, iters, jac_mode, cores):
    """One oracle solve; returns (summary, seconds of the LM loop incl. every linearisation, seconds of one-time structure set-up)."""
    from multiview_b200 import _abi
    import oracle.binding  # noqa: F401  the checker's binding (cpu_baseline / reference arm only)
    lib = C.CDLL(os.path.join(ROOT, "oracle", "liboracle.so"))
    b = _abi.Binding(lib, "oracle_", pr)
    t0 = time.time()
    s, _ = b.solve(_abi.default_options(max_num_iterations=iters), jac_mode=jac_mode, num_threads=cores)
    wall = time.time() - t0
    b.close()
    setup = float(s.seconds_build)          # Solver construction + block indexing (the oracle reports it here)
    return s, max(wall - setup, 1e-9), setup


def cpu_baseline(config_n, full_obs, frac, jac_mode=0, iters=5, cores=None):
    """Times the oracle's Ceres-style LM (jac_mode 0: Ceres-rule numeric-diff Jacobians as the reference incurs them; 1: forward-mode
    analytic) on a bounded sample and scales to the full workload by observation count.  One-time structure analysis is excluded on
    both arms (the GPU arm is timed on a second solve with its structure tables cached)."""
    from multiview_b200 import synth
    cores = cores or host_cores()
    pr = synth.config(config_n, scale=frac)
    s, dt, setup = _oracle_solve(pr, iters, jac_mode, cores)
    it_per_s_sample = s.num_iterations / dt
    ratio = pr.num_obs / full_obs
    how = "Ceres-rule numeric-diff" if jac_mode == 0 else "analytic (forward-mode dual)"
    return {"value": it_per_s_sample * ratio, "unit": "LM iterations/s", "cores": cores, "kind": "port", "jacobians": how,
            "sample": f"{pr.name} at scale {frac:g}: {pr.num_obs} observations / {pr.num_landmarks} landmarks, {s.num_iterations} LM iterations in "
                      f"{dt:.2f} s ({how} Jacobians, OpenMP {cores} threads; {setup:.2f} s of one-time structure set-up excluded)"
                      + ("" if abs(ratio - 1) < 1e-12 else f"; scaled to the full workload by observation count x{ratio:.3g}"),
            "sample_seconds": dt, "sample_lm_iterations_per_s": it_per_s_sample, "same_config": abs(ratio - 1) < 1e-12}


def run_reference(args):
    """The reference arm: the reference's own CPU algorithm (oracle port of the functors + Ceres 1.14's numeric-diff LM with
    ITERATIVE_SCHUR, all host threads) on the SAME config at FULL size.  Rank 0 only."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from multiview_b200 import synth
    t0 = time.time()
    cores = host_cores()
    pr = synth.config(args.config, scale=args.scale)
    snap = pr.snapshot()
    # a short untimed solve: warms the pages / thread pool and measures the per-iteration cost, so that the timed solve can be
    # bounded to a few minutes whatever the box's core count
    w_it = max(1, min(args.warmup, 2))
    s_w, dt_w, _ = _oracle_solve(pr, w_it, 0, cores)
    per_it = dt_w / max(1, s_w.num_iterations)
    pr.restore(snap)
    budget_s = float(os.environ.get("MVBA_REF_BUDGET_S", "200"))
    steps = max(1, args.steps)
    steps_timed = int(max(min(3, steps), min(steps, budget_s / max(per_it, 1e-9))))
    s, dt, setup = _oracle_solve(pr, steps_timed, 0, cores)
    n_it = max(1, s.num_iterations)
    value = n_it / dt
    cb = {"value": value, "unit": "LM iterations/s", "cores": cores, "kind": "port", "jacobians": "Ceres-rule numeric-diff", "same_config": args.scale == 1.0,
          "sample": f"{pr.name} at scale {args.scale:g} (the whole workload): {pr.num_obs} observations / {pr.num_landmarks} landmarks, {n_it} LM iterations in {dt:.2f} s "
                    f"after {s_w.num_iterations} untimed ones (Ceres-rule numeric-diff Jacobians, OpenMP {cores} threads; {setup:.2f} s of one-time structure set-up excluded)",
          "sample_seconds": dt, "steps_timed": n_it}
    line = {"metric": "LM iterations/s", "value": value, "unit": "LM iterations/s", "n_gpus": args.gpus, "steps": n_it, "warmup": s_w.num_iterations,
            "ms_per_step": 1e3 / value if value > 0 else None, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic", "impl": "reference",
            "config": {"workload": pr.name, "ref_frames": len(pr.ref_stamps), "cameras": len(pr.cam_sensor), "landmarks": pr.num_landmarks, "observations": pr.num_obs,
                       "scale": args.scale, "note": "Ceres-equivalent CPU restatement (Ceres itself unavailable on this box)"},
            "cpu_baseline": cb,
            "e2e": {"value": value, "unit": "LM iterations/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "solve": {"initial_cost": s.initial_cost, "final_cost": s.final_cost, "successful_steps": s.num_successful_steps, "pcg_iterations": s.num_linear_iterations},
            "gpu_launches": 0, "wall_s": time.time() - t0}
    print(json.dumps(line), flush=True)


def run_ours(args):
    import torch
    import torch.distributed as dist
    from multiview_b200 import _abi, synth
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    lib = _abi.load_product()
    pr_full = synth.config(args.config, scale=args.scale)
    # landmark ranges cut after ordering the tracks by the earliest image that sees them (each rank touches ~1/N of the images;
    # measured at N = 2 and 8, profiles/r01d_scale_*); MVBA_SHARD_LOCALITY=0 cuts the ranges in input order instead
    locality = os.environ.get("MVBA_SHARD_LOCALITY", "1") == "1"
    pr = pr_full.shard(rank, world, locality=locality) if world > 1 else pr_full
    g = _abi.Binding(lib, "mvba_", pr, device=local)
    if world > 1:
        idbuf = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            raw = (C.c_uint8 * 128)()
            assert lib.mvba_nccl_unique_id(raw) == 0
            idbuf.copy_(torch.tensor(list(raw), dtype=torch.uint8))
        dist.broadcast(idbuf, 0)
        raw = (C.c_uint8 * 128)(*idbuf.cpu().tolist())
        g._ck(lib.mvba_comm_init(g.h, rank, world, raw))
    snap = pr.snapshot()
    opts = dict(preconditioner=args.preconditioner)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # warm-up: W LM iterations from the initial point (also builds the reduced-system structure)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()   # before the warm-up: nvidia-smi takes ~0.1 s to come up and would otherwise sit inside the timed region
    s_w, _ = g.solve(_abi.default_options(max_num_iterations=max(args.warmup, 3), **opts))
    pr.restore(snap)
    barrier()
    sampler.mark()
    t0 = time.perf_counter()
    s, log = g.solve(_abi.default_options(max_num_iterations=args.steps, **opts))
    wall_solve = time.perf_counter() - t0
    barrier()
    wall = time.perf_counter() - t0
    if os.environ.get("MVBA_BENCH_DEBUG"):
        print(f"[rank {rank}] python wall around solve {wall_solve:.4f} s, with barrier {wall:.4f} s, mvba seconds_total {s.seconds_total:.4f}, "
              f"seconds_device {s.seconds_device:.4f}", file=sys.stderr, flush=True)
    clocks = sampler.stop() if rank == 0 else None
    su, sp = C.c_double(), C.c_double()
    lib.mvba_host_seconds(g.h, C.byref(su), C.byref(sp))
    steps = s.num_iterations
    times = torch.tensor([s.seconds_device, wall, s.seconds_build_kernel, s.seconds_build, s.seconds_linear, s.seconds_update], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(times, op=dist.ReduceOp.MAX)
    t_dev, t_wall, t_bk, t_build, t_lin, t_upd = times.tolist()
    # roofline leg: the fused build kernel, CUDA events around its launches inside the timed solve
    b_rj, b_schur, pfree = algorithmic_bytes(pr)
    per_launch_s = t_bk / max(s.num_build_launches, 1)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    achieved = (b_rj + b_schur) / per_launch_s / 1e9
    # DRAM traffic of one build (pass A + pass B kernels) from the committed `ncu --set full` capture of this workload
    traffic = None
    try:
        tr = json.load(open(os.path.join(ROOT, "profiles", "traffic_build.json")))
        if tr.get("workload") == pr_full.name and world == 1 and args.scale == 1.0:
            traffic = float(tr["dram_bytes_per_build"])
    except Exception:
        pass
    # the real bound of the fused build is FP64 work, not HBM: flop executed per build (ncu SASS counts, profiles/*_fp64_build.json)
    # over the measured launch time, against the measured FP64 peak of this GPU (tools/fp64_peak.cu -> profiles/r02_fp64_peak.json)
    fp64 = None
    try:
        fb = json.load(open(os.path.join(ROOT, "profiles", "r02f_fp64_build.json")))
        pk = json.load(open(os.path.join(ROOT, "profiles", "r02_fp64_peak.json")))
        if fb.get("workload") == pr_full.name and world == 1 and args.scale == 1.0:
            peak_tf = max(pk["dfma_only"]["total_tflops"], pk["dmma_only"]["total_tflops"], pk["dfma_and_dmma"]["total_tflops"])
            # the build after a rejected step keeps pass A's sums (same point, new damping): those launches run pass B only
            flop_a = sum(v["flop"] for k, v in fb["kernels"].items() if "tile2" not in k)
            reused = max(0, steps - s.num_successful_steps - (0 if log and log[-1]["step_accepted"] else 1))
            flop_launch = fb["fp64_flop_per_build"] - flop_a * reused / max(s.num_build_launches, 1)
            ach_tf = flop_launch / per_launch_s / 1e12
            fp64 = {"achieved_tflops": ach_tf, "peak_tflops": peak_tf, "frac": ach_tf / peak_tf, "flop_per_launch": flop_launch,
                    "launches_without_pass_a": reused,
                    "source": "profiles/r02f_fp64_build.json (ncu SASS counts) / profiles/r02_fp64_peak.json (measured DFMA+DMMA peak)"}
    except Exception:
        pass
    # timing rule: inputs larger than L2, or say that they are not.  Per rank and per build pass the kernels stream the packed observation
    # records (16 B), the landmarks (24 B) and the static round tables of the tile kernel (about 5x the records, profiles/traffic_build.json)
    in_mb = (pr.num_obs * 16 + pr.num_landmarks * 24) / 1e6
    l2_note = (f"inputs per rank and pass: {in_mb:.0f} MB of observation records + landmarks, ~{6 * in_mb:.0f} MB with the static round tables; "
               + ("larger than the 126 MB L2, no flush" if 6 * in_mb > 126 else "these fit the 126 MB L2 and are not flushed between iterations (secondary line: the headline workload is config 5 on one GPU)"))
    plan = g.build_plan()
    kernel_names = ("camq_kernel + camfold_kernel" if plan["pass_a_accumulators"] else "camera_pass_kernel") + " + " + \
                   ("tile2_kernel" if plan["pass_b"] == 2 else "tile_kernel")
    h2d = sum(a.nbytes for a in pr.param_arrays().values()); d2h = h2d
    line = {"metric": "LM iterations/s", "value": steps / t_dev, "unit": "LM iterations/s", "n_gpus": world, "steps": steps, "warmup": max(args.warmup, 3),
            "ms_per_step": 1e3 * t_dev / steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": pr_full.name, "ref_frames": len(pr_full.ref_stamps), "cameras": len(pr_full.cam_sensor), "landmarks": pr_full.num_landmarks,
                       "observations": pr_full.num_obs, "parallelism": f"landmark-sharded x{world}" + ("" if world == 1 or not locality else " (image-ordered)"), "l2": l2_note,
                       "scale": args.scale, "preconditioner": args.preconditioner, "mean_free_params_per_obs": pfree},
            "evals_per_s": pr_full.num_obs * (s.num_jacobian_evals) / t_dev,
            "e2e": {"value": steps / t_wall, "unit": "LM iterations/s", "h2d_bytes_per_step": h2d / steps, "d2h_bytes_per_step": d2h / steps,
                    "note": "mvba_solve() on host parameter arrays; upload + write-back amortised over the steps of one solve"},
            "gpu_launches": int(s.num_kernel_launches),
            "phases_s": {"build": t_build, "build_kernel": t_bk, "linear": t_lin, "update": t_upd, "device_total": t_dev},
            "solve": {"initial_cost": s.initial_cost, "final_cost": s.final_cost, "successful_steps": s.num_successful_steps,
                      "pcg_iterations": s.num_linear_iterations, "jacobian_evals": s.num_jacobian_evals},
            "roofline": {"bound": "hbm", "kernel": f"fused build = {kernel_names} (residual + Jacobian + J'J + Schur elimination)", "achieved": achieved, "peak": peak,
                         "note": "achieved = algorithmic bytes of the UNFUSED north-star kernels (SURVEY 8(d)) / measured time of the fused kernels; the fused kernels' own bound is FP64 issue, see fp64",
                         "fp64": fp64,
                         "peak_source": "MEASURED_PEAKS.json hbm_gbs" if peaks else "fallback 6.65 TB/s", "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic, "algorithmic_bytes_per_launch": b_rj + b_schur, "launch_ms": 1e3 * per_launch_s, "launches": s.num_build_launches},
            "setup_s": {"problem_assembly": su.value, "structure_tables": sp.value,
                        "note": "one-time host work before the first LM iteration of a pass (rank 0), outside every timed region"},
            "clocks": clocks}
    if rank == 0:
        if not args.no_cpu_baseline and world == 1:
            frac = (args.cpu_sample if args.cpu_sample is not None else CPU_SAMPLE[args.config]) * args.scale
            line["cpu_baseline"] = cpu_baseline(args.config, pr_full.num_obs, frac, jac_mode=0)
            line["cpu_baseline_analytic"] = cpu_baseline(args.config, pr_full.num_obs, frac, jac_mode=1)
        print(json.dumps(line), flush=True)
    g.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)
