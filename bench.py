#!/usr/bin/env python
"""bench.py -- headline benchmark of the FES hot path (BASELINE.json configs[1]).

Workload: Ding2003 with fatigue, B trajectories PER GPU (default 2^20) x the 300 s / 33 Hz fatigue protocol
(1 s on, 1 s off; ns = 9900 intervals x S = 10 RK4 steps, window N = 10), fp64, with the chained forward
sensitivities d x(t) / d x0 (5 columns).  One "step" = one full pass of that protocol over the batch.
Metric: trajectory-steps/s (one trajectory-step = one RK4 step of one trajectory including its 5 tangent columns).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--batch B] [--final-time T] [--impl reference]

N > 1 is launched by torchrun (one rank per GPU, weak scaling: every rank integrates its own B trajectories,
no data-path collective; results are all-gathered over NCCL inside the timed region).
"""
import argparse
import json
import os
import statistics
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

METRIC = "fatigue_ode_trajectory_steps_per_s_with_sensitivities"
UNIT = "trajectory-steps/s"


def flop_survey_per_step(nx=5, n_dir=5):
    """SURVEY.md section 8d, DENSE tangent formulation: 4*(24 + 8*[n_dir>0] + 23*n_dir) + 13*nx*(1+n_dir)
    (+,-,*,/ = 1, FMA = 2, exp = 1) = 978 for the fatigue model with the 5 x0 columns."""
    return 4 * (24 + (8 if n_dir else 0) + 23 * n_dir) + 13 * nx * (1 + n_dir)


def flop_alg_per_step():
    """ALGORITHMIC flops of one RK4 trajectory-step of the kernel's own (reduced) formulation, Ding2003 + fatigue with
    the 5 x0 columns -- DESIGN.md section 4 derives it line by line (+,-,*,reciprocal = 1, FMA = 2):
      nominal stage   13 (F' with the shared reciprocal; Cn' is never formed) + 11 (Jacobian entries incl. cW)
                                                                              = 24   x 4 stages =  96
      nominal Cn      Cn(stage) = Cn dl_s + E psi_s from the stage table: 3 x 3 stages + 3 after the step =  12
      nominal update  F: 2 FMA x 3 stages + 1 FMA                                                  =  14
      nominal fatigue A, Tau1, Km reduced to the scalar Wn within the step (fes_core.cuh, rk_step_scheme):
                      Wn' 2 + Wn accumulate 2 + Wn stage 1 + 3 stage states x 2 FMA = 17 per middle stage,
                      14 first stage, 4 last stage, 12 to rebuild the three states                 =  64
      reduced columns advance through the shared 2 x 2 RK4 step propagator (fes_core.cuh, PROP2):
                      Lambda_s / Phi recursion from cF, cW at the four stages (2+3+9+6+10+6+10, Phi sums 14)  =  60
                      forcing weights S = sum_s p_s X_s: Cn0 profile 15, shared p_s m_s 5, A0 / Tau10 / Km0 12 each =  56
                      columns: Phi v 6 x 5 + forcing 2 FMA x 2 components x 4                      =  46
                      shared modes mu, nu: one multiply each after the step                        =   2
    Total 350 (the shipped SASS executes 375: the reciprocal's three Newton FMAs count as one flop here).  It is
    lower than SURVEY 8d's 978 because the fatigue block of every tangent column collapses to one scalar W
    (fes_core.cuh, dir_reduced), Cn rides on the tabulated modes and the (vF, W) rows of all columns share one
    step propagator; the survey figure is reported beside it.  (Stage-by-stage columns, before PROP2: 466.)"""
    return 96 + 12 + 14 + 64 + 60 + 56 + 46 + 2


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons of one GPU during the timed region (NVML)."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.max_mhz, self._stop_evt = index, [], set(), None, threading.Event()
        try:
            import pynvml

            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if self.nv is None:
            return
        nv = self.nv
        names = {
            "hw_slowdown": getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8),
            "hw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40),
            "sw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20),
            "sw_power_cap": getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4),
            "hw_power_brake_slowdown": getattr(nv, "nvmlClocksEventReasonHwPowerBrakeSlowdown", 0x80),
        }
        while not self._stop_evt.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    mask = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    mask = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for k, bit in names.items():
                    if mask & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            self._stop_evt.wait(0.2)

    def stop(self):
        self._stop_evt.set()
        self.join(timeout=2)
        return {"sm_mhz": statistics.median(self.samples) if self.samples else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples)}


def physical_gpu_index(local_index):
    vis = os.environ.get("CUDA_VISIBLE_DEVICES")
    if vis:
        try:
            return int(vis.split(",")[local_index])
        except Exception:
            return local_index
    return local_index


# ---------------------------------------------------------------------------------------------------------
def cpu_reference_rate(proto, n_traj, n_intervals, threads, repeat=1):
    """Times the C restatement of the reference formulation (oracle/fes_oracle.c: N-term exp sum per RHS,
    dual-number tangents = casadi forward AD) on `threads` host cores.  Returns (steps/s, seconds)."""
    import numpy as np

    from cocofest_b200 import protocols
    from oracle import fes_oracle_c as foc

    params, x0 = protocols.fatigue_population(n_traj)
    p, x = params.numpy().T.copy(), x0.numpy().T.copy()
    win = np.ascontiguousarray(proto["win"][: n_intervals + 1])[None]
    best = None
    for _ in range(repeat):
        t0 = time.perf_counter()
        foc.ivp(0, True, 4, proto["N"], n_intervals, 10, proto["dt"], 0.020, 1.04, p, x, win, None, seed_x0=True,
                n_threads=threads)
        el = time.perf_counter() - t0
        best = el if best is None else min(best, el)
    return n_traj * n_intervals * 10 / best, best


def host_threads():
    """Host cores this process may use.  torchrun exports OMP_NUM_THREADS=1, which would silently turn the CPU arm
    into a single-thread run: the oracle is told the thread count explicitly instead."""
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def size_cpu_sample(proto, threads, target_s):
    """Picks (trajectories, intervals) of the same workload worth about target_s seconds of CPU work."""
    n0, k0 = 2 * max(threads, 1), min(proto["ns"], 132)
    rate, _ = cpu_reference_rate(proto, n0, k0, threads, repeat=2)
    steps = rate * target_s
    n_int = min(proto["ns"], 660)            # first 20 s of the protocol
    n_traj = max(threads, int(steps / (n_int * 10) / max(threads, 1)) * max(threads, 1))
    return n_traj, n_int


def run_reference(args):
    """--impl reference: the reference's CPU path (port: casadi/bioptim are not installable)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from cocofest_b200 import protocols
    from oracle import fes_oracle_c as foc

    proto = protocols.fatigue_protocol(final_time=args.final_time)
    threads = host_threads()
    n_traj, n_int = size_cpu_sample(proto, threads, target_s=8.0)
    for _ in range(args.warmup):
        cpu_reference_rate(proto, max(threads, n_traj // 8), n_int, threads)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        cpu_reference_rate(proto, n_traj, n_int, threads)
    el = time.perf_counter() - t0
    value = args.steps * n_traj * n_int * 10 / el
    sample = f"{n_traj} trajectories x first {n_int} intervals ({n_int * 10} RK4 steps) of the workload per step"
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * el / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": workload_config(args, proto),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def workload_config(args, proto):
    return {
        "workload": "BASELINE configs[1]: Ding2003+fatigue, 33 Hz 1s-on/1s-off protocol, "
                    + ("fp32 mode" if getattr(args, "dtype", "f64") == "f32" else "fp64") + ", d x(T)/d x0 sensitivities",
        "model": "DingModelFrequencyWithFatigue", "batch_per_gpu": args.batch, "final_time_s": args.final_time,
        "n_shooting": proto["ns"], "n_integration_steps": 10, "sum_stim_truncation": proto["N"], "scheme": "RK4",
        "sensitivity_columns": 5, "trace": "all 5 states at every node for the first 1024 trajectories",
        "parallelism": f"dp{args.gpus} (batch axis sharded, no data-path collective; NCCL all-gather of results)",
        "l2": "256 MiB scratch buffer rewritten between timed iterations",
    }


# ---------------------------------------------------------------------------------------------------------
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--batch", type=int, default=1 << 20, help="trajectories per GPU")
    ap.add_argument("--final-time", type=int, default=300, help="protocol length in seconds")
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--dtype", default="f64", choices=["f64", "f32"],
                    help="arithmetic type of the path (f32 = north_star's 1e-4 mode, reported against the FFMA rate)")
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3
    if args.impl == "reference":
        run_reference(args)
        return

    import torch
    import torch.distributed as dist

    from cocofest_b200 import _lib, fes, protocols, sharding

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: cocofest_b200 has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    L = _lib.lib()

    proto = protocols.fatigue_protocol(final_time=args.final_time)
    B, ns, S = args.batch, proto["ns"], 10
    trace_B = min(1024, B)
    # host-resident inputs (pinned) for the end-to-end arm; device-resident copies for the kernel arm
    f32 = args.dtype == "f32"
    tdt, esz = (torch.float32, 4) if f32 else (torch.float64, 8)
    params_h, x0_h = protocols.fatigue_population(B, offset=rank)
    params_h, x0_h = params_h.to(tdt).pin_memory(), x0_h.to(tdt).pin_memory()
    win_h = torch.tensor(proto["win"]).pin_memory()
    params_d, x0_d, win_d = params_h.to(dev), x0_h.to(dev), win_h.to(dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    units_per_step = B * ns * S

    def device_step():
        res = fes.ivp(fes.FES_DING2003, True, params_d, x0_d, win_d, proto["dt"], n_steps=S, n_shooting=ns,
                      dirs=fes.FES_DIR_X0, trace="nodes", trace_batch=trace_B)
        if world > 1:
            full = sharding.gather_last(torch.cat([res.x_final[:, None], res.sens_final], dim=1))
            return res, full
        return res, None

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        device_step()
        flush.fill_(1)
    fp64_peak = fes.probe_fma_peak(f32, launches=5, device=dev)  # FFMA rate in f32 mode

    # ---- kernel arm: inputs resident in HBM ----------------------------------------------------------------
    sampler = ClockSampler(physical_gpu_index(local))
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    barrier()
    launches0 = L.fes_launch_count()
    sampler.start()
    t_start = torch.cuda.Event(enable_timing=True)
    t_end = torch.cuda.Event(enable_timing=True)
    t_start.record()
    for k in range(args.steps):
        ev[k][0].record()
        res, _ = device_step()
        ev[k][1].record()
        flush.fill_(k)
    t_end.record()
    barrier()
    clocks = sampler.stop()
    launches = L.fes_launch_count() - launches0
    elapsed_ms = t_start.elapsed_time(t_end)
    kern_ms = [a.elapsed_time(b) for a, b in ev]  # ivp kernel (+ its tiny table kernel, + gather at N > 1)
    t = torch.tensor([elapsed_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    elapsed_ms = float(t.item())
    value = world * units_per_step * args.steps / (elapsed_ms * 1e-3)

    # ---- end-to-end arm: public API with HOST buffers (H2D + kernels + D2H inside the timed region) -----
    e2e = None
    if not args.no_e2e:
        def e2e_step(out=None):
            r = fes.ivp(fes.FES_DING2003, True, params_h, x0_h, win_h, proto["dt"], n_steps=S, n_shooting=ns,
                        dirs=fes.FES_DIR_X0, trace="nodes", trace_batch=trace_B, out=out)
            return r

        r = None
        for _ in range(2):  # allocates the pinned result buffers once and warms the stream-ordered device pool
            r = e2e_step(r)
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            r = e2e_step(r)  # results land in the same pinned host buffers, as a solver loop would keep them
        torch.cuda.synchronize()
        el = time.perf_counter() - t0
        t = torch.tensor([el], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        h2d = params_h.numel() * esz + x0_h.numel() * esz + ns * proto["N"] * 8
        d2h = (r.x_final.numel() + r.sens_final.numel() + r.trace.numel()) * esz
        e2e = {"value": world * units_per_step * args.steps / float(t.item()), "unit": UNIT,
               "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
               "api": f"cocofest_b200.fes.ivp on pinned host tensors -> fes_ivp_host_{args.dtype} (C ABI)"}

    # ---- roofline of the dominant kernel (fes_ivp_kernel<double, ding2003, fatigue, X0>) -----------------
    flop = flop_alg_per_step()
    kernel_s = statistics.mean(kern_ms) * 1e-3
    achieved = flop * units_per_step / kernel_s / 1e12
    static = {}
    try:  # static SASS count of the RK4 path + dram traffic of one ncu --set full capture (scripts/sass_count.py, profiles/)
        static = json.loads((ROOT / "profiles" / "r01_ivp_x0_static.json").read_text())
    except Exception:
        pass
    inst = None if f32 else static.get("fp64_inst_per_rk4_step")  # the SASS count is of the fp64 kernel
    traffic = None
    cap = static.get("ncu_full_capture", {})
    if not f32 and cap.get("batch") == B and cap.get("final_time_s") == args.final_time:
        traffic = cap.get("dram_bytes")
    hbm_bytes = (params_d.numel() + 2 * x0_d.numel() + 25 * B + (ns + 1) * 5 * trace_B) * esz
    peaks = {}
    try:
        peaks = json.loads((ROOT / "MEASURED_PEAKS.json").read_text())
    except Exception:
        pass
    roofline = {
        "bound": "fp32" if f32 else "fp64", "achieved": achieved, "peak": fp64_peak, "unit": "TFLOP/s", "frac": achieved / fp64_peak if fp64_peak else None,
        "traffic": traffic,
        "kernel": f"fes_ivp_kernel<{'float' if f32 else 'double'}, DING2003, fatigue, X0>", "kernel_ms": statistics.mean(kern_ms),
        "flop_alg_per_trajectory_step": flop,
        "survey_dense_formulation": {"flop_per_trajectory_step": flop_survey_per_step(5, 5),
                                     "equivalent_tflops": flop_survey_per_step(5, 5) * units_per_step / kernel_s / 1e12,
                                     "note": "SURVEY 8d counts the dense 5-component tangent; the kernel propagates the "
                                             "exact same tangent through 2 components per column"},
        "fp64_pipe": None if not inst else {
            "inst_per_trajectory_step": inst, "issue_rate_ginst_s": inst * units_per_step / kernel_s / 1e9,
            "peak_ginst_s": fp64_peak * 1e3 / 2, "frac": inst * units_per_step / kernel_s / 1e9 / (fp64_peak * 1e3 / 2),
            "note": "DFMA+DMUL+DADD of the RK4 path counted in the shipped SASS; every one occupies a DFMA issue slot"},
        "peak_source": ("FFMA" if f32 else "DFMA") + " probe (fes_probe_fma_peak) measured in this run on this GPU; "
                       "MEASURED_PEAKS.json holds no fp64 / fp32 figure",
        "hbm": {"algorithmic_bytes": hbm_bytes, "achieved_gbs": hbm_bytes / kernel_s / 1e9, "peak_gbs": peaks.get("hbm_gbs"),
                "note": "state, parameters and tangents stay in registers: HBM is idle by design"},
    }

    # ---- CPU baseline: the oracle's reference-formulation port on the host cores (rank 0, N = 1 only) ----
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        from oracle import fes_oracle_c as foc

        threads = host_threads()
        n_traj, n_int = size_cpu_sample(proto, threads, target_s=12.0)
        rate, secs = cpu_reference_rate(proto, n_traj, n_int, threads)
        rate1, secs1 = cpu_reference_rate(proto, max(1, n_traj // max(threads, 1) // 4), n_int, 1)  # scalar row, SURVEY 8d (ii)
        cpu = {"value": rate, "unit": UNIT, "cores": threads, "kind": "port", "seconds": secs,
               "single_thread": {"value": rate1, "seconds": secs1},
               "sample": f"{n_traj} trajectories x first {n_int} intervals ({n_int * 10} RK4 steps) with d x/d x0, "
                         "reference formulation (N-term exp sum per RHS, dual-number tangents), OpenMP over trajectories"}

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": elapsed_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": args.dtype, "data": "synthetic", "config": workload_config(args, proto), "clocks": clocks,
            "e2e": e2e, "gpu_launches": int(launches), "roofline": roofline, "cpu_baseline": cpu,
            "final_force_mean": float(res.x_final[1].mean().item()),
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
