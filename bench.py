#!/usr/bin/env python
"""bench.py -- headline benchmark of the FES hot path (BASELINE.json configs[1]).

Workload: Ding2003 with fatigue, B trajectories PER GPU (default 2^20) x the 300 s / 33 Hz fatigue protocol
(1 s on, 1 s off; ns = 9900 intervals x S = 10 RK4 steps, window N = 10), fp64, with the chained forward
sensitivities d x(t) / d x0 (5 columns).  One "step" = one full pass of that protocol over the batch.
Metric: trajectory-steps/s (one trajectory-step = one RK4 step of one trajectory including its 5 tangent columns).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--batch B] [--final-time T] [--impl reference]

N > 1 is launched by torchrun (one rank per GPU, weak scaling: every rank integrates its own B trajectories,
no data-path collective; results are all-gathered over NCCL inside the timed region).  The same line also carries
  "strong"   (N > 1) the SAME B trajectories split over the N ranks, timed the same way;
  "configs"  (N = 1) device-timed sub-records of BASELINE configs 1, 3, 4, 4c, 5 and of the headline workload in fp32,
             with roofline fractions from the instruction counts of the shipped SASS (scripts/bench_configs.py);
  "e2e.breakdown_ms"  the phases of the host-buffer call (allocation, H2D, kernels, D2H) as the library measured them.
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
    dual-number tangents = casadi forward AD) on `threads` host cores, in its TIMED build (-O3 -march=native, compiled
    here, on the machine that runs the benchmark; the strict -ffp-contract=off build stays the parity oracle).
    Returns (steps/s, seconds)."""
    import numpy as np

    from cocofest_b200 import protocols
    from oracle import fes_oracle_c as foc

    cdll, _ = foc.timed_lib()
    params, x0 = protocols.fatigue_population(n_traj)
    p, x = params.numpy().T.copy(), x0.numpy().T.copy()
    win = np.ascontiguousarray(proto["win"][: n_intervals + 1])[None]
    best = None
    for _ in range(repeat):
        t0 = time.perf_counter()
        foc.ivp(0, True, 4, proto["N"], n_intervals, 10, proto["dt"], 0.020, 1.04, p, x, win, None, seed_x0=True,
                n_threads=threads, cdll=cdll)
        el = time.perf_counter() - t0
        best = el if best is None else min(best, el)
    return n_traj * n_intervals * 10 / best, best


def cpu_hoisted_rate(proto, n_traj, n_intervals, threads):
    """The KERNEL's own formulation (hoisted pulse sum, reduced tangents, step propagator) on the host cores: the kernel
    source cocofest_b200/csrc/fes_core.cuh compiled by g++ (tests/host_emul/fes_emul.cpp, -O3 -march=native -fopenmp,
    built here) and driven by a loop over trajectories.  Beside the reference-formulation port it separates what the
    formulation buys from what the hardware buys.  Returns (steps/s, seconds, flags)."""
    import ctypes
    import subprocess

    import numpy as np

    from cocofest_b200 import _lib, fes, protocols

    src = ROOT / "tests" / "host_emul"
    out = src / "_timed"
    out.mkdir(exist_ok=True)
    so = out / "libfes_emul_timed.so"
    flags = ["-O3", "-march=native", "-fopenmp", "-std=c++17", "-fPIC", "-Wno-unknown-pragmas"]
    cxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
    subprocess.run([cxx, *flags, f"-I{ROOT / 'include'}", f"-I{ROOT / 'cocofest_b200' / 'csrc'}", "-shared", "-o", str(so),
                    str(src / "fes_emul.cpp"), "-lm"], check=True)
    emul = _lib._bind(ctypes.CDLL(str(so)))
    try:
        ctypes.CDLL("libgomp.so.1").omp_set_num_threads(int(threads))   # torchrun exports OMP_NUM_THREADS=1
    except OSError:
        pass
    params, x0 = protocols.fatigue_population(n_traj)
    p, x = np.ascontiguousarray(params.numpy()), np.ascontiguousarray(x0.numpy())   # [14, B], [5, B]: the library's layout
    win = np.ascontiguousarray(proto["win"][: n_intervals + 1])
    xf, sf = np.empty((5, n_traj)), np.empty((5, 5, n_traj))
    desc = fes._desc(fes.FES_DING2003, True, 4, proto["N"], n_intervals, 10, fes.FES_DIR_X0, fes.FES_WIN_SHARED, n_traj,
                     proto["dt"], 0.0, None, fes.R0_KM_RELATIONSHIP)
    ptr = lambda a: ctypes.c_void_p(a.ctypes.data)
    buf = fes.FesIvpBuffers(ptr(p), ptr(x), ptr(win), None, ptr(xf), None, ptr(sf), None, None, 0)
    best = None
    for _ in range(2):
        t0 = time.perf_counter()
        rc = emul.fes_ivp_host_f64(ctypes.byref(desc), ctypes.byref(buf), None)
        el = time.perf_counter() - t0
        if rc != 0:
            raise RuntimeError(f"host build of the kernel source returned {rc}")
        best = el if best is None else min(best, el)
    return n_traj * n_intervals * 10 / best, best, " ".join(flags)


def host_threads():
    """Host cores this process may use.  torchrun exports OMP_NUM_THREADS=1, which would silently turn the CPU arm
    into a single-thread run: the oracle is told the thread count explicitly instead."""
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def size_cpu_sample(proto, threads, target_s):
    """Picks (trajectories, intervals) of the same workload worth about target_s seconds of CPU work."""
    n0, k0 = 8 * max(threads, 1), min(proto["ns"], 330)   # large enough that thread start-up does not set the rate
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
    sample = (f"a SAMPLE of the workload named in config: {n_traj} trajectories x the first {n_int} intervals ({n_int * 10} RK4 "
              f"steps) per step, reference formulation with d x/d x0, timed build ({foc.timed_lib()[1]})")
    cfg = workload_config(args, proto)
    cfg["reference_arm_sample"] = sample
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * el / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": cfg,
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
    ap.add_argument("--no-configs", action="store_true", help="skip the sub-records of the other BASELINE configurations")
    ap.add_argument("--no-strong", action="store_true", help="skip the strong-scaling sub-record (N > 1)")
    ap.add_argument("--dtype", default="f64", choices=["f64", "f32"],
                    help="arithmetic type of the path (f32 = north_star's 1e-4 mode, reported against the FFMA rate)")
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3
    if args.impl == "reference":
        run_reference(args)
        return
    # stdout carries exactly ONE line, the JSON record: whatever libraries print there (NCCL's version banner under
    # torchrun) goes to stderr instead -- file descriptor 1 is pointed at stderr and the record is written to the real one
    sys.stdout.flush()
    real_stdout = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)

    import torch
    import torch.distributed as dist

    from cocofest_b200 import _lib, fes, protocols, sharding
    from scripts import bench_configs as bc

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

    def integrate(p, x, tB):
        return fes.ivp(fes.FES_DING2003, True, p, x, win_d, proto["dt"], n_steps=S, n_shooting=ns,
                       dirs=fes.FES_DIR_X0, trace="nodes", trace_batch=tB)

    def device_step():
        res = integrate(params_d, x0_d, trace_B)
        if world > 1:
            full = sharding.gather_last(torch.cat([res.x_final[:, None], res.sens_final], dim=1), batch=B * world)
            return res, full
        return res, None

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    for _ in range(args.warmup):
        device_step()
        flush.fill_(1)
    fma_probe = fes.probe_fma_peak(f32, launches=5, device=dev)  # FFMA rate in f32 mode

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
    kern_ms = [a.elapsed_time(b) for a, b in ev]  # ivp kernel (+ its tiny table kernel, + gather at N > 1)
    elapsed_ms = max_over_ranks(t_start.elapsed_time(t_end))
    value = world * units_per_step * args.steps / (elapsed_ms * 1e-3)

    # ---- strong scaling: the SAME B trajectories split over the N ranks (BASELINE configs[1]: "1M ... on 1-8 x B200") ----
    strong = None
    if world > 1 and not args.no_strong:
        lo, hi = sharding.shard_bounds(B, rank, world)
        # every rank takes its contiguous block of rank 0's population (same seeds on every rank)
        ps_h, xs_h = protocols.fatigue_population(B, offset=0)
        ps_d, xs_d = ps_h[:, lo:hi].to(tdt).to(dev).contiguous(), xs_h[:, lo:hi].to(tdt).to(dev).contiguous()
        tB_s = max(0, min(trace_B, hi) - min(trace_B, lo))   # the first 1024 trajectories of the batch, wherever they live

        def strong_step():
            r = integrate(ps_d, xs_d, tB_s)
            return sharding.gather_last(torch.cat([r.x_final[:, None], r.sens_final], dim=1), batch=B)

        n_strong = max(args.steps, 4)
        for _ in range(args.warmup):
            strong_step()
            flush.fill_(1)
        barrier()
        s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s0.record()
        for k in range(n_strong):
            full = strong_step()
            flush.fill_(k)
        s1.record()
        barrier()
        strong_ms = max_over_ranks(s0.elapsed_time(s1)) / n_strong
        strong_value = units_per_step / (strong_ms * 1e-3)
        # resident blocks of the launch: 64 threads per block, as many blocks per SM as the register file holds
        strong = {
            "value": strong_value, "unit": UNIT, "ms_per_step": strong_ms, "steps": n_strong, "batch_total": B,
            "batch_per_gpu": hi - lo, "scaling": "strong",
            "efficiency_vs_this_runs_weak_rate": strong_value / value,
            "note": "same 2^20-trajectory job split over the ranks; results all-gathered to every rank inside the timed "
                    "region; efficiency = strong rate / the weak rate measured above on the same GPUs (per-GPU work is 1/N, "
                    "so the last wave of resident blocks is the limiter, not NCCL)",
            "checksum_force_mean": float(full[1, 0].mean().item()),
        }
        del ps_d, xs_d, full

    # ---- end-to-end arm: public API with HOST buffers (H2D + kernels + D2H inside the timed region) -----
    e2e = None
    if not args.no_e2e:
        def e2e_step(out=None):
            r = fes.ivp(fes.FES_DING2003, True, params_h, x0_h, win_h, proto["dt"], n_steps=S, n_shooting=ns,
                        dirs=fes.FES_DIR_X0, trace="nodes", trace_batch=trace_B, out=out)
            return r

        r = None
        for _ in range(2):  # allocates the pinned result buffers once and fills the library's staging pool
            r = e2e_step(r)
        barrier()
        phases = []
        t0 = time.perf_counter()
        for _ in range(args.steps):
            r = e2e_step(r)  # results land in the same pinned host buffers, as a solver loop would keep them
            phases.append(fes.last_host_timing())
        torch.cuda.synchronize()
        el = max_over_ranks(time.perf_counter() - t0)
        h2d = params_h.numel() * esz + x0_h.numel() * esz + ns * proto["N"] * 8
        d2h = (r.x_final.numel() + r.sens_final.numel() + r.trace.numel()) * esz
        mean = lambda k: statistics.mean(ph[k] for ph in phases)
        e2e = {"value": world * units_per_step * args.steps / el, "unit": UNIT,
               "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
               "ratio_to_device_value": world * units_per_step * args.steps / el / value,
               "breakdown_ms": {"staging_alloc": mean("alloc_ms"), "h2d": mean("h2d_ms"), "kernels": mean("kernels_ms"),
                                "d2h": mean("d2h_ms"), "call_total": mean("total_ms"),
                                "wall_per_step": 1e3 * el / args.steps,
                                "note": "rank 0's phases as measured inside the library (fes_last_host_timing: CUDA events on "
                                        "the call's stream, host clock for allocation and total); staging memory comes from a "
                                        "library-owned pool that stays mapped between calls"},
               "api": f"cocofest_b200.fes.ivp on pinned host tensors -> fes_ivp_host_{args.dtype} (C ABI)"}

    # ---- roofline of the dominant kernel (fes_ivp_kernel<T, ding2003, fatigue, X0>) -----------------------
    flop = flop_alg_per_step()
    kernel_s = statistics.mean(kern_ms) * 1e-3
    achieved = flop * units_per_step / kernel_s / 1e12
    counts, counts_src = bc.load_sass_counts()
    kname = bc.kernel_name("ivp", args.dtype, fes.FES_DING2003, True, fes.FES_DIR_X0)
    kc = counts.get(kname) if counts else None
    nominal = bc.nominal_peaks(dev, clocks["sm_max_mhz"])           # at the maximum SM clock: the larger denominator
    nominal_run = bc.nominal_peaks(dev, clocks["sm_mhz"]) if clocks.get("sm_mhz") else nominal   # at the clock this run held
    peak_nominal = nominal["fp32_tflops" if f32 else "fp64_tflops"]
    peak = max(peak_nominal, fma_probe)
    traffic, traffic_src = None, "no ncu --set full capture of this build and workload under profiles/"
    try:  # dram bytes of one ncu --set full capture (scripts/ncu_traffic.py), valid only for the build it was taken from
        cap = json.loads((ROOT / "profiles" / "r02_ivp_x0_traffic.json").read_text())
        if cap.get("dtype", "f64") == args.dtype and cap.get("batch") == B and cap.get("final_time_s") == args.final_time:
            # tied to the device code of THIS kernel (sha256 of its SASS, scripts/sass_count.py): other kernels of the
            # library may come and go without orphaning the capture
            if kc and cap.get("kernel_sass_sha256") and cap.get("kernel_sass_sha256") == kc.get("sass_sha256"):
                traffic, traffic_src = cap.get("dram_bytes"), "profiles/r02_ivp_x0_traffic.json (ncu --set full of the same device code: sha256 of the kernel's SASS matches)"
            else:
                traffic_src = "profiles/r02_ivp_x0_traffic.json was captured from another build of the library: dropped"
    except Exception:
        pass
    hbm_bytes = (params_d.numel() + 2 * x0_d.numel() + 25 * B + (ns + 1) * 5 * trace_B) * esz
    peaks = {}
    try:
        peaks = json.loads((ROOT / "MEASURED_PEAKS.json").read_text())
    except Exception:
        pass
    pipe = None
    if kc:
        inst = kc["fp32_inst" if f32 else "fp64_inst"]
        rate = inst * units_per_step / kernel_s
        pipe = {"inst_per_trajectory_step": inst, "flop_executed_per_trajectory_step": kc["fp32_flop" if f32 else "fp64_flop"],
                "issue_rate_ginst_s": rate / 1e9,
                "peak_ginst_s_nominal": nominal["fp32_inst_s" if f32 else "fp64_inst_s"] / 1e9,
                "frac_of_nominal": rate / nominal["fp32_inst_s" if f32 else "fp64_inst_s"],
                "frac_of_nominal_at_run_clock": rate / nominal_run["fp32_inst_s" if f32 else "fp64_inst_s"],
                "frac_of_probe": rate / (fma_probe * 1e12 / 2) if fma_probe else None,
                "registers": kc.get("registers"), "stack_bytes": kc.get("stack_bytes"),
                "note": "FMA + MUL + ADD of the RK4 path counted in the shipped SASS; every one holds an issue slot of the pipe"}
    roofline = {
        "bound": "fp32" if f32 else "fp64", "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak if peak else None,
        "traffic": traffic, "traffic_source": traffic_src,
        "kernel": kname, "kernel_ms": statistics.mean(kern_ms),
        "flop_alg_per_trajectory_step": flop,
        "peaks": {"nominal_tflops": peak_nominal, "nominal_is": f"{nominal['sms']} SMs x {128 if f32 else 64} lanes x 2 x "
                  f"{nominal['sm_mhz']} MHz (ncu's pipe peak x 2 flop)", "probe_tflops": fma_probe,
                  "frac_vs_nominal": achieved / peak_nominal, "frac_vs_probe": achieved / fma_probe if fma_probe else None,
                  "used": "nominal" if peak_nominal >= fma_probe else "probe"},
        "pipe": pipe, "static_counts_source": counts_src,
        "survey_dense_formulation": {"flop_per_trajectory_step": flop_survey_per_step(5, 5),
                                     "note": "SURVEY 8d counts the dense 5-component tangent; the kernel propagates the exact "
                                             "same tangent through 2 components per column and does not execute that count"},
        "peak_source": "the LARGER of the nominal pipe rate at the maximum SM clock and the " + ("FFMA" if f32 else "DFMA")
                       + " probe (fes_probe_fma_peak) of this run; MEASURED_PEAKS.json holds no fp64 / fp32 figure",
        "hbm": {"algorithmic_bytes": hbm_bytes, "achieved_gbs": hbm_bytes / kernel_s / 1e9, "peak_gbs": peaks.get("hbm_gbs"),
                "note": "state, parameters and tangents stay in registers: HBM is idle by design"},
    }

    # ---- the other BASELINE configurations and the fp32 mode, device-timed (rank 0, N = 1 only) ------------
    configs, configs_meta = None, None
    if world == 1 and not args.no_configs:
        del params_d, x0_d
        torch.cuda.empty_cache()
        wanted = ["0", "1", "3", "4", "4c", "5", "5c", "5h", "5pf"]   # 5pf: the parameter columns of a fatigue model (reduced form, 2 x 2 propagator)
        configs, configs_meta = bc.run_configs(wanted, dev, "f64", steps=3, warmup=3, flush=flush)
        other = "f32" if not f32 else "f64"   # the headline workload in the other arithmetic type
        r2, _ = bc.run_configs(["2x0"], dev, other, steps=2, warmup=2, flush=flush, t2=args.final_time)
        configs += r2
        r32, _ = bc.run_configs(["3", "4c", "5"], dev, "f32", steps=3, warmup=3, flush=flush)
        configs += r32

    # ---- CPU baseline: the oracle's reference-formulation port on the host cores (rank 0, N = 1 only) ----
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        from oracle import fes_oracle_c as foc

        threads = host_threads()
        n_traj, n_int = size_cpu_sample(proto, threads, target_s=10.0)
        rate, secs = cpu_reference_rate(proto, n_traj, n_int, threads)
        rate1, secs1 = cpu_reference_rate(proto, max(1, n_traj // max(threads, 1) // 4), n_int, 1)  # scalar row, SURVEY 8d (ii)
        cpu = {"value": rate, "unit": UNIT, "cores": threads, "kind": "port", "seconds": secs,
               "build": foc.timed_lib()[1] + " (compiled on this machine; the parity oracle is the strict -O2 -ffp-contract=off build)",
               "single_thread": {"value": rate1, "seconds": secs1},
               "sample": f"{n_traj} trajectories x first {n_int} intervals ({n_int * 10} RK4 steps) with d x/d x0, "
                         "reference formulation (N-term exp sum per RHS, dual-number tangents), OpenMP over trajectories"}
        try:
            n_h = max(threads, int(rate * 40 / max(rate1, 1.0)) // max(threads, 1) * max(threads, 1))
            rate_h, secs_h, flags_h = cpu_hoisted_rate(proto, min(n_h * 8, 1 << 16), n_int, threads)
            cpu["hoisted"] = {"value": rate_h, "unit": UNIT, "cores": threads, "seconds": secs_h, "build": flags_h,
                              "what": "the kernel's own formulation (fes_core.cuh: hoisted pulse sum, reduced tangents, step "
                                      "propagator) compiled for the host: GPU / hoisted = hardware, hoisted / port = formulation"}
        except Exception as e:  # the row is informative; the baseline above stands without it
            cpu["hoisted"] = {"unavailable": f"{type(e).__name__}: {e}"}

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": elapsed_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": args.dtype, "data": "synthetic", "config": workload_config(args, proto), "clocks": clocks,
            "e2e": e2e, "gpu_launches": int(launches), "roofline": roofline, "cpu_baseline": cpu, "strong": strong,
            "configs": configs, "configs_meta": configs_meta,
            "final_force_mean": float(res.x_final[1].mean().item()),
        }
        real_stdout.write(json.dumps(line) + "\n")
        real_stdout.flush()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
