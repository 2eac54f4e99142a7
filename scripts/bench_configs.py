#!/usr/bin/env python
"""The BASELINE.json configurations other than the headline one (configs 1, 3, 4, 5) plus the
headline workload in its forward-only and fp32 forms, as timed sub-records.  bench.py imports `run_configs` and puts
the records under "configs" of its JSON line; run directly it prints one JSON line per configuration:

    python scripts/bench_configs.py [--configs 1,2fwd,3,4,4b,4c,5,5pf] [--scale 1.0] [--steps 3] [--warmup 3] [--dtype f64]

Inputs are resident in HBM, timing is by CUDA events on torch's current stream (the stream the library launches on), a
256 MiB buffer is rewritten between timed iterations to flush L2.  Roofline fractions use the instruction counts of the
SHIPPED SASS (cocofest_b200/csrc/sass_counts.json, written by the Makefile right after the link and refused when the
library's sha256 differs) against the nominal pipe rate SMs x 64 (fp64) | 128 (fp32) lanes x SM clock -- a fraction can
therefore not exceed 1.
"""
import argparse
import hashlib
import json
import statistics
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))

import torch  # noqa: E402

from cocofest_b200 import _lib, fes, protocols  # noqa: E402

UNIT = "trajectory-steps/s"
KIND_NAME = {fes.FES_DING2003: "DING2003", fes.FES_DING2007: "DING2007", fes.FES_HMED2018: "HMED2018"}


def load_sass_counts():
    """(counts by kernel name, reason) -- counts is None with a stated reason when the file is missing or does not
    describe the library that is loaded."""
    path = _lib.LIB_PATH.parent / "sass_counts.json"
    if not path.exists():
        return None, f"{path.name} not found next to the library (make -C cocofest_b200/csrc writes it)"
    data = json.loads(path.read_text())
    sha = hashlib.sha256(_lib.LIB_PATH.read_bytes()).hexdigest()
    if data.get("lib_sha256") != sha:
        return None, f"{path.name} describes another build (sha256 {str(data.get('lib_sha256'))[:12]}.. != loaded {sha[:12]}..)"
    return data["kernels"], f"{path.name}, sha256 {sha[:16]}.. = the loaded {_lib.LIB_PATH.name}"


def kernel_name(entry, dtype, kind, fatigue=None, groups=None):
    t = "double" if dtype in ("f64", torch.float64) else "float"
    if entry == "ident":   # shared stimulation train and controls: the variant that stages them in shared memory
        return f"fes_ident_kernel<{t}, {KIND_NAME[kind]}, staged>"
    bits = ((1, "X0"), (2, "U"), (32, "C"), (4, "P"), (16, "PF"), (8, "W"))
    g = "|".join(n for b, n in bits if groups & b) or "0"
    return f"fes_{entry}_kernel<{t}, {KIND_NAME[kind]}, {'fatigue' if fatigue else 'no_fatigue'}, {g}>"


def nominal_peaks(dev, sm_mhz=None):
    """Nominal pipe rates of this GPU at `sm_mhz` (default: its maximum SM clock): thread-instructions/s of the fp64 pipe
    (64 lanes per SM and clock = ncu's sm__inst_executed_pipe_fp64 peak) and of the fp32 FMA pipe (128 lanes)."""
    sms = torch.cuda.get_device_properties(dev).multi_processor_count
    if sm_mhz is None:
        try:
            import pynvml

            pynvml.nvmlInit()
            sm_mhz = pynvml.nvmlDeviceGetMaxClockInfo(pynvml.nvmlDeviceGetHandleByIndex(dev.index or 0), pynvml.NVML_CLOCK_SM)
        except Exception:
            sm_mhz = 1965
    hz = sm_mhz * 1e6
    return {"sms": sms, "sm_mhz": sm_mhz, "fp64_inst_s": sms * 64 * hz, "fp32_inst_s": sms * 128 * hz,
            "fp64_tflops": sms * 64 * 2 * hz / 1e12, "fp32_tflops": sms * 128 * 2 * hz / 1e12}


def timed(fn, steps, warmup, flush):
    for _ in range(warmup):
        out = fn()
        flush.fill_(1)
    torch.cuda.synchronize()
    ms = []
    for k in range(steps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        out = fn()
        e1.record()
        flush.fill_(k)
        torch.cuda.synchronize()
        ms.append(e0.elapsed_time(e1))
    return out, ms


def record(name, kernel, units, ms, out_bytes, in_bytes, dtype, counts, peaks, hbm_peak_gbs, extra):
    """One sub-record.  roofline: executed flops of the shipped SASS (2 per FMA, 1 per MUL / ADD) per trajectory-step x
    trajectory-steps / time against the nominal pipe peak; issue_frac: the same in issue slots (every FMA / MUL / ADD holds
    one lane of the pipe for one clock) -- the two agree up to the FMA share of the mix."""
    t = statistics.mean(ms) * 1e-3
    rec = {"config": name, "value": units / t, "unit": UNIT, "ms": 1e3 * t, "ms_all": [round(m, 4) for m in ms], "dtype": dtype,
           "trajectory_steps": units, "kernel": kernel}
    f64 = dtype == "f64"
    k = counts.get(kernel) if counts else None
    hbm = {"algorithmic_bytes": in_bytes + out_bytes, "achieved_gbs": (in_bytes + out_bytes) / t / 1e9, "peak_gbs": hbm_peak_gbs,
           "frac": (in_bytes + out_bytes) / t / 1e9 / hbm_peak_gbs if hbm_peak_gbs else None}
    if k:
        inst = k["fp64_inst"] if f64 else k["fp32_inst"]
        flop = k["fp64_flop"] if f64 else k["fp32_flop"]
        peak_t = peaks["fp64_tflops"] if f64 else peaks["fp32_tflops"]
        peak_i = peaks["fp64_inst_s"] if f64 else peaks["fp32_inst_s"]
        ach = flop * units / t / 1e12
        rec["roofline"] = {"bound": "fp64" if f64 else "fp32", "achieved": ach, "peak": peak_t, "unit": "TFLOP/s", "frac": ach / peak_t,
                           "issue_frac": inst * units / t / peak_i, "inst_per_trajectory_step": inst,
                           "flop_executed_per_trajectory_step": flop, "registers": k.get("registers"),
                           "stack_bytes": k.get("stack_bytes")}
        if hbm["frac"] and hbm["frac"] > rec["roofline"]["issue_frac"]:
            rec["roofline"]["bound"] = "hbm"
    else:
        rec["roofline"] = None
    rec["hbm"] = hbm
    rec.update(extra)
    return rec


def run_configs(wanted, dev, dtype="f64", steps=3, warmup=3, scale=1.0, flush=None, t2=300):
    """Runs the configurations in `wanted` (strings: "0", "1", "2fwd", "2x0", "2pf", "3", "4", "4b", "4c", "5", "5pf") on `dev`; returns
    (records, meta)."""
    dt_t = torch.float64 if dtype == "f64" else torch.float32
    esz = 8 if dtype == "f64" else 4
    if flush is None:
        flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    counts, counts_src = load_sass_counts()
    peaks = nominal_peaks(dev)
    hbm_peak = None
    try:
        hbm_peak = json.loads((ROOT / "MEASURED_PEAKS.json").read_text()).get("hbm_gbs")
    except Exception:
        pass
    S = 10
    recs = []
    add = lambda *a: recs.append(record(*a[:7], counts, peaks, hbm_peak, a[7]))

    if "0" in wanted:
        # BASELINE configs[0]: the README example itself -- ONE trajectory through the reference-named class, wall clock of
        # construction and of integrate() (pageable host arrays in, dict of numpy arrays out), as a user would time it.  The
        # reference's own figures for this call: 0.105 - 0.19 s per integrate(), 2.9 - 10.8 s per construction.
        import time as _time

        import cocofest_b200 as cf

        def once():
            model = cf.DingModelFrequencyWithFatigue(stim_time=[k / 10 for k in range(10)], sum_stim_truncation=20)
            t0 = _time.perf_counter()
            ivp = cf.IvpFes({"model": model}, {"final_time": 1, "ode_solver": cf.OdeSolver.RK4(n_integration_steps=10)}, device=dev)
            t1 = _time.perf_counter()
            res, _ = ivp.integrate()
            return t1 - t0, _time.perf_counter() - t1, float(res["F"][0, -1])

        for _ in range(3):
            once()
        runs = [once() for _ in range(20)]
        build_ms = 1e3 * statistics.median(r[0] for r in runs)
        integ_ms = 1e3 * statistics.median(r[1] for r in runs)
        recs.append({"config": "0: README IvpFes example, ONE trajectory (n_stim = 10, final_time = 1 s, 100 RK4 steps), wall clock per call",
                     "value": 1e3 / integ_ms, "unit": "integrate() calls/s", "ms": integ_ms, "construction_ms": build_ms, "dtype": "f64",
                     "F_end": runs[0][2], "roofline": None, "hbm": {"frac": None},
                     "reference": "0.105 - 0.19 s per integrate(), 2.9 - 10.8 s per IvpFes(...) construction (BASELINE.md section 1)"})

    if "1" in wanted:
        # README IVP: DingModelFrequencyWithFatigue, stim 0..0.9 s at 10 Hz, N = 20, T = 1 -> ns = 10, all 101 points
        B = max(1024, int((1 << 20) * scale))
        stim = [k / 10 for k in range(10)]
        ns = fes.prepare_n_shooting(stim, 1.0)
        win, _ = fes.stim_windows(stim, ns, 1.0, 20)
        win = torch.tensor(win, device=dev)
        p = torch.tensor(fes.default_params(fes.FES_DING2003), device=dev).to(dt_t)
        x0 = torch.zeros((5, B), dtype=dt_t, device=dev)
        x0[2], x0[3], x0[4] = p[0], p[2], p[1]
        # IvpPlan = the built problem (what IvpFes(...) is to the reference); the timed call is one integrate()
        fn = fes.IvpPlan(fes.FES_DING2003, True, p, x0, win, 1.0 / ns, n_steps=S, n_shooting=ns, trace="steps", dtype=dt_t)
        out, ms = timed(fn, steps, warmup, flush)
        # this configuration is a pure WRITE stream; MEASURED_PEAKS.json's figure is a copy (reads + writes): measure what
        # a plain fill of a buffer of the same size reaches on this device, as the denominator that applies
        probe = torch.empty_like(out.trace)
        _, ms_fill = timed(lambda: probe.fill_(1.0), steps, warmup, flush)
        fill_gbs = probe.numel() * esz / (statistics.mean(ms_fill) * 1e-3) / 1e9
        del probe
        add("1: README IvpFes DingModelFrequencyWithFatigue x B, all 101 points x 5 states written",
            kernel_name("ivp", dtype, fes.FES_DING2003, True, 0), B * ns * S, ms,
            out.trace.numel() * esz + out.x_final.numel() * esz, x0.numel() * esz, dtype,
            {"batch": B, "n_shooting": ns, "F_end": float(out.x_final[1, 0]),
             "write_stream": {"fill_of_the_same_size_gbs": fill_gbs, "what": "torch fill_ of a buffer as large as the trace: the write-only rate of this device"}})
        del out, x0

    for tag, dirs in (("2fwd", 0), ("2x0", fes.FES_DIR_X0), ("2pf", fes.FES_DIR_P | fes.FES_DIR_PF)):
        if tag not in wanted:
            continue
        # BASELINE configs[1]: forward only (as the config line states it) / with d x/d x0 (the headline, e.g. in fp32)
        B = max(1024, int((1 << 20) * scale))
        T = max(2, t2)
        proto = protocols.fatigue_protocol(final_time=T)
        p, x0 = protocols.fatigue_population(B, device=dev)
        p, x0 = p.to(dt_t), x0.to(dt_t)
        win = torch.tensor(proto["win"], device=dev)
        ns = proto["ns"]
        fn = lambda: fes.ivp(fes.FES_DING2003, True, p, x0, win, proto["dt"], n_steps=S, n_shooting=ns, trace="nodes",
                             trace_batch=1024, dirs=dirs, dtype=dt_t)
        out, ms = timed(fn, max(1, min(steps, 2)), max(1, min(warmup, 2)), flush)
        ob = (out.trace.numel() + out.x_final.numel() + (out.sens_final.numel() if dirs else 0)) * esz
        what = {0: "forward only", fes.FES_DIR_X0: "with d x/d x0"}.get(dirs, "with d x/d (4 model + 4 fatigue parameters)")
        add(f"2: Ding2003+fatigue 33 Hz protocol, {T} s, 1M trajectories, " + what,
            kernel_name("ivp", dtype, fes.FES_DING2003, True, dirs), B * ns * S, ms, ob, (p.numel() + x0.numel()) * esz, dtype,
            {"batch": B, "n_shooting": ns, "final_time_s": T, "F_end_mean": float(out.x_final[1].mean())})
        del out, p, x0

    if "3" in wanted:
        B = max(1024, int((1 << 22) * scale))
        proto, p, x0, ctrl = protocols.ding2007_sweep(B, device=dev)
        p, x0, ctrl = p.to(dt_t), x0.to(dt_t), ctrl.to(dt_t)
        win = torch.tensor(proto["win"], device=dev)
        ns = proto["ns"]
        fn = fes.IvpPlan(fes.FES_DING2007, False, p, x0, win, proto["dt"], n_steps=S, ctrl=ctrl, n_shooting=ns,
                         dirs=fes.FES_DIR_P, trace="nodes", sens_trace=True, dtype=dt_t)
        out, ms = timed(fn, steps, warmup, flush)
        ob = (out.trace.numel() + out.sens_trace.numel() + out.x_final.numel() + out.sens_final.numel()) * esz
        add("3: Ding2007 pulse-width sweep, 4M parameter sets, F at 21 nodes + d F/d (a_scale, km_rest, tau1_rest, tau2, pd0, pdt)",
            kernel_name("ivp", dtype, fes.FES_DING2007, False, fes.FES_DIR_P), B * ns * S, ms, ob, (p.numel() + x0.numel()) * esz,
            dtype, {"batch": B, "n_shooting": ns, "columns": list(out.columns)})
        del out, p, x0

    if {"4", "4b", "4c"} & set(wanted):
        B = max(1024, int((1 << 18) * scale))
        proto, p, x0, ctrl = protocols.hmed2018_candidates(B, device=dev)
        p, x0, ctrl = p.to(dt_t), x0.to(dt_t), ctrl.to(dt_t)
        win = torch.tensor(proto["win"], device=dev)
        ns = proto["ns"]
        if "4" in wanted:
            fn = fes.IvpPlan(fes.FES_HMED2018, False, p, x0, win, proto["dt"], n_steps=S, ctrl=ctrl, n_shooting=ns,
                             dirs=fes.FES_DIR_P, trace="nodes", sens_trace=True, dtype=dt_t)
            out, ms = timed(fn, steps, warmup, flush)
            ob = (out.trace.numel() + out.sens_trace.numel() + out.x_final.numel() + out.sens_final.numel()) * esz
            add("4: Hmed2018, 256K candidate vectors, residual inputs: F at 21 nodes + d F/d 8 parameters",
                kernel_name("ivp", dtype, fes.FES_HMED2018, False, fes.FES_DIR_P), B * ns * S, ms, ob,
                (p.numel() + x0.numel()) * esz, dtype, {"batch": B, "n_shooting": ns, "columns": list(out.columns)})
        # the same candidates through the fused identification kernel: cost + gradient + Gauss-Newton matrix folded on
        # chip (no [21 x 8] Jacobian in HBM), then 10 Levenberg-Marquardt iterations on device
        F_true = fes.ivp(fes.FES_HMED2018, False, torch.tensor(fes.default_params(fes.FES_HMED2018), device=dev).to(dt_t),
                         x0[:, :1].contiguous(), win, proto["dt"], n_steps=S, ctrl=ctrl, n_shooting=ns, trace="nodes", dtype=dt_t)
        target = F_true.trace[:, 1, 0].contiguous()
        lo = {k: v[0] for k, v in protocols.HMED_BOUNDS.items()}
        hi = {k: v[1] for k, v in protocols.HMED_BOUNDS.items()}
        for tag, iters in (("4b", 0), ("4c", 10)):
            if tag not in wanted:
                continue
            fn = lambda: fes.identify(fes.FES_HMED2018, p, win, proto["dt"], target, n_steps=S, ctrl=ctrl, lower=lo, upper=hi,
                                      max_iter=iters, dtype=dt_t)
            out, ms = timed(fn, steps, warmup, flush)
            evals = int(out.evaluations.sum().item())
            ob = (out.params.numel() + out.cost.numel() + out.grad.numel() + out.hess.numel()) * esz
            add(f"{tag}: Hmed2018 fes_identify, 256K candidates, max_iter={iters} (residuals + Jacobians folded into cost / gradient / "
                "Gauss-Newton matrix on chip" + (", bounded LM loop on device)" if iters else ")"),
                kernel_name("ident", dtype, fes.FES_HMED2018), evals * ns * S, ms, ob, (p.numel() + x0.numel()) * esz, dtype,
                {"batch": B, "n_shooting": ns, "evaluations": evals, "median_cost": float(out.cost.median().item()),
                 "best_cost": float(out.cost[out.best()].item())})
        del p, x0

    if {"5", "5pf", "5c", "5h"} & set(wanted):
        B = max(1024, int((1 << 16) * scale))
        proto, p, X, U = protocols.nmpc_scenarios(B, device=dev)
        p, X, U = p.to(dt_t), X.to(dt_t), U.to(dt_t)
        win = torch.tensor(proto["win"], device=dev)
        ns = proto["ns"]
        if "5" in wanted:
            # a built problem evaluated once per solver iteration: the plan is what prepare_ocp is to the reference (set-up and
            # window tables once), the timed call is one Ipopt iteration's worth of defects and Jacobian blocks
            plan = fes.ShootPlan(fes.FES_DING2007, True, p, win, proto["dt"], ns, B, n_steps=S, ctrl=U, want=("x", "u"), dtype=dt_t)
            fn = lambda: plan(X, U)
            out, ms = timed(fn, steps, warmup, flush)
            ob = (out.G.numel() + out.Jx.numel() + out.Ju.numel()) * esz
            one_shot, ms1 = timed(lambda: fes.shoot(fes.FES_DING2007, True, p, X, win, proto["dt"], n_steps=S, ctrl=U, want=("x", "u"),
                                                    dtype=dt_t), steps, warmup, flush)
            assert torch.equal(one_shot.Jx, out.Jx) and torch.equal(one_shot.G, out.G)
            add("5: NMPC cyclic horizon, 64K scenarios x 33 intervals, defects + dPhi/dx + dPhi/dpw (ShootPlan call per iteration)",
                kernel_name("shoot", dtype, fes.FES_DING2007, True, fes.FES_DIR_X0 | fes.FES_DIR_U), B * ns * S, ms, ob,
                (X.numel() + U.numel()) * esz, dtype, {"scenarios": B, "n_shooting": ns, "interval_integrations": B * ns,
                                                       "one_shot_shoot_call_ms": statistics.mean(ms1)})
        if "5c" in wanted:
            # the same horizon with the force factor fl * fv + fp of every interval as an input (the FES-MSK seam) and d Phi / d factor
            g = torch.Generator(device="cpu").manual_seed(6)
            fc = torch.empty((ns, B), dtype=torch.float64).uniform_(0.5, 1.2, generator=g).to(dev).to(dt_t)
            ext = torch.cat([U[:ns], fc[:, None, :]], dim=1).contiguous()
            plan_c = fes.ShootPlan(fes.FES_DING2007, True, p, win, proto["dt"], ns, B, n_steps=S, ctrl=ext, dtype=dt_t, force_factor=True)
            out, ms = timed(lambda: plan_c(X, ext), steps, warmup, flush)
            ob = (out.G.numel() + out.Jx.numel() + out.Ju.numel() + out.Jc.numel()) * esz
            add("5c: the same horizon with the force factor fl*fv + fp per interval as an input, + dPhi/dfactor (ShootPlan call)",
                kernel_name("shoot", dtype, fes.FES_DING2007, True, fes.FES_DIR_X0 | fes.FES_DIR_U | fes.FES_DIR_C), B * ns * S, ms, ob,
                (X.numel() + ext.numel()) * esz, dtype, {"scenarios": B, "n_shooting": ns})
        if "5h" in wanted and dtype == "f64":
            # one exact-Hessian solver iteration's worth of second derivatives: H_k = sum_r lambda_kr d2 Phi_kr / dz dz for every
            # interval (21 pairs of (x, pw) each), one launch of the second-order tangent kernel (fes_hess.cuh); beside it the
            # route it replaces, central differences of the exact Jacobian (2 nz = 12 ShootPlan calls)
            g = torch.Generator(device="cpu").manual_seed(7)
            lam = torch.empty((ns, 5, B), dtype=torch.float64).normal_(generator=g).to(dev)
            fn = lambda: fes.shoot_hessian(fes.FES_DING2007, True, p, X, win, proto["dt"], n_steps=S, ctrl=U, multipliers=lam, second=False)
            out, ms = timed(fn, steps, warmup, flush)
            plan_fd = fes.ShootPlan(fes.FES_DING2007, True, p, win, proto["dt"], ns, B, n_steps=S, ctrl=U, want=("x", "u"), defects=False, dtype=dt_t)

            def fd_route():
                for _ in range(12):
                    r = plan_fd(X, U)
                return r
            _, ms_fd = timed(fd_route, steps, warmup, flush)
            # static counts of the hyper-dual step are per (pair, step): scaled to the config's unit (one RK4 step of an interval with
            # all its pairs) so that the issue fraction reads like the other records'
            kname, pairs = "fes_hess_kernel<DING2007, fatigue, X0|U>", out.H.shape[1]
            c2 = None
            if counts and kname in counts:
                c2 = {kname: {**counts[kname], **{f: counts[kname][f] * pairs for f in ("fp64_inst", "fp64_flop", "fp32_inst", "fp32_flop")}}}
            recs.append(record(
                "5h: the same horizon, exact Hessian blocks sum_r lambda_r d2Phi_r/dz dz (21 pairs per interval) in one launch: second-order tangents, one thread per (scenario, interval, pair)",
                kname, B * ns * S, ms, out.H.numel() * 8, (X.numel() + U.numel() + lam.numel()) * 8, dtype, c2, peaks, hbm_peak,
                {"scenarios": B, "n_shooting": ns, "pairs": pairs, "pair_steps_per_s": B * ns * S * pairs / (statistics.mean(ms) * 1e-3),
                 "counts_note": "inst_per_trajectory_step = 21 pairs x the per-pair RK4 step of the shipped SASS",
                 "finite_difference_route_ms": statistics.mean(ms_fd),
                 "finite_difference_route": "12 ShootPlan calls (central differences of the exact Jacobian, ~1e-10), arithmetic of the differences not included"}))
        if "5pf" in wanted:
            # the dense parameter columns of a fatigue model (what a fatigue-parameter identification would differentiate)
            fn = lambda: fes.shoot(fes.FES_DING2007, True, p, X, win, proto["dt"], n_steps=S, ctrl=U, want=("p", "pf"),
                                   defects=False, dtype=dt_t)
            out, ms = timed(fn, steps, warmup, flush)
            add("5pf: the same horizon, dPhi/d(6 model + 4 fatigue parameters): reduced parameter columns through the 2 x 2 step propagator",
                kernel_name("shoot", dtype, fes.FES_DING2007, True, fes.FES_DIR_P | fes.FES_DIR_PF), B * ns * S, ms,
                out.Jp.numel() * esz, (X.numel() + U.numel()) * esz, dtype, {"scenarios": B, "n_shooting": ns})
    meta = {"sass_counts": counts_src, "nominal_peaks": peaks}
    return recs, meta


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--configs", default="0,1,3,4,4b,4c,5")
    ap.add_argument("--scale", type=float, default=1.0)
    ap.add_argument("--t2", type=int, default=300, help="protocol length of config 2 in seconds")
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--dtype", default="f64", choices=["f64", "f32"])
    args = ap.parse_args()
    if not torch.cuda.is_available():
        raise SystemExit("needs a CUDA device: cocofest_b200 has no CPU fallback")
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    recs, meta = run_configs(args.configs.split(","), dev, args.dtype, args.steps, args.warmup, args.scale, t2=args.t2)
    for r in recs:
        print(json.dumps(r), flush=True)
    meta["gpu_launches"] = int(_lib.lib().fes_launch_count())
    meta["fma_probe_tflops"] = fes.probe_fma_peak(args.dtype == "f32", launches=5, device=dev)
    print(json.dumps(meta), flush=True)


if __name__ == "__main__":
    main()
