#!/usr/bin/env python
"""Times the hot path on the BASELINE.json configurations other than the headline one (configs 1, 3, 4, 5;
SURVEY.md section 8d) on one B200 and prints one JSON line per configuration.

    python scripts/bench_configs.py [--configs 1,3,4,5] [--scale 1.0] [--steps 3] [--warmup 3] [--dtype f64]

Inputs are resident in HBM, timing is by CUDA events on torch's current stream (the stream the library
launches on), a 256 MiB buffer is rewritten between timed iterations to flush L2.  `--scale` shrinks the batch
(smoke runs); the committed numbers under profiles/ use scale 1.
"""
import argparse
import json
import statistics
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))

import torch  # noqa: E402

from cocofest_b200 import _lib, fes, protocols  # noqa: E402

UNIT = "trajectory-steps/s"


def flop_alg(nx, n_dir):
    """4*(24 + 8*[n_dir>0] + 23*n_dir) + 13*nx*(1+n_dir) for the fatigue model; the two-state
    models drop the three fatigue rows from the RHS (24 -> 15) as the survey's nx = 2 figure does."""
    rhs = 24 if nx == 5 else 15
    return 4 * (rhs + (8 if n_dir else 0) + 23 * n_dir) + 13 * nx * (1 + n_dir)


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


def line(name, units, ms, nx, nd, out_bytes, in_bytes, peak, dtype, extra):
    t = statistics.mean(ms) * 1e-3
    fl = flop_alg(nx, nd)
    return {
        "config": name, "metric": "trajectory_steps_per_s", "value": units / t, "unit": UNIT, "ms_per_step": 1e3 * t,
        "ms_all": ms, "dtype": dtype, "trajectory_steps": units, "sensitivity_columns": nd,
        "roofline": {"bound": "fp64" if dtype == "f64" else "fp32", "flop_alg_per_trajectory_step": fl,
                     "achieved": fl * units / t / 1e12, "peak": peak, "unit": "TFLOP/s",
                     "frac": fl * units / t / 1e12 / peak if peak else None,
                     "hbm_bytes": in_bytes + out_bytes, "hbm_gbs": (in_bytes + out_bytes) / t / 1e9},
        **extra,
    }


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--configs", default="1,3,4,5")
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
    dt_t = torch.float64 if args.dtype == "f64" else torch.float32
    esz = 8 if args.dtype == "f64" else 4
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    peak = fes.probe_fma_peak(args.dtype == "f32", launches=5, device=dev)
    wanted = [int(c) for c in args.configs.split(",")]
    S = 10

    if 1 in wanted:
        # README IVP: DingModelFrequencyWithFatigue, stim 0..0.9 s at 10 Hz, N = 20, T = 1 -> ns = 10, all 101 points
        B = max(1024, int((1 << 20) * args.scale))
        stim = [k / 10 for k in range(10)]
        ns = fes.prepare_n_shooting(stim, 1.0)
        win, _ = fes.stim_windows(stim, ns, 1.0, 20)
        win = torch.tensor(win, device=dev)
        p = torch.tensor(fes.default_params(fes.FES_DING2003), device=dev).to(dt_t)
        x0 = torch.zeros((5, B), dtype=dt_t, device=dev)
        x0[2], x0[3], x0[4] = p[0], p[2], p[1]
        fn = lambda: fes.ivp(fes.FES_DING2003, True, p, x0, win, 1.0 / ns, n_steps=S, n_shooting=ns, trace="steps", dtype=dt_t)
        out, ms = timed(fn, args.steps, args.warmup, flush)
        print(json.dumps(line("1: README IvpFes DingModelFrequencyWithFatigue, B replicated, all 101 points x 5 states",
                              B * ns * S, ms, 5, 0, out.trace.numel() * esz + out.x_final.numel() * esz, x0.numel() * esz,
                              peak, args.dtype, {"batch": B, "n_shooting": ns, "F_end": float(out.x_final[1, 0])})), flush=True)
        del out, x0

    if 2 in wanted:
        # BASELINE configs[1] as stated: forward only (no sensitivities), F at every node for the first 1024
        # trajectories, final state for all; --t2 shortens the protocol for sweeps
        B = 1 << 20
        T = max(2, args.t2)
        proto = protocols.fatigue_protocol(final_time=T)
        p, x0 = protocols.fatigue_population(B, device=dev)
        p, x0 = p.to(dt_t), x0.to(dt_t)
        win = torch.tensor(proto["win"], device=dev)
        ns = proto["ns"]
        fn = lambda: fes.ivp(fes.FES_DING2003, True, p, x0, win, proto["dt"], n_steps=S, n_shooting=ns, trace="nodes",
                             trace_batch=1024, dtype=dt_t)
        out, ms = timed(fn, args.steps, args.warmup, flush)
        print(json.dumps(line(f"2: Ding2003+fatigue 33 Hz protocol, {T} s, forward only (no sensitivities), 1M trajectories",
                              B * ns * S, ms, 5, 0, (out.trace.numel() + out.x_final.numel()) * esz,
                              (p.numel() + x0.numel()) * esz, peak, args.dtype,
                              {"batch": B, "n_shooting": ns, "final_time_s": T, "F_end_mean": float(out.x_final[1].mean())})),
              flush=True)
        del out, p, x0

    if 3 in wanted:
        B = max(1024, int((1 << 22) * args.scale))
        proto, p, x0, ctrl = protocols.ding2007_sweep(B, device=dev)
        p, x0, ctrl = p.to(dt_t), x0.to(dt_t), ctrl.to(dt_t)
        win = torch.tensor(proto["win"], device=dev)
        ns = proto["ns"]
        fn = lambda: fes.ivp(fes.FES_DING2007, False, p, x0, win, proto["dt"], n_steps=S, ctrl=ctrl, n_shooting=ns,
                             dirs=fes.FES_DIR_P, trace="nodes", sens_trace=True, dtype=dt_t)
        out, ms = timed(fn, args.steps, args.warmup, flush)
        ob = (out.trace.numel() + out.sens_trace.numel() + out.x_final.numel() + out.sens_final.numel()) * esz
        print(json.dumps(line("3: Ding2007 pulse-width sweep, F at 21 nodes + d F/d (a_scale, km_rest, tau1_rest, tau2, pd0, pdt)",
                              B * ns * S, ms, 2, 6, ob, (p.numel() + x0.numel()) * esz, peak, args.dtype,
                              {"batch": B, "n_shooting": ns, "columns": list(out.columns)})), flush=True)
        del out, p, x0

    if 4 in wanted:
        B = max(1024, int((1 << 18) * args.scale))
        proto, p, x0, ctrl = protocols.hmed2018_candidates(B, device=dev)
        p, x0, ctrl = p.to(dt_t), x0.to(dt_t), ctrl.to(dt_t)
        win = torch.tensor(proto["win"], device=dev)
        ns = proto["ns"]
        fn = lambda: fes.ivp(fes.FES_HMED2018, False, p, x0, win, proto["dt"], n_steps=S, ctrl=ctrl, n_shooting=ns,
                             dirs=fes.FES_DIR_P, trace="nodes", sens_trace=True, dtype=dt_t)
        out, ms = timed(fn, args.steps, args.warmup, flush)
        ob = (out.trace.numel() + out.sens_trace.numel() + out.x_final.numel() + out.sens_final.numel()) * esz
        print(json.dumps(line("4: Hmed2018 identification candidates, F at 21 nodes + d F/d 8 parameters",
                              B * ns * S, ms, 2, 8, ob, (p.numel() + x0.numel()) * esz, peak, args.dtype,
                              {"batch": B, "n_shooting": ns, "columns": list(out.columns)})), flush=True)
        # the same candidates through the fused identification kernel: cost + gradient + Gauss-Newton matrix folded in
        # registers (no [21 x 8] Jacobian in HBM), then 10 Levenberg-Marquardt iterations on device
        F_true = fes.ivp(fes.FES_HMED2018, False, torch.tensor(fes.default_params(fes.FES_HMED2018), device=dev).to(dt_t),
                         x0[:, :1].contiguous(), win, proto["dt"], n_steps=S, ctrl=ctrl, n_shooting=ns, trace="nodes", dtype=dt_t)
        target = F_true.trace[:, 1, 0].contiguous()
        lo = {k: v[0] for k, v in protocols.HMED_BOUNDS.items()}
        hi = {k: v[1] for k, v in protocols.HMED_BOUNDS.items()}
        for iters in (0, 10):
            fn = lambda: fes.identify(fes.FES_HMED2018, p, win, proto["dt"], target, n_steps=S, ctrl=ctrl, lower=lo, upper=hi,
                                      max_iter=iters, dtype=dt_t)
            out, ms = timed(fn, args.steps, args.warmup, flush)
            evals = int(out.evaluations.sum().item())
            ob = (out.params.numel() + out.cost.numel() + out.grad.numel() + out.hess.numel()) * esz
            print(json.dumps(line(f"4{'b' if iters == 0 else 'c'}: Hmed2018 fes_identify, max_iter={iters} "
                                  "(cost + gradient + Gauss-Newton matrix folded on device" + (", LM loop on device)" if iters else ")"),
                                  evals * ns * S, ms, 2, 8, ob, (p.numel() + x0.numel()) * esz, peak, args.dtype,
                                  {"batch": B, "n_shooting": ns, "evaluations": evals,
                                   "median_cost": float(out.cost.median().item()), "best_cost": float(out.cost[out.best()].item())})),
                  flush=True)
        del out, p, x0

    if 5 in wanted:
        B = max(1024, int((1 << 16) * args.scale))
        proto, p, X, U = protocols.nmpc_scenarios(B, device=dev)
        p, X, U = p.to(dt_t), X.to(dt_t), U.to(dt_t)
        win = torch.tensor(proto["win"], device=dev)
        ns = proto["ns"]
        fn = lambda: fes.shoot(fes.FES_DING2007, True, p, X, win, proto["dt"], n_steps=S, ctrl=U, want=("x", "u"), dtype=dt_t)
        out, ms = timed(fn, args.steps, args.warmup, flush)
        ob = (out.G.numel() + out.Jx.numel() + out.Ju.numel()) * esz
        print(json.dumps(line("5: NMPC cyclic horizon, DingModelPulseWidthFrequencyWithFatigue, defects + dPhi/dx + dPhi/dpw",
                              B * ns * S, ms, 5, 6, ob, (X.numel() + U.numel()) * esz, peak, args.dtype,
                              {"scenarios": B, "n_shooting": ns, "interval_integrations": B * ns})), flush=True)
    print(json.dumps({"gpu_launches": int(_lib.lib().fes_launch_count()), "fma_peak_tflops": peak, "dtype": args.dtype}), flush=True)


if __name__ == "__main__":
    main()
