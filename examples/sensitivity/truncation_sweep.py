"""Summation-truncation sensitivity sweep on the GPU.

Reference: examples/sensitivity/truncation/sensitivity_analysis.py:11-132 -- for every pulse mode (single, doublet,
triplet), every stimulation frequency i = 1..100 Hz over 1 s and every truncation j = 1..i it builds one
``IvpFes(DingModelFrequencyWithFatigue(sum_stim_truncation=j))`` (3-11 s of casadi graph construction per point,
BASELINE.md section 1), integrates it (0.105-0.19 s) and keeps the final F, Cn, A, Km, Tau1: 15 150 points.

Here the points of a pulse mode that share the node grid (``n_shooting`` = ``OcpFes.prepare_n_shooting`` of the
millisecond-rounded pulse times, a divisor of 1000) are ONE launch of ``fes.ivp`` with per-trajectory windows: a window
of truncation j is the window of the largest truncation of the group left-padded with the reference's own pad time
-10 000 000 (ding2003.py:394-397).  The pads contribute exp(-1e7 / tauc) = 0 to both window sums, so a padded window
gives the same bits as the short one (tests/test_truncation_sweep.py).  The output pickles have the reference's keys.

    python examples/sensitivity/truncation_sweep.py [--max-stim 100] [--modes single,doublet,triplet] [--out-dir .]
"""
import argparse
import pickle
import sys
import time
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[2]))

from cocofest_b200 import DingModelFrequencyWithFatigue, fes  # noqa: E402

COEFFICIENT = {"single": 1, "doublet": 2, "triplet": 3}


def sweep_points(mode, min_stim=1, max_stim=100, final_time=1.0):
    """The reference's loop nest for one pulse mode: [(i, j, stim_time, n_shooting)], i = frequency, j = truncation."""
    pts = []
    for i in range(min_stim, max_stim + 1):
        base = list(np.round([k / i for k in range(i)], 3))        # ivp_fes.py:399-401: pulse times rounded to the millisecond
        stim = fes.pulse_mode_times(base, mode)                    # ivp_fes.py:232-256
        ns = fes.prepare_n_shooting(stim, final_time)              # fes_ocp.py:192-222
        for j in range(1, i + 1):
            pts.append((i, j, stim, ns))
    return pts


def padded_windows(points, final_time=1.0):
    """Per-trajectory windows [ns + 1, N_max, B] of points that share n_shooting (each left-padded to the widest)."""
    ns = points[0][3]
    n_max = max(j for _, j, _, _ in points)
    win = np.full((ns + 1, n_max, len(points)), float(fes.PAD_TIME))
    widest = {}   # the window of truncation j is the last j slots of the train's widest window: one builder call per train
    for b, (i, j, stim, _) in enumerate(points):
        if i not in widest:
            widest[i] = fes.stim_windows(stim, ns, final_time, max(q[1] for q in points if q[0] == i))[0]
        win[:, n_max - j:, b] = widest[i][:, widest[i].shape[1] - j:]
    return win


def run_mode(mode, max_stim=100, device="cuda", n_steps=10, final_time=1.0, max_batch_bytes=6 << 30):
    """All points of one pulse mode.  Returns (reference-format dictionary, seconds spent in the launches)."""
    model = DingModelFrequencyWithFatigue()
    p = torch.tensor(model.param_vector(), device=device)
    rest = torch.tensor(np.asarray(model.standard_rest_values(), dtype=np.float64).ravel(), device=device)
    pts = sweep_points(mode, 1, max_stim, final_time)
    finals = {}
    groups = {}
    for pt in pts:
        groups.setdefault(pt[3], []).append(pt)
    spent = 0.0
    for ns, group in sorted(groups.items()):
        group.sort(key=lambda q: q[1])                              # similar truncations together: narrower pads per chunk
        per_traj = (ns + 1) * max(q[1] for q in group) * 8
        chunk = max(1, int(max_batch_bytes // per_traj))
        for lo in range(0, len(group), chunk):
            part = group[lo:lo + chunk]
            win = torch.tensor(padded_windows(part, final_time), device=device)
            x0 = rest[:, None].repeat(1, len(part)).contiguous()
            t0 = time.perf_counter()
            res = fes.ivp(fes.FES_DING2003, True, p, x0, win, final_time / ns, n_steps=n_steps, n_shooting=ns)
            xf = res.x_final.cpu().numpy()
            spent += time.perf_counter() - t0
            for b, (i, j, _, _) in enumerate(part):
                finals[(i, j)] = xf[:, b]
    rows = lambda k: [[float(finals[(i, j)][k]) for j in range(1, i + 1)] for i in range(1, max_stim + 1)]
    out = {
        "parameter_list": [[i, j] for i in range(1, max_stim + 1) for j in range(1, i + 1)],
        "calcium_total_results": rows(0), "force_total_results": rows(1), "a_total_results": rows(2),
        "tau1_total_results": rows(3), "km_total_results": rows(4),
        # the reference times every point on its own; here a launch integrates a whole group, the totals are what compares
        "computations_time": [spent], "computations_time_avg": [spent / len(pts)], "creation_ocp_time": [0.0],
        "repetition": 1, "launch_groups": {int(k): len(v) for k, v in groups.items()},
    }
    return out, spent


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--max-stim", type=int, default=100)
    ap.add_argument("--modes", default="single,doublet,triplet")
    ap.add_argument("--out-dir", default=".")
    ap.add_argument("--n-integration-steps", type=int, default=10, help="IvpFes default: OdeSolver.RK4(n_integration_steps=10), ivp_fes.py:134")
    args = ap.parse_args()
    total_pts, total_s = 0, 0.0
    for mode in args.modes.split(","):
        t0 = time.perf_counter()
        out, spent = run_mode(mode, args.max_stim, n_steps=args.n_integration_steps)
        wall = time.perf_counter() - t0
        n = len(out["parameter_list"])
        total_pts, total_s = total_pts + n, total_s + wall
        with open(Path(args.out_dir) / f"truncation_{mode}.pkl", "wb") as f:
            pickle.dump(out, f)
        print(f"{mode}: {n} (frequency, truncation) points in {len(out['launch_groups'])} node grids, "
              f"{spent:.3f} s in the launches, {wall:.3f} s with window building "
              f"(reference: {n} x (3-11 s build + 0.105-0.19 s integration), BASELINE.md section 1)")
    print(f"total: {total_pts} points in {total_s:.2f} s = {1e3 * total_s / max(total_pts, 1):.3f} ms per point")


if __name__ == "__main__":
    main()
