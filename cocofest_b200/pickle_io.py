"""On-disk formats of the reference, so that existing data sets feed the batched GPU path and its results can be
read by the reference's plotting scripts.

* identification data: ``{"time", "force", "stim_time"}`` + ``"pulse_width"`` (Ding2007) or ``"pulse_intensity"``
  (Hmed2018), one pickle per recorded train (identification/identification_method.py:26-41,
  examples/identification/*.py:43-47);
* solution export: the dictionary of ``SolutionToPickle.pickle`` (result/pickle.py:25-36): ``time``, ``states``
  (key -> 1-D array), ``control``, ``parameters``, ``parameters_bounds``, ``time_to_optimize``, ``bio_model_path``,
  with duplicated node times removed as ``remove_duplicates`` does (:43-58).
"""
from __future__ import annotations

import pickle

import numpy as np


def save_identification_data(path, time, force, stim_time, pulse_width=None, pulse_intensity=None):
    data = {"time": np.asarray(time), "force": np.asarray(force), "stim_time": list(stim_time)}
    if pulse_width is not None:
        data["pulse_width"] = list(pulse_width)
    if pulse_intensity is not None:
        data["pulse_intensity"] = list(pulse_intensity)
    with open(path, "wb") as f:
        pickle.dump(data, f)
    return data


def load_identification_data(path):
    with open(path, "rb") as f:
        data = pickle.load(f)
    for key in ("time", "force", "stim_time"):
        if key not in data:
            raise KeyError(f"{path}: identification pickles need the key '{key}'")
    return data


def batch_targets(paths, n_shooting, final_time):
    """Resamples several recorded trains (same protocol, e.g. one per subject) at the node grid and stacks them as
    the ``target [ns+1, B]`` tensor layout of ``fes.identify``.  Returns (target ndarray, stim_time of the first)."""
    cols, stim = [], None
    for p in paths:
        d = load_identification_data(p)
        t0 = d["stim_time"][0]
        time = np.ravel(np.asarray(d["time"], dtype=np.float64)) - (t0 if t0 != 0 else 0.0)
        cols.append(np.interp(np.linspace(0, final_time, n_shooting + 1), time, np.ravel(np.asarray(d["force"], dtype=np.float64))))
        if stim is None:
            stim = [s - t0 for s in d["stim_time"]]
    return np.stack(cols, axis=1), stim


class SolutionToPickle:
    """result/pickle.py:6-41 for results of the GPU path.

    solution: dict with ``time`` [npts], ``states`` {key: [1, npts] | [npts]}, optional ``control`` {key: array},
    ``parameters`` {key: value}, ``parameters_bounds`` {key: (min, max)}, ``time_to_optimize``."""

    def __init__(self, solution: dict, file_name: str, path: str):
        self.sol, self.file_name, self.path = solution, file_name, path

    def remove_duplicates(self, time):
        time = np.asarray(time).reshape(-1)
        _, idx = np.unique(time, return_index=True)
        states = {}
        for key, val in self.sol["states"].items():
            val = np.asarray(val)
            states[key] = val[0][idx] if (val.ndim == 2 and val.shape[0] == 1) else (val[idx] if val.ndim == 1 else val[:, idx])
        return time[idx], states

    def pickle(self):
        time, states = self.remove_duplicates(self.sol["time"])
        dictionary = {
            "time": time, "states": states, "control": self.sol.get("control"),
            "parameters": self.sol.get("parameters"), "parameters_bounds": self.sol.get("parameters_bounds", {}),
            "time_to_optimize": self.sol.get("time_to_optimize"), "bio_model_path": self.sol.get("bio_model_path"),
        }
        with open(self.path + self.file_name, "wb") as file:
            pickle.dump(dictionary, file)
        return print(f"Solution values has been exported in pickle format in {self.path + self.file_name}")
