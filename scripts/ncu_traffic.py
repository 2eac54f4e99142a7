#!/usr/bin/env python
"""Writes the DRAM traffic of one `ncu --set full` capture of the bench kernel to profiles/r02_ivp_x0_traffic.json together
with the sha256 of the device code (SASS dump) of the library it was captured from (cocofest_b200/csrc/sass_counts.json);
bench.py reports it as `roofline.traffic` only while the loaded library holds that device code.
Usage: ncu_traffic.py report.ncu-rep summary.tsv [batch] [final_time_s] [dtype]"""
import csv
import io
import json
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
OUT = ROOT / "profiles" / "r02_ivp_x0_traffic.json"
rep, summary = sys.argv[1], sys.argv[2]
batch = int(sys.argv[3]) if len(sys.argv) > 3 else 1 << 20
final_time = int(sys.argv[4]) if len(sys.argv) > 4 else 300
dtype = sys.argv[5] if len(sys.argv) > 5 else "f64"
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
rows = list(csv.reader(io.StringIO(raw[raw.index('"ID"'):])))
hdr, units, vals = rows[0], rows[1], rows[2]


def get(name, scale=None):
    i = hdr.index(name)
    v = float(vals[i].replace(",", ""))
    u = units[i].lower()
    if scale == "bytes":
        v *= {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9, "tbyte": 1e12}[u]
    if scale == "seconds":
        v *= {"ns": 1e-9, "us": 1e-6, "ms": 1e-3, "s": 1.0, "second": 1.0, "nsecond": 1e-9, "usecond": 1e-6, "msecond": 1e-3}[u]
    return v


rd, wr = get("dram__bytes_read.sum", "bytes"), get("dram__bytes_write.sum", "bytes")
counts = json.loads((ROOT / "cocofest_b200" / "csrc" / "sass_counts.json").read_text())
sha = counts["sass_sha256"]
kernel = f"fes_ivp_kernel<{'double' if dtype == 'f64' else 'float'}, DING2003, fatigue, X0>"
ksha = counts["kernels"][kernel]["sass_sha256"]   # the device code of the captured kernel alone: what bench.py compares
d = {
    "report": f"{rep} (summary: {summary})", "batch": batch, "final_time_s": final_time, "dtype": dtype, "sass_sha256": sha, "kernel": kernel, "kernel_sass_sha256": ksha,
    "dram_bytes_read": int(rd), "dram_bytes_write": int(wr), "dram_bytes": int(rd + wr),
    "gpu_time_duration_s": get("gpu__time_duration.sum", "seconds"),
    "fp64_pipe_active_pct": get("sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active"),
    "registers": int(get("launch__registers_per_thread")),
    "waves_per_sm": round(get("launch__waves_per_multiprocessor"), 2),
}
OUT.write_text(json.dumps(d, indent=1) + "\n")
print(json.dumps(d))
