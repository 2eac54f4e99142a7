#!/usr/bin/env python
"""Counts the fp64 pipe instructions (DFMA / DMUL / DADD) and MUFU of the RK4 path in the shipped SASS of the bench
kernel fes_ivp_kernel<double, DING2003, fatigue, X0> and writes profiles/r01_ivp_x0_static.json (merging, so the ncu
traffic entry written by scripts/ncu_traffic.py survives).  The RK4 path is the largest straight-line segment of the
innermost loop (one uniform branch per step selects the scheme, fes_core.cuh rk_step)."""
import json
import re
import subprocess
import sys
from collections import Counter
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
OBJ = ROOT / "cocofest_b200" / "csrc" / "build" / "fes_inst_k0f1.o"
SYM = "_ZN3fes14fes_ivp_kernelIdLi0ELb1ELi1EEEvNS_7IvpArgsIT_EE"
OUT = ROOT / "profiles" / "r01_ivp_x0_static.json"

sass = subprocess.run(["cuobjdump", "-sass", "-fun", SYM, str(OBJ)], capture_output=True, text=True, check=True).stdout
ins = []
for line in sass.splitlines():
    m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", line)
    if m:
        ins.append((int(m.group(1), 16), m.group(2).strip()))
fp = re.compile(r"\b(DFMA|DMUL|DADD|MUFU\.\w+)")
inner = None
for addr, text in ins:  # innermost loop that holds the arithmetic = smallest backward-branch body with > 200 fp64 ops
    m = re.search(r"BRA(?:\.U)?\s+(?:!?U?P\d,\s*)?(0x[0-9a-f]+)", text)
    if m and int(m.group(1), 16) < addr:
        body = [x for x in ins if int(m.group(1), 16) <= x[0] <= addr]
        if sum(1 for x in body if fp.search(x[1])) > 200 and (inner is None or len(body) < len(inner)):
            inner = body
segments, cur = [], []
for addr, text in inner:
    cur.append(text)
    if "BRA" in text:
        segments.append(cur)
        cur = []
best = max(segments, key=lambda seg: sum(1 for t in seg if fp.search(t)))
count = Counter(fp.search(t).group(1) for t in best if fp.search(t))
regs = re.search(SYM + r".*?Used (\d+) registers", (OBJ.parent / "fes_inst_k0f1.ptxas.log").read_text(), re.S)
data = json.loads(OUT.read_text()) if OUT.exists() else {}
data.update({
    "kernel": "fes_ivp_kernel<double, DING2003, fatigue, X0>",
    "rk4_path_instructions": len(best), "rk4_path_mix": dict(count),
    "fp64_inst_per_rk4_step": count["DFMA"] + count["DMUL"] + count["DADD"],
    "flop_executed_per_rk4_step": 2 * count["DFMA"] + count["DMUL"] + count["DADD"],
    "registers": int(regs.group(1)) if regs else None,
})
OUT.write_text(json.dumps(data, indent=1) + "\n")
print(json.dumps(data))
