#!/usr/bin/env python
"""fp64-pipe instruction count of the RK4 path (largest straight-line segment of the innermost arithmetic loop) of
any kernel of an object file.  Usage: sass_path.py build/fes_inst_k2f0.o <mangled-symbol-substring>"""
import re
import subprocess
import sys
from collections import Counter

obj, pat = sys.argv[1], sys.argv[2]
syms = subprocess.run(["cuobjdump", "-elf", obj], capture_output=True, text=True).stdout
names = sorted(set(re.findall(r"\.text\.(_Z\w+)", syms)))
fp = re.compile(r"\b(DFMA|DMUL|DADD|FFMA|FMUL|FADD|MUFU\.\w+)")
for sym in names:
    if pat not in sym:
        continue
    sass = subprocess.run(["cuobjdump", "-sass", "-fun", sym, obj], capture_output=True, text=True).stdout
    segs, cur = [], []
    for line in sass.splitlines():
        m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", line)
        if not m:
            continue
        cur.append(m.group(2).strip())
        if "BRA" in cur[-1] or "EXIT" in cur[-1] or "RET" in cur[-1]:
            segs.append(cur)
            cur = []
    best = max(segs, key=lambda seg: sum(1 for t in seg if fp.search(t)))
    c = Counter(fp.search(t).group(1) for t in best if fp.search(t))
    print(sym[:60], len(best), dict(c), "fp:", sum(v for k, v in c.items() if not k.startswith("MUFU")))
