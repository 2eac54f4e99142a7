#!/usr/bin/env python
"""Static instruction counts of the RK4 path of EVERY kernel in the shipped library, tied to the library by its sha256.

    python scripts/sass_count.py [--lib cocofest_b200/csrc/libfes_b200.so] [--out cocofest_b200/csrc/sass_counts.json]

Run by cocofest_b200/csrc/Makefile right after the link, so the counts always describe the .so next to them; bench.py
recomputes the library's sha256 and refuses the counts when it differs (`sass_sha256`, the hash of the device code alone, is
what survives a re-link of the same sources and ties ncu captures to a build).  For every kernel: the RK4 path is the largest
straight-line segment (between branches) of the function by floating-point instruction count -- one uniform branch
per step selects the scheme (fes_core.cuh rk_step), so the whole RK4 step of a trajectory is ONE basic block -- and the
record holds its instruction mix (DFMA/DMUL/DADD, FFMA/FMUL/FADD and their packed f32x2 forms, MUFU), the function's
registers / stack frame (cuobjdump -res-usage) and its local-memory instructions (spills, LamCache).  Counts are per RK4
step: a path that holds 4 k reciprocals (four per step, one per stage) is k steps unrolled by the compiler.
"""
import argparse
import hashlib
import json
import re
import subprocess
from collections import Counter
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
FP = re.compile(r"\b(DFMA|DMUL|DADD|FFMA2|FMUL2|FADD2|FFMA|FMUL|FADD|MUFU\.\w+)\b")
KIND = {"0": "DING2003", "1": "DING2007", "2": "HMED2018"}
GROUP_BITS = ((1, "X0"), (2, "U"), (32, "C"), (4, "P"), (16, "PF"), (8, "W"))


def pretty(sym):
    """_ZN3fes14fes_ivp_kernelIdLi0ELb1ELi1EEEvNS_7IvpArgsIT_EE -> fes_ivp_kernel<double, DING2003, fatigue, X0>"""
    h = re.match(r"_ZN3fes\d+fes_hess_kernelILi(\d)ELb([01])ELb([01])EEE", sym)
    if h:   # fes_hess_kernel<KIND, FAT, FC>: the second-order tangent kernel (fp64 only)
        return f"fes_hess_kernel<{KIND.get(h.group(1), h.group(1))}, {'fatigue' if h.group(2) == '1' else 'no_fatigue'}, {'X0|U|C' if h.group(3) == '1' else 'X0|U'}>"
    m = re.match(r"_ZN3fes\d+(fes_\w+_kernel)I([df])(?:Li(\d)E)?(?:Lb([01])E)?(?:Li(\d+)E)?", sym)
    if not m:
        return sym
    name, t, kind, fat, groups = m.groups()
    parts = ["double" if t == "d" else "float"]
    if kind is not None:
        parts.append(KIND.get(kind, kind))
    if fat is not None and name == "fes_ident_kernel":   # fes_ident_kernel<T, KIND, STAGED>
        parts.append("staged" if fat == "1" else "strided")
    elif fat is not None:
        parts.append("fatigue" if fat == "1" else "no_fatigue")
    if groups is not None:
        g = int(groups)
        parts.append("|".join(n for b, n in GROUP_BITS if g & b) or "0")
    return f"{name}<{', '.join(parts)}>"


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--lib", default=str(ROOT / "cocofest_b200" / "csrc" / "libfes_b200.so"))
    ap.add_argument("--out", default=str(ROOT / "cocofest_b200" / "csrc" / "sass_counts.json"))
    args = ap.parse_args()
    lib = Path(args.lib)
    sha = hashlib.sha256(lib.read_bytes()).hexdigest()
    sass = subprocess.run(["cuobjdump", "-sass", str(lib)], capture_output=True, text=True, check=True).stdout
    # the file hash changes with every link (build id); the device code does not: two builds of the same sources give the
    # same SASS dump, which is what measurements taken with one build (ncu traffic) are tied to
    sass_sha = hashlib.sha256(sass.encode()).hexdigest()
    res = subprocess.run(["cuobjdump", "-res-usage", str(lib)], capture_output=True, text=True, check=True).stdout
    usage = {}
    for m in re.finditer(r"Function (\S+):\s*\n\s*REG:(\d+) STACK:(\d+) SHARED:(\d+) LOCAL:(\d+)", res):
        usage[m.group(1)] = {"registers": int(m.group(2)), "stack_bytes": int(m.group(3)), "local_bytes": int(m.group(5))}
    kernels = {}
    for block in sass.split("Function : ")[1:]:
        sym = block.split("\n", 1)[0].strip()
        segs, cur, ldl, stl = [], [], 0, 0
        for line in block.splitlines():
            m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(.*?);", line)
            if not m:
                continue
            text = m.group(1).strip()
            cur.append(text)
            ldl += bool(re.search(r"\bLDL\b", text))
            stl += bool(re.search(r"\bSTL\b", text))
            if re.search(r"\b(BRA|EXIT|RET|BRX|JMP)\b", text):
                segs.append(cur)
                cur = []
        if cur:
            segs.append(cur)
        best = max(segs, key=lambda seg: sum(1 for t in seg if FP.search(t)))
        mix = Counter(FP.search(t).group(1) for t in best if FP.search(t))
        # every RK4 step holds exactly four reciprocals (one per stage, fes_core.cuh nominal_rhs): a path with 8 of them
        # is two steps unrolled by the compiler (the small no-sensitivity kernels) -- counts below are PER STEP
        rcp = sum(v for k, v in mix.items() if k.startswith("MUFU.RCP"))
        # (the second-order kernel keeps the two reciprocals of F' apart: eight per step, fes_hess.cuh)
        steps = max(1, rcp // (8 if "fes_hess_kernel" in sym else 4))
        if steps > 1:
            mix = Counter({k: v // steps for k, v in mix.items()})
        fp64 = mix["DFMA"] + mix["DMUL"] + mix["DADD"]
        fp32 = mix["FFMA"] + mix["FMUL"] + mix["FADD"]
        fp32x2 = mix["FFMA2"] + mix["FMUL2"] + mix["FADD2"]
        rec = {
            "symbol": sym, "path_instructions": len(best) // steps, "steps_in_path": steps, "path_mix": dict(mix),
            "fp64_inst": fp64, "fp64_flop": 2 * mix["DFMA"] + mix["DMUL"] + mix["DADD"],
            "fp32_inst": fp32 + fp32x2,
            "fp32_flop": 2 * mix["FFMA"] + mix["FMUL"] + mix["FADD"] + 2 * (2 * mix["FFMA2"] + mix["FMUL2"] + mix["FADD2"]),
            "mufu": sum(v for k, v in mix.items() if k.startswith("MUFU")),
            "local_loads": ldl, "local_stores": stl,
            # the device code of THIS kernel alone: what a measurement of the kernel (ncu traffic) is tied to, so that adding
            # or changing another kernel of the library does not orphan it
            "sass_sha256": hashlib.sha256(block.encode()).hexdigest(),
        }
        rec.update(usage.get(sym, {}))
        kernels[pretty(sym)] = rec
    out = {"lib": lib.name, "lib_sha256": sha, "sass_sha256": sass_sha, "how": "scripts/sass_count.py: largest straight-line segment by fp instructions "
           "= the RK4 step of one trajectory (all tangent columns included)", "kernels": kernels}
    Path(args.out).write_text(json.dumps(out, indent=1, sort_keys=True) + "\n")
    print(f"{len(kernels)} kernels, lib sha256 {sha[:16]}.. -> {args.out}")


if __name__ == "__main__":
    main()
