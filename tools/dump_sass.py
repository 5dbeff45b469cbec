#!/usr/bin/env python
"""Writes one SASS listing per kernel of libblfik_cuda.so into profiles/sass/ (instruction text only, no encodings),
plus an opcode histogram that shows the FP64 (DFMA/DMUL/DADD), FP64-tensor (DMMA) and shared-memory (LDS/STS) mix."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "bipedal-locomotion-framework_b200", "libblfik_cuda.so")
OUT = os.path.join(ROOT, "profiles", "sass")
KEEP = sys.argv[1:] or ["assemble_kernelILi18", "assemble_kernelILi32", "solve_kernelILi29ELi18", "solve_kernelILi29ELi32", "fast_kernelILi29ELi18",
                       "blfik_tick_kernel", "blfik_tick_eq_kernel", "blfik_kinematics_kernel", "blfik_integrate_kernel",
                       "blfik_integrate_acceleration_kernel", "blfik_dynamics_kernel", "blfik_bias_kernel",
                       "blfik_spline_coefficients_kernel", "blfik_cubic_spline_coefficients_kernel", "blfik_evaluate_setpoint_kernel"]


def main():
    os.makedirs(OUT, exist_ok=True)
    txt = subprocess.run(["cuobjdump", "-sass", SO], capture_output=True, text=True).stdout.splitlines()
    name, body = None, {}
    for ln in txt:
        m = re.search(r"Function : (\S+)", ln)
        if m:
            name = m.group(1)
            body[name] = []
            continue
        m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
        if m and name:
            body[name].append(f"/*{m.group(1)}*/ {m.group(2).strip()} ;")
    for fn, lines in body.items():
        if not any(k in fn for k in KEEP):
            continue
        short = re.sub(r"[^A-Za-z0-9_]", "_", subprocess.run(["c++filt", fn], capture_output=True, text=True).stdout.strip().split("(")[0])[:80]
        hist = collections.Counter(re.sub(r"\..*", "", l.split("*/ ")[1].split()[0] if not l.split("*/ ")[1].startswith("@") else l.split("*/ ")[1].split()[1]) for l in lines)
        with open(os.path.join(OUT, short + ".sass"), "w") as f:
            f.write(f"// {fn}\n// {len(lines)} instructions; opcode histogram: " + ", ".join(f"{k} {v}" for k, v in hist.most_common(24)) + "\n")
            f.write("\n".join(lines) + "\n")
        print(short, len(lines), {k: hist[k] for k in ("DFMA", "DMUL", "DADD", "DMMA", "LDS", "STS", "SHFL", "LDG", "MUFU") if k in hist})


if __name__ == "__main__":
    main()
