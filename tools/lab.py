#!/usr/bin/env python3
"""Integer-pipe lab: prints every ahv_int_peak mode (development tool)."""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from agent_hypervisor_b200 import Context
ctx = Context(0)
names = {0: "SHF+LOP3 (ALU only)", 1: "SHA rounds const KW x2 states", 2: "compress (compiler adds)", 3: "IMAD only",
         4: "SHF+LOP3+2 IMAD", 5: "IMAD.WIDE(+add)", 6: "IMAD.HI only", 7: "SHF+LOP3+IMAD.WIDE(+add)",
         8: "compress adds->IMAD", 9: "compress adds->IMAD, sigmas->WIDE/HI"}
out = {}
for mode in range(10):
    iters = 30000 if mode in (0, 3, 4, 5, 6, 7) else 3000
    best = 0
    for _ in range(3):
        v, ms = ctx.int_peak(mode, iters)
        best = max(best, v)
    out[mode] = best
    extra = f"  = {best/1400/1e9:.2f} Gcomp/s" if mode in (2, 8, 9) else ""
    print(f"mode {mode} {names[mode]:42s}: {best/1e12:7.3f} Tinstr/s{extra}", flush=True)
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "lab.json"), "w"))
