#!/usr/bin/env python3
"""Add-site mask sweep for the SHA-256 compression (development tool)."""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from agent_hypervisor_b200 import Context
ctx = Context(0)
masks = ["000 native", "010 d", "070 d,S0+Maj,t1+t2", "07E round all but K", "380 sched", "3F0 sched+d,S0Maj,t1t2",
         "BFE all but K", "BFF all", "390 sched+d", "00E t1 chain", "FFE all but K + wide", "400 wide only",
         "060 S0Maj,t1t2", "3E0 sched+S0Maj,t1t2", "200 one sched", "180 two sched", "1000 shr->IMAD.HI", "1380 shr->HI + sched"]
for mode, nm in ((0, "ALU"), (3, "IMAD reg"), (4, "2ALU+2IMAD reg"), (10, "2ALU+2IMAD imm"), (11, "3ALU+1IMAD imm")):
    v, ms = ctx.int_peak(mode, 30000)
    print(f"mode {mode} {nm}: {v/1e12:.3f} Tinstr/s", flush=True)
out = {}
for base in (100,):
    for i, name in enumerate(masks):
        best = 0
        for _ in range(2):
            v, ms = ctx.int_peak(base + i, 2000)
            best = max(best, v)
        out[f"{base+i}"] = best
        print(f"mode {base+i} nst={base//100} mask {name:28s}: {best/1400/1e9:6.2f} Gcomp/s", flush=True)
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "lab2.json"), "w"))
