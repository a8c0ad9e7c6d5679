#!/usr/bin/env python3
"""Sweep of equivalent scaled-sum formulations (sha256_compress_search) - development tool."""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from agent_hypervisor_b200 import Context
ctx = Context(0)
res = []
v, _ = ctx.int_peak(15, 3000)
print(f"current production form (mode 15): {v/1400/1e9:.2f} Gcompr/s", flush=True)
for A in range(4):
    for B in range(2):
        for C in range(2):
            for D in range(3):
                mode = 300 + ((A * 2 + B) * 2 + C) * 3 + D
                best = max(ctx.int_peak(mode, 2000)[0] for _ in range(2))
                res.append((best / 1400 / 1e9, A, B, C, D))
res.sort(reverse=True)
for g, A, B, C, D in res:
    print(f"A={A} B={B} C={C} D={D}: {g:.2f} Gcompr/s", flush=True)
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(res, open(os.path.join(ROOT, "gpurun_out", "lab4.json"), "w"))
