#!/usr/bin/env python3
"""Occupancy sensitivity of the compression microbenchmarks (development tool)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from agent_hypervisor_b200 import Context
ctx = Context(0)
for ctas in (1, 2, 3, 4, 5, 6, 8):
    ctx.set_option("peak_ctas", ctas)
    row = []
    for mode in (2, 15, 21):
        v = max(ctx.int_peak(mode, 2000)[0] for _ in range(2))
        row.append(f"mode {mode}: {v/1400/1e9:6.2f}")
    print(f"{ctas} CTAs x 256 threads per SM ({ctas*2} warps/SMSP): " + "  ".join(row) + "  Gcompr/s", flush=True)
