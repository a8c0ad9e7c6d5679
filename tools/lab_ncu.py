#!/usr/bin/env python3
"""Runs selected ahv_int_peak modes once each (short) so that ncu can capture them."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from agent_hypervisor_b200 import Context
ctx = Context(0)
for mode in (0, 3, 4, 1, 2, 8, 9):
    iters = 4000 if mode in (0, 3, 4) else 400
    v, ms = ctx.int_peak(mode, iters)
    print(mode, v / 1e12, ms, flush=True)
