#!/usr/bin/env python3
"""Chained capture (ahv_chain_capture_dev) on device-resident config-2 sessions: time per launch for a few
session counts (development / measurement tool)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, workloads
from agent_hypervisor_b200 import Context
ctx = Context(0)
for S, L in ((10000, 1000), (100000, 100), (1000, 1000)):
    n = S * L
    a = ctx.host_alloc(n * 256)
    offs, hexpos, rp = workloads.config2_fill(a, 0, S, L)
    d = {k: ctx.dev_alloc(v) for k, v in dict(arena=n * 256 + 16, offs=offs.nbytes, hp=hexpos.nbytes, rp=rp.nbytes,
                                              digs=32 * n, par=32 * n, kind=n).items()}
    ctx.h2d(d["offs"], offs); ctx.h2d(d["hp"], hexpos); ctx.h2d(d["rp"], rp)
    for lanes in (0, 32, 24, 17, 16, 8, 4, 2, 1):                       # sessions per warp (0 = the library's choice)
        ctx.set_option("chain_lanes", lanes)
        best = 1e9
        for rep in range(3):
            ctx.h2d(d["arena"], a); ctx.sync()
            ctx.timer_start(1)
            ctx.chain_capture_dev(d["arena"], n * 256, d["offs"], d["hp"], d["rp"], S, n, d["digs"], d["par"], d["kind"])
            ctx.timer_stop(1); best = min(best, ctx.timer_ms(1))
        print(f"S={S} L={L} lanes/warp={lanes or 'auto'}: chain capture {best:.2f} ms ({n / best / 1e3:.1f} M deltas/s); "
              f"parent field at byte {int(hexpos[1])} of 256", flush=True)
    ctx.set_option("chain_lanes", 0)
    for p in d.values(): ctx.dev_free(p)
    ctx.host_free(a)
