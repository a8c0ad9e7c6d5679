#!/usr/bin/env python3
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from agent_hypervisor_b200 import Context
ctx = Context(0)
for mode, nm in ((0, "ALU only"), (2, "compress native"), (12, "scaled rounds+sched"), (13, "scaled rounds, native sched"),
                 (14, "scaled rounds (K native), native sched"), (15, "scaled rounds (K native) + sched"),
                 (16, "scaled rounds (K const-bank), native sched"), (17, "scaled rounds (K const-bank) + sched"),
                 (21, "mode 15 + plain shifts as IMAD.HI (immediate)"), (22, "mode 15 with h+K+w as one IADD3"),
                 (23, "mode 15 + IMAD.HI shifts + IADD3 h+K+w"), (18, "rt-zero addend: rounds (K native) + sched"), (19, "rt-zero addend: rounds (K native), native sched")):
    best = 0
    for _ in range(3):
        v, ms = ctx.int_peak(mode, int(os.environ.get('LAB_ITERS', '3000')))
        best = max(best, v)
    print(f"mode {mode} {nm}: {best/1e12:.3f} Tinstr/s = {best/1400/1e9:.2f} Gcomp/s", flush=True)

v, ms = ctx.int_peak(20, 20000)
adds = 20000 * 16
info = ctx.device_info()
print(f"dependent DADD chain: {ms*1e-3/adds*1e9:.2f} ns per addition = {ms*1e-3/adds*info['sm_clock_khz']*1e3:.1f} cycles at {info['sm_clock_khz']/1e3:.0f} MHz", flush=True)
