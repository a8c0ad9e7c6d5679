#!/usr/bin/env python3
"""Where does a single-session compute_merkle_root / capture spend its time? (development tool)"""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from agent_hypervisor_b200 import get_context
ctx = get_context()
def med(fn, reps=200):
    for _ in range(10): fn()
    ts = []
    for _ in range(reps):
        t0 = time.perf_counter(); fn(); ts.append(time.perf_counter() - t0)
    ts.sort(); return ts[len(ts) // 2] * 1e6
rng = np.random.default_rng(0)
for n in (2, 10, 100, 1000):
    digs = rng.integers(0, 256, (n, 32), dtype=np.uint8); rp = np.array([0, n], np.uint64)
    hexes = [bytes(d).hex() for d in digs]
    t_all = med(lambda: ctx.merkle_roots(digs, rp))
    t_hex = med(lambda: np.frombuffer(bytes.fromhex("".join(hexes)), dtype=np.uint8).reshape(-1, 32))
    d_dig = ctx.dev_alloc(32 * n); d_rp = ctx.dev_alloc(16); d_r = ctx.dev_alloc(32); d_e = ctx.dev_alloc(1)
    ctx.h2d(d_dig, digs); ctx.h2d(d_rp, rp); ctx.sync()
    def dev():
        ctx.merkle_roots_dev(d_dig, d_rp, 1, n, n, d_r, d_e); ctx.sync()
    t_dev = med(dev)
    msg = rng.integers(0, 256, 300, dtype=np.uint8); off = np.array([0, 300], np.uint64)
    t_sha = med(lambda: ctx.sha256_batch(msg, off))
    print(f"n={n:5d}: merkle_roots host call {t_all:7.1f} us | device-only (launches+sync) {t_dev:7.1f} us | hex->bytes {t_hex:6.1f} us | one 300-B sha256_batch {t_sha:6.1f} us", flush=True)

# one capture: the batch entry (two copies + kernel) against ahv_sha256_one (message in the launch, digest via
# mapped host memory), and the whole DeltaEngine.capture call
from agent_hypervisor_b200 import DeltaEngine, VFSChange
m300 = bytes(rng.integers(0, 256, 300, dtype=np.uint8))
a300 = np.frombuffer(m300, np.uint8); o300 = np.array([0, 300], np.uint64)
t_batch = med(lambda: ctx.sha256_batch(a300, o300), 2000)
t_one = med(lambda: ctx.sha256_one(m300), 2000)
eng = DeltaEngine("bench-session")
ch = [VFSChange(path="/f", operation="add", content_hash="h")]
t_cap = med(lambda: eng.capture("did:mesh:a", ch), 2000)
import hashlib
t_hl = med(lambda: hashlib.sha256(m300).digest(), 2000)
print(f"one 300-B message: sha256_batch {t_batch:.1f} us | sha256_one {t_one:.1f} us | hashlib {t_hl:.2f} us | DeltaEngine.capture {t_cap:.1f} us", flush=True)
for comp in (0, 1):
    ctx.set_option("compress", comp)
    for ln in (64, 300, 1000, 3000):
        mm = bytes(rng.integers(0, 256, ln, dtype=np.uint8))
        print(f"compress={comp} sha256_one {ln:5d} B: {med(lambda: ctx.sha256_one(mm), 1000):.1f} us", flush=True)
ctx.set_option("compress", 1)
