"""The library's own collective (csrc/ahv_comm.cu): result gathering in a world of one on the driver's single
GPU, and - where the box has two GPUs - two ranks started as separate processes whose NCCL id travels through
the launcher-side file exchange of agent_hypervisor_b200.comm."""

import json
import os
import subprocess
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_gather_results_world_of_one():
    from agent_hypervisor_b200 import get_context, sharding
    ctx = get_context()
    assert ctx.comm_info()["world"] == 1
    rng = np.random.default_rng(5)
    for S in (1, 7, 16, 1000, 10_001):
        roots = rng.integers(0, 256, (S, 32), dtype=np.uint8)
        verdict = (rng.random(S) < 0.8).astype(np.uint8)
        stride = ctx.gather_stride(S)
        assert stride % 16 == 0 and 33 * S <= stride < 33 * S + 16
        d_r, d_v, d_all = ctx.dev_alloc(32 * S), ctx.dev_alloc(S), ctx.dev_alloc(stride)
        try:
            ctx.h2d(d_r, roots); ctx.h2d(d_v, verdict); ctx.memset(d_all, 0xA5, stride)
            ctx.gather_results_dev(d_r, d_v, S, d_all)
            got = np.empty(stride, np.uint8)
            ctx.d2h(got, d_all); ctx.sync()
        finally:
            for p in (d_r, d_v, d_all):
                ctx.dev_free(p)
        r, v = sharding.unpack_gathered(got, S, 1)
        assert np.array_equal(r, roots) and np.array_equal(v, verdict) and not got[33 * S:].any()
    x = np.arange(12, dtype=np.float64).reshape(3, 4)
    assert np.array_equal(ctx.comm_all_gather(x), x[None])
    ctx.barrier()


_WORKER = r"""
import json, os, sys
import numpy as np
sys.path.insert(0, sys.argv[1])
from agent_hypervisor_b200 import Context, sharding
from agent_hypervisor_b200 import comm as ahv_comm
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
ctx = Context(int(os.environ["LOCAL_RANK"]))
ahv_comm.init_from_env(ctx, rank, world, tag=os.environ["AHV_TEST_TAG"])
S = 1234
rng = np.random.default_rng(100 + rank)
roots = rng.integers(0, 256, (S, 32), dtype=np.uint8)
verdict = (rng.random(S) < 0.7).astype(np.uint8)
stride = ctx.gather_stride(S)
d_r, d_v, d_all = ctx.dev_alloc(32 * S), ctx.dev_alloc(S), ctx.dev_alloc(world * stride)
ctx.h2d(d_r, roots); ctx.h2d(d_v, verdict)
for _ in range(3):
    ctx.gather_results_dev(d_r, d_v, S, d_all)
got = np.empty(world * stride, np.uint8)
ctx.d2h(got, d_all); ctx.sync()
r, v = sharding.unpack_gathered(got, S, world)
ok = True
for q in range(world):
    g = np.random.default_rng(100 + q)
    wr = g.integers(0, 256, (S, 32), dtype=np.uint8); wv = (g.random(S) < 0.7).astype(np.uint8)
    ok = ok and np.array_equal(r[q * S:(q + 1) * S], wr) and np.array_equal(v[q * S:(q + 1) * S], wv)
allr = ctx.comm_all_gather(np.array([rank, 10 * rank], np.int64))
ok = ok and allr.tolist() == [[q, 10 * q] for q in range(world)]
# host-side session gather through the same communicator
a_r, a_v = sharding.gather_session_results(roots, verdict, world * S, rank, world, gather_fn=ctx.comm_all_gather)
ok = ok and np.array_equal(a_r, r) and np.array_equal(a_v, v)
ctx.barrier()
print(json.dumps({"rank": rank, "ok": bool(ok), "comm": ctx.comm_info()}), flush=True)
ctx.comm_destroy()
"""


def test_two_ranks_nccl_in_the_library():
    import ctypes
    try:
        n = ctypes.c_int(0)
        cudart = ctypes.CDLL("libcudart.so")
        cudart.cudaGetDeviceCount(ctypes.byref(n))
        ndev = n.value
    except OSError:
        out = subprocess.run(["nvidia-smi", "-L"], capture_output=True, text=True).stdout
        ndev = out.count("GPU ")
    if ndev < 2:
        pytest.skip("needs two GPUs (run with gpurun --gpus 2)")
    tag = f"pytest_{os.getpid()}"
    procs = []
    for r in range(2):
        env = dict(os.environ, RANK=str(r), WORLD_SIZE="2", LOCAL_RANK=str(r), AHV_TEST_TAG=tag, MASTER_PORT="0")
        procs.append(subprocess.Popen([sys.executable, "-c", _WORKER, ROOT], env=env, stdout=subprocess.PIPE,
                                      stderr=subprocess.PIPE, text=True))
    outs = [p.communicate(timeout=300) for p in procs]
    for p, (o, e) in zip(procs, outs):
        assert p.returncode == 0, (o, e[-2000:])
        res = json.loads(o.strip().splitlines()[-1])
        assert res["ok"] and res["comm"]["world"] == 2 and res["comm"]["nccl_version"] > 0
