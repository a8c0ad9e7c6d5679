"""N>1 host logic on CPU: world_size-2 (and 3) gloo process groups.  Hashing is done by the oracle
here (test infrastructure); on the GPU box the same helpers call the CUDA context."""

import os
import socket

import numpy as np
import pytest

from agent_hypervisor_b200 import sharding
from oracle import coracle as co


def test_shard_range_covers_everything():
    for n in (0, 1, 7, 10, 10000, 12345):
        for world in (1, 2, 3, 8):
            spans = [sharding.shard_range(n, world, r) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(spans[i][1] == spans[i + 1][0] for i in range(world - 1))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1


@pytest.mark.parametrize("n,k", [(1, 0), (2, 1), (37, 3), (64, 3), (65, 3), (1000, 5), (1001, 5), (4097, 6), (5000, 4)])
def test_subtree_plan_matches_whole_tree_single_process(n, k):
    """Sub-roots at 2^k-aligned cuts + lifted tail + top finish == the reference's whole-tree root."""
    rng = np.random.default_rng(n)
    digs = rng.integers(0, 256, (n, 32), dtype=np.uint8)
    want, _ = co.merkle_roots(digs, np.array([0, n], np.uint64))
    for world in (1, 2, 3, 8):
        plan = sharding.plan_tree(n, world, k)
        subs = []
        for r in range(world):
            lo, hi = plan["ranges"][r]
            first = lo << k
            subs.append(sharding.local_subroots(digs[first: min(hi << k, n)], first, plan, r, co.merkle_roots))
        root = sharding.finish_tree(np.concatenate(subs), co.merkle_roots)
        assert root == bytes(want[0]), (n, k, world)


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _gloo_gather(world):
    """The transport of the CPU tests: a gloo all-gather with the signature sharding's gather_fn has
    (on the GPU box the default is the library's own NCCL all-gather, Context.comm_all_gather)."""
    import torch
    import torch.distributed as dist

    def gather(local: np.ndarray) -> np.ndarray:
        t = torch.from_numpy(np.ascontiguousarray(local))
        out = [torch.empty_like(t) for _ in range(world)]
        dist.all_gather(out, t)
        return np.stack([o.numpy() for o in out])
    return gather


def _worker(rank, world, port, n, n_sessions, q):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    gather = _gloo_gather(world)
    try:
        rng = np.random.default_rng(99)
        digs = rng.integers(0, 256, (n, 32), dtype=np.uint8)            # every rank can regenerate the input
        plan = sharding.plan_tree(n, world)
        lo, hi = plan["ranges"][rank]
        mine = digs[lo << plan["k"]: min(hi << plan["k"], n)]
        root, _ = sharding.sharded_tree_root(mine, n, rank, world, merkle_fn=co.merkle_roots, gather_fn=gather)
        # session sharding: each rank "computes" roots/verdicts of its own sessions, then one gather
        sess_roots = rng.integers(0, 256, (n_sessions, 32), dtype=np.uint8)
        sess_verd = (rng.random(n_sessions) < 0.9).astype(np.uint8)
        s0, s1 = sharding.shard_range(n_sessions, world, rank)
        all_r, all_v = sharding.gather_session_results(sess_roots[s0:s1], sess_verd[s0:s1], n_sessions, rank, world,
                                                            gather_fn=gather)
        ok = np.array_equal(all_r, sess_roots) and np.array_equal(all_v, sess_verd)
        q.put((rank, root, ok))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_gloo_sharded_tree_and_session_gather(world):
    import torch.multiprocessing as mp
    n, n_sessions = 10_007, 101
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n, n_sessions, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=180) for _ in procs]
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    rng = np.random.default_rng(99)
    digs = rng.integers(0, 256, (n, 32), dtype=np.uint8)
    want, _ = co.merkle_roots(digs, np.array([0, n], np.uint64))
    for rank, root, ok in res:
        assert root == bytes(want[0]) and ok


def test_slash_batch_shards_reassemble_to_the_whole():
    """Slashing shards by session with no exchange: every rank's slice run on its own (here by the C
    restatement) and placed back gives exactly the whole batch's scores, flags and records."""
    import recipes
    from helpers import same_f64, slash_batch
    from oracle.liability_oracle import CLIP_DT, SLASH_DT
    cases = [recipes.random_slash_case(seed) for seed in range(97)]
    b = slash_batch(cases)
    args = ("rowptr", "voucher", "active", "expiry", "now", "session_agent_rowptr", "event_rowptr", "event_agent",
            "event_sigma", "event_omega", "scores", "event_depth")
    want = co.slash_cascade(*[b[k] for k in args])
    keys = ("scores", "touched", "active", "slash_log", "clip_log", "slash_count", "clip_count")
    for world in (1, 2, 3, 8):
        A, E, S, NE = len(b["scores"]), len(b["voucher"]), len(cases), len(b["event_agent"])
        whole = dict(scores=np.zeros(A), touched=np.zeros(A, np.uint8), active=np.zeros(E, np.uint8),
                     slash_log=np.zeros(NE + E, SLASH_DT), clip_log=np.zeros(E, CLIP_DT), slash_count=np.zeros(S, np.uint32),
                     clip_count=np.zeros(S, np.uint32), status=np.ones(S, np.uint8))
        covered = 0
        for rank in range(world):
            local, place = sharding.shard_slash_batch(b, rank, world)
            out = dict(zip(keys, co.slash_cascade(*[local[k] for k in args])))
            out["status"] = np.zeros(len(out["slash_count"]), np.uint8)
            sharding.place_slash_results(whole, out, place)
            covered += place["sessions"][1] - place["sessions"][0]
        assert covered == S and not whole["status"].any()
        srp, erp, rp = b["session_agent_rowptr"], b["event_rowptr"], b["rowptr"]
        assert np.array_equal(np.nan_to_num(whole["scores"], nan=-1.0), np.nan_to_num(want[0], nan=-1.0))
        for k, w in zip(keys[1:3] + keys[5:], (want[1], want[2], want[5], want[6])):
            assert np.array_equal(whole[k], w), k
        for s in range(S):
            e0 = int(rp[int(srp[s])])
            for log, cnt, off in (("slash_log", want[5], int(erp[s]) + e0), ("clip_log", want[6], e0)):
                got = whole[log][off: off + int(cnt[s])]
                ref = want[3 if log == "slash_log" else 4][off: off + int(cnt[s])]
                for f in got.dtype.names:
                    same = same_f64(got[f], ref[f]) if got.dtype[f].kind == "f" else np.array_equal(got[f], ref[f])
                    assert same, (world, s, log, f)
