"""Multi-GPU sharding of the audit path: one process per GPU (any launcher, e.g. torchrun), the one
collective - an all-gather of roots / sub-roots - issued by the library's own NCCL communicator
(``ahv_comm_*`` in the C ABI, ``agent_hypervisor_b200.comm``).

Two schemes:

* **Sessions** are independent trees, so ranks own contiguous session ranges and nothing is
  exchanged until the results are collected: one all-gather of 32-byte roots and 1-byte verdicts.
* **One giant tree** (a single chain of n leaves) is cut at multiples of 2^k leaves.  Every full
  sub-tree has exactly 2^k leaves, so no odd-node duplication happens inside it and its root is
  the level-k node of the global tree.  Only the ragged tail is special: its nodes are always the
  last of their level, and the global level width there is odd, so the reference's rule
  (audit/delta.py:129) duplicates it at every level up to k ("lifting").  The ceil(n/2^k) sub-roots
  are all-gathered and every rank finishes the top of the tree.

Hashing is done by the callable passed in (default: the GPU context's ``merkle_roots``) and the exchange
by ``gather_fn(local[cap, width]) -> [world, cap, width]`` (default: the context's NCCL all-gather), so
the planning and reassembly logic can be exercised on CPU with any other transport (the tests use a
world-size-2/3 gloo group).
"""

from __future__ import annotations

from typing import Callable, Optional

import numpy as np


def shard_range(n_items: int, world: int, rank: int) -> tuple[int, int]:
    """Contiguous, balanced [lo, hi) of n_items for this rank (first n_items % world ranks get one more)."""
    base, extra = divmod(n_items, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def subtree_log2(n_leaves: int, world: int, min_subtrees_per_rank: int = 8, max_log2: int = 24) -> int:
    """k such that there are at least world * min_subtrees_per_rank sub-trees of 2^k leaves (k >= 0)."""
    k = max_log2
    while k > 0 and (n_leaves + (1 << k) - 1) >> k < world * min_subtrees_per_rank:
        k -= 1
    return k


def plan_tree(n_leaves: int, world: int, k: Optional[int] = None) -> dict:
    """Sub-tree layout of one n-leaf tree: k, number of sub-trees, and each rank's sub-tree range."""
    if k is None:
        k = subtree_log2(n_leaves, world)
    n_sub = (n_leaves + (1 << k) - 1) >> k
    ranges = [shard_range(n_sub, world, r) for r in range(world)]
    return {"k": k, "n_sub": n_sub, "ranges": ranges, "n_leaves": n_leaves}


def _default_merkle() -> Callable:
    from ._native import get_context
    return get_context().merkle_roots


def local_subroots(digests: np.ndarray, first_leaf: int, plan: dict, rank: int,
                   merkle_fn: Optional[Callable] = None) -> np.ndarray:
    """Level-k nodes of this rank's sub-trees.  ``digests`` are the rank's leaves, leaf ``first_leaf``
    (= range start * 2^k) first.  Returns uint8[n_local_sub, 32]."""
    merkle_fn = merkle_fn or _default_merkle()
    k, n = plan["k"], plan["n_leaves"]
    lo, hi = plan["ranges"][rank]
    if hi <= lo:
        return np.zeros((0, 32), np.uint8)
    size = 1 << k
    assert first_leaf == lo * size
    n_local = min(hi * size, n) - first_leaf
    digests = np.ascontiguousarray(digests, dtype=np.uint8).reshape(-1, 32)[:n_local]
    rowptr = np.minimum(np.arange(hi - lo + 1, dtype=np.uint64) * np.uint64(size), np.uint64(n_local))
    roots, _ = merkle_fn(digests, rowptr)
    roots = np.array(roots, dtype=np.uint8).reshape(-1, 32)
    # the ragged tail (only the globally last sub-tree can be short): lift it to level k
    tail = int(rowptr[-1] - rowptr[-2])
    if tail < size and plan["n_sub"] > 1:
        depth = 0
        while (1 << depth) < tail:
            depth += 1
        x = roots[-1:].copy()
        for _ in range(k - depth):
            pair = np.concatenate([x, x])                      # H(hex(x) + hex(x)), audit/delta.py:129-130
            x, _ = merkle_fn(pair, np.array([0, 2], dtype=np.uint64))
            x = np.array(x, dtype=np.uint8).reshape(1, 32)
        roots[-1] = x[0]
    return roots


def finish_tree(subroots: np.ndarray, merkle_fn: Optional[Callable] = None) -> Optional[bytes]:
    """Top of the tree over all sub-roots in order (the ordinary rule of audit/delta.py:125-134)."""
    merkle_fn = merkle_fn or _default_merkle()
    subroots = np.ascontiguousarray(subroots, dtype=np.uint8).reshape(-1, 32)
    if len(subroots) == 0:
        return None
    roots, _ = merkle_fn(subroots, np.array([0, len(subroots)], dtype=np.uint64))
    return bytes(np.array(roots, dtype=np.uint8).reshape(-1, 32)[0])


# ---------------------------------------------------------------------------- collectives
def _default_gather() -> Callable:
    from ._native import get_context
    return get_context().comm_all_gather


def all_gather_bytes(local: np.ndarray, counts: list[int], gather_fn: Optional[Callable] = None) -> np.ndarray:
    """All-gather rows of a uint8 [rows, width] array whose row counts per rank are ``counts``
    (known to every rank from the plan).  One collective; padded to the largest count."""
    gather_fn = gather_fn or _default_gather()
    world = len(counts)
    width = local.shape[1] if local.ndim == 2 else 1
    cap = max(counts) if counts else 0
    buf = np.zeros((cap, width), dtype=np.uint8)
    if len(local):
        buf[: len(local)] = np.ascontiguousarray(local, dtype=np.uint8).reshape(len(local), width)
    out = np.asarray(gather_fn(buf), dtype=np.uint8).reshape(world, cap, width)
    return np.concatenate([out[r, : counts[r]] for r in range(world)], axis=0)


def sharded_tree_root(local_leaf_digests: np.ndarray, n_leaves: int, rank: int, world: int,
                      k: Optional[int] = None, merkle_fn: Optional[Callable] = None, gather_fn: Optional[Callable] = None):
    """Root of one n-leaf tree whose leaves are spread over the ranks by ``plan_tree``."""
    plan = plan_tree(n_leaves, world, k)
    lo, hi = plan["ranges"][rank]
    sub = local_subroots(local_leaf_digests, lo << plan["k"], plan, rank, merkle_fn)
    counts = [h - l for (l, h) in plan["ranges"]]
    allsub = all_gather_bytes(sub, counts, gather_fn) if world > 1 else sub
    return finish_tree(allsub, merkle_fn), plan


def gather_session_results(local_roots: np.ndarray, local_verdicts: np.ndarray, n_sessions: int, rank: int,
                           world: int, gather_fn: Optional[Callable] = None):
    """Roots [S,32] and verdicts [S] of all sessions, sessions sharded by ``shard_range``: ONE all-gather of
    33 bytes per session (host arrays; the device-resident form is ``Context.gather_results_dev``)."""
    counts = [shard_range(n_sessions, world, r)[1] - shard_range(n_sessions, world, r)[0] for r in range(world)]
    if world == 1:
        return local_roots, local_verdicts
    packed = np.concatenate([np.ascontiguousarray(local_roots, dtype=np.uint8).reshape(-1, 32),
                             np.ascontiguousarray(local_verdicts, dtype=np.uint8).reshape(-1, 1)], axis=1)
    allp = all_gather_bytes(packed, counts, gather_fn)
    return allp[:, :32].copy(), allp[:, 32].copy()


def unpack_gathered(all_blocks: np.ndarray, n_sessions_per_rank: int, world: int):
    """Split the buffer ``ahv_gather_results_dev`` fills (copied to the host) into roots [world*S, 32] and
    verdicts [world*S]: rank r's block is S*32 root bytes, then S verdict bytes, padded to 16 bytes."""
    S = n_sessions_per_rank
    stride = (33 * S + 15) & ~15
    blocks = np.asarray(all_blocks, dtype=np.uint8).reshape(world, stride)
    roots = blocks[:, : 32 * S].reshape(world * S, 32).copy()
    verdicts = blocks[:, 32 * S: 33 * S].reshape(world * S).copy()
    return roots, verdicts


# ---------------------------------------------------------------------------- liability batches
def shard_slash_batch(b: dict, rank: int, world: int) -> tuple[dict, dict]:
    """This rank's part of a slashing batch (the arrays of ``ahv_slash_batch``; SURVEY.md 8e: shard by
    session, no data-path collective).  Sessions own contiguous agent rows, records and events, so a
    contiguous session range is a contiguous slice of every array; indices are rebased to the slice.
    Returns (local batch, placement) where placement says where the rank's results sit in the whole."""
    S = len(b["session_agent_rowptr"]) - 1
    s0, s1 = shard_range(S, world, rank)
    srp = np.asarray(b["session_agent_rowptr"], dtype=np.uint64)
    erp = np.asarray(b["event_rowptr"], dtype=np.uint64)
    rp = np.asarray(b["rowptr"], dtype=np.uint32)
    a0, a1 = int(srp[s0]), int(srp[s1])
    e0, e1 = int(rp[a0]), int(rp[a1])
    v0, v1 = int(erp[s0]), int(erp[s1])
    local = dict(
        rowptr=(rp[a0:a1 + 1] - np.uint32(e0)).astype(np.uint32),
        voucher=(np.asarray(b["voucher"], dtype=np.uint32)[e0:e1] - np.uint32(a0)).astype(np.uint32),
        active=np.asarray(b["active"], dtype=np.uint8)[e0:e1], expiry=np.asarray(b["expiry"], dtype=np.int64)[e0:e1],
        now=b["now"], session_agent_rowptr=srp[s0:s1 + 1] - np.uint64(a0), event_rowptr=erp[s0:s1 + 1] - np.uint64(v0),
        event_agent=(np.asarray(b["event_agent"], dtype=np.uint32)[v0:v1] - np.uint32(a0)).astype(np.uint32),
        event_sigma=np.asarray(b["event_sigma"], dtype=np.float64)[v0:v1],
        event_omega=np.asarray(b["event_omega"], dtype=np.float64)[v0:v1],
        scores=np.asarray(b["scores"], dtype=np.float64)[a0:a1])
    if b.get("event_depth") is not None:
        local["event_depth"] = np.asarray(b["event_depth"], dtype=np.uint8)[v0:v1]
    return local, dict(sessions=(s0, s1), agents=(a0, a1), records=(e0, e1), events=(v0, v1))


def place_slash_results(whole: dict, out: dict, placement: dict) -> None:
    """Write one rank's outputs (dict as returned by ``Context.slash_cascade``) into whole-batch arrays
    (same keys), turning the rank-local indices of the records back into whole-batch ones."""
    (s0, s1), (a0, a1), (e0, e1), (v0, v1) = (placement[k] for k in ("sessions", "agents", "records", "events"))
    whole["scores"][a0:a1] = out["scores"]
    whole["touched"][a0:a1] = out["touched"]
    whole["active"][e0:e1] = out["active"]
    whole["slash_count"][s0:s1] = out["slash_count"]
    whole["clip_count"][s0:s1] = out["clip_count"]
    whole["status"][s0:s1] = out["status"]
    sl = np.array(out["slash_log"], copy=True)
    sl["agent"] += np.uint32(a0)
    sl["event"] += np.uint32(v0)
    cl = np.array(out["clip_log"], copy=True)
    cl["edge"] += np.uint32(e0)
    # a session's slash records start at event_rowptr[s] + rowptr[a_lo]: the rank's block moves by v0 + e0
    whole["slash_log"][v0 + e0: v0 + e0 + len(sl)] = sl
    whole["clip_log"][e0: e0 + len(cl)] = cl
