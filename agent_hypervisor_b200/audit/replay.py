"""Replay / import of stored sessions held as columns: pack on the host cores while the GPU audits.

What an auditor does to check an archive with the reference is, per session, rebuild every
``SemanticDelta``, call ``verify_chain`` (audit/delta.py:136-152: one ``json.dumps`` + SHA-256 per
delta) and compare ``compute_merkle_root`` (:117-134) with the committed root
(audit/commitment.py:59-64; pass the archive's stored roots with each chunk to have that compare done).  Here the archive arrives in chunks of columns; chunk k+1 is serialised
by the C++ packer (``ahv_pack_deltas``, all host threads) into one pinned buffer set while chunk k,
in the other set, goes through ``ahv_audit_sessions`` (upload, leaf hashes, trees, link check,
results back).  Both calls release the GIL, so one helper thread is enough.
"""

from __future__ import annotations

from concurrent.futures import ThreadPoolExecutor
from typing import Iterable, Iterator, Optional

import numpy as np

from .. import _native as nat
from .packing import pack_columns


class _BufferSet:
    def __init__(self, ctx: nat.Context, arena_bytes: int, max_deltas: int, max_sessions: int) -> None:
        n, S = max_deltas, max_sessions
        self.arena = ctx.host_alloc(arena_bytes)
        self.offsets = ctx.host_alloc(8 * (n + 1), np.uint64)
        self.parent_pos = ctx.host_alloc(4 * max(n, 1), np.uint32)
        self.rowptr = ctx.host_alloc(8 * (S + 1), np.uint64)
        self.out = dict(digests=ctx.host_alloc(32 * max(n, 1)), roots=ctx.host_alloc(32 * max(S, 1)),
                        empty=ctx.host_alloc(max(S, 1)), verdict=ctx.host_alloc(max(S, 1)),
                        first_bad=ctx.host_alloc(8 * max(S, 1), np.int64), root_match=ctx.host_alloc(max(S, 1)))

    def arrays(self):
        return [self.arena, self.offsets, self.parent_pos, self.rowptr, *self.out.values()]


class ReplayPipeline:
    """Two pinned buffer sets sized for the largest chunk: ``arena_bytes`` of pre-images, ``max_deltas``
    deltas, ``max_sessions`` sessions."""

    def __init__(self, arena_bytes: int, max_deltas: int, max_sessions: int, ctx: Optional[nat.Context] = None) -> None:
        self.ctx = ctx or nat.get_context()
        self.limits = (int(arena_bytes), int(max_deltas), int(max_sessions))
        self._sets = [_BufferSet(self.ctx, *self.limits) for _ in range(2)]
        self._pool = ThreadPoolExecutor(max_workers=1, thread_name_prefix="ahv-pack")

    def close(self) -> None:
        self._pool.shutdown(wait=True)
        for bs in self._sets:
            for a in bs.arrays():
                self.ctx.host_free(a)
        self._sets = []

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def _pack(self, k: int, cols: dict, session_rowptr, committed_roots=None) -> tuple:
        bs = self._sets[k % 2]
        n = int(cols["n"])
        rp = np.asarray(session_rowptr, dtype=np.uint64)
        S = len(rp) - 1
        if n > self.limits[1] or S > self.limits[2] or int(rp[-1]) != n:
            raise ValueError(f"chunk {k}: {n} deltas / {S} sessions exceed the pipeline's buffers "
                             f"({self.limits[1]} / {self.limits[2]}), or session_rowptr does not end at n")
        arena, offsets, ppos = pack_columns(cols, arena=bs.arena, offsets=bs.offsets, parent_pos=bs.parent_pos)
        bs.rowptr[: S + 1] = rp
        exp = None
        if committed_roots is not None:
            exp = np.ascontiguousarray(committed_roots, dtype=np.uint8).reshape(S, 32)
        return n, S, arena, offsets, ppos, exp

    def run(self, chunks: Iterable[tuple], copy: bool = True) -> Iterator[dict]:
        """chunks: iterable of ``(cols, session_rowptr)`` or ``(cols, session_rowptr, committed_roots)``
        (``cols`` as for ``pack_columns``; ``committed_roots`` = the archive's stored Merkle roots, uint8 [S,32]).
        Yields, per chunk, dict(digests [n,32], roots hex|None per session, verdict bool[S], first_bad int64[S],
        root_match bool[S] against the committed roots, or None when the chunk carried none).
        With ``copy=False`` the arrays are views into the pinned buffers, valid until the chunk after the
        next one is requested."""
        it = iter(chunks)
        k = 0
        try:
            nxt = next(it)
        except StopIteration:
            return
        self.ctx.set_option("blocking_sync", 1)     # the host cores belong to the packer while the GPU works
        try:
            yield from self._run(it, nxt, copy)
        finally:
            self.ctx.set_option("blocking_sync", 0)

    def _run(self, it, nxt, copy: bool) -> Iterator[dict]:
        k = 0
        fut = self._pool.submit(self._pack, k, *nxt)
        while fut is not None:
            n, S, arena, offsets, ppos, exp = fut.result()
            bs = self._sets[k % 2]
            try:
                nxt = next(it)
                fut = self._pool.submit(self._pack, k + 1, *nxt)      # packs while the GPU works on chunk k
            except StopIteration:
                fut = None
            out = {key: (v[: 32 * n].reshape(n, 32) if key == "digests" else
                         v[: 32 * S].reshape(S, 32) if key == "roots" else v[:S]) for key, v in bs.out.items()}
            o = self.ctx.audit_sessions(arena, offsets, None, None, bs.rowptr[: S + 1], parent_pos=ppos, out=out,
                                        expected_roots=exp)
            res = dict(digests=o["digests"], first_bad=o["first_bad"], verdict=o["verdict"].astype(bool),
                       roots=[None if o["empty"][s] else o["roots"][s].tobytes().hex() for s in range(S)],
                       root_match=(o["root_match"].astype(bool) & ~o["empty"].astype(bool)) if exp is not None else None)
            if copy:
                res["digests"] = np.array(res["digests"]); res["first_bad"] = np.array(res["first_bad"])
            yield res
            k += 1
