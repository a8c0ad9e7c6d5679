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


def _roots_hex(roots, empty, S: int) -> list:
    """[S] hex roots (None for an empty session): one hex conversion for the whole chunk, then slices."""
    h = np.ascontiguousarray(roots).reshape(-1)[: 32 * S].tobytes().hex()
    e = np.asarray(empty)[:S].astype(bool).tolist()
    return [None if e[s] else h[64 * s: 64 * s + 64] for s in range(S)]


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
                       roots=_roots_hex(o["roots"], o["empty"], S),
                       root_match=(o["root_match"].astype(bool) & ~o["empty"].astype(bool)) if exp is not None else None)
            if copy:
                res["digests"] = np.array(res["digests"]); res["first_bad"] = np.array(res["first_bad"])
            yield res
            k += 1


class StagedColumns:
    """One chunk of columns already laid out in pinned host memory (``packing.stage_columns``), ready to be
    uploaded as it is: what an archive reader that fills pinned buffers directly would hand to ``DeviceReplay``.
    Made by ``DeviceReplay.prestage``; freed with ``close``."""

    def __init__(self, ctx: nat.Context, cols: dict, session_rowptr, committed_roots=None) -> None:
        from .packing import column_bytes, stage_columns
        self.ctx = ctx
        self.n = int(cols["n"])
        rp = np.asarray(session_rowptr, dtype=np.uint64)
        self.S = len(rp) - 1
        if int(rp[-1]) != self.n:
            raise ValueError("session_rowptr does not end at n")
        self.buf = ctx.host_alloc(column_bytes(cols))
        self.layout, self.used = stage_columns(cols, self.buf)
        self.rp = ctx.host_alloc(8 * (self.S + 1), np.uint64)
        self.rp[:] = rp
        self.maxl = max(int(np.diff(rp.astype(np.int64)).max()) if self.S else 0, 1)
        self.roots = None
        if committed_roots is not None:
            self.roots = ctx.host_alloc(32 * max(self.S, 1))
            self.roots[: 32 * self.S] = np.ascontiguousarray(committed_roots, dtype=np.uint8).reshape(-1)

    def close(self) -> None:
        for a in (self.buf, self.rp, self.roots):
            if a is not None:
                self.ctx.host_free(a)
        self.buf = self.rp = self.roots = None


class DeviceReplay:
    """Replay with the canonical packer ON THE DEVICE (``ahv_pack_deltas_dev``): a chunk's COLUMNS are staged in one
    pinned buffer, uploaded in one copy on the context's second stream (overlapping the previous chunk's kernels),
    serialised by one thread per delta straight into the arena the hashing kernels read, then committed and
    verified (``ahv_extract_parents_dev`` -> ``ahv_commit_dev`` -> ``ahv_verify_chains_dev``), and only digests,
    roots and verdicts travel back.  No host core serialises anything; the upload is the fields (< 200 bytes per
    delta) instead of the pre-images (~330 bytes).  Same results as ``ReplayPipeline``.

    Sizes: ``arena_bytes`` of pre-images, ``max_deltas`` deltas, ``max_sessions`` sessions and ``column_bytes``
    of staged columns (``packing.column_bytes``) for the largest chunk."""

    def __init__(self, arena_bytes: int, max_deltas: int, max_sessions: int, column_bytes: int,
                 ctx: Optional[nat.Context] = None) -> None:
        self.ctx = ctx or nat.get_context()
        c = self.ctx
        n, S = max(int(max_deltas), 1), max(int(max_sessions), 1)
        self.limits = (int(arena_bytes), n, S, int(column_bytes))
        self._stage = [c.host_alloc(int(column_bytes)) for _ in range(2)]
        self._d_cols = [c.dev_alloc(int(column_bytes)) for _ in range(2)]
        self._h_rp = [c.host_alloc(8 * (S + 1), np.uint64) for _ in range(2)]
        self._h_exp = [c.host_alloc(32 * S) for _ in range(2)]
        self._d_rp = [c.dev_alloc(8 * (S + 1)) for _ in range(2)]
        self._d_exp = [c.dev_alloc(32 * S) for _ in range(2)]
        sizes = dict(arena=int(arena_bytes) + 64, offs=8 * (n + 1), ppos=4 * n, par=32 * n, kind=n, digs=32 * n, roots=32 * S,
                     roots_v=32 * S, empty=S, verdict=S, fb=8 * S, match=S)
        self._d = {k: c.dev_alloc(v) for k, v in sizes.items()}
        self._out = [dict(digests=c.host_alloc(32 * n), roots=c.host_alloc(32 * S), empty=c.host_alloc(S), verdict=c.host_alloc(S),
                          first_bad=c.host_alloc(8 * S, np.int64), root_match=c.host_alloc(S)) for _ in range(2)]

    def close(self) -> None:
        c = self.ctx
        c.sync()
        for a in self._stage + self._h_rp + self._h_exp + [v for o in self._out for v in o.values()]:
            c.host_free(a)
        for p in self._d_cols + self._d_rp + self._d_exp + list(self._d.values()):
            c.dev_free(p)
        self._stage = []

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def prestage(self, cols: dict, session_rowptr, committed_roots=None) -> StagedColumns:
        return StagedColumns(self.ctx, cols, session_rowptr, committed_roots)

    def _stage_chunk(self, k: int, cols, session_rowptr=None, committed_roots=None):
        """Host side of chunk k: columns into the pinned staging buffer (unless they already are in one), upload
        started on the second stream."""
        from .packing import device_columns_struct, stage_columns
        c, b = self.ctx, k % 2
        if isinstance(cols, StagedColumns):
            sc = cols
            if sc.n > self.limits[1] or sc.S > self.limits[2] or sc.used > self.limits[3]:
                raise ValueError(f"chunk {k} exceeds the pipeline's buffers")
            c.upload_async(self._d_cols[b], sc.buf, sc.used)
            c.upload_async(self._d_rp[b], sc.rp, 8 * (sc.S + 1))
            if sc.roots is not None:
                c.upload_async(self._d_exp[b], sc.roots, 32 * sc.S)
            return dict(n=sc.n, S=sc.S, b=b, st=device_columns_struct(sc.n, sc.layout, self._d_cols[b]),
                        has_exp=sc.roots is not None, maxl=sc.maxl)
        n = int(cols["n"])
        rp = np.asarray(session_rowptr, dtype=np.uint64)
        S = len(rp) - 1
        if n > self.limits[1] or S > self.limits[2] or int(rp[-1]) != n:
            raise ValueError(f"chunk {k}: {n} deltas / {S} sessions exceed the pipeline's buffers "
                             f"({self.limits[1]} / {self.limits[2]}), or session_rowptr does not end at n")
        layout, used = stage_columns(cols, self._stage[b])
        self._h_rp[b][: S + 1] = rp
        c.upload_async(self._d_cols[b], self._stage[b], used)
        c.upload_async(self._d_rp[b], self._h_rp[b], 8 * (S + 1))
        has_exp = committed_roots is not None
        if has_exp:
            self._h_exp[b][: 32 * S] = np.ascontiguousarray(committed_roots, dtype=np.uint8).reshape(-1)
            c.upload_async(self._d_exp[b], self._h_exp[b], 32 * S)
        maxl = int(np.diff(rp.astype(np.int64)).max()) if S else 0
        return dict(n=n, S=S, b=b, st=device_columns_struct(n, layout, self._d_cols[b]), has_exp=has_exp, maxl=max(maxl, 1))

    def _enqueue(self, ch: dict) -> None:
        """Device side of a staged chunk: pack, parents, commit, verify, results back - all asynchronous."""
        c, d, b, n, S = self.ctx, self._d, ch["b"], ch["n"], ch["S"]
        o = self._out[b]
        c.upload_fence()                                           # the kernels below wait for this chunk's uploads
        if n == 0:
            return
        cap = self.limits[0]
        c.pack_deltas_dev(ch["st"], d["offs"], d["arena"], cap, d["ppos"])
        c.extract_parents_dev(d["arena"], cap, d["offs"], 0, d["ppos"], n, d["par"], d["kind"])
        c.commit_dev(d["arena"], cap, d["offs"], 0, self._d_rp[b], S, n, 0, ch["maxl"], d["digs"], d["roots"], d["empty"])
        c.verify_chains_dev(d["arena"], cap, d["offs"], 0, d["par"], d["kind"], self._d_rp[b], S, n, 0, ch["maxl"],
                            self._d_exp[b] if ch["has_exp"] else d["roots"], d["verdict"], d["fb"], 0, d["roots_v"], d["match"])
        lib, h, vp = c.lib, c.h, nat._vp
        for key, ptr, nb in (("digests", d["digs"], 32 * n), ("roots", d["roots"], 32 * S), ("empty", d["empty"], S),
                             ("verdict", d["verdict"], S), ("first_bad", d["fb"], 8 * S), ("root_match", d["match"], S)):
            c._check(lib.ahv_memcpy_d2h(h, vp(o[key].ctypes.data), vp(ptr), nb), "ahv_memcpy_d2h")

    # completion of a chunk = a CUDA event on the main stream (two of the context's timer slots, one per buffer set)
    _EVENT_SLOT = 60

    def _mark(self, ch: dict) -> None:
        self.ctx.timer_stop(self._EVENT_SLOT + ch["b"])

    def _wait(self, ch: dict) -> None:
        c = self.ctx
        c._check(c.lib.ahv_timer_wait(c.h, self._EVENT_SLOT + ch["b"]), "ahv_timer_wait")

    def run(self, chunks: Iterable[tuple], copy: bool = True) -> Iterator[dict]:
        """chunks: iterable of ``(cols, session_rowptr)`` / ``(cols, session_rowptr, committed_roots)`` as for
        ``ReplayPipeline.run``, or of ``StagedColumns`` (columns already in pinned memory: no host copy at all);
        yields the same per-chunk dicts.  While the GPU works on chunk k the host stages chunk k+1, its upload runs on
        the second stream and its kernels are enqueued behind chunk k's, so the GPU never waits for the host between
        chunks.  With ``copy=False`` the arrays are views into pinned buffers, valid until the next chunk is requested."""
        it = (ch if isinstance(ch, tuple) else (ch,) for ch in chunks)
        try:
            cur = self._stage_chunk(0, *next(it))
        except StopIteration:
            return
        k = 0
        self._enqueue(cur)
        self._mark(cur)
        while cur is not None:
            # chunk k+1 is staged, uploaded (second stream) and ENQUEUED behind chunk k before chunk k is waited for:
            # the GPU goes from one chunk's kernels straight to the next one's (the work buffers are reused in
            # stream order; the pinned result sets and the column buffers alternate)
            try:
                nxt = self._stage_chunk(k + 1, *next(it))
                self._enqueue(nxt)
                self._mark(nxt)
            except StopIteration:
                nxt = None
            self._wait(cur)                                        # chunk k's results are in its pinned output set
            n, S, o = cur["n"], cur["S"], self._out[cur["b"]]
            if n == 0:
                res = dict(digests=np.zeros((0, 32), np.uint8), first_bad=np.full(S, -1, np.int64), verdict=np.ones(S, bool),
                           roots=[None] * S, root_match=np.zeros(S, bool) if cur["has_exp"] else None)
            else:
                empty = o["empty"][:S].astype(bool)
                roots = o["roots"][: 32 * S].reshape(S, 32)
                res = dict(digests=o["digests"][: 32 * n].reshape(n, 32), first_bad=o["first_bad"][:S],
                           verdict=o["verdict"][:S].astype(bool),
                           roots=_roots_hex(roots, empty, S),
                           root_match=(o["root_match"][:S].astype(bool) & ~empty) if cur["has_exp"] else None)
                if copy:
                    res["digests"] = np.array(res["digests"]); res["first_bad"] = np.array(res["first_bad"])
            yield res
            cur = nxt
            k += 1
