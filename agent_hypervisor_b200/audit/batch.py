"""Batched audit over many sessions: what ``Hypervisor.terminate_session`` (core.py:204) and an
auditor calling ``verify_chain`` + ``CommitmentEngine.verify`` do one session at a time in the
reference, done for all sessions in one GPU pass each.
"""

from __future__ import annotations

from typing import Optional, Sequence

import numpy as np

from .. import _native as nat
from .packing import PackedSessions, pack_sessions


class BatchAuditor:
    def __init__(self, ctx: Optional[nat.Context] = None) -> None:
        self.ctx = ctx or nat.get_context()

    def pack(self, engines: Sequence) -> PackedSessions:
        """engines: DeltaEngine objects (or lists of SemanticDelta)."""
        return pack_sessions([getattr(e, "_deltas", e) for e in engines])

    def commit(self, packed: PackedSessions):
        """-> (leaf digests [n,32], roots as hex or None per session)."""
        fixed, uni = packed.uniform_shape()
        digs, roots, empty = self.ctx.commit(
            packed.arena, None if fixed else packed.offsets, None if uni else packed.session_rowptr,
            fixed_len=fixed if fixed else 0, uniform_leaves=uni, n_sessions=packed.n_sessions)
        return digs, [None if empty[s] else roots[s].tobytes().hex() for s in range(packed.n_sessions)]

    def roots_of_engines(self, engines: Sequence) -> list:
        """``[e.compute_merkle_root() for e in engines]`` (audit/delta.py:117-134: trees over the STORED
        ``delta_hash`` strings, nothing re-hashed) with one tree launch for all engines whose hashes are
        well-formed digests; an engine holding any other string goes through its own method."""
        from .delta import _LOWER_HEX
        roots: list = [None] * len(engines)
        blobs, rowptr, where = [], [0], []
        for i, e in enumerate(engines):
            hs = [d.delta_hash for d in getattr(e, "_deltas", e)]
            joined = "".join(hs) if all(type(h) is str and len(h) == 64 for h in hs) else None
            if joined is not None and (not hs or _LOWER_HEX.fullmatch(joined)):
                blobs.append(joined)
                rowptr.append(rowptr[-1] + len(hs))
                where.append(i)
            else:
                roots[i] = e.compute_merkle_root()
        if where:
            digs = np.frombuffer(bytes.fromhex("".join(blobs)), dtype=np.uint8)
            r, empty = self.ctx.merkle_roots(digs, np.array(rowptr, dtype=np.uint64))
            for k, i in enumerate(where):
                roots[i] = None if empty[k] else r[k].tobytes().hex()
        return roots

    def commit_engines(self, engines: Sequence, commitment_engine=None, participants=None) -> list:
        """Roots for many DeltaEngines at once; also refreshes nothing (pure read), like
        compute_merkle_root.  Optionally records them in a CommitmentEngine."""
        packed = self.pack(engines)
        _, roots = self.commit(packed)
        if commitment_engine is not None:
            for e, r in zip(engines, roots):
                if r is not None:
                    commitment_engine.commit(e.session_id, r, list(participants or []), len(e._deltas))
        return roots

    @staticmethod
    def _roots_array(roots_hex: Sequence[Optional[str]], S: int):
        """hex roots (None / anything that is not a 64-hex digest = no usable commitment) -> ([S,32] bytes, known[S])."""
        if len(roots_hex) != S:
            raise ValueError(f"{len(roots_hex)} expected roots for {S} sessions")
        exp = np.zeros((S, 32), dtype=np.uint8)
        known = np.zeros(S, dtype=bool)
        for s, r in enumerate(roots_hex):
            if isinstance(r, str) and len(r) == 64:
                try:
                    exp[s] = np.frombuffer(bytes.fromhex(r), dtype=np.uint8)
                    known[s] = r == r.lower()            # hexdigest() is lowercase; any other string never matches
                except ValueError:
                    pass
        return exp, known

    def verify(self, packed: PackedSessions, expected_roots_hex: Optional[Sequence[Optional[str]]] = None):
        """-> dict(verdict bool[S], first_bad int[S], digests, root_match bool[S] or None)."""
        fixed, uni = packed.uniform_shape()
        exp = None
        if expected_roots_hex is not None:
            exp, known = self._roots_array(expected_roots_hex, packed.n_sessions)
        res = self.ctx.verify_chains(
            packed.arena, None if fixed else packed.offsets, packed.parents, packed.parent_kind,
            None if uni else packed.session_rowptr, fixed_len=fixed if fixed else 0, uniform_leaves=uni,
            n_sessions=packed.n_sessions, expected_roots=exp)
        if exp is not None:
            res["root_match"] = (res["root_match"].astype(bool)) & known   # commitment.py:61-63
        res["verdict"] = res["verdict"].astype(bool)
        return res

    def audit(self, packed: PackedSessions, committed_roots_hex: Optional[Sequence[Optional[str]]] = None):
        """Commit + verify on ONE upload (``ahv_audit_sessions``): what an auditor replaying stored sessions
        needs - every Merkle root (core.py:204-207), every verify_chain verdict (audit/delta.py:136-152) and,
        when the archive's stored commitments are given (``committed_roots_hex``, one per session, None where
        there is none), ``CommitmentEngine.verify`` of each rebuilt root against them (audit/commitment.py:59-64).
        The parents are read out of the pre-images on the device when the packer recorded their positions.
        -> dict(digests, roots hex|None per session, verdict bool[S], first_bad int[S], root_match bool[S] or
        None when no committed roots were given)."""
        fixed, uni = packed.uniform_shape()
        by_pos = packed.parent_pos is not None
        exp = known = None
        if committed_roots_hex is not None:
            exp, known = self._roots_array(committed_roots_hex, packed.n_sessions)
        o = self.ctx.audit_sessions(
            packed.arena, None if fixed else packed.offsets,
            None if by_pos else packed.parents, None if by_pos else packed.parent_kind,
            None if uni else packed.session_rowptr, fixed_len=fixed if fixed else 0, uniform_leaves=uni,
            n_sessions=packed.n_sessions, parent_pos=packed.parent_pos if by_pos else None, expected_roots=exp)
        o["roots"] = [None if o["empty"][s] else o["roots"][s].tobytes().hex() for s in range(packed.n_sessions)]
        o["verdict"] = o["verdict"].astype(bool)
        o["root_match"] = (o["root_match"].astype(bool) & known) if exp is not None else None
        return o

    def verify_engines(self, engines: Sequence) -> list[bool]:
        """verify_chain for many DeltaEngines, including the reference's delta_hash write-back."""
        packed = self.pack(engines)
        res = self.verify(packed)
        rp = packed.session_rowptr
        for s, e in enumerate(engines):
            deltas = getattr(e, "_deltas", e)
            last = len(deltas) - 1 if res["verdict"][s] else int(res["first_bad"][s])
            base = int(rp[s])
            for i in range(last + 1):
                deltas[i].delta_hash = res["digests"][base + i].tobytes().hex()
        return [bool(v) for v in res["verdict"]]
