"""ReplayPipeline (columns -> C++ packer on a helper thread -> ahv_audit_sessions, double-buffered):
same digests, roots and verdicts as the per-session reference semantics (oracle) on every chunk."""

from __future__ import annotations

import hashlib

import numpy as np
import pytest

import recipes
from oracle import audit_oracle as ao

pytestmark = pytest.mark.gpu


def _chunk(seeds, tamper=None):
    """Real SemanticDelta-shaped sessions (recipes.random_session) with properly chained parents."""
    from agent_hypervisor_b200 import SemanticDelta, VFSChange
    from agent_hypervisor_b200.audit.packing import canonical_preimage, columns_from_deltas
    deltas, rowptr, want_roots, want_verdict = [], [0], [], []
    for si, seed in enumerate(seeds):
        parent, hashes, ok = None, [], True
        for i, f in enumerate(recipes.random_session(seed)):
            d = SemanticDelta(delta_id=f["delta_id"], turn_id=f["turn_id"], session_id=f["session_id"], agent_did=f["agent_did"],
                              timestamp=f["timestamp"], changes=[VFSChange(**c) for c in f["changes"]], parent_hash=parent)
            if tamper == (si, i):
                d.parent_hash = "f" * 64               # a broken link (only visible when i > 0 or parent was None)
                ok = False
            h = hashlib.sha256(canonical_preimage(d)).hexdigest()
            hashes.append(h); parent = h
            deltas.append(d)
        rowptr.append(len(deltas))
        want_roots.append(ao.merkle_root_hex(hashes))             # audit/delta.py:117-134 on the hex strings
        want_verdict.append(ok)
    return columns_from_deltas(deltas), np.array(rowptr, np.uint64), deltas, want_roots, want_verdict


def test_pipeline_matches_per_chunk_results():
    from agent_hypervisor_b200 import BatchAuditor, ReplayPipeline
    from agent_hypervisor_b200.audit.packing import canonical_preimage, pack_sessions
    chunks = [_chunk(range(10 * k, 10 * k + 10), tamper=(3, 1) if k == 2 else None) for k in range(5)]
    chunks.append(_chunk([7]))                                   # a one-session chunk last
    n_max = max(int(c[0]["n"]) for c in chunks)
    bytes_max = max(sum(len(canonical_preimage(d)) for d in c[2]) for c in chunks)
    with ReplayPipeline(bytes_max + 64, n_max, 10) as pipe:
        results = list(pipe.run((c[0], c[1]) for c in chunks))
        assert len(results) == len(chunks)
        auditor = BatchAuditor()
        for (cols, rp, deltas, want_roots, want_verdict), got in zip(chunks, results):
            sessions = [deltas[int(rp[s]):int(rp[s + 1])] for s in range(len(rp) - 1)]
            ref = auditor.audit(pack_sessions(sessions))          # same library, Python packer, one call
            assert np.array_equal(got["digests"], ref["digests"]) and got["roots"] == ref["roots"]
            assert list(got["verdict"]) == list(ref["verdict"]) and np.array_equal(got["first_bad"], ref["first_bad"])
            # and the reference's own semantics, by hand
            for i, d in enumerate(deltas):
                assert got["digests"][i].tobytes().hex() == hashlib.sha256(canonical_preimage(d)).hexdigest()
            assert [bool(v) for v in got["verdict"]] == want_verdict and got["roots"] == want_roots
            assert got["root_match"] is None and ref["root_match"] is None      # no stored commitments were given
        # the archive's stored commitments travel with each chunk: session 4 of chunk 1 has a wrong one on file
        def with_roots(k, c):
            stored = np.stack([np.frombuffer(bytes.fromhex(r or "00" * 32), np.uint8) for r in c[3]]).copy()
            if k == 1:
                stored[4, 0] ^= 1
            return c[0], c[1], stored
        checked = list(pipe.run(with_roots(k, c) for k, c in enumerate(chunks)))
        for k, got in enumerate(checked):
            # an empty session has no root (None) and never matches a stored commitment (audit/commitment.py:61-63)
            assert list(got["root_match"]) == [r is not None and not (k == 1 and s == 4) for s, r in enumerate(got["roots"])]
            assert got["roots"] == results[k]["roots"]
        ref = auditor.audit(pack_sessions([chunks[1][2][int(chunks[1][1][s]):int(chunks[1][1][s + 1])] for s in range(10)]),
                            committed_roots_hex=[r if s != 4 else "0" * 64 for s, r in enumerate(chunks[1][3])])
        assert list(ref["root_match"]) == [r is not None and s != 4 for s, r in enumerate(chunks[1][3])]
        # views instead of copies, and a chunk too large for the buffers
        views = list(pipe.run([(chunks[0][0], chunks[0][1])], copy=False))
        assert np.array_equal(views[0]["digests"], results[0]["digests"])
        big = _chunk(range(100, 112))
        with pytest.raises(ValueError):
            list(pipe.run([(big[0], big[1])]))
    with ReplayPipeline(64, 1, 1) as empty:
        assert list(empty.run([])) == []


def test_device_packer_is_byte_identical_to_the_host_packer(audit_golden):
    """ahv_pack_deltas_dev (one thread per delta, same emitter source) against ahv_pack_deltas and against the real
    reference's json.dumps on the golden random sessions (escapes, surrogate pairs, None fields, odd timestamps)."""
    from agent_hypervisor_b200 import get_context
    from agent_hypervisor_b200.audit.packing import (canonical_preimage, column_bytes, columns_from_deltas,
                                                     device_columns_struct, pack_columns, stage_columns)
    from agent_hypervisor_b200 import SemanticDelta, VFSChange
    ctx = get_context()
    rs = audit_golden["random_sessions"]
    deltas = []
    for r in rs[:400]:
        parent = None
        for f in recipes.random_session(r["seed"]):
            d = SemanticDelta(delta_id=f["delta_id"], turn_id=f["turn_id"], session_id=f["session_id"], agent_did=f["agent_did"],
                              timestamp=f["timestamp"], changes=[VFSChange(**c) for c in f["changes"]], parent_hash=parent)
            parent = hashlib.sha256(canonical_preimage(d)).hexdigest()
            deltas.append(d)
    # a few parents that are not digests, and a None in the middle
    deltas[5].parent_hash = "not-a-digest é\U0001f600"; deltas[9].parent_hash = None; deltas[11].parent_hash = "A" * 64
    cols = columns_from_deltas(deltas)
    n = cols["n"]
    want_arena, want_offs, want_pos = pack_columns(cols, want_parent_pos=True)
    assert bytes(want_arena) == b"".join(canonical_preimage(d) for d in deltas)
    stage = ctx.host_alloc(column_bytes(cols))
    layout, used = stage_columns(cols, stage)
    d_cols, d_offs, d_pos = ctx.dev_alloc(used), ctx.dev_alloc(8 * (n + 1)), ctx.dev_alloc(4 * n)
    d_arena = ctx.dev_alloc(len(want_arena) + 64)
    try:
        ctx.h2d(d_cols, stage[:used])
        st = device_columns_struct(n, layout, d_cols)
        nbytes = ctx.pack_deltas_dev(st, d_offs, 0, 0, d_pos, want_bytes=True)                  # size query
        assert nbytes == len(want_arena)
        ctx.pack_deltas_dev(st, d_offs, d_arena, len(want_arena), d_pos)
        got_arena, got_offs, got_pos = np.empty(len(want_arena), np.uint8), np.empty(n + 1, np.uint64), np.empty(n, np.uint32)
        ctx.d2h(got_arena, d_arena); ctx.d2h(got_offs, d_offs); ctx.d2h(got_pos, d_pos); ctx.sync()
        assert np.array_equal(got_offs, want_offs) and np.array_equal(got_pos, want_pos)
        assert np.array_equal(got_arena, want_arena)
        # an arena that is too small: nothing written, reported by the next sync
        from agent_hypervisor_b200 import NativeError
        ctx.memset(d_arena, 0x5A, len(want_arena))
        ctx.pack_deltas_dev(st, d_offs, d_arena, len(want_arena) - 1, d_pos)
        with pytest.raises(NativeError, match="arena_capacity"):
            ctx.sync()
        ctx.d2h(got_arena, d_arena); ctx.sync()
        assert (got_arena == 0x5A).all()
    finally:
        ctx.host_free(stage)
        for p in (d_cols, d_offs, d_pos, d_arena):
            ctx.dev_free(p)


def test_device_packer_long_batch_offsets():
    """More deltas than the packer's one-CTA prefix scan handles (8 tiles of 4096): the three-launch scan must give
    the host packer's offsets and bytes - 70 001 benchmark-shaped deltas with digits of every width in their ids."""
    from agent_hypervisor_b200 import get_context
    from agent_hypervisor_b200.audit.packing import (_str_col, column_bytes, device_columns_struct, pack_columns,
                                                     stage_columns)
    ctx = get_context()
    n = 70001
    idx = np.arange(n)
    cols = {
        "n": n,
        "delta_id": _str_col([f"delta:{i}" for i in idx]),
        "session_id": _str_col([f"session:{i // 1000:08d}" for i in idx]),
        "agent_did": _str_col(["did:mesh:a" + "x" * (i % 7) for i in idx]),
        "turn_id": (idx % 1000 + 1).astype(np.int64),
        "timestamp_us": (1_767_225_600_000_000 + idx * 1_000_003).astype(np.int64),
        "timestamp_utcoffset_s": np.zeros(n, np.int32),
        "parent_kind": np.where(idx % 1000 == 0, 0, 1).astype(np.uint8),
        "parent_digest": np.random.default_rng(5).integers(0, 256, 32 * n, dtype=np.uint8),
        "parent_other": _str_col([""] * n),
        "change_rowptr": np.arange(n + 1, dtype=np.uint64),
        "change_path": _str_col([f"/f{i}" for i in idx]),
        "change_op": _str_col(["add"] * n),
        "change_content": _str_col([f"h{i}" for i in idx]),
        "change_prev": _str_col([None] * n),
    }
    want_arena, want_offs, want_pos = pack_columns(cols, want_parent_pos=True)
    stage = ctx.host_alloc(column_bytes(cols))
    layout, used = stage_columns(cols, stage)
    d_cols, d_offs, d_pos = ctx.dev_alloc(used), ctx.dev_alloc(8 * (n + 1)), ctx.dev_alloc(4 * n)
    d_arena = ctx.dev_alloc(len(want_arena) + 64)
    try:
        ctx.h2d(d_cols, stage[:used])
        st = device_columns_struct(n, layout, d_cols)
        ctx.pack_deltas_dev(st, d_offs, d_arena, len(want_arena), d_pos)
        got_arena, got_offs = np.empty(len(want_arena), np.uint8), np.empty(n + 1, np.uint64)
        ctx.d2h(got_arena, d_arena); ctx.d2h(got_offs, d_offs); ctx.sync()
        assert np.array_equal(got_offs, want_offs)
        assert np.array_equal(got_arena, want_arena)
    finally:
        ctx.host_free(stage)
        for p in (d_cols, d_offs, d_pos, d_arena):
            ctx.dev_free(p)


def test_device_replay_matches_the_host_packer_pipeline():
    from agent_hypervisor_b200 import DeviceReplay, ReplayPipeline
    from agent_hypervisor_b200.audit.packing import canonical_preimage, column_bytes
    chunks = [_chunk(range(10 * k, 10 * k + 10), tamper=(3, 1) if k == 2 else None) for k in range(5)]
    chunks.append(_chunk([7]))
    n_max = max(int(c[0]["n"]) for c in chunks)
    bytes_max = max(sum(len(canonical_preimage(d)) for d in c[2]) for c in chunks)
    col_max = max(column_bytes(c[0]) for c in chunks)

    def with_roots(k, c):
        stored = np.stack([np.frombuffer(bytes.fromhex(r or "00" * 32), np.uint8) for r in c[3]]).copy()
        if k == 1:
            stored[4, 0] ^= 1
        return c[0], c[1], stored
    with ReplayPipeline(bytes_max + 64, n_max, 10) as host, DeviceReplay(bytes_max + 64, n_max, 10, col_max) as dev:
        for feed in (lambda: ((c[0], c[1]) for c in chunks), lambda: (with_roots(k, c) for k, c in enumerate(chunks))):
            want = list(host.run(feed()))
            got = list(dev.run(feed()))
            assert len(got) == len(want) == len(chunks)
            for w, g, c in zip(want, got, chunks):
                assert np.array_equal(g["digests"], w["digests"]) and g["roots"] == w["roots"] == c[3]
                assert list(g["verdict"]) == list(w["verdict"]) == c[4] and np.array_equal(g["first_bad"], w["first_bad"])
                assert (g["root_match"] is None) == (w["root_match"] is None)
                if g["root_match"] is not None:
                    assert list(g["root_match"]) == list(w["root_match"])
        # columns that already sit in pinned memory: same results, no host copy
        staged = [dev.prestage(*with_roots(k, c)) for k, c in enumerate(chunks)]
        again = list(dev.run(staged))
        for w, g in zip(got, again):
            assert np.array_equal(g["digests"], w["digests"]) and g["roots"] == w["roots"] and list(g["root_match"]) == list(w["root_match"])
        for sc in staged:
            sc.close()
        assert list(dev.run([])) == []
        big = _chunk(range(100, 112))
        with pytest.raises(ValueError):
            list(dev.run([(big[0], big[1])]))
