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
