"""CPU oracle for the audit / liability hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``agent_hypervisor_b200/`` may import
this package.  Only ``tests/``, ``__graft_entry__.smoke()`` and the
``cpu_baseline`` / ``--impl reference`` legs of ``bench.py`` use it, and there
only as the checker (or as the timed CPU baseline), never as the product path.

Two restatements of the reference algorithm live here:

* ``oracle.audit_oracle`` / ``oracle.liability_oracle`` - Python, built on
  ``hashlib`` (the primitive the reference itself calls at
  ``src/hypervisor/audit/delta.py:63,130``) and plain float arithmetic.
* ``oracle/c/ahv_oracle.c`` - plain C (own FIPS 180-4 SHA-256), compiled by
  ``oracle/Makefile`` to ``oracle/_build/libahv_oracle.so`` and wrapped by
  ``oracle.coracle``.

Parity is PINNED: both are checked in ``tests/test_oracle_golden.py`` against
golden vectors produced by importing the real reference
(``tests/golden/make_golden.py``, run in the build container where
``/root/reference`` is mounted) and against NIST SHA-256 known answers.
"""
