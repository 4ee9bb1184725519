"""CPU oracle for the robust-aggregation hot path.  TEST INFRASTRUCTURE ONLY.

This package restates, on the CPU, what the reference's defenses compute
(/root/reference/src/defenses/*.py; every function cites the lines it follows).
It exists so the CUDA path can be checked against something that does not need
/root/reference at run time (the GPU box does not have it).

Rules (enforced by tests/test_layout.py):
  * only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
    ``--impl reference`` legs may import anything from here;
  * nothing under multimodal-fl-security_b200/ imports it: the product path has
    no CPU fallback and raises when libflagg.so is missing.

Parity pinning: the reference ships NO golden vectors or known-answer tests for
this path (its tests/test_defenses.py asserts only qualitative facts).  The oracle
is therefore pinned against outputs of the reference itself, run in the build
container by tests/golden/make_golden.py (committed, together with the .npz
fixtures it wrote).  tests/test_oracle_golden.py re-checks oracle == fixtures on
every CPU test run.
"""
from . import ref_port, oracle64  # noqa: F401
