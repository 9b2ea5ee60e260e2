"""ORACLE -- TEST INFRASTRUCTURE ONLY.

CPU restatement of the reference's beam-search decode path.  Only `tests/`,
`__graft_entry__.smoke()` and `bench.py`'s cpu_baseline / `--impl reference`
legs may import anything from this package; the product package
`imagetranslate_b200` never does (tests/test_boundary.py enforces it).

Parity status: the reference's own tests hold NO golden vectors for this path
 -- "parity unpinned" by reference tests.  The oracle is
pinned instead against the reference code itself, executed in this container
through `oracle/reference_loader.py`; its outputs are committed as fixtures
under tests/golden/ by `oracle/make_golden.py`.
"""
