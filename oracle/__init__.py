"""CPU oracle for the MAAVSS training hot path -- TEST INFRASTRUCTURE ONLY.

This package is a CPU restatement (torch-CPU / numpy / plain C) of the reference
algorithms on the hot path.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl reference``
legs may import it, and only as the *checker* (or the timed CPU baseline) -- never
as the product.  The product (``maavss_b200``) never imports ``oracle`` and raises if
its CUDA library is missing.

Parity pinning: the reference ships no tests and no golden vectors,
so the oracle is pinned against *outputs of the reference code itself*, produced in
the authoring container by ``tests/golden/make_golden.py`` (which imports
``/root/reference`` under a small shim) and committed as ``tests/golden/*.npz``.
``tests/test_oracle_golden.py`` checks every oracle function against those files.
"""
