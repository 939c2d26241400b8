"""CPU oracle for the ESPnet2 Conformer CTC/attention train step.

TEST INFRASTRUCTURE ONLY.  This package is a plain fp32/fp64 CPU restatement of
the reference's algorithm for the hot path (each function cites the reference
file:line it follows).  Only ``tests/``, ``__graft_entry__.smoke()`` and the
``cpu_baseline`` / ``--impl reference`` legs of ``bench.py`` may import it, and
there only as the checker / the timed CPU baseline -- never as the product.
The product package ``espnet_slurp_b200`` must not import anything from here.

Parity pinning: the reference pins no numeric values on this path (its tests are
shape/smoke tests), so the oracle is pinned against outputs
of the reference itself, imported in the build container through
``oracle/ref_shim.py`` and frozen by ``oracle/make_golden.py`` into
``tests/golden/*.npz``; ``tests/test_oracle_golden.py`` replays them.
"""
