"""CPU oracle for the vae_npvc VQ-VAE hot path.  TEST INFRASTRUCTURE ONLY.

Nothing under ``vae_npvc_b200/`` may import this package.  Only ``tests/``,
``__graft_entry__.smoke()`` and the ``cpu_baseline`` / ``--impl reference`` legs of
``bench.py`` use it, and only as the checker or as the timed CPU baseline.

Parity status: PINNED against the reference's own Python modules imported from
``/root/reference`` in the build container (``oracle/make_golden.py`` writes the
fixtures under ``tests/golden/``; ``tests/test_oracle_golden.py`` re-checks them with no
access to the reference).  The reference itself ships no golden vectors.
"""
