"""TEST INFRASTRUCTURE ONLY -- the CPU oracle for the Genie 2 reverse-diffusion hot path.

Nothing in the product package (``genie2_b200``) imports this directory.  Only
``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import it, and only as the checker / CPU baseline,
never as the thing shipped.

Parity pinning: the restatement in :mod:`oracle.genie_oracle` is checked against
outputs of the unmodified reference (``/root/reference``, imported read-only by
``tests/golden/make_golden.py``) committed as fixtures under ``tests/golden/``.
"""
