"""CPU oracle for the fly_bProject hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``oracle/`` is part of the product:
only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline /
``--impl reference`` legs may import it, and only as the checker or as the
timed CPU baseline.  The product path (``fly_bproject_b200``) never imports it
and fails loudly when the CUDA library is missing.

Parity status: the reference ships no tests or golden vectors,
so the oracle is pinned against outputs of the reference's own code executed
in the build container (``oracle/ref_harness.py`` imports ``/root/reference``
unmodified; ``tests/golden/make_golden.py`` froze the vectors).  The
un-vendored Isaac Gym helpers (``isaacgym.torch_utils``; no version pinned by
the reference) are restated from their published formulas in
``oracle/isaac_math.py`` and additionally checked against scipy rotations.
"""
