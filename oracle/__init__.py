"""CPU oracle for the FD score-matching hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``oracle/`` is part of the product: only
``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import it, and there only as the checker or the
timed CPU baseline.  The product package (``fd-scorematching_b200/``) never
imports this package and fails loudly when its CUDA library is missing.

Parity status: PINNED.  The reference ships no golden vectors, so
the restatement here is pinned against outputs of the reference's own functions,
executed in the build container by ``oracle/make_golden.py`` (which imports
``/root/reference`` unmodified, replays injected noise through
``torch.randn_like``) and committed under ``tests/golden/``.
"""
