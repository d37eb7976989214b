"""CPU oracle for the kws-b200 hot path.  TEST INFRASTRUCTURE ONLY.

Nothing under ``oracle/`` is product code.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline / ``--impl reference``
legs may import it, and there only as the checker or as the timed CPU baseline.
The product path (``keyword-spotting_b200/``) never imports this package and
fails loudly when its CUDA library is missing.
"""
