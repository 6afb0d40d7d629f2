"""TEST INFRASTRUCTURE ONLY — CPU restatement of the reference's correlation hot path.

Nothing under ``oracle/`` is part of the product.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline / ``--impl reference``
legs may import it, and only as the checker or the timed CPU baseline.
The shipped package (``jax_raft_b200``) never imports this directory.
"""
