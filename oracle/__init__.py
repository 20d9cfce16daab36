"""CPU oracle for the flow hot path — TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline /
``--impl reference`` legs may import this package.  The product package
``normalizing_flows_b200`` must never import it (tests/test_no_oracle_leak.py
checks this).
"""
