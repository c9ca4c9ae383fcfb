"""oracle/ -- TEST INFRASTRUCTURE ONLY.

CPU restatements of the reference's hot path (the checker).  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline / ``--impl reference``
legs may import anything from here; the product package ``haet_b200`` never does.
"""
