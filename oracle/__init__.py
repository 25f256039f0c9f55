"""ORACLE — test infrastructure only (CPU restatements of the reference's hot path).

Nothing under `image_super_resolution_b200/` may import from here; only tests/, the smoke
check in `__graft_entry__.py` and bench.py's cpu_baseline / `--impl reference` legs do.
"""
