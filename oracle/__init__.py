"""ORACLE package: CPU restatements of the reference hot path.  Test infrastructure only.

Importable from `tests/`, `__graft_entry__.smoke()` and `bench.py`'s CPU-baseline / `--impl reference` legs, and
from nowhere else.  See the header of each module for the reference lines it follows and how it is pinned.
"""
