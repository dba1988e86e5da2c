"""oracle/ — TEST INFRASTRUCTURE, not product code.

CPU restatements of the reference's algorithms for the off-grid attention path.  Only tests/, __graft_entry__.smoke()
and bench.py's cpu_baseline / `--impl reference` legs may import this package; nothing under
ml-autofocusformermod_b200/ does (tests/test_layout.py enforces it).
"""
