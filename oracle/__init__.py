"""ORACLE — test infrastructure only (CPU restatements of the reference's algorithms).
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import
this package; the product (plr2_b200/) never does."""
