"""oracle/ — TEST INFRASTRUCTURE ONLY.

CPU restatement of the reference's pairwise + cluster path (oracle/pairwise_oracle.c,
oracle/cluster_oracle.py) plus drivers for the prebuilt unmodified reference (oracle/reftools.py).
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import
this package — and only as the checker.  The product (kspider_b200/) never imports it.
"""
