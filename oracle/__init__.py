"""CPU oracle for the Minimal_RAG retrieval hot path -- TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this package; the product (the CUDA library and its host mirror)
never does.  PARITY UNPINNED: see oracle/ragindex_oracle.c header.
"""
