"""Scratch: few queries at once -- per-query scans against one tensor-core batch (RAGSTORE_BATCH_MINQ=99 forces scans)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from quick_batch import run
if __name__ == "__main__":
    for n in (10000, 100000, 1000000, 4000000):
        for nq in (2, 3, 4, 8, 15):
            run(n, 768, nq, 10, iters=7)
