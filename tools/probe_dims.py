"""Scan bandwidth across embedding widths (the reference takes whatever the model returns: 384, 768, 1024, 1536, 3072, 4096 ...).
usage: python tools/probe_dims.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ragstore_b200 as rb
from quick_bench import run

if __name__ == "__main__":
    F, B = rb.RAGSTORE_F32, rb.RAGSTORE_BF16
    for (d, t) in [(384, F), (768, F), (1024, F), (1536, F), (2048, F), (3072, F), (4096, F), (1536, B), (3072, B), (4096, B), (1000, F), (100, F)]:
        n = int(6e9 // (d * (4 if t == F else 2)))       # ~6 GB per corpus: far beyond L2
        run(n, d, 10, t)
