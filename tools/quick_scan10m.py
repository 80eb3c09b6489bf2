import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__)))); sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import ragstore_b200 as rb
from quick_bench import run
n = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
run(n, 768, 10, rb.RAGSTORE_F32, iters=30)
