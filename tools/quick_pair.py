"""Scratch: CTA-pair (cta_group::2) kernel against the single-CTA kernel, 256 queries."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from quick_batch import run
if __name__ == "__main__":
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 4000000
    which = sys.argv[2] if len(sys.argv) > 2 else "all"
    if which in ("all", "f32"):
        run(n, 1024, 256, 10, iters=5)
    if which in ("all", "bf16"):
        run(n, 768, 256, 10, iters=5, bf16=True)
    if which == "all":
        run(n, 1024, 200, 10, iters=3)
