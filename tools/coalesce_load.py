"""Closed-loop load on ragstore_search_coalesced: T threads, each asking one query at a time (the reference's request
handlers).  Prints queries/s, mean latency and how many corpus sweeps the library needed."""
import sys, os, time, threading
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import numpy as np
import torch
import ragstore_b200 as rb
from ragstore_b200 import native
from quick_bench import fill

def run(n, d, threads, seconds=1.0, window=None):
    st = rb.RagStore(d, n, rb.RAGSTORE_F32)
    fill(st, n, d, 1234)
    if window is not None:
        st.set_coalescing(window, 256)
    qd = torch.empty((256, d), dtype=torch.float32, device="cuda"); native.synth_fill_device(qd, 0, 256, d, d, 5678)
    qs = qd.cpu().numpy()
    st.search_coalesced(qs[0], 10)
    i0 = st.info()
    counts, lat, worst = [0] * threads, [0.0] * threads, [0.0] * threads
    stop = time.perf_counter() + seconds
    def worker(t):
        torch.cuda.set_device(0)
        i = t
        while True:
            t0 = time.perf_counter()
            if t0 >= stop:
                break
            st.search_coalesced(qs[i % 256], 10)
            dt1 = time.perf_counter() - t0
            lat[t] += dt1
            worst[t] = max(worst[t], dt1)
            counts[t] += 1
            i += threads
    ths = [threading.Thread(target=worker, args=(t,)) for t in range(threads)]
    t0 = time.perf_counter()
    for t in ths: t.start()
    for t in ths: t.join()
    dt = time.perf_counter() - t0
    i1 = st.info()
    total = sum(counts)
    sweeps = (i1["batch_launches"] - i0["batch_launches"]) + (i1["scan_launches"] - i0["scan_launches"])
    print(f"n={n} d={d} threads={threads} window={'default (0)' if window is None else window}: {total/dt:.0f} queries/s, "
          f"mean latency {sum(lat)/max(total,1)*1e3:.3f} ms, worst single call {max(worst)*1e3:.3f} ms, {total/max(sweeps,1):.1f} queries per corpus sweep", flush=True)
    st.close()

if __name__ == "__main__":
    for n in (10000, 1000000, 4000000):
        for th in (1, 4, 16, 64):
            run(n, 768, th)
        run(n, 768, 16, window=0)
        run(n, 768, 16, window=150)
