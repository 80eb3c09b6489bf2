"""Turns the ncu outputs brought back in gpurun_out/ into the small committed summaries under profiles/."""
import collections, csv, json, subprocess, sys

def launch_list(src, dst, title):
    rows = [r for r in csv.reader(open(src)) if r and not r[0].startswith("==")]
    hdr, data = rows[0], rows[1:]
    ik, iv, iu = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    agg = collections.OrderedDict()
    for r in data:
        v = float(r[iv].replace(",", "")); u = r[iu]
        ns = v * 1e3 if u in ("us", "usecond") else v * 1e6 if u in ("ms", "msecond") else v
        a = agg.setdefault(r[ik][:110], [0, 0.0]); a[0] += 1; a[1] += ns
    tot = sum(a[1] for a in agg.values())
    lines = ["# " + title, "# ncu --metrics gpu__time_duration.sum --clock-control none (cold-cache, serialised: compare SHARES)",
             "kernel,launches,total_us,avg_us,share"]
    for k, (n, ns) in sorted(agg.items(), key=lambda t: -t[1][1]):
        lines.append(f'"{k}",{n},{ns/1e3:.1f},{ns/1e3/n:.1f},{ns/tot:.4f}')
    open(dst, "w").write("\n".join(lines) + "\n")
    print("\n".join(lines))

WANT = ["Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__bytes_read.sum.per_second",
        "dram__bytes_write.sum.per_second", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "launch__registers_per_thread",
        "launch__grid_size", "launch__block_size", "launch__shared_mem_per_block_dynamic", "launch__cluster_size",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__m_xbar2l1tex_read_bytes.sum", "sm__cycles_elapsed.avg.per_second", "smsp__inst_executed.sum",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "TPC.TriageCompute.sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed"]

def full(rep, dst):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units, vals = rows[0], rows[1], rows[2]
    d = {h: [v, u] for h, u, v in zip(hdr, units, vals) if h in WANT}
    json.dump(d, open(dst, "w"), indent=1)
    for k, v in d.items():
        print(" ", k, v)
    return d

if __name__ == "__main__":
    launch_list("gpurun_out/launches_r1b.csv", "profiles/r1_launches_config2.csv",
                "ncu launch list, round 1: python bench.py --steps 20 --warmup 3 --no-extra (config2: 1M x 768 fp32, single query, top-10)")
    launch_list("gpurun_out/launches_r1b_config3.csv", "profiles/r1_launches_config3.csv",
                "ncu launch list, round 1 (-k regex:batch|prep): python bench.py --workload config3 --rows 4000000 --steps 3 --warmup 3 --no-extra (4M x 1024 fp32, 256 queries)")
    print("== scan"); d = full("gpurun_out/prof_r1b_config2.ncu-rep", "profiles/r1_ncu_full_config2_summary.json")
    print("== k1"); full("gpurun_out/prof_r1b_k1.ncu-rep", "profiles/r1_ncu_full_k1_summary.json")
    def gb(x):
        v, u = x; v = float(v); return v * {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1}.get(u, 1)
    t = gb(d["dram__bytes_read.sum"]) + gb(d["dram__bytes_write.sum"])
    json.dump({"config2:1000000": int(t), "_source": "profiles/r1_ncu_full_config2_summary.json: dram__bytes_read.sum + dram__bytes_write.sum of one scan_topk_kernel launch"},
              open("profiles/traffic.json", "w"), indent=1)
    print("traffic", int(t))
