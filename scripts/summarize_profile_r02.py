"""Turns the scratch output of scripts/gpu_r2_profiles.sh (gpurun_out/<tag>/) into the tracked summaries under
profiles/r02/: bench lines, ncu launch lists (+ the source hash they belong to), launch shares, and selected
metrics of the full ncu captures.   usage: python scripts/summarize_profile_r02.py <tag> <suffix>"""
import collections, csv, glob, os, shutil, subprocess, sys

tag, suf = sys.argv[1], sys.argv[2]
src, rdir = os.path.join("gpurun_out", tag), os.path.join("profiles", "r02")
os.makedirs(rdir, exist_ok=True)
KEYS = ['Block Size', 'Grid Size', 'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'gpu__dram_throughput.avg.pct',
        'sm__throughput.avg.pct', 'sm__warps_active.avg.pct', 'launch__registers_per_thread', 'launch__occupancy_limit', 'smsp__inst_executed.sum',
        'sm__pipe_alu_cycles_active.avg.pct', 'sm__pipe_fma_cycles_active.avg.pct', 'smsp__issue_active.avg.pct', 'stalled_barrier_per_issue',
        'stalled_long_scoreboard_per_issue', 'stalled_math_pipe_throttle_per_issue', 'stalled_no_instruction_per_issue', 'stalled_short_scoreboard_per_issue',
        'stalled_wait_per_issue', 'dram__cycles_active.avg.pct', 'launch__shared_mem_per_block', 'sm__warps_active.avg.per_cycle_active',
        'lts__t_bytes.sum', 'lts__throughput.avg.pct']
for rep, out in [("prof_c4.ncu-rep", f"ncu_full_c4_N262144_M5_summary_{suf}.csv"), ("prof_c2.ncu-rep", f"ncu_full_c2_N65536_M3_summary_{suf}.csv"),
                 ("prof_c5n3.ncu-rep", f"ncu_full_c5_nsga3_N65536_M10_summary_{suf}.csv")]:
    path = os.path.join(src, rep)
    if not os.path.exists(path):
        continue
    raw = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    keep = [h for h in hdr if any(k in h for k in KEYS)]
    with open(os.path.join(rdir, out), "w") as f:
        w = csv.writer(f)
        w.writerow(["metric", "unit"] + [r[idx["Kernel Name"]].split("(")[0] for r in rows[2:]])
        for h in keep:
            w.writerow([h, units[idx[h]]] + [r[idx[h]] for r in rows[2:]])
    print("wrote", out)
for path in sorted(glob.glob(os.path.join(src, f"launches_*_{suf}.csv"))):
    base = os.path.basename(path)
    shutil.copy(path, os.path.join(rdir, base))
    shutil.copy(path + ".meta.json", os.path.join(rdir, base + ".meta.json"))
    lr = list(csv.reader(l for l in open(path) if l.startswith('"')))
    h = lr[0]; ki = h.index("Kernel Name"); vi = h.index("Metric Value"); mi = h.index("Metric Name")
    agg = collections.defaultdict(collections.Counter)
    for r in lr[1:]:
        k = r[ki].split("(")[0]
        agg[k][r[mi]] += float(r[vi].replace(",", ""))
        agg[k]["n"] += r[mi] == "gpu__time_duration.sum"
    T = sum(v["gpu__time_duration.sum"] for v in agg.values())
    out = os.path.join(rdir, base.replace("launches_", "launch_share_").replace(".csv", ".txt"))
    with open(out, "w") as f:
        f.write("ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none: "
                "python scripts/profile_step.py (2 steps of nsga2_select; cold-cache serialised times: compare SHARES)\n")
        for k, v in sorted(agg.items(), key=lambda kv: -kv[1]["gpu__time_duration.sum"]):
            f.write(f"{k[:50]:50s} launches={int(v['n']):4d} total_us={v['gpu__time_duration.sum']/1e3:10.1f} share={v['gpu__time_duration.sum']/T*100:5.1f}% "
                    f"dram_read_GB={v['dram__bytes_read.sum']/1e9:8.3f} dram_write_GB={v['dram__bytes_write.sum']/1e9:8.3f}\n")
    print(open(out).read())
for path in sorted(glob.glob(os.path.join(src, "bench_*.json"))):
    shutil.copy(path, os.path.join(rdir, os.path.basename(path).replace(".json", f"_{suf}.json")))
