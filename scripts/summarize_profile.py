"""Turns the scratch output of scripts/gpu_profile.sh (gpurun_out/<tag>/) into the tracked summaries
under profiles/<round>/ (launch shares, per-kernel ncu metrics, bench lines).
usage: python scripts/summarize_profile.py <tag> <round_dir> <suffix>"""
import collections, csv, os, shutil, subprocess, sys

tag, rdir, suf = sys.argv[1], sys.argv[2], sys.argv[3]
src = os.path.join("gpurun_out", tag)
os.makedirs(rdir, exist_ok=True)
raw = subprocess.run(["ncu", "-i", os.path.join(src, "prof_c2.ncu-rep"), "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units = rows[0], rows[1]
keys = ['Block Size', 'Grid Size', 'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'gpu__dram_throughput.avg.pct',
        'sm__throughput.avg.pct', 'sm__warps_active.avg.pct', 'launch__registers_per_thread', 'launch__occupancy_limit', 'smsp__inst_executed.sum',
        'sm__pipe_alu_cycles_active.avg.pct', 'sm__pipe_fma_cycles_active.avg.pct', 'smsp__issue_active.avg.pct', 'stalled_barrier_per_issue',
        'stalled_long_scoreboard_per_issue', 'stalled_math_pipe_throttle_per_issue', 'stalled_no_instruction_per_issue', 'stalled_short_scoreboard_per_issue',
        'dram__cycles_active.avg.pct', 'launch__shared_mem_per_block_static', 'sm__warps_active.avg.per_cycle_active']
keep = [h for h in hdr if any(k in h for k in keys)]
idx = {h: i for i, h in enumerate(hdr)}
with open(os.path.join(rdir, f"ncu_full_c2_N65536_M3_summary_{suf}.csv"), "w") as f:
    w = csv.writer(f)
    w.writerow(["metric", "unit"] + [r[idx["Kernel Name"]].split("(")[0] for r in rows[2:]])
    for h in keep:
        w.writerow([h, units[idx[h]]] + [r[idx[h]] for r in rows[2:]])
for name, out in [("launches_c4.csv", f"launch_share_c4_N262144_M5_{suf}.txt"), ("launches_c2.csv", f"launch_share_c2_N65536_M3_{suf}.txt")]:
    lr = list(csv.reader(l for l in open(os.path.join(src, name)) if l.startswith('"')))
    h = lr[0]; ki = h.index("Kernel Name"); vi = h.index("Metric Value"); mi = h.index("Metric Name")
    agg = collections.defaultdict(collections.Counter)
    for r in lr[1:]:
        k = r[ki].split("(")[0]
        agg[k][r[mi]] += float(r[vi].replace(",", ""))
        agg[k]["n"] += r[mi] == "gpu__time_duration.sum"
    T = sum(v["gpu__time_duration.sum"] for v in agg.values())
    with open(os.path.join(rdir, out), "w") as f:
        f.write("ncu --metrics gpu__time_duration.sum[,dram__bytes_read.sum,dram__bytes_write.sum] --clock-control none: "
                "python scripts/profile_step.py (2 steps of nsga2_select; cold-cache serialised times: compare SHARES)\n")
        for k, v in sorted(agg.items(), key=lambda kv: -kv[1]["gpu__time_duration.sum"]):
            f.write(f"{k[:50]:50s} launches={int(v['n']):4d} total_us={v['gpu__time_duration.sum']/1e3:10.1f} share={v['gpu__time_duration.sum']/T*100:5.1f}% "
                    f"dram_read_GB={v['dram__bytes_read.sum']/1e9:8.3f} dram_write_GB={v['dram__bytes_write.sum']/1e9:8.3f}\n")
    print(open(os.path.join(rdir, out)).read())
for w in ("c4", "c2", "c5", "c3"):
    shutil.copy(os.path.join(src, f"bench_{w}.json"), os.path.join(rdir, f"bench_{w}_{suf}.json"))
shutil.copy(os.path.join(src, "launches_c4.csv"), os.path.join(rdir, f"launches_c4_N262144_M5_{suf}.csv"))
