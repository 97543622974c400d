"""Turns the raw ncu outputs of a gpurun call (gpurun_out/) into the tracked summaries under profiles/.
  python tools/profile_summary.py <tag> "<what was run>"
Reads gpurun_out/launches.csv (ncu --metrics gpu__time_duration.sum launch list) and
gpurun_out/prof_top.ncu-rep (ncu --set full of the dominant kernels)."""
import collections, csv, gzip, json, os, re, shutil, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag, what = sys.argv[1], sys.argv[2]
G, P = os.path.join(ROOT, "gpurun_out"), os.path.join(ROOT, "profiles")

rows = [r for r in csv.reader(open(os.path.join(G, "launches.csv"))) if len(r) > 5]
hdr = rows[0]
ki, vi, ui, gi = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit"), hdr.index("Grid Size")
agg = collections.OrderedDict()
for r in rows[1:]:
    n = re.sub(r"\(.*", "", r[ki])
    v = float(r[vi].replace(",", "")) / (1000.0 if r[ui].startswith("n") else 1.0)
    a = agg.setdefault(n, [0, 0.0])
    a[0] += 1
    a[1] += v
tot = sum(a[1] for a in agg.values())
with open(os.path.join(P, "%s_launch_summary.txt" % tag), "w") as f:
    f.write("# ncu --metrics gpu__time_duration.sum --clock-control none -c 3000, %s\n" % what)
    f.write("# per-launch times are cold-cache and serialised: compare SHARES with bench.py's roofline.kernel_share_of_step\n")
    for n, a in sorted(agg.items(), key=lambda x: -x[1][1]):
        f.write("%-62s n=%5d total_us=%11.1f share=%.4f mean_us=%9.1f\n" % (n[:62], a[0], a[1], a[1] / tot, a[1] / a[0]))
with open(os.path.join(G, "launches.csv"), "rb") as src, gzip.open(os.path.join(P, "%s_launches_ncu.csv.gz" % tag), "wb") as dst:
    shutil.copyfileobj(src, dst)

raw = subprocess.run(["ncu", "-i", os.path.join(G, "prof_top.ncu-rep"), "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units = rows[0], rows[1]
idx = {h: i for i, h in enumerate(hdr)}
want = ["gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "smsp__warps_eligible.avg.per_cycle_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
        "smsp__sass_inst_executed_op_shared_ld.sum", "smsp__sass_inst_executed_op_global_ld.sum", "smsp__sass_inst_executed_op_global_st.sum"]
traffic = {}
def gb(v, u):
    v = float(v)
    return v * {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}.get(u, 1.0)
with open(os.path.join(P, "%s_ncu_full_summary.txt" % tag), "w") as f:
    f.write("# ncu --set full --clock-control none --import-source on, %s\n" % what)
    for r in rows[2:]:
        name = re.sub(r"\(.*", "", r[idx["Kernel Name"]])
        f.write("--- %s\n" % name)
        for w in want:
            if w in idx:
                f.write("    %-72s %s %s\n" % (w, r[idx[w]], units[idx[w]]))
        b = gb(r[idx["dram__bytes_read.sum"]], units[idx["dram__bytes_read.sum"]]) + gb(r[idx["dram__bytes_write.sum"]], units[idx["dram__bytes_write.sum"]])
        key = "vn_fused" if "k_vn_fused" in name else "cn_minsum" if "k_cn_" in name else name
        traffic[key] = traffic.get(key, 0.0) + b
json.dump({"_what": "dram__bytes_read.sum + dram__bytes_write.sum per launch (cn_minsum = the exact-degree launches of one pass), " + what,
           **{k: round(v) for k, v in traffic.items()}}, open(os.path.join(P, "traffic.json"), "w"), indent=1)
print(open(os.path.join(P, "%s_launch_summary.txt" % tag)).read()[:1500])
print(json.load(open(os.path.join(P, "traffic.json"))))
