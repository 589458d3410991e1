"""Summarise ncu outputs brought back in gpurun_out/ into small tracked files under profiles/.
usage: python scripts/summarize_ncu.py <round tag> [launches.csv] [report.ncu-rep]"""
import csv, subprocess, sys, collections, io, re
tag = sys.argv[1]
launches = sys.argv[2] if len(sys.argv) > 2 else "gpurun_out/launches.csv"
rep = sys.argv[3] if len(sys.argv) > 3 else None
rows = [r for r in csv.reader(l for l in open(launches) if not l.startswith("=="))]
hdr = rows[0]
ki, vi = hdr.index("Kernel Name"), hdr.index("Metric Value")
agg = collections.OrderedDict()
for r in rows[1:]:
    if len(r) <= vi:
        continue
    name = re.sub(r"\(.*", "", r[ki]).replace("void ", "").replace("rn::", "")
    name = re.sub(r"<.*", "", name)
    t = float(r[vi].replace(",", ""))
    a = agg.setdefault(name, [0, 0.0])
    a[0] += 1; a[1] += t
tot = sum(a[1] for a in agg.values())
unit = rows[1][hdr.index("Metric Unit")]
with open("profiles/%s_launch_summary.md" % tag, "w") as f:
    f.write("# %s: kernel launch list of one profiled window (ncu --metrics gpu__time_duration.sum --clock-control none)\n\n" % tag)
    f.write("Per-launch times under ncu are cold-cache and serialised: compare SHARES, not absolutes.\n\n")
    f.write("| kernel | launches | total %s | share |\n|---|---:|---:|---:|\n" % unit)
    for k, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        f.write("| %s | %d | %.1f | %.1f%% |\n" % (k, c, t, 100 * t / tot))
    f.write("| **total** | %d | %.1f | 100%% |\n" % (sum(a[0] for a in agg.values()), tot))
print(open("profiles/%s_launch_summary.md" % tag).read())
if rep:
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rr = list(csv.reader(io.StringIO(out)))
    h, u = rr[0], rr[1]
    keep = ["Kernel Name", "launch__grid_size", "launch__registers_per_thread", "gpu__time_duration.sum",
            "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
            "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
            "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
            "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
            "lts__t_sectors_srcunit_tex_op_read.sum", "lts__t_sectors_srcunit_tex_op_read_lookup_hit.sum",
            "smsp__cycles_active.avg"]
    idx = [h.index(k) for k in keep if k in h]
    with open("profiles/%s_ncu_full_selected.csv" % tag, "w") as f:
        w = csv.writer(f)
        w.writerow([h[i] for i in idx]); w.writerow([u[i] for i in idx])
        for r in rr[2:]:
            w.writerow([r[i] for i in idx])
    print(open("profiles/%s_ncu_full_selected.csv" % tag).read())
