"""Summarise ncu outputs brought back in gpurun_out/ into small tracked files under profiles/.
usage: python scripts/summarize_ncu.py <round tag> [launches.csv] [report.ncu-rep]"""
import csv, subprocess, sys, collections, io, re, json
tag = sys.argv[1]
launches = sys.argv[2] if len(sys.argv) > 2 else "gpurun_out/launches.csv"
rep = sys.argv[3] if len(sys.argv) > 3 else None
rows = [r for r in csv.reader(l for l in open(launches) if not l.startswith("=="))]
hdr = rows[0]
ii, ki, mi, vi = hdr.index("ID"), hdr.index("Kernel Name"), hdr.index("Metric Name"), hdr.index("Metric Value")
ui = hdr.index("Metric Unit")
# one record per launch ID: duration (ns) and, when collected, DRAM bytes
launch = collections.OrderedDict()
for r in rows[1:]:
    if len(r) <= vi:
        continue
    name = re.sub(r"\(.*", "", r[ki]).replace("void ", "").replace("rn::", "")
    name = re.sub(r"<.*", "", name)
    rec = launch.setdefault(r[ii], dict(name=name, t=0.0, rd=0.0, wr=0.0))
    v = float(r[vi].replace(",", ""))
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1.0, "us": 1e3, "ms": 1e6}.get(r[ui], 1.0)
    if r[mi].startswith("gpu__time_duration"):
        rec["t"] = v * scale
    elif r[mi].startswith("dram__bytes_read"):
        rec["rd"] = v * scale
    elif r[mi].startswith("dram__bytes_write"):
        rec["wr"] = v * scale
seq = list(launch.values())
# keep ONE training step: the launches between the last two optimizer kernels
opt = [i for i, r in enumerate(seq) if r["name"].startswith("sgd_momentum")]
if len(opt) >= 2:
    seq = seq[opt[-2] + 1:opt[-1] + 1]
agg = collections.OrderedDict()
for r in seq:
    a = agg.setdefault(r["name"], [0, 0.0, 0.0, 0.0])
    a[0] += 1; a[1] += r["t"]; a[2] += r["rd"]; a[3] += r["wr"]
tot = sum(a[1] for a in agg.values())
with open("profiles/%s_launch_summary.md" % tag, "w") as f:
    f.write("# %s: kernel launches of ONE training step (ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,"
            "dram__bytes_write.sum --clock-control none, un-captured launch list)\n\n" % tag)
    f.write("Per-launch times under ncu are cold-cache and serialised: compare SHARES, not absolutes.\n\n")
    f.write("| kernel | launches | total us | share | DRAM read MB / launch | DRAM write MB / launch |\n|---|---:|---:|---:|---:|---:|\n")
    for k, (c, t, rd, wr) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        f.write("| %s | %d | %.1f | %.1f%% | %.1f | %.1f |\n" % (k, c, t / 1e3, 100 * t / tot, rd / c / 1e6, wr / c / 1e6))
    f.write("| **total** | %d | %.1f | 100%% | | |\n" % (sum(a[0] for a in agg.values()), tot / 1e3))
print(open("profiles/%s_launch_summary.md" % tag).read())
# DRAM bytes per launch by bench.py family name (roofline.traffic)
fam = {"conv_wgrad_tc_kernel": "conv_wgrad_tc", "bn_bwd_reduce_kernel": "bn_bwd_reduce", "bn_bwd_apply_kernel": "bn_bwd_apply",
       "bn_apply_kernel": "bn_apply", "bn_stats_kernel": "bn_stats", "sgd_momentum_kernel": "sgd_momentum_fused"}
import json
traffic = {}
for k, (c, t, rd, wr) in agg.items():
    if k in fam and rd + wr > 0:
        traffic[fam[k]] = round((rd + wr) / c)
# conv_tc_kernel serves fprop (before the loss kernel) and dgrad (after it); bench.py counts one launch per LAYER
# (a stride-2 dgrad is four kernel launches), so divide the family's bytes by bench.py's launch count
cut = next((i for i, r in enumerate(seq) if r["name"].startswith("softmax_xent")), None)
if cut is not None:
    try:
        per_step = {k["kernel"]: k["launches"] for k in json.load(open("profiles/%s_bench_n1.json" % tag))["kernels"]}
    except Exception:
        per_step = {}
    for famname, part in (("conv_fprop_tc", seq[:cut]), ("conv_dgrad_tc", seq[cut:])):
        rs = [r for r in part if r["name"] == "conv_tc_kernel"]
        if rs:
            traffic[famname] = round(sum(r["rd"] + r["wr"] for r in rs) / per_step.get(famname, len(rs)))
if "conv_wgrad_tc" in traffic and "wgrad_reduce_kernel" in agg:      # the family's launch = GEMM + partial reduction
    a = agg["wgrad_reduce_kernel"]
    traffic["conv_wgrad_tc"] += round((a[2] + a[3]) / agg["conv_wgrad_tc_kernel"][0])
if traffic:
    json.dump(traffic, open("profiles/traffic.json", "w"), indent=1)
    print("traffic.json", traffic)
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
