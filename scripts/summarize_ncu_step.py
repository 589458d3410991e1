"""Summarise gpurun_out/<tag>_ncu_step.csv (scripts/ncu_step.sh) into tracked files under profiles/:
  <tag>_ncu_conv_all.csv    one row per convolution kernel launch of ONE training step (fprop / dgrad / wgrad):
                            duration, DRAM bytes, tensor-pipe activity, L2 -> SM bytes, L2 throughput, grid, layer
  <tag>_launch_summary.md   per kernel family: launches, time share, DRAM bytes, tensor-pipe %
  traffic.json              per-launch DRAM bytes and tensor-pipe % per bench.py family + the sources digest
usage: python scripts/summarize_ncu_step.py <tag> [layers.txt from bench.py --dump-layers]"""
import collections, csv, json, os, re, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
tag = sys.argv[1]
layers_file = sys.argv[2] if len(sys.argv) > 2 else None
rows = [r for r in csv.reader(l for l in open(os.path.join(ROOT, "gpurun_out", tag + "_ncu_step.csv")) if l.startswith('"'))]
hdr = rows[0]
I = {h: i for i, h in enumerate(hdr)}
launch = collections.OrderedDict()
for r in rows[1:]:
    if len(r) <= I["Metric Value"]:
        continue
    name = re.sub(r"\(.*", "", r[I["Kernel Name"]]).replace("void ", "").replace("rn::", "")
    rec = launch.setdefault(r[I["ID"]], dict(name=name, grid=r[I["Grid Size"]], block=r[I["Block Size"]]))
    v = float(r[I["Metric Value"]].replace(",", "") or 0)
    u = r[I["Metric Unit"]]
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1.0, "us": 1e3, "ms": 1e6, "usecond": 1e3, "nsecond": 1.0}.get(u, 1.0)
    rec[r[I["Metric Name"]]] = v * scale
seq = list(launch.values())
opt = [i for i, r in enumerate(seq) if r["name"].startswith("sgd_momentum")]
assert len(opt) >= 2, "need two optimizer launches to cut out one step"
step = seq[opt[-2] + 1:opt[-1] + 1]
# kernels after the step's last optimizer launch of the previous step include stem_weights of that step: drop leading non-step kernels
while step and not step[0]["name"].startswith(("cast_kernel", "s2d_pack", "vgg", "cifar")):
    step.pop(0)
T, RD, WR = "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum"
TP = "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active"
TPE = "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed"
XB, LTS, IPC = "l1tex__m_xbar2l1tex_read_bytes.sum", "lts__throughput.avg.pct_of_peak_sustained_elapsed", "sm__inst_executed.avg.per_cycle_elapsed"
cut = next(i for i, r in enumerate(step) if r["name"].startswith("softmax_xent"))
# layer names for the conv launches, from the executor's launch order (bench.py --dump-layers)
names = {"fprop": [], "dgrad": [], "wgrad": []}
if layers_file:
    for l in open(layers_file):
        m = re.match(r"conv_(fprop|dgrad|wgrad)_tc:(\S+)\s+[\d.]+ ms\s+[\d.]+ TFLOP/s\s+(in .*?\))\s+\[(.*)\]", l)
        if m:
            subs = re.search(r"subs=(\d+)", m.group(4))
            names[m.group(1)] += [(m.group(2), m.group(3), m.group(4))] * (int(subs.group(1)) if subs else 1)
out = []
cnt = {"fprop": 0, "dgrad": 0, "wgrad": 0}
for i, r in enumerate(step):
    n = r["name"]
    if n.startswith("conv_tc_kernel"):
        fam = "fprop" if i < cut else "dgrad"
    elif n.startswith("conv_wgrad_tc_kernel"):
        fam = "wgrad"
    else:
        continue
    k = cnt[fam]
    cnt[fam] += 1
    nm = names[fam][k] if k < len(names[fam]) else ("", "", "")
    out.append([fam, nm[0], nm[1], n, r["grid"], "%.2f" % (r.get(T, 0) / 1e3), "%.1f" % (r.get(RD, 0) / 1e6), "%.1f" % (r.get(WR, 0) / 1e6),
                "%.1f" % r.get(TP, 0), "%.1f" % r.get(TPE, 0), "%.1f" % (r.get(XB, 0) / 1e6), "%.1f" % r.get(LTS, 0),
                "%.2f" % r.get(IPC, 0), nm[2]])
with open(os.path.join(ROOT, "profiles", tag + "_ncu_conv_all.csv"), "w", newline="") as f:
    w = csv.writer(f)
    w.writerow(["gemm", "layer", "shape", "kernel", "grid", "duration_us", "dram_read_MB", "dram_write_MB", "tensor_pipe_pct_of_active",
                "tensor_pipe_pct_of_elapsed", "l2_to_sm_MB", "lts_throughput_pct", "sm_ipc", "plan"])
    w.writerows(out)
agg = collections.OrderedDict()
for i, r in enumerate(step):
    n = re.sub(r"<.*", "", r["name"])
    if n == "conv_tc_kernel":
        n = "conv_tc_kernel (fprop)" if i < cut else "conv_tc_kernel (dgrad)"
    a = agg.setdefault(n, [0, 0.0, 0.0, 0.0, 0.0])
    a[0] += 1; a[1] += r.get(T, 0); a[2] += r.get(RD, 0); a[3] += r.get(WR, 0); a[4] += r.get(TP, 0) * r.get(T, 0)
tot = sum(a[1] for a in agg.values())
with open(os.path.join(ROOT, "profiles", tag + "_launch_summary.md"), "w") as f:
    f.write("# %s: every kernel launch of ONE training step (ResNet-50 v2, 224x224, batch 128, bf16) under ncu\n\n" % tag)
    f.write("`ncu --metrics gpu__time_duration.sum,dram__bytes_*,sm__pipe_tensor_cycles_active,l1tex__m_xbar2l1tex_read_bytes,"
            "lts__throughput --clock-control none` on the un-captured launch list (scripts/ncu_step.sh).  Per-launch times under ncu "
            "are cold-cache and serialised: compare SHARES, not absolutes.  Tensor-pipe %% = sm__pipe_tensor_cycles_active of "
            "active cycles, time-weighted over the family's launches.  Per-launch rows: %s_ncu_conv_all.csv.\n\n" % tag)
    f.write("| kernel | launches | total us | share | DRAM read MB / launch | DRAM write MB / launch | tensor pipe % |\n|---|---:|---:|---:|---:|---:|---:|\n")
    for k, (c, t, rd, wr, tp) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        f.write("| %s | %d | %.1f | %.1f%% | %.1f | %.1f | %s |\n" % (k, c, t / 1e3, 100 * t / tot, rd / c / 1e6, wr / c / 1e6,
                                                                 ("%.1f" % (tp / t)) if tp > 0 else ""))
    f.write("| **total** | %d | %.1f | 100%% | %.2f GB read | %.2f GB written | |\n" % (
        sum(a[0] for a in agg.values()), tot / 1e3, sum(a[2] for a in agg.values()) / 1e9, sum(a[3] for a in agg.values()) / 1e9))
    # tensor pipe by layer class
    f.write("\n## Convolution launches by class (time-weighted tensor-pipe %, of active cycles)\n\n| class | launches | us | tensor pipe % | L2->SM MB / launch |\n|---|---:|---:|---:|---:|\n")
    cls = collections.OrderedDict()
    for o in out:
        shape = o[2]
        m = re.search(r"k(\d) s(\d)", shape)
        cin = re.search(r"in \(\d+, \d+, \d+, (\d+)\)", shape)
        pair = "pair" if "pair=1" in o[13] else ""
        key = "%s %sx%s s%s %s" % (o[0], m.group(1), m.group(1), m.group(2), pair) if m else o[0]
        c = cls.setdefault(key, [0, 0.0, 0.0, 0.0])
        c[0] += 1; c[1] += float(o[5]); c[2] += float(o[8]) * float(o[5]); c[3] += float(o[10])
    for k, (c, t, tp, xb) in cls.items():
        f.write("| %s | %d | %.1f | %.1f | %.1f |\n" % (k, c, t, tp / t if t else 0, xb / c))
print(open(os.path.join(ROOT, "profiles", tag + "_launch_summary.md")).read())
import bench
fam_of = {"conv_wgrad_tc_kernel": "conv_wgrad_tc", "bn_bwd_reduce_kernel": "bn_bwd_reduce", "bn_bwd_apply_kernel": "bn_bwd_apply",
          "bn_apply_kernel": "bn_apply", "bn_stats_kernel": "bn_stats", "sgd_momentum_kernel": "sgd_momentum_fused",
          "conv_tc_kernel (fprop)": "conv_fprop_tc", "conv_tc_kernel (dgrad)": "conv_dgrad_tc"}
layer_launches = {"conv_fprop_tc": len(set(n[0] for n in names["fprop"])) or None, "conv_dgrad_tc": len(set(n[0] for n in names["dgrad"])) or None}
traffic, pipe = {}, {}
for k, (c, t, rd, wr, tp) in agg.items():
    if k in fam_of:
        per = layer_launches.get(fam_of[k]) or c          # bench.py counts one launch per LAYER (a stride-2 dgrad = 4 kernels)
        traffic[fam_of[k]] = round((rd + wr) / per)
        if tp > 0:
            pipe[fam_of[k]] = round(tp / t, 1)
json.dump({"sources_sha256": bench.sources_digest(), "captured_by": "scripts/ncu_step.sh + scripts/summarize_ncu_step.py " + tag,
           "workload": "imagenet50", "batch": 128,
           "dram_bytes_per_launch": traffic, "tensor_pipe_pct": pipe}, open(os.path.join(ROOT, "profiles", "traffic.json"), "w"), indent=1)
print(json.dumps(json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))), indent=1))
