#!/usr/bin/env python
"""Benchmark of the synchronous data-parallel ResNet training step (BASELINE.json metric:
ResNet-50 train images/sec at 1/2/4/8 B200; conv tensor-pipe utilisation).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
                    [--workload imagenet50|cifar50|cifar20|wrn28_10] [--global-batch G] [--sustained-steps S]

`--global-batch G` fixes the GLOBAL batch (per-GPU batch = G / N, "scaling": "strong") — BASELINE config 2, the
reference's headline experiment: the CIFAR ResNet-50 at global batch 128 over 1/2/4/8 workers.

N > 1 is launched by the driver under torchrun (one rank per GPU, NCCL).  One "step" is one
full training step (forward, backward, gradient all-reduce, fused update) on synthetic data.
Rank 0 prints ONE JSON line.  See DESIGN.md "Measurement" for how every field is produced.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (dataset, image size, classes, per-GPU batch for weak scaling, weight decay)
    "imagenet50": dict(dataset="imagenet", size=224, classes=1000, batch=128, wd=1e-4,
                       desc="ResNet-50 v2 bottleneck [3,4,6,3], ImageNet-shape 224x224, 128 images/GPU (C3)"),
    "cifar50": dict(dataset="cifar10", size=32, classes=10, batch=128, wd=2e-4,
                    desc="cifar10_resnet_v2_generator(50) basic blocks, 32x32, 128 images/GPU (C2 net)"),
    # upstream-style builders (HParams.num_residual_units given), SURVEY section 8 configs C1 and C4
    "cifar20": dict(dataset="cifar10", size=32, classes=10, batch=128, wd=2e-4,
                    upstream=dict(num_residual_units=3, use_bottleneck=False, relu_leakiness=0.1), filters=None,
                    desc="upstream ResNet-20 (_residual units 3x3, filters 16/16/32/64), CIFAR-10, 128 images/GPU (C1)"),
    "wrn28_10": dict(dataset="cifar100", size=32, classes=100, batch=128, wd=5e-4,
                     upstream=dict(num_residual_units=4, use_bottleneck=False, relu_leakiness=0.1),
                     filters=[16, 160, 320, 640],
                     desc="upstream WRN-28-10 (filters 16/160/320/640), CIFAR-100, 128 images/GPU (C4)"),
}
# derived images/s from the reference README: ImageNet ResNet-50, 128/worker
PUBLISHED = {("imagenet50", 1): 123.0, ("imagenet50", 8): 952.0, ("cifar50", 1): 1784.0}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(hbm=d["hbm_gbs"], tf_burst=d["bf16_tflops"], tf_sust=d.get("bf16_tflops_sustained", d["bf16_tflops"]),
                    src="measured")
    return dict(hbm=6650.0, tf_burst=1590.0, tf_sust=1400.0, src="fallback")


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index = index
        self.samples = []
        self._halt = threading.Event()

    def run(self):
        while not self._halt.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                f = [x.strip() for x in out.strip().split(",")]
                if len(f) >= 7:
                    self.samples.append(f)
            except Exception:
                pass
            self._halt.wait(0.2)

    def stop(self):
        self._halt.set()
        self.join(timeout=6)
        sm = sorted(float(s[0]) for s in self.samples if s[0].replace(".", "").isdigit())
        mx = [float(s[1]) for s in self.samples if s[1].replace(".", "").isdigit()]
        reasons = []
        for i, name in enumerate(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")):
            if any(s[3 + i].lower().startswith("active") for s in self.samples):
                reasons.append(name)
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(self.samples)}


def sources_digest():
    """Digest of csrc/ + the header (what build.py stamps): ties an ncu-derived profiles/traffic.json to the code."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("_rn_build", os.path.join(ROOT, "distributed-tensorflow-resnet_b200", "build.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod.digest()


def conv_flops(graph):
    """Algorithmic FLOPs: 2*N*Ho*Wo*Cout*k^2*Cin per GEMM; dgrad skipped where the
    input needs no gradient (the stem)."""
    out = {}
    for n in graph.nodes:
        if n.op == "conv2d":
            x, y = n.inputs[0], n.outputs[0]
            f = 2.0 * y.numel * n.attrs["k"] ** 2 * x.shape[3]
            out[y.name] = (f, x.requires_grad)
    return out


def conv_layerwise_roofline(ex, graph, per_launch_ms, pk):
    """Every convolution launch against ITS OWN bound: max(FLOPs / measured bf16 peak, algorithmic bytes / measured copy
    bandwidth) — most 1x1 layers of ResNet-50 are HBM-bound, so a tensor-only roofline understates them.  Bytes: the
    activation tensors the launch must read and write once (+ the shortcut of a fused add, + the BN input a fused
    dgrad re-reads); weights and split-K partials are not counted."""
    s = 2 if ex.bf16 else 4
    plans = ex.describe_plans()
    tot = {}
    for n in graph.nodes:
        if n.op != "conv2d":
            continue
        x, y = n.inputs[0], n.outputs[0]
        fl = 2.0 * y.numel * n.attrs["k"] ** 2 * x.shape[3]
        xin = x.numel * s if x.shape[3] >= 16 else y.numel // y.shape[3] * 16 * s      # stem: the packed 16-channel view
        for fam in ("conv_fprop_tc", "conv_dgrad_tc", "conv_wgrad_tc", "conv_fprop", "conv_dgrad", "conv_wgrad"):
            ms = per_launch_ms.get(fam + ":" + y.name)
            if ms is None:
                continue
            by = xin + y.numel * s
            f = getattr(n, "fuse", None)
            if "fprop" in fam and f is not None and f["residual"] is not None:
                by += y.numel * s
            if "dgrad" in fam and "fused=001" in plans.get(fam + ":" + y.name, ""):
                by += x.numel * s
            ideal = max(fl / (pk["tf_burst"] * 1e9), by / (pk["hbm"] * 1e6))           # ms
            base = fam.replace("_tc", "")
            t = tot.setdefault(base, [0.0, 0.0, 0, 0])
            t[0] += ideal
            t[1] += ms
            t[2] += 1
            t[3] += 1 if fl / (pk["tf_burst"] * 1e9) > by / (pk["hbm"] * 1e6) else 0
    out = {k: {"ideal_ms": round(v[0], 3), "measured_ms": round(v[1], 3), "frac": round(v[0] / v[1], 3), "launches": v[2],
               "tensor_bound_launches": v[3]} for k, v in tot.items()}
    i, m = sum(v[0] for v in tot.values()), sum(v[1] for v in tot.values())
    out["all"] = {"ideal_ms": round(i, 3), "measured_ms": round(m, 3), "frac": round(i / m, 3) if m else None}
    return out


def decode_throughput(pipe, pk):
    """The CIFAR decode kernel on the whole record file in ONE launch (50,000 images): algorithmic bytes (record read +
    image + one-hot written) over the CUDA-event time, against the measured copy bandwidth."""
    import torch
    from distributed_tensorflow_resnet_b200._lib import lib, check, ptr, RN_BF16, RN_F32
    if not pipe.resident:
        return None
    n = pipe.n
    img = torch.empty(n, 32, 32, 3, dtype=pipe.images.dtype, device=pipe.images.device)
    oh = torch.empty(n, pipe.classes, dtype=torch.float32, device=img.device)
    odt = RN_BF16 if img.dtype == torch.bfloat16 else RN_F32
    st = torch.cuda.current_stream()
    best = None
    for i in range(4):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(st)
        check(lib.rn_cifar_decode(ptr(pipe.records), n, pipe.record_bytes, pipe.label_offset, pipe.classes, None, n, 1, pipe.pad,
                                  pipe.seed & 0xFFFFFFFF, 0, i, None, 0, ptr(img), odt, ptr(oh), None, None, st.cuda_stream),
              "rn_cifar_decode")
        e1.record(st)
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        best = ms if best is None or ms < best else best
    by = n * (pipe.record_bytes + 3072 * img.element_size() + 4 * pipe.classes)
    return {"images": n, "ms": round(best, 4), "bytes": by, "gbs": round(by / best / 1e6, 1),
            "frac_of_hbm_peak": round(by / best / 1e6 / pk["hbm"], 4), "images_per_s": round(n / best * 1e3)}


def family(name):
    return name.split(":")[0]


def hbm_bytes(name, ex, graph):
    """Algorithmic HBM bytes of one launch of the HBM-bound families."""
    fam, _, key = name.partition(":")
    s = 2 if ex.bf16 else 4
    if fam in ("bn_stats", "bn_apply", "bn_infer", "bn_bwd_reduce", "bn_bwd_apply"):
        node = [n for n in graph.nodes if n.op == "bn_act" and n.attrs["scope"] == key][0]
        E = node.inputs[0].numel
        fused = ex._fused_add_for(node) is not None
        if fam == "bn_stats":
            return E * s * (3 if fused else 1)
        if fam in ("bn_apply", "bn_infer", "bn_bwd_reduce"):
            return 2 * E * s
        slot = node.inputs[0].gradslot
        return (4 if len(slot.members) > 1 else 3) * E * s
    if fam == "sgd_momentum_fused":
        return ex.n_flat * (20 + (2 if ex.bf16 else 0))
    if fam == "cifar_decode":
        # record bytes read + image (fp32 12,288 B / bf16 6,144 B) + one-hot (4*classes) written, per image
        pipe = ex._bench_pipe
        return pipe.B * (pipe.record_bytes + 3072 * (2 if pipe.images.dtype == __import__("torch").bfloat16 else 4)
                         + 4 * pipe.classes)
    return None


def profile_step(ex, graph, iters=3):
    """Per-launch CUDA-event timing of the un-captured launch list (same stream the kernels run on)."""
    import torch
    lists = [("fwd", ex.fwd), ("bwd", ex.bwd), ("upd", ex.update)]
    pipe = getattr(ex, "_bench_pipe", None)
    if pipe is not None:       # the record decode runs once per step in front of the captured graph
        lists.insert(0, ("input", [("cifar_decode:batch", lambda st_: pipe.next_batch())]))
    st = torch.cuda.current_stream()
    acc = {}
    for it in range(iters + 1):
        evs = []
        for _, lst in lists:
            for name, fn in lst:
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record(st)
                fn(st.cuda_stream)
                b.record(st)
                evs.append((name, a, b))
        torch.cuda.synchronize()
        if it == 0:
            continue        # warm-up
        for name, a, b in evs:
            acc.setdefault(name, []).append(a.elapsed_time(b))
    return {k: sum(v) / len(v) for k, v in acc.items()}


def roofline_report(ex, graph, per_launch_ms, pk, clocks=None, workload=None, batch=None):
    cf = conv_flops(graph)
    fams = {}
    for name, ms in per_launch_ms.items():
        fam, _, key = name.partition(":")
        d = fams.setdefault(fam, dict(ms=0.0, flops=0.0, bytes=0.0, launches=0))
        d["ms"] += ms
        d["launches"] += 1
        base = fam.replace("_tc", "")
        if base in ("conv_fprop", "conv_dgrad", "conv_wgrad") and key in cf:
            d["flops"] += cf[key][0]
        else:
            b = hbm_bytes(name, ex, graph)
            if b:
                d["bytes"] += b
    total = sum(d["ms"] for d in fams.values())
    table = []
    for fam, d in sorted(fams.items(), key=lambda kv: -kv[1]["ms"]):
        row = dict(kernel=fam, launches=d["launches"], ms=round(d["ms"], 4), share=round(d["ms"] / total, 4))
        if d["flops"]:
            row.update(bound="tensor", achieved=round(d["flops"] / d["ms"] / 1e9, 2), unit="TFLOP/s",
                       frac=round(d["flops"] / d["ms"] / 1e9 / pk["tf_sust"], 4))
        elif d["bytes"]:
            row.update(bound="hbm", achieved=round(d["bytes"] / d["ms"] / 1e6, 1), unit="GB/s",
                       frac=round(d["bytes"] / d["ms"] / 1e6 / pk["hbm"], 4))
        table.append(row)
    top = next((r for r in table if "bound" in r), None)
    roof = None
    traffic = None
    tensor_pipe = None
    stale = None
    tp = os.path.join(ROOT, "profiles", "traffic.json")     # per-launch DRAM bytes / tensor-pipe % from the ncu captures
    if top and os.path.exists(tp):
        tj = json.load(open(tp))
        if (tj.get("workload"), tj.get("batch")) != (workload, batch):
            stale = "profiles/traffic.json was captured on %s at batch %s: traffic not reported for this workload" % (
                tj.get("workload"), tj.get("batch"))
        elif tj.get("sources_sha256") == sources_digest():
            traffic = tj.get("dram_bytes_per_launch", {}).get(top["kernel"])
            tensor_pipe = tj.get("tensor_pipe_pct")
        else:
            stale = "profiles/traffic.json was captured from other sources (digest mismatch): traffic not reported"
    if top:
        tensor = top["bound"] == "tensor"
        # The per-launch times come from a short un-captured burst; which measured peak applies depends on the
        # clock the GPU actually held (MEASURED_PEAKS: burst figure at max clock, sustained under a long load).
        at_max = bool(clocks and clocks.get("sm_mhz") and clocks.get("sm_max_mhz") and
                      clocks["sm_mhz"] >= 0.97 * clocks["sm_max_mhz"])
        peak = (pk["tf_burst"] if at_max else pk["tf_sust"]) if tensor else pk["hbm"]
        roof = dict(kernel=top["kernel"], bound=top["bound"], achieved=top["achieved"], peak=peak, unit=top["unit"],
                    frac=round(top["achieved"] / peak, 4), traffic=traffic,
                    peak_source=pk["src"] + ((" burst bf16 (SM clock at max during the timed region)" if at_max
                                              else " sustained bf16") if tensor else " copy"),
                    share_of_step=top["share"], launches_per_step=top["launches"])
        if tensor:
            roof["frac_of_burst_peak"] = round(top["achieved"] / pk["tf_burst"], 4)
            roof["frac_of_sustained_peak"] = round(top["achieved"] / pk["tf_sust"], 4)
        if tensor_pipe:
            roof["tensor_pipe_pct"] = tensor_pipe
        if stale:
            roof["traffic_note"] = stale
    return roof, table, total


def ddp_check(ex, model, world, rank, dev):
    """Values, not just speed, at N > 1: (1) after the timed steps every rank holds bit-identical
    weights, momentum and BN moving statistics... of the replicas that must agree: W and Mom (BN statistics are
    rank-local by design, the reference has no sync-BN); (2) two steps of a tiny net (CIFAR ResNet-8, fp32) equal
    the CPU oracle's step on the MEAN of the R per-replica gradients (SyncReplicasOptimizer / Horovod semantics,
    models/resnet_model.py:104-120), on both the eager and the segmented-CUDA-graph path; (3) a checkpoint written
    by the chief and resumed on every rank leaves all ranks on identical weights and the same global_step."""
    import tempfile
    import torch
    import torch.distributed as dist
    from distributed_tensorflow_resnet_b200 import resnet_model, FLAGS, train_utils
    out = {}

    def same_everywhere(t):
        h = torch.stack([t.view(torch.int32).to(torch.int64).sum(), t.double().abs().sum().view(torch.int64)])
        g = [torch.empty_like(h) for _ in range(world)]
        dist.all_gather(g, h)
        return all(torch.equal(g[0], x) for x in g)

    out["ranks_identical"] = bool(same_everywhere(ex.W) and same_everywhere(ex.Mom))
    saved = dict(FLAGS.__dict__)
    worst = 0.0
    try:
        B = 8
        g = torch.Generator().manual_seed(5)
        x = torch.randn(world * B, 32, 32, 3, generator=g)
        y = torch.eye(10)[torch.arange(world * B) % 10]
        for use_graph in (False, True):
            FLAGS.reset()
            FLAGS.update(compute_dtype="f32", use_cuda_graph=use_graph, dataset="cifar10", resnet_size=8, seed=3,
                         job_name="worker", sync_replicas=True, replicas_to_aggregate=world, task_index=rank)
            X = x[rank * B:(rank + 1) * B].to(dev).contiguous()
            Y = y[rank * B:(rank + 1) * B].to(dev).contiguous()
            hps = resnet_model.HParams(num_classes=10, lrn_rate=0.1, weight_decay_rate=2e-4, optimizer='mom')
            m = resnet_model.ResNet(hps, X, Y, 'train')
            m.build_graph()
            e2 = m.executor
            params = {k: v for k, v in e2.state_dict().items() if not k.endswith("/Momentum") and k != "global_step"}
            for _ in range(2):
                m.train_op()
            torch.cuda.synchronize()
            out["ranks_identical"] = bool(out["ranks_identical"] and same_everywhere(e2.W) and same_everywhere(e2.Mom))
            if rank == 0:
                from oracle import resnet_oracle as ro           # the checker (test infrastructure), rank 0 only
                net = ro.OfficialResNetV2("cifar10", 8, 10, dtype=torch.float64, params=params)
                tr = ro.Trainer(net, 2e-4, "mom")
                for _ in range(2):
                    tr.step([(x[r * B:(r + 1) * B].double(), y[r * B:(r + 1) * B].double()) for r in range(world)], 0.1)
                for v in m.graph.trainable_variables():
                    got = e2.var_view(v.name).double().cpu()
                    want = net.params[v.name]
                    worst = max(worst, float((got - want).abs().max() / (want.abs().max() + 1e-30)))
            if use_graph:
                # (3) chief saves, every rank builds a DIFFERENTLY seeded model and resumes
                d = [tempfile.mkdtemp() if rank == 0 else None]
                dist.broadcast_object_list(d, src=0)
                if rank == 0:
                    train_utils.save_checkpoint(m, d[0], m.global_step)
                dist.barrier()
                FLAGS.update(seed=100 + rank)
                m2 = resnet_model.ResNet(hps, X, Y, 'train')
                m2.build_graph()
                train_utils.resume(m2, d[0])
                steps = torch.tensor([m2.global_step], device=dev)
                gs = [torch.empty_like(steps) for _ in range(world)]
                dist.all_gather(gs, steps)
                out["resume_ranks_identical"] = bool(same_everywhere(m2.executor.W) and same_everywhere(m2.executor.Mom)
                                                     and same_everywhere(m2.executor.S)
                                                     and torch.equal(m2.executor.W, e2.W)
                                                     and all(int(s_.item()) == m.global_step for s_ in gs))
                m2.executor.close()
            e2.close()
    finally:
        FLAGS.__dict__.clear()
        FLAGS.__dict__.update(saved)
    out["oracle_rel_err"] = worst if rank == 0 else None
    out["what"] = ("W / Mom checksums all-gathered after the timed steps; 2 steps of CIFAR ResNet-8 fp32 vs the CPU oracle's "
                   "mean-of-%d-replicas step (eager and CUDA-graph paths); chief checkpoint resumed on all ranks" % world)
    return out


def cpu_reference_step(workload, batch, steps, warmup, threads=None):
    """The CPU restatement of the reference's step (oracle, torch-CPU oneDNN, fp32): the reference's
    own TensorFlow CPU path is unavailable on this box (TensorFlow is not installed)."""
    import torch
    from oracle import resnet_oracle as ro
    wl = WORKLOADS[workload]
    torch.set_num_threads(threads or os.cpu_count())
    torch.set_flush_denormal(True)
    g = torch.Generator().manual_seed(1000)
    x = torch.randn(batch, wl["size"], wl["size"], 3, generator=g)
    y = torch.eye(wl["classes"])[torch.randint(0, wl["classes"], (batch,), generator=g)]
    if wl.get("upstream"):
        net = ro.UpstreamResNet(wl["classes"], filters=wl["filters"], dtype=torch.float32, seed=0, **wl["upstream"])
    else:
        net = ro.OfficialResNetV2(wl["dataset"], 50, wl["classes"], dtype=torch.float32, seed=0)
    tr = ro.Trainer(net, wl["wd"], "mom")
    times = []
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        tr.step([(x, y)], 0.1)
        if i >= warmup:
            times.append(time.perf_counter() - t0)
    dt = sum(times) / len(times)
    return batch / dt, dt, torch.get_num_threads()


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    wl = WORKLOADS[args.workload]
    batch = 32 if args.workload == "imagenet50" else 128      # a bounded sample: seconds per step on the host cores
    steps = max(1, min(args.steps, 5))
    warm = 1
    ips, dt, threads = cpu_reference_step(args.workload, batch, steps, warm)
    line = {
        "impl": "reference", "metric": "train images/sec", "value": round(ips, 3), "unit": "images/s",
        "n_gpus": args.gpus, "steps": steps, "warmup": warm, "ms_per_step": round(dt * 1e3, 2),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": args.workload, "description": wl["desc"], "sample_batch": batch},
        "cpu_baseline": {"value": round(ips, 3), "unit": "images/s", "cores": threads, "kind": "port",
                         "sample": "%d timed steps of a %d-image batch of the same net/shape (oracle, torch-CPU fp32); "
                                   "the reference's TF CPU path is unavailable on this box" % (steps, batch)},
        "e2e": {"value": round(ips, 3), "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


def run_ours(args):
    import torch
    import torch.distributed as dist
    from distributed_tensorflow_resnet_b200 import resnet_model, FLAGS, _lib
    wl = WORKLOADS[args.workload]
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    if world != args.gpus and rank == 0:
        sys.stderr.write("warning: --gpus %d but WORLD_SIZE %d\n" % (args.gpus, world))
    strong = args.global_batch > 0
    if strong:
        if args.global_batch % world:
            raise SystemExit("--global-batch %d is not divisible by %d ranks" % (args.global_batch, world))
        B = args.global_batch // world
    else:
        B = args.batch or wl["batch"]
    dev = torch.device("cuda", local)
    dt = torch.bfloat16 if args.dtype == "bf16" else torch.float32
    FLAGS.reset()
    FLAGS.update(dataset=wl["dataset"], compute_dtype=args.dtype, use_cuda_graph=not args.no_graph, resnet_size=50,
                 seed=0, sync_replicas=True, job_name="worker" if world > 1 else None, replicas_to_aggregate=world)
    cifar = wl["dataset"] in ("cifar10", "cifar100")
    pipe = None
    if cifar:
        # the reference's input path: binary CIFAR records -> rn_cifar_decode (crop / flip / standardise) every step;
        # synthetic records, resident in HBM for `value`, in pinned host memory for `e2e`
        from distributed_tensorflow_resnet_b200 import cifar_input
        images, labels = cifar_input.input_fn(True, "synthetic:50000", B, dataset=wl["dataset"], seed=0)
        pipe = images._rn_pipeline
        host_records = torch.from_numpy(cifar_input.synthetic_records(50000, wl["dataset"], 0)).pin_memory()
    else:
        gen = torch.Generator().manual_seed(1000 + rank)
        host_img = torch.randn(B, wl["size"], wl["size"], 3, generator=gen).pin_memory()        # fp32, as tf.data yields
        lab = torch.randint(0, wl["classes"], (B,), generator=gen)
        host_lab = torch.zeros(B, wl["classes"]).pin_memory()
        host_lab[torch.arange(B), lab] = 1.0
        images = torch.empty(B, wl["size"], wl["size"], 3, dtype=torch.float32, device=dev)   # model casts to bf16 itself
        labels = torch.empty(B, wl["classes"], dtype=torch.float32, device=dev)
        images.copy_(host_img)
        labels.copy_(host_lab)
    hkw = dict(num_classes=wl["classes"], lrn_rate=0.1, weight_decay_rate=wl["wd"], optimizer='mom')
    if wl.get("upstream"):
        hkw.update(batch_size=B, min_lrn_rate=1e-4, **wl["upstream"])
    hps = resnet_model.HParams(**hkw)
    model = resnet_model.ResNet(hps, images, labels, 'train')
    if wl.get("filters"):
        model.filters = wl["filters"]
    model.build_graph()
    ex = model.executor
    ex._bench_pipe = pipe
    model.lrn_rate = 0.1
    pk = peaks()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident throughput -----------------------------------------------------
    for _ in range(args.warmup):
        model.train_op()
    barrier()
    sampler = ClockSampler(local) if rank == 0 else None
    if sampler:
        sampler.start()
    l0 = _lib.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for _ in range(args.steps):
        model.train_op()
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1)
    clocks = sampler.stop() if sampler else None
    launched_live = _lib.launch_count() - l0
    t = torch.tensor([ms], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = t.item()
    value = world * B * args.steps / (ms / 1e3)

    # ---- end to end through the public API with host buffers ------------------------------
    # (directly after the device-resident region, so both are measured in the same clock / thermal state; the
    #  sustained leg below then shows what seconds of back-to-back steps do to the clocks)
    out_host = torch.zeros(2, dtype=torch.float32).pin_memory()
    if cifar:
        # the record file stays in pinned HOST memory: every step gathers its B records on the host, copies them
        # H2D (plus the B int32 record indices) and decodes them on the GPU in front of the training step
        pipe.set_resident(False)

        def e2e_steps(k):
            for i in range(k):
                model.train_op()
                out_host.copy_(model._loss[1].buf, non_blocking=True)      # D2H of the step's loss / precision
            torch.cuda.current_stream().synchronize()
            return float(out_host[0])
        h2d = B * pipe.record_bytes + B * 4
    else:
        # the reference's ImageNet input path: decoded uint8 images of arbitrary sizes sit in pinned HOST memory; every
        # step gathers its B images, copies them H2D (~55 MB) and runs the vgg_preprocessing kernel (resize, crop, flip,
        # mean subtraction) straight into the model's input tensor, then the training step, then the loss comes back
        from distributed_tensorflow_resnet_b200 import vgg_preprocessing
        ipipe = vgg_preprocessing.ImagenetPipeline(B, image_size=wl["size"], num_classes=wl["classes"], pool=2 * B, seed=rank,
                                                   resident=False, bind=(images, labels))
        h2d_acc = [0, 0]

        def e2e_steps(k):
            for i in range(k):
                ipipe.next_batch()
                h2d_acc[0] += ipipe.h2d_bytes_last
                h2d_acc[1] += 1
                model.train_op()
                out_host.copy_(model._loss[1].buf, non_blocking=True)      # D2H of the step's loss / precision
            torch.cuda.current_stream().synchronize()
            return float(out_host[0])
        h2d = None
    e2e_steps(2)
    barrier()
    e0.record()
    last_loss = e2e_steps(args.steps)
    e1.record()
    barrier()
    t = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_ms = t.item()
    e2e_value = world * B * args.steps / (e2e_ms / 1e3)
    d2h = 8
    if h2d is None:
        h2d = int(h2d_acc[0] / max(h2d_acc[1], 1))
    if cifar:
        pipe.set_resident(True)

    # ---- sustained leg: seconds of back-to-back steps with the clock sampler running ------------
    sustained = None
    if args.sustained_steps > 0:
        sampler = ClockSampler(local) if rank == 0 else None
        if sampler:
            sampler.start()
        barrier()
        e0.record()
        for _ in range(args.sustained_steps):
            model.train_op()
        e1.record()
        barrier()
        t = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        sclk = sampler.stop() if sampler else None
        sustained = {"steps": args.sustained_steps, "seconds": round(t.item() / 1e3, 3),
                     "ms_per_step": round(t.item() / args.sustained_steps, 4),
                     "value": round(world * B * args.sustained_steps / (t.item() / 1e3), 2), "unit": "images/s",
                     "clocks": sclk}

    # ---- gradient aggregation alone (C5): all-reduce of the flat fp32 fusion buffer, bus bandwidth
    allreduce = None
    if world > 1:
        nbytes = ex.Gd.numel() * 4
        for _ in range(3):
            dist.all_reduce(ex.Gd)
        barrier()
        e0.record()
        for _ in range(10):
            dist.all_reduce(ex.Gd)
        e1.record()
        barrier()
        t = torch.tensor([e0.elapsed_time(e1) / 10], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ar_ms = t.item()
        allreduce = {"bytes": nbytes, "ms": round(ar_ms, 4), "replicas_to_aggregate": world,
                     "bus_gbs": round(2.0 * (world - 1) / world * nbytes / ar_ms / 1e6, 1), "nvlink_nominal_gbs": 900.0,
                     "buckets_in_step": len(ex.buckets)}
        ex.Gd.zero_()

    ddp = ddp_check(ex, model, world, rank, dev) if world > 1 else None

    line = None
    if rank == 0:
        # ---- per-kernel roofline (un-captured launch list, CUDA events per launch) ------------
        per_launch = profile_step(ex, model.graph, iters=2)
        if args.dump_layers:
            cf = conv_flops(model.graph)
            plans = ex.describe_plans()
            with open(args.dump_layers, "w") as f:
                for name, ms_ in per_launch.items():
                    fam, _, key = name.partition(":")
                    extra = ""
                    if key in cf and fam.startswith("conv_"):
                        n_ = [n for n in model.graph.nodes if n.op == "conv2d" and n.outputs[0].name == key][0]
                        extra = " %8.1f TFLOP/s  in %s k%d s%d -> %s" % (cf[key][0] / ms_ / 1e9, n_.inputs[0].shape,
                                                                         n_.attrs["k"], n_.attrs["stride"], n_.outputs[0].shape)
                    else:
                        b_ = hbm_bytes(name, ex, model.graph)
                        if b_:
                            extra = " %8.1f GB/s" % (b_ / ms_ / 1e6)
                    f.write("%-44s %9.4f ms%s%s\n" % (name, ms_, extra, ("  [" + plans[name] + "]") if name in plans else ""))
        roof, table, prof_total = roofline_report(ex, model.graph, per_launch, pk, clocks, args.workload, B)
        cpu = None
        if not args.no_cpu_baseline and world == 1:      # rank 0 at N = 1 only: at N > 1 the other ranks' GPUs would idle
            cb = 8 if args.workload == "imagenet50" else 64
            ips, cdt, threads = cpu_reference_step(args.workload, cb, 2, 1)
            cpu = {"value": round(ips, 3), "unit": "images/s", "cores": threads, "kind": "port",
                   "sample": "2 timed steps of a %d-image batch of the same net/shape (oracle, torch-CPU fp32, %.1f s/step); "
                             "reference TF CPU path unavailable on this box" % (cb, cdt)}
        conv_ms = sum(r["ms"] for r in table if r["kernel"].startswith("conv_"))
        conv_fl = sum(f * (3 if rg else 2) for f, rg in conv_flops(model.graph).values())
        line = {
            "metric": "train images/sec", "value": round(value, 2), "unit": "images/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(ms / args.steps, 4),
            "higher_is_better": True, "scaling": "strong" if strong else "weak",
            "vs_baseline": (round(value / PUBLISHED[(args.workload, world)], 3)
                            if (args.workload, world) in PUBLISHED and B == wl["batch"] and not strong else None),
            "dtype": args.dtype, "data": "synthetic",
            "config": {"workload": args.workload, "description": wl["desc"], "per_gpu_batch": B,
                       "global_batch": B * world, "parallelism": "dp%d" % world,
                       "optimizer": "momentum 0.9 + L2 wd %g, lr 0.1" % wl["wd"],
                       "cuda_graph": not args.no_graph,
                       "cache": "working set (activations + gradients, GBs) far exceeds the 126 MB L2; no explicit flush"
                                if not cifar else "working set %d MB per step vs 126 MB L2; no explicit flush (steady-state training loop)"
                                % (sum(t_.buf.numel() * t_.buf.element_size() for t_ in model.graph.tensors.values() if t_.buf is not None) >> 20),
                       "input": ("synthetic CIFAR records through rn_cifar_decode every step (HBM-resident for value, "
                                 "pinned host memory for e2e)") if cifar else
                                ("value: one synthetic standard-normal fp32 batch resident in HBM; e2e: synthetic uint8 images of "
                                 "random sizes in pinned host memory -> H2D -> rn_vgg_preprocess every step"),
                       "conv_impl": sorted({family(n) for n, _ in ex.fwd + ex.bwd if n.startswith("conv_")})},
            "e2e": {"value": round(e2e_value, 2), "unit": "images/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": round(e2e_ms / args.steps, 4), "last_loss": last_loss},
            "gpu_launches": int(ex.count_launches() * args.steps),
            "gpu_launches_outside_graph": int(launched_live),
            "clocks": clocks,
            "roofline": roof,
            "allreduce": allreduce,
            "ddp_check": ddp,
            "sustained": sustained,
            "conv": {"tflops_all_conv_kernels": round(conv_fl / conv_ms / 1e9, 2) if conv_ms else None,
                     "frac_of_sustained_peak": round(conv_fl / conv_ms / 1e9 / pk["tf_sust"], 4) if conv_ms else None,
                     "frac_of_burst_peak": round(conv_fl / conv_ms / 1e9 / pk["tf_burst"], 4) if conv_ms else None,
                     "gflop_per_step": round(conv_fl / 1e9, 2),
                     "layerwise_roofline": conv_layerwise_roofline(ex, model.graph, per_launch, pk)},
            "cifar_decode": decode_throughput(pipe, pk) if cifar else None,
            "kernels": table,
            "profiled_step_ms": round(prof_total, 3),
            "cpu_baseline": cpu,
            "loss": {"xent": ex.scalars()[0], "cost": ex.scalars()[2]},
        }
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if line is not None:
        print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="imagenet50", choices=sorted(WORKLOADS))
    ap.add_argument("--dtype", default="bf16", choices=["bf16", "f32"])
    ap.add_argument("--batch", type=int, default=0, help="per-GPU batch (default: the workload's)")
    ap.add_argument("--global-batch", type=int, default=0,
                    help="fixed GLOBAL batch split over the ranks (strong scaling; BASELINE config 2 uses 128)")
    ap.add_argument("--sustained-steps", type=int, default=-1,
                    help="extra timed leg of this many back-to-back steps (default: ~3 s worth; 0 disables)")
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--dump-layers", default="", help="write the per-launch timing table to this file")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "ours":
        args.warmup = 3
    if args.sustained_steps < 0:
        args.sustained_steps = 300 if args.workload in ("imagenet50", "wrn28_10") else 1000
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
