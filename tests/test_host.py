"""CPU tests of the host side: C-ABI exports, graph builders vs the reference fixtures, flags,
trainer plumbing, and the world_size-2 aggregation semantics over gloo."""
import ctypes
import os
import re
import subprocess
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    """A TCP port nobody is listening on right now (a fixed pid-derived port collides when two torchrun
    rendezvous of one pytest process follow each other closely)."""
    import socket
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def test_cabi_library_exports_every_declared_symbol():
    from distributed_tensorflow_resnet_b200 import _lib
    hdr = open(os.path.join(ROOT, "include", "rn_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(rn_[a-z0-9_]+)\s*\(", hdr))
    declared -= {"rn_status", "rn_dtype"}
    assert len(declared) >= 30
    so = ctypes.CDLL(_lib.SO_PATH)
    for name in sorted(declared):
        assert hasattr(so, name), "symbol %s declared in include/rn_b200.h is not exported" % name
    assert declared == set(_lib.SIGNATURES), (declared ^ set(_lib.SIGNATURES))
    assert so.rn_abi_version() == _lib.ABI_VERSION == 5


def test_cabi_argument_errors_without_gpu():
    """Argument validation happens before any CUDA call, so it can be exercised on CPU."""
    from distributed_tensorflow_resnet_b200._lib import lib, ConvDesc, last_error
    d = ConvDesc(1, 8, 8, 4, 8, 8, 4, 3, 1, 1, 7)       # bad dtype
    assert lib.rn_conv2d_fprop(ctypes.byref(d), 1, 1, 1, None) == -1 and "dtype" in last_error()
    d = ConvDesc(1, 8, 8, 4, 9, 9, 4, 3, 1, 0, 0)       # output does not fit
    assert lib.rn_conv2d_fprop(ctypes.byref(d), 1, 1, 1, None) == -1
    assert lib.rn_bn_stats_fwd(None, None, None, 10, 8, 0, 1e-5, 0.9, 1, None, None, None, None, None, 0, None) == -1
    assert lib.rn_cifar_decode(1, 10, 3000, 0, 10, None, 4, 1, 4, 0, 0, 0, None, 0, 1, 0, None, None, None, None) == -1
    assert "record" in last_error()
    assert lib.rn_sgd_momentum_fused(1, 1, 1, None, 100, None, 1, 1.0, 0.0, 0.9, 1, None, None, 0, None) == -1


def test_fusion_support_predicates_on_the_c3_shapes():
    """Which convolutions of ResNet-50 / 224x224 may carry a fused epilogue is pure host logic (no CUDA call):
    fprop fusion needs 64-channel store groups, dgrad fusion additionally stride 1 and <= 1024 channels, the
    space-to-depth stem a stride-2 conv over <= 4 channels with a 7x7 or 8x8 kernel."""
    from distributed_tensorflow_resnet_b200._lib import lib, ConvDesc, RN_BF16, RN_F32
    B = ctypes.byref

    def desc(h, cin, cout, k, s, dt=RN_BF16):
        ho = h // s
        return ConvDesc(128, h, h, cin, ho, ho, cout, k, s, (k - 1) // 2, dt)
    assert lib.rn_convtc_fused_supported(B(desc(56, 64, 256, 1, 1))) == 1         # conv3 + shortcut + statistics
    assert lib.rn_convtc_fused_supported(B(desc(56, 64, 64, 3, 1))) == 1
    assert lib.rn_convtc_fused_supported(B(desc(28, 256, 128, 3, 2))) == 1
    assert lib.rn_convtc_fused_supported(B(desc(32, 16, 16, 3, 1))) == 1           # CIFAR width: 16-channel store groups
    assert lib.rn_convtc_fused_supported(B(desc(56, 64, 64, 3, 1, RN_F32))) == 0   # fp32 verification mode: no tensor cores
    assert lib.rn_convtc_fused_dgrad_supported(B(desc(56, 256, 64, 1, 1))) == 1    # dx = gradient of a 256-channel BN
    assert lib.rn_convtc_fused_dgrad_supported(B(desc(14, 1024, 256, 1, 1))) == 1
    assert lib.rn_convtc_fused_dgrad_supported(B(desc(7, 2048, 512, 1, 1))) == 0   # constants table would not fit
    assert lib.rn_convtc_fused_dgrad_supported(B(desc(56, 128, 128, 3, 2))) == 0   # stride-2 dgrad: four sub-GEMMs
    stem = ConvDesc(128, 224, 224, 3, 112, 112, 64, 7, 2, 3, RN_BF16)
    assert lib.rn_stem_s2d_supported(B(stem)) == 1
    h2, w2, t = ctypes.c_int32(), ctypes.c_int32(), ctypes.c_int32()
    assert lib.rn_stem_s2d_geometry(B(stem), B(h2), B(w2), B(t)) == 0 and (h2.value, w2.value, t.value) == (115, 115, 4)
    assert lib.rn_stem_s2d_supported(B(ConvDesc(128, 32, 32, 3, 32, 32, 16, 3, 1, 1, RN_BF16))) == 0   # CIFAR stem: unfold path
    assert lib.rn_convtc_stat_rows(B(stem)) == 148
    assert lib.rn_dense_bwd_scratch_bytes(128, 2048, 1001) > 0 and lib.rn_dense_bwd_scratch_bytes(8, 64, 10) == 0


def _build(dataset, size, shape, classes):
    from distributed_tensorflow_resnet_b200 import graph as G, resnet_model_official as off
    g = G.reset_default_graph()
    x = g.placeholder(shape, G.ACT, "X")
    if dataset == "cifar":
        logits = off.cifar10_resnet_v2_generator(size, classes, "channels_first")(x, True)
    else:
        logits = off.imagenet_resnet_v2(size, classes, "channels_last")(x, True)
    return g, logits


def test_cifar_graph_matches_reference_fixtures():
    from oracle import fixtures
    g, logits = _build("cifar", 50, (128, 32, 32, 3), 10)
    fg, tg = fixtures.frozen_graph(), fixtures.train_graph()
    tv = g.trainable_variables()
    assert [v.name for v in tv] == tg["trainable_variables"]
    assert sum(v.size for v in tv) == 758618 == fg["num_trainable"]
    shapes = {v["name"]: v["shape"] for v in tg["variables"]}
    for v in g.variables.values():
        assert list(v.shape) == shapes[v.name], v.name
    assert all(v.decay for v in tv)                      # L2 over all 152 variables
    # op sequence and conv attributes
    mine = [n for n in g.nodes if n.op in ("conv2d", "bn_act", "add")]
    ref = [o for o in fg["ops"] if o["op"] in ("Conv2D", "FusedBatchNorm", "Add")]
    assert len(mine) == len(ref) == 52 + 49 + 24
    pads = {o["name"]: o["attrs"]["paddings"] for o in fg["ops"] if o["op"] == "Pad"}
    for n, o in zip(mine, ref):
        assert {"conv2d": "Conv2D", "bn_act": "FusedBatchNorm", "add": "Add"}[n.op] == o["op"]
        if n.op == "conv2d":
            assert n.outputs[0].name == o["name"]
            assert list(n.vars["kernel"].shape) == o["attrs"]["kernel_shape"]
            assert n.attrs["stride"] == o["attrs"]["strides"][1]
            if n.attrs["stride"] > 1:
                p = pads[o["inputs"][0]][1][0]
                assert n.attrs["pad"] == p and o["attrs"]["padding"] == "VALID"
        if n.op == "bn_act":
            assert abs(n.attrs["eps"] - o["attrs"]["epsilon"]) < 1e-12 and n.attrs["decay"] == 0.997
        if n.op == "add":       # (main path, shortcut) operand order and wiring
            a, b = n.inputs
            assert [a.name, b.name] == o["inputs"], (n.outputs[0].name, o)
    assert logits.shape == (128, 10)
    for v in tv:
        init = tg["initializers"][v.name]
        if init["kind"] == "truncated_normal":
            assert abs(v.init["stddev"] - init["stddev"]) < 1e-6 * init["stddev"]
        elif init["kind"] == "random_uniform":
            assert abs(v.init["limit"] - init["max"]) < 1e-6
        else:
            assert v.init["kind"] == init["kind"]


def test_imagenet_graph_shapes():
    g, logits = _build("imagenet", 50, (2, 224, 224, 3), 1000)
    tv = g.trainable_variables()
    assert len(tv) == 153 and sum(v.size for v in tv) == 25549352
    convs = [n for n in g.nodes if n.op == "conv2d"]
    assert len(convs) == 53
    stem = convs[0]
    assert stem.attrs == dict(k=7, stride=2, pad=3) and stem.outputs[0].shape == (2, 112, 112, 64)
    mp = [n for n in g.nodes if n.op == "maxpool"][0]
    assert mp.outputs[0].shape == (2, 56, 56, 64) and mp.attrs["pad"] == 0      # TF SAME: 0 before, 1 after
    # stride lives on the 3x3 of the first bottleneck of stages 2-4
    s2 = [n for n in convs if n.attrs["stride"] == 2][1:]
    assert sorted(n.attrs["k"] for n in s2) == [1, 1, 1, 3, 3, 3]
    fwd_flops = sum(2 * n.outputs[0].numel * n.attrs["k"] ** 2 * n.inputs[0].shape[3] for n in convs) / 2
    assert abs(fwd_flops / 1e9 - 8.178) < 0.01           # 8.178 GFLOP / image
    with pytest.raises(ValueError):
        _build("imagenet", 51, (2, 224, 224, 3), 10)
    with pytest.raises(ValueError):
        _build("cifar", 51, (2, 32, 32, 3), 10)


def test_hparams_forms_and_flags():
    from distributed_tensorflow_resnet_b200 import resnet_model, FLAGS
    h4 = resnet_model.HParams(num_classes=10, lrn_rate=0.1, weight_decay_rate=2e-4, optimizer='mom')
    assert h4.num_residual_units is None and h4.optimizer == 'mom'
    h9 = resnet_model.HParams(batch_size=128, num_classes=10, min_lrn_rate=0.0001, lrn_rate=0.1,
                              num_residual_units=5, use_bottleneck=False, weight_decay_rate=0.0002,
                              relu_leakiness=0.1, optimizer='mom')     # resnet_single.py:194-202
    assert h9.num_residual_units == 5 and h9.relu_leakiness == 0.1
    FLAGS.reset()
    assert FLAGS.dataset == "cifar10" and FLAGS.sync_replicas is True and FLAGS.use_horovod is False


def test_trainer_flag_surface_and_schedules():
    import argparse
    from distributed_tensorflow_resnet_b200 import resnet_cifar_train as ct, resnet_imagenet_train as it
    p = argparse.ArgumentParser()
    ct.define_flags(p)
    a = p.parse_args(["--job_name", "worker", "--task_index", "1", "--ps_hosts", "a:1", "--worker_hosts", "b:2,c:3",
                      "--sync_replicas", "True", "--data_format", "channels_last", "--num_gpus", "1",
                      "--batch_size", "64", "--train_steps", "10", "--use_horovod", "true"])
    assert a.batch_size == 64 and a.use_horovod is True and a.task_index == 1
    d = p.parse_args([])
    assert (d.dataset, d.mode, d.image_size, d.eval_batch_count, d.train_steps, d.batch_size) == \
        ("cifar10", "train", 32, 50, 2000, 32)
    assert d.learning_rate == 0.01 and d.sync_replicas is True and d.data_format == "channels_first"
    assert [ct.cifar_learning_rate(s) for s in (0, 39999, 40000, 60000, 80000)] == [0.1, 0.1, 0.01, 0.001, 0.0001]
    assert abs(it.imagenet_learning_rate(6239) - (0.1 + 0.3 * 6239 / 6240.0)) < 1e-12
    assert it.imagenet_learning_rate(6240) == 0.4 and abs(it.imagenet_learning_rate(99840) - 4e-4) < 1e-12
    assert ct.main(["--job_name", "ps"]) is None          # parameter servers exit at once


def test_record_order_sharding():
    """rank r takes records r, r+R, ... of the epoch's permutation; shards are disjoint and cover it."""
    from distributed_tensorflow_resnet_b200.cifar_input import RecordPipeline
    orders = []
    for r in range(2):
        p = RecordPipeline.__new__(RecordPipeline)
        p.shuffle, p.seed, p.n, p.rank, p.world = True, 3, 100, r, 2
        orders.append(p._epoch_order(0))
    assert len(orders[0]) == len(orders[1]) == 50
    assert sorted(np.concatenate(orders).tolist()) == list(range(100))
    p.shuffle = False
    assert p._epoch_order(5).tolist() == list(range(1, 100, 2))


_GLOO_WORKER = r'''
import os, sys, torch, torch.distributed as dist
sys.path.insert(0, sys.argv[1])
from oracle import resnet_oracle as ro
from distributed_tensorflow_resnet_b200.parallel import plan_buckets, allreduce_buckets
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
torch.manual_seed(0)
net = ro.OfficialResNetV2("cifar10", 8, 10, dtype=torch.float64, seed=0)
x = torch.randn(4, 32, 32, 3, dtype=torch.float64); y = torch.eye(10, dtype=torch.float64)[torch.arange(4)]
outs = [ro.loss_and_grads(net, x[2*r:2*r+2], y[2*r:2*r+2], 2e-4, True)["grads"] for r in range(world)]
want = ro.aggregate_mean(outs)
# flat fusion buffer of THIS rank's gradients, 64-element aligned offsets, bucketed all-reduce (sum), then * 1/R
names = list(outs[rank].keys()); offs = []; o = 0
for n in names:
    offs.append(o); o += (outs[rank][n].numel() + 63) // 64 * 64
flat = torch.zeros(o, dtype=torch.float64)
for n, of in zip(names, offs):
    flat[of:of + outs[rank][n].numel()] = outs[rank][n].reshape(-1)
ready = list(range(len(names), 0, -1))
buckets = plan_buckets(offs, o, ready, cap_elems=4096, total_launches=len(names))
assert buckets[0][1] == o and buckets[-1][0] == 0 and all(a[0] == b[1] for a, b in zip(buckets, buckets[1:]))
allreduce_buckets(flat, buckets, None)
flat *= 1.0 / world
for n, of in zip(names, offs):
    got = flat[of:of + want[n].numel()].reshape(want[n].shape)
    assert torch.allclose(got, want[n], rtol=1e-12, atol=1e-14), n
print("rank", rank, "ok", len(buckets))
'''


def test_world_size_2_aggregation_over_gloo(tmp_path):
    """SyncReplicas `replicas_to_aggregate=R` == mean of R gradients: bucketed SUM all-reduce of the
    flat fusion buffer times 1/R equals the oracle's aggregate_mean (gloo, 2 processes, CPU)."""
    script = tmp_path / "w.py"
    script.write_text(_GLOO_WORKER)
    port = _free_port()
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
           "--master-addr", "127.0.0.1", "--master-port", str(port), str(script), ROOT]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    assert "rank 0 ok" in r.stdout and "rank 1 ok" in r.stdout


def test_tf_checkpoint_index_reader_on_the_reference_fixture():
    """tf_io's LevelDB-table / BundleEntryProto reader against the reference's own checkpoint index
    (test/resnet50-cifar-ckpt-20190218/model.ckpt-107738.index, committed unchanged under tests/golden/):
    403 tensor entries, every block checksum valid, names / shapes == the training graph fixture, data
    offsets contiguous in sorted-name order (how BundleWriter lays the shard out)."""
    import json
    from distributed_tensorflow_resnet_b200 import tf_io
    idx = tf_io.read_checkpoint_index(os.path.join(ROOT, "tests", "golden", "ref_model.ckpt-107738"))
    hdr = idx.pop("")
    assert hdr["num_shards"] == 1 and len(idx) == 403
    g = json.load(open(os.path.join(ROOT, "tests", "golden", "cifar50_train_graph.json")))
    shapes = {v["name"]: v["shape"] for v in g["variables"]}
    assert set(shapes) <= set(idx) | {"global_step"} or set(idx) >= set(shapes)
    for name, e in idx.items():
        if name in shapes:
            assert e["shape"] == shapes[name], name
    for name in g["trainable_variables"]:
        assert name in idx and name + "/Momentum" in idx and idx[name + "/Momentum"]["shape"] == idx[name]["shape"]
    assert idx["global_step"]["dtype"] == tf_io.DT_INT64 and idx["global_step"]["shape"] == []
    off = 0
    for name in sorted(idx):
        assert idx[name]["offset"] == off and idx[name]["shard_id"] == 0, name
        off += idx[name]["size"]
    assert idx["conv2d/kernel"]["size"] == 3 * 3 * 3 * 16 * 4
    # the frozen GraphDef of the reference, when the reference tree is present (build container only)
    pb = "/root/reference/test/resnet50-cifar-ckpt-20190218/resnet50_cifar_frozen_model_eval.pb"
    if os.path.exists(pb):
        w = tf_io.read_frozen_graph(pb)
        z = np.load(os.path.join(ROOT, "tests", "golden", "cifar50_frozen_weights.npz"))
        assert set(w) == set(z.files) and all(np.array_equal(w[k], z[k]) for k in z.files)


def test_tf_checkpoint_and_frozen_graph_round_trip(tmp_path):
    from distributed_tensorflow_resnet_b200 import tf_io
    assert tf_io.crc32c(b"123456789") == 0xE3069283 and tf_io.crc32c(bytes(32)) == 0x8A9136AA     # RFC 3720 vectors
    assert tf_io.crc32c(b"6789", tf_io.crc32c(b"12345")) == 0xE3069283                              # incremental
    rng = np.random.default_rng(0)
    t = {"conv2d/kernel": rng.standard_normal((3, 3, 3, 16)).astype(np.float32),
         "conv2d/kernel/Momentum": rng.standard_normal((3, 3, 3, 16)).astype(np.float32),
         "dense/bias": np.arange(10, dtype=np.float32), "global_step": np.int64(107738)}
    t.update({"batch_normalization_%d/gamma" % i: rng.standard_normal(64).astype(np.float32) for i in range(400)})
    prefix = str(tmp_path / "model.ckpt-107738")
    tf_io.write_checkpoint(prefix, t)
    assert sorted(os.listdir(tmp_path)) == ["model.ckpt-107738.data-00000-of-00001", "model.ckpt-107738.index"]
    r = tf_io.read_checkpoint(prefix)
    assert set(r) == set(t) and all(np.array_equal(r[k], t[k]) and r[k].shape == np.shape(t[k]) for k in t)
    raw = bytearray(open(prefix + ".data-00000-of-00001", "rb").read())
    raw[100] ^= 1
    open(prefix + ".data-00000-of-00001", "wb").write(raw)
    with pytest.raises(ValueError, match="checksum"):
        tf_io.read_checkpoint(prefix)
    pb = str(tmp_path / "frozen.pb")
    wts = {k: v for k, v in t.items() if k != "global_step"}
    tf_io.write_frozen_graph(pb, wts)
    back = tf_io.read_frozen_graph(pb)
    assert set(back) == set(wts) and all(np.array_equal(back[k], wts[k]) for k in wts)


def test_summary_writer_produces_tensorboard_events(tmp_path):
    from distributed_tensorflow_resnet_b200.train_utils import SummaryWriter, read_event_scalars
    w = SummaryWriter(str(tmp_path))
    w.add_scalars(100, {"cross_entropy": 2.25, "cost": 2.5, "learning_rate": 0.1, "Precision": 0.125})
    w.add_scalars(200, {"Precision": 0.25, "Best Precision": 0.25})
    w.close()
    assert any("tfevents" in f for f in os.listdir(tmp_path))
    got = read_event_scalars(str(tmp_path))
    assert (100, "cost", 2.5) in got and (200, "Best Precision", 0.25) in got and len(got) == 6
    lines = open(tmp_path / "summaries.jsonl").read().strip().split("\n")
    assert len(lines) == 2 and '"Precision": 0.125' in lines[0]


def test_aggregation_disabled_with_several_ranks_is_refused(tmp_path):
    """A process group with R > 1 ranks and both aggregation flags off would train R diverging replicas."""
    script = tmp_path / "w.py"
    script.write_text(r'''
import os, sys, torch, torch.distributed as dist
sys.path.insert(0, sys.argv[1])
dist.init_process_group("gloo")
from distributed_tensorflow_resnet_b200 import resnet_model, FLAGS
FLAGS.reset(); FLAGS.update(sync_replicas=False, use_horovod=False)
m = resnet_model.ResNet(resnet_model.HParams(10, 0.1, 2e-4, 'mom'), torch.zeros(2, 32, 32, 3), torch.zeros(2, 10), 'train')
try:
    m._world(); print("NO_ERROR")
except ValueError as e:
    print("REFUSED" if "aggregation is off" in str(e) else "WRONG " + str(e))
FLAGS.update(sync_replicas=True, replicas_to_aggregate=1)
try:
    m._world(); print("NO_ERROR")
except ValueError as e:
    print("HINT" if "--nproc-per-node k" in str(e) else "WRONG " + str(e))
dist.destroy_process_group()
''')
    port = _free_port()
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr",
           "127.0.0.1", "--master-port", str(port), str(script), ROOT]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=240)
    assert r.returncode == 0 and r.stdout.count("REFUSED") == 2 and r.stdout.count("HINT") == 2, r.stdout + r.stderr[-3000:]


def test_reference_entry_point_names_exist():
    """Every trainer / eval script name of the reference (resnet_cifar_main.py, resnet_cifar_train_horovod.py,
    resnet_imagenet_main.py, resnet_imagenet_horovod_train.py, resnet_imagenet_eval.py, ...) is a module of the
    package with a main(argv); the Horovod-named ones preset the all-reduce flags, --job_name=ps exits at once."""
    import importlib
    from distributed_tensorflow_resnet_b200 import FLAGS
    for name in ("resnet_cifar_train", "resnet_cifar_main", "resnet_cifar_train_horovod", "resnet_cifar_eval",
                 "resnet_imagenet_train", "resnet_imagenet_main", "resnet_imagenet_horovod_train", "resnet_imagenet_eval",
                 "resnet_single", "resnet_cifar_frozen_model", "resnet_cifar_predict_from_pd"):
        mod = importlib.import_module("distributed_tensorflow_resnet_b200." + name)
        assert callable(getattr(mod, "main", None)) or callable(getattr(mod, "evaluate", None)), name
    FLAGS.reset()
    hv = importlib.import_module("distributed_tensorflow_resnet_b200.resnet_cifar_train_horovod")
    assert hv.main(["--job_name", "ps"]) is None                  # parsed the preset flags, then left (no parameter servers)
    assert FLAGS.use_horovod and FLAGS.variable_update == "horovod"
    FLAGS.reset()
