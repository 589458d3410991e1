"""GPU parity tests of the whole training step through the reference-facing API
(HParams / ResNet / build_graph / train_op) against the CPU oracle: per-layer activations and
gradients by TF op name, parameter gradients, optimizer result, BN moving statistics, the
reference's real trained weights in inference mode, and the loss trajectory."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from util import rel_err, TOL  # noqa: E402
import distributed_tensorflow_resnet_b200 as pkg  # noqa: E402
from distributed_tensorflow_resnet_b200 import resnet_model, FLAGS, cifar_input  # noqa: E402
from oracle import resnet_oracle as ro  # noqa: E402
from oracle import cifar_oracle as co  # noqa: E402
from oracle import fixtures  # noqa: E402


def _mk_model(hps, x, y, mode="train", **flags):
    FLAGS.reset()
    FLAGS.update(use_cuda_graph=False, **flags)
    dt = torch.bfloat16 if FLAGS.compute_dtype == "bf16" else torch.float32
    X = x.to(dt).cuda().contiguous()
    Y = y.float().cuda().contiguous()
    m = resnet_model.ResNet(hps, X, Y, mode)
    return m


def _oracle_for(kind, hps, params, dtype=torch.float64, **kw):
    if kind == "official_cifar":
        return ro.OfficialResNetV2("cifar10", kw.get("resnet_size", 50), hps.num_classes, dtype=dtype, params=params)
    if kind == "official_imagenet":
        return ro.OfficialResNetV2("imagenet", kw.get("resnet_size", 50), hps.num_classes, dtype=dtype, params=params)
    return ro.UpstreamResNet(hps.num_classes, hps.num_residual_units, hps.use_bottleneck, hps.relu_leakiness,
                             dtype=dtype, params=params, filters=kw.get("filters"))


def _op_name_map(model):
    """product tensor name -> oracle tape name (identical except the BN pre-activation)."""
    return {t.name: t.name for t in model.graph.tensors.values()}


CASES = [
    ("official_cifar", dict(dataset="cifar10", resnet_size=14), (8, 32, 32, 3), 10, None),
    ("official_imagenet", dict(dataset="imagenet", resnet_size=50), (2, 64, 64, 3), 16, None),
    ("upstream_basic", dict(dataset="cifar10"), (8, 32, 32, 3), 10, dict(num_residual_units=2, use_bottleneck=False, relu_leakiness=0.1)),
    ("upstream_bottleneck", dict(dataset="cifar10"), (8, 32, 32, 3), 10, dict(num_residual_units=1, use_bottleneck=True, relu_leakiness=0.1)),
]


@pytest.mark.parametrize("mode", ["f32", "bf16"])
@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_training_step_parity(case, mode):
    kind, flags, shape, classes, legacy = case
    g = torch.Generator().manual_seed(11)
    x = torch.randn(*shape, generator=g)
    lab = torch.arange(shape[0]) % classes
    y = torch.eye(classes)[lab]
    hkw = dict(num_classes=classes, lrn_rate=0.1, weight_decay_rate=2e-4, optimizer='mom')
    if legacy:
        hkw.update(batch_size=shape[0], min_lrn_rate=1e-4, **legacy)
    hps = resnet_model.HParams(**hkw)
    model = _mk_model(hps, x, y, compute_dtype=mode, seed=3, **flags)
    model.build_graph()
    ex = model.executor
    params = {k: v for k, v in ex.state_dict().items() if not k.endswith("/Momentum") and k != "global_step"}
    if mode == "bf16":      # the CUDA path multiplies bf16-rounded weights and inputs
        x_or = x.to(torch.bfloat16).float()
    else:
        x_or = x
    net = _oracle_for(kind, hps, params, **flags)
    out = ro.loss_and_grads(net, x_or.double(), y.double(), 2e-4, True)
    assert list(net.params.keys()) == [v.name for v in model.graph.trainable_variables()]
    assert [net.decay_mask[n] for n in net.params] == [v.decay for v in model.graph.trainable_variables()]
    model.train_op()
    torch.cuda.synchronize()
    tol = TOL[mode]
    # --- per-layer activations (by TF op name)
    checked = 0
    worst = 0.0
    for name, t in model.graph.tensors.items():
        if name in net.acts and t.buf is not None and name not in ("X", "Y"):
            e = rel_err(t.buf.float(), net.acts[name].val)
            worst = max(worst, e)
            assert e < tol, (name, e)
            checked += 1
    assert checked >= 10
    xent, prec, cost = ex.scalars()
    assert abs(xent - float(out["xent"])) < (1e-4 if mode == "f32" else 3e-2) * abs(float(out["xent"]))
    assert abs(cost - float(out["cost"])) < (1e-4 if mode == "f32" else 3e-2) * abs(float(out["cost"]))
    # --- per-layer activation gradients
    gchecked = 0
    for name, t in model.graph.tensors.items():
        slot = getattr(t, "gradslot", None)
        if slot is None or name not in net.acts or net.acts[name].grad is None:
            continue
        if len(slot.members) > 1 and slot.members[-1] is not t:
            continue          # shared residual-stream buffer: holds the gradient of its LAST member
        e = rel_err(slot.buf.float(), net.acts[name].grad)
        assert e < tol, (name, e)
        gchecked += 1
    assert gchecked >= 5
    # --- parameter gradients (pre-aggregation, without the wd term: G holds the raw backprop sum)
    for v in model.graph.trainable_variables():
        gw = ex.var_view(v.name, "G")
        want = net.param_grads[v.name]
        e = rel_err(gw, want)
        assert e < tol, (v.name, e)
    # --- optimizer result and BN moving statistics
    ro.apply_bn_updates(net, 0.997 if not legacy else 0.9)
    opt = ro.MomentumSGD(net, "mom")
    if mode == "f32":
        opt.apply(net, out["grads"], 0.1)
        for v in model.graph.trainable_variables():
            assert rel_err(ex.var_view(v.name), net.params[v.name]) < 1e-5, v.name
            assert rel_err(ex.var_view(v.name, "M"), opt.slots[v.name]) < 1e-4, v.name
        for v in model.graph.variables.values():
            if not v.trainable:
                assert rel_err(ex.var_view(v.name), net.state[v.name]) < 1e-4, v.name
    assert model.global_step == 1


def test_frozen_reference_weights_inference():
    """Known-answer test with the REAL trained weights of the reference
    (test/.../resnet50_cifar_frozen_model_eval.pb -> tests/golden/cifar50_frozen_weights.npz):
    eval-mode forward (BN moving statistics) of the product path == oracle, fp32."""
    w = fixtures.frozen_weights()
    rec = co.synthetic_records(64, "cifar10", 0)
    img, onehot, lab, _ = co.decode(rec, np.arange(64), train=False)
    x, y = torch.from_numpy(img), torch.from_numpy(onehot)
    hps = resnet_model.HParams(num_classes=10, lrn_rate=0.1, weight_decay_rate=2e-4, optimizer='mom')
    model = _mk_model(hps, x, y, mode="eval", compute_dtype="f32", dataset="cifar10", resnet_size=50)
    model.init_values = w
    model.build_graph(istrain=False)
    model.eval_op()
    torch.cuda.synchronize()
    net = ro.OfficialResNetV2("cifar10", 50, 10, dtype=torch.float64, params=w)
    lg = net.forward(x.double(), False)
    assert rel_err(model.logits, lg.val) < 1e-4
    _, probs = net.softmax_xent(lg, y.double())
    assert rel_err(model.predictions, probs) < 1e-4
    assert torch.equal(model.predictions.argmax(1).cpu(), probs.argmax(1))
    for name in ("conv2d_51/Conv2D", "Relu_48", "add_23"):
        assert rel_err(model.graph.tensors[name].buf, net.acts[name].val) < 1e-4


@pytest.mark.parametrize("use_graph", [False, True])
def test_loss_trajectory_resnet20(use_graph):
    """C1: ResNet-20 (upstream builders, 3 units), momentum SGD lr 0.1 wd 2e-4, synthetic CIFAR
    records through the GPU decode kernel; loss trajectory vs the fp32 CPU oracle fed the oracle's
    own decode of the same records and draws.  Batch 32 x 40 steps keeps the CPU side short; the
    fp32 paths agree to 1e-3 after 40 steps of chaotic SGD dynamics."""
    B, steps = 32, 40
    FLAGS.reset()
    FLAGS.update(compute_dtype="f32", use_cuda_graph=use_graph, dataset="cifar10", seed=5)
    hps = resnet_model.HParams(batch_size=B, num_classes=10, min_lrn_rate=1e-4, lrn_rate=0.1, num_residual_units=3,
                               use_bottleneck=False, weight_decay_rate=2e-4, relu_leakiness=0.1, optimizer='mom')
    images, labels = cifar_input.input_fn(True, "synthetic:2048", B, dataset="cifar10", seed=5)
    pipe = images._rn_pipeline
    model = resnet_model.ResNet(hps, images, labels, "train")
    model.build_graph()
    ex = model.executor
    params = {k: v for k, v in ex.state_dict().items() if not k.endswith("/Momentum") and k != "global_step"}
    net = ro.UpstreamResNet(10, 3, False, 0.1, dtype=torch.float32, params=params)
    tr = ro.Trainer(net, 2e-4, "mom", bn_decay=0.9)
    rec = co.synthetic_records(2048, "cifar10", 5)
    # replay the pipeline's index stream on the host
    shadow = cifar_input.RecordPipeline.__new__(cifar_input.RecordPipeline)
    shadow.__dict__.update(shuffle=True, seed=5, n=2048, rank=0, world=1, B=B, epoch=0, _cursor=0)
    shadow._order = shadow._epoch_order(0)
    got, want = [], []
    for s in range(steps):
        idx = shadow.next_indices()
        img, onehot, _, _ = co.decode(rec, idx, train=True, pad=4, epoch=shadow.epoch, seed=5)
        o = tr.step([(torch.from_numpy(img), torch.from_numpy(onehot))], 0.1)
        want.append(float(o[0]["cost"]))
        model.lrn_rate = 0.1
        model.train_op()
        got.append(model.cost)
    got, want = np.array(got), np.array(want)
    assert np.all(np.isfinite(got))
    assert np.abs(got[:5] - want[:5]).max() < 1e-4 * want[0]
    assert np.abs(got - want).max() < 2e-2 * want.max(), (got[-5:], want[-5:])
    assert want[-1] < want[0]            # it trains


def test_sgd_optimizer_and_eval_mode_errors():
    hps = resnet_model.HParams(num_classes=10, lrn_rate=0.05, weight_decay_rate=1e-4, optimizer='sgd')
    x = torch.randn(4, 32, 32, 3)
    y = torch.eye(10)[torch.arange(4)]
    model = _mk_model(hps, x, y, compute_dtype="f32", dataset="cifar10", resnet_size=8, seed=1)
    model.build_graph()
    ex = model.executor
    w0 = ex.W.clone()
    model.train_op()
    torch.cuda.synchronize()
    g = ex.Gd + 1e-4 * w0                 # sgd: w -= lr * (g + wd w)
    assert rel_err(ex.W, w0 - 0.05 * g) < 1e-6
    assert ex.Mom.abs().max().item() == 0
    with pytest.raises(ValueError):
        resnet_model.ResNet(resnet_model.HParams(10, 0.1, 1e-4, 'adam'), x.cuda(), y.cuda(), 'train').build_graph()
    FLAGS.update(dataset="mnist")
    with pytest.raises(ValueError):
        resnet_model.ResNet(hps, x.cuda(), y.cuda(), 'train').build_graph()
    FLAGS.reset()


def test_checkpoint_roundtrip(tmp_path):
    from distributed_tensorflow_resnet_b200 import resnet_cifar_train as ct
    hps = resnet_model.HParams(num_classes=10, lrn_rate=0.1, weight_decay_rate=2e-4, optimizer='mom')
    x, y = torch.randn(4, 32, 32, 3), torch.eye(10)[torch.arange(4)]
    m1 = _mk_model(hps, x, y, compute_dtype="bf16", dataset="cifar10", resnet_size=8, seed=1)
    m1.build_graph()
    for _ in range(3):
        m1.train_op()
    ct.save_checkpoint(m1, str(tmp_path), m1.global_step)
    names = set(m1.executor.state_dict().keys())
    assert "conv2d/kernel" in names and "conv2d/kernel/Momentum" in names and "batch_normalization/moving_mean" in names
    m2 = _mk_model(hps, x, y, compute_dtype="bf16", dataset="cifar10", resnet_size=8, seed=99)
    m2.build_graph()
    assert ct.restore_latest(m2, str(tmp_path)) == "model.ckpt-3.npz"
    assert m2.global_step == 3
    assert torch.equal(m1.executor.W, m2.executor.W) and torch.equal(m1.executor.Mom, m2.executor.Mom)
    assert torch.equal(m1.executor.S, m2.executor.S) and torch.equal(m1.executor.Wb, m2.executor.Wb)
    m1.train_op(); m2.train_op()
    torch.cuda.synchronize()
    assert torch.equal(m1.executor.W, m2.executor.W)           # bit-identical continuation (deterministic kernels)
