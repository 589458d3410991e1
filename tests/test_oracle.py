"""CPU tests: the oracle against the reference's golden fixtures and known-answer vectors."""
import math

import numpy as np
import pytest
import torch

from oracle import cifar_oracle as co
from oracle import fixtures
from oracle import resnet_oracle as ro


def test_philox_known_answers():
    # Random123 kat_vectors for philox4x32-10
    kat = [((0, 0, 0, 0), (0, 0), (0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8)),
           ((0xffffffff,) * 4, (0xffffffff, 0xffffffff), (0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd)),
           ((0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344), (0xa4093822, 0x299f31d0),
            (0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1))]
    for ctr, key, want in kat:
        got = co.philox4x32_10(np.array([ctr], dtype=np.uint32), key)[0]
        assert tuple(int(x) for x in got) == want


def test_frozen_graph_structure_matches_oracle_enumeration():
    fg = fixtures.frozen_graph()
    assert fg["op_histogram"]["Conv2D"] == 52 and fg["op_histogram"]["FusedBatchNorm"] == 49
    assert fg["op_histogram"]["Add"] == 24 and fg["op_histogram"]["Pad"] == 4
    assert fg["num_trainable"] == 758618
    w = fixtures.frozen_weights()
    net = ro.OfficialResNetV2("cifar10", 50, 10, dtype=torch.float64, params=w)
    net.forward(torch.zeros(1, 32, 32, 3, dtype=torch.float64), False)
    names = [o["name"] for o in fg["ops"] if o["op"] in ("Conv2D", "FusedBatchNorm", "Relu", "Add", "AvgPool", "BiasAdd")]
    assert names == [n for n in net.acts if n != "X"]
    # Add wiring: (main path conv, shortcut) exactly as the fixture lists them
    for o in fg["ops"]:
        if o["op"] == "FusedBatchNorm":
            assert abs(o["attrs"]["epsilon"] - ro.BN_EPS_OFFICIAL) < 1e-12
    assert len(net.params) == 152 and sum(v.numel() for v in net.params.values()) == 758618
    # kernel shapes / strides / paddings
    convs = [o for o in fg["ops"] if o["op"] == "Conv2D"]
    pads = {o["name"]: o for o in fg["ops"] if o["op"] == "Pad"}
    for o in convs:
        k = o["attrs"]["kernel_shape"]
        assert list(net.params[o["name"].replace("/Conv2D", "/kernel")].shape) == k
        s = o["attrs"]["strides"][1]
        if s == 1:
            assert o["attrs"]["padding"] == "SAME"
        else:
            assert o["attrs"]["padding"] == "VALID" and o["inputs"][0] in pads
            p = (k[0] - 1) // 2
            assert pads[o["inputs"][0]]["attrs"]["paddings"] == [[0, 0], [p, p], [p, p], [0, 0]]


def test_frozen_weights_inference_is_sane():
    """The trained weights must classify: logits of the trained net on standardised noise are
    finite and the BN moving stats are used (inference mode)."""
    w = fixtures.frozen_weights()
    net = ro.OfficialResNetV2("cifar10", 50, 10, dtype=torch.float32, params=w)
    x = torch.randn(8, 32, 32, 3, generator=torch.Generator().manual_seed(0))
    lg = net.forward(x, False).val
    assert torch.isfinite(lg).all() and lg.abs().max() < 100


def test_training_graph_constants():
    tg = fixtures.train_graph()
    assert tg["tensorflow_version"] == "1.12.0"
    assert tg["apply_momentum"]["count"] == 152 and tg["apply_momentum"]["use_nesterov"] == [False]
    assert abs(tg["apply_momentum"]["momentum_const"][0] - 0.9) < 1e-7
    assert tg["l2loss_count"] == 152                      # weight decay over ALL trainable variables
    assert abs(tg["weight_decay_consts"][0] - 2e-4) < 1e-9
    assert tg["assign_sub_count"] == 98
    assert abs(tg["bn_moving_average_decay"][0] - ro.BN_DECAY_OFFICIAL) < 1e-7
    assert tg["fused_batch_norm"]["is_training"] == [True]
    # per-variable gradient = AddN(wd * v, backprop)
    assert sum(tg["grad_addn_structures"].values()) == 152
    net = ro.OfficialResNetV2("cifar10", 50, 10, dtype=torch.float32, seed=0)
    net.forward(torch.zeros(1, 32, 32, 3), True)
    assert list(net.params.keys()) == tg["trainable_variables"]
    assert all(net.decay_mask.values())
    # initialiser constants
    for name, init in tg["initializers"].items():
        if init["kind"] == "truncated_normal":
            k = net.params[name].shape
            want = math.sqrt(1.0 / (k[0] * k[1] * k[2])) / 0.87962566103423978
            assert abs(init["stddev"] - want) / want < 1e-6
        if init["kind"] == "random_uniform":
            k = net.params[name].shape
            assert abs(init["max"] - math.sqrt(6.0 / (k[0] + k[1]))) < 1e-6


def test_input_function_chain_constants():
    """The decode oracle follows the verbatim tf.data function bodies of the fixture."""
    tg = fixtures.train_graph()
    fn = {k: v for k, v in tg["functions"].items() if "wrapper" in k}
    parse = [f for f in fn.values() if any(n["op"] == "DecodeRaw" for n in f["nodes"])][0]
    prep = [f for f in fn.values() if any(n["op"] == "RandomUniformInt" for n in f["nodes"])][0]
    consts = {n["name"]: n.get("value") for n in parse["nodes"] if n["op"] == "Const"}
    assert consts["one_hot/depth"] == 10 and consts["strided_slice_1/stack_1"] == [3073]
    assert consts["Reshape/shape"] == [3, 32, 32] and consts["transpose/perm"] == [1, 2, 0]
    c2 = {n["name"]: n.get("value") for n in prep["nodes"] if n["op"] == "Const"}
    assert c2["resize_image_with_crop_or_pad/pad_to_bounding_box/stack"] == [0, 0, 4, 4, 4, 4, 0, 0]
    assert c2["random_crop/Shape"] == [40, 40, 3] and c2["random_crop/size"] == [32, 32, 3]
    assert c2["random_crop/random_uniform/max"] == 2147483647
    assert c2["random_flip_left_right/Less/y"] == 0.5
    ops = [n["op"] for n in prep["nodes"]]
    for needed in ("FloorMod", "ReverseV2", "Relu", "Sqrt", "Rsqrt", "Maximum", "RealDiv"):
        assert needed in ops
    # draws -> indices mapping
    raw = np.array([[2147483647, 9, 0, 0x00400000], [5, 2147483650, 0, 0x003fffff]], dtype=np.uint32)
    oy, ox, fl = co.draws_to_indices(raw, 4)
    assert list(oy) == [0, 5] and list(ox) == [0, 3]      # 2^31-1 wraps to 0; (2^31+2) % (2^31-1) = 3
    assert list(fl) == [0, 1]                              # u = 0.5 -> no flip; just below -> flip


def test_decode_oracle_properties():
    rec = co.synthetic_records(64, "cifar10", 3)
    idx = np.arange(32)
    img, onehot, lab, (oy, ox, fl) = co.decode(rec, idx, train=True, pad=4, epoch=1, seed=7)
    assert img.shape == (32, 32, 32, 3) and onehot.shape == (32, 10)
    assert np.all(onehot.argmax(1) == rec[idx, 0])
    assert np.allclose(img.reshape(32, -1).mean(1), 0, atol=1e-5)
    assert oy.min() >= 0 and oy.max() <= 8 and ox.max() <= 8
    # explicit draws reproduce, flip is an exact mirror
    a, *_ = co.decode(rec, idx, train=True, pad=4, crop_yx=(oy, ox), flip=fl)
    assert np.array_equal(a, img)
    b, *_ = co.decode(rec, idx, train=True, pad=4, crop_yx=(oy, ox), flip=1 - fl)
    assert np.allclose(b[:, :, ::-1], img, atol=1e-6)
    # eval mode = standardise only; cifar100 label offset
    e, *_ = co.decode(rec, idx, train=False)
    x = rec[idx, 1:].reshape(32, 3, 32, 32).transpose(0, 2, 3, 1).astype(np.float64)
    want = (x - x.mean((1, 2, 3), keepdims=True)) / np.maximum(x.std((1, 2, 3), keepdims=True), 1 / math.sqrt(3072))
    assert np.allclose(e, want, atol=1e-5)
    r100 = co.synthetic_records(16, "cifar100", 0)
    _, oh, lab100, _ = co.decode(r100, np.arange(16), dataset="cifar100", train=False)
    assert oh.shape == (16, 100) and np.all(lab100 == r100[:, 1])
    # constant image: variance 0 -> divide by 1/sqrt(3072), result exactly 0
    z = np.zeros((1, 3073), np.uint8)
    z[0, 1:] = 77
    c, *_ = co.decode(z, [0], train=False)
    assert np.all(c == 0)
    with pytest.raises(ValueError):
        co.decode(rec, idx, dataset="mnist")


@pytest.mark.parametrize("net_kind", ["official_cifar", "official_imagenet", "upstream_basic", "upstream_bottleneck"])
def test_oracle_backward_matches_autograd(net_kind):
    torch.manual_seed(0)
    dt = torch.float64
    if net_kind == "official_cifar":
        mk = lambda p=None: ro.OfficialResNetV2("cifar10", 8, 10, dtype=dt, seed=1, params=p)
        x = torch.randn(4, 32, 32, 3, dtype=dt)
        classes = 10
    elif net_kind == "official_imagenet":
        mk = lambda p=None: ro.OfficialResNetV2("imagenet", 50, 12, dtype=dt, seed=1, params=p)
        x = torch.randn(2, 64, 64, 3, dtype=dt)
        classes = 12
    elif net_kind == "upstream_basic":
        mk = lambda p=None: ro.UpstreamResNet(10, 1, False, 0.1, dtype=dt, seed=1, params=p)
        x = torch.randn(4, 32, 32, 3, dtype=dt)
        classes = 10
    else:
        mk = lambda p=None: ro.UpstreamResNet(10, 1, True, 0.1, dtype=dt, seed=1, params=p)
        x = torch.randn(4, 32, 32, 3, dtype=dt)
        classes = 10
    y = torch.eye(classes, dtype=dt)[torch.arange(x.shape[0]) % classes]
    net = mk()
    out = ro.loss_and_grads(net, x, y, 2e-4, True)
    ps = {n: v.clone().requires_grad_(True) for n, v in net.params.items()}
    net2 = mk({})
    net2.params = type(net.params)(ps)
    net2.decay_mask = dict(net.decay_mask)
    net2.state = {k: v.clone() for k, v in net.state.items()}
    lg = net2.forward(x, True)
    xe, _ = net2.softmax_xent(lg, y)
    cost = xe + 2e-4 * sum((v * v).sum() / 2 for n, v in ps.items() if net.decay_mask[n])
    cost.backward()
    for n in ps:
        g = ps[n].grad
        assert (g - out["grads"][n]).abs().max() <= 1e-9 * (1 + g.abs().max()), n
    assert abs(float(cost.detach()) - float(out["cost"])) < 1e-12


def test_optimizer_aggregation_and_schedules():
    net = ro.OfficialResNetV2("cifar10", 8, 10, dtype=torch.float64, seed=0)
    tr = ro.Trainer(net, 2e-4, "mom")
    g = torch.Generator().manual_seed(0)
    x = torch.randn(8, 32, 32, 3, dtype=torch.float64, generator=g)
    y = torch.eye(10, dtype=torch.float64)[torch.arange(8) % 10]
    w0 = {n: v.clone() for n, v in net.params.items()}
    outs = tr.step([(x[:4], y[:4]), (x[4:], y[4:])], 0.1)
    mean = ro.aggregate_mean([o["grads"] for o in outs])
    for n in w0:      # first momentum step: m = g, w -= lr * g
        assert torch.allclose(net.params[n], w0[n] - 0.1 * mean[n], atol=1e-12)
    # moving stats moved by (1 - 0.997) towards replica 0's batch stats
    mm = net.state["batch_normalization/moving_mean"]
    assert mm.abs().max() > 0 and mm.abs().max() < 1
    assert ro.cifar_lr(0) == 0.1 and ro.cifar_lr(40000) == 0.01 and ro.cifar_lr(60000) == 0.001 and ro.cifar_lr(80000) == 0.0001
    assert abs(ro.imagenet_lr(0) - 0.1) < 1e-12 and abs(ro.imagenet_lr(3120) - 0.25) < 1e-12
    assert ro.imagenet_lr(6240) == 0.4 and abs(ro.imagenet_lr(37440) - 0.04) < 1e-12
    assert abs(ro.imagenet_lr(74880) - 0.004) < 1e-12 and abs(ro.imagenet_lr(99840) - 0.0004) < 1e-12


def test_parameter_counts():
    """C1 ResNet-20 269,722; C2 758,618; C2-bottleneck 497,690; C3 25,549,352."""
    def count(net, x):
        net.forward(x, True)
        return sum(v.numel() for v in net.params.values())
    x = torch.zeros(1, 32, 32, 3)
    assert count(ro.UpstreamResNet(10, 3, False, 0.1), x) == 269722
    assert count(ro.OfficialResNetV2("cifar10", 50, 10), x) == 758618
    assert count(ro.UpstreamResNet(10, 5, True, 0.1), x) == 497690
    assert count(ro.OfficialResNetV2("imagenet", 50, 1000), torch.zeros(1, 64, 64, 3)) == 25549352
