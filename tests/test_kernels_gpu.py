"""GPU parity tests: every kernel, called through the C ABI, against the CPU oracle on the same
seeded inputs.  Tolerances are the north-star's: 1e-4 relative (fp32 verification mode),
2e-2 (bf16 mode); integer / index work is bit-exact."""
import ctypes as C

import numpy as np
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

from util import DT, TOL, rel_err, dev, stream, lib, check, ptr, ConvDesc, RN_F32, RN_BF16  # noqa: E402
from oracle import resnet_oracle as ro  # noqa: E402
from oracle import cifar_oracle as co  # noqa: E402


def _q(t, mode):
    """Round a reference tensor the way the CUDA path stores it."""
    return t.to(torch.bfloat16).to(t.dtype) if mode == "bf16" else t


# ------------------------------------------------------------------------------------------ conv
CONV_CASES = [
    # N, H, W, Cin, Cout, k, stride, pad, (Ho, Wo)
    (4, 32, 32, 3, 16, 3, 1, 1, (32, 32)),       # CIFAR stem
    (4, 16, 16, 16, 32, 3, 2, 1, (8, 8)),        # official 3x3 s2: symmetric pad + VALID
    (4, 16, 16, 16, 32, 3, 2, 0, (8, 8)),        # upstream 3x3 s2: TF-SAME (pad 0 before, 1 after)
    (4, 16, 16, 16, 32, 1, 2, 0, (8, 8)),        # 1x1 s2 projection
    (2, 8, 8, 64, 64, 3, 1, 1, (8, 8)),
    (2, 28, 28, 24, 40, 1, 1, 0, (28, 28)),      # ragged channel counts / tile edges
    (2, 30, 30, 3, 64, 7, 2, 3, (15, 15)),       # ImageNet stem shape, small
    (1, 7, 7, 128, 96, 3, 1, 1, (7, 7)),
    (3, 9, 11, 5, 7, 3, 1, 1, (9, 11)),          # nothing aligned
]


def _oracle_conv(x, w, k, stride, pad, out_hw):
    net = ro.Net(dtype=torch.float64, params={"w": w.numpy()})
    xv = ro.Var(x.double())
    y = net.conv2d(xv, "w", k, x.shape[-1], w.shape[-1], stride, pad, out_hw, None, "y")
    return net, xv, y


@pytest.mark.parametrize("mode", ["f32", "bf16"])
@pytest.mark.parametrize("case", CONV_CASES)
def test_conv_fprop_dgrad_wgrad(case, mode):
    N, H, W, Cin, Cout, k, s, pad, (Ho, Wo) = case
    tdt, rdt = DT[mode]
    g = torch.Generator().manual_seed(hash(case) % 2**31)
    x = _q(torch.randn(N, H, W, Cin, generator=g), mode)
    w = _q(torch.randn(k, k, Cin, Cout, generator=g) / (k * k * Cin) ** 0.5, mode)
    dy = _q(torch.randn(N, Ho, Wo, Cout, generator=g), mode)
    net, xv, yv = _oracle_conv(x, w, k, s, pad, (Ho, Wo))
    yv.grad = dy.double()
    net.backward()
    d = ConvDesc(N, H, W, Cin, Ho, Wo, Cout, k, s, pad, rdt)
    X, Wd, DY = dev(x, tdt), dev(w, tdt), dev(dy, tdt)
    Y = torch.empty(N, Ho, Wo, Cout, dtype=tdt, device="cuda")
    check(lib.rn_conv2d_fprop(C.byref(d), ptr(X), ptr(Wd), ptr(Y), stream()), "fprop")
    assert rel_err(Y.float(), yv.val) < TOL[mode]
    DX = torch.full((N, H, W, Cin), 7.0, dtype=tdt, device="cuda")
    check(lib.rn_conv2d_dgrad(C.byref(d), ptr(DY), ptr(Wd), ptr(DX), 0.0, stream()), "dgrad")
    assert rel_err(DX.float(), xv.grad) < TOL[mode]
    base = _q(torch.randn(N, H, W, Cin, generator=g), mode)
    DX2 = dev(base, tdt)
    check(lib.rn_conv2d_dgrad(C.byref(d), ptr(DY), ptr(Wd), ptr(DX2), 1.0, stream()), "dgrad beta=1")
    assert rel_err(DX2.float(), xv.grad + base.double()) < TOL[mode]
    nb = lib.rn_conv2d_wgrad_scratch_bytes(C.byref(d))
    scratch = torch.zeros(max(nb, 16), dtype=torch.uint8, device="cuda")
    DW = torch.full((k, k, Cin, Cout), 3.0, dtype=torch.float32, device="cuda")
    check(lib.rn_conv2d_wgrad(C.byref(d), ptr(X), ptr(DY), ptr(DW), ptr(scratch), nb, stream()), "wgrad")
    assert rel_err(DW, net.param_grads["w"]) < (1e-4 if mode == "f32" else 1e-3)   # inputs already rounded
    if nb > 0:
        assert lib.rn_conv2d_wgrad(C.byref(d), ptr(X), ptr(DY), ptr(DW), ptr(scratch), nb - 1, stream()) == -2


def test_conv_linearity_at_full_size():
    """Size-independent property at the ImageNet stage-3 shape (b128 is too slow for the CPU
    oracle): conv(a*x1 + x2) == a*conv(x1) + conv(x2) in fp32 verification mode."""
    N, H, Cin, Cout = 16, 14, 256, 256
    d = ConvDesc(N, H, H, Cin, H, H, Cout, 3, 1, 1, RN_F32)
    g = torch.Generator().manual_seed(5)
    x1, x2 = dev(torch.randn(N, H, H, Cin, generator=g)), dev(torch.randn(N, H, H, Cin, generator=g))
    w = dev(torch.randn(3, 3, Cin, Cout, generator=g) / 48.0)
    ys = []
    for x in (x1, x2, 0.5 * x1 + x2):
        y = torch.empty(N, H, H, Cout, device="cuda")
        check(lib.rn_conv2d_fprop(C.byref(d), ptr(x.contiguous()), ptr(w), ptr(y), stream()), "fprop")
        ys.append(y)
    assert rel_err(ys[2], 0.5 * ys[0] + ys[1]) < 1e-5


# ------------------------------------------------------------------------------------------ batch norm
BN_CASES = [(128 * 8 * 8, 64), (4 * 32 * 32, 16), (2 * 7 * 7, 2048), (3 * 5 * 5, 160), (1000, 8), (37, 320)]


def _bn_oracle(x, addend, gamma, beta, eps, decay, unbiased, leak, dy):
    net = ro.Net(dtype=torch.float64, params={"s/gamma": gamma.numpy(), "s/beta": beta.numpy(),
                                              "s/moving_mean": np.full(gamma.shape, 0.25),
                                              "s/moving_variance": np.full(gamma.shape, 1.5)})
    net._bn_decay_flag = True
    a = ro.Var(x.double())
    z = a
    if addend is not None:
        b = ro.Var(addend.double())
        z = net.add(a, b, "z")
    y = net.batch_norm_act(z, "s", True, eps, decay, unbiased, leak, ("pre", "y"))
    y.grad = dy.double()
    net.backward()
    ro.apply_bn_updates(net, decay)
    return net, a, z, y


@pytest.mark.parametrize("mode", ["f32", "bf16"])
@pytest.mark.parametrize("fuse_add", [False, True])
@pytest.mark.parametrize("case", BN_CASES)
def test_bn_forward_backward(case, fuse_add, mode):
    M, Cc = case
    if mode == "bf16" and Cc % 8:
        pytest.skip("bf16 vector path needs C % 8 == 0")
    tdt, rdt = DT[mode]
    g = torch.Generator().manual_seed(M * 131 + Cc)
    x = _q(torch.randn(M, Cc, generator=g) * 1.7 + 0.3, mode)
    addend = _q(torch.randn(M, Cc, generator=g), mode) if fuse_add else None
    gamma = torch.rand(Cc, generator=g) + 0.5
    beta = torch.randn(Cc, generator=g) * 0.2
    dy = _q(torch.randn(M, Cc, generator=g), mode)
    eps, decay, leak = 1.001e-5, 0.997, (0.0 if Cc != 160 else 0.1)
    z_ref = _q(x + addend, mode) if fuse_add else x
    # oracle on the value the CUDA path actually normalises (z as stored)
    net, a, z, y = _bn_oracle(z_ref, None, gamma, beta, eps, decay, True, leak, dy)
    X, DY = dev(x, tdt), dev(dy, tdt)
    A = dev(addend, tdt) if fuse_add else None
    Z = torch.empty(M, Cc, dtype=tdt, device="cuda") if fuse_add else None
    Gm, Bt = dev(gamma), dev(beta)
    mean, rstd = torch.empty(Cc, device="cuda"), torch.empty(Cc, device="cuda")
    mm, mv = torch.full((Cc,), 0.25, device="cuda"), torch.full((Cc,), 1.5, device="cuda")
    nb = lib.rn_bn_scratch_bytes(M, Cc)
    scratch = torch.zeros(nb, dtype=torch.uint8, device="cuda")
    for _ in range(2):      # twice: the scratch ticket counters must be left clean
        mm.fill_(0.25); mv.fill_(1.5)
        check(lib.rn_bn_stats_fwd(ptr(X), ptr(A), ptr(Z), M, Cc, rdt, eps, decay, 1, ptr(mean), ptr(rstd), ptr(mm),
                                  ptr(mv), ptr(scratch), nb, stream()), "bn_stats")
    zin = Z if fuse_add else X
    if fuse_add:
        assert torch.equal(Z.float().cpu(), z_ref.float())
    x2 = z_ref.double()
    assert rel_err(mean, x2.mean(0)) < 1e-5
    assert rel_err(rstd, 1 / torch.sqrt(x2.var(0, unbiased=False) + eps)) < 1e-5
    assert rel_err(mm, net.state["s/moving_mean"]) < 1e-5 and rel_err(mv, net.state["s/moving_variance"]) < 1e-5
    Y = torch.empty(M, Cc, dtype=tdt, device="cuda")
    check(lib.rn_bn_apply_act_fwd(ptr(zin), ptr(Y), M, Cc, rdt, ptr(Gm), ptr(Bt), ptr(mean), ptr(rstd), 0, eps, leak,
                                  stream()), "bn_apply")
    assert rel_err(Y.float(), y.val) < TOL[mode]
    dG, dB = torch.empty(Cc, device="cuda"), torch.empty(Cc, device="cuda")
    check(lib.rn_bn_act_bwd_reduce(ptr(DY), ptr(zin), M, Cc, rdt, ptr(Gm), ptr(Bt), ptr(mean), ptr(rstd), leak, ptr(dG),
                                   ptr(dB), ptr(scratch), nb, stream()), "bn_bwd_reduce")
    tolg = 1e-4 if mode == "f32" else 2e-2
    assert rel_err(dG, net.param_grads["s/gamma"]) < tolg and rel_err(dB, net.param_grads["s/beta"]) < tolg
    DX = torch.empty(M, Cc, dtype=tdt, device="cuda")
    check(lib.rn_bn_act_bwd_apply(ptr(DY), ptr(zin), ptr(DX), None, M, Cc, rdt, ptr(Gm), ptr(Bt), ptr(mean), ptr(rstd),
                                  ptr(dG), ptr(dB), leak, stream()), "bn_bwd_apply")
    assert rel_err(DX.float(), a.grad) < TOL[mode]
    # in-place accumulate (the residual-stream gradient): dx += bn_dx
    base = _q(torch.randn(M, Cc, generator=g), mode)
    ACC = dev(base, tdt)
    check(lib.rn_bn_act_bwd_apply(ptr(DY), ptr(zin), ptr(ACC), ptr(ACC), M, Cc, rdt, ptr(Gm), ptr(Bt), ptr(mean),
                                  ptr(rstd), ptr(dG), ptr(dB), leak, stream()), "bn_bwd_apply add")
    assert rel_err(ACC.float(), a.grad + base.double()) < TOL[mode]


def test_bn_inference_mode_and_errors():
    M, Cc = 512, 32
    g = torch.Generator().manual_seed(1)
    x, gamma, beta = torch.randn(M, Cc, generator=g), torch.rand(Cc, generator=g) + 0.5, torch.randn(Cc, generator=g)
    mm, mv = torch.randn(Cc, generator=g), torch.rand(Cc, generator=g) + 0.1
    want = F.relu((x.double() - mm) / torch.sqrt(mv.double() + 1.001e-5) * gamma + beta)
    Y = torch.empty(M, Cc, device="cuda")
    X, Gm, Bt, MM, MV = dev(x), dev(gamma), dev(beta), dev(mm), dev(mv)      # keep the device buffers alive
    check(lib.rn_bn_apply_act_fwd(ptr(X), ptr(Y), M, Cc, RN_F32, ptr(Gm), ptr(Bt), ptr(MM), ptr(MV), 1, 1.001e-5, 0.0,
                                  stream()), "bn_infer")
    assert rel_err(Y, want) < 1e-5
    assert lib.rn_bn_apply_act_fwd(ptr(Y), ptr(Y), M, 30, RN_BF16, 1, 1, 1, 1, 0, 0.0, 0.0, stream()) == -4
    small = torch.zeros(64, dtype=torch.uint8, device="cuda")
    assert lib.rn_bn_stats_fwd(ptr(Y), None, None, M, Cc, RN_F32, 1e-5, 0.9, 1, ptr(Y), ptr(Y), None, None, ptr(small),
                               64, stream()) == -2


def test_bn_statistics_at_full_size():
    """Largest C3 BN tensor (128x56x56x256 bf16, 205 MB): statistics of a tensor with known
    per-channel mean/variance, and idempotence of the scratch protocol."""
    M, Cc = 128 * 56 * 56, 256
    ch = torch.arange(Cc, device="cuda", dtype=torch.float32)
    x = (torch.randn(M, Cc, device="cuda") * (1 + ch / 256) + ch / 64).to(torch.bfloat16)
    mean, rstd = torch.empty(Cc, device="cuda"), torch.empty(Cc, device="cuda")
    nb = lib.rn_bn_scratch_bytes(M, Cc)
    scratch = torch.zeros(nb, dtype=torch.uint8, device="cuda")
    outs = []
    for _ in range(2):
        check(lib.rn_bn_stats_fwd(ptr(x), None, None, M, Cc, RN_BF16, 1e-5, 0.9, 1, ptr(mean), ptr(rstd), None, None,
                                  ptr(scratch), nb, stream()), "bn_stats")
        outs.append((mean.clone(), rstd.clone()))
    assert torch.equal(outs[0][0], outs[1][0]) and torch.equal(outs[0][1], outs[1][1])   # deterministic
    xf = x.float()
    assert rel_err(mean, xf.mean(0, dtype=torch.float64)) < 1e-5
    assert rel_err(rstd, 1 / torch.sqrt(xf.double().var(0, unbiased=False) + 1e-5)) < 1e-5


# ------------------------------------------------------------------------------------------ pooling
@pytest.mark.parametrize("mode", ["f32", "bf16"])
@pytest.mark.parametrize("shape", [(2, 112, 112, 8), (3, 15, 17, 5), (1, 4, 4, 3)])
def test_maxpool(shape, mode):
    N, H, W, Cc = shape
    tdt, rdt = DT[mode]
    g = torch.Generator().manual_seed(H)
    x = _q(torch.randn(N, H, W, Cc, generator=g), mode)
    net = ro.Net(dtype=torch.float64)
    xv = ro.Var(x.double())
    y = net.max_pool_3x3_s2_same(xv, "mp")
    Ho, Wo = y.val.shape[1:3]
    dy = _q(torch.randn(N, Ho, Wo, Cc, generator=g), mode)
    y.grad = dy.double()
    net.backward()
    pad = ro.tf_same_pad(H, 3, 2)[1]
    X, DY = dev(x, tdt), dev(dy, tdt)
    Y = torch.empty(N, Ho, Wo, Cc, dtype=tdt, device="cuda")
    AM = torch.empty(N, Ho, Wo, Cc, dtype=torch.uint8, device="cuda")
    check(lib.rn_maxpool3x3s2_fwd(ptr(X), ptr(Y), ptr(AM), N, H, W, Cc, Ho, Wo, pad, rdt, stream()), "maxpool")
    assert torch.equal(Y.float().cpu().double(), y.val)            # a max is exact
    DX = torch.empty(N, H, W, Cc, dtype=tdt, device="cuda")
    check(lib.rn_maxpool3x3s2_bwd(ptr(DY), ptr(AM), ptr(DX), N, H, W, Cc, Ho, Wo, pad, rdt, 0.0, stream()), "maxpool bwd")
    assert rel_err(DX.float(), xv.grad) < TOL[mode]


@pytest.mark.parametrize("cfg", [(4, 8, 8, 64, 8, 1, 64, 0, "f32", "f32"), (4, 7, 7, 96, 7, 1, 96, 0, "bf16", "f32"),
                                 (2, 16, 16, 16, 2, 2, 32, 8, "bf16", "bf16"), (2, 8, 8, 16, 1, 1, 160, 72, "f32", "f32")])
def test_avgpool_and_channel_pad(cfg):
    N, H, W, Cc, k, s, Cout, cb, im, om = cfg
    g = torch.Generator().manual_seed(k * 7 + s)
    x = _q(torch.randn(N, H, W, Cc, generator=g), im)
    net = ro.Net(dtype=torch.float64)
    xv = ro.Var(x.double())
    y = net.avg_pool(xv, k, s, "ap")
    if Cout != Cc:
        y = net.pad_channels(y, cb, Cout - Cc - cb, "pad")
    dy = _q(torch.randn(*y.val.shape, generator=g), om)
    y.grad = dy.double()
    net.backward()
    X, DY = dev(x, DT[im][0]), dev(dy, DT[om][0])
    Y = torch.empty(*y.val.shape, dtype=DT[om][0], device="cuda")
    check(lib.rn_avgpool_fwd(ptr(X), ptr(Y), N, H, W, Cc, k, s, Cout, cb, DT[im][1], DT[om][1], stream()), "avgpool")
    assert rel_err(Y.float(), y.val) < TOL[om]
    DX = torch.empty(N, H, W, Cc, dtype=DT[im][0], device="cuda")
    check(lib.rn_avgpool_bwd(ptr(DY), ptr(DX), N, H, W, Cc, k, s, Cout, cb, DT[im][1], DT[om][1], 0.0, stream()), "avgpool bwd")
    assert rel_err(DX.float(), xv.grad) < TOL[im]


# ------------------------------------------------------------------------------------------ head
@pytest.mark.parametrize("cfg", [(128, 64, 10), (16, 2048, 1000), (7, 33, 5)])
def test_dense_and_softmax_xent(cfg):
    B, D, K = cfg
    g = torch.Generator().manual_seed(B)
    x, w, b = torch.randn(B, D, generator=g), torch.randn(D, K, generator=g) / D ** 0.5, torch.randn(K, generator=g)
    lab = torch.randint(0, K, (B,), generator=g)
    onehot = torch.eye(K)[lab]
    net = ro.Net(dtype=torch.float64, params={"w": w.numpy(), "b": b.numpy()})
    xv = ro.Var(x.double())
    lg = net.dense(xv, "w", "b", D, K, None, "logits")
    loss, probs = net.softmax_xent(lg, onehot.double())
    net.backward()
    X, Wd, Bd, OH = dev(x), dev(w), dev(b), dev(onehot)
    L = torch.empty(B, K, device="cuda")
    check(lib.rn_dense_fwd(ptr(X), ptr(Wd), ptr(Bd), ptr(L), B, D, K, None, 0, stream()), "dense")
    assert rel_err(L, lg.val) < 1e-4
    nbs = lib.rn_dense_scratch_bytes(B, D, K)
    if nbs:      # deterministic split-K path (long reductions)
        sc = torch.empty(nbs, dtype=torch.uint8, device="cuda")
        L2 = torch.empty(B, K, device="cuda")
        check(lib.rn_dense_fwd(ptr(X), ptr(Wd), ptr(Bd), ptr(L2), B, D, K, ptr(sc), nbs, stream()), "dense split-K")
        assert rel_err(L2, lg.val) < 1e-4
    P, DL, S = torch.empty(B, K, device="cuda"), torch.empty(B, K, device="cuda"), torch.empty(2, device="cuda")
    check(lib.rn_softmax_xent(ptr(L), ptr(OH), B, K, ptr(P), ptr(DL), ptr(S), stream()), "xent")
    assert rel_err(P, probs) < 1e-4 and rel_err(DL, lg.grad) < 1e-4
    assert abs(S[0].item() - float(loss)) < 1e-4 * abs(float(loss))
    prec = (probs.argmax(1) == lab).double().mean().item()
    assert abs(S[1].item() - prec) < 1e-6
    DW, DB, DX = torch.empty(D, K, device="cuda"), torch.empty(K, device="cuda"), torch.empty(B, D, device="cuda")
    check(lib.rn_dense_bwd(ptr(X), ptr(Wd), ptr(DL), ptr(DW), ptr(DB), ptr(DX), B, D, K, None, 0, stream()), "dense bwd")
    assert rel_err(DW, net.param_grads["w"]) < 1e-4 and rel_err(DB, net.param_grads["b"]) < 1e-4
    assert rel_err(DX, xv.grad) < 1e-4
    nbb = lib.rn_dense_bwd_scratch_bytes(B, D, K)
    if nbb:      # deterministic split-K path for dx (reduction over the classes)
        sc = torch.empty(nbb, dtype=torch.uint8, device="cuda")
        DX2 = torch.full((B, D), float("nan"), device="cuda")
        check(lib.rn_dense_bwd(ptr(X), ptr(Wd), ptr(DL), None, None, ptr(DX2), B, D, K, ptr(sc), nbb, stream()), "dense bwd split-K")
        assert rel_err(DX2, xv.grad) < 1e-4


# ------------------------------------------------------------------------------------------ optimizer
@pytest.mark.parametrize("kind", ["mom", "sgd"])
@pytest.mark.parametrize("n", [64, 64 * 1000, 25549376])
def test_fused_optimizer(n, kind):
    """g' = g/R + wd*mask*w ; m = 0.9 m + g' ; w -= lr m — against the oracle's MomentumSGD on the
    mean of R gradients, at sizes up to the ResNet-50 fusion buffer (25.5 M floats)."""
    R, wd, lr, mom = 4, 2e-4, 0.1, 0.9
    g = torch.Generator(device="cuda").manual_seed(n % 1000)
    w = torch.randn(n, device="cuda", generator=g)
    gs = torch.randn(n, device="cuda", generator=g) * R          # the all-reduced SUM
    m = torch.randn(n, device="cuda", generator=g) * 0.1
    mask = (torch.arange(n // 64, device="cuda") % 3 != 0).to(torch.uint8)
    dec = mask.repeat_interleave(64).double()
    wd_ = w.double(); gd = gs.double() / R + wd * dec * wd_
    if kind == "mom":
        m_ref = mom * m.double() + gd
        w_ref = wd_ - lr * m_ref
    else:
        m_ref = m.double()
        w_ref = wd_ - lr * gd
    ss_ref = (dec * wd_ * wd_).sum().item()
    wb = torch.empty(n, dtype=torch.bfloat16, device="cuda")
    hyper = torch.tensor([lr, 0, 0, 0], device="cuda")
    ss = torch.zeros(1, device="cuda")
    nb = lib.rn_optimizer_scratch_bytes(n)
    scratch = torch.zeros(nb, dtype=torch.uint8, device="cuda")
    check(lib.rn_sgd_momentum_fused(ptr(w), ptr(gs), ptr(m), ptr(wb), n, ptr(mask), ptr(hyper), 1.0 / R, wd, mom,
                                    1 if kind == "mom" else 0, ptr(ss), ptr(scratch), nb, stream()), "opt")
    assert rel_err(w, w_ref) < 1e-6 and rel_err(m, m_ref) < 1e-6
    assert torch.equal(wb, w.to(torch.bfloat16))
    assert abs(ss.item() - ss_ref) <= 1e-5 * ss_ref
    ss2 = torch.zeros(1, device="cuda")
    check(lib.rn_l2_sumsq(ptr(w), n, ptr(mask), ptr(ss2), ptr(scratch), nb, stream()), "l2")
    assert abs(ss2.item() - (dec * w.double() ** 2).sum().item()) <= 1e-5 * ss_ref
    assert lib.rn_sgd_momentum_fused(ptr(w), ptr(gs), ptr(m), None, n + 1, None, ptr(hyper), 1.0, 0.0, 0.9, 1, None,
                                     None, 0, stream()) == -1


# ------------------------------------------------------------------------------------------ decode
@pytest.mark.parametrize("dataset,pad", [("cifar10", 4), ("cifar10", 2), ("cifar100", 4)])
def test_cifar_decode_bit_exact_indices(dataset, pad):
    n, B = 512, 128
    rec = co.synthetic_records(n, dataset, seed=1)
    rec[3, -3072:] = 200                    # constant image: variance 0 edge case
    idx = np.random.Generator(np.random.PCG64(2)).permutation(n)[:B].astype(np.int32)
    idx[0] = 3
    classes = 10 if dataset == "cifar10" else 100
    off = 0 if dataset == "cifar10" else 1
    R = dev(torch.from_numpy(rec))
    I = dev(torch.from_numpy(idx))
    img = torch.empty(B, 32, 32, 3, device="cuda")
    oh = torch.empty(B, classes, device="cuda")
    lab = torch.empty(B, dtype=torch.int32, device="cuda")
    dr = torch.empty(B, 3, dtype=torch.int32, device="cuda")
    for epoch in (0, 5):
        want, woh, wlab, (oy, ox, fl) = co.decode(rec, idx, dataset=dataset, train=True, pad=pad, epoch=epoch, seed=(7 << 32) | 9)
        check(lib.rn_cifar_decode(ptr(R), n, rec.shape[1], off, classes, ptr(I), B, 1, pad, 9, 7, epoch, None,
                                  0, ptr(img), RN_F32, ptr(oh), ptr(lab), ptr(dr), stream()), "decode")
        d = dr.cpu().numpy()
        assert np.array_equal(d[:, 0], oy) and np.array_equal(d[:, 1], ox) and np.array_equal(d[:, 2], fl)  # bit-exact
        assert np.array_equal(lab.cpu().numpy(), wlab) and np.array_equal(oh.cpu().numpy(), woh)
        got = img.cpu().numpy()
        # the uint8 -> float gather is exact: same zero pattern (padding) and same ordering of values
        assert np.array_equal(got == got.min(axis=(1, 2, 3), keepdims=True), want == want.min(axis=(1, 2, 3), keepdims=True))
        assert np.abs(got - want).max() <= 1e-6 * np.abs(want).max()
    # staged batch (only the B records of the batch, e.g. just copied from host memory): identical images / draws
    Rb = dev(torch.from_numpy(np.ascontiguousarray(rec[idx])))
    img2 = torch.empty_like(img)
    dr2 = torch.empty_like(dr)
    check(lib.rn_cifar_decode(ptr(Rb), B, rec.shape[1], off, classes, ptr(I), B, 1, pad, 9, 7, 5, None,
                              1, ptr(img2), RN_F32, ptr(oh), ptr(lab), ptr(dr2), stream()), "decode staged")
    assert torch.equal(img2, img) and torch.equal(dr2, dr)
    # explicit draws, bf16 output, eval mode
    draws = np.stack([oy, ox, fl], 1).astype(np.int32)
    D = dev(torch.from_numpy(draws))
    imgb = torch.empty(B, 32, 32, 3, dtype=torch.bfloat16, device="cuda")
    check(lib.rn_cifar_decode(ptr(R), n, rec.shape[1], off, classes, ptr(I), B, 1, pad, 0, 0, 0, ptr(D), 0, ptr(imgb),
                              RN_BF16, ptr(oh), None, None, stream()), "decode draws")
    assert torch.equal(imgb.cpu(), torch.from_numpy(want).to(torch.bfloat16).cpu()) or \
        np.abs(imgb.float().cpu().numpy() - want).max() < 2e-2
    ev, *_ = co.decode(rec, idx, dataset=dataset, train=False)
    check(lib.rn_cifar_decode(ptr(R), n, rec.shape[1], off, classes, ptr(I), B, 0, pad, 0, 0, 0, None, 0, ptr(img),
                              RN_F32, ptr(oh), None, None, stream()), "decode eval")
    assert np.abs(img.cpu().numpy() - ev).max() <= 1e-6 * np.abs(ev).max()
    assert np.all(img[0].cpu().numpy() == 0)            # constant image: variance 0 -> exactly 0
    # empty batch is a no-op
    check(lib.rn_cifar_decode(ptr(R), n, rec.shape[1], off, classes, ptr(I), 0, 1, pad, 0, 0, 0, None, 0, ptr(img),
                              RN_F32, None, None, None, stream()), "decode empty")


def test_decode_full_epoch_checksum():
    """All 50,000 records in one launch: labels round-trip and every image is standardised
    (mean 0, unit variance) — size-independent properties at the full dataset size."""
    n = 50000
    rec = co.synthetic_records(n, "cifar10", 0)
    R = dev(torch.from_numpy(rec))
    img = torch.empty(n, 32, 32, 3, device="cuda")
    lab = torch.empty(n, dtype=torch.int32, device="cuda")
    check(lib.rn_cifar_decode(ptr(R), n, 3073, 0, 10, None, n, 1, 4, 0, 0, 0, None, 0, ptr(img), RN_F32, None, ptr(lab),
                              None, stream()), "decode")
    assert torch.equal(lab.cpu(), torch.from_numpy(rec[:, 0].astype(np.int32)))
    flat = img.view(n, -1)
    assert flat.mean(1).abs().max().item() < 1e-5
    assert (flat.var(1, unbiased=False) - 1).abs().max().item() < 1e-4


# ------------------------------------------------------------------------------------------ utilities
def test_cast_add_pad():
    g = torch.Generator().manual_seed(0)
    x = torch.randn(1000, 3, generator=g)
    X = dev(x)
    B16 = torch.empty(1000, 3, dtype=torch.bfloat16, device="cuda")
    check(lib.rn_cast(ptr(X), ptr(B16), 3000, RN_F32, RN_BF16, stream()), "cast")
    assert torch.equal(B16.cpu(), x.to(torch.bfloat16))
    P = torch.empty(1000, 8, dtype=torch.bfloat16, device="cuda")
    check(lib.rn_pad_channels(ptr(X), ptr(P), 1000, 3, 8, RN_F32, RN_BF16, stream()), "pad")
    assert torch.equal(P[:, :3].cpu(), x.to(torch.bfloat16)) and P[:, 3:].abs().max().item() == 0
    a, b = torch.randn(1003, generator=g), torch.randn(1003, generator=g)
    A, Bd = dev(a), dev(b)
    Y = torch.empty(1003, device="cuda")
    check(lib.rn_add(ptr(A), ptr(Bd), ptr(Y), 1003, RN_F32, stream()), "add")
    assert torch.equal(Y.cpu(), a + b)
    check(lib.rn_add_inplace(ptr(A), ptr(Bd), 1003, RN_F32, stream()), "add_inplace")
    assert torch.equal(A.cpu(), a + b)
    check(lib.rn_cast(ptr(X), ptr(B16), 0, RN_F32, RN_BF16, stream()), "cast empty")
