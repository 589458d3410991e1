"""Step-through verifier: executes the executor's launch list ONE KERNEL AT A TIME and checks every
launch, in situ, against the CPU oracle's restatement of that op applied to the buffers the kernel
actually read (downloaded to fp64).  This is the "per-layer activations and gradients" criterion
of the north star without error accumulation across layers: fp32 verification mode within 1e-4,
bf16 mode within 2e-2 (max-norm relative).  It covers whole networks (all conv shapes, every
fused add / in-place gradient accumulation, both builder families) at shapes up to the real
ImageNet ResNet-50 layer sizes."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from util import rel_err, TOL  # noqa: E402
from distributed_tensorflow_resnet_b200 import resnet_model, FLAGS  # noqa: E402
from oracle import resnet_oracle as ro  # noqa: E402


def _d(t):
    return t.detach().double().cpu()


def _var(t):
    return ro.Var(_d(t))


class Verifier:
    def __init__(self, model, mode):
        self.m = model
        self.ex = model.executor
        self.tol = TOL[mode]
        self.bf16 = mode == "bf16"
        self.worst = {}
        self.count = 0

    def q(self, t):
        return t.to(torch.bfloat16).double() if self.bf16 else t

    def check(self, what, got, want, tol=None):
        e = rel_err(_d(got), want)
        fam = what.split(":")[0]
        self.worst[fam] = max(self.worst.get(fam, 0.0), e)
        self.count += 1
        assert e < (tol or self.tol), (what, e)

    def weight(self, n):
        """the weight values the conv kernels multiply with (bf16 copy in bf16 mode)"""
        w = _d(self.ex.var_view(n.vars["kernel"].name))
        return self.q(w)

    # ---------------------------------------------------------------- forward
    def verify_forward(self, name, node):
        ex = self.ex
        net = ro.Net(dtype=torch.float64)
        fam = name.split(":")[0]
        if node is None:
            return
        if node.op == "conv2d":
            x, y = node.inputs[0], node.outputs[0]
            a = node.attrs
            net._given = {"w": self.weight(node).numpy()}
            # (a BN output that exists only inside the consuming 1x1 convolutions is re-created with rn_bn_apply_act_fwd:
            #  the check is then "conv with the on-the-fly normalisation == bn_apply followed by the oracle's conv")
            out = net.conv2d(_var(ex.materialize(x)), "w", a["k"], x.shape[3], y.shape[3], a["stride"], a["pad"], y.shape[1:3], None, "y")
            f = getattr(node, "fuse", None)
            if f is None:
                self.check(name, y.buf, out.val)
            else:       # fused epilogue: the launch wrote conv (+ shortcut) into the BN input, never y itself
                want = out.val + (_d(f["residual"].buf) if f["residual"] is not None else 0.0)
                self.check(name + "/fused", f["out"].buf, want)
        elif node.op == "bn_act":
            x, y = node.inputs[0], node.outputs[0]
            a, v = node.attrs, node.vars
            C = x.shape[-1]
            add = ex._fused_add_for(node)
            if fam in ("bn_stats", "bn_finalize"):
                if add is not None and fam == "bn_stats":
                    z = _d(add.inputs[0].buf) + _d(add.inputs[1].buf)
                    self.check(name + "/z", x.buf, z)
                x2 = _d(x.buf).reshape(-1, C)
                self.check(name + "/mean", node.saved[0], x2.mean(0), 1e-5)
                self.check(name + "/rstd", node.saved[1], 1 / torch.sqrt(x2.var(0, unbiased=False) + a["eps"]), 1e-5)
            else:
                x2 = _d(x.buf).reshape(-1, C)
                g, b = _d(ex.var_view(v["gamma"].name)), _d(ex.var_view(v["beta"].name))
                if a["is_training"]:
                    mean, rstd = _d(node.saved[0]), _d(node.saved[1])
                    # (also here: the statistics may have been finalised by the producing convolution's last CTA,
                    #  in which case no bn_stats / bn_finalize launch exists to hang the check on)
                    self.check(name + "/mean", node.saved[0], x2.mean(0), 1e-5)
                    self.check(name + "/rstd", node.saved[1], 1 / torch.sqrt(x2.var(0, unbiased=False) + a["eps"]), 1e-5)
                else:
                    mean = _d(ex.var_view(v["moving_mean"].name))
                    rstd = 1 / torch.sqrt(_d(ex.var_view(v["moving_variance"].name)) + a["eps"])
                pre = (x2 - mean) * rstd * g + b
                want = torch.where(pre < 0, a["leak"] * pre, pre).reshape(x.shape)
                self.check(name, ex.materialize(y), want)
        elif node.op == "add":
            self.check(name, node.outputs[0].buf, _d(node.inputs[0].buf) + _d(node.inputs[1].buf))
        elif node.op == "maxpool":
            out = net.max_pool_3x3_s2_same(_var(node.inputs[0].buf), "mp")
            self.check(name, node.outputs[0].buf, out.val, 1e-7)
        elif node.op == "avgpool":
            a = node.attrs
            x, y = node.inputs[0], node.outputs[0]
            out = net.avg_pool(_var(x.buf), a["k"], a["s"], "ap")
            if y.shape[3] != x.shape[3]:
                out = net.pad_channels(out, a["c_before"], y.shape[3] - x.shape[3] - a["c_before"], "pad")
            self.check(name, y.buf, out.val)
        elif node.op == "dense":
            x, y = node.inputs[0], node.outputs[0]
            w, b = _d(ex.var_view(node.vars["kernel"].name)), _d(ex.var_view(node.vars["bias"].name))
            self.check(name, y.buf, _d(x.buf).reshape(x.shape[0], -1) @ w + b, 1e-4)
        elif node.op == "softmax_xent":
            lg, lab = _var(node.inputs[0].buf), _d(node.inputs[1].buf)
            loss, probs = net.softmax_xent(lg, lab)
            net.backward()
            self.check(name + "/probs", node.outputs[0].buf, probs, 1e-4)
            self.check(name + "/dlogits", node.dlogits, lg.grad, 1e-4)
            assert abs(node.outputs[1].buf[0].item() - float(loss)) < 1e-4 * abs(float(loss))

    # ---------------------------------------------------------------- backward
    def run_backward_launch(self, name, fn, st):
        ex = self.ex
        meta = ex.meta.get(name)
        fam = name.split(":")[0]
        if meta is None or fam in ("grad_copy", "grad_add"):
            fn(st)
            return
        node = meta["node"]
        dy = _d(meta["dy"]) if meta.get("dy") is not None else None
        dx_buf = meta.get("dx")
        beta = meta.get("beta", 0.0)
        before = _d(dx_buf) if (dx_buf is not None and beta != 0.0) else None
        fn(st)
        torch.cuda.synchronize()
        net = ro.Net(dtype=torch.float64)

        def expect(dx):
            return dx + before if before is not None else dx

        if fam in ("conv_dgrad", "conv_dgrad_tc", "conv_wgrad", "conv_wgrad_tc"):
            x, y = node.inputs[0], node.outputs[0]
            a = node.attrs
            net._given = {"w": self.weight(node).numpy()}
            xv = _var(ex.materialize(x))
            out = net.conv2d(xv, "w", a["k"], x.shape[3], y.shape[3], a["stride"], a["pad"], y.shape[1:3], None, "y")
            out.grad = dy.reshape(y.shape)
            net.backward()
            if "dgrad" in fam:
                self.check(name, dx_buf, expect(xv.grad))
            else:
                # fp32 accumulation over up to 1.6 M pixels of products of (bf16) values
                self.check(name, ex.var_view(node.vars["kernel"].name, "G"), net.param_grads["w"],
                           1e-4 if not self.bf16 else 2e-3)
        elif fam in ("bn_bwd_reduce", "bn_bwd_finalize", "bn_bwd_apply"):
            x, y = node.inputs[0], node.outputs[0]
            a, v = node.attrs, node.vars
            C = x.shape[-1]
            M = x.numel // C
            x2 = _d(x.buf).reshape(M, C)
            g_, b_ = _d(ex.var_view(v["gamma"].name)), _d(ex.var_view(v["beta"].name))
            mean, rstd = _d(node.saved[0]), _d(node.saved[1])
            # the mask is defined by the value the forward kernel produced (fp32 fma form)
            af = (g_ * rstd).float()
            bf = (b_.float() - mean.float() * af)
            pre = torch.addcmul(bf, x2.float(), af).double()
            xhat = (x2 - mean) * rstd
            g = dy.reshape(M, C)
            g = torch.where(pre < 0, a["leak"] * g, g)
            dbeta, dgamma = g.sum(0), (g * xhat).sum(0)
            if True:        # after every BN-backward launch: dgamma / dbeta may come from a dgrad's folded finalisation
                scale = max(float(g.abs().sum(0).max()), 1e-30)     # conditioning: sums of signed terms
                e1 = (_d(ex.var_view(v["beta"].name, "G")) - dbeta).abs().max().item() / scale
                e2 = (_d(ex.var_view(v["gamma"].name, "G")) - dgamma).abs().max().item() / max(float((g * xhat).abs().sum(0).max()), 1e-30)
                self.worst[fam] = max(self.worst.get(fam, 0.0), e1, e2)
                self.count += 1
                assert e1 < 1e-5 and e2 < 1e-5, (name, e1, e2)
            if fam == "bn_bwd_apply":
                dgk, dbk = _d(ex.var_view(v["gamma"].name, "G")), _d(ex.var_view(v["beta"].name, "G"))
                dx = g_ * rstd * (g - dbk / M - xhat * dgk / M)
                self.check(name, dx_buf, expect(dx.reshape(x.shape)))
        elif fam == "dense_bwd":
            x = node.inputs[0]
            w = _d(ex.var_view(node.vars["kernel"].name))
            x2 = _d(x.buf).reshape(x.shape[0], -1)
            self.check(name + "/db", ex.var_view(node.vars["bias"].name, "G"), dy.sum(0), 1e-4)
            if dx_buf is not None:
                self.check(name + "/dx", dx_buf, (dy @ w.t()).reshape(x.shape), 1e-4)
        elif fam == "dense_wgrad":
            x2 = _d(node.inputs[0].buf).reshape(node.inputs[0].shape[0], -1)
            self.check(name + "/dw", ex.var_view(node.vars["kernel"].name, "G"), x2.t() @ dy, 1e-4)
        elif fam == "avgpool_bwd":
            a = node.attrs
            x, y = node.inputs[0], node.outputs[0]
            xv = _var(x.buf)
            out = net.avg_pool(xv, a["k"], a["s"], "ap")
            if y.shape[3] != x.shape[3]:
                out = net.pad_channels(out, a["c_before"], y.shape[3] - x.shape[3] - a["c_before"], "pad")
            out.grad = dy.reshape(y.shape)
            net.backward()
            self.check(name, dx_buf, expect(xv.grad))
        elif fam == "maxpool_bwd":
            xv = _var(node.inputs[0].buf)
            out = net.max_pool_3x3_s2_same(xv, "mp")
            out.grad = dy.reshape(node.outputs[0].shape)
            net.backward()
            self.check(name, dx_buf, expect(xv.grad))
        else:
            raise AssertionError("unverified launch " + name)

    def run(self):
        ex = self.ex
        st = torch.cuda.current_stream().cuda_stream
        for name, fn in ex.fwd:
            fn(st)
            torch.cuda.synchronize()
            self.verify_forward(name, ex.meta[name]["node"])
        for name, fn in ex.bwd:
            self.run_backward_launch(name, fn, st)
        torch.cuda.synchronize()
        return self


def _build(kind, mode, shape, classes, flags, legacy=None, filters=None):
    g = torch.Generator().manual_seed(7)
    x = torch.randn(*shape, generator=g)
    y = torch.eye(classes)[torch.arange(shape[0]) % classes]
    hkw = dict(num_classes=classes, lrn_rate=0.1, weight_decay_rate=2e-4, optimizer='mom')
    if legacy:
        hkw.update(batch_size=shape[0], min_lrn_rate=1e-4, **legacy)
    FLAGS.reset()
    FLAGS.update(use_cuda_graph=False, compute_dtype=mode, seed=3, **flags)
    dt = torch.bfloat16 if mode == "bf16" else torch.float32
    m = resnet_model.ResNet(resnet_model.HParams(**hkw), x.to(dt).cuda().contiguous(), y.cuda().contiguous(), "train")
    if filters:
        m.filters = filters
    m.build_graph()
    return m


NETS = [
    ("official_cifar50", dict(dataset="cifar10", resnet_size=50), (16, 32, 32, 3), 10, None, None),
    ("official_imagenet50_96px", dict(dataset="imagenet", resnet_size=50), (4, 96, 96, 3), 100, None, None),
    ("official_imagenet18", dict(dataset="imagenet", resnet_size=18), (4, 64, 64, 3), 10, None, None),
    ("upstream_resnet20", dict(dataset="cifar10"), (16, 32, 32, 3), 10, dict(num_residual_units=3, use_bottleneck=False, relu_leakiness=0.1), None),
    ("upstream_bottleneck", dict(dataset="cifar10"), (8, 32, 32, 3), 10, dict(num_residual_units=2, use_bottleneck=True, relu_leakiness=0.1), None),
    ("upstream_wrn_narrow", dict(dataset="cifar100"), (4, 32, 32, 3), 100, dict(num_residual_units=1, use_bottleneck=False, relu_leakiness=0.1), [16, 40, 80, 160]),
]


@pytest.mark.parametrize("mode", ["f32", "bf16"])
@pytest.mark.parametrize("net", NETS, ids=[n[0] for n in NETS])
def test_every_launch_against_oracle(net, mode):
    name, flags, shape, classes, legacy, filters = net
    m = _build(name, mode, shape, classes, flags, legacy, filters)
    v = Verifier(m, mode).run()
    assert v.count > 50
    print("\n%s %s: %d checks; worst per family: %s" % (name, mode, v.count,
          ", ".join("%s %.1e" % kv for kv in sorted(v.worst.items()))))


def test_bn_operand_fusion_is_used_and_bit_identical(monkeypatch):
    """ResNet-50 (bottleneck) with RN_BN_OPERAND_FUSION=1 (opt-in): the BN outputs read only by 1x1 convolutions are
    not materialised (32 of the 49 bn_apply launches disappear), and training steps give BIT-IDENTICAL weights,
    gradients and BN statistics to the default pipeline, which stores every BN output."""
    res = []
    monkeypatch.setenv("RN_TC_PAIR", "0")        # identical tilings, so the per-CTA statistics rows add up in the same order
    for off in (False, True):
        if off:
            monkeypatch.delenv("RN_BN_OPERAND_FUSION")
        else:
            monkeypatch.setenv("RN_BN_OPERAND_FUSION", "1")
        m = _build("c3", "bf16", (4, 96, 96, 3), 100, dict(dataset="imagenet", resnet_size=50))
        ex = m.executor
        napply = sum(1 for nm, _ in ex.fwd if nm.startswith("bn_apply"))
        nvirtual = sum(1 for n in m.graph.nodes if n.op == "bn_act" and getattr(n, "virtual", False))
        for _ in range(2):
            m.train_op()
        torch.cuda.synchronize()
        res.append((napply, nvirtual, ex.W.clone(), ex.Gd.clone(), ex.S.clone(), ex.scalars()))
    (a0, v0, W0, G0, S0, sc0), (a1, v1, W1, G1, S1, sc1) = res
    assert v1 == 0 and a1 == 49 and v0 == 32 and a0 == 49 - 32
    assert torch.equal(W0, W1) and torch.equal(G0, G1) and torch.equal(S0, S1) and sc0 == sc1


@pytest.mark.parametrize("mode", ["apply", "conv"])
@pytest.mark.parametrize("net", ["c3", "cifar"])
def test_bn_finalize_modes_match_the_standalone_kernels(monkeypatch, mode, net):
    """RN_BN_FINALIZE=apply / conv (opt-in): the per-CTA sums of the fused convolution epilogues are finalised inside
    the BN apply kernels / by the convolution's last CTA instead of rn_bn_finalize / rn_bn_bwd_finalize launches.  The
    finalize launches disappear and two training steps reproduce the default pipeline (every variant sums the same
    fp32 rows in fp64, so weights and statistics agree to rounding of the last bit)."""
    res = []
    for m_ in ("kernel", mode):
        monkeypatch.setenv("RN_BN_FINALIZE", m_)
        if net == "c3":
            m = _build("c3", "bf16", (4, 96, 96, 3), 100, dict(dataset="imagenet", resnet_size=50))
        else:
            m = _build("c1", "bf16", (16, 32, 32, 3), 10, dict(dataset="cifar10", resnet_size=20))
        ex = m.executor
        nfin = sum(1 for nm, _ in ex.fwd if nm.startswith("bn_finalize"))
        for _ in range(2):
            m.train_op()
        torch.cuda.synchronize()
        res.append((nfin, ex.count_launches(), ex.W.clone(), ex.Gd.clone(), ex.S.clone(), ex.scalars()))
    (f0, l0, W0, G0, S0, sc0), (f1, l1, W1, G1, S1, sc1) = res
    assert f0 > 0 and f1 < f0 and l1 < l0
    if mode == "apply":
        assert f1 == 0
    for a, b in ((W0, W1), (G0, G1), (S0, S1)):
        assert torch.isfinite(b).all()
        assert (a.double() - b.double()).abs().max().item() <= 1e-6 * max(1.0, a.abs().max().item())
    assert abs(sc0[2] - sc1[2]) <= 1e-6 * abs(sc0[2])


def test_imagenet_resnet50_full_resolution_layers():
    """The real C3 layer shapes (224x224 input -> 112/56/28/14/7 feature maps) at batch 2, bf16."""
    m = _build("c3", "bf16", (2, 224, 224, 3), 1000, dict(dataset="imagenet", resnet_size=50))
    v = Verifier(m, "bf16").run()
    assert v.count > 300
