"""CPU oracle (torch-CPU, fp32 or fp64) of the reference's ResNet training step.

TEST INFRASTRUCTURE — see oracle/__init__.py.  All `file:line` citations are relative to
/root/reference.  Activations are NHWC, conv kernels HWIO, exactly as TF stores them.

The oracle keeps an explicit tape with hand-written backward formulas so that per-layer
activations *and* gradients can be compared with the CUDA path by TF op name
(`conv2d_3/Conv2D`, `Relu_2`, `add_1`, ...).  tests/test_oracle.py cross-checks the
hand-written backward against torch.autograd.
"""
from __future__ import annotations

import math
import os
from collections import OrderedDict

import numpy as np
import torch
import torch.nn.functional as F

# resnet_model_official.py:37-38.  tf.layers.batch_normalization(fused=True) passes
# epsilon=1e-5 through FusedBatchNorm, whose kernel requires epsilon > 1e-5 and therefore the
# layer bumps it to 1.001e-5 (fixture: epsilon=1.0009999641624745e-05 on all 49 nodes).
BN_DECAY_OFFICIAL = 0.997
BN_EPS_OFFICIAL = 1.001e-5
# upstream tensorflow/models research/resnet/resnet_model.py (un-vendored)
BN_DECAY_UPSTREAM = 0.9
BN_EPS_UPSTREAM = 0.001


class Var:
    """A tape value: NHWC activation (or [N, D] matrix) with an accumulated gradient.
    `q` is the storage rounding of the CUDA path being checked (identity, or bf16 round-trip)."""
    __slots__ = ("val", "grad", "name", "q")

    def __init__(self, val, name=None, q=None):
        self.val = val
        self.grad = None
        self.name = name
        self.q = q

    def add_grad(self, g):
        g = g if self.grad is None else self.grad + g
        self.grad = self.q(g) if self.q is not None else g


def tf_same_pad(size, k, s):
    """TF 'SAME' padding: (out, pad_before) for one spatial axis."""
    out = -(-size // s)
    total = max((out - 1) * s + k - size, 0)
    return out, total // 2


class Net:
    """Holds parameters (TF names, creation order), BN moving stats and the tape."""

    def __init__(self, dtype=torch.float32, seed=0, params=None, storage=None):
        """storage=torch.bfloat16 emulates the bf16 mode of the CUDA path: activations, activation
        gradients and conv weights are rounded to bf16 wherever that path stores them in HBM
        (arithmetic stays in `dtype`), so the comparison isolates kernel error from the
        precision choice itself."""
        self.dtype = dtype
        self.storage = storage
        self.params = OrderedDict()      # trainable, creation order == tf.trainable_variables()
        self.param_grads = OrderedDict()
        self.state = OrderedDict()       # moving_mean / moving_variance
        self.decay_mask = OrderedDict()  # name -> bool (weight decay applies)
        self.tape = []
        self.acts = OrderedDict()
        self._counts = {}
        self._gen = torch.Generator().manual_seed(seed)
        self._given = params             # optional dict name -> array (overrides initialisers)
        self._scope = ""
        self.bn_updates = OrderedDict()  # name -> (batch_mean, batch_var_for_moving)

    # ---------------------------------------------------------------- variables
    def _uniq(self, base):
        i = self._counts.get(base, 0)
        self._counts[base] = i + 1
        return base if i == 0 else "%s_%d" % (base, i)

    def get_param(self, name, shape, init, decay=True):
        if name not in self.params:
            if self._given is not None and name in self._given:
                v = torch.as_tensor(np.asarray(self._given[name]), dtype=self.dtype).reshape(shape).clone()
            else:
                v = init(shape).to(self.dtype)
            self.params[name] = v
            self.decay_mask[name] = decay
        return self.params[name]

    def get_state(self, name, shape, value):
        if name not in self.state:
            if self._given is not None and name in self._given:
                v = torch.as_tensor(np.asarray(self._given[name]), dtype=self.dtype).reshape(shape).clone()
            else:
                v = torch.full(shape, float(value), dtype=self.dtype)
            self.state[name] = v
        return self.state[name]

    def begin_pass(self):
        self.tape = []
        self.acts = OrderedDict()
        self._counts = {}
        self.param_grads = OrderedDict()
        self.bn_updates = OrderedDict()

    def _round(self, t):
        return t if self.storage is None else t.to(self.storage).to(self.dtype)

    def _rec(self, var, name, stored=True):
        """Register an op output; `stored` = the CUDA path keeps it in the activation dtype."""
        var.name = name
        if stored and self.storage is not None:
            var.val = self._round(var.val)
            var.q = self._round
        self.acts[name] = var
        return var

    def _pgrad(self, name, g):
        self.param_grads[name] = g if name not in self.param_grads else self.param_grads[name] + g

    # ---------------------------------------------------------------- initialisers
    def init_variance_scaling(self, shape):
        """tf.variance_scaling_initializer() defaults (resnet_model_official.py:90): scale=1,
        fan_in, truncated normal with stddev sqrt(1/fan_in)/0.87962566 (fixture: 0.218786 for
        3x3x3x16)."""
        fan_in = shape[0] * shape[1] * shape[2]
        std = math.sqrt(1.0 / fan_in) / 0.87962566103423978
        t = torch.empty(shape, dtype=torch.float64)
        torch.nn.init.trunc_normal_(t, 0.0, std, -2 * std, 2 * std, generator=self._gen)
        return t

    def init_glorot_uniform(self, shape):
        """tf.layers.dense default kernel initialiser (fixture: +-0.284747 for 64x10)."""
        lim = math.sqrt(6.0 / (shape[0] + shape[1]))
        return (torch.rand(shape, dtype=torch.float64, generator=self._gen) * 2 - 1) * lim

    # ---------------------------------------------------------------- ops
    def conv2d(self, x, w_name, k, cin, cout, stride, pad_before, out_hw, init, opname):
        """Conv2D NHWC/HWIO, no bias.  `pad_before` zero rows/cols precede the image; the trailing
        pad is whatever `out_hw` needs (zero fill).  Follows tf.layers.conv2d as called at
        resnet_model_official.py:87-91 and tf.nn.conv2d at upstream `_conv`."""
        w = self._round(self.get_param(w_name, (k, k, cin, cout), init, decay=True))
        xv = x.val
        N, H, W, C = xv.shape
        Ho, Wo = out_hw
        pad_after_h = max((Ho - 1) * stride + k - H - pad_before, 0)
        pad_after_w = max((Wo - 1) * stride + k - W - pad_before, 0)
        xp = F.pad(xv.permute(0, 3, 1, 2), (pad_before, pad_after_w, pad_before, pad_after_h))
        wt = w.permute(3, 2, 0, 1)  # OIHW
        # fp64 has no oneDNN convolution (torch falls back to a naive loop, ~20 s per ResNet-20 step): the
        # same contraction as unfold + MKL dgemm is ~30x faster and differs only in fp64 summation order
        gemm = xv.dtype == torch.float64 and os.environ.get("RN_ORACLE_NATIVE_CONV64") is None
        if gemm:
            cols = F.unfold(xp, k, stride=stride)                       # [N, C*k*k, L], rows ordered (c, r, s)
            Hf = (xp.shape[2] - k) // stride + 1
            Wf = (xp.shape[3] - k) // stride + 1
            w2 = wt.reshape(cout, cin * k * k)
            y = torch.matmul(w2, cols).reshape(N, cout, Hf, Wf)[:, :, :Ho, :Wo]
        else:
            y = F.conv2d(xp, wt, stride=stride)[:, :, :Ho, :Wo]
        out = self._rec(Var(y.permute(0, 2, 3, 1).contiguous()), opname)

        def bwd():
            g = out.grad.permute(0, 3, 1, 2)
            # pad the cropped rows/cols of the VALID output back with zeros, then the native
            # (oneDNN) backward kernels: dgrad = transposed conv, wgrad = correlation of x and dy
            full = ((xp.shape[2] - k) // stride + 1, (xp.shape[3] - k) // stride + 1)
            if tuple(full) != (Ho, Wo):
                g = F.pad(g, (0, full[1] - Wo, 0, full[0] - Ho))
            if gemm:
                g2 = g.reshape(N, cout, full[0] * full[1])
                gw = torch.einsum("nol,nkl->ok", g2, cols).reshape(cout, cin, k, k)
                gx = F.fold(torch.matmul(w2.t(), g2), (xp.shape[2], xp.shape[3]), k, stride=stride)
            else:
                gx, gw, _ = torch.ops.aten.convolution_backward(
                    g.contiguous(), xp, wt, None, [stride, stride], [0, 0], [1, 1], False, [0, 0], 1, [True, True, False])
            gx = gx[:, :, pad_before:pad_before + H, pad_before:pad_before + W]
            x.add_grad(gx.permute(0, 2, 3, 1).contiguous())
            self._pgrad(w_name, gw.permute(2, 3, 1, 0).contiguous())
        self.tape.append(bwd)
        return out

    def batch_norm_act(self, x, scope, is_training, eps, decay, unbiased_moving_var, leak, opnames):
        """Batch norm + (leaky) ReLU.
        Official: resnet_model_official.py:41-50 -> FusedBatchNorm(is_training) + Relu; moving
        stats `mov -= (mov - batch)*(1-decay)` with the UNBIASED batch variance (fixture:
        AssignSub(mov, (mov - batch)*(1-0.997))).
        Upstream `_batch_norm`: tf.nn.moments (biased var) for both normalisation
        and the 0.9 moving average, eps 1e-3; `_relu`: where(x<0, leak*x, x)."""
        xv = x.val
        C = xv.shape[-1]
        mk_gamma = lambda: self.get_param(scope + "/gamma", (C,), lambda s: torch.ones(s, dtype=torch.float64), decay=self._bn_decay_flag)
        mk_beta = lambda: self.get_param(scope + "/beta", (C,), lambda s: torch.zeros(s, dtype=torch.float64), decay=self._bn_decay_flag)
        if self._beta_first:      # upstream `_batch_norm` creates beta before gamma; tf.layers the reverse
            beta = mk_beta()
            gamma = mk_gamma()
        else:
            gamma = mk_gamma()
            beta = mk_beta()
        mm = self.get_state(scope + "/moving_mean", (C,), 0.0)
        mv = self.get_state(scope + "/moving_variance", (C,), 1.0)
        M = xv.numel() // C
        x2 = xv.reshape(M, C)
        if is_training:
            mean = x2.mean(0)
            var = ((x2 - mean) ** 2).mean(0)
            mov_var = var * (M / max(M - 1, 1)) if unbiased_moving_var else var
            self.bn_updates[scope] = (mean.clone(), mov_var.clone())
        else:
            mean, var = mm, mv
        rstd = 1.0 / torch.sqrt(var + eps)
        xhat = (x2 - mean) * rstd
        pre = (xhat * gamma + beta).reshape(xv.shape)
        bn_out = self._rec(Var(pre), opnames[0], stored=False)
        act = torch.where(pre < 0, leak * pre, pre)
        out = self._rec(Var(act), opnames[1])
        g_name, b_name = scope + "/gamma", scope + "/beta"

        def bwd():
            g = out.grad
            if bn_out.grad is not None:
                raise RuntimeError("bn pre-activation consumed twice")
            g = torch.where(pre < 0, leak * g, g).reshape(M, C)
            bn_out.grad = g.reshape(xv.shape)
            dbeta = g.sum(0)
            dgamma = (g * xhat).sum(0)
            if is_training:
                dx = gamma * rstd * (g - dbeta / M - xhat * dgamma / M)
            else:
                dx = gamma * rstd * g
            x.add_grad(dx.reshape(xv.shape))
            self._pgrad(g_name, dgamma)
            self._pgrad(b_name, dbeta)
        self.tape.append(bwd)
        return out

    def add(self, a, b, opname):
        out = self._rec(Var(a.val + b.val), opname)

        def bwd():
            a.add_grad(out.grad)
            b.add_grad(out.grad)
        self.tape.append(bwd)
        return out

    def max_pool_3x3_s2_same(self, x, opname):
        """tf.layers.max_pooling2d(pool_size=3, strides=2, padding='SAME'),
        resnet_model_official.py:314-316.  Padding cells never win (TF pads with -inf)."""
        xv = x.val
        N, H, W, C = xv.shape
        Ho, ph = tf_same_pad(H, 3, 2)
        Wo, pw = tf_same_pad(W, 3, 2)
        pah = max((Ho - 1) * 2 + 3 - H - ph, 0)
        paw = max((Wo - 1) * 2 + 3 - W - pw, 0)
        xp = F.pad(xv.permute(0, 3, 1, 2), (pw, paw, ph, pah), value=float("-inf"))
        y, idx = F.max_pool2d(xp, 3, 2, return_indices=True)
        out = self._rec(Var(y.permute(0, 2, 3, 1).contiguous()), opname)

        def bwd():
            g = out.grad.permute(0, 3, 1, 2).contiguous()
            # overlapping windows: one input cell can win several windows -> accumulate
            Hp, Wp = xp.shape[-2:]
            gx = torch.zeros((N, C, Hp * Wp), dtype=g.dtype)
            gx.scatter_add_(2, idx.reshape(N, C, -1), g.reshape(N, C, -1))
            gx = gx.reshape(N, C, Hp, Wp)[:, :, ph:ph + H, pw:pw + W]
            x.add_grad(gx.permute(0, 2, 3, 1).contiguous())
        self.tape.append(bwd)
        return out

    def avg_pool(self, x, k, s, opname, stored=True):
        """VALID average pooling, window k stride s (global pool when k == H; the upstream
        `_residual` shortcut uses k == s == stride)."""
        xv = x.val
        y = F.avg_pool2d(xv.permute(0, 3, 1, 2), k, s)
        out = self._rec(Var(y.permute(0, 2, 3, 1).contiguous()), opname, stored=stored)
        N, H, W, C = xv.shape

        def bwd():
            g = out.grad.permute(0, 3, 1, 2)
            Ho, Wo = g.shape[-2:]
            gx = torch.zeros((N, C, H, W), dtype=xv.dtype)
            for i in range(k):
                for j in range(k):
                    gx[:, :, i:i + (Ho - 1) * s + 1:s, j:j + (Wo - 1) * s + 1:s] += g / (k * k)
            x.add_grad(gx.permute(0, 2, 3, 1).contiguous())
        self.tape.append(bwd)
        return out

    def pad_channels(self, x, before, after, opname):
        out = self._rec(Var(F.pad(x.val, (before, after))), opname)
        C = x.val.shape[-1]

        def bwd():
            x.add_grad(out.grad[..., before:before + C].contiguous())
        self.tape.append(bwd)
        return out

    def dense(self, x, w_name, b_name, din, dout, w_init, opname, decay_bias=True):
        """tf.layers.dense (resnet_model_official.py:274,343) / upstream `_fully_connected`."""
        w = self.get_param(w_name, (din, dout), w_init, decay=True)
        b = self.get_param(b_name, (dout,), lambda s: torch.zeros(s, dtype=torch.float64), decay=decay_bias)
        x2 = x.val.reshape(x.val.shape[0], -1)
        out = self._rec(Var(x2 @ w + b), opname, stored=False)

        def bwd():
            g = out.grad
            self._pgrad(w_name, x2.t() @ g)
            self._pgrad(b_name, g.sum(0))
            x.add_grad((g @ w.t()).reshape(x.val.shape))
        self.tape.append(bwd)
        return out

    def softmax_xent(self, logits, onehot):
        """tf.losses.softmax_cross_entropy (models/resnet_model.py:80-83): per-row
        -sum(y*log_softmax), reduced Sum/num_present == mean over the batch (fixture loss_nodes).
        Upstream: reduce_mean(softmax_cross_entropy_with_logits) — the same value."""
        z = logits.val
        lsm = z - torch.logsumexp(z, dim=1, keepdim=True)
        rows = -(onehot * lsm).sum(1)
        B = z.shape[0]
        loss = rows.sum() / B
        sm = torch.exp(lsm)

        def bwd():
            logits.add_grad((sm * onehot.sum(1, keepdim=True) - onehot) / B)
        self.tape.append(bwd)
        return loss, sm

    def backward(self):
        for fn in reversed(self.tape):
            fn()

    _bn_decay_flag = True
    _beta_first = False


# ======================================================================================
# Networks
# ======================================================================================
class OfficialResNetV2(Net):
    """resnet_model_official.py:41-366 — the net the reference's trainers actually build
    (models/resnet_model.py:72-77, resnet_size=50 hard-coded)."""

    def __init__(self, dataset="cifar10", resnet_size=50, num_classes=10, **kw):
        super().__init__(**kw)
        self.dataset = dataset
        self.resnet_size = resnet_size
        self.num_classes = num_classes
        self._bn_decay_flag = True   # models/resnet_model.py:88-89 decays ALL trainable vars

    # -- builders (names as in the reference) ------------------------------------------
    def batch_norm_relu(self, x, is_training):                      # :41-50
        scope = self._uniq("batch_normalization")
        relu = self._uniq("Relu")
        return self.batch_norm_act(x, scope, is_training, BN_EPS_OFFICIAL, BN_DECAY_OFFICIAL, True, 0.0,
                                   (scope + "/FusedBatchNorm", relu))

    def conv2d_fixed_padding(self, x, filters, kernel_size, strides):  # :53-91
        scope = self._uniq("conv2d")
        N, H, W, C = x.val.shape
        if strides > 1:   # fixed_padding + VALID
            pad = (kernel_size - 1) // 2
            Ho = (H + kernel_size - 1 - kernel_size) // strides + 1
            Wo = (W + kernel_size - 1 - kernel_size) // strides + 1
        else:             # SAME, stride 1
            pad = (kernel_size - 1) // 2
            Ho, Wo = H, W
        return self.conv2d(x, scope + "/kernel", kernel_size, C, filters, strides, pad, (Ho, Wo),
                           self.init_variance_scaling, scope + "/Conv2D")

    def building_block(self, x, filters, is_training, projection, strides):  # :94-130
        shortcut = x
        x = self.batch_norm_relu(x, is_training)
        if projection is not None:
            shortcut = projection(x)
        x = self.conv2d_fixed_padding(x, filters, 3, strides)
        x = self.batch_norm_relu(x, is_training)
        x = self.conv2d_fixed_padding(x, filters, 3, 1)
        return self.add(x, shortcut, self._uniq("add"))

    def bottleneck_block(self, x, filters, is_training, projection, strides):  # :133-175
        shortcut = x
        x = self.batch_norm_relu(x, is_training)
        if projection is not None:
            shortcut = projection(x)
        x = self.conv2d_fixed_padding(x, filters, 1, 1)
        x = self.batch_norm_relu(x, is_training)
        x = self.conv2d_fixed_padding(x, filters, 3, strides)
        x = self.batch_norm_relu(x, is_training)
        x = self.conv2d_fixed_padding(x, 4 * filters, 1, 1)
        return self.add(x, shortcut, self._uniq("add"))

    def block_layer(self, x, filters, bottleneck, blocks, strides, is_training):  # :178-214
        fn = self.bottleneck_block if bottleneck else self.building_block
        filters_out = 4 * filters if bottleneck else filters
        x = fn(x, filters, is_training, lambda t: self.conv2d_fixed_padding(t, filters_out, 1, strides), strides)
        for _ in range(1, blocks):
            x = fn(x, filters, is_training, None, 1)
        return x

    def forward(self, images, is_training=True):
        self.begin_pass()
        x = self._rec(Var(images.to(self.dtype)), "X")
        if self.dataset.startswith("cifar"):                          # :217-278
            if self.resnet_size % 6 != 2:
                raise ValueError("resnet_size must be 6n + 2:", self.resnet_size)
            n = (self.resnet_size - 2) // 6
            x = self.conv2d_fixed_padding(x, 16, 3, 1)
            x = self.block_layer(x, 16, False, n, 1, is_training)
            x = self.block_layer(x, 32, False, n, 2, is_training)
            x = self.block_layer(x, 64, False, n, 2, is_training)
            x = self.batch_norm_relu(x, is_training)
            x = self.avg_pool(x, 8, 1, "average_pooling2d/AvgPool", stored=False)
            feat = 64
        else:                                                           # :281-366
            table = {18: (False, [2, 2, 2, 2]), 34: (False, [3, 4, 6, 3]), 50: (True, [3, 4, 6, 3]),
                     101: (True, [3, 4, 23, 3]), 152: (True, [3, 8, 36, 3]), 200: (True, [3, 24, 36, 3])}
            if self.resnet_size not in table:
                raise ValueError("Not a valid resnet_size:", self.resnet_size)
            bott, layers = table[self.resnet_size]
            x = self.conv2d_fixed_padding(x, 64, 7, 2)
            x = self.max_pool_3x3_s2_same(x, "max_pooling2d/MaxPool")
            for i, (f, s) in enumerate(zip([64, 128, 256, 512], [1, 2, 2, 2])):
                x = self.block_layer(x, f, bott, layers[i], s, is_training)
            x = self.batch_norm_relu(x, is_training)
            x = self.avg_pool(x, x.val.shape[1], 1, "average_pooling2d/AvgPool", stored=False)
            feat = 2048 if bott else 512
        logits = self.dense(x, "dense/kernel", "dense/bias", feat, self.num_classes,
                            self.init_glorot_uniform, "dense/BiasAdd", decay_bias=True)
        return logits


class UpstreamResNet(Net):
    """The north-star-named builders `_conv/_batch_norm/_relu/_residual/_bottleneck_residual/
    _fully_connected` of upstream tensorflow/models research/resnet/resnet_model.py — NOT in
    /root/reference (only call-site evidence: resnet_single.py:53-56,194-202).  Restated from
    the published upstream file as summarised in SURVEY.md a23.  PARITY UNPINNED."""

    def __init__(self, num_classes=10, num_residual_units=5, use_bottleneck=False, relu_leakiness=0.1,
                 filters=None, **kw):
        super().__init__(**kw)
        self.num_classes = num_classes
        self.units = num_residual_units
        self.use_bottleneck = use_bottleneck
        self.leak = relu_leakiness
        self.filters = filters or ([16, 64, 128, 256] if use_bottleneck else [16, 16, 32, 64])
        self._bn_decay_flag = False   # `_decay` only decays names containing 'DW'
        self._beta_first = True

    def _init_conv(self, shape):
        n = shape[0] * shape[1] * shape[3]
        return torch.randn(shape, dtype=torch.float64, generator=self._gen) * math.sqrt(2.0 / n)

    def _init_fc(self, shape):
        # tf.uniform_unit_scaling_initializer(factor=1.0): U(-sqrt(3/dim_in), +sqrt(3/dim_in))
        lim = math.sqrt(3.0 / shape[0])
        return (torch.rand(shape, dtype=torch.float64, generator=self._gen) * 2 - 1) * lim

    def _conv(self, name, x, k, cin, cout, stride):
        N, H, W, C = x.val.shape
        Ho, ph = tf_same_pad(H, k, stride)
        Wo, _ = tf_same_pad(W, k, stride)
        return self.conv2d(x, name + "/DW", k, cin, cout, stride, ph, (Ho, Wo), self._init_conv, name + "/Conv2D")

    def _bn_relu(self, scope, x, is_training):
        return self.batch_norm_act(x, scope, is_training, BN_EPS_UPSTREAM, BN_DECAY_UPSTREAM, False, self.leak,
                                   (scope + "/batchnorm", scope + "/leaky_relu"))

    def _residual(self, s, x, cin, cout, stride, activate_before_residual, is_training):
        if activate_before_residual:
            x = self._bn_relu(s + "/shared_activation/init_bn", x, is_training)
            orig = x
        else:
            orig = x
            x = self._bn_relu(s + "/residual_only_activation/init_bn", x, is_training)
        x = self._conv(s + "/sub1/conv1", x, 3, cin, cout, stride)
        x = self._bn_relu(s + "/sub2/bn2", x, is_training)
        x = self._conv(s + "/sub2/conv2", x, 3, cout, cout, 1)
        if cin != cout:
            if stride > 1:      # (avg_pool + channel pad is ONE kernel in the CUDA path: one rounding)
                orig = self.avg_pool(orig, stride, stride, s + "/sub_add/AvgPool", stored=False)
            orig = self.pad_channels(orig, (cout - cin) // 2, (cout - cin) // 2, s + "/sub_add/Pad")
        return self.add(x, orig, s + "/sub_add/add")

    def _bottleneck_residual(self, s, x, cin, cout, stride, activate_before_residual, is_training):
        if activate_before_residual:
            x = self._bn_relu(s + "/common_bn_relu/init_bn", x, is_training)
            orig = x
        else:
            orig = x
            x = self._bn_relu(s + "/residual_bn_relu/init_bn", x, is_training)
        x = self._conv(s + "/sub1/conv1", x, 1, cin, cout // 4, stride)
        x = self._bn_relu(s + "/sub2/bn2", x, is_training)
        x = self._conv(s + "/sub2/conv2", x, 3, cout // 4, cout // 4, 1)
        x = self._bn_relu(s + "/sub3/bn3", x, is_training)
        x = self._conv(s + "/sub3/conv3", x, 1, cout // 4, cout, 1)
        if cin != cout:
            orig = self._conv(s + "/sub_add/project", orig, 1, cin, cout, stride)
        return self.add(x, orig, s + "/sub_add/add")

    def forward(self, images, is_training=True):
        self.begin_pass()
        x = self._rec(Var(images.to(self.dtype)), "X")
        f = self.filters
        x = self._conv("init/init_conv", x, 3, x.val.shape[-1], f[0], 1)
        res = self._bottleneck_residual if self.use_bottleneck else self._residual
        strides = [1, 2, 2]
        abr = [True, False, False]
        for stage in range(3):
            x = res("unit_%d_0" % (stage + 1), x, f[stage], f[stage + 1], strides[stage], abr[stage], is_training)
            for i in range(1, self.units):
                x = res("unit_%d_%d" % (stage + 1, i), x, f[stage + 1], f[stage + 1], 1, False, is_training)
        x = self._bn_relu("unit_last/final_bn", x, is_training)
        x = self.avg_pool(x, x.val.shape[1], 1, "unit_last/global_avg_pool", stored=False)
        return self.dense(x, "logit/DW", "logit/biases", f[3], self.num_classes, self._init_fc,
                          "logit/xw_plus_b", decay_bias=False)


# ======================================================================================
# Loss, optimizer, aggregation, schedules
# ======================================================================================
def loss_and_grads(net, images, onehot, weight_decay, is_training=True):
    """models/resnet_model.py:69-91 + autodiff (`optimizer.minimize`, :125).
    cost = xent + wd * sum_v l2_loss(v) (l2_loss = sum(v^2)/2) over net.decay_mask; the
    per-variable gradient is AddN(wgrad, wd*v) (fixture grad_addn_structures)."""
    logits = net.forward(images, is_training)
    onehot = onehot.to(net.dtype)
    xent, probs = net.softmax_xent(logits, onehot)
    l2 = sum((v * v).sum() / 2 for n, v in net.params.items() if net.decay_mask[n])
    cost = xent + weight_decay * l2
    net.backward()
    grads = OrderedDict()
    for n, v in net.params.items():
        g = net.param_grads[n]
        if net.decay_mask[n]:
            g = g + weight_decay * v
        grads[n] = g
    return {"logits": logits.val, "predictions": probs, "xent": xent, "cost": cost, "grads": grads}


def apply_bn_updates(net, decay):
    """UPDATE_OPS (models/resnet_model.py:123-125): mov -= (mov - batch) * (1 - decay)."""
    for scope, (mean, var) in net.bn_updates.items():
        net.state[scope + "/moving_mean"] -= (net.state[scope + "/moving_mean"] - mean) * (1 - decay)
        net.state[scope + "/moving_variance"] -= (net.state[scope + "/moving_variance"] - var) * (1 - decay)


class MomentumSGD:
    """tf.train.MomentumOptimizer(lr, 0.9) / GradientDescentOptimizer (models/resnet_model.py:98-101):
    ApplyMomentum(use_nesterov=False): m = momentum*m + g ; w -= lr*m."""

    def __init__(self, net, optimizer="mom", momentum=0.9):
        self.kind = optimizer
        self.momentum = momentum
        self.slots = OrderedDict((n, torch.zeros_like(v)) for n, v in net.params.items())

    def apply(self, net, grads, lr):
        for n, v in net.params.items():
            g = grads[n]
            if self.kind == "mom":
                self.slots[n].mul_(self.momentum).add_(g)
                v.sub_(lr * self.slots[n])
            elif self.kind == "sgd":
                v.sub_(lr * g)
            else:
                raise ValueError("optimizer must be 'sgd' or 'mom'")


def aggregate_mean(list_of_grads):
    """SyncReplicasOptimizer (models/resnet_model.py:104-116) ConditionalAccumulator.take_grad(R):
    the arithmetic mean of the R accepted gradients; identical to hvd.DistributedOptimizer's
    averaged allreduce (:118-120)."""
    R = len(list_of_grads)
    out = OrderedDict()
    for n in list_of_grads[0]:
        acc = list_of_grads[0][n].clone()
        for g in list_of_grads[1:]:
            acc = acc + g[n]
        out[n] = acc / R
    return out


def cifar_lr(train_step):
    """_LearningRateSetterHook.after_run, resnet_cifar_train.py:301-311."""
    if train_step < 40000:
        return 0.1
    if train_step < 60000:
        return 0.01
    if train_step < 80000:
        return 0.001
    return 0.0001


def imagenet_lr(train_step):
    """_LearningRateSetterHook.after_run, resnet_imagenet_train.py:236-260 (Intel-Caffe 8-node)."""
    if train_step < 6240:
        return 0.1 + 0.3 * train_step / 6240.0
    if train_step < 37440:
        return 0.4
    if train_step < 74880:
        return 0.4 * 0.1
    if train_step < 99840:
        return 0.4 * 0.01
    return 0.4 * 0.001


class Trainer:
    """One synchronous data-parallel step over R replicas (each with its own BN statistics —
    the reference has no sync-BN; under Horovod rank 0's moving stats are the checkpointed ones)."""

    def __init__(self, net, weight_decay, optimizer="mom", bn_decay=BN_DECAY_OFFICIAL):
        self.net = net
        self.wd = weight_decay
        self.opt = None
        self.kind = optimizer
        self.bn_decay = bn_decay
        self.global_step = 0

    def step(self, shards, lr):
        """shards: list of (images, onehot) — one per replica; replica 0's BN updates are kept."""
        outs = []
        bn0 = None
        for r, (img, lab) in enumerate(shards):
            o = loss_and_grads(self.net, img, lab, self.wd, True)
            outs.append(o)
            if r == 0:
                bn0 = OrderedDict(self.net.bn_updates)
        if self.opt is None:
            self.opt = MomentumSGD(self.net, self.kind)
        g = aggregate_mean([o["grads"] for o in outs])
        self.net.bn_updates = bn0
        apply_bn_updates(self.net, self.bn_decay)
        self.opt.apply(self.net, g, lr)
        self.global_step += 1
        return outs
