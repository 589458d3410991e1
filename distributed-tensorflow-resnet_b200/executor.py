"""Static executor: turns a `graph.Graph` into flat HBM buffers and an ordered list of CUDA
kernel launches (forward, backward, aggregation, optimizer) through the C ABI of librn_b200.so.

PyTorch is only the shell here: `torch.empty` for device memory, `data_ptr()` for addresses,
`torch.cuda.current_stream()` for the stream handle, `torch.cuda.CUDAGraph` to replay the launch
list, and `torch.distributed` (NCCL) for the gradient all-reduce.  No torch operator computes
anything on the hot path, and there is no CPU fallback.

HBM layout
  W   fp32 [n]   all trainable variables, TF creation order, each start aligned to 64 elements
  G   fp32 [n]   gradients, same offsets (the fusion buffer that is all-reduced, SUM over ranks)
  Mom fp32 [n]   momentum slots, same offsets
  Wb  bf16 [n]   bf16 copy of W (bf16 mode), same offsets, refreshed by the optimizer kernel
  S   fp32       non-trainable state: BN moving_mean / moving_variance
  per-BN saved batch mean / rstd, activations (NHWC) and their gradients: one buffer each.
"""
import ctypes as C
import os
import math
from collections import OrderedDict

import numpy as np
import torch

from . import graph as G
from . import _lib
from ._lib import lib, check, ptr, ConvDesc, RN_F32, RN_BF16

GRANULE = 64


def _align(n, a):
    return (n + a - 1) // a * a


def init_variable(var, gen):
    """Initialisers pinned by the training-graph fixture (tests/golden/cifar50_train_graph.json):
    conv kernels truncated normal (2 sigma) with stddev sqrt(1/fan_in)/0.8796; dense kernel
    Glorot uniform; gamma / moving_variance ones; the rest zeros.  Upstream builders use
    random_normal / uniform_unit_scaling."""
    k = var.init["kind"]
    if k == "zeros":
        return torch.zeros(var.shape, dtype=torch.float32)
    if k == "ones":
        return torch.ones(var.shape, dtype=torch.float32)
    if k == "truncated_normal":
        std = var.init["stddev"]
        t = torch.empty(var.shape, dtype=torch.float32)
        torch.nn.init.trunc_normal_(t, 0.0, std, -2 * std, 2 * std, generator=gen)
        return t
    if k == "normal":
        return torch.randn(var.shape, dtype=torch.float32, generator=gen) * var.init["stddev"]
    if k == "uniform":
        lim = var.init["limit"]
        return (torch.rand(var.shape, dtype=torch.float32, generator=gen) * 2 - 1) * lim
    raise ValueError("unknown initialiser %r" % k)


class _Slot:
    """A gradient buffer, possibly shared by several tensors (the residual-stream gradient)."""

    def __init__(self, buf):
        self.buf = buf
        self.written = False
        self.members = []


class Executor:
    def __init__(self, graph, *, device=None, dtype="bf16", weight_decay=0.0, optimizer="mom", momentum=0.9,
                 lrn_rate=0.1, train=True, seed=0, process_group=None, world_size=1, rank=0,
                 bucket_bytes=32 << 20, use_tensor_cores=True, loss=None, init_values=None):
        if not torch.cuda.is_available():
            raise _lib.RnError("no CUDA device: this framework has no CPU path")
        self.g = graph
        self.dev = torch.device(device if device is not None else "cuda:%d" % torch.cuda.current_device())
        self.bf16 = dtype in ("bf16", torch.bfloat16)
        self.adt = torch.bfloat16 if self.bf16 else torch.float32
        self.rn_dt = RN_BF16 if self.bf16 else RN_F32
        self.wd = float(weight_decay)
        self.opt_kind = optimizer
        if optimizer not in ("sgd", "mom"):
            raise ValueError("optimizer must be 'sgd' or 'mom'")
        self.momentum = float(momentum)
        self.train = train
        self.pg = process_group
        self.world = int(world_size)
        self.rank = int(rank)
        self.use_tc = bool(use_tensor_cores) and self.bf16
        import os as _os
        self.overlap_wgrad = train and not _os.environ.get("RN_NO_OVERLAP")
        self.wg_stream = torch.cuda.Stream(device=self.dev) if self.overlap_wgrad else None
        self.loss = loss   # (probs, scalars) tensors from graph.softmax_xent
        self.global_step = 0
        self.launches_per_step = 0
        self._graph_exec = None
        self._plans = []
        self._keep = []
        self.meta = {}      # launch name -> tensors it reads / writes (for the step-through verifier in tests/)
        self._layout(seed, init_values)
        self._alloc_activations()
        self._build_forward()
        if train:
            self._build_backward()
            self._build_update(bucket_bytes)
        self.set_lr(lrn_rate)
        if self.world > 1:
            self.broadcast_from_rank0()

    # ------------------------------------------------------------------ parameters
    def _layout(self, seed, init_values):
        tv = self.g.trainable_variables()
        off = 0
        for v in tv:
            v.offset = off
            off += _align(v.size, GRANULE)
        self.n_flat = max(off, GRANULE)
        sv = [v for v in self.g.variables.values() if not v.trainable]
        soff = 0
        for v in sv:
            v.offset = soff
            soff += _align(v.size, 4)
        dev = self.dev
        self.W = torch.zeros(self.n_flat, dtype=torch.float32, device=dev)
        self.S = torch.zeros(max(soff, 4), dtype=torch.float32, device=dev)
        self.Wb = torch.zeros(self.n_flat, dtype=torch.bfloat16, device=dev) if self.bf16 else None
        if self.train:
            self.Gd = torch.zeros(self.n_flat, dtype=torch.float32, device=dev)
            self.Mom = torch.zeros(self.n_flat, dtype=torch.float32, device=dev)
        mask = np.zeros(self.n_flat // GRANULE, dtype=np.uint8)
        gen = torch.Generator().manual_seed(seed)
        hostW = torch.zeros(self.n_flat, dtype=torch.float32)
        hostS = torch.zeros(self.S.numel(), dtype=torch.float32)
        for v in self.g.variables.values():
            if init_values is not None and v.name in init_values:
                val = torch.as_tensor(np.asarray(init_values[v.name]), dtype=torch.float32).reshape(v.shape)
            else:
                val = init_variable(v, gen)
            if v.trainable:
                hostW[v.offset:v.offset + v.size] = val.reshape(-1)
                if v.decay:
                    mask[v.offset // GRANULE:(v.offset + _align(v.size, GRANULE)) // GRANULE] = 1
            else:
                hostS[v.offset:v.offset + v.size] = val.reshape(-1)
        self.W.copy_(hostW)
        self.S.copy_(hostS)
        if self.Wb is not None:
            self.Wb.copy_(hostW.to(torch.bfloat16))   # set-up only; per step the optimizer kernel refreshes it
        self.all_decay = bool(mask.all())
        self.mask = None if self.all_decay else torch.from_numpy(mask).to(dev)
        self.hyper = torch.zeros(4, dtype=torch.float32, device=dev)
        self.sumsq = torch.zeros(1, dtype=torch.float32, device=dev)
        nbytes = max(lib.rn_bn_scratch_bytes(1, 8), lib.rn_optimizer_scratch_bytes(self.n_flat))
        self.scratch = torch.zeros(nbytes, dtype=torch.uint8, device=dev)     # ticket counters start at 0
        self.scratch_bytes = nbytes

    def var_view(self, name, which="W"):
        v = self.g.variables[name]
        base = {"W": self.W, "G": getattr(self, "Gd", None), "M": getattr(self, "Mom", None), "S": self.S}[
            which if v.trainable else "S"]
        return base[v.offset:v.offset + v.size].view(v.shape)

    def _wptr(self, v, which="W"):
        base = {"W": self.W, "G": getattr(self, "Gd", None), "S": self.S, "Wb": self.Wb}[which]
        return base.data_ptr() + v.offset * base.element_size()

    def state_dict(self):
        """All variables by TF name (weights, BN moving stats) + momentum slots + global_step."""
        out = OrderedDict()
        for v in self.g.variables.values():
            out[v.name] = self.var_view(v.name).detach().cpu().numpy().copy()
            if v.trainable and self.train:
                out[v.name + "/Momentum"] = self.var_view(v.name, "M").detach().cpu().numpy().copy()
        out["global_step"] = np.int64(self.global_step)
        return out

    def load_state_dict(self, sd):
        for v in self.g.variables.values():
            if v.name in sd:
                self.var_view(v.name).copy_(torch.as_tensor(np.asarray(sd[v.name]), dtype=torch.float32).reshape(v.shape))
            if v.trainable and self.train and v.name + "/Momentum" in sd:
                self.var_view(v.name, "M").copy_(torch.as_tensor(np.asarray(sd[v.name + "/Momentum"]),
                                                                 dtype=torch.float32).reshape(v.shape))
        if "global_step" in sd:
            self.global_step = int(sd["global_step"])
        self.refresh_shadow_weights()

    def refresh_shadow_weights(self):
        """Rebuild the bf16 operand copies from W (set-up / checkpoint restore only)."""
        st = torch.cuda.current_stream().cuda_stream
        if self.Wb is not None:
            check(lib.rn_cast(ptr(self.W), ptr(self.Wb), self.n_flat, RN_F32, RN_BF16, st), "rn_cast")
        for fn in getattr(self, "_relayout", []):
            fn(st)

    # ------------------------------------------------------------------ activations
    def _alloc_activations(self):
        for t in self.g.tensors.values():
            dt = self.adt if t.dtype == G.ACT else torch.float32
            bound = getattr(t, "bound", None)
            if bound is not None and bound.dtype == dt and bound.is_contiguous() and bound.device == self.dev:
                t.buf = bound
                t.cast_from = None
            else:
                t.buf = torch.zeros(t.shape, dtype=dt, device=self.dev)
                t.cast_from = bound
        # requires_grad propagation
        for t in self.g.tensors.values():
            t.requires_grad = False
        for n in self.g.nodes:
            rg = any(v.trainable for v in n.vars.values()) or any(t.requires_grad for t in n.inputs)
            for t in n.outputs:
                t.requires_grad = rg

    # ------------------------------------------------------------------ forward
    def _desc(self, n):
        x, y = n.inputs[0], n.outputs[0]
        N, H, W, Cin = x.shape
        _, Ho, Wo, Cout = y.shape
        return ConvDesc(N, H, W, Cin, Ho, Wo, Cout, n.attrs["k"], n.attrs["stride"], n.attrs["pad"], self.rn_dt)

    def _fused_add_for(self, bn_node):
        """The `add` node whose output this training-mode BN normalises, if the add can be folded
        into the statistics pass (BN must be the first reader of the sum)."""
        x = bn_node.inputs[0]
        p = x.producer
        if p is not None and p.op == "add" and bn_node.attrs["is_training"] and x.consumers and x.consumers[0] is bn_node:
            return p
        return None

    def _fusable_conv(self, t):
        """The conv node producing tensor t if its tensor-core fprop can carry a fused epilogue."""
        p = t.producer
        if p is None or p.op != "conv2d" or hasattr(p, "fuse") or not self.use_tc or not self.bf16:
            return None
        d = self._desc(p)
        if self._is_stem(p) or not lib.rn_convtc_fused_supported(C.byref(d)):
            return None
        return p

    def _plan_epilogue_fusion(self):
        """Decide which convolutions carry a fused epilogue (RN_NO_EPILOGUE_FUSION=1 disables it):
          * a training-mode BN fed directly by a conv: the conv epilogue accumulates the per-channel sums, the
            statistics pass over the tensor is replaced by rn_bn_finalize;
          * a training-mode BN fed by `conv + shortcut` (the bottleneck's residual add): the conv epilogue adds
            the shortcut tile, stores the sum and accumulates its statistics — the conv output itself is never
            materialised (nothing in backward reads it)."""
        if os.environ.get("RN_NO_EPILOGUE_FUSION"):
            return
        for n in self.g.nodes:
            if n.op != "bn_act" or not n.attrs["is_training"]:
                continue
            x = n.inputs[0]
            add = self._fused_add_for(n)
            if add is not None:
                for i in (0, 1):
                    t, other = add.inputs[i], add.inputs[1 - i]
                    cv = self._fusable_conv(t)
                    if cv is not None and len(t.consumers) == 1 and t.shape == x.shape and other.shape == x.shape:
                        cv.fuse = dict(out=x, residual=other, bn=n)
                        n.stats_from = cv
                        break
            else:
                cv = self._fusable_conv(x)
                if cv is not None:
                    cv.fuse = dict(out=x, residual=None, bn=n)
                    n.stats_from = cv

    def _plan_operand_fusion(self):
        """Training-mode BN+ReLU nodes whose output is read ONLY by 1x1 tensor-core convolutions (conv1 / conv3 of a
        bottleneck block, projection shortcuts) are not materialised: those convolutions read the BN's INPUT and
        normalise it in shared memory (rn_convtc_plan_create_bnin), in fprop and in wgrad, so the bn_apply pass and the
        activated tensor disappear from HBM.  OPT-IN (RN_BN_OPERAND_FUSION=1): measured on B200 (ResNet-50 / 224 /
        batch 128, profiles/r02_bn_operand_ab.txt) the step gets SLOWER, 9.56 -> 9.93 ms — 32 bn_apply launches (0.76 ms)
        disappear, but rewriting every operand tile in shared memory costs the 1x1 fprops +0.32 ms and the 1x1 wgrads
        +0.80 ms: the transform adds a shared-memory read and write of every operand byte to kernels that are already
        bound by shared-memory operand traffic (eight transform warps instead of four made it slower still, 10.23 ms)."""
        if os.environ.get("RN_BN_OPERAND_FUSION") != "1" or not (self.use_tc and self.bf16):
            return
        for n in self.g.nodes:
            if n.op != "bn_act" or not n.attrs["is_training"]:
                continue
            y = n.outputs[0]
            if not y.consumers:
                continue
            ok = True
            for c in y.consumers:
                if c.op != "conv2d" or c.inputs[0] is not y or self._is_stem(c):
                    ok = False
                    break
                d = self._desc(c)
                if not (lib.rn_convtc_bnin_supported(C.byref(d), 0) and (not self.train or lib.rn_convtc_bnin_supported(C.byref(d), 2))):
                    ok = False
                    break
            n.virtual = ok

    def _bn_operand(self, t):
        """(buffer to read, BN node or None): the BN input + the node when `t` is a BN output that is not materialised."""
        p = t.producer
        if p is not None and p.op == "bn_act" and getattr(p, "virtual", False):
            return p.inputs[0].buf, p
        return t.buf, None

    def _bnin_args(self, bn):
        v = bn.vars
        return (self._wptr(v["gamma"]), self._wptr(v["beta"]), bn.saved[0].data_ptr(), bn.saved[1].data_ptr(), bn.attrs["leak"])

    def _elide_cast(self, t, src):
        """An fp32 image batch bound to a bf16 placeholder needs no cast launch when its one consumer is a stem
        convolution (space-to-depth pack or unfold: both read either dtype).  RN_NO_CAST_ELISION=1 keeps the cast."""
        if os.environ.get("RN_NO_CAST_ELISION", "0") not in ("", "0") or src.dtype != torch.float32 or t.buf.dtype != torch.bfloat16:
            return False
        if not (src.is_contiguous() and src.device == self.dev and len(t.consumers) == 1):
            return False
        n = t.consumers[0]
        return n.op == "conv2d" and n.inputs[0] is t and self._is_stem(n)

    def materialize(self, t):
        """The values of tensor `t`: its buffer, or — for a BN output that only exists inside the consuming
        convolutions' shared memory — what rn_bn_apply_act_fwd stores for it (verification / debugging only)."""
        if getattr(t, "cast_elided", False):
            return t.cast_from.to(t.buf.dtype)
        if t.buf is not None:
            return t.buf
        xb, bn = self._bn_operand(t)
        if bn is None:
            raise ValueError("tensor %s has no buffer" % t.name)
        out = torch.empty_like(xb)
        C_ = xb.shape[-1]
        a = bn.attrs
        check(lib.rn_bn_apply_act_fwd(ptr(xb), ptr(out), xb.numel() // C_, C_, self.rn_dt, *self._bnin_args(bn)[:4], 0, a["eps"],
                                      a["leak"], torch.cuda.current_stream().cuda_stream), "rn_bn_apply_act_fwd")
        return out

    def _build_forward(self):
        fwd = []
        g = self.g
        self._saved = {}
        fused_adds = set()
        for n in g.nodes:
            if n.op == "bn_act":
                a = self._fused_add_for(n)
                if a is not None:
                    fused_adds.add(id(a))
        self._fused_adds = fused_adds
        self._plan_epilogue_fusion()
        self._plan_operand_fusion()
        for t in g.placeholders.values():
            src = getattr(t, "cast_from", None)
            if src is not None and self._elide_cast(t, src):
                # the only reader is the stem convolution's pack / unfold kernel, which converts on the fly (same
                # round-to-nearest): the bf16 copy of the image batch is never written
                t.cast_elided = True
                self._keep.append(src)
                continue
            if src is not None:
                sdt = RN_BF16 if src.dtype == torch.bfloat16 else RN_F32
                ddt = RN_BF16 if t.buf.dtype == torch.bfloat16 else RN_F32
                if src.dtype not in (torch.float32, torch.bfloat16):
                    raise ValueError("placeholder %s: unsupported dtype %s" % (t.name, src.dtype))
                fwd.append(("cast:" + t.name, lambda s, a=ptr(src), b=ptr(t.buf), k=t.numel, sd=sdt, dd=ddt:
                            check(lib.rn_cast(a, b, k, sd, dd, s), "rn_cast")))
                self._keep.append(src)
        for n in g.nodes:
            if n.op == "conv2d":
                if getattr(n, "fuse", None) and n.fuse["residual"] is not None:
                    continue            # launched where its add stood: the shortcut may be produced after this node
                fwd.append(self._conv_fprop(n))
            elif n.op == "bn_act":
                fwd.extend(self._bn_fwd(n))
            elif n.op == "add":
                if id(n) in fused_adds:
                    for t in n.inputs:
                        cv = t.producer
                        if cv is not None and getattr(cv, "fuse", None) and cv.fuse["out"] is n.outputs[0]:
                            item = self._conv_fprop(cv)
                            fwd.append(item)
                            self.meta[item[0]] = dict(node=cv)
                    continue
                a, b = n.inputs
                y = n.outputs[0]
                dt = RN_BF16 if y.buf.dtype == torch.bfloat16 else RN_F32
                fwd.append(("add:" + y.name, lambda s, pa=ptr(a.buf), pb=ptr(b.buf), py=ptr(y.buf), k=y.numel, dt=dt:
                            check(lib.rn_add(pa, pb, py, k, dt, s), "rn_add")))
            elif n.op == "maxpool":
                x, y = n.inputs[0], n.outputs[0]
                n.argmax = torch.zeros(y.shape, dtype=torch.uint8, device=self.dev)
                N, H, W, Cc = x.shape
                fwd.append(("maxpool:" + y.name,
                            lambda s, px=ptr(x.buf), py=ptr(y.buf), pa=ptr(n.argmax), d=(N, H, W, Cc, y.shape[1], y.shape[2], n.attrs["pad"]):
                            check(lib.rn_maxpool3x3s2_fwd(px, py, pa, *d, self.rn_dt, s), "rn_maxpool3x3s2_fwd")))
            elif n.op == "avgpool":
                x, y = n.inputs[0], n.outputs[0]
                N, H, W, Cc = x.shape
                idt = RN_BF16 if x.buf.dtype == torch.bfloat16 else RN_F32
                odt = RN_BF16 if y.buf.dtype == torch.bfloat16 else RN_F32
                fwd.append(("avgpool:" + y.name,
                            lambda s, px=ptr(x.buf), py=ptr(y.buf), d=(N, H, W, Cc, n.attrs["k"], n.attrs["s"], y.shape[3], n.attrs["c_before"], idt, odt):
                            check(lib.rn_avgpool_fwd(px, py, *d, s), "rn_avgpool_fwd")))
            elif n.op == "dense":
                x, y = n.inputs[0], n.outputs[0]
                if x.buf.dtype != torch.float32:
                    raise ValueError("dense expects fp32 features (use avg_pool(out_dtype=F32))")
                B = x.shape[0]
                D = x.numel // B
                K = y.shape[1]
                nb = lib.rn_dense_scratch_bytes(B, D, K)
                n.scratch = torch.zeros(max(nb, 16), dtype=torch.uint8, device=self.dev)
                fwd.append(("dense:" + y.name,
                            lambda s, px=ptr(x.buf), pw=self._wptr(n.vars["kernel"]), pb=self._wptr(n.vars["bias"]), py=ptr(y.buf), d=(B, D, K), sc=(ptr(n.scratch), nb):
                            check(lib.rn_dense_fwd(px, pw, pb, py, *d, sc[0], sc[1], s), "rn_dense_fwd")))
            elif n.op == "softmax_xent":
                logits, labels = n.inputs
                probs, scal = n.outputs
                B, K = logits.shape
                if self.train and logits.requires_grad:
                    n.dlogits = torch.zeros(logits.shape, dtype=torch.float32, device=self.dev)
                else:
                    n.dlogits = None
                fwd.append(("softmax_xent",
                            lambda s, pl=ptr(logits.buf), py=ptr(labels.buf), pp=ptr(probs.buf), pd=ptr(n.dlogits), ps=ptr(scal.buf), d=(B, K):
                            check(lib.rn_softmax_xent(pl, py, d[0], d[1], pp, pd, ps, s), "rn_softmax_xent")))
            else:
                raise ValueError("unknown op " + n.op)
            for nm, _ in fwd:
                if nm not in self.meta:
                    self.meta[nm] = dict(node=None if nm.startswith("cast:") else n)
        self.fwd = fwd

    def _conv_impl(self, n, kind):
        d = self._desc(n)
        if self.use_tc and lib.rn_convtc_supported(C.byref(d), kind):
            return "tc"
        return "simt"

    def _is_stem(self, n):
        """Convolutions over fewer than 16 channels (the image): unfolded once, then 1x1 tensor-core GEMMs."""
        d = self._desc(n)
        if not (self.use_tc and d.Cin < 16):
            return False
        kpad = _align(d.k * d.k * d.Cin, 32)
        d1 = ConvDesc(d.N, d.Ho, d.Wo, kpad, d.Ho, d.Wo, d.Cout, 1, 1, 0, self.rn_dt)
        return bool(lib.rn_convtc_supported(C.byref(d1), 0))

    def _stem_s2d(self, n):
        """Stride-2 stem (ImageNet 7x7/2): space-to-depth packing + 4 vertical taps over 64 channels instead
        of the unfolded matrix (rn_stem_s2d_* in include/rn_b200.h).  RN_STEM_IM2COL=1 forces the unfold path."""
        d = self._desc(n)
        return (self._is_stem(n) and not os.environ.get("RN_STEM_IM2COL")
                and bool(lib.rn_stem_s2d_supported(C.byref(d))))

    def _stem_s2d_setup(self, n):
        if hasattr(n, "_s2d"):
            return n._s2d
        d = self._desc(n)
        kv = n.vars["kernel"]
        H2, W2, T = C.c_int32(), C.c_int32(), C.c_int32()
        check(lib.rn_stem_s2d_geometry(C.byref(d), C.byref(H2), C.byref(W2), C.byref(T)), "rn_stem_s2d_geometry")
        u = torch.zeros(d.N, H2.value, W2.value, 16, dtype=torch.bfloat16, device=self.dev)
        wp = torch.zeros(T.value, 64, d.Cout, dtype=torch.bfloat16, device=self.dev)
        fn = (lambda s, d=d, a=self._wptr(kv, "Wb"), b=ptr(wp):
              check(lib.rn_stem_s2d_weights(C.byref(d), a, b, s), "rn_stem_s2d_weights"))
        if not hasattr(self, "_relayout"):
            self._relayout = []
        self._relayout.append(fn)
        fn(torch.cuda.current_stream().cuda_stream)
        self._keep += [d, u, wp]
        n._s2d = dict(d=d, u=u, wp=wp, taps=T.value)
        return n._s2d

    def _s2d_plan(self, d, kind, u, b, out):
        nbytes = lib.rn_stem_s2d_scratch_bytes(C.byref(d), kind)
        scratch = torch.zeros(max(nbytes, 16), dtype=torch.uint8, device=self.dev)
        self._keep.append(scratch)
        plan = C.c_void_p()
        check(lib.rn_stem_s2d_plan_create(C.byref(d), kind, ptr(u), ptr(b), ptr(out), ptr(scratch), nbytes,
                                          C.byref(plan)), "rn_stem_s2d_plan_create")
        self._plans.append(plan)
        fn = lambda s, p=plan: check(lib.rn_convtc_run(p, 0.0, s), "rn_convtc_run")   # noqa: E731
        fn.plan = plan
        return fn

    def _stem_setup(self, n):
        if hasattr(n, "_stem"):
            return n._stem
        d = self._desc(n)
        kv = n.vars["kernel"]
        rows = d.k * d.k * d.Cin
        kpad = _align(rows, 32)
        d1 = ConvDesc(d.N, d.Ho, d.Wo, kpad, d.Ho, d.Wo, d.Cout, 1, 1, 0, self.rn_dt)
        cols = torch.zeros(d.N * d.Ho * d.Wo, kpad, dtype=torch.bfloat16, device=self.dev)
        # [Kpad][Cout] bf16: the first k*k*Cin rows are the HWIO kernel (refreshed each step), the rest stay 0
        wf = torch.zeros(kpad, d.Cout, dtype=torch.bfloat16, device=self.dev)
        fn = (lambda s, a=self._wptr(kv, "Wb"), b=ptr(wf), k=rows * d.Cout:
              check(lib.rn_cast(a, b, k, RN_BF16, RN_BF16, s), "rn_cast"))
        if not hasattr(self, "_relayout"):
            self._relayout = []
        self._relayout.append(fn)
        fn(torch.cuda.current_stream().cuda_stream)
        self._keep += [d, d1, cols, wf]
        n._stem = dict(d=d, d1=d1, cols=cols, wf=wf, rows=rows, kpad=kpad)
        return n._stem

    def _plan_raw(self, d, kind, a_buf, b_buf, out):
        nbytes = lib.rn_convtc_scratch_bytes(C.byref(d), kind)
        scratch = torch.zeros(max(nbytes, 16), dtype=torch.uint8, device=self.dev)
        self._keep.append(scratch)
        plan = C.c_void_p()
        outp = out if isinstance(out, int) else ptr(out)
        check(lib.rn_convtc_plan_create(C.byref(d), kind, ptr(a_buf), ptr(b_buf), outp, ptr(scratch), nbytes,
                                        C.byref(plan)), "rn_convtc_plan_create")
        self._plans.append(plan)
        fn = lambda s, p=plan, beta=0.0: check(lib.rn_convtc_run(p, beta, s), "rn_convtc_run")   # noqa: E731
        fn.plan = plan
        return fn

    def _conv_fprop(self, n):
        x, y = n.inputs[0], n.outputs[0]
        d = self._desc(n)
        self._keep.append(d)
        kv = n.vars["kernel"]
        xbuf, bn_in = self._bn_operand(x)
        if bn_in is not None:
            f = getattr(n, "fuse", None)
            res_p = st_p = None
            out_buf = y.buf
            if f is not None:
                rows = lib.rn_convtc_stat_rows(C.byref(d))
                f["partials"] = torch.zeros(rows, d.Cout, 2, dtype=torch.float32, device=self.dev)
                res_p = ptr(f["residual"].buf) if f["residual"] is not None else None
                st_p = ptr(f["partials"])
                out_buf = f["out"].buf
            nrows = C.c_int32(0)
            plan = C.c_void_p()
            check(lib.rn_convtc_plan_create_bnin(C.byref(d), 0, ptr(xbuf), *self._bnin_args(bn_in), self._wptr(kv, "Wb"),
                                                 ptr(out_buf), res_p, st_p, C.byref(nrows), None, 0, C.byref(plan)),
                  "rn_convtc_plan_create_bnin")
            self._plans.append(plan)
            if f is not None:
                f["nrows"] = nrows.value
                f["plan"] = plan
                if f["residual"] is not None:
                    y.buf = None
            fn = lambda s, p=plan: check(lib.rn_convtc_run(p, 0.0, s), "rn_convtc_run")   # noqa: E731
            fn.plan = plan
            return ("conv_fprop_tc:" + y.name, fn)
        if self._stem_s2d(n):
            st = self._stem_s2d_setup(n)
            xsrc = x.cast_from if getattr(x, "cast_elided", False) else x.buf
            sdt = RN_BF16 if xsrc.dtype == torch.bfloat16 else RN_F32
            pack = (lambda s, d=d, px=ptr(xsrc), pu=ptr(st["u"]), sdt=sdt:
                    check(lib.rn_stem_s2d_pack(C.byref(d), px, sdt, pu, s), "rn_stem_s2d_pack"))
            run = self._s2d_plan(d, 0, st["u"], st["wp"], y.buf)
            return ("conv_fprop_tc:" + y.name, lambda s, pack=pack, run=run: (pack(s), run(s)))
        if self._is_stem(n):
            st = self._stem_setup(n)
            xsrc = x.cast_from if getattr(x, "cast_elided", False) else x.buf
            sdt = RN_BF16 if xsrc.dtype == torch.bfloat16 else RN_F32
            im2col = (lambda s, a=(ptr(xsrc), ptr(st["cols"]), d.N, d.H, d.W, d.Cin, d.Ho, d.Wo, d.k, d.stride, d.pad, st["kpad"], sdt):
                      check(lib.rn_im2col(*a, s), "rn_im2col"))
            run = self._plan_raw(st["d1"], 0, st["cols"], st["wf"], y.buf)
            return ("conv_fprop_tc:" + y.name, lambda s, im2col=im2col, run=run: (im2col(s), run(s)))
        if hasattr(n, "fuse"):
            f = n.fuse
            rows = lib.rn_convtc_stat_rows(C.byref(d))
            f["partials"] = torch.zeros(rows, d.Cout, 2, dtype=torch.float32, device=self.dev)
            nrows = C.c_int32(0)
            plan = C.c_void_p()
            check(lib.rn_convtc_plan_create_fused(C.byref(d), ptr(x.buf), self._wptr(kv, "Wb"), ptr(f["out"].buf),
                                                  ptr(f["residual"].buf) if f["residual"] is not None else None,
                                                  ptr(f["partials"]), C.byref(nrows), C.byref(plan)),
                  "rn_convtc_plan_create_fused")
            f["nrows"] = nrows.value
            f["plan"] = plan
            self._plans.append(plan)
            if f["residual"] is not None:
                y.buf = None            # conv + shortcut goes straight into the sum: the conv output is never materialised
            fn = lambda s, p=plan: check(lib.rn_convtc_run(p, 0.0, s), "rn_convtc_run")   # noqa: E731
            fn.plan = plan
            return ("conv_fprop_tc:" + y.name, fn)
        if self._conv_impl(n, 0) == "tc":
            return ("conv_fprop_tc:" + y.name, self._tc_plan(n, d, 0, x.buf, y.buf))
        w = self._wptr(kv, "Wb" if self.bf16 else "W")
        return ("conv_fprop:" + y.name, lambda s, d=d, px=ptr(x.buf), py=ptr(y.buf), w=w:
                check(lib.rn_conv2d_fprop(C.byref(d), px, w, py, s), "rn_conv2d_fprop"))

    # folded finalisation -------------------------------------------------------------------
    def _finalize_mode(self):
        """Where the per-CTA sums of a fused convolution epilogue become BN constants (RN_BN_FINALIZE):
        'kernel' — a standalone rn_bn_finalize / rn_bn_bwd_finalize launch (default: the measured winner);
        'apply'  — inside the BN apply kernel that follows (every CTA sums the rows of its channels): 87 fewer launches
                   per ResNet-50 step, but the dependent loads at the head of every apply CTA cost what the launch did;
        'conv'   — by the convolution's last CTA, up to RN_TC_FOLD_MAX_C channels (slower: fence + ticket + one CTA's loads).
        profiles/r02_finalize_ab.txt has the A/B."""
        return os.environ.get("RN_BN_FINALIZE", "kernel")

    def _fold_ok(self, channels):
        return self._finalize_mode() == "conv" and channels <= int(os.environ.get("RN_TC_FOLD_MAX_C", "256"))

    def _ticket(self):
        """One zero-initialised int32 per folded plan (the last-CTA ticket), carved from a shared device array."""
        if getattr(self, "_tickets", None) is None or self._ticket_next == self._tickets.numel():
            self._tickets = torch.zeros(1024, dtype=torch.int32, device=self.dev)
            self._keep.append(self._tickets)
            self._ticket_next = 0
        p = self._tickets.data_ptr() + 4 * self._ticket_next
        self._ticket_next += 1
        return p

    # tensor-core plans -------------------------------------------------------------------
    def _tc_plan(self, n, d, kind, a_buf, out_ptr_or_buf, b_buf=None):
        """Tensor-core plan for one conv GEMM.  fprop and dgrad both read the conv kernel from `Wb`, the
        flat bf16 copy of the HWIO weights that the fused optimizer kernel refreshes every step (fprop takes
        it as an MN-major operand, dgrad as a K-major one): no per-layer re-layout."""
        if kind in (0, 1):
            b = self._wptr(n.vars["kernel"], "Wb")
        else:
            b = ptr(b_buf)
        nbytes = lib.rn_convtc_scratch_bytes(C.byref(d), kind)
        scratch = torch.zeros(max(nbytes, 16), dtype=torch.uint8, device=self.dev)
        self._keep.append(scratch)
        plan = C.c_void_p()
        outp = out_ptr_or_buf if isinstance(out_ptr_or_buf, int) else ptr(out_ptr_or_buf)
        check(lib.rn_convtc_plan_create(C.byref(d), kind, ptr(a_buf), b, outp, ptr(scratch), nbytes,
                                        C.byref(plan)), "rn_convtc_plan_create")
        self._plans.append(plan)
        fn = lambda s, p=plan, beta=0.0: check(lib.rn_convtc_run(p, beta, s), "rn_convtc_run")   # noqa: E731
        fn.plan = plan
        return fn

    def _bn_fwd(self, n):
        x, y = n.inputs[0], n.outputs[0]
        C_ = x.shape[-1]
        M = x.numel // C_
        a = n.attrs
        v = n.vars
        out = []
        if a["is_training"]:
            saved = torch.zeros(2, C_, dtype=torch.float32, device=self.dev)
            n.saved = saved
            mean_p, rstd_p = saved[0].data_ptr(), saved[1].data_ptr()
            add = self._fused_add_for(n)
            if add is not None:
                px, pa, pz = ptr(add.inputs[0].buf), ptr(add.inputs[1].buf), ptr(x.buf)
            else:
                px, pa, pz = ptr(x.buf), None, None
            mm, mv = self._wptr(v["moving_mean"], "S"), self._wptr(v["moving_variance"], "S")
            if not self.train:
                mm = mv = None       # forward-only with batch statistics: leave moving stats alone
            cv = getattr(n, "stats_from", None)
            fold_apply = cv is not None and self._finalize_mode() == "apply" and not getattr(n, "virtual", False)
            if fold_apply:
                pass        # the apply pass below sums the convolution's per-CTA rows itself
            elif cv is not None and self._fold_ok(C_) and not cv.fuse.get("folded"):
                # ... and its last CTA also turns the per-CTA sums into mean / rstd / moving averages: no launch here
                cv.fuse["folded"] = True
                check(lib.rn_convtc_plan_fold_bn_finalize(cv.fuse["plan"], M, a["eps"], a["decay"], int(a["unbiased"]),
                                                          mean_p, rstd_p, mm, mv, self._ticket()),
                      "rn_convtc_plan_fold_bn_finalize")
            elif cv is not None:
                # the producing conv's epilogue already summed the stored tensor (and added the shortcut)
                out.append(("bn_finalize:" + a["scope"],
                            lambda s, args=(ptr(cv.fuse["partials"]), cv.fuse["nrows"], M, C_, a["eps"], a["decay"],
                                            int(a["unbiased"]), mean_p, rstd_p, mm, mv):
                            check(lib.rn_bn_finalize(*args, s), "rn_bn_finalize")))
            else:
                out.append(("bn_stats:" + a["scope"],
                            lambda s, args=(px, pa, pz, M, C_, self.rn_dt, a["eps"], a["decay"], int(a["unbiased"]), mean_p, rstd_p, mm, mv):
                            check(lib.rn_bn_stats_fwd(*args, ptr(self.scratch), self.scratch_bytes, s), "rn_bn_stats_fwd")))
            if getattr(n, "virtual", False):
                y.buf = None         # normalised inside the consuming 1x1 convolutions' shared memory, never stored
            elif fold_apply:
                out.append(("bn_apply:" + a["scope"],
                            lambda s, args=(ptr(x.buf), ptr(y.buf), M, C_, self.rn_dt, self._wptr(v["gamma"]), self._wptr(v["beta"]),
                                            ptr(cv.fuse["partials"]), cv.fuse["nrows"], a["eps"], a["decay"], int(a["unbiased"]),
                                            mean_p, rstd_p, mm, mv, a["leak"]):
                            check(lib.rn_bn_apply_act_fwd_partials(*args, s), "rn_bn_apply_act_fwd_partials")))
            else:
                out.append(("bn_apply:" + a["scope"],
                            lambda s, args=(ptr(x.buf), ptr(y.buf), M, C_, self.rn_dt, self._wptr(v["gamma"]), self._wptr(v["beta"]), mean_p, rstd_p, 0, a["eps"], a["leak"]):
                            check(lib.rn_bn_apply_act_fwd(*args, s), "rn_bn_apply_act_fwd")))
        else:
            out.append(("bn_infer:" + a["scope"],
                        lambda s, args=(ptr(x.buf), ptr(y.buf), M, C_, self.rn_dt, self._wptr(v["gamma"]), self._wptr(v["beta"]),
                                        self._wptr(v["moving_mean"], "S"), self._wptr(v["moving_variance"], "S"), 1, a["eps"], a["leak"]):
                        check(lib.rn_bn_apply_act_fwd(*args, s), "rn_bn_apply_act_fwd")))
        return out

    # ------------------------------------------------------------------ backward
    def _slot_of(self, t):
        return getattr(t, "gradslot", None)

    def _tdtype(self, t):
        return self.adt if t.dtype == G.ACT else torch.float32

    def _new_slot(self, t):
        s = _Slot(torch.zeros(t.shape, dtype=self._tdtype(t), device=self.dev))
        s.members.append(t)
        t.gradslot = s
        return s

    def _safe_inplace(self, t, slot, done):
        return all((m is t) or (m.producer is None) or (id(m.producer) in done) for m in slot.members)

    def _target(self, t, done, bwd):
        """Where a backward kernel must put d(loss)/d(t): returns (buffer, beta)."""
        s = self._slot_of(t)
        if s is None:
            s = self._new_slot(t)
        if not s.written:
            s.written = True
            return s.buf, 0.0
        if not self._safe_inplace(t, s, done):
            # another reader of the shared buffer has not run yet: give t a private copy
            ns = _Slot(torch.zeros(t.shape, dtype=self._tdtype(t), device=self.dev))
            ns.members.append(t)
            ns.written = True
            dt = RN_BF16 if ns.buf.dtype == torch.bfloat16 else RN_F32
            bwd.append(("grad_copy:" + t.name, lambda st, a=ptr(s.buf), b=ptr(ns.buf), k=t.numel, dt=dt:
                        check(lib.rn_cast(a, b, k, dt, dt, st), "rn_cast")))
            s.members.remove(t)
            t.gradslot = ns
            s = ns
        return s.buf, 1.0

    def _route(self, t, src_slot, done, bwd):
        """d(t) += d(src): alias when t has no gradient yet, else accumulate."""
        if not t.requires_grad:
            return
        s = self._slot_of(t)
        if s is None:
            t.gradslot = src_slot
            src_slot.members.append(t)
            return
        buf, beta = self._target(t, done, bwd)
        dt = RN_BF16 if buf.dtype == torch.bfloat16 else RN_F32
        if beta == 0.0:
            bwd.append(("grad_copy:" + t.name, lambda st, a=ptr(src_slot.buf), b=ptr(buf), k=t.numel, dt=dt:
                        check(lib.rn_cast(a, b, k, dt, dt, st), "rn_cast")))
        else:
            bwd.append(("grad_add:" + t.name, lambda st, a=ptr(buf), b=ptr(src_slot.buf), k=t.numel, dt=dt:
                        check(lib.rn_add_inplace(a, b, k, dt, st), "rn_add_inplace")))

    def _build_backward(self):
        bwd = []
        done = set()
        self.grad_ready_after = {}   # variable name -> index in bwd after which its gradient is final
        for n in reversed(self.g.nodes):
            self._node_backward(n, bwd, done)
            done.add(id(n))
            for v in n.vars.values():
                if v.trainable:
                    self.grad_ready_after[v.name] = len(bwd)
        self.bwd = bwd

    def _emit(self, bwd, name, fn, **meta):
        bwd.append((name, fn))
        self.meta[name] = meta

    def _node_backward(self, n, bwd, done):
        if n.op == "softmax_xent":
            logits = n.inputs[0]
            if n.dlogits is not None:
                s = _Slot(n.dlogits)
                s.written = True
                s.members.append(logits)
                logits.gradslot = s
            return
        y = n.outputs[0]
        ys = self._slot_of(y)
        if ys is None or not ys.written:
            if any(v.trainable for v in n.vars.values()):
                raise RuntimeError("node %s/%s has trainable variables but no gradient reaches it" % (n.op, y.name))
            return
        dy = ys.buf
        if n.op == "dense":
            x = n.inputs[0]
            B = x.shape[0]
            D = x.numel // B
            K = y.shape[1]
            dxp = None
            if x.requires_grad:
                dxb, beta = self._target(x, done, bwd)
                if beta != 0.0:
                    raise NotImplementedError("dense input gradient fan-in")
                dxp = ptr(dxb)
            # input gradient (and the cheap bias gradient) on the main stream; the weight gradient is only needed
            # by the update, so it is forked to the side stream together with the conv weight gradients
            nbb = lib.rn_dense_bwd_scratch_bytes(B, D, K)
            bsc = torch.zeros(max(nbb, 16), dtype=torch.uint8, device=self.dev)
            self._keep.append(bsc)
            self._emit(bwd, "dense_bwd:" + y.name,
                       lambda s, args=(ptr(x.buf), self._wptr(n.vars["kernel"]), ptr(dy), None,
                                       self._wptr(n.vars["bias"], "G"), dxp, B, D, K, ptr(bsc), nbb):
                       check(lib.rn_dense_bwd(*args, s), "rn_dense_bwd"),
                       node=n, dy=dy, dx=dxb if x.requires_grad else None, beta=0.0)
            self._emit(bwd, "dense_wgrad:" + y.name,
                       lambda s, args=(ptr(x.buf), self._wptr(n.vars["kernel"]), ptr(dy), self._wptr(n.vars["kernel"], "G"),
                                       None, None, B, D, K, None, 0):
                       check(lib.rn_dense_bwd(*args, s), "rn_dense_bwd"),
                       node=n, dy=dy)
        elif n.op == "avgpool":
            x = n.inputs[0]
            if not x.requires_grad:
                return
            N, H, W, Cc = x.shape
            dxb, beta = self._target(x, done, bwd)
            idt = RN_BF16 if x.buf.dtype == torch.bfloat16 else RN_F32
            odt = RN_BF16 if y.buf.dtype == torch.bfloat16 else RN_F32
            self._emit(bwd, "avgpool_bwd:" + y.name,
                       lambda s, args=(ptr(dy), ptr(dxb), N, H, W, Cc, n.attrs["k"], n.attrs["s"], y.shape[3], n.attrs["c_before"], idt, odt, beta):
                       check(lib.rn_avgpool_bwd(*args, s), "rn_avgpool_bwd"), node=n, dy=dy, dx=dxb, beta=beta)
        elif n.op == "maxpool":
            x = n.inputs[0]
            if not x.requires_grad:
                return
            N, H, W, Cc = x.shape
            dxb, beta = self._target(x, done, bwd)
            self._emit(bwd, "maxpool_bwd:" + y.name,
                       lambda s, args=(ptr(dy), ptr(n.argmax), ptr(dxb), N, H, W, Cc, y.shape[1], y.shape[2], n.attrs["pad"], self.rn_dt, beta):
                       check(lib.rn_maxpool3x3s2_bwd(*args, s), "rn_maxpool3x3s2_bwd"), node=n, dy=dy, dx=dxb, beta=beta)
        elif n.op == "add":
            for t in n.inputs:
                self._route(t, ys, done, bwd)
        elif n.op == "bn_act":
            x = n.inputs[0]
            a, v = n.attrs, n.vars
            if not a["is_training"]:
                raise NotImplementedError("backward through inference-mode batch norm")
            C_ = x.shape[-1]
            M = x.numel // C_
            mean_p, rstd_p = n.saved[0].data_ptr(), n.saved[1].data_ptr()
            gp, bp = self._wptr(v["gamma"]), self._wptr(v["beta"])
            dgp, dbp = self._wptr(v["gamma"], "G"), self._wptr(v["beta"], "G")
            bf = getattr(n, "bwd_from", None)
            fold_apply = bf is not None and self._finalize_mode() == "apply" and x.requires_grad
            if fold_apply:
                pass        # bn_bwd_apply below sums the dgrad's per-CTA rows itself and writes dgamma / dbeta
            elif bf is not None and self._fold_ok(C_):
                check(lib.rn_convtc_plan_fold_bn_bwd_finalize(bf["plan"], dgp, dbp, self._ticket()),
                      "rn_convtc_plan_fold_bn_bwd_finalize")
            elif bf is not None:
                self._emit(bwd, "bn_bwd_finalize:" + a["scope"],
                           lambda s, args=(ptr(bf["partials"]), bf["nrows"], C_, dgp, dbp):
                           check(lib.rn_bn_bwd_finalize(*args, s), "rn_bn_bwd_finalize"), node=n, dy=dy)
            else:
                self._emit(bwd, "bn_bwd_reduce:" + a["scope"],
                           lambda s, args=(ptr(dy), ptr(x.buf), M, C_, self.rn_dt, gp, bp, mean_p, rstd_p, a["leak"], dgp, dbp):
                           check(lib.rn_bn_act_bwd_reduce(*args, ptr(self.scratch), self.scratch_bytes, s), "rn_bn_act_bwd_reduce"),
                           node=n, dy=dy)
            if x.requires_grad:
                dxb, beta = self._target(x, done, bwd)
                addp = ptr(dxb) if beta != 0.0 else None
                if fold_apply:
                    self._emit(bwd, "bn_bwd_apply:" + a["scope"],
                               lambda s, args=(ptr(dy), ptr(x.buf), ptr(dxb), addp, M, C_, self.rn_dt, gp, bp, mean_p, rstd_p,
                                               ptr(bf["partials"]), bf["nrows"], dgp, dbp, a["leak"]):
                               check(lib.rn_bn_act_bwd_apply_partials(*args, s), "rn_bn_act_bwd_apply_partials"),
                               node=n, dy=dy, dx=dxb, beta=beta)
                else:
                    self._emit(bwd, "bn_bwd_apply:" + a["scope"],
                               lambda s, args=(ptr(dy), ptr(x.buf), ptr(dxb), addp, M, C_, self.rn_dt, gp, bp, mean_p, rstd_p, dgp, dbp, a["leak"]):
                               check(lib.rn_bn_act_bwd_apply(*args, s), "rn_bn_act_bwd_apply"), node=n, dy=dy, dx=dxb, beta=beta)
        elif n.op == "conv2d":
            x = n.inputs[0]
            d = self._desc(n)
            self._keep.append(d)
            kv = n.vars["kernel"]
            if x.requires_grad:
                dxb, beta = self._target(x, done, bwd)
                bn = x.producer
                if (beta == 0.0 and self.bf16 and self._conv_impl(n, 1) == "tc" and bn is not None and bn.op == "bn_act"
                        and bn.attrs["is_training"] and len(x.consumers) == 1 and hasattr(bn, "saved")
                        and not os.environ.get("RN_NO_BNBWD_FUSION")
                        and lib.rn_convtc_fused_dgrad_supported(C.byref(d))):
                    # this dgrad produces the complete incoming gradient of a BN+ReLU: its epilogue also accumulates
                    # that BN's backward reductions, so the BN needs no pass over (dy, x) before its dx kernel
                    rows = lib.rn_convtc_stat_rows(C.byref(d))
                    part = torch.zeros(rows, d.Cin, 2, dtype=torch.float32, device=self.dev)
                    nrows = C.c_int32(0)
                    plan = C.c_void_p()
                    bv = bn.vars
                    check(lib.rn_convtc_plan_create_fused_dgrad(
                        C.byref(d), ptr(dy), self._wptr(kv, "Wb"), ptr(dxb), ptr(bn.inputs[0].buf),
                        self._wptr(bv["gamma"]), self._wptr(bv["beta"]), bn.saved[0].data_ptr(), bn.saved[1].data_ptr(),
                        bn.attrs["leak"], ptr(part), C.byref(nrows), C.byref(plan)), "rn_convtc_plan_create_fused_dgrad")
                    self._plans.append(plan)
                    bn.bwd_from = dict(partials=part, nrows=nrows.value, plan=plan)
                    fn = lambda s, p=plan: check(lib.rn_convtc_run(p, 0.0, s), "rn_convtc_run")   # noqa: E731
                    fn.plan = plan
                    self._emit(bwd, "conv_dgrad_tc:" + y.name, fn, node=n, dy=dy, dx=dxb, beta=0.0)
                elif self._conv_impl(n, 1) == "tc":
                    run = self._tc_plan(n, d, 1, dy, dxb)
                    fn = lambda s, run=run, beta=beta: run(s, beta=beta)   # noqa: E731
                    fn.plan = run.plan
                    self._emit(bwd, "conv_dgrad_tc:" + y.name, fn, node=n, dy=dy, dx=dxb, beta=beta)
                else:
                    w = self._wptr(kv, "Wb" if self.bf16 else "W")
                    self._emit(bwd, "conv_dgrad:" + y.name, lambda s, d=d, pdy=ptr(dy), w=w, pdx=ptr(dxb), beta=beta:
                               check(lib.rn_conv2d_dgrad(C.byref(d), pdy, w, pdx, beta, s), "rn_conv2d_dgrad"),
                               node=n, dy=dy, dx=dxb, beta=beta)
            if self._stem_s2d(n):
                st = self._stem_s2d_setup(n)
                tmp = torch.zeros(st["taps"] * 64, d.Cout, dtype=torch.float32, device=self.dev)
                self._keep.append(tmp)
                run = self._s2d_plan(d, 2, st["u"], dy, tmp)
                up = (lambda s, d=d, a=ptr(tmp), b=self._wptr(kv, "G"):
                      check(lib.rn_stem_s2d_wgrad_unpack(C.byref(d), a, b, s), "rn_stem_s2d_wgrad_unpack"))
                self._emit(bwd, "conv_wgrad_tc:" + y.name, lambda s, run=run, up=up: (run(s), up(s)), node=n, dy=dy)
            elif self._is_stem(n) and lib.rn_convtc_supported(C.byref(self._stem_setup(n)["d1"]), 2):
                st = self._stem_setup(n)
                tmp = torch.zeros(st["kpad"], d.Cout, dtype=torch.float32, device=self.dev)
                self._keep.append(tmp)
                run = self._plan_raw(st["d1"], 2, st["cols"], dy, tmp)
                cp = (lambda s, a=ptr(tmp), b=self._wptr(kv, "G"), k=st["rows"] * d.Cout:
                      check(lib.rn_cast(a, b, k, RN_F32, RN_F32, s), "rn_cast"))
                self._emit(bwd, "conv_wgrad_tc:" + y.name, lambda s, run=run, cp=cp: (run(s), cp(s)), node=n, dy=dy)
            elif self._bn_operand(x)[1] is not None:
                xbuf, bn_in = self._bn_operand(x)
                nbytes = lib.rn_convtc_scratch_bytes(C.byref(d), 2)
                scratch = torch.zeros(max(nbytes, 16), dtype=torch.uint8, device=self.dev)
                self._keep.append(scratch)
                plan = C.c_void_p()
                check(lib.rn_convtc_plan_create_bnin(C.byref(d), 2, ptr(xbuf), *self._bnin_args(bn_in), ptr(dy),
                                                     self._wptr(kv, "G"), None, None, None, ptr(scratch), nbytes,
                                                     C.byref(plan)), "rn_convtc_plan_create_bnin")
                self._plans.append(plan)
                run = lambda s, p=plan: check(lib.rn_convtc_run(p, 0.0, s), "rn_convtc_run")   # noqa: E731
                run.plan = plan
                self._emit(bwd, "conv_wgrad_tc:" + y.name, run, node=n, dy=dy)
            elif self._conv_impl(n, 2) == "tc":
                run = self._tc_plan(n, d, 2, x.buf, self._wptr(kv, "G"), b_buf=dy)
                self._emit(bwd, "conv_wgrad_tc:" + y.name, run, node=n, dy=dy)
            else:
                nb = lib.rn_conv2d_wgrad_scratch_bytes(C.byref(d))
                if not hasattr(self, "_wg_scratch") or self._wg_scratch.numel() < nb:
                    self._wg_scratch = torch.zeros(max(nb, 16), dtype=torch.uint8, device=self.dev)
                    self._wg_rebind = True
                self._emit(bwd, "conv_wgrad:" + y.name, ("wgrad", d, ptr(x.buf), ptr(dy), self._wptr(kv, "G")), node=n, dy=dy)
        else:
            raise ValueError("unknown op " + n.op)

    def _finalize_wgrad(self):
        """SIMT wgrad launches share one split-K scratch buffer whose final size is only known
        after the whole backward list is built."""
        out = []
        sp, sb = ptr(self._wg_scratch) if hasattr(self, "_wg_scratch") else None, \
            self._wg_scratch.numel() if hasattr(self, "_wg_scratch") else 0
        for name, fn in self.bwd:
            if isinstance(fn, tuple) and fn[0] == "wgrad":
                _, d, px, pdy, pdw = fn
                fn = (lambda s, d=d, px=px, pdy=pdy, pdw=pdw:
                      check(lib.rn_conv2d_wgrad(C.byref(d), px, pdy, pdw, sp, sb, s), "rn_conv2d_wgrad"))
            out.append((name, fn))
        self.bwd = out

    # ------------------------------------------------------------------ aggregation + optimizer
    def _build_update(self, bucket_bytes):
        self._finalize_wgrad()
        # buckets: contiguous ranges of the flat gradient buffer, filled from the END (gradients
        # become final in reverse creation order), each reduced as soon as it is complete.
        tv = self.g.trainable_variables()
        self.buckets = []
        if self.world > 1 and tv:
            from .parallel import plan_buckets
            self.buckets = plan_buckets([v.offset for v in tv], self.n_flat,
                                        [self.grad_ready_after[v.name] for v in tv],
                                        max(bucket_bytes // 4, GRANULE), len(self.bwd))
            self.comm_stream = torch.cuda.Stream(device=self.dev)
        use_mom = 1 if self.opt_kind == "mom" else 0
        inv_R = 1.0 / float(self.world)
        self.update = [("sgd_momentum_fused",
                        lambda s: check(lib.rn_sgd_momentum_fused(
                            ptr(self.W), ptr(self.Gd), ptr(self.Mom), ptr(self.Wb), self.n_flat, ptr(self.mask),
                            ptr(self.hyper), inv_R, self.wd, self.momentum, use_mom, ptr(self.sumsq),
                            ptr(self.scratch), self.scratch_bytes, s), "rn_sgd_momentum_fused"))]
        for i, fn in enumerate(getattr(self, "_relayout", [])):     # stem only: copy of the kernel into its K-padded operand
            self.update.append(("stem_weights:%d" % i, fn))

    def set_lr(self, lr):
        """The reference feeds the learning rate every step (resnet_cifar_train.py:297-299); here
        it lives in device memory and is re-uploaded when the value changes: one async H2D copy from
        a ring of pinned staging words, so a schedule that changes the rate on every step (the
        ImageNet warm-up, resnet_imagenet_train.py:236-260) never stalls the host on the stream."""
        lr = float(lr)
        if getattr(self, "lrn_rate", None) == lr:
            return
        self.lrn_rate = lr
        ring = getattr(self, "_hyper_ring", None)
        if ring is None:
            ring = self._hyper_ring = [[torch.zeros(4, dtype=torch.float32).pin_memory(), None] for _ in range(32)]
            self._hyper_slot = 0
        slot = ring[self._hyper_slot]
        self._hyper_slot = (self._hyper_slot + 1) % len(ring)
        if slot[1] is not None:
            slot[1].synchronize()           # only blocks if the GPU is 32 rate changes behind the host
        slot[0][0] = lr
        st = torch.cuda.current_stream()
        self.hyper.copy_(slot[0], non_blocking=True)
        slot[1] = torch.cuda.Event()
        slot[1].record(st)

    def broadcast_from_rank0(self):
        """hvd.BroadcastGlobalVariablesHook(0) (resnet_cifar_train_horovod.py:313): weights, BN
        moving stats and momentum slots from rank 0 (one flat ncclBroadcast each)."""
        import torch.distributed as dist
        for buf in (self.W, self.S) + ((self.Mom,) if self.train else ()):
            dist.broadcast(buf, src=dist.get_global_rank(self.pg, 0) if self.pg is not None else 0, group=self.pg)
        self.refresh_shadow_weights()

    # ------------------------------------------------------------------ running
    def _run_list(self, lst, st):
        for _, fn in lst:
            fn(st)

    def _run_overlapped(self, lst, cur):
        """Run a launch list on stream `cur`, with every weight-gradient kernel forked onto a side stream:
        a conv layer's wgrad and dgrad are independent, so the wgrad GEMMs (tensor-bound, one CTA per SM)
        run concurrently with the batch-norm backward kernels (HBM-bound) that follow the dgrad.  A wgrad
        reads dy; the main stream therefore waits for it before any kernel that overwrites that buffer
        (the in-place residual-stream accumulation) and at the end of the list."""
        if not self.overlap_wgrad:
            return self._run_list(lst, cur.cuda_stream)
        side = self.wg_stream
        pending = []
        for name, fn in lst:
            m = self.meta.get(name)
            if name.startswith(("conv_wgrad", "dense_wgrad")):
                ev = torch.cuda.Event()
                ev.record(cur)
                side.wait_event(ev)
                fn(side.cuda_stream)
                done = torch.cuda.Event()
                done.record(side)
                pending.append((m["dy"].data_ptr(), done))
                continue
            if pending:
                dx = m.get("dx") if m else None
                if m is None or name.startswith("grad_"):
                    for _, e in pending:
                        cur.wait_event(e)
                    pending = []
                elif dx is not None:
                    p = dx.data_ptr()
                    keep = []
                    for q, e in pending:
                        if q == p:
                            cur.wait_event(e)
                        else:
                            keep.append((q, e))
                    pending = keep
            fn(cur.cuda_stream)
        for _, e in pending:
            cur.wait_event(e)

    def forward(self):
        st = torch.cuda.current_stream().cuda_stream
        self._run_list(self.fwd, st)

    def _step_launches(self):
        import torch.distributed as dist
        cur = torch.cuda.current_stream()
        st = cur.cuda_stream
        for lst, ready in self._segments():
            self._run_overlapped(lst, cur)
            for bi in ready:
                self._reduce_bucket(bi, cur, dist)
        if self.world > 1:
            cur.wait_stream(self.comm_stream)
        self._run_list(self.update, st)

    def _reduce_bucket(self, bi, cur, dist):
        lo, hi, _ = self.buckets[bi]
        self.comm_stream.wait_stream(cur)
        with torch.cuda.stream(self.comm_stream):
            dist.all_reduce(self.Gd[lo:hi], op=dist.ReduceOp.SUM, group=self.pg)

    def count_launches(self):
        """Kernel launches of one step, counted by the library itself during an un-captured run."""
        if not self.launches_per_step:
            snap = (self.W.clone(), self.S.clone()) + ((self.Mom.clone(),) if self.train else ())
            l0 = _lib.launch_count()
            st = torch.cuda.current_stream().cuda_stream
            self._run_list(self.fwd, st)
            if self.train:
                self._run_list(self.bwd, st)
                self._run_list(self.update, st)
            torch.cuda.synchronize(self.dev)
            self.launches_per_step = _lib.launch_count() - l0
            self.W.copy_(snap[0]); self.S.copy_(snap[1])
            if self.train:
                self.Mom.copy_(snap[2])
            self.refresh_shadow_weights()
        return self.launches_per_step

    def _segments(self):
        """Launch list of one step cut where gradient buckets become complete:
        [(launches, [bucket indices to all-reduce afterwards])...], then the update list."""
        segs = []
        if self.world > 1 and self.buckets:
            start = 0
            bi = 0
            # buckets are ordered by readiness (monotone `ready`)
            cuts = sorted(set(b[2] for b in self.buckets))
            prev = 0
            for c in cuts:
                lst = (self.fwd if prev == 0 else []) + self.bwd[prev:c]
                ready = [i for i, b in enumerate(self.buckets) if b[2] == c]
                segs.append((lst, ready))
                prev = c
            if prev < len(self.bwd):
                segs.append((self.bwd[prev:], []))
        else:
            segs.append((self.fwd + (self.bwd if self.train else []), []))
        return segs

    def capture(self):
        """Capture the step in CUDA graphs.  Single GPU: one graph for the whole step.  Multi-GPU: one
        graph per backward segment between gradient-bucket boundaries plus one for the update; the NCCL
        all-reduces are issued eagerly on the side stream between the segment launches (ordering by
        stream events), so communication still overlaps the remaining backward segments."""
        if self._graph_exec is not None:
            return
        torch.cuda.synchronize(self.dev)
        side = torch.cuda.Stream(device=self.dev)
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):      # warm-up outside capture (lazy module loading, NCCL set-up)
            if self.train:
                snap = (self.W.clone(), self.Mom.clone(), self.S.clone(), None if self.Wb is None else self.Wb.clone())
                self._step_launches()
                torch.cuda.synchronize(self.dev)
                self.W.copy_(snap[0]); self.Mom.copy_(snap[1]); self.S.copy_(snap[2])
                if self.Wb is not None:
                    self.Wb.copy_(snap[3])
                for fn in getattr(self, "_relayout", []):
                    fn(side.cuda_stream)
            else:
                self.forward()
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize(self.dev)
        graphs = []
        pool = None
        lists = [(lst, ready) for lst, ready in self._segments()] if self.train else [(self.fwd, [])]
        if self.train:
            lists.append((self.update, None))
        # No garbage collection while a stream is capturing: if the cyclic collector happens to free an OLDER model's
        # CUDAGraph objects in the middle of this capture, their cudaGraphExecDestroy is "not permitted when stream is
        # capturing" and invalidates it (torch.cuda.graph no longer collects before capturing by itself).
        import gc
        gc.collect()
        gc_was_on = gc.isenabled()
        gc.disable()
        try:
            for lst, ready in lists:
                gexec = torch.cuda.CUDAGraph()
                with torch.cuda.graph(gexec, pool=pool):
                    if ready is None:
                        self._run_list(lst, torch.cuda.current_stream().cuda_stream)
                    else:
                        self._run_overlapped(lst, torch.cuda.current_stream())
                pool = gexec.pool()
                graphs.append((gexec, ready))
        finally:
            if gc_was_on:
                gc.enable()
        self._graph_exec = graphs

    def _replay(self):
        import torch.distributed as dist
        cur = torch.cuda.current_stream()
        for gexec, ready in self._graph_exec:
            if ready is None:                      # the update graph waits for every all-reduce
                if self.world > 1:
                    cur.wait_stream(self.comm_stream)
                gexec.replay()
                continue
            gexec.replay()
            for bi in ready:
                self._reduce_bucket(bi, cur, dist)

    def step(self):
        """One synchronous training step (== sess.run(train_op), resnet_cifar_train.py:343-344)."""
        if not self.train:
            raise RuntimeError("executor was built with train=False")
        if self._graph_exec is not None:
            self._replay()
        else:
            self._step_launches()
        self.global_step += 1

    def run_forward(self):
        if self._graph_exec is not None and not self.train:
            self._replay()
        else:
            self.forward()

    # ------------------------------------------------------------------ results
    def scalars(self):
        """(xent, precision, cost) of the last step — one small D2H read (synchronises)."""
        if self.loss is None:
            raise RuntimeError("graph has no loss")
        xs = self.loss[1].buf.tolist()
        cost = xs[0] + self.wd * 0.5 * float(self.sumsq.item()) if self.train else xs[0]
        return xs[0], xs[1], cost

    def describe_plans(self):
        """launch name -> one-line description of its tensor-core plan (tile, pipeline depth, CTA pair, fusions)."""
        out = OrderedDict()
        for name, fn in self.fwd + (self.bwd if self.train else []):
            plan = getattr(fn, "plan", None)
            if plan is not None:
                out[name] = _lib.describe_plan(plan)
        return out

    def close(self):
        for p in self._plans:
            lib.rn_convtc_plan_destroy(p)
        self._plans = []
