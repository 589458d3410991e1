"""Debug helper: per-tensor forward/backward errors of the product path vs the oracle."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch
from util import rel_err
from distributed_tensorflow_resnet_b200 import resnet_model, FLAGS
from oracle import resnet_oracle as ro
import test_model_gpu as T

kind_sel, mode = sys.argv[1], sys.argv[2]
case = [c for c in T.CASES if c[0] == kind_sel][0]
kind, flags, shape, classes, legacy = case
g = torch.Generator().manual_seed(11)
x = torch.randn(*shape, generator=g)
y = torch.eye(classes)[torch.arange(shape[0]) % classes]
hkw = dict(num_classes=classes, lrn_rate=0.1, weight_decay_rate=2e-4, optimizer='mom')
if legacy:
    hkw.update(batch_size=shape[0], min_lrn_rate=1e-4, **legacy)
hps = resnet_model.HParams(**hkw)
model = T._mk_model(hps, x, y, compute_dtype=mode, seed=3, **flags)
model.build_graph()
ex = model.executor
params = {k: v for k, v in ex.state_dict().items() if not k.endswith("/Momentum") and k != "global_step"}
net = T._oracle_for(kind, hps, params, storage=torch.bfloat16 if mode == "bf16" else None, **flags)
out = ro.loss_and_grads(net, x.double(), y.double(), 2e-4, True)
ex.forward(); torch.cuda.synchronize()
print("== after forward only")
for name, t in model.graph.tensors.items():
    if name in net.acts and name not in ("X", "Y"):
        print("  %-45s %.3e" % (name, rel_err(t.buf.float(), net.acts[name].val)))
model.train_op(); torch.cuda.synchronize()
print("== after full step")
for name, t in model.graph.tensors.items():
    if name in net.acts and name not in ("X", "Y"):
        e = rel_err(t.buf.float(), net.acts[name].val)
        slot = getattr(t, "gradslot", None)
        ge = None
        if slot is not None and net.acts[name].grad is not None and (len(slot.members) == 1 or slot.members[-1] is t):
            ge = rel_err(slot.buf.float(), net.acts[name].grad)
        print("  %-45s act %.3e  grad %s" % (name, e, "%.3e" % ge if ge is not None else "-"))
for v in model.graph.trainable_variables():
    print("  G %-45s %.3e" % (v.name, rel_err(ex.var_view(v.name, "G"), net.param_grads[v.name])))

# ---- drill into the last BN backward
bn = [n for n in model.graph.nodes if n.op == "bn_act"][-1]
xz = bn.inputs[0].buf.double().cpu().reshape(-1, bn.inputs[0].shape[-1])
yslot = bn.outputs[0].gradslot.buf.double().cpu().reshape(xz.shape)
mean, rstd = bn.saved[0].double().cpu(), bn.saved[1].double().cpu()
gam = ex.var_view(bn.vars["gamma"].name).double().cpu(); bet = ex.var_view(bn.vars["beta"].name).double().cpu()
a = gam * rstd; b = bet - mean * a
pre_k = xz * a + b
oname = bn.outputs[0].name
pre_o = net.acts[[k for k in net.acts if k.endswith("FusedBatchNorm") or k.endswith("batchnorm")][-1]].val.reshape(xz.shape)
print("mask disagreements kernel-formula vs oracle:", int(((pre_k < 0) != (pre_o < 0)).sum()), "of", pre_k.numel())
print("min |pre| oracle:", float(pre_o.abs().min()), " #|pre|<1e-5:", int((pre_o.abs() < 1e-5).sum()))
leak = bn.attrs["leak"]
g_k = torch.where(pre_k < 0, leak * yslot, yslot)
db_k = g_k.sum(0)
db_dev = ex.var_view(bn.vars["beta"].name, "G").double().cpu()
db_o = net.param_grads[bn.vars["beta"].name]
print("dbeta: kernel vs own-formula fp64 %.3e ; kernel vs oracle %.3e ; own-formula vs oracle %.3e" %
      (rel_err(db_dev, db_k), rel_err(db_dev, db_o), rel_err(db_k, db_o)))
print("max|dbeta| %.3e  max|g| %.3e  sum|g| max over channels %.3e" % (float(db_o.abs().max()), float(g_k.abs().max()), float(g_k.abs().sum(0).max())))
var = xz.var(0, unbiased=False)
print("channels with tiny variance (<1e-10):", int((var < 1e-10).sum()), " min var %.3e" % float(var.min()))
