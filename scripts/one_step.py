"""Run a few UN-CAPTURED training steps of a bench workload (for ncu: every kernel is an ordinary launch).
usage: python scripts/one_step.py [workload=imagenet50] [steps=3] [batch]
Prints the number of kernel launches per step (the -s / -c arguments for ncu)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from bench import WORKLOADS
from distributed_tensorflow_resnet_b200 import resnet_model, FLAGS, _lib
wl_name = sys.argv[1] if len(sys.argv) > 1 else "imagenet50"
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
wl = WORKLOADS[wl_name]
B = int(sys.argv[3]) if len(sys.argv) > 3 else wl["batch"]
FLAGS.reset()
FLAGS.update(dataset=wl["dataset"], compute_dtype="bf16", use_cuda_graph=False, resnet_size=50, seed=0)
g = torch.Generator().manual_seed(1000)
images = torch.randn(B, wl["size"], wl["size"], 3, generator=g).cuda()
labels = torch.eye(wl["classes"])[torch.randint(0, wl["classes"], (B,), generator=g)].cuda()
hkw = dict(num_classes=wl["classes"], lrn_rate=0.1, weight_decay_rate=wl["wd"], optimizer='mom')
if wl.get("upstream"):
    hkw.update(batch_size=B, min_lrn_rate=1e-4, **wl["upstream"])
model = resnet_model.ResNet(resnet_model.HParams(**hkw), images, labels, 'train')
if wl.get("filters"):
    model.filters = wl["filters"]
model.build_graph()
for _ in range(max(steps - 1, 0)):      # warm-up steps stay outside the profiled range
    model.train_op()
torch.cuda.synchronize()
torch.cuda.profiler.start()            # ncu --profile-from-start off: only the launches of the LAST step (+ one optimizer
l0 = _lib.launch_count()               # launch before it, see below) are profiled
ex = model.executor
ex._run_list(ex.update[:1], torch.cuda.current_stream().cuda_stream)      # marks the start of the step for the summariser
model.train_op()
torch.cuda.synchronize()
torch.cuda.profiler.stop()
steps = 1
print("launches per step (library count): %d; loss %.4f" % ((_lib.launch_count() - l0) // steps, model.cost))
