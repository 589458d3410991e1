"""Two-GPU synchronous data-parallel step (NCCL): SyncReplicasOptimizer / Horovod semantics —
every rank ends the step with identical weights equal to one momentum step on the MEAN of the R
per-replica gradients (oracle: Trainer.step over both shards).  Needs >= 2 GPUs (skipped otherwise;
the world_size-2 plumbing is also covered on CPU over gloo in test_host.py)."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    """A TCP port nobody is listening on right now (a fixed pid-derived port collides when two torchrun
    rendezvous of one pytest process follow each other closely)."""
    import socket
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]

_WORKER = r'''
import os, sys, numpy as np, torch, torch.distributed as dist
sys.path.insert(0, sys.argv[1]); sys.path.insert(0, os.path.join(sys.argv[1], "tests"))
from distributed_tensorflow_resnet_b200 import resnet_model, FLAGS
from distributed_tensorflow_resnet_b200.resnet_cifar_train import init_distributed
from oracle import resnet_oracle as ro
rank, world = init_distributed()
use_graph = sys.argv[2] == "graph"
FLAGS.reset()
FLAGS.update(compute_dtype="f32", use_cuda_graph=use_graph, dataset="cifar10", resnet_size=8, seed=3,
             job_name="worker", sync_replicas=True, replicas_to_aggregate=world, task_index=rank)
g = torch.Generator().manual_seed(5)
B = 8
x = torch.randn(world * B, 32, 32, 3, generator=g); y = torch.eye(10)[torch.arange(world * B) % 10]
X = x[rank * B:(rank + 1) * B].cuda().contiguous(); Y = y[rank * B:(rank + 1) * B].cuda().contiguous()
hps = resnet_model.HParams(num_classes=10, lrn_rate=0.1, weight_decay_rate=2e-4, optimizer='mom')
m = resnet_model.ResNet(hps, X, Y, 'train'); m.build_graph()
ex = m.executor
assert ex.world == world and len(ex.buckets) >= 1
params = {k: v for k, v in ex.state_dict().items() if not k.endswith("/Momentum") and k != "global_step"}
w0 = ex.W.clone()
gathered = [torch.empty_like(w0) for _ in range(world)]
dist.all_gather(gathered, w0)
assert all(torch.equal(gathered[0], t) for t in gathered), "initial broadcast from rank 0 failed"
for _ in range(2):
    m.train_op()
torch.cuda.synchronize()
dist.all_gather(gathered, ex.W)
assert all(torch.equal(gathered[0], t) for t in gathered), "ranks diverged"
if rank == 0:
    net = ro.OfficialResNetV2("cifar10", 8, 10, dtype=torch.float64, params=params)
    tr = ro.Trainer(net, 2e-4, "mom")
    for _ in range(2):
        tr.step([(x[r * B:(r + 1) * B].double(), y[r * B:(r + 1) * B].double()) for r in range(world)], 0.1)
    worst = 0.0
    for v in m.graph.trainable_variables():
        got = ex.var_view(v.name).double().cpu(); want = net.params[v.name]
        worst = max(worst, float((got - want).abs().max() / (want.abs().max() + 1e-30)))
    print("worst relative weight error after 2 steps:", worst)
    assert worst < 1e-4
    print("DDP_OK")
dist.barrier(); dist.destroy_process_group()
'''


@pytest.mark.parametrize("mode", ["eager", "graph"])
def test_two_gpu_step_matches_oracle_mean_of_replicas(tmp_path, mode):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    script = tmp_path / "w.py"
    script.write_text(_WORKER)
    port = _free_port()
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
           "--master-addr", "127.0.0.1", "--master-port", str(port), str(script), ROOT, mode]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=150)
    assert r.returncode == 0 and "DDP_OK" in r.stdout, r.stdout[-3000:] + r.stderr[-5000:]
