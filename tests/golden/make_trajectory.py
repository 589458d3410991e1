#!/usr/bin/env python
"""Golden loss trajectory for BASELINE.json config C1 / north_star "same loss trajectory over the
first 200 steps on a fixed seed": upstream ResNet-20 (3 units, filters 16/16/32/64, leakiness 0.1),
batch 128, momentum SGD lr 0.1, wd 2e-4 (resnet_single.py:194-202 hyper-parameters), synthetic CIFAR-10
records through the oracle's decode (pad 4, crop, flip, standardise; seed 5).

Runs the CPU ORACLE (oracle/resnet_oracle.py) twice — fp64 (truth) and fp32 (the TF-CPU stand-in) —
for 200 steps each; ~25 minutes on 8 cores, which is why the result is committed instead of being
recomputed inside the GPU test:

  c1_trajectory.json        cost / xent per step for both precisions + the configuration
  c1_trajectory_init.npz    the initial variables (what Executor._layout draws for seed 5), so the
                            GPU run starts from bit-identical weights whatever the box's RNG does

Usage:  python tests/golden/make_trajectory.py [steps]
"""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
OUT = os.path.dirname(os.path.abspath(__file__))
B, SEED, NREC = 128, 5, 8192


def initial_values():
    from distributed_tensorflow_resnet_b200 import resnet_model, FLAGS, graph as G
    from distributed_tensorflow_resnet_b200.executor import init_variable
    FLAGS.reset()
    FLAGS.update(compute_dtype="f32", dataset="cifar10", seed=SEED)
    hps = resnet_model.HParams(batch_size=B, num_classes=10, min_lrn_rate=1e-4, lrn_rate=0.1, num_residual_units=3,
                               use_bottleneck=False, weight_decay_rate=2e-4, relu_leakiness=0.1, optimizer='mom')
    m = resnet_model.ResNet(hps, torch.zeros(B, 32, 32, 3), torch.zeros(B, 10), "train")
    m.graph = G.reset_default_graph()
    m._build_model(True)
    gen = torch.Generator().manual_seed(SEED)
    return {v.name: init_variable(v, gen).numpy() for v in m.graph.variables.values()}


def index_stream(steps):
    order = lambda e: np.random.Generator(np.random.PCG64([SEED, e])).permutation(NREC).astype(np.int32)
    epoch, cur, cursor = 0, order(0), 0
    for _ in range(steps):
        out = np.empty(B, np.int32)
        filled = 0
        while filled < B:
            take = min(B - filled, len(cur) - cursor)
            out[filled:filled + take] = cur[cursor:cursor + take]
            filled += take
            cursor += take
            if cursor >= len(cur):
                epoch += 1
                cur, cursor = order(epoch), 0
        # the epoch that keys the crop / flip draws is the one current AFTER the batch was assembled
        yield out, epoch


def main():
    from oracle import resnet_oracle as ro, cifar_oracle as co
    steps = int(sys.argv[1]) if len(sys.argv) > 1 else 200
    torch.set_num_threads(os.cpu_count())
    init = initial_values()
    np.savez_compressed(os.path.join(OUT, "c1_trajectory_init.npz"), **init)
    rec = co.synthetic_records(NREC, "cifar10", SEED)
    res = {}
    for tag, dt in (("f32", torch.float32), ("f64", torch.float64)):
        net = ro.UpstreamResNet(10, 3, False, 0.1, dtype=dt, params={k: v for k, v in init.items()})
        tr = ro.Trainer(net, 2e-4, "mom", bn_decay=0.9)
        cost, xent = [], []
        t0 = time.time()
        for s, (idx, epoch) in enumerate(index_stream(steps)):
            img, onehot, _, _ = co.decode(rec, idx, train=True, pad=4, epoch=epoch, seed=SEED)
            o = tr.step([(torch.from_numpy(img).to(dt), torch.from_numpy(onehot).to(dt))], 0.1)[0]
            cost.append(float(o["cost"]))
            xent.append(float(o["xent"]))
            if s % 10 == 0:
                print(tag, s, cost[-1], "%.0f s" % (time.time() - t0), flush=True)
        res[tag] = dict(cost=cost, xent=xent)
    json.dump(dict(config=dict(net="upstream ResNet-20 (num_residual_units=3, relu_leakiness=0.1)", batch=B, steps=steps,
                               lr=0.1, weight_decay=2e-4, momentum=0.9, seed=SEED, records=NREC, pad=4,
                               bn_decay=0.9, torch=torch.__version__),
                   f64=res["f64"], f32=res["f32"]),
              open(os.path.join(OUT, "c1_trajectory.json"), "w"), indent=1)
    print("written")


if __name__ == "__main__":
    main()
