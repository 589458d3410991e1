"""Where does the end-to-end step spend its extra time?  ResNet-50 / 224 / batch 128 on cuda:0; CUDA-event time per
step of (a) the training step alone, (b) + rn_vgg_preprocess from an HBM-resident pool, (c) + H2D of the batch from
pinned host memory, (d) as (c) but without the per-step D2H of the loss; and the host time of next_batch()."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from distributed_tensorflow_resnet_b200 import resnet_model, FLAGS, vgg_preprocessing

torch.cuda.set_device(0)
dev = torch.device("cuda", 0)
B, S, K = 128, 224, 1001
FLAGS.reset()
FLAGS.update(dataset="imagenet", compute_dtype="bf16", use_cuda_graph=True, resnet_size=50, seed=0)
images = torch.randn(B, S, S, 3, device=dev)
labels = torch.zeros(B, K, device=dev); labels[:, 1] = 1
model = resnet_model.ResNet(resnet_model.HParams(num_classes=K, lrn_rate=0.1, weight_decay_rate=1e-4, optimizer='mom'), images, labels, 'train')
model.build_graph()
model.lrn_rate = 0.1
out_host = torch.zeros(2).pin_memory()
pipes = {r: vgg_preprocessing.ImagenetPipeline(B, image_size=S, num_classes=K, pool=2 * B, seed=0, resident=r, bind=(images, labels))
         for r in (True, False)}


def run(n, pipe, d2h):
    for _ in range(n):
        if pipe is not None:
            pipe.next_batch()
        model.train_op()
        if d2h:
            out_host.copy_(model._loss[1].buf, non_blocking=True)
    torch.cuda.current_stream().synchronize()


for name, pipe, d2h in (("step only", None, False), ("step + D2H loss", None, True), ("resident pool + preprocess", pipes[True], True),
                        ("host pool + H2D + preprocess", pipes[False], True), ("host pool, no D2H", pipes[False], False)):
    run(5, pipe, d2h)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    e0.record()
    run(30, pipe, d2h)
    e1.record()
    torch.cuda.synchronize()
    print("%-34s %.4f ms/step (events)  %.4f ms/step (wall)" % (name, e0.elapsed_time(e1) / 30, (time.perf_counter() - t0) * 1e3 / 30))
for r in (True, False):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(20):
        pipes[r].next_batch()
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    print("next_batch(resident=%s): host %.3f ms per call" % (r, (t1 - t0) * 1e3 / 20))
