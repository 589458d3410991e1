"""Times rn_vgg_preprocess alone (pool resident in HBM, 128 images -> 224x224x3 bf16) with CUDA events."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from distributed_tensorflow_resnet_b200 import vgg_preprocessing, FLAGS

FLAGS.reset(); FLAGS.update(compute_dtype="bf16")
torch.cuda.set_device(0)
B = 128
pipe = vgg_preprocessing.ImagenetPipeline(B, image_size=224, num_classes=1001, pool=2 * B, seed=0, resident=True)
for _ in range(5):
    pipe.next_batch()
torch.cuda.synchronize()
best = 1e9
for rep in range(5):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        pipe.next_batch()
    e1.record()
    torch.cuda.synchronize()
    best = min(best, e0.elapsed_time(e1) / 20)
src = float(pipe.pool.host_sizes.mean()) * B
out = B * 224 * 224 * 3 * 2
print("vgg_preprocess: %.4f ms per 128-image batch (includes 2 small H2D table copies + a memset); source %.1f MB, output %.1f MB -> %.0f GB/s of (source + output)"
      % (best, src / 1e6, out / 1e6, (src + out) / best / 1e6))
