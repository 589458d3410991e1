"""ImageNet evaluation loop with the surface of the reference's resnet_imagenet_eval.py (flags :30-80,
evaluate() :169-230, main() :233-250): `input_fn(False, eval_data_path, batch_size)` — here decoded uint8 images of
arbitrary sizes through the GPU restatement of `vgg_preprocessing.preprocess_for_eval` (resize the short side to 256,
central 224 crop, mean subtraction) — `ResNet.build_graph(False)` (BN on the moving statistics), then per round: load
the newest checkpoint of --train_dir, run --eval_batch_count batches, report loss / Precision / Best Precision (also as
TensorBoard scalars in --eval_dir), sleep 60 s unless --eval_once.

    python -m distributed_tensorflow_resnet_b200.resnet_imagenet_eval --train_dir /tmp/run --eval_once true
"""
import argparse
import sys

import torch

from . import resnet_model, resnet_cifar_eval, vgg_preprocessing
from . import resnet_cifar_train as _cifar
from .flags import FLAGS
from .resnet_imagenet_train import _NUM_CLASSES, _WEIGHT_DECAY


def evaluate(hps, out=sys.stdout, poll_secs=60, max_rounds=None):
    dt = torch.bfloat16 if FLAGS.compute_dtype == "bf16" else torch.float32
    inputs = vgg_preprocessing.input_fn(False, FLAGS.eval_data_path or "synthetic:1024", FLAGS.batch_size, 1,
                                        image_size=FLAGS.image_size, num_classes=hps.num_classes, seed=FLAGS.seed, dtype=dt)
    return resnet_cifar_eval.evaluate(hps, out=out, poll_secs=poll_secs, max_rounds=max_rounds, inputs=inputs)


def main(argv=None):
    parser = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    _cifar.define_flags(parser)
    parser.add_argument('--num_epochs', type=int, default=1)
    parser.set_defaults(dataset='imagenet', mode='eval', image_size=224, batch_size=128)       # reference :31-32, :53
    args = parser.parse_args(argv)
    FLAGS.update(**vars(args))
    hps = resnet_model.HParams(num_classes=_NUM_CLASSES, lrn_rate=0.1, weight_decay_rate=_WEIGHT_DECAY, optimizer='mom')
    return evaluate(hps)


if __name__ == '__main__':
    main()
