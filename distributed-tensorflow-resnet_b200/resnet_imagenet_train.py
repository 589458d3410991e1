"""ImageNet-shape trainer with the flags / loop / LR schedule of the reference's
resnet_imagenet_train.py (flags :33-99, train() :206-281, main() :283-338).  There is no dataset
or network here, so the TFRecord + JPEG + vgg_preprocessing input (:105-187) is replaced by
synthetic decoded uint8 images of random sizes run through the GPU restatement of vgg_preprocessing
(`--train_data_path synthetic[:N]`, the default), or by one fixed standard-normal batch (`--train_data_path fixed`).

    torchrun --nproc-per-node 8 -m distributed_tensorflow_resnet_b200.resnet_imagenet_train \
        --batch_size 128 --train_steps 100
"""
import argparse
import os
import sys

import torch

from . import resnet_model
from . import resnet_cifar_train as _cifar
from .flags import FLAGS

_WEIGHT_DECAY = 1e-4       # reference :288-291
_NUM_CLASSES = 1000


def imagenet_learning_rate(train_step):
    """_LearningRateSetterHook.after_run (reference :236-260): Intel-Caffe 8-node schedule —
    warm-up 0.1 -> 0.4 over 6240 steps, then x0.1 at 37440 / 74880 / 99840."""
    if train_step < 6240:
        return 0.1 + 0.3 * train_step / 6240.0
    elif train_step < 37440:
        return 0.4
    elif train_step < 74880:
        return 0.4 * 0.1
    elif train_step < 99840:
        return 0.4 * 0.01
    return 0.4 * 0.001


def synthetic_batch(batch_size, image_size, num_classes, rank, dtype, device):
    """Standard-normal images (seed 1000 + rank) and random labels, generated once."""
    gen = torch.Generator().manual_seed(1000 + rank)
    img = torch.randn(batch_size, image_size, image_size, 3, generator=gen)
    lab = torch.randint(0, num_classes, (batch_size,), generator=gen)
    onehot = torch.zeros(batch_size, num_classes)
    onehot[torch.arange(batch_size), lab] = 1.0
    return img.to(device=device, dtype=dtype), onehot.to(device)


def train(hps, out=sys.stdout, log_every=40):
    rank, world = _cifar.init_distributed()
    dev = torch.device("cuda", torch.cuda.current_device())
    dt = torch.bfloat16 if FLAGS.compute_dtype == "bf16" else torch.float32
    if str(FLAGS.train_data_path).startswith("synthetic"):
        # the reference's input_fn (:150-187): decoded images of arbitrary sizes -> vgg_preprocessing on the GPU
        from . import vgg_preprocessing
        images, labels = vgg_preprocessing.input_fn(True, FLAGS.train_data_path, FLAGS.batch_size, FLAGS.num_epochs,
                                                    image_size=FLAGS.image_size, num_classes=hps.num_classes,
                                                    seed=FLAGS.seed, dtype=dt)
    else:       # 'fixed': one standard-normal batch resident on the device (kernel benchmarking)
        images, labels = synthetic_batch(FLAGS.batch_size, FLAGS.image_size, hps.num_classes, rank, dt, dev)
    model = resnet_model.ResNet(hps, images, labels, FLAGS.mode)
    model.build_graph()
    # _LearningRateSetterHook.begin sets 0.4 for the first run (reference :241); hooks as the CIFAR trainer
    return _cifar.run_training_loop(model, 0.4, imagenet_learning_rate, world, out, log_every=log_every)


def main(argv=None):
    parser = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    _cifar.define_flags(parser)
    parser.add_argument('--num_epochs', type=int, default=90)
    parser.set_defaults(dataset='imagenet', image_size=224, batch_size=128, train_steps=200)
    args = parser.parse_args(argv)
    FLAGS.update(**vars(args))
    if FLAGS.job_name == 'ps':
        print("job_name=ps: there are no parameter servers in the NCCL build; exiting.")
        return None
    FLAGS.replicas_to_aggregate = int(os.environ.get("WORLD_SIZE", "1"))   # reference :307-308
    hps = resnet_model.HParams(num_classes=_NUM_CLASSES, lrn_rate=0.1, weight_decay_rate=_WEIGHT_DECAY, optimizer='mom')
    return train(hps)


if __name__ == '__main__':
    main()
