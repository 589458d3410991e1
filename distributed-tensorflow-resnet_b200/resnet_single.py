"""The reference's single-process script, resnet_single.py — the one that names the north-star builders: the upstream
tensorflow/models ResNet (`_conv / _batch_norm / _relu / _residual / _bottleneck_residual / _fully_connected`) built
from the nine-field HParams of :194-202 (batch 128 for training, 100 for evaluation, 5 residual units per stage, no
bottleneck, leaky ReLU 0.1, weight decay 2e-4 on the `DW` variables, momentum SGD), the CIFAR record input of
cifar_input.build_input, `_LearningRateSetterHook` 0.1 / 0.01 @40k / 0.001 @60k / 1e-4 @80k (:95-110), summaries every
100 steps and a precision log every 100 (:76-93).  `--log_root` is the parent of train_dir / eval_dir (:44-46).

    python -m distributed_tensorflow_resnet_b200.resnet_single --train_data_path synthetic --train_steps 200
"""
import argparse
import os

from . import resnet_model
from . import resnet_cifar_train as _t
from .flags import FLAGS


def hparams(mode, num_classes):
    """resnet_single.py:184-202."""
    return resnet_model.HParams(batch_size=128 if mode == 'train' else 100, num_classes=num_classes, min_lrn_rate=0.0001,
                                lrn_rate=0.1, num_residual_units=5, use_bottleneck=False, weight_decay_rate=0.0002,
                                relu_leakiness=0.1, optimizer='mom')


def main(argv=None):
    parser = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    _t.define_flags(parser)
    parser.set_defaults(batch_size=None)
    args = parser.parse_args(argv)
    FLAGS.update(**vars(args))
    if FLAGS.dataset == 'cifar10':
        num_classes = 10
    elif FLAGS.dataset == 'cifar100':
        num_classes = 100
    else:
        raise ValueError('Not supported dataset %s' % FLAGS.dataset)
    if FLAGS.log_root:                       # the checkpoints live in log_root/train, the eval summaries in log_root/eval
        FLAGS.train_dir = FLAGS.train_dir or os.path.join(FLAGS.log_root, "train")
        FLAGS.eval_dir = FLAGS.eval_dir or os.path.join(FLAGS.log_root, "eval")
        FLAGS.log_dir = FLAGS.log_dir or FLAGS.train_dir          # SummarySaverHook(output_dir=FLAGS.train_dir), :76-81
    hps = hparams(FLAGS.mode, num_classes)
    if FLAGS.batch_size is None:
        FLAGS.batch_size = hps.batch_size
    FLAGS.replicas_to_aggregate = int(os.environ.get("WORLD_SIZE", "1"))
    if FLAGS.mode == 'train':
        return _t.train(hps, log_every=100)
    if FLAGS.mode == 'eval':
        from . import resnet_cifar_eval
        return resnet_cifar_eval.evaluate(hps)
    raise ValueError("--mode must be train or eval")


if __name__ == '__main__':
    main()
