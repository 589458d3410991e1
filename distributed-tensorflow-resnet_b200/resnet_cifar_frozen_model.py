"""The reference's resnet_cifar_frozen_model.py: restore the newest checkpoint of --train_dir into the eval graph and
write it as a frozen GraphDef — every variable a `Const` node under its TF name — next to it
(`<train_dir>/frozen_model.pb`, :100-125), then run one batch through the frozen weights and print prediction / label
/ precision (:161-200).  The file is a GraphDef protobuf that the reference's `tf.import_graph_def` reads and that
`ResNet.load_frozen_graph` / resnet_cifar_predict_from_pd load back.

    python -m distributed_tensorflow_resnet_b200.resnet_cifar_frozen_model --train_dir /tmp/run [--output model.pb]
"""
import argparse
import os
import sys

from . import cifar_input, resnet_model, tf_io
from . import resnet_cifar_train as _t
from .flags import FLAGS
from .train_utils import latest_checkpoint, load_checkpoint_file


def predict_batch(model, labels, out=sys.stdout):
    """One batch through the eval graph: (predicted classes, true classes, precision), printed as the reference does."""
    model.eval_op()
    prediction = model.predictions.argmax(1)
    truth = labels.argmax(1)
    precision = float((prediction == truth).float().mean().item())
    out.write("The prediction:{0}\nThe label:{1}\nThe precision:{2}\n".format(prediction.tolist(), truth.tolist(), precision))
    return prediction.cpu().numpy(), truth.cpu().numpy(), precision


def build_eval_model(batch_size):
    num_classes = 10 if FLAGS.dataset == 'cifar10' else 100
    images, labels = cifar_input.build_input(FLAGS.dataset, FLAGS.eval_data_path or "synthetic:10000", batch_size, 'eval',
                                             seed=FLAGS.seed)
    hps = resnet_model.HParams(num_classes=num_classes, lrn_rate=0.1, weight_decay_rate=2e-4, optimizer='mom')
    model = resnet_model.ResNet(hps, images, labels, 'eval')
    return model, labels


def main(argv=None, out=sys.stdout):
    parser = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    _t.define_flags(parser)
    parser.add_argument('--output', default='', help='frozen GraphDef to write (default <train_dir>/frozen_model.pb)')
    parser.set_defaults(mode='eval', batch_size=100)
    args = parser.parse_args(argv)
    FLAGS.update(**vars(args))
    path = latest_checkpoint(FLAGS.train_dir) if FLAGS.train_dir else None
    if path is None:
        raise IOError("no checkpoint in --train_dir %r" % FLAGS.train_dir)
    model, labels = build_eval_model(FLAGS.batch_size)
    model.build_graph(False)
    model.executor.load_state_dict(load_checkpoint_file(path))
    sd = model.executor.state_dict()
    tensors = {k: v for k, v in sd.items() if not k.endswith("/Momentum") and k != "global_step"}
    pb = FLAGS.output or os.path.join(FLAGS.train_dir, "frozen_model.pb")
    tf_io.write_frozen_graph(pb, tensors)
    out.write("%d ops in the final graph: %s\n" % (len(tensors), pb))
    res = predict_batch(model, labels, out)
    return dict(path=pb, tensors=len(tensors), precision=res[2], prediction=res[0], label=res[1])


if __name__ == '__main__':
    main()
