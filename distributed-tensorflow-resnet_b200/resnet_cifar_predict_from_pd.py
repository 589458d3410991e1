"""The reference's resnet_cifar_predict_from_pd.py (:66-103): load a frozen GraphDef (`.pb`, e.g. the reference's own
test/resnet50-cifar-ckpt-20190218/resnet50_cifar_frozen_model_eval.pb or one written by resnet_cifar_frozen_model),
run a batch of CIFAR records through it — BN on the stored moving statistics — and print prediction / label /
precision.

    python -m distributed_tensorflow_resnet_b200.resnet_cifar_predict_from_pd --frozen_model model.pb --EVAL_NUM 100
"""
import argparse
import sys

from . import resnet_cifar_train as _t
from .flags import FLAGS
from .resnet_cifar_frozen_model import build_eval_model, predict_batch


def main(argv=None, out=sys.stdout):
    parser = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    _t.define_flags(parser)
    parser.add_argument('--frozen_model', required=True, help='frozen GraphDef (.pb)')
    parser.add_argument('--EVAL_NUM', type=int, default=100, help='The number of images we want to predict (resnet_cifar_predict.py:23)')
    parser.set_defaults(mode='eval')
    args = parser.parse_args(argv)
    FLAGS.update(**vars(args))
    model, labels = build_eval_model(FLAGS.EVAL_NUM)
    names = model.load_frozen_graph(FLAGS.frozen_model)          # initial values of the graph about to be built
    model.build_graph(False)
    prediction, truth, precision = predict_batch(model, labels, out)
    return dict(loaded=len(names), precision=precision, prediction=prediction, label=truth)


if __name__ == '__main__':
    main()
