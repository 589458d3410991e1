"""Evaluation loop with the surface of the reference's resnet_cifar_eval.py (evaluate() :60-143):
batch size 100, `cifar_input.build_input(..., mode='eval')`, `ResNet.build_graph(False)` (BN uses the
moving statistics), restore the latest checkpoint of --train_dir, run --eval_batch_count batches,
report `Precision` and `Best Precision` (TensorBoard events in --eval_dir), repeat every 60 s unless
--eval_once.

Checkpoints may be this build's `.npz`, a TF-V2 bundle (`model.ckpt-N.index` + data) written by
either this build or the reference, or a frozen GraphDef `.pb` (resnet_cifar_predict_from_pd.py).

    python -m distributed_tensorflow_resnet_b200.resnet_cifar_train --mode eval --train_dir /tmp/run \
        --eval_data_path synthetic:10000 --eval_batch_count 50 --eval_once true
"""
import sys
import time

import numpy as np
import torch

from . import cifar_input, resnet_model
from .flags import FLAGS
from .train_utils import latest_checkpoint, load_checkpoint_file, SummaryWriter


def evaluate(hps, out=sys.stdout, poll_secs=60, max_rounds=None, inputs=None):
    """`inputs` = (images, labels) device tensors refilled by `images._rn_pipeline` (seek / next_batch); default: the
    CIFAR eval pipeline at batch 100 (reference :62).  resnet_imagenet_eval passes the vgg_preprocessing eval input."""
    if inputs is None:
        batch_size = 100                                          # reference :62
        inputs = cifar_input.build_input(FLAGS.dataset, FLAGS.eval_data_path or "synthetic:10000", batch_size,
                                         FLAGS.mode, seed=FLAGS.seed)
    images, labels = inputs
    pipe = images._rn_pipeline
    model = resnet_model.ResNet(hps, images, labels, FLAGS.mode)
    model.build_graph(False)
    writer = SummaryWriter(FLAGS.eval_dir) if FLAGS.eval_dir else None
    best_precision = 0.0
    result = None
    rounds = 0
    while True:
        path = latest_checkpoint(FLAGS.train_dir) if FLAGS.train_dir else None
        if path is None:
            out.write("No model to eval yet at %s\n" % FLAGS.train_dir)
            if FLAGS.eval_once:
                break
            time.sleep(poll_secs)
            continue
        out.write("Loading checkpoint %s\n" % path)
        model.executor.load_state_dict(load_checkpoint_file(path))
        train_step = model.executor.global_step
        total_prediction, correct_prediction = 0, 0
        loss = 0.0
        pipe.seek(0)                                              # every round evaluates the same batches
        for b in range(FLAGS.eval_batch_count):
            if b:
                pipe.next_batch()
            model.eval_op()
            truth = labels.argmax(1)
            predictions = model.predictions.argmax(1)
            correct_prediction += int((truth == predictions).sum().item())
            total_prediction += int(predictions.shape[0])
            loss = model.executor.scalars()[0]
        precision = 1.0 * correct_prediction / total_prediction
        best_precision = max(precision, best_precision)
        if writer is not None:
            writer.add_scalars(train_step, {"Precision": precision})
            writer.add_scalars(train_step, {"Best Precision": best_precision})
            writer.add_scalars(train_step, {"cross_entropy": loss, "cost": loss})
            writer.flush()
        out.write("loss: %.3f, precision: %.3f, best precision: %.3f\n" % (loss, precision, best_precision))
        out.flush()
        result = dict(loss=loss, precision=precision, best_precision=best_precision, global_step=train_step,
                      total=total_prediction)
        rounds += 1
        if FLAGS.eval_once or (max_rounds is not None and rounds >= max_rounds):
            break
        time.sleep(poll_secs)
    if writer is not None:
        writer.close()
    return result
