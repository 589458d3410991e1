"""CIFAR trainer with the flag surface and loop of the reference's resnet_cifar_train.py
(flags :32-87, train() :251-344, main() :346-406), running the B200 training step.

Launch one process per GPU (torchrun); `--worker_hosts` / `--ps_hosts` / `--job_name` are
accepted for command-line compatibility: the number of ranks replaces len(worker_hosts) as
replicas_to_aggregate (the reference overwrites the flag the same way, :375-376), `--job_name=ps`
exits at once (there are no parameter servers), and gradients are aggregated by an NCCL allreduce.

    python -m distributed_tensorflow_resnet_b200.resnet_cifar_train --train_data_path synthetic \
        --batch_size 128 --train_steps 200
"""
import argparse
import os
import sys
import time

import numpy as np
import torch

from . import cifar_input, resnet_model
from .flags import FLAGS, str2bool

_WEIGHT_DECAY = 2e-4      # reference :98-100


def define_flags(parser):
    a = parser.add_argument
    a('--dataset', default='cifar10', help='cifar10 or cifar100.')
    a('--mode', default='train', help='train or eval.')
    a('--train_data_path', default='synthetic', help="Filepattern for training data ('synthetic' = generated records).")
    a('--eval_data_path', default='', help='Filepattern for eval data')
    a('--image_size', type=int, default=32, help='Image side length.')
    a('--train_dir', default='', help='Directory to keep training outputs (checkpoints).')
    a('--eval_dir', default='', help='Directory to keep eval outputs.')
    a('--eval_batch_count', type=int, default=50, help='Number of batches to eval.')
    a('--eval_once', type=str2bool, default=False, help='Whether evaluate the model only once.')
    a('--log_dir', default='', help='Directory for summaries.')
    a('--log_root', default='', help='(legacy resnet_single.py) parent of train_dir/eval_dir.')
    a('--num_gpus', type=int, default=0, help='Accepted; one process drives one GPU.')
    a('--task_index', type=int, default=None, help='Worker task index (default: RANK).')
    a('--replicas_to_aggregate', type=int, default=None, help='Forced to the number of workers.')
    a('--train_steps', type=int, default=2000, help='Number of (global) training steps to perform')
    a('--batch_size', type=int, default=32, help='Training batch size (per worker)')
    a('--learning_rate', type=float, default=0.01, help='Learning rate (unused, as in the reference)')
    a('--sync_replicas', type=str2bool, default=True, help='Synchronous aggregation (async is not supported)')
    a('--existing_servers', type=str2bool, default=False, help='Accepted and ignored.')
    a('--ps_hosts', default='localhost:2222', help='Accepted and ignored (no parameter servers).')
    a('--worker_hosts', default='', help='Accepted; the rank count comes from the launcher.')
    a('--job_name', default=None, help='worker or ps (ps exits immediately)')
    a('--data_format', default='channels_first', help='Accepted; kernels are NHWC.')
    a('--num_intra_threads', type=int, default=0)
    a('--num_inter_threads', type=int, default=0)
    a('--use_horovod', type=str2bool, default=False, help='Allreduce aggregation (the only mode here).')
    a('--variable_update', default='parameter_server',
      choices=['independent', 'parameter_server', 'replicated', 'distributed_replicated',
               'distributed_all_reduce', 'collective_all_reduce', 'horovod'])
    # additions of this build
    a('--compute_dtype', default='bf16', choices=['bf16', 'f32'], help='f32 = verification mode')
    a('--use_cuda_graph', type=str2bool, default=True)
    a('--resnet_size', type=int, default=50)
    a('--seed', type=int, default=0)
    a('--save_checkpoint_steps', type=int, default=1000, help='chief checkpoint cadence (reference: 1000)')
    a('--checkpoint_format', default='npz', choices=['npz', 'tf'],
      help="npz = one numpy archive per checkpoint; tf = TF-V2 bundle (.index/.data) under the reference's tensor names")


def cifar_learning_rate(train_step):
    """_LearningRateSetterHook.after_run (reference :301-311)."""
    if train_step < 40000:
        return 0.1
    elif train_step < 60000:
        return 0.01
    elif train_step < 80000:
        return 0.001
    return 0.0001


def init_distributed():
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world > 1 and not dist.is_initialized():
        local = int(os.environ.get("LOCAL_RANK", "0"))
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    return (dist.get_rank(), dist.get_world_size()) if dist.is_initialized() else (0, 1)


from .train_utils import save_checkpoint, restore_latest, resume, SummaryWriter  # noqa: E402,F401


def run_training_loop(model, lr_begin, lr_fn, world, out, log_every=20, summary_every=100, checkpoint_every=None):
    """The body of `with MonitoredTrainingSession(...) as mon_sess: while not mon_sess.should_stop():
    mon_sess.run(model.train_op)` with the reference's hooks (resnet_cifar_train.py:275-344):
    LoggingTensorHook every 20 iterations {step, loss, precision, lr}; SummarySaverHook every 100
    steps on the chief (cross_entropy, cost, learning_rate, Precision -> TensorBoard events in
    --log_dir); _LearningRateSetterHook (rate for the NEXT run from the step this run returned);
    CheckpointSaverHook every 1000 steps and at the end on the chief; StopAtStepHook(train_steps).
    A run started on a --train_dir that already holds a checkpoint resumes from it on every rank."""
    import torch.distributed as dist
    rank = dist.get_rank() if dist.is_available() and dist.is_initialized() else 0
    is_chief = rank == 0
    restored = resume(model, FLAGS.train_dir) if FLAGS.train_dir else None
    if restored and is_chief:
        out.write("restored %s (global_step %d)\n" % (restored, model.global_step))
    lrn_rate = lr_fn(model.global_step) if restored else lr_begin     # _LearningRateSetterHook.begin
    writer = SummaryWriter(FLAGS.log_dir) if (FLAGS.log_dir and is_chief) else None
    fmt = getattr(FLAGS, "checkpoint_format", "npz")
    if checkpoint_every is None:
        checkpoint_every = int(getattr(FLAGS, "save_checkpoint_steps", 1000) or 1000)
    t0 = time.time()
    it = 0
    last_saved = model.global_step if restored else -1
    while model.global_step < FLAGS.train_steps:
        model.lrn_rate = lrn_rate
        model.train_op()
        it += 1
        step = model.global_step
        want_log = it % log_every == 0 or step >= FLAGS.train_steps
        want_summ = writer is not None and (it == 1 or step % summary_every == 0)
        if want_log or want_summ:
            xent, prec, cost = model.executor.scalars()
            if want_log:
                dt = time.time() - t0
                out.write("step = %d, loss = %.6f, precision = %.4f, lr = %g (%.1f images/sec)\n" %
                          (step, cost, prec, lrn_rate, it * FLAGS.batch_size * world / max(dt, 1e-9)))
                out.flush()
            if want_summ:
                writer.add_scalars(step, {"cross_entropy": xent, "cost": cost, "learning_rate": lrn_rate,
                                          "Precision": prec})
                writer.flush()
        lrn_rate = lr_fn(step)
        if FLAGS.train_dir and is_chief and step % checkpoint_every == 0:
            save_checkpoint(model, FLAGS.train_dir, step, fmt)
            last_saved = step
    torch.cuda.synchronize()
    if FLAGS.train_dir and is_chief and last_saved != model.global_step:
        save_checkpoint(model, FLAGS.train_dir, model.global_step, fmt)
    if writer is not None:
        writer.close()
    return model


def train(hps, lr_fn=cifar_learning_rate, log_every=20, out=sys.stdout):
    """Training loop (reference :251-344)."""
    rank, world = init_distributed()
    images, labels = cifar_input.input_fn(True, FLAGS.train_data_path, FLAGS.batch_size, num_epochs=1000,
                                          dataset=FLAGS.dataset, seed=FLAGS.seed)
    model = resnet_model.ResNet(hps, images, labels, FLAGS.mode)
    model.build_graph()
    return run_training_loop(model, 0.1, lr_fn, world, out, log_every=log_every)


def main(argv=None):
    parser = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    define_flags(parser)
    args = parser.parse_args(argv)
    FLAGS.update(**vars(args))
    if FLAGS.job_name == 'ps':
        print("job_name=ps: there are no parameter servers in the NCCL build; exiting.")
        return None
    if FLAGS.job_name == "":
        raise ValueError("Must specify an explicit `job_name`")
    if not FLAGS.sync_replicas:
        print("warning: --sync_replicas=False (asynchronous PS updates) is not supported; running synchronously")
        FLAGS.sync_replicas = True
    if FLAGS.dataset == 'cifar10':
        num_classes = 10
    elif FLAGS.dataset == 'cifar100':
        num_classes = 100
    else:
        raise ValueError('Not supported dataset %s' % FLAGS.dataset)
    world = int(os.environ.get("WORLD_SIZE", "1"))
    FLAGS.replicas_to_aggregate = world          # reference :375-376
    if FLAGS.task_index is None:
        FLAGS.task_index = int(os.environ.get("RANK", "0"))
    hps = resnet_model.HParams(num_classes=num_classes, lrn_rate=0.1, weight_decay_rate=_WEIGHT_DECAY, optimizer='mom')
    if FLAGS.mode == 'train':
        return train(hps)
    if FLAGS.mode == 'eval':
        from . import resnet_cifar_eval
        return resnet_cifar_eval.evaluate(hps)
    raise ValueError("--mode must be train or eval")


if __name__ == '__main__':
    main()
