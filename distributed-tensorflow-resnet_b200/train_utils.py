"""Session plumbing shared by the trainers and the eval loop: what MonitoredTrainingSession and its
hooks do around `sess.run(train_op)` in the reference (resnet_cifar_train.py:275-344).

  * SummaryWriter      — SummarySaverHook / tf.summary.FileWriter: TensorBoard event files
                         (`events.out.tfevents.*`, scalar summaries) plus a line-per-record
                         `summaries.jsonl` copy that needs no TensorBoard to read.
  * save_checkpoint / latest_checkpoint / restore_latest
                       — CheckpointSaverHook + the session's auto-restore from `checkpoint_dir`:
                         `model.ckpt-<step>.npz` holding every variable under its TF name, the
                         `<var>/Momentum` slots and `global_step`, and TF's `checkpoint` state file.
  * resume             — restore on EVERY rank, then re-synchronise the replicas from rank 0
                         (weights, moving statistics, momentum, global_step) and move the input
                         pipeline to the batch the interrupted run would have read next.
"""
import json
import os
import time

import numpy as np
import torch


class SummaryWriter(object):
    """Scalar summaries as TensorBoard events (reference: SummarySaverHook(save_steps=100,
    output_dir=FLAGS.log_dir, ...), resnet_cifar_train.py:275-280; tf.summary.FileWriter in
    resnet_cifar_eval.py:69)."""

    def __init__(self, logdir, jsonl=True):
        from tensorboard.summary.writer.event_file_writer import EventFileWriter
        os.makedirs(logdir, exist_ok=True)
        self.logdir = logdir
        self._w = EventFileWriter(logdir)
        self._jsonl = open(os.path.join(logdir, "summaries.jsonl"), "a") if jsonl else None

    def add_scalars(self, step, scalars):
        from tensorboard.compat.proto import event_pb2, summary_pb2
        summ = summary_pb2.Summary(value=[summary_pb2.Summary.Value(tag=k, simple_value=float(v))
                                          for k, v in scalars.items()])
        self._w.add_event(event_pb2.Event(wall_time=time.time(), step=int(step), summary=summ))
        if self._jsonl is not None:
            rec = {"step": int(step)}
            rec.update({k: float(v) for k, v in scalars.items()})
            self._jsonl.write(json.dumps(rec) + "\n")

    def flush(self):
        self._w.flush()
        if self._jsonl is not None:
            self._jsonl.flush()

    def close(self):
        self.flush()
        self._w.close()
        if self._jsonl is not None:
            self._jsonl.close()
            self._jsonl = None


def read_event_scalars(logdir):
    """[(step, tag, value)] of every scalar in the event files of `logdir` (tests, eval tooling)."""
    from tensorboard.backend.event_processing.event_file_loader import EventFileLoader
    out = []
    for f in sorted(os.listdir(logdir)):
        if "tfevents" not in f:
            continue
        for ev in EventFileLoader(os.path.join(logdir, f)).Load():
            for v in ev.summary.value:
                if v.HasField("simple_value"):
                    out.append((int(ev.step), v.tag, float(v.simple_value)))
                elif v.HasField("tensor") and v.tensor.float_val:
                    out.append((int(ev.step), v.tag, float(v.tensor.float_val[0])))
    return out


def save_checkpoint(model, train_dir, step, fmt="npz"):
    """Chief-only CheckpointSaverHook (save_checkpoint_steps=1000, resnet_cifar_train.py:318).
    fmt "npz": one numpy archive; fmt "tf": a TF-V2 bundle (`model.ckpt-N.index` + `.data-00000-of-00001`)
    under the reference's tensor names, readable by tf.train.Saver.restore."""
    os.makedirs(train_dir, exist_ok=True)
    sd = model.executor.state_dict()
    if fmt == "tf":
        from . import tf_io
        path = os.path.join(train_dir, "model.ckpt-%d" % step)
        tf_io.write_checkpoint(path, sd)
    else:
        path = os.path.join(train_dir, "model.ckpt-%d.npz" % step)
        tmp = path + ".tmp.npz"
        np.savez(tmp, **sd)
        os.replace(tmp, path)                   # a killed run never leaves a half-written checkpoint behind
    with open(os.path.join(train_dir, "checkpoint.tmp"), "w") as f:
        f.write('model_checkpoint_path: "%s"\n' % os.path.basename(path))
    os.replace(os.path.join(train_dir, "checkpoint.tmp"), os.path.join(train_dir, "checkpoint"))
    return path


def latest_checkpoint(train_dir):
    """tf.train.get_checkpoint_state(train_dir).model_checkpoint_path, or None."""
    ck = os.path.join(train_dir, "checkpoint")
    if not os.path.exists(ck):
        return None
    name = open(ck).read().split('"')[1]
    path = name if os.path.isabs(name) else os.path.join(train_dir, name)
    return path if (os.path.exists(path) or os.path.exists(path + ".index")) else None


def load_checkpoint_file(path):
    """name -> array for a checkpoint written by save_checkpoint (.npz), a frozen GraphDef (.pb) or a
    TF-V2 checkpoint prefix (`model.ckpt-N` with .index / .data-* beside it)."""
    from . import tf_io
    if path.endswith(".npz"):
        z = np.load(path)
        return {k: z[k] for k in z.files}
    if path.endswith(".pb"):
        return tf_io.read_frozen_graph(path)
    if os.path.exists(path + ".index"):
        return tf_io.read_checkpoint(path)
    raise IOError("unrecognised checkpoint %r" % path)


def restore_latest(model, train_dir):
    path = latest_checkpoint(train_dir)
    if path is None:
        return None
    model.executor.load_state_dict(load_checkpoint_file(path))
    model.global_step = model.executor.global_step
    return os.path.basename(path)


def resume(model, train_dir):
    """What MonitoredTrainingSession does when `checkpoint_dir` already holds a checkpoint
    (resnet_cifar_train.py:330-344), for R replicas: every rank loads the chief's file, rank 0 then
    re-broadcasts weights / moving statistics / momentum and the step counter (so a lagging file
    system cannot leave replicas on different states), and the record pipeline is moved to the batch
    that step `global_step` reads.  Returns the checkpoint name or None."""
    import torch.distributed as dist
    restored = restore_latest(model, train_dir) if train_dir else None
    ex = model.executor
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        flag = torch.tensor([ex.global_step, 1 if restored else 0], dtype=torch.int64, device=ex.dev)
        dist.broadcast(flag, src=0)
        step0, any_restored = int(flag[0].item()), bool(flag[1].item())
        if any_restored:
            ex.broadcast_from_rank0()
            ex.global_step = step0
            model.global_step = step0
            if not restored:
                restored = "(state received from rank 0)"
    if restored:
        pipe = getattr(model._images, "_rn_pipeline", None)
        if pipe is not None:
            pipe.seek(model.global_step)
            model._pulled_first = False
    return restored
