"""A tiny stand-in for tf.app.flags: one process-wide FLAGS namespace that the model wrapper
reads (the reference's resnet_model.py reads FLAGS.dataset / data_format / job_name /
sync_replicas / replicas_to_aggregate / task_index / use_horovod globally,
models/resnet_model.py:71-77,104-131)."""
import argparse


class _Flags(object):
    _defaults = dict(
        dataset="cifar10", mode="train", data_format="channels_first", job_name=None, sync_replicas=True,
        replicas_to_aggregate=None, task_index=None, use_horovod=False,
        # B200 build additions (no reference counterpart)
        compute_dtype="bf16", use_cuda_graph=True, resnet_size=50, seed=0,
    )

    def __init__(self):
        self.__dict__.update(self._defaults)

    def reset(self):
        self.__dict__.clear()
        self.__dict__.update(self._defaults)

    def update(self, **kw):
        self.__dict__.update(kw)


FLAGS = _Flags()


def str2bool(v):
    if isinstance(v, bool):
        return v
    if v.lower() in ("true", "t", "1", "yes", "y"):
        return True
    if v.lower() in ("false", "f", "0", "no", "n"):
        return False
    raise argparse.ArgumentTypeError("boolean expected, got %r" % v)
