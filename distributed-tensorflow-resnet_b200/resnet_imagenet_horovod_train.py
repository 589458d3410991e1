"""The reference's Horovod ImageNet trainer, resnet_imagenet_horovod_train.py: the trainer of resnet_imagenet_train.py
with `--use_horovod true --variable_update horovod` preset (averaged all-reduce, broadcast from rank 0, rank-0
checkpoints).  One process per GPU (torchrun instead of mpirun)."""
from . import resnet_imagenet_train as _t


def main(argv=None):
    import sys
    argv = list(sys.argv[1:] if argv is None else argv)
    return _t.main(["--use_horovod", "true", "--variable_update", "horovod"] + argv)


if __name__ == '__main__':
    main()
