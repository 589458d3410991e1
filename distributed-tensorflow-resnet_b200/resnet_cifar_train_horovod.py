"""The reference's Horovod CIFAR trainer, resnet_cifar_train_horovod.py (hvd.DistributedOptimizer = averaged all-reduce
of every gradient, `BroadcastGlobalVariablesHook(0)`, checkpoints on rank 0 only, :296-330): the same trainer with
`--use_horovod true --variable_update horovod` preset.  One process per GPU (torchrun instead of mpirun)."""
from . import resnet_cifar_train as _t


def main(argv=None):
    import sys
    argv = list(sys.argv[1:] if argv is None else argv)
    return _t.main(["--use_horovod", "true", "--variable_update", "horovod"] + argv)


if __name__ == '__main__':
    main()
