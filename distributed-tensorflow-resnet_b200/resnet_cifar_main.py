"""The reference's combined CIFAR entry point, resnet_cifar_main.py: the trainer of resnet_cifar_train.py plus
`--variable_update` (:91-97; parameter_server and horovod are the two the reference implements — both are the NCCL
all-reduce here) and `--mode eval` running the polling `evaluate()` (:361-421).  Same flags, same loop:

    torchrun --nproc-per-node 8 -m distributed_tensorflow_resnet_b200.resnet_cifar_main --variable_update horovod ...
    python -m distributed_tensorflow_resnet_b200.resnet_cifar_main --mode eval --train_dir /tmp/run --eval_once true
"""
from .resnet_cifar_train import main, train, define_flags, cifar_learning_rate   # noqa: F401

if __name__ == '__main__':
    main()
