"""The reference's combined ImageNet entry point, resnet_imagenet_main.py: `--mode train` is the trainer of
resnet_imagenet_train.py (with `--variable_update`, :98-104), `--mode eval` the polling evaluate() (:393-457)."""
import sys

from . import resnet_imagenet_train, resnet_imagenet_eval


def main(argv=None):
    argv = list(sys.argv[1:] if argv is None else argv)
    mode = "train"
    for i, a in enumerate(argv):
        if a == "--mode" and i + 1 < len(argv):
            mode = argv[i + 1]
        elif a.startswith("--mode="):
            mode = a.split("=", 1)[1]
    return resnet_imagenet_eval.main(argv) if mode == "eval" else resnet_imagenet_train.main(argv)


if __name__ == '__main__':
    main()
