"""B200-native synchronous data-parallel ResNet training step (drop-in for the hot path of
michaelwfc/distributed-tensorflow-resnet).  Importing this package loads librn_b200.so and
raises if it is missing: there is no CPU or eager fallback."""
from . import _lib            # noqa: F401  (fails loudly when the CUDA library is absent)
from . import graph, flags, resnet_model_official, resnet_model, executor  # noqa: F401
from .resnet_model import HParams, ResNet  # noqa: F401
from .flags import FLAGS  # noqa: F401

__all__ = ["HParams", "ResNet", "FLAGS", "graph", "resnet_model", "resnet_model_official", "executor"]
