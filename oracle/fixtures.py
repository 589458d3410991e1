"""Loader for the committed golden fixtures (tests/golden/, made by tests/golden/make_golden.py
from the reference's test/resnet50-cifar-ckpt-20190218/ files).  TEST INFRASTRUCTURE."""
import json
import os

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def frozen_graph():
    return json.load(open(os.path.join(GOLDEN, "cifar50_frozen_graph.json")))


def train_graph():
    return json.load(open(os.path.join(GOLDEN, "cifar50_train_graph.json")))


def frozen_weights():
    """dict TF name -> float32 array: 152 trainable tensors (758,618 values) + 98 moving stats."""
    z = np.load(os.path.join(GOLDEN, "cifar50_frozen_weights.npz"))
    return {k: z[k] for k in z.files}
