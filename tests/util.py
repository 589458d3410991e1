"""Helpers shared by the tests."""
import ctypes as C

import numpy as np
import torch

import distributed_tensorflow_resnet_b200 as pkg
from distributed_tensorflow_resnet_b200 import _lib
from distributed_tensorflow_resnet_b200._lib import lib, check, ptr, ConvDesc, RN_F32, RN_BF16

DT = {"f32": (torch.float32, RN_F32), "bf16": (torch.bfloat16, RN_BF16)}
# north_star tolerances: 1e-4 relative in fp32 verification mode, 2e-2 in bf16 mode
TOL = {"f32": 1e-4, "bf16": 2e-2}


def stream():
    return torch.cuda.current_stream().cuda_stream


def rel_err(got, want):
    """max-norm relative error (SURVEY 8(c) parity protocol)."""
    got = torch.as_tensor(got).double().cpu()
    want = torch.as_tensor(want).double().cpu()
    denom = want.abs().max().item()
    if denom == 0:
        return got.abs().max().item()
    return (got - want).abs().max().item() / denom


def dev(t, dtype=None):
    t = torch.as_tensor(t)
    if dtype is not None:
        t = t.to(dtype)
    return t.contiguous().cuda()
