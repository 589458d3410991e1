#!/usr/bin/env python
"""Generate the committed golden fixtures from the reference's TF-1.12 graph fixtures.

Runs ONLY in the build container (needs /root/reference and tensorboard's proto
stubs).  The outputs are committed so that the GPU box (which has no
/root/reference) can pin the oracle and the CUDA path against them:

  cifar50_frozen_graph.json    op-by-op structure of
                               test/resnet50-cifar-ckpt-20190218/resnet50_cifar_frozen_model_eval.pb
  cifar50_frozen_weights.npz   the 758,618 trained weights + 49x2 moving stats (by TF name)
  cifar50_train_graph.json     training MetaGraphDef (model.ckpt-107738.meta): variables,
                               optimizer/loss/BN-update constants, initialisers, and the verbatim
                               tf.data function bodies (decode / pad / crop / flip / standardise)
  ref_model.ckpt-107738.index  the reference's TF-V2 checkpoint index file as it is (binary, 14 KB)

Usage:  python tests/golden/make_golden.py
"""
import collections
import json
import os
import sys

import numpy as np
from tensorboard.compat.proto import graph_pb2, meta_graph_pb2

REF = "/root/reference/test/resnet50-cifar-ckpt-20190218"
OUT = os.path.dirname(os.path.abspath(__file__))

_DT = {1: np.float32, 3: np.int32, 9: np.int64, 10: np.bool_, 2: np.float64, 4: np.uint8}


def tensor_to_numpy(t):
    shape = [d.size for d in t.tensor_shape.dim]
    if t.dtype not in _DT:
        return None
    dt = _DT[t.dtype]
    if t.tensor_content:
        return np.frombuffer(t.tensor_content, dtype=dt).reshape(shape).copy()
    for field in ("float_val", "int_val", "int64_val", "bool_val", "double_val"):
        v = list(getattr(t, field))
        if v:
            a = np.array(v, dtype=dt)
            n = int(np.prod(shape)) if shape else 1
            if a.size == 1 and n > 1:
                a = np.full(n, a[0], dtype=dt)
            return a.reshape(shape)
    return np.zeros(shape, dtype=dt)


def attr_to_py(a):
    k = a.WhichOneof("value")
    if k == "list":
        if a.list.i:
            return [int(x) for x in a.list.i]
        if a.list.s:
            return [x.decode() for x in a.list.s]
        return []
    if k == "s":
        return a.s.decode(errors="replace")
    if k == "i":
        return int(a.i)
    if k == "f":
        return float(a.f)
    if k == "b":
        return bool(a.b)
    if k == "type":
        return int(a.type)
    if k == "shape":
        return [d.size for d in a.shape.dim]
    return None


def frozen():
    g = graph_pb2.GraphDef()
    g.ParseFromString(open(os.path.join(REF, "resnet50_cifar_frozen_model_eval.pb"), "rb").read())
    nodes = {n.name: n for n in g.node}

    def resolve(name):
        name = name.split(":")[0].lstrip("^")
        n = nodes[name]
        while n.op == "Identity" and not name.endswith(("initial_conv", "block_layer1", "block_layer2",
                                                         "block_layer3", "final_avg_pool", "final_dense")):
            name = n.input[0].split(":")[0]
            n = nodes[name]
        # also see through the named tf.identity markers
        while n.op == "Identity":
            name = n.input[0].split(":")[0]
            n = nodes[name]
        return name

    weights = {}
    ops = []
    for n in g.node:
        if n.op == "Const":
            arr = tensor_to_numpy(n.attr["value"].tensor)
            if arr is not None and arr.dtype == np.float32 and "/" in n.name:
                weights[n.name] = arr
            continue
        if n.op in ("Identity",):
            continue
        rec = {"name": n.name, "op": n.op, "inputs": [resolve(i) for i in n.input]}
        attrs = {}
        for k in ("strides", "padding", "data_format", "epsilon", "is_training", "ksize", "dilations",
                  "transpose_a", "transpose_b", "use_cudnn_on_gpu"):
            if k in n.attr:
                attrs[k] = attr_to_py(n.attr[k])
        if n.op == "Placeholder":
            attrs["shape"] = attr_to_py(n.attr["shape"])
        if n.op == "Pad":
            attrs["paddings"] = tensor_to_numpy(nodes[resolve(n.input[1])].attr["value"].tensor).tolist()
        if n.op in ("Conv2D", "MatMul"):
            attrs["kernel_shape"] = list(weights[resolve(n.input[1])].shape)
        if n.op == "Reshape":
            attrs["shape"] = tensor_to_numpy(nodes[resolve(n.input[1])].attr["value"].tensor).tolist()
        rec["attrs"] = attrs
        ops.append(rec)
    hist = collections.Counter(n.op for n in g.node)
    json.dump({"source": "resnet50_cifar_frozen_model_eval.pb", "op_histogram": dict(hist), "ops": ops,
               "num_trainable": int(sum(v.size for k, v in weights.items()
                                        if not k.endswith(("moving_mean", "moving_variance"))))},
              open(os.path.join(OUT, "cifar50_frozen_graph.json"), "w"), indent=1)
    np.savez_compressed(os.path.join(OUT, "cifar50_frozen_weights.npz"), **weights)
    print("frozen:", len(ops), "ops,", len(weights), "weight tensors")


def train_meta():
    m = meta_graph_pb2.MetaGraphDef()
    m.ParseFromString(open(os.path.join(REF, "model.ckpt-107738.meta"), "rb").read())
    g = m.graph_def
    nodes = {n.name: n for n in g.node}
    hist = collections.Counter(n.op for n in g.node)

    def shape_of_var(n):
        return [d.size for d in n.attr["shape"].shape.dim]

    variables = [{"name": n.name, "shape": shape_of_var(n), "device": n.device}
                 for n in g.node if n.op == "VariableV2"]
    # trainable variable names in collection order
    trainable = []
    for b in m.collection_def["trainable_variables"].bytes_list.value:
        # VariableDef.variable_name is field 1 (length-delimited)
        ln = b[1]
        trainable.append(b[2:2 + ln].decode().split(":")[0])

    def const_val(name):
        n = nodes[name.split(":")[0]]
        return tensor_to_numpy(n.attr["value"].tensor)

    apply_mom = [n for n in g.node if n.op == "ApplyMomentum"]
    mom = {
        "count": len(apply_mom),
        "use_nesterov": sorted({bool(n.attr["use_nesterov"].b) for n in apply_mom}),
        "use_locking": sorted({bool(n.attr["use_locking"].b) for n in apply_mom}),
        "momentum_const": sorted({float(const_val(n.input[4])) for n in apply_mom}),
        "lr_input": sorted({n.input[2] for n in apply_mom}),
        "grad_input_ops": sorted({nodes[n.input[3].split(":")[0]].op for n in apply_mom}),
    }
    lr_nodes = {i: (nodes[i.split(":")[0]].op, float(const_val(i)) if nodes[i.split(":")[0]].op == "Const" else None)
                for i in mom["lr_input"]}
    # weight decay: cost = xent + wd * AddN(L2Loss...)
    l2 = [n.name for n in g.node if n.op == "L2Loss"]
    wd_consts = []
    for n in g.node:
        if n.op == "Mul" and any(nodes[i.split(":")[0]].op == "AddN" for i in n.input) and "gradients" not in n.name:
            for i in n.input:
                if nodes[i.split(":")[0]].op == "Const":
                    wd_consts.append(float(const_val(i)))
    # BN moving-average update
    assign_sub = [n for n in g.node if n.op == "AssignSub"]
    decays = set()
    for n in assign_sub:
        # AssignSub(mov, Mul(Sub(mov, batch_stat), Sub(1.0, decay)))
        mul = nodes[n.input[1].split(":")[0]]
        one_minus = nodes[mul.input[1].split(":")[0]]
        decays.add(float(const_val(one_minus.input[1])))
        assert float(const_val(one_minus.input[0])) == 1.0
    bn = [n for n in g.node if n.op == "FusedBatchNorm"]
    bn_info = {"count": len(bn), "epsilon": sorted({float(n.attr["epsilon"].f) for n in bn}),
               "is_training": sorted({bool(n.attr["is_training"].b) for n in bn}),
               "data_format": sorted({n.attr["data_format"].s.decode() for n in bn})}
    convs = [n for n in g.node if n.op == "Conv2D"]
    conv_info = [{"name": n.name, "strides": attr_to_py(n.attr["strides"]), "padding": n.attr["padding"].s.decode(),
                  "data_format": n.attr["data_format"].s.decode(), "device": n.device} for n in convs]
    # initialisers: truncated_normal stddev per conv kernel, dense uniform limits
    inits = {}
    for n in g.node:
        if n.op == "Const" and n.name.endswith("Initializer/truncated_normal/stddev"):
            inits[n.name.split("/Initializer")[0]] = {"kind": "truncated_normal", "stddev": float(const_val(n.name))}
        if n.op == "Const" and n.name.endswith("Initializer/random_uniform/max"):
            inits[n.name.split("/Initializer")[0]] = {"kind": "random_uniform", "max": float(const_val(n.name))}
        if n.op == "Const" and n.name.endswith(("Initializer/ones", "Initializer/zeros")):
            inits[n.name.split("/Initializer")[0]] = {"kind": n.name.rsplit("/", 1)[1]}
    # per-variable gradient structure: ApplyMomentum grad input -> AddN inputs' ops
    addn_structs = collections.Counter()
    for n in apply_mom:
        gn = nodes[n.input[3].split(":")[0]]
        if gn.op == "AddN":
            addn_structs[tuple(sorted(nodes[i.split(":")[0]].op + ":" + i.split(":")[0].rsplit("/", 1)[-1]
                                      for i in gn.input))] += 1
    # loss reduction
    loss_nodes = [(n.name, n.op, list(n.input)) for n in g.node if n.name.startswith("softmax_cross_entropy_loss/")
                  and n.op in ("Sum", "RealDiv", "Select", "Equal", "Greater", "Mul", "SoftmaxCrossEntropyWithLogits")]

    funcs = {}
    for f in g.library.function:
        body = []
        for n in f.node_def:
            rec = {"name": n.name, "op": n.op, "inputs": list(n.input)}
            if n.op == "Const":
                v = tensor_to_numpy(n.attr["value"].tensor)
                rec["value"] = v.tolist() if v is not None else None
            for k in ("out_type", "little_endian", "seed", "seed2", "DstT", "SrcT", "axis"):
                if k in n.attr:
                    rec[k] = attr_to_py(n.attr[k])
            body.append(rec)
        funcs[f.signature.name] = {"inputs": [a.name for a in f.signature.input_arg],
                                   "outputs": [a.name for a in f.signature.output_arg],
                                   "ret": dict(f.ret), "nodes": body}
    json.dump({
        "source": "model.ckpt-107738.meta",
        "tensorflow_version": m.meta_info_def.tensorflow_version,
        "tensorflow_git_version": m.meta_info_def.tensorflow_git_version,
        "num_nodes": len(g.node),
        "op_histogram": dict(hist),
        "variables": variables,
        "trainable_variables": trainable,
        "apply_momentum": mom,
        "lr_nodes": lr_nodes,
        "l2loss_count": len(l2),
        "weight_decay_consts": sorted(set(wd_consts)),
        "assign_sub_count": len(assign_sub),
        "bn_moving_average_decay": sorted(decays),
        "fused_batch_norm": bn_info,
        "conv2d": conv_info,
        "initializers": inits,
        "grad_addn_structures": {"|".join(k): v for k, v in addn_structs.items()},
        "loss_nodes": loss_nodes,
        "functions": funcs,
    }, open(os.path.join(OUT, "cifar50_train_graph.json"), "w"), indent=1)
    print("train meta:", len(variables), "variables,", len(trainable), "trainable,", len(funcs), "functions")


def checkpoint_index():
    """The reference's own TF-V2 checkpoint index (14 KB, 403 tensor entries: every variable, its
    Momentum slot, global_step, ...).  The .data shards are not in the reference tree, so the index is
    what pins the bundle reader of distributed-tensorflow-resnet_b200/tf_io.py (names, shapes, dtypes,
    offsets, block checksums)."""
    import shutil
    shutil.copyfile(os.path.join(REF, "model.ckpt-107738.index"), os.path.join(OUT, "ref_model.ckpt-107738.index"))
    print("checkpoint index copied")


if __name__ == "__main__":
    if not os.path.isdir(REF):
        sys.exit("reference fixtures not found at " + REF)
    frozen()
    train_meta()
    checkpoint_index()
