"""Symbolic graph of the training step (the TF-1.x "build a graph, then run train_op" model of
the reference, resnet_cifar_train.py:256-259 / models/resnet_model.py:61-67).

Builders (resnet_model_official.py names) append nodes to the default Graph and return symbolic
Tensors; `executor.Executor` turns the graph into a static list of CUDA kernel launches.
Variables are named exactly as TF names them (`conv2d_3/kernel`, `batch_normalization_7/gamma`,
`dense/bias`, ...) so checkpoints / fixtures map one-to-one.
"""
from collections import OrderedDict

ACT = "act"    # activation storage dtype (fp32 in verification mode, bf16 otherwise)
F32 = "f32"


class Tensor:
    """A symbolic NHWC activation or [N, D] matrix."""

    def __init__(self, graph, shape, dtype, name, producer=None):
        self.graph = graph
        self.shape = tuple(int(s) for s in shape)
        self.dtype = dtype
        self.name = name
        self.producer = producer
        self.consumers = []
        self.buf = None       # bound by the executor (torch tensor)
        self.gradbuf = None

    def get_shape(self):
        return self.shape

    @property
    def numel(self):
        n = 1
        for s in self.shape:
            n *= s
        return n

    def __repr__(self):
        return "<Tensor %s %s %s>" % (self.name, self.shape, self.dtype)

    def __add__(self, other):          # `inputs + shortcut`, resnet_model_official.py:130,175
        return self.graph.add(self, other)


class Variable:
    def __init__(self, name, shape, init, trainable=True, decay=True):
        self.name = name
        self.shape = tuple(int(s) for s in shape)
        self.init = init               # dict: kind + parameters (see executor.init_variable)
        self.trainable = trainable
        self.decay = decay             # included in the L2 weight-decay term
        self.offset = None             # element offset in the flat buffer (executor)
        self.size = 1
        for s in self.shape:
            self.size *= s


class Node:
    def __init__(self, op, inputs, outputs, attrs=None, variables=None):
        self.op = op
        self.inputs = inputs
        self.outputs = outputs
        self.attrs = attrs or {}
        self.vars = variables or {}
        for t in inputs:
            t.consumers.append(self)
        for t in outputs:
            t.producer = self


class Graph:
    def __init__(self):
        self.nodes = []
        self.tensors = OrderedDict()
        self.variables = OrderedDict()   # creation order == tf.global_variables() order
        self._counts = {}
        self.placeholders = OrderedDict()

    # ---- naming (TF's "name, name_1, name_2" uniquification) ----
    def unique(self, base):
        i = self._counts.get(base, 0)
        self._counts[base] = i + 1
        return base if i == 0 else "%s_%d" % (base, i)

    def tensor(self, shape, dtype, name, producer=None):
        if name in self.tensors:
            name = self.unique(name)
        t = Tensor(self, shape, dtype, name, producer)
        self.tensors[name] = t
        return t

    def variable(self, name, shape, init, trainable=True, decay=True):
        if name in self.variables:
            v = self.variables[name]
            if v.shape != tuple(shape):
                raise ValueError("variable %s exists with shape %s, requested %s" % (name, v.shape, tuple(shape)))
            return v
        v = Variable(name, shape, init, trainable, decay)
        self.variables[name] = v
        return v

    def trainable_variables(self):
        return [v for v in self.variables.values() if v.trainable]

    def add_node(self, node):
        self.nodes.append(node)
        return node

    # ---- ops ----
    def placeholder(self, shape, dtype, name):
        t = self.tensor(shape, dtype, name)
        self.placeholders[name] = t
        return t

    def conv2d(self, x, kernel_var, k, stride, pad, out_hw, opname):
        N, H, W, Cin = x.shape
        Cout = kernel_var.shape[3]
        y = self.tensor((N, out_hw[0], out_hw[1], Cout), ACT, opname)
        self.add_node(Node("conv2d", [x], [y], dict(k=k, stride=stride, pad=pad), dict(kernel=kernel_var)))
        return y

    def bn_act(self, x, scope, gamma, beta, mmean, mvar, eps, decay, unbiased, leak, is_training, out_name):
        y = self.tensor(x.shape, ACT, out_name)
        self.add_node(Node("bn_act", [x], [y],
                           dict(scope=scope, eps=eps, decay=decay, unbiased=unbiased, leak=leak,
                                is_training=is_training),
                           dict(gamma=gamma, beta=beta, moving_mean=mmean, moving_variance=mvar)))
        return y

    def add(self, a, b, name=None):
        if a.shape != b.shape:
            raise ValueError("add: shapes differ %s vs %s" % (a.shape, b.shape))
        y = self.tensor(a.shape, a.dtype, name or self.unique("add"))
        self.add_node(Node("add", [a, b], [y]))
        return y

    def max_pool_3x3_s2(self, x, pad, out_hw, name):
        N, H, W, C = x.shape
        y = self.tensor((N, out_hw[0], out_hw[1], C), ACT, name)
        self.add_node(Node("maxpool", [x], [y], dict(pad=pad)))
        return y

    def avg_pool(self, x, k, s, name, c_out=None, c_before=0, out_dtype=ACT):
        N, H, W, C = x.shape
        Ho, Wo = (H - k) // s + 1, (W - k) // s + 1
        y = self.tensor((N, Ho, Wo, c_out or C), out_dtype, name)
        self.add_node(Node("avgpool", [x], [y], dict(k=k, s=s, c_before=c_before)))
        return y

    def dense(self, x, w, b, name):
        N = x.shape[0]
        D = x.numel // N
        if D != w.shape[0]:
            raise ValueError("dense: input features %d != kernel rows %d" % (D, w.shape[0]))
        y = self.tensor((N, w.shape[1]), F32, name)
        self.add_node(Node("dense", [x], [y], {}, dict(kernel=w, bias=b)))
        return y

    def softmax_xent(self, logits, labels, name="softmax_cross_entropy_loss"):
        probs = self.tensor(logits.shape, F32, "Softmax")
        scal = self.tensor((2,), F32, name + "/value")   # [mean xent, precision]
        self.add_node(Node("softmax_xent", [logits, labels], [probs, scal]))
        return probs, scal


_default = [Graph()]


def get_default_graph():
    return _default[-1]


def reset_default_graph():
    _default[-1] = Graph()
    return _default[-1]


def tf_same_pad(size, k, s):
    """TF 'SAME': (output size, pad_before) along one axis."""
    out = -(-size // s)
    total = max((out - 1) * s + k - size, 0)
    return out, total // 2
