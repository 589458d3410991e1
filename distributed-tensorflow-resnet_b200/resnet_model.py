"""ResNet model wrapper with the reference's surface (/root/reference/models/resnet_model.py):

    hps = HParams(num_classes=10, lrn_rate=0.1, weight_decay_rate=2e-4, optimizer='mom')
    model = ResNet(hps, images, labels, 'train'); model.build_graph()
    while ...: model.train_op()            # == sess.run(model.train_op)

`images` [B,H,W,3] and `labels` [B,num_classes] (one-hot fp32) are device tensors that the input
pipeline refills in place every step (the analogue of the tf.data iterator's outputs).
`HParams` accepts the reference's 4 fields and the legacy 9-field form of resnet_single.py:194-202
(`batch_size, min_lrn_rate, num_residual_units, use_bottleneck, relu_leakiness`).  When
`num_residual_units` is given the upstream builders `_conv/_batch_norm/_relu/_residual/
_bottleneck_residual/_fully_connected` are used (tensorflow/models research/resnet — the module the
reference's legacy script was written against); otherwise the official v2 generators,
exactly as models/resnet_model.py:72-77 does with resnet_size=50.
"""
import math
from collections import namedtuple

import torch

from . import graph as G
from . import resnet_model_official
from .executor import Executor
from .flags import FLAGS

_HParams = namedtuple('HParams',
                      'num_classes, lrn_rate, weight_decay_rate, optimizer, '
                      'batch_size, min_lrn_rate, num_residual_units, use_bottleneck, relu_leakiness')


def HParams(num_classes, lrn_rate, weight_decay_rate, optimizer, batch_size=None, min_lrn_rate=None,
            num_residual_units=None, use_bottleneck=False, relu_leakiness=0.0):
    return _HParams(num_classes, lrn_rate, weight_decay_rate, optimizer, batch_size, min_lrn_rate,
                    num_residual_units, use_bottleneck, relu_leakiness)


class _Inert(object):
    """Stand-in for the SyncReplicasOptimizer session plumbing (hooks, init ops, queue runners):
    with NCCL allreduce there is no token queue or chief thread to start."""

    def __init__(self, name):
        self.name = name

    def __call__(self, *a, **k):
        return None

    def __repr__(self):
        return "<inert %s>" % self.name


class ResNet(object):
    """ResNet model."""

    def __init__(self, hps, images, labels, mode):
        """Args as the reference (models/resnet_model.py:45-59): hps, images
        [batch, size, size, 3], labels [batch, num_classes], mode 'train' | 'eval'."""
        self.hps = hps
        self._images = images
        self.labels = labels
        self.mode = mode
        self._extra_train_ops = []
        self.executor = None
        self._pulled_first = False      # build_input already decoded the first batch

    # ------------------------------------------------------------------ graph
    def build_graph(self, istrain=True):
        """Build the whole graph (reference :61-67) and compile it into CUDA launches."""
        self.graph = G.reset_default_graph()
        self.global_step = 0
        self._build_model(istrain)
        self._build_train_op() if self.mode == 'train' else self._build_eval_op(istrain)
        self.summaries = ['cross_entropy', 'cost', 'learning_rate']
        return self

    def _placeholders(self):
        g = self.graph
        x = g.placeholder(tuple(self._images.shape), G.ACT, "X")
        x.bound = self._images
        y = g.placeholder(tuple(self.labels.shape), G.F32, "Y")
        y.bound = self.labels
        return x, y

    def _build_model(self, istrain):
        """Reference :69-91: network -> logits -> softmax / cross-entropy; the weight-decay term of
        `cost` is evaluated inside the fused optimizer kernel."""
        x, y = self._placeholders()
        self._upstream = self.hps.num_residual_units is not None
        if self._upstream:
            logits = self._build_upstream(x, istrain)
        elif FLAGS.dataset in ('cifar10', 'cifar100'):
            network = resnet_model_official.cifar10_resnet_v2_generator(
                resnet_size=FLAGS.resnet_size, num_classes=self.hps.num_classes, data_format=FLAGS.data_format)
            logits = network(x, istrain)
        elif FLAGS.dataset == 'imagenet':
            network = resnet_model_official.imagenet_resnet_v2(
                resnet_size=FLAGS.resnet_size, num_classes=self.hps.num_classes, data_format=FLAGS.data_format)
            logits = network(x, istrain)
        else:
            raise ValueError('Not supported dataset %s' % FLAGS.dataset)
        self._logits = logits
        self._loss = self.graph.softmax_xent(logits, y)

    # ---- upstream (north-star-named) builders -------------------------------------------
    def _stride_arr(self, stride):
        return [1, stride, stride, 1]

    def _conv(self, name, x, filter_size, in_filters, out_filters, strides):
        """tf.nn.conv2d(x, DW, strides, padding='SAME'); DW ~ N(0, sqrt(2/(k*k*out)))."""
        stride = strides[1] if isinstance(strides, (list, tuple)) else int(strides)
        g = x.graph
        N, H, W, C = x.shape
        if C != in_filters:
            raise ValueError("_conv %s: input has %d channels, in_filters=%d" % (name, C, in_filters))
        n = filter_size * filter_size * out_filters
        kernel = g.variable(name + "/DW", (filter_size, filter_size, in_filters, out_filters),
                            dict(kind="normal", stddev=math.sqrt(2.0 / n)))
        Ho, ph = G.tf_same_pad(H, filter_size, stride)
        Wo, _ = G.tf_same_pad(W, filter_size, stride)
        return g.conv2d(x, kernel, filter_size, stride, ph, (Ho, Wo), name + "/Conv2D")

    def _batch_norm(self, name, x, leakiness=None):
        """tf.nn.moments + tf.nn.batch_normalization(eps=0.001), moving averages with decay 0.9 of
        the (biased) batch moments.  The activation that always follows (`_relu`) is fused: pass
        `leakiness`, or call `_relu` on the result."""
        g = x.graph
        C = x.shape[-1]
        beta = g.variable(name + "/beta", (C,), dict(kind="zeros"), decay=False)
        gamma = g.variable(name + "/gamma", (C,), dict(kind="ones"), decay=False)
        mm = g.variable(name + "/moving_mean", (C,), dict(kind="zeros"), trainable=False, decay=False)
        mv = g.variable(name + "/moving_variance", (C,), dict(kind="ones"), trainable=False, decay=False)
        leak = 1.0 if leakiness is None else float(leakiness)     # leak 1.0 == identity activation
        y = g.bn_act(x, name, gamma, beta, mm, mv, 0.001, 0.9, False, leak, self.mode == 'train' and self._istrain,
                     name + "/leaky_relu")
        y._pending_bn = y.producer if leakiness is None else None
        return y

    def _relu(self, x, leakiness=0.0):
        """tf.where(x < 0, leakiness * x, x).  Directly after `_batch_norm` it is folded into the
        BN kernel (no extra HBM pass)."""
        node = getattr(x, "_pending_bn", None)
        if node is None:
            raise NotImplementedError("_relu is only supported directly after _batch_norm")
        node.attrs["leak"] = float(leakiness)
        x._pending_bn = None
        return x

    def _bn_relu(self, name, x):
        return self._relu(self._batch_norm(name, x), self.hps.relu_leakiness)

    def _residual(self, x, in_filter, out_filter, stride, activate_before_residual=False, scope=""):
        g = x.graph
        s = stride[1] if isinstance(stride, (list, tuple)) else int(stride)
        if activate_before_residual:
            x = self._bn_relu(scope + "shared_activation/init_bn", x)
            orig_x = x
        else:
            orig_x = x
            x = self._bn_relu(scope + "residual_only_activation/init_bn", x)
        x = self._conv(scope + "sub1/conv1", x, 3, in_filter, out_filter, s)
        x = self._bn_relu(scope + "sub2/bn2", x)
        x = self._conv(scope + "sub2/conv2", x, 3, out_filter, out_filter, 1)
        if in_filter != out_filter:
            pad = (out_filter - in_filter) // 2
            # avg_pool(stride, stride, 'VALID') + zero-pad of the channel axis, one kernel
            orig_x = g.avg_pool(orig_x, s, s, scope + "sub_add/Pad", c_out=out_filter, c_before=pad)
        return g.add(x, orig_x, scope + "sub_add/add")

    def _bottleneck_residual(self, x, in_filter, out_filter, stride, activate_before_residual=False, scope=""):
        g = x.graph
        s = stride[1] if isinstance(stride, (list, tuple)) else int(stride)
        if activate_before_residual:
            x = self._bn_relu(scope + "common_bn_relu/init_bn", x)
            orig_x = x
        else:
            orig_x = x
            x = self._bn_relu(scope + "residual_bn_relu/init_bn", x)
        x = self._conv(scope + "sub1/conv1", x, 1, in_filter, out_filter // 4, s)
        x = self._bn_relu(scope + "sub2/bn2", x)
        x = self._conv(scope + "sub2/conv2", x, 3, out_filter // 4, out_filter // 4, 1)
        x = self._bn_relu(scope + "sub3/bn3", x)
        x = self._conv(scope + "sub3/conv3", x, 1, out_filter // 4, out_filter, 1)
        if in_filter != out_filter:
            orig_x = self._conv(scope + "sub_add/project", orig_x, 1, in_filter, out_filter, s)
        return g.add(x, orig_x, scope + "sub_add/add")

    def _fully_connected(self, x, out_dim):
        """reshape [batch, -1]; DW ~ uniform_unit_scaling(1.0); biases 0; xw_plus_b."""
        g = x.graph
        D = x.numel // x.shape[0]
        w = g.variable("logit/DW", (D, out_dim), dict(kind="uniform", limit=math.sqrt(3.0 / D)))
        b = g.variable("logit/biases", (out_dim,), dict(kind="zeros"), decay=False)
        return g.dense(x, w, b, "logit/xw_plus_b")

    def _global_avg_pool(self, x):
        return x.graph.avg_pool(x, x.shape[1], 1, "unit_last/global_avg_pool", out_dtype=G.F32)

    def _build_upstream(self, x, istrain):
        hps = self.hps
        self._istrain = istrain
        filters = getattr(self, "filters", None) or ([16, 64, 128, 256] if hps.use_bottleneck else [16, 16, 32, 64])
        res = self._bottleneck_residual if hps.use_bottleneck else self._residual
        x = self._conv("init/init_conv", x, 3, x.shape[-1], filters[0], self._stride_arr(1))
        strides = [1, 2, 2]
        abr = [True, False, False]
        for st in range(3):
            x = res(x, filters[st], filters[st + 1], self._stride_arr(strides[st]), abr[st], scope="unit_%d_0/" % (st + 1))
            for i in range(1, hps.num_residual_units):
                x = res(x, filters[st + 1], filters[st + 1], self._stride_arr(1), False, scope="unit_%d_%d/" % (st + 1, i))
        x = self._bn_relu("unit_last/final_bn", x)
        x = self._global_avg_pool(x)
        return self._fully_connected(x, hps.num_classes)

    # ------------------------------------------------------------------ train / eval ops
    def _world(self):
        """SyncReplicasOptimizer(replicas_to_aggregate=R, total_num_replicas=R) / hvd: the mean of
        exactly R gradients per step, R = number of ranks (reference :104-120; main() overwrites
        replicas_to_aggregate with the worker count, resnet_cifar_train.py:375-376)."""
        import torch.distributed as dist
        if not (dist.is_available() and dist.is_initialized()):
            return 1, 0
        world = dist.get_world_size()
        if not (FLAGS.sync_replicas or FLAGS.use_horovod):
            if world > 1:
                # the reference's asynchronous PS mode is not carried over; R independent, silently
                # diverging replicas would be the result of honouring the flags literally
                raise ValueError("a process group with %d ranks is initialised but gradient aggregation is off "
                                 "(sync_replicas=False and use_horovod=False): set FLAGS.sync_replicas=True (or "
                                 "use_horovod=True), or run one process" % world)
            return 1, 0
        R = FLAGS.replicas_to_aggregate
        if R is not None and int(R) != world:
            raise ValueError("replicas_to_aggregate=%s must equal the number of ranks (%d): the reference always "
                             "sets it to the worker count (resnet_cifar_train.py:375-376).  To aggregate over "
                             "k < %d replicas launch k processes (torchrun --nproc-per-node k): the communicator, "
                             "and with it the 1/k of the mean, is the set of ranks that exist"
                             % (R, world, world))
        return world, dist.get_rank()

    def _make_executor(self, train):
        world, rank = self._world()
        self.executor = Executor(
            self.graph, dtype=FLAGS.compute_dtype, weight_decay=self.hps.weight_decay_rate,
            optimizer=self.hps.optimizer, momentum=0.9, lrn_rate=self.hps.lrn_rate, train=train, seed=FLAGS.seed,
            world_size=world, rank=rank, loss=self._loss, init_values=getattr(self, "init_values", None))
        return self.executor

    def _build_train_op(self):
        """Reference :93-144."""
        if self.hps.optimizer not in ('sgd', 'mom'):
            raise ValueError("optimizer must be 'sgd' or 'mom'")
        ex = self._make_executor(True)
        self._lrn_rate = self.hps.lrn_rate
        if FLAGS.use_cuda_graph:
            ex.capture()
        self.train_op = self._train_step
        is_sync = FLAGS.job_name is not None and FLAGS.sync_replicas
        for name in ("replicas_hook", "local_init_op", "ready_for_local_init_op", "chief_queue_runner", "sync_init_op"):
            setattr(self, name, _Inert(name) if is_sync else None)
        self.replicas_hook = _Inert("replicas_hook")     # the serial path also lists it (reference bug F10)
        self.init_op = _Inert("init_op")

    def _build_eval_op(self, istrain):
        ex = self._make_executor(False)
        self.train_op = None
        self.eval_op = ex.run_forward
        self.init_op = _Inert("init_op")

    def _train_step(self):
        pipe = getattr(self._images, "_rn_pipeline", None)
        if pipe is not None and self._pulled_first:
            pipe.next_batch()           # the dequeue that train_op depends on in the reference graph
        self._pulled_first = True
        self.executor.step()
        self.global_step = self.executor.global_step

    # ------------------------------------------------------------------ reference artefacts
    def load_frozen_graph(self, path):
        """Weights and BN moving statistics from a frozen GraphDef written by the reference
        (resnet_cifar_frozen_model.py; consumed by resnet_cifar_predict_from_pd.py:66-103 through
        import_graph_def), matched by TF variable name.  Before build_graph() the values become the
        initial values; afterwards they are loaded into the running executor.  Returns the names."""
        from . import tf_io
        w = tf_io.read_frozen_graph(path)
        if self.executor is None:
            self.init_values = dict(getattr(self, "init_values", None) or {}, **w)
        else:
            missing = [v.name for v in self.graph.variables.values() if v.name not in w]
            if missing:
                raise KeyError("%s lacks %d variables of this model, e.g. %s" % (path, len(missing), missing[:3]))
            self.executor.load_state_dict(w)
        return sorted(w)

    def load_checkpoint(self, path):
        """`saver.restore(sess, path)`: an .npz of this build, a TF-V2 bundle prefix, or a frozen .pb."""
        from .train_utils import load_checkpoint_file
        self.executor.load_state_dict(load_checkpoint_file(path))
        self.global_step = self.executor.global_step

    # ------------------------------------------------------------------ fetchable values
    @property
    def lrn_rate(self):
        return self._lrn_rate

    @lrn_rate.setter
    def lrn_rate(self, v):
        self._lrn_rate = float(v)
        if self.executor is not None:
            self.executor.set_lr(v)

    @property
    def predictions(self):
        return self._loss[0].buf

    @property
    def logits(self):
        return self._logits.buf

    @property
    def cost(self):
        return self.executor.scalars()[2]

    @property
    def cross_entropy(self):
        return self.executor.scalars()[0]

    @property
    def precision(self):
        return self.executor.scalars()[1]
