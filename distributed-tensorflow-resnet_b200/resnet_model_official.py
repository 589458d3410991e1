"""Pre-activation ResNet v2 builders with the reference's names and signatures
(/root/reference/resnet_model_official.py:41-366), emitting nodes of the B200 graph instead of
TensorFlow ops.  Tensors are NHWC internally for every `data_format` (the reference transposes
to NCHW only to please cuDNN/MKL, :244-248; the result is layout-independent).

Differences from the TF original that a user can observe: none in values; `data_format` is
accepted and ignored; the returned `model(inputs, is_training)` takes a symbolic
`graph.Tensor` (or a torch CUDA tensor, which is wrapped in a placeholder).
"""
import math

from . import graph as G

_BATCH_NORM_DECAY = 0.997
_BATCH_NORM_EPSILON = 1e-5
# tf.layers.batch_normalization(fused=True) bumps 1e-5 to 1.001e-5 for the fused kernel
# (fixture: FusedBatchNorm epsilon = 1.0009999641624745e-05 on all 49 nodes).
_FUSED_BATCH_NORM_EPSILON = 1.001e-5


def _as_tensor(inputs, name="X"):
    if isinstance(inputs, G.Tensor):
        return inputs
    g = G.get_default_graph()
    t = g.placeholder(tuple(inputs.shape), G.ACT, name)
    t.bound = inputs
    return t


def batch_norm_relu(inputs, is_training, data_format=None):
    """Performs a batch normalization followed by a ReLU (reference :41-50)."""
    g = inputs.graph
    scope = g.unique("batch_normalization")
    C = inputs.shape[-1]
    gamma = g.variable(scope + "/gamma", (C,), dict(kind="ones"))
    beta = g.variable(scope + "/beta", (C,), dict(kind="zeros"))
    mm = g.variable(scope + "/moving_mean", (C,), dict(kind="zeros"), trainable=False, decay=False)
    mv = g.variable(scope + "/moving_variance", (C,), dict(kind="ones"), trainable=False, decay=False)
    return g.bn_act(inputs, scope, gamma, beta, mm, mv, _FUSED_BATCH_NORM_EPSILON, _BATCH_NORM_DECAY, True, 0.0,
                    bool(is_training), g.unique("Relu"))


def fixed_padding(inputs, kernel_size, data_format=None):
    """Reference :53-77 pads explicitly with tf.pad.  Here padding is folded into the conv
    kernels' out-of-bounds zero fill, so this only annotates the tensor with the pad that the
    following VALID convolution must assume."""
    pad_total = kernel_size - 1
    inputs._fixed_pad = (pad_total // 2, pad_total - pad_total // 2)
    return inputs


def conv2d_fixed_padding(inputs, filters, kernel_size, strides, data_format=None):
    """Strided 2-D convolution with explicit padding (reference :80-91): SAME for stride 1,
    symmetric fixed_padding + VALID for stride > 1; no bias; variance_scaling initialiser."""
    g = inputs.graph
    N, H, W, Cin = inputs.shape
    scope = g.unique("conv2d")
    pad = (kernel_size - 1) // 2
    if strides > 1:
        fixed_padding(inputs, kernel_size, data_format)
        Ho = (H + kernel_size - 1 - kernel_size) // strides + 1
        Wo = (W + kernel_size - 1 - kernel_size) // strides + 1
    else:
        Ho, Wo = H, W
    fan_in = kernel_size * kernel_size * Cin
    kernel = g.variable(scope + "/kernel", (kernel_size, kernel_size, Cin, filters),
                        dict(kind="truncated_normal", stddev=math.sqrt(1.0 / fan_in) / 0.87962566103423978))
    return g.conv2d(inputs, kernel, kernel_size, strides, pad, (Ho, Wo), scope + "/Conv2D")


def building_block(inputs, filters, is_training, projection_shortcut, strides, data_format=None):
    """Standard building block with BN before convolutions (reference :94-130)."""
    shortcut = inputs
    inputs = batch_norm_relu(inputs, is_training, data_format)
    # the projection shortcut comes after the first batch norm and ReLU
    if projection_shortcut is not None:
        shortcut = projection_shortcut(inputs)
    inputs = conv2d_fixed_padding(inputs, filters, 3, strides, data_format)
    inputs = batch_norm_relu(inputs, is_training, data_format)
    inputs = conv2d_fixed_padding(inputs, filters, 3, 1, data_format)
    return inputs + shortcut


def bottleneck_block(inputs, filters, is_training, projection_shortcut, strides, data_format=None):
    """Bottleneck block variant, stride on the 3x3 (reference :133-175)."""
    shortcut = inputs
    inputs = batch_norm_relu(inputs, is_training, data_format)
    if projection_shortcut is not None:
        shortcut = projection_shortcut(inputs)
    inputs = conv2d_fixed_padding(inputs, filters, 1, 1, data_format)
    inputs = batch_norm_relu(inputs, is_training, data_format)
    inputs = conv2d_fixed_padding(inputs, filters, 3, strides, data_format)
    inputs = batch_norm_relu(inputs, is_training, data_format)
    inputs = conv2d_fixed_padding(inputs, 4 * filters, 1, 1, data_format)
    return inputs + shortcut


def block_layer(inputs, filters, block_fn, blocks, strides, is_training, name, data_format=None):
    """One layer of blocks (reference :178-214): first block projects + strides."""
    filters_out = 4 * filters if block_fn is bottleneck_block else filters

    def projection_shortcut(inputs):
        return conv2d_fixed_padding(inputs, filters_out, 1, strides, data_format)

    inputs = block_fn(inputs, filters, is_training, projection_shortcut, strides, data_format)
    for _ in range(1, blocks):
        inputs = block_fn(inputs, filters, is_training, None, 1, data_format)
    inputs.alias = name          # tf.identity(inputs, name)
    return inputs


def _head(inputs, pool, feat, num_classes):
    g = inputs.graph
    inputs = g.avg_pool(inputs, pool, 1, "average_pooling2d/AvgPool", out_dtype=G.F32)
    lim = math.sqrt(6.0 / (feat + num_classes))
    w = g.variable("dense/kernel", (feat, num_classes), dict(kind="uniform", limit=lim))
    b = g.variable("dense/bias", (num_classes,), dict(kind="zeros"))
    return g.dense(inputs, w, b, "dense/BiasAdd")


def cifar10_resnet_v2_generator(resnet_size, num_classes, data_format=None):
    """Generator for CIFAR-10 ResNet v2 models (reference :217-278)."""
    if resnet_size % 6 != 2:
        raise ValueError('resnet_size must be 6n + 2:', resnet_size)
    num_blocks = (resnet_size - 2) // 6

    def model(inputs, is_training):
        inputs = _as_tensor(inputs)
        inputs = conv2d_fixed_padding(inputs, 16, 3, 1, data_format)
        inputs = block_layer(inputs, 16, building_block, num_blocks, 1, is_training, 'block_layer1', data_format)
        inputs = block_layer(inputs, 32, building_block, num_blocks, 2, is_training, 'block_layer2', data_format)
        inputs = block_layer(inputs, 64, building_block, num_blocks, 2, is_training, 'block_layer3', data_format)
        inputs = batch_norm_relu(inputs, is_training, data_format)
        return _head(inputs, 8, 64, num_classes)

    return model


def imagenet_resnet_v2_generator(block_fn, layers, num_classes, data_format=None):
    """Generator for ImageNet ResNet v2 models (reference :281-347)."""

    def model(inputs, is_training):
        inputs = _as_tensor(inputs)
        g = inputs.graph
        inputs = conv2d_fixed_padding(inputs, 64, 7, 2, data_format)
        N, H, W, C = inputs.shape
        Ho, ph = G.tf_same_pad(H, 3, 2)
        Wo, _ = G.tf_same_pad(W, 3, 2)
        inputs = g.max_pool_3x3_s2(inputs, ph, (Ho, Wo), "max_pooling2d/MaxPool")
        for i, (f, s) in enumerate(zip((64, 128, 256, 512), (1, 2, 2, 2))):
            inputs = block_layer(inputs, f, block_fn, layers[i], s, is_training, 'block_layer%d' % (i + 1),
                                 data_format)
        inputs = batch_norm_relu(inputs, is_training, data_format)
        return _head(inputs, inputs.shape[1], 512 if block_fn is building_block else 2048, num_classes)

    return model


def imagenet_resnet_v2(resnet_size, num_classes, data_format=None):
    """Returns the ResNet model for a given size and number of output classes (reference :350-366)."""
    model_params = {
        18: {'block': building_block, 'layers': [2, 2, 2, 2]},
        34: {'block': building_block, 'layers': [3, 4, 6, 3]},
        50: {'block': bottleneck_block, 'layers': [3, 4, 6, 3]},
        101: {'block': bottleneck_block, 'layers': [3, 4, 23, 3]},
        152: {'block': bottleneck_block, 'layers': [3, 8, 36, 3]},
        200: {'block': bottleneck_block, 'layers': [3, 24, 36, 3]},
    }
    if resnet_size not in model_params:
        raise ValueError('Not a valid resnet_size:', resnet_size)
    params = model_params[resnet_size]
    return imagenet_resnet_v2_generator(params['block'], params['layers'], num_classes, data_format)
