"""ImageNet input preprocessing on the GPU with the entry points of the reference's
vgg_preprocessing.py (:284-353), operating on a BATCH of decoded uint8 images of arbitrary sizes:

    batch = ImageBatch(list_of_HWC_uint8_arrays)                 # packed into one device buffer
    x = preprocess_image(batch, 224, 224, is_training=True)      # -> [B, 224, 224, 3] device tensor

One `rn_vgg_preprocess` launch does image/255, the aspect-preserving bilinear resize
(align_corners=False) to a random smallest side in [256, 512], the random 224x224 crop, the random
left-right flip and the mean subtraction; eval mode resizes to `resize_side` and crops centrally.
JPEG decoding stays outside (the reference decodes with tf.image.decode_image; there is no dataset here).

`ImagenetPipeline` is the synthetic stand-in for the reference's TFRecord input_fn
(resnet_imagenet_train.py:150-187): a pool of seeded random images of random sizes, resident in HBM or in
pinned host memory, preprocessed into the model's input tensor every step.
"""
import numpy as np
import torch

from ._lib import lib, check, ptr, RN_F32, RN_BF16, RnError
from .flags import FLAGS

_R_MEAN = 123.68 / 255
_G_MEAN = 116.78 / 255
_B_MEAN = 103.94 / 255

_RESIZE_SIDE_MIN = 256
_RESIZE_SIDE_MAX = 512


class ImageBatch(object):
    """Decoded HWC uint8 images packed back to back (+ the byte offset and (height, width) of each)."""

    def __init__(self, images, device=None, pinned=False):
        self.n = len(images)
        hw = np.array([[im.shape[0], im.shape[1]] for im in images], dtype=np.int32).reshape(-1, 2)
        sizes = hw[:, 0].astype(np.int64) * hw[:, 1] * 3
        off = np.zeros(self.n, dtype=np.int64)
        if self.n:
            off[1:] = np.cumsum(sizes)[:-1]
        flat = np.concatenate([np.ascontiguousarray(im, dtype=np.uint8).reshape(-1) for im in images]) if self.n else \
            np.zeros(0, np.uint8)
        self.host_hw, self.host_off, self.host_sizes = hw, off, sizes
        self.nbytes = int(flat.size)
        if pinned:
            self.host = torch.from_numpy(flat).pin_memory()
            self.data = None
        else:
            dev = torch.device(device if device is not None else "cuda:%d" % torch.cuda.current_device())
            self.host = None
            self.data = torch.from_numpy(flat).to(dev)
            self.hw = torch.from_numpy(hw).to(dev)
            self.offsets = torch.from_numpy(off).to(dev)


def _run(data, offsets, hw, B, out, output_height, output_width, side_min, side_max, train, seed, epoch, index, draws_in,
         draws_out, err, labels=None, classes=0, onehot=None):
    odt = RN_BF16 if out.dtype == torch.bfloat16 else RN_F32
    check(lib.rn_vgg_preprocess(ptr(data), ptr(offsets), ptr(hw), B, output_height, output_width, side_min, side_max,
                                int(train), seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF, epoch & 0xFFFFFFFF,
                                ptr(index), ptr(draws_in), ptr(out), odt, ptr(draws_out), ptr(err),
                                ptr(labels), classes, ptr(onehot),
                                torch.cuda.current_stream().cuda_stream), "rn_vgg_preprocess")


def _preprocess(batch, output_height, output_width, train, side_min, side_max, seed, epoch, index, draws, out, dtype,
                check_errors=True):
    if batch.data is None:
        raise RnError("ImageBatch lives in host memory: copy it to the device first (ImagenetPipeline does)")
    dev = batch.data.device
    if out is None:
        if dtype is None:
            dtype = torch.bfloat16 if FLAGS.compute_dtype == "bf16" else torch.float32
        out = torch.empty(batch.n, output_height, output_width, 3, dtype=dtype, device=dev)
    used = torch.zeros(batch.n, 4, dtype=torch.int32, device=dev)
    err = torch.zeros(1, dtype=torch.int32, device=dev)
    idx = None if index is None else torch.as_tensor(np.asarray(index, dtype=np.int32)).to(dev)
    din = None if draws is None else torch.as_tensor(np.asarray(draws, dtype=np.int32)).to(dev).contiguous()
    _run(batch.data, batch.offsets, batch.hw, batch.n, out, output_height, output_width, side_min, side_max, train, seed,
         epoch, idx, din, used, err)
    out._rn_draws = used
    if check_errors and int(err.item()):
        raise ValueError("Crop size greater than the image size. (image %d)" % (int(err.item()) - 1))
    return out


def preprocess_for_train(image, output_height, output_width, resize_side_min=_RESIZE_SIDE_MIN,
                         resize_side_max=_RESIZE_SIDE_MAX, *, seed=0, epoch=0, index=None, draws=None, out=None, dtype=None):
    """Reference :284-314 on an ImageBatch.  The draws used (side, offset_y, offset_x, flip) per image are
    attached to the result as `._rn_draws`."""
    return _preprocess(image, output_height, output_width, True, resize_side_min, resize_side_max, seed, epoch, index,
                       draws, out, dtype)


def preprocess_for_eval(image, output_height, output_width, resize_side, *, out=None, dtype=None):
    """Reference :317-333: resize to `resize_side`, central crop, mean subtraction."""
    return _preprocess(image, output_height, output_width, False, resize_side, resize_side, 0, 0, None, None, out, dtype)


def preprocess_image(image, output_height, output_width, is_training=False, resize_side_min=_RESIZE_SIDE_MIN,
                     resize_side_max=_RESIZE_SIDE_MAX, **kw):
    """Reference :336-353."""
    if is_training:
        return preprocess_for_train(image, output_height, output_width, resize_side_min, resize_side_max, **kw)
    return preprocess_for_eval(image, output_height, output_width, resize_side_min, **kw)


def synthetic_images(n, seed=0, lo=240, hi=520):
    """n seeded random uint8 images with heights / widths in [lo, hi] (there is no dataset or network here)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    sizes = rng.integers(lo, hi + 1, size=(n, 2))
    return [rng.integers(0, 256, size=(int(h), int(w), 3), dtype=np.uint8) for h, w in sizes]


class ImagenetPipeline(object):
    """Synthetic stand-in for input_fn(is_training, data_dir, batch_size) of resnet_imagenet_train.py:150-187:
    owns a pool of decoded images and the two device tensors the model is bound to (images [B,S,S,3], one-hot
    labels [B,1000]); `next_batch()` picks the next B images (a per-epoch permutation, rank r takes images
    r, r+R, ...) and preprocesses them with one kernel launch.  resident=False keeps the pool in pinned host
    memory and copies each batch's images H2D first (what a loader fed by a JPEG decoder would do)."""

    def __init__(self, batch_size, image_size=224, num_classes=1000, pool=512, train=True, seed=0, dtype=None,
                 device=None, rank=0, world=1, resident=True, images=None, labels=None, bind=None):
        dev = torch.device(device if device is not None else "cuda:%d" % torch.cuda.current_device())
        self.B, self.S, self.classes = int(batch_size), int(image_size), int(num_classes)
        self.train, self.seed, self.rank, self.world = bool(train), int(seed), rank, world
        imgs = images if images is not None else synthetic_images(pool, seed)
        self.n = len(imgs)
        rng = np.random.Generator(np.random.PCG64([seed, 7]))
        self.label_ids = np.asarray(labels if labels is not None else rng.integers(0, num_classes, self.n), dtype=np.int64)
        self.pool = ImageBatch(imgs, dev, pinned=not resident)
        self.resident = bool(resident)
        if dtype is None:
            dtype = torch.bfloat16 if FLAGS.compute_dtype == "bf16" else torch.float32
        if bind is not None:          # write into tensors a model is already bound to
            self.images, self.labels = bind
            assert tuple(self.images.shape) == (self.B, self.S, self.S, 3) and tuple(self.labels.shape) == (self.B, self.classes)
        else:
            self.images = torch.zeros(self.B, self.S, self.S, 3, dtype=dtype, device=dev)
            self.labels = torch.zeros(self.B, self.classes, dtype=torch.float32, device=dev)
        self.draws = torch.zeros(self.B, 4, dtype=torch.int32, device=dev)
        self._err = torch.zeros(1, dtype=torch.int32, device=dev)
        max_bytes = int(np.sort(self.pool.host_sizes)[::-1][:self.B].sum()) if self.n else 0
        nring = 4
        # per-step tables (byte offsets int64; height/width, image ids, labels int32) and, for a host-resident pool,
        # the image bytes: a ring of pinned staging buffers with device copies, so the host only waits for the GPU
        # when it is `nring` batches ahead
        B = self.B
        self._off_host = [torch.zeros(B, dtype=torch.int64).pin_memory() for _ in range(nring)]
        self._off_dev = [torch.zeros(B, dtype=torch.int64, device=dev) for _ in range(nring)]
        self._i32_host = [torch.zeros(4 * B, dtype=torch.int32).pin_memory() for _ in range(nring)]   # hw[B][2], ids[B], labels[B]
        self._i32_dev = [torch.zeros(4 * B, dtype=torch.int32, device=dev) for _ in range(nring)]
        self._evt = [None] * nring
        if not self.resident:
            self._img_dev = [torch.zeros(max_bytes, dtype=torch.uint8, device=dev) for _ in range(nring)]
            self._copy_stream = torch.cuda.Stream(device=dev)
        self._slot = 0
        self.epoch = 0
        self._cursor = 0
        self._order = self._epoch_order(0)
        self.h2d_bytes_last = 0
        self.images._rn_pipeline = self
        self.labels._rn_pipeline = self

    def _epoch_order(self, epoch):
        order = np.random.Generator(np.random.PCG64([self.seed, epoch])).permutation(self.n) if self.train \
            else np.arange(self.n)
        return order[self.rank::self.world].astype(np.int64)

    def next_indices(self):
        out = np.empty(self.B, dtype=np.int64)
        filled = 0
        while filled < self.B:
            take = min(self.B - filled, len(self._order) - self._cursor)
            out[filled:filled + take] = self._order[self._cursor:self._cursor + take]
            filled += take
            self._cursor += take
            if self._cursor >= len(self._order):
                self.epoch += 1
                self._order = self._epoch_order(self.epoch)
                self._cursor = 0
        return out

    def seek(self, batches):
        self.epoch, self._cursor = 0, 0
        self._order = self._epoch_order(0)
        for _ in range(int(batches)):
            self.next_indices()
        return self.next_batch()

    def next_batch(self):
        idx = self.next_indices()
        epoch = self.epoch
        st = torch.cuda.current_stream()
        k = self._slot
        self._slot = (k + 1) % len(self._off_host)
        if self._evt[k] is not None:
            self._evt[k].synchronize()
        B = self.B
        p = self.pool
        sizes = p.host_sizes[idx]
        off_h = self._off_host[k].numpy()
        i32 = self._i32_host[k].numpy()
        if self.resident:
            off_h[:] = p.host_off[idx]
            data = p.data
            self.h2d_bytes_last = off_h.nbytes + i32.nbytes
        else:
            off_h[0] = 0
            off_h[1:] = np.cumsum(sizes)[:-1]
            # H2D of this batch's raw images straight from the pinned pool (one async copy per image, no host-side
            # gather), on a copy stream so that it overlaps the previous step's kernels
            total = int(off_h[-1] + sizes[-1])
            dst = self._img_dev[k]
            cs = self._copy_stream
            # (slot k is free: the host waited above for the event recorded after its previous reader, 4 batches ago)
            with torch.cuda.stream(cs):
                for i in range(B):
                    a, n_ = int(p.host_off[idx[i]]), int(sizes[i])
                    dst[int(off_h[i]):int(off_h[i]) + n_].copy_(p.host[a:a + n_], non_blocking=True)
            st.wait_stream(cs)
            data = dst
            self.h2d_bytes_last = total + off_h.nbytes + i32.nbytes
        i32[0:2 * B] = p.host_hw[idx].reshape(-1)
        i32[2 * B:3 * B] = idx
        i32[3 * B:4 * B] = self.label_ids[idx]
        self._off_dev[k].copy_(self._off_host[k], non_blocking=True)
        self._i32_dev[k].copy_(self._i32_host[k], non_blocking=True)
        t = self._i32_dev[k]
        self._err.zero_()              # a 4-byte memset
        _run(data, self._off_dev[k], t[0:2 * B], B, self.images, self.S, self.S, _RESIZE_SIDE_MIN, _RESIZE_SIDE_MAX,
             self.train, self.seed, epoch, t[2 * B:3 * B], None, self.draws, self._err, t[3 * B:4 * B], self.classes,
             self.labels)
        self._evt[k] = torch.cuda.Event()
        self._evt[k].record(st)
        return self.images, self.labels

    def set_resident(self, resident):
        resident = bool(resident)
        if resident == self.resident:
            return
        raise RnError("choose resident / host-resident when the pipeline is built")


def input_fn(is_training, data_dir, batch_size, num_epochs=1, **kw):
    """Reference resnet_imagenet_train.py:150-187.  `data_dir` = 'synthetic' or 'synthetic:N' (N pooled images);
    there are no TFRecord / JPEG readers in this build."""
    import torch.distributed as dist
    if not (isinstance(data_dir, str) and data_dir.startswith("synthetic")):
        raise IOError("ImageNet TFRecord input is not available in this build (no dataset, no JPEG decoder): "
                      "use --train_data_path synthetic[:N]")
    pool = int(data_dir.split(":")[1]) if ":" in data_dir else 512
    rank, world = (dist.get_rank(), dist.get_world_size()) if dist.is_available() and dist.is_initialized() else (0, 1)
    p = ImagenetPipeline(batch_size, pool=pool, train=is_training, rank=rank, world=world, **kw)
    p.num_epochs = num_epochs
    p.next_batch()
    return p.images, p.labels
