"""CIFAR binary-record input on the GPU, with the reference's entry points:

    images, labels = cifar_input.build_input(dataset, data_path, batch_size, mode)   # cifar_input.py:25
    images, labels = cifar_input.input_fn(is_training, data_dir, batch_size, num_epochs)  # resnet_cifar_train.py:204

Both return two DEVICE tensors (images [B,32,32,3], labels [B,classes] one-hot fp32) that are
refilled in place by one `rn_cifar_decode` launch per step — the analogue of the queue-runner /
tf.data iterator outputs.  `ResNet.train_op` pulls the next batch automatically (the dequeue is a
dependency of train_op in the reference's graph).  The whole record file set lives in HBM
(50,000 x 3,073 B = 154 MB).

There is no network in this environment, so `data_path` may be the literal string
"synthetic" or "synthetic:N" (N seeded pseudo-random records); real
`data_batch_*.bin` / `train.bin` files are read when they exist.
"""
import glob
import os

import numpy as np
import torch

from ._lib import lib, check, ptr, RN_F32, RN_BF16
from .flags import FLAGS


def synthetic_records(n, dataset="cifar10", seed=0):
    rng = np.random.Generator(np.random.PCG64(seed))
    if dataset == "cifar10":
        rec = rng.integers(0, 256, size=(n, 3073), dtype=np.uint8)
        rec[:, 0] = np.arange(n) % 10
    elif dataset == "cifar100":
        rec = rng.integers(0, 256, size=(n, 3074), dtype=np.uint8)
        rec[:, 0] = np.arange(n) % 20
        rec[:, 1] = np.arange(n) % 100
    else:
        raise ValueError('Not supported dataset %s' % dataset)
    return rec


def _load_records(dataset, data_path, record_bytes, seed):
    if isinstance(data_path, np.ndarray):
        return np.ascontiguousarray(data_path, dtype=np.uint8).reshape(-1, record_bytes)
    if data_path.startswith("synthetic"):
        n = int(data_path.split(":")[1]) if ":" in data_path else 50000
        return synthetic_records(n, dataset, seed)
    files = sorted(glob.glob(data_path))
    if not files and os.path.isdir(data_path):
        files = sorted(glob.glob(os.path.join(data_path, "*.bin")))
    if not files:
        raise IOError("no CIFAR record files match %r (use 'synthetic' for generated records)" % data_path)
    chunks = []
    for f in files:
        raw = np.fromfile(f, dtype=np.uint8)
        if raw.size % record_bytes:
            raise ValueError("%s: size %d is not a multiple of the record size %d" % (f, raw.size, record_bytes))
        chunks.append(raw.reshape(-1, record_bytes))
    return np.concatenate(chunks, 0)


class RecordPipeline(object):
    """Owns the device-resident records and the two output tensors."""

    def __init__(self, dataset, data_path, batch_size, train, pad, *, seed=0, dtype=None, device=None,
                 rank=0, world=1, shuffle=True, resident=True):
        if dataset == 'cifar10':
            self.label_offset, self.classes = 0, 10
        elif dataset == 'cifar100':
            self.label_offset, self.classes = 1, 100
        else:
            raise ValueError('Not supported dataset %s' % dataset)
        self.record_bytes = 1 + self.label_offset + 3072
        self.B = int(batch_size)
        self.train = bool(train)
        self.pad = int(pad)
        self.seed = int(seed)
        self.rank, self.world = rank, world
        self.shuffle = shuffle and self.train
        dev = torch.device(device if device is not None else "cuda:%d" % torch.cuda.current_device())
        host = _load_records(dataset, data_path, self.record_bytes, seed)
        self.n = host.shape[0]
        # resident: the whole record file lives in HBM (CIFAR: 154 MB).  Otherwise it stays in pinned host
        # memory and every batch's records are gathered on the host and copied H2D in front of the decode.
        self.resident = bool(resident)
        self.records = torch.from_numpy(host).to(dev) if self.resident else None
        self.host_records = None if self.resident else torch.from_numpy(host).pin_memory()
        self._stage_host = [torch.zeros(self.B, self.record_bytes, dtype=torch.uint8).pin_memory() for _ in range(4)]
        self._stage_dev = [torch.zeros(self.B, self.record_bytes, dtype=torch.uint8, device=dev) for _ in range(4)]
        self._stage_evt = [None] * 4
        self._stage = 0
        if dtype is None:
            dtype = torch.bfloat16 if FLAGS.compute_dtype == "bf16" else torch.float32
        self.images = torch.zeros(self.B, 32, 32, 3, dtype=dtype, device=dev)
        self.labels = torch.zeros(self.B, self.classes, dtype=torch.float32, device=dev)
        self.label_ids = torch.zeros(self.B, dtype=torch.int32, device=dev)
        self.draws = torch.zeros(self.B, 3, dtype=torch.int32, device=dev)
        self.indices = torch.zeros(self.B, dtype=torch.int32, device=dev)
        # ring of pinned staging buffers for the per-step index upload (no host sync per step)
        self._idx_host = [torch.zeros(self.B, dtype=torch.int32).pin_memory() for _ in range(8)]
        self._idx_evt = [None] * 8
        self._ring = 0
        self.epoch = 0
        self._cursor = 0
        self._order = self._epoch_order(0)
        self.images._rn_pipeline = self
        self.labels._rn_pipeline = self

    def _epoch_order(self, epoch):
        if self.shuffle:   # dataset.shuffle(buffer_size=50000): a full-epoch permutation
            order = np.random.Generator(np.random.PCG64([self.seed, epoch])).permutation(self.n)
        else:
            order = np.arange(self.n)
        return order[self.rank::self.world].astype(np.int32)   # rank r takes records r, r+R, ...

    def next_indices(self):
        out = np.empty(self.B, dtype=np.int32)
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
        """Position the stream as if `batches` batches had been consumed and decode the next one
        (checkpoint resume: the record order is a pure function of (seed, epoch), so replaying the
        index stream on the host reproduces exactly what the interrupted run would have read)."""
        self.epoch = 0
        self._cursor = 0
        self._order = self._epoch_order(0)
        for _ in range(int(batches)):
            self.next_indices()
        return self.next_batch()

    def next_batch(self, indices=None, draws=None):
        """Decode the next batch into self.images / self.labels (one H2D of B int32 + one launch)."""
        idx = self.next_indices() if indices is None else np.asarray(indices, dtype=np.int32)
        epoch = self.epoch
        st = torch.cuda.current_stream()
        slot = self._ring
        self._ring = (slot + 1) % len(self._idx_host)
        if self._idx_evt[slot] is not None:
            self._idx_evt[slot].synchronize()        # only blocks if the GPU is 8 steps behind
        self._idx_host[slot].copy_(torch.from_numpy(idx))
        self.indices.copy_(self._idx_host[slot], non_blocking=True)
        self._idx_evt[slot] = torch.cuda.Event()
        self._idx_evt[slot].record(st)
        din = None
        if draws is not None:
            self._draws_in = torch.as_tensor(np.asarray(draws, dtype=np.int32)).to(self.records.device)
            din = ptr(self._draws_in)
        odt = RN_BF16 if self.images.dtype == torch.bfloat16 else RN_F32
        if self.resident:
            rec, nrec, staged = self.records, self.n, 0
        else:
            k = self._stage
            self._stage = (k + 1) % len(self._stage_host)
            if self._stage_evt[k] is not None:
                self._stage_evt[k].synchronize()     # only blocks if the GPU is 4 batches behind the host
            torch.index_select(self.host_records, 0, torch.from_numpy(idx.astype(np.int64)), out=self._stage_host[k])
            self._stage_dev[k].copy_(self._stage_host[k], non_blocking=True)        # H2D, B x record_bytes
            self._stage_evt[k] = torch.cuda.Event()
            self._stage_evt[k].record(st)
            rec, nrec, staged = self._stage_dev[k], self.B, 1
        check(lib.rn_cifar_decode(ptr(rec), nrec, self.record_bytes, self.label_offset, self.classes,
                                  ptr(self.indices), self.B, int(self.train), self.pad,
                                  self.seed & 0xFFFFFFFF, (self.seed >> 32) & 0xFFFFFFFF, epoch & 0xFFFFFFFF, din,
                                  staged, ptr(self.images), odt, ptr(self.labels), ptr(self.label_ids), ptr(self.draws),
                                  st.cuda_stream), "rn_cifar_decode")
        return self.images, self.labels

    def set_resident(self, resident):
        """Move the record file between HBM and pinned host memory (same batches, same draws either way)."""
        resident = bool(resident)
        if resident == self.resident:
            return
        if resident:
            self.records = self.host_records.to(self.images.device)
            self.host_records = None
        else:
            self.host_records = self.records.cpu().pin_memory()
            self.records = None
        self.resident = resident


def _dist():
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


def build_input(dataset, data_path, batch_size, mode, **kw):
    """Reference cifar_input.py:25-119: train = pad to 36x36, random crop, flip, standardise;
    eval = standardise only."""
    rank, world = _dist()
    p = RecordPipeline(dataset, data_path, batch_size, mode == 'train', 2, rank=rank, world=world, **kw)
    p.next_batch()
    return p.images, p.labels


def input_fn(is_training, data_dir, batch_size, num_epochs=1, dataset='cifar10', **kw):
    """Reference resnet_cifar_train.py:204-247 (tf.data variant the trainers use): pad to 40x40."""
    if isinstance(data_dir, str) and not data_dir.startswith("synthetic"):
        sub = os.path.join(data_dir, 'cifar-10-batches-bin')
        base = sub if os.path.isdir(sub) else data_dir
        data_dir = os.path.join(base, 'data_batch_*.bin' if is_training else 'test_batch.bin')
    rank, world = _dist()
    p = RecordPipeline(dataset, data_dir, batch_size, is_training, 4, rank=rank, world=world, **kw)
    p.num_epochs = num_epochs
    p.next_batch()
    return p.images, p.labels
