"""CPU oracle (numpy) of the reference's CIFAR binary-record input path.

TEST INFRASTRUCTURE — see oracle/__init__.py.  Follows
  * cifar_input.py:25-119  (`build_input`: label_offset 0/1, pad to 36x36, crop, flip,
    per-image standardisation, dense one-hot), and
  * resnet_cifar_train.py:158-247 (`parse_record` / `preprocess_image`: pad to 40x40), whose exact
    TF op chain is pinned by the tf.data function bodies in tests/golden/cifar50_train_graph.json:
      DecodeRaw(uint8) -> rec[0] -> int32 -> OneHot(depth, 1.0, 0.0)
      rec[1:3073] -> [3,32,32] -> transpose [1,2,0] -> float32
      Pad 4/4 -> 40x40 ; offset = RandomUniformInt([3], 0, 2^31-1) floormod [9,9,1] ; Slice 32x32x3
      u = RandomUniform[0,1) ; flip W iff u < 0.5
      mean = Mean(x) ; var = relu(Mean(x^2) - mean^2) ; y = (x - mean) / max(sqrt(var), rsqrt(3072))

Randomness: the fixture's seeds are (0, 0) = non-deterministic in TF, so "bit-exact crop/flip"
means: given the same raw 32-bit draws, the same offsets and flip bits.  The raw draws are
specified here as Philox4x32-10 with key (seed_lo, seed_hi) and counter
(record_index, epoch, 0, 0): words 0..2 feed RandomUniformInt (TF: lo + u32 % (hi - lo)), word 3
feeds RandomUniform (TF Uint32ToFloat: bitcast((u32 & 0x7fffff) | 0x3f800000) - 1.0).
"""
import numpy as np

PHILOX_M0 = np.uint64(0xD2511F53)
PHILOX_M1 = np.uint64(0xCD9E8D57)
PHILOX_W0 = np.uint32(0x9E3779B9)
PHILOX_W1 = np.uint32(0xBB67AE85)
MASK32 = np.uint64(0xFFFFFFFF)


def philox4x32_10(counter, key):
    """counter: uint32 [..., 4]; key: (k0, k1) -> uint32 [..., 4].  Standard Philox4x32-10."""
    c = np.array(counter, dtype=np.uint32, copy=True)
    c0, c1, c2, c3 = (c[..., i].astype(np.uint64) for i in range(4))
    k0 = np.uint32(key[0])
    k1 = np.uint32(key[1])
    with np.errstate(over="ignore"):
        for _ in range(10):
            p0 = PHILOX_M0 * c0
            p1 = PHILOX_M1 * c2
            hi0, lo0 = p0 >> np.uint64(32), p0 & MASK32
            hi1, lo1 = p1 >> np.uint64(32), p1 & MASK32
            n0 = (hi1 ^ c1 ^ np.uint64(k0)) & MASK32
            n1 = lo1
            n2 = (hi0 ^ c3 ^ np.uint64(k1)) & MASK32
            n3 = lo0
            c0, c1, c2, c3 = n0, n1, n2, n3
            k0 = np.uint32(k0 + PHILOX_W0)
            k1 = np.uint32(k1 + PHILOX_W1)
    return np.stack([c0, c1, c2, c3], axis=-1).astype(np.uint32)


def raw_draws(record_index, epoch, seed):
    ctr = np.zeros((len(record_index), 4), dtype=np.uint32)
    ctr[:, 0] = np.asarray(record_index, dtype=np.uint32)
    ctr[:, 1] = np.uint32(epoch)
    return philox4x32_10(ctr, (seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF))


def draws_to_indices(raw, pad):
    """raw uint32 [B,4] -> (crop_y, crop_x, flip) exactly as the fixture's op chain."""
    span = np.uint32(2 * pad + 1)                       # random_crop/add: shape - size + 1
    r = raw[:, :3] % np.uint32(2147483647)              # RandomUniformInt(0, 2^31-1)
    oy = (r[:, 0] % span).astype(np.int32)              # FloorMod [span, span, 1]
    ox = (r[:, 1] % span).astype(np.int32)
    u = ((raw[:, 3] & np.uint32(0x7FFFFF)) | np.uint32(0x3F800000)).view(np.float32) - np.float32(1.0)
    flip = (u < np.float32(0.5)).astype(np.uint8)
    return oy, ox, flip


def decode(records, record_index, *, dataset="cifar10", train=True, pad=4, epoch=0, seed=0,
           crop_yx=None, flip=None):
    """records: uint8 [n_records, record_bytes]; record_index: which records form the batch.
    Returns images float32 [B,32,32,3], one-hot float32 [B,classes], labels, (oy, ox, flip)."""
    if dataset == "cifar10":
        label_offset, classes = 0, 10
    elif dataset == "cifar100":
        label_offset, classes = 1, 100
    else:
        raise ValueError("Not supported dataset %s" % dataset)
    record_bytes = 1 + label_offset + 3072
    records = np.asarray(records, dtype=np.uint8).reshape(-1, record_bytes)
    idx = np.asarray(record_index, dtype=np.int64)
    B = len(idx)
    rec = records[idx]
    labels = rec[:, label_offset].astype(np.int32)
    img = rec[:, label_offset + 1:].reshape(B, 3, 32, 32).transpose(0, 2, 3, 1).astype(np.float32)
    if train:
        if crop_yx is None:
            oy, ox, fl = draws_to_indices(raw_draws(idx, epoch, seed), pad)
        else:
            oy, ox = np.asarray(crop_yx[0], np.int32), np.asarray(crop_yx[1], np.int32)
            fl = np.asarray(flip, np.uint8)
        padded = np.zeros((B, 32 + 2 * pad, 32 + 2 * pad, 3), dtype=np.float32)
        padded[:, pad:pad + 32, pad:pad + 32] = img
        out = np.empty((B, 32, 32, 3), dtype=np.float32)
        for b in range(B):
            c = padded[b, oy[b]:oy[b] + 32, ox[b]:ox[b] + 32]
            out[b] = c[:, ::-1] if fl[b] else c
        img = out
    else:
        oy = ox = np.zeros(B, np.int32)
        fl = np.zeros(B, np.uint8)
    flat = img.reshape(B, -1).astype(np.float64)
    mean = flat.mean(1)
    var = np.maximum((flat * flat).mean(1) - mean * mean, 0.0)
    adj = np.maximum(np.sqrt(var), 1.0 / np.sqrt(3072.0))
    y = ((img.astype(np.float64) - mean[:, None, None, None]) / adj[:, None, None, None]).astype(np.float32)
    onehot = np.zeros((B, classes), dtype=np.float32)
    onehot[np.arange(B), labels] = 1.0
    return y, onehot, labels, (oy, ox, fl)


def synthetic_records(n, dataset="cifar10", seed=0):
    """PCG64(seed) uniform uint8 pixels; label byte i % 10 (CIFAR-100: coarse i%20,
    fine i%100)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    if dataset == "cifar10":
        rec = rng.integers(0, 256, size=(n, 3073), dtype=np.uint8)
        rec[:, 0] = np.arange(n) % 10
    else:
        rec = rng.integers(0, 256, size=(n, 3074), dtype=np.uint8)
        rec[:, 0] = np.arange(n) % 20
        rec[:, 1] = np.arange(n) % 100
    return rec
