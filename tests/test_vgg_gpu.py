"""rn_vgg_preprocess (ImageNet input preprocessing on the GPU) against the numpy float32 oracle of
vgg_preprocessing.py:284-333 on synthetic uint8 images of random sizes: the drawn resize side, crop
offsets and flip bits must be bit-exact; the pixels must agree to 1e-5 (they are computed with the same
float32 operations in the same order, so in practice they are identical)."""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from util import stream, lib, check, ptr, RN_F32, RN_BF16  # noqa: E402
from distributed_tensorflow_resnet_b200 import vgg_preprocessing as vp, FLAGS  # noqa: E402
from oracle import vgg_oracle as vo  # noqa: E402


def _images(n, seed):
    ims = vo.synthetic_images(n, seed)
    # edge cases: portrait, landscape, square, smallest side already 256, exact integer scale, tiny image (upscaled)
    rng = np.random.default_rng(seed + 100)
    for hw in ((500, 375), (375, 500), (300, 300), (256, 341), (128, 512), (64, 80), (224, 224), (1000, 257)):
        ims.append(rng.integers(0, 256, size=hw + (3,), dtype=np.uint8))
    return ims


def test_oracle_and_product_generate_the_same_synthetic_images():
    a, b = vo.synthetic_images(6, 3), vp.synthetic_images(6, 3)
    assert all(np.array_equal(x, y) for x, y in zip(a, b))


@pytest.mark.parametrize("epoch", [0, 3])
def test_train_preprocessing_matches_oracle(epoch):
    ims = _images(24, 1)
    seed = (5 << 32) | 11
    batch = vp.ImageBatch(ims)
    ids = np.arange(len(ims), dtype=np.int32)[::-1].copy()         # image ids key the draws, not the batch position
    out = vp.preprocess_for_train(batch, 224, 224, seed=seed, epoch=epoch, index=ids, dtype=torch.float32)
    torch.cuda.synchronize()
    want_d = vo.draws(ids, epoch, seed, [i.shape[0] for i in ims], [i.shape[1] for i in ims])
    got_d = out._rn_draws.cpu().numpy()
    assert np.array_equal(got_d, want_d)                            # side, crop offsets, flip: bit-exact
    assert got_d[:, 0].min() >= 256 and got_d[:, 0].max() <= 512 and set(got_d[:, 3]) == {0, 1}
    got = out.cpu().numpy()
    worst = 0.0
    for i, im in enumerate(ims):
        want = vo.preprocess(im, want_d[i], train=True)
        worst = max(worst, float(np.abs(got[i] - want).max()))
    assert worst <= 1e-5, worst
    # explicit draws reproduce the same pixels; bf16 output is the rounded fp32 output
    out2 = vp.preprocess_for_train(batch, 224, 224, draws=want_d, dtype=torch.bfloat16)
    assert torch.equal(out2.cpu(), out.cpu().to(torch.bfloat16))


def test_eval_preprocessing_matches_oracle():
    ims = _images(8, 2)
    out = vp.preprocess_image(vp.ImageBatch(ims), 224, 224, is_training=False, dtype=torch.float32)
    torch.cuda.synchronize()
    got = out.cpu().numpy()
    for i, im in enumerate(ims):
        want = vo.preprocess(im, train=False, resize_side=256)
        assert np.abs(got[i] - want).max() <= 1e-5
    d = out._rn_draws.cpu().numpy()
    assert (d[:, 0] == 256).all() and (d[:, 3] == 0).all()


def test_argument_errors_and_small_images():
    from distributed_tensorflow_resnet_b200._lib import last_error
    ims = _images(2, 3)
    b = vp.ImageBatch(ims)
    out = torch.empty(b.n, 224, 224, 3, device="cuda")
    # a resize side below the crop size is the reference's 'Crop size greater than the image size.'
    rc = lib.rn_vgg_preprocess(ptr(b.data), ptr(b.offsets), ptr(b.hw), b.n, 224, 224, 200, 512, 1, 0, 0, 0, None, None,
                               ptr(out), RN_F32, None, None, None, 0, None, stream())
    assert rc == -1 and "Crop size" in last_error()
    assert lib.rn_vgg_preprocess(None, ptr(b.offsets), ptr(b.hw), b.n, 224, 224, 256, 512, 1, 0, 0, 0, None, None, ptr(out),
                                 RN_F32, None, None, None, 0, None, stream()) == -1
    # explicit draws that push the crop outside the resized image are reported through the error flag
    bad = np.array([[256, 400, 0, 0]] * b.n, dtype=np.int32)
    with pytest.raises(ValueError, match="Crop size"):
        vp.preprocess_for_train(b, 224, 224, draws=bad, dtype=torch.float32)
    # empty batch
    check(lib.rn_vgg_preprocess(ptr(b.data), ptr(b.offsets), ptr(b.hw), 0, 224, 224, 256, 512, 1, 0, 0, 0, None, None,
                                ptr(out), RN_F32, None, None, None, 0, None, stream()), "empty")


@pytest.mark.parametrize("resident", [True, False])
def test_imagenet_pipeline_batches_match_oracle(resident):
    """The synthetic input_fn: per-epoch permutation, rank sharding, labels one-hot, images from the kernel — the
    HBM-resident pool and the pinned-host pool (H2D of the raw images every step) give identical batches."""
    FLAGS.reset()
    B, pool, seed = 8, 20, 4
    ims = vp.synthetic_images(pool, seed)
    p = vp.ImagenetPipeline(B, pool=pool, seed=seed, dtype=torch.float32, resident=resident, rank=1, world=2)
    shadow_order = lambda e: np.random.Generator(np.random.PCG64([seed, e])).permutation(pool)[1::2]   # noqa: E731
    cur, cursor, epoch = shadow_order(0), 0, 0
    for step in range(4):
        p.next_batch()
        torch.cuda.synchronize()
        idx = []
        while len(idx) < B:
            take = min(B - len(idx), len(cur) - cursor)
            idx += list(cur[cursor:cursor + take])
            cursor += take
            if cursor >= len(cur):
                epoch += 1
                cur, cursor = shadow_order(epoch), 0
        d = vo.draws(np.array(idx), epoch, seed, [ims[i].shape[0] for i in idx], [ims[i].shape[1] for i in idx])
        assert np.array_equal(p.draws.cpu().numpy(), d), step
        got = p.images.cpu().numpy()
        for j, i in enumerate(idx):
            assert np.abs(got[j] - vo.preprocess(ims[i], d[j])).max() <= 1e-5
        lab = p.labels.cpu().numpy()
        assert (lab.sum(1) == 1).all() and np.array_equal(lab.argmax(1), p.label_ids[idx])
        assert p.h2d_bytes_last > (0 if resident else sum(ims[i].size for i in idx))


def test_full_batch_properties():
    """BASELINE size (128 images per step): every output finite, inside [-mean, 1 - mean] per channel (a convex
    combination of pixels / 255 minus the channel mean), and a left-right flip of the draws mirrors the output."""
    ims = vp.synthetic_images(128, 9)
    b = vp.ImageBatch(ims)
    out = vp.preprocess_for_train(b, 224, 224, seed=1, dtype=torch.float32)
    d = out._rn_draws.cpu().numpy().copy()
    x = out.cpu().numpy()
    assert np.isfinite(x).all()
    for c in range(3):
        assert x[..., c].min() >= -float(vo.MEANS[c]) - 1e-6 and x[..., c].max() <= 1.0 - float(vo.MEANS[c]) + 1e-6
    d[:, 3] ^= 1
    mirrored = vp.preprocess_for_train(b, 224, 224, draws=d, dtype=torch.float32).cpu().numpy()
    assert np.array_equal(mirrored, x[:, :, ::-1])
