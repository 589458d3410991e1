"""CPU oracle (numpy, float32) of the reference's ImageNet preprocessing.

TEST INFRASTRUCTURE — see oracle/__init__.py.  Restates vgg_preprocessing.py of the reference:
  * preprocess_for_train :284-314 — resize_side = random_uniform([], 256, 512 + 1, int32);
    _aspect_preserving_resize; _random_crop([image], 224, 224); to_float; random_flip_left_right;
    _mean_image_subtraction with [_R_MEAN, _G_MEAN, _B_MEAN] = [123.68, 116.78, 103.94] / 255 (:37-39)
  * preprocess_for_eval :317-333 — resize to resize_side (256), _central_crop, mean subtraction
  * _smallest_size_at_least :229-256 — float32 scale = side / width if height > width else side / height;
    new_height = to_int32(height * scale), new_width = to_int32(width * scale)
  * _aspect_preserving_resize :259-281 — tf.image.resize_bilinear(align_corners=False)
as called from record_parser (resnet_imagenet_train.py:137-147) on
tf.image.convert_image_dtype(decode_image(...), tf.float32), i.e. uint8 * float32(1/255).

tf.image.resize_bilinear (TF 1.x CPU kernel, align_corners=False) is third-party code absent from the
reference tree; its published algorithm (tensorflow/core/kernels/resize_bilinear_op.cc,
compute_interpolation_weights + compute_lerp) is restated in `resize_bilinear`:
  scale = in_size / float(out_size); in = i * scale; lower = int(in); upper = min(lower + 1, in_size - 1);
  lerp = in - lower;  top = tl + (tr - tl) * x_lerp; bottom = bl + (br - bl) * x_lerp; out = top + (bottom - top) * y_lerp
all in float32.  The reference holds no golden images for this path: parity unpinned numerically.

Random draws (the reference uses TF's unseeded generator): Philox4x32-10 keyed by the seed with counter
(image id, epoch, 1, 0) — the same specification the CUDA kernel implements (include/rn_b200.h).
"""
import numpy as np

from .cifar_oracle import philox4x32_10

MEANS = np.array([123.68 / 255, 116.78 / 255, 103.94 / 255], dtype=np.float32)
F = np.float32


def synthetic_images(n, seed=0, lo=240, hi=520):
    """n decoded uint8 images of random sizes (height, width in [lo, hi]), PCG64(seed)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    sizes = rng.integers(lo, hi + 1, size=(n, 2))
    return [rng.integers(0, 256, size=(int(h), int(w), 3), dtype=np.uint8) for h, w in sizes]


def smallest_size_at_least(height, width, smallest_side):
    h, w, s = F(height), F(width), F(smallest_side)
    scale = s / w if height > width else s / h          # float32 division
    return int(h * scale), int(w * scale)                # float32 product, truncation


def draws(image_ids, epoch, seed, heights, widths, out_h=224, out_w=224, side_min=256, side_max=512):
    """(side, offset_y, offset_x, flip) per image, from the raw Philox words."""
    ctr = np.zeros((len(image_ids), 4), dtype=np.uint32)
    ctr[:, 0] = np.asarray(image_ids, dtype=np.uint32)
    ctr[:, 1] = np.uint32(epoch)
    ctr[:, 2] = 1
    raw = philox4x32_10(ctr, (seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF))
    out = np.zeros((len(image_ids), 4), dtype=np.int32)
    for i in range(len(image_ids)):
        side = side_min + int(raw[i, 0] % np.uint32(side_max - side_min + 1))
        nh, nw = smallest_size_at_least(heights[i], widths[i], side)
        oy = int(raw[i, 1] % np.uint32(nh - out_h + 1))
        ox = int(raw[i, 2] % np.uint32(nw - out_w + 1))
        u = ((raw[i, 3] & np.uint32(0x7FFFFF)) | np.uint32(0x3F800000)).view(np.float32) - F(1.0)
        out[i] = (side, oy, ox, int(u < F(0.5)))
    return out


def resize_bilinear(img, out_h, out_w):
    """img float32 [H, W, C] -> [out_h, out_w, C], align_corners=False, float32 arithmetic in TF's order."""
    h, w, _ = img.shape
    hs, ws = F(h) / F(out_h), F(w) / F(out_w)
    in_y = np.arange(out_h, dtype=np.float32) * hs
    in_x = np.arange(out_w, dtype=np.float32) * ws
    top = in_y.astype(np.int64)
    left = in_x.astype(np.int64)
    bot = np.minimum(top + 1, h - 1)
    right = np.minimum(left + 1, w - 1)
    ly = (in_y - top.astype(np.float32))[:, None, None]
    lx = (in_x - left.astype(np.float32))[None, :, None]
    tl, tr = img[top][:, left], img[top][:, right]
    bl, br = img[bot][:, left], img[bot][:, right]
    tp = tl + (tr - tl) * lx
    bt = bl + (br - bl) * lx
    return (tp + (bt - tp) * ly).astype(np.float32)


def preprocess(image_u8, draw=None, train=True, out_h=224, out_w=224, resize_side=256):
    """One image.  draw = (side, offset_y, offset_x, flip) for the training path."""
    img = image_u8.astype(np.float32) * F(1.0 / 255.0)              # convert_image_dtype
    h, w, _ = img.shape
    side = int(draw[0]) if train else resize_side
    nh, nw = smallest_size_at_least(h, w, side)
    res = resize_bilinear(img, nh, nw)
    if train:
        oy, ox, flip = int(draw[1]), int(draw[2]), int(draw[3])
    else:
        oy, ox, flip = (nh - out_h) // 2, (nw - out_w) // 2, 0
    if nh < out_h or nw < out_w:
        raise ValueError("Crop size greater than the image size.")
    crop = res[oy:oy + out_h, ox:ox + out_w]
    if flip:
        crop = crop[:, ::-1]
    return (crop - MEANS[None, None, :]).astype(np.float32)
