// ImageNet input preprocessing on the GPU: the arithmetic of the reference's vgg_preprocessing.py
// (preprocess_for_train :284-314, preprocess_for_eval :317-333) applied to decoded uint8 images of
// arbitrary sizes that sit packed in one device buffer (JPEG decoding is not part of this path).
//
//   image/255 (convert_image_dtype, resnet_imagenet_train.py:141)
//   -> aspect-preserving bilinear resize, align_corners=False, smallest side = a random integer in
//      [side_min, side_max] (train) or side_min (eval)                    (_aspect_preserving_resize :259-281)
//   -> random (train) / central (eval) crop out_h x out_w                 (_random_crop :88-168, _central_crop :171-193)
//   -> random left-right flip (train)                                     (tf.image.random_flip_left_right :312)
//   -> per-channel mean subtraction, means 123.68/255, 116.78/255, 103.94/255   (:37-39, :196-226)
//
// Only the out_h x out_w window of the resized image is ever computed: an output pixel gathers its four
// source pixels straight from the uint8 image.  All float arithmetic is done with explicitly rounded
// single-precision operations in TensorFlow's order (no FMA contraction), so the pixels equal a
// float32 restatement bit for bit.  One CTA per (image, strip of output rows).
#include "rn_common.cuh"

namespace rn {

__device__ __forceinline__ void philox4x32_10_v(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0,
                                                uint32_t k1, uint32_t (&out)[4]) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
    const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
    const uint32_t n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
    c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
  out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

constexpr int VGG_THREADS = 256;
constexpr int VGG_ROWS = 8;         // output rows per CTA

struct VggGeom { int side, new_h, new_w, oy, ox, flip; };

// _smallest_size_at_least (:229-256): float32 scale = side / min(h, w); new sizes truncate h*scale, w*scale
__device__ __forceinline__ void vgg_resized_size(int h, int w, int side, int* nh, int* nw) {
  const float fh = (float)h, fw = (float)w, fs = (float)side;
  const float scale = h > w ? __fdiv_rn(fs, fw) : __fdiv_rn(fs, fh);
  *nh = (int)__fmul_rn(fh, scale);
  *nw = (int)__fmul_rn(fw, scale);
}

template <typename T>
__global__ void __launch_bounds__(VGG_THREADS)
vgg_preprocess_kernel(const uint8_t* __restrict__ images, const int64_t* __restrict__ offsets,
                      const int32_t* __restrict__ hw, int out_h, int out_w, int side_min, int side_max, int train,
                      uint32_t seed_lo, uint32_t seed_hi, uint32_t epoch, const int32_t* __restrict__ index,
                      const int32_t* __restrict__ draws_in, T* __restrict__ out, int32_t* __restrict__ draws_out,
                      float mean_r, float mean_g, float mean_b, int* __restrict__ error_flag,
                      const int32_t* __restrict__ labels, int classes, float* __restrict__ out_onehot) {
  const int b = blockIdx.x;
  const int h = hw[2 * b], w = hw[2 * b + 1];
  __shared__ VggGeom g;
  if (threadIdx.x == 0) {
    int side = side_min, oy, ox, flip = 0;
    int nh, nw;
    if (draws_in) {
      side = draws_in[4 * b];
      vgg_resized_size(h, w, side, &nh, &nw);
      oy = draws_in[4 * b + 1]; ox = draws_in[4 * b + 2]; flip = draws_in[4 * b + 3];
    } else if (train) {
      // Philox4x32-10, key (seed), counter (image id, epoch, 1, 0): word 0 -> side, 1 -> row offset, 2 -> column
      // offset (tf.random_uniform over an integer range is minval + u32 % range), 3 -> flip iff uniform < 0.5
      uint32_t r[4];
      philox4x32_10_v(index ? (uint32_t)index[b] : (uint32_t)b, epoch, 1u, 0u, seed_lo, seed_hi, r);
      side = side_min + (int)(r[0] % (uint32_t)(side_max - side_min + 1));
      vgg_resized_size(h, w, side, &nh, &nw);
      oy = nh >= out_h ? (int)(r[1] % (uint32_t)(nh - out_h + 1)) : 0;
      ox = nw >= out_w ? (int)(r[2] % (uint32_t)(nw - out_w + 1)) : 0;
      const float u = __uint_as_float((r[3] & 0x7fffffu) | 0x3f800000u) - 1.0f;
      flip = u < 0.5f ? 1 : 0;
    } else {
      vgg_resized_size(h, w, side, &nh, &nw);
      oy = (nh - out_h) / 2;          // _central_crop
      ox = (nw - out_w) / 2;
    }
    // the reference asserts 'Crop size greater than the image size.'
    if (nh < out_h || nw < out_w || oy < 0 || ox < 0 || oy + out_h > nh || ox + out_w > nw) {
      if (error_flag) atomicExch(error_flag, b + 1);
      oy = ox = 0;
    }
    g.side = side; g.new_h = nh; g.new_w = nw; g.oy = oy; g.ox = ox; g.flip = flip;
    if (draws_out && blockIdx.y == 0) {
      draws_out[4 * b] = side; draws_out[4 * b + 1] = oy; draws_out[4 * b + 2] = ox; draws_out[4 * b + 3] = flip;
    }
  }
  __syncthreads();
  if (out_onehot && blockIdx.y == 0) {          // tf.one_hot(label, _LABEL_CLASSES), resnet_imagenet_train.py:144-147
    const int label = labels[b];
    for (int k = threadIdx.x; k < classes; k += VGG_THREADS) out_onehot[(int64_t)b * classes + k] = (k == label) ? 1.f : 0.f;
  }
  const int nh = g.new_h, nw = g.new_w, oy = g.oy, ox = g.ox, flip = g.flip;
  // tf.image.resize_bilinear, align_corners=False (legacy scaler): in = out_index * (in_size / out_size)
  const float hs = __fdiv_rn((float)h, (float)nh), ws = __fdiv_rn((float)w, (float)nw);
  const uint8_t* img = images + offsets[b];
  const float inv255 = 1.0f / 255.0f;
  const int y0 = blockIdx.y * VGG_ROWS;
  const int total = VGG_ROWS * out_w;
  T* ob = out + (int64_t)b * out_h * out_w * 3;
  for (int i = threadIdx.x; i < total; i += VGG_THREADS) {
    const int yo = y0 + i / out_w, xo = i % out_w;
    if (yo >= out_h) break;
    const int yr = oy + yo;                                  // row / column in the resized image
    const int xr = ox + (flip ? out_w - 1 - xo : xo);
    const float in_y = __fmul_rn((float)yr, hs), in_x = __fmul_rn((float)xr, ws);
    const float fy = floorf(in_y), fx = floorf(in_x);
    // compute_interpolation_weights of TF 1.x: lower = (int)in, upper = min(lower + 1, in_size - 1), lerp = in - lower
    const int top = (int)fy, left = (int)fx;
    const int bot = min(top + 1, h - 1), right = min(left + 1, w - 1);
    const float ly = __fsub_rn(in_y, fy), lx = __fsub_rn(in_x, fx);
    const uint8_t* ptl = img + ((int64_t)top * w + left) * 3;
    const uint8_t* ptr_ = img + ((int64_t)top * w + right) * 3;
    const uint8_t* pbl = img + ((int64_t)bot * w + left) * 3;
    const uint8_t* pbr = img + ((int64_t)bot * w + right) * 3;
    const float mean[3] = {mean_r, mean_g, mean_b};
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const float tl = __fmul_rn((float)ptl[c], inv255), tr = __fmul_rn((float)ptr_[c], inv255);
      const float bl = __fmul_rn((float)pbl[c], inv255), br = __fmul_rn((float)pbr[c], inv255);
      const float tp = __fadd_rn(tl, __fmul_rn(__fsub_rn(tr, tl), lx));
      const float bt = __fadd_rn(bl, __fmul_rn(__fsub_rn(br, bl), lx));
      const float v = __fadd_rn(tp, __fmul_rn(__fsub_rn(bt, tp), ly));
      ob[((int64_t)yo * out_w + xo) * 3 + c] = from_f<T>(__fsub_rn(v, mean[c]));
    }
  }
}

}  // namespace rn
using namespace rn;

extern "C" int rn_vgg_preprocess(const uint8_t* images, const int64_t* offsets, const int32_t* hw, int32_t B,
                                 int32_t out_h, int32_t out_w, int32_t side_min, int32_t side_max, int32_t train,
                                 uint32_t seed_lo, uint32_t seed_hi, uint32_t epoch, const int32_t* index,
                                 const int32_t* draws_in, void* out, int32_t out_dtype, int32_t* draws_out,
                                 int32_t* error_flag, const int32_t* labels, int32_t classes, float* out_onehot,
                                 void* stream) {
  RN_REQUIRE(images && offsets && hw && out && B >= 0, "vgg_preprocess: NULL pointer or negative batch");
  RN_REQUIRE(out_h > 0 && out_w > 0 && side_min >= 1 && side_max >= side_min, "vgg_preprocess: bad sizes");
  RN_REQUIRE(side_min >= out_h && side_min >= out_w,
             "vgg_preprocess: resize side %d is smaller than the crop %dx%d (the reference asserts "
             "'Crop size greater than the image size.')", side_min, out_h, out_w);
  RN_REQUIRE(out_onehot == nullptr || (labels != nullptr && classes > 0), "vgg_preprocess: one-hot output needs labels and classes");
  if (B == 0) return RN_OK;
  // vgg_preprocessing.py:37-39 (python doubles rounded to float32 when TF subtracts them from the float image)
  const float mr = (float)(123.68 / 255), mg = (float)(116.78 / 255), mb = (float)(103.94 / 255);
  dim3 grid(B, (out_h + VGG_ROWS - 1) / VGG_ROWS);
  cudaStream_t st = as_stream(stream);
  if (out_dtype == RN_F32)
    vgg_preprocess_kernel<float><<<grid, VGG_THREADS, 0, st>>>(images, offsets, hw, out_h, out_w, side_min, side_max, train,
                                                              seed_lo, seed_hi, epoch, index, draws_in, (float*)out,
                                                              draws_out, mr, mg, mb, error_flag, labels, classes, out_onehot);
  else if (out_dtype == RN_BF16)
    vgg_preprocess_kernel<__nv_bfloat16><<<grid, VGG_THREADS, 0, st>>>(images, offsets, hw, out_h, out_w, side_min, side_max,
                                                                      train, seed_lo, seed_hi, epoch, index, draws_in,
                                                                      (__nv_bfloat16*)out, draws_out, mr, mg, mb, error_flag, labels, classes, out_onehot);
  else
    return fail(RN_ERR_INVALID, "vgg_preprocess: bad out_dtype %d", out_dtype);
  return check_launch("vgg_preprocess_kernel");
}
