// CIFAR binary-record decode + pad / random crop / random flip / per-image standardisation.
// One warp per image (eight images per CTA): the record is staged in shared memory with 16-byte
// loads, the crop/flip is a shared-memory gather, the statistics are exact integer sums
// (sum <= 783,360 and sum of squares <= 199,756,800 both fit int32), and the HWC output is
// written with 16-byte stores.  Algorithmic HBM bytes per image: record_bytes read +
// 12,288 (fp32) or 6,144 (bf16) + 4*classes written.
#include "rn_common.cuh"

namespace rn {

__device__ __forceinline__ void philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0,
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

constexpr int DEC_WARPS = 8;                 // images per CTA: one warp decodes one image
constexpr int DEC_THREADS = DEC_WARPS * 32;
constexpr int DEC_SLOT = 3104;               // bytes of shared memory per image: 3074-byte record + up to 15 bytes of lead-in, 16-aligned

// One WARP per image.  The record (3073 / 3074 bytes at an arbitrary byte address) is copied into shared memory with
// 16-byte loads of the aligned chunks that enclose it (single bytes at the very end of the record buffer, where the
// enclosing chunk would reach past the allocation); pad / crop / flip is a gather from shared memory; the statistics
// are exact int32 sums (sum <= 783,360; sum of squares <= 199,756,800) reduced with warp shuffles; each lane then
// writes 8 consecutive output elements at a time (one or two 16-byte stores).  No CTA-wide barrier.
template <typename T>
__global__ void __launch_bounds__(DEC_THREADS)
cifar_decode_kernel(const uint8_t* __restrict__ records, int64_t n_records, int record_bytes, int label_offset,
                    int classes, const int32_t* __restrict__ indices, int B, int train, int pad, uint32_t seed_lo,
                    uint32_t seed_hi, uint32_t epoch, const int32_t* __restrict__ draws_in, int staged,
                    T* __restrict__ out_img, float* __restrict__ out_onehot, int32_t* __restrict__ out_labels,
                    int32_t* __restrict__ draws_out) {
  __shared__ __align__(16) uint8_t spx[DEC_WARPS][DEC_SLOT];
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  const int b = blockIdx.x * DEC_WARPS + warp;
  if (b >= B) return;                                    // whole warps leave together; nothing below synchronises the CTA
  const int64_t rec = indices ? (int64_t)indices[b] : (int64_t)b;
  // staged: `records` holds just this batch (row b), copied from host memory; `rec` still keys the random draws
  const uint8_t* base = records + (staged ? (int64_t)b : rec) * record_bytes;
  const uint8_t* end = records + n_records * (int64_t)record_bytes;
  const int mis = (int)(reinterpret_cast<uintptr_t>(base) & 15);
  const uint8_t* al = base - mis;
  const int chunks = (mis + record_bytes + 15) / 16;
  uint8_t* sp = spx[warp];
  for (int c = lane; c < chunks; c += 32) {
    const uint8_t* src = al + 16 * c;
    if (src + 16 <= end) {
      *reinterpret_cast<uint4*>(sp + 16 * c) = __ldg(reinterpret_cast<const uint4*>(src));
    } else {
      for (int i = 0; i < 16; ++i) sp[16 * c + i] = (src + i < end) ? src[i] : (uint8_t)0;
    }
  }
  int oy = pad, ox = pad, flip = 0;  // identity window when not training
  if (train) {
    if (draws_in) {
      oy = draws_in[b * 3]; ox = draws_in[b * 3 + 1]; flip = draws_in[b * 3 + 2];
    } else {
      uint32_t r[4];
      philox4x32_10((uint32_t)rec, epoch, 0u, 0u, seed_lo, seed_hi, r);
      const uint32_t span = 2u * (uint32_t)pad + 1u;
      oy = (int)((r[0] % 2147483647u) % span);
      ox = (int)((r[1] % 2147483647u) % span);
      const float u = __uint_as_float((r[3] & 0x7fffffu) | 0x3f800000u) - 1.0f;
      flip = u < 0.5f ? 1 : 0;
    }
  }
  __syncwarp();
  const uint8_t* px = sp + mis + label_offset + 1;        // CHW planar pixels
  const int label = sp[mis + label_offset];
  // output element e of the [32][32][3] HWC image: c = e % 3, x = (e / 3) % 32, y = e / 96
  auto pixel = [&](int e) -> int {
    const int c = e % 3, xo = (e / 3) % 32, yo = e / 96;
    const int xs = flip ? 31 - xo : xo;
    const int hy = yo + oy - pad, wx = xs + ox - pad;      // position in the un-padded image
    return (hy >= 0 && hy < 32 && wx >= 0 && wx < 32) ? (int)px[c * 1024 + hy * 32 + wx] : 0;
  };
  // gather once: the 96 values of this lane stay in 24 registers (four bytes each) for the output pass
  uint32_t pk[24];
  int s = 0, q = 0;
#pragma unroll
  for (int j = 0; j < 12; ++j) {
    const int e0 = (lane + 32 * j) * 8;
    uint32_t lo = 0, hi = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int v = pixel(e0 + i);
      s += v;
      q += v * v;
      if (i < 4) lo |= (uint32_t)v << (8 * i); else hi |= (uint32_t)v << (8 * (i - 4));
    }
    pk[2 * j] = lo;
    pk[2 * j + 1] = hi;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    s += __shfl_xor_sync(0xffffffffu, s, o);
    q += __shfl_xor_sync(0xffffffffu, q, o);
  }
  // per_image_standardization: mean, var = relu(E[x^2] - mean^2), divide by max(sqrt(var), 1/sqrt(3072));
  // (v - mean) / adj == (3072 v - S) / (3072 adj): the numerator is an exact integer (< 2^24)
  const double mean = (double)s / 3072.0;
  double var = (double)q / 3072.0 - mean * mean;
  if (var < 0.0) var = 0.0;
  double adj = sqrt(var);
  const double min_adj = 1.0 / sqrt(3072.0);
  if (adj < min_adj) adj = min_adj;
  const float scale = (float)(1.0 / (3072.0 * adj));
  if (lane == 0) {
    if (out_labels) out_labels[b] = label;
    if (draws_out) { draws_out[b * 3] = oy; draws_out[b * 3 + 1] = ox; draws_out[b * 3 + 2] = flip; }
  }
  T* o = out_img + (int64_t)b * 3072;
#pragma unroll
  for (int j = 0; j < 12; ++j) {
    const int e0 = (lane + 32 * j) * 8;
    float f[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int v = (int)((pk[2 * j + (i >> 2)] >> (8 * (i & 3))) & 0xffu);
      f[i] = (float)(v * 3072 - s) * scale;
    }
    if (sizeof(T) == 2) {
      Vec<__nv_bfloat16> v;
      v.pack(f);
      v.store(reinterpret_cast<__nv_bfloat16*>(o) + e0);
    } else {
      float* of = reinterpret_cast<float*>(o) + e0;
      *reinterpret_cast<float4*>(of) = make_float4(f[0], f[1], f[2], f[3]);
      *reinterpret_cast<float4*>(of + 4) = make_float4(f[4], f[5], f[6], f[7]);
    }
  }
  if (out_onehot)
    for (int k = lane; k < classes; k += 32) out_onehot[(int64_t)b * classes + k] = (k == label) ? 1.f : 0.f;
}

}  // namespace rn
using namespace rn;

extern "C" int rn_cifar_decode(const uint8_t* records, int64_t n_records, int32_t record_bytes, int32_t label_offset,
                               int32_t classes, const int32_t* indices, int32_t B, int32_t train, int32_t pad,
                               uint32_t seed_lo, uint32_t seed_hi, uint32_t epoch, const int32_t* draws_in,
                               int32_t staged, void* out_img, int32_t out_dtype, float* out_onehot,
                               int32_t* out_labels, int32_t* draws_out, void* stream) {
  RN_REQUIRE(records && out_img && B >= 0 && n_records > 0, "cifar_decode: bad arguments");
  RN_REQUIRE(record_bytes == 1 + label_offset + 3072 && (label_offset == 0 || label_offset == 1),
             "cifar_decode: record_bytes=%d label_offset=%d is not a CIFAR record", record_bytes, label_offset);
  RN_REQUIRE(pad >= 0 && pad <= 16 && classes > 0, "cifar_decode: bad pad/classes");
  RN_REQUIRE(indices != nullptr || B <= n_records, "cifar_decode: B > n_records without indices");
  RN_REQUIRE(!staged || B <= n_records, "cifar_decode: a staged batch must hold B records");
  if (B == 0) return RN_OK;
  cudaStream_t st = as_stream(stream);
  if (out_dtype == RN_F32)
    cifar_decode_kernel<float><<<(B + DEC_WARPS - 1) / DEC_WARPS, DEC_THREADS, 0, st>>>(
        records, n_records, record_bytes, label_offset, classes, indices, B, train, pad, seed_lo, seed_hi, epoch, draws_in,
                                                         staged, (float*)out_img, out_onehot, out_labels, draws_out);
  else if (out_dtype == RN_BF16)
    cifar_decode_kernel<__nv_bfloat16><<<(B + DEC_WARPS - 1) / DEC_WARPS, DEC_THREADS, 0, st>>>(
        records, n_records, record_bytes, label_offset, classes, indices, B, train, pad, seed_lo, seed_hi, epoch,
                                                                 draws_in, staged, (__nv_bfloat16*)out_img, out_onehot,
                                                                 out_labels, draws_out);
  else
    return fail(RN_ERR_INVALID, "cifar_decode: bad out_dtype %d", out_dtype);
  return check_launch("cifar_decode_kernel");
}
