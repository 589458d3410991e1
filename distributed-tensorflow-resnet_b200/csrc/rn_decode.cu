// CIFAR binary-record decode + pad / random crop / random flip / per-image standardisation.
// One CTA per image: the 3072 pixel bytes are staged in shared memory with coalesced byte
// loads, the crop/flip is a shared-memory gather, the statistics are exact integer sums
// (sum <= 783,360 and sum of squares <= 199,756,800 both fit int32), and the HWC output is
// written with contiguous stores.  Algorithmic HBM bytes per image: record_bytes read +
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

constexpr int DEC_THREADS = 256;

template <typename T>
__global__ void __launch_bounds__(DEC_THREADS)
cifar_decode_kernel(const uint8_t* __restrict__ records, int64_t n_records, int record_bytes, int label_offset,
                    int classes, const int32_t* __restrict__ indices, int train, int pad, uint32_t seed_lo,
                    uint32_t seed_hi, uint32_t epoch, const int32_t* __restrict__ draws_in, int staged,
                    T* __restrict__ out_img, float* __restrict__ out_onehot, int32_t* __restrict__ out_labels,
                    int32_t* __restrict__ draws_out) {
  __shared__ uint8_t px[3072];
  __shared__ int red[2][DEC_THREADS / 32];
  __shared__ float s_scale;
  __shared__ int s_sum;
  const int b = blockIdx.x;
  const int64_t rec = indices ? (int64_t)indices[b] : (int64_t)b;
  // staged: `records` holds just this batch (row b), copied from host memory; `rec` still keys the random draws
  const uint8_t* base = records + (staged ? (int64_t)b : rec) * record_bytes;
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
  for (int i = threadIdx.x; i < 3072; i += DEC_THREADS) px[i] = base[label_offset + 1 + i];
  const int label = base[label_offset];
  __syncthreads();
  // each thread owns output elements e = threadIdx.x + 256*j of the [32][32][3] HWC image
  int vals[12];
  int s = 0, q = 0;
#pragma unroll
  for (int j = 0; j < 12; ++j) {
    const int e = threadIdx.x + DEC_THREADS * j;
    const int c = e % 3, xo = (e / 3) % 32, yo = e / 96;
    const int xs = flip ? 31 - xo : xo;
    const int hy = yo + oy - pad, wx = xs + ox - pad;  // position in the un-padded image
    int v = 0;
    if (hy >= 0 && hy < 32 && wx >= 0 && wx < 32) v = px[c * 1024 + hy * 32 + wx];
    vals[j] = v;
    s += v;
    q += v * v;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    s += __shfl_xor_sync(0xffffffffu, s, o);
    q += __shfl_xor_sync(0xffffffffu, q, o);
  }
  if (threadIdx.x % 32 == 0) { red[0][threadIdx.x / 32] = s; red[1][threadIdx.x / 32] = q; }
  __syncthreads();
  if (threadIdx.x == 0) {
    int S = 0, Q = 0;
    for (int i = 0; i < DEC_THREADS / 32; ++i) { S += red[0][i]; Q += red[1][i]; }
    // per_image_standardization: mean, var = relu(E[x^2] - mean^2), divide by max(sqrt(var), 1/sqrt(3072))
    const double mean = (double)S / 3072.0;
    double var = (double)Q / 3072.0 - mean * mean;
    if (var < 0.0) var = 0.0;
    double adj = sqrt(var);
    const double min_adj = 1.0 / sqrt(3072.0);
    if (adj < min_adj) adj = min_adj;
    // (v - mean) / adj == (3072 v - S) / (3072 adj): the numerator is an exact integer (< 2^24)
    s_sum = S;
    s_scale = (float)(1.0 / (3072.0 * adj));
    if (out_labels) out_labels[b] = label;
    if (draws_out) { draws_out[b * 3] = oy; draws_out[b * 3 + 1] = ox; draws_out[b * 3 + 2] = flip; }
  }
  __syncthreads();
  const int S = s_sum;
  const float scale = s_scale;
  T* o = out_img + (int64_t)b * 3072;
#pragma unroll
  for (int j = 0; j < 12; ++j) o[threadIdx.x + DEC_THREADS * j] = from_f<T>((float)(vals[j] * 3072 - S) * scale);
  if (out_onehot)
    for (int k = threadIdx.x; k < classes; k += DEC_THREADS) out_onehot[(int64_t)b * classes + k] = (k == label) ? 1.f : 0.f;
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
    cifar_decode_kernel<float><<<B, DEC_THREADS, 0, st>>>(records, n_records, record_bytes, label_offset, classes,
                                                         indices, train, pad, seed_lo, seed_hi, epoch, draws_in,
                                                         staged, (float*)out_img, out_onehot, out_labels, draws_out);
  else if (out_dtype == RN_BF16)
    cifar_decode_kernel<__nv_bfloat16><<<B, DEC_THREADS, 0, st>>>(records, n_records, record_bytes, label_offset,
                                                                 classes, indices, train, pad, seed_lo, seed_hi, epoch,
                                                                 draws_in, staged, (__nv_bfloat16*)out_img, out_onehot,
                                                                 out_labels, draws_out);
  else
    return fail(RN_ERR_INVALID, "cifar_decode: bad out_dtype %d", out_dtype);
  return check_launch("cifar_decode_kernel");
}
