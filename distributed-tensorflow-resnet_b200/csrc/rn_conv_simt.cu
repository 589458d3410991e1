// CUDA-core (FFMA) implicit-GEMM convolution: fprop, dgrad, wgrad.
// This is the fp32 VERIFICATION path (activations/gradients within 1e-4 of the CPU oracle) and
// the fallback for shapes the tcgen05 path does not take.  fp32 accumulation for both storage
// types.  64x64 output tile, 16-deep K slices through shared memory, 4x4 register micro-tile.
#include "rn_common.cuh"

namespace rn {

struct ConvGeom {
  int N, H, W, Cin, Ho, Wo, Cout, k, stride, pad;
};

constexpr int CT = 256;

// fprop : out[m=(n,ho,wo)][co]  = sum_{kk=(r,s,ci)} in(n, ho*st-pad+r, wo*st-pad+s, ci) * w[kk][co]
// dgrad : out[m=(n,h,w)][ci]    = sum_{kk=(r,s,co)} dy(n, (h+pad-r)/st, (w+pad-s)/st, co) * w[(r,s)][ci][co]
template <typename T, bool DGRAD>
__global__ void __launch_bounds__(CT)
conv_igemm_kernel(ConvGeom g, const T* __restrict__ in, const T* __restrict__ w, T* out, float beta) {
  __shared__ float As[16][64 + 4];
  __shared__ float Bs[16][64 + 4];
  const int tid = threadIdx.x;
  const int OH = DGRAD ? g.H : g.Ho, OW = DGRAD ? g.W : g.Wo;       // rows of the GEMM
  const int IH = DGRAD ? g.Ho : g.H, IW = DGRAD ? g.Wo : g.W;       // gathered tensor
  const int CI = DGRAD ? g.Cout : g.Cin;                            // reduction channels
  const int CO = DGRAD ? g.Cin : g.Cout;                            // GEMM N
  const int64_t M = (int64_t)g.N * OH * OW;
  const int Kd = g.k * g.k * CI;
  const int64_t m0 = (int64_t)blockIdx.x * 64;
  const int n0 = blockIdx.y * 64;
  // A loader: this thread owns reduction column kl and rows ml + 16*j
  const int kl = tid % 16, ml = tid / 16;
  int rn_[4], rh[4], rw[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int64_t m = m0 + ml + 16 * j;
    if (m < M) {
      const int ww = (int)(m % OW);
      const int64_t t = m / OW;
      rw[j] = ww; rh[j] = (int)(t % OH); rn_[j] = (int)(t / OH);
    } else {
      rn_[j] = -1; rh[j] = 0; rw[j] = 0;
    }
  }
  const int nl = tid % 64, k2 = tid / 64;  // B loader
  const int tr = tid / 16, tc = tid % 16;
  float acc[4][4] = {};
  for (int k0 = 0; k0 < Kd; k0 += 16) {
    {
      const int kk = k0 + kl;
      const bool kval = kk < Kd;
      const int tap = kval ? kk / CI : 0;
      const int c = kk - tap * CI;
      const int r = tap / g.k, s = tap - r * g.k;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float v = 0.f;
        if (kval && rn_[j] >= 0) {
          int ih, iw;
          bool ok;
          if (!DGRAD) {
            ih = rh[j] * g.stride - g.pad + r;
            iw = rw[j] * g.stride - g.pad + s;
            ok = ih >= 0 && ih < IH && iw >= 0 && iw < IW;
          } else {
            const int th = rh[j] + g.pad - r, tw = rw[j] + g.pad - s;
            ok = th >= 0 && tw >= 0 && (th % g.stride == 0) && (tw % g.stride == 0);
            ih = th / g.stride;
            iw = tw / g.stride;
            ok = ok && ih < IH && iw < IW;
          }
          if (ok) v = to_f<T>(in[(((int64_t)rn_[j] * IH + ih) * IW + iw) * CI + c]);
        }
        As[kl][ml + 16 * j] = v;
      }
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int kk = k0 + k2 + 4 * j;
      const int n = n0 + nl;
      float v = 0.f;
      if (kk < Kd && n < CO) {
        if (!DGRAD) {
          v = to_f<T>(w[(int64_t)kk * g.Cout + n]);
        } else {
          const int tap = kk / CI, co = kk - tap * CI;
          v = to_f<T>(w[((int64_t)tap * g.Cin + n) * g.Cout + co]);
        }
      }
      Bs[k2 + 4 * j][nl] = v;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < 16; ++kk) {
      float a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) { a[i] = As[kk][tr * 4 + i]; b[i] = Bs[kk][tc * 4 + i]; }
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int64_t m = m0 + tr * 4 + i;
    if (m >= M) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int n = n0 + tc * 4 + j;
      if (n >= CO) continue;
      const int64_t o = m * CO + n;
      const float v = beta != 0.f ? to_f<T>(out[o]) + acc[i][j] : acc[i][j];
      out[o] = from_f<T>(v);
    }
  }
}

// wgrad: dw[row=(r,s,ci)][co] = sum_{p=(n,ho,wo)} x(n, ho*st-pad+r, wo*st-pad+s, ci) * dy[p][co]
// split-K over blockIdx.z pixel ranges; partial tiles go to `part` ([splits][rows][Cout]) or
// straight to dw when there is a single split.
template <typename T>
__global__ void __launch_bounds__(CT)
conv_wgrad_kernel(ConvGeom g, const T* __restrict__ x, const T* __restrict__ dy, float* __restrict__ part,
                  int64_t pix_per_split) {
  __shared__ float As[16][64 + 4];
  __shared__ float Bs[16][64 + 4];
  const int tid = threadIdx.x;
  const int rows = g.k * g.k * g.Cin;
  const int64_t P = (int64_t)g.N * g.Ho * g.Wo;
  const int r0 = blockIdx.x * 64, n0 = blockIdx.y * 64;
  const int64_t p_begin = (int64_t)blockIdx.z * pix_per_split;
  int64_t p_end = p_begin + pix_per_split;
  if (p_end > P) p_end = P;
  const int el = tid % 64, kp = tid / 64;  // loader: element (row or co) el, pixel slots kp + 4*j
  const int row = r0 + el;
  const bool rval = row < rows;
  const int tap = rval ? row / g.Cin : 0;
  const int ci = row - tap * g.Cin;
  const int fr = tap / g.k, fs = tap - fr * g.k;
  const int co_l = n0 + el;
  const int tr = tid / 16, tc = tid % 16;
  float acc[4][4] = {};
  for (int64_t p0 = p_begin; p0 < p_end; p0 += 16) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int64_t p = p0 + kp + 4 * j;
      float va = 0.f, vb = 0.f;
      if (p < p_end) {
        const int wo = (int)(p % g.Wo);
        const int64_t t = p / g.Wo;
        const int ho = (int)(t % g.Ho);
        const int n = (int)(t / g.Ho);
        if (rval) {
          const int ih = ho * g.stride - g.pad + fr, iw = wo * g.stride - g.pad + fs;
          if (ih >= 0 && ih < g.H && iw >= 0 && iw < g.W)
            va = to_f<T>(x[(((int64_t)n * g.H + ih) * g.W + iw) * g.Cin + ci]);
        }
        if (co_l < g.Cout) vb = to_f<T>(dy[p * g.Cout + co_l]);
      }
      As[kp + 4 * j][el] = va;
      Bs[kp + 4 * j][el] = vb;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < 16; ++kk) {
      float a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) { a[i] = As[kk][tr * 4 + i]; b[i] = Bs[kk][tc * 4 + i]; }
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }
  float* o = part + (int64_t)blockIdx.z * rows * g.Cout;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int rr = r0 + tr * 4 + i;
    if (rr >= rows) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int n = n0 + tc * 4 + j;
      if (n < g.Cout) o[(int64_t)rr * g.Cout + n] = acc[i][j];
    }
  }
}

__global__ void splitk_reduce_kernel(const float* __restrict__ part, float* __restrict__ out, int64_t n, int splits) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    float s = 0.f;
    for (int z = 0; z < splits; ++z) s += part[(int64_t)z * n + i];
    out[i] = s;
  }
}

static int check_desc(const rn_conv_desc* d, const char* who) {
  if (!d) return fail(RN_ERR_INVALID, "%s: NULL desc", who);
  if (d->N <= 0 || d->H <= 0 || d->W <= 0 || d->Cin <= 0 || d->Ho <= 0 || d->Wo <= 0 || d->Cout <= 0 || d->k <= 0 ||
      d->stride <= 0 || d->pad < 0)
    return fail(RN_ERR_INVALID, "%s: non-positive dimension", who);
  if (d->dtype != RN_F32 && d->dtype != RN_BF16) return fail(RN_ERR_INVALID, "%s: bad dtype %d", who, d->dtype);
  // every output must see at least the (possibly padded) start of its window
  if ((d->Ho - 1) * d->stride - d->pad >= d->H || (d->Wo - 1) * d->stride - d->pad >= d->W)
    return fail(RN_ERR_INVALID, "%s: output %dx%d does not fit input %dx%d (k=%d s=%d pad=%d)", who, d->Ho, d->Wo, d->H,
                d->W, d->k, d->stride, d->pad);
  return RN_OK;
}

static ConvGeom geom(const rn_conv_desc* d) {
  return ConvGeom{d->N, d->H, d->W, d->Cin, d->Ho, d->Wo, d->Cout, d->k, d->stride, d->pad};
}

static int wgrad_splits(const rn_conv_desc* d) {
  const int rows = d->k * d->k * d->Cin;
  const int64_t tiles = (int64_t)((rows + 63) / 64) * ((d->Cout + 63) / 64);
  const int64_t P = (int64_t)d->N * d->Ho * d->Wo;
  int64_t s = (RN_NUM_SMS * 4 + tiles - 1) / tiles;
  const int64_t max_by_work = (P + 511) / 512;
  if (s > max_by_work) s = max_by_work;
  if (s > 128) s = 128;
  if (s < 1) s = 1;
  return (int)s;
}

}  // namespace rn
using namespace rn;

extern "C" int rn_conv2d_fprop(const rn_conv_desc* d, const void* x, const void* w, void* y, void* stream) {
  int rc = check_desc(d, "conv2d_fprop");
  if (rc) return rc;
  RN_REQUIRE(x && w && y, "conv2d_fprop: NULL pointer");
  const int64_t M = (int64_t)d->N * d->Ho * d->Wo;
  dim3 grid((unsigned)((M + 63) / 64), (d->Cout + 63) / 64);
  if (d->dtype == RN_F32)
    conv_igemm_kernel<float, false><<<grid, CT, 0, as_stream(stream)>>>(geom(d), (const float*)x, (const float*)w,
                                                                        (float*)y, 0.f);
  else
    conv_igemm_kernel<__nv_bfloat16, false><<<grid, CT, 0, as_stream(stream)>>>(
        geom(d), (const __nv_bfloat16*)x, (const __nv_bfloat16*)w, (__nv_bfloat16*)y, 0.f);
  return check_launch("conv_igemm_kernel<fprop>");
}

extern "C" int rn_conv2d_dgrad(const rn_conv_desc* d, const void* dy, const void* w, void* dx, float beta,
                               void* stream) {
  int rc = check_desc(d, "conv2d_dgrad");
  if (rc) return rc;
  RN_REQUIRE(dy && w && dx, "conv2d_dgrad: NULL pointer");
  const int64_t M = (int64_t)d->N * d->H * d->W;
  dim3 grid((unsigned)((M + 63) / 64), (d->Cin + 63) / 64);
  if (d->dtype == RN_F32)
    conv_igemm_kernel<float, true><<<grid, CT, 0, as_stream(stream)>>>(geom(d), (const float*)dy, (const float*)w,
                                                                       (float*)dx, beta);
  else
    conv_igemm_kernel<__nv_bfloat16, true><<<grid, CT, 0, as_stream(stream)>>>(
        geom(d), (const __nv_bfloat16*)dy, (const __nv_bfloat16*)w, (__nv_bfloat16*)dx, beta);
  return check_launch("conv_igemm_kernel<dgrad>");
}

extern "C" size_t rn_conv2d_wgrad_scratch_bytes(const rn_conv_desc* d) {
  if (!d) return 0;
  const int s = wgrad_splits(d);
  if (s <= 1) return 0;
  return (size_t)s * d->k * d->k * d->Cin * d->Cout * sizeof(float);
}

extern "C" int rn_conv2d_wgrad(const rn_conv_desc* d, const void* x, const void* dy, float* dw, void* scratch,
                               size_t scratch_bytes, void* stream) {
  int rc = check_desc(d, "conv2d_wgrad");
  if (rc) return rc;
  RN_REQUIRE(x && dy && dw, "conv2d_wgrad: NULL pointer");
  const int splits = wgrad_splits(d);
  const size_t need = rn_conv2d_wgrad_scratch_bytes(d);
  if (need > 0 && (scratch == nullptr || scratch_bytes < need))
    return fail(RN_ERR_SCRATCH, "conv2d_wgrad: scratch %zu < %zu bytes", scratch_bytes, need);
  const int rows = d->k * d->k * d->Cin;
  const int64_t P = (int64_t)d->N * d->Ho * d->Wo;
  int64_t pps = (P + splits - 1) / splits;
  pps = (pps + 15) / 16 * 16;
  dim3 grid((rows + 63) / 64, (d->Cout + 63) / 64, splits);
  float* part = splits > 1 ? (float*)scratch : dw;
  cudaStream_t st = as_stream(stream);
  if (d->dtype == RN_F32)
    conv_wgrad_kernel<float><<<grid, CT, 0, st>>>(geom(d), (const float*)x, (const float*)dy, part, pps);
  else
    conv_wgrad_kernel<__nv_bfloat16><<<grid, CT, 0, st>>>(geom(d), (const __nv_bfloat16*)x, (const __nv_bfloat16*)dy,
                                                         part, pps);
  rc = check_launch("conv_wgrad_kernel");
  if (rc) return rc;
  if (splits > 1) {
    const int64_t n = (int64_t)rows * d->Cout;
    splitk_reduce_kernel<<<grid_for(n, 256), 256, 0, st>>>(part, dw, n, splits);
    rc = check_launch("splitk_reduce_kernel");
  }
  return rc;
}
