// Pooling kernels (HBM-bound, NHWC; one thread per output/input element, channels fastest so
// every warp touches contiguous memory).
#include "rn_common.cuh"

namespace rn {

// 3x3 stride-2 max pool.  Padding cells never win (TF pads with -inf).  The first maximum in
// row-major window order wins ties; its window position (0..8) is kept for the backward pass.
template <typename T>
__global__ void maxpool_fwd_kernel(const T* __restrict__ x, T* __restrict__ y, uint8_t* __restrict__ argmax, int N,
                                   int H, int W, int C, int Ho, int Wo, int pad) {
  const int64_t total = (int64_t)N * Ho * Wo * C;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % C);
    int64_t p = i / C;
    const int wo = (int)(p % Wo);
    p /= Wo;
    const int ho = (int)(p % Ho);
    const int n = (int)(p / Ho);
    float best = -INFINITY;
    int bi = 0;
    bool any = false;
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      const int h = ho * 2 - pad + r;
      if (h < 0 || h >= H) continue;
#pragma unroll
      for (int s = 0; s < 3; ++s) {
        const int w = wo * 2 - pad + s;
        if (w < 0 || w >= W) continue;
        const float v = to_f<T>(x[(((int64_t)n * H + h) * W + w) * C + c]);
        if (!any || v > best) { best = v; bi = r * 3 + s; any = true; }
      }
    }
    y[i] = from_f<T>(best);
    argmax[i] = (uint8_t)bi;
  }
}

// Gather form of MaxPoolGrad: every input element sums the gradients of the (at most 2x2)
// windows whose recorded argmax points at it.  No atomics, deterministic.
template <typename T>
__global__ void maxpool_bwd_kernel(const T* __restrict__ dy, const uint8_t* __restrict__ argmax, T* dx, int N, int H,
                                   int W, int C, int Ho, int Wo, int pad, float beta) {
  const int64_t total = (int64_t)N * H * W * C;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % C);
    int64_t p = i / C;
    const int w = (int)(p % W);
    p /= W;
    const int h = (int)(p % H);
    const int n = (int)(p / H);
    float acc = 0.f;
    // windows with ho*2 - pad + r == h, r in [0,3)
    for (int r = 0; r < 3; ++r) {
      const int t = h + pad - r;
      if (t < 0 || (t & 1)) continue;
      const int ho = t >> 1;
      if (ho >= Ho) continue;
      for (int s = 0; s < 3; ++s) {
        const int u = w + pad - s;
        if (u < 0 || (u & 1)) continue;
        const int wo = u >> 1;
        if (wo >= Wo) continue;
        const int64_t o = (((int64_t)n * Ho + ho) * Wo + wo) * C + c;
        if (argmax[o] == (uint8_t)(r * 3 + s)) acc += to_f<T>(dy[o]);
      }
    }
    dx[i] = from_f<T>(beta != 0.f ? to_f<T>(dx[i]) + acc : acc);
  }
}

// 128-bit vector versions (C % Vec<T>::N == 0): one thread per pixel x channel group, flat grid-stride index with
// 32-bit divisions.  Every thread issues ALL its loads (clamped addresses, so they are unconditional) before it
// consumes any: with the boundary tests in front of each load the compiler kept one load in flight per thread and
// the kernels ran at a third of the HBM bandwidth (ResNet-50 stem, 128 x 112 x 112 x 64: 109 / 144 us).
template <typename T>
__global__ void __launch_bounds__(256)
maxpool_fwd_vec_kernel(const T* __restrict__ x, T* __restrict__ y, uint8_t* __restrict__ argmax, int N,
                       int H, int W, int C, int Ho, int Wo, int pad) {
  constexpr int V = Vec<T>::N;
  const int CV = C / V;
  const int per_row = Wo * CV;
  const int64_t total = (int64_t)N * Ho * per_row;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int row = (int)(i / per_row), e = (int)(i - (int64_t)row * per_row);
    const int cg = e % CV, wo = e / CV;
    const int ho = row % Ho, n = row / Ho;
    bool vh[3], vw[3];
    int hh[3], ww[3];
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      const int h = ho * 2 - pad + r, w = wo * 2 - pad + r;
      vh[r] = h >= 0 && h < H;
      vw[r] = w >= 0 && w < W;
      hh[r] = min(max(h, 0), H - 1);
      ww[r] = min(max(w, 0), W - 1);
    }
    Vec<T> v[9];
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
      for (int q = 0; q < 3; ++q)
        v[r * 3 + q].load(x + (((int64_t)n * H + hh[r]) * W + ww[q]) * C + (int64_t)cg * V);
    float best[V];
    int bi[V];
    bool any = false;
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
      for (int q = 0; q < 3; ++q) {
        if (!(vh[r] && vw[q])) continue;
        float f[V];
        v[r * 3 + q].unpack(f);
#pragma unroll
        for (int j = 0; j < V; ++j)
          if (!any || f[j] > best[j]) { best[j] = f[j]; bi[j] = r * 3 + q; }
        any = true;
      }
    Vec<T> o;
    o.pack(best);
    const int64_t ob = i * V;                     // = (((n * Ho + ho) * Wo + wo) * C + cg * V
    o.store(y + ob);
    uint8_t am[V];
#pragma unroll
    for (int j = 0; j < V; ++j) am[j] = (uint8_t)bi[j];
    if (V == 8) *reinterpret_cast<uint2*>(argmax + ob) = *reinterpret_cast<const uint2*>(am);
    else        *reinterpret_cast<uint32_t*>(argmax + ob) = *reinterpret_cast<const uint32_t*>(am);
  }
}

// Backward, gather form over 2 x 2 input CELLS: the input pixels (h + pad, w + pad) in {2m, 2m+1} x {2u, 2u+1} lie
// in the windows {m-1, m} x {u-1, u} only, and those four windows hold exactly nine (window, position) pairs that
// point into the cell.  One thread per (cell, channel group): four (dy, argmax) loads feed four dx stores, i.e.
// 24 bytes loaded per 16 bytes stored instead of 96 for a thread per input pixel (every window is read by 4 threads,
// not 9), which is what bounded the kernel (L2 -> SM traffic).  No atomics, deterministic.
template <typename T>
__global__ void __launch_bounds__(256)
maxpool_bwd_vec_kernel(const T* __restrict__ dy, const uint8_t* __restrict__ argmax, T* dx, int N, int H,
                       int W, int C, int Ho, int Wo, int pad, float beta) {
  constexpr int V = Vec<T>::N;
  const int CV = C / V;
  const int Hc = (H + pad + 1) / 2, Wc = (W + pad + 1) / 2;          // cells
  const int per_row = Wc * CV;
  const int64_t total = (int64_t)N * Hc * per_row;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int row = (int)(i / per_row), e = (int)(i - (int64_t)row * per_row);
    const int cg = e % CV, u = e / CV;
    const int m = row % Hc, n = row / Hc;
    Vec<T> g[4];
    uint2 amv[4];
    bool ok[4];
#pragma unroll
    for (int a = 0; a < 2; ++a) {
      const int ho = m - 1 + a;
      const bool okh = ho >= 0 && ho < Ho;
      const int hoc = min(max(ho, 0), Ho - 1);
#pragma unroll
      for (int b = 0; b < 2; ++b) {
        const int wo = u - 1 + b;
        ok[a * 2 + b] = okh && wo >= 0 && wo < Wo;
        const int woc = min(max(wo, 0), Wo - 1);
        const int64_t ob = (((int64_t)n * Ho + hoc) * Wo + woc) * C + (int64_t)cg * V;
        g[a * 2 + b].load(dy + ob);
        if (V == 8) amv[a * 2 + b] = *reinterpret_cast<const uint2*>(argmax + ob);
        else        amv[a * 2 + b] = make_uint2(*reinterpret_cast<const uint32_t*>(argmax + ob), 0u);
      }
    }
    float f[4][V];
#pragma unroll
    for (int q = 0; q < 4; ++q) g[q].unpack(f[q]);
    // cell pixel (i2, j2): window (a, b) reaches it at position r = i2 + 2 * (1 - a) (< 3), s = j2 + 2 * (1 - b) (< 3)
#pragma unroll
    for (int i2 = 0; i2 < 2; ++i2) {
      const int h = 2 * m + i2 - pad;
      if (h < 0 || h >= H) continue;
#pragma unroll
      for (int j2 = 0; j2 < 2; ++j2) {
        const int w = 2 * u + j2 - pad;
        if (w < 0 || w >= W) continue;
        float acc[V];
#pragma unroll
        for (int j = 0; j < V; ++j) acc[j] = 0.f;
#pragma unroll
        for (int a = 0; a < 2; ++a) {
          const int r = i2 + 2 * (1 - a);
          if (r > 2) continue;
#pragma unroll
          for (int b = 0; b < 2; ++b) {
            const int sft = j2 + 2 * (1 - b);
            if (sft > 2 || !ok[a * 2 + b]) continue;
            const uint32_t me = (uint32_t)(r * 3 + sft);
#pragma unroll
            for (int j = 0; j < V; ++j) {
              const uint32_t word = j < 4 ? amv[a * 2 + b].x : amv[a * 2 + b].y;
              if (((word >> (8 * (j & 3))) & 0xffu) == me) acc[j] += f[a * 2 + b][j];
            }
          }
        }
        const int64_t xb = (((int64_t)n * H + h) * W + w) * C + (int64_t)cg * V;
        if (beta != 0.f) {
          Vec<T> o;
          float t[V];
          o.load(dx + xb);
          o.unpack(t);
#pragma unroll
          for (int j = 0; j < V; ++j) acc[j] += t[j];
        }
        Vec<T> o;
        o.pack(acc);
        o.store(dx + xb);
      }
    }
  }
}

template <typename TI, typename TO>
__global__ void avgpool_fwd_kernel(const TI* __restrict__ x, TO* __restrict__ y, int N, int H, int W, int C, int k,
                                   int s, int Ho, int Wo, int Cout, int cb) {
  const int64_t total = (int64_t)N * Ho * Wo * Cout;
  const float inv = 1.f / (float)(k * k);
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int co = (int)(i % Cout);
    int64_t p = i / Cout;
    const int wo = (int)(p % Wo);
    p /= Wo;
    const int ho = (int)(p % Ho);
    const int n = (int)(p / Ho);
    const int c = co - cb;
    float acc = 0.f;
    if (c >= 0 && c < C) {
      for (int r = 0; r < k; ++r)
        for (int q = 0; q < k; ++q)
          acc += to_f<TI>(x[(((int64_t)n * H + ho * s + r) * W + wo * s + q) * C + c]);
      acc *= inv;
    }
    y[i] = from_f<TO>(acc);
  }
}

template <typename TI, typename TO>
__global__ void avgpool_bwd_kernel(const TO* __restrict__ dy, TI* dx, int N, int H, int W, int C, int k, int s, int Ho,
                                   int Wo, int Cout, int cb, float beta) {
  const int64_t total = (int64_t)N * H * W * C;
  const float inv = 1.f / (float)(k * k);
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % C);
    int64_t p = i / C;
    const int w = (int)(p % W);
    p /= W;
    const int h = (int)(p % H);
    const int n = (int)(p / H);
    float acc = 0.f;
    // windows (ho, wo) with ho*s <= h < ho*s + k
    const int ho_lo = h - k + 1 <= 0 ? 0 : (h - k + 1 + s - 1) / s;
    const int wo_lo = w - k + 1 <= 0 ? 0 : (w - k + 1 + s - 1) / s;
    for (int ho = ho_lo; ho < Ho && ho * s <= h; ++ho)
      for (int wo = wo_lo; wo < Wo && wo * s <= w; ++wo)
        acc += to_f<TO>(dy[(((int64_t)n * Ho + ho) * Wo + wo) * Cout + cb + c]);
    acc *= inv;
    dx[i] = from_f<TI>(beta != 0.f ? to_f<TI>(dx[i]) + acc : acc);
  }
}

// 8 output channels per thread (bf16 x; C / Cout / cb multiples of 8, so a group is either all padding or all data)
template <typename TO>
__global__ void avgpool_fwd_vec_kernel(const __nv_bfloat16* __restrict__ x, TO* __restrict__ y, int N, int H, int W, int C,
                                       int k, int s, int Ho, int Wo, int Cout, int cb) {
  const int CV = Cout / 8;
  const int64_t total = (int64_t)N * Ho * Wo * CV;
  const float inv = 1.f / (float)(k * k);
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int cg = (int)(i % CV);
    int64_t p = i / CV;
    const int wo = (int)(p % Wo);
    p /= Wo;
    const int ho = (int)(p % Ho);
    const int n = (int)(p / Ho);
    const int c = cg * 8 - cb;
    float acc[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = 0.f;
    if (c >= 0 && c < C) {
      for (int r = 0; r < k; ++r)
        for (int q = 0; q < k; ++q) {
          Vec<__nv_bfloat16> v;
          float f[8];
          v.load(x + (((int64_t)n * H + ho * s + r) * W + wo * s + q) * C + c);
          v.unpack(f);
#pragma unroll
          for (int j = 0; j < 8; ++j) acc[j] += f[j];
        }
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] *= inv;
    }
    if (sizeof(TO) == 4) {
      float4* o = reinterpret_cast<float4*>(y + i * 8);
      o[0] = make_float4(acc[0], acc[1], acc[2], acc[3]);
      o[1] = make_float4(acc[4], acc[5], acc[6], acc[7]);
    } else {
      Vec<__nv_bfloat16> o;
      o.pack(acc);
      o.store(reinterpret_cast<__nv_bfloat16*>(y) + i * 8);
    }
  }
}

// 8 channels per thread (bf16 dx, C / Cout / cb multiples of 8): one 128-bit store per pixel x channel group
template <typename TO>
__global__ void avgpool_bwd_vec_kernel(const TO* __restrict__ dy, __nv_bfloat16* dx, int N, int H, int W, int C, int k,
                                       int s, int Ho, int Wo, int Cout, int cb, float beta) {
  const int CV = C / 8;
  const int64_t total = (int64_t)N * H * W * CV;
  const float inv = 1.f / (float)(k * k);
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int cg = (int)(i % CV);
    int64_t p = i / CV;
    const int w = (int)(p % W);
    p /= W;
    const int h = (int)(p % H);
    const int n = (int)(p / H);
    float acc[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = 0.f;
    const int ho_lo = h - k + 1 <= 0 ? 0 : (h - k + 1 + s - 1) / s;
    const int wo_lo = w - k + 1 <= 0 ? 0 : (w - k + 1 + s - 1) / s;
    for (int ho = ho_lo; ho < Ho && ho * s <= h; ++ho)
      for (int wo = wo_lo; wo < Wo && wo * s <= w; ++wo) {
        const TO* src = dy + (((int64_t)n * Ho + ho) * Wo + wo) * Cout + cb + cg * 8;
        float f[8];
        if (sizeof(TO) == 4) {
          const float4 a = *reinterpret_cast<const float4*>(src), b = *reinterpret_cast<const float4*>(src + 4);
          f[0] = a.x; f[1] = a.y; f[2] = a.z; f[3] = a.w; f[4] = b.x; f[5] = b.y; f[6] = b.z; f[7] = b.w;
        } else {
          Vec<__nv_bfloat16> v;
          v.load(reinterpret_cast<const __nv_bfloat16*>(src));
          v.unpack(f);
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] += f[j];
      }
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] *= inv;
    Vec<__nv_bfloat16> o;
    if (beta != 0.f) {
      float f[8];
      o.load(dx + i * 8);
      o.unpack(f);
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] += f[j];
    }
    o.pack(acc);
    o.store(dx + i * 8);
  }
}

}  // namespace rn
using namespace rn;

extern "C" int rn_maxpool3x3s2_fwd(const void* x, void* y, uint8_t* argmax, int32_t N, int32_t H, int32_t W, int32_t C,
                                   int32_t Ho, int32_t Wo, int32_t pad, int32_t dtype, void* stream) {
  RN_REQUIRE(x && y && argmax && N > 0 && H > 0 && W > 0 && C > 0 && Ho > 0 && Wo > 0 && pad >= 0 && pad <= 2,
             "maxpool_fwd: bad arguments");
  RN_REQUIRE((Ho - 1) * 2 - pad < H && (Wo - 1) * 2 - pad < W, "maxpool_fwd: empty window");
  const int g = grid_for((int64_t)N * Ho * Wo * C, 256);
  const bool al = (((uintptr_t)x | (uintptr_t)y) & 15) == 0 && ((uintptr_t)argmax & 7) == 0;
  if (dtype == RN_F32 && C % 4 == 0 && al)
    maxpool_fwd_vec_kernel<float><<<grid_for((int64_t)N * Ho * Wo * C / 4, 256, 16), 256, 0, as_stream(stream)>>>(
        (const float*)x, (float*)y, argmax, N, H, W, C, Ho, Wo, pad);
  else if (dtype == RN_BF16 && C % 8 == 0 && al)
    maxpool_fwd_vec_kernel<__nv_bfloat16><<<grid_for((int64_t)N * Ho * Wo * C / 8, 256, 16), 256, 0, as_stream(stream)>>>(
        (const __nv_bfloat16*)x, (__nv_bfloat16*)y, argmax, N, H, W, C, Ho, Wo, pad);
  else if (dtype == RN_F32)
    maxpool_fwd_kernel<float><<<g, 256, 0, as_stream(stream)>>>((const float*)x, (float*)y, argmax, N, H, W, C, Ho, Wo, pad);
  else if (dtype == RN_BF16)
    maxpool_fwd_kernel<__nv_bfloat16><<<g, 256, 0, as_stream(stream)>>>((const __nv_bfloat16*)x, (__nv_bfloat16*)y,
                                                                        argmax, N, H, W, C, Ho, Wo, pad);
  else
    return fail(RN_ERR_INVALID, "maxpool_fwd: bad dtype %d", dtype);
  return check_launch("maxpool_fwd_kernel");
}

extern "C" int rn_maxpool3x3s2_bwd(const void* dy, const uint8_t* argmax, void* dx, int32_t N, int32_t H, int32_t W,
                                   int32_t C, int32_t Ho, int32_t Wo, int32_t pad, int32_t dtype, float beta,
                                   void* stream) {
  RN_REQUIRE(dy && dx && argmax && N > 0 && H > 0 && W > 0 && C > 0 && Ho > 0 && Wo > 0, "maxpool_bwd: bad arguments");
  const int g = grid_for((int64_t)N * H * W * C, 256);
  const bool al = (((uintptr_t)dy | (uintptr_t)dx) & 15) == 0 && ((uintptr_t)argmax & 7) == 0;
  if (dtype == RN_F32 && C % 4 == 0 && al)
    maxpool_bwd_vec_kernel<float><<<grid_for((int64_t)N * ((H + pad + 1) / 2) * ((W + pad + 1) / 2) * (C / 4), 256, 16), 256, 0, as_stream(stream)>>>(
        (const float*)dy, argmax, (float*)dx, N, H, W, C, Ho, Wo, pad, beta);
  else if (dtype == RN_BF16 && C % 8 == 0 && al)
    maxpool_bwd_vec_kernel<__nv_bfloat16><<<grid_for((int64_t)N * ((H + pad + 1) / 2) * ((W + pad + 1) / 2) * (C / 8), 256, 16), 256, 0, as_stream(stream)>>>(
        (const __nv_bfloat16*)dy, argmax, (__nv_bfloat16*)dx, N, H, W, C, Ho, Wo, pad, beta);
  else if (dtype == RN_F32)
    maxpool_bwd_kernel<float><<<g, 256, 0, as_stream(stream)>>>((const float*)dy, argmax, (float*)dx, N, H, W, C, Ho, Wo,
                                                                pad, beta);
  else if (dtype == RN_BF16)
    maxpool_bwd_kernel<__nv_bfloat16><<<g, 256, 0, as_stream(stream)>>>((const __nv_bfloat16*)dy, argmax,
                                                                        (__nv_bfloat16*)dx, N, H, W, C, Ho, Wo, pad, beta);
  else
    return fail(RN_ERR_INVALID, "maxpool_bwd: bad dtype %d", dtype);
  return check_launch("maxpool_bwd_kernel");
}

extern "C" int rn_avgpool_fwd(const void* x, void* y, int32_t N, int32_t H, int32_t W, int32_t C, int32_t k, int32_t s,
                              int32_t Cout, int32_t cb, int32_t idt, int32_t odt, void* stream) {
  RN_REQUIRE(x && y && N > 0 && C > 0 && k > 0 && s > 0 && H >= k && W >= k && cb >= 0 && Cout >= C + cb,
             "avgpool_fwd: bad arguments");
  const int Ho = (H - k) / s + 1, Wo = (W - k) / s + 1;
  const int g = grid_for((int64_t)N * Ho * Wo * Cout, 256);
  cudaStream_t st = as_stream(stream);
  if (idt == RN_BF16 && C % 8 == 0 && Cout % 8 == 0 && cb % 8 == 0 && (((uintptr_t)x | (uintptr_t)y) & 15) == 0 &&
      (odt == RN_F32 || odt == RN_BF16)) {
    const int gv = grid_for((int64_t)N * Ho * Wo * (Cout / 8), 256);
    if (odt == RN_F32)
      avgpool_fwd_vec_kernel<float><<<gv, 256, 0, st>>>((const __nv_bfloat16*)x, (float*)y, N, H, W, C, k, s, Ho, Wo, Cout, cb);
    else
      avgpool_fwd_vec_kernel<__nv_bfloat16><<<gv, 256, 0, st>>>((const __nv_bfloat16*)x, (__nv_bfloat16*)y, N, H, W, C, k, s,
                                                                Ho, Wo, Cout, cb);
    return check_launch("avgpool_fwd_vec_kernel");
  }
  if (idt == RN_F32 && odt == RN_F32)
    avgpool_fwd_kernel<float, float><<<g, 256, 0, st>>>((const float*)x, (float*)y, N, H, W, C, k, s, Ho, Wo, Cout, cb);
  else if (idt == RN_BF16 && odt == RN_F32)
    avgpool_fwd_kernel<__nv_bfloat16, float><<<g, 256, 0, st>>>((const __nv_bfloat16*)x, (float*)y, N, H, W, C, k, s, Ho,
                                                               Wo, Cout, cb);
  else if (idt == RN_BF16 && odt == RN_BF16)
    avgpool_fwd_kernel<__nv_bfloat16, __nv_bfloat16><<<g, 256, 0, st>>>((const __nv_bfloat16*)x, (__nv_bfloat16*)y, N, H,
                                                                       W, C, k, s, Ho, Wo, Cout, cb);
  else
    return fail(RN_ERR_INVALID, "avgpool_fwd: bad dtypes %d -> %d", idt, odt);
  return check_launch("avgpool_fwd_kernel");
}

extern "C" int rn_avgpool_bwd(const void* dy, void* dx, int32_t N, int32_t H, int32_t W, int32_t C, int32_t k, int32_t s,
                              int32_t Cout, int32_t cb, int32_t idt, int32_t odt, float beta, void* stream) {
  RN_REQUIRE(dy && dx && N > 0 && C > 0 && k > 0 && s > 0 && H >= k && W >= k && cb >= 0 && Cout >= C + cb,
             "avgpool_bwd: bad arguments");
  const int Ho = (H - k) / s + 1, Wo = (W - k) / s + 1;
  const int g = grid_for((int64_t)N * H * W * C, 256);
  cudaStream_t st = as_stream(stream);
  const bool vec = idt == RN_BF16 && C % 8 == 0 && Cout % 8 == 0 && cb % 8 == 0 &&
                   (((uintptr_t)dy | (uintptr_t)dx) & 15) == 0;
  if (vec) {
    const int gv = grid_for((int64_t)N * H * W * (C / 8), 256);
    if (odt == RN_F32)
      avgpool_bwd_vec_kernel<float><<<gv, 256, 0, st>>>((const float*)dy, (__nv_bfloat16*)dx, N, H, W, C, k, s, Ho, Wo,
                                                        Cout, cb, beta);
    else if (odt == RN_BF16)
      avgpool_bwd_vec_kernel<__nv_bfloat16><<<gv, 256, 0, st>>>((const __nv_bfloat16*)dy, (__nv_bfloat16*)dx, N, H, W, C,
                                                                k, s, Ho, Wo, Cout, cb, beta);
    else
      return fail(RN_ERR_INVALID, "avgpool_bwd: bad dtypes %d -> %d", idt, odt);
    return check_launch("avgpool_bwd_vec_kernel");
  }
  if (idt == RN_F32 && odt == RN_F32)
    avgpool_bwd_kernel<float, float><<<g, 256, 0, st>>>((const float*)dy, (float*)dx, N, H, W, C, k, s, Ho, Wo, Cout, cb, beta);
  else if (idt == RN_BF16 && odt == RN_F32)
    avgpool_bwd_kernel<__nv_bfloat16, float><<<g, 256, 0, st>>>((const float*)dy, (__nv_bfloat16*)dx, N, H, W, C, k, s,
                                                               Ho, Wo, Cout, cb, beta);
  else if (idt == RN_BF16 && odt == RN_BF16)
    avgpool_bwd_kernel<__nv_bfloat16, __nv_bfloat16><<<g, 256, 0, st>>>((const __nv_bfloat16*)dy, (__nv_bfloat16*)dx, N,
                                                                       H, W, C, k, s, Ho, Wo, Cout, cb, beta);
  else
    return fail(RN_ERR_INVALID, "avgpool_bwd: bad dtypes %d -> %d", idt, odt);
  return check_launch("avgpool_bwd_kernel");
}
