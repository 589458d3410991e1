// Classifier head (dense + softmax cross-entropy), and the fused
// mean / weight-decay / momentum update over the flat fp32 parameter buffer.
#include "rn_common.cuh"

namespace rn {

// ------------------------------------------------------------------------------------------
// Small generic fp32 GEMM, C[i,j] (+)= sum_k A(i,k) * B(k,j) (+ bias[j]) with arbitrary strides.
// 64x64 tile, 16-deep, 256 threads, 4x4 micro-tile.  The head is <0.1 % of the step's FLOPs.
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
sgemm_strided_kernel(const float* __restrict__ A, int64_t a_rs, int64_t a_cs, const float* __restrict__ Bm, int64_t b_rs,
                     int64_t b_cs, const float* __restrict__ bias, float* __restrict__ Cm, int M, int N, int K) {
  __shared__ float As[16][64 + 4];
  __shared__ float Bs[16][64 + 4];
  const int tid = threadIdx.x;
  const int tr = tid / 16, tc = tid % 16;
  const int m0 = blockIdx.y * 64, n0 = blockIdx.x * 64;
  float acc[4][4] = {};
  float ra[4], rb[4];          // register prefetch of the next k-step (see sgemm_splitk_kernel)
  auto fetch = [&](int k0) {
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      // consecutive threads follow whichever index is contiguous in memory (either operand may be transposed)
      const int e = tid + t * 256;
      const int kk = a_cs == 1 ? e % 16 : e / 64, mm = a_cs == 1 ? e / 16 : e % 64;
      const int gm = m0 + mm, gk = k0 + kk;
      ra[t] = (gm < M && gk < K) ? A[gm * a_rs + gk * a_cs] : 0.f;
      const int nn = b_cs == 1 ? e % 64 : e / 16, k2 = b_cs == 1 ? e / 64 : e % 16;
      const int gn = n0 + nn, gk2 = k0 + k2;
      rb[t] = (gn < N && gk2 < K) ? Bm[gk2 * b_rs + gn * b_cs] : 0.f;
    }
  };
  fetch(0);
  for (int k0 = 0; k0 < K; k0 += 16) {
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      const int e = tid + t * 256;
      const int kk = a_cs == 1 ? e % 16 : e / 64, mm = a_cs == 1 ? e / 16 : e % 64;
      As[kk][mm] = ra[t];
      const int nn = b_cs == 1 ? e % 64 : e / 16, k2 = b_cs == 1 ? e / 64 : e % 16;
      Bs[k2][nn] = rb[t];
    }
    __syncthreads();
    if (k0 + 16 < K) fetch(k0 + 16);
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
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int gm = m0 + tr * 4 + i, gn = n0 + tc * 4 + j;
      if (gm < M && gn < N) Cm[(int64_t)gm * N + gn] = acc[i][j] + (bias ? bias[gn] : 0.f);
    }
}

// Split-K flavour for GEMMs with few output tiles and a long reduction (the ImageNet classifier:
// 128 x 1000 outputs, K = 2048): blockIdx.z takes a K slice and writes a partial tile; a second pass sums the
// slices in a fixed order and adds the bias (deterministic).
__global__ void __launch_bounds__(256)
sgemm_splitk_kernel(const float* __restrict__ A, int64_t a_rs, int64_t a_cs, const float* __restrict__ Bm, int64_t b_rs,
                    int64_t b_cs, float* __restrict__ part, int M, int N, int K, int kslice) {
  __shared__ float As[16][64 + 4];
  __shared__ float Bs[16][64 + 4];
  const int tid = threadIdx.x;
  const int tr = tid / 16, tc = tid % 16;
  const int m0 = blockIdx.y * 64, n0 = blockIdx.x * 64;
  const int kb = blockIdx.z * kslice;
  const int ke = min(K, kb + kslice);
  float acc[4][4] = {};
  // the next k-step's operands are fetched into registers while this one is multiplied (the loop used to be
  // load -> barrier -> multiply -> barrier with nothing in flight during the multiply); same summation order
  float ra[4], rb[4];
  auto fetch = [&](int k0) {
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      // consecutive threads follow whichever index is contiguous in memory (either operand may be transposed)
      const int e = tid + t * 256;
      const int kk = a_cs == 1 ? e % 16 : e / 64, mm = a_cs == 1 ? e / 16 : e % 64;
      const int gm = m0 + mm, gk = k0 + kk;
      ra[t] = (gm < M && gk < ke) ? A[gm * a_rs + gk * a_cs] : 0.f;
      const int nn = b_cs == 1 ? e % 64 : e / 16, k2 = b_cs == 1 ? e / 64 : e % 16;
      const int gn = n0 + nn, gk2 = k0 + k2;
      rb[t] = (gn < N && gk2 < ke) ? Bm[gk2 * b_rs + gn * b_cs] : 0.f;
    }
  };
  fetch(kb);
  for (int k0 = kb; k0 < ke; k0 += 16) {
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      const int e = tid + t * 256;
      const int kk = a_cs == 1 ? e % 16 : e / 64, mm = a_cs == 1 ? e / 16 : e % 64;
      As[kk][mm] = ra[t];
      const int nn = b_cs == 1 ? e % 64 : e / 16, k2 = b_cs == 1 ? e / 64 : e % 16;
      Bs[k2][nn] = rb[t];
    }
    __syncthreads();
    if (k0 + 16 < ke) fetch(k0 + 16);
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
  float* o = part + (int64_t)blockIdx.z * M * N;
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int gm = m0 + tr * 4 + i, gn = n0 + tc * 4 + j;
      if (gm < M && gn < N) o[(int64_t)gm * N + gn] = acc[i][j];
    }
}

__global__ void splitk_bias_reduce_kernel(const float* __restrict__ part, const float* __restrict__ bias,
                                          float* __restrict__ out, int M, int N, int splits) {
  const int64_t n = (int64_t)M * N;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    float s = bias ? bias[i % N] : 0.f;
    for (int z = 0; z < splits; ++z) s += part[(int64_t)z * n + i];
    out[i] = s;
  }
}

__global__ void colsum_kernel(const float* __restrict__ dy, float* __restrict__ db, int B, int K) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= K) return;
  float s = 0.f;
  for (int i = 0; i < B; ++i) s += dy[(int64_t)i * K + j];
  db[j] = s;
}

static int sgemm(const float* A, int64_t a_rs, int64_t a_cs, const float* B, int64_t b_rs, int64_t b_cs,
                 const float* bias, float* C, int M, int N, int K, cudaStream_t st) {
  dim3 grid((N + 63) / 64, (M + 63) / 64);
  sgemm_strided_kernel<<<grid, 256, 0, st>>>(A, a_rs, a_cs, B, b_rs, b_cs, bias, C, M, N, K);
  return check_launch("sgemm_strided_kernel");
}

// ------------------------------------------------------------------------------------------
// softmax + cross-entropy + precision, one CTA, one warp per row, fixed-order final sum.
// ------------------------------------------------------------------------------------------
// One warp per row, 8 rows per CTA; per-row loss / hit flags go to a small static device buffer and the last
// CTA (atomic ticket) sums them in a fixed order.  The buffer is shared by all calls on a device: two
// rn_softmax_xent calls running CONCURRENTLY on different streams would race on it (calls on one stream,
// or ordered by events, are fine).
constexpr int XENT_MAX_ROWS = 4096;
__device__ float g_xent_rows[2 * XENT_MAX_ROWS];
__device__ unsigned g_xent_ticket;

__global__ void __launch_bounds__(256)
softmax_xent_kernel(const float* __restrict__ logits, const float* __restrict__ onehot, int B, int K,
                    float* __restrict__ probs, float* __restrict__ dlogits, float* __restrict__ out) {
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  const float invB = 1.f / (float)B;
  const int r = blockIdx.x * 8 + warp;
  if (r < B) {
    const float* z = logits + (int64_t)r * K;
    const float* y = onehot + (int64_t)r * K;
    float mx = -INFINITY;
    int amx = 0x7fffffff;
    float ymx = -INFINITY;
    int aymx = 0x7fffffff;
    float ysum = 0.f;
    for (int j = lane; j < K; j += 32) {
      const float v = z[j];
      if (v > mx) { mx = v; amx = j; }
      const float yv = y[j];
      if (yv > ymx) { ymx = yv; aymx = j; }
      ysum += yv;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float om = __shfl_xor_sync(0xffffffffu, mx, o);
      const int oa = __shfl_xor_sync(0xffffffffu, amx, o);
      if (om > mx || (om == mx && oa < amx)) { mx = om; amx = oa; }
      const float oy = __shfl_xor_sync(0xffffffffu, ymx, o);
      const int oya = __shfl_xor_sync(0xffffffffu, aymx, o);
      if (oy > ymx || (oy == ymx && oya < aymx)) { ymx = oy; aymx = oya; }
    }
    ysum = warp_sum(ysum);
    float se = 0.f;
    for (int j = lane; j < K; j += 32) se += expf(z[j] - mx);
    se = warp_sum(se);
    const float lse = mx + logf(se);
    float loss = 0.f;
    for (int j = lane; j < K; j += 32) {
      const float lsm = z[j] - lse;
      const float p = expf(lsm);
      const float yv = y[j];
      loss -= yv * lsm;
      if (probs) probs[(int64_t)r * K + j] = p;
      if (dlogits) dlogits[(int64_t)r * K + j] = (p * ysum - yv) * invB;
    }
    loss = warp_sum(loss);
    if (lane == 0) { g_xent_rows[r] = loss; g_xent_rows[XENT_MAX_ROWS + r] = (amx == aymx) ? 1.f : 0.f; }
  }
  __shared__ bool is_last;
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    is_last = atomicAdd(&g_xent_ticket, 1u) == gridDim.x - 1;
    __threadfence();
  }
  __syncthreads();
  if (!is_last || warp != 0) return;
  // fixed order: lane l sums rows l, l+32, ...; then a shuffle tree
  double s = 0.0, c = 0.0;
  for (int i = lane; i < B; i += 32) { s += (double)__ldcg(&g_xent_rows[i]); c += (double)__ldcg(&g_xent_rows[XENT_MAX_ROWS + i]); }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    s += __shfl_xor_sync(0xffffffffu, s, o);
    c += __shfl_xor_sync(0xffffffffu, c, o);
  }
  if (lane == 0) {
    out[0] = (float)(s / B);
    out[1] = (float)(c / B);
    g_xent_ticket = 0;
  }
}

// ------------------------------------------------------------------------------------------
// fused aggregation-mean + L2 weight decay + momentum / SGD update.
// 20 B per parameter of compulsory HBM traffic (read g, w, m; write w, m) + 2 B for the bf16 copy.
// ------------------------------------------------------------------------------------------
constexpr int OPT_THREADS = 256;
constexpr int OPT_MAX_BLOCKS = RN_NUM_SMS * 8;

template <bool UPDATE>
__global__ void __launch_bounds__(OPT_THREADS)
sgd_momentum_kernel(float* __restrict__ w, const float* __restrict__ g, float* __restrict__ m,
                    __nv_bfloat16* __restrict__ wb, int64_t n4, const uint8_t* __restrict__ mask,
                    const float* __restrict__ hyper, float inv_R, float wd, float momentum, int use_momentum,
                    float* sumsq_out, unsigned* counter, float* partials) {
  float lr = 0.f;
  if (UPDATE) lr = hyper[0];
  float ss = 0.f;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
    const float dec = mask ? (float)mask[i >> 4] : 1.f;  // 16 float4 per 64-parameter granule
    float4 wv = reinterpret_cast<const float4*>(w)[i];
    ss += dec * (wv.x * wv.x + wv.y * wv.y + wv.z * wv.z + wv.w * wv.w);
    if (UPDATE) {
      const float4 gv = __ldcs(reinterpret_cast<const float4*>(g) + i);
      const float wdd = wd * dec;
      float4 gg = make_float4(fmaf(gv.x, inv_R, wdd * wv.x), fmaf(gv.y, inv_R, wdd * wv.y),
                              fmaf(gv.z, inv_R, wdd * wv.z), fmaf(gv.w, inv_R, wdd * wv.w));
      if (use_momentum) {
        float4 mv = reinterpret_cast<const float4*>(m)[i];
        mv = make_float4(fmaf(momentum, mv.x, gg.x), fmaf(momentum, mv.y, gg.y), fmaf(momentum, mv.z, gg.z),
                         fmaf(momentum, mv.w, gg.w));
        reinterpret_cast<float4*>(m)[i] = mv;
        gg = mv;
      }
      wv = make_float4(wv.x - lr * gg.x, wv.y - lr * gg.y, wv.z - lr * gg.z, wv.w - lr * gg.w);
      reinterpret_cast<float4*>(w)[i] = wv;
      if (wb) {
        __nv_bfloat162 lo = __floats2bfloat162_rn(wv.x, wv.y), hi = __floats2bfloat162_rn(wv.z, wv.w);
        uint2 pk = make_uint2(*reinterpret_cast<uint32_t*>(&lo), *reinterpret_cast<uint32_t*>(&hi));
        reinterpret_cast<uint2*>(wb)[i] = pk;
      }
    }
  }
  if (sumsq_out == nullptr) return;
  __shared__ float wsum[OPT_THREADS / 32];
  ss = warp_sum(ss);
  if (threadIdx.x % 32 == 0) wsum[threadIdx.x / 32] = ss;
  __syncthreads();
  __shared__ bool is_last;
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int i = 0; i < OPT_THREADS / 32; ++i) t += wsum[i];
    partials[blockIdx.x] = t;
    __threadfence();
    is_last = (atomicAdd(counter, 1u) == gridDim.x - 1);
  }
  __syncthreads();
  if (!is_last) return;
  if (threadIdx.x == 0) {
    __threadfence();
    double t = 0.0;
    for (unsigned b = 0; b < gridDim.x; ++b) t += (double)__ldcg(partials + b);
    sumsq_out[0] = (float)t;
    *counter = 0;
  }
}

}  // namespace rn
using namespace rn;

extern "C" size_t rn_dense_scratch_bytes(int32_t B, int32_t D, int32_t K) {
  const int tiles = ((K + 63) / 64) * ((B + 63) / 64);
  int splits = (2 * RN_NUM_SMS + tiles - 1) / tiles;
  if (splits > D / 64) splits = D / 64;
  if (splits < 2) return 0;
  return (size_t)splits * B * K * sizeof(float);
}

extern "C" int rn_dense_fwd(const float* x, const float* w, const float* b, float* y, int32_t B, int32_t D, int32_t K,
                            void* scratch, size_t scratch_bytes, void* stream) {
  RN_REQUIRE(x && w && y && B > 0 && D > 0 && K > 0, "dense_fwd: bad arguments");
  const size_t need = rn_dense_scratch_bytes(B, D, K);
  if (need == 0 || scratch == nullptr || scratch_bytes < need)      // small problem (or no scratch): one pass
    return sgemm(x, D, 1, w, K, 1, b, y, B, K, D, as_stream(stream));
  const int splits = (int)(need / ((size_t)B * K * sizeof(float)));
  int kslice = (D + splits - 1) / splits;
  kslice = (kslice + 15) / 16 * 16;
  dim3 grid((K + 63) / 64, (B + 63) / 64, splits);
  sgemm_splitk_kernel<<<grid, 256, 0, as_stream(stream)>>>(x, D, 1, w, K, 1, (float*)scratch, B, K, D, kslice);
  int rc = check_launch("sgemm_splitk_kernel");
  if (rc) return rc;
  splitk_bias_reduce_kernel<<<grid_for((int64_t)B * K, 256), 256, 0, as_stream(stream)>>>((const float*)scratch, b, y,
                                                                                         B, K, splits);
  return check_launch("splitk_bias_reduce_kernel");
}

extern "C" size_t rn_dense_bwd_scratch_bytes(int32_t B, int32_t D, int32_t K) {
  const int tiles = ((D + 63) / 64) * ((B + 63) / 64);
  int splits = (2 * RN_NUM_SMS + tiles - 1) / tiles;
  if (splits > K / 64) splits = K / 64;
  if (splits < 2) return 0;
  return (size_t)splits * B * D * sizeof(float);
}

extern "C" int rn_dense_bwd(const float* x, const float* w, const float* dy, float* dw, float* db, float* dx, int32_t B,
                            int32_t D, int32_t K, void* scratch, size_t scratch_bytes, void* stream) {
  RN_REQUIRE(x && w && dy && B > 0 && D > 0 && K > 0, "dense_bwd: bad arguments");
  cudaStream_t st = as_stream(stream);
  int rc;
  if (dw) {  // dw[D,K] = x^T [D,B] . dy [B,K]
    rc = sgemm(x, 1, D, dy, K, 1, nullptr, dw, D, K, B, st);
    if (rc) return rc;
  }
  if (db) {
    colsum_kernel<<<(K + 127) / 128, 128, 0, st>>>(dy, db, B, K);
    rc = check_launch("colsum_kernel");
    if (rc) return rc;
  }
  if (dx) {  // dx[B,D] = dy [B,K] . w^T [K,D]; with scratch: deterministic split-K over the classes
    const size_t need = rn_dense_bwd_scratch_bytes(B, D, K);
    if (need == 0 || scratch == nullptr || scratch_bytes < need) {
      rc = sgemm(dy, K, 1, w, 1, K, nullptr, dx, B, D, K, st);
      if (rc) return rc;
    } else {
      const int splits = (int)(need / ((size_t)B * D * sizeof(float)));
      int kslice = (K + splits - 1) / splits;
      kslice = (kslice + 15) / 16 * 16;
      dim3 grid((D + 63) / 64, (B + 63) / 64, splits);
      sgemm_splitk_kernel<<<grid, 256, 0, st>>>(dy, K, 1, w, 1, K, (float*)scratch, B, D, K, kslice);
      rc = check_launch("sgemm_splitk_kernel");
      if (rc) return rc;
      splitk_bias_reduce_kernel<<<grid_for((int64_t)B * D, 256), 256, 0, st>>>((const float*)scratch, nullptr, dx, B, D,
                                                                             splits);
      rc = check_launch("splitk_bias_reduce_kernel");
      if (rc) return rc;
    }
  }
  return RN_OK;
}

extern "C" int rn_softmax_xent(const float* logits, const float* onehot, int32_t B, int32_t K, float* probs,
                               float* dlogits, float* out_scalars, void* stream) {
  RN_REQUIRE(logits && onehot && out_scalars && B > 0 && K > 0, "softmax_xent: bad arguments");
  RN_REQUIRE(B <= 4096, "softmax_xent: batch %d > 4096 rows per call", B);
  softmax_xent_kernel<<<(B + 7) / 8, 256, 0, as_stream(stream)>>>(logits, onehot, B, K, probs, dlogits, out_scalars);
  return check_launch("softmax_xent_kernel");
}

extern "C" size_t rn_optimizer_scratch_bytes(int64_t n) {
  (void)n;
  return 256 + (size_t)OPT_MAX_BLOCKS * sizeof(float);
}

static int opt_common(bool update, float* w, const float* g, float* m, void* wb, int64_t n, const uint8_t* mask,
                      const float* hyper, float inv_R, float wd, float momentum, int use_momentum, float* sumsq,
                      void* scratch, size_t scratch_bytes, void* stream) {
  RN_REQUIRE(w && n > 0 && n % 64 == 0, "optimizer: n=%lld must be a positive multiple of 64", (long long)n);
  RN_REQUIRE((((uintptr_t)w | (uintptr_t)g | (uintptr_t)m) & 15) == 0, "optimizer: buffers must be 16-byte aligned");
  if (update) RN_REQUIRE(g && hyper && (m || !use_momentum), "optimizer: NULL pointer");
  if (sumsq && (scratch == nullptr || scratch_bytes < rn_optimizer_scratch_bytes(n)))
    return fail(RN_ERR_SCRATCH, "optimizer: scratch %zu < %zu bytes", scratch_bytes, rn_optimizer_scratch_bytes(n));
  const int64_t n4 = n / 4;
  int grid = (int)((n4 + OPT_THREADS * 4 - 1) / (OPT_THREADS * 4));
  if (grid > OPT_MAX_BLOCKS) grid = OPT_MAX_BLOCKS;
  if (grid < 1) grid = 1;
  unsigned* counter = (unsigned*)scratch;
  float* partials = scratch ? (float*)((char*)scratch + 256) : nullptr;
  if (update)
    sgd_momentum_kernel<true><<<grid, OPT_THREADS, 0, as_stream(stream)>>>(
        w, g, m, (__nv_bfloat16*)wb, n4, mask, hyper, inv_R, wd, momentum, use_momentum, sumsq, counter, partials);
  else
    sgd_momentum_kernel<false><<<grid, OPT_THREADS, 0, as_stream(stream)>>>(
        w, nullptr, nullptr, nullptr, n4, mask, nullptr, 0.f, 0.f, 0.f, 0, sumsq, counter, partials);
  return check_launch("sgd_momentum_kernel");
}

extern "C" int rn_sgd_momentum_fused(float* w, const float* g, float* m, void* w_bf16, int64_t n,
                                     const uint8_t* granule_mask, const float* hyper, float inv_R, float wd,
                                     float momentum, int32_t use_momentum, float* sumsq_out, void* scratch,
                                     size_t scratch_bytes, void* stream) {
  return opt_common(true, w, g, m, w_bf16, n, granule_mask, hyper, inv_R, wd, momentum, use_momentum, sumsq_out,
                    scratch, scratch_bytes, stream);
}

extern "C" int rn_l2_sumsq(const float* w, int64_t n, const uint8_t* granule_mask, float* out, void* scratch,
                           size_t scratch_bytes, void* stream) {
  RN_REQUIRE(out, "l2_sumsq: out is NULL");
  return opt_common(false, const_cast<float*>(w), nullptr, nullptr, nullptr, n, granule_mask, nullptr, 0.f, 0.f, 0.f, 0,
                    out, scratch, scratch_bytes, stream);
}
