// Batch-norm (+ leaky ReLU, + residual add) forward / backward: HBM-streaming kernels.
//
// x is viewed as [M = N*H*W rows, C channels] (NHWC).  Every thread owns ONE 128-bit column
// group (4 fp32 or 8 bf16 channels) and walks down the rows, so per-channel constants live in
// registers and every warp reads up to 512 contiguous bytes per row.  Per-channel reductions:
// registers -> warp shuffles (when several rows share a warp) -> shared memory -> one fp32
// partial per CTA in caller scratch -> the last CTA of each channel chunk (atomic ticket)
// finalises in fp64 in a fixed order (deterministic).
//
// Algorithmic HBM bytes (s = bytes/element, E = M*C):
//   stats            : E*s            (+2*E*s when the residual add is fused: read addend, write z)
//   apply+act        : 2*E*s
//   bwd reduce       : 2*E*s          (dy, x)
//   bwd apply        : 3*E*s          (+E*s with the shortcut-gradient addend)
#include "rn_common.cuh"
#include <stdlib.h>

namespace rn {

constexpr int BN_THREADS = 256;
constexpr int BN_MAX_CHUNKS = 64;       // counters reserved at the head of scratch
constexpr int BN_MAX_G = RN_NUM_SMS * 4;  // row-CTAs per channel chunk (upper bound)

struct BnLayout {
  int CV;     // column groups per row
  int CW;     // column groups per CTA
  int RY;     // rows per CTA iteration
  int gridY;  // channel chunks
  int G;      // row CTAs
  int il;     // 1: CTAs take row blocks round-robin (b, b+G, ...) instead of one contiguous range each
};

template <typename T>
static BnLayout bn_layout(int64_t M, int C, int waves, int max_cw = 32, int64_t bytes_per_cta = 65536) {
  BnLayout L;
  constexpr int V = Vec<T>::N;
  L.CV = C / V;
  // Reducing kernels use narrow channel chunks (max_cw = 8 column groups): more chunks -> the serial
  // last-CTA reduction runs in parallel across chunks and each CTA covers 4x more rows per iteration
  // (fewer dependent load batches on small tensors).
  if ((L.CV + max_cw - 1) / max_cw > BN_MAX_CHUNKS) max_cw = 32;
  L.CW = L.CV < max_cw ? L.CV : max_cw;
  L.RY = BN_THREADS / L.CW;
  L.gridY = (L.CV + L.CW - 1) / L.CW;
  // at least ~64 KB of the tensor per CTA (per channel chunk): small tensors get few CTAs, which keeps
  // the serial last-CTA reduction over the per-CTA partials short
  const int64_t chunk_bytes = M * (int64_t)L.CW * V * (int64_t)sizeof(T);
  int64_t want = (chunk_bytes + bytes_per_cta - 1) / bytes_per_cta;
  int64_t cap = (int64_t)RN_NUM_SMS * waves / L.gridY;
  if (cap < 1) cap = 1;
  if (cap > BN_MAX_G) cap = BN_MAX_G;
  L.G = (int)(want < 1 ? 1 : (want > cap ? cap : want));
  L.il = 0;
  return L;
}

// Reduce NV*2 per-thread accumulators over the rows of the CTA; thread (ty == 0) of each column
// group ends up with the CTA total.  smem: [RY][CW][2*V] floats.
template <int V>
__device__ __forceinline__ void cta_reduce_cols(float (&a)[V], float (&b)[V], int tx, int ty, int CW, int RY,
                                                bool active, float* smem) {
  // warp-shuffle stage: when CW is a power of two below 32, lanes l and l ^ (CW*j) hold the
  // same column group for different rows of the same warp.
  const bool pow2 = (CW & (CW - 1)) == 0;
  int rows_in_warp = 1;
  if (pow2 && CW < 32) {
    rows_in_warp = 32 / CW;
    for (int o = 16; o >= CW; o >>= 1) {
#pragma unroll
      for (int i = 0; i < V; ++i) {
        a[i] += __shfl_xor_sync(0xffffffffu, a[i], o);
        b[i] += __shfl_xor_sync(0xffffffffu, b[i], o);
      }
    }
  }
  const bool writer = active && (ty % rows_in_warp == 0);
  if (writer) {
    float* p = smem + ((size_t)(ty / rows_in_warp) * CW + tx) * 2 * V;
#pragma unroll
    for (int i = 0; i < V; ++i) { p[i] = a[i]; p[V + i] = b[i]; }
  }
  __syncthreads();
  if (active && ty == 0) {
    const int nrow = (RY + rows_in_warp - 1) / rows_in_warp;
    for (int r = 1; r < nrow; ++r) {
      const float* p = smem + ((size_t)r * CW + tx) * 2 * V;
#pragma unroll
      for (int i = 0; i < V; ++i) { a[i] += p[i]; b[i] += p[V + i]; }
    }
  }
}


// Last-CTA finalisation: sum `nparts` per-CTA partials ([2][chunkC] floats each) in a FIXED order.
// Thread t owns four adjacent channels (one 128-bit load per partial and statistic) and one of
// S = 1024 / chunkC slices of the partial list (slice j sums partials j, j+S, ...).  Loads are issued in
// batches before any arithmetic: the L2 ("strong") loads keep program order, so interleaving them
// with the conversions would leave one load in flight and make the tail latency-bound.  The slices
// are then combined in order through shared memory.  Results are valid for threads < chunkC.
constexpr int BN_FIN_DOUBLES = 2 * 1024;
__device__ __forceinline__ void final_reduce(const float* base, unsigned nparts, int chunkC, double* sh, double& out0,
                                             double& out1) {
  const int q4 = chunkC / 4;                         // float4 columns per statistic
  const int slices = BN_THREADS / q4;                // >= 4 (chunkC <= 256)
  const int c4 = threadIdx.x % q4;
  const int slice = threadIdx.x / q4;
  double a0[4] = {0, 0, 0, 0}, a1[4] = {0, 0, 0, 0};
  if (slice < slices) {
    unsigned b = slice;
    constexpr int U = 4;
    for (; b + (unsigned)((U - 1) * slices) < nparts; b += (unsigned)(U * slices)) {
      float4 t0[U], t1[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const float4* q = reinterpret_cast<const float4*>(base + (size_t)(b + u * slices) * 2 * chunkC);
        t0[u] = __ldcg(q + c4);
        t1[u] = __ldcg(q + q4 + c4);
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        a0[0] += (double)t0[u].x; a0[1] += (double)t0[u].y; a0[2] += (double)t0[u].z; a0[3] += (double)t0[u].w;
        a1[0] += (double)t1[u].x; a1[1] += (double)t1[u].y; a1[2] += (double)t1[u].z; a1[3] += (double)t1[u].w;
      }
    }
    for (; b < nparts; b += slices) {
      const float4* q = reinterpret_cast<const float4*>(base + (size_t)b * 2 * chunkC);
      const float4 t0 = __ldcg(q + c4), t1 = __ldcg(q + q4 + c4);
      a0[0] += (double)t0.x; a0[1] += (double)t0.y; a0[2] += (double)t0.z; a0[3] += (double)t0.w;
      a1[0] += (double)t1.x; a1[1] += (double)t1.y; a1[2] += (double)t1.z; a1[3] += (double)t1.w;
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      sh[(size_t)slice * chunkC + c4 * 4 + j] = a0[j];
      sh[1024 + (size_t)slice * chunkC + c4 * 4 + j] = a1[j];
    }
  }
  __syncthreads();
  double s0 = 0.0, s1 = 0.0;
  if ((int)threadIdx.x < chunkC) {
    for (int j = 0; j < slices; ++j) {
      s0 += sh[(size_t)j * chunkC + threadIdx.x];
      s1 += sh[1024 + (size_t)j * chunkC + threadIdx.x];
    }
  }
  out0 = s0;
  out1 = s1;
}

// layout knobs of the two reducing kernels (tunable from the environment for experiments)
static int bn_reduce_cw() {
  static int v = -1;
  if (v < 0) { const char* e = getenv("RN_BN_CW"); v = e ? atoi(e) : 8; if (v < 1 || v > 32) v = 8; }
  return v;
}
static int bn_apply_cw() {
  static int v = -1;
  if (v < 0) { const char* e = getenv("RN_BN_APPLY_CW"); v = e ? atoi(e) : 16; if (v < 1 || v > 32) v = 16; }
  return v;
}
// CTAs per SM that are actually resident: the statistics kernel (<= 64 registers) fits 4 x 256 threads, the
// backward reduction (80 registers) only 3 — a 4th "wave" would run as a tail after the first three finish.
static int bn_reduce_waves(int dflt) {
  static int v = -1;
  if (v < 0) { const char* e = getenv("RN_BN_WAVES"); v = e ? atoi(e) : 0; if (v < 0 || v > 16) v = 0; }
  return v ? v : dflt;
}
static int bn_interleave() {
  static int v = -1;
  if (v < 0) { const char* e = getenv("RN_BN_IL"); v = e ? atoi(e) : 1; }
  return v;
}
static int64_t bn_reduce_cta_bytes() {
  static int64_t v = -1;
  if (v < 0) { const char* e = getenv("RN_BN_CTA_KB"); v = (e ? atoi(e) : 64) * 1024; if (v < 4096) v = 65536; }
  return v;
}

// ------------------------------------------------------------------------------------------
// forward statistics (+ fused residual add)
// ------------------------------------------------------------------------------------------
template <typename T, bool ADD, int U>
__global__ void __launch_bounds__(BN_THREADS)
bn_stats_kernel(const T* __restrict__ x, const T* __restrict__ addend, T* __restrict__ sum_out, int64_t M, int C,
                BnLayout L, float eps, float decay, int unbiased, float* __restrict__ mean_out,
                float* __restrict__ rstd_out, float* moving_mean, float* moving_var, unsigned* counters,
                float* partials) {
  constexpr int V = Vec<T>::N;
  pdl_trigger();
  pdl_wait();
  extern __shared__ float smem[];
  const int tx = threadIdx.x % L.CW, ty = threadIdx.x / L.CW;
  const int cg = blockIdx.y * L.CW + tx;
  const bool active = ty < L.RY && cg < L.CV;
  float s[V], q[V];
#pragma unroll
  for (int i = 0; i < V; ++i) { s[i] = 0.f; q[i] = 0.f; }
  if (active) {
    const int64_t stride = L.RY;
    const int64_t col = (int64_t)cg * V;
    const int64_t rows_per_cta = ((M + gridDim.x - 1) / gridDim.x + L.RY - 1) / L.RY * L.RY;
    const int64_t M_end = L.il ? M : min(M, (int64_t)(blockIdx.x + 1) * rows_per_cta);
    int64_t r = (L.il ? (int64_t)blockIdx.x * (U * stride) : (int64_t)blockIdx.x * rows_per_cta) + ty;
    const int64_t step = L.il ? (int64_t)gridDim.x * (U * stride) : (int64_t)U * stride;
    // U = independent 128-bit loads in flight per thread and stream
    for (; r < M_end; r += step) {
      if (r + (U - 1) * stride < M_end) {
        Vec<T> v[U], w[U];
#pragma unroll
        for (int u = 0; u < U; ++u) v[u].load(x + (r + u * stride) * C + col);
        if (ADD) {
#pragma unroll
          for (int u = 0; u < U; ++u) w[u].load(addend + (r + u * stride) * C + col);
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
          float f[V];
          v[u].unpack(f);
          if (ADD) {
            float g[V];
            w[u].unpack(g);
#pragma unroll
            for (int i = 0; i < V; ++i) f[i] += g[i];
            Vec<T> o;
            o.pack(f);
            o.store(sum_out + (r + u * stride) * C + col);
            o.unpack(f);  // statistics of the value as stored (matters in bf16 mode)
          }
#pragma unroll
          for (int i = 0; i < V; ++i) { s[i] += f[i]; q[i] = fmaf(f[i], f[i], q[i]); }
        }
      } else {
        for (int64_t rr = r; rr < M_end && rr < r + U * stride; rr += stride) {
          Vec<T> v;
          float f[V];
          const int64_t off = rr * C + col;
          v.load(x + off);
          v.unpack(f);
          if (ADD) {
            Vec<T> w;
            float g[V];
            w.load(addend + off);
            w.unpack(g);
#pragma unroll
            for (int i = 0; i < V; ++i) f[i] += g[i];
            Vec<T> o;
            o.pack(f);
            o.store(sum_out + off);
            o.unpack(f);
          }
#pragma unroll
          for (int i = 0; i < V; ++i) { s[i] += f[i]; q[i] = fmaf(f[i], f[i], q[i]); }
        }
      }
    }
  }
  cta_reduce_cols<V>(s, q, tx, ty, L.CW, L.RY, active, smem);
  const int chunkC = L.CW * V;
  float* my = partials + ((size_t)blockIdx.y * gridDim.x + blockIdx.x) * 2 * chunkC;
  if (active && ty == 0) {
#pragma unroll
    for (int i = 0; i < V; ++i) { my[tx * V + i] = s[i]; my[chunkC + tx * V + i] = q[i]; }
  }
  // Publish the partials: barrier (CTA-scope ordering of every thread's stores), then ONE gpu-scope
  // fence by the ticket thread (fences are cumulative).  A __threadfence() by all 256 threads costs
  // ~20 us per resident CTA on this part (each one invalidates L1) and dominated small layers.
  __shared__ bool is_last;
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    unsigned t = atomicAdd(&counters[blockIdx.y], 1u);
    is_last = (t == gridDim.x - 1);
    __threadfence();
  }
  __syncthreads();
  if (!is_last) return;
  __shared__ double fin[BN_FIN_DOUBLES];
  double S, Q;
  final_reduce(partials + (size_t)blockIdx.y * gridDim.x * 2 * chunkC, gridDim.x, chunkC, fin, S, Q);
  const int c_local = threadIdx.x;
  const int c = blockIdx.y * chunkC + c_local;
  if (c_local < chunkC && c < C) {
    const double mean = S / (double)M;
    double var = Q / (double)M - mean * mean;
    if (var < 0.0) var = 0.0;
    mean_out[c] = (float)mean;
    rstd_out[c] = (float)(1.0 / sqrt(var + (double)eps));
    if (moving_mean != nullptr) {
      const double mv = unbiased ? var * ((double)M / (double)(M > 1 ? M - 1 : 1)) : var;
      const float om = 1.f - decay;
      moving_mean[c] -= (moving_mean[c] - (float)mean) * om;
      moving_var[c] -= (moving_var[c] - (float)mv) * om;
    }
  }
  if (threadIdx.x == 0) counters[blockIdx.y] = 0;  // leave scratch clean for the next launch
}

// ------------------------------------------------------------------------------------------
// Consumer-side finalisation.  A convolution with a fused epilogue leaves a table partials[rows][C][2] of per-CTA
// sums (rows <= 148).  Instead of one more tiny kernel between two big ones (measured: ~5 us of drain + launch + ramp
// per BN on ResNet-50), every CTA of the apply kernel that follows sums the table for ITS <= 64 channels itself:
// thread -> (channel pair, row slice), 16-byte loads, fp64, slices combined in slice order — the same fixed order in
// every CTA, so all CTAs derive identical constants; the CTA with blockIdx.x == 0 of each channel chunk also writes
// them out (mean / rstd / moving averages, or dgamma / dbeta) for the kernels that need them later.
// ------------------------------------------------------------------------------------------
struct BnFold {
  const float* partials;
  int rows;
  double M;
  float eps, decay;
  int unbiased;
  float *o0, *o1;      // forward: mean, rstd;  backward: dgamma, dbeta
  float *mm, *mv;      // forward: moving mean / variance or NULL
};
constexpr int BN_FOLD_CW = 8;          // column groups per CTA of a folded apply kernel (64 bf16 channels: 32 pairs x 8 slices)
constexpr int BN_FOLD_BATCH = 10;      // loads in flight per thread

// All BN_THREADS threads call this.  Returns true for the threads that own a channel pair (slice 0): s = (S0, Q0, S1, Q1)
// of channels c0 + 2 * pair, c0 + 2 * pair + 1.  sh: BN_THREADS * 4 doubles.
__device__ __forceinline__ bool cta_sum_partial_rows(const float* __restrict__ partials, int rows, int C, int c0, int nch,
                                                     double* sh, double (&s)[4], int* pair) {
  const int np = nch >> 1;
  const int S = BN_THREADS / np;
  const int pr = threadIdx.x % np, slice = threadIdx.x / np;
  s[0] = s[1] = s[2] = s[3] = 0.0;
  if (slice < S) {
    const float4* q = reinterpret_cast<const float4*>(partials + (size_t)2 * c0) + pr;
    const size_t rstride = (size_t)C >> 1;                       // float4 per row
    int z = slice;
    for (; z + (BN_FOLD_BATCH - 1) * S < rows; z += BN_FOLD_BATCH * S) {
      float4 v[BN_FOLD_BATCH];
#pragma unroll
      for (int u = 0; u < BN_FOLD_BATCH; ++u) v[u] = __ldcg(q + (size_t)(z + u * S) * rstride);
#pragma unroll
      for (int u = 0; u < BN_FOLD_BATCH; ++u) { s[0] += (double)v[u].x; s[1] += (double)v[u].y; s[2] += (double)v[u].z; s[3] += (double)v[u].w; }
    }
    float4 v[BN_FOLD_BATCH];
#pragma unroll
    for (int u = 0; u < BN_FOLD_BATCH; ++u) {
      const int zz = z + u * S;
      v[u] = zz < rows ? __ldcg(q + (size_t)zz * rstride) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
#pragma unroll
    for (int u = 0; u < BN_FOLD_BATCH; ++u) { s[0] += (double)v[u].x; s[1] += (double)v[u].y; s[2] += (double)v[u].z; s[3] += (double)v[u].w; }
  }
  if (S > 1) {
    double* mine = sh + 4 * threadIdx.x;
    mine[0] = s[0]; mine[1] = s[1]; mine[2] = s[2]; mine[3] = s[3];
    __syncthreads();
    if (slice == 0)
      for (int j = 1; j < S; ++j) {
        const double* o = sh + 4 * (j * np + pr);
        s[0] += o[0]; s[1] += o[1]; s[2] += o[2]; s[3] += o[3];
      }
  }
  *pair = pr;
  return slice == 0;
}

// ------------------------------------------------------------------------------------------
// forward apply + activation
// ------------------------------------------------------------------------------------------
template <typename T, bool FOLD>
__global__ void __launch_bounds__(BN_THREADS)
bn_apply_kernel(const T* __restrict__ x, T* __restrict__ y, int64_t M, int C, BnLayout L,
                const float* __restrict__ gamma, const float* __restrict__ beta, const float* __restrict__ mean,
                const float* __restrict__ rstd, int stats_are_moving, float eps, float leak, BnFold fold) {
  constexpr int V = Vec<T>::N;
  pdl_trigger();
  pdl_wait();
  const int tx = threadIdx.x % L.CW, ty = threadIdx.x / L.CW;
  const int cg = blockIdx.y * L.CW + tx;
  __shared__ double fold_sh[FOLD ? BN_THREADS * 4 : 1];
  __shared__ float fold_ab[FOLD ? BN_FOLD_CW * V * 2 : 1];
  if (FOLD) {
    const int c0 = blockIdx.y * L.CW * V;
    const int nch = min(L.CW * V, C - c0);
    double s[4];
    int pr;
    if (cta_sum_partial_rows(fold.partials, fold.rows, C, c0, nch, fold_sh, s, &pr)) {
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int c = c0 + 2 * pr + h;
        const double mu = s[2 * h] / fold.M;
        double var = s[2 * h + 1] / fold.M - mu * mu;
        if (var < 0.0) var = 0.0;
        const float mf = (float)mu, rf = (float)(1.0 / sqrt(var + (double)fold.eps));
        if (blockIdx.x == 0)
          bn_finalize_channel(s[2 * h], s[2 * h + 1], fold.M, fold.eps, fold.decay, fold.unbiased, c, fold.o0, fold.o1, fold.mm, fold.mv);
        bn_affine(gamma[c], beta[c], mf, rf, &fold_ab[2 * (2 * pr + h)], &fold_ab[2 * (2 * pr + h) + 1]);
      }
    }
    __syncthreads();
  }
  if (!(ty < L.RY && cg < L.CV)) return;
  float a[V], b[V];
#pragma unroll
  for (int i = 0; i < V; ++i) {
    if (FOLD) {
      a[i] = fold_ab[2 * (tx * V + i)];
      b[i] = fold_ab[2 * (tx * V + i) + 1];
    } else {
      const int c = cg * V + i;
      const float rs = stats_are_moving ? rsqrtf(rstd[c] + eps) : rstd[c];
      bn_affine(gamma[c], beta[c], mean[c], rs, &a[i], &b[i]);
    }
  }
  const int64_t stride = L.RY;
  const int64_t col = (int64_t)cg * V;
  const int64_t rows_per_cta = ((M + gridDim.x - 1) / gridDim.x + L.RY - 1) / L.RY * L.RY;
  const int64_t M_end = min(M, (int64_t)(blockIdx.x + 1) * rows_per_cta);
  int64_t r = (int64_t)blockIdx.x * rows_per_cta + ty;
  constexpr int U = 4;
  for (; r + (U - 1) * stride < M_end; r += U * stride) {
    Vec<T> v[U];
#pragma unroll
    for (int u = 0; u < U; ++u) v[u].load_stream(x + (r + u * stride) * C + col);
#pragma unroll
    for (int u = 0; u < U; ++u) {
      float f[V];
      v[u].unpack(f);
#pragma unroll
      for (int i = 0; i < V; ++i) f[i] = act_fn(fmaf(f[i], a[i], b[i]), leak);
      v[u].pack(f);
      v[u].store(y + (r + u * stride) * C + col);
    }
  }
  for (; r < M_end; r += stride) {
    const int64_t off = r * C + col;
    Vec<T> v;
    float f[V];
    v.load_stream(x + off);
    v.unpack(f);
#pragma unroll
    for (int i = 0; i < V; ++i) f[i] = act_fn(fmaf(f[i], a[i], b[i]), leak);
    v.pack(f);
    v.store(y + off);
  }
}

// ------------------------------------------------------------------------------------------
// backward: reductions
// ------------------------------------------------------------------------------------------
template <typename T, int U, bool CS>
__global__ void __launch_bounds__(BN_THREADS)
bn_bwd_reduce_kernel(const T* __restrict__ dy, const T* __restrict__ x, int64_t M, int C, BnLayout L,
                     const float* __restrict__ gamma, const float* __restrict__ beta, const float* __restrict__ mean,
                     const float* __restrict__ rstd, float leak, float* __restrict__ dgamma, float* __restrict__ dbeta,
                     unsigned* counters, float* partials) {
  constexpr int V = Vec<T>::N;
  pdl_trigger();
  pdl_wait();
  extern __shared__ float smem[];
  const int tx = threadIdx.x % L.CW, ty = threadIdx.x / L.CW;
  const int cg = blockIdx.y * L.CW + tx;
  const bool active = ty < L.RY && cg < L.CV;
  float sb[V], sg[V], a[V], b[V], mu[V], rs[V];
#pragma unroll
  for (int i = 0; i < V; ++i) { sb[i] = 0.f; sg[i] = 0.f; }
  if (active) {
#pragma unroll
    for (int i = 0; i < V; ++i) {
      const int c = cg * V + i;
      rs[i] = rstd[c];
      mu[i] = mean[c];
      a[i] = gamma[c] * rs[i];
      b[i] = beta[c] - mu[i] * a[i];
    }
    const int64_t stride = L.RY;
    const int64_t col = (int64_t)cg * V;
    const int64_t rows_per_cta = ((M + gridDim.x - 1) / gridDim.x + L.RY - 1) / L.RY * L.RY;
    const int64_t M_end = L.il ? M : min(M, (int64_t)(blockIdx.x + 1) * rows_per_cta);
    int64_t r = (L.il ? (int64_t)blockIdx.x * (U * stride) : (int64_t)blockIdx.x * rows_per_cta) + ty;
    const int64_t step = L.il ? (int64_t)gridDim.x * (U * stride) : (int64_t)U * stride;
    for (; r < M_end; r += step) {
      if (r + (U - 1) * stride < M_end) {
        Vec<T> vx[U], vg[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
          if (CS) {
            vx[u].load_stream(x + (r + u * stride) * C + col);
            vg[u].load_stream(dy + (r + u * stride) * C + col);
          } else {
            vx[u].load(x + (r + u * stride) * C + col);
            vg[u].load(dy + (r + u * stride) * C + col);
          }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
          float fx[V], fg[V];
          vx[u].unpack(fx);
          vg[u].unpack(fg);
#pragma unroll
          for (int i = 0; i < V; ++i) {
            const float pre = fmaf(fx[i], a[i], b[i]);
            const float g = pre < 0.f ? leak * fg[i] : fg[i];
            sb[i] += g;
            sg[i] = fmaf(g, (fx[i] - mu[i]) * rs[i], sg[i]);
          }
        }
      } else {
        for (int64_t rr = r; rr < M_end && rr < r + U * stride; rr += stride) {
          const int64_t off = rr * C + col;
          Vec<T> vx, vg;
          float fx[V], fg[V];
          vx.load(x + off);
          vg.load(dy + off);
          vx.unpack(fx);
          vg.unpack(fg);
#pragma unroll
          for (int i = 0; i < V; ++i) {
            const float pre = fmaf(fx[i], a[i], b[i]);
            const float g = pre < 0.f ? leak * fg[i] : fg[i];
            sb[i] += g;
            sg[i] = fmaf(g, (fx[i] - mu[i]) * rs[i], sg[i]);
          }
        }
      }
    }
  }
  cta_reduce_cols<V>(sb, sg, tx, ty, L.CW, L.RY, active, smem);
  const int chunkC = L.CW * V;
  float* my = partials + ((size_t)blockIdx.y * gridDim.x + blockIdx.x) * 2 * chunkC;
  if (active && ty == 0) {
#pragma unroll
    for (int i = 0; i < V; ++i) { my[tx * V + i] = sb[i]; my[chunkC + tx * V + i] = sg[i]; }
  }
  // Publish the partials: barrier (CTA-scope ordering of every thread's stores), then ONE gpu-scope
  // fence by the ticket thread (fences are cumulative).  A __threadfence() by all 256 threads costs
  // ~20 us per resident CTA on this part (each one invalidates L1) and dominated small layers.
  __shared__ bool is_last;
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    unsigned t = atomicAdd(&counters[blockIdx.y], 1u);
    is_last = (t == gridDim.x - 1);
    __threadfence();
  }
  __syncthreads();
  if (!is_last) return;
  __shared__ double fin[BN_FIN_DOUBLES];
  double B, G;
  final_reduce(partials + (size_t)blockIdx.y * gridDim.x * 2 * chunkC, gridDim.x, chunkC, fin, B, G);
  const int c_local = threadIdx.x;
  const int c = blockIdx.y * chunkC + c_local;
  if (c_local < chunkC && c < C) {
    dbeta[c] = (float)B;
    dgamma[c] = (float)G;
  }
  if (threadIdx.x == 0) counters[blockIdx.y] = 0;
}

// ------------------------------------------------------------------------------------------
// backward: dx
// ------------------------------------------------------------------------------------------
template <typename T, bool ADD, bool FOLD>
__global__ void __launch_bounds__(BN_THREADS)
bn_bwd_apply_kernel(const T* __restrict__ dy, const T* __restrict__ x, T* dx, const T* dx_addend, int64_t M, int C,
                    BnLayout L, const float* __restrict__ gamma, const float* __restrict__ beta,
                    const float* __restrict__ mean, const float* __restrict__ rstd, const float* __restrict__ dgamma,
                    const float* __restrict__ dbeta, float leak, BnFold fold) {
  constexpr int V = Vec<T>::N;
  pdl_trigger();
  pdl_wait();
  const int tx = threadIdx.x % L.CW, ty = threadIdx.x / L.CW;
  const int cg = blockIdx.y * L.CW + tx;
  __shared__ double fold_sh[FOLD ? BN_THREADS * 4 : 1];
  __shared__ float fold_k[FOLD ? BN_FOLD_CW * V * 2 : 1];          // (dbeta, dgamma) per channel of the chunk
  if (FOLD) {
    const int c0 = blockIdx.y * L.CW * V;
    const int nch = min(L.CW * V, C - c0);
    double s[4];
    int pr;
    if (cta_sum_partial_rows(fold.partials, fold.rows, C, c0, nch, fold_sh, s, &pr)) {
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int cl = 2 * pr + h;
        const float db = (float)s[2 * h], dg = (float)s[2 * h + 1];   // (sum g, sum g * xhat), as bn_bwd_finalize_kernel stores them
        fold_k[2 * cl] = db; fold_k[2 * cl + 1] = dg;
        if (blockIdx.x == 0) { fold.o1[c0 + cl] = db; fold.o0[c0 + cl] = dg; }
      }
    }
    __syncthreads();
  }
  if (!(ty < L.RY && cg < L.CV)) return;
  float a[V], b[V], mu[V], rs[V], k1[V], k2[V];
  const float invM = 1.f / (float)M;
#pragma unroll
  for (int i = 0; i < V; ++i) {
    const int c = cg * V + i;
    rs[i] = rstd[c];
    mu[i] = mean[c];
    a[i] = gamma[c] * rs[i];
    b[i] = beta[c] - mu[i] * a[i];
    k1[i] = (FOLD ? fold_k[2 * (tx * V + i)] : dbeta[c]) * invM;
    k2[i] = (FOLD ? fold_k[2 * (tx * V + i) + 1] : dgamma[c]) * invM;
  }
  const int64_t stride = L.RY;
  const int64_t col = (int64_t)cg * V;
  const int64_t rows_per_cta = ((M + gridDim.x - 1) / gridDim.x + L.RY - 1) / L.RY * L.RY;
  const int64_t M_end = min(M, (int64_t)(blockIdx.x + 1) * rows_per_cta);
  int64_t r = (int64_t)blockIdx.x * rows_per_cta + ty;
  constexpr int U = 2;
  for (; r + (U - 1) * stride < M_end; r += U * stride) {
    Vec<T> vx[U], vg[U], va[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      vx[u].load_stream(x + (r + u * stride) * C + col);
      vg[u].load_stream(dy + (r + u * stride) * C + col);
      if (ADD) va[u].load(dx_addend + (r + u * stride) * C + col);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      float fx[V], fg[V], fo[V];
      vx[u].unpack(fx);
      vg[u].unpack(fg);
      if (ADD) va[u].unpack(fo);
#pragma unroll
      for (int i = 0; i < V; ++i) {
        const float pre = fmaf(fx[i], a[i], b[i]);
        const float g = pre < 0.f ? leak * fg[i] : fg[i];
        const float xhat = (fx[i] - mu[i]) * rs[i];
        const float d = a[i] * (g - k1[i] - xhat * k2[i]);
        fo[i] = ADD ? fo[i] + d : d;
      }
      Vec<T> vo;
      vo.pack(fo);
      vo.store(dx + (r + u * stride) * C + col);
    }
  }
  for (; r < M_end; r += stride) {
    const int64_t off = r * C + col;
    Vec<T> vx, vg;
    float fx[V], fg[V], fo[V];
    vx.load_stream(x + off);
    vg.load_stream(dy + off);
    vx.unpack(fx);
    vg.unpack(fg);
    if (ADD) {
      Vec<T> va;
      va.load(dx_addend + off);
      va.unpack(fo);
    }
#pragma unroll
    for (int i = 0; i < V; ++i) {
      const float pre = fmaf(fx[i], a[i], b[i]);
      const float g = pre < 0.f ? leak * fg[i] : fg[i];
      const float xhat = (fx[i] - mu[i]) * rs[i];
      const float d = a[i] * (g - k1[i] - xhat * k2[i]);
      fo[i] = ADD ? fo[i] + d : d;
    }
    Vec<T> vo;
    vo.pack(fo);
    vo.store(dx + off);
  }
}

template <typename T>
static int check_bn_args(int64_t M, int C) {
  if (M <= 0 || C <= 0) return fail(RN_ERR_INVALID, "bn: M=%lld C=%d must be positive", (long long)M, C);
  if (C % Vec<T>::N != 0) return fail(RN_ERR_UNSUPPORTED, "bn: C=%d must be a multiple of %d", C, Vec<T>::N);
  if ((C / Vec<T>::N + 31) / 32 > BN_MAX_CHUNKS) return fail(RN_ERR_UNSUPPORTED, "bn: C=%d too large", C);
  return RN_OK;
}

static size_t bn_scratch_bytes(int64_t M, int C) {
  (void)M;
  (void)C;
  // 256 B of ticket counters + one [2][<=256 channels] fp32 partial per CTA; the layouts above
  // never launch more than BN_MAX_G + BN_MAX_CHUNKS CTAs in total.
  return 256 + (size_t)(BN_MAX_G + BN_MAX_CHUNKS) * 2 * 256 * sizeof(float);
}

template <typename T>
static int bn_stats_launch(const void* x, const void* addend, void* sum_out, int64_t M, int C, float eps, float decay,
                           int unbiased, float* mean, float* rstd, float* mm, float* mv, void* scratch,
                           size_t scratch_bytes, cudaStream_t st) {
  int rc = check_bn_args<T>(M, C);
  if (rc) return rc;
  BnLayout L = bn_layout<T>(M, C, bn_reduce_waves(4), bn_reduce_cw(), bn_reduce_cta_bytes());
  L.il = bn_interleave();
  const size_t need = 256 + (size_t)L.gridY * L.G * 2 * L.CW * Vec<T>::N * sizeof(float);
  if (scratch == nullptr || scratch_bytes < need)
    return fail(RN_ERR_SCRATCH, "bn_stats: scratch %zu < %zu bytes", scratch_bytes, need);
  unsigned* counters = (unsigned*)scratch;
  float* partials = (float*)((char*)scratch + 256);
  dim3 grid(L.G, L.gridY);
  size_t smem = (size_t)L.RY * L.CW * 2 * Vec<T>::N * sizeof(float);
  static int u_plain = -1, u_add = -1;
  if (u_plain < 0) { const char* e = getenv("RN_BN_STATS_U"); u_plain = e ? atoi(e) : 2; }
  if (u_add < 0) { const char* e = getenv("RN_BN_STATS_ADD_U"); u_add = e ? atoi(e) : 4; }
#define RN_LAUNCH_STATS(ADDV, UU)                                                                              \
  launch_pdl(bn_stats_kernel<T, ADDV, UU>, grid, BN_THREADS, smem, st, (const T*)x, (const T*)addend, (T*)sum_out, M, \
             C, L, eps, decay, unbiased, mean, rstd, mm, mv, counters, partials)
  if (addend != nullptr) {
    RN_REQUIRE(sum_out != nullptr, "bn_stats: addend given but sum_out is NULL");
    switch (u_add) {
      case 2: RN_LAUNCH_STATS(true, 2); break;
      case 4: RN_LAUNCH_STATS(true, 4); break;
      default: RN_LAUNCH_STATS(true, 1); break;
    }
  } else {
    switch (u_plain) {
      case 1: RN_LAUNCH_STATS(false, 1); break;
      case 4: RN_LAUNCH_STATS(false, 4); break;
      case 8: RN_LAUNCH_STATS(false, 8); break;
      default: RN_LAUNCH_STATS(false, 2); break;
    }
  }
#undef RN_LAUNCH_STATS
  return check_launch("bn_stats_kernel");
}

}  // namespace rn

using namespace rn;

// Finalisation of statistics whose per-CTA sums were produced by a convolution epilogue
// (rn_convtc_plan_create_fused): partials[nparts][C][2] = (sum, sum of squares) per channel.
// A CTA handles 8 channels with 32 thread slices over the partial rows: every thread issues its <= 5 loads
// (148 rows / 32 slices) at once, so the kernel is one memory round trip; slices are then combined in a fixed
// order in fp64.
namespace rn {
constexpr int FIN_CH = 8, FIN_SLICES = 32;
__global__ void __launch_bounds__(256)
bn_finalize_kernel(const float* __restrict__ partials, int nparts, int64_t M, int C, float eps, float decay, int unbiased,
                   float* __restrict__ mean_out, float* __restrict__ rstd_out, float* moving_mean, float* moving_var) {
  pdl_trigger();
  pdl_wait();
  __shared__ double sh[2][FIN_SLICES][FIN_CH];
  const int cl = threadIdx.x % FIN_CH, slice = threadIdx.x / FIN_CH;
  const int c = blockIdx.x * FIN_CH + cl;
  double s = 0.0, q = 0.0;
  if (c < C) {
    const float2* p = reinterpret_cast<const float2*>(partials) + c;
    int z = slice;
    for (; z + 3 * FIN_SLICES < nparts; z += 4 * FIN_SLICES) {
      const float2 a = __ldcg(p + (size_t)z * C), b = __ldcg(p + (size_t)(z + FIN_SLICES) * C);
      const float2 d = __ldcg(p + (size_t)(z + 2 * FIN_SLICES) * C), e = __ldcg(p + (size_t)(z + 3 * FIN_SLICES) * C);
      s += (double)a.x; q += (double)a.y;
      s += (double)b.x; q += (double)b.y;
      s += (double)d.x; q += (double)d.y;
      s += (double)e.x; q += (double)e.y;
    }
    float2 t[3];
#pragma unroll
    for (int u = 0; u < 3; ++u) {
      const int zz = z + u * FIN_SLICES;
      t[u] = zz < nparts ? __ldcg(p + (size_t)zz * C) : make_float2(0.f, 0.f);
    }
#pragma unroll
    for (int u = 0; u < 3; ++u) { s += (double)t[u].x; q += (double)t[u].y; }
  }
  sh[0][slice][cl] = s;
  sh[1][slice][cl] = q;
  __syncthreads();
  if (slice == 0 && c < C) {
    double S = 0.0, Q = 0.0;
    for (int j = 0; j < FIN_SLICES; ++j) { S += sh[0][j][cl]; Q += sh[1][j][cl]; }
    bn_finalize_channel(S, Q, (double)M, eps, decay, unbiased, c, mean_out, rstd_out, moving_mean, moving_var);
  }
}
}  // namespace rn

extern "C" int rn_bn_finalize(const float* partials, int32_t nparts, int64_t M, int32_t C, float eps, float decay,
                              int32_t unbiased, float* mean, float* rstd, float* moving_mean, float* moving_var,
                              void* stream) {
  RN_REQUIRE(partials && mean && rstd && nparts > 0 && M > 0 && C > 0, "bn_finalize: bad arguments");
  RN_REQUIRE((moving_mean == nullptr) == (moving_var == nullptr), "bn_finalize: moving_mean/var must both be set or NULL");
  RN_REQUIRE(((uintptr_t)partials & 7) == 0, "bn_finalize: partials must be 8-byte aligned");
  launch_pdl(rn::bn_finalize_kernel, (C + rn::FIN_CH - 1) / rn::FIN_CH, 256, 0, rn::as_stream(stream), partials, (int)nparts, M, (int)C, eps,
             decay, (int)unbiased, mean, rstd, moving_mean, moving_var);
  return rn::check_launch("bn_finalize_kernel");
}

// dgamma / dbeta from the per-CTA sums a fused dgrad epilogue produced (rn_convtc_plan_create_fused_dgrad):
// partials[nparts][C][2] = (sum g, sum g * xhat).  Same shape of work as bn_finalize_kernel.
namespace rn {
__global__ void __launch_bounds__(256)
bn_bwd_finalize_kernel(const float* __restrict__ partials, int nparts, int C, float* __restrict__ dgamma,
                       float* __restrict__ dbeta) {
  pdl_trigger();
  pdl_wait();
  __shared__ double sh[2][FIN_SLICES][FIN_CH];
  const int cl = threadIdx.x % FIN_CH, slice = threadIdx.x / FIN_CH;
  const int c = blockIdx.x * FIN_CH + cl;
  double s = 0.0, q = 0.0;
  if (c < C) {
    const float2* p = reinterpret_cast<const float2*>(partials) + c;
    for (int z0 = slice; z0 < nparts; z0 += 4 * FIN_SLICES) {
      float2 t[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int zz = z0 + u * FIN_SLICES;
        t[u] = zz < nparts ? __ldcg(p + (size_t)zz * C) : make_float2(0.f, 0.f);
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) { s += (double)t[u].x; q += (double)t[u].y; }
    }
  }
  sh[0][slice][cl] = s;
  sh[1][slice][cl] = q;
  __syncthreads();
  if (slice == 0 && c < C) {
    double S = 0.0, Q = 0.0;
    for (int j = 0; j < FIN_SLICES; ++j) { S += sh[0][j][cl]; Q += sh[1][j][cl]; }
    dbeta[c] = (float)S;
    dgamma[c] = (float)Q;
  }
}
}  // namespace rn

extern "C" int rn_bn_bwd_finalize(const float* partials, int32_t nparts, int32_t C, float* dgamma, float* dbeta,
                                  void* stream) {
  RN_REQUIRE(partials && dgamma && dbeta && nparts > 0 && C > 0, "bn_bwd_finalize: bad arguments");
  RN_REQUIRE(((uintptr_t)partials & 7) == 0, "bn_bwd_finalize: partials must be 8-byte aligned");
  launch_pdl(rn::bn_bwd_finalize_kernel, (C + rn::FIN_CH - 1) / rn::FIN_CH, 256, 0, rn::as_stream(stream), partials,
             (int)nparts, (int)C, dgamma, dbeta);
  return rn::check_launch("bn_bwd_finalize_kernel");
}

extern "C" size_t rn_bn_scratch_bytes(int64_t M, int32_t C) { return bn_scratch_bytes(M, C); }

extern "C" int rn_bn_stats_fwd(const void* x, const void* addend, void* sum_out, int64_t M, int32_t C, int32_t dtype,
                               float eps, float decay, int32_t unbiased, float* mean, float* rstd, float* moving_mean,
                               float* moving_var, void* scratch, size_t scratch_bytes, void* stream) {
  RN_REQUIRE(x && mean && rstd, "bn_stats: NULL pointer");
  RN_REQUIRE((((uintptr_t)x | (uintptr_t)addend | (uintptr_t)sum_out) & 15) == 0, "bn_stats: x / addend / sum_out must be 16-byte aligned");
  RN_REQUIRE((moving_mean == nullptr) == (moving_var == nullptr), "bn_stats: moving_mean/var must both be set or NULL");
  if (dtype == RN_F32)
    return bn_stats_launch<float>(x, addend, sum_out, M, C, eps, decay, unbiased, mean, rstd, moving_mean, moving_var,
                                  scratch, scratch_bytes, as_stream(stream));
  if (dtype == RN_BF16)
    return bn_stats_launch<__nv_bfloat16>(x, addend, sum_out, M, C, eps, decay, unbiased, mean, rstd, moving_mean,
                                          moving_var, scratch, scratch_bytes, as_stream(stream));
  return fail(RN_ERR_INVALID, "bn_stats: bad dtype %d", dtype);
}

template <typename T>
static int bn_apply_launch(const void* x, void* y, int64_t M, int C, const float* gamma, const float* beta,
                           const float* mean, const float* rstd, int moving, float eps, float leak, cudaStream_t st,
                           const BnFold* fold = nullptr) {
  int rc = check_bn_args<T>(M, C);
  if (rc) return rc;
  static int awaves = -1;
  if (awaves < 0) { const char* e = getenv("RN_BN_APPLY_WAVES"); awaves = e ? atoi(e) : 4; }   // 56 registers: 4 resident CTAs per SM
  if (fold) {
    BnLayout L = bn_layout<T>(M, C, awaves, BN_FOLD_CW, 16384);
    launch_pdl(bn_apply_kernel<T, true>, dim3(L.G, L.gridY), BN_THREADS, 0, st, (const T*)x, (T*)y, M, C, L, gamma, beta,
               mean, rstd, 0, eps, leak, *fold);
    return check_launch("bn_apply_kernel (folded finalisation)");
  }
  BnLayout L = bn_layout<T>(M, C, awaves, bn_apply_cw(), 16384);
  launch_pdl(bn_apply_kernel<T, false>, dim3(L.G, L.gridY), BN_THREADS, 0, st, (const T*)x, (T*)y, M, C, L, gamma, beta, mean,
             rstd, moving, eps, leak, BnFold{});
  return check_launch("bn_apply_kernel");
}

extern "C" int rn_bn_apply_act_fwd(const void* x, void* y, int64_t M, int32_t C, int32_t dtype, const float* gamma,
                                   const float* beta, const float* mean, const float* rstd, int32_t stats_are_moving,
                                   float eps, float leak, void* stream) {
  RN_REQUIRE(x && y && gamma && beta && mean && rstd, "bn_apply: NULL pointer");
  RN_REQUIRE((((uintptr_t)x | (uintptr_t)y) & 15) == 0, "bn_apply: x / y must be 16-byte aligned");
  if (dtype == RN_F32)
    return bn_apply_launch<float>(x, y, M, C, gamma, beta, mean, rstd, stats_are_moving, eps, leak, as_stream(stream));
  if (dtype == RN_BF16)
    return bn_apply_launch<__nv_bfloat16>(x, y, M, C, gamma, beta, mean, rstd, stats_are_moving, eps, leak,
                                          as_stream(stream));
  return fail(RN_ERR_INVALID, "bn_apply: bad dtype %d", dtype);
}

template <typename T>
static int bn_bwd_reduce_launch(const void* dy, const void* x, int64_t M, int C, const float* gamma, const float* beta,
                                const float* mean, const float* rstd, float leak, float* dgamma, float* dbeta,
                                void* scratch, size_t scratch_bytes, cudaStream_t st) {
  int rc = check_bn_args<T>(M, C);
  if (rc) return rc;
  BnLayout L = bn_layout<T>(M, C, bn_reduce_waves(3), bn_reduce_cw(), bn_reduce_cta_bytes());
  L.il = bn_interleave();
  const size_t need = 256 + (size_t)L.gridY * L.G * 2 * L.CW * Vec<T>::N * sizeof(float);
  if (scratch == nullptr || scratch_bytes < need)
    return fail(RN_ERR_SCRATCH, "bn_bwd_reduce: scratch %zu < %zu bytes", scratch_bytes, need);
  size_t smem = (size_t)L.RY * L.CW * 2 * Vec<T>::N * sizeof(float);
  static int variant = -1;
  if (variant < 0) { const char* e = getenv("RN_BN_VARIANT"); variant = e ? atoi(e) : 0; }
#define RN_LAUNCH_REDUCE(UU, CC)                                                                              \
  launch_pdl(bn_bwd_reduce_kernel<T, UU, CC>, dim3(L.G, L.gridY), BN_THREADS, smem, st, (const T*)dy, (const T*)x, M,   \
             C, L, gamma, beta, mean, rstd, leak, dgamma, dbeta, (unsigned*)scratch, (float*)((char*)scratch + 256))
  // streaming (evict-first) loads measured +13 % on tensors that exceed L2; smaller tensors keep the
  // default policy so that the dx pass that follows finds x and dy in L2
  const bool big = (size_t)M * C * sizeof(T) > ((size_t)48 << 20);
  switch (variant) {
    case 1: RN_LAUNCH_REDUCE(2, true); break;
    case 2: RN_LAUNCH_REDUCE(4, false); break;
    case 3: RN_LAUNCH_REDUCE(4, true); break;
    case 4: RN_LAUNCH_REDUCE(1, false); break;
    case 5: RN_LAUNCH_REDUCE(2, false); break;
    default:
      if (big) RN_LAUNCH_REDUCE(2, true);
      else RN_LAUNCH_REDUCE(2, false);
      break;
  }
#undef RN_LAUNCH_REDUCE
  return check_launch("bn_bwd_reduce_kernel");
}

extern "C" int rn_bn_act_bwd_reduce(const void* dy, const void* x, int64_t M, int32_t C, int32_t dtype,
                                    const float* gamma, const float* beta, const float* mean, const float* rstd,
                                    float leak, float* dgamma, float* dbeta, void* scratch, size_t scratch_bytes,
                                    void* stream) {
  RN_REQUIRE(dy && x && gamma && beta && mean && rstd && dgamma && dbeta, "bn_bwd_reduce: NULL pointer");
  RN_REQUIRE((((uintptr_t)dy | (uintptr_t)x) & 15) == 0, "bn_bwd_reduce: dy / x must be 16-byte aligned");
  if (dtype == RN_F32)
    return bn_bwd_reduce_launch<float>(dy, x, M, C, gamma, beta, mean, rstd, leak, dgamma, dbeta, scratch,
                                       scratch_bytes, as_stream(stream));
  if (dtype == RN_BF16)
    return bn_bwd_reduce_launch<__nv_bfloat16>(dy, x, M, C, gamma, beta, mean, rstd, leak, dgamma, dbeta, scratch,
                                               scratch_bytes, as_stream(stream));
  return fail(RN_ERR_INVALID, "bn_bwd_reduce: bad dtype %d", dtype);
}

template <typename T>
static int bn_bwd_apply_launch(const void* dy, const void* x, void* dx, const void* addend, int64_t M, int C,
                               const float* gamma, const float* beta, const float* mean, const float* rstd,
                               const float* dgamma, const float* dbeta, float leak, cudaStream_t st,
                               const BnFold* fold = nullptr) {
  int rc = check_bn_args<T>(M, C);
  if (rc) return rc;
  static int bwaves = -1, bkb = -1;
  if (bwaves < 0) { const char* e = getenv("RN_BN_BWD_APPLY_WAVES"); bwaves = e ? atoi(e) : 3; }   // 80 registers: 3 resident CTAs per SM = one full wave
  if (bkb < 0) { const char* e = getenv("RN_BN_BWD_APPLY_KB"); bkb = e ? atoi(e) : 16; }
  BnLayout L = bn_layout<T>(M, C, bwaves, fold ? BN_FOLD_CW : bn_apply_cw(), (int64_t)bkb * 1024);
  dim3 grid(L.G, L.gridY);
  const BnFold f = fold ? *fold : BnFold{};
#define RN_LAUNCH_BWD_APPLY(ADD_, FOLD_)                                                                              \
  launch_pdl(bn_bwd_apply_kernel<T, ADD_, FOLD_>, grid, BN_THREADS, 0, st, (const T*)dy, (const T*)x, (T*)dx,        \
             (const T*)addend, M, C, L, gamma, beta, mean, rstd, dgamma, dbeta, leak, f)
  if (addend) { if (fold) RN_LAUNCH_BWD_APPLY(true, true); else RN_LAUNCH_BWD_APPLY(true, false); }
  else        { if (fold) RN_LAUNCH_BWD_APPLY(false, true); else RN_LAUNCH_BWD_APPLY(false, false); }
#undef RN_LAUNCH_BWD_APPLY
  return check_launch("bn_bwd_apply_kernel");
}

extern "C" int rn_bn_act_bwd_apply(const void* dy, const void* x, void* dx, const void* dx_addend, int64_t M, int32_t C,
                                   int32_t dtype, const float* gamma, const float* beta, const float* mean,
                                   const float* rstd, const float* dgamma, const float* dbeta, float leak,
                                   void* stream) {
  RN_REQUIRE(dy && x && dx && gamma && beta && mean && rstd && dgamma && dbeta, "bn_bwd_apply: NULL pointer");
  RN_REQUIRE((((uintptr_t)dy | (uintptr_t)x | (uintptr_t)dx | (uintptr_t)dx_addend) & 15) == 0,
             "bn_bwd_apply: dy / x / dx / dx_addend must be 16-byte aligned");
  if (dtype == RN_F32)
    return bn_bwd_apply_launch<float>(dy, x, dx, dx_addend, M, C, gamma, beta, mean, rstd, dgamma, dbeta, leak,
                                      as_stream(stream));
  if (dtype == RN_BF16)
    return bn_bwd_apply_launch<__nv_bfloat16>(dy, x, dx, dx_addend, M, C, gamma, beta, mean, rstd, dgamma, dbeta, leak,
                                              as_stream(stream));
  return fail(RN_ERR_INVALID, "bn_bwd_apply: bad dtype %d", dtype);
}

// The apply passes with the finalisation of a fused convolution epilogue's partial sums folded in (see BnFold).
extern "C" int rn_bn_apply_act_fwd_partials(const void* x, void* y, int64_t M, int32_t C, int32_t dtype, const float* gamma,
                                            const float* beta, const float* partials, int32_t nparts, float eps,
                                            float decay, int32_t unbiased, float* mean, float* rstd, float* moving_mean,
                                            float* moving_var, float leak, void* stream) {
  RN_REQUIRE(x && y && gamma && beta && partials && mean && rstd && nparts > 0, "bn_apply_partials: bad arguments");
  RN_REQUIRE((((uintptr_t)x | (uintptr_t)y | (uintptr_t)partials) & 15) == 0, "bn_apply_partials: x / y / partials must be 16-byte aligned");
  RN_REQUIRE((moving_mean == nullptr) == (moving_var == nullptr), "bn_apply_partials: moving_mean/var must both be set or NULL");
  const BnFold f{partials, nparts, (double)M, eps, decay, unbiased, mean, rstd, moving_mean, moving_var};
  if (dtype == RN_F32)
    return bn_apply_launch<float>(x, y, M, C, gamma, beta, nullptr, nullptr, 0, eps, leak, as_stream(stream), &f);
  if (dtype == RN_BF16)
    return bn_apply_launch<__nv_bfloat16>(x, y, M, C, gamma, beta, nullptr, nullptr, 0, eps, leak, as_stream(stream), &f);
  return fail(RN_ERR_INVALID, "bn_apply_partials: bad dtype %d", dtype);
}

extern "C" int rn_bn_act_bwd_apply_partials(const void* dy, const void* x, void* dx, const void* dx_addend, int64_t M,
                                            int32_t C, int32_t dtype, const float* gamma, const float* beta,
                                            const float* mean, const float* rstd, const float* partials, int32_t nparts,
                                            float* dgamma, float* dbeta, float leak, void* stream) {
  RN_REQUIRE(dy && x && dx && gamma && beta && mean && rstd && partials && dgamma && dbeta && nparts > 0,
             "bn_bwd_apply_partials: bad arguments");
  RN_REQUIRE((((uintptr_t)dy | (uintptr_t)x | (uintptr_t)dx | (uintptr_t)dx_addend | (uintptr_t)partials) & 15) == 0,
             "bn_bwd_apply_partials: dy / x / dx / dx_addend / partials must be 16-byte aligned");
  const BnFold f{partials, nparts, (double)M, 0.f, 0.f, 0, dgamma, dbeta, nullptr, nullptr};
  if (dtype == RN_F32)
    return bn_bwd_apply_launch<float>(dy, x, dx, dx_addend, M, C, gamma, beta, mean, rstd, nullptr, nullptr, leak,
                                      as_stream(stream), &f);
  if (dtype == RN_BF16)
    return bn_bwd_apply_launch<__nv_bfloat16>(dy, x, dx, dx_addend, M, C, gamma, beta, mean, rstd, nullptr, nullptr, leak,
                                              as_stream(stream), &f);
  return fail(RN_ERR_INVALID, "bn_bwd_apply_partials: bad dtype %d", dtype);
}

// ------------------------------------------------------------------------------------------
// rn_bn_stats_fwd_noalloc — the forward statistics pass for callers that cannot hand us scratch.
//
// rn_bn_stats_fwd needs a caller-allocated scratch buffer for its per-CTA partial sums and ticket
// counters.  Some hosts have no way to provide one: a plugin loaded into a process whose framework
// has already claimed (nearly) all device memory for its own pool (TensorFlow 1.x, the reference's
// runtime, does this by default, so a cudaMalloc from a side library fails), or a latency-critical
// loop that must not risk cudaMalloc's implicit device-wide synchronisation.  For them the scratch
// lives in the fixed device array below, which is part of the library image: every GPU that loads
// librn_b200.so holds these 512 KiB until the process exits, whether or not the entry is used.
//
// The array's name is only a host-side handle; its device address is looked up once per device
// (with that device current) through cudaGetSymbolAddress on the FIRST call made on that device —
// make that first call during set-up, before any stream capture is begun.  On EVERY call the device
// that owns `stream` must be the current device: the array used is the current device's.  The entry point calls
// no cudaMalloc / cudaFree and no allocating library routine.
//
// All calls on a device share the one array and it keeps whatever the previous call left behind:
// the ticket counters are cleared at the start of every call (stream-ordered memset), the partial
// sums are fully overwritten before they are read.  Two calls that overlap on the GPU (different
// streams, or different host threads without ordering) race on it — serialise them.
// Limits: C <= RN_BN_NOALLOC_MAX_C channels (any M); larger C is rejected with RN_ERR_INVALID.
// ------------------------------------------------------------------------------------------
#include <mutex>

#define RN_BN_NOALLOC_SCRATCH_BYTES (512 * 1024)
#define RN_BN_NOALLOC_MAX_C 4096
#define RN_BN_NOALLOC_MAX_DEVICES 64

static __device__ __align__(256) unsigned char g_bn_noalloc_scratch[RN_BN_NOALLOC_SCRATCH_BYTES];

namespace rn {
static std::mutex g_noalloc_mu;
static void* g_noalloc_dev_ptr[RN_BN_NOALLOC_MAX_DEVICES] = {nullptr};

// Device address of the static scratch on the CURRENT device (resolved once per device).
static int bn_noalloc_scratch_ptr(void** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return fail(RN_ERR_CUDA, "bn_stats_noalloc: cudaGetDevice: %s", cudaGetErrorString(e));
  if (dev < 0 || dev >= RN_BN_NOALLOC_MAX_DEVICES)
    return fail(RN_ERR_UNSUPPORTED, "bn_stats_noalloc: device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lock(g_noalloc_mu);
  if (g_noalloc_dev_ptr[dev] == nullptr) {
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_bn_noalloc_scratch);
    if (e != cudaSuccess || p == nullptr)
      return fail(RN_ERR_CUDA, "bn_stats_noalloc: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
    g_noalloc_dev_ptr[dev] = p;
  }
  *out = g_noalloc_dev_ptr[dev];
  return RN_OK;
}

template <typename T>
static int bn_stats_noalloc_launch(const void* x, const void* addend, void* sum_out, int64_t M, int C, float eps,
                                   float decay, int unbiased, float* mean, float* rstd, float* mm, float* mv,
                                   cudaStream_t st) {
  int rc = check_bn_args<T>(M, C);
  if (rc) return rc;
  BnLayout L = bn_layout<T>(M, C, 4);
  // shrink the number of row CTAs until the partial sums fit the fixed array
  const size_t per_cta = (size_t)2 * L.CW * Vec<T>::N * sizeof(float);
  const size_t max_g = (RN_BN_NOALLOC_SCRATCH_BYTES - 256) / ((size_t)L.gridY * per_cta);
  if (max_g < 1) return fail(RN_ERR_INVALID, "bn_stats_noalloc: C=%d does not fit the static scratch", C);
  if ((size_t)L.G > max_g) L.G = (int)max_g;
  void* base = nullptr;
  rc = bn_noalloc_scratch_ptr(&base);
  if (rc) return rc;
  cudaError_t e = cudaMemsetAsync(base, 0, 256, st);   // ticket counters must start at zero
  if (e != cudaSuccess) return fail(RN_ERR_CUDA, "bn_stats_noalloc: cudaMemsetAsync: %s", cudaGetErrorString(e));
  unsigned* counters = (unsigned*)base;
  float* partials = (float*)((char*)base + 256);
  dim3 grid(L.G, L.gridY);
  const size_t smem = (size_t)L.RY * L.CW * 2 * Vec<T>::N * sizeof(float);
  if (addend != nullptr)
    bn_stats_kernel<T, true, 1><<<grid, BN_THREADS, smem, st>>>((const T*)x, (const T*)addend, (T*)sum_out, M, C, L, eps,
                                                             decay, unbiased, mean, rstd, mm, mv, counters, partials);
  else
    bn_stats_kernel<T, false, 2><<<grid, BN_THREADS, smem, st>>>((const T*)x, nullptr, nullptr, M, C, L, eps, decay,
                                                              unbiased, mean, rstd, mm, mv, counters, partials);
  return check_launch("bn_stats_kernel (static scratch)");
}
}  // namespace rn

extern "C" int rn_bn_stats_fwd_noalloc(const void* x, const void* addend, void* sum_out, int64_t M, int32_t C,
                                       int32_t dtype, float eps, float decay, int32_t unbiased, float* mean,
                                       float* rstd, float* moving_mean, float* moving_var, void* stream) {
  RN_REQUIRE(x && mean && rstd, "bn_stats_noalloc: NULL pointer");
  RN_REQUIRE((((uintptr_t)x | (uintptr_t)addend | (uintptr_t)sum_out) & 15) == 0,
             "bn_stats_noalloc: x / addend / sum_out must be 16-byte aligned");
  RN_REQUIRE((moving_mean == nullptr) == (moving_var == nullptr),
             "bn_stats_noalloc: moving_mean/var must both be set or NULL");
  RN_REQUIRE(addend == nullptr || sum_out != nullptr, "bn_stats_noalloc: addend given but sum_out is NULL");
  if (C > RN_BN_NOALLOC_MAX_C)
    return fail(RN_ERR_INVALID, "bn_stats_noalloc: C=%d exceeds the static-scratch limit of %d channels; use "
                "rn_bn_stats_fwd with caller scratch", C, RN_BN_NOALLOC_MAX_C);
  if (dtype == RN_F32)
    return bn_stats_noalloc_launch<float>(x, addend, sum_out, M, C, eps, decay, unbiased, mean, rstd, moving_mean,
                                          moving_var, as_stream(stream));
  if (dtype == RN_BF16)
    return bn_stats_noalloc_launch<__nv_bfloat16>(x, addend, sum_out, M, C, eps, decay, unbiased, mean, rstd,
                                                  moving_mean, moving_var, as_stream(stream));
  return fail(RN_ERR_INVALID, "bn_stats_noalloc: bad dtype %d", dtype);
}
