// tcgen05 / TMA / TMEM implicit-GEMM convolution for sm_100a (bf16 operands, fp32 accumulation).
//
// fprop and dgrad are the same GEMM:   OUT[pixel, col] = sum_{tap, c} ACT[pixel shifted by tap, c] * WT[tap][col][c]
//   fprop: ACT = x,  WT = w as [tap][Cout][Cin], OUT = y           (tap shift  s - pad, r - pad)
//   dgrad: ACT = dy, WT = w as [tap][Cin][Cout], OUT = dx          (tap shift  pad - s, pad - r)
// A CTA owns a 128-pixel x BLOCK_N tile.  The 128 pixels are a BW x BH x BN box of the output
// (w fastest), so the A operand of every tap is ONE 4-D TMA box load of the activation tensor at
// shifted coordinates; rows that fall outside the image are zero-filled by TMA, which is exactly
// the reference's fixed_padding / SAME semantics (resnet_model_official.py:53-91) — no padded
// copy, no im2col buffer.  The K loop runs over (tap, 64-channel chunk); operands land in
// 128B-swizzled K-major shared-memory tiles that tcgen05.mma reads through matrix descriptors;
// the fp32 accumulator lives in tensor memory (double-buffered: the epilogue of tile i overlaps
// the MMAs of tile i+1); the epilogue converts to bf16, stages 64-channel groups in swizzled
// shared memory and writes them with TMA box stores (out-of-image rows are clipped by TMA).
//
// Warp roles (192 threads): warp 0 = TMA producer, warp 1 = MMA issuer + TMEM allocator,
// warps 2-5 = epilogue (one TMEM lane quadrant each).  Persistent grid: one CTA per SM.
#include "rn_common.cuh"
#include "rn_tc_common.cuh"

#include <string.h>
#include <stdlib.h>
#include <vector>

namespace rn {
using namespace tc;

constexpr int TC_THREADS = 192;           // wgrad kernel: producer, MMA issuer, 4 epilogue warps
constexpr int TC_EPI_WARPS = 8;           // fprop / dgrad kernel: two epilogue warps per TMEM lane quadrant (half a row each)
constexpr int TC_EPI_THREADS = TC_EPI_WARPS * 32;
constexpr int TC_CONV_THREADS = 64 + TC_EPI_THREADS;
constexpr int TC_MAX_TAPS = 52;
constexpr int TC_BLOCK_M = 128;

struct TcConvParams {
  CUtensorMap a_map[4];
  CUtensorMap b_map;
  CUtensorMap o_map;
  int tiles_w, tiles_h, tiles_n;  // pixel tiles along W, H, N
  int BW, BH, BN;                 // pixel box (BW*BH*BN <= 128)
  int n_tiles;                    // column tiles
  int block_n;                    // UMMA N
  int kc;                         // reduction elements per stage (64 / 32 / 16)
  int k_chunks;                   // ceil(Cred / kc)
  int taps;
  int stages;
  int gw;                         // channels per epilogue store group (64 / 32 / 16)
  uint32_t a_bytes, b_bytes;      // stage buffer sizes (1024-aligned)
  uint32_t tx_bytes;              // bytes TMA delivers per stage
  uint32_t tmem_cols;
  int accumulate;                 // 1: OUT += result (TMA reduce-add), 0: OUT = result
  int acc_m;                      // pixel tiles per CTA iteration (they share every B tile): 1, 2 or 4
  int tstages;                    // TMEM accumulator sets (2 = epilogue overlaps the next tile's MMAs)
  int b_mn;                       // 1: B operand is MN-major (fprop reads the HWIO weights as they are)
  int b_boxes;                    // MN-major: TMA boxes per stage and CTA (block_n / gw, halved for a CTA pair), each kc rows of gw channels
  uint32_t b_box_bytes;
  int pair;                       // 1: launched as clusters of two CTAs issuing tcgen05.mma.cta_group::2 (M = 256: this CTA's 128
                                  // pixels and its peer's; each CTA loads HALF of the B tile)
  // fused epilogue (fprop, 64-channel store groups only)
  int fuse_stats;                 // per-channel sum / sum of squares of the STORED bf16 output -> stat_partials[CTA][ncols][2]
  int fuse_add;                   // OUT = bf16(bf16(result) + residual): the residual tile arrives through r_map (warp 6)
  int fuse_bnbwd;                 // dgrad: the output is dy of a BN+ReLU whose input x arrives through r_map; the epilogue
                                  // accumulates that BN's backward reductions (sum g, sum g*xhat; g = dy masked by the ReLU)
  const float *bn_gamma, *bn_beta, *bn_mean, *bn_rstd;
  float bn_leak;
  int res_slots;                  // tiles in the residual / BN-input ring (16 KB each)
  int ncols_total;
  int oW, oH, oN;                 // output extents (rows of a tile that fall outside are not part of the tensor)
  float* stat_partials;
  // Folded finalisation: the CTA that finishes last (atomic ticket, reset for the next launch) sums the rows of the
  // partial table in fp64 in a fixed order and writes what bn_finalize_kernel (1) / bn_bwd_finalize_kernel (2) would.
  int fin_mode;                   // 0 none, 1 BN statistics (mean, rstd, moving averages), 2 BN backward (dgamma, dbeta)
  int fin_unbiased;
  int* fin_ticket;
  float *fin_o0, *fin_o1;         // mode 1: mean, rstd; mode 2: dgamma, dbeta
  float *fin_mm, *fin_mv;         // mode 1: moving mean / variance (or NULL)
  double fin_M;
  float fin_eps, fin_decay;
  CUtensorMap r_map;
  int8_t tap_map[TC_MAX_TAPS], tap_dw[TC_MAX_TAPS], tap_dh[TC_MAX_TAPS], tap_b[TC_MAX_TAPS];
  // Halo mode (3x3, stride 1): a pipeline stage is one HORIZONTAL tap (and k-chunk).  Its A box carries BH + 2 image
  // rows, loaded once, and the three vertical taps read it at row offsets 0, BW, 2*BW through the matrix descriptor
  // (start address + the descriptor's base-offset field when the offset is not a whole 1024-byte swizzle atom);
  // the stage also holds the three weight tiles.  A traffic per output tile drops from 9 to 3 * (BH + 2) / BH boxes.
  int halo;                       // 0 / 1
  // Operand transform (1x1 fprop): the A operand is the INPUT of a training-mode batch_norm_relu; four extra warps
  // turn every landed A tile into bf16(act(a*x + b)) in place (a = gamma*rstd, b = beta - mean*a, exactly the
  // arithmetic of bn_apply_kernel) before the MMA warp may read it, so the normalised tensor is never written to HBM.
  int xf;                         // 0 / 1
  int xf_channels;                // channels of the BN (= reduction length of the GEMM)
  const float *xf_gamma, *xf_beta, *xf_mean, *xf_rstd;
  float xf_leak;
  int8_t halo_dw[3];              // per stage group: horizontal shift of the box
  int8_t halo_dh0;                // vertical shift of the box origin (the topmost tap)
  int8_t halo_row[3][3];          // [group][j]: first row of the box the j-th vertical tap of the group reads, in units of BW rows
  int8_t halo_b[3][3];            // [group][j]: weight tap index
};

// Last-CTA finalisation of the partial table [rows][C][2] by the 256 epilogue threads: a thread owns a channel PAIR
// (one 16-byte load per row) and a slice of the rows (rows slice, slice + S, ...; eight loads in flight), the slices
// are combined in slice order through shared memory (`sh`: 256 x 4 doubles).
__device__ __forceinline__ void fold_finalize(const TcConvParams& p, int te, int rows, double* sh) {
  const int C2 = p.ncols_total >> 1;
  const float4* tab = reinterpret_cast<const float4*>(p.stat_partials);
  const int lanes = C2 < TC_EPI_THREADS ? C2 : TC_EPI_THREADS;
  const int S = TC_EPI_THREADS / lanes;
  const int cpl = te % lanes, slice = te / lanes;
  for (int cp0 = 0; cp0 < C2; cp0 += lanes) {
    const int cp = cp0 + cpl;
    double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
    if (slice < S && cp < C2) {
      const float4* q = tab + cp;
      int z = slice;
      for (; z + 7 * S < rows; z += 8 * S) {
        float4 v[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) v[u] = __ldcg(q + (size_t)(z + u * S) * C2);
#pragma unroll
        for (int u = 0; u < 8; ++u) { a0 += (double)v[u].x; a1 += (double)v[u].y; a2 += (double)v[u].z; a3 += (double)v[u].w; }
      }
      for (; z < rows; z += S) {
        const float4 v = __ldcg(q + (size_t)z * C2);
        a0 += (double)v.x; a1 += (double)v.y; a2 += (double)v.z; a3 += (double)v.w;
      }
    }
    if (S > 1) {                    // C <= 256: a single pass of the outer loop
      sh[4 * te] = a0; sh[4 * te + 1] = a1; sh[4 * te + 2] = a2; sh[4 * te + 3] = a3;
      asm volatile("bar.sync 1, 256;" ::: "memory");
      if (slice == 0)
        for (int j = 1; j < S; ++j) {
          const double* o = sh + 4 * (j * lanes + cpl);
          a0 += o[0]; a1 += o[1]; a2 += o[2]; a3 += o[3];
        }
    }
    if (slice == 0 && cp < C2) {
      const int c = 2 * cp;
      if (p.fin_mode == 1) {
        bn_finalize_channel(a0, a1, p.fin_M, p.fin_eps, p.fin_decay, p.fin_unbiased, c, p.fin_o0, p.fin_o1, p.fin_mm, p.fin_mv);
        bn_finalize_channel(a2, a3, p.fin_M, p.fin_eps, p.fin_decay, p.fin_unbiased, c + 1, p.fin_o0, p.fin_o1, p.fin_mm, p.fin_mv);
      } else {                      // (sum g, sum g * xhat) -> dbeta, dgamma
        p.fin_o1[c] = (float)a0; p.fin_o0[c] = (float)a1;
        p.fin_o1[c + 1] = (float)a2; p.fin_o0[c + 1] = (float)a3;
      }
    }
  }
}

constexpr int TC_RES_SLOTS_MAX = 6;        // upper bound of TcConvParams::res_slots (tiles of 16 KB in flight)
constexpr int TC_THREADS_RES = TC_CONV_THREADS + 32;

// FUSED = false compiles the plain epilogue only (fprop / dgrad without fusion keep their registers and schedule).
// PAIR = true: the grid is made of 2-CTA clusters; the pair works on two neighbouring groups of pixel tiles with ONE
// stream of tcgen05.mma.cta_group::2 instructions issued by the leader CTA (see rn_tc_common.cuh).
// HALO = true: 3x3 halo mode (see TcConvParams::halo); compiled separately so that the per-tap issue loops of the
// other layers stay as tight as they were.
// Operand transform of one 16-byte chunk (8 bf16 channels of one pixel row): bf16(act(a*x + b)) with the arithmetic of
// bn_apply_kernel (fp32 fma per element, one rounding to bf16).  ab = (a0, b0, a1, b1, ...).  For a plain ReLU the
// max is taken on the packed bf16 pair after rounding — rounding is monotonic and keeps zero, so
// max(round(v), 0) == round(max(v, 0)) bit for bit — which halves the activation instructions.
__device__ __forceinline__ uint4 bn_transform_chunk(const uint4 v, const float (&ab)[16], float leak) {
  const uint32_t w[4] = {v.x, v.y, v.z, v.w};
  uint32_t o[4];
#pragma unroll
  for (int e = 0; e < 4; ++e) {
    const float2 x = make_float2(__uint_as_float(w[e] << 16), __uint_as_float(w[e] & 0xffff0000u));
    const float2 y = __ffma2_rn(x, make_float2(ab[4 * e], ab[4 * e + 2]), make_float2(ab[4 * e + 1], ab[4 * e + 3]));
    if (leak == 0.f) {
      const __nv_bfloat162 h = __hmax2(__floats2bfloat162_rn(y.x, y.y), __floats2bfloat162_rn(0.f, 0.f));
      o[e] = *reinterpret_cast<const uint32_t*>(&h);
    } else {
      const __nv_bfloat162 h = __floats2bfloat162_rn(act_fn(y.x, leak), act_fn(y.y, leak));
      o[e] = *reinterpret_cast<const uint32_t*>(&h);
    }
  }
  return make_uint4(o[0], o[1], o[2], o[3]);
}

// XF = true: operand transform (see TcConvParams::xf): TC_XF_WARPS more warps, one more barrier per stage.
constexpr int TC_XF_WARPS = 4;            // measured: 8 warps are slower (9.93 -> 10.23 ms per ResNet-50 step): the transform is bound by
                                           // shared-memory bandwidth (it adds a read and a write of every operand byte), not by issue slots
constexpr int TC_XF_ROWS = TC_XF_WARPS * 4;  // rows covered by one pass of the transform threads (8 threads per 128-byte row)
template <bool FUSED, bool PAIR, bool HALO, bool XF>
__global__ void __launch_bounds__((FUSED ? TC_THREADS_RES : TC_CONV_THREADS) + (XF ? TC_XF_WARPS * 32 : 0), 1)
conv_tc_kernel(const __grid_constant__ TcConvParams p) {
  static_assert(!(XF && (PAIR || HALO)), "the operand transform is built for single-CTA 1x1 layers");
  const uint32_t cta_rank = PAIR ? cluster_ctarank() : 0u;
  const bool leader = cta_rank == 0u;
  const bool fuse_add = FUSED && p.fuse_add;
  const bool fuse_bnbwd = FUSED && p.fuse_bnbwd;
  const bool fuse_stats = FUSED && (p.fuse_stats || p.fuse_bnbwd);      // both keep per-channel running sums
  const bool use_ring = fuse_add || fuse_bnbwd;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base_u32 = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem = smem_raw + (base_u32 - smem_u32(smem_raw));
  constexpr int b_per_stage = HALO ? 3 : 1;                                    // weight tiles per pipeline stage
  const uint32_t stage_bytes = (uint32_t)p.acc_m * p.a_bytes + (uint32_t)b_per_stage * p.b_bytes;
  const int row_bytes_out = p.gw * 2;
  uint8_t* epi_buf = smem + (size_t)p.stages * stage_bytes;                  // 2 x 16 KB
  uint8_t* res_buf = epi_buf + 2 * 16384;                                     // res_slots x 16 KB (fused add / fused BN backward)
  uint8_t* after_res = res_buf + (use_ring ? p.res_slots * 16384 : 0);
  float* cs = reinterpret_cast<float*>(after_res);                            // 2 x [8 warps][32 column pairs][4] (fused stats)
  float* pacc = cs + 2048;                                                    // [ncols_total][2]
  float* ctab = pacc + 2 * p.ncols_total;                                     // [ncols_total][4] = a, b, mean, rstd (fused BN backward)
  uint8_t* after_stats = after_res + (fuse_stats ? 8192 + (size_t)p.ncols_total * 8 : 0) +
                         (fuse_bnbwd ? (size_t)p.ncols_total * 16 : 0);
  float* xtab = reinterpret_cast<float*>(after_stats);                        // [xf_channels][2] = a, b (operand transform)
  if (XF) after_stats += (size_t)p.xf_channels * 8;
  uint64_t* bars = reinterpret_cast<uint64_t*>(after_stats);
  uint64_t* full = bars;                  // [stages]
  uint64_t* empty = bars + p.stages;      // [stages]
  uint64_t* tfull = bars + 2 * p.stages;  // [2]
  uint64_t* tempty = tfull + 2;           // [2]
  uint64_t* rfull = tempty + 2;           // [TC_RES_SLOTS_MAX]
  uint64_t* rempty = rfull + TC_RES_SLOTS_MAX;
  uint64_t* xfull = rempty + TC_RES_SLOTS_MAX;      // [stages] (operand transform: A tile rewritten, MMA may read)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(xfull + (XF ? p.stages : 0));

  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  if (warp == 0 && lane == 0) {
    for (int i = 0; i < p.stages; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
    if (XF) for (int i = 0; i < p.stages; ++i) mbar_init(&xfull[i], TC_XF_WARPS);
    // pair: the leader's tempty collects the epilogue warps of BOTH CTAs (the peer's arrive remotely)
    for (int i = 0; i < 2; ++i) { mbar_init(&tfull[i], 1); mbar_init(&tempty[i], PAIR ? 2 * TC_EPI_WARPS : TC_EPI_WARPS); }
    for (int i = 0; i < TC_RES_SLOTS_MAX; ++i) { mbar_init(&rfull[i], 1); mbar_init(&rempty[i], TC_EPI_WARPS); }
    fence_barrier_init();
    tma_prefetch_desc(&p.a_map[0]);
    tma_prefetch_desc(&p.b_map);
    tma_prefetch_desc(&p.o_map);
    if (use_ring) tma_prefetch_desc(&p.r_map);
  }
  if (warp == 1) {
    if (PAIR) { tmem_alloc_pair(tmem_slot, p.tmem_cols); tmem_relinquish_pair(); }
    else      { tmem_alloc(tmem_slot, p.tmem_cols); tmem_relinquish(); }
  }
  if (fuse_stats && warp >= 2 && warp < 2 + TC_EPI_WARPS)
    for (int i = threadIdx.x - 64; i < 2 * p.ncols_total; i += TC_EPI_THREADS) pacc[i] = 0.f;
  tc_fence_before();
  if (PAIR) cluster_sync_all();       // the peer's barriers and tensor memory exist before anything targets them
  else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  // everything above (barriers, TMEM, descriptor prefetch) overlapped the previous kernel's tail
  pdl_trigger();
  pdl_wait();
  if (fuse_bnbwd && warp >= 2 && warp < 2 + TC_EPI_WARPS) {
    // per-channel constants of the BN whose backward sums this dgrad accumulates (read by the epilogue warps only,
    // ordered by their first named barrier)
    for (int c = threadIdx.x - 64; c < p.ncols_total; c += TC_EPI_THREADS) {
      const float rs = p.bn_rstd[c], mu = p.bn_mean[c];
      const float a = p.bn_gamma[c] * rs;
      *reinterpret_cast<float4*>(ctab + 4 * c) = make_float4(a, p.bn_beta[c] - mu * a, mu, rs);
    }
  }

  // work item = (group of acc_m consecutive pixel tiles, column tile); the acc_m tiles share each B tile
  // pair: a work item covers 2 * acc_m pixel tiles (acc_m per CTA); tiles past the end have out-of-range
  // coordinates (TMA zero-fills their loads and clips their stores), so both CTAs always run the same pipeline
  const int m_tiles = p.tiles_w * p.tiles_h * p.tiles_n;
  const int m_groups = (m_tiles + p.acc_m - 1) / p.acc_m;
  const int total = PAIR ? ((m_groups + 1) / 2) * p.n_tiles : m_groups * p.n_tiles;
  const int wstart = PAIR ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;
  const int wstep = PAIR ? (int)(gridDim.x >> 1) : (int)gridDim.x;
  const int stage_groups = HALO ? 3 : p.taps;                             // A boxes per k-chunk: horizontal taps / taps
  const int k_iters = stage_groups * p.k_chunks;
  const uint32_t a_tx = (uint32_t)(p.BW * (HALO ? p.BH + 2 : p.BH) * p.BN) * (uint32_t)(p.kc * 2);
  const uint32_t b_tx = (uint32_t)b_per_stage * (uint32_t)p.block_n * (uint32_t)(p.kc * 2);   // whole B tiles of a stage (both halves of a pair)

  if (warp == 0) {
    // ===================== TMA producer (whole warp loops, one elected lane issues) =====================
    int stage = 0;
    uint32_t phase = 0;
    for (int t = wstart; t < total; t += wstep) {
      const int nt = t % p.n_tiles;
      const int mg = PAIR ? (t / p.n_tiles) * 2 + (int)cta_rank : t / p.n_tiles;
      int nvalid = m_tiles - mg * p.acc_m;
      if (nvalid > p.acc_m || PAIR) nvalid = p.acc_m;
      int w0[4], h0[4], n0[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        int mt = mg * p.acc_m + j;
        w0[j] = (mt % p.tiles_w) * p.BW;
        mt /= p.tiles_w;
        h0[j] = (mt % p.tiles_h) * p.BH;
        n0[j] = (mt / p.tiles_h) * p.BN;
      }
      // pair: only the leader arms its barrier, with the bytes of BOTH CTAs (the peer's loads complete on it too)
      const uint32_t tx = PAIR ? 2u * (uint32_t)p.acc_m * a_tx + b_tx : (uint32_t)nvalid * a_tx + b_tx;
      for (int tap = 0; tap < stage_groups; ++tap) {
        const CUtensorMap* am = &p.a_map[HALO ? 0 : p.tap_map[tap]];
        const int dw = HALO ? p.halo_dw[tap] : p.tap_dw[tap], dh = HALO ? p.halo_dh0 : p.tap_dh[tap];
        for (int kk = 0; kk < p.k_chunks; ++kk) {
          mbar_wait(&empty[stage], phase ^ 1);
          if (elect_one()) {
            uint8_t* sa = smem + (size_t)stage * stage_bytes;
            if (!PAIR || leader) mbar_expect_tx(&full[stage], tx);
#pragma unroll
            for (int j = 0; j < 4; ++j)
              if (j < nvalid) {
                if (PAIR) tma_load_4d_pair(sa + (size_t)j * p.a_bytes, am, &full[stage], kk * p.kc, w0[j] + dw, h0[j] + dh, n0[j]);
                else      tma_load_4d(sa + (size_t)j * p.a_bytes, am, &full[stage], kk * p.kc, w0[j] + dw, h0[j] + dh, n0[j]);
              }
            for (int q = 0; q < b_per_stage; ++q) {
              uint8_t* sb = sa + (size_t)p.acc_m * p.a_bytes + (size_t)q * p.b_bytes;
              const int tb = HALO ? p.halo_b[tap][q] : p.tap_b[tap];
              if (p.b_mn) {
                // one op: (gw couts, kc cin rows, b_boxes cout blocks) lands as b_boxes consecutive MN-major boxes
                if (PAIR) tma_load_4d_pair(sb, &p.b_map, &full[stage], 0, kk * p.kc, (2 * nt + (int)cta_rank) * p.b_boxes, tb);
                else      tma_load_4d(sb, &p.b_map, &full[stage], 0, kk * p.kc, nt * p.b_boxes, tb);
              } else {
                if (PAIR) tma_load_3d_pair(sb, &p.b_map, &full[stage], kk * p.kc, nt * p.block_n + (int)cta_rank * (p.block_n / 2), tb);
                else      tma_load_3d(sb, &p.b_map, &full[stage], kk * p.kc, nt * p.block_n, tb);
              }
            }
          }
          __syncwarp();
          if (++stage == p.stages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (whole warp loops, one elected lane issues; pair: the leader CTA only) ==========
    if (!PAIR || leader) {
    const uint32_t idesc = make_idesc_bf16(PAIR ? 2 * TC_BLOCK_M : TC_BLOCK_M, p.block_n, 0, p.b_mn);
    const uint32_t rowB = (uint32_t)p.gw * 2, ltB = swizzle_layout_type(p.gw * 2);
    const int row_bytes = p.kc * 2;
    const uint32_t lt = swizzle_layout_type(row_bytes);
    const uint32_t sbo = 8u * row_bytes;
    const int ksteps = p.kc / 16;
    // Descriptors differ only in their 14-bit start-address field (bytes >> 4); every address lies inside this
    // CTA's shared-memory window (< 256 KB), so stepping is a plain add on the low word.  (Pair: the same offsets
    // address the peer's shared memory — both CTAs use one layout.)
    const uint64_t da0 = make_smem_desc(base_u32, 16, sbo, lt);
    // K-major B: +32 bytes per K step inside the swizzled row.  MN-major B (rows = input channels):
    // +16 rows per K step, LBO = next 64-column box, SBO = 8 rows.
    const uint64_t db0 = p.b_mn ? make_smem_desc(base_u32 + (uint32_t)p.acc_m * p.a_bytes, p.b_box_bytes, 8 * rowB, ltB)
                                : make_smem_desc(base_u32 + (uint32_t)p.acc_m * p.a_bytes, 16, sbo, lt);
    const uint32_t a_step = 2, b_step = p.b_mn ? (16u * rowB) >> 4 : 2u;
    const uint32_t stage16 = stage_bytes >> 4, a16 = p.a_bytes >> 4;
    int stage = 0;
    uint32_t phase = 0;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int t = wstart; t < total; t += wstep) {
      const int mg = t / p.n_tiles;
      int nvalid = m_tiles - mg * p.acc_m;
      if (nvalid > p.acc_m || PAIR) nvalid = p.acc_m;
      mbar_wait(&tempty[acc], acc_phase ^ 1);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + (uint32_t)(acc * p.acc_m * p.block_n);
      for (int it = 0; it < k_iters; ++it) {
        mbar_wait(XF ? &xfull[stage] : &full[stage], phase);
        tc_fence_after();
        if (elect_one()) {
          const uint64_t da_s = da0 + (uint64_t)((uint32_t)stage * stage16);
          const uint64_t db_s = db0 + (uint64_t)((uint32_t)stage * stage16);
          const int grp = HALO ? it / p.k_chunks : 0;
          for (int q = 0; q < b_per_stage; ++q) {
            // halo mode: vertical tap q reads the A box from row halo_row * BW on: only the start-address field moves
            // (16-byte units).  Measured on B200: the 128-byte swizzle is a function of the absolute shared-memory
            // address bits, so a start that is not a whole 1024-byte atom needs NO base-offset (bits 49-51 stay 0;
            // setting them to (row & 7) produced wrong results for every BW that is not a multiple of 8).
            const uint32_t rows = HALO ? (uint32_t)p.halo_row[grp][q] * (uint32_t)p.BW : 0u;
            const uint64_t a_off = (uint64_t)(rows * (uint32_t)row_bytes >> 4);
            const uint64_t db_q = db_s + (uint64_t)((uint32_t)q * (p.b_bytes >> 4));
            for (int j = 0; j < nvalid; ++j) {
              const uint64_t da_j = da_s + (uint64_t)((uint32_t)j * a16) + a_off;
              const uint32_t d = d_tmem + (uint32_t)(j * p.block_n);
#pragma unroll 4
              for (int k = 0; k < ksteps; ++k) {
                const uint32_t accf = (it > 0 || q > 0 || k > 0) ? 1u : 0u;
                if (PAIR) umma_bf16_pair(d, da_j + (uint64_t)(k * a_step), db_q + (uint64_t)(k * b_step), idesc, accf);
                else      umma_bf16(d, da_j + (uint64_t)(k * a_step), db_q + (uint64_t)(k * b_step), idesc, accf);
              }
            }
          }
          // frees the smem slot (in both CTAs of a pair) once these MMAs have read it
          if (PAIR) umma_commit_pair(&empty[stage]); else umma_commit(&empty[stage]);
          if (it == k_iters - 1) {                                       // accumulators complete
            if (PAIR) umma_commit_pair(&tfull[acc]); else umma_commit(&tfull[acc]);
          }
        }
        __syncwarp();
        if (++stage == p.stages) { stage = 0; phase ^= 1; }
      }
      if (++acc == p.tstages) { acc = 0; acc_phase ^= 1; }
    }
    }
  } else if (warp < 2 + TC_EPI_WARPS) {
    // ===================== epilogue (warps 2..9) =====================
    // Two warps per TMEM lane quadrant: a thread owns one pixel row and HALF of each 64-channel store group.
    const int quad = warp & 3;                      // TMEM lane quadrant this warp may access
    const int half = (warp - 2) >> 2;               // which 32 channels of a 64-channel group
    const int ew = warp - 2;                        // 0..7
    const int row = quad * 32 + lane;
    const int te = threadIdx.x - 64;                // 0..255
    const int box_rows = p.BW * p.BH * p.BN;
    int acc = 0;
    uint32_t acc_phase = 0;
    int buf = 0;
    int rslot = 0;
    uint32_t rphase = 0;
    int csb = 0, pend_col = -1;                     // fused statistics: cs buffer in use, column base awaiting its combine
    const int groups = p.block_n / p.gw;
    // Fused statistics, second half: add the eight partial sums of the PREVIOUS store group (cs[csb ^ 1], written
    // before this group's first barrier) into the CTA's running per-channel totals.  Thread te < 2 * gw owns
    // (column te % gw, statistic te / gw); fixed order.
    const int gw_lg = p.gw == 64 ? 6 : (p.gw == 32 ? 5 : 4);            // store groups are 16, 32 or 64 channels wide
    auto combine = [&](int col0, int which) {
      if (te >= 2 * p.gw) return;
      const float* c = cs + which * 1024;
      const int col = te & (p.gw - 1), st = te >> gw_lg;
      float v = 0.f;
#pragma unroll
      for (int q = 0; q < TC_EPI_WARPS; ++q) v += c[(q * 32 + (col >> 1)) * 4 + st * 2 + (col & 1)];
      pacc[(col0 + col) * 2 + st] += v;
    };
    const uint32_t tempty_leader[2] = {PAIR ? mapa_u32(&tempty[0], 0) : 0u, PAIR ? mapa_u32(&tempty[1], 0) : 0u};
    for (int t = wstart; t < total; t += wstep) {
      const int nt = t % p.n_tiles;
      const int mg = PAIR ? (t / p.n_tiles) * 2 + (int)cta_rank : t / p.n_tiles;
      int nvalid = m_tiles - mg * p.acc_m;
      if (nvalid > p.acc_m || PAIR) nvalid = p.acc_m;
      mbar_wait(&tfull[acc], acc_phase);
      tc_fence_after();
      for (int j = 0; j < nvalid; ++j) {
        int mt = mg * p.acc_m + j;
        const int w0 = (mt % p.tiles_w) * p.BW;
        mt /= p.tiles_w;
        const int h0 = (mt % p.tiles_h) * p.BH;
        const int n0 = (mt / p.tiles_h) * p.BN;
        // rows of this quadrant whose output pixel exists: a k > 1 convolution produces non-zero values for
        // pixels just outside the image (TMA clips them on store; the statistics must skip them too)
        uint32_t row_mask = 0;
        if (fuse_stats) {
          const int bw = row % p.BW, bh = (row / p.BW) % p.BH, bn = row / (p.BW * p.BH);
          const bool ok = row < box_rows && w0 + bw < p.oW && h0 + bh < p.oH && n0 + bn < p.oN;
          row_mask = __ballot_sync(0xffffffffu, ok);
        }
        for (int g = 0; g < groups; ++g) {
          const uint32_t taddr = tmem_base + ((uint32_t)(quad * 32) << 16) +
                                 (uint32_t)((acc * p.acc_m + j) * p.block_n + g * p.gw);
          // 64-column groups: the TMEM load is issued before waiting for the staging buffer, so its latency
          // overlaps the barrier; then one wait, 4 x (4 cvt + one 128-bit shared store)
          uint32_t va[32];
          if (p.gw == 64) tmem_ld_32x32(taddr + half * 32, va);
          // the TMA store that last read this staging buffer must have finished reading it
          if (warp == 2) {
            if (elect_one()) tma_store_wait_read<1>();
            __syncwarp();
          }
          asm volatile("bar.sync 1, 256;" ::: "memory");
          uint8_t* sb = epi_buf + buf * 16384;
          if (p.gw == 64) {
            if (fuse_stats && pend_col >= 0) combine(pend_col, csb ^ 1);
            tmem_ld_wait();
            const uint32_t rbase = (uint32_t)row * 128u + (uint32_t)half * 64u;
            const uint8_t* rb = res_buf + rslot * 16384;
            if (fuse_add) mbar_wait(&rfull[rslot], rphase);
#pragma unroll
            for (int c = 0; c < 4; ++c) {
              uint32_t pk[4];
#pragma unroll
              for (int i = 0; i < 4; ++i) {
                __nv_bfloat162 h = __floats2bfloat162_rn(__uint_as_float(va[8 * c + 2 * i]), __uint_as_float(va[8 * c + 2 * i + 1]));
                pk[i] = *reinterpret_cast<uint32_t*>(&h);
              }
              const uint32_t so = swizzle_offset(rbase + 16u * c, 128);
              if (fuse_add) {
                // z = bf16(bf16(conv) + residual): the same two roundings as a stored conv output followed by the
                // add.  One packed bf16x2 add per channel pair: the sum of two bf16 values rounded once to bf16
                // equals their fp32 sum rounded to bf16 (the fp32 add is exact unless one addend is below
                // 2^-16 of the other, where both round to the larger one).
                const uint4 rv = *reinterpret_cast<const uint4*>(rb + so);
                const uint32_t rw[4] = {rv.x, rv.y, rv.z, rv.w};
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                  const __nv_bfloat162 h = __hadd2(*reinterpret_cast<const __nv_bfloat162*>(&pk[i]),
                                                   *reinterpret_cast<const __nv_bfloat162*>(&rw[i]));
                  pk[i] = *reinterpret_cast<const uint32_t*>(&h);
                }
              }
              *reinterpret_cast<uint4*>(sb + so) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
            }
            if (fuse_add) {
              __syncwarp();
              if (lane == 0) mbar_arrive(&rempty[rslot]);
              if (++rslot == p.res_slots) { rslot = 0; rphase ^= 1; }
            }
          } else {
            // narrow store groups (16 / 32 channels): one warp per quadrant converts whole rows, 16 columns at a time
            if (fuse_stats && pend_col >= 0) combine(pend_col, csb ^ 1);
            const uint8_t* rb = res_buf + rslot * 16384;
            if (half == 0) {
              if (fuse_add) mbar_wait(&rfull[rslot], rphase);
              for (int c0 = 0; c0 < p.gw; c0 += 16) {
                uint32_t v[16];
                tmem_ld_32x16(taddr + c0, v);
                tmem_ld_wait();
                uint32_t pk[8];
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                  __nv_bfloat162 h = __floats2bfloat162_rn(__uint_as_float(v[2 * i]), __uint_as_float(v[2 * i + 1]));
                  pk[i] = *reinterpret_cast<uint32_t*>(&h);
                }
                // 16 channels = two 16-byte chunks of this row, XOR-swizzled like the TMA store expects
                const uint32_t off0 = (uint32_t)row * row_bytes_out + (uint32_t)c0 * 2;
                const uint32_t so0 = swizzle_offset(off0, row_bytes_out), so1 = swizzle_offset(off0 + 16, row_bytes_out);
                if (fuse_add) {                     // z = bf16(bf16(conv) + residual), as in the 64-channel path
                  const uint4 r0 = *reinterpret_cast<const uint4*>(rb + so0), r1 = *reinterpret_cast<const uint4*>(rb + so1);
                  const uint32_t rw[8] = {r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w};
#pragma unroll
                  for (int i = 0; i < 8; ++i) {
                    const __nv_bfloat162 h = __hadd2(*reinterpret_cast<const __nv_bfloat162*>(&pk[i]),
                                                     *reinterpret_cast<const __nv_bfloat162*>(&rw[i]));
                    pk[i] = *reinterpret_cast<const uint32_t*>(&h);
                  }
                }
                *reinterpret_cast<uint4*>(sb + so0) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
                *reinterpret_cast<uint4*>(sb + so1) = make_uint4(pk[4], pk[5], pk[6], pk[7]);
              }
            }
            if (fuse_add) {                         // every epilogue warp releases the ring slot (the barrier counts all eight)
              __syncwarp();
              if (lane == 0) mbar_arrive(&rempty[rslot]);
              if (++rslot == p.res_slots) { rslot = 0; rphase ^= 1; }
            }
          }
          fence_proxy_async();
          asm volatile("bar.sync 1, 256;" ::: "memory");
          if (warp == 2) {
            // the same (deterministically elected) lane issues, commits and later waits on the bulk stores
            if (elect_one()) {
              if (p.accumulate) tma_reduce_add_4d(&p.o_map, sb, nt * p.block_n + g * p.gw, w0, h0, n0);
              else              tma_store_4d(&p.o_map, sb, nt * p.block_n + g * p.gw, w0, h0, n0);
              tma_store_commit();
            }
            __syncwarp();
          }
          if (fuse_bnbwd && p.gw == 64) {
            // BN backward reductions of the staged dy group (barrier above): g = dy where the forward pre-activation
            // a*x + b was >= 0, leak*dy elsewhere; sums of g and g*xhat over this warp's 16 rows, one channel pair
            // per lane.  x comes from the tile ring (same swizzled layout as the staging buffer).
            mbar_wait(&rfull[rslot], rphase);
            const uint8_t* rb = res_buf + rslot * 16384;
            const int cpair = nt * p.block_n + g * 64 + 2 * lane;
            const float4 k0 = *reinterpret_cast<const float4*>(ctab + 4 * cpair);
            const float4 k1 = *reinterpret_cast<const float4*>(ctab + 4 * cpair + 4);
            const uint32_t hmask = (row_mask >> (half * 16)) & 0xffffu;
            float s0 = 0.f, s1 = 0.f, q0 = 0.f, q1 = 0.f;
#pragma unroll
            for (int r = 0; r < 16; ++r) {
              const uint32_t so = swizzle_offset((uint32_t)(quad * 32 + half * 16 + r) * 128u + (uint32_t)lane * 4u, 128);
              uint32_t wg = *reinterpret_cast<const uint32_t*>(sb + so);
              uint32_t wx = *reinterpret_cast<const uint32_t*>(rb + so);
              wg = ((hmask >> r) & 1u) ? wg : 0u;                  // rows outside the tensor contribute +0
              wx = ((hmask >> r) & 1u) ? wx : 0u;                  // (rows past the pixel box are stale smem: 0 * NaN)
              const float g0 = __uint_as_float(wg << 16), g1 = __uint_as_float(wg & 0xffff0000u);
              const float x0 = __uint_as_float(wx << 16), x1 = __uint_as_float(wx & 0xffff0000u);
              const float m0 = fmaf(x0, k0.x, k0.y) < 0.f ? p.bn_leak * g0 : g0;
              const float m1 = fmaf(x1, k1.x, k1.y) < 0.f ? p.bn_leak * g1 : g1;
              s0 += m0; s1 += m1;
              q0 = fmaf(m0, (x0 - k0.z) * k0.w, q0);
              q1 = fmaf(m1, (x1 - k1.z) * k1.w, q1);
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&rempty[rslot]);
            if (++rslot == p.res_slots) { rslot = 0; rphase ^= 1; }
            *reinterpret_cast<float4*>(cs + csb * 1024 + (ew * 32 + lane) * 4) = make_float4(s0, s1, q0, q1);
            pend_col = nt * p.block_n + g * 64;
            csb ^= 1;
          } else if (fuse_stats && p.gw == 64) {
            // first half of the fused statistics: the staged group is complete (barrier above); this warp sums 16 of
            // its quadrant's rows, one column pair per lane (a row is the 32 lanes' 128 bytes: conflict-free under
            // the swizzle), value and value^2 of the STORED bf16.  All loads independent; rows outside the tensor
            // contribute +0 (they may hold anything).
            float2 sv = make_float2(0.f, 0.f), qv = make_float2(0.f, 0.f);     // packed fp32x2 accumulators
            const uint32_t hmask = (row_mask >> (half * 16)) & 0xffffu;
            if (hmask == 0xffffu) {                                            // warp-uniform: all 16 rows exist
#pragma unroll
              for (int r = 0; r < 16; ++r) {
                const uint32_t w = *reinterpret_cast<const uint32_t*>(
                    sb + swizzle_offset((uint32_t)(quad * 32 + half * 16 + r) * 128u + (uint32_t)lane * 4u, 128));
                const float2 f = make_float2(__uint_as_float(w << 16), __uint_as_float(w & 0xffff0000u));
                sv = __fadd2_rn(sv, f);
                qv = __ffma2_rn(f, f, qv);
              }
            } else {
#pragma unroll
              for (int r = 0; r < 16; ++r) {
                uint32_t w = *reinterpret_cast<const uint32_t*>(
                    sb + swizzle_offset((uint32_t)(quad * 32 + half * 16 + r) * 128u + (uint32_t)lane * 4u, 128));
                w = ((hmask >> r) & 1u) ? w : 0u;
                const float2 f = make_float2(__uint_as_float(w << 16), __uint_as_float(w & 0xffff0000u));
                sv = __fadd2_rn(sv, f);
                qv = __ffma2_rn(f, f, qv);
              }
            }
            const float s0 = sv.x, s1 = sv.y, q0 = qv.x, q1 = qv.y;
            *reinterpret_cast<float4*>(cs + csb * 1024 + (ew * 32 + lane) * 4) = make_float4(s0, s1, q0, q1);
            pend_col = nt * p.block_n + g * 64;
            csb ^= 1;
          } else if (fuse_stats) {
            // Narrow store groups (16 / 32 channels: the CIFAR widths).  A row of the staged group is 32 / 64 bytes, so a
            // warp's 32 lanes cover gw/2 channel pairs x (64/gw) rows at a time: lane = (row lane, channel pair); each
            // warp sums 16 rows of its quadrant, the row lanes are folded with shuffles, lanes < gw/2 publish the sums.
            const int npairs = p.gw >> 1, rl = 32 / npairs;
            const int cp = lane % npairs, ro = lane / npairs;
            const uint32_t hmask = (row_mask >> (half * 16)) & 0xffffu;
            const uint32_t rowB = (uint32_t)row_bytes_out;
            float s0 = 0.f, s1 = 0.f, q0 = 0.f, q1 = 0.f;
            if (fuse_bnbwd) {
              mbar_wait(&rfull[rslot], rphase);
              const uint8_t* rb = res_buf + rslot * 16384;
              const int cpair = nt * p.block_n + g * p.gw + 2 * cp;
              const float4 k0 = *reinterpret_cast<const float4*>(ctab + 4 * cpair);
              const float4 k1 = *reinterpret_cast<const float4*>(ctab + 4 * cpair + 4);
              for (int rr = ro; rr < 16; rr += rl) {
                const uint32_t so = swizzle_offset((uint32_t)(quad * 32 + half * 16 + rr) * rowB + (uint32_t)cp * 4u, row_bytes_out);
                uint32_t wg = *reinterpret_cast<const uint32_t*>(sb + so);
                uint32_t wx = *reinterpret_cast<const uint32_t*>(rb + so);
                wg = ((hmask >> rr) & 1u) ? wg : 0u;
                wx = ((hmask >> rr) & 1u) ? wx : 0u;
                const float g0 = __uint_as_float(wg << 16), g1 = __uint_as_float(wg & 0xffff0000u);
                const float x0 = __uint_as_float(wx << 16), x1 = __uint_as_float(wx & 0xffff0000u);
                const float m0 = fmaf(x0, k0.x, k0.y) < 0.f ? p.bn_leak * g0 : g0;
                const float m1 = fmaf(x1, k1.x, k1.y) < 0.f ? p.bn_leak * g1 : g1;
                s0 += m0; s1 += m1;
                q0 = fmaf(m0, (x0 - k0.z) * k0.w, q0);
                q1 = fmaf(m1, (x1 - k1.z) * k1.w, q1);
              }
              __syncwarp();
              if (lane == 0) mbar_arrive(&rempty[rslot]);
              if (++rslot == p.res_slots) { rslot = 0; rphase ^= 1; }
            } else {
              for (int rr = ro; rr < 16; rr += rl) {
                uint32_t w = *reinterpret_cast<const uint32_t*>(
                    sb + swizzle_offset((uint32_t)(quad * 32 + half * 16 + rr) * rowB + (uint32_t)cp * 4u, row_bytes_out));
                w = ((hmask >> rr) & 1u) ? w : 0u;
                const float f0 = __uint_as_float(w << 16), f1 = __uint_as_float(w & 0xffff0000u);
                s0 += f0; s1 += f1;
                q0 = fmaf(f0, f0, q0); q1 = fmaf(f1, f1, q1);
              }
            }
            for (int o = npairs; o < 32; o <<= 1) {               // fold the row lanes (fixed order)
              s0 += __shfl_xor_sync(0xffffffffu, s0, o); s1 += __shfl_xor_sync(0xffffffffu, s1, o);
              q0 += __shfl_xor_sync(0xffffffffu, q0, o); q1 += __shfl_xor_sync(0xffffffffu, q1, o);
            }
            if (lane < npairs)
              *reinterpret_cast<float4*>(cs + csb * 1024 + (ew * 32 + cp) * 4) = make_float4(s0, s1, q0, q1);
            pend_col = nt * p.block_n + g * p.gw;
            csb ^= 1;
          }
          buf ^= 1;
        }
      }
      // all TMEM reads of this accumulator set are complete (tcgen05.wait::ld above)
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if (PAIR) mbar_arrive_cluster(tempty_leader[acc]);     // the leader's MMA warp waits for both CTAs' epilogues
        else mbar_arrive(&tempty[acc]);
      }
      if (++acc == p.tstages) { acc = 0; acc_phase ^= 1; }
    }
    if (fuse_stats) {
      // last group's partial sums, then this CTA's totals go to its row of the partial table (the finalize
      // kernel sums the <= 148 rows in fp64 in a fixed order)
      asm volatile("bar.sync 1, 256;" ::: "memory");
      if (pend_col >= 0) combine(pend_col, csb ^ 1);
      asm volatile("bar.sync 1, 256;" ::: "memory");
      float* dst = p.stat_partials + (size_t)blockIdx.x * 2 * p.ncols_total;
      for (int i = te; i < 2 * p.ncols_total; i += TC_EPI_THREADS) dst[i] = pacc[i];
      if (p.fin_mode) {
        __threadfence();                                   // this CTA's row is visible device-wide before its ticket
        asm volatile("bar.sync 1, 256;" ::: "memory");
        int* flag = reinterpret_cast<int*>(cs);
        if (te == 0) {
          const int last = atomicAdd(p.fin_ticket, 1) == (int)gridDim.x - 1;
          if (last) atomicExch(p.fin_ticket, 0);           // every CTA has drawn its ticket: ready for the next launch
          *flag = last;
        }
        asm volatile("bar.sync 1, 256;" ::: "memory");
        const int last = *flag;
        asm volatile("bar.sync 1, 256;" ::: "memory");     // the flag has been read before `cs` is reused
        if (last) {
          __threadfence();
          fold_finalize(p, te, (int)gridDim.x, reinterpret_cast<double*>(cs));
        }
      }
    }
    if (warp == 2) {
      if (elect_one()) tma_store_wait<0>();
      __syncwarp();
    }
  } else if (XF && warp >= (int)(blockDim.x / 32) - TC_XF_WARPS) {
    // ===================== operand transform (the last four warps) =====================
    const int tt = (warp - ((int)(blockDim.x / 32) - TC_XF_WARPS)) * 32 + lane;       // 0..127
    for (int c = tt; c < p.xf_channels; c += TC_XF_WARPS * 32) {
      float a, b;
      bn_affine(p.xf_gamma[c], p.xf_beta[c], p.xf_mean[c], p.xf_rstd[c], &a, &b);
      xtab[2 * c] = a; xtab[2 * c + 1] = b;
    }
    asm volatile("bar.sync 2, %0;" ::"n"(TC_XF_WARPS * 32) : "memory");
    // a thread owns one 16-byte chunk column (8 channels: constants in registers) of rows r0, r0 + TC_XF_ROWS, ...; the
    // 128-byte swizzle keeps logical chunk lc of row r at physical chunk lc ^ (r & 7)
    const int lc = tt & 7, r0 = tt >> 3;
    const float leak = p.xf_leak;
    int stage = 0;
    uint32_t phase = 0;
    for (int t = wstart; t < total; t += wstep) {
      const int mg = t / p.n_tiles;
      int nvalid = m_tiles - mg * p.acc_m;
      if (nvalid > p.acc_m) nvalid = p.acc_m;
      for (int it = 0; it < k_iters; ++it) {
        const int kk = it % p.k_chunks;
        float ab[16];
        const float4* src = reinterpret_cast<const float4*>(xtab + 2 * (kk * 64 + lc * 8));
#pragma unroll
        for (int i = 0; i < 4; ++i) { const float4 v = src[i]; ab[4 * i] = v.x; ab[4 * i + 1] = v.y; ab[4 * i + 2] = v.z; ab[4 * i + 3] = v.w; }
        mbar_wait(&full[stage], phase);
        uint8_t* sa = smem + (size_t)stage * stage_bytes;
        for (int j = 0; j < nvalid; ++j) {
          uint8_t* tile = sa + (size_t)j * p.a_bytes;
#pragma unroll 4
          for (int r = r0; r < TC_BLOCK_M; r += TC_XF_ROWS) {
            uint4* q = reinterpret_cast<uint4*>(tile + r * 128 + ((lc ^ (r & 7)) << 4));
            *q = bn_transform_chunk(*q, ab, leak);
          }
        }
        fence_proxy_async();                 // generic-proxy writes -> visible to the tensor core's (async proxy) reads
        __syncwarp();
        if (lane == 0) mbar_arrive(&xfull[stage]);
        if (++stage == p.stages) { stage = 0; phase ^= 1; }
      }
    }
  } else {
    // ===================== residual-tile producer (warp 10; only launched with the fused add) =====================
    const uint32_t rtx = (uint32_t)(p.BW * p.BH * p.BN) * (uint32_t)(p.gw * 2);
    const int groups = p.block_n / p.gw;
    int rslot = 0;
    uint32_t rphase = 0;
    for (int t = wstart; t < total; t += wstep) {
      const int nt = t % p.n_tiles;
      const int mg = PAIR ? (t / p.n_tiles) * 2 + (int)cta_rank : t / p.n_tiles;
      int nvalid = m_tiles - mg * p.acc_m;
      if (nvalid > p.acc_m || PAIR) nvalid = p.acc_m;
      for (int j = 0; j < nvalid; ++j) {
        int mt = mg * p.acc_m + j;
        const int w0 = (mt % p.tiles_w) * p.BW;
        mt /= p.tiles_w;
        const int h0 = (mt % p.tiles_h) * p.BH;
        const int n0 = (mt / p.tiles_h) * p.BN;
        for (int g = 0; g < groups; ++g) {
          mbar_wait(&rempty[rslot], rphase ^ 1);
          if (elect_one()) {
            mbar_expect_tx(&rfull[rslot], rtx);
            tma_load_4d(res_buf + rslot * 16384, &p.r_map, &rfull[rslot], nt * p.block_n + g * p.gw, w0, h0, n0);
          }
          __syncwarp();
          if (++rslot == p.res_slots) { rslot = 0; rphase ^= 1; }
        }
      }
    }
  }
  tc_fence_before();
  if (PAIR) cluster_sync_all();      // neither CTA may leave while the pair's MMAs / remote arrivals can still touch it
  else __syncthreads();
  if (warp == 1) {
    if (PAIR) tmem_dealloc_pair(tmem_base, p.tmem_cols);
    else tmem_dealloc(tmem_base, p.tmem_cols);
  }
}

// ------------------------------------------------------------------------------------------
// wgrad:  dW[(tap, ci), co] = sum_pixels  X[pixel shifted by tap, ci] * dY[pixel, co]
// Both operands have the reduction (pixel) dimension as their STRIDED dimension in NHWC, i.e. they
// are "MN-major" UMMA operands: a TMA box (64 channels x KP pixels) lands as KP rows of 128 bytes,
// which is exactly the canonical 128B-swizzled MN-major layout (8 pixel rows per 1024-byte atom,
// SBO = 1024; consecutive 64-channel boxes LBO apart).  No transpose pass over HBM.
// A CTA accumulates up to four 128 x BLOCK_N fp32 tiles in tensor memory (512 columns) that share
// the dY tile: A rows are (tap, 64-channel block) pairs, so for 64-channel layers two taps ride in
// one M = 128 instruction.  Split-K over pixel tiles; partial tiles go to caller scratch and are
// summed in a fixed order (deterministic).
// ------------------------------------------------------------------------------------------
struct TcWgradParams {
  CUtensorMap x_map[4];
  CUtensorMap dy_map;
  float* partial;                 // [splits][taps*Cin][Cout]
  int Cin, Cout, taps, cin_blocks, RB;
  int runA;                       // consecutive cin blocks fetched by one A op (5-D box); A ops per stage = rb_per_cta / runA
  int bwA, bwB;                   // channels per A / B box (64, 32 or 16 -> 128 / 64 / 32-byte swizzled rows)
  int blocksA;                    // A boxes per 128-row accumulator (128 / bwA)
  uint32_t a_box_bytes, b_box_bytes;
  int acc;                        // accumulators per CTA; row blocks per CTA = blocksA * acc
  int row_groups, block_n, n_tiles;
  int PW, PH, PN, tiles_w, tiles_h, tiles_n, k_tiles;
  int splits, tiles_per_split;
  int kp, stages;
  uint32_t stage_bytes, tmem_cols;
  int8_t tap_map[TC_MAX_TAPS], tap_dw[TC_MAX_TAPS], tap_dh[TC_MAX_TAPS];
  int pair;                       // 1: clusters of two CTAs, M = 256 (row_groups then counts PAIRS of row groups)
  // operand transform (see TcConvParams::xf): X is the input of a training-mode BN+ReLU, normalised in shared memory
  int xf;
  const float *xf_gamma, *xf_beta, *xf_mean, *xf_rstd;
  float xf_leak;
};

// PAIR = true: clusters of two CTAs and tcgen05.mma.cta_group::2 (M = 256): the CTAs of a pair own neighbouring row
// groups of dW (each keeps its 128-row accumulators in its own tensor memory) and every CTA loads HALF of the dY tile,
// so the operand bytes fetched from shared memory per MAC drop by a third (see the fprop / dgrad kernel).
template <bool XF, bool PAIR>
__global__ void __launch_bounds__(TC_THREADS + (XF ? TC_XF_WARPS * 32 : 0), 1)
conv_wgrad_tc_kernel(const __grid_constant__ TcWgradParams p) {
  static_assert(!(XF && PAIR), "the operand transform is built for single CTAs");
  const uint32_t cta_rank = PAIR ? cluster_ctarank() : 0u;
  const bool leader = cta_rank == 0u;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base_u32 = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem = smem_raw + (base_u32 - smem_u32(smem_raw));
  uint8_t* epi_slab = smem + (size_t)p.stages * p.stage_bytes;                 // 4 epilogue warps x 4 KB
  uint64_t* bars = reinterpret_cast<uint64_t*>(epi_slab + 16384);
  uint64_t* full = bars;
  uint64_t* empty = bars + p.stages;
  uint64_t* tfull = bars + 2 * p.stages;
  uint64_t* tempty = tfull + 1;
  uint64_t* xfull = tempty + 1;                                               // [stages] (operand transform)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(xfull + (XF ? p.stages : 0));
  float* xtab = reinterpret_cast<float*>(epi_slab + 16384 + 1024);            // [Cin][2] = a, b (operand transform)
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  if (warp == 0 && lane == 0) {
    for (int i = 0; i < p.stages; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
    if (XF) for (int i = 0; i < p.stages; ++i) mbar_init(&xfull[i], TC_XF_WARPS);
    mbar_init(tfull, 1);
    mbar_init(tempty, PAIR ? 8 : 4);          // pair: the leader's barrier collects the epilogue warps of both CTAs
    fence_barrier_init();
    tma_prefetch_desc(&p.x_map[0]);
    tma_prefetch_desc(&p.dy_map);
  }
  if (warp == 1) {
    if (PAIR) { tmem_alloc_pair(tmem_slot, p.tmem_cols); tmem_relinquish_pair(); }
    else      { tmem_alloc(tmem_slot, p.tmem_cols); tmem_relinquish(); }
  }
  tc_fence_before();
  if (PAIR) cluster_sync_all();       // the peer's barriers and tensor memory exist before anything targets them
  else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_trigger();
  pdl_wait();
  // pair: p.row_groups counts PAIRS of row groups; CTA `cta_rank` of a pair works on row group 2 * rgp + cta_rank (a
  // group past the end has no valid row blocks: nothing is loaded for it and nothing stored)
  const int total = p.row_groups * p.n_tiles * p.splits;
  const int wstart = PAIR ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;
  const int wstep = PAIR ? (int)(gridDim.x >> 1) : (int)gridDim.x;
  const int rb_per_cta = p.blocksA * p.acc;
  const int nb = PAIR ? p.block_n / p.bwB / 2 : p.block_n / p.bwB;      // dY boxes this CTA loads per stage

  if (warp == 0) {
    // TMA producer: the whole warp loops, one elected lane issues
    int stage = 0;
    uint32_t phase = 0;
    const uint32_t b_tx = (uint32_t)nb * p.b_box_bytes;
    for (int it = wstart; it < total; it += wstep) {
      const int z = it % p.splits;
      const int nt = (it / p.splits) % p.n_tiles;
      const int rgp = it / (p.splits * p.n_tiles);
      const int rg = PAIR ? 2 * rgp + (int)cta_rank : rgp;
      int valid = min(max(p.RB - rg * rb_per_cta, 0), rb_per_cta);
      // pair: only the leader arms its barrier, with the bytes of BOTH CTAs (the peer's loads complete on it too)
      const int valid_peer = PAIR ? min(max(p.RB - (2 * rgp + 1 - (int)cta_rank) * rb_per_cta, 0), rb_per_cta) : 0;
      const uint32_t tx = (uint32_t)(valid + valid_peer) * p.a_box_bytes + (PAIR ? 2u * b_tx : b_tx);
      const int kt_end = min((z + 1) * p.tiles_per_split, p.k_tiles);
      for (int kt = z * p.tiles_per_split; kt < kt_end; ++kt) {
        int q = kt;
        const int w0 = (q % p.tiles_w) * p.PW;
        q /= p.tiles_w;
        const int h0 = (q % p.tiles_h) * p.PH;
        const int n0 = (q / p.tiles_h) * p.PN;
        mbar_wait(&empty[stage], phase ^ 1);
        if (elect_one()) {
          uint8_t* sa = smem + (size_t)stage * p.stage_bytes;
          if (!PAIR || leader) mbar_expect_tx(&full[stage], tx);
          int rb = rg * rb_per_cta;
          int tap = rb / p.cin_blocks, cb = rb - tap * p.cin_blocks;
          for (int j = 0; j < valid; j += p.runA) {      // `valid` is a multiple of runA (whole runs only)
            if (PAIR) tma_load_5d_pair(sa + (size_t)j * p.a_box_bytes, &p.x_map[p.tap_map[tap]], &full[stage], 0,
                                       w0 + p.tap_dw[tap], h0 + p.tap_dh[tap], n0, cb);
            else      tma_load_5d(sa + (size_t)j * p.a_box_bytes, &p.x_map[p.tap_map[tap]], &full[stage], 0,
                                  w0 + p.tap_dw[tap], h0 + p.tap_dh[tap], n0, cb);
            cb += p.runA;
            if (cb >= p.cin_blocks) { cb -= p.cin_blocks; ++tap; }   // runA divides cin_blocks
          }
          uint8_t* sbm = sa + (size_t)rb_per_cta * p.a_box_bytes;
          if (PAIR) tma_load_5d_pair(sbm, &p.dy_map, &full[stage], 0, w0, h0, n0, (2 * nt + (int)cta_rank) * nb);
          else      tma_load_5d(sbm, &p.dy_map, &full[stage], 0, w0, h0, n0, nt * nb);
        }
        __syncwarp();
        if (++stage == p.stages) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    // MMA issuer: the whole warp loops, one elected lane issues (pair: the leader CTA only)
    if (!PAIR || leader) {
    const uint32_t idesc = make_idesc_bf16(PAIR ? 2 * TC_BLOCK_M : TC_BLOCK_M, p.block_n, 1, 1);
    const int ksteps = p.kp / 16;
    const uint32_t rowA = (uint32_t)p.bwA * 2, rowB = (uint32_t)p.bwB * 2;
    const uint32_t ltA = swizzle_layout_type(rowA), ltB = swizzle_layout_type(rowB);
    // MN-major operands: LBO = distance between consecutive channel boxes, SBO = 8 pixel rows; descriptors differ
    // only in the start-address field (bytes >> 4)
    const uint64_t da0 = make_smem_desc(base_u32, p.a_box_bytes, 8 * rowA, ltA);
    const uint64_t db0 = make_smem_desc(base_u32 + (uint32_t)rb_per_cta * p.a_box_bytes, p.b_box_bytes, 8 * rowB, ltB);
    const uint32_t a_step = (16u * rowA) >> 4, b_step = (16u * rowB) >> 4;
    const uint32_t stage16 = p.stage_bytes >> 4, acc16 = ((uint32_t)p.blocksA * p.a_box_bytes) >> 4;
    int stage = 0;
    uint32_t phase = 0, tphase = 0;
    for (int it = wstart; it < total; it += wstep) {
      const int z = it % p.splits;
      const int rgp = it / (p.splits * p.n_tiles);
      const int rg = PAIR ? 2 * rgp : rgp;              // the leader's row group is the fuller one of a pair
      int valid = p.RB - rg * rb_per_cta;
      if (valid > rb_per_cta) valid = rb_per_cta;
      const int nacc = (valid + p.blocksA - 1) / p.blocksA;
      mbar_wait(tempty, tphase ^ 1);
      tc_fence_after();
      const int kt_begin = z * p.tiles_per_split, kt_end = min((z + 1) * p.tiles_per_split, p.k_tiles);
      for (int kt = kt_begin; kt < kt_end; ++kt) {
        mbar_wait(XF ? &xfull[stage] : &full[stage], phase);
        tc_fence_after();
        if (elect_one()) {
          const uint64_t da_s = da0 + (uint64_t)((uint32_t)stage * stage16);
          const uint64_t db_s = db0 + (uint64_t)((uint32_t)stage * stage16);
          for (int a = 0; a < nacc; ++a) {
            const uint64_t da_a = da_s + (uint64_t)((uint32_t)a * acc16);
            const uint32_t d = tmem_base + (uint32_t)(a * p.block_n);
#pragma unroll 4
            for (int ks = 0; ks < ksteps; ++ks) {
              const uint32_t accf = (kt > kt_begin || ks > 0) ? 1u : 0u;
              if (PAIR) umma_bf16_pair(d, da_a + (uint64_t)(ks * a_step), db_s + (uint64_t)(ks * b_step), idesc, accf);
              else      umma_bf16(d, da_a + (uint64_t)(ks * a_step), db_s + (uint64_t)(ks * b_step), idesc, accf);
            }
          }
          // frees the stage (in both CTAs of a pair) once these MMAs have read it
          if (PAIR) umma_commit_pair(&empty[stage]); else umma_commit(&empty[stage]);
          if (kt == kt_end - 1) { if (PAIR) umma_commit_pair(tfull); else umma_commit(tfull); }
        }
        __syncwarp();
        if (++stage == p.stages) { stage = 0; phase ^= 1; }
      }
      tphase ^= 1;
    }
    }
  } else if (XF && warp >= 6) {
    // ===================== operand transform (warps 6..): X tile -> bf16(act(a*x + b)) in place =====================
    const int tt = (warp - 6) * 32 + lane;
    for (int c = tt; c < p.Cin; c += TC_XF_WARPS * 32) {
      float a, b;
      bn_affine(p.xf_gamma[c], p.xf_beta[c], p.xf_mean[c], p.xf_rstd[c], &a, &b);
      xtab[2 * c] = a; xtab[2 * c + 1] = b;
    }
    asm volatile("bar.sync 2, %0;" ::"n"(TC_XF_WARPS * 32) : "memory");
    const int lc = tt & 7, r0 = tt >> 3;
    const float leak = p.xf_leak;
    int stage = 0;
    uint32_t phase = 0;
    for (int it = wstart; it < total; it += wstep) {
      const int z = it % p.splits;
      const int rg = it / (p.splits * p.n_tiles);
      int valid = p.RB - rg * rb_per_cta;
      if (valid > rb_per_cta) valid = rb_per_cta;
      const int kt_end = min((z + 1) * p.tiles_per_split, p.k_tiles);
      for (int kt = z * p.tiles_per_split; kt < kt_end; ++kt) {
        mbar_wait(&full[stage], phase);
        uint8_t* sa = smem + (size_t)stage * p.stage_bytes;
        for (int j = 0; j < valid; ++j) {
          const int cb = (rg * rb_per_cta + j) % p.cin_blocks;             // 64-channel block of this box
          float ab[16];
          const float4* src = reinterpret_cast<const float4*>(xtab + 2 * (cb * 64 + lc * 8));
#pragma unroll
          for (int i = 0; i < 4; ++i) { const float4 v = src[i]; ab[4 * i] = v.x; ab[4 * i + 1] = v.y; ab[4 * i + 2] = v.z; ab[4 * i + 3] = v.w; }
          uint8_t* tile = sa + (size_t)j * p.a_box_bytes;
          for (int r = r0; r < p.kp; r += TC_XF_ROWS) {
            uint4* q = reinterpret_cast<uint4*>(tile + r * 128 + ((lc ^ (r & 7)) << 4));
            *q = bn_transform_chunk(*q, ab, leak);
          }
        }
        fence_proxy_async();
        __syncwarp();
        if (lane == 0) mbar_arrive(&xfull[stage]);
        if (++stage == p.stages) { stage = 0; phase ^= 1; }
      }
    }
  } else {
    const int quad = warp & 3;
    const int row = quad * 32 + lane;
    uint32_t tphase = 0;
    const int64_t rows_total = (int64_t)p.taps * p.Cin;
    const uint32_t tempty_leader = PAIR ? mapa_u32(tempty, 0) : 0u;
    for (int it = wstart; it < total; it += wstep) {
      const int z = it % p.splits;
      const int nt = (it / p.splits) % p.n_tiles;
      const int rgp = it / (p.splits * p.n_tiles);
      const int rg = PAIR ? 2 * rgp + (int)cta_rank : rgp;
      mbar_wait(tfull, tphase);
      tc_fence_after();
      for (int a = 0; a < p.acc; ++a) {
        const int rb0 = rg * rb_per_cta + p.blocksA * a;
        if (rb0 >= p.RB) break;                                   // uniform: this accumulator was not computed
        const int rb = rb0 + row / p.bwA;
        const bool ok = rb < p.RB;
        const int tap = ok ? rb / p.cin_blocks : 0;
        const int cb = rb - tap * p.cin_blocks;
        const int64_t grow = (int64_t)tap * p.Cin + cb * p.bwA + (row % p.bwA);
        float* dst = p.partial + ((int64_t)z * rows_total + grow) * p.Cout + (int64_t)nt * p.block_n;
        const uint32_t taddr = tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(a * p.block_n);
        // A thread holds one ROW of the accumulator; storing it directly would make every store instruction touch
        // 32 different rows (32 half-used sectors).  The warp transposes 32 rows x 32 columns through its own 4 KB
        // slab of shared memory (16-byte chunks XOR-swizzled by row) so that each store instruction writes four
        // complete 128-byte lines.
        uint8_t* slab = epi_slab + (warp - 2) * 4096;
        const unsigned long long dst_u = reinterpret_cast<unsigned long long>(dst);
        int c0 = 0;
        for (; c0 + 32 <= p.block_n; c0 += 32) {
          uint32_t v[32];
          tmem_ld_32x32(taddr + c0, v);
          tmem_ld_wait();
          __syncwarp();                                  // the slab's previous contents have been read
#pragma unroll
          for (int c = 0; c < 8; ++c)
            *reinterpret_cast<uint4*>(slab + lane * 128 + ((c ^ (lane & 7)) << 4)) =
                make_uint4(v[4 * c], v[4 * c + 1], v[4 * c + 2], v[4 * c + 3]);
          __syncwarp();
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const int r = i * 4 + (lane >> 3), c = lane & 7;
            const uint4 val = *reinterpret_cast<const uint4*>(slab + r * 128 + ((c ^ (r & 7)) << 4));
            const unsigned long long d = __shfl_sync(0xffffffffu, dst_u, r);
            const int okr = __shfl_sync(0xffffffffu, (int)ok, r);
            if (okr) *reinterpret_cast<uint4*>(reinterpret_cast<float*>(d) + c0 + 4 * c) = val;
          }
        }
        for (; c0 < p.block_n; c0 += 16) {               // column tail (BLOCK_N a multiple of 16 only)
          uint32_t v[16];
          tmem_ld_32x16(taddr + c0, v);
          tmem_ld_wait();
          if (ok) {
#pragma unroll
            for (int i = 0; i < 4; ++i)
              *reinterpret_cast<uint4*>(dst + c0 + 4 * i) = make_uint4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if (PAIR) mbar_arrive_cluster(tempty_leader);     // the leader's MMA warp waits for both CTAs' epilogues
        else mbar_arrive(tempty);
      }
      tphase ^= 1;
    }
  }
  tc_fence_before();
  if (PAIR) cluster_sync_all();      // neither CTA may leave while the pair's MMAs / remote arrivals can still touch it
  else __syncthreads();
  if (warp == 1) {
    if (PAIR) tmem_dealloc_pair(tmem_base, p.tmem_cols);
    else tmem_dealloc(tmem_base, p.tmem_cols);
  }
}

// Sum of the split-K partials in a fixed order.  A CTA covers 256/S consecutive float4 outputs with S thread
// slices; slice s adds partials s, s+S, ... (independent loads, several in flight), then the slices are combined
// in slice order through shared memory.  With one thread per output the 100+ dependent-latency loads of a
// 148-way split took longer than the GEMM itself.
__global__ void __launch_bounds__(256)
wgrad_reduce_kernel(const float* __restrict__ part, float* __restrict__ out, int64_t n4, int splits, int S) {
  __shared__ float4 sh[256];
  pdl_trigger();
  pdl_wait();
  const int per = 256 / S;
  const int o = threadIdx.x % per, slice = threadIdx.x / per;
  const int64_t i = (int64_t)blockIdx.x * per + o;
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  if (i < n4) {
    const float4* p = reinterpret_cast<const float4*>(part) + i;
    int z = slice;
    for (; z + 3 * S < splits; z += 4 * S) {
      const float4 a = __ldcg(p + (int64_t)z * n4), b = __ldcg(p + (int64_t)(z + S) * n4);
      const float4 c = __ldcg(p + (int64_t)(z + 2 * S) * n4), d = __ldcg(p + (int64_t)(z + 3 * S) * n4);
      acc.x += a.x; acc.y += a.y; acc.z += a.z; acc.w += a.w;
      acc.x += b.x; acc.y += b.y; acc.z += b.z; acc.w += b.w;
      acc.x += c.x; acc.y += c.y; acc.z += c.z; acc.w += c.w;
      acc.x += d.x; acc.y += d.y; acc.z += d.z; acc.w += d.w;
    }
    for (; z < splits; z += S) {
      const float4 a = __ldcg(p + (int64_t)z * n4);
      acc.x += a.x; acc.y += a.y; acc.z += a.z; acc.w += a.w;
    }
  }
  if (S == 1) {
    if (i < n4) reinterpret_cast<float4*>(out)[i] = acc;
    return;
  }
  sh[threadIdx.x] = acc;
  __syncthreads();
  if (slice == 0 && i < n4) {
    for (int q = 1; q < S; ++q) {
      const float4 t = sh[q * per + o];
      acc.x += t.x; acc.y += t.y; acc.z += t.z; acc.w += t.w;
    }
    reinterpret_cast<float4*>(out)[i] = acc;
  }
}

// ------------------------------------------------------------------------------------------
// weight re-layout: fp32 HWIO -> bf16 [tap][Cout][Cin] (fprop B operand) and bf16 [tap][Cin][Cout] (dgrad)
// ------------------------------------------------------------------------------------------
__global__ void weights_to_bf16_kernel(const float* __restrict__ w, int taps, int Cin, int Cout,
                                       __nv_bfloat16* __restrict__ wf, __nv_bfloat16* __restrict__ wd) {
  const int64_t n = (int64_t)taps * Cin * Cout;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int co = (int)(i % Cout);
    const int64_t r = i / Cout;
    const int ci = (int)(r % Cin);
    const int tap = (int)(r / Cin);
    const __nv_bfloat16 v = __float2bfloat16_rn(w[i]);
    if (wd) wd[i] = v;
    if (wf) wf[((int64_t)tap * Cout + co) * Cin + ci] = v;
  }
}

// All conv kernels of the net in ONE launch: a device table of jobs, blockIdx.y = job.
struct RelayoutJob {
  const float* w;
  __nv_bfloat16* wf;
  __nv_bfloat16* wd;
  int taps, Cin, Cout, kpad;      // kpad > 0: stem layout bf16 [Cout][kpad] from [taps*Cin][Cout]
};
__global__ void weights_to_bf16_batched_kernel(const RelayoutJob* __restrict__ jobs) {
  const RelayoutJob j = jobs[blockIdx.y];
  if (j.kpad > 0) {
    const int rows = j.taps * j.Cin;
    const int64_t total = (int64_t)j.Cout * j.kpad;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
      const int kk = (int)(i % j.kpad);
      const int co = (int)(i / j.kpad);
      j.wf[i] = __float2bfloat16_rn(kk < rows ? j.w[(int64_t)kk * j.Cout + co] : 0.f);
    }
    return;
  }
  const int64_t n = (int64_t)j.taps * j.Cin * j.Cout;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int co = (int)(i % j.Cout);
    const int64_t r = i / j.Cout;
    const int ci = (int)(r % j.Cin);
    const int tap = (int)(r / j.Cin);
    const __nv_bfloat16 v = __float2bfloat16_rn(j.w[i]);
    if (j.wd) j.wd[i] = v;
    if (j.wf) j.wf[((int64_t)tap * j.Cout + co) * j.Cin + ci] = v;
  }
}

// ------------------------------------------------------------------------------------------
// host side: tensor maps and plans
// ------------------------------------------------------------------------------------------
typedef void (*conv_tc_fn)(TcConvParams);
static conv_tc_fn conv_tc_variant(bool fused, bool pair, bool halo, bool xf) {
  if (xf) return fused ? conv_tc_kernel<true, false, false, true> : conv_tc_kernel<false, false, false, true>;
  if (halo) {
    if (fused) return pair ? conv_tc_kernel<true, true, true, false> : conv_tc_kernel<true, false, true, false>;
    return pair ? conv_tc_kernel<false, true, true, false> : conv_tc_kernel<false, false, true, false>;
  }
  if (fused) return pair ? conv_tc_kernel<true, true, false, false> : conv_tc_kernel<true, false, false, false>;
  return pair ? conv_tc_kernel<false, true, false, false> : conv_tc_kernel<false, false, false, false>;
}
static cudaError_t set_conv_tc_smem_attr() {
  for (int v = 0; v < 16; ++v) {
    if ((v & 8) && (v & 6)) continue;                  // the operand transform has no pair / halo variants
    cudaError_t e = cudaFuncSetAttribute(conv_tc_variant(v & 1, v & 2, v & 4, v & 8), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(227 * 1024));
    if (e != cudaSuccess) return e;
  }
  return cudaSuccess;
}

// Launch with the programmatic-serialization attribute and, for CTA pairs, a cluster dimension of two.
template <typename... KArgs, typename... Args>
static void launch_conv(void (*kernel)(KArgs...), int grid, int block, size_t smem, cudaStream_t st, bool pair, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(block);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  int n = 0;
  if (pdl_enabled()) {
    attr[n].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[n].val.programmaticStreamSerializationAllowed = 1;
    ++n;
  }
  if (pair) {
    attr[n].id = cudaLaunchAttributeClusterDimension;
    attr[n].val.clusterDim.x = 2;
    attr[n].val.clusterDim.y = 1;
    attr[n].val.clusterDim.z = 1;
    ++n;
  }
  cfg.attrs = attr;
  cfg.numAttrs = n;
  cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);      // errors surface through check_launch()
}

// CTA pairs (tcgen05 .cta_group::2): RN_TC_PAIR=0 turns them off, =1 uses them wherever the tile allows (default:
// where the weight tile is the larger share of the operand traffic, see create_sub)
static int pair_mode() {
  const char* e = getenv("RN_TC_PAIR");          // read on every plan creation: tests toggle it in-process
  return e ? atoi(e) : -1;
}

typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static PFN_encodeTiled get_encode() {
  static PFN_encodeTiled fn = nullptr;
  if (fn) return fn;
  void* p = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || p == nullptr)
    return nullptr;
  fn = reinterpret_cast<PFN_encodeTiled>(p);
  return fn;
}

static CUtensorMapSwizzle swz_mode(int row_bytes) {
  return row_bytes == 128 ? CU_TENSOR_MAP_SWIZZLE_128B
                          : (row_bytes == 64 ? CU_TENSOR_MAP_SWIZZLE_64B
                                             : (row_bytes == 32 ? CU_TENSOR_MAP_SWIZZLE_32B : CU_TENSOR_MAP_SWIZZLE_NONE));
}

// bf16 tensor map of rank `rank`: dims[0] is the contiguous dimension; strides_bytes[i] is the byte stride of dim i+1.
static int encode_map(CUtensorMap* m, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                      const uint32_t* box, int row_bytes, const char* what) {
  PFN_encodeTiled enc = get_encode();
  if (!enc) return fail(RN_ERR_CUDA, "convtc: cuTensorMapEncodeTiled is not available");
  cuuint64_t gd[5], gs[4];
  cuuint32_t bx[5], es[5];
  for (int i = 0; i < rank; ++i) { gd[i] = dims[i]; bx[i] = box[i]; es[i] = 1; }
  for (int i = 0; i + 1 < rank; ++i) gs[i] = strides_bytes[i];
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, const_cast<void*>(base), gd, gs, bx, es,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, swz_mode(row_bytes), CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(RN_ERR_CUDA, "convtc: cuTensorMapEncodeTiled(%s) failed with %d", what, (int)r);
  return RN_OK;
}

// Pixel box: BW x BH x BN <= 128 output pixels, chosen to minimise the number of tiles.
static void choose_pixel_box(int Wo, int Ho, int N, int* BW, int* BH, int* BN) {
  long best_tiles = -1;
  int bw = 1, bh = 1, bn = 1;
  for (int w = 1; w <= Wo && w <= 128; ++w) {
    for (int h = 1; h <= Ho && w * h <= 128; ++h) {
      int n = 128 / (w * h);
      if (n > N) n = N;
      if (n < 1) continue;
      const long tiles = (long)((Wo + w - 1) / w) * ((Ho + h - 1) / h) * ((N + n - 1) / n);
      const bool better = best_tiles < 0 || tiles < best_tiles || (tiles == best_tiles && w > bw);
      if (better) { best_tiles = tiles; bw = w; bh = h; bn = n; }
    }
  }
  *BW = bw; *BH = bh; *BN = bn;
}

// Halo mode box: one image (BN = 1), BW x BH <= 128 pixels.  Maximise the fraction of the 128 MMA rows that are real
// output pixels, then minimise the rows loaded per tile, (BH + 2) * BW.
static double choose_halo_box(int Wo, int Ho, int* BW, int* BH) {
  double best_eff = -1.0;
  long best_rows = 0;
  int bw = 1, bh = 1;
  for (int w = 1; w <= Wo && w <= 128; ++w)
    for (int h = 2; h <= Ho && w * h <= 128; ++h) {
      const long tiles = (long)((Wo + w - 1) / w) * ((Ho + h - 1) / h);
      const double eff = (double)Wo * Ho / ((double)tiles * 128.0);
      const long rows = (long)(h + 2) * w * tiles;
      if (eff > best_eff + 1e-9 || (eff > best_eff - 1e-9 && rows < best_rows)) { best_eff = eff; best_rows = rows; bw = w; bh = h; }
    }
  if (const char* e = getenv("RN_TC_HALO_BW")) {           // probing knobs: force the box
    const int w = atoi(e);
    const char* e2 = getenv("RN_TC_HALO_BH");
    int h = e2 ? atoi(e2) : 128 / (w > 0 ? w : 1);
    if (h > Ho) h = Ho;
    if (w >= 1 && h >= 2 && w * h <= 128) {
      bw = w; bh = h;
      best_eff = (double)Wo * Ho / ((double)((Wo + w - 1) / w) * ((Ho + h - 1) / h) * 128.0);
    }
  }
  *BW = bw; *BH = bh;
  return best_eff;
}

static int halo_mode() {
  const char* e = getenv("RN_TC_HALO");          // 1 = wherever possible; default off (measured: no gain, see create_sub)
  return e ? atoi(e) : 0;
}

static int max_block_n() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("RN_TC_MAX_BLOCK_N");     // tuning knob
    v = e ? atoi(e) : 256;
    if (v < 16 || v > 256) v = 256;
  }
  return v;
}
static int choose_block_n(int ncols) {
  for (int b = max_block_n() / 16 * 16; b >= 16; b -= 16)
    if (ncols % b == 0) return b;
  return 0;
}


// Power-of-two pixel box with exactly `kp` pixels (the K tile of wgrad must be completely written by
// TMA: out-of-image pixels are zero-filled, but rows the box does not cover would be stale smem).
static void choose_k_box(int Wo, int Ho, int N, int kp, int* PW, int* PH, int* PN) {
  long best = -1;
  int bw = 1, bh = 1, bn = kp;
  for (int w = 1; w <= kp; w <<= 1)
    for (int h = 1; w * h <= kp; h <<= 1) {
      const int n = kp / (w * h);
      const long tiles = (long)((Wo + w - 1) / w) * ((Ho + h - 1) / h) * ((N + n - 1) / n);
      if (best < 0 || tiles < best || (tiles == best && w > bw)) { best = tiles; bw = w; bh = h; bn = n; }
    }
  *PW = bw; *PH = bh; *PN = bn;
}

// Activation tensor map(s).  stride 1: one map of [N,H,W,C].  stride 2: one map per (row, col) parity
// of the input — a strided sub-view with base offset (ph*W + pw)*C and doubled W/H strides — so that a
// tap (r, s) reads a dense box of view (r-pad)&1, (s-pad)&1 at offset floor((r-pad)/2), floor((s-pad)/2).
static int encode_activation_maps(CUtensorMap* maps, const void* base, int C, int W, int H, int N, int stride,
                                  const uint32_t* box, int row_bytes) {
  if (stride == 1) {
    const uint64_t dims[4] = {(uint64_t)C, (uint64_t)W, (uint64_t)H, (uint64_t)N};
    const uint64_t str[3] = {(uint64_t)C * 2, (uint64_t)W * C * 2, (uint64_t)H * W * C * 2};
    return encode_map(&maps[0], base, 4, dims, str, box, row_bytes, "activation");
  }
  for (int ph = 0; ph < 2; ++ph)
    for (int pw = 0; pw < 2; ++pw) {
      const int W2 = (W - pw + 1) / 2, H2 = (H - ph + 1) / 2;
      if (W2 <= 0 || H2 <= 0) continue;
      const uint64_t dims[4] = {(uint64_t)C, (uint64_t)W2, (uint64_t)H2, (uint64_t)N};
      const uint64_t str[3] = {(uint64_t)2 * C * 2, (uint64_t)2 * W * C * 2, (uint64_t)H * W * C * 2};
      const char* b = (const char*)base + ((size_t)ph * W + pw) * C * 2;
      int rc = encode_map(&maps[ph * 2 + pw], b, 4, dims, str, box, row_bytes, "activation parity view");
      if (rc) return rc;
    }
  return RN_OK;
}

// Same views with the channel axis split as (bw channels inside a block, ..., block index) so that one 5-D
// box fetches `nblk` consecutive channel blocks: each block lands as its own [pixels][bw] swizzled tile.
static int encode_activation_maps_5d(CUtensorMap* maps, const void* base, int C, int W, int H, int N, int stride, int bw,
                                     int pw_, int ph_, int pn_, int nblk) {
  const int nph = stride == 1 ? 1 : 2;
  for (int ph = 0; ph < nph; ++ph)
    for (int pw = 0; pw < nph; ++pw) {
      const int W2 = stride == 1 ? W : (W - pw + 1) / 2, H2 = stride == 1 ? H : (H - ph + 1) / 2;
      if (W2 <= 0 || H2 <= 0) continue;
      const uint64_t st = (uint64_t)stride;
      const uint64_t dims[5] = {(uint64_t)bw, (uint64_t)W2, (uint64_t)H2, (uint64_t)N, (uint64_t)(C / bw)};
      const uint64_t str[4] = {st * C * 2, st * W * C * 2, (uint64_t)H * W * C * 2, (uint64_t)bw * 2};
      const uint32_t box[5] = {(uint32_t)bw, (uint32_t)pw_, (uint32_t)ph_, (uint32_t)pn_, (uint32_t)nblk};
      const char* b = (const char*)base + ((size_t)ph * W + pw) * C * 2;
      int rc = encode_map(&maps[ph * 2 + pw], b, 5, dims, str, box, bw * 2, "activation (5-D)");
      if (rc) return rc;
    }
  return RN_OK;
}

// An explicitly strided [N,H,W,C] bf16 view (strides in bytes; the W stride may be SMALLER than C*2, i.e. the
// "channel" rows of neighbouring pixels overlap: the space-to-depth stem reads a 4-pixel window per row).
struct ActView { int C, W, H, N; uint64_t sw, sh, sn; };

static int encode_view(CUtensorMap* m, const void* base, const ActView& v, const uint32_t* box, int row_bytes) {
  const uint64_t dims[4] = {(uint64_t)v.C, (uint64_t)v.W, (uint64_t)v.H, (uint64_t)v.N};
  const uint64_t str[3] = {v.sw, v.sh, v.sn};
  return encode_map(m, base, 4, dims, str, box, row_bytes, "activation view");
}
static int encode_view_5d(CUtensorMap* m, const void* base, const ActView& v, int bw, int pw_, int ph_, int pn_, int nblk) {
  const uint64_t dims[5] = {(uint64_t)bw, (uint64_t)v.W, (uint64_t)v.H, (uint64_t)v.N, (uint64_t)(v.C / bw)};
  const uint64_t str[4] = {v.sw, v.sh, v.sn, (uint64_t)bw * 2};
  const uint32_t box[5] = {(uint32_t)bw, (uint32_t)pw_, (uint32_t)ph_, (uint32_t)pn_, (uint32_t)nblk};
  return encode_map(m, base, 5, dims, str, box, bw * 2, "activation view (5-D)");
}

// tap (r, s) of a forward-style gather at `stride`: which map and which box offset
static void forward_tap(int r, int s, int pad, int stride, int8_t* map, int8_t* dw, int8_t* dh) {
  if (stride == 1) { *map = 0; *dw = (int8_t)(s - pad); *dh = (int8_t)(r - pad); return; }
  const int ah = r - pad, aw = s - pad;
  const int ph = ah & 1, pw = aw & 1;
  *map = (int8_t)(ph * 2 + pw);
  *dh = (int8_t)((ah - ph) / 2);
  *dw = (int8_t)((aw - pw) / 2);
}

struct WgradShape {
  int pair;
  int block_n, n_tiles, acc, cin_blocks, RB, row_groups, kp, stages, PW, PH, PN, tiles_w, tiles_h, tiles_n, k_tiles;
  int splits, tiles_per_split, bwA, bwB, blocksA, runA;
  uint32_t a_box_bytes, b_box_bytes, stage_bytes, tmem_cols;
  size_t smem, scratch;
};

static int box_width(int c) { return c % 64 == 0 ? 64 : (c % 32 == 0 ? 32 : (c % 16 == 0 ? 16 : 0)); }

// CTA pairs for wgrad: RN_WG_PAIR=0 off, =1 wherever the dY tile splits in two, default see wgrad_shape.
static int wgrad_pair_mode() {
  const char* e = getenv("RN_WG_PAIR");            // read on every plan creation: tests toggle it in-process
  return e ? atoi(e) : -1;
}

static bool wgrad_shape(const rn_conv_desc* d, WgradShape* w, bool allow_pair = true) {
  w->pair = 0;
  w->bwA = box_width(d->Cin);
  w->block_n = choose_block_n(d->Cout);
  if (w->block_n == 0) return false;
  // dY boxes must tile one BLOCK_N column tile exactly: 320 / 640 output channels (WRN-28-10) run as BLOCK_N = 160
  // = five 32-channel boxes, although the tensor itself would allow 64-channel boxes
  w->bwB = box_width(w->block_n);
  if (!w->bwA || !w->bwB || d->Cout % w->bwB) return false;
  w->n_tiles = d->Cout / w->block_n;
  w->blocksA = 128 / w->bwA;
  w->cin_blocks = d->Cin / w->bwA;
  w->RB = d->k * d->k * w->cin_blocks;
  const int nb = w->block_n / w->bwB;
  static int f_acc = -1, f_kp = -1;
  if (f_acc < 0) { const char* e = getenv("RN_WG_ACC"); f_acc = e ? atoi(e) : 0; }   // tuning knobs
  if (f_kp < 0) { const char* e = getenv("RN_WG_KP"); f_kp = e ? atoi(e) : 0; }
  // Choose (accumulators per CTA, pixels per stage).  Measured on B200: a stage costs ~300 cycles per TMA
  // op issued, so the winner is the configuration with the fewest TMA ops per MAC that still leaves >= 3
  // pipeline stages in shared memory (ties: fewer bytes per MAC).
  double best = 1e30;
  bool found = false;
  const int need_acc = (w->RB + w->blocksA - 1) / w->blocksA;
  for (int acc = 1; acc <= 4; acc <<= 1) {
    if (acc * w->block_n > 512) break;
    if (acc > 1 && acc / 2 >= need_acc) break;
    if (f_acc && acc != f_acc) continue;
    const int rbpc = w->blocksA * acc;
    int run;
    if (w->cin_blocks >= rbpc) run = (w->cin_blocks % rbpc == 0) ? rbpc : 1;
    else run = (rbpc % w->cin_blocks == 0) ? w->cin_blocks : 1;
    for (int kp = 128; kp >= 32; kp >>= 1) {
      if (f_kp && kp != f_kp) continue;
      const uint32_t abox = (uint32_t)kp * w->bwA * 2, bbox = (uint32_t)kp * w->bwB * 2;
      if (abox % 1024 || bbox % 1024) continue;            // swizzle atoms must stay 1024-byte aligned
      const size_t stage = (size_t)rbpc * abox + (size_t)nb * bbox;
      const int stages = (int)((204 * 1024) / stage);    // 227 KB - epilogue transpose slabs, barriers, alignment
      if (stages < 3) continue;
      const double ops = (double)rbpc / run + 1.0;
      const double macs = (double)acc * kp * w->block_n;
      const double cost = ops / macs * 1e6 + (double)stage / macs * 1e-3;
      if (cost < best) {
        best = cost; found = true;
        w->acc = acc; w->kp = kp; w->runA = run;
        w->a_box_bytes = abox; w->b_box_bytes = bbox; w->stage_bytes = (uint32_t)stage;
        w->stages = stages > 8 ? 8 : stages;
      }
    }
  }
  if (!found) return false;
  w->row_groups = (w->RB + w->blocksA * w->acc - 1) / (w->blocksA * w->acc);
  int sms = RN_NUM_SMS;
  {
    const int pm = wgrad_pair_mode();
    // default: wherever the tile allows it, unless an odd number of row groups would leave a tenth or more of the
    // pairs' tensor-core time on an empty row group.  Measured on ResNet-50 (profiles/r02_wgrad_pair_ab.txt): the
    // gain comes as much from the halved split-K factor (half the partial tiles written and re-read) as from the
    // cheaper operand fetch: 1x1 256 -> 1024 at 14x14 -11 %, 1x1/2 256 -> 512 -14 %, 3x3 256 -> 256 (9 row groups) +5 %.
    const int rg = (w->row_groups < 2 && w->acc >= 2) ? 2 : w->row_groups;
    const bool want = pm == 1 || (pm < 0 && (rg % 2 == 0 || rg >= 11));
    // a pair needs two row groups: when one CTA's accumulators cover all rows, each CTA of the pair takes half of them
    if (allow_pair && want && pm != 0 && nb % 2 == 0 && w->row_groups < 2 && w->acc >= 2 &&
        w->RB > w->blocksA * (w->acc / 2)) {
      w->acc /= 2;
      const int rbpc = w->blocksA * w->acc;
      if (w->cin_blocks >= rbpc) w->runA = (w->cin_blocks % rbpc == 0) ? rbpc : 1;
      else w->runA = (rbpc % w->cin_blocks == 0) ? w->cin_blocks : 1;
      w->row_groups = (w->RB + rbpc - 1) / rbpc;
    }
    if (allow_pair && want && pm != 0 && nb % 2 == 0 && w->row_groups >= 2) {
      w->pair = 1;
      w->row_groups = (w->row_groups + 1) / 2;         // pairs of row groups
      sms = RN_NUM_SMS / 2;                            // work items are executed by pairs of CTAs
      const size_t stage = (size_t)w->blocksA * w->acc * w->a_box_bytes + (size_t)(nb / 2) * w->b_box_bytes;
      const int stages = (int)((204 * 1024) / stage);
      w->stage_bytes = (uint32_t)stage;
      w->stages = stages > 8 ? 8 : stages;
    }
  }
  w->smem = (size_t)w->stages * w->stage_bytes + 16384 + 2048;
  uint32_t cols = 32;
  while (cols < (uint32_t)(w->acc * w->block_n)) cols <<= 1;
  w->tmem_cols = cols;
  choose_k_box(d->Wo, d->Ho, d->N, w->kp, &w->PW, &w->PH, &w->PN);
  w->tiles_w = (d->Wo + w->PW - 1) / w->PW;
  w->tiles_h = (d->Ho + w->PH - 1) / w->PH;
  w->tiles_n = (d->N + w->PN - 1) / w->PN;
  w->k_tiles = w->tiles_w * w->tiles_h * w->tiles_n;
  // split-K factor: as many work items as fit in ONE wave of the persistent grid (rounding up would leave a
  // second, almost empty wave that doubles the kernel time)
  const int base = w->row_groups * w->n_tiles;
  int splits = sms / base;
  const int max_splits = w->k_tiles / 8 > 0 ? w->k_tiles / 8 : 1;
  if (splits > max_splits) splits = max_splits;
  if (splits < 1) splits = 1;
  w->tiles_per_split = (w->k_tiles + splits - 1) / splits;
  w->splits = (w->k_tiles + w->tiles_per_split - 1) / w->tiles_per_split;
  w->scratch = w->splits > 1 ? (size_t)w->splits * d->k * d->k * d->Cin * d->Cout * sizeof(float) : 0;
  return true;
}

}  // namespace rn

using namespace rn;

struct rn_convtc_plan {
  int kind;
  int nsub;                 // GEMM sub-problems (4 output-parity classes for a stride-2 dgrad, else 1)
  TcConvParams p[4];
  int grid[4];
  size_t smem_sub[4];
  void* clear_ptr;          // output to zero first (stride-2 dgrad whose taps do not reach every parity)
  size_t clear_bytes;
  TcWgradParams w;
  int wgrid;
  size_t wsmem;
  float* dw;
  int64_t dw_elems;
};

extern "C" int rn_convtc_supported(const rn_conv_desc* d, int kind) {
  if (!d || d->dtype != RN_BF16) return 0;
  if (kind < 0 || kind > 2) return 0;
  if (d->stride != 1 && d->stride != 2) return 0;
  if (d->k * d->k > TC_MAX_TAPS || d->pad > 8) return 0;
  if (d->Cin % 8 || d->Cout % 8) return 0;
  if (kind == 2) {
    WgradShape w;
    return wgrad_shape(d, &w) ? 1 : 0;
  }
  const int cred = kind == 0 ? d->Cin : d->Cout;
  const int ncols = kind == 0 ? d->Cout : d->Cin;
  if (cred < 16 || choose_block_n(ncols) == 0) return 0;
  return 1;
}

extern "C" size_t rn_convtc_scratch_bytes(const rn_conv_desc* d, int kind) {
  if (!d || kind != 2) return 0;
  WgradShape w, w1;
  if (!wgrad_shape(d, &w)) return 0;
  if (w.pair && wgrad_shape(d, &w1, false) && w1.scratch > w.scratch) return w1.scratch;   // a plan with the operand transform never pairs
  return w.scratch;
}

// One GEMM sub-problem.  `oW,oH` and `ostr` describe the (possibly strided) view of the output that this
// sub-problem writes; taps[] lists (map, dw, dh, weight tap).
struct TapSpec { int8_t map, dw, dh, b; };
// dgrad whose output is the dy of a BN+ReLU: the epilogue accumulates that BN's backward reductions
struct BnBwdOpts { const void* x; const float *gamma, *beta, *mean, *rstd; float leak; };
// fprop / wgrad whose activation operand is the INPUT of a training-mode BN+ReLU, normalised on the fly
struct BnOperand { const float *gamma, *beta, *mean, *rstd; float leak; };

static int create_sub(const rn_conv_desc* d, int kind, const void* a, const void* b, const void* out_base,
                      int oW, int oH, const uint64_t* ostr, const TapSpec* taps, int ntaps, TcConvParams* pp,
                      int* grid, size_t* smem_bytes, const ActView* view = nullptr, int b_taps = 0,
                      const void* residual = nullptr, float* stat_partials = nullptr, const BnBwdOpts* bnbwd = nullptr,
                      const BnOperand* xf = nullptr) {
  TcConvParams& p = *pp;
  memset(&p, 0, sizeof(p));
  if (xf) {
    if (kind != 0 || ntaps != 1 || view != nullptr || (d->Cin % 64) != 0)
      return fail(RN_ERR_UNSUPPORTED, "convtc: the BN operand transform needs a 1x1 fprop over a multiple of 64 channels");
    p.xf = 1; p.xf_channels = d->Cin;
    p.xf_gamma = xf->gamma; p.xf_beta = xf->beta; p.xf_mean = xf->mean; p.xf_rstd = xf->rstd; p.xf_leak = xf->leak;
  }
  const int cred = kind == 0 ? d->Cin : d->Cout;
  const int ncols = kind == 0 ? d->Cout : d->Cin;
  const int aW = kind == 0 ? d->W : d->Wo, aH = kind == 0 ? d->H : d->Ho;   // activation tensor that is gathered
  p.kc = cred >= 64 ? 64 : (cred >= 32 ? 32 : 16);
  p.k_chunks = (cred + p.kc - 1) / p.kc;
  p.block_n = choose_block_n(ncols);
  p.n_tiles = ncols / p.block_n;
  p.gw = p.block_n % 64 == 0 ? 64 : (p.block_n % 32 == 0 ? 32 : 16);
  choose_pixel_box(oW, oH, d->N, &p.BW, &p.BH, &p.BN);
  p.taps = ntaps;
  for (int t = 0; t < ntaps; ++t) {
    p.tap_map[t] = taps[t].map; p.tap_dw[t] = taps[t].dw; p.tap_dh[t] = taps[t].dh; p.tap_b[t] = taps[t].b;
  }
  // Halo mode: a 3x3 stencil at stride 1 (all taps on map 0, shifts in {-1, 0, 1}^2), 64-channel k-chunks.
  {
    bool stencil = ntaps == 9 && view == nullptr && p.kc == 64;
    for (int t = 0; t < ntaps && stencil; ++t)
      stencil = taps[t].map == 0 && taps[t].dw >= -1 && taps[t].dw <= 1 && taps[t].dh >= -1 && taps[t].dh <= 1;
    const int mode = halo_mode();
    if (stencil && mode != 0) {
      int hw_ = 1, hh_ = 2;
      const double eff = choose_halo_box(oW, oH, &hw_, &hh_);
      const long tiles_now = (long)((oW + p.BW - 1) / p.BW) * ((oH + p.BH - 1) / p.BH) * ((d->N + p.BN - 1) / p.BN);
      const double eff_now = (double)oW * oH * d->N / ((double)tiles_now * 128.0);
      // A rows fetched per output pixel: 9 per-tap boxes, 3 * (BH + 2) / BH in halo mode.  MEASURED on B200
      // (profiles/r02_halo_sweep.txt): no layer of ResNet-50 gets faster — 56x56 64->64: 0.0573 ms per-tap vs 0.0603 ms
      // best halo box (8x16), 28x28 128->128: 0.0389 vs 0.0451, 14x14 256->256: 0.0379 vs 0.0388 — because these
      // kernels are not bound by L2 -> shared-memory traffic but by the rate at which tcgen05.mma fetches its operands
      // from shared memory (~45 B/clk/SM), which the halo does not change, while a one-image box
      // wastes an eighth of the MMA rows.  Halo mode therefore stays opt-in (RN_TC_HALO=1).
      (void)eff; (void)eff_now;
      if (mode == 1) {
        p.halo = 1;
        p.BW = hw_; p.BH = hh_; p.BN = 1;
        // group the taps by horizontal shift (first-seen order); inside a group the vertical taps keep weight order
        int ng = 0;
        int8_t seen[3] = {0, 0, 0};
        for (int t = 0; t < ntaps; ++t) {
          int g = -1;
          for (int q = 0; q < ng; ++q) if (p.halo_dw[q] == taps[t].dw) g = q;
          if (g < 0) { g = ng++; p.halo_dw[g] = taps[t].dw; seen[g] = 0; }
          p.halo_row[g][seen[g]] = (int8_t)(taps[t].dh + 1);
          p.halo_b[g][seen[g]] = taps[t].b;
          seen[g]++;
        }
        p.halo_dh0 = -1;
      }
    }
  }
  p.tiles_w = (oW + p.BW - 1) / p.BW;
  p.tiles_h = (oH + p.BH - 1) / p.BH;
  p.tiles_n = (d->N + p.BN - 1) / p.BN;
  const int row_bytes = p.kc * 2;
  const uint32_t box_rows = (uint32_t)(p.BW * p.BH * p.BN);
  // halo mode: the box holds BH + 2 rows, and the MMA of the LAST vertical tap reads 128 rows from row 2 * BW on
  const uint32_t a_rows = p.halo ? (uint32_t)(2 * p.BW) + TC_BLOCK_M : TC_BLOCK_M;
  p.a_bytes = (a_rows * row_bytes + 1023u) & ~1023u;
  // CTA pair: one tcgen05.mma.cta_group::2 stream over two neighbouring pixel-tile groups; each CTA stages half of
  // the B tile.  Needs two whole 64-column halves (MN-major boxes) / 8-row groups (K-major) and full swizzle atoms.
  {
    const int mode = pair_mode();
    const bool can = p.block_n % 128 == 0 && p.kc == 64 && p.gw == 64;
    // Measured per layer on B200 (ResNet-50 / 224, profiles/r02_pair_ab.txt): pairs win 4-33 % where the GEMM is
    // operand-traffic bound — N = 256 tiles with a long reduction (3x3 over >= 256 channels, 1x1 over >= 1024) —
    // and lose 10-50 % on the HBM-bound large-image layers (short reductions, one or two k-chunks per tile), where
    // running two CTAs in lock step only removes slack.  Default: N = 256 and taps * channels >= 1024.
    p.pair = can && !p.xf && (mode == 1 || (mode != 0 && p.block_n >= 256 && (long)ntaps * cred >= 1024)) ? 1 : 0;
  }
  const int b_rows = p.pair ? p.block_n / 2 : p.block_n;           // B rows (columns of the output) staged per CTA
  p.b_bytes = ((uint32_t)b_rows * row_bytes + 1023u) & ~1023u;
  p.tx_bytes = box_rows * row_bytes + (uint32_t)p.block_n * row_bytes;
  // pixel tiles per CTA iteration: every extra tile re-uses the B tile (less L2 -> SM traffic per FLOP,
  // which is what bounds these kernels), limited by the 512 TMEM columns, by shared memory (>= 3 stages)
  // and by keeping every SM busy
  const long m_tiles = (long)p.tiles_w * p.tiles_h * p.tiles_n;
  // epilogue staging, barriers, alignment slack; fused epilogue: residual tile ring, statistics scratch
  const bool want_fused = residual != nullptr || stat_partials != nullptr;
  if (bnbwd ? (kind != 1 || stat_partials == nullptr || residual != nullptr) : (want_fused && kind != 0))
    return fail(RN_ERR_UNSUPPORTED, "convtc: fused add / statistics belong to fprop, fused BN backward sums to dgrad");
  p.fuse_add = residual != nullptr;
  p.fuse_bnbwd = bnbwd != nullptr;
  p.fuse_stats = stat_partials != nullptr && !bnbwd;
  if (bnbwd) {
    residual = bnbwd->x;                     // the BN input rides the residual-tile ring
    p.bn_gamma = bnbwd->gamma; p.bn_beta = bnbwd->beta; p.bn_mean = bnbwd->mean; p.bn_rstd = bnbwd->rstd;
    p.bn_leak = bnbwd->leak;
  }
  p.ncols_total = ncols;
  p.oW = oW; p.oH = oH; p.oN = d->N;
  p.stat_partials = stat_partials;
  {
    // measured on B200 (ResNet-50 step): 2 slots 9.71 ms, 3 slots 9.81 ms, 4 slots 9.90 ms — the shared memory is
    // worth more as an extra operand pipeline stage than as a deeper tile ring
    static int slots = -1;
    if (slots < 0) { const char* e = getenv("RN_TC_RES_SLOTS"); slots = e ? atoi(e) : 2; if (slots < 2 || slots > TC_RES_SLOTS_MAX) slots = 2; }
    p.res_slots = slots;
  }
  const size_t fixed = 2 * 16384 + 1024 + 1024 + ((p.fuse_add || p.fuse_bnbwd) ? (size_t)p.res_slots * 16384 : 0) +
                       ((p.fuse_stats || p.fuse_bnbwd) ? 8192 + (size_t)ncols * 8 : 0) +
                       (p.fuse_bnbwd ? (size_t)ncols * 16 : 0) + (p.xf ? (size_t)p.xf_channels * 8 : 0);
  const size_t budget = 227 * 1024;
  int acc_m = 1;
  {
    static int forced = -1;
    if (forced < 0) { const char* e = getenv("RN_TC_ACC_M"); forced = e ? atoi(e) : 0; }
    for (int cand = 4; cand >= 2; cand >>= 1) {
      if (forced && cand != forced) continue;
      if (cand * p.block_n > 512) continue;
      if ((budget - fixed) / ((size_t)cand * p.a_bytes + (size_t)(p.halo ? 3 : 1) * p.b_bytes) < 3) continue;
      // measured on B200: sharing B across two pixel tiles pays (+10..18 %) only while TMEM can still be
      // double-buffered (2 * cand * BLOCK_N <= 512) and every SM keeps at least one work item
      if (!forced && (2 * cand * p.block_n > 512 || cand > 2)) continue;
      if (!forced && ((m_tiles + cand - 1) / cand) * p.n_tiles < RN_NUM_SMS) continue;
      if (p.pair && cand > 1) continue;          // pairs already double the pixels per B tile
      acc_m = cand;
      break;
    }
  }
  p.acc_m = acc_m;
  p.tstages = (2 * acc_m * p.block_n <= 512) ? 2 : 1;
  uint32_t cols = 32;
  while (cols < (uint32_t)(p.tstages * acc_m * p.block_n)) cols <<= 1;
  p.tmem_cols = cols;
  const size_t stage_sz = (size_t)acc_m * p.a_bytes + (size_t)(p.halo ? 3 : 1) * p.b_bytes;
  int stages = (int)((budget - fixed) / stage_sz);
  if (stages > 8) stages = 8;
  if (stages < 2) return fail(RN_ERR_UNSUPPORTED, "convtc: tile does not fit shared memory");
  p.stages = stages;
  *smem_bytes = fixed + (size_t)stages * stage_sz;
  int rc;
  const uint32_t abox[4] = {(uint32_t)p.kc, (uint32_t)p.BW, (uint32_t)(p.halo ? p.BH + 2 : p.BH), (uint32_t)p.BN};
  if (b_taps == 0) b_taps = d->k * d->k;
  rc = view ? encode_view(&p.a_map[0], a, *view, abox, row_bytes)
            : encode_activation_maps(p.a_map, a, cred, aW, aH, d->N, kind == 0 ? d->stride : 1, abox, row_bytes);
  if (rc) return rc;
  if (kind == 0) {
    // fprop: bf16 HWIO [tap][Cin][Cout] read as an MN-major B operand: dims (Cout, Cin, taps)
    p.b_mn = 1;
    p.b_boxes = b_rows / p.gw;
    p.b_box_bytes = (uint32_t)p.kc * p.gw * 2;
    // dims (gw couts inside a block, cin, cout block, tap): one TMA op fetches all cout blocks of the tile
    const uint64_t dims[4] = {(uint64_t)p.gw, (uint64_t)cred, (uint64_t)(ncols / p.gw), (uint64_t)b_taps};
    const uint64_t str[3] = {(uint64_t)ncols * 2, (uint64_t)p.gw * 2, (uint64_t)ncols * cred * 2};
    const uint32_t box[4] = {(uint32_t)p.gw, (uint32_t)p.kc, (uint32_t)p.b_boxes, 1u};
    rc = encode_map(&p.b_map, b, 4, dims, str, box, p.gw * 2, "weights (MN-major)");
    if (rc) return rc;
  } else {
    // dgrad: the same bf16 HWIO tensor is K-major for this GEMM (K = Cout contiguous): dims (cred, ncols, taps)
    const uint64_t dims[3] = {(uint64_t)cred, (uint64_t)ncols, (uint64_t)b_taps};
    const uint64_t str[2] = {(uint64_t)cred * 2, (uint64_t)ncols * cred * 2};
    const uint32_t box[3] = {(uint32_t)p.kc, (uint32_t)b_rows, 1u};
    rc = encode_map(&p.b_map, b, 3, dims, str, box, row_bytes, "weights");
    if (rc) return rc;
  }
  {  // output view: dims (ncols, oW, oH, N)
    const uint64_t dims[4] = {(uint64_t)ncols, (uint64_t)oW, (uint64_t)oH, (uint64_t)d->N};
    const uint32_t box[4] = {(uint32_t)p.gw, (uint32_t)p.BW, (uint32_t)p.BH, (uint32_t)p.BN};
    rc = encode_map(&p.o_map, out_base, 4, dims, ostr, box, p.gw * 2, "output");
    if (rc) return rc;
    if (p.fuse_add || p.fuse_bnbwd) {
      rc = encode_map(&p.r_map, residual, 4, dims, ostr, box, p.gw * 2, "residual");
      if (rc) return rc;
    }
  }
  const long groups = (m_tiles + acc_m - 1) / acc_m;
  if (p.pair) {
    const long total = ((groups + 1) / 2) * p.n_tiles;                 // work items of a pair
    *grid = 2 * (int)(total < RN_NUM_SMS / 2 ? total : RN_NUM_SMS / 2);
  } else {
    const long total = groups * p.n_tiles;
    *grid = (int)(total < RN_NUM_SMS ? total : RN_NUM_SMS);
  }
  return RN_OK;
}

static int create_gemm_plan(const rn_conv_desc* d, int kind, const void* a, const void* b, void* out,
                            rn_convtc_plan* pl) {
  cudaError_t e = set_conv_tc_smem_attr();
  if (e != cudaSuccess) return fail(RN_ERR_CUDA, "convtc: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
  TapSpec taps[TC_MAX_TAPS];
  pl->clear_ptr = nullptr;
  pl->clear_bytes = 0;
  if (kind == 0 || d->stride == 1) {
    const int ncols = kind == 0 ? d->Cout : d->Cin;
    const int oW = kind == 0 ? d->Wo : d->W, oH = kind == 0 ? d->Ho : d->H;
    for (int r = 0; r < d->k; ++r)
      for (int s = 0; s < d->k; ++s) {
        TapSpec& t = taps[r * d->k + s];
        t.b = (int8_t)(r * d->k + s);
        if (kind == 0) forward_tap(r, s, d->pad, d->stride, &t.map, &t.dw, &t.dh);
        else { t.map = 0; t.dw = (int8_t)(d->pad - s); t.dh = (int8_t)(d->pad - r); }
      }
    const uint64_t ostr[3] = {(uint64_t)ncols * 2, (uint64_t)oW * ncols * 2, (uint64_t)oH * oW * ncols * 2};
    pl->nsub = 1;
    return create_sub(d, kind, a, b, out, oW, oH, ostr, taps, d->k * d->k, &pl->p[0], &pl->grid[0], &pl->smem_sub[0]);
  }
  // stride-2 dgrad: dx[2i+ph, 2j+pw] = sum over taps with r = ph+pad, s = pw+pad (mod 2) of
  //   dy[i + (ph+pad-r)/2, j + (pw+pad-s)/2] * w[r,s]  — one dense GEMM per output parity class, written
  //   through a strided view of dx.
  pl->nsub = 0;
  bool some_empty = false;
  for (int ph = 0; ph < 2; ++ph)
    for (int pw = 0; pw < 2; ++pw) {
      const int W2 = (d->W - pw + 1) / 2, H2 = (d->H - ph + 1) / 2;
      if (W2 <= 0 || H2 <= 0) continue;
      int nt = 0;
      for (int r = 0; r < d->k; ++r)
        for (int s = 0; s < d->k; ++s) {
          if (((ph + d->pad - r) & 1) || ((pw + d->pad - s) & 1)) continue;
          TapSpec& t = taps[nt++];
          t.map = 0; t.b = (int8_t)(r * d->k + s);
          t.dh = (int8_t)((ph + d->pad - r) / 2);     // exact: the numerator is even
          t.dw = (int8_t)((pw + d->pad - s) / 2);
        }
      if (nt == 0) { some_empty = true; continue; }
      const uint64_t ostr[3] = {(uint64_t)2 * d->Cin * 2, (uint64_t)2 * d->W * d->Cin * 2, (uint64_t)d->H * d->W * d->Cin * 2};
      const char* base = (const char*)out + ((size_t)ph * d->W + pw) * d->Cin * 2;
      const int i = pl->nsub;
      int rc = create_sub(d, 1, a, b, base, W2, H2, ostr, taps, nt, &pl->p[i], &pl->grid[i], &pl->smem_sub[i]);
      if (rc) return rc;
      pl->nsub++;
    }
  if (some_empty) { pl->clear_ptr = out; pl->clear_bytes = (size_t)d->N * d->H * d->W * d->Cin * 2; }
  return RN_OK;
}

static int create_wgrad_plan(const rn_conv_desc* d, const void* x, const void* dy, void* dw, void* scratch,
                             size_t scratch_bytes, rn_convtc_plan* pl, const ActView* view = nullptr,
                             const TapSpec* taps = nullptr, const BnOperand* xf = nullptr) {
  WgradShape ws;
  if (!wgrad_shape(d, &ws, xf == nullptr)) return fail(RN_ERR_UNSUPPORTED, "convtc wgrad: shape not supported");
  if (ws.scratch > 0 && (scratch == nullptr || scratch_bytes < ws.scratch))
    return fail(RN_ERR_SCRATCH, "convtc wgrad: scratch %zu < %zu bytes", scratch_bytes, ws.scratch);
  TcWgradParams& w = pl->w;
  w.Cin = d->Cin; w.Cout = d->Cout; w.taps = d->k * d->k; w.cin_blocks = ws.cin_blocks; w.RB = ws.RB;
  w.acc = ws.acc; w.row_groups = ws.row_groups; w.block_n = ws.block_n; w.n_tiles = ws.n_tiles;
  w.PW = ws.PW; w.PH = ws.PH; w.PN = ws.PN; w.tiles_w = ws.tiles_w; w.tiles_h = ws.tiles_h; w.tiles_n = ws.tiles_n;
  w.k_tiles = ws.k_tiles; w.splits = ws.splits; w.tiles_per_split = ws.tiles_per_split; w.kp = ws.kp;
  w.stages = ws.stages; w.stage_bytes = ws.stage_bytes; w.tmem_cols = ws.tmem_cols;
  w.bwA = ws.bwA; w.bwB = ws.bwB; w.blocksA = ws.blocksA; w.a_box_bytes = ws.a_box_bytes; w.b_box_bytes = ws.b_box_bytes;
  w.runA = ws.runA;
  w.pair = ws.pair;
  if (xf) {
    if (d->k != 1 || ws.bwA != 64 || view != nullptr)
      return fail(RN_ERR_UNSUPPORTED, "convtc wgrad: the BN operand transform needs a 1x1 conv over a multiple of 64 channels");
    w.xf = 1;
    w.xf_gamma = xf->gamma; w.xf_beta = xf->beta; w.xf_mean = xf->mean; w.xf_rstd = xf->rstd; w.xf_leak = xf->leak;
    // the per-channel constants live behind the barriers: give up pipeline stages if they do not fit
    while ((size_t)w.stages * w.stage_bytes + 16384 + 2048 + (size_t)d->Cin * 8 > (size_t)227 * 1024 && w.stages > 2) --w.stages;
    if ((size_t)w.stages * w.stage_bytes + 16384 + 2048 + (size_t)d->Cin * 8 > (size_t)227 * 1024)
      return fail(RN_ERR_UNSUPPORTED, "convtc wgrad: the BN operand transform does not fit shared memory for this shape");
  }
  w.partial = ws.splits > 1 ? (float*)scratch : (float*)dw;
  for (int r = 0; r < d->k; ++r)
    for (int s = 0; s < d->k; ++s)
      forward_tap(r, s, d->pad, d->stride, &w.tap_map[r * d->k + s], &w.tap_dw[r * d->k + s], &w.tap_dh[r * d->k + s]);
  if (taps)
    for (int t = 0; t < d->k * d->k; ++t) { w.tap_map[t] = taps[t].map; w.tap_dw[t] = taps[t].dw; w.tap_dh[t] = taps[t].dh; }
  int rc = view ? encode_view_5d(&w.x_map[0], x, *view, ws.bwA, ws.PW, ws.PH, ws.PN, ws.runA)
                : encode_activation_maps_5d(w.x_map, x, d->Cin, d->W, d->H, d->N, d->stride, ws.bwA, ws.PW, ws.PH, ws.PN, ws.runA);
  if (rc) return rc;
  rc = encode_activation_maps_5d(&w.dy_map, dy, d->Cout, d->Wo, d->Ho, d->N, 1, ws.bwB, ws.PW, ws.PH, ws.PN,
                                 ws.pair ? ws.block_n / ws.bwB / 2 : ws.block_n / ws.bwB);   // pair: each CTA loads half of the tile's columns
  if (rc) return rc;
  const long total = (long)ws.row_groups * ws.n_tiles * ws.splits;
  if (ws.pair) pl->wgrid = 2 * (int)(total < RN_NUM_SMS / 2 ? total : RN_NUM_SMS / 2);
  else pl->wgrid = (int)(total < RN_NUM_SMS ? total : RN_NUM_SMS);
  pl->wsmem = xf ? (size_t)w.stages * w.stage_bytes + 16384 + 2048 + (size_t)d->Cin * 8 : ws.smem;
  pl->dw = (float*)dw;
  pl->dw_elems = (int64_t)d->k * d->k * d->Cin * d->Cout;
  cudaError_t e = cudaFuncSetAttribute(conv_wgrad_tc_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(227 * 1024));
  if (e == cudaSuccess)
    e = cudaFuncSetAttribute(conv_wgrad_tc_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(227 * 1024));
  if (e == cudaSuccess)
    e = cudaFuncSetAttribute(conv_wgrad_tc_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(227 * 1024));
  if (e != cudaSuccess) return fail(RN_ERR_CUDA, "convtc: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
  return RN_OK;
}

extern "C" int rn_convtc_plan_create(const rn_conv_desc* d, int kind, const void* a, const void* b, void* out,
                                     void* scratch, size_t scratch_bytes, rn_convtc_plan** plan) {
  RN_REQUIRE(d && a && b && out && plan, "convtc_plan_create: NULL pointer");
  if (!rn_convtc_supported(d, kind)) return fail(RN_ERR_UNSUPPORTED, "convtc: shape/kind not supported");
  RN_REQUIRE((((uintptr_t)a | (uintptr_t)b | (uintptr_t)out) & 15) == 0, "convtc: pointers must be 16-byte aligned");
  rn_convtc_plan* pl = new rn_convtc_plan();
  memset(pl, 0, sizeof(*pl));
  pl->kind = kind;
  const int rc = kind == 2 ? create_wgrad_plan(d, a, b, out, scratch, scratch_bytes, pl)
                           : create_gemm_plan(d, kind, a, b, out, pl);
  if (rc) { delete pl; return rc; }
  *plan = pl;
  return RN_OK;
}

extern "C" int rn_convtc_fused_supported(const rn_conv_desc* d) {
  // store groups of 64 channels (ImageNet widths) or 16 / 32 (CIFAR / WRN widths)
  if (!rn_convtc_supported(d, 0) || d->Cin < 16) return 0;
  return (choose_block_n(d->Cout) > 0 && d->Cout <= 4096) ? 1 : 0;
}

extern "C" int rn_convtc_stat_rows(const rn_conv_desc* d) {
  (void)d;
  return RN_NUM_SMS;            // the persistent grid never exceeds one CTA per SM
}

extern "C" int rn_convtc_plan_create_fused(const rn_conv_desc* d, const void* x, const void* w_bf16, void* out,
                                           const void* residual, float* stat_partials, int32_t* stat_rows,
                                           rn_convtc_plan** plan) {
  RN_REQUIRE(d && x && w_bf16 && out && plan, "convtc_plan_create_fused: NULL pointer");
  RN_REQUIRE(residual || stat_partials, "convtc_plan_create_fused: nothing to fuse");
  if (!rn_convtc_fused_supported(d)) return fail(RN_ERR_UNSUPPORTED, "convtc: fused epilogue not supported for this shape");
  RN_REQUIRE((((uintptr_t)x | (uintptr_t)w_bf16 | (uintptr_t)out | (uintptr_t)residual) & 15) == 0,
             "convtc: pointers must be 16-byte aligned");
  cudaError_t e = set_conv_tc_smem_attr();
  if (e != cudaSuccess) return fail(RN_ERR_CUDA, "convtc: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
  rn_convtc_plan* pl = new rn_convtc_plan();
  memset(pl, 0, sizeof(*pl));
  pl->kind = 0;
  pl->nsub = 1;
  TapSpec taps[TC_MAX_TAPS];
  for (int r = 0; r < d->k; ++r)
    for (int s = 0; s < d->k; ++s) {
      TapSpec& t = taps[r * d->k + s];
      t.b = (int8_t)(r * d->k + s);
      forward_tap(r, s, d->pad, d->stride, &t.map, &t.dw, &t.dh);
    }
  const uint64_t ostr[3] = {(uint64_t)d->Cout * 2, (uint64_t)d->Wo * d->Cout * 2, (uint64_t)d->Ho * d->Wo * d->Cout * 2};
  const int rc = create_sub(d, 0, x, w_bf16, out, d->Wo, d->Ho, ostr, taps, d->k * d->k, &pl->p[0], &pl->grid[0],
                            &pl->smem_sub[0], nullptr, 0, residual, stat_partials);
  if (rc) { delete pl; return rc; }
  if (stat_rows) *stat_rows = pl->grid[0];
  *plan = pl;
  return RN_OK;
}

extern "C" int rn_convtc_fused_dgrad_supported(const rn_conv_desc* d) {
  if (!rn_convtc_supported(d, 1) || d->stride != 1) return 0;
  return (choose_block_n(d->Cin) > 0 && d->Cin <= 1024) ? 1 : 0;
}

extern "C" int rn_convtc_plan_create_fused_dgrad(const rn_conv_desc* d, const void* dy, const void* w_bf16, void* dx,
                                                 const void* bn_x, const float* gamma, const float* beta,
                                                 const float* mean, const float* rstd, float leak,
                                                 float* sum_partials, int32_t* rows, rn_convtc_plan** plan) {
  RN_REQUIRE(d && dy && w_bf16 && dx && bn_x && gamma && beta && mean && rstd && sum_partials && plan,
             "convtc_plan_create_fused_dgrad: NULL pointer");
  if (!rn_convtc_fused_dgrad_supported(d))
    return fail(RN_ERR_UNSUPPORTED, "convtc: fused BN-backward dgrad not supported for this shape");
  RN_REQUIRE((((uintptr_t)dy | (uintptr_t)w_bf16 | (uintptr_t)dx | (uintptr_t)bn_x) & 15) == 0,
             "convtc: pointers must be 16-byte aligned");
  cudaError_t e = set_conv_tc_smem_attr();
  if (e != cudaSuccess) return fail(RN_ERR_CUDA, "convtc: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
  rn_convtc_plan* pl = new rn_convtc_plan();
  memset(pl, 0, sizeof(*pl));
  pl->kind = 1;
  pl->nsub = 1;
  TapSpec taps[TC_MAX_TAPS];
  for (int r = 0; r < d->k; ++r)
    for (int s = 0; s < d->k; ++s) {
      TapSpec& t = taps[r * d->k + s];
      t.b = (int8_t)(r * d->k + s);
      t.map = 0; t.dw = (int8_t)(d->pad - s); t.dh = (int8_t)(d->pad - r);
    }
  const uint64_t ostr[3] = {(uint64_t)d->Cin * 2, (uint64_t)d->W * d->Cin * 2, (uint64_t)d->H * d->W * d->Cin * 2};
  const BnBwdOpts o = {bn_x, gamma, beta, mean, rstd, leak};
  const int rc = create_sub(d, 1, dy, w_bf16, dx, d->W, d->H, ostr, taps, d->k * d->k, &pl->p[0], &pl->grid[0],
                            &pl->smem_sub[0], nullptr, 0, nullptr, sum_partials, &o);
  if (rc) { delete pl; return rc; }
  if (rows) *rows = pl->grid[0];
  *plan = pl;
  return RN_OK;
}

// ---- convolutions that read the INPUT of a training-mode batch_norm_relu and normalise it on the fly -------------
extern "C" int rn_convtc_bnin_supported(const rn_conv_desc* d, int kind) {
  if (!d || (kind != 0 && kind != 2) || d->k != 1 || d->pad != 0 || d->Cin % 64 != 0) return 0;
  if (!rn_convtc_supported(d, kind)) return 0;
  if (kind == 2) {
    WgradShape w;
    return wgrad_shape(d, &w) && w.bwA == 64 ? 1 : 0;
  }
  return 1;
}

extern "C" int rn_convtc_plan_create_bnin(const rn_conv_desc* d, int kind, const void* bn_x, const float* gamma,
                                          const float* beta, const float* mean, const float* rstd, float leak,
                                          const void* b, void* out, const void* residual, float* stat_partials,
                                          int32_t* stat_rows, void* scratch, size_t scratch_bytes, rn_convtc_plan** plan) {
  RN_REQUIRE(d && bn_x && gamma && beta && mean && rstd && b && out && plan, "convtc_plan_create_bnin: NULL pointer");
  if (!rn_convtc_bnin_supported(d, kind))
    return fail(RN_ERR_UNSUPPORTED, "convtc: the BN operand transform needs a 1x1 convolution over a multiple of 64 channels");
  RN_REQUIRE((((uintptr_t)bn_x | (uintptr_t)b | (uintptr_t)out | (uintptr_t)residual) & 15) == 0,
             "convtc: pointers must be 16-byte aligned");
  RN_REQUIRE(kind == 0 || (residual == nullptr && stat_partials == nullptr), "convtc_plan_create_bnin: fused epilogues belong to fprop");
  if ((residual || stat_partials) && !rn_convtc_fused_supported(d))
    return fail(RN_ERR_UNSUPPORTED, "convtc: fused epilogue not supported for this shape");
  const BnOperand xf = {gamma, beta, mean, rstd, leak};
  rn_convtc_plan* pl = new rn_convtc_plan();
  memset(pl, 0, sizeof(*pl));
  pl->kind = kind;
  int rc;
  if (kind == 2) {
    rc = create_wgrad_plan(d, bn_x, b, out, scratch, scratch_bytes, pl, nullptr, nullptr, &xf);
  } else {
    cudaError_t e = set_conv_tc_smem_attr();
    if (e != cudaSuccess) { delete pl; return fail(RN_ERR_CUDA, "convtc: cudaFuncSetAttribute: %s", cudaGetErrorString(e)); }
    pl->nsub = 1;
    TapSpec tap;
    tap.b = 0;
    forward_tap(0, 0, 0, d->stride, &tap.map, &tap.dw, &tap.dh);
    const uint64_t ostr[3] = {(uint64_t)d->Cout * 2, (uint64_t)d->Wo * d->Cout * 2, (uint64_t)d->Ho * d->Wo * d->Cout * 2};
    rc = create_sub(d, 0, bn_x, b, out, d->Wo, d->Ho, ostr, &tap, 1, &pl->p[0], &pl->grid[0], &pl->smem_sub[0], nullptr, 0,
                    residual, stat_partials, nullptr, &xf);
    if (rc == RN_OK && stat_rows) *stat_rows = pl->grid[0];
  }
  if (rc) { delete pl; return rc; }
  *plan = pl;
  return RN_OK;
}

extern "C" int rn_convtc_run(rn_convtc_plan* plan, float beta, void* stream) {
  RN_REQUIRE(plan, "convtc_run: NULL plan");
  RN_REQUIRE(beta == 0.f || beta == 1.f, "convtc_run: beta must be 0 or 1");
  cudaStream_t st = as_stream(stream);
  if (plan->kind == 2) {
    if (beta != 0.f) return fail(RN_ERR_UNSUPPORTED, "convtc_run: wgrad overwrites its output (beta must be 0)");
    if (plan->w.xf)        launch_pdl(conv_wgrad_tc_kernel<true, false>, plan->wgrid, TC_THREADS + TC_XF_WARPS * 32, plan->wsmem, st, plan->w);
    else if (plan->w.pair) launch_conv(conv_wgrad_tc_kernel<false, true>, plan->wgrid, TC_THREADS, plan->wsmem, st, true, plan->w);
    else                   launch_pdl(conv_wgrad_tc_kernel<false, false>, plan->wgrid, TC_THREADS, plan->wsmem, st, plan->w);
    int rc = check_launch("conv_wgrad_tc_kernel");
    if (rc) return rc;
    if (plan->w.splits > 1) {
      const int64_t n4 = plan->dw_elems / 4;
      const int sp = plan->w.splits;
      const int S = sp <= 4 ? 1 : (sp <= 16 ? 4 : (sp <= 48 ? 16 : 32));
      const int64_t blocks = (n4 + 256 / S - 1) / (256 / S);
      launch_pdl(wgrad_reduce_kernel, (unsigned)blocks, 256, 0, st, (const float*)plan->w.partial, plan->dw, n4, sp, S);
      rc = check_launch("wgrad_reduce_kernel");
    }
    return rc;
  }
  if (beta != 0.f && (plan->p[0].fuse_add || plan->p[0].fuse_stats || plan->p[0].fuse_bnbwd))
    return fail(RN_ERR_UNSUPPORTED, "convtc_run: a plan with a fused epilogue overwrites its output (beta must be 0)");
  if (beta == 0.f && plan->clear_ptr) {
    cudaError_t e = cudaMemsetAsync(plan->clear_ptr, 0, plan->clear_bytes, st);
    if (e != cudaSuccess) return fail(RN_ERR_CUDA, "convtc_run: cudaMemsetAsync: %s", cudaGetErrorString(e));
  }
  for (int i = 0; i < plan->nsub; ++i) {
    TcConvParams p = plan->p[i];
    p.accumulate = beta != 0.f ? 1 : 0;
    const bool fused = p.fuse_add || p.fuse_stats || p.fuse_bnbwd;
    const int threads = (p.fuse_add || p.fuse_bnbwd) ? TC_THREADS_RES : TC_CONV_THREADS;
    launch_conv(conv_tc_variant(fused, p.pair != 0, p.halo != 0, p.xf != 0), plan->grid[i], threads + (p.xf ? TC_XF_WARPS * 32 : 0),
                plan->smem_sub[i], st, p.pair != 0, p);
    int rc = check_launch("conv_tc_kernel");
    if (rc) return rc;
  }
  return RN_OK;
}

// Fold the finalisation of a fused epilogue's partial rows into the convolution kernel itself (see TcConvParams::fin_mode).
// The table is <= 148 rows of C (sum, sum) pairs; one CTA sums it in a few microseconds while C is small, which is
// cheaper than one more dependent launch, so the caller decides per layer (the standalone kernels spread wide tables
// over many CTAs).  `ticket` is one zero-initialised int32 in device memory owned by the caller, private to this plan.
extern "C" int rn_convtc_plan_fold_bn_finalize(rn_convtc_plan* plan, int64_t M, float eps, float decay, int32_t unbiased,
                                               float* mean, float* rstd, float* moving_mean, float* moving_var,
                                               int32_t* ticket) {
  RN_REQUIRE(plan && mean && rstd && ticket && M > 0, "convtc_plan_fold_bn_finalize: bad arguments");
  RN_REQUIRE((moving_mean == nullptr) == (moving_var == nullptr), "convtc_plan_fold_bn_finalize: moving_mean/var must both be set or NULL");
  if (plan->kind != 0 || plan->nsub != 1 || !plan->p[0].fuse_stats)
    return fail(RN_ERR_UNSUPPORTED, "convtc_plan_fold_bn_finalize: the plan has no fused statistics epilogue");
  TcConvParams& p = plan->p[0];
  p.fin_mode = 1;
  p.fin_ticket = ticket;
  p.fin_o0 = mean; p.fin_o1 = rstd; p.fin_mm = moving_mean; p.fin_mv = moving_var;
  p.fin_M = (double)M; p.fin_eps = eps; p.fin_decay = decay; p.fin_unbiased = unbiased;
  return RN_OK;
}

extern "C" int rn_convtc_plan_fold_bn_bwd_finalize(rn_convtc_plan* plan, float* dgamma, float* dbeta, int32_t* ticket) {
  RN_REQUIRE(plan && dgamma && dbeta && ticket, "convtc_plan_fold_bn_bwd_finalize: bad arguments");
  if (plan->kind != 1 || plan->nsub != 1 || !plan->p[0].fuse_bnbwd)
    return fail(RN_ERR_UNSUPPORTED, "convtc_plan_fold_bn_bwd_finalize: the plan has no fused BN-backward epilogue");
  TcConvParams& p = plan->p[0];
  p.fin_mode = 2;
  p.fin_ticket = ticket;
  p.fin_o0 = dgamma; p.fin_o1 = dbeta;
  return RN_OK;
}

extern "C" int rn_convtc_plan_describe(const rn_convtc_plan* plan, char* buf, size_t n) {
  RN_REQUIRE(plan && buf && n > 0, "convtc_plan_describe: bad arguments");
  if (plan->kind == 2) {
    const TcWgradParams& w = plan->w;
    snprintf(buf, n, "wgrad block_n=%d acc=%d kp=%d stages=%d splits=%d grid=%d pair=%d bn_operand=%d", w.block_n, w.acc, w.kp, w.stages,
             w.splits, plan->wgrid, w.pair, w.xf);
    return RN_OK;
  }
  const TcConvParams& p = plan->p[0];
  snprintf(buf, n, "%s subs=%d block_n=%d acc_m=%d box=%dx%dx%d kc=%d taps=%d stages=%d tstages=%d pair=%d grid=%d fused=%d%d%d halo=%d fold=%d",
           plan->kind == 0 ? "fprop" : "dgrad", plan->nsub, p.block_n, p.acc_m, p.BW, p.BH, p.BN, p.kc, p.taps, p.stages,
           p.tstages, p.pair, plan->grid[0], p.fuse_add, p.fuse_stats, p.fuse_bnbwd, p.halo, p.fin_mode);
  if (p.xf) strncat(buf, " bn_operand=1", n - strlen(buf) - 1);
  return RN_OK;
}

extern "C" int rn_convtc_plan_destroy(rn_convtc_plan* plan) {
  delete plan;
  return RN_OK;
}

extern "C" int rn_weights_to_bf16(const float* w_hwio, int k, int Cin, int Cout, void* w_fprop, void* w_dgrad,
                                  void* stream) {
  RN_REQUIRE(w_hwio && k > 0 && Cin > 0 && Cout > 0, "weights_to_bf16: bad arguments");
  const int64_t n = (int64_t)k * k * Cin * Cout;
  weights_to_bf16_kernel<<<grid_for(n, 256), 256, 0, as_stream(stream)>>>(w_hwio, k * k, Cin, Cout,
                                                                         (__nv_bfloat16*)w_fprop, (__nv_bfloat16*)w_dgrad);
  return check_launch("weights_to_bf16_kernel");
}

// ------------------------------------------------------------------------------------------
// Stem (first) convolution: 3 input channels cannot feed TMA rows (16-byte granularity) or UMMA K
// steps, so the image is unfolded once per step into bf16 rows of Kpad = roundup(k*k*Cin, 64)
// columns ordered (r, s, ci) — the HWIO flattening — and the GEMM kernels above run it as a 1x1
// convolution with Kpad input channels (forward and wgrad; the image needs no dgrad).
// ------------------------------------------------------------------------------------------
namespace rn {
constexpr int IM2COL_MAX_K = 512;
template <typename T>
__global__ void __launch_bounds__(256)
im2col_kernel(const T* __restrict__ x, __nv_bfloat16* __restrict__ cols, int N, int H, int W, int Cin, int Ho, int Wo,
              int k, int stride, int pad, int Kpad) {
  // per-column lookup: input offset relative to the window origin, and the (r, s) of the tap
  __shared__ int s_off[IM2COL_MAX_K];
  __shared__ int8_t s_r[IM2COL_MAX_K], s_s[IM2COL_MAX_K];
  const int K = k * k * Cin;
  for (int kk = threadIdx.x; kk < Kpad; kk += blockDim.x) {
    if (kk < K) {
      const int tap = kk / Cin, c = kk - tap * Cin;
      const int r = tap / k, sft = tap - r * k;
      s_off[kk] = (r * W + sft) * Cin + c;
      s_r[kk] = (int8_t)r;
      s_s[kk] = (int8_t)sft;
    } else {
      s_off[kk] = 0;
      s_r[kk] = -128;          // padding column
      s_s[kk] = 0;
    }
  }
  __syncthreads();
  const int chunks = Kpad / 8;
  const int64_t total = (int64_t)N * Ho * Wo * chunks;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int ch = (int)(i % chunks);
    int64_t p = i / chunks;
    const int wo = (int)(p % Wo);
    p /= Wo;
    const int ho = (int)(p % Ho);
    const int n = (int)(p / Ho);
    const int h0 = ho * stride - pad, w0 = wo * stride - pad;
    const int64_t base = (((int64_t)n * H + h0) * W + w0) * Cin;
    float f[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int kk = ch * 8 + j;
      const int r = s_r[kk];
      const int h = h0 + r, w = w0 + s_s[kk];
      float v = 0.f;
      if (r >= 0 && h >= 0 && h < H && w >= 0 && w < W) v = to_f<T>(x[base + s_off[kk]]);
      f[j] = v;
    }
    Vec<__nv_bfloat16> o;
    o.pack(f);
    o.store(cols + i * 8);
  }
}

// fp32 HWIO [rows = k*k*Cin][Cout] -> bf16 [Cout][Kpad] (K-major B operand, zero padded)
__global__ void stem_weights_kernel(const float* __restrict__ w, int rows, int Cout, int Kpad,
                                    __nv_bfloat16* __restrict__ wf) {
  const int64_t total = (int64_t)Cout * Kpad;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int kk = (int)(i % Kpad);
    const int co = (int)(i / Kpad);
    wf[i] = __float2bfloat16_rn(kk < rows ? w[(int64_t)kk * Cout + co] : 0.f);
  }
}
}  // namespace rn

extern "C" int rn_im2col(const void* x, void* cols, int32_t N, int32_t H, int32_t W, int32_t Cin, int32_t Ho, int32_t Wo,
                         int32_t k, int32_t stride, int32_t pad, int32_t Kpad, int32_t src_dtype, void* stream) {
  RN_REQUIRE(x && cols && N > 0 && H > 0 && W > 0 && Cin > 0 && Ho > 0 && Wo > 0 && k > 0 && stride > 0 && pad >= 0,
             "im2col: bad arguments");
  RN_REQUIRE(Kpad % 8 == 0 && Kpad >= k * k * Cin && Kpad <= IM2COL_MAX_K,
             "im2col: Kpad=%d must be a multiple of 8, >= k*k*Cin and <= 512", Kpad);
  const int64_t total = (int64_t)N * Ho * Wo * (Kpad / 8);
  const int g = grid_for(total, 256, 16);
  if (src_dtype == RN_BF16)
    im2col_kernel<__nv_bfloat16><<<g, 256, 0, as_stream(stream)>>>((const __nv_bfloat16*)x, (__nv_bfloat16*)cols, N, H,
                                                                  W, Cin, Ho, Wo, k, stride, pad, Kpad);
  else if (src_dtype == RN_F32)
    im2col_kernel<float><<<g, 256, 0, as_stream(stream)>>>((const float*)x, (__nv_bfloat16*)cols, N, H, W, Cin, Ho, Wo,
                                                          k, stride, pad, Kpad);
  else
    return fail(RN_ERR_INVALID, "im2col: bad src_dtype %d", src_dtype);
  return check_launch("im2col_kernel");
}

extern "C" int rn_stem_weights(const float* w_hwio, int32_t rows, int32_t Cout, int32_t Kpad, void* w_fprop,
                               void* stream) {
  RN_REQUIRE(w_hwio && w_fprop && rows > 0 && Cout > 0 && Kpad >= rows, "stem_weights: bad arguments");
  stem_weights_kernel<<<grid_for((int64_t)Cout * Kpad, 256), 256, 0, as_stream(stream)>>>(w_hwio, rows, Cout, Kpad,
                                                                                         (__nv_bfloat16*)w_fprop);
  return check_launch("stem_weights_kernel");
}

extern "C" size_t rn_relayout_job_bytes(void) { return sizeof(rn::RelayoutJob); }

extern "C" int rn_weights_to_bf16_batched(const void* jobs_device, int32_t njobs, void* stream) {
  RN_REQUIRE(jobs_device && njobs > 0 && njobs <= 65535, "weights_to_bf16_batched: bad arguments");
  weights_to_bf16_batched_kernel<<<dim3(16, njobs), 256, 0, as_stream(stream)>>>((const rn::RelayoutJob*)jobs_device);
  return check_launch("weights_to_bf16_batched_kernel");
}

// ------------------------------------------------------------------------------------------
// Stride-2 stem (ImageNet 7x7/2 over the 3-channel image, resnet_model_official.py:310) without an
// unfolded matrix.  The padded image is packed 2x2 -> channels ("space to depth"):
//     U[n, h2, w2, (dy*2 + dx)*Cin + c] = x[n, 2*h2 + dy - pad, 2*w2 + dx - pad, c]      (16 bf16 = 32 bytes per pixel)
// so the k x k / 2 convolution becomes a T x T / 1 one, T = ceil(k/2), over U.  A TMA view of U whose
// W stride (32 B) is smaller than its 128-byte rows hands the kernels, per output pixel and kernel row
// kh2, the 4 horizontally neighbouring packed pixels as ONE 64-channel row: the layer runs as T
// vertical taps over 64 channels (fprop K = T*64; wgrad 2 accumulators), reading U (tens of MB,
// L2-resident) instead of a [pixels][k*k*Cin] matrix.
// ------------------------------------------------------------------------------------------
namespace rn {
struct S2dGeom { int T, H2, W2, Cp; };
static bool s2d_geometry(const rn_conv_desc* d, S2dGeom* g) {
  if (!d || d->stride != 2 || d->Cin < 1 || d->Cin > 4 || d->k < 2 || d->k > 8 || d->pad < 0) return false;
  g->T = (d->k + 1) / 2;
  g->H2 = d->Ho + g->T - 1;
  g->W2 = d->Wo + 3;            // every row reads a 4-pixel window
  g->Cp = 16;
  return true;
}

template <typename T, int CIN>
__global__ void __launch_bounds__(256)
s2d_pack_kernel(const T* __restrict__ x, __nv_bfloat16* __restrict__ u, int N, int H, int W, int H2, int W2, int pad) {
  const int64_t total = (int64_t)N * H2 * W2;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int w2 = (int)(i % W2);
    int64_t q = i / W2;
    const int h2 = (int)(q % H2);
    const int n = (int)(q / H2);
    float lo[8], hi[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) { lo[j] = 0.f; hi[j] = 0.f; }
#pragma unroll
    for (int dy = 0; dy < 2; ++dy) {
      const int h = 2 * h2 + dy - pad;
#pragma unroll
      for (int dx = 0; dx < 2; ++dx) {
        const int w = 2 * w2 + dx - pad;
        if (h >= 0 && h < H && w >= 0 && w < W) {
          const T* src = x + (((int64_t)n * H + h) * W + w) * CIN;
#pragma unroll
          for (int c = 0; c < CIN; ++c) {
            const int j = (dy * 2 + dx) * CIN + c;       // compile-time after unrolling
            const float v = to_f<T>(src[c]);
            if (j < 8) lo[j] = v; else hi[j - 8] = v;
          }
        }
      }
    }
    Vec<__nv_bfloat16> o;
    o.pack(lo);
    o.store(u + i * 16);
    o.pack(hi);
    o.store(u + i * 16 + 8);
  }
}

template <typename T>
static void s2d_pack_launch(const T* x, __nv_bfloat16* u, const rn_conv_desc* d, const S2dGeom& g, cudaStream_t st) {
  const int64_t total = (int64_t)d->N * g.H2 * g.W2;
  const int grid = grid_for(total, 256, 16);
  switch (d->Cin) {
    case 1: s2d_pack_kernel<T, 1><<<grid, 256, 0, st>>>(x, u, d->N, d->H, d->W, g.H2, g.W2, d->pad); break;
    case 2: s2d_pack_kernel<T, 2><<<grid, 256, 0, st>>>(x, u, d->N, d->H, d->W, g.H2, g.W2, d->pad); break;
    case 3: s2d_pack_kernel<T, 3><<<grid, 256, 0, st>>>(x, u, d->N, d->H, d->W, g.H2, g.W2, d->pad); break;
    default: s2d_pack_kernel<T, 4><<<grid, 256, 0, st>>>(x, u, d->N, d->H, d->W, g.H2, g.W2, d->pad); break;
  }
}

// bf16 HWIO [k][k][Cin][Cout] -> bf16 [T][64 = kw2*16 + (dy*2+dx)*Cin + c][Cout]  (zero where kh or kw >= k)
__global__ void s2d_weights_kernel(const __nv_bfloat16* __restrict__ w, __nv_bfloat16* __restrict__ w2, int k, int Cin,
                                   int Cout, int T) {
  const int64_t total = (int64_t)T * 64 * Cout;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int co = (int)(i % Cout);
    const int r = (int)((i / Cout) % 64);
    const int kh2 = (int)(i / ((int64_t)Cout * 64));
    const int kw2 = r / 16, c16 = r % 16;
    const int q = c16 / Cin, c = c16 % Cin;
    const int kh = 2 * kh2 + (q >> 1), kw = 2 * kw2 + (q & 1);
    __nv_bfloat16 v = __float2bfloat16_rn(0.f);
    if (q < 4 && kh < k && kw < k) v = w[(((int64_t)kh * k + kw) * Cin + c) * Cout + co];
    w2[i] = v;
  }
}

// fp32 [T][64][Cout] packed gradient -> fp32 HWIO [k][k][Cin][Cout]
__global__ void s2d_wgrad_unpack_kernel(const float* __restrict__ dw2, float* __restrict__ dw, int k, int Cin, int Cout) {
  const int64_t total = (int64_t)k * k * Cin * Cout;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int co = (int)(i % Cout);
    int64_t q = i / Cout;
    const int c = (int)(q % Cin);
    q /= Cin;
    const int kw = (int)(q % k), kh = (int)(q / k);
    const int r = (kw >> 1) * 16 + ((kh & 1) * 2 + (kw & 1)) * Cin + c;
    dw[i] = dw2[((int64_t)(kh >> 1) * 64 + r) * Cout + co];
  }
}

// the T-tap, 64-channel stand-in geometry handed to the GEMM plan builders
static void s2d_fake_desc(const rn_conv_desc* d, const S2dGeom& g, rn_conv_desc* f, ActView* v, TapSpec* taps) {
  *f = *d;
  f->H = g.H2; f->W = d->Wo; f->Cin = 64; f->stride = 1; f->pad = 0; f->k = 2;   // k*k = 4 >= T tap slots
  v->C = 64; v->W = d->Wo; v->H = g.H2; v->N = d->N;
  v->sw = (uint64_t)g.Cp * 2;
  v->sh = (uint64_t)g.W2 * g.Cp * 2;
  v->sn = (uint64_t)g.H2 * g.W2 * g.Cp * 2;
  for (int t = 0; t < 4; ++t) { taps[t].map = 0; taps[t].dw = 0; taps[t].dh = (int8_t)t; taps[t].b = (int8_t)t; }
}
}  // namespace rn

extern "C" int rn_stem_s2d_supported(const rn_conv_desc* d) {
  S2dGeom g;
  if (!s2d_geometry(d, &g) || d->dtype != RN_BF16) return 0;
  if (g.T != 4) return 0;                        // the plan builders below carry exactly four tap slots (k = 7 or 8)
  if (d->Cout % 8 || choose_block_n(d->Cout) == 0) return 0;
  return 1;
}

extern "C" int rn_stem_s2d_geometry(const rn_conv_desc* d, int32_t* H2, int32_t* W2, int32_t* taps) {
  S2dGeom g;
  if (!s2d_geometry(d, &g)) return fail(RN_ERR_UNSUPPORTED, "stem_s2d: needs stride 2, Cin <= 4, k <= 8");
  if (H2) *H2 = g.H2;
  if (W2) *W2 = g.W2;
  if (taps) *taps = g.T;
  return RN_OK;
}

extern "C" int rn_stem_s2d_pack(const rn_conv_desc* d, const void* x, int32_t src_dtype, void* u, void* stream) {
  S2dGeom g;
  RN_REQUIRE(x && u, "stem_s2d_pack: NULL pointer");
  if (!s2d_geometry(d, &g)) return fail(RN_ERR_UNSUPPORTED, "stem_s2d: needs stride 2, Cin <= 4, k <= 8");
  if (src_dtype == RN_BF16) s2d_pack_launch((const __nv_bfloat16*)x, (__nv_bfloat16*)u, d, g, as_stream(stream));
  else if (src_dtype == RN_F32) s2d_pack_launch((const float*)x, (__nv_bfloat16*)u, d, g, as_stream(stream));
  else return fail(RN_ERR_INVALID, "stem_s2d_pack: bad src_dtype %d", src_dtype);
  return check_launch("s2d_pack_kernel");
}

extern "C" int rn_stem_s2d_weights(const rn_conv_desc* d, const void* w_hwio_bf16, void* w2, void* stream) {
  S2dGeom g;
  RN_REQUIRE(w_hwio_bf16 && w2, "stem_s2d_weights: NULL pointer");
  if (!s2d_geometry(d, &g)) return fail(RN_ERR_UNSUPPORTED, "stem_s2d: needs stride 2, Cin <= 4, k <= 8");
  const int64_t total = (int64_t)g.T * 64 * d->Cout;
  s2d_weights_kernel<<<grid_for(total, 256), 256, 0, as_stream(stream)>>>((const __nv_bfloat16*)w_hwio_bf16,
                                                                         (__nv_bfloat16*)w2, d->k, d->Cin, d->Cout, g.T);
  return check_launch("s2d_weights_kernel");
}

extern "C" int rn_stem_s2d_wgrad_unpack(const rn_conv_desc* d, const float* dw2, float* dw_hwio, void* stream) {
  S2dGeom g;
  RN_REQUIRE(dw2 && dw_hwio, "stem_s2d_wgrad_unpack: NULL pointer");
  if (!s2d_geometry(d, &g)) return fail(RN_ERR_UNSUPPORTED, "stem_s2d: needs stride 2, Cin <= 4, k <= 8");
  const int64_t total = (int64_t)d->k * d->k * d->Cin * d->Cout;
  s2d_wgrad_unpack_kernel<<<grid_for(total, 256), 256, 0, as_stream(stream)>>>(dw2, dw_hwio, d->k, d->Cin, d->Cout);
  return check_launch("s2d_wgrad_unpack_kernel");
}

extern "C" size_t rn_stem_s2d_scratch_bytes(const rn_conv_desc* d, int kind) {
  S2dGeom g;
  if (!rn_stem_s2d_supported(d) || kind != 2 || !s2d_geometry(d, &g)) return 0;
  rn_conv_desc f; ActView v; TapSpec taps[4];
  s2d_fake_desc(d, g, &f, &v, taps);
  WgradShape w;
  return wgrad_shape(&f, &w) ? w.scratch : 0;
}

extern "C" int rn_stem_s2d_plan_create(const rn_conv_desc* d, int kind, const void* u, const void* b, void* out,
                                       void* scratch, size_t scratch_bytes, rn_convtc_plan** plan) {
  RN_REQUIRE(d && u && b && out && plan, "stem_s2d_plan_create: NULL pointer");
  RN_REQUIRE(kind == 0 || kind == 2, "stem_s2d_plan_create: kind must be 0 (fprop) or 2 (wgrad)");
  if (!rn_stem_s2d_supported(d)) return fail(RN_ERR_UNSUPPORTED, "stem_s2d: shape not supported");
  RN_REQUIRE((((uintptr_t)u | (uintptr_t)b | (uintptr_t)out) & 15) == 0, "stem_s2d: pointers must be 16-byte aligned");
  S2dGeom g;
  s2d_geometry(d, &g);
  rn_conv_desc f; ActView v; TapSpec taps[4];
  s2d_fake_desc(d, g, &f, &v, taps);
  rn_convtc_plan* pl = new rn_convtc_plan();
  memset(pl, 0, sizeof(*pl));
  pl->kind = kind;
  int rc;
  if (kind == 0) {
    cudaError_t e = set_conv_tc_smem_attr();
    if (e != cudaSuccess) { delete pl; return fail(RN_ERR_CUDA, "stem_s2d: cudaFuncSetAttribute: %s", cudaGetErrorString(e)); }
    const uint64_t ostr[3] = {(uint64_t)d->Cout * 2, (uint64_t)d->Wo * d->Cout * 2, (uint64_t)d->Ho * d->Wo * d->Cout * 2};
    pl->nsub = 1;
    rc = create_sub(&f, 0, u, b, out, d->Wo, d->Ho, ostr, taps, g.T, &pl->p[0], &pl->grid[0], &pl->smem_sub[0], &v, g.T);
  } else {
    rc = create_wgrad_plan(&f, u, b, out, scratch, scratch_bytes, pl, &v, taps);
  }
  if (rc) { delete pl; return rc; }
  *plan = pl;
  return RN_OK;
}
