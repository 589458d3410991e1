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
#include <vector>

namespace rn {
using namespace tc;

constexpr int TC_THREADS = 192;
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
  int8_t tap_map[TC_MAX_TAPS], tap_dw[TC_MAX_TAPS], tap_dh[TC_MAX_TAPS], tap_b[TC_MAX_TAPS];
};

__global__ void __launch_bounds__(TC_THREADS, 1) conv_tc_kernel(const __grid_constant__ TcConvParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base_u32 = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem = smem_raw + (base_u32 - smem_u32(smem_raw));
  const uint32_t stage_bytes = p.a_bytes + p.b_bytes;
  const int row_bytes_out = p.gw * 2;
  uint8_t* epi_buf = smem + (size_t)p.stages * stage_bytes;                  // 2 x 16 KB
  uint64_t* bars = reinterpret_cast<uint64_t*>(epi_buf + 2 * 16384);
  uint64_t* full = bars;                  // [stages]
  uint64_t* empty = bars + p.stages;      // [stages]
  uint64_t* tfull = bars + 2 * p.stages;  // [2]
  uint64_t* tempty = tfull + 2;           // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);

  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  if (warp == 0 && lane == 0) {
    for (int i = 0; i < p.stages; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&tfull[i], 1); mbar_init(&tempty[i], 4); }
    fence_barrier_init();
    tma_prefetch_desc(&p.a_map[0]);
    tma_prefetch_desc(&p.b_map);
    tma_prefetch_desc(&p.o_map);
  }
  if (warp == 1) {
    tmem_alloc(tmem_slot, p.tmem_cols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  const int m_tiles = p.tiles_w * p.tiles_h * p.tiles_n;
  const int total = m_tiles * p.n_tiles;
  const int k_iters = p.taps * p.k_chunks;

  if (warp == 0) {
    if (lane == 0) {
      // ===================== TMA producer =====================
      int stage = 0;
      uint32_t phase = 0;
      for (int t = blockIdx.x; t < total; t += gridDim.x) {
        const int nt = t % p.n_tiles;
        int mt = t / p.n_tiles;
        const int w0 = (mt % p.tiles_w) * p.BW;
        mt /= p.tiles_w;
        const int h0 = (mt % p.tiles_h) * p.BH;
        const int n0 = (mt / p.tiles_h) * p.BN;
        for (int tap = 0; tap < p.taps; ++tap) {
          const CUtensorMap* am = &p.a_map[p.tap_map[tap]];
          const int cw = w0 + p.tap_dw[tap], ch = h0 + p.tap_dh[tap];
          const int tb = p.tap_b[tap];
          for (int kk = 0; kk < p.k_chunks; ++kk) {
            mbar_wait(&empty[stage], phase ^ 1);
            uint8_t* sa = smem + (size_t)stage * stage_bytes;
            mbar_expect_tx(&full[stage], p.tx_bytes);
            tma_load_4d(sa, am, &full[stage], kk * p.kc, cw, ch, n0);
            tma_load_3d(sa + p.a_bytes, &p.b_map, &full[stage], kk * p.kc, nt * p.block_n, tb);
            if (++stage == p.stages) { stage = 0; phase ^= 1; }
          }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      // ===================== MMA issuer =====================
      const uint32_t idesc = make_idesc_bf16(TC_BLOCK_M, p.block_n, 0, 0);
      const int row_bytes = p.kc * 2;
      const uint32_t lt = swizzle_layout_type(row_bytes);
      const uint32_t sbo = 8u * row_bytes;
      const int ksteps = p.kc / 16;
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int t = blockIdx.x; t < total; t += gridDim.x) {
        mbar_wait(&tempty[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)(acc * p.block_n);
        for (int it = 0; it < k_iters; ++it) {
          mbar_wait(&full[stage], phase);
          tc_fence_after();
          const uint32_t a_addr = base_u32 + (uint32_t)stage * stage_bytes;
          const uint32_t b_addr = a_addr + p.a_bytes;
          for (int k = 0; k < ksteps; ++k) {
            const uint64_t da = make_smem_desc(a_addr + k * 32, 16, sbo, lt);
            const uint64_t db = make_smem_desc(b_addr + k * 32, 16, sbo, lt);
            umma_bf16(d_tmem, da, db, idesc, (it > 0 || k > 0) ? 1u : 0u);
          }
          umma_commit(&empty[stage]);      // frees the smem slot once these MMAs have read it
          if (++stage == p.stages) { stage = 0; phase ^= 1; }
        }
        umma_commit(&tfull[acc]);          // accumulator complete
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
    }
  } else {
    // ===================== epilogue (warps 2..5) =====================
    const int quad = warp & 3;                      // TMEM lane quadrant this warp may access
    const int row = quad * 32 + lane;
    const bool issuer = (threadIdx.x == 64);
    int acc = 0;
    uint32_t acc_phase = 0;
    int buf = 0;
    const int groups = p.block_n / p.gw;
    for (int t = blockIdx.x; t < total; t += gridDim.x) {
      const int nt = t % p.n_tiles;
      int mt = t / p.n_tiles;
      const int w0 = (mt % p.tiles_w) * p.BW;
      mt /= p.tiles_w;
      const int h0 = (mt % p.tiles_h) * p.BH;
      const int n0 = (mt / p.tiles_h) * p.BN;
      mbar_wait(&tfull[acc], acc_phase);
      tc_fence_after();
      for (int g = 0; g < groups; ++g) {
        // the TMA store that last read this staging buffer must have finished reading it
        if (issuer) tma_store_wait_read<1>();
        asm volatile("bar.sync 1, 128;" ::: "memory");
        uint8_t* sb = epi_buf + buf * 16384;
        const uint32_t taddr = tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(acc * p.block_n + g * p.gw);
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
          *reinterpret_cast<uint4*>(sb + swizzle_offset(off0, row_bytes_out)) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
          *reinterpret_cast<uint4*>(sb + swizzle_offset(off0 + 16, row_bytes_out)) = make_uint4(pk[4], pk[5], pk[6], pk[7]);
        }
        fence_proxy_async();
        asm volatile("bar.sync 1, 128;" ::: "memory");
        if (issuer) {
          tma_store_4d(&p.o_map, sb, nt * p.block_n + g * p.gw, w0, h0, n0);
          tma_store_commit();
        }
        buf ^= 1;
      }
      // all TMEM reads of this accumulator are complete (tcgen05.wait::ld above)
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty[acc]);
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
    if (issuer) tma_store_wait<0>();
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, p.tmem_cols);
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

// ------------------------------------------------------------------------------------------
// host side: tensor maps and plans
// ------------------------------------------------------------------------------------------
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

static int choose_block_n(int ncols) {
  for (int b = 256; b >= 16; b -= 16)
    if (ncols % b == 0) return b;
  return 0;
}

}  // namespace rn

using namespace rn;

struct rn_convtc_plan {
  int kind;
  TcConvParams p;
  int grid;
  size_t smem;
};

extern "C" int rn_convtc_supported(const rn_conv_desc* d, int kind) {
  if (!d || d->dtype != RN_BF16) return 0;
  if (kind != 0 && kind != 1) return 0;
  if (d->stride != 1) return 0;
  if (d->k * d->k > TC_MAX_TAPS) return 0;
  if (d->Cin % 8 || d->Cout % 8) return 0;
  const int cred = kind == 0 ? d->Cin : d->Cout;
  const int ncols = kind == 0 ? d->Cout : d->Cin;
  if (cred < 16 || choose_block_n(ncols) == 0) return 0;
  if (d->pad > 8 || d->k > 2 * d->pad + 1 + 8) return 0;
  return 1;
}

extern "C" size_t rn_convtc_scratch_bytes(const rn_conv_desc*, int) { return 0; }

extern "C" int rn_convtc_plan_create(const rn_conv_desc* d, int kind, const void* a, const void* b, void* out,
                                     void* scratch, size_t scratch_bytes, rn_convtc_plan** plan) {
  (void)scratch; (void)scratch_bytes;
  RN_REQUIRE(d && a && b && out && plan, "convtc_plan_create: NULL pointer");
  if (!rn_convtc_supported(d, kind)) return fail(RN_ERR_UNSUPPORTED, "convtc: shape/kind not supported");
  RN_REQUIRE((((uintptr_t)a | (uintptr_t)b | (uintptr_t)out) & 15) == 0, "convtc: pointers must be 16-byte aligned");
  rn_convtc_plan* pl = new rn_convtc_plan();
  pl->kind = kind;
  TcConvParams& p = pl->p;
  memset(&p, 0, sizeof(p));
  // GEMM view: rows = pixels of the OUTPUT tensor of this GEMM, reduction over (tap, cred), columns = ncols
  const int cred = kind == 0 ? d->Cin : d->Cout;
  const int ncols = kind == 0 ? d->Cout : d->Cin;
  const int aW = kind == 0 ? d->W : d->Wo, aH = kind == 0 ? d->H : d->Ho;   // activation tensor that is gathered
  const int oW = kind == 0 ? d->Wo : d->W, oH = kind == 0 ? d->Ho : d->H;   // tensor that is written
  p.kc = cred >= 64 ? 64 : (cred >= 32 ? 32 : 16);
  p.k_chunks = (cred + p.kc - 1) / p.kc;
  p.block_n = choose_block_n(ncols);
  p.n_tiles = ncols / p.block_n;
  p.gw = p.block_n % 64 == 0 ? 64 : (p.block_n % 32 == 0 ? 32 : 16);
  choose_pixel_box(oW, oH, d->N, &p.BW, &p.BH, &p.BN);
  p.tiles_w = (oW + p.BW - 1) / p.BW;
  p.tiles_h = (oH + p.BH - 1) / p.BH;
  p.tiles_n = (d->N + p.BN - 1) / p.BN;
  p.taps = d->k * d->k;
  for (int r = 0; r < d->k; ++r)
    for (int s = 0; s < d->k; ++s) {
      const int t = r * d->k + s;
      p.tap_map[t] = 0;
      p.tap_b[t] = (int8_t)t;
      if (kind == 0) { p.tap_dw[t] = (int8_t)(s - d->pad); p.tap_dh[t] = (int8_t)(r - d->pad); }
      else           { p.tap_dw[t] = (int8_t)(d->pad - s); p.tap_dh[t] = (int8_t)(d->pad - r); }
    }
  const int row_bytes = p.kc * 2;
  const uint32_t box_rows = (uint32_t)(p.BW * p.BH * p.BN);
  p.a_bytes = ((uint32_t)TC_BLOCK_M * row_bytes + 1023u) & ~1023u;
  p.b_bytes = ((uint32_t)p.block_n * row_bytes + 1023u) & ~1023u;
  p.tx_bytes = box_rows * row_bytes + (uint32_t)p.block_n * row_bytes;
  uint32_t cols = 32;
  while (cols < 2u * p.block_n) cols <<= 1;
  p.tmem_cols = cols;
  const size_t fixed = 2 * 16384 + 1024 + 1024;     // epilogue staging, barriers, alignment slack
  const size_t budget = 227 * 1024;
  int stages = (int)((budget - fixed) / (p.a_bytes + p.b_bytes));
  if (stages > 8) stages = 8;
  if (stages < 2) { delete pl; return fail(RN_ERR_UNSUPPORTED, "convtc: tile does not fit shared memory"); }
  p.stages = stages;
  pl->smem = fixed + (size_t)stages * (p.a_bytes + p.b_bytes);
  int rc;
  {  // activation: dims (C, W, H, N)
    const uint64_t dims[4] = {(uint64_t)cred, (uint64_t)aW, (uint64_t)aH, (uint64_t)d->N};
    const uint64_t str[3] = {(uint64_t)cred * 2, (uint64_t)aW * cred * 2, (uint64_t)aH * aW * cred * 2};
    const uint32_t box[4] = {(uint32_t)p.kc, (uint32_t)p.BW, (uint32_t)p.BH, (uint32_t)p.BN};
    rc = encode_map(&p.a_map[0], a, 4, dims, str, box, row_bytes, "activation");
    if (rc) { delete pl; return rc; }
  }
  {  // weights: dims (cred, ncols, taps)
    const uint64_t dims[3] = {(uint64_t)cred, (uint64_t)ncols, (uint64_t)p.taps};
    const uint64_t str[2] = {(uint64_t)cred * 2, (uint64_t)ncols * cred * 2};
    const uint32_t box[3] = {(uint32_t)p.kc, (uint32_t)p.block_n, 1u};
    rc = encode_map(&p.b_map, b, 3, dims, str, box, row_bytes, "weights");
    if (rc) { delete pl; return rc; }
  }
  {  // output: dims (ncols, W, H, N)
    const uint64_t dims[4] = {(uint64_t)ncols, (uint64_t)oW, (uint64_t)oH, (uint64_t)d->N};
    const uint64_t str[3] = {(uint64_t)ncols * 2, (uint64_t)oW * ncols * 2, (uint64_t)oH * oW * ncols * 2};
    const uint32_t box[4] = {(uint32_t)p.gw, (uint32_t)p.BW, (uint32_t)p.BH, (uint32_t)p.BN};
    rc = encode_map(&p.o_map, out, 4, dims, str, box, p.gw * 2, "output");
    if (rc) { delete pl; return rc; }
  }
  const long total = (long)p.tiles_w * p.tiles_h * p.tiles_n * p.n_tiles;
  pl->grid = (int)(total < RN_NUM_SMS ? total : RN_NUM_SMS);
  cudaError_t e = cudaFuncSetAttribute(conv_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(227 * 1024));
  if (e != cudaSuccess) { delete pl; return fail(RN_ERR_CUDA, "convtc: cudaFuncSetAttribute: %s", cudaGetErrorString(e)); }
  *plan = pl;
  return RN_OK;
}

extern "C" int rn_convtc_run(rn_convtc_plan* plan, float beta, void* stream) {
  RN_REQUIRE(plan, "convtc_run: NULL plan");
  if (beta != 0.f) return fail(RN_ERR_UNSUPPORTED, "convtc_run: accumulation (beta=1) is not supported by this kernel");
  conv_tc_kernel<<<plan->grid, TC_THREADS, plan->smem, as_stream(stream)>>>(plan->p);
  return check_launch("conv_tc_kernel");
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
