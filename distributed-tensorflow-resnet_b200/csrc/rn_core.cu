// Library state (error string, launch counter) and small element-wise utilities.
#include "rn_common.cuh"
#include <string.h>

namespace rn {
thread_local char g_err[512] = "";
std::atomic<uint64_t> g_launches{0};

template <typename S, typename D>
__global__ void cast_kernel(const S* __restrict__ src, D* __restrict__ dst, int64_t n) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    dst[i] = from_f<D>(to_f<S>(src[i]));
}

template <typename T>
__global__ void add_inplace_kernel(T* dst, const T* __restrict__ src, int64_t nvec, int64_t n) {
  constexpr int V = Vec<T>::N;
  const int64_t gid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, gsz = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = gid; i < nvec; i += gsz) {
    Vec<T> a, b;
    float fa[V], fb[V];
    a.load(dst + i * V);
    b.load(src + i * V);
    a.unpack(fa);
    b.unpack(fb);
#pragma unroll
    for (int j = 0; j < V; ++j) fa[j] += fb[j];
    a.pack(fa);
    a.store(dst + i * V);
  }
  for (int64_t i = nvec * V + gid; i < n; i += gsz) dst[i] = from_f<T>(to_f<T>(dst[i]) + to_f<T>(src[i]));
}

template <typename T>
__global__ void add_kernel(const T* __restrict__ a_, const T* __restrict__ b_, T* __restrict__ y, int64_t nvec,
                           int64_t n) {
  constexpr int V = Vec<T>::N;
  const int64_t gid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, gsz = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = gid; i < nvec; i += gsz) {
    Vec<T> a, b;
    float fa[V], fb[V];
    a.load(a_ + i * V);
    b.load(b_ + i * V);
    a.unpack(fa);
    b.unpack(fb);
#pragma unroll
    for (int j = 0; j < V; ++j) fa[j] += fb[j];
    a.pack(fa);
    a.store(y + i * V);
  }
  for (int64_t i = nvec * V + gid; i < n; i += gsz) y[i] = from_f<T>(to_f<T>(a_[i]) + to_f<T>(b_[i]));
}

template <typename S, typename D>
__global__ void pad_channels_kernel(const S* __restrict__ src, D* __restrict__ dst, int64_t pixels, int Cs, int Cd) {
  const int64_t total = pixels * Cd;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t p = i / Cd;
    const int c = (int)(i - p * Cd);
    dst[i] = from_f<D>(c < Cs ? to_f<S>(src[p * Cs + c]) : 0.f);
  }
}
}  // namespace rn

using namespace rn;

extern "C" const char* rn_last_error(void) { return g_err; }
extern "C" int rn_abi_version(void) { return RN_ABI_VERSION; }
extern "C" uint64_t rn_launch_count(void) { return g_launches.load(); }

extern "C" int rn_cast(const void* src, void* dst, int64_t n, int32_t sd, int32_t dd, void* stream) {
  RN_REQUIRE(src && dst && n >= 0, "cast: bad arguments");
  if (n == 0) return RN_OK;
  const int g = grid_for(n, 256);
  cudaStream_t st = as_stream(stream);
  if (sd == RN_F32 && dd == RN_BF16)
    cast_kernel<float, __nv_bfloat16><<<g, 256, 0, st>>>((const float*)src, (__nv_bfloat16*)dst, n);
  else if (sd == RN_BF16 && dd == RN_F32)
    cast_kernel<__nv_bfloat16, float><<<g, 256, 0, st>>>((const __nv_bfloat16*)src, (float*)dst, n);
  else if (sd == RN_F32 && dd == RN_F32)
    cast_kernel<float, float><<<g, 256, 0, st>>>((const float*)src, (float*)dst, n);
  else if (sd == RN_BF16 && dd == RN_BF16)
    cast_kernel<__nv_bfloat16, __nv_bfloat16><<<g, 256, 0, st>>>((const __nv_bfloat16*)src, (__nv_bfloat16*)dst, n);
  else
    return fail(RN_ERR_INVALID, "cast: bad dtypes %d -> %d", sd, dd);
  return check_launch("cast_kernel");
}

extern "C" int rn_add_inplace(void* dst, const void* src, int64_t n, int32_t dtype, void* stream) {
  RN_REQUIRE(dst && src && n >= 0, "add_inplace: bad arguments");
  if (n == 0) return RN_OK;
  cudaStream_t st = as_stream(stream);
  const bool al = (((uintptr_t)dst | (uintptr_t)src) & 15) == 0;
  if (dtype == RN_F32) {
    int64_t nv = al ? n / 4 : 0;
    add_inplace_kernel<float><<<grid_for(n / 4 + 1, 256), 256, 0, st>>>((float*)dst, (const float*)src, nv, n);
  } else if (dtype == RN_BF16) {
    int64_t nv = al ? n / 8 : 0;
    add_inplace_kernel<__nv_bfloat16><<<grid_for(n / 8 + 1, 256), 256, 0, st>>>((__nv_bfloat16*)dst,
                                                                              (const __nv_bfloat16*)src, nv, n);
  } else {
    return fail(RN_ERR_INVALID, "add_inplace: bad dtype %d", dtype);
  }
  return check_launch("add_inplace_kernel");
}

extern "C" int rn_add(const void* a, const void* b, void* y, int64_t n, int32_t dtype, void* stream) {
  RN_REQUIRE(a && b && y && n >= 0, "add: bad arguments");
  if (n == 0) return RN_OK;
  cudaStream_t st = as_stream(stream);
  const bool al = (((uintptr_t)a | (uintptr_t)b | (uintptr_t)y) & 15) == 0;
  if (dtype == RN_F32) {
    int64_t nv = al ? n / 4 : 0;
    add_kernel<float><<<grid_for(n / 4 + 1, 256), 256, 0, st>>>((const float*)a, (const float*)b, (float*)y, nv, n);
  } else if (dtype == RN_BF16) {
    int64_t nv = al ? n / 8 : 0;
    add_kernel<__nv_bfloat16><<<grid_for(n / 8 + 1, 256), 256, 0, st>>>((const __nv_bfloat16*)a, (const __nv_bfloat16*)b,
                                                                       (__nv_bfloat16*)y, nv, n);
  } else {
    return fail(RN_ERR_INVALID, "add: bad dtype %d", dtype);
  }
  return check_launch("add_kernel");
}

extern "C" int rn_pad_channels(const void* src, void* dst, int64_t pixels, int32_t Cs, int32_t Cd, int32_t sd,
                               int32_t dd, void* stream) {
  RN_REQUIRE(src && dst && pixels >= 0 && Cs > 0 && Cd >= Cs, "pad_channels: bad arguments");
  if (pixels == 0) return RN_OK;
  cudaStream_t st = as_stream(stream);
  const int g = grid_for(pixels * Cd, 256);
  if (sd == RN_F32 && dd == RN_BF16)
    pad_channels_kernel<float, __nv_bfloat16><<<g, 256, 0, st>>>((const float*)src, (__nv_bfloat16*)dst, pixels, Cs, Cd);
  else if (sd == RN_F32 && dd == RN_F32)
    pad_channels_kernel<float, float><<<g, 256, 0, st>>>((const float*)src, (float*)dst, pixels, Cs, Cd);
  else if (sd == RN_BF16 && dd == RN_BF16)
    pad_channels_kernel<__nv_bfloat16, __nv_bfloat16><<<g, 256, 0, st>>>((const __nv_bfloat16*)src, (__nv_bfloat16*)dst,
                                                                        pixels, Cs, Cd);
  else
    return fail(RN_ERR_INVALID, "pad_channels: bad dtypes %d -> %d", sd, dd);
  return check_launch("pad_channels_kernel");
}

// ------------------------------------------------------------------------------------------
// CRC-32C (Castagnoli, reflected polynomial 0x82F63B78), slicing-by-8, host code: the checksum of
// TensorFlow's checkpoint bundles (tensor data and index-table blocks), used by tf_io.py.
// ------------------------------------------------------------------------------------------
namespace rn {
static uint32_t g_crc_tab[8][256];
static std::atomic<int> g_crc_ready{0};
static void crc32c_init() {
  if (g_crc_ready.load(std::memory_order_acquire)) return;
  uint32_t t[8][256];
  for (uint32_t i = 0; i < 256; ++i) {
    uint32_t c = i;
    for (int k = 0; k < 8; ++k) c = (c & 1u) ? (c >> 1) ^ 0x82F63B78u : c >> 1;
    t[0][i] = c;
  }
  for (uint32_t i = 0; i < 256; ++i)
    for (int s = 1; s < 8; ++s) t[s][i] = (t[s - 1][i] >> 8) ^ t[0][t[s - 1][i] & 0xffu];
  memcpy(g_crc_tab, t, sizeof(t));          // idempotent: concurrent initialisers write identical bytes
  g_crc_ready.store(1, std::memory_order_release);
}
}  // namespace rn

extern "C" uint32_t rn_crc32c(const void* data, size_t n, uint32_t crc) {
  rn::crc32c_init();
  const uint8_t* p = (const uint8_t*)data;
  uint32_t c = ~crc;
  while (n && ((uintptr_t)p & 7)) { c = rn::g_crc_tab[0][(c ^ *p++) & 0xffu] ^ (c >> 8); --n; }
  while (n >= 8) {
    uint64_t w;
    memcpy(&w, p, 8);
    w ^= c;
    c = rn::g_crc_tab[7][w & 0xff] ^ rn::g_crc_tab[6][(w >> 8) & 0xff] ^ rn::g_crc_tab[5][(w >> 16) & 0xff] ^
        rn::g_crc_tab[4][(w >> 24) & 0xff] ^ rn::g_crc_tab[3][(w >> 32) & 0xff] ^ rn::g_crc_tab[2][(w >> 40) & 0xff] ^
        rn::g_crc_tab[1][(w >> 48) & 0xff] ^ rn::g_crc_tab[0][(w >> 56) & 0xff];
    p += 8; n -= 8;
  }
  while (n--) c = rn::g_crc_tab[0][(c ^ *p++) & 0xffu] ^ (c >> 8);
  return ~c;
}
