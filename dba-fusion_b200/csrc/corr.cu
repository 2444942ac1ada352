// Correlation-volume lookup (corr_index_forward/backward) for sm_100a.
//
// Layout: volume [n, h1*w1, h2, w2]: every source pixel owns a contiguous h2*w2 "slab"; the
// lookup reads an 8x8 window of that slab and writes a 7x7 bilinear window into
// corr [n, 7(x offset), 7(y offset), h1*w1].  One thread = one source pixel, consecutive
// threads = consecutive pixels, so coordinate loads and all 49 output stores are coalesced.
// The 8 taps of a window row are 16 contiguous bytes at 2-byte alignment: they are fetched
// with two aligned 16-byte loads and realigned with 64-bit funnel shifts (2 loads per row
// instead of 8 scalar ones; each warp-wide load touches 32 different lines, so the number of
// load instructions is what bounds the L1 pipe here).
//
// fp16 arithmetic reproduces the reference's rounding sequence exactly
// (src/correlation_kernels.cu:43-65: weights rounded to half, every product and every
// accumulate rounded to half, taps visited x-offset outer / y-offset inner).  The reference
// does each at::Half operation as "convert to float, operate, round to half"; for + and * that
// double rounding is innocuous (24 >= 2*11 + 2 significand bits), so native mul.rn.f16x2 /
// add.rn.f16x2 give the same bits.  Two neighbouring x offsets share one packed accumulator;
// every accumulator still receives its four contributions in the reference's order.
#include "common.cuh"

namespace dbaf {
namespace {

constexpr int kThreads = 128;
constexpr int kMaxLevels = 8;

struct LevelPtrs {
  const void* vol[kMaxLevels];
};

__device__ __forceinline__ float round_h(float x) { return __half2float(__float2half_rn(x)); }

// floor(x) as an int the way the reference converts it (saturating, NaN -> 0), then clamped to
// +-2^24 so that the window arithmetic below can never overflow.  Any base beyond the clamp is
// out of bounds for every tap either way, so results are unchanged.
__device__ __forceinline__ int window_base(float floored, int r) {
  const int v = (int)floored;
  return max(-(1 << 24), min(1 << 24, v)) - r;
}

// non-contractable packed half arithmetic (.rn forbids fusing the pair into an fma)
__device__ __forceinline__ unsigned int hmul2_rn(unsigned int a, unsigned int b) {
  unsigned int d;
  asm("mul.rn.f16x2 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b));
  return d;
}
__device__ __forceinline__ unsigned int hadd2_rn(unsigned int a, unsigned int b) {
  unsigned int d;
  asm("add.rn.f16x2 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b));
  return d;
}
__device__ __forceinline__ unsigned int bcast_h(float w) {
  const unsigned int h = __half_as_ushort(__float2half_rn(w));
  return h | (h << 16);
}

__device__ __forceinline__ unsigned long long funnel64(unsigned long long lo, unsigned long long hi, int sh) {
  return sh ? (lo >> sh) | (hi << (64 - sh)) : lo;
}

// Loads the 8 fp16 taps x = xs..xs+7 of one slab row into t[0..3] (packed pairs).
// `row` points at element 0 of the slab row; `row_elem` is its element index in the whole volume.
__device__ __forceinline__ void load_row8_h(const __half* __restrict__ vol, long long row_elem, int xs,
                                            long long total_elems, bool vec_ok, int w2,
                                            unsigned int t[4]) {
  const long long first = row_elem + xs;
  const long long c0 = first & ~7LL;
  const int a = (int)(first - c0);
  if (vec_ok && c0 >= 0 && c0 + 16 <= total_elems) {
    const uint4 A = __ldg(reinterpret_cast<const uint4*>(vol + c0));
    uint4 B = make_uint4(0, 0, 0, 0);
    if (a) B = __ldg(reinterpret_cast<const uint4*>(vol + c0 + 8));
    const unsigned long long q0 = ((unsigned long long)A.y << 32) | A.x;
    const unsigned long long q1 = ((unsigned long long)A.w << 32) | A.z;
    const unsigned long long q2 = ((unsigned long long)B.y << 32) | B.x;
    const unsigned long long q3 = ((unsigned long long)B.w << 32) | B.z;
    const int sh = (a & 3) * 16;
    const bool up = a >= 4;
    const unsigned long long lo = funnel64(up ? q1 : q0, up ? q2 : q1, sh);
    const unsigned long long hi = funnel64(up ? q2 : q1, up ? q3 : q2, sh);
    t[0] = (unsigned int)lo; t[1] = (unsigned int)(lo >> 32);
    t[2] = (unsigned int)hi; t[3] = (unsigned int)(hi >> 32);
  } else {
    const unsigned short* v16 = reinterpret_cast<const unsigned short*>(vol);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int x0 = xs + 2 * i, x1 = x0 + 1;
      const unsigned int lo = (x0 >= 0 && x0 < w2) ? __ldg(v16 + row_elem + x0) : 0u;
      const unsigned int hi = (x1 >= 0 && x1 < w2) ? __ldg(v16 + row_elem + x1) : 0u;
      t[i] = lo | (hi << 16);
    }
  }
}

// One launch covers `levels` pyramid levels (blockIdx.z = level*n + edge).  For the plain
// corr_index_forward call levels == 1 and ch_total == rd*rd.
template <typename T, int R>
__global__ void __launch_bounds__(kThreads)
k_corr_lookup(LevelPtrs lv, const float* __restrict__ coords, int n, int HW1, int h2_0, int w2_0,
              int ch_total, T* __restrict__ out, const int* __restrict__ slot, int n_vol) {
  // slot != null: edge e reads volume slot[e] of a pool of n_vol volumes (CorrPool); else its own, n_vol == n
  constexpr int RD = 2 * R + 1;
  const int level = blockIdx.z / n;
  const int e = blockIdx.z - level * n;
  const int p = blockIdx.x * kThreads + threadIdx.x;
  if (p >= HW1) return;
  const int h2 = h2_0 >> level, w2 = w2_0 >> level;
  const float scale = 1.0f / (float)(1 << level);
  const float x0 = __ldg(coords + ((size_t)e * 2 + 0) * HW1 + p) * scale;
  const float y0 = __ldg(coords + ((size_t)e * 2 + 1) * HW1 + p) * scale;
  const float fx0 = floorf(x0), fy0 = floorf(y0);
  const float dx = x0 - fx0, dy = y0 - fy0;
  int bx = window_base(fx0, R), by = window_base(fy0, R);

  // a slot outside the pool (a caller's bookkeeping error) reads as an all-zero volume instead of foreign memory
  const int vslot = slot ? min(max(slot[e], 0), n_vol - 1) : e;
  const bool bad_slot = slot != nullptr && slot[e] != vslot;
  const long long slab = ((long long)vslot * HW1 + p) * (long long)(h2 * w2);
  const long long total = (long long)n_vol * HW1 * (long long)(h2 * w2);
  if (bad_slot) { bx = 1 << 24; by = 1 << 24; }      // every tap out of bounds: zeros
  // selected with compares so that the pointer table stays in the parameter bank (no local copy)
  const void* volp = lv.vol[0];
#pragma unroll
  for (int l = 1; l < kMaxLevels; ++l)
    if (level == l) volp = lv.vol[l];
  T* o = out + ((size_t)e * ch_total + (size_t)level * RD * RD) * HW1 + p;

  if constexpr (sizeof(T) == 2) {
    const __half* vol = static_cast<const __half*>(volp);
    const bool vec_ok = (reinterpret_cast<size_t>(vol) & 15) == 0;
    // weights rounded to half, as scalar_t(dx*dy) does; both lanes of a packed register carry the same weight
    const unsigned int w11 = bcast_h(dx * dy), w10 = bcast_h(dx * (1.0f - dy));
    const unsigned int w01 = bcast_h((1.0f - dx) * dy), w00 = bcast_h((1.0f - dx) * (1.0f - dy));
    static_assert(RD + 1 == 8, "row loader is written for radius 3");
    unsigned int taps[RD + 1][(RD + 1) / 2];   // [window row][x pair]: lanes = x offsets 2k, 2k+1
    unsigned int rowmask = 0;
#pragma unroll
    for (int j = 0; j < RD + 1; ++j) {
      const int y1 = by + j;
      const bool rok = y1 >= 0 && y1 < h2;
      if (rok) {
        rowmask |= 1u << j;
        load_row8_h(vol, slab + (long long)y1 * w2, bx, total, vec_ok, w2, taps[j]);
      } else {
#pragma unroll
        for (int i = 0; i < (RD + 1) / 2; ++i) taps[j][i] = 0u;
      }
    }
    // lane masks of the taps inside the image: aligned pairs (x = 2k, 2k+1) and shifted pairs (2k+1, 2k+2)
    unsigned int m_al[4], m_sh[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const unsigned int c0 = (bx + 2 * k >= 0 && bx + 2 * k < w2) ? 0xffffu : 0u;
      const unsigned int c1 = (bx + 2 * k + 1 >= 0 && bx + 2 * k + 1 < w2) ? 0xffffu : 0u;
      const unsigned int c2 = (bx + 2 * k + 2 >= 0 && bx + 2 * k + 2 < w2) ? 0xffffu : 0u;
      m_al[k] = c0 | (c1 << 16);
      m_sh[k] = c1 | (c2 << 16);
    }
    unsigned short* oh = reinterpret_cast<unsigned short*>(o);
    // output (x offset a, y offset b) = (((0 + w00 s[a][b]) + w01 s[a][b+1]) + w10 s[a+1][b]) + w11 s[a+1][b+1],
    // a contribution is skipped (not added as zero) when its tap lies outside the image
#pragma unroll
    for (int b = 0; b < RD; ++b) {
      const unsigned int r0 = ((rowmask >> b) & 1u) ? 0xffffffffu : 0u;
      const unsigned int r1 = ((rowmask >> (b + 1)) & 1u) ? 0xffffffffu : 0u;
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const unsigned int a0 = taps[b][k], a1 = taps[b + 1][k];
        const unsigned int s0 = __byte_perm(a0, k < 3 ? taps[b][k + 1] : 0u, 0x5432);
        const unsigned int s1 = __byte_perm(a1, k < 3 ? taps[b + 1][k + 1] : 0u, 0x5432);
        unsigned int acc = 0u, m, r;
        m = m_al[k] & r0; r = hadd2_rn(acc, hmul2_rn(a0, w00)); acc = (r & m) | (acc & ~m);
        m = m_al[k] & r1; r = hadd2_rn(acc, hmul2_rn(a1, w01)); acc = (r & m) | (acc & ~m);
        m = m_sh[k] & r0; r = hadd2_rn(acc, hmul2_rn(s0, w10)); acc = (r & m) | (acc & ~m);
        m = m_sh[k] & r1; r = hadd2_rn(acc, hmul2_rn(s1, w11)); acc = (r & m) | (acc & ~m);
        oh[(size_t)((2 * k) * RD + b) * HW1] = (unsigned short)(acc & 0xffffu);
        if (2 * k + 1 < RD) oh[(size_t)((2 * k + 1) * RD + b) * HW1] = (unsigned short)(acc >> 16);
      }
    }
  } else {
    const float* vol = static_cast<const float*>(volp);
    float acc[RD][RD];
#pragma unroll
    for (int a = 0; a < RD; ++a)
#pragma unroll
      for (int b = 0; b < RD; ++b) acc[a][b] = 0.f;
    const float w11 = dx * dy, w10 = dx * (1.0f - dy);
    const float w01 = (1.0f - dx) * dy, w00 = (1.0f - dx) * (1.0f - dy);
#pragma unroll
    for (int i = 0; i < RD + 1; ++i) {
      const int x1 = bx + i;
#pragma unroll
      for (int j = 0; j < RD + 1; ++j) {
        const int y1 = by + j;
        const bool inb = x1 >= 0 && x1 < w2 && y1 >= 0 && y1 < h2;
        const float s = inb ? __ldg(vol + slab + (long long)y1 * w2 + x1) : 0.f;
        if (i > 0 && j > 0) acc[i - 1][j - 1] = inb ? fmaf(s, w11, acc[i - 1][j - 1]) : acc[i - 1][j - 1];
        if (i > 0 && j < RD) acc[i - 1][j] = inb ? fmaf(s, w10, acc[i - 1][j]) : acc[i - 1][j];
        if (i < RD && j > 0) acc[i][j - 1] = inb ? fmaf(s, w01, acc[i][j - 1]) : acc[i][j - 1];
        if (i < RD && j < RD) acc[i][j] = inb ? fmaf(s, w00, acc[i][j]) : acc[i][j];
      }
    }
#pragma unroll
    for (int a = 0; a < RD; ++a)
#pragma unroll
      for (int b = 0; b < RD; ++b) o[(size_t)(a * RD + b) * HW1] = acc[a][b];
  }

}

// Generic-radius fallback (any r): same arithmetic, taps loaded one by one, outputs
// accumulated in global memory order-preserving (one thread owns its outputs).
template <typename T>
__global__ void __launch_bounds__(kThreads)
k_corr_lookup_any(const T* __restrict__ vol, const float* __restrict__ coords, int HW1, int h2, int w2,
                  int r, T* __restrict__ out) {
  const int e = blockIdx.z;
  const int p = blockIdx.x * kThreads + threadIdx.x;
  if (p >= HW1) return;
  const int rd = 2 * r + 1;
  const float x0 = __ldg(coords + ((size_t)e * 2 + 0) * HW1 + p);
  const float y0 = __ldg(coords + ((size_t)e * 2 + 1) * HW1 + p);
  const float fx0 = floorf(x0), fy0 = floorf(y0);
  const float dx = x0 - fx0, dy = y0 - fy0;
  const int bx = window_base(fx0, r), by = window_base(fy0, r);
  const bool half = sizeof(T) == 2;
  float w11 = dx * dy, w10 = dx * (1.0f - dy), w01 = (1.0f - dx) * dy, w00 = (1.0f - dx) * (1.0f - dy);
  if (half) { w11 = round_h(w11); w10 = round_h(w10); w01 = round_h(w01); w00 = round_h(w00); }
  const T* slab = vol + ((size_t)e * HW1 + p) * (size_t)(h2 * w2);
  T* o = out + (size_t)e * rd * rd * HW1 + p;
  for (int c = 0; c < rd * rd; ++c) o[(size_t)c * HW1] = T(0.f);
  auto upd = [&](int a, int b, float s, float w) {
    T* q = o + (size_t)(a * rd + b) * HW1;
    const float cur = (float)*q;
    *q = half ? T(round_h(cur + round_h(s * w))) : T(fmaf(s, w, cur));
  };
  for (int i = 0; i <= rd; ++i)
    for (int j = 0; j <= rd; ++j) {
      const int x1 = bx + i, y1 = by + j;
      if (x1 < 0 || x1 >= w2 || y1 < 0 || y1 >= h2) continue;
      const float s = (float)slab[(size_t)y1 * w2 + x1];
      if (i > 0 && j > 0) upd(i - 1, j - 1, s, w11);
      if (i > 0 && j < rd) upd(i - 1, j, s, w10);
      if (i < rd && j > 0) upd(i, j - 1, s, w01);
      if (i < rd && j < rd) upd(i, j, s, w00);
    }
}

// Adjoint: volume_grad is zero-filled by the caller side (memset), then every pixel writes
// the (2r+2)^2 window of its own slab.  src/correlation_kernels.cu:73-124.
template <typename T>
__global__ void __launch_bounds__(kThreads)
k_corr_lookup_bwd(const float* __restrict__ coords, const T* __restrict__ cg, int HW1, int h2, int w2,
                  int r, T* __restrict__ vg) {
  const int e = blockIdx.z;
  const int p = blockIdx.x * kThreads + threadIdx.x;
  if (p >= HW1) return;
  const int rd = 2 * r + 1;
  const float x0 = __ldg(coords + ((size_t)e * 2 + 0) * HW1 + p);
  const float y0 = __ldg(coords + ((size_t)e * 2 + 1) * HW1 + p);
  const float fx0 = floorf(x0), fy0 = floorf(y0);
  const float dx = x0 - fx0, dy = y0 - fy0;
  const int bx = window_base(fx0, r), by = window_base(fy0, r);
  const bool half = sizeof(T) == 2;
  float w11 = dx * dy, w10 = dx * (1.0f - dy), w01 = (1.0f - dx) * dy, w00 = (1.0f - dx) * (1.0f - dy);
  if (half) { w11 = round_h(w11); w10 = round_h(w10); w01 = round_h(w01); w00 = round_h(w00); }
  const T* g0 = cg + (size_t)e * rd * rd * HW1 + p;
  T* slab = vg + ((size_t)e * HW1 + p) * (size_t)(h2 * w2);
  auto madd = [&](float g, int a, int b, float w) {
    const float c = (float)g0[(size_t)(a * rd + b) * HW1];
    return half ? round_h(g + round_h(c * w)) : fmaf(c, w, g);
  };
  for (int i = 0; i <= rd; ++i)
    for (int j = 0; j <= rd; ++j) {
      const int x1 = bx + i, y1 = by + j;
      if (x1 < 0 || x1 >= w2 || y1 < 0 || y1 >= h2) continue;
      float g = 0.f;
      if (i > 0 && j > 0) g = madd(g, i - 1, j - 1, w11);
      if (i > 0 && j < rd) g = madd(g, i - 1, j, w10);
      if (i < rd && j > 0) g = madd(g, i, j - 1, w01);
      if (i < rd && j < rd) g = madd(g, i, j, w00);
      slab[(size_t)y1 * w2 + x1] = T(g);
    }
}

template <typename T>
int launch_lookup(const LevelPtrs& lv, int levels, const float* coords, int n, int HW1, int h2, int w2,
                  int radius, T* out, cudaStream_t st, const int* slot = nullptr, int n_vol = -1) {
  if (n_vol < 0) n_vol = n;
  dim3 grid((HW1 + kThreads - 1) / kThreads, 1, n * levels);
  if (radius == 3) {
    k_corr_lookup<T, 3><<<grid, kThreads, 0, st>>>(lv, coords, n, HW1, h2, w2, levels * 49, out, slot, n_vol);
  } else {
    // generic radius: one level per launch
    const int rd = 2 * radius + 1;
    for (int l = 0; l < levels; ++l) {
      DBAF_REQUIRE(levels == 1, "pyramid lookup supports radius 3 only");
      dim3 g1((HW1 + kThreads - 1) / kThreads, 1, n);
      k_corr_lookup_any<T><<<g1, kThreads, 0, st>>>(static_cast<const T*>(lv.vol[l]), coords, HW1, h2, w2, radius, out);
      (void)rd;
    }
  }
  DBAF_LAUNCH_CHECK("k_corr_lookup");
  return DBAF_OK;
}

}  // namespace
}  // namespace dbaf

using namespace dbaf;

extern "C" int dbaf_corr_index_forward(const void* volume, int dtype, const float* coords, int n, int h1,
                                       int w1, int h2, int w2, int radius, void* corr, void* stream) {
  DBAF_REQUIRE(dtype == DBAF_F16 || dtype == DBAF_F32, "dtype must be DBAF_F16 or DBAF_F32");
  DBAF_REQUIRE(n >= 0 && h1 > 0 && w1 > 0 && h2 >= 0 && w2 >= 0 && radius >= 0, "bad sizes");
  if (n == 0) return DBAF_OK;
  LevelPtrs lv{};
  lv.vol[0] = volume;
  if (dtype == DBAF_F16)
    return launch_lookup<__half>(lv, 1, coords, n, h1 * w1, h2, w2, radius, static_cast<__half*>(corr), (cudaStream_t)stream);
  return launch_lookup<float>(lv, 1, coords, n, h1 * w1, h2, w2, radius, static_cast<float*>(corr), (cudaStream_t)stream);
}

extern "C" int dbaf_corr_lookup_pyramid(const void* const* volumes_host, int levels, int dtype,
                                        const float* coords, int n, int h1, int w1, int h2, int w2,
                                        int radius, void* out, void* stream) {
  DBAF_REQUIRE(dtype == DBAF_F16 || dtype == DBAF_F32, "dtype must be DBAF_F16 or DBAF_F32");
  DBAF_REQUIRE(levels >= 1 && levels <= kMaxLevels, "levels out of range");
  DBAF_REQUIRE(radius == 3, "pyramid lookup supports radius 3 only");
  DBAF_REQUIRE(n >= 0 && h1 > 0 && w1 > 0 && h2 >= 0 && w2 >= 0, "bad sizes");
  if (n == 0) return DBAF_OK;
  LevelPtrs lv{};
  for (int l = 0; l < levels; ++l) lv.vol[l] = volumes_host[l];
  if (dtype == DBAF_F16)
    return launch_lookup<__half>(lv, levels, coords, n, h1 * w1, h2, w2, radius, static_cast<__half*>(out), (cudaStream_t)stream);
  return launch_lookup<float>(lv, levels, coords, n, h1 * w1, h2, w2, radius, static_cast<float*>(out), (cudaStream_t)stream);
}

// The same lookup over a slot pool: edge e reads volume slots[e] of pools holding `capacity` volumes per level
// (slots are not range-checked here: dbaf_corr_volume_pyramid_pool checked them when it filled the pool).
extern "C" int dbaf_corr_lookup_pyramid_pool(const void* const* pool_host, int levels, int dtype, const float* coords,
                                             const int32_t* slots, int n, int capacity, int h1, int w1, int h2, int w2,
                                             int radius, void* out, void* stream) {
  DBAF_REQUIRE(dtype == DBAF_F16 || dtype == DBAF_F32, "dtype must be DBAF_F16 or DBAF_F32");
  DBAF_REQUIRE(levels >= 1 && levels <= kMaxLevels, "levels out of range");
  DBAF_REQUIRE(radius == 3, "pyramid lookup supports radius 3 only");
  DBAF_REQUIRE(n >= 0 && capacity > 0 && h1 > 0 && w1 > 0 && h2 >= 0 && w2 >= 0 && slots != nullptr, "bad sizes");
  if (n == 0) return DBAF_OK;
  LevelPtrs lv{};
  for (int l = 0; l < levels; ++l) lv.vol[l] = pool_host[l];
  if (dtype == DBAF_F16)
    return launch_lookup<__half>(lv, levels, coords, n, h1 * w1, h2, w2, radius, static_cast<__half*>(out), (cudaStream_t)stream,
                                 slots, capacity);
  return launch_lookup<float>(lv, levels, coords, n, h1 * w1, h2, w2, radius, static_cast<float*>(out), (cudaStream_t)stream, slots,
                              capacity);
}

extern "C" int dbaf_corr_index_backward(const float* coords, const void* corr_grad, int dtype, int n,
                                        int h1, int w1, int h2, int w2, int radius, void* volume_grad,
                                        void* stream) {
  DBAF_REQUIRE(dtype == DBAF_F16 || dtype == DBAF_F32, "dtype must be DBAF_F16 or DBAF_F32");
  DBAF_REQUIRE(n >= 0 && h1 > 0 && w1 > 0 && h2 >= 0 && w2 >= 0 && radius >= 0, "bad sizes");
  if (n == 0) return DBAF_OK;
  cudaStream_t st = (cudaStream_t)stream;
  const size_t esz = dtype == DBAF_F16 ? 2 : 4;
  const size_t bytes = (size_t)n * h1 * w1 * h2 * w2 * esz;
  DBAF_CUDA(cudaMemsetAsync(volume_grad, 0, bytes, st));
  dim3 grid((h1 * w1 + kThreads - 1) / kThreads, 1, n);
  if (dtype == DBAF_F16)
    k_corr_lookup_bwd<__half><<<grid, kThreads, 0, st>>>(coords, static_cast<const __half*>(corr_grad), h1 * w1, h2, w2, radius, static_cast<__half*>(volume_grad));
  else
    k_corr_lookup_bwd<float><<<grid, kThreads, 0, st>>>(coords, static_cast<const float*>(corr_grad), h1 * w1, h2, w2, radius, static_cast<float*>(volume_grad));
  DBAF_LAUNCH_CHECK("k_corr_lookup_bwd");
  return DBAF_OK;
}
