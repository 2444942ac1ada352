// On-the-fly correlation (altcorr_forward/backward): the same 7x7 bilinear windows as the
// volume lookup, but every tap is a 128-channel dot product taken straight from the feature
// maps.  Features are pixel-major ([B,H,W,C], 512 B per pixel at C=128), so each tap is read
// with 16-byte loads; the source pixel's own feature chunk stays in registers.
// Reference: src/altcorr_kernel.cu:27-149 (forward), 152-286 (backward).
#include "common.cuh"

namespace dbaf {
namespace {

constexpr int kThreads = 128;
constexpr int kChunk = 32;   // CHANNEL_STRIDE of the reference: partial dots over 32 channels

// see corr.cu: saturating float->int like the reference, clamped so window arithmetic cannot overflow
__device__ __forceinline__ int window_base(float floored, int r) {
  const int v = (int)floored;
  return max(-(1 << 24), min(1 << 24, v)) - r;
}

template <int R>
__global__ void __launch_bounds__(kThreads)
k_altcorr_fwd(const float* __restrict__ f1, const float* __restrict__ f2,
              const float* __restrict__ coords, int S, int H1, int W1, int H2, int W2, int C,
              float* __restrict__ corr) {
  constexpr int RD = 2 * R + 1;
  const int HW1 = H1 * W1;
  const int bs = blockIdx.y;             // b*S + s
  const int b = bs / S;
  const int p = blockIdx.x * kThreads + threadIdx.x;
  if (p >= HW1) return;
  const float x = __ldg(coords + ((size_t)bs * HW1 + p) * 2 + 0);
  const float y = __ldg(coords + ((size_t)bs * HW1 + p) * 2 + 1);
  const float fx0 = floorf(x), fy0 = floorf(y);
  const float dx = x - fx0, dy = y - fy0;
  const int bx = window_base(fx0, R), by = window_base(fy0, R);
  const float wnw = dy * dx, wne = dy * (1.0f - dx), wsw = (1.0f - dy) * dx, wse = (1.0f - dy) * (1.0f - dx);

  float acc[RD][RD];   // [iy][ix]
#pragma unroll
  for (int a = 0; a < RD; ++a)
#pragma unroll
    for (int c = 0; c < RD; ++c) acc[a][c] = 0.f;

  const float* f1p = f1 + ((size_t)b * HW1 + p) * C;
  const float* f2b = f2 + (size_t)b * H2 * W2 * C;

  for (int c0 = 0; c0 < C; c0 += kChunk) {
    float4 a4[kChunk / 4];
#pragma unroll
    for (int k = 0; k < kChunk / 4; ++k) a4[k] = __ldg(reinterpret_cast<const float4*>(f1p + c0) + k);
#pragma unroll
    for (int iy = 0; iy < RD + 1; ++iy) {
      const int h2 = by + iy;
#pragma unroll
      for (int ix = 0; ix < RD + 1; ++ix) {
        const int w2 = bx + ix;
        float s = 0.f;
        if (h2 >= 0 && h2 < H2 && w2 >= 0 && w2 < W2) {
          const float4* g = reinterpret_cast<const float4*>(f2b + ((size_t)h2 * W2 + w2) * C + c0);
#pragma unroll
          for (int k = 0; k < kChunk / 4; ++k) {
            const float4 v = __ldg(g + k);
            s = fmaf(a4[k].x, v.x, s);
            s = fmaf(a4[k].y, v.y, s);
            s = fmaf(a4[k].z, v.z, s);
            s = fmaf(a4[k].w, v.w, s);
          }
        }
        if (iy > 0 && ix > 0) acc[iy - 1][ix - 1] += s * wnw;
        if (iy > 0 && ix < RD) acc[iy - 1][ix] += s * wne;
        if (iy < RD && ix > 0) acc[iy][ix - 1] += s * wsw;
        if (iy < RD && ix < RD) acc[iy][ix] += s * wse;
      }
    }
  }
  float* o = corr + (size_t)bs * RD * RD * HW1 + p;
#pragma unroll
  for (int iy = 0; iy < RD; ++iy)
#pragma unroll
    for (int ix = 0; ix < RD; ++ix) o[(size_t)(iy + RD * ix) * HW1] = acc[iy][ix];
}

// Backward: one CTA per source pixel, one thread per channel; f2 gradients are scattered with
// coalesced atomics (neighbouring pixels' windows overlap), f1 gradients accumulate in a register.
__global__ void __launch_bounds__(kThreads)
k_altcorr_bwd(const float* __restrict__ f1, const float* __restrict__ f2,
              const float* __restrict__ coords, const float* __restrict__ cg, int S, int H1, int W1,
              int H2, int W2, int C, int r, float* __restrict__ f1g, float* __restrict__ f2g) {
  const int HW1 = H1 * W1;
  const int p = blockIdx.x;
  const int b = blockIdx.y;
  const int rd = 2 * r + 1;
  for (int c = threadIdx.x; c < C; c += kThreads) {
    const float a = __ldg(f1 + ((size_t)b * HW1 + p) * C + c);
    float ga = 0.f;
    for (int s = 0; s < S; ++s) {
      const size_t bs = (size_t)b * S + s;
      const float x = __ldg(coords + (bs * HW1 + p) * 2 + 0);
      const float y = __ldg(coords + (bs * HW1 + p) * 2 + 1);
      const float fx0 = floorf(x), fy0 = floorf(y);
      const float dx = x - fx0, dy = y - fy0;
      const int bx = window_base(fx0, r), by = window_base(fy0, r);
      const float* g0 = cg + bs * rd * rd * HW1 + p;
      for (int iy = 0; iy <= rd; ++iy)
        for (int ix = 0; ix <= rd; ++ix) {
          const int h2 = by + iy, w2 = bx + ix;
          if (h2 < 0 || h2 >= H2 || w2 < 0 || w2 >= W2) continue;
          float g = 0.f;
          if (iy > 0 && ix > 0) g += __ldg(g0 + (size_t)((iy - 1) + rd * (ix - 1)) * HW1) * dy * dx;
          if (iy > 0 && ix < rd) g += __ldg(g0 + (size_t)((iy - 1) + rd * ix) * HW1) * dy * (1.0f - dx);
          if (iy < rd && ix > 0) g += __ldg(g0 + (size_t)(iy + rd * (ix - 1)) * HW1) * (1.0f - dy) * dx;
          if (iy < rd && ix < rd) g += __ldg(g0 + (size_t)(iy + rd * ix) * HW1) * (1.0f - dy) * (1.0f - dx);
          const size_t q = (((size_t)b * H2 + h2) * W2 + w2) * C + c;
          ga = fmaf(g, __ldg(f2 + q), ga);
          atomicAdd(f2g + q, g * a);
        }
    }
    f1g[((size_t)b * HW1 + p) * C + c] = ga;
  }
}

// Generic-radius forward (any r, any C): plain loops, same arithmetic.
__global__ void __launch_bounds__(kThreads)
k_altcorr_fwd_any(const float* __restrict__ f1, const float* __restrict__ f2,
                  const float* __restrict__ coords, int S, int H1, int W1, int H2, int W2, int C, int r,
                  float* __restrict__ corr) {
  const int HW1 = H1 * W1;
  const int bs = blockIdx.y;
  const int b = bs / S;
  const int p = blockIdx.x * kThreads + threadIdx.x;
  if (p >= HW1) return;
  const int rd = 2 * r + 1;
  const float x = __ldg(coords + ((size_t)bs * HW1 + p) * 2 + 0);
  const float y = __ldg(coords + ((size_t)bs * HW1 + p) * 2 + 1);
  const float fx0 = floorf(x), fy0 = floorf(y);
  const float dx = x - fx0, dy = y - fy0;
  const int bx = window_base(fx0, r), by = window_base(fy0, r);
  float* o = corr + (size_t)bs * rd * rd * HW1 + p;
  for (int c = 0; c < rd * rd; ++c) o[(size_t)c * HW1] = 0.f;
  const float* f1p = f1 + ((size_t)b * HW1 + p) * C;
  for (int c0 = 0; c0 < C; c0 += kChunk) {
    const int c1 = min(C, c0 + kChunk);
    for (int iy = 0; iy <= rd; ++iy)
      for (int ix = 0; ix <= rd; ++ix) {
        const int h2 = by + iy, w2 = bx + ix;
        float s = 0.f;
        if (h2 >= 0 && h2 < H2 && w2 >= 0 && w2 < W2) {
          const float* g = f2 + (((size_t)b * H2 + h2) * W2 + w2) * C;
          for (int k = c0; k < c1; ++k) s = fmaf(__ldg(f1p + k), __ldg(g + k), s);
        }
        if (iy > 0 && ix > 0) o[(size_t)((iy - 1) + rd * (ix - 1)) * HW1] += s * (dy * dx);
        if (iy > 0 && ix < rd) o[(size_t)((iy - 1) + rd * ix) * HW1] += s * (dy * (1.0f - dx));
        if (iy < rd && ix > 0) o[(size_t)(iy + rd * (ix - 1)) * HW1] += s * ((1.0f - dy) * dx);
        if (iy < rd && ix < rd) o[(size_t)(iy + rd * ix) * HW1] += s * ((1.0f - dy) * (1.0f - dx));
      }
  }
}

}  // namespace
}  // namespace dbaf

using namespace dbaf;

extern "C" int dbaf_altcorr_forward(const float* fmap1, const float* fmap2, const float* coords, int B,
                                    int S, int H1, int W1, int H2, int W2, int C, int radius,
                                    float* corr, void* stream) {
  DBAF_REQUIRE(B >= 0 && S >= 0 && H1 > 0 && W1 > 0 && H2 >= 0 && W2 >= 0 && C > 0 && radius >= 0, "bad sizes");
  if (B == 0 || S == 0) return DBAF_OK;
  const int HW1 = H1 * W1;
  dim3 grid((HW1 + kThreads - 1) / kThreads, B * S);
  const bool fast = radius == 3 && C % kChunk == 0 &&
                    ((reinterpret_cast<size_t>(fmap1) | reinterpret_cast<size_t>(fmap2)) & 15) == 0;
  if (fast)
    k_altcorr_fwd<3><<<grid, kThreads, 0, (cudaStream_t)stream>>>(fmap1, fmap2, coords, S, H1, W1, H2, W2, C, corr);
  else
    k_altcorr_fwd_any<<<grid, kThreads, 0, (cudaStream_t)stream>>>(fmap1, fmap2, coords, S, H1, W1, H2, W2, C, radius, corr);
  DBAF_LAUNCH_CHECK("k_altcorr_fwd");
  return DBAF_OK;
}

extern "C" int dbaf_altcorr_backward(const float* fmap1, const float* fmap2, const float* coords,
                                     const float* corr_grad, int B, int S, int H1, int W1, int H2, int W2,
                                     int C, int radius, float* fmap1_grad, float* fmap2_grad,
                                     float* coords_grad, void* stream) {
  DBAF_REQUIRE(B >= 0 && S >= 0 && H1 > 0 && W1 > 0 && H2 >= 0 && W2 >= 0 && C > 0 && radius >= 0, "bad sizes");
  if (B == 0) return DBAF_OK;
  cudaStream_t st = (cudaStream_t)stream;
  DBAF_CUDA(cudaMemsetAsync(fmap2_grad, 0, (size_t)B * H2 * W2 * C * 4, st));
  if (coords_grad) DBAF_CUDA(cudaMemsetAsync(coords_grad, 0, (size_t)B * S * H1 * W1 * 2 * 4, st));
  dim3 grid(H1 * W1, B);
  k_altcorr_bwd<<<grid, kThreads, 0, st>>>(fmap1, fmap2, coords, corr_grad, S, H1, W1, H2, W2, C, radius, fmap1_grad, fmap2_grad);
  DBAF_LAUNCH_CHECK("k_altcorr_bwd");
  return DBAF_OK;
}
