// Reprojection utilities: projmap, frame_distance, iproj, depth_filter,
// plus the library's error plumbing.  All four are streaming kernels over [frames|edges] x pixels:
// pixel index fastest, one thread per pixel, so every global access is a coalesced 128 B line.
#include "common.cuh"

#include <cstdlib>
#include <cstring>

namespace dbaf {

static thread_local std::string g_error;

void set_error(const std::string& msg) { g_error = msg; }

int cuda_fail(cudaError_t e, const char* what) {
  g_error = std::string(what) + ": " + cudaGetErrorString(e);
  return DBAF_ECUDA;
}

namespace {
bool env_is(const char* name, const char* value) {
  const char* v = std::getenv(name);
  return v != nullptr && std::strcmp(v, value) == 0;
}
}  // namespace

Options& options() {
  static Options o = [] {
    Options x{};
    x.schur_simt = env_is("DBAF_SCHUR_IMPL", "simt");
    x.band_single = env_is("DBAF_BAND_IMPL", "single");
    x.band_twisted = env_is("DBAF_BAND_IMPL", "twisted");
    x.band_smem = env_is("DBAF_BAND_SWEEP", "smem");
    x.volume_simt = env_is("DBAF_VOLUME_IMPL", "simt");
    const char* d = std::getenv("DBAF_BAND_DOMAINS");
    x.band_domains = d ? std::atoi(d) : 0;
    const char* r = std::getenv("DBAF_EIJ_RECOMPUTE");
    x.eij_recompute = r ? std::atoi(r) : -1;
    return x;
  }();
  return o;
}

namespace {

constexpr int kThreads = 256;

// projmap_kernel, src/droid_kernels.cu:471-560: no stereo special case here.
__global__ void __launch_bounds__(kThreads)
k_projmap(const float* __restrict__ poses, const float* __restrict__ disps,
          const float* __restrict__ intr, const int64_t* __restrict__ ii,
          const int64_t* __restrict__ jj, int HW, int wd, float* __restrict__ coords,
          float* __restrict__ valid) {
  const int e = blockIdx.y;
  const int p = blockIdx.x * kThreads + threadIdx.x;
  if (p >= HW) return;
  const int ix = (int)ii[e], jx = (int)jj[e];
  const float fx = __ldg(intr), fy = __ldg(intr + 1), cx = __ldg(intr + 2), cy = __ldg(intr + 3);
  const Se3 G = load_relative(poses, ix, jx);
  const int i = p / wd, j = p - i * wd;
  const float u = (float)j, v = (float)i;
  const float Xi[4] = {(u - cx) / fx, (v - cy) / fy, 1.0f, __ldg(disps + (size_t)ix * HW + p)};
  float Xj[4];
  act(G, Xi, Xj);
  float cu = u, cv = v;
  if (Xj[2] > 0.01f) {
    cu = fx * (Xj[0] / Xj[2]) + cx;
    cv = fy * (Xj[1] / Xj[2]) + cy;
  }
  float* c = coords + ((size_t)e * HW + p) * 3;
  c[0] = cu;
  c[1] = cv;
  c[2] = 0.0f;
  valid[(size_t)e * HW + p] = Xj[2] > kMinDepth ? 1.0f : 0.0f;
}

// DepthVideo.reproject -> pops.projective_transform without Jacobians (dbaf/depth_video.py:221-229,
// dbaf/geom/projective_ops.py:96-125,38-50): X0 = iproj(disps[ii]) with the intrinsics of frame ii, Gij = Tj Ti^-1
// (stereo edge ii == jj: t = (-0.1, 0, 0), identity rotation), X1 = Gij X0, pinhole projection with the
// intrinsics of frame jj where Z < 0.1 is replaced by 1, valid = X1.z > 0.2 (MIN_DEPTH of geom, not the 0.25
// of the CUDA backend).  The reference builds this from ~10 elementwise torch ops over poses repeated per
// pixel; here it is one pass: 4 B read and 12 B written per pixel-edge.
__global__ void __launch_bounds__(kThreads)
k_reproject(const float* __restrict__ poses, const float* __restrict__ disps, const float* __restrict__ intr,
            int intr_stride, const int64_t* __restrict__ ii, const int64_t* __restrict__ jj, int HW, int wd,
            float2* __restrict__ coords, float* __restrict__ valid) {
  const int e = blockIdx.y;
  const int p = blockIdx.x * kThreads + threadIdx.x;
  if (p >= HW) return;
  const int ix = (int)ii[e], jx = (int)jj[e];
  const float* Ki = intr + (size_t)ix * intr_stride;
  const float* Kj = intr + (size_t)jx * intr_stride;
  Se3 G;
  if (ix == jx) G = stereo_baseline();
  else G = load_relative(poses, ix, jx);
  const int i = p / wd, j = p - i * wd;
  const float Xi[4] = {((float)j - __ldg(Ki + 2)) / __ldg(Ki), ((float)i - __ldg(Ki + 3)) / __ldg(Ki + 1), 1.0f,
                       __ldg(disps + (size_t)ix * HW + p)};
  float Xj[4];
  act(G, Xi, Xj);
  const float z = Xj[2] < 0.1f ? 1.0f : Xj[2];
  const float d = 1.0f / z;
  coords[(size_t)e * HW + p] = make_float2(__ldg(Kj) * (Xj[0] * d) + __ldg(Kj + 2), __ldg(Kj + 1) * (Xj[1] * d) + __ldg(Kj + 3));
  valid[(size_t)e * HW + p] = Xj[2] > 0.2f ? 1.0f : 0.0f;
}

// frame_distance_kernel, src/droid_kernels.cu:562-702.  One CTA per pair; three block sums.
__global__ void __launch_bounds__(kThreads)
k_frame_distance(const float* __restrict__ poses, const float* __restrict__ disps,
                 const float* __restrict__ intr, const int64_t* __restrict__ ii,
                 const int64_t* __restrict__ jj, int HW, int wd, float beta,
                 float* __restrict__ dist) {
  const int e = blockIdx.x;
  const int ix = (int)ii[e], jx = (int)jj[e];
  const float fx = __ldg(intr), fy = __ldg(intr + 1), cx = __ldg(intr + 2), cy = __ldg(intr + 3);
  const Se3 G = load_relative(poses, ix, jx);
  float accum = 0.f, valid = 0.f, total = 0.f;
  for (int p = threadIdx.x; p < HW; p += kThreads) {
    const int i = p / wd, j = p - i * wd;
    const float u = (float)j, v = (float)i;
    const float Xi[4] = {(u - cx) / fx, (v - cy) / fy, 1.0f, __ldg(disps + (size_t)ix * HW + p)};
    float Xj[4];
    act(G, Xi, Xj);
    float du = fx * (Xj[0] / Xj[2]) + cx - u;
    float dv = fy * (Xj[1] / Xj[2]) + cy - v;
    float d = sqrtf(du * du + dv * dv);
    total += beta;
    if (Xj[2] > kMinDepth) { accum += beta * d; valid += beta; }
    // translation-only flow
    Xj[0] = Xi[0] + Xi[3] * G.t[0];
    Xj[1] = Xi[1] + Xi[3] * G.t[1];
    Xj[2] = Xi[2] + Xi[3] * G.t[2];
    du = fx * (Xj[0] / Xj[2]) + cx - u;
    dv = fy * (Xj[1] / Xj[2]) + cy - v;
    d = sqrtf(du * du + dv * dv);
    total += (1.0f - beta);
    if (Xj[2] > kMinDepth) { accum += (1.0f - beta) * d; valid += (1.0f - beta); }
  }
  __shared__ float red[3][kThreads / 32];
  accum = warp_sum(accum);
  valid = warp_sum(valid);
  total = warp_sum(total);
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  if (lane == 0) { red[0][wid] = accum; red[1][wid] = valid; red[2][wid] = total; }
  __syncthreads();
  if (threadIdx.x == 0) {
    float a = 0.f, va = 0.f, t = 0.f;
#pragma unroll
    for (int w = 0; w < kThreads / 32; ++w) { a += red[0][w]; va += red[1][w]; t += red[2][w]; }
    dist[e] = (va / (t + 1e-8f) < 0.75f) ? 1000.0f : a / va;
  }
}

// iproj_kernel, src/droid_kernels.cu:824-895
__global__ void __launch_bounds__(kThreads)
k_iproj(const float* __restrict__ poses, const float* __restrict__ disps,
        const float* __restrict__ intr, int HW, int wd, float* __restrict__ points) {
  const int f = blockIdx.y;
  const int p = blockIdx.x * kThreads + threadIdx.x;
  if (p >= HW) return;
  const float fx = __ldg(intr), fy = __ldg(intr + 1), cx = __ldg(intr + 2), cy = __ldg(intr + 3);
  Se3 T;
#pragma unroll
  for (int k = 0; k < 3; ++k) T.t[k] = __ldg(poses + 7 * f + k);
#pragma unroll
  for (int k = 0; k < 4; ++k) T.q[k] = __ldg(poses + 7 * f + 3 + k);
  const int i = p / wd, j = p - i * wd;
  const float Xi[4] = {((float)j - cx) / fx, ((float)i - cy) / fy, 1.0f, __ldg(disps + (size_t)f * HW + p)};
  float Xj[4];
  act(T, Xi, Xj);
  float* o = points + ((size_t)f * HW + p) * 3;
  o[0] = Xj[0] / Xj[3];
  o[1] = Xj[1] / Xj[3];
  o[2] = Xj[2] / Xj[3];
}

// depth_filter_kernel, src/droid_kernels.cu:706-820.  The reference launches one CTA per
// (frame, neighbour, tile) and counts with atomicAdd; here one thread owns a pixel and walks
// the six neighbours (ix-1..ix-3, ix+3..ix+5) itself, so the count is a plain store.
__global__ void __launch_bounds__(kThreads)
k_depth_filter(const float* __restrict__ poses, const float* __restrict__ disps,
               const float* __restrict__ intr, const int64_t* __restrict__ inds,
               const float* __restrict__ thresh, int num, int ht, int wd,
               float* __restrict__ counter) {
  const int b = blockIdx.y;
  const int HW = ht * wd;
  const int p = blockIdx.x * kThreads + threadIdx.x;
  if (p >= HW) return;
  const int ix = (int)inds[b];
  const float t = __ldg(thresh + b);
  const float fx = __ldg(intr), fy = __ldg(intr + 1), cx = __ldg(intr + 2), cy = __ldg(intr + 3);
  const int i = p / wd, j = p - i * wd;
  const float di = __ldg(disps + (size_t)ix * HW + p);
  const float Xi[4] = {((float)j - cx) / fx, ((float)i - cy) / fy, 1.0f, di};
  float count = 0.f;
#pragma unroll 1
  for (int nb = 0; nb < 6; ++nb) {
    const int jx = nb < 3 ? ix - nb - 1 : ix + nb;
    if (jx < 0 || jx >= num) continue;
    const Se3 G = load_relative(poses, ix, jx);
    float Xj[4];
    act(G, Xi, Xj);
    const float uj = fx * (Xj[0] / Xj[2]) + cx;
    const float vj = fy * (Xj[1] / Xj[2]) + cy;
    const float dj = Xj[3] / Xj[2];
    const float fu = floorf(uj), fv = floorf(vj);
    // compare as floats first so NaN / huge values never reach the int conversion
    if (!(fu >= 0.f && fv >= 0.f && fu < (float)(wd - 1) && fv < (float)(ht - 1))) continue;
    const int u0 = (int)fu, v0 = (int)fv;
    const float* D = disps + (size_t)jx * HW;
    const float d00 = __ldg(D + v0 * wd + u0), d01 = __ldg(D + v0 * wd + u0 + 1);
    const float d10 = __ldg(D + (v0 + 1) * wd + u0), d11 = __ldg(D + (v0 + 1) * wd + u0 + 1);
    // the reference evaluates these differences in double (1.0 / dj with a double literal)
    const double inv = 1.0 / (double)dj, td = (double)t;
    if (fabs(inv - 1.0 / (double)d00) < td || fabs(inv - 1.0 / (double)d01) < td ||
        fabs(inv - 1.0 / (double)d10) < td || fabs(inv - 1.0 / (double)d11) < td)
      count += 1.0f;
  }
  counter[(size_t)b * HW + p] = count;
}

}  // namespace
}  // namespace dbaf

using namespace dbaf;

extern "C" int dbaf_version(void) { return 100; }
extern "C" int dbaf_ba_layout_bytes(void) { return (int)sizeof(dbaf_ba_layout); }

extern "C" const char* dbaf_last_error(void) { return g_error.c_str(); }

extern "C" int dbaf_projmap(const float* poses, const float* disps, const float* intrinsics,
                            const int64_t* ii, const int64_t* jj, int E, int ht, int wd,
                            float* coords, float* valid, void* stream) {
  DBAF_REQUIRE(E >= 0 && ht > 0 && wd > 0, "bad sizes");
  if (E == 0) return DBAF_OK;
  const int HW = ht * wd;
  dim3 grid((HW + kThreads - 1) / kThreads, E);
  k_projmap<<<grid, kThreads, 0, (cudaStream_t)stream>>>(poses, disps, intrinsics, ii, jj, HW, wd, coords, valid);
  DBAF_LAUNCH_CHECK("k_projmap");
  return DBAF_OK;
}

extern "C" int dbaf_reproject(const float* poses, const float* disps, const float* intrinsics, int per_frame_intrinsics,
                              const int64_t* ii, const int64_t* jj, int E, int ht, int wd, float* coords,
                              float* valid, void* stream) {
  DBAF_REQUIRE(E >= 0 && ht > 0 && wd > 0, "bad sizes");
  DBAF_REQUIRE((reinterpret_cast<uintptr_t>(coords) & 7) == 0, "coords must be 8-byte aligned");
  if (E == 0) return DBAF_OK;
  const int HW = ht * wd;
  dim3 grid((HW + kThreads - 1) / kThreads, E);
  k_reproject<<<grid, kThreads, 0, (cudaStream_t)stream>>>(poses, disps, intrinsics, per_frame_intrinsics ? 4 : 0, ii, jj, HW,
                                                          wd, reinterpret_cast<float2*>(coords), valid);
  DBAF_LAUNCH_CHECK("k_reproject");
  return DBAF_OK;
}

extern "C" int dbaf_frame_distance(const float* poses, const float* disps, const float* intrinsics,
                                   const int64_t* ii, const int64_t* jj, int M, int ht, int wd,
                                   float beta, float* dist, void* stream) {
  DBAF_REQUIRE(M >= 0 && ht > 0 && wd > 0, "bad sizes");
  if (M == 0) return DBAF_OK;
  k_frame_distance<<<M, kThreads, 0, (cudaStream_t)stream>>>(poses, disps, intrinsics, ii, jj, ht * wd, wd, beta, dist);
  DBAF_LAUNCH_CHECK("k_frame_distance");
  return DBAF_OK;
}

extern "C" int dbaf_iproj(const float* poses, const float* disps, const float* intrinsics, int N,
                          int ht, int wd, float* points, void* stream) {
  DBAF_REQUIRE(N >= 0 && ht > 0 && wd > 0, "bad sizes");
  if (N == 0) return DBAF_OK;
  const int HW = ht * wd;
  dim3 grid((HW + kThreads - 1) / kThreads, N);
  k_iproj<<<grid, kThreads, 0, (cudaStream_t)stream>>>(poses, disps, intrinsics, HW, wd, points);
  DBAF_LAUNCH_CHECK("k_iproj");
  return DBAF_OK;
}

extern "C" int dbaf_depth_filter(const float* poses, const float* disps, const float* intrinsics,
                                 const int64_t* ix, const float* thresh, int M, int N, int ht, int wd,
                                 float* counter, void* stream) {
  DBAF_REQUIRE(M >= 0 && N > 0 && ht > 0 && wd > 0, "bad sizes");
  if (M == 0) return DBAF_OK;
  const int HW = ht * wd;
  dim3 grid((HW + kThreads - 1) / kThreads, M);
  k_depth_filter<<<grid, kThreads, 0, (cudaStream_t)stream>>>(poses, disps, intrinsics, ix, thresh, N, ht, wd, counter);
  DBAF_LAUNCH_CHECK("k_depth_filter");
  return DBAF_OK;
}

// ------------------------------------------------------------------ probes ----
namespace dbaf {
namespace {
// write-only stream of 128-bit stores, one 512-byte line group per warp instruction
__global__ void __launch_bounds__(256) k_fill128(uint4* __restrict__ dst, size_t n16, uint4 val) {
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += stride) dst[i] = val;
}
// the store pattern of the correlation-volume kernel: a CTA owns 128 rows of `pitch` bytes and writes them in chunks
// of `seg` bytes per row (row after row inside a chunk), chunk after chunk
__global__ void __launch_bounds__(256) k_fill_tiled(uint4* __restrict__ dst, size_t n_tiles, int pitch16, int seg16, uint4 val) {
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  for (size_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    uint4* base = dst + tile * 128 * (size_t)pitch16;
    for (int c0 = 0; c0 < pitch16; c0 += seg16)
      for (int r = wid; r < 128; r += 8)
        for (int q = lane; q < seg16; q += 32) base[(size_t)r * pitch16 + c0 + q] = val;
  }
}
}  // namespace
}  // namespace dbaf

extern "C" int dbaf_probe_fill(void* dst, size_t bytes, int kind, void* stream) {
  DBAF_REQUIRE(dst != nullptr && (reinterpret_cast<uintptr_t>(dst) & 15) == 0, "dst must be 16-byte aligned");
  DBAF_REQUIRE(kind >= 0 && kind <= 5, "kind: 0 = cudaMemsetAsync, 1 = 128-bit store kernel, 2..5 = tiled store patterns");
  cudaStream_t st = (cudaStream_t)stream;
  if (kind == 0) {
    DBAF_CUDA(cudaMemsetAsync(dst, 1, bytes, st));
  } else if (kind == 1) {
    k_fill128<<<148 * 8, 256, 0, st>>>(static_cast<uint4*>(dst), bytes / 16, make_uint4(1u, 2u, 3u, 4u));
    DBAF_LAUNCH_CHECK("k_fill128");
  } else {
    // 128 rows x 8 KB tiles written in 512 B (2), 1 KB (3), 2 KB (4) or 8 KB (5) pieces per row
    const int pitch16 = 8192 / 16, seg16 = (kind == 2 ? 512 : kind == 3 ? 1024 : kind == 4 ? 2048 : 8192) / 16;
    k_fill_tiled<<<148 * 2, 256, 0, st>>>(static_cast<uint4*>(dst), bytes / (128 * 8192), pitch16, seg16, make_uint4(1u, 2u, 3u, 4u));
    DBAF_LAUNCH_CHECK("k_fill_tiled");
  }
  return DBAF_OK;
}

// ------------------------------------------------------------------ options ----
extern "C" int dbaf_set_option(const char* key, const char* value) {
  DBAF_REQUIRE(key != nullptr && value != nullptr, "null key or value");
  Options& o = options();
  const std::string k(key), v(value);
  if (k == "schur_impl") {
    DBAF_REQUIRE(v == "auto" || v == "simt", "schur_impl: auto | simt");
    o.schur_simt = v == "simt";
  } else if (k == "band_impl") {
    DBAF_REQUIRE(v == "auto" || v == "single" || v == "twisted", "band_impl: auto | single | twisted");
    o.band_single = v == "single";
    o.band_twisted = v == "twisted";
  } else if (k == "band_sweep") {
    DBAF_REQUIRE(v == "auto" || v == "smem", "band_sweep: auto | smem");
    o.band_smem = v == "smem";
  } else if (k == "band_domains") {
    o.band_domains = std::atoi(value);
  } else if (k == "volume_impl") {
    DBAF_REQUIRE(v == "auto" || v == "simt", "volume_impl: auto | simt");
    o.volume_simt = v == "simt";
  } else if (k == "eij") {
    DBAF_REQUIRE(v == "auto" || v == "store" || v == "recompute", "eij: auto | store | recompute");
    o.eij_recompute = v == "auto" ? -1 : v == "recompute";
  } else {
    DBAF_REQUIRE(false, "unknown option '" + k + "'");
  }
  return DBAF_OK;
}
