// All-pairs correlation volume + 4-level pyramid (CorrBlock.__init__,
// dbaf/modules/corr.py:24-38,63-71).
//
//   L0[e, p1, p2] = rn16( sum_c f1[e,c,p1] * f2[e,c,p2] / 16 )        (fp32 accumulate)
//   L(l+1) = rn16( ((a+b)+c)+d ) / 4 ) over 2x2 blocks of the fp16 L(l)  (floor sizes)
//
// This file holds the entry point, the pyramid kernel for arbitrary shapes and a plain SIMT
// GEMM that serves as the in-library cross-check of the tcgen05 kernel (volume_tc.cu).
#include "common.cuh"

#include <cstdlib>
#include <cstring>

namespace dbaf {

// volume_tc.cu
int corr_volume_tc(const __half* f1, int n1, const __half* f2, int n2, const int* idx1, const int* idx2, int n, int C,
                   int ht, int wd, int levels, __half* const* out, void* workspace, size_t workspace_bytes,
                   cudaStream_t st, int* levels_done, const int* out_slot, int n_out, const int* bad_flag);
size_t corr_volume_tc_workspace(int n1, int n2, int C, int ht, int wd);

namespace {

constexpr int kTile = 64;

// 64x64 output tile per CTA, 256 threads x (4x4), K staged 32 channels at a time.
__global__ void __launch_bounds__(256)
k_volume_simt(const __half* __restrict__ f1, const __half* __restrict__ f2, const int* __restrict__ idx1,
              const int* __restrict__ idx2, int C, int HW, __half* __restrict__ out, const int* __restrict__ out_slot,
              const int* __restrict__ bad_flag) {
  __shared__ float As[32][kTile + 1];
  __shared__ float Bs[32][kTile + 1];
  if (bad_flag && *bad_flag) return;                 // an index outside its stack: nothing is read or written
  const int e = blockIdx.z;
  const int m0 = blockIdx.y * kTile, n0 = blockIdx.x * kTile;
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  const __half* A = f1 + (size_t)(idx1 ? idx1[e] : e) * C * HW;
  const __half* B = f2 + (size_t)(idx2 ? idx2[e] : e) * C * HW;
  float acc[4][4] = {};
  for (int c0 = 0; c0 < C; c0 += 32) {
    for (int t = threadIdx.x; t < 32 * kTile; t += 256) {
      const int c = t / kTile, x = t - c * kTile;
      const bool cok = c0 + c < C;
      As[c][x] = (cok && m0 + x < HW) ? __half2float(__float2half_rn(__half2float(A[(size_t)(c0 + c) * HW + m0 + x]) * 0.25f)) : 0.f;
      Bs[c][x] = (cok && n0 + x < HW) ? __half2float(__float2half_rn(__half2float(B[(size_t)(c0 + c) * HW + n0 + x]) * 0.25f)) : 0.f;
    }
    __syncthreads();
#pragma unroll 8
    for (int c = 0; c < 32; ++c) {
      float a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) { a[i] = As[c][ty * 4 + i]; b[i] = Bs[c][tx * 4 + i]; }
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int m = m0 + ty * 4 + i;
    if (m >= HW) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int nn = n0 + tx * 4 + j;
      if (nn < HW) out[((size_t)(out_slot ? out_slot[e] : e) * HW + m) * HW + nn] = __float2half_rn(acc[i][j]);   // operands were divided by 4 in fp16, corr.py:67-68
    }
  }
}

__device__ __forceinline__ __half pool4(const __half* s, int w, int y, int x) {
  const float a = __half2float(s[(2 * y) * w + 2 * x]), b = __half2float(s[(2 * y) * w + 2 * x + 1]);
  const float c = __half2float(s[(2 * y + 1) * w + 2 * x]), d = __half2float(s[(2 * y + 1) * w + 2 * x + 1]);
  return __float2half_rn((((a + b) + c) + d) * 0.25f);
}

// Slabs (source pixels) are pooled by groups of `gthreads` threads that never meet the rest of the CTA: a whole
// 128-thread CTA per slab when the slab is a level-0 map (thousands of values), one warp per slab when it is a
// level-1 map of a few hundred values (then four slabs are in flight per CTA and the barriers are warp barriers).
// The slab and every pooled level live in the group's part of shared memory.
__device__ __forceinline__ void group_sync(int gthreads) {
  if (gthreads == 32) __syncwarp();
  else __syncthreads();
}

__global__ void __launch_bounds__(128)
k_pyramid(const __half* __restrict__ L0, long long slabs, int h2, int w2, int levels,
          __half* __restrict__ L1, __half* __restrict__ L2, __half* __restrict__ L3, int gthreads, int group_halves,
          const int* __restrict__ slot, int hw_src, long long total_slabs, const int* __restrict__ bad_flag) {
  if (bad_flag && *bad_flag) return;
  // slot != null: logical slab s = (edge, source pixel) lives at physical slab slot[edge] * hw_src + pixel of buffers
  // that hold total_slabs slabs
  extern __shared__ __half sm[];
  const int s0 = h2 * w2;
  const int h1 = h2 >> 1, w1 = w2 >> 1, hq = h2 >> 2, wq = w2 >> 2, he = h2 >> 3, we = w2 >> 3;
  const int groups = 128 / gthreads, grp = threadIdx.x / gthreads, gt = threadIdx.x - grp * gthreads;
  // the slab is fetched with aligned 16-byte loads whatever its own alignment (h2*w2 is odd for 55 x 55): the
  // raw copy starts up to 7 halves before the slab, `a0` points at the slab inside it
  __half* raw = sm + (size_t)grp * group_halves;
  __half* a1 = raw + ((s0 + 15) & ~7);
  __half* a2 = a1 + h1 * w1;
  const bool vec = (reinterpret_cast<size_t>(L0) & 15) == 0;
  for (long long sl = (long long)blockIdx.x * groups + grp; sl < slabs; sl += (long long)gridDim.x * groups) {
    const long long s = slot ? (long long)slot[sl / hw_src] * hw_src + sl % hw_src : sl;
    const long long first = s * s0;
    const int off = vec ? (int)(first & 7) : 0;
    const __half* a0 = raw + off;
    if (vec) {
      const uint4* src = reinterpret_cast<const uint4*>(L0 + (first - off));
      const int nvec = (off + s0 + 7) >> 3;
      // the last vector may reach past the end of the whole volume: it is then read element-wise
      const bool tail_ok = first - off + (long long)nvec * 8 <= total_slabs * (long long)s0;
      for (int t = gt; t < nvec - (tail_ok ? 0 : 1); t += gthreads) reinterpret_cast<uint4*>(raw)[t] = __ldg(src + t);
      if (!tail_ok)
        for (int t = (nvec - 1) * 8 + gt; t < off + s0; t += gthreads) raw[t] = L0[first - off + t];
    } else {
      for (int t = gt; t < s0; t += gthreads) raw[t] = L0[first + t];
    }
    group_sync(gthreads);
    for (int t = gt; t < h1 * w1; t += gthreads) {
      const __half v = pool4(a0, w2, t / w1, t % w1);
      a1[t] = v;
      L1[s * (h1 * w1) + t] = v;
    }
    group_sync(gthreads);
    if (levels > 2) {
      for (int t = gt; t < hq * wq; t += gthreads) {
        const __half v = pool4(a1, w1, t / wq, t % wq);
        a2[t] = v;
        L2[s * (hq * wq) + t] = v;
      }
      group_sync(gthreads);
      if (levels > 3)
        for (int t = gt; t < he * we; t += gthreads) L3[s * (he * we) + t] = pool4(a2, wq, t / we, t % we);
    }
    group_sync(gthreads);
  }
}

}  // namespace
}  // namespace dbaf

using namespace dbaf;

namespace dbaf {
namespace {

__global__ void k_check_index(const int* __restrict__ idx, int n, int limit, int* __restrict__ bad) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t < n && (idx[t] < 0 || idx[t] >= limit)) atomicExch(bad, 1);
}

int volume_pyramid(const __half* f1, int n1, const __half* f2, int n2, const int* idx1, const int* idx2, int n, int C,
                   int ht, int wd, int levels, void* const* out_host, void* workspace, size_t workspace_bytes,
                   cudaStream_t st, const int* out_slot = nullptr, int n_out = -1, const int* bad_flag = nullptr) {
  const int HW = ht * wd;
  if (n_out < 0) n_out = n;
  __half* outs[4] = {nullptr, nullptr, nullptr, nullptr};
  for (int l = 0; l < levels; ++l) outs[l] = static_cast<__half*>(out_host[l]);
  // option volume_impl=simt forces the cross-check kernels (tests use it to validate the
  // tensor-core path against an independent implementation inside the same library).
  const bool force_simt = options().volume_simt != 0;
  int done = 0;
  if (!force_simt) {
    const int rc = corr_volume_tc(f1, n1, f2, n2, idx1, idx2, n, C, ht, wd, levels, outs, workspace, workspace_bytes, st,
                                  &done, out_slot, n_out, bad_flag);
    if (rc != DBAF_OK) return rc;
    if (done >= levels) return DBAF_OK;
  }
  if (done == 0) {
    dim3 grid((HW + kTile - 1) / kTile, (HW + kTile - 1) / kTile, n);
    k_volume_simt<<<grid, 256, 0, st>>>(f1, f2, idx1, idx2, C, HW, outs[0], out_slot, bad_flag);
    DBAF_LAUNCH_CHECK("k_volume_simt");
  }
  if (levels > std::max(done, 1)) {
    // remaining levels from the deepest one already written: level 0 (SIMT / 256-pixel chunks) or level 1
    const int src = std::max(done, 1) - 1;
    const int hs = ht >> src, wsz = wd >> src, HWs = hs * wsz;
    const int gthreads = HWs <= 1024 ? 32 : 128;      // small (level-1) slabs: one warp each
    const int group_halves = ((((HWs + 15) & ~7) + (hs / 2) * (wsz / 2) + (hs / 4) * (wsz / 4)) + 7) & ~7;   // 16-byte multiple
    const size_t smem = (size_t)group_halves * (128 / gthreads) * 2 + 16;
    DBAF_REQUIRE(smem <= 200 * 1024, "feature map too large for the pyramid kernel");
    static bool attr[64] = {};                        // per device
    int dev_id = 0;
    DBAF_CUDA(cudaGetDevice(&dev_id));
    if (!attr[dev_id & 63]) {
      DBAF_CUDA(cudaFuncSetAttribute(k_pyramid, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
      attr[dev_id & 63] = true;
    }
    const long long slabs = (long long)n * HW;
    const int blocks = (int)std::min<long long>((slabs + 128 / gthreads - 1) / (128 / gthreads), 148LL * 16);
    k_pyramid<<<blocks, 128, smem, st>>>(outs[src], slabs, hs, wsz, levels - src, outs[src + 1], src + 2 < 4 ? outs[src + 2] : nullptr,
                                         src + 3 < 4 ? outs[src + 3] : nullptr, gthreads, group_halves, out_slot, HW,
                                         (long long)n_out * HW, bad_flag);
    DBAF_LAUNCH_CHECK("k_pyramid");
  }
  return DBAF_OK;
}

}  // namespace
}  // namespace dbaf

extern "C" size_t dbaf_corr_volume_workspace_bytes(int n, int C, int ht, int wd) {
  return corr_volume_tc_workspace(n, n, C, ht, wd);
}

extern "C" int dbaf_corr_volume_pyramid(const void* fmap1, const void* fmap2, int n, int C, int ht, int wd,
                                        int levels, void* const* out_host, void* workspace,
                                        size_t workspace_bytes, void* stream) {
  DBAF_REQUIRE(n >= 0 && C > 0 && ht > 0 && wd > 0, "bad sizes");
  DBAF_REQUIRE(levels >= 1 && levels <= 4, "levels must be 1..4");
  DBAF_REQUIRE(out_host != nullptr, "null output list");
  if (n == 0) return DBAF_OK;
  return volume_pyramid(static_cast<const __half*>(fmap1), n, static_cast<const __half*>(fmap2), n, nullptr, nullptr, n,
                        C, ht, wd, levels, out_host, workspace, workspace_bytes, (cudaStream_t)stream);
}

extern "C" size_t dbaf_corr_volume_indexed_workspace_bytes(int n1, int n2, int C, int ht, int wd) {
  return corr_volume_tc_workspace(n1, n2, C, ht, wd) + 256;   // + the index check flag
}

extern "C" int dbaf_corr_volume_pyramid_indexed(const void* fmaps1, int n1, const void* fmaps2, int n2,
                                                const int32_t* ii, const int32_t* jj, int n_edges, int C, int ht,
                                                int wd, int levels, void* const* out_host, void* workspace,
                                                size_t workspace_bytes, void* stream) {
  DBAF_REQUIRE(n1 > 0 && n2 > 0 && n_edges >= 0 && C > 0 && ht > 0 && wd > 0, "bad sizes");
  DBAF_REQUIRE(levels >= 1 && levels <= 4, "levels must be 1..4");
  DBAF_REQUIRE(out_host != nullptr && ii != nullptr && jj != nullptr, "null output list or index array");
  if (n_edges == 0) return DBAF_OK;
  cudaStream_t st = (cudaStream_t)stream;
  // An index outside its frame stack would read foreign memory.  It is checked on the device before anything is
  // computed: the check raises a flag in the last 256 bytes of the workspace and every later kernel of this call
  // returns at once when it is set; nothing synchronises here, dbaf_corr_volume_status() reads the flag.
  DBAF_REQUIRE(workspace != nullptr && workspace_bytes >= 256, "workspace too small (dbaf_corr_volume_indexed_workspace_bytes)");
  int* flag = reinterpret_cast<int*>((reinterpret_cast<uintptr_t>(workspace) + workspace_bytes - 64) & ~(uintptr_t)63);
  DBAF_CUDA(cudaMemsetAsync(flag, 0, sizeof(int), st));
  k_check_index<<<(n_edges + 255) / 256, 256, 0, st>>>(ii, n_edges, n1, flag);
  k_check_index<<<(n_edges + 255) / 256, 256, 0, st>>>(jj, n_edges, n2, flag);
  DBAF_LAUNCH_CHECK("k_check_index");
  return volume_pyramid(static_cast<const __half*>(fmaps1), n1, static_cast<const __half*>(fmaps2), n2, ii, jj, n_edges,
                        C, ht, wd, levels, out_host, workspace, workspace_bytes - 256, st, nullptr, -1, flag);
}

extern "C" int dbaf_corr_volume_status(const void* workspace, size_t workspace_bytes, int32_t* status_host, void* stream) {
  DBAF_REQUIRE(workspace != nullptr && workspace_bytes >= 256 && status_host != nullptr, "null pointer or workspace too small");
  const int* flag = reinterpret_cast<const int*>((reinterpret_cast<uintptr_t>(workspace) + workspace_bytes - 64) & ~(uintptr_t)63);
  DBAF_CUDA(cudaMemcpyAsync(status_host, flag, sizeof(int32_t), cudaMemcpyDeviceToHost, (cudaStream_t)stream));
  return DBAF_OK;
}

// ---------------------------------------------------------------- slot pool (CorrBlock bookkeeping) ----
extern "C" int dbaf_corr_volume_pyramid_pool(const void* fmaps1, int n1, const void* fmaps2, int n2,
                                             const int32_t* ii, const int32_t* jj, const int32_t* slots, int n_edges,
                                             int capacity, int C, int ht, int wd, int levels, void* const* pool_host,
                                             void* workspace, size_t workspace_bytes, void* stream) {
  DBAF_REQUIRE(n1 > 0 && n2 > 0 && n_edges >= 0 && capacity > 0 && C > 0 && ht > 0 && wd > 0, "bad sizes");
  DBAF_REQUIRE(levels >= 1 && levels <= 4, "levels must be 1..4");
  DBAF_REQUIRE(pool_host != nullptr && ii != nullptr && jj != nullptr && slots != nullptr, "null pointer");
  if (n_edges == 0) return DBAF_OK;
  cudaStream_t st = (cudaStream_t)stream;
  DBAF_REQUIRE(workspace != nullptr && workspace_bytes >= 256, "workspace too small");
  int* flag = reinterpret_cast<int*>((reinterpret_cast<uintptr_t>(workspace) + workspace_bytes - 64) & ~(uintptr_t)63);
  DBAF_CUDA(cudaMemsetAsync(flag, 0, sizeof(int), st));
  k_check_index<<<(n_edges + 255) / 256, 256, 0, st>>>(ii, n_edges, n1, flag);
  k_check_index<<<(n_edges + 255) / 256, 256, 0, st>>>(jj, n_edges, n2, flag);
  k_check_index<<<(n_edges + 255) / 256, 256, 0, st>>>(slots, n_edges, capacity, flag);
  DBAF_LAUNCH_CHECK("k_check_index");
  return volume_pyramid(static_cast<const __half*>(fmaps1), n1, static_cast<const __half*>(fmaps2), n2, ii, jj, n_edges,
                        C, ht, wd, levels, pool_host, workspace, workspace_bytes - 256, st, slots, capacity, flag);
}
