// Frame-graph edge selection on the device: the proximity / non-maximum-suppression sweep
// of CovisibleGraph.add_proximity_factors (dbaf/covisible_graph.py:357-441) and the duplicate filter of
// CovisibleGraph.__filter_repeated_edges (:61-72).  The reference runs both as Python loops with one `.item()`
// (a device synchronisation) per candidate, per suppression stamp and per edge; here each is one small kernel
// and the host reads a single count at the end.  Integer work: results are bit-exact with the reference.
#include "common.cuh"

#include <cmath>

namespace dbaf {
namespace {

constexpr int kGraphThreads = 256;

struct ProxParams {
  int n_cand, cc, n_ex, t0, t1, t, rad, nms, stereo, max_factors, es_cap;
  float thresh;
};

// d[(i1-t0) W + (j1-t1)] = inf for the diamond |di| + |dj| <= max(min(|i-j|-2, nms), 0) around (i, j)  (:386-394)
__device__ __forceinline__ void stamp_one(float* d, const ProxParams& p, int i, int j, int q) {
  const int side = 2 * p.nms + 1;
  const int di = q / side - p.nms, dj = q % side - p.nms;
  const int lim = max(min(abs(i - j) - 2, p.nms), 0);
  if (abs(di) + abs(dj) > lim) return;
  const int i1 = i + di, j1 = j + dj;
  if (i1 >= p.t0 && i1 < p.t && j1 >= p.t1 && j1 < p.t) d[(i1 - p.t0) * (p.t - p.t1) + (j1 - p.t1)] = INFINITY;
}

// number of edges the unconditional part of the list (:396-406) holds for source frames t0 .. i-1
__device__ __host__ inline int base_count_before(int i, int t0, int rad, int stereo) {
  int n = 0;
  for (int a = t0; a < i; ++a) n += (stereo ? 1 : 0) + 2 * (a - max(a - rad - 1, 0));
  return n;
}

__global__ void __launch_bounds__(kGraphThreads)
k_proximity_select(const float* __restrict__ d1, const float* __restrict__ d2, const int64_t* __restrict__ cand_ii,
                   const int64_t* __restrict__ cand_jj, const int64_t* __restrict__ ex_ii,
                   const int64_t* __restrict__ ex_jj, ProxParams p, float* __restrict__ d, int64_t* __restrict__ es_ii,
                   int64_t* __restrict__ es_jj, int* __restrict__ n_es_out) {
  const int tid = threadIdx.x, nt = blockDim.x, lane = tid & 31, wid = tid >> 5;
  const int W = p.t - p.t1;
  __shared__ float s_val[kGraphThreads / 32];
  __shared__ int s_idx[kGraphThreads / 32];
  __shared__ int s_pick, s_n;

  // distances (bidirectional mean like DepthVideo.distance, dbaf/depth_video.py:255-261) and the static masks (:379-381)
  for (int k = tid; k < p.n_cand; k += nt) {
    float v = d2 ? 0.5f * (d1[k] + d2[k]) : d1[k];
    if (cand_ii[k] - p.rad < cand_jj[k]) v = INFINITY;
    if (v > 100.0f) v = INFINITY;
    d[k] = v;
  }
  __syncthreads();
  // suppression around the edges the graph already holds (:383-394)
  const int side2 = (2 * p.nms + 1) * (2 * p.nms + 1);
  for (int w = tid; w < p.n_ex * side2; w += nt) stamp_one(d, p, (int)ex_ii[w / side2], (int)ex_jj[w / side2], w % side2);
  __syncthreads();
  // unconditional edges: stereo self edge and the temporal neighbours within rad + 1 (:396-406)
  for (int i = p.t0 + tid; i < p.t; i += nt) {
    int o = base_count_before(i, p.t0, p.rad, p.stereo);
    if (p.stereo) {
      es_ii[o] = i;
      es_jj[o] = i;
      ++o;
      int idx = (i - p.t0) * W + (i - p.t1);
      if (idx < 0) idx += p.n_cand;                      // the reference indexes a tensor: negative wraps
      if (idx >= 0 && idx < p.n_cand) d[idx] = INFINITY;
    }
    for (int j = max(i - p.rad - 1, 0); j < i; ++j) {
      es_ii[o] = i; es_jj[o] = j; ++o;
      es_ii[o] = j; es_jj[o] = i; ++o;
      const int idx = (i - p.t0) * W + (j - p.t1);
      if (idx >= 0 && idx < p.n_cand) d[idx] = INFINITY;
    }
  }
  if (tid == 0) s_n = base_count_before(p.t, p.t0, p.rad, p.stereo);
  __syncthreads();
  // greedy sweep in ascending distance (:408-433): the smallest live entry of the grid is taken, its neighbourhood
  // suppressed, until nothing at or below the threshold is left or the list is longer than max_factors
  for (;;) {
    float bv = INFINITY;
    int bi = -1;
    for (int k = tid; k < p.cc; k += nt) {
      const float v = d[k];
      if (v < bv) { bv = v; bi = k; }                    // strictly smaller: the lowest index wins a tie; NaN never wins
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
      const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
      if (oi >= 0 && (bi < 0 || ov < bv || (ov == bv && oi < bi))) { bv = ov; bi = oi; }
    }
    if (lane == 0) { s_val[wid] = bv; s_idx[wid] = bi; }
    __syncthreads();
    if (tid == 0) {
      for (int w = 1; w < nt / 32; ++w) {
        const float ov = s_val[w];
        const int oi = s_idx[w];
        if (oi >= 0 && (bi < 0 || ov < bv || (ov == bv && oi < bi))) { bv = ov; bi = oi; }
      }
      int pick = -1;
      if (bi >= 0 && !(bv > p.thresh) && !(s_n > p.max_factors)) pick = bi;
      s_pick = pick;
      if (pick >= 0) {
        const int i = (int)cand_ii[pick], j = (int)cand_jj[pick];
        if (s_n + 2 <= p.es_cap) {
          es_ii[s_n] = i; es_jj[s_n] = j;
          es_ii[s_n + 1] = j; es_jj[s_n + 1] = i;
        }
        s_n += 2;
      }
    }
    __syncthreads();
    const int pick = s_pick;
    if (pick < 0) break;
    if (tid < side2) stamp_one(d, p, (int)cand_ii[pick], (int)cand_jj[pick], tid);
    if (tid == 0) d[pick] = INFINITY;                    // (covered by the stamp's centre; kept for nms = 0 off-grid cases)
    __syncthreads();
  }
  if (tid == 0) {
    // NaN distances sort behind everything in the reference's argsort and pass its `d > thresh` test: they are taken
    // last, in index order, under the same length limit
    for (int k = 0; k < p.cc && !(s_n > p.max_factors); ++k) {
      if (d[k] != d[k]) {
        const int i = (int)cand_ii[k], j = (int)cand_jj[k];
        if (s_n + 2 <= p.es_cap) {
          es_ii[s_n] = i; es_jj[s_n] = j;
          es_ii[s_n + 1] = j; es_jj[s_n + 1] = i;
        }
        s_n += 2;
        for (int q = 0; q < side2; ++q) stamp_one(d, p, i, j, q);
      }
    }
    // the closest "skip" edge, if any is closer than the threshold (:435-439)
    if (p.n_cand > p.cc) {
      int b = -1;
      float v = 0.f;
      for (int k = p.cc; k < p.n_cand; ++k) {            // first minimum, NaN last (the order of a stable argsort)
        const float x = d[k];
        if (b < 0) { b = k; v = x; continue; }
        const bool xn = x != x, vn = v != v;
        if ((vn && !xn) || (!xn && !vn && x < v)) { b = k; v = x; }
      }
      if (b >= 0 && v < p.thresh && v > 0.0f) {
        const int i = (int)cand_ii[b], j = (int)cand_jj[b];
        if (s_n + 2 <= p.es_cap) {
          es_ii[s_n] = i; es_jj[s_n] = j;
          es_ii[s_n + 1] = j; es_jj[s_n + 1] = i;
        }
        s_n += 2;
      }
    }
    *n_es_out = s_n;
  }
}

// keep[k] = (ii[k], jj[k]) not among the existing edges; survivors compacted in order (:61-72)
__global__ void __launch_bounds__(kGraphThreads)
k_filter_edges(const int64_t* __restrict__ ii, const int64_t* __restrict__ jj, const int* __restrict__ n_dev, int n_host,
               const int64_t* __restrict__ ex_ii, const int64_t* __restrict__ ex_jj, int n_ex,
               int64_t* __restrict__ out_ii, int64_t* __restrict__ out_jj, int* __restrict__ n_out) {
  const int tid = threadIdx.x, nt = blockDim.x, lane = tid & 31, wid = tid >> 5;
  const int n = n_dev ? min(*n_dev, n_host) : n_host;
  __shared__ int s_warp[kGraphThreads / 32];
  __shared__ int s_base;
  if (tid == 0) s_base = 0;
  __syncthreads();
  for (int base = 0; base < n; base += nt) {
    const int k = base + tid;
    bool keep = false;
    long long a = 0, b = 0;
    if (k < n) {
      a = ii[k];
      b = jj[k];
      keep = true;
      for (int e = 0; e < n_ex; ++e)
        if (ex_ii[e] == a && ex_jj[e] == b) { keep = false; break; }
    }
    const unsigned m = __ballot_sync(0xffffffffu, keep);
    if (lane == 0) s_warp[wid] = __popc(m);
    __syncthreads();
    int off = s_base;
    for (int w = 0; w < wid; ++w) off += s_warp[w];
    if (keep) {
      const int o = off + __popc(m & ((1u << lane) - 1u));
      out_ii[o] = a;
      out_jj[o] = b;
    }
    __syncthreads();
    if (tid == 0) {
      int tot = 0;
      for (int w = 0; w < nt / 32; ++w) tot += s_warp[w];
      s_base += tot;
    }
    __syncthreads();
  }
  if (tid == 0) *n_out = s_base;
}

}  // namespace
}  // namespace dbaf

using namespace dbaf;

extern "C" int dbaf_proximity_base_count(int t0, int t, int rad, int stereo) {
  return base_count_before(t, t0, rad, stereo);
}

extern "C" int dbaf_proximity_select(const float* d1, const float* d2, const int64_t* cand_ii, const int64_t* cand_jj,
                                     int n_cand, int cc, const int64_t* ex_ii, const int64_t* ex_jj, int n_existing,
                                     int t0, int t1, int t, int rad, int nms, float thresh, int stereo, int max_factors,
                                     float* d_work, int64_t* es_ii, int64_t* es_jj, int es_capacity, int32_t* n_es,
                                     void* stream) {
  DBAF_REQUIRE(t0 >= 0 && t1 >= 0 && t >= t0 && t >= t1, "need 0 <= t0, t1 <= t");
  DBAF_REQUIRE(rad >= 0 && nms >= 0 && nms <= 7, "need rad >= 0 and 0 <= nms <= 7");
  DBAF_REQUIRE(cc == (t - t0) * (t - t1) && n_cand >= cc, "cc must be (t-t0)*(t-t1) and n_cand >= cc");
  DBAF_REQUIRE(n_cand == 0 || (d1 && cand_ii && cand_jj && d_work), "null candidate arrays");
  DBAF_REQUIRE(n_existing == 0 || (ex_ii && ex_jj), "null existing-edge arrays");
  DBAF_REQUIRE(es_ii && es_jj && n_es, "null output");
  const int base = base_count_before(t, t0, rad, stereo);
  DBAF_REQUIRE(es_capacity >= base + 2 * cc + 2, "es_capacity must be at least dbaf_proximity_base_count() + 2*cc + 2");
  ProxParams p{n_cand, cc, n_existing, t0, t1, t, rad, nms, stereo ? 1 : 0, max_factors, es_capacity, thresh};
  k_proximity_select<<<1, kGraphThreads, 0, (cudaStream_t)stream>>>(d1, d2, cand_ii, cand_jj, ex_ii, ex_jj, p, d_work, es_ii,
                                                                     es_jj, n_es);
  DBAF_LAUNCH_CHECK("k_proximity_select");
  return DBAF_OK;
}

extern "C" int dbaf_filter_repeated_edges(const int64_t* ii, const int64_t* jj, const int32_t* n_dev, int n_max,
                                          const int64_t* ex_ii, const int64_t* ex_jj, int n_existing, int64_t* out_ii,
                                          int64_t* out_jj, int32_t* n_out, void* stream) {
  DBAF_REQUIRE(n_max >= 0 && n_existing >= 0, "negative size");
  DBAF_REQUIRE(n_out != nullptr && (n_max == 0 || (ii && jj && out_ii && out_jj)), "null pointer");
  DBAF_REQUIRE(n_existing == 0 || (ex_ii && ex_jj), "null existing-edge arrays");
  k_filter_edges<<<1, kGraphThreads, 0, (cudaStream_t)stream>>>(ii, jj, n_dev, n_max, ex_ii, ex_jj, n_existing, out_ii, out_jj,
                                                                 n_out);
  DBAF_LAUNCH_CHECK("k_filter_edges");
  return DBAF_OK;
}
