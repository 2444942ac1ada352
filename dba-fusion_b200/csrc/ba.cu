// Dense bundle adjustment for sm_100a: graph bookkeeping, per-depth-frame linearisation
// (Jacobians, pose Hessian blocks, depth blocks), Schur complement, float64 Cholesky,
// depth back-substitution and SE3 / disparity retraction.  Everything stays on the device;
// the host only enqueues kernels.
//
// Replaces ba_cuda / ba_cuda_extend / BACore of the reference (src/droid_kernels.cu:1394-1956).
// Data layout in HBM (one caller-owned workspace, see dbaf_ba_layout):
//   H  float64 [6P,6P] + v float64 [6P]   reduced camera system, accumulated with atomics
//   Eij float32 [E,6,HW], Ei float32 [P,6,HW]   pose-depth blocks, pixel index fastest
//   Q, w float32 [K,HW]                    inverse depth Hessian and depth gradient per depth frame
//   kx, kk_exp, csr_*, srow_*  int32       graph tables
#include "common.cuh"
#include "tc_ptx.cuh"

#include <cstdlib>
#include <cstring>

#include <algorithm>
#include <atomic>

namespace dbaf {

namespace {

constexpr int kThreads = 256;
constexpr int kWarps = kThreads / 32;
constexpr int kSmemSolveMaxN = 160;   // (161*160 + 160)*8 B = 207 KB of shared memory

inline int64_t align_up(int64_t x, int64_t a) { return (x + a - 1) / a * a; }

struct Ws {
  int* info;
  double* H;
  double* v;
  float* dx;
  int* kx;
  int* kk_exp;
  int* csr_ptr;
  int* csr_idx;
  int* srow_ptr;
  int* srow_idx;
  int* frame2k;
  int* cursor;
  float* Q;
  float* w;
  float* Ei;
  float* Eij;
  double* Lfac;
  int* fused;     // [Kcap] 1: the frame went through k_linschur_tc (no Eij; the back-substitution recomputes it)
  float* exf;     // [E][28] relative-pose record of every edge of such a frame (LsEdge), read by k_backsub_rc
  const float** rc_ptrs;  // [2] weights and intrinsics pointers of the last fused linearisation
  int Kcap;
};

void fill_layout(int N, int E, int t0, int t1, int ht, int wd, dbaf_ba_layout* L) {
  const int64_t P = t1 - t0, HW = (int64_t)ht * wd, n = 6 * P;
  const int64_t Kcap = std::min<int64_t>(N, P + E);
  int64_t off = 0;
  auto take = [&](int64_t bytes) { int64_t o = off; off = align_up(off + bytes, 256); return o; };
  L->info = take(16 * 4);
  L->H = take((n * n + n) * 8);
  L->v = L->H + n * n * 8;
  L->dx = take(std::max<int64_t>(1, n) * 4);
  L->Lfac = take(n > kSmemSolveMaxN ? ((n + 1) * n + n) * 8 : 8);
  L->kx = take(Kcap * 4);
  L->kk_exp = take((P + E) * 4);
  L->csr_ptr = take((Kcap + 1) * 4);
  L->csr_idx = take(std::max<int64_t>(1, E) * 4);
  L->srow_ptr = take((Kcap + 1) * 4);
  L->srow_idx = take((P + E) * 4);
  L->frame2k = take((int64_t)N * 4);
  L->cursor = take((Kcap + 1) * 4);
  L->Q = take(Kcap * HW * 4);
  L->w = take(Kcap * HW * 4);
  L->Ei = take(P * 6 * HW * 4);
  L->Eij = take((int64_t)E * 6 * HW * 4);
  L->fused = take(Kcap * 4);
  L->exf = take(std::max<int64_t>(1, E) * 28 * 4);
  L->rc_ptrs = take(2 * 8);
  L->Kcap = Kcap;
  L->total_bytes = off;
}

Ws view(void* base, const dbaf_ba_layout& L) {
  char* b = static_cast<char*>(base);
  Ws w;
  w.info = (int*)(b + L.info);
  w.H = (double*)(b + L.H);
  w.v = (double*)(b + L.v);
  w.dx = (float*)(b + L.dx);
  w.kx = (int*)(b + L.kx);
  w.kk_exp = (int*)(b + L.kk_exp);
  w.csr_ptr = (int*)(b + L.csr_ptr);
  w.csr_idx = (int*)(b + L.csr_idx);
  w.srow_ptr = (int*)(b + L.srow_ptr);
  w.srow_idx = (int*)(b + L.srow_idx);
  w.frame2k = (int*)(b + L.frame2k);
  w.cursor = (int*)(b + L.cursor);
  w.Q = (float*)(b + L.Q);
  w.w = (float*)(b + L.w);
  w.Ei = (float*)(b + L.Ei);
  w.Eij = (float*)(b + L.Eij);
  w.Lfac = (double*)(b + L.Lfac);
  w.fused = (int*)(b + L.fused);
  w.exf = (float*)(b + L.exf);
  w.rc_ptrs = (const float**)(b + L.rc_ptrs);
  w.Kcap = (int)L.Kcap;
  return w;
}

// ------------------------------------------------------------------ tables ----

// In-place exclusive scan of data[0..n) by one 1024-thread block; returns the total.
__device__ int block_exclusive_scan(int* data, int n) {
  __shared__ int warp_tot[32];
  __shared__ int carry;
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  if (tid == 0) carry = 0;
  __syncthreads();
  for (int base = 0; base < n; base += 1024) {
    const int i = base + tid;
    const int x = i < n ? data[i] : 0;
    int incl = x;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int y = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += y;
    }
    if (lane == 31) warp_tot[wid] = incl;
    __syncthreads();
    if (wid == 0) {
      const int t = warp_tot[lane];
      int s = t;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int y = __shfl_up_sync(0xffffffffu, s, o);
        if (lane >= o) s += y;
      }
      warp_tot[lane] = s - t;
    }
    __syncthreads();
    const int excl = incl - x + warp_tot[wid] + carry;
    if (i < n) data[i] = excl;
    __syncthreads();
    if (tid == 1023) carry = excl + x;
    __syncthreads();
  }
  return carry;
}

// One block builds every table of the window: kx / kk_exp (sorted unique of
// cat(arange(t0,t1), ii) and its inverse), edges grouped by source frame in ascending edge
// order, and per depth frame the rows of E that enter the Schur complement.
__global__ void __launch_bounds__(1024) k_tables(const int64_t* __restrict__ ii,
                                                 const int64_t* __restrict__ jj, int E, int N,
                                                 int t0, int t1, Ws ws) {
  const int tid = threadIdx.x;
  const int P = t1 - t0;
  int* f2k = ws.frame2k;
  if (tid < 16) ws.info[tid] = 0;
  if (ws.fused) {
    for (int q = tid; q < ws.Kcap; q += 1024) ws.fused[q] = 0;
  }
  for (int f = tid; f < N; f += 1024) f2k[f] = (f >= t0 && f < t1) ? 1 : 0;
  __syncthreads();
  for (int e = tid; e < E; e += 1024) {
    const long long i = ii[e], j = jj[e];
    if (i < 0 || i >= N || j < 0 || j >= N) ws.info[2] = 1;
    else f2k[i] = 1;
  }
  __syncthreads();
  // flags -> exclusive ranks in place (absent frames become -1); kx lists the present frames
  __shared__ int K_s;
  {
    __shared__ int warp_tot[32];
    __shared__ int carry;
    const int lane = tid & 31, wid = tid >> 5;
    if (tid == 0) carry = 0;
    __syncthreads();
    for (int base = 0; base < N; base += 1024) {
      const int f = base + tid;
      const int x = f < N ? f2k[f] : 0;
      int incl = x;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int y = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += y;
      }
      if (lane == 31) warp_tot[wid] = incl;
      __syncthreads();
      if (wid == 0) {
        const int t = warp_tot[lane];
        int s = t;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          const int y = __shfl_up_sync(0xffffffffu, s, o);
          if (lane >= o) s += y;
        }
        warp_tot[lane] = s - t;
      }
      __syncthreads();
      const int excl = incl - x + warp_tot[wid] + carry;
      if (f < N) {
        f2k[f] = x ? excl : -1;
        if (x && excl < ws.Kcap) ws.kx[excl] = f;
      }
      __syncthreads();
      if (tid == 1023) carry = excl + x;
      __syncthreads();
    }
    if (tid == 0) K_s = carry;
    __syncthreads();
  }
  const int K = K_s;
  // out-degree per depth frame
  for (int k = tid; k <= K; k += 1024) ws.csr_ptr[k] = 0;
  __syncthreads();
  for (int e = tid; e < E; e += 1024) {
    const long long i = ii[e], j = jj[e];
    if (i >= 0 && i < N && j >= 0 && j < N) atomicAdd(&ws.csr_ptr[f2k[i]], 1);
  }
  __syncthreads();
  const int n_valid = block_exclusive_scan(ws.csr_ptr, K);
  if (tid == 0) ws.csr_ptr[K] = n_valid;
  for (int k = tid; k < K; k += 1024) ws.cursor[k] = 0;
  __syncthreads();
  for (int e = tid; e < E; e += 1024) {
    const long long i = ii[e], j = jj[e];
    if (i >= 0 && i < N && j >= 0 && j < N) {
      const int k = f2k[i];
      const int pos = atomicAdd(&ws.cursor[k], 1);
      ws.csr_idx[ws.csr_ptr[k] + pos] = e;
    }
  }
  __syncthreads();
  // ascending edge order inside each segment (insertion sort; segments are short)
  for (int k = tid; k < K; k += 1024) {
    const int lo = ws.csr_ptr[k], hi = ws.csr_ptr[k + 1];
    for (int a = lo + 1; a < hi; ++a) {
      const int x = ws.csr_idx[a];
      int b = a - 1;
      while (b >= lo && ws.csr_idx[b] > x) { ws.csr_idx[b + 1] = ws.csr_idx[b]; --b; }
      ws.csr_idx[b + 1] = x;
    }
  }
  // inverse map of the unique
  for (int n = tid; n < P + E; n += 1024) {
    int k = -1;
    if (n < P) k = f2k[t0 + n];
    else {
      const long long i = ii[n - P];
      if (i >= 0 && i < N) k = f2k[i];
    }
    ws.kk_exp[n] = k;
  }
  __syncthreads();
  // rows of E per depth frame whose pose lies inside the window: the self row (row index
  // frame-t0) and one row per out-edge (row index P+e) with t0 <= jj[e] < t1.
  for (int k = tid; k <= K; k += 1024) ws.srow_ptr[k] = 0;
  __syncthreads();
  for (int k = tid; k < K; k += 1024) {
    const int f = ws.kx[k];
    int c = (f >= t0 && f < t1) ? 1 : 0;
    for (int a = ws.csr_ptr[k]; a < ws.csr_ptr[k + 1]; ++a) {
      const long long j = jj[ws.csr_idx[a]];
      c += (j >= t0 && j < t1) ? 1 : 0;
    }
    ws.srow_ptr[k] = c;
  }
  __syncthreads();
  const int n_rows = block_exclusive_scan(ws.srow_ptr, K);
  if (tid == 0) { ws.srow_ptr[K] = n_rows; ws.info[4] = n_rows; ws.info[0] = K; }
  __syncthreads();
  for (int k = tid; k < K; k += 1024) {
    const int f = ws.kx[k];
    int o = ws.srow_ptr[k];
    if (f >= t0 && f < t1) ws.srow_idx[o++] = f - t0;
    for (int a = ws.csr_ptr[k]; a < ws.csr_ptr[k + 1]; ++a) {
      const int e = ws.csr_idx[a];
      const long long j = jj[e];
      if (j >= t0 && j < t1) ws.srow_idx[o++] = P + e;
    }
  }
}

// -------------------------------------------------------------- linearise ----

// value index l in [0,78) <-> lower triangle (n,m), m<=n, of the 12x12 edge Hessian,
// enumerated row by row; [78,84) = gradient wrt pose i, [84,90) = gradient wrt pose j.
__device__ __forceinline__ void tri_unpack(int l, int& n, int& m) {
  n = (int)((sqrtf(8.0f * l + 1.0f) - 1.0f) * 0.5f);
  while ((n + 1) * (n + 2) / 2 <= l) ++n;
  while (n * (n + 1) / 2 > l) --n;
  m = l - n * (n + 1) / 2;
}

// (the linearisation kernel lives in ba_lin2.cuh)

// ------------------------------------------------------------------ Schur ----

// One CTA = depth frame k x pixel tile; each warp owns row pairs (a>=b) of the frame's slab
// and accumulates the 6x6 block  sum_px E_a Q E_b^T  (plus E_a Q w on the diagonal pairs),
// then subtracts it from H (both triangles) / v.
__global__ void __launch_bounds__(kThreads)
k_schur(const int64_t* __restrict__ jj, int t0, int P, int HW, int tile_px, int min_rows, int k_lo, int k_hi, Ws ws) {
  const int k = k_lo + blockIdx.y;
  const int K = ws.info[0];
  if (k >= K || (k_hi >= 0 && k >= k_hi)) return;
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int r_lo = ws.srow_ptr[k];
  const int m = ws.srow_ptr[k + 1] - r_lo;
  if (m < min_rows) return;
  const int pairs = m * (m + 1) / 2;
  const int p_lo = blockIdx.x * tile_px, p_hi = min(HW, p_lo + tile_px);
  const int n6 = 6 * P;
  const float* Qk = ws.Q + (size_t)k * HW;
  const float* wk = ws.w + (size_t)k * HW;

  for (int q = wid; q < pairs; q += kWarps) {
    int a, b;
    tri_unpack(q, a, b);
    const int na = ws.srow_idx[r_lo + a], nb = ws.srow_idx[r_lo + b];
    const float* Ea = na < P ? ws.Ei + (size_t)na * 6 * HW : ws.Eij + (size_t)(na - P) * 6 * HW;
    const float* Eb = nb < P ? ws.Ei + (size_t)nb * 6 * HW : ws.Eij + (size_t)(nb - P) * 6 * HW;
    const int pa = na < P ? na : (int)jj[na - P] - t0;
    const int pb = nb < P ? nb : (int)jj[nb - P] - t0;
    const bool diag = a == b;

    float acc[64];
#pragma unroll
    for (int l = 0; l < 64; ++l) acc[l] = 0.f;
    for (int p = p_lo + lane; p < p_hi; p += 32) {
      const float qv = __ldg(Qk + p);
      float ea[6], eb[6];
#pragma unroll
      for (int c = 0; c < 6; ++c) {
        ea[c] = Ea[(size_t)c * HW + p] * qv;
        eb[c] = Eb[(size_t)c * HW + p];
      }
#pragma unroll
      for (int r = 0; r < 6; ++r)
#pragma unroll
        for (int c = 0; c < 6; ++c) acc[r * 6 + c] += ea[r] * eb[c];
      if (diag) {
        const float wv = __ldg(wk + p);
#pragma unroll
        for (int r = 0; r < 6; ++r) acc[36 + r] += ea[r] * wv;
      }
    }
    warp_reduce32(acc, lane);
    warp_reduce32(acc + 32, lane);
    // lane l now holds value l in acc[0] and value 32+l in acc[32]
    {
      const int l = lane;
      const int r = l / 6, c = l % 6;
      const double val = -(double)acc[0];
      atomicAdd(ws.H + (size_t)(6 * pa + r) * n6 + 6 * pb + c, val);
      if (!diag) atomicAdd(ws.H + (size_t)(6 * pb + c) * n6 + 6 * pa + r, val);
    }
    if (lane < 4) {
      const int l = 32 + lane;
      const int r = l / 6, c = l % 6;
      const double val = -(double)acc[32];
      atomicAdd(ws.H + (size_t)(6 * pa + r) * n6 + 6 * pb + c, val);
      if (!diag) atomicAdd(ws.H + (size_t)(6 * pb + c) * n6 + 6 * pa + r, val);
    } else if (diag && lane < 10) {
      atomicAdd(ws.v + 6 * pa + (lane - 4), -(double)acc[32]);
    }
  }
}

// ------------------------------------------------------------------ solve ----

// In-place Cholesky of the n x n matrix L (leading dimension ld) with the right-hand side
// carried as row n (so y = L^-1 b falls out of the factorisation), then back-substitution.
// One CTA.  Returns false when a pivot is <= 0.
__device__ bool cholesky_solve(double* L, int ld, int n, double* x) {
  const int tid = threadIdx.x, nt = blockDim.x;
  __shared__ int fail;
  if (tid == 0) fail = 0;
  __syncthreads();
  for (int j = 0; j < n; ++j) {
    const double piv = L[(size_t)j * ld + j];
    if (piv <= 0.0) {
      if (tid == 0) fail = 1;
      break;   // uniform: every thread reads the same pivot
    }
    const double inv = 1.0 / sqrt(piv);
    __syncthreads();
    // scale column j (rows j..n, row n is the rhs)
    for (int i = j + 1 + tid; i <= n; i += nt) L[(size_t)i * ld + j] *= inv;
    if (tid == 0) L[(size_t)j * ld + j] = piv * inv;
    __syncthreads();
    // trailing update of rows j+1..n (lower triangle) : L[i][c] -= L[i][j]*L[c][j]
    const int rem = n - j;           // rows j+1..n  -> rem rows, row n included
    const int total = rem * (rem + 1) / 2;
    for (int t = tid; t < total; t += nt) {
      int a, b;
      tri_unpack(t, a, b);           // a>=b in [0,rem)
      const int i = j + 1 + a, c = j + 1 + b;
      if (c < n) L[(size_t)i * ld + c] -= L[(size_t)i * ld + j] * L[(size_t)c * ld + j];
    }
    __syncthreads();
  }
  __syncthreads();
  if (fail) return false;
  // row n now holds y (forward substitution done column by column above); solve L^T x = y
  for (int i = tid; i < n; i += nt) x[i] = L[(size_t)n * ld + i];
  __syncthreads();
  for (int j = n - 1; j >= 0; --j) {
    if (tid == 0) x[j] /= L[(size_t)j * ld + j];
    __syncthreads();
    const double xj = x[j];
    for (int i = tid; i < j; i += nt) x[i] -= L[(size_t)j * ld + i] * xj;
    __syncthreads();
  }
  return true;
}

// L = H (+ prior), damped; Cholesky in shared memory when it fits, else in the workspace.
__global__ void __launch_bounds__(kThreads)
k_solve(int n, float lm, float ep, const double* __restrict__ prior, int prior_rows, int prior_cols,
        int use_smem, Ws ws) {
  extern __shared__ double sm[];
  const int tid = threadIdx.x, nt = blockDim.x;
  const int ld = n;
  double* L = use_smem ? sm : ws.Lfac;                 // (n+1) x n, row n = rhs
  double* x = use_smem ? sm + (size_t)(n + 1) * n : ws.Lfac + (size_t)(n + 1) * n;
  for (int t = tid; t < n * n; t += nt) {
    const int i = t / n, c = t - i * n;
    if (c > i) continue;                               // lower triangle only
    double a = ws.H[t];
    if (prior) {
      // ba_cuda_extend: L = Ad + Adprior, Adprior = Ad overwritten by the prior on its block
      a += (i < prior_rows && c < prior_cols) ? prior[(size_t)i * prior_cols + c] : ws.H[t];
    }
    if (i == c) a += (double)ep + (double)lm * a;
    L[(size_t)i * ld + c] = a;
  }
  for (int i = tid; i < n; i += nt) L[(size_t)n * ld + i] = ws.v[i];
  __syncthreads();
  const bool ok = cholesky_solve(L, ld, n, x);
  __syncthreads();
  for (int i = tid; i < n; i += nt) ws.dx[i] = ok ? (float)x[i] : 0.0f;
  if (!ok && tid == 0) atomicAdd(&ws.info[1], 1);
}

// ---------------------------------------------------------------- retract ----

__global__ void k_pose_retract(float* __restrict__ poses, const float* __restrict__ dx, int t0, int P,
                               float* __restrict__ dx_out) {
  for (int k = threadIdx.x; k < P; k += blockDim.x) {
    float xi[6], T[7];
#pragma unroll
    for (int n = 0; n < 6; ++n) xi[n] = dx[6 * k + n];
    if (poses) {
#pragma unroll
      for (int n = 0; n < 7; ++n) T[n] = poses[7 * (t0 + k) + n];
      retract_pose(xi, T);
#pragma unroll
      for (int n = 0; n < 7; ++n) poses[7 * (t0 + k) + n] = T[n];
    }
    if (dx_out) {
#pragma unroll
      for (int n = 0; n < 6; ++n) dx_out[6 * k + n] = xi[n];
    }
  }
}

// Dense frame graphs (global BA): frames with 8..21 row blocks go to the tensor-core Schur kernel; those with at
// most 20 out-edges among them can be linearised and Schur-complemented in one kernel (ba_linschur_tc.cuh).
constexpr int kStcMaxM = 21;                 // 6 * 21 = 126 rows + w + q = 128
constexpr int kStcMinM = 8;
constexpr int kLsMaxDeg = 20;                // out-edges per frame (120 operand rows)
__device__ __forceinline__ bool ls_eligible(const Ws& ws, int k, int min_m) {
  const int m = ws.srow_ptr[k + 1] - ws.srow_ptr[k];
  const int deg = ws.csr_ptr[k + 1] - ws.csr_ptr[k];
  return m >= min_m && m <= kStcMaxM && deg >= 4 && deg <= kLsMaxDeg;
}

#include "ba_fast.cuh"
#include "ba_banded.cuh"
#include "ba_banded_reg.cuh"
#include "ba_spike.cuh"
#include "ba_lin2.cuh"
#include "ba_schur_tc.cuh"
#include "ba_linschur_tc.cuh"

int check_dims(int N, int E, int t0, int t1, int ht, int wd) {
  DBAF_REQUIRE(N > 0 && E >= 0 && ht > 0 && wd > 0, "non-positive size");
  DBAF_REQUIRE(t0 >= 0 && t1 >= t0 && t1 <= N, "need 0 <= t0 <= t1 <= n_frames");
  return DBAF_OK;
}

}  // namespace
}  // namespace dbaf

using namespace dbaf;

extern "C" int dbaf_ba_workspace_layout(int n_frames, int n_edges, int t0, int t1, int ht, int wd,
                                        dbaf_ba_layout* out) {
  DBAF_REQUIRE(out != nullptr, "null layout");
  if (int rc = check_dims(n_frames, n_edges, t0, t1, ht, wd)) return rc;
  fill_layout(n_frames, n_edges, t0, t1, ht, wd, out);
  return DBAF_OK;
}

extern "C" int dbaf_ba_prepare(void* workspace, const int64_t* ii, const int64_t* jj, int E, int N,
                               int t0, int t1, int ht, int wd, void* stream) {
  DBAF_REQUIRE(workspace != nullptr, "null workspace");
  if (int rc = check_dims(N, E, t0, t1, ht, wd)) return rc;
  dbaf_ba_layout L;
  fill_layout(N, E, t0, t1, ht, wd, &L);
  Ws ws = view(workspace, L);
  k_tables<<<1, 1024, 0, (cudaStream_t)stream>>>(ii, jj, E, N, t0, t1, ws);
  DBAF_LAUNCH_CHECK("k_tables");
  return DBAF_OK;
}

namespace {
int pick_pxt(int Kcap, int HW) {
  // keep at least ~2 CTAs per SM busy; two pixels per thread amortise the per-edge reduction.  Four pixels per
  // thread were measured slower everywhere since the next edge's inputs are prefetched into registers
  // (170 registers, one CTA per SM): C5 linearise+Schur 2.82 ms with 4, 2.56 ms with 2, 3.19 ms with 1.
  const long ctas1 = (long)Kcap * ((HW + 255) / 256);
  if (ctas1 >= 148L * 6) return 2;
  return 1;
}
}  // namespace

namespace {
// mean out-degree >= 7 over the window (frames with >= 8 row blocks exist), whole 32-pixel slices
bool dense_graph(int E, int Kcap, int HW) { return options().schur_simt == 0 && E >= 7 * Kcap && HW % 32 == 0; }

// Schur complement of the depth frames [k_lo, k_lo + k_cap): H -= E Q E^T, v -= E Q w from the E, Q, w in the workspace.
int launch_schur(Ws ws, const int64_t* jj, int E, int t0, int P, int HW, int k_lo, int k_hi, int k_cap, cudaStream_t st,
                 bool fused = false) {
  static bool schur_attr[64] = {};
  int dev_id = 0;
  DBAF_CUDA(cudaGetDevice(&dev_id));
  if (!schur_attr[dev_id & 63]) {   // once per device: the attribute call costs ~10 us of host time
    DBAF_CUDA(cudaFuncSetAttribute(k_schur_smem, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSchurSmemBytes));
    schur_attr[dev_id & 63] = true;
  }
  // dense graphs (global BA: ~20 out-edges per frame) go to the tensor-core kernel, whose cost does not grow
  // with the square of the row count; dbaf_set_option("schur_impl", "simt") keeps everything on the fp32 SIMT kernels
  const bool simt_only = options().schur_simt != 0;
  // mean out-degree >= 7, i.e. frames with >= 8 row blocks exist; whole 32-pixel slices for the 16-byte copies
  const bool use_tc = fused || (!simt_only && E >= 7 * k_cap && HW % 32 == 0);
  const int tc_min_m = kStcMinM;
  if (use_tc && !fused) {      // (fused: k_linschur_tc has already subtracted the Schur terms of the frames it took)
    static bool tc_attr[64] = {};
    if (!tc_attr[dev_id & 63]) {
      DBAF_CUDA(cudaFuncSetAttribute(k_schur_tc, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kStcSmemBytes));
      tc_attr[dev_id & 63] = true;
    }
    // one CTA per frame when there are enough frames to fill the machine, else split the pixels
    int splits = std::max(1, std::min((2 * 148 + k_cap - 1) / k_cap, (HW + 255) / 256));
    const int St = (((HW + splits - 1) / splits) + 31) & ~31;
    splits = (HW + St - 1) / St;
    k_schur_tc<<<dim3(splits, k_cap), kStcThreads, kStcSmemBytes, st>>>(jj, t0, P, HW, St, tc_min_m, k_lo, k_hi, ws);
    DBAF_LAUNCH_CHECK("k_schur_tc");
  }
  // (fused: only the few frames k_linschur_tc left -- short out-edge lists at the ends of the chain -- remain; one CTA
  // walks all pixel tiles of such a frame, so that the ~1000 frames that return at once cost 1000 CTAs, not 16 000)
  const int S = fused ? HW : 256;
  dim3 sgrid((HW + S - 1) / S, k_cap);
  k_schur_smem<<<sgrid, kThreads, kSchurSmemBytes, st>>>(jj, t0, P, HW, S, k_lo, k_hi, use_tc ? tc_min_m : 1,
                                                         use_tc ? kStcMaxM : 0, fused ? 1 : 0, ws);
  DBAF_LAUNCH_CHECK("k_schur_smem");
  if (E + 1 > kSchurMaxRows) {
    // a depth frame can only have more than kSchurMaxRows rows if the graph has that many edges
    const int tile_px = fused ? HW : 1024;
    dim3 ogrid((HW + tile_px - 1) / tile_px, k_cap);
    k_schur<<<ogrid, kThreads, 0, st>>>(jj, t0, P, HW, tile_px, kSchurMaxRows + 1, k_lo, k_hi, ws);
    DBAF_LAUNCH_CHECK("k_schur");
  }
  return DBAF_OK;
}
}  // namespace

extern "C" int dbaf_ba_linearize(void* workspace, const float* poses, const float* disps,
                                 const float* intrinsics, const float* disps_sens,
                                 const float* targets, const float* weights, const float* eta,
                                 int eta_rows, const int64_t* ii, const int64_t* jj, int E, int N,
                                 int t0, int t1, int ht, int wd, float alpha, int motion_only,
                                 int zero_first, int k_lo, int k_hi, void* stream) {
  (void)ii;
  DBAF_REQUIRE(workspace != nullptr, "null workspace");
  if (int rc = check_dims(N, E, t0, t1, ht, wd)) return rc;
  DBAF_REQUIRE(motion_only == 1 || eta_rows >= 1, "eta needs at least one row");
  DBAF_REQUIRE(motion_only >= 0 && motion_only <= 2, "motion_only must be 0, 1 or 2");
  dbaf_ba_layout L;
  fill_layout(N, E, t0, t1, ht, wd, &L);
  Ws ws = view(workspace, L);
  cudaStream_t st = (cudaStream_t)stream;
  const int P = t1 - t0, HW = ht * wd;
  const int64_t n = 6 * (int64_t)P;
  if (zero_first) DBAF_CUDA(cudaMemsetAsync(ws.H, 0, (n * n + n) * 8, st));
  if (P == 0 && motion_only == 1) return DBAF_OK;
  if (k_lo < 0) k_lo = 0;
  const int k_cap = (k_hi >= 0 ? std::min(k_hi, ws.Kcap) : ws.Kcap) - k_lo;
  if (k_cap <= 0) return DBAF_OK;
  // Dense graphs: eligible frames are linearised and Schur-complemented in one kernel that keeps the pose-depth
  // blocks on the SM (dbaf_set_option("eij", "store") keeps the two-kernel path that writes them to the workspace).
  // The decision uses the window's totals, not this call's frame range, so that a window processed in several
  // ranges (dbaf_b200.global_ba) takes the same path in each.
  const auto al16 = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
  const bool fused = motion_only == 0 && P > 0 && dense_graph(E, ws.Kcap, HW) && options().eij_recompute != 0 &&
                     ht <= kLsTabMax && wd <= kLsTabMax &&
                     al16(disps) && al16(disps_sens) && al16(targets) && al16(weights) && al16(eta);
  if (fused) {
    static bool ls_attr[64] = {};
    int dev_id = 0;
    DBAF_CUDA(cudaGetDevice(&dev_id));
    if (!ls_attr[dev_id & 63]) {
      DBAF_CUDA(cudaFuncSetAttribute(k_linschur_tc, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kLsSmemBytes));
      ls_attr[dev_id & 63] = true;
    }
    // Pixel ranges per frame: the kernel runs one CTA per SM, in rounds of 148; a CTA that takes 1/s of a frame costs
    // 1/s + ~0.07 of a whole-frame CTA (prologue, epilogue sums and atomics), so the split that minimises
    // rounds x cost per round is taken.  It is derived from the WINDOW's frame count, not from this call's range: the
    // per-CTA sums then do not depend on how the window is sharded, and N ranks reproduce one rank bit for bit.
    int splits = 1;
    {
      double best = 1e30;
      const int s_max = std::max(1, std::min(8, HW / 256));
      for (int s_ = 1; s_ <= s_max; ++s_) {
        const double cost = (double)((ws.Kcap * s_ + 147) / 148) * (1.0 / s_ + 0.07);
        if (cost < best - 1e-9) { best = cost; splits = s_; }
      }
    }
    const int St = (((HW + splits - 1) / splits) + 31) & ~31;
    splits = (HW + St - 1) / St;
    k_depth_diag<<<dim3((HW / 4 + 127) / 128, k_cap), 128, 0, st>>>(poses, disps, intrinsics, disps_sens, weights, eta,
                                                                   eta_rows, jj, t0, P, HW, wd, alpha, kStcMinM, k_lo,
                                                                   k_hi, ws);
    DBAF_LAUNCH_CHECK("k_depth_diag");
    k_linschur_tc<<<dim3(splits, k_cap), kLsThreads, kLsSmemBytes, st>>>(disps, intrinsics, disps_sens, targets, weights, t0,
                                                                      P, HW, wd, alpha, St, kStcMinM, k_lo, k_hi, ws);
    DBAF_LAUNCH_CHECK("k_linschur_tc");
  }
  const int skip_min_m = fused ? kStcMinM : 0;
  const int pxt = pick_pxt(k_cap, HW);
  dim3 grid((HW + 256 * pxt - 1) / (256 * pxt), k_cap);
#define DBAF_LIN(PX, MO)                                                                       \
  k_linearize2<PX, MO><<<grid, kThreads, 0, st>>>(poses, disps, intrinsics, disps_sens, targets, \
                                                 weights, eta, eta_rows, jj, t0, P, HW, wd,    \
                                                 alpha, k_lo, k_hi, skip_min_m, ws)
  if (motion_only == 1) {
    if (pxt == 2) DBAF_LIN(2, true); else DBAF_LIN(1, true);
  } else {
    if (pxt == 2) DBAF_LIN(2, false); else DBAF_LIN(1, false);
  }
#undef DBAF_LIN
  DBAF_LAUNCH_CHECK("k_linearize2");
  if (motion_only == 0 && P > 0) {
    if (int rc = launch_schur(ws, jj, E, t0, P, HW, k_lo, k_hi, k_cap, st, fused)) return rc;
  }
  return DBAF_OK;
}

// The Schur stage of dbaf_ba_linearize on its own: H -= E Q E^T, v -= E Q w from whatever E, Q, w the workspace
// holds (diagnostics and tests: lets a caller feed the Schur kernels with chosen blocks).
extern "C" int dbaf_ba_schur(void* workspace, const int64_t* jj, int E, int N, int t0, int t1, int ht, int wd,
                             int k_lo, int k_hi, void* stream) {
  DBAF_REQUIRE(workspace != nullptr, "null workspace");
  if (int rc = check_dims(N, E, t0, t1, ht, wd)) return rc;
  dbaf_ba_layout L;
  fill_layout(N, E, t0, t1, ht, wd, &L);
  Ws ws = view(workspace, L);
  if (k_lo < 0) k_lo = 0;
  const int k_cap = (k_hi >= 0 ? std::min(k_hi, ws.Kcap) : ws.Kcap) - k_lo;
  if (k_cap <= 0 || t1 == t0) return DBAF_OK;
  return launch_schur(ws, jj, E, t0, t1 - t0, ht * wd, k_lo, k_hi, k_cap, (cudaStream_t)stream);
}

namespace {
// number of domains the banded solve of P poses with block half-bandwidth B is cut into (1: one sweep, 2: two sweeps
// meeting at one separator, >= 3: ba_spike.cuh), before the check against the size of the factorisation scratch
int banded_domains(int P, int B) {
  const bool single = options().band_single != 0;
  const bool smem_sweep = B > kBandRegMaxB || options().band_smem != 0;
  int D = 1;
  if (!single && !options().band_twisted && !smem_sweep && B >= 1) {
    const int n_min = B + 2;                                   // interiors longer than the band
    const int d_max = (P + B) / (B + n_min);
    // (cyclic reduction over the separators: the chain costs log2(D) levels, so interiors of ~2 bands are long enough)
    const int d_want = options().band_domains > 0 ? options().band_domains : (P + B) / (3 * std::max(B, 4));
    D = std::min({d_want, d_max, kSpikeMaxD});
    if (D < 3) D = 1;
  }
  if (D < 3) D = (!single && P >= kBandTwistMinP && P >= 3 * B + 2) ? 2 : 1;
  return D;
}

int launch_solve(void* workspace, int t0, int t1, float lm, float ep, const double* prior, int prior_rows,
                 int prior_cols, int rezero, cudaStream_t st, int band_blocks = 0) {
  // Only info, H, v, dx and Lfac are touched; their offsets depend on P alone.
  dbaf_ba_layout L;
  fill_layout(1, 0, 0, t1 - t0, 1, 1, &L);
  Ws ws = view(workspace, L);
  const int P = t1 - t0, n = 6 * P;
  if (n == 0) return DBAF_OK;
  if (n <= kSmemSolveMaxN) {
    const int ld = n | 1;
    const size_t smem = ((size_t)(n + 1) * ld + 2 * n) * sizeof(double);
    // static shared memory of the kernel counts against the 227 KB opt-in limit
    static bool solve_attr[64] = {};
    int dev_id = 0;
    DBAF_CUDA(cudaGetDevice(&dev_id));
    if (!solve_attr[dev_id & 63]) {
      DBAF_CUDA(cudaFuncSetAttribute(k_solve6, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024));
      solve_attr[dev_id & 63] = true;
    }
    k_solve6<<<1, kThreads, smem, st>>>(P, lm, ep, prior, prior_rows, prior_cols, rezero, ws);
    DBAF_LAUNCH_CHECK("k_solve6");
  } else if (band_blocks > 0 && band_blocks <= kBandMaxB && prior == nullptr) {
    const int B = std::min(band_blocks, P - 1);
    const int NB = B + 2;
    const size_t smem = std::max((size_t)NB * NB * 36 + NB * 6 + 72 + 12 + 8, (size_t)band_back_smem_doubles(B)) * sizeof(double);
    static bool band_attr[64] = {};
    int dev_id = 0;
    DBAF_CUDA(cudaGetDevice(&dev_id));
    if (!band_attr[dev_id & 63]) {
      DBAF_CUDA(cudaFuncSetAttribute(k_solve_banded, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024));
      band_attr[dev_id & 63] = true;
    }
    BandPlan plan{};
    const bool single = options().band_single != 0;              // the one-CTA sweep (tests)
    const bool smem_sweep = B > kBandRegMaxB || options().band_smem != 0;
    // ---- how the chain is cut: D domains (ba_spike.cuh), two sweeps meeting at one separator, or one sweep
    SpikePlan sp{};
    int D = banded_domains(P, B);
    const long long lfac_cap = (long long)(n + 1) * n + n;
    if (D >= 3) {
      const int NS = 6 * B, ld = spike_ld(NS);
      auto even = [](long long x) { return (x + 1) & ~1LL; };
      long long off = 0;
      auto take = [&](long long doubles) { const long long o = off; off = even(off + doubles); return o; };
      const int interior = P - (D - 1) * B;
      int first = 0;
      sp.D = D; sp.B = B; sp.G = std::max(1, std::min(8, NS / 6));
      // cyclic reduction over the separators 1 .. D-1: level l eliminates s with s mod 2^(l+1) == 2^l while more than
      // one multiple of 2^l is left; the last one (sp.mid) is solved densely
      sp.nlev = 0;
      while ((D - 1) >> sp.nlev > 1) ++sp.nlev;      // multiples of 2^l in [1, D-1]: (D-1) >> l
      sp.mid = 1 << sp.nlev;
      for (int d = 0; d < D; ++d) {
        const int nd = interior / D + (d < interior % D ? 1 : 0);
        BandPart& pt = plan.part[d];
        pt.Plog = nd + B;
        pt.n_elim = nd;
        if (d < D - 1) {
          pt.rev = 0;
          pt.base = first;
          pt.sep_first = first + nd;
          sp.m[d + 1] = first + nd;
        } else {
          pt.rev = 1;
          pt.base = P - 1;
          pt.sep_first = sp.m[D - 1];
        }
        pt.store = take(band_store_doubles(pt.Plog, B));
        pt.dump = take(band_dump_doubles(B));
        first += nd + B;
      }
      for (int d = 1; d <= D - 2; ++d) {
        sp.W[d] = take((long long)plan.part[d].n_elim * 6 * NS);
        sp.Roff[d] = take((long long)NS * NS);
        sp.SLL[d] = take((long long)NS * NS + NS);
      }
      for (int q = 1; q <= D - 1; ++q) {
        sp.Lf[q] = take((long long)(NS + 2) * ld);
        sp.X[q][0] = take((long long)NS * NS);
        sp.X[q][1] = take((long long)NS * NS);
      }
      const long long ublk = (long long)NS * NS + NS;            // (even)
      sp.Ubase = take(2 * (long long)(D - 1) * ublk);
      sp.Ubytes = 2 * (long long)(D - 1) * ublk * 8;
      for (int q = 1; q <= D - 1; ++q) {
        sp.U[q][0] = sp.Ubase + (2 * (long long)(q - 1)) * ublk;
        sp.U[q][1] = sp.Ubase + (2 * (long long)(q - 1) + 1) * ublk;
      }
      for (int l = 1; l <= sp.nlev && l < kSpikeMaxLev; ++l)
        for (int a = 1 << l; a + (1 << l) <= D - 1; a += 1 << l) sp.Cl[l][a] = take((long long)NS * NS);
      sp.xsep = take((long long)D * NS);
      plan.flags = take(D + 4);
      plan.twisted = 1;
      if (off > lfac_cap) D = 1;                                 // does not fit the factorisation scratch: fewer sweeps
    }
    if (D < 3) {
      plan = BandPlan{};
      if (!single && P >= kBandTwistMinP && P >= 3 * B + 2) {
        // two sweeps meeting at the separator [m, m+B)
        const int m = (P - B) / 2;
        plan.twisted = 1;
        plan.m = m;
        plan.sep = band_store_doubles(m + B, B) + band_store_doubles(P - m, B);
        plan.flags = plan.sep + 2 * band_dump_doubles(B) + 6 * B;
        plan.part[0] = BandPart{0, m + B, m, 0, m, 0, plan.sep};
        plan.part[1] = BandPart{1, P - m, P - m - B, P - 1, m, band_store_doubles(m + B, B), plan.sep + band_dump_doubles(B)};
        D = 2;
      } else {
        plan.part[0] = BandPart{0, P, P, 0, 0, 0, 0};
        plan.part[1] = plan.part[0];
        D = 1;
      }
    }
    if (D >= 3) {
      DBAF_CUDA(cudaMemsetAsync(ws.Lfac + plan.flags, 0, (size_t)(D + 4) * 8, st));
      DBAF_CUDA(cudaMemsetAsync(ws.Lfac + sp.Ubase, 0, (size_t)sp.Ubytes, st));     // the separators' update accumulators
    }
    // B <= 20: the sweep that keeps the window in registers; option band_sweep=smem keeps the shared-memory window
    if (smem_sweep) {
      k_solve_banded<<<D, kBandThreads, smem, st>>>(P, B, lm, ep, rezero, plan, ws);
      DBAF_LAUNCH_CHECK("k_solve_banded");
    } else {
      const size_t rsm = std::max((size_t)NB * (2 * kPnS + 6) + 84 + 4 + (size_t)NB * (NB - 1) * 18 + NB * 42,
                                  (size_t)band_back_smem_doubles(B)) * sizeof(double);
      static bool reg_attr[64] = {};
      if (!reg_attr[dev_id & 63]) {
        DBAF_CUDA(cudaFuncSetAttribute(k_solve_banded_reg, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024));
        reg_attr[dev_id & 63] = true;
      }
      k_solve_banded_reg<<<D, kBandThreads, rsm, st>>>(P, B, lm, ep, plan, ws);
      DBAF_LAUNCH_CHECK("k_solve_banded_reg");
    }
    if (D >= 3) {
      const int NS = 6 * B, ld = spike_ld(NS);
      static bool spike_attr[64] = {};
      if (!spike_attr[dev_id & 63]) {
        DBAF_CUDA(cudaFuncSetAttribute(k_spike_fwd, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
        DBAF_CUDA(cudaFuncSetAttribute(k_cr_a, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        DBAF_CUDA(cudaFuncSetAttribute(k_cr_b, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        DBAF_CUDA(cudaFuncSetAttribute(k_red_mid, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        DBAF_CUDA(cudaFuncSetAttribute(k_cr_back, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        DBAF_CUDA(cudaFuncSetAttribute(k_spike_back, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
        spike_attr[dev_id & 63] = true;
      }
      const int n_mid = D - 2;
      const int ncg = (NS + 29) / 30;
      const size_t fsm = ((size_t)kSpikeRing * ((B + 1) * 36 + 6) + (size_t)(B + 1) * 192 + 4 * 192) * sizeof(double);
      {
        // may start while the sweeps are still running and follows them through their progress words
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(ncg, n_mid);
        cfg.blockDim = dim3(128);
        cfg.dynamicSmemBytes = fsm;
        cfg.stream = st;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = attr;
        cfg.numAttrs = 1;
        DBAF_CUDA(cudaLaunchKernelEx(&cfg, k_spike_fwd, P, B, plan, sp, ws));
      }
      DBAF_LAUNCH_CHECK("k_spike_fwd");
      const int gt = (NS + kGramTile - 1) / kGramTile;
      k_spike_gram<<<dim3(gt * (gt + 1) / 2, n_mid), 256, 0, st>>>(B, plan, sp, ws);
      DBAF_LAUNCH_CHECK("k_spike_gram");
      const size_t asm_ = ((size_t)(NS + 1) * ld + NS + (size_t)2 * (kSpikeThreads / 32) * ld) * sizeof(double);   // L, 1/diag, per warp: x row + A row
      const int rps = (NS + sp.G - 1) / sp.G;
      const size_t bsm = ((size_t)NS * ld + (size_t)NS * (rps | 1)) * sizeof(double);   // Q^T and this CTA's rows of P, transposed
      auto n_elim = [&](int level) { const int h = 1 << level; return (D - 1 + h) / (2 * h); };   // s = h, 3h, 5h, ... <= D-1
      for (int level = 0; level < sp.nlev; ++level) {
        k_cr_a<<<dim3(sp.G, 2 * n_elim(level)), kSpikeThreads, asm_, st>>>(P, B, lm, ep, level, plan, sp, ws);
        DBAF_LAUNCH_CHECK("k_cr_a");
        k_cr_b<<<dim3(sp.G, 3 * n_elim(level)), kSpikeThreads, bsm, st>>>(B, level, plan, sp, ws);
        DBAF_LAUNCH_CHECK("k_cr_b");
      }
      const size_t msm = ((size_t)(NS + 1) * ld + 2 * NS) * sizeof(double);
      k_red_mid<<<1, kSpikeThreads, msm, st>>>(P, B, lm, ep, plan, sp, ws);
      DBAF_LAUNCH_CHECK("k_red_mid");
      for (int level = sp.nlev - 1; level >= 0; --level) {
        k_cr_back<<<n_elim(level), kSpikeThreads, ((size_t)NS * ld + 4 * NS) * sizeof(double), st>>>(B, level, plan, sp, ws);
        DBAF_LAUNCH_CHECK("k_cr_back");
      }
      int nd_max = 0;
      for (int d = 1; d <= D - 2; ++d) nd_max = std::max(nd_max, plan.part[d].n_elim);
      k_spike_apply<<<dim3((6 * nd_max + 7) / 8, n_mid), 256, 0, st>>>(B, plan, sp, ws);
      DBAF_LAUNCH_CHECK("k_spike_apply");
      k_spike_back<<<D, 64, band_back_smem_doubles(B) * sizeof(double), st>>>(P, B, plan, sp, ws);
      DBAF_LAUNCH_CHECK("k_spike_back");
    } else if (D == 2) {
      const int ns = 6 * B;
      const size_t ssm = ((size_t)(ns + 1) * (ns | 1) + 2 * ns) * sizeof(double);
      static bool sep_attr[64] = {};
      if (!sep_attr[dev_id & 63]) {
        DBAF_CUDA(cudaFuncSetAttribute(k_band_sep, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024));
        sep_attr[dev_id & 63] = true;
      }
      k_band_sep<<<1, kThreads, ssm, st>>>(P, B, lm, ep, plan, ws);
      DBAF_LAUNCH_CHECK("k_band_sep");
      static bool back_attr[64] = {};
      if (!back_attr[dev_id & 63]) {
        DBAF_CUDA(cudaFuncSetAttribute(k_band_back, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
        back_attr[dev_id & 63] = true;
      }
      k_band_back<<<2, 64, band_back_smem_doubles(B) * sizeof(double), st>>>(P, B, plan, ws);
      DBAF_LAUNCH_CHECK("k_band_back");
    }
    if (rezero) DBAF_CUDA(cudaMemsetAsync(ws.H, 0, ((size_t)n * n + n) * 8, st));
  } else {
    k_solve<<<1, kThreads, 0, st>>>(n, lm, ep, prior, prior_rows, prior_cols, 0, ws);
    DBAF_LAUNCH_CHECK("k_solve");
    if (rezero) DBAF_CUDA(cudaMemsetAsync(ws.H, 0, ((size_t)n * n + n) * 8, st));
  }
  return DBAF_OK;
}
}  // namespace

namespace dbaf {
namespace {
// one CTA per scalar row: the entries of H within `B` pose blocks of the diagonal, and v (row n)
__global__ void __launch_bounds__(128) k_zero_band(double* __restrict__ H, double* __restrict__ v, int n, int B) {
  const int r = blockIdx.x;
  if (r == n) {
    for (int c = threadIdx.x; c < n; c += 128) v[c] = 0.0;
    return;
  }
  const int blk = r / 6;
  const int c_lo = max(0, (blk - B) * 6), c_hi = min(n, (blk + B + 1) * 6);
  double* row = H + (size_t)r * n;
  for (int c = c_lo + threadIdx.x; c < c_hi; c += 128) row[c] = 0.0;
}
}  // namespace
}  // namespace dbaf

namespace dbaf {
namespace {
// H rows <-> a dense [n][6 (2B + 1)] strip (row r keeps the columns of pose blocks r/6 - B .. r/6 + B), v behind it
template <bool PACK>
__global__ void __launch_bounds__(256) k_band_strip(double* __restrict__ H, double* __restrict__ v, int n, int B,
                                                    double* __restrict__ strip) {
  const int r = blockIdx.x;
  const int W = 6 * (2 * B + 1);
  if (r == n) {
    for (int c = threadIdx.x; c < n; c += 256) {
      if (PACK) strip[(size_t)n * W + c] = v[c]; else v[c] = strip[(size_t)n * W + c];
    }
    return;
  }
  const int c0 = (r / 6 - B) * 6;
  double* row = H + (size_t)r * n;
  double* srow = strip + (size_t)r * W;
  for (int j = threadIdx.x; j < W; j += 256) {
    const int c = c0 + j;
    const bool in = c >= 0 && c < n;
    if (PACK) srow[j] = in ? row[c] : 0.0;
    else if (in) row[c] = srow[j];
  }
}
}  // namespace
}  // namespace dbaf

extern "C" size_t dbaf_ba_band_strip_doubles(int n_poses, int band_blocks) {
  const size_t n = 6 * (size_t)n_poses;
  return n * 6 * (2 * (size_t)band_blocks + 1) + n;
}

extern "C" int dbaf_ba_band_strip(void* workspace, int t0, int t1, int band_blocks, double* strip, int pack, void* stream) {
  DBAF_REQUIRE(workspace != nullptr && strip != nullptr, "null pointer");
  DBAF_REQUIRE(t1 >= t0 && band_blocks >= 0, "bad sizes");
  dbaf_ba_layout L;
  fill_layout(1, 0, 0, t1 - t0, 1, 1, &L);
  Ws ws = view(workspace, L);
  const int n = 6 * (t1 - t0);
  if (n == 0) return DBAF_OK;
  if (pack) k_band_strip<true><<<n + 1, 256, 0, (cudaStream_t)stream>>>(ws.H, ws.v, n, band_blocks, strip);
  else k_band_strip<false><<<n + 1, 256, 0, (cudaStream_t)stream>>>(ws.H, ws.v, n, band_blocks, strip);
  DBAF_LAUNCH_CHECK("k_band_strip");
  return DBAF_OK;
}

extern "C" int dbaf_ba_zero_system(void* workspace, int t0, int t1, int band_blocks, void* stream) {
  DBAF_REQUIRE(workspace != nullptr, "null workspace");
  DBAF_REQUIRE(t1 >= t0, "t1 < t0");
  dbaf_ba_layout L;
  fill_layout(1, 0, 0, t1 - t0, 1, 1, &L);
  Ws ws = view(workspace, L);
  const int P = t1 - t0, n = 6 * P;
  if (n == 0) return DBAF_OK;
  cudaStream_t st = (cudaStream_t)stream;
  if (band_blocks <= 0 || 2 * band_blocks + 1 >= P) {
    DBAF_CUDA(cudaMemsetAsync(ws.H, 0, ((size_t)n * n + n) * 8, st));
    return DBAF_OK;
  }
  k_zero_band<<<n + 1, 128, 0, st>>>(ws.H, ws.v, n, band_blocks);
  DBAF_LAUNCH_CHECK("k_zero_band");
  return DBAF_OK;
}

extern "C" int dbaf_ba_solve(void* workspace, int t0, int t1, float lm, float ep, const double* prior,
                             int prior_rows, int prior_cols, void* stream) {
  DBAF_REQUIRE(workspace != nullptr, "null workspace");
  DBAF_REQUIRE(t1 >= t0, "t1 < t0");
  return launch_solve(workspace, t0, t1, lm, ep, prior, prior_rows, prior_cols, 0, (cudaStream_t)stream);
}

extern "C" int dbaf_ba_solve_banded(void* workspace, int t0, int t1, float lm, float ep, int band_blocks,
                                    void* stream) {
  DBAF_REQUIRE(workspace != nullptr, "null workspace");
  DBAF_REQUIRE(t1 >= t0, "t1 < t0");
  DBAF_REQUIRE(band_blocks >= 0, "negative band");
  return launch_solve(workspace, t0, t1, lm, ep, nullptr, 0, 0, 0, (cudaStream_t)stream, band_blocks);
}

extern "C" int dbaf_ba_solve_banded_launches(int n_poses, int band_blocks) {
  const int P = n_poses, n = 6 * P;
  if (n == 0) return 0;
  if (n <= kSmemSolveMaxN || band_blocks <= 0 || band_blocks > kBandMaxB) return 1;
  const int B = std::min(band_blocks, P - 1);
  const int D = banded_domains(P, B);
  if (D == 1) return 1;
  if (D == 2) return 3;                                        // two sweeps, separator, two back-substitutions
  int nlev = 0;
  while ((D - 1) >> nlev > 1) ++nlev;
  return 1 + 2 + 2 * nlev + 1 + nlev + 1 + 1;      // sweeps, spike + Gram, levels, last block, level back-substitutions, apply, back
}

extern "C" int dbaf_ba_retract(void* workspace, float* poses, float* disps, const float* dx_in,
                               const int64_t* ii, const int64_t* jj, int E, int N, int t0, int t1,
                               int ht, int wd, int update_poses, int update_disps, int k_lo, int k_hi,
                               float* dx_out, float* dz_out, void* stream) {
  (void)ii;
  DBAF_REQUIRE(workspace != nullptr, "null workspace");
  if (int rc = check_dims(N, E, t0, t1, ht, wd)) return rc;
  dbaf_ba_layout L;
  fill_layout(N, E, t0, t1, ht, wd, &L);
  Ws ws = view(workspace, L);
  cudaStream_t st = (cudaStream_t)stream;
  const int P = t1 - t0, HW = ht * wd;
  const float* dx = dx_in ? dx_in : ws.dx;
  if (k_lo < 0) k_lo = 0;
  const int k_cap = (k_hi >= 0 ? std::min(k_hi, ws.Kcap) : ws.Kcap) - k_lo;
  const bool pose_work = P > 0 && (update_poses || dx_out);
  if (update_disps && k_cap > 0) {
    dim3 grid((HW + kThreads - 1) / kThreads, k_cap);
    k_backsub_fused<<<grid, kThreads, 0, st>>>(jj, dx, disps, dz_out, update_poses ? poses : nullptr,
                                               pose_work ? dx_out : nullptr, t0, P, pose_work ? P : 0, HW,
                                               k_lo, k_hi, ws);
    DBAF_LAUNCH_CHECK("k_backsub_fused");
    if (E >= 7 * ws.Kcap && HW % 32 == 0) {
      // frames whose pose-depth blocks were not stored (flagged on the device by k_linschur_tc)
      k_backsub_rc<<<dim3((HW / 4 + 127) / 128, k_cap), 128, 0, st>>>(dx, disps, dz_out, t0, P, HW, wd, k_lo, k_hi, ws);
      DBAF_LAUNCH_CHECK("k_backsub_rc");
    }
  } else if (pose_work) {
    k_pose_retract<<<1, 128, 0, st>>>(update_poses ? poses : nullptr, dx, t0, P, dx_out);
    DBAF_LAUNCH_CHECK("k_pose_retract");
  }
  return DBAF_OK;
}

extern "C" int dbaf_ba(void* workspace, float* poses, float* disps, const float* intrinsics,
                       const float* disps_sens, const float* targets, const float* weights,
                       const float* eta, int eta_rows, const int64_t* ii, const int64_t* jj, int E,
                       int N, int t0, int t1, int ht, int wd, int iterations, float lm, float ep,
                       int motion_only, float* dx_out, float* dz_out, void* stream) {
  if (int rc = dbaf_ba_prepare(workspace, ii, jj, E, N, t0, t1, ht, wd, stream)) return rc;
  for (int it = 0; it < iterations; ++it) {
    if (int rc = dbaf_ba_linearize(workspace, poses, disps, intrinsics, disps_sens, targets, weights,
                                   eta, eta_rows, ii, jj, E, N, t0, t1, ht, wd, 0.05f, motion_only,
                                   it == 0 ? 1 : 0, 0, -1, stream))
      return rc;
    if (int rc = launch_solve(workspace, t0, t1, lm, ep, nullptr, 0, 0, it + 1 < iterations, (cudaStream_t)stream)) return rc;
    const bool last = it == iterations - 1;
    if (int rc = dbaf_ba_retract(workspace, poses, disps, nullptr, ii, jj, E, N, t0, t1, ht, wd, 1,
                                 motion_only ? 0 : 1, 0, -1, last ? dx_out : nullptr,
                                 last ? dz_out : nullptr, stream))
      return rc;
  }
  return DBAF_OK;
}

// ------------------------------------------------------- H/v hand-off to the factor graph ----
// BA2GTSAM, dbaf/depth_video.py:20-29: H' = J^T H J, v' = J^T v with J = blockdiag(A), A a 6x6 matrix (the
// row-swapped -Ad(Tbc^-1) that carries camera-frame tangents to the body frame).  The reference does this with
// two dense 6P x 6P host matmuls per iteration; per 6x6 block it is A^T H_IK A, done here on the device on the
// way out, so that one copy delivers the system in the form the factor graph wants.
namespace dbaf {
namespace {
__global__ void __launch_bounds__(64)
k_congruence(const double* __restrict__ H, const double* __restrict__ v, const double* __restrict__ A, int P,
             double* __restrict__ Ho, double* __restrict__ vo) {
  __shared__ double As[36], Hb[36], Tb[36];
  const int I = blockIdx.y, K = blockIdx.x, t = threadIdx.x;
  const int n = 6 * P;
  if (t < 36) {
    As[t] = A[t];
    Hb[t] = H[(size_t)(6 * I + t / 6) * n + 6 * K + t % 6];
  }
  __syncthreads();
  if (t < 36) {                        // T = H_IK A
    const int r = t / 6, c = t % 6;
    double s = 0.0;
#pragma unroll
    for (int b = 0; b < 6; ++b) s += Hb[r * 6 + b] * As[b * 6 + c];
    Tb[t] = s;
  }
  __syncthreads();
  if (t < 36) {                        // A^T T
    const int r = t / 6, c = t % 6;
    double s = 0.0;
#pragma unroll
    for (int a = 0; a < 6; ++a) s += As[a * 6 + r] * Tb[a * 6 + c];
    Ho[(size_t)(6 * I + r) * n + 6 * K + c] = s;
  } else if (K == 0 && t < 42) {
    const int r = t - 36;
    double s = 0.0;
#pragma unroll
    for (int a = 0; a < 6; ++a) s += As[a * 6 + r] * v[6 * I + a];
    vo[6 * I + r] = s;
  }
}
}  // namespace
}  // namespace dbaf

extern "C" int dbaf_ba_congruence(void* workspace, int t0, int t1, const double* A, double* H_out, double* v_out,
                                  void* stream) {
  DBAF_REQUIRE(workspace != nullptr && A != nullptr && H_out != nullptr && v_out != nullptr, "null pointer");
  DBAF_REQUIRE(t1 >= t0, "t1 < t0");
  const int P = t1 - t0;
  if (P == 0) return DBAF_OK;
  dbaf_ba_layout L;
  fill_layout(1, 0, 0, P, 1, 1, &L);          // H and v offsets depend on P alone
  Ws ws = view(workspace, L);
  k_congruence<<<dim3(P, P), 64, 0, (cudaStream_t)stream>>>(ws.H, ws.v, A, P, H_out, v_out);
  DBAF_LAUNCH_CHECK("k_congruence");
  return DBAF_OK;
}

// ------------------------------------------------------- host-buffer entry point ----
namespace {
struct HostStage {
  size_t poses, disps, intr, sens, targets, weights, eta, ii, jj, dx, total;
};
HostStage host_stage_layout(int N, int E, int t0, int t1, int ht, int wd, int eta_rows) {
  HostStage L{};
  size_t off = 0;
  auto take = [&](size_t bytes) { const size_t o = off; off += (bytes + 255) & ~(size_t)255; return o; };
  const size_t HW = (size_t)ht * wd;
  L.poses = take((size_t)N * 7 * 4);
  L.disps = take((size_t)N * HW * 4);
  L.intr = take(16);
  L.sens = take((size_t)N * HW * 4);
  L.targets = take((size_t)E * 2 * HW * 4);
  L.weights = take((size_t)E * 2 * HW * 4);
  L.eta = take((size_t)std::max(eta_rows, 1) * HW * 4);
  L.ii = take((size_t)E * 8);
  L.jj = take((size_t)E * 8);
  L.dx = take((size_t)std::max(t1 - t0, 1) * 6 * 4);
  L.total = off;
  return L;
}
}  // namespace

extern "C" size_t dbaf_ba_host_staging_bytes(int n_frames, int n_edges, int t0, int t1, int ht, int wd, int eta_rows) {
  return host_stage_layout(n_frames, n_edges, t0, t1, ht, wd, eta_rows).total;
}

extern "C" int dbaf_ba_host(void* workspace, void* staging, size_t staging_bytes, float* poses_h, float* disps_h,
                            const float* intrinsics_h, const float* disps_sens_h, const float* targets_h,
                            const float* weights_h, const float* eta_h, int eta_rows, const int64_t* ii_h,
                            const int64_t* jj_h, int E, int N, int t0, int t1, int ht, int wd, int iterations,
                            float lm, float ep, int motion_only, float* dx_out_h, int32_t* info_out_h, void* stream) {
  DBAF_REQUIRE(workspace != nullptr && staging != nullptr, "null workspace or staging buffer");
  if (int rc = check_dims(N, E, t0, t1, ht, wd)) return rc;
  DBAF_REQUIRE(poses_h && disps_h && intrinsics_h && disps_sens_h && ii_h && jj_h && (E == 0 || (targets_h && weights_h)),
               "null host input");
  DBAF_REQUIRE(motion_only == 1 || (eta_h != nullptr && eta_rows >= 1), "eta needs at least one row");
  const HostStage L = host_stage_layout(N, E, t0, t1, ht, wd, eta_rows);
  DBAF_REQUIRE(staging_bytes >= L.total, "staging buffer smaller than dbaf_ba_host_staging_bytes()");
  cudaStream_t st = (cudaStream_t)stream;
  char* d = static_cast<char*>(staging);
  const size_t HW = (size_t)ht * wd;
  const int P = t1 - t0;
  auto up = [&](size_t off, const void* src, size_t bytes) {
    return bytes == 0 ? cudaSuccess : cudaMemcpyAsync(d + off, src, bytes, cudaMemcpyHostToDevice, st);
  };
  DBAF_CUDA(up(L.poses, poses_h, (size_t)N * 7 * 4));
  DBAF_CUDA(up(L.disps, disps_h, (size_t)N * HW * 4));
  DBAF_CUDA(up(L.intr, intrinsics_h, 16));
  DBAF_CUDA(up(L.sens, disps_sens_h, (size_t)N * HW * 4));
  DBAF_CUDA(up(L.targets, targets_h, (size_t)E * 2 * HW * 4));
  DBAF_CUDA(up(L.weights, weights_h, (size_t)E * 2 * HW * 4));
  if (eta_h != nullptr) DBAF_CUDA(up(L.eta, eta_h, (size_t)eta_rows * HW * 4));
  DBAF_CUDA(up(L.ii, ii_h, (size_t)E * 8));
  DBAF_CUDA(up(L.jj, jj_h, (size_t)E * 8));
  float* poses_d = reinterpret_cast<float*>(d + L.poses);
  float* disps_d = reinterpret_cast<float*>(d + L.disps);
  float* dx_d = reinterpret_cast<float*>(d + L.dx);
  if (int rc = dbaf_ba(workspace, poses_d, disps_d, reinterpret_cast<const float*>(d + L.intr),
                       reinterpret_cast<const float*>(d + L.sens), reinterpret_cast<const float*>(d + L.targets),
                       reinterpret_cast<const float*>(d + L.weights), reinterpret_cast<const float*>(d + L.eta), eta_rows,
                       reinterpret_cast<const int64_t*>(d + L.ii), reinterpret_cast<const int64_t*>(d + L.jj), E, N, t0, t1,
                       ht, wd, iterations, lm, ep, motion_only, dx_out_h ? dx_d : nullptr, nullptr, stream))
    return rc;
  if (iterations > 0) {
    if (P > 0) DBAF_CUDA(cudaMemcpyAsync(poses_h + (size_t)t0 * 7, poses_d + (size_t)t0 * 7, (size_t)P * 7 * 4, cudaMemcpyDeviceToHost, st));
    if (!motion_only) DBAF_CUDA(cudaMemcpyAsync(disps_h, disps_d, (size_t)N * HW * 4, cudaMemcpyDeviceToHost, st));
    if (dx_out_h && P > 0) DBAF_CUDA(cudaMemcpyAsync(dx_out_h, dx_d, (size_t)P * 6 * 4, cudaMemcpyDeviceToHost, st));
    // the status words sit at the start of the workspace (dbaf_ba_layout.info == 0)
    if (info_out_h) DBAF_CUDA(cudaMemcpyAsync(info_out_h, workspace, 16 * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
  }
  return DBAF_OK;
}

// ------------------------------------------------------- allocation-free graph tables ----
//
// dbaf_ba_tables_noalloc: the graph bookkeeping kernel (k_tables) for hosts that must not
// allocate device memory around the call.  Two kinds of caller are in that position: a
// real-time tracking thread that has handed its whole memory budget out at start-up and may not
// touch an allocator afterwards (cudaMalloc takes a driver lock and synchronises the whole
// device, which stalls the camera pipeline running next to it), and a plugin whose host
// application owns every buffer and only passes pointers in.  Such a caller brings the inputs
// (ii, jj) and the outputs; the intermediate tables the kernel needs live in the fixed array
// below, which is part of the library image: every GPU that loads libdbaf_b200.so holds these
// 1.5 MiB until the process exits.
//
// Limits (checked, DBAF_EINVAL with a message otherwise): n_frames <= 16384, n_edges <= 262144.
//
// The first call on each device looks up the device address of the array
// (cudaGetSymbolAddress, a host-side set-up step, not an allocation).  Make that first call
// before starting any stream capture.  All calls on one device share the array and it keeps
// whatever the previous call left in it; the kernel re-initialises every word it later reads,
// so stale contents are harmless, but two calls that overlap on the GPU (different streams)
// race on it -- serialise them.
namespace dbaf {
namespace {

constexpr int kNoAllocMaxFrames = 16384;
constexpr int kNoAllocMaxEdges = 262144;
// frame2k[N] + cursor[N+1] + srow_ptr[N+1] + srow_idx[N+E] int32, rounded up
constexpr size_t kNoAllocScratchBytes = 1536 * 1024;
static_assert(4ull * (4ull * kNoAllocMaxFrames + kNoAllocMaxEdges + 2) + 3 * 256 <= kNoAllocScratchBytes,
              "static scratch too small for the advertised limits");

static __device__ __align__(256) unsigned char g_tables_scratch[kNoAllocScratchBytes];

constexpr int kMaxDevices = 64;
std::atomic<unsigned char*> g_tables_scratch_addr[kMaxDevices];   // device address per device, filled on first use

// The array's name is only a host-side handle here; the kernel needs the address the symbol has
// on the device that is current right now.
int tables_scratch_for_current_device(unsigned char** out) {
  int dev = 0;
  DBAF_CUDA(cudaGetDevice(&dev));
  DBAF_REQUIRE(dev >= 0 && dev < kMaxDevices, "device ordinal out of range");
  unsigned char* addr = g_tables_scratch_addr[dev].load(std::memory_order_acquire);
  if (addr == nullptr) {
    // two host threads may both get here; they resolve the same address, so the second store is harmless
    void* p = nullptr;
    DBAF_CUDA(cudaGetSymbolAddress(&p, g_tables_scratch));
    DBAF_REQUIRE(p != nullptr, "cudaGetSymbolAddress returned a null address");
    addr = static_cast<unsigned char*>(p);
    g_tables_scratch_addr[dev].store(addr, std::memory_order_release);
  }
  *out = addr;
  return DBAF_OK;
}

}  // namespace
}  // namespace dbaf

extern "C" int dbaf_ba_tables_noalloc(const int64_t* ii, const int64_t* jj, int n_edges, int n_frames,
                                      int t0, int t1, int32_t* info, int32_t* kx, int32_t* kk_exp,
                                      int32_t* csr_ptr, int32_t* csr_idx, void* stream) {
  if (int rc = check_dims(n_frames, n_edges, t0, t1, 1, 1)) return rc;
  DBAF_REQUIRE(n_frames <= kNoAllocMaxFrames,
               "n_frames exceeds the 16384-frame limit of the allocation-free entry point");
  DBAF_REQUIRE(n_edges <= kNoAllocMaxEdges,
               "n_edges exceeds the 262144-edge limit of the allocation-free entry point");
  DBAF_REQUIRE(info && kx && kk_exp && csr_ptr && (csr_idx || n_edges == 0), "null output pointer");
  unsigned char* scratch = nullptr;
  if (int rc = tables_scratch_for_current_device(&scratch)) return rc;
  const int P = t1 - t0;
  const int Kcap = std::min(n_frames, P + n_edges);
  auto carve = [&](size_t count) {
    unsigned char* p = scratch;
    scratch += align_up((int64_t)count * 4, 256);
    return reinterpret_cast<int*>(p);
  };
  Ws ws{};
  ws.info = info;
  ws.kx = kx;
  ws.kk_exp = kk_exp;
  ws.csr_ptr = csr_ptr;
  ws.csr_idx = csr_idx;
  ws.frame2k = carve(n_frames);
  ws.cursor = carve(Kcap + 1);
  ws.srow_ptr = carve(Kcap + 1);
  ws.srow_idx = carve(P + n_edges);
  ws.Kcap = Kcap;
  // k_tables writes every table word before reading it (flags, counters and cursors are reset at
  // the top of the kernel), so nothing from an earlier call can leak into this one.
  k_tables<<<1, 1024, 0, (cudaStream_t)stream>>>(ii, jj, n_edges, n_frames, t0, t1, ws);
  DBAF_LAUNCH_CHECK("k_tables (noalloc)");
  return DBAF_OK;
}
