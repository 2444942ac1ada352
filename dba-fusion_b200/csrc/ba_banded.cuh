// Block-banded float64 Cholesky for large windows (global / backend BA), included by ba.cu inside
// namespace dbaf::{anonymous} after ba_fast.cuh.
//
// A temporal keyframe graph whose edges reach at most `reach` frames couples poses at most
// B = 2*reach blocks apart in the reduced camera system (the Schur complement joins every pair of
// poses that observe the same depth frame).  The factorisation then only ever touches a
// (B+1) x (B+1)-block window that slides down the diagonal: one CTA keeps that window in shared
// memory (circularly indexed, so sliding costs nothing), streams the next block row in from the
// dense H while it works, and writes each finished panel column to a compact band store that the
// back-substitution walks in reverse.  6000 x 6000 with B = 20: 1000 steps of ~45k FMAs instead
// of 72 GFLOP dense.  The reference solves this system on one CPU core with Eigen::SimplicialLLT
// (src/droid_kernels.cu:1248-1269).
#pragma once

constexpr int kBandThreads = 512;
constexpr int kBandMaxB = 25;      // (B+2)^2 blocks * 288 B must fit in shared memory

// 1/sqrt(d) from the single-precision hardware estimate plus two Newton steps in double (relative
// error ~2e-16); the library rsqrt(double) carries special-case handling that sits on the critical
// path of the pivot chain.
__device__ __forceinline__ double rsqrt_nr(double d) {
  if (!(d > 1e-30 && d < 1e30)) return rsqrt(d);
  double y = (double)rsqrtf((float)d);
  y = y * (1.5 - 0.5 * d * y * y);
  y = y * (1.5 - 0.5 * d * y * y);
  return y;
}

__device__ __forceinline__ double warp_sum_d(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Lband: [P][B+1][36] (block d = I-J of column J, row-major 6x6), then dinv [6P], then y [6P].
__global__ void __launch_bounds__(kBandThreads)
k_solve_banded(int P, int B, float lm, float ep, int rezero, Ws ws) {
  extern __shared__ double smb[];
  const int tid = threadIdx.x, nt = blockDim.x, lane = tid & 31, wid = tid >> 5;
  const int n = 6 * P;
  const int NB = B + 2;
  double* Wd = smb;                                  // NB x NB blocks of 36
  double* yw = Wd + (size_t)NB * NB * 36;            // NB x 6 (right-hand side window, later x window)
  double* Ljj = yw + NB * 6;                         // 36
  double* dj = Ljj + 72;                             // 2 x 6 (1 / diag), double-buffered like Ljj (2 x 36)
  double* sred = dj + 12;                            // 8 (back-substitution partial sums)
  __shared__ int fail;
  double* Lband = ws.Lfac;
  double* dinv_g = Lband + (size_t)P * (B + 1) * 36;
  double* y_g = dinv_g + n;
  auto slot = [&](int I, int K) { return Wd + ((size_t)(I % NB) * NB + (K % NB)) * 36; };
  const double epd = (double)ep, lmd = (double)lm;

  // element e of block row I: blocks K = max(0, I-B) .. I, then the 6 rhs entries
  auto row_elems = [&](int I) { return (min(I, B) + 1) * 36 + 6; };
  auto fetch = [&](int I, int e) -> double {
    const int nb = min(I, B) + 1;
    if (e < nb * 36) {
      const int b = e / 36, rc = e - b * 36, r = rc / 6, c = rc - r * 6;
      const int K = I - (nb - 1) + b;
      double a = ws.H[(size_t)(6 * I + r) * n + 6 * K + c];
      if (K == I && r == c) a += epd + lmd * a;
      return a;
    }
    return ws.v[6 * I + (e - nb * 36)];
  };
  auto place = [&](int I, int e, double a) {
    const int nb = min(I, B) + 1;
    if (e < nb * 36) {
      const int b = e / 36;
      slot(I, I - (nb - 1) + b)[e - b * 36] = a;
    } else {
      yw[(I % NB) * 6 + (e - nb * 36)] = a;
    }
  };

  // Factorises the (already updated) diagonal block J held in the window, forward-solves its
  // right-hand side and publishes L_JJ / 1/diag into buffer J & 1.  One thread, all in registers.
  // `upd` = the panel block L_{J,J-1} whose outer product still has to be subtracted (look-ahead
  // call from step J-1), or nullptr.
  auto factor_diag = [&](int J, const double* upd) {
    double* A = slot(J, J);
    double a[6][6];
#pragma unroll
    for (int r = 0; r < 6; ++r)
#pragma unroll
      for (int c = 0; c < 6; ++c)
        if (c <= r) a[r][c] = A[r * 6 + c];
    if (upd) {
      double l[6][6];
#pragma unroll
      for (int r = 0; r < 6; ++r)
#pragma unroll
        for (int c = 0; c < 6; ++c) l[r][c] = upd[r * 6 + c];
#pragma unroll
      for (int r = 0; r < 6; ++r)
#pragma unroll
        for (int c = 0; c < 6; ++c)
          if (c <= r) {
            double s = 0.0;
#pragma unroll
            for (int k = 0; k < 6; ++k) s += l[r][k] * l[c][k];
            a[r][c] -= s;
          }
    }
    bool bad = false;
    double di[6];
#pragma unroll
    for (int c = 0; c < 6; ++c) {
      double d = a[c][c];
#pragma unroll
      for (int k = 0; k < 6; ++k)
        if (k < c) d -= a[c][k] * a[c][k];
      if (d <= 0.0) { bad = true; d = 1.0; }
      const double inv = rsqrt_nr(d);
      a[c][c] = d * inv;
      di[c] = inv;
#pragma unroll
      for (int r = 0; r < 6; ++r)
        if (r > c) {
          double s = a[r][c];
#pragma unroll
          for (int k = 0; k < 6; ++k)
            if (k < c) s -= a[r][k] * a[c][k];
          a[r][c] = s * inv;
        }
    }
    if (bad) fail = 1;
    double y[6];
#pragma unroll
    for (int r = 0; r < 6; ++r) {
      double s = yw[(J % NB) * 6 + r];
#pragma unroll
      for (int k = 0; k < 6; ++k)
        if (k < r) s -= a[r][k] * y[k];
      y[r] = s * di[r];
    }
    double* Lj = Ljj + (J & 1) * 36;
    double* dd = dj + (J & 1) * 6;
#pragma unroll
    for (int r = 0; r < 6; ++r) {
      yw[(J % NB) * 6 + r] = y[r];
      y_g[6 * J + r] = y[r];
      dd[r] = di[r];
      dinv_g[6 * J + r] = di[r];
#pragma unroll
      for (int c = 0; c < 6; ++c) {
        const double l = c <= r ? a[r][c] : 0.0;
        Lj[r * 6 + c] = l;
        Lband[((size_t)J * (B + 1)) * 36 + r * 6 + c] = l;
      }
    }
  };

  if (tid == 0) fail = 0;
  for (int I = 0; I <= min(B, P - 1); ++I)
    for (int e = tid; e < row_elems(I); e += nt) place(I, e, fetch(I, e));
  __syncthreads();
  if (tid == 0) factor_diag(0, nullptr);
  __syncthreads();

  const int nwork = nt - 32;                         // the last warp runs the look-ahead factorisation
  for (int J = 0; J < P; ++J) {
    if (fail) break;
    const int last = min(J + B, P - 1);              // last block row inside the band of column J
    const int nbelow = last - J;
    const double* Lj = Ljj + (J & 1) * 36;
    const double* dd = dj + (J & 1) * 6;
    // the block row that enters the window after this step is requested now and stored at the end
    const int Inew = J + B + 1;
    double pre[2] = {0.0, 0.0};
    if (Inew < P) {
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        const int e = tid + u * nt;
        if (e < row_elems(Inew)) pre[u] = fetch(Inew, e);
      }
    }
    // (2) panel: every row (I, r) below the diagonal block: X L_JJ^T = A, and its rhs entry
    for (int t = tid; t < nbelow * 6; t += nt) {
      const int I = J + 1 + t / 6, r = t % 6;
      double* A = slot(I, J) + r * 6;
      double x[6];
#pragma unroll
      for (int c = 0; c < 6; ++c) {
        double s = A[c];
#pragma unroll
        for (int k = 0; k < 6; ++k)
          if (k < c) s -= x[k] * Lj[c * 6 + k];
        x[c] = s * dd[c];
      }
      double yy = yw[(I % NB) * 6 + r];
#pragma unroll
      for (int c = 0; c < 6; ++c) {
        A[c] = x[c];
        Lband[((size_t)J * (B + 1) + (I - J)) * 36 + r * 6 + c] = x[c];
        yy -= x[c] * yw[(J % NB) * 6 + c];
      }
      yw[(I % NB) * 6 + r] = yy;
    }
    __syncthreads();
    // (3) trailing update inside the window: A_IK -= L_IJ L_KJ^T for J < K <= I <= last.  4 threads
    //     share a row (I, r) and split the K range.  The next diagonal block (J+1, J+1) is left to the
    //     last warp, which updates AND factorises it at the same time (look-ahead), so the serial
    //     rsqrt chain of the factorisation overlaps the bulk of the update.
    if (tid < nwork) {
      // one thread = half (3 rows) of one 6x6 block (I, K): 54 shared-memory loads feed 108 FMAs
      const int jm = J % NB;
      const int npairs = nbelow * (nbelow + 1) / 2;
      for (int t = tid; t < 2 * npairs; t += nwork) {
        const int pair = t >> 1, half = t & 1;
        if (pair == 0) continue;                     // block (J+1, J+1): look-ahead warp
        int Io, Ko;
        tri_unpack(pair, Io, Ko);
        int im = jm + 1 + Io, km = jm + 1 + Ko;
        if (im >= NB) im -= NB;
        if (km >= NB) km -= NB;
        const double* Li = Wd + ((size_t)im * NB + jm) * 36 + half * 18;
        const double* Lk = Wd + ((size_t)km * NB + jm) * 36;
        double* A = Wd + ((size_t)im * NB + km) * 36 + half * 18;
        double li[3][6], acc[3][6];
#pragma unroll
        for (int r = 0; r < 3; ++r)
#pragma unroll
          for (int k = 0; k < 6; ++k) li[r][k] = Li[r * 6 + k];
#pragma unroll
        for (int c = 0; c < 6; ++c) {
          double lk[6];
#pragma unroll
          for (int k = 0; k < 6; ++k) lk[k] = Lk[c * 6 + k];
#pragma unroll
          for (int r = 0; r < 3; ++r) {
            double sum = 0.0;
#pragma unroll
            for (int k = 0; k < 6; ++k) sum += li[r][k] * lk[k];
            acc[r][c] = sum;
          }
        }
#pragma unroll
        for (int r = 0; r < 3; ++r)
#pragma unroll
          for (int c = 0; c < 6; ++c) A[r * 6 + c] -= acc[r][c];
      }
    } else if (tid == nwork && J + 1 < P) {
      factor_diag(J + 1, nbelow > 0 ? slot(J + 1, J) : nullptr);
    }
    // incoming block row (its slots are unused during this step: NB = B + 2)
    if (Inew < P) {
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        const int e = tid + u * nt;
        if (e < row_elems(Inew)) place(Inew, e, pre[u]);
      }
    }
    __syncthreads();
  }
  __syncthreads();
  const bool ok = fail == 0;
  if (ok) {
    // back-substitution L^T x = y, bottom up; yw becomes the window of already-solved x blocks
    for (int J = P - 1; J >= 0; --J) {
      const int nbelow = min(J + B, P - 1) - J;
      // s_c = y_J[c] - sum_{d=1..nbelow} sum_r L_{J+d,J}[r][c] * x_{J+d}[r]; warp c owns component c
      if (wid < 6) {
        double s = 0.0;
        for (int t = lane; t < nbelow * 6; t += 32) {
          const int d = 1 + t / 6, r = t % 6;
          s += Lband[((size_t)J * (B + 1) + d) * 36 + r * 6 + wid] * yw[((J + d) % NB) * 6 + r];
        }
        s = warp_sum_d(s);
        if (lane == 0) sred[wid] = y_g[6 * J + wid] - s;
      }
      __syncthreads();
      if (tid == 0) {
        const double* L = Lband + ((size_t)J * (B + 1)) * 36;
        double x[6];
#pragma unroll
        for (int c = 5; c >= 0; --c) {
          double s = sred[c];
#pragma unroll
          for (int k = 0; k < 6; ++k)
            if (k > c) s -= L[k * 6 + c] * x[k];
          x[c] = s * dinv_g[6 * J + c];
        }
#pragma unroll
        for (int c = 0; c < 6; ++c) {
          yw[(J % NB) * 6 + c] = x[c];
          ws.dx[6 * J + c] = (float)x[c];
        }
      }
      __syncthreads();
    }
  } else {
    for (int i = tid; i < n; i += nt) ws.dx[i] = 0.0f;
    if (tid == 0) atomicAdd(&ws.info[1], 1);
  }
  (void)rezero;
}
