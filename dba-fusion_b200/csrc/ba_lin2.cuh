// Second-generation linearisation kernel, included by ba.cu inside namespace dbaf::{anonymous}.
//
// The pose-i Jacobian of a residual is a fixed linear image of its pose-j Jacobian for all pixels of
// an edge:  Ji = -M Jj  with  M = [[R^T, 0], [-R^T [t]x, R^T]]  (adjSE3 of the reference,
// src/droid_kernels.cu:82-97,369-370, written as a matrix).  So per pixel only the 21 unique terms of
// Hjj = sum w Jj Jj^T and the 6 of vj are accumulated (27 values through one 31-shuffle transposed
// warp reduction instead of 90 through three), and once per edge
//   Hii = M Hjj M^T,  Hij = -M Hjj,  Hji = Hij^T,  vi = -M vj
// are formed in float64 from the block sums.  The per-pixel pose-i rows of E follow the same way,
// Ei_px = -M Ej_px.  The per-edge relative pose and M are computed once per CTA into shared memory.
#pragma once

struct EdgeXf {
  float t[3], q[4];     // Gij
  float Rt[9];          // R(qij)^T, row-major
  float Nm[9];          // -R^T [t]x
  int jx;
  int stereo;
};

__device__ __forceinline__ void edge_xf(const float* __restrict__ poses, int ix, int jx, EdgeXf& X) {
  const bool stereo = ix == jx;
  const Se3 G = stereo ? stereo_baseline() : load_relative(poses, ix, jx);
#pragma unroll
  for (int k = 0; k < 3; ++k) X.t[k] = G.t[k];
#pragma unroll
  for (int k = 0; k < 4; ++k) X.q[k] = G.q[k];
  X.jx = jx;
  X.stereo = stereo ? 1 : 0;
  // M = [[R^T, 0], [N, R^T]]: column c of R^T is the inverse rotation of e_c, column c of N the inverse
  // rotation of e_c x t (the cross term of the reference's adjSE3)
  const float qinv[4] = {-G.q[0], -G.q[1], -G.q[2], G.q[3]};
  const float ex[3][3] = {{1.f, 0.f, 0.f}, {0.f, 1.f, 0.f}, {0.f, 0.f, 1.f}};
  const float ct[3][3] = {{0.f, -G.t[2], G.t[1]}, {G.t[2], 0.f, -G.t[0]}, {-G.t[1], G.t[0], 0.f}};
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    float y[3], z[3];
    rotate(qinv, ex[c], y);
    rotate(qinv, ct[c], z);
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      X.Rt[r * 3 + c] = y[r];
      X.Nm[r * 3 + c] = z[r];
    }
  }
}

constexpr int kLinChunk = 32;   // edges whose transforms are staged per pass

template <int PXT, bool MOTION_ONLY>
__global__ void __launch_bounds__(kThreads)
k_linearize2(const float* __restrict__ poses, const float* __restrict__ disps,
             const float* __restrict__ intr, const float* __restrict__ sens,
             const float* __restrict__ targets, const float* __restrict__ weights,
             const float* __restrict__ eta, int eta_rows, const int64_t* __restrict__ jj,
             int t0, int P, int HW, int wd, float alpha, int k_lo, int k_hi, int skip_min_m, Ws ws) {
  const int k = k_lo + blockIdx.y;
  const int K = ws.info[0];
  if (k >= K || (k_hi >= 0 && k >= k_hi)) return;
  if (skip_min_m > 0 && ls_eligible(ws, k, skip_min_m)) return;   // taken by k_linschur_tc
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  if (!MOTION_ONLY && blockIdx.x == 0 && tid == 0) ws.fused[k] = 0;   // this frame's Eij is (again) in the workspace
  const int ix = ws.kx[k];
  const float fx = __ldg(intr), fy = __ldg(intr + 1), cx = __ldg(intr + 2), cy = __ldg(intr + 3);
  const int n6 = 6 * P;
  const int self = ix - t0;
  const bool has_self = self >= 0 && self < P;

  __shared__ EdgeXf sx[kLinChunk];
  __shared__ float part[kLinChunk][kWarps][28];      // per-warp sums of the 27 terms of every edge of the chunk
  __shared__ float tot[kLinChunk][28];
  __shared__ float Tm[kLinChunk][36];
  __shared__ int sedge[kLinChunk + 1];               // edge ids of the chunk and the first one of the next chunk

  int px[PXT];
  bool ok[PXT];
  float X0[PXT], X1[PXT], d0[PXT];
  float accC[PXT], accW[PXT], accE[PXT][6];
#pragma unroll
  for (int s = 0; s < PXT; ++s) {
    px[s] = (blockIdx.x * PXT + s) * kThreads + tid;
    ok[s] = px[s] < HW;
    const int p = ok[s] ? px[s] : 0;
    const int i = p / wd, j = p - i * wd;
    X0[s] = ((float)j - cx) / fx;
    X1[s] = ((float)i - cy) / fy;
    d0[s] = ok[s] ? __ldg(disps + (size_t)ix * HW + p) : 0.f;
    accC[s] = 0.f; accW[s] = 0.f;
#pragma unroll
    for (int c = 0; c < 6; ++c) accE[s][c] = 0.f;
  }

  const int e_lo = ws.csr_ptr[k], e_hi = ws.csr_ptr[k + 1];
  // targets and weights of the next edge are requested while the current edge is being worked on (the
  // kernel is otherwise exposed to one full memory latency per edge): tw[s] = {wu, wv, tu, tv}
  float4 tw_next[PXT];
  auto fetch_edge = [&](int pos, int e) {
    if (pos < e_hi) {
#pragma unroll
      for (int s = 0; s < PXT; ++s) {
        const size_t o = ((size_t)e * 2) * HW + (ok[s] ? px[s] : 0);
        tw_next[s] = make_float4(__ldg(weights + o), __ldg(weights + o + HW), __ldg(targets + o), __ldg(targets + o + HW));
      }
    }
  };
  fetch_edge(e_lo, e_lo < e_hi ? ws.csr_idx[e_lo] : 0);
  for (int base = e_lo; base < e_hi; base += kLinChunk) {
    const int cnt = min(kLinChunk, e_hi - base);
    __syncthreads();                                   // previous chunk's transforms no longer in use
    if (tid <= cnt && base + tid < e_hi) {
      const int eid = ws.csr_idx[base + tid];
      sedge[tid] = eid;
      if (tid < cnt) edge_xf(poses, ix, (int)jj[eid], sx[tid]);
    }
    __syncthreads();
    for (int a = 0; a < cnt; ++a) {
      const int e = sedge[a];
      const EdgeXf& X = sx[a];
      const bool stereo = X.stereo != 0;
      Se3 G;
#pragma unroll
      for (int q = 0; q < 3; ++q) G.t[q] = X.t[q];
#pragma unroll
      for (int q = 0; q < 4; ++q) G.q[q] = X.q[q];

      float h[32];     // [0,21): lower triangle of Hjj row by row, [21,27): vj
#pragma unroll
      for (int l = 0; l < 32; ++l) h[l] = 0.f;
      float4 tw_cur[PXT];
#pragma unroll
      for (int s = 0; s < PXT; ++s) tw_cur[s] = tw_next[s];
      fetch_edge(base + a + 1, sedge[a + 1]);

#pragma unroll
      for (int s = 0; s < PXT; ++s) {
        if (!ok[s]) continue;
        const int p = px[s];
        const float Xi[4] = {X0[s], X1[s], 1.0f, d0[s]};
        float Xj[4];
        act(G, Xi, Xj);
        const float x = Xj[0], y = Xj[1], hh = Xj[3];
        const bool near = Xj[2] < kMinDepth;
        const float d = near ? 0.0f : 1.0f / Xj[2];
        const float d2 = d * d;
        float wu = near ? 0.0f : (float)(0.001 * (double)tw_cur[s].x);
        float wv = near ? 0.0f : (float)(0.001 * (double)tw_cur[s].y);
        const float ru = tw_cur[s].z - (fx * d * x + cx);
        const float rv = tw_cur[s].w - (fy * d * y + cy);
        // Ju[1] and Jv[0] are structural zeros (src/droid_kernels.cu:356-361,389-394): their terms are skipped
        const float Ju[6] = {fx * (hh * d), 0.f, fx * (-x * hh * d2), fx * (-x * y * d2),
                             fx * (1.0f + x * x * d2), fx * (-y * d)};
        const float Jv[6] = {0.f, fy * (hh * d), fy * (-y * hh * d2), fy * (-1.0f - y * y * d2),
                             fy * (x * y * d2), fy * (x * d)};
        const float Jzu = fx * (G.t[0] * d - G.t[2] * (x * d2));
        const float Jzv = fy * (G.t[1] * d - G.t[2] * (y * d2));
        if (!MOTION_ONLY) {
          accC[s] += wu * Jzu * Jzu + wv * Jzv * Jzv;
          accW[s] += wu * ru * Jzu + wv * rv * Jzv;
        }
        if (stereo) { wu = 0.f; wv = 0.f; }            // stereo edges feed the depth blocks only
        {
          float wJu[6], wJv[6];
#pragma unroll
          for (int n = 0; n < 6; ++n) { wJu[n] = wu * Ju[n]; wJv[n] = wv * Jv[n]; }
          int l = 0;
#pragma unroll
          for (int n = 0; n < 6; ++n)
#pragma unroll
            for (int m = 0; m <= n; ++m) {
              if (n != 1 && m != 1) h[l] = fmaf(wJu[n], Ju[m], h[l]);
              if (n != 0 && m != 0) h[l] = fmaf(wJv[n], Jv[m], h[l]);
              ++l;
            }
#pragma unroll
          for (int n = 0; n < 6; ++n) {
            if (n != 1) h[21 + n] = fmaf(wJu[n], ru, h[21 + n]);
            if (n != 0) h[21 + n] = fmaf(wJv[n], rv, h[21 + n]);
          }
        }
        if (!MOTION_ONLY) {
          const float wzu = wu * Jzu, wzv = wv * Jzv;
          float Ej[6];
          float* Eo = ws.Eij + (size_t)e * 6 * HW + p;
#pragma unroll
          for (int n = 0; n < 6; ++n) {
            Ej[n] = n == 0 ? wzu * Ju[0] : n == 1 ? wzv * Jv[1] : wzu * Ju[n] + wzv * Jv[n];
            Eo[(size_t)n * HW] = Ej[n];
          }
          if (has_self) {
            // Ei_px = -M Ej_px
#pragma unroll
            for (int r = 0; r < 3; ++r) {
              const float top = X.Rt[r * 3] * Ej[0] + X.Rt[r * 3 + 1] * Ej[1] + X.Rt[r * 3 + 2] * Ej[2];
              const float bot = X.Rt[r * 3] * Ej[3] + X.Rt[r * 3 + 1] * Ej[4] + X.Rt[r * 3 + 2] * Ej[5] +
                                X.Nm[r * 3] * Ej[0] + X.Nm[r * 3 + 1] * Ej[1] + X.Nm[r * 3 + 2] * Ej[2];
              accE[s][r] -= top;
              accE[s][3 + r] -= bot;
            }
          }
        }
      }

      // per-warp sums of the 27 terms (no block barrier inside the edge loop)
      warp_reduce32(h, lane);
      if (lane < 27) part[a][wid][lane] = h[0];
    }
    // ---- chunk epilogue, all threads: block sums, T = M Hjj, then the four pose blocks and two gradients
    __syncthreads();
    for (int o = tid; o < cnt * 27; o += kThreads) {
      const int a = o / 27, l = o - a * 27;
      float sum = 0.f;
#pragma unroll
      for (int w = 0; w < kWarps; ++w) sum += part[a][w][l];
      tot[a][l] = sum;
    }
    __syncthreads();
    auto Hjj = [&](int a, int r, int c) -> float {       // symmetric access into the packed triangle
      const int hi = r > c ? r : c, lo = r > c ? c : r;
      return tot[a][hi * (hi + 1) / 2 + lo];
    };
    auto Mm = [&](int a, int r, int c) -> float {        // M = [[Rt, 0], [Nm, Rt]]
      const EdgeXf& X = sx[a];
      if (r < 3) return c < 3 ? X.Rt[r * 3 + c] : 0.f;
      return c < 3 ? X.Nm[(r - 3) * 3 + c] : X.Rt[(r - 3) * 3 + (c - 3)];
    };
    for (int o = tid; o < cnt * 36; o += kThreads) {
      const int a = o / 36, rc = o - a * 36, r = rc / 6, c = rc - r * 6;
      float val = 0.f;
#pragma unroll
      for (int kk = 0; kk < 6; ++kk) val += Mm(a, r, kk) * Hjj(a, kk, c);
      Tm[a][rc] = val;
    }
    __syncthreads();
    // 156 outputs per edge: [0,36) Hii = T M^T, [36,72) Hij = -T, [72,108) Hji = -T^T, [108,144) Hjj,
    //                       [144,150) vi = -M vj, [150,156) vj
    for (int o = tid; o < cnt * 156; o += kThreads) {
      const int a = o / 156, q = o - a * 156;
      const int pi = ix - t0, pj = sx[a].jx - t0;
      const bool iok = pi >= 0 && pi < P, jok = pj >= 0 && pj < P;
      float val = 0.f;
      if (q < 144) {
        const int blk = q / 36, rc = q - blk * 36, r = rc / 6, c = rc - r * 6;
        if (blk == 3) val = Hjj(a, r, c);
        else if (blk == 1) val = -Tm[a][r * 6 + c];
        else if (blk == 2) val = -Tm[a][c * 6 + r];
        else {
#pragma unroll
          for (int ll = 0; ll < 6; ++ll) val += Tm[a][r * 6 + ll] * Mm(a, c, ll);
        }
        const bool rok = (blk < 2) ? iok : jok, cok = (blk == 0 || blk == 2) ? iok : jok;
        if (rok && cok) {
          const int row = 6 * ((blk < 2) ? pi : pj) + r;
          const int col = 6 * ((blk == 0 || blk == 2) ? pi : pj) + c;
          atomicAdd(ws.H + (size_t)row * n6 + col, (double)val);
        }
      } else {
        const int r = (q - 144) % 6;
        if (q < 150) {
#pragma unroll
          for (int kk = 0; kk < 6; ++kk) val -= Mm(a, r, kk) * tot[a][21 + kk];
          if (iok) atomicAdd(ws.v + 6 * pi + r, (double)val);
        } else {
          if (jok) atomicAdd(ws.v + 6 * pj + r, (double)tot[a][21 + r]);
        }
      }
    }
  }

  if (!MOTION_ONLY) {
    const int er = eta_rows == 1 ? 0 : min(k, eta_rows - 1);
#pragma unroll
    for (int s = 0; s < PXT; ++s) {
      if (!ok[s]) continue;
      const int p = px[s];
      const float sd = __ldg(sens + (size_t)ix * HW + p);
      const float m = sd > 0.f ? 1.0f : 0.0f;
      const float C = accC[s] + m * alpha + (1.0f - m) * __ldg(eta + (size_t)er * HW + p);
      const float wk = accW[s] - m * alpha * (d0[s] - sd);
      ws.Q[(size_t)k * HW + p] = 1.0f / C;
      ws.w[(size_t)k * HW + p] = wk;
      if (has_self) {
#pragma unroll
        for (int n = 0; n < 6; ++n) ws.Ei[((size_t)self * 6 + n) * HW + p] = accE[s][n];
      }
    }
    if (eta_rows != 1 && eta_rows != K && blockIdx.x == 0 && tid == 0) ws.info[3] = 1;
  }
}
