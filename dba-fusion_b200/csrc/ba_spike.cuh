// Domain-decomposed block-banded Cholesky solve, included by ba.cu after ba_banded_reg.cuh.
//
// The twisted factorisation runs two sweeps; a chain of P block columns still costs P/2 dependent steps on two
// SMs.  Here the chain is cut into D domains by D-1 separators of B blocks each (B = block half-bandwidth):
//
//     [ I_0 | S_1 | I_1 | S_2 | ... | S_{D-1} | I_{D-1} ]
//
// Interiors are not coupled to each other, so D sweeps (the same kernels as before, one CTA each) eliminate them
// at the same time: I_0 .. I_{D-2} top down, running into the separator on their right, I_{D-1} bottom up, running
// into S_{D-1}.  A sweep's continuation into the separator it meets gives that separator's update for free (the
// "dump" of the twisted scheme).  A middle domain I_d also touches the separator S_d on its LEFT through its first
// B block rows; eliminating I_d couples S_d to itself and to S_{d+1} through the whole domain -- the "spike"
//     W_d = G_d^-1 A(I_d, S_d)              (G_d the Cholesky factor of the interior, n_d x B blocks, dense)
// which a second kernel computes by a blocked forward substitution with 6B right-hand sides from the band store of
// the sweep; continuing that substitution through the rows of S_{d+1} yields the coupling block
//     R(S_{d+1}, S_d) = - L(S_{d+1}, I_d) W_d,
// a third kernel forms  W_d^T W_d  and  W_d^T y_d.  What is left is a block-tridiagonal SPD system over the
// separators (6B x 6B dense blocks), solved by block cyclic reduction: at level l every separator s with
// s mod 2^(l+1) == 2^l is eliminated at the same time (its neighbours s -+ 2^l survive): every CTA of a group
// factorises the diagonal block redundantly in shared memory (no communication) and solves its slice of rows of
//     Xa = A(a,s) L_s^-T,   Xb = A(b,s) L_s^-T;
// after one kernel boundary the updates  Xa Xa^T, Xb Xb^T  of the neighbours' diagonal blocks and their new coupling
// -Xb Xa^T are formed.  log2(D) levels instead of D/2, so the chain can be cut into 16 domains at the depth 10 had
// with an elimination from both ends.  The last separator is solved densely, the levels are back-substituted in
// reverse, the interiors' right-hand sides are corrected by  y_d -= W_d x(S_d), and D back-substitutions
// (k_spike_back, the streaming two-warp routine of ba_banded.cuh) finish the solve.  Same arithmetic as a sequential
// Cholesky of a reordered matrix; replaces Eigen::SimplicialLLT of the reference (src/droid_kernels.cu:1248-1269).
#pragma once

constexpr int kSpikeMaxD = kBandMaxParts;
constexpr int kSpikeThreads = 512;   // 4 warps per scheduler: the dense kernels are chains of short shared-memory round trips

constexpr int kSpikeMaxLev = 5;   // cyclic-reduction levels (16 domains: 15 separators, 3 elimination levels + the last block)
struct SpikePlan {
  int D, B, mid, G;               // domains, band, LAST separator of the cyclic reduction, CTAs per group in the reduced levels
  int nlev;                       // elimination levels of the cyclic reduction over the separators
  int m[kSpikeMaxD];              // first physical block of separator s, s = 1 .. D-1
  long long W[kSpikeMaxD];        // middle domain d: spike [n_d][6][NS] doubles
  long long Roff[kSpikeMaxD];     // middle domain d: coupling block rows S_{d+1}, cols S_d, [NS][NS] (level-0 coupling of s = d)
  long long SLL[kSpikeMaxD];      // middle domain d: W^T W [NS][NS], then W^T y [NS]
  long long Lf[kSpikeMaxD];       // separator s: factor [NS][ld], then y [NS], then 1/diag [NS]
  long long X[kSpikeMaxD][2];     // eliminated separator s: Xa = A(s-h,s) L_s^-T and Xb = A(s+h,s) L_s^-T, [NS][NS] each
  long long U[kSpikeMaxD][2];     // separator s: updates received from its left / right neighbours (accumulated over the
                                  // levels; cleared by a memset per solve), [NS][NS] then rhs part [NS]
  long long Cl[kSpikeMaxLev][kSpikeMaxD];   // level l >= 1, separator a: coupling A(a + 2^l, a), [NS][NS] (rows: a + 2^l)
  long long Ubase, Ubytes;        // the U blocks as one span (offset in doubles, size in bytes)
  long long xsep;                 // [D][NS]: separator solutions, entry s
};
// flags (int32, at plan.flags): [0, D) sweeps, [D] reduced system, [D + 1] overall, [D + 2, 2 D + 2) progress of the sweeps

__host__ __device__ inline int spike_ld(int NS) { return NS | 1; }

// --------------------------------------------------------------------------------------------- spike ----
// grid (column groups, middle domains), 128 threads.  lane = right-hand-side column, warp w takes the panel
// blocks L(I, I-o) with o % 4 == w.  Block rows of L come from the sweep's band store through a 4-deep cp.async
// ring; the last B rows of W live in a shared-memory window.
constexpr int kSpikeRing = 5;     // block rows in flight: three ahead, one in use, one still read by the row solver

__global__ void __launch_bounds__(128)
k_spike_fwd(int Pphys, int B, BandPlan plan, SpikePlan sp, Ws ws) {
  extern __shared__ double sms[];
  const int d = 1 + blockIdx.y;
  const BandPart pt = plan.part[d];
  const int tid = threadIdx.x, lane = tid & 31, wq = tid >> 5;
  const int NS = 6 * B, n = 6 * Pphys;
  const int ncg = gridDim.x;
  const int NC = (NS + ncg - 1) / ncg;              // columns of this CTA: [c0, c0 + nc)
  const int c0 = blockIdx.x * NC;
  const int nc = min(NC, NS - c0);
  const bool live = lane < nc;
  const int nd = pt.n_elim, rows = pt.Plog;         // interior rows, then the B rows of the separator on the right
  const int RD = (B + 1) * 36, SD = RD + 6;         // ring stage: block row of L | 1/diag
  const int WB = B + 1;                             // window slots
  double* ring = sms;                               // [kSpikeRing][SD]
  double* win = ring + kSpikeRing * SD;             // [WB][6][32]
  double* part = win + WB * 192;                    // [4][6][32]
  const double* Lband = ws.Lfac + pt.store;
  const double* dinv_g = Lband + (size_t)pt.Plog * RD;
  double* Wg = ws.Lfac + sp.W[d];
  double* Ro = ws.Lfac + sp.Roff[d];
  const int mL = sp.m[d];                           // first physical block of the separator on the left

  // The sweep of this domain may still be running (this kernel is launched with programmatic stream serialisation):
  // block row I of its band store is complete once its progress word has passed min(I, nd - 1).
  const volatile int* progress = reinterpret_cast<const volatile int*>(ws.Lfac + plan.flags) + sp.D + 2 + d;
  auto request = [&](int I) {
    if (I < rows) {
      const int need = min(I, nd - 1) + 1;
      while (*progress < need) __nanosleep(100);
      __threadfence();
      const unsigned dst = (unsigned)__cvta_generic_to_shared(ring + (I % kSpikeRing) * SD);
      const double* src = Lband + (size_t)I * RD;
      for (int q = tid; q < RD / 2; q += 128)
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + q * 16), "l"(src + 2 * q) : "memory");
      if (tid < 3 && I < nd)
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + (RD + 2 * tid) * 8), "l"(dinv_g + 6 * I + 2 * tid) : "memory");
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
  for (int I = 0; I < 3; ++I) request(I);
  for (int I = 0; I < rows; ++I) {
    request(I + 3);
    asm volatile("cp.async.wait_group 3;" ::: "memory");          // everything but the 3 newest groups: row I is in
    // right-hand side of this block row: A(I, S_left), non-zero only within the band (I < B)
    double a[6] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
    // (block row I of the domain lies B + I - cb blocks below column block cb of the separator: only the blocks
    //  within the band exist -- what lies beyond it is never written, nor cleared by dbaf_ba_zero_system)
    if (wq == 0 && live && I < B && I < nd && B + I - (c0 + lane) / 6 <= B) {
      const int pr = pt.rev ? pt.base - I : pt.base + I;
      const double* src = ws.H + (size_t)(6 * pr) * n + 6 * mL + c0 + lane;
#pragma unroll
      for (int r = 0; r < 6; ++r) a[r] = src[(size_t)r * n];
    }
    __syncthreads();                                  // ring stage visible; previous row's window entry visible
    const double* st = ring + (I % kSpikeRing) * SD;
    double acc[6] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
    // separator rows only see the interior: o >= I - nd + 1
    const int o_lo = max(1, I - nd + 1), o_hi = min(B, I);
    for (int o = o_lo + ((wq - o_lo) & 3); o <= o_hi; o += 4) {
      const double* Lt = st + o * 36;                 // transposed panel block: Lt[c * 6 + r] = L(I, I-o)[r][c]
      const double* wv = win + ((I - o) % WB) * 192 + lane;
#pragma unroll
      for (int c = 0; c < 6; ++c) {
        const double x = wv[c * 32];
        const double2 l0 = *reinterpret_cast<const double2*>(Lt + c * 6);
        const double2 l1 = *reinterpret_cast<const double2*>(Lt + c * 6 + 2);
        const double2 l2 = *reinterpret_cast<const double2*>(Lt + c * 6 + 4);
        acc[0] = fma(l0.x, x, acc[0]);
        acc[1] = fma(l0.y, x, acc[1]);
        acc[2] = fma(l1.x, x, acc[2]);
        acc[3] = fma(l1.y, x, acc[3]);
        acc[4] = fma(l2.x, x, acc[4]);
        acc[5] = fma(l2.y, x, acc[5]);
      }
    }
#pragma unroll
    for (int r = 0; r < 6; ++r) part[(wq * 6 + r) * 32 + lane] = acc[r];
    __syncthreads();
    if (wq == 0 && live) {
      double x[6];
#pragma unroll
      for (int r = 0; r < 6; ++r)
        x[r] = a[r] - (part[r * 32 + lane] + part[(6 + r) * 32 + lane] + part[(12 + r) * 32 + lane] + part[(18 + r) * 32 + lane]);
      if (I < nd) {
        const double* Ld = st;                        // diagonal factor, row-major lower
        const double* di = st + RD;
#pragma unroll
        for (int c = 0; c < 6; ++c) {
          x[c] *= di[c];
#pragma unroll
          for (int r = 0; r < 6; ++r)
            if (r > c) x[r] = fma(-Ld[r * 6 + c], x[c], x[r]);
        }
        double* wo = win + (I % WB) * 192 + lane;
#pragma unroll
        for (int r = 0; r < 6; ++r) {
          wo[r * 32] = x[r];
          Wg[((size_t)I * 6 + r) * NS + c0 + lane] = x[r];
        }
      } else {
#pragma unroll
        for (int r = 0; r < 6; ++r) Ro[((size_t)(I - nd) * 6 + r) * NS + c0 + lane] = x[r];
      }
    }
  }
  asm volatile("cp.async.wait_group 0;" ::: "memory");
}

// W^T W (lower tiles and their mirror images) and W^T y of a middle domain.  grid (10 tiles, middle domains)
constexpr int kGramTile = 32;
__global__ void __launch_bounds__(256)
k_spike_gram(int B, BandPlan plan, SpikePlan sp, Ws ws) {
  __shared__ double Wi[32][kGramTile + 1], Wj[32][kGramTile + 1], ys[32];
  const int d = 1 + blockIdx.y;
  const BandPart pt = plan.part[d];
  const int NS = 6 * B;
  const int nt = (NS + kGramTile - 1) / kGramTile;
  int ti = 0, rest = blockIdx.x;                     // lower-triangular tile (ti >= tj)
  while (rest > ti) { rest -= ti + 1; ++ti; }
  const int tj = rest;
  if (ti >= nt) return;
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int K = 6 * pt.n_elim;
  const double* Wg = ws.Lfac + sp.W[d];
  const double* y_g = ws.Lfac + pt.store + (size_t)pt.Plog * (B + 1) * 36 + 6 * pt.Plog;
  double* S = ws.Lfac + sp.SLL[d];
  double acc[2][2] = {{0.0, 0.0}, {0.0, 0.0}};
  double gy[2] = {0.0, 0.0};
  const int i0 = ti * kGramTile, j0 = tj * kGramTile;
  for (int k0 = 0; k0 < K; k0 += 32) {
    __syncthreads();
    for (int e = tid; e < 32 * kGramTile; e += 256) {
      const int kk = e / kGramTile, c = e - kk * kGramTile;
      const bool okk = k0 + kk < K;
      Wi[kk][c] = (okk && i0 + c < NS) ? Wg[(size_t)(k0 + kk) * NS + i0 + c] : 0.0;
      Wj[kk][c] = (okk && j0 + c < NS) ? Wg[(size_t)(k0 + kk) * NS + j0 + c] : 0.0;
    }
    if (tid < 32) ys[tid] = k0 + tid < K ? y_g[k0 + tid] : 0.0;
    __syncthreads();
    const int kn = min(32, K - k0);
    for (int kk = 0; kk < kn; ++kk) {
      const double a0 = Wi[kk][ty], a1 = Wi[kk][ty + 16], b0 = Wj[kk][tx], b1 = Wj[kk][tx + 16];
      acc[0][0] = fma(a0, b0, acc[0][0]);
      acc[0][1] = fma(a0, b1, acc[0][1]);
      acc[1][0] = fma(a1, b0, acc[1][0]);
      acc[1][1] = fma(a1, b1, acc[1][1]);
      if (tj == 0 && tx == 0) {
        const double yv = ys[kk];
        gy[0] = fma(a0, yv, gy[0]);
        gy[1] = fma(a1, yv, gy[1]);
      }
    }
  }
#pragma unroll
  for (int u = 0; u < 2; ++u)
#pragma unroll
    for (int v = 0; v < 2; ++v) {
      const int i = i0 + ty + 16 * u, j = j0 + tx + 16 * v;
      if (i < NS && j < NS) {
        S[(size_t)i * NS + j] = acc[u][v];
        if (ti != tj) S[(size_t)j * NS + i] = acc[u][v];
      }
    }
  if (tj == 0 && tx == 0) {
#pragma unroll
    for (int u = 0; u < 2; ++u)
      if (i0 + ty + 16 * u < NS) S[(size_t)NS * NS + i0 + ty + 16 * u] = gy[u];
  }
}

// --------------------------------------------------------------------------- reduced (separator) system ----
// Entry (i, j), j <= i, of the current diagonal block of separator s, and entry i of its right-hand side:
// what the sweeps left, minus the spike term of the domain on its right, minus the updates of the chain so far.
struct SepAsm {
  const double* dumpA;      // top-down sweep of domain s-1
  const double* dumpR;      // bottom-up sweep (s == D-1 only), else nullptr
  const double* sll;        // W^T W of domain s (middle domains), else nullptr
  const double* u0;         // update from side 0, else nullptr
  const double* u1;         // update from side 1, else nullptr
  int m;                    // first physical block of the separator
};
__device__ __forceinline__ SepAsm sep_asm(int s, const BandPlan& plan, const SpikePlan& sp, const Ws& ws) {
  SepAsm a;
  a.dumpA = ws.Lfac + plan.part[s - 1].dump;
  a.dumpR = s == sp.D - 1 ? ws.Lfac + plan.part[sp.D - 1].dump : nullptr;
  a.sll = s <= sp.D - 2 ? ws.Lfac + sp.SLL[s] : nullptr;
  a.u0 = ws.Lfac + sp.U[s][0];                         // (zero until a neighbour has been eliminated)
  a.u1 = ws.Lfac + sp.U[s][1];
  a.m = sp.m[s];
  return a;
}
// L (lower triangle, leading dimension ld) and the right-hand side in row NS, whole CTA; one simple pass per term
__device__ void sep_assemble(const SepAsm& a, double* L, int ld, int B, int n, float lm, float ep, const Ws& ws) {
  const int tid = threadIdx.x, nt = blockDim.x;
  const int NS = 6 * B;
  // blocks (bi >= bj) of the top-down sweep's view of the separator, 36 consecutive doubles each.  One CTA pulls
  // ~115 KB per term through a ~1 us latency: what counts is the number of bytes in flight, so every thread issues
  // eight 16-byte loads (pairs of one block row) before it touches the first value
  const int n_pairs = B * B * 18;
  for (int q0 = tid; q0 < n_pairs; q0 += 8 * nt) {
    double2 v[8];
    int dst[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const int q = q0 + u * nt;
      dst[u] = -1;
      v[u] = make_double2(0.0, 0.0);
      if (q >= n_pairs) continue;
      const int e = 2 * q;
      const int blk = e / 36, rc = e - blk * 36, bi = blk / B, bj = blk - bi * B;
      const int r = rc / 6, c = rc - 6 * r;          // c even: (c, c + 1) lie in one block row
      if (bi < bj || (bi == bj && c > r)) continue;  // (strictly above the diagonal; c == r keeps the first of the pair)
      dst[u] = (6 * bi + r) * ld + 6 * bj + c;
      v[u] = *reinterpret_cast<const double2*>(a.dumpA + e);
      if (a.dumpR) {
        // the bottom-up sweep works on the block-reversed ordering: physical block (bi, bj), bi >= bj, is the transpose
        // of its logical block (B-1-bj, B-1-bi); both sweeps carry the damped A_ss, counted once
        const double* dr = a.dumpR + (size_t)((B - 1 - bj) * B + (B - 1 - bi)) * 36;
        const double* hr = ws.H + (size_t)(6 * (a.m + bi) + r) * n + 6 * (a.m + bj) + c;
        double h0 = hr[0], h1 = hr[1];
        if (bi == bj && r == c) h0 += (double)ep + (double)lm * h0;
        if (bi == bj && r == c + 1) h1 += (double)ep + (double)lm * h1;
        v[u].x += dr[bi == bj ? r * 6 + c : c * 6 + r] - h0;
        v[u].y += dr[bi == bj ? r * 6 + c + 1 : (c + 1) * 6 + r] - h1;
      }
    }
#pragma unroll
    for (int u = 0; u < 8; ++u)
      if (dst[u] >= 0) {
        L[dst[u]] = v[u].x;
        // second element of the pair: inside the lower triangle unless the pair straddles the diagonal of a diagonal block
        const int e = 2 * (q0 + u * nt);
        const int blk = e / 36, rc = e - blk * 36, bi = blk / B, bj = blk - bi * B;
        const int r = rc / 6, c = rc - 6 * r;
        if (!(bi == bj && c + 1 > r)) L[dst[u] + 1] = v[u].y;
      }
  }
  for (int i = tid; i < NS; i += nt) {
    double v = a.dumpA[(size_t)B * B * 36 + i];
    if (a.dumpR) {
      const int bi = i / 6, r = i - 6 * bi;
      v += a.dumpR[(size_t)B * B * 36 + (B - 1 - bi) * 6 + r] - ws.v[6 * a.m + i];
    }
    L[(size_t)NS * ld + i] = v;
  }
  __syncthreads();
  // the spike term and the chain updates: one warp per row (row NS = right-hand-side part behind each matrix), all
  // loads of a pass issued before the first use
  for (int i = tid >> 5; i <= NS; i += nt >> 5) {
    const int jn = i < NS ? i + 1 : NS;
    for (int j0 = tid & 31; j0 < jn; j0 += 128) {
      double t[4] = {0.0, 0.0, 0.0, 0.0};
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int j = j0 + 32 * u;
        if (j < jn) {
          const size_t o = (size_t)i * NS + j;
          double x = 0.0;
          if (a.sll) x += a.sll[o];
          if (a.u0) x += a.u0[o];
          if (a.u1) x += a.u1[o];
          t[u] = x;
        }
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int j = j0 + 32 * u;
        if (j < jn) L[(size_t)i * ld + j] -= t[u];
      }
    }
  }
  __syncthreads();
}

// cyclic reduction, level l (h = 2^l): the e-th separator eliminated at this level and its two neighbours
__device__ __forceinline__ void cr_step(const SpikePlan& sp, int level, int e, int& s, int& a, int& b) {
  const int h = 1 << level;
  s = h + e * 2 * h;
  a = s - h >= 1 ? s - h : 0;                 // 0: no such neighbour
  b = s + h <= sp.D - 1 ? s + h : 0;
}
// coupling block A(x + 2^l, x) at level l: rows = unknowns of x + 2^l, columns = unknowns of x
__device__ __forceinline__ const double* cr_coupling(const SpikePlan& sp, const Ws& ws, int level, int x) {
  return ws.Lfac + (level == 0 ? sp.Roff[x] : sp.Cl[level][x]);
}

// level, part a: factor the diagonal block of an eliminated separator s (every CTA of its two groups, redundantly),
// then this CTA's rows of X = A(t,s) L_s^-T for the group's neighbour t and of the right-hand-side update X y_s.
// grid (G, 2 * eliminated separators): y = 2 e + side, side 0: t = s - h, side 1: t = s + h
__global__ void __launch_bounds__(kSpikeThreads)
k_cr_a(int Pphys, int B, float lm, float ep, int level, BandPlan plan, SpikePlan sp, Ws ws) {
  extern __shared__ double smr[];
  __shared__ int fail;
  const int side = blockIdx.y & 1;
  int s, na, nb;
  cr_step(sp, level, blockIdx.y >> 1, s, na, nb);
  const int t = side == 0 ? na : nb;
  if (t == 0) return;                                // this neighbour does not exist
  const bool keeper = blockIdx.x == 0 && (side == 0 || na == 0);   // the one CTA that stores the factor
  const int tid = threadIdx.x, nt = blockDim.x, lane = tid & 31, wid = tid >> 5;
  const int NS = 6 * B, n = 6 * Pphys, ld = spike_ld(NS);
  double* L = smr;                                   // (NS + 1) x ld
  double* dinv = L + (size_t)(NS + 1) * ld;          // NS
  double* xs = dinv + NS;                            // per warp: the row of X being built and the row of A, [2][ld]
  const SepAsm as = sep_asm(s, plan, sp, ws);
  sep_assemble(as, L, ld, B, n, lm, ep, ws);
  const bool ok = cholesky6_factor(L, ld, B, dinv, &fail);
  __syncthreads();
  int* flags = reinterpret_cast<int*>(ws.Lfac + plan.flags);
  if (!ok) {
    if (tid == 0 && blockIdx.x == 0) atomicExch(&flags[sp.D], 1);
    return;
  }
  double* Lf = ws.Lfac + sp.Lf[s];
  if (keeper) {                                       // keep the factor, y and 1/diag for the back-substitution
    for (int e = tid; e < (NS + 1) * ld; e += nt) Lf[e] = L[e];   // (the unused upper triangle travels along)
    for (int i = tid; i < NS; i += nt) Lf[(size_t)(NS + 1) * ld + i] = dinv[i];
  }
  // rows [i_lo, i_hi) of X: x L^T = a, a = row of A(t,s).  One warp per row.  The row of A is brought into shared
  // memory up front (one exposed latency instead of one per 6-wide block); for block J the 24 lanes (part, column)
  // split the dot products  sum_k x[k] L[c][k]  four ways, lanes 0..5 finish the 6x6 triangular solve through
  // shuffles.
  const int rps = (NS + gridDim.x - 1) / gridDim.x;
  const int i_lo = blockIdx.x * rps, i_hi = min(NS, i_lo + rps);
  // A(t,s)[i][c]: t = s + h reads the coupling of s as stored (rows: t), t = s - h the transpose of the coupling of t
  const double* Cts = cr_coupling(sp, ws, level, side == 1 ? s : t);
  const int nw = nt >> 5;
  const int part = lane / 6, cw = lane - 6 * part;   // part 0..3 take part in the dots (lanes 0..23)
  double* Xg = ws.Lfac + sp.X[s][side];
  double* Ug = ws.Lfac + sp.U[t][side == 0 ? 1 : 0];  // t = s - h receives from its right, t = s + h from its left
  for (int i = i_lo + wid; i < i_hi; i += nw) {
    double* xr = xs + (size_t)wid * 2 * ld;            // x (being built), then the row of A behind it
    double* ar = xr + ld;
    for (int c = lane; c < NS; c += 32) ar[c] = side == 1 ? Cts[(size_t)i * NS + c] : Cts[(size_t)c * NS + i];
    __syncwarp();
    for (int J = 0; J < B; ++J) {
      const int c = 6 * J + cw;
      double sum = 0.0;
      if (part < 4) {
        const double* lc = L + (size_t)c * ld;
        double s0 = 0.0, s1 = 0.0;
        int k = part;
        for (; k + 4 < 6 * J; k += 8) {
          s0 = fma(xr[k], lc[k], s0);
          s1 = fma(xr[k + 4], lc[k + 4], s1);
        }
        if (k < 6 * J) s0 = fma(xr[k], lc[k], s0);
        sum = s0 + s1;
      }
      // lanes 0..5 collect the four parts (sources 6, 12, 18 lanes up)
      sum += __shfl_down_sync(0xffffffffu, sum, 12);
      sum += __shfl_down_sync(0xffffffffu, sum, 6);
      const bool act = lane < 6;
      sum = act ? ar[c] - sum : 0.0;
      // finish the 6x6 triangular solve of the block on lanes 0..5: column sweep through shuffles
#pragma unroll
      for (int c2 = 0; c2 < 6; ++c2) {
        double xc = sum * (act ? dinv[6 * J + c2] : 0.0);
        xc = __shfl_sync(0xffffffffu, xc, c2);                       // x[c2] of this row, final
        if (act && cw > c2) sum = fma(-xc, L[(size_t)(6 * J + cw) * ld + 6 * J + c2], sum);
        if (act && cw == c2) sum = xc;
      }
      if (act) xr[6 * J + cw] = sum;
      __syncwarp();
    }
    // the row of X to global, and its entry of the right-hand-side update X y_s (accumulated: one writer per level)
    double dot = 0.0;
    for (int k = lane; k < NS; k += 32) {
      const double x = xr[k];
      Xg[(size_t)i * NS + k] = x;
      dot = fma(x, L[(size_t)NS * ld + k], dot);
    }
    dot = warp_sum_d(dot);
    if (lane == 0) Ug[(size_t)NS * NS + i] += dot;
    __syncwarp();
  }
}

// level, part b: this CTA's rows of one of the three products of an eliminated separator s:
//   which 0:  U(a)[right] += Xa Xa^T (lower triangle)      which 1:  U(b)[left] += Xb Xb^T (lower triangle)
//   which 2:  coupling A(b,a) of the next level = -Xb Xa^T (all of it)
// grid (G, 3 * eliminated separators): y = 3 e + which
__global__ void __launch_bounds__(kSpikeThreads)
k_cr_b(int B, int level, BandPlan plan, SpikePlan sp, Ws ws) {
  extern __shared__ double smx[];
  const int which = blockIdx.y % 3;
  int s, na, nb;
  cr_step(sp, level, blockIdx.y / 3, s, na, nb);
  if ((which == 0 && na == 0) || (which == 1 && nb == 0) || (which == 2 && (na == 0 || nb == 0))) return;
  const int tid = threadIdx.x, nt = blockDim.x;
  const int NS = 6 * B, ld = spike_ld(NS);
  const int* flags = reinterpret_cast<const int*>(ws.Lfac + plan.flags);
  if (flags[sp.D] != 0) return;
  const double* Qg = ws.Lfac + sp.X[s][which == 1 ? 1 : 0];    // right factor (transposed in the product)
  const double* Pg = ws.Lfac + sp.X[s][which == 0 ? 0 : 1];    // left factor: rows of the result
  // Q transposed: smx[k][j], so that the lanes of a warp (consecutive columns j) read consecutive words
  for (int e = tid; e < NS * NS; e += nt) {
    const int i = e / NS, k = e - i * NS;
    smx[(size_t)k * ld + i] = Qg[e];
  }
  const int rps = (NS + gridDim.x - 1) / gridDim.x;
  const int i_lo = blockIdx.x * rps, i_hi = min(NS, i_lo + rps);
  // this CTA's rows of P, transposed as well: prow[k][r], r = local row
  double* prow = smx + (size_t)NS * ld;
  const int pld = rps | 1;
  for (int e = tid; e < (i_hi - i_lo) * NS; e += nt) {
    const int r = e / NS, k = e - r * NS;
    prow[(size_t)k * pld + r] = Pg[(size_t)(i_lo + r) * NS + k];
  }
  __syncthreads();
  const bool full = which == 2;
  double* Og = which == 0 ? ws.Lfac + sp.U[na][1] : which == 1 ? ws.Lfac + sp.U[nb][0] : ws.Lfac + sp.Cl[level + 1][na];
  // a warp task = 3 rows x 32 columns: per k one wavefront for the columns, three broadcasts for the rows, 96 FMAs
  const int lane = tid & 31, wid = tid >> 5, nw = nt >> 5;
  const int n_rt = (i_hi - i_lo + 2) / 3, n_ct = (NS + 31) / 32;
  for (int task = wid; task < n_rt * n_ct; task += nw) {
    const int rt = task / n_ct, ct = task - rt * n_ct;
    const int i0 = i_lo + 3 * rt, j = ct * 32 + lane;
    if (!full && ct * 32 > min(i0 + 2, i_hi - 1)) continue;      // the whole tile lies above the diagonal
    const int ra = i0 - i_lo, rb = min(i0 + 1, i_hi - 1) - i_lo, rc = min(i0 + 2, i_hi - 1) - i_lo;
    const int jj = min(j, NS - 1);
    double s0 = 0.0, s1 = 0.0, s2 = 0.0;
#pragma unroll 4
    for (int k = 0; k < NS; ++k) {
      const double xj = smx[(size_t)k * ld + jj];
      const double* pr = prow + (size_t)k * pld;
      s0 = fma(pr[ra], xj, s0);
      s1 = fma(pr[rb], xj, s1);
      s2 = fma(pr[rc], xj, s2);
    }
    if (j < NS) {
      const int ia = i0, ib = i0 + 1, ic = i0 + 2;
      if (full) {
        Og[(size_t)ia * NS + j] = -s0;
        if (ib < i_hi) Og[(size_t)ib * NS + j] = -s1;
        if (ic < i_hi) Og[(size_t)ic * NS + j] = -s2;
      } else {
        if (j <= ia) Og[(size_t)ia * NS + j] += s0;
        if (ib < i_hi && j <= ib) Og[(size_t)ib * NS + j] += s1;
        if (ic < i_hi && j <= ic) Og[(size_t)ic * NS + j] += s2;
      }
    }
  }
}

// the last separator of the cyclic reduction (sp.mid): dense solve; decides whether the whole step is kept.  one CTA
__global__ void __launch_bounds__(kSpikeThreads)
k_red_mid(int Pphys, int B, float lm, float ep, BandPlan plan, SpikePlan sp, Ws ws) {
  extern __shared__ double smm[];
  __shared__ int fail;
  const int tid = threadIdx.x, nt = blockDim.x;
  const int NS = 6 * B, n = 6 * Pphys, ld = spike_ld(NS), s = sp.mid;
  double* L = smm;
  double* x = L + (size_t)(NS + 1) * ld;
  double* dinv = x + NS;
  const SepAsm as = sep_asm(s, plan, sp, ws);
  sep_assemble(as, L, ld, B, n, lm, ep, ws);
  bool ok = cholesky6_solve(L, ld, B, x, dinv, &fail);
  __syncthreads();
  int* flags = reinterpret_cast<int*>(ws.Lfac + plan.flags);
  for (int q = 0; q <= sp.D; ++q) ok = ok && flags[q] == 0;
  double* xsep = ws.Lfac + sp.xsep + (size_t)s * NS;
  for (int i = tid; i < NS; i += nt) {
    xsep[i] = ok ? x[i] : 0.0;
    ws.dx[6 * sp.m[s] + i] = ok ? (float)x[i] : 0.0f;
  }
  if (tid == 0) {
    flags[sp.D + 1] = ok ? 0 : 1;
    if (!ok) atomicAdd(&ws.info[1], 1);
  }
}

// back-substitution of one level of the cyclic reduction: x_s = L_s^-T (y_s - Xa^T x_a - Xb^T x_b) for every
// separator eliminated at that level (its neighbours are solved already).  grid: eliminated separators, one CTA each
__global__ void __launch_bounds__(kSpikeThreads)
k_cr_back(int B, int level, BandPlan plan, SpikePlan sp, Ws ws) {
  extern __shared__ double smk2[];
  const int tid = threadIdx.x, nt = blockDim.x;
  const int NS = 6 * B, ld = spike_ld(NS);
  const int* flags = reinterpret_cast<const int*>(ws.Lfac + plan.flags);
  const bool ok = flags[sp.D + 1] == 0;
  int s, na, nb;
  cr_step(sp, level, blockIdx.x, s, na, nb);
  double* L = smk2;                                  // NS x ld
  double* x = L + (size_t)NS * ld;
  double* dinv = x + NS;
  double* xa = dinv + NS;                            // solutions of the two neighbours
  double* xb = xa + NS;
  double* xs_g = ws.Lfac + sp.xsep + (size_t)s * NS;
  if (!ok) {
    for (int i = tid; i < NS; i += nt) {
      xs_g[i] = 0.0;
      ws.dx[6 * sp.m[s] + i] = 0.0f;
    }
    return;
  }
  const double* Lf = ws.Lfac + sp.Lf[s];
  // 117 KB through one CTA: eight 16-byte loads in flight per thread (the offsets of Lf are even, ld is odd: the
  // factor is copied as a flat array)
  {
    const int n2 = (NS * ld) >> 1;
    const double2* src = reinterpret_cast<const double2*>(Lf);
    for (int q0 = tid; q0 < n2; q0 += 8 * nt) {
      double2 v[8];
#pragma unroll
      for (int u = 0; u < 8; ++u)
        if (q0 + u * nt < n2) v[u] = src[q0 + u * nt];
#pragma unroll
      for (int u = 0; u < 8; ++u)
        if (q0 + u * nt < n2) { L[2 * (q0 + u * nt)] = v[u].x; L[2 * (q0 + u * nt) + 1] = v[u].y; }
    }
    if (((NS * ld) & 1) && tid == 0) L[NS * ld - 1] = Lf[NS * ld - 1];
  }
  for (int i = tid; i < NS; i += nt) {
    dinv[i] = Lf[(size_t)(NS + 1) * ld + i];
    x[i] = Lf[(size_t)NS * ld + i];
    xa[i] = na ? (ws.Lfac + sp.xsep + (size_t)na * NS)[i] : 0.0;
    xb[i] = nb ? (ws.Lfac + sp.xsep + (size_t)nb * NS)[i] : 0.0;
  }
  __syncthreads();
  {
    // z = y - Xa^T x_a - Xb^T x_b: thread (k, part) sums a slice of the rows, partial sums meet in shared memory
    const double* Xa = ws.Lfac + sp.X[s][0];
    const double* Xb = ws.Lfac + sp.X[s][1];
    const int parts = max(1, nt / NS), k = tid % NS, part = tid / NS;
    double sum = 0.0;
    if (part < parts) {
      const int per = (NS + parts - 1) / parts;
      const int a0 = part * per, a1 = min(NS, a0 + per);
      if (na) {
#pragma unroll 8
        for (int i = a0; i < a1; ++i) sum = fma(Xa[(size_t)i * NS + k], xa[i], sum);
      }
      if (nb) {
#pragma unroll 8
        for (int i = a0; i < a1; ++i) sum = fma(Xb[(size_t)i * NS + k], xb[i], sum);
      }
    }
    for (int q = 0; q < parts; ++q) {
      if (part == q) x[k] -= sum;
      __syncthreads();
    }
  }
  cholesky6_back(L, ld, B, x, dinv);
  for (int i = tid; i < NS; i += nt) {
    xs_g[i] = x[i];
    ws.dx[6 * sp.m[s] + i] = (float)x[i];
  }
}

// y_d -= W_d x(S_d) for the middle domains.  grid (row groups, middle domains), one warp per scalar row
__global__ void __launch_bounds__(256)
k_spike_apply(int B, BandPlan plan, SpikePlan sp, Ws ws) {
  const int d = 1 + blockIdx.y;
  const BandPart pt = plan.part[d];
  const int NS = 6 * B;
  const int lane = threadIdx.x & 31;
  const int row = blockIdx.x * 8 + (threadIdx.x >> 5);
  if (row >= 6 * pt.n_elim) return;
  const int* flags = reinterpret_cast<const int*>(ws.Lfac + plan.flags);
  if (flags[sp.D + 1] != 0) return;
  const double* Wg = ws.Lfac + sp.W[d] + (size_t)row * NS;
  const double* xs = ws.Lfac + sp.xsep + (size_t)d * NS;
  double sum = 0.0;
  for (int c = lane; c < NS; c += 32) sum = fma(Wg[c], xs[c], sum);
  sum = warp_sum_d(sum);
  if (lane == 0) {
    double* y_g = ws.Lfac + pt.store + (size_t)pt.Plog * (B + 1) * 36 + 6 * pt.Plog;
    y_g[row] -= sum;
  }
}

// the D interior back-substitutions.  grid D, 64 threads
__global__ void __launch_bounds__(64)
k_spike_back(int Pphys, int B, BandPlan plan, SpikePlan sp, Ws ws) {
  extern __shared__ double smk3[];
  const int d = blockIdx.x;
  const BandPart pt = plan.part[d];
  const int* flags = reinterpret_cast<const int*>(ws.Lfac + plan.flags);
  if (flags[sp.D + 1] != 0) {
    for (int I = threadIdx.x; I < pt.n_elim * 6; I += blockDim.x) {
      const int J = I / 6;
      ws.dx[6 * (pt.rev ? pt.base - J : pt.base + J) + I % 6] = 0.0f;
    }
    return;
  }
  const int s = d == sp.D - 1 ? sp.D - 1 : d + 1;     // the separator this part's sweep ran into
  band_backsub(Pphys, B, pt, pt.sep_first, ws.Lfac + sp.xsep + (size_t)s * 6 * B, smk3, ws);
}
