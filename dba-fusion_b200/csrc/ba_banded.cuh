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
//
// Long chains are factorised from both ends at once ("twisted" factorisation): CTA 0 eliminates blocks
// 0 .. m-1 top down, CTA 1 eliminates blocks P-1 .. m+B bottom up (the same sweep on the block-reversed
// ordering, read straight from the symmetric H).  The two parts are not coupled to each other (they are
// more than B blocks apart), both only touch the B-block separator [m, m+B) in between.  Each CTA leaves
// its updated copy of the separator in global memory; k_band_sep adds them (S = S_top + S_bottom - A_ss),
// solves the dense 6B x 6B system with the shared-memory Cholesky of ba_fast.cuh, and k_band_back runs the
// two back-substitutions outwards in parallel.  Same arithmetic, half the serial pivot chain.
#pragma once

constexpr int kBandThreads = 512;
constexpr int kBandTwistMinP = 64;   // below this the single sweep is just as fast

// One sweep of the (possibly reversed) chain: logical block I is physical block  rev ? P-1-I : I.
constexpr int kBandMaxParts = 16;
struct BandPart {
  int rev;               // 1: bottom-up sweep
  int Plog;              // logical size of this part's leading submatrix (eliminated blocks + separator)
  int n_elim;            // number of block columns this sweep eliminates
  int base;              // physical block of logical block 0: logical I is physical base + I (base - I when rev)
  int sep_first;         // first physical block of the separator this sweep runs into
  long long store;       // offset (doubles) of this part's band store inside ws.Lfac
  long long dump;        // offset (doubles) of this part's view of its separator: B*B*36 window dump + 6B rhs dump
};
struct BandPlan {
  int twisted;           // 0: one CTA, everything (including the back-substitution) in the sweep kernel;
                         // 1: several sweeps that leave their separator views for a second stage (two sweeps meeting
                         //    at one separator: k_band_sep; more: the domain decomposition of ba_spike.cuh)
  int m;                 // two sweeps: first physical block of the separator
  BandPart part[kBandMaxParts];
  long long sep;         // two sweeps: offset (doubles) of the exchange area inside ws.Lfac:
                         // 2 x (B*B*36 window dump + 6B rhs dump), 6B solution, 4 flags
  long long flags;       // offset (doubles) of the int32 failure flags, one per sweep (then stage-specific ones)
};
__host__ __device__ inline long long band_store_doubles(int Plog, int B) { return (long long)Plog * (B + 1) * 36 + 12LL * Plog; }
__host__ __device__ inline long long band_dump_doubles(int B) { return (long long)B * B * 36 + 6LL * B; }
constexpr int kBandMaxB = 25;      // (B+2)^2 blocks * 288 B must fit in shared memory

__device__ __forceinline__ double warp_sum_d(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Back-substitution L^T x = y of one part, bottom up, right-looking, on TWO warps: once x_J is known every
// pending right-hand side above it is updated, y_I -= L_{J,I}^T x_J (I = J-B .. J-1).  A step is a short serial
// 6x6 solve on lane 0 of the consumer warp plus 6B dot products of length 6 spread over its lanes.  Everything
// a step reads (block row J of L, 1/diag, the y row entering the window) is brought into a kBackDepth-deep
// shared-memory ring by the producer warp with 16-byte asynchronous copies that signal a per-slot mbarrier when
// they land; the consumer frees a slot with a second mbarrier.  The copies (~750 cycles of issue per block row)
// and the latency of the band store are thus off the serial chain.  Rows J >= n_elim are not solved: their x
// comes from `xsep` (the separator solution, physical order, separator-local) and only the update is applied.
// Call with the first 64 threads of the CTA; `sm` needs band_back_smem_doubles(B) doubles.
constexpr int kBackDepth = 8;
__host__ __device__ inline int band_back_stage_doubles(int B) { return (B + 1) * 36 + 12; }   // L row (d = 0..B) | dinv | y_in
__host__ __device__ inline int band_back_smem_doubles(int B) {
  return kBackDepth * band_back_stage_doubles(B) + (B + 2) * 6 + 6 * B + 8 + 2 * kBackDepth;   // ring | yw | xsep | xs | barriers
}

__device__ void band_backsub(int Pphys, int B, const BandPart& pt, int m, const double* __restrict__ xsep_g,
                             double* sm, Ws ws) {
  const int lane = threadIdx.x & 31, role = (threadIdx.x >> 5) & 1;      // role 0: consumer, 1: producer
  const int P = pt.Plog, NB = B + 2, SD = band_back_stage_doubles(B), RD = (B + 1) * 36;
  const double* Lband = ws.Lfac + pt.store;
  const double* dinv_g = Lband + (size_t)P * (B + 1) * 36;
  const double* y_g = dinv_g + 6 * P;
  double* ring = sm;
  double* yw = ring + kBackDepth * SD;
  double* xsep = yw + NB * 6;
  double* xs = xsep + 6 * B;
  uint64_t* full = reinterpret_cast<uint64_t*>(xs + 8);
  uint64_t* empty = full + kBackDepth;
  auto phys = [&](int I) { return pt.rev ? pt.base - I : pt.base + I; };
  (void)Pphys;

  if (threadIdx.x == 0) {
    for (int i = 0; i < kBackDepth; ++i) {
      mbar_init(smem_u32(&full[i]), 32);             // every producer lane arrives when its copies have landed
      mbar_init(smem_u32(&empty[i]), 1);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  // consumer set-up: window rows P-1 .. P-B (row P-1-B arrives with the first stage), separator solution
  if (role == 0) {
    for (int e = lane; e < B * 6; e += 32) {
      const int I = P - 1 - e / 6;
      if (I >= 0) yw[(I % NB) * 6 + e % 6] = I < pt.n_elim ? y_g[6 * I + e % 6] : 0.0;
    }
    if (xsep_g != nullptr)
      for (int e = lane; e < 6 * B; e += 32) xsep[e] = xsep_g[e];
  }
  asm volatile("bar.sync 1, 64;" ::: "memory");      // the two warps only (the rest of the CTA may be gone)

  if (role == 1) {
    // ---------------------------------------------------------------- producer
    // stage for step J: block row J of the band store (L_JJ, L_{J,J-1}, .. L_{J,J-B}; entries whose column was
    // never eliminated hold garbage and are skipped by the consumer), then 1/diag(J) and y of row J-B
    const unsigned ring_s = (unsigned)__cvta_generic_to_shared(ring) + lane * 16;
    const int npiece = RD / 2;                       // 16-byte pieces of a block row of L
    int slot = 0, use = 0;
    for (int J = P - 1; J >= 0; --J) {
      mbar_wait(smem_u32(&empty[slot]), (use + 1) & 1);
      const unsigned dst = ring_s + (unsigned)(slot * SD * 8);
      const double* src = Lband + (size_t)J * RD + 2 * lane;
      for (int q = 0; q * 32 + lane < npiece; ++q)
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + q * 512), "l"(src + q * 64) : "memory");
      const unsigned dtail = (unsigned)__cvta_generic_to_shared(ring + slot * SD + RD);
      if (lane < 3)
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dtail + 16 * lane),
                     "l"(dinv_g + 6 * (J < pt.n_elim ? J : 0) + 2 * lane) : "memory");
      if (lane >= 3 && lane < 6) {
        const int I = J - B, ok = I >= 0 && I < pt.n_elim;
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dtail + 16 * lane),
                     "l"(y_g + 6 * (ok ? I : 0) + 2 * (lane - 3)), "r"(ok ? 16 : 0) : "memory");
      }
      asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(smem_u32(&full[slot])) : "memory");
      if (++slot == kBackDepth) { slot = 0; ++use; }
    }
    asm volatile("cp.async.wait_all;" ::: "memory");
    return;
  }

  // ------------------------------------------------------------------ consumer
  // per-lane constants of the update: item u is (block offset o, component c) = divmod(lane + 32 u, 6)
  int it_o[5], it_c[5];
#pragma unroll
  for (int u = 0; u < 5; ++u) {
    it_o[u] = (lane + 32 * u) / 6;
    it_c[u] = (lane + 32 * u) - 6 * it_o[u];
  }
  int jr = (P - 1) % NB;                             // J mod NB, kept incrementally
  int jb = ((P - 1 - B) % NB + NB) % NB;             // (J - B) mod NB
  int slot = 0, use = 0;
#ifdef DBAF_BAND_PROFILE
  long long bt[4] = {0, 0, 0, 0}, b0, b1;
#define BK_TICK(i) do { b1 = clock64(); bt[i] += b1 - b0; b0 = b1; } while (0)
#else
#define BK_TICK(i) do {} while (0)
#endif
  for (int J = P - 1; J >= 0; --J) {
#ifdef DBAF_BAND_PROFILE
    b0 = clock64();
#endif
    mbar_wait(smem_u32(&full[slot]), use & 1);
    __syncwarp();                                    // also orders the previous step's yw updates before lane 0 reads
    BK_TICK(1);
    const double* st = ring + slot * SD;
    if (lane < 6 && J - B >= 0) yw[jb * 6 + lane] = st[RD + 6 + lane];
    if (lane == 0) {
      double x[6];
      if (J >= pt.n_elim) {
#pragma unroll
        for (int c = 0; c < 6; ++c) x[c] = xsep[6 * (phys(J) - m) + c];
      } else {
        const double* L = st;
        const double* di = st + RD;
#pragma unroll
        for (int r = 0; r < 6; ++r) x[r] = yw[jr * 6 + r];
        // column sweep: as soon as x[c] is final the rows above it are corrected, so the serial chain per
        // unknown is one multiply and one fused multiply-add
#pragma unroll
        for (int c = 5; c >= 0; --c) {
          x[c] *= di[c];
#pragma unroll
          for (int k = 0; k < 6; ++k)
            if (k < c) x[k] -= L[c * 6 + k] * x[c];
        }
        float* dxo = ws.dx + 6 * phys(J);
#pragma unroll
        for (int c = 0; c < 6; ++c) dxo[c] = (float)x[c];
      }
#pragma unroll
      for (int c = 0; c < 6; ++c) xs[c] = x[c];
    }
    __syncwarp();
    BK_TICK(2);
    double xv[6];
#pragma unroll
    for (int r = 0; r < 6; ++r) xv[r] = xs[r];
    const int nab = min(B, J);                       // rows above J inside the band
    const int first_elim = J - pt.n_elim;            // items with o < first_elim belong to separator x separator blocks
    // fixed trip count (6B <= 150 items over 32 lanes) so that the five dot products overlap; the panel blocks are
    // stored transposed (see the sweeps), so a dot product reads six contiguous doubles
#pragma unroll
    for (int u = 0; u < 5; ++u) {
      const int o = it_o[u], c = it_c[u];
      if (o < nab && o >= first_elim) {
        const double2* q = reinterpret_cast<const double2*>(st + (o + 1) * 36 + c * 6);
        const double2 q0 = q[0], q1 = q[1], q2 = q[2];
        double sum = q0.x * xv[0];
        sum = fma(q0.y, xv[1], sum);
        sum = fma(q1.x, xv[2], sum);
        sum = fma(q1.y, xv[3], sum);
        sum = fma(q2.x, xv[4], sum);
        sum = fma(q2.y, xv[5], sum);
        int row = jr - 1 - o;
        if (row < 0) row += NB;
        yw[row * 6 + c] -= sum;
      }
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(smem_u32(&empty[slot]));      // the producer may refill this slot
    jr = jr == 0 ? NB - 1 : jr - 1;
    jb = jb == 0 ? NB - 1 : jb - 1;
    if (++slot == kBackDepth) { slot = 0; ++use; }
    BK_TICK(3);
  }
#ifdef DBAF_BAND_PROFILE
  if (lane == 0 && pt.rev == 0) for (int i = 0; i < 4; ++i) ws.info[6 + i] = (int)(bt[i] >> 10);
#endif
}

// Band store of a part: [Plog][B+1][36] (entry (I, d) = L_{I, I-d}, row-major 6x6: a block row of L is one
// contiguous run, which is what the back-substitution streams; off-diagonal blocks are stored TRANSPOSED so that
// its dot products read contiguous doubles), then dinv [6 Plog],
// then y [6 Plog].
__global__ void __launch_bounds__(kBandThreads)
k_solve_banded(int Pphys, int B, float lm, float ep, int rezero, BandPlan plan, Ws ws) {
  extern __shared__ double smb[];
  const int tid = threadIdx.x, nt = blockDim.x, lane = tid & 31, wid = tid >> 5;
  (void)lane; (void)wid;
  const BandPart pt = plan.part[blockIdx.x];
  const int P = pt.Plog;                             // logical size seen by this sweep
  const int n = 6 * Pphys;                           // leading dimension of H
  const int NB = B + 2;
  double* Wd = smb;                                  // NB x NB blocks of 36
  double* yw = Wd + (size_t)NB * NB * 36;            // NB x 6 (right-hand side window, later x window)
  double* Ljj = yw + NB * 6;                         // 36
  double* dj = Ljj + 72;                             // 2 x 6 (1 / diag), double-buffered like Ljj (2 x 36)
  __shared__ int fail;
  double* Lband = ws.Lfac + pt.store;
  double* dinv_g = Lband + (size_t)P * (B + 1) * 36;
  double* y_g = dinv_g + 6 * P;
  // logical -> physical block index: p0 + sd * I
  const int p0 = pt.base, sd = pt.rev ? -1 : 1;
  auto slot = [&](int I, int K) { return Wd + ((size_t)(I % NB) * NB + (K % NB)) * 36; };
  const double epd = (double)ep, lmd = (double)lm;

  // element e of block row I: blocks K = max(0, I-B) .. I, then the 6 rhs entries
  auto row_elems = [&](int I) { return (min(I, B) + 1) * 36 + 6; };
  auto fetch = [&](int I, int e) -> double {
    const int nb = min(I, B) + 1;
    if (e < nb * 36) {
      const int b = e / 36, rc = e - b * 36, r = rc / 6, c = rc - r * 6;
      const int K = I - (nb - 1) + b;
      double a = ws.H[(size_t)(6 * (p0 + sd * I) + r) * n + 6 * (p0 + sd * K) + c];
      if (K == I && r == c) a += epd + lmd * a;
      return a;
    }
    return ws.v[6 * (p0 + sd * I) + (e - nb * 36)];
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

  // Factorises diagonal block J, forward-solves its right-hand side and publishes L_JJ and 1/diag into
  // buffer J & 1.  Runs on one whole warp: 21 lanes first apply the pending rank-6 update of the block
  // (`upd` = panel block L_{J,J-1}, or nullptr) in parallel, lane 0 then runs the serial pivot chain in
  // registers, and all lanes copy the results to the band store in global memory.
#ifdef DBAF_BAND_PROFILE
  long long fd[5] = {0, 0, 0, 0, 0};
#endif
  auto factor_diag = [&](int J, const double* upd, int ln, bool factorise) {
#ifdef DBAF_BAND_PROFILE
    long long f0 = clock64(), f1;
#define FD_TICK(i) do { f1 = clock64(); fd[i] += f1 - f0; f0 = f1; } while (0)
#else
#define FD_TICK(i) do {} while (0)
#endif
    double* A = slot(J, J);
    if (upd && ln < 21) {
      int r, c;
      tri_unpack(ln, r, c);
      double sum = 0.0;
#pragma unroll
      for (int k = 0; k < 6; ++k) sum += upd[r * 6 + k] * upd[c * 6 + k];
      A[r * 6 + c] -= sum;
    }
    __syncwarp();
    FD_TICK(0);
    if (!factorise) return;                           // first separator block: updated, never eliminated here
    double* Lj = Ljj + (J & 1) * 36;
    double* dd = dj + (J & 1) * 6;
    if (ln == 0) {
      double a[6][6];
#pragma unroll
      for (int r = 0; r < 6; ++r)
#pragma unroll
        for (int c = 0; c < 6; ++c)
          if (c <= r) a[r][c] = A[r * 6 + c];
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
            double sum = a[r][c];
#pragma unroll
            for (int k = 0; k < 6; ++k)
              if (k < c) sum -= a[r][k] * a[c][k];
            a[r][c] = sum * inv;
          }
      }
      if (bad) fail = 1;
      FD_TICK(1);
      double y[6];
#pragma unroll
      for (int r = 0; r < 6; ++r) {
        double sum = yw[(J % NB) * 6 + r];
#pragma unroll
        for (int k = 0; k < 6; ++k)
          if (k < r) sum -= a[r][k] * y[k];
        y[r] = sum * di[r];
      }
#pragma unroll
      for (int r = 0; r < 6; ++r) {
        yw[(J % NB) * 6 + r] = y[r];
        dd[r] = di[r];
#pragma unroll
        for (int c = 0; c < 6; ++c) Lj[r * 6 + c] = c <= r ? a[r][c] : 0.0;
      }
      FD_TICK(2);
    }
    __syncwarp();
    FD_TICK(3);
    // band store (needed again only by the back-substitution)
    for (int e = ln; e < 36; e += 32) Lband[((size_t)J * (B + 1)) * 36 + e] = Lj[e];
    if (ln < 6) {
      y_g[6 * J + ln] = yw[(J % NB) * 6 + ln];
      dinv_g[6 * J + ln] = dd[ln];
    }
    FD_TICK(4);
  };

  if (tid == 0) fail = 0;
  for (int I = 0; I <= min(B, P - 1); ++I)
    for (int e = tid; e < row_elems(I); e += nt) place(I, e, fetch(I, e));
  __syncthreads();
  if (tid >= nt - 32 && pt.n_elim > 0) factor_diag(0, nullptr, tid - (nt - 32), true);
  __syncthreads();

  const int nwork = nt - 32;                         // the last warp runs the look-ahead factorisation
  // Incoming rows are always full (B+1 blocks + rhs) inside the loop, so each thread's two elements keep a
  // fixed position inside the row: only the row index moves (pointer += 6n+6 per step).
  int pf_kind[2], pf_b[2], pf_rc[2];
  long long pf_off[2];
  const long long pf_stride = (long long)sd * (6LL * n + 6);
#pragma unroll
  for (int u = 0; u < 2; ++u) {
    const int e = tid + u * nt;
    pf_kind[u] = 0;
    pf_b[u] = pf_rc[u] = 0;
    pf_off[u] = 0;
    if (e < (B + 1) * 36) {
      const int bb = e / 36, rc = e - bb * 36, r = rc / 6, c = rc - r * 6;
      pf_kind[u] = (bb == B && r == c) ? 2 : 1;     // 2 = diagonal entry (damped)
      pf_b[u] = bb;
      pf_rc[u] = rc;
      // H[(6 phys(I) + r) n + 6 phys(I - B + bb) + c] = pf_off + I * pf_stride
      pf_off[u] = (long long)r * n + c + (6LL * n + 6) * p0 + 6LL * sd * (bb - B);
    } else if (e < (B + 1) * 36 + 6) {
      pf_kind[u] = 3;                                // right-hand side entry
      pf_rc[u] = e - (B + 1) * 36;
    }
  }
  int jm = 0;                                        // J % NB, kept incrementally
#ifdef DBAF_BAND_PROFILE
  long long tp0 = 0, tp1 = 0, tp2 = 0, tp3 = 0, tla = 0, c0, c1;
#define BAND_TICK(var) do { c1 = clock64(); var += c1 - c0; c0 = c1; } while (0)
#else
#define BAND_TICK(var) do {} while (0)
#endif
  for (int J = 0; J < pt.n_elim; ++J) {
    if (fail) break;
#ifdef DBAF_BAND_PROFILE
    c0 = clock64();
#endif
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
        if (pf_kind[u] == 3) pre[u] = ws.v[6 * (p0 + sd * Inew) + pf_rc[u]];
        else if (pf_kind[u]) pre[u] = ws.H[pf_off[u] + (long long)Inew * pf_stride];
      }
    }
    BAND_TICK(tp0);
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
        Lband[((size_t)I * (B + 1) + (I - J)) * 36 + c * 6 + r] = x[c];     // transposed: see band_backsub
        yy -= x[c] * yw[(J % NB) * 6 + c];
      }
      yw[(I % NB) * 6 + r] = yy;
    }
    BAND_TICK(tp1);
    __syncthreads();
    BAND_TICK(tp2);
    // (3) trailing update inside the window: A_IK -= L_IJ L_KJ^T for J < K <= I <= last.  4 threads
    //     share a row (I, r) and split the K range.  The next diagonal block (J+1, J+1) is left to the
    //     last warp, which updates AND factorises it at the same time (look-ahead), so the serial
    //     rsqrt chain of the factorisation overlaps the bulk of the update.
    if (tid < nwork) {
      // one thread = half (3 rows) of one 6x6 block (I, K): 54 shared-memory loads feed 108 FMAs
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
    } else if (J + 1 < P) {
      factor_diag(J + 1, nbelow > 0 ? slot(J + 1, J) : nullptr, tid - nwork, J + 1 < pt.n_elim);
    }
    BAND_TICK(tp3);
    // incoming block row (its slots are unused during this step: NB = B + 2)
    if (Inew < P) {
      int rs = jm - 1;                               // (J + B + 1) % NB
      if (rs < 0) rs += NB;
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        if (pf_kind[u] == 3) {
          yw[rs * 6 + pf_rc[u]] = pre[u];
        } else if (pf_kind[u]) {
          int cs = jm + 1 + pf_b[u];                 // (Inew - B + b) % NB
          if (cs >= NB) cs -= NB;
          const double a = pf_kind[u] == 2 ? pre[u] + (epd + lmd * pre[u]) : pre[u];
          Wd[((size_t)rs * NB + cs) * 36 + pf_rc[u]] = a;
        }
      }
    }
    if (++jm == NB) jm = 0;
    __syncthreads();
    BAND_TICK(tla);
  }
#ifdef DBAF_BAND_PROFILE
  if (tid == 0 || tid == nwork) {
    const int o = tid == 0 ? 6 : 11;      // info[6..10]: thread 0, info[11..15]: look-ahead thread (kilo-cycles)
    ws.info[o] = (int)(tp0 >> 10); ws.info[o + 1] = (int)(tp1 >> 10); ws.info[o + 2] = (int)(tp2 >> 10);
    ws.info[o + 3] = (int)(tp3 >> 10); ws.info[o + 4] = (int)(tla >> 10);
    if (tid == nwork) for (int i = 0; i < 5; ++i) ws.info[11 + i] = (int)(fd[i] >> 10);
  }
#endif
  __syncthreads();
  const bool ok = fail == 0;
  if (plan.twisted) {
    // leave this sweep's view of the separator (blocks [n_elim, Plog) of the window and their rhs) for k_band_sep
    double* dump = ws.Lfac + pt.dump;
    for (int e = tid; e < B * B * 36; e += nt) {
      const int blk = e / 36, ai = blk / B, ak = blk - ai * B;
      if (ai >= ak) dump[e] = slot(pt.n_elim + ai, pt.n_elim + ak)[e - blk * 36];
    }
    for (int e = tid; e < 6 * B; e += nt) dump[(size_t)B * B * 36 + e] = yw[((pt.n_elim + e / 6) % NB) * 6 + e % 6];
    if (tid == 0) reinterpret_cast<int*>(ws.Lfac + plan.flags)[blockIdx.x] = ok ? 0 : 1;
    return;
  }
  if (ok) {
    // the window is no longer needed: its shared memory becomes the back-substitution's staging ring
    if (tid < 64) band_backsub(Pphys, B, pt, 0, nullptr, smb, ws);
  } else {
    for (int i = tid; i < n; i += nt) ws.dx[i] = 0.0f;
    if (tid == 0) atomicAdd(&ws.info[1], 1);
  }
  (void)rezero;
}

// Separator of the twisted factorisation: S = S_top + S_bottom - A_ss (A_ss damped like everywhere else),
// rhs likewise, dense Cholesky in shared memory, x_sep to global memory and to dx.
__global__ void __launch_bounds__(kThreads)
k_band_sep(int Pphys, int B, float lm, float ep, BandPlan plan, Ws ws) {
  extern __shared__ double sms[];
  __shared__ int fail;
  const int tid = threadIdx.x, nt = blockDim.x;
  const int n = 6 * Pphys, ns = 6 * B, m = plan.m;
  const int ld = ns | 1;
  double* L = sms;                                   // (ns+1) x ld
  double* x = L + (size_t)(ns + 1) * ld;             // ns
  double* dinv = x + ns;                             // ns
  const double* top = ws.Lfac + plan.sep;
  const double* bot = top + band_dump_doubles(B);
  double* xsep = ws.Lfac + plan.sep + 2 * band_dump_doubles(B);
  int* flags = reinterpret_cast<int*>(xsep + ns);
  const bool sweeps_ok = flags[0] == 0 && flags[1] == 0;
  for (int e = tid; e < B * B * 36; e += nt) {
    const int blk = e / 36, a = blk / B, b = blk - a * B;
    if (a < b) continue;
    const int rc = e - blk * 36, r = rc / 6, c = rc - r * 6;
    if (a == b && c > r) continue;
    // the bottom sweep works on the block-reversed ordering: physical block (m+a, m+b), a >= b, is its upper
    // triangle, i.e. the transpose of its logical block (B-1-b, B-1-a)
    double h = ws.H[(size_t)(6 * (m + a) + r) * n + 6 * (m + b) + c];
    if (a == b && r == c) h += (double)ep + (double)lm * h;
    // (diagonal blocks are symmetric in themselves; only their lower triangle is kept up to date by the sweeps)
    const int brc = a == b ? r * 6 + c : c * 6 + r;
    L[(size_t)(6 * a + r) * ld + 6 * b + c] =
        top[(size_t)(a * B + b) * 36 + r * 6 + c] + bot[(size_t)((B - 1 - b) * B + (B - 1 - a)) * 36 + brc] - h;
  }
  for (int e = tid; e < ns; e += nt) {
    const int a = e / 6, r = e - a * 6;
    L[(size_t)ns * ld + e] = top[(size_t)B * B * 36 + e] + bot[(size_t)B * B * 36 + (B - 1 - a) * 6 + r] - ws.v[6 * m + e];
  }
  __syncthreads();
  const bool ok = cholesky6_solve(L, ld, B, x, dinv, &fail) && sweeps_ok;
  __syncthreads();
  for (int e = tid; e < ns; e += nt) {
    xsep[e] = ok ? x[e] : 0.0;
    ws.dx[6 * m + e] = ok ? (float)x[e] : 0.0f;
  }
  if (tid == 0) {
    flags[2] = ok ? 0 : 1;
    if (!ok) atomicAdd(&ws.info[1], 1);
  }
}

__global__ void __launch_bounds__(64)
k_band_back(int Pphys, int B, BandPlan plan, Ws ws) {
  extern __shared__ double smk[];
  const BandPart pt = plan.part[blockIdx.x];
  const double* xsep = ws.Lfac + plan.sep + 2 * band_dump_doubles(B);
  const int* flags = reinterpret_cast<const int*>(xsep + 6 * B);
  if (flags[2] != 0) {
    // a pivot failed somewhere: the whole step is dropped (dx = 0), like the single sweep does
    for (int I = threadIdx.x; I < pt.n_elim * 6; I += blockDim.x) {
      const int J = I / 6;
      ws.dx[6 * (pt.rev ? pt.base - J : pt.base + J) + I % 6] = 0.0f;
    }
    return;
  }
  band_backsub(Pphys, B, pt, pt.sep_first, xsep, smk, ws);
}
