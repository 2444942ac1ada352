// Banded sweep with the window held in REGISTERS, included by ba.cu after ba_banded.cuh.
//
// k_solve_banded keeps the (B+2)^2-block window in shared memory and spends ~4 000 of its ~5 300 cycles per
// block column moving blocks through it (54 + 36 eight-byte accesses per 108 FMAs).  Here every thread owns one
// half (3 x 6) of one window block for as long as that block is alive and keeps it in registers; shared memory
// only holds the current panel column (B blocks), the diagonal factor and the right-hand-side window.
//
// Ownership is static.  A block (I, K), I >= K, belongs to the slot pair {I mod NB, K mod NB} (NB = B + 2), and
// because the window spans fewer than NB block rows exactly one of the slots (a, b), (b, a) is alive at a time:
// the NB (NB - 1) / 2 pairs x 2 halves plus NB diagonal blocks give NB^2 <= 512 threads for B <= 20, each with at
// most one live block and therefore one role per step:
//   * its block is in column J            -> panel: X L_JJ^T = A for its 3 rows, publish to the panel buffer
//   * its block lies in the trailing part -> update: C -= L_I L_K^T from the panel buffer (54 loads, 108 FMAs)
//   * its block row enters the window     -> load the block from H straight into the registers
//   * diagonal block of column J + 1      -> look-ahead: update, factorise, forward-solve the rhs, publish L_JJ
// Two barriers per block column.  Same band store, same twisted plan, same back-substitution as k_solve_banded.
#pragma once

constexpr int kBandRegMaxB = 20;     // (B + 2)^2 <= 512 threads
constexpr int kPnS = 38;             // block stride of the panel buffers: 16-byte aligned rows, conflict-free 128-bit reads

__global__ void __launch_bounds__(kBandThreads)
k_solve_banded_reg(int Pphys, int B, float lm, float ep, BandPlan plan, Ws ws) {
  extern __shared__ double smr[];
  __shared__ int fail;
  const int tid = threadIdx.x, nt = blockDim.x;
  const BandPart pt = plan.part[blockIdx.x];
  const int P = pt.Plog;
  const int n = 6 * Pphys;
  const int NB = B + 2;
  // panel blocks are kPnS = 38 doubles apart: lanes that read different blocks hit different banks, and every
  // block starts on a 16-byte boundary so that two doubles come with one load
  double* Pn = smr;                       // NB x 38: panel blocks L_{I,J} of the current column, by I mod NB
  double* Praw = Pn + NB * kPnS;          // NB x 38: the same blocks before the triangular solve
  double* yw = Praw + NB * kPnS;          // NB x 6: right-hand-side window
  double* Ljj = yw + NB * 6;              // 2 x 36
  double* dj = Ljj + 72;                  // 2 x 6
  // landing area of the block row that enters the window: one slot per task (18 doubles, 36 for a diagonal task).
  // The copies are asynchronous global->shared ones on purpose: a plain load into the owner's registers would
  // make every other lane of its warp wait for it at the next use of the same register names.
  double* Inb = dj + 12 + 4;
  double* Lband = ws.Lfac + pt.store;
  double* dinv_g = Lband + (size_t)P * (B + 1) * 36;
  double* y_g = dinv_g + 6 * P;
  // Domain-decomposed solve: the spike kernel of this domain (k_spike_fwd, launched with programmatic stream
  // serialisation) may start now and follow this sweep row by row: progress[d] = number of finished block rows of the
  // band store (the flag words were cleared by a memset ahead of this kernel).
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  volatile int* progress = (plan.twisted && gridDim.x >= 3)      // (D = gridDim.x domains; flag words: see ba_spike.cuh)
                               ? reinterpret_cast<volatile int*>(ws.Lfac + plan.flags) + gridDim.x + 2 + blockIdx.x : nullptr;
  const int p0 = pt.base, sd = pt.rev ? -1 : 1;
  const double epd = (double)ep, lmd = (double)lm;

  // ---- static task of this thread
  const int n_off = NB * (NB - 1);
  int ta = -1, tb = -1, half = 0;
  bool diag = false;
  if (tid < n_off) {
    int rest = tid;
    for (int a = 0; a < NB - 1; ++a) {
      const int cnt = NB - 1 - a;          // partners b > a; half 0 first, then half 1: a warp shares (a, half)
      if (rest < 2 * cnt) {
        ta = a;
        half = rest >= cnt;
        tb = a + 1 + (rest - half * cnt);
        break;
      }
      rest -= 2 * cnt;
    }
  } else if (tid < n_off + NB) {
    diag = true;
    ta = tb = tid - n_off;
  }
  // one register set serves both kinds of task: an off-diagonal task keeps rows 3*half .. +2 of its live block in
  // C[0..18), a diagonal task the lower triangle of its block, row by row, in D[0..21)
  double R[21];
  double* const C = R;
  double* const D = R;
#pragma unroll
  for (int i = 0; i < 21; ++i) R[i] = 0.0;

  auto off_of = [&](int res, int jm) { const int o = res - jm; return o < 0 ? o + NB : o; };
  auto load_off = [&](int I, int K) {      // block (I, K), this thread's three rows
    const double* src = ws.H + (size_t)(6 * (p0 + sd * I) + 3 * half) * n + 6 * (p0 + sd * K);
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
      for (int c = 0; c < 6; ++c) C[r * 6 + c] = src[(size_t)r * n + c];
  };
  auto load_diag = [&](int I) {            // lower triangle of (I, I), damped, and the rhs row
    const double* src = ws.H + (size_t)(6 * (p0 + sd * I)) * n + 6 * (p0 + sd * I);
    int l = 0;
#pragma unroll
    for (int r = 0; r < 6; ++r)
#pragma unroll
      for (int c = 0; c <= r; ++c) {
        double a = src[(size_t)r * n + c];
        if (r == c) a += epd + lmd * a;
        D[l++] = a;
      }
#pragma unroll
    for (int r = 0; r < 6; ++r) yw[(I % NB) * 6 + r] = ws.v[6 * (p0 + sd * I) + r];
  };
  // D -= p p^T (lower triangle), p = panel block of this block row
  auto update_diag = [&](const double* p) {
#pragma unroll
    for (int k = 0; k < 6; ++k) {
      double pk[6];
#pragma unroll
      for (int r = 0; r < 6; ++r) pk[r] = p[r * 6 + k];
      int l = 0;
#pragma unroll
      for (int r = 0; r < 6; ++r)
#pragma unroll
        for (int c = 0; c <= r; ++c) D[l++] -= pk[r] * pk[c];
    }
  };
  // factorise D in place (right-looking, registers), forward-solve the rhs of block row J, publish everything
  auto factor_publish = [&](int J) {
#define DL(r, c) D[(r) * ((r) + 1) / 2 + (c)]
    double di[6];
    bool bad = false;
#pragma unroll
    for (int c = 0; c < 6; ++c) {
      double d = DL(c, c);
      if (d <= 0.0) { bad = true; d = 1.0; }
      const double inv = rsqrt_nr(d);
      di[c] = inv;
      DL(c, c) = d * inv;
#pragma unroll
      for (int r = 0; r < 6; ++r)
        if (r > c) DL(r, c) *= inv;
#pragma unroll
      for (int r = 0; r < 6; ++r)
#pragma unroll
        for (int k = 0; k < 6; ++k)
          if (r > c && k > c && k <= r) DL(r, k) -= DL(r, c) * DL(k, c);
    }
    if (bad) fail = 1;
    double y[6];
#pragma unroll
    for (int r = 0; r < 6; ++r) y[r] = yw[(J % NB) * 6 + r];
#pragma unroll
    for (int c = 0; c < 6; ++c) {
      y[c] *= di[c];
#pragma unroll
      for (int r = 0; r < 6; ++r)
        if (r > c) y[r] -= DL(r, c) * y[c];
    }
    // shared memory only: the band-store copy is written by idle threads at the top of step J
    double* Lj = Ljj + (J & 1) * 36;
    double* dd = dj + (J & 1) * 6;
#pragma unroll
    for (int r = 0; r < 6; ++r) {
      yw[(J % NB) * 6 + r] = y[r];
      dd[r] = di[r];
#pragma unroll
      for (int c = 0; c < 6; ++c) Lj[r * 6 + c] = c <= r ? DL(r, c) : 0.0;
    }
#undef DL
  };

  // asynchronous request of this thread's part of block row R (it lands in the thread's slot of Inb)
  auto request_off = [&](int R, int K) {
    const double* src = ws.H + (size_t)(6 * (p0 + sd * R) + 3 * half) * n + 6 * (p0 + sd * K);
    const unsigned dst = (unsigned)__cvta_generic_to_shared(Inb + tid * 18);
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
      for (int q = 0; q < 3; ++q)
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + (r * 6 + 2 * q) * 8), "l"(src + (size_t)r * n + 2 * q) : "memory");
  };
  auto request_diag = [&](int R) {          // full 6x6 block and the 6 rhs entries behind it
    const double* src = ws.H + (size_t)(6 * (p0 + sd * R)) * n + 6 * (p0 + sd * R);
    const unsigned dst = (unsigned)__cvta_generic_to_shared(Inb + n_off * 18 + ta * 42);
#pragma unroll
    for (int r = 0; r < 6; ++r)
#pragma unroll
      for (int q = 0; q < 3; ++q)
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + (r * 6 + 2 * q) * 8), "l"(src + (size_t)r * n + 2 * q) : "memory");
#pragma unroll
    for (int q = 0; q < 3; ++q)
      asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + (36 + 2 * q) * 8), "l"(ws.v + 6 * (p0 + sd * R) + 2 * q) : "memory");
  };
  auto pickup = [&]() {
    if (diag) {
      const double* in = Inb + n_off * 18 + ta * 42;
      int l = 0;
#pragma unroll
      for (int r = 0; r < 6; ++r)
#pragma unroll
        for (int c = 0; c <= r; ++c) {
          double a = in[r * 6 + c];
          if (r == c) a += epd + lmd * a;
          D[l++] = a;
        }
#pragma unroll
      for (int r = 0; r < 6; ++r) yw[ta * 6 + r] = in[36 + r];
    } else {
      const double* in = Inb + tid * 18;
#pragma unroll
      for (int i = 0; i < 18; ++i) C[i] = in[i];
    }
  };
  auto stage_raw = [&](int res_row) {       // hand a block of the next panel column to the row solvers
#pragma unroll
    for (int i = 0; i < 18; ++i) Praw[res_row * kPnS + half * 18 + i] = C[i];
  };

  // ---- initial window: block rows 0 .. min(B, P-1) directly, block row B+1 requested (picked up after step 0)
  if (tid == 0) fail = 0;
  const int top = min(B, P - 1);
  int pend = 0;                            // steps until this thread's pending request is picked up
  if (ta >= 0) {
    if (diag) {
      if (ta <= top) load_diag(ta);
      if (ta == B + 1 && B + 1 < P) { request_diag(B + 1); pend = 1; }
    } else {
      if (tb <= top) load_off(tb, ta);     // at J = 0 residues are the block indices themselves, tb > ta
      if (tb == B + 1 && ta >= 1 && B + 1 < P) { request_off(B + 1, ta); pend = 1; }
      if (ta == 0 && tb <= top) stage_raw(tb);         // column 0 of the panel
    }
  }
  asm volatile("cp.async.commit_group;" ::: "memory");
  __syncthreads();
  if (diag && ta == 0 && pt.n_elim > 0) factor_publish(0);
  __syncthreads();

  int jm = 0;
#ifdef DBAF_BAND_PROFILE
  __shared__ long long prof[8];
  if (tid < 8) prof[tid] = 0;
  long long q0, q1;
#define RP_TICK(i) do { q1 = clock64(); if (tid == 0) prof[i] += q1 - q0; q0 = q1; } while (0)
#else
#define RP_TICK(i) do {} while (0)
#endif
  for (int J = 0; J < pt.n_elim; ++J) {
    if (fail) break;
#ifdef DBAF_BAND_PROFILE
    q0 = clock64();
#endif
    const int nbelow = min(J + B, P - 1) - J;
    const double* Lj = Ljj + (J & 1) * 36;
    const double* dd = dj + (J & 1) * 6;
    // (1) row solve of the panel column J (its blocks were staged in Praw at the end of the previous step):
    //     6 nbelow consecutive threads, one row each, X L_JJ^T = A; the task-less threads behind the task
    //     threads write the diagonal factor of this column to the band store
    if (tid < nbelow * 6) {
      const int o = tid / 6, r = tid - o * 6;
      const int I = J + 1 + o;
      int rs = jm + 1 + o;
      if (rs >= NB) rs -= NB;
      const double* A = Praw + rs * kPnS + r * 6;
      // column sweep: as soon as x[c] is final the later entries are corrected, so the serial chain per unknown
      // is one multiply and one fused multiply-add (12 dependent operations instead of 21)
      double x[6], lj[15], yj[6];
#pragma unroll
      for (int c = 0; c < 6; ++c) {
        x[c] = A[c];
        yj[c] = yw[jm * 6 + c];
      }
      {
        int l = 0;
#pragma unroll
        for (int c = 1; c < 6; ++c)
#pragma unroll
          for (int k = 0; k < c; ++k) lj[l++] = Lj[c * 6 + k];
      }
      double yy = 0.0;
#pragma unroll
      for (int c = 0; c < 6; ++c) {
        x[c] *= dd[c];
#pragma unroll
        for (int k = c + 1; k < 6; ++k) x[k] = fma(-x[c], lj[k * (k - 1) / 2 + c], x[k]);
        yy = fma(x[c], yj[c], yy);
      }
#pragma unroll
      for (int c = 0; c < 6; ++c) {
        Pn[rs * kPnS + r * 6 + c] = x[c];
        Lband[((size_t)I * (B + 1) + (I - J)) * 36 + c * 6 + r] = x[c];     // transposed: see band_backsub
      }
      yw[rs * 6 + r] -= yy;
    } else if (tid >= nt - 32) {
      const int u = tid - (nt - 32);       // NB^2 <= 484 leaves the last warp without tasks
      for (int e = u; e < 36; e += 32) Lband[((size_t)J * (B + 1)) * 36 + e] = Lj[e];
      if (u < 6) {
        dinv_g[6 * J + u] = dd[u];
        y_g[6 * J + u] = yw[jm * 6 + u];
      }
    }
    RP_TICK(0);
    __syncthreads();
    RP_TICK(1);
    // block row J of the band store is complete (its diagonal block and 1/diag were the last pieces): tell the
    // spike kernel.  The last warp has no task in the trailing update.
    if (progress != nullptr && tid == nt - 32) {
      __threadfence();
      *progress = J + 1;
    }
    // (2) trailing update, look-ahead factorisation of (J+1, J+1), requests for the block row two steps ahead
    int oi = -1, ok = -1, ri = 0, rk = 0, oa = -1;
    if (ta >= 0) {
      oa = off_of(ta, jm);
      if (!diag) {
        const int ob = off_of(tb, jm);
        if (oa > ob) { oi = oa; ok = ob; ri = ta; rk = tb; } else { oi = ob; ok = oa; ri = tb; rk = ta; }
      }
    }
    const int R2 = J + B + 2;
    if (R2 < P) {
      // the owners are the tasks that hold the (finished) column J: their registers and slots are free
      if (!diag && ok == 0 && oi >= 2) { request_off(R2, J + oi); pend = 2; }
      if (diag && oa == 0) { request_diag(R2); pend = 2; }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
    if (!diag) {
      if (ok >= 1 && oi <= nbelow) {
        // C -= L_I L_K^T for this thread's three rows: 27 two-double loads feed 108 fused multiply-adds
        const double2* Li = reinterpret_cast<const double2*>(Pn + ri * kPnS + half * 18);
        const double2* Lk = reinterpret_cast<const double2*>(Pn + rk * kPnS);
        double li[18];
#pragma unroll
        for (int i = 0; i < 9; ++i) {
          const double2 v = Li[i];
          li[2 * i] = -v.x;
          li[2 * i + 1] = -v.y;
        }
#pragma unroll
        for (int c = 0; c < 6; ++c) {
          double lk[6];
#pragma unroll
          for (int k = 0; k < 3; ++k) {
            const double2 v = Lk[c * 3 + k];
            lk[2 * k] = v.x;
            lk[2 * k + 1] = v.y;
          }
#pragma unroll
          for (int r = 0; r < 3; ++r)
#pragma unroll
            for (int k = 0; k < 6; ++k) C[r * 6 + c] = fma(li[r * 6 + k], lk[k], C[r * 6 + c]);
        }
      }
    } else if (oa >= 1 && oa <= nbelow) {
      update_diag(Pn + ta * kPnS);
#ifdef DBAF_BAND_PROFILE
      const long long l1 = clock64();
#endif
      if (oa == 1 && J + 1 < pt.n_elim) factor_publish(J + 1);
#ifdef DBAF_BAND_PROFILE
      if (oa == 1) prof[5] += clock64() - l1;
#endif
    }
    // (3) last step's request has landed: pick it up; then stage the blocks of the next panel column
    asm volatile("cp.async.wait_group 1;" ::: "memory");
    if (pend && --pend == 0) pickup();
    if (!diag && ok == 1 && oi >= 2 && J + 1 < pt.n_elim && J + oi <= P - 1) stage_raw(ri);
    if (++jm == NB) jm = 0;
    RP_TICK(2);
    __syncthreads();
    RP_TICK(3);
  }
  asm volatile("cp.async.wait_group 0;" ::: "memory");
  __syncthreads();
#ifdef DBAF_BAND_PROFILE
  if (tid < 6 && pt.rev == 0) ws.info[10 + tid] = (int)(prof[tid] >> 10);
#endif
  const bool ok_all = fail == 0;
  if (progress != nullptr && tid == 0) {         // a failed sweep must not leave its spike kernel waiting
    __threadfence();
    *progress = 1 << 30;
  }
  if (plan.twisted) {
    // this sweep's view of the separator [n_elim, Plog): window blocks and rhs, for k_band_sep
    double* dump = ws.Lfac + pt.dump;
    if (ta >= 0) {
      const int oa = off_of(ta, jm), ob = off_of(tb, jm);
      if (oa < B && ob < B) {
        if (diag) {
          int l = 0;
#pragma unroll
          for (int r = 0; r < 6; ++r)
#pragma unroll
            for (int c = 0; c <= r; ++c) dump[(size_t)(oa * B + oa) * 36 + r * 6 + c] = D[l++];
        } else {
          const int ai = max(oa, ob), ak = min(oa, ob);
#pragma unroll
          for (int r = 0; r < 3; ++r)
#pragma unroll
            for (int c = 0; c < 6; ++c) dump[(size_t)(ai * B + ak) * 36 + (3 * half + r) * 6 + c] = C[r * 6 + c];
        }
      }
    }
    for (int e = tid; e < 6 * B; e += nt) dump[(size_t)B * B * 36 + e] = yw[((pt.n_elim + e / 6) % NB) * 6 + e % 6];
    if (tid == 0) reinterpret_cast<int*>(ws.Lfac + plan.flags)[blockIdx.x] = ok_all ? 0 : 1;
    return;
  }
  if (ok_all) {
    if (tid < 64) band_backsub(Pphys, B, pt, 0, nullptr, smr, ws);
  } else {
    for (int i = tid; i < n; i += nt) ws.dx[i] = 0.0f;
    if (tid == 0) atomicAdd(&ws.info[1], 1);
  }
}
