// Linearisation and Schur complement of a depth frame in ONE kernel, included by ba.cu inside namespace
// dbaf::{anonymous} after ba_schur_tc.cuh (whose tile constants, instruction descriptor and barrier scheme it
// shares).  Dense frame graphs only (global BA: 8..21 row blocks, at most 20 out-edges per frame).
//
// k_linearize2 + k_schur_tc write the pose-depth blocks Eij (24 B per pixel-edge, 2 GB at the 1000-keyframe size)
// to HBM once and read them twice (Schur complement, back-substitution).  Here they never leave the SM: producer
// warps turn targets / weights / disparity (20 B per pixel-edge) straight into the two TF32 operand tiles of
//     S = X X^T,  X = E sqrt(q)                                    (ba_schur_tc.cuh)
// and accumulate the pose blocks on the way; the back-substitution (k_backsub_rc) recomputes E from the weights
// alone (8 B per pixel-edge).  What the reference does in projective_transform_kernel, accum_cuda and
// EEt6x6 / Ev6x1 (src/droid_kernels.cu:220-468, 993-1043, 1046-1138).
//
// Two kernels.  k_depth_diag first forms the depth diagonal C = sum_e w Jz^2 (+ prior / eta) and q = 1 / C from the
// weights alone and writes the per-edge relative-pose records: q scales the operand tiles, so it has to be complete
// before the first tile is written (as a pass inside the producer warps it could not hide its loads: DESIGN.md 3).
// k_linschur_tc (384 threads, one CTA per SM): warps 0-3 drain the TMEM accumulators into fp32 running sums (one
// thread per row of S), warp 4 issues the MMAs, warps 6-11 are producers.  Producer p owns operand stage p; the
// 32-pixel slices are dealt to the producers in a fixed round of 14 (see c_ls_order).  Inside the warp, lane =
// (group g = lane / 8, quad l8 = lane % 8): a thread works on the four consecutive pixels 4 l8 .. 4 l8 + 3 of the
// slice (one 16-byte chunk of an operand row) and group g on the edges a = g, g + 4, ... of the frame, so every
// global load and every shared-memory store is 128 bits wide.  Per edge: residuals and Jacobians, E sqrt(q) split
// into TF32 hi / lo and stored into the operand rows 6 a .. 6 a + 5, the 21 + 6 terms of Hjj / vj summed over the
// thread's four pixels, reduced over the eight lanes of the group (transposed butterfly, 28 shuffles for 4 edges x
// 32 pixels) and added to the warp's own partial sums in shared memory.
// The self rows (pose-i side of every edge, Ei = -sum_e M_e Ej_e) are not multiplied: their blocks of S are linear
// images of the edge blocks, S[self][f] = -sum_e M_e S[e][f], formed from the finished S in the epilogue.
#pragma once

constexpr int kLsProducers = 6;                       // producer warps = operand stages
constexpr int kLsDrainWarps = 4;
constexpr int kLsFirstProducer = 6;                   // warps 0-3 drain, 4 MMA, 5 spare, 6.. producers
constexpr int kLsThreads = (kLsFirstProducer + kLsProducers) * 32;
constexpr int kLsTabMax = 512;                        // frames up to 512 x 512 (at 1/8 resolution) use the coordinate tables
constexpr int kLsWRow = 126;                          // operand row that carries w sqrt(q)
constexpr int kLsSld = 129;                           // leading dimension of S in shared memory (epilogue)

struct LsEdge {          // 28 words, 16-byte aligned: what the kernels need of an edge
  float q[4];            // rotation of G_ij (x y z w)
  float t[3];            // translation of G_ij
  int flags;             // bit 0: stereo edge (depth blocks only)
  float Rt[9];           // R(q_ij)^T, row-major
  float Nm[9];           // -R^T [t]x: M = [[Rt, 0], [Nm, Rt]] carries a pose-j Jacobian to pose i (negated)
  int jx;                // target frame
  int prow;              // pose index of the target (jx - t0) if inside the window, else -1
};
constexpr int kLsEdgeWords = 28;
static_assert(sizeof(LsEdge) == kLsEdgeWords * 4, "LsEdge is copied with 16-byte loads");

constexpr size_t kLsPartBytes = (size_t)kLsProducers * kLsMaxDeg * 32 * 4;
constexpr size_t kLsSmemBytes = (size_t)kLsProducers * kStcStageBytes + 1024 /*align*/;      // dynamic part
static_assert(kLsSmemBytes + kLsMaxDeg * sizeof(LsEdge) + kLsPartBytes + kLsMaxDeg * 8 + 256 <= 227 * 1024,
              "k_linschur_tc shared memory (dynamic + static)");
// epilogue scratch aliases the operand stages
static_assert((size_t)128 * kLsSld * 4 + (size_t)kLsMaxDeg * (28 + 36) * 4 + (size_t)(kLsMaxDeg + 1) * 36 * 4 <=
              (size_t)kLsProducers * kStcStageBytes, "epilogue scratch");

// Slices are dealt to the producers in a fixed round of 14: producers 2 and 3 (warps 8, 9) sit on schedulers that
// hold no second producer and run ~1.4x faster than the other four, so they take three slices per round, the others
// two.  Fixed, so that the sums do not depend on timing.
constexpr int kLsRound = 14;
__constant__ int c_ls_order[kLsRound] = {2, 3, 0, 1, 2, 3, 4, 5, 2, 3, 0, 1, 4, 5};      // slice position -> producer
__constant__ int c_ls_rank[kLsRound] = {0, 0, 0, 0, 1, 1, 0, 0, 2, 2, 1, 1, 1, 1};       // how many earlier positions of the round it had
__constant__ int c_ls_count[kLsProducers] = {2, 2, 3, 3, 2, 2};                          // slices per round
__constant__ int c_ls_pos[kLsProducers][3] = {{2, 10, 0}, {3, 11, 0}, {0, 4, 8}, {1, 5, 9}, {6, 12, 0}, {7, 13, 0}};

constexpr int kLsBarOpFull = 0, kLsBarOpEmpty = 6, kLsBarAccFull = 12, kLsBarAccRead = 16;

// waiting roles (MMA issuer, drain warps) back off between polls: a spinning warp takes issue slots from the producers
static __device__ __forceinline__ void mbar_wait_sleep(uint32_t bar, uint32_t parity, unsigned ns) {
  uint32_t done = 0;
  for (;;) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(bar), "r"(parity)
        : "memory");
    if (done) break;
    __nanosleep(ns);
  }
}

// 16 consecutive TMEM columns of this thread's lane, twice, one wait
static __device__ __forceinline__ void tc_ld16x2(uint32_t ta, uint32_t tb, uint32_t* a, uint32_t* b) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%32];\n\t"
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%33];\n\t"
      "tcgen05.wait::ld.sync.aligned;"
      : "=r"(a[0]), "=r"(a[1]), "=r"(a[2]), "=r"(a[3]), "=r"(a[4]), "=r"(a[5]), "=r"(a[6]), "=r"(a[7]), "=r"(a[8]), "=r"(a[9]),
        "=r"(a[10]), "=r"(a[11]), "=r"(a[12]), "=r"(a[13]), "=r"(a[14]), "=r"(a[15]), "=r"(b[0]), "=r"(b[1]), "=r"(b[2]), "=r"(b[3]),
        "=r"(b[4]), "=r"(b[5]), "=r"(b[6]), "=r"(b[7]), "=r"(b[8]), "=r"(b[9]), "=r"(b[10]), "=r"(b[11]), "=r"(b[12]), "=r"(b[13]),
        "=r"(b[14]), "=r"(b[15])
      : "r"(ta), "r"(tb)
      : "memory");
}

// geometry of one pixel against one edge: everything the two passes share
struct LsPix {
  float d, xd, yd, gu, gv;   // 1/z (0 behind the near plane), x/z, y/z, fx/z, fy/z
  float Jzu, Jzv;
  bool near;
};
// Transformed points of a thread's four pixels.  Fast path: three FMAs per coordinate with the rotation matrix.  The
// near-plane test z < 0.25 must fall the same way as in the reference for every pixel (a pixel on the other side of
// the threshold changes its whole contribution), so whenever one of the four depths comes within 1e-3 of the
// threshold the four points are redone with the reference's own formula (actSE3, src/droid_kernels.cu:73-80, through
// `act`), whose z is bit-identical to the reference's.  Rt: R^T row-major, i.e. R[r][c] = Rt[3 c + r].
__device__ __forceinline__ void ls_points4(const float* Rt, const float* q, const float* t, const float* X0, const float* X1,
                                           const float* d0, float (*Y)[3]) {
  bool close = false;
#pragma unroll
  for (int s = 0; s < 4; ++s) {
    Y[s][0] = fmaf(Rt[0], X0[s], fmaf(Rt[3], X1[s], fmaf(t[0], d0[s], Rt[6])));
    Y[s][1] = fmaf(Rt[1], X0[s], fmaf(Rt[4], X1[s], fmaf(t[1], d0[s], Rt[7])));
    Y[s][2] = fmaf(Rt[2], X0[s], fmaf(Rt[5], X1[s], fmaf(t[2], d0[s], Rt[8])));
    close = close || fabsf(Y[s][2] - kMinDepth) < 1e-3f;
  }
  if (close) {
    Se3 G;
#pragma unroll
    for (int c = 0; c < 3; ++c) G.t[c] = t[c];
#pragma unroll
    for (int c = 0; c < 4; ++c) G.q[c] = q[c];
#pragma unroll
    for (int s = 0; s < 4; ++s) {
      const float Xi[4] = {X0[s], X1[s], 1.0f, d0[s]};
      float Xj[4];
      act(G, Xi, Xj);
      Y[s][0] = Xj[0]; Y[s][1] = Xj[1]; Y[s][2] = Xj[2];
    }
  }
}
__device__ __forceinline__ LsPix ls_project(const float* Yp, const float* t, float fx, float fy) {
  LsPix o;
  const float Y0 = Yp[0], Y1 = Yp[1], Y2 = Yp[2];
  o.near = Y2 < kMinDepth;
  o.d = o.near ? 0.0f : __fdividef(1.0f, Y2);
  o.xd = Y0 * o.d;
  o.yd = Y1 * o.d;
  o.gu = fx * o.d;
  o.gv = fy * o.d;
  o.Jzu = o.gu * fmaf(-t[2], o.xd, t[0]);
  o.Jzv = o.gv * fmaf(-t[2], o.yd, t[1]);
  return o;
}

// First half of the fused path: per-edge relative-pose records of the frame and the depth diagonal
//     C = sum_e w Jz^2 + (sensor-depth prior | eta),   q = 1 / C             (src/droid_kernels.cu:363-367, 1475-1478)
// from the weights alone (8 B per pixel-edge): q scales the operand tiles of k_linschur_tc, so it has to be complete
// before the first tile is written.  grid (HW / 512, K), 128 threads, four pixels per thread.
__global__ void __launch_bounds__(128)
k_depth_diag(const float* __restrict__ poses, const float* __restrict__ disps, const float* __restrict__ intr,
             const float* __restrict__ sens, const float* __restrict__ weights, const float* __restrict__ eta,
             int eta_rows, const int64_t* __restrict__ jj, int t0, int P, int HW, int wd, float alpha, int min_m,
             int k_lo, int k_hi, Ws ws) {
  const int k = k_lo + blockIdx.y;
  const int K = ws.info[0];
  if (k >= K || (k_hi >= 0 && k >= k_hi)) return;
  if (!ls_eligible(ws, k, min_m)) return;
  __shared__ LsEdge sx[kLsMaxDeg];
  __shared__ size_t eoff[kLsMaxDeg];
  const int tid = threadIdx.x;
  const int e_lo = ws.csr_ptr[k], deg = ws.csr_ptr[k + 1] - e_lo;
  const int ix = ws.kx[k];
  const float fx = __ldg(intr), fy = __ldg(intr + 1), cx = __ldg(intr + 2), cy = __ldg(intr + 3);
  if (tid < deg) {
    const int e = ws.csr_idx[e_lo + tid];
    const int jx = (int)jj[e];
    EdgeXf X;
    edge_xf(poses, ix, jx, X);
    LsEdge L;
#pragma unroll
    for (int q = 0; q < 9; ++q) { L.Rt[q] = X.Rt[q]; L.Nm[q] = X.Nm[q]; }
#pragma unroll
    for (int q = 0; q < 3; ++q) L.t[q] = X.t[q];
#pragma unroll
    for (int q = 0; q < 4; ++q) L.q[q] = X.q[q];
    L.jx = jx;
    L.flags = X.stereo;
    L.prow = (jx >= t0 && jx < t0 + P) ? jx - t0 : -1;
    sx[tid] = L;
    eoff[tid] = (size_t)e * 2 * HW;
    if (blockIdx.x == 0) {                 // k_linschur_tc reads the records; so does k_backsub_rc, after the poses have moved
      float4* dst = reinterpret_cast<float4*>(ws.exf + (size_t)e * kLsEdgeWords);
      const float4* src = reinterpret_cast<const float4*>(&L);
#pragma unroll
      for (int q = 0; q < kLsEdgeWords / 4; ++q) dst[q] = src[q];
    }
  }
  if (blockIdx.x == 0 && tid == 0) {
    ws.fused[k] = 1;
    ws.rc_ptrs[0] = weights;
    ws.rc_ptrs[1] = intr;
    if (eta_rows != 1 && eta_rows != K) ws.info[3] = 1;
  }
  __syncthreads();
  const int px0 = (blockIdx.x * 128 + tid) * 4;
  if (px0 >= HW) return;
  float X0[4], X1[4];
#pragma unroll
  for (int s = 0; s < 4; ++s) {
    const int pp = px0 + s;
    const int i = pp / wd, j = pp - i * wd;
    X0[s] = ((float)j - cx) / fx;
    X1[s] = ((float)i - cy) / fy;
  }
  const int er = eta_rows == 1 ? 0 : min(k, eta_rows - 1);
  const float4 d4 = __ldg(reinterpret_cast<const float4*>(disps + (size_t)ix * HW + px0));
  const float4 s4 = __ldg(reinterpret_cast<const float4*>(sens + (size_t)ix * HW + px0));
  const float4 e4 = __ldg(reinterpret_cast<const float4*>(eta + (size_t)er * HW + px0));
  const float d0[4] = {d4.x, d4.y, d4.z, d4.w};
  const float sd[4] = {s4.x, s4.y, s4.z, s4.w};
  const float et[4] = {e4.x, e4.y, e4.z, e4.w};
  float C[4] = {0.f, 0.f, 0.f, 0.f};
  auto wptr = [&](int a, int comp) {
    return reinterpret_cast<const float4*>(weights + eoff[min(a, deg - 1)] + (size_t)comp * HW + px0);
  };
  // two edges in flight ahead of the one being worked on
  float4 wu_a = __ldg(wptr(0, 0)), wv_a = __ldg(wptr(0, 1)), wu_b = __ldg(wptr(1, 0)), wv_b = __ldg(wptr(1, 1));
  for (int a = 0; a < deg; ++a) {
    const float4 wu4 = wu_a, wv4 = wv_a;
    wu_a = wu_b; wv_a = wv_b;
    if (a + 2 < deg) { wu_b = __ldg(wptr(a + 2, 0)); wv_b = __ldg(wptr(a + 2, 1)); }
    const LsEdge& X = sx[a];
    const float wu[4] = {wu4.x, wu4.y, wu4.z, wu4.w}, wv[4] = {wv4.x, wv4.y, wv4.z, wv4.w};
    float Y[4][3];
    ls_points4(X.Rt, X.q, X.t, X0, X1, d0, Y);
#pragma unroll
    for (int s = 0; s < 4; ++s) {
      const LsPix o = ls_project(Y[s], X.t, fx, fy);
      const float a_u = o.near ? 0.f : 0.001f * wu[s], a_v = o.near ? 0.f : 0.001f * wv[s];
      C[s] = fmaf(a_u * o.Jzu, o.Jzu, fmaf(a_v * o.Jzv, o.Jzv, C[s]));
    }
  }
  float q[4];
#pragma unroll
  for (int s = 0; s < 4; ++s) {
    const float m = sd[s] > 0.f ? 1.0f : 0.0f;
    q[s] = 1.0f / (C[s] + m * alpha + (1.0f - m) * et[s]);
  }
  *reinterpret_cast<float4*>(ws.Q + (size_t)k * HW + px0) = make_float4(q[0], q[1], q[2], q[3]);
}

__global__ void __launch_bounds__(kLsThreads, 1)
k_linschur_tc(const float* __restrict__ disps, const float* __restrict__ intr, const float* __restrict__ sens,
              const float* __restrict__ targets, const float* __restrict__ weights, int t0, int P, int HW, int wd,
              float alpha, int S, int min_m, int k_lo, int k_hi, Ws ws) {
  const int k = k_lo + blockIdx.y;
  const int K = ws.info[0];
  if (k >= K || (k_hi >= 0 && k >= k_hi)) return;
  if (!ls_eligible(ws, k, min_m)) return;                     // uniform for the CTA: nothing allocated yet
  const int s_lo = blockIdx.x * S, s_hi = min(HW, s_lo + S);  // HW and S are multiples of 32 (host)
  if (s_lo >= s_hi) return;
  const int e_lo = ws.csr_ptr[k], deg = ws.csr_ptr[k + 1] - e_lo;
  const int ix = ws.kx[k];
  const int self = ix - t0;
  const bool has_self = self >= 0 && self < P;

  extern __shared__ uint8_t ls_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(ls_raw) + 1023) & ~(uintptr_t)1023);
  // static shared arrays: the compiler sees the address space and emits LDS / STS (through a pointer derived from the
  // re-aligned dynamic buffer it emits generic loads, which share a scoreboard with the global prefetches)
  __shared__ __align__(16) LsEdge sx[kLsMaxDeg];
  __shared__ __align__(16) float part[kLsProducers * kLsMaxDeg * 32];       // [producer][edge][32]
  __shared__ uint64_t bars[24];
  __shared__ uint32_t tmem_slot[1];
  __shared__ float xtab[kLsTabMax], ytab[kLsTabMax];         // (j - cx) / fx and (i - cy) / fy, divided once per CTA like the reference does per pixel
  __shared__ size_t eoff[kLsMaxDeg];                          // element offset of the edge's [2][HW] block

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int n_chunks = (s_hi - s_lo) / 32;
  const float fx = __ldg(intr), fy = __ldg(intr + 1), cx = __ldg(intr + 2), cy = __ldg(intr + 3);

  // ---------------------------------------------------------------- prologue
  for (int q = tid; q < wd; q += kLsThreads) xtab[q] = ((float)q - cx) / fx;
  for (int q = tid; q < HW / wd; q += kLsThreads) ytab[q] = ((float)q - cy) / fy;
  if (tid < deg) {
    const int e = ws.csr_idx[e_lo + tid];
    const float4* src = reinterpret_cast<const float4*>(ws.exf + (size_t)e * kLsEdgeWords);   // written by k_depth_diag
    float4* dst = reinterpret_cast<float4*>(&sx[tid]);
#pragma unroll
    for (int q = 0; q < kLsEdgeWords / 4; ++q) dst[q] = src[q];
    eoff[tid] = (size_t)e * 2 * HW;
  }
  // operand rows that no producer writes (beyond 6 deg, except the w row) must read as zero
  for (int i = tid; i < kLsProducers * kStcStageBytes / 16; i += kLsThreads) reinterpret_cast<uint4*>(smem)[i] = make_uint4(0, 0, 0, 0);
  for (int i = tid; i < kLsProducers * kLsMaxDeg * 32; i += kLsThreads) part[i] = 0.f;
  if (tid == 0) {
    for (int i = 0; i < kLsProducers; ++i) {
      mbar_init(smem_u32(&bars[kLsBarOpFull + i]), 1);
      mbar_init(smem_u32(&bars[kLsBarOpEmpty + i]), 1);
    }
    for (int i = 0; i < 4; ++i) {
      mbar_init(smem_u32(&bars[kLsBarAccFull + i]), 1);
      mbar_init(smem_u32(&bars[kLsBarAccRead + i]), kLsDrainWarps);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 4) tc_alloc(smem_u32(tmem_slot), 512);
  fence_async_smem();                                         // the zeroes above are read by the MMA unit
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  float* Ssm = reinterpret_cast<float*>(smem);                 // epilogue scratch over the operand stages: S [128][kLsSld]
  float* tot = Ssm + 128 * kLsSld;                             // [deg][28]
  float* Tm = tot + kLsMaxDeg * 28;                            // [deg][36]
  float* Tf = Tm + kLsMaxDeg * 36;                             // [deg + 1][36]: sum_e M_e S[e][f], last: the w column
  if (warp >= kLsFirstProducer) {
    // ================================================================ producers
    const int p = warp - kLsFirstProducer;
    const int g = lane >> 3, l8 = lane & 7;
    const uint32_t stage = smem_u32(smem) + (uint32_t)p * kStcStageBytes;   // shared-window address of the hi tile
    float* mypart = &part[p * (kLsMaxDeg * 32)];
    const int n_it = (deg + 3) >> 2;
    // byte offset of this thread's 16-byte chunk inside operand row r: r * 128 + ((l8 ^ (r & 7)) << 4)
    auto row_off = [&](int r) -> uint32_t { return (uint32_t)r * 128u + (((uint32_t)l8 ^ ((uint32_t)r & 7u)) << 4); };
    // x -> (hi, lo): hi = x rounded to TF32 (round to nearest, ties away from zero, as cvt.rna.tf32.f32 does; done on
    // the bit pattern: two integer instructions instead of the conversion's five), lo = x - hi exactly
    auto store_split_at = [&](uint32_t dst, const float* v) {
      float hi[4];
#pragma unroll
      for (int s = 0; s < 4; ++s) hi[s] = __uint_as_float((__float_as_uint(v[s]) + 0x1000u) & 0xffffe000u);
      asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(dst), "f"(hi[0]), "f"(hi[1]), "f"(hi[2]), "f"(hi[3]) : "memory");
      asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(dst + (uint32_t)kStcTile), "f"(v[0] - hi[0]), "f"(v[1] - hi[1]),
                   "f"(v[2] - hi[2]), "f"(v[3] - hi[3]) : "memory");
    };
    // Loads are software-pipelined across every boundary: the next edge of the group is requested while the current
    // one is worked on, the pixel rows and first edge of the warp's next slice before this slice's tail.  eoff[a] =
    // element offset of edge a's [2][HW] block in targets / weights.
    const float* wbase = weights + s_lo + 4 * l8;
    const float* tbase = targets + s_lo + 4 * l8;
    auto ld4 = [](const float* q) { return __ldg(reinterpret_cast<const float4*>(q)); };
    auto edge_off = [&](int a) -> size_t { return eoff[min(a, deg - 1)]; };
    float4 d4, s4, q4, wu_n, wv_n, tu_n, tv_n;
    auto slice_loads = [&](int c) {
      const int px0 = s_lo + 32 * c + 4 * l8;
      d4 = ld4(disps + (size_t)ix * HW + px0);
      s4 = ld4(sens + (size_t)ix * HW + px0);
      q4 = *reinterpret_cast<const float4*>(ws.Q + (size_t)k * HW + px0);     // written by k_depth_diag
      const size_t o = edge_off(g) + 32 * c;
      wu_n = ld4(wbase + o);
      wv_n = ld4(wbase + o + HW);
      tu_n = ld4(tbase + o);
      tv_n = ld4(tbase + o + HW);
    };
    const int my_count = c_ls_count[p];
    const int my_pos[3] = {c_ls_pos[p][0], c_ls_pos[p][1], c_ls_pos[p][2]};
    auto my_slice = [&](int use) -> int {                       // this producer's use-th slice
      const int r = use / my_count, k = use - r * my_count;
      return kLsRound * r + (k == 0 ? my_pos[0] : k == 1 ? my_pos[1] : my_pos[2]);
    };
    if (my_slice(0) < n_chunks) slice_loads(my_slice(0));
    int use = 0;
    for (int c = my_slice(0); c < n_chunks; c = my_slice(++use)) {
      const int px0 = s_lo + 32 * c + 4 * l8;
      const int pc = 32 * c;
      float X0[4], X1[4];
#pragma unroll
      for (int s = 0; s < 4; ++s) {
        const int pp = px0 + s;
        const int i = pp / wd, j = pp - i * wd;
        X0[s] = xtab[j];
        X1[s] = ytab[i];
      }
      const float d0[4] = {d4.x, d4.y, d4.z, d4.w};
      const float sd[4] = {s4.x, s4.y, s4.z, s4.w};
      const float sq[4] = {sqrtf(q4.x), sqrtf(q4.y), sqrtf(q4.z), sqrtf(q4.w)};

      float W[4] = {0.f, 0.f, 0.f, 0.f};
      const int c_next = my_slice(use + 1);
      for (int it = 0; it < n_it; ++it) {
        const int a = 4 * it + g;
        const float4 wu4 = wu_n, wv4 = wv_n, tu4 = tu_n, tv4 = tv_n;
        if (it + 1 < n_it) {
          const size_t o = edge_off(a + 4) + pc;
          wu_n = ld4(wbase + o); wv_n = ld4(wbase + o + HW);
          tu_n = ld4(tbase + o); tv_n = ld4(tbase + o + HW);
        } else if (c_next < n_chunks) {
          slice_loads(c_next);                                 // pixel rows and first edge of this warp's next slice
        }
        float h[32];
#pragma unroll
        for (int l = 0; l < 32; ++l) h[l] = 0.f;
        if (a < deg) {
          const LsEdge& X = sx[a];
          const bool stereo = (X.flags & 1) != 0;
          const float wu[4] = {wu4.x, wu4.y, wu4.z, wu4.w}, wv[4] = {wv4.x, wv4.y, wv4.z, wv4.w};
          const float tu[4] = {tu4.x, tu4.y, tu4.z, tu4.w}, tv[4] = {tv4.x, tv4.y, tv4.z, tv4.w};
          float Es[6][4];
          float Y[4][3];
          ls_points4(X.Rt, X.q, X.t, X0, X1, d0, Y);
#pragma unroll
          for (int s = 0; s < 4; ++s) {
            const LsPix o = ls_project(Y[s], X.t, fx, fy);
            float a_u = o.near ? 0.f : 0.001f * wu[s], a_v = o.near ? 0.f : 0.001f * wv[s];
            const float bu = fx * o.xd, bv = fy * o.yd;
            const float ru = tu[s] - (bu + cx), rv = tv[s] - (bv + cy);
            W[s] = fmaf(a_u * o.Jzu, ru, fmaf(a_v * o.Jzv, rv, W[s]));
            if (stereo) { a_u = 0.f; a_v = 0.f; }              // stereo edges feed the depth blocks only
            // Ju = fx (h d, 0, -x h d^2, -x y d^2, 1 + x^2 d^2, -y d),  Jv = fy (0, h d, -y h d^2, -1 - y^2 d^2, x y d^2, x d)
            // (src/droid_kernels.cu:356-361, 389-394) with x d, y d, h d formed once
            const float Ju0 = o.gu * d0[s], Jv1 = o.gv * d0[s];
            const float Ju[6] = {Ju0, 0.f, -Ju0 * o.xd, -bu * o.yd, fmaf(bu, o.xd, fx), -fx * o.yd};
            const float Jv[6] = {0.f, Jv1, -Jv1 * o.yd, -fmaf(bv, o.yd, fy), bv * o.xd, fy * o.xd};
            float wJu[6], wJv[6];
#pragma unroll
            for (int n = 0; n < 6; ++n) { wJu[n] = a_u * Ju[n]; wJv[n] = a_v * Jv[n]; }
            int l = 0;
#pragma unroll
            for (int n = 0; n < 6; ++n)
#pragma unroll
              for (int m2 = 0; m2 <= n; ++m2) {
                if (n != 1 && m2 != 1) h[l] = fmaf(wJu[n], Ju[m2], h[l]);
                if (n != 0 && m2 != 0) h[l] = fmaf(wJv[n], Jv[m2], h[l]);
                ++l;
              }
#pragma unroll
            for (int n = 0; n < 6; ++n) {
              if (n != 1) h[21 + n] = fmaf(wJu[n], ru, h[21 + n]);
              if (n != 0) h[21 + n] = fmaf(wJv[n], rv, h[21 + n]);
            }
            const float zu = o.Jzu * sq[s], zv = o.Jzv * sq[s];   // E sqrt(q) = (w Jz sqrt(q)) J
#pragma unroll
            for (int n = 0; n < 6; ++n)
              Es[n][s] = n == 0 ? zu * wJu[0] : n == 1 ? zv * wJv[1] : fmaf(zu, wJu[n], zv * wJv[n]);
          }
          // the MMAs that read this stage one round ago have retired (asked only now: the first edge's arithmetic
          // overlaps the wait; every group has an edge in round 0, frames here have at least seven)
          if (it == 0) mbar_wait(smem_u32(&bars[kLsBarOpEmpty + p]), (use + 1) & 1);
#pragma unroll
          for (int n = 0; n < 6; ++n) store_split_at(stage + row_off(6 * a + n), Es[n]);
        }
        // sums over the eight lanes of the group: lane l8 ends up with terms 4 l8 .. 4 l8 + 3
#pragma unroll
        for (int half = 4, cnt = 16; half >= 1; half >>= 1, cnt >>= 1) {
          const bool up = (l8 & half) != 0;
#pragma unroll
          for (int i = 0; i < cnt; ++i) {
            const float send = up ? h[i] : h[i + cnt];
            const float keep = up ? h[i + cnt] : h[i];
            h[i] = keep + __shfl_xor_sync(0xffffffffu, send, half);
          }
        }
        if (a < deg) {
          float4* acc = reinterpret_cast<float4*>(mypart + a * 32 + 4 * l8);
          float4 v = *acc;
          v.x += h[0]; v.y += h[1]; v.z += h[2]; v.w += h[3];
          *acc = v;
        }
      }
      // w of the frame: sum over the groups, sensor-depth prior, then its operand row and the global copy
#pragma unroll
      for (int s = 0; s < 4; ++s) {
        W[s] += __shfl_xor_sync(0xffffffffu, W[s], 8);
        W[s] += __shfl_xor_sync(0xffffffffu, W[s], 16);
        if (sd[s] > 0.f) W[s] -= alpha * (d0[s] - sd[s]);
      }
      if (g == 0) {
        const float wsq[4] = {W[0] * sq[0], W[1] * sq[1], W[2] * sq[2], W[3] * sq[3]};
        store_split_at(stage + row_off(kLsWRow), wsq);
        *reinterpret_cast<float4*>(ws.w + (size_t)k * HW + px0) = make_float4(W[0], W[1], W[2], W[3]);
      }
      fence_async_smem();                                      // generic-proxy stores -> visible to the MMA unit
      __syncwarp();
      if (lane == 0) mbar_arrive(smem_u32(&bars[kLsBarOpFull + p]));
    }
  } else if (warp == 4) {
    if (lane == 0) {
      // ============================================================== MMA issuer
      const uint64_t desc0 = umma_desc_k128(smem_u32(smem));
      for (int c = 0; c < n_chunks; ++c) {
        const int rr = c / kLsRound, pos = c - rr * kLsRound;
        const int st = c_ls_order[pos], acc_i = c & 3;
        const int fill = rr * c_ls_count[st] + c_ls_rank[pos];        // how often stage st has been filled before
        if (c >= 4) mbar_wait_sleep(smem_u32(&bars[kLsBarAccRead + acc_i]), ((c - 4) >> 2) & 1, 200);
        mbar_wait_sleep(smem_u32(&bars[kLsBarOpFull + st]), fill & 1, 200);
        tc_fence_after();
        const uint64_t dhi = desc0 + (uint64_t)((st * kStcStageBytes) >> 4);
        const uint64_t dlo = dhi + (uint64_t)(kStcTile >> 4);
        const uint32_t acc = tmem_base + (uint32_t)(acc_i * 128);
#pragma unroll
        for (int k8 = 0; k8 < 4; ++k8) {
          tc_mma_tf32(acc, dlo + 2 * k8, dhi + 2 * k8, kStcInstr, k8 ? 1u : 0u);
          tc_mma_tf32(acc, dhi + 2 * k8, dlo + 2 * k8, kStcInstr, 1u);
        }
#pragma unroll
        for (int k8 = 0; k8 < 4; ++k8) tc_mma_tf32(acc, dhi + 2 * k8, dhi + 2 * k8, kStcInstr, 1u);
        tc_commit(smem_u32(&bars[kLsBarOpEmpty + st]));
        tc_commit(smem_u32(&bars[kLsBarAccFull + acc_i]));
      }
    }
  } else if (warp < kLsDrainWarps) {
    // ================================================================ drain: fresh accumulator per slice, summed
    // here with round-to-nearest operations (see ba_schur_tc.cuh for why, and for the factor 1 + 2^-23)
    float run[128];                                           // row `tid` of S, all 128 columns
#pragma unroll
    for (int i = 0; i < 128; ++i) run[i] = 0.f;
    for (int c = 0; c < n_chunks; ++c) {
      mbar_wait_sleep(smem_u32(&bars[kLsBarAccFull + (c & 3)]), (c >> 2) & 1, 1000);
      tc_fence_after();
      const uint32_t t = tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)((c & 3) * 128);
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        uint32_t a0[16], a1[16];
        tc_ld16x2(t + 32 * q, t + 32 * q + 16, a0, a1);
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          run[32 * q + i] = fmaf(__uint_as_float(a0[i]), 1.00000011920928955f, run[32 * q + i]);
          run[32 * q + 16 + i] = fmaf(__uint_as_float(a1[i]), 1.00000011920928955f, run[32 * q + 16 + i]);
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(smem_u32(&bars[kLsBarAccRead + (c & 3)]));
    }
    // every MMA has retired (this warp has read the last accumulator) and the producers wrote their last slice
    // before that: the operand stages are free, S goes where they were
#pragma unroll
    for (int b = 0; b < 128; ++b) Ssm[tid * kLsSld + b] = run[b];
  }

  // ------------------------------------------------------------------ epilogue
  tc_fence_before();
  __syncthreads();
  for (int o = tid; o < deg * 27; o += kLsThreads) {
    const int a = o / 27, l = o - a * 27;
    float sum = 0.f;
#pragma unroll
    for (int q = 0; q < kLsProducers; ++q) sum += part[(q * kLsMaxDeg + a) * 32 + l];
    tot[a * 28 + l] = sum;
  }
  __syncthreads();
  auto Hjj = [&](int a, int r, int c) -> float {
    const int hi = r > c ? r : c, lo = r > c ? c : r;
    return tot[a * 28 + hi * (hi + 1) / 2 + lo];
  };
  auto Mm = [&](int a, int r, int c) -> float {               // M = [[Rt, 0], [Nm, Rt]]
    const LsEdge& X = sx[a];
    if (r < 3) return c < 3 ? X.Rt[r * 3 + c] : 0.f;
    return c < 3 ? X.Nm[(r - 3) * 3 + c] : X.Rt[(r - 3) * 3 + (c - 3)];
  };
  for (int o = tid; o < deg * 36; o += kLsThreads) {
    const int a = o / 36, rc = o - a * 36, r = rc / 6, c = rc - r * 6;
    float val = 0.f;
#pragma unroll
    for (int kk = 0; kk < 6; ++kk) val += Mm(a, r, kk) * Hjj(a, kk, c);
    Tm[a * 36 + rc] = val;
  }
  // Tf[f] = sum_e M_e S[e][f] (f = deg: the w column, stored as a 6-vector in the first column)
  if (has_self) {
    for (int o = tid; o < (deg + 1) * 36; o += kLsThreads) {
      const int f = o / 36, rc = o - f * 36, r = rc / 6, c = rc - r * 6;
      if (f == deg && c != 0) continue;
      const int col = f == deg ? kLsWRow : 6 * f + c;
      float val = 0.f;
      for (int e = 0; e < deg; ++e) {
#pragma unroll
        for (int kk = 0; kk < 6; ++kk) val += Mm(e, r, kk) * Ssm[(6 * e + kk) * kLsSld + col];
      }
      Tf[f * 36 + rc] = val;
    }
  }
  __syncthreads();
  const int n6 = 6 * P;
  // pose blocks of the edges: [0,36) Hii = T M^T, [36,72) Hij = -T, [72,108) Hji = -T^T, [108,144) Hjj,
  //                           [144,150) vi = -M vj, [150,156) vj          (as in k_linearize2)
  for (int o = tid; o < deg * 156; o += kLsThreads) {
    const int a = o / 156, q = o - a * 156;
    const int pi = self, pj = sx[a].jx - t0;
    const bool iok = has_self, jok = pj >= 0 && pj < P;
    float val = 0.f;
    if (q < 144) {
      const int blk = q / 36, rc = q - blk * 36, r = rc / 6, c = rc - r * 6;
      if (blk == 3) val = Hjj(a, r, c);
      else if (blk == 1) val = -Tm[a * 36 + r * 6 + c];
      else if (blk == 2) val = -Tm[a * 36 + c * 6 + r];
      else {
#pragma unroll
        for (int ll = 0; ll < 6; ++ll) val += Tm[a * 36 + r * 6 + ll] * Mm(a, c, ll);
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
        for (int kk = 0; kk < 6; ++kk) val -= Mm(a, r, kk) * tot[a * 28 + 21 + kk];
        if (iok) atomicAdd(ws.v + 6 * pi + r, (double)val);
      } else {
        if (jok) atomicAdd(ws.v + 6 * pj + r, (double)tot[a * 28 + 21 + r]);
      }
    }
  }
  // Schur complement, edge rows: H[pose_a][pose_b] -= S[a][b] (lower triangle to both sides), v[pose_a] -= S[a][w]
  const int nr = 6 * deg;
  for (int o = tid; o < nr * (nr + 1); o += kLsThreads) {
    const int a = o / (nr + 1), b = o - a * (nr + 1);
    const int pa = sx[a / 6].prow;
    if (pa < 0) continue;
    const int ra6 = 6 * pa + a % 6;
    if (b == nr) {
      atomicAdd(ws.v + ra6, -(double)Ssm[a * kLsSld + kLsWRow]);
    } else if (b <= a) {
      const int pb = sx[b / 6].prow;
      if (pb < 0) continue;
      const int cb6 = 6 * pb + b % 6;
      const double val = -(double)Ssm[a * kLsSld + b];
      atomicAdd(ws.H + (size_t)ra6 * n6 + cb6, val);
      if (b < a) atomicAdd(ws.H + (size_t)cb6 * n6 + ra6, val);
    }
  }
  if (has_self) {
    // self rows: S[self][f] = -Tf[f], S[self][self] = sum_f Tf[f] M_f^T, sv[self] = -Tf[deg]
    for (int o = tid; o < deg * 36; o += kLsThreads) {
      const int f = o / 36, rc = o - f * 36, r = rc / 6, c = rc - r * 6;
      const int pf = sx[f].prow;
      if (pf < 0) continue;
      const double val = (double)Tf[f * 36 + rc];               // -(-Tf)
      atomicAdd(ws.H + (size_t)(6 * self + r) * n6 + 6 * pf + c, val);
      atomicAdd(ws.H + (size_t)(6 * pf + c) * n6 + 6 * self + r, val);
    }
    for (int o = tid; o < 36; o += kLsThreads) {
      const int r = o / 6, c = o - r * 6;
      float val = 0.f;
      for (int f = 0; f < deg; ++f) {
#pragma unroll
        for (int kk = 0; kk < 6; ++kk) val += Tf[f * 36 + r * 6 + kk] * Mm(f, c, kk);
      }
      atomicAdd(ws.H + (size_t)(6 * self + r) * n6 + 6 * self + c, -(double)val);
    }
    if (tid < 6) atomicAdd(ws.v + 6 * self + tid, (double)Tf[deg * 36 + tid * 6]);
  }
  __syncthreads();
  if (warp == 4) {
    tc_fence_after();
    tc_dealloc(tmem_base, 512);
  }
}

// ------------------------------------------------------------------------------------------------------------
// Depth back-substitution of the frames k_linschur_tc took: dz = q (w - sum_e Ej_e . g_e) with the per-pixel Ej
// recomputed from the weights and the edge records, and g_e = dx[pose j] - M_e^T dx[pose i] (the self row
// Ei = -sum_e M_e Ej_e folded in).  Rows whose pose index is 0 or outside the window are skipped like in the
// reference (src/droid_kernels.cu:1150).  The pose update of the iteration stays in k_backsub_fused, which
// skips these frames.  grid (HW / 512, K), 128 threads, four pixels per thread.
struct RcEdge { float q[4]; float t[3]; int live; float Rt[9]; float g[6]; int eid; };
__global__ void __launch_bounds__(128)
k_backsub_rc(const float* __restrict__ dx, float* __restrict__ disps, float* __restrict__ dz_out, int t0, int P, int HW,
             int wd, int k_lo, int k_hi, Ws ws) {
  const int k = k_lo + blockIdx.y;
  const int K = ws.info[0];
  if (k >= K || (k_hi >= 0 && k >= k_hi)) return;
  if (ws.fused[k] == 0) return;
  __shared__ RcEdge se[kLsMaxDeg];
  const int tid = threadIdx.x;
  const int e_lo = ws.csr_ptr[k], deg = ws.csr_ptr[k + 1] - e_lo;
  const int f = ws.kx[k];
  const int self = f - t0;
  const float* __restrict__ weights = ws.rc_ptrs[0];
  const float* __restrict__ intr = ws.rc_ptrs[1];
  const float fx = __ldg(intr), fy = __ldg(intr + 1), cx = __ldg(intr + 2), cy = __ldg(intr + 3);
  if (tid < deg) {
    const int e = ws.csr_idx[e_lo + tid];
    const LsEdge* L = reinterpret_cast<const LsEdge*>(ws.exf + (size_t)e * kLsEdgeWords);
    RcEdge R;
#pragma unroll
    for (int q = 0; q < 4; ++q) R.q[q] = L->q[q];
#pragma unroll
    for (int q = 0; q < 3; ++q) R.t[q] = L->t[q];
#pragma unroll
    for (int q = 0; q < 9; ++q) R.Rt[q] = L->Rt[q];
    const int pj = L->prow;
    float xi[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f}, g[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    if (pj > 0 && pj < P) {
#pragma unroll
      for (int n = 0; n < 6; ++n) g[n] = __ldg(dx + 6 * pj + n);
    }
    if (self > 0 && self < P) {
#pragma unroll
      for (int n = 0; n < 6; ++n) xi[n] = __ldg(dx + 6 * self + n);
      // M^T x = (Rt^T x[0:3] + Nm^T x[3:6], Rt^T x[3:6])
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        g[c] -= L->Rt[c] * xi[0] + L->Rt[3 + c] * xi[1] + L->Rt[6 + c] * xi[2] +
                L->Nm[c] * xi[3] + L->Nm[3 + c] * xi[4] + L->Nm[6 + c] * xi[5];
        g[3 + c] -= L->Rt[c] * xi[3] + L->Rt[3 + c] * xi[4] + L->Rt[6 + c] * xi[5];
      }
    }
#pragma unroll
    for (int n = 0; n < 6; ++n) R.g[n] = g[n];
    R.eid = e;
    R.live = (L->flags & 1) ? 0 : 1;                            // stereo edges have no pose blocks
    se[tid] = R;
  }
  __syncthreads();
  const int px0 = (blockIdx.x * 128 + tid) * 4;
  if (px0 >= HW) return;                                        // HW is a multiple of 32
  float X0[4], X1[4];
#pragma unroll
  for (int s = 0; s < 4; ++s) {
    const int pp = px0 + s;
    const int i = pp / wd, j = pp - i * wd;
    X0[s] = ((float)j - cx) / fx;
    X1[s] = ((float)i - cy) / fy;
  }
  const float4 d4 = *reinterpret_cast<const float4*>(disps + (size_t)f * HW + px0);
  const float d0[4] = {d4.x, d4.y, d4.z, d4.w};
  float acc[4] = {0.f, 0.f, 0.f, 0.f};
  auto wptr = [&](int a, int comp) {
    return reinterpret_cast<const float4*>(weights + ((size_t)se[min(a, deg - 1)].eid * 2 + comp) * HW + px0);
  };
  float4 wu_n = __ldg(wptr(0, 0)), wv_n = __ldg(wptr(0, 1));
  for (int a = 0; a < deg; ++a) {
    const float4 wu4 = wu_n, wv4 = wv_n;
    if (a + 1 < deg) { wu_n = __ldg(wptr(a + 1, 0)); wv_n = __ldg(wptr(a + 1, 1)); }
    const RcEdge& X = se[a];
    if (!X.live) continue;
    const float wu[4] = {wu4.x, wu4.y, wu4.z, wu4.w}, wv[4] = {wv4.x, wv4.y, wv4.z, wv4.w};
    float Y[4][3];
    ls_points4(X.Rt, X.q, X.t, X0, X1, d0, Y);
#pragma unroll
    for (int s = 0; s < 4; ++s) {
      const LsPix o = ls_project(Y[s], X.t, fx, fy);
      const float a_u = o.near ? 0.f : 0.001f * wu[s], a_v = o.near ? 0.f : 0.001f * wv[s];
      const float bu = fx * o.xd, bv = fy * o.yd;
      const float Ju0 = o.gu * d0[s], Jv1 = o.gv * d0[s];
      const float ju = Ju0 * X.g[0] - Ju0 * o.xd * X.g[2] - bu * o.yd * X.g[3] + fmaf(bu, o.xd, fx) * X.g[4] - fx * o.yd * X.g[5];
      const float jv = Jv1 * X.g[1] - Jv1 * o.yd * X.g[2] - fmaf(bv, o.yd, fy) * X.g[3] + bv * o.xd * X.g[4] + fy * o.xd * X.g[5];
      acc[s] = fmaf(a_u * o.Jzu, ju, fmaf(a_v * o.Jzv, jv, acc[s]));
    }
  }
  if (ws.info[3] != 0) return;                                   // eta row mismatch: nothing is written (see k_backsub_fused)
  const float4 q4 = *reinterpret_cast<const float4*>(ws.Q + (size_t)k * HW + px0);
  const float4 w4 = *reinterpret_cast<const float4*>(ws.w + (size_t)k * HW + px0);
  const float4 dz = make_float4(q4.x * (w4.x - acc[0]), q4.y * (w4.y - acc[1]), q4.z * (w4.z - acc[2]), q4.w * (w4.w - acc[3]));
  if (dz_out) *reinterpret_cast<float4*>(dz_out + (size_t)k * HW + px0) = dz;
  if (disps) {
    *reinterpret_cast<float4*>(disps + (size_t)f * HW + px0) = make_float4(d0[0] + dz.x, d0[1] + dz.y, d0[2] + dz.z, d0[3] + dz.w);
  }
}
