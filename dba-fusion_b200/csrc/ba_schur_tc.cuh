// Schur complement on the tensor cores, included by ba.cu inside namespace dbaf::{anonymous}.
//
// For a depth frame k with m row blocks (6m scalar rows e_a of E that touch its depths) the Schur term is
//     S[a][b] = sum_px (e_a[px] q[px]) e_b[px],      sv[a] = sum_px (e_a[px] q[px]) w[px]
// (computed as X X^T with X = E sqrt(q), w sqrt(q) riding along as one more row: q = 1/C > 0)
// i.e. one (6m x HW) x (HW x 6m) product per frame: GEMM-shaped, K = HW.  The reference forms it block by
// block (EEt6x6_kernel / Ev6x1_kernel, src/droid_kernels.cu:1046-1138) in fp32.  Here one CTA owns a frame
// (x a pixel range) and runs it as a 128 x 128 x K TF32 product with fp32 accumulation in TMEM:
//   * rows of E are contiguous in pixels, so both operands are K-major as they lie in HBM;
//   * w rides along as row 126 of the B operand, so sv falls out as column 126 of the same product;
//   * fp32 accuracy is kept with the usual three-product split  x = hi + lo  (hi = x rounded to TF32,
//     lo = x - hi exactly):  X X^T ~ Xhi Xhi^T + Xhi Xlo^T + Xlo Xhi^T,  error ~2^-22 per product (tests compare
//     against the fp32 SIMT kernel and the reference to 1e-4 normwise like every other path).
// One warp streams 32-pixel slices of the 128 rows from HBM with 16-byte asynchronous copies (four slices =
// 64 KB in flight per SM); eight transform warps scale them by sqrt(q), split them and store the two operand tiles
// (Xhi, Xlo; 16 KB each, 128-byte swizzled; both are A and B operand) of a 4-stage ring; one thread issues
// 12 tcgen05.mma.kind::tf32 (M = N = 128, K = 8) per slice into one of four TMEM accumulators.  The tensor core
// accumulates with truncation, so each slice starts a fresh accumulator and the producer warps add the
// finished slices into fp32 registers (round to nearest) while the next loads are in flight.  At the end
// the sums are subtracted from the dense fp64 H, v with atomics (same addresses as k_schur_smem).
// Frames with fewer than `min_m` or more than 21 row blocks are left to the SIMT kernels.
#pragma once   // needs tc_ptx.cuh, included by ba.cu at file scope

constexpr int kStcOpStages = 4;             // operand stages = TMEM accumulators (4 x 128 columns): slice c uses slot c & 3
constexpr int kStcRawStages = 4;
constexpr int kStcTile = 128 * 128;          // bytes of one tile: 128 rows x 32 floats
constexpr int kStcStageBytes = 2 * kStcTile; // Xhi, Xlo of X = E sqrt(q): both operands of the product
constexpr int kStcXformWarps = 8;
constexpr int kStcThreads = (kStcXformWarps + 2) * 32;   // + MMA issuer warp + loader warp
constexpr size_t kStcSmemBytes = (size_t)kStcOpStages * kStcStageBytes + (size_t)kStcRawStages * kStcTile +
                                 1024 /*align*/ + 128 * 8 /*row sources*/ + 32 * 4 /*row poses*/ + 256 /*barriers*/;
// kind::tf32: D = F32, A = B = TF32, both K-major, M = N = 128
constexpr uint32_t kStcInstr = (1u << 4) | (2u << 7) | (2u << 10) | ((128u >> 3) << 17) | ((128u >> 4) << 24);

__device__ __forceinline__ float tf32_rna(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return __uint_as_float(r);
}
// barrier slots
constexpr int kBarRawFull = 0, kBarRawEmpty = 4, kBarOpFull = 8, kBarOpEmpty = 12, kBarAccFull = 16, kBarAccRead = 20;

__global__ void __launch_bounds__(kStcThreads, 1)
k_schur_tc(const int64_t* __restrict__ jj, int t0, int P, int HW, int S, int min_m, int k_lo, int k_hi, Ws ws) {
  const int k = k_lo + blockIdx.y;
  const int K = ws.info[0];
  if (k >= K || (k_hi >= 0 && k >= k_hi)) return;
  const int r_lo = ws.srow_ptr[k];
  const int m = ws.srow_ptr[k + 1] - r_lo;
  if (m < min_m || m > kStcMaxM) return;                 // uniform for the CTA: nothing allocated yet
  const int s_lo = blockIdx.x * S, s_hi = min(HW, s_lo + S);   // HW and S are multiples of 32 (host)
  if (s_lo >= s_hi) return;

  extern __shared__ uint8_t stc_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(stc_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* raw = smem + kStcOpStages * kStcStageBytes;                    // [4][128 rows][32 floats], linear
  const float** rowsrc = reinterpret_cast<const float**>(raw + kStcRawStages * kStcTile);   // [128]
  int* rowp = reinterpret_cast<int*>(rowsrc + 128);                        // [32]
  uint64_t* bars = reinterpret_cast<uint64_t*>(rowp + 32);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 24);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int rows6 = 6 * m;
  const int n_chunks = (s_hi - s_lo) / 32;

  if (tid < 128) {
    // scalar row r of the operands: component r % 6 of row block r / 6; row 126 = w; row 127 carries q
    const float* src = nullptr;
    if (tid < rows6) {
      const int n = ws.srow_idx[r_lo + tid / 6];
      src = (n < P ? ws.Ei + (size_t)n * 6 * HW : ws.Eij + (size_t)(n - P) * 6 * HW) + (size_t)(tid % 6) * HW;
      if (tid % 6 == 0) rowp[tid / 6] = n < P ? n : (int)jj[n - P] - t0;
    } else if (tid == 126) {
      src = ws.w + (size_t)k * HW;
    } else if (tid == 127) {
      src = ws.Q + (size_t)k * HW;
    }
    rowsrc[tid] = src;
  }
  // rows without a source are never written by the copy engine: zero them once in every raw stage
  for (int i = tid; i < kStcRawStages * kStcTile / 16; i += kStcThreads) reinterpret_cast<uint4*>(raw)[i] = make_uint4(0, 0, 0, 0);
  if (tid == 0) {
    for (int i = 0; i < kStcRawStages; ++i) {
      mbar_init(smem_u32(&bars[kBarRawFull + i]), 32);     // every lane of the loader warp
      mbar_init(smem_u32(&bars[kBarRawEmpty + i]), kStcXformWarps);
    }
    for (int i = 0; i < kStcOpStages; ++i) {
      mbar_init(smem_u32(&bars[kBarOpFull + i]), kStcXformWarps);
      mbar_init(smem_u32(&bars[kBarOpEmpty + i]), 1);
      mbar_init(smem_u32(&bars[kBarAccFull + i]), 1);
      mbar_init(smem_u32(&bars[kBarAccRead + i]), kStcXformWarps);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == kStcXformWarps) tc_alloc(smem_u32(tmem_slot), 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == kStcXformWarps + 1) {
    // ------------------------------------------------------------ loader warp: HBM -> raw fp32 slices
    // 16-byte asynchronous copies (LDGSTS): one warp instruction moves 4 rows x 128 B, 32 instructions a
    // slice, and up to four slices (64 KB) are in flight per SM without holding registers
    const int piece = lane & 7, rsub = lane >> 3;
    // this lane's 32 row sources live in registers: a shared-memory lookup per copy would queue behind the
    // copies already in the load/store pipe
    const float* rsrc[32];
    uint32_t valid = 0;
#pragma unroll
    for (int j = 0; j < 32; ++j) {
      rsrc[j] = rowsrc[4 * j + rsub];
      if (rsrc[j] != nullptr) valid |= 1u << j;
      else rsrc[j] = ws.Q;                          // never dereferenced
      rsrc[j] += s_lo + piece * 4;
    }
    for (int c = 0; c < n_chunks; ++c) {
      const int rs = c % kStcRawStages, use = c / kStcRawStages;
      mbar_wait(smem_u32(&bars[kBarRawEmpty + rs]), (use + 1) & 1);
      const uint32_t dst = smem_u32(raw + rs * kStcTile) + (uint32_t)piece * 16u;
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        if ((valid >> j) & 1u)
          asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + (uint32_t)(4 * j + rsub) * 128u), "l"(rsrc[j] + c * 32)
                       : "memory");
      }
      // arrives on the slice's barrier once all of this thread's copies above have landed
      asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(smem_u32(&bars[kBarRawFull + rs])) : "memory");
    }
  } else if (warp == kStcXformWarps) {
    if (lane == 0) {
      // ---------------------------------------------------------- MMA issuer
      const uint64_t desc0 = umma_desc_k128(smem_u32(smem));
      for (int c = 0; c < n_chunks; ++c) {
        const int st = c & 3;
        if (c >= 4) mbar_wait(smem_u32(&bars[kBarAccRead + st]), ((c - 4) >> 2) & 1);   // accumulator st has been read
        mbar_wait(smem_u32(&bars[kBarOpFull + st]), (c >> 2) & 1);
        tc_fence_after();
        // shared-memory descriptors are linear in the address (start field = address >> 4): one is built before the
        // loop, the other 31 are integer adds, so that the issuing thread keeps the tensor pipe fed
        const uint64_t dhi = desc0 + (uint64_t)((st * kStcStageBytes) >> 4);
        const uint64_t dlo = dhi + (uint64_t)(kStcTile >> 4);
        const uint32_t acc = tmem_base + (uint32_t)(st * 128);
        // the small cross terms first, while the accumulator is still small, then the four hi x hi products
#pragma unroll
        for (int k8 = 0; k8 < 4; ++k8) {
          tc_mma_tf32(acc, dlo + 2 * k8, dhi + 2 * k8, kStcInstr, k8 ? 1u : 0u);
          tc_mma_tf32(acc, dhi + 2 * k8, dlo + 2 * k8, kStcInstr, 1u);
        }
#pragma unroll
        for (int k8 = 0; k8 < 4; ++k8) tc_mma_tf32(acc, dhi + 2 * k8, dhi + 2 * k8, kStcInstr, 1u);
        tc_commit(smem_u32(&bars[kBarOpEmpty + st]));      // operand stage free
        tc_commit(smem_u32(&bars[kBarAccFull + st]));      // accumulator complete
      }
    }
  } else {
    // ------------------------------------------------------------ transform warps: raw slice -> split operand
    // tiles; they also own the running sums: thread (qd, lane, half) keeps row qd*32+lane, columns half*64..+63
    const int qd = warp & 3;                       // TMEM lane quarter this warp may read
    const int half = warp >> 2;
    const int rg = tid >> 3, piece = tid & 7;      // transform role: rows rg + 32 i, pixels 4 piece .. +3
    float run[64];
#pragma unroll
    for (int i = 0; i < 64; ++i) run[i] = 0.f;
    // The tensor core adds into its fp32 accumulator with truncation, so a chain over all of K drifts
    // (-4e-5 relative at K = 4096, scripts/experiments/tf32_accum_probe.cu).  Every slice therefore gets a fresh
    // accumulator (12 MMAs) and the slices are summed here with round-to-nearest operations.  Inside a slice the
    // four hi x hi accumulations still truncate (partial sums at 1/4 .. 4/4 of the slice sum lose half an ulp
    // each on average, always toward zero: -1.2e-7 relative on this data); the factor 1 + 2^-23 folded into the
    // FMA gives that expected loss back, leaving a zero-mean error below that of a sequential fp32 chain.
    auto drain = [&](int c) {
      mbar_wait(smem_u32(&bars[kBarAccFull + (c & 3)]), (c >> 2) & 1);
      tc_fence_after();
      const uint32_t t = tmem_base + ((uint32_t)(qd * 32) << 16) + (uint32_t)((c & 3) * 128 + half * 64);
      uint32_t d0[32], d1[32];
      tc_ld32x2(t, t + 32, d0, d1);
#pragma unroll
      for (int i = 0; i < 32; ++i) {
        run[i] = fmaf(__uint_as_float(d0[i]), 1.00000011920928955f, run[i]);
        run[32 + i] = fmaf(__uint_as_float(d1[i]), 1.00000011920928955f, run[32 + i]);
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(smem_u32(&bars[kBarAccRead + (c & 3)]));
    };
#ifdef DBAF_STC_PROFILE
    long long sp[6] = {0, 0, 0, 0, 0, 0}, p0, p1;
#define SP_TICK(i) do { p1 = clock64(); sp[i] += p1 - p0; p0 = p1; } while (0)
#else
#define SP_TICK(i) do {} while (0)
#endif
    for (int c = 0; c < n_chunks; ++c) {
      const int rs = c % kStcRawStages, st = c & 3;
#ifdef DBAF_STC_PROFILE
      p0 = clock64();
#endif
      mbar_wait(smem_u32(&bars[kBarRawFull + rs]), (c / kStcRawStages) & 1);
      SP_TICK(0);
      // thread = (row group, 16-byte piece): 4 pixels of rows rg, rg+32, rg+64, rg+96.  A warp reads and writes
      // 4 whole rows per instruction (512 contiguous bytes in, 4 x 128 swizzled bytes out): no bank conflicts
      const uint8_t* rw = raw + rs * kStcTile;
      const float4 q4r = *reinterpret_cast<const float4*>(rw + 127 * 128 + piece * 16);
      // S = (E sqrt(q)) (E sqrt(q))^T: one scaled matrix serves as both operands (and S is symmetric by construction)
      const float4 q4 = make_float4(sqrtf(q4r.x), sqrtf(q4r.y), sqrtf(q4r.z), sqrtf(q4r.w));
      float4 e4[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) e4[i] = *reinterpret_cast<const float4*>(rw + (rg + 32 * i) * 128 + piece * 16);
      __syncwarp();
      if (lane == 0) mbar_arrive(smem_u32(&bars[kBarRawEmpty + rs]));
      SP_TICK(1);
      if (c >= 3) drain(c - 3);                                // three slices behind: its MMAs have long retired
      SP_TICK(2);
      mbar_wait(smem_u32(&bars[kBarOpEmpty + st]), ((c >> 2) + 1) & 1);   // the MMAs that read this stage have retired
      SP_TICK(3);
      uint8_t* base = smem + st * kStcStageBytes;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int r = rg + 32 * i;
        const uint32_t off = (uint32_t)r * 128u + (((uint32_t)piece ^ ((uint32_t)r & 7u)) << 4);
        float4 ev = e4[i];
        if (r == 127) ev = make_float4(0.f, 0.f, 0.f, 0.f);             // row 127 only transports q
        const float4 aq = make_float4(ev.x * q4.x, ev.y * q4.y, ev.z * q4.z, ev.w * q4.w);
        const float4 ahi = make_float4(tf32_rna(aq.x), tf32_rna(aq.y), tf32_rna(aq.z), tf32_rna(aq.w));
        *reinterpret_cast<float4*>(base + off) = ahi;
        *reinterpret_cast<float4*>(base + kStcTile + off) = make_float4(aq.x - ahi.x, aq.y - ahi.y, aq.z - ahi.z, aq.w - ahi.w);
      }
      SP_TICK(4);
      fence_async_smem();                                      // generic-proxy stores -> visible to the MMA unit
      __syncwarp();
      if (lane == 0) mbar_arrive(smem_u32(&bars[kBarOpFull + st]));
      SP_TICK(5);
    }
#ifdef DBAF_STC_PROFILE
    if (tid == 0 && blockIdx.x == 0 && blockIdx.y == 500) for (int i = 0; i < 6; ++i) ws.info[10 + i] = (int)(sp[i] / n_chunks);
#endif
    for (int c = max(0, n_chunks - 3); c < n_chunks; ++c) drain(c);

    // ------------------------------------------------------------ epilogue: H -= S, v -= sv
    const int a = qd * 32 + lane;                  // accumulator row = scalar row of E q
    const int n6 = 6 * P;
    if (a < rows6) {
      const int pa = rowp[a / 6];
      const int ra6 = 6 * pa + a % 6;
      double* hrow = ws.H + (size_t)ra6 * n6;
      // lower triangle (b <= a) written to both sides, like the SIMT path
#pragma unroll
      for (int i = 0; i < 64; ++i) {
        const int b = half * 64 + i;
        if (b <= a) {
          const int cb6 = 6 * rowp[b / 6] + b % 6;
          const double val = -(double)run[i];
          atomicAdd(hrow + cb6, val);
          if (b < a) atomicAdd(ws.H + (size_t)cb6 * n6 + ra6, val);
        } else if (b == 126) {
          atomicAdd(ws.v + ra6, -(double)run[i]);
        }
      }
    }
  }
  __syncwarp();
  tc_fence_before();
  __syncthreads();
  if (warp == kStcXformWarps) {
    tc_fence_after();
    tc_dealloc(tmem_base, 512);
  }
}
