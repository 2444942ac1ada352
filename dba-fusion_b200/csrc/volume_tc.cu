// All-pairs correlation volume on the 5th-generation tensor cores (tcgen05), sm_100a.
//
//   L0[e, p1, p2] = rn16( (1/16) * sum_c f1[e, c, p1] * f2[e, c, p2] )          fp16 in, fp32 accumulate
//
// One CTA = one edge x 128 source pixels (UMMA M = 128).  The A tile (128 px x 128 ch, 32 KB) is
// loaded once by TMA; the CTA then walks the target image in chunks of 256 pixels (UMMA N = 256):
// TMA brings the B chunk (64 KB) into a single shared-memory buffer, one thread issues 8
// tcgen05.mma (K = 16 each) into one of two 256-column TMEM accumulator stages, and four epilogue
// warps drain the other stage (tcgen05.ld, scale, round to fp16, store).  The epilogue is the long
// pole (the kernel is bound by the HBM write of the volume: 128 FLOP per output byte is below the
// machine's ~220 FLOP/B ridge), so a single B buffer is enough: it is free again as soon as the MMAs
// of the chunk have committed, long before the epilogue of that chunk is finished.
//
// With 64-pixel-wide frames a 256-pixel chunk is exactly 4 target rows, and the 2x2 / 4x4 / 8x8
// average-pool pyramid of dbaf/modules/corr.py:35-38 is produced in the same epilogue from registers
// (level 3 needs 8 rows = two chunks; the level-2 row of the even chunk is carried in registers).
// The pooled levels are computed from the fp16-rounded level below, in the reference's order
// ((a+b)+c)+d, *0.25, one rounding), so they are bit-identical to avg_pool2d of our level 0.
//
// Operands are K-major (pixel-major, 128 channels contiguous = two 128-byte swizzle rows per pixel);
// a pre-pass transposes the channel-major feature maps of DepthVideo into that layout once per call.
#include "common.cuh"
#include "tc_ptx.cuh"

#include <cuda.h>

#include <algorithm>

namespace dbaf {
namespace {

constexpr int kBM = 128;          // source pixels per CTA (UMMA M)
constexpr int kBN = 256;          // target pixels per chunk (UMMA N)
constexpr int kC = 128;           // channels (K)
constexpr int kKBlock = 64;       // channels per 128-byte swizzle row
constexpr int kSmemA = kBM * kC * 2;                 // 32 KB
constexpr int kSmemB = kBN * kC * 2;                 // 64 KB
constexpr int kStageWarp = 3 * 4096;                 // per fused epilogue warp: 2 level-0 boxes + 1 level-1 box, 32 px x 128 B each
// generic shapes stage a 32 x n_mma tile per epilogue warp with an odd word pitch; eight warps fit up to a pitch of 121
constexpr int kGenPitchMax = 121;
constexpr int kSmemStage = 8 * 32 * kGenPitchMax * 4;    // 121 KB (the fused kernel uses 8 x 12 KB of it)
// pipeline depth.  Fused (64-wide frames): chunks of 128 target pixels (two image rows), two B buffers of 32 KB and
// four 128-column accumulators, so that the next B tile is on its way while the tensor core and the epilogue work.
// Generic shapes: chunks of up to 256 pixels, one B buffer, two 256-column accumulators.
template <bool FUSE> struct Pipe {
  static constexpr int kNB = FUSE ? 2 : 1;           // B buffers
  static constexpr int kNAcc = FUSE ? 4 : 2;         // TMEM accumulator stages
  static constexpr int kAccCols = FUSE ? 128 : 256;  // columns per stage
};
constexpr int kBarA = 0, kBarBFull = 1, kBarBFree = 3, kBarAccFull = 5 /* [epilogue set][stage] */, kBarAccFree = 13, kNumBars = 17;
constexpr int kSmemBytes = kSmemA + kSmemB + kSmemStage + 1024 /*alignment slack*/ + 256 /*barriers*/;
// warp 0: TMA, warp 1: MMA, warp 2: TMEM alloc, warps 4-7 and 8-11: two sets of epilogue warps that take the chunks
// in turn (fused: groups of four chunks; generic shapes: every other chunk)
constexpr int kThreadsTc = 384;

// instruction descriptor, kind::f16: D = F32 (bit 4), A = B = F16, both K-major, N >> 3 at bit 17, M >> 4 at bit 24

__device__ __forceinline__ uint32_t pack_h2(__half a, __half b) {
  return (uint32_t)__half_as_ushort(a) | ((uint32_t)__half_as_ushort(b) << 16);
}
__device__ __forceinline__ __half pool_h(__half a, __half b, __half c, __half d) {
  return __float2half_rn((((__half2float(a) + __half2float(b)) + __half2float(c)) + __half2float(d)) * 0.25f);
}

struct Pyr {
  __half* l[4];
};

// ------------------------------------------------------------------ transpose pre-pass ----
// [n, C, HW] (channel-major) -> [n, HW, C] (pixel-major): 32 px x 64 ch tiles through shared memory
__global__ void __launch_bounds__(256) k_to_pixel_major(const __half* __restrict__ src, __half* __restrict__ dst, int C, int HW,
                                                        float scale) {
  __shared__ __half tile[64][33];
  const int e = blockIdx.z, p0 = blockIdx.x * 32, c0 = blockIdx.y * 64;
  const __half* s = src + (size_t)e * C * HW;
  __half* d = dst + (size_t)e * HW * C;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  for (int c = ty; c < 64; c += 8) {
    const int p = p0 + tx;
    // scale is a power of two: exact in fp16 (barring subnormals), like the reference's two /4.0
    tile[c][tx] = (p < HW && c0 + c < C) ? __float2half_rn(__half2float(s[(size_t)(c0 + c) * HW + p]) * scale) : __float2half(0.f);
  }
  __syncthreads();
  // each thread writes a pair of channels (4 B) so that a warp writes 128 contiguous bytes of one pixel
  for (int p = ty; p < 32; p += 8) {
    if (p0 + p < HW && c0 + 2 * tx + 1 < C) {
      const uint32_t v = pack_h2(tile[2 * tx][p], tile[2 * tx + 1][p]);
      *reinterpret_cast<uint32_t*>(d + (size_t)(p0 + p) * C + c0 + 2 * tx) = v;
    }
  }
}

// ------------------------------------------------------------------ main kernel ----
template <bool FUSE>
__global__ void __launch_bounds__(kThreadsTc, 1)
k_volume_tc(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapB,
            const __grid_constant__ CUtensorMap mapO0, const __grid_constant__ CUtensorMap mapO1,
            int HW, int m_tiles, Pyr out, const int* __restrict__ idx1, const int* __restrict__ idx2,
            int chunk_px, int n_mma, int kr, int wd, const int* __restrict__ out_slot, const int* __restrict__ bad_flag,
            int n_sets) {
  // n_sets: generic shapes only -- 2 when the staged tiles of eight epilogue warps fit into shared memory, else 1
  if (bad_flag && *bad_flag) return;                 // an edge index outside its frame stack (checked by k_check_index)
  // chunk_px: target pixels a chunk advances by; n_mma: UMMA N (chunk_px rounded up to 16, <= 256);
  // kr > 0 (generic shapes only): a chunk is 2*kr whole image rows of width wd, and level 1 is pooled from it here
  extern __shared__ uint8_t smem_raw[];
  // 1 KB alignment as an offset into the array (not through an integer cast), so that the compiler keeps
  // treating everything below as shared memory and emits LDS/STS rather than generic accesses
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* sA = smem;
  uint8_t* sB = smem + kSmemA;
  uint8_t* sStage = smem + kSmemA + kSmemB;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kSmemA + kSmemB + kSmemStage);
  // bars[kBarA]: A landed, [kBarBFull + b]: B buffer b landed, [kBarBFree + b]: B buffer b free,
  // [kBarAccFull + s]: accumulator stage s complete, [kBarAccFree + s]: accumulator stage s drained
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + kNumBars + 1);
  constexpr int NB = Pipe<FUSE>::kNB, NACC = Pipe<FUSE>::kNAcc, ACC_COLS = Pipe<FUSE>::kAccCols;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int e = blockIdx.x / m_tiles;
  const int m0 = (blockIdx.x - e * m_tiles) * kBM;
  const int n_chunks = (HW + chunk_px - 1) / chunk_px;
  // The CTAs of an edge all walk the same target frame.  Started together at chunk 0 they would ask the same few L2
  // slices for the same B tile at the same time; each M tile therefore starts at its own chunk (a multiple of four:
  // the fused pooling works on groups of four chunks) and wraps around.
  const int rot = (((blockIdx.x - e * m_tiles) * 4) % n_chunks) & ~3;
  auto data_chunk = [&](int k) { const int cd = k + rot; return cd >= n_chunks ? cd - n_chunks : cd; };
  // operand rows: the edge's own maps, or frames picked by index (video.fmaps[ii] / [jj], covisible_graph.py:129-130)
  const int ea = idx1 ? idx1[e] : e, eb = idx2 ? idx2[e] : e;
  // where the edge's volume goes: its own position, or a slot of a larger pool (CorrPool)
  const int eo = out_slot ? out_slot[e] : e;

  if (threadIdx.x == 0) {
    mbar_init(smem_u32(&bars[kBarA]), 1);
    for (int b = 0; b < NB; ++b) {
      mbar_init(smem_u32(&bars[kBarBFull + b]), 1);
      mbar_init(smem_u32(&bars[kBarBFree + b]), 1);
    }
    for (int a = 0; a < NACC; ++a) {
      mbar_init(smem_u32(&bars[kBarAccFull + a]), 1);
      mbar_init(smem_u32(&bars[kBarAccFull + 4 + a]), 1);
      mbar_init(smem_u32(&bars[kBarAccFree + a]), 4);   // one arrival per epilogue warp of the set that owns the chunk
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 2) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(512)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
    // ---------------------------------------------------------------- TMA producer
    asm volatile("prefetch.tensormap [%0];" ::"l"(&mapA) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&mapB) : "memory");
    const uint32_t barA = smem_u32(&bars[kBarA]);
    mbar_expect_tx(barA, kSmemA);
    tma_load_2d(smem_u32(sA), &mapA, barA, 0, ea * HW + m0);
    tma_load_2d(smem_u32(sA) + kBM * 128, &mapA, barA, kKBlock, ea * HW + m0);
#ifdef DBAF_VOL_PROFILE
    long long pp = 0;
#endif
    for (int c = 0; c < n_chunks; ++c) {
      const int b = c % NB, use = c / NB;
      const uint32_t barB = smem_u32(&bars[kBarBFull + b]);
      const uint32_t dstB = smem_u32(sB) + (uint32_t)b * (kSmemB / NB);
#ifdef DBAF_VOL_PROFILE
      const long long p0c = clock64();
#endif
      mbar_wait(smem_u32(&bars[kBarBFree + b]), (use + 1) & 1);
#ifdef DBAF_VOL_PROFILE
      pp += clock64() - p0c;
      if (FUSE && blockIdx.x == 700 && c == n_chunks - 1) printf("k_volume_tc producer: wait B buffer free %lld cycles per chunk\n", pp / n_chunks);
#endif
#ifdef DBAF_VOL_NOLOAD
      if (c >= NB) { mbar_arrive(barB); continue; }             // experiment: the B tiles are loaded once
#endif
      mbar_expect_tx(barB, (uint32_t)n_mma * 256u);
      const int cd = data_chunk(c);
      tma_load_2d(dstB, &mapB, barB, 0, eb * HW + cd * chunk_px);
      tma_load_2d(dstB + n_mma * 128, &mapB, barB, kKBlock, eb * HW + cd * chunk_px);
    }
    }
    __syncwarp();
  } else if (warp == 1) {
    if (lane == 0) {
    // ---------------------------------------------------------------- MMA issuer
    mbar_wait(smem_u32(&bars[kBarA]), 0);
    const uint32_t idesc = (1u << 4) | ((uint32_t)(n_mma >> 3) << 17) | ((uint32_t)(kBM >> 4) << 24);
#ifdef DBAF_VOL_PROFILE
    long long mp[3] = {0, 0, 0}, m0c, m1c;
#define MP_TICK(i) do { m1c = clock64(); mp[i] += m1c - m0c; m0c = m1c; } while (0)
#else
#define MP_TICK(i) do {} while (0)
#endif
    for (int c = 0; c < n_chunks; ++c) {
      const int s = c % NACC, u = c / NACC, b = c % NB;
#ifdef DBAF_VOL_PROFILE
      m0c = clock64();
#endif
      mbar_wait(smem_u32(&bars[kBarAccFree + s]), (u + 1) & 1);      // epilogue has drained this accumulator stage
      MP_TICK(0);
      mbar_wait(smem_u32(&bars[kBarBFull + b]), (c / NB) & 1);
      MP_TICK(1);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + (uint32_t)(s * ACC_COLS);
      const uint32_t bB = smem_u32(sB) + (uint32_t)b * (kSmemB / NB);
#pragma unroll
      for (int kb = 0; kb < 2; ++kb) {
#pragma unroll
        for (int k4 = 0; k4 < 4; ++k4) {
          const uint64_t ad = umma_desc_k128(smem_u32(sA) + kb * (kBM * 128) + k4 * 32);
          const uint64_t bd = umma_desc_k128(bB + kb * (n_mma * 128) + k4 * 32);
          tc_mma_f16(d_tmem, ad, bd, idesc, (kb | k4) ? 1u : 0u);
        }
      }
      tc_commit(smem_u32(&bars[kBarBFree + b]));     // B buffer may be overwritten once these MMAs are done
      // accumulator stage s is complete: tell the epilogue warp set that owns this chunk (each set has its own
      // barriers, so that it sees their phases one by one whatever the other set does)
      tc_commit(smem_u32(&bars[kBarAccFull + (FUSE ? ((c >> 2) & 1) : (n_sets == 2 ? (c & 1) : 0)) * 4 + s]));
      MP_TICK(2);
    }
#ifdef DBAF_VOL_PROFILE
    if (FUSE && blockIdx.x == 700)
      printf("k_volume_tc MMA thread, cycles per chunk: wait accumulator drained %lld, wait B tile %lld, issue %lld\n",
             mp[0] / n_chunks, mp[1] / n_chunks, mp[2] / n_chunks);
#endif
    }
    __syncwarp();
  } else if (warp >= 4) {
    // ---------------------------------------------------------------- epilogue
    const int q = (warp - 4) & 3;                     // TMEM lane quarter == warp % 4
    const int eset = (warp - 4) >> 2;                 // fused: which set of epilogue warps (groups of four chunks alternate)
    const int m = m0 + q * 32 + lane;                 // this thread's source pixel
    const bool mok = m < HW;
    const size_t row = (size_t)eo * HW + (mok ? m : 0);
    uint32_t carry1[16];                              // level-1 row of the even chunk (FUSE)
    uint32_t carry2[8];                               // level-2 row of the even chunk pair (FUSE)
#pragma unroll
    for (int i = 0; i < 16; ++i) carry1[i] = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) carry2[i] = 0;
#ifdef DBAF_VOL_PROFILE
    long long vp[4] = {0, 0, 0, 0}, v0, v1;
#define VP_TICK(i) do { v1 = clock64(); vp[i] += v1 - v0; v0 = v1; } while (0)
#else
#define VP_TICK(i) do {} while (0)
#endif
    for (int c = 0; c < n_chunks; ++c) {
      const int s = c % NACC, u = c / NACC;
#ifdef DBAF_VOL_PROFILE
      v0 = clock64();
#endif
      // fused: four chunks (eight target rows, one level-3 row) form a group and the two warp sets take the groups
      // in turn
      if (FUSE && ((c >> 2) & 1) != eset) continue;
      if (!FUSE && (n_sets == 2 ? (c & 1) != eset : eset != 0)) continue;    // generic: every other chunk (= one stage each)
      const int cd = data_chunk(c);                    // which chunk of the target frame this pipeline step holds
      mbar_wait(smem_u32(&bars[kBarAccFull + eset * 4 + s]), (FUSE ? (c >> 3) : (n_sets == 2 ? (c >> 1) : u)) & 1);
      tc_fence_after();
      VP_TICK(0);
      const uint32_t t0 = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(s * ACC_COLS);
      if constexpr (FUSE) {
        // chunk = target rows 2c, 2c+1 of a 64-wide frame = level-1 row c.  The stores are issued by different lanes
        // (bulk groups are per thread): lane 0 / lane 1 own the two level-0 boxes of a chunk, lane 4 the level-1
        // box, which is filled by two consecutive chunks.
        uint8_t* stage = sStage + (eset * 4 + q) * kStageWarp;
        uint8_t* b0 = stage;
        uint8_t* b1 = b0 + 4096;
        uint8_t* l1box = stage + 2 * 4096;
#ifdef DBAF_VOL_TMA_L0
        if (lane < 2) bulk_wait_read0();                        // this lane's box of the previous chunk has left shared memory
#endif
        if (lane == 4 && (c & 1) == 0) bulk_wait_read0();       // the level-1 box of the previous chunk pair
        __syncwarp();
        VP_TICK(1);
        uint32_t l1row[16];                                     // level-1 row c (32 halves)
#pragma unroll
        for (int xb = 0; xb < 2; ++xb) {
          uint32_t r0[32], r1[32];
          tc_ld32x2(t0 + (uint32_t)(xb * 32), t0 + (uint32_t)(64 + xb * 32), r0, r1);
          // fp32 -> fp16 two at a time; the pooled value is built from the ROUNDED level-0 values
          __half2 p0[16], p1[16];
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            p0[i] = __floats2half2_rn(__uint_as_float(r0[2 * i]), __uint_as_float(r0[2 * i + 1]));
            p1[i] = __floats2half2_rn(__uint_as_float(r1[2 * i]), __uint_as_float(r1[2 * i + 1]));
          }
          {
            const uint32_t* w0 = reinterpret_cast<const uint32_t*>(p0);
            const uint32_t* w1 = reinterpret_cast<const uint32_t*>(p1);
#pragma unroll
            for (int g = 0; g < 4; ++g) {
              *swz(b0, lane, xb * 4 + g) = make_uint4(w0[4 * g], w0[4 * g + 1], w0[4 * g + 2], w0[4 * g + 3]);
              *swz(b1, lane, xb * 4 + g) = make_uint4(w1[4 * g], w1[4 * g + 1], w1[4 * g + 2], w1[4 * g + 3]);
            }
          }
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            // level-1 columns 2i and 2i+1 of this half row: ((a+b)+c)+d, *0.25, one rounding
            const float2 a0 = __half22float2(p0[2 * i]), a1 = __half22float2(p1[2 * i]);
            const float2 c0 = __half22float2(p0[2 * i + 1]), c1 = __half22float2(p1[2 * i + 1]);
            const float lo = (((a0.x + a0.y) + a1.x) + a1.y) * 0.25f;
            const float hi = (((c0.x + c0.y) + c1.x) + c1.y) * 0.25f;
            const __half2 pr = __floats2half2_rn(lo, hi);
            l1row[xb * 8 + i] = *reinterpret_cast<const uint32_t*>(&pr);
          }
        }
        // this accumulator stage has been read completely
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(smem_u32(&bars[kBarAccFree + s]));
        if (out.l[1]) {
#pragma unroll
          for (int g = 0; g < 4; ++g)
            *swz(l1box, lane, (c & 1) * 4 + g) = make_uint4(l1row[4 * g], l1row[4 * g + 1], l1row[4 * g + 2], l1row[4 * g + 3]);
        }
        VP_TICK(2);
        const int r0g = eo * HW + m0 + q * 32;
#ifdef DBAF_VOL_TMA_L0
        // hand the staged tiles to the TMA store engine
        fence_async_smem();
        __syncwarp();
        if (lane < 2) {
          tma_store_2d(&mapO0, smem_u32(lane == 0 ? b0 : b1), cd * 128 + lane * 64, r0g);
          bulk_commit();
        } else if (lane == 4 && (c & 1) && out.l[1]) {
          tma_store_2d(&mapO1, smem_u32(l1box), (cd >> 1) * 64, r0g);
          bulk_commit();
        }
#else
        // Level 0 leaves through the load/store unit: the two boxes hold target rows 2c and 2c+1, which are adjacent
        // in the volume (256 contiguous bytes per source pixel), so half a warp writes one source pixel's 256 bytes
        // with 16-byte stores and a warp instruction covers two source pixels.  (Measured on this part: 128-byte TMA
        // store boxes top out near 4.1 TB/s chip-wide, plain 16-byte stores in this pattern reach 5.8 TB/s.)
        if ((c & 1) && out.l[1]) fence_async_smem();            // the level-1 box still goes through TMA
        __syncwarp();
        {
          const int half_id = lane >> 4, piece = lane & 15;     // which of the two source pixels, which 16 bytes of 256
          const uint8_t* box = (piece & 8) ? b1 : b0;
          const int nrow = min(32, HW - (m0 + q * 32));
          __half* dst0 = out.l[0] + (size_t)r0g * (size_t)HW + (size_t)cd * 128 + piece * 8;
#pragma unroll 4
          for (int r = half_id; r < nrow; r += 2) {
            const uint4 v = *reinterpret_cast<const uint4*>(box + r * 128 + (((piece & 7) ^ (r & 7)) << 4));
#ifndef DBAF_VOL_NOSTORE
            *reinterpret_cast<uint4*>(dst0 + (size_t)r * HW) = v;
#endif
          }
        }
        if (lane == 4 && (c & 1) && out.l[1]) {
          tma_store_2d(&mapO1, smem_u32(l1box), (cd >> 1) * 64, r0g);
          bulk_commit();
        }
        __syncwarp();                                           // the boxes may be overwritten by the next chunk
#endif
        if ((c & 1) == 0) {
#pragma unroll
          for (int i = 0; i < 16; ++i) carry1[i] = l1row[i];
        } else {
          // level 2: row c/2 (16 values) from level-1 rows c-1, c
          uint32_t l2[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const uint32_t a0 = carry1[2 * i], a1 = carry1[2 * i + 1];
            const uint32_t b0v = l1row[2 * i], b1v = l1row[2 * i + 1];
            const __half x = pool_h(__ushort_as_half(a0 & 0xffff), __ushort_as_half(a0 >> 16),
                                    __ushort_as_half(b0v & 0xffff), __ushort_as_half(b0v >> 16));
            const __half y = pool_h(__ushort_as_half(a1 & 0xffff), __ushort_as_half(a1 >> 16),
                                    __ushort_as_half(b1v & 0xffff), __ushort_as_half(b1v >> 16));
            l2[i] = pack_h2(x, y);
          }
          const int c2 = cd >> 1;
          if (mok && out.l[2]) {
            uint4* d = reinterpret_cast<uint4*>(out.l[2] + row * (size_t)(HW / 16) + (size_t)c2 * 16);
            d[0] = make_uint4(l2[0], l2[1], l2[2], l2[3]);
            d[1] = make_uint4(l2[4], l2[5], l2[6], l2[7]);
          }
          if ((c2 & 1) == 0) {
#pragma unroll
            for (int i = 0; i < 8; ++i) carry2[i] = l2[i];
          } else if (mok && out.l[3]) {
            uint32_t l3[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              const uint32_t a0 = carry2[2 * i], a1 = carry2[2 * i + 1];
              const uint32_t b0v = l2[2 * i], b1v = l2[2 * i + 1];
              const __half x = pool_h(__ushort_as_half(a0 & 0xffff), __ushort_as_half(a0 >> 16),
                                      __ushort_as_half(b0v & 0xffff), __ushort_as_half(b0v >> 16));
              const __half y = pool_h(__ushort_as_half(a1 & 0xffff), __ushort_as_half(a1 >> 16),
                                      __ushort_as_half(b1v & 0xffff), __ushort_as_half(b1v >> 16));
              l3[i] = pack_h2(x, y);
            }
            *reinterpret_cast<uint4*>(out.l[3] + row * (size_t)(HW / 64) + (size_t)(c2 >> 1) * 8) =
                make_uint4(l3[0], l3[1], l3[2], l3[3]);
          }
        }
        VP_TICK(3);
      } else {
        // generic shapes: a chunk is 2*kr whole target-image rows (or 256 consecutive pixels when a row pair does
        // not fit); level 0 is written here, level 1 is pooled from the staged tile.  A row of the volume is
        // HW halves, in general neither 16- nor 4-byte aligned (55 x 55, 28 x 107), so the lane-per-row
        // accumulator layout is turned around through shared memory: the warp's 32 x 256 tile is staged with a
        // 129-word row pitch (conflict-free both ways) and written out row by row, every store instruction
        // covering 64 contiguous bytes of one row.
        const int n0 = cd * chunk_px;
        const int nb = (n_mma + 31) >> 5;                // 32-column blocks of the accumulator that hold data
        const int tp = 16 * nb + 1;                      // tile pitch in words (odd: conflict-free both ways)
        uint32_t* tile = reinterpret_cast<uint32_t*>(sStage) + (size_t)(eset * 4 + q) * 32 * tp;      // [32][tp] words
        __syncwarp();                                    // the previous chunk's rows have been read out
#pragma unroll 1
        for (int b = 0; b < nb; ++b) {
          float v[32];
          tc_ld32(t0 + (uint32_t)(b * 32), v);
#pragma unroll
          for (int i = 0; i < 16; ++i)
            tile[lane * tp + b * 16 + i] = pack_h2(__float2half_rn(v[2 * i]), __float2half_rn(v[2 * i + 1]));
        }
        // the accumulator stage is free as soon as it sits in shared memory
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(smem_u32(&bars[kBarAccFree + s]));
        const int ncol = min(chunk_px, HW - n0);
        const int mrow0 = m0 + q * 32;
        // (all eight loads issued before the first store, and plain shared-memory loads: as generic loads with a
        //  store behind each one this loop cost 20 k cycles per chunk)
        const int nrow = min(32, HW - mrow0);
#pragma unroll 2
        for (int r = 0; r < nrow; ++r) {
          unsigned short* dst = reinterpret_cast<unsigned short*>(out.l[0]) + ((size_t)eo * HW + mrow0 + r) * (size_t)HW + n0 + lane;
          unsigned short v[kBN / 32];
          const unsigned short* src = reinterpret_cast<const unsigned short*>(tile) + r * (2 * tp) + lane;
#pragma unroll
          for (int k = 0; k < kBN / 32; ++k) v[k] = k < nb ? src[k * 32] : (unsigned short)0;
#pragma unroll
          for (int k = 0; k < kBN / 32; ++k)
            if (k * 32 + lane < ncol) dst[k * 32] = v[k];
        }
        if (kr > 0 && out.l[1] != nullptr) {
          // level 1 (dbaf/modules/corr.py:35-38: avg_pool2d 2x2, floor sizes) from the fp16 tile: the chunk holds
          // ncol / wd whole image rows starting at an even row, so every 2x2 window it touches is complete.
          // ((a+b)+c)+d, *0.25, one rounding, like the pooling kernel and the fused 64-wide path.
          // A chunk yields at most kr * (wd/2) <= 64 outputs per source pixel: two per lane, decoded once.
          const int w1 = wd >> 1, hw1 = ((HW / wd) >> 1) * w1;
          const int nout = ((ncol / wd) >> 1) * w1;
          int colA[2], offA[2];
#pragma unroll
          for (int t = 0; t < 2; ++t) {
            const int j = lane + 32 * t;
            const int y1 = j / w1, x1 = j - y1 * w1;
            colA[t] = j < nout ? 2 * y1 * wd + 2 * x1 : 0;      // even: the upper pair is one aligned word
            offA[t] = j < nout ? j : -1;
          }
          const unsigned short* t16 = reinterpret_cast<const unsigned short*>(tile);
          unsigned short* l1base = reinterpret_cast<unsigned short*>(out.l[1]) + ((size_t)eo * HW + mrow0) * (size_t)hw1 +
                                   (size_t)cd * kr * w1;
#pragma unroll 4
          for (int r = 0; r < nrow; ++r) {
            const unsigned short* row16 = t16 + r * (2 * tp);
            const uint32_t* row32 = tile + r * tp;
#pragma unroll
            for (int t = 0; t < 2; ++t) {
              const uint32_t ab = row32[colA[t] >> 1];
              const unsigned short cc = row16[colA[t] + wd], dd = row16[colA[t] + wd + 1];
              const __half pooled = pool_h(__ushort_as_half((unsigned short)(ab & 0xffffu)), __ushort_as_half((unsigned short)(ab >> 16)),
                                           __ushort_as_half(cc), __ushort_as_half(dd));
              if (offA[t] >= 0) l1base[(size_t)r * hw1 + offA[t]] = __half_as_ushort(pooled);
            }
          }
        }
      }
    }
#ifdef DBAF_VOL_PROFILE
    if (FUSE && warp == 4 && lane == 0 && blockIdx.x == 700)
      printf("k_volume_tc epilogue warp, cycles per chunk: wait accumulator %lld, wait store read-out %lld, tmem->fp16->stage %lld, stores + levels 2,3 %lld\n",
             vp[0] / n_chunks, vp[1] / n_chunks, vp[2] / n_chunks, vp[3] / n_chunks);
#endif
  }

  if (FUSE && warp >= 4 && (lane < 2 || lane == 4)) bulk_wait0();   // (no-op for lanes that never committed a group)
  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512) : "memory");
  }
}

// ------------------------------------------------------------------ host side ----
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_fn() {
  // resolved through the runtime so the library has no link-time dependency on libcuda
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

// generic 2-D fp16 row-major matrix [rows, cols], box = 64 columns (128 B, swizzled) x box_rows rows
int make_map2(CUtensorMap* map, const __half* base, size_t rows, size_t cols, int box_rows) {
  EncodeTiledFn fn = encode_fn();
  DBAF_REQUIRE(fn != nullptr, "cuTensorMapEncodeTiled is not available from the driver");
  const cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  const cuuint64_t strides[1] = {(cuuint64_t)cols * 2};
  const cuuint32_t box[2] = {64u, (cuuint32_t)box_rows};
  const cuuint32_t estr[2] = {1, 1};
  const CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<__half*>(base), dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  DBAF_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled (output) failed");
  return DBAF_OK;
}

// [rows, 128] fp16 pixel-major matrix, box = 64 channels (128 B, swizzled) x box_rows pixels
int make_map(CUtensorMap* map, const __half* base, size_t rows, int box_rows) {
  EncodeTiledFn fn = encode_fn();
  DBAF_REQUIRE(fn != nullptr, "cuTensorMapEncodeTiled is not available from the driver");
  const cuuint64_t dims[2] = {(cuuint64_t)kC, (cuuint64_t)rows};
  const cuuint64_t strides[1] = {(cuuint64_t)kC * 2};
  const cuuint32_t box[2] = {(cuuint32_t)kKBlock, (cuuint32_t)box_rows};
  const cuuint32_t estr[2] = {1, 1};
  const CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<__half*>(base), dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  DBAF_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed");
  return DBAF_OK;
}

}  // namespace

size_t corr_volume_tc_workspace(int n1, int n2, int C, int ht, int wd) {
  // pixel-major copies of both sets of feature maps
  return (size_t)(n1 + n2) * ht * wd * C * sizeof(__half) + 512;
}

// f1 [n1,C,ht,wd], f2 [n2,C,ht,wd]; edge e correlates f1[idx1[e]] with f2[idx2[e]] (idx null: e itself).
int corr_volume_tc(const __half* f1, int n1, const __half* f2, int n2, const int* idx1, const int* idx2, int n, int C,
                   int ht, int wd, int levels, __half* const* out, void* workspace, size_t workspace_bytes,
                   cudaStream_t st, int* levels_done, const int* out_slot, int n_out, const int* bad_flag) {
  // out_slot / n_out: edge e is written to slot out_slot[e] of outputs that hold n_out volumes (null: e itself, n)
  *levels_done = 0;
  const int HW = ht * wd;
  if (C != kC) return DBAF_OK;                              // the SIMT kernel takes other channel counts
  if (workspace == nullptr || workspace_bytes < corr_volume_tc_workspace(n1, n2, C, ht, wd)) return DBAF_OK;
  if ((size_t)std::max(n, n_out) * HW >= (size_t)1 << 31 || (size_t)std::max(n1, n2) * HW >= (size_t)1 << 31) return DBAF_OK;
  __half* a_t = reinterpret_cast<__half*>((reinterpret_cast<uintptr_t>(workspace) + 255) & ~(uintptr_t)255);
  __half* b_t = a_t + (size_t)n1 * HW * C;
  const bool shared_src = f1 == f2 && n1 == n2;             // one set of frames on both sides: transpose it once
  k_to_pixel_major<<<dim3((HW + 31) / 32, C / 64, n1), 256, 0, st>>>(f1, a_t, C, HW, 0.25f);   // fmap / 4.0, corr.py:67-68
  if (shared_src) b_t = a_t;
  else k_to_pixel_major<<<dim3((HW + 31) / 32, C / 64, n2), 256, 0, st>>>(f2, b_t, C, HW, 0.25f);
  DBAF_LAUNCH_CHECK("k_to_pixel_major");
  CUtensorMap mapA, mapB;
  if (int rc = make_map(&mapA, a_t, (size_t)n1 * HW, kBM)) return rc;
  const bool fuse = wd == 64 && ht % 8 == 0;
  // generic shapes: a chunk of 2*kr whole image rows (so that level 1 can be pooled from it) when a row pair fits
  // into one UMMA (N <= 256), else 256 consecutive pixels and level 0 only
  int kr = 0, chunk_px = kBN, n_mma = kBN;
  if (fuse) chunk_px = n_mma = Pipe<true>::kAccCols;        // two rows of a 64-wide frame per chunk
  if (!fuse && 2 * wd <= kBN && levels > 1) {
    kr = kBN / (2 * wd);
    chunk_px = 2 * wd * kr;
    n_mma = (chunk_px + 15) & ~15;
  }
  if (int rc = make_map(&mapB, b_t, (size_t)n2 * HW, n_mma)) return rc;
  const int m_tiles = (HW + kBM - 1) / kBM;
  CUtensorMap mapO0 = mapA, mapO1 = mapA;     // only the fused kernel stores through tensor maps
  if (fuse) {
    if (int rc = make_map2(&mapO0, out[0], (size_t)n_out * HW, (size_t)HW, 32)) return rc;
    mapO1 = mapO0;
    if (levels > 1)
      if (int rc = make_map2(&mapO1, out[1], (size_t)n_out * HW, (size_t)HW / 4, 32)) return rc;
  }
  Pyr pyr;
  for (int l = 0; l < 4; ++l) pyr.l[l] = (fuse && l < levels) ? out[l] : nullptr;
  pyr.l[0] = out[0];
  if (kr > 0) pyr.l[1] = out[1];
  static bool attr[64][2] = {};                       // per device: the attribute belongs to the device's copy of the kernel
  int dev_id = 0;
  DBAF_CUDA(cudaGetDevice(&dev_id));
  if (!attr[dev_id & 63][fuse]) {
    if (fuse) DBAF_CUDA(cudaFuncSetAttribute(k_volume_tc<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
    else DBAF_CUDA(cudaFuncSetAttribute(k_volume_tc<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
    attr[dev_id & 63][fuse] = true;
  }
  // generic shapes: two sets of epilogue warps when their staged tiles fit (pitch 16 * ceil(n_mma / 32) + 1 words)
  const int n_sets = (16 * ((n_mma + 31) / 32) + 1 <= kGenPitchMax) ? 2 : 1;
  if (fuse) k_volume_tc<true><<<n * m_tiles, kThreadsTc, kSmemBytes, st>>>(mapA, mapB, mapO0, mapO1, HW, m_tiles, pyr, idx1, idx2, chunk_px, n_mma, kr, wd, out_slot, bad_flag, 2);
  else k_volume_tc<false><<<n * m_tiles, kThreadsTc, kSmemBytes, st>>>(mapA, mapB, mapO0, mapO1, HW, m_tiles, pyr, idx1, idx2, chunk_px, n_mma, kr, wd, out_slot, bad_flag, n_sets);
  DBAF_LAUNCH_CHECK("k_volume_tc");
  // fused: all requested levels are written; generic with row-pair chunks: levels 0 and 1; otherwise level 0 only.
  // The caller pools the rest.
  *levels_done = fuse ? levels : (kr > 0 ? 2 : 1);
  return DBAF_OK;
}

}  // namespace dbaf
