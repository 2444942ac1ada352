// Numerical probe: how accurate is a long tcgen05 kind::tf32 accumulation chain?
// D[128x128] = A[128xK] * B[128xK]^T with positive entries, compared with float64 on the host.
// Variants: 1 = inputs pre-rounded to TF32, one product; 3 / 4 = hi/lo split with 3 / 4 products;
// drain = 0: one chain over all of K in TMEM; drain = n: accumulator read back every n 32-column slices
// and summed in fp32 registers (round-to-nearest adds).
// build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -I dba-fusion_b200/csrc \
//        scripts/experiments/tf32_accum_probe.cu -o scripts/experiments/tf32_probe.bin
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <vector>
#include <cuda_runtime.h>
#include "tc_ptx.cuh"
using namespace dbaf;

constexpr int kTile = 128 * 128;
constexpr uint32_t kInstr = (1u << 4) | (2u << 7) | (2u << 10) | ((128u >> 3) << 17) | ((128u >> 4) << 24);

__device__ __forceinline__ float rna(float x) { uint32_t r; asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x)); return __uint_as_float(r); }

__global__ void __launch_bounds__(128) probe(const float* A, const float* B, int K, int variant, int drain, float* D, float corr_ulps) {
  extern __shared__ uint8_t raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~(uintptr_t)1023);
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if (tid == 0) { mbar_init(smem_u32(&bar), 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
  if (warp == 0) tc_alloc(smem_u32(&slot), 128);
  tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t tmem = slot;
  float run[128];
  for (int i = 0; i < 128; ++i) run[i] = 0.f;
  int phase = 0, in_chain = 0;
  const int chunks = K / 32;
  for (int c = 0; c < chunks; ++c) {
    for (int i = 0; i < 32; ++i) {
      const int r = warp + 4 * i;
      const uint32_t off = (uint32_t)r * 128u + ((((uint32_t)lane >> 2) ^ ((uint32_t)r & 7u)) << 4) + (((uint32_t)lane & 3u) << 2);
      const float a = A[(size_t)r * K + c * 32 + lane], b = B[(size_t)r * K + c * 32 + lane];
      const float ah = rna(a), bh = rna(b);
      *reinterpret_cast<float*>(smem + off) = ah;
      *reinterpret_cast<float*>(smem + kTile + off) = rna(a - ah);
      *reinterpret_cast<float*>(smem + 2 * kTile + off) = bh;
      *reinterpret_cast<float*>(smem + 3 * kTile + off) = rna(b - bh);
    }
    fence_async_smem();
    __syncthreads();
    if (tid == 0) {
      tc_fence_after();
      const uint32_t base = smem_u32(smem);
      for (int k8 = 0; k8 < 4; ++k8) {
        const uint64_t ahi = umma_desc_k128(base + k8 * 32), alo = umma_desc_k128(base + kTile + k8 * 32);
        const uint64_t bhi = umma_desc_k128(base + 2 * kTile + k8 * 32), blo = umma_desc_k128(base + 3 * kTile + k8 * 32);
        uint32_t acc = (in_chain | k8) ? 1u : 0u;
        if (variant == 4) { tc_mma_tf32(tmem, alo, blo, kInstr, acc); acc = 1u; }
        if (variant >= 3 && variant != 5) { tc_mma_tf32(tmem, alo, bhi, kInstr, acc); acc = 1u; tc_mma_tf32(tmem, ahi, blo, kInstr, acc); }
        if (variant != 5) tc_mma_tf32(tmem, ahi, bhi, kInstr, acc);
      }
      if (variant == 5) {
        for (int k8 = 0; k8 < 4; ++k8) {
          const uint64_t ahi = umma_desc_k128(base + k8 * 32), alo = umma_desc_k128(base + kTile + k8 * 32);
          const uint64_t bhi = umma_desc_k128(base + 2 * kTile + k8 * 32), blo = umma_desc_k128(base + 3 * kTile + k8 * 32);
          tc_mma_tf32(tmem, alo, bhi, kInstr, (in_chain | k8) ? 1u : 0u);
          tc_mma_tf32(tmem, ahi, blo, kInstr, 1u);
        }
        for (int k8 = 0; k8 < 4; ++k8)
          tc_mma_tf32(tmem, umma_desc_k128(base + k8 * 32), umma_desc_k128(base + 2 * kTile + k8 * 32), kInstr, 1u);
      }
      tc_commit(smem_u32(&bar));
    }
    mbar_wait(smem_u32(&bar), phase); phase ^= 1;
    tc_fence_after();
    ++in_chain;
    if ((drain > 0 && in_chain == drain) || c == chunks - 1) {
      for (int cb = 0; cb < 4; ++cb) {
        float d[32];
        tc_ld32(tmem + ((uint32_t)(warp * 32) << 16) + cb * 32, d);
        for (int i = 0; i < 32; ++i) {
          float v = d[i];
          if (variant == 5) {
            // expected loss of the four truncating hi x hi accumulations (partial sums ~ i/4 of the slice sum):
            // half an ulp each, ulp(v) = 2^(exponent - 23)
            const float ulp = __uint_as_float((__float_as_uint(v) & 0x7f800000u)) * 1.1920929e-07f;
            v += copysignf(corr_ulps * ulp, v);
          }
          run[cb * 32 + i] += v;
        }
      }
      tc_fence_before();
      in_chain = 0;
    }
    __syncthreads();
  }
  for (int i = 0; i < 128; ++i) D[(size_t)tid * 128 + i] = run[i];
  tc_fence_before(); __syncthreads();
  if (warp == 0) { tc_fence_after(); tc_dealloc(tmem, 128); }
}

int main() {
  for (int K : {768, 4096, 16384}) {
    std::vector<float> A((size_t)128 * K), B((size_t)128 * K);
    srand(1);
    for (auto& x : A) x = 0.5f + (float)rand() / RAND_MAX;
    for (auto& x : B) x = 0.5f + (float)rand() / RAND_MAX;
    std::vector<double> ref(128 * 128);
    for (int i = 0; i < 128; ++i) for (int j = 0; j < 128; ++j) {
      double s = 0; for (int k = 0; k < K; ++k) s += (double)A[(size_t)i * K + k] * B[(size_t)j * K + k];
      ref[i * 128 + j] = s;
    }
    // fp32 round-to-nearest FMA chain for comparison
    double fbias = 0, frms = 0;
    for (int i = 0; i < 128; i += 8) for (int j = 0; j < 128; j += 8) {
      float s = 0; for (int k = 0; k < K; ++k) s = fmaf(A[(size_t)i * K + k], B[(size_t)j * K + k], s);
      const double r = (s - ref[i * 128 + j]) / ref[i * 128 + j]; fbias += r; frms += r * r;
    }
    printf("K=%d fp32-fma-chain bias %.3e rms %.3e\n", K, fbias / 256, sqrt(frms / 256));
    float *dA, *dB, *dD;
    cudaMalloc(&dA, A.size() * 4); cudaMalloc(&dB, B.size() * 4); cudaMalloc(&dD, 128 * 128 * 4);
    cudaMemcpy(dA, A.data(), A.size() * 4, cudaMemcpyHostToDevice);
    cudaMemcpy(dB, B.data(), B.size() * 4, cudaMemcpyHostToDevice);
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 4 * kTile + 1024);
    for (int pass = 0; pass < 16; ++pass) {
      const int variant = pass < 12 ? (pass < 4 ? 1 : pass < 8 ? 3 : 4) : 5;
      const int drain = pass < 12 ? (pass % 4 == 0 ? 0 : pass % 4 == 1 ? 1 : pass % 4 == 2 ? 4 : 16) : 1;
      const float corr = pass < 12 ? 0.f : (pass == 12 ? 0.f : pass == 13 ? 1.0f : pass == 14 ? 1.25f : 1.5f);
      probe<<<1, 128, 4 * kTile + 1024>>>(dA, dB, K, variant, drain, dD, corr);
      std::vector<float> D(128 * 128);
      cudaError_t e = cudaMemcpy(D.data(), dD, D.size() * 4, cudaMemcpyDeviceToHost);
      if (e != cudaSuccess) { printf("cuda error %s\n", cudaGetErrorString(e)); return 1; }
      double bias = 0, rms = 0;
      for (int i = 0; i < 128 * 128; ++i) { const double r = (D[i] - ref[i]) / ref[i]; bias += r; rms += r * r; }
      printf("K=%d variant=%d drain=%d corr=%.2f  bias %.3e  rms %.3e\n", K, variant, drain, corr, bias / 16384, sqrt(rms / 16384));
    }
    cudaFree(dA); cudaFree(dB); cudaFree(dD);
  }
  return 0;
}
