// How fast is the FP64 pipe of one SM?  One CTA, W warps, each thread runs `chains` independent DFMA chains.
// build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a scripts/experiments/fp64_rate_probe.cu -o scripts/experiments/fp64_probe.bin
#include <cstdio>
#include <cuda_runtime.h>
template <int CH>
__global__ void probe(double* out, long long* cyc, int iters) {
  double a[CH];
  for (int i = 0; i < CH; ++i) a[i] = threadIdx.x * 1e-3 + i;
  const double m = 1.0000001, c = 1e-9;
  __syncthreads();
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it)
#pragma unroll
    for (int i = 0; i < CH; ++i) a[i] = fma(a[i], m, c);
  const long long t1 = clock64();
  double s = 0;
  for (int i = 0; i < CH; ++i) s += a[i];
  out[threadIdx.x] = s;
  if (threadIdx.x == 0) *cyc = t1 - t0;
}
int main() {
  double* out; long long* cyc; cudaMalloc(&out, 1024 * 8); cudaMalloc(&cyc, 8);
  const int iters = 2000;
  for (int warps : {1, 4, 8, 16}) {
    long long h;
    probe<1><<<1, warps * 32>>>(out, cyc, iters); cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
    printf("warps=%2d chains=1 : %.1f cycles per dependent DFMA\n", warps, (double)h / iters);
    probe<8><<<1, warps * 32>>>(out, cyc, iters); cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
    printf("warps=%2d chains=8 : %.2f cycles per warp-DFMA per scheduler, %.1f DFMA lanes/clk/SM\n", warps,
           (double)h / (iters * 8.0 * (warps > 4 ? warps / 4.0 : 1.0)), iters * 8.0 * warps * 32 / (double)h);
  }
  return 0;
}
