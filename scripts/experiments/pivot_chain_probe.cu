// Latency of one Cholesky pivot step as a dependent chain, on one thread: the recipe used by factor_diag6
// (float rsqrt seed + one Newton step, scale, update) against reciprocal-based variants.
// nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o pivot_chain_probe.bin pivot_chain_probe.cu && ./pivot_chain_probe.bin
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ double rsqrt_nr(double d) {
  const double y0 = (double)rsqrtf((float)d);
  const double t = d * y0;
  const double e = fma(-t, y0, 1.0);
  return fma(0.5 * y0, e, y0);
}
__device__ __forceinline__ double rcp_seed32(double d) {        // float seed + two Newton steps
  const double r0 = (double)__frcp_rn((float)d);
  const double e0 = fma(-d, r0, 1.0);
  const double r1 = fma(r0, e0, r0);
  const double e1 = fma(-d, r1, 1.0);
  return fma(r1, e1, r1);
}
__device__ __forceinline__ double rcp_mufu64(double d) {        // MUFU.RCP64H seed + two Newton steps
  double r0;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r0) : "d"(d));
  const double e0 = fma(-d, r0, 1.0);
  const double r1 = fma(r0, e0, r0);
  const double e1 = fma(-d, r1, 1.0);
  return fma(r1, e1, r1);
}

__device__ __forceinline__ double rsqrt_nr64(double d) {       // MUFU.RSQ64H seed + one Newton step (what the kernels use)
  double y0;
  asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y0) : "d"(d));
  const double t = d * y0;
  const double e = fma(-t, y0, 1.0);
  return fma(0.5 * y0, e, y0);
}

template <int MODE>
__global__ void k(double* out, long long* cyc, int iters, double a0, double b0) {
  double d = a0;
  const double b = b0;
  const long long t0 = clock64();
  for (int i = 0; i < iters; ++i) {
    if (MODE == 0) {            // cholesky: inv = rsqrt(d); l = b * inv; d' = a - l * l
      const double inv = rsqrt_nr(d);
      const double l = b * inv;
      d = fma(-l, l, a0);
    } else if (MODE == 1) {     // ldl, float seed: r = 1/d; d' = a - (b*b) * r
      const double r = rcp_seed32(d);
      d = fma(-(b * b), r, a0);
    } else if (MODE == 2) {     // ldl, MUFU.RCP64H seed
      const double r = rcp_mufu64(d);
      d = fma(-(b * b), r, a0);
    } else if (MODE == 3) {     // plain division
      d = a0 - (b * b) / d;
    } else if (MODE == 4) {     // one DFMA (latency reference)
      d = fma(d, 0.999999, b);
    } else if (MODE == 6) {     // cholesky, rsqrt.approx.f64 seed + 1 Newton
      const double inv = rsqrt_nr64(d);
      const double l = b * inv;
      d = fma(-l, l, a0);
    } else {                    // library rsqrt
      const double inv = rsqrt(d);
      const double l = b * inv;
      d = fma(-l, l, a0);
    }
  }
  const long long t1 = clock64();
  out[0] = d;
  cyc[0] = t1 - t0;
}

int main() {
  double* out; long long* cyc;
  cudaMalloc(&out, 8); cudaMalloc(&cyc, 8);
  const int iters = 20000;
  const char* names[7] = {"cholesky pivot (float rsqrt seed + 1 Newton)", "LDL pivot (float rcp seed + 2 Newton)",
                          "LDL pivot (rcp.approx.f64 seed + 2 Newton)", "LDL pivot (IEEE division)", "one dependent DFMA",
                          "cholesky pivot (library rsqrt)", "cholesky pivot (rsqrt.approx.f64 seed + 1 Newton)"};
  for (int m = 0; m < 7; ++m) {
    for (int rep = 0; rep < 2; ++rep) {
      switch (m) {
        case 0: k<0><<<1, 1>>>(out, cyc, iters, 4.0, 1.0); break;
        case 1: k<1><<<1, 1>>>(out, cyc, iters, 4.0, 1.0); break;
        case 2: k<2><<<1, 1>>>(out, cyc, iters, 4.0, 1.0); break;
        case 3: k<3><<<1, 1>>>(out, cyc, iters, 4.0, 1.0); break;
        case 4: k<4><<<1, 1>>>(out, cyc, iters, 4.0, 1.0); break;
        case 5: k<5><<<1, 1>>>(out, cyc, iters, 4.0, 1.0); break;
        default: k<6><<<1, 1>>>(out, cyc, iters, 4.0, 1.0); break;
      }
      cudaDeviceSynchronize();
    }
    long long c; double d;
    cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
    cudaMemcpy(&d, out, 8, cudaMemcpyDeviceToHost);
    printf("%-48s %7.1f cycles per step (result %.15g)\n", names[m], (double)c / iters, d);
  }
  return 0;
}
