// Shared device helpers for libdbaf_b200: SE3 algebra on (t, q=xyzw) poses, warp/block
// reductions, error plumbing.  sm_100a only.
#pragma once

#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include <stdio.h>
#include <string>

#include "../../include/dbaf_b200.h"

namespace dbaf {

// ---------------------------------------------------------------- errors ----
void set_error(const std::string& msg);
int cuda_fail(cudaError_t e, const char* what);

#define DBAF_REQUIRE(cond, msg)                                              \
  do {                                                                       \
    if (!(cond)) {                                                           \
      ::dbaf::set_error(std::string(__func__) + ": " + (msg));               \
      return DBAF_EINVAL;                                                    \
    }                                                                        \
  } while (0)

#define DBAF_LAUNCH_CHECK(what)                                              \
  do {                                                                       \
    cudaError_t e__ = cudaGetLastError();                                    \
    if (e__ != cudaSuccess) return ::dbaf::cuda_fail(e__, what);             \
  } while (0)

#define DBAF_CUDA(call)                                                      \
  do {                                                                       \
    cudaError_t e__ = (call);                                                \
    if (e__ != cudaSuccess) return ::dbaf::cuda_fail(e__, #call);            \
  } while (0)

// ---------------------------------------------------------------- options ----
// Kernel-selection switches (test and profiling aids).  Read once from the environment when the
// library is first used (DBAF_SCHUR_IMPL, DBAF_BAND_IMPL, DBAF_BAND_SWEEP, DBAF_VOLUME_IMPL,
// DBAF_BAND_DOMAINS, DBAF_STRICT), changed afterwards only through dbaf_set_option().
struct Options {
  int schur_simt;      // 1: keep the Schur complement on the fp32 SIMT kernels
  int band_single;     // 1: one-CTA banded sweep (no twisted / domain-decomposed factorisation)
  int band_twisted;    // 1: two-CTA twisted factorisation instead of the domain decomposition
  int band_smem;       // 1: banded sweep with the window in shared memory
  int band_domains;    // > 0: number of domains of the decomposed banded solve
  int volume_simt;     // 1: SIMT cross-check kernels for the correlation volume
  int eij_recompute;   // -1: automatic, 0: always materialise Eij, 1: always recompute it
};
Options& options();

constexpr float kMinDepth = 0.25f;   // MIN_DEPTH, src/droid_kernels.cu:29

// ------------------------------------------------------------------ SE3 ----
// rotate / act / relative / adjoint_T / so3_exp / se3_exp / cross_into / retract_pose below restate the closed-form
// quaternion algebra of the reference, src/droid_kernels.cu:61-178 (actSO3, actSE3, relSE3, adjSE3, expSO3, expSE3,
// crossInplace, retrSE3), with the same operation order, series constants and small-angle thresholds: the 1e-6 pose
// parity depends on evaluating these few formulas the way the reference does.
struct Se3 {
  float t[3];
  float q[4];  // x y z w
};

// v' = v + w*(2 q x v) + q x (2 q x v)
__device__ __forceinline__ void rotate(const float* q, const float* X, float* Y) {
  const float ux = 2.0f * (q[1] * X[2] - q[2] * X[1]);
  const float uy = 2.0f * (q[2] * X[0] - q[0] * X[2]);
  const float uz = 2.0f * (q[0] * X[1] - q[1] * X[0]);
  Y[0] = X[0] + q[3] * ux + (q[1] * uz - q[2] * uy);
  Y[1] = X[1] + q[3] * uy + (q[2] * ux - q[0] * uz);
  Y[2] = X[2] + q[3] * uz + (q[0] * uy - q[1] * ux);
}

// Y = (R X[0:3] + X[3] t, X[3])
__device__ __forceinline__ void act(const Se3& G, const float* X, float* Y) {
  rotate(G.q, X, Y);
  Y[3] = X[3];
  Y[0] += X[3] * G.t[0];
  Y[1] += X[3] * G.t[1];
  Y[2] += X[3] * G.t[2];
}

// Gij = Tj * Ti^-1 from raw pose rows
__device__ __forceinline__ Se3 relative(const float* Pi, const float* Pj) {
  const float* ti = Pi;
  const float* qi = Pi + 3;
  const float* tj = Pj;
  const float* qj = Pj + 3;
  Se3 G;
  G.q[0] = -qj[3] * qi[0] + qj[0] * qi[3] - qj[1] * qi[2] + qj[2] * qi[1];
  G.q[1] = -qj[3] * qi[1] + qj[1] * qi[3] - qj[2] * qi[0] + qj[0] * qi[2];
  G.q[2] = -qj[3] * qi[2] + qj[2] * qi[3] - qj[0] * qi[1] + qj[1] * qi[0];
  G.q[3] = qj[3] * qi[3] + qj[0] * qi[0] + qj[1] * qi[1] + qj[2] * qi[2];
  float r[3];
  rotate(G.q, ti, r);
  G.t[0] = tj[0] - r[0];
  G.t[1] = tj[1] - r[1];
  G.t[2] = tj[2] - r[2];
  return G;
}

__device__ __forceinline__ Se3 stereo_baseline() {
  Se3 G;
  G.t[0] = -0.1f; G.t[1] = 0.f; G.t[2] = 0.f;
  G.q[0] = 0.f; G.q[1] = 0.f; G.q[2] = 0.f; G.q[3] = 1.f;
  return G;
}

__device__ __forceinline__ Se3 load_relative(const float* poses, int ix, int jx) {
  float Pi[7], Pj[7];
#pragma unroll
  for (int k = 0; k < 7; ++k) {
    Pi[k] = __ldg(poses + 7 * ix + k);
    Pj[k] = __ldg(poses + 7 * jx + k);
  }
  return relative(Pi, Pj);
}

// Dual adjoint that carries a pose-j Jacobian row to pose i (negated by the caller).
__device__ __forceinline__ void adjoint_T(const Se3& G, const float* X, float* Y) {
  const float qinv[4] = {-G.q[0], -G.q[1], -G.q[2], G.q[3]};
  rotate(qinv, X, Y);
  rotate(qinv, X + 3, Y + 3);
  const float u[3] = {G.t[2] * X[1] - G.t[1] * X[2],
                      G.t[0] * X[2] - G.t[2] * X[0],
                      G.t[1] * X[0] - G.t[0] * X[1]};
  float v[3];
  rotate(qinv, u, v);
  Y[3] += v[0];
  Y[4] += v[1];
  Y[5] += v[2];
}

__device__ __forceinline__ void so3_exp(const float* phi, float* q) {
  const float th2 = phi[0] * phi[0] + phi[1] * phi[1] + phi[2] * phi[2];
  const float th4 = th2 * th2;
  const float th = sqrtf(th2);
  float imag, real;
  if (th2 < 1e-8f) {
    imag = 0.5f - (1.0f / 48.0f) * th2 + (1.0f / 3840.0f) * th4;
    real = 1.0f - (1.0f / 8.0f) * th2 + (1.0f / 384.0f) * th4;
  } else {
    imag = sinf(0.5f * th) / th;
    real = cosf(0.5f * th);
  }
  q[0] = imag * phi[0];
  q[1] = imag * phi[1];
  q[2] = imag * phi[2];
  q[3] = real;
}

__device__ __forceinline__ void cross_into(const float* a, float* b) {
  const float x = a[1] * b[2] - a[2] * b[1];
  const float y = a[2] * b[0] - a[0] * b[2];
  const float z = a[0] * b[1] - a[1] * b[0];
  b[0] = x; b[1] = y; b[2] = z;
}

__device__ __forceinline__ Se3 se3_exp(const float* xi) {
  Se3 G;
  so3_exp(xi + 3, G.q);
  float tau[3] = {xi[0], xi[1], xi[2]};
  const float phi[3] = {xi[3], xi[4], xi[5]};
  const float th2 = phi[0] * phi[0] + phi[1] * phi[1] + phi[2] * phi[2];
  const float th = sqrtf(th2);
  G.t[0] = tau[0]; G.t[1] = tau[1]; G.t[2] = tau[2];
  if (th > 1e-4f) {
    const float a = (1.0f - cosf(th)) / th2;
    cross_into(phi, tau);
    G.t[0] += a * tau[0]; G.t[1] += a * tau[1]; G.t[2] += a * tau[2];
    const float b = (th - sinf(th)) / (th * th2);
    cross_into(phi, tau);
    G.t[0] += b * tau[0]; G.t[1] += b * tau[1]; G.t[2] += b * tau[2];
  }
  return G;
}

// pose <- Exp(xi) * pose
__device__ __forceinline__ void retract_pose(const float* xi, float* P) {
  const Se3 d = se3_exp(xi);
  const float* q = P + 3;
  float q1[4], t1[3];
  q1[0] = d.q[3] * q[0] + d.q[0] * q[3] + d.q[1] * q[2] - d.q[2] * q[1];
  q1[1] = d.q[3] * q[1] + d.q[1] * q[3] + d.q[2] * q[0] - d.q[0] * q[2];
  q1[2] = d.q[3] * q[2] + d.q[2] * q[3] + d.q[0] * q[1] - d.q[1] * q[0];
  q1[3] = d.q[3] * q[3] - d.q[0] * q[0] - d.q[1] * q[1] - d.q[2] * q[2];
  rotate(d.q, P, t1);
  P[0] = t1[0] + d.t[0];
  P[1] = t1[1] + d.t[1];
  P[2] = t1[2] + d.t[2];
  P[3] = q1[0]; P[4] = q1[1]; P[5] = q1[2]; P[6] = q1[3];
}

// ------------------------------------------------------------ reductions ----
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Transposed warp reduction: every lane brings 32 values v[0..31]; afterwards lane l
// holds in v[0] the sum over all 32 lanes of value l.  31 shuffles instead of 160.
__device__ __forceinline__ void warp_reduce32(float* v, int lane) {
#pragma unroll
  for (int half = 16; half >= 1; half >>= 1) {
    const bool up = (lane & half) != 0;
#pragma unroll
    for (int i = 0; i < half; ++i) {
      const float send = up ? v[i] : v[i + half];
      const float keep = up ? v[i + half] : v[i];
      v[i] = keep + __shfl_xor_sync(0xffffffffu, send, half);
    }
  }
}

}  // namespace dbaf
