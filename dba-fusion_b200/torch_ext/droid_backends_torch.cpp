// `droid_backends` as a compiled torch extension over the C ABI of libdbaf_b200.so: the binding a maintainer of
// DBA-Fusion would keep in src/droid.cpp (PYBIND11_MODULE at src/droid.cpp:297-316), with every `*_cuda` call replaced
// by the dbaf_* entry point that include/dbaf_b200.h names for it.  Same function names, positional signatures,
// in-place contract and error behaviour (TORCH_CHECK -> RuntimeError) as the reference; torch supplies tensors and the
// current stream, nothing else.  Built by dba-fusion_b200/torch_ext/build.py (host compiler only: there is no device
// code in this file).
#include <ATen/cuda/CUDAContext.h>
#include <torch/extension.h>

#include <algorithm>
#include <vector>

#include "dbaf_b200.h"

#define CHECK_CONTIGUOUS(x) TORCH_CHECK(x.is_contiguous(), #x " must be contiguous")
#define CHECK_INPUT(x) CHECK_CONTIGUOUS(x)          // src/droid.cpp:105-106
#define DBAF_CALL(expr) TORCH_CHECK((expr) == DBAF_OK, #expr ": ", dbaf_last_error())

namespace {

void* stream() { return (void*)at::cuda::getCurrentCUDAStream().stream(); }

const float* f32(const torch::Tensor& t, const char* name) {
  TORCH_CHECK(t.is_cuda() && t.scalar_type() == torch::kFloat32, name, " must be a float32 CUDA tensor");
  return t.data_ptr<float>();
}
const int64_t* i64(const torch::Tensor& t, const char* name) {
  TORCH_CHECK(t.is_cuda() && t.scalar_type() == torch::kInt64, name, " must be an int64 CUDA tensor");
  return t.data_ptr<int64_t>();
}

// dimensions, workspace and eta bookkeeping shared by ba / ba_extend / BACore
struct Window {
  int N = 0, E = 0, ht = 0, wd = 0, t0 = 0, t1 = 0, P = 0, HW = 0, eta_rows = 1, K = 0;
  dbaf_ba_layout L{};
  torch::Tensor ws, eta;

  void setup(const torch::Tensor& poses, const torch::Tensor& disps, const torch::Tensor& disps_sens,
             const torch::Tensor& targets, const torch::Tensor& weights, const torch::Tensor& eta_in,
             const torch::Tensor& ii, const torch::Tensor& jj, int t0_, int t1_, bool need_eta) {
    TORCH_CHECK(disps.dim() == 3 && poses.dim() == 2 && poses.size(1) == 7, "poses must be [N,7] and disps [N,ht,wd]");
    N = (int)disps.size(0); ht = (int)disps.size(1); wd = (int)disps.size(2);
    E = (int)ii.numel(); t0 = t0_; t1 = t1_; P = t1 - t0; HW = ht * wd;
    TORCH_CHECK(jj.numel() == E && targets.numel() == (int64_t)E * 2 * HW && weights.numel() == targets.numel(),
                "targets/weights must be [E,2,ht,wd] with E = len(ii) = len(jj)");
    TORCH_CHECK(poses.size(0) >= t1 && disps_sens.sizes() == disps.sizes() && 0 <= t0 && t0 <= t1 && t1 <= N,
                "need 0 <= t0 <= t1 <= N and matching poses/disps/disps_sens");
    eta = eta_in.contiguous();
    eta_rows = std::max<int64_t>(1, eta.numel() / HW);
    if (need_eta) TORCH_CHECK((int64_t)eta_rows * HW == eta.numel(), "eta must be [K,ht,wd]");
    DBAF_CALL(dbaf_ba_workspace_layout(N, E, t0, t1, ht, wd, &L));
    K = eta_rows > 1 ? eta_rows : (int)L.Kcap;          // dz is sized so that the device can never write past it
    ws = torch::empty({L.total_bytes}, torch::TensorOptions().dtype(torch::kUInt8).device(poses.device()));
  }
  int dz_rows() const { return std::max<int>(K, (int)L.Kcap); }
  double* Hd() const { return reinterpret_cast<double*>(ws.data_ptr<uint8_t>() + L.H); }
};

// undamped reduced system -> the caller's float64 CPU tensors (src/droid_kernels.cu:1632-1640, 1889-1897)
void export_system(const Window& w, torch::Tensor H, torch::Tensor v) {
  TORCH_CHECK(!H.is_cuda() && !v.is_cuda() && H.scalar_type() == torch::kFloat64 && v.scalar_type() == torch::kFloat64,
              "H and v must be float64 CPU tensors");
  const int64_t n = 6 * (int64_t)w.P;
  TORCH_CHECK(H.dim() == 2 && H.size(0) <= n && H.size(1) <= n, "H is larger than the reduced camera system");
  auto host = torch::empty({n * n + n}, torch::TensorOptions().dtype(torch::kFloat64).pinned_memory(true));
  auto st = at::cuda::getCurrentCUDAStream();
  C10_CUDA_CHECK(cudaMemcpyAsync(host.data_ptr<double>(), w.Hd(), (size_t)(n * n + n) * 8, cudaMemcpyDeviceToHost, st.stream()));
  C10_CUDA_CHECK(cudaStreamSynchronize(st.stream()));
  const int64_t r = H.size(0), c = H.size(1);
  H.copy_(host.slice(0, 0, n * n).view({n, n}).slice(0, 0, r).slice(1, 0, c));
  v.slice(0, 0, r).copy_(host.slice(0, n * n, n * n + r));
}

}  // namespace

std::vector<torch::Tensor> ba(torch::Tensor poses, torch::Tensor disps, torch::Tensor intrinsics, torch::Tensor disps_sens,
                              torch::Tensor targets, torch::Tensor weights, torch::Tensor eta, torch::Tensor ii,
                              torch::Tensor jj, const int t0, const int t1, const int iterations, const float lm,
                              const float ep, const bool motion_only) {
  CHECK_INPUT(targets); CHECK_INPUT(weights); CHECK_INPUT(poses); CHECK_INPUT(disps); CHECK_INPUT(intrinsics);
  CHECK_INPUT(disps_sens); CHECK_INPUT(ii); CHECK_INPUT(jj);
  Window w;
  w.setup(poses, disps, disps_sens, targets, weights, eta, ii, jj, t0, t1, !motion_only);
  torch::Tensor dx, dz;
  if (iterations <= 0) return {dx, dz};
  dx = torch::empty({w.P, 6}, poses.options());
  if (!motion_only) dz = torch::empty({w.dz_rows(), w.HW}, poses.options());
  DBAF_CALL(dbaf_ba(w.ws.data_ptr(), poses.data_ptr<float>(), disps.data_ptr<float>(), f32(intrinsics, "intrinsics"),
                    f32(disps_sens, "disps_sens"), f32(targets, "targets"), f32(weights, "weights"), f32(w.eta, "eta"),
                    w.eta_rows, i64(ii, "ii"), i64(jj, "jj"), w.E, w.N, t0, t1, w.ht, w.wd, iterations, lm, ep,
                    motion_only ? 1 : 0, dx.data_ptr<float>(), motion_only ? nullptr : dz.data_ptr<float>(), stream()));
  if (!motion_only) dz = dz.slice(0, 0, w.eta_rows > 1 ? w.eta_rows : w.dz_rows());
  return {dx, dz};
}

std::vector<torch::Tensor> ba_extend(torch::Tensor poses, torch::Tensor disps, torch::Tensor intrinsics,
                                     torch::Tensor disps_sens, torch::Tensor targets, torch::Tensor weights,
                                     torch::Tensor eta, torch::Tensor ii, torch::Tensor jj, torch::Tensor H, torch::Tensor v,
                                     torch::Tensor A_prior, const int t0, const int t1, const int iterations, const float lm,
                                     const float ep, const bool motion_only, const bool skip_solve) {
  CHECK_INPUT(targets); CHECK_INPUT(weights); CHECK_INPUT(poses); CHECK_INPUT(disps); CHECK_INPUT(intrinsics);
  CHECK_INPUT(disps_sens); CHECK_INPUT(ii); CHECK_INPUT(jj); CHECK_INPUT(H); CHECK_INPUT(v); CHECK_INPUT(A_prior);
  Window w;
  w.setup(poses, disps, disps_sens, targets, weights, eta, ii, jj, t0, t1, !motion_only);
  torch::Tensor dx, dz, prior;
  if (iterations <= 0) return {dx, dz};
  DBAF_CALL(dbaf_ba_prepare(w.ws.data_ptr(), i64(ii, "ii"), i64(jj, "jj"), w.E, w.N, t0, t1, w.ht, w.wd, stream()));
  for (int it = 0; it < iterations; ++it) {
    DBAF_CALL(dbaf_ba_linearize(w.ws.data_ptr(), f32(poses, "poses"), f32(disps, "disps"), f32(intrinsics, "intrinsics"),
                                f32(disps_sens, "disps_sens"), f32(targets, "targets"), f32(weights, "weights"),
                                f32(w.eta, "eta"), w.eta_rows, i64(ii, "ii"), i64(jj, "jj"), w.E, w.N, t0, t1, w.ht, w.wd,
                                0.05f, motion_only ? 1 : 0, 1, 0, -1, stream()));
    if (!motion_only) {
      export_system(w, H, v);
      if (skip_solve) return {dx, dz};
      if (!prior.defined()) {
        TORCH_CHECK(A_prior.scalar_type() == torch::kFloat64 && A_prior.dim() == 2, "A_prior must be a 2-D float64 tensor");
        // a non-null pointer even for an empty prior: the reference then still solves with Ad + Adprior = 2 Ad (:1622-1644)
        prior = A_prior.numel() > 0 ? A_prior.to(poses.device()).contiguous()
                                    : torch::zeros({1}, torch::TensorOptions().dtype(torch::kFloat64).device(poses.device()));
      }
    }
    DBAF_CALL(dbaf_ba_solve(w.ws.data_ptr(), t0, t1, lm, ep, motion_only ? nullptr : prior.data_ptr<double>(),
                            motion_only ? 0 : (int)A_prior.size(0), motion_only ? 0 : (int)A_prior.size(1), stream()));
    dx = torch::empty({w.P, 6}, poses.options());
    if (!motion_only) dz = torch::empty({w.dz_rows(), w.HW}, poses.options());
    DBAF_CALL(dbaf_ba_retract(w.ws.data_ptr(), poses.data_ptr<float>(), disps.data_ptr<float>(), nullptr, i64(ii, "ii"),
                              i64(jj, "jj"), w.E, w.N, t0, t1, w.ht, w.wd, 1, motion_only ? 0 : 1, 0, -1,
                              dx.data_ptr<float>(), motion_only ? nullptr : dz.data_ptr<float>(), stream()));
    if (!motion_only) dz = dz.slice(0, 0, w.eta_rows > 1 ? w.eta_rows : w.dz_rows());
  }
  return {dx, dz};
}

torch::Tensor frame_distance(torch::Tensor poses, torch::Tensor disps, torch::Tensor intrinsics, torch::Tensor ii,
                             torch::Tensor jj, const float beta) {
  CHECK_INPUT(poses); CHECK_INPUT(disps); CHECK_INPUT(intrinsics); CHECK_INPUT(ii); CHECK_INPUT(jj);
  auto dist = torch::empty({ii.numel()}, poses.options());
  DBAF_CALL(dbaf_frame_distance(f32(poses, "poses"), f32(disps, "disps"), f32(intrinsics, "intrinsics"), i64(ii, "ii"),
                                i64(jj, "jj"), (int)ii.numel(), (int)disps.size(1), (int)disps.size(2), beta,
                                dist.data_ptr<float>(), stream()));
  return dist;
}

std::vector<torch::Tensor> projmap(torch::Tensor poses, torch::Tensor disps, torch::Tensor intrinsics, torch::Tensor ii,
                                   torch::Tensor jj) {
  CHECK_INPUT(poses); CHECK_INPUT(disps); CHECK_INPUT(intrinsics); CHECK_INPUT(ii); CHECK_INPUT(jj);
  const int64_t E = ii.numel(), ht = disps.size(1), wd = disps.size(2);
  auto coords = torch::empty({E, ht, wd, 3}, poses.options());
  auto valid = torch::empty({E, ht, wd, 1}, poses.options());
  DBAF_CALL(dbaf_projmap(f32(poses, "poses"), f32(disps, "disps"), f32(intrinsics, "intrinsics"), i64(ii, "ii"),
                         i64(jj, "jj"), (int)E, (int)ht, (int)wd, coords.data_ptr<float>(), valid.data_ptr<float>(), stream()));
  return {coords, valid};
}

torch::Tensor iproj(torch::Tensor poses, torch::Tensor disps, torch::Tensor intrinsics) {
  CHECK_INPUT(poses); CHECK_INPUT(disps); CHECK_INPUT(intrinsics);
  auto pts = torch::empty({disps.size(0), disps.size(1), disps.size(2), 3}, disps.options());
  DBAF_CALL(dbaf_iproj(f32(poses, "poses"), f32(disps, "disps"), f32(intrinsics, "intrinsics"), (int)disps.size(0),
                       (int)disps.size(1), (int)disps.size(2), pts.data_ptr<float>(), stream()));
  return pts;
}

torch::Tensor depth_filter(torch::Tensor poses, torch::Tensor disps, torch::Tensor intrinsics, torch::Tensor ix,
                           torch::Tensor thresh) {
  CHECK_INPUT(poses); CHECK_INPUT(disps); CHECK_INPUT(intrinsics); CHECK_INPUT(ix); CHECK_INPUT(thresh);
  auto counter = torch::empty({ix.numel(), disps.size(1), disps.size(2)}, disps.options());
  DBAF_CALL(dbaf_depth_filter(f32(poses, "poses"), f32(disps, "disps"), f32(intrinsics, "intrinsics"), i64(ix, "ix"),
                              f32(thresh, "thresh"), (int)ix.numel(), (int)disps.size(0), (int)disps.size(1),
                              (int)disps.size(2), counter.data_ptr<float>(), stream()));
  return counter;
}

static int corr_dtype(const torch::Tensor& t, const char* name) {
  TORCH_CHECK(t.is_cuda(), name, " must be a CUDA tensor");
  if (t.scalar_type() == torch::kFloat16) return DBAF_F16;
  TORCH_CHECK(t.scalar_type() == torch::kFloat32, name, ": only float16 and float32 are supported");
  return DBAF_F32;
}

std::vector<torch::Tensor> corr_index_forward(torch::Tensor volume, torch::Tensor coords, int radius) {
  CHECK_INPUT(volume); CHECK_INPUT(coords);
  const int dt = corr_dtype(volume, "volume");
  const int64_t n = volume.size(0), h1 = volume.size(1), w1 = volume.size(2), h2 = volume.size(3), w2 = volume.size(4);
  const int64_t rd = 2 * radius + 1;
  auto corr = torch::empty({n, rd, rd, h1, w1}, volume.options());
  DBAF_CALL(dbaf_corr_index_forward(volume.data_ptr(), dt, f32(coords, "coords"), (int)n, (int)h1, (int)w1, (int)h2, (int)w2,
                                    radius, corr.data_ptr(), stream()));
  return {corr};
}

std::vector<torch::Tensor> corr_index_backward(torch::Tensor volume, torch::Tensor coords, torch::Tensor corr_grad, int radius) {
  CHECK_INPUT(volume); CHECK_INPUT(coords); CHECK_INPUT(corr_grad);
  const int dt = corr_dtype(volume, "volume");
  TORCH_CHECK(corr_grad.scalar_type() == volume.scalar_type(), "corr_grad must have the dtype of volume");
  auto grad = torch::empty_like(volume);
  DBAF_CALL(dbaf_corr_index_backward(f32(coords, "coords"), corr_grad.data_ptr(), dt, (int)volume.size(0), (int)volume.size(1),
                                     (int)volume.size(2), (int)volume.size(3), (int)volume.size(4), radius, grad.data_ptr(),
                                     stream()));
  return {grad};
}

std::vector<torch::Tensor> altcorr_forward(torch::Tensor fmap1, torch::Tensor fmap2, torch::Tensor coords, int radius) {
  CHECK_INPUT(fmap1); CHECK_INPUT(fmap2); CHECK_INPUT(coords);
  const int64_t B = coords.size(0), S = coords.size(1), H1 = coords.size(2), W1 = coords.size(3);
  const int64_t H2 = fmap2.size(1), W2 = fmap2.size(2), C = fmap2.size(3), rd = 2 * radius + 1;
  auto corr = torch::empty({B, S, rd * rd, H1, W1}, fmap1.options());
  DBAF_CALL(dbaf_altcorr_forward(f32(fmap1, "fmap1"), f32(fmap2, "fmap2"), f32(coords, "coords"), (int)B, (int)S, (int)H1,
                                 (int)W1, (int)H2, (int)W2, (int)C, radius, corr.data_ptr<float>(), stream()));
  return {corr};
}

std::vector<torch::Tensor> altcorr_backward(torch::Tensor fmap1, torch::Tensor fmap2, torch::Tensor coords,
                                            torch::Tensor corr_grad, int radius) {
  CHECK_INPUT(fmap1); CHECK_INPUT(fmap2); CHECK_INPUT(coords); CHECK_INPUT(corr_grad);
  const int64_t B = coords.size(0), S = coords.size(1), H1 = coords.size(2), W1 = coords.size(3);
  const int64_t H2 = fmap2.size(1), W2 = fmap2.size(2), C = fmap2.size(3);
  auto g1 = torch::empty_like(fmap1), g2 = torch::empty_like(fmap2), gc = torch::empty_like(coords);
  DBAF_CALL(dbaf_altcorr_backward(f32(fmap1, "fmap1"), f32(fmap2, "fmap2"), f32(coords, "coords"), f32(corr_grad, "corr_grad"),
                                  (int)B, (int)S, (int)H1, (int)W1, (int)H2, (int)W2, (int)C, radius, g1.data_ptr<float>(),
                                  g2.data_ptr<float>(), gc.data_ptr<float>(), stream()));
  return {g1, g2, gc};
}

// src/bacore.h:4-70, src/droid_kernels.cu:1786-1956.  (In a namespace of its own and registered module-local: the
// reference extension, when loaded into the same process, registers a global C++ type of the same name.)
namespace dbaf_torch {
class BACore {
 public:
  void init(torch::Tensor _poses, torch::Tensor _disps, torch::Tensor _intrinsics, torch::Tensor _disps_sens,
            torch::Tensor _targets, torch::Tensor _weights, torch::Tensor _eta, torch::Tensor _ii, torch::Tensor _jj,
            const int t0, const int t1, const int /*iterations*/, const float lm_, const float ep_, const bool /*motion_only*/) {
    // `iterations` and `motion_only` are accepted and ignored, as in the reference (:1802-1814)
    poses = _poses; disps = _disps; intrinsics = _intrinsics; disps_sens = _disps_sens; targets = _targets;
    weights = _weights; ii = _ii; jj = _jj; lm = lm_; ep = ep_;
    CHECK_INPUT(targets); CHECK_INPUT(weights); CHECK_INPUT(poses); CHECK_INPUT(disps); CHECK_INPUT(intrinsics);
    CHECK_INPUT(disps_sens); CHECK_INPUT(ii); CHECK_INPUT(jj);
    w.setup(poses, disps, disps_sens, targets, weights, _eta, ii, jj, t0, t1, true);
    DBAF_CALL(dbaf_ba_prepare(w.ws.data_ptr(), i64(ii, "ii"), i64(jj, "jj"), w.E, w.N, t0, t1, w.ht, w.wd, stream()));
  }
  void hessian(torch::Tensor H, torch::Tensor v) {
    DBAF_CALL(dbaf_ba_linearize(w.ws.data_ptr(), f32(poses, "poses"), f32(disps, "disps"), f32(intrinsics, "intrinsics"),
                                f32(disps_sens, "disps_sens"), f32(targets, "targets"), f32(weights, "weights"),
                                f32(w.eta, "eta"), w.eta_rows, i64(ii, "ii"), i64(jj, "jj"), w.E, w.N, w.t0, w.t1, w.ht, w.wd,
                                0.001f /* src/droid_kernels.cu:1872 */, 0, 1, 0, -1, stream()));
    export_system(w, H, v);
  }
  void optimize(torch::Tensor H, torch::Tensor v) {
    const int64_t n = 6 * (int64_t)w.P;
    TORCH_CHECK(H.numel() == n * n && v.numel() == n, "optimize expects the full reduced system");
    auto hv = torch::cat({H.reshape({-1}).to(torch::kFloat64), v.reshape({-1}).to(torch::kFloat64)}).to(poses.device());
    C10_CUDA_CHECK(cudaMemcpyAsync(w.Hd(), hv.data_ptr<double>(), (size_t)(n * n + n) * 8, cudaMemcpyDeviceToDevice,
                                   at::cuda::getCurrentCUDAStream().stream()));
    DBAF_CALL(dbaf_ba_solve(w.ws.data_ptr(), w.t0, w.t1, lm, ep, nullptr, 0, 0, stream()));
  }
  std::vector<torch::Tensor> retract(torch::Tensor _dx) {
    TORCH_CHECK(!_dx.is_cuda() && _dx.scalar_type() == torch::kFloat64, "dx must be a float64 CPU tensor");
    auto dxd = _dx.reshape({-1}).slice(0, 0, 6 * (int64_t)w.P).to(torch::kFloat32).to(poses.device()).contiguous();
    dx = torch::empty({w.P, 6}, poses.options());
    dz = torch::empty({w.dz_rows(), w.HW}, poses.options());
    DBAF_CALL(dbaf_ba_retract(w.ws.data_ptr(), poses.data_ptr<float>(), disps.data_ptr<float>(), dxd.data_ptr<float>(),
                              i64(ii, "ii"), i64(jj, "jj"), w.E, w.N, w.t0, w.t1, w.ht, w.wd, 1, 1, 0, -1,
                              dx.data_ptr<float>(), dz.data_ptr<float>(), stream()));
    dz = dz.slice(0, 0, w.eta_rows > 1 ? w.eta_rows : w.dz_rows());
    return {dx, dz};
  }

 private:
  torch::Tensor poses, disps, intrinsics, disps_sens, targets, weights, ii, jj, dx, dz;
  float lm = 0.f, ep = 0.f;
  Window w;
};
}  // namespace dbaf_torch
using dbaf_torch::BACore;

PYBIND11_MODULE(TORCH_EXTENSION_NAME, m) {
  TORCH_CHECK(dbaf_ba_layout_bytes() == (int)sizeof(dbaf_ba_layout),
              "droid_backends_torch was compiled against another include/dbaf_b200.h than libdbaf_b200.so: rebuild it "
              "(python dba-fusion_b200/torch_ext/build.py --force)");
  m.def("ba", &ba, "bundle adjustment");
  m.def("ba_extend", &ba_extend, "bundle adjustment (extended)");
  m.def("frame_distance", &frame_distance, "frame_distance");
  m.def("projmap", &projmap, "projmap");
  m.def("depth_filter", &depth_filter, "depth_filter");
  m.def("iproj", &iproj, "back projection");
  m.def("altcorr_forward", &altcorr_forward, "ALTCORR forward");
  m.def("altcorr_backward", &altcorr_backward, "ALTCORR backward");
  m.def("corr_index_forward", &corr_index_forward, "INDEX forward");
  m.def("corr_index_backward", &corr_index_backward, "INDEX backward");
  py::class_<BACore>(m, "BACore", py::module_local()).def(py::init<>())
      .def("init", &BACore::init)
      .def("hessian", &BACore::hessian)
      .def("optimize", &BACore::optimize)
      .def("retract", &BACore::retract);
}
