// Thin torch C++-extension front end over the C ABI (north_star: "a thin torch C++-extension C-ABI that keeps the
// existing nn.Module forward signatures").  It does what ops.py does through ctypes -- checks, output allocation,
// current stream -- in C++, for the operators the reference's eager LowCNN.forward calls every step
// (models/LocalCostVolume/baseline.py:114-134), so the host cost per call drops from ~13 us (Python + ctypes) to ~3 us.
// Every function forwards to an `extern "C"` entry of libsfb200.so (include/stereoformer_b200.h); there is no compute here.
#include <c10/cuda/CUDAGuard.h>
#include <c10/cuda/CUDAStream.h>
#include <torch/extension.h>

#include "../../include/stereoformer_b200.h"

namespace {

void chk(const at::Tensor& t, const char* name, at::ScalarType dt = at::kFloat) {
  TORCH_CHECK(t.is_cuda(), "Not implemented on the CPU (", name, " is a CPU tensor; stereoformer_b200 has no CPU path)");
  TORCH_CHECK(t.scalar_type() == dt, name, " must be ", c10::toString(dt));
  TORCH_CHECK(t.is_contiguous(), name, " must be contiguous");
}

void rc_check(int rc, const char* fn) {
  if (rc == 0) return;
  const char* msg = sf_last_error();
  TORCH_CHECK_VALUE(rc != 1, fn, ": ", msg);
  TORCH_CHECK(false, fn, " failed (code ", rc, "): ", msg);
}

void* cur_stream() { return static_cast<void*>(c10::cuda::getCurrentCUDAStream().stream()); }

at::Tensor corr_volume(const at::Tensor& left, const at::Tensor& right, int64_t max_disp, int64_t groups, int64_t algo,
                       c10::optional<at::Tensor> out_) {
  chk(left, "left_feature"); chk(right, "right_feature");
  TORCH_CHECK_VALUE(left.dim() == 4 && left.sizes() == right.sizes(), "feature shapes differ or are not 4-D");
  const c10::cuda::CUDAGuard guard(left.device());
  const int64_t B = left.size(0), C = left.size(1), H = left.size(2), W = left.size(3);
  at::Tensor out = out_.has_value() ? *out_
                                    : (groups == 1 ? at::empty({B, max_disp, H, W}, left.options())
                                                   : at::empty({B, groups, max_disp, H, W}, left.options()));
  if (out_.has_value()) {
    chk(out, "out");
    TORCH_CHECK_VALUE(out.numel() == B * groups * max_disp * H * W, "out has the wrong size");
  }
  rc_check(sf_corr_volume_f32(left.data_ptr<float>(), right.data_ptr<float>(), out.data_ptr<float>(), (int)B, (int)C,
                              (int)H, (int)W, (int)max_disp, (int)groups, (int)algo, cur_stream()), "sf_corr_volume_f32");
  return out;
}

at::Tensor soft_argmin(const at::Tensor& cv, bool match_similarity) {
  chk(cv, "cost_volume");
  TORCH_CHECK_VALUE(cv.dim() == 4, "cost_volume must be 4-D");
  const c10::cuda::CUDAGuard guard(cv.device());
  at::Tensor out = at::empty({cv.size(0), cv.size(2), cv.size(3)}, cv.options());
  rc_check(sf_soft_argmin_f32(cv.data_ptr<float>(), out.data_ptr<float>(), nullptr, (int)cv.size(0), (int)cv.size(1),
                              (int)cv.size(2), (int)cv.size(3), match_similarity ? 1 : 0, cur_stream()), "sf_soft_argmin_f32");
  return out;
}

at::Tensor local_window(const at::Tensor& cv, c10::optional<at::Tensor> disp, c10::optional<at::Tensor> lo,
                        c10::optional<at::Tensor> up, int64_t mode, double radius, double gamma, int64_t n_points,
                        bool consider_valid, int64_t div_mode, c10::optional<at::Tensor> out_) {
  chk(cv, "old_cost_volume");
  TORCH_CHECK_VALUE(cv.dim() == 4, "old_cost_volume must be 4-D");
  const int64_t B = cv.size(0), D = cv.size(1), H = cv.size(2), W = cv.size(3);
  const float *dp = nullptr, *lp = nullptr, *upp = nullptr;
  if (disp.has_value()) { chk(*disp, "cur_disparity"); TORCH_CHECK_VALUE(disp->numel() == B * H * W, "cur_disparity must have B*H*W elements"); dp = disp->data_ptr<float>(); }
  if (lo.has_value()) {
    chk(*lo, "lower bound");
    TORCH_CHECK_VALUE(lo->numel() == (mode == SF_WIN_HEAD ? 2 : 1) * B * H * W, "lower bound has the wrong number of elements");
    lp = lo->data_ptr<float>();
  }
  if (up.has_value()) { chk(*up, "upper bound"); TORCH_CHECK_VALUE(up->numel() == B * H * W, "upper bound must have B*H*W elements"); upp = up->data_ptr<float>(); }
  const c10::cuda::CUDAGuard guard(cv.device());
  at::Tensor out = out_.has_value() ? *out_ : at::empty({B, 1, H, W}, cv.options());
  if (out_.has_value()) { chk(out, "out"); TORCH_CHECK_VALUE(out.numel() == B * H * W, "out must have B*H*W elements"); }
  rc_check(sf_local_window_f32(cv.data_ptr<float>(), dp, lp, upp, out.data_ptr<float>(), nullptr, nullptr, nullptr, nullptr,
                               nullptr, (int)B, (int)D, (int)H, (int)W, (int)mode, (float)radius, (float)gamma,
                               (int)n_points, consider_valid ? 1 : 0, (int)div_mode, cur_stream()), "sf_local_window_f32");
  return out;
}

at::Tensor upsample_convex(const at::Tensor& disp, const at::Tensor& mask, int64_t factor, double mask_scale) {
  chk(disp, "disp"); chk(mask, "mask");
  TORCH_CHECK_VALUE(disp.dim() == 4 && mask.dim() == 4, "disp / mask must be 4-D");
  const int64_t N = disp.size(0), H = disp.size(2), W = disp.size(3);
  TORCH_CHECK_VALUE(mask.size(0) == N && mask.size(1) == 9 * factor * factor && mask.size(2) == H && mask.size(3) == W,
                    "mask must be [N, 9*F*F, H, W]");
  const c10::cuda::CUDAGuard guard(disp.device());
  at::Tensor out = at::empty({N, 1, H * factor, W * factor}, disp.options());
  rc_check(sf_upsample_convex_scaled_f32(disp.data_ptr<float>(), mask.data_ptr<float>(), (float)mask_scale,
                                         out.data_ptr<float>(), (int)N, (int)H, (int)W, (int)factor, cur_stream()),
           "sf_upsample_convex_scaled_f32");
  return out;
}

}  // namespace

PYBIND11_MODULE(TORCH_EXTENSION_NAME, m) {
  m.doc() = "stereoformer_b200: torch C++ front end over the C ABI of libsfb200.so";
  m.def("corr_volume", &corr_volume, py::arg("left"), py::arg("right"), py::arg("max_disp"), py::arg("groups") = 1,
        py::arg("algo") = 0, py::arg("out") = py::none());
  m.def("soft_argmin", &soft_argmin, py::arg("cv"), py::arg("match_similarity") = true);
  m.def("local_window", &local_window, py::arg("cv"), py::arg("disp") = py::none(), py::arg("lo") = py::none(),
        py::arg("up") = py::none(), py::arg("mode") = 0, py::arg("radius") = 0.0, py::arg("gamma") = 1.0,
        py::arg("n_points") = 10, py::arg("consider_valid") = false, py::arg("div_mode") = 0, py::arg("out") = py::none());
  m.def("upsample_convex", &upsample_convex, py::arg("disp"), py::arg("mask"), py::arg("factor") = 8,
        py::arg("mask_scale") = 1.0);
}
