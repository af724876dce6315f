#include <torch/all.h>
#include <torch/library.h>
namespace {
torch::Tensor f1(torch::Tensor& a, int64_t const& id, int64_t size_m, bool flag) {
  TORCH_CHECK(a.dim() == 2 && a.size(0) == size_m, "Shape mismatch: a.size(0) = ", a.size(0), ", size_m = ", size_m);
  torch::Tensor c = torch::empty({size_m, 4}, a.options());
  return c;
}
}
TORCH_LIBRARY_FRAGMENT(_mini, ops) {
  ops.def("f1(Tensor a, int id, SymInt size_m, bool flag) -> Tensor", {at::Tag::needs_fixed_stride_order});
}
TORCH_LIBRARY_IMPL(_mini, CPU, ops) { ops.impl("f1", &f1); }
