// Oracle build aid only (NOT product code). Force-included when compiling the
// unmodified reference gn_kernels.cu against torch >= 2.11, which no longer
// ships <torch/linalg.h>: the reference calls torch::linalg::{linalg_norm,norm}
// only for its convergence test (gn_kernels.cu:986,1927,2284,2693).
#pragma once
#include <torch/extension.h>
#include <optional>
namespace torch { namespace linalg {
inline at::Tensor linalg_norm(const at::Tensor& self, const std::optional<c10::Scalar>& ord,
                              at::OptionalIntArrayRef dim, bool keepdim,
                              std::optional<at::ScalarType> dtype) {
  return at::linalg_norm(self, ord, dim, keepdim, dtype);
}
inline at::Tensor norm(const at::Tensor& self, const std::optional<c10::Scalar>& ord,
                       at::OptionalIntArrayRef dim, bool keepdim,
                       std::optional<at::ScalarType> dtype) {
  return at::linalg_norm(self, ord, dim, keepdim, dtype);
}
}}
