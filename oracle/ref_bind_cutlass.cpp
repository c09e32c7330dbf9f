// TEST INFRASTRUCTURE ONLY (oracle).  Binding shim that exposes the reference's own Blackwell
// CUTLASS kernels (compiled for sm_100a from the sources where they lie under
// /root/reference/csrc/quantization/cutlass_w8a8 -- nothing is copied) under the private torch
// library name `_ref_cutlass`.  The arch-dispatch layer (scaled_mm_entry.cu:169-240) is bypassed:
// the sm100 entry points it would select are bound directly.
// Schemas: /root/reference/csrc/torch_bindings.cpp:422-426 (cutlass_scaled_mm), :450-456
// (cutlass_moe_mm), :465-471 (get_cutlass_moe_mm_data).
#include <torch/all.h>
#include <torch/library.h>

#include <optional>

namespace vllm {
// csrc/quantization/cutlass_w8a8/c3x/scaled_mm_kernels.hpp:34-38
void cutlass_scaled_mm_sm100_fp8(torch::Tensor& out, torch::Tensor const& a, torch::Tensor const& b,
                                 torch::Tensor const& a_scales, torch::Tensor const& b_scales,
                                 std::optional<torch::Tensor> const& bias);
}  // namespace vllm
// csrc/quantization/cutlass_w8a8/moe/grouped_mm_c3x_sm100.cu:130-140
void cutlass_moe_mm_sm100(torch::Tensor& out_tensors, torch::Tensor const& a_tensors,
                          torch::Tensor const& b_tensors, torch::Tensor const& a_scales,
                          torch::Tensor const& b_scales, torch::Tensor const& expert_offsets,
                          torch::Tensor const& problem_sizes, torch::Tensor const& a_strides,
                          torch::Tensor const& b_strides, torch::Tensor const& c_strides,
                          bool per_act_token, bool per_out_ch);
// csrc/quantization/cutlass_w8a8/moe/moe_data.cu:107-116
void get_cutlass_moe_mm_data_caller(const torch::Tensor& topk_ids, torch::Tensor& expert_offsets,
                                    torch::Tensor& problem_sizes1, torch::Tensor& problem_sizes2,
                                    torch::Tensor& input_permutation,
                                    torch::Tensor& output_permutation, const int64_t num_experts,
                                    const int64_t n, const int64_t k,
                                    const std::optional<torch::Tensor>& blockscale_offsets);

static void ref_cutlass_scaled_mm(torch::Tensor& out, torch::Tensor const& a, torch::Tensor const& b,
                                  torch::Tensor const& a_scales, torch::Tensor const& b_scales,
                                  std::optional<torch::Tensor> const& bias) {
  vllm::cutlass_scaled_mm_sm100_fp8(out, a, b, a_scales, b_scales, bias);
}

TORCH_LIBRARY(_ref_cutlass, m) {
  m.def(
      "cutlass_scaled_mm(Tensor! out, Tensor a, Tensor b, Tensor a_scales, Tensor b_scales,"
      " Tensor? bias) -> ()");
  m.impl("cutlass_scaled_mm", torch::kCUDA, &ref_cutlass_scaled_mm);
  m.def(
      "cutlass_moe_mm(Tensor! out_tensors, Tensor a_tensors, Tensor b_tensors, Tensor a_scales,"
      " Tensor b_scales, Tensor expert_offsets, Tensor problem_sizes, Tensor a_strides,"
      " Tensor b_strides, Tensor c_strides, bool per_act_token, bool per_out_ch) -> ()");
  m.impl("cutlass_moe_mm", torch::kCUDA, &cutlass_moe_mm_sm100);
  m.def(
      "get_cutlass_moe_mm_data(Tensor topk_ids, Tensor! expert_offsets, Tensor! problem_sizes1,"
      " Tensor! problem_sizes2, Tensor! input_permutation, Tensor! output_permutation,"
      " int num_experts, int n, int k, Tensor? blockscale_offsets) -> ()");
  m.impl("get_cutlass_moe_mm_data", torch::kCUDA, &get_cutlass_moe_mm_data_caller);
}
