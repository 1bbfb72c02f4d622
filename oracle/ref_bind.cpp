// Binding for the REFERENCE's own CUDA PQ (AQLM) kernels, compiled unmodified from
// /root/reference/csrc/quantization/aqlm/gemm_kernels.cu by oracle/Makefile.
// TEST INFRASTRUCTURE ONLY (the checker and the same-box "kernel to beat"): nothing in the
// product package links or loads this.  Schemas are the reference's, torch_bindings.cpp:161-173;
// prototypes are the reference's, csrc/ops.h:122-131.  Registered under the namespace `pqref`
// so that it can never shadow the product's `pq::` ops.
#include <torch/all.h>
#include <torch/library.h>

#include <optional>
#include <vector>

torch::Tensor aqlm_gemm(const torch::Tensor& input, const torch::Tensor& codes,
                        const torch::Tensor& codebooks, const torch::Tensor& scales,
                        const std::vector<int64_t>& codebook_partition_sizes,
                        const std::optional<torch::Tensor>& bias);

torch::Tensor aqlm_dequant(const torch::Tensor& codes, const torch::Tensor& codebooks,
                           const std::vector<int64_t>& codebook_partition_sizes);

TORCH_LIBRARY(pqref, ops) {
  ops.def(
      "aqlm_gemm(Tensor input, Tensor codes, Tensor codebooks, "
      "Tensor scales, int[] codebook_partition_sizes, Tensor? bias) -> Tensor");
  ops.impl("aqlm_gemm", torch::kCUDA, &aqlm_gemm);
  ops.def(
      "aqlm_dequant(Tensor codes, Tensor codebooks, "
      "int[] codebook_partition_sizes) -> Tensor");
  ops.impl("aqlm_dequant", torch::kCUDA, &aqlm_dequant);
}
