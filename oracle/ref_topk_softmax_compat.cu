// TEST INFRASTRUCTURE ONLY (oracle).  Compiles the reference's csrc/moe/topk_softmax_kernels.cu
// where it lies.  The only adaptation: the reference spells the int64 scalar type
// `at::ScalarType::Int64` (topk_softmax_kernels.cu:544), which torch 2.11 names `Long`; the
// headers are included first so the macro cannot touch them.
#include <torch/all.h>
#include <ATen/cuda/CUDAContext.h>
#include <c10/cuda/CUDAGuard.h>
#include <cub/util_type.cuh>
#include <cub/cub.cuh>
#define Int64 Long
#include "moe/topk_softmax_kernels.cu"
