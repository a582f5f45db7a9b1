// Test infrastructure (oracle/): a stand-in for the un-vendored third-party header
// "pytorch_extension_utils.h" (flashinfer csrc, fetched by the reference's cmake build; absent from
// /root/reference), defining only the four argument-check macros that
// python/sglang/kernels/aot/csrc/attention/merge_attn_states.cu:186-205 uses, so that the reference kernel
// compiles from the source where it lies.  Written for this repo; not a copy of the original header.
#pragma once
#include <torch/all.h>

#define CHECK_INPUT(x)                                                  \
  TORCH_CHECK((x).is_cuda(), #x " must be a CUDA tensor");              \
  TORCH_CHECK((x).is_contiguous(), #x " must be contiguous")
#define CHECK_EQ(a, b) TORCH_CHECK((a) == (b), "CHECK_EQ(" #a ", " #b ") failed")
#define CHECK_DIM(d, x) TORCH_CHECK((x).dim() == (d), #x " must be a " #d "D tensor")
#define CHECK_SHAPE(a, b) TORCH_CHECK((a).sizes() == (b).sizes(), #a " and " #b " must have the same shape")
