// Test infrastructure (oracle/): exposes the reference's merge_state_v2
// (python/sglang/kernels/aot/csrc/attention/merge_attn_states.cu:186-206, compiled from where it lies by
// oracle/make_ref.sh) as torch.ops.b200ref.merge_state_v2, the way the reference registers it as
// torch.ops.sgl_kernel.merge_state_v2 (kernels/aot/csrc/common_extension.cc).
#include <torch/all.h>
#include <torch/library.h>

void merge_state_v2(at::Tensor v_a, at::Tensor s_a, at::Tensor v_b, at::Tensor s_b, at::Tensor v_merged,
                    at::Tensor s_merged);

TORCH_LIBRARY(b200ref, m) {
  m.def("merge_state_v2(Tensor v_a, Tensor s_a, Tensor v_b, Tensor s_b, Tensor! v_merged, Tensor! s_merged) -> ()");
  m.impl("merge_state_v2", torch::kCUDA, &merge_state_v2);
}
