// Host <-> device staging of a step's inputs / result by a KERNEL instead of copy-engine memcpy nodes.
//
// A decode step driven from host buffers inside one CUDA graph (SGLang's decode runner,
// decode_cuda_graph_runner.py:1148-1363: static buffers refilled from the scheduler's pinned tensors, then replay)
// holds two H2D memcpy nodes, four kernels and one D2H memcpy node.  Every memcpy <-> kernel edge is a copy-engine /
// SM semaphore hand-over, and which engine a node lands on is decided at instantiation: the same graph came out at
// 225 us in one process and 275 us in the next.  Pinned host memory is mapped into the device address space (UVA), so
// a kernel can do the copy itself: the graph is then kernel nodes only, in one queue.
//
// Loads from the host side bypass the caches (ld.volatile); 16 bytes per thread, all loads of a thread issued before
// the first store (PCIe read latency is microseconds: bytes in flight are what make the bandwidth).
#include "common.cuh"

namespace b200attn {

constexpr int kStageThreads = 256;
constexpr int kStageUnroll = 4;

__device__ __forceinline__ uint4 ld_stage(const uint4* p) {
  uint4 v;
  asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(p));
  return v;
}

__global__ void __launch_bounds__(kStageThreads)
stage_copy_kernel(uint4* __restrict__ d0, const uint4* __restrict__ s0, int64_t n0, uint4* __restrict__ d1,
                  const uint4* __restrict__ s1, int64_t n1) {
  pdl_wait();
  pdl_trigger();
  const int64_t stride = (int64_t)gridDim.x * kStageThreads;
  const int64_t n = n0 + n1;  // one index space over both segments (16-byte units)
  for (int64_t base = (int64_t)blockIdx.x * kStageThreads + threadIdx.x; base < n; base += stride * kStageUnroll) {
    uint4 v[kStageUnroll];
#pragma unroll
    for (int u = 0; u < kStageUnroll; ++u) {
      const int64_t i = base + u * stride;
      if (i < n) v[u] = i < n0 ? ld_stage(s0 + i) : ld_stage(s1 + (i - n0));
    }
#pragma unroll
    for (int u = 0; u < kStageUnroll; ++u) {
      const int64_t i = base + u * stride;
      if (i < n) {
        if (i < n0) d0[i] = v[u];
        else d1[i - n0] = v[u];
      }
    }
  }
}

}  // namespace b200attn

extern "C" int b200attn_stage_copy(void* dst0, const void* src0, int64_t bytes0, void* dst1, const void* src1,
                                   int64_t bytes1, void* stream) {
  using namespace b200attn;
  B2_CHECK_ARG(bytes0 >= 0 && bytes1 >= 0, "stage_copy: negative size");
  if (bytes0 + bytes1 == 0) return B200ATTN_OK;
  B2_CHECK_ARG((bytes0 == 0 || (dst0 && src0)) && (bytes1 == 0 || (dst1 && src1)), "stage_copy: null pointer");
  B2_CHECK_ARG(bytes0 % 16 == 0 && bytes1 % 16 == 0, "stage_copy: sizes must be multiples of 16 bytes");
  B2_CHECK_ARG((((uintptr_t)dst0 | (uintptr_t)src0 | (uintptr_t)dst1 | (uintptr_t)src1) & 15) == 0,
               "stage_copy: buffers must be 16-byte aligned");
  const int64_t n = (bytes0 + bytes1) / 16;
  int64_t blocks = (n + kStageThreads * kStageUnroll - 1) / (kStageThreads * kStageUnroll);
  const int cap = 2 * sm_count();
  if (blocks > cap) blocks = cap;
  B2_CUDA(launch_pdl(stage_copy_kernel, dim3((unsigned)blocks), dim3(kStageThreads), 0, (cudaStream_t)stream,
                     reinterpret_cast<uint4*>(dst0), reinterpret_cast<const uint4*>(src0), bytes0 / 16,
                     reinterpret_cast<uint4*>(dst1), reinterpret_cast<const uint4*>(src1), bytes1 / 16));
  return B200ATTN_OK;
}
