// Sum all-reduce of the o_proj output over NVLink SHARP (NVLS): the one collective on the attention path
// (RowParallelLinear after RadixAttention -> GroupCoordinator.all_reduce,
//  python/sglang/srt/distributed/parallel_state.py:648-757; the reference's symmetric-memory communicator is
//  device_communicators/torch_symm_mem.py:34-200: copy into the symmetric buffer, torch's multimem op, copy out).
//
// ONE launch per rank does all of it.  The payload lives in a symmetric allocation (every rank maps every peer's
// instance, and one MULTICAST address maps all of them: a multimem.ld_reduce through it returns the sum computed
// inside the NVSwitch, a multimem.st through it writes every rank's instance).  The payload is cut into one chunk
// per CTA; CTA b of every rank owns chunk b in all three phases, so the only cross-GPU synchronisation is between
// the CTAs b of the ranks:
//   A  copy the local input's chunk b into the local instance                       fence.sys, barrier_b
//   B  rank r: ld_reduce the r-th 1/N of chunk b through the switch, multimem.st it  fence.sys, barrier_b
//   C  copy chunk b of the local instance to the local output
// Two-shot (reduce-scatter + all-gather in the switch): every byte crosses each NVLink once in each direction.
// Messages up to kArSmallMax bytes take the ONE-shot form instead (one barrier, no multimem.st, no copy-out pass):
//   A' copy the local input's 32 KB piece b into piece b of one HALF of a small dedicated region   fence.sys, barrier_b
//   B' ld_reduce piece b through the switch straight into the local output
// The two halves alternate per call (a per-CTA call counter in the flag block picks it, so a replayed CUDA graph
// alternates too): a peer may still be reading piece b of the previous call while this call stages its input.  Why
// that is enough: a rank's kernels run in stream order, so a peer that has signalled barrier_b of call k has finished
// every read of call k-1; this CTA waited for that signal in call k before its next write to the same half in k+1.
// barrier_b: slot [b][src] in each rank's flag block holds the number of the last barrier src's CTA b has reached
// (see ar_barrier).  All CTAs of all ranks must be co-resident: grid <= 128.
#include "common.cuh"

namespace b200attn {

constexpr int kArMaxBlocks = 128;
constexpr int kArMaxWorld = 16;
constexpr int kArThreads = 512;
constexpr int kArUnroll = 8;  // 16-byte accesses in flight per thread
constexpr int64_t kArPiece = 32768;                   // one-shot: bytes per CTA, fixed (piece b is always CTA b's)
constexpr int64_t kArSmallHalf = kArMaxBlocks * kArPiece;  // 4 MB
constexpr int64_t kArSmallMax = 1 << 20;              // largest message that takes the one-shot form

#ifdef AR_TRACE
// dev-only (libb200attn_trace.so, tools/allreduce_nvls_check.py --trace): globaltimer stamps per CTA and phase
__device__ long long g_ar_trace[kArMaxBlocks * 8];
#define ATRACE(slot)                                                     \
  if (threadIdx.x == 0) {                                                \
    long long gt_;                                                       \
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gt_));              \
    g_ar_trace[blockIdx.x * 8 + (slot)] = gt_;                           \
  }
#else
#define ATRACE(slot)
#endif

struct ArParams {
  const uint4* in;    // local input or null (the caller wrote the payload itself, e.g. as a GEMM output)
  uint4* out;         // local output or null (the result is read from the payload)
  uint4* local;       // this rank's payload instance
  uint4* mc;          // the multicast address of the payload
  uint4* small_local; // one-shot region (two halves of kArSmallHalf bytes), this rank's instance / multicast address
  uint4* small_mc;
  uint32_t* flags[kArMaxWorld];  // flags[r]: rank r's flag block: [kArMaxBlocks][kArMaxWorld] slots, [kArMaxBlocks] counters
  int rank, world, dtype;
  int64_t n16;        // payload size in 16-byte units
};

// barrier_b without a remote round trip: the signal is a POSTED release store of this CTA's barrier number into the
// peer's slot, the wait polls the local slot until it reaches that number.  The number of CTA b's last barrier lives
// in the local flag block (after the slots); it is read ONCE when the kernel starts (ar_epoch_begin, overlapped with
// the staging copy) and written back when it ends; every rank runs the same sequence of calls with the same grids,
// so the CTAs b of all ranks count in step.
__device__ __forceinline__ void ar_epoch_begin(const ArParams& p, uint32_t* s_epoch) {
  if (threadIdx.x == 0) *s_epoch = p.flags[p.rank][kArMaxBlocks * kArMaxWorld + blockIdx.x];
}
__device__ __forceinline__ void ar_epoch_end(const ArParams& p, const uint32_t* s_epoch) {
  if (threadIdx.x == 0) p.flags[p.rank][kArMaxBlocks * kArMaxWorld + blockIdx.x] = *s_epoch;
}
__device__ __forceinline__ void ar_barrier(const ArParams& p, uint32_t* s_epoch) {
  __threadfence_system();
  __syncthreads();  // also orders ar_epoch_begin's shared store before the reads below
  const int t = threadIdx.x;
  const uint32_t e = *s_epoch + 1u;
  if (t < p.world) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p.flags[t] + (blockIdx.x * kArMaxWorld + p.rank)), "r"(e)
                 : "memory");
    const uint32_t* mine = p.flags[p.rank] + (blockIdx.x * kArMaxWorld + t);
    uint32_t v;
    do {
      asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(mine) : "memory");
    } while ((int32_t)(v - e) < 0);
  }
  __syncthreads();
  if (t == 0) *s_epoch = e;
}

template <int kDtype>
__device__ __forceinline__ uint4 mc_ld_reduce(const uint4* addr) {
  uint4 v;
  if constexpr (kDtype == B200ATTN_BF16) {
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0, %1, %2, %3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(addr) : "memory");
  } else if constexpr (kDtype == B200ATTN_FP16) {
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.f16x2 {%0, %1, %2, %3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(addr) : "memory");
  } else {
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(addr) : "memory");
  }
  return v;
}
__device__ __forceinline__ void mc_st(uint4* addr, const uint4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "r"(v.x), "r"(v.y),
               "r"(v.z), "r"(v.w) : "memory");
}

__device__ __forceinline__ void ar_copy(uint4* dst, const uint4* src, int64_t c0, int64_t c1) {
  for (int64_t i = c0 + threadIdx.x; i < c1; i += kArUnroll * kArThreads) {
    uint4 v[kArUnroll];
#pragma unroll
    for (int u = 0; u < kArUnroll; ++u)
      if (i + u * kArThreads < c1) v[u] = src[i + u * kArThreads];
#pragma unroll
    for (int u = 0; u < kArUnroll; ++u)
      if (i + u * kArThreads < c1) dst[i + u * kArThreads] = v[u];
  }
}

template <int kDtype>
__global__ void __launch_bounds__(kArThreads) allreduce_nvls_kernel(const ArParams p) {
  __shared__ uint32_t s_epoch;
  pdl_wait();
  pdl_trigger();
  ar_epoch_begin(p, &s_epoch);
  const int64_t per_block = (p.n16 + gridDim.x - 1) / gridDim.x;
  const int64_t c0 = (int64_t)blockIdx.x * per_block;
  const int64_t c1 = c0 + per_block < p.n16 ? c0 + per_block : p.n16;
  // A: stage the input
  ATRACE(0);
  if (p.in != nullptr) ar_copy(p.local, p.in, c0, c1);
  ATRACE(1);
  ar_barrier(p, &s_epoch);
  ATRACE(2);
  // B: this rank's share of the chunk, reduced in the switch and written to every instance
  {
    const int64_t n = c1 > c0 ? c1 - c0 : 0;
    const int64_t per_rank = (n + p.world - 1) / p.world;
    const int64_t s0 = c0 + p.rank * per_rank;
    const int64_t s1 = s0 + per_rank < c1 ? s0 + per_rank : c1;
    // kArUnroll reductions in flight per thread before the first store (an NVSwitch round trip is ~2 us)
    for (int64_t i = s0 + threadIdx.x; i < s1; i += kArUnroll * kArThreads) {
      uint4 v[kArUnroll];
#pragma unroll
      for (int u = 0; u < kArUnroll; ++u)
        if (i + u * kArThreads < s1) v[u] = mc_ld_reduce<kDtype>(p.mc + i + u * kArThreads);
#pragma unroll
      for (int u = 0; u < kArUnroll; ++u)
        if (i + u * kArThreads < s1) mc_st(p.mc + i + u * kArThreads, v[u]);
    }
  }
  ATRACE(3);
  ar_barrier(p, &s_epoch);
  ar_epoch_end(p, &s_epoch);
  ATRACE(4);
  // C: hand the result over
  if (p.out != nullptr) ar_copy(p.out, p.local, c0, c1);
  ATRACE(5);
}

template <int kDtype>
__global__ void __launch_bounds__(kArThreads) allreduce_nvls_oneshot_kernel(const ArParams p) {
  __shared__ uint32_t s_half, s_epoch;
  pdl_wait();
  pdl_trigger();
  ar_epoch_begin(p, &s_epoch);
  if (threadIdx.x == 0) {
    uint32_t* calls = p.flags[p.rank] + kArMaxBlocks * kArMaxWorld + kArMaxBlocks + blockIdx.x;
    const uint32_t k = *calls;
    *calls = k + 1u;
    s_half = k & 1u;
  }
  __syncthreads();
  const int64_t units = kArPiece / 16;
  const int64_t c0 = (int64_t)blockIdx.x * units;
  const int64_t c1 = c0 + units < p.n16 ? c0 + units : p.n16;
  const int64_t off = (int64_t)s_half * (kArSmallHalf / 16);
  ar_copy(p.small_local + off, p.in, c0, c1);  // same index in the input and in the half
  ar_barrier(p, &s_epoch);
  ar_epoch_end(p, &s_epoch);
  for (int64_t i = c0 + threadIdx.x; i < c1; i += kArUnroll * kArThreads) {
    uint4 v[kArUnroll];
#pragma unroll
    for (int u = 0; u < kArUnroll; ++u)
      if (i + u * kArThreads < c1) v[u] = mc_ld_reduce<kDtype>(p.small_mc + off + i + u * kArThreads);
#pragma unroll
    for (int u = 0; u < kArUnroll; ++u)
      if (i + u * kArThreads < c1) p.out[i + u * kArThreads] = v[u];
  }
}

}  // namespace b200attn

extern "C" int64_t b200attn_allreduce_small_bytes(void) { return 2 * b200attn::kArSmallHalf; }

extern "C" int64_t b200attn_allreduce_flag_bytes(void) {
  return 16384;  // [kArMaxBlocks][kArMaxWorld] slots + 2 x [kArMaxBlocks] counters (barriers, one-shot calls), rounded up
}

extern "C" int b200attn_allreduce_nvls(const void* in, void* out, void* local_payload, void* mc_payload,
                                       void* local_small, void* mc_small, void* const* flag_ptrs, int32_t rank,
                                       int32_t world, int64_t bytes, int dtype, void* stream) {
  using namespace b200attn;
  B2_CHECK_ARG(bytes >= 0, "allreduce: negative size");
  B2_CHECK_ARG(world >= 1 && world <= kArMaxWorld && rank >= 0 && rank < world, "allreduce: bad rank %d / world %d",
               rank, world);
  if (bytes == 0) return B200ATTN_OK;
  B2_CHECK_ARG(local_payload && mc_payload && flag_ptrs, "allreduce: null pointer");
  B2_CHECK_ARG(bytes % 16 == 0, "allreduce: size must be a multiple of 16 bytes");
  B2_CHECK_ARG((((uintptr_t)in | (uintptr_t)out | (uintptr_t)local_payload | (uintptr_t)mc_payload) & 15) == 0,
               "allreduce: buffers must be 16-byte aligned");
  if (dtype != B200ATTN_BF16 && dtype != B200ATTN_FP16 && dtype != B200ATTN_FP32)
    B2_UNSUPPORTED("allreduce: dtype %d", dtype);
  ArParams p;
  p.in = reinterpret_cast<const uint4*>(in);
  p.out = reinterpret_cast<uint4*>(out);
  p.local = reinterpret_cast<uint4*>(local_payload);
  p.mc = reinterpret_cast<uint4*>(mc_payload);
  p.small_local = reinterpret_cast<uint4*>(local_small);
  p.small_mc = reinterpret_cast<uint4*>(mc_small);
  for (int r = 0; r < kArMaxWorld; ++r) p.flags[r] = r < world ? reinterpret_cast<uint32_t*>(flag_ptrs[r]) : nullptr;
  for (int r = 0; r < world; ++r) B2_CHECK_ARG(p.flags[r] != nullptr, "allreduce: null flag block of rank %d", r);
  p.rank = rank;
  p.world = world;
  p.dtype = dtype;
  p.n16 = bytes / 16;
  cudaStream_t st = (cudaStream_t)stream;
  // one-shot limit: B200ATTN_AR_ONESHOT_MAX (bytes, read once; 0 forces the two-shot form) overrides the default
  static const int64_t oneshot_max = [] {
    const char* e = getenv("B200ATTN_AR_ONESHOT_MAX");
    return e != nullptr ? (int64_t)atoll(e) : kArSmallMax;
  }();
  if (in != nullptr && out != nullptr && local_small != nullptr && mc_small != nullptr && bytes <= oneshot_max &&
      bytes <= kArSmallMax && bytes * world <= 4 * kArSmallMax) {  // every rank pulls the whole message through the switch: small ones only
    B2_CHECK_ARG((((uintptr_t)local_small | (uintptr_t)mc_small) & 15) == 0, "allreduce: small region alignment");
    const dim3 g1((unsigned)((bytes + kArPiece - 1) / kArPiece)), b1(kArThreads);
    if (dtype == B200ATTN_BF16) {
      B2_CUDA(launch_pdl(allreduce_nvls_oneshot_kernel<B200ATTN_BF16>, g1, b1, 0, st, p));
    } else if (dtype == B200ATTN_FP16) {
      B2_CUDA(launch_pdl(allreduce_nvls_oneshot_kernel<B200ATTN_FP16>, g1, b1, 0, st, p));
    } else {
      B2_CUDA(launch_pdl(allreduce_nvls_oneshot_kernel<B200ATTN_FP32>, g1, b1, 0, st, p));
    }
    return B200ATTN_OK;
  }
  // one CTA per 32 KB, at most kArMaxBlocks (all of them must be resident on every rank at the same time)
  int64_t blocks = (bytes + 32767) / 32768;
  const int sms = sm_count();
  if (blocks > sms) blocks = sms;
  if (blocks > kArMaxBlocks) blocks = kArMaxBlocks;
  const dim3 grid((unsigned)blocks), block(kArThreads);
  if (dtype == B200ATTN_BF16) {
    B2_CUDA(launch_pdl(allreduce_nvls_kernel<B200ATTN_BF16>, grid, block, 0, st, p));
  } else if (dtype == B200ATTN_FP16) {
    B2_CUDA(launch_pdl(allreduce_nvls_kernel<B200ATTN_FP16>, grid, block, 0, st, p));
  } else {
    B2_CUDA(launch_pdl(allreduce_nvls_kernel<B200ATTN_FP32>, grid, block, 0, st, p));
  }
  return B200ATTN_OK;
}

#ifdef AR_TRACE
extern "C" int b200attn_allreduce_trace_read(void* dst_host, int64_t bytes) {
  return cudaMemcpyFromSymbol(dst_host, b200attn::g_ar_trace, (size_t)bytes) == cudaSuccess ? 0 : 1;
}
#endif
