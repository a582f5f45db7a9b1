// Shared helpers for the b200attn kernels (sm_100a only).
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_fp8.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <cstdarg>
#include <cstdio>
#include <cstdlib>

#include "../../include/b200attn.h"

namespace b200attn {

// ---- error plumbing -------------------------------------------------------
void set_error(const char* fmt, ...);

#define B2_CHECK_ARG(cond, ...)               \
  do {                                        \
    if (!(cond)) {                            \
      ::b200attn::set_error(__VA_ARGS__);     \
      return B200ATTN_ERR_INVALID;            \
    }                                         \
  } while (0)

#define B2_UNSUPPORTED(...)                   \
  do {                                        \
    ::b200attn::set_error(__VA_ARGS__);       \
    return B200ATTN_ERR_UNSUPPORTED;          \
  } while (0)

#define B2_CUDA(expr)                                                         \
  do {                                                                        \
    cudaError_t _e = (expr);                                                  \
    if (_e != cudaSuccess) {                                                  \
      ::b200attn::set_error("%s failed: %s (%s:%d)", #expr,                   \
                            cudaGetErrorString(_e), __FILE__, __LINE__);      \
      return B200ATTN_ERR_CUDA;                                               \
    }                                                                         \
  } while (0)

#define B2_LAUNCH_CHECK()                                                     \
  do {                                                                        \
    cudaError_t _e = cudaGetLastError();                                      \
    if (_e != cudaSuccess) {                                                  \
      ::b200attn::set_error("kernel launch failed: %s (%s:%d)",               \
                            cudaGetErrorString(_e), __FILE__, __LINE__);      \
      return B200ATTN_ERR_CUDA;                                               \
    }                                                                         \
  } while (0)

int sm_count();

// ---- device helpers -------------------------------------------------------
__device__ __forceinline__ int64_t load_index(const void* p, int dtype, int64_t i) {
  return dtype == B200ATTN_I64 ? reinterpret_cast<const int64_t*>(p)[i]
                               : (int64_t) reinterpret_cast<const int32_t*>(p)[i];
}

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

// 16-byte async global->shared copy; src_bytes < 16 zero-fills the remainder.
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, int src_bytes) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(dst), "l"(src),
               "r"(src_bytes));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;\n" ::"n"(N));
}

__device__ __forceinline__ void ldmatrix_x4(uint32_t& r0, uint32_t& r1, uint32_t& r2,
                                            uint32_t& r3, uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];\n"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3)
               : "r"(addr));
}
__device__ __forceinline__ void ldmatrix_x4_trans(uint32_t& r0, uint32_t& r1, uint32_t& r2,
                                                  uint32_t& r3, uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];\n"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3)
               : "r"(addr));
}

// D(16x8,f32) += A(16x16) * B(16x8); bf16 or fp16 operands.
template <bool kBf16>
__device__ __forceinline__ void mma_16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0,
                                          uint32_t b1) {
  if constexpr (kBf16) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, "
        "{%8,%9}, {%0,%1,%2,%3};\n"
        : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
  } else {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, "
        "{%8,%9}, {%0,%1,%2,%3};\n"
        : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
  }
}

template <bool kBf16>
__device__ __forceinline__ uint32_t pack2(float lo, float hi) {
  if constexpr (kBf16) {
    __nv_bfloat162 t = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&t);
  } else {
    __half2 t = __floats2half2_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&t);
  }
}

template <bool kBf16>
__device__ __forceinline__ float2 unpack2(uint32_t x) {
  if constexpr (kBf16) {
    return __bfloat1622float2(*reinterpret_cast<__nv_bfloat162*>(&x));
  } else {
    return __half22float2(*reinterpret_cast<__half2*>(&x));
  }
}

// float -> float8_e4m3fn with c10's semantics (round to nearest even; what would round above 448 becomes the
// NaN pattern 0x7f, no saturation): 464 is the midpoint of 448 and the non-existent 480, ties go to even (448).
__device__ __forceinline__ uint8_t to_e4m3_c10(float x) {
  if (!(fabsf(x) <= 464.f)) return (uint8_t)(0x7f | ((__float_as_uint(x) >> 24) & 0x80));
  return (uint8_t)__nv_cvt_float_to_fp8(x, __NV_SATFINITE, __NV_E4M3);
}
template <bool kBf16>
__device__ __forceinline__ float round_to_src(float x) {
  if constexpr (kBf16) return __bfloat162float(__float2bfloat16_rn(x));
  return __half2float(__float2half_rn(x));
}
// one 16-bit element -> e4m3 byte, the pool's rounding order (memory_pool.py:2365-2371): true division by the
// scale rounded to the source dtype, then the cast; scale <= 0 = no division
template <bool kBf16>
__device__ __forceinline__ uint8_t quant_e4m3(uint16_t bits, float scale) {
  float y = unpack2<kBf16>((uint32_t)bits).x;
  if (scale > 0.f) y = round_to_src<kBf16>(__fdiv_rn(y, scale));
  return to_e4m3_c10(y);
}

__device__ __forceinline__ float fast_exp2(float x) {
  float y;
  asm volatile("ex2.approx.ftz.f32 %0, %1;\n" : "=f"(y) : "f"(x));
  return y;
}

constexpr float kLog2e = 1.4426950408889634f;
constexpr float kLn2 = 0.6931471805599453f;

// Split-KV policy shared by the decode kernel and its merge: a pure function of
// the kv length so that graph replay needs no host recomputation.
//   nsplit = clamp(ceil(L / min_chunk), 1, max_splits)
//   chunk  = round_up(ceil(L / nsplit), tile);  nsplit = ceil(L / chunk)
__host__ __device__ __forceinline__ void split_policy(int L, int max_splits, int min_chunk,
                                                      int tile, int& nsplit, int& chunk) {
  if (L <= 0) {
    nsplit = 1;
    chunk = tile;
    return;
  }
  int n = (L + min_chunk - 1) / min_chunk;
  n = n < 1 ? 1 : (n > max_splits ? max_splits : n);
  int c = (L + n - 1) / n;
  c = (c + tile - 1) / tile * tile;
  chunk = c;
  nsplit = (L + c - 1) / c;
}

}  // namespace b200attn

// Programmatic dependent launch (PDL).  A kernel launched through launch_pdl() may be scheduled while the previous
// kernel of the stream is still draining (as soon as all of that kernel's CTAs have executed pdl_trigger() or
// exited); it must execute pdl_wait() before its first global-memory access: the wait returns when the previous
// grid has completed and its writes are visible.  What overlaps is launch latency and CTA start-up, nothing else.
// B200ATTN_PDL=0 falls back to plain stream order.
#ifdef __CUDACC__
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;\n" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;\n" ::: "memory"); }

inline bool pdl_enabled() {
  static const bool on = [] { const char* e = getenv("B200ATTN_PDL"); return e == nullptr || atoi(e) != 0; }();
  return on;
}
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                              Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}
#endif

