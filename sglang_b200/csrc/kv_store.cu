// (a1) KV store: scatter new K/V rows into the paged pool through out_cache_loc.
//
// Behaviour follows the reference store path
//   python/sglang/srt/mem_cache/memory_pool.py:141-192 (_set_kv_buffer_impl)
//   python/sglang/kernels/jit/csrc/elementwise/kvcache.cuh:186-218 (store_kvcache)
// re-designed for B200: one warp moves one (token, K|V) row with 16-byte vector
// accesses, all loads of a row batch issued before the stores (MLP >= 4 per lane),
// streaming cache hints (rows are written once and read by later kernels from L2/HBM).
// The copy is dtype-agnostic raw bytes => bit-exact by construction.
#include "common.cuh"

namespace b200attn {

template <typename Vec>
__device__ __forceinline__ Vec ld_stream(const Vec* p) {
  return *p;
}
template <>
__device__ __forceinline__ uint4 ld_stream<uint4>(const uint4* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];\n"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}

template <typename Vec, int kWarpsPerBlock>
__global__ void __launch_bounds__(kWarpsPerBlock * 32)
    kv_store_kernel(const uint8_t* __restrict__ k, const uint8_t* __restrict__ v,
                    uint8_t* __restrict__ k_cache, uint8_t* __restrict__ v_cache,
                    const void* __restrict__ loc, int loc_dtype, int64_t T, int k_chunks,
                    int v_chunks, int64_t k_src_stride, int64_t v_src_stride,
                    int64_t k_cache_stride, int64_t v_cache_stride, int64_t size_limit,
                    int64_t skip_index, int32_t* __restrict__ oob_count) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = (int64_t)blockIdx.x * kWarpsPerBlock + (threadIdx.x >> 5);
  if (warp >= 2 * T) return;
  const int64_t t = warp >> 1;
  const bool is_v = warp & 1;
  const int64_t slot = load_index(loc, loc_dtype, t);
  if (slot == skip_index) return;  // kvcache.cuh:215 (reserved padding slot)
  if (slot < 0 || slot >= size_limit) {
    if (lane == 0 && oob_count != nullptr) atomicAdd(oob_count, 1);
    return;
  }
  const Vec* src = reinterpret_cast<const Vec*>(is_v ? v + t * v_src_stride : k + t * k_src_stride);
  Vec* dst = reinterpret_cast<Vec*>(is_v ? v_cache + slot * v_cache_stride
                                         : k_cache + slot * k_cache_stride);
  const int n = is_v ? v_chunks : k_chunks;
  int i = lane;
  for (; i + 96 < n; i += 128) {
    Vec a = ld_stream(src + i), b = ld_stream(src + i + 32), c = ld_stream(src + i + 64),
        d = ld_stream(src + i + 96);
    dst[i] = a;
    dst[i + 32] = b;
    dst[i + 64] = c;
    dst[i + 96] = d;
  }
  for (; i < n; i += 32) dst[i] = ld_stream(src + i);
}

template <typename Vec>
static int launch_store(const void* k, const void* v, void* k_cache, void* v_cache,
                        const void* loc, int loc_dtype, int64_t T, int64_t k_row_bytes,
                        int64_t v_row_bytes, int64_t ks, int64_t vs, int64_t kcs, int64_t vcs,
                        int64_t size_limit, int64_t skip_index, int32_t* oob, cudaStream_t st) {
  constexpr int kWarps = 8;
  const int64_t warps = 2 * T;
  const int64_t blocks = (warps + kWarps - 1) / kWarps;
  B2_CHECK_ARG(blocks < (1ll << 31), "kv_store: T too large");
  kv_store_kernel<Vec, kWarps><<<(unsigned)blocks, kWarps * 32, 0, st>>>(
      (const uint8_t*)k, (const uint8_t*)v, (uint8_t*)k_cache, (uint8_t*)v_cache, loc, loc_dtype,
      T, (int)(k_row_bytes / sizeof(Vec)), (int)(v_row_bytes / sizeof(Vec)), ks, vs, kcs, vcs,
      size_limit, skip_index, oob);
  B2_LAUNCH_CHECK();
  return B200ATTN_OK;
}

}  // namespace b200attn

extern "C" int b200attn_kv_store(const void* k, const void* v, void* k_cache, void* v_cache,
                                 const void* loc, int loc_dtype, int64_t T, int64_t k_row_bytes,
                                 int64_t v_row_bytes, int64_t k_src_stride_bytes,
                                 int64_t v_src_stride_bytes, int64_t k_cache_stride_bytes,
                                 int64_t v_cache_stride_bytes, int64_t size_limit,
                                 int64_t skip_index, int32_t* oob_count, void* stream) {
  using namespace b200attn;
  B2_CHECK_ARG(T >= 0, "kv_store: negative T");
  if (T == 0) return B200ATTN_OK;
  B2_CHECK_ARG(k && v && k_cache && v_cache && loc, "kv_store: null pointer");
  B2_CHECK_ARG(loc_dtype == B200ATTN_I32 || loc_dtype == B200ATTN_I64, "kv_store: bad loc dtype");
  B2_CHECK_ARG(k_row_bytes > 0 && v_row_bytes > 0, "kv_store: empty rows");
  B2_CHECK_ARG(k_src_stride_bytes >= k_row_bytes && v_src_stride_bytes >= v_row_bytes &&
                   k_cache_stride_bytes >= k_row_bytes && v_cache_stride_bytes >= v_row_bytes,
               "kv_store: stride smaller than row");
  // widest vector that divides every row size, stride and base address
  uint64_t bits = (uint64_t)k_row_bytes | (uint64_t)v_row_bytes | (uint64_t)k_src_stride_bytes |
                  (uint64_t)v_src_stride_bytes | (uint64_t)k_cache_stride_bytes |
                  (uint64_t)v_cache_stride_bytes | (uint64_t)(uintptr_t)k | (uint64_t)(uintptr_t)v |
                  (uint64_t)(uintptr_t)k_cache | (uint64_t)(uintptr_t)v_cache;
  cudaStream_t st = (cudaStream_t)stream;
#define B2_STORE(VEC)                                                                          \
  return launch_store<VEC>(k, v, k_cache, v_cache, loc, loc_dtype, T, k_row_bytes, v_row_bytes, \
                           k_src_stride_bytes, v_src_stride_bytes, k_cache_stride_bytes,        \
                           v_cache_stride_bytes, size_limit, skip_index, oob_count, st)
  if ((bits & 15) == 0) B2_STORE(uint4);
  if ((bits & 7) == 0) B2_STORE(uint2);
  if ((bits & 3) == 0) B2_STORE(uint32_t);
  if ((bits & 1) == 0) B2_STORE(uint16_t);
  B2_STORE(uint8_t);
#undef B2_STORE
}
