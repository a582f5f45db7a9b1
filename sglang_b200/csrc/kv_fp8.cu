// (f4) FP8 (e4m3) KV cache: quantising store and the dequantising gather used by extend.
//
// Store semantics follow the reference pool
//   python/sglang/srt/mem_cache/memory_pool.py:2365-2375
//     if cache_k.dtype != self.dtype: cache_k.div_(k_scale); cache_k = cache_k.to(self.dtype)
//     (then viewed as uint8, the store dtype of float8 pools)
// i.e. per element, in this order: true division by the scale IN THE INPUT DTYPE (the fp32 quotient
// is rounded to bf16/fp16 first), then the c10 float -> float8_e4m3fn conversion (round to nearest
// even; anything that would round above 448 becomes the NaN pattern 0x7f, there is no saturation).
// The kernel reproduces both roundings, so the stored bytes are bit-exact against
// `(k / k_scale).to(torch.float8_e4m3fn).view(torch.uint8)` on the CPU.
//
// One warp owns one (token, K|V) row; a lane converts 8 elements per step (16 B in, 8 B out).
#include "common.cuh"

namespace b200attn {

template <bool kBf16>
__device__ __forceinline__ uint32_t quant4(uint32_t w0, uint32_t w1, float scale, bool has_scale) {
  const float2 a = unpack2<kBf16>(w0), b = unpack2<kBf16>(w1);
  float x[4] = {a.x, a.y, b.x, b.y};
  uint32_t out = 0;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float y = x[i];
    if (has_scale) y = round_to_src<kBf16>(__fdiv_rn(y, scale));
    out |= (uint32_t)to_e4m3_c10(y) << (8 * i);
  }
  return out;
}

template <bool kBf16, bool kVec, int kWarpsPerBlock>
__global__ void __launch_bounds__(kWarpsPerBlock * 32)
    kv_store_fp8_kernel(const uint16_t* __restrict__ k, const uint16_t* __restrict__ v,
                        uint8_t* __restrict__ k_cache, uint8_t* __restrict__ v_cache,
                        const void* __restrict__ loc, int loc_dtype, int64_t T, int k_elems, int v_elems,
                        int64_t k_src_stride, int64_t v_src_stride, int64_t k_cache_stride,
                        int64_t v_cache_stride, float k_scale, float v_scale, int64_t size_limit,
                        int64_t skip_index, int32_t* __restrict__ oob_count) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = (int64_t)blockIdx.x * kWarpsPerBlock + (threadIdx.x >> 5);
  if (warp >= 2 * T) return;
  const int64_t t = warp >> 1;
  const bool is_v = warp & 1;
  const int64_t slot = load_index(loc, loc_dtype, t);
  if (slot == skip_index) return;
  if (slot < 0 || slot >= size_limit) {
    if (lane == 0 && oob_count != nullptr) atomicAdd(oob_count, 1);
    return;
  }
  const uint16_t* src = is_v ? v + t * v_src_stride : k + t * k_src_stride;
  uint8_t* dst = is_v ? v_cache + slot * v_cache_stride : k_cache + slot * k_cache_stride;
  const int n = is_v ? v_elems : k_elems;
  const float scale = is_v ? v_scale : k_scale;
  const bool has_scale = scale > 0.f;
  if constexpr (kVec) {
    for (int i = lane * 8; i < n; i += 256) {
      const uint4 w = *reinterpret_cast<const uint4*>(src + i);
      uint2 o;
      o.x = quant4<kBf16>(w.x, w.y, scale, has_scale);
      o.y = quant4<kBf16>(w.z, w.w, scale, has_scale);
      *reinterpret_cast<uint2*>(dst + i) = o;
    }
  } else {
    for (int i = lane; i < n; i += 32) {
      const float2 f = unpack2<kBf16>((uint32_t)src[i]);
      float y = f.x;
      if (has_scale) y = round_to_src<kBf16>(__fdiv_rn(y, scale));
      dst[i] = to_e4m3_c10(y);
    }
  }
}

// out_k[i] = T(k8[kv_indices[i]] * k_mul), out_v[i] = T(v8[kv_indices[i]] * v_mul): the extend kernel's
// K/V operands are 16-bit, so an fp8 pool is gathered once per forward into a contiguous 16-bit scratch
// (slot ids become 0..n-1: every K/V tile then takes the kernel's contiguous TMA-box path).  With
// k_mul = 1 the K rows are exact (every e4m3 value is a bf16/fp16 value) and k_descale rides on sm_scale.
template <bool kBf16, int kWarpsPerBlock>
__global__ void __launch_bounds__(kWarpsPerBlock * 32)
    kv_dequant_gather_kernel(const uint8_t* __restrict__ k_cache, const uint8_t* __restrict__ v_cache,
                             int64_t cache_stride, const int32_t* __restrict__ kv_indices, int64_t n,
                             int row_elems, uint16_t* __restrict__ out_k, uint16_t* __restrict__ out_v,
                             int64_t out_stride, float k_mul, float v_mul, const int32_t* __restrict__ n_dev) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = (int64_t)blockIdx.x * kWarpsPerBlock + (threadIdx.x >> 5);
  if (n_dev != nullptr) n = min(n, (int64_t)__ldg(n_dev));  // graph replay: the live row count is device data
  if (warp >= 2 * n) return;
  const int64_t i = warp >> 1;
  const bool is_v = warp & 1;
  const int64_t slot = kv_indices[i];
  const uint8_t* src = (is_v ? v_cache : k_cache) + slot * cache_stride;
  uint16_t* dst = (is_v ? out_v : out_k) + i * out_stride;
  const float mul = is_v ? v_mul : k_mul;
  for (int e = lane * 16; e < row_elems; e += 512) {
    const uint4 w = *reinterpret_cast<const uint4*>(src + e);
    const uint32_t ws[4] = {w.x, w.y, w.z, w.w};
    uint32_t o[8];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const __half2_raw lo = __nv_cvt_fp8x2_to_halfraw2((__nv_fp8x2_storage_t)(ws[j] & 0xffff), __NV_E4M3);
      const __half2_raw hi = __nv_cvt_fp8x2_to_halfraw2((__nv_fp8x2_storage_t)(ws[j] >> 16), __NV_E4M3);
      const float2 a = __half22float2(*reinterpret_cast<const __half2*>(&lo));
      const float2 b = __half22float2(*reinterpret_cast<const __half2*>(&hi));
      o[2 * j] = pack2<kBf16>(a.x * mul, a.y * mul);
      o[2 * j + 1] = pack2<kBf16>(b.x * mul, b.y * mul);
    }
    *reinterpret_cast<uint4*>(dst + e) = make_uint4(o[0], o[1], o[2], o[3]);
    *reinterpret_cast<uint4*>(dst + e + 8) = make_uint4(o[4], o[5], o[6], o[7]);
  }
}

// ---- de-duplicating variant: a slot referenced by several requests (radix-shared prefix) is dequantised ONCE ----
// first[slot] = smallest row i of kv_indices that references it (atomicMin over a pool-sized int32 map that the
// launcher resets to 0x7f7f7f7f); row i then reads r = first[slot], publishes out_indices[i] = r and only the
// owner row (r == i) converts.  The extend kernel reads the scratch through out_indices, so the shared prefix is
// also shared in L2, exactly as with a 16-bit pool.
__global__ void kv_first_ref_kernel(const int32_t* __restrict__ kv_indices, int64_t n,
                                    const int32_t* __restrict__ n_dev, int32_t* __restrict__ first,
                                    int64_t num_slots) {
  if (n_dev != nullptr) n = min(n, (int64_t)__ldg(n_dev));
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t slot = kv_indices[i];
    if (slot >= 0 && slot < num_slots) atomicMin(first + slot, (int32_t)i);
  }
}

template <bool kBf16, int kWarpsPerBlock>
__global__ void __launch_bounds__(kWarpsPerBlock * 32)
    kv_dequant_gather_unique_kernel(const uint8_t* __restrict__ k_cache, const uint8_t* __restrict__ v_cache,
                                    int64_t cache_stride, const int32_t* __restrict__ kv_indices, int64_t n,
                                    int row_elems, uint16_t* __restrict__ out_k, uint16_t* __restrict__ out_v,
                                    int64_t out_stride, float k_mul, float v_mul, const int32_t* __restrict__ n_dev,
                                    const int32_t* __restrict__ first, int64_t num_slots,
                                    int32_t* __restrict__ out_indices) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = (int64_t)blockIdx.x * kWarpsPerBlock + (threadIdx.x >> 5);
  if (n_dev != nullptr) n = min(n, (int64_t)__ldg(n_dev));
  if (warp >= 2 * n) return;
  const int64_t i = warp >> 1;
  const bool is_v = warp & 1;
  const int64_t slot = kv_indices[i];
  const bool in_range = slot >= 0 && slot < num_slots;
  const int64_t r = in_range ? (int64_t)first[slot] : i;
  if (!is_v && lane == 0) out_indices[i] = (int32_t)r;
  if (r != i || !in_range) return;  // another row owns this slot's 16-bit copy (out-of-range slots stay unread)
  const uint8_t* src = (is_v ? v_cache : k_cache) + slot * cache_stride;
  uint16_t* dst = (is_v ? out_v : out_k) + i * out_stride;
  const float mul = is_v ? v_mul : k_mul;
  for (int e = lane * 16; e < row_elems; e += 512) {
    const uint4 w = *reinterpret_cast<const uint4*>(src + e);
    const uint32_t ws[4] = {w.x, w.y, w.z, w.w};
    uint32_t o[8];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const __half2_raw lo = __nv_cvt_fp8x2_to_halfraw2((__nv_fp8x2_storage_t)(ws[j] & 0xffff), __NV_E4M3);
      const __half2_raw hi = __nv_cvt_fp8x2_to_halfraw2((__nv_fp8x2_storage_t)(ws[j] >> 16), __NV_E4M3);
      const float2 a = __half22float2(*reinterpret_cast<const __half2*>(&lo));
      const float2 b = __half22float2(*reinterpret_cast<const __half2*>(&hi));
      o[2 * j] = pack2<kBf16>(a.x * mul, a.y * mul);
      o[2 * j + 1] = pack2<kBf16>(b.x * mul, b.y * mul);
    }
    *reinterpret_cast<uint4*>(dst + e) = make_uint4(o[0], o[1], o[2], o[3]);
    *reinterpret_cast<uint4*>(dst + e + 8) = make_uint4(o[4], o[5], o[6], o[7]);
  }
}

}  // namespace b200attn

extern "C" int b200attn_kv_store_fp8(const void* k, const void* v, void* k_cache, void* v_cache,
                                     const void* loc, int loc_dtype, int64_t T, int64_t k_row_elems,
                                     int64_t v_row_elems, int64_t k_src_stride, int64_t v_src_stride,
                                     int64_t k_cache_stride, int64_t v_cache_stride, int src_dtype,
                                     float k_scale, float v_scale, int64_t size_limit,
                                     int64_t skip_index, int32_t* oob_count, void* stream) {
  using namespace b200attn;
  B2_CHECK_ARG(T >= 0, "kv_store_fp8: negative T");
  if (T == 0) return B200ATTN_OK;
  B2_CHECK_ARG(k && v && k_cache && v_cache && loc, "kv_store_fp8: null pointer");
  B2_CHECK_ARG(loc_dtype == B200ATTN_I32 || loc_dtype == B200ATTN_I64, "kv_store_fp8: bad loc dtype");
  B2_CHECK_ARG(k_row_elems > 0 && v_row_elems > 0 && k_row_elems < (1ll << 31) && v_row_elems < (1ll << 31),
               "kv_store_fp8: bad row size");
  B2_CHECK_ARG(k_src_stride >= k_row_elems && v_src_stride >= v_row_elems &&
                   k_cache_stride >= k_row_elems && v_cache_stride >= v_row_elems,
               "kv_store_fp8: stride smaller than row");
  if (src_dtype != B200ATTN_BF16 && src_dtype != B200ATTN_FP16) B2_UNSUPPORTED("kv_store_fp8: src dtype %d", src_dtype);
  const uint64_t ebits = (uint64_t)k_row_elems | (uint64_t)v_row_elems | (uint64_t)k_src_stride |
                         (uint64_t)v_src_stride | (uint64_t)k_cache_stride | (uint64_t)v_cache_stride;
  const uint64_t sbits = (uint64_t)(uintptr_t)k | (uint64_t)(uintptr_t)v;
  const uint64_t dbits = (uint64_t)(uintptr_t)k_cache | (uint64_t)(uintptr_t)v_cache;
  const bool vec = (ebits & 7) == 0 && (sbits & 15) == 0 && (dbits & 7) == 0;
  constexpr int kWarps = 8;
  const int64_t blocks = (2 * T + kWarps - 1) / kWarps;
  B2_CHECK_ARG(blocks < (1ll << 31), "kv_store_fp8: T too large");
  cudaStream_t st = (cudaStream_t)stream;
#define B2_ST8(BF, VEC)                                                                              \
  kv_store_fp8_kernel<BF, VEC, kWarps><<<(unsigned)blocks, kWarps * 32, 0, st>>>(                    \
      (const uint16_t*)k, (const uint16_t*)v, (uint8_t*)k_cache, (uint8_t*)v_cache, loc, loc_dtype,  \
      T, (int)k_row_elems, (int)v_row_elems, k_src_stride, v_src_stride, k_cache_stride,             \
      v_cache_stride, k_scale, v_scale, size_limit, skip_index, oob_count)
  if (src_dtype == B200ATTN_BF16) {
    if (vec) B2_ST8(true, true); else B2_ST8(true, false);
  } else {
    if (vec) B2_ST8(false, true); else B2_ST8(false, false);
  }
#undef B2_ST8
  B2_LAUNCH_CHECK();
  return B200ATTN_OK;
}

extern "C" int b200attn_kv_dequant_gather(const void* k_cache, const void* v_cache, int64_t cache_stride,
                                          const int32_t* kv_indices, int64_t n, int64_t row_elems,
                                          void* out_k, void* out_v, int64_t out_stride, int out_dtype,
                                          float k_mul, float v_mul, const int32_t* n_dev, void* stream) {
  using namespace b200attn;
  B2_CHECK_ARG(n >= 0, "kv_dequant_gather: negative n");
  if (n == 0) return B200ATTN_OK;
  B2_CHECK_ARG(k_cache && v_cache && kv_indices && out_k && out_v, "kv_dequant_gather: null pointer");
  B2_CHECK_ARG(row_elems > 0 && row_elems % 16 == 0 && row_elems < (1ll << 31) && cache_stride % 16 == 0 &&
                   out_stride % 8 == 0 && cache_stride >= row_elems && out_stride >= row_elems,
               "kv_dequant_gather: rows must be multiples of 16 elements with 16-byte aligned strides");
  B2_CHECK_ARG((((uintptr_t)k_cache | (uintptr_t)v_cache | (uintptr_t)out_k | (uintptr_t)out_v) & 15) == 0,
               "kv_dequant_gather: tensors must be 16-byte aligned");
  if (out_dtype != B200ATTN_BF16 && out_dtype != B200ATTN_FP16) B2_UNSUPPORTED("kv_dequant_gather: dtype %d", out_dtype);
  constexpr int kWarps = 8;
  const int64_t blocks = (2 * n + kWarps - 1) / kWarps;
  B2_CHECK_ARG(blocks < (1ll << 31), "kv_dequant_gather: n too large");
  cudaStream_t st = (cudaStream_t)stream;
  if (out_dtype == B200ATTN_BF16) {
    kv_dequant_gather_kernel<true, kWarps><<<(unsigned)blocks, kWarps * 32, 0, st>>>(
        (const uint8_t*)k_cache, (const uint8_t*)v_cache, cache_stride, kv_indices, n, (int)row_elems,
        (uint16_t*)out_k, (uint16_t*)out_v, out_stride, k_mul, v_mul, n_dev);
  } else {
    kv_dequant_gather_kernel<false, kWarps><<<(unsigned)blocks, kWarps * 32, 0, st>>>(
        (const uint8_t*)k_cache, (const uint8_t*)v_cache, cache_stride, kv_indices, n, (int)row_elems,
        (uint16_t*)out_k, (uint16_t*)out_v, out_stride, k_mul, v_mul, n_dev);
  }
  B2_LAUNCH_CHECK();
  return B200ATTN_OK;
}

extern "C" int b200attn_kv_dequant_gather_unique(const void* k_cache, const void* v_cache, int64_t cache_stride,
                                                 const int32_t* kv_indices, int64_t n, int64_t row_elems,
                                                 void* out_k, void* out_v, int64_t out_stride, int out_dtype,
                                                 float k_mul, float v_mul, const int32_t* n_dev,
                                                 int32_t* first_ws, int64_t num_slots, int32_t* out_indices,
                                                 void* stream) {
  using namespace b200attn;
  B2_CHECK_ARG(n >= 0 && n < (1ll << 31), "kv_dequant_gather_unique: bad n");
  if (n == 0) return B200ATTN_OK;
  B2_CHECK_ARG(k_cache && v_cache && kv_indices && out_k && out_v && first_ws && out_indices && num_slots > 0,
               "kv_dequant_gather_unique: null pointer");
  B2_CHECK_ARG(row_elems > 0 && row_elems % 16 == 0 && row_elems < (1ll << 31) && cache_stride % 16 == 0 &&
                   out_stride % 8 == 0 && cache_stride >= row_elems && out_stride >= row_elems,
               "kv_dequant_gather_unique: rows must be multiples of 16 elements with 16-byte aligned strides");
  B2_CHECK_ARG((((uintptr_t)k_cache | (uintptr_t)v_cache | (uintptr_t)out_k | (uintptr_t)out_v) & 15) == 0,
               "kv_dequant_gather_unique: tensors must be 16-byte aligned");
  if (out_dtype != B200ATTN_BF16 && out_dtype != B200ATTN_FP16) B2_UNSUPPORTED("kv_dequant_gather_unique: dtype %d", out_dtype);
  cudaStream_t st = (cudaStream_t)stream;
  B2_CUDA(cudaMemsetAsync(first_ws, 0x7f, (size_t)num_slots * sizeof(int32_t), st));
  {
    const int64_t blocks = (n + 255) / 256;
    kv_first_ref_kernel<<<(unsigned)(blocks < 4096 ? blocks : 4096), 256, 0, st>>>(kv_indices, n, n_dev, first_ws,
                                                                                   num_slots);
    B2_LAUNCH_CHECK();
  }
  constexpr int kWarps = 8;
  const int64_t blocks = (2 * n + kWarps - 1) / kWarps;
  B2_CHECK_ARG(blocks < (1ll << 31), "kv_dequant_gather_unique: n too large");
  if (out_dtype == B200ATTN_BF16) {
    kv_dequant_gather_unique_kernel<true, kWarps><<<(unsigned)blocks, kWarps * 32, 0, st>>>(
        (const uint8_t*)k_cache, (const uint8_t*)v_cache, cache_stride, kv_indices, n, (int)row_elems,
        (uint16_t*)out_k, (uint16_t*)out_v, out_stride, k_mul, v_mul, n_dev, first_ws, num_slots, out_indices);
  } else {
    kv_dequant_gather_unique_kernel<false, kWarps><<<(unsigned)blocks, kWarps * 32, 0, st>>>(
        (const uint8_t*)k_cache, (const uint8_t*)v_cache, cache_stride, kv_indices, n, (int)row_elems,
        (uint16_t*)out_k, (uint16_t*)out_v, out_stride, k_mul, v_mul, n_dev, first_ws, num_slots, out_indices);
  }
  B2_LAUNCH_CHECK();
  return B200ATTN_OK;
}
