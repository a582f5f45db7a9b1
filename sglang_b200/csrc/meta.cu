// (a3) ForwardBatch metadata: kv_indptr / kv_indices / qo_indptr and the page table.
//
// Integer bookkeeping, bit-exact against the reference builders
//   python/sglang/kernels/ops/kvcache/kv_indices.py:8-44
//   python/sglang/srt/layers/attention/triton_backend.py:438-456
//   python/sglang/kernels/ops/kvcache/trtllm_mha_page_table.py:40-120
// ONE launch, no host sync, static shapes => CUDA-graph replayable: a (chunks, B) grid copies
// req_to_token rows into the flat index list, 4 indices per thread, coalesced on both sides; every
// CTA derives its request's offset itself (block reduction over the lengths of the requests before
// it: B loads, far cheaper than a second dependent launch) and the first CTA of a request also
// publishes kv_indptr / qo_indptr.  Batches above kFusedMaxB keep the scan + gather pair.
// An optional slot map translates every slot id on the way out (hybrid sliding-window pools:
// full -> SWA pool index, python/sglang/srt/mem_cache/swa_memory_pool.py:228-231).
#include "common.cuh"

namespace b200attn {

constexpr int kScanThreads = 1024;

__global__ void __launch_bounds__(kScanThreads)
    indptr_scan_kernel(const void* __restrict__ kv_lens, int kv_dtype,
                       const void* __restrict__ qo_lens, int qo_dtype, int B,
                       int32_t* __restrict__ kv_indptr, int32_t* __restrict__ qo_indptr) {
  __shared__ int32_t warp_sums[2][32];
  __shared__ int32_t carry[2];
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  if (tid == 0) {
    carry[0] = carry[1] = 0;
    kv_indptr[0] = 0;
    if (qo_indptr) qo_indptr[0] = 0;
  }
  __syncthreads();
  for (int base = 0; base < B; base += kScanThreads) {
    const int i = base + tid;
    int32_t x[2];
    x[0] = (i < B) ? (int32_t)load_index(kv_lens, kv_dtype, i) : 0;
    x[1] = (i < B && qo_lens) ? (int32_t)load_index(qo_lens, qo_dtype, i) : 0;
#pragma unroll
    for (int w = 0; w < 2; ++w) {
      int32_t s = x[w];
#pragma unroll
      for (int off = 1; off < 32; off <<= 1) {
        int32_t n = __shfl_up_sync(0xffffffffu, s, off);
        if (lane >= off) s += n;
      }
      if (lane == 31) warp_sums[w][wid] = s;
      x[w] = s;
    }
    __syncthreads();
    if (wid == 0) {
#pragma unroll
      for (int w = 0; w < 2; ++w) {
        int32_t s = warp_sums[w][lane];
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
          int32_t n = __shfl_up_sync(0xffffffffu, s, off);
          if (lane >= off) s += n;
        }
        warp_sums[w][lane] = s;  // inclusive scan over warps
      }
    }
    __syncthreads();
    const int32_t kv_incl = x[0] + (wid ? warp_sums[0][wid - 1] : 0) + carry[0];
    const int32_t qo_incl = x[1] + (wid ? warp_sums[1][wid - 1] : 0) + carry[1];
    if (i < B) {
      kv_indptr[i + 1] = kv_incl;
      if (qo_indptr) qo_indptr[i + 1] = qo_incl;
    }
    __syncthreads();
    if (tid == kScanThreads - 1) {
      carry[0] = kv_incl;
      carry[1] = qo_incl;
    }
    __syncthreads();
  }
}

constexpr int kGatherThreads = 256;
constexpr int kGatherPerThread = 4;
constexpr int kGatherChunk = kGatherThreads * kGatherPerThread;

template <typename OutT>
__device__ __forceinline__ void gather_row(const int32_t* __restrict__ req_to_token, int64_t stride,
                                           const void* __restrict__ req_pool_indices, int req_dtype,
                                           const void* __restrict__ kv_start, int start_dtype, int b,
                                           int32_t beg, int32_t len, OutT* __restrict__ kv_indices,
                                           int64_t capacity, const void* __restrict__ slot_map,
                                           int map_dtype) {
  const int32_t first = blockIdx.x * kGatherChunk;
  if (first >= len) return;
  const int64_t req = load_index(req_pool_indices, req_dtype, b);
  const int64_t start = kv_start ? load_index(kv_start, start_dtype, b) : 0;
  const int32_t* __restrict__ row = req_to_token + req * stride + start;
  int32_t vals[kGatherPerThread];
#pragma unroll
  for (int j = 0; j < kGatherPerThread; ++j) {
    const int32_t p = first + j * kGatherThreads + threadIdx.x;
    vals[j] = (p < len) ? __ldg(row + p) : 0;
  }
#pragma unroll
  for (int j = 0; j < kGatherPerThread; ++j) {
    const int32_t p = first + j * kGatherThreads + threadIdx.x;
    if (p < len && (int64_t)beg + p < capacity) {
      const int64_t slot = slot_map ? load_index(slot_map, map_dtype, vals[j]) : (int64_t)vals[j];
      kv_indices[(int64_t)beg + p] = (OutT)slot;
    }
  }
}

// out[i] = slot_map[loc[i]] (the SWA translation of out_cache_loc, swa_memory_pool.py:228-231)
__global__ void __launch_bounds__(256)
    translate_slots_kernel(const void* __restrict__ loc, int loc_dtype, const void* __restrict__ slot_map,
                           int map_dtype, int64_t n, int64_t* __restrict__ out) {
  const int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (i < n) out[i] = load_index(slot_map, map_dtype, load_index(loc, loc_dtype, i));
}

template <typename OutT>
__global__ void __launch_bounds__(kGatherThreads)
    kv_indices_gather_kernel(const int32_t* __restrict__ req_to_token, int64_t stride,
                             const void* __restrict__ req_pool_indices, int req_dtype,
                             const void* __restrict__ kv_start, int start_dtype,
                             const int32_t* __restrict__ kv_indptr,
                             OutT* __restrict__ kv_indices, int64_t capacity,
                             const void* __restrict__ slot_map, int map_dtype) {
  const int b = blockIdx.y;
  const int32_t beg = kv_indptr[b];
  const int32_t len = kv_indptr[b + 1] - beg;
  gather_row<OutT>(req_to_token, stride, req_pool_indices, req_dtype, kv_start, start_dtype, b, beg, len,
                   kv_indices, capacity, slot_map, map_dtype);
}

// Fused variant: offsets by block reduction, indptr published by the request's first CTA.
constexpr int kFusedMaxB = 4096;

template <typename OutT>
__global__ void __launch_bounds__(kGatherThreads)
    kv_meta_fused_kernel(const int32_t* __restrict__ req_to_token, int64_t stride,
                         const void* __restrict__ req_pool_indices, int req_dtype,
                         const void* __restrict__ kv_lens, int kv_dtype,
                         const void* __restrict__ kv_start, int start_dtype,
                         const void* __restrict__ qo_lens, int qo_dtype,
                         int32_t* __restrict__ kv_indptr, int32_t* __restrict__ qo_indptr,
                         OutT* __restrict__ kv_indices, int64_t capacity,
                         const void* __restrict__ slot_map, int map_dtype) {
  const int b = blockIdx.y;
  const int32_t len = (int32_t)load_index(kv_lens, kv_dtype, b);
  const int32_t first = blockIdx.x * kGatherChunk;
  if (blockIdx.x > 0 && first >= len) return;  // only a request's first CTA has work besides the gather
  __shared__ int32_t red[2][kGatherThreads / 32];
  int32_t s_kv = 0, s_qo = 0;
  const bool want_qo = blockIdx.x == 0 && qo_lens != nullptr;
  for (int i = threadIdx.x; i < b; i += kGatherThreads) {
    s_kv += (int32_t)load_index(kv_lens, kv_dtype, i);
    if (want_qo) s_qo += (int32_t)load_index(qo_lens, qo_dtype, i);
  }
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) {
    s_kv += __shfl_xor_sync(0xffffffffu, s_kv, off);
    s_qo += __shfl_xor_sync(0xffffffffu, s_qo, off);
  }
  if ((threadIdx.x & 31) == 0) {
    red[0][threadIdx.x >> 5] = s_kv;
    red[1][threadIdx.x >> 5] = s_qo;
  }
  __syncthreads();
  int32_t beg = 0, qbeg = 0;
#pragma unroll
  for (int w = 0; w < kGatherThreads / 32; ++w) {
    beg += red[0][w];
    qbeg += red[1][w];
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    if (b == 0) {
      kv_indptr[0] = 0;
      if (qo_indptr) qo_indptr[0] = 0;
    }
    kv_indptr[b + 1] = beg + len;
    if (qo_indptr) qo_indptr[b + 1] = qbeg + (int32_t)load_index(qo_lens, qo_dtype, b);
  }
  if (kv_indices == nullptr) return;
  gather_row<OutT>(req_to_token, stride, req_pool_indices, req_dtype, kv_start, start_dtype, b, beg, len,
                   kv_indices, capacity, slot_map, map_dtype);
}

__global__ void __launch_bounds__(256)
    page_table_kernel(const int32_t* __restrict__ req_to_token, int64_t stride,
                      const void* __restrict__ req_pool_indices, int req_dtype,
                      const void* __restrict__ seq_lens, int len_dtype, int page_size,
                      int32_t* __restrict__ page_table, int64_t pt_stride, int max_pages) {
  const int b = blockIdx.y;
  const int64_t len = load_index(seq_lens, len_dtype, b);
  int64_t pages = (len + page_size - 1) / page_size;
  if (pages > max_pages) pages = max_pages;
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= pages) return;
  const int64_t req = load_index(req_pool_indices, req_dtype, b);
  const int32_t slot = __ldg(req_to_token + req * stride + (int64_t)j * page_size);
  page_table[b * pt_stride + j] = slot / page_size;
}

// one item per QUERY TOKEN of an extend batch (the generic-head-dim extend path runs them through the decode
// kernel): the request the token belongs to and the number of leading keys it sees
__global__ void __launch_bounds__(256)
    token_items_kernel(const int32_t* __restrict__ qo_indptr, const int32_t* __restrict__ kv_indptr, int B, int64_t T,
                       int causal, int32_t* __restrict__ req_ids, int32_t* __restrict__ item_lens) {
  const int64_t t = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (t >= T) return;
  if (t >= qo_indptr[B]) {  // padded token of a static buffer
    req_ids[t] = 0;
    item_lens[t] = 0;
    return;
  }
  int lo = 0, hi = B;  // largest b with qo_indptr[b] <= t
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if (qo_indptr[mid] <= t) lo = mid; else hi = mid;
  }
  const int q0 = qo_indptr[lo], E = qo_indptr[lo + 1] - q0;
  const int L = kv_indptr[lo + 1] - kv_indptr[lo];
  req_ids[t] = lo;
  item_lens[t] = causal ? L - E + (int)(t - q0) + 1 : L;
}

}  // namespace b200attn

extern "C" int b200attn_build_token_items(const int32_t* qo_indptr, const int32_t* kv_indptr, int32_t B, int64_t T,
                                          int causal, int32_t* req_ids, int32_t* item_lens, void* stream) {
  using namespace b200attn;
  B2_CHECK_ARG(B >= 0 && T >= 0, "build_token_items: negative size");
  if (T == 0 || B == 0) return B200ATTN_OK;
  B2_CHECK_ARG(qo_indptr && kv_indptr && req_ids && item_lens, "build_token_items: null pointer");
  token_items_kernel<<<(unsigned)((T + 255) / 256), 256, 0, (cudaStream_t)stream>>>(qo_indptr, kv_indptr, B, T,
                                                                                    causal, req_ids, item_lens);
  B2_LAUNCH_CHECK();
  return B200ATTN_OK;
}

extern "C" int b200attn_build_kv_indices(
    const int32_t* req_to_token, int64_t req_to_token_stride, const void* req_pool_indices,
    int req_pool_dtype, const void* kv_lens, int kv_lens_dtype, const void* kv_start,
    int kv_start_dtype, const void* qo_lens, int qo_lens_dtype, int32_t B, int32_t* kv_indptr,
    void* kv_indices, int kv_indices_dtype, int64_t kv_indices_capacity, int32_t* qo_indptr,
    const void* slot_map, int slot_map_dtype, void* stream) {
  using namespace b200attn;
  B2_CHECK_ARG(B >= 0, "build_kv_indices: negative batch");
  B2_CHECK_ARG(kv_indptr != nullptr, "build_kv_indices: kv_indptr is null");
  B2_CHECK_ARG(B == 0 || (req_to_token && req_pool_indices && kv_lens),
               "build_kv_indices: null input");
  B2_CHECK_ARG((qo_lens == nullptr) == (qo_indptr == nullptr),
               "build_kv_indices: qo_lens and qo_indptr must be given together");
  B2_CHECK_ARG(kv_indices_dtype == B200ATTN_I32 || kv_indices_dtype == B200ATTN_I64,
               "build_kv_indices: bad kv_indices dtype");
  B2_CHECK_ARG(slot_map == nullptr || slot_map_dtype == B200ATTN_I32 || slot_map_dtype == B200ATTN_I64,
               "build_kv_indices: bad slot_map dtype");
  cudaStream_t st = (cudaStream_t)stream;
  const bool want_indices = kv_indices != nullptr && kv_indices_capacity > 0;
  // grid.x is sized from the row width (an upper bound on any request's length),
  // so no device->host read of the lengths is needed.
  int64_t chunks = want_indices ? (req_to_token_stride + kGatherChunk - 1) / kGatherChunk : 1;
  if (chunks < 1) chunks = 1;
  B2_CHECK_ARG(chunks <= 2147483647ll && B <= 65535, "build_kv_indices: grid too large");
  if (B > 0 && B <= kFusedMaxB) {
    dim3 grid((unsigned)chunks, (unsigned)B);
    if (kv_indices_dtype == B200ATTN_I32) {
      kv_meta_fused_kernel<int32_t><<<grid, kGatherThreads, 0, st>>>(
          req_to_token, req_to_token_stride, req_pool_indices, req_pool_dtype, kv_lens, kv_lens_dtype,
          kv_start, kv_start_dtype, qo_lens, qo_lens_dtype, kv_indptr, qo_indptr,
          want_indices ? (int32_t*)kv_indices : nullptr, kv_indices_capacity, slot_map, slot_map_dtype);
    } else {
      kv_meta_fused_kernel<int64_t><<<grid, kGatherThreads, 0, st>>>(
          req_to_token, req_to_token_stride, req_pool_indices, req_pool_dtype, kv_lens, kv_lens_dtype,
          kv_start, kv_start_dtype, qo_lens, qo_lens_dtype, kv_indptr, qo_indptr,
          want_indices ? (int64_t*)kv_indices : nullptr, kv_indices_capacity, slot_map, slot_map_dtype);
    }
    B2_LAUNCH_CHECK();
    return B200ATTN_OK;
  }
  indptr_scan_kernel<<<1, kScanThreads, 0, st>>>(kv_lens, kv_lens_dtype, qo_lens, qo_lens_dtype,
                                                 B, kv_indptr, qo_indptr);
  B2_LAUNCH_CHECK();
  if (B == 0 || !want_indices) return B200ATTN_OK;
  dim3 grid((unsigned)chunks, (unsigned)B);
  if (kv_indices_dtype == B200ATTN_I32) {
    kv_indices_gather_kernel<int32_t><<<grid, kGatherThreads, 0, st>>>(
        req_to_token, req_to_token_stride, req_pool_indices, req_pool_dtype, kv_start,
        kv_start_dtype, kv_indptr, (int32_t*)kv_indices, kv_indices_capacity, slot_map, slot_map_dtype);
  } else {
    kv_indices_gather_kernel<int64_t><<<grid, kGatherThreads, 0, st>>>(
        req_to_token, req_to_token_stride, req_pool_indices, req_pool_dtype, kv_start,
        kv_start_dtype, kv_indptr, (int64_t*)kv_indices, kv_indices_capacity, slot_map, slot_map_dtype);
  }
  B2_LAUNCH_CHECK();
  return B200ATTN_OK;
}

extern "C" int b200attn_translate_slots(const void* loc, int loc_dtype, const void* slot_map,
                                        int slot_map_dtype, int64_t n, int64_t* out, void* stream) {
  using namespace b200attn;
  B2_CHECK_ARG(n >= 0, "translate_slots: negative size");
  if (n == 0) return B200ATTN_OK;
  B2_CHECK_ARG(loc && slot_map && out, "translate_slots: null pointer");
  translate_slots_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      loc, loc_dtype, slot_map, slot_map_dtype, n, out);
  B2_LAUNCH_CHECK();
  return B200ATTN_OK;
}

extern "C" int b200attn_build_page_table(const int32_t* req_to_token,
                                         int64_t req_to_token_stride,
                                         const void* req_pool_indices, int req_pool_dtype,
                                         const void* seq_lens, int seq_lens_dtype, int32_t B,
                                         int32_t page_size, int32_t* page_table,
                                         int64_t page_table_stride, int32_t max_pages,
                                         void* stream) {
  using namespace b200attn;
  B2_CHECK_ARG(B >= 0 && page_size >= 1 && max_pages >= 0, "build_page_table: bad sizes");
  if (B == 0 || max_pages == 0) return B200ATTN_OK;
  B2_CHECK_ARG(req_to_token && req_pool_indices && seq_lens && page_table,
               "build_page_table: null pointer");
  B2_CHECK_ARG(B <= 65535, "build_page_table: batch too large");
  dim3 grid((unsigned)((max_pages + 255) / 256), (unsigned)B);
  page_table_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(
      req_to_token, req_to_token_stride, req_pool_indices, req_pool_dtype, seq_lens,
      seq_lens_dtype, page_size, page_table, page_table_stride, max_pages);
  B2_LAUNCH_CHECK();
  return B200ATTN_OK;
}
