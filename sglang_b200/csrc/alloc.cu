// (f3) Paged KV allocator on the device: out_cache_loc for an extend / decode batch from the free-page list.
//
// Behaviour follows
//   python/sglang/kernels/ops/memory/allocator.py:15-103   (alloc_extend_kernel)
//   python/sglang/kernels/ops/memory/allocator.py:106-140  (alloc_decode_kernel)
//   python/sglang/srt/mem_cache/allocator/paged.py:44-102  (alloc_extend_naive, the oracle)
// Per request b (prefix P, new length S, last_loc = slot of token P-1):
//   part 1: the tail of the current partial page       -> last_loc + 1 + j
//   part 2: whole fresh pages from free_pages           -> page * page_size + j
//   part 3: the head of one more fresh page (S % page)  -> page * page_size + j
// with fresh pages taken in request order (exclusive prefix sum of new-page counts) and outputs packed in
// request order (exclusive prefix sum of extend lengths).  Pure integer work, bit-exact.
//
// One CTA per request.  The two exclusive prefix sums are recomputed per CTA by a block reduction over the
// requests before it (B <= a few thousand => at most a few loads per thread, all L2 hits): no workspace, no second
// launch, no host sync, CUDA-graph capturable.  Output stores are coalesced 8-byte (or 4-byte) writes.
#include "common.cuh"

namespace b200attn {

__device__ __forceinline__ void store_index(void* p, int dtype, int64_t i, int64_t v) {
  if (dtype == B200ATTN_I64) reinterpret_cast<int64_t*>(p)[i] = v;
  else reinterpret_cast<int32_t*>(p)[i] = (int32_t)v;
}

// block-wide sum of two int64 values (blockDim.x == 256)
__device__ __forceinline__ void block_sum2(int64_t& a, int64_t& b) {
  __shared__ int64_t sa[8], sb[8];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    a += __shfl_xor_sync(0xffffffffu, a, o);
    b += __shfl_xor_sync(0xffffffffu, b, o);
  }
  const int w = threadIdx.x >> 5;
  if ((threadIdx.x & 31) == 0) {
    sa[w] = a;
    sb[w] = b;
  }
  __syncthreads();
  a = 0;
  b = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    a += sa[i];
    b += sb[i];
  }
}

__global__ void __launch_bounds__(256)
    alloc_extend_kernel(const void* __restrict__ pre_lens, const void* __restrict__ seq_lens,
                        const void* __restrict__ last_loc, int len_dtype,
                        const void* __restrict__ free_pages, int page_dtype, void* __restrict__ out,
                        int out_dtype, int page_size, int64_t num_free, int32_t* __restrict__ status) {
  const int b = blockIdx.x;
  int64_t tok_before = 0, pages_before = 0;
  for (int i = threadIdx.x; i < b; i += blockDim.x) {
    const int64_t s = load_index(seq_lens, len_dtype, i), p = load_index(pre_lens, len_dtype, i);
    tok_before += s - p;
    pages_before += (s + page_size - 1) / page_size - (p + page_size - 1) / page_size;
  }
  block_sum2(tok_before, pages_before);
  const int64_t S = load_index(seq_lens, len_dtype, b), P = load_index(pre_lens, len_dtype, b);
  const int64_t my_pages = (S + page_size - 1) / page_size - (P + page_size - 1) / page_size;
  if (pages_before + my_pages > num_free) {  // paged.py:214-215: caller returns None; nothing is written
    if (threadIdx.x == 0 && status != nullptr) atomicExch(status, 1);
    return;
  }
  const int64_t p_up = (P + page_size - 1) / page_size * page_size;
  const int64_t n1 = (S < p_up ? S : p_up) - P;
  const int64_t n2 = n1 + P == S ? 0 : S / page_size * page_size - p_up;
  const int64_t n3 = S - P - n1 - n2;
  const int64_t ll = load_index(last_loc, len_dtype, b);
  for (int64_t j = threadIdx.x; j < S - P; j += blockDim.x) {
    int64_t v;
    if (j < n1) {
      v = ll + 1 + j;
    } else if (j < n1 + n2) {
      const int64_t o = j - n1;
      v = load_index(free_pages, page_dtype, pages_before + o / page_size) * page_size + o % page_size;
    } else {
      v = load_index(free_pages, page_dtype, pages_before + my_pages - 1) * page_size + (j - n1 - n2);
    }
    store_index(out, out_dtype, tok_before + j, v);
  }
  (void)n3;
}

__global__ void __launch_bounds__(256)
    alloc_decode_kernel(const void* __restrict__ seq_lens, const void* __restrict__ last_loc,
                        int len_dtype, const void* __restrict__ free_pages, int page_dtype,
                        void* __restrict__ out, int out_dtype, int page_size, int B, int64_t num_free,
                        int32_t* __restrict__ status) {
  // one thread per request; new-page prefix count by a block scan over 256-request chunks
  __shared__ int warp_tot[8];
  __shared__ int64_t carry_s;
  if (threadIdx.x == 0) carry_s = 0;
  __syncthreads();
  for (int base = 0; base < B; base += blockDim.x) {
    const int b = base + threadIdx.x;
    int need = 0;
    int64_t S = 0;
    if (b < B) {
      S = load_index(seq_lens, len_dtype, b);
      need = (int)((S + page_size - 1) / page_size - (S - 1 + page_size - 1) / page_size);
    }
    int incl = need;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, incl, o);
      if ((threadIdx.x & 31) >= o) incl += t;
    }
    const int w = threadIdx.x >> 5;
    if ((threadIdx.x & 31) == 31) warp_tot[w] = incl;
    __syncthreads();
    int woff = 0, tot = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (i < w) woff += warp_tot[i];
      tot += warp_tot[i];
    }
    const int64_t before = carry_s + woff + incl - need;
    if (b < B) {
      if (need == 0) {
        store_index(out, out_dtype, b, load_index(last_loc, len_dtype, b) + 1);
      } else if (before < num_free) {
        store_index(out, out_dtype, b, load_index(free_pages, page_dtype, before) * page_size);
      } else if (status != nullptr) {
        atomicExch(status, 1);
      }
    }
    __syncthreads();
    if (threadIdx.x == 0) carry_s += tot;
    __syncthreads();
  }
}

}  // namespace b200attn

extern "C" int b200attn_alloc_extend(const void* prefix_lens, const void* seq_lens,
                                     const void* last_loc, int len_dtype, const void* free_pages,
                                     int page_dtype, int64_t num_free, void* out_indices,
                                     int out_dtype, int32_t B, int32_t page_size, int32_t* status,
                                     void* stream) {
  using namespace b200attn;
  B2_CHECK_ARG(B >= 0 && page_size >= 1, "alloc_extend: bad B / page_size");
  if (B == 0) return B200ATTN_OK;
  B2_CHECK_ARG(prefix_lens && seq_lens && last_loc && free_pages && out_indices,
               "alloc_extend: null pointer");
  B2_CHECK_ARG((len_dtype | page_dtype | out_dtype) >> 1 == 0, "alloc_extend: bad index dtype");
  alloc_extend_kernel<<<B, 256, 0, (cudaStream_t)stream>>>(prefix_lens, seq_lens, last_loc, len_dtype,
                                                           free_pages, page_dtype, out_indices,
                                                           out_dtype, page_size, num_free, status);
  B2_LAUNCH_CHECK();
  return B200ATTN_OK;
}

extern "C" int b200attn_alloc_decode(const void* seq_lens, const void* last_loc, int len_dtype,
                                     const void* free_pages, int page_dtype, int64_t num_free,
                                     void* out_indices, int out_dtype, int32_t B, int32_t page_size,
                                     int32_t* status, void* stream) {
  using namespace b200attn;
  B2_CHECK_ARG(B >= 0 && page_size >= 1, "alloc_decode: bad B / page_size");
  if (B == 0) return B200ATTN_OK;
  B2_CHECK_ARG(seq_lens && last_loc && free_pages && out_indices, "alloc_decode: null pointer");
  B2_CHECK_ARG((len_dtype | page_dtype | out_dtype) >> 1 == 0, "alloc_decode: bad index dtype");
  alloc_decode_kernel<<<1, 256, 0, (cudaStream_t)stream>>>(seq_lens, last_loc, len_dtype, free_pages,
                                                           page_dtype, out_indices, out_dtype,
                                                           page_size, B, num_free, status);
  B2_LAUNCH_CHECK();
  return B200ATTN_OK;
}
