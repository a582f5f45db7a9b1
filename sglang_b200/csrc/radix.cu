// (f3) Radix prefix match on the device.
//
// Replaces the lookup side of python/sglang/srt/mem_cache/radix_cache.py (RadixCache.match_prefix :377-423,
// insert :425-445, and the slot release of evict :593-621) for the attention path: given the token ids of a batch
// of requests it returns, per request, the KV slot indices of the longest cached prefix, page-granular, bit-exactly
// what the reference's tree returns.
//
// The reference walks a pointer-linked tree on the CPU, node by node.  Here the tree is a PAGE TRIE stored in a
// hash table in HBM: one entry per cached page = (parent entry, the page's `page` token ids, the page's `page`
// KV slots, the hash of the whole token prefix that ends with this page).  Node splits of the reference are
// invisible at this granularity (a node is a run of pages).  A lookup is PARALLEL over pages:
//
//   * lane j owns page j of the request: polynomial page hash (a_j, b_j) of its tokens, an affine warp scan turns
//     them into the prefix hashes h_0..h_31 of 32 pages at once (h_j = h_{j-1} * P^page + poly(page j), mod 2^64);
//   * every lane probes the open-addressing table for ITS prefix hash (independent L2 round trips, not a chain),
//     compares the page's tokens, and the chain is verified with one shuffle: entry_j.parent == entry_{j-1};
//     the match length is the number of leading lanes that pass.  Exact: a hit needs the tokens of the page AND the
//     parent entry to agree, so by induction the whole prefix agrees; the hash only finds candidates.
//
// insert: one warp walks the batch request by request (stream-ordered with the lookups, off the hot path), reuses
// the same probe, and appends the unmatched pages (existing pages keep their slots, as the reference's
// _insert_helper does; the matched length is returned so the caller frees the duplicates).  erase_slots: the
// allocator hands over the slots of evicted leaves; a reverse map KV page -> entry finds and unlinks them.
#include "common.cuh"

namespace b200attn {

constexpr uint64_t kRadixP = 0x9E3779B97F4A7C15ull;   // odd multiplier of the polynomial hash
constexpr uint64_t kRadixSeed = 0xC2B2AE3D27D4EB4Full;
constexpr uint64_t kEmpty = 0ull, kTomb = 1ull;

__device__ __forceinline__ uint64_t radix_key(uint64_t h) { return h < 2 ? h + 2 : h; }
__device__ __forceinline__ uint32_t radix_bucket(uint64_t h, int tcap) {
  h ^= h >> 32;
  h *= 0xD6E8FEB86659FD93ull;
  h ^= h >> 32;
  return (uint32_t)h & (uint32_t)(tcap - 1);
}

// (a, b) of one page: h' = h * a + b
__device__ __forceinline__ void page_poly(const void* toks, int tok_dtype, int64_t off, int page, uint64_t& a,
                                          uint64_t& b) {
  a = 1;
  b = 0;
  for (int i = 0; i < page; ++i) {
    const uint64_t t = (uint64_t)load_index(toks, tok_dtype, off + i) + 0x632BE59BD9B4E019ull;
    b = b * kRadixP + t;
    a *= kRadixP;
  }
}

// inclusive affine scan over the warp: lane j gets the composition of pages 0..j
__device__ __forceinline__ void affine_scan(uint64_t& a, uint64_t& b, int lane) {
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint64_t pa = __shfl_up_sync(0xffffffffu, a, o);
    const uint64_t pb = __shfl_up_sync(0xffffffffu, b, o);
    if (lane >= o) {
      b = pb * a + b;  // (pa, pb) first, then (a, b)
      a = pa * a;
    }
  }
}

// entry holding prefix hash h whose page tokens equal toks[off .. off + page), or -1
__device__ __forceinline__ int radix_probe(const b200attn_radix_t& r, uint64_t h, const void* toks, int tok_dtype,
                                           int64_t off) {
  const uint64_t key = radix_key(h);
  uint32_t idx = radix_bucket(h, r.tcap);
  for (int step = 0; step < r.tcap; ++step) {
    const uint64_t k = __ldcg(reinterpret_cast<const unsigned long long*>(r.tkeys) + idx);
    if (k == kEmpty) return -1;
    if (k == key) {
      const int e = __ldcg(r.tvals + idx);
      if (e > 0 && __ldcg(reinterpret_cast<const unsigned long long*>(r.ehash) + e) == h) {
        bool same = true;
        for (int i = 0; i < r.page && same; ++i)
          same = (int64_t)__ldcg(r.tokens + (int64_t)e * r.page + i) == load_index(toks, tok_dtype, off + i);
        if (same) return e;
      }
    }
    idx = (idx + 1) & (uint32_t)(r.tcap - 1);
  }
  return -1;
}

// One warp matches one request: returns the number of matched pages; optionally writes the slots.  `h_io` /
// `prev_io` leave the hash / entry at the end of the matched prefix (the insert path continues from there).
__device__ __forceinline__ int radix_match_warp(const b200attn_radix_t& r, const void* toks, int tok_dtype,
                                                int64_t beg, int n_pages, int64_t* out, int lane) {
  uint64_t h = kRadixSeed;
  int prev = 0;  // root
  int matched = 0;
  for (int base = 0; base < n_pages; base += 32) {
    const int j = base + lane;
    const bool valid = j < n_pages;
    uint64_t a = 1, b = 0;
    if (valid) page_poly(toks, tok_dtype, beg + (int64_t)j * r.page, r.page, a, b);
    affine_scan(a, b, lane);
    const uint64_t hj = h * a + b;
    const int e = valid ? radix_probe(r, hj, toks, tok_dtype, beg + (int64_t)j * r.page) : -1;
    int pe = __shfl_up_sync(0xffffffffu, e, 1);
    if (lane == 0) pe = prev;
    const bool ok = valid && e > 0 && pe >= 0 && __ldcg(r.parent + e) == pe;
    const unsigned bad = ~__ballot_sync(0xffffffffu, ok);
    const int lead = bad ? __ffs(bad) - 1 : 32;  // leading lanes whose chain holds
    if (out != nullptr && lane < lead) {
      for (int i = 0; i < r.page; ++i) out[(int64_t)j * r.page + i] = __ldcg(r.values + (int64_t)e * r.page + i);
    }
    matched += lead;
    if (lead < 32) break;
    h = __shfl_sync(0xffffffffu, hj, 31);
    prev = __shfl_sync(0xffffffffu, e, 31);
  }
  return matched;
}

__global__ void __launch_bounds__(32)
    radix_match_kernel(const b200attn_radix_t r, const void* toks, int tok_dtype, const int64_t* indptr,
                       int64_t* out, int64_t out_stride, int32_t* out_len) {
  const int b = blockIdx.x, lane = threadIdx.x;
  const int64_t beg = indptr[b];
  const int n_pages = (int)((indptr[b + 1] - beg) / r.page);  // the key is truncated to whole pages
  const int m = radix_match_warp(r, toks, tok_dtype, beg, n_pages, out + (int64_t)b * out_stride, lane);
  if (lane == 0) out_len[b] = m * r.page;
}


// One warp, request after request (inserts of one batch may share new prefixes: serial order = the reference's).
__global__ void __launch_bounds__(32)
    radix_insert_kernel(const b200attn_radix_t r, const void* toks, int tok_dtype, const int64_t* indptr,
                        const int64_t* values, int B, int32_t* out_prefix_len) {
  const int lane = threadIdx.x;
  for (int b = 0; b < B; ++b) {
    const int64_t beg = indptr[b];
    const int n_pages = (int)((indptr[b + 1] - beg) / r.page);
    uint64_t h = kRadixSeed;
    int prev = 0;
    int matched = 0;
    bool matching = true;
    for (int base = 0; base < n_pages; base += 32) {
      const int j = base + lane;
      const bool valid = j < n_pages;
      uint64_t a = 1, bb = 0;
      if (valid) page_poly(toks, tok_dtype, beg + (int64_t)j * r.page, r.page, a, bb);
      affine_scan(a, bb, lane);
      const uint64_t hj = h * a + bb;
      int e = -1;
      int lead = 0;
      if (matching) {
        e = valid ? radix_probe(r, hj, toks, tok_dtype, beg + (int64_t)j * r.page) : -1;
        int pe = __shfl_up_sync(0xffffffffu, e, 1);
        if (lane == 0) pe = prev;
        const bool ok = valid && e > 0 && pe >= 0 && __ldcg(r.parent + e) == pe;
        const unsigned bad = ~__ballot_sync(0xffffffffu, ok);
        lead = bad ? __ffs(bad) - 1 : 32;
        matched += lead;
        if (lead < 32) matching = false;
      }
      // lanes [lead, 32) with a valid page are NEW pages: allocate entry ids (free list first, then fresh ones)
      const bool fresh = valid && lane >= lead;
      const unsigned fmask = __ballot_sync(0xffffffffu, fresh);
      const int need = __popc(fmask);
      int id = e;  // matched lanes keep the entry they found
      if (need > 0) {
        int free_top = 0, bump = 0;
        if (lane == 0) {
          free_top = r.ctrs[1];
          bump = r.ctrs[0];
          const int from_free = min(need, free_top);
          const int from_bump = need - from_free;
          if (bump + from_bump > r.cap) {
            r.ctrs[3] = 1;  // out of entries: the caller sized `cap` too small
          } else {
            r.ctrs[1] = free_top - from_free;
            r.ctrs[0] = bump + from_bump;
            r.ctrs[2] += need;
          }
        }
        free_top = __shfl_sync(0xffffffffu, free_top, 0);
        bump = __shfl_sync(0xffffffffu, bump, 0);
        const int overflow = __shfl_sync(0xffffffffu, lane == 0 ? r.ctrs[3] : 0, 0);
        if (overflow) break;
        if (fresh) {
          const int rank = __popc(fmask & ((1u << lane) - 1u));
          const int from_free = min(need, free_top);
          id = rank < from_free ? r.freelist[free_top - 1 - rank] : bump + (rank - from_free);
        }
      }
      int pid = __shfl_up_sync(0xffffffffu, id, 1);
      if (lane == 0) pid = prev;
      if (fresh) {
        r.parent[id] = pid;
        reinterpret_cast<unsigned long long*>(r.ehash)[id] = hj;
        for (int i = 0; i < r.page; ++i) {
          r.tokens[(int64_t)id * r.page + i] = (int32_t)load_index(toks, tok_dtype, beg + (int64_t)j * r.page + i);
          r.values[(int64_t)id * r.page + i] = values[beg + (int64_t)j * r.page + i];
        }
        const int64_t kvpage = values[beg + (int64_t)j * r.page] / r.page;
        if (kvpage >= 0 && kvpage < r.num_pages) r.rev[kvpage] = id;
        __threadfence();  // the entry is complete before the table points at it
        const unsigned long long key = radix_key(hj);
        uint32_t idx = radix_bucket(hj, r.tcap);
        for (int step = 0; step < r.tcap; ++step) {
          unsigned long long* slot = reinterpret_cast<unsigned long long*>(r.tkeys) + idx;
          const unsigned long long k = *reinterpret_cast<volatile unsigned long long*>(slot);
          if (k == kEmpty || k == kTomb) {
            if (atomicCAS(slot, k, key) == k) {
              r.tvals[idx] = id;
              break;
            }
            continue;  // lost the slot to a lane of this warp: look at it again
          }
          idx = (idx + 1) & (uint32_t)(r.tcap - 1);
        }
      }
      __threadfence();
      __syncwarp();
      h = __shfl_sync(0xffffffffu, hj, 31);
      prev = __shfl_sync(0xffffffffu, id, 31);
    }
    if (lane == 0) out_prefix_len[b] = matched * r.page;
    __syncwarp();
  }
}

// slots[] = KV slots released by the allocator (the evicted leaves' values): entries whose first slot is among them
// are unlinked from the table and returned to the free list.
__global__ void __launch_bounds__(256)
    radix_erase_kernel(const b200attn_radix_t r, const void* slots, int slot_dtype, int64_t n) {
  const int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (i >= n) return;
  const int64_t slot = load_index(slots, slot_dtype, i);
  if (slot < 0 || slot % r.page != 0) return;
  const int64_t kvpage = slot / r.page;
  if (kvpage >= r.num_pages) return;
  const int e = r.rev[kvpage];
  if (e <= 0 || r.values[(int64_t)e * r.page] != slot) return;
  const unsigned long long h = reinterpret_cast<unsigned long long*>(r.ehash)[e];
  const unsigned long long key = radix_key(h);
  uint32_t idx = radix_bucket(h, r.tcap);
  for (int step = 0; step < r.tcap; ++step) {
    const unsigned long long k = reinterpret_cast<unsigned long long*>(r.tkeys)[idx];
    if (k == kEmpty) break;
    if (k == key && r.tvals[idx] == e) {
      reinterpret_cast<unsigned long long*>(r.tkeys)[idx] = kTomb;
      r.tvals[idx] = 0;
      break;
    }
    idx = (idx + 1) & (uint32_t)(r.tcap - 1);
  }
  r.rev[kvpage] = -1;
  r.parent[e] = -2;  // dead: no child can chain to it
  r.freelist[atomicAdd(&r.ctrs[1], 1)] = e;
  atomicSub(&r.ctrs[2], 1);
}

static int radix_check(const b200attn_radix_t* r, const char* what) {
  B2_CHECK_ARG(r != nullptr && r->parent && r->tokens && r->values && r->ehash && r->tkeys && r->tvals && r->rev &&
                   r->freelist && r->ctrs,
               "%s: null table pointer", what);
  B2_CHECK_ARG(r->cap > 1 && r->page >= 1 && r->num_pages >= 1 && r->tcap >= 2 * r->cap &&
                   (r->tcap & (r->tcap - 1)) == 0,
               "%s: bad table sizes (tcap must be a power of two >= 2 * cap)", what);
  return B200ATTN_OK;
}

}  // namespace b200attn

extern "C" int b200attn_radix_match(const b200attn_radix_t* table, const void* tokens, int tokens_dtype,
                                    const int64_t* indptr, int32_t B, int64_t* out_indices, int64_t out_stride,
                                    int32_t* out_len, void* stream) {
  using namespace b200attn;
  int rc = radix_check(table, "radix_match");
  if (rc) return rc;
  B2_CHECK_ARG(B >= 0, "radix_match: negative batch");
  if (B == 0) return B200ATTN_OK;
  B2_CHECK_ARG(tokens && indptr && out_indices && out_len, "radix_match: null pointer");
  radix_match_kernel<<<(unsigned)B, 32, 0, (cudaStream_t)stream>>>(*table, tokens, tokens_dtype, indptr, out_indices,
                                                                    out_stride, out_len);
  B2_LAUNCH_CHECK();
  return B200ATTN_OK;
}

extern "C" int b200attn_radix_insert(const b200attn_radix_t* table, const void* tokens, int tokens_dtype,
                                     const int64_t* indptr, const int64_t* values, int32_t B,
                                     int32_t* out_prefix_len, void* stream) {
  using namespace b200attn;
  int rc = radix_check(table, "radix_insert");
  if (rc) return rc;
  B2_CHECK_ARG(B >= 0, "radix_insert: negative batch");
  if (B == 0) return B200ATTN_OK;
  B2_CHECK_ARG(tokens && indptr && values && out_prefix_len, "radix_insert: null pointer");
  radix_insert_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(*table, tokens, tokens_dtype, indptr, values, B,
                                                          out_prefix_len);
  B2_LAUNCH_CHECK();
  return B200ATTN_OK;
}

extern "C" int b200attn_radix_erase_slots(const b200attn_radix_t* table, const void* slots, int slots_dtype,
                                          int64_t n, void* stream) {
  using namespace b200attn;
  int rc = radix_check(table, "radix_erase_slots");
  if (rc) return rc;
  B2_CHECK_ARG(n >= 0, "radix_erase_slots: negative size");
  if (n == 0) return B200ATTN_OK;
  B2_CHECK_ARG(slots != nullptr, "radix_erase_slots: null pointer");
  radix_erase_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(*table, slots, slots_dtype, n);
  B2_LAUNCH_CHECK();
  return B200ATTN_OK;
}
