/*
 * b200attn.h — C ABI of the B200-native (sm_100a) attention hot path.
 *
 * This is the drop-in boundary for SGLang's attention path.  Every entry point
 * names the reference interface it replaces (paths relative to the SGLang tree).
 *
 * Conventions (all entry points):
 *   - extern "C", plain pointers and sizes; no torch / C++ types.
 *   - return 0 on success, nonzero B200ATTN_ERR_* on failure; the message of the
 *     last failure on the calling thread is b200attn_last_error().
 *   - every launcher takes the CUDA stream as an opaque `void*` (cudaStream_t);
 *     nothing allocates, nothing synchronises, nothing is copied to the host:
 *     all launchers are CUDA-graph capturable.
 *   - all pointers are DEVICE pointers unless the name ends in `_host`.
 *   - strides are in ELEMENTS of the tensor's dtype unless the name says bytes.
 *   - caller owns every buffer, including workspaces (b200attn_*_workspace_*).
 */
#ifndef B200ATTN_H_
#define B200ATTN_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200ATTN_OK 0
#define B200ATTN_ERR_INVALID 1     /* bad argument (shape, alignment, dtype) */
#define B200ATTN_ERR_UNSUPPORTED 2 /* valid but not implemented (head_dim, ...) */
#define B200ATTN_ERR_CUDA 3        /* a CUDA runtime / driver call failed      */

/* element types of q/k/v/o and of the KV pool */
#define B200ATTN_BF16 0
#define B200ATTN_FP16 1
#define B200ATTN_FP32 2 /* merge_state only */

/* index dtypes */
#define B200ATTN_I32 0
#define B200ATTN_I64 1

/* ABI version: bumped whenever a signature below changes. */
int b200attn_abi_version(void);

/* Message of the last error raised on this thread ("" if none). */
const char* b200attn_last_error(void);

/* Number of SMs of the current device (grid sizing for callers; 148 on B200). */
int b200attn_sm_count(void);

/* --------------------------------------------------------------------------
 * (a1) KV store.  Replaces MHATokenToKVPool.set_kv_buffer -> _set_kv_buffer_impl
 *   python/sglang/srt/mem_cache/memory_pool.py:141-192, 2331-2456
 * and the JIT kernel it calls,
 *   python/sglang/kernels/jit/csrc/elementwise/kvcache.cuh:186-318 (store_kvcache)
 *   python/sglang/kernels/ops/kvcache/kvcache.py:56-110 (store_cache wrapper).
 *
 *   k_cache[loc[t]] = k[t];  v_cache[loc[t]] = v[t]   for t in [0, T)
 *
 * Rows are raw bytes (bit-exact copy, dtype agnostic).  `skip_index` mirrors the
 * reference kernel's `reserved_skip_index` (kvcache.cuh:215): rows whose loc ==
 * skip_index are not written; pass -1 to write every row (the naive
 * `k_cache[loc] = k` branch, memory_pool.py:190-192).  Slots outside
 * [0, size_limit) are never written; if `oob_count` is non-null the number of
 * such rows is atomically added to it (the reference device-asserts instead,
 * kvcache.cuh:207-209).
 * -------------------------------------------------------------------------- */
int b200attn_kv_store(const void* k, const void* v, void* k_cache, void* v_cache,
                      const void* loc, int loc_dtype, int64_t T,
                      int64_t k_row_bytes, int64_t v_row_bytes,
                      int64_t k_src_stride_bytes, int64_t v_src_stride_bytes,
                      int64_t k_cache_stride_bytes, int64_t v_cache_stride_bytes,
                      int64_t size_limit, int64_t skip_index, int32_t* oob_count,
                      void* stream);

/* --------------------------------------------------------------------------
 * (a3) Metadata.  Replaces TritonAttnBackend._fill_kv_indptr_and_indices
 *   python/sglang/srt/layers/attention/triton_backend.py:438-456
 *   python/sglang/kernels/ops/kvcache/kv_indices.py:8-44
 *     (create_flashinfer_kv_indices_triton)
 * and the oracle's gather req_to_token[req, start:end]
 *   python/sglang/srt/layers/attention/torch_native_backend.py:138-139,239-240.
 *
 *   kv_indptr[0] = 0; kv_indptr[b+1] = kv_indptr[b] + kv_lens[b]
 *   kv_indices[kv_indptr[b] + j] = req_to_token[req_pool_indices[b], kv_start[b] + j]
 *   qo_indptr (optional): same exclusive scan over qo_lens.
 *
 * kv_lens / qo_lens / req_pool_indices / kv_start may be int32 or int64
 * (`*_dtype`); kv_start and qo_lens/qo_indptr may be null.  kv_indices is int32
 * or int64 (`kv_indices_dtype`); kv_indptr / qo_indptr are int32.
 * `kv_indices_capacity` bounds the writes (entries beyond it are dropped).
 * `slot_map` (nullable, int32 / int64 `slot_map_dtype`): every slot id is replaced by
 * slot_map[slot] on the way out (hybrid sliding-window pools: the full -> SWA index
 * mapping of python/sglang/srt/mem_cache/swa_memory_pool.py:228-231).
 * One launch for B <= 4096 (each CTA reduces the lengths before its request itself).
 * -------------------------------------------------------------------------- */
int b200attn_build_kv_indices(const int32_t* req_to_token, int64_t req_to_token_stride,
                              const void* req_pool_indices, int req_pool_dtype,
                              const void* kv_lens, int kv_lens_dtype,
                              const void* kv_start, int kv_start_dtype,
                              const void* qo_lens, int qo_lens_dtype,
                              int32_t B,
                              int32_t* kv_indptr, void* kv_indices, int kv_indices_dtype,
                              int64_t kv_indices_capacity,
                              int32_t* qo_indptr, const void* slot_map, int slot_map_dtype,
                              void* stream);

/* Host <-> device staging by a kernel: dst0[0:bytes0] = src0, dst1[0:bytes1] = src1 in ONE launch (either segment
 * may be empty).  Either side of a segment may be PINNED HOST memory (mapped under UVA): this is how a step driven
 * from host buffers inside a CUDA graph refills its static device buffers and returns its result without memcpy
 * nodes - the role of the buffer refills in python/sglang/srt/model_executor/cuda_graph_runner.py (replay_prepare)
 * and runner_utils/buffers.py.  16-byte aligned buffers, sizes multiples of 16. */
int b200attn_stage_copy(void* dst0, const void* src0, int64_t bytes0, void* dst1, const void* src1,
                        int64_t bytes1, void* stream);

/* --------------------------------------------------------------------------
 * (f4) The collective after attention: sum all-reduce of the o_proj output over NVLS (in-switch reduction), one
 * launch per rank.  Replaces GroupCoordinator.all_reduce for this message
 * (python/sglang/srt/distributed/parallel_state.py:648-757) the way the reference's symmetric-memory communicator
 * does (device_communicators/torch_symm_mem.py:34-200: copy in, torch's multimem op, copy out = three launches).
 *   local_payload / mc_payload: this rank's instance and the MULTICAST address of the same symmetric allocation
 *   (torch.distributed._symmetric_memory: buffer_ptrs[rank] / multicast_ptr, plus the payload offset);
 *   flag_ptrs: HOST array of `world` device pointers, flag_ptrs[r] = rank r's flag block as mapped on THIS rank
 *   (b200attn_allreduce_flag_bytes() bytes each, zeroed once; self-resetting);
 *   in / out: local tensors staged into / out of the payload by the same kernel; either may be null (the producer
 *   wrote the payload itself / the consumer reads it).  bytes % 16 == 0, dtype BF16 / FP16 / FP32.
 *   local_small / mc_small (nullable): a second region of b200attn_allreduce_small_bytes() bytes in the same
 *   symmetric allocation; with it, messages up to 1 MB (and bytes * world <= 4 MB) with both `in` and `out` take the
 *   one-shot form: stage-in, ONE barrier, every rank reduces the whole message through the switch into `out`.
 * Every rank of the group must launch it with the same arguments' shape; at most 128 CTAs, all co-resident. */
int64_t b200attn_allreduce_flag_bytes(void);
int64_t b200attn_allreduce_small_bytes(void);
int b200attn_allreduce_nvls(const void* in, void* out, void* local_payload, void* mc_payload, void* local_small,
                            void* mc_small, void* const* flag_ptrs, int32_t rank, int32_t world, int64_t bytes,
                            int dtype, void* stream);

/* out[i] = slot_map[loc[i]] (int64 out): KVWriteLoc.swa_loc of a hybrid sliding-window pool,
 *   python/sglang/srt/mem_cache/memory_pool.py:1550-1590, swa_memory_pool.py:228-231,
 *   triton_backend.py:978-982. */
int b200attn_translate_slots(const void* loc, int loc_dtype, const void* slot_map,
                             int slot_map_dtype, int64_t n, int64_t* out, void* stream);

/* One decode ITEM per query token of an extend batch: req_ids[t] = the request of token t, item_lens[t] = the
 * number of leading keys it sees (causal: prefix + j + 1, else the request's full length; 0 for tokens past
 * qo_indptr[B]).  Feeds b200attn_decode (`req_ids`, `item_lens`) for head dims the tcgen05 extend kernel does not
 * take (torch_native_backend.py:61-174 semantics, one query row at a time). */
int b200attn_build_token_items(const int32_t* qo_indptr, const int32_t* kv_indptr, int32_t B, int64_t T,
                               int causal, int32_t* req_ids, int32_t* item_lens, void* stream);

/* Page-table form.  Replaces build_trtllm_mha_page_table
 *   python/sglang/kernels/ops/kvcache/trtllm_mha_page_table.py:40-120
 *   page_table[b, j] = req_to_token[req_pool_indices[b], j*page_size] / page_size
 * for j < ceil(seq_lens[b] / page_size); other entries are left untouched
 * (same as the reference, whose programs past the page count are guarded out). */
int b200attn_build_page_table(const int32_t* req_to_token, int64_t req_to_token_stride,
                              const void* req_pool_indices, int req_pool_dtype,
                              const void* seq_lens, int seq_lens_dtype, int32_t B,
                              int32_t page_size, int32_t* page_table,
                              int64_t page_table_stride, int32_t max_pages,
                              void* stream);

/* --------------------------------------------------------------------------
 * (a4) Paged GQA split-KV decode.  Replaces forward_decode of
 *   python/sglang/srt/layers/attention/torch_native_backend.py:176-277,338-398
 *   (oracle) and the production kernels
 *   python/sglang/kernels/ops/attention/decode_attention.py:539-775,910-994,1163-1239.
 *
 *   o[r, h, :] = softmax(sm_scale * q[r, h, :] . K_r[:, h/G, :]^T) . V_r[:, h/G, :]
 *   K_r = k_cache[kv_indices[kv_indptr[r] : kv_indptr[r+1]]]   (token-granular slots)
 *
 * q: [B, Hq, D] (row stride q_stride, head stride D); o likewise.
 * k_cache/v_cache: [slots, Hkv, D] NHD (token stride kv_stride, head stride D).
 * `q_rows` (nullable int32[B]): row of q/o used by request r (identity if null);
 * `req_ids` (nullable int32[B]): kv_indptr entry used by work item r (identity
 * if null) — together they let a MIXED batch route its extend_len==1 rows here.
 * `lse` (nullable float[B*Hq], indexed like o rows): natural-log LSE.
 * Workspace: b200attn_decode_workspace_bytes(B, Hq, D, max_splits) bytes, only
 * touched when max_splits > 1: a FIXED-SIZE header of per-unit arrival counters
 * (256 KiB = one int32 per (request, kv head group), at most 65536 units) followed by
 * the fp32 partials.  The split partials are merged INSIDE the kernel by the last CTA
 * of a unit to arrive (no second launch).  The workspace must be ZERO-INITIALISED ONCE
 * by the caller before its first use; every launch leaves the counters at zero again,
 * and because the header's size and position do not depend on (B, Hq, D, max_splits)
 * ONE buffer may be reused across launches of any shape (also by b200attn_decode_sk),
 * as long as launches that share it are stream-ordered.  The split policy is a pure function of
 * (kv length, max_splits, min_chunk) evaluated on the device, so CUDA-graph
 * replay needs no host-side recomputation.
 * `tail_begin` >= 0 restricts splitting to the units (request-major, then kv head
 * group) of rank >= tail_begin: the caller passes the number of units that fill whole
 * waves of resident CTAs (b200attn_decode_ctas_per_sm per SM), so that only the last, partial wave is split
 * `max_splits` ways and no SM idles for a whole unit (a function of B and the head
 * counts only => graph-safe).  -1: every unit uses the length-based policy.
 * Head dims: any multiple of 8 up to 256 (rows are D wide; the kernel pads to 64 / 128 / 256 on the way into shared
 * memory).  `item_lens` (nullable int32[B]): item r sees only the first item_lens[r] keys of its request
 * (see b200attn_build_token_items).
 * `warps` = 4 or 2 warps per CTA (16 keys per warp and tile): 4 streams fastest per CTA (steady state,
 * small batches); 2 puts four CTAs on an SM, which turns a batch of 297..592 units (cfg2: 512) into ONE
 * wave of equally fast CTAs instead of a full wave plus a partial one.
 *
 * Feature flags shared with b200attn_extend (triton_backend.py:1405-1416,
 * decode_attention.py:986-988, extend_attention.py:480-486,546-547,721-723):
 *   window_left >= 0 : a query at position q_id sees only keys kv_id >= q_id -
 *                      window_left (window_left + 1 keys; decode: q_id = L - 1);
 *                      whole tiles outside the window are never loaded.  -1 = off.
 *   logit_cap  > 0   : scores become logit_cap * tanh(sm_scale * qk / logit_cap).  0 = off.
 *   sinks (nullable float[Hq]): exp(sinks[h] - max) is added to the softmax
 *                      denominator of q head h (attention sinks).
 * -------------------------------------------------------------------------- */
int64_t b200attn_decode_workspace_bytes(int32_t B, int32_t Hq, int32_t D, int32_t max_splits);

/* Resident CTAs per SM of the kernel b200attn_decode launches with `warps` warps per CTA (4: 64-key tiles, 2 CTAs
 * per SM; 2: 64-thread CTAs with 32-key tiles, FOUR per SM = 592 resident slots on B200) or of
 * b200attn_decode_fp8kv (kv_fp8 = 1): the caller's split / tail-split wave arithmetic needs it. */
int b200attn_decode_ctas_per_sm(int kv_fp8, int warps);

/* b200attn_decode / b200attn_decode_fp8kv are launched with programmatic stream serialization: the grid may be
 * scheduled while the previous kernel of the stream drains and orders itself with griddepcontrol.wait before its
 * first global access.  Environment B200ATTN_PDL=0 restores plain stream order. */

int b200attn_decode(const void* q, int64_t q_stride, const void* k_cache,
                    const void* v_cache, int64_t kv_stride, void* o, int64_t o_stride,
                    float* lse, const int32_t* kv_indptr, const int32_t* kv_indices,
                    const int32_t* q_rows, const int32_t* req_ids, int32_t B,
                    int32_t Hq, int32_t Hkv, int32_t D, int dtype, float sm_scale,
                    int32_t max_splits, int32_t min_chunk, int32_t tail_begin, int32_t warps,
                    int32_t window_left, float logit_cap, const float* sinks,
                    const int32_t* item_lens, void* workspace, int64_t workspace_bytes, void* stream);

/* b200attn_decode without a caller workspace and without any device allocation.  NO COUNTERPART IN THE REFERENCE:
 * added for callers that cannot call cudaMalloc (code under CUDA graph
 * capture, tight serving loops, harnesses that forbid it).  Same arguments and results as b200attn_decode minus
 * `workspace` / `workspace_bytes`: the split-KV scratch is a file-scope `static __device__` array inside the library,
 * b200attn_decode_noalloc_max_workspace_bytes() = 256 KiB of counters + 8 MiB of partials, resident on every GPU
 * that loads the library until the process exits.  Shapes whose b200attn_decode_workspace_bytes(B, Hq, D,
 * max_splits) exceeds it are rejected with B200ATTN_ERR_INVALID and a message that says so (max_splits = 1 always
 * fits: it needs no scratch).  The entry point calls no cudaMalloc / cudaFree / allocating routine; it clears the
 * arrival counters it will use on `stream` before the launch.  All calls share the one array: calls that OVERLAP
 * on the GPU (different streams, graphs in flight at the same time) race on it - serialise them, or use
 * b200attn_decode with a workspace per stream. */
int64_t b200attn_decode_noalloc_max_workspace_bytes(void);
int b200attn_decode_noalloc(const void* q, int64_t q_stride, const void* k_cache, const void* v_cache,
                            int64_t kv_stride, void* o, int64_t o_stride, float* lse,
                            const int32_t* kv_indptr, const int32_t* kv_indices, const int32_t* q_rows,
                            const int32_t* req_ids, int32_t B, int32_t Hq, int32_t Hkv, int32_t D, int dtype,
                            float sm_scale, int32_t max_splits, int32_t min_chunk, int32_t tail_begin,
                            int32_t warps, int32_t window_left, float logit_cap, const float* sinks,
                            const int32_t* item_lens, void* stream);

/* --------------------------------------------------------------------------
 * (a4) Decode, flat tile-stream scheduling.  Same contract and arithmetic as
 * b200attn_decode (16-bit pools), but the batch is ONE stream of 64-key tiles cut
 * into equal contiguous ranges over 2 CTAs per SM, so every CTA moves the same
 * number of KV bytes whatever the batch shape; there is no host-side split policy
 * (`min_chunk` only bounds how finely short batches are cut).  Units cut across
 * CTAs are merged by the last piece to arrive.  Workspace:
 * b200attn_decode_sk_workspace_bytes(B, Hq, Hkv, D) bytes, zero-initialised once by
 * the caller (arrival counters; every launch leaves them at zero).
 * -------------------------------------------------------------------------- */
int64_t b200attn_decode_sk_workspace_bytes(int32_t B, int32_t Hq, int32_t Hkv, int32_t D);

int b200attn_decode_sk(const void* q, int64_t q_stride, const void* k_cache, const void* v_cache,
                       int64_t kv_stride, void* o, int64_t o_stride, float* lse,
                       const int32_t* kv_indptr, const int32_t* kv_indices, const int32_t* q_rows,
                       const int32_t* req_ids, int32_t B, int32_t Hq, int32_t Hkv, int32_t D,
                       int dtype, float sm_scale, int32_t min_chunk, int32_t window_left,
                       float logit_cap, const float* sinks, void* workspace,
                       int64_t workspace_bytes, void* stream);

/* --------------------------------------------------------------------------
 * (a4) Decode, warp-specialised persistent kernel fed by TMA (16-bit pools; the
 * default decode path of B200AttnBackend).  Same contract and arithmetic as
 * b200attn_decode_sk; the tile stream is cut by a guided schedule (one static chunk
 * of half the fair share per CTA, then ever smaller chunks drawn from a global
 * ticket, so CTAs that stream faster take more); one producer warp per CTA keeps a
 * 3-stage ring full through the copy engine (16-row tiled boxes when every aligned
 * 16-key group sits in 16 consecutive pool slots, tile::gather4 otherwise) and runs
 * ahead across unit boundaries, four consumer warps do the mma.sync math.
 * `num_slots` = rows of the pool (gather rows past it are zero-filled).
 * `new_k` / `new_v` (nullable, [B, Hkv, D] with token stride `new_kv_stride`
 * elements): the step's new K/V rows; the kernel writes row b into slot
 * kv_indices[kv_indptr[b+1] - 1] of the pool before reading it, replacing the
 * separate b200attn_kv_store launch of a decode layer-call (memory_pool.py:2331-2456:
 * store before read).  Rows whose slot == skip_slot are not written (-1: write all).
 * Needs identity q_rows / req_ids.  k_cache / v_cache are written only in that case.
 * Workspace: b200attn_decode_ws_workspace_bytes(Hq, Hkv, D) bytes, zero-initialised once
 * (fixed-size counter header first, like b200attn_decode).
 * -------------------------------------------------------------------------- */
int64_t b200attn_decode_ws_workspace_bytes(int32_t Hq, int32_t Hkv, int32_t D);

int b200attn_decode_ws(const void* q, int64_t q_stride, void* k_cache, void* v_cache,
                       int64_t kv_stride, int64_t num_slots, void* o, int64_t o_stride, float* lse,
                       const int32_t* kv_indptr, const int32_t* kv_indices, const int32_t* q_rows,
                       const int32_t* req_ids, int32_t B, int32_t Hq, int32_t Hkv, int32_t D,
                       int dtype, float sm_scale, int32_t min_chunk, int32_t window_left,
                       float logit_cap, const float* sinks, const void* new_k, const void* new_v,
                       int64_t new_kv_stride, int64_t skip_slot, void* workspace,
                       int64_t workspace_bytes, void* stream);

/* b200attn_decode_sk over an e4m3 byte pool (the byte path of b200attn_decode_fp8kv, 3 CTAs per SM). */
int b200attn_decode_sk_fp8kv(const void* q, int64_t q_stride, const void* k_cache, const void* v_cache,
                             int64_t kv_stride, void* o, int64_t o_stride, float* lse,
                             const int32_t* kv_indptr, const int32_t* kv_indices,
                             const int32_t* q_rows, const int32_t* req_ids, int32_t B, int32_t Hq,
                             int32_t Hkv, int32_t D, int dtype, float sm_scale, float k_descale,
                             float v_descale, int32_t min_chunk, int32_t window_left, float logit_cap,
                             const float* sinks, void* workspace, int64_t workspace_bytes,
                             void* stream);

/* --------------------------------------------------------------------------
 * (f4) FP8 (e4m3) KV cache.  Pool dtype float8_e4m3fn stored as bytes
 *   python/sglang/srt/mem_cache/memory_pool.py:2365-2375 (div by k_scale/v_scale in
 *     the input dtype, .to(float8_e4m3fn), uint8 view)
 *   python/sglang/srt/layers/attention/triton_backend.py:1344-1365 (store with
 *     layer.k_scale), :1418-1420 (k_descale / v_descale = layer.k_scale_float).
 *
 * b200attn_kv_store_fp8: k_cache[loc[t]] = e4m3(round_src(k[t] / k_scale)) (same for v),
 *   bit-exact against `(k / k_scale).to(torch.float8_e4m3fn)` on the CPU (true division,
 *   quotient rounded to the source dtype, c10 conversion: RNE, > 464 -> 0x7f, no
 *   saturation).  k_scale / v_scale <= 0: no division.  Row sizes and strides are in
 *   ELEMENTS; src_dtype is B200ATTN_BF16 or B200ATTN_FP16; loc / skip_index /
 *   size_limit / oob_count as in b200attn_kv_store.
 *
 * b200attn_decode_fp8kv: b200attn_decode over a byte pool (`kv_stride` in elements =
 *   bytes, multiple of 16).  q / o stay `dtype` (bf16 or fp16); scores are
 *   q . k8 * k_descale * sm_scale, the output is (softmax . v8) * v_descale.  The mma
 *   runs in f16 (e4m3 -> f16 is exact), accumulation and softmax in fp32.  Same
 *   workspace, split policy and feature flags as b200attn_decode.
 *
 * b200attn_kv_dequant_gather: out_k[i] = dtype(k8[kv_indices[i]] * k_mul), out_v[i] =
 *   dtype(v8[kv_indices[i]] * v_mul) for i in [0, n): the 16-bit K/V operand scratch the
 *   extend kernel reads when the pool is fp8 (call b200attn_extend on the scratch with
 *   kv_indices = 0..n-1 and sm_scale * k_descale when k_mul = 1).  `n_dev` (nullable, device int32): the live
 *   row count min(n, *n_dev) is read on the device, so one captured launch sized for the static buffers serves
 *   every replay (pass &kv_indptr[B]).
 * -------------------------------------------------------------------------- */
int b200attn_kv_store_fp8(const void* k, const void* v, void* k_cache, void* v_cache,
                          const void* loc, int loc_dtype, int64_t T, int64_t k_row_elems,
                          int64_t v_row_elems, int64_t k_src_stride, int64_t v_src_stride,
                          int64_t k_cache_stride, int64_t v_cache_stride, int src_dtype,
                          float k_scale, float v_scale, int64_t size_limit,
                          int64_t skip_index, int32_t* oob_count, void* stream);

int b200attn_decode_fp8kv(const void* q, int64_t q_stride, const void* k_cache,
                          const void* v_cache, int64_t kv_stride, void* o,
                          int64_t o_stride, float* lse, const int32_t* kv_indptr,
                          const int32_t* kv_indices, const int32_t* q_rows,
                          const int32_t* req_ids, int32_t B, int32_t Hq, int32_t Hkv,
                          int32_t D, int dtype, float sm_scale, float k_descale,
                          float v_descale, int32_t max_splits, int32_t min_chunk,
                          int32_t tail_begin, int32_t window_left, float logit_cap,
                          const float* sinks, void* workspace, int64_t workspace_bytes,
                          void* stream);

int b200attn_kv_dequant_gather(const void* k_cache, const void* v_cache, int64_t cache_stride,
                               const int32_t* kv_indices, int64_t n, int64_t row_elems,
                               void* out_k, void* out_v, int64_t out_stride, int out_dtype,
                               float k_mul, float v_mul, const int32_t* n_dev, void* stream);

/* De-duplicating form: a slot referenced by several rows of kv_indices (a radix-shared prefix) is converted once.
 * `first_ws` is an int32[num_slots] scratch the call resets itself; out_indices[i] (int32[n]) receives the scratch
 * row that holds row i's slot (its first referencing row): pass it to b200attn_extend as kv_indices over the scratch.
 * Rows of out_k / out_v that are not an owner row are left untouched. */
int b200attn_kv_dequant_gather_unique(const void* k_cache, const void* v_cache, int64_t cache_stride,
                                      const int32_t* kv_indices, int64_t n, int64_t row_elems,
                                      void* out_k, void* out_v, int64_t out_stride, int out_dtype,
                                      float k_mul, float v_mul, const int32_t* n_dev,
                                      int32_t* first_ws, int64_t num_slots, int32_t* out_indices,
                                      void* stream);

/* --------------------------------------------------------------------------
 * (a5) Ragged extend / prefill over cached prefix + new tokens.  Replaces
 *   python/sglang/srt/layers/attention/torch_native_backend.py:61-174,279-336
 *   (oracle) and python/sglang/kernels/ops/attention/extend_attention.py:304-923,
 *   963-1410 (unified single-pass variant).
 *
 * Request b owns q rows [qo_indptr[b], qo_indptr[b+1]) and keys
 * kv_indices[kv_indptr[b] : kv_indptr[b+1]] (prefix AND the new tokens, which
 * the caller has already stored into the pool — single KV source, like the
 * oracle, torch_native_backend.py:298-301,139-141).  With E = #q rows and
 * L = #keys, query j sits at absolute position L - E + j and attends keys
 * 0..L-E+j when `causal`, else all L keys (ENCODER_ONLY / bidirectional).
 * `causal` is a bit field: bit 0 = causal mask (B200ATTN_EXT_CAUSAL); bit 1 =
 * skip requests with exactly one query row (B200ATTN_EXT_SKIP_SINGLE): the
 * running-decode rows of a MIXED batch (schedule_batch.py:2758-2788), which the
 * caller sends through b200attn_decode (q_rows / req_ids) instead.
 *
 * The QK^T and PV contractions run on tcgen05 tensor cores with TMEM
 * accumulators; K/V tiles are fetched by TMA.  `page_size` is accepted for interface
 * parity and IGNORED: the kernel gathers token-granular slots and discovers per tile
 * whether every aligned 16-key group sits in 16 consecutive pool slots (true under the
 * paged allocator for page_size >= 16, paged.py:163-166) -> 16-row tiled TMA boxes;
 * otherwise (page_size 1, fragmented or ragged tiles) tile::gather4 rows.
 * `workspace` (nullable; b200attn_extend_workspace_bytes = 64 bytes, zeroed ONCE by the caller): the kernel's work
 * items are handed out by a device ticket counter that lives at the start of the workspace and is left at zero by
 * every launch; launches that share a workspace must not overlap (the same rule as b200attn_decode's).  Without a
 * workspace a process-wide pool of 1024 counter pairs is used round-robin: fine for launches in one stream, and
 * for concurrent streams as long as fewer than 1024 extend launches separate two overlapping ones.
 * `lse` nullable float[T*Hq] natural log.  `max_qo_len_host` / `total_q_host`
 * are host-side upper bounds used only to size the grid (graph-safe).
 * `custom_mask` (nullable uint8, nonzero = visible) with `mask_indptr` int32[B+1]:
 * TARGET_VERIFY tree masks of speculative decoding (extend_attention.py:467-479,
 * triton_backend.py:911-917): request b owns the row-major block
 * custom_mask[mask_indptr[b] + j * L_b + n]; `mask_skip_prefix` keeps the prefix
 * keys (n < L_b - E_b) always visible.  Combine with causal = 0.
 * -------------------------------------------------------------------------- */
#define B200ATTN_EXT_CAUSAL 1
#define B200ATTN_EXT_SKIP_SINGLE 2

int64_t b200attn_extend_workspace_bytes(int32_t B, int64_t total_q, int32_t Hq);

int b200attn_extend(const void* q, int64_t q_stride, const void* k_cache,
                    const void* v_cache, int64_t kv_stride, int64_t num_slots,
                    void* o, int64_t o_stride, float* lse, const int32_t* qo_indptr,
                    const int32_t* kv_indptr, const int32_t* kv_indices, int32_t B,
                    int64_t total_q_host, int32_t max_qo_len_host, int32_t Hq,
                    int32_t Hkv, int32_t D, int32_t page_size, int dtype, int causal,
                    float sm_scale, int32_t window_left, float logit_cap,
                    const float* sinks, const uint8_t* custom_mask, const int32_t* mask_indptr,
                    int mask_skip_prefix, void* workspace, int64_t workspace_bytes,
                    void* stream);

/* Extend over a float8_e4m3fn KV pool (memory_pool.py:2365-2375), in the kernel: k_cache / v_cache point at e4m3
 * bytes, `kv_stride_bytes` is the byte stride between slots.  What the reference's Triton kernels compute for such
 * a pool (extend_attention.py:523 `tl.dot(q.to(k.dtype), k)`, :596 `p = p.to(v.dtype)`; k / v descales as in
 * triton_backend.py:1418-1420): Q is cast to e4m3 in shared memory, P to e4m3 in tensor memory, both contractions
 * run as tcgen05.mma.kind::f8f6f4 on the pool's bytes, k_descale rides on the softmax scale and v_descale on the
 * output.  head_dim 64 or 128 (64-byte / 128-byte e4m3 rows); q / o stay bf16 / fp16 (`dtype`).  Everything else
 * as b200attn_extend. */
int b200attn_extend_fp8kv(const void* q, int64_t q_stride, const void* k_cache,
                          const void* v_cache, int64_t kv_stride_bytes, int64_t num_slots,
                          void* o, int64_t o_stride, float* lse, const int32_t* qo_indptr,
                          const int32_t* kv_indptr, const int32_t* kv_indices, int32_t B,
                          int64_t total_q_host, int32_t max_qo_len_host, int32_t Hq,
                          int32_t Hkv, int32_t D, int32_t page_size, int dtype, int causal,
                          float sm_scale, float k_descale, float v_descale, int32_t window_left,
                          float logit_cap, const float* sinks, const uint8_t* custom_mask,
                          const int32_t* mask_indptr, int mask_skip_prefix, void* workspace,
                          int64_t workspace_bytes, void* stream);

/* --------------------------------------------------------------------------
 * (f3) Radix prefix match on the device.  Replaces, for the attention path, the lookup side of
 *   python/sglang/srt/mem_cache/radix_cache.py: RadixCache.match_prefix (:377-423), insert (:425-445) and the
 *   slot release of evict (:593-621; which leaves to evict stays the scheduler's LRU policy).
 * The cache is a page trie in a caller-owned hash table (all device memory, zero-initialised once except
 * `rev` = -1 and ctrs[0] = 1): entry e > 0 = one cached page; entry 0 = the root.
 *   match : per request b, tokens[indptr[b] : indptr[b+1]) truncated to whole pages ->
 *           out_indices[b * out_stride + i] (int64 KV slots of the longest cached prefix), out_len[b] tokens.
 *           One warp per request, all pages of a request probed in parallel; bit-exact vs match_prefix.
 *   insert: request after request; pages already cached keep their slots, the rest are appended with
 *           values[indptr[b] + i]; out_prefix_len[b] = tokens that were already cached (InsertResult.prefix_len).
 *           ctrs[3] becomes 1 when `cap` entries do not suffice.
 *   erase_slots: KV slots handed back by the allocator (values of evicted leaves) -> their entries are unlinked.
 * tokens / slots are int32 or int64 (`*_dtype`).
 * -------------------------------------------------------------------------- */
typedef struct {
  int32_t* parent;    /* [cap]          parent entry (0 = root, -2 = dead)                      */
  int32_t* tokens;    /* [cap * page]   token ids of the entry's page                            */
  int64_t* values;    /* [cap * page]   KV slots of the entry's page                             */
  uint64_t* ehash;    /* [cap]          hash of the token prefix that ends with this page        */
  uint64_t* tkeys;    /* [tcap]         open addressing: 0 empty, 1 tombstone, else the hash     */
  int32_t* tvals;     /* [tcap]         entry id                                                 */
  int32_t* rev;       /* [num_pages]    KV page -> entry that holds it, -1 none                  */
  int32_t* freelist;  /* [cap]          recycled entry ids                                       */
  int32_t* ctrs;      /* [4]            next fresh id (init 1), free-list top, live entries, overflow flag */
  int32_t cap, tcap, page, num_pages;
} b200attn_radix_t;

int b200attn_radix_match(const b200attn_radix_t* table, const void* tokens, int tokens_dtype,
                         const int64_t* indptr, int32_t B, int64_t* out_indices, int64_t out_stride,
                         int32_t* out_len, void* stream);
int b200attn_radix_insert(const b200attn_radix_t* table, const void* tokens, int tokens_dtype,
                          const int64_t* indptr, const int64_t* values, int32_t B,
                          int32_t* out_prefix_len, void* stream);
int b200attn_radix_erase_slots(const b200attn_radix_t* table, const void* slots, int slots_dtype,
                               int64_t n, void* stream);

/* --------------------------------------------------------------------------
 * (a6) merge_state.  Replaces merge_state / merge_state_v2
 *   python/sglang/srt/layers/attention/merge_state.py:26-46
 *   python/sglang/kernels/aot/csrc/attention/merge_attn_states.cu:31-206
 *
 *   m = max(sa, sb); wa = e^(sa-m); wb = e^(sb-m)
 *   v = (wa*va + wb*vb) / (wa + wb);  s = log(wa + wb) + m      (natural log)
 *   +inf LSE is treated as -inf (merge_attn_states.cu:66-67).
 *
 * v_*: [T, H, D] contiguous, dtype bf16/fp16/fp32; s_*: float [T, H];
 * s_out nullable.  In-place (v_out == v_a, s_out == s_a) is allowed.
 * -------------------------------------------------------------------------- */
int b200attn_merge_state(const void* v_a, const float* s_a, const void* v_b,
                         const float* s_b, void* v_out, float* s_out, int64_t T,
                         int32_t H, int32_t D, int dtype, void* stream);

/* --------------------------------------------------------------------------
 * (f1) Fused RoPE / M-RoPE + KV store.  Replaces, in one pass over q/k/v,
 *   RotaryEmbedding.forward_native / forward_cuda
 *     python/sglang/srt/layers/rotary_embedding/base.py:236-276, 365-436
 *   MRotaryEmbedding.forward_native (sectioned and interleaved M-RoPE)
 *     python/sglang/srt/layers/rotary_embedding/mrope.py:47-51, 173-217
 *   apply_rope_with_cos_sin_cache_inplace(..., fused_args=FusedSetKVBufferArg)
 *     python/sglang/kernels/ops/attention/rope.py:98-113, 183-225
 *     python/sglang/kernels/jit/csrc/elementwise/rope.cuh:56-110, 178-260
 *   followed by the (a1) store of the rotated k and of v.
 *
 *   q[t,h,:rot], k[t,h,:rot] rotated IN PLACE by cos/sin of positions[t]
 *   (neox: pairs (i, i+rot/2); gpt-j: pairs (2i, 2i+1)); columns >= rot pass through.
 *   If k_cache != NULL: k_cache[loc[t]] = rotated k[t]; v_cache[loc[t]] = v[t]
 *   with the (a1) skip_index / size_limit / oob_count semantics.
 *
 * cos_sin_cache: [cache_rows, rot] = (cos | sin) halves per row, fp32 or the element
 * dtype (base.py:173-182).  positions: [T] (pos_axis_stride = 0, mrope_section_host
 * NULL) or [3, T] with `pos_axis_stride` elements between the t/h/w rows (>= T, or 0
 * when the three rows alias one [T] vector, i.e. an expanded text-only batch) and
 * mrope_section_host = 3 HOST ints summing to rot/2 (captured by value at launch).
 * Out-of-range positions are clamped to [0, cache_rows).
 *
 * mode: B200ATTN_ROPE_F32    fp32 products and sum, one rounding (rope.cuh:96-99)
 *       B200ATTN_ROPE_NATIVE cos/sin and every intermediate rounded to the element
 *                            dtype (torch bf16 arithmetic of utils.py:50-58)
 * Both are bit-exact against the oracle restatement of the respective reference path.
 * -------------------------------------------------------------------------- */
#define B200ATTN_ROPE_F32 0
#define B200ATTN_ROPE_NATIVE 1
int b200attn_rope_store(void* q, void* k, const void* v, void* k_cache, void* v_cache,
                        const void* loc, int loc_dtype, const void* cos_sin_cache,
                        int cache_dtype, int64_t cache_rows, const void* positions,
                        int pos_dtype, int64_t pos_axis_stride,
                        const int32_t* mrope_section_host, int mrope_interleaved, int64_t T,
                        int32_t Hq, int32_t Hkv, int32_t D, int32_t Dv, int32_t rotary_dim,
                        int is_neox, int64_t q_stride, int64_t k_stride, int64_t v_stride,
                        int64_t k_cache_stride, int64_t v_cache_stride, int64_t size_limit,
                        int64_t skip_index, int32_t* oob_count, int elem_dtype, int mode,
                        void* stream);

/* (f1 + f4) b200attn_rope_store over an e4m3 pool: same arguments plus the pool's scales.  q/k are rotated in
 * place exactly as b200attn_rope_store does; the pool rows are BYTES (cache strides in elements = bytes) and
 * receive e4m3(round_src(k_rot / k_scale)) and e4m3(round_src(v / v_scale)) - the bytes b200attn_kv_store_fp8
 * would write after the rotation (python/sglang/kernels/ops/kvcache/fused_fp8_qkv_kv_cache.py is the
 * reference's fused fp8 cache write).  k_scale / v_scale <= 0: no division.  k_cache must be non-null. */
int b200attn_rope_store_fp8(void* q, void* k, const void* v, void* k_cache, void* v_cache,
                            const void* loc, int loc_dtype, const void* cos_sin_cache,
                            int cache_dtype, int64_t cache_rows, const void* positions,
                            int pos_dtype, int64_t pos_axis_stride,
                            const int32_t* mrope_section_host, int mrope_interleaved,
                            int64_t T, int32_t Hq, int32_t Hkv, int32_t D, int32_t Dv,
                            int32_t rotary_dim, int is_neox, int64_t q_stride,
                            int64_t k_stride, int64_t v_stride, int64_t k_cache_stride,
                            int64_t v_cache_stride, int64_t size_limit, int64_t skip_index,
                            int32_t* oob_count, int elem_dtype, int mode, float k_scale,
                            float v_scale, void* stream);

/* --------------------------------------------------------------------------
 * (f3) Paged KV allocator on the device.  Replaces the Triton kernels
 *   alloc_extend_kernel  python/sglang/kernels/ops/memory/allocator.py:15-103
 *   alloc_decode_kernel  python/sglang/kernels/ops/memory/allocator.py:106-140
 * launched by PagedTokenToKVPoolAllocator.alloc_extend / alloc_decode
 *   python/sglang/srt/mem_cache/allocator/paged.py:172-260
 * (oracle: alloc_extend_naive, paged.py:44-102).
 *
 * extend: for request b (prefix P_b, new length S_b, last_loc_b = slot of token P_b-1)
 *   out[start_b + j] = slot of token P_b + j, j < S_b - P_b, start_b = sum_{i<b}(S_i - P_i):
 *   first the rest of the current partial page (last_loc + 1 + j), then whole pages
 *   popped in order from free_pages, then the head of one more page.
 * decode: S_b - 1 tokens are cached; out[b] = last_loc_b + 1 unless S_b - 1 is a
 *   multiple of page_size, in which case the next free page is opened.
 * The caller advances its free list by the number of new pages (host arithmetic on
 * the CPU copies of the lengths, as the reference does).  If fewer than the needed
 * pages are free, *status is set to 1 (when non-NULL) and the affected rows are left
 * unwritten; the reference returns None in that case (paged.py:214-215).
 * prefix_lens/seq_lens/last_loc share len_dtype; free_pages and out have their own.
 * -------------------------------------------------------------------------- */
int b200attn_alloc_extend(const void* prefix_lens, const void* seq_lens, const void* last_loc,
                          int len_dtype, const void* free_pages, int page_dtype, int64_t num_free,
                          void* out_indices, int out_dtype, int32_t B, int32_t page_size,
                          int32_t* status, void* stream);
int b200attn_alloc_decode(const void* seq_lens, const void* last_loc, int len_dtype,
                          const void* free_pages, int page_dtype, int64_t num_free,
                          void* out_indices, int out_dtype, int32_t B, int32_t page_size,
                          int32_t* status, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200ATTN_H_ */
