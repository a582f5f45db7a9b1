// Shared by the decode kernels (decode.cu: split-KV and fp8 variants; decode_sk.cu: flat tile-stream variant).
#pragma once
#include "common.cuh"

namespace b200attn {

constexpr int kDecTile = 64;     // keys per pipeline stage
constexpr int kDecStages = 3;    // cp.async ring depth
constexpr int kDecThreads = 128; // 4 warps x 16 keys
constexpr int kDecMaxG = 16;     // q heads per CTA (M of the mma)
// Split workspaces start with a FIXED-SIZE header of per-unit arrival counters (one int32 per (request, kv head
// group)), so their address does not depend on the launch shape: a caller may reuse one zero-initialised buffer
// across launches with different B / max_splits / D without stale partials ever aliasing a counter.
constexpr int kDecMaxUnits = 1 << 16;
constexpr int64_t kDecCounterBytes = (int64_t)kDecMaxUnits * 4;

struct DecodeParams {
  const void* q;
  const void* k_cache;
  const void* v_cache;
  void* o;
  float* lse;
  const int32_t* kv_indptr;
  const int32_t* kv_indices;
  const int32_t* q_rows;
  const int32_t* req_ids;
  float* ws_o;    // [B, Hq, max_splits, D]
  float* ws_lse;  // [B, Hq, max_splits]   (log2 domain)
  int* ws_cnt;    // [B, Hkv * groups] arrival counters of the split CTAs of a unit; zero on entry, zero on exit
  int64_t q_stride, kv_stride, o_stride;
  int B, Hq, Hkv, G, groups_per_kvh;
  int max_splits, min_chunk;
  float scale_log2;
  // feature flags: triton_backend.py:1405-1416, decode_attention.py:986-988
  int window_left;      // >= 0: only keys kv_id >= q_id - window_left (q_id = L - 1); < 0: off
  float cap_pre;        // logit cap c > 0: x = c * tanh(s * sm_scale / c); cap_pre = sm_scale / c
  float cap_post;       // c * log2(e); 0 = no cap
  const float* sinks;   // per q head sink logit (natural units) added to the softmax denominator, or null
  // Tail splitting: units (request, kv head group) are launched in rank order, `tail_begin` of them fill whole
  // waves of resident CTAs and stay unsplit; the remainder would leave SMs idle for up to a whole unit, so only
  // THOSE units are split `max_splits` ways.  tail_begin < 0: every unit uses the length-based policy.
  int tail_begin;
  float out_scale;  // fp8 KV only: v_descale * 2^8 applied to the normalised output (decode_fp8_kernel)
  // 16-bit split kernel only.  Dv: the REAL head dim (multiple of 8, <= the kernel's padded template D): rows of q /
  // k / v / o are Dv wide, columns >= Dv are zero-filled on the way into shared memory / registers and never
  // stored (head dims 80, 96, ... run on the D = 128 instantiation, 16, 32, ... on D = 64).
  int Dv;
  // item_lens (nullable int32[B]): number of leading keys of request req_ids[item] that item sees, instead of all
  // of them - one item per QUERY TOKEN of an extend batch (causal: prefix + j + 1), the generic-head-dim extend path
  const int32_t* item_lens;
};

// split cap of one unit (shared by the attention and the merge kernel)
__device__ __forceinline__ int unit_max_splits(const DecodeParams& p, int item, int hg) {
  if (p.tail_begin < 0) return p.max_splits;
  return item * (p.Hkv * p.groups_per_kvh) + hg >= p.tail_begin ? p.max_splits : 1;
}

// keys of a request that the (single) query row sees: [kv_lo, L)
__device__ __forceinline__ int window_lo(int L, int window_left) {
  return window_left >= 0 ? max(0, L - 1 - window_left) : 0;
}
__device__ __forceinline__ float tanh_approx(float x) {
  float y;
  asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

template <int D>
__device__ __forceinline__ uint32_t swz_off(int row, int chunk) {
  // byte offset of 16-byte chunk `chunk` of row `row` in a [rows][D] 2-byte tile;
  // chunks are XOR-swizzled with the row so that ldmatrix (8 rows x 16 B) is conflict free.
  constexpr int kChunks = D / 8;
  return (uint32_t)(row * (D * 2) + (((chunk ^ (row & 7)) & (kChunks - 1)) << 4));
}

// kFp8HwCvt: cvt.rn.f16x2.e4m3x2 (SASS F2FP.F16.E4M3.UNPACK_B, one instruction per pair, true values) instead of
// the 4-instruction bit trick (value * 2^-8).  kFp8Bias is the factor the scales must carry.
#ifndef B200ATTN_FP8_HWCVT
#define B200ATTN_FP8_HWCVT 1
#endif
constexpr bool kFp8HwCvt = B200ATTN_FP8_HWCVT != 0;
constexpr float kFp8Bias = kFp8HwCvt ? 1.f : 256.f;

__device__ __forceinline__ uint32_t e4m3x2_lo(uint32_t w) {  // bytes 0,1 -> f16x2
  uint32_t r;
  asm("{ .reg .b16 lo, hi; mov.b32 {lo, hi}, %1; cvt.rn.f16x2.e4m3x2 %0, lo; }" : "=r"(r) : "r"(w));
  return r;
}
__device__ __forceinline__ uint32_t e4m3x2_hi(uint32_t w) {  // bytes 2,3 -> f16x2
  uint32_t r;
  asm("{ .reg .b16 lo, hi; mov.b32 {lo, hi}, %1; cvt.rn.f16x2.e4m3x2 %0, hi; }" : "=r"(r) : "r"(w));
  return r;
}
__device__ __forceinline__ uint32_t e4m3x2_trick(uint32_t w, uint32_t sel) {
  const uint32_t t = __byte_perm(w, 0u, sel);  // the two selected bytes in the high byte of each half
  return (t & 0x80008000u) | ((t & 0x7f007f00u) >> 1);
}
// K word (4 consecutive d of one key): {d0,d1} and {d2,d3}
__device__ __forceinline__ void e4m3x4_pairs(uint32_t w, uint32_t& a, uint32_t& b) {
  if constexpr (kFp8HwCvt) {
    a = e4m3x2_lo(w);
    b = e4m3x2_hi(w);
  } else {
    a = e4m3x2_trick(w, 0x1404);
    b = e4m3x2_trick(w, 0x3424);
  }
}
// V word (byte0 = V[k0][d], byte1 = V[k0][d+1], byte2 = V[k0+1][d], byte3 = V[k0+1][d+1]): {k0,k0+1} at d and at d+1
__device__ __forceinline__ void e4m3x4_cross(uint32_t w, uint32_t& even, uint32_t& odd) {
  if constexpr (kFp8HwCvt) {
    const uint32_t t = __byte_perm(w, 0u, 0x3120);
    even = e4m3x2_lo(t);
    odd = e4m3x2_hi(t);
  } else {
    even = e4m3x2_trick(w, 0x2404);
    odd = e4m3x2_trick(w, 0x3414);
  }
}

template <int D>
__device__ __forceinline__ uint32_t swz8_off(int row, int chunk) {
  // byte offset of 16-byte chunk `chunk` of row `row` in a [rows][D] 1-byte tile (row pitch D bytes)
  constexpr int kChunks = D / 16;
  const int x = (kChunks >= 8) ? (row & 7) : ((row >> 1) & (kChunks - 1));
  return (uint32_t)(row * D + ((chunk ^ x) << 4));
}

#ifndef B200ATTN_FP8_STAGES
#define B200ATTN_FP8_STAGES 4
#endif
#ifndef B200ATTN_FP8_CTAS
#define B200ATTN_FP8_CTAS 3
#endif
constexpr int kDec8Stages = B200ATTN_FP8_STAGES;  // 4 x 16 KB per CTA, 3 CTAs per SM: 144 KB of loads in flight per SM

}  // namespace b200attn
