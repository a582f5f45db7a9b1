// (a5) Ragged extend / prefill attention on tcgen05 tensor cores (sm_100a).
//
// Semantics follow the oracle python/sglang/srt/layers/attention/torch_native_backend.py:61-174
// (prefix + new tokens all read from the pool through req_to_token; query j of a request
// with L keys and E query rows sits at absolute position L-E+j; causal or not).
//
// One CTA = TWO 128-row query tiles that share one kv head of one request, so every K/V tile
// fetched into shared memory feeds two S = Q K^T and two O += P V contractions.  The two tiles
// are consecutive entries of the request's "virtual tile" list v = qtile * G + g (g = q head
// inside the GQA group): for even G they are two q heads over the same query rows, for odd G
// (and MHA) a pair may straddle two row tiles.  Each tile has its own length bookkeeping.
//
// K/V arrive in 128-key smem tiles; S/P live in TMEM as two 64-key halves per tile (double buffer),
// so S(c+1) is computed while the softmax warpgroup still works on chunk c.
//
//   warps 0-3   softmax of tile 0      thread = query row = TMEM lane, per 64-key chunk c: tcgen05.ld
//   warps 4-7   softmax of tile 1      S, masking, exp2 online softmax (packed f32x2 FMAs, 3-input max,
//                                      40 % of the exp2 on the FMA pipe by polynomial, lazy rescale of O
//                                      in TMEM when the row max grew by > 2^8), P written back as bf16
//                                      INTO its S half (tcgen05.st), so the second GEMM reads its A
//                                      operand from TMEM; epilogue O / l.
// The kernel is persistent: one CTA per SM draws work items from the item list (item = (request, kv head,
// tile pair), heavy causal pairs first; two static items, then tickets), and every role keeps its pipeline state
// (K/V ring position, S/P half parity, Q/O phases) across items, so the TMA producers run ahead
// into the next item while the softmax warps still finish the current one.
//
//   warp 8      TMA producer: K tiles + the item scheduler (draws tickets, publishes descriptors)
//   warp 9      TMA producer: Q and V tiles.  Per 128-key tile a producer warp reads the 128 slot
//               ids and checks that every aligned 16-key group
//               sits in 16 consecutive pool slots (true under the paged allocator for page_size >=
//               16): yes -> 16-row tiled TMA boxes; no / ragged tail -> tile::gather4 (4 arbitrary
//               pool rows per instruction, OOB rows zero-filled).  Same 128B-swizzled layout.
//   warp 10/11  MMA issuers of tile 0 / tile 1 (one elected lane each; a single thread sustains only
//               ~1 UTCHMMA per 45 cycles).  Per chunk c: wait P_i(c); O_i += P_i(c) V_c (4 x M128 N128
//               K16, A from TMEM); S_i(c+2) = Q_i K_{c+2}^T (8 x M128 N64 K16) into the half P_i(c) just
//               vacated: correctness relies on the tensor pipe executing one thread's tcgen05.mma in
//               issue order (PV_i(c) reads P before QK_i(c+2) overwrites it).
// TMEM (512 columns): S0/P0 [0,128)  S1/P1 [128,256)  O0 [256,384)  O1 [384,512).
// e4m3 pools (kFp8): the same kernel on the pool's bytes - K/V tiles of 64 / 128-byte rows, Q rewritten as e4m3 by
// the softmax threads, P as e4m3 in TMEM, tcgen05.mma.kind::f8f6f4 (see the template's comment).
// Algorithmic FLOPs per (request, q head): 4 * D * sum_j (#visible keys of row j).
#include "sm100.cuh"

#include <atomic>
#include <mutex>
#include <type_traits>

namespace b200attn {

constexpr int kExtM = 128;        // query rows per tile (two tiles per CTA)
constexpr int kExtN = 128;        // keys per K/V smem tile (one TMA stage)
constexpr int kExtC = 64;         // keys per softmax / MMA chunk (S/P double-buffered in two 64-column halves)
constexpr int kExtStages = 2;     // K and V ring depth
constexpr int kExtThreads = 384;  // 2 softmax warpgroups + {K producer, V producer, MMA, idle}
constexpr int kExtBoxRows = 16;   // rows of one tiled K/V TMA box
#ifndef EXT_RESCALE_THRESHOLD
#define EXT_RESCALE_THRESHOLD 8.f
#endif
constexpr float kRescaleThreshold = EXT_RESCALE_THRESHOLD;  // log2 domain
constexpr int kRegsSoftmax = 208, kRegsOther = 80;
constexpr int kExtRing = 8;        // work-item descriptor ring (scheduler -> every role)
constexpr int kExtSchedSlots = 1024;  // launches in flight that may hold a ticket counter pair at the same time
constexpr int kExtOSlabCols = 32; // epilogue: O leaves through shared memory in 128 x 32 slabs (TMA store)

struct ExtendParams {
  void* o;
  float* lse;
  const int32_t* qo_indptr;
  const int32_t* kv_indptr;
  const int32_t* kv_indices;
  int64_t o_stride;
  int B, Hq, Hkv, G, num_slots;
  int causal;
  int skip_single;  // requests with one query row are left to the decode kernel
  float scale_log2;
  // feature flags: extend_attention.py:480-486,546-547,721-723
  int window_left;     // >= 0: query q_id sees only keys kv_id >= q_id - window_left; < 0: off
  float cap_pre;       // logit cap c > 0: s <- tanh(s * cap_pre) / cap_pre with cap_pre = sm_scale / c; 0 = off
  const float* sinks;  // per q head sink logit (natural units) added to the softmax denominator, or null
  int o_tma;           // full output tiles leave through shared memory + TMA store (o is 16-byte aligned)
  // TARGET_VERIFY custom masks (extend_attention.py:467-479,614-628): request b owns the row-major uint8 block
  // custom_mask[mask_indptr[b] + j * L_b + n] (query j, key n; nonzero = visible); with mask_skip_prefix the
  // prefix keys (n < L - E) are always visible
  const uint8_t* custom_mask;
  const int32_t* mask_indptr;
  int mask_skip_prefix;
  int sched_slot;  // this launch's {ticket, CTAs done} pair in g_ext_sched ...
  int* sched_ptr;  // ... or, when the caller passed a workspace, the pair at its start (nothing shared across callers)
  float out_scale;  // multiplies the normalised output: v_descale of an e4m3 pool, else 1
#ifdef EXT_TRACE
  long long* trace;  // dev-only timeline of CTA (0,0,0): [role 0..2][step 0..31][8]
#endif
};

// shared memory carve-up (bytes, from a 1024-aligned base)
template <int D>
struct ExtSmem {
  static constexpr int kTile = kExtN * D * 2;  // one 128 x D operand tile (bf16)
  static constexpr int kHalf = kExtN * 128;    // one 64-column half: 128 rows x 128 B
  static constexpr int kQ = 0;                 // [2] query tiles
  static constexpr int kK = kQ + 2 * kTile;
  static constexpr int kV = kK + kExtStages * kTile;
  static constexpr int kOSlab = kExtM * kExtOSlabCols * 2;                 // one 128 x 32 output slab (8 KB)
  static constexpr int kO = kV + kExtStages * kTile;                       // [2 tiles][2 buffers] output slabs
  static constexpr int kBar = kO + 4 * kOSlab;
  static constexpr int kItems = kBar + 512;                                // ItemDesc[kExtRing]
  static constexpr int kCum = kItems + kExtRing * 32;                      // int[kMaxB + 1] pair prefix sums
  // requests per launch (the host slices larger batches); D = 128 leaves room for 128 (227 KB shared memory)
  static constexpr int kMaxB = D == 128 ? 128 : 2048;
  static constexpr int kTotal = kCum + (kMaxB + 1) * 4 + 12;
};

// One work item = two query tiles of one (request, kv head); E < 0 = no more work for this CTA.
struct ItemDesc {
  int q_beg, E, kv_beg, L, kvh, pair, pad0, pad1;
};

// what every role derives from an ItemDesc
struct ItemTiles {
  int h[2], m0[2], rows[2], kvend[2], nc[2];  // per tile: q head, first row, valid rows, keys, 64-key chunks
  int prefix, kv_end, ntiles, nc_max;
  int kbase;  // first key any row of the pair can see, rounded down to a K/V tile (0 without a sliding window)
};

enum ExtBar {
  kBarQ = 0,                            // [2]
  kBarKFull = kBarQ + 2,                // [stages]
  kBarKEmpty = kBarKFull + kExtStages,
  kBarVFull = kBarKEmpty + kExtStages,
  kBarVEmpty = kBarVFull + kExtStages,
  kBarSFull = kBarVEmpty + kExtStages,  // [2 tiles][2 halves]  MMA -> softmax_i : S_i(c) is in TMEM
  kBarPFull = kBarSFull + 4,            // [2][2]  softmax_i -> MMA : P_i(c) is in TMEM, O_i rescaled
  kBarPVDone = kBarPFull + 4,           // [2]  MMA -> softmax_i : PV_i(c) done (only waited on before a rescale)
  kBarOFinal = kBarPVDone + 2,          // [2]  MMA -> softmax_i : last PV_i of the item done
  kBarQEmpty = kBarOFinal + 2,          // [2]  MMA -> producer : last QK_i of the item done, Q_i tile is free
  kBarItemFull = kBarQEmpty + 2,        // [ring]  scheduler (K producer warp) -> every role : descriptor n is written
  kBarItemEmpty = kBarItemFull + kExtRing,  // [ring]  every other role -> scheduler : descriptor n has been read
  kBarQ8 = kBarItemEmpty + kExtRing,    // [2]  softmax_i -> MMA_i (e4m3 pools): the e4m3 copy of Q_i is in shared memory
  kBarCount = kBarQ8 + 2
};
static_assert(kBarCount * 8 <= 512, "barrier block");

// {next ticket, CTAs that drew their last ticket} per launch slot.  Zero at module load; the last CTA of a launch
// to run out of work re-arms its pair, so a replayed CUDA graph finds it zero again.
__device__ int g_ext_sched[kExtSchedSlots * 2];

// O_i *= alpha (this thread's row; D fp32 columns of TMEM).  Rare (row max grew by more than the lazy-rescale
// threshold) and deliberately NOT inlined: the softmax chunk loop stays short.
template <int D>
__device__ __noinline__ void rescale_o(uint32_t col_o, float alpha) {
  using namespace sm100;
#pragma unroll
  for (int q = 0; q < D / 32; ++q) {
    uint32_t v[32];
    tmem_ld32(col_o + q * 32, v);
    tmem_wait_ld();
#pragma unroll
    for (int e = 0; e < 32; ++e) v[e] = __float_as_uint(__uint_as_float(v[e]) * alpha);
    tmem_st32(col_o + q * 32, v);
  }
}

// ---- packed fp32x2 helpers (FFMA2 / FADD2 / FMNMX3 on sm_100) ----
__device__ __forceinline__ uint64_t f2_pack(float lo, float hi) {
  uint64_t r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void f2_unpack(uint64_t v, float& lo, float& hi) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ uint64_t f2_fma(uint64_t a, uint64_t b, uint64_t c) {
  uint64_t r;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
  return r;
}
__device__ __forceinline__ uint64_t f2_add(uint64_t a, uint64_t b) {
  uint64_t r;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ uint64_t f2_mul(uint64_t a, uint64_t b) {
  uint64_t r;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ float fmax3(float a, float b, float c) {
  float r;
  asm("max.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c));
  return r;
}
#ifdef EXT_TRACE
__device__ __forceinline__ long long clk() {
  long long t;
  asm volatile("mov.u64 %0, %%clock64;" : "=l"(t)::"memory");
  return t;
}
#define TRACE(role, step, slot)                                                              \
  if (trace_on && (step) < 32) p.trace[((role) * 32 + (step)) * 8 + (slot)] = clk()
#else
#define TRACE(role, step, slot)
#endif
__device__ __forceinline__ float tanh_approx_f(float x) {
  float y;
  asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// exp2 of two values on the FMA pipe (no MUFU): round-to-nearest range reduction x = j + f,
// f in [-0.5, 0.5], degree-3 minimax polynomial for 2^f (max relative error 7.5e-5, far below the
// bf16 rounding of P), then j is added into the exponent field.  x is clamped at -125 so that
// masked (-inf) and far-below-max scores give ~1e-38 instead of wrapping the exponent.
#ifndef EXT_EMU_PERIOD
#define EXT_EMU_PERIOD 5
#endif
#ifndef EXT_EMU_COUNT
#define EXT_EMU_COUNT 2
#endif
constexpr int kEmuPeriod = EXT_EMU_PERIOD, kEmuCount = EXT_EMU_COUNT;  // pairs q with q % 5 < 2 are emulated (40 %)
__device__ __forceinline__ void exp2_emu2(float x0, float x1, float& p0, float& p1) {
  constexpr float kMagic = 12582912.f;  // 1.5 * 2^23: low mantissa bits of (x + kMagic) = round(x)
  x0 = fmaxf(x0, -125.f);
  x1 = fmaxf(x1, -125.f);
  const uint64_t x = f2_pack(x0, x1);
  const uint64_t t = f2_add(x, f2_pack(kMagic, kMagic));
  const uint64_t fl = f2_add(t, f2_pack(-kMagic, -kMagic));
  const uint64_t f = f2_fma(fl, f2_pack(-1.f, -1.f), x);
  uint64_t r = f2_fma(f2_pack(0.055171459913253784f, 0.055171459913253784f), f,
                      f2_pack(0.2426108568906784f, 0.2426108568906784f));
  r = f2_fma(r, f, f2_pack(0.6932609677314758f, 0.6932609677314758f));
  r = f2_fma(r, f, f2_pack(0.9999281167984009f, 0.9999281167984009f));
  float t0, t1, r0, r1;
  f2_unpack(t, t0, t1);
  f2_unpack(r, r0, r1);
  p0 = __uint_as_float(__float_as_uint(r0) + (__float_as_uint(t0) << 23));
  p1 = __uint_as_float(__float_as_uint(r1) + (__float_as_uint(t1) << 23));
}

__device__ __forceinline__ void derive_tiles(const ItemDesc& d, const ExtendParams& p, ItemTiles& t) {
  const int num_qtiles = (d.E + kExtM - 1) / kExtM;
  const int nv = num_qtiles * p.G;  // virtual tiles v = qtile * G + g of this (request, kv head)
  t.prefix = d.L - d.E;
  int lo = d.L;
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    const int v = 2 * d.pair + i;
    const bool ok = v < nv;
    t.h[i] = d.kvh * p.G + (ok ? v % p.G : 0);
    t.m0[i] = (ok ? v / p.G : 0) * kExtM;
    t.rows[i] = ok ? min(kExtM, d.E - t.m0[i]) : 0;
    t.kvend[i] = !ok ? 0 : (p.causal ? min(d.L, t.prefix + t.m0[i] + t.rows[i]) : d.L);
    if (ok) lo = min(lo, t.prefix + t.m0[i] - p.window_left);
  }
  t.kbase = p.window_left >= 0 ? max(0, lo) / kExtN * kExtN : 0;
#pragma unroll
  for (int i = 0; i < 2; ++i) t.nc[i] = max(0, t.kvend[i] - t.kbase + kExtC - 1) / kExtC;
  t.kv_end = max(t.kvend[0], t.kvend[1]);
  t.ntiles = max(0, t.kv_end - t.kbase + kExtN - 1) / kExtN;
  t.nc_max = max(t.nc[0], t.nc[1]);
}

// kFp8: the KV pool holds e4m3 bytes (memory_pool.py:2365-2375).  Q is converted to e4m3 in shared memory and P is
// written to TMEM as e4m3, and both contractions run as tcgen05.mma.kind::f8f6f4 (K = 32 per instruction) straight
// on the pool's bytes: what the reference's Triton kernels compute for such a pool (extend_attention.py:523
// `tl.dot(q.to(k.dtype), k)`, :596 `p = p.to(v.dtype)`), with k_descale in the softmax scale and v_descale in the
// output scale (triton_backend.py:1418-1420).  No dequantised scratch copy of the pool rows.  A K/V row is exactly
// one swizzle row: 128 bytes at head_dim 128, 64 bytes (SWIZZLE_64B maps and descriptors) at head_dim 64.
template <int D, bool kBf16, bool kFp8 = false>
__global__ void __launch_bounds__(kExtThreads, 1)
    extend_kernel(const __grid_constant__ CUtensorMap tm_q, const __grid_constant__ CUtensorMap tm_k,
                  const __grid_constant__ CUtensorMap tm_v, const __grid_constant__ CUtensorMap tm_kbox,
                  const __grid_constant__ CUtensorMap tm_vbox, const __grid_constant__ CUtensorMap tm_o,
                  const ExtendParams p) {
  using namespace sm100;
  using S = ExtSmem<D>;
  constexpr int kHalves = D / 64;  // 64-column halves of the head dim (16-bit tiles: Q always, K/V of 16-bit pools)
  // K / V / e4m3-Q tile row pitch: 128 bytes (16-bit tiles in 64-column halves; e4m3 rows of head_dim 128), or the
  // 64-byte e4m3 rows of head_dim 64 (SWIZZLE_64B tensor maps and matrix descriptors)
  constexpr int kRowB = kFp8 ? D : 128;
  constexpr bool kSw64 = kFp8 && D == 64;
  constexpr int kHalvesKV = kFp8 ? 1 : kHalves;                  // 128-byte column groups of a K/V row
  constexpr uint32_t kKVTileBytes = kFp8 ? kExtN * D : S::kTile;  // bytes of one K or V tile as loaded
  // e4m3 copy of Q tile i: the upper (unused) half of K ring stage i
  constexpr uint32_t kQ8Off = S::kK + kExtN * D;
  constexpr uint32_t kTmemCols = 512;
  constexpr uint32_t kColS = 0, kColO = 256;  // + i * 128

  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem0 = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem = smem_raw + (smem0 - smem_u32(smem_raw));
  const uint32_t bar0 = smem0 + S::kBar;
  auto bar = [&](int i) { return bar0 + 8u * i; };
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + S::kBar + 8 * kBarCount);
  ItemDesc* items = reinterpret_cast<ItemDesc*>(smem + S::kItems);
  int* cum = reinterpret_cast<int*>(smem + S::kCum);  // cum[b] = tile pairs of requests < b

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#ifdef EXT_TRACE
  const bool trace_on = p.trace != nullptr && blockIdx.x == 0 &&
                        (threadIdx.x == 0 || threadIdx.x == 128 || (warp >= 10 && lane == 0));
#endif

#ifdef EXT_TRACE
  if (p.trace != nullptr && blockIdx.x == 0 && threadIdx.x == 0) {
    long long gt;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gt));
    p.trace[(3 * 32 + 0) * 8 + 0] = clk();
    p.trace[(3 * 32 + 0) * 8 + 1] = gt;
  }
  if (p.trace != nullptr && threadIdx.x == 0) {  // every CTA: start / end wall time (tools/trace_extend_ctas.py)
    long long gt;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gt));
    p.trace[6 * 32 * 8 + blockIdx.x * 2] = gt;
  }
#endif
  if (threadIdx.x == 0) {
    for (int s = 0; s < kExtStages; ++s) {
      mbar_init(bar(kBarKFull + s), 1);
      mbar_init(bar(kBarKEmpty + s), 2);  // one release per MMA issuer
      mbar_init(bar(kBarVFull + s), 1);
      mbar_init(bar(kBarVEmpty + s), 2);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(bar(kBarQ + i), 1);
      for (int hb = 0; hb < 2; ++hb) {
        mbar_init(bar(kBarSFull + i * 2 + hb), 1);
        mbar_init(bar(kBarPFull + i * 2 + hb), 128);
      }
      mbar_init(bar(kBarPVDone + i), 1);
      mbar_init(bar(kBarOFinal + i), 1);
      mbar_init(bar(kBarQEmpty + i), 1);
    }
    for (int i = 0; i < 2; ++i) mbar_init(bar(kBarQ8 + i), 128);
    for (int s = 0; s < kExtRing; ++s) {
      mbar_init(bar(kBarItemFull + s), 1);
      mbar_init(bar(kBarItemEmpty + s), 11);  // V producer, two MMA issuers, eight softmax warps
    }
    mbar_fence_init();
  }
  if (warp == 10) tmem_alloc<kTmemCols>(smem_u32(tmem_slot));
  if (warp == 8 && lane == 0) {
    tma_prefetch_desc(&tm_k);
    tma_prefetch_desc(&tm_kbox);
  }
  if (warp == 9 && lane == 0) {
    tma_prefetch_desc(&tm_q);
    tma_prefetch_desc(&tm_v);
    tma_prefetch_desc(&tm_vbox);
  }
  // ---- tile-pair prefix sums over the batch (every CTA builds the same table) ----
  for (int b = threadIdx.x; b < p.B; b += kExtThreads) {
    const int E = p.qo_indptr[b + 1] - p.qo_indptr[b];
    const int nv = ((E + kExtM - 1) / kExtM) * p.G;
    cum[b + 1] = (E <= 0 || (p.skip_single && E == 1)) ? 0 : (nv + 1) >> 1;
  }
  if (threadIdx.x == 0) cum[0] = 0;
  __syncthreads();
  if (warp == 0) {
    int carry = 0;
    for (int base = 0; base < p.B; base += 32) {
      int v = (base + lane < p.B) ? cum[base + lane + 1] : 0;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int u = __shfl_up_sync(0xffffffffu, v, o);
        if (lane >= o) v += u;
      }
      if (base + lane < p.B) cum[base + lane + 1] = v + carry;
      carry += __shfl_sync(0xffffffffu, v, 31);
    }
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  const int total_items = cum[p.B] * p.Hkv;
  // Dynamic schedule: CTA x starts on items x and 2 * grid - 1 - x, every further item is a ticket drawn from a
  // device counter, in list order (request, kv head, heavy pairs first: neighbours in time share a kv head's K/V
  // tiles in L2, and the last tickets are the light pairs).  A static split left the CTAs finishing 4 % (32 x 2.5 K prefill) to 16 %
  // (16 x 1280, 28/4 heads) apart: per-item cost is not proportional to its keys and SMs do not stream at one rate.
  // The K producer warp is the scheduler: it draws the ticket of item n+2 when it starts item n (the atomic's latency
  // hides behind the item's K loads) and publishes that descriptor when the loads are issued, one whole item before
  // any role needs it (published one item ahead only, Q(n+1) landed ~1.4 us late: 1131 -> 1092 TFLOP/s on cfg3).
  int* const sched = p.sched_ptr != nullptr ? p.sched_ptr : g_ext_sched + 2 * p.sched_slot;
  auto describe = [&](int id) {  // id -> (request, kv head, pair)
    ItemDesc d;
    d.q_beg = d.kv_beg = d.L = d.kvh = d.pair = d.pad0 = d.pad1 = 0;
    d.E = -1;
    if (id >= total_items) return d;
    int lo = 0, hi = p.B;  // largest b with cum[b] * Hkv <= id
    while (hi - lo > 1) {
      const int mid = (lo + hi) >> 1;
      if (cum[mid] * p.Hkv <= id) lo = mid; else hi = mid;
    }
    const int b = lo;
    const int np = cum[b + 1] - cum[b];
    const int rem = id - cum[b] * p.Hkv;
    d.q_beg = p.qo_indptr[b];
    d.E = p.qo_indptr[b + 1] - d.q_beg;
    d.kv_beg = p.kv_indptr[b];
    d.L = p.kv_indptr[b + 1] - d.kv_beg;
    d.kvh = rem / np;
    d.pair = np - 1 - rem % np;
    d.pad0 = p.custom_mask != nullptr ? p.mask_indptr[b] : 0;  // offset of this request's mask block
    return d;
  };
  // scheduler side: wait until every role has read the descriptor that lived in this ring slot, then overwrite it
  auto publish = [&](int n, const ItemDesc& d) {
    const int slot = n % kExtRing;
    if (n >= kExtRing) mbar_wait(bar(kBarItemEmpty + slot), ((n / kExtRing) + 1) & 1);
    if (lane == 0) {
      items[slot] = d;
      mbar_arrive(bar(kBarItemFull + slot));  // release: the descriptor is visible to whoever sees the phase flip
    }
    __syncwarp();
  };
  // reader side (whole warps): descriptor n, acknowledged as soon as it sits in registers
  auto acquire = [&](int n) {
    const int slot = n % kExtRing;
    mbar_wait(bar(kBarItemFull + slot), (n / kExtRing) & 1);
    const ItemDesc d = items[slot];
    __syncwarp();
    if (lane == 0) mbar_arrive(bar(kBarItemEmpty + slot));
    return d;
  };

  // ---- per-role pipeline state, carried across items and rounds ----
  int kt = 0;              // K/V tiles consumed / produced so far (ring position)
  int g_my = 0;            // softmax WG i / issuer i: 64-key chunks of tile i so far (S/P half parity, PVDone phase)
  int nq_my = 0;           // softmax WG i / issuer i: items in which tile i was valid (Q / OFinal phases)
  int nq_p[2] = {0, 0};    // K producer: same count per tile (Q / QEmpty phases)

  // Role branch outermost (ptxas sizes registers per setmaxnreg region); both sides run the same
  // round loop and meet at the two barriers of every round.
  if (warp >= 8) {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;\n" ::"n"(kRegsOther));
    {
      if (warp == 8 || warp == 9) {
        // ===================== TMA producers (warp 8: Q + K, warp 9: V) =====================
        const bool is_k = (warp == 8);
        const CUtensorMap* tm_g = is_k ? &tm_k : &tm_v;
        const CUtensorMap* tm_b = is_k ? &tm_kbox : &tm_vbox;
        const uint32_t tile0 = smem0 + (is_k ? S::kK : S::kV);
        const int bar_full = is_k ? kBarKFull : kBarVFull, bar_empty = is_k ? kBarKEmpty : kBarVEmpty;
        ItemDesc d;
        int ticket = 0;
        if (is_k) {
          publish(0, describe((int)blockIdx.x));
          publish(1, describe(2 * (int)gridDim.x - 1 - (int)blockIdx.x));
        }
        for (int n = 0;; ++n) {
          if (is_k) {
            d = items[n % kExtRing];  // written by this warp's lane 0 (program order + __syncwarp)
            if (d.E < 0) break;
            if (lane == 0) ticket = atomicAdd(sched, 1);  // item n + 2; needed only after this item's loads are issued
          } else {
            d = acquire(n);
            if (d.E < 0) break;
          }
          ItemTiles t;
          derive_tiles(d, p, t);
          // The Q tiles are loaded by the V producer: the K producer must NOT wait for the previous item's last
          // QK (QEmpty) — it runs ahead into this item's first K tiles (ring permitting) while the previous item
          // drains, so QK(0) can issue as soon as Q lands.  V tiles are needed a softmax later than K tiles, so the
          // V producer has the slack to sit on that barrier.
          if (!is_k) {
#pragma unroll
            for (int i = 0; i < 2; ++i) {
              if (t.nc[i] == 0) continue;
              if (lane == 0) {
                if (nq_p[i] > 0) mbar_wait(bar(kBarQEmpty + i), (nq_p[i] - 1) & 1);  // previous item's QK_i done
                mbar_arrive_expect_tx(bar(kBarQ + i), S::kTile);
#pragma unroll
                for (int hf = 0; hf < kHalves; ++hf)
                  tma_load_2d(smem0 + S::kQ + i * S::kTile + hf * S::kHalf, &tm_q, bar(kBarQ + i),
                              t.h[i] * D + hf * 64, d.q_beg + t.m0[i]);
              }
              ++nq_p[i];
            }
            __syncwarp();
          }
          const int32_t* idx = p.kv_indices + d.kv_beg;
          const int col = kFp8 ? d.kvh * (D / 2) : d.kvh * D;  // the tensor maps count 16-bit columns
          int r[4];
          auto load_slots = [&](int j) {
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              const int key = t.kbase + j * kExtN + lane * 4 + i;
              r[i] = (j < t.ntiles && key < t.kv_end) ? __ldg(idx + key) : p.num_slots;  // OOB row -> zeros
            }
          };
          load_slots(0);
          for (int j = 0; j < t.ntiles; ++j, ++kt) {
            const int st = kt % kExtStages;
            const uint32_t dst = tile0 + st * S::kTile;
            // every aligned 16-key group in 16 consecutive slots?  (4 lanes x 4 keys per group)
            const int g0 = __shfl_sync(0xffffffffu, r[0], lane & ~3);
            bool contig = (r[1] == r[0] + 1) && (r[2] == r[0] + 2) && (r[3] == r[0] + 3) &&
                          (r[0] == g0 + 4 * (lane & 3));
            contig = __all_sync(0xffffffffu, contig) && (t.kbase + (j + 1) * kExtN <= t.kv_end);
            if (kt >= kExtStages) mbar_wait(bar(bar_empty + st), ((kt / kExtStages) + 1) & 1);
            if (lane == 0) mbar_arrive_expect_tx(bar(bar_full + st), kKVTileBytes);
            __syncwarp();
            if (contig) {
              int rows[kExtN / kExtBoxRows];
#pragma unroll
              for (int g = 0; g < kExtN / kExtBoxRows; ++g) rows[g] = __shfl_sync(0xffffffffu, r[0], g * 4);
              if (lane == 0) {
#pragma unroll
                for (int g = 0; g < kExtN / kExtBoxRows; ++g)
#pragma unroll
                  for (int hf = 0; hf < kHalvesKV; ++hf)
                    tma_load_2d(dst + hf * S::kHalf + g * kExtBoxRows * kRowB, tm_b, bar(bar_full + st),
                                col + hf * 64, rows[g]);
              }
            } else {
#pragma unroll
              for (int hf = 0; hf < kHalvesKV; ++hf)
                tma_gather4(dst + hf * S::kHalf + lane * 4 * kRowB, tm_g, bar(bar_full + st), col + hf * 64,
                            r[0], r[1], r[2], r[3]);
            }
            load_slots(j + 1);
          }
          if (is_k) publish(n + 2, describe(2 * (int)gridDim.x + __shfl_sync(0xffffffffu, ticket, 0)));
        }
        // out of work: the last CTA of the launch to get here re-arms the counter pair
        if (is_k && lane == 0 && atomicAdd(sched + 1, 1) == (int)gridDim.x - 1) {
          sched[0] = 0;
          sched[1] = 0;
        }
      } else {
        // ===================== MMA issuers (warp 10: tile 0, warp 11: tile 1) =====================
        // One elected thread per tile runs the whole loop (waits, tcgen05.mma, commits): a single thread
        // sustains only one UTCHMMA per ~45 cycles, so each tile gets its own issuing thread.  Every
        // value that feeds a descriptor is made provably warp-uniform (shuffle broadcast) so the compiler
        // keeps descriptors in uniform registers instead of wrapping each UTCHMMA in an elect loop.
        // (the format fields read 0 = f16 under kind::f16 and 0 = e4m3 under kind::f8f6f4)
        constexpr uint32_t idesc_qk = idesc_f16(kFp8 ? 0 : (kBf16 ? 1 : 0), kExtM, kExtC, 0, 0);
        constexpr uint32_t idesc_pv = idesc_f16(kFp8 ? 0 : (kBf16 ? 1 : 0), kExtM, D, 0, 1);
        const int i = __shfl_sync(0xffffffffu, warp - 10, 0);
        const uint32_t tmem_u = __shfl_sync(0xffffffffu, tmem, 0);
        const uint32_t smem_u = __shfl_sync(0xffffffffu, smem0, 0);
        const uint32_t barb = smem_u + S::kBar;
        // descriptor = constant high word | (address >> 4) in the low 14 bits
        constexpr uint64_t kDescKMajor =  // Q, K: K-major, 8-row swizzle atoms
            kSw64 ? smem_desc_sw64_hi(16, 512) : smem_desc_sw128_hi(16, 1024);
        constexpr uint64_t kDescVMn =     // V: MN-major
            kSw64 ? smem_desc_sw64_hi(S::kHalf, 512) : smem_desc_sw128_hi(S::kHalf, 1024);
        auto desc = [](uint64_t hi, uint32_t addr) { return hi | (uint64_t)((addr & 0x3FFFF) >> 4); };
        const uint32_t qa = kFp8 ? smem_u + kQ8Off + i * S::kTile : smem_u + S::kQ + i * S::kTile;
        const uint32_t ts = tmem_u + kColS + i * 128, to = tmem_u + kColO + i * 128;
        for (int n = 0;; ++n) {
          const ItemDesc d = acquire(n);
          if (d.E < 0) break;
          const int it = n & 31;
          (void)it;
          ItemTiles t;
          derive_tiles(d, p, t);
          const int nc_i = __shfl_sync(0xffffffffu, i == 0 ? t.nc[0] : t.nc[1], 0);
          const int nc = __shfl_sync(0xffffffffu, t.nc_max, 0);
          const int ntl = __shfl_sync(0xffffffffu, t.ntiles, 0);
          const int ktl = __shfl_sync(0xffffffffu, kt, 0);
          const int g = __shfl_sync(0xffffffffu, g_my, 0);
          const int nqi = __shfl_sync(0xffffffffu, nq_my, 0);
          if (elect_one()) {
            // S_i[(g + c) & 1] = Q_i K_c^T over the 64 keys of chunk c (both operands K-major from smem)
            auto issue_qk = [&](int c) {
              const uint32_t ka =
                  smem_u + S::kK + ((ktl + (c >> 1)) % kExtStages) * S::kTile + (c & 1) * (kExtC * kRowB);
              const int hb = (g + c) & 1;
#pragma unroll
#ifdef EXT_ABL_NOQK
              for (int ks = 0; ks < 1; ++ks) {
#else
              for (int ks = 0; ks < (kFp8 ? D / 32 : D / 16); ++ks) {
#endif
                const uint32_t off = (ks / 4) * S::kHalf + (ks % 4) * 32;  // 32 bytes per step either way
                if constexpr (kFp8)
                  umma_ss_f8(ts + hb * kExtC, desc(kDescKMajor, qa + off), desc(kDescKMajor, ka + off), idesc_qk,
                             ks > 0 ? 1u : 0u);
                else
                  umma_ss(ts + hb * kExtC, desc(kDescKMajor, qa + off), desc(kDescKMajor, ka + off), idesc_qk,
                          ks > 0 ? 1u : 0u);
              }
              umma_commit(barb + 8u * (kBarSFull + i * 2 + hb));
              if (c == nc_i - 1) umma_commit(barb + 8u * (kBarQEmpty + i));  // Q_i tile is free
            };
            // O_i (+)= P_i(c) V_c  (A = bf16 P in its S half of TMEM, B = V MN-major from smem)
            auto issue_pv = [&](int c) {
              const uint32_t va =
                  smem_u + S::kV + ((ktl + (c >> 1)) % kExtStages) * S::kTile + (c & 1) * (kExtC * kRowB);
              const int hb = (g + c) & 1;
#pragma unroll
#ifdef EXT_ABL_NOPV
              for (int ks = 0; ks < 1; ++ks) {
#else
              for (int ks = 0; ks < (kFp8 ? kExtC / 32 : kExtC / 16); ++ks) {
#endif
                // 16 keys per step = 8 packed TMEM columns of A; 2 KB down the V tile; d-halves 16 KB apart
                // (e4m3: 32 keys per step = 8 columns of four; 4 KB down the tile; one 128-byte group of d)
                if constexpr (kFp8)
                  umma_ts_f8(to, ts + hb * kExtC + ks * 8, desc(kDescVMn, va + ks * 32 * kRowB), idesc_pv,
                             (c > 0 || ks > 0) ? 1u : 0u);
                else
                  umma_ts(to, ts + hb * kExtC + ks * 8, desc(kDescVMn, va + ks * 2048), idesc_pv,
                          (c > 0 || ks > 0) ? 1u : 0u);
              }
              umma_commit(barb + 8u * (kBarPVDone + i));
              if (c == nc_i - 1) umma_commit(barb + 8u * (kBarOFinal + i));
            };
            if (i == 0) { TRACE(5, it, 0); }
            if (nc_i > 0) mbar_wait(barb + 8u * ((kFp8 ? kBarQ8 : kBarQ) + i), nqi & 1);
            if (i == 0) { TRACE(5, it, 1); }
            mbar_wait(barb + 8u * (kBarKFull + ktl % kExtStages), (ktl / kExtStages) & 1);
            tc_fence_after();
            if (i == 0) { TRACE(5, it, 2); }
            if (nc_i > 0) issue_qk(0);
            if (nc_i > 1) issue_qk(1);
            umma_commit(barb + 8u * (kBarKEmpty + ktl % kExtStages));
            // loops over the ITEM's chunk count so that both issuers release every K/V stage
            for (int c = 0; c < nc; ++c) {
              const int tt = ktl + (c >> 1), st = tt % kExtStages, st1 = (tt + 1) % kExtStages;
              if ((c & 1) == 0) {
                mbar_wait(barb + 8u * (kBarVFull + st), (tt / kExtStages) & 1);
                if (c + 2 < nc) mbar_wait(barb + 8u * (kBarKFull + st1), ((tt + 1) / kExtStages) & 1);
              }
              if (c < nc_i) {
                TRACE(2, c, 3 * i + 0);
                mbar_wait(barb + 8u * (kBarPFull + i * 2 + ((g + c) & 1)), ((g + c) >> 1) & 1);
                tc_fence_after();
                TRACE(2, c, 3 * i + 1);
                issue_pv(c);
#ifdef EXT_SERIAL_PVQK
                // checker build (libb200attn_serial.so, tests/test_extend_determinism_gpu.py): PV_i(c) has COMPLETED
                // before QK_i(c+2) may overwrite the TMEM columns it reads P from; the product build relies on the
                // tensor pipe executing one thread's tcgen05.mma in issue order instead
                if (c + 2 < nc_i) mbar_wait(barb + 8u * (kBarPVDone + i), (g + c) & 1);
#endif
                if (c + 2 < nc_i) issue_qk(c + 2);
                TRACE(2, c, 3 * i + 2);
              }
              if ((c & 1) == 1 || c == nc - 1) umma_commit(barb + 8u * (kBarVEmpty + st));
              if ((c & 1) == 1 && c + 1 < nc) umma_commit(barb + 8u * (kBarKEmpty + st1));
            }
            if (i == 0) { TRACE(5, it, 3); }
          }
          __syncwarp();
          kt += ntl;
          g_my += nc_i;
          nq_my += nc_i > 0 ? 1 : 0;
        }
      }
    }
  } else {
    asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;\n" ::"n"(kRegsSoftmax));
    {
      // ===================== softmax + epilogue (thread = query row of tile i) =====================
      using T = typename std::conditional<kBf16, __nv_bfloat16, __half>::type;
      const int i = warp >> 2;            // tile of this warpgroup
      const int row = threadIdx.x & 127;  // == TMEM lane
      const uint32_t lane_addr = tmem + ((uint32_t)((warp & 3) * 32) << 16);
      const uint32_t col_s = lane_addr + kColS + i * 128, col_o = lane_addr + kColO + i * 128;
      const uint64_t scale2 = f2_pack(p.scale_log2, p.scale_log2);
      for (int n = 0;; ++n) {
        const ItemDesc d = acquire(n);
        if (d.E < 0) break;
        const int it = n & 31;
        (void)it;
        ItemTiles t;
        derive_tiles(d, p, t);
        const int nc_i = i == 0 ? t.nc[0] : t.nc[1];
        if (nc_i == 0) continue;
        if (i == 0) { TRACE(4, it, 0); }
        if constexpr (kFp8) {
          // Q_i arrived as 16-bit (TMA, two 64-column halves of 128-byte swizzled rows): this thread rewrites its
          // row as e4m3 into the K-major e4m3 tile the QK instruction reads (one 128-byte swizzled row).
          mbar_wait(bar(kBarQ + i), nq_my & 1);
          const uint8_t* qrow = smem + S::kQ + i * S::kTile + row * 128;
          uint8_t* q8row = smem + kQ8Off + i * S::kTile + row * kRowB;
          const int swz8 = kSw64 ? (row >> 1) & 3 : row & 7;  // 16-byte chunk XOR of the e4m3 row's swizzle mode
#pragma unroll
          for (int j = 0; j < D / 16; ++j) {  // 16 columns -> one 16-byte chunk of the e4m3 row
            uint32_t w[4];
#pragma unroll
            for (int hc = 0; hc < 2; ++hc) {  // two 8-column 16-byte chunks of the 16-bit row
              const int c16 = 2 * (j % 4) + hc;
              const uint4 v = *reinterpret_cast<const uint4*>(qrow + (j / 4) * S::kHalf + ((c16 ^ (row & 7)) << 4));
              const uint32_t in[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
              for (int k2 = 0; k2 < 2; ++k2) {
                const float2 f0 = unpack2<kBf16>(in[2 * k2]), f1 = unpack2<kBf16>(in[2 * k2 + 1]);
                w[2 * hc + k2] = (uint32_t)__nv_cvt_float2_to_fp8x2(f0, __NV_SATFINITE, __NV_E4M3) |
                                 ((uint32_t)__nv_cvt_float2_to_fp8x2(f1, __NV_SATFINITE, __NV_E4M3) << 16);
              }
            }
            *reinterpret_cast<uint4*>(q8row + ((j ^ swz8) << 4)) = make_uint4(w[0], w[1], w[2], w[3]);
          }
          fence_proxy_async_smem();  // generic-proxy stores -> visible to the tensor core's async-proxy reads
          mbar_arrive(bar(kBarQ8 + i));
        }
        const int m0 = i == 0 ? t.m0[0] : t.m0[1];
        const int kvend_i = i == 0 ? t.kvend[0] : t.kvend[1];
        // keys visible to this row: [0, limit)
        int limit = p.causal ? min(d.L, t.prefix + m0 + row + 1) : d.L;
        limit = max(1, min(limit, kvend_i));
        // sliding window: keys below `lower` are masked (whole tiles below kbase are never loaded)
        const int lower = p.window_left >= 0 ? min(limit - 1, max(0, t.prefix + m0 + row - p.window_left)) : 0;
        float m_used = -INFINITY;  // reference max (log2 domain) the running sums are expressed in
        float l_run = 0.f;
        // One 64-key chunk.  Two copies of this body exist: kEdge = false is the straight-line interior chunk (no
        // mask, cap or window code at all: ~400 instructions, fits the instruction cache of a sub-partition);
        // kEdge = true carries every masking block.  With one copy, an interior chunk branched over four fully
        // unrolled 64-element mask blocks and paid an instruction-fetch stall after each jump (ncu: ~15 % of the
        // softmax warps' samples were no_inst / branch_resolving on those BSSY / BSYNC / BRA).
        auto chunk = [&](auto edge_tag, const int c) {
          constexpr bool kEdge = decltype(edge_tag)::value;
          const int gc = g_my + c;                          // chunk count of this tile since kernel start
          const uint32_t col_c = col_s + (gc & 1) * kExtC;  // S/P half of this chunk
          TRACE(i, c, 0);
          mbar_wait(bar(kBarSFull + i * 2 + (gc & 1)), (gc >> 1) & 1);
          tc_fence_after();
          TRACE(i, c, 1);
          if (i == 0 && c == 0) { TRACE(4, it, 1); }
          float s[kExtC];
#pragma unroll
          for (int q = 0; q < kExtC / 32; ++q) {
            uint32_t v[32];
            tmem_ld32(col_c + q * 32, v);
#pragma unroll
            for (int e = 0; e < 32; ++e) s[q * 32 + e] = __uint_as_float(v[e]);
          }
          tmem_wait_ld();
          TRACE(i, c, 2);
          const int n0 = t.kbase + c * kExtC;
          if constexpr (kEdge) {
          if (p.cap_pre > 0.f) {  // logit cap (uniform branch): s <- c/sm_scale * tanh(s * sm_scale / c)
            const float inv_pre = 1.f / p.cap_pre;
#pragma unroll
            for (int e = 0; e < kExtC; ++e) s[e] = tanh_approx_f(s[e] * p.cap_pre) * inv_pre;
          }
          if (n0 + kExtC > limit) {
#pragma unroll
            for (int e = 0; e < kExtC; ++e)
              if (n0 + e >= limit) s[e] = -INFINITY;
          }
          if (n0 < lower) {
#pragma unroll
            for (int e = 0; e < kExtC; ++e)
              if (n0 + e < lower) s[e] = -INFINITY;
          }
          if (p.custom_mask != nullptr && row < (i == 0 ? t.rows[0] : t.rows[1])) {  // speculative-decoding tree masks
            const uint8_t* mrow = p.custom_mask + (int64_t)d.pad0 + (int64_t)(m0 + row) * d.L;
            const int first = p.mask_skip_prefix ? t.prefix : 0;
#pragma unroll  // full unroll: s[] must stay in registers
            for (int e = 0; e < kExtC; ++e) {
              const int key = n0 + e;
              if (key >= first && key < limit && __ldg(mrow + key) == 0) s[e] = -INFINITY;
            }
          }
          }  // kEdge
          float mx4[4] = {s[0], s[1], s[2], s[3]};
#pragma unroll
          for (int e = 4; e < kExtC; e += 8) {
            mx4[0] = fmax3(mx4[0], s[e], s[e + 1]);
            mx4[1] = fmax3(mx4[1], s[e + 2], s[e + 3]);
            if (e + 4 < kExtC) {
              mx4[2] = fmax3(mx4[2], s[e + 4], s[e + 5]);
              mx4[3] = fmax3(mx4[3], s[e + 6], s[e + 7]);
            }
          }
          const float mx = fmaxf(fmaxf(mx4[0], mx4[1]), fmaxf(mx4[2], mx4[3])) * p.scale_log2;
          float alpha = 1.f;
          bool rescale = false;
          if (m_used == -INFINITY) {
            m_used = mx;  // first chunk with a visible key (a sliding window can mask whole leading chunks)
          } else if (mx > m_used + kRescaleThreshold) {
            alpha = ex2_approx(m_used - mx);
            m_used = mx;
            rescale = true;
          }
          if (__any_sync(0xffffffffu, rescale)) {
            // rare: O_i must be quiescent, i.e. PV_i(c-1) (possibly still in flight) has completed
            mbar_wait(bar(kBarPVDone + i), (gc - 1) & 1);
            tc_fence_after();
            l_run *= alpha;
            rescale_o<D>(col_o, alpha);  // out of line: keeps the hot loop compact
          }
          TRACE(i, c, 3);
          // p = exp2(s * scale - m); packed bf16 pairs go back into the first 32 columns of this S half
          const float m_eff = (m_used == -INFINITY) ? 0.f : m_used;  // all keys so far masked: p = exp2(-inf) = 0
          const uint64_t negm2 = f2_pack(-m_eff, -m_eff);
          uint64_t sum2[2] = {0ull, 0ull};
          uint32_t pk[kFp8 ? 16 : 32];
#pragma unroll
          for (int e = 0; e < kExtC; e += 2) {
            float x0, x1, p0, p1;
            f2_unpack(f2_fma(f2_pack(s[e], s[e + 1]), scale2, negm2), x0, x1);
#ifdef EXT_ABL_NOEXP
            if (true) {
              p0 = x0;
              p1 = x1;
            } else
#endif
            if (((e >> 1) % kEmuPeriod) < kEmuCount) {
              exp2_emu2(x0, x1, p0, p1);  // FMA-pipe polynomial: takes load off the 16-lane MUFU
            } else {
              p0 = ex2_approx(x0);
              p1 = ex2_approx(x1);
            }
            sum2[(e >> 1) & 1] = f2_add(sum2[(e >> 1) & 1], f2_pack(p0, p1));
            if constexpr (kFp8) {  // four e4m3 per TMEM column (p <= 2^8 under the lazy rescale: inside e4m3's range)
              const uint32_t h = __nv_cvt_float2_to_fp8x2(make_float2(p0, p1), __NV_SATFINITE, __NV_E4M3);
              if ((e & 2) == 0) pk[e >> 2] = h;
              else pk[e >> 2] |= h << 16;
            } else {
              pk[e >> 1] = pack2<kBf16>(p0, p1);
            }
          }
          if constexpr (kFp8) tmem_st16(col_c, pk);
          else tmem_st32(col_c, pk);
          float a0, a1;
          f2_unpack(f2_add(sum2[0], sum2[1]), a0, a1);
          l_run += a0 + a1;
          TRACE(i, c, 4);
          tmem_wait_st();
          tc_fence_before();
          mbar_arrive(bar(kBarPFull + i * 2 + (gc & 1)));
          TRACE(i, c, 5);
        };
        const bool any_feature = p.cap_pre > 0.f || p.custom_mask != nullptr;
        // warp-uniform: no row of this warp needs a mask in chunk c
        auto interior = [&](int c) {
          const int n0 = t.kbase + c * kExtC;
          return __all_sync(0xffffffffu, n0 + kExtC <= limit && n0 >= lower) && !any_feature;
        };
        for (int c = 0; c < nc_i;) {
          if (interior(c)) {
            do {  // runs of interior chunks stay inside this short loop (one near back edge)
              chunk(std::false_type{}, c);
              ++c;
            } while (c < nc_i && interior(c));
          } else {
            chunk(std::true_type{}, c);
            ++c;
          }
        }
        // ---- epilogue of this item (the issuer already runs QK of the next one) ----
        if (i == 0) { TRACE(4, it, 2); }
        mbar_wait(bar(kBarOFinal + i), nq_my & 1);
        tc_fence_after();
        if (i == 0) { TRACE(4, it, 3); }
        const int h = i == 0 ? t.h[0] : t.h[1];
        if (p.sinks != nullptr) l_run += ex2_approx(__ldg(p.sinks + h) * kLog2e - m_used);
        const float inv = p.out_scale / l_run;
        const int rows_i = i == 0 ? t.rows[0] : t.rows[1];
        const bool valid = row < rows_i;
        if (rows_i == kExtM && p.o_tma) {
          // Full tile: O leaves through shared memory and the copy engine.  A thread owns a ROW, so direct stores
          // are 32 scattered 16-byte pieces per warp instruction (measured: 4.7 k cycles per item, LSU-bound);
          // here each WARP writes its 32 rows x 32 columns to a 64B-swizzled staging block (conflict free) and its
          // lane 0 issues a 2-D TMA store of that block.  Two blocks per warp: slab j waits only for the store of
          // slab j-2, and nothing is synchronised wider than the warp (it was two 128-thread named barriers per slab).
          const int wq = warp & 3;  // warp within the warpgroup = 32-row quarter of the tile
          const uint32_t obuf = smem0 + S::kO + i * 2 * S::kOSlab + wq * (32 * 64);
#pragma unroll
          for (int c = 0; c < D / kExtOSlabCols; ++c) {
            uint32_t v[32];
            tmem_ld32(col_o + c * 32, v);
            if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 1;\n" ::: "memory");
            tmem_wait_ld();
            __syncwarp();  // block (c & 1) of this warp is free again
            const uint32_t dst = obuf + (c & 1) * S::kOSlab + lane * 64;
#pragma unroll
            for (int e = 0; e < 32; e += 8) {
              uint4 w;
              w.x = pack2<kBf16>(__uint_as_float(v[e]) * inv, __uint_as_float(v[e + 1]) * inv);
              w.y = pack2<kBf16>(__uint_as_float(v[e + 2]) * inv, __uint_as_float(v[e + 3]) * inv);
              w.z = pack2<kBf16>(__uint_as_float(v[e + 4]) * inv, __uint_as_float(v[e + 5]) * inv);
              w.w = pack2<kBf16>(__uint_as_float(v[e + 6]) * inv, __uint_as_float(v[e + 7]) * inv);
              const uint32_t chunk = (uint32_t)(e >> 3) ^ ((uint32_t)(lane >> 1) & 3u);  // CU_TENSOR_MAP_SWIZZLE_64B
              asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};\n" ::"r"(dst + chunk * 16), "r"(w.x),
                           "r"(w.y), "r"(w.z), "r"(w.w)
                           : "memory");
            }
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) {
              asm volatile(
                  "cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];\n" ::"l"(&tm_o),
                  "r"(obuf + (c & 1) * S::kOSlab), "r"(h * D + c * kExtOSlabCols), "r"(d.q_beg + m0 + wq * 32)
                  : "memory");
              asm volatile("cp.async.bulk.commit_group;\n" ::: "memory");
            }
          }
        } else {
          T* dst = reinterpret_cast<T*>(p.o) + (int64_t)(d.q_beg + m0 + row) * p.o_stride + (int64_t)h * D;
#pragma unroll
          for (int c = 0; c < D / 32; ++c) {
            uint32_t v[32];
            tmem_ld32(col_o + c * 32, v);
            tmem_wait_ld();
            if (valid) {
#pragma unroll
              for (int e = 0; e < 32; e += 8) {
                uint4 w;
                w.x = pack2<kBf16>(__uint_as_float(v[e]) * inv, __uint_as_float(v[e + 1]) * inv);
                w.y = pack2<kBf16>(__uint_as_float(v[e + 2]) * inv, __uint_as_float(v[e + 3]) * inv);
                w.z = pack2<kBf16>(__uint_as_float(v[e + 4]) * inv, __uint_as_float(v[e + 5]) * inv);
                w.w = pack2<kBf16>(__uint_as_float(v[e + 6]) * inv, __uint_as_float(v[e + 7]) * inv);
                *reinterpret_cast<uint4*>(dst + c * 32 + e) = w;
              }
            }
          }
        }
        if (valid && p.lse != nullptr)
          p.lse[(int64_t)(d.q_beg + m0 + row) * p.Hq + h] = (m_used + log2f(l_run)) * kLn2;
        tc_fence_before();
        if (i == 0) { TRACE(4, it, 4); }
        g_my += nc_i;
        ++nq_my;
      }
    }
    // the staged output blocks must have left shared memory (and be globally visible) before the CTA retires
    if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;\n" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
#ifdef EXT_TRACE
  if (p.trace != nullptr && blockIdx.x == 0 && threadIdx.x == 0) {
    long long gt;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gt));
    p.trace[(3 * 32 + 0) * 8 + 2] = clk();
    p.trace[(3 * 32 + 0) * 8 + 3] = gt;
  }
  if (p.trace != nullptr && threadIdx.x == 0) {
    long long gt;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gt));
    p.trace[6 * 32 * 8 + blockIdx.x * 2 + 1] = gt;
  }
#endif
  if (warp == 10) {
    tc_fence_after();
    tmem_dealloc<kTmemCols>(tmem);
  }
}

}  // namespace b200attn

// ------------------------------------------------------------------ host side
namespace b200attn {

static std::atomic<unsigned> g_sched_next{0};

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (fn == nullptr) {
    void* sym = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &qres) ==
            cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(sym);
  }
  return fn;
}

int make_tensor_map_2d(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols,
                       uint64_t row_stride_bytes, uint32_t box_cols, uint32_t box_rows,
                       CUtensorMapSwizzle swizzle) {
  EncodeTiledFn fn = get_encode_fn();
  if (fn == nullptr) {
    set_error("cuTensorMapEncodeTiled is not available from the driver");
    return B200ATTN_ERR_CUDA;
  }
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {row_stride_bytes};
  cuuint32_t box[2] = {box_cols, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_UINT16, 2, const_cast<void*>(base), dims, strides,
                  box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed with CUresult %d (rows=%llu cols=%llu stride=%llu)",
              (int)r, (unsigned long long)rows, (unsigned long long)cols,
              (unsigned long long)row_stride_bytes);
    return B200ATTN_ERR_CUDA;
  }
  return B200ATTN_OK;
}

// Tensor maps of long-lived buffers (the KV pool) depend only on (base, shape, stride, box): they are encoded once
// and kept - the encode is a driver call of a few microseconds, paid per layer-call otherwise.  Small FIFO cache,
// mutex-protected; entries are returned by pointer and stay valid for the life of the process.
int tensor_map_cached(const CUtensorMap** out, const void* base, uint64_t rows, uint64_t cols,
                      uint64_t row_stride_bytes, uint32_t box_cols, uint32_t box_rows, CUtensorMapSwizzle swizzle) {
  struct Entry {
    const void* base;
    uint64_t rows, cols, stride;
    uint32_t bc, br;
    int sw;
    CUtensorMap map;
  };
  constexpr int kCap = 256;
  static Entry* entries = new Entry[kCap];
  static int count = 0, next = 0;
  static std::mutex mu;
  std::lock_guard<std::mutex> lock(mu);
  for (int i = 0; i < count; ++i) {
    const Entry& e = entries[i];
    if (e.base == base && e.rows == rows && e.cols == cols && e.stride == row_stride_bytes && e.bc == box_cols &&
        e.br == box_rows && e.sw == (int)swizzle) {
      *out = &e.map;
      return B200ATTN_OK;
    }
  }
  Entry fresh;
  int rc = make_tensor_map_2d(&fresh.map, base, rows, cols, row_stride_bytes, box_cols, box_rows, swizzle);
  if (rc) return rc;
  fresh.base = base; fresh.rows = rows; fresh.cols = cols; fresh.stride = row_stride_bytes;
  fresh.bc = box_cols; fresh.br = box_rows; fresh.sw = (int)swizzle;
  // entries handed out earlier must stay valid: once the table is full the oldest slot is overwritten IN PLACE only
  // when no launch can still be reading it - tensor maps are passed to kernels BY VALUE (__grid_constant__), so a
  // returned pointer is dereferenced on the host before the launcher returns, under this same call
  const int slot = count < kCap ? count++ : (next++ % kCap);
  entries[slot] = fresh;
  *out = &entries[slot].map;
  return B200ATTN_OK;
}

template <int D, bool kBf16, bool kFp8 = false>
static int launch_extend(const CUtensorMap& tq, const CUtensorMap& tk, const CUtensorMap& tv,
                         const CUtensorMap& tkb, const CUtensorMap& tvb, const CUtensorMap& to,
                         const ExtendParams& p, int max_qtiles, cudaStream_t st) {
  constexpr int kSmem = ExtSmem<D>::kTotal + 1024;  // + alignment slack
  static bool configured = false;
  if (!configured) {
    B2_CUDA(cudaFuncSetAttribute(extend_kernel<D, kBf16, kFp8>,
                                 cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem));
    configured = true;
  }
  // Persistent: one CTA per SM, items (request, kv head, tile pair) drawn in list order; consecutive items
  // share a kv head's K/V tiles, run at the same time on neighbouring SMs and hit them in L2.  The host only
  // knows an upper bound of the item count (sizes the grid, graph-safe).  Each launch gets its own ticket counter
  // pair: launches that overlap in time (other streams) are independent as long as fewer than kExtSchedSlots
  // launches separate them.
  const long long upper = (long long)((max_qtiles * p.G + 1) / 2) * p.Hkv;
  const int sms = sm_count();
  constexpr int kExtMaxB = ExtSmem<D>::kMaxB;
  for (int b0 = 0; b0 < p.B; b0 += kExtMaxB) {  // the in-kernel prefix table holds kMaxB requests
    ExtendParams ps = p;
    ps.B = p.B - b0 < kExtMaxB ? p.B - b0 : kExtMaxB;
    ps.qo_indptr = p.qo_indptr + b0;
    ps.kv_indptr = p.kv_indptr + b0;
    const long long items = upper * ps.B;
    dim3 grid((unsigned)(items < sms ? items : sms));
    ps.sched_slot = (int)(g_sched_next.fetch_add(1u) % (unsigned)kExtSchedSlots);
    extend_kernel<D, kBf16, kFp8><<<grid, kExtThreads, kSmem, st>>>(tq, tk, tv, tkb, tvb, to, ps);
    B2_LAUNCH_CHECK();
  }
  return B200ATTN_OK;
}

}  // namespace b200attn

// the launch's two ticket counters (zeroed once by the caller, left at zero by every launch)
extern "C" int64_t b200attn_extend_workspace_bytes(int32_t, int64_t, int32_t) { return 64; }

// kv_fp8: k_cache / v_cache hold e4m3 bytes, kv_stride counts BYTES per slot row (16-bit pools: elements)
static int extend_impl(const void* q, int64_t q_stride, const void* k_cache,
                       const void* v_cache, int64_t kv_stride, int64_t num_slots, void* o,
                       int64_t o_stride, float* lse, const int32_t* qo_indptr,
                       const int32_t* kv_indptr, const int32_t* kv_indices, int32_t B,
                       int64_t total_q_host, int32_t max_qo_len_host, int32_t Hq,
                       int32_t Hkv, int32_t D, int32_t page_size, int dtype, int causal,
                       float sm_scale, int32_t window_left, float logit_cap,
                       const float* sinks, const uint8_t* custom_mask,
                       const int32_t* mask_indptr, int mask_skip_prefix, void* workspace,
                       int64_t workspace_bytes, void* stream, bool kv_fp8, float k_descale, float v_descale) {
  using namespace b200attn;
  B2_CHECK_ARG(logit_cap >= 0.f, "extend: negative logit_cap");
  (void)page_size;  // rows are gathered token-granular: any page size, no contiguity assumption
  B2_CHECK_ARG(B >= 0 && total_q_host >= 0 && max_qo_len_host >= 0, "extend: negative size");
  if (B == 0 || total_q_host == 0 || max_qo_len_host == 0) return B200ATTN_OK;
  B2_CHECK_ARG(q && k_cache && v_cache && o && qo_indptr && kv_indptr && kv_indices,
               "extend: null pointer");
  B2_CHECK_ARG(Hq > 0 && Hkv > 0 && Hq % Hkv == 0, "extend: Hq=%d must be a multiple of Hkv=%d",
               Hq, Hkv);
  B2_CHECK_ARG(Hkv <= 65535, "extend: kv head count too large");
  B2_CHECK_ARG(num_slots > 0 && num_slots < (1ll << 31), "extend: bad num_slots");
  B2_CHECK_ARG(q_stride % 8 == 0 && kv_stride % (kv_fp8 ? 16 : 8) == 0 && o_stride % 8 == 0,
               "extend: strides must keep rows 16-byte aligned");
  B2_CHECK_ARG(!kv_fp8 || (k_descale > 0.f && v_descale > 0.f), "extend: fp8 descales must be positive");
  B2_CHECK_ARG((((uintptr_t)q | (uintptr_t)k_cache | (uintptr_t)v_cache | (uintptr_t)o) & 15) == 0,
               "extend: tensors must be 16-byte aligned");
  if (dtype != B200ATTN_BF16 && dtype != B200ATTN_FP16) B2_UNSUPPORTED("extend: dtype %d", dtype);
  if (D != 64 && D != 128) B2_UNSUPPORTED("extend: head_dim %d (supported: 64, 128)", D);
  const bool sw64 = kv_fp8 && D == 64;  // 64-byte e4m3 rows
  const uint32_t kv_box_cols = sw64 ? 32 : 64;
  const CUtensorMapSwizzle kv_swz = sw64 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_128B;
  // the pool maps count 16-bit columns: an e4m3 row of Hkv * D bytes is Hkv * D / 2 of them
  const uint64_t kv_cols = kv_fp8 ? (uint64_t)Hkv * D / 2 : (uint64_t)Hkv * D;
  const uint64_t kv_row_bytes = kv_fp8 ? (uint64_t)kv_stride : (uint64_t)kv_stride * 2;
  // q / o are activations (a new address every call): encoded per call; the four pool maps are cached
  CUtensorMap tq, to;
  const CUtensorMap *ptk, *ptv, *ptkb, *ptvb;
  int rc = make_tensor_map_2d(&tq, q, (uint64_t)total_q_host, (uint64_t)Hq * D,
                              (uint64_t)q_stride * 2, 64, kExtM);
  if (rc) return rc;
  rc = tensor_map_cached(&ptk, k_cache, (uint64_t)num_slots, kv_cols, kv_row_bytes, kv_box_cols, 1, kv_swz);
  if (rc) return rc;
  rc = tensor_map_cached(&ptv, v_cache, (uint64_t)num_slots, kv_cols, kv_row_bytes, kv_box_cols, 1, kv_swz);
  if (rc) return rc;
  rc = tensor_map_cached(&ptkb, k_cache, (uint64_t)num_slots, kv_cols, kv_row_bytes, kv_box_cols,
                         kExtBoxRows, kv_swz);
  if (rc) return rc;
  rc = tensor_map_cached(&ptvb, v_cache, (uint64_t)num_slots, kv_cols, kv_row_bytes, kv_box_cols,
                         kExtBoxRows, kv_swz);
  if (rc) return rc;
  const CUtensorMap &tk = *ptk, &tv = *ptv, &tkb = *ptkb, &tvb = *ptvb;
  rc = make_tensor_map_2d(&to, o, (uint64_t)total_q_host, (uint64_t)Hq * D, (uint64_t)o_stride * 2,
                          kExtOSlabCols, 32, CU_TENSOR_MAP_SWIZZLE_64B);  // one warp's 32 rows x 32 columns
  if (rc) return rc;
  ExtendParams p;
  p.o_tma = 1;
  if (const char* e = getenv("B200ATTN_EXT_OTMA")) p.o_tma = atoi(e) != 0;  // dev ablation hook
  p.o = o;
  p.lse = lse;
  p.qo_indptr = qo_indptr;
  p.kv_indptr = kv_indptr;
  p.kv_indices = kv_indices;
  p.o_stride = o_stride;
  p.B = B;
  p.Hq = Hq;
  p.Hkv = Hkv;
  p.G = Hq / Hkv;
  p.num_slots = (int)num_slots;
  p.causal = (causal & B200ATTN_EXT_CAUSAL) ? 1 : 0;
  p.skip_single = (causal & B200ATTN_EXT_SKIP_SINGLE) ? 1 : 0;
  if (kv_fp8) sm_scale *= k_descale;  // scores come out of the tensor core in units of k_descale
  p.scale_log2 = sm_scale * kLog2e;
  p.out_scale = kv_fp8 ? v_descale : 1.f;
  p.window_left = window_left;
  p.cap_pre = logit_cap > 0.f ? sm_scale / logit_cap : 0.f;
  p.sinks = sinks;
  B2_CHECK_ARG(custom_mask == nullptr || mask_indptr != nullptr, "extend: custom_mask needs mask_indptr");
  p.custom_mask = custom_mask;
  p.mask_indptr = mask_indptr;
  p.mask_skip_prefix = mask_skip_prefix;
#ifdef EXT_TRACE
  p.trace = workspace_bytes >= (6 * 32 * 8 + 2 * 160) * 8 ? (long long*)workspace : nullptr;
  p.sched_ptr = nullptr;  // the dev build's workspace is the trace buffer
#else
  // Ticket counters: the caller's workspace when there is one (launches that share a workspace are serialised by
  // the workspace contract, so nothing is shared between independent callers / streams / graphs); otherwise a
  // process-wide pool of kExtSchedSlots pairs handed out round-robin.
  B2_CHECK_ARG(workspace == nullptr || workspace_bytes < 8 || ((uintptr_t)workspace & 3) == 0,
               "extend: workspace must be 4-byte aligned");
  p.sched_ptr = (workspace != nullptr && workspace_bytes >= 8) ? reinterpret_cast<int*>(workspace) : nullptr;
#endif
  const int max_qtiles = (max_qo_len_host + kExtM - 1) / kExtM;
  cudaStream_t st = (cudaStream_t)stream;
  if (kv_fp8 && D == 128)
    return dtype == B200ATTN_BF16 ? launch_extend<128, true, true>(tq, tk, tv, tkb, tvb, to, p, max_qtiles, st)
                                  : launch_extend<128, false, true>(tq, tk, tv, tkb, tvb, to, p, max_qtiles, st);
  if (kv_fp8)
    return dtype == B200ATTN_BF16 ? launch_extend<64, true, true>(tq, tk, tv, tkb, tvb, to, p, max_qtiles, st)
                                  : launch_extend<64, false, true>(tq, tk, tv, tkb, tvb, to, p, max_qtiles, st);
  if (D == 128)
    return dtype == B200ATTN_BF16 ? launch_extend<128, true>(tq, tk, tv, tkb, tvb, to, p, max_qtiles, st)
                                  : launch_extend<128, false>(tq, tk, tv, tkb, tvb, to, p, max_qtiles, st);
  return dtype == B200ATTN_BF16 ? launch_extend<64, true>(tq, tk, tv, tkb, tvb, to, p, max_qtiles, st)
                                : launch_extend<64, false>(tq, tk, tv, tkb, tvb, to, p, max_qtiles, st);
}

extern "C" int b200attn_extend(const void* q, int64_t q_stride, const void* k_cache,
                               const void* v_cache, int64_t kv_stride, int64_t num_slots, void* o,
                               int64_t o_stride, float* lse, const int32_t* qo_indptr,
                               const int32_t* kv_indptr, const int32_t* kv_indices, int32_t B,
                               int64_t total_q_host, int32_t max_qo_len_host, int32_t Hq,
                               int32_t Hkv, int32_t D, int32_t page_size, int dtype, int causal,
                               float sm_scale, int32_t window_left, float logit_cap,
                               const float* sinks, const uint8_t* custom_mask,
                               const int32_t* mask_indptr, int mask_skip_prefix, void* workspace,
                               int64_t workspace_bytes, void* stream) {
  return extend_impl(q, q_stride, k_cache, v_cache, kv_stride, num_slots, o, o_stride, lse, qo_indptr, kv_indptr,
                     kv_indices, B, total_q_host, max_qo_len_host, Hq, Hkv, D, page_size, dtype, causal, sm_scale,
                     window_left, logit_cap, sinks, custom_mask, mask_indptr, mask_skip_prefix, workspace,
                     workspace_bytes, stream, false, 1.f, 1.f);
}

extern "C" int b200attn_extend_fp8kv(const void* q, int64_t q_stride, const void* k_cache,
                                     const void* v_cache, int64_t kv_stride_bytes, int64_t num_slots, void* o,
                                     int64_t o_stride, float* lse, const int32_t* qo_indptr,
                                     const int32_t* kv_indptr, const int32_t* kv_indices, int32_t B,
                                     int64_t total_q_host, int32_t max_qo_len_host, int32_t Hq,
                                     int32_t Hkv, int32_t D, int32_t page_size, int dtype, int causal,
                                     float sm_scale, float k_descale, float v_descale, int32_t window_left,
                                     float logit_cap, const float* sinks, const uint8_t* custom_mask,
                                     const int32_t* mask_indptr, int mask_skip_prefix, void* workspace,
                                     int64_t workspace_bytes, void* stream) {
  return extend_impl(q, q_stride, k_cache, v_cache, kv_stride_bytes, num_slots, o, o_stride, lse, qo_indptr,
                     kv_indptr, kv_indices, B, total_q_host, max_qo_len_host, Hq, Hkv, D, page_size, dtype, causal,
                     sm_scale, window_left, logit_cap, sinks, custom_mask, mask_indptr, mask_skip_prefix, workspace,
                     workspace_bytes, stream, true, k_descale, v_descale);
}
