// (a4) Paged GQA split-KV flash-decode for B200 (sm_100a).
//
// Semantics follow the oracle python/sglang/srt/layers/attention/torch_native_backend.py:
// 176-277 (per request: gather K/V rows through req_to_token, softmax(q K^T * scale) V,
// non-causal, GQA head h -> kv head h / G).  The design is HBM-first:
//
//   * one CTA owns (request, kv head, kv split): all G query heads of the kv head
//     share every K/V byte that is fetched -> each KV byte crosses HBM exactly once;
//   * K/V rows are gathered token-granular (page_size agnostic) with 16-byte
//     cp.async.cg into a 3-stage shared-memory ring, 32 KB per stage and two
//     resident CTAs per SM => 128 KB of loads in flight per SM;
//   * the G x D query block is padded to the M=16 of mma.sync.m16n8k16; each of the
//     four warps takes 16 keys of a 64-key tile, S = Q K^T and O += P V run on the
//     tensor pipe straight out of XOR-swizzled shared memory (ldmatrix / .trans),
//     the online softmax runs in fp32 with quad shuffles;
//   * warps are merged through shared memory, splits through a fp32 workspace and
//     a tiny merge kernel.  The split policy is a device-side pure function of
//     the kv length, so a captured CUDA graph replays with new lengths untouched.
//
// Algorithmic bytes per (request, kv head): 2 * L * D * sizeof(T) (K and V rows)
// + 4 * L (int32 slot ids).
#include <mutex>
#include "decode.cuh"

namespace b200attn {

// NW = warps per CTA (16 keys per warp and tile).  NW = 4: 64-key tiles, 3 x 32 KB ring, 2 CTAs per SM (the original
// shape).  NW = 2: 64-thread CTAs, 32-key tiles, 3 x 16 KB ring, FOUR CTAs per SM: the same 128 KB of loads in
// flight per SM, but 592 resident slots instead of 296 - cfg2's 512 units become ONE wave (no 73 %-full second
// wave, every CTA streams at the same rate from start to end) and small batches can be split twice as finely
// before a second wave forms.
#ifdef DEC_TRACE
// dev-only (libb200attn_trace.so, tools/trace_decode.py): per-CTA wall-clock stamps of decode_kernel
__device__ long long g_dec_trace[8 * 16384];
#define DTRACE(slot)                                                                                       \
  if (threadIdx.x == 0) {                                                                                  \
    const int cta_ = blockIdx.x + gridDim.x * (blockIdx.y + gridDim.y * blockIdx.z);                       \
    long long gt_;                                                                                         \
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gt_));                                                \
    if (cta_ < 16384) g_dec_trace[cta_ * 8 + (slot)] = gt_;                                                \
  }
#else
#define DTRACE(slot)
#endif

// kCluster: the max_splits CTAs of a unit form ONE thread-block cluster (cluster dim x = grid dim x = 2 / 4) and
// the split merge goes through distributed shared memory: every CTA stores its normalised partial into the leader's
// (split 0) shared memory with st.shared::cluster, one cluster barrier, the leader merges and writes the output.  No
// global partials, no device-scope fence, no ticket atomic, no L2 round trip: ~2 us off a split decode
// (B=16 x 4096: 48.7 -> 47.2 us).  Two cluster barriers per CTA: #1 "my K/V ring is dead" (the partial region of
// the leader overlaps its ring), #2 "my partial is in the leader's memory".
constexpr int kDecClusterMax = 8;
__device__ __forceinline__ void cluster_arrive() { asm volatile("barrier.cluster.arrive.release.aligned;\n" ::: "memory"); }
__device__ __forceinline__ void cluster_wait() { asm volatile("barrier.cluster.wait.acquire.aligned;\n" ::: "memory"); }

constexpr int kMergeGroup = 8;  // partials the last arriver of a split unit loads per L2 round trip

template <int D, bool kBf16, int NW, bool kCluster = false>
__global__ void __launch_bounds__(32 * NW, NW == 2 ? 4 : 2) decode_kernel(const DecodeParams p) {
  constexpr int kDecTile = 16 * NW;     // shadows the 4-warp constants of decode.cuh
  constexpr int kDecThreads = 32 * NW;
  using T = typename std::conditional<kBf16, __nv_bfloat16, __half>::type;
  constexpr int kChunks = D / 8;                   // 16-byte chunks per row
  constexpr int kTileBytes = kDecTile * D * 2;     // one operand, one stage
  constexpr int kStageBytes = 2 * kTileBytes;      // K + V
  constexpr int kLoadsPerThread = kDecTile * kChunks / kDecThreads;  // per operand
  constexpr int kRowsPerPass = kDecThreads / kChunks;
  extern __shared__ __align__(128) uint8_t smem[];

  const int split = blockIdx.x;
  const int hg = blockIdx.y;  // kv head * groups_per_kvh + group
  const int item = blockIdx.z;
  const int kvh = hg / p.groups_per_kvh;
  const int grp = hg - kvh * p.groups_per_kvh;
  DTRACE(0);
  pdl_wait();     // everything below reads what earlier kernels of the stream wrote (indices, the new K/V rows)
  pdl_trigger();  // the next kernel may be scheduled into SMs as CTAs of this grid retire
  DTRACE(1);
  const int req = p.req_ids ? p.req_ids[item] : item;
  const int qrow = p.q_rows ? p.q_rows[item] : item;
  int kv_beg = p.kv_indptr[req];
  int L = p.kv_indptr[req + 1] - kv_beg;
  if (p.item_lens != nullptr) L = min(L, max(0, p.item_lens[item]));
  const int Dv = p.Dv;
  {
    const int lo = window_lo(L, p.window_left);  // sliding window: drop the keys before it
    kv_beg += lo;
    L -= lo;
  }
  int nsplit, chunk;
  split_policy(L, unit_max_splits(p, item, hg), p.min_chunk, kDecTile, nsplit, chunk);
  if (split >= nsplit) {
    if constexpr (kCluster) {  // nothing to do for this unit, but the cluster's barriers count every CTA
      cluster_arrive();
      cluster_wait();
      cluster_arrive();
      cluster_wait();
    }
    return;
  }
  const int k_lo = split * chunk;
  const int k_hi = min(L, k_lo + chunk);
  const int ntiles = (k_hi - k_lo + kDecTile - 1) / kDecTile;

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int h0 = kvh * p.G + grp * kDecMaxG;             // first q head of this CTA
  const int gcount = min(kDecMaxG, p.G - grp * kDecMaxG); // valid q heads (rows of M)

  // ---- producer state -------------------------------------------------------
  const int ld_chunk = tid % kChunks;
  const int ld_row0 = tid / kChunks;
  const bool ld_col_ok = ld_chunk * 8 < Dv;  // chunks past the real head dim are zero-filled
  const uint8_t* kc = reinterpret_cast<const uint8_t*>(p.k_cache) + (int64_t)kvh * Dv * 2 + (ld_col_ok ? ld_chunk * 16 : 0);
  const uint8_t* vc = reinterpret_cast<const uint8_t*>(p.v_cache) + (int64_t)kvh * Dv * 2 + (ld_col_ok ? ld_chunk * 16 : 0);
  const int64_t row_bytes = p.kv_stride * 2;
  const int32_t* idx = p.kv_indices + kv_beg;
  const uint32_t smem_base = smem_u32(smem);

  int32_t slots[kLoadsPerThread];  // slot ids of the NEXT tile to issue
  auto fetch_slots = [&](int tile) {
    const int base = k_lo + tile * kDecTile + ld_row0;
#pragma unroll
    for (int i = 0; i < kLoadsPerThread; ++i) {
      const int key = base + i * kRowsPerPass;
      slots[i] = (tile < ntiles && key < k_hi) ? __ldg(idx + key) : -1;
    }
  };
  auto issue_tile_from = [&](const int32_t (&sl)[kLoadsPerThread], int stage) {
    const uint32_t ks = smem_base + stage * kStageBytes;
    const uint32_t vs = ks + kTileBytes;
#pragma unroll
    for (int i = 0; i < kLoadsPerThread; ++i) {
      const int row = ld_row0 + i * kRowsPerPass;
      const uint32_t off = swz_off<D>(row, ld_chunk);
      const bool ok = sl[i] >= 0 && ld_col_ok;
      const int64_t g = ok ? (int64_t)sl[i] * row_bytes : 0;
      cp_async16(ks + off, kc + g, ok ? 16 : 0);  // invalid rows / columns are zero-filled
      cp_async16(vs + off, vc + g, ok ? 16 : 0);
    }
  };
  auto issue_tile = [&](int stage) { issue_tile_from(slots, stage); };

  // Prologue, ordered for memory-level parallelism (a warp issues in order: one dependent use stalls every load
  // behind it).  The slot ids of ALL prologue tiles and of the first steady-state tile go out first, then the Q
  // fragment loads, and only then the first instruction that needs a slot id.  (Fetching tile s+1's ids after
  // issuing tile s serialised two L2 round trips: 4.1 us from CTA start to the first tile, tools/trace_decode.py.)
  int32_t pslots[kDecStages - 1][kLoadsPerThread];
#pragma unroll
  for (int s = 0; s < kDecStages - 1; ++s) {
    fetch_slots(s);
#pragma unroll
    for (int i = 0; i < kLoadsPerThread; ++i) pslots[s][i] = slots[i];
  }
  fetch_slots(kDecStages - 1);
  // ---- Q fragments (A operand, 16 x D), rows >= gcount are zero -----------
  uint32_t qf[D / 16][4];
  {
    const T* qbase = reinterpret_cast<const T*>(p.q) + (int64_t)qrow * p.q_stride + (int64_t)h0 * Dv;
    const int r0 = lane >> 2, c0 = (lane & 3) * 2;
#pragma unroll
    for (int ks = 0; ks < D / 16; ++ks) {
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int r = r0 + (j & 1) * 8;
        const int c = ks * 16 + c0 + (j >> 1) * 8;
        qf[ks][j] = (r < gcount && c < Dv) ? *reinterpret_cast<const uint32_t*>(qbase + (int64_t)r * Dv + c) : 0u;
      }
    }
  }

#pragma unroll
  for (int s = 0; s < kDecStages - 1; ++s) {
    if (s < ntiles) issue_tile_from(pslots[s], s);
    cp_async_commit();
  }

  // ---- per-warp online-softmax state ---------------------------------------
  float o_acc[D / 8][4];
#pragma unroll
  for (int n = 0; n < D / 8; ++n) o_acc[n][0] = o_acc[n][1] = o_acc[n][2] = o_acc[n][3] = 0.f;
  float m_run[2] = {-INFINITY, -INFINITY};  // rows lane/4 and lane/4 + 8
  float l_run[2] = {0.f, 0.f};              // per-thread partial row sums

  const int lm = lane >> 3, lr = lane & 7;  // ldmatrix: matrix id / row in matrix

  DTRACE(2);
  for (int t = 0; t < ntiles; ++t) {
    cp_async_wait<kDecStages - 2>();
    __syncthreads();
    if (t == 0) { DTRACE(3); }
    {
      const int nt = t + kDecStages - 1;
      if (nt < ntiles) issue_tile(nt % kDecStages);
      cp_async_commit();
      fetch_slots(nt + 1);
    }
    const uint32_t ks = smem_base + (t % kDecStages) * kStageBytes;
    const uint32_t vs = ks + kTileBytes;
    const int key0 = k_lo + t * kDecTile + warp * 16;  // first key of this warp
    if (key0 < k_hi) {                                 // warp-uniform
      // S = Q K^T for 16 keys (two 8-key n-blocks)
      float s[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
      {
        const int krow = warp * 16 + (lm >> 1) * 8 + lr;
#pragma unroll
        for (int ksp = 0; ksp < D / 16; ++ksp) {
          uint32_t b0, b1, b2, b3;
          ldmatrix_x4(b0, b1, b2, b3, ks + swz_off<D>(krow, 2 * ksp + (lm & 1)));
          mma_16816<kBf16>(s[0], qf[ksp], b0, b1);
          mma_16816<kBf16>(s[1], qf[ksp], b2, b3);
        }
      }
      // scale, mask the ragged tail, online softmax (rows r and r+8 of the quad)
      const int kcol = key0 + (lane & 3) * 2;
      float tmax[2] = {-INFINITY, -INFINITY};
#pragma unroll
      for (int nb = 0; nb < 2; ++nb) {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int key = kcol + nb * 8 + (j & 1);
          float x = p.cap_post > 0.f ? p.cap_post * tanh_approx(s[nb][j] * p.cap_pre) : s[nb][j] * p.scale_log2;
          x = (key < k_hi) ? x : -INFINITY;
          s[nb][j] = x;
          tmax[j >> 1] = fmaxf(tmax[j >> 1], x);
        }
      }
      float alpha[2];
#pragma unroll
      for (int r = 0; r < 2; ++r) {
        tmax[r] = fmaxf(tmax[r], __shfl_xor_sync(0xffffffffu, tmax[r], 1));
        tmax[r] = fmaxf(tmax[r], __shfl_xor_sync(0xffffffffu, tmax[r], 2));
        const float m_new = fmaxf(m_run[r], tmax[r]);  // finite: key0 < k_hi => >= 1 valid key
        alpha[r] = fast_exp2(m_run[r] - m_new);
        m_run[r] = m_new;
        l_run[r] *= alpha[r];
      }
      uint32_t pf[4];
#pragma unroll
      for (int nb = 0; nb < 2; ++nb) {
        const float p0 = fast_exp2(s[nb][0] - m_run[0]);
        const float p1 = fast_exp2(s[nb][1] - m_run[0]);
        const float p2 = fast_exp2(s[nb][2] - m_run[1]);
        const float p3 = fast_exp2(s[nb][3] - m_run[1]);
        l_run[0] += p0 + p1;
        l_run[1] += p2 + p3;
        pf[nb * 2 + 0] = pack2<kBf16>(p0, p1);
        pf[nb * 2 + 1] = pack2<kBf16>(p2, p3);
      }
      // O = alpha * O + P V
      const int vrow = warp * 16 + (lm & 1) * 8 + lr;
#pragma unroll
      for (int np = 0; np < D / 16; ++np) {
        uint32_t b0, b1, b2, b3;
        ldmatrix_x4_trans(b0, b1, b2, b3, vs + swz_off<D>(vrow, 2 * np + (lm >> 1)));
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          float(&acc)[4] = o_acc[2 * np + h];
          acc[0] *= alpha[0];
          acc[1] *= alpha[0];
          acc[2] *= alpha[1];
          acc[3] *= alpha[1];
        }
        mma_16816<kBf16>(o_acc[2 * np], pf, b0, b1);
        mma_16816<kBf16>(o_acc[2 * np + 1], pf, b2, b3);
      }
    }
  }
  cp_async_wait<0>();
  __syncthreads();  // ring is dead; reuse it for the cross-warp merge
  if constexpr (kCluster) cluster_arrive();  // #1
  DTRACE(4);

  // ---- merge the four warps (fp32, shared memory) ---------------------------
#pragma unroll
  for (int r = 0; r < 2; ++r) {
    l_run[r] += __shfl_xor_sync(0xffffffffu, l_run[r], 1);
    l_run[r] += __shfl_xor_sync(0xffffffffu, l_run[r], 2);
  }
  float* sm_o = reinterpret_cast<float*>(smem);           // [NW][16][D]
  float* sm_m = sm_o + NW * 16 * D;                       // [NW][16]
  float* sm_l = sm_m + NW * 16;                           // [NW][16]
  {
    const int r0 = lane >> 2, c0 = (lane & 3) * 2;
#pragma unroll
    // only the gcount (<= 16) real q heads are written and read back: with G = 4 that is a quarter of the tile
    if (r0 < gcount) {
#pragma unroll
      for (int n = 0; n < D / 8; ++n)
        *reinterpret_cast<float2*>(sm_o + (warp * 16 + r0) * D + n * 8 + c0) = make_float2(o_acc[n][0], o_acc[n][1]);
    }
    if (r0 + 8 < gcount) {
#pragma unroll
      for (int n = 0; n < D / 8; ++n)
        *reinterpret_cast<float2*>(sm_o + (warp * 16 + r0 + 8) * D + n * 8 + c0) =
            make_float2(o_acc[n][2], o_acc[n][3]);
    }
    if ((lane & 3) == 0) {
      sm_m[warp * 16 + r0] = m_run[0];
      sm_m[warp * 16 + r0 + 8] = m_run[1];
      sm_l[warp * 16 + r0] = l_run[0];
      sm_l[warp * 16 + r0 + 8] = l_run[1];
    }
  }
  __syncthreads();
  const bool direct = (nsplit == 1);
  // leader's partial region: [kDecClusterMax][16][D] fp32 + [kDecClusterMax][16] LSE, after the local merge scratch
  constexpr uint32_t kClOff = (NW * 16 * D + 2 * NW * 16) * 4;
  uint32_t cl_base = 0;
  if constexpr (kCluster) {
    cluster_wait();  // #1: every ring of the cluster is dead, the leader's memory may be written
    asm volatile("mapa.shared::cluster.u32 %0, %1, 0;\n" : "=r"(cl_base) : "r"(smem_base + kClOff));
  }
  for (int e = tid; e < gcount * (D / 4); e += kDecThreads) {
    const int r = e / (D / 4), c = (e - r * (D / 4)) * 4;
    float m = -INFINITY;
#pragma unroll
    for (int w = 0; w < NW; ++w) m = fmaxf(m, sm_m[w * 16 + r]);
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    float l = 0.f;
#pragma unroll
    for (int w = 0; w < NW; ++w) {
      const float mw = sm_m[w * 16 + r];
      const float f = (mw == -INFINITY) ? 0.f : fast_exp2(mw - m);
      const float4 x = *reinterpret_cast<const float4*>(sm_o + (w * 16 + r) * D + c);
      acc.x += f * x.x;
      acc.y += f * x.y;
      acc.z += f * x.z;
      acc.w += f * x.w;
      l += f * sm_l[w * 16 + r];
    }
    if (direct && p.sinks != nullptr && l > 0.f) l += fast_exp2(__ldg(p.sinks + h0 + r) * kLog2e - m);
    const float inv = (l > 0.f) ? 1.f / l : 0.f;
    acc.x *= inv;
    acc.y *= inv;
    acc.z *= inv;
    acc.w *= inv;
    const float lse2 = (l > 0.f) ? m + log2f(l) : -INFINITY;
    const int h = h0 + r;
    if (direct) {
      T* dst = reinterpret_cast<T*>(p.o) + (int64_t)qrow * p.o_stride + (int64_t)h * Dv + c;
      uint2 pk;
      pk.x = pack2<kBf16>(acc.x, acc.y);
      pk.y = pack2<kBf16>(acc.z, acc.w);
      if (c < Dv) *reinterpret_cast<uint2*>(dst) = pk;
      if (p.lse != nullptr && c == 0) p.lse[(int64_t)qrow * p.Hq + h] = lse2 * kLn2;
    } else if constexpr (kCluster) {
      const uint32_t dst = cl_base + (uint32_t)(((split * 16 + r) * D + c) * 4);
      asm volatile("st.shared::cluster.v4.f32 [%0], {%1, %2, %3, %4};\n" ::"r"(dst), "f"(acc.x), "f"(acc.y),
                   "f"(acc.z), "f"(acc.w) : "memory");
      if (c == 0)
        asm volatile("st.shared::cluster.f32 [%0], %1;\n" ::"r"(cl_base + (uint32_t)((kDecClusterMax * 16 * D +
                                                                                     split * 16 + r) * 4)),
                     "f"(lse2) : "memory");
    } else {
      const int64_t w = ((int64_t)item * p.Hq + h) * p.max_splits + split;
      *reinterpret_cast<float4*>(p.ws_o + w * D + c) = acc;
      if (c == 0) p.ws_lse[w] = lse2;
    }
  }
  DTRACE(5);
  if constexpr (kCluster) {
    cluster_arrive();  // #2 (release: the partial stores above)
    cluster_wait();
    DTRACE(6);
    if (direct || split != 0) return;
    const float* cl_o = reinterpret_cast<const float*>(smem + kClOff);
    const float* cl_lse = cl_o + kDecClusterMax * 16 * D;
    for (int e = tid; e < gcount * (D / 4); e += kDecThreads) {
      const int r = e / (D / 4), c = (e - r * (D / 4)) * 4;
      const int h = h0 + r;
      float m = -INFINITY;
      for (int sp = 0; sp < nsplit; ++sp) m = fmaxf(m, cl_lse[sp * 16 + r]);
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
      float l = 0.f;
      for (int sp = 0; sp < nsplit; ++sp) {
        const float ls = cl_lse[sp * 16 + r];
        const float f = (ls == -INFINITY) ? 0.f : fast_exp2(ls - m);
        const float4 x = *reinterpret_cast<const float4*>(cl_o + (sp * 16 + r) * D + c);
        acc.x += f * x.x;
        acc.y += f * x.y;
        acc.z += f * x.z;
        acc.w += f * x.w;
        l += f;
      }
      if (p.sinks != nullptr && l > 0.f) l += fast_exp2(__ldg(p.sinks + h) * kLog2e - m);
      const float inv = (l > 0.f) ? 1.f / l : 0.f;
      T* dst = reinterpret_cast<T*>(p.o) + (int64_t)qrow * p.o_stride + (int64_t)h * Dv + c;
      uint2 pk;
      pk.x = pack2<kBf16>(acc.x * inv, acc.y * inv);
      pk.y = pack2<kBf16>(acc.z * inv, acc.w * inv);
      if (c < Dv) *reinterpret_cast<uint2*>(dst) = pk;
      if (p.lse != nullptr && c == 0)
        p.lse[(int64_t)qrow * p.Hq + h] = (l > 0.f) ? (m + log2f(l)) * kLn2 : -INFINITY;
    }
    DTRACE(7);
    return;
  }
  if (direct) return;
  // ---- split merge by the LAST CTA of the unit to arrive (no second launch) ----
  // writers: partials -> fence -> barrier -> one relaxed atomic ticket; the CTA that draws the last ticket sees
  // every partial (fence after the ticket) and reads them through L2 (ld.global.cg, never the read-only path).
  __shared__ int s_last;
  __syncthreads();  // every thread's partial stores happen-before thread 0's fence (cumulativity carries them)
  if (tid == 0) {
    int* cnt = p.ws_cnt + (int64_t)item * (p.Hkv * p.groups_per_kvh) + hg;
    __threadfence();
    const int ticket = atomicAdd(cnt, 1);
    s_last = (ticket == nsplit - 1);
    if (s_last) {
      *cnt = 0;          // self-cleaning: the next launch finds zeros again
      __threadfence();   // acquire side: the other CTAs' partials, ordered before their tickets, are visible below
    }
  }
  __syncthreads();
  DTRACE(6);
  if (!s_last) return;
  for (int e = tid; e < gcount * (D / 4); e += kDecThreads) {
    const int r = e / (D / 4), c = (e - r * (D / 4)) * 4;
    const int h = h0 + r;
    const int64_t w0 = ((int64_t)item * p.Hq + h) * p.max_splits;
    // one pass, kMergeGroup partials in flight at a time (all loads of a group are issued before the first use),
    // running max carried across groups: an up-to-8-way split costs ONE L2 round trip instead of two dependent passes
    float m = -INFINITY;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    float l = 0.f;
    for (int sp0 = 0; sp0 < nsplit; sp0 += kMergeGroup) {
      float ls[kMergeGroup];
      float4 x[kMergeGroup];
#pragma unroll
      for (int u = 0; u < kMergeGroup; ++u) {
        const bool ok = sp0 + u < nsplit;
        ls[u] = ok ? __ldcg(p.ws_lse + w0 + sp0 + u) : -INFINITY;
        x[u] = ok ? __ldcg(reinterpret_cast<const float4*>(p.ws_o + (w0 + sp0 + u) * D + c))
                  : make_float4(0.f, 0.f, 0.f, 0.f);
      }
      float mg = ls[0];
#pragma unroll
      for (int u = 1; u < kMergeGroup; ++u) mg = fmaxf(mg, ls[u]);
      if (mg > m) {  // rescale what has been accumulated so far
        const float f0 = (m == -INFINITY) ? 0.f : fast_exp2(m - mg);
        acc.x *= f0;
        acc.y *= f0;
        acc.z *= f0;
        acc.w *= f0;
        l *= f0;
        m = mg;
      }
#pragma unroll
      for (int u = 0; u < kMergeGroup; ++u) {
        const float f = (ls[u] == -INFINITY) ? 0.f : fast_exp2(ls[u] - m);
        acc.x += f * x[u].x;
        acc.y += f * x[u].y;
        acc.z += f * x[u].z;
        acc.w += f * x[u].w;
        l += f;
      }
    }
    if (p.sinks != nullptr && l > 0.f) l += fast_exp2(__ldg(p.sinks + h) * kLog2e - m);
    const float inv = (l > 0.f) ? 1.f / l : 0.f;
    T* dst = reinterpret_cast<T*>(p.o) + (int64_t)qrow * p.o_stride + (int64_t)h * Dv + c;
    uint2 pk;
    pk.x = pack2<kBf16>(acc.x * inv, acc.y * inv);
    pk.y = pack2<kBf16>(acc.z * inv, acc.w * inv);
    if (c < Dv) *reinterpret_cast<uint2*>(dst) = pk;
    if (p.lse != nullptr && c == 0)
      p.lse[(int64_t)qrow * p.Hq + h] = (l > 0.f) ? (m + log2f(l)) * kLn2 : -INFINITY;
  }
  DTRACE(7);
}

// ---------------------------------------------------------------------------------------------------
// FP8 (e4m3) KV cache variant (pool dtype float8_e4m3fn stored as bytes,
// memory_pool.py:2365-2375; k/v descale as in triton_backend.py:1418-1420).
//
// Same CTA shape and split machinery as decode_kernel; what changes is the byte path:
//   * a K/V row is D BYTES, a 64-key tile 64*D bytes per operand, so the ring is 6 stages deep for the
//     same 96 KB of loads in flight per CTA;
//   * ldmatrix (b16) moves byte PAIRS; the contraction index of QK^T is permuted so that the four
//     bytes a lane receives are its k-slots {0,1} and {8,9} of one m16n8k16 (the Q fragment is loaded
//     with the same permutation), and for PV the transposed b16 load hands a lane V[key, d..d+1] for
//     two keys, which two byte permutes turn into the B fragments of an "even d" and an "odd d" mma;
//     a lane ends up owning 4 consecutive output dims per 16-d chunk;
//   * e4m3 -> f16 is exact and branch-free: (b & 0x7f) << 7 reinterpreted as f16 is value * 2^-8
//     (denormals included), so the mma runs in f16 (bf16 queries are converted once; P in f16) and the
//     2^8 goes into the score scale with k_descale, and into the output scale with v_descale.
template <int D, bool kBf16>
__global__ void __launch_bounds__(kDecThreads, B200ATTN_FP8_CTAS) decode_fp8_kernel(const DecodeParams p) {
  using T = typename std::conditional<kBf16, __nv_bfloat16, __half>::type;
  constexpr int kChunks = D / 16;                  // 16-byte chunks per row
  constexpr int kTileBytes = kDecTile * D;         // one operand, one stage
  constexpr int kStageBytes = 2 * kTileBytes;      // K + V
  constexpr int kLoadsPerThread = kDecTile * kChunks / kDecThreads;
  constexpr int kRowsPerPass = kDecThreads / kChunks;
  extern __shared__ __align__(128) uint8_t smem[];

  const int split = blockIdx.x;
  const int hg = blockIdx.y;
  const int item = blockIdx.z;
  const int kvh = hg / p.groups_per_kvh;
  const int grp = hg - kvh * p.groups_per_kvh;
  pdl_wait();     // everything below reads what earlier kernels of the stream wrote (indices, the new K/V rows)
  pdl_trigger();  // the next kernel may be scheduled into SMs as CTAs of this grid retire
  const int req = p.req_ids ? p.req_ids[item] : item;
  const int qrow = p.q_rows ? p.q_rows[item] : item;
  int kv_beg = p.kv_indptr[req];
  int L = p.kv_indptr[req + 1] - kv_beg;
  {
    const int lo = window_lo(L, p.window_left);
    kv_beg += lo;
    L -= lo;
  }
  int nsplit, chunk;
  split_policy(L, unit_max_splits(p, item, hg), p.min_chunk, kDecTile, nsplit, chunk);
  if (split >= nsplit) return;
  const int k_lo = split * chunk;
  const int k_hi = min(L, k_lo + chunk);
  const int ntiles = (k_hi - k_lo + kDecTile - 1) / kDecTile;

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int h0 = kvh * p.G + grp * kDecMaxG;
  const int gcount = min(kDecMaxG, p.G - grp * kDecMaxG);

  const int ld_chunk = tid % kChunks;
  const int ld_row0 = tid / kChunks;
  const uint8_t* kc8 = reinterpret_cast<const uint8_t*>(p.k_cache) + (int64_t)kvh * D + ld_chunk * 16;
  const uint8_t* vc8 = reinterpret_cast<const uint8_t*>(p.v_cache) + (int64_t)kvh * D + ld_chunk * 16;
  const int64_t row_bytes = p.kv_stride;
  const int32_t* idx = p.kv_indices + kv_beg;
  const uint32_t smem_base = smem_u32(smem);

  int32_t slots[kLoadsPerThread];
  auto fetch_slots = [&](int tile) {
    const int base = k_lo + tile * kDecTile + ld_row0;
#pragma unroll
    for (int i = 0; i < kLoadsPerThread; ++i) {
      const int key = base + i * kRowsPerPass;
      slots[i] = (tile < ntiles && key < k_hi) ? __ldg(idx + key) : -1;
    }
  };
  auto issue_tile_from = [&](const int32_t (&sl)[kLoadsPerThread], int stage) {
    const uint32_t ks = smem_base + stage * kStageBytes;
    const uint32_t vs = ks + kTileBytes;
#pragma unroll
    for (int i = 0; i < kLoadsPerThread; ++i) {
      const int row = ld_row0 + i * kRowsPerPass;
      const uint32_t off = swz8_off<D>(row, ld_chunk);
      const bool ok = sl[i] >= 0;
      const int64_t g = ok ? (int64_t)sl[i] * row_bytes : 0;
      cp_async16(ks + off, kc8 + g, ok ? 16 : 0);
      cp_async16(vs + off, vc8 + g, ok ? 16 : 0);
    }
  };
  auto issue_tile = [&](int stage) { issue_tile_from(slots, stage); };

  // prologue ordered for memory-level parallelism, as in decode_kernel: every slot-id load first, then Q, then uses
  int32_t pslots[kDec8Stages - 1][kLoadsPerThread];
#pragma unroll
  for (int s = 0; s < kDec8Stages - 1; ++s) {
    fetch_slots(s);
#pragma unroll
    for (int i = 0; i < kLoadsPerThread; ++i) pslots[s][i] = slots[i];
  }
  fetch_slots(kDec8Stages - 1);
  // ---- Q fragments in f16, contraction index permuted: chunk kc, lane quad slot t holds d = 16kc + 4t + {0,1}
  // (k-slots 2t,2t+1) and d = 16kc + 4t + {2,3} (k-slots 2t+8, 2t+9) -----------------------------------------
  uint32_t qf[D / 16][4];
  {
    const T* qbase = reinterpret_cast<const T*>(p.q) + (int64_t)qrow * p.q_stride + (int64_t)h0 * D;
    const int r0 = lane >> 2, c0 = (lane & 3) * 4;
#pragma unroll
    for (int kc = 0; kc < D / 16; ++kc) {
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int r = r0 + (j & 1) * 8;
        const int c = kc * 16 + c0 + (j >> 1) * 2;
        uint32_t w = (r < gcount) ? *reinterpret_cast<const uint32_t*>(qbase + (int64_t)r * D + c) : 0u;
        if constexpr (kBf16) {
          const float2 f = unpack2<true>(w);
          w = pack2<false>(f.x, f.y);
        }
        qf[kc][j] = w;
      }
    }
  }

#pragma unroll
  for (int s = 0; s < kDec8Stages - 1; ++s) {
    if (s < ntiles) issue_tile_from(pslots[s], s);
    cp_async_commit();
  }

  // o_acc[2c + e]: chunk c (16 output dims), e = 0: dims 16c + 4t + {0,2}, e = 1: dims 16c + 4t + {1,3}
  float o_acc[D / 8][4];
#pragma unroll
  for (int n = 0; n < D / 8; ++n) o_acc[n][0] = o_acc[n][1] = o_acc[n][2] = o_acc[n][3] = 0.f;
  float m_run[2] = {-INFINITY, -INFINITY};
  float l_run[2] = {0.f, 0.f};

  const int lm = lane >> 3, lr = lane & 7;

  for (int t = 0; t < ntiles; ++t) {
    cp_async_wait<kDec8Stages - 2>();
    __syncthreads();
    {
      const int nt = t + kDec8Stages - 1;
      if (nt < ntiles) issue_tile(nt % kDec8Stages);
      cp_async_commit();
      fetch_slots(nt + 1);
    }
    const uint32_t ks = smem_base + (t % kDec8Stages) * kStageBytes;
    const uint32_t vs = ks + kTileBytes;
    const int key0 = k_lo + t * kDecTile + warp * 16;
    if (key0 < k_hi) {
      float s[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
      {
        const int krow = warp * 16 + (lm >> 1) * 8 + lr;
#pragma unroll
        for (int kp = 0; kp < D / 32; ++kp) {
          uint32_t m0, m1, m2, m3;  // (keys 0-7, chunk 2kp), (keys 0-7, 2kp+1), (keys 8-15, 2kp), (keys 8-15, 2kp+1)
          ldmatrix_x4(m0, m1, m2, m3, ks + swz8_off<D>(krow, 2 * kp + (lm & 1)));
          uint32_t a, b;
          e4m3x4_pairs(m0, a, b);
          mma_16816<false>(s[0], qf[2 * kp], a, b);
          e4m3x4_pairs(m1, a, b);
          mma_16816<false>(s[0], qf[2 * kp + 1], a, b);
          e4m3x4_pairs(m2, a, b);
          mma_16816<false>(s[1], qf[2 * kp], a, b);
          e4m3x4_pairs(m3, a, b);
          mma_16816<false>(s[1], qf[2 * kp + 1], a, b);
        }
      }
      const int kcol = key0 + (lane & 3) * 2;
      float tmax[2] = {-INFINITY, -INFINITY};
#pragma unroll
      for (int nb = 0; nb < 2; ++nb) {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int key = kcol + nb * 8 + (j & 1);
          float x = p.cap_post > 0.f ? p.cap_post * tanh_approx(s[nb][j] * p.cap_pre) : s[nb][j] * p.scale_log2;
          x = (key < k_hi) ? x : -INFINITY;
          s[nb][j] = x;
          tmax[j >> 1] = fmaxf(tmax[j >> 1], x);
        }
      }
      float alpha[2];
#pragma unroll
      for (int r = 0; r < 2; ++r) {
        tmax[r] = fmaxf(tmax[r], __shfl_xor_sync(0xffffffffu, tmax[r], 1));
        tmax[r] = fmaxf(tmax[r], __shfl_xor_sync(0xffffffffu, tmax[r], 2));
        const float m_new = fmaxf(m_run[r], tmax[r]);
        alpha[r] = fast_exp2(m_run[r] - m_new);
        m_run[r] = m_new;
        l_run[r] *= alpha[r];
      }
      uint32_t pf[4];
#pragma unroll
      for (int nb = 0; nb < 2; ++nb) {
        const float p0 = fast_exp2(s[nb][0] - m_run[0]);
        const float p1 = fast_exp2(s[nb][1] - m_run[0]);
        const float p2 = fast_exp2(s[nb][2] - m_run[1]);
        const float p3 = fast_exp2(s[nb][3] - m_run[1]);
        l_run[0] += p0 + p1;
        l_run[1] += p2 + p3;
        pf[nb * 2 + 0] = pack2<false>(p0, p1);
        pf[nb * 2 + 1] = pack2<false>(p2, p3);
      }
      const int vrow = warp * 16 + (lm & 1) * 8 + lr;
      const bool rescale = __any_sync(0xffffffffu, alpha[0] != 1.f || alpha[1] != 1.f);
#pragma unroll
      for (int np = 0; np < D / 32; ++np) {
        uint32_t m0, m1, m2, m3;  // (keys 0-7, chunk 2np), (keys 8-15, 2np), (keys 0-7, 2np+1), (keys 8-15, 2np+1)
        ldmatrix_x4_trans(m0, m1, m2, m3, vs + swz8_off<D>(vrow, 2 * np + (lm >> 1)));
        if (rescale) {  // warp-uniform: the running max moved in this tile
#pragma unroll
          for (int h = 0; h < 4; ++h) {
            float(&acc)[4] = o_acc[4 * np + h];
            acc[0] *= alpha[0];
            acc[1] *= alpha[0];
            acc[2] *= alpha[1];
            acc[3] *= alpha[1];
          }
        }
        // a lane's word: byte0 = V[k0][d], byte1 = V[k0][d+1], byte2 = V[k0+1][d], byte3 = V[k0+1][d+1]
        uint32_t e0, o0, e1, o1;
        e4m3x4_cross(m0, e0, o0);
        e4m3x4_cross(m1, e1, o1);
        mma_16816<false>(o_acc[4 * np + 0], pf, e0, e1);
        mma_16816<false>(o_acc[4 * np + 1], pf, o0, o1);
        e4m3x4_cross(m2, e0, o0);
        e4m3x4_cross(m3, e1, o1);
        mma_16816<false>(o_acc[4 * np + 2], pf, e0, e1);
        mma_16816<false>(o_acc[4 * np + 3], pf, o0, o1);
      }
    }
  }
  cp_async_wait<0>();
  __syncthreads();

#pragma unroll
  for (int r = 0; r < 2; ++r) {
    l_run[r] += __shfl_xor_sync(0xffffffffu, l_run[r], 1);
    l_run[r] += __shfl_xor_sync(0xffffffffu, l_run[r], 2);
  }
  float* sm_o = reinterpret_cast<float*>(smem);
  float* sm_m = sm_o + 4 * 16 * D;
  float* sm_l = sm_m + 64;
  {
    const int r0 = lane >> 2, c0 = (lane & 3) * 4;
#pragma unroll
    for (int c = 0; c < D / 16; ++c) {  // only the gcount real q-head rows are read back
      float* dst = sm_o + (warp * 16 + r0) * D + c * 16 + c0;
      if (r0 < gcount)
        *reinterpret_cast<float4*>(dst) =
            make_float4(o_acc[2 * c][0], o_acc[2 * c + 1][0], o_acc[2 * c][1], o_acc[2 * c + 1][1]);
      if (r0 + 8 < gcount)
        *reinterpret_cast<float4*>(dst + 8 * D) =
            make_float4(o_acc[2 * c][2], o_acc[2 * c + 1][2], o_acc[2 * c][3], o_acc[2 * c + 1][3]);
    }
    if ((lane & 3) == 0) {
      sm_m[warp * 16 + r0] = m_run[0];
      sm_m[warp * 16 + r0 + 8] = m_run[1];
      sm_l[warp * 16 + r0] = l_run[0];
      sm_l[warp * 16 + r0 + 8] = l_run[1];
    }
  }
  __syncthreads();
  const bool direct = (nsplit == 1);
  for (int e = tid; e < gcount * (D / 4); e += kDecThreads) {
    const int r = e / (D / 4), c = (e - r * (D / 4)) * 4;
    float m = -INFINITY;
#pragma unroll
    for (int w = 0; w < 4; ++w) m = fmaxf(m, sm_m[w * 16 + r]);
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    float l = 0.f;
#pragma unroll
    for (int w = 0; w < 4; ++w) {
      const float mw = sm_m[w * 16 + r];
      const float f = (mw == -INFINITY) ? 0.f : fast_exp2(mw - m);
      const float4 x = *reinterpret_cast<const float4*>(sm_o + (w * 16 + r) * D + c);
      acc.x += f * x.x;
      acc.y += f * x.y;
      acc.z += f * x.z;
      acc.w += f * x.w;
      l += f * sm_l[w * 16 + r];
    }
    if (direct && p.sinks != nullptr && l > 0.f) l += fast_exp2(__ldg(p.sinks + h0 + r) * kLog2e - m);
    const float inv = (l > 0.f) ? p.out_scale / l : 0.f;  // v_descale * 2^8 rides on the normalisation
    acc.x *= inv;
    acc.y *= inv;
    acc.z *= inv;
    acc.w *= inv;
    const float lse2 = (l > 0.f) ? m + log2f(l) : -INFINITY;
    const int h = h0 + r;
    if (direct) {
      T* dst = reinterpret_cast<T*>(p.o) + (int64_t)qrow * p.o_stride + (int64_t)h * D + c;
      uint2 pk;
      pk.x = pack2<kBf16>(acc.x, acc.y);
      pk.y = pack2<kBf16>(acc.z, acc.w);
      *reinterpret_cast<uint2*>(dst) = pk;
      if (p.lse != nullptr && c == 0) p.lse[(int64_t)qrow * p.Hq + h] = lse2 * kLn2;
    } else {
      const int64_t w = ((int64_t)item * p.Hq + h) * p.max_splits + split;
      *reinterpret_cast<float4*>(p.ws_o + w * D + c) = acc;
      if (c == 0) p.ws_lse[w] = lse2;
    }
  }
  if (direct) return;
  __shared__ int s_last;
  __syncthreads();  // as in decode_kernel: one thread fences on either side of the ticket
  if (tid == 0) {
    int* cnt = p.ws_cnt + (int64_t)item * (p.Hkv * p.groups_per_kvh) + hg;
    __threadfence();
    const int ticket = atomicAdd(cnt, 1);
    s_last = (ticket == nsplit - 1);
    if (s_last) {
      *cnt = 0;
      __threadfence();
    }
  }
  __syncthreads();
  if (!s_last) return;
  for (int e = tid; e < gcount * (D / 4); e += kDecThreads) {
    const int r = e / (D / 4), c = (e - r * (D / 4)) * 4;
    const int h = h0 + r;
    const int64_t w0 = ((int64_t)item * p.Hq + h) * p.max_splits;
    float m = -INFINITY;  // one pass, kMergeGroup partials per L2 round trip (see decode_kernel)
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    float l = 0.f;
    for (int sp0 = 0; sp0 < nsplit; sp0 += kMergeGroup) {
      float ls[kMergeGroup];
      float4 x[kMergeGroup];
#pragma unroll
      for (int u = 0; u < kMergeGroup; ++u) {
        const bool ok = sp0 + u < nsplit;
        ls[u] = ok ? __ldcg(p.ws_lse + w0 + sp0 + u) : -INFINITY;
        x[u] = ok ? __ldcg(reinterpret_cast<const float4*>(p.ws_o + (w0 + sp0 + u) * D + c))
                  : make_float4(0.f, 0.f, 0.f, 0.f);
      }
      float mg = ls[0];
#pragma unroll
      for (int u = 1; u < kMergeGroup; ++u) mg = fmaxf(mg, ls[u]);
      if (mg > m) {
        const float f0 = (m == -INFINITY) ? 0.f : fast_exp2(m - mg);
        acc.x *= f0;
        acc.y *= f0;
        acc.z *= f0;
        acc.w *= f0;
        l *= f0;
        m = mg;
      }
#pragma unroll
      for (int u = 0; u < kMergeGroup; ++u) {
        const float f = (ls[u] == -INFINITY) ? 0.f : fast_exp2(ls[u] - m);
        acc.x += f * x[u].x;
        acc.y += f * x[u].y;
        acc.z += f * x[u].z;
        acc.w += f * x[u].w;
        l += f;
      }
    }
    if (p.sinks != nullptr && l > 0.f) l += fast_exp2(__ldg(p.sinks + h) * kLog2e - m);
    const float inv = (l > 0.f) ? 1.f / l : 0.f;  // the partials already carry out_scale
    T* dst = reinterpret_cast<T*>(p.o) + (int64_t)qrow * p.o_stride + (int64_t)h * D + c;
    uint2 pk;
    pk.x = pack2<kBf16>(acc.x * inv, acc.y * inv);
    pk.y = pack2<kBf16>(acc.z * inv, acc.w * inv);
    *reinterpret_cast<uint2*>(dst) = pk;
    if (p.lse != nullptr && c == 0)
      p.lse[(int64_t)qrow * p.Hq + h] = (l > 0.f) ? (m + log2f(l)) * kLn2 : -INFINITY;
  }
}

template <int D, bool kBf16>
static int launch_decode_fp8(const DecodeParams& p, cudaStream_t st) {
  constexpr int kRing = kDec8Stages * 2 * kDecTile * D;
  constexpr int kMerge = (4 * 16 * D + 128) * 4;
  constexpr int kSmem = kRing > kMerge ? kRing : kMerge;
  static bool configured = false;
  if (!configured) {
    B2_CUDA(cudaFuncSetAttribute(decode_fp8_kernel<D, kBf16>,
                                 cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem));
    B2_CUDA(cudaFuncSetAttribute(decode_fp8_kernel<D, kBf16>,
                                 cudaFuncAttributePreferredSharedMemoryCarveout, 100));
    configured = true;
  }
  dim3 grid((unsigned)p.max_splits, (unsigned)(p.Hkv * p.groups_per_kvh), (unsigned)p.B);
  B2_CUDA(launch_pdl(decode_fp8_kernel<D, kBf16>, grid, dim3(kDecThreads), kSmem, st, p));
  return B200ATTN_OK;
}

template <int D, bool kBf16, int NW>
static int launch_decode_nw(const DecodeParams& p, cudaStream_t st) {
  constexpr int kRing = kDecStages * 2 * (16 * NW) * D * 2;
  constexpr int kMerge = (NW * 16 * D + 2 * NW * 16) * 4;
  constexpr int kSmem = kRing > kMerge ? kRing : kMerge;
  static bool configured = false;  // idempotent attribute; races are benign
  if (!configured) {
    B2_CUDA(cudaFuncSetAttribute(decode_kernel<D, kBf16, NW>,
                                 cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem));
    B2_CUDA(cudaFuncSetAttribute(decode_kernel<D, kBf16, NW>,
                                 cudaFuncAttributePreferredSharedMemoryCarveout, 100));
    configured = true;
  }
  dim3 grid((unsigned)p.max_splits, (unsigned)(p.Hkv * p.groups_per_kvh), (unsigned)p.B);
  B2_CUDA(launch_pdl(decode_kernel<D, kBf16, NW>, grid, dim3(32 * NW), kSmem, st, p));
  return B200ATTN_OK;
}

// 4-warp kernel with the split merge through cluster shared memory.  Used for max_splits = 2 / 4 (measured, same
// process A/B with B200ATTN_DEC_CLUSTER=0: B=16 x 4096 48.7 -> 47.2 us, B=8 x 8192 48.4 -> 47.1 us); clusters of 8 are
// SLOWER than the workspace path (52.2 -> 60.5 us): the scheduler packs a cluster's CTAs two per SM inside one GPC
// and leaves other SMs empty, and a lightly loaded grid is bound by per-SM load bandwidth.
template <int D, bool kBf16>
static int launch_decode_cluster(const DecodeParams& p, cudaStream_t st) {
  constexpr int NW = 4;
  constexpr int kRing = kDecStages * 2 * (16 * NW) * D * 2;
  constexpr int kMerge = (NW * 16 * D + 2 * NW * 16) * 4 + (kDecClusterMax * 16 * D + kDecClusterMax * 16) * 4;
  constexpr int kSmem = kRing > kMerge ? kRing : kMerge;
  static bool configured = false;
  if (!configured) {
    B2_CUDA(cudaFuncSetAttribute(decode_kernel<D, kBf16, NW, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem));
    B2_CUDA(cudaFuncSetAttribute(decode_kernel<D, kBf16, NW, true>, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
    configured = true;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)p.max_splits, (unsigned)(p.Hkv * p.groups_per_kvh), (unsigned)p.B);
  cfg.blockDim = dim3(32 * NW);
  cfg.dynamicSmemBytes = kSmem;
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
  attr[1].id = cudaLaunchAttributeClusterDimension;
  attr[1].val.clusterDim.x = (unsigned)p.max_splits;
  attr[1].val.clusterDim.y = 1;
  attr[1].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 2;
  B2_CUDA(cudaLaunchKernelEx(&cfg, decode_kernel<D, kBf16, NW, true>, p));
  return B200ATTN_OK;
}

static bool decode_cluster_enabled() {
  static const bool on = [] { const char* e = getenv("B200ATTN_DEC_CLUSTER"); return e == nullptr || atoi(e) != 0; }();
  return on;
}

template <int D, bool kBf16>
static int launch_decode(const DecodeParams& p, int warps, cudaStream_t st) {
  // not with a tail split (tail_begin >= 0): there most units are NOT split and their second CTA would sit in the
  // cluster barrier, holding an SM slot, until its partner has streamed the whole unit (cfg5 shard: 0.62 -> 0.85 ms)
  if (warps == 4 && (p.max_splits == 2 || p.max_splits == 4) && p.tail_begin < 0 && decode_cluster_enabled())
    return launch_decode_cluster<D, kBf16>(p, st);
  return warps == 2 ? launch_decode_nw<D, kBf16, 2>(p, st) : launch_decode_nw<D, kBf16, 4>(p, st);
}

}  // namespace b200attn

// resident CTAs per SM of the kernel b200attn_decode / b200attn_decode_fp8kv will launch (the wave arithmetic of the
// host-side split heuristics needs it)
extern "C" int b200attn_decode_ctas_per_sm(int kv_fp8, int warps) {
  return kv_fp8 ? B200ATTN_FP8_CTAS : (warps == 2 ? 4 : 2);  // head dims above 128: two (always the 2-warp kernel)
}

extern "C" int64_t b200attn_decode_workspace_bytes(int32_t B, int32_t Hq, int32_t D,
                                                   int32_t max_splits) {
  if (max_splits <= 1) return 0;
  D = D <= 64 ? 64 : (D <= 128 ? 128 : 256);  // the kernels run on the padded head dim
  // fixed-size counter header (kDecCounterBytes) + fp32 partials and their LSE
  return b200attn::kDecCounterBytes + (int64_t)B * Hq * max_splits * (D + 1) * 4;
}

static int decode_impl(const void* q, int64_t q_stride, const void* k_cache,
                       const void* v_cache, int64_t kv_stride, void* o, int64_t o_stride,
                       float* lse, const int32_t* kv_indptr, const int32_t* kv_indices,
                       const int32_t* q_rows, const int32_t* req_ids, int32_t B,
                       int32_t Hq, int32_t Hkv, int32_t D, int dtype, float sm_scale,
                       int32_t max_splits, int32_t min_chunk, int32_t tail_begin,
                       int32_t window_left,
                       float logit_cap, const float* sinks, void* workspace,
                       int64_t workspace_bytes, void* stream, bool kv_fp8, float k_descale,
                       float v_descale, int32_t warps = 4, const int32_t* item_lens = nullptr) {
  using namespace b200attn;
  B2_CHECK_ARG(B >= 0, "decode: negative batch");
  B2_CHECK_ARG(logit_cap >= 0.f, "decode: negative logit_cap");
  if (B == 0) return B200ATTN_OK;
  B2_CHECK_ARG(q && k_cache && v_cache && o && kv_indptr && kv_indices, "decode: null pointer");
  B2_CHECK_ARG(Hq > 0 && Hkv > 0 && Hq % Hkv == 0, "decode: Hq=%d must be a multiple of Hkv=%d",
               Hq, Hkv);
  B2_CHECK_ARG(B <= 65535 && Hq <= 65535, "decode: grid too large");
  B2_CHECK_ARG(max_splits >= 1 && max_splits <= 1024 && min_chunk >= 1, "decode: bad split policy");
  B2_CHECK_ARG(warps == 2 || warps == 4, "decode: warps must be 2 or 4");
  B2_CHECK_ARG(q_stride % 8 == 0 && kv_stride % (kv_fp8 ? 16 : 8) == 0 && o_stride % 4 == 0,
               "decode: strides must keep rows 16-byte aligned");
  B2_CHECK_ARG(!kv_fp8 || (k_descale > 0.f && v_descale > 0.f), "decode: fp8 descales must be positive");
  B2_CHECK_ARG((((uintptr_t)q | (uintptr_t)k_cache | (uintptr_t)v_cache | (uintptr_t)o) & 15) == 0,
               "decode: tensors must be 16-byte aligned");
  if (dtype != B200ATTN_BF16 && dtype != B200ATTN_FP16) B2_UNSUPPORTED("decode: dtype %d", dtype);
  // 16-bit pools: any head dim that is a multiple of 8 up to 256 runs on the padded 64 / 128 / 256 instantiation
  // (256: Gemma-style heads; 64-thread CTAs with 32-key tiles only, two per SM - the 96 KB ring sets that)
  const int Dv = D;
  if (kv_fp8 ? (D != 64 && D != 128) : (D < 8 || D > 256 || D % 8 != 0))
    B2_UNSUPPORTED("decode: head_dim %d (supported: multiples of 8 up to 256; e4m3 pools: 64, 128)", D);
  B2_CHECK_ARG(!kv_fp8 || item_lens == nullptr, "decode: item_lens is not supported over e4m3 pools");
  D = D <= 64 ? 64 : (D <= 128 ? 128 : 256);
  const int64_t need = b200attn_decode_workspace_bytes(B, Hq, D, max_splits);
  B2_CHECK_ARG(need == 0 || (workspace && workspace_bytes >= need),
               "decode: workspace too small (%lld < %lld)", (long long)workspace_bytes,
               (long long)need);
  B2_CHECK_ARG(need == 0 || (int64_t)B * Hkv * ((Hq / Hkv + kDecMaxG - 1) / kDecMaxG) <= kDecMaxUnits,
               "decode: more than %d units with max_splits > 1", kDecMaxUnits);
  DecodeParams p;
  p.q = q;
  p.k_cache = k_cache;
  p.v_cache = v_cache;
  p.o = o;
  p.lse = lse;
  p.kv_indptr = kv_indptr;
  p.kv_indices = kv_indices;
  p.q_rows = q_rows;
  p.req_ids = req_ids;
  p.ws_cnt = need > 0 ? reinterpret_cast<int*>(workspace) : nullptr;
  p.ws_o = need > 0 ? reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(workspace) + kDecCounterBytes) : nullptr;
  p.ws_lse = p.ws_o ? p.ws_o + (int64_t)B * Hq * max_splits * D : nullptr;
  p.q_stride = q_stride;
  p.kv_stride = kv_stride;
  p.o_stride = o_stride;
  p.B = B;
  p.Hq = Hq;
  p.Hkv = Hkv;
  p.G = Hq / Hkv;
  p.groups_per_kvh = (p.G + kDecMaxG - 1) / kDecMaxG;
  p.max_splits = max_splits;
  p.min_chunk = min_chunk;
  // fp8 KV: the bit-trick e4m3 -> f16 conversion yields value * 2^-8 (kFp8Bias = 2^8 then rides on both scales)
  if (kv_fp8) sm_scale *= k_descale * kFp8Bias;
  p.out_scale = kv_fp8 ? v_descale * kFp8Bias : 1.f;
  p.scale_log2 = sm_scale * kLog2e;
  p.window_left = window_left;
  p.cap_pre = logit_cap > 0.f ? sm_scale / logit_cap : 0.f;
  p.cap_post = logit_cap > 0.f ? logit_cap * kLog2e : 0.f;
  p.sinks = sinks;
  p.tail_begin = max_splits > 1 ? tail_begin : -1;
  p.Dv = Dv;
  p.item_lens = item_lens;
  cudaStream_t st = (cudaStream_t)stream;
  if (kv_fp8) {
    if (D == 128) {
      return dtype == B200ATTN_BF16 ? launch_decode_fp8<128, true>(p, st) : launch_decode_fp8<128, false>(p, st);
    }
    return dtype == B200ATTN_BF16 ? launch_decode_fp8<64, true>(p, st) : launch_decode_fp8<64, false>(p, st);
  }
  if (D == 256) {
    return dtype == B200ATTN_BF16 ? launch_decode_nw<256, true, 2>(p, st) : launch_decode_nw<256, false, 2>(p, st);
  }
  if (D == 128) {
    return dtype == B200ATTN_BF16 ? launch_decode<128, true>(p, warps, st) : launch_decode<128, false>(p, warps, st);
  }
  return dtype == B200ATTN_BF16 ? launch_decode<64, true>(p, warps, st) : launch_decode<64, false>(p, warps, st);
}

extern "C" int b200attn_decode(const void* q, int64_t q_stride, const void* k_cache,
                               const void* v_cache, int64_t kv_stride, void* o, int64_t o_stride,
                               float* lse, const int32_t* kv_indptr, const int32_t* kv_indices,
                               const int32_t* q_rows, const int32_t* req_ids, int32_t B,
                               int32_t Hq, int32_t Hkv, int32_t D, int dtype, float sm_scale,
                               int32_t max_splits, int32_t min_chunk, int32_t tail_begin, int32_t warps,
                               int32_t window_left,
                               float logit_cap, const float* sinks, const int32_t* item_lens, void* workspace,
                               int64_t workspace_bytes, void* stream) {
  return decode_impl(q, q_stride, k_cache, v_cache, kv_stride, o, o_stride, lse, kv_indptr, kv_indices,
                     q_rows, req_ids, B, Hq, Hkv, D, dtype, sm_scale, max_splits, min_chunk, tail_begin,
                     window_left, logit_cap, sinks, workspace, workspace_bytes, stream, false, 1.f, 1.f, warps,
                     item_lens);
}

extern "C" int b200attn_decode_fp8kv(const void* q, int64_t q_stride, const void* k_cache,
                                     const void* v_cache, int64_t kv_stride, void* o,
                                     int64_t o_stride, float* lse, const int32_t* kv_indptr,
                                     const int32_t* kv_indices, const int32_t* q_rows,
                                     const int32_t* req_ids, int32_t B, int32_t Hq, int32_t Hkv,
                                     int32_t D, int dtype, float sm_scale, float k_descale,
                                     float v_descale, int32_t max_splits, int32_t min_chunk,
                                     int32_t tail_begin, int32_t window_left, float logit_cap,
                                     const float* sinks, void* workspace, int64_t workspace_bytes,
                                     void* stream) {
  return decode_impl(q, q_stride, k_cache, v_cache, kv_stride, o, o_stride, lse, kv_indptr, kv_indices,
                     q_rows, req_ids, B, Hq, Hkv, D, dtype, sm_scale, max_splits, min_chunk, tail_begin,
                     window_left, logit_cap, sinks, workspace, workspace_bytes, stream, true, k_descale,
                     v_descale);
}

// ---------------------------------------------------------------------------------------------------
// b200attn_decode_noalloc: b200attn_decode for callers that cannot allocate device memory (CUDA graph capture, tight
// serving loops, harnesses that forbid it).  No counterpart in the reference.
// The split-KV workspace (arrival counters + fp32 partials) is this file-scope device array instead of a
// caller buffer: kDecStaticWsBytes per GPU that loads the library, held until the process exits.
//   * In HOST code the array's name is only a stand-in for the symbol, not a device address: the address is asked
//     for once per device, with that device current, through cudaGetSymbolAddress and cached.
//   * Every call shares the array and finds the previous call's partials in it; the arrival counters of the units
//     of this call are cleared on the stream before the launch (they must start at zero).  Calls that OVERLAP on
//     the GPU (two streams, two graphs in flight) race on it: serialise them or use b200attn_decode.
//   * Nothing here calls cudaMalloc / cudaFree or any allocating routine.
namespace b200attn {
constexpr int64_t kDecStaticPartialBytes = 8ll << 20;  // fp32 partials + LSE: B * Hq * max_splits * (Dpad + 1) * 4 <= 8 MiB
constexpr int64_t kDecStaticWsBytes = kDecCounterBytes + kDecStaticPartialBytes;
static __device__ __align__(256) unsigned char g_dec_static_ws[kDecStaticWsBytes];

constexpr int kMaxDevices = 64;
static std::mutex g_static_ws_mu;
static void* g_static_ws_ptr[kMaxDevices] = {};  // device address of g_dec_static_ws on device i, 0 = not asked yet

static int static_ws_address(void** out) {
  int dev = 0;
  B2_CUDA(cudaGetDevice(&dev));
  B2_CHECK_ARG(dev >= 0 && dev < kMaxDevices, "decode_noalloc: device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lock(g_static_ws_mu);
  if (g_static_ws_ptr[dev] == nullptr) {
    void* ptr = nullptr;
    B2_CUDA(cudaGetSymbolAddress(&ptr, g_dec_static_ws));  // the current device's instance
    g_static_ws_ptr[dev] = ptr;
  }
  *out = g_static_ws_ptr[dev];
  return B200ATTN_OK;
}
}  // namespace b200attn

extern "C" int64_t b200attn_decode_noalloc_max_workspace_bytes(void) { return b200attn::kDecStaticWsBytes; }

extern "C" int b200attn_decode_noalloc(const void* q, int64_t q_stride, const void* k_cache, const void* v_cache,
                                       int64_t kv_stride, void* o, int64_t o_stride, float* lse,
                                       const int32_t* kv_indptr, const int32_t* kv_indices, const int32_t* q_rows,
                                       const int32_t* req_ids, int32_t B, int32_t Hq, int32_t Hkv, int32_t D,
                                       int dtype, float sm_scale, int32_t max_splits, int32_t min_chunk,
                                       int32_t tail_begin, int32_t warps, int32_t window_left, float logit_cap,
                                       const float* sinks, const int32_t* item_lens, void* stream) {
  using namespace b200attn;
  B2_CHECK_ARG(B >= 0 && Hq > 0 && Hkv > 0 && D > 0 && max_splits >= 1, "decode_noalloc: bad shape");
  const int64_t need = b200attn_decode_workspace_bytes(B, Hq, D, max_splits);
  B2_CHECK_ARG(need <= kDecStaticWsBytes,
               "decode_noalloc: this shape needs a %lld-byte split workspace, the built-in one holds %lld "
               "(B * Hq * max_splits * (head_dim padded to 64/128/256 + 1) * 4 <= %lld): lower max_splits or call "
               "b200attn_decode with a caller workspace",
               (long long)need, (long long)kDecStaticWsBytes, (long long)kDecStaticPartialBytes);
  void* ws = nullptr;
  if (need > 0) {
    const int rc = static_ws_address(&ws);
    if (rc != B200ATTN_OK) return rc;
    // the counters this launch will use start at zero whatever an earlier (possibly aborted) call left behind
    const int64_t units = (int64_t)B * Hkv * ((Hq / Hkv + kDecMaxG - 1) / kDecMaxG);
    B2_CHECK_ARG(units <= kDecMaxUnits, "decode_noalloc: more than %d units with max_splits > 1", kDecMaxUnits);
    B2_CUDA(cudaMemsetAsync(ws, 0, (size_t)units * sizeof(int), (cudaStream_t)stream));
  }
  return decode_impl(q, q_stride, k_cache, v_cache, kv_stride, o, o_stride, lse, kv_indptr, kv_indices, q_rows,
                     req_ids, B, Hq, Hkv, D, dtype, sm_scale, max_splits, min_chunk, tail_begin, window_left,
                     logit_cap, sinks, ws, need > 0 ? kDecStaticWsBytes : 0, stream, false, 1.f, 1.f, warps, item_lens);
}

#ifdef DEC_TRACE
extern "C" int b200attn_decode_trace_read(void* dst_host, int64_t bytes) {
  return cudaMemcpyFromSymbol(dst_host, b200attn::g_dec_trace, (size_t)bytes) == cudaSuccess ? 0 : 1;
}
#endif
