// (a4) Paged GQA decode, flat tile-stream ("stream-K") scheduling.
//
// Same arithmetic, shared-memory ring and mma path as decode_kernel (decode.cu); what changes is WHO computes
// WHAT.  decode_kernel launches one CTA per (request, kv head group, split): 512 units on 296 resident slots
// leave a 73 %-full last wave (cfg2), 128 units x 2 splits put two CTAs on 108 SMs and one on 40 (B=16 x 4K).
// Here the whole batch is one stream of 64-key tiles, unit-major:
//
//     g = HG * (tiles of the requests before i) + hg * tiles_i + t        (unit = (request i, kv head group hg))
//
// and CTA c of N (= 2 per SM) owns the contiguous range [c*T/N, (c+1)*T/N): every CTA moves the same number of
// KV bytes (+-1 tile) whatever the batch shape.  A CTA walks its range segment by segment (a segment = its tiles
// of one unit).  A segment covering a whole unit writes the output directly; a partial one writes an fp32
// partial into one of the CTA's TWO workspace slots (slot 1: the piece holding the unit's first tile, which is
// necessarily the CTA's last segment; slot 0: any other piece, necessarily its first segment) and takes a ticket
// on the unit's arrival counter; the CTA drawing the last ticket merges the pieces of CTAs c_first..c_last, both
// computed arithmetically from the partition.  Everything (tile counts, partition, piece counts) is a pure
// function of kv_indptr read on the device: one captured launch replays over any batch, no host split policy.
#include "decode.cuh"

namespace b200attn {

struct SkItem {
  int kv_beg, L, tiles;
};

__device__ __forceinline__ SkItem sk_item(const DecodeParams& p, int item) {
  const int req = p.req_ids ? p.req_ids[item] : item;
  int kv_beg = p.kv_indptr[req];
  int L = p.kv_indptr[req + 1] - kv_beg;
  const int lo = window_lo(L, p.window_left);
  SkItem it;
  it.kv_beg = kv_beg + lo;
  it.L = L - lo;
  it.tiles = max(1, (it.L + kDecTile - 1) / kDecTile);  // an empty request still owns one (fully masked) tile
  return it;
}

// CTA that owns global tile g under the partition start(c) = floor(c * T / N)
__device__ __forceinline__ int sk_owner(int64_t g, int64_t T, int N) {
  return (int)(((g + 1) * N - 1) / T);
}

// kFp8: e4m3 pool (rows of D bytes, the byte path of decode_fp8_kernel: permuted-k ldmatrix, hardware e4m3 -> f16,
// f16 mma, 4 stages x 3 CTAs per SM); otherwise the 16-bit path of decode_kernel.
template <int D, bool kBf16, bool kFp8>
__global__ void __launch_bounds__(kDecThreads, kFp8 ? B200ATTN_FP8_CTAS : 2)
    decode_sk_kernel(const DecodeParams p, const int min_tiles) {
  using T = typename std::conditional<kBf16, __nv_bfloat16, __half>::type;
  constexpr int kEsz = kFp8 ? 1 : 2;                 // pool bytes per element
  constexpr int kChunks = D * kEsz / 16;             // 16-byte chunks per row
  constexpr int kTileBytes = kDecTile * D * kEsz;
  constexpr int kStageBytes = 2 * kTileBytes;
  constexpr int kStages = kFp8 ? kDec8Stages : kDecStages;
  constexpr bool kMmaBf16 = kBf16 && !kFp8;          // the fp8 path multiplies in f16
  constexpr int kLoadsPerThread = kDecTile * kChunks / kDecThreads;
  constexpr int kRowsPerPass = kDecThreads / kChunks;
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ int64_t s_red[4];
  __shared__ int s_i0;
  __shared__ int64_t s_pre0;
  __shared__ int s_last;

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int HG = p.Hkv * p.groups_per_kvh;

  // ---- partition: total tiles, this CTA's range, the request its range starts in ------------------------
  int64_t mine = 0;
  for (int i = tid; i < p.B; i += kDecThreads) mine += sk_item(p, i).tiles;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) mine += __shfl_xor_sync(0xffffffffu, mine, o);
  if (lane == 0) s_red[warp] = mine;
  __syncthreads();
  const int64_t total = (s_red[0] + s_red[1] + s_red[2] + s_red[3]) * HG;
  // never cut finer than min_tiles per CTA: short batches use fewer CTAs
  const int N = (int)max((int64_t)1, min((int64_t)gridDim.x, total / min_tiles));
  const int c = blockIdx.x;
  if (c >= N) return;
  const int64_t g0 = (int64_t)c * total / N;
  const int64_t g1 = (int64_t)(c + 1) * total / N;
  if (g0 >= g1) return;
  {
    // chunked block scan over the requests until the one holding tile g0 is found
    int64_t base = 0;  // tiles of the requests before this chunk (same value in every thread)
    for (int i0 = 0; i0 < p.B; i0 += kDecThreads) {
      const int i = i0 + tid;
      const int64_t t = i < p.B ? sk_item(p, i).tiles : 0;
      int64_t incl = t;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int64_t y = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += y;
      }
      __syncthreads();  // s_red reuse
      if (lane == 31) s_red[warp] = incl;
      __syncthreads();
      int64_t woff = 0;
      for (int w = 0; w < warp; ++w) woff += s_red[w];
      const int64_t excl = base + woff + incl - t;
      if (i < p.B && excl * HG <= g0 && g0 < (excl + t) * HG) {
        s_i0 = i;
        s_pre0 = excl;
      }
      const int64_t chunk_total = s_red[0] + s_red[1] + s_red[2] + s_red[3];
      base += chunk_total;
      if (base * HG > g0) break;  // uniform: found in this chunk
    }
    __syncthreads();
  }
  int item = s_i0;
  int64_t pre = s_pre0;  // tiles of the requests before `item`
  SkItem it = sk_item(p, item);

  const int ld_chunk = tid % kChunks;
  const int ld_row0 = tid / kChunks;
  const int64_t row_bytes = p.kv_stride * kEsz;
  const uint32_t smem_base = smem_u32(smem);
  const int lm = lane >> 3, lr = lane & 7;
  float* sm_o = reinterpret_cast<float*>(smem);  // [4][16][D]   (aliases the ring between segments)
  float* sm_m = sm_o + 4 * 16 * D;               // [4][16]
  float* sm_l = sm_m + 64;                       // [4][16]

  int64_t cur = g0;
  while (cur < g1) {
    // ---- segment: tiles [t_lo, t_hi) of unit (item, hg) -------------------------------------------------
    const int64_t ubase = pre * HG;
    const int hg = (int)((cur - ubase) / it.tiles);
    const int64_t us = ubase + (int64_t)hg * it.tiles;  // unit's first global tile
    const int64_t ue = us + it.tiles;
    const int t_lo = (int)(cur - us);
    const int64_t seg_end = min(g1, ue);
    const int t_hi = t_lo + (int)(seg_end - cur);
    const bool whole = (t_lo == 0 && t_hi == it.tiles);
    const int kvh = hg / p.groups_per_kvh;
    const int grp = hg - kvh * p.groups_per_kvh;
    const int qrow = p.q_rows ? p.q_rows[item] : item;
    const int h0 = kvh * p.G + grp * kDecMaxG;
    const int gcount = min(kDecMaxG, p.G - grp * kDecMaxG);
    const int k_lo = t_lo * kDecTile;
    const int k_hi = min(it.L, t_hi * kDecTile);
    const int ntiles = t_hi - t_lo;

    uint32_t qf[D / 16][4];
    {
      const T* qbase = reinterpret_cast<const T*>(p.q) + (int64_t)qrow * p.q_stride + (int64_t)h0 * D;
      const int r0 = lane >> 2, c0 = (lane & 3) * (kFp8 ? 4 : 2);
#pragma unroll
      for (int ks = 0; ks < D / 16; ++ks) {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int r = r0 + (j & 1) * 8;
          const int cc = ks * 16 + c0 + (j >> 1) * (kFp8 ? 2 : 8);  // fp8: contraction index permuted (decode.cu)
          uint32_t w = (r < gcount) ? *reinterpret_cast<const uint32_t*>(qbase + (int64_t)r * D + cc) : 0u;
          if constexpr (kFp8 && kBf16) {
            const float2 f = unpack2<true>(w);
            w = pack2<false>(f.x, f.y);
          }
          qf[ks][j] = w;
        }
      }
    }
    const uint8_t* kc = reinterpret_cast<const uint8_t*>(p.k_cache) + (int64_t)kvh * D * kEsz + ld_chunk * 16;
    const uint8_t* vc = reinterpret_cast<const uint8_t*>(p.v_cache) + (int64_t)kvh * D * kEsz + ld_chunk * 16;
    const int32_t* idx = p.kv_indices + it.kv_beg;

    int32_t slots[kLoadsPerThread];
    auto fetch_slots = [&](int tile) {
      const int base = k_lo + tile * kDecTile + ld_row0;
#pragma unroll
      for (int i = 0; i < kLoadsPerThread; ++i) {
        const int key = base + i * kRowsPerPass;
        slots[i] = (tile < ntiles && key < k_hi) ? __ldg(idx + key) : -1;
      }
    };
    auto issue_tile = [&](int stage) {
      const uint32_t ks = smem_base + stage * kStageBytes;
      const uint32_t vs = ks + kTileBytes;
#pragma unroll
      for (int i = 0; i < kLoadsPerThread; ++i) {
        const int row = ld_row0 + i * kRowsPerPass;
        const uint32_t off = kFp8 ? swz8_off<D>(row, ld_chunk) : swz_off<D>(row, ld_chunk);
        const bool ok = slots[i] >= 0;
        const int64_t g = ok ? (int64_t)slots[i] * row_bytes : 0;
        cp_async16(ks + off, kc + g, ok ? 16 : 0);
        cp_async16(vs + off, vc + g, ok ? 16 : 0);
      }
    };

    fetch_slots(0);
#pragma unroll
    for (int s = 0; s < kStages - 1; ++s) {
      if (s < ntiles) issue_tile(s);
      cp_async_commit();
      fetch_slots(s + 1);
    }

    float o_acc[D / 8][4];
#pragma unroll
    for (int n = 0; n < D / 8; ++n) o_acc[n][0] = o_acc[n][1] = o_acc[n][2] = o_acc[n][3] = 0.f;
    float m_run[2] = {-INFINITY, -INFINITY};
    float l_run[2] = {0.f, 0.f};

    for (int t = 0; t < ntiles; ++t) {
      cp_async_wait<kStages - 2>();
      __syncthreads();
      {
        const int nt = t + kStages - 1;
        if (nt < ntiles) issue_tile(nt % kStages);
        cp_async_commit();
        fetch_slots(nt + 1);
      }
      const uint32_t ks = smem_base + (t % kStages) * kStageBytes;
      const uint32_t vs = ks + kTileBytes;
      const int key0 = k_lo + t * kDecTile + warp * 16;
      if (key0 < k_hi) {
        float s[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
        {
          const int krow = warp * 16 + (lm >> 1) * 8 + lr;
          if constexpr (kFp8) {
#pragma unroll
            for (int kp = 0; kp < D / 32; ++kp) {
              uint32_t m0, m1, m2, m3, a, b;
              ldmatrix_x4(m0, m1, m2, m3, ks + swz8_off<D>(krow, 2 * kp + (lm & 1)));
              e4m3x4_pairs(m0, a, b);
              mma_16816<false>(s[0], qf[2 * kp], a, b);
              e4m3x4_pairs(m1, a, b);
              mma_16816<false>(s[0], qf[2 * kp + 1], a, b);
              e4m3x4_pairs(m2, a, b);
              mma_16816<false>(s[1], qf[2 * kp], a, b);
              e4m3x4_pairs(m3, a, b);
              mma_16816<false>(s[1], qf[2 * kp + 1], a, b);
            }
          } else {
#pragma unroll
            for (int ksp = 0; ksp < D / 16; ++ksp) {
              uint32_t b0, b1, b2, b3;
              ldmatrix_x4(b0, b1, b2, b3, ks + swz_off<D>(krow, 2 * ksp + (lm & 1)));
              mma_16816<kMmaBf16>(s[0], qf[ksp], b0, b1);
              mma_16816<kMmaBf16>(s[1], qf[ksp], b2, b3);
            }
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
          pf[nb * 2 + 0] = pack2<kMmaBf16>(p0, p1);
          pf[nb * 2 + 1] = pack2<kMmaBf16>(p2, p3);
        }
        const int vrow = warp * 16 + (lm & 1) * 8 + lr;
        if constexpr (kFp8) {
          const bool rescale = __any_sync(0xffffffffu, alpha[0] != 1.f || alpha[1] != 1.f);
#pragma unroll
          for (int np = 0; np < D / 32; ++np) {
            uint32_t m0, m1, m2, m3;
            ldmatrix_x4_trans(m0, m1, m2, m3, vs + swz8_off<D>(vrow, 2 * np + (lm >> 1)));
            if (rescale) {
#pragma unroll
              for (int h = 0; h < 4; ++h) {
                float(&acc)[4] = o_acc[4 * np + h];
                acc[0] *= alpha[0];
                acc[1] *= alpha[0];
                acc[2] *= alpha[1];
                acc[3] *= alpha[1];
              }
            }
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
        } else {
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
            mma_16816<kMmaBf16>(o_acc[2 * np], pf, b0, b1);
            mma_16816<kMmaBf16>(o_acc[2 * np + 1], pf, b2, b3);
          }
        }
      }
    }
    cp_async_wait<0>();
    __syncthreads();  // ring is dead until the next segment's prologue

    // ---- merge the four warps ------------------------------------------------------------------------------
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      l_run[r] += __shfl_xor_sync(0xffffffffu, l_run[r], 1);
      l_run[r] += __shfl_xor_sync(0xffffffffu, l_run[r], 2);
    }
    {
      const int r0 = lane >> 2;
      if constexpr (kFp8) {  // a lane owns 4 consecutive dims per 16-d chunk (decode_fp8_kernel's layout)
        const int c0 = (lane & 3) * 4;
#pragma unroll
        for (int cch = 0; cch < D / 16; ++cch) {
          float* dst = sm_o + (warp * 16 + r0) * D + cch * 16 + c0;
          *reinterpret_cast<float4*>(dst) =
              make_float4(o_acc[2 * cch][0], o_acc[2 * cch + 1][0], o_acc[2 * cch][1], o_acc[2 * cch + 1][1]);
          *reinterpret_cast<float4*>(dst + 8 * D) =
              make_float4(o_acc[2 * cch][2], o_acc[2 * cch + 1][2], o_acc[2 * cch][3], o_acc[2 * cch + 1][3]);
        }
      } else {
        const int c0 = (lane & 3) * 2;
#pragma unroll
        for (int n = 0; n < D / 8; ++n) {
          float* dst = sm_o + (warp * 16 + r0) * D + n * 8 + c0;
          *reinterpret_cast<float2*>(dst) = make_float2(o_acc[n][0], o_acc[n][1]);
          *reinterpret_cast<float2*>(dst + 8 * D) = make_float2(o_acc[n][2], o_acc[n][3]);
        }
      }
      if ((lane & 3) == 0) {
        sm_m[warp * 16 + r0] = m_run[0];
        sm_m[warp * 16 + r0 + 8] = m_run[1];
        sm_l[warp * 16 + r0] = l_run[0];
        sm_l[warp * 16 + r0 + 8] = l_run[1];
      }
    }
    __syncthreads();
    const int wslot = c * 2 + (t_lo == 0 ? 1 : 0);  // partial pieces only
    for (int e = tid; e < gcount * (D / 4); e += kDecThreads) {
      const int r = e / (D / 4), cc = (e - r * (D / 4)) * 4;
      float m = -INFINITY;
#pragma unroll
      for (int w = 0; w < 4; ++w) m = fmaxf(m, sm_m[w * 16 + r]);
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
      float l = 0.f;
#pragma unroll
      for (int w = 0; w < 4; ++w) {
        const float mw = sm_m[w * 16 + r];
        const float f = (mw == -INFINITY) ? 0.f : fast_exp2(mw - m);
        const float4 x = *reinterpret_cast<const float4*>(sm_o + (w * 16 + r) * D + cc);
        acc.x += f * x.x;
        acc.y += f * x.y;
        acc.z += f * x.z;
        acc.w += f * x.w;
        l += f * sm_l[w * 16 + r];
      }
      if (whole && p.sinks != nullptr && l > 0.f) l += fast_exp2(__ldg(p.sinks + h0 + r) * kLog2e - m);
      const float inv = (l > 0.f) ? p.out_scale / l : 0.f;  // out_scale: v_descale for e4m3 pools, else 1
      acc.x *= inv;
      acc.y *= inv;
      acc.z *= inv;
      acc.w *= inv;
      const float lse2 = (l > 0.f) ? m + log2f(l) : -INFINITY;
      const int h = h0 + r;
      if (whole) {
        T* dst = reinterpret_cast<T*>(p.o) + (int64_t)qrow * p.o_stride + (int64_t)h * D + cc;
        uint2 pk;
        pk.x = pack2<kBf16>(acc.x, acc.y);
        pk.y = pack2<kBf16>(acc.z, acc.w);
        *reinterpret_cast<uint2*>(dst) = pk;
        if (p.lse != nullptr && cc == 0) p.lse[(int64_t)qrow * p.Hq + h] = lse2 * kLn2;
      } else {
        const int64_t w = (int64_t)wslot * kDecMaxG + r;
        *reinterpret_cast<float4*>(p.ws_o + w * D + cc) = acc;
        if (cc == 0) p.ws_lse[w] = lse2;
      }
    }
    if (!whole) {
      // ---- arrival ticket; the last piece of the unit to arrive merges all of them -------------------------
      const int c_first = sk_owner(us, total, N);
      const int c_last = sk_owner(ue - 1, total, N);
      __threadfence();
      __syncthreads();
      if (tid == 0) {
        int* cnt = p.ws_cnt + (int64_t)item * HG + hg;
        const int ticket = atomicAdd(cnt, 1);
        s_last = (ticket == c_last - c_first);
        if (s_last) *cnt = 0;  // self-cleaning
      }
      __syncthreads();
      if (s_last) {
        __threadfence();
        for (int e = tid; e < gcount * (D / 4); e += kDecThreads) {
          const int r = e / (D / 4), cc = (e - r * (D / 4)) * 4;
          const int h = h0 + r;
          float m = -INFINITY;
          for (int pc = c_first; pc <= c_last; ++pc) {
            const int64_t w = (int64_t)(pc * 2 + (pc == c_first ? 1 : 0)) * kDecMaxG + r;
            m = fmaxf(m, __ldcg(p.ws_lse + w));
          }
          float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
          float l = 0.f;
          for (int pc = c_first; pc <= c_last; ++pc) {
            const int64_t w = (int64_t)(pc * 2 + (pc == c_first ? 1 : 0)) * kDecMaxG + r;
            const float ls = __ldcg(p.ws_lse + w);
            const float f = (ls == -INFINITY) ? 0.f : fast_exp2(ls - m);
            const float4 x = __ldcg(reinterpret_cast<const float4*>(p.ws_o + w * D + cc));
            acc.x += f * x.x;
            acc.y += f * x.y;
            acc.z += f * x.z;
            acc.w += f * x.w;
            l += f;
          }
          if (p.sinks != nullptr && l > 0.f) l += fast_exp2(__ldg(p.sinks + h) * kLog2e - m);
          const float inv = (l > 0.f) ? 1.f / l : 0.f;
          T* dst = reinterpret_cast<T*>(p.o) + (int64_t)qrow * p.o_stride + (int64_t)h * D + cc;
          uint2 pk;
          pk.x = pack2<kBf16>(acc.x * inv, acc.y * inv);
          pk.y = pack2<kBf16>(acc.z * inv, acc.w * inv);
          *reinterpret_cast<uint2*>(dst) = pk;
          if (p.lse != nullptr && cc == 0)
            p.lse[(int64_t)qrow * p.Hq + h] = (l > 0.f) ? (m + log2f(l)) * kLn2 : -INFINITY;
        }
      }
    }
    __syncthreads();  // sm_o / sm_m / s_last are reused by the next segment
    // ---- advance ---------------------------------------------------------------------------------------------
    cur = seg_end;
    if (cur == ue && hg == HG - 1 && cur < g1) {
      pre += it.tiles;
      ++item;
      it = sk_item(p, item);
    }
  }
}

template <int D, bool kBf16, bool kFp8>
static int launch_decode_sk(const DecodeParams& p, int ctas, int min_tiles, cudaStream_t st) {
  constexpr int kRing = (kFp8 ? kDec8Stages : kDecStages) * 2 * kDecTile * D * (kFp8 ? 1 : 2);
  constexpr int kMerge = (4 * 16 * D + 128) * 4;
  constexpr int kSmem = kRing > kMerge ? kRing : kMerge;
  static bool configured = false;
  if (!configured) {
    B2_CUDA(cudaFuncSetAttribute(decode_sk_kernel<D, kBf16, kFp8>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem));
    B2_CUDA(cudaFuncSetAttribute(decode_sk_kernel<D, kBf16, kFp8>, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
    configured = true;
  }
  decode_sk_kernel<D, kBf16, kFp8><<<ctas, kDecThreads, kSmem, st>>>(p, min_tiles);
  B2_LAUNCH_CHECK();
  return B200ATTN_OK;
}

static int sk_ctas(bool fp8) { return (fp8 ? B200ATTN_FP8_CTAS : 2) * sm_count(); }
static int sk_ctas_cap() { return sk_ctas(true) > sk_ctas(false) ? sk_ctas(true) : sk_ctas(false); }  // workspace layout

}  // namespace b200attn

extern "C" int64_t b200attn_decode_sk_workspace_bytes(int32_t B, int32_t Hq, int32_t Hkv, int32_t D) {
  using namespace b200attn;
  if (B <= 0 || Hq <= 0 || Hkv <= 0) return 0;
  const int groups = (Hq / Hkv + kDecMaxG - 1) / kDecMaxG;
  // two partial slots per CTA ([16][D] fp32 + [16] lse) + one arrival counter per unit
  (void)groups;
  return kDecCounterBytes + (int64_t)sk_ctas_cap() * 2 * kDecMaxG * (D + 1) * 4;  // fixed counter header first
}

static int decode_sk_impl(const void* q, int64_t q_stride, const void* k_cache, const void* v_cache,
                          int64_t kv_stride, void* o, int64_t o_stride, float* lse,
                          const int32_t* kv_indptr, const int32_t* kv_indices, const int32_t* q_rows,
                          const int32_t* req_ids, int32_t B, int32_t Hq, int32_t Hkv, int32_t D,
                          int dtype, float sm_scale, int32_t min_chunk, int32_t window_left,
                          float logit_cap, const float* sinks, void* workspace,
                          int64_t workspace_bytes, void* stream, bool fp8, float k_descale, float v_descale) {
  using namespace b200attn;
  B2_CHECK_ARG(B >= 0, "decode_sk: negative batch");
  B2_CHECK_ARG(logit_cap >= 0.f, "decode_sk: negative logit_cap");
  if (B == 0) return B200ATTN_OK;
  B2_CHECK_ARG(q && k_cache && v_cache && o && kv_indptr && kv_indices, "decode_sk: null pointer");
  B2_CHECK_ARG(Hq > 0 && Hkv > 0 && Hq % Hkv == 0, "decode_sk: Hq=%d must be a multiple of Hkv=%d", Hq, Hkv);
  B2_CHECK_ARG(min_chunk >= 1, "decode_sk: bad min_chunk");
  B2_CHECK_ARG(q_stride % 8 == 0 && kv_stride % (fp8 ? 16 : 8) == 0 && o_stride % 4 == 0,
               "decode_sk: strides must keep rows 16-byte aligned");
  B2_CHECK_ARG(!fp8 || (k_descale > 0.f && v_descale > 0.f), "decode_sk: fp8 descales must be positive");
  B2_CHECK_ARG((((uintptr_t)q | (uintptr_t)k_cache | (uintptr_t)v_cache | (uintptr_t)o) & 15) == 0,
               "decode_sk: tensors must be 16-byte aligned");
  if (dtype != B200ATTN_BF16 && dtype != B200ATTN_FP16) B2_UNSUPPORTED("decode_sk: dtype %d", dtype);
  if (D != 64 && D != 128) B2_UNSUPPORTED("decode_sk: head_dim %d (supported: 64, 128)", D);
  const int64_t need = b200attn_decode_sk_workspace_bytes(B, Hq, Hkv, D);
  B2_CHECK_ARG(workspace && workspace_bytes >= need, "decode_sk: workspace too small (%lld < %lld)",
               (long long)workspace_bytes, (long long)need);
  const int ctas = sk_ctas(fp8);
  const int cap = sk_ctas_cap();
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
  B2_CHECK_ARG((int64_t)B * Hkv * ((Hq / Hkv + kDecMaxG - 1) / kDecMaxG) <= kDecMaxUnits,
               "decode_sk: more than %d units", kDecMaxUnits);
  p.ws_cnt = reinterpret_cast<int*>(workspace);
  p.ws_o = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(workspace) + kDecCounterBytes);
  p.ws_lse = p.ws_o + (int64_t)cap * 2 * kDecMaxG * D;
  p.q_stride = q_stride;
  p.kv_stride = kv_stride;
  p.o_stride = o_stride;
  p.B = B;
  p.Hq = Hq;
  p.Hkv = Hkv;
  p.G = Hq / Hkv;
  p.groups_per_kvh = (p.G + kDecMaxG - 1) / kDecMaxG;
  p.max_splits = 1;
  p.min_chunk = min_chunk;
  if (fp8) sm_scale *= k_descale * kFp8Bias;
  p.scale_log2 = sm_scale * kLog2e;
  p.window_left = window_left;
  p.cap_pre = logit_cap > 0.f ? sm_scale / logit_cap : 0.f;
  p.cap_post = logit_cap > 0.f ? logit_cap * kLog2e : 0.f;
  p.sinks = sinks;
  p.tail_begin = -1;
  p.out_scale = fp8 ? v_descale * kFp8Bias : 1.f;
  const int min_tiles = (min_chunk + kDecTile - 1) / kDecTile;
  cudaStream_t st = (cudaStream_t)stream;
#define B2_SK(DD, FP8)                                                                          \
  return dtype == B200ATTN_BF16 ? launch_decode_sk<DD, true, FP8>(p, ctas, min_tiles, st)      \
                                : launch_decode_sk<DD, false, FP8>(p, ctas, min_tiles, st)
  if (fp8) {
    if (D == 128) B2_SK(128, true);
    B2_SK(64, true);
  }
  if (D == 128) B2_SK(128, false);
  B2_SK(64, false);
#undef B2_SK
}

extern "C" int b200attn_decode_sk(const void* q, int64_t q_stride, const void* k_cache, const void* v_cache,
                                  int64_t kv_stride, void* o, int64_t o_stride, float* lse,
                                  const int32_t* kv_indptr, const int32_t* kv_indices, const int32_t* q_rows,
                                  const int32_t* req_ids, int32_t B, int32_t Hq, int32_t Hkv, int32_t D,
                                  int dtype, float sm_scale, int32_t min_chunk, int32_t window_left,
                                  float logit_cap, const float* sinks, void* workspace,
                                  int64_t workspace_bytes, void* stream) {
  return decode_sk_impl(q, q_stride, k_cache, v_cache, kv_stride, o, o_stride, lse, kv_indptr, kv_indices, q_rows,
                        req_ids, B, Hq, Hkv, D, dtype, sm_scale, min_chunk, window_left, logit_cap, sinks, workspace,
                        workspace_bytes, stream, false, 1.f, 1.f);
}

extern "C" int b200attn_decode_sk_fp8kv(const void* q, int64_t q_stride, const void* k_cache, const void* v_cache,
                                        int64_t kv_stride, void* o, int64_t o_stride, float* lse,
                                        const int32_t* kv_indptr, const int32_t* kv_indices,
                                        const int32_t* q_rows, const int32_t* req_ids, int32_t B, int32_t Hq,
                                        int32_t Hkv, int32_t D, int dtype, float sm_scale, float k_descale,
                                        float v_descale, int32_t min_chunk, int32_t window_left, float logit_cap,
                                        const float* sinks, void* workspace, int64_t workspace_bytes,
                                        void* stream) {
  return decode_sk_impl(q, q_stride, k_cache, v_cache, kv_stride, o, o_stride, lse, kv_indptr, kv_indices, q_rows,
                        req_ids, B, Hq, Hkv, D, dtype, sm_scale, min_chunk, window_left, logit_cap, sinks, workspace,
                        workspace_bytes, stream, true, k_descale, v_descale);
}
