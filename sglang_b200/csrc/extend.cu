// (a5) Ragged extend / prefill attention on tcgen05 tensor cores (sm_100a).
//
// Semantics follow the oracle python/sglang/srt/layers/attention/torch_native_backend.py:61-174
// (prefix + new tokens all read from the pool through req_to_token; query j of a request
// with L keys and E query rows sits at absolute position L-E+j; causal or not).
//
// One CTA = one 128-row query tile of one (request, q head):
//   warp 4/5    TMA producers (4: Q + K tiles, 5: V tiles).  Per 128-key step the warp reads the
//               128 slot ids and checks on the fly that every aligned 16-key group sits in 16
//               consecutive pool slots (true under the paged allocator for page_size >= 16):
//                 yes -> 16 tiled TMA boxes (16 rows x 64 cols) from one elected lane;
//                 no / ragged last step -> cp.async.bulk.tensor tile::gather4 (4 pool rows per
//                 instruction, OOB rows zero-filled): any page size, no contiguity assumption.
//               Both land in the same 128B-swizzled K-major layout.
//   warp 6      MMA issuer (one elected lane): S = Q K^T as 8 x tcgen05.mma M128 N128 K16 into a
//               double-buffered TMEM accumulator, O += P V as 8 x tcgen05.mma (V is the MN-major
//               B operand straight from the gathered tile).  QK^T of step j+1 is issued before
//               P V of step j so the tensor pipe works under the softmax.
//   warps 0-3   softmax: thread = query row (TMEM lane); tcgen05.ld S, causal / ragged masking,
//               exp2 online softmax with lazy rescaling (O in TMEM is only rescaled when the row
//               max grew by > 2^8), P written as bf16 into a 128B-swizzled smem tile for the
//               second GEMM; epilogue O / l -> global.
// TMEM: S0 [0,128) S1 [128,256) O [256,384) fp32 columns.
// Algorithmic FLOPs per (request, q head): 4 * D * sum_j (#visible keys of row j).
#include "sm100.cuh"

namespace b200attn {

constexpr int kExtM = 128;        // query rows per CTA
constexpr int kExtN = 128;        // keys per step
constexpr int kExtStages = 2;     // K and V ring depth
constexpr int kExtThreads = 224;  // 4 softmax warps + 2 producers + MMA
constexpr int kExtBoxRows = 16;   // rows of one tiled K/V TMA box
constexpr float kRescaleThreshold = 8.f;  // log2 domain

struct ExtendParams {
  void* o;
  float* lse;
  const int32_t* qo_indptr;
  const int32_t* kv_indptr;
  const int32_t* kv_indices;
  int64_t o_stride;
  int B, Hq, Hkv, G, num_slots;
  int causal;
  float scale_log2;
};

// shared memory carve-up (bytes, from a 1024-aligned base)
template <int D>
struct ExtSmem {
  static constexpr int kTile = kExtN * D * 2;  // one 128 x D operand tile (bf16)
  static constexpr int kHalf = kExtN * 128;    // one 64-column half: 128 rows x 128 B
  static constexpr int kQ = 0;
  static constexpr int kK = kQ + kTile;
  static constexpr int kV = kK + kExtStages * kTile;
  static constexpr int kP = kV + kExtStages * kTile;
  static constexpr int kBar = kP + kExtM * kExtN * 2;
  static constexpr int kTotal = kBar + 256;
};

enum ExtBar {
  kBarQ = 0,
  kBarKFull = 1,                       // [stages]
  kBarKEmpty = kBarKFull + kExtStages,
  kBarVFull = kBarKEmpty + kExtStages,
  kBarVEmpty = kBarVFull + kExtStages,
  kBarSFull = kBarVEmpty + kExtStages,  // [2]
  kBarSEmpty = kBarSFull + 2,           // [2]
  kBarPFull = kBarSEmpty + 2,
  kBarPVDone = kBarPFull + 1,
  kBarCount = kBarPVDone + 1
};

template <int D, bool kBf16>
__global__ void __launch_bounds__(kExtThreads, 1)
    extend_kernel(const __grid_constant__ CUtensorMap tm_q, const __grid_constant__ CUtensorMap tm_k,
                  const __grid_constant__ CUtensorMap tm_v, const __grid_constant__ CUtensorMap tm_kbox,
                  const __grid_constant__ CUtensorMap tm_vbox, const ExtendParams p) {
  using namespace sm100;
  using S = ExtSmem<D>;
  constexpr int kHalves = D / 64;  // 64-column halves of the head dim
  constexpr uint32_t kTmemCols = 512;
  constexpr uint32_t kColS = 0, kColO = 256;

  const int h = blockIdx.x;
  const int b = blockIdx.z;
  const int q_beg = p.qo_indptr[b];
  const int E = p.qo_indptr[b + 1] - q_beg;
  const int num_qtiles = (E + kExtM - 1) / kExtM;
  if ((int)blockIdx.y >= num_qtiles) return;
  const int qt = num_qtiles - 1 - blockIdx.y;  // heavy (late) causal tiles first
  const int m0 = qt * kExtM;
  const int rows_valid = min(kExtM, E - m0);
  const int kv_beg = p.kv_indptr[b];
  const int L = p.kv_indptr[b + 1] - kv_beg;
  const int prefix = L - E;
  const int kv_end = p.causal ? min(L, prefix + m0 + rows_valid) : L;
  const int ntiles = (kv_end + kExtN - 1) / kExtN;
  const int kvh = h / p.G;

  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem0 = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem = smem_raw + (smem0 - smem_u32(smem_raw));
  const uint32_t bar0 = smem0 + S::kBar;
  auto bar = [&](int i) { return bar0 + 8u * i; };
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + S::kBar + 8 * kBarCount);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    mbar_init(bar(kBarQ), 1);
    for (int s = 0; s < kExtStages; ++s) {
      mbar_init(bar(kBarKFull + s), 1);
      mbar_init(bar(kBarKEmpty + s), 1);
      mbar_init(bar(kBarVFull + s), 1);
      mbar_init(bar(kBarVEmpty + s), 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(bar(kBarSFull + s), 1);
      mbar_init(bar(kBarSEmpty + s), 128);
    }
    mbar_init(bar(kBarPFull), 128);
    mbar_init(bar(kBarPVDone), 1);
    mbar_fence_init();
  }
  if (warp == 6) tmem_alloc<kTmemCols>(smem_u32(tmem_slot));
  if (warp == 4 && lane == 0) {
    tma_prefetch_desc(&tm_q);
    tma_prefetch_desc(&tm_k);
    tma_prefetch_desc(&tm_kbox);
  }
  if (warp == 5 && lane == 0) {
    tma_prefetch_desc(&tm_v);
    tma_prefetch_desc(&tm_vbox);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  if (warp == 4 || warp == 5) {
    // ===================== TMA producers (warp 4: Q + K, warp 5: V) =====================
    const bool is_k = (warp == 4);
    if (is_k && lane == 0) {
      mbar_arrive_expect_tx(bar(kBarQ), S::kTile);
#pragma unroll
      for (int hf = 0; hf < kHalves; ++hf)
        tma_load_2d(smem0 + S::kQ + hf * S::kHalf, &tm_q, bar(kBarQ), h * D + hf * 64, q_beg + m0);
    }
    const int32_t* idx = p.kv_indices + kv_beg;
    const int col = kvh * D;
    const CUtensorMap* tm_g = is_k ? &tm_k : &tm_v;
    const CUtensorMap* tm_b = is_k ? &tm_kbox : &tm_vbox;
    const uint32_t tile0 = smem0 + (is_k ? S::kK : S::kV);
    const int bar_full = is_k ? kBarKFull : kBarVFull, bar_empty = is_k ? kBarKEmpty : kBarVEmpty;
    int r[4];
    auto load_slots = [&](int j) {
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int key = j * kExtN + lane * 4 + i;
        r[i] = (j < ntiles && key < kv_end) ? __ldg(idx + key) : p.num_slots;  // OOB row -> zero fill
      }
    };
    load_slots(0);
    for (int j = 0; j < ntiles; ++j) {
      const int st = j % kExtStages;
      const uint32_t dst = tile0 + st * S::kTile;
      // every aligned 16-key group in 16 consecutive slots?  (4 lanes x 4 keys per group)
      const int g0 = __shfl_sync(0xffffffffu, r[0], lane & ~3);
      bool contig = (r[1] == r[0] + 1) && (r[2] == r[0] + 2) && (r[3] == r[0] + 3) &&
                    (r[0] == g0 + 4 * (lane & 3));
      contig = __all_sync(0xffffffffu, contig) && ((j + 1) * kExtN <= kv_end);
      if (j >= kExtStages) mbar_wait(bar(bar_empty + st), ((j / kExtStages) + 1) & 1);
      if (lane == 0) mbar_arrive_expect_tx(bar(bar_full + st), S::kTile);
      __syncwarp();
      if (contig) {
        int rows[kExtN / kExtBoxRows];
#pragma unroll
        for (int g = 0; g < kExtN / kExtBoxRows; ++g) rows[g] = __shfl_sync(0xffffffffu, r[0], g * 4);
        if (lane == 0) {
#pragma unroll
          for (int g = 0; g < kExtN / kExtBoxRows; ++g)
#pragma unroll
            for (int hf = 0; hf < kHalves; ++hf)
              tma_load_2d(dst + hf * S::kHalf + g * kExtBoxRows * 128, tm_b, bar(bar_full + st),
                          col + hf * 64, rows[g]);
        }
      } else {
#pragma unroll
        for (int hf = 0; hf < kHalves; ++hf)
          tma_gather4(dst + hf * S::kHalf + lane * 4 * 128, tm_g, bar(bar_full + st), col + hf * 64,
                      r[0], r[1], r[2], r[3]);
      }
      load_slots(j + 1);
    }
  } else if (warp == 6) {
    // ===================== MMA issuer =====================
    constexpr uint32_t idesc_qk = idesc_f16(kBf16 ? 1 : 0, kExtM, kExtN, 0, 0);
    constexpr uint32_t idesc_pv = idesc_f16(kBf16 ? 1 : 0, kExtM, D, 0, 1);
    mbar_wait(bar(kBarQ), 0);
    tc_fence_after();
    auto issue_pv = [&](int j) {
      const int st = j % kExtStages;
      mbar_wait(bar(kBarPFull), j & 1);
      mbar_wait(bar(kBarVFull + st), (j / kExtStages) & 1);
      tc_fence_after();
      if (lane == 0) {
#pragma unroll
        for (int ks = 0; ks < kExtN / 16; ++ks) {
          // A = P (K-major, keys): 64-key halves of 128 rows x 128 B
          const uint64_t a = smem_desc_sw128(
              smem0 + S::kP + (ks / 4) * (kExtM * 128) + (ks % 4) * 32, 16, 1024);
          // B = V (MN-major): 16 keys per step = 2 KB down the tile; d-halves 16 KB apart
          const uint64_t bd = smem_desc_sw128(smem0 + S::kV + st * S::kTile + ks * 2048,
                                              S::kHalf, 1024);
          umma_ss(tmem + kColO, a, bd, idesc_pv, (j > 0 || ks > 0) ? 1u : 0u);
        }
        umma_commit(bar(kBarVEmpty + st));
        umma_commit(bar(kBarPVDone));
      }
      __syncwarp();
    };
    for (int j = 0; j < ntiles; ++j) {
      const int st = j % kExtStages, sb = j & 1;
      mbar_wait(bar(kBarKFull + st), (j / kExtStages) & 1);
      if (j >= 2) mbar_wait(bar(kBarSEmpty + sb), ((j / 2) + 1) & 1);
      tc_fence_after();
      if (lane == 0) {
#pragma unroll
        for (int ks = 0; ks < D / 16; ++ks) {
          const uint32_t off = (ks / 4) * S::kHalf + (ks % 4) * 32;
          const uint64_t a = smem_desc_sw128(smem0 + S::kQ + off, 16, 1024);
          const uint64_t bd = smem_desc_sw128(smem0 + S::kK + st * S::kTile + off, 16, 1024);
          umma_ss(tmem + kColS + sb * kExtN, a, bd, idesc_qk, ks > 0 ? 1u : 0u);
        }
        umma_commit(bar(kBarKEmpty + st));
        umma_commit(bar(kBarSFull + sb));
      }
      __syncwarp();
      if (j > 0) issue_pv(j - 1);
    }
    issue_pv(ntiles - 1);
  } else {
    // ===================== softmax + epilogue (thread = query row) =====================
    using T = typename std::conditional<kBf16, __nv_bfloat16, __half>::type;
    const int row = threadIdx.x;  // 0..127 == TMEM lane
    const uint32_t lane_addr = tmem + ((uint32_t)(warp * 32) << 16);
    // keys visible to this row: [0, limit)
    int limit = p.causal ? min(L, prefix + m0 + row + 1) : L;
    limit = max(1, min(limit, kv_end));
    float m_used = -INFINITY;  // reference max (log2 domain) the running sums are expressed in
    float l_run = 0.f;
    uint8_t* p_row = smem + S::kP + row * 128;
    for (int j = 0; j < ntiles; ++j) {
      const int sb = j & 1;
      mbar_wait(bar(kBarSFull + sb), (j / 2) & 1);
      tc_fence_after();
      float s[kExtN];
#pragma unroll
      for (int c = 0; c < kExtN / 32; ++c) {
        uint32_t v[32];
        tmem_ld32(lane_addr + kColS + sb * kExtN + c * 32, v);
        tmem_wait_ld();
#pragma unroll
        for (int i = 0; i < 32; ++i) s[c * 32 + i] = __uint_as_float(v[i]);
      }
      tc_fence_before();
      mbar_arrive(bar(kBarSEmpty + sb));
      const int n0 = j * kExtN;
      if (n0 + kExtN > limit) {
#pragma unroll
        for (int i = 0; i < kExtN; ++i)
          if (n0 + i >= limit) s[i] = -INFINITY;
      }
      float mx4[4] = {s[0], s[1], s[2], s[3]};
#pragma unroll
      for (int i = 4; i < kExtN; i += 4) {
        mx4[0] = fmaxf(mx4[0], s[i]);
        mx4[1] = fmaxf(mx4[1], s[i + 1]);
        mx4[2] = fmaxf(mx4[2], s[i + 2]);
        mx4[3] = fmaxf(mx4[3], s[i + 3]);
      }
      float mx = fmaxf(fmaxf(mx4[0], mx4[1]), fmaxf(mx4[2], mx4[3])) * p.scale_log2;
      float alpha = 1.f;
      bool rescale = false;
      if (m_used == -INFINITY) {
        m_used = (mx == -INFINITY) ? 0.f : mx;  // j == 0 (limit >= 1 => finite)
      } else if (mx > m_used + kRescaleThreshold) {
        alpha = fast_exp2(m_used - mx);
        m_used = mx;
        rescale = true;
      }
      uint32_t pk[kExtN / 2];
      float sum4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int i = 0; i < kExtN; i += 2) {
        const float p0 = fast_exp2(fmaf(s[i], p.scale_log2, -m_used));
        const float p1 = fast_exp2(fmaf(s[i + 1], p.scale_log2, -m_used));
        sum4[(i / 2) & 3] += p0 + p1;
        pk[i / 2] = pack2<kBf16>(p0, p1);
      }
      l_run = l_run * alpha + ((sum4[0] + sum4[1]) + (sum4[2] + sum4[3]));
      // P buffer and O are free once P V of the previous step has completed
      if (j > 0) {
        mbar_wait(bar(kBarPVDone), (j - 1) & 1);
        tc_fence_after();
        if (__any_sync(0xffffffffu, rescale)) {
#pragma unroll
          for (int c = 0; c < D / 32; ++c) {
            uint32_t v[32];
            tmem_ld32(lane_addr + kColO + c * 32, v);
            tmem_wait_ld();
#pragma unroll
            for (int i = 0; i < 32; ++i) v[i] = __float_as_uint(__uint_as_float(v[i]) * alpha);
            tmem_st32(lane_addr + kColO + c * 32, v);
          }
          tmem_wait_st();
        }
      }
      // P[row][key] -> K-major 128B-swizzled smem (two 64-key halves)
#pragma unroll
      for (int c = 0; c < kExtN / 8; ++c) {
        const int half = c / 8, chunk = c % 8;
        uint4 val = make_uint4(pk[c * 4], pk[c * 4 + 1], pk[c * 4 + 2], pk[c * 4 + 3]);
        *reinterpret_cast<uint4*>(p_row + half * (kExtM * 128) + ((chunk ^ (row & 7)) << 4)) = val;
      }
      fence_proxy_async_smem();
      tc_fence_before();
      mbar_arrive(bar(kBarPFull));
    }
    // ---- epilogue ----
    mbar_wait(bar(kBarPVDone), (ntiles - 1) & 1);
    tc_fence_after();
    const float inv = 1.f / l_run;
    const bool valid = row < rows_valid;
    T* dst = reinterpret_cast<T*>(p.o) + (int64_t)(q_beg + m0 + row) * p.o_stride + (int64_t)h * D;
#pragma unroll
    for (int c = 0; c < D / 32; ++c) {
      uint32_t v[32];
      tmem_ld32(lane_addr + kColO + c * 32, v);
      tmem_wait_ld();
      if (valid) {
#pragma unroll
        for (int i = 0; i < 32; i += 8) {
          uint4 w;
          w.x = pack2<kBf16>(__uint_as_float(v[i]) * inv, __uint_as_float(v[i + 1]) * inv);
          w.y = pack2<kBf16>(__uint_as_float(v[i + 2]) * inv, __uint_as_float(v[i + 3]) * inv);
          w.z = pack2<kBf16>(__uint_as_float(v[i + 4]) * inv, __uint_as_float(v[i + 5]) * inv);
          w.w = pack2<kBf16>(__uint_as_float(v[i + 6]) * inv, __uint_as_float(v[i + 7]) * inv);
          *reinterpret_cast<uint4*>(dst + c * 32 + i) = w;
        }
      }
    }
    if (valid && p.lse != nullptr)
      p.lse[(int64_t)(q_beg + m0 + row) * p.Hq + h] = (m_used + log2f(l_run)) * kLn2;
    tc_fence_before();
  }
  __syncthreads();
  if (warp == 6) {
    tc_fence_after();
    tmem_dealloc<kTmemCols>(tmem);
  }
}

}  // namespace b200attn

// ------------------------------------------------------------------ host side
namespace b200attn {

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
                       uint64_t row_stride_bytes, uint32_t box_cols, uint32_t box_rows) {
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
                  box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed with CUresult %d (rows=%llu cols=%llu stride=%llu)",
              (int)r, (unsigned long long)rows, (unsigned long long)cols,
              (unsigned long long)row_stride_bytes);
    return B200ATTN_ERR_CUDA;
  }
  return B200ATTN_OK;
}

template <int D, bool kBf16>
static int launch_extend(const CUtensorMap& tq, const CUtensorMap& tk, const CUtensorMap& tv,
                         const CUtensorMap& tkb, const CUtensorMap& tvb, const ExtendParams& p,
                         int max_qtiles, cudaStream_t st) {
  constexpr int kSmem = ExtSmem<D>::kTotal + 1024;  // + alignment slack
  static bool configured = false;
  if (!configured) {
    B2_CUDA(cudaFuncSetAttribute(extend_kernel<D, kBf16>,
                                 cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem));
    configured = true;
  }
  dim3 grid((unsigned)p.Hq, (unsigned)max_qtiles, (unsigned)p.B);
  extend_kernel<D, kBf16><<<grid, kExtThreads, kSmem, st>>>(tq, tk, tv, tkb, tvb, p);
  B2_LAUNCH_CHECK();
  return B200ATTN_OK;
}

}  // namespace b200attn

extern "C" int64_t b200attn_extend_workspace_bytes(int32_t, int64_t, int32_t) { return 0; }

extern "C" int b200attn_extend(const void* q, int64_t q_stride, const void* k_cache,
                               const void* v_cache, int64_t kv_stride, int64_t num_slots, void* o,
                               int64_t o_stride, float* lse, const int32_t* qo_indptr,
                               const int32_t* kv_indptr, const int32_t* kv_indices, int32_t B,
                               int64_t total_q_host, int32_t max_qo_len_host, int32_t Hq,
                               int32_t Hkv, int32_t D, int32_t page_size, int dtype, int causal,
                               float sm_scale, void* workspace, int64_t workspace_bytes,
                               void* stream) {
  using namespace b200attn;
  (void)workspace;
  (void)workspace_bytes;
  (void)page_size;  // rows are gathered token-granular: any page size, no contiguity assumption
  B2_CHECK_ARG(B >= 0 && total_q_host >= 0 && max_qo_len_host >= 0, "extend: negative size");
  if (B == 0 || total_q_host == 0 || max_qo_len_host == 0) return B200ATTN_OK;
  B2_CHECK_ARG(q && k_cache && v_cache && o && qo_indptr && kv_indptr && kv_indices,
               "extend: null pointer");
  B2_CHECK_ARG(Hq > 0 && Hkv > 0 && Hq % Hkv == 0, "extend: Hq=%d must be a multiple of Hkv=%d",
               Hq, Hkv);
  B2_CHECK_ARG(B <= 65535, "extend: batch too large");
  B2_CHECK_ARG(num_slots > 0 && num_slots < (1ll << 31), "extend: bad num_slots");
  B2_CHECK_ARG(q_stride % 8 == 0 && kv_stride % 8 == 0 && o_stride % 8 == 0,
               "extend: strides must keep rows 16-byte aligned");
  B2_CHECK_ARG((((uintptr_t)q | (uintptr_t)k_cache | (uintptr_t)v_cache | (uintptr_t)o) & 15) == 0,
               "extend: tensors must be 16-byte aligned");
  if (dtype != B200ATTN_BF16 && dtype != B200ATTN_FP16) B2_UNSUPPORTED("extend: dtype %d", dtype);
  if (D != 64 && D != 128) B2_UNSUPPORTED("extend: head_dim %d (supported: 64, 128)", D);
  CUtensorMap tq, tk, tv, tkb, tvb;
  int rc = make_tensor_map_2d(&tq, q, (uint64_t)total_q_host, (uint64_t)Hq * D,
                              (uint64_t)q_stride * 2, 64, kExtM);
  if (rc) return rc;
  rc = make_tensor_map_2d(&tk, k_cache, (uint64_t)num_slots, (uint64_t)Hkv * D,
                          (uint64_t)kv_stride * 2, 64, 1);
  if (rc) return rc;
  rc = make_tensor_map_2d(&tv, v_cache, (uint64_t)num_slots, (uint64_t)Hkv * D,
                          (uint64_t)kv_stride * 2, 64, 1);
  if (rc) return rc;
  rc = make_tensor_map_2d(&tkb, k_cache, (uint64_t)num_slots, (uint64_t)Hkv * D,
                          (uint64_t)kv_stride * 2, 64, kExtBoxRows);
  if (rc) return rc;
  rc = make_tensor_map_2d(&tvb, v_cache, (uint64_t)num_slots, (uint64_t)Hkv * D,
                          (uint64_t)kv_stride * 2, 64, kExtBoxRows);
  if (rc) return rc;
  ExtendParams p;
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
  p.causal = causal ? 1 : 0;
  p.scale_log2 = sm_scale * kLog2e;
  const int max_qtiles = (max_qo_len_host + kExtM - 1) / kExtM;
  cudaStream_t st = (cudaStream_t)stream;
  if (D == 128)
    return dtype == B200ATTN_BF16 ? launch_extend<128, true>(tq, tk, tv, tkb, tvb, p, max_qtiles, st)
                                  : launch_extend<128, false>(tq, tk, tv, tkb, tvb, p, max_qtiles, st);
  return dtype == B200ATTN_BF16 ? launch_extend<64, true>(tq, tk, tv, tkb, tvb, p, max_qtiles, st)
                                : launch_extend<64, false>(tq, tk, tv, tkb, tvb, p, max_qtiles, st);
}
