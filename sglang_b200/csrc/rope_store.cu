// (f1) Fused RoPE / M-RoPE + KV store: rotate q and k in place and scatter the rotated k and the
// untouched v rows into the paged pool in ONE pass over q/k/v (the step immediately before attention).
//
// Behaviour follows
//   python/sglang/srt/layers/rotary_embedding/base.py:236-276   (RotaryEmbedding.forward_native)
//   python/sglang/srt/layers/rotary_embedding/utils.py:36-62    (apply_rotary_emb, neox / gpt-j pairing)
//   python/sglang/srt/layers/rotary_embedding/mrope.py:47-51,173-217 (M-RoPE section / interleaved select)
//   python/sglang/kernels/ops/attention/rope.py:183-225 + jit/csrc/elementwise/rope.cuh:56-110,178-260
//     (apply_rope_with_cos_sin_cache_inplace, fused K/V cache store = FusedSetKVBufferArg)
//   python/sglang/kernels/jit/csrc/elementwise/kvcache.cuh:186-218 (store semantics, reserved slot skip)
//
// Two arithmetic modes, each bit-exact against its oracle restatement:
//   B200ATTN_ROPE_F32    : cos/sin read as stored (fp32 cache), x widened to fp32,
//                          o = rn(rn(x1*cos) - rn(x2*sin)) in fp32, ONE rounding to the element type
//                          (the production CUDA kernel's arithmetic, rope.cuh:96-99, without FMA contraction)
//   B200ATTN_ROPE_NATIVE : cos/sin first rounded to the element type, every product and the sum rounded to the
//                          element type (what torch does for bf16 tensors in forward_native: utils.py:50-58)
//
// HBM-bound: per token (Hq+2*Hkv)*D*2 B read + (Hq+Hkv)*D*2 B written in place + 2*Hkv*D*2 B into the pool.
// One thread owns VEC (<= 8) rotation pairs of one head: two 16-byte loads (x1 | x2 halves for neox, 32
// contiguous bytes for gpt-j), the cos/sin chunk through the read-only path (shared by every head of the
// token => L1/L2 resident), two 16-byte stores in place and, for k, two more into the pool row.  Pass-through
// columns of k (rotary_dim < head_dim) and all of v are plain 16-byte copies done by the trailing work items.
#include <cstdlib>

#include "common.cuh"

namespace b200attn {

struct RopeParams {
  uint8_t* q;
  uint8_t* k;
  const uint8_t* v;
  uint8_t* k_cache;
  uint8_t* v_cache;
  const void* loc;
  const void* cos_sin;
  const void* pos;
  int64_t pos_axis_stride;  // elements between the t/h/w rows of positions[3,T]; 0 = 1-D positions
  int64_t T, cache_rows;
  int64_t q_stride, k_stride, v_stride, kc_stride, vc_stride;  // elements
  int64_t size_limit, skip_index;
  int32_t* oob_count;
  int Hq, Hkv, D, Dv, rot;
  int sec0, sec1, sec2;  // M-RoPE sections (pairs per axis); sec0 < 0 = none
  int loc_dtype, pos_dtype;
  int interleaved, neox, store;
  int cache_vec16;  // cos/sin rows and halves are 16-byte aligned
  int abl_copy_only;  // dev ablation (B200ATTN_ROPE_ABL=copy): staged copy without the rotation
  // e4m3 pool (b200attn_rope_store_fp8): rows are bytes, rotated k and v are quantised on the way (register kernel)
  int pool_fp8;
  float k_scale, v_scale;  // <= 0: no division
};

template <bool kBf16>
__device__ __forceinline__ float to_f(uint16_t x) {
  if constexpr (kBf16) return __uint_as_float((uint32_t)x << 16);
  return __half2float(__ushort_as_half(x));
}
template <bool kBf16>
__device__ __forceinline__ uint16_t from_f(float x) {
  if constexpr (kBf16) return __bfloat16_as_ushort(__float2bfloat16_rn(x));
  return __half_as_ushort(__float2half_rn(x));
}
template <bool kBf16>
__device__ __forceinline__ float round_elem(float x) {
  return to_f<kBf16>(from_f<kBf16>(x));
}

template <int VEC>
struct VecOf;
template <>
struct VecOf<8> { using type = uint4; };
template <>
struct VecOf<4> { using type = uint2; };
template <>
struct VecOf<2> { using type = uint32_t; };
template <>
struct VecOf<1> { using type = uint16_t; };

template <int VEC>
union Pack {
  typename VecOf<VEC>::type v;
  uint16_t e[VEC];
  uint32_t w[VEC >= 2 ? VEC / 2 : 1];
};

// packed fp32x2 arithmetic (FMUL2 / FADD2 on sm_100): same round-to-nearest per lane as the scalar ops
__device__ __forceinline__ uint64_t rp_pack(float lo, float hi) {
  uint64_t r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void rp_unpack(uint64_t v, float& lo, float& hi) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ uint64_t rp_mul(uint64_t a, uint64_t b) {
  uint64_t r;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ uint64_t rp_add(uint64_t a, uint64_t b) {
  uint64_t r;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}

// axis of rotation pair i under M-RoPE (mrope.py:47-51 interleaved, :190-199 sectioned)
__device__ __forceinline__ int mrope_axis(const RopeParams& p, int i) {
  if (p.interleaved) {
    const int m = i % 3;
    if (m == 1 && i < p.sec1 * 3) return 1;
    if (m == 2 && i < p.sec2 * 3) return 2;
    return 0;
  }
  return i < p.sec0 ? 0 : (i < p.sec0 + p.sec1 ? 1 : 2);
}

// cos/sin of VEC consecutive pairs of one cache row; 16-byte loads when the row base allows it
template <int VEC, bool kBf16, bool kCacheF32>
__device__ __forceinline__ void load_cos_sin(const RopeParams& p, int64_t pos, int i0, int half,
                                             float (&cs)[VEC], float (&sn)[VEC]) {
  if constexpr (kCacheF32) {
    const float* cr = reinterpret_cast<const float*>(p.cos_sin) + pos * p.rot;
    if constexpr (VEC % 4 == 0) {
      if (p.cache_vec16) {
#pragma unroll
        for (int j = 0; j < VEC; j += 4) {
          const float4 c4 = __ldg(reinterpret_cast<const float4*>(cr + i0 + j));
          const float4 s4 = __ldg(reinterpret_cast<const float4*>(cr + half + i0 + j));
          cs[j] = c4.x; cs[j + 1] = c4.y; cs[j + 2] = c4.z; cs[j + 3] = c4.w;
          sn[j] = s4.x; sn[j + 1] = s4.y; sn[j + 2] = s4.z; sn[j + 3] = s4.w;
        }
        return;
      }
    }
#pragma unroll
    for (int j = 0; j < VEC; ++j) {
      cs[j] = __ldg(cr + i0 + j);
      sn[j] = __ldg(cr + half + i0 + j);
    }
  } else {
    const uint16_t* cr = reinterpret_cast<const uint16_t*>(p.cos_sin) + pos * p.rot;
    if constexpr (VEC == 4) {
      if (p.cache_vec16) {  // rows / halves 16-byte aligned and i0 % 4 == 0 => 8-byte aligned
        Pack<4> c4, s4;
        c4.v = __ldg(reinterpret_cast<const uint2*>(cr + i0));
        s4.v = __ldg(reinterpret_cast<const uint2*>(cr + half + i0));
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          cs[j] = to_f<kBf16>(c4.e[j]);
          sn[j] = to_f<kBf16>(s4.e[j]);
        }
        return;
      }
    }
#pragma unroll
    for (int j = 0; j < VEC; ++j) {
      cs[j] = to_f<kBf16>(__ldg(cr + i0 + j));
      sn[j] = to_f<kBf16>(__ldg(cr + half + i0 + j));
    }
  }
}

// slot of token t for the pool writes; false = skip (reserved slot, kvcache.cuh:215, or out of range)
__device__ __forceinline__ bool store_slot(const RopeParams& p, int64_t t, bool count_oob,
                                           int64_t& slot) {
  slot = load_index(p.loc, p.loc_dtype, t);
  if (slot == p.skip_index) return false;
  if (slot < 0 || slot >= p.size_limit) {
    if (count_oob && p.oob_count != nullptr) atomicAdd(p.oob_count, 1);
    return false;
  }
  return true;
}

template <int VEC, bool kBf16, bool kCacheF32>
__device__ __forceinline__ void token_cos_sin(const RopeParams& p, int64_t t, int i0, int half,
                                              float (&cs)[VEC], float (&sn)[VEC]) {
  if (p.sec0 < 0) {
    int64_t pos = load_index(p.pos, p.pos_dtype, t);
    pos = pos < 0 ? 0 : (pos >= p.cache_rows ? p.cache_rows - 1 : pos);
    load_cos_sin<VEC, kBf16, kCacheF32>(p, pos, i0, half, cs, sn);
  } else if (!p.interleaved && mrope_axis(p, i0) == mrope_axis(p, i0 + VEC - 1)) {
    // the whole chunk lies in one section: one position, vector loads
    const int ax = mrope_axis(p, i0);
    int64_t pos = load_index(p.pos, p.pos_dtype, ax * p.pos_axis_stride + t);
    pos = pos < 0 ? 0 : (pos >= p.cache_rows ? p.cache_rows - 1 : pos);
    load_cos_sin<VEC, kBf16, kCacheF32>(p, pos, i0, half, cs, sn);
  } else {
    int64_t pos3[3];
#pragma unroll
    for (int ax = 0; ax < 3; ++ax) {
      int64_t pos = load_index(p.pos, p.pos_dtype, ax * p.pos_axis_stride + t);
      pos3[ax] = pos < 0 ? 0 : (pos >= p.cache_rows ? p.cache_rows - 1 : pos);
    }
#pragma unroll
    for (int j = 0; j < VEC; ++j) {
      const int ax = mrope_axis(p, i0 + j);
      const int64_t pos = ax == 0 ? pos3[0] : (ax == 1 ? pos3[1] : pos3[2]);
      if constexpr (kCacheF32) {
        const float* cr = reinterpret_cast<const float*>(p.cos_sin) + pos * p.rot;
        cs[j] = __ldg(cr + i0 + j);
        sn[j] = __ldg(cr + half + i0 + j);
      } else {
        const uint16_t* cr = reinterpret_cast<const uint16_t*>(p.cos_sin) + pos * p.rot;
        cs[j] = to_f<kBf16>(__ldg(cr + i0 + j));
        sn[j] = to_f<kBf16>(__ldg(cr + half + i0 + j));
      }
    }
  }
}

// rotate pairs [j0, j0+G) of the thread IN PLACE in (a | b); element j: neox -> x1 = a.e[j], x2 = b.e[j];
// gpt-j -> the 2*VEC contiguous values (a | b) hold pairs (2j, 2j+1)
template <int VEC, int G, bool kBf16, bool kNative>
__device__ __forceinline__ void rotate(Pack<VEC>& a, Pack<VEC>& b, int j0, const float (&cs)[G],
                                       const float (&sn)[G], bool neox) {
#pragma unroll
  for (int g = 0; g < G; ++g) {
    const int j = j0 + g;
    const int e1 = 2 * j, e2 = 2 * j + 1;
    uint16_t r1, r2;
    if (neox) {
      r1 = a.e[j];
      r2 = b.e[j];
    } else {
      r1 = e1 < VEC ? a.e[e1 % VEC] : b.e[e1 % VEC];
      r2 = e2 < VEC ? a.e[e2 % VEC] : b.e[e2 % VEC];
    }
    const float x1 = to_f<kBf16>(r1), x2 = to_f<kBf16>(r2);
    float c_ = cs[g], s_ = sn[g];
    float o1, o2;
    if constexpr (kNative) {
      c_ = round_elem<kBf16>(c_);
      s_ = round_elem<kBf16>(s_);
      o1 = __fsub_rn(round_elem<kBf16>(__fmul_rn(x1, c_)), round_elem<kBf16>(__fmul_rn(x2, s_)));
      o2 = __fadd_rn(round_elem<kBf16>(__fmul_rn(x2, c_)), round_elem<kBf16>(__fmul_rn(x1, s_)));
    } else {
      o1 = __fsub_rn(__fmul_rn(x1, c_), __fmul_rn(x2, s_));
      o2 = __fadd_rn(__fmul_rn(x2, c_), __fmul_rn(x1, s_));
    }
    const uint16_t w1 = from_f<kBf16>(o1), w2 = from_f<kBf16>(o2);
    if (neox) {
      a.e[j] = w1;
      b.e[j] = w2;
    } else {
      if (e1 < VEC) a.e[e1 % VEC] = w1; else b.e[e1 % VEC] = w1;
      if (e2 < VEC) a.e[e2 % VEC] = w2; else b.e[e2 % VEC] = w2;
    }
  }
}

// grid = (ceil(T / TOK), ceil(per_tok / 256)).  A thread owns work item `it` of TOK consecutive tokens and issues
// all of their 16-byte loads before the first dependent instruction (TOK * 32 B in flight per thread).
// VEC 16-bit elements of `val` -> VEC e4m3 bytes at dst (dst is VEC-byte aligned by the launcher's vector choice)
template <int VEC, bool kBf16>
__device__ __forceinline__ void store_e4m3(uint8_t* dst, const typename VecOf<VEC>::type& val, float scale) {
  const uint16_t* e = reinterpret_cast<const uint16_t*>(&val);
  uint8_t out[VEC];
#pragma unroll
  for (int i = 0; i < VEC; ++i) out[i] = quant_e4m3<kBf16>(e[i], scale);
  if constexpr (VEC == 8) *reinterpret_cast<uint2*>(dst) = *reinterpret_cast<const uint2*>(out);
  else if constexpr (VEC == 4) *reinterpret_cast<uint32_t*>(dst) = *reinterpret_cast<const uint32_t*>(out);
  else if constexpr (VEC == 2) *reinterpret_cast<uint16_t*>(dst) = *reinterpret_cast<const uint16_t*>(out);
  else *dst = out[0];
}

template <int VEC, bool kBf16, bool kCacheF32, bool kNative, int TOK>
__global__ void __launch_bounds__(256, TOK == 1 ? 6 : 4) rope_store_kernel(const RopeParams p) {
  using V = typename VecOf<VEC>::type;
  const int half = p.rot >> 1;
  const int rot_items = half / VEC;                 // per head
  const int pass_items = (p.D - p.rot) / VEC;       // per k head (pool copy only)
  const int v_items = p.Dv / VEC;                   // per v head
  const int nq = p.Hq * rot_items, nk = p.Hkv * rot_items;
  const int nkp = p.store ? p.Hkv * pass_items : 0;
  const int nv = p.store ? p.Hkv * v_items : 0;
  const int per_tok = nq + nk + nkp + nv;
  const int64_t t0 = (int64_t)blockIdx.x * TOK;
  int it = blockIdx.y * blockDim.x + threadIdx.x;
  if (it >= per_tok) return;

  if (it >= nq + nk) {  // plain copies into the pool: k pass-through columns, then v
    it -= nq + nk;
    const bool is_kp = it < nkp;
    int64_t col, sstride, dstride;
    const uint8_t* sbase;
    uint8_t* dbase;
    if (is_kp) {
      const int h = it / pass_items, c = it - h * pass_items;
      col = (int64_t)h * p.D + p.rot + c * VEC;
      sbase = p.k; dbase = p.k_cache; sstride = p.k_stride; dstride = p.kc_stride;
    } else {
      col = (int64_t)(it - nkp) * VEC;
      sbase = p.v; dbase = p.v_cache; sstride = p.v_stride; dstride = p.vc_stride;
    }
    V val[TOK];
    int64_t slot[TOK];
    bool ok[TOK];
#pragma unroll
    for (int u = 0; u < TOK; ++u) {
      const int64_t t = t0 + u;
      ok[u] = t < p.T && store_slot(p, t, false, slot[u]);
      if (ok[u]) val[u] = *reinterpret_cast<const V*>(sbase + (t * sstride + col) * 2);
    }
#pragma unroll
    for (int u = 0; u < TOK; ++u) {
      if (!ok[u]) continue;
      if (p.pool_fp8) store_e4m3<VEC, kBf16>(dbase + slot[u] * dstride + col, val[u], is_kp ? p.k_scale : p.v_scale);
      else *reinterpret_cast<V*>(dbase + (slot[u] * dstride + col) * 2) = val[u];
    }
    return;
  }

  const bool is_k = it >= nq;
  if (is_k) it -= nq;
  const int h = it / rot_items, c = it - h * rot_items;
  const int i0 = c * VEC;  // first rotation pair of this thread
  // neox: pair i = (x[i], x[i+half]);  gpt-j: pair i = (x[2i], x[2i+1])
  const int off_a = (p.neox ? i0 : 2 * i0) * 2, off_b = (p.neox ? i0 + half : 2 * i0 + VEC) * 2;
  uint8_t* const base = is_k ? p.k : p.q;
  const int64_t stride = is_k ? p.k_stride : p.q_stride;
  Pack<VEC> a[TOK], b[TOK];
#pragma unroll
  for (int u = 0; u < TOK; ++u) {
    const int64_t t = t0 + u;
    if (t < p.T) {
      const uint8_t* row = base + (t * stride + (int64_t)h * p.D) * 2;
      a[u].v = *reinterpret_cast<const V*>(row + off_a);
      b[u].v = *reinterpret_cast<const V*>(row + off_b);
    }
  }
#pragma unroll
  for (int u = 0; u < TOK; ++u) {
    const int64_t t = t0 + u;
    if (t >= p.T) break;
    constexpr int G = VEC < 4 ? VEC : 4;  // cos/sin live 4 pairs at a time (register budget)
#pragma unroll
    for (int j0 = 0; j0 < VEC; j0 += G) {
      float cs[G], sn[G];
      token_cos_sin<G, kBf16, kCacheF32>(p, t, i0 + j0, half, cs, sn);
      rotate<VEC, G, kBf16, kNative>(a[u], b[u], j0, cs, sn, p.neox);
    }
    uint8_t* row = base + (t * stride + (int64_t)h * p.D) * 2;
    *reinterpret_cast<V*>(row + off_a) = a[u].v;
    *reinterpret_cast<V*>(row + off_b) = b[u].v;
    int64_t slot;
    if (is_k && p.store && store_slot(p, t, it == 0, slot)) {
      if (p.pool_fp8) {
        uint8_t* crow = p.k_cache + slot * p.kc_stride + (int64_t)h * p.D;
        store_e4m3<VEC, kBf16>(crow + off_a / 2, a[u].v, p.k_scale);
        store_e4m3<VEC, kBf16>(crow + off_b / 2, b[u].v, p.k_scale);
      } else {
        uint8_t* crow = p.k_cache + (slot * p.kc_stride + (int64_t)h * p.D) * 2;
        *reinterpret_cast<V*>(crow + off_a) = a[u].v;
        *reinterpret_cast<V*>(crow + off_b) = b[u].v;
      }
    }
  }
}


// the (clamped) cache rows of token t: one for 1-D positions, three (t/h/w) for M-RoPE
__device__ __forceinline__ void resolve_pos(const RopeParams& p, int64_t t, int64_t (&pos3)[3]) {
  if (p.sec0 < 0) {
    int64_t pos = load_index(p.pos, p.pos_dtype, t);
    pos = pos < 0 ? 0 : (pos >= p.cache_rows ? p.cache_rows - 1 : pos);
    pos3[0] = pos3[1] = pos3[2] = pos;
  } else {
#pragma unroll
    for (int ax = 0; ax < 3; ++ax) {
      int64_t pos = load_index(p.pos, p.pos_dtype, ax * p.pos_axis_stride + t);
      pos3[ax] = pos < 0 ? 0 : (pos >= p.cache_rows ? p.cache_rows - 1 : pos);
    }
  }
}

template <int VEC, bool kBf16, bool kCacheF32>
__device__ __forceinline__ void cos_sin_at(const RopeParams& p, const int64_t (&pos3)[3], int i0, int half,
                                           float (&cs)[VEC], float (&sn)[VEC]) {
  if (p.sec0 < 0) {
    load_cos_sin<VEC, kBf16, kCacheF32>(p, pos3[0], i0, half, cs, sn);
  } else if (!p.interleaved && mrope_axis(p, i0) == mrope_axis(p, i0 + VEC - 1)) {
    const int ax = mrope_axis(p, i0);
    load_cos_sin<VEC, kBf16, kCacheF32>(p, ax == 0 ? pos3[0] : (ax == 1 ? pos3[1] : pos3[2]), i0, half, cs, sn);
  } else {
#pragma unroll
    for (int j = 0; j < VEC; ++j) {
      const int ax = mrope_axis(p, i0 + j);
      const int64_t pos = ax == 0 ? pos3[0] : (ax == 1 ? pos3[1] : pos3[2]);
      if constexpr (kCacheF32) {
        const float* cr = reinterpret_cast<const float*>(p.cos_sin) + pos * p.rot;
        cs[j] = __ldg(cr + i0 + j);
        sn[j] = __ldg(cr + half + i0 + j);
      } else {
        const uint16_t* cr = reinterpret_cast<const uint16_t*>(p.cos_sin) + pos * p.rot;
        cs[j] = to_f<kBf16>(__ldg(cr + i0 + j));
        sn[j] = to_f<kBf16>(__ldg(cr + half + i0 + j));
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------------------
// Bulk-copy pipeline variant (large batches, every row 16-byte aligned).  The register kernel above is latency
// bound (ncu: DRAM 46 %, L1 36 %, nothing saturated): its bytes in flight are capped by registers.  Here the
// copy engine moves whole token rows: cp.async.bulk global -> shared (q | k | v row of a token, one mbarrier
// per ring stage), all threads rotate q and k IN PLACE in shared memory, then four bulk stores write the q row
// and the k row back in place and the k and v rows into the pool slot.  Bytes in flight are a function of the
// ring depth, not of registers; v never enters a register.
constexpr int kRopeBulkThreads = 256;

__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n" ::"r"(dst),
      "l"(src), "r"(bytes), "r"(bar)
      : "memory");
}
__device__ __forceinline__ void bulk_s2g(void* dst, uint32_t src, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;\n" ::"l"(dst), "r"(src),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;\n" ::: "memory"); }
template <int N>
__device__ __forceinline__ void bulk_wait_read() {
  asm volatile("cp.async.bulk.wait_group.read %0;\n" ::"n"(N) : "memory");
}
__device__ __forceinline__ void rope_mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void rope_mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void rope_mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n.reg .pred p;\nWAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\nbra WAIT_%=;\nDONE_%=:\n}\n" ::"r"(bar),
      "r"(parity)
      : "memory");
}

// cos/sin of G consecutive pairs from the cache rows staged in shared memory ([1 or 3][rot] of the cache dtype)
template <int G, bool kBf16, bool kCacheF32>
__device__ __forceinline__ void cos_sin_smem(const RopeParams& p, const uint8_t* rows, int i0, int half,
                                             float (&cs)[G], float (&sn)[G]) {
  constexpr int kEsz = kCacheF32 ? 4 : 2;
  const int row_bytes = p.rot * kEsz;
  const bool uniform = p.sec0 < 0 || (!p.interleaved && mrope_axis(p, i0) == mrope_axis(p, i0 + G - 1));
  if (uniform) {
    const uint8_t* r = rows + (p.sec0 < 0 ? 0 : mrope_axis(p, i0)) * row_bytes;
    if constexpr (kCacheF32 && G == 4) {
      const float4 c4 = *reinterpret_cast<const float4*>(r + i0 * 4);
      const float4 s4 = *reinterpret_cast<const float4*>(r + (half + i0) * 4);
      cs[0] = c4.x; cs[1] = c4.y; cs[2] = c4.z; cs[3] = c4.w;
      sn[0] = s4.x; sn[1] = s4.y; sn[2] = s4.z; sn[3] = s4.w;
      return;
    }
#pragma unroll
    for (int j = 0; j < G; ++j) {
      if constexpr (kCacheF32) {
        cs[j] = reinterpret_cast<const float*>(r)[i0 + j];
        sn[j] = reinterpret_cast<const float*>(r)[half + i0 + j];
      } else {
        cs[j] = to_f<kBf16>(reinterpret_cast<const uint16_t*>(r)[i0 + j]);
        sn[j] = to_f<kBf16>(reinterpret_cast<const uint16_t*>(r)[half + i0 + j]);
      }
    }
    return;
  }
#pragma unroll
  for (int j = 0; j < G; ++j) {
    const uint8_t* r = rows + mrope_axis(p, i0 + j) * row_bytes;
    if constexpr (kCacheF32) {
      cs[j] = reinterpret_cast<const float*>(r)[i0 + j];
      sn[j] = reinterpret_cast<const float*>(r)[half + i0 + j];
    } else {
      cs[j] = to_f<kBf16>(reinterpret_cast<const uint16_t*>(r)[i0 + j]);
      sn[j] = to_f<kBf16>(reinterpret_cast<const uint16_t*>(r)[half + i0 + j]);
    }
  }
}

// Stage layout: [q row | k row | v row | cos/sin cache row(s) of the token].  The cache rows ride along with the
// token (the producer thread knows the position three tokens ahead), so the rotation reads cos/sin from shared
// memory: no global-memory latency on the per-token critical path of the CTA.
template <bool kBf16, bool kCacheF32, bool kNative>
__global__ void __launch_bounds__(kRopeBulkThreads, 4)
    rope_store_bulk_kernel(const RopeParams p, const int stages) {
  constexpr int VEC = 8;
  using V = uint4;
  extern __shared__ __align__(128) uint8_t rope_smem[];
  const uint32_t q_bytes = (uint32_t)p.Hq * p.D * 2, k_bytes = (uint32_t)p.Hkv * p.D * 2;
  const uint32_t v_bytes = p.store ? (uint32_t)p.Hkv * p.Dv * 2 : 0;
  const uint32_t cs_row = (uint32_t)p.rot * (kCacheF32 ? 4 : 2);
  const int n_cs = p.sec0 < 0 ? 1 : 3;
  const uint32_t cs_off = q_bytes + k_bytes + v_bytes;
  const uint32_t stage_bytes = cs_off + n_cs * cs_row;
  uint8_t* ring = rope_smem + 128;  // first 128 B: mbarriers
  const uint32_t bar0 = smem_u32(rope_smem);
  const uint32_t ring0 = smem_u32(ring);
  const int tid = threadIdx.x;
  if (tid == 0) {
    for (int s = 0; s < stages; ++s) rope_mbar_init(bar0 + 8 * s, 1);
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
  }
  __syncthreads();
  // tokens of this CTA: blockIdx.x, blockIdx.x + gridDim.x, ...
  const int64_t n_my = p.T > blockIdx.x ? (p.T - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
  auto token = [&](int64_t i) { return (int64_t)blockIdx.x + i * gridDim.x; };
  auto issue_load = [&](int64_t i, const int64_t (&pos3)[3]) {  // thread 0 only
    const int s = (int)(i % stages);
    const int64_t t = token(i);
    const uint32_t dst = ring0 + s * stage_bytes, bar = bar0 + 8 * s;
    rope_mbar_expect_tx(bar, stage_bytes);
    bulk_g2s(dst, p.q + t * p.q_stride * 2, q_bytes, bar);
    bulk_g2s(dst + q_bytes, p.k + t * p.k_stride * 2, k_bytes, bar);
    if (v_bytes) bulk_g2s(dst + q_bytes + k_bytes, p.v + t * p.v_stride * 2, v_bytes, bar);
    for (int a = 0; a < n_cs; ++a)
      bulk_g2s(dst + cs_off + a * cs_row, reinterpret_cast<const uint8_t*>(p.cos_sin) + pos3[a] * cs_row, cs_row, bar);
  };
  int64_t pos_pref[3] = {0, 0, 0};  // thread 0: cache rows of the next token to load (fetched one iteration early)
  if (tid == 0) {
    for (int64_t i = 0; i < stages - 1 && i < n_my; ++i) {
      resolve_pos(p, token(i), pos_pref);
      issue_load(i, pos_pref);
    }
    if (stages - 1 < n_my) resolve_pos(p, token(stages - 1), pos_pref);
  }

  const int half = p.rot >> 1;
  const int rot_items = half / VEC;
  const int n_rot = (p.Hq + p.Hkv) * rot_items;  // q heads then k heads: rows are adjacent in the stage
  for (int64_t i = 0; i < n_my; ++i) {
    const int s = (int)(i % stages);
    const int64_t t = token(i);
    uint8_t* st = ring + s * stage_bytes;
    int64_t pos_next[3] = {0, 0, 0};
    if (tid == 0 && i + stages < n_my) resolve_pos(p, token(i + stages), pos_next);  // in flight during the rotation
    rope_mbar_wait(bar0 + 8 * s, (uint32_t)((i / stages) & 1));
    for (int it = tid; it < (p.abl_copy_only ? 0 : n_rot); it += blockDim.x) {
      const int h = it / rot_items, c = it - h * rot_items;  // h indexes q heads then k heads
      const int i0 = c * VEC;
      uint8_t* row = st + (size_t)h * p.D * 2;
      V* pa = reinterpret_cast<V*>(row + (p.neox ? i0 : 2 * i0) * 2);
      V* pb = reinterpret_cast<V*>(row + (p.neox ? i0 + half : 2 * i0 + VEC) * 2);
      Pack<VEC> a, b;
      a.v = *pa;
      b.v = *pb;
#pragma unroll
      for (int j0 = 0; j0 < VEC; j0 += 4) {
        float cs[4], sn[4];
        cos_sin_smem<4, kBf16, kCacheF32>(p, st + cs_off, i0 + j0, half, cs, sn);
        rotate<VEC, 4, kBf16, kNative>(a, b, j0, cs, sn, p.neox);
      }
      *pa = a.v;
      *pb = b.v;
    }
    asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");  // generic writes -> visible to the bulk stores
    __syncthreads();
    if (tid == 0) {
      const uint32_t src = ring0 + s * stage_bytes;
      bulk_s2g(p.q + t * p.q_stride * 2, src, q_bytes);
      bulk_s2g(p.k + t * p.k_stride * 2, src + q_bytes, k_bytes);
      int64_t slot;
      if (p.store && store_slot(p, t, true, slot)) {
        bulk_s2g(p.k_cache + slot * p.kc_stride * 2, src + q_bytes, k_bytes);
        bulk_s2g(p.v_cache + slot * p.vc_stride * 2, src + q_bytes + k_bytes, v_bytes);
      }
      bulk_commit();
      // the stage of token i-1 is re-filled next: all but the group just committed must have read their smem
      bulk_wait_read<1>();
      if (i + stages - 1 < n_my) issue_load(i + stages - 1, pos_pref);
      pos_pref[0] = pos_next[0];
      pos_pref[1] = pos_next[1];
      pos_pref[2] = pos_next[2];
    }
  }
  if (tid == 0) bulk_wait_read<0>();  // smem must outlive the last stores
}

// Returns B200ATTN_OK after launching, or -1 when the shape is not eligible (caller falls back).
static int launch_rope_bulk(const RopeParams& p, int elem_dtype, int cache_dtype, int mode, cudaStream_t st) {
  const int64_t cs_row = (int64_t)p.rot * (cache_dtype == B200ATTN_FP32 ? 4 : 2);
  if (!p.cache_vec16 || cs_row % 16 != 0) return -1;  // the cache rows are bulk-copied too
  const int64_t stage_bytes =
      ((int64_t)p.Hq * p.D + (int64_t)p.Hkv * p.D + (p.store ? (int64_t)p.Hkv * p.Dv : 0)) * 2 +
      (p.sec0 < 0 ? 1 : 3) * cs_row;
  if (stage_bytes > 96 * 1024 || p.T >= (1ll << 31)) return -1;
  // measured (tools/bench_rope.py, T = 131072): 128 threads x 2 stages x 8 CTAs/SM 6.24 TB/s; 256 threads x 4 stages
  // x 4 CTAs/SM 4.92 TB/s: many small CTAs interleave their load -> rotate -> store chains better
  int budget_kb = 27, threads = 128;
  if (const char* e = getenv("B200ATTN_ROPE_STAGE_KB")) budget_kb = atoi(e);   // dev sweep hooks
  if (const char* e = getenv("B200ATTN_ROPE_THREADS")) threads = atoi(e);
  if (threads != 64 && threads != 128 && threads != 256) threads = 128;
  int stages = (int)((int64_t)budget_kb * 1024 / stage_bytes);
  stages = stages < 2 ? 2 : (stages > 8 ? 8 : stages);
  const int smem = 128 + stages * (int)stage_bytes;
  const int sms = sm_count();
  const bool bf16 = elem_dtype == B200ATTN_BF16;
  const bool cf32 = cache_dtype == B200ATTN_FP32;
  const bool native = mode == B200ATTN_ROPE_NATIVE;
#define B2_ROPE_BULK(BF, CF, NA)                                                                      \
  if (bf16 == BF && cf32 == CF && native == NA) {                                                     \
    B2_CUDA(cudaFuncSetAttribute(rope_store_bulk_kernel<BF, CF, NA>,                                  \
                                 cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));           \
    int occ = 0; /* persistent grid = exactly the resident CTAs (registers may bind before smem) */   \
    B2_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, rope_store_bulk_kernel<BF, CF, NA>,   \
                                                          threads, smem));                            \
    if (occ < 1) return -1;                                                                           \
    const int64_t full = (int64_t)sms * occ;                                                          \
    const int64_t grid = p.T < full ? p.T : full;                                                     \
    rope_store_bulk_kernel<BF, CF, NA><<<(unsigned)grid, threads, smem, st>>>(p, stages);             \
    B2_LAUNCH_CHECK();                                                                                \
    return B200ATTN_OK;                                                                               \
  }
  B2_ROPE_BULK(true, true, false)
  B2_ROPE_BULK(true, true, true)
  B2_ROPE_BULK(true, false, false)
  B2_ROPE_BULK(true, false, true)
  B2_ROPE_BULK(false, true, false)
  B2_ROPE_BULK(false, true, true)
  B2_ROPE_BULK(false, false, false)
  B2_ROPE_BULK(false, false, true)
#undef B2_ROPE_BULK
  return -1;
}

template <int VEC>
static int launch_rope(const RopeParams& p, int elem_dtype, int cache_dtype, int mode,
                       cudaStream_t st) {
  const int half = p.rot / 2;
  const int64_t per_tok = (int64_t)(p.Hq + p.Hkv) * (half / VEC) +
                          (p.store ? (int64_t)p.Hkv * ((p.D - p.rot) / VEC + p.Dv / VEC) : 0);
  B2_CHECK_ARG(p.T < (1ll << 31) && per_tok < 65535ll * 256, "rope_store: too many work items");
  // measured (tools/bench_rope.py, T = 131072): 2 tokens per thread 4.56 TB/s, 1 token 3.99, 4 tokens 3.48;
  // small batches keep one token per thread so that every SM gets work
  int tok = p.T * ((per_tok + 255) / 256) >= 8 * 148 * 2 ? 2 : 1;
  if (const char* e = getenv("B200ATTN_ROPE_TOK")) tok = atoi(e) == 2 ? 2 : 1;  // dev ablation hook
  const dim3 blocks((unsigned)((p.T + tok - 1) / tok), (unsigned)((per_tok + 255) / 256));
  const bool bf16 = elem_dtype == B200ATTN_BF16;
  const bool cf32 = cache_dtype == B200ATTN_FP32;
  const bool native = mode == B200ATTN_ROPE_NATIVE;
#define B2_ROPE(BF, CF, NA)                                                                  \
  if (bf16 == BF && cf32 == CF && native == NA) {                                            \
    if (tok == 2) rope_store_kernel<VEC, BF, CF, NA, 2><<<blocks, 256, 0, st>>>(p);          \
    else rope_store_kernel<VEC, BF, CF, NA, 1><<<blocks, 256, 0, st>>>(p);                   \
    B2_LAUNCH_CHECK();                                                                       \
    return B200ATTN_OK;                                                                      \
  }
  B2_ROPE(true, true, false)
  B2_ROPE(true, true, true)
  B2_ROPE(true, false, false)
  B2_ROPE(true, false, true)
  B2_ROPE(false, true, false)
  B2_ROPE(false, true, true)
  B2_ROPE(false, false, false)
  B2_ROPE(false, false, true)
#undef B2_ROPE
  return B200ATTN_ERR_INVALID;
}

}  // namespace b200attn

static int rope_store_impl(void* q, void* k, const void* v, void* k_cache, void* v_cache,
                           const void* loc, int loc_dtype, const void* cos_sin_cache,
                           int cache_dtype, int64_t cache_rows, const void* positions,
                           int pos_dtype, int64_t pos_axis_stride,
                           const int32_t* mrope_section_host, int mrope_interleaved,
                           int64_t T, int32_t Hq, int32_t Hkv, int32_t D, int32_t Dv,
                           int32_t rotary_dim, int is_neox, int64_t q_stride,
                           int64_t k_stride, int64_t v_stride, int64_t k_cache_stride,
                           int64_t v_cache_stride, int64_t size_limit, int64_t skip_index,
                           int32_t* oob_count, int elem_dtype, int mode, void* stream, bool pool_fp8,
                           float k_scale, float v_scale) {
  using namespace b200attn;
  B2_CHECK_ARG(T >= 0, "rope_store: negative T");
  if (T == 0) return B200ATTN_OK;
  B2_CHECK_ARG(q && k && cos_sin_cache && positions, "rope_store: null pointer");
  B2_CHECK_ARG(elem_dtype == B200ATTN_BF16 || elem_dtype == B200ATTN_FP16,
               "rope_store: q/k must be bf16 or fp16");
  B2_CHECK_ARG(cache_dtype == B200ATTN_FP32 || cache_dtype == elem_dtype,
               "rope_store: cos_sin_cache must be fp32 or the element dtype");
  B2_CHECK_ARG(mode == B200ATTN_ROPE_F32 || mode == B200ATTN_ROPE_NATIVE, "rope_store: bad mode");
  B2_CHECK_ARG(Hq > 0 && Hkv > 0 && D > 0 && rotary_dim > 0 && rotary_dim <= D &&
                   (rotary_dim & 1) == 0,
               "rope_store: bad head geometry (Hq %d Hkv %d D %d rot %d)", Hq, Hkv, D, rotary_dim);
  B2_CHECK_ARG(cache_rows > 0, "rope_store: empty cos_sin_cache");
  B2_CHECK_ARG(q_stride >= (int64_t)Hq * D && k_stride >= (int64_t)Hkv * D,
               "rope_store: q/k stride smaller than row");
  const bool store = k_cache != nullptr;
  if (store) {
    B2_CHECK_ARG(v && v_cache && loc && Dv > 0, "rope_store: store needs v, v_cache, loc");
    B2_CHECK_ARG(loc_dtype == B200ATTN_I32 || loc_dtype == B200ATTN_I64, "rope_store: bad loc dtype");
    B2_CHECK_ARG(v_stride >= (int64_t)Hkv * Dv && k_cache_stride >= (int64_t)Hkv * D &&
                     v_cache_stride >= (int64_t)Hkv * Dv,
                 "rope_store: stride smaller than row");
  }
  RopeParams p{};
  p.q = (uint8_t*)q; p.k = (uint8_t*)k; p.v = (const uint8_t*)v;
  p.k_cache = (uint8_t*)k_cache; p.v_cache = (uint8_t*)v_cache;
  p.loc = loc; p.cos_sin = cos_sin_cache; p.pos = positions;
  p.pos_axis_stride = pos_axis_stride; p.T = T; p.cache_rows = cache_rows;
  p.q_stride = q_stride; p.k_stride = k_stride; p.v_stride = v_stride;
  p.kc_stride = k_cache_stride; p.vc_stride = v_cache_stride;
  p.size_limit = size_limit; p.skip_index = skip_index; p.oob_count = oob_count;
  p.Hq = Hq; p.Hkv = Hkv; p.D = D; p.Dv = store ? Dv : 0; p.rot = rotary_dim;
  p.loc_dtype = loc_dtype; p.pos_dtype = pos_dtype;
  p.interleaved = mrope_interleaved; p.neox = is_neox; p.store = store;
  p.pool_fp8 = pool_fp8 && store; p.k_scale = k_scale; p.v_scale = v_scale;
  if (mrope_section_host != nullptr) {
    p.sec0 = mrope_section_host[0]; p.sec1 = mrope_section_host[1]; p.sec2 = mrope_section_host[2];
    B2_CHECK_ARG(p.sec0 >= 0 && p.sec1 >= 0 && p.sec2 >= 0 &&
                     p.sec0 + p.sec1 + p.sec2 == rotary_dim / 2,
                 "rope_store: mrope_section must sum to rotary_dim/2");
    // 0 = the three axes alias one row (text-only batches: positions.unsqueeze(0).expand(3, -1))
    B2_CHECK_ARG(pos_axis_stride >= T || pos_axis_stride == 0, "rope_store: positions[3,T] axis stride < T");
  } else {
    p.sec0 = -1;
    B2_CHECK_ARG(pos_axis_stride == 0, "rope_store: 2-D positions need mrope_section");
  }
  B2_CHECK_ARG(pos_dtype == B200ATTN_I32 || pos_dtype == B200ATTN_I64, "rope_store: bad pos dtype");
  // widest vector (in elements) dividing the rotation half, the pass-through width, every stride and
  // every base address
  uint64_t bits = (uint64_t)(rotary_dim / 2) | (uint64_t)D | (uint64_t)q_stride | (uint64_t)k_stride |
                  ((uint64_t)(uintptr_t)q >> 1) | ((uint64_t)(uintptr_t)k >> 1);
  B2_CHECK_ARG((((uintptr_t)q | (uintptr_t)k) & 1) == 0, "rope_store: unaligned q/k");
  if (store) {
    bits |= (uint64_t)Dv | (uint64_t)(D - rotary_dim) | (uint64_t)v_stride | (uint64_t)k_cache_stride |
            (uint64_t)v_cache_stride | ((uint64_t)(uintptr_t)v >> 1) |
            ((uint64_t)(uintptr_t)k_cache >> (pool_fp8 ? 0 : 1)) | ((uint64_t)(uintptr_t)v_cache >> (pool_fp8 ? 0 : 1));
  } else {
    bits |= (uint64_t)(D - rotary_dim);
  }
  {
    const int64_t esz = cache_dtype == B200ATTN_FP32 ? 4 : 2;
    p.cache_vec16 = ((uintptr_t)cos_sin_cache % 16 == 0) && ((rotary_dim / 2) * esz % 16 == 0);
  }
  {
    const char* e = getenv("B200ATTN_ROPE_ABL");
    p.abl_copy_only = e != nullptr && e[0] == 'c';
  }
  cudaStream_t st = (cudaStream_t)stream;
  if ((bits & 7) == 0) {
    // bulk pipeline: every row start must be 16-byte aligned for cp.async.bulk (bits covers strides / bases) and
    // the pass-through columns ride along in the row; small batches keep the register kernel (more CTAs)
    const char* e = getenv("B200ATTN_ROPE_BULK");  // dev ablation hook: 0 = never, 1 = always
    const bool want = (e ? atoi(e) != 0 : T >= 2 * sm_count()) && !p.pool_fp8;  // bulk stores move 16-bit rows only
    if (want && (!store || (D - rotary_dim) % 8 == 0)) {
      const int rc = launch_rope_bulk(p, elem_dtype, cache_dtype, mode, st);
      if (rc >= 0) return rc;
    }
    return launch_rope<8>(p, elem_dtype, cache_dtype, mode, st);
  }
  if ((bits & 3) == 0) return launch_rope<4>(p, elem_dtype, cache_dtype, mode, st);
  if ((bits & 1) == 0) return launch_rope<2>(p, elem_dtype, cache_dtype, mode, st);
  return launch_rope<1>(p, elem_dtype, cache_dtype, mode, st);
}

extern "C" int b200attn_rope_store(void* q, void* k, const void* v, void* k_cache, void* v_cache,
                                   const void* loc, int loc_dtype, const void* cos_sin_cache,
                                   int cache_dtype, int64_t cache_rows, const void* positions,
                                   int pos_dtype, int64_t pos_axis_stride,
                                   const int32_t* mrope_section_host, int mrope_interleaved,
                                   int64_t T, int32_t Hq, int32_t Hkv, int32_t D, int32_t Dv,
                                   int32_t rotary_dim, int is_neox, int64_t q_stride,
                                   int64_t k_stride, int64_t v_stride, int64_t k_cache_stride,
                                   int64_t v_cache_stride, int64_t size_limit, int64_t skip_index,
                                   int32_t* oob_count, int elem_dtype, int mode, void* stream) {
  return rope_store_impl(q, k, v, k_cache, v_cache, loc, loc_dtype, cos_sin_cache, cache_dtype, cache_rows,
                         positions, pos_dtype, pos_axis_stride, mrope_section_host, mrope_interleaved, T, Hq, Hkv,
                         D, Dv, rotary_dim, is_neox, q_stride, k_stride, v_stride, k_cache_stride, v_cache_stride,
                         size_limit, skip_index, oob_count, elem_dtype, mode, stream, false, 0.f, 0.f);
}

extern "C" int b200attn_rope_store_fp8(void* q, void* k, const void* v, void* k_cache, void* v_cache,
                                       const void* loc, int loc_dtype, const void* cos_sin_cache,
                                       int cache_dtype, int64_t cache_rows, const void* positions,
                                       int pos_dtype, int64_t pos_axis_stride,
                                       const int32_t* mrope_section_host, int mrope_interleaved,
                                       int64_t T, int32_t Hq, int32_t Hkv, int32_t D, int32_t Dv,
                                       int32_t rotary_dim, int is_neox, int64_t q_stride,
                                       int64_t k_stride, int64_t v_stride, int64_t k_cache_stride,
                                       int64_t v_cache_stride, int64_t size_limit, int64_t skip_index,
                                       int32_t* oob_count, int elem_dtype, int mode, float k_scale,
                                       float v_scale, void* stream) {
  if (k_cache == nullptr) {
    b200attn::set_error("rope_store_fp8: the e4m3 pool is required");
    return B200ATTN_ERR_INVALID;
  }
  return rope_store_impl(q, k, v, k_cache, v_cache, loc, loc_dtype, cos_sin_cache, cache_dtype, cache_rows,
                         positions, pos_dtype, pos_axis_stride, mrope_section_host, mrope_interleaved, T, Hq, Hkv,
                         D, Dv, rotary_dim, is_neox, q_stride, k_stride, v_stride, k_cache_stride, v_cache_stride,
                         size_limit, skip_index, oob_count, elem_dtype, mode, stream, true, k_scale, v_scale);
}
