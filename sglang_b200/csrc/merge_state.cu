// (a6) merge_state: LSE-weighted merge of two partial attention states.
//
// Arithmetic follows python/sglang/kernels/aot/csrc/attention/merge_attn_states.cu:58-104
// (natural-log LSE, +inf treated as -inf).  One thread owns one 16-byte pack of
// the head vector; the (token, head) LSE pair is read once per thread through
// the read-only path.  Pure bandwidth: 3 * T*H*D*sizeof(T) bytes.
#include "common.cuh"

namespace b200attn {

template <typename T>
struct Pack16;
template <>
struct Pack16<__nv_bfloat16> {
  static constexpr int N = 8;
};
template <>
struct Pack16<__half> {
  static constexpr int N = 8;
};
template <>
struct Pack16<float> {
  static constexpr int N = 4;
};

template <typename T>
__device__ __forceinline__ float to_f(T x);
template <>
__device__ __forceinline__ float to_f(__nv_bfloat16 x) {
  return __bfloat162float(x);
}
template <>
__device__ __forceinline__ float to_f(__half x) {
  return __half2float(x);
}
template <>
__device__ __forceinline__ float to_f(float x) {
  return x;
}
template <typename T>
__device__ __forceinline__ T from_f(float x);
template <>
__device__ __forceinline__ __nv_bfloat16 from_f(float x) {
  return __float2bfloat16_rn(x);
}
template <>
__device__ __forceinline__ __half from_f(float x) {
  return __float2half_rn(x);
}
template <>
__device__ __forceinline__ float from_f(float x) {
  return x;
}

template <typename T>
__global__ void __launch_bounds__(256)
    merge_state_kernel(const T* __restrict__ va, const float* __restrict__ sa,
                       const T* __restrict__ vb, const float* __restrict__ sb, T* vo, float* so,
                       int64_t rows /* T*H */, int packs_per_row) {
  constexpr int N = Pack16<T>::N;
  const int64_t gid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= rows * packs_per_row) return;
  const int64_t row = gid / packs_per_row;
  const int pk = (int)(gid - row * packs_per_row);
  float a = __ldg(sa + row), b = __ldg(sb + row);
  if (isinf(a) && a > 0) a = -INFINITY;  // merge_attn_states.cu:66-67
  if (isinf(b) && b > 0) b = -INFINITY;
  const float m = fmaxf(a, b);
  // same operations in the same order as the reference kernel (merge_attn_states.cu:71-78,95: expf, two divisions
  // by the sum, p * p_scale + (s * s_scale)), so that the outputs agree bit for bit with merge_state_v2 on every
  // row it defines; both states empty (where the reference computes -inf - -inf = NaN): zeros and LSE -inf
  float wa, wb, ca, cb;
  if (m == -INFINITY) {
    wa = wb = ca = cb = 0.f;
  } else {
    wa = expf(a - m);
    wb = expf(b - m);
    const float tot = wa + wb;
    ca = wa / tot;
    cb = wb / tot;
  }
  const int64_t off = row * (int64_t)packs_per_row * N + (int64_t)pk * N;
  const uint4 pa = *reinterpret_cast<const uint4*>(va + off);
  const uint4 pb = *reinterpret_cast<const uint4*>(vb + off);
  uint4 po;
  const T* ea = reinterpret_cast<const T*>(&pa);
  const T* eb = reinterpret_cast<const T*>(&pb);
  T* eo = reinterpret_cast<T*>(&po);
#pragma unroll
  for (int i = 0; i < N; ++i) eo[i] = from_f<T>(to_f(ea[i]) * ca + (to_f(eb[i]) * cb));
  *reinterpret_cast<uint4*>(vo + off) = po;
  if (so != nullptr && pk == 0) so[row] = (m == -INFINITY) ? -INFINITY : logf(wa + wb) + m;
}

template <typename T>
static int launch_merge(const void* va, const float* sa, const void* vb, const float* sb,
                        void* vo, float* so, int64_t rows, int D, cudaStream_t st) {
  const int packs = D / Pack16<T>::N;
  const int64_t threads = rows * packs;
  const int64_t blocks = (threads + 255) / 256;
  B2_CHECK_ARG(blocks < (1ll << 31), "merge_state: too many rows");
  merge_state_kernel<T><<<(unsigned)blocks, 256, 0, st>>>((const T*)va, sa, (const T*)vb, sb,
                                                          (T*)vo, so, rows, packs);
  B2_LAUNCH_CHECK();
  return B200ATTN_OK;
}

}  // namespace b200attn

extern "C" int b200attn_merge_state(const void* v_a, const float* s_a, const void* v_b,
                                    const float* s_b, void* v_out, float* s_out, int64_t T,
                                    int32_t H, int32_t D, int dtype, void* stream) {
  using namespace b200attn;
  B2_CHECK_ARG(T >= 0 && H > 0 && D > 0, "merge_state: bad shape");
  if (T == 0) return B200ATTN_OK;
  B2_CHECK_ARG(v_a && s_a && v_b && s_b && v_out, "merge_state: null pointer");
  const int n = (dtype == B200ATTN_FP32) ? 4 : 8;
  B2_CHECK_ARG(D % n == 0, "merge_state: head_dim %d must be a multiple of %d", D, n);
  B2_CHECK_ARG((((uintptr_t)v_a | (uintptr_t)v_b | (uintptr_t)v_out) & 15) == 0,
               "merge_state: v tensors must be 16-byte aligned");
  cudaStream_t st = (cudaStream_t)stream;
  const int64_t rows = T * H;
  switch (dtype) {
    case B200ATTN_BF16:
      return launch_merge<__nv_bfloat16>(v_a, s_a, v_b, s_b, v_out, s_out, rows, D, st);
    case B200ATTN_FP16:
      return launch_merge<__half>(v_a, s_a, v_b, s_b, v_out, s_out, rows, D, st);
    case B200ATTN_FP32:
      return launch_merge<float>(v_a, s_a, v_b, s_b, v_out, s_out, rows, D, st);
    default:
      B2_CHECK_ARG(false, "merge_state: bad dtype %d", dtype);
  }
}
