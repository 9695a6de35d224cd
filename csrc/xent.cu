// Fused softmax cross-entropy forward + backward over [rows, vocab] logits.
//
// Replaces: ATen log_softmax + nll_loss (reference examples/tutorials/mnist_pytorch/model.py:34,
// train.py:43; the LM loss inside HF GPT-2 and GPT-NeoX).  Forward reads the logits exactly once
// (online max/sum), emits per-row loss and log-sum-exp.  Backward recomputes the probabilities
// from logits + lse and may write dlogits *in place* over the logits, so no [rows, vocab]
// probability tensor is ever materialised.
#include "common.cuh"

namespace {

constexpr int kXT = 512;

template <typename T> struct Ld8;
template <> struct Ld8<__nv_bfloat16> {
  static constexpr int kPerVec = 8;
  static __device__ __forceinline__ void ld(const __nv_bfloat16* p, float* o) {
    uint4 u = ldg_stream_u4(reinterpret_cast<const uint4*>(p));
    o[0] = bf16lo(u.x); o[1] = bf16hi(u.x); o[2] = bf16lo(u.y); o[3] = bf16hi(u.y);
    o[4] = bf16lo(u.z); o[5] = bf16hi(u.z); o[6] = bf16lo(u.w); o[7] = bf16hi(u.w);
  }
  static __device__ __forceinline__ void st(__nv_bfloat16* p, const float* o) {
    uint4 u = make_uint4(pack_bf16(o[0], o[1]), pack_bf16(o[2], o[3]), pack_bf16(o[4], o[5]), pack_bf16(o[6], o[7]));
    *reinterpret_cast<uint4*>(p) = u;
  }
  static __device__ __forceinline__ float ld1(const __nv_bfloat16* p) { return __bfloat162float(*p); }
  static __device__ __forceinline__ void st1(__nv_bfloat16* p, float v) { *p = __float2bfloat16(v); }
};
template <> struct Ld8<float> {
  static constexpr int kPerVec = 4;
  static __device__ __forceinline__ void ld(const float* p, float* o) {
    float4 u = ldg_stream(reinterpret_cast<const float4*>(p));
    o[0] = u.x; o[1] = u.y; o[2] = u.z; o[3] = u.w;
  }
  static __device__ __forceinline__ void st(float* p, const float* o) {
    *reinterpret_cast<float4*>(p) = make_float4(o[0], o[1], o[2], o[3]);
  }
  static __device__ __forceinline__ float ld1(const float* p) { return *p; }
  static __device__ __forceinline__ void st1(float* p, float v) { *p = v; }
};

__device__ __forceinline__ void online_merge(float& m, float& s, float m2, float s2) {
  float mn = fmaxf(m, m2);
  if (mn == -INFINITY) { m = mn; s = 0.f; return; }
  s = s * __expf(m - mn) + s2 * __expf(m2 - mn);
  m = mn;
}
__device__ __forceinline__ void online_add(float& m, float& s, float x) {
  if (x > m) { s = s * __expf(m - x) + 1.f; m = x; }
  else if (m != -INFINITY) s += __expf(x - m);
}

template <typename T>
__global__ void __launch_bounds__(kXT)
xent_fwd_kernel(const T* __restrict__ logits, const int64_t* __restrict__ target, float* __restrict__ loss,
                float* __restrict__ lse_out, int64_t rows, int vocab, int64_t ignore_index) {
  __shared__ float sm[32], ss[32];
  constexpr int V = Ld8<T>::kPerVec;
  for (int64_t row = blockIdx.x; row < rows; row += gridDim.x) {
    const T* p = logits + row * (int64_t)vocab;
    // peel to a 16-byte boundary so the body can use 128-bit loads even when vocab is odd
    int head = (int)(((16 - ((uintptr_t)p & 15)) & 15) / sizeof(T));
    if (head > vocab) head = vocab;
    const int nvec = (vocab - head) / V;
    const int tail0 = head + nvec * V;
    float m = -INFINITY, s = 0.f;
    if (threadIdx.x < head) online_add(m, s, Ld8<T>::ld1(p + threadIdx.x));
    for (int i = threadIdx.x; i < nvec; i += kXT) {
      float v[V];
      Ld8<T>::ld(p + head + i * V, v);
      float vm = v[0];
#pragma unroll
      for (int k = 1; k < V; ++k) vm = fmaxf(vm, v[k]);
      if (vm > m) { s *= __expf(m - vm); m = vm; }
      const float mm = (m == -INFINITY) ? 0.f : m;
#pragma unroll
      for (int k = 0; k < V; ++k) s += __expf(v[k] - mm);
    }
    if (tail0 + threadIdx.x < vocab) online_add(m, s, Ld8<T>::ld1(p + tail0 + threadIdx.x));
    // block merge of (m, s)
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      float m2 = __shfl_xor_sync(0xffffffffu, m, o), s2 = __shfl_xor_sync(0xffffffffu, s, o);
      online_merge(m, s, m2, s2);
    }
    __syncthreads();
    if ((threadIdx.x & 31) == 0) { sm[threadIdx.x >> 5] = m; ss[threadIdx.x >> 5] = s; }
    __syncthreads();
    if (threadIdx.x < 32) {
      m = (threadIdx.x < kXT / 32) ? sm[threadIdx.x] : -INFINITY;
      s = (threadIdx.x < kXT / 32) ? ss[threadIdx.x] : 0.f;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        float m2 = __shfl_xor_sync(0xffffffffu, m, o), s2 = __shfl_xor_sync(0xffffffffu, s, o);
        online_merge(m, s, m2, s2);
      }
      if (threadIdx.x == 0) {
        float lse = m + __logf(s);
        int64_t t = target[row];
        lse_out[row] = lse;
        loss[row] = (t == ignore_index || t < 0 || t >= vocab) ? 0.f : (lse - Ld8<T>::ld1(p + t));
      }
    }
  }
}

template <typename T>
__global__ void __launch_bounds__(kXT)
xent_bwd_kernel(const T* logits, const int64_t* __restrict__ target, const float* __restrict__ lse,
                const float* __restrict__ grad_scale, T* dlogits, int64_t rows, int vocab, int64_t ignore_index) {
  constexpr int V = Ld8<T>::kPerVec;
  const float gs = *grad_scale;
  for (int64_t row = blockIdx.x; row < rows; row += gridDim.x) {
    const T* p = logits + row * (int64_t)vocab;
    T* d = dlogits + row * (int64_t)vocab;
    const int64_t t = target[row];
    const bool ign = (t == ignore_index || t < 0 || t >= vocab);
    const float l = lse[row];
    const float g = ign ? 0.f : gs;
    int head = (int)(((16 - ((uintptr_t)p & 15)) & 15) / sizeof(T));
    if (head > vocab) head = vocab;
    const int nvec = (vocab - head) / V;
    const int tail0 = head + nvec * V;
    if (threadIdx.x < head) {
      int c = threadIdx.x;
      Ld8<T>::st1(d + c, (__expf(Ld8<T>::ld1(p + c) - l) - (c == t ? 1.f : 0.f)) * g);
    }
    for (int i = threadIdx.x; i < nvec; i += kXT) {
      float v[V];
      const int c0 = head + i * V;
      Ld8<T>::ld(p + c0, v);
#pragma unroll
      for (int k = 0; k < V; ++k) v[k] = (__expf(v[k] - l) - ((c0 + k) == t ? 1.f : 0.f)) * g;
      Ld8<T>::st(d + c0, v);
    }
    if (tail0 + threadIdx.x < vocab) {
      int c = tail0 + threadIdx.x;
      Ld8<T>::st1(d + c, (__expf(Ld8<T>::ld1(p + c) - l) - (c == t ? 1.f : 0.f)) * g);
    }
  }
}

inline int xent_grid(int64_t rows) {
  int64_t g = rows < (int64_t)kNumSMs * 4 ? rows : (int64_t)kNumSMs * 4;
  return (int)(g < 1 ? 1 : g);
}

}  // namespace

DETB200_API int detb200_xent_fwd(const void* logits, const int64_t* target, float* loss, float* lse, int64_t rows,
                                 int vocab, int64_t ignore_index, int dtype, cudaStream_t stream) {
  if (rows <= 0) return 0;
  if (!logits || !target || !loss || !lse || vocab < 1) { detb200_set_error("xent_fwd: bad argument"); return DETB200_ERR_BAD_ARG; }
  if (dtype == 1) xent_fwd_kernel<__nv_bfloat16><<<xent_grid(rows), kXT, 0, stream>>>((const __nv_bfloat16*)logits, target, loss, lse, rows, vocab, ignore_index);
  else xent_fwd_kernel<float><<<xent_grid(rows), kXT, 0, stream>>>((const float*)logits, target, loss, lse, rows, vocab, ignore_index);
  DETB200_LAUNCH_CHECK();
  return 0;
}

// dlogits may alias logits (in-place backward).  grad_scale: device scalar = dL/dloss_mean / n_valid.
DETB200_API int detb200_xent_bwd(const void* logits, const int64_t* target, const float* lse, const float* grad_scale,
                                 void* dlogits, int64_t rows, int vocab, int64_t ignore_index, int dtype,
                                 cudaStream_t stream) {
  if (rows <= 0) return 0;
  if (!logits || !target || !lse || !grad_scale || !dlogits) { detb200_set_error("xent_bwd: bad argument"); return DETB200_ERR_BAD_ARG; }
  if (dtype == 1) xent_bwd_kernel<__nv_bfloat16><<<xent_grid(rows), kXT, 0, stream>>>((const __nv_bfloat16*)logits, target, lse, grad_scale, (__nv_bfloat16*)dlogits, rows, vocab, ignore_index);
  else xent_bwd_kernel<float><<<xent_grid(rows), kXT, 0, stream>>>((const float*)logits, target, lse, grad_scale, (float*)dlogits, rows, vocab, ignore_index);
  DETB200_LAUNCH_CHECK();
  return 0;
}
