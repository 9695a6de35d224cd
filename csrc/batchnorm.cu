// Training-mode BatchNorm for channels-last activations viewed as [M = N*H*W, C] (SURVEY K14 row:
// "Convs/BN/ReLU ... optional"; measured on B200 to be 46 % of the ResNet-18/CIFAR step when left
// to ATen's batch_norm_*_channels_last kernels, see profiles/r1_launches_resnet_bs256_graph_baseline.csv).
//
// Replaces: ATen batch_norm_collect_statistics / transform_input / backward_reduce /
// backward_elemt reached from nn.BatchNorm2d in the user's model under the reference harness.
//
// Forward  = stats (one read of x, per-CTA partial sums; the last CTA to finish merges the partials
//            with Chan's parallel-variance formula, updates running stats and emits per-channel
//            scale/shift) + apply (x*scale+shift, second read usually served by the 126 MB L2).
// Backward = reduce (sum dy, sum dy*xhat -> dgamma, dbeta and per-channel coefficients) + apply.
// All loads are 128-bit; the channel group of a thread is fixed, so per-channel values live in
// registers for the whole grid-stride loop.
#include "common.cuh"

namespace {

constexpr int kBT = 256;

template <typename T> struct Vec;
template <> struct Vec<__nv_bfloat16> {
  static constexpr int N = 8;
  static __device__ __forceinline__ void ld(const __nv_bfloat16* p, float* o) {
    uint4 u = *reinterpret_cast<const uint4*>(p);
    o[0] = bf16lo(u.x); o[1] = bf16hi(u.x); o[2] = bf16lo(u.y); o[3] = bf16hi(u.y);
    o[4] = bf16lo(u.z); o[5] = bf16hi(u.z); o[6] = bf16lo(u.w); o[7] = bf16hi(u.w);
  }
  static __device__ __forceinline__ void st(__nv_bfloat16* p, const float* o) {
    *reinterpret_cast<uint4*>(p) = make_uint4(pack_bf16(o[0], o[1]), pack_bf16(o[2], o[3]), pack_bf16(o[4], o[5]), pack_bf16(o[6], o[7]));
  }
};
template <> struct Vec<float> {
  static constexpr int N = 4;
  static __device__ __forceinline__ void ld(const float* p, float* o) {
    float4 u = *reinterpret_cast<const float4*>(p);
    o[0] = u.x; o[1] = u.y; o[2] = u.z; o[3] = u.w;
  }
  static __device__ __forceinline__ void st(float* p, const float* o) {
    *reinterpret_cast<float4*>(p) = make_float4(o[0], o[1], o[2], o[3]);
  }
};

struct BnParams {
  const float* gamma; const float* beta;
  float* running_mean; float* running_var; long long* num_batches_tracked;
  float* mean; float* invstd;     // saved for backward, [C]
  float* scale; float* shift;     // y = x*scale + shift, [C]
  float momentum; float eps;
};

// Column reduction skeleton shared by forward-stats and backward-reduce.  Each thread owns one
// vector of channels (cg) and walks rows ty, ty+rows_per_iter, ...; K accumulators per channel.
// CTA partial sums are atomically added into `part` ([K][C] floats); returns true in the last CTA to finish.
// `load(r, c0, regs)` fetches one row's vectors into registers (NT tensors x V floats) and
// `accum(regs, acc)` folds them into the K x V accumulators.  The row loop is unrolled by kU with
// all loads issued before any use so every thread keeps kU*NT 128-bit requests in flight (the
// loop is otherwise pure load->FMA latency).
template <typename T, int K, int NT, typename Load, typename Accum>
__device__ __forceinline__ bool column_reduce(int64_t M, int C, float* __restrict__ part,
                                              unsigned int* __restrict__ ticket, Load load, Accum accum) {
  constexpr int V = Vec<T>::N;
  constexpr int kU = 8;
  extern __shared__ float smem[];   // [ty][K*V*cthreads]
  const int cthreads = C / V;
  const int ty_n = kBT / cthreads;
  const int cg = threadIdx.x % cthreads;
  const int ty = threadIdx.x / cthreads;
  float acc[K][V];
#pragma unroll
  for (int k = 0; k < K; ++k)
#pragma unroll
    for (int j = 0; j < V; ++j) acc[k][j] = 0.f;
  {
    const int64_t stride = (int64_t)gridDim.x * ty_n;
    int64_t r = (int64_t)blockIdx.x * ty_n + ty;
    float regs[kU][NT][V];
    for (; r + (kU - 1) * stride < M; r += kU * stride) {
#pragma unroll
      for (int u = 0; u < kU; ++u) load(r + u * stride, cg * V, regs[u]);
#pragma unroll
      for (int u = 0; u < kU; ++u) accum(regs[u], acc);
    }
    for (; r < M; r += stride) {
      load(r, cg * V, regs[0]);
      accum(regs[0], acc);
    }
  }
  // reduce across ty through shared memory, then ONE atomicAdd per (k, channel) per CTA into the
  // global accumulators `part` ([K][C] floats, zero on entry).  The last CTA to finish (ticket)
  // returns true and owns the finalisation; it also re-zeroes the accumulators for the next call.
  const int width = K * V * cthreads;
#pragma unroll
  for (int k = 0; k < K; ++k)
#pragma unroll
    for (int j = 0; j < V; ++j) smem[ty * width + k * (V * cthreads) + cg * V + j] = acc[k][j];
  __syncthreads();
  for (int i = threadIdx.x; i < width; i += kBT) {
    float sum = 0.f;
    for (int t = 0; t < ty_n; ++t) sum += smem[t * width + i];
    atomicAdd(part + i, sum);                       // i = k*C + c  (V*cthreads == C)
  }
  __shared__ bool last;
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned int t = atomicAdd(ticket, 1u);
    last = (t == gridDim.x - 1);
    if (last) *ticket = 0u;
  }
  __syncthreads();
  if (last) __threadfence();
  return last;
}

template <typename T>
__global__ void __launch_bounds__(kBT)
bn_stats_kernel(const T* __restrict__ x, int64_t M, int C, float* __restrict__ part,
                unsigned int* __restrict__ ticket, BnParams p) {
  constexpr int V = Vec<T>::N;
  // Shifted moments: accumulate sum(x - pivot) and sum((x - pivot)^2) with pivot = running mean
  // (an estimate of the batch mean), so var = E[d^2] - E[d]^2 has no catastrophic cancellation in
  // fp32 even when |mean| >> std.  Every CTA reads the same pivot: running_mean is only rewritten
  // by the last CTA after all others are done.
  const int cg = threadIdx.x % (C / V);
  float pv[V];
#pragma unroll
  for (int j = 0; j < V; ++j) pv[j] = p.running_mean ? p.running_mean[cg * V + j] : 0.f;
  bool last = column_reduce<T, 2, 1>(M, C, part, ticket,
      [&](int64_t r, int c0, float (*regs)[V]) { Vec<T>::ld(x + r * C + c0, regs[0]); },
      [&](float (*regs)[V], float (*acc)[V]) {
#pragma unroll
        for (int j = 0; j < V; ++j) { float d = regs[0][j] - pv[j]; acc[0][j] += d; acc[1][j] = fmaf(d, d, acc[1][j]); }
      });
  if (!last) return;
  const float invM = 1.f / (float)M;
  for (int c = threadIdx.x; c < C; c += kBT) {
    const float s0 = __ldcg(part + c), s1 = __ldcg(part + C + c);
    part[c] = 0.f;
    part[C + c] = 0.f;
    const float pivot = p.running_mean ? p.running_mean[c] : 0.f;
    const float dm = s0 * invM;
    const float mean = pivot + dm;
    const float var = fmaxf(s1 * invM - dm * dm, 0.f);
    const float inv = rsqrtf(var + p.eps);
    p.mean[c] = mean;
    p.invstd[c] = inv;
    const float gm = p.gamma ? p.gamma[c] : 1.f, bta = p.beta ? p.beta[c] : 0.f;
    p.scale[c] = gm * inv;
    p.shift[c] = bta - mean * gm * inv;
    if (p.running_mean) {
      const float unbiased = M > 1 ? var * ((float)M / (float)(M - 1)) : var;
      p.running_mean[c] = (1.f - p.momentum) * pivot + p.momentum * mean;
      p.running_var[c] = (1.f - p.momentum) * p.running_var[c] + p.momentum * unbiased;
    }
  }
  if (threadIdx.x == 0 && p.num_batches_tracked) *p.num_batches_tracked += 1;
}

template <typename T, bool kRelu, bool kAdd = false>
__global__ void __launch_bounds__(kBT)
bn_apply_kernel(const T* __restrict__ x, T* __restrict__ y, int64_t nvec, int C,
                const float* __restrict__ scale, const float* __restrict__ shift, const T* __restrict__ res = nullptr) {
  constexpr int V = Vec<T>::N;
  const int cthreads = C / V;
  const int64_t tid = (int64_t)blockIdx.x * kBT + threadIdx.x;
  const int c0 = (int)(tid % cthreads) * V;   // fixed: the grid stride is a multiple of cthreads
  float sc[V], sh[V];
#pragma unroll
  for (int j = 0; j < V; ++j) { sc[j] = scale[c0 + j]; sh[j] = shift[c0 + j]; }
  const int64_t stride = (int64_t)gridDim.x * kBT;
  for (int64_t i = tid; i < nvec; i += stride) {
    float v[V], rr[V];
    Vec<T>::ld(x + i * V, v);
    if (kAdd) Vec<T>::ld(res + i * V, rr);
#pragma unroll
    for (int j = 0; j < V; ++j) {
      v[j] = fmaf(v[j], sc[j], sh[j]);
      if (kAdd) v[j] += rr[j];
      if (kRelu) v[j] = fmaxf(v[j], 0.f);
    }
    Vec<T>::st(y + i * V, v);
  }
}

template <typename T, bool kMask = false>
__global__ void __launch_bounds__(kBT)
bn_bwd_reduce_kernel(const T* __restrict__ dy, const T* __restrict__ x, const T* __restrict__ yout, int64_t M, int C,
                     const float* __restrict__ mean, const float* __restrict__ invstd,
                     const float* __restrict__ gamma, float* __restrict__ part,
                     unsigned int* __restrict__ ticket, float* __restrict__ dgamma, float* __restrict__ dbeta,
                     float* __restrict__ coef /* [3][C]: a, b, k */) {
  constexpr int V = Vec<T>::N;
  const int cg = threadIdx.x % (C / V);
  float mu[V], is[V];
#pragma unroll
  for (int j = 0; j < V; ++j) { mu[j] = mean[cg * V + j]; is[j] = invstd[cg * V + j]; }
  // kMask: the forward fused a ReLU after the affine (+ residual); its gradient gate is (output > 0)
  bool last = column_reduce<T, 2, kMask ? 3 : 2>(M, C, part, ticket,
      [&](int64_t r, int c0, float (*regs)[V]) {
        Vec<T>::ld(dy + r * C + c0, regs[0]);
        Vec<T>::ld(x + r * C + c0, regs[1]);
        if (kMask) Vec<T>::ld(yout + r * C + c0, regs[2]);
      },
      [&](float (*regs)[V], float (*acc)[V]) {
#pragma unroll
        for (int j = 0; j < V; ++j) {
          const float g = (kMask && !(regs[kMask ? 2 : 0][j] > 0.f)) ? 0.f : regs[0][j];
          acc[0][j] += g;
          acc[1][j] = fmaf(g, (regs[1][j] - mu[j]) * is[j], acc[1][j]);
        }
      });
  if (!last) return;
  const float invM = 1.f / (float)M;
  for (int c = threadIdx.x; c < C; c += kBT) {
    const float s0 = __ldcg(part + c), s1 = __ldcg(part + C + c);
    part[c] = 0.f;
    part[C + c] = 0.f;
    if (dbeta) dbeta[c] = s0;
    if (dgamma) dgamma[c] = s1;
    const float gm = gamma ? gamma[c] : 1.f;
    coef[c] = gm * invstd[c];                  // a
    coef[C + c] = s0 * invM;                   // mean(dy)
    coef[2 * C + c] = s1 * invM * invstd[c];   // mean(dy*xhat)*invstd, multiplies (x-mean)
  }
}

template <typename T, bool kMask = false>
__global__ void __launch_bounds__(kBT)
bn_bwd_apply_kernel(const T* __restrict__ dy, const T* __restrict__ x, T* __restrict__ dx, int64_t nvec, int C,
                    const float* __restrict__ mean, const float* __restrict__ coef, const T* __restrict__ yout = nullptr,
                    T* __restrict__ dres = nullptr) {
  constexpr int V = Vec<T>::N;
  const int cthreads = C / V;
  const int64_t tid = (int64_t)blockIdx.x * kBT + threadIdx.x;
  const int c0 = (int)(tid % cthreads) * V;
  float a[V], b[V], k[V], mu[V];
#pragma unroll
  for (int j = 0; j < V; ++j) { a[j] = coef[c0 + j]; b[j] = coef[C + c0 + j]; k[j] = coef[2 * C + c0 + j]; mu[j] = mean[c0 + j]; }
  const int64_t stride = (int64_t)gridDim.x * kBT;
  for (int64_t i = tid; i < nvec; i += stride) {
    float g[V], v[V];
    Vec<T>::ld(dy + i * V, g);
    Vec<T>::ld(x + i * V, v);
    if (kMask) {
      float yo[V];
      Vec<T>::ld(yout + i * V, yo);
#pragma unroll
      for (int j = 0; j < V; ++j) if (!(yo[j] > 0.f)) g[j] = 0.f;
      if (dres != nullptr) Vec<T>::st(dres + i * V, g);     // gradient of the residual branch = gated dy
    }
#pragma unroll
    for (int j = 0; j < V; ++j) g[j] = a[j] * (g[j] - b[j] - (v[j] - mu[j]) * k[j]);
    Vec<T>::st(dx + i * V, g);
  }
}

inline int reduce_grid(int64_t M, int C, int V) {
  int ty_n = kBT / (C / V);
  int64_t g = (M + (int64_t)ty_n * 8 - 1) / ((int64_t)ty_n * 8);   // >= 8 rows per thread (one unrolled sweep)
  if (g > 4 * kNumSMs) g = 4 * kNumSMs;
  if (g < 1) g = 1;
  return (int)g;
}
inline int apply_grid(int64_t nvec) {
  int64_t g = (nvec + (int64_t)kBT * 4 - 1) / ((int64_t)kBT * 4);
  if (g > 8 * kNumSMs) g = 8 * kNumSMs;
  if (g < 1) g = 1;
  return (int)g;
}
inline bool shape_ok(int C, int V) { return C % V == 0 && (C / V) <= kBT && kBT % (C / V) == 0; }

}  // namespace

#define DETB200_BN_MAX_GRID (4 * 148)
#define DETB200_BN_MAX_C 2048

// Workspace (bytes) for one BatchNorm call: global accumulators [2][C] floats + a ticket (all zero between calls).
DETB200_API int64_t detb200_bn_workspace_bytes(int C) { return (int64_t)2 * C * 4 + 256; }

// dtype 0 fp32 / 1 bf16.  x,y: [M, C] channels-last view.  gamma/beta/running_*/num_batches_tracked nullable.
// mean, invstd, scale, shift: fp32 [C] outputs.  workspace must be zero-initialised once (ticket).
DETB200_API int detb200_bn_fwd_train(const void* x, void* y, int64_t M, int C, const float* gamma, const float* beta,
                                     float* running_mean, float* running_var, long long* num_batches_tracked,
                                     float* mean, float* invstd, float* scale, float* shift, float momentum, float eps,
                                     int relu, void* workspace, int dtype, cudaStream_t stream) {
  const int V = dtype == 1 ? 8 : 4;
  if (M <= 0) return 0;
  if (!shape_ok(C, V)) { detb200_set_error("bn: C=%d unsupported for this dtype", C); return DETB200_ERR_UNSUPPORTED; }
  float* part = (float*)workspace;
  unsigned int* ticket = (unsigned int*)((char*)workspace + (int64_t)2 * C * 4);
  BnParams p{gamma, beta, running_mean, running_var, num_batches_tracked, mean, invstd, scale, shift, momentum, eps};
  int rg = reduce_grid(M, C, V);
  size_t smem = (size_t)kBT * 2 * V * sizeof(float);
  int64_t nvec = M * C / V;
  int ag = apply_grid(nvec);
  if (dtype == 1) {
    bn_stats_kernel<__nv_bfloat16><<<rg, kBT, smem, stream>>>((const __nv_bfloat16*)x, M, C, part, ticket, p);
    if (relu) bn_apply_kernel<__nv_bfloat16, true><<<ag, kBT, 0, stream>>>((const __nv_bfloat16*)x, (__nv_bfloat16*)y, nvec, C, scale, shift);
    else bn_apply_kernel<__nv_bfloat16, false><<<ag, kBT, 0, stream>>>((const __nv_bfloat16*)x, (__nv_bfloat16*)y, nvec, C, scale, shift);
  } else {
    bn_stats_kernel<float><<<rg, kBT, smem, stream>>>((const float*)x, M, C, part, ticket, p);
    if (relu) bn_apply_kernel<float, true><<<ag, kBT, 0, stream>>>((const float*)x, (float*)y, nvec, C, scale, shift);
    else bn_apply_kernel<float, false><<<ag, kBT, 0, stream>>>((const float*)x, (float*)y, nvec, C, scale, shift);
  }
  DETB200_LAUNCH_CHECK();
  return 0;
}

// y = x*scale + shift with caller-provided per-channel fp32 scale/shift (eval mode).
DETB200_API int detb200_bn_apply(const void* x, void* y, int64_t M, int C, const float* scale, const float* shift,
                                 int relu, int dtype, cudaStream_t stream) {
  const int V = dtype == 1 ? 8 : 4;
  if (M <= 0) return 0;
  if (!shape_ok(C, V)) { detb200_set_error("bn: C=%d unsupported for this dtype", C); return DETB200_ERR_UNSUPPORTED; }
  int64_t nvec = M * C / V;
  int ag = apply_grid(nvec);
  if (dtype == 1) {
    if (relu) bn_apply_kernel<__nv_bfloat16, true><<<ag, kBT, 0, stream>>>((const __nv_bfloat16*)x, (__nv_bfloat16*)y, nvec, C, scale, shift);
    else bn_apply_kernel<__nv_bfloat16, false><<<ag, kBT, 0, stream>>>((const __nv_bfloat16*)x, (__nv_bfloat16*)y, nvec, C, scale, shift);
  } else {
    if (relu) bn_apply_kernel<float, true><<<ag, kBT, 0, stream>>>((const float*)x, (float*)y, nvec, C, scale, shift);
    else bn_apply_kernel<float, false><<<ag, kBT, 0, stream>>>((const float*)x, (float*)y, nvec, C, scale, shift);
  }
  DETB200_LAUNCH_CHECK();
  return 0;
}

// dgamma/dbeta: fp32 [C] (nullable).  coef: fp32 [3*C] scratch.  dx may alias dy.
DETB200_API int detb200_bn_bwd(const void* dy, const void* x, void* dx, int64_t M, int C, const float* gamma,
                               const float* mean, const float* invstd, float* dgamma, float* dbeta, float* coef,
                               void* workspace, int dtype, cudaStream_t stream) {
  const int V = dtype == 1 ? 8 : 4;
  if (M <= 0) return 0;
  if (!shape_ok(C, V)) { detb200_set_error("bn: C=%d unsupported for this dtype", C); return DETB200_ERR_UNSUPPORTED; }
  float* part = (float*)workspace;
  unsigned int* ticket = (unsigned int*)((char*)workspace + (int64_t)2 * C * 4);
  int rg = reduce_grid(M, C, V);
  size_t smem = (size_t)kBT * 2 * V * sizeof(float);
  int64_t nvec = M * C / V;
  int ag = apply_grid(nvec);
  if (dtype == 1) {
    bn_bwd_reduce_kernel<__nv_bfloat16><<<rg, kBT, smem, stream>>>((const __nv_bfloat16*)dy, (const __nv_bfloat16*)x, nullptr, M, C, mean, invstd, gamma, part, ticket, dgamma, dbeta, coef);
    bn_bwd_apply_kernel<__nv_bfloat16><<<ag, kBT, 0, stream>>>((const __nv_bfloat16*)dy, (const __nv_bfloat16*)x, (__nv_bfloat16*)dx, nvec, C, mean, coef);
  } else {
    bn_bwd_reduce_kernel<float><<<rg, kBT, smem, stream>>>((const float*)dy, (const float*)x, nullptr, M, C, mean, invstd, gamma, part, ticket, dgamma, dbeta, coef);
    bn_bwd_apply_kernel<float><<<ag, kBT, 0, stream>>>((const float*)dy, (const float*)x, (float*)dx, nvec, C, mean, coef);
  }
  DETB200_LAUNCH_CHECK();
  return 0;
}

// Fused variants.  Forward: y = relu?(bn(x) + residual?).  Backward: `y` (the forward output) gates dy by (y > 0) when
// the forward applied a ReLU; `dres` (nullable) receives the gated dy = gradient of the residual branch.
DETB200_API int detb200_bn_fwd_train_fused(const void* x, const void* residual, void* y, int64_t M, int C, const float* gamma,
                                           const float* beta, float* running_mean, float* running_var, long long* num_batches_tracked,
                                           float* mean, float* invstd, float* scale, float* shift, float momentum, float eps, int relu,
                                           void* workspace, int dtype, cudaStream_t stream) {
  if (residual == nullptr)
    return detb200_bn_fwd_train(x, y, M, C, gamma, beta, running_mean, running_var, num_batches_tracked, mean, invstd, scale, shift, momentum,
                                eps, relu, workspace, dtype, stream);
  const int V = dtype == 1 ? 8 : 4;
  if (M <= 0) return 0;
  if (!shape_ok(C, V)) { detb200_set_error("bn: C=%d unsupported for this dtype", C); return DETB200_ERR_UNSUPPORTED; }
  float* part = (float*)workspace;
  unsigned int* ticket = (unsigned int*)((char*)workspace + (int64_t)2 * C * 4);
  BnParams p{gamma, beta, running_mean, running_var, num_batches_tracked, mean, invstd, scale, shift, momentum, eps};
  const int rg = reduce_grid(M, C, V);
  const size_t smem = (size_t)kBT * 2 * V * sizeof(float);
  const int64_t nvec = M * C / V;
  const int ag = apply_grid(nvec);
  if (dtype == 1) {
    bn_stats_kernel<__nv_bfloat16><<<rg, kBT, smem, stream>>>((const __nv_bfloat16*)x, M, C, part, ticket, p);
    if (relu) bn_apply_kernel<__nv_bfloat16, true, true><<<ag, kBT, 0, stream>>>((const __nv_bfloat16*)x, (__nv_bfloat16*)y, nvec, C, scale, shift, (const __nv_bfloat16*)residual);
    else bn_apply_kernel<__nv_bfloat16, false, true><<<ag, kBT, 0, stream>>>((const __nv_bfloat16*)x, (__nv_bfloat16*)y, nvec, C, scale, shift, (const __nv_bfloat16*)residual);
  } else {
    bn_stats_kernel<float><<<rg, kBT, smem, stream>>>((const float*)x, M, C, part, ticket, p);
    if (relu) bn_apply_kernel<float, true, true><<<ag, kBT, 0, stream>>>((const float*)x, (float*)y, nvec, C, scale, shift, (const float*)residual);
    else bn_apply_kernel<float, false, true><<<ag, kBT, 0, stream>>>((const float*)x, (float*)y, nvec, C, scale, shift, (const float*)residual);
  }
  DETB200_LAUNCH_CHECK();
  return 0;
}

DETB200_API int detb200_bn_bwd_fused(const void* dy, const void* x, const void* y, void* dx, void* dres, int64_t M, int C, const float* gamma,
                                     const float* mean, const float* invstd, float* dgamma, float* dbeta, float* coef, void* workspace,
                                     int dtype, cudaStream_t stream) {
  if (y == nullptr) return detb200_bn_bwd(dy, x, dx, M, C, gamma, mean, invstd, dgamma, dbeta, coef, workspace, dtype, stream);
  const int V = dtype == 1 ? 8 : 4;
  if (M <= 0) return 0;
  if (!shape_ok(C, V)) { detb200_set_error("bn: C=%d unsupported for this dtype", C); return DETB200_ERR_UNSUPPORTED; }
  float* part = (float*)workspace;
  unsigned int* ticket = (unsigned int*)((char*)workspace + (int64_t)2 * C * 4);
  const int rg = reduce_grid(M, C, V);
  const size_t smem = (size_t)kBT * 2 * V * sizeof(float);
  const int64_t nvec = M * C / V;
  const int ag = apply_grid(nvec);
  if (dtype == 1) {
    bn_bwd_reduce_kernel<__nv_bfloat16, true><<<rg, kBT, smem, stream>>>((const __nv_bfloat16*)dy, (const __nv_bfloat16*)x, (const __nv_bfloat16*)y, M, C, mean, invstd, gamma, part, ticket, dgamma, dbeta, coef);
    bn_bwd_apply_kernel<__nv_bfloat16, true><<<ag, kBT, 0, stream>>>((const __nv_bfloat16*)dy, (const __nv_bfloat16*)x, (__nv_bfloat16*)dx, nvec, C, mean, coef, (const __nv_bfloat16*)y, (__nv_bfloat16*)dres);
  } else {
    bn_bwd_reduce_kernel<float, true><<<rg, kBT, smem, stream>>>((const float*)dy, (const float*)x, (const float*)y, M, C, mean, invstd, gamma, part, ticket, dgamma, dbeta, coef);
    bn_bwd_apply_kernel<float, true><<<ag, kBT, 0, stream>>>((const float*)dy, (const float*)x, (float*)dx, nvec, C, mean, coef, (const float*)y, (float*)dres);
  }
  DETB200_LAUNCH_CHECK();
  return 0;
}
