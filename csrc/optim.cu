// Fused flat-buffer optimizers and gradient-norm kernels.
//
// Replaces: torch/apex/DeepSpeed multi-tensor optimizer steps called at reference
// harness/determined/pytorch/_pytorch_context.py:899-907, the per-parameter grad.div_ at
// :805-812, and clip_grad_norm_ at :885-888.
//
// Design: parameters, gradients and optimizer state of one param group live in flat fp32 buffers
// (determined_b200/parallel/flat.py).  One launch updates the whole group: it reads p,g,m,v once
// and writes p,m,v once plus g=0 (folding optimizer.zero_grad() into the same pass), i.e. the HBM
// floor of 32 B/element.  All step-varying scalars (lr, step count, clip/unscale coefficient,
// found-inf skip flag) are read from device memory so the launch is CUDA-graph capturable.
#include "common.cuh"

namespace {

constexpr int kThreads = 256;
constexpr int kVec = 4;

struct Dyn {
  const float* lr;          // required
  const int* step;          // required for adam / sgd first-step logic (already incremented)
  const float* grad_scale;  // optional multiplier applied to g (clip coef x unscale)
  const int* skip;          // optional: non-zero => leave p/state untouched (found inf/nan)
};

__device__ __forceinline__ bool dyn_skip(const Dyn& d) { return d.skip != nullptr && *d.skip != 0; }

// ---------------------------------------------------------------- Adam / AdamW
template <bool kDecoupled>
__global__ void __launch_bounds__(kThreads)
adam_kernel(float* __restrict__ p, float* __restrict__ g, float* __restrict__ m, float* __restrict__ v,
            __nv_bfloat16* __restrict__ p_lowp, int64_t n, Dyn dyn, float beta1, float beta2, float eps,
            float wd, float static_scale, int zero_grad, int maximize) {
  const bool skip = dyn_skip(dyn);
  const float lr = *dyn.lr;
  const int step = *dyn.step;
  const float gs = static_scale * (dyn.grad_scale ? *dyn.grad_scale : 1.f) * (maximize ? -1.f : 1.f);
  // per-thread scalars in double: the fast-math powf has ~1e-4 relative error in 1-beta2^step over the first steps,
  // which made the fused Adam drift from torch.optim.Adam beyond fp32 rounding
  const float bc1 = (float)(1.0 - pow((double)beta1, (double)step));
  const float bc2 = (float)(1.0 - pow((double)beta2, (double)step));
  const float step_size = lr / bc1;
  const float inv_sqrt_bc2 = rsqrtf(bc2);
  const int64_t nvec = n / kVec;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += stride) {
    float4 gv = ldg_stream(reinterpret_cast<const float4*>(g) + i);
    if (!skip) {
      float4 pv = ldg_stream(reinterpret_cast<const float4*>(p) + i);
      float4 mv = ldg_stream(reinterpret_cast<const float4*>(m) + i);
      float4 vv = ldg_stream(reinterpret_cast<const float4*>(v) + i);
      float* pp = &pv.x; float* gg = &gv.x; float* mm = &mv.x; float* vvv = &vv.x;
#pragma unroll
      for (int k = 0; k < kVec; ++k) {
        float gk = gg[k] * gs;
        float pk = pp[k];
        if (kDecoupled) pk *= (1.f - lr * wd); else gk = fmaf(wd, pk, gk);
        float mk = fmaf(beta1, mm[k], (1.f - beta1) * gk);
        float vk = fmaf(beta2, vvv[k], (1.f - beta2) * gk * gk);
        float denom = sqrtf(vk) * inv_sqrt_bc2 + eps;
        pk = pk - step_size * __fdividef(mk, denom);
        pp[k] = pk; mm[k] = mk; vvv[k] = vk;
      }
      stg_stream(reinterpret_cast<float4*>(p) + i, pv);
      stg_stream(reinterpret_cast<float4*>(m) + i, mv);
      stg_stream(reinterpret_cast<float4*>(v) + i, vv);
      if (p_lowp) {
        uint2 o = make_uint2(pack_bf16(pv.x, pv.y), pack_bf16(pv.z, pv.w));
        *(reinterpret_cast<uint2*>(p_lowp) + i) = o;
      }
    }
    if (zero_grad) stg_stream(reinterpret_cast<float4*>(g) + i, make_float4(0.f, 0.f, 0.f, 0.f));
  }
  // tail (n % 4) handled by the first few threads of block 0
  if (blockIdx.x == 0 && threadIdx.x < (n - nvec * kVec)) {
    const int64_t i = nvec * kVec + threadIdx.x;
    float gk = g[i] * gs;
    if (!skip) {
      float pk = p[i];
      if (kDecoupled) pk *= (1.f - lr * wd); else gk = fmaf(wd, pk, gk);
      float mk = fmaf(beta1, m[i], (1.f - beta1) * gk);
      float vk = fmaf(beta2, v[i], (1.f - beta2) * gk * gk);
      pk = pk - step_size * __fdividef(mk, sqrtf(vk) * inv_sqrt_bc2 + eps);
      p[i] = pk; m[i] = mk; v[i] = vk;
      if (p_lowp) p_lowp[i] = __float2bfloat16(pk);
    }
    if (zero_grad) g[i] = 0.f;
  }
}

// ---------------------------------------------------------------- SGD (+momentum, nesterov)
__global__ void __launch_bounds__(kThreads)
sgd_kernel(float* __restrict__ p, float* __restrict__ g, float* __restrict__ buf, int64_t n, Dyn dyn,
           float momentum, float dampening, float wd, int nesterov, float static_scale, int zero_grad,
           int maximize) {
  const bool skip = dyn_skip(dyn);
  const float lr = *dyn.lr;
  const bool first = (dyn.step != nullptr) && (*dyn.step <= 1);
  const float gs = static_scale * (dyn.grad_scale ? *dyn.grad_scale : 1.f) * (maximize ? -1.f : 1.f);
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const int64_t nvec = n / kVec;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += stride) {
    float4 gv = ldg_stream(reinterpret_cast<const float4*>(g) + i);
    if (!skip) {
      float4 pv = ldg_stream(reinterpret_cast<const float4*>(p) + i);
      float4 bv = make_float4(0.f, 0.f, 0.f, 0.f);
      if (buf && !first) bv = ldg_stream(reinterpret_cast<const float4*>(buf) + i);
      float* pp = &pv.x; float* gg = &gv.x; float* bb = &bv.x;
#pragma unroll
      for (int k = 0; k < kVec; ++k) {
        float gk = fmaf(wd, pp[k], gg[k] * gs);
        if (buf) {
          float bk = first ? gk : fmaf(momentum, bb[k], (1.f - dampening) * gk);
          bb[k] = bk;
          gk = nesterov ? fmaf(momentum, bk, gk) : bk;
        }
        pp[k] = fmaf(-lr, gk, pp[k]);
      }
      stg_stream(reinterpret_cast<float4*>(p) + i, pv);
      if (buf) stg_stream(reinterpret_cast<float4*>(buf) + i, bv);
    }
    if (zero_grad) stg_stream(reinterpret_cast<float4*>(g) + i, make_float4(0.f, 0.f, 0.f, 0.f));
  }
  if (blockIdx.x == 0 && threadIdx.x < (n - nvec * kVec)) {
    const int64_t i = nvec * kVec + threadIdx.x;
    if (!skip) {
      float gk = fmaf(wd, p[i], g[i] * gs);
      if (buf) {
        float bk = first ? gk : fmaf(momentum, buf[i], (1.f - dampening) * gk);
        buf[i] = bk;
        gk = nesterov ? fmaf(momentum, bk, gk) : bk;
      }
      p[i] = fmaf(-lr, gk, p[i]);
    }
    if (zero_grad) g[i] = 0.f;
  }
}

// ---------------------------------------------------------------- Adadelta (MNIST tutorial optimizer)
__global__ void __launch_bounds__(kThreads)
adadelta_kernel(float* __restrict__ p, float* __restrict__ g, float* __restrict__ sq, float* __restrict__ acc,
                int64_t n, Dyn dyn, float rho, float eps, float wd, float static_scale, int zero_grad,
                int maximize) {
  const bool skip = dyn_skip(dyn);
  const float lr = *dyn.lr;
  const float gs = static_scale * (dyn.grad_scale ? *dyn.grad_scale : 1.f) * (maximize ? -1.f : 1.f);
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    float gk = g[i] * gs;
    if (!skip) {
      float pk = p[i];
      gk = fmaf(wd, pk, gk);
      float s = fmaf(rho, sq[i], (1.f - rho) * gk * gk);
      float a = acc[i];
      float delta = sqrtf(a + eps) * rsqrtf(s + eps) * gk;
      a = fmaf(rho, a, (1.f - rho) * delta * delta);
      p[i] = fmaf(-lr, delta, pk);
      sq[i] = s; acc[i] = a;
    }
    if (zero_grad) g[i] = 0.f;
  }
}

// ---------------------------------------------------------------- step counter tick
__global__ void tick_kernel(int* step, const int* skip) {
  if (threadIdx.x == 0 && blockIdx.x == 0) {
    if (skip == nullptr || *skip == 0) *step += 1;
  }
}

// ---------------------------------------------------------------- sum of squares (+ non-finite flag)
// Two-stage reduction.  Stage 1: each block writes one partial into `partials` (scratch memory,
// >= gridDim.x floats).  Stage 2: the last block to arrive (ticket counter in scratch) sums the
// partials in a fixed order -> deterministic result, no float atomics.
__global__ void __launch_bounds__(kThreads)
sumsq_kernel(const float* __restrict__ x, int64_t n, float* __restrict__ partials,
             unsigned int* __restrict__ ticket, float* __restrict__ out, int* __restrict__ nonfinite,
             int accumulate) {
  __shared__ float red[32];
  __shared__ bool is_last;
  float acc = 0.f;
  bool bad = false;
  const int64_t nvec = n / kVec;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += stride) {
    float4 v = ldg_stream(reinterpret_cast<const float4*>(x) + i);
    acc = fmaf(v.x, v.x, acc); acc = fmaf(v.y, v.y, acc);
    acc = fmaf(v.z, v.z, acc); acc = fmaf(v.w, v.w, acc);
  }
  if (blockIdx.x == 0 && threadIdx.x < (n - nvec * kVec)) {
    float v = x[nvec * kVec + threadIdx.x];
    acc = fmaf(v, v, acc);
  }
  bad = !isfinite(acc);
  float tot = block_sum(acc, red);
  if (threadIdx.x == 0) {
    partials[blockIdx.x] = tot;
    __threadfence();
    unsigned int t = atomicAdd(ticket, 1u);
    is_last = (t == gridDim.x - 1);
  }
  if (nonfinite != nullptr && __syncthreads_or(bad) && threadIdx.x == 0) atomicOr(nonfinite, 1);
  __syncthreads();
  if (is_last) {
    __threadfence();
    float s = 0.f;
    for (int i = threadIdx.x; i < gridDim.x; i += blockDim.x) s += __ldcg(partials + i);
    s = block_sum(s, red);
    if (threadIdx.x == 0) {
      *out = accumulate ? (*out + s) : s;
      *ticket = 0u;  // re-arm for the next launch
    }
  }
}

// coef = extra_scale * min(1, max_norm / (sqrt(sum terms) + 1e-6)); norm_out = sqrt(sum terms)
__global__ void clip_coef_kernel(const float* __restrict__ terms, int nterms, float max_norm, float extra_scale,
                                 float* __restrict__ coef_out, float* __restrict__ norm_out) {
  if (threadIdx.x == 0 && blockIdx.x == 0) {
    float s = 0.f;
    for (int i = 0; i < nterms; ++i) s += terms[i];
    float norm = sqrtf(s);
    float c = (max_norm > 0.f) ? fminf(1.f, max_norm / (norm + 1e-6f)) : 1.f;
    if (!isfinite(norm)) c = 1.f;  // torch semantics with error_if_nonfinite=False: coef becomes nan; we keep grads as is
    *coef_out = c * extra_scale;
    if (norm_out) *norm_out = norm;
  }
}

__global__ void __launch_bounds__(kThreads)
scale_kernel(float* __restrict__ x, int64_t n, const float* __restrict__ dyn_scale, float static_scale) {
  const float s = static_scale * (dyn_scale ? *dyn_scale : 1.f);
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) x[i] *= s;
}

inline int optim_grid(int64_t n) { return grid_for(n, kThreads, kVec * 4, kNumSMs * 8); }

}  // namespace

DETB200_API int detb200_optim_tick(int* step, const int* skip, cudaStream_t stream) {
  tick_kernel<<<1, 32, 0, stream>>>(step, skip);
  DETB200_LAUNCH_CHECK();
  return 0;
}

DETB200_API int detb200_adam_flat(float* p, float* g, float* m, float* v, void* p_lowp_bf16, int64_t n,
                                  const float* lr, const int* step, const float* grad_scale, const int* skip,
                                  float beta1, float beta2, float eps, float weight_decay, int decoupled_wd,
                                  float static_grad_scale, int zero_grad, int maximize, cudaStream_t stream) {
  if (n <= 0) return 0;
  if (!p || !g || !m || !v || !lr || !step) { detb200_set_error("adam_flat: null pointer"); return DETB200_ERR_BAD_ARG; }
  if ((((uintptr_t)p | (uintptr_t)g | (uintptr_t)m | (uintptr_t)v) & 15) != 0) {
    detb200_set_error("adam_flat: buffers must be 16-byte aligned"); return DETB200_ERR_BAD_ARG;
  }
  Dyn d{lr, step, grad_scale, skip};
  int grid = optim_grid(n);
  if (decoupled_wd)
    adam_kernel<true><<<grid, kThreads, 0, stream>>>(p, g, m, v, (__nv_bfloat16*)p_lowp_bf16, n, d, beta1, beta2, eps,
                                                     weight_decay, static_grad_scale, zero_grad, maximize);
  else
    adam_kernel<false><<<grid, kThreads, 0, stream>>>(p, g, m, v, (__nv_bfloat16*)p_lowp_bf16, n, d, beta1, beta2, eps,
                                                      weight_decay, static_grad_scale, zero_grad, maximize);
  DETB200_LAUNCH_CHECK();
  return 0;
}

DETB200_API int detb200_sgd_flat(float* p, float* g, float* momentum_buf, int64_t n, const float* lr,
                                 const int* step, const float* grad_scale, const int* skip, float momentum,
                                 float dampening, float weight_decay, int nesterov, float static_grad_scale,
                                 int zero_grad, int maximize, cudaStream_t stream) {
  if (n <= 0) return 0;
  if (!p || !g || !lr) { detb200_set_error("sgd_flat: null pointer"); return DETB200_ERR_BAD_ARG; }
  if (momentum_buf && !step) { detb200_set_error("sgd_flat: momentum needs a step counter"); return DETB200_ERR_BAD_ARG; }
  Dyn d{lr, step, grad_scale, skip};
  sgd_kernel<<<optim_grid(n), kThreads, 0, stream>>>(p, g, momentum_buf, n, d, momentum, dampening, weight_decay,
                                                     nesterov, static_grad_scale, zero_grad, maximize);
  DETB200_LAUNCH_CHECK();
  return 0;
}

DETB200_API int detb200_adadelta_flat(float* p, float* g, float* square_avg, float* acc_delta, int64_t n,
                                      const float* lr, const float* grad_scale, const int* skip, float rho, float eps,
                                      float weight_decay, float static_grad_scale, int zero_grad, int maximize,
                                      cudaStream_t stream) {
  if (n <= 0) return 0;
  if (!p || !g || !square_avg || !acc_delta || !lr) { detb200_set_error("adadelta_flat: null pointer"); return DETB200_ERR_BAD_ARG; }
  Dyn d{lr, nullptr, grad_scale, skip};
  adadelta_kernel<<<grid_for(n, kThreads, 4, kNumSMs * 8), kThreads, 0, stream>>>(
      p, g, square_avg, acc_delta, n, d, rho, eps, weight_decay, static_grad_scale, zero_grad, maximize);
  DETB200_LAUNCH_CHECK();
  return 0;
}

// Scratch layout expected by detb200_sumsq: `scratch` points at (DETB200_SUMSQ_MAX_BLOCKS floats of partials)
// followed by one uint32 ticket that must be zero before the first call (the kernel re-arms it).
#define DETB200_SUMSQ_MAX_BLOCKS 1184

DETB200_API int64_t detb200_sumsq_scratch_bytes() { return (int64_t)(DETB200_SUMSQ_MAX_BLOCKS + 1) * 4; }

DETB200_API int detb200_sumsq(const float* x, int64_t n, void* scratch, float* out, int* nonfinite_flag,
                              int accumulate, cudaStream_t stream) {
  if (!x || !scratch || !out || n < 0) { detb200_set_error("sumsq: bad argument"); return DETB200_ERR_BAD_ARG; }
  if (((uintptr_t)x & 15) != 0) { detb200_set_error("sumsq: x must be 16-byte aligned"); return DETB200_ERR_BAD_ARG; }
  float* partials = (float*)scratch;
  unsigned int* ticket = (unsigned int*)(partials + DETB200_SUMSQ_MAX_BLOCKS);
  int grid = grid_for(n, kThreads, kVec * 4, DETB200_SUMSQ_MAX_BLOCKS);
  sumsq_kernel<<<grid, kThreads, 0, stream>>>(x, n, partials, ticket, out, nonfinite_flag, accumulate);
  DETB200_LAUNCH_CHECK();
  return 0;
}

DETB200_API int detb200_clip_coef(const float* sumsq_terms, int nterms, float max_norm, float extra_scale,
                                  float* coef_out, float* norm_out, cudaStream_t stream) {
  if (!sumsq_terms || !coef_out || nterms < 1) { detb200_set_error("clip_coef: bad argument"); return DETB200_ERR_BAD_ARG; }
  clip_coef_kernel<<<1, 32, 0, stream>>>(sumsq_terms, nterms, max_norm, extra_scale, coef_out, norm_out);
  DETB200_LAUNCH_CHECK();
  return 0;
}

DETB200_API int detb200_scale_flat(float* x, int64_t n, const float* dyn_scale, float static_scale, cudaStream_t stream) {
  if (n <= 0) return 0;
  scale_kernel<<<grid_for(n, kThreads, 4, kNumSMs * 8), kThreads, 0, stream>>>(x, n, dyn_scale, static_scale);
  DETB200_LAUNCH_CHECK();
  return 0;
}

// ---------------------------------------------------------------------------------------------
// detb200_sumsq_noalloc — the same sum-of-squares reduction for callers that cannot obtain device
// memory of their own for the scratch area.
//
// Who that is: a process that has already handed the whole HBM to a fixed pool or caching
// allocator (a fresh cudaMalloc would fail or fragment the pool); a plain C caller that links this
// library without any allocator of its own; and code on a latency-critical path that must not
// pay cudaMalloc/cudaFree's device-wide synchronisation (it would stall the communication stream
// that is overlapping backward).  For them the partial-sum area and the ticket live in a
// file-scope __device__ array below, so the call needs nothing beyond its input and outputs.
//
// Costs and rules that come with that choice:
//   * every GPU whose context loads this library carries the array (DETB200_SUMSQ_NOALLOC_SCRATCH_BYTES
//     = 4864 bytes) until the process exits;
//   * n must be <= DETB200_SUMSQ_NOALLOC_MAX_ELEMS (2^30 floats, 4 GiB): the array holds a fixed
//     1184 fp32 partials and the variant is only validated up to that length; larger inputs are
//     refused with DETB200_ERR_TOO_LARGE rather than silently losing accuracy;
//   * the array is shared by every call on that device and keeps whatever the previous call left,
//     so the ticket word is cleared on the caller's stream before each launch, and two calls that
//     can overlap on the GPU (different streams, no ordering between them) race on it — order them;
//   * the first call on a device looks the array's device address up (cudaGetSymbolAddress) with
//     that device current; make that first call outside of any stream capture.
// It never calls cudaMalloc, cudaFree or any allocating library routine.
// ---------------------------------------------------------------------------------------------
#define DETB200_SUMSQ_NOALLOC_MAX_ELEMS (1ll << 30)
#define DETB200_SUMSQ_NOALLOC_SCRATCH_BYTES (((DETB200_SUMSQ_MAX_BLOCKS + 1) * 4 + 255) / 256 * 256)
#define DETB200_NOALLOC_MAX_DEVICES 64

static __device__ __align__(256) unsigned char g_sumsq_noalloc_scratch[DETB200_SUMSQ_NOALLOC_SCRATCH_BYTES];

#include <mutex>

namespace {
std::mutex g_noalloc_mu;
void* g_noalloc_ptr[DETB200_NOALLOC_MAX_DEVICES] = {nullptr};

// In host code `g_sumsq_noalloc_scratch` is only a handle for the runtime, not an address the GPU
// can use; the real device address is per device and must be asked for with that device current.
int noalloc_scratch_for_current_device(void** out) {
  int dev = -1;
  DETB200_CHECK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= DETB200_NOALLOC_MAX_DEVICES) {
    detb200_set_error("sumsq_noalloc: device index %d outside [0,%d)", dev, DETB200_NOALLOC_MAX_DEVICES);
    return DETB200_ERR_UNSUPPORTED;
  }
  std::lock_guard<std::mutex> lock(g_noalloc_mu);
  if (g_noalloc_ptr[dev] == nullptr) {
    void* p = nullptr;
    DETB200_CHECK(cudaGetSymbolAddress(&p, g_sumsq_noalloc_scratch));
    if (p == nullptr) { detb200_set_error("sumsq_noalloc: cudaGetSymbolAddress returned NULL"); return DETB200_ERR_UNSUPPORTED; }
    g_noalloc_ptr[dev] = p;
  }
  *out = g_noalloc_ptr[dev];
  return 0;
}
}  // namespace

DETB200_API int64_t detb200_sumsq_noalloc_max_elems() { return DETB200_SUMSQ_NOALLOC_MAX_ELEMS; }

DETB200_API int detb200_sumsq_noalloc(const float* x, int64_t n, float* out, int* nonfinite_flag, int accumulate,
                                      cudaStream_t stream) {
  if (!x || !out || n < 0) { detb200_set_error("sumsq_noalloc: bad argument"); return DETB200_ERR_BAD_ARG; }
  if (((uintptr_t)x & 15) != 0) { detb200_set_error("sumsq_noalloc: x must be 16-byte aligned"); return DETB200_ERR_BAD_ARG; }
  if (n > DETB200_SUMSQ_NOALLOC_MAX_ELEMS) {
    detb200_set_error("sumsq_noalloc: n=%lld exceeds the built-in scratch limit of %lld elements; use detb200_sumsq with "
                      "caller-provided scratch", (long long)n, (long long)DETB200_SUMSQ_NOALLOC_MAX_ELEMS);
    return DETB200_ERR_TOO_LARGE;
  }
  // The scratch address belongs to the CURRENT device; a stream of another device would make the launch fail with a
  // generic error (or, with peer access enabled, reduce into the wrong GPU's array).  Refuse that up front.
  {
    int cur = -1, sdev = -1;
    if (cudaGetDevice(&cur) == cudaSuccess && cudaStreamGetDevice(stream, &sdev) == cudaSuccess) {
      if (sdev != cur) {
        detb200_set_error("sumsq_noalloc: stream belongs to device %d but device %d is current; make the stream's device "
                          "current before the call", sdev, cur);
        return DETB200_ERR_BAD_ARG;
      }
    } else {
      cudaGetLastError();  // the query is advisory: fall through to the launch, which reports real problems
    }
  }
  void* scratch = nullptr;
  int rc = noalloc_scratch_for_current_device(&scratch);
  if (rc) return rc;
  float* partials = (float*)scratch;
  unsigned int* ticket = (unsigned int*)(partials + DETB200_SUMSQ_MAX_BLOCKS);
  // the array remembers the previous call: the ticket must read 0 when the kernel starts
  DETB200_CHECK(cudaMemsetAsync(ticket, 0, sizeof(unsigned int), stream));
  int grid = grid_for(n, kThreads, kVec * 4, DETB200_SUMSQ_MAX_BLOCKS);
  sumsq_kernel<<<grid, kThreads, 0, stream>>>(x, n, partials, ticket, out, nonfinite_flag, accumulate);
  DETB200_LAUNCH_CHECK();
  return 0;
}
