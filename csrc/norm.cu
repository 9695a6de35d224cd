// LayerNorm / RMSNorm forward + backward, bf16 or fp32 I/O with fp32 statistics.
//
// Replaces: ATen / apex FusedLayerNorm / Megatron fused LN used by the reference's GPT-2 and
// GPT-NeoX examples (examples/deepspeed/gpt_neox/Dockerfile:11 builds Megatron fused_kernels).
//
// One CTA (256 threads) per row at a time; each thread owns P column *pairs* that stay in
// registers from the load to the store, so a row is read from HBM exactly once in the forward and
// x,dy exactly once in the backward.  dgamma/dbeta are accumulated per thread across the CTA's
// rows, written as per-CTA partials to a caller-provided workspace and reduced by a second kernel.
#include "common.cuh"

namespace {

constexpr int kT = 256;

template <typename T> struct Pair;
template <> struct Pair<float> {
  static __device__ __forceinline__ float2 ld(const float* p, int64_t pair) { return reinterpret_cast<const float2*>(p)[pair]; }
  static __device__ __forceinline__ void st(float* p, int64_t pair, float2 v) { reinterpret_cast<float2*>(p)[pair] = v; }
};
template <> struct Pair<__nv_bfloat16> {
  static __device__ __forceinline__ float2 ld(const __nv_bfloat16* p, int64_t pair) {
    uint32_t u = reinterpret_cast<const uint32_t*>(p)[pair];
    return make_float2(bf16lo(u), bf16hi(u));
  }
  static __device__ __forceinline__ void st(__nv_bfloat16* p, int64_t pair, float2 v) {
    reinterpret_cast<uint32_t*>(p)[pair] = pack_bf16(v.x, v.y);
  }
};

template <typename T, int P, bool kRms>
__global__ void __launch_bounds__(kT)
norm_fwd_kernel(const T* __restrict__ x, const T* __restrict__ gamma, const T* __restrict__ beta,
                T* __restrict__ y, float* __restrict__ mean_out, float* __restrict__ rstd_out,
                int64_t rows, int cols, float eps) {
  __shared__ float red[32];
  const int npairs = cols >> 1;
  float2 gmm[P], bta[P];
#pragma unroll
  for (int k = 0; k < P; ++k) {
    int c = threadIdx.x + k * kT;
    gmm[k] = (c < npairs) ? Pair<T>::ld(gamma, c) : make_float2(0.f, 0.f);
    bta[k] = (!kRms && beta != nullptr && c < npairs) ? Pair<T>::ld(beta, c) : make_float2(0.f, 0.f);
  }
  const float inv_n = 1.f / (float)cols;
  for (int64_t row = blockIdx.x; row < rows; row += gridDim.x) {
    const T* xr = x + row * (int64_t)cols;
    float2 v[P];
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < P; ++k) {
      int c = threadIdx.x + k * kT;
      v[k] = (c < npairs) ? Pair<T>::ld(xr, c) : make_float2(0.f, 0.f);
      s += kRms ? (v[k].x * v[k].x + v[k].y * v[k].y) : (v[k].x + v[k].y);
    }
    float mean = 0.f, rstd;
    if (kRms) {
      rstd = rsqrtf(block_sum(s, red) * inv_n + eps);
    } else {
      mean = block_sum(s, red) * inv_n;
      float q = 0.f;
#pragma unroll
      for (int k = 0; k < P; ++k) {
        int c = threadIdx.x + k * kT;
        if (c < npairs) { float a = v[k].x - mean, b = v[k].y - mean; q += a * a + b * b; }
      }
      rstd = rsqrtf(block_sum(q, red) * inv_n + eps);
    }
    if (threadIdx.x == 0) {
      if (mean_out && !kRms) mean_out[row] = mean;
      if (rstd_out) rstd_out[row] = rstd;
    }
    T* yr = y + row * (int64_t)cols;
#pragma unroll
    for (int k = 0; k < P; ++k) {
      int c = threadIdx.x + k * kT;
      if (c < npairs) {
        float2 o;
        o.x = (v[k].x - mean) * rstd * gmm[k].x + bta[k].x;
        o.y = (v[k].y - mean) * rstd * gmm[k].y + bta[k].y;
        Pair<T>::st(yr, c, o);
      }
    }
  }
}

template <typename T, int P, bool kRms>
__global__ void __launch_bounds__(kT)
norm_bwd_kernel(const T* __restrict__ dy, const T* __restrict__ x, const T* __restrict__ gamma,
                const float* __restrict__ mean_in, const float* __restrict__ rstd_in, T* __restrict__ dx,
                float* __restrict__ dgamma_part, float* __restrict__ dbeta_part, int64_t rows, int cols) {
  __shared__ float red[32];
  const int npairs = cols >> 1;
  float2 gmm[P], dg[P], db[P];
#pragma unroll
  for (int k = 0; k < P; ++k) {
    int c = threadIdx.x + k * kT;
    gmm[k] = (c < npairs) ? Pair<T>::ld(gamma, c) : make_float2(0.f, 0.f);
    dg[k] = make_float2(0.f, 0.f);
    db[k] = make_float2(0.f, 0.f);
  }
  const float inv_n = 1.f / (float)cols;
  for (int64_t row = blockIdx.x; row < rows; row += gridDim.x) {
    const float mean = kRms ? 0.f : mean_in[row];
    const float rstd = rstd_in[row];
    const T* xr = x + row * (int64_t)cols;
    const T* gr = dy + row * (int64_t)cols;
    float2 xh[P], g[P];
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int k = 0; k < P; ++k) {
      int c = threadIdx.x + k * kT;
      if (c < npairs) {
        float2 xv = Pair<T>::ld(xr, c), gv = Pair<T>::ld(gr, c);
        xh[k] = make_float2((xv.x - mean) * rstd, (xv.y - mean) * rstd);
        dg[k].x += gv.x * xh[k].x; dg[k].y += gv.y * xh[k].y;
        db[k].x += gv.x; db[k].y += gv.y;
        g[k] = make_float2(gv.x * gmm[k].x, gv.y * gmm[k].y);
        s1 += g[k].x + g[k].y;
        s2 += g[k].x * xh[k].x + g[k].y * xh[k].y;
      } else {
        xh[k] = g[k] = make_float2(0.f, 0.f);
      }
    }
    const float m1 = kRms ? 0.f : block_sum(s1, red) * inv_n;
    const float m2 = block_sum(s2, red) * inv_n;
    T* dr = dx + row * (int64_t)cols;
#pragma unroll
    for (int k = 0; k < P; ++k) {
      int c = threadIdx.x + k * kT;
      if (c < npairs) {
        float2 o;
        o.x = (g[k].x - m1 - xh[k].x * m2) * rstd;
        o.y = (g[k].y - m1 - xh[k].y * m2) * rstd;
        Pair<T>::st(dr, c, o);
      }
    }
  }
  float* gp = dgamma_part + (int64_t)blockIdx.x * cols;
  float* bp = (kRms || dbeta_part == nullptr) ? nullptr : dbeta_part + (int64_t)blockIdx.x * cols;
#pragma unroll
  for (int k = 0; k < P; ++k) {
    int c = threadIdx.x + k * kT;
    if (c < npairs) {
      reinterpret_cast<float2*>(gp)[c] = dg[k];
      if (bp) reinterpret_cast<float2*>(bp)[c] = db[k];
    }
  }
}

// out[c] = sum_b part[b, c]  (written in T)
template <typename T>
__global__ void colsum_kernel(const float* __restrict__ part, int nblocks, int cols, T* __restrict__ out) {
  int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= cols) return;
  float s = 0.f;
  for (int b = 0; b < nblocks; ++b) s += part[(int64_t)b * cols + c];
  out[c] = (T)s;
}
template <>
__global__ void colsum_kernel<__nv_bfloat16>(const float* __restrict__ part, int nblocks, int cols,
                                             __nv_bfloat16* __restrict__ out) {
  int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= cols) return;
  float s = 0.f;
  for (int b = 0; b < nblocks; ++b) s += part[(int64_t)b * cols + c];
  out[c] = __float2bfloat16(s);
}

inline int pick_p(int cols) {
  int npairs = cols / 2;
  int p = (npairs + kT - 1) / kT;
  if (p <= 1) return 1;
  if (p <= 2) return 2;
  if (p <= 4) return 4;
  if (p <= 8) return 8;
  if (p <= 16) return 16;
  return 0;
}

template <typename T, bool kRms>
int launch_fwd(const void* x, const void* gamma, const void* beta, void* y, float* mean, float* rstd, int64_t rows,
               int cols, float eps, cudaStream_t st) {
  int grid = (int)(rows < (int64_t)kNumSMs * 8 ? rows : (int64_t)kNumSMs * 8);
#define FWD(PV) norm_fwd_kernel<T, PV, kRms><<<grid, kT, 0, st>>>((const T*)x, (const T*)gamma, (const T*)beta, (T*)y, mean, rstd, rows, cols, eps)
  switch (pick_p(cols)) {
    case 1: FWD(1); break; case 2: FWD(2); break; case 4: FWD(4); break;
    case 8: FWD(8); break; case 16: FWD(16); break;
    default: detb200_set_error("norm: cols=%d unsupported (max 8192)", cols); return DETB200_ERR_UNSUPPORTED;
  }
#undef FWD
  return 0;
}

template <typename T, bool kRms>
int launch_bwd(const void* dy, const void* x, const void* gamma, const float* mean, const float* rstd, void* dx,
               void* dgamma, void* dbeta, float* ws, int grid, int64_t rows, int cols, cudaStream_t st) {
  float* gpart = ws;
  float* bpart = kRms ? nullptr : ws + (int64_t)grid * cols;
#define BWD(PV) norm_bwd_kernel<T, PV, kRms><<<grid, kT, 0, st>>>((const T*)dy, (const T*)x, (const T*)gamma, mean, rstd, (T*)dx, gpart, bpart, rows, cols)
  switch (pick_p(cols)) {
    case 1: BWD(1); break; case 2: BWD(2); break; case 4: BWD(4); break;
    case 8: BWD(8); break; case 16: BWD(16); break;
    default: detb200_set_error("norm: cols=%d unsupported (max 8192)", cols); return DETB200_ERR_UNSUPPORTED;
  }
#undef BWD
  int cb = (cols + 255) / 256;
  colsum_kernel<T><<<cb, 256, 0, st>>>(gpart, grid, cols, (T*)dgamma);
  if (!kRms && dbeta) colsum_kernel<T><<<cb, 256, 0, st>>>(bpart, grid, cols, (T*)dbeta);
  return 0;
}

inline int bwd_grid(int64_t rows) {
  int64_t g = rows < (int64_t)kNumSMs * 2 ? rows : (int64_t)kNumSMs * 2;
  return (int)(g < 1 ? 1 : g);
}

}  // namespace

// dtype: 0 = fp32, 1 = bf16.  rms: 0 = LayerNorm, 1 = RMSNorm (beta/mean ignored).
DETB200_API int detb200_norm_fwd(const void* x, const void* gamma, const void* beta, void* y, float* mean,
                                 float* rstd, int64_t rows, int cols, float eps, int dtype, int rms,
                                 cudaStream_t stream) {
  if (rows <= 0) return 0;
  if (cols % 2 != 0 || cols < 2) { detb200_set_error("norm: cols must be even"); return DETB200_ERR_UNSUPPORTED; }
  int rc;
  if (dtype == 1) rc = rms ? launch_fwd<__nv_bfloat16, true>(x, gamma, beta, y, mean, rstd, rows, cols, eps, stream)
                           : launch_fwd<__nv_bfloat16, false>(x, gamma, beta, y, mean, rstd, rows, cols, eps, stream);
  else rc = rms ? launch_fwd<float, true>(x, gamma, beta, y, mean, rstd, rows, cols, eps, stream)
                : launch_fwd<float, false>(x, gamma, beta, y, mean, rstd, rows, cols, eps, stream);
  if (rc) return rc;
  DETB200_LAUNCH_CHECK();
  return 0;
}

// Workspace (fp32 elements) the backward needs: per-CTA partial dgamma (+ dbeta).
DETB200_API int64_t detb200_norm_bwd_workspace_elems(int64_t rows, int cols, int rms) {
  return (int64_t)bwd_grid(rows) * cols * (rms ? 1 : 2);
}

DETB200_API int detb200_norm_bwd(const void* dy, const void* x, const void* gamma, const float* mean,
                                 const float* rstd, void* dx, void* dgamma, void* dbeta, float* workspace,
                                 int64_t rows, int cols, int dtype, int rms, cudaStream_t stream) {
  if (rows <= 0) return 0;
  if (cols % 2 != 0 || cols < 2) { detb200_set_error("norm: cols must be even"); return DETB200_ERR_UNSUPPORTED; }
  if (!workspace) { detb200_set_error("norm_bwd: workspace required"); return DETB200_ERR_BAD_ARG; }
  int grid = bwd_grid(rows);
  int rc;
  if (dtype == 1) rc = rms ? launch_bwd<__nv_bfloat16, true>(dy, x, gamma, mean, rstd, dx, dgamma, dbeta, workspace, grid, rows, cols, stream)
                           : launch_bwd<__nv_bfloat16, false>(dy, x, gamma, mean, rstd, dx, dgamma, dbeta, workspace, grid, rows, cols, stream);
  else rc = rms ? launch_bwd<float, true>(dy, x, gamma, mean, rstd, dx, dgamma, dbeta, workspace, grid, rows, cols, stream)
                : launch_bwd<float, false>(dy, x, gamma, mean, rstd, dx, dgamma, dbeta, workspace, grid, rows, cols, stream);
  if (rc) return rc;
  DETB200_LAUNCH_CHECK();
  return 0;
}
