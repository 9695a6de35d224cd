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

// ---- v2: 16-byte vector kernels, one warp (or a team of 2/4/8 warps) per row ----------------------------
// A thread keeps up to 4 vectors of 8 elements of its row in registers; statistics are shifted moments
// about the row's first element, so one shuffle/smem reduction round gives mean and variance.
template <typename T> struct Vec8;
template <> struct Vec8<__nv_bfloat16> {
  struct Raw { uint4 u; };
  static __device__ __forceinline__ Raw ld_raw(const __nv_bfloat16* p) { Raw r; r.u = ldg_stream_u4(reinterpret_cast<const uint4*>(p)); return r; }
  static __device__ __forceinline__ void unpack(const Raw& r, float* v) {
    v[0] = bf16lo(r.u.x); v[1] = bf16hi(r.u.x); v[2] = bf16lo(r.u.y); v[3] = bf16hi(r.u.y);
    v[4] = bf16lo(r.u.z); v[5] = bf16hi(r.u.z); v[6] = bf16lo(r.u.w); v[7] = bf16hi(r.u.w);
  }
  static __device__ __forceinline__ void ld(const __nv_bfloat16* p, float* v) {
    const uint4 u = ldg_stream_u4(reinterpret_cast<const uint4*>(p));
    v[0] = bf16lo(u.x); v[1] = bf16hi(u.x); v[2] = bf16lo(u.y); v[3] = bf16hi(u.y);
    v[4] = bf16lo(u.z); v[5] = bf16hi(u.z); v[6] = bf16lo(u.w); v[7] = bf16hi(u.w);
  }
  static __device__ __forceinline__ void ld_cached(const __nv_bfloat16* p, float* v) {
    const uint4 u = __ldg(reinterpret_cast<const uint4*>(p));
    v[0] = bf16lo(u.x); v[1] = bf16hi(u.x); v[2] = bf16lo(u.y); v[3] = bf16hi(u.y);
    v[4] = bf16lo(u.z); v[5] = bf16hi(u.z); v[6] = bf16lo(u.w); v[7] = bf16hi(u.w);
  }
  static __device__ __forceinline__ void st(__nv_bfloat16* p, const float* v) {
    *reinterpret_cast<uint4*>(p) = make_uint4(pack_bf16(v[0], v[1]), pack_bf16(v[2], v[3]), pack_bf16(v[4], v[5]), pack_bf16(v[6], v[7]));
  }
};
template <> struct Vec8<float> {
  struct Raw { float4 a, b; };
  static __device__ __forceinline__ Raw ld_raw(const float* p) {
    Raw r; r.a = ldg_stream(reinterpret_cast<const float4*>(p)); r.b = ldg_stream(reinterpret_cast<const float4*>(p) + 1); return r;
  }
  static __device__ __forceinline__ void unpack(const Raw& r, float* v) {
    v[0] = r.a.x; v[1] = r.a.y; v[2] = r.a.z; v[3] = r.a.w; v[4] = r.b.x; v[5] = r.b.y; v[6] = r.b.z; v[7] = r.b.w;
  }
  static __device__ __forceinline__ void ld(const float* p, float* v) {
    const float4 a = ldg_stream(reinterpret_cast<const float4*>(p)), b = ldg_stream(reinterpret_cast<const float4*>(p) + 1);
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
  }
  static __device__ __forceinline__ void ld_cached(const float* p, float* v) {
    const float4 a = __ldg(reinterpret_cast<const float4*>(p)), b = __ldg(reinterpret_cast<const float4*>(p) + 1);
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
  }
  static __device__ __forceinline__ void st(float* p, const float* v) {
    reinterpret_cast<float4*>(p)[0] = make_float4(v[0], v[1], v[2], v[3]);
    reinterpret_cast<float4*>(p)[1] = make_float4(v[4], v[5], v[6], v[7]);
  }
};

constexpr int kVPT = 4;

// sum (a, b) over the row team; kWPR == 1 is shuffle-only, larger teams go through parity-buffered smem
template <int kWPR>
__device__ __forceinline__ void team_sum2(float& a, float& b, float2 (*red)[8], int parity) {
  a = warp_sum(a);
  b = warp_sum(b);
  if (kWPR > 1) {
    const int w = threadIdx.x >> 5;
    if ((threadIdx.x & 31) == 0) red[parity][w] = make_float2(a, b);
    __syncthreads();
    a = 0.f; b = 0.f;
#pragma unroll
    for (int i = 0; i < kWPR; ++i) { a += red[parity][i].x; b += red[parity][i].y; }
  }
}

template <typename T, int kWPR, bool kRms>
__global__ void __launch_bounds__(kWPR == 1 ? 128 : 32 * kWPR)
norm_fwd_v2(const T* __restrict__ x, const T* __restrict__ gamma, const T* __restrict__ beta, T* __restrict__ y,
            float* __restrict__ mean_out, float* __restrict__ rstd_out, int64_t rows, int cols, float eps) {
  __shared__ float2 red[2][8];
  constexpr int kTT = 32 * kWPR;                       // threads in a row team
  const int teams = kWPR == 1 ? 4 : 1;
  const int tid = kWPR == 1 ? (threadIdx.x & 31) : threadIdx.x;
  const int team = kWPR == 1 ? (threadIdx.x >> 5) : 0;
  const int nvec = cols >> 3;
  const float inv_n = 1.f / (float)cols;
  int parity = 0;
  const int64_t row0 = (int64_t)blockIdx.x * teams + team, stride = (int64_t)gridDim.x * teams;
  typename Vec8<T>::Raw raw[kVPT];
  if (row0 < rows) {
#pragma unroll
    for (int k = 0; k < kVPT; ++k) {
      const int j = tid + k * kTT;
      if (j < nvec) raw[k] = Vec8<T>::ld_raw(x + row0 * (int64_t)cols + j * 8);
    }
  }
  for (int64_t row = row0; row < rows; row += stride, parity ^= 1) {
    const T* xr = x + row * (int64_t)cols;
    float v[kVPT][8];
#pragma unroll
    for (int k = 0; k < kVPT; ++k)
      if (tid + k * kTT < nvec) Vec8<T>::unpack(raw[k], v[k]);
    if (row + stride < rows) {                          // next row's loads fly while this one is reduced and stored
#pragma unroll
      for (int k = 0; k < kVPT; ++k) {
        const int j = tid + k * kTT;
        if (j < nvec) raw[k] = Vec8<T>::ld_raw(xr + stride * (int64_t)cols + j * 8);
      }
    }
    const float pivot = kRms ? 0.f : (float)__ldg(xr);
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int k = 0; k < kVPT; ++k) {
      if (tid + k * kTT < nvec) {
#pragma unroll
        for (int i = 0; i < 8; ++i) { const float d = v[k][i] - pivot; s1 += d; s2 += d * d; }
      }
    }
    team_sum2<kWPR>(s1, s2, red, parity);
    const float m1 = s1 * inv_n;
    const float mean = kRms ? 0.f : pivot + m1;
    const float var = kRms ? s2 * inv_n : fmaxf(s2 * inv_n - m1 * m1, 0.f);
    const float rstd = rsqrtf(var + eps);
    if (tid == 0) {
      if (mean_out && !kRms) mean_out[row] = mean;
      if (rstd_out) rstd_out[row] = rstd;
    }
    T* yr = y + row * (int64_t)cols;
#pragma unroll
    for (int k = 0; k < kVPT; ++k) {
      const int j = tid + k * kTT;
      if (j < nvec) {
        float g[8], b[8], o[8];
        Vec8<T>::ld_cached(gamma + j * 8, g);
        if (!kRms && beta != nullptr) Vec8<T>::ld_cached(beta + j * 8, b);
#pragma unroll
        for (int i = 0; i < 8; ++i) o[i] = (v[k][i] - mean) * rstd * g[i] + ((!kRms && beta != nullptr) ? b[i] : 0.f);
        Vec8<T>::st(yr + j * 8, o);
      }
    }
  }
}

template <typename T, int kWPR, bool kRms>
__global__ void __launch_bounds__(kWPR == 1 ? 128 : 32 * kWPR)
norm_bwd_v2(const T* __restrict__ dy, const T* __restrict__ x, const T* __restrict__ gamma, const float* __restrict__ mean_in,
            const float* __restrict__ rstd_in, T* __restrict__ dx, float* __restrict__ dgamma_part, float* __restrict__ dbeta_part,
            int64_t rows, int cols) {
  __shared__ float2 red[2][8];
  extern __shared__ float comb[];                      // kWPR == 1: [2][cols] to merge the 4 warps of a CTA
  constexpr int kTT = 32 * kWPR;
  const int teams = kWPR == 1 ? 4 : 1;
  const int tid = kWPR == 1 ? (threadIdx.x & 31) : threadIdx.x;
  const int team = kWPR == 1 ? (threadIdx.x >> 5) : 0;
  const int nvec = cols >> 3;
  const float inv_n = 1.f / (float)cols;
  float dg[kVPT][8], db[kVPT][8];
#pragma unroll
  for (int k = 0; k < kVPT; ++k)
#pragma unroll
    for (int i = 0; i < 8; ++i) { dg[k][i] = 0.f; db[k][i] = 0.f; }
  int parity = 0;
  const int64_t row0 = (int64_t)blockIdx.x * teams + team, stride = (int64_t)gridDim.x * teams;
  typename Vec8<T>::Raw rx[kVPT], rg[kVPT];
  if (row0 < rows) {
#pragma unroll
    for (int k = 0; k < kVPT; ++k) {
      const int j = tid + k * kTT;
      if (j < nvec) { rx[k] = Vec8<T>::ld_raw(x + row0 * (int64_t)cols + j * 8); rg[k] = Vec8<T>::ld_raw(dy + row0 * (int64_t)cols + j * 8); }
    }
  }
  for (int64_t row = row0; row < rows; row += stride, parity ^= 1) {
    const float mean = kRms ? 0.f : __ldg(mean_in + row);
    const float rstd = __ldg(rstd_in + row);
    const T* xr = x + row * (int64_t)cols;
    const T* gr = dy + row * (int64_t)cols;
    float xh[kVPT][8], g[kVPT][8];
#pragma unroll
    for (int k = 0; k < kVPT; ++k)
      if (tid + k * kTT < nvec) { Vec8<T>::unpack(rx[k], xh[k]); Vec8<T>::unpack(rg[k], g[k]); }
    if (row + stride < rows) {
#pragma unroll
      for (int k = 0; k < kVPT; ++k) {
        const int j = tid + k * kTT;
        if (j < nvec) { rx[k] = Vec8<T>::ld_raw(xr + stride * (int64_t)cols + j * 8); rg[k] = Vec8<T>::ld_raw(gr + stride * (int64_t)cols + j * 8); }
      }
    }
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int k = 0; k < kVPT; ++k) {
      const int j = tid + k * kTT;
      if (j < nvec) {
        float gm[8];
        Vec8<T>::ld_cached(gamma + j * 8, gm);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const float h = (xh[k][i] - mean) * rstd;
          const float d = g[k][i];
          dg[k][i] += d * h;
          db[k][i] += d;
          const float gg = d * gm[i];
          xh[k][i] = h;
          g[k][i] = gg;
          s1 += gg;
          s2 += gg * h;
        }
      }
    }
    team_sum2<kWPR>(s1, s2, red, parity);
    const float m1 = kRms ? 0.f : s1 * inv_n;
    const float m2 = s2 * inv_n;
    T* dr = dx + row * (int64_t)cols;
#pragma unroll
    for (int k = 0; k < kVPT; ++k) {
      const int j = tid + k * kTT;
      if (j < nvec) {
        float o[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) o[i] = (g[k][i] - m1 - xh[k][i] * m2) * rstd;
        Vec8<T>::st(dr + j * 8, o);
      }
    }
  }
  float* gp = dgamma_part + (int64_t)blockIdx.x * cols;
  float* bp = (kRms || dbeta_part == nullptr) ? nullptr : dbeta_part + (int64_t)blockIdx.x * cols;
  if (kWPR == 1) {
    for (int c = threadIdx.x; c < 2 * cols; c += blockDim.x) comb[c] = 0.f;
    __syncthreads();
#pragma unroll
    for (int k = 0; k < kVPT; ++k) {
      const int j = tid + k * kTT;
      if (j < nvec) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          atomicAdd(&comb[j * 8 + i], dg[k][i]);
          if (bp) atomicAdd(&comb[cols + j * 8 + i], db[k][i]);
        }
      }
    }
    __syncthreads();
    for (int c = threadIdx.x; c < cols; c += blockDim.x) {
      gp[c] = comb[c];
      if (bp) bp[c] = comb[cols + c];
    }
  } else {
#pragma unroll
    for (int k = 0; k < kVPT; ++k) {
      const int j = tid + k * kTT;
      if (j < nvec) {
        reinterpret_cast<float4*>(gp + j * 8)[0] = make_float4(dg[k][0], dg[k][1], dg[k][2], dg[k][3]);
        reinterpret_cast<float4*>(gp + j * 8)[1] = make_float4(dg[k][4], dg[k][5], dg[k][6], dg[k][7]);
        if (bp) {
          reinterpret_cast<float4*>(bp + j * 8)[0] = make_float4(db[k][0], db[k][1], db[k][2], db[k][3]);
          reinterpret_cast<float4*>(bp + j * 8)[1] = make_float4(db[k][4], db[k][5], db[k][6], db[k][7]);
        }
      }
    }
  }
}

// out[c] = sum_b part[b, c]: 32 columns x 8 row groups per CTA, coalesced 128-byte rows
template <typename T>
__global__ void __launch_bounds__(256) colsum_v2(const float* __restrict__ part, int nblocks, int cols, T* __restrict__ out) {
  __shared__ float sm[8][33];
  const int cx = threadIdx.x & 31, by = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + cx;
  float s = 0.f;
  if (c < cols)
    for (int b = by; b < nblocks; b += 8) s += part[(int64_t)b * cols + c];
  sm[by][cx] = s;
  __syncthreads();
  if (by == 0 && c < cols) {
    float t = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) t += sm[i][cx];
    out[c] = (T)t;
  }
}

inline int pick_wpr(int cols) { return cols <= 1024 ? 1 : cols <= 2048 ? 2 : cols <= 4096 ? 4 : 8; }
inline bool v2_ok(int cols) { return cols % 8 == 0 && cols >= 8 && cols <= 8192; }
inline int bwd_grid_v2(int64_t rows, int cols) {
  const int teams = pick_wpr(cols) == 1 ? 4 : 1;
  int64_t want = (rows + teams - 1) / teams;
  int64_t cap = (int64_t)kNumSMs * 3;
  return (int)(want < cap ? (want < 1 ? 1 : want) : cap);
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
  if (v2_ok(cols)) {
    const int wpr = pick_wpr(cols);
    const int teams = wpr == 1 ? 4 : 1;
    int64_t want = (rows + teams - 1) / teams;
    const int64_t cap = (int64_t)kNumSMs * 16;
    const int grid = (int)(want < cap ? want : cap);
#define FWD2(W) norm_fwd_v2<T, W, kRms><<<grid, W == 1 ? 128 : 32 * W, 0, st>>>((const T*)x, (const T*)gamma, (const T*)beta, (T*)y, mean, rstd, rows, cols, eps)
    switch (wpr) { case 1: FWD2(1); break; case 2: FWD2(2); break; case 4: FWD2(4); break; default: FWD2(8); break; }
#undef FWD2
    return 0;
  }
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
  if (v2_ok(cols)) {
    const int wpr = pick_wpr(cols);
    const size_t sm = wpr == 1 ? (size_t)2 * cols * sizeof(float) : 0;
#define BWD2(W) norm_bwd_v2<T, W, kRms><<<grid, W == 1 ? 128 : 32 * W, sm, st>>>((const T*)dy, (const T*)x, (const T*)gamma, mean, rstd, (T*)dx, gpart, bpart, rows, cols)
    switch (wpr) { case 1: BWD2(1); break; case 2: BWD2(2); break; case 4: BWD2(4); break; default: BWD2(8); break; }
#undef BWD2
    const int cb2 = (cols + 31) / 32;
    colsum_v2<T><<<cb2, 256, 0, st>>>(gpart, grid, cols, (T*)dgamma);
    if (!kRms && dbeta) colsum_v2<T><<<cb2, 256, 0, st>>>(bpart, grid, cols, (T*)dbeta);
    return 0;
  }
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
  const int grid = v2_ok(cols) ? bwd_grid_v2(rows, cols) : bwd_grid(rows);
  return (int64_t)grid * cols * (rms ? 1 : 2);
}

DETB200_API int detb200_norm_bwd(const void* dy, const void* x, const void* gamma, const float* mean,
                                 const float* rstd, void* dx, void* dgamma, void* dbeta, float* workspace,
                                 int64_t rows, int cols, int dtype, int rms, cudaStream_t stream) {
  if (rows <= 0) return 0;
  if (cols % 2 != 0 || cols < 2) { detb200_set_error("norm: cols must be even"); return DETB200_ERR_UNSUPPORTED; }
  if (!workspace) { detb200_set_error("norm_bwd: workspace required"); return DETB200_ERR_BAD_ARG; }
  int grid = v2_ok(cols) ? bwd_grid_v2(rows, cols) : bwd_grid(rows);
  int rc;
  if (dtype == 1) rc = rms ? launch_bwd<__nv_bfloat16, true>(dy, x, gamma, mean, rstd, dx, dgamma, dbeta, workspace, grid, rows, cols, stream)
                           : launch_bwd<__nv_bfloat16, false>(dy, x, gamma, mean, rstd, dx, dgamma, dbeta, workspace, grid, rows, cols, stream);
  else rc = rms ? launch_bwd<float, true>(dy, x, gamma, mean, rstd, dx, dgamma, dbeta, workspace, grid, rows, cols, stream)
                : launch_bwd<float, false>(dy, x, gamma, mean, rstd, dx, dgamma, dbeta, workspace, grid, rows, cols, stream);
  if (rc) return rc;
  DETB200_LAUNCH_CHECK();
  return 0;
}
