// Small fused element-wise / reduction kernels of the transformer path (bf16 I/O, fp32 math).
//
//   rotary_qkv     rotate the first `rot` channels of Q and K *in place* inside a packed [B,S,3,H,D]
//                  projection (NeoX "rotate half"); `conj` applies the inverse rotation = the backward.
//   colsum_bf16    out[n] (+)= sum_m x[m, n]  (bias gradients), fp32 accumulation, coalesced rows.
//   cast_bf16      fp32 -> bf16 copy (weight shadows for the tcgen05 GEMMs).
//   add_bf16       y = a + b (+ c)  residual joins of the parallel-residual block.
// Replaces chains of ATen elementwise/cat/neg/mul/reduce kernels (K11-K13 glue in SURVEY.md §2.7).
#include "common.cuh"

namespace {

// one thread = 8 consecutive channel pairs (x[i..i+8), x[i+half..i+half+8)) of one (b, s, which, h)
__global__ void __launch_bounds__(256)
rotary_qkv_kernel(__nv_bfloat16* __restrict__ qkv, const float* __restrict__ cos_t, const float* __restrict__ sin_t, int64_t total_vec,
                  int S, int H, int D, int half, int64_t sb, int64_t ss, int64_t sw, int64_t sh, float sign) {
  const int vec_per_head = half >> 3;
  for (int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; v < total_vec; v += (int64_t)gridDim.x * blockDim.x) {
    const int iv = (int)(v % vec_per_head);
    int64_t t = v / vec_per_head;
    const int h = (int)(t % H); t /= H;
    const int which = (int)(t % 2); t /= 2;                    // 0 = q, 1 = k
    const int s = (int)(t % S);
    const int64_t b = t / S;
    __nv_bfloat16* p = qkv + b * sb + (int64_t)s * ss + (int64_t)which * sw + (int64_t)h * sh + iv * 8;
    uint4 u1 = *reinterpret_cast<const uint4*>(p), u2 = *reinterpret_cast<const uint4*>(p + half);
    const float4 c0 = *reinterpret_cast<const float4*>(cos_t + (int64_t)s * half + iv * 8), c1 = *reinterpret_cast<const float4*>(cos_t + (int64_t)s * half + iv * 8 + 4);
    const float4 s0 = *reinterpret_cast<const float4*>(sin_t + (int64_t)s * half + iv * 8), s1 = *reinterpret_cast<const float4*>(sin_t + (int64_t)s * half + iv * 8 + 4);
    const float c[8] = {c0.x, c0.y, c0.z, c0.w, c1.x, c1.y, c1.z, c1.w};
    const float sn[8] = {s0.x * sign, s0.y * sign, s0.z * sign, s0.w * sign, s1.x * sign, s1.y * sign, s1.z * sign, s1.w * sign};
    const uint32_t* a = reinterpret_cast<const uint32_t*>(&u1);
    const uint32_t* bq = reinterpret_cast<const uint32_t*>(&u2);
    uint32_t o1[4], o2[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float x1a = bf16lo(a[i]), x1b = bf16hi(a[i]), x2a = bf16lo(bq[i]), x2b = bf16hi(bq[i]);
      o1[i] = pack_bf16(x1a * c[2 * i] - x2a * sn[2 * i], x1b * c[2 * i + 1] - x2b * sn[2 * i + 1]);
      o2[i] = pack_bf16(x2a * c[2 * i] + x1a * sn[2 * i], x2b * c[2 * i + 1] + x1b * sn[2 * i + 1]);
    }
    *reinterpret_cast<uint4*>(p) = make_uint4(o1[0], o1[1], o1[2], o1[3]);
    *reinterpret_cast<uint4*>(p + half) = make_uint4(o2[0], o2[1], o2[2], o2[3]);
  }
}

// grid (ceil(N/128), row_chunks); 8 warps per CTA, lane = 4 columns (8-byte loads), 4 rows in flight per warp
__global__ void __launch_bounds__(256)
colsum_bf16_kernel(const __nv_bfloat16* __restrict__ x, float* __restrict__ out, int64_t M, int N, int64_t ld) {
  __shared__ float4 sm[8][32];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int col = blockIdx.x * 128 + lane * 4;
  const int64_t rows_per = (M + gridDim.y - 1) / gridDim.y;
  const int64_t r0 = (int64_t)blockIdx.y * rows_per, r1 = (r0 + rows_per < M) ? r0 + rows_per : M;
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  if (col < N) {
    int64_t r = r0 + w;
    for (; r + 24 < r1; r += 32) {
      uint2 u[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) u[k] = *reinterpret_cast<const uint2*>(x + (r + 8 * k) * ld + col);
#pragma unroll
      for (int k = 0; k < 4; ++k) { acc.x += bf16lo(u[k].x); acc.y += bf16hi(u[k].x); acc.z += bf16lo(u[k].y); acc.w += bf16hi(u[k].y); }
    }
    for (; r < r1; r += 8) {
      const uint2 u = *reinterpret_cast<const uint2*>(x + r * ld + col);
      acc.x += bf16lo(u.x); acc.y += bf16hi(u.x); acc.z += bf16lo(u.y); acc.w += bf16hi(u.y);
    }
  }
  sm[w][lane] = acc;
  __syncthreads();
  if (w == 0 && col < N) {
    float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int i = 0; i < 8; ++i) { t.x += sm[i][lane].x; t.y += sm[i][lane].y; t.z += sm[i][lane].z; t.w += sm[i][lane].w; }
    atomicAdd(out + col, t.x);
    atomicAdd(out + col + 1, t.y);
    atomicAdd(out + col + 2, t.z);
    atomicAdd(out + col + 3, t.w);
  }
}

__global__ void __launch_bounds__(256) cast_bf16_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ y, int64_t n8) {
  for (int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; v < n8; v += (int64_t)gridDim.x * blockDim.x) {
    const float4 a = ldg_stream(reinterpret_cast<const float4*>(x) + 2 * v), b = ldg_stream(reinterpret_cast<const float4*>(x) + 2 * v + 1);
    reinterpret_cast<uint4*>(y)[v] = make_uint4(pack_bf16(a.x, a.y), pack_bf16(a.z, a.w), pack_bf16(b.x, b.y), pack_bf16(b.z, b.w));
  }
}

__global__ void __launch_bounds__(256)
add_bf16_kernel(const __nv_bfloat16* __restrict__ a, const __nv_bfloat16* __restrict__ b, const __nv_bfloat16* __restrict__ c,
                __nv_bfloat16* __restrict__ y, int64_t n8) {
  for (int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; v < n8; v += (int64_t)gridDim.x * blockDim.x) {
    const uint4 ua = ldg_stream_u4(reinterpret_cast<const uint4*>(a) + v), ub = ldg_stream_u4(reinterpret_cast<const uint4*>(b) + v);
    uint4 uc = make_uint4(0, 0, 0, 0);
    if (c != nullptr) uc = ldg_stream_u4(reinterpret_cast<const uint4*>(c) + v);
    const uint32_t* pa = reinterpret_cast<const uint32_t*>(&ua);
    const uint32_t* pb = reinterpret_cast<const uint32_t*>(&ub);
    const uint32_t* pc = reinterpret_cast<const uint32_t*>(&uc);
    uint32_t o[4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
      o[i] = pack_bf16(bf16lo(pa[i]) + bf16lo(pb[i]) + bf16lo(pc[i]), bf16hi(pa[i]) + bf16hi(pb[i]) + bf16hi(pc[i]));
    reinterpret_cast<uint4*>(y)[v] = make_uint4(o[0], o[1], o[2], o[3]);
  }
}

// Multi-tensor gradient accumulation: dst[k] (fp32, inside the flat gradient buffer) += src[k] (bf16 or fp32) for up to
// kAccumMax tensors in one launch.  The table travels by value in the kernel parameters (no device-side table to keep in
// sync, and CUDA-graph capture records it).  blockIdx.y picks the tensor, blockIdx.x strides over it.
constexpr int kAccumMax = 48;
struct AccumTable {
  const void* src[kAccumMax];
  float* dst[kAccumMax];
  int64_t n[kAccumMax];
  unsigned char is_bf16[kAccumMax];
};

__global__ void __launch_bounds__(256) accum_multi_kernel(AccumTable t) {
  const int k = blockIdx.y;
  const int64_t n = t.n[k];
  float* __restrict__ dst = t.dst[k];
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t.is_bf16[k]) {
    const __nv_bfloat16* __restrict__ src = reinterpret_cast<const __nv_bfloat16*>(t.src[k]);
    const bool vec = ((reinterpret_cast<uintptr_t>(src) & 15u) == 0) && ((reinterpret_cast<uintptr_t>(dst) & 15u) == 0);
    const int64_t n8 = vec ? (n >> 3) : 0;
    for (int64_t v = tid; v < n8; v += stride) {
      const uint4 u = ldg_stream_u4(reinterpret_cast<const uint4*>(src) + v);
      float4 a = reinterpret_cast<const float4*>(dst)[2 * v], b = reinterpret_cast<const float4*>(dst)[2 * v + 1];
      a.x += bf16lo(u.x); a.y += bf16hi(u.x); a.z += bf16lo(u.y); a.w += bf16hi(u.y);
      b.x += bf16lo(u.z); b.y += bf16hi(u.z); b.z += bf16lo(u.w); b.w += bf16hi(u.w);
      reinterpret_cast<float4*>(dst)[2 * v] = a;
      reinterpret_cast<float4*>(dst)[2 * v + 1] = b;
    }
    for (int64_t i = n8 * 8 + tid; i < n; i += stride) dst[i] += __bfloat162float(src[i]);
  } else {
    const float* __restrict__ src = reinterpret_cast<const float*>(t.src[k]);
    const bool vec = ((reinterpret_cast<uintptr_t>(src) & 15u) == 0) && ((reinterpret_cast<uintptr_t>(dst) & 15u) == 0);
    const int64_t n4 = vec ? (n >> 2) : 0;
    for (int64_t v = tid; v < n4; v += stride) {
      const float4 s4 = ldg_stream(reinterpret_cast<const float4*>(src) + v);
      float4 a = reinterpret_cast<const float4*>(dst)[v];
      a.x += s4.x; a.y += s4.y; a.z += s4.z; a.w += s4.w;
      reinterpret_cast<float4*>(dst)[v] = a;
    }
    for (int64_t i = n4 * 4 + tid; i < n; i += stride) dst[i] += src[i];
  }
}

}  // namespace

// dst[k][0:n[k]) += src[k][0:n[k]) for k < count.  srcs / dsts / numels / is_bf16 are HOST arrays; dst is fp32, src is
// bf16 (is_bf16[k] != 0) or fp32.  Replaces the one-add-kernel-per-parameter gradient accumulation of autograd
// (AccumulateGrad) and the bf16 -> fp32 gradient casts of autocast: reference `loss.backward()` under
// harness/determined/pytorch/_pytorch_context.py:786-803.  Any count: launched in groups of 48 tensors.
DETB200_API int detb200_accum_multi(const void* const* srcs, float* const* dsts, const int64_t* numels, const unsigned char* is_bf16,
                                    int count, cudaStream_t stream) {
  if (count <= 0) return 0;
  if (!srcs || !dsts || !numels || !is_bf16) { detb200_set_error("accum_multi: null table"); return DETB200_ERR_BAD_ARG; }
  for (int base = 0; base < count; base += kAccumMax) {
    AccumTable t;
    const int m = (count - base < kAccumMax) ? count - base : kAccumMax;
    int64_t nmax = 0;
    for (int k = 0; k < m; ++k) {
      if (!srcs[base + k] || !dsts[base + k] || numels[base + k] < 0) { detb200_set_error("accum_multi: bad entry %d", base + k); return DETB200_ERR_BAD_ARG; }
      t.src[k] = srcs[base + k]; t.dst[k] = dsts[base + k]; t.n[k] = numels[base + k]; t.is_bf16[k] = is_bf16[base + k];
      if (t.n[k] > nmax) nmax = t.n[k];
    }
    for (int k = m; k < kAccumMax; ++k) { t.src[k] = nullptr; t.dst[k] = nullptr; t.n[k] = 0; t.is_bf16[k] = 0; }
    // enough CTAs along x for the largest tensor at 8 elements per thread per trip, capped so that a group fills the GPU a few times
    int gx = grid_for(nmax, 256, 16, (kNumSMs * 8 + m - 1) / m);
    accum_multi_kernel<<<dim3(gx, m), 256, 0, stream>>>(t);
    DETB200_LAUNCH_CHECK();
  }
  return 0;
}

// qkv: bf16 packed projection addressed as (b, s, which in {q,k,v}, h, d) with element strides sb, ss, sw, sh (d contiguous).
// cos/sin: fp32 [S, rot/2].  rot/2 must be a multiple of 8.  conj != 0 applies the inverse rotation (the backward).
DETB200_API int detb200_rotary_qkv(void* qkv, const float* cos_t, const float* sin_t, int64_t B, int S, int H, int D, int rot, int64_t sb,
                                   int64_t ss, int64_t sw, int64_t sh, int conj, cudaStream_t stream) {
  const int half = rot / 2;
  if (rot <= 0) return 0;
  if (half % 8 != 0 || rot > D || (sb % 8) || (ss % 8) || (sw % 8) || (sh % 8) || (reinterpret_cast<uintptr_t>(qkv) & 15u)) {
    detb200_set_error("rotary_qkv: rot/2 (%d) must be a multiple of 8, rot <= D, strides multiples of 8 and a 16-byte aligned base", half);
    return DETB200_ERR_UNSUPPORTED;
  }
  const int64_t total = B * S * 2 * H * (half / 8);
  const int grid = grid_for(total, 256, 1, kNumSMs * 8);
  rotary_qkv_kernel<<<grid, 256, 0, stream>>>(reinterpret_cast<__nv_bfloat16*>(qkv), cos_t, sin_t, total, S, H, D, half, sb, ss, sw, sh,
                                              conj ? -1.f : 1.f);
  DETB200_LAUNCH_CHECK();
  return 0;
}

// out (fp32 [N]) = (accumulate ? out : 0) + column sums of x (bf16 [M, N], row stride ld elements, N % 4 == 0).
DETB200_API int detb200_colsum_bf16(const void* x, float* out, int64_t M, int N, int64_t ld, int accumulate, cudaStream_t stream) {
  if (N <= 0) return 0;
  if ((N & 3) || (ld & 3) || (reinterpret_cast<uintptr_t>(x) & 7u)) {
    detb200_set_error("colsum_bf16: N and ld must be multiples of 4 and x 8-byte aligned");
    return DETB200_ERR_UNSUPPORTED;
  }
  if (!accumulate) DETB200_CHECK(cudaMemsetAsync(out, 0, sizeof(float) * (size_t)N, stream));
  if (M <= 0) return 0;
  const int gx = (N + 127) / 128;
  int gy = (kNumSMs * 6 + gx - 1) / gx;
  if (gy > (M + 63) / 64) gy = (int)((M + 63) / 64);
  if (gy < 1) gy = 1;
  colsum_bf16_kernel<<<dim3(gx, gy), 256, 0, stream>>>(reinterpret_cast<const __nv_bfloat16*>(x), out, M, N, ld);
  DETB200_LAUNCH_CHECK();
  return 0;
}

DETB200_API int detb200_cast_bf16(const float* x, void* y, int64_t n, cudaStream_t stream) {
  if (n <= 0) return 0;
  if ((n & 7) || (reinterpret_cast<uintptr_t>(x) & 15u) || (reinterpret_cast<uintptr_t>(y) & 15u)) {
    detb200_set_error("cast_bf16: n must be a multiple of 8 and pointers 16-byte aligned");
    return DETB200_ERR_UNSUPPORTED;
  }
  cast_bf16_kernel<<<grid_for(n / 8, 256, 2, kNumSMs * 8), 256, 0, stream>>>(x, reinterpret_cast<__nv_bfloat16*>(y), n / 8);
  DETB200_LAUNCH_CHECK();
  return 0;
}

// y = a + b (+ c when non-null); all bf16, n a multiple of 8, 16-byte aligned; y may alias an input.
DETB200_API int detb200_add_bf16(const void* a, const void* b, const void* c, void* y, int64_t n, cudaStream_t stream) {
  if (n <= 0) return 0;
  if (n & 7) {
    detb200_set_error("add_bf16: n must be a multiple of 8");
    return DETB200_ERR_UNSUPPORTED;
  }
  add_bf16_kernel<<<grid_for(n / 8, 256, 2, kNumSMs * 8), 256, 0, stream>>>(reinterpret_cast<const __nv_bfloat16*>(a), reinterpret_cast<const __nv_bfloat16*>(b),
                                                                          reinterpret_cast<const __nv_bfloat16*>(c), reinterpret_cast<__nv_bfloat16*>(y), n / 8);
  DETB200_LAUNCH_CHECK();
  return 0;
}
