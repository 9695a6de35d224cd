// bf16 GEMM on the 5th-generation tensor cores (tcgen05.mma, accumulators in TMEM, operands staged
// by TMA), persistent and warp-specialised:
//
//   warp 0      TMA producer       global -> 128B-swizzled smem ring (kStages deep)
//   warp 1      MMA issuer         one elected lane issues tcgen05.mma; tcgen05.commit frees ring slots
//                                  and publishes finished accumulators (TMEM is double-buffered)
//   warps 2..5  epilogue           tcgen05.ld -> bias / activation / accumulate -> global
//
// C[M,N] (+)= op(A)[M,K] * op(B)[N,K]^T.  Each operand may be "K-major" (the reduction index is the
// contiguous one: A[M,K], B[N,K] — the nn.Linear forward layout) or "MN-major" (A stored [K,M],
// B stored [K,N]) so that the backward GEMMs (dX = dY*W, dW = dY^T*X) need no transposes.
// Optional A-readiness flags let a concurrently running gather (csrc/gemm_comm.cu) feed the GEMM
// shard by shard.  No counterpart in the reference (its GEMMs are whatever torch calls).
#include "sm100.cuh"
#include <cudaTypedefs.h>
#include <mutex>
#include <unordered_map>
#include <string>
#include <cstring>

using namespace sm100;

namespace {

constexpr int kBlockM = 128;
constexpr int kBlockK = 64;                 // 64 bf16 = one 128-byte swizzle row
constexpr int kUmmaK = 16;
constexpr int kThreads = 192;
constexpr int kEpiThreads = 128;
constexpr int kStageBufBytes = 32 * 128;      // one TMA-store box: 32 rows x 128 B

enum Epilogue : int { EPI_NONE = 0, EPI_GELU_ERF = 1, EPI_GELU_TANH = 2, EPI_RELU = 3, EPI_DGELU_ERF = 4, EPI_DGELU_TANH = 5, EPI_DRELU = 6 };

struct GemmArgs {
  void* C;
  const void* bias;
  void* aux;                 // fwd activations: pre-activation output (bf16, optional); backward: pre-activation input
  int M, N, K;
  int64_t ldc, ldaux;
  int out_fp32, bias_fp32, epilogue, accumulate;
  // optional shard-readiness protocol for A rows (all-gather fused GEMM)
  const unsigned int* a_ready;   // [n_shards] monotonically increasing counters (device memory)
  unsigned int ready_value;
  int rows_per_shard;
  int m_block_rotation;          // first m-block to work on (own shard first)
};

template <int kBlockN> struct Cfg {
  static constexpr int kABytes = kBlockM * kBlockK * 2;
  static constexpr int kBBytes = kBlockN * kBlockK * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kStages = (kBlockN == 256) ? 4 : 6;
  static constexpr int kTmemCols = 2 * kBlockN;                    // two accumulator stages
  static constexpr int kSmemBytes = kStages * kStageBytes + 4 * 2 * kStageBufBytes + 1024 /*align slack*/ + 256 /*barriers*/;
};

__device__ __forceinline__ float gelu_erf(float x) { return 0.5f * x * (1.f + erff(x * 0.70710678118654752f)); }
__device__ __forceinline__ float dgelu_erf(float x) {
  return 0.5f * (1.f + erff(x * 0.70710678118654752f)) + x * 0.39894228040143268f * __expf(-0.5f * x * x);
}
__device__ __forceinline__ float gelu_tanh(float x) {
  const float u = 0.7978845608028654f * (x + 0.044715f * x * x * x);
  return 0.5f * x * (1.f + tanhf(u));
}
__device__ __forceinline__ float dgelu_tanh(float x) {
  const float u = 0.7978845608028654f * (x + 0.044715f * x * x * x);
  const float t = tanhf(u);
  return 0.5f * (1.f + t) + 0.5f * x * (1.f - t * t) * 0.7978845608028654f * (1.f + 3.f * 0.044715f * x * x);
}

__device__ __forceinline__ unsigned int ld_acquire_gpu(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

template <int kBlockN, bool kAMN, bool kBMN>
__global__ void __launch_bounds__(kThreads, 1)
gemm_bf16_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                 const __grid_constant__ CUtensorMap tmap_c, const __grid_constant__ CUtensorMap tmap_aux, const GemmArgs args) {
  using C = Cfg<kBlockN>;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t stage_base = smem_base + C::kStages * C::kStageBytes;     // epilogue staging: 4 warps x 2 x 4 KiB
  const uint32_t bar_base = stage_base + 4 * 2 * kStageBufBytes;
  // barrier layout: full[kStages], empty[kStages], tmem_full[2], tmem_empty[2], then the TMEM base address word
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (C::kStages + s); };
  auto tfull_bar = [&](int s) { return bar_base + 8u * (2 * C::kStages + s); };
  auto tempty_bar = [&](int s) { return bar_base + 8u * (2 * C::kStages + 2 + s); };
  const uint32_t tmem_slot = bar_base + 8u * (2 * C::kStages + 4);
  uint8_t* smem_aligned = smem_raw + (smem_base - smem_u32(smem_raw));
  volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(smem_aligned + C::kStages * C::kStageBytes + 4 * 2 * kStageBufBytes + 8 * (2 * C::kStages + 4));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int mt = (args.M + kBlockM - 1) / kBlockM, nt = (args.N + kBlockN - 1) / kBlockN;
  const int num_tiles = mt * nt;
  const int kb_count = (args.K + kBlockK - 1) / kBlockK;

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmap_a);
    prefetch_tmap(&tmap_b);
    prefetch_tmap(&tmap_c);
    if (args.aux != nullptr && args.epilogue >= EPI_GELU_ERF && args.epilogue <= EPI_RELU) prefetch_tmap(&tmap_aux);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < C::kStages; ++s) {
      mbar_init(full_bar(s), 1);
      mbar_init(empty_bar(s), 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(tfull_bar(s), 1);
      mbar_init(tempty_bar(s), kEpiThreads);
    }
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, C::kTmemCols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;

  if (warp == 0) {
    // ================= TMA producer =================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
        int m_blk = t % mt;
        const int n_blk = t / mt;
        if (args.m_block_rotation) m_blk = (m_blk + args.m_block_rotation) % mt;
        if (args.a_ready != nullptr) {
          // rows of this tile may span two shards; wait for both
          const int s0 = (m_blk * kBlockM) / args.rows_per_shard;
          const int last_row = min(m_blk * kBlockM + kBlockM, args.M) - 1;
          const int s1 = last_row / args.rows_per_shard;
          const uint64_t t0 = globaltimer_ns();
          for (int s = s0; s <= s1; ++s) {
            while (ld_acquire_gpu(args.a_ready + s) < args.ready_value) {
              if (globaltimer_ns() - t0 > 10000000000ull) {
                printf("detb200: gemm timed out waiting for A shard %d\n", s);
                __trap();
              }
              __nanosleep(64);
            }
          }
          fence_proxy_async();   // generic-proxy writes of the gather -> async-proxy (TMA) reads
        }
        for (int kb = 0; kb < kb_count; ++kb) {
          mbar_wait(empty_bar(stage), phase ^ 1u);
          mbar_expect_tx(full_bar(stage), C::kStageBytes);
          const uint32_t sa = smem_base + stage * C::kStageBytes;
          const uint32_t sb = sa + C::kABytes;
          if (!kAMN) {
            tma_load_2d(sa, &tmap_a, full_bar(stage), kb * kBlockK, m_blk * kBlockM);
          } else {
#pragma unroll
            for (int j = 0; j < kBlockM / 64; ++j)
              tma_load_2d(sa + j * (64 * kBlockK * 2), &tmap_a, full_bar(stage), m_blk * kBlockM + j * 64, kb * kBlockK);
          }
          if (!kBMN) {
            tma_load_2d(sb, &tmap_b, full_bar(stage), kb * kBlockK, n_blk * kBlockN);
          } else {
#pragma unroll
            for (int j = 0; j < kBlockN / 64; ++j)
              tma_load_2d(sb + j * (64 * kBlockK * 2), &tmap_b, full_bar(stage), n_blk * kBlockN + j * 64, kb * kBlockK);
          }
          if (++stage == C::kStages) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    // ================= MMA issuer =================
    if (lane == 0) {
      constexpr uint32_t idesc = idesc_bf16_f32(kBlockM, kBlockN, kAMN ? 1 : 0, kBMN ? 1 : 0);
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
        mbar_wait(tempty_bar(acc), acc_phase ^ 1u);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * kBlockN;
        for (int kb = 0; kb < kb_count; ++kb) {
          mbar_wait(full_bar(stage), phase);
          tc_fence_after();
          const uint32_t sa = smem_base + stage * C::kStageBytes;
          const uint32_t sb = sa + C::kABytes;
#pragma unroll
          for (int k = 0; k < kBlockK / kUmmaK; ++k) {
            // K-major: advance 16 elements (32 B) inside the swizzle row; SBO = 8 rows * 128 B.
            // MN-major: advance two 8-row k-groups (2 KiB); LBO = one 64-wide MN box (64 k-rows * 128 B).
            const uint64_t da = kAMN ? smem_desc_sw128(sa + k * 2048, 64 * kBlockK * 2, 1024) : smem_desc_sw128(sa + k * 32, 16, 1024);
            const uint64_t db = kBMN ? smem_desc_sw128(sb + k * 2048, 64 * kBlockK * 2, 1024) : smem_desc_sw128(sb + k * 32, 16, 1024);
            mma_f16_ss(d_tmem, da, db, idesc, (kb | k) != 0 ? 1u : 0u);
          }
          tc_commit(empty_bar(stage));                     // smem slot reusable once these MMAs retire
          if (kb == kb_count - 1) tc_commit(tfull_bar(acc));  // accumulator complete
          if (++stage == C::kStages) { stage = 0; phase ^= 1u; }
        }
        if (++acc == 2) { acc = 0; acc_phase ^= 1u; }
      }
    }
  } else {
    // ================= epilogue (warps 2..5) =================
    // Each warp owns one TMEM lane quarter (32 accumulator rows).  Per 64-column chunk it reads
    // TMEM, applies the fused element-wise work, stages 32 rows x 128 B in 128B-swizzled smem and
    // lets TMA write (or reduce-add) the box to global: coalesced, asynchronous, tails clipped by
    // the tensor map.  Two staging buffers per warp keep one store in flight while the next is built.
    const int q = warp & 3;                      // TMEM lane quarter this warp may read
    const int ew = warp - 2;
    const int row_in_tile = q * 32 + lane;
    const uint32_t stg0 = stage_base + (uint32_t)ew * (2 * kStageBufBytes);
    int buf = 0;
    auto emit = [&](const CUtensorMap* tm, const uint32_t* packed, int c0, int r0, bool reduce_add) {
      if (lane == 0) tma_store_wait_read<1>();       // the buffer used two stores ago has been read out
      __syncwarp();
      const uint32_t dst = stg0 + (uint32_t)buf * kStageBufBytes + (uint32_t)lane * 128u;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const uint32_t a = dst + (uint32_t)((j ^ (lane & 7)) << 4);
        asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(a), "r"(packed[4 * j]), "r"(packed[4 * j + 1]), "r"(packed[4 * j + 2]),
                     "r"(packed[4 * j + 3]) : "memory");
      }
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) {
        if (reduce_add) tma_reduce_add_2d(tm, stg0 + (uint32_t)buf * kStageBufBytes, c0, r0);
        else tma_store_2d(tm, stg0 + (uint32_t)buf * kStageBufBytes, c0, r0);
        tma_store_commit();
      }
      buf ^= 1;
    };
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
      int m_blk = t % mt;
      const int n_blk = t / mt;
      if (args.m_block_rotation) m_blk = (m_blk + args.m_block_rotation) % mt;
      mbar_wait(tfull_bar(acc), acc_phase);
      tc_fence_after();
      const int gm0 = m_blk * kBlockM + q * 32;          // first row of this warp's box
      const int gm = m_blk * kBlockM + row_in_tile;
      const uint32_t t_row = tmem_base + ((uint32_t)(q * 32) << 16) + acc * kBlockN;
#pragma unroll 1
      for (int c = 0; c < kBlockN; c += 64) {
        const int gn = n_blk * kBlockN + c;
        if (gn >= args.N) break;                          // warp-uniform
        uint32_t r[64];
        tmem_ld_32x32(t_row + c, r);
        tmem_ld_32x32(t_row + c + 32, r + 32);
        tmem_ld_wait();
        if (c + 64 >= kBlockN || gn + 64 >= args.N) {     // last chunk of the tile: accumulator drained
          tc_fence_before();
          mbar_arrive(tempty_bar(acc));
        }
        float* v = reinterpret_cast<float*>(r);
        if (args.bias != nullptr) {
          if (args.bias_fp32) {
            const float* b = reinterpret_cast<const float*>(args.bias) + gn;
#pragma unroll
            for (int j = 0; j < 64; ++j) if (gn + j < args.N) v[j] += __ldg(b + j);
          } else {
            const __nv_bfloat16* b = reinterpret_cast<const __nv_bfloat16*>(args.bias) + gn;
#pragma unroll
            for (int j = 0; j < 64; ++j) if (gn + j < args.N) v[j] += __bfloat162float(b[j]);
          }
        }
        uint32_t packed[32];
        if (args.epilogue != EPI_NONE) {
          if (args.epilogue <= EPI_RELU) {
            if (args.aux != nullptr) {                    // keep the pre-activation for backward
#pragma unroll
              for (int j = 0; j < 32; ++j) packed[j] = pack_bf16(v[2 * j], v[2 * j + 1]);
              emit(&tmap_aux, packed, gn, gm0, false);
            }
            if (args.epilogue == EPI_GELU_ERF) {
#pragma unroll
              for (int j = 0; j < 64; ++j) v[j] = gelu_erf(v[j]);
            } else if (args.epilogue == EPI_GELU_TANH) {
#pragma unroll
              for (int j = 0; j < 64; ++j) v[j] = gelu_tanh(v[j]);
            } else {
#pragma unroll
              for (int j = 0; j < 64; ++j) v[j] = fmaxf(v[j], 0.f);
            }
          } else {
            const bool row_ok = gm < args.M;
            const __nv_bfloat16* ap = reinterpret_cast<const __nv_bfloat16*>(args.aux) + (int64_t)gm * args.ldaux + gn;
            const bool full = gn + 64 <= args.N;
#pragma unroll
            for (int j = 0; j < 64; j += 8) {
              float pre[8];
              if (row_ok && full) {
                const uint4 u = *reinterpret_cast<const uint4*>(ap + j);
                pre[0] = bf16lo(u.x); pre[1] = bf16hi(u.x); pre[2] = bf16lo(u.y); pre[3] = bf16hi(u.y);
                pre[4] = bf16lo(u.z); pre[5] = bf16hi(u.z); pre[6] = bf16lo(u.w); pre[7] = bf16hi(u.w);
              } else {
#pragma unroll
                for (int i = 0; i < 8; ++i) pre[i] = (row_ok && gn + j + i < args.N) ? __bfloat162float(ap[j + i]) : 0.f;
              }
              if (args.epilogue == EPI_DGELU_ERF) {
#pragma unroll
                for (int i = 0; i < 8; ++i) v[j + i] *= dgelu_erf(pre[i]);
              } else if (args.epilogue == EPI_DGELU_TANH) {
#pragma unroll
                for (int i = 0; i < 8; ++i) v[j + i] *= dgelu_tanh(pre[i]);
              } else {
#pragma unroll
                for (int i = 0; i < 8; ++i) v[j + i] *= pre[i] > 0.f ? 1.f : 0.f;
              }
            }
          }
        }
        if (args.out_fp32) {
          emit(&tmap_c, r, gn, gm0, args.accumulate != 0);
          if (gn + 32 < args.N) emit(&tmap_c, r + 32, gn + 32, gm0, args.accumulate != 0);
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j) packed[j] = pack_bf16(v[2 * j], v[2 * j + 1]);
          emit(&tmap_c, packed, gn, gm0, args.accumulate != 0);
        }
      }
      if (++acc == 2) { acc = 0; acc_phase ^= 1u; }
    }
    if (lane == 0) tma_store_wait<0>();
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, C::kTmemCols);
  }
}

// ---- host side ------------------------------------------------------------------------------------
PFN_cuTensorMapEncodeTiled_v12000 g_encode = nullptr;
std::once_flag g_encode_once;

void resolve_encode() {
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
    g_encode = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(fn);
}

struct TmapKey {
  const void* base; uint64_t inner, outer, stride; uint32_t bi, bo; uint32_t fp32, pad;
  bool operator==(const TmapKey& o) const { return std::memcmp(this, &o, sizeof(TmapKey)) == 0; }
};
struct TmapHash {
  size_t operator()(const TmapKey& k) const {
    uint64_t h = 1469598103934665603ull;
    const uint64_t v[7] = {(uint64_t)k.base, k.inner, k.outer, k.stride, k.bi, k.bo, k.fp32};
    for (uint64_t x : v) { h ^= x; h *= 1099511628211ull; }
    return (size_t)h;
  }
};
std::mutex g_tmap_mu;
std::unordered_map<TmapKey, CUtensorMap, TmapHash> g_tmap_cache;

template <int kBlockN, bool kAMN, bool kBMN>
int launch(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& tc, const CUtensorMap& taux, const GemmArgs& a, int sms,
           cudaStream_t stream) {
  using C = Cfg<kBlockN>;
  auto* kern = gemm_bf16_kernel<kBlockN, kAMN, kBMN>;
  // per device and cheap; not a stream operation, so it is legal during graph capture
  DETB200_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::kSmemBytes));
  const int mt = (a.M + kBlockM - 1) / kBlockM, nt = (a.N + kBlockN - 1) / kBlockN;
  const int grid = mt * nt < sms ? mt * nt : sms;
  kern<<<grid, kThreads, C::kSmemBytes, stream>>>(ta, tb, tc, taux, a);
  DETB200_LAUNCH_CHECK();
  return 0;
}

}  // namespace

int detb200_make_tmap_2d(CUtensorMap* out, int fp32, const void* base, uint64_t inner, uint64_t outer, uint64_t row_stride_bytes,
                         uint32_t box_inner, uint32_t box_outer) {
  std::call_once(g_encode_once, resolve_encode);
  if (g_encode == nullptr) {
    detb200_set_error("cuTensorMapEncodeTiled is not available from this driver");
    return DETB200_ERR_UNSUPPORTED;
  }
  TmapKey key;
  std::memset(&key, 0, sizeof(key));
  key.base = base; key.inner = inner; key.outer = outer; key.stride = row_stride_bytes; key.bi = box_inner; key.bo = box_outer; key.fp32 = fp32 ? 1u : 0u;
  {
    std::lock_guard<std::mutex> g(g_tmap_mu);
    auto it = g_tmap_cache.find(key);
    if (it != g_tmap_cache.end()) { *out = it->second; return 0; }
  }
  if ((reinterpret_cast<uintptr_t>(base) & 15u) != 0 || (row_stride_bytes & 15u) != 0) {
    detb200_set_error("TMA operand must be 16-byte aligned with a 16-byte-multiple row stride (base %p, stride %llu)", base,
                      (unsigned long long)row_stride_bytes);
    return DETB200_ERR_BAD_ARG;
  }
  const cuuint64_t dims[2] = {inner, outer};
  const cuuint64_t strides[1] = {row_stride_bytes};
  const cuuint32_t box[2] = {box_inner, box_outer};
  const cuuint32_t estr[2] = {1, 1};
  const CUresult r = g_encode(out, fp32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                              CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                              CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    detb200_set_error("cuTensorMapEncodeTiled failed with CUresult %d (inner %llu outer %llu stride %llu box %ux%u)", (int)r,
                      (unsigned long long)inner, (unsigned long long)outer, (unsigned long long)row_stride_bytes, box_inner, box_outer);
    return DETB200_ERR_BAD_ARG;
  }
  std::lock_guard<std::mutex> g(g_tmap_mu);
  if (g_tmap_cache.size() > 4096) g_tmap_cache.clear();
  g_tmap_cache.emplace(key, *out);
  return 0;
}

// Shared by the plain and the gather-fused entry points.
int detb200_gemm_bf16_impl(const void* A, const void* B, void* C, const void* bias, void* aux, int64_t M, int64_t N, int64_t K, int64_t lda,
                           int64_t ldb, int64_t ldc, int64_t ldaux, int a_mn_major, int b_mn_major, int out_fp32, int bias_fp32,
                           int epilogue, int accumulate, const unsigned int* a_ready, unsigned int ready_value, int rows_per_shard,
                           int m_block_rotation, int max_ctas, cudaStream_t stream) {
  if (M <= 0 || N <= 0 || K <= 0) return 0;
  if (M > INT32_MAX || N > INT32_MAX || K > INT32_MAX) {
    detb200_set_error("gemm: dimensions above 2^31 are not supported");
    return DETB200_ERR_TOO_LARGE;
  }
  if (epilogue < 0 || epilogue > EPI_DRELU || (epilogue > EPI_RELU && aux == nullptr)) {
    detb200_set_error("gemm: bad epilogue %d (backward-activation epilogues need the saved pre-activation)", epilogue);
    return DETB200_ERR_BAD_ARG;
  }
  const int elem_c = out_fp32 ? 4 : 2;
  if ((reinterpret_cast<uintptr_t>(C) & 15u) || ((ldc * elem_c) & 15) || (aux && ((reinterpret_cast<uintptr_t>(aux) & 15u) || ((ldaux * 2) & 15)))) {
    detb200_set_error("gemm: C/aux must be 16-byte aligned with 16-byte-multiple row strides");
    return DETB200_ERR_BAD_ARG;
  }
  const int bn = (N > 128) ? 256 : 128;
  CUtensorMap ta, tb;
  int rc;
  if (!a_mn_major) rc = detb200_make_tmap_2d(&ta, 0, A, (uint64_t)K, (uint64_t)M, (uint64_t)lda * 2, kBlockK, kBlockM);
  else rc = detb200_make_tmap_2d(&ta, 0, A, (uint64_t)M, (uint64_t)K, (uint64_t)lda * 2, 64, kBlockK);
  if (rc) return rc;
  if (!b_mn_major) rc = detb200_make_tmap_2d(&tb, 0, B, (uint64_t)K, (uint64_t)N, (uint64_t)ldb * 2, kBlockK, bn);
  else rc = detb200_make_tmap_2d(&tb, 0, B, (uint64_t)N, (uint64_t)K, (uint64_t)ldb * 2, 64, kBlockK);
  if (rc) return rc;
  // output boxes: 32 rows x 128 bytes (64 bf16 or 32 fp32 columns)
  CUtensorMap tc, taux;
  rc = detb200_make_tmap_2d(&tc, out_fp32, C, (uint64_t)N, (uint64_t)M, (uint64_t)ldc * elem_c, out_fp32 ? 32 : 64, 32);
  if (rc) return rc;
  if (aux != nullptr && epilogue >= EPI_GELU_ERF && epilogue <= EPI_RELU) {
    rc = detb200_make_tmap_2d(&taux, 0, aux, (uint64_t)N, (uint64_t)M, (uint64_t)ldaux * 2, 64, 32);
    if (rc) return rc;
  } else {
    taux = tc;
  }
  GemmArgs g;
  g.C = C; g.bias = bias; g.aux = aux; g.M = (int)M; g.N = (int)N; g.K = (int)K; g.ldc = ldc; g.ldaux = ldaux;
  g.out_fp32 = out_fp32; g.bias_fp32 = bias_fp32; g.epilogue = epilogue; g.accumulate = accumulate;
  g.a_ready = a_ready; g.ready_value = ready_value; g.rows_per_shard = rows_per_shard > 0 ? rows_per_shard : (int)M;
  g.m_block_rotation = m_block_rotation;
  int dev = 0, sms = kNumSMs;
  DETB200_CHECK(cudaGetDevice(&dev));
  DETB200_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  if (max_ctas > 0 && max_ctas < sms) sms = max_ctas;
#define DETB200_GEMM_DISPATCH(BN)                                                                         \
  do {                                                                                                    \
    if (!a_mn_major && !b_mn_major) return launch<BN, false, false>(ta, tb, tc, taux, g, sms, stream);              \
    if (!a_mn_major && b_mn_major) return launch<BN, false, true>(ta, tb, tc, taux, g, sms, stream);                \
    if (a_mn_major && !b_mn_major) return launch<BN, true, false>(ta, tb, tc, taux, g, sms, stream);                \
    return launch<BN, true, true>(ta, tb, tc, taux, g, sms, stream);                                                \
  } while (0)
  if (bn == 256) DETB200_GEMM_DISPATCH(256);
  DETB200_GEMM_DISPATCH(128);
#undef DETB200_GEMM_DISPATCH
}

DETB200_API int detb200_gemm_bf16(const void* A, const void* B, void* C, const void* bias, void* aux, int64_t M, int64_t N, int64_t K,
                                  int64_t lda, int64_t ldb, int64_t ldc, int64_t ldaux, int a_mn_major, int b_mn_major, int out_fp32,
                                  int bias_fp32, int epilogue, int accumulate, cudaStream_t stream) {
  return detb200_gemm_bf16_impl(A, B, C, bias, aux, M, N, K, lda, ldb, ldc, ldaux, a_mn_major, b_mn_major, out_fp32, bias_fp32, epilogue,
                                accumulate, nullptr, 0, 0, 0, 0, stream);
}
