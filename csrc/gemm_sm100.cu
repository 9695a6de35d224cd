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
#include <cstdlib>

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
  int m_block_rotation;          // first block (of the sharded dimension) to work on: own shard first
  int gather_mode;               // which dimension the readiness counters shard: 0 = rows of A (M), 1 = rows of a K-major B (N),
                                 // 2 = rows of an MN-major B (K: the k-loop is rotated instead of the tile walk)
  // optional fused reduce-scatter: tiles of row shard d are stored (bf16) into rank d's receive slot
  int rs_world, rs_rank;
  unsigned int rs_epoch;
  unsigned int* rs_entered;            // my ctrl: entered[r] written by rank r at kernel start
  unsigned int* rs_peer_count[8];      // rank d's counter for partials coming from me
};

struct TmapArray { CUtensorMap m[8]; };

// Control block of the fused GEMM+collective kernels (one per rank, symmetric memory, zero-initialised).
struct GemmCtrl {
  unsigned int entered[16];     // entered[r]: rank r has started call number `epoch` (its earlier work is complete)
  unsigned int a_ready[16];     // a_ready[s]: arrivals for gathered shard s (monotonic)
  unsigned int rs_count[16];    // rs_count[s]: partial tiles received from rank s (monotonic)
  unsigned int pad[16];
};

template <int kBlockN, bool kTwoCta = false> struct Cfg {
  static constexpr int kBRows = kTwoCta ? kBlockN / 2 : kBlockN;      // rows of B this CTA stages (a CTA pair splits N)
  static constexpr int kABytes = kBlockM * kBlockK * 2;
  static constexpr int kBBytes = kBRows * kBlockK * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kStages = (kStageBytes > 32768) ? 4 : 6;
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

__device__ __forceinline__ unsigned int ld_acquire_sys_u32(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys_u32(unsigned int* p, unsigned int v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void red_add_release_sys(unsigned int* p, unsigned int v) {
  asm volatile("red.release.sys.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
// spin until *p - target >= 0 (wrap-safe), trapping after `ns`
__device__ __forceinline__ void wait_counter_sys(const unsigned int* p, unsigned int target, const char* what, int idx) {
  if ((int)(ld_acquire_sys_u32(p) - target) >= 0) return;
  const uint64_t t0 = globaltimer_ns();
  while ((int)(ld_acquire_sys_u32(p) - target) < 0) {
    if (globaltimer_ns() - t0 > 10000000000ull) {
      printf("detb200: timed out waiting for %s[%d] (have %u want %u) block %d\n", what, idx, ld_acquire_sys_u32(p), target, blockIdx.x);
      __trap();
    }
    __nanosleep(100);
  }
}

// Shard-major tile walk: t = (shard_idx * nt + n_blk) * mb_per_shard + m_local, shards rotated by `rotation` m-blocks.
__device__ __forceinline__ void tile_coords(int t, int mt, int nt, int mb_per_shard, int rotation, int& m_blk, int& n_blk) {
  const int tiles_per_shard = mb_per_shard * nt;
  const int shard_idx = t / tiles_per_shard;
  const int r = t - shard_idx * tiles_per_shard;
  n_blk = r / mb_per_shard;
  m_blk = shard_idx * mb_per_shard + (r - n_blk * mb_per_shard) + rotation;
  if (m_blk >= mt) m_blk -= mt;
}

// kRS: the epilogue ships bf16 partial tiles to the rank owning the row shard instead of writing C.
template <int kBlockN, bool kAMN, bool kBMN, bool kRS, bool kTwoCta = false>
__device__ __forceinline__ void gemm_body(const CUtensorMap& tmap_a, const CUtensorMap& tmap_b, const CUtensorMap& tmap_c,
                                          const CUtensorMap& tmap_aux, const CUtensorMap* tmap_rs, const GemmArgs& args, const int cta,
                                          const int ncta) {
  using C = Cfg<kBlockN, kTwoCta>;
  const uint32_t cta_rank = kTwoCta ? cluster_ctarank() : 0u;
  const bool leader = cta_rank == 0;
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
  // a CTA pair walks 256-row tiles; each CTA of the pair owns 128 of the rows
  constexpr int kTileM = kTwoCta ? 2 * kBlockM : kBlockM;
  const int mt = (args.M + kTileM - 1) / kTileM, nt = (args.N + kBlockN - 1) / kBlockN;
  const int num_tiles = mt * nt;
  const int kb_count = (args.K + kBlockK - 1) / kBlockK;
  // tiles are walked shard by shard (all column blocks of a row shard before the next shard) so that a fused
  // collective can overlap: the plain GEMM is the one-shard case
  const bool sharded = !kTwoCta && args.rows_per_shard > 0;
  const int mb_per_shard = (sharded && args.gather_mode == 0 && args.rows_per_shard < args.M) ? args.rows_per_shard / kBlockM : mt;
  const int nb_per_shard = (sharded && args.gather_mode == 1 && args.rows_per_shard < args.N) ? args.rows_per_shard / kBlockN : nt;
  const int kb_per_shard = (sharded && args.gather_mode == 2 && args.rows_per_shard < args.K) ? args.rows_per_shard / kBlockK : kb_count;
  auto coords = [&](int t, int& m_blk, int& n_blk) {
    if (sharded && args.gather_mode == 1) tile_coords(t, nt, mt, nb_per_shard, args.m_block_rotation, n_blk, m_blk);   // walk N shard by shard
    else tile_coords(t, mt, nt, mb_per_shard, args.gather_mode == 0 ? args.m_block_rotation : 0, m_blk, n_blk);
  };

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
      mbar_init(tempty_bar(s), kTwoCta ? 2 * kEpiThreads : kEpiThreads);     // the leader's copy collects both CTAs' epilogues
    }
    fence_mbar_init();
  }
  if (warp == 1) {
    if (kTwoCta) tmem_alloc_2sm(tmem_slot, C::kTmemCols);
    else tmem_alloc(tmem_slot, C::kTmemCols);
  }
  tc_fence_before();
  if (kTwoCta) cluster_sync();
  else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;

  if (warp == 0) {
    // ================= TMA producer =================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int t = cta; t < num_tiles; t += ncta) {
        int m_blk, n_blk;
        coords(t, m_blk, n_blk);
        if (kTwoCta) m_blk = 2 * m_blk + (int)cta_rank;
        if (args.a_ready != nullptr && args.gather_mode != 2) {
          // rows of this tile may span two shards; wait for both
          const int first = args.gather_mode == 0 ? m_blk * kBlockM : n_blk * kBlockN;
          const int last = (args.gather_mode == 0 ? min(first + kBlockM, args.M) : min(first + kBlockN, args.N)) - 1;
          for (int s = first / args.rows_per_shard; s <= last / args.rows_per_shard; ++s) wait_counter_sys(args.a_ready + s, args.ready_value, "a_ready", s);
          fence_proxy_async();   // async-proxy writes of the gather (remote) -> async-proxy (TMA) reads after the acquire
        }
        for (int kb_i = 0; kb_i < kb_count; ++kb_i) {
          int kb = kb_i;
          if (args.gather_mode == 2 && args.a_ready != nullptr) {
            // K-sharded operand: start the reduction at the own shard and take remote k-ranges as they arrive
            kb = kb_i + args.m_block_rotation;
            if (kb >= kb_count) kb -= kb_count;
            if (kb_i == 0 || kb % kb_per_shard == 0) {
              wait_counter_sys(args.a_ready + kb / kb_per_shard, args.ready_value, "a_ready(k)", kb / kb_per_shard);
              fence_proxy_async();
            }
          }
          mbar_wait(empty_bar(stage), phase ^ 1u);
          // pair mode: both CTAs' loads are credited to the leader's barrier, which expects the bytes of both
          if (!kTwoCta || leader) mbar_expect_tx(full_bar(stage), C::kStageBytes * (kTwoCta ? 2 : 1));
          const uint32_t sa = smem_base + stage * C::kStageBytes;
          const uint32_t sb = sa + C::kABytes;
          const int b_row0 = n_blk * kBlockN + (int)cta_rank * C::kBRows;
          auto load = [&](uint32_t dst, const CUtensorMap* tm, int c0, int c1) {
            if (kTwoCta) tma_load_2d_2sm(dst, tm, full_bar(stage), c0, c1);
            else tma_load_2d(dst, tm, full_bar(stage), c0, c1);
          };
          if (!kAMN) {
            load(sa, &tmap_a, kb * kBlockK, m_blk * kBlockM);
          } else {
#pragma unroll
            for (int j = 0; j < kBlockM / 64; ++j) load(sa + j * (64 * kBlockK * 2), &tmap_a, m_blk * kBlockM + j * 64, kb * kBlockK);
          }
          if (!kBMN) {
            load(sb, &tmap_b, kb * kBlockK, b_row0);
          } else {
#pragma unroll
            for (int j = 0; j < C::kBRows / 64; ++j) load(sb + j * (64 * kBlockK * 2), &tmap_b, b_row0 + j * 64, kb * kBlockK);
          }
          if (++stage == C::kStages) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    // ================= MMA issuer =================
    if (lane == 0 && (!kTwoCta || leader)) {
      constexpr uint32_t idesc = idesc_bf16_f32(kTileM, kBlockN, kAMN ? 1 : 0, kBMN ? 1 : 0);
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int t = cta; t < num_tiles; t += ncta) {
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
            if (kTwoCta) mma_f16_ss_2sm(d_tmem, da, db, idesc, (kb | k) != 0 ? 1u : 0u);
            else mma_f16_ss(d_tmem, da, db, idesc, (kb | k) != 0 ? 1u : 0u);
          }
          if (kTwoCta) {
            tc_commit_2sm(empty_bar(stage), 3);                       // frees the slot in both CTAs
            if (kb == kb_count - 1) tc_commit_2sm(tfull_bar(acc), 3);  // both epilogues may read their half
          } else {
            tc_commit(empty_bar(stage));                     // smem slot reusable once these MMAs retire
            if (kb == kb_count - 1) tc_commit(tfull_bar(acc));  // accumulator complete
          }
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
    for (int t = cta; t < num_tiles; t += ncta) {
      int m_blk, n_blk;
      coords(t, m_blk, n_blk);
      if (kTwoCta) m_blk = 2 * m_blk + (int)cta_rank;
      mbar_wait(tfull_bar(acc), acc_phase);
      tc_fence_after();
      int gm0 = m_blk * kBlockM + q * 32;                // first row of this warp's box
      const int gm = m_blk * kBlockM + row_in_tile;
      const CUtensorMap* tm_out = &tmap_c;
      int shard = 0;
      if (kRS) {
        shard = (m_blk * kBlockM) / args.rows_per_shard;
        tm_out = tmap_rs + shard;
        gm0 -= shard * args.rows_per_shard;
        // the owner must have left its previous reduction before its receive slots are overwritten
        if (shard != args.rs_rank) wait_counter_sys(args.rs_entered + shard, args.rs_epoch, "entered", shard);
      }
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
          if (kTwoCta) mbar_arrive_cluster(tempty_bar(acc), 0);
          else mbar_arrive(tempty_bar(acc));
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
        if (!kRS && args.out_fp32) {
          emit(tm_out, r, gn, gm0, args.accumulate != 0);
          if (gn + 32 < args.N) emit(tm_out, r + 32, gn + 32, gm0, args.accumulate != 0);
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j) packed[j] = pack_bf16(v[2 * j], v[2 * j + 1]);
          emit(tm_out, packed, gn, gm0, !kRS && args.accumulate != 0);
        }
      }
      if (kRS) {
        // partial rows of this warp have fully landed at the owner -> count one arrival there
        if (lane == 0) {
          tma_store_wait<0>();
          fence_proxy_async();
          red_add_release_sys(args.rs_peer_count[shard], 1u);
        }
        __syncwarp();
      }
      if (++acc == 2) { acc = 0; acc_phase ^= 1u; }
    }
    if (lane == 0) tma_store_wait<0>();
  }

  tc_fence_before();
  if (kTwoCta) cluster_sync();
  else __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    if (kTwoCta) tmem_dealloc_2sm(tmem_base, C::kTmemCols);
    else tmem_dealloc(tmem_base, C::kTmemCols);
  }
}

// CTA-pair variant: two SMs of a TPC share one 256x256 tile (cta_group::2); each stages half of A's rows and half of B's,
// which cuts shared-memory traffic per SM by a third and leaves room for six pipeline stages.
template <bool kAMN, bool kBMN>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1)
gemm2_bf16_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                  const __grid_constant__ CUtensorMap tmap_c, const __grid_constant__ CUtensorMap tmap_aux, const GemmArgs args) {
  gemm_body<256, kAMN, kBMN, false, true>(tmap_a, tmap_b, tmap_c, tmap_aux, nullptr, args, blockIdx.x >> 1, gridDim.x >> 1);
}

template <int kBlockN, bool kAMN, bool kBMN>
__global__ void __launch_bounds__(kThreads, 1)
gemm_bf16_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                 const __grid_constant__ CUtensorMap tmap_c, const __grid_constant__ CUtensorMap tmap_aux, const GemmArgs args) {
  gemm_body<kBlockN, kAMN, kBMN, false>(tmap_a, tmap_b, tmap_c, tmap_aux, nullptr, args, blockIdx.x, gridDim.x);
}

// ---- all-gather fused GEMM --------------------------------------------------------------------------
// The first n_push CTAs push this rank's row shard of A into every peer's copy of the gather buffer (posted NVLink
// stores) and count arrivals in the peer's a_ready[rank]; the other n_gemm CTAs run the GEMM over the gathered A,
// starting with the own shard and consuming remote shards in the order they are sent.
struct AgArgs {
  uint8_t* peer_a[16];         // base of the [T, lda] gather buffer on every rank (peer-mapped)
  GemmCtrl* peer_ctrl[16];
  int rank, world, n_gemm;
  int rows_per_shard;
  int64_t row_bytes, ld_bytes;
  unsigned int epoch;
};

// One thread per pusher CTA drives a ring of bulk copies: global (local HBM) -> smem -> peer global over NVLink.
// kPushBufs x 16 KiB in flight per CTA; no registers or generic-proxy traffic involved.
constexpr int kPushBufs = 12;
constexpr int kPushLookahead = 8;
constexpr uint32_t kPushChunk = 16384;

__device__ __forceinline__ void ag_pusher(const AgArgs& ag, const int c, const int nc) {
  extern __shared__ uint8_t smem_raw[];
  GemmCtrl* me = ag.peer_ctrl[ag.rank];
  if (c == 0 && threadIdx.x < ag.world && (int)threadIdx.x != ag.rank)
    st_release_sys_u32(&ag.peer_ctrl[threadIdx.x]->entered[ag.rank], ag.epoch);   // my shard is final, my old gather data is dead
  if (threadIdx.x != 0) return;
  red_add_release_sys(&me->a_ready[ag.rank], 1u);                                  // own shard needs no copy
  const uint32_t buf0 = (smem_u32(smem_raw) + 127u) & ~127u;
  const uint32_t bar0 = buf0 + kPushBufs * kPushChunk;
  for (int b = 0; b < kPushBufs; ++b) mbar_init(bar0 + 8u * b, 1);
  fence_mbar_init();
  // the shard as a list of contiguous pieces: whole shard if rows are dense, else one piece per row
  const bool dense = ag.ld_bytes == ag.row_bytes;
  const int64_t piece_bytes = dense ? (int64_t)ag.rows_per_shard * ag.row_bytes : ag.row_bytes;
  const int64_t n_pieces = dense ? 1 : ag.rows_per_shard;
  const int64_t chunks_per_piece = (piece_bytes + kPushChunk - 1) / kPushChunk;
  const int64_t n_chunks = n_pieces * chunks_per_piece;
  const int64_t per = (n_chunks + nc - 1) / nc;
  const int64_t lo = (int64_t)c * per, hi = (lo + per < n_chunks) ? lo + per : n_chunks;
  const int64_t shard_off = (int64_t)ag.rank * ag.rows_per_shard * ag.ld_bytes;
  auto locate = [&](int64_t ch, int64_t& off, uint32_t& bytes) {
    const int64_t piece = ch / chunks_per_piece, k = ch - piece * chunks_per_piece;
    const int64_t in_piece = k * kPushChunk;
    off = shard_off + piece * ag.ld_bytes + in_piece;
    const int64_t left = piece_bytes - in_piece;
    bytes = (uint32_t)(left < (int64_t)kPushChunk ? left : (int64_t)kPushChunk);
  };
  uint32_t round = 0;                 // chunks issued so far (ring position / parity source), continues across peers
  for (int i = 1; i < ag.world; ++i) {
    const int p = (ag.rank - i + ag.world) % ag.world;         // peer p consumes my shard i-th
    wait_counter_sys(&me->entered[p], ag.epoch, "entered", p);
    const uint8_t* src = ag.peer_a[ag.rank];
    uint8_t* dst = ag.peer_a[p];
    const int64_t n = hi > lo ? hi - lo : 0;
    for (int64_t j = 0; j < n + kPushLookahead; ++j) {
      if (j < n) {
        const uint32_t slot = (round + (uint32_t)j) % kPushBufs;
        tma_store_wait_read<kPushBufs - kPushLookahead - 1>();            // the store that last used this slot has drained it
        int64_t off; uint32_t bytes;
        locate(lo + j, off, bytes);
        mbar_expect_tx(bar0 + 8u * slot, bytes);
        bulk_load(buf0 + slot * kPushChunk, src + off, bytes, bar0 + 8u * slot);
      }
      if (j >= kPushLookahead) {
        const uint32_t seq = round + (uint32_t)(j - kPushLookahead);
        const uint32_t slot = seq % kPushBufs;
        mbar_wait(bar0 + 8u * slot, (seq / kPushBufs) & 1u);
        int64_t off; uint32_t bytes;
        locate(lo + j - kPushLookahead, off, bytes);
        bulk_store(dst + off, buf0 + slot * kPushChunk, bytes);
        tma_store_commit();
      } else {
        tma_store_commit();                                               // keep group numbering aligned with the ring
      }
    }
    round += (uint32_t)n;
    tma_store_wait<0>();                                                  // my slice has landed at peer p
    __threadfence_system();
    red_add_release_sys(&ag.peer_ctrl[p]->a_ready[ag.rank], 1u);
  }
}

template <int kBlockN, bool kBMN>
__global__ void __launch_bounds__(kThreads, 1)
ag_gemm_bf16_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                    const __grid_constant__ CUtensorMap tmap_c, const __grid_constant__ CUtensorMap tmap_aux, const GemmArgs args,
                    const AgArgs ag) {
  // Pushers take the lowest block indices: they are what the *peers* wait for, so they must be resident even when something
  // else (an NCCL kernel on another stream) holds a few SMs and part of this grid has to queue.  A GEMM CTA that queues
  // only delays its own tiles.
  const int n_push = (int)gridDim.x - ag.n_gemm;
  if ((int)blockIdx.x < n_push) ag_pusher(ag, blockIdx.x, n_push);
  else gemm_body<kBlockN, false, kBMN, false>(tmap_a, tmap_b, tmap_c, tmap_aux, nullptr, args, blockIdx.x - n_push, ag.n_gemm);
}

// ---- GEMM fused reduce-scatter ------------------------------------------------------------------------
// Every rank multiplies its K-slice; the bf16 partial tile of row shard d goes straight from the epilogue's
// staging smem into rank d's receive slot [me] (TMA store over NVLink).  Once a rank has received all
// partials of its shard (rs_count), all of its CTAs sum the W slots in fp32 (+bias) into the output.
struct RsArgs {
  const __nv_bfloat16* recv;       // my receive slots [world][rows_per_shard][N]
  void* out;                       // [rows_per_shard, N]
  const void* bias;
  int64_t ldo;
  int out_fp32, bias_fp32;
  GemmCtrl* peer_ctrl[8];
  unsigned int count_target;       // rs_count[s] value that means "all partials of this call from s are here"
};

template <int kBlockN, bool kAMN, bool kBMN>
__global__ void __launch_bounds__(kThreads, 1)
gemm_rs_bf16_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                    const __grid_constant__ TmapArray tmap_rs, const GemmArgs args, const RsArgs rs) {
  GemmCtrl* me = rs.peer_ctrl[args.rs_rank];
  if (blockIdx.x == 0 && threadIdx.x < args.rs_world && (int)threadIdx.x != args.rs_rank)
    st_release_sys_u32(&rs.peer_ctrl[threadIdx.x]->entered[args.rs_rank], args.rs_epoch);
  gemm_body<kBlockN, kAMN, kBMN, true>(tmap_a, tmap_b, tmap_rs.m[0], tmap_rs.m[0], tmap_rs.m, args, blockIdx.x, gridDim.x);
  // ---- reduction of my shard ----
  if (threadIdx.x < args.rs_world) wait_counter_sys(&me->rs_count[threadIdx.x], rs.count_target, "rs_count", threadIdx.x);
  __syncthreads();
  const int N = args.N, rows = args.rows_per_shard;
  const int64_t vec_per_row = N >> 3;
  const int64_t total = (int64_t)rows * vec_per_row;
  const int64_t slot = (int64_t)rows * N;
  for (int64_t v = (int64_t)blockIdx.x * kThreads + threadIdx.x; v < total; v += (int64_t)gridDim.x * kThreads) {
    const int64_t row = v / vec_per_row;
    const int col = (int)(v % vec_per_row) * 8;
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int s = 0; s < args.rs_world; ++s) {
      const uint4 u = ldg_stream_u4(reinterpret_cast<const uint4*>(rs.recv + s * slot + row * N + col));
      acc[0] += bf16lo(u.x); acc[1] += bf16hi(u.x); acc[2] += bf16lo(u.y); acc[3] += bf16hi(u.y);
      acc[4] += bf16lo(u.z); acc[5] += bf16hi(u.z); acc[6] += bf16lo(u.w); acc[7] += bf16hi(u.w);
    }
    if (rs.bias != nullptr) {
#pragma unroll
      for (int j = 0; j < 8; ++j)
        acc[j] += rs.bias_fp32 ? reinterpret_cast<const float*>(rs.bias)[col + j] : __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(rs.bias)[col + j]);
    }
    if (rs.out_fp32) {
      float* o = reinterpret_cast<float*>(rs.out) + row * rs.ldo + col;
      *reinterpret_cast<float4*>(o) = make_float4(acc[0], acc[1], acc[2], acc[3]);
      *reinterpret_cast<float4*>(o + 4) = make_float4(acc[4], acc[5], acc[6], acc[7]);
    } else {
      __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(rs.out) + row * rs.ldo + col;
      *reinterpret_cast<uint4*>(o) = make_uint4(pack_bf16(acc[0], acc[1]), pack_bf16(acc[2], acc[3]), pack_bf16(acc[4], acc[5]), pack_bf16(acc[6], acc[7]));
    }
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

template <bool kAMN, bool kBMN>
int launch2(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& tc, const CUtensorMap& taux, const GemmArgs& a, int sms,
            cudaStream_t stream) {
  using C = Cfg<256, true>;
  auto* kern = gemm2_bf16_kernel<kAMN, kBMN>;
  DETB200_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::kSmemBytes));
  const int mt2 = (a.M + 2 * kBlockM - 1) / (2 * kBlockM), nt = (a.N + 255) / 256;
  int pairs = sms / 2;
  if (mt2 * nt < pairs) pairs = mt2 * nt;
  kern<<<2 * pairs, kThreads, C::kSmemBytes, stream>>>(ta, tb, tc, taux, a);
  DETB200_LAUNCH_CHECK();
  return 0;
}

bool two_cta_enabled() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("DETB200_GEMM_2CTA");
    v = (e == nullptr || e[0] != '0') ? 1 : 0;
  }
  return v == 1;
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

int detb200_make_tmap_4d_bf16(CUtensorMap* out, const void* base, const uint64_t dims[4], const uint64_t strides[3], const uint32_t box[4]) {
  std::call_once(g_encode_once, resolve_encode);
  if (g_encode == nullptr) {
    detb200_set_error("cuTensorMapEncodeTiled is not available from this driver");
    return DETB200_ERR_UNSUPPORTED;
  }
  if ((reinterpret_cast<uintptr_t>(base) & 15u) != 0 || (strides[0] & 15u) || (strides[1] & 15u) || (strides[2] & 15u)) {
    detb200_set_error("TMA operand must be 16-byte aligned with 16-byte-multiple strides");
    return DETB200_ERR_BAD_ARG;
  }
  const cuuint64_t d[4] = {dims[0], dims[1], dims[2], dims[3]};
  const cuuint64_t st[3] = {strides[0], strides[1], strides[2]};
  const cuuint32_t bx[4] = {box[0], box[1], box[2], box[3]};
  const cuuint32_t estr[4] = {1, 1, 1, 1};
  const CUresult r = g_encode(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), d, st, bx, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                              CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    detb200_set_error("cuTensorMapEncodeTiled(4d) failed with CUresult %d", (int)r);
    return DETB200_ERR_BAD_ARG;
  }
  return 0;
}

// fp32, no swizzle (dense [box3][box2][box1][box0] image in smem): used for TMA reduce-add into fp32 accumulators
int detb200_make_tmap_4d_f32_plain(CUtensorMap* out, const void* base, const uint64_t dims[4], const uint64_t strides[3], const uint32_t box[4]) {
  std::call_once(g_encode_once, resolve_encode);
  if (g_encode == nullptr) {
    detb200_set_error("cuTensorMapEncodeTiled is not available from this driver");
    return DETB200_ERR_UNSUPPORTED;
  }
  const cuuint64_t d[4] = {dims[0], dims[1], dims[2], dims[3]};
  const cuuint64_t st[3] = {strides[0], strides[1], strides[2]};
  const cuuint32_t bx[4] = {box[0], box[1], box[2], box[3]};
  const cuuint32_t estr[4] = {1, 1, 1, 1};
  const CUresult r = g_encode(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, const_cast<void*>(base), d, st, bx, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                              CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    detb200_set_error("cuTensorMapEncodeTiled(4d f32) failed with CUresult %d", (int)r);
    return DETB200_ERR_BAD_ARG;
  }
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
  // CTA pairs (cta_group::2) for anything with at least one full 256x256 tile; the plain kernel otherwise and for the
  // shard-ordered (gather-fused) walks
  const bool two_cta = bn == 256 && M > kBlockM && a_ready == nullptr && m_block_rotation == 0 && max_ctas <= 0 && two_cta_enabled();
  CUtensorMap ta, tb;
  int rc;
  if (!a_mn_major) rc = detb200_make_tmap_2d(&ta, 0, A, (uint64_t)K, (uint64_t)M, (uint64_t)lda * 2, kBlockK, kBlockM);
  else rc = detb200_make_tmap_2d(&ta, 0, A, (uint64_t)M, (uint64_t)K, (uint64_t)lda * 2, 64, kBlockK);
  if (rc) return rc;
  if (!b_mn_major) rc = detb200_make_tmap_2d(&tb, 0, B, (uint64_t)K, (uint64_t)N, (uint64_t)ldb * 2, kBlockK, two_cta ? bn / 2 : bn);
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
  if (two_cta) {
    if (!a_mn_major && !b_mn_major) return launch2<false, false>(ta, tb, tc, taux, g, sms, stream);
    if (!a_mn_major && b_mn_major) return launch2<false, true>(ta, tb, tc, taux, g, sms, stream);
    if (a_mn_major && !b_mn_major) return launch2<true, false>(ta, tb, tc, taux, g, sms, stream);
    return launch2<true, true>(ta, tb, tc, taux, g, sms, stream);
  }
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

// ---- fused collectives: host entry points -----------------------------------------------------------
DETB200_API int64_t detb200_gemm_ctrl_bytes(void) { return (int64_t)sizeof(GemmCtrl); }

namespace {
int sm_count(int* out) {
  int dev = 0;
  DETB200_CHECK(cudaGetDevice(&dev));
  DETB200_CHECK(cudaDeviceGetAttribute(out, cudaDevAttrMultiProcessorCount, dev));
  return 0;
}
}  // namespace

// Gather-fused GEMM.  `peer_g[r]` is rank r's copy of the *gathered* operand's buffer (row-sharded: rank r owns rows
// [r*rows/world, ...)); `other` is the local operand.
//   gather_b = 0: the gathered operand is A ([T, ld_g], K-major): sequence-parallel activations.   other = B.
//   gather_b = 1: the gathered operand is B: ZeRO-3 style weight shards.  gathered_mn_major = 0: B is [N, K] sharded over N
//                 (forward, Y = X W^T); 1: B is stored [K, N] and sharded over its rows = the reduction dimension (dX = dY W):
//                 the k-loop then starts at the own shard.  other = A ([M, K], K-major).
DETB200_API int detb200_ag_gemm_bf16_ex(void* const* peer_g, void* const* peer_ctrl, int rank, int world, const void* other, void* C,
                                        const void* bias, void* aux, int64_t M, int64_t N, int64_t K, int64_t ld_g, int64_t ld_other,
                                        int64_t ldc, int64_t ldaux, int gather_b, int gathered_mn_major, int other_mn_major, int out_fp32,
                                        int bias_fp32, int epilogue, unsigned int epoch, unsigned int ready_value, int n_push_ctas,
                                        cudaStream_t stream) {
  const int bn = (N > 128) ? 256 : 128;
  const int mode = !gather_b ? 0 : (gathered_mn_major ? 2 : 1);
  const int64_t g_rows = mode == 0 ? M : mode == 1 ? N : K;                   // rows of the gathered operand's storage
  const int64_t g_cols = mode == 2 ? N : K;                                   // its contiguous dimension
  const int align = mode == 0 ? kBlockM : mode == 1 ? bn : kBlockK;
  if (world < 1 || world > 16 || rank < 0 || rank >= world || g_rows % world != 0 || (g_rows / world) % align != 0) {
    detb200_set_error("ag_gemm: world %d rank %d: the gathered operand has %lld rows; rows per rank must be a multiple of %d", world, rank,
                      (long long)g_rows, align);
    return DETB200_ERR_BAD_ARG;
  }
  if (epilogue < 0 || epilogue > EPI_DRELU || (epilogue > EPI_RELU && aux == nullptr) || (gather_b && other_mn_major) ||
      (!gather_b && gathered_mn_major)) {
    detb200_set_error("ag_gemm: unsupported combination (derivative epilogues need aux; the A operand must be K-major)");
    return DETB200_ERR_BAD_ARG;
  }
  const int elem_c = out_fp32 ? 4 : 2;
  const int rows = (int)(g_rows / world);
  const void* A = gather_b ? other : peer_g[rank];
  const void* B = gather_b ? peer_g[rank] : other;
  const int64_t lda = gather_b ? ld_other : ld_g, ldb = gather_b ? ld_g : ld_other;
  const int b_mn = gather_b ? gathered_mn_major : other_mn_major;
  CUtensorMap ta, tb, tc, taux;
  int rc = detb200_make_tmap_2d(&ta, 0, A, (uint64_t)K, (uint64_t)M, (uint64_t)lda * 2, kBlockK, kBlockM);
  if (rc) return rc;
  if (!b_mn) rc = detb200_make_tmap_2d(&tb, 0, B, (uint64_t)K, (uint64_t)N, (uint64_t)ldb * 2, kBlockK, bn);
  else rc = detb200_make_tmap_2d(&tb, 0, B, (uint64_t)N, (uint64_t)K, (uint64_t)ldb * 2, 64, kBlockK);
  if (rc) return rc;
  rc = detb200_make_tmap_2d(&tc, out_fp32, C, (uint64_t)N, (uint64_t)M, (uint64_t)ldc * elem_c, out_fp32 ? 32 : 64, 32);
  if (rc) return rc;
  if (aux != nullptr && epilogue >= EPI_GELU_ERF && epilogue <= EPI_RELU) {
    rc = detb200_make_tmap_2d(&taux, 0, aux, (uint64_t)N, (uint64_t)M, (uint64_t)ldaux * 2, 64, 32);
    if (rc) return rc;
  } else {
    taux = tc;
  }
  int sms = kNumSMs;
  if ((rc = sm_count(&sms))) return rc;
  if (world == 1) n_push_ctas = 0;
  else if (n_push_ctas <= 0) n_push_ctas = 4;
  if (n_push_ctas >= sms) n_push_ctas = sms / 2;
  GemmCtrl* me = reinterpret_cast<GemmCtrl*>(peer_ctrl[rank]);
  GemmArgs g;
  std::memset(&g, 0, sizeof(g));
  g.C = C; g.bias = bias; g.aux = aux; g.M = (int)M; g.N = (int)N; g.K = (int)K; g.ldc = ldc; g.ldaux = ldaux;
  g.out_fp32 = out_fp32; g.bias_fp32 = bias_fp32; g.epilogue = epilogue; g.accumulate = 0;
  g.a_ready = world > 1 ? me->a_ready : nullptr; g.ready_value = ready_value; g.rows_per_shard = rows;
  g.gather_mode = mode;
  g.m_block_rotation = rank * (rows / align);
  AgArgs ag;
  std::memset(&ag, 0, sizeof(ag));
  for (int r = 0; r < world; ++r) { ag.peer_a[r] = reinterpret_cast<uint8_t*>(peer_g[r]); ag.peer_ctrl[r] = reinterpret_cast<GemmCtrl*>(peer_ctrl[r]); }
  ag.rank = rank; ag.world = world; ag.rows_per_shard = rows; ag.row_bytes = g_cols * 2; ag.ld_bytes = ld_g * 2; ag.epoch = epoch;
  const int mt = (int)((M + kBlockM - 1) / kBlockM), nt = (int)((N + bn - 1) / bn);
  int n_gemm = sms - n_push_ctas;
  if (mt * nt < n_gemm) n_gemm = mt * nt;
  ag.n_gemm = n_gemm;
  const int grid = n_gemm + n_push_ctas;
#define DETB200_AG_LAUNCH(BN, BMN)                                                                                   \
  do {                                                                                                               \
    auto* kern = ag_gemm_bf16_kernel<BN, BMN>;                                                                       \
    DETB200_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg<BN>::kSmemBytes));     \
    kern<<<grid, kThreads, Cfg<BN>::kSmemBytes, stream>>>(ta, tb, tc, taux, g, ag);                                  \
  } while (0)
  if (bn == 256) { if (b_mn) DETB200_AG_LAUNCH(256, true); else DETB200_AG_LAUNCH(256, false); }
  else { if (b_mn) DETB200_AG_LAUNCH(128, true); else DETB200_AG_LAUNCH(128, false); }
#undef DETB200_AG_LAUNCH
  DETB200_LAUNCH_CHECK();
  return 0;
}

DETB200_API int detb200_ag_gemm_bf16(void* const* peer_a, void* const* peer_ctrl, int rank, int world, const void* B, void* C,
                                     const void* bias, void* aux, int64_t T, int64_t N, int64_t K, int64_t lda, int64_t ldb, int64_t ldc,
                                     int64_t ldaux, int b_mn_major, int out_fp32, int bias_fp32, int epilogue, unsigned int epoch,
                                     unsigned int ready_value, int n_push_ctas, cudaStream_t stream) {
  return detb200_ag_gemm_bf16_ex(peer_a, peer_ctrl, rank, world, B, C, bias, aux, T, N, K, lda, ldb, ldc, ldaux, 0, 0, b_mn_major, out_fp32,
                                 bias_fp32, epilogue, epoch, ready_value, n_push_ctas, stream);
}

DETB200_API int detb200_gemm_rs_bf16(const void* A, const void* B, void* const* peer_recv, void* const* peer_ctrl, int rank, int world,
                                     void* out, const void* bias, int64_t T, int64_t N, int64_t K, int64_t lda, int64_t ldb, int64_t ldo,
                                     int a_mn_major, int b_mn_major, int out_fp32, int bias_fp32, unsigned int epoch,
                                     unsigned int count_base, cudaStream_t stream) {
  if (world < 1 || world > 8 || rank < 0 || rank >= world || T % world != 0 || (T / world) % kBlockM != 0 || N % 8 != 0) {
    detb200_set_error("gemm_rs: world %d (max 8) rank %d T %lld N %lld: rows per rank must be a multiple of %d and N of 8", world, rank,
                      (long long)T, (long long)N, kBlockM);
    return DETB200_ERR_BAD_ARG;
  }
  const int bn = (N > 128) ? 256 : 128;
  const int rows = (int)(T / world);
  CUtensorMap ta, tb;
  int rc;
  if (!a_mn_major) rc = detb200_make_tmap_2d(&ta, 0, A, (uint64_t)K, (uint64_t)T, (uint64_t)lda * 2, kBlockK, kBlockM);
  else rc = detb200_make_tmap_2d(&ta, 0, A, (uint64_t)T, (uint64_t)K, (uint64_t)lda * 2, 64, kBlockK);
  if (rc) return rc;
  if (!b_mn_major) rc = detb200_make_tmap_2d(&tb, 0, B, (uint64_t)K, (uint64_t)N, (uint64_t)ldb * 2, kBlockK, bn);
  else rc = detb200_make_tmap_2d(&tb, 0, B, (uint64_t)N, (uint64_t)K, (uint64_t)ldb * 2, 64, kBlockK);
  if (rc) return rc;
  TmapArray trs;
  std::memset(&trs, 0, sizeof(trs));
  for (int d = 0; d < world; ++d) {
    uint8_t* base = reinterpret_cast<uint8_t*>(peer_recv[d]) + (int64_t)rank * rows * N * 2;     // slot [me] on rank d
    rc = detb200_make_tmap_2d(&trs.m[d], 0, base, (uint64_t)N, (uint64_t)rows, (uint64_t)N * 2, 64, 32);
    if (rc) return rc;
  }
  for (int d = world; d < 8; ++d) trs.m[d] = trs.m[0];
  GemmArgs g;
  std::memset(&g, 0, sizeof(g));
  g.M = (int)T; g.N = (int)N; g.K = (int)K; g.rows_per_shard = rows;
  g.m_block_rotation = ((rank + 1) % world) * (rows / kBlockM);      // remote shards first, own shard last
  g.rs_world = world; g.rs_rank = rank; g.rs_epoch = epoch;
  GemmCtrl* me = reinterpret_cast<GemmCtrl*>(peer_ctrl[rank]);
  g.rs_entered = me->entered;
  RsArgs rs;
  std::memset(&rs, 0, sizeof(rs));
  for (int d = 0; d < world; ++d) {
    rs.peer_ctrl[d] = reinterpret_cast<GemmCtrl*>(peer_ctrl[d]);
    g.rs_peer_count[d] = &rs.peer_ctrl[d]->rs_count[rank];
  }
  rs.recv = reinterpret_cast<const __nv_bfloat16*>(peer_recv[rank]);
  rs.out = out; rs.bias = bias; rs.ldo = ldo; rs.out_fp32 = out_fp32; rs.bias_fp32 = bias_fp32;
  const int mt = (int)(T / kBlockM), nt = (int)((N + bn - 1) / bn);
  rs.count_target = count_base + (unsigned int)((rows / kBlockM) * nt * 4);
  int sms = kNumSMs;
  if ((rc = sm_count(&sms))) return rc;
  const int grid = mt * nt < sms ? mt * nt : sms;
#define DETB200_RS_LAUNCH(BN, AMN, BMN)                                                                              \
  do {                                                                                                               \
    auto* kern = gemm_rs_bf16_kernel<BN, AMN, BMN>;                                                                  \
    DETB200_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg<BN>::kSmemBytes));     \
    kern<<<grid, kThreads, Cfg<BN>::kSmemBytes, stream>>>(ta, tb, trs, g, rs);                                       \
  } while (0)
#define DETB200_RS_DISPATCH(BN)                                                        \
  do {                                                                                 \
    if (!a_mn_major && !b_mn_major) DETB200_RS_LAUNCH(BN, false, false);               \
    else if (!a_mn_major && b_mn_major) DETB200_RS_LAUNCH(BN, false, true);            \
    else if (a_mn_major && !b_mn_major) DETB200_RS_LAUNCH(BN, true, false);            \
    else DETB200_RS_LAUNCH(BN, true, true);                                            \
  } while (0)
  if (bn == 256) DETB200_RS_DISPATCH(256); else DETB200_RS_DISPATCH(128);
#undef DETB200_RS_DISPATCH
#undef DETB200_RS_LAUNCH
  DETB200_LAUNCH_CHECK();
  return 0;
}

// arrivals one gemm_rs call adds to each rs_count entry of the shard owner (for the caller's running total)
DETB200_API int detb200_gemm_rs_arrivals(int64_t T, int64_t N, int world) {
  const int bn = (N > 128) ? 256 : 128;
  return (int)((T / world / kBlockM) * ((N + bn - 1) / bn) * 4);
}
