// Flash attention forward on tcgen05: S = Q K^T and O += P V both run on the 5th-gen tensor cores,
// S/P/O live in tensor memory, K/V blocks are streamed by TMA, the online softmax runs in registers.
//
//   warp 0      TMA producer   Q tile once, then (K_j, V_j) pairs into a 2-deep smem ring
//   warp 1      MMA issuer     QK(j+1) is issued as soon as the softmax has *read* S(j), so it overlaps the
//                              exponentials; PV(j) follows when P(j) has been written to TMEM
//   warps 2..5  softmax        one thread per query row: tcgen05.ld S -> running max / sum (lazy rescale of O
//                              only when the max moved by > 2^8) -> P as packed bf16 -> tcgen05.st
//
// Two CTAs are resident per SM (96 KiB smem, 256 TMEM columns each) so one CTA's softmax overlaps the other's
// MMAs.  Layout-agnostic: Q/K/V/O are addressed as (batch, seq, head, dim) with arbitrary strides (dim contiguous),
// which covers [B,H,S,D], [B,S,H,D] and packed QKV.  GQA via h_kv = h / (H / H_kv).
// Replaces: the attention of the GPT-2 / GPT-NeoX examples (K13; Megatron softmax+bmm or flash-attn there).
#include "sm100.cuh"

using namespace sm100;

namespace {

constexpr int kBM = 128;          // query rows per CTA
constexpr int kBN = 64;           // keys per block
constexpr int kAThreads = 192;
constexpr int kStagesKV = 2;
constexpr float kRescaleThreshold = 8.0f;   // log2 units

struct AttnArgs {
  __nv_bfloat16* O;
  float* lse;                      // [B, H, S_q]
  int B, H, H_kv, S_q, S_kv;
  int64_t o_sb, o_ss, o_sh;        // element strides of O
  float scale_log2;                // softmax scale * log2(e)
  int causal;
};

template <int D> struct ACfg {
  static constexpr int kQBytes = kBM * D * 2;
  static constexpr int kKVBytes = kBN * D * 2;           // one of K or V
  static constexpr int kStageBytes = 2 * kKVBytes;
  static constexpr int kSmemBytes = kQBytes + kStagesKV * kStageBytes + 1024 + 128;
  static constexpr int kColS = 0, kColP = kBN, kColO = 128;
  static constexpr int kTmemCols = 256;
};

template <int D>
__global__ void __launch_bounds__(kAThreads, 2)
attn_fwd_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_k,
                const __grid_constant__ CUtensorMap tmap_v, const AttnArgs args) {
  using C = ACfg<D>;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t sQ = smem_base;
  const uint32_t sKV = smem_base + C::kQBytes;
  const uint32_t bar_base = sKV + kStagesKV * C::kStageBytes;
  const uint32_t q_full = bar_base;
  auto kv_full = [&](int s) { return bar_base + 8u * (1 + s); };
  auto kv_empty = [&](int s) { return bar_base + 8u * (1 + kStagesKV + s); };
  const uint32_t s_full = bar_base + 8u * (1 + 2 * kStagesKV);
  const uint32_t s_empty = s_full + 8u, p_full = s_full + 16u, o_done = s_full + 24u;
  const uint32_t tmem_slot = s_full + 32u;
  volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_slot - smem_u32(smem_raw)));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int q_blocks = (args.S_q + kBM - 1) / kBM;
  const int q_blk = q_blocks - 1 - (int)blockIdx.x;               // longest (causal) tiles first
  const int q0 = q_blk * kBM;
  const int bh = blockIdx.y, b = bh / args.H, h = bh % args.H;
  const int h_kv = h / (args.H / args.H_kv);
  const int causal_off = args.S_kv - args.S_q;
  int n_kv = (args.S_kv + kBN - 1) / kBN;
  if (args.causal) {
    const int lim = (q0 + kBM + causal_off + kBN - 1) / kBN;
    n_kv = lim < n_kv ? lim : n_kv;
    if (n_kv < 1) n_kv = 1;
  }

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmap_q);
    prefetch_tmap(&tmap_k);
    prefetch_tmap(&tmap_v);
  }
  if (warp == 1 && lane == 0) {
    mbar_init(q_full, 1);
    for (int s = 0; s < kStagesKV; ++s) { mbar_init(kv_full(s), 1); mbar_init(kv_empty(s), 1); }
    mbar_init(s_full, 1);
    mbar_init(s_empty, 128);
    mbar_init(p_full, 128);
    mbar_init(o_done, 1);
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, C::kTmemCols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;
  const uint32_t tS = tmem_base + C::kColS, tP = tmem_base + C::kColP, tO = tmem_base + C::kColO;

  if (warp == 0) {
    if (lane == 0) {
      mbar_expect_tx(q_full, C::kQBytes);
#pragma unroll
      for (int hh = 0; hh < D / 64; ++hh) tma_load_4d(sQ + hh * (kBM * 128), &tmap_q, q_full, hh * 64, h, q0, b);
      for (int j = 0; j < n_kv; ++j) {
        const int st = j % kStagesKV;
        mbar_wait(kv_empty(st), ((j / kStagesKV) & 1) ^ 1u);
        mbar_expect_tx(kv_full(st), C::kStageBytes);
        const uint32_t sK = sKV + st * C::kStageBytes, sV = sK + C::kKVBytes;
#pragma unroll
        for (int hh = 0; hh < D / 64; ++hh) {
          tma_load_4d(sK + hh * (kBN * 128), &tmap_k, kv_full(st), hh * 64, h_kv, j * kBN, b);
          tma_load_4d(sV + hh * (kBN * 128), &tmap_v, kv_full(st), hh * 64, h_kv, j * kBN, b);
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc_qk = idesc_bf16_f32(kBM, kBN, 0, 0);
      constexpr uint32_t idesc_pv = idesc_bf16_f32(kBM, D, 0, 1);
      auto issue_qk = [&](int j) {
        const uint32_t sK = sKV + (j % kStagesKV) * C::kStageBytes;
#pragma unroll
        for (int k = 0; k < D / 16; ++k) {
          const uint64_t da = smem_desc_sw128(sQ + (k / 4) * (kBM * 128) + (k % 4) * 32, 16, 1024);
          const uint64_t db = smem_desc_sw128(sK + (k / 4) * (kBN * 128) + (k % 4) * 32, 16, 1024);
          mma_f16_ss(tS, da, db, idesc_qk, k != 0 ? 1u : 0u);
        }
        tc_commit(s_full);
      };
      mbar_wait(q_full, 0);
      mbar_wait(kv_full(0), 0);
      tc_fence_after();
      issue_qk(0);
      for (int j = 0; j < n_kv; ++j) {
        if (j + 1 < n_kv) {
          mbar_wait(kv_full((j + 1) % kStagesKV), ((j + 1) / kStagesKV) & 1);
          mbar_wait(s_empty, j & 1);                  // softmax(j) holds S(j) in registers
          tc_fence_after();
          issue_qk(j + 1);
        }
        mbar_wait(p_full, j & 1);
        tc_fence_after();
        const uint32_t sV = sKV + (j % kStagesKV) * C::kStageBytes + C::kKVBytes;
#pragma unroll
        for (int k = 0; k < kBN / 16; ++k) {
          // V block is [keys, D] with D contiguous: MN-major B; 64-wide d groups are kBN*128 B apart
          const uint64_t dv = smem_desc_sw128(sV + k * 2048, kBN * 128, 1024);
          mma_f16_ts(tO, tP + k * 8, dv, idesc_pv, (j | k) != 0 ? 1u : 0u);
        }
        tc_commit(kv_empty(j % kStagesKV));
        tc_commit(o_done);
      }
    }
  } else {
    const int q = warp & 3;
    const int row = q * 32 + lane;
    const int qi = q0 + row;
    const uint32_t lane_off = (uint32_t)(q * 32) << 16;
    float m_used = -INFINITY, l = 0.f;
    for (int j = 0; j < n_kv; ++j) {
      mbar_wait(s_full, j & 1);
      tc_fence_after();
      uint32_t r[64];
      tmem_ld_32x32(tS + lane_off, r);
      tmem_ld_32x32(tS + lane_off + 32, r + 32);
      tmem_ld_wait();
      tc_fence_before();
      mbar_arrive(s_empty);
      float* s = reinterpret_cast<float*>(r);
      const int kv0 = j * kBN;
      const bool need_mask = (args.causal && kv0 + kBN - 1 > q0 + causal_off) || (kv0 + kBN > args.S_kv);
      if (need_mask) {
        const int lim = args.causal ? min(qi + causal_off, args.S_kv - 1) : args.S_kv - 1;   // last visible key
#pragma unroll
        for (int c = 0; c < 64; ++c) if (kv0 + c > lim) s[c] = -INFINITY;
      }
      float mx = s[0];
#pragma unroll
      for (int c = 1; c < 64; ++c) mx = fmaxf(mx, s[c]);
      const float m_new = fmaxf(m_used, mx * args.scale_log2);
      const bool need = m_new > m_used + kRescaleThreshold;       // also true for the first finite block
      const bool any = __any_sync(0xffffffffu, need);
      if (j > 0) {
        mbar_wait(o_done, (j - 1) & 1);                           // PV(j-1) retired: P reusable, O stable
        tc_fence_after();
      }
      if (any) {
        const float alpha = need ? exp2f(m_used - m_new) : 1.f;   // exp2(-inf) = 0 on the first block
        if (need) { m_used = m_new; l *= alpha; }
        if (j > 0) {
#pragma unroll
          for (int c = 0; c < D; c += 32) {
            uint32_t o[32];
            tmem_ld_32x32(tO + lane_off + c, o);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
            tmem_st_32x32(tO + lane_off + c, o);
          }
        }
      }
      uint32_t pk[32];
      float sum = 0.f;
      const float mm = (m_used == -INFINITY) ? 0.f : m_used;       // fully masked row so far: p = exp2(-inf) = 0
#pragma unroll
      for (int c = 0; c < 64; c += 2) {
        const float p0 = exp2f(s[c] * args.scale_log2 - mm), p1 = exp2f(s[c + 1] * args.scale_log2 - mm);
        sum += p0 + p1;
        pk[c >> 1] = pack_bf16(p0, p1);
      }
      l += sum;
      tmem_st_32x32(tP + lane_off, pk);
      tmem_st_wait();
      tc_fence_before();
      mbar_arrive(p_full);
    }
    mbar_wait(o_done, (n_kv - 1) & 1);
    tc_fence_after();
    const float inv_l = l > 0.f ? 1.f / l : 0.f;
    const bool row_ok = qi < args.S_q;
    __nv_bfloat16* op = args.O + (int64_t)b * args.o_sb + (int64_t)qi * args.o_ss + (int64_t)h * args.o_sh;
#pragma unroll
    for (int c = 0; c < D; c += 32) {
      uint32_t o[32];
      tmem_ld_32x32(tO + lane_off + c, o);
      tmem_ld_wait();
      if (row_ok) {
#pragma unroll
        for (int i = 0; i < 32; i += 8) {
          uint4 u;
          u.x = pack_bf16(__uint_as_float(o[i]) * inv_l, __uint_as_float(o[i + 1]) * inv_l);
          u.y = pack_bf16(__uint_as_float(o[i + 2]) * inv_l, __uint_as_float(o[i + 3]) * inv_l);
          u.z = pack_bf16(__uint_as_float(o[i + 4]) * inv_l, __uint_as_float(o[i + 5]) * inv_l);
          u.w = pack_bf16(__uint_as_float(o[i + 6]) * inv_l, __uint_as_float(o[i + 7]) * inv_l);
          *reinterpret_cast<uint4*>(op + c + i) = u;
        }
      }
    }
    if (row_ok && args.lse != nullptr)
      args.lse[((int64_t)b * args.H + h) * args.S_q + qi] = (l > 0.f) ? m_used * 0.6931471805599453f + logf(l) : -INFINITY;
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, C::kTmemCols);
  }
}

template <int D>
int launch_fwd(const CUtensorMap& tq, const CUtensorMap& tk, const CUtensorMap& tv, const AttnArgs& a, cudaStream_t stream) {
  using C = ACfg<D>;
  auto* kern = attn_fwd_kernel<D>;
  DETB200_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::kSmemBytes));
  dim3 grid((a.S_q + kBM - 1) / kBM, a.B * a.H);
  kern<<<grid, kAThreads, C::kSmemBytes, stream>>>(tq, tk, tv, a);
  DETB200_LAUNCH_CHECK();
  return 0;
}

}  // namespace

// q/k/v/o: bf16, logical shape (B, S, H, D) with element strides (sb, ss, sh) and D contiguous.
// lse (nullable): fp32 [B, H, S_q], natural-log sum-exp of the scaled scores (needed by the backward).
DETB200_API int detb200_attn_fwd(const void* q, const void* k, const void* v, void* o, float* lse, int B, int H, int H_kv, int S_q,
                                 int S_kv, int D, const int64_t* q_strides, const int64_t* k_strides, const int64_t* v_strides,
                                 const int64_t* o_strides, float scale, int causal, cudaStream_t stream) {
  if (D != 64 && D != 128) {
    detb200_set_error("attn_fwd: head dim %d unsupported (64 or 128)", D);
    return DETB200_ERR_UNSUPPORTED;
  }
  if (H_kv <= 0 || H % H_kv != 0 || B <= 0 || S_q <= 0 || S_kv <= 0) {
    detb200_set_error("attn_fwd: bad shape B %d H %d H_kv %d S_q %d S_kv %d", B, H, H_kv, S_q, S_kv);
    return DETB200_ERR_BAD_ARG;
  }
  if ((reinterpret_cast<uintptr_t>(o) & 15u) || (o_strides[0] & 7) || (o_strides[1] & 7) || (o_strides[2] & 7)) {
    detb200_set_error("attn_fwd: O must be 16-byte aligned with strides that are multiples of 8 elements");
    return DETB200_ERR_BAD_ARG;
  }
  CUtensorMap tq, tk, tv;
  auto mk = [&](CUtensorMap* m, const void* base, int heads, int S, const int64_t* st, uint32_t rows) {
    const uint64_t dims[4] = {(uint64_t)D, (uint64_t)heads, (uint64_t)S, (uint64_t)B};
    const uint64_t strides[3] = {(uint64_t)st[2] * 2, (uint64_t)st[1] * 2, (uint64_t)st[0] * 2};
    const uint32_t box[4] = {64, 1, rows, 1};
    return detb200_make_tmap_4d_bf16(m, base, dims, strides, box);
  };
  int rc;
  if ((rc = mk(&tq, q, H, S_q, q_strides, kBM))) return rc;
  if ((rc = mk(&tk, k, H_kv, S_kv, k_strides, kBN))) return rc;
  if ((rc = mk(&tv, v, H_kv, S_kv, v_strides, kBN))) return rc;
  AttnArgs a;
  a.O = reinterpret_cast<__nv_bfloat16*>(o); a.lse = lse; a.B = B; a.H = H; a.H_kv = H_kv; a.S_q = S_q; a.S_kv = S_kv;
  a.o_sb = o_strides[0]; a.o_ss = o_strides[1]; a.o_sh = o_strides[2];
  a.scale_log2 = scale * 1.4426950408889634f; a.causal = causal;
  return D == 128 ? launch_fwd<128>(tq, tk, tv, a, stream) : launch_fwd<64>(tq, tk, tv, a, stream);
}
