// Flash attention forward on tcgen05: S = Q K^T and O += P V both run on the 5th-gen tensor cores,
// S/P/O live in tensor memory, K/V blocks are streamed by TMA, the online softmax runs in registers.
//
//   warp 0      TMA producer   Q tile once, then (K_j, V_j) pairs into a 2-deep smem ring
//   warp 1      MMA issuer     QK(j+1) is issued as soon as the softmax has *read* S(j), so it overlaps the
//                              exponentials; PV(j) follows when P(j) has been written to TMEM
//   warps 2..5  softmax        one thread per query row: tcgen05.ld S -> running max / sum (lazy rescale of O
//                              only when the max moved by > 2^8) -> P as packed bf16 -> tcgen05.st into one of two P
//                              buffers, so P(j) is produced while PV(j-1) is still running
//
// Two CTAs are resident per SM (96 KiB smem, 256 TMEM columns each) so one CTA's softmax overlaps the other's
// MMAs.  Layout-agnostic: Q/K/V/O are addressed as (batch, seq, head, dim) with arbitrary strides (dim contiguous),
// which covers [B,H,S,D], [B,S,H,D] and packed QKV.  GQA via h_kv = h / (H / H_kv).
// Replaces: the attention of the GPT-2 / GPT-NeoX examples (K13; Megatron softmax+bmm or flash-attn there).
#include "sm100.cuh"
#include <cstdlib>

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
  long long* prof;                 // debug (DETB200_ATTN_PROF=1): cycle stamps of CTA (0,0)'s first softmax warp
  int delay_v_ns;                  // test hook (detb200_attn_debug_delay_v): hold every V tile back, as a congested HBM/NVLink would
};

int g_delay_v_ns = 0;

// 2^x on the FMA / integer pipes (no MUFU): round-to-nearest split x = n + f with the 1.5*2^23 trick, cubic for 2^f on
// [-0.5, 0.5] (relative error 6e-4, below bf16 resolution), n added straight into the exponent field.
__device__ __forceinline__ float exp2_fma(float x) {
  x = fmaxf(x, -125.f);
  const float t = x + 12582912.f;
  const float f = x - (t - 12582912.f);
  float p = fmaf(f, 0.0555041087f, 0.2402265070f);
  p = fmaf(p, f, 0.6931471806f);
  p = fmaf(p, f, 1.0f);
  return __int_as_float(__float_as_int(p) + (__float_as_int(t) << 23));
}

template <int D> struct ACfg {
  static constexpr int kQBytes = kBM * D * 2;
  static constexpr int kKVBytes = kBN * D * 2;           // one of K or V
  static constexpr int kStageBytes = 2 * kKVBytes;
  static constexpr int kSmemBytes = kQBytes + kStagesKV * kStageBytes + 1024 + 256;
  static constexpr int kColS = 0, kColP = kBN, kColO = 128;
  static constexpr int kTmemCols = 256;
};

// kPolyOf4: of every 4 exponentials, this many take the FMA path (MUFU.EX2 is 16/clk/SM on B200)
template <int D, int kPolyOf4>
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
  // K and V have their own full/empty barriers: a K slot is free as soon as its QK MMA retired (one block earlier than
  // the V slot), so K(j+2) is in flight while block j is still in the softmax
  auto k_full = [&](int s) { return bar_base + 8u * (1 + s); };
  auto k_empty = [&](int s) { return bar_base + 8u * (1 + kStagesKV + s); };
  auto v_full = [&](int s) { return bar_base + 8u * (1 + 2 * kStagesKV + s); };
  auto v_empty = [&](int s) { return bar_base + 8u * (1 + 3 * kStagesKV + s); };
  const uint32_t s_full = bar_base + 8u * (1 + 4 * kStagesKV);
  // p_full and o_done are barrier PAIRS indexed by j & 1.  With one p_full barrier the softmax can complete P(j) and P(j+1)
  // (S(j+1) is issued before the MMA thread waits for P(j)) while the MMA thread still waits for V(j) to land: two phases
  // ahead of a parity wait is indistinguishable from none, and the kernel hangs -- seen under NVLink/HBM contention only.
  const uint32_t s_empty = s_full + 8u, p_full = s_full + 16u, o_done = s_full + 32u;
  const uint32_t tmem_slot = s_full + 48u;
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
    for (int s = 0; s < kStagesKV; ++s) { mbar_init(k_full(s), 1); mbar_init(k_empty(s), 1); mbar_init(v_full(s), 1); mbar_init(v_empty(s), 1); }
    mbar_init(s_full, 1);
    mbar_init(s_empty, 128);
    mbar_init(p_full, 128);
    mbar_init(p_full + 8u, 128);
    mbar_init(o_done, 1);
    mbar_init(o_done + 8u, 1);
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
      auto load_k = [&](int j) {
        const int st = j % kStagesKV;
        mbar_wait(k_empty(st), ((j / kStagesKV) & 1) ^ 1u);
        mbar_expect_tx(k_full(st), C::kKVBytes);
        const uint32_t sK = sKV + st * C::kStageBytes;
#pragma unroll
        for (int hh = 0; hh < D / 64; ++hh) tma_load_4d(sK + hh * (kBN * 128), &tmap_k, k_full(st), hh * 64, h_kv, j * kBN, b);
      };
      auto load_v = [&](int j) {
        const int st = j % kStagesKV;
        mbar_wait(v_empty(st), ((j / kStagesKV) & 1) ^ 1u);
        if (args.delay_v_ns > 0) {
          const uint64_t t0 = globaltimer_ns();
          while (globaltimer_ns() - t0 < (uint64_t)args.delay_v_ns) __nanosleep(200);
        }
        mbar_expect_tx(v_full(st), C::kKVBytes);
        const uint32_t sV = sKV + st * C::kStageBytes + C::kKVBytes;
#pragma unroll
        for (int hh = 0; hh < D / 64; ++hh) tma_load_4d(sV + hh * (kBN * 128), &tmap_v, v_full(st), hh * 64, h_kv, j * kBN, b);
      };
      // K runs one block ahead of V: K(j+1) only waits for QK(j-1), V(j) for PV(j-2); neither wait blocks the other
      load_k(0);
      for (int j = 0; j < n_kv; ++j) {
        if (j + 1 < n_kv) load_k(j + 1);
        load_v(j);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc_qk = idesc_bf16_f32(kBM, kBN, 0, 0);
      constexpr uint32_t idesc_pv = idesc_bf16_f32(kBM, D, 0, 1);
      const uint64_t dq_base = smem_desc_sw128(sQ, 16, 1024);
      const uint64_t dk_base0 = smem_desc_sw128(sKV, 16, 1024);
      const uint64_t dv_base0 = smem_desc_sw128(sKV + C::kKVBytes, kBN * 128, 1024);      // MN-major: LBO = one 64-wide d group
      auto issue_qk = [&](int j) {
        // descriptors differ from their base only by a compile-time constant in the (16-byte granular) address field: one
        // add per MMA keeps the single issuing thread far ahead of the tensor pipe
        const uint64_t db0 = dk_base0 + (uint64_t)((j % kStagesKV) * (C::kStageBytes >> 4));
#pragma unroll
        for (int k = 0; k < D / 16; ++k) {
          const uint64_t da = dq_base + (uint64_t)(((k / 4) * (kBM * 128) + (k % 4) * 32) >> 4);
          const uint64_t db = db0 + (uint64_t)(((k / 4) * (kBN * 128) + (k % 4) * 32) >> 4);
          mma_f16_ss(tS, da, db, idesc_qk, k != 0 ? 1u : 0u);
        }
        tc_commit(s_full);
      };
      mbar_wait(q_full, 0);
      mbar_wait(k_full(0), 0);
      tc_fence_after();
      issue_qk(0);
      tc_commit(k_empty(0));
      for (int j = 0; j < n_kv; ++j) {
        if (j + 1 < n_kv) {
          mbar_wait(k_full((j + 1) % kStagesKV), ((j + 1) / kStagesKV) & 1);
          mbar_wait(s_empty, j & 1);                  // softmax(j) holds S(j) in registers
          tc_fence_after();
          issue_qk(j + 1);
          tc_commit(k_empty((j + 1) % kStagesKV));
        }
        mbar_wait(v_full(j % kStagesKV), (j / kStagesKV) & 1);
        mbar_wait(p_full + 8u * (j & 1), (j >> 1) & 1);
        tc_fence_after();
        const uint64_t dv0 = dv_base0 + (uint64_t)((j % kStagesKV) * (C::kStageBytes >> 4));
#pragma unroll
        for (int k = 0; k < kBN / 16; ++k) {
          // V block is [keys, D] with D contiguous: MN-major B; 64-wide d groups are kBN*128 B apart
          const uint64_t dv = dv0 + (uint64_t)((k * 2048) >> 4);
          mma_f16_ts(tO, tP + (j & 1) * (kBN / 2) + k * 8, dv, idesc_pv, (j | k) != 0 ? 1u : 0u);
        }
        tc_commit(v_empty(j % kStagesKV));
        tc_commit(o_done + 8u * (j & 1));
      }
    }
  } else {
    const int q = warp & 3;
    const int row = q * 32 + lane;
    const int qi = q0 + row;
    const uint32_t lane_off = (uint32_t)(q * 32) << 16;
    float m_used = -INFINITY, l = 0.f;
    // PV(i) commits to o_done[i & 1].  A tcgen05.commit covers every MMA issued before it, and QK(j) is issued after
    // PV(j-2): once S(j) has been observed, PV(j-2) has retired and the P buffer (j & 1) is free without any further wait.
    // Only a rescale of O needs PV(j-1); its barrier cannot be more than one phase ahead (PV(j+1) needs this thread's
    // P(j+1)), so a direct parity wait is unambiguous.
    auto wait_pv_done = [&](int i) {
      mbar_wait(o_done + 8u * (i & 1), (i >> 1) & 1);
      tc_fence_after();
    };
    const bool prof = args.prof != nullptr && blockIdx.x == 0 && blockIdx.y == 0 && warp == 2 && lane == 0;
    for (int j = 0; j < n_kv; ++j) {
      if (prof && j < 8) args.prof[j * 8 + 0] = clock64();
      mbar_wait(s_full, j & 1);
      tc_fence_after();
      if (prof && j < 8) args.prof[j * 8 + 1] = clock64();
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
      // P is double-buffered: P(j) may be written while PV(j-1) still runs; only PV(j-2), the previous user of this
      // buffer, must have retired.  Rescaling O additionally needs PV(j-1) to be done with it.
      if (prof && j < 8) args.prof[j * 8 + 2] = clock64();
      if (any && j > 0) wait_pv_done(j - 1);
      if (prof && j < 8) args.prof[j * 8 + 4] = clock64();
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
        const float x0 = s[c] * args.scale_log2 - mm, x1 = s[c + 1] * args.scale_log2 - mm;
        const float p0 = ((c & 3) < kPolyOf4) ? exp2_fma(x0) : exp2f(x0);
        const float p1 = (((c + 1) & 3) < kPolyOf4) ? exp2_fma(x1) : exp2f(x1);
        sum += p0 + p1;
        pk[c >> 1] = pack_bf16(p0, p1);
      }
      l += sum;
      if (prof && j < 8) args.prof[j * 8 + 5] = clock64() + (long long)(sum == 12345.f);
      tmem_st_32x32(tP + (j & 1) * (kBN / 2) + lane_off, pk);
      tmem_st_wait();
      tc_fence_before();
      mbar_arrive(p_full + 8u * (j & 1));
      if (prof && j < 8) args.prof[j * 8 + 3] = clock64();
    }
    wait_pv_done(n_kv - 1);
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

template <int D, int kPoly>
int launch_fwd_p(const CUtensorMap& tq, const CUtensorMap& tk, const CUtensorMap& tv, const AttnArgs& a, cudaStream_t stream) {
  using C = ACfg<D>;
  auto* kern = attn_fwd_kernel<D, kPoly>;
  DETB200_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::kSmemBytes));
  dim3 grid((a.S_q + kBM - 1) / kBM, a.B * a.H);
  kern<<<grid, kAThreads, C::kSmemBytes, stream>>>(tq, tk, tv, a);
  DETB200_LAUNCH_CHECK();
  return 0;
}

template <int D>
int launch_fwd(const CUtensorMap& tq, const CUtensorMap& tk, const CUtensorMap& tv, const AttnArgs& a, cudaStream_t stream) {
  static const int poly = [] { const char* e = getenv("DETB200_ATTN_POLY"); return e ? atoi(e) : 0; }();
  if (poly == 1) return launch_fwd_p<D, 1>(tq, tk, tv, a, stream);
  if (poly == 2) return launch_fwd_p<D, 2>(tq, tk, tv, a, stream);
  return launch_fwd_p<D, 0>(tq, tk, tv, a, stream);
}

}  // namespace

// Test hook: delay every V tile of the forward kernel by `ns` nanoseconds (0 = off).  Reorders the pipeline the way a
// congested memory system does; tests/test_attention_gpu.py uses it to pin the barrier protocol.
DETB200_API int detb200_attn_debug_delay_v(int ns) {
  g_delay_v_ns = ns < 0 ? 0 : ns;
  return 0;
}

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
  a.prof = nullptr;
  a.delay_v_ns = g_delay_v_ns;
  static const bool want_prof = getenv("DETB200_ATTN_PROF") != nullptr;
  if (want_prof) {
    static long long* dbuf = nullptr;
    if (dbuf == nullptr) DETB200_CHECK(cudaMalloc(&dbuf, 64 * sizeof(long long)));
    DETB200_CHECK(cudaMemsetAsync(dbuf, 0, 64 * sizeof(long long), stream));
    a.prof = dbuf;
    rc = D == 128 ? launch_fwd<128>(tq, tk, tv, a, stream) : launch_fwd<64>(tq, tk, tv, a, stream);
    long long h[64];
    DETB200_CHECK(cudaMemcpyAsync(h, dbuf, sizeof(h), cudaMemcpyDeviceToHost, stream));
    DETB200_CHECK(cudaStreamSynchronize(stream));
    for (int j = 0; j < 8 && h[j * 8 + 3] != 0; ++j)
      fprintf(stderr, "attn_prof j=%d wait_s=%lld ld+max=%lld wait_pv=%lld rescale+exp=%lld st+arrive=%lld loop=%lld\n", j, h[j * 8 + 1] - h[j * 8],
              h[j * 8 + 2] - h[j * 8 + 1], h[j * 8 + 4] - h[j * 8 + 2], h[j * 8 + 5] - h[j * 8 + 4], h[j * 8 + 3] - h[j * 8 + 5],
              j ? h[j * 8] - h[(j - 1) * 8] : 0);
    return rc;
  }
  return D == 128 ? launch_fwd<128>(tq, tk, tv, a, stream) : launch_fwd<64>(tq, tk, tv, a, stream);
}

// =====================================================================================================
// Backward.  One CTA owns a block of 128 keys of one (batch, head) and walks the query blocks (64 rows)
// that attend to it.  All five GEMMs of a (key block, query block) pair run on tcgen05; the accumulators
// dK, dV stay in tensor memory for the whole walk:
//
//   S^T  = K Q_i^T          dP^T = V dO_i^T                      (A, B from smem, K-major)
//   P^T  = exp2(S^T*c - lse_i)      dS^T = P^T o (dP^T - delta_i) * scale       (8 compute warps, one key row per thread)
//   dV  += P^T dO_i         dK  += dS^T Q_i                      (A = bf16 P^T / dS^T in TMEM, B = the same smem tiles read MN-major)
//   dQ_i^T = K^T dS^T                                            (A = K read MN-major, B = dS^T staged in smem) -> fp32 atomics
//
// TMEM columns: S^T 64 | dP^T 64 | P^T 32 | dS^T 32 | dQ^T 64 | dV 128 | dK 128 = 512.
// S^T/dP^T of block i+1 are issued as soon as the compute warps hold block i in registers.
namespace {

constexpr int kBwdKeys = 128;
constexpr int kBwdQ = 64;
constexpr int kBwdThreads = 320;            // warp 0 TMA, warp 1 MMA, warps 2..9 compute
constexpr int kBwdCompute = 256;
constexpr int kBwdQStages = 3;              // (Q_i, dO_i) ring: a slot is refilled only after dQ(i) retired, so two slots expose the load latency

struct AttnBwdArgs {
  const float* lse;       // [B, H, S]
  const float* delta;     // [B, H, S]
  float* dq_acc;          // [B, S, H, D] fp32, zero-initialised
  __nv_bfloat16* dK;
  __nv_bfloat16* dV;
  int B, H, S;
  int64_t dk_sb, dk_ss, dk_sh, dv_sb, dv_ss, dv_sh;
  float scale, scale_log2;
  int causal;
  long long* prof;        // debug (DETB200_ATTN_PROF=1): cycle stamps of CTA (0,0)'s first compute warp
};

template <int kD> struct BCfg {
  static constexpr int D = kD;
  static constexpr int kKBytes = kBwdKeys * D * 2;          // 32 KiB (K or V)
  static constexpr int kQBytes = kBwdQ * D * 2;             // 16 KiB (Q_i or dO_i)
  static constexpr int kdSBytes = kBwdKeys * kBwdQ * 2;     // 16 KiB
  static constexpr int kdQBytes = kBwdQ * kD * 4;           // fp32 dQ tile staged for the TMA reduce-add
  static constexpr int kSmemBytes = 2 * kKBytes + 2 * kBwdQStages * kQBytes + kdSBytes + kdQBytes + 1024 /*align*/ + kBwdQStages * 2 * kBwdQ * 4 /*lse2, delta per ring slot*/ + 256;
  static constexpr int tS = 0, tdP = 64, tP = 128, tdS = 160, tdQ = 192, tdV = 256, tdK = 256 + kD;
};

template <int kD>
__global__ void __launch_bounds__(kBwdThreads, 1)
attn_bwd_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_k,
                const __grid_constant__ CUtensorMap tmap_v, const __grid_constant__ CUtensorMap tmap_do,
                const __grid_constant__ CUtensorMap tmap_dq, const __grid_constant__ CUtensorMap tmap_lse2,
                const __grid_constant__ CUtensorMap tmap_delta, const AttnBwdArgs args) {
  using C = BCfg<kD>;
  constexpr int D = C::D;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t sK = smem_base, sV = sK + C::kKBytes;
  const uint32_t sQ0 = sV + C::kKBytes;                       // Q stage s at sQ0 + s*2*kQBytes, dO right after it
  const uint32_t sdS = sQ0 + 2 * kBwdQStages * C::kQBytes;
  const uint32_t sdQ = sdS + C::kdSBytes;                     // [64 q][D] fp32, dense
  const uint32_t sStat = sdQ + C::kdQBytes;                   // per ring slot: float lse2[64], delta[64] (TMA-loaded with Q / dO)
  const uint32_t bar_base = sStat + kBwdQStages * 2 * kBwdQ * 4;
  const uint32_t kv_full = bar_base;
  auto qdo_full = [&](int s) { return bar_base + 8u * (1 + s); };
  auto qdo_empty = [&](int s) { return bar_base + 8u * (1 + kBwdQStages + s); };
  constexpr uint32_t kB0 = 1 + 2 * kBwdQStages;
  const uint32_t sp_full = bar_base + 8u * kB0, sp_empty = bar_base + 8u * (kB0 + 1), pds_full = bar_base + 8u * (kB0 + 2),
                 mma_done = bar_base + 8u * (kB0 + 3), dq_empty = bar_base + 8u * (kB0 + 4);
  const uint32_t tmem_slot = bar_base + 8u * (kB0 + 5);
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
  float* stat = reinterpret_cast<float*>(smem_gen + (sStat - smem_base));
  volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(smem_gen + (tmem_slot - smem_base));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int k_blk = blockIdx.x, k0 = k_blk * kBwdKeys;
  const int bh = blockIdx.y, b = bh / args.H, h = bh % args.H;
  const int nq_total = (args.S + kBwdQ - 1) / kBwdQ;
  const int i_min = args.causal ? (k0 / kBwdQ) : 0;            // first query block that can see these keys
  const int n_it = nq_total - i_min;

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmap_q); prefetch_tmap(&tmap_k); prefetch_tmap(&tmap_v); prefetch_tmap(&tmap_do); prefetch_tmap(&tmap_dq); prefetch_tmap(&tmap_lse2); prefetch_tmap(&tmap_delta);
  }
  if (warp == 1 && lane == 0) {
    mbar_init(kv_full, 1);
    for (int s = 0; s < kBwdQStages; ++s) { mbar_init(qdo_full(s), 1); mbar_init(qdo_empty(s), 1); }
    mbar_init(sp_full, 1);
    mbar_init(sp_empty, kBwdCompute);
    mbar_init(pds_full, kBwdCompute);
    mbar_init(mma_done, 1);
    mbar_init(dq_empty, kBwdCompute);
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;

  if (n_it > 0) {
    if (warp == 0) {
      if (lane == 0) {
        mbar_expect_tx(kv_full, 2 * C::kKBytes);
#pragma unroll
        for (int hh = 0; hh < D / 64; ++hh) {
          tma_load_4d(sK + hh * (kBwdKeys * 128), &tmap_k, kv_full, hh * 64, h, k0, b);
          tma_load_4d(sV + hh * (kBwdKeys * 128), &tmap_v, kv_full, hh * 64, h, k0, b);
        }
        for (int it = 0; it < n_it; ++it) {
          const int st = it % kBwdQStages, q0 = (i_min + it) * kBwdQ;
          mbar_wait(qdo_empty(st), ((it / kBwdQStages) & 1) ^ 1u);
          mbar_expect_tx(qdo_full(st), 2 * C::kQBytes + 2 * kBwdQ * 4);
          const uint32_t sQ = sQ0 + st * 2 * C::kQBytes, sdO = sQ + C::kQBytes;
          // the block's softmax statistics ride along (rows beyond S are zero-filled, which makes their dS vanish)
          tma_load_4d(sStat + st * (2 * kBwdQ * 4), &tmap_lse2, qdo_full(st), q0, bh, 0, 0);
          tma_load_4d(sStat + st * (2 * kBwdQ * 4) + kBwdQ * 4, &tmap_delta, qdo_full(st), q0, bh, 0, 0);
#pragma unroll
          for (int hh = 0; hh < D / 64; ++hh) {
            tma_load_4d(sQ + hh * (kBwdQ * 128), &tmap_q, qdo_full(st), hh * 64, h, q0, b);
            tma_load_4d(sdO + hh * (kBwdQ * 128), &tmap_do, qdo_full(st), hh * 64, h, q0, b);
          }
        }
      }
    } else if (warp == 1) {
      if (lane == 0) {
        constexpr uint32_t idesc_s = idesc_bf16_f32(128, kBwdQ, 0, 0);       // S^T, dP^T: [128 keys x 64 q]
        constexpr uint32_t idesc_dkv = idesc_bf16_f32(128, D, 0, 1);         // dV, dK: A in TMEM, B MN-major
        constexpr uint32_t idesc_dq = idesc_bf16_f32(128, kBwdQ, 1, 1);      // dQ^T: A = K MN-major, B = dS^T MN-major
        // base descriptors; every MMA adds a compile-time constant to the address field (see the forward kernel)
        const uint64_t dK_k = smem_desc_sw128(sK, 16, 1024), dV_k = smem_desc_sw128(sV, 16, 1024);                  // K-major A
        const uint64_t dK_mn = smem_desc_sw128(sK, D == 128 ? kBwdKeys * 128 : 0, 1024);                          // MN-major A (M = d)
        const uint64_t ddS_mn = smem_desc_sw128(sdS, kBwdKeys * 128, 1024);
        const uint64_t dQ_k0 = smem_desc_sw128(sQ0, 16, 1024), ddO_k0 = smem_desc_sw128(sQ0 + C::kQBytes, 16, 1024);
        const uint64_t dQ_mn0 = smem_desc_sw128(sQ0, kBwdQ * 128, 1024), ddO_mn0 = smem_desc_sw128(sQ0 + C::kQBytes, kBwdQ * 128, 1024);
        constexpr uint32_t kSlot16 = (2 * C::kQBytes) >> 4;                                   // ring slot pitch in descriptor units
        auto issue_s_dp = [&](int it) {
          const uint64_t slot = (uint64_t)((it % kBwdQStages) * kSlot16);
          const uint64_t q0d = dQ_k0 + slot, do0d = ddO_k0 + slot;
#pragma unroll
          for (int k = 0; k < D / 16; ++k) {
            const uint64_t da = dK_k + (uint64_t)(((k / 4) * (kBwdKeys * 128) + (k % 4) * 32) >> 4);
            const uint64_t db = q0d + (uint64_t)(((k / 4) * (kBwdQ * 128) + (k % 4) * 32) >> 4);
            mma_f16_ss(tmem_base + C::tS, da, db, idesc_s, k != 0 ? 1u : 0u);
          }
#pragma unroll
          for (int k = 0; k < D / 16; ++k) {
            const uint64_t da = dV_k + (uint64_t)(((k / 4) * (kBwdKeys * 128) + (k % 4) * 32) >> 4);
            const uint64_t db = do0d + (uint64_t)(((k / 4) * (kBwdQ * 128) + (k % 4) * 32) >> 4);
            mma_f16_ss(tmem_base + C::tdP, da, db, idesc_s, k != 0 ? 1u : 0u);
          }
          tc_commit(sp_full);
        };
        const bool mprof = args.prof != nullptr && blockIdx.x == 0 && blockIdx.y == 0;
        mbar_wait(kv_full, 0);
        mbar_wait(qdo_full(0), 0);
        tc_fence_after();
        issue_s_dp(0);
        for (int it = 0; it < n_it; ++it) {
          if (it + 1 < n_it) {
            mbar_wait(qdo_full((it + 1) % kBwdQStages), ((it + 1) / kBwdQStages) & 1);
            mbar_wait(sp_empty, it & 1);
            tc_fence_after();
            if (mprof && it < 4) args.prof[64 + it * 4 + 0] = clock64();
            issue_s_dp(it + 1);
            if (mprof && it < 4) args.prof[64 + it * 4 + 1] = clock64();
          }
          mbar_wait(pds_full, it & 1);
          tc_fence_after();
          if (mprof && it < 4) args.prof[64 + it * 4 + 2] = clock64();
          const uint64_t slot_mn = (uint64_t)((it % kBwdQStages) * kSlot16);
#pragma unroll
          for (int k = 0; k < kBwdQ / 16; ++k) {       // reduction over the 64 queries
            const uint64_t ddo = ddO_mn0 + slot_mn + (uint64_t)((k * 2048) >> 4);
            mma_f16_ts(tmem_base + C::tdV, tmem_base + C::tP + k * 8, ddo, idesc_dkv, (it | k) != 0 ? 1u : 0u);
          }
#pragma unroll
          for (int k = 0; k < kBwdQ / 16; ++k) {
            const uint64_t dq = dQ_mn0 + slot_mn + (uint64_t)((k * 2048) >> 4);
            mma_f16_ts(tmem_base + C::tdK, tmem_base + C::tdS + k * 8, dq, idesc_dkv, (it | k) != 0 ? 1u : 0u);
          }
          if (it > 0) {
            mbar_wait(dq_empty, (it - 1) & 1);
            tc_fence_after();
          }
#pragma unroll
          for (int k = 0; k < kBwdKeys / 16; ++k) {    // reduction over the 128 keys
            const uint64_t da = dK_mn + (uint64_t)((k * 2048) >> 4);
            const uint64_t db = ddS_mn + (uint64_t)((k * 2048) >> 4);
            mma_f16_ss(tmem_base + C::tdQ, da, db, idesc_dq, k != 0 ? 1u : 0u);
          }
          tc_commit(qdo_empty(it % kBwdQStages));
          tc_commit(mma_done);
          if (mprof && it < 4) args.prof[64 + it * 4 + 3] = clock64();
        }
      }
    } else {
      const int cw = warp - 2;                    // 0..7
      const int q = warp & 3;                     // TMEM lane quarter
      const int half = cw >> 2;                   // which 32 of the 64 query columns
      const int row = q * 32 + lane;              // key row inside the block / d index for dQ^T
      const int key = k0 + row;
      const int ct = threadIdx.x - 64;            // 0..255
      const uint32_t lane_off = (uint32_t)(q * 32) << 16;

      // dQ^T of block `it`: TMEM -> registers -> transposed into a dense [q][d] fp32 tile in smem -> one asynchronous TMA
      // reduce-add into the global fp32 accumulator (the L2 does the atomics; the compute threads do not wait for them)
      auto drain_dq = [&](int it) {
        uint32_t r[32];
        tmem_ld_32x32(tmem_base + C::tdQ + lane_off + half * 32, r);
        tmem_ld_wait();
        tc_fence_before();
        mbar_arrive(dq_empty);
        if (ct == 0) tma_store_wait_read<0>();          // the previous reduce has read the staging tile
        named_bar_sync(2, kBwdCompute);
        if (row < D) {                                  // D == 64: TMEM lanes 64..127 hold duplicates of the padded MMA rows
          const uint32_t dst = sdQ + (uint32_t)((half * 32) * D + row) * 4u;
#pragma unroll
          for (int c = 0; c < 32; ++c) asm volatile("st.shared.b32 [%0], %1;" ::"r"(dst + (uint32_t)(c * D * 4)), "r"(r[c]) : "memory");
        }
        fence_proxy_async_smem();
        named_bar_sync(2, kBwdCompute);
        if (ct == 0) {
          tma_reduce_add_4d(&tmap_dq, sdQ, 0, h, (i_min + it) * kBwdQ, b);
          tma_store_commit();
        }
      };

      const bool prof = args.prof != nullptr && blockIdx.x == 0 && blockIdx.y == 0 && warp == 2 && lane == 0;
      for (int it = 0; it < n_it; ++it) {
        if (prof && it < 8) args.prof[it * 8 + 0] = clock64();
        const int q0 = (i_min + it) * kBwdQ;
        const float* st_l = stat + (it % kBwdQStages) * 2 * kBwdQ;
        if (prof && it < 8) args.prof[it * 8 + 1] = clock64();
        mbar_wait(sp_full, it & 1);
        mbar_wait(qdo_full(it % kBwdQStages), (it / kBwdQStages) & 1);   // already complete (S(it) consumed this slot); orders the stats reads
        tc_fence_after();
        if (prof && it < 8) args.prof[it * 8 + 2] = clock64();
        uint32_t rs[32], rp[32];
        tmem_ld_32x32(tmem_base + C::tS + lane_off + half * 32, rs);
        tmem_ld_32x32(tmem_base + C::tdP + lane_off + half * 32, rp);
        tmem_ld_wait();
        tc_fence_before();
        mbar_arrive(sp_empty);
        uint32_t pP[16], pS[16];
        // Rows of keys beyond S need no mask: their K and V rows were zero-filled by TMA, so whatever P they get multiplies zeros
        // in dQ and lands in dK/dV rows that are never stored.  Only tiles crossing the causal diagonal mask, per element.
        const bool diag = args.causal && (q0 < k0 + kBwdKeys);
        const float4* l4 = reinterpret_cast<const float4*>(st_l + half * 32);
        const float4* d4 = reinterpret_cast<const float4*>(st_l + kBwdQ + half * 32);
#pragma unroll
        for (int c = 0; c < 32; c += 4) {
          const float4 L = l4[c >> 2], Dl = d4[c >> 2];
          const float lv[4] = {L.x, L.y, L.z, L.w}, dv[4] = {Dl.x, Dl.y, Dl.z, Dl.w};
          float pq[4], dq[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            float pv = exp2f(__uint_as_float(rs[c + e]) * args.scale_log2 - lv[e]);
            if (diag && key > q0 + half * 32 + c + e) pv = 0.f;
            pq[e] = pv;
            dq[e] = pv * (__uint_as_float(rp[c + e]) - dv[e]) * args.scale;
          }
          pP[c >> 1] = pack_bf16(pq[0], pq[1]);
          pP[(c >> 1) + 1] = pack_bf16(pq[2], pq[3]);
          pS[c >> 1] = pack_bf16(dq[0], dq[1]);
          pS[(c >> 1) + 1] = pack_bf16(dq[2], dq[3]);
        }
        if (prof && it < 8) args.prof[it * 8 + 3] = clock64() + (long long)(pP[0] == 0x12345678u && pS[3] == 77u);
        if (it > 0) {
          mbar_wait(mma_done, (it - 1) & 1);      // dV/dK/dQ of the previous block retired: P/dS buffers free, dQ^T ready
          tc_fence_after();
          if (prof && it < 8) args.prof[it * 8 + 4] = clock64();
          drain_dq(it - 1);
        }
        if (prof && it < 8) args.prof[it * 8 + 5] = clock64();
        tmem_st_32x16(tmem_base + C::tP + lane_off + half * 16, pP);
        tmem_st_32x16(tmem_base + C::tdS + lane_off + half * 16, pS);
        {
          const uint32_t dst = sdS + (uint32_t)row * 128u;
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const uint32_t a = dst + (uint32_t)(((half * 4 + j) ^ (row & 7)) << 4);
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(a), "r"(pS[4 * j]), "r"(pS[4 * j + 1]), "r"(pS[4 * j + 2]), "r"(pS[4 * j + 3]) : "memory");
          }
        }
        tmem_st_wait();
        fence_proxy_async_smem();
        tc_fence_before();
        mbar_arrive(pds_full);
        if (prof && it < 8) args.prof[it * 8 + 6] = clock64();
      }
      mbar_wait(mma_done, (n_it - 1) & 1);
      tc_fence_after();
      drain_dq(n_it - 1);
      if (ct == 0) tma_store_wait<0>();
      // dV, dK: this thread owns key row `row`, columns [D/2 * half, D/2 * half + D/2)
      const bool row_ok = key < args.S;
#pragma unroll
      for (int which = 0; which < 2; ++which) {
        __nv_bfloat16* base = which == 0 ? args.dV + (int64_t)b * args.dv_sb + (int64_t)key * args.dv_ss + (int64_t)h * args.dv_sh
                                         : args.dK + (int64_t)b * args.dk_sb + (int64_t)key * args.dk_ss + (int64_t)h * args.dk_sh;
        const uint32_t tcol = tmem_base + (which == 0 ? C::tdV : C::tdK) + lane_off + half * (D / 2);
#pragma unroll
        for (int c = 0; c < D / 2; c += 32) {
          uint32_t r[32];
          tmem_ld_32x32(tcol + c, r);
          tmem_ld_wait();
          if (row_ok) {
#pragma unroll
            for (int i = 0; i < 32; i += 8) {
              uint4 u;
              u.x = pack_bf16(__uint_as_float(r[i]), __uint_as_float(r[i + 1]));
              u.y = pack_bf16(__uint_as_float(r[i + 2]), __uint_as_float(r[i + 3]));
              u.z = pack_bf16(__uint_as_float(r[i + 4]), __uint_as_float(r[i + 5]));
              u.w = pack_bf16(__uint_as_float(r[i + 6]), __uint_as_float(r[i + 7]));
              *reinterpret_cast<uint4*>(base + half * (D / 2) + c + i) = u;
            }
          }
        }
      }
    }
  } else if (warp >= 2) {
    // keys nobody attends to (cannot happen with S_q == S_kv, kept for safety): zero gradients
    const int row = (warp & 3) * 32 + lane, half = (warp - 2) >> 2, key = k0 + row;
    if (key < args.S) {
      for (int c = 0; c < kD / 2; c += 8) {
        *reinterpret_cast<uint4*>(args.dV + (int64_t)b * args.dv_sb + (int64_t)key * args.dv_ss + (int64_t)h * args.dv_sh + half * (kD / 2) + c) = make_uint4(0, 0, 0, 0);
        *reinterpret_cast<uint4*>(args.dK + (int64_t)b * args.dk_sb + (int64_t)key * args.dk_ss + (int64_t)h * args.dk_sh + half * (kD / 2) + c) = make_uint4(0, 0, 0, 0);
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

// delta[b,h,s] = sum_d O * dO   (one warp per row; lanes cover D in 4-element pieces)
__global__ void __launch_bounds__(256)
attn_delta_kernel(const __nv_bfloat16* __restrict__ o, const __nv_bfloat16* __restrict__ d_o, const float* __restrict__ lse, float* __restrict__ delta,
                  float* __restrict__ lse2, int B, int H, int S, int S_pad, int D,
                  int64_t o_sb, int64_t o_ss, int64_t o_sh, int64_t do_sb, int64_t do_ss, int64_t do_sh) {
  const int lane = threadIdx.x & 31;
  const int64_t rows = (int64_t)B * H * S;
  for (int64_t r = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5); r < rows; r += (int64_t)gridDim.x * 8) {
    const int s = (int)(r % S);
    const int h = (int)((r / S) % H);
    const int64_t b = r / ((int64_t)S * H);
    float v = 0.f;
    if (lane * 4 < D) {
      const uint2 a = *reinterpret_cast<const uint2*>(o + b * o_sb + (int64_t)s * o_ss + (int64_t)h * o_sh + lane * 4);
      const uint2 g = *reinterpret_cast<const uint2*>(d_o + b * do_sb + (int64_t)s * do_ss + (int64_t)h * do_sh + lane * 4);
      v = bf16lo(a.x) * bf16lo(g.x) + bf16hi(a.x) * bf16hi(g.x) + bf16lo(a.y) * bf16lo(g.y) + bf16hi(a.y) * bf16hi(g.y);
    }
    v = warp_sum(v);
    if (lane == 0) {                      // r enumerates (b, h, s); the statistics rows are padded to S_pad
      const int64_t w = (r / S) * S_pad + s;
      delta[w] = v;
      lse2[w] = lse[r] * 1.4426950408889634f;
    }
  }
}

// dq (bf16, strided) = dq_acc (fp32 [B,S,H,D] contiguous)
__global__ void __launch_bounds__(256)
attn_dq_convert_kernel(const float* __restrict__ acc, __nv_bfloat16* __restrict__ dq, int64_t n8, int S, int H, int D, int64_t sb, int64_t ss, int64_t sh) {
  for (int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; v < n8; v += (int64_t)gridDim.x * blockDim.x) {
    const float4 a = ldg_stream(reinterpret_cast<const float4*>(acc) + 2 * v), c = ldg_stream(reinterpret_cast<const float4*>(acc) + 2 * v + 1);
    const int d8 = (int)(v % (D / 8));
    int64_t t = v / (D / 8);
    const int h = (int)(t % H); t /= H;
    const int s = (int)(t % S);
    const int64_t b = t / S;
    *reinterpret_cast<uint4*>(dq + b * sb + (int64_t)s * ss + (int64_t)h * sh + d8 * 8) =
        make_uint4(pack_bf16(a.x, a.y), pack_bf16(a.z, a.w), pack_bf16(c.x, c.y), pack_bf16(c.z, c.w));
  }
}

}  // namespace

// Backward of detb200_attn_fwd for H_kv == H, S_q == S_kv, D in {64, 128}.  All tensors bf16 (B, S, H, D)-addressed with element
// strides {batch, seq, head}.  `delta` (fp32 [2,B,H,ceil64(S)]: row-dot(O,dO) and lse*log2e) and `dq_acc` (fp32 [B,S,H,D]) are caller-provided scratch; dq_acc is
// zeroed here.  dq/dk/dv may be slices of one packed buffer.
DETB200_API int detb200_attn_bwd(const void* q, const void* k, const void* v, const void* o, const void* d_o, const float* lse, float* delta,
                                 float* dq_acc, void* dq, void* dk, void* dv, int B, int H, int S, int D, const int64_t* q_st,
                                 const int64_t* k_st, const int64_t* v_st, const int64_t* o_st, const int64_t* do_st, const int64_t* dq_st,
                                 const int64_t* dk_st, const int64_t* dv_st, float scale, int causal, cudaStream_t stream) {
  if (D != 128 && D != 64) {
    detb200_set_error("attn_bwd: head dim %d unsupported (64 or 128)", D);
    return DETB200_ERR_UNSUPPORTED;
  }
  for (const int64_t* st : {o_st, do_st, dq_st, dk_st, dv_st})
    if ((st[0] & 7) || (st[1] & 7) || (st[2] & 7)) {
      detb200_set_error("attn_bwd: strides must be multiples of 8 elements");
      return DETB200_ERR_BAD_ARG;
    }
  CUtensorMap tq, tk, tv, tdo;
  auto mk = [&](CUtensorMap* m, const void* base, const int64_t* st, uint32_t rows) {
    const uint64_t dims[4] = {(uint64_t)D, (uint64_t)H, (uint64_t)S, (uint64_t)B};
    const uint64_t strides[3] = {(uint64_t)st[2] * 2, (uint64_t)st[1] * 2, (uint64_t)st[0] * 2};
    const uint32_t box[4] = {64, 1, rows, 1};
    return detb200_make_tmap_4d_bf16(m, base, dims, strides, box);
  };
  int rc;
  if ((rc = mk(&tq, q, q_st, kBwdQ))) return rc;
  if ((rc = mk(&tdo, d_o, do_st, kBwdQ))) return rc;
  if ((rc = mk(&tk, k, k_st, kBwdKeys))) return rc;
  if ((rc = mk(&tv, v, v_st, kBwdKeys))) return rc;
  CUtensorMap tdq;
  {
    const uint64_t dims[4] = {(uint64_t)D, (uint64_t)H, (uint64_t)S, (uint64_t)B};
    const uint64_t strides[3] = {(uint64_t)D * 4, (uint64_t)H * D * 4, (uint64_t)S * H * D * 4};
    const uint32_t box[4] = {(uint32_t)D, 1, kBwdQ, 1};
    if ((rc = detb200_make_tmap_4d_f32_plain(&tdq, dq_acc, dims, strides, box))) return rc;
  }
  const int64_t rows = (int64_t)B * H * S;
  const int S_pad = (S + 63) / 64 * 64;                          // row pitch of the statistics (TMA needs 16-byte multiples)
  const int64_t stat_elems = (int64_t)B * H * S_pad;
  CUtensorMap tl2, tdl;
  {
    const uint64_t dims[4] = {(uint64_t)S, (uint64_t)B * H, 1, 1};
    const uint64_t strides[3] = {(uint64_t)S_pad * 4, (uint64_t)stat_elems * 4, (uint64_t)stat_elems * 4};
    const uint32_t box[4] = {kBwdQ, 1, 1, 1};
    if ((rc = detb200_make_tmap_4d_f32_plain(&tdl, delta, dims, strides, box))) return rc;
    if ((rc = detb200_make_tmap_4d_f32_plain(&tl2, delta + stat_elems, dims, strides, box))) return rc;
  }
  attn_delta_kernel<<<grid_for(rows, 8, 1, kNumSMs * 16), 256, 0, stream>>>(reinterpret_cast<const __nv_bfloat16*>(o), reinterpret_cast<const __nv_bfloat16*>(d_o),
                                                                         lse, delta, delta + stat_elems, B, H, S, S_pad, D, o_st[0], o_st[1], o_st[2], do_st[0], do_st[1], do_st[2]);
  DETB200_LAUNCH_CHECK();
  DETB200_CHECK(cudaMemsetAsync(dq_acc, 0, sizeof(float) * (size_t)rows * D, stream));
  AttnBwdArgs a;
  a.lse = lse; a.delta = delta; a.dq_acc = dq_acc; a.dK = reinterpret_cast<__nv_bfloat16*>(dk); a.dV = reinterpret_cast<__nv_bfloat16*>(dv);
  a.B = B; a.H = H; a.S = S;
  a.dk_sb = dk_st[0]; a.dk_ss = dk_st[1]; a.dk_sh = dk_st[2]; a.dv_sb = dv_st[0]; a.dv_ss = dv_st[1]; a.dv_sh = dv_st[2];
  a.scale = scale; a.scale_log2 = scale * 1.4426950408889634f; a.causal = causal;
  a.prof = nullptr;
  static const bool want_prof = getenv("DETB200_ATTN_PROF") != nullptr;
  static long long* dbuf = nullptr;
  if (want_prof) {
    if (dbuf == nullptr) DETB200_CHECK(cudaMalloc(&dbuf, 96 * sizeof(long long)));
    DETB200_CHECK(cudaMemsetAsync(dbuf, 0, 96 * sizeof(long long), stream));
    a.prof = dbuf;
  }
  dim3 grid((S + kBwdKeys - 1) / kBwdKeys, B * H);
  if (D == 128) {
    DETB200_CHECK(cudaFuncSetAttribute(attn_bwd_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, BCfg<128>::kSmemBytes));
    attn_bwd_kernel<128><<<grid, kBwdThreads, BCfg<128>::kSmemBytes, stream>>>(tq, tk, tv, tdo, tdq, tl2, tdl, a);
  } else {
    DETB200_CHECK(cudaFuncSetAttribute(attn_bwd_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, BCfg<64>::kSmemBytes));
    attn_bwd_kernel<64><<<grid, kBwdThreads, BCfg<64>::kSmemBytes, stream>>>(tq, tk, tv, tdo, tdq, tl2, tdl, a);
  }
  DETB200_LAUNCH_CHECK();
  if (want_prof) {
    long long hbuf[96];
    DETB200_CHECK(cudaMemcpyAsync(hbuf, dbuf, sizeof(hbuf), cudaMemcpyDeviceToHost, stream));
    DETB200_CHECK(cudaStreamSynchronize(stream));
    const long long t0 = hbuf[0];
    for (int j = 1; j < 4; ++j)
      fprintf(stderr, "attn_bwd_abs it=%d compute: top=%lld sp_seen=%lld math_done=%lld mma_seen=%lld drained=%lld pds_arrive=%lld | mma: issue_sdp(it+1)=%lld..%lld pds_seen=%lld issued_all=%lld\n",
              j, hbuf[j * 8] - t0, hbuf[j * 8 + 2] - t0, hbuf[j * 8 + 3] - t0, hbuf[j * 8 + 4] - t0, hbuf[j * 8 + 5] - t0, hbuf[j * 8 + 6] - t0,
              hbuf[64 + j * 4] - t0, hbuf[64 + j * 4 + 1] - t0, hbuf[64 + j * 4 + 2] - t0, hbuf[64 + j * 4 + 3] - t0);
    for (int j = 0; j < 8 && hbuf[j * 8 + 6] != 0; ++j)
      fprintf(stderr, "attn_bwd_prof it=%d stage+bar=%lld wait_sp=%lld ld+math=%lld wait_mma=%lld drain_dq=%lld st+arrive=%lld loop=%lld\n", j,
              hbuf[j * 8 + 1] - hbuf[j * 8], hbuf[j * 8 + 2] - hbuf[j * 8 + 1], hbuf[j * 8 + 3] - hbuf[j * 8 + 2], j ? hbuf[j * 8 + 4] - hbuf[j * 8 + 3] : 0,
              j ? hbuf[j * 8 + 5] - hbuf[j * 8 + 4] : 0, hbuf[j * 8 + 6] - hbuf[j * 8 + 5], j ? hbuf[j * 8] - hbuf[(j - 1) * 8] : 0);
  }
  const int64_t n8 = rows * D / 8;
  attn_dq_convert_kernel<<<grid_for(n8, 256, 2, kNumSMs * 8), 256, 0, stream>>>(dq_acc, reinterpret_cast<__nv_bfloat16*>(dq), n8, S, H, D, dq_st[0], dq_st[1], dq_st[2]);
  DETB200_LAUNCH_CHECK();
  return 0;
}
