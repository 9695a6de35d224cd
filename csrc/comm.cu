// In-kernel collectives over NVLink 5 / NVSwitch peer memory, fused with the adjacent math
// (work-list rows R1, R3, R9).
//
// Replaces: the NCCL all-reduce that torch DDP / Horovod issue for the reference harness
// (harness/determined/pytorch/_pytorch_context.py:296-299, :357-368), the per-parameter
// grad.div_ (:805-812), GradScaler.unscale_ + inf check (:886-887), the norm pass of
// clip_grad_norm_ (:885-888), and DeepSpeed's ZeRO reduce-scatter + FusedAdam + all-gather
// (harness/determined/pytorch/deepspeed/_deepspeed_trial.py:364-499 user engine.step()).
//
// Memory model.  Every rank owns a "symmetric" allocation (plain cudaMalloc, exported with
// cudaIpcGetMemHandle and opened by the peers) holding its flat gradient buffer, and a small
// control block:
//     uint32 start[MAXR]   written by peer r at the beginning of a collective  (epoch value)
//     uint32 done [MAXR]   written by peer r when it finished touching OUR memory
//     float  stats[2][MAXR] per-peer running sum of squares [0][r] and non-finite count [1][r] for this step
//     uint32 epoch, blocks_done; float acc[2]   (local only)
// Flags carry a monotonically increasing epoch, so nothing is ever reset and back-to-back
// collectives (one per bucket) cannot confuse each other.
//
// Two-shot all-reduce: rank r reduces slice r of the bucket by reading all peers' buffers
// (16-byte loads over NVLink), applies scale = 1/(world*agg) [* 1/loss_scale], accumulates
// sum-of-squares and a non-finite flag, and stores the result back into *every* peer's buffer
// (16-byte remote stores).  Per GPU that is (W-1)/W*N bytes in and out: the same wire volume as a
// ring, but two NVSwitch hops of latency instead of 2(W-1).
#include "common.cuh"
#include <string.h>

#define DETB200_MAX_RANKS 16

namespace {

constexpr int kCommThreads = 512;

struct Ctrl {
  uint32_t start[DETB200_MAX_RANKS];
  uint32_t done[DETB200_MAX_RANKS];
  float stats[DETB200_MAX_RANKS * 2];
  uint32_t epoch;
  uint32_t blocks_done;
  float acc[2];
  uint32_t timed_out;
  uint32_t pad[3];
};

struct Peers {
  float* buf[DETB200_MAX_RANKS];
  Ctrl* ctrl[DETB200_MAX_RANKS];
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float4 ld_peer(const float4* p) {
  float4 r;
  asm volatile("ld.volatile.global.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p) : "memory");
  return r;
}
__device__ __forceinline__ void st_peer(float4* p, const float4& v) {
  asm volatile("st.volatile.global.v4.f32 [%0], {%1,%2,%3,%4};"
               :: "l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

// Spin with a deadline: a peer that never arrives (crashed rank, diverged control flow) must not
// hang the GPU forever.  On timeout the kernel records it and traps; the host sees a launch
// failure and the trial restarts from its last checkpoint (SURVEY 5.3 "new build").
__device__ __forceinline__ void wait_flag(const uint32_t* flag, uint32_t epoch, Ctrl* me, long long timeout_cycles) {
  long long t0 = clock64();
  while ((int32_t)(ld_acquire_sys(flag) - epoch) < 0) {
    if (timeout_cycles > 0 && clock64() - t0 > timeout_cycles) {
      me->timed_out = epoch;
      __threadfence_system();
      __trap();
    }
    __nanosleep(20);
  }
}

// phase bits: 1 = start barrier, 2 = reduce+scatter my slice, 4 = end barrier.  All bits set is the
// normal multi-GPU launch.  Single-GPU tests emulate W ranks by launching phase 2 once per
// emulated rank (kernels that spin on each other must not share a GPU).
__global__ void __launch_bounds__(kCommThreads)
allreduce_fused_kernel(Peers peers, int rank, int world, int64_t off, int64_t n, float scale,
                       const float* __restrict__ inv_loss_scale, int* __restrict__ found_inf,
                       int first_of_step, int phases, long long timeout_cycles) {
  __shared__ float red[32];
  __shared__ uint32_t s_epoch;
  __shared__ bool s_last;
  Ctrl* me = peers.ctrl[rank];
  if (threadIdx.x == 0) s_epoch = me->epoch + 1;
  __syncthreads();
  const uint32_t epoch = s_epoch;

  if (phases & 1) {
    if (threadIdx.x < world && threadIdx.x != rank) {
      if (blockIdx.x == 0) st_release_sys(&peers.ctrl[threadIdx.x]->start[rank], epoch);
      wait_flag(&me->start[threadIdx.x], epoch, me, timeout_cycles);
    }
    __syncthreads();
  }

  float sq = 0.f;
  int bad = 0;
  if (phases & 2) {
    const float s = scale * (inv_loss_scale ? *inv_loss_scale : 1.f);
    // slice boundaries in float4 units; the bucket is padded to a multiple of 4*world elements
    const int64_t nvec = n >> 2;
    const int64_t per = (nvec + world - 1) / world;
    const int64_t v0 = per * rank;
    const int64_t v1 = (v0 + per < nvec) ? v0 + per : nvec;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = v0 + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < v1; i += stride) {
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
      // issue all peer loads before consuming them: W independent NVLink requests in flight
      float4 v[DETB200_MAX_RANKS];
#pragma unroll
      for (int r = 0; r < DETB200_MAX_RANKS; ++r)
        if (r < world) v[r] = ld_peer(reinterpret_cast<const float4*>(peers.buf[(rank + r) % world] + off) + i);
#pragma unroll
      for (int r = 0; r < DETB200_MAX_RANKS; ++r)
        if (r < world) { acc.x += v[r].x; acc.y += v[r].y; acc.z += v[r].z; acc.w += v[r].w; }
      acc.x *= s; acc.y *= s; acc.z *= s; acc.w *= s;
      float q = acc.x * acc.x + acc.y * acc.y + acc.z * acc.z + acc.w * acc.w;
      bad |= !isfinite(q);
      sq += q;
#pragma unroll
      for (int r = 0; r < DETB200_MAX_RANKS; ++r)
        if (r < world) st_peer(reinterpret_cast<float4*>(peers.buf[(rank + r) % world] + off) + i, acc);
    }
    float tot = block_sum(sq, red);
    int anybad = __syncthreads_or(bad);
    if (threadIdx.x == 0) {
      atomicAdd(&me->acc[0], tot);
      if (anybad) atomicAdd(&me->acc[1], 1.f);
    }
  }

  // ---- completion: last block publishes stats + "done", then waits for every peer's "done"
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence_system();
    uint32_t t = atomicAdd(&me->blocks_done, 1u);
    s_last = (t == gridDim.x - 1);
  }
  __syncthreads();
  if (!s_last) return;
  __threadfence_system();
  if (phases & 4) {
    const float a0 = __ldcg(&me->acc[0]);
    const float a1 = __ldcg(&me->acc[1]);
    // read my running totals BEFORE any thread overwrites them (thread `rank` writes to `me`)
    const float base0 = first_of_step ? 0.f : __ldcg(&me->stats[rank]);
    const float base1 = first_of_step ? 0.f : __ldcg(&me->stats[DETB200_MAX_RANKS + rank]);
    __syncthreads();
    if (threadIdx.x < world) {
      Ctrl* pc = peers.ctrl[threadIdx.x];
      // running totals for this step: overwrite my slot on every peer
      float* slot = &pc->stats[rank];
      asm volatile("st.volatile.global.f32 [%0], %1;" :: "l"(slot), "f"(base0 + a0) : "memory");
      asm volatile("st.volatile.global.f32 [%0], %1;" :: "l"(slot + DETB200_MAX_RANKS), "f"(base1 + a1) : "memory");
      __threadfence_system();
      if (threadIdx.x != rank) {
        st_release_sys(&pc->done[rank], epoch);
        wait_flag(&me->done[threadIdx.x], epoch, me, timeout_cycles);
      }
    }
    __syncthreads();
    if (found_inf != nullptr && threadIdx.x == 0) {
      float nbad = 0.f;
      for (int r = 0; r < world; ++r) nbad += me->stats[DETB200_MAX_RANKS + r];
      if (nbad > 0.f) *found_inf = 1;
    }
    if (threadIdx.x == 0) {
      me->acc[0] = 0.f; me->acc[1] = 0.f;
      me->blocks_done = 0u;
      __threadfence();
      me->epoch = epoch;
    }
  } else if (threadIdx.x == 0) {
    // emulation mode: keep acc for inspection by the host test, re-arm the ticket only
    me->blocks_done = 0u;
  }
}

// ------------------------------------------------------------------------------------------
// ZeRO path: reduce-scatter fused with the partitioned Adam(W) update and the parameter all-gather.
// Rank r owns slice r of every bucket: fp32 master weights + moments exist only for that slice.
// One kernel: read slice r of all peers' gradient buffers -> sum*scale -> AdamW on the owned shard
// -> write the updated weights (bf16 and/or fp32) into EVERY peer's parameter buffer.  The reduced
// gradient shard never goes back to HBM.  With clipping enabled the caller runs it as two phases
// (mode 1: reduce + sum-of-squares into `gshard`; mode 2: Adam from `gshard` + push weights).
// ------------------------------------------------------------------------------------------
struct ParamPeers {
  void* p_full[DETB200_MAX_RANKS];  // each peer's full (gathered) parameter buffer: bf16 or fp32
};

template <bool kLowp>
__global__ void __launch_bounds__(kCommThreads)
rs_adam_ag_kernel(Peers peers, ParamPeers ppeers, int rank, int world, int64_t off, int64_t n,
                  float* __restrict__ master, float* __restrict__ m, float* __restrict__ v,
                  float* __restrict__ gshard, float scale, const float* __restrict__ lr_p,
                  const int* __restrict__ step_p, const float* __restrict__ grad_scale_p,
                  const int* __restrict__ skip_p, float beta1, float beta2, float eps, float wd, int mode,
                  int first_of_step, int phases, long long timeout_cycles) {
  __shared__ float red[32];
  __shared__ uint32_t s_epoch;
  __shared__ bool s_last;
  Ctrl* me = peers.ctrl[rank];
  if (threadIdx.x == 0) s_epoch = me->epoch + 1;
  __syncthreads();
  const uint32_t epoch = s_epoch;
  if (phases & 1) {
    if (threadIdx.x < world && threadIdx.x != rank) {
      if (blockIdx.x == 0) st_release_sys(&peers.ctrl[threadIdx.x]->start[rank], epoch);
      wait_flag(&me->start[threadIdx.x], epoch, me, timeout_cycles);
    }
    __syncthreads();
  }
  float sq = 0.f;
  int bad = 0;
  if (phases & 2) {
    const bool do_reduce = (mode == 0 || mode == 1);
    const bool do_update = (mode == 0 || mode == 2);
    const bool skip = skip_p != nullptr && *skip_p != 0;
    const float lr = lr_p ? *lr_p : 0.f;
    const int step = step_p ? *step_p : 1;
    const float gs = grad_scale_p ? *grad_scale_p : 1.f;
    const float bc1 = 1.f - __powf(beta1, (float)step);
    const float bc2 = 1.f - __powf(beta2, (float)step);
    const float step_size = lr / bc1;
    const float inv_sqrt_bc2 = rsqrtf(bc2);
    const int64_t nvec = n >> 2;
    const int64_t per = (nvec + world - 1) / world;   // shard length in float4
    const int64_t v0 = per * rank;
    const int64_t v1 = (v0 + per < nvec) ? v0 + per : nvec;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = v0 + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < v1; i += stride) {
      const int64_t li = i - v0;  // index into the owned shard
      float4 g;
      if (do_reduce) {
        float4 vv[DETB200_MAX_RANKS];
#pragma unroll
        for (int r = 0; r < DETB200_MAX_RANKS; ++r)
          if (r < world) vv[r] = ld_peer(reinterpret_cast<const float4*>(peers.buf[(rank + r) % world] + off) + i);
        g = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int r = 0; r < DETB200_MAX_RANKS; ++r)
          if (r < world) { g.x += vv[r].x; g.y += vv[r].y; g.z += vv[r].z; g.w += vv[r].w; }
        g.x *= scale; g.y *= scale; g.z *= scale; g.w *= scale;
        float q = g.x * g.x + g.y * g.y + g.z * g.z + g.w * g.w;
        bad |= !isfinite(q);
        sq += q;
        if (mode == 1) reinterpret_cast<float4*>(gshard)[li] = g;
      } else {
        g = reinterpret_cast<const float4*>(gshard)[li];
      }
      if (do_update && !skip) {
        float4 pv = reinterpret_cast<float4*>(master)[li];
        float4 mv = reinterpret_cast<float4*>(m)[li];
        float4 v2 = reinterpret_cast<float4*>(v)[li];
        float* pp = &pv.x; float* gg = &g.x; float* mm = &mv.x; float* vq = &v2.x;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          float gk = gg[k] * gs;
          float pk = pp[k] * (1.f - lr * wd);
          float mk = fmaf(beta1, mm[k], (1.f - beta1) * gk);
          float vk = fmaf(beta2, vq[k], (1.f - beta2) * gk * gk);
          pk -= step_size * __fdividef(mk, sqrtf(vk) * inv_sqrt_bc2 + eps);
          pp[k] = pk; mm[k] = mk; vq[k] = vk;
        }
        reinterpret_cast<float4*>(master)[li] = pv;
        reinterpret_cast<float4*>(m)[li] = mv;
        reinterpret_cast<float4*>(v)[li] = v2;
        // all-gather fused into the epilogue: push the fresh weights to every peer
        if (kLowp) {
          uint2 o = make_uint2(pack_bf16(pv.x, pv.y), pack_bf16(pv.z, pv.w));
#pragma unroll
          for (int r = 0; r < DETB200_MAX_RANKS; ++r)
            if (r < world) {
              uint2* dst = reinterpret_cast<uint2*>((__nv_bfloat16*)ppeers.p_full[(rank + r) % world] + off) + i;
              asm volatile("st.volatile.global.v2.u32 [%0], {%1,%2};" :: "l"(dst), "r"(o.x), "r"(o.y) : "memory");
            }
        } else {
#pragma unroll
          for (int r = 0; r < DETB200_MAX_RANKS; ++r)
            if (r < world) st_peer(reinterpret_cast<float4*>((float*)ppeers.p_full[(rank + r) % world] + off) + i, pv);
        }
      }
    }
    float tot = block_sum(sq, red);
    int anybad = __syncthreads_or(bad);
    if (threadIdx.x == 0 && do_reduce) {
      atomicAdd(&me->acc[0], tot);
      if (anybad) atomicAdd(&me->acc[1], 1.f);
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence_system();
    uint32_t t = atomicAdd(&me->blocks_done, 1u);
    s_last = (t == gridDim.x - 1);
  }
  __syncthreads();
  if (!s_last) return;
  __threadfence_system();
  if (phases & 4) {
    const float a0 = __ldcg(&me->acc[0]);
    const float a1 = __ldcg(&me->acc[1]);
    const float base0 = first_of_step ? 0.f : __ldcg(&me->stats[rank]);
    const float base1 = first_of_step ? 0.f : __ldcg(&me->stats[DETB200_MAX_RANKS + rank]);
    __syncthreads();
    if (threadIdx.x < world) {
      Ctrl* pc = peers.ctrl[threadIdx.x];
      if (mode != 2) {
        float* slot = &pc->stats[rank];
        asm volatile("st.volatile.global.f32 [%0], %1;" :: "l"(slot), "f"(base0 + a0) : "memory");
        asm volatile("st.volatile.global.f32 [%0], %1;" :: "l"(slot + DETB200_MAX_RANKS), "f"(base1 + a1) : "memory");
        __threadfence_system();
      }
      if (threadIdx.x != rank) {
        st_release_sys(&pc->done[rank], epoch);
        wait_flag(&me->done[threadIdx.x], epoch, me, timeout_cycles);
      }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      me->acc[0] = 0.f; me->acc[1] = 0.f;
      me->blocks_done = 0u;
      __threadfence();
      me->epoch = epoch;
    }
  } else if (threadIdx.x == 0) {
    me->blocks_done = 0u;
  }
}

// Broadcast rank `root`'s buffer to all peers (parameter sync at start; reference
// _pytorch_trial.py:602-606 hvd.broadcast_parameters / DDP ctor broadcast).  Pull model: every
// non-root rank reads the root's buffer.
__global__ void __launch_bounds__(kCommThreads)
broadcast_kernel(Peers peers, int rank, int world, int root, int64_t off, int64_t n, int phases,
                 long long timeout_cycles) {
  __shared__ uint32_t s_epoch;
  __shared__ bool s_last;
  Ctrl* me = peers.ctrl[rank];
  if (threadIdx.x == 0) s_epoch = me->epoch + 1;
  __syncthreads();
  const uint32_t epoch = s_epoch;
  if (phases & 1) {
    if (threadIdx.x < world && threadIdx.x != rank) {
      if (blockIdx.x == 0) st_release_sys(&peers.ctrl[threadIdx.x]->start[rank], epoch);
      wait_flag(&me->start[threadIdx.x], epoch, me, timeout_cycles);
    }
    __syncthreads();
  }
  if ((phases & 2) && rank != root) {
    const int64_t nvec = n >> 2;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    const float4* src = reinterpret_cast<const float4*>(peers.buf[root] + off);
    float4* dst = reinterpret_cast<float4*>(peers.buf[rank] + off);
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += stride) dst[i] = ld_peer(src + i);
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence_system();
    uint32_t t = atomicAdd(&me->blocks_done, 1u);
    s_last = (t == gridDim.x - 1);
  }
  __syncthreads();
  if (!s_last) return;
  if (phases & 4) {
    if (threadIdx.x < world && threadIdx.x != rank) {
      st_release_sys(&peers.ctrl[threadIdx.x]->done[rank], epoch);
      wait_flag(&me->done[threadIdx.x], epoch, me, timeout_cycles);
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    me->blocks_done = 0u;
    __threadfence();
    if (phases & 4) me->epoch = epoch;
  }
}

int fill_peers(Peers* P, void* const* bufs, void* const* ctrls, int world) {
  if (world < 1 || world > DETB200_MAX_RANKS) {
    detb200_set_error("world size %d outside [1,%d]", world, DETB200_MAX_RANKS);
    return DETB200_ERR_BAD_ARG;
  }
  for (int r = 0; r < world; ++r) {
    if (!bufs[r] || !ctrls[r]) { detb200_set_error("null peer pointer for rank %d", r); return DETB200_ERR_BAD_ARG; }
    P->buf[r] = (float*)bufs[r];
    P->ctrl[r] = (Ctrl*)ctrls[r];
  }
  return 0;
}

inline int comm_grid(int64_t n, int world) {
  int64_t per = (n / 4 + world - 1) / world;
  int g = grid_for(per, kCommThreads, 2, kNumSMs);
  return g;
}

}  // namespace

// ---------------------------------------------------------------------------------------------
// Symmetric memory plumbing (host side).  Plain cudaMalloc + legacy CUDA IPC: the handle is 64
// opaque bytes that ranks exchange through torch.distributed's object all-gather.
// ---------------------------------------------------------------------------------------------
DETB200_API int64_t detb200_ctrl_bytes() { return (int64_t)sizeof(Ctrl); }
DETB200_API int detb200_ipc_handle_bytes() { return (int)sizeof(cudaIpcMemHandle_t); }

DETB200_API int detb200_symm_alloc(void** ptr, int64_t bytes) {
  if (!ptr || bytes <= 0) { detb200_set_error("symm_alloc: bad argument"); return DETB200_ERR_BAD_ARG; }
  DETB200_CHECK(cudaMalloc(ptr, (size_t)bytes));
  DETB200_CHECK(cudaMemset(*ptr, 0, (size_t)bytes));
  DETB200_CHECK(cudaDeviceSynchronize());
  return 0;
}
DETB200_API int detb200_symm_free(void* ptr) {
  if (ptr) DETB200_CHECK(cudaFree(ptr));
  return 0;
}
DETB200_API int detb200_ipc_export(void* ptr, void* handle_out) {
  DETB200_CHECK(cudaIpcGetMemHandle((cudaIpcMemHandle_t*)handle_out, ptr));
  return 0;
}
DETB200_API int detb200_ipc_open(const void* handle, void** ptr_out) {
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, sizeof(h));
  DETB200_CHECK(cudaIpcOpenMemHandle(ptr_out, h, cudaIpcMemLazyEnablePeerAccess));
  return 0;
}
DETB200_API int detb200_ipc_close(void* ptr) {
  DETB200_CHECK(cudaIpcCloseMemHandle(ptr));
  return 0;
}
DETB200_API int detb200_enable_peer_access(int peer_device) {
  int dev = 0;
  DETB200_CHECK(cudaGetDevice(&dev));
  if (dev == peer_device) return 0;
  int can = 0;
  DETB200_CHECK(cudaDeviceCanAccessPeer(&can, dev, peer_device));
  if (!can) { detb200_set_error("device %d cannot access peer %d", dev, peer_device); return DETB200_ERR_UNSUPPORTED; }
  cudaError_t e = cudaDeviceEnablePeerAccess(peer_device, 0);
  if (e == cudaErrorPeerAccessAlreadyEnabled) { cudaGetLastError(); return 0; }
  DETB200_CHECK(e);
  return 0;
}

// stats_out[0] = sum of squares over the whole (reduced, scaled) step so far, stats_out[1] = non-finite count
DETB200_API int detb200_comm_read_stats(const void* ctrl, int world, float* stats_out, cudaStream_t stream) {
  float h[DETB200_MAX_RANKS * 2];
  DETB200_CHECK(cudaMemcpyAsync(h, ((const Ctrl*)ctrl)->stats, sizeof(float) * 2 * DETB200_MAX_RANKS, cudaMemcpyDeviceToHost, stream));
  DETB200_CHECK(cudaStreamSynchronize(stream));
  stats_out[0] = stats_out[1] = 0.f;
  for (int r = 0; r < world; ++r) { stats_out[0] += h[r]; stats_out[1] += h[DETB200_MAX_RANKS + r]; }
  return 0;
}
// Device pointer to the statistics table: float[2][MAX_RANKS] = {sumsq per rank, nonfinite per rank} for consumption by
// detb200_clip_coef without a host round trip.
DETB200_API void* detb200_comm_stats_ptr(void* ctrl) { return (void*)((Ctrl*)ctrl)->stats; }

DETB200_API int detb200_allreduce_fused(void* const* peer_bufs, void* const* peer_ctrls, int rank, int world,
                                        int64_t offset_elems, int64_t n_elems, float scale,
                                        const float* inv_loss_scale, int* found_inf, int first_of_step,
                                        int phases, double timeout_s, cudaStream_t stream) {
  Peers P;
  int rc = fill_peers(&P, peer_bufs, peer_ctrls, world);
  if (rc) return rc;
  if (rank < 0 || rank >= world) { detb200_set_error("bad rank %d", rank); return DETB200_ERR_BAD_ARG; }
  if (n_elems % 4 != 0 || offset_elems % 4 != 0) {
    detb200_set_error("allreduce_fused: offset/length must be multiples of 4 floats"); return DETB200_ERR_BAD_ARG;
  }
  if (n_elems == 0) return 0;
  long long cycles = (long long)(timeout_s * 1.9e9);
  allreduce_fused_kernel<<<comm_grid(n_elems, world), kCommThreads, 0, stream>>>(
      P, rank, world, offset_elems, n_elems, scale, inv_loss_scale, found_inf, first_of_step, phases, cycles);
  DETB200_LAUNCH_CHECK();
  return 0;
}

DETB200_API int detb200_rs_adam_ag(void* const* peer_grad_bufs, void* const* peer_ctrls, void* const* peer_param_bufs,
                                   int param_is_bf16, int rank, int world, int64_t offset_elems, int64_t n_elems,
                                   float* master_shard, float* m_shard, float* v_shard, float* grad_shard,
                                   float scale, const float* lr, const int* step, const float* grad_scale,
                                   const int* skip, float beta1, float beta2, float eps, float weight_decay,
                                   int mode, int first_of_step, int phases, double timeout_s, cudaStream_t stream) {
  Peers P;
  int rc = fill_peers(&P, peer_grad_bufs, peer_ctrls, world);
  if (rc) return rc;
  ParamPeers PP;
  for (int r = 0; r < world; ++r) PP.p_full[r] = peer_param_bufs[r];
  if (n_elems % (4 * world) != 0 || offset_elems % 4 != 0) {
    detb200_set_error("rs_adam_ag: bucket length must be a multiple of 4*world floats"); return DETB200_ERR_BAD_ARG;
  }
  if ((mode == 1 || mode == 2) && !grad_shard) { detb200_set_error("rs_adam_ag: two-phase mode needs grad_shard"); return DETB200_ERR_BAD_ARG; }
  if (n_elems == 0) return 0;
  long long cycles = (long long)(timeout_s * 1.9e9);
  int grid = comm_grid(n_elems, world);
  if (param_is_bf16)
    rs_adam_ag_kernel<true><<<grid, kCommThreads, 0, stream>>>(P, PP, rank, world, offset_elems, n_elems, master_shard,
        m_shard, v_shard, grad_shard, scale, lr, step, grad_scale, skip, beta1, beta2, eps, weight_decay, mode,
        first_of_step, phases, cycles);
  else
    rs_adam_ag_kernel<false><<<grid, kCommThreads, 0, stream>>>(P, PP, rank, world, offset_elems, n_elems, master_shard,
        m_shard, v_shard, grad_shard, scale, lr, step, grad_scale, skip, beta1, beta2, eps, weight_decay, mode,
        first_of_step, phases, cycles);
  DETB200_LAUNCH_CHECK();
  return 0;
}

DETB200_API int detb200_broadcast(void* const* peer_bufs, void* const* peer_ctrls, int rank, int world, int root,
                                  int64_t offset_elems, int64_t n_elems, int phases, double timeout_s,
                                  cudaStream_t stream) {
  Peers P;
  int rc = fill_peers(&P, peer_bufs, peer_ctrls, world);
  if (rc) return rc;
  if (n_elems % 4 != 0 || offset_elems % 4 != 0) { detb200_set_error("broadcast: length must be a multiple of 4 floats"); return DETB200_ERR_BAD_ARG; }
  if (n_elems == 0) return 0;
  long long cycles = (long long)(timeout_s * 1.9e9);
  broadcast_kernel<<<grid_for(n_elems / 4, kCommThreads, 2, kNumSMs), kCommThreads, 0, stream>>>(
      P, rank, world, root, offset_elems, n_elems, phases, cycles);
  DETB200_LAUNCH_CHECK();
  return 0;
}
