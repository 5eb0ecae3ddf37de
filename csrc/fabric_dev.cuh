// Device-side pieces of the fused federated aggregation shared by csrc/fabric.cu (the stand-alone
// aggregation kernel) and csrc/dec_sweep.cu (the d-theta sweep whose spare CTAs run the beta part of the
// aggregation over NVLink while the others keep the tensor cores busy).
#pragma once
#include <cuda_bf16.h>
#include <stdio.h>
#include <cuda.h>
#include "ptx.cuh"
#include "opt_math.cuh"
#include "../include/gfedntm_kernels.h"

namespace gf {

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float4 ld_peer_f4(const float* p) {
  float4 v;
  asm volatile("ld.relaxed.sys.global.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(p)
               : "memory");
  return v;
}
__device__ __forceinline__ float ld_peer_f1(const float* p) {
  float v;
  asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_peer_f4(float* p, float4 v) {
  asm volatile("st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z),
               "f"(v.w)
               : "memory");
}

// NVLS multicast stores: the address is the multicast alias of a symmetric buffer; the switch delivers
// the store to every rank of the group (including the issuing one).
__device__ __forceinline__ void mc_st_f4(float* p, float4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z),
               "f"(v.w)
               : "memory");
}
__device__ __forceinline__ void mc_st_b64(void* p, uint2 v) {
  const unsigned long long q = ((unsigned long long)v.y << 32) | v.x;
  asm volatile("multimem.st.relaxed.sys.global.b64 [%0], %1;" ::"l"(p), "l"(q) : "memory");
}
__device__ __forceinline__ void mc_st_b128(void* p, uint4 v) {
  mc_st_b64(p, make_uint2(v.x, v.y));
  mc_st_b64(reinterpret_cast<uint8_t*>(p) + 8, make_uint2(v.z, v.w));
}
// store helpers used by every aggregation path: multicast when the alias is present, else one store per rank
__device__ __forceinline__ void fed_push_f4(float* const* per_rank, float* mc, long long off, float4 v, int world) {
  if (mc != nullptr) { mc_st_f4(mc + off, v); return; }
  for (int c = 0; c < world; ++c) st_peer_f4(per_rank[c] + off, v);
}
__device__ __forceinline__ void fed_push_b64(void* const* per_rank, void* mc, size_t elem_off_bf16, uint2 v, int world) {
  if (mc != nullptr) { mc_st_b64(reinterpret_cast<__nv_bfloat16*>(mc) + elem_off_bf16, v); return; }
  for (int c = 0; c < world; ++c)
    if (per_rank[c] != nullptr) *reinterpret_cast<uint2*>(reinterpret_cast<__nv_bfloat16*>(per_rank[c]) + elem_off_bf16) = v;
}
__device__ __forceinline__ void fed_push_b128(void* const* per_rank, void* mc, size_t elem_off_bf16, uint4 v, int world) {
  if (mc != nullptr) { mc_st_b128(reinterpret_cast<__nv_bfloat16*>(mc) + elem_off_bf16, v); return; }
  for (int c = 0; c < world; ++c)
    if (per_rank[c] != nullptr) *reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(per_rank[c]) + elem_off_bf16) = v;
}

typedef OptH AdamF;  // (name kept from when Adam was the only solver; see opt_math.cuh)
__device__ __forceinline__ float adam_upd(float p, float g, float& m, float& v, const AdamF& h) {
  return opt_update(p, g, m, v, h);
}

// A peer that died (out of memory, crashed process) would otherwise leave every other rank spinning on its
// flag for ever, i.e. a hung GPU.  After kFedTimeoutNs without the flag the kernel traps: the process gets
// a CUDA error and the job fails instead of hanging.
constexpr unsigned long long kFedTimeoutNs = 120ull * 1000ull * 1000ull * 1000ull;
__device__ __forceinline__ unsigned long long fed_now_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ void fed_wait_flag(const uint32_t* f, uint32_t epoch, int peer, int phase) {
  const unsigned long long t0 = fed_now_ns();
  unsigned int polls = 0;
  while (ld_acquire_sys(f) < epoch) {
    __nanosleep(64);
    if ((++polls & 0x3fffu) == 0u && fed_now_ns() - t0 > kFedTimeoutNs) {
      printf("[gfedntm] aggregation: no flag from rank %d (phase %d, epoch %u) after 120 s - aborting\n", peer, phase,
             epoch);
      __trap();
    }
  }
}

// all threads of the block return once every peer's flag of `phase` has reached `epoch`
__device__ __forceinline__ void wait_peers(const GfFabricRank& me, int world, int phase, uint32_t epoch) {
  if (threadIdx.x < (unsigned)world)
    fed_wait_flag(me.flags[me.rank] + phase * GF_MAX_RANKS + threadIdx.x, epoch, (int)threadIdx.x, phase);
  __syncthreads();
}


// 256 threads of the calling block (threadIdx.x < 256) synchronise on named barrier 2
__device__ __forceinline__ void fed_sync256() { asm volatile("bar.sync 2, 256;" ::: "memory"); }

// Adam constants for step `t` (bias corrections on the device: graph replays read t from a counter)
__device__ __forceinline__ AdamF fed_adam_consts(const GfFedArgs& a, long long t) {
  AdamF h;
  h.lr = a.lr; h.b1 = a.b1; h.b2 = a.b2; h.eps = a.eps; h.wd = a.wd; h.solver = a.solver;
  h.bc1 = 1.f - powf(h.b1, (float)t);
  h.bc2_sqrt = sqrtf(1.f - powf(h.b2, (float)t));
  return h;
}

// beta [K,V] part of the aggregation for the word-tile range owned by `me.rank`: reduce the peers'
// gradients (P2P loads), Adam on the owner's shard, push fp32 beta and both bf16 operand shadows to every
// rank (P2P stores).  Called by threads 0..255 of `n_cta` cooperating blocks (`cta` = index among them);
// `tile_v` is 32*132 floats of shared memory.
__device__ __forceinline__ void fed_beta_phase(const GfFedArgs& a, const GfFabricRank& me, const AdamF& h, int cta,
                                               int n_cta, float* tile_v) {
  const int world = a.world, rank = me.rank;
  float(*tile)[33] = reinterpret_cast<float(*)[33]>(tile_v);  // 32 x 32 view used by the scalar fallback
  if (a.K > 0 && a.V % 4 == 0 && a.ldb % 4 == 0 && a.Kp % 16 == 0) {
    // vectorised: 32 topics x 128 words per tile, float4 peer loads (all peers of a row in flight
    // together), float4 / 8-byte / 32-byte stores to every rank
    float(*tv)[132] = reinterpret_cast<float(*)[132]>(tile_v);
    const int vt = (a.V + 127) / 128, kt = (a.K + 31) / 32;
    const int per = (vt + world - 1) / world;
    const int t_lo = rank * per, t_hi = (t_lo + per < vt) ? t_lo + per : vt;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const long long n_tiles = (long long)(t_hi > t_lo ? t_hi - t_lo : 0) * kt;
    for (long long t = cta; t < n_tiles; t += n_cta) {
      const int v0 = (t_lo + (int)(t / kt)) * 128, k0 = (int)(t % kt) * 32;
      // All peer loads of the thread's four rows are issued before the first use (one NVLink round trip
      // per tile instead of four); more than four peers are reduced in two groups to bound registers.
      const int vv = v0 + tx * 4;
      float4 gsum[4];
      bool live[4];
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        gsum[r] = make_float4(0.f, 0.f, 0.f, 0.f);
        live[r] = (k0 + ty + r * 8 < a.K) && (vv < a.V);
      }
#pragma unroll
      for (int c0 = 0; c0 < GF_MAX_RANKS; c0 += 4) {
        if (c0 < world) {
          float4 in[4][4];
#pragma unroll
          for (int r = 0; r < 4; ++r) {
            const long long ip = a.beta_off + (long long)(k0 + ty + r * 8) * a.V + vv;
#pragma unroll
            for (int cc = 0; cc < 4; ++cc)
              if (live[r] && c0 + cc < world) in[r][cc] = ld_peer_f4((a.mode == 0 ? me.G[c0 + cc] : me.P[c0 + cc]) + ip);
          }
#pragma unroll
          for (int r = 0; r < 4; ++r) {
#pragma unroll
            for (int cc = 0; cc < 4; ++cc)
              if (live[r] && c0 + cc < world) {
                const float w = (a.mode == 0) ? 1.f : a.w[c0 + cc];
                gsum[r].x = fmaf(w, in[r][cc].x, gsum[r].x); gsum[r].y = fmaf(w, in[r][cc].y, gsum[r].y);
                gsum[r].z = fmaf(w, in[r][cc].z, gsum[r].z); gsum[r].w = fmaf(w, in[r][cc].w, gsum[r].w);
              }
          }
        }
      }
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        const int k = k0 + ty + r * 8;
        float4 val = make_float4(0.f, 0.f, 0.f, 0.f);
        if (live[r]) {
          const long long ip = a.beta_off + (long long)k * a.V + vv;
          if (a.mode == 0) {
            val = *reinterpret_cast<const float4*>(me.P[rank] + ip);
            float4 m = *reinterpret_cast<const float4*>(me.M + ip);
            float4 v2 = *reinterpret_cast<const float4*>(me.V2 + ip);
            val.x = adam_upd(val.x, gsum[r].x, m.x, v2.x, h);
            val.y = adam_upd(val.y, gsum[r].y, m.y, v2.y, h);
            val.z = adam_upd(val.z, gsum[r].z, m.z, v2.z, h);
            val.w = adam_upd(val.w, gsum[r].w, m.w, v2.w, h);
            *reinterpret_cast<float4*>(me.M + ip) = m;
            *reinterpret_cast<float4*>(me.V2 + ip) = v2;
          } else {
            val = gsum[r];
          }
          __align__(8) __nv_bfloat16 o[4] = {__float2bfloat16(val.x), __float2bfloat16(val.y),
                                            __float2bfloat16(val.z), __float2bfloat16(val.w)};
          fed_push_f4(me.P, me.P_mc, ip, val, world);
          fed_push_b64(me.beta_bf, me.beta_bf_mc, (size_t)k * a.ldb + vv, *reinterpret_cast<const uint2*>(o), world);
        }
        *reinterpret_cast<float4*>(&tv[ty + r * 8][tx * 4]) = val;
      }
      fed_sync256();
      const int wv = threadIdx.x >> 1, half = threadIdx.x & 1;
      const int vw = v0 + wv, kb = k0 + half * 16;
      if (vw < a.V && kb < a.Kp) {
        __align__(16) __nv_bfloat16 o[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) o[j] = __float2bfloat16(tv[half * 16 + j][wv]);
        fed_push_b128(me.betaT_bf, me.betaT_bf_mc, (size_t)vw * a.Kp + kb, reinterpret_cast<const uint4*>(o)[0], world);
        fed_push_b128(me.betaT_bf, me.betaT_bf_mc, (size_t)vw * a.Kp + kb + 8, reinterpret_cast<const uint4*>(o)[1], world);
      }
      fed_sync256();
    }
  } else if (a.K > 0) {
    const int vt = (a.V + 31) / 32, kt = (a.K + 31) / 32;
    const int per = (vt + world - 1) / world;
    const int t_lo = rank * per, t_hi = (t_lo + per < vt) ? t_lo + per : vt;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const long long n_tiles = (long long)(t_hi > t_lo ? t_hi - t_lo : 0) * kt;
    for (long long t = cta; t < n_tiles; t += n_cta) {
      const int v0 = (t_lo + (int)(t / kt)) * 32, k0 = (int)(t % kt) * 32;
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        const int k = k0 + ty + r * 8, vv = v0 + tx;
        float val = 0.f;
        if (k < a.K && vv < a.V) {
          const long long ip = a.beta_off + (long long)k * a.V + vv;
          if (a.mode == 0) {
            float g = 0.f;
            for (int c = 0; c < world; ++c) g += ld_peer_f1(me.G[c] + ip);
            float m = me.M[ip], v2 = me.V2[ip];
            val = adam_upd(me.P[rank][ip], g, m, v2, h);
            me.M[ip] = m;
            me.V2[ip] = v2;
          } else {
            for (int c = 0; c < world; ++c) val = fmaf(a.w[c], ld_peer_f1(me.P[c] + ip), val);
          }
          const __nv_bfloat16 bv = __float2bfloat16(val);
          for (int c = 0; c < world; ++c) {
            me.P[c][ip] = val;
            if (me.beta_bf[c] != nullptr)
              reinterpret_cast<__nv_bfloat16*>(me.beta_bf[c])[(size_t)k * a.ldb + vv] = bv;
          }
        }
        tile[ty + r * 8][tx] = val;
      }
      fed_sync256();
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        const int vv = v0 + ty + r * 8, k = k0 + tx;
        if (vv < a.V && k < a.K) {
          const __nv_bfloat16 bv = __float2bfloat16(tile[tx][ty + r * 8]);
          for (int c = 0; c < world; ++c)
            if (me.betaT_bf[c] != nullptr)
              reinterpret_cast<__nv_bfloat16*>(me.betaT_bf[c])[(size_t)vv * a.Kp + k] = bv;
        }
      }
      fed_sync256();
    }
  }

}

// ---------------------------------------------------------------------------------------------------
// Overlapped beta phase (gradient mode), run by CTAs of the d-theta sweep (csrc/dec_sweep.cu): the CTAs
// set aside for it from the start, and every sweep CTA once it has run out of sweep work.
//
// Exchange format: every rank's beta gradient is bf16 [K, ld_gb] (GF_COL_GRAD2 writes it directly, the client
// weight is folded into the backward pass).  The owner of a word-tile range
//   - pulls the peers' 32-topic x 128-word tiles with ONE 2-D TMA load per peer and tile (plain
//     cp.async.bulk.tensor on a tensor map of that peer's buffer, no multicast) into a shared-memory ring,
//   - sums them with its own gradient in fp32 and applies Adam to ITS fp32 master copy and moments (which
//     never leave the owner: gf_fed_gather_beta collects the master copy when somebody needs all of it),
//   - pushes only the bf16 operand shadows beta_bf / betaT_bf to every rank (multimem.st on the NVLS
//     alias, or one P2P store per rank).
// Word tiles are handed out by an atomic counter (me.agg_counter), so any number of CTAs can work on them
// and latecomers simply take what is left.
//   threads 0..255   consumers
//   thread  256      producer (claims word tiles, issues the TMA loads)
// `smem` must be 1024-byte aligned; layout: [2 x tile 16x132 f32][pointer table][transposed staging 128 x (Kp+8) bf16]
// [stages x n_remote x 8 KB][item table][mbarriers].  Requires V % 8 == 0, ldb % 4 == 0, Kp % 16 == 0.
// ---------------------------------------------------------------------------------------------------
struct FedTmaps {
  CUtensorMap m[GF_MAX_RANKS];  // every rank's bf16 beta gradient [K, ld_gb], boxes of kFedRows topics x 128 words
  CUtensorMap loc[3];           // this rank's fp32 beta, first and second optimiser moment [K, V], same boxes
};
constexpr int kFedRows = 16;                            // topics per work item
constexpr int kFedGradBytes = kFedRows * 128 * 2;       // one rank's bf16 gradient tile (4 KB)
constexpr int kFedStateBytes = kFedRows * 128 * 4;      // one fp32 state tile (8 KB)
constexpr int kFedMaxStages = 8;

// Shared-memory stage of one work item (kFedRows topics x 128 words): [P][M][V2] fp32, own gradient, the peers'
// gradients.  EVERY operand arrives by TMA, issued stages ahead by the producer thread, so the 256 consumer
// threads never wait on a global load: with the operands loaded into registers by the consumers themselves the
// loop ran at ~0.5 element/ns per SM (one exposed HBM round trip per item and no second CTA on the SM to hide
// it), which is what the exposed part of the aggregation was made of.
static __device__ __noinline__ void fed_beta_phase_tma(const GfFedArgs& a, const GfFabricRank& me, const AdamF& h,
                                                   const FedTmaps& tm, uint8_t* smem, uint32_t smem_bytes) {
  const int world = a.world, rank = me.rank, n_remote = world - 1;
  float* tile_v = reinterpret_cast<float*>(smem);
  float(*tv)[132] = reinterpret_cast<float(*)[132]>(tile_v);  // kFedRows x 128 (+4 pad): transposition buffer
  // transposed-shadow staging: 128 words x Kp topics bf16 (rows padded by 8), flushed as whole rows
  const int ldt = a.Kp + 8;
  // [0, 16896): two transposition buffers of 16 x 132 floats; [16896, 17408): the peers' shadow pointers
  void** psh = reinterpret_cast<void**>(smem + 16896);
  constexpr uint32_t kHead = 17408;
  __nv_bfloat16* tT = reinterpret_cast<__nv_bfloat16*>(smem + kHead);
  const uint32_t tT_bytes = ((uint32_t)(128 * ldt * 2) + 1023u) & ~1023u;
  uint8_t* ring = smem + kHead + tT_bytes;
  const uint32_t stage_bytes = 3u * kFedStateBytes + (uint32_t)world * kFedGradBytes;
  int stages = (int)((smem_bytes - kHead - tT_bytes - 512) / stage_bytes);
  if (stages > kFedMaxStages) stages = kFedMaxStages;
  if (stages < 1) stages = 1;  // (the launcher sizes the shared memory so that this never triggers)
  int* item_wt = reinterpret_cast<int*>(ring + (size_t)stages * stage_bytes);
  int* item_k = item_wt + kFedMaxStages;
  uint64_t* full = reinterpret_cast<uint64_t*>(item_k + kFedMaxStages);
  uint64_t* empty = full + kFedMaxStages;
  if (threadIdx.x == 0) {
    for (int i = 0; i < stages; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], 1);
    }
    mbar_fence_init();
  }
  for (int i = threadIdx.x; i < 128 * ldt / 2; i += blockDim.x) reinterpret_cast<uint32_t*>(tT)[i] = 0u;
  if (threadIdx.x < GF_MAX_RANKS) {
    psh[threadIdx.x] = me.beta_bf[threadIdx.x];
    psh[GF_MAX_RANKS + threadIdx.x] = me.betaT_bf[threadIdx.x];
  }
  __syncthreads();  // all threads of the CTA call this function

  const int vt = (a.V + 127) / 128, kt = (a.K + kFedRows - 1) / kFedRows;
  const int per = (vt + world - 1) / world;
  const int t_lo = rank * per, t_hi = (t_lo + per < vt) ? t_lo + per : vt;
  const int n_wt = t_hi > t_lo ? t_hi - t_lo : 0;  // word tiles owned by this rank; a CTA takes whole word tiles

  if (threadIdx.x == 256) {
    // ---------------- producer
    uint32_t it = 0;
    for (;;) {
      const int wt = (int)atomicAdd(me.agg_counter, 1u);
      const bool done = wt >= n_wt;
      const int v0 = (t_lo + wt) * 128;
      for (int kti = 0; kti < (done ? 1 : kt); ++kti, ++it) {
        const int st = it % stages;
        mbar_wait(&empty[st], ((it / stages) & 1) ^ 1);
        item_wt[st] = done ? -1 : wt;
        item_k[st] = kti;
        if (done) {
          mbar_arrive(&full[st]);
        } else {
          uint8_t* sb = ring + (size_t)st * stage_bytes;
          mbar_arrive_expect_tx(&full[st], stage_bytes);  // out-of-range rows / words are zero-filled and counted
          const int k0 = kti * kFedRows;
          tma_load_2d(sb, &tm.loc[0], &full[st], v0, k0);
          tma_load_2d(sb + kFedStateBytes, &tm.loc[1], &full[st], v0, k0);
          tma_load_2d(sb + 2 * kFedStateBytes, &tm.loc[2], &full[st], v0, k0);
          tma_load_2d(sb + 3 * kFedStateBytes, &tm.m[rank], &full[st], v0, k0);
          int slot = 1;
          for (int c = 0; c < world; ++c) {
            if (c == rank) continue;
            tma_load_2d(sb + 3 * kFedStateBytes + (size_t)slot * kFedGradBytes, &tm.m[c], &full[st], v0, k0);
            ++slot;
          }
        }
      }
      if (done) break;
    }
  } else if (threadIdx.x < 256) {
    // ---------------- consumers: thread (ty, tx) owns words 4 tx .. 4 tx + 3 of topics ty and ty + 8 of the item
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    // Everything the loop needs from the argument blocks is copied into registers / shared memory first: both
    // live in global memory, and with stores to global memory in the loop the compiler has to reload every field
    // (sizes, strides, the pointer tables) in front of each use otherwise.
    const int K = a.K, V = a.V, Kp = a.Kp, ldb = a.ldb;
    const long long beta_off = a.beta_off;
    float* const Pm = me.P[rank];
    float* const Mm = me.M;
    float* const Vm = me.V2;
    void* const bf_mc = me.beta_bf_mc;
    void* const bT_mc = me.betaT_bf_mc;
    void* const* bf_tab = psh;
    void* const* bT_tab = psh + GF_MAX_RANKS;
    const int wv = threadIdx.x >> 1, half = threadIdx.x & 1;
    // optimiser constants in registers (`h` is a reference into the caller's frame); Adam without weight decay, the
    // usual case, gets a path with the two constant divisions hoisted and no solver dispatch per element.  With
    // eight consumer warps per SM this loop is bound by instruction issue, so every instruction per element counts.
    const AdamF hl = h;
    const OptFast hf = opt_fast(hl);
    auto upd = [&](float p, float g, float& m, float& v) -> float {
      return hf.on ? opt_update_fast(p, g, m, v, hf) : opt_update(p, g, m, v, hl);
    };
    const uint32_t ring_s = smem_u32(ring), tv_s = smem_u32(tile_v), tT_s = smem_u32(tT);
    const int cpr = Kp >> 3;  // 16-byte chunks per row of the transposed shadow
    const int rw0 = (int)threadIdx.x / cpr, ch0 = (int)threadIdx.x % cpr, dq = 256 / cpr, dr = 256 % cpr;
    for (uint32_t it = 0;; ++it) {
      const int st = it % stages;
      mbar_wait(&full[st], (it / stages) & 1);
      const int wt = item_wt[st], kti = item_k[st];
      if (wt < 0) break;
      const int v0 = (t_lo + wt) * 128;
      const int vv = v0 + tx * 4;
      const int k0 = kti * kFedRows;
      const uint32_t sb = ring_s + (uint32_t)st * stage_bytes;                    // shared-space addresses: LDS / STS,
      const uint32_t tvb = tv_s + (it & 1) * (uint32_t)(kFedRows * 132 * 4);     // not generic loads and stores
#pragma unroll
      for (int r = 0; r < kFedRows / 8; ++r) {
        const int row = ty + r * 8, k = k0 + row;
        float4 val = make_float4(0.f, 0.f, 0.f, 0.f);
        if (k < K && vv < V) {
          const uint32_t o4 = (uint32_t)row * 512u + (uint32_t)tx * 16u;  // fp32 tiles: 512-byte rows
          const uint32_t o2 = (uint32_t)row * 256u + (uint32_t)tx * 8u;   // bf16 tiles: 256-byte rows
          val = lds_f4(sb + o4);
          float4 mm = lds_f4(sb + kFedStateBytes + o4);
          float4 v2 = lds_f4(sb + 2 * kFedStateBytes + o4);
          float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
          for (int c = 0; c < world; ++c) {  // own tile first, then the peers' in rank order
            const float2 xr = lds_f2(sb + 3 * kFedStateBytes + (uint32_t)c * kFedGradBytes + o2);
            const uint32_t x0 = __float_as_uint(xr.x), x1 = __float_as_uint(xr.y);  // 4 bf16: widen by shifting
            g.x += __uint_as_float(x0 << 16); g.y += __uint_as_float(x0 & 0xffff0000u);
            g.z += __uint_as_float(x1 << 16); g.w += __uint_as_float(x1 & 0xffff0000u);
          }
          const long long ip = beta_off + (long long)k * V + vv;
          val.x = upd(val.x, g.x, mm.x, v2.x);
          val.y = upd(val.y, g.y, mm.y, v2.y);
          val.z = upd(val.z, g.z, mm.z, v2.z);
          val.w = upd(val.w, g.w, mm.w, v2.w);
          *reinterpret_cast<float4*>(Pm + ip) = val;  // master copy: owner only
          *reinterpret_cast<float4*>(Mm + ip) = mm;
          *reinterpret_cast<float4*>(Vm + ip) = v2;
          uint2 o;
          o.x = pack_bf16(val.x, val.y);
          o.y = pack_bf16(val.z, val.w);
          fed_push_b64(bf_tab, bf_mc, (size_t)k * ldb + vv, o, world);
        }
        sts_u4(tvb + (uint32_t)(row * 132 + tx * 4) * 4u, __float_as_uint(val.x), __float_as_uint(val.y),
               __float_as_uint(val.z), __float_as_uint(val.w));
      }
      fed_sync256();  // the item's values are in tvb; every consumer is done with the stage and its item entry
      if (threadIdx.x == 0) mbar_arrive(&empty[st]);
      // transpose this 16-topic slab into the row-major [word][topic] staging tile: thread = (word, 8 topics).
      // (tvb is rewritten two items later, i.e. after the next item's barrier, which every thread reaches only
      // after this read.)
      const int kb = k0 + half * 8;
      if (kb < Kp) {
        uint32_t o[4];
#pragma unroll
        for (int j = 0; j < 8; j += 2) {
          float lo, hi;
          asm volatile("ld.shared.f32 %0, [%1];" : "=f"(lo) : "r"(tvb + (uint32_t)((half * 8 + j) * 132 + wv) * 4u));
          asm volatile("ld.shared.f32 %0, [%1];" : "=f"(hi) : "r"(tvb + (uint32_t)((half * 8 + j + 1) * 132 + wv) * 4u));
          o[j >> 1] = pack_bf16(lo, hi);
        }
        sts_u4(tT_s + (uint32_t)(wv * ldt + kb) * 2u, o[0], o[1], o[2], o[3]);
      }
      if (kti + 1 == kt) {
        fed_sync256();
        // flush the transposed shadow of this word tile: whole rows of Kp bf16, 16-byte pieces, consecutive
        // threads -> consecutive addresses (large NVLink packets instead of one 16-byte store per topic slab)
        // (global offsets are linear in idx; the padded staging rows need (row, chunk), stepped without a division)
        int rw = rw0, ch = ch0;
        for (int idx = threadIdx.x; idx < 128 * cpr; idx += 256) {
          if (v0 + rw < V) {
            uint4 q;
            asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];"
                         : "=r"(q.x), "=r"(q.y), "=r"(q.z), "=r"(q.w)
                         : "r"(tT_s + (uint32_t)(rw * ldt + ch * 8) * 2u));
            fed_push_b128(bT_tab, bT_mc, (size_t)v0 * Kp + (size_t)idx * 8, q, world);
          }
          rw += dq;
          ch += dr;
          if (ch >= cpr) { ch -= cpr; ++rw; }
        }
        fed_sync256();
      }
    }
  }
}

}  // namespace gf
