// Fused federated aggregation over NVLink / NVSwitch peer memory (no NCCL on this path).
//
// Replaces the reference's per-minibatch gRPC gather -> numpy weighted mean -> gRPC broadcast
// (src/federation/server.py:442-521) and the clients' optimizer.step()/load_state_dict
// (federated_avitm.py:78, federated_model.py:117-131) with ONE kernel per rank:
//
//   phase A  signal "my gradients are ready" to every peer, wait for all peers        (flags)
//   phase B  for the slice of the flat parameter space this rank owns:
//              g = sum_c G_c[i]            P2P loads from every peer (reduce-scatter; the client
//                                          weight n_c / sum n is already folded into G_c)
//              Adam(m, v, p)               sharded optimiser state, fp32 master copy
//              P_c[i] = p  for every c     P2P stores (all-gather), plus for beta the bf16
//                                          operand shadows beta_bf / betaT_bf the next forward's
//                                          tensor-core sweeps read
//            mode PARAM: p = sum_c w_c P_c[i] instead (reference-parity parameter averaging)
//            BatchNorm running statistics are averaged with the client weights in both modes.
//   phase C  signal completion, wait until every peer has finished pushing into this rank.
//
// Memory ordering: data stores -> __threadfence_system() -> st.release.sys flag; consumers poll
// with ld.acquire.sys.  Flags carry a monotonically increasing epoch, so they never need a reset.
//
// Single-GPU emulation (tests): the same kernel is launched once with gridDim.y = world and one
// GfFabricRank context per y-slice, so all "ranks" are co-resident blocks of one grid.
#include "ptx.cuh"
#include "tmap.h"
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

struct AdamF {
  float lr, b1, b2, eps, bc1, bc2_sqrt, wd;
};
__device__ __forceinline__ float adam_upd(float p, float g, float& m, float& v, const AdamF& h) {
  if (h.wd != 0.f) g = fmaf(h.wd, p, g);
  m = fmaf(h.b1, m, (1.f - h.b1) * g);
  v = fmaf(h.b2, v, (1.f - h.b2) * g * g);
  const float denom = sqrtf(v) / h.bc2_sqrt + h.eps;
  return p - (h.lr / h.bc1) * (m / denom);
}

// all threads of the block return once every peer's flag of `phase` has reached `epoch`
__device__ __forceinline__ void wait_peers(const GfFabricRank& me, int world, int phase, uint32_t epoch) {
  if (threadIdx.x < (unsigned)world) {
    const uint32_t* f = me.flags[me.rank] + phase * GF_MAX_RANKS + threadIdx.x;
    while (ld_acquire_sys(f) < epoch) {
      __nanosleep(64);
    }
  }
  __syncthreads();
}

__global__ void __launch_bounds__(256) fed_aggregate_kernel(const GfFedArgs a, AdamF h) {
  __shared__ __align__(16) float tile_v[32 * 132];  // 32 x 128 (+4 pad) tile of the vectorised beta phase
  float(*tile)[33] = reinterpret_cast<float(*)[33]>(tile_v);  // 32 x 32 view used by the scalar fallback
  __shared__ int is_last;
  const GfFabricRank& me = a.ranks[blockIdx.y];
  const int world = a.world, rank = me.rank;
  uint32_t epoch = (uint32_t)a.epoch;
  if (a.step_ctr != nullptr) {  // graph replay: epoch and Adam step come from the device counters
    const long long t = a.step_ctr[1];
    epoch = (uint32_t)t;
    h.bc1 = 1.f - powf(h.b1, (float)t);
    h.bc2_sqrt = sqrtf(1.f - powf(h.b2, (float)t));
  }

  // ---------------- phase A: publish "gradients ready", wait for everyone
  if (blockIdx.x == 0 && threadIdx.x < (unsigned)world) {
    __threadfence_system();
    st_release_sys(me.flags[threadIdx.x] + 0 * GF_MAX_RANKS + rank, epoch);
  }
  wait_peers(me, world, 0, epoch);

  const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long nthr = (long long)gridDim.x * blockDim.x;

  // ---------------- phase B.1: flat region [0, flat_n), float4 granularity, owner = contiguous slice
  {
    const long long n4 = (a.flat_n + 3) / 4;
    const long long per = (n4 + world - 1) / world;
    const long long lo = rank * per, hi = (lo + per < n4) ? lo + per : n4;
    for (long long q = lo + tid; q < hi; q += nthr) {
      const long long i = q * 4;
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
      float4 in[GF_MAX_RANKS];
      // issue every peer load before the first use so all of them are in flight together
#pragma unroll
      for (int c = 0; c < GF_MAX_RANKS; ++c)
        if (c < world) in[c] = ld_peer_f4((a.mode == 0 ? me.G[c] : me.P[c]) + i);
      if (a.mode == 0) {
#pragma unroll
        for (int c = 0; c < GF_MAX_RANKS; ++c)
          if (c < world) { acc.x += in[c].x; acc.y += in[c].y; acc.z += in[c].z; acc.w += in[c].w; }
        float4 p = *reinterpret_cast<const float4*>(me.P[rank] + i);
        float4 m = *reinterpret_cast<const float4*>(me.M + i);
        float4 v = *reinterpret_cast<const float4*>(me.V2 + i);
        p.x = adam_upd(p.x, acc.x, m.x, v.x, h);
        p.y = adam_upd(p.y, acc.y, m.y, v.y, h);
        p.z = adam_upd(p.z, acc.z, m.z, v.z, h);
        p.w = adam_upd(p.w, acc.w, m.w, v.w, h);
        *reinterpret_cast<float4*>(me.M + i) = m;
        *reinterpret_cast<float4*>(me.V2 + i) = v;
        acc = p;
      } else {
#pragma unroll
        for (int c = 0; c < GF_MAX_RANKS; ++c)
          if (c < world) {
            const float w = a.w[c];
            acc.x = fmaf(w, in[c].x, acc.x); acc.y = fmaf(w, in[c].y, acc.y);
            acc.z = fmaf(w, in[c].z, acc.z); acc.w = fmaf(w, in[c].w, acc.w);
          }
      }
      for (int c = 0; c < world; ++c) st_peer_f4(me.P[c] + i, acc);
    }
  }

  // ---------------- phase B.2: BatchNorm running statistics (weighted mean in both modes)
  {
    const long long n4 = (a.s_n + 3) / 4;
    const long long per = (n4 + world - 1) / world;
    const long long lo = rank * per, hi = (lo + per < n4) ? lo + per : n4;
    for (long long q = lo + tid; q < hi; q += nthr) {
      const long long i = q * 4;
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int c = 0; c < world; ++c) {
        const float4 s = ld_peer_f4(me.S[c] + i);
        const float w = a.w[c];
        acc.x = fmaf(w, s.x, acc.x); acc.y = fmaf(w, s.y, acc.y);
        acc.z = fmaf(w, s.z, acc.z); acc.w = fmaf(w, s.w, acc.w);
      }
      for (int c = 0; c < world; ++c) st_peer_f4(me.S[c] + i, acc);
    }
  }

  // ---------------- phase B.3: beta [K,V] in 32x32 tiles, owner = contiguous range of word tiles;
  //                  writes fp32 beta, bf16 beta_bf [Kp,ldb] and bf16 betaT_bf [V,Kp] to every rank
  if (a.K > 0 && a.V % 4 == 0 && a.ldb % 4 == 0 && a.Kp % 16 == 0) {
    // vectorised: 32 topics x 128 words per tile, float4 peer loads (all peers of a row in flight
    // together), float4 / 8-byte / 32-byte stores to every rank
    float(*tv)[132] = reinterpret_cast<float(*)[132]>(tile_v);
    const int vt = (a.V + 127) / 128, kt = (a.K + 31) / 32;
    const int per = (vt + world - 1) / world;
    const int t_lo = rank * per, t_hi = (t_lo + per < vt) ? t_lo + per : vt;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const long long n_tiles = (long long)(t_hi > t_lo ? t_hi - t_lo : 0) * kt;
    for (long long t = blockIdx.x; t < n_tiles; t += gridDim.x) {
      const int v0 = (t_lo + (int)(t / kt)) * 128, k0 = (int)(t % kt) * 32;
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        const int k = k0 + ty + r * 8, vv = v0 + tx * 4;
        float4 val = make_float4(0.f, 0.f, 0.f, 0.f);
        if (k < a.K && vv < a.V) {
          const long long ip = a.beta_off + (long long)k * a.V + vv;
          float4 in[GF_MAX_RANKS];
#pragma unroll
          for (int c = 0; c < GF_MAX_RANKS; ++c)
            if (c < world) in[c] = ld_peer_f4((a.mode == 0 ? me.G[c] : me.P[c]) + ip);
          if (a.mode == 0) {
            float4 gsum = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
            for (int c = 0; c < GF_MAX_RANKS; ++c)
              if (c < world) { gsum.x += in[c].x; gsum.y += in[c].y; gsum.z += in[c].z; gsum.w += in[c].w; }
            val = *reinterpret_cast<const float4*>(me.P[rank] + ip);
            float4 m = *reinterpret_cast<const float4*>(me.M + ip);
            float4 v2 = *reinterpret_cast<const float4*>(me.V2 + ip);
            val.x = adam_upd(val.x, gsum.x, m.x, v2.x, h);
            val.y = adam_upd(val.y, gsum.y, m.y, v2.y, h);
            val.z = adam_upd(val.z, gsum.z, m.z, v2.z, h);
            val.w = adam_upd(val.w, gsum.w, m.w, v2.w, h);
            *reinterpret_cast<float4*>(me.M + ip) = m;
            *reinterpret_cast<float4*>(me.V2 + ip) = v2;
          } else {
#pragma unroll
            for (int c = 0; c < GF_MAX_RANKS; ++c)
              if (c < world) {
                const float w = a.w[c];
                val.x = fmaf(w, in[c].x, val.x); val.y = fmaf(w, in[c].y, val.y);
                val.z = fmaf(w, in[c].z, val.z); val.w = fmaf(w, in[c].w, val.w);
              }
          }
          __align__(8) __nv_bfloat16 o[4] = {__float2bfloat16(val.x), __float2bfloat16(val.y),
                                            __float2bfloat16(val.z), __float2bfloat16(val.w)};
          for (int c = 0; c < world; ++c) {
            st_peer_f4(me.P[c] + ip, val);
            if (me.beta_bf[c] != nullptr)
              *reinterpret_cast<uint2*>(reinterpret_cast<__nv_bfloat16*>(me.beta_bf[c]) + (size_t)k * a.ldb + vv) =
                  *reinterpret_cast<const uint2*>(o);
          }
        }
        *reinterpret_cast<float4*>(&tv[ty + r * 8][tx * 4]) = val;
      }
      __syncthreads();
      const int wv = threadIdx.x >> 1, half = threadIdx.x & 1;
      const int vv = v0 + wv, kb = k0 + half * 16;
      if (vv < a.V && kb < a.Kp) {
        __align__(16) __nv_bfloat16 o[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) o[j] = __float2bfloat16(tv[half * 16 + j][wv]);
        for (int c = 0; c < world; ++c)
          if (me.betaT_bf[c] != nullptr) {
            uint4* dst = reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(me.betaT_bf[c]) + (size_t)vv * a.Kp + kb);
            dst[0] = reinterpret_cast<const uint4*>(o)[0];
            dst[1] = reinterpret_cast<const uint4*>(o)[1];
          }
      }
      __syncthreads();
    }
  } else if (a.K > 0) {
    const int vt = (a.V + 31) / 32, kt = (a.K + 31) / 32;
    const int per = (vt + world - 1) / world;
    const int t_lo = rank * per, t_hi = (t_lo + per < vt) ? t_lo + per : vt;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const long long n_tiles = (long long)(t_hi > t_lo ? t_hi - t_lo : 0) * kt;
    for (long long t = blockIdx.x; t < n_tiles; t += gridDim.x) {
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
      __syncthreads();
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
      __syncthreads();
    }
  }

  // ---------------- phase C: all my pushes are done -> tell the peers, wait for theirs
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int prev = atomicAdd(me.done_counter, 1u);
    is_last = (prev == gridDim.x - 1);
  }
  __syncthreads();
  if (is_last) {
    if (threadIdx.x == 0) *me.done_counter = 0;  // ready for the next launch
    if (threadIdx.x < (unsigned)world) {
      __threadfence_system();
      st_release_sys(me.flags[threadIdx.x] + 1 * GF_MAX_RANKS + rank, epoch);
    }
    wait_peers(me, world, 1, epoch);
  }
}

}  // namespace gf

using namespace gf;

extern "C" int gf_fed_aggregate(const GfFedArgs* a, void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  GF_ARG_CHECK(a->world >= 1 && a->world <= GF_MAX_RANKS, "world size must be in [1,8]");
  GF_ARG_CHECK(a->n_rank_ctx == 1 || a->n_rank_ctx == a->world, "n_rank_ctx must be 1 or world");
  GF_ARG_CHECK(a->mode == 0 || a->mode == 1, "mode");
  GF_ARG_CHECK(a->flat_n % 4 == 0 && a->beta_off % 4 == 0 && a->s_n % 4 == 0,
               "flat_n, beta_off and s_n must be multiples of 4 (16-byte vectors)");
  GF_ARG_CHECK(a->mode == 1 || a->step >= 1 || a->step_ctr != nullptr, "Adam step must be >= 1");
  AdamF h;
  h.lr = a->lr; h.b1 = a->b1; h.b2 = a->b2; h.eps = a->eps; h.wd = a->wd;
  const int step = a->step >= 1 ? a->step : 1;
  h.bc1 = 1.f - powf(a->b1, (float)step);
  h.bc2_sqrt = sqrtf(1.f - powf(a->b2, (float)step));
  int blocks = 2 * num_sms();
  if (a->n_rank_ctx > 1) {
    // emulation: every rank's blocks must be co-resident (they wait on each other's flags)
    blocks = num_sms() / a->n_rank_ctx;
    if (blocks < 1) blocks = 1;
  }
  dim3 grid(blocks, a->n_rank_ctx);
  fed_aggregate_kernel<<<grid, 256, 0, st>>>(*a, h);
  GF_LAUNCHED();
  return GF_OK;
}
