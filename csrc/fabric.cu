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
#include "fabric_dev.cuh"

namespace gf {

__global__ void __launch_bounds__(256) fed_aggregate_kernel(const GfFedArgs a, AdamF h) {
  __shared__ __align__(16) float tile_v[32 * 132];  // 32 x 128 (+4 pad) tile of the vectorised beta phase
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

  const int parts = a.parts ? a.parts : 3;  // bit 0: flat region + BatchNorm statistics, bit 1: beta

  // ---------------- phase B.1: flat region [0, flat_n), float4 granularity, owner = contiguous slice
  if (parts & 1) {
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
      fed_push_f4(me.P, me.P_mc, i, acc, world);
    }
  }

  // ---------------- phase B.2: BatchNorm running statistics (weighted mean in both modes)
  if (parts & 1) {
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
      fed_push_f4(me.S, me.S_mc, i, acc, world);
    }
  }

  // ---------------- phase B.3: beta [K,V], owner = contiguous range of word tiles; writes fp32 beta,
  //                  bf16 beta_bf [Kp,ldb] and bf16 betaT_bf [V,Kp] to every rank (fabric_dev.cuh)
  if (parts & 2) fed_beta_phase(a, me, h, (int)blockIdx.x, (int)gridDim.x, tile_v);

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

// Raises flag `phase` of this rank on every peer ("my beta gradient is final and I no longer read my
// beta shadows"): lets the peers' overlapped beta aggregation (d-theta sweep, csrc/dec_sweep.cu) start.
__global__ void fed_signal_kernel(const GfFedArgs a, int phase) {
  const GfFabricRank& me = a.ranks[0];
  uint32_t epoch = (uint32_t)a.epoch;
  if (a.step_ctr != nullptr) epoch = (uint32_t)a.step_ctr[1];
  if (threadIdx.x < (unsigned)a.world) {
    __threadfence_system();
    st_release_sys(me.flags[threadIdx.x] + phase * GF_MAX_RANKS + me.rank, epoch);
  }
}

}  // namespace gf

using namespace gf;

extern "C" int gf_fed_signal(const GfFedArgs* a, int phase, void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  GF_ARG_CHECK(a->world >= 1 && a->world <= GF_MAX_RANKS && a->n_rank_ctx == 1, "world / n_rank_ctx");
  GF_ARG_CHECK(phase >= 2 && phase < 8, "phase slots 0 and 1 belong to gf_fed_aggregate");
  fed_signal_kernel<<<1, 32, 0, st>>>(*a, phase);
  GF_LAUNCHED();
  return GF_OK;
}

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
