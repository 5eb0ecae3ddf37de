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
#include <string.h>

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

  // ---------------- phase B.1: flat region [0, flat_n) as a list of ranges (one range = everything when the
  // caller gave none), float4 granularity.  A shared range is split into contiguous owner slices; an unshared
  // range stays client-local: in gradient mode every rank applies its own Adam step to all of it (with the
  // client weight taken out of the gradient again), in parameter mode it is left alone.
  if (parts & 1) {
    const int n_ranges = a.n_ranges > 0 ? a.n_ranges : 1;
    for (int ri = 0; ri < n_ranges; ++ri) {
      const long long r_off = a.n_ranges > 0 ? a.ranges[ri].off : 0;
      const long long r_len = a.n_ranges > 0 ? a.ranges[ri].len : a.flat_n;
      const int kind = a.n_ranges > 0 ? a.ranges[ri].shared : 1;
      // kind 2 (replicated): G already holds the global gradient on every rank -> the same local Adam everywhere
      const bool shared = (kind == 1) || (kind >= 2 && a.mode != 0);
      const long long n4 = (r_len + 3) / 4;
      if (!shared) {
        if (a.mode != 0 || kind == 3) continue;  // kind 3: the caller applies the (replicated) update itself
        const float gs = (kind == 2) ? 1.0f : a.local_gscale;
        for (long long q = tid; q < n4; q += nthr) {
          const long long i = r_off + q * 4;
          float4 g = *reinterpret_cast<const float4*>(me.G[rank] + i);
          float4 p = *reinterpret_cast<const float4*>(me.P[rank] + i);
          float4 m = *reinterpret_cast<const float4*>(me.M + i);
          float4 v = *reinterpret_cast<const float4*>(me.V2 + i);
          p.x = adam_upd(p.x, g.x * gs, m.x, v.x, h);
          p.y = adam_upd(p.y, g.y * gs, m.y, v.y, h);
          p.z = adam_upd(p.z, g.z * gs, m.z, v.z, h);
          p.w = adam_upd(p.w, g.w * gs, m.w, v.w, h);
          *reinterpret_cast<float4*>(me.M + i) = m;
          *reinterpret_cast<float4*>(me.V2 + i) = v;
          *reinterpret_cast<float4*>(me.P[rank] + i) = p;
        }
        continue;
      }
      const long long per = (n4 + world - 1) / world;
      const long long lo = rank * per, hi = (lo + per < n4) ? lo + per : n4;
      for (long long q = lo + tid; q < hi; q += nthr) {
        const long long i = r_off + q * 4;
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
  }

  // ---------------- phase B.2: BatchNorm running statistics (weighted mean in both modes).  The buffer is
  // {mu mean, mu var, sigma mean, sigma var}[K] then {beta mean, beta var}[V]; bit b of s_mask = block b shared
  // (an unshared statistic keeps the client's own value).
  if (parts & 1) {
    const long long n4 = (a.s_n + 3) / 4;
    const long long per = (n4 + world - 1) / world;
    const long long lo = rank * per, hi = (lo + per < n4) ? lo + per : n4;
    const bool all_shared = (a.n_ranges == 0) || ((a.s_mask & 63) == 63);
    for (long long q = lo + tid; q < hi; q += nthr) {
      const long long i = q * 4;
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int c = 0; c < world; ++c) {
        const float4 s = ld_peer_f4(me.S[c] + i);
        const float w = a.w[c];
        acc.x = fmaf(w, s.x, acc.x); acc.y = fmaf(w, s.y, acc.y);
        acc.z = fmaf(w, s.z, acc.z); acc.w = fmaf(w, s.w, acc.w);
      }
      if (all_shared) {
        fed_push_f4(me.S, me.S_mc, i, acc, world);
      } else {
        const float vals[4] = {acc.x, acc.y, acc.z, acc.w};
        for (int e = 0; e < 4; ++e) {
          const long long j = i + e;
          const int blk = j < 4LL * a.K ? (int)(j / (a.K > 0 ? a.K : 1)) : (j < 4LL * a.K + a.V ? 4 : 5);
          if (j < 4LL * a.K + 2LL * a.V && ((a.s_mask >> blk) & 1))
            for (int c = 0; c < world; ++c) me.S[c][j] = vals[e];
        }
      }
    }
  }

  // ---------------- phase B.3: beta [K,V], owner = contiguous range of word tiles; writes fp32 beta,
  //                  bf16 beta_bf [Kp,ldb] and bf16 betaT_bf [V,Kp] to every rank (fabric_dev.cuh)
  if ((parts & 2) && (a.n_ranges == 0 || a.beta_shared)) fed_beta_phase(a, me, h, (int)blockIdx.x, (int)gridDim.x, tile_v);

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
  if (phase == GF_FED_PHASE_BETA_READY && threadIdx.x == 0 && me.agg_counter != nullptr) *me.agg_counter = 0u;
  if (threadIdx.x < (unsigned)a.world) {
    __threadfence_system();
    st_release_sys(me.flags[threadIdx.x] + phase * GF_MAX_RANKS + me.rank, epoch);
  }
}

// One-shot small all-reduce: push x into my slot on every rank, flag, wait, sum the slots in rank order.
__global__ void __launch_bounds__(256) fed_allreduce_small_kernel(const GfFedArgs a, float* __restrict__ x, long long n) {
  __shared__ int is_last;
  const GfFabricRank& me = a.ranks[0];
  const int world = a.world, rank = me.rank;
  uint32_t epoch = (uint32_t)a.epoch;
  if (a.step_ctr != nullptr) epoch = (uint32_t)a.step_ctr[1];
  const long long n4 = n / 4;
  const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x, nthr = (long long)gridDim.x * blockDim.x;
  for (long long q = tid; q < n4; q += nthr) {
    const float4 v = *reinterpret_cast<const float4*>(x + q * 4);
    for (int c = 0; c < world; ++c) st_peer_f4(me.AR[c] + (long long)rank * me.ar_cap + q * 4, v);
  }
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int prev = atomicAdd(me.done_counter, 1u);
    is_last = (prev == gridDim.x - 1);
  }
  __syncthreads();
  if (is_last) {
    if (threadIdx.x == 0) *me.done_counter = 0;
    if (threadIdx.x < (unsigned)world) {
      __threadfence_system();
      st_release_sys(me.flags[threadIdx.x] + GF_FED_PHASE_SMALL_AR * GF_MAX_RANKS + rank, epoch);
    }
  }
  wait_peers(me, world, GF_FED_PHASE_SMALL_AR, epoch);
  for (long long q = tid; q < n4; q += nthr) {
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int c = 0; c < world; ++c) {
      const float4 v = ld_peer_f4(me.AR[rank] + (long long)c * me.ar_cap + q * 4);
      acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
    }
    *reinterpret_cast<float4*>(x + q * 4) = acc;
  }
}

// Collects the fp32 master copy of beta: every word-tile range this rank does not own is read from its
// owner (the overlapped aggregation updates the master copy on the owner only).
__global__ void __launch_bounds__(256) fed_gather_beta_kernel(const GfFedArgs a) {
  const GfFabricRank& me = a.ranks[0];
  const int world = a.world, rank = me.rank;
  const int vt = (a.V + 127) / 128;
  const int per = (vt + world - 1) / world;
  const long long v4 = a.V / 4;  // V % 4 == 0 (checked by the launcher)
  const long long n = (long long)a.K * v4;
  for (long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x; q < n; q += (long long)gridDim.x * blockDim.x) {
    const long long k = q / v4, c4 = q - k * v4;
    int owner = (int)((c4 * 4) / 128) / per;
    if (owner >= world) owner = world - 1;
    if (owner == rank) continue;
    const long long ip = a.beta_off + k * a.V + c4 * 4;
    *reinterpret_cast<float4*>(me.P[rank] + ip) = ld_peer_f4(me.P[owner] + ip);
  }
}

}  // namespace gf

using namespace gf;

extern "C" int gf_fed_signal(const GfFedArgs* a, int phase, void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  GF_ARG_CHECK(a->world >= 1 && a->world <= GF_MAX_RANKS && a->n_rank_ctx == 1, "world / n_rank_ctx");
  GF_ARG_CHECK(phase >= 2 && phase < GF_FED_FLAG_SLOTS, "phase slots 0 and 1 belong to gf_fed_aggregate");
  fed_signal_kernel<<<1, 32, 0, st>>>(*a, phase);
  GF_LAUNCHED();
  return GF_OK;
}

extern "C" int gf_fed_make_tmaps(const GfFedArgs* a, void* out_) {
  GF_ARG_CHECK(a != nullptr && out_ != nullptr && a->n_rank_ctx == 1, "args");
  GF_ARG_CHECK((reinterpret_cast<uintptr_t>(out_) & 63) == 0, "the tensor-map buffer must be 64-byte aligned");
  const GfFabricRank& me = a->ranks[0];
  GF_ARG_CHECK(me.ld_gb % 8 == 0 && me.ld_gb >= a->V, "ld_gb must be a multiple of 8 (16-byte rows) and >= V");
  FedTmaps* out = reinterpret_cast<FedTmaps*>(out_);
  memset(out, 0, sizeof(FedTmaps));
  PFN_encodeTiled fn = get_encode_fn();
  GF_ARG_CHECK(fn != nullptr, "cuTensorMapEncodeTiled not available");
  for (int c = 0; c < a->world; ++c) {
    if (me.Gb[c] == nullptr) continue;
    cuuint64_t dims[2] = {(cuuint64_t)a->V, (cuuint64_t)a->K};
    cuuint64_t strides[1] = {(cuuint64_t)me.ld_gb * 2};
    cuuint32_t box[2] = {128, (cuuint32_t)kFedRows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = fn(&out->m[c], CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, me.Gb[c], dims, strides, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
      fprintf(stderr, "[gfedntm] cuTensorMapEncodeTiled failed (%d) for the beta gradient of rank %d\n", (int)r, c);
      return GF_ERR_DRIVER;
    }
  }
  // this rank's fp32 master copy of beta and its two optimiser moments ([K, V] inside the flat buffers)
  GF_ARG_CHECK(a->V % 4 == 0 && a->beta_off % 4 == 0, "V and beta_off must be multiples of 4");
  float* loc[3] = {me.P[me.rank] + a->beta_off, me.M + a->beta_off, me.V2 + a->beta_off};
  for (int i = 0; i < 3; ++i) {
    cuuint64_t dims[2] = {(cuuint64_t)a->V, (cuuint64_t)a->K};
    cuuint64_t strides[1] = {(cuuint64_t)a->V * 4};
    cuuint32_t box[2] = {128, (cuuint32_t)kFedRows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = fn(&out->loc[i], CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, loc[i], dims, strides, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
      fprintf(stderr, "[gfedntm] cuTensorMapEncodeTiled failed (%d) for the local beta state %d\n", (int)r, i);
      return GF_ERR_DRIVER;
    }
  }
  return GF_OK;
}

extern "C" int gf_fed_allreduce_small(const GfFedArgs* a, float* x, long long n, void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  GF_ARG_CHECK(a->world >= 1 && a->world <= GF_MAX_RANKS && a->n_rank_ctx == 1, "world / n_rank_ctx");
  const GfFabricRank& me = a->ranks[0];
  GF_ARG_CHECK(n > 0 && n % 4 == 0 && n <= me.ar_cap && me.AR[0] != nullptr && me.ar_cap % 4 == 0,
               "n must be a multiple of 4 and fit the staging area");
  GF_ARG_CHECK((reinterpret_cast<uintptr_t>(x) & 15) == 0, "x must be 16-byte aligned");
  if (a->world == 1) return GF_OK;
  long long blocks = (n / 4 + 255) / 256;
  if (blocks > 64) blocks = 64;  // every block waits on the peers' flags: all of them must be co-resident
  fed_allreduce_small_kernel<<<(int)blocks, 256, 0, st>>>(*a, x, n);
  GF_LAUNCHED();
  return GF_OK;
}

extern "C" int gf_fed_gather_beta(const GfFedArgs* a, void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  GF_ARG_CHECK(a->world >= 1 && a->world <= GF_MAX_RANKS && a->n_rank_ctx == 1, "world / n_rank_ctx");
  GF_ARG_CHECK(a->V % 4 == 0 && a->beta_off % 4 == 0, "V and beta_off must be multiples of 4");
  if (a->world == 1 || a->K == 0) return GF_OK;
  fed_gather_beta_kernel<<<4 * num_sms(), 256, 0, st>>>(*a);
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
  GF_ARG_CHECK(a->n_ranges >= 0 && a->n_ranges <= GF_FED_MAX_RANGES, "n_ranges");
  for (int i = 0; i < a->n_ranges; ++i)
    GF_ARG_CHECK(a->ranges[i].off % 4 == 0 && a->ranges[i].len % 4 == 0 && a->ranges[i].off >= 0 &&
                     a->ranges[i].off + a->ranges[i].len <= a->flat_n,
                 "ranges must be 16-byte aligned pieces of [0, flat_n)");
  AdamF h;
  h.lr = a->lr; h.b1 = a->b1; h.b2 = a->b2; h.eps = a->eps; h.wd = a->wd; h.solver = a->solver;
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
