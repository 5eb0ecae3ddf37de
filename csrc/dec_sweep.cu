// Decoder sweeps for ProdLDA-type models on sm_100a (tcgen05 + TMEM + TMA).
//
// The decoder of the reference (src/models/base/pytorchavitm/avitm_network/decoder_network.py:121-126,
// loss in avitm.py:203-229) is  softmax_V( BatchNorm_V( theta . beta ) )  followed by the BoW
// negative log-likelihood.  The [B,V] logits never exist in HBM here: every sweep recomputes a
// 128x128 logits tile with one short tcgen05.mma K-loop (K = n_components) into TMEM and consumes
// it in the epilogue warps.
//
// One kernel skeleton, two operand families:
//   ROW family: stationary A = 128 documents of theta_c [B,Kp]  (K-major),
//               streamed  B = 128-word tiles of beta [Kp,V]      (MN-major for MMA1).
//               TMEM lane = document  -> per-document reductions are thread-local.
//   COL family: stationary A = 128-word slab of beta [Kp,V]      (MN-major),
//               streamed  B = 128-document tiles of theta_c      (K-major for MMA1).
//               TMEM lane = word      -> per-word (BatchNorm) reductions are thread-local.
// Modes with a second MMA write the epilogue result P (bf16, 128x128) to shared memory and
// accumulate   D2 += P . (streamed tile)   in a persistent TMEM accumulator; the streamed tile's
// shared-memory image is reused by MMA2 with the other major-ness, so beta and theta_c each live
// in ONE global layout:
//   ROW_GRAD : D2[doc, topic]  = dLogit . beta^T      (d theta)
//   COL_GRAD : D2[word, topic] = dLogit^T . theta_c   (d beta, transposed)
//
// Warp roles (384 threads): warp 0 = TMA producer, warp 1 = MMA issuer (one thread, event driven),
// warp 2 = TMEM allocator, warps 4..11 = epilogue (warp w owns TMEM lanes 32*(w%4)..+31; warps 4..7 take
// the first 64 columns of every tile, warps 8..11 the second 64).
//
// The "folded" modes (ROW_LSE2, COL_GRAD2, ROW_GRAD2, COL_STATS2) carry every epilogue constant in the
// GEMM operands (csrc/fold_ops.cu) and need no side band; the earlier side-band modes are kept for small
// problems, where their shorter launch sequence wins.  With a federation context, the last CTAs of a
// ROW_GRAD2 grid run the beta part of the federated aggregation over NVLink instead of the sweep
// (csrc/fabric_dev.cuh).
#include <stdlib.h>
#include "ptx.cuh"
#include "tmap.h"
#include "../include/gfedntm_kernels.h"
#include "fabric_dev.cuh"

namespace gf {

enum : int {
  ROW_LOGITS = 0,  // out[b,v] = theta_c[b].beta[:,v]                      (debug / inference)
  ROW_LSE = 1,     // per (doc, chunk) online log-sum-exp partials of BN'd logits
  ROW_GRAD = 2,    // d theta (dense part)
  COL_STATS = 3,   // per-word second moment over the batch -> BatchNorm rstd, running stats
  COL_GRAD = 4,    // pass 1: BN-backward column sums; pass 2: d beta (dense part)
  COL_LOGITS = 5,  // out[b,v] via the COL family                          (debug)
  // "Folded" modes: every per-word and per-document constant of the epilogue is carried by the GEMM
  // itself (beta' = beta*rstd*log2e plus bias rows, theta_c^T plus constant-one rows and -e_b rows,
  // e_b = lse2_b - log2 S_b), so the accumulator IS the exponent and the epilogue needs no side band:
  ROW_LSE2 = 6,    // online (max2,sum) of 2^D1
  COL_GRAD2 = 7,   // ONE pass: P = 2^D1 (= S_b p_bv), D2 = P^T-free d beta accumulation, BN-backward
                   // constants from D2 in the drain, (C beta) correction from two pseudo-document tiles
  ROW_GRAD2 = 8,   // P = 2^D1, d theta += ln2 * P . beta'^T
  // Batch variance of the centred logits WITHOUT a pass over the batch:
  //   sum_b yc_bv^2 = beta_v^T C beta_v,  C = theta_c^T theta_c  (K x K, streamed as hi + lo bf16 tiles)
  COL_STATS2 = 9
};

constexpr int kTile = 128;
// (64-wide streamed tiles for the gradient modes were measured slower at V=200k, K=200: ROW_GRAD
// 1034 vs 826 us, COL_GRAD 1829 vs 1556 us -- per-tile hand-off overhead dominates.)
constexpr int kBN = 128;  // streamed-tile width
constexpr int kThreads = 384;     // warp 0 TMA, 1 MMA, 2 TMEM alloc, 3 idle, 4..11 epilogue
constexpr int kEpiThreads = 256;  // two epilogue warps per SM sub-partition: each takes half of a tile's columns
constexpr int kMaxStages = 6;
constexpr float kLog2e = 1.4426950408889634f;
constexpr float kLn2 = 0.6931471805599453f;

struct SweepParams {
  int mode;
  int B, V, K, Kp;
  int n_stat, n_stream, n_chunks, n_pass, stages;
  int p_bufs;               // 1 or 2 shared-memory P tiles (gradient modes)
  int n_ext, ext_tile0;     // COL_GRAD2: pseudo-document tiles (C = theta_c^T theta_c) streamed after the batch
                            // COL_STATS2: the streamed tiles are 2*n_ext C tiles (hi then lo) from ext_tile0
  const __nv_bfloat16* betaT_bf;  // COL_GRAD2: [V, Kp] unscaled beta^T (drain)
  __nv_bfloat16* beta3;     // COL_GRAD2 out: [Kp, ld3] rows k<K = gamma_v*beta[k,v], row K = alpha_v
  int ld3;
  const GfFedArgs* fed;     // ROW_GRAD2: device copy of the aggregation arguments (or null)
  int fed_ctas;             // number of trailing CTAs that run the beta aggregation instead of the sweep
  uint32_t smem_bytes;      // dynamic shared memory of the launch (the aggregation CTAs use all of it)
  int debug;                // profiling experiments: bit0 = epilogue skips TMEM loads/math, bit1 = skip MMA1 issue
  const float4* sb_stream;  // side band per streamed index (padded to a multiple of 128)
  // generic
  float* out;               // LOGITS modes: [B,V] fp32
  // ROW_LSE
  float2* lse_part;         // [B, n_chunks] (max in log2 units, sum)
  // ROW_GRAD
  const float2* doc_const;  // [B] {lse*log2e, S*ln2}
  float* dtheta;            // [B, Kp] fp32 (atomicAdd)
  // COL_*
  float4* colc;             // [Vpad] {rstd*log2e, bias, alpha, gamma}: written by COL_STATS / COL_GRAD
  const float* theta_mean;  // [K] batch mean of theta (COL_STATS, for running_mean)
  const __nv_bfloat16* beta_bf;  // [Kp, ldb] bf16 shadow (COL_STATS mean, epilogue reads)
  int ldb;
  float* running_mean;      // [V] (may be null)
  float* running_var;       // [V]
  float bn_eps, bn_momentum;
  int use_running;          // eval mode: normalise with running stats (COL_STATS only)
  const float* asp;         // [V] sparse part of sum_b dy
  const float* csp;         // [V] sparse part of sum_b dy*y
  float* dbeta;             // [K, ldg] fp32 gradient (COL_GRAD output)
  int ldg;
  __nv_bfloat16* dbeta_bf;  // COL_GRAD2: bf16 [K, ldgb] gradient instead of `dbeta` (federated exchange format)
  int ldgb;
  const float* corr;        // COL_GRAD: [V, Kp] sparse correction to d beta^T (may be null); COL_GRAD2: [Kp, ldc]
  int ldc;
  const int* colptr;        // [V+1] CSC pointer of the batch (with corr)
};

constexpr int kBarRegion = 3072;  // mbarriers (512 B) + the two 1 KB epilogue exchange arrays
constexpr int kSbStagesMax = 4;  // side-band ring depth (2 KB per entry); 2 in the gradient modes

__host__ __device__ constexpr bool mode_is_row(int m) {
  return m == ROW_LOGITS || m == ROW_LSE || m == ROW_GRAD || m == ROW_LSE2 || m == ROW_GRAD2;
}
__host__ __device__ constexpr bool mode_has_mma2(int m) { return m == ROW_GRAD || m == COL_GRAD || m == ROW_GRAD2 || m == COL_GRAD2; }
__host__ __device__ constexpr bool mode_has_sb(int m) { return m == ROW_LSE || m == ROW_GRAD || m == COL_GRAD; }
// side-band ring depth: 4 for ROW_LSE, 2 where the P tiles need the shared memory, none in the other modes
__host__ __device__ constexpr int mode_sb_stages(int m) { return mode_has_sb(m) ? (mode_has_mma2(m) ? 2 : kSbStagesMax) : 0; }

struct Smem {
  uint8_t* a;
  uint8_t* b0;         // operand stages: b0 + i * bstride   (address arithmetic, not a pointer table: a table
  uint32_t bstride;    //  indexed at run time lives in local memory and costs the single MMA thread a load per use)
  float4* sb0;         // side-band ring: sb0 + i * kBN
  uint8_t* p0;         // P tiles: p0 + i * 32 KB
  __device__ __forceinline__ uint8_t* b(int i) const { return b0 + (uint32_t)i * bstride; }
  __device__ __forceinline__ float4* sb(int i) const { return sb0 + i * kBN; }
  __device__ __forceinline__ uint8_t* p(int i) const { return p0 + (uint32_t)i * (kBN * 256u); }
  uint64_t* a_full;
  uint64_t* a_empty;
  uint64_t* b_full;    // [kMaxStages]  operand tile landed
  uint64_t* b_empty;   // [kMaxStages]  operand tile no longer read by any MMA
  uint64_t* sb_full;   // [kSbStages]   side band landed
  uint64_t* sb_empty;  // [kSbStages]   side band consumed by the epilogue
  uint64_t* d1_full;   // [2]
  uint64_t* d1_empty;  // [2]
  uint64_t* p_full;    // [2]
  uint64_t* p_empty;   // [2]
  uint64_t* d2_full;
  uint64_t* d2_empty;
  uint32_t* tmem_ptr;
  float2* xa;          // [128] epilogue half 1 -> half 0 partial accumulators
  float2* xb;          // [128] epilogue half 0 -> half 1 finalised per-word constants
};

// Both operands are stored "topics x contiguous axis" in global memory:
//   theta_T [Kp, B]  (documents contiguous)      beta [Kp, V]  (words contiguous)
// so a 128-wide tile of either is two TMA boxes of {64 columns x Kp rows} (Kp*128 bytes each,
// 128B-swizzled rows).  For MMA1 both tiles are MN-major operands (rows = contraction index);
// for MMA2 the streamed tile's very same bytes are a K-major B operand (rows = N index = topics).
template <int kMode, bool kFed = false>
__global__ void __launch_bounds__(kThreads, 1)
sweep_kernel(const __grid_constant__ CUtensorMap tm_theta, const __grid_constant__ CUtensorMap tm_beta,
             const __grid_constant__ FedTmaps fed_tm, const SweepParams p) {
  constexpr bool kRow = mode_is_row(kMode);
  constexpr bool kHasMMA2 = mode_has_mma2(kMode);
  constexpr bool kSb = mode_has_sb(kMode);
  constexpr bool kExt = (kMode == COL_GRAD2 || kMode == ROW_GRAD2);
  constexpr bool kEpiHoldsA = (kMode == COL_GRAD2);  // the drain reads the stationary tile from shared memory
  constexpr bool kFolded = (kMode == ROW_LSE2 || kMode == COL_GRAD2 || kMode == ROW_GRAD2);
  constexpr int kSbStages = mode_sb_stages(kMode);
  extern __shared__ uint8_t smem_raw[];
  uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  constexpr uint32_t kSbBytes = kBN * sizeof(float4);
  constexpr uint32_t kPBytes = kBN * 256u;          // P tile: 128 rows x kBN bf16
  const int stages = p.stages;
  const int Kp = p.Kp;
  const uint32_t tile_bytes = (uint32_t)Kp * 256u;  // stationary tile: 2 boxes of Kp rows x 128 B
  const uint32_t stile_bytes = (uint32_t)Kp * 2u * kBN;  // streamed tile: kBN/64 boxes

  Smem s;
  s.a = base;
  uint8_t* cur = base + tile_bytes;
  s.b0 = cur;
  s.bstride = stile_bytes;
  cur += (uint32_t)stages * stile_bytes;
  s.p0 = cur;
  if (kHasMMA2) cur += kPBytes * p.p_bufs;  // double-buffered when it fits: epilogue(t+1) overlaps MMA2(t)
  s.sb0 = reinterpret_cast<float4*>(cur);
  cur += kSbBytes * kSbStages;
  uint64_t* bars = reinterpret_cast<uint64_t*>(cur);
  s.a_full = bars + 0;
  s.a_empty = bars + 1;
  s.b_full = bars + 2;
  s.b_empty = s.b_full + kMaxStages;
  s.sb_full = s.b_empty + kMaxStages;
  s.sb_empty = s.sb_full + kSbStagesMax;
  s.d1_full = s.sb_empty + kSbStagesMax;
  s.d1_empty = s.d1_full + 2;
  s.p_full = s.d1_empty + 2;
  s.p_empty = s.p_full + 2;
  s.d2_full = s.p_empty + 2;
  s.d2_empty = s.d2_full + 1;
  s.tmem_ptr = reinterpret_cast<uint32_t*>(s.d2_empty + 1);
  s.xa = reinterpret_cast<float2*>(bars + 64);
  s.xb = s.xa + kTile;

  const uint32_t warp = threadIdx.x >> 5;
  const uint32_t lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tm_theta);
    tma_prefetch_desc(&tm_beta);
  }
  if (warp == 1 && lane == 0) {
    mbar_init(s.a_full, 1);
    mbar_init(s.a_empty, kEpiHoldsA ? 1 + kEpiThreads : 1);
    for (int i = 0; i < kMaxStages; ++i) {
      mbar_init(&s.b_full[i], 1);
      mbar_init(&s.b_empty[i], 1);  // tcgen05.commit of the last MMA that reads the tile
    }
    for (int i = 0; i < kSbStagesMax; ++i) {
      mbar_init(&s.sb_full[i], 1);
      mbar_init(&s.sb_empty[i], kEpiThreads);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&s.d1_full[i], 1);
      mbar_init(&s.d1_empty[i], kEpiThreads);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&s.p_full[i], kEpiThreads);
      mbar_init(&s.p_empty[i], 1);
    }
    mbar_init(s.d2_full, 1);
    mbar_init(s.d2_empty, kEpiThreads);
    mbar_fence_init();
  }
  constexpr uint32_t kTmemCols = kHasMMA2 ? 512 : 256;
  if (warp == 2) tmem_alloc<kTmemCols>(s.tmem_ptr);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *s.tmem_ptr;

  const int total_work = p.n_stat * p.n_chunks;
  const int n_ksteps = Kp >> 4;
  const CUtensorMap* tm_stat = kRow ? &tm_theta : &tm_beta;
  const CUtensorMap* tm_strm = kRow ? &tm_beta : &tm_theta;
  const uint32_t blk = (uint32_t)Kp * 128u;  // bytes of one 64-column block

  // ROW_GRAD2 with a federation context: the last fed_ctas CTAs of the grid overlap the beta part of the
  // federated aggregation (NVLink peer loads / stores) with the sweep the other CTAs are running.
  int n_sweep = (int)gridDim.x;
  bool agg_role = false;
  if (kFed) {
    n_sweep = (int)gridDim.x - p.fed_ctas;
    agg_role = (int)blockIdx.x >= n_sweep;
  }

  if (kFed && agg_role) {
    // (runs the aggregation after the role dispatch below, together with the sweep CTAs that are done)
  } else if (warp == 0) {
    // ============================ TMA producer ============================
    if (lane == 0) {
      uint32_t a_it = 0, b_it = 0, sb_it = 0;
      for (int w = blockIdx.x; w < total_work; w += n_sweep) {
        const int stat = w / p.n_chunks, chunk = w - stat * p.n_chunks;
        const int t0 = (int)(((long long)chunk * p.n_stream) / p.n_chunks);
        const int t1 = (int)(((long long)(chunk + 1) * p.n_stream) / p.n_chunks);
        mbar_wait(s.a_empty, (a_it & 1) ^ 1);
        mbar_arrive_expect_tx(s.a_full, tile_bytes);
        tma_load_2d(s.a, tm_stat, s.a_full, stat * kTile, 0);
        tma_load_2d(s.a + blk, tm_stat, s.a_full, stat * kTile + 64, 0);
        ++a_it;
        for (int pass = 0; pass < p.n_pass; ++pass) {
          for (int t = t0; t < t1; ++t) {
            const int st = b_it % stages;
            mbar_wait(&s.b_empty[st], ((b_it / stages) & 1) ^ 1);
            mbar_arrive_expect_tx(&s.b_full[st], stile_bytes);
            const int tc = (kMode == COL_STATS2) ? (p.ext_tile0 + t) : t;
            tma_load_2d(s.b(st), tm_strm, &s.b_full[st], tc * kBN, 0);
            tma_load_2d(s.b(st) + blk, tm_strm, &s.b_full[st], tc * kBN + 64, 0);
            if (kSb) {
              // side band: 128 float4 = 2 KB, plain bulk copy with its own full/empty pair (its ring
              // slot is only reused after the epilogue released it, so the parity can never alias)
              const int sbs = sb_it % (kSb ? kSbStages : 1);
              mbar_wait(&s.sb_empty[sbs], ((sb_it / (kSb ? kSbStages : 1)) & 1) ^ 1);
              mbar_arrive_expect_tx(&s.sb_full[sbs], kSbBytes);
              asm volatile(
                  "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n" ::
                      "r"(smem_u32(s.sb(sbs))),
                  "l"(reinterpret_cast<uint64_t>(p.sb_stream + (size_t)t * kBN)), "r"(kSbBytes),
                  "r"(smem_u32(&s.sb_full[sbs]))
                  : "memory");
              ++sb_it;
            }
            ++b_it;
          }
          if (kExt && pass == p.n_pass - 1) {
            // COL_GRAD2: pseudo-document tiles (columns of C = theta_c^T theta_c, appended to theta_c^T);
            // ROW_GRAD2: pseudo-word tiles (M hi/lo, appended to beta'), first vocabulary chunk only
            const int n_x = (kMode == ROW_GRAD2 && chunk != 0) ? 0 : p.n_ext;
            for (int i = 0; i < n_x; ++i) {
              const int st = b_it % stages;
              mbar_wait(&s.b_empty[st], ((b_it / stages) & 1) ^ 1);
              mbar_arrive_expect_tx(&s.b_full[st], stile_bytes);
              tma_load_2d(s.b(st), tm_strm, &s.b_full[st], (p.ext_tile0 + i) * kBN, 0);
              tma_load_2d(s.b(st) + blk, tm_strm, &s.b_full[st], (p.ext_tile0 + i) * kBN + 64, 0);
              ++b_it;
            }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ============================ MMA issuer ============================
    if (lane == 0) {
      const uint32_t idesc1 = umma_idesc_bf16(128, kBN, 1u, 1u);
      const uint32_t idesc2 = umma_idesc_bf16(128, (uint32_t)Kp, 0u, 0u);
      const uint32_t a_base = smem_u32(s.a);
      uint32_t a_it = 0, b_it = 0, d1_it = 0, p_it = 0, d2_it = 0;

      auto issue_mma2 = [&](int st, bool first) {
        const uint32_t pb = (p.p_bufs == 2) ? (p_it & 1) : 0;  // caller observed p_full[pb]
        tc_fence_after();
        const uint32_t p_base = smem_u32(s.p(pb));
        const uint32_t b_base = smem_u32(s.b(st));
#pragma unroll
        for (int j = 0; j < kBN / 16; ++j) {  // contraction over the 128 streamed columns, 16 per step
          const uint64_t da = umma_smem_desc(p_base + (j >> 2) * 16384u + (j & 3) * 32u, 16, 1024);
          const uint64_t db = umma_smem_desc(b_base + (j >> 2) * blk + (j & 3) * 32u, 16, 1024);
          if (!(p.debug & 4)) umma_f16(tmem + 256, da, db, idesc2, (first && j == 0) ? 0u : 1u);
        }
        umma_commit(&s.p_empty[pb]);
        umma_commit(&s.b_empty[st]);
        ++p_it;
      };

      for (int w = blockIdx.x; w < total_work; w += n_sweep) {
        const int stat = w / p.n_chunks, chunk = w - stat * p.n_chunks;
        const int t0 = (int)(((long long)chunk * p.n_stream) / p.n_chunks);
        const int t1 = (int)(((long long)(chunk + 1) * p.n_stream) / p.n_chunks);
        mbar_wait(s.a_full, a_it & 1);
        for (int pass = 0; pass < p.n_pass; ++pass) {
          const bool m2 = kHasMMA2 && (pass == p.n_pass - 1) && !(p.debug & 8);
          // ---- event-driven issue.  Two independent streams of work share this thread: MMA1 of the
          // next tile (needs its operands landed and a free accumulator buffer) and, in a gradient
          // pass, MMA2 of the oldest tile whose epilogue has published P (which releases that tile's
          // stage).  Polling both instead of a fixed order lets MMA1 run ahead of the epilogue ->
          // MMA2 chain, so no iteration waits for a full epilogue round trip.  (One MMA1 site and one
          // MMA2 site on purpose: with several inlined copies of the predicated k-step sequence the
          // gradient sweeps returned wrong results for Kp > 64.)
          if (m2) mbar_wait(s.d2_empty, (d2_it & 1) ^ 1);
          const uint32_t b_it0 = b_it;
          const int t1x = t1 + ((kExt && m2 && !(kMode == ROW_GRAD2 && chunk != 0)) ? p.n_ext : 0);  // pseudo tiles: MMA1 only
          int n1 = t0, n2 = m2 ? t0 : t1;
          while (n1 < t1x || n2 < t1) {
            if (kHasMMA2 && n2 < n1) {
              const uint32_t pb = (p.p_bufs == 2) ? (p_it & 1) : 0;
              if (mbar_test_wait(&s.p_full[pb], ((p.p_bufs == 2) ? (p_it >> 1) : p_it) & 1)) {
                issue_mma2((int)((b_it0 + (uint32_t)(n2 - t0)) % stages), n2 == t0);
                ++n2;
              }
            }
            if (n1 < t1x) {
              const uint32_t bi = b_it0 + (uint32_t)(n1 - t0);
              const int st = bi % stages;
              const uint32_t buf = d1_it & 1;
              if (mbar_test_wait(&s.b_full[st], (bi / stages) & 1) &&
                  mbar_test_wait(&s.d1_empty[buf], ((d1_it >> 1) & 1) ^ 1)) {
                tc_fence_after();
                const uint32_t b_base = smem_u32(s.b(st));
                const uint32_t d_addr = tmem + buf * kBN;
#pragma unroll
                for (int j = 0; j < 15; ++j) {
                  if (j < n_ksteps && !(p.debug & 2))
                    umma_f16(d_addr, umma_smem_desc(a_base + j * 2048u, blk, 1024),
                             umma_smem_desc(b_base + j * 2048u, blk, 1024), idesc1, j > 0 ? 1u : 0u);
                }
                umma_commit(&s.d1_full[buf]);
                ++d1_it;
                if (!m2 || n1 >= t1) umma_commit(&s.b_empty[st]);  // no MMA2: free once MMA1 has read it
                ++n1;
              }
            }
          }
          b_it += (uint32_t)(t1x - t0);
          if (m2) {
            umma_commit(s.d2_full);
            ++d2_it;
          }
        }
        umma_commit(s.a_empty);
        ++a_it;
      }
      if (kFed) {
        // This CTA goes on to the aggregation and reuses its shared memory: wait for the asynchronous
        // arrivals of the commits nobody else waits for (the last operand stages, P tiles and stationary tile).
        const uint32_t nb = b_it < (uint32_t)stages ? b_it : (uint32_t)stages;
        for (uint32_t i = b_it - nb; i < b_it; ++i) mbar_wait(&s.b_empty[i % stages], (i / stages) & 1);
        const uint32_t np = p_it < (uint32_t)p.p_bufs ? p_it : (uint32_t)p.p_bufs;
        for (uint32_t j = p_it - np; j < p_it; ++j)
          mbar_wait(&s.p_empty[(p.p_bufs == 2) ? (j & 1) : 0], ((p.p_bufs == 2) ? (j >> 1) : j) & 1);
        if (a_it > 0) mbar_wait(s.a_empty, (a_it - 1) & 1);
      }
    }
  } else if (kMode == COL_GRAD2 && warp == 3) {
    // ============================ drain prefetch (otherwise idle warp) ============================
    // The drain of a word slab reads the slab's sparse correction (Kp rows x 512 bytes of a 166 MB array written
    // just before this kernel, mostly out of L2 by now) and a few per-word scalars, in dependent steps with the
    // tensor pipe idle.  Pull those lines into L2 while the slab's main loop runs.
    uint32_t pf_it = 0;
    for (int w = blockIdx.x; w < total_work; w += n_sweep, ++pf_it) {
      // pace with the producer: slab n is fetched when the drain of slab n-1 has ended (same wait as the TMA warp)
      mbar_wait(s.a_empty, (pf_it & 1) ^ 1);
      const int stat = w / p.n_chunks;
      const size_t w0 = (size_t)stat * kTile;
      if (p.corr != nullptr) {
        for (int i = (int)lane; i < p.K * 4; i += 32) {
          const size_t col = w0 + (size_t)(i & 3) * 32;
          if (col < (size_t)p.V)
            asm volatile("prefetch.global.L2 [%0];" ::"l"(p.corr + (size_t)(i >> 2) * p.ldc + col));
        }
      }
      if (lane < 4 && w0 + lane * 32 < (size_t)p.V) {
        asm volatile("prefetch.global.L2 [%0];" ::"l"(p.asp + w0 + lane * 32));
        asm volatile("prefetch.global.L2 [%0];" ::"l"(p.csp + w0 + lane * 32));
        asm volatile("prefetch.global.L2 [%0];" ::"l"(p.colc + w0 + lane * 32));
        if (p.colptr != nullptr) asm volatile("prefetch.global.L2 [%0];" ::"l"(p.colptr + w0 + lane * 32));
      }
    }
  } else if (warp >= 4) {
    // ============================ epilogue ============================
    // Two warps per TMEM lane quarter (a warp may only touch lanes 32*(warp%4)..+31): warps 4..7 take
    // the first half of every tile's columns, warps 8..11 the second half.  One epilogue warp per
    // sub-partition was latency bound (MUFU / TMEM-load / LDS chains with nothing to overlap).
    const uint32_t q = warp & 3;
    const uint32_t h = (warp - 4) >> 2;
    const uint32_t row = q * 32 + lane;  // TMEM lane == tile row
    auto epi_sync = [] { asm volatile("bar.sync 1, %0;" ::"n"(kEpiThreads) : "memory"); };
    const uint32_t lane_addr = tmem + ((q * 32) << 16);
    uint32_t b_it = 0, d1_it = 0, p_it = 0, d2_it = 0;

    for (int w = blockIdx.x; w < total_work; w += n_sweep) {
      const int stat = w / p.n_chunks, chunk = w - stat * p.n_chunks;
      const int t0 = (int)(((long long)chunk * p.n_stream) / p.n_chunks);
      const int t1 = (int)(((long long)(chunk + 1) * p.n_stream) / p.n_chunks);
      const int grow = stat * kTile + (int)row;  // global doc (ROW) or word (COL) index

      // ---- per-work-item thread state
      float acc0 = 0.f, acc1 = 0.f;   // ROW_LSE: (max2, sum); COL: (a, c') / sumsq
      float c_r2 = 0.f, c_alpha = 0.f, c_gamma = 0.f, c_rstd = 0.f, c_bias = 0.f;
      float d_lse2 = 0.f, d_sp = 0.f;
      if (kMode == ROW_LSE || kMode == ROW_LSE2) acc0 = -INFINITY;
      if (kMode == ROW_GRAD && grow < p.B) {
        const float2 dc = p.doc_const[grow];
        d_lse2 = dc.x;
        d_sp = dc.y;
      }
      if ((kMode == COL_GRAD || kMode == COL_GRAD2) && grow < p.V) {
        const float4 cc = p.colc[grow];
        c_r2 = cc.x;
        c_bias = cc.y;
        c_rstd = cc.x * kLn2;
      }

      for (int pass = 0; pass < p.n_pass; ++pass) {
        const bool m2 = kHasMMA2 && (pass == p.n_pass - 1) && !(p.debug & 8);
        for (int t = t0; t < t1; ++t) {
          const int sbs = kSb ? (int)(b_it % (kSb ? kSbStages : 1)) : 0;
          const uint32_t buf = d1_it & 1;
          if (kSb) mbar_wait(&s.sb_full[sbs], (b_it / (kSb ? kSbStages : 1)) & 1);
          mbar_wait(&s.d1_full[buf], (d1_it >> 1) & 1);
          tc_fence_after();
          const uint32_t pb = (p.p_bufs == 2) ? (p_it & 1) : 0;
          const uint32_t pbuf = smem_u32(s.p(pb));
          if (m2) mbar_wait(&s.p_empty[pb], (((p.p_bufs == 2) ? (p_it >> 1) : p_it) & 1) ^ 1);
          const uint32_t sb = smem_u32(s.sb(sbs));  // shared-space address of this tile's side band
          const uint32_t taddr = lane_addr + buf * kBN;

          // consume one 32-column chunk `c` of the accumulator held in v[]
          auto process = [&](const int c, float (&v)[32]) {
            const int col0 = t * kBN + c * 32;  // global streamed index of v[0]
            if (kMode == ROW_LOGITS) {
              if (grow < p.B) {
                float* o = p.out + (size_t)grow * p.V + col0;
#pragma unroll
                for (int i = 0; i < 32; ++i)
                  if (col0 + i < p.V) o[i] = v[i];
              }
            } else if (kMode == COL_LOGITS) {
              if (grow < p.V) {
#pragma unroll
                for (int i = 0; i < 32; ++i)
                  if (col0 + i < p.B) p.out[(size_t)(col0 + i) * p.V + grow] = v[i];
              }
            } else if (kMode == ROW_LSE) {
              float cmax = -INFINITY;
#pragma unroll
              for (int i = 0; i < 32; ++i) {
                const float2 cc = lds_f2(sb + (c * 32 + i) * 16);
                v[i] = fmaf(v[i], cc.x, cc.y);
                cmax = fmaxf(cmax, v[i]);
              }
              const float nm = fmaxf(acc0, cmax);
              if (nm > -INFINITY) {
                float sum = acc1 * ex2_approx(acc0 - nm);
#pragma unroll
                for (int i = 0; i < 32; ++i) sum += ex2_approx(v[i] - nm);
                acc1 = sum;
                acc0 = nm;
              }
            } else if (kMode == ROW_GRAD) {
              // dl = rstd*S*p - alpha - gamma*yc ; p = 2^(yc*r2 + bias - lse2)
              uint32_t pk[16];
#pragma unroll
              for (int i = 0; i < 32; i += 2) {
                const float4 c0 = lds_f4(sb + (c * 32 + i) * 16);
                const float4 c1 = lds_f4(sb + (c * 32 + i + 1) * 16);
                const float p0 = ex2_approx(fmaf(v[i], c0.x, c0.y) - d_lse2);
                const float p1 = ex2_approx(fmaf(v[i + 1], c1.x, c1.y) - d_lse2);
                const float g0 = fmaf(-c0.w, v[i], fmaf(c0.x * d_sp, p0, -c0.z));
                const float g1 = fmaf(-c1.w, v[i + 1], fmaf(c1.x * d_sp, p1, -c1.z));
                pk[i >> 1] = pack_bf16(g0, g1);
              }
              // P tile, K-major SW128: chunk = (c>>1), slots (c&1)*4 .. +3 of row `row`
              const uint32_t prow = pbuf + (c >> 1) * 16384 + row * 128;
#pragma unroll
              for (int g = 0; g < 4; ++g) {
                const uint32_t slot = ((c & 1) * 4 + g) ^ (row & 7);
                sts_u4(prow + slot * 16, pk[g * 4], pk[g * 4 + 1], pk[g * 4 + 2], pk[g * 4 + 3]);
              }
            } else if (kMode == ROW_LSE2) {
              // the accumulator already is the exponent (log2 units): online max / sum only
              float cmax = v[0];
#pragma unroll
              for (int i = 1; i < 32; ++i) cmax = fmaxf(cmax, v[i]);
              const float nm = fmaxf(acc0, cmax);
              float sum = acc1 * ex2_approx(acc0 - nm);
#pragma unroll
              for (int i = 0; i < 32; ++i) sum += ex2_approx(v[i] - nm);
              acc1 = sum;
              acc0 = nm;
            } else if (kMode == COL_GRAD2 || kMode == ROW_GRAD2) {
              // P = 2^D1 = S_b * p_bv  (D1 = BN'd logit in log2 units minus e_b = lse2_b - log2 S_b)
              uint32_t pk[16];
#pragma unroll
              for (int i = 0; i < 32; i += 2) {
                const float p0 = ex2_approx(v[i]);
                const float p1 = ex2_approx(v[i + 1]);
                if (kMode == COL_GRAD2) acc0 += p0 + p1;  // sum_b dy (dense part)
                pk[i >> 1] = pack_bf16(p0, p1);
              }
              const uint32_t prow = pbuf + (c >> 1) * 16384 + row * 128;
#pragma unroll
              for (int g = 0; g < 4; ++g) {
                const uint32_t slot = ((c & 1) * 4 + g) ^ (row & 7);
                sts_u4(prow + slot * 16, pk[g * 4], pk[g * 4 + 1], pk[g * 4 + 2], pk[g * 4 + 3]);
              }
            } else if (kMode == COL_STATS2) {
              // v[i] = (C beta)[k, word] for k = k0 + i; accumulate beta[k,word] * v[i]
              const int k0 = (t % p.n_ext) * kBN + c * 32;
              if (grow < p.V) {
                const __nv_bfloat16* brow = p.betaT_bf + (size_t)grow * Kp;
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                  if (k0 + u * 8 < Kp) {
                    const uint4 q = __ldg(reinterpret_cast<const uint4*>(brow + k0 + u * 8));
                    const __nv_bfloat162* h2 = reinterpret_cast<const __nv_bfloat162*>(&q);
#pragma unroll
                    for (int w2 = 0; w2 < 4; ++w2) {
                      const float2 f = __bfloat1622float2(h2[w2]);
                      acc0 = fmaf(f.x, v[u * 8 + 2 * w2], acc0);
                      acc0 = fmaf(f.y, v[u * 8 + 2 * w2 + 1], acc0);
                    }
                  }
                }
              }
            } else if (kMode == COL_STATS) {
#pragma unroll
              for (int i = 0; i < 32; ++i) acc0 = fmaf(v[i], v[i], acc0);
            } else if (kMode == COL_GRAD) {
              if (!m2) {
                // pass 1: a += S*p ; c' += S*p*yc
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                  const float2 dc = lds_f2(sb + (c * 32 + i) * 16);  // {lse2_b, S_b, -, -}
                  const float d = dc.y * ex2_approx(fmaf(v[i], c_r2, c_bias - dc.x));
                  acc0 += d;
                  acc1 = fmaf(d, v[i], acc1);
                }
              } else {
                uint32_t pk[16];
#pragma unroll
                for (int i = 0; i < 32; i += 2) {
                  const float2 d0 = lds_f2(sb + (c * 32 + i) * 16);
                  const float2 d1 = lds_f2(sb + (c * 32 + i + 1) * 16);
                  const float p0 = ex2_approx(fmaf(v[i], c_r2, c_bias - d0.x));
                  const float p1 = ex2_approx(fmaf(v[i + 1], c_r2, c_bias - d1.x));
                  const float g0 = fmaf(-c_gamma, v[i], fmaf(c_rstd * d0.y, p0, -c_alpha));
                  const float g1 = fmaf(-c_gamma, v[i + 1], fmaf(c_rstd * d1.y, p1, -c_alpha));
                  pk[i >> 1] = pack_bf16(g0, g1);
                }
                const uint32_t prow = pbuf + (c >> 1) * 16384 + row * 128;
#pragma unroll
                for (int g = 0; g < 4; ++g) {
                  const uint32_t slot = ((c & 1) * 4 + g) ^ (row & 7);
                  sts_u4(prow + slot * 16, pk[g * 4], pk[g * 4 + 1], pk[g * 4 + 2], pk[g * 4 + 3]);
                }
              }
            }
          };

          // TMEM -> registers two chunks ahead of their use
          float va[32], vb[32];
          if (p.debug & 1) {  // profiling experiment: no TMEM loads / math / P writes
            tc_fence_before();
            mbar_arrive(&s.d1_empty[buf]);
            if (kSb) mbar_arrive(&s.sb_empty[sbs]);
            ++d1_it;
            ++b_it;
            if (m2) {
              fence_proxy_async_smem();
              mbar_arrive(&s.p_full[pb]);
              ++p_it;
            }
            continue;
          }
          tmem_ld32(taddr + h * 64, va);
          tmem_ld32(taddr + h * 64 + 32, vb);
          tmem_ld_wait();
          tc_fence_before();
          mbar_arrive(&s.d1_empty[buf]);  // accumulator buffer can be overwritten by the next MMA1
          process(2 * h, va);
          process(2 * h + 1, vb);
          if (kSb) mbar_arrive(&s.sb_empty[sbs]);  // side band of this tile fully consumed
          ++d1_it;
          ++b_it;
          if (m2) {
            fence_proxy_async_smem();
            mbar_arrive(&s.p_full[pb]);
            ++p_it;
          }
        }

        if (kMode == ROW_GRAD2 && m2 && chunk == 0) {
          // pseudo-word tiles: D1[doc, k] = a_k + (theta_c M)[doc,k] (hi tiles then lo tiles); subtract
          for (int x = 0; x < p.n_ext; ++x) {
            const uint32_t buf = d1_it & 1;
            mbar_wait(&s.d1_full[buf], (d1_it >> 1) & 1);
            tc_fence_after();
            float va[32], vb[32];
            tmem_ld32(lane_addr + buf * kBN + h * 64, va);
            tmem_ld32(lane_addr + buf * kBN + h * 64 + 32, vb);
            tmem_ld_wait();
            tc_fence_before();
            mbar_arrive(&s.d1_empty[buf]);
            ++d1_it;
            const int k0 = (x % ((Kp + kBN - 1) / kBN)) * kBN + (int)h * 64;
            if (grow < p.B) {
#pragma unroll
              for (int i = 0; i < 32; ++i) {
                if (k0 + i < p.K) atomicAdd(&p.dtheta[(size_t)grow * Kp + k0 + i], -va[i]);
                if (k0 + 32 + i < p.K) atomicAdd(&p.dtheta[(size_t)grow * Kp + k0 + 32 + i], -vb[i]);
              }
            }
          }
        }
        // ---- end of a pass
        if (kMode == COL_GRAD && !m2) {
          // finalise BN-backward column means, publish the per-word constants
          if (h == 1) s.xa[row] = make_float2(acc0, acc1);
          epi_sync();
          if (h == 0) {
            const float2 o = s.xa[row];
            acc0 += o.x;
            acc1 += o.y;
            if (grow < p.V) {
              const float invB = 1.0f / (float)p.B;
              const float abar = (acc0 - p.asp[grow]) * invB;
              const float cbar = (c_rstd * acc1 - p.csp[grow]) * invB;
              c_alpha = c_rstd * abar;
              c_gamma = c_rstd * c_rstd * cbar;
              p.colc[grow] = make_float4(c_r2, c_bias, c_alpha, c_gamma);
            }
            s.xb[row] = make_float2(c_alpha, c_gamma);
          }
          epi_sync();
          if (h == 1) {
            const float2 o = s.xb[row];
            c_alpha = o.x;
            c_gamma = o.y;
          }
        }
        if (m2) {
          mbar_wait(s.d2_full, d2_it & 1);
          tc_fence_after();
          const bool has_corr = (kMode == COL_GRAD || kMode == COL_GRAD2) && p.corr != nullptr && grow < p.V &&
                                (p.colptr[grow + 1] > p.colptr[grow]);
          if (kMode == COL_GRAD2) {
            // The two pseudo-document tiles (C = theta_c^T theta_c, bf16) are parked in the D1 buffers:
            // D1x[word, k] = sum_j beta'[j,word] C[j,k] = r2_word * (C beta)[k,word].
            const uint32_t d1_base = d1_it;
            for (int i = 0; i < p.n_ext; ++i) mbar_wait(&s.d1_full[(d1_base + i) & 1], ((d1_base + i) >> 1) & 1);
            tc_fence_after();
            const bool vw = grow < p.V;
            const int nchunk = (Kp + 31) >> 5;
            // beta'[k, word] of this thread's word from the stationary tile (MN-major, 128B swizzle):
            // box = word>>6, row k, 16-byte chunk ((word&63)>>3) ^ (k&7), element word&7
            const uint32_t a_word = smem_u32(s.a) + (row >> 6) * blk + (row & 7) * 2;
            const uint32_t a_chunk = (row & 63) >> 3;
            auto beta_s = [&](int k) -> float {
              uint16_t raw;
              asm("ld.shared.u16 %0, [%1];" : "=h"(raw) : "r"(a_word + (uint32_t)k * 128u + ((a_chunk ^ ((uint32_t)k & 7u)) << 4)));
              return __uint_as_float((uint32_t)raw << 16);
            };
            // ---- drain pass 1: r2 * c' = sum_k beta'[k,word] * D2[word,k]   (c' = sum_b dy*yc, dense part)
            float cp = 0.f;
            for (int c = (int)h; c < nchunk; c += 2) {
              float v[32];
              tmem_ld32(lane_addr + 256 + c * 32, v);
              tmem_ld_wait();
#pragma unroll
              for (int i = 0; i < 32; ++i)
                if (c * 32 + i < p.K) cp = fmaf(beta_s(c * 32 + i), v[i], cp);
            }
            const float inv_r2 = (c_r2 > 0.f) ? 1.0f / c_r2 : 0.f;
            cp *= inv_r2;
            if (h == 1) s.xa[row] = make_float2(acc0, cp);
            epi_sync();
            if (h == 0) {
              const float2 o = s.xa[row];
              acc0 += o.x;
              cp += o.y;
              if (vw) {
                const float invB = 1.0f / (float)p.B;
                const float abar = (acc0 - p.asp[grow]) * invB;
                const float cbar = (c_rstd * cp - p.csp[grow]) * invB;
                c_alpha = c_rstd * abar;
                c_gamma = c_rstd * c_rstd * cbar;
                p.colc[grow] = make_float4(c_r2, c_bias, c_alpha, c_gamma);
                if (p.beta3 != nullptr) p.beta3[(size_t)p.K * p.ld3 + grow] = __float2bfloat16(c_alpha);
              }
              s.xb[row] = make_float2(c_alpha, c_gamma);
            }
            epi_sync();
            if (h == 1) {
              const float2 o = s.xb[row];
              c_alpha = o.x;
              c_gamma = o.y;
            }
            // ---- drain pass 2: d beta[k,word] = rstd * D2 - gamma * (C beta)[k,word] - sparse correction
            const float g_cb = c_gamma * inv_r2;
            for (int c = (int)h; c < nchunk && !(p.debug & 16); c += 2) {
              float v[32], e[32];
              tmem_ld32(lane_addr + 256 + c * 32, v);
              tmem_ld32(lane_addr + ((d1_base + (uint32_t)(c >> 2)) & 1) * kBN + (c & 3) * 32, e);
              tmem_ld_wait();
              if (vw) {
                // all loads first (non-coherent path): a load behind a store to a possibly aliasing
                // pointer would otherwise serialise the chunk on memory latency
                float cr[32], bs[32];
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                  const int k = c * 32 + i;
                  cr[i] = (has_corr && k < p.K) ? __ldg(&p.corr[(size_t)k * p.ldc + grow]) : 0.f;
                  bs[i] = (k < p.K) ? beta_s(k) : 0.f;
                }
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                  const int k = c * 32 + i;
                  if (k < p.K) {
                    const float gval = fmaf(c_rstd, v[i], -g_cb * e[i]) - cr[i];
                    if (p.dbeta_bf != nullptr) p.dbeta_bf[(size_t)k * p.ldgb + grow] = __float2bfloat16(gval);
                    else p.dbeta[(size_t)k * p.ldg + grow] = gval;
                    if (p.beta3 != nullptr) p.beta3[(size_t)k * p.ld3 + grow] = __float2bfloat16(g_cb * bs[i]);
                  }
                }
              }
            }
            mbar_arrive(s.a_empty);  // the stationary tile may be overwritten now
            tc_fence_before();
            for (int i = 0; i < p.n_ext; ++i) mbar_arrive(&s.d1_empty[(d1_base + i) & 1]);
            d1_it += (uint32_t)p.n_ext;
          }
          for (int c = (int)h; c * 32 < Kp && !(p.debug & 16) && kMode != COL_GRAD2; c += 2) {  // halves interleave the chunks
            float v[32];
            tmem_ld32(lane_addr + 256 + c * 32, v);
            tmem_ld_wait();
            if (kMode == ROW_GRAD || kMode == ROW_GRAD2) {
              if (grow < p.B) {
                const float sc = (kMode == ROW_GRAD2) ? kLn2 : 1.0f;  // rstd = r2 * ln2 is folded into beta'
                const int klim = (kMode == ROW_GRAD2) ? p.K : Kp;
#pragma unroll
                for (int i = 0; i < 32; ++i)
                  if (c * 32 + i < klim) atomicAdd(&p.dtheta[(size_t)grow * Kp + c * 32 + i], sc * v[i]);
              }
            } else {  // COL_GRAD: d beta[k, word]
              if (grow < p.V) {
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                  const int k = c * 32 + i;
                  if (k < p.K) {
                    float g = v[i];
                    if (has_corr) g -= __ldg(&p.corr[(size_t)grow * Kp + k]);
                    p.dbeta[(size_t)k * p.ldg + grow] = g;
                  }
                }
              }
            }
          }
          tc_fence_before();
          mbar_arrive(s.d2_empty);
          ++d2_it;
        }
      }

      // ---- end of work item
      if (kMode == ROW_LSE || kMode == ROW_LSE2 || kMode == COL_STATS || kMode == COL_STATS2) {  // merge the two column halves of every row
        if (h == 1) s.xa[row] = make_float2(acc0, acc1);
        epi_sync();
        if (h == 0) {
          const float2 o = s.xa[row];
          if (kMode == ROW_LSE || kMode == ROW_LSE2) {
            const float nm = fmaxf(acc0, o.x);
            acc1 = (nm > -INFINITY) ? acc1 * ex2_approx(acc0 - nm) + o.y * ex2_approx(o.x - nm) : 0.f;
            acc0 = nm;
          } else {
            acc0 += o.x;
          }
        }
        epi_sync();  // xa may be rewritten by the next work item only after it has been read
      }
      if (kMode == ROW_LSE || kMode == ROW_LSE2) {
        if (h == 0 && grow < p.B) p.lse_part[(size_t)grow * p.n_chunks + chunk] = make_float2(acc0, acc1);
      } else if ((kMode == COL_STATS || kMode == COL_STATS2) && h == 0) {
        if (grow < p.V) {
          float mean = 0.f;
          if (p.theta_mean != nullptr) {
            for (int k = 0; k < p.K; ++k)
              mean = fmaf(p.theta_mean[k], __bfloat162float(p.beta_bf[(size_t)k * p.ldb + grow]), mean);
          }
          const float var = fmaxf(acc0, 0.f) / (float)p.B;  // biased batch variance of the centred logits
          float rstd, bias;
          if (p.use_running) {
            rstd = rsqrtf(p.running_var[grow] + p.bn_eps);
            bias = -p.running_mean[grow] * rstd * kLog2e;
          } else {
            rstd = rsqrtf(var + p.bn_eps);
            bias = 0.f;
            if (p.running_mean != nullptr) {
              const float mom = p.bn_momentum;
              const float unb = p.B > 1 ? var * (float)p.B / (float)(p.B - 1) : var;
              p.running_mean[grow] = (1.f - mom) * p.running_mean[grow] + mom * mean;
              p.running_var[grow] = (1.f - mom) * p.running_var[grow] + mom * unb;
            }
          }
          p.colc[grow] = make_float4(rstd * kLog2e, bias, 0.f, 0.f);
        } else if (grow < p.n_stat * kTile) {
          p.colc[grow] = make_float4(0.f, -INFINITY, 0.f, 0.f);  // padded words never contribute
        }
      }
    }
  }

  // ---- ROW_GRAD2 with a federation context: beta part of the aggregation.  The CTAs set aside for it come
  // straight here; a sweep CTA arrives once its share of the sweep is done (every MMA it issued has completed
  // and every tile it loaded has been consumed, so its shared memory is free) and takes the word tiles the
  // atomic counter still has.
  if constexpr (kFed) {
    tc_fence_before();
    __syncthreads();
    if (threadIdx.x == 0) {  // the sweep's mbarriers die here: the aggregation lays its own over the same bytes
      for (int i = 0; i < 32; ++i)  // a_full .. d2_empty are 32 consecutive 64-bit words
        asm volatile("mbarrier.inval.shared::cta.b64 [%0];" ::"r"(smem_u32(s.a_full + i)) : "memory");
    }
    const GfFedArgs& fa = *p.fed;
    const GfFabricRank& me = fa.ranks[0];
    const long long tstep = fa.step_ctr[1];
    if (threadIdx.x < (unsigned)fa.world) {  // every peer's beta gradient final, its shadows no longer read
      fed_wait_flag(me.flags[me.rank] + GF_FED_PHASE_BETA_READY * GF_MAX_RANKS + threadIdx.x, (uint32_t)tstep,
                    (int)threadIdx.x, GF_FED_PHASE_BETA_READY);
    }
    __syncthreads();
    asm volatile("fence.proxy.async;" ::: "memory");  // peer data is read, shared memory rewritten, by the async proxy
    const AdamF hh = fed_adam_consts(fa, tstep);
    fed_beta_phase_tma(fa, me, hh, fed_tm, base, p.smem_bytes - 1024);
    __threadfence_system();  // the completion handshake is phase C of the closing gf_fed_aggregate(parts=1)
  }

  // ---- teardown
  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc<kTmemCols>(tmem);
}

// Shared-memory plan of one mode (host side; mirrors the carve-up at the top of the kernel).
struct SmemPlan {
  int stages, p_bufs;
  size_t bytes;
};
static bool plan_smem(int mode, int Kp, int max_smem, SmemPlan* out) {
  const bool m2 = mode_has_mma2(mode);
  const size_t tile = (size_t)Kp * 256, stile = (size_t)Kp * 2 * kBN, ptile = (size_t)kBN * 256;
  const size_t base_fixed = 1024 + tile + (size_t)mode_sb_stages(mode) * kBN * sizeof(float4) + kBarRegion;
  int p_bufs = m2 ? 2 : 0;
  if (m2 && (size_t)max_smem < base_fixed + 2 * ptile + 2 * stile) p_bufs = 1;  // large Kp: single P tile
  const size_t fixed = base_fixed + ptile * (size_t)p_bufs;
  if ((size_t)max_smem < fixed + 2 * stile) return false;
  int stages = (int)(((size_t)max_smem - fixed) / stile);
  if (stages > kMaxStages) stages = kMaxStages;
  out->stages = stages;
  out->p_bufs = p_bufs;
  out->bytes = fixed + (size_t)stages * stile;
  return true;
}

template <int kMode>
static int launch_sweep(const CUtensorMap& tm_theta, const CUtensorMap& tm_beta, const FedTmaps& fed_tm, SweepParams& p,
                        cudaStream_t stream) {
  int max_smem = 0, dev = 0;
  GF_CUDA_CHECK(cudaGetDevice(&dev));
  GF_CUDA_CHECK(cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
  SmemPlan plan;
  GF_ARG_CHECK(plan_smem(kMode, p.Kp, max_smem, &plan),
               "n_components too large for the fused decoder sweeps (need Kp <= 240)");
  p.stages = plan.stages;
  p.p_bufs = plan.p_bufs;
  // aggregation CTAs stage (world-1) peer tiles of 16 KB per pipeline stage: give them the whole SM
  if (kMode == ROW_GRAD2 && p.fed != nullptr) plan.bytes = (size_t)max_smem;
  p.smem_bytes = (uint32_t)plan.bytes;
  // the aggregation role is a separate instantiation: its code in the same kernel costs the plain sweep ~25 %
  auto kern = (kMode == ROW_GRAD2 && p.fed != nullptr) ? sweep_kernel<kMode, kMode == ROW_GRAD2> : sweep_kernel<kMode, false>;
  GF_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)plan.bytes));
  const int total = p.n_stat * p.n_chunks;
  const int fed_ctas = (kMode == ROW_GRAD2 && p.fed != nullptr) ? p.fed_ctas : 0;
  int grid = num_sms() - fed_ctas;  // 12 warps with up to 168 registers each: one CTA per SM
  GF_ARG_CHECK(grid >= 1, "fed_ctas leaves no CTA for the sweep");
  if (grid > total) grid = total;
  if (grid < 1) grid = 1;
  grid += fed_ctas;
  kern<<<grid, kThreads, plan.bytes, stream>>>(tm_theta, tm_beta, fed_tm, p);
  GF_LAUNCHED();
  return GF_OK;
}

static int dispatch_sweep(const CUtensorMap& tm_theta, const CUtensorMap& tm_beta, const FedTmaps& fed_tm,
                          SweepParams& p, cudaStream_t stream) {
  switch (p.mode) {
    case ROW_LOGITS: return launch_sweep<ROW_LOGITS>(tm_theta, tm_beta, fed_tm, p, stream);
    case ROW_LSE: return launch_sweep<ROW_LSE>(tm_theta, tm_beta, fed_tm, p, stream);
    case ROW_GRAD: return launch_sweep<ROW_GRAD>(tm_theta, tm_beta, fed_tm, p, stream);
    case COL_STATS: return launch_sweep<COL_STATS>(tm_theta, tm_beta, fed_tm, p, stream);
    case COL_STATS2: return launch_sweep<COL_STATS2>(tm_theta, tm_beta, fed_tm, p, stream);
    case COL_GRAD: return launch_sweep<COL_GRAD>(tm_theta, tm_beta, fed_tm, p, stream);
    case COL_LOGITS: return launch_sweep<COL_LOGITS>(tm_theta, tm_beta, fed_tm, p, stream);
    case ROW_LSE2: return launch_sweep<ROW_LSE2>(tm_theta, tm_beta, fed_tm, p, stream);
    case COL_GRAD2: return launch_sweep<COL_GRAD2>(tm_theta, tm_beta, fed_tm, p, stream);
    case ROW_GRAD2: return launch_sweep<ROW_GRAD2>(tm_theta, tm_beta, fed_tm, p, stream);
    default: break;
  }
  fprintf(stderr, "[gfedntm] dec sweep: unsupported mode %d\n", p.mode);
  return GF_ERR_ARG;
}

}  // namespace gf

using namespace gf;

// Vocabulary split for the ROW modes: the smallest chunk count whose (document tiles x chunks) work
// items fill the persistent grid in whole waves (>= 95 % wave efficiency), so no SM idles in the tail.
extern "C" int gf_dec_row_chunks(int mode, int B, int V, int Kp) {
  (void)mode;
  (void)Kp;
  return gf_dec_row_chunks_grid(B, V, num_sms());
}

extern "C" int gf_num_sms(void) { return num_sms(); }

extern "C" int gf_dec_row_chunks_grid(int B, int V, int grid) {
  const int doc_tiles = (B + kTile - 1) / kTile, word_tiles = (V + kBN - 1) / kBN;
  if (grid < 1) grid = 1;
  int best = 1;
  double best_eff = 0.0;
  const int cmax = word_tiles < 4 * grid ? word_tiles : 4 * grid;
  for (int c = 1; c <= cmax; ++c) {
    const long long w = (long long)doc_tiles * c;
    const long long waves = (w + grid - 1) / grid;
    const double eff = (double)w / (double)(waves * grid);
    if (eff > best_eff + 1e-9) {
      best_eff = eff;
      best = c;
    }
    if (w >= grid && eff >= 0.95) return c;
  }
  return best;
}

extern "C" int gf_dec_sweep(const GfSweepArgs* a, void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  GF_ARG_CHECK(a != nullptr, "null args");
  GF_ARG_CHECK(a->Kp % 16 == 0 && a->Kp >= 16 && a->Kp <= 240, "Kp must be a multiple of 16 in [16,240]");
  GF_ARG_CHECK(a->K <= a->Kp && a->K > 0, "K");
  GF_ARG_CHECK(a->ld_theta % 8 == 0 && a->ld_beta % 8 == 0, "bf16 leading dims must be multiples of 8");
  GF_ARG_CHECK(a->mode >= 0 && a->mode <= 9, "mode");
  // Tensor extents along the contiguous axis: B / V by default (columns beyond read as zero); the
  // folded modes pass the padded widths so padded documents / words read their masking rows.
  const int theta_cols = a->theta_cols > 0 ? a->theta_cols : a->B;
  const int beta_cols = a->beta_cols > 0 ? a->beta_cols : a->V;
  GF_ARG_CHECK(a->ld_theta >= theta_cols && a->ld_beta >= beta_cols && theta_cols >= a->B && beta_cols >= a->V,
               "leading dims too small");
  CUtensorMap tm_theta, tm_beta;
  // theta_c^T [Kp, ld_theta] bf16: box = 64 documents x Kp topics
  int rc = make_tmap_bf16_2d(&tm_theta, a->theta_bf, (uint64_t)theta_cols, (uint64_t)a->Kp,
                             (uint64_t)a->ld_theta * 2, 64, (uint32_t)a->Kp);
  if (rc) return rc;
  // beta [Kp, ld_beta] bf16: box = 64 words x Kp topics
  rc = make_tmap_bf16_2d(&tm_beta, a->beta_bf, (uint64_t)beta_cols, (uint64_t)a->Kp,
                         (uint64_t)a->ld_beta * 2, 64, (uint32_t)a->Kp);
  if (rc) return rc;

  SweepParams p{};
  p.mode = a->mode;
  p.B = a->B; p.V = a->V; p.K = a->K; p.Kp = a->Kp;
  const bool row = mode_is_row(a->mode);
  const int bn = kBN;
  const int doc_tiles = (a->B + kTile - 1) / kTile, word_tiles = (a->V + kTile - 1) / kTile;
  p.n_stat = row ? doc_tiles : word_tiles;
  p.n_stream = row ? (a->V + bn - 1) / bn : (a->B + bn - 1) / bn;
  p.n_chunks = row ? a->n_chunks : 1;
  GF_ARG_CHECK(p.n_chunks >= 1 && p.n_chunks <= p.n_stream, "n_chunks");
  p.n_pass = (a->mode == COL_GRAD) ? 2 : 1;
  p.sb_stream = reinterpret_cast<const float4*>(a->sb_stream);
  GF_ARG_CHECK(p.sb_stream != nullptr || !mode_has_sb(a->mode), "sb_stream (padded side band) is required");
  if (a->mode == COL_STATS2) {
    p.n_ext = (a->Kp + kBN - 1) / kBN;
    GF_ARG_CHECK(a->ext_col0 % kBN == 0 && a->ext_col0 >= a->B && a->ext_col0 + 2 * p.n_ext * kBN <= theta_cols,
                 "ext_col0: the hi and lo C blocks must be tile aligned and inside theta_cols");
    p.ext_tile0 = a->ext_col0 / kBN;
    p.n_stream = 2 * p.n_ext;
    GF_ARG_CHECK(a->betaT_bf != nullptr, "betaT_bf");
    p.betaT_bf = reinterpret_cast<const __nv_bfloat16*>(a->betaT_bf);
  }
  if (a->mode == COL_GRAD2) {
    p.n_ext = (a->Kp + kBN - 1) / kBN;
    GF_ARG_CHECK(a->ext_col0 % kBN == 0 && a->ext_col0 >= a->B && a->ext_col0 + p.n_ext * kBN <= theta_cols,
                 "ext_col0: pseudo-document block must be tile aligned and inside theta_cols");
    p.ext_tile0 = a->ext_col0 / kBN;
    GF_ARG_CHECK(a->betaT_bf != nullptr, "betaT_bf");
    p.betaT_bf = reinterpret_cast<const __nv_bfloat16*>(a->betaT_bf);
    p.beta3 = reinterpret_cast<__nv_bfloat16*>(a->beta3);
    p.ld3 = a->ld_beta3;
    p.ldc = a->ld_corr;
    GF_ARG_CHECK(a->corr == nullptr || a->ld_corr >= a->V, "ld_corr");
  }
  if (a->mode == ROW_GRAD2) {
    // pseudo-word tiles: M hi then M lo (tile aligned) behind the padded vocabulary of beta'
    p.n_ext = 2 * ((a->Kp + kBN - 1) / kBN);
    GF_ARG_CHECK(a->ext_col0 % kBN == 0 && a->ext_col0 >= a->V && a->ext_col0 + p.n_ext * kBN <= beta_cols,
                 "ext_col0: the M hi/lo blocks must be tile aligned and inside beta_cols");
    p.ext_tile0 = a->ext_col0 / kBN;
    p.fed = reinterpret_cast<const GfFedArgs*>(a->fed);
    p.fed_ctas = a->fed != nullptr ? a->fed_ctas : 0;
    GF_ARG_CHECK(a->fed == nullptr || (a->fed_ctas >= 0 && a->fed_ctas < num_sms()), "fed_ctas");
    GF_ARG_CHECK(a->fed == nullptr || a->fed_tmaps != nullptr, "fed_tmaps (gf_fed_make_tmaps) is required with fed");
    GF_ARG_CHECK(a->fed == nullptr || (a->V % 8 == 0 && a->Kp % 16 == 0),
                 "the overlapped aggregation needs V % 8 == 0 (use gf_fed_aggregate(parts=3) otherwise)");
  }
  p.out = a->out;
  p.lse_part = reinterpret_cast<float2*>(a->lse_part);
  p.doc_const = reinterpret_cast<const float2*>(a->doc_const);
  p.dtheta = a->dtheta;
  p.colc = reinterpret_cast<float4*>(a->colc);
  p.theta_mean = a->theta_mean;
  p.beta_bf = reinterpret_cast<const __nv_bfloat16*>(a->beta_bf);
  p.ldb = a->ld_beta;
  p.running_mean = a->running_mean;
  p.running_var = a->running_var;
  p.bn_eps = a->bn_eps;
  p.bn_momentum = a->bn_momentum;
  p.use_running = a->use_running;
  p.asp = a->asp; p.csp = a->csp;
  p.dbeta = a->dbeta; p.ldg = a->ld_dbeta;
  p.dbeta_bf = reinterpret_cast<__nv_bfloat16*>(a->dbeta_bf); p.ldgb = a->ld_dbeta_bf;
  GF_ARG_CHECK(a->dbeta_bf == nullptr || a->mode == COL_GRAD2, "dbeta_bf is a COL_GRAD2 output");
  p.corr = a->corr; p.colptr = a->colptr;
  {
    const char* dbg = getenv("GFEDNTM_SWEEP_DEBUG");
    p.debug = dbg ? atoi(dbg) : 0;
  }
  static const FedTmaps kNoTmaps{};
  const FedTmaps* fed_tm = (a->mode == ROW_GRAD2 && a->fed != nullptr) ? reinterpret_cast<const FedTmaps*>(a->fed_tmaps)
                                                                      : &kNoTmaps;
  return dispatch_sweep(tm_theta, tm_beta, *fed_tm, p, stream);
}
