// Split-V Gram GEMM on tcgen05:  partial[g][k,j] = sum_{v in CTA g's words} x[j,v] * y[k,v]  (stored transposed).
//
// Used by the folded ProdLDA decoder backward (csrc/fold_ops.cu header): with x = beta3 (rows
// gamma_v*beta[j,v], row K = alpha_v, written by the COL_GRAD2 drain) and y = the unscaled bf16 beta
// shadow, the sum of the partials is M = beta diag(gamma) beta^T with a = beta.alpha in row K; the
// BatchNorm-backward part of d theta is then  -(a + theta_c M): gf_gram_finish stores M (hi/lo bf16) as
// pseudo-word tiles of beta', so the MMA1 of GF_ROW_GRAD2 produces that term.
// Ref math: autograd of decoder_network.py:123-124 (BatchNorm1d over the batch of theta.beta).
//
// Both operands are [Kp, V] bf16 with the vocabulary contiguous, i.e. K-major for a contraction
// over words: a tile is one TMA box of {64 words x Kp rows} (128-byte swizzled rows).  The
// accumulator is 2 x (128 x Kp) fp32 in TMEM (rows 0..127 and 128..255 of x).
#include "ptx.cuh"
#include "tmap.h"
#include "../include/gfedntm_kernels.h"

namespace gf {

constexpr int kGramThreads = 256;
constexpr int kGramStages = 3;
constexpr uint32_t kGramOpBytes = 256 * 128;  // one operand stage: 256 rows x 64 words bf16

__global__ void __launch_bounds__(kGramThreads, 1)
gram_kernel(const __grid_constant__ CUtensorMap tm_x, const __grid_constant__ CUtensorMap tm_y,
            float* __restrict__ partial, int n_tiles, int Kp) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* xs[kGramStages];
  uint8_t* ys[kGramStages];
  for (int i = 0; i < kGramStages; ++i) {
    xs[i] = base + (size_t)i * 2 * kGramOpBytes;
    ys[i] = xs[i] + kGramOpBytes;
  }
  uint64_t* bars = reinterpret_cast<uint64_t*>(base + (size_t)kGramStages * 2 * kGramOpBytes);
  uint64_t* full = bars;
  uint64_t* empty = bars + kGramStages;
  uint64_t* done = bars + 2 * kGramStages;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 2 * kGramStages + 1);

  const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tm_x);
    tma_prefetch_desc(&tm_y);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < kGramStages; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], 1);
    }
    mbar_init(done, 1);
    mbar_fence_init();
  }
  if (warp == 2) tmem_alloc<512>(tmem_ptr);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_ptr;

  const int g = blockIdx.x, G = gridDim.x;
  const int t0 = (int)(((long long)g * n_tiles) / G), t1 = (int)(((long long)(g + 1) * n_tiles) / G);
  const uint32_t op_bytes = (uint32_t)Kp * 128u;

  if (warp == 0) {
    if (lane == 0) {
      uint32_t it = 0;
      for (int t = t0; t < t1; ++t, ++it) {
        const int st = it % kGramStages;
        mbar_wait(&empty[st], ((it / kGramStages) & 1) ^ 1);
        mbar_arrive_expect_tx(&full[st], 2 * op_bytes);
        tma_load_2d(xs[st], &tm_x, &full[st], t * 64, 0);
        tma_load_2d(ys[st], &tm_y, &full[st], t * 64, 0);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      const uint32_t idesc = umma_idesc_bf16(128, (uint32_t)Kp, 0u, 0u);
      uint32_t it = 0;
      for (int t = t0; t < t1; ++t, ++it) {
        const int st = it % kGramStages;
        mbar_wait(&full[st], (it / kGramStages) & 1);
        tc_fence_after();
        const uint32_t xb = smem_u32(xs[st]), yb = smem_u32(ys[st]);
#pragma unroll
        for (int j = 0; j < 4; ++j) {  // 64 words = 4 k-steps of 16
          const uint64_t db = umma_smem_desc(yb + j * 32u, 16, 1024);
          umma_f16(tmem, umma_smem_desc(xb + j * 32u, 16, 1024), db, idesc, (t > t0 || j > 0) ? 1u : 0u);
          if (Kp > 128)
            umma_f16(tmem + 256, umma_smem_desc(xb + 16384u + j * 32u, 16, 1024), db, idesc,
                     (t > t0 || j > 0) ? 1u : 0u);
        }
        umma_commit(&empty[st]);
      }
      umma_commit(done);
    }
  } else if (warp >= 4) {
    const uint32_t q = warp - 4;
    mbar_wait(done, 0);
    tc_fence_after();
    float* out = partial + (size_t)g * Kp * Kp;
    for (int half = 0; half * 128 < Kp; ++half) {
      const int j = half * 128 + (int)(q * 32 + lane);
      for (int c = 0; c * 32 < Kp; ++c) {
        float v[32];
        tmem_ld32(tmem + ((q * 32) << 16) + half * 256 + c * 32, v);
        tmem_ld_wait();
        if (j < Kp) {  // transposed store: lanes (rows j) are contiguous -> coalesced
#pragma unroll
          for (int i = 0; i < 32; ++i)
            if (c * 32 + i < Kp) out[(size_t)(c * 32 + i) * Kp + j] = v[i];
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc<512>(tmem);
}

}  // namespace gf

using namespace gf;

extern "C" int gf_gram_parts(void) { return num_sms(); }

extern "C" int gf_gram(const void* x_bf, int ldx, const void* y_bf, int ldy, int V, int Kp, float* partial,
                       void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  GF_ARG_CHECK(Kp % 16 == 0 && Kp >= 16 && Kp <= 256, "Kp");
  GF_ARG_CHECK(ldx % 8 == 0 && ldy % 8 == 0 && ldx >= V && ldy >= V, "leading dims");
  CUtensorMap tm_x, tm_y;
  int rc = make_tmap_bf16_2d(&tm_x, x_bf, (uint64_t)V, (uint64_t)Kp, (uint64_t)ldx * 2, 64, (uint32_t)Kp);
  if (rc) return rc;
  rc = make_tmap_bf16_2d(&tm_y, y_bf, (uint64_t)V, (uint64_t)Kp, (uint64_t)ldy * 2, 64, (uint32_t)Kp);
  if (rc) return rc;
  const int n_tiles = (V + 63) / 64;
  // every CTA must own at least one tile (its accumulator is otherwise never written): the caller sums
  // gf_gram_parts() partials, so surplus slots are zero-filled instead of shrinking the grid
  const int parts = num_sms();
  int grid = parts < n_tiles ? parts : n_tiles;
  if (grid < parts)
    GF_CUDA_CHECK(cudaMemsetAsync(partial + (size_t)grid * Kp * Kp, 0, (size_t)(parts - grid) * Kp * Kp * sizeof(float),
                                  stream));
  const size_t smem = 1024 + (size_t)kGramStages * 2 * kGramOpBytes + 256;
  GF_CUDA_CHECK(cudaFuncSetAttribute(gram_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  gram_kernel<<<grid, kGramThreads, smem, stream>>>(tm_x, tm_y, partial, n_tiles, Kp);
  GF_LAUNCHED();
  return GF_OK;
}
