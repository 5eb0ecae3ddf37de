// Small CUDA-core kernels around the "folded" decoder sweeps (csrc/dec_sweep.cu, modes *2):
// they put every per-word / per-document constant of the ProdLDA decoder epilogue into the GEMM
// operands, so the tcgen05 accumulator is directly the exponent of the softmax term.
//
// Math (ref: avitm_network/decoder_network.py:121-126 and the autograd of avitm.py:225):
//   z_bv   = log2e * BN(theta_c . beta)_bv = sum_k theta_c[b,k] (beta[k,v] r2_v) + bias_v
//   P_bv   = 2^(z_bv - e_b) = S_b p_bv,  e_b = lse2_b - log2 S_b        (dense part of dL/dy)
//   d beta[k,v]  = rstd_v D2[v,k] - gamma_v (C beta)[k,v] - sparse part,  D2 = P^T theta_c, C = theta_c^T theta_c
//   d theta[b,k] = ln2 (P beta'^T)[b,k] - a_k - (theta_c M)[b,k] - sparse part,
//                  M = beta diag(gamma) beta^T,  a = beta alpha
// (the alpha term of d beta vanishes because theta_c is centred over the batch).
#include <cuda_bf16.h>
#include "ptx.cuh"
#include "tmap.h"
#include "../include/gfedntm_kernels.h"

namespace gf {

constexpr float kMaskExp = -30000.f;  // exponent of padded documents / words: 2^-30000 == 0

// theta_c (bf16) in two layouts, plus the constant rows of the transposed one.
__global__ void theta_center_fold_kernel(const float* __restrict__ theta, const float* __restrict__ colsum,
                                         __nv_bfloat16* __restrict__ out, __nv_bfloat16* __restrict__ outT,
                                         float* __restrict__ mean_out, int B, int K, int ld, int ldT, int Bpad,
                                         int center) {
  __shared__ float tile[32][33];
  const int b0 = blockIdx.x * 32, k0 = blockIdx.y * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;  // 32 x 8
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    const int b = b0 + ty + r * 8, k = k0 + tx;
    float v = 0.f;
    if (b < B && k < K) {
      const float m = center ? colsum[k] / (float)B : 0.f;
      v = theta[(size_t)b * K + k] - m;
      if (b == 0 && mean_out != nullptr) mean_out[k] = m;
    }
    if (b < B && k < ld) out[(size_t)b * ld + k] = __float2bfloat16(v);
    tile[ty + r * 8][tx] = v;
  }
  __syncthreads();
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    const int k = k0 + ty + r * 8, b = b0 + tx;
    if (k < ld && b < Bpad) {
      float v = tile[tx][ty + r * 8];             // zero for k >= K or b >= B
      if (k == K || k == K + 1) v = 1.f;          // multiplies the bias rows of beta'
      if (k == K + 2 && b >= B) v = kMaskExp;     // padded documents: P = 0
      outT[(size_t)k * ldT + b] = __float2bfloat16(v);
    }
  }
}

// beta' = scaled operand shadow (see header).  One thread per 8 consecutive words and a group of
// kScaleRows topic rows (the per-word constants are read once per group).
constexpr int kScaleRows = 32;
__global__ void __launch_bounds__(256) scale_beta_kernel(const __nv_bfloat16* __restrict__ beta, int ldb,
                                                         const float4* __restrict__ colc,
                                                         __nv_bfloat16* __restrict__ out, int ldo, int K, int Kp,
                                                         int V, int cols) {
  const int v0 = (blockIdx.x * blockDim.x + threadIdx.x) * 8;
  if (v0 >= cols) return;
  float r2[8], bias[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int v = v0 + i;
    float4 cc = make_float4(0.f, 0.f, 0.f, 0.f);
    if (v < V) cc = colc[v];
    r2[i] = cc.x;
    bias[i] = cc.y;
  }
  const int k_begin = blockIdx.y * kScaleRows, k_end = min(Kp, k_begin + kScaleRows);
#pragma unroll 4
  for (int k = k_begin; k < k_end; ++k) {
    __align__(16) __nv_bfloat16 o[8];
    if (k < K) {
      uint4 q = make_uint4(0, 0, 0, 0);
      if (v0 < V) q = *reinterpret_cast<const uint4*>(beta + (size_t)k * ldb + v0);  // ldb % 8 == 0: in-bounds
      const __nv_bfloat16* h = reinterpret_cast<const __nv_bfloat16*>(&q);
#pragma unroll
      for (int i = 0; i < 8; ++i) o[i] = __float2bfloat16((v0 + i < V) ? __bfloat162float(h[i]) * r2[i] : 0.f);
    } else {
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        float x = 0.f;
        if (v0 + i < V) {
          const __nv_bfloat16 hi = __float2bfloat16(bias[i]);
          if (k == K) x = __bfloat162float(hi);
          else if (k == K + 1) x = bias[i] - __bfloat162float(hi);
          else if (k == K + 2 || k == K + 3) x = 1.f;
        } else if (k == K) {
          x = kMaskExp;
        }
        o[i] = __float2bfloat16(x);
      }
    }
    *reinterpret_cast<uint4*>(out + (size_t)k * ldo + v0) = *reinterpret_cast<const uint4*>(o);
  }
}

// C[j,k] = sum_b theta_c[b,j] theta_c[b,k]: 32x32 output tile per block, the batch split over
// gridDim.z and combined with fp32 atomics in a zero-initialised scratch matrix.
__global__ void __launch_bounds__(256) theta_gram_kernel(const __nv_bfloat16* __restrict__ th, int ld, int B, int K,
                                                         float* __restrict__ cf) {
  __shared__ float As[64][33], Bs[64][33];
  const int j0 = blockIdx.y * 32, k0 = blockIdx.x * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;  // thread: k = k0+tx, j = j0 + ty*4 + r
  const int per = (B + gridDim.z - 1) / gridDim.z;
  const int b_begin = blockIdx.z * per, b_end = min(B, b_begin + per);
  float acc[4] = {0.f, 0.f, 0.f, 0.f};
  for (int b0 = b_begin; b0 < b_end; b0 += 64) {
#pragma unroll
    for (int r = 0; r < 8; ++r) {
      const int bb = ty * 8 + r, b = b0 + bb;
      As[bb][tx] = (b < b_end && j0 + tx < K) ? __bfloat162float(th[(size_t)b * ld + j0 + tx]) : 0.f;
      Bs[bb][tx] = (b < b_end && k0 + tx < K) ? __bfloat162float(th[(size_t)b * ld + k0 + tx]) : 0.f;
    }
    __syncthreads();
#pragma unroll 8
    for (int bb = 0; bb < 64; ++bb) {
      const float bk = Bs[bb][tx];
#pragma unroll
      for (int r = 0; r < 4; ++r) acc[r] = fmaf(As[bb][ty * 4 + r], bk, acc[r]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    const int j = j0 + ty * 4 + r, k = k0 + tx;
    if (j < K && k < K) atomicAdd(&cf[(size_t)j * K + k], acc[r]);
  }
}

// bf16 hi / lo split of C into the pseudo-document blocks of theta_c^T; re-zeroes the scratch.
__global__ void theta_gram_store_kernel(float* __restrict__ cf, int K, __nv_bfloat16* __restrict__ outT, int ldT,
                                        int col_hi, int col_lo) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= K * K) return;
  const int j = idx / K, k = idx - j * K;
  const float c = cf[idx];
  cf[idx] = 0.f;
  const __nv_bfloat16 hi = __float2bfloat16(c);
  outT[(size_t)j * ldT + col_hi + k] = hi;
  outT[(size_t)j * ldT + col_lo + k] = __float2bfloat16(c - __bfloat162float(hi));
}

// M[j,k] = sum_g partialT[g][k,j]  (gf_gram stores its partials transposed so its stores coalesce), rows
// j <= K, written as bf16 hi / lo into the pseudo-word blocks of beta'.
__global__ void gram_finish_kernel(const float* __restrict__ partial, int n_parts, int Kp, int K,
                                   __nv_bfloat16* __restrict__ out, int ldo, int col_hi, int col_lo) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (K + 1) * K) return;
  const int k = idx / (K + 1), j = idx - k * (K + 1);  // j fastest: coalesced reads of the transposed partials
  float acc = 0.f;
#pragma unroll 4
  for (int g = 0; g < n_parts; ++g) acc += partial[(size_t)g * Kp * Kp + (size_t)k * Kp + j];
  const __nv_bfloat16 hi = __float2bfloat16(acc);
  out[(size_t)j * ldo + col_hi + k] = hi;
  out[(size_t)j * ldo + col_lo + k] = __float2bfloat16(acc - __bfloat162float(hi));
}

}  // namespace gf

using namespace gf;

extern "C" int gf_theta_center_fold(const float* theta, const float* colsum, void* out_bf16, void* outT_bf16,
                                    float* mean_out, int B, int K, int ld, int ldT, int Bpad, int center,
                                    void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  GF_ARG_CHECK(ld >= K + 4, "ld must leave 4 spare rows above K");
  GF_ARG_CHECK(Bpad >= B && Bpad <= ldT, "Bpad");
  dim3 grid((Bpad + 31) / 32, (ld + 31) / 32);
  theta_center_fold_kernel<<<grid, 256, 0, st>>>(theta, colsum, reinterpret_cast<__nv_bfloat16*>(out_bf16),
                                                 reinterpret_cast<__nv_bfloat16*>(outT_bf16), mean_out, B, K, ld,
                                                 ldT, Bpad, center);
  GF_LAUNCHED();
  return GF_OK;
}

extern "C" int gf_scale_beta(const void* beta_bf, int ld_beta, const void* colc, void* out, int ld_out, int K,
                             int Kp, int V, int cols, void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  GF_ARG_CHECK(Kp >= K + 4 && cols % 8 == 0 && ld_out >= cols && ld_beta % 8 == 0 && ld_out % 8 == 0, "scale_beta dims");
  dim3 grid((cols / 8 + 255) / 256, (Kp + kScaleRows - 1) / kScaleRows);
  scale_beta_kernel<<<grid, 256, 0, st>>>(reinterpret_cast<const __nv_bfloat16*>(beta_bf), ld_beta,
                                          reinterpret_cast<const float4*>(colc),
                                          reinterpret_cast<__nv_bfloat16*>(out), ld_out, K, Kp, V, cols);
  GF_LAUNCHED();
  return GF_OK;
}

extern "C" int gf_theta_gram(const void* theta_bf, int ld, int B, int K, void* outT_bf16, int ldT, int ext_col0,
                             int lo_col0, float* c_scratch, void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  GF_ARG_CHECK(K <= 256 && c_scratch != nullptr, "K / scratch");
  int nz = (B + 255) / 256;
  if (nz > 32) nz = 32;
  dim3 grid((K + 31) / 32, (K + 31) / 32, nz);
  theta_gram_kernel<<<grid, 256, 0, st>>>(reinterpret_cast<const __nv_bfloat16*>(theta_bf), ld, B, K, c_scratch);
  GF_LAUNCHED();
  theta_gram_store_kernel<<<(K * K + 255) / 256, 256, 0, st>>>(c_scratch, K, reinterpret_cast<__nv_bfloat16*>(outT_bf16),
                                                               ldT, ext_col0, lo_col0);
  GF_LAUNCHED();
  return GF_OK;
}

extern "C" int gf_gram_finish(const float* partial, int n_parts, int K, int Kp, void* out_bf16, int ld_out,
                              int col_hi, int col_lo, void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  GF_ARG_CHECK(K < Kp && Kp <= 256, "K/Kp");
  const int n = (K + 1) * K;
  gram_finish_kernel<<<(n + 255) / 256, 256, 0, st>>>(partial, n_parts, Kp, K, reinterpret_cast<__nv_bfloat16*>(out_bf16),
                                                      ld_out, col_hi, col_lo);
  GF_LAUNCHED();
  return GF_OK;
}
