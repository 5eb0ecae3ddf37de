// Sparse (bag-of-words) side of the decoder and of the first encoder layer, the per-step
// CSR->CSC transpose of the minibatch, and the NeuralLDA (model_type='LDA') decoder which is
// sparse end to end.
//
// Ref: avitm_network/avitm.py:225 (RL = -sum x log(p+1e-10)), decoder_network.py:121-132,
// and the autograd of those ops.
#include "ptx.cuh"
#include "tmap.h"
#include "../include/gfedntm_kernels.h"

namespace gf {

constexpr float kLog2e_s = 1.4426950408889634f;
constexpr float kLn2_s = 0.6931471805599453f;
constexpr int kMaxKp = 256;

// ---------------------------------------------------------------------------------------------
// CSR -> CSC of one minibatch (counting sort by word id; order inside a column is arbitrary).
// ---------------------------------------------------------------------------------------------
__global__ void csc_hist_kernel(const int* __restrict__ indices, int nnz, int* __restrict__ count) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < nnz) atomicAdd(&count[indices[i] + 1], 1);
}
// In-place inclusive scan of data[0..n) (data[0] == 0 on entry => CSC column pointer), two kernels:
// per-block sums of 4096-element segments, then every block adds the sums of its predecessors to a
// local scan.  (A single-block version took 95 us at V = 200k.)
constexpr int kScanSeg = 4096;
__global__ void __launch_bounds__(1024) scan_sums_kernel(const int* __restrict__ data, int n, int* __restrict__ sums) {
  __shared__ int warp_tot[32];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int i0 = blockIdx.x * kScanSeg + tid * 4;
  int s = 0;
#pragma unroll
  for (int j = 0; j < 4; ++j) s += (i0 + j < n) ? data[i0 + j] : 0;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if (lane == 0) warp_tot[warp] = s;
  __syncthreads();
  if (warp == 0) {
    int w = warp_tot[lane];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) w += __shfl_xor_sync(0xffffffffu, w, o);
    if (lane == 0) sums[blockIdx.x] = w;
  }
}
__global__ void __launch_bounds__(1024) scan_apply_kernel(int* __restrict__ data, int n, const int* __restrict__ sums) {
  __shared__ int warp_tot[32];
  __shared__ int carry_s;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  // carry = sum of the preceding segments
  int c = 0;
  for (int b = tid; b < (int)blockIdx.x; b += 1024) c += sums[b];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
  if (lane == 0) warp_tot[warp] = c;
  __syncthreads();
  if (warp == 0) {
    int w = warp_tot[lane];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) w += __shfl_xor_sync(0xffffffffu, w, o);
    if (lane == 0) carry_s = w;
  }
  __syncthreads();
  const int carry = carry_s;
  __syncthreads();
  int v[4], s = 0;
  const int i0 = blockIdx.x * kScanSeg + tid * 4;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    v[j] = (i0 + j < n) ? data[i0 + j] : 0;
    s += v[j];
  }
  int inc = s;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int t = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += t;
  }
  if (lane == 31) warp_tot[warp] = inc;
  __syncthreads();
  if (warp == 0) {
    int w = warp_tot[lane];
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, w, o);
      if (lane >= o) w += t;
    }
    warp_tot[lane] = w;
  }
  __syncthreads();
  int run = carry + (warp > 0 ? warp_tot[warp - 1] : 0) + inc - s;  // exclusive prefix of this thread
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    run += v[j];
    if (i0 + j < n) data[i0 + j] = run;
  }
}
// scratch: >= ceil(n / 4096) ints, fully overwritten
static int launch_scan(int* data, int n, int* scratch, cudaStream_t st) {
  const int nb = (n + kScanSeg - 1) / kScanSeg;
  scan_sums_kernel<<<nb, 1024, 0, st>>>(data, n, scratch);
  GF_LAUNCHED();
  scan_apply_kernel<<<nb, 1024, 0, st>>>(data, n, scratch);
  GF_LAUNCHED();
  return GF_OK;
}
__global__ void csc_scatter_kernel(const int* __restrict__ indptr, const int* __restrict__ indices, int B,
                                   const int* __restrict__ colptr, int* __restrict__ cursor,
                                   int* __restrict__ csc_row, int* __restrict__ csc_src) {
  // one warp per document
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= B) return;
  const int e0 = indptr[warp], e1 = indptr[warp + 1];
  for (int e = e0 + lane; e < e1; e += 32) {
    const int v = indices[e];
    const int pos = colptr[v] + atomicAdd(&cursor[v], 1);
    csc_row[pos] = warp;
    csc_src[pos] = e;
  }
}

// ---------------------------------------------------------------------------------------------
// ProdLDA decoder, sparse part (one warp per document):
//   lse_b from the ROW_LSE partials, then for every non-zero (b,v,x):
//     y = BN'd logit, p = softmax prob, RL -= x log(p+1e-10), t = g*x*p/(p+1e-10)
//   S_b = sum t; asp[v] += t; csp[v] += t*y; tnnz[e] = rstd_v * t;
//   dtheta[b,:] = -sum rstd_v t beta[:,v]   (plain store: initialises the d theta buffer)
// ---------------------------------------------------------------------------------------------
// kWpd warps per document (8 warps per block, 8 / kWpd documents per block): the non-zeros of a document are
// dealt to its warps in rounds of 32, so a document's dependent gather chain is kWpd times shorter and the
// small-batch configurations (64 documents) still put a few hundred warps on the machine.  Partial RL / S /
// d theta of a document's warps are combined in shared memory (fixed order: deterministic).
template <int kWpd>
__global__ void __launch_bounds__(256) dec_sparse_doc_kernel(const GfDecSparseArgs a) {
  constexpr int kDpb = 8 / kWpd;
  const int wib = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int dslot = wib / kWpd, sub = wib % kWpd;
  const int b = blockIdx.x * kDpb + dslot;
  const bool live = b < a.B;
  const int Kp = a.Kp;
  const float2* lse_part = reinterpret_cast<const float2*>(a.lse_part);
  const __nv_bfloat16* theta_bf = reinterpret_cast<const __nv_bfloat16*>(a.theta_bf);
  const __nv_bfloat16* betaT_bf = reinterpret_cast<const __nv_bfloat16*>(a.betaT_bf);
  const float4* colc = reinterpret_cast<const float4*>(a.colc);
  float2* doc_const = reinterpret_cast<float2*>(a.doc_const);
  float4* sb_doc = reinterpret_cast<float4*>(a.sb_doc);
  __shared__ float th_s[kDpb][kMaxKp];
  __shared__ float dred[8][kMaxKp];
  __shared__ float2 sred[8];
  // combine log-sum-exp partials (every warp of the document: cheap, avoids a broadcast)
  float lse2 = 0.f;
  if (live) {
    float m = -INFINITY;
    for (int c = lane; c < a.n_chunks; c += 32) m = fmaxf(m, lse_part[(size_t)b * a.n_chunks + c].x);
    m = warp_max(m);
    float ssum = 0.f;
    for (int c = lane; c < a.n_chunks; c += 32) {
      const float2 pr = lse_part[(size_t)b * a.n_chunks + c];
      if (pr.x > -INFINITY) ssum += pr.y * exp2f(pr.x - m);
    }
    ssum = warp_sum(ssum);
    lse2 = m + log2f(ssum);
    // theta_c row of this document in shared memory (read by every lane in phase 1)
    for (int k = sub * 32 + lane; k < Kp; k += 32 * kWpd) th_s[dslot][k] = __bfloat162float(theta_bf[(size_t)b * a.ld_theta + k]);
  }
  __syncthreads();
  float dth[kMaxKp / 32];
#pragma unroll
  for (int c = 0; c < kMaxKp / 32; ++c) dth[c] = 0.f;
  float rl = 0.f, S = 0.f;
  if (live) {
    const int e0 = a.indptr[b], e1 = a.indptr[b + 1];
    for (int base = e0 + 32 * sub; base < e1; base += 32 * kWpd) {
      // ---- phase 1: one non-zero per lane (32 independent row gathers in flight)
      const int e = base + lane;
      const bool valid = e < e1;
      const int v = valid ? a.indices[e] : 0;
      const float x = valid ? a.counts[e] : 0.f;
      const uint4* br = reinterpret_cast<const uint4*>(betaT_bf + (size_t)v * Kp);
      float dot = 0.f;
#pragma unroll 4
      for (int k8 = 0; k8 < (Kp >> 3); ++k8) {
        const uint4 q = __ldg(br + k8);
        const __nv_bfloat162* h2 = reinterpret_cast<const __nv_bfloat162*>(&q);
        const float* t = &th_s[dslot][k8 * 8];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const float2 f = __bfloat1622float2(h2[u]);
          dot = fmaf(t[2 * u], f.x, dot);
          dot = fmaf(t[2 * u + 1], f.y, dot);
        }
      }
      const float4 cc = colc[v];
      const float t2 = fmaf(dot, cc.x, cc.y);
      const float p = exp2f(t2 - lse2);
      const float y = t2 * kLn2_s;
      const float rstd = cc.x * kLn2_s;
      float w = 0.f;
      if (valid) {
        rl -= x * __logf(p + 1e-10f);
        const float t = a.gscale * x * p / (p + 1e-10f);
        S += t;
        w = rstd * t;
        if (a.asp != nullptr) {
          atomicAdd(&a.asp[v], t);
          atomicAdd(&a.csp[v], t * y);
          a.tnnz[e] = w;
        }
      }
      // ---- phase 2: d theta[b,:] -= sum_e w_e * beta[:, v_e]   (lanes over topics, rows now cached)
      if (a.dtheta != nullptr) {
        const int n = min(32, e1 - base);
        for (int i = 0; i < n; ++i) {
          const int vi = __shfl_sync(0xffffffffu, v, i);
          const float wi = __shfl_sync(0xffffffffu, w, i);
          const __nv_bfloat16* bri = betaT_bf + (size_t)vi * Kp;
#pragma unroll
          for (int c = 0; c < kMaxKp / 32; ++c) {
            const int k = c * 32 + lane;
            if (k < Kp) dth[c] = fmaf(-wi, __bfloat162float(bri[k]), dth[c]);
          }
        }
      }
    }
  }
  rl = warp_sum(rl);
  S = warp_sum(S);
  if (kWpd > 1) {
    if (lane == 0) sred[wib] = make_float2(rl, S);
    if (a.dtheta != nullptr) {
#pragma unroll
      for (int c = 0; c < kMaxKp / 32; ++c) {
        const int k = c * 32 + lane;
        if (k < Kp) dred[wib][k] = dth[c];
      }
    }
    __syncthreads();
    if (sub == 0) {
      rl = 0.f; S = 0.f;
#pragma unroll
      for (int u = 0; u < kWpd; ++u) { rl += sred[wib + u].x; S += sred[wib + u].y; }
      if (a.dtheta != nullptr) {
#pragma unroll
        for (int c = 0; c < kMaxKp / 32; ++c) {
          const int k = c * 32 + lane;
          float acc = 0.f;
          if (k < Kp)
            for (int u = 0; u < kWpd; ++u) acc += dred[wib + u][k];
          dth[c] = acc;
        }
      }
    }
  }
  if (!live || sub != 0) return;
  if (a.dtheta != nullptr) {
#pragma unroll
    for (int c = 0; c < kMaxKp / 32; ++c) {
      const int k = c * 32 + lane;
      if (k < Kp) a.dtheta[(size_t)b * Kp + k] = dth[c];
    }
  }
  if (lane == 0) {
    a.rl[b] = rl;
    if (a.loss != nullptr) atomicAdd(a.loss, (double)rl);
    if (doc_const != nullptr) doc_const[b] = make_float2(lse2, S * kLn2_s);
    if (sb_doc != nullptr) sb_doc[b] = make_float4(lse2, S, 0.f, 0.f);
    if (a.thetaT_bf != nullptr) {
      // folded sweeps: 2^(z - e) = S_b * p_bv with e = lse2 - log2 S; -e as a hi/lo bf16 pair
      __nv_bfloat16* tT = reinterpret_cast<__nv_bfloat16*>(a.thetaT_bf);
      const float ne = (S > 0.f) ? (log2f(S) - lse2) : -30000.f;
      const __nv_bfloat16 hi = __float2bfloat16(ne);
      tT[(size_t)(a.K + 2) * a.ldT + b] = hi;
      tT[(size_t)(a.K + 3) * a.ldT + b] = __float2bfloat16(ne - __bfloat162float(hi));
    }
  }
}

// corr[v, :] = sum over the column's entries of tnnz[e] * theta_c[b, :]   (one warp per word)
__global__ void __launch_bounds__(256) csc_corr_kernel(const int* __restrict__ colptr, const int* __restrict__ csc_row,
                                                       const int* __restrict__ csc_src, const float* __restrict__ tnnz,
                                                       const __nv_bfloat16* __restrict__ theta_bf, int ld_theta,
                                                       float* __restrict__ corr, int V, int Kp) {
  const int v = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (v >= V) return;
  const int e0 = colptr[v], e1 = colptr[v + 1];
  if (e0 == e1) return;
  float acc[kMaxKp / 32];
#pragma unroll
  for (int c = 0; c < kMaxKp / 32; ++c) acc[c] = 0.f;
  for (int e = e0; e < e1; ++e) {
    const int b = csc_row[e];
    const float w = tnnz[csc_src[e]];
    const __nv_bfloat16* tr = theta_bf + (size_t)b * ld_theta;
#pragma unroll
    for (int c = 0; c < kMaxKp / 32; ++c) {
      const int k = c * 32 + lane;
      if (k < Kp) acc[c] = fmaf(w, __bfloat162float(tr[k]), acc[c]);
    }
  }
#pragma unroll
  for (int c = 0; c < kMaxKp / 32; ++c) {
    const int k = c * 32 + lane;
    if (k < Kp) corr[(size_t)v * Kp + k] = acc[c];
  }
}

// Same sums, written transposed (corrT[k, v]) so that a reader whose threads own consecutive words
// gets coalesced rows: a block covers 32 consecutive words (4 per warp), stages the [32 x Kp] block in
// shared memory and writes 128-byte rows.  Words without entries are not written.
__global__ void __launch_bounds__(256) csc_corr_t_kernel(const int* __restrict__ colptr, const int* __restrict__ csc_row,
                                                         const int* __restrict__ csc_src, const float* __restrict__ tnnz,
                                                         const __nv_bfloat16* __restrict__ theta_bf, int ld_theta,
                                                         float* __restrict__ corrT, int ldc, int V, int Kp) {
  __shared__ float tile[32][kMaxKp + 1];
  __shared__ int nonempty[32];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int v0 = blockIdx.x * 32;
  for (int i = 0; i < 4; ++i) {
    const int wl = warp * 4 + i, v = v0 + wl;
    int e0 = 0, e1 = 0;
    if (v < V) {
      e0 = colptr[v];
      e1 = colptr[v + 1];
    }
    if (lane == 0) nonempty[wl] = e1 > e0;
    if (e1 == e0) continue;
    float acc[kMaxKp / 32];
#pragma unroll
    for (int c = 0; c < kMaxKp / 32; ++c) acc[c] = 0.f;
    for (int e = e0; e < e1; ++e) {
      const int b = csc_row[e];
      const float w = tnnz[csc_src[e]];
      const __nv_bfloat16* tr = theta_bf + (size_t)b * ld_theta;
#pragma unroll
      for (int c = 0; c < kMaxKp / 32; ++c) {
        const int k = c * 32 + lane;
        if (k < Kp) acc[c] = fmaf(w, __bfloat162float(tr[k]), acc[c]);
      }
    }
#pragma unroll
    for (int c = 0; c < kMaxKp / 32; ++c) {
      const int k = c * 32 + lane;
      if (k < Kp) tile[wl][k] = acc[c];
    }
  }
  __syncthreads();
  const int v = v0 + lane;
  if (v < V && nonempty[lane]) {
    for (int k = warp; k < Kp; k += 8) corrT[(size_t)k * ldc + v] = tile[lane][k];
  }
}

// out[v, :] (+)= scale * sum over the column's entries of val[e] * dense[b, :]  (one warp per word).
// Used for d W1^T (val = counts, dense = d pre_0) and by the NeuralLDA decoder.
__global__ void __launch_bounds__(256) csc_spmm_kernel(const int* __restrict__ colptr, const int* __restrict__ csc_row,
                                                       const int* __restrict__ csc_src, const float* __restrict__ val,
                                                       const float* __restrict__ dense, int ldd,
                                                       float* __restrict__ out, int ldo, int V, int N, int accumulate) {
  const int v = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (v >= V) return;
  const int e0 = colptr[v], e1 = colptr[v + 1];
  float acc[kMaxKp / 32];
#pragma unroll
  for (int c = 0; c < kMaxKp / 32; ++c) acc[c] = 0.f;
  for (int e = e0; e < e1; ++e) {
    const int b = csc_row[e];
    const float w = val[csc_src[e]];
    const float* dr = dense + (size_t)b * ldd;
#pragma unroll
    for (int c = 0; c < kMaxKp / 32; ++c) {
      const int j = c * 32 + lane;
      if (j < N) acc[c] = fmaf(w, __ldg(dr + j), acc[c]);
    }
  }
#pragma unroll
  for (int c = 0; c < kMaxKp / 32; ++c) {
    const int j = c * 32 + lane;
    if (j < N) {
      float* o = out + (size_t)v * ldo + j;
      *o = accumulate ? (*o + acc[c]) : acc[c];
    }
  }
}

// ---------------------------------------------------------------------------------------------
// NeuralLDA ('LDA') decoder:  beta' = softmax_V(BN_K(beta)),  word_dist = theta . beta'.
// Only the non-zero words of each document are ever needed.
// ---------------------------------------------------------------------------------------------
// Stage 1 (one thread per word): BatchNorm over the K topic rows of beta for this word column.
__global__ void lda_beta_bn_kernel(const float* __restrict__ beta, int ldb, int K, int V, float eps, float momentum,
                                   int training, float* __restrict__ running_mean, float* __restrict__ running_var,
                                   float* __restrict__ xhat /*[K,V]*/, float* __restrict__ col_rstd /*[V]*/) {
  const int v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v >= V) return;
  float mean, var;
  if (training) {
    float s = 0.f;
    for (int k = 0; k < K; ++k) s += beta[(size_t)k * ldb + v];
    mean = s / K;
    float q = 0.f;
    for (int k = 0; k < K; ++k) { const float d = beta[(size_t)k * ldb + v] - mean; q = fmaf(d, d, q); }
    var = q / K;
    if (running_mean != nullptr) {
      const float unb = K > 1 ? var * K / (float)(K - 1) : var;
      running_mean[v] = (1.f - momentum) * running_mean[v] + momentum * mean;
      running_var[v] = (1.f - momentum) * running_var[v] + momentum * unb;
    }
  } else {
    mean = running_mean[v];
    var = running_var[v];
  }
  const float rstd = rsqrtf(var + eps);
  col_rstd[v] = rstd;
  for (int k = 0; k < K; ++k) xhat[(size_t)k * V + v] = (beta[(size_t)k * ldb + v] - mean) * rstd;
}
// Stage 2 (one block per topic row): row soft-max over the vocabulary -> beta'.
__device__ __forceinline__ float block_sum(float v, float* red) {
  v = warp_sum(v);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  float r = (threadIdx.x < (blockDim.x >> 5)) ? red[threadIdx.x] : 0.f;
  r = warp_sum(r);
  __syncthreads();
  if (threadIdx.x == 0) red[0] = r;
  __syncthreads();
  r = red[0];
  __syncthreads();
  return r;
}
__device__ __forceinline__ float block_max(float v, float* red) {
  v = warp_max(v);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  float r = (threadIdx.x < (blockDim.x >> 5)) ? red[threadIdx.x] : -INFINITY;
  r = warp_max(r);
  __syncthreads();
  if (threadIdx.x == 0) red[0] = r;
  __syncthreads();
  r = red[0];
  __syncthreads();
  return r;
}
__global__ void __launch_bounds__(1024) lda_row_softmax_kernel(const float* __restrict__ xhat, float* __restrict__ bp,
                                                               int V) {
  __shared__ float red[32];
  const int k = blockIdx.x;
  const float* x = xhat + (size_t)k * V;
  float m = -INFINITY;
  for (int v = threadIdx.x; v < V; v += blockDim.x) m = fmaxf(m, x[v]);
  m = block_max(m, red);
  float s = 0.f;
  for (int v = threadIdx.x; v < V; v += blockDim.x) s += __expf(x[v] - m);
  s = block_sum(s, red);
  const float inv = 1.f / s;
  for (int v = threadIdx.x; v < V; v += blockDim.x) bp[(size_t)k * V + v] = __expf(x[v] - m) * inv;
}
// Stage 3: wd = theta . beta'[:,v] at the non-zeros; RL; d theta; g[e] = -gscale * x / (wd + 1e-10) kept for the
// d beta' scatter.  kWpd warps per document (8 at small batches, where one warp per document left most SMs idle),
// four non-zeros in flight per warp: a column of beta' [K,V] is K different cache lines, and one non-zero at a
// time was a chain of nnz dependent L2 round trips (118 us at B=64, K=50, V=20k).
template <int kWpd>
__global__ void __launch_bounds__(256) lda_doc_kernel(const int* __restrict__ indptr, const int* __restrict__ indices,
                                                      const float* __restrict__ counts, const float* __restrict__ theta,
                                                      const float* __restrict__ bp /*[K,V]*/, int B, int K, int V,
                                                      float gscale, float* __restrict__ rl, double* __restrict__ loss,
                                                      float* __restrict__ dtheta, int ld_dtheta,
                                                      float* __restrict__ gnnz) {
  constexpr int kDpb = 8 / kWpd;
  constexpr int kC = kMaxKp / 32;
  const int wib = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int dslot = wib / kWpd, sub = wib % kWpd;
  const int b = blockIdx.x * kDpb + dslot;
  const bool live = b < B;
  __shared__ float dred[8][kMaxKp];
  __shared__ float rred[8];
  float th[kC], dth[kC];
#pragma unroll
  for (int c = 0; c < kC; ++c) {
    const int k = c * 32 + lane;
    th[c] = (live && k < K) ? theta[(size_t)b * K + k] : 0.f;
    dth[c] = 0.f;
  }
  float r = 0.f;
  if (live) {
    const int e0 = indptr[b], e1 = indptr[b + 1];
    for (int base = e0 + sub * 4; base < e1; base += 4 * kWpd) {
      int v[4];
      float x[4], bv[4][kC], dot[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const bool ok = base + u < e1;
        v[u] = ok ? indices[base + u] : 0;
        x[u] = ok ? counts[base + u] : 0.f;
      }
#pragma unroll
      for (int u = 0; u < 4; ++u)
#pragma unroll
        for (int c = 0; c < kC; ++c) {
          const int k = c * 32 + lane;
          bv[u][c] = (k < K) ? __ldg(bp + (size_t)k * V + v[u]) : 0.f;
        }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        float d = 0.f;
#pragma unroll
        for (int c = 0; c < kC; ++c) d = fmaf(th[c], bv[u][c], d);
        dot[u] = warp_sum(d);
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        if (base + u < e1) {
          r -= x[u] * __logf(dot[u] + 1e-10f);
          const float g = -gscale * x[u] / (dot[u] + 1e-10f);
          if (lane == 0 && gnnz != nullptr) gnnz[base + u] = g;
#pragma unroll
          for (int c = 0; c < kC; ++c) dth[c] = fmaf(g, bv[u][c], dth[c]);
        }
      }
    }
  }
  if (kWpd > 1) {
    if (lane == 0) rred[wib] = r;
#pragma unroll
    for (int c = 0; c < kC; ++c) dred[wib][c * 32 + lane] = dth[c];
    __syncthreads();
    if (sub == 0) {
      r = 0.f;
#pragma unroll
      for (int u = 0; u < kWpd; ++u) r += rred[wib + u];
#pragma unroll
      for (int c = 0; c < kC; ++c) {
        float acc = 0.f;
#pragma unroll
        for (int u = 0; u < kWpd; ++u) acc += dred[wib + u][c * 32 + lane];
        dth[c] = acc;
      }
    }
  }
  if (!live || sub != 0) return;
  if (dtheta != nullptr) {
#pragma unroll
    for (int c = 0; c < kC; ++c) {
      const int k = c * 32 + lane;
      if (k < ld_dtheta) dtheta[(size_t)b * ld_dtheta + k] = (k < K) ? dth[c] : 0.f;
    }
  }
  if (lane == 0) {
    rl[b] = r;
    if (loss != nullptr) atomicAdd(loss, (double)r);
  }
}
// Stage 4 (one block per topic row): soft-max backward along V given d beta' (sparse, scattered
// through the CSC of the batch into dbp [V,K] beforehand):  dxhat = bp * (dbp - sum_v dbp*bp).
__global__ void __launch_bounds__(1024) lda_row_softmax_bwd_kernel(const float* __restrict__ bp, const float* __restrict__ dbpT,
                                                                   int ldt, float* __restrict__ dxhat, int K, int V) {
  __shared__ float red[32];
  const int k = blockIdx.x;
  float s = 0.f;
  for (int v = threadIdx.x; v < V; v += blockDim.x) s = fmaf(dbpT[(size_t)v * ldt + k], bp[(size_t)k * V + v], s);
  const float dot = block_sum(s, red);
  for (int v = threadIdx.x; v < V; v += blockDim.x) {
    const float p = bp[(size_t)k * V + v];
    dxhat[(size_t)k * V + v] = p * (dbpT[(size_t)v * ldt + k] - dot);
  }
}
// Stage 5 (one thread per word): BatchNorm-over-K backward -> d beta[:, v].
__global__ void lda_beta_bn_bwd_kernel(const float* __restrict__ dxhat, const float* __restrict__ xhat,
                                       const float* __restrict__ col_rstd, float* __restrict__ dbeta, int ldg, int K,
                                       int V, int training) {
  const int v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v >= V) return;
  const float rstd = col_rstd[v];
  float s1 = 0.f, s2 = 0.f;
  if (training) {
    for (int k = 0; k < K; ++k) {
      const float d = dxhat[(size_t)k * V + v];
      s1 += d;
      s2 = fmaf(d, xhat[(size_t)k * V + v], s2);
    }
    s1 /= K;
    s2 /= K;
  }
  for (int k = 0; k < K; ++k) {
    const float d = dxhat[(size_t)k * V + v];
    dbeta[(size_t)k * ldg + v] = rstd * (d - s1 - xhat[(size_t)k * V + v] * s2);
  }
}

// ---------------------------------------------------------------------------------------------
// Row top-k of a [K,V] matrix (k <= 32), one block per row.  Ref: avitm.py:574 (torch.topk).
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) row_topk_kernel(const float* __restrict__ mat, int ld, int V, int topk,
                                                       int* __restrict__ out_idx, float* __restrict__ out_val) {
  __shared__ float sval[256];
  __shared__ int sidx[256];
  __shared__ int chosen[32];
  const int row = blockIdx.x;
  const float* x = mat + (size_t)row * ld;
  for (int t = 0; t < topk; ++t) {
    float best = -INFINITY;
    int bi = 0x7fffffff;
    for (int v = threadIdx.x; v < V; v += blockDim.x) {
      bool used = false;
      for (int j = 0; j < t; ++j) used |= (chosen[j] == v);
      const float val = x[v];
      if (!used && (val > best || (val == best && v < bi))) { best = val; bi = v; }
    }
    sval[threadIdx.x] = best;
    sidx[threadIdx.x] = bi;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
      if (threadIdx.x < o) {
        const float ov = sval[threadIdx.x + o];
        const int oi = sidx[threadIdx.x + o];
        if (ov > sval[threadIdx.x] || (ov == sval[threadIdx.x] && oi < sidx[threadIdx.x])) {
          sval[threadIdx.x] = ov;
          sidx[threadIdx.x] = oi;
        }
      }
      __syncthreads();
    }
    if (threadIdx.x == 0) {
      chosen[t] = sidx[0];
      out_idx[row * topk + t] = sidx[0];
      if (out_val != nullptr) out_val[row * topk + t] = sval[0];
    }
    __syncthreads();
  }
}

}  // namespace gf

using namespace gf;

extern "C" int gf_csr_to_csc(const int* indptr, const int* indices, int B, int V, int nnz, int* colptr,
                             int* cursor_scratch, int* csc_row, int* csc_src, void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  GF_ARG_CHECK(B >= 0 && V > 0 && nnz >= 0, "sizes");
  GF_CUDA_CHECK(cudaMemsetAsync(colptr, 0, sizeof(int) * (size_t)(V + 1), st));
  if (nnz > 0) {
    csc_hist_kernel<<<(nnz + 255) / 256, 256, 0, st>>>(indices, nnz, colptr);
    GF_LAUNCHED();
  }
  { const int rc = launch_scan(colptr, V + 1, cursor_scratch, st); if (rc) return rc; }
  GF_CUDA_CHECK(cudaMemsetAsync(cursor_scratch, 0, sizeof(int) * (size_t)V, st));
  if (nnz > 0) {
    csc_scatter_kernel<<<(B * 32 + 255) / 256, 256, 0, st>>>(indptr, indices, B, colptr, cursor_scratch, csc_row, csc_src);
    GF_LAUNCHED();
  }
  return GF_OK;
}

namespace gf {
__global__ void csc_hist_dyn_kernel(const int* __restrict__ indptr, int B, const int* __restrict__ indices,
                                    int* __restrict__ count) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < indptr[B]) atomicAdd(&count[indices[i] + 1], 1);
}
}  // namespace gf

extern "C" int gf_csr_to_csc_dyn(const int* indptr, const int* indices, int B, int V, int nnz_capacity, int* colptr,
                                 int* cursor_scratch, int* csc_row, int* csc_src, void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  GF_ARG_CHECK(B >= 0 && V > 0 && nnz_capacity >= 0, "sizes");
  GF_CUDA_CHECK(cudaMemsetAsync(colptr, 0, sizeof(int) * (size_t)(V + 1), st));
  if (nnz_capacity > 0) {
    gf::csc_hist_dyn_kernel<<<(nnz_capacity + 255) / 256, 256, 0, st>>>(indptr, B, indices, colptr);
    GF_LAUNCHED();
  }
  { const int rc = launch_scan(colptr, V + 1, cursor_scratch, st); if (rc) return rc; }
  GF_CUDA_CHECK(cudaMemsetAsync(cursor_scratch, 0, sizeof(int) * (size_t)V, st));
  if (nnz_capacity > 0) {
    csc_scatter_kernel<<<(B * 32 + 255) / 256, 256, 0, st>>>(indptr, indices, B, colptr, cursor_scratch, csc_row, csc_src);
    GF_LAUNCHED();
  }
  return GF_OK;
}

extern "C" int gf_dec_sparse_doc(const GfDecSparseArgs* a, void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  GF_ARG_CHECK(a->Kp <= kMaxKp && a->Kp % 16 == 0, "Kp");
  // warps per document: 8 at small batches (64 documents would otherwise be 64 warps on 148 SMs); one at large
  // batches, where 4 per document measured slower (200 vs 171 us at B=4096: the grid no longer fits one wave)
  if (a->B <= 512) dec_sparse_doc_kernel<8><<<a->B, 256, 0, st>>>(*a);
  else dec_sparse_doc_kernel<1><<<(a->B + 7) / 8, 256, 0, st>>>(*a);
  GF_LAUNCHED();
  return GF_OK;
}

extern "C" int gf_csc_corr(const int* colptr, const int* csc_row, const int* csc_src, const float* tnnz,
                           const void* theta_bf, int ld_theta, float* corr, int V, int Kp, void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  GF_ARG_CHECK(Kp <= kMaxKp, "Kp");
  csc_corr_kernel<<<((size_t)V * 32 + 255) / 256, 256, 0, st>>>(colptr, csc_row, csc_src, tnnz,
                                                                reinterpret_cast<const __nv_bfloat16*>(theta_bf), ld_theta,
                                                                corr, V, Kp);
  GF_LAUNCHED();
  return GF_OK;
}

extern "C" int gf_csc_corr_t(const int* colptr, const int* csc_row, const int* csc_src, const float* tnnz,
                             const void* theta_bf, int ld_theta, float* corrT, int ldc, int V, int Kp, void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  GF_ARG_CHECK(Kp <= kMaxKp && ldc >= V, "Kp / ldc");
  csc_corr_t_kernel<<<(V + 31) / 32, 256, 0, st>>>(colptr, csc_row, csc_src, tnnz,
                                                   reinterpret_cast<const __nv_bfloat16*>(theta_bf), ld_theta, corrT,
                                                   ldc, V, Kp);
  GF_LAUNCHED();
  return GF_OK;
}

extern "C" int gf_csc_spmm(const int* colptr, const int* csc_row, const int* csc_src, const float* val,
                           const float* dense, int ldd, float* out, int ldo, int V, int N, int accumulate,
                           void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  GF_ARG_CHECK(N <= kMaxKp, "N must be <= 256");
  csc_spmm_kernel<<<((size_t)V * 32 + 255) / 256, 256, 0, st>>>(colptr, csc_row, csc_src, val, dense, ldd, out, ldo, V,
                                                                N, accumulate);
  GF_LAUNCHED();
  return GF_OK;
}

extern "C" int gf_lda_beta_fwd(const float* beta, int ldb, int K, int V, float eps, float momentum, int training,
                               float* running_mean, float* running_var, float* xhat, float* col_rstd, float* bp,
                               void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  lda_beta_bn_kernel<<<(V + 255) / 256, 256, 0, st>>>(beta, ldb, K, V, eps, momentum, training, running_mean,
                                                      running_var, xhat, col_rstd);
  GF_LAUNCHED();
  lda_row_softmax_kernel<<<K, 1024, 0, st>>>(xhat, bp, V);
  GF_LAUNCHED();
  return GF_OK;
}

extern "C" int gf_lda_doc(const int* indptr, const int* indices, const float* counts, const float* theta,
                          const float* bp, int B, int K, int V, float gscale, float* rl, double* loss, float* dtheta,
                          int ld_dtheta, float* gnnz, void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  GF_ARG_CHECK(K <= kMaxKp && ld_dtheta <= kMaxKp, "K");
  if (B <= 512)
    lda_doc_kernel<8><<<B, 256, 0, st>>>(indptr, indices, counts, theta, bp, B, K, V, gscale, rl, loss,
                                        dtheta, ld_dtheta, gnnz);
  else
    lda_doc_kernel<1><<<(B + 7) / 8, 256, 0, st>>>(indptr, indices, counts, theta, bp, B, K, V, gscale, rl, loss,
                                                       dtheta, ld_dtheta, gnnz);
  GF_LAUNCHED();
  return GF_OK;
}

extern "C" int gf_lda_beta_bwd(const float* bp, const float* dbpT, int ldt, const float* xhat, const float* col_rstd,
                               float* dxhat, float* dbeta, int ldg, int K, int V, int training, void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  lda_row_softmax_bwd_kernel<<<K, 1024, 0, st>>>(bp, dbpT, ldt, dxhat, K, V);
  GF_LAUNCHED();
  lda_beta_bn_bwd_kernel<<<(V + 255) / 256, 256, 0, st>>>(dxhat, xhat, col_rstd, dbeta, ldg, K, V, training);
  GF_LAUNCHED();
  return GF_OK;
}

extern "C" int gf_row_topk(const float* mat, int ld, int rows, int V, int topk, int* out_idx, float* out_val,
                           void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  GF_ARG_CHECK(topk >= 1 && topk <= 32 && topk <= V, "topk must be in [1,32]");
  row_topk_kernel<<<rows, 256, 0, st>>>(mat, ld, V, topk, out_idx, out_val);
  GF_LAUNCHED();
  return GF_OK;
}

// ---------------------------------------------------------------------------------------------
// gf_csr_to_csc_static — the minibatch transpose for callers that must not allocate.
//
// gf_csr_to_csc needs V ints of scratch (the per-word write cursors) and leaves it to the caller to
// allocate them.  Some callers cannot: a host application that hands the whole device to its own
// memory pool at start-up has nothing left for cudaMalloc to give out, and a caller inside a
// latency-critical loop cannot afford cudaMalloc/cudaFree because both synchronise the device and
// stall every other stream.  For them the cursors live in the fixed array below, which is part of
// the library image: each GPU that loads the library carries these 4 MiB until the process exits.
//
// Rules that follow from sharing one array:
//   * V must not exceed GF_CSC_STATIC_MAX_V; larger vocabularies are refused (use gf_csr_to_csc).
//   * The array keeps whatever the previous call left in it.  The cursors have to start at zero,
//     so every call clears the first V of them on `stream` before the kernels run.
//   * All calls on one GPU use the same array.  Two calls whose kernels can be in flight at the same
//     time (different streams, or different host threads on different streams) race on it; issue them
//     on one stream or order them with events.
//   * The first call on each device looks up the array's device address (cudaGetSymbolAddress, with
//     that device current) and remembers it.  Make that first call during set-up, before any stream
//     capture is started on the device.
// No cudaMalloc, cudaFree or allocating library routine is reached from this entry point.
// ---------------------------------------------------------------------------------------------
#include <mutex>

static __device__ __align__(256) int g_csc_static_cursor[GF_CSC_STATIC_MAX_V];

namespace {
constexpr int kMaxCscDevices = 64;
int* g_csc_cursor_addr[kMaxCscDevices] = {nullptr};  // device address of the array, per device ordinal
std::mutex g_csc_cursor_mu;
}  // namespace

extern "C" int gf_csr_to_csc_static(const int* indptr, const int* indices, int B, int V, int nnz, int* colptr,
                                    int* csc_row, int* csc_src, void* stream_) {
  if (V > GF_CSC_STATIC_MAX_V) {
    fprintf(stderr,
            "[gfedntm] gf_csr_to_csc_static: V=%d exceeds the built-in scratch (GF_CSC_STATIC_MAX_V=%d); "
            "call gf_csr_to_csc with caller-allocated scratch instead\n",
            V, (int)GF_CSC_STATIC_MAX_V);
    return GF_ERR_ARG;
  }
  int dev = -1;
  GF_CUDA_CHECK(cudaGetDevice(&dev));
  GF_ARG_CHECK(dev >= 0 && dev < kMaxCscDevices, "device ordinal out of range for gf_csr_to_csc_static");
  int* cursor = nullptr;
  {
    std::lock_guard<std::mutex> lock(g_csc_cursor_mu);
    cursor = g_csc_cursor_addr[dev];
    if (cursor == nullptr) {
      // In host code `g_csc_static_cursor` is only a handle for this lookup, never an address to pass
      // to a kernel: the kernels get the device address resolved here, for the current device.
      GF_CUDA_CHECK(cudaGetSymbolAddress(reinterpret_cast<void**>(&cursor), g_csc_static_cursor));
      g_csc_cursor_addr[dev] = cursor;
    }
  }
  // Same launches as the allocating-caller entry point; it zeroes colptr and the first V cursors itself.
  return gf_csr_to_csc(indptr, indices, B, V, nnz, colptr, cursor, csc_row, csc_src, stream_);
}
