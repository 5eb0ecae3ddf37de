// Encoder-side kernels (inference network + latent layer), forward and backward.
//
// Ref forward: src/models/base/pytorchavitm/avitm_network/inference_network.py:76-85 and
// decoder_network.py:102-118; CTM variants contextualized_topic_models/ctm_network/
// inference_network.py:79-94,160-193.  Ref KL: avitm_network/avitm.py:207-220.
//
// The bag-of-words input is consumed as CSR (never densified): the first layer is a sparse
// row-gather over W1^T [V,H0]; everything after it is [B,H]/[B,K]-sized and runs in one block
// per 8 documents.  Batch-coupled steps (BatchNorm over the batch) split the work into
// kernels at the points where a batch-wide reduction must be complete.
#include "ptx.cuh"
#include "tmap.h"
#include "../include/gfedntm_kernels.h"

namespace gf {

constexpr int kMaxH = 256;      // widest supported hidden layer / topic count
constexpr int kDocsPerBlock = 8;

enum { ACT_SOFTPLUS = 0, ACT_RELU, ACT_SIGMOID, ACT_TANH, ACT_LEAKYRELU, ACT_RRELU, ACT_ELU, ACT_SELU };

__device__ __forceinline__ float act_fwd(int act, float x, float slope) {
  switch (act) {
    case ACT_SOFTPLUS: return x > 20.f ? x : log1pf(__expf(x));
    case ACT_RELU: return fmaxf(x, 0.f);
    case ACT_SIGMOID: return 1.f / (1.f + __expf(-x));
    case ACT_TANH: return tanhf(x);
    case ACT_LEAKYRELU: return x > 0.f ? x : 0.01f * x;
    case ACT_RRELU: return x >= 0.f ? x : slope * x;
    case ACT_ELU: return x > 0.f ? x : (__expf(x) - 1.f);
    case ACT_SELU: return 1.0507009873554805f * (x > 0.f ? x : 1.6732632423543772f * (__expf(x) - 1.f));
  }
  return x;
}
__device__ __forceinline__ float act_bwd(int act, float x, float slope) {  // d act / d x at pre-activation x
  switch (act) {
    case ACT_SOFTPLUS: return x > 20.f ? 1.f : 1.f / (1.f + __expf(-x));
    case ACT_RELU: return x > 0.f ? 1.f : 0.f;
    case ACT_SIGMOID: { float s = 1.f / (1.f + __expf(-x)); return s * (1.f - s); }
    case ACT_TANH: { float t = tanhf(x); return 1.f - t * t; }
    case ACT_LEAKYRELU: return x > 0.f ? 1.f : 0.01f;
    case ACT_RRELU: return x >= 0.f ? 1.f : slope;
    case ACT_ELU: return x > 0.f ? 1.f : __expf(x);
    case ACT_SELU: return 1.0507009873554805f * (x > 0.f ? 1.f : 1.6732632423543772f * __expf(x));
  }
  return 1.f;
}
// RReLU slope: U(1/8, 1/3) in training (regenerated identically in backward), mean in eval.
__device__ __forceinline__ float rrelu_slope(int act, int training, uint64_t seed, uint64_t layer, uint64_t idx) {
  if (act != ACT_RRELU) return 0.f;
  if (!training) return 0.5f * (0.125f + 1.f / 3.f);
  Philox ph(seed);
  uint4 r = ph(idx, 0x52520000ull + layer);
  return 0.125f + (1.f / 3.f - 0.125f) * u32_to_unit(r.x);
}

// ---------------------------------------------------------------------------------------------
// enc_fwd: CSR gather + optional dense input + hidden layers + enc-dropout + mu/sigma heads.
// ---------------------------------------------------------------------------------------------
// seed of this step: host seed advanced by the device-side step counter (graph-replay safe)
__device__ __forceinline__ unsigned long long step_seed(unsigned long long seed, const long long* ctr) {
  return ctr != nullptr ? seed + (unsigned long long)ctr[0] * 0x9E3779B97F4A7C15ull : seed;
}

// Block-wide small product for the kDocsPerBlock documents of a block:
//   out[r][m] += sum_{k < Kred} w(k, m) * in[r][k]        (out zeroed by the caller, shared memory)
// Thread = (output m, slice of the reduction): every weight element is loaded ONCE for all documents of the block
// and the loads of a slice are independent of each other (four in flight), instead of one thread per (document,
// output) walking the whole reduction with a weight load in front of every FMA -- at batch 64 these chains of
// dependent L2 round trips were most of enc_fwd / enc_bwd.  Partial sums meet in shared-memory atomics.
template <typename WFn>
__device__ __forceinline__ void block_matvec(float* out, int ldo, const float* in, int ldi, int M, int Kred, WFn w) {
  const int G = max(1, (int)blockDim.x / M);
  const int g = (int)threadIdx.x / M, m = (int)threadIdx.x - g * M;
  if (g >= G) return;
  const int per = (Kred + G - 1) / G;
  const int k0 = g * per, k1 = min(Kred, k0 + per);
  float acc[kDocsPerBlock];
#pragma unroll
  for (int r = 0; r < kDocsPerBlock; ++r) acc[r] = 0.f;
  int k = k0;
  for (; k + 4 <= k1; k += 4) {
    const float w0 = w(k, m), w1 = w(k + 1, m), w2 = w(k + 2, m), w3 = w(k + 3, m);
#pragma unroll
    for (int r = 0; r < kDocsPerBlock; ++r) {
      const float* x = in + r * ldi + k;
      acc[r] = fmaf(w3, x[3], fmaf(w2, x[2], fmaf(w1, x[1], fmaf(w0, x[0], acc[r]))));
    }
  }
  for (; k < k1; ++k) {
    const float w0 = w(k, m);
#pragma unroll
    for (int r = 0; r < kDocsPerBlock; ++r) acc[r] = fmaf(w0, in[r * ldi + k], acc[r]);
  }
  if (k1 > k0) {
#pragma unroll
    for (int r = 0; r < kDocsPerBlock; ++r) atomicAdd(&out[r * ldo + m], acc[r]);
  }
}

__global__ void __launch_bounds__(kDocsPerBlock * 32)
enc_fwd_kernel(const GfEncArgs a_) {
  // (the argument block stays in the constant bank: a modified local copy would live on the stack, and every
  //  a.H[l] / a.pre[l] / a.Wh[l] with a run-time l would be a local-memory load in front of its use)
  const GfEncArgs& a = a_;
  const unsigned long long seed_ = step_seed(a_.seed, a_.step_ctr);
  __shared__ float hbuf[2][kDocsPerBlock][kMaxH];
  extern __shared__ float wsm[];  // optional staging of the head weights, rows padded to HL+1 (a.w_smem)
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b0 = blockIdx.x * kDocsPerBlock;
  const int b = b0 + warp;
  const int H0 = a.H[0];

  // ---- layer 0: sparse gather (+ dense part)
  if (b < a.B) {
    float acc[kMaxH / 32];
#pragma unroll
    for (int c = 0; c < kMaxH / 32; ++c) {
      const int j = c * 32 + lane;
      acc[c] = (j < H0) ? a.b1[j] : 0.f;
    }
    if (a.indptr != nullptr) {
      // 32 non-zeros per round: every lane fetches one (word, count) pair, then the warp walks them
      // with the row gathers of 8 words in flight at a time (independent loads, no serial chain)
      const int e0 = a.indptr[b], e1 = a.indptr[b + 1];
      for (int base = e0; base < e1; base += 32) {
        const int e = base + lane;
        const int v_l = (e < e1) ? a.indices[e] : 0;
        const float x_l = (e < e1) ? a.counts[e] : 0.f;
        const int n = min(32, e1 - base);
        for (int i0 = 0; i0 < n; i0 += 8) {
          float w[8][kMaxH / 32];
          float xs[8];
#pragma unroll
          for (int u = 0; u < 8; ++u) {
            const int vi = __shfl_sync(0xffffffffu, v_l, (i0 + u) & 31);
            xs[u] = (i0 + u < n) ? __shfl_sync(0xffffffffu, x_l, (i0 + u) & 31) : 0.f;
            const float* wr = a.W1T + (size_t)vi * H0;
#pragma unroll
            for (int c = 0; c < kMaxH / 32; ++c) {
              const int j = c * 32 + lane;
              w[u][c] = (j < H0) ? __ldg(wr + j) : 0.f;
            }
          }
#pragma unroll
          for (int u = 0; u < 8; ++u)
#pragma unroll
            for (int c = 0; c < kMaxH / 32; ++c) acc[c] = fmaf(xs[u], w[u][c], acc[c]);
        }
      }
    }
    if (a.pre0_ext != nullptr) {
      // dense part (contextual embedding) already contracted on the tensor cores (gf_ctm_rc)
#pragma unroll
      for (int c = 0; c < kMaxH / 32; ++c) {
        const int j = c * 32 + lane;
        if (j < H0) acc[c] += __ldg(a.pre0_ext + (size_t)b * H0 + j);
      }
    } else if (a.xdense != nullptr) {
      // dense part (contextual embedding): 32 inputs per round (one per lane), weight rows of 8 inputs in
      // flight at a time -- a one-input-at-a-time loop was a serial chain of C dependent L2 round trips
      const float* xr = a.xdense + (size_t)b * a.C;
      for (int base = 0; base < a.C; base += 32) {
        const float x_l = (base + lane < a.C) ? __ldg(xr + base + lane) : 0.f;
        const int n = min(32, a.C - base);
        for (int i0 = 0; i0 < n; i0 += 8) {
          float w[8][kMaxH / 32];
          float xs[8];
#pragma unroll
          for (int u = 0; u < 8; ++u) {
            xs[u] = (i0 + u < n) ? __shfl_sync(0xffffffffu, x_l, (i0 + u) & 31) : 0.f;
            const int ci = min(base + i0 + u, a.C - 1);
            const float* wr = a.WdT + (size_t)ci * H0;
#pragma unroll
            for (int c = 0; c < kMaxH / 32; ++c) {
              const int j = c * 32 + lane;
              w[u][c] = (j < H0) ? __ldg(wr + j) : 0.f;
            }
          }
#pragma unroll
          for (int u = 0; u < 8; ++u)
#pragma unroll
            for (int c = 0; c < kMaxH / 32; ++c) acc[c] = fmaf(xs[u], w[u][c], acc[c]);
        }
      }
    }
#pragma unroll
    for (int c = 0; c < kMaxH / 32; ++c) {
      const int j = c * 32 + lane;
      if (j < H0) {
        a.pre[0][(size_t)b * H0 + j] = acc[c];
        const float sl = rrelu_slope(a.act, a.training, seed_, 0, (uint64_t)b * H0 + j);
        hbuf[0][warp][j] = act_fwd(a.act, acc[c], sl);
      }
    }
  } else {
    for (int j = lane; j < kMaxH; j += 32) hbuf[0][warp][j] = 0.f;  // rows past the batch: defined inputs below
  }
  for (int idx = threadIdx.x; idx < kDocsPerBlock * kMaxH; idx += blockDim.x) (&hbuf[1][0][0])[idx] = 0.f;
  __syncthreads();

  // ---- hidden layers 1..L-1: block_matvec accumulates W h into the (zeroed) other buffer, then bias + activation
  int cur = 0;
  for (int l = 1; l < a.L; ++l) {
    const int M = a.H[l - 1], N = a.H[l];
    const float* W = a.Wh[l - 1];
    const float* bias = a.bh[l - 1];
    block_matvec(&hbuf[cur ^ 1][0][0], kMaxH, &hbuf[cur][0][0], kMaxH, N, M,
                 [&](int m, int n) { return __ldg(W + (size_t)n * M + m); });
    __syncthreads();
    for (int idx = threadIdx.x; idx < kDocsPerBlock * kMaxH; idx += blockDim.x) {
      const int r = idx / kMaxH, n = idx - r * kMaxH;
      hbuf[cur][r][n] = 0.f;  // accumulator of the next layer
      if (n < N) {
        float h = 0.f;
        if (b0 + r < a.B) {
          const float s = hbuf[cur ^ 1][r][n] + bias[n];
          a.pre[l][(size_t)(b0 + r) * N + n] = s;
          const float sl = rrelu_slope(a.act, a.training, seed_, l, (uint64_t)(b0 + r) * N + n);
          h = act_fwd(a.act, s, sl);
        }
        hbuf[cur ^ 1][r][n] = h;
      }
    }
    __syncthreads();
    cur ^= 1;
  }

  // ---- encoder dropout on the last hidden activation
  const int HL = a.H[a.L - 1];
  for (int idx = threadIdx.x; idx < kDocsPerBlock * HL; idx += blockDim.x) {
    const int r = idx / HL, n = idx - r * HL;
    const int bb = b0 + r;
    if (bb >= a.B) continue;
    float m = 1.f;
    if (a.training && a.enc_drop_p > 0.f) {
      if (a.mask_enc_in != nullptr) {
        m = a.mask_enc_in[(size_t)bb * HL + n];
      } else {
        Philox ph(seed_);
        const uint4 rr = ph((uint64_t)bb * HL + n, 0x454e4300ull);
        m = (u32_to_unit(rr.x) > a.enc_drop_p) ? 1.f / (1.f - a.enc_drop_p) : 0.f;
      }
    }
    if (a.mask_enc != nullptr) a.mask_enc[(size_t)bb * HL + n] = m;
    hbuf[cur][r][n] *= m;
  }
  __syncthreads();

  // ---- heads: mu_pre / ls_pre, plus BatchNorm partial sums.
  // Thread = one output n2 for ALL documents of the block: the weight row is read once (from the padded
  // shared-memory copy when it fits: consecutive threads -> consecutive rows, stride HL+1 words, no bank
  // conflicts; a thread-per-row walk over global memory touched 32 sectors per load), the BatchNorm sums of
  // the block's documents are combined in registers -> kDocsPerBlock times fewer fp64 atomics.
  const int K = a.K;
  const int ldw = HL + 1;
  if (a.w_smem) {
    // (four independent loads in flight per thread: one load -> one store at a time was a chain of L2 round trips)
    const int n_w = K * HL;  // Wmu and Wsig are contiguous [K, HL] each
    for (int i0 = threadIdx.x; i0 < 2 * n_w; i0 += 4 * blockDim.x) {
      float v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int idx = i0 + u * (int)blockDim.x;
        v[u] = (idx < 2 * n_w) ? __ldg(idx < n_w ? a.Wmu + idx : a.Wsig + (idx - n_w)) : 0.f;
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int idx = i0 + u * (int)blockDim.x;
        if (idx < 2 * n_w) {
          const int n2 = idx / HL, m = idx - n2 * HL;
          wsm[n2 * ldw + m] = v[u];
        }
      }
    }
    __syncthreads();
  }
  for (int n2 = threadIdx.x; n2 < 2 * K; n2 += blockDim.x) {
    const bool is_sig = n2 >= K;
    const int n = is_sig ? n2 - K : n2;
    const float* wg = (is_sig ? a.Wsig : a.Wmu) + (size_t)n * HL;
    const float* ws = wsm + n2 * ldw;
    const float bias = (is_sig ? a.bsig : a.bmu)[n];
    float s[kDocsPerBlock];
#pragma unroll
    for (int r = 0; r < kDocsPerBlock; ++r) s[r] = bias;
    for (int m = 0; m < HL; ++m) {
      const float w = a.w_smem ? ws[m] : __ldg(wg + m);
#pragma unroll
      for (int r = 0; r < kDocsPerBlock; ++r) s[r] = fmaf(w, hbuf[cur][r][m], s[r]);
    }
    double s1 = 0.0, s2 = 0.0;
#pragma unroll
    for (int r = 0; r < kDocsPerBlock; ++r) {
      if (b0 + r < a.B) {
        (is_sig ? a.ls_pre : a.mu_pre)[(size_t)(b0 + r) * K + n] = s[r];
        s1 += (double)s[r];
        s2 += (double)s[r] * (double)s[r];
      }
    }
    if (a.bn_sums != nullptr) {
      atomicAdd(&a.bn_sums[(is_sig ? 2 * K : 0) + n], s1);
      atomicAdd(&a.bn_sums[(is_sig ? 3 * K : K) + n], s2);
    }
  }
}

// ---------------------------------------------------------------------------------------------
// latent_fwd: BatchNorm(mu), BatchNorm(log sigma^2), reparameterise, softmax, theta dropout, KL.
// One warp per document; lanes stride over topics.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) latent_fwd_kernel(const GfLatentArgs a_) {
  // (the argument block stays in the constant bank: a modified local copy would live on the stack, and every
  //  a.H[l] / a.pre[l] / a.Wh[l] with a run-time l would be a local-memory load in front of its use)
  const GfLatentArgs& a = a_;
  const unsigned long long seed_ = step_seed(a_.seed, a_.step_ctr);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b = blockIdx.x * (blockDim.x >> 5) + warp;
  const int K = a.K;
  const double invB = 1.0 / (double)a.B;

  // running statistics (one block, after the sums are final because this kernel follows enc_fwd)
  if (a.training && blockIdx.x == 0 && a.mu_rm != nullptr) {
    for (int k = threadIdx.x; k < 2 * K; k += blockDim.x) {
      const bool sg = k >= K;
      const int kk = sg ? k - K : k;
      const double mean = a.bn_sums[(sg ? 2 * K : 0) + kk] * invB;
      double var = a.bn_sums[(sg ? 3 * K : K) + kk] * invB - mean * mean;
      if (var < 0) var = 0;
      const double unb = a.B > 1 ? var * (double)a.B / (double)(a.B - 1) : var;
      float* rm = sg ? a.ls_rm : a.mu_rm;
      float* rv = sg ? a.ls_rv : a.mu_rv;
      rm[kk] = (1.f - a.bn_momentum) * rm[kk] + a.bn_momentum * (float)mean;
      rv[kk] = (1.f - a.bn_momentum) * rv[kk] + a.bn_momentum * (float)unb;
    }
  }
  // BatchNorm constants of both heads once per block (shared memory) instead of four fp64 loads and the fp64
  // arithmetic per element; column sums of theta are combined per block before they go to global memory
  // (one atomic per topic and block instead of one per element on only K addresses).
  __shared__ float s_stat[4][kMaxH];
  __shared__ float s_cs[kMaxH];
  for (int k = threadIdx.x; k < K; k += blockDim.x) {
    float mmean, mrstd, smean, srstd;
    if (a.training) {
      const double m1 = a.bn_sums[k] * invB;
      double v1 = a.bn_sums[K + k] * invB - m1 * m1;
      const double m2 = a.bn_sums[2 * K + k] * invB;
      double v2 = a.bn_sums[3 * K + k] * invB - m2 * m2;
      if (v1 < 0) v1 = 0;
      if (v2 < 0) v2 = 0;
      mmean = (float)m1; mrstd = rsqrtf((float)v1 + a.bn_eps);
      smean = (float)m2; srstd = rsqrtf((float)v2 + a.bn_eps);
    } else {
      mmean = a.mu_rm[k]; mrstd = rsqrtf(a.mu_rv[k] + a.bn_eps);
      smean = a.ls_rm[k]; srstd = rsqrtf(a.ls_rv[k] + a.bn_eps);
    }
    s_stat[0][k] = mmean; s_stat[1][k] = mrstd; s_stat[2][k] = smean; s_stat[3][k] = srstd;
    s_cs[k] = 0.f;
  }
  __syncthreads();
  const bool live = b < a.B;

  float z[kMaxH / 32], mu[kMaxH / 32], lv[kMaxH / 32], ep[kMaxH / 32];
  float zmax = -INFINITY;
#pragma unroll
  for (int c = 0; c < kMaxH / 32; ++c) {
    const int k = c * 32 + lane;
    z[c] = -INFINITY; mu[c] = 0.f; lv[c] = 0.f; ep[c] = 0.f;
    if (live && k < K) {
      mu[c] = (a.mu_pre[(size_t)b * K + k] - s_stat[0][k]) * s_stat[1][k];
      lv[c] = (a.ls_pre[(size_t)b * K + k] - s_stat[2][k]) * s_stat[3][k];
      float e;
      if (a.eps_in != nullptr) {
        e = a.eps_in[(size_t)b * K + k];
      } else if (a.sample) {
        Philox ph(seed_);
        const uint4 rr = ph((uint64_t)b * K + k, 0x45505300ull);
        e = box_muller(rr.x, rr.y).x;
      } else {
        e = 0.f;
      }
      ep[c] = e;
      z[c] = mu[c] + e * __expf(0.5f * lv[c]);
      zmax = fmaxf(zmax, z[c]);
    }
  }
  zmax = warp_max(zmax);
  float zs = 0.f;
#pragma unroll
  for (int c = 0; c < kMaxH / 32; ++c) {
    const int k = c * 32 + lane;
    if (live && k < K) { z[c] = __expf(z[c] - zmax); zs += z[c]; }
  }
  zs = warp_sum(zs);
  const float inv = 1.f / zs;
  float kl = 0.f;
#pragma unroll
  for (int c = 0; c < kMaxH / 32; ++c) {
    const int k = c * 32 + lane;
    if (live && k < K) {
      const float th = z[c] * inv;
      float m = 1.f;
      if (a.training && a.theta_drop_p > 0.f) {
        if (a.mask_theta_in != nullptr) {
          m = a.mask_theta_in[(size_t)b * K + k];
        } else {
          Philox ph(seed_);
          const uint4 rr = ph((uint64_t)b * K + k, 0x54484500ull);
          m = (u32_to_unit(rr.x) > a.theta_drop_p) ? 1.f / (1.f - a.theta_drop_p) : 0.f;
        }
      }
      const size_t o = (size_t)b * K + k;
      a.mu[o] = mu[c];
      a.logvar[o] = lv[c];
      a.eps[o] = ep[c];
      a.theta_s[o] = th;
      a.mask_theta[o] = m;
      const float thd = th * m;
      a.theta[o] = thd;
      if (a.theta_colsum != nullptr) atomicAdd(&s_cs[k], thd);
      const float pv = a.prior_var[k], pm = a.prior_mean[k];
      const float var = __expf(lv[c]);
      const float d = pm - mu[c];
      kl += var / pv + d * d / pv - 1.f + __logf(pv) - lv[c];
    }
  }
  kl = 0.5f * warp_sum(kl);
  if (live && lane == 0) {
    a.kl[b] = kl;
    if (a.loss != nullptr) atomicAdd(a.loss, (double)(kl * a.kl_weight));
  }
  if (a.theta_colsum != nullptr) {
    __syncthreads();
    for (int k = threadIdx.x; k < K; k += blockDim.x) atomicAdd(&a.theta_colsum[k], s_cs[k]);
  }
}

// theta_c (bf16) = theta - column mean (training) or theta (eval), written in two layouts:
// out [B, ld] (zero padded columns) and outT [Kp = ld rows, ldT] (documents contiguous).
__global__ void theta_center_kernel(const float* __restrict__ theta, const float* __restrict__ colsum,
                                    __nv_bfloat16* __restrict__ out, __nv_bfloat16* __restrict__ outT,
                                    float* __restrict__ mean_out, int B, int K, int ld, int ldT, int center) {
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
    if (k < ld && b < B) outT[(size_t)k * ldT + b] = __float2bfloat16(tile[tx][ty + r * 8]);
  }
}

// ---------------------------------------------------------------------------------------------
// latent_bwd: d theta -> (dropout, softmax, reparameterisation, KL) -> d mu_bn, d logvar_bn,
// prior gradients and the BatchNorm-backward batch sums.  One warp per document.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) latent_bwd_kernel(const GfLatentBwdArgs a) {
  // per-block partial sums of the six per-topic reductions (8 documents per block): one global atomic per
  // block and topic instead of one per document and topic
  __shared__ float red[6][kMaxH];
  for (int i = threadIdx.x; i < 6 * kMaxH; i += blockDim.x) (&red[0][0])[i] = 0.f;
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b = blockIdx.x * 8 + warp;
  const int K = a.K;
  if (b < a.B) {
    float g[kMaxH / 32], ths[kMaxH / 32];
    float dot = 0.f;
#pragma unroll
    for (int c = 0; c < kMaxH / 32; ++c) {
      const int k = c * 32 + lane;
      g[c] = 0.f; ths[c] = 0.f;
      if (k < K) {
        const size_t o = (size_t)b * K + k;
        g[c] = a.dtheta[(size_t)b * a.ld_dtheta + k] * a.mask_theta[o];
        ths[c] = a.theta_s[o];
        dot = fmaf(g[c], ths[c], dot);
      }
    }
    dot = warp_sum(dot);
#pragma unroll
    for (int c = 0; c < kMaxH / 32; ++c) {
      const int k = c * 32 + lane;
      if (k < K) {
        const size_t o = (size_t)b * K + k;
        const float dz = ths[c] * (g[c] - dot);
        const float mu = a.mu[o], lv = a.logvar[o], ep = a.eps[o];
        const float pv = a.prior_var[k], pm = a.prior_mean[k];
        const float var = __expf(lv), sd = __expf(0.5f * lv);
        const float w = a.kl_weight * a.gscale;
        const float dmu = dz + w * (mu - pm) / pv;
        const float dlv = dz * 0.5f * ep * sd + w * 0.5f * (var / pv - 1.f);
        a.dmu[o] = dmu;
        a.dlv[o] = dlv;
        // BatchNorm backward needs sum_b d and sum_b d*xhat per topic
        atomicAdd(&red[0][k], dmu);
        atomicAdd(&red[1][k], dmu * mu);
        atomicAdd(&red[2][k], dlv);
        atomicAdd(&red[3][k], dlv * lv);
        if (a.d_prior_mean != nullptr) {
          const float d = pm - mu;
          atomicAdd(&red[4][k], w * d / pv);
          atomicAdd(&red[5][k], w * 0.5f * (-var / (pv * pv) - d * d / (pv * pv) + 1.f / pv));
        }
      }
    }
  }
  __syncthreads();
  for (int k = threadIdx.x; k < K; k += blockDim.x) {
    atomicAdd(&a.bn_bsums[k], red[0][k]);
    atomicAdd(&a.bn_bsums[K + k], red[1][k]);
    atomicAdd(&a.bn_bsums[2 * K + k], red[2][k]);
    atomicAdd(&a.bn_bsums[3 * K + k], red[3][k]);
    if (a.d_prior_mean != nullptr) {
      atomicAdd(&a.d_prior_mean[k], red[4][k]);
      atomicAdd(&a.d_prior_var[k], red[5][k]);
    }
  }
}

// ---------------------------------------------------------------------------------------------
// enc_bwd: BatchNorm backward for both heads, head / hidden-layer weight gradients, and the
// gradient w.r.t. the first-layer pre-activation (consumed by the CSC scatter kernel).
// Weight gradients are accumulated per block in registers/shared memory, then added with atomics.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kDocsPerBlock * 32)
enc_bwd_kernel(const GfEncBwdArgs a_) {
  // (the argument block stays in the constant bank: a modified local copy would live on the stack, and every
  //  a.H[l] / a.pre[l] / a.Wh[l] with a run-time l would be a local-memory load in front of its use)
  const GfEncBwdArgs& a = a_;
  const unsigned long long seed_ = step_seed(a_.seed, a_.step_ctr);
  __shared__ float hact[kDocsPerBlock][kMaxH];   // forward activation of the layer below
  __shared__ float dcur[kDocsPerBlock][kMaxH];   // gradient w.r.t. current layer output
  __shared__ float dhead[kDocsPerBlock][2 * kMaxH];  // d mu_pre | d ls_pre
  const int b0 = blockIdx.x * kDocsPerBlock;
  const int K = a.K, L = a.L;
  const int HL = a.H[L - 1];
  const float invB = 1.f / (float)a.B;

  // ---- BN backward on the heads: dpre = rstd * (d - mean(d) - xhat * mean(d*xhat))
  for (int idx = threadIdx.x; idx < kDocsPerBlock * 2 * K; idx += blockDim.x) {
    const int r = idx / (2 * K), n2 = idx - r * 2 * K;
    const int bb = b0 + r;
    float v = 0.f;
    if (bb < a.B) {
      const bool sg = n2 >= K;
      const int k = sg ? n2 - K : n2;
      const size_t o = (size_t)bb * K + k;
      const float d = (sg ? a.dlv : a.dmu)[o];
      const float xh = (sg ? a.logvar : a.mu)[o];
      const double m1 = a.bn_sums[(sg ? 2 * K : 0) + k] / (double)a.B;
      double var = a.bn_sums[(sg ? 3 * K : K) + k] / (double)a.B - m1 * m1;
      if (var < 0) var = 0;
      const float rstd = rsqrtf((float)var + a.bn_eps);
      const float s1 = a.bn_bsums[(sg ? 2 * K : 0) + k] * invB;
      const float s2 = a.bn_bsums[(sg ? 3 * K : K) + k] * invB;
      v = rstd * (d - s1 - xh * s2);
    }
    dhead[r][n2] = v;
  }
  // ---- forward activation of the last hidden layer (post act, post dropout)
  for (int idx = threadIdx.x; idx < kDocsPerBlock * HL; idx += blockDim.x) {
    const int r = idx / HL, n = idx - r * HL;
    const int bb = b0 + r;
    float h = 0.f;
    if (bb < a.B) {
      const size_t o = (size_t)bb * HL + n;
      const float sl = rrelu_slope(a.act, 1, seed_, L - 1, o);
      h = act_fwd(a.act, a.pre[L - 1][o], sl) * a.mask_enc[o];
    }
    hact[r][n] = h;
    dcur[r][n] = 0.f;  // accumulator of the d h_last product below
  }
  __syncthreads();

  // ---- head weight / bias gradients
  for (int idx = threadIdx.x; idx < 2 * K * HL; idx += blockDim.x) {
    const int n2 = idx / HL, m = idx - n2 * HL;
    float s = 0.f;
#pragma unroll
    for (int r = 0; r < kDocsPerBlock; ++r) s = fmaf(dhead[r][n2], hact[r][m], s);
    const bool sg = n2 >= K;
    atomicAdd(&(sg ? a.dWsig : a.dWmu)[(size_t)(sg ? n2 - K : n2) * HL + m], s);
  }
  for (int n2 = threadIdx.x; n2 < 2 * K; n2 += blockDim.x) {
    float s = 0.f;
#pragma unroll
    for (int r = 0; r < kDocsPerBlock; ++r) s += dhead[r][n2];
    const bool sg = n2 >= K;
    atomicAdd(&(sg ? a.dbsig : a.dbmu)[sg ? n2 - K : n2], s);
  }
  // ---- d h_last = Wmu^T dmu_pre + Wsig^T dls_pre, then dropout mask
  block_matvec(&dcur[0][0], kMaxH, &dhead[0][0], 2 * kMaxH, HL, 2 * K, [&](int k, int m) {
    return __ldg((k < K ? a.Wmu + (size_t)k * HL : a.Wsig + (size_t)(k - K) * HL) + m);
  });
  __syncthreads();
  for (int idx = threadIdx.x; idx < kDocsPerBlock * HL; idx += blockDim.x) {
    const int r = idx / HL, m = idx - r * HL;
    const int bb = b0 + r;
    dcur[r][m] = (bb < a.B) ? dcur[r][m] * a.mask_enc[(size_t)bb * HL + m] : 0.f;
  }
  __syncthreads();

  // ---- hidden layers, top down
  for (int l = L - 1; l >= 1; --l) {
    const int N = a.H[l], M = a.H[l - 1];
    // dpre_l = dcur * act'(pre_l)   (in place)
    for (int idx = threadIdx.x; idx < kDocsPerBlock * N; idx += blockDim.x) {
      const int r = idx / N, n = idx - r * N;
      const int bb = b0 + r;
      float v = 0.f;
      if (bb < a.B) {
        const size_t o = (size_t)bb * N + n;
        const float sl = rrelu_slope(a.act, 1, seed_, l, o);
        v = dcur[r][n] * act_bwd(a.act, a.pre[l][o], sl);
      }
      dcur[r][n] = v;
    }
    // activation of layer l-1
    for (int idx = threadIdx.x; idx < kDocsPerBlock * M; idx += blockDim.x) {
      const int r = idx / M, m = idx - r * M;
      const int bb = b0 + r;
      float h = 0.f;
      if (bb < a.B) {
        const size_t o = (size_t)bb * M + m;
        const float sl = rrelu_slope(a.act, 1, seed_, l - 1, o);
        h = act_fwd(a.act, a.pre[l - 1][o], sl);
      }
      hact[r][m] = h;
      dhead[r][m] = 0.f;  // (dhead is scratch from here on: accumulator of d h_{l-1})
    }
    __syncthreads();
    for (int idx = threadIdx.x; idx < N * M; idx += blockDim.x) {
      const int n = idx / M, m = idx - n * M;
      float s = 0.f;
#pragma unroll
      for (int r = 0; r < kDocsPerBlock; ++r) s = fmaf(dcur[r][n], hact[r][m], s);
      atomicAdd(&a.dWh[l - 1][idx], s);
    }
    for (int n = threadIdx.x; n < N; n += blockDim.x) {
      float s = 0.f;
#pragma unroll
      for (int r = 0; r < kDocsPerBlock; ++r) s += dcur[r][n];
      atomicAdd(&a.dbh[l - 1][n], s);
    }
    // d h_{l-1} = W_l^T dpre_l  -> stage in dhead (reused as scratch), then copy to dcur
    {
      const float* Wl = a.Wh[l - 1];
      block_matvec(&dhead[0][0], 2 * kMaxH, &dcur[0][0], kMaxH, M, N,
                   [&](int n, int m) { return __ldg(Wl + (size_t)n * M + m); });
    }
    __syncthreads();
    for (int idx = threadIdx.x; idx < kDocsPerBlock * M; idx += blockDim.x) {
      const int r = idx / M, m = idx - r * M;
      dcur[r][m] = dhead[r][m];
    }
    __syncthreads();
  }

  // ---- layer 0: d pre_0 = dcur * act'(pre_0); bias gradient; write for the scatter kernels
  const int H0 = a.H[0];
  for (int idx = threadIdx.x; idx < kDocsPerBlock * H0; idx += blockDim.x) {
    const int r = idx / H0, n = idx - r * H0;
    const int bb = b0 + r;
    float v = 0.f;
    if (bb < a.B) {
      const size_t o = (size_t)bb * H0 + n;
      const float sl = rrelu_slope(a.act, 1, seed_, 0, o);
      v = dcur[r][n] * act_bwd(a.act, a.pre[0][o], sl);
      a.dpre0[o] = v;
    }
    dcur[r][n] = v;
  }
  __syncthreads();
  for (int n = threadIdx.x; n < H0; n += blockDim.x) {
    float s = 0.f;
#pragma unroll
    for (int r = 0; r < kDocsPerBlock; ++r) s += dcur[r][n];
    atomicAdd(&a.db1[n], s);
  }
  // dense first-layer input (CTM): d WdT[c, j] += sum_b x[b,c] * dpre0[b,j]
  if (a.xdense != nullptr && a.dWdT != nullptr) {
    for (int idx = threadIdx.x; idx < a.C * H0; idx += blockDim.x) {
      const int c = idx / H0, j = idx - c * H0;
      float s = 0.f;
#pragma unroll
      for (int r = 0; r < kDocsPerBlock; ++r)
        if (b0 + r < a.B) s = fmaf(__ldg(a.xdense + (size_t)(b0 + r) * a.C + c), dcur[r][j], s);
      atomicAdd(&a.dWdT[idx], s);
    }
  }
}

// ---------------------------------------------------------------------------------------------
// bow_gather: first-layer contribution of the sparse bag of words, out[b,:] (+)= sum_e x_e * W1T[v_e,:].
// The gather is a chain of dependent DRAM round trips per warp (indices, then the rows); with one warp per
// document it dominated the encoder at every batch size.  Here `wpd` warps share a document (rounds of 32
// non-zeros dealt round-robin), each keeps kRows weight rows in flight, and the partial sums meet in shared
// memory in a fixed order.  Ref: inference_network.py:78 (x . W_in^T for the BoW part of the input).
// ---------------------------------------------------------------------------------------------
template <int kNC>
__global__ void __launch_bounds__(256)
bow_gather_kernel(const int* __restrict__ indptr, const int* __restrict__ indices, const float* __restrict__ counts,
                  const float* __restrict__ W1T, int H0, int B, float* __restrict__ out, int accumulate, int wpd) {
  constexpr int kRows = kNC <= 2 ? 16 : (kNC <= 4 ? 8 : 4);
  __shared__ float red[8][kNC * 32];
  const int wib = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int dpb = 8 / wpd, dslot = wib / wpd, sub = wib - dslot * wpd;
  const int b = blockIdx.x * dpb + dslot;
  float acc[kNC];
#pragma unroll
  for (int c = 0; c < kNC; ++c) acc[c] = 0.f;
  if (b < B) {
    const int e0 = indptr[b], e1 = indptr[b + 1];
    for (int base = e0 + 32 * sub; base < e1; base += 32 * wpd) {
      const int e = base + lane;
      const int v_l = (e < e1) ? indices[e] : 0;
      const float x_l = (e < e1) ? counts[e] : 0.f;
      const int n = min(32, e1 - base);
      for (int i0 = 0; i0 < n; i0 += kRows) {
        float w[kRows][kNC];
        float xs[kRows];
#pragma unroll
        for (int u = 0; u < kRows; ++u) {
          const int vi = __shfl_sync(0xffffffffu, v_l, (i0 + u) & 31);
          xs[u] = (i0 + u < n) ? __shfl_sync(0xffffffffu, x_l, (i0 + u) & 31) : 0.f;
          const float* wr = W1T + (size_t)vi * H0;
#pragma unroll
          for (int c = 0; c < kNC; ++c) {
            const int j = c * 32 + lane;
            w[u][c] = (j < H0) ? __ldg(wr + j) : 0.f;
          }
        }
#pragma unroll
        for (int u = 0; u < kRows; ++u)
#pragma unroll
          for (int c = 0; c < kNC; ++c) acc[c] = fmaf(xs[u], w[u][c], acc[c]);
      }
    }
  }
#pragma unroll
  for (int c = 0; c < kNC; ++c) red[wib][c * 32 + lane] = acc[c];
  __syncthreads();
  if (b < B && sub == 0) {
#pragma unroll
    for (int c = 0; c < kNC; ++c) {
      const int j = c * 32 + lane;
      if (j < H0) {
        float t = 0.f;
        for (int u = 0; u < wpd; ++u) t += red[wib + u][j];
        float* o = out + (size_t)b * H0 + j;
        *o = accumulate ? *o + t : t;
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------
// Supervised CTM (label_size > 0).  Ref: the one-hot labels are concatenated to the encoder input
// (ctm_network/inference_network.py:79-94,178-183), a Linear(K -> L) head classifies theta and the mean
// cross-entropy joins the loss (decoding_network.py:84-85,156-159, ctm.py:292-296).
//   small_linear_fwd    out[b,:] (+)= sum_l x[b,l] W[l,:]          label rows of the first layer
//   small_linear_wgrad  dW[l,:]   = sum_b x[b,l] g[b,:]
//   label_head          logits = theta Wc^T + bc, loss += CE / B, d theta += ..., d Wc, d bc (one warp per document)
// ---------------------------------------------------------------------------------------------
__global__ void small_linear_fwd_kernel(const float* __restrict__ x, const float* __restrict__ W, float* __restrict__ out,
                                        int B, int L, int H, int accumulate) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= B * H) return;
  const int b = idx / H, h = idx - b * H;
  float s = 0.f;
  for (int l = 0; l < L; ++l) s = fmaf(x[(size_t)b * L + l], __ldg(W + (size_t)l * H + h), s);
  out[idx] = accumulate ? out[idx] + s : s;
}
__global__ void small_linear_wgrad_kernel(const float* __restrict__ x, const float* __restrict__ g, float* __restrict__ dW,
                                          int B, int L, int H) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= L * H) return;
  const int l = idx / H, h = idx - l * H;
  float s = 0.f;
  for (int b = 0; b < B; ++b) s = fmaf(x[(size_t)b * L + l], g[(size_t)b * H + h], s);
  dW[idx] = s;
}
constexpr int kMaxLabels = 256;
__global__ void __launch_bounds__(256)
label_head_kernel(const float* __restrict__ theta, const float* __restrict__ Wc, const float* __restrict__ bc,
                  const float* __restrict__ labels, int B, int K, int L, float gscale, double* __restrict__ loss,
                  float* __restrict__ dtheta, int ld_dtheta, float* __restrict__ dWc, float* __restrict__ dbc) {
  __shared__ float dl_s[8][kMaxLabels];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b = blockIdx.x * 8 + warp;
  if (b >= B) return;
  const float* th = theta + (size_t)b * K;
  // logits (lanes over labels), arg-max of the one-hot target row
  float mx = -INFINITY, tgt_val = -INFINITY;
  int tgt = 0;
  for (int l = lane; l < L; l += 32) {
    float s = bc[l];
    for (int k = 0; k < K; ++k) s = fmaf(th[k], __ldg(Wc + (size_t)l * K + k), s);
    dl_s[warp][l] = s;
    mx = fmaxf(mx, s);
    const float y = labels[(size_t)b * L + l];
    if (y > tgt_val) { tgt_val = y; tgt = l; }
  }
  mx = warp_max(mx);
  for (int o = 16; o > 0; o >>= 1) {  // arg-max across lanes (first maximum wins, like torch.argmax)
    const float ov = __shfl_xor_sync(0xffffffffu, tgt_val, o);
    const int ot = __shfl_xor_sync(0xffffffffu, tgt, o);
    if (ov > tgt_val || (ov == tgt_val && ot < tgt)) { tgt_val = ov; tgt = ot; }
  }
  float se = 0.f;
  for (int l = lane; l < L; l += 32) se += __expf(dl_s[warp][l] - mx);
  se = warp_sum(se);
  __syncwarp();
  const float lse = mx + __logf(se);
  if (lane == 0 && loss != nullptr) atomicAdd(loss, (double)((lse - dl_s[warp][tgt]) / (float)B));
  if (dtheta == nullptr) return;
  const float sc = gscale / (float)B;
  __syncwarp();
  for (int l = lane; l < L; l += 32) {
    const float d = (__expf(dl_s[warp][l] - lse) - (l == tgt ? 1.f : 0.f)) * sc;
    dl_s[warp][l] = d;
    atomicAdd(&dbc[l], d);
  }
  __syncwarp();
  for (int k = lane; k < K; k += 32) {
    float acc = 0.f;
    const float t = th[k];
    for (int l = 0; l < L; ++l) {
      const float d = dl_s[warp][l];
      acc = fmaf(d, __ldg(Wc + (size_t)l * K + k), acc);
      atomicAdd(&dWc[(size_t)l * K + k], d * t);
    }
    dtheta[(size_t)b * ld_dtheta + k] += acc;
  }
}

// ---------------------------------------------------------------------------------------------
// theta_mc: Monte-Carlo document-topic inference in ONE launch.  Ref: avitm.py:470-523 (n_samples full passes of
// get_theta, averaged) + federated_model.py:170-173 (entries < thr zeroed, rows L1-renormalised).  The encoder
// pass does not depend on the draw, so only the reparameterised soft-max is repeated: one warp per document
// keeps mu / sigma in registers, loops over the draws and applies the threshold / renormalisation epilogue.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
theta_mc_kernel(const float* __restrict__ mu_pre, const float* __restrict__ ls_pre, const float* __restrict__ mu_rm,
                const float* __restrict__ mu_rv, const float* __restrict__ ls_rm, const float* __restrict__ ls_rv,
                float bn_eps, int B, int K, int n_samples, unsigned long long seed, float thr, float* __restrict__ out) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b = blockIdx.x * 8 + warp;
  if (b >= B) return;
  float mu[kMaxH / 32], sd[kMaxH / 32], acc[kMaxH / 32];
#pragma unroll
  for (int c = 0; c < kMaxH / 32; ++c) {
    const int k = c * 32 + lane;
    mu[c] = 0.f; sd[c] = 0.f; acc[c] = 0.f;
    if (k < K) {
      mu[c] = (mu_pre[(size_t)b * K + k] - mu_rm[k]) * rsqrtf(mu_rv[k] + bn_eps);
      const float lv = (ls_pre[(size_t)b * K + k] - ls_rm[k]) * rsqrtf(ls_rv[k] + bn_eps);
      sd[c] = __expf(0.5f * lv);
    }
  }
  Philox ph(seed);
  for (int s = 0; s < n_samples; ++s) {
    float z[kMaxH / 32];
    float zmax = -INFINITY;
#pragma unroll
    for (int c = 0; c < kMaxH / 32; ++c) {
      const int k = c * 32 + lane;
      z[c] = -INFINITY;
      if (k < K) {
        const uint4 rr = ph((uint64_t)b * K + k, 0x4d430000ull + (uint64_t)s);
        z[c] = fmaf(box_muller(rr.x, rr.y).x, sd[c], mu[c]);
        zmax = fmaxf(zmax, z[c]);
      }
    }
    zmax = warp_max(zmax);
    float zs = 0.f;
#pragma unroll
    for (int c = 0; c < kMaxH / 32; ++c) {
      const int k = c * 32 + lane;
      if (k < K) { z[c] = __expf(z[c] - zmax); zs += z[c]; }
    }
    zs = warp_sum(zs);
    const float inv = 1.f / zs;
#pragma unroll
    for (int c = 0; c < kMaxH / 32; ++c)
      if (c * 32 + lane < K) acc[c] = fmaf(z[c], inv, acc[c]);
  }
  const float inv_n = 1.f / (float)n_samples;
  float tot = 0.f;
#pragma unroll
  for (int c = 0; c < kMaxH / 32; ++c) {
    acc[c] *= inv_n;
    if (thr > 0.f && acc[c] < thr) acc[c] = 0.f;
    if (c * 32 + lane < K) tot += acc[c];
  }
  if (thr > 0.f) {
    tot = warp_sum(tot);
    const float r = tot > 0.f ? 1.f / tot : 1.f;
#pragma unroll
    for (int c = 0; c < kMaxH / 32; ++c) acc[c] *= r;
  }
#pragma unroll
  for (int c = 0; c < kMaxH / 32; ++c) {
    const int k = c * 32 + lane;
    if (k < K) out[(size_t)b * K + k] = acc[c];
  }
}

// ctr[0] += inc_rng (every forward pass), ctr[1] += inc_opt (every optimiser step)
__global__ void step_begin_kernel(long long* ctr, int inc_rng, int inc_opt) {
  if (threadIdx.x == 0 && blockIdx.x == 0) {
    ctr[0] += inc_rng;
    ctr[1] += inc_opt;
  }
}

// The same plus the step's zero fills (accumulators, small gradient region, column sums) in ONE launch: at batch
// 64 a training step is ~25 short kernels, and four separate fills were a sixth of them.
struct ZeroList {
  void* p[GF_ZERO_MAX];
  long long bytes[GF_ZERO_MAX];
};
__global__ void __launch_bounds__(256) step_begin_zero_kernel(long long* ctr, int inc_rng, int inc_opt, ZeroList z) {
  if (threadIdx.x == 0 && blockIdx.x == 0 && ctr != nullptr) {
    ctr[0] += inc_rng;
    ctr[1] += inc_opt;
  }
  const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x, nthr = (long long)gridDim.x * blockDim.x;
#pragma unroll
  for (int i = 0; i < GF_ZERO_MAX; ++i) {
    if (z.p[i] == nullptr) continue;
    const long long n16 = z.bytes[i] >> 4;
    uint4* p16 = reinterpret_cast<uint4*>(z.p[i]);
    for (long long q = tid; q < n16; q += nthr) p16[q] = make_uint4(0u, 0u, 0u, 0u);
    uint32_t* p4 = reinterpret_cast<uint32_t*>(z.p[i]) + (n16 << 2);
    const long long n4 = (z.bytes[i] & 15) >> 2;
    if (tid < n4) p4[tid] = 0u;
  }
}

}  // namespace gf

using namespace gf;

extern "C" int gf_step_begin_zero(long long* ctr, int inc_rng, int inc_opt, void* const* ptrs, const long long* bytes,
                                  int n, void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  GF_ARG_CHECK(n >= 0 && n <= GF_ZERO_MAX, "at most GF_ZERO_MAX buffers");
  ZeroList z{};
  long long most = 0;
  for (int i = 0; i < n; ++i) {
    GF_ARG_CHECK((reinterpret_cast<uintptr_t>(ptrs[i]) & 15) == 0 && (bytes[i] & 3) == 0, "16-byte aligned buffers, sizes % 4");
    z.p[i] = bytes[i] > 0 ? ptrs[i] : nullptr;
    z.bytes[i] = bytes[i];
    if (bytes[i] > most) most = bytes[i];
  }
  long long blocks = (most / 16 + 255) / 256;
  if (blocks < 1) blocks = 1;
  if (blocks > 2 * num_sms()) blocks = 2 * num_sms();
  step_begin_zero_kernel<<<(int)blocks, 256, 0, st>>>(ctr, inc_rng, inc_opt, z);
  GF_LAUNCHED();
  return GF_OK;
}

extern "C" int gf_step_begin(long long* ctr, int inc_rng, int inc_opt, void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  gf::step_begin_kernel<<<1, 32, 0, st>>>(ctr, inc_rng, inc_opt);
  GF_LAUNCHED();
  return GF_OK;
}

extern "C" int gf_enc_fwd(const GfEncArgs* a, void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  GF_ARG_CHECK(a->L >= 1 && a->L <= GF_MAX_HIDDEN, "number of hidden layers");
  for (int l = 0; l < a->L; ++l) GF_ARG_CHECK(a->H[l] >= 1 && a->H[l] <= kMaxH, "hidden size must be <= 256");
  GF_ARG_CHECK(a->K >= 1 && a->K <= kMaxH, "n_components must be <= 256");
  const int grid = (a->B + kDocsPerBlock - 1) / kDocsPerBlock;
  GfEncArgs aa = *a;
  const size_t wbytes = (size_t)2 * a->K * (a->H[a->L - 1] + 1) * sizeof(float);
  aa.w_smem = wbytes <= 160 * 1024;  // head weights staged in shared memory when they fit
  const size_t dyn = aa.w_smem ? wbytes : 0;
  if (dyn > 0) GF_CUDA_CHECK(cudaFuncSetAttribute(enc_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn));
  enc_fwd_kernel<<<grid, kDocsPerBlock * 32, dyn, st>>>(aa);
  GF_LAUNCHED();
  return GF_OK;
}

extern "C" int gf_latent_fwd(const GfLatentArgs* a, void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  GF_ARG_CHECK(a->K >= 1 && a->K <= kMaxH, "n_components must be <= 256");
  GF_ARG_CHECK(!a->training || a->B >= 2, "training-mode BatchNorm needs at least 2 documents");
  // four documents per block: 78 registers per thread allow 3 x 256 threads per SM, so 8-document blocks left
  // the second wave of a 4096-document batch mostly empty (1.15 waves)
  const int grid = (a->B + 3) / 4;
  latent_fwd_kernel<<<grid, 128, 0, st>>>(*a);
  GF_LAUNCHED();
  return GF_OK;
}

extern "C" int gf_theta_center(const float* theta, const float* colsum, void* out_bf16, void* outT_bf16,
                               float* mean_out, int B, int K, int ld, int ldT, int center, void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  dim3 grid((B + 31) / 32, (ld + 31) / 32);
  theta_center_kernel<<<grid, 256, 0, st>>>(theta, colsum, reinterpret_cast<__nv_bfloat16*>(out_bf16),
                                            reinterpret_cast<__nv_bfloat16*>(outT_bf16), mean_out, B, K, ld, ldT,
                                            center);
  GF_LAUNCHED();
  return GF_OK;
}

extern "C" int gf_latent_bwd(const GfLatentBwdArgs* a, void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  const int grid = (a->B + 7) / 8;
  latent_bwd_kernel<<<grid, 256, 0, st>>>(*a);
  GF_LAUNCHED();
  return GF_OK;
}

extern "C" int gf_enc_bwd(const GfEncBwdArgs* a, void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  const int grid = (a->B + kDocsPerBlock - 1) / kDocsPerBlock;
  enc_bwd_kernel<<<grid, kDocsPerBlock * 32, 0, st>>>(*a);
  GF_LAUNCHED();
  return GF_OK;
}

extern "C" int gf_theta_mc(const float* mu_pre, const float* ls_pre, const float* mu_rm, const float* mu_rv,
                           const float* ls_rm, const float* ls_rv, float bn_eps, int B, int K, int n_samples,
                           unsigned long long seed, float thr, float* out, void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  GF_ARG_CHECK(K >= 1 && K <= kMaxH && B >= 1 && n_samples >= 1, "shapes");
  theta_mc_kernel<<<(B + 7) / 8, 256, 0, st>>>(mu_pre, ls_pre, mu_rm, mu_rv, ls_rm, ls_rv, bn_eps, B, K, n_samples, seed,
                                               thr, out);
  GF_LAUNCHED();
  return GF_OK;
}

extern "C" int gf_bow_gather(const int* indptr, const int* indices, const float* counts, const float* W1T, int H0, int B,
                             float* out, int accumulate, void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  GF_ARG_CHECK(H0 >= 1 && H0 <= kMaxH && B >= 1, "shapes");
  const int wpd = B <= 512 ? 8 : 4;
  const int grid = (B * wpd + 7) / 8;
  const int nc = (H0 + 31) / 32;
  if (nc <= 1) bow_gather_kernel<1><<<grid, 256, 0, st>>>(indptr, indices, counts, W1T, H0, B, out, accumulate, wpd);
  else if (nc <= 2) bow_gather_kernel<2><<<grid, 256, 0, st>>>(indptr, indices, counts, W1T, H0, B, out, accumulate, wpd);
  else if (nc <= 4) bow_gather_kernel<4><<<grid, 256, 0, st>>>(indptr, indices, counts, W1T, H0, B, out, accumulate, wpd);
  else bow_gather_kernel<8><<<grid, 256, 0, st>>>(indptr, indices, counts, W1T, H0, B, out, accumulate, wpd);
  GF_LAUNCHED();
  return GF_OK;
}

extern "C" int gf_small_linear_fwd(const float* x, const float* W, float* out, int B, int L, int H, int accumulate,
                                   void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  GF_ARG_CHECK(B >= 1 && L >= 1 && H >= 1, "shapes");
  small_linear_fwd_kernel<<<(B * H + 255) / 256, 256, 0, st>>>(x, W, out, B, L, H, accumulate);
  GF_LAUNCHED();
  return GF_OK;
}
extern "C" int gf_small_linear_wgrad(const float* x, const float* g, float* dW, int B, int L, int H, void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  GF_ARG_CHECK(B >= 1 && L >= 1 && H >= 1, "shapes");
  small_linear_wgrad_kernel<<<(L * H + 255) / 256, 256, 0, st>>>(x, g, dW, B, L, H);
  GF_LAUNCHED();
  return GF_OK;
}
extern "C" int gf_label_head(const float* theta, const float* Wc, const float* bc, const float* labels, int B, int K, int L,
                             float gscale, double* loss, float* dtheta, int ld_dtheta, float* dWc, float* dbc,
                             void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  GF_ARG_CHECK(B >= 1 && K >= 1 && L >= 1 && L <= kMaxLabels, "label_size must be <= 256");
  GF_ARG_CHECK(dtheta == nullptr || (dWc != nullptr && dbc != nullptr), "gradient outputs");
  label_head_kernel<<<(B + 7) / 8, 256, 0, st>>>(theta, Wc, bc, labels, B, K, L, gscale, loss, dtheta, ld_dtheta, dWc, dbc);
  GF_LAUNCHED();
  return GF_OK;
}
