// Fused optimiser kernels.  Ref: optimizer.step() in federated_avitm.py:78 / avitm.py:141-143
// (torch.optim.Adam, betas=(momentum, 0.99)), plus the bf16 operand shadows the tensor-core
// decoder sweeps read (beta [Kp,ldb] and beta^T [V,Kp]).
#include <stdlib.h>
#include "ptx.cuh"
#include "opt_math.cuh"
#include "tmap.h"
#include "../include/gfedntm_kernels.h"

namespace gf {

struct AdamH {
  float lr, b1, b2, eps, bc1, bc2_sqrt, wd;  // bc1 = 1-b1^t, bc2_sqrt = sqrt(1-b2^t)
  int solver;                               // GfSolver (opt_math.cuh)
  const long long* step_ctr;                // optional: t = step_ctr[1] read on the device
};
__device__ __forceinline__ AdamH resolve_step(AdamH h) {
  if (h.step_ctr != nullptr) {
    const float t = (float)h.step_ctr[1];
    h.bc1 = 1.f - powf(h.b1, t);
    h.bc2_sqrt = sqrtf(1.f - powf(h.b2, t));
  }
  return h;
}

// per-kernel update context: the generic solver constants plus the hoisted Adam fast path (opt_math.cuh)
struct AdamCtx {
  OptH o;
  OptFast f;
};
__device__ __forceinline__ AdamCtx adam_ctx(const AdamH& h) {
  AdamCtx c;
  c.o.lr = h.lr; c.o.b1 = h.b1; c.o.b2 = h.b2; c.o.eps = h.eps; c.o.bc1 = h.bc1; c.o.bc2_sqrt = h.bc2_sqrt; c.o.wd = h.wd;
  c.o.solver = h.solver;
  c.f = opt_fast(c.o);
  return c;
}
__device__ __forceinline__ float adam_one(float p, float g, float& m, float& v, const AdamCtx& c) {
  return c.f.on ? opt_update_fast(p, g, m, v, c.f) : opt_update(p, g, m, v, c.o);
}

__global__ void __launch_bounds__(256) adam_flat_kernel(float* __restrict__ p, const float* __restrict__ g,
                                                        float* __restrict__ m, float* __restrict__ v, long long n,
                                                        AdamH h_, float gscale) {
  const AdamCtx h = adam_ctx(resolve_step(h_));
  const long long i0 = ((long long)blockIdx.x * blockDim.x + threadIdx.x) * 4;
  if (i0 + 3 < n && ((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(g) |
                      reinterpret_cast<uintptr_t>(m) | reinterpret_cast<uintptr_t>(v)) & 15) == 0) {
    float4 pp = *reinterpret_cast<float4*>(p + i0);
    const float4 gg = *reinterpret_cast<const float4*>(g + i0);
    float4 mm = *reinterpret_cast<float4*>(m + i0);
    float4 vv = *reinterpret_cast<float4*>(v + i0);
    pp.x = adam_one(pp.x, gg.x * gscale, mm.x, vv.x, h);
    pp.y = adam_one(pp.y, gg.y * gscale, mm.y, vv.y, h);
    pp.z = adam_one(pp.z, gg.z * gscale, mm.z, vv.z, h);
    pp.w = adam_one(pp.w, gg.w * gscale, mm.w, vv.w, h);
    *reinterpret_cast<float4*>(p + i0) = pp;
    *reinterpret_cast<float4*>(m + i0) = mm;
    *reinterpret_cast<float4*>(v + i0) = vv;
  } else {
    for (long long i = i0; i < i0 + 4 && i < n; ++i) {
      float mm = m[i], vv = v[i];
      p[i] = adam_one(p[i], g[i] * gscale, mm, vv, h);
      m[i] = mm;
      v[i] = vv;
    }
  }
}

// beta [K, ldp] fp32 (+ Adam) -> beta_bf [Kp, ldb] bf16 and betaT_bf [V, Kp] bf16.
// 32x32 tiles; the transposed shadow goes through shared memory so both writes are coalesced.
__global__ void __launch_bounds__(256) adam_beta_kernel(float* __restrict__ p, const float* __restrict__ g,
                                                        float* __restrict__ m, float* __restrict__ v, int K, int V,
                                                        int ldp, int ldg, __nv_bfloat16* __restrict__ beta_bf, int ldb,
                                                        __nv_bfloat16* __restrict__ betaT_bf, int Kp, AdamH h_,
                                                        float gscale, int do_update) {
  const AdamCtx h = adam_ctx(resolve_step(h_));
  __shared__ float tile[32][33];
  const int v0 = blockIdx.x * 32, k0 = blockIdx.y * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;  // 32 x 8
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    const int k = k0 + ty + r * 8, vv = v0 + tx;
    float val = 0.f;
    if (k < K && vv < V) {
      const size_t ip = (size_t)k * ldp + vv;
      val = p[ip];
      if (do_update) {
        float mm = m[ip], v2 = v[ip];
        val = adam_one(val, g[(size_t)k * ldg + vv] * gscale, mm, v2, h);
        p[ip] = val;
        m[ip] = mm;
        v[ip] = v2;
      }
      beta_bf[(size_t)k * ldb + vv] = __float2bfloat16(val);
    }
    tile[ty + r * 8][tx] = val;
  }
  __syncthreads();
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    const int vv = v0 + ty + r * 8, k = k0 + tx;
    if (vv < V && k < K) betaT_bf[(size_t)vv * Kp + k] = __float2bfloat16(tile[tx][ty + r * 8]);
  }
}

// Vectorised variant (V, ldp, ldg multiples of 4): 32 topics x 128 words per block, float4 global
// accesses, 8-byte bf16 stores for beta_bf and 32-byte rows for the transposed shadow.
__global__ void __launch_bounds__(256) adam_beta_vec_kernel(float* __restrict__ p, const float* __restrict__ g,
                                                            float* __restrict__ m, float* __restrict__ v, int K, int V,
                                                            int ldp, int ldg, __nv_bfloat16* __restrict__ beta_bf,
                                                            int ldb, __nv_bfloat16* __restrict__ betaT_bf, int Kp,
                                                            AdamH h_, float gscale, int do_update) {
  const AdamCtx h = adam_ctx(resolve_step(h_));
  __shared__ __align__(16) float tile[32][132];
  const int v0 = blockIdx.x * 128, k0 = blockIdx.y * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    const int k = k0 + ty + r * 8, vv = v0 + tx * 4;
    float4 val = make_float4(0.f, 0.f, 0.f, 0.f);
    if (k < K && vv < V) {
      const size_t ip = (size_t)k * ldp + vv;
      val = *reinterpret_cast<const float4*>(p + ip);
      if (do_update) {
        float4 gg = *reinterpret_cast<const float4*>(g + (size_t)k * ldg + vv);
        float4 mm = *reinterpret_cast<const float4*>(m + ip);
        float4 v2 = *reinterpret_cast<const float4*>(v + ip);
        val.x = adam_one(val.x, gg.x * gscale, mm.x, v2.x, h);
        val.y = adam_one(val.y, gg.y * gscale, mm.y, v2.y, h);
        val.z = adam_one(val.z, gg.z * gscale, mm.z, v2.z, h);
        val.w = adam_one(val.w, gg.w * gscale, mm.w, v2.w, h);
        *reinterpret_cast<float4*>(p + ip) = val;
        *reinterpret_cast<float4*>(m + ip) = mm;
        *reinterpret_cast<float4*>(v + ip) = v2;
      }
      __align__(8) __nv_bfloat16 o[4] = {__float2bfloat16(val.x), __float2bfloat16(val.y), __float2bfloat16(val.z),
                                        __float2bfloat16(val.w)};
      *reinterpret_cast<uint2*>(beta_bf + (size_t)k * ldb + vv) = *reinterpret_cast<const uint2*>(o);
    }
    *reinterpret_cast<float4*>(&tile[ty + r * 8][tx * 4]) = val;
  }
  __syncthreads();
  // transposed shadow: thread -> (word, 16-topic half); rows of 16 bf16 = 32 bytes
  const int wv = threadIdx.x >> 1, half = threadIdx.x & 1;
  const int vv = v0 + wv, kb = k0 + half * 16;
  if (vv < V && kb < Kp) {
    __align__(16) __nv_bfloat16 o[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) o[j] = __float2bfloat16(tile[half * 16 + j][wv]);
    uint4* dst = reinterpret_cast<uint4*>(betaT_bf + (size_t)vv * Kp + kb);
    dst[0] = reinterpret_cast<const uint4*>(o)[0];
    dst[1] = reinterpret_cast<const uint4*>(o)[1];
  }
}

// "Background" variant for the step where this update runs NEXT TO the d-theta sweep on a second stream: no
// shared memory and <= 64 registers x 128 threads per block, so two of these blocks fit beside a sweep CTA (384
// threads x 128 registers, ~224 KB of shared memory) on every SM; a fixed grid of a few blocks per SM loops over the
// (128 words x 16 topics) items.  A thread owns one word: its 4-byte accesses coalesce across the warp, and its 16
// topics of the transposed shadow leave as one 32-byte row piece.  Streaming (evict-first) accesses keep the
// 1.3 GB this kernel moves from pushing the sweep's operands out of L2.  The launcher asks for the maximum
// shared-memory carve-out although the kernel uses none: an SM runs two kernels side by side only under one
// L1 / shared-memory split, and the sweep needs all of it as shared memory.
__global__ void __launch_bounds__(128, 8) adam_beta_bg_kernel(float* __restrict__ p, const float* __restrict__ g,
                                                              float* __restrict__ m, float* __restrict__ v, int K, int V,
                                                              int ldp, int ldg, __nv_bfloat16* __restrict__ beta_bf,
                                                              int ldb, __nv_bfloat16* __restrict__ betaT_bf, int Kp,
                                                              AdamH h_, float gscale) {
  const AdamH h = resolve_step(h_);
  OptH oh;
  oh.lr = h.lr; oh.b1 = h.b1; oh.b2 = h.b2; oh.eps = h.eps; oh.bc1 = h.bc1; oh.bc2_sqrt = h.bc2_sqrt; oh.wd = h.wd;
  oh.solver = h.solver;
  const OptFast hf = opt_fast(oh);
  const int n_vt = (V + 127) / 128, n_kt = Kp / 16;
  const long long n_items = (long long)n_vt * n_kt;
  for (long long it = blockIdx.x; it < n_items; it += gridDim.x) {
    const int vv = (int)(it / n_kt) * 128 + (int)threadIdx.x, k0 = (int)(it % n_kt) * 16;
    if (vv >= V) continue;
    uint32_t packed[8];
#pragma unroll
    for (int half = 0; half < 2; ++half) {
      float pv[8], gv[8], mv[8], sv[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int k = k0 + half * 8 + j;
        pv[j] = 0.f; gv[j] = 0.f; mv[j] = 0.f; sv[j] = 0.f;
        if (k < K) {
          const size_t ip = (size_t)k * ldp + vv;
          pv[j] = __ldcs(p + ip);
          gv[j] = __ldcs(g + (size_t)k * ldg + vv);
          mv[j] = __ldcs(m + ip);
          sv[j] = __ldcs(v + ip);
        }
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int k = k0 + half * 8 + j;
        if (k < K) {
          const size_t ip = (size_t)k * ldp + vv;
          pv[j] = hf.on ? opt_update_fast(pv[j], gv[j] * gscale, mv[j], sv[j], hf)
                        : opt_update(pv[j], gv[j] * gscale, mv[j], sv[j], oh);
          __stcs(p + ip, pv[j]);
          __stcs(m + ip, mv[j]);
          __stcs(v + ip, sv[j]);
          beta_bf[(size_t)k * ldb + vv] = __float2bfloat16(pv[j]);
        }
      }
#pragma unroll
      for (int j = 0; j < 8; j += 2) packed[half * 4 + (j >> 1)] = pack_bf16(pv[j], pv[j + 1]);
    }
    uint4* dst = reinterpret_cast<uint4*>(betaT_bf + (size_t)vv * Kp + k0);
    dst[0] = make_uint4(packed[0], packed[1], packed[2], packed[3]);
    dst[1] = make_uint4(packed[4], packed[5], packed[6], packed[7]);
  }
}

}  // namespace gf

using namespace gf;

static AdamH make_h(float lr, float b1, float b2, float eps, float wd, int step, const long long* step_ctr, int solver = 0) {
  AdamH h;
  h.lr = lr; h.b1 = b1; h.b2 = b2; h.eps = eps; h.wd = wd; h.step_ctr = step_ctr;
  h.bc1 = 1.f - powf(b1, (float)step);
  h.bc2_sqrt = sqrtf(1.f - powf(b2, (float)step));
  h.solver = solver;
  return h;
}

extern "C" int gf_adam_flat(float* p, const float* g, float* m, float* v, long long n, float lr, float b1, float b2,
                            float eps, float wd, int step, float gscale, const long long* step_ctr, int solver, void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  if (n <= 0) return GF_OK;
  GF_ARG_CHECK(solver >= 0 && solver <= 4, "solver");
  GF_ARG_CHECK(step >= 1 || step_ctr != nullptr, "Adam step must be >= 1");
  if (step < 1) step = 1;
  const long long thr = (n + 3) / 4;
  adam_flat_kernel<<<(unsigned)((thr + 255) / 256), 256, 0, st>>>(p, g, m, v, n, make_h(lr, b1, b2, eps, wd, step, step_ctr, solver), gscale);
  GF_LAUNCHED();
  return GF_OK;
}

extern "C" int gf_adam_beta(float* p, const float* g, float* m, float* v, int K, int V, int ldp, int ldg,
                            void* beta_bf, int ldb, void* betaT_bf, int Kp, float lr, float b1, float b2, float eps,
                            float wd, int step, float gscale, int do_update, const long long* step_ctr, int solver,
                            void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  GF_ARG_CHECK(!do_update || step >= 1 || step_ctr != nullptr, "Adam step must be >= 1");
  if (step < 1) step = 1;
  if (V % 4 == 0 && ldp % 4 == 0 && ldg % 4 == 0 && ldb % 4 == 0 && Kp % 16 == 0) {
    dim3 gridv((V + 127) / 128, (K + 31) / 32);
    adam_beta_vec_kernel<<<gridv, 256, 0, st>>>(p, g, m, v, K, V, ldp, ldg, reinterpret_cast<__nv_bfloat16*>(beta_bf), ldb,
                                                reinterpret_cast<__nv_bfloat16*>(betaT_bf), Kp,
                                                make_h(lr, b1, b2, eps, wd, do_update ? step : 1, do_update ? step_ctr : nullptr, solver),
                                                gscale, do_update);
    GF_LAUNCHED();
    return GF_OK;
  }
  dim3 grid((V + 31) / 32, (K + 31) / 32);
  adam_beta_kernel<<<grid, 256, 0, st>>>(p, g, m, v, K, V, ldp, ldg, reinterpret_cast<__nv_bfloat16*>(beta_bf), ldb,
                                         reinterpret_cast<__nv_bfloat16*>(betaT_bf), Kp,
                                         make_h(lr, b1, b2, eps, wd, do_update ? step : 1, do_update ? step_ctr : nullptr, solver), gscale,
                                         do_update);
  GF_LAUNCHED();
  return GF_OK;
}

// Same update as gf_adam_beta for a caller that overlaps it with the d-theta sweep on another stream (see
// adam_beta_bg_kernel): needs Kp % 16 == 0; the whole update (do_update) only.
extern "C" int gf_adam_beta_bg(float* p, const float* g, float* m, float* v, int K, int V, int ldp, int ldg,
                               void* beta_bf, int ldb, void* betaT_bf, int Kp, float lr, float b1, float b2, float eps,
                               float wd, int step, float gscale, const long long* step_ctr, int solver, int blocks_per_sm,
                               void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  GF_ARG_CHECK(step >= 1 || step_ctr != nullptr, "Adam step must be >= 1");
  GF_ARG_CHECK(Kp % 16 == 0 && K <= Kp, "Kp");
  GF_ARG_CHECK(solver >= 0 && solver <= 4, "solver");
  if (step < 1) step = 1;
  if (blocks_per_sm < 1) blocks_per_sm = 2;
  int dev = 0, sms = 0;
  GF_CUDA_CHECK(cudaGetDevice(&dev));
  GF_CUDA_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  GF_CUDA_CHECK(cudaFuncSetAttribute(adam_beta_bg_kernel, cudaFuncAttributePreferredSharedMemoryCarveout,
                                     (int)cudaSharedmemCarveoutMaxShared));
  adam_beta_bg_kernel<<<sms * blocks_per_sm, 128, 0, st>>>(p, g, m, v, K, V, ldp, ldg, reinterpret_cast<__nv_bfloat16*>(beta_bf),
                                                          ldb, reinterpret_cast<__nv_bfloat16*>(betaT_bf), Kp,
                                                          make_h(lr, b1, b2, eps, wd, step, step_ctr, solver), gscale);
  GF_LAUNCHED();
  return GF_OK;
}
