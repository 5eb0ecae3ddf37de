// Skinny tcgen05 GEMMs of the CombinedTM encoder (and of the dense contextual input of ZeroShotTM).
//
// Reference: ctm_network/inference_network.py:160-193 -- xb = adapt_bert(x_bert) ([B,C] x [C,V]), concatenated
// with the BoW and fed to input_layer ([B,2V] x [2V,H0]) with no non-linearity in between.  The engine
// collapses the two layers: with A = adapt_bert.weight [V,C], ab = adapt_bert.bias [V] and
// W = input_layer.weight[:, V:2V]^T [V,H0],
//     x_bert . A^T . W  =  x_bert . M^T,      M^T = A^T W  [C,H0],     b_eff = b1 + ab . W,
// and the exact gradients are
//     dW = A . dM^T + ab (x) dsum   [V,H0],   dA = W . dM   [V,C],     d ab = W . dsum   [V]
// (dM^T [C,H0] and dsum = d b1 come out of the small encoder backward).  With the bias folded in as column C of
// A' = [A | ab] and row C of dM'^T = [dM^T ; dsum] these are three plain products, all of them V-sliceable:
//
//   gf_ctm_vc   out[m,n] += sum_r X[r,m] Y[r,n]     contraction over the ROWS (vocabulary) of both operands
//                                                   (both MN-major for the MMA): M'^T = A'^T W, split over V
//   gf_ctm_rc   out[r,n]  = sum_k X[r,k] Z[k,n]     A K-major, B MN-major:  dW = A' dM'^T
//   gf_ctm_rr   out[r,c]  = sum_k X[r,k] Z[c,k]     both K-major:           dA' = W dM'   (persistent over row tiles)
//
// Operands are bf16 hi + lo shadows of the fp32 tensors (x = hi + lo to 16 mantissa bits): every product is
// three MMAs, hi.hi + lo.hi + hi.lo, accumulated in fp32 in TMEM -- fp32-level accuracy on the bf16 tensor pipe
// (these GEMMs are bound by the 154 MB of A, not by the MMA count).  Tiles arrive by TMA straight from the
// shadows: [rows, 64-element 128-byte swizzled boxes]; the same global layout serves as K-major or MN-major
// operand (the byte image is identical, only the descriptor differs), so A' and W each have ONE shadow layout.
#include "ptx.cuh"
#include "opt_math.cuh"
#include "tmap.h"
#include "../include/gfedntm_kernels.h"

namespace gf {

constexpr int kCtmThreads = 256;  // warp 0 TMA, 1 MMA, 2 TMEM alloc, 4..7 epilogue

struct CtmParams {
  float* out;       // fp32 result
  float* extra;     // rr: column `n_valid` of the result goes here (d ab), may be null
  int ldo;
  int rows;         // vc: contraction length R;  rc / rr: number of result rows
  int m_total;      // vc: result rows (columns of X)
  int n_valid;      // valid result columns
  int nb;           // 64-column blocks of the narrow operand (Np = 64 nb)
  int k_total;      // rc: contraction length (columns of X, rows of Z)
  int cbw;          // rr: width of one column block (multiple of 16, <= 256)
  int n_tiles;      // rr: row tiles
  int stages;
  int tma_store;    // rr: the result leaves through shared memory and TMA stores (needs ldo % 4 == 0)
};

struct CtmSmem {
  uint64_t* full;
  uint64_t* empty;
  uint64_t* aux;    // done / accumulator barriers
  uint32_t* tmem_ptr;
};

__device__ __forceinline__ uint8_t* ctm_align(uint8_t* p) {
  return reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(p) + 1023) & ~uintptr_t(1023));
}

// -------------------------------------------------------------------------------------------------------
// vc: out[m0 + lane, n] += sum_{r in this CTA's range} X[r, m0 + lane] * Y[r, n]
// grid = (m blocks of 128, row splits).  A stage holds 64 rows: X hi / lo as two 64-column boxes each, Y hi / lo
// as nb boxes each.
// -------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kCtmThreads, 1)
ctm_vc_kernel(const __grid_constant__ CUtensorMap tm_xh, const __grid_constant__ CUtensorMap tm_xl,
              const __grid_constant__ CUtensorMap tm_yh, const __grid_constant__ CUtensorMap tm_yl, const CtmParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* base = ctm_align(smem_raw);
  const int nb = p.nb, stages = p.stages;
  const uint32_t x_bytes = 2u * 8192u, y_bytes = (uint32_t)nb * 8192u;
  const uint32_t stage_bytes = 2 * x_bytes + 2 * y_bytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(base + (size_t)stages * stage_bytes);
  uint64_t* full = bars;
  uint64_t* empty = bars + 8;
  uint64_t* done = bars + 16;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 17);
  const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tm_xh); tma_prefetch_desc(&tm_xl); tma_prefetch_desc(&tm_yh); tma_prefetch_desc(&tm_yl);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < stages; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
    mbar_init(done, 1);
    mbar_fence_init();
  }
  if (warp == 2) tmem_alloc<256>(tmem_ptr);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_ptr;

  const int m0 = blockIdx.x * 128;
  const int n_chunks = (p.rows + 63) / 64;
  const int c0 = (int)(((long long)blockIdx.y * n_chunks) / gridDim.y);
  const int c1 = (int)(((long long)(blockIdx.y + 1) * n_chunks) / gridDim.y);

  if (warp == 0) {
    if (lane == 0) {
      uint32_t it = 0;
      for (int c = c0; c < c1; ++c, ++it) {
        const int st = it % stages;
        uint8_t* sb = base + (size_t)st * stage_bytes;
        mbar_wait(&empty[st], ((it / stages) & 1) ^ 1);
        mbar_arrive_expect_tx(&full[st], stage_bytes);
        for (int h = 0; h < 2; ++h) {
          tma_load_2d(sb + h * 8192, &tm_xh, &full[st], m0 + h * 64, c * 64);
          tma_load_2d(sb + x_bytes + h * 8192, &tm_xl, &full[st], m0 + h * 64, c * 64);
        }
        for (int b = 0; b < nb; ++b) {
          tma_load_2d(sb + 2 * x_bytes + b * 8192, &tm_yh, &full[st], b * 64, c * 64);
          tma_load_2d(sb + 2 * x_bytes + y_bytes + b * 8192, &tm_yl, &full[st], b * 64, c * 64);
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      const uint32_t idesc = umma_idesc_bf16(128, (uint32_t)nb * 64u, 1u, 1u);
      uint32_t it = 0;
      for (int c = c0; c < c1; ++c, ++it) {
        const int st = it % stages;
        mbar_wait(&full[st], (it / stages) & 1);
        tc_fence_after();
        const uint32_t xh = smem_u32(base + (size_t)st * stage_bytes), xl = xh + x_bytes;
        const uint32_t yh = xh + 2 * x_bytes, yl = yh + y_bytes;
#pragma unroll
        for (int j = 0; j < 4; ++j) {  // 64 rows = 4 k-steps of 16 rows (2 KB each in every 64-column box)
          const uint64_t dxh = umma_smem_desc(xh + j * 2048u, 8192, 1024), dxl = umma_smem_desc(xl + j * 2048u, 8192, 1024);
          const uint64_t dyh = umma_smem_desc(yh + j * 2048u, 8192, 1024), dyl = umma_smem_desc(yl + j * 2048u, 8192, 1024);
          umma_f16(tmem, dxh, dyh, idesc, (c > c0 || j > 0) ? 1u : 0u);
          umma_f16(tmem, dxl, dyh, idesc, 1u);
          umma_f16(tmem, dxh, dyl, idesc, 1u);
        }
        umma_commit(&empty[st]);
      }
      umma_commit(done);
    }
  } else if (warp >= 4) {
    const uint32_t q = warp - 4;
    if (c1 > c0) {
      mbar_wait(done, 0);
      tc_fence_after();
      const int m = m0 + (int)(q * 32 + lane);
      for (int c = 0; c * 32 < nb * 64; ++c) {
        float v[32];
        tmem_ld32(tmem + ((q * 32) << 16) + c * 32, v);
        tmem_ld_wait();
        if (m < p.m_total) {
#pragma unroll
          for (int i = 0; i < 32; ++i)
            if (c * 32 + i < p.n_valid) atomicAdd(&p.out[(size_t)m * p.ldo + c * 32 + i], v[i]);
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc<256>(tmem);
}

// -------------------------------------------------------------------------------------------------------
// rc: out[r0 + lane, n] = sum_k X[r0 + lane, k] * Z[k, n]     one 128-row tile per CTA, k in stages of 64
// -------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kCtmThreads, 1)
ctm_rc_kernel(const __grid_constant__ CUtensorMap tm_xh, const __grid_constant__ CUtensorMap tm_xl,
              const __grid_constant__ CUtensorMap tm_zh, const __grid_constant__ CUtensorMap tm_zl, const CtmParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* base = ctm_align(smem_raw);
  const int nb = p.nb, stages = p.stages;
  const uint32_t x_bytes = 16384u, z_bytes = (uint32_t)nb * 8192u;
  const uint32_t stage_bytes = 2 * x_bytes + 2 * z_bytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(base + (size_t)stages * stage_bytes);
  uint64_t* full = bars;
  uint64_t* empty = bars + 8;
  uint64_t* done = bars + 16;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 17);
  const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tm_xh); tma_prefetch_desc(&tm_xl); tma_prefetch_desc(&tm_zh); tma_prefetch_desc(&tm_zl);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < stages; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
    mbar_init(done, 1);
    mbar_fence_init();
  }
  if (warp == 2) tmem_alloc<256>(tmem_ptr);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_ptr;
  const int r0 = blockIdx.x * 128;
  const int n_k = (p.k_total + 63) / 64;

  if (warp == 0) {
    if (lane == 0) {
      for (int c = 0; c < n_k; ++c) {
        const int st = c % stages;
        uint8_t* sb = base + (size_t)st * stage_bytes;
        mbar_wait(&empty[st], ((c / stages) & 1) ^ 1);
        mbar_arrive_expect_tx(&full[st], stage_bytes);
        tma_load_2d(sb, &tm_xh, &full[st], c * 64, r0);
        tma_load_2d(sb + x_bytes, &tm_xl, &full[st], c * 64, r0);
        for (int b = 0; b < nb; ++b) {
          tma_load_2d(sb + 2 * x_bytes + b * 8192, &tm_zh, &full[st], b * 64, c * 64);
          tma_load_2d(sb + 2 * x_bytes + z_bytes + b * 8192, &tm_zl, &full[st], b * 64, c * 64);
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      const uint32_t idesc = umma_idesc_bf16(128, (uint32_t)nb * 64u, 0u, 1u);
      for (int c = 0; c < n_k; ++c) {
        const int st = c % stages;
        mbar_wait(&full[st], (c / stages) & 1);
        tc_fence_after();
        const uint32_t xh = smem_u32(base + (size_t)st * stage_bytes), xl = xh + x_bytes;
        const uint32_t zh = xh + 2 * x_bytes, zl = zh + z_bytes;
#pragma unroll
        for (int j = 0; j < 4; ++j) {  // 64 contraction elements = 4 k-steps of 16 (32 bytes along the X rows)
          const uint64_t dxh = umma_smem_desc(xh + j * 32u, 16, 1024), dxl = umma_smem_desc(xl + j * 32u, 16, 1024);
          const uint64_t dzh = umma_smem_desc(zh + j * 2048u, 8192, 1024), dzl = umma_smem_desc(zl + j * 2048u, 8192, 1024);
          umma_f16(tmem, dxh, dzh, idesc, (c > 0 || j > 0) ? 1u : 0u);
          umma_f16(tmem, dxl, dzh, idesc, 1u);
          umma_f16(tmem, dxh, dzl, idesc, 1u);
        }
        umma_commit(&empty[st]);
      }
      umma_commit(done);
    }
  } else if (warp >= 4) {
    const uint32_t q = warp - 4;
    mbar_wait(done, 0);
    tc_fence_after();
    const int r = r0 + (int)(q * 32 + lane);
    for (int c = 0; c * 32 < nb * 64; ++c) {
      float v[32];
      tmem_ld32(tmem + ((q * 32) << 16) + c * 32, v);
      tmem_ld_wait();
      if (r < p.rows) {
#pragma unroll
        for (int i = 0; i < 32; ++i)
          if (c * 32 + i < p.n_valid) p.out[(size_t)r * p.ldo + c * 32 + i] = v[i];
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc<256>(tmem);
}

// -------------------------------------------------------------------------------------------------------
// rr: out[r, cb*cbw + j] = sum_k X[r,k] * Z[cb*cbw + j, k]    (K = 64 nb, everything in one stage)
// grid = (column blocks, P); CTA (cb, pidx) keeps its Z block resident and walks the row tiles pidx, pidx + P,
// ... with a double-buffered 128 x cbw accumulator, so the MMAs of tile t+1 run under the stores of tile t.
// -------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kCtmThreads, 1)
ctm_rr_kernel(const __grid_constant__ CUtensorMap tm_xh, const __grid_constant__ CUtensorMap tm_xl,
              const __grid_constant__ CUtensorMap tm_zh, const __grid_constant__ CUtensorMap tm_zl,
              const __grid_constant__ CUtensorMap tm_out, const CtmParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* base = ctm_align(smem_raw);
  const int nb = p.nb, stages = p.stages, cbw = p.cbw;
  const uint32_t zbox = (uint32_t)cbw * 128u, z_bytes = (uint32_t)nb * zbox;  // one Z copy (hi or lo)
  const uint32_t x_bytes = (uint32_t)nb * 16384u, stage_bytes = 2 * x_bytes;
  uint8_t* zs = base;
  uint8_t* xs = ctm_align(base + 2 * z_bytes);
  uint8_t* ebuf = ctm_align(xs + (size_t)stages * stage_bytes);  // 4 warps x 2 slots x (32 rows x 128 B), TMA-store staging
  uint64_t* bars = reinterpret_cast<uint64_t*>(ebuf + 32768);
  uint64_t* full = bars;        // [stages]
  uint64_t* empty = bars + 8;   // [stages]
  uint64_t* zfull = bars + 16;
  uint64_t* d_full = bars + 17;   // [2]
  uint64_t* d_empty = bars + 19;  // [2]
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 21);
  const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tm_xh); tma_prefetch_desc(&tm_xl); tma_prefetch_desc(&tm_zh); tma_prefetch_desc(&tm_zl);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < stages; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
    mbar_init(zfull, 1);
    for (int i = 0; i < 2; ++i) { mbar_init(&d_full[i], 1); mbar_init(&d_empty[i], 128); }
    mbar_fence_init();
  }
  if (warp == 2) tmem_alloc<512>(tmem_ptr);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_ptr;
  const int cb = blockIdx.x, c_base = cb * cbw;
  const int P = gridDim.y;

  if (warp == 0) {
    if (lane == 0) {
      mbar_arrive_expect_tx(zfull, 2 * z_bytes);
      for (int b = 0; b < nb; ++b) {
        tma_load_2d(zs + b * zbox, &tm_zh, zfull, b * 64, c_base);
        tma_load_2d(zs + z_bytes + b * zbox, &tm_zl, zfull, b * 64, c_base);
      }
      uint32_t it = 0;
      for (int t = blockIdx.y; t < p.n_tiles; t += P, ++it) {
        const int st = it % stages;
        uint8_t* sb = xs + (size_t)st * stage_bytes;
        mbar_wait(&empty[st], ((it / stages) & 1) ^ 1);
        mbar_arrive_expect_tx(&full[st], stage_bytes);
        for (int b = 0; b < nb; ++b) {
          tma_load_2d(sb + b * 16384, &tm_xh, &full[st], b * 64, t * 128);
          tma_load_2d(sb + x_bytes + b * 16384, &tm_xl, &full[st], b * 64, t * 128);
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      const uint32_t idesc = umma_idesc_bf16(128, (uint32_t)cbw, 0u, 0u);
      mbar_wait(zfull, 0);
      const uint32_t zh = smem_u32(zs), zl = zh + z_bytes;
      uint32_t it = 0;
      for (int t = blockIdx.y; t < p.n_tiles; t += P, ++it) {
        const int st = it % stages;
        const uint32_t buf = it & 1;
        mbar_wait(&full[st], (it / stages) & 1);
        mbar_wait(&d_empty[buf], ((it >> 1) & 1) ^ 1);
        tc_fence_after();
        const uint32_t xh = smem_u32(xs + (size_t)st * stage_bytes), xl = xh + x_bytes;
        const uint32_t d = tmem + buf * 256;
        for (int j = 0; j < 4 * nb; ++j) {
          const uint32_t xo = (uint32_t)(j >> 2) * 16384u + (uint32_t)(j & 3) * 32u;
          const uint32_t zo = (uint32_t)(j >> 2) * zbox + (uint32_t)(j & 3) * 32u;
          const uint64_t dxh = umma_smem_desc(xh + xo, 16, 1024), dxl = umma_smem_desc(xl + xo, 16, 1024);
          const uint64_t dzh = umma_smem_desc(zh + zo, 16, 1024), dzl = umma_smem_desc(zl + zo, 16, 1024);
          umma_f16(d, dxh, dzh, idesc, j > 0 ? 1u : 0u);
          umma_f16(d, dxl, dzh, idesc, 1u);
          umma_f16(d, dxh, dzl, idesc, 1u);
        }
        umma_commit(&empty[st]);
        umma_commit(&d_full[buf]);
      }
    }
  } else if (warp >= 4) {
    const uint32_t q = warp - 4;
    uint32_t it = 0, n_st = 0;
    for (int t = blockIdx.y; t < p.n_tiles; t += P, ++it) {
      const uint32_t buf = it & 1;
      mbar_wait(&d_full[buf], (it >> 1) & 1);
      tc_fence_after();
      const int r = t * 128 + (int)(q * 32 + lane);
      for (int c = 0; c * 32 < cbw; ++c) {
        float v[32];
        tmem_ld32(tmem + ((q * 32) << 16) + buf * 256 + c * 32, v);
        tmem_ld_wait();
        if (p.tma_store) {
          // registers -> 128-byte-swizzled staging tile (32 rows x 32 columns) -> one TMA store per warp and
          // chunk: the vocabulary rows of the result are 3 KB apart, so direct stores would touch 32 lines each
          const int col0 = c_base + c * 32;
          const uint32_t stg = smem_u32(ebuf + q * 8192 + (n_st & 1) * 4096);
          if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");  // slot read out two stores ago
          __syncwarp();
#pragma unroll
          for (int j = 0; j < 8; ++j)
            asm volatile("st.shared.v4.f32 [%0], {%1,%2,%3,%4};" ::"r"(stg + lane * 128 + ((j ^ (lane & 7)) << 4)),
                         "f"(v[4 * j]), "f"(v[4 * j + 1]), "f"(v[4 * j + 2]), "f"(v[4 * j + 3])
                         : "memory");
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) {
            asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%1, %2}], [%3];" ::"l"(
                             reinterpret_cast<uint64_t>(&tm_out)),
                         "r"(col0), "r"(t * 128 + (int)q * 32), "r"(stg)
                         : "memory");
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
          }
          ++n_st;
          if (p.extra != nullptr && r < p.rows && p.n_valid >= col0 && p.n_valid < col0 + 32) p.extra[r] = v[p.n_valid - col0];
        } else if (r < p.rows) {
          const int col0 = c_base + c * 32;
          float* o = p.out + (size_t)r * p.ldo + col0;
          if (col0 + 32 <= p.n_valid && c * 32 + 32 <= cbw && (p.ldo & 3) == 0) {
#pragma unroll
            for (int i = 0; i < 32; i += 4) *reinterpret_cast<float4*>(o + i) = make_float4(v[i], v[i + 1], v[i + 2], v[i + 3]);
          } else {
#pragma unroll
            for (int i = 0; i < 32; ++i) {
              if (c * 32 + i < cbw) {
                if (col0 + i < p.n_valid) o[i] = v[i];
                else if (col0 + i == p.n_valid && p.extra != nullptr) p.extra[r] = v[i];
              }
            }
          }
        }
      }
      tc_fence_before();
      mbar_arrive(&d_empty[buf]);
    }
    if (p.tma_store && lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");  // stores complete
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc<512>(tmem);
}

// -------------------------------------------------------------------------------------------------------
// bf16 hi / lo shadows:  hi = bf16(x), lo = bf16(x - hi)
// -------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void split_bf16(float x, __nv_bfloat16& hi, __nv_bfloat16& lo) {
  hi = __float2bfloat16(x);
  lo = __float2bfloat16(x - __bfloat162float(hi));
}

// out_hi / out_lo [rows (+1), ldo] <- src [rows, cols] (row pitch ld_src); column `cols` <- extra[row] (optional,
// e.g. the bias folded in as one more column); row `rows` <- tail[0..cols) (optional); the rest is zeroed.
__global__ void ctm_split_kernel(const float* __restrict__ src, long long ld_src, const float* __restrict__ extra,
                                 const float* __restrict__ tail, int rows, int cols, __nv_bfloat16* __restrict__ out_hi,
                                 __nv_bfloat16* __restrict__ out_lo, int ldo) {
  const long long n = (long long)(rows + (tail != nullptr ? 1 : 0)) * ldo;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / ldo;
    const int c = (int)(i - r * ldo);
    float x = 0.f;
    if (r == rows) x = (c < cols) ? tail[c] : 0.f;
    else if (c < cols) x = src[r * ld_src + c];
    else if (c == cols && extra != nullptr) x = extra[r];
    __nv_bfloat16 h, l;
    split_bf16(x, h, l);
    out_hi[i] = h;
    out_lo[i] = l;
  }
}

// out [rows, ld] <- 0 except row `bias_row` <- bias[0..cols)   (pre-load of the vc accumulation target: the
// effective first-layer bias b1 + ab.W accumulates into the row that the folded bias column produces)
__global__ void ctm_init_out_kernel(float* __restrict__ out, int rows, int cols, int ld, int bias_row,
                                    const float* __restrict__ bias) {
  const int n = rows * ld;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const int r = i / ld, c = i - r * ld;
    out[i] = (r == bias_row && bias != nullptr && c < cols) ? bias[c] : 0.f;
  }
}

// Adam on a row-major [rows, cols] parameter block (+ one optional per-row scalar, e.g. a bias folded in as one
// more column) fused with the refresh of its bf16 hi / lo operand shadows [rows, ldo]; do_update = 0 only
// refreshes the shadows.  One thread per 8 consecutive shadow columns (16-byte hi / lo stores).
struct AdamS {
  float lr, b1, b2, eps, bc1, bc2_sqrt, wd, gscale;
  int solver;
  const long long* step_ctr;
};
__device__ __forceinline__ float adam_s(float p, float g, float& m, float& v, const AdamS& h) {
  OptH o;
  o.lr = h.lr; o.b1 = h.b1; o.b2 = h.b2; o.eps = h.eps; o.bc1 = h.bc1; o.bc2_sqrt = h.bc2_sqrt; o.wd = h.wd;
  o.solver = h.solver;
  return opt_update(p, g * h.gscale, m, v, o);
}
__global__ void __launch_bounds__(256)
adam_split_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m, float* __restrict__ v,
                  long long ld, int rows, int cols, float* __restrict__ pc, const float* __restrict__ gc,
                  float* __restrict__ mc, float* __restrict__ vc, __nv_bfloat16* __restrict__ out_hi,
                  __nv_bfloat16* __restrict__ out_lo, int ldo, AdamS h, int do_update) {
  if (h.step_ctr != nullptr) {
    const float t = (float)h.step_ctr[1];
    h.bc1 = 1.f - powf(h.b1, t);
    h.bc2_sqrt = sqrtf(1.f - powf(h.b2, t));
  }
  const int cpr = ldo >> 3;  // 8-column chunks per shadow row
  const long long n = (long long)rows * cpr;
  const bool vec = (cols % 8 == 0) && (ld % 4 == 0) &&
                   ((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(g) | reinterpret_cast<uintptr_t>(m) |
                     reinterpret_cast<uintptr_t>(v)) & 15) == 0;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / cpr;
    const int c0 = (int)(i - r * cpr) * 8;
    float x[8];
    if (vec && c0 + 8 <= cols) {
      float* pr = p + r * ld + c0;
      float4 a0 = *reinterpret_cast<const float4*>(pr), a1 = *reinterpret_cast<const float4*>(pr + 4);
      x[0] = a0.x; x[1] = a0.y; x[2] = a0.z; x[3] = a0.w; x[4] = a1.x; x[5] = a1.y; x[6] = a1.z; x[7] = a1.w;
      if (do_update) {
        const float* gr = g + r * ld + c0;
        float* mr = m + r * ld + c0;
        float* vr = v + r * ld + c0;
        const float4 g0 = *reinterpret_cast<const float4*>(gr), g1 = *reinterpret_cast<const float4*>(gr + 4);
        float4 m0 = *reinterpret_cast<const float4*>(mr), m1 = *reinterpret_cast<const float4*>(mr + 4);
        float4 v0 = *reinterpret_cast<const float4*>(vr), v1 = *reinterpret_cast<const float4*>(vr + 4);
        x[0] = adam_s(x[0], g0.x, m0.x, v0.x, h); x[1] = adam_s(x[1], g0.y, m0.y, v0.y, h);
        x[2] = adam_s(x[2], g0.z, m0.z, v0.z, h); x[3] = adam_s(x[3], g0.w, m0.w, v0.w, h);
        x[4] = adam_s(x[4], g1.x, m1.x, v1.x, h); x[5] = adam_s(x[5], g1.y, m1.y, v1.y, h);
        x[6] = adam_s(x[6], g1.z, m1.z, v1.z, h); x[7] = adam_s(x[7], g1.w, m1.w, v1.w, h);
        *reinterpret_cast<float4*>(pr) = make_float4(x[0], x[1], x[2], x[3]);
        *reinterpret_cast<float4*>(pr + 4) = make_float4(x[4], x[5], x[6], x[7]);
        *reinterpret_cast<float4*>(mr) = m0; *reinterpret_cast<float4*>(mr + 4) = m1;
        *reinterpret_cast<float4*>(vr) = v0; *reinterpret_cast<float4*>(vr + 4) = v1;
      }
    } else {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int c = c0 + j;
        float val = 0.f;
        if (c < cols) {
          const long long ip = r * ld + c;
          val = p[ip];
          if (do_update) {
            float mm = m[ip], vv = v[ip];
            val = adam_s(val, g[ip], mm, vv, h);
            p[ip] = val; m[ip] = mm; v[ip] = vv;
          }
        } else if (c == cols && pc != nullptr) {
          val = pc[r];
          if (do_update) {
            float mm = mc[r], vv = vc[r];
            val = adam_s(val, gc[r], mm, vv, h);
            pc[r] = val; mc[r] = mm; vc[r] = vv;
          }
        }
        x[j] = val;
      }
    }
    __align__(16) __nv_bfloat16 hi[8], lo[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) split_bf16(x[j], hi[j], lo[j]);
    *reinterpret_cast<uint4*>(out_hi + r * ldo + c0) = *reinterpret_cast<const uint4*>(hi);
    *reinterpret_cast<uint4*>(out_lo + r * ldo + c0) = *reinterpret_cast<const uint4*>(lo);
  }
}

static int ctm_max_smem() {
  int max_smem = 0, dev = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
  return max_smem;
}

}  // namespace gf

using namespace gf;

extern "C" int gf_ctm_split(const float* src, long long ld_src, const float* extra, const float* tail, int rows, int cols,
                            void* out_hi, void* out_lo, int ldo, void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  GF_ARG_CHECK(ldo >= cols + (extra != nullptr ? 1 : 0) && rows > 0, "ldo");
  const long long n = (long long)(rows + (tail != nullptr ? 1 : 0)) * ldo;
  int blocks = (int)((n + 255) / 256 < 8 * (long long)num_sms() ? (n + 255) / 256 : 8 * (long long)num_sms());
  ctm_split_kernel<<<blocks, 256, 0, stream>>>(src, ld_src, extra, tail, rows, cols,
                                               reinterpret_cast<__nv_bfloat16*>(out_hi),
                                               reinterpret_cast<__nv_bfloat16*>(out_lo), ldo);
  GF_LAUNCHED();
  return GF_OK;
}

extern "C" int gf_ctm_init_out(float* out, int rows, int cols, int ld, int bias_row, const float* bias, void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  const int n = rows * ld;
  ctm_init_out_kernel<<<(n + 255) / 256, 256, 0, stream>>>(out, rows, cols, ld, bias_row, bias);
  GF_LAUNCHED();
  return GF_OK;
}

// out[m, n] += sum_r X[r, m] * Y[r, n];  X hi/lo [R, ldx] bf16 (m_total valid columns), Y hi/lo [R, 64 nb] bf16.
extern "C" int gf_ctm_vc(const void* x_hi, const void* x_lo, int ldx, int m_total, const void* y_hi, const void* y_lo,
                         int nb, int R, float* out, int ldo, int n_valid, void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  GF_ARG_CHECK(nb >= 1 && nb <= 4 && ldx % 8 == 0 && m_total <= ldx && n_valid <= nb * 64 && R > 0, "shapes");
  CUtensorMap txh, txl, tyh, tyl;
  int rc;
  if ((rc = make_tmap_bf16_2d(&txh, x_hi, (uint64_t)ldx, (uint64_t)R, (uint64_t)ldx * 2, 64, 64))) return rc;
  if ((rc = make_tmap_bf16_2d(&txl, x_lo, (uint64_t)ldx, (uint64_t)R, (uint64_t)ldx * 2, 64, 64))) return rc;
  if ((rc = make_tmap_bf16_2d(&tyh, y_hi, (uint64_t)nb * 64, (uint64_t)R, (uint64_t)nb * 128, 64, 64))) return rc;
  if ((rc = make_tmap_bf16_2d(&tyl, y_lo, (uint64_t)nb * 64, (uint64_t)R, (uint64_t)nb * 128, 64, 64))) return rc;
  CtmParams p{};
  p.out = out; p.ldo = ldo; p.rows = R; p.m_total = m_total; p.n_valid = n_valid; p.nb = nb;
  const size_t stage = 2 * 16384 + 2 * (size_t)nb * 8192;
  int stages = (int)(((size_t)ctm_max_smem() - 2048) / stage);
  if (stages > 6) stages = 6;
  GF_ARG_CHECK(stages >= 2, "shared memory");
  p.stages = stages;
  const size_t smem = 1024 + stages * stage + 256;
  const int mblocks = (m_total + 127) / 128;
  const int chunks = (R + 63) / 64;
  int splits = num_sms() / mblocks;
  if (splits < 1) splits = 1;
  if (splits > chunks) splits = chunks;
  GF_CUDA_CHECK(cudaFuncSetAttribute(ctm_vc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  ctm_vc_kernel<<<dim3(mblocks, splits), kCtmThreads, smem, stream>>>(txh, txl, tyh, tyl, p);
  GF_LAUNCHED();
  return GF_OK;
}

// out[r, n] = sum_k X[r, k] * Z[k, n];  X hi/lo [R, ldx] bf16 (k_total valid columns), Z hi/lo [k rows, 64 nb] bf16.
extern "C" int gf_ctm_rc(const void* x_hi, const void* x_lo, int ldx, int k_total, const void* z_hi, const void* z_lo,
                         int nb, int z_rows, int R, float* out, int ldo, int n_valid, void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  GF_ARG_CHECK(nb >= 1 && nb <= 4 && ldx % 8 == 0 && k_total <= ldx && n_valid <= nb * 64 && R > 0 && z_rows >= 1,
               "shapes");
  CUtensorMap txh, txl, tzh, tzl;
  int rc;
  if ((rc = make_tmap_bf16_2d(&txh, x_hi, (uint64_t)k_total, (uint64_t)R, (uint64_t)ldx * 2, 64, 128))) return rc;
  if ((rc = make_tmap_bf16_2d(&txl, x_lo, (uint64_t)k_total, (uint64_t)R, (uint64_t)ldx * 2, 64, 128))) return rc;
  if ((rc = make_tmap_bf16_2d(&tzh, z_hi, (uint64_t)nb * 64, (uint64_t)z_rows, (uint64_t)nb * 128, 64, 64))) return rc;
  if ((rc = make_tmap_bf16_2d(&tzl, z_lo, (uint64_t)nb * 64, (uint64_t)z_rows, (uint64_t)nb * 128, 64, 64))) return rc;
  CtmParams p{};
  p.out = out; p.ldo = ldo; p.rows = R; p.n_valid = n_valid; p.nb = nb; p.k_total = k_total;
  const size_t stage = 2 * 16384 + 2 * (size_t)nb * 8192;
  int stages = (int)(((size_t)ctm_max_smem() - 2048) / stage);
  if (stages > 6) stages = 6;
  GF_ARG_CHECK(stages >= 2, "shared memory");
  p.stages = stages;
  const size_t smem = 1024 + stages * stage + 256;
  GF_CUDA_CHECK(cudaFuncSetAttribute(ctm_rc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  ctm_rc_kernel<<<(R + 127) / 128, kCtmThreads, smem, stream>>>(txh, txl, tzh, tzl, p);
  GF_LAUNCHED();
  return GF_OK;
}

// out[r, c] = sum_k X[r, k] * Z[c, k] for c < n_valid, extra[r] = column n_valid (optional);
// X hi/lo [R, 64 nb] bf16, Z hi/lo [z_rows >= n_valid (+1), 64 nb] bf16.
extern "C" int gf_ctm_rr(const void* x_hi, const void* x_lo, const void* z_hi, const void* z_lo, int nb, int z_rows, int R,
                         float* out, int ldo, int n_valid, float* extra, void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  GF_ARG_CHECK(nb >= 1 && nb <= 4 && R > 0 && z_rows >= n_valid + (extra != nullptr ? 1 : 0), "shapes");
  const int cols = n_valid + (extra != nullptr ? 1 : 0);
  const int n_cb = (cols + 255) / 256;
  int cbw = ((cols + n_cb - 1) / n_cb + 31) / 32 * 32;  // whole 32-column epilogue chunks
  if (cbw > 256) cbw = 256;
  CUtensorMap txh, txl, tzh, tzl;
  int rc;
  if ((rc = make_tmap_bf16_2d(&txh, x_hi, (uint64_t)nb * 64, (uint64_t)R, (uint64_t)nb * 128, 64, 128))) return rc;
  if ((rc = make_tmap_bf16_2d(&txl, x_lo, (uint64_t)nb * 64, (uint64_t)R, (uint64_t)nb * 128, 64, 128))) return rc;
  if ((rc = make_tmap_bf16_2d(&tzh, z_hi, (uint64_t)nb * 64, (uint64_t)z_rows, (uint64_t)nb * 128, 64, (uint32_t)cbw))) return rc;
  if ((rc = make_tmap_bf16_2d(&tzl, z_lo, (uint64_t)nb * 64, (uint64_t)z_rows, (uint64_t)nb * 128, 64, (uint32_t)cbw))) return rc;
  CtmParams p{};
  p.out = out; p.extra = extra; p.ldo = ldo; p.rows = R; p.n_valid = n_valid; p.nb = nb; p.cbw = cbw;
  p.n_tiles = (R + 127) / 128;
  CUtensorMap tout{};
  p.tma_store = (ldo % 4 == 0) && ((reinterpret_cast<uintptr_t>(out) & 15) == 0);
  if (p.tma_store) {  // fp32 [R, n_valid] with 32 x 32 boxes, 128-byte swizzle; stores clip at the tensor bounds
    PFN_encodeTiled fn = get_encode_fn();
    GF_ARG_CHECK(fn != nullptr, "cuTensorMapEncodeTiled not available");
    cuuint64_t dims[2] = {(cuuint64_t)n_valid, (cuuint64_t)R};
    cuuint64_t strides[1] = {(cuuint64_t)ldo * 4};
    cuuint32_t box[2] = {32, 32};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = fn(&tout, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, out, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                    CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) p.tma_store = 0;
  }
  const size_t z_bytes = 2 * (size_t)nb * cbw * 128, stage = 2 * (size_t)nb * 16384;
  int stages = (int)(((size_t)ctm_max_smem() - 4096 - 32768 - z_bytes) / stage);
  if (stages > 4) stages = 4;
  GF_ARG_CHECK(stages >= 1, "shared memory");
  p.stages = stages;
  const size_t smem = 3072 + z_bytes + stages * stage + 32768 + 256;
  int P = num_sms() / n_cb;
  if (P < 1) P = 1;
  if (P > p.n_tiles) P = p.n_tiles;
  GF_CUDA_CHECK(cudaFuncSetAttribute(ctm_rr_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  ctm_rr_kernel<<<dim3(n_cb, P), kCtmThreads, smem, stream>>>(txh, txl, tzh, tzl, tout, p);
  GF_LAUNCHED();
  return GF_OK;
}

extern "C" int gf_adam_split(float* p, const float* g, float* m, float* v, long long ld, int rows, int cols, float* p_col,
                             const float* g_col, float* m_col, float* v_col, void* out_hi, void* out_lo, int ldo, float lr,
                             float b1, float b2, float eps, float wd, int step, float gscale, int do_update,
                             const long long* step_ctr, int solver, void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  GF_ARG_CHECK(rows > 0 && ldo % 8 == 0 && ldo >= cols + (p_col != nullptr ? 1 : 0), "ldo");
  GF_ARG_CHECK(!do_update || (g != nullptr && m != nullptr && v != nullptr), "gradient / moments");
  GF_ARG_CHECK(!do_update || p_col == nullptr || (g_col != nullptr && m_col != nullptr && v_col != nullptr),
               "gradient / moments of the extra column");
  GF_ARG_CHECK(!do_update || step >= 1 || step_ctr != nullptr, "Adam step must be >= 1");
  AdamS h;
  h.lr = lr; h.b1 = b1; h.b2 = b2; h.eps = eps; h.wd = wd; h.gscale = gscale; h.step_ctr = step_ctr; h.solver = solver;
  const int t = step >= 1 ? step : 1;
  h.bc1 = 1.f - powf(b1, (float)t);
  h.bc2_sqrt = sqrtf(1.f - powf(b2, (float)t));
  const long long n = (long long)rows * (ldo / 8);
  long long blocks = (n + 255) / 256;
  if (blocks > 16LL * num_sms()) blocks = 16LL * num_sms();
  adam_split_kernel<<<(int)blocks, 256, 0, stream>>>(p, g, m, v, ld, rows, cols, p_col, g_col, m_col, v_col,
                                                     reinterpret_cast<__nv_bfloat16*>(out_hi),
                                                     reinterpret_cast<__nv_bfloat16*>(out_lo), ldo, h, do_update);
  GF_LAUNCHED();
  return GF_OK;
}
