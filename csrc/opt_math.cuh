// One parameter update of the supported solvers, shared by every kernel that steps parameters (flat Adam, the beta
// kernels, the fused Adam + shadow refresh and the federated aggregation).  Two state buffers per parameter (m, v)
// hold whatever the solver needs.  Ref: the optimiser factory of the reference, avitm.py:141-153 -- torch.optim
// Adam(betas=(momentum, 0.99)), SGD(momentum), Adagrad, Adadelta, RMSprop(momentum) with torch's defaults.
#pragma once
#include <cuda_runtime.h>

namespace gf {

enum : int { GF_SOLVER_ADAM = 0, GF_SOLVER_SGD = 1, GF_SOLVER_ADAGRAD = 2, GF_SOLVER_ADADELTA = 3, GF_SOLVER_RMSPROP = 4 };

struct OptH {
  float lr, b1, b2, eps, bc1, bc2_sqrt, wd;  // adam: bc1 = 1-b1^t, bc2_sqrt = sqrt(1-b2^t); others: unused
  int solver;
};

//   adam      m = b1 m + (1-b1) g,  v = b2 v + (1-b2) g^2,  p -= lr/bc1 * m / (sqrt(v)/bc2_sqrt + eps)
//   sgd       m = b1 m + g                                   p -= lr m                (b1 = momentum)
//   adagrad   v += g^2                                       p -= lr g / (sqrt(v) + eps)
//   adadelta  v = b2 v + (1-b2) g^2, d = sqrt(m+eps)/sqrt(v+eps) g, m = b2 m + (1-b2) d^2,  p -= lr d   (b2 = rho)
//   rmsprop   v = b2 v + (1-b2) g^2, a = g / (sqrt(v)+eps);  b1 > 0: m = b1 m + a, p -= lr m;  else p -= lr a
__device__ __forceinline__ float opt_update(float p, float g, float& m, float& v, const OptH& h) {
  if (h.wd != 0.f) g = fmaf(h.wd, p, g);
  switch (h.solver) {
    case GF_SOLVER_SGD:
      m = fmaf(h.b1, m, g);
      return p - h.lr * m;
    case GF_SOLVER_ADAGRAD:
      v = fmaf(g, g, v);
      return p - h.lr * g / (sqrtf(v) + h.eps);
    case GF_SOLVER_ADADELTA: {
      v = fmaf(h.b2, v, (1.f - h.b2) * g * g);
      const float d = sqrtf(m + h.eps) / sqrtf(v + h.eps) * g;
      m = fmaf(h.b2, m, (1.f - h.b2) * d * d);
      return p - h.lr * d;
    }
    case GF_SOLVER_RMSPROP: {
      v = fmaf(h.b2, v, (1.f - h.b2) * g * g);
      const float a = g / (sqrtf(v) + h.eps);
      if (h.b1 > 0.f) {
        m = fmaf(h.b1, m, a);
        return p - h.lr * m;
      }
      return p - h.lr * a;
    }
    default: {
      // Adam.  sqrt / divisions through the approximate SFU forms (2 ulp): with IEEE sqrtf and three IEEE divisions
      // the update was ~70 instructions per element, which made the overlapped aggregation loop (8 warps per SM)
      // instruction-bound at ~0.5 element/ns per SM; the relative error of 1e-7 is far below the test tolerances.
      m = fmaf(h.b1, m, (1.f - h.b1) * g);
      v = fmaf(h.b2, v, (1.f - h.b2) * g * g);
      float sq;
      asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(sq) : "f"(v));
      const float denom = fmaf(sq, __fdividef(1.f, h.bc2_sqrt), h.eps);
      return fmaf(-__fdividef(h.lr, h.bc1), __fdividef(m, denom), p);
    }
  }
}

// Adam without weight decay (the usual case) with the per-step constants hoisted: the same operations in the same
// order as the Adam branch of opt_update (bitwise identical results), ~12 instructions per element instead of ~50
// with the solver dispatch and the two constant divisions.  Kernels whose update loop is bound by instruction issue
// (the overlapped aggregation, the background beta step) test `on` once and use this.
struct OptFast {
  float b1, omb1, b2, omb2, inv_bc2, nlr, eps;
  bool on;
};
__device__ __forceinline__ OptFast opt_fast(const OptH& h) {
  OptFast f;
  f.b1 = h.b1; f.omb1 = 1.f - h.b1; f.b2 = h.b2; f.omb2 = 1.f - h.b2; f.eps = h.eps;
  f.inv_bc2 = __fdividef(1.f, h.bc2_sqrt);
  f.nlr = -__fdividef(h.lr, h.bc1);
  f.on = (h.solver == GF_SOLVER_ADAM) && (h.wd == 0.f);
  return f;
}
__device__ __forceinline__ float opt_update_fast(float p, float g, float& m, float& v, const OptFast& f) {
  m = fmaf(f.b1, m, f.omb1 * g);
  v = fmaf(f.b2, v, f.omb2 * g * g);
  float sq;
  asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(sq) : "f"(v));
  return fmaf(f.nlr, __fdividef(m, fmaf(sq, f.inv_bc2, f.eps)), p);
}

}  // namespace gf
