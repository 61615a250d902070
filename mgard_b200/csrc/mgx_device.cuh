// Device helpers shared by the mgard_b200 kernels.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

// Explicitly rounded arithmetic: the reference builds with MGARD_X_FMA
// undefined, i.e. every product and sum below is rounded on its own.  Using the
// *_rn intrinsics keeps nvcc from contracting a*b+c into an FMA.
__device__ __forceinline__ float mul_rn(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float add_rn(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float sub_rn(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ float div_rn(float a, float b) { return __fdiv_rn(a, b); }
__device__ __forceinline__ double mul_rn(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ double add_rn(double a, double b) { return __dadd_rn(a, b); }
__device__ __forceinline__ double sub_rn(double a, double b) { return __dsub_rn(a, b); }
__device__ __forceinline__ double div_rn(double a, double b) { return __ddiv_rn(a, b); }

// lerp, reference Coefficient/GPKFunctor.h:13-25 (non-FMA branch):
//   r = v0 + v0 * t * -1;  r = r + t * v1;
// (x * -1 is an exact negation, so v0 + (v0*t)*-1 == v0 - v0*t bit for bit.)
template <typename T> __device__ __forceinline__ T mgx_lerp(T v0, T v1, T t) {
  T r = sub_rn(v0, mul_rn(v0, t));
  return add_rn(r, mul_rn(t, v1));
}

// mass_trans, reference Correction/LPKFunctor.h:13-67 (non-FMA branch)
template <typename T>
__device__ __forceinline__ T mgx_mass_trans(T a, T b, T c, T d, T e, T h1, T h2,
                                            T h3, T h4) {
  T s12 = add_rn(h1, h2), s34 = add_rn(h3, h4), s23 = add_rn(h2, h3);
  T r1 = (s12 != (T)0) ? div_rn(h1, s12) : (T)0;
  T r4 = (s34 != (T)0) ? div_rn(h4, s34) : (T)0;
  T h1_6 = div_rn(h1, (T)6), h2_6 = div_rn(h2, (T)6), h3_6 = div_rn(h3, (T)6),
    h4_6 = div_rn(h4, (T)6);
  T tb = add_rn(add_rn(mul_rn(a, h1_6), mul_rn(b, div_rn(s12, (T)3))),
                mul_rn(c, h2_6));
  T tc = add_rn(add_rn(mul_rn(b, h2_6), mul_rn(c, div_rn(s23, (T)3))),
                mul_rn(d, h3_6));
  T td = add_rn(add_rn(mul_rn(c, h3_6), mul_rn(d, div_rn(s34, (T)3))),
                mul_rn(e, h4_6));
  return add_rn(tc, add_rn(mul_rn(tb, r1), mul_rn(td, r4)));
}

// lwqzk quantise, reference Quantization/LinearQuantization.hpp:96-111:
//   q = (int64) copysign(0.5 + |t * quantizer * volume|, t) + dict/2
__device__ __forceinline__ long long mgx_quantize(float t, float qv, float vol,
                                                  long long shift) {
  float a = fabsf(mul_rn(mul_rn(t, qv), vol));
  float r = add_rn(0.5f, a);
  long long q = (long long)r;
  return (t < 0.f ? -q : q) + shift;
}
__device__ __forceinline__ long long mgx_quantize(double t, double qv, double vol,
                                                  long long shift) {
  double a = fabs(mul_rn(mul_rn(t, qv), vol));
  double r = add_rn(0.5, a);
  long long q = (long long)r;
  return (t < 0. ? -q : q) + shift;
}

// Shared-memory histogram update.  Smooth fields put most of a warp on the
// same symbol (dict/2), so a warp whose valid lanes all agree issues one
// atomicAdd of the population count instead of 32 colliding ones.  Must be
// called by all 32 lanes of the warp.
__device__ __forceinline__ void mgx_hist_add(uint32_t *hist, uint32_t sym,
                                             bool valid) {
  const unsigned mask = __ballot_sync(0xffffffffu, valid);
  if (mask == 0)
    return;
  const int leader = __ffs(mask) - 1;
  const uint32_t lead_sym = __shfl_sync(0xffffffffu, sym, leader);
  const bool uniform = __all_sync(0xffffffffu, !valid || sym == lead_sym);
  if (uniform) {
    if ((int)(threadIdx.x & 31) == leader)
      atomicAdd(&hist[lead_sym], (uint32_t)__popc(mask));
  } else if (valid) {
    atomicAdd(&hist[sym], 1u);
  }
}

static inline unsigned mgx_grid_for(uint64_t total, unsigned block,
                                    unsigned max_blocks) {
  uint64_t g = (total + block - 1) / block;
  if (g < 1)
    g = 1;
  if (g > max_blocks)
    g = max_blocks;
  return (unsigned)g;
}
