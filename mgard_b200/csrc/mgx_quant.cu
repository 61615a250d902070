// Norm reduction and level-wise linear quantisation kernels (sm_100a).
//
// reference: CompressionLowLevel/NormCalculator.hpp:13-80 (norm),
//            Quantization/LinearQuantization.hpp:19-145 (lwqzk kernel),
//            Quantization/LinearQuantization.hpp:234-297 (CalcQuantizers, host).
#include <algorithm>
#include <cmath>
#include <limits>

#include <type_traits>

#include "mgx_device.cuh"
#include "mgx_internal.h"

namespace {

// ---- norm -------------------------------------------------------------------
// One pass over the input: max|u| (L-inf) or sum u^2 accumulated in fp64.  The
// partial lands in d_scalars[MGX_SC_NORM] (bit pattern of a double; for the max
// the non-negative double ordering equals the uint64 ordering).
template <typename T, bool LINF>
__global__ void __launch_bounds__(512)
    k_norm(const T *__restrict__ v, uint64_t n, unsigned long long *out,
           double *__restrict__ block_part) {
  // 16-byte loads, four in flight per thread; |max| of floats is exact in T,
  // sums of squares accumulate in double
  constexpr int VEC = 16 / sizeof(T);
  typedef typename std::conditional<sizeof(T) == 4, float4, double2>::type V;
  double acc = 0.0;
  T tmax = 0;
  const uint64_t nvec = (((uintptr_t)v & 15) == 0) ? n / VEC : 0;
  const V *vv = reinterpret_cast<const V *>(v);
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
  auto take = [&](const V &q) {
    const T *e = reinterpret_cast<const T *>(&q);
#pragma unroll
    for (int k = 0; k < VEC; k++) {
      if (LINF)
        tmax = fmax(tmax, fabs(e[k]));
      else
        acc += (double)e[k] * (double)e[k];
    }
  };
  for (; i + 3 * stride < nvec; i += 4 * stride) {
    const V a = vv[i], b = vv[i + stride], c = vv[i + 2 * stride],
            d = vv[i + 3 * stride];
    take(a);
    take(b);
    take(c);
    take(d);
  }
  for (; i < nvec; i += stride)
    take(vv[i]);
  for (uint64_t j = nvec * VEC + blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
       j < n; j += stride) {
    const T x = v[j];
    if (LINF)
      tmax = fmax(tmax, fabs(x));
    else
      acc += (double)x * (double)x;
  }
  if (LINF)
    acc = (double)tmax;
  for (int o = 16; o > 0; o >>= 1) {
    double other = __shfl_xor_sync(0xffffffffu, acc, o);
    acc = LINF ? fmax(acc, other) : acc + other;
  }
  __shared__ double s[16];
  int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0)
    s[w] = acc;
  __syncthreads();
  if (w == 0) {
    acc = lane < (int)(blockDim.x >> 5) ? s[lane] : 0.0;
    for (int o = 8; o > 0; o >>= 1) {
      double other = __shfl_xor_sync(0xffffffffu, acc, o);
      acc = LINF ? fmax(acc, other) : acc + other;
    }
    if (lane == 0) {
      if (LINF)
        atomicMax(out, (unsigned long long)__double_as_longlong(acc));
      else
        block_part[blockIdx.x] = acc; // summed in a fixed order by k_norm_final

    }
  }
}

// sum of the per-block partial sums in a fixed order: the L2 / s-norm comes out
// bit-identical from run to run (atomicAdd of doubles would not)
__global__ void __launch_bounds__(1024)
    k_norm_final(const double *__restrict__ part, int n, double *out) {
  __shared__ double s[1024];
  s[threadIdx.x] = (int)threadIdx.x < n ? part[threadIdx.x] : 0.0;
  __syncthreads();
  for (int o = 512; o > 0; o >>= 1) {
    if ((int)threadIdx.x < o)
      s[threadIdx.x] += s[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0)
    *out = s[0];
}

// ---- quantiser tables ---------------------------------------------------------
template <typename T> struct QuantArgs {
  T q[MGX_MAX_L];   // per-level quantiser (reciprocal for quantise)
  T vol[MGX_MAX_L]; // per-level volume factor (1 when s == inf)
  uint32_t n[MGX_MAX_L][MGX_MAX_D]; // level shapes (level marks)
  uint32_t shape[MGX_MAX_D];
  int L;
  int calc_vol;
  long long shift; // dict_size / 2
  uint32_t dict;
  uint64_t total;
};

// LinearQuantizer::CalcQuantizers, LinearQuantization.hpp:234-297
// (decomposition_type::MultiDim); tol / s / norm arrive converted to T exactly
// as Compressor::Quantize passes them (Compressor.hpp:139-146).
template <typename T>
void fill_quant_args(const mgx_ctx *ctx, QuantArgs<T> *a, int rel, double tol_d,
                     double s_d, double norm_d, bool reciprocal) {
  const MgxLevels &lv = ctx->lv;
  const MgxHostTables<T> &tb =
      *reinterpret_cast<const MgxHostTables<T> *>(
          mgx_is_f32(ctx) ? (const void *)&ctx->tbf : (const void *)&ctx->tbd);
  const int L = lv.L, D = lv.D;
  T tol = (T)tol_d, norm = (T)norm_d;
  T s = std::isinf(s_d) ? std::numeric_limits<T>::infinity() : (T)s_d;
  bool linf = std::isinf(s_d);
  bool orth = !(linf && D <= 3); // infer_orthogonal_projection, Compressor.h:38-40
  double abs_tol = tol;
  if (rel)
    abs_tol *= norm;
  abs_tol *= 2;
  size_t dof = lv.N;
  for (int l = 0; l < L + 1; l++) {
    T qv;
    if (linf) {
      if (!orth)
        qv = (T)((abs_tol) / (uint64_t)(L + 1));
      else
        qv = (T)((abs_tol) / ((uint64_t)(L + 1) * (1 + std::pow(3, D))));
    } else {
      qv = (T)((abs_tol) / (std::exp2(s * l) * std::sqrt((double)dof)));
    }
    if (reciprocal)
      qv = 1.0f / qv;
    a->q[l] = qv;
    a->vol[l] = linf ? (T)1 : (reciprocal ? tb.vol[l] : tb.volr[l]);
  }
  for (int l = 0; l <= L; l++)
    for (int d = 0; d < MGX_MAX_D; d++)
      a->n[l][d] = (uint32_t)lv.n[l][d];
  for (int d = 0; d < MGX_MAX_D; d++)
    a->shape[d] = (uint32_t)lv.shape[d];
  a->L = L;
  a->calc_vol = !linf;
  a->shift = (long long)(ctx->dict / 2);
  a->dict = ctx->dict;
  a->total = lv.N;
}

// level of the node at linear position it: max_d level_marks[d][p_d]
// (Hierarchy.hpp:340-360, LinearQuantization.hpp:74-79)
template <typename T>
__device__ __forceinline__ int node_level(const QuantArgs<T> &a, uint64_t it) {
  int level = 0;
  uint64_t r = it;
#pragma unroll
  for (int d = MGX_MAX_D - 1; d >= 0; d--) {
    uint32_t p = (uint32_t)(r % a.shape[d]);
    r /= a.shape[d];
    int m = 0;
    while (m < a.L && p >= a.n[m][d])
      m++;
    level = max(level, m);
  }
  return level;
}

template <typename T>
__global__ void __launch_bounds__(256)
    k_quantize(const T *__restrict__ c, long long *__restrict__ codes,
               const __grid_constant__ QuantArgs<T> a) {
  for (uint64_t it = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
       it < a.total; it += (uint64_t)gridDim.x * blockDim.x) {
    int level = a.calc_vol ? node_level(a, it) : 0;
    codes[it] = mgx_quantize(c[it], a.q[level], a.vol[level], a.shift);
  }
}

template <typename T>
__global__ void __launch_bounds__(256)
    k_dequantize(const long long *__restrict__ codes, T *__restrict__ c,
                 const __grid_constant__ QuantArgs<T> a) {
  for (uint64_t it = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
       it < a.total; it += (uint64_t)gridDim.x * blockDim.x) {
    int level = a.calc_vol ? node_level(a, it) : 0;
    long long q = codes[it] - a.shift;
    c[it] = mul_rn(mul_rn(a.q[level], a.vol[level]), (T)q);
  }
}

// quantise straight to 16-bit dictionary symbols + shared-memory histogram +
// outlier list (fuses lwqzk, OutlierSeparator.hpp:17-120 and Histogram.hpp)
template <typename T>
__global__ void __launch_bounds__(256)
    k_quantize_sym(const T *__restrict__ c, uint16_t *__restrict__ sym,
                   uint32_t *__restrict__ g_freq,
                   unsigned long long *__restrict__ out_count,
                   unsigned long long *__restrict__ out_idx,
                   long long *__restrict__ out_val, unsigned long long out_cap,
                   const __grid_constant__ QuantArgs<T> a) {
  extern __shared__ uint32_t s_hist[];
  for (uint32_t i = threadIdx.x; i < a.dict; i += blockDim.x)
    s_hist[i] = 0;
  __syncthreads();
  const uint64_t step = (uint64_t)gridDim.x * blockDim.x;
  const uint64_t rounds = (a.total + step - 1) / step;
  for (uint64_t k = 0; k < rounds; k++) {
    uint64_t it = k * step + blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
    bool valid = it < a.total;
    uint32_t sv = 0;
    if (valid) {
      int level = a.calc_vol ? node_level(a, it) : 0;
      long long q = mgx_quantize(c[it], a.q[level], a.vol[level], a.shift);
      if (q < 0 || q >= (long long)a.dict) {
        unsigned long long slot = atomicAdd(out_count, 1ull);
        if (slot < out_cap) {
          out_idx[slot] = it;
          out_val[slot] = q;
        }
        q = 0;
      }
      sv = (uint32_t)q;
      sym[it] = (uint16_t)sv;
    }
    mgx_hist_add(s_hist, sv, valid);
  }
  __syncthreads();
  for (uint32_t i = threadIdx.x; i < a.dict; i += blockDim.x) {
    uint32_t v = s_hist[i];
    if (v)
      atomicAdd(&g_freq[i], v);
  }
}

template <typename T>
__global__ void __launch_bounds__(256)
    k_dequantize_sym(const uint16_t *__restrict__ sym, T *__restrict__ c,
                     const __grid_constant__ QuantArgs<T> a) {
  for (uint64_t it = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
       it < a.total; it += (uint64_t)gridDim.x * blockDim.x) {
    int level = a.calc_vol ? node_level(a, it) : 0;
    long long q = (long long)sym[it] - a.shift;
    c[it] = mul_rn(mul_rn(a.q[level], a.vol[level]), (T)q);
  }
}

// Row-wise variant: one block per run of rows of the contiguous dimension; the
// level of the outer coordinates is found once per row, the level of the
// column by descending the level shapes (half the columns stop at the first
// test), so there is no per-element division.
template <typename T>
__global__ void __launch_bounds__(256)
    k_dequantize_sym_rows(const uint16_t *__restrict__ sym, T *__restrict__ c,
                          const __grid_constant__ QuantArgs<T> a, uint64_t rows) {
  const uint32_t nf = a.shape[MGX_MAX_D - 1];
  for (uint64_t row = blockIdx.x; row < rows; row += gridDim.x) {
    int row_level = 0;
    if (a.calc_vol) {
      uint64_t r = row;
#pragma unroll
      for (int d = MGX_MAX_D - 2; d >= 0; d--) {
        uint32_t p = (uint32_t)(r % a.shape[d]);
        r /= a.shape[d];
        int m = 0;
        while (m < a.L && p >= a.n[m][d])
          m++;
        row_level = max(row_level, m);
      }
    }
    const uint16_t *srow = sym + row * nf;
    T *crow = c + row * nf;
    for (uint32_t x = threadIdx.x; x < nf; x += blockDim.x) {
      int level = 0;
      if (a.calc_vol) {
        int m = a.L;
        while (m > 0 && x < a.n[m - 1][MGX_MAX_D - 1])
          m--;
        level = max(row_level, m);
      }
      long long q = (long long)srow[x] - a.shift;
      crow[x] = mul_rn(mul_rn(a.q[level], a.vol[level]), (T)q);
    }
  }
}

// outliers carry the true code: overwrite the dequantised value
template <typename T>
__global__ void __launch_bounds__(256)
    k_dequantize_outliers(const unsigned long long *__restrict__ idx,
                          const long long *__restrict__ val, uint64_t n_out,
                          T *__restrict__ c,
                          const __grid_constant__ QuantArgs<T> a) {
  for (uint64_t j = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; j < n_out;
       j += (uint64_t)gridDim.x * blockDim.x) {
    uint64_t it = idx[j];
    int level = a.calc_vol ? node_level(a, it) : 0;
    long long q = val[j] - a.shift;
    c[it] = mul_rn(mul_rn(a.q[level], a.vol[level]), (T)q);
  }
}

// level marks regardless of the norm type (node_level() skips them for L-inf)
template <typename T>
__device__ __forceinline__ int node_level_always(const QuantArgs<T> &a, uint64_t it) {
  int level = 0;
  uint64_t r = it;
#pragma unroll
  for (int d = MGX_MAX_D - 1; d >= 0; d--) {
    uint32_t p = (uint32_t)(r % a.shape[d]);
    r /= a.shape[d];
    int m = 0;
    while (m < a.L && p >= a.n[m][d])
      m++;
    level = max(level, m);
  }
  return level;
}

template <typename T>
__global__ void __launch_bounds__(256)
    k_outlier_max_level(const unsigned long long *__restrict__ idx, uint64_t n_out,
                        const __grid_constant__ QuantArgs<T> a,
                        unsigned long long *out) {
  int lv = 0;
  for (uint64_t j = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; j < n_out;
       j += (uint64_t)gridDim.x * blockDim.x)
    lv = max(lv, node_level_always(a, idx[j]));
  for (int o = 16; o > 0; o >>= 1)
    lv = max(lv, __shfl_xor_sync(0xffffffffu, lv, o));
  if ((threadIdx.x & 31) == 0 && lv > 0)
    atomicMax(out, (unsigned long long)lv);
}

// dequantise the box [0, n[level][d]) of the reordered array only
template <typename T>
__global__ void __launch_bounds__(256)
    k_dequantize_sym_box(const uint16_t *__restrict__ sym, T *__restrict__ c,
                         const __grid_constant__ QuantArgs<T> a, int level,
                         uint64_t total) {
  for (uint64_t it = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; it < total;
       it += (uint64_t)gridDim.x * blockDim.x) {
    uint64_t r = it, o = 0, st = 1;
    int lvl = 0;
#pragma unroll
    for (int d = MGX_MAX_D - 1; d >= 0; d--) {
      const uint32_t p = (uint32_t)(r % a.n[level][d]);
      r /= a.n[level][d];
      o += (uint64_t)p * st;
      st *= a.shape[d];
      int m = 0;
      while (m < a.L && p >= a.n[m][d])
        m++;
      lvl = max(lvl, m);
    }
    if (!a.calc_vol)
      lvl = 0;
    long long q = (long long)sym[o] - a.shift;
    c[o] = mul_rn(mul_rn(a.q[lvl], a.vol[lvl]), (T)q);
  }
}

} // namespace

void mgx_level_quantizers(const mgx_ctx *ctx, int rel, double tol, double s,
                          double norm, bool reciprocal, double *q, double *vol) {
  if (mgx_is_f32(ctx)) {
    QuantArgs<float> a;
    fill_quant_args<float>(ctx, &a, rel, tol, s, norm, reciprocal);
    for (int l = 0; l <= ctx->lv.L; l++) {
      q[l] = a.q[l];
      vol[l] = a.vol[l];
    }
  } else {
    QuantArgs<double> a;
    fill_quant_args<double>(ctx, &a, rel, tol, s, norm, reciprocal);
    for (int l = 0; l <= ctx->lv.L; l++) {
      q[l] = a.q[l];
      vol[l] = a.vol[l];
    }
  }
}

int mgxk_norm_partial(mgx_ctx *ctx, const void *d_in, int linf) {
  unsigned long long *out =
      (unsigned long long *)(ctx->d_scalars + MGX_SC_NORM);
  MGX_CUDA(cudaMemsetAsync(out, 0, 8, ctx->stream));
  uint64_t n = ctx->lv.N;
  unsigned grid = mgx_grid_for(n, 512 * 8, std::min(ctx->num_sms * 4, 1024));
  double *part = ctx->d_norm_part;
  if (mgx_is_f32(ctx)) {
    if (linf)
      k_norm<float, true><<<grid, 512, 0, ctx->stream>>>((const float *)d_in, n, out,
                                                        part);
    else
      k_norm<float, false><<<grid, 512, 0, ctx->stream>>>((const float *)d_in, n, out,
                                                         part);
  } else {
    if (linf)
      k_norm<double, true><<<grid, 512, 0, ctx->stream>>>((const double *)d_in, n, out,
                                                         part);
    else
      k_norm<double, false><<<grid, 512, 0, ctx->stream>>>((const double *)d_in, n,
                                                          out, part);
  }
  ctx->launches++;
  if (!linf) {
    k_norm_final<<<1, 1024, 0, ctx->stream>>>(part, (int)grid, (double *)out);
    ctx->launches++;
  }
  MGX_CUDA(cudaGetLastError());
  return MGX_SUCCESS;
}

template <typename T>
static int quantize_t(mgx_ctx *ctx, const void *d_coeff, int rel, double tol,
                      double s, double norm, int64_t *d_codes) {
  QuantArgs<T> a;
  fill_quant_args<T>(ctx, &a, rel, tol, s, norm, true);
  k_quantize<T><<<mgx_grid_for(a.total, 256, ctx->num_sms * 16), 256, 0,
                  ctx->stream>>>((const T *)d_coeff, (long long *)d_codes, a);
  ctx->launches++;
  MGX_CUDA(cudaGetLastError());
  return MGX_SUCCESS;
}

template <typename T>
static int dequantize_t(mgx_ctx *ctx, const int64_t *d_codes, int rel,
                        double tol, double s, double norm, void *d_coeff) {
  QuantArgs<T> a;
  fill_quant_args<T>(ctx, &a, rel, tol, s, norm, false);
  k_dequantize<T><<<mgx_grid_for(a.total, 256, ctx->num_sms * 16), 256, 0,
                    ctx->stream>>>((const long long *)d_codes, (T *)d_coeff, a);
  ctx->launches++;
  MGX_CUDA(cudaGetLastError());
  return MGX_SUCCESS;
}

int mgxk_quantize(mgx_ctx *ctx, const void *d_coeff, int rel, double tol,
                  double s, double norm, int64_t *d_codes) {
  return mgx_is_f32(ctx)
             ? quantize_t<float>(ctx, d_coeff, rel, tol, s, norm, d_codes)
             : quantize_t<double>(ctx, d_coeff, rel, tol, s, norm, d_codes);
}

int mgxk_dequantize(mgx_ctx *ctx, const int64_t *d_codes, int rel, double tol,
                    double s, double norm, void *d_coeff) {
  return mgx_is_f32(ctx)
             ? dequantize_t<float>(ctx, d_codes, rel, tol, s, norm, d_coeff)
             : dequantize_t<double>(ctx, d_codes, rel, tol, s, norm, d_coeff);
}

template <typename T>
static int quantize_sym_t(mgx_ctx *ctx, const void *d_coeff, int rel,
                          double tol, double s, double norm) {
  QuantArgs<T> a;
  fill_quant_args<T>(ctx, &a, rel, tol, s, norm, true);
  size_t smem = (size_t)ctx->dict * 4;
  MGX_CUDA(cudaFuncSetAttribute(k_quantize_sym<T>,
                                cudaFuncAttributeMaxDynamicSharedMemorySize,
                                (int)smem));
  unsigned grid = mgx_grid_for(a.total, 256 * 16, ctx->num_sms * 4);
  k_quantize_sym<T><<<grid, 256, smem, ctx->stream>>>(
      (const T *)d_coeff, ctx->d_sym, ctx->d_freq,
      (unsigned long long *)(ctx->d_scalars + MGX_SC_OUTLIERS),
      (unsigned long long *)ctx->d_out_idx, (long long *)ctx->d_out_val,
      ctx->out_cap, a);
  ctx->launches++;
  MGX_CUDA(cudaGetLastError());
  return MGX_SUCCESS;
}

int mgxk_quantize_sym(mgx_ctx *ctx, const void *d_coeff, int rel, double tol,
                      double s, const double *, double norm) {
  return mgx_is_f32(ctx) ? quantize_sym_t<float>(ctx, d_coeff, rel, tol, s, norm)
                         : quantize_sym_t<double>(ctx, d_coeff, rel, tol, s, norm);
}

template <typename T>
static int dequantize_sym_t(mgx_ctx *ctx, int rel, double tol, double s,
                            double norm, void *d_coeff,
                            const uint64_t *d_oidx, const int64_t *d_oval,
                            uint64_t n_out) {
  QuantArgs<T> a;
  fill_quant_args<T>(ctx, &a, rel, tol, s, norm, false);
  const uint64_t nf = ctx->lv.shape[MGX_MAX_D - 1];
  if (nf >= 128) {
    const uint64_t rows = a.total / nf;
    k_dequantize_sym_rows<T><<<(unsigned)std::min<uint64_t>(rows, ctx->num_sms * 16),
                               256, 0, ctx->stream>>>(ctx->d_sym, (T *)d_coeff, a,
                                                      rows);
  } else {
    k_dequantize_sym<T><<<mgx_grid_for(a.total, 256, ctx->num_sms * 16), 256, 0,
                          ctx->stream>>>(ctx->d_sym, (T *)d_coeff, a);
  }
  ctx->launches++;
  if (n_out) {
    k_dequantize_outliers<T><<<mgx_grid_for(n_out, 256, ctx->num_sms * 4), 256,
                               0, ctx->stream>>>(
        (const unsigned long long *)d_oidx, (const long long *)d_oval, n_out,
        (T *)d_coeff, a);
    ctx->launches++;
  }
  MGX_CUDA(cudaGetLastError());
  return MGX_SUCCESS;
}

int mgxk_dequantize_sym_out(mgx_ctx *ctx, int rel, double tol, double s,
                            double norm, void *d_coeff, const uint64_t *d_oidx,
                            const int64_t *d_oval, uint64_t n_out) {
  return mgx_is_f32(ctx) ? dequantize_sym_t<float>(ctx, rel, tol, s, norm, d_coeff,
                                                   d_oidx, d_oval, n_out)
                         : dequantize_sym_t<double>(ctx, rel, tol, s, norm,
                                                    d_coeff, d_oidx, d_oval, n_out);
}

template <typename T>
static int outlier_max_level_t(mgx_ctx *ctx, const uint64_t *d_oidx, uint64_t n_out,
                               int *max_level) {
  QuantArgs<T> a;
  fill_quant_args<T>(ctx, &a, 0, 1.0, 0.0, 1.0, false);
  unsigned long long *out = (unsigned long long *)(ctx->d_scalars + MGX_SC_NZ);
  MGX_CUDA(cudaMemsetAsync(out, 0, 8, ctx->stream));
  k_outlier_max_level<T><<<mgx_grid_for(n_out, 256, ctx->num_sms * 4), 256, 0,
                           ctx->stream>>>((const unsigned long long *)d_oidx, n_out,
                                          a, out);
  ctx->launches++;
  MGX_CUDA(cudaMemcpyAsync(&ctx->h_scalars[MGX_SC_NZ], out, 8, cudaMemcpyDeviceToHost,
                           ctx->stream));
  MGX_CUDA(cudaStreamSynchronize(ctx->stream));
  *max_level = (int)ctx->h_scalars[MGX_SC_NZ];
  return MGX_SUCCESS;
}

int mgxk_outlier_max_level(mgx_ctx *ctx, const uint64_t *d_oidx, uint64_t n_out,
                           int *max_level) {
  return mgx_is_f32(ctx) ? outlier_max_level_t<float>(ctx, d_oidx, n_out, max_level)
                         : outlier_max_level_t<double>(ctx, d_oidx, n_out, max_level);
}

template <typename T>
static int dequantize_box_t(mgx_ctx *ctx, int level, int rel, double tol, double s,
                            double norm, void *d_coeff, const uint64_t *d_oidx,
                            const int64_t *d_oval, uint64_t n_out) {
  QuantArgs<T> a;
  fill_quant_args<T>(ctx, &a, rel, tol, s, norm, false);
  uint64_t total = 1;
  for (int d = 0; d < MGX_MAX_D; d++)
    total *= ctx->lv.n[level][d];
  k_dequantize_sym_box<T><<<mgx_grid_for(total, 256, ctx->num_sms * 8), 256, 0,
                            ctx->stream>>>(ctx->d_sym, (T *)d_coeff, a, level, total);
  ctx->launches++;
  if (n_out) {
    k_dequantize_outliers<T><<<mgx_grid_for(n_out, 256, ctx->num_sms * 4), 256, 0,
                               ctx->stream>>>((const unsigned long long *)d_oidx,
                                              (const long long *)d_oval, n_out,
                                              (T *)d_coeff, a);
    ctx->launches++;
  }
  MGX_CUDA(cudaGetLastError());
  return MGX_SUCCESS;
}

int mgxk_dequantize_box(mgx_ctx *ctx, int level, int rel, double tol, double s,
                        double norm, void *d_coeff, const uint64_t *d_oidx,
                        const int64_t *d_oval, uint64_t n_out) {
  return mgx_is_f32(ctx) ? dequantize_box_t<float>(ctx, level, rel, tol, s, norm,
                                                   d_coeff, d_oidx, d_oval, n_out)
                         : dequantize_box_t<double>(ctx, level, rel, tol, s, norm,
                                                    d_coeff, d_oidx, d_oval, n_out);
}
