// Multilevel decomposition / recomposition kernels (sm_100a).
//
// Two families:
//  (1) level kernels, any D <= 5, with or without the orthogonal correction:
//      coefficients+reorder, mass-matrix x restriction sweeps, batched Thomas
//      solves, coarse add/subtract, coefficient restore.  They implement what
//      the reference spreads over GpkReo/GpkRev, Lpk1-3, Ipk1-3 and Lwpk
//      (include/mgard-x/DataRefactoring/MultiDimension/**) but are written as
//      flat, dimension-generic gather kernels rather than as ports of those
//      functors.
//  (2) the fused hierarchical path (mgx_hier.cu) for the L-inf, D <= 3 case.
//
// Rounding: every floating-point expression uses explicit *_rn intrinsics in
// the reference's operation order (MGARD_X_FMA undefined => no contraction),
// so coefficients are bit-identical to the SERIAL oracle.
#include <algorithm>

#include "mgx_device.cuh"
#include "mgx_internal.h"

namespace {

struct LevelArgs {
  uint32_t nf[MGX_MAX_D];  // fine (level l) shape
  uint32_t nc[MGX_MAX_D];  // coarse (level l-1) shape
  uint64_t st[MGX_MAX_D];  // strides of the full array
  uint32_t roff[MGX_MAX_D]; // offset of ratio[l][d] in the table buffer
  uint64_t total;          // prod nf
  int half;                // all ratios are exactly 0.5
};

// position p (reordered layout) -> interleaved index q, odd flag
__device__ __forceinline__ void pos_to_node(uint32_t p, uint32_t nf, uint32_t nc,
                                            uint32_t &q, bool &odd) {
  if (p < nc) {
    q = (p == nc - 1) ? nf - 1 : 2 * p;
    odd = false;
  } else {
    q = 2 * (p - nc) + 1;
    odd = true;
  }
}

// One level of coefficients: src = interleaved level box, dst = reordered.
// (reference GpkReo3D / GpkReo<D>, GridProcessingKernel3D.hpp:20-1228)
template <typename T>
__global__ void __launch_bounds__(256)
    k_level_coeff(const T *__restrict__ src, T *__restrict__ dst,
                  const T *__restrict__ tables, LevelArgs a) {
  for (uint64_t it = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
       it < a.total; it += (uint64_t)gridDim.x * blockDim.x) {
    uint64_t r = it;
    uint32_t q[MGX_MAX_D];
    uint64_t so = 0, dof = 0;
    int dims[MGX_MAX_D], nodd = 0;
#pragma unroll
    for (int d = MGX_MAX_D - 1; d >= 0; d--) {
      uint32_t p = (uint32_t)(r % a.nf[d]);
      r /= a.nf[d];
      bool odd;
      pos_to_node(p, a.nf[d], a.nc[d], q[d], odd);
      dof += (uint64_t)p * a.st[d];
      so += (uint64_t)q[d] * a.st[d];
      if (odd)
        dims[nodd++] = d; // dims[0] = innermost odd dim
    }
    T val = src[so];
    if (nodd > 0) {
      T buf[1 << MGX_MAX_D];
      int cnt = 1 << nodd;
      // bit j of m selects lo/hi along dims[j] (bit 0 = innermost)
      for (int m = 0; m < cnt; m++) {
        uint64_t o = so;
        for (int j = 0; j < nodd; j++)
          o = ((m >> j) & 1) ? o + a.st[dims[j]] : o - a.st[dims[j]];
        buf[m] = src[o];
      }
      for (int j = 0; j < nodd; j++) { // innermost dimension first
        T t = a.half ? (T)0.5
                     : tables[a.roff[dims[j]] + q[dims[j]] - 1];
        cnt >>= 1;
        for (int m = 0; m < cnt; m++)
          buf[m] = mgx_lerp(buf[2 * m], buf[2 * m + 1], t);
      }
      val = sub_rn(val, buf[0]);
    }
    dst[dof] = val;
  }
}

// Inverse: src = reordered level box (coarse values final), dst = interleaved.
// (reference GpkRev3D, GridProcessingKernel3D.hpp:1230-2412)
template <typename T>
__global__ void __launch_bounds__(256)
    k_level_restore(const T *__restrict__ src, T *__restrict__ dst,
                    const T *__restrict__ tables, LevelArgs a) {
  for (uint64_t it = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
       it < a.total; it += (uint64_t)gridDim.x * blockDim.x) {
    uint64_t r = it;
    uint32_t q[MGX_MAX_D];
    uint64_t so = 0, dof = 0;
    int dims[MGX_MAX_D], nodd = 0;
    // reordered offset of the node with every odd dim moved to its lower
    // coarse neighbour: (q-1)/2 ; the upper neighbour is one position further
    uint64_t base = 0;
#pragma unroll
    for (int d = MGX_MAX_D - 1; d >= 0; d--) {
      uint32_t p = (uint32_t)(r % a.nf[d]);
      r /= a.nf[d];
      bool odd;
      pos_to_node(p, a.nf[d], a.nc[d], q[d], odd);
      so += (uint64_t)p * a.st[d];
      dof += (uint64_t)q[d] * a.st[d];
      if (odd) {
        dims[nodd++] = d;
        base += (uint64_t)((q[d] - 1) / 2) * a.st[d];
      } else {
        base += (uint64_t)p * a.st[d];
      }
    }
    T val = src[so];
    if (nodd > 0) {
      T buf[1 << MGX_MAX_D];
      int cnt = 1 << nodd;
      for (int m = 0; m < cnt; m++) {
        uint64_t o = base;
        for (int j = 0; j < nodd; j++)
          if ((m >> j) & 1) {
            // upper neighbour q+1: position (q+1)/2, which never is the
            // kept last node of an even-sized level (q+1 <= nf-2)
            o += a.st[dims[j]];
          }
        buf[m] = src[o];
      }
      for (int j = 0; j < nodd; j++) {
        T t = a.half ? (T)0.5
                     : tables[a.roff[dims[j]] + q[dims[j]] - 1];
        cnt >>= 1;
        for (int m = 0; m < cnt; m++)
          buf[m] = mgx_lerp(buf[2 * m], buf[2 * m + 1], t);
      }
      val = add_rn(val, buf[0]);
    }
    dst[dof] = val;
  }
}

// dst[box] = src[box] over a level box with the full-array strides.
template <typename T>
__global__ void __launch_bounds__(256)
    k_copy_box(const T *__restrict__ src, T *__restrict__ dst, LevelArgs a) {
  for (uint64_t it = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
       it < a.total; it += (uint64_t)gridDim.x * blockDim.x) {
    uint64_t r = it, o = 0;
#pragma unroll
    for (int d = MGX_MAX_D - 1; d >= 0; d--) {
      o += (r % a.nf[d]) * a.st[d];
      r /= a.nf[d];
    }
    dst[o] = src[o];
  }
}


// ---- correction: mass-matrix multiply fused with restriction ---------------
// One sweep along dimension `dim` (reference Lpk1/2/3Reo,
// LinearProcessingKernel3D.hpp:27-1097 and LPKFunctor.h:62-67).  Input is a
// box of shape cur[] addressed with in_st[]; along `dim` it is in the
// reordered layout (nc coarse values then nf-nc coefficients).  Output is
// compact with cur[dim] replaced by nc.  When zero_box is set the all-coarse
// corner [0,zc[d]) of the input reads as 0 (the "coefficient-only" field,
// LinearProcessingKernel3D.hpp:75-81).
struct LpkArgs {
  uint32_t cur[MGX_MAX_D];
  uint64_t in_st[MGX_MAX_D];
  uint32_t zc[MGX_MAX_D];
  int zero_box;
  int dim;
  uint32_t nf, nc;
  uint32_t dist_off;
  uint64_t total; // prod(cur) with cur[dim] = nc
};

template <typename T>
__global__ void __launch_bounds__(256)
    k_lpk(const T *__restrict__ in, T *__restrict__ out,
          const T *__restrict__ tables, LpkArgs a) {
  const T *dist = tables + a.dist_off;
  for (uint64_t it = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
       it < a.total; it += (uint64_t)gridDim.x * blockDim.x) {
    uint64_t r = it, base = 0;
    uint32_t i = 0;
    bool others_coarse = true;
#pragma unroll
    for (int d = MGX_MAX_D - 1; d >= 0; d--) {
      uint32_t n = (d == a.dim) ? a.nc : a.cur[d];
      uint32_t p = (uint32_t)(r % n);
      r /= n;
      if (d == a.dim) {
        i = p;
      } else {
        base += (uint64_t)p * a.in_st[d];
        if (p >= a.zc[d])
          others_coarse = false;
      }
    }
    const uint64_t sd = a.in_st[a.dim];
    const int64_t k = 2 * (int64_t)i;
    const int64_t wmax = 2 * ((int64_t)a.nc - 1);
    const int64_t ncoef = (int64_t)a.nf - (int64_t)a.nc;
    T w[5], h[4];
#pragma unroll
    for (int j = 0; j < 5; j++) {
      int64_t idx = k - 2 + j;
      T v = 0;
      if (idx >= 0 && idx <= wmax) {
        if ((idx & 1) == 0) {
          uint32_t pos = (uint32_t)(idx >> 1);
          bool z = a.zero_box && others_coarse && pos < a.zc[a.dim];
          v = z ? (T)0 : in[base + pos * sd];
        } else {
          int64_t jj = (idx - 1) >> 1;
          if (jj < ncoef)
            v = in[base + (uint64_t)(a.nc + jj) * sd];
        }
      }
      w[j] = v;
    }
#pragma unroll
    for (int j = 0; j < 4; j++) {
      int64_t idx = k - 2 + j;
      h[j] = (idx >= 0 && idx < (int64_t)a.nf) ? dist[idx] : (T)0;
    }
    out[it] = mgx_mass_trans(w[0], w[1], w[2], w[3], w[4], h[0], h[1], h[2],
                             h[3]);
  }
}

// ---- correction: batched Thomas solves --------------------------------------
// reference Ipk1/2/3Reo (IterativeProcessingKernel3D.hpp:28-1169) with
// tridiag_forward2 / tridiag_backward2 (IPKFunctor.h:13-50):
//   forward : x[i] = x[i] - x[i-1] * (am[i] / bm[i])
//   backward: x[i] = (x[i] - am[i+1] * x[i+1]) / bm[i+1]
// Strided variant: the solve dimension is not the contiguous one, consecutive
// threads own consecutive `inner` indices (coalesced), the recurrence value
// stays in a register.
template <typename T>
__global__ void __launch_bounds__(128)
    k_ipk_strided(T *__restrict__ x, const T *__restrict__ am,
                  const T *__restrict__ bm, uint32_t n, uint64_t inner,
                  uint64_t lines) {
  for (uint64_t ln = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
       ln < lines; ln += (uint64_t)gridDim.x * blockDim.x) {
    uint64_t o = ln / inner, in = ln % inner;
    T *p = x + o * n * inner + in;
    T prev = 0;
    for (uint32_t i = 0; i < n; i++) {
      T curr = p[(uint64_t)i * inner];
      prev = sub_rn(curr, mul_rn(prev, div_rn(am[i], bm[i])));
      p[(uint64_t)i * inner] = prev;
    }
    prev = 0;
    for (uint32_t j = 0; j < n; j++) {
      uint32_t i = n - 1 - j;
      T curr = p[(uint64_t)i * inner];
      prev = div_rn(sub_rn(curr, mul_rn(am[n - j], prev)), bm[n - j]);
      p[(uint64_t)i * inner] = prev;
    }
  }
}

// Contiguous variant: the solve runs along the fastest dimension.  One warp
// owns 32 lines; 32x32 tiles are moved with coalesced row accesses and
// transposed through shared memory so that lane r walks line r.
template <typename T>
__global__ void __launch_bounds__(128)
    k_ipk_contig(T *__restrict__ x, const T *__restrict__ am,
                 const T *__restrict__ bm, uint32_t n, uint64_t lines) {
  __shared__ T tile[4][32][33];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  T(*t)[33] = tile[warp];
  const uint64_t nwarps = (uint64_t)gridDim.x * 4;
  for (uint64_t l0 = (blockIdx.x * 4ull + warp) * 32; l0 < lines;
       l0 += nwarps * 32) {
    const uint32_t nl = (uint32_t)min((uint64_t)32, lines - l0);
    T prev = 0;
    for (uint32_t c0 = 0; c0 < n; c0 += 32) {
      const uint32_t ncol = min(32u, n - c0);
      for (uint32_t r = 0; r < nl; r++)
        if ((uint32_t)lane < ncol)
          t[r][lane] = x[(l0 + r) * n + c0 + lane];
      __syncwarp();
      if ((uint32_t)lane < nl)
        for (uint32_t j = 0; j < ncol; j++) {
          uint32_t i = c0 + j;
          prev = sub_rn(t[lane][j], mul_rn(prev, div_rn(am[i], bm[i])));
          t[lane][j] = prev;
        }
      __syncwarp();
      for (uint32_t r = 0; r < nl; r++)
        if ((uint32_t)lane < ncol)
          x[(l0 + r) * n + c0 + lane] = t[r][lane];
      __syncwarp();
    }
    prev = 0;
    const uint32_t nch = (n + 31) / 32;
    for (uint32_t cc = 0; cc < nch; cc++) {
      const uint32_t c0 = (nch - 1 - cc) * 32;
      const uint32_t ncol = min(32u, n - c0);
      for (uint32_t r = 0; r < nl; r++)
        if ((uint32_t)lane < ncol)
          t[r][lane] = x[(l0 + r) * n + c0 + lane];
      __syncwarp();
      if ((uint32_t)lane < nl)
        for (uint32_t jj = 0; jj < ncol; jj++) {
          uint32_t j = ncol - 1 - jj;
          uint32_t i = c0 + j;
          prev = div_rn(sub_rn(t[lane][j], mul_rn(am[i + 1], prev)), bm[i + 1]);
          t[lane][j] = prev;
        }
      __syncwarp();
      for (uint32_t r = 0; r < nl; r++)
        if ((uint32_t)lane < ncol)
          x[(l0 + r) * n + c0 + lane] = t[r][lane];
      __syncwarp();
    }
  }
}

// v[coarse box] (+/-)= corr (compact), reference Lwpk ADD / SUBTRACT
// (CopyND/LevelwiseProcessingKernel.hpp:21-127)
template <typename T>
__global__ void __launch_bounds__(256)
    k_apply_correction(T *__restrict__ v, const T *__restrict__ corr,
                       LevelArgs a, int sign) {
  for (uint64_t it = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
       it < a.total; it += (uint64_t)gridDim.x * blockDim.x) {
    uint64_t r = it, o = 0;
#pragma unroll
    for (int d = MGX_MAX_D - 1; d >= 0; d--) {
      o += (r % a.nc[d]) * a.st[d];
      r /= a.nc[d];
    }
    v[o] = sign > 0 ? add_rn(v[o], corr[it]) : sub_rn(v[o], corr[it]);
  }
}

} // namespace

// ---- host drivers ------------------------------------------------------------
namespace {

LevelArgs make_level_args(const mgx_ctx *ctx, int l) {
  LevelArgs a;
  const MgxTableOffsets &off =
      mgx_is_f32(ctx) ? ctx->tbf.off : ctx->tbd.off;
  a.total = 1;
  for (int d = 0; d < MGX_MAX_D; d++) {
    a.nf[d] = (uint32_t)ctx->lv.n[l][d];
    a.nc[d] = (uint32_t)ctx->lv.n[l - 1][d];
    a.st[d] = ctx->lv.stride[d];
    a.roff[d] = off.ratio[l][d];
    a.total *= a.nf[d];
  }
  a.half = mgx_is_f32(ctx) ? ctx->tbf.ratios_half : ctx->tbd.ratios_half;
  return a;
}

template <typename T>
int level_correction(mgx_ctx *ctx, int l, T *v, int sign) {
  const MgxLevels &lv = ctx->lv;
  const MgxTableOffsets &off = mgx_is_f32(ctx) ? ctx->tbf.off : ctx->tbd.off;
  const T *tables = (const T *)ctx->d_tables;
  const int pad = MGX_MAX_D - lv.D;
  const unsigned maxb = ctx->num_sms * 16;
  T *bufs[2] = {(T *)ctx->d_c0, (T *)ctx->d_c1};
  const T *in = v;
  LpkArgs a;
  for (int d = 0; d < MGX_MAX_D; d++) {
    a.cur[d] = (uint32_t)lv.n[l][d];
    a.in_st[d] = lv.stride[d];
    a.zc[d] = (uint32_t)lv.n[l - 1][d];
  }
  a.zero_box = 1;
  int which = 0;
  for (int d = MGX_MAX_D - 1; d >= pad; d--) {
    a.dim = d;
    a.nf = (uint32_t)lv.n[l][d];
    a.nc = (uint32_t)lv.n[l - 1][d];
    a.dist_off = off.dist[l][d];
    a.total = 1;
    for (int e = 0; e < MGX_MAX_D; e++)
      a.total *= (e == d) ? a.nc : a.cur[e];
    T *out = bufs[which];
    k_lpk<T><<<mgx_grid_for(a.total, 256, maxb), 256, 0, ctx->stream>>>(
        in, out, tables, a);
    ctx->launches++;
    // the output becomes the (compact) input of the next sweep
    a.cur[d] = a.nc;
    uint64_t st = 1;
    for (int e = MGX_MAX_D - 1; e >= 0; e--) {
      a.in_st[e] = st;
      st *= a.cur[e];
    }
    a.zero_box = 0;
    in = out;
    which ^= 1;
  }
  T *corr = const_cast<T *>(in);
  for (int d = MGX_MAX_D - 1; d >= pad; d--) {
    uint32_t n = (uint32_t)lv.n[l - 1][d];
    uint64_t inner = 1, outer = 1;
    for (int e = d + 1; e < MGX_MAX_D; e++)
      inner *= lv.n[l - 1][e];
    for (int e = 0; e < d; e++)
      outer *= lv.n[l - 1][e];
    const T *am = tables + off.am[l - 1][d], *bm = tables + off.bm[l - 1][d];
    uint64_t lines = outer * inner;
    if (inner == 1) {
      uint64_t groups = (lines + 127) / 128;
      k_ipk_contig<T><<<(unsigned)std::min<uint64_t>(groups, maxb), 128, 0,
                        ctx->stream>>>(corr, am, bm, n, lines);
    } else {
      k_ipk_strided<T><<<mgx_grid_for(lines, 128, maxb), 128, 0, ctx->stream>>>(
          corr, am, bm, n, inner, lines);
    }
    ctx->launches++;
  }
  LevelArgs la = make_level_args(ctx, l);
  la.total = 1;
  for (int d = 0; d < MGX_MAX_D; d++)
    la.total *= la.nc[d];
  k_apply_correction<T><<<mgx_grid_for(la.total, 256, maxb), 256, 0,
                          ctx->stream>>>(v, corr, la, sign);
  ctx->launches++;
  return MGX_SUCCESS;
}

template <typename T> int decompose_t(mgx_ctx *ctx, T *v, int orthogonal) {
  const T *tables = (const T *)ctx->d_tables;
  T *w = (T *)ctx->d_w;
  const unsigned maxb = ctx->num_sms * 16;
  for (int l = ctx->lv.L; l > 0; l--) {
    LevelArgs a = make_level_args(ctx, l);
    unsigned g = mgx_grid_for(a.total, 256, maxb);
    k_copy_box<T><<<g, 256, 0, ctx->stream>>>(v, w, a);
    k_level_coeff<T><<<g, 256, 0, ctx->stream>>>(w, v, tables, a);
    ctx->launches += 2;
    if (orthogonal)
      level_correction<T>(ctx, l, v, +1);
  }
  return MGX_SUCCESS;
}

template <typename T> int recompose_t(mgx_ctx *ctx, T *v, int orthogonal) {
  const T *tables = (const T *)ctx->d_tables;
  T *w = (T *)ctx->d_w;
  const unsigned maxb = ctx->num_sms * 16;
  for (int l = 1; l <= ctx->lv.L; l++) {
    if (orthogonal)
      level_correction<T>(ctx, l, v, -1);
    LevelArgs a = make_level_args(ctx, l);
    unsigned g = mgx_grid_for(a.total, 256, maxb);
    k_level_restore<T><<<g, 256, 0, ctx->stream>>>(v, w, tables, a);
    k_copy_box<T><<<g, 256, 0, ctx->stream>>>(w, v, a);
    ctx->launches += 2;
  }
  return MGX_SUCCESS;
}

} // namespace

// multi_dimension::decompose / recompose, reference
// MultiDimension/DataRefactoring.hpp:24-317
int mgxk_decompose(mgx_ctx *ctx, void *d_data, int orthogonal) {
  int rc = mgx_ensure_workspaces(ctx, true, false);
  if (rc)
    return rc;
  rc = mgx_is_f32(ctx) ? decompose_t<float>(ctx, (float *)d_data, orthogonal)
                       : decompose_t<double>(ctx, (double *)d_data, orthogonal);
  MGX_CUDA(cudaGetLastError());
  return rc;
}

// Hierarchical (no correction) recompose of levels 1..levels on a COMPACT array
// holding the level-`levels` box; used by the fused inverse for the coarse
// lattice.  d_v: in/out, d_w: scratch of the same size.
template <typename T>
static int recompose_compact_t(mgx_ctx *ctx, T *v, T *w, int levels) {
  const T *tables = (const T *)ctx->d_tables;
  const unsigned maxb = ctx->num_sms * 16;
  uint64_t st[MGX_MAX_D];
  uint64_t acc = 1;
  for (int d = MGX_MAX_D - 1; d >= 0; d--) {
    st[d] = acc;
    acc *= ctx->lv.n[levels][d];
  }
  for (int l = 1; l <= levels; l++) {
    LevelArgs a = make_level_args(ctx, l);
    for (int d = 0; d < MGX_MAX_D; d++)
      a.st[d] = st[d];
    unsigned g = mgx_grid_for(a.total, 256, maxb);
    k_level_restore<T><<<g, 256, 0, ctx->stream>>>(v, w, tables, a);
    k_copy_box<T><<<g, 256, 0, ctx->stream>>>(w, v, a);
    ctx->launches += 2;
  }
  return MGX_SUCCESS;
}

int mgxk_recompose_compact(mgx_ctx *ctx, void *d_v, void *d_w, int levels) {
  int rc = mgx_is_f32(ctx)
               ? recompose_compact_t<float>(ctx, (float *)d_v, (float *)d_w, levels)
               : recompose_compact_t<double>(ctx, (double *)d_v, (double *)d_w, levels);
  MGX_CUDA(cudaGetLastError());
  return rc;
}

int mgxk_recompose(mgx_ctx *ctx, void *d_data, int orthogonal) {
  int rc = mgx_ensure_workspaces(ctx, true, false);
  if (rc)
    return rc;
  rc = mgx_is_f32(ctx) ? recompose_t<float>(ctx, (float *)d_data, orthogonal)
                       : recompose_t<double>(ctx, (double *)d_data, orthogonal);
  MGX_CUDA(cudaGetLastError());
  return rc;
}
