// Row-streaming level kernels for the orthogonal-projection path (any D <= 5):
// the path taken for L2 / s-norm bounds and for D > 3 (Compressor.h:38-40).
//
// One level l of multi_dimension::decompose (MultiDimension/DataRefactoring.hpp:
// 24-177) is, in the reference, CopyND + GpkReo + Lpk1..3 + Ipk1..3 + Lwpk(ADD),
// each a full pass over the level box.  Here a level is
//
//   k_level_forward : interleaved level-l nodes -> coefficients (written once, as
//                     values or -- fused with lwqzk / outliers / histogram -- as
//                     16-bit symbols), coarse nodal values -> compact box A, and
//                     the first mass-matrix x restriction sweep (along the
//                     contiguous dimension) -> R, all from one read of the level
//   k_lpk_mid       : mass x restriction along each remaining dimension
//   k_ipk_*         : batched Thomas solves; the last one adds the correction
//                     into A (the reference's separate Lwpk ADD pass)
//
// and the next level reads A.  No level box is copied, nothing is done in
// place, and every kernel walks rows of the contiguous dimension with the lanes
// of a warp on consecutive elements; index arithmetic is per row, not per
// element.  recompose mirrors it (k_lpk_first, ..., k_level_restore).
//
// Arithmetic is the reference's non-FMA order (see mgx_device.cuh), so outputs
// are bit-identical to the SERIAL reference and to the gather kernels in
// mgx_refactor.cu that this file supersedes for orthogonal = 1.
#include <algorithm>
#include <cmath>

#include "mgx_device.cuh"
#include "mgx_internal.h"

namespace {

constexpr int OUTER = MGX_MAX_D - 1; // dimensions other than the contiguous one
constexpr int FD = MGX_MAX_D - 1;    // index of the contiguous dimension

struct LvGeom {
  uint32_t nf[MGX_MAX_D]; // level-l shape
  uint32_t nc[MGX_MAX_D]; // level-(l-1) shape
  uint64_t vst[MGX_MAX_D]; // strides of the full array (reordered layout)
  uint64_t sst[MGX_MAX_D]; // strides of the interleaved level-l node array
  uint64_t ast[MGX_MAX_D]; // compact strides of the level-(l-1) box
  uint32_t roff[MGX_MAX_D]; // ratio tables of level l
  uint32_t dist_f;          // dist table of the contiguous dim, level l
  uint64_t rows;            // prod nf[0..3]
  uint32_t rows_per_block;
  int half;                 // every ratio is exactly 0.5
};

template <typename T> struct QuantOut {
  uint16_t *sym;
  uint32_t *g_freq;
  unsigned long long *out_count;
  unsigned long long *out_idx;
  long long *out_val;
  unsigned long long out_cap;
  T qv, vol;
  long long shift;
  uint32_t dict;
};

// ---- mass_trans with its data-independent part hoisted -----------------------
// (Correction/LPKFunctor.h:13-67, non-FMA branch; same operations as
// mgx_mass_trans, the nine quotients are formed once per output index)
template <typename T> struct MassK {
  T h1_6, h2_6, h3_6, h4_6, s12_3, s23_3, s34_3, r1, r4;
};

template <typename T>
__device__ __forceinline__ MassK<T> mass_consts(const T *__restrict__ dist,
                                                int64_t k, uint32_t nf) {
  T h[4];
#pragma unroll
  for (int j = 0; j < 4; j++) {
    int64_t idx = k - 2 + j;
    h[j] = (idx >= 0 && idx < (int64_t)nf) ? dist[idx] : (T)0;
  }
  MassK<T> m;
  T s12 = add_rn(h[0], h[1]), s34 = add_rn(h[2], h[3]), s23 = add_rn(h[1], h[2]);
  m.r1 = (s12 != (T)0) ? div_rn(h[0], s12) : (T)0;
  m.r4 = (s34 != (T)0) ? div_rn(h[3], s34) : (T)0;
  m.h1_6 = div_rn(h[0], (T)6);
  m.h2_6 = div_rn(h[1], (T)6);
  m.h3_6 = div_rn(h[2], (T)6);
  m.h4_6 = div_rn(h[3], (T)6);
  m.s12_3 = div_rn(s12, (T)3);
  m.s23_3 = div_rn(s23, (T)3);
  m.s34_3 = div_rn(s34, (T)3);
  return m;
}

template <typename T>
__device__ __forceinline__ T mass_apply(const MassK<T> &m, T a, T b, T c, T d,
                                        T e) {
  T tb = add_rn(add_rn(mul_rn(a, m.h1_6), mul_rn(b, m.s12_3)), mul_rn(c, m.h2_6));
  T tc = add_rn(add_rn(mul_rn(b, m.h2_6), mul_rn(c, m.s23_3)), mul_rn(d, m.h3_6));
  T td = add_rn(add_rn(mul_rn(c, m.h3_6), mul_rn(d, m.s34_3)), mul_rn(e, m.h4_6));
  return add_rn(tc, add_rn(mul_rn(tb, m.r1), mul_rn(td, m.r4)));
}

// ---- row walker ----------------------------------------------------------------
// A block owns a contiguous range of rows of the level box in REORDERED row
// order (p[d] < nc[d]: coarse node, else coefficient).  For the current row:
//   q[d]    interleaved index of the row along d, odd[d] = it is a coefficient row
//   nodd    number of odd outer dimensions; lo[j]/hi[j] element offsets of the
//           two neighbour rows along the j-th odd dimension (innermost first),
//           tt[j] its interpolation ratio
template <typename T> struct Row {
  uint32_t p[OUTER];
  uint64_t so, vo, ao; // offsets of the row in the node array / full array / A
  int nodd;
  int64_t lo[OUTER], hi[OUTER];
  T tt[OUTER];
};

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

template <typename T>
__device__ __forceinline__ void row_first(Row<T> &r, const LvGeom &g,
                                          uint64_t row) {
#pragma unroll
  for (int d = OUTER - 1; d >= 0; d--) {
    r.p[d] = (uint32_t)(row % g.nf[d]);
    row /= g.nf[d];
  }
}

template <typename T>
__device__ __forceinline__ void row_next(Row<T> &r, const LvGeom &g) {
  bool carry = true;
#pragma unroll
  for (int d = OUTER - 1; d >= 0; d--) {
    if (carry) {
      r.p[d]++;
      carry = r.p[d] == g.nf[d];
      if (carry)
        r.p[d] = 0;
    }
  }
}

// FORWARD: neighbours are rows of the interleaved node array (q -/+ 1);
// otherwise rows of the compact coarse box A (positions (q-1)/2 and (q-1)/2+1).
template <typename T, bool FORWARD>
__device__ __forceinline__ void row_setup(Row<T> &r, const LvGeom &g,
                                          const T *__restrict__ tables) {
  r.so = 0;
  r.vo = 0;
  r.ao = 0;
  r.nodd = 0;
#pragma unroll
  for (int j = 0; j < OUTER; j++) {
    r.lo[j] = 0;
    r.hi[j] = 0;
    r.tt[j] = (T)0.5;
  }
#pragma unroll
  for (int d = OUTER - 1; d >= 0; d--) {
    uint32_t q;
    bool odd;
    pos_to_node(r.p[d], g.nf[d], g.nc[d], q, odd);
    r.so += (uint64_t)q * g.sst[d];
    r.vo += (uint64_t)r.p[d] * g.vst[d];
    if (odd) {
      T t = g.half ? (T)0.5 : tables[g.roff[d] + q - 1];
      int64_t l, h;
      if (FORWARD) {
        l = -(int64_t)g.sst[d];
        h = (int64_t)g.sst[d];
      } else {
        r.ao += (uint64_t)((q - 1) / 2) * g.ast[d];
        l = 0;
        h = (int64_t)g.ast[d];
      }
#pragma unroll
      for (int j = 0; j < OUTER; j++)
        if (j == r.nodd) {
          r.lo[j] = l;
          r.hi[j] = h;
          r.tt[j] = t;
        }
      r.nodd++;
    } else {
      r.ao += (uint64_t)r.p[d] * g.ast[d];
    }
  }
}

// Interpolant of the 2^K neighbour rows at one even column (a) and at the odd
// column to its right (b, interpolated along the contiguous dim FIRST, then
// along the odd outer dims innermost to outermost -- the reference's f, c, r
// order, GridProcessingKernel3D.hpp).
template <typename T, int K>
__device__ __forceinline__ void corners(const T *__restrict__ base, int64_t off,
                                        const Row<T> &r, uint32_t e0, uint32_t e1,
                                        bool has_odd, T tf, T &a, T &b) {
  if constexpr (K == 0) {
    T x0 = base[off + e0];
    a = x0;
    b = has_odd ? mgx_lerp(x0, base[off + e1], tf) : (T)0;
  } else {
    T a0, b0, a1, b1;
    corners<T, K - 1>(base, off + r.lo[K - 1], r, e0, e1, has_odd, tf, a0, b0);
    corners<T, K - 1>(base, off + r.hi[K - 1], r, e0, e1, has_odd, tf, a1, b1);
    a = mgx_lerp(a0, a1, r.tt[K - 1]);
    b = mgx_lerp(b0, b1, r.tt[K - 1]);
  }
}

// coefficient -> float store, or -> symbol + histogram + outlier list
// (lwqzk LinearQuantization.hpp:96-111, OutlierSeparator.hpp:17-120,
// Histogram.hpp).  Called by all lanes of a warp.
template <typename T, bool QUANT>
__device__ __forceinline__ void emit(T c, bool valid, uint64_t idx,
                                     T *__restrict__ v, const QuantOut<T> &qo,
                                     uint32_t *s_hist) {
  if constexpr (!QUANT) {
    if (valid)
      v[idx] = c;
  } else {
    uint32_t sv = 0;
    if (valid) {
      long long q = mgx_quantize(c, qo.qv, qo.vol, qo.shift);
      if (q < 0 || q >= (long long)qo.dict) {
        unsigned long long slot = atomicAdd(qo.out_count, 1ull);
        if (slot < qo.out_cap) {
          qo.out_idx[slot] = idx;
          qo.out_val[slot] = q;
        }
        q = 0;
      }
      sv = (uint32_t)q;
      qo.sym[idx] = (uint16_t)sv;
    }
    mgx_hist_add(s_hist, sv, valid);
  }
}

// ---- forward level kernel ------------------------------------------------------
// Lane l of warp-column wc owns "pair" i = 30*wc - 1 + l: the coarse-along-f
// position i and the coefficient-along-f position nc_f + i of every row.  Lanes
// 0 and 31 are halo pairs (computed, not written) so that the restriction
// stencil (pairs i-1, i, i+1) is served by two shuffles.
constexpr int FW_WARPS = 8;
constexpr int FW_OWN = 30;

template <typename T, bool QUANT, int K>
__device__ __forceinline__ void forward_row(
    const T *__restrict__ src, T *__restrict__ v, T *__restrict__ A,
    T *__restrict__ R, const Row<T> &r, const LvGeom &g, const QuantOut<T> &qo,
    uint32_t *s_hist, const MassK<T> &mk, int64_t i, bool pair_ok, bool has_odd,
    bool owned, uint32_t qe, T tf, uint64_t row) {
  const uint32_t ncf = g.nc[FD];
  T ce = 0, co = 0;
  if (pair_ok) {
    const T *srow = src + r.so;
    T ve = srow[qe];
    T vo = has_odd ? srow[2 * i + 1] : (T)0;
    if constexpr (K == 0) {
      if (owned)
        A[r.ao + i] = ve; // coarse node of the next level
      if (has_odd)
        co = sub_rn(vo, mgx_lerp(ve, srow[2 * i + 2], tf));
    } else {
      T a, b;
      corners<T, K>(src, (int64_t)r.so, r, qe, (uint32_t)(2 * i + 2), has_odd, tf,
                    a, b);
      ce = sub_rn(ve, a);
      if (has_odd)
        co = sub_rn(vo, b);
    }
  }
  if (K > 0)
    emit<T, QUANT>(ce, owned && pair_ok, r.vo + (uint64_t)i, v, qo, s_hist);
  emit<T, QUANT>(co, owned && has_odd, r.vo + ncf + (uint64_t)i, v, qo, s_hist);
  // first restriction sweep (Lpk1Reo3D, LinearProcessingKernel3D.hpp:27-372):
  // the all-coarse corner of the coefficient field reads as zero (ce = 0, K = 0)
  T cem1 = __shfl_up_sync(0xffffffffu, ce, 1);
  T com1 = __shfl_up_sync(0xffffffffu, co, 1);
  T cep1 = __shfl_down_sync(0xffffffffu, ce, 1);
  if (R != nullptr && owned && pair_ok)
    R[row * ncf + (uint64_t)i] = mass_apply(mk, cem1, com1, ce, co, cep1);
}

template <typename T, bool QUANT>
__global__ void __launch_bounds__(FW_WARPS * 32)
    k_level_forward(const T *__restrict__ src, T *__restrict__ v,
                    T *__restrict__ A, T *__restrict__ R,
                    const T *__restrict__ tables,
                    const __grid_constant__ LvGeom g,
                    const __grid_constant__ QuantOut<T> qo) {
  extern __shared__ uint32_t s_hist[];
  if (QUANT) {
    for (uint32_t k = threadIdx.x; k < qo.dict; k += blockDim.x)
      s_hist[k] = 0;
    __syncthreads();
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t ncf = g.nc[FD], nff = g.nf[FD], ncoef = nff - ncf;
  const int64_t i = (int64_t)(blockIdx.x * FW_WARPS + warp) * FW_OWN - 1 + lane;
  const bool pair_ok = i >= 0 && i < (int64_t)ncf;
  const bool has_odd = i >= 0 && i < (int64_t)ncoef;
  const bool owned = lane >= 1 && lane <= FW_OWN;
  const uint32_t qe = pair_ok ? ((i == (int64_t)ncf - 1) ? nff - 1 : (uint32_t)(2 * i)) : 0;
  const T tf = (g.half || !has_odd) ? (T)0.5 : tables[g.roff[FD] + 2 * i];
  MassK<T> mk = mass_consts<T>(tables + g.dist_f, pair_ok ? 2 * i : 0, nff);
  const uint64_t row0 = (uint64_t)blockIdx.y * g.rows_per_block;
  const uint64_t row1 = min(g.rows, row0 + g.rows_per_block);
  // whole warp-columns past the end have nothing to do (but must still take
  // part in the histogram flush below)
  const bool warp_live = (int64_t)(blockIdx.x * FW_WARPS + warp) * FW_OWN < (int64_t)ncf;
  if (warp_live && row0 < row1) {
    Row<T> r;
    row_first(r, g, row0);
    for (uint64_t row = row0; row < row1; row++) {
      row_setup<T, true>(r, g, tables);
      switch (r.nodd) {
      case 0:
        forward_row<T, QUANT, 0>(src, v, A, R, r, g, qo, s_hist, mk, i, pair_ok,
                                 has_odd, owned, qe, tf, row);
        break;
      case 1:
        forward_row<T, QUANT, 1>(src, v, A, R, r, g, qo, s_hist, mk, i, pair_ok,
                                 has_odd, owned, qe, tf, row);
        break;
      case 2:
        forward_row<T, QUANT, 2>(src, v, A, R, r, g, qo, s_hist, mk, i, pair_ok,
                                 has_odd, owned, qe, tf, row);
        break;
      case 3:
        forward_row<T, QUANT, 3>(src, v, A, R, r, g, qo, s_hist, mk, i, pair_ok,
                                 has_odd, owned, qe, tf, row);
        break;
      default:
        forward_row<T, QUANT, 4>(src, v, A, R, r, g, qo, s_hist, mk, i, pair_ok,
                                 has_odd, owned, qe, tf, row);
        break;
      }
      row_next(r, g);
    }
  }
  if (QUANT) {
    __syncthreads();
    for (uint32_t k = threadIdx.x; k < qo.dict; k += blockDim.x) {
      uint32_t c = s_hist[k];
      if (c)
        atomicAdd(&qo.g_freq[k], c);
    }
  }
}

// ---- first restriction sweep from stored coefficients (recompose) ------------------
template <typename T>
__global__ void __launch_bounds__(256)
    k_lpk_first(const T *__restrict__ v, T *__restrict__ R,
                const T *__restrict__ tables, const __grid_constant__ LvGeom g) {
  const uint32_t ncf = g.nc[FD], nff = g.nf[FD], ncoef = nff - ncf;
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (int64_t)ncf)
    return;
  MassK<T> mk = mass_consts<T>(tables + g.dist_f, 2 * i, nff);
  const uint64_t row0 = (uint64_t)blockIdx.y * g.rows_per_block;
  const uint64_t row1 = min(g.rows, row0 + g.rows_per_block);
  if (row0 >= row1)
    return;
  Row<T> r;
  row_first(r, g, row0);
  for (uint64_t row = row0; row < row1; row++) {
    bool coarse_row = true;
    uint64_t vo = 0;
#pragma unroll
    for (int d = 0; d < OUTER; d++) {
      coarse_row = coarse_row && r.p[d] < g.nc[d];
      vo += (uint64_t)r.p[d] * g.vst[d];
    }
    const T *vr = v + vo;
    T a = 0, b = 0, c = 0, d = 0, e = 0;
    if (!coarse_row) {
      a = i >= 1 ? vr[i - 1] : (T)0;
      c = vr[i];
      e = i + 1 < (int64_t)ncf ? vr[i + 1] : (T)0;
    }
    b = (i >= 1 && i - 1 < (int64_t)ncoef) ? vr[ncf + i - 1] : (T)0;
    d = i < (int64_t)ncoef ? vr[ncf + i] : (T)0;
    R[row * ncf + (uint64_t)i] = mass_apply(mk, a, b, c, d, e);
    row_next(r, g);
  }
}

// ---- mass x restriction along a non-contiguous dimension -----------------------
// in : (outer, nf [nc coarse | nf-nc coefficients], inner) compact
// out: (outer, nc, inner) compact.  A block owns one output index i along the
// dimension (its nine quotients are formed once) and streams (outer, inner).
// (Lpk2Reo3D / Lpk3Reo3D, LinearProcessingKernel3D.hpp:374-1097)
template <typename T>
__global__ void __launch_bounds__(256)
    k_lpk_mid(const T *__restrict__ in, T *__restrict__ out,
              const T *__restrict__ dist, uint32_t outer, uint32_t nf, uint32_t nc,
              uint64_t inner) {
  const uint32_t i = blockIdx.x;
  const uint32_t ncoef = nf - nc;
  const MassK<T> mk = mass_consts<T>(dist, 2 * (int64_t)i, nf);
  const bool ha = i >= 1, hb = i >= 1 && i - 1 < ncoef, hd = i < ncoef,
             he = i + 1 < nc;
  for (uint32_t o = blockIdx.z; o < outer; o += gridDim.z) {
    const T *ib = in + (uint64_t)o * nf * inner;
    T *ob = out + ((uint64_t)o * nc + i) * inner;
    for (uint64_t x = (uint64_t)blockIdx.y * blockDim.x + threadIdx.x; x < inner;
         x += (uint64_t)gridDim.y * blockDim.x) {
      T a = ha ? ib[(uint64_t)(i - 1) * inner + x] : (T)0;
      T b = hb ? ib[(uint64_t)(nc + i - 1) * inner + x] : (T)0;
      T c = ib[(uint64_t)i * inner + x];
      T d = hd ? ib[(uint64_t)(nc + i) * inner + x] : (T)0;
      T e = he ? ib[(uint64_t)(i + 1) * inner + x] : (T)0;
      ob[x] = mass_apply(mk, a, b, c, d, e);
    }
  }
}

// ---- batched Thomas solves -------------------------------------------------------
// reference Ipk1/2/3Reo (IterativeProcessingKernel3D.hpp:28-1169) with
// tridiag_forward2 / tridiag_backward2 (IPKFunctor.h:13-50):
//   forward : x[i] = x[i] - x[i-1] * (am[i] / bm[i])        (abr = am/bm table)
//   backward: x[i] = (x[i] - am[i+1] * x[i+1]) / bm[i+1]
// ACC != 0: the backward sweep does not store x but adds (ACC > 0) or subtracts
// (ACC < 0) it into acc at the same compact index -- the reference's separate
// Lwpk ADD / SUBTRACT pass (LevelwiseProcessingKernel.hpp:21-127).
//
// Strided variant (solve dimension not contiguous): a thread owns one line,
// consecutive threads consecutive `inner` indices; loads are issued a batch
// ahead of the dependent chain.
template <typename T, int B>
__global__ void __launch_bounds__(128)
    k_ipk_strided(T *__restrict__ x, const T *__restrict__ am,
                  const T *__restrict__ bm, const T *__restrict__ abr, uint32_t n,
                  uint64_t inner, uint64_t lines, T *__restrict__ acc, int sign) {
  const uint64_t ln = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
  if (ln >= lines)
    return;
  const uint64_t o = ln / inner, in = ln % inner;
  T *p = x + o * n * inner + in;
  T *pa = acc ? acc + o * n * inner + in : nullptr;
  T cur[B], nxt[B];
  T prev = 0;
#pragma unroll
  for (int k = 0; k < B; k++)
    cur[k] = (uint32_t)k < n ? p[(uint64_t)k * inner] : (T)0;
  for (uint32_t i0 = 0; i0 < n; i0 += B) {
#pragma unroll
    for (int k = 0; k < B; k++) {
      uint32_t i = i0 + B + k;
      nxt[k] = i < n ? p[(uint64_t)i * inner] : (T)0;
    }
#pragma unroll
    for (int k = 0; k < B; k++) {
      uint32_t i = i0 + k;
      if (i < n) {
        prev = sub_rn(cur[k], mul_rn(prev, abr[i]));
        p[(uint64_t)i * inner] = prev;
      }
    }
#pragma unroll
    for (int k = 0; k < B; k++)
      cur[k] = nxt[k];
  }
  // backward, from the last element; batch j covers i = n-1-j*B-k
  prev = 0;
  T ca[B], na[B];
#pragma unroll
  for (int k = 0; k < B; k++) {
    bool ok = (uint32_t)k < n;
    uint64_t off = (uint64_t)(n - 1 - (ok ? k : 0)) * inner;
    cur[k] = ok ? p[off] : (T)0;
    ca[k] = (ok && pa) ? pa[off] : (T)0;
  }
  for (uint32_t j0 = 0; j0 < n; j0 += B) {
#pragma unroll
    for (int k = 0; k < B; k++) {
      uint32_t j = j0 + B + k;
      bool ok = j < n;
      uint64_t off = (uint64_t)(n - 1 - (ok ? j : 0)) * inner;
      nxt[k] = ok ? p[off] : (T)0;
      na[k] = (ok && pa) ? pa[off] : (T)0;
    }
#pragma unroll
    for (int k = 0; k < B; k++) {
      uint32_t j = j0 + k;
      if (j < n) {
        uint32_t i = n - 1 - j;
        prev = div_rn(sub_rn(cur[k], mul_rn(am[i + 1], prev)), bm[i + 1]);
        if (pa)
          pa[(uint64_t)i * inner] =
              sign > 0 ? add_rn(ca[k], prev) : sub_rn(ca[k], prev);
        else
          p[(uint64_t)i * inner] = prev;
      }
    }
#pragma unroll
    for (int k = 0; k < B; k++) {
      cur[k] = nxt[k];
      ca[k] = na[k];
    }
  }
}

// Contiguous variant: one warp owns 32 lines; 32x32 tiles are moved with
// coalesced row accesses and transposed through shared memory so that lane r
// walks line r.  The next tile is loaded into registers while the current one
// is being eliminated.
template <typename T>
__global__ void __launch_bounds__(128)
    k_ipk_contig(T *__restrict__ x, const T *__restrict__ am,
                 const T *__restrict__ bm, const T *__restrict__ abr, uint32_t n,
                 uint64_t lines, T *__restrict__ acc, int sign) {
  __shared__ T tile[4][32][33];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  T(*t)[33] = tile[warp];
  const uint64_t l0 = (blockIdx.x * 4ull + warp) * 32;
  if (l0 >= lines)
    return;
  const uint32_t nl = (uint32_t)min((uint64_t)32, lines - l0);
  const uint32_t nch = (n + 31) / 32;
  T *xb = x + l0 * n;
  T *ab = acc ? acc + l0 * n : nullptr;
  T reg[32];
  auto load_tile = [&](uint32_t c0) {
#pragma unroll
    for (uint32_t r = 0; r < 32; r++)
      reg[r] = (r < nl && c0 + lane < n) ? xb[(uint64_t)r * n + c0 + lane] : (T)0;
  };
  // forward
  T prev = 0;
  load_tile(0);
  for (uint32_t ch = 0; ch < nch; ch++) {
    const uint32_t c0 = ch * 32;
    const uint32_t ncol = min(32u, n - c0);
#pragma unroll
    for (uint32_t r = 0; r < 32; r++)
      t[r][lane] = reg[r];
    __syncwarp();
    if (ch + 1 < nch)
      load_tile(c0 + 32);
    if ((uint32_t)lane < nl)
      for (uint32_t j = 0; j < ncol; j++) {
        prev = sub_rn(t[lane][j], mul_rn(prev, abr[c0 + j]));
        t[lane][j] = prev;
      }
    __syncwarp();
    for (uint32_t r = 0; r < nl; r++)
      if ((uint32_t)lane < ncol)
        xb[(uint64_t)r * n + c0 + lane] = t[r][lane];
    __syncwarp();
  }
  // backward
  prev = 0;
  load_tile((nch - 1) * 32);
  for (uint32_t cc = 0; cc < nch; cc++) {
    const uint32_t c0 = (nch - 1 - cc) * 32;
    const uint32_t ncol = min(32u, n - c0);
#pragma unroll
    for (uint32_t r = 0; r < 32; r++)
      t[r][lane] = reg[r];
    __syncwarp();
    if (cc + 1 < nch)
      load_tile(c0 - 32);
    if ((uint32_t)lane < nl)
      for (uint32_t jj = 0; jj < ncol; jj++) {
        uint32_t j = ncol - 1 - jj;
        uint32_t i = c0 + j;
        prev = div_rn(sub_rn(t[lane][j], mul_rn(am[i + 1], prev)), bm[i + 1]);
        t[lane][j] = prev;
      }
    __syncwarp();
    if (ab) {
      for (uint32_t r = 0; r < nl; r++)
        if ((uint32_t)lane < ncol) {
          uint64_t off = (uint64_t)r * n + c0 + lane;
          ab[off] = sign > 0 ? add_rn(ab[off], t[r][lane]) : sub_rn(ab[off], t[r][lane]);
        }
    } else {
      for (uint32_t r = 0; r < nl; r++)
        if ((uint32_t)lane < ncol)
          xb[(uint64_t)r * n + c0 + lane] = t[r][lane];
    }
    __syncwarp();
  }
}

// ---- inverse level kernel ------------------------------------------------------------
// coefficients (v, reordered) + corrected coarse box A -> interleaved level-l
// nodes (GpkRev3D, GridProcessingKernel3D.hpp:1230-2412)
template <typename T, int K>
__device__ __forceinline__ void restore_row(const T *__restrict__ v,
                                            const T *__restrict__ A,
                                            T *__restrict__ dst, const Row<T> &r,
                                            const LvGeom &g, uint32_t i,
                                            bool has_odd, uint32_t qe, T tf) {
  const uint32_t ncf = g.nc[FD];
  T *drow = dst + r.so;
  const T *vr = v + r.vo;
  if constexpr (K == 0) {
    const T *arow = A + r.ao;
    T a0 = arow[i];
    drow[qe] = a0;
    if (has_odd)
      drow[2 * i + 1] = add_rn(vr[ncf + i], mgx_lerp(a0, arow[i + 1], tf));
  } else {
    T a, b;
    corners<T, K>(A, (int64_t)r.ao, r, i, i + 1, has_odd, tf, a, b);
    drow[qe] = add_rn(vr[i], a);
    if (has_odd)
      drow[2 * i + 1] = add_rn(vr[ncf + i], b);
  }
}

template <typename T>
__global__ void __launch_bounds__(256)
    k_level_restore_rows(const T *__restrict__ v, const T *__restrict__ A,
                         T *__restrict__ dst, const T *__restrict__ tables,
                         const __grid_constant__ LvGeom g) {
  const uint32_t ncf = g.nc[FD], nff = g.nf[FD], ncoef = nff - ncf;
  const uint64_t ii = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (ii >= ncf)
    return;
  const uint32_t i = (uint32_t)ii;
  const bool has_odd = i < ncoef;
  const uint32_t qe = (i == ncf - 1) ? nff - 1 : 2 * i;
  const T tf = (g.half || !has_odd) ? (T)0.5 : tables[g.roff[FD] + 2 * i];
  const uint64_t row0 = (uint64_t)blockIdx.y * g.rows_per_block;
  const uint64_t row1 = min(g.rows, row0 + g.rows_per_block);
  if (row0 >= row1)
    return;
  Row<T> r;
  row_first(r, g, row0);
  for (uint64_t row = row0; row < row1; row++) {
    row_setup<T, false>(r, g, tables);
    switch (r.nodd) {
    case 0:
      restore_row<T, 0>(v, A, dst, r, g, i, has_odd, qe, tf);
      break;
    case 1:
      restore_row<T, 1>(v, A, dst, r, g, i, has_odd, qe, tf);
      break;
    case 2:
      restore_row<T, 2>(v, A, dst, r, g, i, has_odd, qe, tf);
      break;
    case 3:
      restore_row<T, 3>(v, A, dst, r, g, i, has_odd, qe, tf);
      break;
    default:
      restore_row<T, 4>(v, A, dst, r, g, i, has_odd, qe, tf);
      break;
    }
    row_next(r, g);
  }
}

// ---- level-0 box: compact A <-> full array (values or symbols) ------------------------
template <typename T, bool QUANT, bool GATHER>
__global__ void __launch_bounds__(256)
    k_coarse_box(T *__restrict__ A, T *__restrict__ v,
                 const __grid_constant__ LvGeom g,
                 const __grid_constant__ QuantOut<T> qo) {
  uint64_t total = 1;
#pragma unroll
  for (int d = 0; d < MGX_MAX_D; d++)
    total *= g.nc[d];
  for (uint64_t it = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; it < total;
       it += (uint64_t)gridDim.x * blockDim.x) {
    uint64_t r = it, o = 0;
#pragma unroll
    for (int d = MGX_MAX_D - 1; d >= 0; d--) {
      o += (r % g.nc[d]) * g.vst[d];
      r /= g.nc[d];
    }
    if constexpr (GATHER) {
      A[it] = v[o];
    } else if constexpr (!QUANT) {
      v[o] = A[it];
    } else {
      long long q = mgx_quantize(A[it], qo.qv, qo.vol, qo.shift);
      if (q < 0 || q >= (long long)qo.dict) {
        unsigned long long slot = atomicAdd(qo.out_count, 1ull);
        if (slot < qo.out_cap) {
          qo.out_idx[slot] = o;
          qo.out_val[slot] = q;
        }
        q = 0;
      }
      qo.sym[o] = (uint16_t)q;
      atomicAdd(&qo.g_freq[(uint32_t)q], 1u);
    }
  }
}

// ---- host side ----------------------------------------------------------------------
template <typename T> const MgxHostTables<T> &host_tables(const mgx_ctx *ctx);
template <> const MgxHostTables<float> &host_tables<float>(const mgx_ctx *ctx) {
  return ctx->tbf;
}
template <> const MgxHostTables<double> &host_tables<double>(const mgx_ctx *ctx) {
  return ctx->tbd;
}

void compact_strides(const uint64_t *n, uint64_t *st) {
  uint64_t acc = 1;
  for (int d = MGX_MAX_D - 1; d >= 0; d--) {
    st[d] = acc;
    acc *= n[d];
  }
}

// geometry of level l; `node_full`: the interleaved level-l node array is the
// full array (l == L), else the compact level-l box
template <typename T>
LvGeom level_geom(const mgx_ctx *ctx, int l, unsigned target_blocks,
                  unsigned col_blocks) {
  const MgxLevels &lv = ctx->lv;
  const MgxTableOffsets &off = host_tables<T>(ctx).off;
  LvGeom g;
  uint64_t cst[MGX_MAX_D], ast[MGX_MAX_D];
  compact_strides(lv.n[l], cst);
  compact_strides(lv.n[l - 1], ast);
  g.rows = 1;
  for (int d = 0; d < MGX_MAX_D; d++) {
    g.nf[d] = (uint32_t)lv.n[l][d];
    g.nc[d] = (uint32_t)lv.n[l - 1][d];
    g.vst[d] = lv.stride[d];
    g.sst[d] = (l == lv.L) ? lv.stride[d] : cst[d];
    g.ast[d] = ast[d];
    g.roff[d] = off.ratio[l][d];
    if (d < FD)
      g.rows *= g.nf[d];
  }
  g.dist_f = off.dist[l][FD];
  g.half = host_tables<T>(ctx).ratios_half ? 1 : 0;
  // rows per block: enough blocks to fill the machine, contiguous row ranges
  uint64_t want_y = std::max<uint64_t>(1, target_blocks / std::max(1u, col_blocks));
  want_y = std::min<uint64_t>(want_y, 65535);
  uint64_t rpb = (g.rows + want_y - 1) / want_y;
  g.rows_per_block = (uint32_t)std::max<uint64_t>(1, std::min<uint64_t>(rpb, 1u << 30));
  return g;
}

unsigned grid_y(const LvGeom &g) {
  return (unsigned)((g.rows + g.rows_per_block - 1) / g.rows_per_block);
}

struct Scratch {
  void *A[2]; // ping-pong compact coarse boxes
  void *R[2]; // ping-pong restriction outputs
};

int get_scratch(mgx_ctx *ctx, Scratch *s) {
  int rc = mgx_ensure_workspaces(ctx, false, false, true);
  if (rc)
    return rc;
  const MgxLevels &lv = ctx->lv;
  size_t need = mgx_elem(ctx);
  for (int d = 0; d < MGX_MAX_D; d++)
    need *= lv.L >= 1 ? lv.n[lv.L - 1][d] : 1;
  if (ctx->coarse_bytes < need) {
    if (ctx->d_coarse0)
      cudaFree(ctx->d_coarse0);
    if (ctx->d_coarse1)
      cudaFree(ctx->d_coarse1);
    ctx->d_coarse0 = ctx->d_coarse1 = nullptr;
    ctx->coarse_bytes = 0;
    MGX_CUDA(cudaMalloc(&ctx->d_coarse0, need));
    MGX_CUDA(cudaMalloc(&ctx->d_coarse1, need));
    ctx->coarse_bytes = need;
  }
  s->A[0] = ctx->d_coarse0;
  s->A[1] = ctx->d_coarse1;
  s->R[0] = ctx->d_c0;
  s->R[1] = ctx->d_c1;
  return MGX_SUCCESS;
}

// restriction sweeps along the non-contiguous dims + all solves of level l.
// R0 holds the output of the first sweep: shape (nf0..nf3, nc4).  The final
// solve accumulates into A (sign +1 decompose, -1 recompose).
template <typename T>
int correction_tail(mgx_ctx *ctx, int l, T *R0, T *R1, T *A, int sign) {
  const MgxLevels &lv = ctx->lv;
  const MgxTableOffsets &off = host_tables<T>(ctx).off;
  const T *tables = (const T *)ctx->d_tables;
  const int pad = MGX_MAX_D - lv.D;
  uint64_t cur[MGX_MAX_D];
  for (int d = 0; d < MGX_MAX_D; d++)
    cur[d] = lv.n[l][d];
  cur[FD] = lv.n[l - 1][FD];
  T *in = R0, *out = R1;
  for (int d = FD - 1; d >= pad; d--) {
    uint64_t outer = 1, inner = 1;
    for (int e = 0; e < d; e++)
      outer *= cur[e];
    for (int e = d + 1; e < MGX_MAX_D; e++)
      inner *= cur[e];
    const uint32_t nf = (uint32_t)lv.n[l][d], nc = (uint32_t)lv.n[l - 1][d];
    unsigned gy = (unsigned)std::min<uint64_t>((inner + 255) / 256, 65535);
    uint64_t want_z = std::max<uint64_t>(1, (uint64_t)ctx->num_sms * 8 / ((uint64_t)nc * gy));
    unsigned gz = (unsigned)std::min<uint64_t>(std::min<uint64_t>(outer, want_z), 65535);
    k_lpk_mid<T><<<dim3(nc, gy, gz), 256, 0, ctx->stream>>>(
        in, out, tables + off.dist[l][d], (uint32_t)outer, nf, nc, inner);
    ctx->launches++;
    cur[d] = nc;
    std::swap(in, out);
  }
  T *corr = in;
  for (int d = FD; d >= pad; d--) {
    const uint32_t n = (uint32_t)lv.n[l - 1][d];
    uint64_t inner = 1, outer = 1;
    for (int e = d + 1; e < MGX_MAX_D; e++)
      inner *= lv.n[l - 1][e];
    for (int e = 0; e < d; e++)
      outer *= lv.n[l - 1][e];
    const T *am = tables + off.am[l - 1][d], *bm = tables + off.bm[l - 1][d],
            *abr = tables + off.abr[l - 1][d];
    const uint64_t lines = outer * inner;
    T *acc = (d == pad) ? A : nullptr;
    if (inner == 1) {
      uint64_t groups = (lines + 127) / 128;
      k_ipk_contig<T><<<(unsigned)groups, 128, 0, ctx->stream>>>(
          corr, am, bm, abr, n, lines, acc, sign);
    } else {
      constexpr int B = sizeof(T) == 4 ? 16 : 8;
      k_ipk_strided<T, B><<<(unsigned)((lines + 127) / 128), 128, 0, ctx->stream>>>(
          corr, am, bm, abr, n, inner, lines, acc, sign);
    }
    ctx->launches++;
  }
  return MGX_SUCCESS;
}

template <typename T>
void fill_quant(const mgx_ctx *ctx, QuantOut<T> *qo, const double *q,
                const double *vol, int level) {
  qo->sym = ctx->d_sym;
  qo->g_freq = ctx->d_freq;
  qo->out_count = (unsigned long long *)(ctx->d_scalars + MGX_SC_OUTLIERS);
  qo->out_idx = (unsigned long long *)ctx->d_out_idx;
  qo->out_val = (long long *)ctx->d_out_val;
  qo->out_cap = ctx->out_cap;
  qo->qv = q ? (T)q[level] : (T)0;
  qo->vol = vol ? (T)vol[level] : (T)0;
  qo->shift = (long long)(ctx->dict / 2);
  qo->dict = ctx->dict;
}

// decompose: src (interleaved nodes of level L, full strides, read only) ->
// coefficients.  QUANT: symbols in ctx->d_sym + histogram + outliers;
// otherwise float coefficients in v (reordered layout).  src != v.
template <typename T, bool QUANT>
int decompose_rows(mgx_ctx *ctx, const T *src, T *v, const double *q,
                   const double *vol) {
  Scratch s;
  int rc = get_scratch(ctx, &s);
  if (rc)
    return rc;
  const MgxLevels &lv = ctx->lv;
  const T *tables = (const T *)ctx->d_tables;
  const size_t smem = QUANT ? (size_t)ctx->dict * 4 : 0;
  if (QUANT)
    MGX_CUDA(cudaFuncSetAttribute(k_level_forward<T, QUANT>,
                                  cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  (int)smem));
  QuantOut<T> qo;
  const T *cur = src;
  int which = 0;
  for (int l = lv.L; l >= 1; l--) {
    const uint32_t ncf = (uint32_t)lv.n[l - 1][FD];
    unsigned gx = (ncf + FW_WARPS * FW_OWN - 1) / (FW_WARPS * FW_OWN);
    LvGeom g = level_geom<T>(ctx, l, ctx->num_sms * 6, gx);
    fill_quant<T>(ctx, &qo, q, vol, l);
    T *A = (T *)s.A[which];
    k_level_forward<T, QUANT><<<dim3(gx, grid_y(g)), FW_WARPS * 32, smem,
                                ctx->stream>>>(cur, v, A, (T *)s.R[0], tables, g, qo);
    ctx->launches++;
    rc = correction_tail<T>(ctx, l, (T *)s.R[0], (T *)s.R[1], A, +1);
    if (rc)
      return rc;
    cur = A;
    which ^= 1;
  }
  // level-0 nodes
  LvGeom g0;
  for (int d = 0; d < MGX_MAX_D; d++) {
    g0.nc[d] = (uint32_t)lv.n[0][d];
    g0.vst[d] = lv.stride[d];
  }
  fill_quant<T>(ctx, &qo, q, vol, 0);
  uint64_t n0 = 1;
  for (int d = 0; d < MGX_MAX_D; d++)
    n0 *= lv.n[0][d];
  const unsigned cb_grid = mgx_grid_for(n0, 256, ctx->num_sms * 4);
  if (lv.L == 0) {
    // nothing was decomposed: the level-0 box is the array itself
    if (QUANT)
      k_coarse_box<T, QUANT, false><<<cb_grid, 256, 0, ctx->stream>>>(
          const_cast<T *>(src), v, g0, qo);
    else
      MGX_CUDA(cudaMemcpyAsync(v, src, lv.N * sizeof(T), cudaMemcpyDeviceToDevice,
                               ctx->stream));
  } else {
    k_coarse_box<T, QUANT, false><<<cb_grid, 256, 0, ctx->stream>>>(
        const_cast<T *>(cur), v, g0, qo);
  }
  ctx->launches++;
  MGX_CUDA(cudaGetLastError());
  return MGX_SUCCESS;
}

// recompose: coefficients v (reordered, read only) -> dst (full array), dst != v
template <typename T> int recompose_rows(mgx_ctx *ctx, const T *v, T *dst) {
  Scratch s;
  int rc = get_scratch(ctx, &s);
  if (rc)
    return rc;
  const MgxLevels &lv = ctx->lv;
  const T *tables = (const T *)ctx->d_tables;
  QuantOut<T> qo;
  fill_quant<T>(ctx, &qo, nullptr, nullptr, 0);
  if (lv.L == 0) {
    MGX_CUDA(cudaMemcpyAsync(dst, v, lv.N * sizeof(T), cudaMemcpyDeviceToDevice,
                             ctx->stream));
    return MGX_SUCCESS;
  }
  LvGeom g0;
  for (int d = 0; d < MGX_MAX_D; d++) {
    g0.nc[d] = (uint32_t)lv.n[0][d];
    g0.vst[d] = lv.stride[d];
  }
  uint64_t n0 = 1;
  for (int d = 0; d < MGX_MAX_D; d++)
    n0 *= lv.n[0][d];
  const unsigned cb_grid = mgx_grid_for(n0, 256, ctx->num_sms * 4);
  int which = 0;
  T *A = (T *)s.A[which];
  k_coarse_box<T, false, true><<<cb_grid, 256, 0, ctx->stream>>>(
      A, const_cast<T *>(v), g0, qo);
  ctx->launches++;
  for (int l = 1; l <= lv.L; l++) {
    const uint32_t ncf = (uint32_t)lv.n[l - 1][FD];
    unsigned gx = (ncf + 255) / 256;
    LvGeom g = level_geom<T>(ctx, l, ctx->num_sms * 6, gx);
    k_lpk_first<T><<<dim3(gx, grid_y(g)), 256, 0, ctx->stream>>>(
        v, (T *)s.R[0], tables, g);
    ctx->launches++;
    rc = correction_tail<T>(ctx, l, (T *)s.R[0], (T *)s.R[1], A, -1);
    if (rc)
      return rc;
    T *out = (l == lv.L) ? dst : (T *)s.A[which ^ 1];
    k_level_restore_rows<T><<<dim3(gx, grid_y(g)), 256, 0, ctx->stream>>>(
        v, A, out, tables, g);
    ctx->launches++;
    A = out;
    which ^= 1;
  }
  MGX_CUDA(cudaGetLastError());
  return MGX_SUCCESS;
}

} // namespace

// multi_dimension::decompose with orthogonal projection, out of place
int mgxk_decompose_rows(mgx_ctx *ctx, const void *d_src, void *d_coeff) {
  return mgx_is_f32(ctx)
             ? decompose_rows<float, false>(ctx, (const float *)d_src,
                                            (float *)d_coeff, nullptr, nullptr)
             : decompose_rows<double, false>(ctx, (const double *)d_src,
                                             (double *)d_coeff, nullptr, nullptr);
}

// decompose + lwqzk + outlier separation + histogram: field -> ctx->d_sym
int mgxk_decompose_quantize_rows(mgx_ctx *ctx, const void *d_src, int rel,
                                 double tol, double s, double norm) {
  double q[MGX_MAX_L], vol[MGX_MAX_L];
  mgx_level_quantizers(ctx, rel, tol, s, norm, true, q, vol);
  return mgx_is_f32(ctx)
             ? decompose_rows<float, true>(ctx, (const float *)d_src, nullptr, q, vol)
             : decompose_rows<double, true>(ctx, (const double *)d_src, nullptr, q,
                                            vol);
}

int mgxk_recompose_rows(mgx_ctx *ctx, const void *d_coeff, void *d_dst) {
  return mgx_is_f32(ctx)
             ? recompose_rows<float>(ctx, (const float *)d_coeff, (float *)d_dst)
             : recompose_rows<double>(ctx, (const double *)d_coeff, (double *)d_dst);
}
