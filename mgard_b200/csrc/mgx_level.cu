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
  uint32_t rows_per_task;   // a warp task = one column chunk x this many rows
  uint32_t col_chunks;      // column chunks per row
  uint64_t tasks;           // col_chunks * ceil(rows / rows_per_task)
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
// A warp owns a contiguous range of rows of the level box in REORDERED row
// order (p[d] < nc[d]: coarse node, else coefficient) and one chunk of columns.
// For the current row:
//   so/vo/ao  offsets of the row in the node array / full array / coarse box A
//   nodd      number of odd (coefficient) outer dimensions; hi[j] element offset
//             to the upper neighbour row along the j-th odd dimension (innermost
//             first), tt[j] its interpolation ratio.  FORWARD: neighbours are
//             rows q -/+ 1 of the interleaved node array (lower = -hi);
//             otherwise rows (q-1)/2 and (q-1)/2 + 1 of A (lower = +0).
// Everything is indexed with compile-time constants so the state stays in
// registers.
template <typename T> struct Row {
  uint64_t so, vo, ao;
  int nodd;
  int64_t hi[OUTER];
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

__device__ __forceinline__ void row_first(uint32_t (&p)[OUTER], const LvGeom &g,
                                          uint64_t row) {
#pragma unroll
  for (int d = OUTER - 1; d >= 0; d--) {
    p[d] = (uint32_t)(row % g.nf[d]);
    row /= g.nf[d];
  }
}

__device__ __forceinline__ void row_next(uint32_t (&p)[OUTER], const LvGeom &g) {
  bool carry = true;
#pragma unroll
  for (int d = OUTER - 1; d >= 0; d--) {
    if (carry) {
      p[d]++;
      carry = p[d] == g.nf[d];
      if (carry)
        p[d] = 0;
    }
  }
}

template <typename T, bool FORWARD, bool SKIP>
__device__ __forceinline__ void row_setup(Row<T> &r, const uint32_t (&p)[OUTER],
                                          const LvGeom &g,
                                          const T *__restrict__ tables) {
  r.so = 0;
  r.vo = 0;
  r.ao = 0;
  r.nodd = 0;
#pragma unroll
  for (int j = 0; j < OUTER; j++) {
    r.hi[j] = 0;
    r.tt[j] = (T)0.5;
  }
  // outermost dimension first, each odd one pushed to the front: the
  // innermost odd dimension ends at index 0
#pragma unroll
  for (int d = 0; d < OUTER; d++) {
    if (SKIP && g.nf[d] == 1)
      continue; // padded dimension (D < 5): contributes nothing (the SKIP
                // variants serve D < 5; with D = 5 the test only costs)
    uint32_t q;
    bool odd;
    pos_to_node(p[d], g.nf[d], g.nc[d], q, odd);
    r.so += (uint64_t)q * g.sst[d];
    r.vo += (uint64_t)p[d] * g.vst[d];
    if (odd) {
#pragma unroll
      for (int j = OUTER - 1; j > 0; j--) {
        r.hi[j] = r.hi[j - 1];
        r.tt[j] = r.tt[j - 1];
      }
      r.hi[0] = (int64_t)(FORWARD ? g.sst[d] : g.ast[d]);
      r.tt[0] = g.half ? (T)0.5 : tables[g.roff[d] + q - 1];
      r.nodd++;
      if (!FORWARD)
        r.ao += (uint64_t)((q - 1) / 2) * g.ast[d];
    } else {
      r.ao += (uint64_t)p[d] * g.ast[d];
    }
  }
}

// Interpolant of the 2^K neighbour rows at one even column (a) and at the odd
// column to its right (b, interpolated along the contiguous dim FIRST, then
// along the odd outer dims innermost to outermost -- the reference's f, c, r
// order, GridProcessingKernel3D.hpp).
template <typename T, int K, bool FORWARD>
__device__ __forceinline__ void corners(const T *__restrict__ base, int64_t off,
                                        const Row<T> &r, uint32_t e0, uint32_t e1,
                                        bool has_odd, T tf, T &a, T &b) {
  if constexpr (K == 0) {
    T x0 = base[off + e0];
    a = x0;
    b = has_odd ? mgx_lerp(x0, base[off + e1], tf) : (T)0;
  } else {
    T a0, b0, a1, b1;
    corners<T, K - 1, FORWARD>(base, FORWARD ? off - r.hi[K - 1] : off, r, e0, e1,
                               has_odd, tf, a0, b0);
    corners<T, K - 1, FORWARD>(base, off + r.hi[K - 1], r, e0, e1, has_odd, tf, a1,
                               b1);
    a = mgx_lerp(a0, a1, r.tt[K - 1]);
    b = mgx_lerp(b0, b1, r.tt[K - 1]);
  }
}

// Forward variant, called by ALL lanes of the warp: a lane loads only its even
// column; the even column to the right of its odd one is the next lane's even
// column and arrives by shuffle (lane 31 is a halo pair whose odd column is
// never used).  Halves the neighbour-row loads.  (The same trick on the
// inverse kernel, which has no halo lanes, measured slightly slower.)
template <typename T, int K>
__device__ __forceinline__ void corners_fw(const T *__restrict__ base, int64_t off,
                                           const Row<T> &r, uint32_t e0, T tf, T &a,
                                           T &b) {
  if constexpr (K == 0) {
    const T x0 = base[off + e0];
    const T x1 = __shfl_down_sync(0xffffffffu, x0, 1);
    a = x0;
    b = mgx_lerp(x0, x1, tf);
  } else {
    T a0, b0, a1, b1;
    corners_fw<T, K - 1>(base, off - r.hi[K - 1], r, e0, tf, a0, b0);
    corners_fw<T, K - 1>(base, off + r.hi[K - 1], r, e0, tf, a1, b1);
    a = mgx_lerp(a0, a1, r.tt[K - 1]);
    b = mgx_lerp(b0, b1, r.tt[K - 1]);
  }
}

// L2 -> L1 prefetch of what the NEXT row will load (the row loop is one
// dependent load -> compute -> store chain per warp; without this every row
// pays a full memory latency)
__device__ __forceinline__ void prefetch_l1(const void *p) {
  asm volatile("prefetch.global.L1 [%0];" ::"l"(p));
}
template <typename T, int K, bool FORWARD>
__device__ __forceinline__ void prefetch_corners(const T *__restrict__ base,
                                                 int64_t off, const Row<T> &r) {
  if constexpr (K == 0) {
    prefetch_l1(base + off);
  } else {
    prefetch_corners<T, K - 1, FORWARD>(base, FORWARD ? off - r.hi[K - 1] : off, r);
    prefetch_corners<T, K - 1, FORWARD>(base, off + r.hi[K - 1], r);
  }
}
template <typename T, bool FORWARD>
__device__ __forceinline__ void prefetch_row(const T *__restrict__ base, int64_t off,
                                             const Row<T> &r) {
  switch (r.nodd) {
  case 0:
    break;
  case 1:
    prefetch_corners<T, 1, FORWARD>(base, off, r);
    break;
  case 2:
    prefetch_corners<T, 2, FORWARD>(base, off, r);
    break;
  case 3:
    prefetch_corners<T, 3, FORWARD>(base, off, r);
    break;
  default:
    prefetch_corners<T, 4, FORWARD>(base, off, r);
    break;
  }
}

// coefficient -> float store, or -> symbol + histogram + outlier list
// (lwqzk LinearQuantization.hpp:96-111, OutlierSeparator.hpp:17-120,
// Histogram.hpp).  Called by all lanes of a warp.
__device__ __forceinline__ int to_int_sat(float r) { return __float2int_rz(r); }
__device__ __forceinline__ int to_int_sat(double r) { return __double2int_rz(r); }

template <typename T, bool QUANT>
__device__ __forceinline__ void emit(T c, bool valid, uint64_t idx,
                                     T *__restrict__ v, const QuantOut<T> &qo,
                                     uint32_t *s_hist) {
  if constexpr (!QUANT) {
    if (valid)
      v[idx] = c;
  } else {
    if (valid) {
      // 32-bit fast path: the saturating conversion sends every magnitude that
      // does not fit an int out of the dictionary range, where the exact
      // 64-bit code is formed for the outlier list
      const T r = add_rn((T)0.5, fabs(mul_rn(mul_rn(c, qo.qv), qo.vol)));
      const int qi = to_int_sat(r);
      int sv = (c < (T)0 ? -qi : qi) + (int)qo.shift;
      if (__builtin_expect((unsigned)sv >= qo.dict, 0)) {
        const long long q = mgx_quantize(c, qo.qv, qo.vol, qo.shift);
        unsigned long long slot = atomicAdd(qo.out_count, 1ull);
        if (slot < qo.out_cap) {
          qo.out_idx[slot] = idx;
          qo.out_val[slot] = q;
        }
        sv = 0;
      }
      qo.sym[idx] = (uint16_t)sv;
      atomicAdd(&s_hist[sv], 1u); // ATOMS.POPC.INC merges lanes on one bin
    }
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
  // every lane loads (lanes without a pair read column 0 and are ignored), so
  // that the shuffles below are warp-wide
  T ce = 0, co = 0;
  {
    const T *srow = src + r.so;
    const T ve = srow[qe];
    const T vo = has_odd ? srow[2 * i + 1] : (T)0;
    if constexpr (K == 0) {
      const T e2 = __shfl_down_sync(0xffffffffu, ve, 1); // node 2i + 2
      if (owned && pair_ok)
        A[r.ao + i] = ve; // coarse node of the next level
      if (has_odd)
        co = sub_rn(vo, mgx_lerp(ve, e2, tf));
    } else {
      T a, b;
      corners_fw<T, K>(src, (int64_t)r.so, r, qe, tf, a, b);
      if (pair_ok)
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

template <typename T, bool QUANT, int MINB, bool SKIP>
__global__ void __launch_bounds__(FW_WARPS * 32, MINB)
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
  const uint64_t task = (uint64_t)blockIdx.x * FW_WARPS + warp;
  if (task < g.tasks) {
    const uint32_t cw = (uint32_t)(task % g.col_chunks);
    const uint64_t rg = task / g.col_chunks;
    const int64_t i = (int64_t)cw * FW_OWN - 1 + lane;
    const bool pair_ok = i >= 0 && i < (int64_t)ncf;
    const bool has_odd = i >= 0 && i < (int64_t)ncoef;
    const bool owned = lane >= 1 && lane <= FW_OWN;
    const uint32_t qe =
        pair_ok ? ((i == (int64_t)ncf - 1) ? nff - 1 : (uint32_t)(2 * i)) : 0;
    const T tf = (g.half || !has_odd) ? (T)0.5 : tables[g.roff[FD] + 2 * i];
    const MassK<T> mk = mass_consts<T>(tables + g.dist_f, pair_ok ? 2 * i : 0, nff);
    const uint64_t row0 = rg * g.rows_per_task;
    const uint64_t row1 = min(g.rows, row0 + g.rows_per_task);
    uint32_t p[OUTER];
    row_first(p, g, row0);
    Row<T> r;
    row_setup<T, true, SKIP>(r, p, g, tables);
    for (uint64_t row = row0; row < row1; row++) {
      const bool more = row + 1 < row1;
      Row<T> nx = r;
      if (more) {
        row_next(p, g);
        row_setup<T, true, SKIP>(nx, p, g, tables);
        if (pair_ok) {
          prefetch_l1(src + nx.so + qe);
          prefetch_row<T, true>(src, (int64_t)nx.so + qe, nx);
        }
      }
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
      r = nx;
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

// Where recompose reads a level's coefficients from: the float array of the
// staged API, or -- on the fused decompress path -- the 16-bit symbols, which are
// dequantised on the fly (lwqzk inverse, LinearQuantization.hpp:113-128:
// (quantizer * volume) * (T)(code - dict/2), a per-level scalar times the code).
template <typename T, bool SYM> struct Coef {
  const T *v;
  const uint16_t *sym;
  T dq;      // quantizer * volume of this level
  int shift; // dict / 2
  __device__ __forceinline__ T at(uint64_t idx) const {
    if constexpr (SYM)
      return mul_rn(dq, (T)((int)sym[idx] - shift));
    else
      return v[idx];
  }
  __device__ __forceinline__ void prefetch(uint64_t idx) const {
    if constexpr (SYM)
      prefetch_l1(sym + idx);
    else
      prefetch_l1(v + idx);
  }
};

// ---- first restriction sweep from stored coefficients (recompose) ------------------
template <typename T, bool SYM>
__global__ void __launch_bounds__(256)
    k_lpk_first(const Coef<T, SYM> cf, T *__restrict__ R,
                const T *__restrict__ tables, const __grid_constant__ LvGeom g) {
  const uint32_t ncf = g.nc[FD], nff = g.nf[FD], ncoef = nff - ncf;
  const int lane = threadIdx.x & 31;
  const uint64_t task = (uint64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (task >= g.tasks)
    return;
  const uint32_t cw = (uint32_t)(task % g.col_chunks);
  const uint64_t rg = task / g.col_chunks;
  const int64_t i = (int64_t)cw * 32 + lane;
  if (i >= (int64_t)ncf)
    return;
  MassK<T> mk = mass_consts<T>(tables + g.dist_f, 2 * i, nff);
  const uint64_t row0 = rg * g.rows_per_task;
  const uint64_t row1 = min(g.rows, row0 + g.rows_per_task);
  uint32_t p[OUTER];
  row_first(p, g, row0);
  for (uint64_t row = row0; row < row1; row++) {
    bool coarse_row = true;
    uint64_t vo = 0;
#pragma unroll
    for (int d = 0; d < OUTER; d++) {
      coarse_row = coarse_row && p[d] < g.nc[d];
      vo += (uint64_t)p[d] * g.vst[d];
    }
    row_next(p, g);
    if (row + 1 < row1) {
      uint64_t vn = 0;
#pragma unroll
      for (int d = 0; d < OUTER; d++)
        vn += (uint64_t)p[d] * g.vst[d];
      cf.prefetch(vn + i);
      if (i < (int64_t)ncoef)
        cf.prefetch(vn + ncf + i);
    }
    T a = 0, b = 0, c = 0, d = 0, e = 0;
    if (!coarse_row) {
      a = i >= 1 ? cf.at(vo + i - 1) : (T)0;
      c = cf.at(vo + i);
      e = i + 1 < (int64_t)ncf ? cf.at(vo + i + 1) : (T)0;
    }
    b = (i >= 1 && i - 1 < (int64_t)ncoef) ? cf.at(vo + ncf + i - 1) : (T)0;
    d = i < (int64_t)ncoef ? cf.at(vo + ncf + i) : (T)0;
    R[row * ncf + (uint64_t)i] = mass_apply(mk, a, b, c, d, e);
  }
}

// ---- mass x restriction along a non-contiguous dimension -----------------------
// in : (outer, nf [nc coarse | nf-nc coefficients], inner) compact
// out: (outer, nc, inner) compact.  A block owns one output index i along the
// dimension (its nine quotients are formed once) and streams (outer, inner).
// (Lpk2Reo3D / Lpk3Reo3D, LinearProcessingKernel3D.hpp:374-1097)
template <typename T, int U>
__global__ void __launch_bounds__(256)
    k_lpk_mid(const T *__restrict__ in, T *__restrict__ out,
              const T *__restrict__ dist, uint32_t outer, uint32_t nf, uint32_t nc,
              uint64_t inner) {
  const uint32_t i = blockIdx.x;
  const uint32_t ncoef = nf - nc;
  const MassK<T> mk = mass_consts<T>(dist, 2 * (int64_t)i, nf);
  const bool ha = i >= 1, hb = i >= 1 && i - 1 < ncoef, hd = i < ncoef,
             he = i + 1 < nc;
  // U `outer` slices per iteration: 5 U independent loads in flight per thread
  for (uint32_t o0 = blockIdx.z * U; o0 < outer; o0 += gridDim.z * U) {
    for (uint64_t x = (uint64_t)blockIdx.y * blockDim.x + threadIdx.x; x < inner;
         x += (uint64_t)gridDim.y * blockDim.x) {
      T a[U], b[U], c[U], d[U], e[U];
#pragma unroll
      for (int u = 0; u < U; u++) {
        const uint32_t o = min(o0 + u, outer - 1);
        const T *ib = in + (uint64_t)o * nf * inner + x;
        a[u] = ha ? ib[(uint64_t)(i - 1) * inner] : (T)0;
        b[u] = hb ? ib[(uint64_t)(nc + i - 1) * inner] : (T)0;
        c[u] = ib[(uint64_t)i * inner];
        d[u] = hd ? ib[(uint64_t)(nc + i) * inner] : (T)0;
        e[u] = he ? ib[(uint64_t)(i + 1) * inner] : (T)0;
      }
#pragma unroll
      for (int u = 0; u < U; u++)
        if (o0 + u < outer)
          out[((uint64_t)(o0 + u) * nc + i) * inner + x] =
              mass_apply(mk, a[u], b[u], c[u], d[u], e[u]);
    }
  }
}

// ---- batched Thomas solves -------------------------------------------------------
// reference Ipk1/2/3Reo (IterativeProcessingKernel3D.hpp:28-1169) with
// tridiag_forward2 / tridiag_backward2 (IPKFunctor.h:13-50):
//   forward : x[i] = x[i] - x[i-1] * (am[i] / bm[i])        (abr = am/bm table)
//   backward: x[i] = (x[i] - am[i+1] * x[i+1]) / bm[i+1]
// ACC != 0: the backward sweep does not store x but adds (sign > 0) or
// subtracts (sign < 0) it into acc at the same compact index -- the reference's
// separate Lwpk ADD / SUBTRACT pass (LevelwiseProcessingKernel.hpp:21-127).
//
// The recurrences are sequential per line, and bit-exactness rules out cyclic
// reduction, so a 2-D 8193^2 level has only ~4k independent chains: the kernels
// are bound by the latency of one chain step.  Two things keep that step short:
//   * data moves through a per-thread cp.async ring (dozens of rows in flight
//     per thread, none of it on the dependent chain);
//   * the division is replaced by an exactly rounded equivalent that uses the
//     precomputed reciprocal y = RN(1/b):
//       fp64: q0 = a*y; two residual corrections q <- fma(fma(-b,q,a), y, q).
//             After the first q is faithful, and then the second is the
//             correctly rounded a/b (Markstein's theorem; y correctly rounded,
//             b's significand not all ones -- checked at plan time; operands
//             outside 1e-270..1e270 take the true division);
//       fp32: RN32((double)a * RN64(1/b)): the product is within 2^-52 of a/b
//             and no quotient of two 24-bit significands lies within 2^-49 of a
//             25-bit midpoint, so the rounding equals that of the true quotient.
//     172 -> ~75 cycles per fp64 step on B200.
// fast quotient (valid when range_bad() is 0 for the dividend) ...
__device__ __forceinline__ double div_fast(double a, double b, double y) {
  double q = __dmul_rn(a, y);
  q = __fma_rn(__fma_rn(-b, q, a), y, q);
  return __fma_rn(__fma_rn(-b, q, a), y, q);
}
__device__ __forceinline__ float div_fast(float a, float, double y) {
  return __double2float_rn(__dmul_rn((double)a, y));
}
// ... and its validity test, integer only so that it stays off the dependent
// fp64 chain: exponent outside [2^-897, 2^897] and not an exact zero
__device__ __forceinline__ unsigned range_bad(double a) {
  const unsigned hi = (unsigned)__double2hiint(a), lo = (unsigned)__double2loint(a);
  const unsigned e = (hi >> 20) & 0x7ffu;
  return ((e - 126u) > 1794u) && (((hi & 0x7fffffffu) | lo) != 0u);
}
__device__ __forceinline__ unsigned range_bad(float) { return 0u; }

// NB steps of the backward recurrence x[i] = (c - am*x[i+1]) / bm, last step
// first in c[]; results in res[].  The fast quotients are verified per batch
// (off the chain) and the batch is redone with true divisions if need be.
template <typename T, int NB>
__device__ __forceinline__ T backward_batch(const T (&c)[NB], const T (&a1)[NB],
                                            const T (&b1)[NB],
                                            const double (&y1)[NB], int cnt,
                                            T prev, bool fast, T (&res)[NB]) {
  const T prev0 = prev;
  unsigned bad = fast ? 0u : 1u;
#pragma unroll
  for (int k = 0; k < NB; k++)
    if (k < cnt) {
      T t = sub_rn(c[k], mul_rn(a1[k], prev));
      bad |= range_bad(t);
      prev = div_fast(t, b1[k], y1[k]);
      res[k] = prev;
    }
  if (bad) {
    prev = prev0;
#pragma unroll
    for (int k = 0; k < NB; k++)
      if (k < cnt) {
        prev = div_rn(sub_rn(c[k], mul_rn(a1[k], prev)), b1[k]);
        res[k] = prev;
      }
  }
  return prev;
}

template <int BYTES>
__device__ __forceinline__ void cp_async(void *smem, const void *gmem) {
  unsigned sa = (unsigned)__cvta_generic_to_shared(smem);
  asm volatile("cp.async.ca.shared.global [%0], [%1], %2;\n" ::"r"(sa), "l"(gmem),
               "n"(BYTES)
               : "memory");
}
__device__ __forceinline__ void cp_commit() {
  asm volatile("cp.async.commit_group;\n" ::: "memory");
}
template <int N> __device__ __forceinline__ void cp_wait() {
  asm volatile("cp.async.wait_group %0;\n" ::"n"(N) : "memory");
}

// Strided variant (solve dimension not contiguous): a thread owns one line,
// consecutive threads consecutive `inner` indices.  Ring: S stages of B rows.
constexpr int IPK_B = 8;

template <typename T, bool ACC, int IPK_S>
__global__ void __launch_bounds__(128)
    k_ipk_strided(T *__restrict__ x, const T *__restrict__ am,
                  const T *__restrict__ bm, const T *__restrict__ abr,
                  const double *__restrict__ rbm, uint32_t n, uint64_t inner,
                  uint64_t lines, T *__restrict__ acc, int sign, int fast) {
  extern __shared__ __align__(16) unsigned char ipk_smem[];
  const uint32_t tpb = blockDim.x, tid = threadIdx.x;
  const uint32_t lane = tid & 31, warp = tid >> 5, nwarp = tpb >> 5;
  uint64_t ln = blockIdx.x * (uint64_t)blockDim.x + tid;
  // lanes past the last line shadow it (no stores) so that warps stay whole
  const bool active = ln < lines;
  if (!active)
    ln = lines - 1;
  // ring of S stages: B rows of x (and of acc), plus the table entries of
  // those B rows per warp -- a table entry is used once per SM, so without
  // the ring every batch would wait on an L1-missing table load
  double *ty = reinterpret_cast<double *>(ipk_smem) + (size_t)warp * IPK_S * IPK_B;
  T *tt = reinterpret_cast<T *>(reinterpret_cast<double *>(ipk_smem) +
                                (size_t)nwarp * IPK_S * IPK_B) +
          (size_t)warp * IPK_S * 2 * IPK_B;
  T *rx = reinterpret_cast<T *>(reinterpret_cast<double *>(ipk_smem) +
                                (size_t)nwarp * IPK_S * IPK_B) +
          (size_t)nwarp * IPK_S * 2 * IPK_B;
  T *ra = rx + (size_t)IPK_S * IPK_B * tpb;
  const uint64_t o = ln / inner, in = ln % inner;
  T *p = x + o * n * inner + in;
  T *pa = ACC ? acc + o * n * inner + in : nullptr;
  const uint32_t nb = (n + IPK_B - 1) / IPK_B;
  // ---- forward ----
  auto issue_f = [&](uint32_t j) {
    if (j < nb) {
      const uint32_t slot = j % IPK_S;
#pragma unroll
      for (int k = 0; k < IPK_B; k++) {
        uint32_t i = j * IPK_B + k;
        if (i < n)
          cp_async<sizeof(T)>(&rx[(slot * IPK_B + k) * tpb + tid],
                              p + (uint64_t)i * inner);
      }
      if (lane < IPK_B && j * IPK_B + lane < n)
        cp_async<sizeof(T)>(&tt[slot * 2 * IPK_B + lane], abr + j * IPK_B + lane);
    }
    cp_commit();
  };
  for (uint32_t j = 0; j < IPK_S - 1; j++)
    issue_f(j);
  T prev = 0;
  for (uint32_t j = 0; j < nb; j++) {
    __syncwarp(); // everyone is done with the slot about to be refilled
    issue_f(j + IPK_S - 1);
    cp_wait<IPK_S - 1>();
    __syncwarp();
    const uint32_t slot = j % IPK_S;
    T c[IPK_B], f[IPK_B];
#pragma unroll
    for (int k = 0; k < IPK_B; k++) {
      c[k] = rx[(slot * IPK_B + k) * tpb + tid];
      f[k] = tt[slot * 2 * IPK_B + k];
    }
#pragma unroll
    for (int k = 0; k < IPK_B; k++) {
      uint32_t i = j * IPK_B + k;
      if (i < n) {
        prev = sub_rn(c[k], mul_rn(prev, f[k]));
        if (active)
          p[(uint64_t)i * inner] = prev;
      }
    }
  }
  cp_wait<0>();
  __threadfence_block(); // the backward sweep re-reads what was just stored
  __syncwarp();
  // ---- backward: batch j covers i = n-1-(j*B+k) ----
  auto issue_b = [&](uint32_t j) {
    if (j < nb) {
      const uint32_t slot = j % IPK_S;
#pragma unroll
      for (int k = 0; k < IPK_B; k++) {
        uint32_t jj = j * IPK_B + k;
        if (jj < n) {
          uint64_t off = (uint64_t)(n - 1 - jj) * inner;
          cp_async<sizeof(T)>(&rx[(slot * IPK_B + k) * tpb + tid], p + off);
          if (ACC)
            cp_async<sizeof(T)>(&ra[(slot * IPK_B + k) * tpb + tid], pa + off);
        }
      }
      const uint32_t jj = j * IPK_B + lane;
      if (lane < IPK_B && jj < n) {
        const uint32_t i1 = n - jj; // i + 1
        cp_async<sizeof(T)>(&tt[slot * 2 * IPK_B + lane], am + i1);
        cp_async<sizeof(T)>(&tt[slot * 2 * IPK_B + IPK_B + lane], bm + i1);
        cp_async<8>(&ty[slot * IPK_B + lane], rbm + i1);
      }
    }
    cp_commit();
  };
  for (uint32_t j = 0; j < IPK_S - 1; j++)
    issue_b(j);
  prev = 0;
  for (uint32_t j = 0; j < nb; j++) {
    __syncwarp();
    issue_b(j + IPK_S - 1);
    cp_wait<IPK_S - 1>();
    __syncwarp();
    const uint32_t slot = j % IPK_S;
    T c[IPK_B], a1[IPK_B], b1[IPK_B], av[IPK_B], res[IPK_B];
    double y1[IPK_B];
    const int cnt = (int)min((uint32_t)IPK_B, n - j * IPK_B);
#pragma unroll
    for (int k = 0; k < IPK_B; k++) {
      c[k] = rx[(slot * IPK_B + k) * tpb + tid];
      av[k] = ACC ? ra[(slot * IPK_B + k) * tpb + tid] : (T)0;
      a1[k] = tt[slot * 2 * IPK_B + k];
      b1[k] = tt[slot * 2 * IPK_B + IPK_B + k];
      y1[k] = ty[slot * IPK_B + k];
    }
    prev = backward_batch<T, IPK_B>(c, a1, b1, y1, cnt, prev, fast != 0, res);
    if (active) {
#pragma unroll
      for (int k = 0; k < IPK_B; k++)
        if (k < cnt) {
          uint32_t i = n - 1 - (j * IPK_B + k);
          if (ACC)
            pa[(uint64_t)i * inner] =
                sign > 0 ? add_rn(av[k], res[k]) : sub_rn(av[k], res[k]);
          else
            p[(uint64_t)i * inner] = res[k];
        }
    }
  }
  cp_wait<0>();
}

// Contiguous variant: one warp owns 32 lines; 32x32 tiles are fetched row-wise
// (coalesced) by cp.async into a ring of shared-memory tiles and walked
// transposed, lane r on line r; results go back row-wise.
constexpr int IPC_S = 4;

template <typename T>
__global__ void __launch_bounds__(32)
    k_ipk_contig(T *__restrict__ x, const T *__restrict__ am,
                 const T *__restrict__ bm, const T *__restrict__ abr,
                 const double *__restrict__ rbm, uint32_t n, uint64_t lines,
                 T *__restrict__ acc, int sign, int fast) {
  __shared__ T tile[IPC_S][32][33];
  __shared__ T tabs[IPC_S][2][32];  // abr | am, bm of the tile's 32 columns
  __shared__ double taby[IPC_S][32]; // 1 / bm
  const int lane = threadIdx.x;
  const uint64_t l0 = (uint64_t)blockIdx.x * 32;
  if (l0 >= lines)
    return;
  const uint32_t nl = (uint32_t)min((uint64_t)32, lines - l0);
  const uint32_t nch = (n + 31) / 32;
  T *xb = x + l0 * n;
  T *ab = acc ? acc + l0 * n : nullptr;
  auto issue = [&](uint32_t ch, bool backward) {
    if (ch < nch) {
      const uint32_t c0 = (backward ? nch - 1 - ch : ch) * 32;
      const uint32_t slot = ch % IPC_S;
      T(*t)[33] = tile[slot];
      if (c0 + lane < n) {
        for (uint32_t r = 0; r < nl; r++)
          cp_async<sizeof(T)>(&t[r][lane], xb + (uint64_t)r * n + c0 + lane);
        if (!backward) {
          cp_async<sizeof(T)>(&tabs[slot][0][lane], abr + c0 + lane);
        } else {
          cp_async<sizeof(T)>(&tabs[slot][0][lane], am + c0 + lane + 1);
          cp_async<sizeof(T)>(&tabs[slot][1][lane], bm + c0 + lane + 1);
          cp_async<8>(&taby[slot][lane], rbm + c0 + lane + 1);
        }
      }
    }
    cp_commit();
  };
  // ---- forward ----
  for (uint32_t ch = 0; ch < IPC_S - 1; ch++)
    issue(ch, false);
  T prev = 0;
  for (uint32_t ch = 0; ch < nch; ch++) {
    issue(ch + IPC_S - 1, false);
    cp_wait<IPC_S - 1>();
    __syncwarp();
    const uint32_t c0 = ch * 32;
    const uint32_t ncol = min(32u, n - c0);
    T(*t)[33] = tile[ch % IPC_S];
    const T *fa = tabs[ch % IPC_S][0];
    if ((uint32_t)lane < nl) {
      if (ncol == 32) {
#pragma unroll 8
        for (uint32_t j = 0; j < 32; j++) {
          prev = sub_rn(t[lane][j], mul_rn(prev, fa[j]));
          t[lane][j] = prev;
        }
      } else {
        for (uint32_t j = 0; j < ncol; j++) {
          prev = sub_rn(t[lane][j], mul_rn(prev, fa[j]));
          t[lane][j] = prev;
        }
      }
    }
    __syncwarp();
    if ((uint32_t)lane < ncol)
      for (uint32_t r = 0; r < nl; r++)
        xb[(uint64_t)r * n + c0 + lane] = t[r][lane];
    __syncwarp();
  }
  cp_wait<0>();
  __threadfence_block();
  __syncwarp();
  // ---- backward ----
  for (uint32_t ch = 0; ch < IPC_S - 1; ch++)
    issue(ch, true);
  prev = 0;
  for (uint32_t ch = 0; ch < nch; ch++) {
    issue(ch + IPC_S - 1, true);
    cp_wait<IPC_S - 1>();
    __syncwarp();
    const uint32_t c0 = (nch - 1 - ch) * 32;
    const uint32_t ncol = min(32u, n - c0);
    const uint32_t slot = ch % IPC_S;
    T(*t)[33] = tile[slot];
    if ((uint32_t)lane < nl) {
      for (uint32_t j0 = 0; j0 < ncol; j0 += 8) {
        // sub-batch covers columns ncol-1-j0 ... down, 8 at a time
        T c[8], a1[8], b1[8], res[8];
        double y1[8];
        const int cnt = (int)min(8u, ncol - j0);
#pragma unroll
        for (int k = 0; k < 8; k++) {
          uint32_t j = k < cnt ? ncol - 1 - j0 - k : 0;
          c[k] = t[lane][j];
          a1[k] = tabs[slot][0][j];
          b1[k] = tabs[slot][1][j];
          y1[k] = taby[slot][j];
        }
        prev = backward_batch<T, 8>(c, a1, b1, y1, cnt, prev, fast != 0, res);
#pragma unroll
        for (int k = 0; k < 8; k++)
          if (k < cnt)
            t[lane][ncol - 1 - j0 - k] = res[k];
      }
    }
    __syncwarp();
    if ((uint32_t)lane < ncol) {
      if (ab) {
        for (uint32_t r = 0; r < nl; r++) {
          uint64_t off = (uint64_t)r * n + c0 + lane;
          ab[off] = sign > 0 ? add_rn(ab[off], t[r][lane]) : sub_rn(ab[off], t[r][lane]);
        }
      } else {
        for (uint32_t r = 0; r < nl; r++)
          xb[(uint64_t)r * n + c0 + lane] = t[r][lane];
      }
    }
    __syncwarp();
  }
  cp_wait<0>();
}

// ---- warp-specialised Thomas solves (scarce chains) ---------------------------------
// When a level has only a few thousand lines (2-D grids), every SM holds a
// single warp of chains and that warp's instruction stream IS the run time:
// measured 424 cycles per element with loads, table fetches, address
// arithmetic and stores issued around a 93-cycle dependent chain.  Here a block
// is three warps around a shared-memory ring of WS_S batches:
//   warp 2 (loader)  : cp.async of the batch's rows (+ acc rows + table
//                      entries) -> FULL[slot] once they have landed;
//   warp 0 (consumer): nothing but LDS -> chain -> STS -> DONE[slot];
//   warp 1 (drainer) : results -> global (added into `acc` on the last solve)
//                      -> FREE[slot], which lets the loader refill it.
// Synchronisation: one named barrier per slot and edge (bar.arrive / bar.sync
// over the two warps involved), no CTA-wide barrier.
constexpr int WS_B = 16, WS_S = 4;
constexpr int WS_FULL = 1, WS_DONE = 1 + WS_S, WS_FREE = 1 + 2 * WS_S; // barrier ids

__device__ __forceinline__ void bar_sync(int id) {
  asm volatile("bar.sync %0, 64;" ::"r"(id) : "memory");
}
__device__ __forceinline__ void bar_arrive(int id) {
  asm volatile("bar.arrive %0, 64;" ::"r"(id) : "memory");
}

template <typename T, bool ACC>
__global__ void __launch_bounds__(96)
    k_ipk_strided_ws(T *__restrict__ x, const T *__restrict__ am,
                     const T *__restrict__ bm, const T *__restrict__ abr,
                     const double *__restrict__ rbm, uint32_t n, uint64_t inner,
                     uint64_t lines, T *__restrict__ acc, int sign, int fast) {
  __shared__ T sx[WS_S][WS_B][32];
  __shared__ T sa[ACC ? WS_S : 1][WS_B][32];
  __shared__ T st[WS_S][2][WS_B];
  __shared__ double sy[WS_S][WS_B];
  const uint32_t lane = threadIdx.x & 31;
  const int role = threadIdx.x >> 5; // 0 consumer, 1 drainer, 2 loader
  uint64_t ln = (uint64_t)blockIdx.x * 32 + lane;
  const bool active = ln < lines;
  if (!active)
    ln = lines - 1;
  const uint64_t o = ln / inner, in = ln % inner;
  T *p = x + o * n * inner + in;
  T *pa = ACC ? acc + o * n * inner + in : nullptr;
  const uint32_t nb = (n + WS_B - 1) / WS_B;
  const int64_t step = (int64_t)inner; // elements between consecutive rows
#pragma unroll 1
  for (int bwd = 0; bwd < 2; bwd++) {
    // row r of the sweep is element (bwd ? n-1-r : r) of the line
    const int64_t dstep = bwd ? -step : step;
    if (role == 2) { // ---- loader ----
      const T *src = bwd ? p + (uint64_t)(n - 1) * inner : p;
      const T *srca = ACC ? (bwd ? pa + (uint64_t)(n - 1) * inner : pa) : nullptr;
      auto issue = [&](uint32_t j) {
        if (j < nb) {
          const uint32_t slot = j % WS_S;
          const bool whole = (j + 1) * WS_B <= n;
#pragma unroll
          for (int k = 0; k < WS_B; k++) {
            if (whole || j * WS_B + k < n) {
              cp_async<sizeof(T)>(&sx[slot][k][lane], src);
              if (ACC && bwd)
                cp_async<sizeof(T)>(&sa[slot][k][lane], srca);
            }
            src += dstep;
            if (ACC)
              srca += dstep;
          }
          const uint32_t r = j * WS_B + lane;
          if (lane < WS_B && r < n) {
            if (!bwd) {
              cp_async<sizeof(T)>(&st[slot][0][lane], abr + r);
            } else {
              const uint32_t i1 = n - r; // i + 1
              cp_async<sizeof(T)>(&st[slot][0][lane], am + i1);
              cp_async<sizeof(T)>(&st[slot][1][lane], bm + i1);
              cp_async<8>(&sy[slot][lane], rbm + i1);
            }
          }
        }
        cp_commit();
      };
      for (uint32_t j = 0; j < WS_S; j++)
        issue(j);
      cp_wait<WS_S - 1>();
      bar_arrive(WS_FULL + 0);
      for (uint32_t j = 0; j < nb; j++) {
        if (j + 1 < nb) {
          cp_wait<WS_S - 2>();
          bar_arrive(WS_FULL + (j + 1) % WS_S);
        }
        bar_sync(WS_FREE + j % WS_S); // batch j drained: its slot is free
        issue(j + WS_S);
      }
      cp_wait<0>();
    } else if (role == 1) { // ---- drainer ----
      T *dst = bwd ? p + (uint64_t)(n - 1) * inner : p;
      T *dsta = ACC ? (bwd ? pa + (uint64_t)(n - 1) * inner : pa) : nullptr;
      for (uint32_t j = 0; j < nb; j++) {
        const uint32_t slot = j % WS_S;
        const bool whole = (j + 1) * WS_B <= n;
        bar_sync(WS_DONE + slot);
#pragma unroll
        for (int k = 0; k < WS_B; k++) {
          if (active && (whole || j * WS_B + k < n)) {
            const T v = sx[slot][k][lane];
            if (ACC && bwd)
              *dsta = sign > 0 ? add_rn(sa[slot][k][lane], v)
                               : sub_rn(sa[slot][k][lane], v);
            else
              *dst = v;
          }
          dst += dstep;
          if (ACC)
            dsta += dstep;
        }
        if (!bwd)
          __threadfence_block(); // the backward sweep's loader re-reads these
        bar_arrive(WS_FREE + slot);
      }
    } else { // ---- consumer ----
      T prev = 0;
      for (uint32_t j = 0; j < nb; j++) {
        const uint32_t slot = j % WS_S;
        const int cnt = (int)min((uint32_t)WS_B, n - j * WS_B);
        bar_sync(WS_FULL + slot);
        if (!bwd) {
          T c[WS_B], f[WS_B];
#pragma unroll
          for (int k = 0; k < WS_B; k++) {
            c[k] = sx[slot][k][lane];
            f[k] = st[slot][0][k];
          }
#pragma unroll
          for (int k = 0; k < WS_B; k++)
            if (k < cnt) {
              prev = sub_rn(c[k], mul_rn(prev, f[k]));
              sx[slot][k][lane] = prev;
            }
        } else {
#pragma unroll
          for (int h = 0; h < WS_B; h += 8) {
            T c[8], a1[8], b1[8], res[8];
            double y1[8];
#pragma unroll
            for (int k = 0; k < 8; k++) {
              c[k] = sx[slot][h + k][lane];
              a1[k] = st[slot][0][h + k];
              b1[k] = st[slot][1][h + k];
              y1[k] = sy[slot][h + k];
            }
            prev = backward_batch<T, 8>(c, a1, b1, y1, cnt - h, prev, fast != 0, res);
#pragma unroll
            for (int k = 0; k < 8; k++)
              if (h + k < cnt)
                sx[slot][h + k][lane] = res[k];
          }
        }
        bar_arrive(WS_DONE + slot);
      }
    }
  }
}

template <typename T>
__global__ void __launch_bounds__(96)
    k_ipk_contig_ws(T *__restrict__ x, const T *__restrict__ am,
                    const T *__restrict__ bm, const T *__restrict__ abr,
                    const double *__restrict__ rbm, uint32_t n, uint64_t lines,
                    T *__restrict__ acc, int sign, int fast) {
  __shared__ T tile[WS_S][32][33];
  __shared__ T tabs[WS_S][2][32];
  __shared__ double taby[WS_S][32];
  const uint32_t lane = threadIdx.x & 31;
  const int role = threadIdx.x >> 5; // 0 consumer, 1 drainer, 2 loader
  const uint64_t l0 = (uint64_t)blockIdx.x * 32;
  const uint32_t nl = (uint32_t)min((uint64_t)32, lines - l0);
  const uint32_t nch = (n + 31) / 32;
  T *xb = x + l0 * n;
  T *ab = acc ? acc + l0 * n : nullptr;
#pragma unroll 1
  for (int bwd = 0; bwd < 2; bwd++) {
    if (role == 2) { // ---- loader ----
      auto issue = [&](uint32_t ch) {
        if (ch < nch) {
          const uint32_t c0 = (bwd ? nch - 1 - ch : ch) * 32;
          const uint32_t slot = ch % WS_S;
          if (c0 + lane < n) {
            const T *src = xb + c0 + lane;
            if (nl == 32) {
#pragma unroll 8
              for (uint32_t r = 0; r < 32; r++, src += n)
                cp_async<sizeof(T)>(&tile[slot][r][lane], src);
            } else {
              for (uint32_t r = 0; r < nl; r++, src += n)
                cp_async<sizeof(T)>(&tile[slot][r][lane], src);
            }
            if (!bwd) {
              cp_async<sizeof(T)>(&tabs[slot][0][lane], abr + c0 + lane);
            } else {
              cp_async<sizeof(T)>(&tabs[slot][0][lane], am + c0 + lane + 1);
              cp_async<sizeof(T)>(&tabs[slot][1][lane], bm + c0 + lane + 1);
              cp_async<8>(&taby[slot][lane], rbm + c0 + lane + 1);
            }
          }
        }
        cp_commit();
      };
      for (uint32_t ch = 0; ch < WS_S; ch++)
        issue(ch);
      cp_wait<WS_S - 1>();
      bar_arrive(WS_FULL + 0);
      for (uint32_t ch = 0; ch < nch; ch++) {
        if (ch + 1 < nch) {
          cp_wait<WS_S - 2>();
          bar_arrive(WS_FULL + (ch + 1) % WS_S);
        }
        bar_sync(WS_FREE + ch % WS_S);
        issue(ch + WS_S);
      }
      cp_wait<0>();
    } else if (role == 1) { // ---- drainer ----
      for (uint32_t ch = 0; ch < nch; ch++) {
        const uint32_t slot = ch % WS_S;
        const uint32_t c0 = (bwd ? nch - 1 - ch : ch) * 32;
        bar_sync(WS_DONE + slot);
        if (c0 + lane < n) {
          if (ab && bwd) {
            T *d = ab + c0 + lane;
            for (uint32_t r = 0; r < nl; r++, d += n)
              *d = sign > 0 ? add_rn(*d, tile[slot][r][lane])
                            : sub_rn(*d, tile[slot][r][lane]);
          } else {
            T *d = xb + c0 + lane;
            if (nl == 32) {
#pragma unroll 8
              for (uint32_t r = 0; r < 32; r++, d += n)
                *d = tile[slot][r][lane];
            } else {
              for (uint32_t r = 0; r < nl; r++, d += n)
                *d = tile[slot][r][lane];
            }
          }
        }
        if (!bwd)
          __threadfence_block();
        bar_arrive(WS_FREE + slot);
      }
    } else { // ---- consumer ----
      T prev = 0;
      for (uint32_t ch = 0; ch < nch; ch++) {
        const uint32_t slot = ch % WS_S;
        const uint32_t c0 = (bwd ? nch - 1 - ch : ch) * 32;
        const uint32_t ncol = min(32u, n - c0);
        bar_sync(WS_FULL + slot);
        T(*t)[33] = tile[slot];
        if (lane < nl) {
          if (!bwd) {
            const T *fa = tabs[slot][0];
            if (ncol == 32) {
#pragma unroll 8
              for (uint32_t j = 0; j < 32; j++) {
                prev = sub_rn(t[lane][j], mul_rn(prev, fa[j]));
                t[lane][j] = prev;
              }
            } else {
              for (uint32_t j = 0; j < ncol; j++) {
                prev = sub_rn(t[lane][j], mul_rn(prev, fa[j]));
                t[lane][j] = prev;
              }
            }
          } else {
            for (uint32_t j0 = 0; j0 < ncol; j0 += 8) {
              T c[8], a1[8], b1[8], res[8];
              double y1[8];
              const int cnt = (int)min(8u, ncol - j0);
#pragma unroll
              for (int k = 0; k < 8; k++) {
                uint32_t j = k < cnt ? ncol - 1 - j0 - k : 0;
                c[k] = t[lane][j];
                a1[k] = tabs[slot][0][j];
                b1[k] = tabs[slot][1][j];
                y1[k] = taby[slot][j];
              }
              prev = backward_batch<T, 8>(c, a1, b1, y1, cnt, prev, fast != 0, res);
#pragma unroll
              for (int k = 0; k < 8; k++)
                if (k < cnt)
                  t[lane][ncol - 1 - j0 - k] = res[k];
            }
          }
        }
        __syncwarp();
        bar_arrive(WS_DONE + slot);
      }
    }
  }
}

// ---- inverse level kernel ------------------------------------------------------------
// coefficients (v, reordered) + corrected coarse box A -> interleaved level-l
// nodes (GpkRev3D, GridProcessingKernel3D.hpp:1230-2412)
template <typename T, int K, bool SYM>
__device__ __forceinline__ void restore_row(const Coef<T, SYM> &cf,
                                            const T *__restrict__ A,
                                            T *__restrict__ dst, const Row<T> &r,
                                            const LvGeom &g, uint32_t i,
                                            bool has_odd, uint32_t qe, T tf) {
  const uint32_t ncf = g.nc[FD];
  T *drow = dst + r.so;
  if constexpr (K == 0) {
    const T *arow = A + r.ao;
    T a0 = arow[i];
    drow[qe] = a0;
    if (has_odd)
      drow[2 * i + 1] = add_rn(cf.at(r.vo + ncf + i), mgx_lerp(a0, arow[i + 1], tf));
  } else {
    T a, b;
    corners<T, K, false>(A, (int64_t)r.ao, r, i, i + 1, has_odd, tf, a, b);
    drow[qe] = add_rn(cf.at(r.vo + i), a);
    if (has_odd)
      drow[2 * i + 1] = add_rn(cf.at(r.vo + ncf + i), b);
  }
}

template <typename T, bool SYM, bool SKIP>
__global__ void __launch_bounds__(256)
    k_level_restore_rows(const Coef<T, SYM> cf, const T *__restrict__ A,
                         T *__restrict__ dst, const T *__restrict__ tables,
                         const __grid_constant__ LvGeom g) {
  const uint32_t ncf = g.nc[FD], nff = g.nf[FD], ncoef = nff - ncf;
  const int lane = threadIdx.x & 31;
  const uint64_t task = (uint64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (task >= g.tasks)
    return;
  const uint32_t cw = (uint32_t)(task % g.col_chunks);
  const uint64_t rg = task / g.col_chunks;
  const uint64_t ii = (uint64_t)cw * 32 + lane;
  if (ii >= ncf)
    return;
  const uint32_t i = (uint32_t)ii;
  const bool has_odd = i < ncoef;
  const uint32_t qe = (i == ncf - 1) ? nff - 1 : 2 * i;
  const T tf = (g.half || !has_odd) ? (T)0.5 : tables[g.roff[FD] + 2 * i];
  const uint64_t row0 = rg * g.rows_per_task;
  const uint64_t row1 = min(g.rows, row0 + g.rows_per_task);
  uint32_t p[OUTER];
  row_first(p, g, row0);
  Row<T> r;
  row_setup<T, false, SKIP>(r, p, g, tables);
  for (uint64_t row = row0; row < row1; row++) {
    const bool more = row + 1 < row1;
    if (more) {
      row_next(p, g);
      Row<T> nx;
      row_setup<T, false, SKIP>(nx, p, g, tables);
      cf.prefetch(nx.vo + i);
      if (has_odd)
        cf.prefetch(nx.vo + ncf + i);
      prefetch_l1(A + nx.ao + i);
      prefetch_row<T, false>(A, (int64_t)nx.ao + i, nx);
    }
    switch (r.nodd) {
    case 0:
      restore_row<T, 0, SYM>(cf, A, dst, r, g, i, has_odd, qe, tf);
      break;
    case 1:
      restore_row<T, 1, SYM>(cf, A, dst, r, g, i, has_odd, qe, tf);
      break;
    case 2:
      restore_row<T, 2, SYM>(cf, A, dst, r, g, i, has_odd, qe, tf);
      break;
    case 3:
      restore_row<T, 3, SYM>(cf, A, dst, r, g, i, has_odd, qe, tf);
      break;
    default:
      restore_row<T, 4, SYM>(cf, A, dst, r, g, i, has_odd, qe, tf);
      break;
    }
    if (more)
      row_setup<T, false, SKIP>(r, p, g, tables);
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
      // v == nullptr: dequantise the level-0 symbols (qo.qv = quantizer * volume)
      A[it] = v ? v[o] : mul_rn(qo.qv, (T)((int)qo.sym[o] - (int)qo.shift));
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
// `cols_per_chunk`: columns (pairs) one warp covers per row
template <typename T>
LvGeom level_geom(const mgx_ctx *ctx, int l, uint32_t cols_per_chunk) {
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
  // warp tasks: enough of them to fill the machine a few times over, each a
  // contiguous run of rows (neighbour rows are then shared through L1)
  g.col_chunks = (g.nc[FD] + cols_per_chunk - 1) / cols_per_chunk;
  const uint64_t want_tasks = (uint64_t)ctx->num_sms * 64 * 3;
  uint64_t groups = std::max<uint64_t>(1, want_tasks / g.col_chunks);
  uint64_t rpt = (g.rows + groups - 1) / groups;
  rpt = std::max<uint64_t>(rpt, std::min<uint64_t>(g.rows, 4));
  g.rows_per_task = (uint32_t)std::min<uint64_t>(rpt, 1u << 30);
  g.tasks = (uint64_t)g.col_chunks * ((g.rows + g.rows_per_task - 1) / g.rows_per_task);
  return g;
}

// ---- partitioned Thomas solves (long, scarce chains) -----------------------------
// A 2-D 8193^2 level has 4097 chains of 4097 steps: one thread per chain leaves the
// GPU at ~10 % of its bandwidth.  Both recurrences contract (|am / bm| ~ 0.27 per
// step), so a chain is cut into P partitions that run concurrently: partition p
// first runs the recurrence over the PT_W elements BEFORE its range from a zero
// state, storing nothing; after PT_W steps the influence of the wrong start is
// 0.27^64 ~ 1e-37 of the neighbouring values -- far below an ulp -- and the state
// it enters its own range with is, bit for bit, the state the sequential sweep has
// there.  That is not assumed but CHECKED: the state partition p entered with is
// compared bitwise with the last state partition p-1 computed; where they differ
// (a decaying tail after a run of zeros, a 1e+250 next to a 1e-250) partition p
// repeats its range from its neighbour's state, and the comparison is repeated
// until every pair agrees (at most P rounds; the first partition is exact, so
// round r fixes partition r at the latest).  The result then IS the sequential
// sweep.  Hence: forward sweep out of place (x -> y, x stays readable for the
// repeats), backward sweep y -> x, and for the accumulating solve a third in-kernel
// pass acc +-= x over the CTA's own chains once they are final.
// A CTA owns 32 chains (lanes) x P partitions (warps).
constexpr int PT_W = 64;       // warm-up steps
constexpr int PT_MAXP = 16;    // partitions (warps) per CTA
#ifndef MGX_PT_MIN_STRIDED
#define MGX_PT_MIN_STRIDED 1024
#endif
constexpr uint32_t PT_MIN_CONTIG = 512;  // shortest chains worth partitioning (measured)
constexpr uint32_t PT_MIN_STRIDED = MGX_PT_MIN_STRIDED;

template <typename T> __device__ __forceinline__ bool bits_differ(T a, T b);
template <> __device__ __forceinline__ bool bits_differ<float>(float a, float b) {
  return __float_as_uint(a) != __float_as_uint(b);
}
template <> __device__ __forceinline__ bool bits_differ<double>(double a, double b) {
  return __double_as_longlong(a) != __double_as_longlong(b);
}

// Repeat `sweep(enter) -> last state` for the partitions whose entering state is not
// the neighbour's last state, until none is left.  nb = index of the neighbour the
// state comes from (p-1 forward, p+1 backward), has_nb = there is one.
template <typename T, class Sweep>
__device__ __forceinline__ void
part_repair(T (&s_edge)[PT_MAXP][32], uint32_t p, uint32_t lane, bool has_nb,
            uint32_t nb, bool live, T enter, T last, Sweep sweep,
            unsigned long long *redo_count) {
  for (;;) {
    s_edge[p][lane] = last;
    __syncthreads();
    const bool need = has_nb && live && bits_differ(enter, s_edge[nb][lane]);
    const T from = has_nb ? s_edge[nb][lane] : enter;
    const int any = __syncthreads_or(need ? 1 : 0);
    if (!any)
      break;
    if (__any_sync(0xffffffffu, need)) {
      // lanes that agree recompute the same bits from the same state
      enter = from;
      last = sweep(enter);
      if (need)
        atomicAdd(redo_count, 1ull);
    }
  }
}

// chain element i of lane's chain lives at base[i * inner]
template <typename T, bool ACC>
__global__ void __launch_bounds__(PT_MAXP * 32)
    k_ipk_part_strided(T *__restrict__ x, T *__restrict__ y, const T *__restrict__ am,
                       const T *__restrict__ bm, const T *__restrict__ abr,
                       const double *__restrict__ rbm, uint32_t n, uint64_t inner,
                       uint64_t lines, uint32_t lp, T *__restrict__ acc, int sign,
                       int fast, unsigned long long *__restrict__ redo_count) {
  __shared__ T s_edge[PT_MAXP][32];
  const uint32_t lane = threadIdx.x & 31, p = threadIdx.x >> 5, P = blockDim.x >> 5;
  uint64_t ln = (uint64_t)blockIdx.x * 32 + lane;
  const bool active = ln < lines;
  if (!active)
    ln = lines - 1; // shadow lanes: compute, never store
  const uint64_t o = ln / inner, in_ = ln % inner;
  const uint64_t base = o * n * inner + in_;
  T *xl = x + base, *yl = y + base;
  T *al = ACC ? acc + base : nullptr;
  const uint32_t s = min(p * lp, n), e = min(s + lp, n);
  // ---- forward: y[i] = x[i] - y[i-1] * abr[i] ----
  // (the rows of the NEXT batch of 8 are requested before the current batch's
  // dependent chain runs: one memory latency per sweep instead of one per batch)
  auto fwd = [&](uint32_t lo, uint32_t hi, T prev, bool store) {
    if (lo >= hi)
      return prev;
    T c[8], cn[8];
#pragma unroll
    for (int k = 0; k < 8; k++)
      c[k] = xl[(uint64_t)min(lo + k, hi - 1) * inner];
    for (uint32_t i0 = lo; i0 < hi; i0 += 8) {
      T f[8];
#pragma unroll
      for (int k = 0; k < 8; k++) {
        cn[k] = xl[(uint64_t)min(i0 + 8 + k, hi - 1) * inner];
        f[k] = __ldg(abr + min(i0 + k, hi - 1));
      }
#pragma unroll
      for (int k = 0; k < 8; k++)
        if (i0 + k < hi) {
          prev = sub_rn(c[k], mul_rn(prev, f[k]));
          if (store && active)
            yl[(uint64_t)(i0 + k) * inner] = prev;
        }
#pragma unroll
      for (int k = 0; k < 8; k++)
        c[k] = cn[k];
    }
    return prev;
  };
  {
    const T enter = fwd(s > (uint32_t)PT_W ? s - PT_W : 0, s, (T)0, false);
    const T last = fwd(s, e, enter, true);
    part_repair<T>(s_edge, p, lane, p > 0, p - 1, s < e, enter, last,
                   [&](T from) { return fwd(s, e, from, true); }, redo_count);
  }
  __syncthreads(); // every y of the CTA's chains is final and visible
  // ---- backward: x[i] = (y[i] - am[i+1] * x[i+1]) / bm[i+1] ----
  auto bwd = [&](uint32_t lo, uint32_t hi, T prev, bool store) {
    // elements hi-1, hi-2, ..., lo; the next batch's rows are requested first
    if (hi <= lo)
      return prev;
    T c[8], cn[8];
#pragma unroll
    for (int k = 0; k < 8; k++)
      c[k] = yl[(uint64_t)(hi - 1 - min((uint32_t)k, hi - lo - 1)) * inner];
    while (hi > lo) {
      const uint32_t cnt = min(8u, hi - lo);
      const uint32_t nhi = hi - cnt; // next batch: nhi-1, nhi-2, ...
      T a1[8], b1[8], res[8];
      double y1[8];
#pragma unroll
      for (int k = 0; k < 8; k++) {
        const uint32_t left = nhi > lo ? nhi - lo : 1u;
        const uint32_t in = nhi > lo ? nhi - 1 - min((uint32_t)k, left - 1) : lo;
        cn[k] = yl[(uint64_t)in * inner];
        const uint32_t i = hi - 1 - min((uint32_t)k, cnt - 1);
        a1[k] = __ldg(am + i + 1);
        b1[k] = __ldg(bm + i + 1);
        y1[k] = __ldg(rbm + i + 1);
      }
      prev = backward_batch<T, 8>(c, a1, b1, y1, (int)cnt, prev, fast != 0, res);
      if (store && active) {
#pragma unroll
        for (int k = 0; k < 8; k++)
          if ((uint32_t)k < cnt)
            xl[(uint64_t)(hi - 1 - k) * inner] = res[k];
      }
#pragma unroll
      for (int k = 0; k < 8; k++)
        c[k] = cn[k];
      hi = nhi;
    }
    return prev;
  };
  {
    const T enter = bwd(e, min(e + (uint32_t)PT_W, n), (T)0, false);
    const T last = bwd(s, e, enter, true);
    part_repair<T>(s_edge, p, lane, p + 1 < P && e < n, p + 1, s < e, enter, last,
                   [&](T from) { return bwd(s, e, from, true); }, redo_count);
  }
  if (ACC) {
    __syncthreads();
    if (active)
      for (uint32_t i0 = s; i0 < e; i0 += 8) { // 16 independent loads in flight
        T av[8], xv[8];
#pragma unroll
        for (int k = 0; k < 8; k++) {
          const uint64_t off = (uint64_t)min(i0 + k, e - 1) * inner;
          av[k] = al[off];
          xv[k] = xl[off];
        }
#pragma unroll
        for (int k = 0; k < 8; k++)
          if (i0 + k < e)
            al[(uint64_t)(i0 + k) * inner] =
                sign > 0 ? add_rn(av[k], xv[k]) : sub_rn(av[k], xv[k]);
      }
  }
}

// Contiguous chains: lanes = 32 chains, a warp moves 32 x 32 tiles of its partition
// through shared memory (coalesced rows in, a lane walks its chain's row, coalesced
// rows out), as k_ipk_contig does for whole chains.  lp and PT_W are multiples of 32.
template <typename T> struct PartTile {
  T t[32][33];
  T ta[32], tb[32];
  double ty[32];
};

template <typename T, bool ACC>
__global__ void __launch_bounds__(PT_MAXP * 32)
    k_ipk_part_contig(T *__restrict__ x, T *__restrict__ y, const T *__restrict__ am,
                      const T *__restrict__ bm, const T *__restrict__ abr,
                      const double *__restrict__ rbm, uint32_t n, uint64_t lines,
                      uint32_t lp, T *__restrict__ acc, int sign, int fast,
                      unsigned long long *__restrict__ redo_count) {
  extern __shared__ __align__(16) unsigned char pt_smem[];
  __shared__ T s_edge[PT_MAXP][32];
  const uint32_t lane = threadIdx.x & 31, p = threadIdx.x >> 5, P = blockDim.x >> 5;
  PartTile<T> &w = reinterpret_cast<PartTile<T> *>(pt_smem)[p];
  const uint64_t l0 = (uint64_t)blockIdx.x * 32;
  const uint32_t nl = (uint32_t)min((uint64_t)32, lines - l0);
  T *xb = x + l0 * n, *yb = y + l0 * n;
  T *ab = ACC ? acc + l0 * n : nullptr;
  const uint32_t s = min(p * lp, n), e = min(s + lp, n);
  // ---- forward over the 32-column tiles [lo, hi) ----
  auto fwd = [&](uint32_t lo, uint32_t hi, T prev, bool store) {
    for (uint32_t c0 = lo; c0 < hi; c0 += 32) {
      const uint32_t ncol = min(32u, n - c0);
      if (lane < ncol) {
        for (uint32_t r = 0; r < nl; r++)
          w.t[r][lane] = xb[(uint64_t)r * n + c0 + lane];
        w.ta[lane] = __ldg(abr + c0 + lane);
      }
      __syncwarp();
      if (lane < nl) {
        if (ncol == 32) {
#pragma unroll 8
          for (uint32_t j = 0; j < 32; j++) {
            prev = sub_rn(w.t[lane][j], mul_rn(prev, w.ta[j]));
            w.t[lane][j] = prev;
          }
        } else {
          for (uint32_t j = 0; j < ncol; j++) {
            prev = sub_rn(w.t[lane][j], mul_rn(prev, w.ta[j]));
            w.t[lane][j] = prev;
          }
        }
      }
      __syncwarp();
      if (store && lane < ncol)
        for (uint32_t r = 0; r < nl; r++)
          yb[(uint64_t)r * n + c0 + lane] = w.t[r][lane];
      __syncwarp();
    }
    return prev;
  };
  {
    const T enter = fwd(s > (uint32_t)PT_W ? s - PT_W : 0, s, (T)0, false);
    const T last = fwd(s, e, enter, true);
    part_repair<T>(s_edge, p, lane, p > 0, p - 1, s < e && lane < nl, enter, last,
                   [&](T from) { return fwd(s, e, from, true); }, redo_count);
  }
  __syncthreads();
  // ---- backward over the tiles of [lo, hi), highest first (lo a multiple of 32) ----
  auto bwd = [&](uint32_t lo, uint32_t hi, T prev, bool store) {
    if (hi <= lo)
      return prev;
    for (uint32_t c0 = ((hi - 1) / 32) * 32 + 32; c0 > lo;) {
      c0 -= 32;
      const uint32_t ncol = min(32u, hi - c0);
      if (lane < ncol) {
        for (uint32_t r = 0; r < nl; r++)
          w.t[r][lane] = yb[(uint64_t)r * n + c0 + lane];
        w.ta[lane] = __ldg(am + c0 + lane + 1);
        w.tb[lane] = __ldg(bm + c0 + lane + 1);
        w.ty[lane] = __ldg(rbm + c0 + lane + 1);
      }
      __syncwarp();
      if (lane < nl) {
        for (uint32_t j0 = 0; j0 < ncol; j0 += 8) {
          T c[8], a1[8], b1[8], res[8];
          double y1[8];
          const int cnt = (int)min(8u, ncol - j0);
#pragma unroll
          for (int k = 0; k < 8; k++) {
            const uint32_t j = k < cnt ? ncol - 1 - j0 - k : 0;
            c[k] = w.t[lane][j];
            a1[k] = w.ta[j];
            b1[k] = w.tb[j];
            y1[k] = w.ty[j];
          }
          prev = backward_batch<T, 8>(c, a1, b1, y1, cnt, prev, fast != 0, res);
#pragma unroll
          for (int k = 0; k < 8; k++)
            if (k < cnt)
              w.t[lane][ncol - 1 - j0 - k] = res[k];
        }
      }
      __syncwarp();
      if (store && lane < ncol)
        for (uint32_t r = 0; r < nl; r++)
          xb[(uint64_t)r * n + c0 + lane] = w.t[r][lane];
      __syncwarp();
    }
    return prev;
  };
  {
    const T enter = bwd(e, min(e + (uint32_t)PT_W, n), (T)0, false);
    const T last = bwd(s, e, enter, true);
    part_repair<T>(s_edge, p, lane, p + 1 < P && e < n, p + 1, s < e && lane < nl,
                   enter, last, [&](T from) { return bwd(s, e, from, true); },
                   redo_count);
  }
  if (ACC) {
    __syncthreads();
    for (uint32_t c0 = s; c0 < e; c0 += 32)
      if (c0 + lane < n)
        for (uint32_t r = 0; r < nl; r++) {
          const uint64_t off = (uint64_t)r * n + c0 + lane;
          ab[off] = sign > 0 ? add_rn(ab[off], xb[off]) : sub_rn(ab[off], xb[off]);
        }
  }
}

unsigned task_grid(const LvGeom &g, unsigned warps_per_block) {
  return (unsigned)((g.tasks + warps_per_block - 1) / warps_per_block);
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
    if (outer >= 4) {
      unsigned gz = (unsigned)std::min<uint64_t>(
          std::min<uint64_t>((outer + 3) / 4, want_z), 65535);
      k_lpk_mid<T, 4><<<dim3(nc, gy, gz), 256, 0, ctx->stream>>>(
          in, out, tables + off.dist[l][d], (uint32_t)outer, nf, nc, inner);
    } else {
      unsigned gz = (unsigned)std::min<uint64_t>(std::min<uint64_t>(outer, want_z), 65535);
      k_lpk_mid<T, 1><<<dim3(nc, gy, gz), 256, 0, ctx->stream>>>(
          in, out, tables + off.dist[l][d], (uint32_t)outer, nf, nc, inner);
    }
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
    const double *rbm = ctx->d_rcp + off.bm[l - 1][d];
    const uint64_t lines = outer * inner;
    const int fast = ctx->fast_div ? 1 : 0;
    T *acc = (d == pad) ? A : nullptr;
    // scarce chains: the producer/consumer pair kernels (one pair per 32 lines)
    const bool scarce = lines < (uint64_t)ctx->num_sms * 128 && !ctx->no_ws;
    // long, scarce chains: partitioned solves (bit-identical, checked in-kernel)
    if (!ctx->no_part && n >= (inner == 1 ? PT_MIN_CONTIG : PT_MIN_STRIDED) &&
        lines < (uint64_t)ctx->num_sms * 512) {
      const uint32_t P = (uint32_t)std::min<uint64_t>(PT_MAXP, n / 128);
      const uint32_t lp = ((n + P - 1) / P + 31) / 32 * 32;
      T *other = corr == R0 ? R1 : R0;
      unsigned long long *redo =
          (unsigned long long *)(ctx->d_scalars + MGX_SC_REDO);
      const unsigned grid = (unsigned)((lines + 31) / 32);
      if (inner == 1) {
        const size_t smem = (size_t)P * sizeof(PartTile<T>);
        auto kern = acc ? k_ipk_part_contig<T, true> : k_ipk_part_contig<T, false>;
        MGX_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      (int)(PT_MAXP * sizeof(PartTile<T>))));
        kern<<<grid, P * 32, smem, ctx->stream>>>(corr, other, am, bm, abr, rbm, n,
                                                  lines, lp, acc, sign, fast, redo);
      } else {
        auto kern = acc ? k_ipk_part_strided<T, true> : k_ipk_part_strided<T, false>;
        kern<<<grid, P * 32, 0, ctx->stream>>>(corr, other, am, bm, abr, rbm, n, inner,
                                               lines, lp, acc, sign, fast, redo);
      }
    } else if (inner == 1 && scarce) {
      k_ipk_contig_ws<T><<<(unsigned)((lines + 31) / 32), 96, 0, ctx->stream>>>(
          corr, am, bm, abr, rbm, n, lines, acc, sign, fast);
    } else if (inner == 1) {
      k_ipk_contig<T><<<(unsigned)((lines + 31) / 32), 32, 0, ctx->stream>>>(
          corr, am, bm, abr, rbm, n, lines, acc, sign, fast);
    } else if (scarce) {
      const unsigned grid = (unsigned)((lines + 31) / 32);
      if (acc)
        k_ipk_strided_ws<T, true><<<grid, 96, 0, ctx->stream>>>(
            corr, am, bm, abr, rbm, n, inner, lines, acc, sign, fast);
      else
        k_ipk_strided_ws<T, false><<<grid, 96, 0, ctx->stream>>>(
            corr, am, bm, abr, rbm, n, inner, lines, nullptr, sign, fast);
    } else {
      // few lines: small blocks, so that every SM gets a chain to work on
      unsigned tpb = 128;
      while (tpb > 32 && lines / tpb < (uint64_t)ctx->num_sms * 2)
        tpb >>= 1;
      const unsigned grid = (unsigned)((lines + tpb - 1) / tpb);
      // deep ring when chains are scarce (latency bound), shallow otherwise
      const bool deep = lines < (uint64_t)ctx->num_sms * 512;
      const size_t stages = deep ? 8 : 2;
      const size_t smem = stages * IPK_B * tpb * sizeof(T) * (acc ? 2 : 1) +
                          (size_t)(tpb / 32) * stages * IPK_B * (8 + 2 * sizeof(T));
#define MGX_IPK_LAUNCH(ACCF, S)                                                  \
  do {                                                                           \
    /* per device: set on every launch (a process may drive several GPUs) */     \
    MGX_CUDA(cudaFuncSetAttribute(k_ipk_strided<T, ACCF, S>,                     \
                                  cudaFuncAttributeMaxDynamicSharedMemorySize,   \
                                  8 * IPK_B * 128 * 8 * 2 + 4 * 8 * IPK_B * 24)); \
    k_ipk_strided<T, ACCF, S><<<grid, tpb, smem, ctx->stream>>>(                 \
        corr, am, bm, abr, rbm, n, inner, lines, acc, sign, fast);               \
  } while (0)
      if (acc && deep)
        MGX_IPK_LAUNCH(true, 8);
      else if (acc)
        MGX_IPK_LAUNCH(true, 2);
      else if (deep)
        MGX_IPK_LAUNCH(false, 8);
      else
        MGX_IPK_LAUNCH(false, 2);
#undef MGX_IPK_LAUNCH
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
  const bool skip = lv.D < MGX_MAX_D; // padded dimensions exist
  const bool three = ctx->fw_minb == 3 || (ctx->fw_minb == 0 && sizeof(T) == 4);
  void (*fwd)(const T *, T *, T *, T *, const T *, const LvGeom, const QuantOut<T>) =
      three ? (skip ? k_level_forward<T, QUANT, 3, true> : k_level_forward<T, QUANT, 3, false>)
            : (skip ? k_level_forward<T, QUANT, 2, true> : k_level_forward<T, QUANT, 2, false>);
  if (QUANT)
    MGX_CUDA(cudaFuncSetAttribute(fwd, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  (int)smem));
  QuantOut<T> qo;
  const T *cur = src;
  int which = 0;
  for (int l = lv.L; l >= 1; l--) {
    LvGeom g = level_geom<T>(ctx, l, FW_OWN);
    fill_quant<T>(ctx, &qo, q, vol, l);
    T *A = (T *)s.A[which];
    // (3 CTAs/SM pays for fp32: 80 regs without spills; fp64 spills at 80)
    fwd<<<task_grid(g, FW_WARPS), FW_WARPS * 32, smem, ctx->stream>>>(
        cur, v, A, (T *)s.R[0], tables, g, qo);
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
// recompose: coefficients -> dst (full array), dst != v.  Levels below
// `sym_from` read the float coefficients in v (reordered, read only); levels
// from `sym_from` on read the symbols in ctx->d_sym and dequantise them on the
// fly with dq[l] (sym_from = L + 1: floats only; 0: symbols only).
template <typename T>
int recompose_rows(mgx_ctx *ctx, const T *v, T *dst, const double *dq,
                   int sym_from) {
  Scratch s;
  int rc = get_scratch(ctx, &s);
  if (rc)
    return rc;
  const MgxLevels &lv = ctx->lv;
  const T *tables = (const T *)ctx->d_tables;
  QuantOut<T> qo;
  fill_quant<T>(ctx, &qo, nullptr, nullptr, 0);
  Coef<T, true> cs;
  cs.v = nullptr;
  cs.sym = ctx->d_sym;
  cs.shift = (int)(ctx->dict / 2);
  cs.dq = (T)0;
  Coef<T, false> cf;
  cf.v = v;
  cf.sym = nullptr;
  cf.shift = 0;
  cf.dq = (T)0;
  LvGeom g0;
  for (int d = 0; d < MGX_MAX_D; d++) {
    g0.nc[d] = (uint32_t)lv.n[0][d];
    g0.vst[d] = lv.stride[d];
  }
  uint64_t n0 = 1;
  for (int d = 0; d < MGX_MAX_D; d++)
    n0 *= lv.n[0][d];
  const unsigned cb_grid = mgx_grid_for(n0, 256, ctx->num_sms * 4);
  const bool sym0 = sym_from <= 0;
  if (sym0)
    qo.qv = (T)dq[0];
  if (lv.L == 0) {
    if (sym0) {
      k_coarse_box<T, false, true><<<cb_grid, 256, 0, ctx->stream>>>(dst, nullptr, g0,
                                                                    qo);
      ctx->launches++;
    } else {
      MGX_CUDA(cudaMemcpyAsync(dst, v, lv.N * sizeof(T), cudaMemcpyDeviceToDevice,
                               ctx->stream));
    }
    return MGX_SUCCESS;
  }
  const bool skip = lv.D < MGX_MAX_D;
  int which = 0;
  T *A = (T *)s.A[which];
  k_coarse_box<T, false, true><<<cb_grid, 256, 0, ctx->stream>>>(
      A, sym0 ? nullptr : const_cast<T *>(v), g0, qo);
  ctx->launches++;
  for (int l = 1; l <= lv.L; l++) {
    LvGeom g = level_geom<T>(ctx, l, 32);
    const bool use_sym = l >= sym_from;
    if (use_sym) {
      cs.dq = (T)dq[l];
      k_lpk_first<T, true><<<task_grid(g, 8), 256, 0, ctx->stream>>>(cs, (T *)s.R[0],
                                                                    tables, g);
    } else {
      k_lpk_first<T, false><<<task_grid(g, 8), 256, 0, ctx->stream>>>(
          cf, (T *)s.R[0], tables, g);
    }
    ctx->launches++;
    rc = correction_tail<T>(ctx, l, (T *)s.R[0], (T *)s.R[1], A, -1);
    if (rc)
      return rc;
    T *out = (l == lv.L) ? dst : (T *)s.A[which ^ 1];
    if (use_sym)
      (skip ? k_level_restore_rows<T, true, true>
            : k_level_restore_rows<T, true, false>)<<<task_grid(g, 8), 256, 0,
                                                      ctx->stream>>>(cs, A, out,
                                                                     tables, g);
    else
      (skip ? k_level_restore_rows<T, false, true>
            : k_level_restore_rows<T, false, false>)<<<task_grid(g, 8), 256, 0,
                                                       ctx->stream>>>(cf, A, out,
                                                                      tables, g);
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
  const int never = ctx->lv.L + 1;
  return mgx_is_f32(ctx)
             ? recompose_rows<float>(ctx, (const float *)d_coeff, (float *)d_dst,
                                     nullptr, never)
             : recompose_rows<double>(ctx, (const double *)d_coeff, (double *)d_dst,
                                      nullptr, never);
}

// dequantise + recompose from the symbols in ctx->d_sym; levels below
// `sym_from` (the ones holding outliers) are read from the float coefficients
// the caller dequantised into d_coeff
int mgxk_recompose_sym_rows(mgx_ctx *ctx, int rel, double tol, double s,
                            double norm, const void *d_coeff, int sym_from,
                            void *d_dst) {
  double q[MGX_MAX_L], vol[MGX_MAX_L], dq[MGX_MAX_L];
  mgx_level_quantizers(ctx, rel, tol, s, norm, false, q, vol);
  for (int l = 0; l <= ctx->lv.L; l++) {
    // (quantizer * volume) rounded in T, as k_dequantize_sym forms it
    if (mgx_is_f32(ctx))
      dq[l] = (double)((float)q[l] * (float)vol[l]);
    else
      dq[l] = q[l] * vol[l];
  }
  return mgx_is_f32(ctx)
             ? recompose_rows<float>(ctx, (const float *)d_coeff, (float *)d_dst, dq,
                                     sym_from)
             : recompose_rows<double>(ctx, (const double *)d_coeff, (double *)d_dst,
                                      dq, sym_from);
}
