// Pieces shared by the fused hierarchical-basis kernels (mgx_hier.cu: LDGSTS-staged
// tiles, 1-3 D; mgx_hier3.cu: TMA-staged, warp-specialised 3-D forward pass).
// Included inside each translation unit; everything lives in an anonymous namespace.
#pragma once
#include <algorithm>
#include <cmath>
#include <limits>
#include <type_traits>

#include "mgx_device.cuh"
#include "mgx_internal.h"

namespace {

constexpr int HIER_THREADS = 256;
constexpr int HIER_WARPS = HIER_THREADS / 32;

template <typename T> struct HierArgs {
  uint32_t n[3];                // nz, ny, nx (padded with 1)
  uint32_t nl[MGX_MAX_L][3];    // level sizes per axis
  uint32_t roff[MGX_MAX_L][3];  // ratio table offsets per level and axis
  int L;
  T qv;                         // quantiser reciprocal (all levels equal, L-inf)
  int shift;                    // dict / 2
  uint32_t dict;
  uint32_t tiles[3];
  uint64_t ntiles;
  uint64_t oct[12][3];            // pass p: byte offsets of the odd x / y / z blocks
  T dq;                          // dequantisation step (inverse)
  const T *coarse;               // inverse: nodal values of the level-(L-PASSES) grid
  uint32_t cdim[3];              // its shape
  unsigned long long n_out;      // inverse: number of outliers (sorted by index)
  uint16_t *sym;                 // symbols (reordered layout)
  unsigned long long *out_count; // outlier list
  unsigned long long *out_idx;
  long long *out_val;
  unsigned long long out_cap;
};

// Tile geometry: 2^LZ x 2^LY x 2^LX nodes; an axis with extent 1 is "flat"
// (no halo, never interpolated).  PASSES = strides 1, 2, .. handled in-tile.
template <int LZ_, int LY_, int LX_> struct TileShape {
  static constexpr int LZ = LZ_, LY = LY_, LX = LX_;
  static constexpr int TZ = 1 << LZ, TY = 1 << LY, TX = 1 << LX;
  static constexpr bool RZ = LZ > 0, RY = LY > 0; // real (non-flat) axes
  static constexpr int HZ = RZ ? TZ + 1 : 1, HY = RY ? TY + 1 : 1;
  static constexpr int PX = TX + 2; // even pitch: (even, odd) pairs stay aligned
  static constexpr int ROWS = HZ * HY;
  static constexpr int PASSES =
      RZ ? (LZ < LY ? (LZ < LX ? LZ : LX) : (LY < LX ? LY : LX))
         : (RY ? (LY < LX ? LY : LX) : LX);
  static constexpr size_t tile_elems = (size_t)ROWS * PX;
};

struct AxisInfo {
  int lvl;       // level at which this index is a coefficient (0 = nodal)
  uint32_t lo, hi; // interpolation neighbours (valid when lvl > 0)
  uint32_t pos;  // position in the reordered layout
  uint32_t j;    // index in the level-lvl grid
};

// Closed form of the reference's recursive "even indices + last node" coarsening
// (SURVEY.md section 9.1): the level-l grid is {multiples of 2^(L-l)} U {n-1}.
__device__ __forceinline__ AxisInfo axis_info(uint32_t i, uint32_t n, int L,
                                              const uint32_t *nl, int ax) {
  AxisInfo a;
  a.lvl = 0;
  a.lo = a.hi = i;
  a.j = 0;
  if (n == 1) {
    a.pos = 0;
  } else if (i == n - 1) {
    a.pos = nl[0 * 3 + ax] - 1;
  } else if (i == 0) {
    a.pos = 0;
  } else {
    int tz = __ffs(i) - 1;
    int lvl = L - tz;
    if (lvl <= 0) {
      a.pos = i >> L;
    } else {
      int sh = L - lvl;
      a.lvl = lvl;
      a.j = i >> sh;
      a.pos = nl[(lvl - 1) * 3 + ax] + ((a.j - 1) >> 1);
      a.lo = i - (1u << sh);
      a.hi = min(i + (1u << sh), n - 1);
    }
  }
  return a;
}

// Reordered position: a level-l coefficient keeps, along the axes in which it
// is a level-l COARSE node, its index in the level-(l-1) grid (coarser levels
// only permute the [0, n_{l-1}) box, which this element is outside of); along
// its odd axes it sits in the coefficient block [n_{l-1}, n_l).
__device__ __forceinline__ uint32_t final_pos(const AxisInfo &a, uint32_t i,
                                              uint32_t n, int l, int L,
                                              const uint32_t *nl, int ax) {
  if (l == 0 || a.lvl == l)
    return a.pos;
  if (n == 1)
    return 0;
  if (i == n - 1)
    return nl[(l - 1) * 3 + ax] - 1;
  return i >> (L - l + 1);
}

// Rare path of the quantiser: the int conversion saturated or the symbol is
// outside the dictionary.  Redo it in 64 bits and, if it really is an outlier,
// append (position, value) to the list and return symbol 0
// (OutlierSeparator.hpp:17-44).
template <typename T>
__device__ __noinline__ uint32_t hier_slow_symbol(T c, T r, const HierArgs<T> &a,
                                                  uint64_t opos) {
  long long q = (long long)r;
  q = (c < (T)0 ? -q : q) + (long long)a.shift;
  if (q >= 0 && q < (long long)a.dict)
    return (uint32_t)q;
  unsigned long long slot = atomicAdd(a.out_count, 1ull);
  if (slot < a.out_cap) {
    a.out_idx[slot] = opos;
    a.out_val[slot] = q;
  }
  return 0;
}

// Per-thread sink: quantise (lwqzk, LinearQuantization.hpp:96-111), separate
// outliers, count the symbol, store it.  Symbols equal to dict/2 (a zero
// coefficient, by far the most frequent one on smooth data) are counted in a
// register, everything else with a shared-memory reduction.
template <typename T> struct Sink {
  const HierArgs<T> &a;
  uint32_t hist_saddr; // shared-window address of the CTA histogram

  // candidate symbol; the (int) conversion saturates, so anything that does
  // not fit an int lands outside [0, dict) after the shift
  __device__ __forceinline__ uint32_t quant(T c) const {
    const T r = add_rn((T)0.5, fabs(mul_rn(c, a.qv)));
    return (uint32_t)((int)copysign(r, c) + a.shift);
  }
  __device__ __forceinline__ uint32_t fix(uint32_t sv, T c, uint16_t *dst) const {
    if (sv >= a.dict) {
      const T r = add_rn((T)0.5, fabs(mul_rn(c, a.qv)));
      sv = hier_slow_symbol<T>(c, r, a, (uint64_t)(dst - a.sym));
    }
    return sv;
  }
  // store + count (ATOMS.POPC.INC merges lanes that hit the same bin)
  __device__ __forceinline__ void store(uint32_t sv, uint16_t *dst) const {
    *dst = (uint16_t)sv;
    asm volatile("red.shared.add.u32 [%0], 1;" ::"r"(hist_saddr + sv * 4u));
  }
  __device__ __forceinline__ void put_at(T c, uint16_t *dst) const {
    store(fix(quant(c), c, dst), dst);
  }
  __device__ __forceinline__ void put(T c, uint64_t opos) const {
    put_at(c, a.sym + opos);
  }
  // N coefficients at once: one range test for the whole group
  template <int N>
  __device__ __forceinline__ void put_group(const T (&c)[N],
                                            uint16_t *const (&dst)[N]) const {
    uint32_t sv[N], worst = 0;
#pragma unroll
    for (int i = 0; i < N; i++) {
      sv[i] = quant(c[i]);
      worst = max(worst, sv[i]);
    }
    if (__builtin_expect(worst >= a.dict, 0)) {
#pragma unroll
      for (int i = 0; i < N; i++)
        sv[i] = fix(sv[i], c[i], dst[i]);
    }
#pragma unroll
    for (int i = 0; i < N; i++)
      store(sv[i], dst[i]);
  }
};

// Generic per-element path (any index, any size): gathers the 2^k neighbours
// from the staged tile when they are inside it, else from global memory.
template <typename T, bool HALF, class TS, class SinkT>
__device__ __forceinline__ void
element_generic(const T *__restrict__ u, const T *tile, const T *__restrict__ tables,
                const HierArgs<T> &a, uint32_t gx, uint32_t gy, uint32_t gz,
                uint32_t x0, uint32_t y0, uint32_t z0, uint32_t ex, uint32_t ey,
                uint32_t ez, const SinkT &sink) {
  const uint32_t nz = a.n[0], ny = a.n[1], nx = a.n[2];
  const uint64_t sy = nx, sz = (uint64_t)nx * ny;
  const uint32_t *nl = &a.nl[0][0];
  const AxisInfo ax = axis_info(gx, nx, a.L, nl, 2);
  const AxisInfo ay = axis_info(gy, ny, a.L, nl, 1);
  const AxisInfo az = axis_info(gz, nz, a.L, nl, 0);
  const int l = max(ax.lvl, max(ay.lvl, az.lvl));
  const bool ix = l > 0 && ax.lvl == l, iy = l > 0 && ay.lvl == l,
             iz = l > 0 && az.lvl == l;
  const uint32_t xl = ix ? ax.lo : gx, xh = ix ? ax.hi : gx;
  const uint32_t yl = iy ? ay.lo : gy, yh = iy ? ay.hi : gy;
  const uint32_t zl = iz ? az.lo : gz, zh = iz ? az.hi : gz;
  T c = tile ? tile[((gz - z0) * TS::HY + (gy - y0)) * TS::PX + (gx - x0)]
             : __ldg(u + gz * sz + gy * sy + gx);
  if (l > 0) {
    T v[8];
    const bool in_tile = tile && xl >= x0 && xh < x0 + ex && yl >= y0 &&
                         yh < y0 + ey && zl >= z0 && zh < z0 + ez;
    if (in_tile) {
      const uint32_t a0 = (zl - z0) * TS::HY, a1 = (zh - z0) * TS::HY;
      const uint32_t b0 = yl - y0, b1 = yh - y0, c0 = xl - x0, c1 = xh - x0;
      v[0] = tile[(a0 + b0) * TS::PX + c0];
      v[1] = tile[(a0 + b0) * TS::PX + c1];
      v[2] = tile[(a0 + b1) * TS::PX + c0];
      v[3] = tile[(a0 + b1) * TS::PX + c1];
      v[4] = tile[(a1 + b0) * TS::PX + c0];
      v[5] = tile[(a1 + b0) * TS::PX + c1];
      v[6] = tile[(a1 + b1) * TS::PX + c0];
      v[7] = tile[(a1 + b1) * TS::PX + c1];
    } else {
      const uint64_t a0 = zl * sz, a1 = zh * sz, b0 = yl * sy, b1 = yh * sy;
      v[0] = __ldg(u + a0 + b0 + xl);
      v[1] = __ldg(u + a0 + b0 + xh);
      v[2] = __ldg(u + a0 + b1 + xl);
      v[3] = __ldg(u + a0 + b1 + xh);
      v[4] = __ldg(u + a1 + b0 + xl);
      v[5] = __ldg(u + a1 + b0 + xh);
      v[6] = __ldg(u + a1 + b1 + xl);
      v[7] = __ldg(u + a1 + b1 + xh);
    }
    // interpolate along f (x), then c (y), then r (z) -- the order of
    // GridProcessingKernel3D.hpp:732-877; ratio index = left neighbour
    T tx_ = (T)0.5, ty_ = (T)0.5, tz_ = (T)0.5;
    if (!HALF) {
      if (ix) tx_ = tables[a.roff[l][2] + ax.j - 1];
      if (iy) ty_ = tables[a.roff[l][1] + ay.j - 1];
      if (iz) tz_ = tables[a.roff[l][0] + az.j - 1];
    }
    T f0 = ix ? mgx_lerp(v[0], v[1], tx_) : v[0];
    T f1 = ix ? mgx_lerp(v[2], v[3], tx_) : v[2];
    T f2 = ix ? mgx_lerp(v[4], v[5], tx_) : v[4];
    T f3 = ix ? mgx_lerp(v[6], v[7], tx_) : v[6];
    T g0 = iy ? mgx_lerp(f0, f1, ty_) : f0;
    T g1 = iy ? mgx_lerp(f2, f3, ty_) : f2;
    T h0 = iz ? mgx_lerp(g0, g1, tz_) : g0;
    c = sub_rn(c, h0);
  }
  const uint32_t px = final_pos(ax, gx, nx, l, a.L, nl, 2);
  const uint32_t py = final_pos(ay, gy, ny, l, a.L, nl, 1);
  const uint32_t pz = final_pos(az, gz, nz, l, a.L, nl, 0);
  sink.put(c, (uint64_t)pz * sz + (uint64_t)py * sy + px);
}

} // namespace
