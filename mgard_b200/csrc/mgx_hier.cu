// Fused hierarchical-basis path (L-inf error control, D <= 3): the reference
// skips the orthogonal correction there (infer_orthogonal_projection,
// CompressionLowLevel/Compressor.h:38-40), so coarse nodes are never modified
// (MultiDimension/DataRefactoring.hpp:93-102) and EVERY level's coefficient is
// a function of the original nodal values only:
//
//   c(x) = u(x) - lerp_r(lerp_c(lerp_f(u at the 2^k level-l neighbours)))
//
// with the neighbours at distance 2^(L-l) (clipped to the last node, which every
// level keeps).  One gather pass over the input therefore replaces the
// reference's per-level CopyND + GpkReo3D chain, the lwqzk quantiser, the
// outlier separation and the histogram (5 full-size passes with an int64 code
// stream) -- HBM traffic: 4 B read + 2 B written per fp32 element.
//
// Tiles of TZ x TY x TX nodes (+1 halo plane on the high side of each axis) are
// staged in shared memory with coalesced row loads; because tile origins are
// multiples of the tile size, all neighbours at distance <= T/2 are inside the
// staged tile, and the few coarser nodes (< 1/512 of the grid) gather from
// global memory / L2 instead.  TMA (cp.async.bulk.tensor) is not usable here:
// the 2^k+1 row pitch (e.g. 1025 * 4 B) is not a multiple of 16 bytes.
#include <algorithm>
#include <cmath>
#include <limits>

#include "mgx_device.cuh"
#include "mgx_internal.h"

namespace {

constexpr int TZ = 8, TY = 8, TX = 64;
constexpr int HZ = TZ + 1, HY = TY + 1, HX = TX + 1;
constexpr int HIER_THREADS = 256;

template <typename T> struct HierArgs {
  uint32_t n[3];                // nz, ny, nx (padded with 1)
  uint32_t nl[MGX_MAX_L][3];    // level sizes per axis
  uint32_t roff[MGX_MAX_L][3];  // ratio table offsets per level and axis
  int L;
  int half;                     // all ratios exactly 0.5
  T qv;                         // quantiser reciprocal (all levels equal, L-inf)
  long long shift;
  uint32_t dict;
  uint32_t tiles[3];
  uint64_t ntiles;
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

template <typename T>
__global__ void __launch_bounds__(HIER_THREADS)
    k_hier_forward(const T *__restrict__ u, uint16_t *__restrict__ sym,
                   uint32_t *__restrict__ g_freq,
                   unsigned long long *__restrict__ out_count,
                   unsigned long long *__restrict__ out_idx,
                   long long *__restrict__ out_val, unsigned long long out_cap,
                   const T *__restrict__ tables,
                   const __grid_constant__ HierArgs<T> a) {
  extern __shared__ __align__(16) unsigned char smem[];
  T *tile = (T *)smem;
  uint32_t *s_hist = (uint32_t *)(smem + sizeof(T) * HZ * HY * HX);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  for (uint32_t i = tid; i < a.dict; i += HIER_THREADS)
    s_hist[i] = 0;
  const uint32_t nz = a.n[0], ny = a.n[1], nx = a.n[2];
  const uint64_t sy = nx, sz = (uint64_t)nx * ny;
  const uint32_t *nl = &a.nl[0][0];

  for (uint64_t t = blockIdx.x; t < a.ntiles; t += gridDim.x) {
    const uint32_t tx = (uint32_t)(t % a.tiles[2]);
    const uint32_t ty = (uint32_t)((t / a.tiles[2]) % a.tiles[1]);
    const uint32_t tzz = (uint32_t)(t / ((uint64_t)a.tiles[2] * a.tiles[1]));
    const uint32_t z0 = tzz * TZ, y0 = ty * TY, x0 = tx * TX;
    const uint32_t ez = min((uint32_t)HZ, nz - z0), ey = min((uint32_t)HY, ny - y0),
                   ex = min((uint32_t)HX, nx - x0);
    __syncthreads(); // previous tile fully consumed (also orders the hist init)
    // stage the tile: one warp per (z,y) row, coalesced along x
    for (uint32_t r = warp; r < ez * ey; r += HIER_THREADS / 32) {
      uint32_t rz = r / ey, ry = r % ey;
      const T *src = u + (uint64_t)(z0 + rz) * sz + (uint64_t)(y0 + ry) * sy + x0;
      T *dst = tile + (rz * HY + ry) * HX;
      for (uint32_t x = lane; x < ex; x += 32)
        dst[x] = src[x];
    }
    __syncthreads();
    for (int k = 0; k < (TZ * TY * TX) / HIER_THREADS; k++) {
      const int id = k * HIER_THREADS + tid;
      const uint32_t lx = id % TX, ly = (id / TX) % TY, lz = id / (TX * TY);
      const uint32_t gx = x0 + lx, gy = y0 + ly, gz = z0 + lz;
      const bool valid = gx < nx && gy < ny && gz < nz;
      uint32_t sv = 0;
      if (valid) {
        const AxisInfo ax = axis_info(gx, nx, a.L, nl, 2);
        const AxisInfo ay = axis_info(gy, ny, a.L, nl, 1);
        const AxisInfo az = axis_info(gz, nz, a.L, nl, 0);
        const int l = max(ax.lvl, max(ay.lvl, az.lvl));
        const bool ix = l > 0 && ax.lvl == l, iy = l > 0 && ay.lvl == l,
                   iz = l > 0 && az.lvl == l;
        const uint32_t xl = ix ? ax.lo : gx, xh = ix ? ax.hi : gx;
        const uint32_t yl = iy ? ay.lo : gy, yh = iy ? ay.hi : gy;
        const uint32_t zl = iz ? az.lo : gz, zh = iz ? az.hi : gz;
        T c = tile[(lz * HY + ly) * HX + lx];
        if (l > 0) {
          T v[8];
          const bool in_tile = xl >= x0 && xh < x0 + ex && yl >= y0 &&
                               yh < y0 + ey && zl >= z0 && zh < z0 + ez;
          if (in_tile) {
            const uint32_t a0 = (zl - z0) * HY, a1 = (zh - z0) * HY;
            const uint32_t b0 = yl - y0, b1 = yh - y0, c0 = xl - x0,
                           c1 = xh - x0;
            v[0] = tile[(a0 + b0) * HX + c0];
            v[1] = tile[(a0 + b0) * HX + c1];
            v[2] = tile[(a0 + b1) * HX + c0];
            v[3] = tile[(a0 + b1) * HX + c1];
            v[4] = tile[(a1 + b0) * HX + c0];
            v[5] = tile[(a1 + b0) * HX + c1];
            v[6] = tile[(a1 + b1) * HX + c0];
            v[7] = tile[(a1 + b1) * HX + c1];
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
          if (!a.half) {
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
        long long q = mgx_quantize(c, a.qv, (T)1, a.shift);
        // reordered position: a level-l coefficient keeps, along the axes in
        // which it is a level-l COARSE node, its index in the level-(l-1)
        // grid (coarser levels only permute the [0, n_{l-1}) box, which this
        // element is outside of); along its odd axes it sits in the
        // coefficient block [n_{l-1}, n_l).
        const uint32_t px = final_pos(ax, gx, nx, l, a.L, nl, 2);
        const uint32_t py = final_pos(ay, gy, ny, l, a.L, nl, 1);
        const uint32_t pz = final_pos(az, gz, nz, l, a.L, nl, 0);
        const uint64_t opos = (uint64_t)pz * sz + (uint64_t)py * sy + px;
        if (q < 0 || q >= (long long)a.dict) {
          unsigned long long slot = atomicAdd(out_count, 1ull);
          if (slot < out_cap) {
            out_idx[slot] = opos;
            out_val[slot] = q;
          }
          q = 0;
        }
        sv = (uint32_t)q;
        sym[opos] = (uint16_t)sv;
      }
      mgx_hist_add(s_hist, sv, valid);
    }
  }
  __syncthreads();
  for (uint32_t i = tid; i < a.dict; i += HIER_THREADS) {
    uint32_t v = s_hist[i];
    if (v)
      atomicAdd(&g_freq[i], v);
  }
}

template <typename T>
int hier_forward_t(mgx_ctx *ctx, const void *d_in, int rel, double tol_d,
                   double norm_d) {
  const MgxLevels &lv = ctx->lv;
  const MgxHostTables<T> &tb = *reinterpret_cast<const MgxHostTables<T> *>(
      mgx_is_f32(ctx) ? (const void *)&ctx->tbf : (const void *)&ctx->tbd);
  HierArgs<T> a;
  for (int d = 0; d < 3; d++) {
    a.n[d] = (uint32_t)lv.shape[2 + d];
    a.tiles[d] = (a.n[d] + (d == 0 ? TZ : d == 1 ? TY : TX) - 1) /
                 (d == 0 ? TZ : d == 1 ? TY : TX);
  }
  for (int l = 0; l <= lv.L; l++)
    for (int d = 0; d < 3; d++) {
      a.nl[l][d] = (uint32_t)lv.n[l][2 + d];
      a.roff[l][d] = tb.off.ratio[l][2 + d];
    }
  a.L = lv.L;
  a.half = tb.ratios_half;
  // CalcQuantizers, LinearQuantization.hpp:234-261 (hierarchical branch)
  T tol = (T)tol_d, norm = (T)norm_d;
  double abs_tol = tol;
  if (rel)
    abs_tol *= norm;
  abs_tol *= 2;
  T qv = (T)((abs_tol) / (uint64_t)(lv.L + 1));
  a.qv = 1.0f / qv;
  a.shift = (long long)(ctx->dict / 2);
  a.dict = ctx->dict;
  a.ntiles = (uint64_t)a.tiles[0] * a.tiles[1] * a.tiles[2];
  size_t smem = sizeof(T) * HZ * HY * HX + (size_t)ctx->dict * 4;
  MGX_CUDA(cudaFuncSetAttribute(k_hier_forward<T>,
                                cudaFuncAttributeMaxDynamicSharedMemorySize,
                                (int)smem));
  int occ = 1;
  MGX_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(
      &occ, k_hier_forward<T>, HIER_THREADS, smem));
  unsigned grid = (unsigned)std::min<uint64_t>(
      a.ntiles, (uint64_t)std::max(occ, 1) * ctx->num_sms);
  k_hier_forward<T><<<grid, HIER_THREADS, smem, ctx->stream>>>(
      (const T *)d_in, ctx->d_sym, ctx->d_freq,
      (unsigned long long *)(ctx->d_scalars + MGX_SC_OUTLIERS),
      (unsigned long long *)ctx->d_out_idx, (long long *)ctx->d_out_val,
      ctx->out_cap, (const T *)ctx->d_tables, a);
  ctx->launches++;
  MGX_CUDA(cudaGetLastError());
  return MGX_SUCCESS;
}

} // namespace

int mgxk_hier_forward(mgx_ctx *ctx, const void *d_in, int rel, double tol, bool,
                      double norm) {
  if (ctx->lv.D > 3) {
    ctx->err = "hierarchical fast path needs D <= 3";
    return MGX_FAILURE;
  }
  return mgx_is_f32(ctx) ? hier_forward_t<float>(ctx, d_in, rel, tol, norm)
                         : hier_forward_t<double>(ctx, d_in, rel, tol, norm);
}
