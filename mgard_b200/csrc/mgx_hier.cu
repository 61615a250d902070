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
// global memory / L2 instead.
//
// Interior tiles run a structured schedule instead of a per-element gather:
// pass p (stride h = 2^p) visits the sub-grid of multiples of h row by row,
// one thread per (even, odd) pair of neighbouring sub-grid nodes along x.  The
// odd node is always a level-(L-p) coefficient, the even one is a coefficient
// iff its row has an odd y or z sub-index -- a per-row (warp-uniform in pass 0)
// property -- and both share the same neighbour-row loads.  Only nodes that are
// multiples of the tile edge on every axis (1/512 in 3-D) fall through to the
// generic per-element path, as do the tiles that touch the last node of an
// axis (where the reference's "even indices + last node" coarsening is not a
// plain stride).  TMA (cp.async.bulk.tensor) is not usable here:
// the 2^k+1 row pitch (e.g. 1025 * 4 B) is not a multiple of 16 bytes.
#include "mgx_hier_common.cuh"

// TMA-staged, warp-specialised 3-D forward pass (mgx_hier3.cu)
template <typename T>
int mgxk_hier_forward3(mgx_ctx *ctx, const void *d_in, const void *args, bool half,
                       bool *used);

namespace {


// One structured pass over an interior tile: stride h = 2^P, coefficients of
// level l = L - P.  Thread <-> (even, odd) pair of sub-grid nodes along x.
template <typename T, bool HALF, class TS, int P>
__device__ __forceinline__ void
tile_pass(const T *tile, const T *__restrict__ tables, const HierArgs<T> &a,
          uint32_t x0, uint32_t y0, uint32_t z0, const Sink<T> &sink) {
  constexpr int H = 1 << P;
  constexpr int PPR_SH = TS::LX - P - 1;               // pairs per row
  constexpr int RY_SH = TS::RY ? TS::LY - P : 0;       // sub-rows along y
  constexpr int RZ_SH = TS::RZ ? TS::LZ - P : 0;       // sub-rows along z
  constexpr int NPAIRS = 1 << (PPR_SH + RY_SH + RZ_SH);
  const int l = a.L - P;
  const uint32_t ncx = a.nl[l - 1][2], ncy = a.nl[l - 1][1], ncz = a.nl[l - 1][0];
  const uint64_t sy = a.n[2], sz = (uint64_t)a.n[2] * a.n[1];
  const uint32_t xb = x0 >> (P + 1);
  for (int id = threadIdx.x; id < NPAIRS; id += HIER_THREADS) {
    const int k = id & ((1 << PPR_SH) - 1);
    const int row = id >> PPR_SH;
    const int ys = row & ((1 << RY_SH) - 1);
    const int zs = row >> RY_SH;
    const int iy = TS::RY ? (ys & 1) : 0, iz = TS::RZ ? (zs & 1) : 0;
    const int ly = ys << P, lz = zs << P, lxe = k << (P + 1);
    const uint32_t jy = (y0 + ly) >> P, jz = (z0 + lz) >> P;
    T tx = (T)0.5, ty = (T)0.5, tz = (T)0.5;
    if (!HALF) {
      tx = tables[a.roff[l][2] + ((x0 + lxe) >> P)];
      if (iy) ty = tables[a.roff[l][1] + jy - 1];
      if (iz) tz = tables[a.roff[l][0] + jz - 1];
    }
    const int yl = ly - (iy << P), yh = ly + (iy << P);
    const int zl = lz - (iz << P), zh = lz + (iz << P);
    const T *pO = tile + (lz * TS::HY + ly) * TS::PX + lxe;
    const T own_e = pO[0], own_o = pO[H];
    const T *pA = tile + (zl * TS::HY + yl) * TS::PX + lxe;
    T gE0 = pA[0];
    T gO0 = mgx_lerp(gE0, pA[2 * H], tx);
    if (iy) {
      const T *pB = tile + (zl * TS::HY + yh) * TS::PX + lxe;
      const T eB = pB[0];
      const T fB = mgx_lerp(eB, pB[2 * H], tx);
      gE0 = mgx_lerp(gE0, eB, ty);
      gO0 = mgx_lerp(gO0, fB, ty);
    }
    T hE = gE0, hO = gO0;
    if (iz) {
      const T *pC = tile + (zh * TS::HY + yl) * TS::PX + lxe;
      T gE1 = pC[0];
      T gO1 = mgx_lerp(gE1, pC[2 * H], tx);
      if (iy) {
        const T *pD = tile + (zh * TS::HY + yh) * TS::PX + lxe;
        const T eD = pD[0];
        const T fD = mgx_lerp(eD, pD[2 * H], tx);
        gE1 = mgx_lerp(gE1, eD, ty);
        gO1 = mgx_lerp(gO1, fD, ty);
      }
      hE = mgx_lerp(gE0, gE1, tz);
      hO = mgx_lerp(gO0, gO1, tz);
    }
    // reordered positions: j odd -> coefficient block, j even -> j / 2
    const uint32_t py = (iy ? ncy : 0u) + (jy >> 1);
    const uint32_t pz = (iz ? ncz : 0u) + (jz >> 1);
    const uint64_t rowpos = (uint64_t)pz * sz + (uint64_t)py * sy + xb + k;
    sink.put(sub_rn(own_o, hO), rowpos + ncx);
    if (iy | iz)
      sink.put(sub_rn(own_e, hE), rowpos);
  }
}

// Pass 0 (stride 1, the finest level: 7/8 of all nodes in 3-D) as a march
// along z.  A thread owns one (even, odd) x pair and one (even, odd) y pair of
// rows and walks the z planes of its half of the tile.  On an even plane Z it
// forms, from the two even rows y = 2a, 2a+2,
//     f0 = lerp_x(u[2k], u[2k+2]) on row 2a      gE = lerp_y(u[2k] rows)
//     f2 = ... on row 2a+2                        gO = lerp_y(f0, f2)
// which are the interpolants of the three coefficients of that plane; the four
// coefficients of the odd plane Z+1 interpolate {u[2k], f0, gE, gO} of planes Z
// and Z+2 along z.  Same operations in the same order (x, then y, then z) as
// the per-element gather, with a third of the shared-memory loads and half of
// the lerps, and every address a compile-time offset from one base pointer.
template <typename T> struct PlaneP {
  T e0, f0, gE, gO;
};

__device__ __forceinline__ uint16_t *byte_off(uint16_t *p, uint64_t bytes) {
  return (uint16_t *)((char *)p + bytes);
}

// The same schedule serves every in-tile stride H = 2^P on the sub-grid of
// multiples of H (fewer threads are busy in the coarser passes).
template <typename T, bool HALF, class TS, int P>
__device__ __forceinline__ void
march3d(const T *tile, const T *__restrict__ tables, const HierArgs<T> &a,
        uint32_t x0, uint32_t y0, uint32_t z0, const Sink<T> &sink) {
  static_assert(TS::TX == 64 && TS::TY == 8 && TS::TZ == 8 && HIER_THREADS == 256,
                "thread mapping below assumes an 8 x 8 x 64 tile");
  constexpr int H = 1 << P;
  constexpr int SX = H, SY = H * TS::PX, SZ = H * TS::HY * TS::PX;
  constexpr int NK = 32 >> P, NA = 4 >> P, NB = 4 >> P; // x pairs, y pairs, z pairs
  constexpr int BPT = P == 0 ? 2 : 1;                   // z pairs per thread
  constexpr int ACTIVE = NK * NA * NB / BPT;
  // coarser passes run on their own warps so that no warp carries them all
  constexpr int FIRST = P == 0 ? 0 : P == 1 ? 64 : 128;
  const int tid = (int)threadIdx.x - FIRST;
  if (ACTIVE < HIER_THREADS && (tid < 0 || tid >= ACTIVE))
    return;
  const int k = tid % NK, ya = (tid / NK) % NA, zb = (tid / (NK * NA)) * BPT;
  const int l = a.L - P;
  const uint64_t sy = a.n[2], sz = (uint64_t)a.n[2] * a.n[1];
  const uint64_t oX = a.oct[P][0], oY = a.oct[P][1], oZ = a.oct[P][2];
  T tx = (T)0.5, ty = (T)0.5;
  if (!HALF) {
    tx = tables[a.roff[l][2] + (x0 >> P) + 2 * k];
    ty = tables[a.roff[l][1] + (y0 >> P) + 2 * ya];
  }
  const T *p = tile + (2 * zb) * SZ + (2 * ya) * SY + (2 * k) * SX;
  uint16_t *o = a.sym + ((uint64_t)((z0 >> (P + 1)) + zb) * sz +
                         (uint64_t)((y0 >> (P + 1)) + ya) * sy + (x0 >> (P + 1)) + k);
  typedef typename std::conditional<sizeof(T) == 4, float2, double2>::type V2;
  auto pair = [&](const T *q) -> V2 { // (even, odd) x nodes of a row
    if constexpr (P == 0) {
      return *reinterpret_cast<const V2 *>(q);
    } else {
      V2 v;
      v.x = q[0];
      v.y = q[SX];
      return v;
    }
  };
  auto plane = [&](const T *q, PlaneP<T> &Q, T &o0) {
    const V2 own = pair(q);
    const T r0 = q[2 * SX], e2 = q[2 * SY], r2 = q[2 * SY + 2 * SX];
    Q.e0 = own.x;
    o0 = own.y;
    Q.f0 = mgx_lerp(own.x, r0, tx);
    const T f2 = mgx_lerp(e2, r2, tx);
    Q.gE = mgx_lerp(own.x, e2, ty);
    Q.gO = mgx_lerp(Q.f0, f2, ty);
  };
  PlaneP<T> A, B;
  T oA, oB;
  plane(p, A, oA);
#pragma unroll
  for (int b = 0; b < BPT; b++) {
    const T *q = p + (2 * b) * SZ;
    uint16_t *ob = o + (uint64_t)b * sz;
    const V2 r1 = pair(q + SY);
    plane(q + 2 * SZ, B, oB);
    T tz = (T)0.5;
    if (!HALF)
      tz = tables[a.roff[l][0] + (z0 >> P) + 2 * (zb + b)];
    const V2 s0 = pair(q + SZ);
    const V2 s1 = pair(q + SZ + SY);
    // even plane: rows 2a (odd x only) and 2a + 1; odd plane between
    // A = P(Z) and B = P(Z + 2H): all four nodes
    const T c[7] = {sub_rn(oA, A.f0),
                    sub_rn(r1.x, A.gE),
                    sub_rn(r1.y, A.gO),
                    sub_rn(s0.x, mgx_lerp(A.e0, B.e0, tz)),
                    sub_rn(s0.y, mgx_lerp(A.f0, B.f0, tz)),
                    sub_rn(s1.x, mgx_lerp(A.gE, B.gE, tz)),
                    sub_rn(s1.y, mgx_lerp(A.gO, B.gO, tz))};
    uint16_t *const dst[7] = {byte_off(ob, oX),
                              byte_off(ob, oY),
                              byte_off(ob, oX + oY),
                              byte_off(ob, oZ),
                              byte_off(ob, oZ + oX),
                              byte_off(ob, oZ + oY),
                              byte_off(ob, oZ + oX + oY)};
    sink.template put_group<7>(c, dst);
    A = B;
    oA = oB;
  }
}

// 2-D analogue: march along y; even rows carry {u[2k], f}, odd rows interpolate
// them along y.
template <typename T, bool HALF, class TS>
__device__ __forceinline__ void
pass0_march2d(const T *tile, const T *__restrict__ tables, const HierArgs<T> &a,
              uint32_t x0, uint32_t y0, const Sink<T> &sink) {
  static_assert(TS::TX == 128 && TS::TY == 32 && !TS::RZ && HIER_THREADS == 256,
                "thread mapping below assumes a 32 x 128 tile");
  constexpr int SY = TS::PX;
  const int k = threadIdx.x & 63, yq = threadIdx.x >> 6; // x pair, y quarter
  const int l = a.L;
  const uint64_t sy = a.n[2];
  const uint64_t oX = a.oct[0][0] / 2, oY = a.oct[0][1] / 2; // elements
  T tx = (T)0.5;
  if (!HALF)
    tx = tables[a.roff[l][2] + x0 + 2 * k];
  const T *p = tile + (8 * yq) * SY + 2 * k;
  uint16_t *o = a.sym + ((uint64_t)((y0 >> 1) + 4 * yq) * sy + (x0 >> 1) + k);
  typedef typename std::conditional<sizeof(T) == 4, float2, double2>::type V2;
  V2 ownA = *reinterpret_cast<const V2 *>(p);
  T fA = mgx_lerp(ownA.x, p[2], tx);
#pragma unroll
  for (int b = 0; b < 4; b++) {
    const T *q = p + (2 * b) * SY;
    uint16_t *ob = o + (uint64_t)b * sy;
    sink.put_at(sub_rn(ownA.y, fA), ob + oX);
    const V2 ownB = *reinterpret_cast<const V2 *>(q + 2 * SY);
    const T fB = mgx_lerp(ownB.x, q[2 * SY + 2], tx);
    T ty = (T)0.5;
    if (!HALF)
      ty = tables[a.roff[l][1] + y0 + 8 * yq + 2 * b];
    const V2 s = *reinterpret_cast<const V2 *>(q + SY);
    sink.put_at(sub_rn(s.x, mgx_lerp(ownA.x, ownB.x, ty)), ob + oY);
    sink.put_at(sub_rn(s.y, mgx_lerp(fA, fB, ty)), ob + oY + oX);
    ownA = ownB;
    fA = fB;
  }
}

template <typename T, bool HALF, class TS, int P>
__device__ __forceinline__ void
run_passes(const T *tile, const T *__restrict__ tables, const HierArgs<T> &a,
           uint32_t x0, uint32_t y0, uint32_t z0, const Sink<T> &sink) {
  if constexpr (P < TS::PASSES) {
    if constexpr (TS::RZ)
      march3d<T, HALF, TS, P>(tile, tables, a, x0, y0, z0, sink);
    else if constexpr (P == 0 && TS::RY)
      pass0_march2d<T, HALF, TS>(tile, tables, a, x0, y0, sink);
    else
      tile_pass<T, HALF, TS, P>(tile, tables, a, x0, y0, z0, sink);
    run_passes<T, HALF, TS, P + 1>(tile, tables, a, x0, y0, z0, sink);
  }
}

// 4- or 8-byte asynchronous global -> shared copy (LDGSTS); rows of a 2^k+1
// grid are only element-aligned, so wider copies / TMA boxes are not possible.
template <typename T>
__device__ __forceinline__ void cp_async_elem(T *smem_dst, const T *gsrc) {
  const uint32_t d = (uint32_t)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.ca.shared.global [%0], [%1], %2;\n" ::"r"(d), "l"(gsrc),
               "n"(sizeof(T)));
}
__device__ __forceinline__ void cp_async_commit() {
  asm volatile("cp.async.commit_group;\n" ::);
}
__device__ __forceinline__ void cp_async_wait_all() {
  asm volatile("cp.async.wait_group 0;\n" ::: "memory");
}

template <typename T, bool HALF, class TS>
__global__ void __launch_bounds__(HIER_THREADS, 3)
    k_hier_forward(const T *__restrict__ u, uint32_t *__restrict__ g_freq,
                   const T *__restrict__ tables,
                   const __grid_constant__ HierArgs<T> a) {
  extern __shared__ __align__(16) unsigned char smem[];
  T *const buf0 = (T *)smem;
  uint32_t *s_hist = (uint32_t *)(smem + 2 * sizeof(T) * TS::tile_elems);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  for (uint32_t i = tid; i < a.dict; i += HIER_THREADS)
    s_hist[i] = 0;
  const uint32_t nz = a.n[0], ny = a.n[1], nx = a.n[2];
  const uint64_t sy = nx, sz = (uint64_t)nx * ny;
  const Sink<T> sink{a, (uint32_t)__cvta_generic_to_shared(s_hist)};

  // tile walk t = blockIdx.x + i * gridDim.x, kept as mixed-radix digits
  const uint32_t T2 = a.tiles[2], T1 = a.tiles[1], T0 = a.tiles[0];
  uint32_t cx = blockIdx.x % T2, cy = (blockIdx.x / T2) % T1,
           cz = blockIdx.x / (T2 * T1);
  const uint32_t dx = gridDim.x % T2, dy = (gridDim.x / T2) % T1,
                 dz = gridDim.x / (T2 * T1);

  // stage the (TZ+1) x (TY+1) x (TX+1) box of a tile (clipped to the grid):
  // one warp per row, element-wide asynchronous copies
  auto issue = [&](T *buf, uint32_t tx, uint32_t ty, uint32_t tz) {
    const uint32_t z0 = tz * TS::TZ, y0 = ty * TS::TY, x0 = tx * TS::TX;
    const T *src0 = u + (uint64_t)z0 * sz + (uint64_t)y0 * sy + x0;
    const bool full = x0 + TS::TX <= nx - 1 && (!TS::RY || y0 + TS::TY <= ny - 1) &&
                      (!TS::RZ || z0 + TS::TZ <= nz - 1);
    auto row = [&](T *dst, const T *src) {
#pragma unroll
      for (int x = 0; x < TS::TX; x += 32)
        cp_async_elem(dst + x + lane, src + x + lane);
      if (lane == 0)
        cp_async_elem(dst + TS::TX, src + TS::TX);
    };
    if (full && TS::RZ && TS::TY == HIER_WARPS) {
      // warp w: row y = w of every plane, then row y = TY of plane w
      const T *src = src0 + warp * sy;
      T *dst = buf + warp * TS::PX;
#pragma unroll
      for (int rz = 0; rz < TS::HZ; rz++)
        row(dst + rz * TS::HY * TS::PX, src + rz * sz);
      row(buf + (warp * TS::HY + TS::TY) * TS::PX, src0 + warp * sz + TS::TY * sy);
      if (warp == 0)
        row(buf + (TS::TZ * TS::HY + TS::TY) * TS::PX,
            src0 + TS::TZ * sz + TS::TY * sy);
    } else if (full && !TS::RZ) {
      for (int r = warp; r < TS::ROWS; r += HIER_WARPS)
        row(buf + r * TS::PX, src0 + r * sy);
    } else {
      const uint32_t ez = min((uint32_t)TS::HZ, nz - z0),
                     ey = min((uint32_t)TS::HY, ny - y0),
                     ex = min((uint32_t)TS::TX + 1, nx - x0);
      for (uint32_t r = warp; r < (uint32_t)TS::ROWS; r += HIER_WARPS) {
        const uint32_t rz = r / TS::HY, ry = r - rz * TS::HY;
        if (rz < ez && ry < ey) {
          const T *src = src0 + rz * sz + ry * sy;
          T *dst = buf + r * TS::PX;
          for (uint32_t x = lane; x < ex; x += 32)
            cp_async_elem(dst + x, src + x);
        }
      }
    }
    cp_async_commit();
  };

  bool have = cz < T0;
  if (have)
    issue(buf0, cx, cy, cz);
  for (int it = 0; have; it ^= 1) {
    // next tile of this CTA
    uint32_t nx_ = cx + dx, ny_ = cy + dy, nz_ = cz + dz;
    if (nx_ >= T2) {
      nx_ -= T2;
      ++ny_;
    }
    if (ny_ >= T1) {
      ny_ -= T1;
      ++nz_;
    }
    const bool have_next = nz_ < T0;
    cp_async_wait_all();
    __syncthreads(); // tile `it` landed; tile `it ^ 1` fully consumed
    if (have_next)
      issue(buf0 + (it ^ 1) * TS::tile_elems, nx_, ny_, nz_);
    const T *tile = buf0 + it * TS::tile_elems;
    const uint32_t z0 = cz * TS::TZ, y0 = cy * TS::TY, x0 = cx * TS::TX;
    // interior tile: the +1 halo exists on every real axis and no node of the
    // tile is the last node of its axis
    const bool regular = a.L >= TS::PASSES && x0 + TS::TX <= nx - 1 &&
                         (!TS::RY || y0 + TS::TY <= ny - 1) &&
                         (!TS::RZ || z0 + TS::TZ <= nz - 1);
    if (regular) {
      run_passes<T, HALF, TS, 0>(tile, tables, a, x0, y0, z0, sink);
      // (nodes that are multiples of 2^PASSES on every axis belong to coarser
      //  levels whose neighbours lie outside the tile: k_hier_deep)
    } else {
      const uint32_t ez = min((uint32_t)TS::HZ, nz - z0),
                     ey = min((uint32_t)TS::HY, ny - y0),
                     ex = min((uint32_t)TS::TX + 1, nx - x0);
      const uint32_t vz = min((uint32_t)TS::TZ, nz - z0),
                     vy = min((uint32_t)TS::TY, ny - y0),
                     vx = min((uint32_t)TS::TX, nx - x0);
      const uint32_t cnt = vz * vy * vx;
      for (uint32_t id = tid; id < cnt; id += HIER_THREADS) {
        const uint32_t lx = id % vx, ly = (id / vx) % vy, lz = id / (vx * vy);
        element_generic<T, HALF, TS>(u, tile, tables, a, x0 + lx, y0 + ly, z0 + lz,
                                     x0, y0, z0, ex, ey, ez, sink);
      }
    }
    cx = nx_;
    cy = ny_;
    cz = nz_;
    have = have_next;
  }
  // flush the CTA histogram
  __syncthreads();
  for (uint32_t i = tid; i < a.dict; i += HIER_THREADS) {
    uint32_t v = s_hist[i];
    if (v)
      atomicAdd(&g_freq[i], v);
  }
}

// The nodes the structured passes leave out: multiples of 2^PASSES on every
// axis inside interior tiles (1/512 of a 3-D grid).  Their neighbours are up to
// half the grid away, so they gather straight from global memory / L2; keeping
// them out of the tile loop keeps that latency off every tile's critical path.
template <typename T, bool HALF, class TS>
__global__ void __launch_bounds__(HIER_THREADS)
    k_hier_deep(const T *__restrict__ u, uint32_t *__restrict__ g_freq,
                const T *__restrict__ tables,
                const __grid_constant__ HierArgs<T> a) {
  extern __shared__ __align__(16) unsigned char smem[];
  uint32_t *s_hist = (uint32_t *)smem;
  for (uint32_t i = threadIdx.x; i < a.dict; i += HIER_THREADS)
    s_hist[i] = 0;
  __syncthreads();
  const Sink<T> sink{a, (uint32_t)__cvta_generic_to_shared(s_hist)};
  constexpr int CX = TS::TX >> TS::PASSES,
                CY = TS::RY ? (TS::TY >> TS::PASSES) : 1,
                CZ = TS::RZ ? (TS::TZ >> TS::PASSES) : 1;
  constexpr int PER_TILE = CX * CY * CZ;
  const uint32_t nz = a.n[0], ny = a.n[1], nx = a.n[2];
  const uint64_t total = a.ntiles * PER_TILE;
  for (uint64_t i = blockIdx.x * (uint64_t)HIER_THREADS + threadIdx.x; i < total;
       i += (uint64_t)gridDim.x * HIER_THREADS) {
    const uint64_t t = i / PER_TILE;
    const int id = (int)(i % PER_TILE);
    const uint32_t tx = (uint32_t)(t % a.tiles[2]);
    const uint32_t ty = (uint32_t)((t / a.tiles[2]) % a.tiles[1]);
    const uint32_t tz = (uint32_t)(t / ((uint64_t)a.tiles[2] * a.tiles[1]));
    const uint32_t z0 = tz * TS::TZ, y0 = ty * TS::TY, x0 = tx * TS::TX;
    const bool regular = a.L >= TS::PASSES && x0 + TS::TX <= nx - 1 &&
                         (!TS::RY || y0 + TS::TY <= ny - 1) &&
                         (!TS::RZ || z0 + TS::TZ <= nz - 1);
    if (!regular)
      continue; // boundary tiles handled all their nodes generically
    const uint32_t lx = (id % CX) << TS::PASSES, ly = ((id / CX) % CY) << TS::PASSES,
                   lz = (id / (CX * CY)) << TS::PASSES;
    element_generic<T, HALF, TS>(u, (const T *)nullptr, tables, a, x0 + lx, y0 + ly,
                                 z0 + lz, x0, y0, z0, 0, 0, 0, sink);
  }
  __syncthreads();
  for (uint32_t i = threadIdx.x; i < a.dict; i += HIER_THREADS) {
    uint32_t v = s_hist[i];
    if (v)
      atomicAdd(&g_freq[i], v);
  }
}

// ================================ inverse =====================================
// Dequantize + recompose for the hierarchical basis in one pass over the symbols
// (2 B read + 4 B written per fp32 node instead of the reference's level-by-level
// CopyND / GpkRev3D chain, CoefficientsRestore3D.hpp, DataRefactoring.hpp:179-317).
//
// u(x) = c(x) + lerp_r(lerp_c(lerp_f(u at the level-l neighbours))) needs the
// RECONSTRUCTED coarser nodes, so levels are sequential.  The level-(L-PASSES)
// grid (1/512 of a 3-D field) is reconstructed first by the level kernels on a
// compact array (mgxk_recompose_compact); a tile then holds its extended box
// (+1 halo plane per axis) in shared memory, fills in that coarse lattice, runs
// the strides 2^(PASSES-1) .. 2 over the whole extended lattice (halo nodes of
// those strides are recomputed by every tile that needs them; the stride-1 halo
// nodes are never needed), and finally the stride-1 nodes of its interior, which
// go straight to global memory.

// value of an outlier (symbol 0 that is in the sorted outlier list), or code 0
template <typename T>
__device__ __noinline__ long long hier_outlier_code(const HierArgs<T> &a,
                                                    uint64_t opos) {
  unsigned long long lo = 0, hi = a.n_out;
  while (lo < hi) {
    const unsigned long long mid = (lo + hi) >> 1;
    if (a.out_idx[mid] < opos)
      lo = mid + 1;
    else
      hi = mid;
  }
  if (lo < a.n_out && a.out_idx[lo] == opos)
    return a.out_val[lo];
  return 0;
}

// lwqzk dequantise (LinearQuantization.hpp:113-126), volume factor 1
template <typename T>
__device__ __forceinline__ T coef_of(const HierArgs<T> &a, uint32_t s,
                                     const uint16_t *at) {
  if (__builtin_expect(s == 0 && a.n_out != 0, 0)) {
    const long long q = hier_outlier_code<T>(a, (uint64_t)(at - a.sym));
    return mul_rn(a.dq, (T)(q - (long long)a.shift));
  }
  return mul_rn(a.dq, (T)((int)s - a.shift));
}

// position of node index i in the level-(L-P) grid (i is a node of that grid)
__device__ __forceinline__ uint32_t coarse_index(uint32_t i, uint32_t n, int P,
                                                 uint32_t cn) {
  return (n > 1 && i == n - 1) ? cn - 1 : (i >> P);
}

// One node of level `want` inside the staged (extended) tile.
template <typename T, bool HALF, class TS>
__device__ __forceinline__ void
inv_node(T *tile, const T *__restrict__ tables, const HierArgs<T> &a, uint32_t gx,
         uint32_t gy, uint32_t gz, uint32_t x0, uint32_t y0, uint32_t z0, int want) {
  const uint32_t nz = a.n[0], ny = a.n[1], nx = a.n[2];
  const uint64_t sy = nx, sz = (uint64_t)nx * ny;
  const uint32_t *nl = &a.nl[0][0];
  const AxisInfo ax = axis_info(gx, nx, a.L, nl, 2);
  const AxisInfo ay = axis_info(gy, ny, a.L, nl, 1);
  const AxisInfo az = axis_info(gz, nz, a.L, nl, 0);
  const int l = max(ax.lvl, max(ay.lvl, az.lvl));
  if (l != want)
    return;
  const bool ix = ax.lvl == l, iy = ay.lvl == l, iz = az.lvl == l;
  const uint32_t xl = (ix ? ax.lo : gx) - x0, xh = (ix ? ax.hi : gx) - x0;
  const uint32_t yl = (iy ? ay.lo : gy) - y0, yh = (iy ? ay.hi : gy) - y0;
  const uint32_t zl = (iz ? az.lo : gz) - z0, zh = (iz ? az.hi : gz) - z0;
  const uint32_t a0 = zl * TS::HY, a1 = zh * TS::HY;
  const T v0 = tile[(a0 + yl) * TS::PX + xl], v1 = tile[(a0 + yl) * TS::PX + xh];
  const T v2 = tile[(a0 + yh) * TS::PX + xl], v3 = tile[(a0 + yh) * TS::PX + xh];
  const T v4 = tile[(a1 + yl) * TS::PX + xl], v5 = tile[(a1 + yl) * TS::PX + xh];
  const T v6 = tile[(a1 + yh) * TS::PX + xl], v7 = tile[(a1 + yh) * TS::PX + xh];
  T tx_ = (T)0.5, ty_ = (T)0.5, tz_ = (T)0.5;
  if (!HALF) {
    if (ix) tx_ = tables[a.roff[l][2] + ax.j - 1];
    if (iy) ty_ = tables[a.roff[l][1] + ay.j - 1];
    if (iz) tz_ = tables[a.roff[l][0] + az.j - 1];
  }
  const T f0 = ix ? mgx_lerp(v0, v1, tx_) : v0;
  const T f1 = ix ? mgx_lerp(v2, v3, tx_) : v2;
  const T f2 = ix ? mgx_lerp(v4, v5, tx_) : v4;
  const T f3 = ix ? mgx_lerp(v6, v7, tx_) : v6;
  const T g0 = iy ? mgx_lerp(f0, f1, ty_) : f0;
  const T g1 = iy ? mgx_lerp(f2, f3, ty_) : f2;
  const T h0 = iz ? mgx_lerp(g0, g1, tz_) : g0;
  const uint32_t px = final_pos(ax, gx, nx, l, a.L, nl, 2);
  const uint32_t py = final_pos(ay, gy, ny, l, a.L, nl, 1);
  const uint32_t pz = final_pos(az, gz, nz, l, a.L, nl, 0);
  const uint16_t *at = a.sym + ((uint64_t)pz * sz + (uint64_t)py * sy + px);
  const T c = coef_of<T>(a, (uint32_t)__ldg(at), at);
  tile[((gz - z0) * TS::HY + (gy - y0)) * TS::PX + (gx - x0)] = add_rn(c, h0);
}

// addresses of the 7 companion symbols of block (bi, bj, bk) of pass P
template <typename T, int P>
__device__ __forceinline__ void
block_sym_ptrs(const HierArgs<T> &a, uint32_t x0, uint32_t y0, uint32_t z0, int bi,
               int bj, int bk, const uint16_t *(&at)[7]) {
  const uint64_t sy = a.n[2], sz = (uint64_t)a.n[2] * a.n[1];
  const uint64_t oX = a.oct[P][0], oY = a.oct[P][1], oZ = a.oct[P][2];
  uint16_t *ob = a.sym + ((uint64_t)((z0 >> (P + 1)) + bk) * sz +
                          (uint64_t)((y0 >> (P + 1)) + bj) * sy +
                          (x0 >> (P + 1)) + bi);
  at[0] = byte_off(ob, oX);
  at[1] = byte_off(ob, oY);
  at[2] = byte_off(ob, oX + oY);
  at[3] = byte_off(ob, oZ);
  at[4] = byte_off(ob, oZ + oX);
  at[5] = byte_off(ob, oZ + oY);
  at[6] = byte_off(ob, oZ + oX + oY);
}

// which of the 7 companions exist in a block with missing odd halves
__device__ __forceinline__ void block_exists(bool hx, bool hy, bool hz, bool (&e)[7]) {
  e[0] = !hx;
  e[1] = !hy;
  e[2] = !hx && !hy;
  e[3] = !hz;
  e[4] = !hx && !hz;
  e[5] = !hy && !hz;
  e[6] = !hx && !hy && !hz;
}

template <typename T, bool HALF, class TS, int P>
__device__ __forceinline__ void
inv_block_core(T *tile, const T *__restrict__ tables, const HierArgs<T> &a,
               uint32_t x0, uint32_t y0, uint32_t z0, int bi, int bj, int bk, bool hx,
               bool hy, bool hz, const uint32_t (&sv)[7],
               const uint16_t *const (&at)[7]) {
  constexpr int H = 1 << P;
  constexpr int SX = H, SY = H * TS::PX, SZ = H * TS::HY * TS::PX;
  const int l = a.L - P;
  bool e[7];
  block_exists(hx, hy, hz, e);
  T tx = (T)0.5, ty = (T)0.5, tz = (T)0.5;
  if (!HALF) {
    if (!hx) tx = tables[a.roff[l][2] + (x0 >> P) + 2 * bi];
    if (!hy) ty = tables[a.roff[l][1] + (y0 >> P) + 2 * bj];
    if (!hz) tz = tables[a.roff[l][0] + (z0 >> P) + 2 * bk];
  }
  T *q = tile + (2 * bk) * SZ + (2 * bj) * SY + (2 * bi) * SX;
  auto plane = [&](const T *r, PlaneP<T> &Q) {
    const T e0 = r[0];
    const T r0 = hx ? e0 : r[2 * SX];
    const T e2 = hy ? e0 : r[2 * SY];
    const T r2 = (hx || hy) ? e0 : r[2 * SY + 2 * SX];
    Q.e0 = e0;
    Q.f0 = mgx_lerp(e0, r0, tx);
    const T f2 = mgx_lerp(e2, r2, tx);
    Q.gE = mgx_lerp(e0, e2, ty);
    Q.gO = mgx_lerp(Q.f0, f2, ty);
  };
  PlaneP<T> A, B;
  plane(q, A);
  if (e[0]) q[SX] = add_rn(coef_of<T>(a, sv[0], at[0]), A.f0);
  if (e[1]) q[SY] = add_rn(coef_of<T>(a, sv[1], at[1]), A.gE);
  if (e[2]) q[SY + SX] = add_rn(coef_of<T>(a, sv[2], at[2]), A.gO);
  if (!hz) {
    plane(q + 2 * SZ, B);
    q[SZ] = add_rn(coef_of<T>(a, sv[3], at[3]), mgx_lerp(A.e0, B.e0, tz));
    if (e[4])
      q[SZ + SX] = add_rn(coef_of<T>(a, sv[4], at[4]), mgx_lerp(A.f0, B.f0, tz));
    if (e[5])
      q[SZ + SY] = add_rn(coef_of<T>(a, sv[5], at[5]), mgx_lerp(A.gE, B.gE, tz));
    if (e[6])
      q[SZ + SY + SX] =
          add_rn(coef_of<T>(a, sv[6], at[6]), mgx_lerp(A.gO, B.gO, tz));
  }
}

// One 2 x 2 x 2 block of the stride-2^P lattice of an interior tile: anchor node
// (2bi, 2bj, 2bk) * 2^P is known, its up to 7 companions are reconstructed with
// the march's operation order (x, then y, then z lerps).  The last block of an
// axis is the +1 halo plane: it has no odd half along that axis (hx / hy / hz).
template <typename T, bool HALF, class TS, int P>
__device__ __forceinline__ void
inv_block(T *tile, const T *__restrict__ tables, const HierArgs<T> &a, uint32_t x0,
          uint32_t y0, uint32_t z0, int bi, int bj, int bk, bool hx, bool hy,
          bool hz) {
  const uint16_t *at[7];
  block_sym_ptrs<T, P>(a, x0, y0, z0, bi, bj, bk, at);
  bool e[7];
  block_exists(hx, hy, hz, e);
  uint32_t sv[7];
#pragma unroll
  for (int i = 0; i < 7; i++)
    sv[i] = e[i] ? (uint32_t)__ldg(at[i]) : 0u; // all loads first
  inv_block_core<T, HALF, TS, P>(tile, tables, a, x0, y0, z0, bi, bj, bk, hx, hy, hz,
                                 sv, at);
}

// pass P over the extended lattice (interior_only: stride-1 pass of tiles whose
// results are copied out of shared memory afterwards)
template <typename T, bool HALF, class TS, int P>
__device__ __forceinline__ void
inv_pass(T *tile, const T *__restrict__ tables, const HierArgs<T> &a, uint32_t x0,
         uint32_t y0, uint32_t z0, bool interior_only) {
  constexpr int NBX = (TS::TX >> (P + 1)) + 1;
  constexpr int NBY = TS::RY ? (TS::TY >> (P + 1)) + 1 : 1;
  constexpr int NBZ = TS::RZ ? (TS::TZ >> (P + 1)) + 1 : 1;
  // spread the coarse passes over different warps
  constexpr int FIRST = (P == 1 || NBX * NBY * NBZ >= HIER_THREADS) ? 0
                        : (P == 2 ? 160 : 224);
  for (int id = (int)threadIdx.x - FIRST; id >= 0 && id < NBX * NBY * NBZ;
       id += HIER_THREADS - FIRST) {
    const int bi = id % NBX, bj = (id / NBX) % NBY, bk = id / (NBX * NBY);
    const bool hx = bi == NBX - 1, hy = !TS::RY || bj == NBY - 1,
               hz = !TS::RZ || bk == NBZ - 1;
    if (interior_only && (hx || (TS::RY && hy) || (TS::RZ && hz)))
      continue;
    inv_block<T, HALF, TS, P>(tile, tables, a, x0, y0, z0, bi, bj, bk, hx, hy, hz);
  }
}

// passes PASSES-1 .. LAST of an interior tile, a barrier before each
template <typename T, bool HALF, class TS, int P, int LAST>
__device__ __forceinline__ void
inv_coarse_passes(T *tile, const T *__restrict__ tables, const HierArgs<T> &a,
                  uint32_t x0, uint32_t y0, uint32_t z0) {
  if constexpr (P >= LAST) {
    __syncthreads();
    inv_pass<T, HALF, TS, P>(tile, tables, a, x0, y0, z0, P == 0);
    inv_coarse_passes<T, HALF, TS, P - 1, LAST>(tile, tables, a, x0, y0, z0);
  }
}

template <typename T, bool HALF, class TS>
__global__ void __launch_bounds__(HIER_THREADS, 5)
    k_hier_inverse(T *__restrict__ out, const T *__restrict__ tables,
                   const __grid_constant__ HierArgs<T> a) {
  extern __shared__ __align__(16) unsigned char smem[];
  T *tile = (T *)smem;
  const int tid = threadIdx.x;
  const uint32_t nz = a.n[0], ny = a.n[1], nx = a.n[2];
  const uint64_t sy = nx, sz = (uint64_t)nx * ny;
  constexpr int P = TS::PASSES;
  const uint32_t T2 = a.tiles[2], T1 = a.tiles[1], T0 = a.tiles[0];
  uint32_t cx = blockIdx.x % T2, cy = (blockIdx.x / T2) % T1,
           cz = blockIdx.x / (T2 * T1);
  const uint32_t dx = gridDim.x % T2, dy = (gridDim.x / T2) % T1,
                 dz = gridDim.x / (T2 * T1);
  while (cz < T0) {
    const uint32_t z0 = cz * TS::TZ, y0 = cy * TS::TY, x0 = cx * TS::TX;
    const bool regular = x0 + TS::TX <= nx - 1 &&
                         (!TS::RY || y0 + TS::TY <= ny - 1) &&
                         (!TS::RZ || z0 + TS::TZ <= nz - 1);
    __syncthreads(); // previous tile no longer read
    if (regular) {
      // stride-1 symbols of this thread's blocks, fetched before anything else
      uint16_t sv[2][7];
      const uint16_t *ob0 = nullptr;
      uint64_t oX = 0, oY = 0, oZ = 0;
      int k = 0, ya = 0, zb = 0;
      if constexpr (TS::RZ) {
        k = tid & 31;
        ya = (tid >> 5) & 3;
        zb = (tid >> 7) * 2;
        oX = a.oct[0][0];
        oY = a.oct[0][1];
        oZ = a.oct[0][2];
        ob0 = a.sym + ((uint64_t)((z0 >> 1) + zb) * sz +
                       (uint64_t)((y0 >> 1) + ya) * sy + (x0 >> 1) + k);
#pragma unroll
        for (int b = 0; b < 2; b++) {
          const uint16_t *ob = ob0 + (uint64_t)b * sz;
          sv[b][0] = __ldg(byte_off((uint16_t *)ob, oX));
          sv[b][1] = __ldg(byte_off((uint16_t *)ob, oY));
          sv[b][2] = __ldg(byte_off((uint16_t *)ob, oX + oY));
          sv[b][3] = __ldg(byte_off((uint16_t *)ob, oZ));
          sv[b][4] = __ldg(byte_off((uint16_t *)ob, oZ + oX));
          sv[b][5] = __ldg(byte_off((uint16_t *)ob, oZ + oY));
          sv[b][6] = __ldg(byte_off((uint16_t *)ob, oZ + oX + oY));
        }
      }
      // coarse lattice (stride 2^P) of the extended tile
      constexpr int CX = (TS::TX >> P) + 1, CY = TS::RY ? (TS::TY >> P) + 1 : 1,
                    CZ = TS::RZ ? (TS::TZ >> P) + 1 : 1;
      for (int id = tid; id < CX * CY * CZ; id += HIER_THREADS) {
        const int i = id % CX, j = (id / CX) % CY, kk = id / (CX * CY);
        const uint32_t gx = x0 + (i << P), gy = y0 + (j << P), gz = z0 + (kk << P);
        const uint32_t ci = coarse_index(gx, nx, P, a.cdim[2]),
                       cj = coarse_index(gy, ny, P, a.cdim[1]),
                       ck = coarse_index(gz, nz, P, a.cdim[0]);
        tile[(((kk << P) * TS::HY) + (j << P)) * TS::PX + (i << P)] =
            a.coarse[((uint64_t)ck * a.cdim[1] + cj) * a.cdim[2] + ci];
      }
      inv_coarse_passes<T, HALF, TS, P - 1, (TS::RZ ? 1 : 0)>(tile, tables, a, x0, y0,
                                                               z0);
      __syncthreads();
      if constexpr (TS::RZ) {
        // stride 1: the forward march run backwards (x, then y, then z lerps of
        // the reconstructed even lattice), results straight to global memory
        constexpr int SY = TS::PX, SZ = TS::HY * TS::PX;
        const int l = a.L;
        T tx = (T)0.5, ty = (T)0.5;
        if (!HALF) {
          tx = tables[a.roff[l][2] + x0 + 2 * k];
          ty = tables[a.roff[l][1] + y0 + 2 * ya];
        }
        const T *p = tile + (2 * zb) * SZ + (2 * ya) * SY + 2 * k;
        T *g = out + ((uint64_t)(z0 + 2 * zb) * sz + (uint64_t)(y0 + 2 * ya) * sy +
                      x0 + 2 * k);
        auto plane = [&](const T *q, PlaneP<T> &Q) {
          const T e0 = q[0], r0 = q[2], e2 = q[2 * SY], r2 = q[2 * SY + 2];
          Q.e0 = e0;
          Q.f0 = mgx_lerp(e0, r0, tx);
          const T f2 = mgx_lerp(e2, r2, tx);
          Q.gE = mgx_lerp(e0, e2, ty);
          Q.gO = mgx_lerp(Q.f0, f2, ty);
        };
        PlaneP<T> A, B;
        plane(p, A);
#pragma unroll
        for (int b = 0; b < 2; b++) {
          const T *q = p + (2 * b) * SZ;
          const uint16_t *ob = ob0 + (uint64_t)b * sz;
          T *gb = g + (uint64_t)(2 * b) * sz;
          plane(q + 2 * SZ, B);
          T tz = (T)0.5;
          if (!HALF)
            tz = tables[a.roff[l][0] + z0 + 2 * (zb + b)];
          gb[0] = A.e0;
          gb[1] = add_rn(coef_of<T>(a, sv[b][0], byte_off((uint16_t *)ob, oX)), A.f0);
          gb[sy] = add_rn(coef_of<T>(a, sv[b][1], byte_off((uint16_t *)ob, oY)), A.gE);
          gb[sy + 1] =
              add_rn(coef_of<T>(a, sv[b][2], byte_off((uint16_t *)ob, oX + oY)), A.gO);
          gb[sz] = add_rn(coef_of<T>(a, sv[b][3], byte_off((uint16_t *)ob, oZ)),
                          mgx_lerp(A.e0, B.e0, tz));
          gb[sz + 1] = add_rn(coef_of<T>(a, sv[b][4], byte_off((uint16_t *)ob, oZ + oX)),
                              mgx_lerp(A.f0, B.f0, tz));
          gb[sz + sy] = add_rn(coef_of<T>(a, sv[b][5], byte_off((uint16_t *)ob, oZ + oY)),
                               mgx_lerp(A.gE, B.gE, tz));
          gb[sz + sy + 1] =
              add_rn(coef_of<T>(a, sv[b][6], byte_off((uint16_t *)ob, oZ + oX + oY)),
                     mgx_lerp(A.gO, B.gO, tz));
          A = B;
        }
      } else {
        // 1-D / 2-D tiles: the stride-1 pass ran in shared memory; copy out
        for (int id = tid; id < TS::TZ * TS::TY * TS::TX; id += HIER_THREADS) {
          const int i = id % TS::TX, j = (id / TS::TX) % TS::TY,
                    kk = id / (TS::TX * TS::TY);
          out[(uint64_t)(z0 + kk) * sz + (uint64_t)(y0 + j) * sy + x0 + i] =
              tile[(kk * TS::HY + j) * TS::PX + i];
        }
      }
    } else {
      // tile touching the last node of an axis: every level by the generic
      // per-node rule on the clipped extended box
      const uint32_t ez = min((uint32_t)TS::HZ, nz - z0),
                     ey = min((uint32_t)TS::HY, ny - y0),
                     ex = min((uint32_t)TS::TX + 1, nx - x0);
      const uint32_t cnt = ez * ey * ex;
      const uint32_t *nl = &a.nl[0][0];
      for (uint32_t id = tid; id < cnt; id += HIER_THREADS) {
        const uint32_t i = id % ex, j = (id / ex) % ey, kk = id / (ex * ey);
        const uint32_t gx = x0 + i, gy = y0 + j, gz = z0 + kk;
        const int l = max(axis_info(gx, nx, a.L, nl, 2).lvl,
                          max(axis_info(gy, ny, a.L, nl, 1).lvl,
                              axis_info(gz, nz, a.L, nl, 0).lvl));
        if (l <= a.L - P) {
          const uint32_t ci = coarse_index(gx, nx, P, a.cdim[2]),
                         cj = coarse_index(gy, ny, P, a.cdim[1]),
                         ck = coarse_index(gz, nz, P, a.cdim[0]);
          tile[(kk * TS::HY + j) * TS::PX + i] =
              a.coarse[((uint64_t)ck * a.cdim[1] + cj) * a.cdim[2] + ci];
        }
      }
      for (int pp = P - 1; pp >= 0; pp--) {
        __syncthreads();
        for (uint32_t id = tid; id < cnt; id += HIER_THREADS) {
          const uint32_t i = id % ex, j = (id / ex) % ey, kk = id / (ex * ey);
          inv_node<T, HALF, TS>(tile, tables, a, x0 + i, y0 + j, z0 + kk, x0, y0, z0,
                                a.L - pp);
        }
      }
      __syncthreads();
      const uint32_t vz = min((uint32_t)TS::TZ, nz - z0),
                     vy = min((uint32_t)TS::TY, ny - y0),
                     vx = min((uint32_t)TS::TX, nx - x0);
      for (uint32_t id = tid; id < vz * vy * vx; id += HIER_THREADS) {
        const uint32_t i = id % vx, j = (id / vx) % vy, kk = id / (vx * vy);
        out[(uint64_t)(z0 + kk) * sz + (uint64_t)(y0 + j) * sy + x0 + i] =
            tile[(kk * TS::HY + j) * TS::PX + i];
      }
    }
    cx += dx;
    cy += dy;
    cz += dz;
    if (cx >= T2) {
      cx -= T2;
      ++cy;
    }
    if (cy >= T1) {
      cy -= T1;
      ++cz;
    }
  }
}

// 3-D specialisation with everything a tile needs from global memory (its
// stride-1 symbols, the symbols of its two coarse passes, its coarse-lattice
// values) fetched into registers ONE TILE AHEAD: the loads of tile i+1 are in
// flight while tile i runs its four barrier-separated phases out of registers and
// shared memory, so no global latency sits on a tile's critical path.
// Static roles: threads 0..152 own the 17x3x3 blocks of pass 1, threads 160..195
// the 9x2x2 blocks of pass 2, threads 224..255 the 9x2x2 coarse-lattice nodes;
// every thread owns two stride-1 blocks (x pair, y pair, z half) of the interior.
template <typename T> struct InvPref {
  uint32_t p0[7]; // stride-1 symbols: low half = first z block, high = second
  uint32_t pc[7]; // symbols of this thread's coarse-pass block
  T deep[2];
};

template <typename T, bool HALF>
__global__ void __launch_bounds__(HIER_THREADS, 4)
    k_hier_inverse3d(T *__restrict__ out, const T *__restrict__ tables,
                     const __grid_constant__ HierArgs<T> a) {
  typedef TileShape<3, 3, 6> TS;
  constexpr int P = TS::PASSES;
  extern __shared__ __align__(16) unsigned char smem[];
  T *tile = (T *)smem;
  const int tid = threadIdx.x;
  const uint32_t nz = a.n[0], ny = a.n[1], nx = a.n[2];
  const uint64_t sy = nx, sz = (uint64_t)nx * ny;
  // static roles
  const int k = tid & 31, ya = (tid >> 5) & 3, zb = (tid >> 7) * 2;
  const bool r1 = tid < 153, r2 = tid >= 160 && tid < 196;
  const int id1 = tid, id2 = tid - 160;
  const int b1i = id1 % 17, b1j = (id1 / 17) % 3, b1k = id1 / 51;
  const int b2i = id2 % 9, b2j = (id2 / 9) % 2, b2k = id2 / 18;
  const bool h1x = b1i == 16, h1y = b1j == 2, h1z = b1k == 2;
  const bool h2x = b2i == 8, h2y = b2j == 1, h2z = b2k == 1;

  auto is_regular = [&](uint32_t x0, uint32_t y0, uint32_t z0) {
    return x0 + TS::TX <= nx - 1 && y0 + TS::TY <= ny - 1 && z0 + TS::TZ <= nz - 1;
  };
  auto deep_value = [&](int id, uint32_t x0, uint32_t y0, uint32_t z0) {
    const int i = id % 9, j = (id / 9) % 2, kk = id / 18;
    const uint32_t ci = coarse_index(x0 + (i << P), nx, P, a.cdim[2]),
                   cj = coarse_index(y0 + (j << P), ny, P, a.cdim[1]),
                   ck = coarse_index(z0 + (kk << P), nz, P, a.cdim[0]);
    return __ldg(a.coarse + ((uint64_t)ck * a.cdim[1] + cj) * a.cdim[2] + ci);
  };
  auto prefetch = [&](uint32_t x0, uint32_t y0, uint32_t z0, InvPref<T> &pf) {
    if (!is_regular(x0, y0, z0))
      return;
    const uint16_t *at[7];
    block_sym_ptrs<T, 0>(a, x0, y0, z0, k, ya, zb, at);
#pragma unroll
    for (int i = 0; i < 7; i++)
      pf.p0[i] = (uint32_t)__ldg(at[i]) | ((uint32_t)__ldg(at[i] + sz) << 16);
    bool e[7];
    if (r1) {
      block_sym_ptrs<T, 1>(a, x0, y0, z0, b1i, b1j, b1k, at);
      block_exists(h1x, h1y, h1z, e);
#pragma unroll
      for (int i = 0; i < 7; i++)
        pf.pc[i] = e[i] ? (uint32_t)__ldg(at[i]) : 0u;
    } else if (r2) {
      block_sym_ptrs<T, 2>(a, x0, y0, z0, b2i, b2j, b2k, at);
      block_exists(h2x, h2y, h2z, e);
#pragma unroll
      for (int i = 0; i < 7; i++)
        pf.pc[i] = e[i] ? (uint32_t)__ldg(at[i]) : 0u;
      // the 36 deep (coarse-lattice) values of the tile are staged by the same
      // 36 threads that run the stride-4 pass on them: the hand-over is then a
      // barrier between their two warps, not a CTA barrier
      pf.deep[0] = deep_value(id2, x0, y0, z0);
    }
  };

  const uint32_t T2 = a.tiles[2], T1 = a.tiles[1], T0 = a.tiles[0];
  uint32_t cx = blockIdx.x % T2, cy = (blockIdx.x / T2) % T1,
           cz = blockIdx.x / (T2 * T1);
  const uint32_t dx = gridDim.x % T2, dy = (gridDim.x / T2) % T1,
                 dz = gridDim.x / (T2 * T1);
  InvPref<T> cur = {}, nxt = {};
  if (cz < T0)
    prefetch(cx * TS::TX, cy * TS::TY, cz * TS::TZ, cur);
  while (cz < T0) {
    const uint32_t z0 = cz * TS::TZ, y0 = cy * TS::TY, x0 = cx * TS::TX;
    uint32_t ncx = cx + dx, ncy = cy + dy, ncz = cz + dz;
    if (ncx >= T2) {
      ncx -= T2;
      ++ncy;
    }
    if (ncy >= T1) {
      ncy -= T1;
      ++ncz;
    }
    if (ncz < T0)
      prefetch(ncx * TS::TX, ncy * TS::TY, ncz * TS::TZ, nxt);
    __syncthreads(); // previous tile no longer read
    if (is_regular(x0, y0, z0)) {
      if (r2) {
        const int i = id2 % 9, j = (id2 / 9) % 2, kk = id2 / 18;
        tile[(((kk << P) * TS::HY) + (j << P)) * TS::PX + (i << P)] = cur.deep[0];
      }
      if (tid >= 160 && tid < 224) // warps 5 and 6 hold all r2 threads
        asm volatile("bar.sync 1, 64;" ::: "memory");
      const uint16_t *at[7];
      if (r2) {
        block_sym_ptrs<T, 2>(a, x0, y0, z0, b2i, b2j, b2k, at);
        inv_block_core<T, HALF, TS, 2>(tile, tables, a, x0, y0, z0, b2i, b2j, b2k, h2x,
                                       h2y, h2z, cur.pc, at);
      }
      __syncthreads();
      if (r1) {
        block_sym_ptrs<T, 1>(a, x0, y0, z0, b1i, b1j, b1k, at);
        inv_block_core<T, HALF, TS, 1>(tile, tables, a, x0, y0, z0, b1i, b1j, b1k, h1x,
                                       h1y, h1z, cur.pc, at);
      }
      __syncthreads();
      // stride 1: the forward march run backwards, results straight to global
      constexpr int SY = TS::PX, SZ = TS::HY * TS::PX;
      const int l = a.L;
      T tx = (T)0.5, ty = (T)0.5;
      if (!HALF) {
        tx = tables[a.roff[l][2] + x0 + 2 * k];
        ty = tables[a.roff[l][1] + y0 + 2 * ya];
      }
      block_sym_ptrs<T, 0>(a, x0, y0, z0, k, ya, zb, at);
      const T *p = tile + (2 * zb) * SZ + (2 * ya) * SY + 2 * k;
      T *g = out + ((uint64_t)(z0 + 2 * zb) * sz + (uint64_t)(y0 + 2 * ya) * sy + x0 +
                    2 * k);
      auto plane = [&](const T *q, PlaneP<T> &Q) {
        const T e0 = q[0], r0 = q[2], e2 = q[2 * SY], r2_ = q[2 * SY + 2];
        Q.e0 = e0;
        Q.f0 = mgx_lerp(e0, r0, tx);
        const T f2 = mgx_lerp(e2, r2_, tx);
        Q.gE = mgx_lerp(e0, e2, ty);
        Q.gO = mgx_lerp(Q.f0, f2, ty);
      };
      PlaneP<T> A, B;
      plane(p, A);
#pragma unroll
      for (int b = 0; b < 2; b++) {
        const T *q = p + (2 * b) * SZ;
        T *gb = g + (uint64_t)(2 * b) * sz;
        plane(q + 2 * SZ, B);
        T tz = (T)0.5;
        if (!HALF)
          tz = tables[a.roff[l][0] + z0 + 2 * (zb + b)];
        uint32_t sv[7];
#pragma unroll
        for (int i = 0; i < 7; i++)
          sv[i] = (cur.p0[i] >> (16 * b)) & 0xffffu;
        const uint64_t bo = (uint64_t)b * sz;
        gb[0] = A.e0;
        gb[1] = add_rn(coef_of<T>(a, sv[0], at[0] + bo), A.f0);
        gb[sy] = add_rn(coef_of<T>(a, sv[1], at[1] + bo), A.gE);
        gb[sy + 1] = add_rn(coef_of<T>(a, sv[2], at[2] + bo), A.gO);
        gb[sz] = add_rn(coef_of<T>(a, sv[3], at[3] + bo), mgx_lerp(A.e0, B.e0, tz));
        gb[sz + 1] = add_rn(coef_of<T>(a, sv[4], at[4] + bo), mgx_lerp(A.f0, B.f0, tz));
        gb[sz + sy] = add_rn(coef_of<T>(a, sv[5], at[5] + bo), mgx_lerp(A.gE, B.gE, tz));
        gb[sz + sy + 1] =
            add_rn(coef_of<T>(a, sv[6], at[6] + bo), mgx_lerp(A.gO, B.gO, tz));
        A = B;
      }
    } else {
      // tile touching the last node of an axis: every level by the generic
      // per-node rule on the clipped extended box
      const uint32_t ez = min((uint32_t)TS::HZ, nz - z0),
                     ey = min((uint32_t)TS::HY, ny - y0),
                     ex = min((uint32_t)TS::TX + 1, nx - x0);
      const uint32_t cnt = ez * ey * ex;
      const uint32_t *nl = &a.nl[0][0];
      for (uint32_t id = tid; id < cnt; id += HIER_THREADS) {
        const uint32_t i = id % ex, j = (id / ex) % ey, kk = id / (ex * ey);
        const uint32_t gx = x0 + i, gy = y0 + j, gz = z0 + kk;
        const int l = max(axis_info(gx, nx, a.L, nl, 2).lvl,
                          max(axis_info(gy, ny, a.L, nl, 1).lvl,
                              axis_info(gz, nz, a.L, nl, 0).lvl));
        if (l <= a.L - P) {
          const uint32_t ci = coarse_index(gx, nx, P, a.cdim[2]),
                         cj = coarse_index(gy, ny, P, a.cdim[1]),
                         ck = coarse_index(gz, nz, P, a.cdim[0]);
          tile[(kk * TS::HY + j) * TS::PX + i] =
              a.coarse[((uint64_t)ck * a.cdim[1] + cj) * a.cdim[2] + ci];
        }
      }
      for (int pp = P - 1; pp >= 0; pp--) {
        __syncthreads();
        for (uint32_t id = tid; id < cnt; id += HIER_THREADS) {
          const uint32_t i = id % ex, j = (id / ex) % ey, kk = id / (ex * ey);
          inv_node<T, HALF, TS>(tile, tables, a, x0 + i, y0 + j, z0 + kk, x0, y0, z0,
                                a.L - pp);
        }
      }
      __syncthreads();
      const uint32_t vz = min((uint32_t)TS::TZ, nz - z0),
                     vy = min((uint32_t)TS::TY, ny - y0),
                     vx = min((uint32_t)TS::TX, nx - x0);
      for (uint32_t id = tid; id < vz * vy * vx; id += HIER_THREADS) {
        const uint32_t i = id % vx, j = (id / vx) % vy, kk = id / (vx * vy);
        out[(uint64_t)(z0 + kk) * sz + (uint64_t)(y0 + j) * sy + x0 + i] =
            tile[(kk * TS::HY + j) * TS::PX + i];
      }
    }
    cur = nxt;
    cx = ncx;
    cy = ncy;
    cz = ncz;
  }
}

// symbols of the coarse box [0, n_{L-P})^3 -> compact array of coefficients
template <typename T>
__global__ void __launch_bounds__(256)
    k_coarse_gather(T *__restrict__ dst, const __grid_constant__ HierArgs<T> a) {
  const uint64_t total = (uint64_t)a.cdim[0] * a.cdim[1] * a.cdim[2];
  const uint64_t sy = a.n[2], sz = (uint64_t)a.n[2] * a.n[1];
  for (uint64_t i = blockIdx.x * 256ull + threadIdx.x; i < total;
       i += (uint64_t)gridDim.x * 256) {
    const uint32_t x = (uint32_t)(i % a.cdim[2]);
    const uint32_t y = (uint32_t)((i / a.cdim[2]) % a.cdim[1]);
    const uint32_t z = (uint32_t)(i / ((uint64_t)a.cdim[2] * a.cdim[1]));
    const uint16_t *at = a.sym + (z * sz + y * sy + x);
    dst[i] = coef_of<T>(a, (uint32_t)__ldg(at), at);
  }
}

// ===================== 3-D inverse, one level per launch =======================
// The tile kernel above keeps three levels in shared memory and pays for it with
// four CTA barriers per tile, a 2.4x redundant stride-2 pass on the tile halo and
// ~60 warp-instructions per 32 nodes.  For 3-D fields the levels are run one launch
// each instead: level l reads the COMPACT nodal array of level l-1 (1/8 of the
// nodes), the 7/8 of level l's symbols that sit in the seven odd blocks of the
// reordered layout, and writes the compact nodal array of level l (the caller's
// buffer for l = L).  No shared memory, no barriers, no halo: 6.25 B per fp32 node
// at the finest level instead of 6, +1/7 of that for all coarser levels together.
//
// In level-l index space the level-(l-1) nodes are the even indices plus the last
// one.  Block b of an axis = coarse node b (level-l index
// 2b, or n_l - 1 for the last) and the odd node 2b+1 when 2b+1 < n_l - 1, whose
// symbol sits at reordered position cn + b.
template <typename T> struct InvLevel {
  const T *coarse; // level l-1: cn[0] x cn[1] x cn[2]
  T *fine;         // level l:   fn[0] x fn[1] x fn[2]
  uint32_t cn[3], fn[3];
  uint32_t roff[3];       // ratio tables of level l per axis
  uint64_t oX, oY, oZ;    // element offsets of the odd x / y / z blocks in sym
};

template <typename T> struct LvlPlane {
  T e0, f0, gE, gO;
};

// x and y lerps of one z plane of a full block: c -> coarse node (bk, bj, bi)
template <typename T, bool HALF>
__device__ __forceinline__ void lvl_plane(const T *__restrict__ c, uint64_t csy, T tx,
                                          T ty, LvlPlane<T> &Q) {
  const T e0 = __ldg(c), r0 = __ldg(c + 1), e2 = __ldg(c + csy),
          r2 = __ldg(c + csy + 1);
  Q.e0 = e0;
  Q.f0 = mgx_lerp(e0, r0, tx);
  const T f2 = mgx_lerp(e2, r2, tx);
  Q.gE = mgx_lerp(e0, e2, ty);
  Q.gO = mgx_lerp(Q.f0, f2, ty);
}

#ifndef MGX_IL_THREADS
#define MGX_IL_THREADS 128
#endif
#ifndef MGX_IL_MINB
#define MGX_IL_MINB (1024 / MGX_IL_THREADS)
#endif
constexpr int IL_THREADS = MGX_IL_THREADS; // 32 x-blocks x 4 y-blocks x IL_ZW z-pairs
constexpr int IL_ZW = IL_THREADS / 128;

template <typename T, bool HALF>
__global__ void __launch_bounds__(IL_THREADS, MGX_IL_MINB)
    k_inv_level3(const __grid_constant__ HierArgs<T> a,
                 const __grid_constant__ InvLevel<T> v, const T *__restrict__ tables) {
  const uint32_t lane = threadIdx.x & 31, wy = (threadIdx.x >> 5) & 3,
                 wz = threadIdx.x >> 7;
  const uint32_t bi = blockIdx.x * 32 + lane, bj = blockIdx.y * 4 + wy,
                 bk0 = blockIdx.z * (2 * IL_ZW) + wz * 2;
  const uint32_t cnz = v.cn[0], cny = v.cn[1], cnx = v.cn[2];
  const uint32_t fnz = v.fn[0], fny = v.fn[1], fnx = v.fn[2];
  if (bi >= cnx || bj >= cny || bk0 >= cnz)
    return;
  const uint64_t sy = a.n[2], sz = (uint64_t)a.n[2] * a.n[1];
  const uint64_t csy = cnx, csz = (uint64_t)cnx * cny;
  const uint64_t fsy = fnx, fsz = (uint64_t)fnx * fny;
  const bool ex = 2 * bi + 1 < fnx - 1, ey = 2 * bj + 1 < fny - 1;
  const bool full = ex && ey && bk0 + 1 < cnz && 2 * (bk0 + 1) + 1 < fnz - 1;
  const bool slow_sym = a.n_out != 0;
  if (full) {
    // two full blocks stacked along z: three coarse planes, 14 symbols, 16 nodes
    T tx = (T)0.5, ty = (T)0.5;
    if (!HALF) {
      tx = tables[v.roff[2] + 2 * bi];
      ty = tables[v.roff[1] + 2 * bj];
    }
    const T *c = v.coarse + ((uint64_t)bk0 * csz + (uint64_t)bj * csy + bi);
    const uint16_t *ob = a.sym + ((uint64_t)bk0 * sz + (uint64_t)bj * sy + bi);
    T *g = v.fine + ((uint64_t)(2 * bk0) * fsz + (uint64_t)(2 * bj) * fsy + 2 * bi);
    uint32_t sv[2][7];
#pragma unroll
    for (int b = 0; b < 2; b++) {
      const uint16_t *o = ob + (uint64_t)b * sz;
      sv[b][0] = __ldg(o + v.oX);
      sv[b][1] = __ldg(o + v.oY);
      sv[b][2] = __ldg(o + v.oX + v.oY);
      sv[b][3] = __ldg(o + v.oZ);
      sv[b][4] = __ldg(o + v.oZ + v.oX);
      sv[b][5] = __ldg(o + v.oZ + v.oY);
      sv[b][6] = __ldg(o + v.oZ + v.oX + v.oY);
    }
    LvlPlane<T> A, B;
    lvl_plane<T, HALF>(c, csy, tx, ty, A);
#pragma unroll
    for (int b = 0; b < 2; b++) {
      lvl_plane<T, HALF>(c + (uint64_t)(b + 1) * csz, csy, tx, ty, B);
      T tz = (T)0.5;
      if (!HALF)
        tz = tables[v.roff[0] + 2 * (bk0 + b)];
      T cf[7];
      uint32_t mn = 0xffffffffu;
#pragma unroll
      for (int i = 0; i < 7; i++) {
        cf[i] = mul_rn(a.dq, (T)((int)sv[b][i] - a.shift));
        mn = min(mn, sv[b][i]);
      }
      if (__builtin_expect(slow_sym && mn == 0, 0)) {
        const uint16_t *o = ob + (uint64_t)b * sz;
        const uint64_t off[7] = {v.oX, v.oY, v.oX + v.oY, v.oZ, v.oZ + v.oX,
                                 v.oZ + v.oY, v.oZ + v.oX + v.oY};
#pragma unroll
        for (int i = 0; i < 7; i++)
          if (sv[b][i] == 0)
            cf[i] = coef_of<T>(a, 0u, o + off[i]);
      }
      T *gb = g + (uint64_t)(2 * b) * fsz;
      gb[0] = A.e0;
      gb[1] = add_rn(cf[0], A.f0);
      gb[fsy] = add_rn(cf[1], A.gE);
      gb[fsy + 1] = add_rn(cf[2], A.gO);
      gb[fsz] = add_rn(cf[3], mgx_lerp(A.e0, B.e0, tz));
      gb[fsz + 1] = add_rn(cf[4], mgx_lerp(A.f0, B.f0, tz));
      gb[fsz + fsy] = add_rn(cf[5], mgx_lerp(A.gE, B.gE, tz));
      gb[fsz + fsy + 1] = add_rn(cf[6], mgx_lerp(A.gO, B.gO, tz));
      A = B;
    }
    return;
  }
  // blocks on the last node of an axis: missing odd halves, last coarse node at n_l - 1
  for (uint32_t bk = bk0; bk < min(bk0 + 2, cnz); bk++) {
    const bool ez = 2 * bk + 1 < fnz - 1;
    const uint32_t fx = bi == cnx - 1 ? fnx - 1 : 2 * bi,
                   fy = bj == cny - 1 ? fny - 1 : 2 * bj,
                   fz = bk == cnz - 1 ? fnz - 1 : 2 * bk;
    T tx = (T)0.5, ty = (T)0.5, tz = (T)0.5;
    if (!HALF) {
      if (ex) tx = tables[v.roff[2] + 2 * bi];
      if (ey) ty = tables[v.roff[1] + 2 * bj];
      if (ez) tz = tables[v.roff[0] + 2 * bk];
    }
    const T *c = v.coarse + ((uint64_t)bk * csz + (uint64_t)bj * csy + bi);
    const uint64_t dx = ex ? 1 : 0, dy = ey ? csy : 0, dz = ez ? csz : 0;
    auto plane = [&](const T *r, LvlPlane<T> &Q) {
      const T e0 = r[0], r0 = r[dx], e2 = r[dy], r2 = r[dy + dx];
      Q.e0 = e0;
      Q.f0 = mgx_lerp(e0, r0, tx);
      const T f2 = mgx_lerp(e2, r2, tx);
      Q.gE = mgx_lerp(e0, e2, ty);
      Q.gO = mgx_lerp(Q.f0, f2, ty);
    };
    LvlPlane<T> A, B;
    plane(c, A);
    const uint16_t *ob = a.sym + ((uint64_t)bk * sz + (uint64_t)bj * sy + bi);
    auto cf = [&](uint64_t off) {
      const uint16_t *at = ob + off;
      return coef_of<T>(a, (uint32_t)__ldg(at), at);
    };
    T *gb = v.fine + ((uint64_t)fz * fsz + (uint64_t)fy * fsy + fx);
    gb[0] = A.e0;
    if (ex) gb[1] = add_rn(cf(v.oX), A.f0);
    if (ey) gb[fsy] = add_rn(cf(v.oY), A.gE);
    if (ex && ey) gb[fsy + 1] = add_rn(cf(v.oX + v.oY), A.gO);
    if (ez) {
      plane(c + dz, B);
      gb[fsz] = add_rn(cf(v.oZ), mgx_lerp(A.e0, B.e0, tz));
      if (ex) gb[fsz + 1] = add_rn(cf(v.oZ + v.oX), mgx_lerp(A.f0, B.f0, tz));
      if (ey) gb[fsz + fsy] = add_rn(cf(v.oZ + v.oY), mgx_lerp(A.gE, B.gE, tz));
      if (ex && ey)
        gb[fsz + fsy + 1] = add_rn(cf(v.oZ + v.oX + v.oY), mgx_lerp(A.gO, B.gO, tz));
    }
  }
}

typedef TileShape<3, 3, 6> Tile3D;  // 8 x 8 x 64
typedef TileShape<0, 5, 7> Tile2D;  // 32 x 128
typedef TileShape<0, 0, 10> Tile1D; // 1024

template <typename T, bool HALF, class TS>
int launch_forward(mgx_ctx *ctx, const void *d_in, HierArgs<T> &a) {
  const int tdim[3] = {TS::TZ, TS::TY, TS::TX};
  for (int d = 0; d < 3; d++)
    a.tiles[d] = (a.n[d] + tdim[d] - 1) / tdim[d];
  a.ntiles = (uint64_t)a.tiles[0] * a.tiles[1] * a.tiles[2];
  size_t smem = 2 * sizeof(T) * TS::tile_elems + (size_t)ctx->dict * 4;
  auto kern = k_hier_forward<T, HALF, TS>;
  MGX_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                (int)smem));
  int occ = 1;
  MGX_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, HIER_THREADS,
                                                         smem));
  unsigned grid = (unsigned)std::min<uint64_t>(
      a.ntiles, (uint64_t)std::max(occ, 1) * ctx->num_sms);
  for (int p = 0; p < 12; p++) {
    const int l = a.L - p;
    const uint64_t sy = a.n[2], sz = (uint64_t)a.n[2] * a.n[1];
    for (int d = 0; d < 3; d++)
      a.oct[p][d] = 0;
    if (l >= 1) {
      a.oct[p][0] = 2ull * a.nl[l - 1][2];
      a.oct[p][1] = 2ull * a.nl[l - 1][1] * sy;
      a.oct[p][2] = 2ull * a.nl[l - 1][0] * sz;
    }
  }
  a.sym = ctx->d_sym;
  a.out_count = (unsigned long long *)(ctx->d_scalars + MGX_SC_OUTLIERS);
  a.out_idx = (unsigned long long *)ctx->d_out_idx;
  a.out_val = (long long *)ctx->d_out_val;
  a.out_cap = ctx->out_cap;
  kern<<<grid, HIER_THREADS, smem, ctx->stream>>>(
      (const T *)d_in, ctx->d_freq, (const T *)ctx->d_tables, a);
  ctx->launches++;
  if (a.L >= TS::PASSES) {
    constexpr int PER_TILE = (TS::TX >> TS::PASSES) *
                             (TS::RY ? (TS::TY >> TS::PASSES) : 1) *
                             (TS::RZ ? (TS::TZ >> TS::PASSES) : 1);
    auto deep = k_hier_deep<T, HALF, TS>;
    const size_t hsmem = (size_t)ctx->dict * 4;
    MGX_CUDA(cudaFuncSetAttribute(deep, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  (int)hsmem));
    const uint64_t work = a.ntiles * PER_TILE;
    unsigned dgrid = mgx_grid_for(work, HIER_THREADS * 4, ctx->num_sms * 4);
    deep<<<dgrid, HIER_THREADS, hsmem, ctx->stream>>>(
        (const T *)d_in, ctx->d_freq, (const T *)ctx->d_tables, a);
    ctx->launches++;
  }
  MGX_CUDA(cudaGetLastError());
  return MGX_SUCCESS;
}

template <typename T, bool HALF>
int launch_forward_shape(mgx_ctx *ctx, const void *d_in, HierArgs<T> &a) {
  if (a.n[0] > 1)
    return launch_forward<T, HALF, Tile3D>(ctx, d_in, a);
  if (a.n[1] > 1)
    return launch_forward<T, HALF, Tile2D>(ctx, d_in, a);
  return launch_forward<T, HALF, Tile1D>(ctx, d_in, a);
}

template <typename T>
void fill_hier_args(mgx_ctx *ctx, HierArgs<T> &a, int rel, double tol_d,
                    double norm_d) {
  const MgxLevels &lv = ctx->lv;
  const MgxHostTables<T> &tb = *reinterpret_cast<const MgxHostTables<T> *>(
      mgx_is_f32(ctx) ? (const void *)&ctx->tbf : (const void *)&ctx->tbd);
  for (int d = 0; d < 3; d++)
    a.n[d] = (uint32_t)lv.shape[2 + d];
  for (int l = 0; l <= lv.L; l++)
    for (int d = 0; d < 3; d++) {
      a.nl[l][d] = (uint32_t)lv.n[l][2 + d];
      a.roff[l][d] = tb.off.ratio[l][2 + d];
    }
  a.L = lv.L;
  // CalcQuantizers, LinearQuantization.hpp:234-261 (hierarchical branch)
  T tol = (T)tol_d, norm = (T)norm_d;
  double abs_tol = tol;
  if (rel)
    abs_tol *= norm;
  abs_tol *= 2;
  T qv = (T)((abs_tol) / (uint64_t)(lv.L + 1));
  a.dq = qv;
  a.qv = 1.0f / qv;
  a.shift = (int)(ctx->dict / 2);
  a.dict = ctx->dict;
  a.coarse = nullptr;
  a.n_out = 0;
  a.cdim[0] = a.cdim[1] = a.cdim[2] = 1;
}

template <typename T>
int hier_forward_t(mgx_ctx *ctx, const void *d_in, int rel, double tol_d,
                   double norm_d) {
  const MgxHostTables<T> &tb = *reinterpret_cast<const MgxHostTables<T> *>(
      mgx_is_f32(ctx) ? (const void *)&ctx->tbf : (const void *)&ctx->tbd);
  HierArgs<T> a;
  fill_hier_args<T>(ctx, a, rel, tol_d, norm_d);
  if (a.n[0] > 1) {
    bool used = false;
    const int rc = mgxk_hier_forward3<T>(ctx, d_in, &a, tb.ratios_half, &used);
    if (rc || used)
      return rc;
  }
  return tb.ratios_half ? launch_forward_shape<T, true>(ctx, d_in, a)
                        : launch_forward_shape<T, false>(ctx, d_in, a);
}

template <typename T, bool HALF, class TS>
int launch_inverse(mgx_ctx *ctx, void *d_out, HierArgs<T> &a) {
  constexpr int P = TS::PASSES;
  if (a.L < P)
    return MGX_FAILURE + 101;
  const int tdim[3] = {TS::TZ, TS::TY, TS::TX};
  for (int d = 0; d < 3; d++)
    a.tiles[d] = (a.n[d] + tdim[d] - 1) / tdim[d];
  a.ntiles = (uint64_t)a.tiles[0] * a.tiles[1] * a.tiles[2];
  for (int p = 0; p < 12; p++) {
    const int l = a.L - p;
    const uint64_t sy = a.n[2], sz = (uint64_t)a.n[2] * a.n[1];
    for (int d = 0; d < 3; d++)
      a.oct[p][d] = 0;
    if (l >= 1) {
      a.oct[p][0] = 2ull * a.nl[l - 1][2];
      a.oct[p][1] = 2ull * a.nl[l - 1][1] * sy;
      a.oct[p][2] = 2ull * a.nl[l - 1][0] * sz;
    }
  }
  // coarse lattice = level-(L-P) grid: gather + dequantise its box, recompose
  // it with the level kernels on a compact array
  const int lc = a.L - P;
  uint64_t celems = 1;
  for (int d = 0; d < 3; d++) {
    a.cdim[d] = a.nl[lc][d];
    celems *= a.cdim[d];
  }
  const size_t need = celems * sizeof(T);
  if (ctx->coarse_bytes < need) {
    if (ctx->d_coarse0)
      cudaFree(ctx->d_coarse0);
    if (ctx->d_coarse1)
      cudaFree(ctx->d_coarse1);
    MGX_CUDA(cudaMalloc(&ctx->d_coarse0, need));
    MGX_CUDA(cudaMalloc(&ctx->d_coarse1, need));
    ctx->coarse_bytes = need;
  }
  k_coarse_gather<T><<<mgx_grid_for(celems, 256, ctx->num_sms * 8), 256, 0,
                       ctx->stream>>>((T *)ctx->d_coarse0, a);
  ctx->launches++;
  int rc = mgxk_recompose_compact(ctx, ctx->d_coarse0, ctx->d_coarse1, lc);
  if (rc)
    return rc;
  a.coarse = (const T *)ctx->d_coarse0;
  const size_t smem = sizeof(T) * TS::tile_elems;
  void (*kern)(T *, const T *, const HierArgs<T>) = k_hier_inverse<T, HALF, TS>;
  if (TS::RZ)
    kern = k_hier_inverse3d<T, HALF>;
  MGX_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                (int)smem));
  int occ = 1;
  MGX_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, HIER_THREADS,
                                                         smem));
  unsigned grid = (unsigned)std::min<uint64_t>(
      a.ntiles, (uint64_t)std::max(occ, 1) * ctx->num_sms);
  kern<<<grid, HIER_THREADS, smem, ctx->stream>>>((T *)d_out,
                                                  (const T *)ctx->d_tables, a);
  ctx->launches++;
  MGX_CUDA(cudaGetLastError());
  return MGX_SUCCESS;
}

// 3-D: level 0 gathered + dequantised from its box, then one k_inv_level3 per level
// through two ping-pong compact arrays (level L-1 and level L-2 sized)
template <typename T, bool HALF>
int launch_inverse_levels(mgx_ctx *ctx, void *d_out, HierArgs<T> &a) {
  const int L = a.L;
  auto elems = [&](int l) {
    return (uint64_t)a.nl[l][0] * a.nl[l][1] * a.nl[l][2];
  };
  const size_t need[2] = {L >= 2 ? elems(L - 2) * sizeof(T) : 0,
                          L >= 1 ? elems(L - 1) * sizeof(T) : 0};
  for (int i = 0; i < 2; i++)
    if (ctx->lvl_bytes[i] < need[i]) {
      if (ctx->d_lvl[i])
        cudaFree(ctx->d_lvl[i]);
      ctx->d_lvl[i] = nullptr;
      ctx->lvl_bytes[i] = 0;
      MGX_CUDA(cudaMalloc(&ctx->d_lvl[i], need[i]));
      ctx->lvl_bytes[i] = need[i];
    }
  auto buf = [&](int l) { return l == L ? (T *)d_out : (T *)ctx->d_lvl[(L - l) & 1]; };
  uint64_t celems = 1;
  for (int d = 0; d < 3; d++) {
    a.cdim[d] = a.nl[0][d];
    celems *= a.cdim[d];
  }
  k_coarse_gather<T><<<mgx_grid_for(celems, 256, ctx->num_sms * 8), 256, 0,
                       ctx->stream>>>(buf(0), a);
  ctx->launches++;
  const uint64_t sy = a.n[2], sz = (uint64_t)a.n[2] * a.n[1];
  for (int l = 1; l <= L; l++) {
    InvLevel<T> v;
    v.coarse = buf(l - 1);
    v.fine = buf(l);
    for (int d = 0; d < 3; d++) {
      v.cn[d] = a.nl[l - 1][d];
      v.fn[d] = a.nl[l][d];
      v.roff[d] = a.roff[l][d];
    }
    v.oX = v.cn[2];
    v.oY = (uint64_t)v.cn[1] * sy;
    v.oZ = (uint64_t)v.cn[0] * sz;
    const dim3 grid((v.cn[2] + 31) / 32, (v.cn[1] + 3) / 4,
                    (v.cn[0] + 2 * IL_ZW - 1) / (2 * IL_ZW));
    if (grid.y > 65535u || grid.z > 65535u)
      return MGX_FAILURE + 102;
    k_inv_level3<T, HALF><<<grid, IL_THREADS, 0, ctx->stream>>>(
        a, v, (const T *)ctx->d_tables);
    ctx->launches++;
  }
  MGX_CUDA(cudaGetLastError());
  return MGX_SUCCESS;
}

template <typename T>
int hier_inverse_t(mgx_ctx *ctx, int rel, double tol_d, double norm_d, void *d_out,
                   const uint64_t *d_oidx, const int64_t *d_oval, uint64_t n_out) {
  const MgxHostTables<T> &tb = *reinterpret_cast<const MgxHostTables<T> *>(
      mgx_is_f32(ctx) ? (const void *)&ctx->tbf : (const void *)&ctx->tbd);
  HierArgs<T> a;
  fill_hier_args<T>(ctx, a, rel, tol_d, norm_d);
  a.sym = ctx->d_sym;
  a.out_idx = (unsigned long long *)d_oidx;
  a.out_val = (long long *)d_oval;
  a.n_out = n_out;
  a.out_count = nullptr;
  a.out_cap = 0;
  if (a.n[0] > 1 && !ctx->inv_tiled)
    return tb.ratios_half ? launch_inverse_levels<T, true>(ctx, d_out, a)
                          : launch_inverse_levels<T, false>(ctx, d_out, a);
  if (a.n[0] > 1)
    return tb.ratios_half ? launch_inverse<T, true, Tile3D>(ctx, d_out, a)
                          : launch_inverse<T, false, Tile3D>(ctx, d_out, a);
  if (a.n[1] > 1)
    return tb.ratios_half ? launch_inverse<T, true, Tile2D>(ctx, d_out, a)
                          : launch_inverse<T, false, Tile2D>(ctx, d_out, a);
  return tb.ratios_half ? launch_inverse<T, true, Tile1D>(ctx, d_out, a)
                        : launch_inverse<T, false, Tile1D>(ctx, d_out, a);
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

int mgxk_hier_inverse(mgx_ctx *ctx, int rel, double tol, double norm, void *d_out,
                      const uint64_t *d_oidx, const int64_t *d_oval,
                      uint64_t n_out) {
  if (ctx->lv.D > 3) {
    ctx->err = "hierarchical fast path needs D <= 3";
    return MGX_FAILURE;
  }
  return mgx_is_f32(ctx)
             ? hier_inverse_t<float>(ctx, rel, tol, norm, d_out, d_oidx, d_oval, n_out)
             : hier_inverse_t<double>(ctx, rel, tol, norm, d_out, d_oidx, d_oval,
                                      n_out);
}
