// 3-D fused hierarchical forward pass, second generation: TMA-staged tiles and
// warp-specialised roles (replaces, for 3-D fields, the LDGSTS-staged
// k_hier_forward of mgx_hier.cu; same outputs bit for bit).
//
// What it computes is unchanged -- every level's coefficient of the hierarchical
// basis from the ORIGINAL nodal values (reference: MultiDimension/
// DataRefactoring.hpp:77-114, Coefficient/GridProcessingKernel3D.hpp:20-1228),
// lwqzk quantisation (Quantization/LinearQuantization.hpp:19-145), outlier
// separation (ParallelHuffman/OutlierSeparator.hpp:17-120) and the histogram
// (ParallelHuffman/Histogram.hpp:15-235) in ONE pass: 4 B read + 2 B written per
// fp32 node.  How it moves and schedules the data is new:
//
// * tiles of 8 x 8 x 64 nodes (+1 halo plane per axis) are fetched by the TMA
//   unit.  Rows of a 2^k+1 grid are only element-aligned, which rules out a
//   tensor map of the array itself (strides must be multiples of 16 bytes), so
//   the array is DESCRIBED as a 2-D tensor of "super rows" of four consecutive
//   rows (4 * nx elements; 16 * nx bytes is always a multiple of 16): row R of
//   the flattened (z, y) index lives at x-offset (R mod 4) * nx of super row
//   R div 4, and the rows R, R + 4, R + 8 of a tile plane are one box.  A box
//   must START on a 16-byte boundary of the tensor's inner dimension (measured
//   on B200: any other start coordinate raises "illegal instruction"), so each
//   box fetches the 16-byte-aligned superset of its row segments and the wanted
//   node x0 lands `shift` elements into the staged row.  The shift of row m of
//   staged plane p is ((p * ny + m) * nx) mod E (E = elements per 16 bytes; tile
//   origins are multiples of 8): it does not depend on the tile, and starting the
//   box A(p) + B(m) elements early (A(p) = p*ny*nx mod E, B(m) = m*nx mod E, no
//   wrap) makes the staged address of a node the SUM of a plane term and a row
//   term.  36 UTMALDG per tile replace 5 300 per-element LDGSTS.
// * one CTA per SM, no CTA barrier in steady state: NP producer warps issue the
//   boxes into a ring of S tile buffers (mbarrier full/empty pairs); G groups of
//   8 warps run the stride-1 pass of one tile each; C further warps run the
//   stride-2 and stride-4 passes of the same tiles.  Tiles are read-only for all
//   roles, so a tile needs no synchronisation between its passes.
// * the CTAs draw their tiles from a device-wide counter, several consecutive
//   tiles per draw (see the producer loop); outliers are staged per warp in shared
//   memory and leave for the global list with one atomic per flush (Sink3).
// * with ratio 0.5 (uniform grids) lerp(a, b) = (a - RN(a/2)) + RN(b/2): the two
//   halves of every node are formed once and shared by all the interpolations
//   that use it -- same roundings as the reference's four-operation lerp
//   (GPKFunctor.h:13-25), 28 instead of 39 flops per 2 x 2 x 2 block.
// * a block whose seven coefficients all quantise to zero (|c| <= cmax, the
//   exact threshold found on the host) writes the symbol dict/2 seven times and
//   skips conversion, range test and histogram; the bin of dict/2 is never
//   counted in the kernels and is restored afterwards as N - sum(other bins).
// * the tile order runs through super blocks of (all x) x 8 x 8 tiles so that the
//   halo planes shared by neighbouring tiles are still in L2 when they are re-read.
//
// Nodes outside the structured passes (multiples of 8 on every axis, and the
// nodes of the clipped tiles that touch the last node of an axis) go through
// k_fwd3_rest, a per-node gather from global memory / L2.
#include <cuda.h>

#include <cstdlib>
#include <cstring>

#include "mgx_hier_common.cuh"

namespace {

// ------------------------------------------------------------------ geometry
template <typename T> struct F3 {
  static constexpr int TZ = 8, TY = 8, TX = 64;
  static constexpr int E = 16 / (int)sizeof(T);       // elements per 16 bytes
  // 65 nodes + up to 2 (E - 1) elements of alignment slack, a multiple of E
  static constexpr int BX = (TX + 1 + 2 * (E - 1) + E - 1) / E * E; // float 72, double 68
  static constexpr int RB = BX * (int)sizeof(T);      // bytes per staged row
  static constexpr int BOX3 = (3 * RB + 127) / 128 * 128; // rows m, m+4, m+8 (m = 0)
  static constexpr int BOX2 = (2 * RB + 127) / 128 * 128; // rows m, m+4     (m = 1, 2, 3)
  static constexpr int PLANE = BOX3 + 3 * BOX2;
  static constexpr int TILE = 9 * PLANE;
  static constexpr uint32_t TX_BYTES = 81u * RB; // bytes the 36 boxes of a tile deliver
  // byte offset of row m (0..8) inside a staged plane
  __host__ __device__ static constexpr int rowoff(int m) {
    return (m & 3) == 0 ? (m >> 2) * RB : BOX3 + ((m & 3) - 1) * BOX2 + (m >> 2) * RB;
  }
};

constexpr int F3_GW = 8; // stride-1 warps per group (one tile per group at a time)
#ifndef F3_NP_
#define F3_NP_ 8
#endif
#ifndef F3_G_
#define F3_G_ 2
#endif
#ifndef F3_C_
#define F3_C_ 4
#endif
#ifndef F3_NP64_
#define F3_NP64_ 5
#endif
constexpr int F3_SMEM_MAX = 227 * 1024;
constexpr int F3_MAX_DICT = 8192;

template <typename T> constexpr int f3_stages() {
  return (F3_SMEM_MAX - F3_MAX_DICT * 4 - 1024) / F3<T>::TILE > 8
             ? 8
             : (F3_SMEM_MAX - F3_MAX_DICT * 4 - 1024) / F3<T>::TILE;
}
// roles per CTA: NP producer warps (36 / NP of a tile's 36 boxes each), G groups of 8
// stride-1 warps, C warps for the stride-2 / stride-4 passes (about 1.6 warp-tiles of
// work per tile against 8 for the stride-1 pass).  fp32: (8, 2, 4) = 28 warps.  With
// the tiles drawn from a counter the CTA waits for its boxes, not for its consumers:
// 8 producer warps deliver a tile in 0.83 us where 4 needed 1.0 (loads alone), and 896
// threads get 72 registers each where 1024 got 64 -- at 29 warps and more the smooth
// field's long path is 25 % slower.  fp64 values take two registers: fewer threads,
// more each -- (5, 2, 3) = 24 warps with 80 registers (641^3: 0.69 ms with 4 producer
// warps, 0.66 with 5, 0.69 with 6 or 8).
template <typename T> struct F3Cfg {
  static constexpr int S = f3_stages<T>();
  static constexpr int NP = sizeof(T) == 4 ? F3_NP_ : F3_NP64_;
  static constexpr int G = sizeof(T) == 4 ? F3_G_ : 2;
  static constexpr int C = sizeof(T) == 4 ? F3_C_ : 3;
  static constexpr int WARPS = NP + G * F3_GW + C;
  static constexpr int THREADS = 32 * WARPS;
  static_assert(G <= S && C <= S, "end markers need a stage each");
};

// per-stage tile descriptor written by the producer
struct F3Info {
  unsigned long long base[3]; // address of the tile's first symbol of pass P (block 0)
  uint32_t x0, y0, z0;
  uint32_t seq; // index of the tile in this CTA's sequence (see full_wait)
};

template <typename T> struct F3Args {
  T cmax;               // |c| <= cmax  <=>  the coefficient quantises to code 0
  uint32_t sh_b, sh_c;  // staging shifts: B(m) = m * sh_b mod E, A(p) = p * sh_c mod E
  uint32_t dbg;         // MGX_F3_DBG (measurements only): 1 = no compute, 2 = no loads
  uint32_t rt[3];       // regular tiles per axis (z, y, x)
  uint32_t sby, sbz;    // super blocks (8 x 8 tiles in y, z)
  uint64_t padded;      // rt[2] * 64 * sby * sbz tile slots
  uint32_t step[4];     // gridDim.x in the mixed radix (rt[2], 64, sby, sbz) of a slot index
  uint32_t chunk;       // tiles per draw of the tile counter (power of two)
};

// ------------------------------------------------------------- PTX wrappers
__device__ __forceinline__ uint32_t saddr(const void *p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar),
               "r"(bytes)
               : "memory");
}
// bounded wait: a protocol error (or a lost TMA transaction) must not hang the
// GPU.  On time-out the caller raises a flag in the context's error word and
// leaves its loop; the host then reports MGX_FAILURE.
__device__ __forceinline__ bool mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t done = 0;
  for (uint32_t spin = 0; !done; ++spin) {
    asm volatile("{\n\t.reg .pred p;\n\t"
                 "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
                 "selp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(done)
                 : "r"(bar), "r"(parity)
                 : "memory");
    if (spin > (1u << 22))
      return false;
  }
  return true;
}
// Consumer side of a full barrier.  The stages of the ring are NOT consumed by
// the same warps in successive rounds (tile j goes to group j mod G and stride-2
// warp j mod C, stage j mod S), so a fast consumer can reach the barrier of its
// next tile while the PREVIOUS round of that stage is still loading; a parity
// wait alone would then return at once (it would be reading the completed round
// before that one) and the warp would re-process the stage's old tile -- seen on
// hardware as double-counted histogram bins and extra empty arrivals.  "Previous
// round still loading" and "own round complete" show the same parity, so the
// producer numbers the tiles: once the descriptor of tile j is in place -- which
// implies that the previous round was consumed and the barrier is in tile j's
// phase -- the parity test means "tile j has landed".  The descriptor is written
// when the loads are ISSUED, a microsecond before they land, so polling it with
// a 200 ns back-off adds no latency and next to no instructions.
__device__ __forceinline__ bool full_wait(uint32_t bar, uint32_t parity,
                                          const volatile uint32_t *seq, uint32_t j) {
  for (uint32_t spin = 0; *seq != j; ++spin) {
    if (spin > (1u << 20))
      return false;
    __nanosleep(200);
  }
  return mbar_wait(bar, parity);
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap *map,
                                            uint32_t bar, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes"
               " [%0], [%1, {%3, %4}], [%2];" ::"r"(dst),
               "l"(map), "r"(bar), "r"(c0), "r"(c1)
               : "memory");
}

// symbol store: the address is assembled from integers, so say "global" explicitly
// (a generic ST would do the address-space check on every store)
__device__ __forceinline__ void stg_u16(uint16_t *p, uint32_t v) {
  asm volatile("st.global.u16 [%0], %1;" ::"l"(p), "h"((uint16_t)v) : "memory");
}

// ---------------------------------------------------------------- quantiser
// Sink of the TMA kernels: like Sink<T> of mgx_hier_common.cuh, but the modal
// symbol dict/2 is never counted (restored by k_hist_fix) and a group of seven
// coefficients that all quantise to zero takes a four-instruction test.
// Outliers are staged per consumer warp in shared memory (F3_OCAP entries of 16 bytes,
// slots handed out by a shared-memory atomic) and moved to the global list by the whole
// warp after it has released a tile: one global atomic per flush.  An atomicAdd on the
// global counter per outlier, with the two stores waiting for its result, put a
// device-wide round trip on the consumer warp's critical path for every coefficient
// group that held one: 60 000 outliers of the NYX field at 1e-5 doubled the kernel time.
#ifndef F3_CHUNK_MAX_
#define F3_CHUNK_MAX_ 32
#endif
// tiles per draw of the device-wide tile counter: a power of two <= 32 (one lane decodes
// one tile), about 1/8 of a CTA's share of the tiles (F3Args::chunk)
constexpr uint32_t F3_CHUNK_MAX = F3_CHUNK_MAX_;
constexpr uint32_t F3_CHUNKQ = 16; // ring of published draws (see the producer loop)
static_assert((F3_CHUNK_MAX & (F3_CHUNK_MAX - 1)) == 0 && F3_CHUNK_MAX <= 32, "");
constexpr uint32_t F3_OCAP = 32;
constexpr uint32_t F3_OFLUSH = 12; // flush at a tile boundary from this fill level

template <typename T>
__device__ __noinline__ void f3_outlier_direct(const HierArgs<T> &a, uint64_t opos,
                                               long long q) {
  const unsigned long long slot = atomicAdd(a.out_count, 1ull);
  if (slot < a.out_cap) {
    a.out_idx[slot] = opos;
    a.out_val[slot] = q;
  }
}

template <typename T> struct Sink3 {
  const HierArgs<T> &a;
  T cmax;
  uint32_t hist_saddr;
  uint32_t obuf_saddr; // this warp's F3_OCAP staged outliers (index, value)
  uint32_t ocnt_saddr; // entries handed out since the last flush (may exceed F3_OCAP)

  // symbol of a coefficient whose candidate left [0, dict): 0 and an entry of the
  // outlier list (hier_slow_symbol of mgx_hier_common.cuh; OutlierSeparator.hpp:17-44)
  __device__ __forceinline__ uint32_t outlier(T c, uint64_t opos) const {
    const T r = add_rn((T)0.5, fabs(mul_rn(c, a.qv)));
    long long q = (long long)r;
    q = (c < (T)0 ? -q : q) + (long long)a.shift;
    if (q >= 0 && q < (long long)a.dict)
      return (uint32_t)q;
    uint32_t slot;
    asm volatile("atom.shared.add.u32 %0, [%1], 1;" : "=r"(slot) : "r"(ocnt_saddr) : "memory");
    if (slot < F3_OCAP)
      asm volatile("st.shared.v2.u64 [%0], {%1, %2};" ::"r"(obuf_saddr + slot * 16u),
                   "l"((unsigned long long)opos), "l"(q)
                   : "memory");
    else
      f3_outlier_direct<T>(a, opos, q); // staging area full: the direct way
    return 0;
  }
  // whole warp, converged, after __syncwarp(): move the staged entries to the list
  __device__ __forceinline__ void flush(int lane, uint32_t level) const {
    uint32_t n;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(n) : "r"(ocnt_saddr) : "memory");
    if (n < level)
      return;
    n = min(n, F3_OCAP);
    unsigned long long base = 0;
    if (lane == 0)
      base = atomicAdd(a.out_count, (unsigned long long)n);
    base = __shfl_sync(0xffffffffu, base, 0);
    if ((uint32_t)lane < n) {
      unsigned long long idx;
      long long q;
      asm volatile("ld.shared.v2.u64 {%0, %1}, [%2];"
                   : "=l"(idx), "=l"(q)
                   : "r"(obuf_saddr + (uint32_t)lane * 16u)
                   : "memory");
      if (base + lane < a.out_cap) {
        a.out_idx[base + lane] = idx;
        a.out_val[base + lane] = q;
      }
    }
    __syncwarp();
    if (lane == 0)
      asm volatile("st.shared.u32 [%0], %1;" ::"r"(ocnt_saddr), "r"(0u) : "memory");
    __syncwarp();
  }

  __device__ __forceinline__ uint32_t quant(T c) const {
    const T r = add_rn((T)0.5, fabs(mul_rn(c, a.qv)));
    return (uint32_t)((int)copysign(r, c) + a.shift);
  }
  __device__ __forceinline__ void slow7(const T (&c)[7], uint16_t *const (&dst)[7]) const {
    uint32_t sv[7], worst = 0;
#pragma unroll
    for (int i = 0; i < 7; i++) {
      sv[i] = quant(c[i]);
      worst = max(worst, sv[i]);
    }
    if (__builtin_expect(worst >= a.dict, 0)) {
#pragma unroll
      for (int i = 0; i < 7; i++)
        if (sv[i] >= a.dict)
          sv[i] = outlier(c[i], (uint64_t)(dst[i] - a.sym));
    }
#pragma unroll
    for (int i = 0; i < 7; i++) {
      stg_u16(dst[i], sv[i]);
      if (sv[i] != (uint32_t)a.shift)
        asm volatile("red.shared.add.u32 [%0], 1;" ::"r"(hist_saddr + sv[i] * 4u));
    }
  }
  __device__ __forceinline__ void put7(const T (&c)[7], uint16_t *const (&dst)[7]) const {
    const T m = fmax(fmax(fmax(fabs(c[0]), fabs(c[1])), fmax(fabs(c[2]), fabs(c[3]))),
                     fmax(fmax(fabs(c[4]), fabs(c[5])), fabs(c[6])));
#ifndef MGX_F3_LANE_ZERO_TEST
    // one branch for the whole warp: a lane of zeros next to a lane of codes takes the
    // long path too (slow7 of seven zeros stores the same seven modal symbols), so every
    // symbol store is issued once, by all lanes, and covers whole 32-byte sectors
    const bool zero = __ballot_sync(__activemask(), !(m <= cmax)) == 0;
#else
    const bool zero = m <= cmax; // (false for NaN)
#endif
    if (zero) {
#pragma unroll
      for (int i = 0; i < 7; i++)
        stg_u16(dst[i], (uint32_t)a.shift);
    } else {
      slow7(c, dst);
    }
  }
};

// Sink for the per-node gather of k_fwd3_rest (same conventions)
template <typename T> struct SinkRest {
  const HierArgs<T> &a;
  uint32_t hist_saddr;
  __device__ __forceinline__ void put(T c, uint64_t opos) const {
    const T r = add_rn((T)0.5, fabs(mul_rn(c, a.qv)));
    uint32_t sv = (uint32_t)((int)copysign(r, c) + a.shift);
    if (sv >= a.dict)
      sv = hier_slow_symbol<T>(c, r, a, opos);
    a.sym[opos] = (uint16_t)sv;
    if (sv != (uint32_t)a.shift)
      asm volatile("red.shared.add.u32 [%0], 1;" ::"r"(hist_saddr + sv * 4u));
  }
};

// ------------------------------------------------------------------- passes
// One even plane of the march at stride H = 2^P: from the nodes (e0, r0) of row
// m0 and (e2, r2) of row m2 (x and x + 2H), the interpolants f0 (odd x), gE (odd
// y), gO (odd x and y) -- plus, with ratio 0.5, the two halves of each value
// that the next interpolation needs: h = RN(v/2), c = v - h.
template <typename T, bool HALF> struct Plane3 {
  T f0, gE, gO;       // interpolants of the plane's own three coefficient classes
  T e0;               // !HALF: the even node itself
  T h[4], c[4];       // HALF: halves of (e0, f0, gE, gO)
};

template <typename T, bool HALF, int P>
__device__ __forceinline__ void
plane3(const char *r0p, const char *r2p, Plane3<T, HALF> &Q, T &odd0, T tx, T ty) {
  constexpr int H = 1 << P;
  // (scalar loads: staged rows carry an element shift, pairs are not 8-byte aligned)
  const T e0 = *reinterpret_cast<const T *>(r0p);
  const T o0 = *reinterpret_cast<const T *>(r0p + H * sizeof(T));
  const T r0 = *reinterpret_cast<const T *>(r0p + 2 * H * sizeof(T));
  const T e2 = *reinterpret_cast<const T *>(r2p);
  const T r2 = *reinterpret_cast<const T *>(r2p + 2 * H * sizeof(T));
  odd0 = o0;
  if constexpr (HALF) {
    const T he0 = mul_rn(e0, (T)0.5), ce0 = sub_rn(e0, he0);
    const T he2 = mul_rn(e2, (T)0.5), ce2 = sub_rn(e2, he2);
    const T f0 = add_rn(ce0, mul_rn((T)0.5, r0));
    const T f2 = add_rn(ce2, mul_rn((T)0.5, r2));
    const T gE = add_rn(ce0, he2);
    const T hf0 = mul_rn(f0, (T)0.5), cf0 = sub_rn(f0, hf0);
    const T gO = add_rn(cf0, mul_rn((T)0.5, f2));
    Q.f0 = f0;
    Q.gE = gE;
    Q.gO = gO;
    Q.h[0] = he0;
    Q.c[0] = ce0;
    Q.h[1] = hf0;
    Q.c[1] = cf0;
    Q.h[2] = mul_rn(gE, (T)0.5);
    Q.c[2] = sub_rn(gE, Q.h[2]);
    Q.h[3] = mul_rn(gO, (T)0.5);
    Q.c[3] = sub_rn(gO, Q.h[3]);
  } else {
    Q.e0 = e0;
    Q.f0 = mgx_lerp(e0, r0, tx);
    const T f2 = mgx_lerp(e2, r2, tx);
    Q.gE = mgx_lerp(e0, e2, ty);
    Q.gO = mgx_lerp(Q.f0, f2, ty);
  }
}

template <typename T, bool HALF>
__device__ __forceinline__ T zlerp(const Plane3<T, HALF> &A, const Plane3<T, HALF> &B,
                                   int i, T tz) {
  if constexpr (HALF) {
    return add_rn(A.c[i], B.h[i]);
  } else {
    const T a = i == 0 ? A.e0 : i == 1 ? A.f0 : i == 2 ? A.gE : A.gO;
    const T b = i == 0 ? B.e0 : i == 1 ? B.f0 : i == 2 ? B.gE : B.gO;
    return mgx_lerp(a, b, tz);
  }
}

// March of one thread along z at stride H = 2^P: NB consecutive 2 x 2 x 2 blocks
// of the stride-H lattice whose anchor is (2k, 2ya, 2zb0) * H.  `tile` points at
// the staged tile; ro[j] = row term of row (2ya + j) * H (row_term below).
template <typename T, bool HALF, int P, int NB>
__device__ __forceinline__ void
march3(const char *tile, const T *__restrict__ tables, const HierArgs<T> &a,
       const Sink3<T> &sink, int k, int ya, int zb0, const int (&ro)[3], uint32_t sh_c,
       unsigned long long base, const uint32_t (&tc)[7], uint32_t x0, uint32_t y0,
       uint32_t z0) {
  typedef F3<T> G;
  typedef typename std::conditional<sizeof(T) == 4, float2, double2>::type V2;
  constexpr int H = 1 << P;
  const int l = a.L - P;
  const uint64_t sz2 = 2ull * a.n[2] * a.n[1]; // bytes between z rows of symbols
  T tx = (T)0.5, ty = (T)0.5;
  if constexpr (!HALF) {
    tx = tables[a.roff[l][2] + (x0 >> P) + 2 * k];
    ty = tables[a.roff[l][1] + (y0 >> P) + 2 * ya];
  }
  auto pair = [&](const char *q) -> V2 { // nodes x and x + H of a row
    V2 v;
    v.x = *reinterpret_cast<const T *>(q);
    v.y = *reinterpret_cast<const T *>(q + H * sizeof(T));
    return v;
  };
  // staged plane i (relative to the march's first plane): plane term of the address
  auto pl = [&](int i) -> int {
    const int pp = (2 * zb0 + i) * H;
    return pp * G::PLANE + (int)((pp * sh_c) & (G::E - 1)) * (int)sizeof(T);
  };
  const char *p = tile + (2 * k * H) * (int)sizeof(T);
  Plane3<T, HALF> A, B;
  T oA, oB;
  plane3<T, HALF, P>(p + pl(0) + ro[0], p + pl(0) + ro[2], A, oA, tx, ty);
#pragma unroll
  for (int b = 0; b < NB; b++) {
    const char *q0 = p + pl(2 * b), *q1 = p + pl(2 * b + 1), *q2 = p + pl(2 * b + 2);
    const V2 r1 = pair(q0 + ro[1]);
    plane3<T, HALF, P>(q2 + ro[0], q2 + ro[2], B, oB, tx, ty);
    T tz = (T)0.5;
    if constexpr (!HALF)
      tz = tables[a.roff[l][0] + (z0 >> P) + 2 * (zb0 + b)];
    const V2 s0 = pair(q1 + ro[0]);
    const V2 s1 = pair(q1 + ro[1]);
    const T c[7] = {sub_rn(oA, A.f0),
                    sub_rn(r1.x, A.gE),
                    sub_rn(r1.y, A.gO),
                    sub_rn(s0.x, zlerp<T, HALF>(A, B, 0, tz)),
                    sub_rn(s0.y, zlerp<T, HALF>(A, B, 1, tz)),
                    sub_rn(s1.x, zlerp<T, HALF>(A, B, 2, tz)),
                    sub_rn(s1.y, zlerp<T, HALF>(A, B, 3, tz))};
    const unsigned long long bb = base + (unsigned long long)b * sz2;
    uint16_t *dst[7];
#pragma unroll
    for (int i = 0; i < 7; i++)
      dst[i] = reinterpret_cast<uint16_t *>(bb + 2ull * tc[i]);
    sink.put7(c, dst);
    A = B;
    oA = oB;
  }
}

// row term of a staged address: offset of row m inside a plane + its shift B(m)
template <typename T> __device__ __forceinline__ int row_term(int m, uint32_t sh_b) {
  return F3<T>::rowoff(m) + (int)((m * sh_b) & (F3<T>::E - 1)) * (int)sizeof(T);
}

// element offsets of the 7 coefficient classes of pass P relative to the block's
// first symbol, plus `te` (the thread's own element offset inside the tile)
template <typename T>
__device__ __forceinline__ void class_offsets(const HierArgs<T> &a, int P, uint32_t te,
                                              uint32_t (&tc)[7]) {
  const uint32_t oX = (uint32_t)(a.oct[P][0] >> 1), oY = (uint32_t)(a.oct[P][1] >> 1),
                 oZ = (uint32_t)(a.oct[P][2] >> 1);
  tc[0] = te + oX;
  tc[1] = te + oY;
  tc[2] = te + oX + oY;
  tc[3] = te + oZ;
  tc[4] = te + oZ + oX;
  tc[5] = te + oZ + oY;
  tc[6] = te + oZ + oX + oY;
}

template <typename T, bool HALF>
__global__ void __launch_bounds__(F3Cfg<T>::THREADS, 1)
    k_fwd3(uint32_t *__restrict__ g_freq, const T *__restrict__ tables,
           const __grid_constant__ HierArgs<T> a, const __grid_constant__ F3Args<T> fa,
           const __grid_constant__ CUtensorMap tm3,
           const __grid_constant__ CUtensorMap tm2) {
  typedef F3<T> G;
  constexpr int S = F3Cfg<T>::S, F3_G = F3Cfg<T>::G, F3_C = F3Cfg<T>::C,
                F3_THREADS = F3Cfg<T>::THREADS;
  extern __shared__ __align__(128) unsigned char smem[];
  unsigned char *tiles = smem;
  uint32_t *s_hist = (uint32_t *)(smem + (size_t)S * G::TILE);
  F3Info *info = (F3Info *)(s_hist + a.dict);
  unsigned long long *bars = (unsigned long long *)(info + S);
  // outlier staging of the consumer warps: F3_OCAP 16-byte entries each, then the counts
  constexpr int F3_NP = F3Cfg<T>::NP;
  constexpr int CW = F3Cfg<T>::WARPS - F3_NP;
  unsigned char *obuf = (unsigned char *)(((uintptr_t)(bars + 2 * S) + 15) & ~(uintptr_t)15);
  uint32_t *ocnt = (uint32_t *)(obuf + (size_t)CW * F3_OCAP * 16);
  unsigned long long *chunkq =
      (unsigned long long *)(((uintptr_t)(ocnt + CW) + 7) & ~(uintptr_t)7);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if (tid < CW)
    ocnt[tid] = 0;
  if (tid < F3_CHUNKQ)
    chunkq[tid] = ~0ull;
  const uint32_t full0 = saddr(bars), empty0 = saddr(bars + S);
  for (uint32_t i = tid; i < a.dict; i += F3_THREADS)
    s_hist[i] = 0;
  if (tid == 0) {
    for (int s = 0; s < S; s++) {
      info[s].seq = 0xffffffffu;
      mbar_init(full0 + 8 * s, 1);
      mbar_init(empty0 + 8 * s, F3_GW + 1);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  const uint32_t ny = a.n[1], nx = a.n[2];
  const uint64_t sy = nx, sz = (uint64_t)nx * ny;

  unsigned long long *const err_word = a.out_count + (MGX_SC_ERR - MGX_SC_OUTLIERS);
  if (warp < F3_NP) {
    // ---- producers: F3_NP warps walk the same tile list; warp w issues 36 / F3_NP of
    // the tile's 36 (plane, row class) boxes, one box per lane (a single thread issuing all 36 boxes
    // of a tile was measured to be the bottleneck of the whole kernel: ~120
    // cycles per UTMALDG with its operand set-up); warp 0 also arms the barrier
    // and writes the tile descriptor
    const uint32_t T2 = fa.rt[2];
    int s = 0;
    uint32_t ph = 0, seq = 0;
    bool ok = true;
    auto next_stage = [&]() {
      ++seq;
      if (++s == S) {
        s = 0;
        ph ^= 1;
      }
    };
    // this lane's box: plane p (0..8), row class m (0..3)
    constexpr int BPW = (36 + F3_NP - 1) / F3_NP;
#ifdef F3_BOX_BLOCKED
    const int op = min(warp * BPW + lane, 35);
    const bool has_op = lane < BPW && warp * BPW + lane < 36;
#else
    // box o goes to warp o mod NP: 5, 5, 5, 5, 4, 4, 4, 4 boxes for NP = 8
    const int op = min(lane * F3_NP + warp, 35);
    const bool has_op = lane < BPW && lane * F3_NP + warp < 36;
#endif
    const int bp = op >> 2, bm_ = op & 3;
    const int bshift = (int)((bp * fa.sh_c) & (G::E - 1)) + (int)((bm_ * fa.sh_b) & (G::E - 1));
    const uint32_t bdst = (uint32_t)(bp * G::PLANE + G::rowoff(bm_));
    const CUtensorMap *const bmap = bm_ == 0 ? &tm3 : &tm2;
#ifdef MGX_F3_STATIC_TILES
    // slot index of this CTA's next tile as mixed-radix digits (x tile, slot in the
    // 8 x 8 super block, super block y, super block z), advanced by gridDim.x per
    // step without divisions
    const uint32_t SB = T2 * 64;
    uint32_t tx = blockIdx.x % T2, w = (blockIdx.x / T2) % 64, by = (blockIdx.x / SB) % fa.sby,
             bz = (blockIdx.x / SB) / fa.sby;
    for (; bz < fa.sbz && ok; tx += fa.step[0], w += fa.step[1], by += fa.step[2],
                              bz += fa.step[3]) {
      if (tx >= T2) {
        tx -= T2;
        ++w;
      }
      if (w >= 64) {
        w -= 64;
        ++by;
      }
      if (by >= fa.sby) {
        by -= fa.sby;
        ++bz;
      }
      if (bz >= fa.sbz)
        break;
      const uint32_t ty = by * 8 + (w & 7), tz = bz * 8 + (w >> 3);
      if (ty >= fa.rt[1] || tz >= fa.rt[0])
        continue;
#else
    // Tiles are handed out by a device-wide counter, fa.chunk consecutive tiles of the
    // super-block order per draw (x tile fastest, then the <= 8 x 8 (y, z) tiles of a
    // super block, then super blocks along y, then z).  The cost of a tile depends on
    // its data (all-zero groups, outliers) and on the SM's distance to the memory
    // partitions: a fixed tile list per CTA left the first CTAs idle for a quarter of
    // the kernel.  Warp 0 draws one chunk ahead and publishes {chunk number, first
    // tile} through a ring; the other producer warps cannot fall more than
    // S + 1 tiles, i.e. F3_CHUNKQ - 1 draws, behind (warp 0 waits for the stage of tile
    // i - S).  One draw costs the producers about 0.4 us (the decode below): 4 tiles per
    // draw were no faster than the fixed lists, 32 are 11 % faster.
    const uint32_t ntile = (uint32_t)a.ntiles;
    const uint32_t rowsz = fa.rt[1] * T2 * 8; // tiles of a full row of super blocks
    unsigned long long *const tile_ctr = a.out_count + (MGX_SC_TILE - MGX_SC_OUTLIERS);
    uint32_t drawn = 0;
    if (warp == 0 && lane == 0)
      drawn = (uint32_t)atomicAdd(tile_ctr, (unsigned long long)fa.chunk);
    for (uint32_t chunk = 0; ok; chunk++) {
      uint32_t t0;
      if (warp == 0) {
        t0 = __shfl_sync(0xffffffffu, drawn, 0);
        if (lane == 0) {
          chunkq[chunk % F3_CHUNKQ] = ((unsigned long long)chunk << 32) | t0;
          if (t0 < ntile) // next draw: its round trip overlaps this chunk's tiles
            drawn = (uint32_t)atomicAdd(tile_ctr, (unsigned long long)fa.chunk);
        }
      } else {
        unsigned long long e;
        uint32_t spins = 0;
        do {
          e = *(volatile unsigned long long *)&chunkq[chunk % F3_CHUNKQ];
        } while ((uint32_t)(e >> 32) != chunk && ++spins < (1u << 26));
        if ((uint32_t)(e >> 32) != chunk) {
          ok = false;
          if (lane == 0)
            atomicOr(err_word, 0x100ull);
          break;
        }
        t0 = (uint32_t)e;
      }
      if (t0 >= ntile)
        break;
      // lane j decodes tile t0 + j (five divisions: once per chunk, not per tile --
      // the producer warps have about 1000 cycles per tile and spend them on the boxes)
      uint32_t dx, dy, dz;
      {
        const uint32_t v = min(t0 + (lane & (fa.chunk - 1)), ntile - 1);
        const uint32_t bz = v / rowsz, r = v - bz * rowsz;
        const uint32_t hz = min(8u, fa.rt[0] - 8 * bz), fb = 8 * hz * T2;
        const uint32_t by = r / fb, r2 = r - by * fb;
        const uint32_t hy = min(8u, fa.rt[1] - 8 * by);
        const uint32_t q = r2 / T2;
        dx = r2 - q * T2;
        dz = bz * 8 + q / hy;
        dy = by * 8 + q % hy;
      }
      const uint32_t nin = min(fa.chunk, ntile - t0);
      for (uint32_t j = 0; j < nin && ok; j++) {
      const uint32_t tx = __shfl_sync(0xffffffffu, dx, j), ty = __shfl_sync(0xffffffffu, dy, j),
                     tz = __shfl_sync(0xffffffffu, dz, j);
#endif
      const uint32_t x0 = tx * G::TX, y0 = ty * G::TY, z0 = tz * G::TZ;
      ok = mbar_wait(empty0 + 8 * s, ph ^ 1);
      if (!ok) {
        if (lane == 0)
          atomicOr(err_word, 0x100ull);
        break;
      }
      const uint32_t bar = full0 + 8 * s;
      if (warp == 0 && lane == 0) {
        F3Info &fi = info[s];
#pragma unroll
        for (int P = 0; P < 3; P++)
          fi.base[P] = (unsigned long long)(a.sym + ((uint64_t)(z0 >> (P + 1)) * sz +
                                                     (uint64_t)(y0 >> (P + 1)) * sy +
                                                     (x0 >> (P + 1))));
        fi.x0 = x0;
        fi.y0 = y0;
        fi.z0 = z0;
        fi.seq = seq;
        // (boxes of the other producer warps may land first: the phase cannot
        // complete before this arrival)
        mbar_expect_tx(bar, G::TX_BYTES);
      }
      if (fa.dbg & 4)
        asm volatile("bar.sync 1, %0;" ::"n"(F3_NP * 32) : "memory");
      if (has_op) {
        const uint32_t R = (z0 + bp) * ny + y0 + bm_;
        // start A(p) + B(m) elements before the node: a 16-byte boundary
        tma_load_2d(saddr(tiles + (size_t)s * G::TILE) + bdst, bmap, bar,
                    (int)((R & 3) * nx + x0) - bshift, (int)(R >> 2));
      }
      __syncwarp();
      next_stage();
    }
#ifndef MGX_F3_STATIC_TILES
    }
#endif
    // end markers: one for every consumer role's next tile slot (on a time-out
    // the markers are written without waiting, so that every role terminates)
    if (warp == 0)
      for (int e = 0; e < (F3_G > F3_C ? F3_G : F3_C); e++) {
        if (ok && !mbar_wait(empty0 + 8 * s, ph ^ 1)) {
          ok = false;
          if (lane == 0)
            atomicOr(err_word, 0x200ull);
        }
        if (lane == 0) {
          info[s].x0 = 0xffffffffu;
          info[s].seq = seq;
          mbar_arrive(full0 + 8 * s);
        }
        __syncwarp();
        next_stage();
      }
  } else {
    const int cw = warp - F3_NP;
    const Sink3<T> sink{a, fa.cmax, saddr(s_hist), saddr(obuf) + (uint32_t)cw * F3_OCAP * 16u,
                        saddr(ocnt + cw)};
    if (cw < F3_G * F3_GW) {
      // ---------------------------------- stride-1 pass: group g, warp iw
      const int g = cw / F3_GW, iw = cw % F3_GW;
      const int k = lane, ya = iw & 3, zb0 = (iw >> 2) * 2;
      const int ro[3] = {row_term<T>(2 * ya, fa.sh_b), row_term<T>(2 * ya + 1, fa.sh_b),
                         row_term<T>(2 * ya + 2, fa.sh_b)};
      uint32_t tc[7];
      class_offsets<T>(a, 0, (uint32_t)(zb0 * sz + ya * sy + k), tc);
      int s = g;
      uint32_t ph = 0;
      for (uint32_t j = g;; j += F3_G) {
        if (!full_wait(full0 + 8 * s, ph, &info[s].seq, j)) {
          if (lane == 0)
            atomicOr(err_word, 0x400ull);
          break;
        }
        const F3Info &fi = info[s];
        const uint32_t x0 = fi.x0;
        if (x0 == 0xffffffffu)
          break;
        if (!(fa.dbg & 1))
        march3<T, HALF, 0, 2>((const char *)tiles + (size_t)s * G::TILE, tables, a, sink, k,
                              ya, zb0, ro, fa.sh_c, fi.base[0], tc, x0, fi.y0, fi.z0);
        __syncwarp();
        if (lane == 0)
          mbar_arrive(empty0 + 8 * s);
        sink.flush(lane, F3_OFLUSH);
        s += F3_G;
        if (s >= S) {
          s -= S;
          ph ^= 1;
        }
      }
    } else {
      // ------------------- stride-2 and stride-4 passes: warp c of F3_C
      const int c = cw - F3_G * F3_GW;
      // stride 2: 16 x 2 x 2 blocks, lane = (k, ya), march over both z blocks
      const int k1 = lane & 15, ya1 = lane >> 4;
      const int ro1[3] = {row_term<T>(4 * ya1, fa.sh_b), row_term<T>(4 * ya1 + 2, fa.sh_b),
                          row_term<T>(4 * ya1 + 4, fa.sh_b)};
      uint32_t tc1[7];
      class_offsets<T>(a, 1, (uint32_t)(ya1 * sy + k1), tc1);
      // stride 4: 8 x 1 x 1 blocks, lanes 0..7
      const int ro2[3] = {row_term<T>(0, fa.sh_b), row_term<T>(4, fa.sh_b),
                          row_term<T>(8, fa.sh_b)};
      uint32_t tc2[7];
      class_offsets<T>(a, 2, (uint32_t)(lane & 7), tc2);
      int s = c % S;
      uint32_t ph = (c / S) & 1;
      for (uint32_t j = c;; j += F3_C) {
        if (!full_wait(full0 + 8 * s, ph, &info[s].seq, j)) {
          if (lane == 0)
            atomicOr(err_word, 0x800ull);
          break;
        }
        const F3Info &fi = info[s];
        const uint32_t x0 = fi.x0;
        if (x0 == 0xffffffffu)
          break;
        const char *tile = (const char *)tiles + (size_t)s * G::TILE;
        if (!(fa.dbg & 1)) {
        march3<T, HALF, 1, 2>(tile, tables, a, sink, k1, ya1, 0, ro1, fa.sh_c, fi.base[1], tc1, x0,
                              fi.y0, fi.z0);
        if (lane < 8)
          march3<T, HALF, 2, 1>(tile, tables, a, sink, lane, 0, 0, ro2, fa.sh_c, fi.base[2], tc2, x0,
                                fi.y0, fi.z0);
        }
        __syncwarp();
        if (lane == 0)
          mbar_arrive(empty0 + 8 * s);
        sink.flush(lane, F3_OFLUSH);
        s += F3_C;
        while (s >= S) {
          s -= S;
          ph ^= 1;
        }
      }
    }
    __syncwarp();
    sink.flush(lane, 1);
  }
  // flush the CTA histogram (the bin of dict/2 is restored by k_hist_fix)
  __syncthreads();
  for (uint32_t i = tid; i < a.dict; i += F3_THREADS) {
    const uint32_t v = s_hist[i];
    if (v)
      atomicAdd(&g_freq[i], v);
  }
}

// Nodes outside the structured passes: (1) inside the regular region, the
// multiples of 8 on every axis (coarser than the in-tile strides; neighbours up
// to half the grid away); (2) every node of the clipped tiles that touch the
// last node of an axis, i.e. the region outside [0, 8*rt0) x [0, 8*rt1) x
// [0, 64*rt2).  Per-node gather from global memory / L2.
template <typename T, bool HALF>
__global__ void __launch_bounds__(HIER_THREADS)
    k_fwd3_rest(const T *__restrict__ u, uint32_t *__restrict__ g_freq,
                const T *__restrict__ tables, const __grid_constant__ HierArgs<T> a,
                const __grid_constant__ F3Args<T> fa) {
  typedef TileShape<3, 3, 6> TS;
  extern __shared__ __align__(16) unsigned char smem[];
  uint32_t *s_hist = (uint32_t *)smem;
  for (uint32_t i = threadIdx.x; i < a.dict; i += HIER_THREADS)
    s_hist[i] = 0;
  __syncthreads();
  const SinkRest<T> sink{a, saddr(s_hist)};
  const uint32_t nz = a.n[0], ny = a.n[1], nx = a.n[2];
  const uint32_t RZ = fa.rt[0] * 8, RY = fa.rt[1] * 8, RX = fa.rt[2] * 64;
  // deep nodes of the regular region, then slabs z >= RZ, y >= RY, x >= RX
  // (N < 2^32 is a launch condition: every count below fits 32 bits, and 32-bit
  // divisions are a fraction of the cost of 64-bit ones on this per-node path)
  const uint32_t nd = fa.rt[0] * fa.rt[1] * (RX / 8);
  const uint32_t nA = (nz - RZ) * ny * nx;
  const uint32_t nB = RZ * (ny - RY) * nx;
  const uint32_t nC = RZ * RY * (nx - RX);
  const uint32_t total = nd + nA + nB + nC;
  const uint32_t stride = gridDim.x * HIER_THREADS;
  for (uint32_t i = blockIdx.x * HIER_THREADS + threadIdx.x; i < total;) {
    uint32_t gx, gy, gz;
    if (i < nd) {
      const uint32_t per = RX / 8;
      gx = (i % per) * 8;
      const uint32_t r = i / per;
      gy = (r % fa.rt[1]) * 8;
      gz = (r / fa.rt[1]) * 8;
    } else if (i < nd + nA) {
      const uint32_t j = i - nd;
      gx = j % nx;
      gy = (j / nx) % ny;
      gz = RZ + j / (nx * ny);
    } else if (i < nd + nA + nB) {
      const uint32_t j = i - nd - nA;
      const uint32_t wy = ny - RY;
      gx = j % nx;
      gy = RY + (j / nx) % wy;
      gz = j / (nx * wy);
    } else {
      const uint32_t j = i - nd - nA - nB;
      const uint32_t wx = nx - RX;
      gx = RX + j % wx;
      gy = (j / wx) % RY;
      gz = j / (wx * RY);
    }
    // (the last block's i + stride may wrap around 2^32: stop on overflow)
    const uint32_t next = i + stride;
    const bool last = next < i;
    i = next;
    element_generic<T, HALF, TS>(u, (const T *)nullptr, tables, a, gx, gy, gz, 0, 0, 0, 0,
                                 0, 0, sink);
    if (last)
      break;
  }
  __syncthreads();
  for (uint32_t i = threadIdx.x; i < a.dict; i += HIER_THREADS) {
    const uint32_t v = s_hist[i];
    if (v)
      atomicAdd(&g_freq[i], v);
  }
}

// freq[dict/2] = N - sum of all other bins (the kernels above never count it)
__global__ void __launch_bounds__(1024)
    k_hist_fix(uint32_t *__restrict__ freq, uint32_t dict, uint32_t modal,
               unsigned long long n) {
  __shared__ unsigned long long part[32];
  unsigned long long s = 0;
  for (uint32_t i = threadIdx.x; i < dict; i += 1024)
    if (i != modal)
      s += freq[i];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1)
    s += __shfl_xor_sync(0xffffffffu, s, o);
  if ((threadIdx.x & 31) == 0)
    part[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned long long t = 0;
    for (int i = 0; i < 32; i++)
      t += part[i];
    freq[modal] = (uint32_t)(n - t);
  }
}

// --------------------------------------------------------------- host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *,
                                  const cuuint64_t *, const cuuint64_t *, const cuuint32_t *,
                                  const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion,
                                  CUtensorMapFloatOOBfill);

EncodeTiledFn encode_tiled_fn() {
  static EncodeTiledFn fn = []() -> EncodeTiledFn {
    void *p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) !=
            cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
      return nullptr;
    return (EncodeTiledFn)p;
  }();
  return fn;
}

// largest |c| that still quantises to code 0 (exact: bisection over the bit
// pattern with the kernel's own arithmetic; host files build with
// -ffp-contract=off).  Negative when even c = 0 does not (non-finite quantiser).
template <typename T> T zero_threshold(T qv) {
  typedef typename std::conditional<sizeof(T) == 4, uint32_t, uint64_t>::type U;
  auto is_zero = [&](T c) {
    volatile T t = c * qv;
    volatile T r = (T)0.5 + std::fabs(t);
    return std::isfinite(r) && (long long)r == 0;
  };
  if (!std::isfinite(qv) || !is_zero((T)0))
    return (T)-1;
  U lo = 0, hi; // lo: is_zero true; hi: false
  const T inf = std::numeric_limits<T>::infinity();
  std::memcpy(&hi, &inf, sizeof(T));
  while (hi - lo > 1) {
    const U mid = lo + (hi - lo) / 2;
    T c;
    std::memcpy(&c, &mid, sizeof(T));
    if (is_zero(c))
      lo = mid;
    else
      hi = mid;
  }
  T c;
  std::memcpy(&c, &lo, sizeof(T));
  return c;
}

template <typename T, bool HALF>
int launch_fwd3(mgx_ctx *ctx, const void *d_in, HierArgs<T> &a, bool *used) {
  typedef F3<T> G;
  *used = false;
  const uint64_t nz = a.n[0], ny = a.n[1], nx = a.n[2];
  const uint64_t N = nz * ny * nx;
  EncodeTiledFn enc = encode_tiled_fn();
  if (!enc || ctx->no_tma || a.L < 3 || nz < 9 || ny < 9 || nx < 65 ||
      ((uintptr_t)d_in & 15) != 0 || N >= (1ull << 32) || ctx->dict > F3_MAX_DICT ||
      4 * nx >= (1ull << 31))
    return MGX_SUCCESS;
  F3Args<T> fa;
  fa.rt[0] = (uint32_t)((nz - 1) / G::TZ);
  fa.rt[1] = (uint32_t)((ny - 1) / G::TY);
  fa.rt[2] = (uint32_t)((nx - 1) / G::TX);
  // The box of the array's very last row is 16-byte granular and may reach up to
  // BX - 65 elements past the end of the array (only when that row is not the
  // fourth of its super row, where the tensor bound clips it).  Harmless as long
  // as those bytes share a page with the last element; else use the LDGSTS kernel.
  if ((uint64_t)(fa.rt[2] - 1) * G::TX + G::BX > nx && fa.rt[1] * (uint64_t)G::TY == ny - 1 &&
      fa.rt[0] * (uint64_t)G::TZ == nz - 1 && ((nz * ny) & 3) != 0) {
    const uintptr_t end = (uintptr_t)d_in + N * sizeof(T);
    if ((end & 4095) == 0 || (end & 4095) > 4096 - 32)
      return MGX_SUCCESS;
  }
  {
    const char *e = getenv("MGX_F3_DBG");
    fa.dbg = e ? (uint32_t)atoi(e) : 0;
  }
  fa.sh_b = (uint32_t)(nx % G::E);
  fa.sh_c = (uint32_t)((ny * nx) % G::E);
  fa.sby = (fa.rt[1] + 7) / 8;
  fa.sbz = (fa.rt[0] + 7) / 8;
  fa.padded = (uint64_t)fa.rt[2] * 64 * fa.sby * fa.sbz;
  fa.cmax = zero_threshold<T>(a.qv);
  if (getenv("MGX_F3_NOZERO"))
    fa.cmax = (T)-1;
  {
    uint64_t g = std::min<uint64_t>((uint64_t)fa.rt[0] * fa.rt[1] * fa.rt[2], (uint64_t)ctx->num_sms);
    fa.step[0] = (uint32_t)(g % fa.rt[2]);
    g /= fa.rt[2];
    fa.step[1] = (uint32_t)(g % 64);
    g /= 64;
    fa.step[2] = (uint32_t)(g % fa.sby);
    fa.step[3] = (uint32_t)(g / fa.sby);
  }
  {
    // tiles per draw: about 1/8 of a CTA's share, so that the last draws level the CTAs
    const uint64_t nt = (uint64_t)fa.rt[0] * fa.rt[1] * fa.rt[2];
    const uint64_t share = nt / (std::min<uint64_t>(nt, (uint64_t)ctx->num_sms) * 8);
    fa.chunk = 1;
    while (fa.chunk * 2 <= share && fa.chunk * 2 <= F3_CHUNK_MAX)
      fa.chunk *= 2;
  }
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
  a.tiles[0] = fa.rt[0];
  a.tiles[1] = fa.rt[1];
  a.tiles[2] = fa.rt[2];
  a.ntiles = (uint64_t)fa.rt[0] * fa.rt[1] * fa.rt[2];
  a.sym = ctx->d_sym;
  a.out_count = (unsigned long long *)(ctx->d_scalars + MGX_SC_OUTLIERS);
  a.out_idx = (unsigned long long *)ctx->d_out_idx;
  a.out_val = (long long *)ctx->d_out_val;
  a.out_cap = ctx->out_cap;

  CUtensorMap tm3, tm2;
  const cuuint64_t dims[2] = {4 * nx, (nz * ny + 3) / 4};
  const cuuint64_t strides[1] = {4 * nx * sizeof(T)};
  const cuuint32_t estr[2] = {1, 1};
  const CUtensorMapDataType dt =
      sizeof(T) == 4 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_FLOAT64;
  for (int rows = 2; rows <= 3; rows++) {
    const cuuint32_t box[2] = {(cuuint32_t)G::BX, (cuuint32_t)rows};
    CUresult r = enc(rows == 3 ? &tm3 : &tm2, dt, 2, const_cast<void *>(d_in), dims, strides,
                     box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                     CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS)
      return MGX_SUCCESS; // fall back to the LDGSTS kernel
  }
  constexpr int S = F3Cfg<T>::S;
  constexpr size_t OSTAGE =
      16 + (size_t)(F3Cfg<T>::WARPS - F3Cfg<T>::NP) * (F3_OCAP * 16 + 4) + 8 + F3_CHUNKQ * 8;
  static_assert(F3_CHUNKQ > F3Cfg<T>::S + 2,
                "the ring of draws must outlast the lag between producer warps");
  static_assert((size_t)S * G::TILE + F3_MAX_DICT * 4 + S * sizeof(F3Info) + 2 * S * 8 + OSTAGE <=
                    (size_t)F3_SMEM_MAX,
                "stages + histogram + outlier staging exceed the CTA's shared memory");
  const size_t smem = (size_t)S * G::TILE + (size_t)ctx->dict * 4 + S * sizeof(F3Info) +
                      2 * S * 8 + OSTAGE;
  auto kern = k_fwd3<T, HALF>;
  MGX_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                (int)smem));
  const unsigned grid = (unsigned)std::min<uint64_t>(a.ntiles, (uint64_t)ctx->num_sms);
  MGX_CUDA(cudaMemsetAsync(ctx->d_scalars + MGX_SC_TILE, 0, 8, ctx->stream)); // tile counter
  kern<<<grid, F3Cfg<T>::THREADS, smem, ctx->stream>>>(ctx->d_freq, (const T *)ctx->d_tables, a,
                                                fa, tm3, tm2);
  ctx->launches++;
  {
    auto rest = k_fwd3_rest<T, HALF>;
    const size_t hsmem = (size_t)ctx->dict * 4;
    MGX_CUDA(cudaFuncSetAttribute(rest, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  (int)hsmem));
    const uint64_t RZ = fa.rt[0] * 8ull, RY = fa.rt[1] * 8ull, RX = fa.rt[2] * 64ull;
    const uint64_t work = N - RZ * RY * RX + a.ntiles * 8;
    const unsigned rgrid = mgx_grid_for(work, HIER_THREADS * 2, ctx->num_sms * 6);
    rest<<<rgrid, HIER_THREADS, hsmem, ctx->stream>>>(
        (const T *)d_in, ctx->d_freq, (const T *)ctx->d_tables, a, fa);
    ctx->launches++;
  }
  k_hist_fix<<<1, 1024, 0, ctx->stream>>>(ctx->d_freq, ctx->dict, ctx->dict / 2, N);
  ctx->launches++;
  MGX_CUDA(cudaGetLastError());
  *used = true;
  return MGX_SUCCESS;
}

} // namespace

// Entry used by mgx_hier.cu: runs the TMA path when the shape / pointer allow
// it (*used = true), else leaves everything to the caller's LDGSTS kernel.
template <typename T>
int mgxk_hier_forward3(mgx_ctx *ctx, const void *d_in, const void *args, bool half,
                       bool *used) {
  HierArgs<T> a = *reinterpret_cast<const HierArgs<T> *>(args);
  return half ? launch_fwd3<T, true>(ctx, d_in, a, used)
              : launch_fwd3<T, false>(ctx, d_in, a, used);
}
template int mgxk_hier_forward3<float>(mgx_ctx *, const void *, const void *, bool, bool *);
template int mgxk_hier_forward3<double>(mgx_ctx *, const void *, const void *, bool, bool *);
