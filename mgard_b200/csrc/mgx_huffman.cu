// Huffman stage on the GPU (sm_100a): symbol staging, histogram, fused
// encode + bit-pack with a decoupled look-back scan, container assembly,
// table-driven decode.
//
// reference: Lossless/ParallelHuffman/Huffman.hpp:157-632 (stage sequencing and
// the "MGXHUFF" container), OutlierSeparator.hpp:17-120, Histogram.hpp,
// ParallelDeflate.hpp:19-357 (group bits -> scans -> pack), Decode.hpp:14-206.
//
// Differences in *how* (never in the bytes produced):
//  * dictionary symbols travel as uint16 (the reference keeps int64);
//  * per-chunk word offsets come from a single-pass decoupled look-back scan
//    fused into the bit-count kernel instead of two device-wide scans;
//  * one warp packs one chunk into shared memory and writes it out coalesced
//    (the reference packs 32 symbols per thread straight to global memory with
//    atomicOr on shared words);
//  * decoding uses a 2^12-entry prefix table in shared memory with the
//    canonical first[]/entry[] search only for longer codes.
#include <algorithm>
#include <limits>
#include <cub/device/device_radix_sort.cuh>

#include "mgx_device.cuh"
#include "mgx_internal.h"

size_t mgx_codebook_ws_bytes(int dict, int P);
cudaError_t mgx_launch_codebook(const uint32_t *d_freq, int dict, int P,
                                uint64_t *d_codebook, uint64_t *d_decodebook,
                                unsigned char *gws, int use_smem, size_t smem,
                                uint64_t *d_scalars, cudaStream_t stream);

namespace {

typedef unsigned long long u64;

// ---- int64 codes <-> 16-bit symbols -------------------------------------------
// OutlierSeparator (SEPARATE) + Histogram in one pass over the int64 codes.
__global__ void __launch_bounds__(256)
    k_codes_to_symbols(const long long *__restrict__ codes, uint64_t n,
                       uint32_t dict, uint16_t *__restrict__ sym,
                       uint32_t *__restrict__ g_freq, u64 *__restrict__ out_count,
                       u64 *__restrict__ out_idx, long long *__restrict__ out_val,
                       u64 out_cap) {
  extern __shared__ uint32_t s_hist[];
  for (uint32_t i = threadIdx.x; i < dict; i += blockDim.x)
    s_hist[i] = 0;
  __syncthreads();
  const uint64_t step = (uint64_t)gridDim.x * blockDim.x;
  const uint64_t rounds = (n + step - 1) / step;
  for (uint64_t k = 0; k < rounds; k++) {
    uint64_t it = k * step + blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
    bool valid = it < n;
    uint32_t sv = 0;
    if (valid) {
      long long q = codes[it];
      if (q < 0 || q >= (long long)dict) {
        u64 slot = atomicAdd(out_count, 1ull);
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
  for (uint32_t i = threadIdx.x; i < dict; i += blockDim.x) {
    uint32_t v = s_hist[i];
    if (v)
      atomicAdd(&g_freq[i], v);
  }
}

__global__ void __launch_bounds__(256)
    k_symbols_to_codes(const uint16_t *__restrict__ sym, uint64_t n,
                       long long *__restrict__ codes) {
  for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n;
       i += (uint64_t)gridDim.x * blockDim.x)
    codes[i] = (long long)sym[i];
}

// OutlierSeparator (RESTORE), OutlierSeparator.hpp:45-51
__global__ void __launch_bounds__(256)
    k_restore_outliers(const u64 *__restrict__ idx,
                       const long long *__restrict__ val, uint64_t n_out,
                       long long *__restrict__ codes) {
  for (uint64_t j = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; j < n_out;
       j += (uint64_t)gridDim.x * blockDim.x)
    codes[idx[j]] = val[j];
}

} // namespace

int mgxk_histogram_reset(mgx_ctx *ctx) {
  MGX_CUDA(cudaMemsetAsync(ctx->d_freq, 0, (size_t)ctx->dict * 4, ctx->stream));
  MGX_CUDA(cudaMemsetAsync(ctx->d_scalars, 0, MGX_SC_COUNT * 8, ctx->stream));
  return MGX_SUCCESS;
}

int mgxk_codes_to_symbols(mgx_ctx *ctx, const int64_t *d_codes, uint64_t n) {
  size_t smem = (size_t)ctx->dict * 4;
  MGX_CUDA(cudaFuncSetAttribute(k_codes_to_symbols,
                                cudaFuncAttributeMaxDynamicSharedMemorySize,
                                (int)smem));
  unsigned grid = mgx_grid_for(n, 256 * 16, ctx->num_sms * 4);
  k_codes_to_symbols<<<grid, 256, smem, ctx->stream>>>(
      (const long long *)d_codes, n, ctx->dict, ctx->d_sym, ctx->d_freq,
      (u64 *)(ctx->d_scalars + MGX_SC_OUTLIERS), (u64 *)ctx->d_out_idx,
      (long long *)ctx->d_out_val, ctx->out_cap);
  ctx->launches++;
  MGX_CUDA(cudaGetLastError());
  return MGX_SUCCESS;
}

int mgxk_symbols_to_codes(mgx_ctx *ctx, uint64_t n, int64_t *d_codes,
                          const uint64_t *d_oidx, const int64_t *d_oval,
                          uint64_t n_out) {
  k_symbols_to_codes<<<mgx_grid_for(n, 256, ctx->num_sms * 16), 256, 0,
                       ctx->stream>>>(ctx->d_sym, n, (long long *)d_codes);
  ctx->launches++;
  if (n_out) {
    k_restore_outliers<<<mgx_grid_for(n_out, 256, ctx->num_sms * 4), 256, 0,
                         ctx->stream>>>((const u64 *)d_oidx,
                                        (const long long *)d_oval, n_out,
                                        (long long *)d_codes);
    ctx->launches++;
  }
  MGX_CUDA(cudaGetLastError());
  return MGX_SUCCESS;
}

// GetCodebook, GetCodebook.hpp:23-145 (see mgx_codebook.cu)
int mgxk_codebook(mgx_ctx *ctx, const uint32_t *d_freq, uint64_t *d_codebook,
                  uint64_t *d_decodebook) {
  int dict = (int)ctx->dict;
  if (dict > 16384) {
    ctx->err = "dictionary sizes above 16384 are not supported";
    return MGX_FAILURE;
  }
  int P = 1;
  while (P < dict)
    P <<= 1;
  size_t ws = mgx_codebook_ws_bytes(dict, P);
  int use_smem = ws <= 200 * 1024;
  // L2-resident fallback workspace (large dicts), owned by the context
  if (!use_smem && ctx->cb_ws_bytes < ws) {
    if (ctx->d_cb_ws)
      cudaFree(ctx->d_cb_ws);
    ctx->d_cb_ws = nullptr;
    ctx->cb_ws_bytes = 0;
    MGX_CUDA(cudaMalloc(&ctx->d_cb_ws, ws));
    ctx->cb_ws_bytes = ws;
  }
  void *gws = ctx->d_cb_ws;
  size_t smem = use_smem ? ws : 0;
  MGX_CUDA(mgx_launch_codebook(d_freq, dict, P, d_codebook, d_decodebook,
                               (unsigned char *)gws, use_smem, smem,
                               ctx->d_scalars, ctx->stream));
  ctx->launches++;
  MGX_CUDA(cudaGetLastError());
  return MGX_SUCCESS;
}

// ============================ encode ==========================================
namespace {

constexpr int ENC_WARPS = 8; // chunks per tile (one warp each)
constexpr u64 FLAG_AGG = 1ull << 62, FLAG_INC = 2ull << 62,
              VAL_MASK = (1ull << 62) - 1;

// Pass 1 (DeflateGroupBits + both ScanSumExtended + DeflateChunkMeta,
// ParallelDeflate.hpp:39-175, Huffman.hpp:243-274): per-chunk bit length,
// ceil(bits/64) words, and the exclusive scan of the word counts through a
// decoupled look-back over tiles of ENC_WARPS chunks.  Grid is persistent and
// fully resident, tiles are taken round-robin, so a tile's predecessors are
// always owned by running CTAs.
__global__ void __launch_bounds__(ENC_WARPS * 32)
    k_chunk_bits(const uint16_t *__restrict__ sym, uint64_t n, uint32_t chunk,
                 uint32_t dict, const u64 *__restrict__ codebook,
                 u64 *__restrict__ chunk_bits, u64 *__restrict__ chunk_off,
                 u64 *__restrict__ tile_state, uint64_t nchunk, uint64_t ntiles,
                 u64 *__restrict__ scalars) {
  extern __shared__ uint8_t s_len[]; // code length per symbol
  __shared__ u64 s_words[ENC_WARPS];
  __shared__ u64 s_prefix;
  for (uint32_t i = threadIdx.x; i < dict; i += blockDim.x)
    s_len[i] = (uint8_t)(codebook[i] >> 56);
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (uint64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const uint64_t c = tile * ENC_WARPS + warp;
    u64 bits = 0;
    if (c < nchunk) {
      const uint64_t b = c * chunk;
      const uint64_t e = min(b + (uint64_t)chunk, n);
      for (uint64_t i = b + lane; i < e; i += 32)
        bits += s_len[sym[i]];
      for (int o = 16; o > 0; o >>= 1)
        bits += __shfl_xor_sync(0xffffffffu, bits, o);
    }
    const u64 words = (bits + 63) >> 6;
    if (lane == 0) {
      s_words[warp] = words;
      if (c < nchunk)
        chunk_bits[c] = bits;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      u64 agg = 0;
      for (int w = 0; w < ENC_WARPS; w++)
        agg += s_words[w];
      u64 prefix = 0;
      if (tile > 0) {
        volatile u64 *vs = tile_state;
        __threadfence();
        vs[tile] = FLAG_AGG | agg;
        int64_t p = (int64_t)tile - 1;
        while (true) {
          u64 v = vs[p];
          if (v & FLAG_INC) {
            prefix += v & VAL_MASK;
            break;
          }
          if (v & FLAG_AGG) {
            prefix += v & VAL_MASK;
            --p;
          }
        }
      }
      __threadfence();
      ((volatile u64 *)tile_state)[tile] = FLAG_INC | (prefix + agg);
      s_prefix = prefix;
      if (tile == ntiles - 1) {
        chunk_off[nchunk] = prefix + agg;
        scalars[MGX_SC_DDATA] = prefix + agg;
      }
    }
    __syncthreads();
    if (lane == 0 && c < nchunk) {
      u64 off = s_prefix;
      for (int w = 0; w < warp; w++)
        off += s_words[w];
      chunk_off[c] = off;
    }
    __syncthreads();
  }
}

// Pass 2 (DeflatePack, ParallelDeflate.hpp:177-266): one warp packs one chunk,
// MSB first, into a zeroed shared-memory image of the chunk's words, then
// streams it to its final place in the container.  Lane l owns the contiguous
// run of ceil(chunk/32) symbols the reference calls a "group".
__global__ void __launch_bounds__(ENC_WARPS * 32)
    k_pack(const uint16_t *__restrict__ sym, uint64_t n, uint32_t chunk,
           const u64 *__restrict__ codebook, const u64 *__restrict__ chunk_off,
           u64 *__restrict__ packed, uint64_t nchunk, uint32_t words_cap) {
  extern __shared__ __align__(16) u64 s_out[]; // ENC_WARPS * words_cap
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  u64 *out = s_out + (size_t)warp * words_cap;
  const uint32_t G = (chunk + 31) / 32;
  const uint64_t nwarps = (uint64_t)gridDim.x * ENC_WARPS;
  for (uint64_t c = (uint64_t)blockIdx.x * ENC_WARPS + warp; c < nchunk;
       c += nwarps) {
    const uint64_t cb = c * chunk;
    const uint64_t ce = min(cb + (uint64_t)chunk, n);
    const uint64_t gb = min(cb + (uint64_t)lane * G, ce);
    const uint64_t ge = min(gb + G, ce);
    const u64 w0 = chunk_off[c], w1 = chunk_off[c + 1];
    const uint32_t nwords = (uint32_t)(w1 - w0);
    for (uint32_t i = lane; i < nwords; i += 32)
      out[i] = 0;
    // group bit count -> exclusive warp scan -> my first bit
    uint32_t gbits = 0;
    for (uint64_t i = gb; i < ge; i++)
      gbits += (uint32_t)(__ldg(&codebook[sym[i]]) >> 56);
    uint32_t incl = gbits;
    for (int o = 1; o < 32; o <<= 1) {
      uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o)
        incl += t;
    }
    uint32_t bit = incl - gbits;
    __syncwarp();
    // append my codes; words I do not fully own are merged with atomicOr
    u64 acc = 0;
    uint32_t word = bit >> 6, fill = bit & 63; // bits already used in `word`
    bool first_word = true;
    const bool left_shared = fill != 0;
    for (uint64_t i = gb; i < ge; i++) {
      u64 cw = __ldg(&codebook[sym[i]]);
      uint32_t len = (uint32_t)(cw >> 56);
      cw &= 0x00ffffffffffffffull;
      uint32_t room = 64 - fill;
      if (len < room) {
        acc |= cw << (room - len);
        fill += len;
      } else {
        uint32_t r = len - room; // bits spilling into the next word
        acc |= cw >> r;
        if (first_word && left_shared)
          atomicOr(&out[word], acc);
        else
          out[word] = acc;
        first_word = false;
        ++word;
        acc = r ? (cw << (64 - r)) : 0;
        fill = r;
      }
    }
    if (fill != 0 && (gb < ge))
      atomicOr(&out[word], acc);
    __syncwarp();
    u64 *dst = packed + w0;
    for (uint32_t i = lane; i < nwords; i += 32)
      dst[i] = out[i];
    __syncwarp();
  }
}

// ---- single-pass encoder ---------------------------------------------------
// DeflateGroupBits + scans + DeflateChunkMeta + DeflatePack in ONE pass over the
// symbols (2 B read per symbol, packed words written once).  Everything is
// warp-granular -- no CTA barrier inside the loop:
//  * a warp owns one chunk at a time (round-robin over a persistent, fully
//    resident grid of one CTA per SM); a lane holds 8 R CONSECUTIVE symbols (the
//    reference's 32-symbol "group" for the default chunk of 1024), fetched with R
//    16-byte loads;
//  * length pass (length_pass): table look-ups in shared memory, four symbols
//    merged per 32-bit accumulator, two accumulators per 64-bit run blob; ONE warp
//    scan per chunk turns the lanes' bit counts into first-bit offsets;
//  * the chunk's word offset comes from a decoupled look-back over tiles of EW
//    chunks (scanner warp), overlapped with the packing of the next two tiles;
//  * a lane appends its blobs in order through LaneSink into a zeroed
//    shared-memory image of the chunk (at most two atomicOr per lane: the units it
//    shares with its neighbours), which is then streamed to its final place in
//    the container.  A chunk whose image would not fit (more than ENC_CAP_BITS bits
//    per symbol on average) is packed straight into global memory instead.
constexpr int ENC_CAP_BITS = 10; // image capacity, bits per symbol on average
constexpr int ENC_IMGS = 3;      // images per packer warp: packers run 2 tiles ahead
constexpr u64 CODE_MASK = 0x00ffffffffffffffull;

__device__ __forceinline__ uint32_t sym_of_dyn(const uint4 &v, int e) {
  const uint32_t w = e < 2 ? v.x : e < 4 ? v.y : e < 6 ? v.z : v.w;
  return (e & 1) ? (w >> 16) : (w & 0xffffu);
}

// Lane-sequential bit sink: a lane appends the codes of its OWN contiguous run of
// symbols, MSB first.  Units it fills from their first bit are stored plainly;
// only the unit it starts in (when it does not start on a unit boundary) and the
// unit it ends in can be shared with neighbouring lanes and are merged with
// atomicOr -- at most two shared-memory atomics per lane and chunk, where the
// per-run placement needed up to three per run of 8 symbols.
struct LaneSink {
  uint32_t *dst; // unit u of the MSB-first stream lives at dst[u ^ 1]
  uint32_t unit, fill, cur;
  bool shared;
  __device__ __forceinline__ void begin(uint32_t *d, uint32_t bit) {
    dst = d;
    unit = bit >> 5;
    fill = bit & 31;
    cur = 0;
    shared = fill != 0;
  }
  __device__ __forceinline__ void flush(uint32_t w) {
    if (shared) {
      atomicOr(&dst[unit ^ 1], w);
      shared = false;
    } else {
      dst[unit ^ 1] = w;
    }
    ++unit;
  }
  // append the low nb bits of a (0 < nb <= 64)
  __device__ __forceinline__ void append(u64 a, uint32_t nb) {
    const u64 A = a << (64 - nb);
    const uint32_t w0 = cur | (uint32_t)(A >> (32 + fill));
    const uint32_t t = fill + nb;
    if (t >= 32) {
      flush(w0);
      const uint32_t w1 = (uint32_t)(A >> fill);
      if (t >= 64) {
        flush(w1);
        cur = (uint32_t)(A << (32 - fill));
        fill = t - 64;
      } else {
        cur = w1;
        fill = t - 32;
      }
    } else {
      cur = w0;
      fill = t;
    }
  }
  __device__ __forceinline__ void end() {
    if (fill)
      atomicOr(&dst[unit ^ 1], cur);
  }
};

__device__ __forceinline__ uint32_t lds_u32(uint32_t saddr) {
  uint32_t v;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(saddr));
  return v;
}

// Length pass over one lane's 8 R consecutive symbols.  s_tab[s] = len << 26 |
// code for codes of at most 26 bits, 0xffffffff otherwise.  Four symbols are
// merged in a 32-bit accumulator, two such halves make the run's 64-bit blob; a
// half whose lengths add up to more than 32 bits (which includes every marker:
// its length field reads 63) sends the run to the per-symbol path.
template <int R, bool FULL>
__device__ __forceinline__ void
length_pass(const uint4 (&v)[R], const int (&cnt)[R], uint32_t tab, u64 (&blob)[R],
            uint32_t (&rbits)[R], unsigned &slow, unsigned &inexact, uint32_t &tot) {
  slow = 0;
  inexact = 0;
  tot = 0;
#pragma unroll
  for (int j = 0; j < R; j++) {
    const uint32_t w[4] = {v[j].x, v[j].y, v[j].z, v[j].w};
    uint32_t acc[2], hb[2];
#pragma unroll
    for (int h = 0; h < 2; h++) {
      uint32_t a = 0, b = 0;
#pragma unroll
      for (int e = 0; e < 4; e++) {
        const int idx = 4 * h + e;
        if (FULL || idx < cnt[j]) {
          const uint32_t ww = w[idx >> 1];
          // byte offset of the symbol's entry: sym * 4
          const uint32_t off = (idx & 1) ? ((ww >> 14) & 0x3fffcu) : ((ww << 2) & 0x3fffcu);
          const uint32_t t = lds_u32(tab + off);
          const uint32_t len = t >> 26;
          b += len;
          a = (a << len) | (t & 0x3ffffffu);
        }
      }
      acc[h] = a;
      hb[h] = b;
    }
    rbits[j] = hb[0] + hb[1];
    blob[j] = ((u64)acc[0] << hb[1]) | acc[1];
    if (max(hb[0], hb[1]) > 32u) {
      slow |= 1u << j;
      if (max(hb[0], hb[1]) >= 63u) // possibly a marker: the sum is not the bit count
        inexact |= 1u << j;
    }
    tot += rbits[j];
  }
}

// exact bit count of a run on the per-symbol path (codes longer than 26 bits come
// from the global codebook)
__device__ __forceinline__ uint32_t slow_run_bits(const uint4 &v, int cnt, uint32_t tab,
                                                  const u64 *__restrict__ codebook) {
  uint32_t rb = 0;
#pragma unroll 1
  for (int e = 0; e < cnt; e++) {
    const uint32_t sy = sym_of_dyn(v, e);
    const uint32_t t = lds_u32(tab + sy * 4u);
    rb += t == 0xffffffffu ? (uint32_t)(__ldg(&codebook[sy]) >> 56) : (t >> 26);
  }
  return rb;
}

// Tiles of ENC_WARPS chunks, one per packer warp, plus a ninth "scanner" warp.
// Nobody meets at a CTA barrier: a packer adds its word count to the tile's
// slot in shared memory and goes on packing; the scanner waits for the eight
// arrivals, publishes the tile aggregate and runs the decoupled look-back over
// tiles (32 predecessors per probe).  Packers are one tile ahead of the scan:
// they stream tile i-1 out of its shared-memory image after packing tile i, so
// the latency of the look-back chain (which couples all resident CTAs) is
// hidden behind a full tile of work.
constexpr int ENC_SLOTS = 8;
// packer warps (= chunks per tile) of the single-pass encoder.  ONE CTA per SM
// with many warps, not several small CTAs: the look-back chain advances about 32
// tiles per L2 round trip, so a wave of tiles (one per resident CTA) costs
// ceil(CTAs / 32) round trips -- measured with 444 CTAs of 8 packers: 5.6 us per
// wave against 1.5 us of packing, the whole kernel waiting for prefixes.
template <int R> struct EncCfg {
  static constexpr int EW = R == 8 ? 16 : 24;
};

// mbarrier (shared::cta) for the scanner -> packers hand-over: a waiting warp is
// suspended by the hardware inside try_wait instead of spinning on a shared-memory
// flag -- measured: __nanosleep(100) in that spin returned after ~10 cycles, and the
// spin was 32 % of all executed instructions of an issue-bound kernel.
__device__ __forceinline__ void enc_mbar_init(uint32_t addr, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(addr), "r"(count)
               : "memory");
}
__device__ __forceinline__ void enc_mbar_arrive(uint32_t addr) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(addr) : "memory");
}
__device__ __forceinline__ bool enc_mbar_try_wait(uint32_t addr, uint32_t parity) {
  uint32_t ok;
  asm volatile("{\n\t.reg .pred p;\n\t"
               "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
               "selp.u32 %0, 1, 0, p;\n\t}"
               : "=r"(ok)
               : "r"(addr), "r"(parity)
               : "memory");
  return ok != 0;
}

template <int R>
__global__ void __launch_bounds__((EncCfg<R>::EW + 1) * 32, 1)
    k_encode(const uint16_t *__restrict__ sym, uint64_t n, uint32_t dict,
             const u64 *__restrict__ codebook, u64 *__restrict__ chunk_bits,
             u64 *__restrict__ chunk_off, u64 *__restrict__ packed,
             uint64_t cap_words, u64 *__restrict__ state, uint64_t nchunk,
             uint64_t ntiles, u64 *__restrict__ scalars) {
  constexpr int CHUNK = 256 * R;
  constexpr int EW = EncCfg<R>::EW;
  constexpr int UNITS = CHUNK * ENC_CAP_BITS / 32; // per-warp image, u32 units
  extern __shared__ __align__(16) unsigned char enc_smem[];
  uint32_t *s_img = (uint32_t *)enc_smem;                  // [ENC_IMGS][EW][UNITS]
  uint32_t *s_tab = s_img + ENC_IMGS * EW * UNITS;  // dict entries
  __shared__ uint32_t s_words[ENC_SLOTS][EW];
  __shared__ uint32_t s_arrived[ENC_SLOTS];
  __shared__ u64 s_prefix[ENC_SLOTS];
  __shared__ __align__(8) u64 s_bar[ENC_SLOTS]; // completes once per tile of the slot
  for (uint32_t i = threadIdx.x; i < dict; i += blockDim.x) {
    const u64 cw = codebook[i];
    s_tab[i] = (cw >> 56) <= 26 ? (uint32_t)(cw >> 56) << 26 | (uint32_t)(cw & 0x3ffffffu)
                                : 0xffffffffu;
  }
  if (threadIdx.x < ENC_SLOTS) {
    s_arrived[threadIdx.x] = 0;
    enc_mbar_init((uint32_t)__cvta_generic_to_shared(&s_bar[threadIdx.x]), 1);
  }
  __syncthreads();
  const uint32_t bar0 = (uint32_t)__cvta_generic_to_shared(&s_bar[0]);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t tab = (uint32_t)__cvta_generic_to_shared(s_tab);
  volatile u64 *vs = state;
  volatile uint32_t *v_arrived = s_arrived;
  volatile u64 *v_prefix = s_prefix;
  const uint64_t t_begin = blockIdx.x, t_end = ntiles, t_step = gridDim.x;
  const u64 region_words = cap_words;

  if (warp == EW) { // ---- scanner ----------------------------------------
    uint32_t iter = 0;
    for (uint64_t tile = t_begin; tile < t_end; tile += t_step, ++iter) {
      const int slot = iter & (ENC_SLOTS - 1);
      while (v_arrived[slot] != EW)
        __nanosleep(40); // leave the issue slots to the packers
      __threadfence_block();
      u64 agg = lane < EW ? ((volatile uint32_t *)s_words[slot])[lane] : 0;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1)
        agg += __shfl_xor_sync(0xffffffffu, agg, o);
      u64 prefix = 0;
      if (tile > 0) {
        if (lane == 0)
          vs[tile] = FLAG_AGG | agg;
        int64_t p = (int64_t)tile - 1;
        while (true) {
          const int64_t idx = p - lane;
          u64 x = 2ull << 62; // before tile 0: inclusive prefix 0
          if (idx >= 0)
            x = vs[idx];
          const unsigned inc = __ballot_sync(0xffffffffu, (x >> 63) != 0);
          const int stop = inc ? __ffs(inc) - 1 : 31; // nearest inclusive prefix
          const unsigned need = stop == 31 ? 0xffffffffu : ((2u << stop) - 1u);
          const unsigned ready = __ballot_sync(0xffffffffu, (x >> 62) != 0);
          if ((ready & need) != need) {
            __nanosleep(20);
            continue;
          }
          u64 part = lane <= stop ? (x & VAL_MASK) : 0;
#pragma unroll
          for (int o = 16; o > 0; o >>= 1)
            part += __shfl_xor_sync(0xffffffffu, part, o);
          prefix += part;
          if (inc)
            break;
          p -= 32;
        }
      }
      if (lane == 0) {
        vs[tile] = FLAG_INC | (prefix + agg);
        if (tile == ntiles - 1)
          scalars[MGX_SC_DDATA] = prefix + agg;
        v_prefix[slot] = prefix;
        v_arrived[slot] = 0;
        enc_mbar_arrive(bar0 + 8u * slot); // release: the prefix is visible after the wait
      }
      __syncwarp();
    }
    return;
  }

  // ---- packers -------------------------------------------------------------------
  // stream a finished tile out of its image once the scanner has its prefix
  auto finish = [&](uint32_t it, uint64_t c, uint32_t words) {
    const int slot = it & (ENC_SLOTS - 1);
    // tile `it` is the (it / ENC_SLOTS)-th use of its slot
    const uint32_t parity = (it / ENC_SLOTS) & 1u;
    while (!enc_mbar_try_wait(bar0 + 8u * slot, parity)) {
    }
    uint32_t before = lane < warp ? ((volatile uint32_t *)s_words[slot])[lane] : 0u;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
      before += __shfl_xor_sync(0xffffffffu, before, o);
    const u64 off = v_prefix[slot] + before;
    if (lane == 0)
      chunk_off[c] = off;
    if (off + words > region_words) {
      if (lane == 0)
        atomicOr((unsigned long long *)&scalars[MGX_SC_ERR], 8ull);
      return ~0ull;
    }
    return off;
  };
  uint32_t iter = 0;
  // packed tiles waiting in their images for the scan: A = tile iter - 2,
  // B = tile iter - 1.  Two tiles of slack absorb the skew between CTAs (the
  // prefix of a tile exists only when every earlier tile has been counted).
  bool pendA = false, pendB = false;
  uint64_t cA = 0, cB = 0;
  uint32_t wA = 0, wB = 0, itA = 0, itB = 0;
  auto drain = [&](uint32_t it, uint64_t c, uint32_t words) {
    const u64 off = finish(it, c, words);
    if (off != ~0ull) {
      const u64 *src =
          (const u64 *)(s_img + ((it % ENC_IMGS) * EW + warp) * UNITS);
      for (uint32_t i = lane; i < words; i += 32)
        packed[off + i] = src[i];
    }
  };
  for (uint64_t tile = t_begin; tile < t_end; tile += t_step, ++iter) {
    const int slot = iter & (ENC_SLOTS - 1);
    uint32_t *img = s_img + ((iter % ENC_IMGS) * EW + warp) * UNITS;
    const uint64_t c = tile * EW + warp;
    const bool have = c < nchunk;
    const uint64_t base = c * CHUNK;
    const bool full = have && base + CHUNK <= n;
    uint4 v[R];
    int cnt[R];
    {
      // pull the chunk this warp packs NEXT into L2 now: a warp works on one
      // chunk at a time, so its 2 B/symbol would otherwise arrive at DRAM
      // latency in front of every tile
      const uint64_t nb = ((tile + t_step) * EW + warp) * (uint64_t)CHUNK;
      const uint64_t pa = nb + (uint64_t)lane * 64; // 128-byte lines of uint16
      if (lane * 64 < CHUNK && pa < n)
        asm volatile("prefetch.global.L2 [%0];" ::"l"(sym + pa));
    }
    // a lane owns 8 R CONSECUTIVE symbols (64 bytes for R = 4): R 16-byte loads
#pragma unroll
    for (int j = 0; j < R; j++) {
      const uint64_t idx = base + (uint64_t)lane * (8 * R) + 8 * j;
      cnt[j] = full ? 8
                    : (!have || idx >= n) ? 0 : (n - idx >= 8 ? 8 : (int)(n - idx));
      v[j] = cnt[j] > 0 ? *(const uint4 *)(sym + idx) : make_uint4(0, 0, 0, 0);
    }
    // length pass: every run of 8 symbols becomes one MSB-first blob
    u64 blob[R];
    uint32_t rbits[R];
    unsigned slow, inexact;
    uint32_t tot;
    if (full)
      length_pass<R, true>(v, cnt, tab, blob, rbits, slow, inexact, tot);
    else
      length_pass<R, false>(v, cnt, tab, blob, rbits, slow, inexact, tot);
    if (__builtin_expect(inexact != 0, 0)) {
#pragma unroll
      for (int j = 0; j < R; j++)
        if ((inexact >> j) & 1u) {
          tot -= rbits[j];
          rbits[j] = slow_run_bits(v[j], cnt[j], tab, codebook);
          tot += rbits[j];
        }
    }
    // one scan per chunk: first bit of this lane's symbols
    uint32_t incl = tot;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o)
        incl += t;
    }
    const uint32_t start = incl - tot;
    const u64 bits = __shfl_sync(0xffffffffu, incl, 31);
    const uint32_t words = (uint32_t)((bits + 63) >> 6);
    const bool fits = 2 * words <= (uint32_t)UNITS;
    if (lane == 0) {
      if (have)
        chunk_bits[c] = bits;
      ((volatile uint32_t *)s_words[slot])[warp] = words;
      __threadfence_block();
      atomicAdd(&s_arrived[slot], 1u);
    }
    // the lane appends its runs in order
    auto emit = [&](uint32_t *dst) {
      LaneSink sk;
      sk.begin(dst, start);
#pragma unroll
      for (int j = 0; j < R; j++) {
        if (__builtin_expect((slow >> j) & 1u, 0)) {
          // more than 32 bits in a group of four symbols: merge each group in 64
          // bits; symbol by symbol only when even that overflows or a code is
          // longer than 26 bits
#pragma unroll 1
          for (int h = 0; h < 2; h++) {
            const int e0 = 4 * h, e1 = min(4 * h + 4, cnt[j]);
            u64 a = 0;
            uint32_t b = 0;
            bool marker = false;
            for (int e = e0; e < e1; e++) {
              const uint32_t t = lds_u32(tab + sym_of_dyn(v[j], e) * 4u);
              marker |= t == 0xffffffffu;
              b += t >> 26;
              a = (a << (t >> 26)) | (t & 0x3ffffffu);
            }
            if (!marker && b <= 64) {
              if (b)
                sk.append(a, b);
            } else {
              for (int e = e0; e < e1; e++) {
                const uint32_t sy = sym_of_dyn(v[j], e);
                const uint32_t t = lds_u32(tab + sy * 4u);
                if (t == 0xffffffffu) {
                  const u64 cw = __ldg(&codebook[sy]);
                  sk.append(cw & CODE_MASK, (uint32_t)(cw >> 56));
                } else if (t >> 26) {
                  sk.append(t & 0x3ffffffu, t >> 26);
                }
              }
            }
          }
        } else if (rbits[j]) {
          sk.append(blob[j], rbits[j]);
        }
      }
      sk.end();
    };
    if (have && fits) {
      for (uint32_t i = lane; i < 2 * words; i += 32)
        img[i] = 0;
      __syncwarp();
      emit(img);
      __syncwarp();
    }
    if (pendA)
      drain(itA, cA, wA);
    pendA = pendB;
    cA = cB;
    wA = wB;
    itA = itB;
    pendB = false;
    if (have && fits) {
      pendB = true;
      cB = c;
      wB = words;
      itB = iter;
    } else if (have) { // oversized chunk: pack straight into the container
      const u64 off = finish(iter, c, words);
      if (off != ~0ull) {
        for (uint32_t i = lane; i < words; i += 32)
          packed[off + i] = 0;
        __syncwarp();
        emit((uint32_t *)(packed + off));
      }
    }
    __syncwarp();
  }
  if (pendA)
    drain(itA, cA, wA);
  if (pendB)
    drain(itB, cB, wB);
}

// scalar header fields of the container (Huffman.hpp:344-395)
struct HuffHeader {
  uint64_t n, nchunk, dbs;
  int32_t dict, chunk;
  uint64_t off_bits, off_offs, off_dbs, off_db, off_ddata, off_packed;
};

__global__ void k_write_header(uint8_t *out, HuffHeader h,
                               const u64 *__restrict__ scalars) {
  if (threadIdx.x == 0 && blockIdx.x == 0) {
    const char sig[8] = {'M', 'G', 'X', 'H', 'U', 'F', 'F', 0};
    for (int i = 0; i < 8; i++)
      out[i] = (uint8_t)sig[i];
    *(u64 *)(out + 8) = h.n;
    *(int32_t *)(out + 16) = h.dict;
    *(int32_t *)(out + 20) = h.chunk;
    *(u64 *)(out + 24) = 2 * h.nchunk;
    *(u64 *)(out + h.off_dbs) = h.dbs;
    *(u64 *)(out + h.off_ddata) = scalars[MGX_SC_DDATA];
  }
}

__global__ void k_write_outlier_count(uint8_t *out, u64 off, u64 count) {
  if (threadIdx.x == 0 && blockIdx.x == 0)
    *(u64 *)(out + off) = count;
}

} // namespace

uint64_t mgx_container_fixed_bytes(uint64_t n, uint32_t dict, uint32_t chunk) {
  uint64_t nchunk = (n - 1) / chunk + 1;
  return 32 + 16 * nchunk + 8 + 8ull * (128 + dict) + 8;
}

// Huffman::CompressPrimary + Serialize, Huffman.hpp:157-434: symbols (d_sym),
// histogram (d_freq) and outliers are already in the context.
int mgxk_encode_container(mgx_ctx *ctx, uint64_t n, uint8_t *d_out, uint64_t cap,
                          uint64_t *out_bytes) {
  const uint32_t chunk = ctx->chunk, dict = ctx->dict;
  const uint64_t nchunk = (n - 1) / chunk + 1;
  uint64_t ntiles = (nchunk + ENC_WARPS - 1) / ENC_WARPS;
  HuffHeader h;
  h.n = n;
  h.nchunk = nchunk;
  h.dict = (int32_t)dict;
  h.chunk = (int32_t)chunk;
  h.dbs = 8ull * (128 + dict);
  h.off_bits = 32;
  h.off_offs = 32 + 8 * nchunk;
  h.off_dbs = 32 + 16 * nchunk;
  h.off_db = h.off_dbs + 8;
  h.off_ddata = h.off_db + h.dbs;
  h.off_packed = h.off_ddata + 8;
  if (h.off_packed + 8 > cap) {
    *out_bytes = h.off_packed + 8;
    return MGX_OUTPUT_TOO_LARGE;
  }
  if (((uintptr_t)d_out & 7) != 0) {
    ctx->err = "output buffer must be 8-byte aligned";
    return MGX_FAILURE;
  }
  int rc = mgxk_codebook(ctx, ctx->d_freq, ctx->d_codebook, ctx->d_decodebook);
  if (rc)
    return rc;
  mgx_tick(ctx, MGX_TICK_C_CODEBOOK);
  MGX_CUDA(cudaMemsetAsync(ctx->d_scan_state, 0, (nchunk + 1) * 8, ctx->stream));
  const bool single_pass =
      chunk == 256 || chunk == 512 || chunk == 1024 || chunk == 2048;
  int occ = 1;
  if (single_pass) {
    // sizes, offsets and packed words in one pass, written straight into the
    // container (the capacity is enforced inside the kernel)
    const int R = (int)(chunk / 256);
    const int ew = R == 8 ? EncCfg<8>::EW : EncCfg<4>::EW;
    ntiles = (nchunk + ew - 1) / ew;
    size_t smem = (size_t)ENC_IMGS * ew * (chunk * ENC_CAP_BITS / 32) * 4 + dict * 4;
    auto kern = R == 1 ? k_encode<1> : R == 2 ? k_encode<2> : R == 4 ? k_encode<4>
                                                                     : k_encode<8>;
    MGX_CUDA(cudaFuncSetAttribute(
        kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    MGX_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(
        &occ, kern, (ew + 1) * 32, smem));
    unsigned grid = (unsigned)std::min<uint64_t>(
        ntiles, (uint64_t)std::max(occ, 1) * ctx->num_sms);
    kern<<<grid, (ew + 1) * 32, smem, ctx->stream>>>(
        ctx->d_sym, n, dict, (const u64 *)ctx->d_codebook,
        (u64 *)(d_out + h.off_bits), (u64 *)(d_out + h.off_offs),
        (u64 *)(d_out + h.off_packed), (cap - h.off_packed) / 8,
        (u64 *)ctx->d_scan_state, nchunk, ntiles, (u64 *)ctx->d_scalars);
    ctx->launches++;
    mgx_tick(ctx, MGX_TICK_C_SIZES);
    mgx_tick(ctx, MGX_TICK_C_PACK);
  } else {
    // pass 1: sizes + offsets
    MGX_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(
        &occ, k_chunk_bits, ENC_WARPS * 32, dict));
    unsigned grid1 = (unsigned)std::min<uint64_t>(
        ntiles, (uint64_t)std::max(occ, 1) * ctx->num_sms);
    k_chunk_bits<<<grid1, ENC_WARPS * 32, dict, ctx->stream>>>(
        ctx->d_sym, n, chunk, dict, (const u64 *)ctx->d_codebook,
        (u64 *)ctx->d_chunk_bits, (u64 *)ctx->d_chunk_off,
        (u64 *)ctx->d_scan_state, nchunk, ntiles, (u64 *)ctx->d_scalars);
    ctx->launches++;
    mgx_tick(ctx, MGX_TICK_C_SIZES);
  }
  // one host round trip: packed size, longest code, outlier count, error flags
  MGX_CUDA(cudaMemcpyAsync(ctx->h_scalars, ctx->d_scalars, MGX_SC_COUNT * 8,
                           cudaMemcpyDeviceToHost, ctx->stream));
  MGX_CUDA(cudaStreamSynchronize(ctx->stream));
  if (ctx->h_scalars[MGX_SC_ERR] & 2) {
    ctx->err = "longest Huffman codeword exceeds 56 bits"; // GetCodebook.hpp:111
    return MGX_FAILURE;
  }
  if (ctx->h_scalars[MGX_SC_ERR] & 0xf00ull) {
    ctx->err = "forward pass: tile pipeline timed out (flags " +
               std::to_string(ctx->h_scalars[MGX_SC_ERR]) + ")";
    return MGX_FAILURE;
  }
  if (ctx->h_scalars[MGX_SC_ERR] & ~8ull) {
    ctx->err = "codebook construction failed";
    return MGX_FAILURE;
  }
  const uint64_t ddata = ctx->h_scalars[MGX_SC_DDATA];
  const uint64_t n_out = ctx->h_scalars[MGX_SC_OUTLIERS];
  const uint32_t maxlen = (uint32_t)ctx->h_scalars[MGX_SC_MAXLEN];
  if (n_out > ctx->out_cap) {
    ctx->err = "outlier workspace overflow";
    return MGX_FAILURE + 100; // caller grows the workspace and retries
  }
  const uint64_t off_outc = h.off_packed + 8 * ddata;
  const uint64_t total = off_outc + 8 + 16 * n_out;
  *out_bytes = total;
  if (total > cap || (ctx->h_scalars[MGX_SC_ERR] & 8)) {
    ctx->err = "compressed stream does not fit the output buffer";
    return MGX_OUTPUT_TOO_LARGE;
  }
  // metadata sections
  k_write_header<<<1, 32, 0, ctx->stream>>>(d_out, h, (const u64 *)ctx->d_scalars);
  ctx->launches++;
  MGX_CUDA(cudaMemcpyAsync(d_out + h.off_db, ctx->d_decodebook, h.dbs,
                           cudaMemcpyDeviceToDevice, ctx->stream));
  if (!single_pass) {
    MGX_CUDA(cudaMemcpyAsync(d_out + h.off_bits, ctx->d_chunk_bits, 8 * nchunk,
                             cudaMemcpyDeviceToDevice, ctx->stream));
    MGX_CUDA(cudaMemcpyAsync(d_out + h.off_offs, ctx->d_chunk_off, 8 * nchunk,
                             cudaMemcpyDeviceToDevice, ctx->stream));
    // pass 2: pack
    uint32_t words_cap = (uint32_t)(((uint64_t)chunk * maxlen + 63) / 64 + 1);
    size_t smem = (size_t)ENC_WARPS * words_cap * 8;
    if (smem > 200 * 1024) {
      ctx->err = "huff_block_size too large for the shared-memory packer";
      return MGX_FAILURE;
    }
    MGX_CUDA(cudaFuncSetAttribute(
        k_pack, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    MGX_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_pack,
                                                           ENC_WARPS * 32, smem));
    unsigned grid2 = (unsigned)std::min<uint64_t>(
        ntiles, (uint64_t)std::max(occ, 1) * ctx->num_sms);
    k_pack<<<grid2, ENC_WARPS * 32, smem, ctx->stream>>>(
        ctx->d_sym, n, chunk, (const u64 *)ctx->d_codebook,
        (const u64 *)ctx->d_chunk_off, (u64 *)(d_out + h.off_packed), nchunk,
        words_cap);
    ctx->launches++;
    mgx_tick(ctx, MGX_TICK_C_PACK);
  }
  // outliers in ascending index order (the SERIAL reference's order)
  k_write_outlier_count<<<1, 32, 0, ctx->stream>>>(d_out, off_outc, n_out);
  ctx->launches++;
  if (n_out > (uint64_t)std::numeric_limits<int>::max()) {
    ctx->err = "more than 2^31 outliers: the bound is far too tight for this dictionary";
    return MGX_FAILURE;
  }
  if (n_out) {
    size_t tmp = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, tmp, (const u64 *)ctx->d_out_idx,
                                    (u64 *)(d_out + off_outc + 8),
                                    (const long long *)ctx->d_out_val,
                                    (long long *)(d_out + off_outc + 8 + 8 * n_out),
                                    (int)n_out, 0, 64, ctx->stream);
    if (tmp > ctx->sort_tmp_bytes) {
      if (ctx->d_sort_tmp)
        cudaFree(ctx->d_sort_tmp);
      MGX_CUDA(cudaMalloc(&ctx->d_sort_tmp, tmp));
      ctx->sort_tmp_bytes = tmp;
    }
    cub::DeviceRadixSort::SortPairs(ctx->d_sort_tmp, tmp,
                                    (const u64 *)ctx->d_out_idx,
                                    (u64 *)(d_out + off_outc + 8),
                                    (const long long *)ctx->d_out_val,
                                    (long long *)(d_out + off_outc + 8 + 8 * n_out),
                                    (int)n_out, 0, 64, ctx->stream);
    ctx->launches += 4;
  }
  MGX_CUDA(cudaGetLastError());
  return MGX_SUCCESS;
}

// ============================ decode ==========================================
namespace {

constexpr int LUT_BITS = 12;

// Prefix table for codes of up to LUT_BITS bits: entry = symbol | len << 16
// (len 0 = longer code, fall back to the canonical search).  Built from the
// container's own first[] / entry[] / keys[] tables (Decode.hpp:60-66): at
// length l the codes are first[l] + j, j = 0 .. entry[l+1]-entry[l]-1, and
// decode to keys[entry[l] + j].
__global__ void __launch_bounds__(256)
    k_build_lut(const u64 *__restrict__ db, uint32_t dict,
                uint32_t *__restrict__ lut) {
  const u64 *first = db, *entry = db + 64, *keys = db + 128;
  for (uint32_t i = threadIdx.x; i < (1u << LUT_BITS); i += blockDim.x)
    lut[i] = 0;
  __syncthreads();
  for (int l = 1; l <= LUT_BITS; l++) {
    u64 f = first[l];
    if (f == ~0ull)
      continue;
    u64 e0 = entry[l];
    u64 cnt = entry[l + 1] - e0; // number of codes of exactly this length
    u64 span = 1ull << (LUT_BITS - l);
    for (u64 j = threadIdx.x; j < cnt * span; j += blockDim.x) {
      u64 code = f + j / span;
      u64 r = e0 + (code - f);
      if (r < dict)
        lut[(code << (LUT_BITS - l)) + (j % span)] =
            (uint32_t)keys[r] | ((uint32_t)l << 16);
    }
  }
}

// DecodeFunctor, Decode.hpp:14-130: one thread per chunk, but table driven and
// with 16-byte symbol stores.
__global__ void __launch_bounds__(128)
    k_decode(const u64 *__restrict__ packed, const u64 *__restrict__ meta,
             uint64_t nchunk, uint32_t chunk, uint64_t n,
             const u64 *__restrict__ db, const uint32_t *__restrict__ g_lut,
             uint16_t *__restrict__ sym, uint32_t dict) {
  __shared__ uint32_t lut[1 << LUT_BITS];
  __shared__ u64 s_first[64], s_entry[64];
  for (uint32_t i = threadIdx.x; i < (1u << LUT_BITS); i += blockDim.x)
    lut[i] = g_lut[i];
  if (threadIdx.x < 64) {
    s_first[threadIdx.x] = db[threadIdx.x];
    s_entry[threadIdx.x] = db[64 + threadIdx.x];
  }
  __syncthreads();
  const u64 *keys = db + 128;
  for (uint64_t c = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; c < nchunk;
       c += (uint64_t)gridDim.x * blockDim.x) {
    const u64 *w = packed + meta[nchunk + c];
    const uint64_t total = meta[c];
    const uint64_t base = c * chunk;
    const uint32_t count = (uint32_t)min((uint64_t)chunk, n - base);
    uint16_t *dst = sym + base;
    // 128-bit window: `hi` holds the next bits MSB-aligned, `lo` the refill
    uint64_t pos = 0;      // bits consumed
    uint32_t widx = 0;     // next word to load
    u64 cur = w[0], nxt = (total > 64) ? w[1] : 0;
    widx = 2;
    uint32_t used = 0;     // bits of `cur` already consumed
    uint32_t k = 0;
    uint16_t stage[8];
    const bool aligned16 = ((base * 2) & 15) == 0;
    while (k < count && pos < total) {
      // peek 64 bits starting at the cursor
      u64 window = used ? ((cur << used) | (nxt >> (64 - used))) : cur;
      uint32_t e = lut[(uint32_t)(window >> (64 - LUT_BITS))];
      uint32_t len = e >> 16, s = e & 0xffff;
      if (len == 0) { // long code: canonical search from LUT_BITS+1 upwards
        uint32_t l = LUT_BITS;
        u64 v = window >> (64 - LUT_BITS);
        do {
          ++l;
          v = window >> (64 - l);
        } while (v < s_first[l] && l < 63);
        s = (uint32_t)keys[min(s_entry[l] + v - s_first[l], (u64)dict - 1)];
        len = l;
      }
      stage[k & 7] = (uint16_t)s;
      ++k;
      if ((k & 7) == 0) {
        if (aligned16) {
          uint4 pk;
          pk.x = stage[0] | ((uint32_t)stage[1] << 16);
          pk.y = stage[2] | ((uint32_t)stage[3] << 16);
          pk.z = stage[4] | ((uint32_t)stage[5] << 16);
          pk.w = stage[6] | ((uint32_t)stage[7] << 16);
          *reinterpret_cast<uint4 *>(dst + k - 8) = pk;
        } else {
#pragma unroll
          for (int j = 0; j < 8; j++)
            dst[k - 8 + j] = stage[j];
        }
      }
      pos += len;
      used += len;
      if (used >= 64) {
        used -= 64;
        cur = nxt;
        nxt = ((uint64_t)widx * 64 < total) ? w[widx] : 0;
        ++widx;
      }
    }
    for (uint32_t j = k & ~7u; j < k; j++)
      dst[j] = stage[j & 7];
  }
}

// Multi-symbol prefix table: entry for a 12-bit window = up to four complete
// codes found in it (smooth fields spend 1-2 bits on most symbols, so one
// look-up usually yields 3-4 symbols).
//   four 13-bit symbols in the low 13 bits of the four 16-bit fields (dictionaries
//   up to 8192; unused fields are 0, so an entry can be OR-ed into a register of
//   packed symbols as it is once its spare bits are masked off); spare bits:
//   13..15 = total length bits 0..2, 29 = total length bit 3, 61..63 = count
//   (0 = first code longer than the window).
constexpr u64 MLUT_SYMS = 0x1fff1fff1fff1fffull;
__global__ void __launch_bounds__(256)
    k_build_mlut(const uint32_t *__restrict__ lut1, u64 *__restrict__ mlut) {
  for (uint32_t idx = blockIdx.x * blockDim.x + threadIdx.x; idx < (1u << LUT_BITS);
       idx += gridDim.x * blockDim.x) {
    uint32_t pos = 0, cnt = 0;
    u64 e = 0;
    while (cnt < 4) {
      const uint32_t sub = (idx << pos) & ((1u << LUT_BITS) - 1);
      const uint32_t t = lut1[sub];
      const uint32_t len = t >> 16;
      if (len == 0 || len > LUT_BITS - pos)
        break; // not fully inside the window
      e |= (u64)(t & 0x1fffu) << (16 * cnt);
      pos += len;
      ++cnt;
    }
    mlut[idx] = e | ((u64)(pos & 7u) << 13) | ((u64)((pos >> 3) & 1u) << 29) |
                ((u64)cnt << 61);
  }
}
__device__ __forceinline__ uint32_t mlut_len(u64 e) {
  const uint32_t lo = (uint32_t)e;
  return ((lo >> 13) & 7u) | ((lo >> 26) & 8u);
}

#ifndef MGX_DEC_THREADS
#define MGX_DEC_THREADS 128
#endif
constexpr int DEC_THREADS = MGX_DEC_THREADS;
constexpr int DEC_RING = 20; // symbols per thread (40 B pitch: 2-way bank conflicts; 7 + 8 + 4-store spill)

#ifndef MGX_DEC_SLOTS
#define MGX_DEC_SLOTS 8
#endif
constexpr int DEC_SLOTS = MGX_DEC_SLOTS;
constexpr int DEC_PERIOD = 16; // iterations between top-ups of the stream ring // 64-bit stream words in flight / staged per thread

// 8-byte asynchronous copy global -> shared (LDGSTS): the data never passes through
// a register, so no register scoreboard ties one lane's fetch to another lane's use
__device__ __forceinline__ void cp_async8(void *smem_dst, const void *gsrc) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(
                   (uint32_t)__cvta_generic_to_shared(smem_dst)),
               "l"(gsrc)
               : "memory");
}
__device__ __forceinline__ void cp_async_commit() {
  asm volatile("cp.async.commit_group;" ::: "memory");
}
template <int N> __device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

// DecodeFunctor, Decode.hpp:14-130: one thread per chunk.  Each lane walks its own
// stream at its own pace, so refills are divergent; a refill fed by an ordinary load
// makes the WHOLE warp wait for that load at the next iteration (the scoreboard is
// per warp and register), i.e. one L2 / DRAM latency per decoded group.  The stream
// therefore arrives by cp.async into a per-lane ring of DEC_SLOTS 64-bit words in
// shared memory, topped up at warp-uniform points one period ahead of use (the
// cp.async group counter is per warp too); the bit buffer refills 32 bits at a time
// from that ring, up to four symbols come out of one
// look-up, and symbols leave through a small per-thread ring as 16-byte stores.
template <bool REGS>
__global__ void __launch_bounds__(DEC_THREADS)
    k_decode_multi(const u64 *__restrict__ packed, const u64 *__restrict__ meta,
                   uint64_t nchunk, uint32_t chunk, uint64_t n,
                   const u64 *__restrict__ db, const u64 *__restrict__ g_mlut,
                   uint16_t *__restrict__ sym, uint32_t dict) {
  __shared__ u64 mlut[1 << LUT_BITS];
  __shared__ u64 s_first[64], s_entry[64];
  __shared__ u64 s_words[DEC_SLOTS * DEC_THREADS]; // slot-major: lanes side by side
  // REGS = false: symbols leave through a per-thread ring in shared memory
  __shared__ __align__(8) uint16_t s_ring[REGS ? 4 : DEC_THREADS * DEC_RING];
  uint16_t *ring = s_ring + (REGS ? 0 : threadIdx.x * DEC_RING);
  for (uint32_t i = threadIdx.x; i < (1u << LUT_BITS); i += blockDim.x)
    mlut[i] = g_mlut[i];
  if (threadIdx.x < 64) {
    s_first[threadIdx.x] = db[threadIdx.x];
    s_entry[threadIdx.x] = db[64 + threadIdx.x];
  }
  __syncthreads();
  const u64 *keys = db + 128;
  u64 *words = s_words + threadIdx.x;
  for (uint64_t c = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; c < nchunk;
       c += (uint64_t)gridDim.x * blockDim.x) {
    const u64 *w = packed + meta[nchunk + c];
    const uint64_t total = meta[c];
    const uint32_t nwords = (uint32_t)((total + 63) >> 6);
    const uint64_t base = c * chunk;
    const uint32_t count = (uint32_t)min((uint64_t)chunk, n - base);
    uint16_t *dst = sym + base;
    // The per-lane ring holds stream words [unit / 2, nissued).  Every DEC_PERIOD
    // iterations -- the same iteration for every lane still decoding -- the copies
    // issued one period ago are waited for (long landed) and each lane tops its ring
    // up to DEC_SLOTS words ahead.  A look-up consumes at most LUT_BITS bits, so a lane
    // uses at most DEC_PERIOD * LUT_BITS / 64 = 3 words per period and always finds
    // its bits among the words that landed before the last top-up.
    uint32_t nissued = 0;
    auto fetch = [&]() {
      if (nissued < nwords)
        cp_async8(words + (nissued & (DEC_SLOTS - 1)) * DEC_THREADS, w + nissued);
      ++nissued;
    };
    auto half = [&](uint32_t u) -> uint32_t { // 32-bit unit u of the stream
      const uint32_t *p32 =
          (const uint32_t *)(words + ((u >> 1) & (DEC_SLOTS - 1)) * DEC_THREADS);
      return p32[(u & 1) ^ 1];
    };
    u64 buf;
    uint32_t nbits, unit;
    auto prime = [&](uint32_t w0, uint32_t r) { // bit cursor at word w0, bit r
      cp_async_wait<0>();
      nissued = w0;
#pragma unroll
      for (int i = 0; i < DEC_SLOTS; i++)
        fetch();
      cp_async_commit();
      cp_async_wait<0>();
      buf = (((u64)half(2 * w0) << 32) | half(2 * w0 + 1)) << r;
      nbits = 64 - r;
      unit = 2 * w0 + 2;
    };
    prime(0, 0);
    uint32_t it = 0;
    // Two look-ups per iteration: with more than 32 bits in the buffer at the top,
    // both windows are complete (2 * LUT_BITS <= 32 - 8) and one 32-bit refill
    // restores the invariant.  Two symbol sinks, chosen by the stream's bits per
    // symbol (measured at 1025^3: 1.61 / 1.71 ms at 1.5 bits, 4.41 / 4.06 ms at 7.4):
    // REGS: a0 | a1 hold the 8 symbols of the next 16-byte store, a2 what spills
    // over; a look-up's four 16-bit fields are OR-ed in at symbol position k with
    // two 64-bit shifts -- no shared-memory traffic for symbols (8 two-byte stores
    // per iteration with bank conflicts otherwise), more ALU work;
    // ring: 20 entries per thread in shared memory, flushed 8 at a time.
    uint32_t produced = 0, written = 0;
    u64 a0 = 0, a1 = 0, a2 = 0;
    auto insert = [&](u64 v, uint32_t k) { // k = 0..7 (REGS) / 0..11 (ring)
      if constexpr (REGS) {
        const uint32_t b = (k & 3u) << 4;
        const u64 lo = v << b;
        const u64 hi = (v >> 1) >> (63u - b);
        if (k & 4u) {
          a1 |= lo;
          a2 |= hi;
        } else {
          a0 |= lo;
          a1 |= hi;
        }
      } else {
        uint16_t *r = ring + k;
        r[0] = (uint16_t)v;
        r[1] = (uint16_t)(v >> 16);
        r[2] = (uint16_t)(v >> 32);
        r[3] = (uint16_t)(v >> 48);
      }
    };
    auto flush8 = [&]() {
      if constexpr (REGS) {
        *reinterpret_cast<uint4 *>(dst + written) = make_uint4(
            (uint32_t)a0, (uint32_t)(a0 >> 32), (uint32_t)a1, (uint32_t)(a1 >> 32));
        a0 = a2;
        a1 = 0;
        a2 = 0;
      } else {
        const uint2 x0 = *reinterpret_cast<const uint2 *>(ring);
        const uint2 x1 = *reinterpret_cast<const uint2 *>(ring + 4);
        *reinterpret_cast<uint4 *>(dst + written) = make_uint4(x0.x, x0.y, x1.x, x1.y);
        *reinterpret_cast<uint2 *>(ring) = *reinterpret_cast<const uint2 *>(ring + 8);
        *reinterpret_cast<uint2 *>(ring + 4) = *reinterpret_cast<const uint2 *>(ring + 12);
      }
      written += 8;
    };
    while (produced < count) {
      if ((it++ & (DEC_PERIOD / 2 - 1)) == 0) {
        cp_async_wait<0>();
#pragma unroll
        for (int i = 0; i < DEC_PERIOD * LUT_BITS / 64; i++)
          if (nissued < (unit >> 1) + DEC_SLOTS)
            fetch();
        cp_async_commit();
      }
      const u64 e1 = mlut[(uint32_t)(buf >> (64 - LUT_BITS))];
      const uint32_t c1 = (uint32_t)(e1 >> 61), len1 = mlut_len(e1);
      if (__builtin_expect(c1 == 0, 0)) {
        // first code is longer than the window: canonical search on the raw
        // words (Decode.hpp:60-66), then re-prime the bit buffer
        uint64_t pos = 32ull * unit - nbits; // bits consumed so far
        const uint32_t wi = (uint32_t)(pos >> 6), used = (uint32_t)(pos & 63);
        const u64 cur = wi < nwords ? w[wi] : 0;
        const u64 nxt = wi + 1 < nwords ? w[wi + 1] : 0;
        const u64 window = used ? ((cur << used) | (nxt >> (64 - used))) : cur;
        uint32_t l = LUT_BITS;
        u64 v;
        do {
          ++l;
          v = window >> (64 - l);
        } while (v < s_first[l] && l < 63);
        insert(keys[min(s_entry[l] + v - s_first[l], (u64)dict - 1)] & 0xffffu,
               produced - written);
        ++produced;
        pos += l;
        prime((uint32_t)(pos >> 6), (uint32_t)(pos & 63));
      } else {
        const u64 b2 = buf << len1;
        const u64 e2 = mlut[(uint32_t)(b2 >> (64 - LUT_BITS))];
        insert(e1 & MLUT_SYMS, produced - written);
        produced += c1;
        if (REGS && produced - written >= 8)
          flush8(); // three registers hold 12 symbols, the ring 20
        // a second window that starts with a long code has count 0 and length 0:
        // nothing is consumed and the next iteration takes the branch above
        const uint32_t c2 = (uint32_t)(e2 >> 61), len2 = mlut_len(e2);
        insert(e2 & MLUT_SYMS, produced - written);
        produced = min(produced + c2, count);
        buf = b2 << len2;
        nbits -= len1 + len2;
      }
      if (nbits <= 32) {
        buf |= (u64)half(unit) << (32 - nbits);
        nbits += 32;
        ++unit;
      }
      if (produced - written >= 8)
        flush8();
    }
    for (uint32_t j = 0; written + j < produced; j++)
      dst[written + j] = REGS ? (uint16_t)((j < 4 ? a0 >> (16 * j) : a1 >> (16 * (j - 4))))
                              : ring[j];
    cp_async_wait<0>(); // the word ring is reused by this thread's next chunk
  }
}

// out[0] = status (0 ok, 1 bad signature, 2 truncated), then n, dict, chunk,
// huffmeta, decodebook bytes, packed words, outlier count
__global__ void k_parse_header(const uint8_t *in, uint64_t bytes, u64 *out) {
  const char sig[7] = {'M', 'G', 'X', 'H', 'U', 'F', 'F'};
  for (int i = 0; i < 9; i++)
    out[i] = 0;
  for (int i = 0; i < 7; i++)
    if (in[i] != (uint8_t)sig[i]) {
      out[0] = 1;
      return;
    }
  const u64 n = *(const u64 *)(in + 8);
  const int32_t dict = *(const int32_t *)(in + 16);
  const int32_t chunk = *(const int32_t *)(in + 20);
  const u64 huffmeta = *(const u64 *)(in + 24);
  out[1] = n;
  out[2] = (u64)(int64_t)dict;
  out[3] = (u64)(int64_t)chunk;
  out[4] = huffmeta;
  if (chunk <= 0 || n == 0) {
    out[0] = 2;
    return;
  }
  const u64 nchunk = (n - 1) / (u64)chunk + 1;
  const u64 off_dbs = 32 + 16 * nchunk;
  if (off_dbs + 8 > bytes) {
    out[0] = 2;
    return;
  }
  const u64 dbs = *(const u64 *)(in + off_dbs);
  out[5] = dbs;
  const u64 off_ddata = off_dbs + 8 + dbs;
  if (dbs > bytes || off_ddata + 8 > bytes) {
    out[0] = 2;
    return;
  }
  const u64 ddata = *(const u64 *)(in + off_ddata);
  out[6] = ddata;
  const u64 off_outc = off_ddata + 8 + 8 * ddata;
  if (ddata > bytes / 8 || off_outc + 8 > bytes) {
    out[0] = 2;
    return;
  }
  out[7] = *(const u64 *)(in + off_outc);
}

// Payload checks the decoder relies on (only the MGARD header carries a CRC; a
// corrupt or hostile container must not cause out-of-bounds accesses).  Driven by
// the parsed header in out[0..7]; flags go to out[8]:
//   1 chunk metadata points outside the packed words     2 outlier index >= n
//   4 outlier indices not ascending (legal: the reference's CUDA / OpenMP back ends
//     append with atomics, OutlierSeparator.hpp:66-97)   8 decodebook key >= dict
__global__ void __launch_bounds__(256)
    k_validate_container(const uint8_t *__restrict__ in, uint64_t bytes, uint64_t n_plan,
                         u64 *__restrict__ out) {
  if (out[0] != 0)
    return;
  const u64 n = out[1], dict = out[2], chunk = out[3], dbs = out[5], ddata = out[6],
            n_out = out[7];
  if (n != n_plan || dbs != 8 * (128 + dict))
    return; // the host rejects these before anything is decoded
  const u64 nchunk = (n - 1) / chunk + 1;
  const u64 *meta = (const u64 *)(in + 32);
  const u64 *keys = (const u64 *)(in + 32 + 16 * nchunk + 8) + 128;
  const u64 off_outc = 32 + 16 * nchunk + 8 + dbs + 8 + 8 * ddata;
  unsigned flags = 0;
  const u64 tid = blockIdx.x * (u64)blockDim.x + threadIdx.x, nth = (u64)gridDim.x * blockDim.x;
  for (u64 c = tid; c < nchunk; c += nth) {
    const u64 bits = meta[c], off = meta[nchunk + c];
    if (bits > chunk * 56 || off > ddata || off + ((bits + 63) >> 6) > ddata)
      flags |= 1;
  }
  for (u64 i = tid; i < dict; i += nth)
    if (keys[i] >= dict)
      flags |= 8;
  if (n_out > n || off_outc + 8 + 16 * n_out > bytes) {
    flags |= 2;
  } else {
    const u64 *idx = (const u64 *)(in + off_outc + 8);
    for (u64 i = tid; i < n_out; i += nth) {
      const u64 v = idx[i];
      if (v >= n)
        flags |= 2;
      if (i + 1 < n_out && idx[i + 1] <= v)
        flags |= 4;
    }
  }
  if (flags)
    atomicOr((unsigned long long *)&out[8], (unsigned long long)flags);
}

struct ParsedHeader {
  uint64_t n, nchunk, dbs, ddata, n_out;
  int32_t dict, chunk;
  uint64_t off_meta, off_db, off_packed, off_oidx, off_oval;
};

} // namespace

// Huffman::Deserialize + DecompressPrimary, Huffman.hpp:453-551.  Symbols land
// in ctx->d_sym; the outlier sections are returned as views into d_in.
int mgxk_decode_container(mgx_ctx *ctx, const uint8_t *d_in, uint64_t bytes,
                          uint64_t n, const uint64_t **d_oidx,
                          const int64_t **d_oval, uint64_t *n_out) {
  if (bytes < 40 || ((uintptr_t)d_in & 7) != 0) {
    ctx->err = "bad Huffman container (size/alignment)";
    return MGX_FAILURE;
  }
  // the section sizes chain through the container (Huffman.hpp:453-518): one
  // single-thread kernel follows the chain, one read-back brings it to the host
  k_parse_header<<<1, 1, 0, ctx->stream>>>(d_in, bytes, (u64 *)ctx->d_scalars);
  k_validate_container<<<ctx->num_sms * 2, 256, 0, ctx->stream>>>(d_in, bytes, n,
                                                                  (u64 *)ctx->d_scalars);
  ctx->launches += 2;
  MGX_CUDA(cudaMemcpyAsync(ctx->h_scalars, ctx->d_scalars, MGX_SC_COUNT * 8,
                           cudaMemcpyDeviceToHost, ctx->stream));
  MGX_CUDA(cudaStreamSynchronize(ctx->stream));
  const uint64_t *hs = ctx->h_scalars;
  if (hs[0] == 1) {
    ctx->err = "Huffman signature mismatch."; // Huffman.hpp:459-461
    return MGX_FAILURE;
  }
  if (hs[0] != 0) {
    ctx->err = "truncated or inconsistent Huffman container";
    return MGX_FAILURE;
  }
  ParsedHeader p;
  p.n = hs[1];
  p.dict = (int32_t)hs[2];
  p.chunk = (int32_t)hs[3];
  const uint64_t huffmeta = hs[4];
  p.dbs = hs[5];
  p.ddata = hs[6];
  p.n_out = hs[7];
  if (p.n != n || p.chunk <= 0 || p.dict <= 0 || p.dict > 65536) {
    ctx->err = "Huffman container does not match the plan";
    return MGX_FAILURE;
  }
  p.nchunk = (p.n - 1) / (uint64_t)p.chunk + 1;
  if (huffmeta != 2 * p.nchunk || p.dbs != 8ull * (128 + (uint64_t)p.dict))
    return MGX_FAILURE;
  p.off_meta = 32;
  p.off_db = 32 + 16 * p.nchunk + 8;
  p.off_packed = p.off_db + p.dbs + 8;
  const uint64_t off_outc = p.off_packed + 8 * p.ddata;
  p.off_oidx = off_outc + 8;
  p.off_oval = p.off_oidx + 8 * p.n_out;
  if (p.off_oval + 8 * p.n_out > bytes)
    return MGX_FAILURE;
  if (hs[8] & (1 | 2 | 8)) {
    ctx->err = "corrupt Huffman container (chunk offsets, outlier indices or decodebook "
               "out of range)";
    return MGX_FAILURE;
  }
  *n_out = p.n_out;
  *d_oidx = (const uint64_t *)(d_in + p.off_oidx);
  *d_oval = (const int64_t *)(d_in + p.off_oval);
  if (hs[8] & 4) {
    // outlier list in append order (a stream of the reference's CUDA / OpenMP back
    // end): the fused inverse looks outliers up by binary search, so sort it
    if (p.n_out > (uint64_t)std::numeric_limits<int>::max()) {
      ctx->err = "more than 2^31 unsorted outliers";
      return MGX_FAILURE;
    }
    if (ctx->out_cap < p.n_out) {
      if (ctx->d_out_idx)
        cudaFree(ctx->d_out_idx);
      if (ctx->d_out_val)
        cudaFree(ctx->d_out_val);
      ctx->d_out_idx = nullptr;
      ctx->d_out_val = nullptr;
      ctx->out_cap = 0;
      MGX_CUDA(cudaMalloc(&ctx->d_out_idx, p.n_out * 8));
      MGX_CUDA(cudaMalloc(&ctx->d_out_val, p.n_out * 8));
      ctx->out_cap = p.n_out;
    }
    size_t tmp = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, tmp, (const u64 *)*d_oidx, (u64 *)ctx->d_out_idx,
                                    (const long long *)*d_oval,
                                    (long long *)ctx->d_out_val, (int)p.n_out, 0, 64,
                                    ctx->stream);
    if (tmp > ctx->sort_tmp_bytes) {
      if (ctx->d_sort_tmp)
        cudaFree(ctx->d_sort_tmp);
      ctx->d_sort_tmp = nullptr;
      ctx->sort_tmp_bytes = 0;
      MGX_CUDA(cudaMalloc(&ctx->d_sort_tmp, tmp));
      ctx->sort_tmp_bytes = tmp;
    }
    cub::DeviceRadixSort::SortPairs(ctx->d_sort_tmp, tmp, (const u64 *)*d_oidx,
                                    (u64 *)ctx->d_out_idx, (const long long *)*d_oval,
                                    (long long *)ctx->d_out_val, (int)p.n_out, 0, 64,
                                    ctx->stream);
    ctx->launches += 4;
    *d_oidx = ctx->d_out_idx;
    *d_oval = ctx->d_out_val;
  }

  const u64 *db = (const u64 *)(d_in + p.off_db);
  k_build_lut<<<1, 256, 0, ctx->stream>>>(db, (uint32_t)p.dict,
                                          (uint32_t *)ctx->d_lut);
  // 16-byte symbol stores need chunks that start on 16-byte boundaries
  if (p.dict <= 8192 && p.chunk % 8 == 0) {
    u64 *mlut = (u64 *)((uint8_t *)ctx->d_lut + 4096 * 4);
    k_build_mlut<<<16, 256, 0, ctx->stream>>>((const uint32_t *)ctx->d_lut, mlut);
    // symbol sink by the stream's density: above ~3 bits per symbol the look-ups
    // yield few symbols each and the shared-memory ring costs more than it saves
    const bool regs = p.ddata * 64 > p.n * 3;
    auto kern = regs ? k_decode_multi<true> : k_decode_multi<false>;
    kern<<<mgx_grid_for(p.nchunk, DEC_THREADS, ctx->num_sms * 16), DEC_THREADS, 0,
           ctx->stream>>>(
        (const u64 *)(d_in + p.off_packed), (const u64 *)(d_in + p.off_meta),
        p.nchunk, (uint32_t)p.chunk, p.n, db, mlut, ctx->d_sym, (uint32_t)p.dict);
    ctx->launches += 3;
  } else {
    k_decode<<<mgx_grid_for(p.nchunk, 128, ctx->num_sms * 16), 128, 0,
               ctx->stream>>>((const u64 *)(d_in + p.off_packed),
                              (const u64 *)(d_in + p.off_meta), p.nchunk,
                              (uint32_t)p.chunk, p.n, db,
                              (const uint32_t *)ctx->d_lut, ctx->d_sym, (uint32_t)p.dict);
    ctx->launches += 2;
  }
  MGX_CUDA(cudaGetLastError());
  return MGX_SUCCESS;
}
