// mgard_b200 internal declarations (not part of the C ABI).
#pragma once
#include <cuda_runtime.h>

#include <cstdint>
#include <string>
#include <vector>

#include "../../include/mgard_b200.h"

#define MGX_MAX_D 5
#define MGX_MAX_L 32 // level index 0..31 (2^31+1 nodes per dimension)

// ---- host-side hierarchy (reference Hierarchy/Hierarchy.hpp:22-268,277-496)
struct MgxLevels {
  int D = 0;                       // user dimensionality
  int L = 0;                       // l_target
  uint64_t shape[MGX_MAX_D];       // padded to 5 dims with leading 1s
  uint64_t stride[MGX_MAX_D];      // element strides of the dense array
  uint64_t n[MGX_MAX_L][MGX_MAX_D]; // level shapes, padded dims stay 1
  uint64_t N = 0;
};

// Offsets (in elements of T) into the single device table buffer.
struct MgxTableOffsets {
  uint32_t dist[MGX_MAX_L][MGX_MAX_D];
  uint32_t ratio[MGX_MAX_L][MGX_MAX_D];
  uint32_t am[MGX_MAX_L][MGX_MAX_D];
  uint32_t bm[MGX_MAX_L][MGX_MAX_D];
  uint32_t abr[MGX_MAX_L][MGX_MAX_D]; // am[i] / bm[i]
};

template <typename T> struct MgxHostTables {
  std::vector<T> buf;       // all per-level per-dim arrays back to back
  MgxTableOffsets off;
  T vol[MGX_MAX_L];         // sqrt(prod_d level volume), lwqzk volume factor
  T volr[MGX_MAX_L];        // same from reciprocal volumes (dequantize)
  bool ratios_half = true;  // every interpolation ratio is exactly 0.5
};

int mgx_build_levels(int D, const uint64_t *shape, uint64_t max_level,
                     MgxLevels *lv);
template <typename T>
void mgx_build_tables(const MgxLevels &lv, MgxHostTables<T> *tb,
                      const void *const *coords = nullptr, bool normalize = true);

// ---- context ---------------------------------------------------------------
struct mgx_ctx {
  int device = 0;
  cudaStream_t stream = nullptr;
  bool own_stream = false;
  int num_sms = 148;
  std::string err;
  uint64_t launches = 0;
  bool legacy_levels = false; // MGX_LEGACY_LEVELS=1: gather kernels of mgx_refactor.cu

  bool planned = false;
  int dtype = MGX_F32;
  uint32_t dict = 8192, chunk = 1024;
  MgxLevels lv;
  MgxHostTables<float> tbf;
  MgxHostTables<double> tbd;
  void *d_tables = nullptr; // T[] copy of tb.buf
  double *d_rcp = nullptr;  // 1.0 / tb.buf[k] in double, same offsets
  bool fast_div = true;
  bool nonuniform = false;  // planned with caller-supplied node coordinates
  bool normalize = true;    // Config::normalize_coordinates of the plan
  int fw_minb = 0;          // MGX_FW_MINB=2|3: force the k_level_forward variant (0 = by dtype)
  bool no_sym_recompose = false; // MGX_NO_SYM_RECOMPOSE=1: dequantise to floats first
  bool no_ws = false;       // MGX_NO_WS=1: single-warp Thomas kernels only (A/B)
  bool no_part = false;     // MGX_NO_PART=1: no partitioned Thomas solves (A/B)
  bool inv_tiled = false;   // MGX_INV_TILED=1: 3-D inverse on the three-level tile kernel (A/B)
  bool no_tma = false;      // MGX_NO_TMA=1: 3-D hierarchical forward pass on the LDGSTS kernel (A/B)

  // workspaces (allocated on first use, sized for the plan)
  void *d_w = nullptr;      // N elements of T (level workspace)
  void *d_c0 = nullptr;     // correction ping buffer
  void *d_c1 = nullptr;     // correction pong buffer
  void *d_v = nullptr;      // N elements of T (compress/decompress scratch)
  uint16_t *d_sym = nullptr; // N 16-bit symbols
  uint32_t *d_freq = nullptr;
  uint64_t *d_codebook = nullptr;
  uint64_t *d_decodebook = nullptr; // 128 + dict
  uint64_t *d_chunk_bits = nullptr; // nchunk
  uint64_t *d_chunk_off = nullptr;  // nchunk + 1
  uint64_t *d_scan_state = nullptr; // decoupled look-back descriptors
  uint64_t *d_out_idx = nullptr;    // outlier positions
  int64_t *d_out_val = nullptr;     // outlier values
  uint64_t out_cap = 0;
  uint64_t *d_scalars = nullptr;    // small device scalars (see MGX_SC_*)
  uint64_t *h_scalars = nullptr;    // pinned mirror
  double *d_norm_part = nullptr;    // per-block partial sums of the norm (1024)
  void *d_lut = nullptr;            // decode look-up table
  void *d_coarse0 = nullptr, *d_coarse1 = nullptr; // fused inverse: coarse lattice
  size_t coarse_bytes = 0;
  void *d_lvl[2] = {nullptr, nullptr}; // 3-D level-by-level inverse: compact levels L-2, L-1
  size_t lvl_bytes[2] = {0, 0};
  void *d_sort_tmp = nullptr;
  void *d_cb_ws = nullptr;          // codebook workspace when it exceeds shared memory
  size_t cb_ws_bytes = 0;
  size_t sort_tmp_bytes = 0;
  uint64_t *d_out_idx2 = nullptr;
  int64_t *d_out_val2 = nullptr;
  void *h_stage = nullptr;          // pinned staging for the *_host calls
  size_t h_stage_bytes = 0;
  void *d_stage_in = nullptr, *d_stage_out = nullptr;
  size_t d_stage_in_bytes = 0, d_stage_out_bytes = 0;

  // optional stage timing (mgx_set_timing): events on ctx->stream
  bool timing = false;
  cudaEvent_t ev[MGX_TICK_COUNT] = {};
  bool ev_set[MGX_TICK_COUNT] = {};
};

static inline void mgx_tick(mgx_ctx *c, int id) {
  if (c->timing) {
    cudaEventRecord(c->ev[id], c->stream);
    c->ev_set[id] = true;
  }
}

// indices into d_scalars / h_scalars (uint64 each; doubles are bit-cast)
enum {
  MGX_SC_OUTLIERS = 0, // outlier counter
  MGX_SC_DDATA = 1,    // packed words
  MGX_SC_NORM = 2,     // double: norm partial (max or sum of squares)
  MGX_SC_MAXLEN = 3,   // longest code length
  MGX_SC_TICKET = 4,   // dynamic block id for the look-back scan
  MGX_SC_ERR = 5,      // device-side error flags
  MGX_SC_NZ = 6,       // number of non-zero symbols
  MGX_SC_REDO = 7,     // partitioned Thomas solves: chains redone sequentially
  MGX_SC_TILE = 8,     // k_fwd3: next tile of the device-wide tile list
  MGX_SC_COUNT = 16
};

#define MGX_CUDA(call)                                                         \
  do {                                                                         \
    cudaError_t e_ = (call);                                                   \
    if (e_ != cudaSuccess) {                                                   \
      ctx->err = std::string(#call) + ": " + cudaGetErrorString(e_);           \
      return MGX_FAILURE;                                                      \
    }                                                                          \
  } while (0)

static inline bool mgx_is_f32(const mgx_ctx *c) { return c->dtype == MGX_F32; }
static inline size_t mgx_elem(const mgx_ctx *c) {
  return c->dtype == MGX_F32 ? 4 : 8;
}

// ---- kernel launchers (mgx_refactor.cu / mgx_quant.cu / mgx_huffman.cu) ----
int mgxk_decompose(mgx_ctx *ctx, void *d_data, int orthogonal);
int mgxk_recompose(mgx_ctx *ctx, void *d_data, int orthogonal);
int mgxk_norm_partial(mgx_ctx *ctx, const void *d_in, int linf);
int mgxk_quantize(mgx_ctx *ctx, const void *d_coeff, int rel, double tol,
                  double s, double norm, int64_t *d_codes);
int mgxk_dequantize(mgx_ctx *ctx, const int64_t *d_codes, int rel, double tol,
                    double s, double norm, void *d_coeff);
// symbols (u16) <-> coefficients, used by the fused compress / decompress
int mgxk_quantize_sym(mgx_ctx *ctx, const void *d_coeff, int rel, double tol,
                      double s, const double *d_norm_or_null, double norm);
int mgxk_dequantize_sym_out(mgx_ctx *ctx, int rel, double tol, double s,
                            double norm, void *d_coeff, const uint64_t *d_oidx,
                            const int64_t *d_oval, uint64_t n_out);
// fused hierarchical (no correction, D <= 3) forward pass: input -> symbols +
// histogram + outliers, quantiser derived on the device from the norm scalar
int mgxk_hier_forward(mgx_ctx *ctx, const void *d_in, int rel, double tol,
                      bool norm_on_device, double norm);
// fused inverse: symbols (ctx->d_sym) + sorted outlier list -> field.  Returns
// MGX_FAILURE + 101 when the shape has too few levels for it (caller falls
// back to dequantise + level-by-level recompose).
int mgxk_hier_inverse(mgx_ctx *ctx, int rel, double tol, double norm, void *d_out,
                      const uint64_t *d_oidx, const int64_t *d_oval,
                      uint64_t n_out);
int mgxk_recompose_compact(mgx_ctx *ctx, void *d_v, void *d_w, int levels);

int mgxk_codes_to_symbols(mgx_ctx *ctx, const int64_t *d_codes, uint64_t n);
int mgxk_histogram_reset(mgx_ctx *ctx);
int mgxk_codebook(mgx_ctx *ctx, const uint32_t *d_freq, uint64_t *d_codebook,
                  uint64_t *d_decodebook);
// symbols + codebook + outliers -> container; n symbols in ctx->d_sym
int mgxk_encode_container(mgx_ctx *ctx, uint64_t n, uint8_t *d_out,
                          uint64_t cap, uint64_t *out_bytes);
// container -> symbols (ctx->d_sym) + outlier list (views into the container)
int mgxk_decode_container(mgx_ctx *ctx, const uint8_t *d_in, uint64_t bytes,
                          uint64_t n, const uint64_t **d_oidx,
                          const int64_t **d_oval, uint64_t *n_out);
int mgxk_symbols_to_codes(mgx_ctx *ctx, uint64_t n, int64_t *d_codes,
                          const uint64_t *d_oidx, const int64_t *d_oval,
                          uint64_t n_out);

int mgx_ensure_workspaces(mgx_ctx *ctx, bool need_w, bool need_v,
                          bool need_c = false);

// row-streaming level kernels for the orthogonal-projection path (mgx_level.cu);
// all out of place: source and destination must differ
int mgxk_decompose_rows(mgx_ctx *ctx, const void *d_src, void *d_coeff);
int mgxk_decompose_quantize_rows(mgx_ctx *ctx, const void *d_src, int rel,
                                 double tol, double s, double norm);
int mgxk_recompose_rows(mgx_ctx *ctx, const void *d_coeff, void *d_dst);
int mgxk_recompose_sym_rows(mgx_ctx *ctx, int rel, double tol, double s,
                            double norm, const void *d_coeff, int sym_from,
                            void *d_dst);
// dequantise only the level-`level` box [0, n_level)^D (+ all outliers, which the
// caller knows to lie inside it) into d_coeff; *max_level = highest level that
// holds an outlier (mgxk_outlier_max_level)
int mgxk_outlier_max_level(mgx_ctx *ctx, const uint64_t *d_oidx, uint64_t n_out,
                           int *max_level);
int mgxk_dequantize_box(mgx_ctx *ctx, int level, int rel, double tol, double s,
                        double norm, void *d_coeff, const uint64_t *d_oidx,
                        const int64_t *d_oval, uint64_t n_out);
// per-level quantiser (reciprocal for quantise) and volume factor, as doubles
// holding the T-precision values (CalcQuantizers, LinearQuantization.hpp:234-297)
void mgx_level_quantizers(const mgx_ctx *ctx, int rel, double tol, double s,
                          double norm, bool reciprocal, double *q, double *vol);
