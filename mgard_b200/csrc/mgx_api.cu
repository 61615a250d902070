// C ABI of mgard_b200 (see include/mgard_b200.h for the reference interface
// each entry point replaces).
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <new>
#include <vector>

#include "mgx_internal.h"

uint64_t mgx_container_fixed_bytes(uint64_t n, uint32_t dict, uint32_t chunk);

namespace {

template <typename P> void free_dev(P *&p) {
  if (p) {
    cudaFree(p);
    p = nullptr;
  }
}

void release_plan(mgx_ctx *ctx) {
  free_dev(ctx->d_tables);
  free_dev(ctx->d_rcp);
  free_dev(ctx->d_w);
  free_dev(ctx->d_c0);
  free_dev(ctx->d_c1);
  free_dev(ctx->d_v);
  free_dev(ctx->d_sym);
  free_dev(ctx->d_freq);
  free_dev(ctx->d_codebook);
  free_dev(ctx->d_decodebook);
  free_dev(ctx->d_chunk_bits);
  free_dev(ctx->d_chunk_off);
  free_dev(ctx->d_scan_state);
  free_dev(ctx->d_out_idx);
  free_dev(ctx->d_out_val);
  free_dev(ctx->d_lut);
  free_dev(ctx->d_coarse0);
  free_dev(ctx->d_coarse1);
  ctx->coarse_bytes = 0;
  for (int i = 0; i < 2; i++) {
    free_dev(ctx->d_lvl[i]);
    ctx->lvl_bytes[i] = 0;
  }
  ctx->out_cap = 0;
  ctx->planned = false;
}

int grow_outliers(mgx_ctx *ctx, uint64_t want) {
  free_dev(ctx->d_out_idx);
  free_dev(ctx->d_out_val);
  MGX_CUDA(cudaMalloc(&ctx->d_out_idx, want * 8));
  MGX_CUDA(cudaMalloc(&ctx->d_out_val, want * 8));
  ctx->out_cap = want;
  return MGX_SUCCESS;
}

bool need_orthogonal(const mgx_ctx *ctx, double s) {
  // infer_orthogonal_projection, CompressionLowLevel/Compressor.h:38-40
  return !(std::isinf(s) && ctx->lv.D <= 3);
}

} // namespace

int mgx_ensure_workspaces(mgx_ctx *ctx, bool need_w, bool need_v, bool need_c) {
  const size_t es = mgx_elem(ctx);
  const MgxLevels &lv = ctx->lv;
  if (need_w && !ctx->d_w)
    MGX_CUDA(cudaMalloc(&ctx->d_w, lv.N * es));
  if ((need_w || need_c) && !ctx->d_c0) {
    // correction ping/pong buffers: outputs of the first and second
    // restriction sweeps at the finest level are the largest ones
    uint64_t c0 = 1, c1 = 1;
    const int pad = MGX_MAX_D - lv.D;
    if (lv.L >= 1) {
      for (int d = 0; d < MGX_MAX_D; d++) {
        uint64_t nf = lv.n[lv.L][d], nc = lv.n[lv.L - 1][d];
        c0 *= (d == MGX_MAX_D - 1) ? nc : nf;
        c1 *= (d >= MGX_MAX_D - 2 && d >= pad) ? nc : nf;
      }
    }
    MGX_CUDA(cudaMalloc(&ctx->d_c0, c0 * es));
    MGX_CUDA(cudaMalloc(&ctx->d_c1, c1 * es));
  }
  if (need_v && !ctx->d_v)
    MGX_CUDA(cudaMalloc(&ctx->d_v, lv.N * es));
  return MGX_SUCCESS;
}

extern "C" {

const char *mgx_version(void) { return "mgard_b200 0.1 (sm_100a)"; }

int mgx_create(mgx_ctx **out, int device, void *cuda_stream) {
  if (!out)
    return MGX_FAILURE;
  *out = nullptr;
  int count = 0;
  if (cudaGetDeviceCount(&count) != cudaSuccess || count <= 0 ||
      device >= count || device < 0)
    return MGX_BACKEND_UNAVAILABLE; // no CPU fallback by design
  if (cudaSetDevice(device) != cudaSuccess)
    return MGX_BACKEND_UNAVAILABLE;
  mgx_ctx *ctx = new (std::nothrow) mgx_ctx();
  if (!ctx)
    return MGX_FAILURE;
  ctx->device = device;
  {
    // A/B switch for measurements and tests: the superseded gather kernels
    const char *e = getenv("MGX_LEGACY_LEVELS");
    ctx->legacy_levels = e && e[0] == '1';
    e = getenv("MGX_FW_MINB");
    if (e && (e[0] == '2' || e[0] == '3'))
      ctx->fw_minb = e[0] - '0';
    e = getenv("MGX_NO_SYM_RECOMPOSE");
    ctx->no_sym_recompose = e && e[0] == '1';
    e = getenv("MGX_NO_WS");
    ctx->no_ws = e && e[0] == '1';
    e = getenv("MGX_NO_TMA");
    ctx->no_tma = e && e[0] == '1';
    e = getenv("MGX_NO_PART");
    ctx->no_part = e && e[0] == '1';
    e = getenv("MGX_INV_TILED");
    ctx->inv_tiled = e && e[0] == '1';
  }
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, device) == cudaSuccess)
    ctx->num_sms = prop.multiProcessorCount;
  if (cuda_stream) {
    ctx->stream = (cudaStream_t)cuda_stream;
  } else {
    if (cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) !=
        cudaSuccess) {
      delete ctx;
      return MGX_FAILURE;
    }
    ctx->own_stream = true;
  }
  if (cudaMalloc(&ctx->d_norm_part, 1024 * 8) != cudaSuccess ||
      cudaMalloc(&ctx->d_scalars, MGX_SC_COUNT * 8) != cudaSuccess ||
      cudaMallocHost(&ctx->h_scalars, MGX_SC_COUNT * 8) != cudaSuccess) {
    delete ctx;
    return MGX_FAILURE;
  }
  cudaMemsetAsync(ctx->d_scalars, 0, MGX_SC_COUNT * 8, ctx->stream);
  *out = ctx;
  return MGX_SUCCESS;
}

void mgx_destroy(mgx_ctx *ctx) {
  if (!ctx)
    return;
  cudaSetDevice(ctx->device);
  cudaStreamSynchronize(ctx->stream);
  for (int i = 0; i < MGX_TICK_COUNT; i++)
    if (ctx->ev[i])
      cudaEventDestroy(ctx->ev[i]);
  release_plan(ctx);
  free_dev(ctx->d_scalars);
  free_dev(ctx->d_norm_part);
  free_dev(ctx->d_sort_tmp);
  free_dev(ctx->d_cb_ws);
  free_dev(ctx->d_stage_in);
  free_dev(ctx->d_stage_out);
  if (ctx->h_scalars)
    cudaFreeHost(ctx->h_scalars);
  if (ctx->h_stage)
    cudaFreeHost(ctx->h_stage);
  if (ctx->own_stream)
    cudaStreamDestroy(ctx->stream);
  delete ctx;
}

const char *mgx_last_error(const mgx_ctx *ctx) {
  return ctx ? ctx->err.c_str() : "null context";
}

uint64_t mgx_launch_count(const mgx_ctx *ctx) { return ctx ? ctx->launches : 0; }

// chains the partitioned Thomas solves had to redo sequentially since the device
// scalars were last reset (start of mgx_compress / mgx_hier_quantize / a decode):
// their warm-up did not reproduce the sequential state bit for bit; 0 in practice
uint64_t mgx_last_redo_count(mgx_ctx *ctx) {
  if (!ctx)
    return 0;
  unsigned long long v = 0;
  if (cudaMemcpyAsync(&v, ctx->d_scalars + MGX_SC_REDO, 8, cudaMemcpyDeviceToHost,
                      ctx->stream) != cudaSuccess ||
      cudaStreamSynchronize(ctx->stream) != cudaSuccess)
    return 0;
  return v;
}

int mgx_last_histogram(mgx_ctx *ctx, uint32_t *h_freq) {
  if (!ctx || !ctx->planned || !ctx->d_freq || !h_freq)
    return MGX_FAILURE;
  MGX_CUDA(cudaMemcpyAsync(h_freq, ctx->d_freq, (size_t)ctx->dict * 4,
                           cudaMemcpyDeviceToHost, ctx->stream));
  MGX_CUDA(cudaStreamSynchronize(ctx->stream));
  return MGX_SUCCESS;
}

int mgx_set_timing(mgx_ctx *ctx, int on) {
  if (!ctx)
    return MGX_FAILURE;
  if (on && !ctx->ev[0])
    for (int i = 0; i < MGX_TICK_COUNT; i++)
      MGX_CUDA(cudaEventCreate(&ctx->ev[i]));
  ctx->timing = on != 0;
  for (int i = 0; i < MGX_TICK_COUNT; i++)
    ctx->ev_set[i] = false;
  return MGX_SUCCESS;
}

int mgx_elapsed_ms(mgx_ctx *ctx, int from_tick, int to_tick, float *ms) {
  if (!ctx || from_tick < 0 || to_tick < 0 || from_tick >= MGX_TICK_COUNT ||
      to_tick >= MGX_TICK_COUNT || !ctx->ev_set[from_tick] ||
      !ctx->ev_set[to_tick])
    return MGX_FAILURE;
  MGX_CUDA(cudaEventSynchronize(ctx->ev[to_tick]));
  MGX_CUDA(cudaEventElapsedTime(ms, ctx->ev[from_tick], ctx->ev[to_tick]));
  return MGX_SUCCESS;
}

int mgx_plan(mgx_ctx *ctx, int D, const uint64_t *shape, int dtype,
             uint32_t dict_size, uint32_t chunk_size, uint64_t max_level) {
  return mgx_plan_grid(ctx, D, shape, dtype, dict_size, chunk_size, max_level, nullptr,
                       1);
}

int mgx_plan_grid(mgx_ctx *ctx, int D, const uint64_t *shape, int dtype,
                  uint32_t dict_size, uint32_t chunk_size, uint64_t max_level,
                  const void *const *coords, int normalize_coordinates) {
  if (!ctx)
    return MGX_FAILURE;
  if (dtype != MGX_F32 && dtype != MGX_F64)
    return MGX_BAD_DATA_TYPE;
  if (D > MGX_MAX_D)
    return MGX_TOO_MANY_DIMENSIONS;
  if (dict_size < 2 || dict_size > 16384 || chunk_size < 1) {
    ctx->err = "unsupported huff_dict_size / huff_block_size";
    return MGX_FAILURE;
  }
  MGX_CUDA(cudaSetDevice(ctx->device));
  MGX_CUDA(cudaStreamSynchronize(ctx->stream));
  release_plan(ctx);
  int rc = mgx_build_levels(D, shape, max_level, &ctx->lv);
  if (rc) {
    ctx->err = "bad shape (every dimension needs at least 3 nodes)";
    return rc;
  }
  ctx->dtype = dtype;
  ctx->dict = dict_size;
  ctx->chunk = chunk_size;
  const void *src;
  size_t bytes;
  if (coords) {
    // node coordinates must increase strictly (Hierarchy(shape, coords, config))
    for (int d = 0; d < D; d++) {
      if (!coords[d]) {
        ctx->err = "coordinate array missing";
        return MGX_FAILURE;
      }
      for (uint64_t i = 0; i + 1 < shape[d]; i++) {
        const double a = dtype == MGX_F32 ? (double)((const float *)coords[d])[i]
                                          : ((const double *)coords[d])[i];
        const double b = dtype == MGX_F32 ? (double)((const float *)coords[d])[i + 1]
                                          : ((const double *)coords[d])[i + 1];
        if (!(b > a)) {
          ctx->err = "node coordinates must be strictly increasing";
          return MGX_FAILURE;
        }
      }
    }
  }
  ctx->nonuniform = coords != nullptr;
  ctx->normalize = normalize_coordinates != 0;
  if (dtype == MGX_F32) {
    mgx_build_tables<float>(ctx->lv, &ctx->tbf, coords, ctx->normalize);
    src = ctx->tbf.buf.data();
    bytes = ctx->tbf.buf.size() * 4;
  } else {
    mgx_build_tables<double>(ctx->lv, &ctx->tbd, coords, ctx->normalize);
    src = ctx->tbd.buf.data();
    bytes = ctx->tbd.buf.size() * 8;
  }
  MGX_CUDA(cudaMalloc(&ctx->d_tables, bytes));
  MGX_CUDA(cudaMemcpyAsync(ctx->d_tables, src, bytes, cudaMemcpyHostToDevice,
                           ctx->stream));
  {
    // double-precision reciprocals of every table entry (used at the bm
    // offsets by the Thomas solves' exactly rounded fast division); the fast
    // path is switched off if a divisor has an all-ones significand, the one
    // case the correction-step theorem excludes
    const size_t cnt = bytes / mgx_elem(ctx);
    std::vector<double> rcp(cnt);
    ctx->fast_div = true;
    for (size_t k = 0; k < cnt; k++) {
      double b = dtype == MGX_F32 ? (double)ctx->tbf.buf[k] : ctx->tbd.buf[k];
      rcp[k] = 1.0 / b;
      if (dtype == MGX_F64) {
        uint64_t bits;
        std::memcpy(&bits, &b, 8);
        if ((bits & 0xFFFFFFFFFFFFFull) == 0xFFFFFFFFFFFFFull)
          ctx->fast_div = false;
      }
    }
    MGX_CUDA(cudaMalloc(&ctx->d_rcp, cnt * 8));
    // (on the context's stream, synchronised below: a cudaMemcpy from pageable memory
    // returns once the bytes are staged, and a non-blocking stream does not wait for it)
    MGX_CUDA(cudaMemcpyAsync(ctx->d_rcp, rcp.data(), cnt * 8, cudaMemcpyHostToDevice,
                             ctx->stream));
    MGX_CUDA(cudaStreamSynchronize(ctx->stream));
  }
  const uint64_t N = ctx->lv.N;
  const uint64_t nchunk = (N - 1) / chunk_size + 1;
  MGX_CUDA(cudaMalloc(&ctx->d_sym, N * 2 + 64));
  MGX_CUDA(cudaMalloc(&ctx->d_freq, (size_t)dict_size * 4));
  MGX_CUDA(cudaMalloc(&ctx->d_codebook, (size_t)dict_size * 8));
  MGX_CUDA(cudaMalloc(&ctx->d_decodebook, (size_t)(128 + dict_size) * 8));
  MGX_CUDA(cudaMalloc(&ctx->d_chunk_bits, nchunk * 8));
  MGX_CUDA(cudaMalloc(&ctx->d_chunk_off, (nchunk + 1) * 8));
  MGX_CUDA(cudaMalloc(&ctx->d_scan_state, (nchunk + 2) * 8));
  MGX_CUDA(cudaMalloc(&ctx->d_lut, 4096 * 4 + 4096 * 8)); // single + multi-symbol
  rc = grow_outliers(ctx, std::max<uint64_t>(1u << 16, N / 64));
  if (rc)
    return rc;
  MGX_CUDA(cudaStreamSynchronize(ctx->stream));
  ctx->planned = true;
  return MGX_SUCCESS;
}

int mgx_levels(const mgx_ctx *ctx, int *l_target, uint64_t *level_shapes) {
  if (!ctx || !ctx->planned)
    return MGX_FAILURE;
  const int pad = MGX_MAX_D - ctx->lv.D;
  if (l_target)
    *l_target = ctx->lv.L;
  if (level_shapes)
    for (int l = 0; l <= ctx->lv.L; l++)
      for (int d = 0; d < ctx->lv.D; d++)
        level_shapes[l * ctx->lv.D + d] = ctx->lv.n[l][d + pad];
  return MGX_SUCCESS;
}

uint64_t mgx_max_compressed_bytes(const mgx_ctx *ctx) {
  if (!ctx || !ctx->planned)
    return 0;
  const uint64_t N = ctx->lv.N;
  // header + worst-case codes (56 bits/symbol) + every element an outlier
  return mgx_container_fixed_bytes(N, ctx->dict, ctx->chunk) + 8 * N + 8 +
         16 * N + 8 * ((N - 1) / ctx->chunk + 1);
}

#define MGX_CHECK_PLAN()                                                       \
  do {                                                                         \
    if (!ctx)                                                                  \
      return MGX_FAILURE;                                                      \
    if (!ctx->planned) {                                                       \
      ctx->err = "mgx_plan has not been called";                               \
      return MGX_FAILURE;                                                      \
    }                                                                          \
    MGX_CUDA(cudaSetDevice(ctx->device));                                      \
  } while (0)

int mgx_norm_partial(mgx_ctx *ctx, const void *d_in, int linf, double *partial) {
  MGX_CHECK_PLAN();
  int rc = mgxk_norm_partial(ctx, d_in, linf);
  if (rc)
    return rc;
  MGX_CUDA(cudaMemcpyAsync(ctx->h_scalars, ctx->d_scalars + MGX_SC_NORM, 8,
                           cudaMemcpyDeviceToHost, ctx->stream));
  MGX_CUDA(cudaStreamSynchronize(ctx->stream));
  memcpy(partial, ctx->h_scalars, 8);
  return MGX_SUCCESS;
}

// NormCalculator.hpp:49-71
double mgx_finish_norm(int dtype, int linf, double partial, uint64_t total_elems,
                       int normalize_coordinates) {
  double norm;
  if (linf) {
    norm = partial;
  } else if (dtype == MGX_F32) {
    float p = (float)partial;
    norm = normalize_coordinates ? std::sqrt(p / total_elems) : std::sqrt(p);
  } else {
    norm = normalize_coordinates ? std::sqrt(partial / total_elems)
                                 : std::sqrt(partial);
  }
  if (norm == 0)
    norm = dtype == MGX_F32 ? (double)std::numeric_limits<float>::epsilon()
                            : std::numeric_limits<double>::epsilon();
  return norm;
}

int mgx_decompose(mgx_ctx *ctx, void *d_data, int orthogonal) {
  MGX_CHECK_PLAN();
  int rc;
  if (orthogonal && !ctx->legacy_levels) {
    // the row-streaming level kernels are out of place: nodes are read from a
    // copy, coefficients land in the caller's array
    rc = mgx_ensure_workspaces(ctx, true, false);
    if (rc)
      return rc;
    MGX_CUDA(cudaMemcpyAsync(ctx->d_w, d_data, ctx->lv.N * mgx_elem(ctx),
                             cudaMemcpyDeviceToDevice, ctx->stream));
    rc = mgxk_decompose_rows(ctx, ctx->d_w, d_data);
  } else {
    rc = mgxk_decompose(ctx, d_data, orthogonal);
  }
  if (rc)
    return rc;
  MGX_CUDA(cudaStreamSynchronize(ctx->stream));
  return MGX_SUCCESS;
}

int mgx_recompose(mgx_ctx *ctx, void *d_data, int orthogonal) {
  MGX_CHECK_PLAN();
  int rc;
  if (orthogonal && !ctx->legacy_levels) {
    rc = mgx_ensure_workspaces(ctx, true, false);
    if (rc)
      return rc;
    rc = mgxk_recompose_rows(ctx, d_data, ctx->d_w);
    if (!rc)
      MGX_CUDA(cudaMemcpyAsync(d_data, ctx->d_w, ctx->lv.N * mgx_elem(ctx),
                               cudaMemcpyDeviceToDevice, ctx->stream));
  } else {
    rc = mgxk_recompose(ctx, d_data, orthogonal);
  }
  if (rc)
    return rc;
  MGX_CUDA(cudaStreamSynchronize(ctx->stream));
  return MGX_SUCCESS;
}

int mgx_quantize(mgx_ctx *ctx, const void *d_coeff, int rel, double tol,
                 double s, double norm, int64_t *d_codes) {
  MGX_CHECK_PLAN();
  int rc = mgxk_quantize(ctx, d_coeff, rel, tol, s, norm, d_codes);
  if (rc)
    return rc;
  MGX_CUDA(cudaStreamSynchronize(ctx->stream));
  return MGX_SUCCESS;
}

int mgx_dequantize(mgx_ctx *ctx, const int64_t *d_codes, int rel, double tol,
                   double s, double norm, void *d_coeff) {
  MGX_CHECK_PLAN();
  int rc = mgxk_dequantize(ctx, d_codes, rel, tol, s, norm, d_coeff);
  if (rc)
    return rc;
  MGX_CUDA(cudaStreamSynchronize(ctx->stream));
  return MGX_SUCCESS;
}

int mgx_huff_codebook(mgx_ctx *ctx, const uint32_t *d_freq, uint64_t *d_codebook,
                      uint64_t *d_decodebook, int *max_len) {
  MGX_CHECK_PLAN();
  MGX_CUDA(cudaMemsetAsync(ctx->d_scalars, 0, MGX_SC_COUNT * 8, ctx->stream));
  int rc = mgxk_codebook(ctx, d_freq, d_codebook, d_decodebook);
  if (rc)
    return rc;
  MGX_CUDA(cudaMemcpyAsync(ctx->h_scalars, ctx->d_scalars, MGX_SC_COUNT * 8,
                           cudaMemcpyDeviceToHost, ctx->stream));
  MGX_CUDA(cudaStreamSynchronize(ctx->stream));
  if (max_len)
    *max_len = (int)ctx->h_scalars[MGX_SC_MAXLEN];
  if (ctx->h_scalars[MGX_SC_ERR]) {
    ctx->err = "codebook construction failed";
    return MGX_FAILURE;
  }
  return MGX_SUCCESS;
}

int mgx_huff_encode(mgx_ctx *ctx, const int64_t *d_codes, uint64_t n,
                    uint8_t *d_out, uint64_t cap, uint64_t *out_bytes) {
  MGX_CHECK_PLAN();
  if (n != ctx->lv.N) {
    ctx->err = "mgx_huff_encode: n must equal the planned element count";
    return MGX_FAILURE;
  }
  mgx_tick(ctx, MGX_TICK_C_NORM);
  for (int attempt = 0; attempt < 2; attempt++) {
    int rc = mgxk_histogram_reset(ctx);
    if (!rc)
      rc = mgxk_codes_to_symbols(ctx, d_codes, n);
    if (!rc)
      rc = mgxk_encode_container(ctx, n, d_out, cap, out_bytes);
    if (rc == MGX_FAILURE + 100 && attempt == 0) {
      rc = grow_outliers(ctx, ctx->h_scalars[MGX_SC_OUTLIERS]);
      if (rc)
        return rc;
      continue;
    }
    if (rc)
      return rc > MGX_BACKEND_UNAVAILABLE ? MGX_FAILURE : rc;
    break;
  }
  mgx_tick(ctx, MGX_TICK_C_END);
  MGX_CUDA(cudaStreamSynchronize(ctx->stream));
  return MGX_SUCCESS;
}

int mgx_huff_decode(mgx_ctx *ctx, const uint8_t *d_in, uint64_t bytes,
                    int64_t *d_codes, uint64_t n) {
  MGX_CHECK_PLAN();
  if (n != ctx->lv.N) {
    ctx->err = "mgx_huff_decode: n must equal the planned element count";
    return MGX_FAILURE;
  }
  const uint64_t *oidx;
  const int64_t *oval;
  uint64_t n_out;
  int rc = mgxk_decode_container(ctx, d_in, bytes, n, &oidx, &oval, &n_out);
  if (!rc)
    rc = mgxk_symbols_to_codes(ctx, n, d_codes, oidx, oval, n_out);
  if (rc)
    return rc;
  MGX_CUDA(cudaStreamSynchronize(ctx->stream));
  return MGX_SUCCESS;
}

int mgx_hier_quantize(mgx_ctx *ctx, const void *d_in, int rel, double tol,
                      double norm, int64_t *d_codes) {
  MGX_CHECK_PLAN();
  int rc = mgxk_histogram_reset(ctx);
  if (!rc)
    rc = mgxk_hier_forward(ctx, d_in, rel, tol, false, norm);
  if (rc)
    return rc;
  MGX_CUDA(cudaMemcpyAsync(ctx->h_scalars, ctx->d_scalars, MGX_SC_COUNT * 8,
                           cudaMemcpyDeviceToHost, ctx->stream));
  MGX_CUDA(cudaStreamSynchronize(ctx->stream));
  if (ctx->h_scalars[MGX_SC_ERR] & 0xf00ull) {
    ctx->err = "forward pass: tile pipeline timed out (flags " +
               std::to_string(ctx->h_scalars[MGX_SC_ERR]) + ")";
    return MGX_FAILURE;
  }
  uint64_t n_out = ctx->h_scalars[MGX_SC_OUTLIERS];
  if (n_out > ctx->out_cap) {
    rc = grow_outliers(ctx, n_out);
    if (!rc)
      rc = mgxk_histogram_reset(ctx);
    if (!rc)
      rc = mgxk_hier_forward(ctx, d_in, rel, tol, false, norm);
    if (rc)
      return rc;
  }
  rc = mgxk_symbols_to_codes(ctx, ctx->lv.N, d_codes, ctx->d_out_idx,
                             ctx->d_out_val, n_out);
  if (rc)
    return rc;
  MGX_CUDA(cudaStreamSynchronize(ctx->stream));
  return MGX_SUCCESS;
}

// Compressor::Compress, Compressor.hpp:204-251
int mgx_compress(mgx_ctx *ctx, const void *d_in, int rel, double tol, double s,
                 double *norm_inout, uint8_t *d_out, uint64_t cap,
                 uint64_t *out_bytes) {
  MGX_CHECK_PLAN();
  const bool linf = std::isinf(s);
  const bool orth = need_orthogonal(ctx, s);
  const uint64_t N = ctx->lv.N;
  double norm = norm_inout ? *norm_inout : 0.0;
  mgx_tick(ctx, MGX_TICK_C_BEGIN);
  if (rel && !(norm > 0)) { // CalculateNorm, Compressor.hpp:122-130
    double partial;
    int rc = mgx_norm_partial(ctx, d_in, linf, &partial);
    if (rc)
      return rc;
    norm = mgx_finish_norm(ctx->dtype, linf, partial, N, ctx->normalize ? 1 : 0);
    if (norm_inout)
      *norm_inout = norm;
  }
  mgx_tick(ctx, MGX_TICK_C_NORM);
  for (int attempt = 0; attempt < 2; attempt++) {
    int rc = mgxk_histogram_reset(ctx);
    if (rc)
      return rc;
    if (!orth) {
      // fused: coefficients of every level + quantise + histogram in one pass
      rc = mgxk_hier_forward(ctx, d_in, rel, tol, false, norm);
    } else if (!ctx->legacy_levels) {
      // level kernels with the quantiser fused into the coefficient pass
      rc = mgxk_decompose_quantize_rows(ctx, d_in, rel, tol, s, norm);
    } else {
      rc = mgx_ensure_workspaces(ctx, true, true);
      if (rc)
        return rc;
      MGX_CUDA(cudaMemcpyAsync(ctx->d_v, d_in, N * mgx_elem(ctx),
                               cudaMemcpyDeviceToDevice, ctx->stream));
      rc = mgxk_decompose(ctx, ctx->d_v, 1);
      if (!rc)
        rc = mgxk_quantize_sym(ctx, ctx->d_v, rel, tol, s, nullptr, norm);
    }
    mgx_tick(ctx, MGX_TICK_C_REFACTOR);
    if (!rc)
      rc = mgxk_encode_container(ctx, N, d_out, cap, out_bytes);
    if (rc == MGX_FAILURE + 100 && attempt == 0) {
      rc = grow_outliers(ctx, ctx->h_scalars[MGX_SC_OUTLIERS]);
      if (rc)
        return rc;
      continue;
    }
    if (rc)
      return rc > MGX_BACKEND_UNAVAILABLE ? MGX_FAILURE : rc;
    break;
  }
  mgx_tick(ctx, MGX_TICK_C_END);
  MGX_CUDA(cudaStreamSynchronize(ctx->stream));
  return MGX_SUCCESS;
}

// Compressor::Decompress, Compressor.hpp:253-286
int mgx_decompress(mgx_ctx *ctx, const uint8_t *d_in, uint64_t bytes, int rel,
                   double tol, double s, double norm, void *d_out) {
  MGX_CHECK_PLAN();
  const bool orth = need_orthogonal(ctx, s);
  const uint64_t N = ctx->lv.N;
  const uint64_t *oidx;
  const int64_t *oval;
  uint64_t n_out;
  mgx_tick(ctx, MGX_TICK_D_BEGIN);
  int rc = mgxk_decode_container(ctx, d_in, bytes, N, &oidx, &oval, &n_out);
  if (rc)
    return rc;
  mgx_tick(ctx, MGX_TICK_D_DECODE);
  rc = MGX_FAILURE + 101;
  if (!orth) // fused dequantise + recompose (hierarchical basis)
    rc = mgxk_hier_inverse(ctx, rel, tol, norm, d_out, oidx, oval, n_out);
  if (rc == MGX_FAILURE + 101 && orth && !ctx->legacy_levels &&
      !ctx->no_sym_recompose) {
    // the level kernels dequantise the symbols on the fly; only the coarse
    // levels that hold outliers (their true codes are in the outlier list) are
    // dequantised to floats first
    int sym_from = 0;
    rc = MGX_SUCCESS;
    if (n_out) {
      int lo = 0;
      rc = mgxk_outlier_max_level(ctx, oidx, n_out, &lo);
      if (!rc)
        rc = mgx_ensure_workspaces(ctx, false, true);
      if (!rc)
        rc = mgxk_dequantize_box(ctx, lo, rel, tol, s, norm, ctx->d_v, oidx, oval,
                                 n_out);
      sym_from = lo + 1;
    }
    mgx_tick(ctx, MGX_TICK_D_DEQUANT);
    if (!rc)
      rc = mgxk_recompose_sym_rows(ctx, rel, tol, s, norm, ctx->d_v, sym_from, d_out);
  } else if (rc == MGX_FAILURE + 101 && orth && !ctx->legacy_levels) {
    rc = mgx_ensure_workspaces(ctx, false, true);
    if (!rc)
      rc = mgxk_dequantize_sym_out(ctx, rel, tol, s, norm, ctx->d_v, oidx, oval,
                                   n_out);
    mgx_tick(ctx, MGX_TICK_D_DEQUANT);
    if (!rc)
      rc = mgxk_recompose_rows(ctx, ctx->d_v, d_out);
  } else if (rc == MGX_FAILURE + 101) {
    rc = mgxk_dequantize_sym_out(ctx, rel, tol, s, norm, d_out, oidx, oval, n_out);
    mgx_tick(ctx, MGX_TICK_D_DEQUANT);
    if (!rc)
      rc = mgxk_recompose(ctx, d_out, orth ? 1 : 0);
  } else {
    mgx_tick(ctx, MGX_TICK_D_DEQUANT);
  }
  if (rc)
    return rc;
  mgx_tick(ctx, MGX_TICK_D_END);
  MGX_CUDA(cudaStreamSynchronize(ctx->stream));
  return MGX_SUCCESS;
}

static int ensure_stage(mgx_ctx *ctx, size_t in_bytes, size_t out_bytes) {
  if (ctx->d_stage_in_bytes < in_bytes) {
    free_dev(ctx->d_stage_in);
    MGX_CUDA(cudaMalloc(&ctx->d_stage_in, in_bytes));
    ctx->d_stage_in_bytes = in_bytes;
  }
  if (ctx->d_stage_out_bytes < out_bytes) {
    free_dev(ctx->d_stage_out);
    MGX_CUDA(cudaMalloc(&ctx->d_stage_out, out_bytes));
    ctx->d_stage_out_bytes = out_bytes;
  }
  return MGX_SUCCESS;
}

// host-pointer path of general_compress_pipeline
// (CompressionHighLevel.hpp:141-183, GPUPipelines.hpp:89-205)
int mgx_compress_host(mgx_ctx *ctx, const void *h_in, int rel, double tol,
                      double s, double *norm_inout, uint8_t *h_out,
                      uint64_t cap, uint64_t *out_bytes) {
  MGX_CHECK_PLAN();
  const size_t in_bytes = ctx->lv.N * mgx_elem(ctx);
  int rc = ensure_stage(ctx, in_bytes, cap);
  if (rc)
    return rc;
  MGX_CUDA(cudaMemcpyAsync(ctx->d_stage_in, h_in, in_bytes,
                           cudaMemcpyHostToDevice, ctx->stream));
  rc = mgx_compress(ctx, ctx->d_stage_in, rel, tol, s, norm_inout,
                    (uint8_t *)ctx->d_stage_out, cap, out_bytes);
  if (rc)
    return rc;
  MGX_CUDA(cudaMemcpyAsync(h_out, ctx->d_stage_out, *out_bytes,
                           cudaMemcpyDeviceToHost, ctx->stream));
  MGX_CUDA(cudaStreamSynchronize(ctx->stream));
  return MGX_SUCCESS;
}

int mgx_decompress_host(mgx_ctx *ctx, const uint8_t *h_in, uint64_t bytes,
                        int rel, double tol, double s, double norm,
                        void *h_out) {
  MGX_CHECK_PLAN();
  const size_t out_bytes = ctx->lv.N * mgx_elem(ctx);
  int rc = ensure_stage(ctx, bytes + 8, out_bytes);
  if (rc)
    return rc;
  MGX_CUDA(cudaMemcpyAsync(ctx->d_stage_in, h_in, bytes, cudaMemcpyHostToDevice,
                           ctx->stream));
  rc = mgx_decompress(ctx, (const uint8_t *)ctx->d_stage_in, bytes, rel, tol, s,
                      norm, ctx->d_stage_out);
  if (rc)
    return rc;
  MGX_CUDA(cudaMemcpyAsync(h_out, ctx->d_stage_out, out_bytes,
                           cudaMemcpyDeviceToHost, ctx->stream));
  MGX_CUDA(cudaStreamSynchronize(ctx->stream));
  return MGX_SUCCESS;
}

} // extern "C"
