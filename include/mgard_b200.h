/* mgard_b200 -- C ABI of the B200-native MGARD-X compression hot path.
 *
 * This is the drop-in boundary: plain pointers and sizes, no C++/torch types.
 * Every entry point names the reference interface (file:line under the
 * reference tree) it stands in for.  All `d_*` pointers are DEVICE pointers on
 * the context's GPU; work is enqueued on the context's CUDA stream and every
 * call that returns a value through a host pointer synchronises that stream
 * before returning (the reference's API is synchronous at return as well,
 * CompressionHighLevel.hpp:47-286).
 *
 * There is no CPU fallback: every function fails with MGX_BACKEND_UNAVAILABLE
 * when no CUDA device can be used.
 */
#ifndef MGARD_B200_H
#define MGARD_B200_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* mirrors mgard_x::compress_status_type, include/mgard-x/Utilities/Types.h:76-83 */
enum mgx_status {
  MGX_SUCCESS = 0,
  MGX_FAILURE = 1,
  MGX_OUTPUT_TOO_LARGE = 2,
  MGX_TOO_MANY_DIMENSIONS = 3,
  MGX_BAD_DATA_TYPE = 4,
  MGX_BACKEND_UNAVAILABLE = 5
};

/* mirrors mgard_x::data_type, Types.h:60 */
enum mgx_dtype { MGX_F32 = 0, MGX_F64 = 1 };

typedef struct mgx_ctx mgx_ctx;

/* Context = one (device, stream) pair with its workspaces and cached hierarchy
 * tables; replaces DeviceRuntime<Dev>::Initialize + the process-global
 * CompressorCache (CompressionLowLevel/CompressorCache.hpp:25-147).
 * cuda_stream may be NULL (a private stream is created). */
int mgx_create(mgx_ctx **ctx, int device, void *cuda_stream);
void mgx_destroy(mgx_ctx *ctx);
const char *mgx_last_error(const mgx_ctx *ctx);
const char *mgx_version(void);

/* Hierarchy<D,T,Dev>(shape, config) + Compressor::Adapt
 * (Hierarchy/Hierarchy.hpp:814-837, CompressionLowLevel/Compressor.hpp:60-84).
 * shape is slowest -> fastest; dict_size/chunk_size are Config::huff_dict_size /
 * huff_block_size (Config/Config.h:17-18); max_level = Config::max_larget_level
 * (UINT64_MAX = unlimited). */
int mgx_plan(mgx_ctx *ctx, int D, const uint64_t *shape, int dtype,
             uint32_t dict_size, uint32_t chunk_size, uint64_t max_level);

/* Hierarchy<D,T,Dev>(shape, coords, config) (Hierarchy/Hierarchy.hpp:840-857): a
 * NON-UNIFORM Cartesian grid.  coords[d] -> shape[d] strictly increasing node
 * coordinates of dimension d in HOST memory, in the data type T; per-level spacing
 * follows coord_to_dist / reduce_dist for Cartesian_Grid_Non_Uniform
 * (Hierarchy.hpp:22-61, 86-140: pairs of cells merged, last cell split in half when
 * the node count is even).  coords == NULL: uniform grid on the unit cube
 * (normalize_coordinates != 0, the default of mgx_plan) or with unit spacing
 * (normalize_coordinates == 0: create_uniform_coords, Hierarchy.hpp:762-782; the L2
 * norm is then sqrt(sum u^2) without the 1/N, NormCalculator.hpp:64). */
int mgx_plan_grid(mgx_ctx *ctx, int D, const uint64_t *shape, int dtype,
                  uint32_t dict_size, uint32_t chunk_size, uint64_t max_level,
                  const void *const *coords, int normalize_coordinates);
/* l_target and the (l_target+1) x D level shapes (Hierarchy.hpp:277-309). */
int mgx_levels(const mgx_ctx *ctx, int *l_target, uint64_t *level_shapes);
/* bytes a caller must provide to mgx_compress / mgx_huff_encode in the worst
 * case (every code an outlier). */
uint64_t mgx_max_compressed_bytes(const mgx_ctx *ctx);

/* norm_calculator (CompressionLowLevel/NormCalculator.hpp:13-80), split so a
 * multi-GPU caller can all-reduce the partial: linf != 0 -> max|u|, else
 * sum u^2 (fp64 accumulation).  mgx_finish_norm applies the sqrt / N /
 * zero->epsilon rules of NormCalculator.hpp:49-71 to a (reduced) partial. */
int mgx_norm_partial(mgx_ctx *ctx, const void *d_in, int linf, double *partial);
double mgx_finish_norm(int dtype, int linf, double partial, uint64_t total_elems,
                       int normalize_coordinates);

/* DataRefactor::Decompose / Recompose, in place, reordered level-blocked layout
 * (DataRefactoring/DataRefactor.hpp:73-130,
 *  MultiDimension/DataRefactoring.hpp:24-317). */
int mgx_decompose(mgx_ctx *ctx, void *d_data, int orthogonal);
int mgx_recompose(mgx_ctx *ctx, void *d_data, int orthogonal);

/* LinearQuantizer::Quantize / Dequantize with the dictionary shift folded in
 * (Quantization/LinearQuantization.hpp:19-145, 315-420).  s = +inf -> L-inf,
 * 0 -> L2, else s-norm; rel != 0 multiplies tol by norm. */
int mgx_quantize(mgx_ctx *ctx, const void *d_coeff, int rel, double tol,
                 double s, double norm, int64_t *d_codes);
int mgx_dequantize(mgx_ctx *ctx, const int64_t *d_codes, int rel, double tol,
                   double s, double norm, void *d_coeff);

/* Huffman::Compress + Serialize / Deserialize + Decompress: the "MGXHUFF"
 * container (Lossless/ParallelHuffman/Huffman.hpp:157-632). */
int mgx_huff_encode(mgx_ctx *ctx, const int64_t *d_codes, uint64_t n,
                    uint8_t *d_out, uint64_t cap, uint64_t *out_bytes);
int mgx_huff_decode(mgx_ctx *ctx, const uint8_t *d_in, uint64_t bytes,
                    int64_t *d_codes, uint64_t n);
/* GetCodebook (ParallelHuffman/GetCodebook.hpp:23-145) on its own: d_freq[dict]
 * -> d_codebook[dict] (len<<56|code) and d_decodebook[128+dict]
 * (first[64] | entry[64] | keys[dict]).  Returns the longest code length
 * through max_len. */
int mgx_huff_codebook(mgx_ctx *ctx, const uint32_t *d_freq, uint64_t *d_codebook,
                      uint64_t *d_decodebook, int *max_len);

/* Compressor<D,T,Dev>::Compress / Decompress
 * (CompressionLowLevel/Compressor.hpp:204-286): norm -> decompose -> quantize
 * -> Huffman -> serialize, fused where the basis allows it.  d_in is NOT
 * modified.  REL: if *norm_inout > 0 it is used as the (already global) norm,
 * otherwise the norm of d_in is computed and returned. */
int mgx_compress(mgx_ctx *ctx, const void *d_in, int rel, double tol, double s,
                 double *norm_inout, uint8_t *d_out, uint64_t cap,
                 uint64_t *out_bytes);
int mgx_decompress(mgx_ctx *ctx, const uint8_t *d_in, uint64_t bytes, int rel,
                   double tol, double s, double norm, void *d_out);

/* The fused fast path of mgx_compress on its own (L-inf, D <= 3): Decompose with
 * the hierarchical basis + Quantize in ONE pass over d_in
 * (DataRefactoring.hpp:77-114 with orthogonal_projection == false, then
 * LinearQuantization.hpp:19-145), returned as the reference's int64 codes so
 * it can be compared with mgx_decompose + mgx_quantize and with the oracle. */
int mgx_hier_quantize(mgx_ctx *ctx, const void *d_in, int rel, double tol,
                      double norm, int64_t *d_codes);

/* Same two calls with HOST buffers: pinned staging + H2D / D2H inside
 * (the host-pointer branch of general_compress_pipeline,
 *  CompressionHighLevel.hpp:141-183, GPUPipelines.hpp:89-205). */
int mgx_compress_host(mgx_ctx *ctx, const void *h_in, int rel, double tol,
                      double s, double *norm_inout, uint8_t *h_out,
                      uint64_t cap, uint64_t *out_bytes);
int mgx_decompress_host(mgx_ctx *ctx, const uint8_t *h_in, uint64_t bytes,
                        int rel, double tol, double s, double norm,
                        void *h_out);

/* ---- host-only helpers (no context, no GPU) -------------------------------- */

/* Level structure of a shape without planning it: l_target and the
 * (l_target+1) x D level shapes (Hierarchy.hpp:277-309).  level_shapes may be
 * NULL.  Returns MGX_TOO_MANY_DIMENSIONS for D > 5. */
int mgx_hierarchy_levels(int D, const uint64_t *shape, uint64_t max_level,
                         int *l_target, uint64_t *level_shapes);

/* MaxDim domain decomposition (DomainDecomposer.hpp:192-223): split the largest
 * dimension (first one on ties) into slabs of *size rows, halving the slab
 * (round up) until there are at least min_subdomains of them.  The reference
 * halves until a slab fits the device; with one slab per GPU the stop rule is
 * the GPU count.  *num = ceil(shape[dim] / size) (the last slab is the
 * leftover, DomainDecomposer.hpp:118-133). */
int mgx_max_dim_decomposition(int D, const uint64_t *shape,
                              uint64_t min_subdomains, int *dim, uint64_t *size,
                              uint64_t *num);
/* shape and offset (along dim) of subdomain id */
int mgx_subdomain_shape(int D, const uint64_t *shape, int dim, uint64_t size,
                        uint64_t id, uint64_t *sub_shape, uint64_t *offset);
/* calc_local_abs_tol (ErrorToleranceCalculator.hpp:126-147), evaluated in the
 * data type's arithmetic like the reference: every subdomain is then compressed
 * with an ABS bound of this value. */
double mgx_local_abs_tol(int dtype, int rel, double norm, double tol, double s,
                         uint64_t num_subdomains);

/* Histogram of the (shifted, outlier-zeroed) symbols of the context's last
 * mgx_compress / mgx_hier_quantize / mgx_huff_encode call: dict_size uint32 counts
 * copied to HOST memory.  Replaces reading Huffman's freq workspace
 * (ParallelHuffman/Histogram.hpp:15-235, HuffmanWorkspace.hpp: freq_subarray);
 * lets stage A13 be parity-checked on its own. */
int mgx_last_histogram(mgx_ctx *ctx, uint32_t *h_freq);

/* Chains that the partitioned Thomas solves have had to redo sequentially since the
 * context's counters were last reset (start of mgx_compress or of a decode; one
 * synchronising 8-byte read) because a partition's warm-up did not reproduce the sequential
 * sweep's state bit for bit (mgx_level.cu: k_ipk_part_*; replaces the one-thread-per-
 * chain Ipk1/2/3Reo, IterativeProcessingKernel3D.hpp:28-1169).  0 in practice. */
uint64_t mgx_last_redo_count(mgx_ctx *ctx);

/* kernels launched by this context since creation (bench.py's gpu_launches) */
uint64_t mgx_launch_count(const mgx_ctx *ctx);

/* Stage timing with CUDA events on the context's stream (what the reference
 * prints through its Timer / log::time calls, Compressor.hpp:125-248,
 * Huffman.hpp:186-330).  Ticks are recorded by mgx_compress / mgx_decompress
 * while timing is on; mgx_elapsed_ms reads the last call's ticks. */
enum mgx_tick_id {
  MGX_TICK_C_BEGIN = 0,    /* mgx_compress entered                          */
  MGX_TICK_C_NORM = 1,     /* after CalculateNorm (if it ran)               */
  MGX_TICK_C_REFACTOR = 2, /* after Decompose + Quantize (+ histogram)      */
  MGX_TICK_C_CODEBOOK = 3, /* after GetCodebook                             */
  MGX_TICK_C_SIZES = 4,    /* after chunk bit counts + offset scan          */
  MGX_TICK_C_PACK = 5,     /* after bit packing                             */
  MGX_TICK_C_END = 6,      /* container complete                            */
  MGX_TICK_D_BEGIN = 8,    /* mgx_decompress entered                        */
  MGX_TICK_D_DECODE = 9,   /* after Huffman decode                          */
  MGX_TICK_D_DEQUANT = 10, /* after Dequantize (+ outlier restore)          */
  MGX_TICK_D_END = 11,     /* after Recompose                               */
  MGX_TICK_COUNT = 16
};
/* Norm reduction WITHOUT any allocation (no cudaMalloc / cudaFree, no context):
 * for callers under CUDA graph capture, tight serving loops, harnesses that
 * forbid allocation.  No counterpart in the reference.
 * Computes what mgx_norm_partial computes (linf != 0:
 * max |u|, else sum u^2 accumulated in fp64) over n elements of `dtype` at
 * d_in and writes it to the DEVICE double d_partial, asynchronously on
 * cuda_stream (a cudaStream_t, or NULL for the default stream) of `device`;
 * the caller synchronises.  Scratch is a 16 KB static __device__ array that
 * every GPU loading the library holds until exit.  Limit: n <= 2^31 elements
 * (2048 blocks x 2^20), larger n returns MGX_OUTPUT_TOO_LARGE.  All calls on a
 * device share the scratch: calls that overlap on the GPU race.  The first call
 * on a device resolves the scratch address and loads the kernels' module: make
 * one eager call per device before calling under stream capture
 * (tests/test_zz_no_alloc.py replays it from a CUDA graph).  On failure a
 * message is copied into err (if err != NULL, capacity err_cap). */
int mgx_norm_partial_noalloc(int device, void *cuda_stream, const void *d_in,
                             int dtype, uint64_t n, int linf, double *d_partial,
                             char *err, uint64_t err_cap);

int mgx_set_timing(mgx_ctx *ctx, int on);
int mgx_elapsed_ms(mgx_ctx *ctx, int from_tick, int to_tick, float *ms);

#ifdef __cplusplus
}
#endif
#endif
