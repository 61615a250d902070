/* TEST INFRASTRUCTURE ONLY -- CPU restatement ("port") of the MGARD-X hot path.
 *
 * Plain-C restatement of the algorithms behind
 *   mgard_x::Compressor<D,T,Dev>::Compress / Decompress
 *   (reference include/mgard-x/CompressionLowLevel/Compressor.hpp:204-286)
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may load
 * this library; the product (mgard_b200/) never does.
 *
 * Parity pin: every entry point is checked against the reference itself
 * (oracle/_ref/libmgardx_ref.so, compiled from /root/reference by
 * oracle/Makefile) in tests/test_oracle_vs_ref.py, and against golden vectors
 * generated from that library (tests/golden/).
 */
#ifndef MGX_ORACLE_H
#define MGX_ORACLE_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MGXO_MAX_D 5
#define MGXO_MAX_L 40

/* Level shapes: Hierarchy::init, reference Hierarchy/Hierarchy.hpp:277-309.
 * level_shapes is row-major [(L+1)][D]; returns L (l_target). */
int mgxo_levels(int D, const uint64_t *shape, uint64_t max_level,
                uint64_t *level_shapes);

/* In-place multilevel decomposition / recomposition into the reordered,
 * level-blocked layout (DataRefactoring.hpp:24-317). */
/* Grid used by every call that follows: coords[d] -> shape[d] node coordinates in
 * the data type of that call (float for *_f32, double for *_f64), or NULL for a
 * uniform grid (unit cube if normalize_coordinates, else unit spacing). */
void mgxo_set_grid(int D, const void *const *coords, int normalize_coordinates);

int mgxo_decompose_f32(int D, const uint64_t *shape, float *v, int orthogonal);
int mgxo_decompose_f64(int D, const uint64_t *shape, double *v, int orthogonal);
int mgxo_recompose_f32(int D, const uint64_t *shape, float *v, int orthogonal);
int mgxo_recompose_f64(int D, const uint64_t *shape, double *v, int orthogonal);

/* norm_calculator, NormCalculator.hpp:13-80 (accumulates in T like SERIAL). */
double mgxo_norm_f32(uint64_t n, const float *v, double s, int normalize);
double mgxo_norm_f64(uint64_t n, const double *v, double s, int normalize);

/* LinearQuantizer::Quantize / Dequantize, LinearQuantization.hpp:19-145,234-420.
 * s = +inf selects L-inf; rel!=0 multiplies tol by norm. */
int mgxo_quantize_f32(int D, const uint64_t *shape, const float *c, int rel,
                      double tol, double s, double norm, uint32_t dict,
                      int64_t *codes);
int mgxo_quantize_f64(int D, const uint64_t *shape, const double *c, int rel,
                      double tol, double s, double norm, uint32_t dict,
                      int64_t *codes);
int mgxo_dequantize_f32(int D, const uint64_t *shape, const int64_t *codes,
                        int rel, double tol, double s, double norm,
                        uint32_t dict, float *c);
int mgxo_dequantize_f64(int D, const uint64_t *shape, const int64_t *codes,
                        int rel, double tol, double s, double norm,
                        uint32_t dict, double *c);

/* GetCodebook (GetCodebook.hpp:23-145): freq[dict] -> codebook[dict]
 * (len<<56|code), decodebook = first[64] | entry[64] | keys[dict] (u64 each).
 * Returns max code length, or -1 when it exceeds 56. */
int mgxo_codebook(const uint32_t *freq, uint32_t dict, uint64_t *codebook,
                  uint64_t *decodebook);

/* Huffman::Compress + Serialize / Deserialize + Decompress
 * (Huffman.hpp:157-632): the "MGXHUFF" container. */
/* reference defect hooks, see mgx_oracle_huffman.inc */
void mgxo_set_oob_value(uint32_t v);
uint64_t mgxo_oob_reads(void);
int mgxo_huffman_compress(const int64_t *codes, uint64_t n, uint32_t dict,
                          uint32_t chunk, uint8_t *out, uint64_t cap,
                          uint64_t *out_bytes);
int mgxo_huffman_decompress(const uint8_t *in, uint64_t bytes, uint64_t n,
                            int64_t *codes);

/* Compressor::Compress / Decompress (Compressor.hpp:204-286). */
int mgxo_compress_f32(int D, const uint64_t *shape, const float *in, int rel,
                      double tol, double s, double *norm_io, uint32_t dict,
                      uint32_t chunk, uint8_t *out, uint64_t cap,
                      uint64_t *out_bytes);
int mgxo_compress_f64(int D, const uint64_t *shape, const double *in, int rel,
                      double tol, double s, double *norm_io, uint32_t dict,
                      uint32_t chunk, uint8_t *out, uint64_t cap,
                      uint64_t *out_bytes);
int mgxo_decompress_f32(int D, const uint64_t *shape, const uint8_t *in,
                        uint64_t bytes, int rel, double tol, double s,
                        double norm, uint32_t dict, float *out);
int mgxo_decompress_f64(int D, const uint64_t *shape, const uint8_t *in,
                        uint64_t bytes, int rel, double tol, double s,
                        double norm, uint32_t dict, double *out);

#ifdef __cplusplus
}
#endif
#endif
