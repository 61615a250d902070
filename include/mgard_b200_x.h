/* mgard_b200 -- C ABI of the HIGH-LEVEL path: what mgard_x::compress /
 * mgard_x::decompress (reference include/compress_x.hpp:31-178) do around the
 * low-level Compressor: self-describing "MGARD" header, MaxDim domain
 * decomposition with [u64 size][payload] records, host- or device-resident
 * buffers, raw fallback when a subdomain does not compress.
 *
 * include/mgard_b200/compress_x.hpp is the C++ face of these entry points
 * (namespace mgard_x, the reference's signatures); nothing but plain pointers
 * and sizes crosses this boundary.
 *
 *   mgx_x_compress / mgx_x_decompress   general_compress / general_decompress
 *                                       (CompressionHighLevel.hpp:47-309, 373-579)
 *                                       + compress_pipeline_gpu /
 *                                       decompress_pipeline_gpu
 *                                       (GPUPipelines.hpp:3-567)
 *   mgx_metadata_serialize / _parse     MetadataBase::Serialize / Deserialize
 *                                       (src/mgard-x/Metadata/Metadata.cpp:259-762,
 *                                        src/mgard.proto)
 *   mgx_crc32                           zlib crc32_z as used by ComputeCRC32
 *                                       (Metadata.cpp:29-41)
 *   mgx_x_release_cache                 release_cache (compress_x.hpp:166)
 *   mgx_x_pin_memory / _unpin / _check  pin_memory ... (compress_x.hpp:170-176)
 *
 * Threading: ONE call at a time per process.  mgx_x_compress / mgx_x_decompress /
 * mgx_x_release_cache serialise on a process-wide mutex (they share the per-device
 * context cache, like the reference's static Cache singletons,
 * CompressionHighLevel.hpp:91-102): concurrent callers are safe but do not overlap.
 * To drive several GPUs at once use one process per GPU (what bench.py and
 * mgard_b200/sharding.py do) or the low-level mgx_ctx API, whose contexts are
 * independent (one context = one device + one stream, one call at a time each).
 * No exception crosses this boundary: C++ exceptions are caught and returned as
 * MGX_FAILURE with the text in mgx_x_last_error().
 */
#ifndef MGARD_B200_X_H
#define MGARD_B200_X_H

#include <stddef.h>
#include <stdint.h>

#include "mgard_b200.h"

#ifdef __cplusplus
extern "C" {
#endif

/* numeric values follow the reference enums they stand for */
enum mgx_device_type { /* RuntimeX/DataTypes.h:106-114 */
  MGX_DEV_AUTO = 0,
  MGX_DEV_SERIAL = 1,
  MGX_DEV_OPENMP = 2,
  MGX_DEV_CUDA = 3,
  MGX_DEV_HIP = 4,
  MGX_DEV_SYCL = 5,
  MGX_DEV_NONE = 6
};
enum mgx_error_bound { MGX_EB_REL = 0, MGX_EB_ABS = 1 }; /* Types.h:34 */
enum mgx_lossless {                                      /* Types.h:36-56 */
  MGX_LL_HUFFMAN = 0,
  MGX_LL_HUFFMAN_LZ4 = 1,
  MGX_LL_HUFFMAN_ZSTD = 2,
  MGX_LL_CPU = 3
};
enum mgx_dd_type { MGX_DD_MAXDIM = 0, MGX_DD_BLOCK = 1, MGX_DD_VARIABLE = 2 };
enum mgx_decomp_type { MGX_DC_MULTIDIM = 0, MGX_DC_SINGLEDIM = 1, MGX_DC_HYBRID = 2 };

/* The fields of mgard_x::Config (Config/Config.h:10-46) this path reads. */
typedef struct mgx_config {
  int dev_type;                  /* mgx_device_type; AUTO and CUDA are served */
  int dev_id;
  int domain_decomposition;      /* mgx_dd_type; MaxDim or Block              */
  int decomposition;             /* mgx_decomp_type; MultiDim only            */
  int lossless;                  /* mgx_lossless; Huffman, Huffman_LZ4, Huffman_Zstd */
  uint32_t huff_dict_size;       /* 8192                                      */
  uint32_t huff_block_size;      /* 1024                                      */
  uint32_t reorder;              /* recorded in the header (0)                */
  int normalize_coordinates;     /* uniform grids: unit cube (1) or unit spacing (0) */
  int auto_pin_host_buffers;     /* cudaHostRegister host buffers for the call */
  int compress_with_dryrun;      /* 1: the reconstruction is written over the INPUT
                                    array (GPUPipelines.hpp:198-202); one-GPU entries */
  int auto_cache_release;
  int log_level;
  uint64_t max_larget_level;     /* UINT64_MAX = no limit                     */
  uint64_t max_memory_footprint; /* UINT64_MAX = no limit (bytes)             */
  uint64_t block_size;           /* Block decomposition: edge of a block (256) */
  int zstd_compress_level;       /* Huffman_Zstd: level of ZSTD_compress (3)   */
  uint32_t lz4_block_size;       /* Huffman_LZ4: bytes per LZ4 block (1 << 15) */
} mgx_config;

void mgx_config_default(mgx_config *cfg); /* Config::Config(), Config.cpp:31-66 */

/* compress(D, dtype, shape, tol, s, mode, in, out, out_size, config,
 *          output_pre_allocated)  -- compress_x.hpp:56-61.
 * `in` may be a host or a device pointer (MemoryManager::IsDevicePointer,
 * CompressionHighLevel.hpp:144).  output_pre_allocated == 0: *out is allocated
 * here with malloc (host input) or cudaMalloc on the input's device (device
 * input), N*sizeof(T)+1e6 bytes, and owned by the caller.  != 0: *out_size
 * carries the capacity in; the compressed size comes back in *out_size. */
int mgx_x_compress(int D, int dtype, const uint64_t *shape, double tol, double s,
                   int mode, const void *in, void **out, size_t *out_size,
                   const mgx_config *cfg, int output_pre_allocated);

/* compress(..., coords, config, output_pre_allocated) -- the NON-UNIFORM overloads,
 * compress_x.hpp:74-100: coords[d] -> shape[d] strictly increasing node coordinates
 * of dimension d in HOST memory, in the data type of the field
 * (CompressionHighLevel.hpp:687-740 reinterprets the byte pointers as float* /
 * double*).  The coordinates are written into the header as doubles
 * (Metadata.hpp:138-153); mgx_x_decompress rebuilds the hierarchy from them, narrowed
 * to float as the reference does (CompressionHighLevel.hpp:443-450). */
int mgx_x_compress_coords(int D, int dtype, const uint64_t *shape, double tol, double s,
                          int mode, const void *in, void **out, size_t *out_size,
                          const void *const *coords, const mgx_config *cfg,
                          int output_pre_allocated);

/* decompress(in, in_size, out, shape, dtype, config, output_pre_allocated)
 * -- compress_x.hpp:137-141.  shape_out (5 entries), D_out, dtype_out may be
 * NULL. */
/* compress on SEVERAL GPUs of this process (no counterpart in the reference, whose
 * compress drives one device, Config::dev_id; this is the subdomain loop of
 * compress_pipeline_gpu, GPUPipelines.hpp:89-205, dealt to `num_devices` devices).
 * `in` and `*out` are HOST buffers.  The domain is cut into at least num_devices
 * MaxDim subdomains (DomainDecomposer.hpp:192-223), devices[k] takes the k-th
 * contiguous block of them on its own host thread, the norm of a REL bound is
 * combined over all subdomains (calc_norm_decomposed,
 * ErrorToleranceCalculator.hpp:61-123) and every subdomain is compressed with the
 * same ABS bound.  The stream is what mgx_x_compress writes for that decomposition,
 * byte for byte: any single GPU (or stock MGARD-X) decodes it. */
int mgx_x_compress_multi(int D, int dtype, const uint64_t *shape, double tol, double s,
                         int mode, const void *in, void **out, size_t *out_size,
                         const int *devices, int num_devices, const mgx_config *cfg,
                         int output_pre_allocated);

/* decompress on SEVERAL GPUs of this process (counterpart of mgx_x_compress_multi):
 * devices[k] decodes the k-th contiguous block of subdomain records of a
 * domain-decomposed stream on its own host thread, straight into the caller's host
 * array.  `in` and `*out` are HOST buffers. */
int mgx_x_decompress_multi(const void *in, size_t in_size, void **out,
                           uint64_t *shape_out, int *D_out, int *dtype_out,
                           const int *devices, int num_devices, const mgx_config *cfg,
                           int output_pre_allocated);

int mgx_x_decompress(const void *in, size_t in_size, void **out,
                     uint64_t *shape_out, int *D_out, int *dtype_out,
                     const mgx_config *cfg, int output_pre_allocated);

int mgx_x_release_cache(const mgx_config *cfg);
int mgx_x_pin_memory(void *ptr, uint64_t bytes, const mgx_config *cfg);
int mgx_x_check_memory_pinned(void *ptr, const mgx_config *cfg);
int mgx_x_unpin_memory(void *ptr, const mgx_config *cfg);
const char *mgx_x_last_error(void);

/* 1 when libzstd.so.1 was found on this machine (lossless = Huffman_Zstd runs
 * ZSTD_compress / ZSTD_decompress on the host, Lossless/Zstd.hpp:72-141), else 0. */
int mgx_x_have_zstd(void);

/* lossless = Huffman_LZ4: the reference's portable LZ4 container
 * (Lossless/LZ4/LZ4.hpp:28-36: "MGXLZ4P\0" | n | chunk_size | nchunks |
 * comp_bytes[nchunks] | packed_bytes | packed[], every chunk an LZ4 block) written
 * and read on the host with liblz4.so.1 (dlopen).  mgx_x_have_lz4: 1 when the
 * library was found.  The two helpers are the stage on host buffers (out = NULL or
 * too small: MGX_OUTPUT_TOO_LARGE with the size needed in *out_bytes). */
int mgx_x_have_lz4(void);
int mgx_x_lz4_wrap(const uint8_t *in, uint64_t n, uint64_t chunk_size, uint8_t *out,
                   uint64_t cap, uint64_t *out_bytes);
int mgx_x_lz4_unwrap(const uint8_t *in, uint64_t bytes, uint8_t *out, uint64_t cap,
                     uint64_t *out_bytes);
/* kernels launched by the high-level path's contexts since the cache was made */
uint64_t mgx_x_launch_count(void);

/* ---- the "MGARD" header (host only, no GPU) -------------------------------- */
typedef struct mgx_metadata {
  uint32_t version[3];      /* Header.mgard_version (the reference writes the
                               FILE version here, Metadata.cpp:287-291)       */
  uint32_t file_version[3]; /* Header.file_format_version (left empty by it)  */
  uint32_t D;
  uint64_t shape[5];
  int nonuniform;           /* Domain.geometry == EXPLICIT_CUBE               */
  int dtype;                /* mgx_dtype                                      */
  int mode;                 /* mgx_error_bound                                */
  int linf;                 /* ErrorControl.norm == L_INFINITY                */
  double s, norm, tol;
  int decomposed;           /* DomainDecomposition.method != NOOP_METHOD      */
  int dd_type;              /* mgx_dd_type                                    */
  uint32_t dd_dim;
  uint64_t dd_size;
  int decomposition;        /* mgx_decomp_type                                */
  uint32_t l_target;
  int lossless;             /* mgx_lossless                                   */
  uint32_t huff_dict_size, huff_block_size, reorder;
  int backend;              /* mgard.pb.Device.Backend (3 = X_CUDA)           */
} mgx_metadata;

/* preamble (5 + 8 + 4 bytes) + proto3 Header.  Returns MGX_OUTPUT_TOO_LARGE
 * (with *bytes = needed) when cap is too small; out may be NULL to size. */
int mgx_metadata_serialize(const mgx_metadata *m, uint8_t *out, uint64_t cap,
                           uint64_t *bytes);
/* Checks signature, CRC32, version, shape/dimension agreement like
 * MetadataBase::Deserialize.  *header_bytes = offset of the first record. */
int mgx_metadata_parse(const uint8_t *in, uint64_t bytes, mgx_metadata *m,
                       uint64_t *header_bytes);
/* bytes of the whole header given its first 13 bytes (DeserializeSize) */
/* Non-uniform grids (Metadata.hpp:130-154, Metadata.cpp:303-321, 557-580): the
 * header of a stream compressed with node coordinates carries them as doubles
 * (Domain.geometry = EXPLICIT_CUBE), all dimensions back to back.
 * mgx_metadata_serialize_coords: as mgx_metadata_serialize, with m->nonuniform != 0
 * and coords -> sum(shape) doubles.  mgx_metadata_coords: copies them out of a
 * stream's header (*count = 0 for a uniform grid). */
int mgx_metadata_serialize_coords(const mgx_metadata *m, const double *coords,
                                  uint8_t *out, uint64_t cap, uint64_t *bytes);
int mgx_metadata_coords(const uint8_t *in, uint64_t bytes, double *coords,
                        uint64_t cap, uint64_t *count);
int mgx_metadata_size(const uint8_t *in, uint64_t bytes, uint64_t *header_bytes);
uint32_t mgx_crc32(const void *data, uint64_t bytes);

/* device bytes one subdomain of this shape needs in the high-level path (the
 * role of DomainDecomposer::EstimateMemoryFootprint, DomainDecomposer.hpp:23-69,
 * for THIS implementation's buffers) */
uint64_t mgx_estimate_footprint(int D, const uint64_t *shape, int dtype, double s,
                                int prefetch);
/* generate_max_dim_domain_decomposition_strategy (DomainDecomposer.hpp:192-223):
 * halve the largest dimension until one slab fits `available` bytes. */
int mgx_max_dim_by_memory(int D, const uint64_t *shape, int dtype, double s,
                          uint64_t available, int *decomposed, int *dim,
                          uint64_t *size, uint64_t *num);

#ifdef __cplusplus
}
#endif
#endif
