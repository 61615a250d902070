// mgard_b200 -- the C++ face of the drop-in: namespace mgard_x with the
// reference's public high-level API (reference include/compress_x.hpp:31-178,
// include/mgard-x/Config/Config.h:10-46, include/mgard-x/Utilities/Types.h:19-85,
// include/mgard-x/RuntimeX/DataTypes.h:106-132).  Header-only: every function
// here marshals its arguments into the C ABI of include/mgard_b200_x.h, so a
// program written against the reference recompiles against this header and
// links -lmgard_b200.
//
// Served: uniform Cartesian grids, D <= 5, float / double, REL / ABS,
// L-inf / L2 / s-norm, Huffman lossless, MultiDim decomposition, MaxDim and Block
// domain decomposition, host or device buffers.  Anything else a Config can ask for
// (Huffman_Zstd / LZ4, SingleDim / Hybrid, Variable,
// ZFP, CPU back ends) is answered with the reference's failure codes -- there
// is no CPU fallback.
#ifndef MGARD_B200_COMPRESS_X_HPP
#define MGARD_B200_COMPRESS_X_HPP

#include <cstddef>
#include <cstdint>
#include <limits>
#include <vector>

#include "../mgard_b200_x.h"

namespace mgard_x {

using SIZE = uint64_t;
using DIM = uint8_t;
using Byte = unsigned char;

enum class device_type : uint8_t { AUTO, SERIAL, OPENMP, CUDA, HIP, SYCL, NONE };
enum class cpu_parallelization_mode : uint8_t { INTRA_BLOCK, INTER_BLOCK };
enum class decomposition_type : uint8_t { MultiDim, SingleDim, Hybrid };
enum class error_bound_type : uint8_t { REL, ABS };
enum class norm_type : uint8_t { L_Inf, L_2 };
enum class lossless_type : uint8_t {
  Huffman,
  Huffman_LZ4,
  Huffman_Zstd,
  CPU_Lossless,
  BlockDelta,
  LZ4,
  ZeroRLE_Rans,
  SymbolRans
};
enum class block_delta_mode_type : uint8_t { Fixed, Delta, Outlier };
enum class data_type : uint8_t { Float, Double };
enum class domain_decomposition_type : uint8_t { MaxDim, Block, Variable };
enum class compressor_type : uint8_t { MGARD, ZFP };
enum class compress_status_type : uint8_t {
  Success,
  Failure,
  OutputTooLargeFailure,
  NotSupportHigherNumberOfDimensionsFailure,
  NotSupportDataTypeFailure,
  BackendNotAvailableFailure
};

// Same members and defaults as the reference's Config (Config.cpp:31-66); the
// ones this path does not read are carried so user code that sets them builds.
struct Config {
  device_type dev_type = device_type::AUTO;
  int dev_id = 0;
  compressor_type compressor = compressor_type::MGARD;
  domain_decomposition_type domain_decomposition = domain_decomposition_type::MaxDim;
  decomposition_type decomposition = decomposition_type::MultiDim;
  double estimate_outlier_ratio = 1.0;
  SIZE huff_dict_size = 8192;
  SIZE huff_block_size = 1024;
  SIZE block_delta_block_size = 256;
  block_delta_mode_type block_delta_mode = block_delta_mode_type::Delta;
  SIZE lz4_block_size = 1 << 15;
  int zstd_compress_level = 3;
  bool normalize_coordinates = true;
  lossless_type lossless = lossless_type::Huffman;
  int reorder = 0;
  int log_level = 1;
  bool prefetch = false;
  bool auto_pin_host_buffers = true;
  SIZE max_larget_level = std::numeric_limits<SIZE>::max();
  SIZE max_memory_footprint = std::numeric_limits<SIZE>::max();
  SIZE total_num_bitplanes = 32;
  SIZE block_size = 256;
  SIZE domain_decomposition_dim = 0;
  std::vector<SIZE> domain_decomposition_sizes;
  bool mdr_adaptive_resolution = false;
  bool adjust_shape = false;
  bool compress_with_dryrun = false;
  int num_local_refactoring_level = 1;
  bool auto_cache_release = false;
  cpu_parallelization_mode cpu_mode = cpu_parallelization_mode::INTER_BLOCK;
  bool mdr_qoi_mode = false;
  int mdr_qoi_num_variables = 3;
  void apply() {}
};

namespace detail {
inline mgx_config to_c(const Config &c) {
  mgx_config m;
  mgx_config_default(&m);
  m.dev_type = (int)c.dev_type;
  m.dev_id = c.dev_id;
  m.domain_decomposition = (int)c.domain_decomposition;
  m.decomposition = (int)c.decomposition;
  // lossless types past CPU_Lossless have no C-side name: anything that is not
  // Huffman, Huffman_LZ4 or Huffman_Zstd is refused by the library
  m.lossless = (int)c.lossless;
  m.zstd_compress_level = c.zstd_compress_level;
  m.lz4_block_size = (uint32_t)c.lz4_block_size;
  m.huff_dict_size = (uint32_t)c.huff_dict_size;
  m.huff_block_size = (uint32_t)c.huff_block_size;
  m.reorder = (uint32_t)c.reorder;
  m.normalize_coordinates = c.normalize_coordinates;
  m.auto_pin_host_buffers = c.auto_pin_host_buffers;
  m.compress_with_dryrun = c.compress_with_dryrun;
  m.auto_cache_release = c.auto_cache_release;
  m.log_level = c.log_level;
  m.max_larget_level = c.max_larget_level;
  m.max_memory_footprint = c.max_memory_footprint;
  m.block_size = c.block_size;
  return m;
}
inline compress_status_type status(int rc) {
  return rc >= 0 && rc <= 5 ? (compress_status_type)rc : compress_status_type::Failure;
}
} // namespace detail

// ---- compress (compress_x.hpp:45-100) ----------------------------------------
inline compress_status_type
compress(DIM D, data_type dtype, std::vector<SIZE> shape, double tol, double s,
         error_bound_type mode, const void *original_data, void *&compressed_data,
         size_t &compressed_size, Config config, bool output_pre_allocated) {
  if (config.compressor != compressor_type::MGARD)
    return compress_status_type::Failure; // ZFP: external compressor not built
  if (shape.size() != D)
    return compress_status_type::Failure;
  const mgx_config c = detail::to_c(config);
  return detail::status(mgx_x_compress(
      D, dtype == data_type::Double ? MGX_F64 : dtype == data_type::Float ? MGX_F32 : -1,
      shape.data(), tol, s, (int)mode, original_data, &compressed_data,
      &compressed_size, &c, output_pre_allocated));
}

inline compress_status_type
compress(DIM D, data_type dtype, std::vector<SIZE> shape, double tol, double s,
         error_bound_type mode, const void *original_data, void *&compressed_data,
         size_t &compressed_size, bool output_pre_allocated) {
  return compress(D, dtype, shape, tol, s, mode, original_data, compressed_data,
                  compressed_size, Config(), output_pre_allocated);
}

// Non-uniform grids (coords overloads, compress_x.hpp:74-100): coords[d] points to
// shape[d] node coordinates of the field's data type.
inline compress_status_type
compress(DIM D, data_type dtype, std::vector<SIZE> shape, double tol, double s,
         error_bound_type mode, const void *original_data, void *&compressed_data,
         size_t &compressed_size, std::vector<const Byte *> coords, Config config,
         bool output_pre_allocated) {
  if (config.compressor != compressor_type::MGARD)
    return compress_status_type::Failure;
  if (shape.size() != D || coords.size() != D)
    return compress_status_type::Failure;
  const mgx_config c = detail::to_c(config);
  const void *cp[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};
  for (DIM d = 0; d < D && d < 5; d++)
    cp[d] = coords[d];
  return detail::status(mgx_x_compress_coords(
      D, dtype == data_type::Double ? MGX_F64 : dtype == data_type::Float ? MGX_F32 : -1,
      shape.data(), tol, s, (int)mode, original_data, &compressed_data,
      &compressed_size, cp, &c, output_pre_allocated));
}
inline compress_status_type
compress(DIM D, data_type dtype, std::vector<SIZE> shape, double tol, double s,
         error_bound_type mode, const void *original_data, void *&compressed_data,
         size_t &compressed_size, std::vector<const Byte *> coords,
         bool output_pre_allocated) {
  return compress(D, dtype, shape, tol, s, mode, original_data, compressed_data,
                  compressed_size, coords, Config(), output_pre_allocated);
}

// ---- decompress (compress_x.hpp:108-160) -----------------------------------------
inline compress_status_type
decompress(const void *compressed_data, size_t compressed_size,
           void *&decompressed_data, std::vector<SIZE> &shape, data_type &dtype,
           Config config, bool output_pre_allocated) {
  const mgx_config c = detail::to_c(config);
  uint64_t shp[5] = {0, 0, 0, 0, 0};
  int D = 0, dt = 0;
  const int rc = mgx_x_decompress(compressed_data, compressed_size, &decompressed_data,
                                  shp, &D, &dt, &c, output_pre_allocated);
  if (D > 0) {
    shape.assign(shp, shp + D);
    dtype = dt == MGX_F64 ? data_type::Double : data_type::Float;
  }
  return detail::status(rc);
}

inline compress_status_type
decompress(const void *compressed_data, size_t compressed_size,
           void *&decompressed_data, std::vector<SIZE> &shape, data_type &dtype,
           bool output_pre_allocated) {
  return decompress(compressed_data, compressed_size, decompressed_data, shape, dtype,
                    Config(), output_pre_allocated);
}

inline compress_status_type
decompress(const void *compressed_data, size_t compressed_size,
           void *&decompressed_data, Config config, bool output_pre_allocated) {
  std::vector<SIZE> shape;
  data_type dtype;
  return decompress(compressed_data, compressed_size, decompressed_data, shape, dtype,
                    config, output_pre_allocated);
}

inline compress_status_type
decompress(const void *compressed_data, size_t compressed_size,
           void *&decompressed_data, bool output_pre_allocated) {
  return decompress(compressed_data, compressed_size, decompressed_data, Config(),
                    output_pre_allocated);
}

// ---- cache and pinned memory (compress_x.hpp:166-176) -----------------------------
inline compress_status_type release_cache(Config config) {
  const mgx_config c = detail::to_c(config);
  return detail::status(mgx_x_release_cache(&c));
}
inline void pin_memory(void *ptr, SIZE num_bytes, Config config) {
  const mgx_config c = detail::to_c(config);
  mgx_x_pin_memory(ptr, num_bytes, &c);
}
inline bool check_memory_pinned(void *ptr, Config config) {
  const mgx_config c = detail::to_c(config);
  return mgx_x_check_memory_pinned(ptr, &c) != 0;
}
inline void unpin_memory(void *ptr, Config config) {
  const mgx_config c = detail::to_c(config);
  mgx_x_unpin_memory(ptr, &c);
}

} // namespace mgard_x

#endif
