// TEST INFRASTRUCTURE ONLY -- never linked into the product library.
//
// Thin C-ABI wrapper around the UNMODIFIED reference sources under
// /root/reference (MGARD-X SERIAL / OPENMP back-ends).  This translation unit
// is compiled once per (REF_D, REF_T) pair by oracle/Makefile; the reference
// headers are used where they lie (nothing is copied into this repository).
// Output goes to oracle/_ref/libmgardx_ref.so, which is the parity oracle for
// tests/ and the "reference" CPU arm of bench.py.
//
// Entry points follow the stage boundaries of
//   include/mgard-x/CompressionLowLevel/Compressor.hpp:204-286
// so each GPU stage can be compared in isolation.
#include <cstdint>
#include <chrono>
#include <cmath>
#include <cstring>
#include <limits>
#include <vector>

#include "mgard-x/RuntimeX/RuntimeX.h"

#include "mgard-x/CompressionLowLevel/Compressor.hpp"
#include "mgard-x/Hierarchy/Hierarchy.hpp"

#include "mgard-x/DataRefactoring/MultiDimension/DataRefactoring.hpp"

#include "mgard-x/DataRefactoring/MultiDimension/Coefficient/CalcCoefficients3D.hpp"
#include "mgard-x/DataRefactoring/MultiDimension/Coefficient/CalcCoefficientsND.hpp"
#include "mgard-x/DataRefactoring/MultiDimension/Coefficient/CoefficientsRestore3D.hpp"
#include "mgard-x/DataRefactoring/MultiDimension/Coefficient/CoefficientsRestoreND.hpp"
#include "mgard-x/DataRefactoring/MultiDimension/CopyND/AddND.hpp"
#include "mgard-x/DataRefactoring/MultiDimension/CopyND/CopyND.hpp"
#include "mgard-x/DataRefactoring/MultiDimension/CopyND/SubtractND.hpp"
#include "mgard-x/DataRefactoring/MultiDimension/Correction/CalcCorrection3D.hpp"
#include "mgard-x/DataRefactoring/MultiDimension/Correction/CalcCorrectionND.hpp"

#include "mgard-x/DataRefactoring/SingleDimension/Coefficient/CalcCoefficients.hpp"
#include "mgard-x/DataRefactoring/SingleDimension/Coefficient/CoefficientsRestore.hpp"
#include "mgard-x/DataRefactoring/SingleDimension/Correction/CalcCorrection.hpp"
#include "mgard-x/DataRefactoring/SingleDimension/DataRefactoring.hpp"

#ifndef REF_D
#error "REF_D must be defined"
#endif
#ifndef REF_T
#error "REF_T must be defined"
#endif
#ifndef REF_SUFFIX
#error "REF_SUFFIX must be defined"
#endif

#define CAT_(a, b) a##b
#define CAT(a, b) CAT_(a, b)
#define FN(name) CAT(name, REF_SUFFIX)

using namespace mgard_x;

extern "C" int ref_get_lossless(void); // ref_wrap_huff.cpp: 0 Huffman, 1 Huffman_LZ4, 2 Huffman_Zstd

namespace {

constexpr DIM D = REF_D;
using T = REF_T;

// grid of the hierarchies built from now on (ref_set_grid_<D><t>): the caller's node
// coordinates -> Hierarchy(shape, coords, config), or uniform with / without
// Config::normalize_coordinates
bool g_nonuniform = false;
bool g_normalize = true;
std::vector<T *> g_coords(D, nullptr);

Config make_config(uint32_t dict_size, uint32_t chunk_size) {
  Config config;
  config.huff_dict_size = dict_size;
  config.huff_block_size = chunk_size;
  config.lossless = ref_get_lossless() == 2   ? lossless_type::Huffman_Zstd
                    : ref_get_lossless() == 1 ? lossless_type::Huffman_LZ4
                                              : lossless_type::Huffman;
  config.log_level = log::ERR;
  config.normalize_coordinates = g_normalize;
  return config;
}

template <typename Dev>
Hierarchy<D, T, Dev> make_hierarchy(std::vector<SIZE> shp, Config config) {
  if (g_nonuniform)
    return Hierarchy<D, T, Dev>(shp, g_coords, config);
  return Hierarchy<D, T, Dev>(shp, config);
}

std::vector<SIZE> make_shape(const uint64_t *shape) {
  std::vector<SIZE> s(D);
  for (DIM d = 0; d < D; d++)
    s[d] = shape[d];
  return s;
}

T s_value(double s) {
  return std::isinf(s) ? std::numeric_limits<T>::infinity() : (T)s;
}

template <typename Dev>
int do_compress(const uint64_t *shape, const T *in, int rel, double tol,
                double s, double *norm_io, uint32_t dict, uint32_t chunk,
                uint8_t *out, uint64_t cap, uint64_t *out_bytes) {
  DeviceRuntime<Dev>::Initialize();
  Config config = make_config(dict, chunk);
  std::vector<SIZE> shp = make_shape(shape);
  Hierarchy<D, T, Dev> hierarchy = make_hierarchy<Dev>(shp, config);
  Compressor<D, T, Dev> compressor;
  compressor.Adapt(hierarchy, config, 0);
  DeviceRuntime<Dev>::SyncQueue(0);
  Array<D, T, Dev> data(shp);
  data.load(in);
  T norm = (T)*norm_io;
  Array<1, Byte, Dev> compressed;
  compressor.Compress(data, rel ? error_bound_type::REL : error_bound_type::ABS,
                      (T)tol, s_value(s), norm, compressed, 0);
  DeviceRuntime<Dev>::SyncQueue(0);
  *norm_io = (double)norm;
  *out_bytes = compressed.shape(0);
  if (compressed.shape(0) > cap)
    return 2;
  std::memcpy(out, compressed.hostCopy(false, 0), compressed.shape(0));
  return 0;
}

template <typename Dev>
int do_decompress(const uint64_t *shape, const uint8_t *in, uint64_t bytes,
                  int rel, double tol, double s, double norm, uint32_t dict,
                  uint32_t chunk, T *out) {
  DeviceRuntime<Dev>::Initialize();
  Config config = make_config(dict, chunk);
  std::vector<SIZE> shp = make_shape(shape);
  Hierarchy<D, T, Dev> hierarchy = make_hierarchy<Dev>(shp, config);
  Compressor<D, T, Dev> compressor;
  compressor.Adapt(hierarchy, config, 0);
  DeviceRuntime<Dev>::SyncQueue(0);
  Array<1, Byte, Dev> compressed({(SIZE)bytes});
  compressed.load(in);
  Array<D, T, Dev> data(shp);
  T n = (T)norm;
  compressor.Decompress(compressed,
                        rel ? error_bound_type::REL : error_bound_type::ABS,
                        (T)tol, s_value(s), n, data, 0);
  DeviceRuntime<Dev>::SyncQueue(0);
  SIZE total = hierarchy.total_num_elems();
  std::memcpy(out, data.hostCopy(false, 0), total * sizeof(T));
  return 0;
}

// Timed run for bench.py: hierarchy + workspace set-up and the host<->Array
// copies stay outside the clock, like the plan / resident-input convention of
// the GPU arm; reps calls of Compress then reps calls of Decompress.
template <typename Dev>
int do_bench(const uint64_t *shape, const T *in, int rel, double tol, double s,
             int reps, double *compress_s, double *decompress_s,
             uint64_t *out_bytes, double *max_abs_err) {
  DeviceRuntime<Dev>::Initialize();
  Config config = make_config(8192, 1024);
  std::vector<SIZE> shp = make_shape(shape);
  Hierarchy<D, T, Dev> hierarchy = make_hierarchy<Dev>(shp, config);
  Compressor<D, T, Dev> compressor;
  compressor.Adapt(hierarchy, config, 0);
  DeviceRuntime<Dev>::SyncQueue(0);
  Array<D, T, Dev> data(shp);
  Array<1, Byte, Dev> compressed;
  Array<D, T, Dev> back(shp);
  T norm = 0;
  const error_bound_type eb =
      rel ? error_bound_type::REL : error_bound_type::ABS;
  for (int r = 0; r < reps; r++) {
    data.load(in); // Decompose works in place
    norm = 0;
    DeviceRuntime<Dev>::SyncQueue(0);
    auto t0 = std::chrono::steady_clock::now();
    compressor.Compress(data, eb, (T)tol, s_value(s), norm, compressed, 0);
    DeviceRuntime<Dev>::SyncQueue(0);
    auto t1 = std::chrono::steady_clock::now();
    compress_s[r] = std::chrono::duration<double>(t1 - t0).count();
  }
  *out_bytes = compressed.shape(0);
  for (int r = 0; r < reps; r++) {
    auto t0 = std::chrono::steady_clock::now();
    compressor.Decompress(compressed, eb, (T)tol, s_value(s), norm, back, 0);
    DeviceRuntime<Dev>::SyncQueue(0);
    auto t1 = std::chrono::steady_clock::now();
    decompress_s[r] = std::chrono::duration<double>(t1 - t0).count();
  }
  const T *b = back.hostCopy(false, 0);
  double e = 0;
  SIZE total = hierarchy.total_num_elems();
  for (SIZE i = 0; i < total; i++) {
    double d = std::fabs((double)b[i] - (double)in[i]);
    if (d > e)
      e = d;
  }
  *max_abs_err = e;
  return 0;
}

} // namespace

extern "C" {

// grid for the calls that follow: coords[d] -> shape[d] node coordinates (T), kept
// alive by the caller; NULL -> uniform grid
void FN(ref_set_grid_)(const void *const *coords, int normalize_coordinates) {
  g_nonuniform = coords != nullptr;
  g_normalize = normalize_coordinates != 0;
  for (DIM d = 0; d < D; d++)
    g_coords[d] = coords ? (T *)coords[d] : nullptr;
}

// DataRefactor::Decompose / Recompose (DataRefactoring/DataRefactor.hpp:73-130)
int FN(ref_decompose_)(const uint64_t *shape, T *data, int orthogonal) {
  using Dev = SERIAL;
  DeviceRuntime<Dev>::Initialize();
  Config config = make_config(8192, 1024);
  std::vector<SIZE> shp = make_shape(shape);
  Hierarchy<D, T, Dev> hierarchy = make_hierarchy<Dev>(shp, config);
  data_refactoring::DataRefactor<D, T, Dev> refactor(hierarchy, config);
  Array<D, T, Dev> arr(shp);
  arr.load(data);
  refactor.Decompose(SubArray(arr), orthogonal != 0, 0);
  DeviceRuntime<Dev>::SyncQueue(0);
  std::memcpy(data, arr.hostCopy(false, 0),
              hierarchy.total_num_elems() * sizeof(T));
  return 0;
}

int FN(ref_recompose_)(const uint64_t *shape, T *data, int orthogonal) {
  using Dev = SERIAL;
  DeviceRuntime<Dev>::Initialize();
  Config config = make_config(8192, 1024);
  std::vector<SIZE> shp = make_shape(shape);
  Hierarchy<D, T, Dev> hierarchy = make_hierarchy<Dev>(shp, config);
  data_refactoring::DataRefactor<D, T, Dev> refactor(hierarchy, config);
  Array<D, T, Dev> arr(shp);
  arr.load(data);
  refactor.Recompose(SubArray(arr), orthogonal != 0, 0);
  DeviceRuntime<Dev>::SyncQueue(0);
  std::memcpy(data, arr.hostCopy(false, 0),
              hierarchy.total_num_elems() * sizeof(T));
  return 0;
}

// LinearQuantizer::Quantize / Dequantize
// (Quantization/LinearQuantization.hpp:315-420)
int FN(ref_quantize_)(const uint64_t *shape, const T *coeff, int rel,
                      double tol, double s, double norm, uint32_t dict,
                      int64_t *codes) {
  using Dev = SERIAL;
  DeviceRuntime<Dev>::Initialize();
  Config config = make_config(dict, 1024);
  std::vector<SIZE> shp = make_shape(shape);
  Hierarchy<D, T, Dev> hierarchy = make_hierarchy<Dev>(shp, config);
  LinearQuantizer<D, T, QUANTIZED_INT, Dev> quantizer(hierarchy, config);
  Array<D, T, Dev> arr(shp);
  arr.load(coeff);
  Array<D, QUANTIZED_INT, Dev> q(shp);
  int dummy = 0;
  T sv = s_value(s);
  quantizer.Quantize(SubArray(arr),
                     rel ? error_bound_type::REL : error_bound_type::ABS,
                     (T)tol, sv, (T)norm, SubArray(q), dummy, 0,
                     infer_orthogonal_projection<D, T>(sv));
  DeviceRuntime<Dev>::SyncQueue(0);
  std::memcpy(codes, q.hostCopy(false, 0),
              hierarchy.total_num_elems() * sizeof(int64_t));
  return 0;
}

int FN(ref_dequantize_)(const uint64_t *shape, const int64_t *codes, int rel,
                        double tol, double s, double norm, uint32_t dict,
                        T *coeff) {
  using Dev = SERIAL;
  DeviceRuntime<Dev>::Initialize();
  Config config = make_config(dict, 1024);
  std::vector<SIZE> shp = make_shape(shape);
  Hierarchy<D, T, Dev> hierarchy = make_hierarchy<Dev>(shp, config);
  LinearQuantizer<D, T, QUANTIZED_INT, Dev> quantizer(hierarchy, config);
  Array<D, T, Dev> arr(shp);
  Array<D, QUANTIZED_INT, Dev> q(shp);
  q.load(codes);
  int dummy = 0;
  T sv = s_value(s);
  quantizer.Dequantize(SubArray(arr),
                       rel ? error_bound_type::REL : error_bound_type::ABS,
                       (T)tol, sv, (T)norm, SubArray(q), dummy, 0,
                       infer_orthogonal_projection<D, T>(sv));
  DeviceRuntime<Dev>::SyncQueue(0);
  std::memcpy(coeff, arr.hostCopy(false, 0),
              hierarchy.total_num_elems() * sizeof(T));
  return 0;
}

// Compressor::Compress / Decompress (CompressionLowLevel/Compressor.hpp:204-286)
// backend: 0 = SERIAL, 1 = OPENMP
int FN(ref_compress_)(int backend, const uint64_t *shape, const T *in, int rel,
                      double tol, double s, double *norm_io, uint32_t dict,
                      uint32_t chunk, uint8_t *out, uint64_t cap,
                      uint64_t *out_bytes) {
  try {
#if MGARD_ENABLE_OPENMP
    if (backend == 1)
      return do_compress<OPENMP>(shape, in, rel, tol, s, norm_io, dict, chunk,
                                 out, cap, out_bytes);
#endif
    return do_compress<SERIAL>(shape, in, rel, tol, s, norm_io, dict, chunk,
                               out, cap, out_bytes);
  } catch (...) {
    return 1;
  }
}

int FN(ref_decompress_)(int backend, const uint64_t *shape, const uint8_t *in,
                        uint64_t bytes, int rel, double tol, double s,
                        double norm, uint32_t dict, uint32_t chunk, T *out) {
  try {
#if MGARD_ENABLE_OPENMP
    if (backend == 1)
      return do_decompress<OPENMP>(shape, in, bytes, rel, tol, s, norm, dict,
                                   chunk, out);
#endif
    return do_decompress<SERIAL>(shape, in, bytes, rel, tol, s, norm, dict,
                                 chunk, out);
  } catch (...) {
    return 1;
  }
}

int FN(ref_bench_)(int backend, const uint64_t *shape, const T *in, int rel,
                   double tol, double s, int reps, double *compress_s,
                   double *decompress_s, uint64_t *out_bytes,
                   double *max_abs_err) {
  try {
#if MGARD_ENABLE_OPENMP
    if (backend == 1)
      return do_bench<OPENMP>(shape, in, rel, tol, s, reps, compress_s,
                              decompress_s, out_bytes, max_abs_err);
#endif
    return do_bench<SERIAL>(shape, in, rel, tol, s, reps, compress_s,
                            decompress_s, out_bytes, max_abs_err);
  } catch (...) {
    return 1;
  }
}

} // extern "C"
