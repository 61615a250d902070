// A user program written against the reference's high-level C++ API
// (namespace mgard_x; compare examples/mgard-x/HighLevelAPIs/Example.cpp:18-36
// and tests/mgard-x/src/test_Compressor.cpp.in:27-43 for the data), compiled
// against include/mgard_b200/compress_x.hpp and linked with libmgard_b200.so.
// Built by __graft_entry__.build(); run by tests/test_highlevel_gpu.py.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <limits>
#include <vector>

#include "mgard_b200/compress_x.hpp"

static int failures = 0;
#define EXPECT(cond)                                                           \
  do {                                                                         \
    if (!(cond)) {                                                             \
      std::printf("FAILED %s:%d  %s  (%s)\n", __FILE__, __LINE__, #cond,       \
                  mgx_x_last_error());                                         \
      failures++;                                                              \
    }                                                                          \
  } while (0)

template <typename T>
static std::vector<T> field(const std::vector<mgard_x::SIZE> &shape) {
  size_t n = 1;
  for (auto s : shape)
    n *= s;
  std::vector<T> v(n);
  for (size_t i = 0; i < n; i++) {
    size_t rest = i;
    double x = 1;
    for (int d = (int)shape.size() - 1; d >= 0; d--) {
      size_t idx = rest % shape[d];
      rest /= shape[d];
      x *= std::sin(2 * 3.14159265358979 * (double)idx / (double)shape[d] + 0.5 * d);
    }
    v[i] = (T)x;
  }
  return v;
}

template <typename T>
static void roundtrip(std::vector<mgard_x::SIZE> shape, mgard_x::data_type dt,
                      double tol, double s, mgard_x::error_bound_type mode,
                      mgard_x::Config config) {
  std::vector<T> in = field<T>(shape);
  void *compressed = nullptr;
  size_t compressed_size = 0;
  auto st = mgard_x::compress((mgard_x::DIM)shape.size(), dt, shape, tol, s, mode,
                              in.data(), compressed, compressed_size, config, false);
  EXPECT(st == mgard_x::compress_status_type::Success);
  if (st != mgard_x::compress_status_type::Success)
    return;
  EXPECT(compressed_size < in.size() * sizeof(T));

  void *out = nullptr;
  std::vector<mgard_x::SIZE> got_shape;
  mgard_x::data_type got_dt;
  st = mgard_x::decompress(compressed, compressed_size, out, got_shape, got_dt, config,
                           false);
  EXPECT(st == mgard_x::compress_status_type::Success);
  EXPECT(got_shape == shape);
  EXPECT(got_dt == dt);
  if (st == mgard_x::compress_status_type::Success) {
    const T *r = (const T *)out;
    double maxerr = 0, maxabs = 0, se = 0, su = 0;
    for (size_t i = 0; i < in.size(); i++) {
      double e = std::fabs((double)r[i] - (double)in[i]);
      maxerr = std::fmax(maxerr, e);
      maxabs = std::fmax(maxabs, std::fabs((double)in[i]));
      se += e * e;
      su += (double)in[i] * (double)in[i];
    }
    const bool rel = mode == mgard_x::error_bound_type::REL;
    if (std::isinf(s))
      EXPECT(maxerr <= tol * (rel ? maxabs : 1.0));
    else
      EXPECT(std::sqrt(se / in.size()) <= tol * (rel ? std::sqrt(su / in.size()) : 1.0));
    std::printf("  %zu-D %s  ratio %.2f  max err %.3g\n", shape.size(),
                sizeof(T) == 4 ? "float" : "double",
                (double)(in.size() * sizeof(T)) / compressed_size, maxerr);
    std::free(out);
  }
  std::free(compressed);
}

int main() {
  const double inf = std::numeric_limits<double>::infinity();
  mgard_x::Config config;
  config.lossless = mgard_x::lossless_type::Huffman;
  config.dev_type = mgard_x::device_type::CUDA;

  roundtrip<float>({65, 70, 130}, mgard_x::data_type::Float, 1e-3, inf,
                   mgard_x::error_bound_type::REL, config);
  roundtrip<double>({257, 300}, mgard_x::data_type::Double, 1e-4, 0,
                    mgard_x::error_bound_type::REL, config);
  roundtrip<float>({100000}, mgard_x::data_type::Float, 1e-2, inf,
                   mgard_x::error_bound_type::ABS, config);
  roundtrip<double>({9, 10, 12, 60}, mgard_x::data_type::Double, 1e-3, 1.0,
                    mgard_x::error_bound_type::ABS, config);

  // a memory budget that forces domain decomposition
  mgard_x::Config small = config;
  small.max_memory_footprint = 12u << 20;
  roundtrip<float>({200, 129, 65}, mgard_x::data_type::Float, 1e-3, inf,
                   mgard_x::error_bound_type::REL, small);

  // pre-allocated output that is too small, unsupported requests
  {
    std::vector<mgard_x::SIZE> shape = {33, 33, 33};
    std::vector<float> in = field<float>(shape);
    std::vector<unsigned char> buf(200);
    void *p = buf.data();
    size_t n = buf.size();
    EXPECT(mgard_x::compress(3, mgard_x::data_type::Float, shape, 1e-3, inf,
                             mgard_x::error_bound_type::REL, in.data(), p, n, config,
                             true) == mgard_x::compress_status_type::OutputTooLargeFailure);
    mgard_x::Config cpu = config;
    cpu.dev_type = mgard_x::device_type::SERIAL;
    void *q = nullptr;
    EXPECT(mgard_x::compress(3, mgard_x::data_type::Float, shape, 1e-3, inf,
                             mgard_x::error_bound_type::REL, in.data(), q, n, cpu,
                             false) == mgard_x::compress_status_type::BackendNotAvailableFailure);
    mgard_x::Config lz4 = config; // (Huffman, Huffman_LZ4, Huffman_Zstd are the types built)
    lz4.lossless = mgard_x::lossless_type::CPU_Lossless;
    EXPECT(mgard_x::compress(3, mgard_x::data_type::Float, shape, 1e-3, inf,
                             mgard_x::error_bound_type::REL, in.data(), q, n, lz4,
                             false) == mgard_x::compress_status_type::Failure);
    std::vector<mgard_x::SIZE> six = {3, 3, 3, 3, 3, 3};
    EXPECT(mgard_x::compress(6, mgard_x::data_type::Float, six, 1e-3, inf,
                             mgard_x::error_bound_type::REL, in.data(), q, n, config, false) ==
           mgard_x::compress_status_type::NotSupportHigherNumberOfDimensionsFailure);
  }

  // pinned-memory helpers and the cache
  {
    std::vector<float> buf(1 << 18);
    EXPECT(!mgard_x::check_memory_pinned(buf.data(), config));
    mgard_x::pin_memory(buf.data(), buf.size() * sizeof(float), config);
    EXPECT(mgard_x::check_memory_pinned(buf.data(), config));
    mgard_x::unpin_memory(buf.data(), config);
    EXPECT(!mgard_x::check_memory_pinned(buf.data(), config));
    EXPECT(mgard_x::release_cache(config) == mgard_x::compress_status_type::Success);
  }

  if (failures) {
    std::printf("%d FAILURES\n", failures);
    return 1;
  }
  std::printf("ALL OK\n");
  return 0;
}
