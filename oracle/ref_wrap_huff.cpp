// TEST INFRASTRUCTURE ONLY -- never linked into the product library.
//
// C-ABI wrapper around the UNMODIFIED reference Huffman stage
//   include/mgard-x/Lossless/ParallelHuffman/Huffman.hpp:553-632
// (SERIAL back-end, Q=uint64, S=int64, H=uint64 exactly as instantiated by
//  CompressionLowLevel/Compressor.h:47-48).  Compiled by oracle/Makefile into
// oracle/_ref/libmgardx_ref.so.
#include <dlfcn.h>
#include <cstdint>
#include <cstring>
#include <vector>

#include "mgard-x/RuntimeX/RuntimeX.h"

#include "mgard-x/Lossless/ParallelHuffman/Huffman.hpp"

using namespace mgard_x;

extern "C" {

// codes: shifted quantised integers (dictionary range is [0, dict)); values
// outside become outliers.  Returns the serialized "MGXHUFF" container.
int ref_huffman_compress(const int64_t *codes, uint64_t n, uint32_t dict,
                         uint32_t chunk, uint8_t *out, uint64_t cap,
                         uint64_t *out_bytes) {
  using Dev = SERIAL;
  try {
    DeviceRuntime<Dev>::Initialize();
    log::level = log::ERR;
    Huffman<QUANTIZED_UNSIGNED_INT, QUANTIZED_INT, HUFFMAN_CODE, Dev> huffman;
    huffman.Resize((SIZE)n, (int)dict, (int)chunk, 1.0, 0);
    DeviceRuntime<Dev>::SyncQueue(0);
    Array<1, QUANTIZED_INT, Dev> data({(SIZE)n});
    data.load(codes);
    Array<1, Byte, Dev> compressed;
    huffman.Compress(data, compressed, 0.0f, 0);
    huffman.Serialize(compressed, 0);
    DeviceRuntime<Dev>::SyncQueue(0);
    *out_bytes = compressed.shape(0);
    if (compressed.shape(0) > cap)
      return 2;
    std::memcpy(out, compressed.hostCopy(false, 0), compressed.shape(0));
    return 0;
  } catch (...) {
    return 1;
  }
}

int ref_huffman_decompress(const uint8_t *in, uint64_t bytes, uint64_t n,
                           uint32_t dict, uint32_t chunk, int64_t *codes) {
  using Dev = SERIAL;
  try {
    DeviceRuntime<Dev>::Initialize();
    log::level = log::ERR;
    Huffman<QUANTIZED_UNSIGNED_INT, QUANTIZED_INT, HUFFMAN_CODE, Dev> huffman;
    huffman.Resize((SIZE)n, (int)dict, (int)chunk, 1.0, 0);
    DeviceRuntime<Dev>::SyncQueue(0);
    Array<1, Byte, Dev> compressed({(SIZE)bytes});
    compressed.load(in);
    Array<1, QUANTIZED_INT, Dev> data({(SIZE)n});
    huffman.Deserialize(compressed, 0);
    huffman.Decompress(compressed, data, 0);
    DeviceRuntime<Dev>::SyncQueue(0);
    std::memcpy(codes, data.hostCopy(false, 0), n * sizeof(int64_t));
    return 0;
  } catch (...) {
    return 1;
  }
}

// zstd is referenced unconditionally by Lossless/Zstd.hpp and Lossless/CPU.hpp.  The
// image has the run-time library (libzstd.so.1) but not its headers: the five symbols
// forward to it through dlopen, so that the reference's own Huffman_Zstd writer and
// reader run in the oracle (ref_set_lossless(2)); without the library they report an
// error and the Huffman path, which never calls them, is unaffected.
static void *zstd_handle() {
  static void *h = dlopen("libzstd.so.1", RTLD_NOW | RTLD_LOCAL);
  return h;
}
static void *zstd_raw(const char *name) {
  void *h = zstd_handle();
  return h ? dlsym(h, name) : nullptr;
}
#define zstd_sym(F, name) ((F)zstd_raw(name))
size_t ZSTD_compress(void *d, size_t dc, const void *s, size_t sc, int level) {
  auto f = zstd_sym(size_t (*)(void *, size_t, const void *, size_t, int), "ZSTD_compress");
  return f ? f(d, dc, s, sc, level) : (size_t)-1;
}
size_t ZSTD_decompress(void *d, size_t dc, const void *s, size_t sc) {
  auto f = zstd_sym(size_t (*)(void *, size_t, const void *, size_t), "ZSTD_decompress");
  return f ? f(d, dc, s, sc) : (size_t)-1;
}
size_t ZSTD_compressBound(size_t n) {
  auto f = zstd_sym(size_t (*)(size_t), "ZSTD_compressBound");
  return f ? f(n) : n;
}
unsigned ZSTD_isError(size_t c) {
  auto f = zstd_sym(unsigned (*)(size_t), "ZSTD_isError");
  return f ? f(c) : 1u;
}
const char *ZSTD_getErrorName(size_t c) {
  auto f = zstd_sym(const char *(*)(size_t), "ZSTD_getErrorName");
  return f ? f(c) : "libzstd.so.1 not found";
}

// lossless_type of the Config the wrappers build (0 Huffman, 1 Huffman_LZ4, 2 Huffman_Zstd)
static int g_ref_lossless = 0;
void ref_set_lossless(int t) { g_ref_lossless = t; }
int ref_get_lossless(void) { return g_ref_lossless; }
int ref_have_zstd(void) { return zstd_handle() != nullptr; }

} // extern "C"
