// mgard_b200 -- the high-level path around the low-level compressor: what the
// reference's general_compress / general_decompress do with a whole dataset
// (CompressionHighLevel.hpp:47-309, 373-579; GPUPipelines.hpp:3-567;
// DomainDecomposer.hpp:118-330; Metadata.cpp:259-762).  Host code only: every
// array operation goes through the low-level C ABI (include/mgard_b200.h).
//
// Design notes (B200):
//   * one compute stream + one copy-in + one copy-out stream per device; the
//     next subdomain's H2D and the previous payload's D2H run under the current
//     subdomain's kernels (two input and two payload buffers);
//   * a device-resident input is never copied when its slab is contiguous
//     (MaxDim along dim 0, or no decomposition): the kernels read it in place;
//   * at most two planned contexts are kept (full slab + leftover), the same
//     working set the reference's hierarchy cache holds for MaxDim.
#include <cuda_runtime.h>
#include <dlfcn.h>

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <map>
#include <condition_variable>
#include <mutex>
#include <thread>
#include <string>
#include <vector>

#include "../../include/mgard_b200_x.h"

namespace {

std::string g_err;
std::mutex g_mu; // the reference's cache is process-global and not
                 // thread-safe; calls are serialised here instead

std::mutex g_err_mu; // mgx_x_compress_multi reports from several threads
int fail(int rc, const std::string &msg) {
  std::lock_guard<std::mutex> lock(g_err_mu);
  g_err = msg;
  return rc;
}

// ---- libzstd, bound at run time -----------------------------------------------
// The second lossless stage of lossless_type::Huffman_Zstd is the host library's
// ZSTD_compress / ZSTD_decompress over the serialised Huffman container
// (Lossless/Zstd.hpp:72-141).  The image ships libzstd.so.1 without its headers:
// the five entry points used are declared here and resolved with dlopen, and a
// machine without the library refuses Huffman_Zstd with an error text.
struct ZstdApi {
  size_t (*compressBound)(size_t) = nullptr;
  size_t (*compress)(void *, size_t, const void *, size_t, int) = nullptr;
  size_t (*decompress)(void *, size_t, const void *, size_t) = nullptr;
  unsigned (*isError)(size_t) = nullptr;
  const char *(*getErrorName)(size_t) = nullptr;
  bool ok = false;
};
const ZstdApi &zstd_api() {
  static ZstdApi api;
  static std::once_flag once;
  std::call_once(once, [] {
    void *h = dlopen("libzstd.so.1", RTLD_NOW | RTLD_LOCAL);
    if (!h)
      h = dlopen("libzstd.so", RTLD_NOW | RTLD_LOCAL);
    if (!h)
      return;
    api.compressBound = (size_t(*)(size_t))dlsym(h, "ZSTD_compressBound");
    api.compress =
        (size_t(*)(void *, size_t, const void *, size_t, int))dlsym(h, "ZSTD_compress");
    api.decompress =
        (size_t(*)(void *, size_t, const void *, size_t))dlsym(h, "ZSTD_decompress");
    api.isError = (unsigned (*)(size_t))dlsym(h, "ZSTD_isError");
    api.getErrorName = (const char *(*)(size_t))dlsym(h, "ZSTD_getErrorName");
    api.ok = api.compressBound && api.compress && api.decompress && api.isError &&
             api.getErrorName;
  });
  return api;
}

// ---- liblz4, bound at run time; the reference's portable LZ4 container -----------
// lossless_type::Huffman_LZ4 is, in this reference, its own nvcomp-free LZ4 stage
// (Lossless/LZ4/LZ4.hpp:28-36): "MGXLZ4P\0" | n | chunk_size | nchunks |
// comp_bytes[nchunks] | packed_bytes | packed[] (all size_t), every chunk of
// chunk_size input bytes an independent LZ4 BLOCK, decoded by output length
// (LZ4Kernels.hpp:202-262).  The blocks here come from liblz4's LZ4_compress_default
// and go through LZ4_decompress_safe: the same block format, not the same match
// finder -- the reference's decoder reads these blocks and this one reads the
// reference's (tests/test_oracle_vs_ref.py), the compressed sizes differ.
struct Lz4Api {
  int (*compress_default)(const char *, char *, int, int) = nullptr;
  int (*decompress_safe)(const char *, char *, int, int) = nullptr;
  int (*compressBound)(int) = nullptr;
  bool ok = false;
};
const Lz4Api &lz4_api() {
  static Lz4Api api;
  static std::once_flag once;
  std::call_once(once, [] {
    void *h = dlopen("liblz4.so.1", RTLD_NOW | RTLD_LOCAL);
    if (!h)
      h = dlopen("liblz4.so", RTLD_NOW | RTLD_LOCAL);
    if (!h)
      return;
    api.compress_default =
        (int (*)(const char *, char *, int, int))dlsym(h, "LZ4_compress_default");
    api.decompress_safe =
        (int (*)(const char *, char *, int, int))dlsym(h, "LZ4_decompress_safe");
    api.compressBound = (int (*)(int))dlsym(h, "LZ4_compressBound");
    api.ok = api.compress_default && api.decompress_safe && api.compressBound;
  });
  return api;
}
const uint8_t LZ4_SIGNATURE[8] = {'M', 'G', 'X', 'L', 'Z', '4', 'P', 0};
void put_u64_at(uint8_t *p, uint64_t v) {
  for (int k = 0; k < 8; k++)
    p[k] = (uint8_t)(v >> (8 * k));
}
uint64_t get_u64_at(const uint8_t *p) {
  uint64_t v = 0;
  for (int k = 0; k < 8; k++)
    v |= (uint64_t)p[k] << (8 * k);
  return v;
}
// -> out = the container of in[0, n)
int lz4_wrap(const uint8_t *in, uint64_t n, uint64_t chunk, std::vector<uint8_t> &out) {
  const Lz4Api &z = lz4_api();
  if (!z.ok)
    return fail(MGX_FAILURE, "Huffman_LZ4 needs liblz4.so.1 on this machine");
  if (n == 0 || chunk == 0 || chunk > 0x7E000000ull)
    return fail(MGX_FAILURE, "LZ4 stage: bad size or chunk size");
  const uint64_t nchunks = (n - 1) / chunk + 1;
  const uint64_t head = 8 + 8 * 3 + 8 * nchunks + 8;
  const uint64_t bound = (uint64_t)z.compressBound((int)chunk);
  out.resize(head + nchunks * bound);
  uint64_t packed = 0;
  for (uint64_t c = 0; c < nchunks; c++) {
    const uint64_t len = std::min<uint64_t>(chunk, n - c * chunk);
    const int got = z.compress_default((const char *)in + c * chunk,
                                       (char *)out.data() + head + packed, (int)len,
                                       (int)bound);
    if (got <= 0)
      return fail(MGX_FAILURE, "LZ4_compress_default failed");
    put_u64_at(out.data() + 32 + 8 * c, (uint64_t)got);
    packed += (uint64_t)got;
  }
  memcpy(out.data(), LZ4_SIGNATURE, 8);
  put_u64_at(out.data() + 8, n);
  put_u64_at(out.data() + 16, chunk);
  put_u64_at(out.data() + 24, nchunks);
  put_u64_at(out.data() + 32 + 8 * nchunks, packed);
  out.resize(head + packed);
  return MGX_SUCCESS;
}
// -> out = the bytes a container holds (at most `limit` of them)
int lz4_unwrap(const uint8_t *in, uint64_t bytes, uint64_t limit, std::vector<uint8_t> &out) {
  const Lz4Api &z = lz4_api();
  if (!z.ok)
    return fail(MGX_FAILURE, "Huffman_LZ4 needs liblz4.so.1 on this machine");
  if (bytes < 40 || memcmp(in, LZ4_SIGNATURE, 8) != 0)
    return fail(MGX_FAILURE, "LZ4 signature mismatch.");
  const uint64_t n = get_u64_at(in + 8), chunk = get_u64_at(in + 16),
                 nchunks = get_u64_at(in + 24);
  if (n == 0 || n > limit || chunk == 0 || chunk > 0x7E000000ull ||
      nchunks != (n - 1) / chunk + 1 || nchunks > (bytes - 40) / 8)
    return fail(MGX_FAILURE, "LZ4 container: bad header");
  const uint64_t head = 8 + 8 * 3 + 8 * nchunks + 8;
  const uint64_t packed = get_u64_at(in + 32 + 8 * nchunks);
  if (packed != bytes - head)
    return fail(MGX_FAILURE, "LZ4 container: size mismatch");
  out.resize(n);
  uint64_t pos = 0;
  for (uint64_t c = 0; c < nchunks; c++) {
    const uint64_t cb = get_u64_at(in + 32 + 8 * c);
    const uint64_t len = std::min<uint64_t>(chunk, n - c * chunk);
    if (cb == 0 || cb > packed - pos || cb > 0x7FFFFFFFull)
      return fail(MGX_FAILURE, "LZ4 container: bad chunk size");
    const int got = z.decompress_safe((const char *)in + head + pos,
                                      (char *)out.data() + c * chunk, (int)cb, (int)len);
    if (got < 0 || (uint64_t)got != len)
      return fail(MGX_FAILURE, "LZ4 container: a chunk does not decode to its length");
    pos += cb;
  }
  if (pos != packed)
    return fail(MGX_FAILURE, "LZ4 container: trailing bytes");
  return MGX_SUCCESS;
}

// ---- CRC-32 (IEEE 802.3, reflected 0xEDB88320) == zlib crc32_z --------------
uint32_t g_crc_table[256];
bool g_crc_ready = false;
void crc_init() {
  for (uint32_t i = 0; i < 256; i++) {
    uint32_t c = i;
    for (int k = 0; k < 8; k++)
      c = (c & 1) ? 0xEDB88320u ^ (c >> 1) : c >> 1;
    g_crc_table[i] = c;
  }
  g_crc_ready = true;
}

// ---- proto3 wire format -------------------------------------------------------
// Emission rules of the protobuf C++ serializer the reference links: fields in
// field-number order, scalar fields equal to their default are skipped (doubles
// compared by bit pattern), repeated scalars packed, sub-messages that were
// touched with mutable_*() are always present even when empty.
struct PbWriter {
  std::vector<uint8_t> b;
  void varint(uint64_t v) {
    while (v >= 0x80) {
      b.push_back((uint8_t)(v | 0x80));
      v >>= 7;
    }
    b.push_back((uint8_t)v);
  }
  void key(uint32_t field, uint32_t wire) { varint(((uint64_t)field << 3) | wire); }
  void put_uint(uint32_t field, uint64_t v) {
    if (v) {
      key(field, 0);
      varint(v);
    }
  }
  void put_double(uint32_t field, double d) {
    uint64_t bits;
    memcpy(&bits, &d, 8);
    if (bits) {
      key(field, 1);
      for (int i = 0; i < 8; i++)
        b.push_back((uint8_t)(bits >> (8 * i)));
    }
  }
  void put_message(uint32_t field, const PbWriter &m) {
    key(field, 2);
    varint(m.b.size());
    b.insert(b.end(), m.b.begin(), m.b.end());
  }
  void put_packed_double(uint32_t field, const double *v, size_t n) {
    if (!n)
      return;
    key(field, 2);
    varint(n * 8);
    for (size_t i = 0; i < n; i++) {
      uint64_t bits;
      memcpy(&bits, &v[i], 8);
      for (int k = 0; k < 8; k++)
        b.push_back((uint8_t)(bits >> (8 * k)));
    }
  }
  void put_packed_uint(uint32_t field, const uint64_t *v, size_t n) {
    if (!n)
      return;
    PbWriter t;
    for (size_t i = 0; i < n; i++)
      t.varint(v[i]);
    put_message(field, t);
  }
};

struct PbReader {
  const uint8_t *p, *e;
  bool varint(uint64_t &v) {
    v = 0;
    for (int shift = 0; shift < 64 && p < e; shift += 7) {
      uint8_t c = *p++;
      v |= (uint64_t)(c & 0x7F) << shift;
      if (!(c & 0x80))
        return true;
    }
    return false;
  }
  bool next(uint32_t &field, uint32_t &wire) {
    uint64_t k;
    if (!varint(k))
      return false;
    field = (uint32_t)(k >> 3);
    wire = (uint32_t)(k & 7);
    return field != 0;
  }
  bool fixed64(uint64_t &v) {
    if (e - p < 8)
      return false;
    v = 0;
    for (int i = 0; i < 8; i++)
      v |= (uint64_t)p[i] << (8 * i);
    p += 8;
    return true;
  }
  bool sub(PbReader &r) {
    uint64_t n;
    if (!varint(n) || (uint64_t)(e - p) < n)
      return false;
    r.p = p;
    r.e = p + n;
    p += n;
    return true;
  }
  bool skip(uint32_t wire) {
    uint64_t v;
    PbReader r;
    switch (wire) {
    case 0:
      return varint(v);
    case 1:
      return fixed64(v);
    case 2:
      return sub(r);
    case 5:
      if (e - p < 4)
        return false;
      p += 4;
      return true;
    default:
      return false;
    }
  }
  bool done() const { return p >= e; }
};

const char SIGNATURE[5] = {'M', 'G', 'A', 'R', 'D'};
const uint64_t PREAMBLE = 5 + 8 + 4;
const uint32_t SW_VERSION_MAJOR = 1;   // MGARD 1.6.0 (reference CMakeLists.txt:13-15)
const uint32_t FILE_VERSION_MAJOR = 1; // file format 1.0.0 (:17-19)

// mgard.pb enum numbers (src/mgard.proto)
enum { PB_ABSOLUTE = 0, PB_RELATIVE = 1 };
enum { PB_L_INFINITY = 0, PB_S_NORM = 1 };
enum { PB_NOOP_METHOD = 0, PB_MAX_DIMENSION = 1, PB_BLOCK = 2, PB_VARIABLE = 3 };
enum { PB_MULTIDIM = 1, PB_ONE_DIM = 2, PB_HYBRID = 3 };
enum { PB_COEFFICIENTWISE_LINEAR = 1, PB_INT64_T = 3 };
enum { PB_CPU_HUFFMAN_ZSTD = 2, PB_X_HUFFMAN = 3, PB_X_HUFFMAN_LZ4 = 4, PB_X_HUFFMAN_ZSTD = 5 };

void put_version(PbWriter &h, uint32_t field, const uint32_t v[3]) {
  PbWriter m;
  m.put_uint(1, v[0]);
  m.put_uint(2, v[1]);
  m.put_uint(3, v[2]);
  h.put_message(field, m);
}

// coords: node coordinates of all dimensions back to back (sum of shape doubles) for
// a non-uniform grid (Domain.geometry = EXPLICIT_CUBE, Metadata.cpp:303-321), else null
std::vector<uint8_t> header_bytes(const mgx_metadata &md, const double *coords = nullptr) {
  PbWriter h;
  put_version(h, 2, md.version);
  put_version(h, 3, md.file_version);
  { // Domain
    PbWriter grid, dom;
    grid.put_uint(1, md.D);
    grid.put_packed_uint(2, md.shape, md.D);
    dom.put_message(2, grid); // topology CARTESIAN_GRID = 0, geometry UNIT_CUBE = 0
    if (md.nonuniform && coords) {
      uint64_t total = 0;
      for (uint32_t d = 0; d < md.D; d++)
        total += md.shape[d];
      PbWriter cube;
      cube.put_packed_double(2, coords, total);
      dom.put_uint(3, 1); // EXPLICIT_CUBE
      dom.put_message(4, cube);
    }
    h.put_message(4, dom);
  }
  { // Dataset
    PbWriter m;
    m.put_uint(1, md.dtype == MGX_F64 ? 1 : 0);
    m.put_uint(2, 1);
    h.put_message(5, m);
  }
  { // ErrorControl
    PbWriter m;
    m.put_uint(1, md.mode == MGX_EB_REL ? PB_RELATIVE : PB_ABSOLUTE);
    m.put_uint(2, md.linf ? PB_L_INFINITY : PB_S_NORM);
    m.put_double(3, md.s);
    if (md.mode == MGX_EB_REL)
      m.put_double(4, md.norm);
    m.put_double(5, md.tol);
    h.put_message(6, m);
  }
  { // DomainDecomposition
    PbWriter m;
    uint32_t method = PB_NOOP_METHOD;
    if (md.decomposed)
      method = md.dd_type == MGX_DD_MAXDIM  ? PB_MAX_DIMENSION
               : md.dd_type == MGX_DD_BLOCK ? PB_BLOCK
                                            : PB_VARIABLE;
    m.put_uint(1, method);
    m.put_uint(2, md.dd_dim);
    m.put_uint(3, md.dd_size);
    h.put_message(7, m);
  }
  { // FunctionDecomposition
    PbWriter m;
    m.put_uint(2, md.decomposition == MGX_DC_MULTIDIM    ? PB_MULTIDIM
                  : md.decomposition == MGX_DC_SINGLEDIM ? PB_ONE_DIM
                                                         : PB_HYBRID);
    m.put_uint(3, md.l_target);
    h.put_message(8, m);
  }
  { // Quantization
    PbWriter m;
    m.put_uint(1, PB_COEFFICIENTWISE_LINEAR);
    m.put_uint(3, PB_INT64_T); // bin_widths PER_COEFFICIENT = 0, little endian
    h.put_message(9, m);
  }
  h.put_message(10, PbWriter()); // BitplaneEncoding: NOOP
  { // Encoding
    PbWriter m;
    m.put_uint(1, md.reorder ? 1 : 0);
    uint32_t comp = PB_X_HUFFMAN;
    if (md.lossless == MGX_LL_HUFFMAN_LZ4)
      comp = PB_X_HUFFMAN_LZ4;
    else if (md.lossless == MGX_LL_HUFFMAN_ZSTD)
      comp = PB_X_HUFFMAN_ZSTD;
    else if (md.lossless == MGX_LL_CPU)
      comp = PB_CPU_HUFFMAN_ZSTD;
    m.put_uint(2, comp);
    if (md.lossless != MGX_LL_CPU) {
      m.put_uint(3, md.huff_dict_size);
      m.put_uint(4, md.huff_block_size);
    }
    h.put_message(11, m);
  }
  { // Device
    PbWriter m;
    m.put_uint(1, (uint32_t)md.backend);
    h.put_message(12, m);
  }
  return h.b;
}

bool parse_version(PbReader r, uint32_t v[3]) {
  uint32_t f, w;
  while (!r.done()) {
    if (!r.next(f, w))
      return false;
    uint64_t x;
    if (w == 0 && f >= 1 && f <= 3) {
      if (!r.varint(x))
        return false;
      v[f - 1] = (uint32_t)x;
    } else if (!r.skip(w))
      return false;
  }
  return true;
}

// repeated uint64, packed or not
bool parse_grid(PbReader r, mgx_metadata &md, uint64_t &count) {
  uint32_t f, w;
  count = 0;
  while (!r.done()) {
    if (!r.next(f, w))
      return false;
    uint64_t x;
    if (f == 1 && w == 0) {
      if (!r.varint(x))
        return false;
      md.D = (uint32_t)x;
    } else if (f == 2 && w == 2) {
      PbReader s;
      if (!r.sub(s))
        return false;
      while (!s.done()) {
        if (!s.varint(x))
          return false;
        if (count < 5)
          md.shape[count] = x;
        count++;
      }
    } else if (f == 2 && w == 0) {
      if (!r.varint(x))
        return false;
      if (count < 5)
        md.shape[count] = x;
      count++;
    } else if (!r.skip(w))
      return false;
  }
  return true;
}

// generic scalar message: varint fields into u[field], fixed64 into d[field]
bool parse_scalars(PbReader r, uint64_t *u, double *d, uint32_t nf) {
  uint32_t f, w;
  while (!r.done()) {
    if (!r.next(f, w))
      return false;
    uint64_t x;
    if (w == 0) {
      if (!r.varint(x))
        return false;
      if (f < nf && u)
        u[f] = x;
    } else if (w == 1) {
      if (!r.fixed64(x))
        return false;
      if (f < nf && d)
        memcpy(&d[f], &x, 8);
    } else if (!r.skip(w))
      return false;
  }
  return true;
}

int parse_header(const uint8_t *p, uint64_t n, mgx_metadata &md,
                 std::vector<double> *coords = nullptr) {
  memset(&md, 0, sizeof(md));
  md.linf = 1; // proto3 default ErrorControl.norm == L_INFINITY
  md.mode = MGX_EB_ABS;
  PbReader r{p, p + n};
  uint32_t f, w;
  uint64_t shape_count = 0;
  uint32_t quant_method = 0, bitplane_method = 0, hierarchy = 0, compressor = 0;
  bool have_norm_kind = false;
  uint64_t norm_kind = 0;
  while (!r.done()) {
    if (!r.next(f, w))
      return fail(MGX_FAILURE, "header parsing encountered read or format error");
    if (w != 2) {
      if (!r.skip(w))
        return fail(MGX_FAILURE, "header parsing encountered read or format error");
      continue;
    }
    PbReader m;
    if (!r.sub(m))
      return fail(MGX_FAILURE, "header parsing encountered read or format error");
    bool ok = true;
    uint64_t u[8] = {};
    double d[8] = {};
    switch (f) {
    case 2:
      ok = parse_version(m, md.version);
      break;
    case 3:
      ok = parse_version(m, md.file_version);
      break;
    case 4: { // Domain
      uint32_t f2, w2;
      while (ok && !m.done()) {
        if (!m.next(f2, w2)) {
          ok = false;
          break;
        }
        if (f2 == 2 && w2 == 2) {
          PbReader g;
          ok = m.sub(g) && parse_grid(g, md, shape_count);
        } else if (f2 == 3 && w2 == 0) {
          uint64_t x;
          ok = m.varint(x);
          md.nonuniform = x == 1;
        } else if (f2 == 4 && w2 == 2) { // ExplicitCubeGeometry
          PbReader cube;
          ok = m.sub(cube);
          uint32_t f3, w3;
          while (ok && !cube.done()) {
            if (!cube.next(f3, w3)) {
              ok = false;
              break;
            }
            uint64_t bits;
            double v;
            if (f3 == 2 && w3 == 2) { // packed doubles
              PbReader pk;
              ok = cube.sub(pk);
              while (ok && !pk.done()) {
                ok = pk.fixed64(bits);
                memcpy(&v, &bits, 8);
                if (ok && coords)
                  coords->push_back(v);
              }
            } else if (f3 == 2 && w3 == 1) {
              ok = cube.fixed64(bits);
              memcpy(&v, &bits, 8);
              if (ok && coords)
                coords->push_back(v);
            } else
              ok = cube.skip(w3);
          }
        } else
          ok = m.skip(w2);
      }
      break;
    }
    case 5:
      ok = parse_scalars(m, u, nullptr, 8);
      md.dtype = u[1] == 1 ? MGX_F64 : MGX_F32;
      break;
    case 6:
      ok = parse_scalars(m, u, d, 8);
      md.mode = u[1] == PB_RELATIVE ? MGX_EB_REL : MGX_EB_ABS;
      norm_kind = u[2];
      have_norm_kind = true;
      md.s = d[3];
      md.norm = d[4];
      md.tol = d[5];
      break;
    case 7:
      ok = parse_scalars(m, u, nullptr, 8);
      md.decomposed = u[1] != PB_NOOP_METHOD;
      md.dd_type = u[1] == PB_BLOCK ? MGX_DD_BLOCK
                   : u[1] == PB_VARIABLE ? MGX_DD_VARIABLE
                                         : MGX_DD_MAXDIM;
      md.dd_dim = (uint32_t)u[2];
      md.dd_size = u[3];
      break;
    case 8:
      ok = parse_scalars(m, u, nullptr, 8);
      hierarchy = (uint32_t)u[2];
      md.l_target = (uint32_t)u[3];
      break;
    case 9:
      ok = parse_scalars(m, u, nullptr, 8);
      quant_method = (uint32_t)u[1];
      break;
    case 10:
      ok = parse_scalars(m, u, nullptr, 8);
      bitplane_method = (uint32_t)u[1];
      break;
    case 11:
      ok = parse_scalars(m, u, nullptr, 8);
      md.reorder = u[1] == 1;
      compressor = (uint32_t)u[2];
      md.huff_dict_size = (uint32_t)u[3];
      md.huff_block_size = (uint32_t)u[4];
      break;
    case 12:
      ok = parse_scalars(m, u, nullptr, 8);
      md.backend = (int)u[1];
      break;
    default:
      break;
    }
    if (!ok)
      return fail(MGX_FAILURE, "header parsing encountered read or format error");
  }
  (void)have_norm_kind;
  // the checks of MetadataBase::Deserialize (Metadata.cpp:519-760)
  if (md.version[0] > SW_VERSION_MAJOR)
    return fail(MGX_FAILURE, "MGARD version mismatch.");
  if (md.file_version[0] > FILE_VERSION_MAJOR)
    return fail(MGX_FAILURE, "MGARD file format version mismatch.");
  if (md.D != shape_count)
    return fail(MGX_FAILURE, "grid shape does not match given dimension.");
  if (md.nonuniform && coords) {
    uint64_t total = 0;
    for (uint32_t d = 0; d < md.D && d < 5; d++)
      total += md.shape[d];
    if (coords->size() != total)
      return fail(MGX_FAILURE,
                  "mismatch between number of node coordinates and grid shape.");
  }
  if (norm_kind == PB_L_INFINITY) {
    md.linf = 1;
    md.s = std::numeric_limits<double>::infinity();
  } else {
    md.linf = 0;
  }
  if (hierarchy == PB_MULTIDIM)
    md.decomposition = MGX_DC_MULTIDIM;
  else if (hierarchy == PB_ONE_DIM)
    md.decomposition = MGX_DC_SINGLEDIM;
  else if (hierarchy == PB_HYBRID)
    md.decomposition = MGX_DC_HYBRID;
  else
    return fail(MGX_FAILURE, "this decomposition hierarchy mismatch the hierarchy "
                             "used in MGARD-X.");
  if ((quant_method != 0) == (bitplane_method != 0))
    return fail(MGX_FAILURE, "cannot determine whether this is compressed or "
                             "refactored data.");
  if (compressor == PB_X_HUFFMAN)
    md.lossless = MGX_LL_HUFFMAN;
  else if (compressor == PB_X_HUFFMAN_LZ4)
    md.lossless = MGX_LL_HUFFMAN_LZ4;
  else if (compressor == PB_X_HUFFMAN_ZSTD)
    md.lossless = MGX_LL_HUFFMAN_ZSTD;
  else if (compressor == PB_CPU_HUFFMAN_ZSTD)
    md.lossless = MGX_LL_CPU;
  else
    return fail(MGX_FAILURE, "unknown lossless compressor type.");
  if (md.backend == 0)
    return fail(MGX_FAILURE, "this data was not compressed with MGARD-X.");
  return MGX_SUCCESS;
}

// ---- where does a pointer live -----------------------------------------------
struct PtrLoc {
  bool device = false;
  int dev = 0;
};
PtrLoc locate(const void *p) {
  PtrLoc loc;
  cudaPointerAttributes at;
  if (cudaPointerGetAttributes(&at, p) == cudaSuccess) {
    if (at.type == cudaMemoryTypeDevice || at.type == cudaMemoryTypeManaged) {
      loc.device = true;
      loc.dev = at.device;
    }
  } else {
    cudaGetLastError(); // plain host memory on old drivers: clear the error
  }
  return loc;
}

bool host_registered(const void *p) {
  cudaPointerAttributes at;
  if (cudaPointerGetAttributes(&at, p) != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  return at.type == cudaMemoryTypeHost;
}

// pin for the duration of a call (CompressionHighLevel.hpp:157-183)
struct ScopedPin {
  void *p = nullptr;
  void pin(const void *ptr, size_t bytes, bool enabled) {
    if (!enabled || !ptr || !bytes || host_registered(ptr))
      return;
    if (cudaHostRegister((void *)ptr, bytes, cudaHostRegisterDefault) == cudaSuccess)
      p = (void *)ptr;
    else
      cudaGetLastError(); // pageable copies still work, only slower
  }
  ~ScopedPin() {
    if (p)
      cudaHostUnregister(p);
  }
};

// ---- per-device cache (role of CompressorCache.hpp:25-147) --------------------
struct PlanEntry {
  std::vector<uint64_t> shape;
  int dtype;
  uint32_t dict, chunk;
  uint64_t max_level;
  mgx_ctx *ctx;
  uint64_t stamp;
  bool normalize = true;
  bool nonuniform = false; // never reused (Hierarchy::can_reuse, Hierarchy.hpp:797-807)
};

struct DeviceCache {
  int dev = 0;
  cudaStream_t compute = nullptr, copy_in = nullptr, copy_out = nullptr;
  cudaEvent_t ev_in[2] = {}, ev_out[2] = {}, ev_free[2] = {};
  std::vector<PlanEntry> plans;
  void *sub[2] = {nullptr, nullptr};
  size_t sub_bytes[2] = {0, 0};
  void *cbuf[2] = {nullptr, nullptr};
  size_t cbuf_bytes[2] = {0, 0};
  std::vector<uint8_t> zin, zout[2]; // host side of the Zstd stage
  uint64_t stamp = 0;
  uint64_t retired_launches = 0;
  bool ready = false;

  size_t held() const {
    return sub_bytes[0] + sub_bytes[1] + cbuf_bytes[0] + cbuf_bytes[1];
  }
};

std::map<int, DeviceCache> g_cache;

#define X_CUDA(call)                                                           \
  do {                                                                         \
    cudaError_t e_ = (call);                                                   \
    if (e_ != cudaSuccess)                                                     \
      return fail(MGX_FAILURE, std::string(#call) + ": " + cudaGetErrorString(e_)); \
  } while (0)

int cache_get(int dev, DeviceCache **out) {
  int count = 0;
  if (cudaGetDeviceCount(&count) != cudaSuccess || dev < 0 || dev >= count) {
    cudaGetLastError();
    return fail(MGX_BACKEND_UNAVAILABLE, "no CUDA device " + std::to_string(dev));
  }
  X_CUDA(cudaSetDevice(dev));
  DeviceCache &c = g_cache[dev];
  if (!c.ready) {
    c.dev = dev;
    X_CUDA(cudaStreamCreateWithFlags(&c.compute, cudaStreamNonBlocking));
    X_CUDA(cudaStreamCreateWithFlags(&c.copy_in, cudaStreamNonBlocking));
    X_CUDA(cudaStreamCreateWithFlags(&c.copy_out, cudaStreamNonBlocking));
    for (int i = 0; i < 2; i++) {
      X_CUDA(cudaEventCreateWithFlags(&c.ev_in[i], cudaEventDisableTiming));
      X_CUDA(cudaEventCreateWithFlags(&c.ev_out[i], cudaEventDisableTiming));
      X_CUDA(cudaEventCreateWithFlags(&c.ev_free[i], cudaEventDisableTiming));
    }
    c.ready = true;
  }
  *out = &c;
  return MGX_SUCCESS;
}

void cache_release(DeviceCache &c) {
  if (!c.ready)
    return;
  cudaSetDevice(c.dev);
  cudaDeviceSynchronize();
  for (auto &p : c.plans) {
    c.retired_launches += mgx_launch_count(p.ctx);
    mgx_destroy(p.ctx);
  }
  c.plans.clear();
  for (int i = 0; i < 2; i++) {
    if (c.sub[i])
      cudaFree(c.sub[i]);
    if (c.cbuf[i])
      cudaFree(c.cbuf[i]);
    c.sub[i] = c.cbuf[i] = nullptr;
    c.sub_bytes[i] = c.cbuf_bytes[i] = 0;
    cudaEventDestroy(c.ev_in[i]);
    cudaEventDestroy(c.ev_out[i]);
    cudaEventDestroy(c.ev_free[i]);
  }
  cudaStreamDestroy(c.compute);
  cudaStreamDestroy(c.copy_in);
  cudaStreamDestroy(c.copy_out);
  c.ready = false;
}

// coords: node coordinates of THIS (sub)domain per dimension (host, type T), or null
int cache_ctx(DeviceCache &c, int D, const uint64_t *shape, int dtype,
              const mgx_config &cfg, mgx_ctx **out,
              const void *const *coords = nullptr) {
  const bool normalize = cfg.normalize_coordinates != 0;
  if (!coords)
    for (auto &p : c.plans)
      if (!p.nonuniform && p.normalize == normalize && p.dtype == dtype &&
          p.dict == cfg.huff_dict_size && p.chunk == cfg.huff_block_size &&
          p.max_level == cfg.max_larget_level && p.shape.size() == (size_t)D &&
          std::equal(p.shape.begin(), p.shape.end(), shape)) {
        p.stamp = ++c.stamp;
        *out = p.ctx;
        return MGX_SUCCESS;
      }
  if (c.plans.size() >= 2) { // evict the least recently used plan
    size_t victim = c.plans[0].stamp < c.plans[1].stamp ? 0 : 1;
    c.retired_launches += mgx_launch_count(c.plans[victim].ctx);
    mgx_destroy(c.plans[victim].ctx);
    c.plans.erase(c.plans.begin() + victim);
  }
  mgx_ctx *ctx = nullptr;
  int rc = mgx_create(&ctx, c.dev, c.compute);
  if (rc)
    return fail(rc, "mgx_create failed");
  rc = mgx_plan_grid(ctx, D, shape, dtype, cfg.huff_dict_size, cfg.huff_block_size,
                     cfg.max_larget_level, coords, normalize ? 1 : 0);
  if (rc) {
    std::string msg = mgx_last_error(ctx);
    mgx_destroy(ctx);
    return fail(rc, "mgx_plan: " + msg);
  }
  PlanEntry e{std::vector<uint64_t>(shape, shape + D), dtype, cfg.huff_dict_size,
              cfg.huff_block_size, cfg.max_larget_level, ctx, ++c.stamp};
  e.normalize = normalize;
  e.nonuniform = coords != nullptr;
  c.plans.push_back(e);
  *out = ctx;
  return MGX_SUCCESS;
}

// coordinates of subdomain i: the decomposed dimension's array starts at the slab's
// first row (DomainDecomposer.hpp:259-289)
struct SubCoords {
  const void *p[5];
  // offs: first node of the subdomain per dimension
  const void *const *get(const void *const *coords, int D, const uint64_t *offs,
                         size_t es) {
    if (!coords)
      return nullptr;
    for (int d = 0; d < D; d++)
      p[d] = (const void *)((const char *)coords[d] + offs[d] * es);
    return p;
  }
};

int ensure_buffer(void *&p, size_t &have, size_t want) {
  if (have >= want)
    return MGX_SUCCESS;
  if (p)
    cudaFree(p);
  p = nullptr;
  have = 0;
  X_CUDA(cudaMalloc(&p, want));
  have = want;
  return MGX_SUCCESS;
}

// ---- the decomposition of one call ----------------------------------------------
struct Slabs {
  int D = 0;
  uint64_t shape[5] = {};
  size_t es = 4;
  bool decomposed = false;
  int type = MGX_DD_MAXDIM; // MGX_DD_MAXDIM: slabs along `dim`; MGX_DD_BLOCK: cubes
  int dim = 0;
  uint64_t size = 0, num = 1; // rows per slab / edge of a block; subdomain count
  uint64_t outer = 1, inner = 1; // MaxDim: products of the dims before / after `dim`
  uint64_t nb[5] = {1, 1, 1, 1, 1}; // Block: blocks per dimension

  bool block() const { return decomposed && type == MGX_DD_BLOCK; }
  void finish() {
    outer = inner = 1;
    if (block()) {
      num = 1;
      for (int d = 0; d < D; d++) {
        nb[d] = (shape[d] - 1) / size + 1;
        num *= nb[d];
      }
      return;
    }
    for (int d = 0; d < dim; d++)
      outer *= shape[d];
    for (int d = dim + 1; d < D; d++)
      inner *= shape[d];
  }
  // shape of subdomain `id` and its offset per dimension
  // (DomainDecomposer::subdomain_shape / dim_subdomain_offset, :95-160)
  void sub_box(uint64_t id, uint64_t *out, uint64_t *offs) const {
    for (int d = 0; d < D; d++) {
      out[d] = shape[d];
      offs[d] = 0;
    }
    if (!decomposed)
      return;
    if (block()) {
      for (int d = D - 1; d >= 0; d--) { // last dimension fastest
        const uint64_t bid = id % nb[d];
        id /= nb[d];
        offs[d] = bid * size;
        out[d] = bid < shape[d] / size ? size : shape[d] % size;
      }
      return;
    }
    uint64_t off = 0;
    mgx_subdomain_shape(D, shape, dim, size, id, out, &off);
    offs[dim] = off;
  }
  // MaxDim view: row offset along `dim`
  void sub_shape(uint64_t id, uint64_t *out, uint64_t *off) const {
    uint64_t offs[5];
    sub_box(id, out, offs);
    *off = block() ? 0 : offs[dim];
  }
  uint64_t elems(uint64_t id) const {
    uint64_t s[5], offs[5], n = 1;
    sub_box(id, s, offs);
    for (int d = 0; d < D; d++)
      n *= s[d];
    return n;
  }
  // a subdomain is one contiguous range of the dense array
  bool contiguous() const { return !decomposed || (!block() && outer == 1); }
};

// DomainDecomposer::copy_subdomain (DomainDecomposer.hpp:628-820) for MaxDim:
// `whole` is the dense array, `slab` a dense sub-array; either side may be host
// or device memory.
int copy_slab(const Slabs &sl, uint64_t id, void *slab, void *whole, bool to_slab,
              cudaStream_t stream) {
  if (sl.block()) {
    // a D-dimensional box: one pitched copy per index of the dimensions before the
    // last two (DomainDecomposer.hpp:714-790)
    uint64_t b[5], o[5];
    sl.sub_box(id, b, o);
    const int D = sl.D;
    const size_t width = (size_t)b[D - 1] * sl.es;
    const size_t pitch = (size_t)sl.shape[D - 1] * sl.es;
    const uint64_t rows = D >= 2 ? b[D - 2] : 1;
    uint64_t planes = 1;
    for (int d = 0; d + 2 < D; d++)
      planes *= b[d];
    for (uint64_t p = 0; p < planes; p++) {
      // multi-index of plane p over dims 0 .. D-3, row-major
      uint64_t rem = p, woff = 0;
      uint64_t idxs[5] = {0, 0, 0, 0, 0};
      for (int d = D - 3; d >= 0; d--) {
        idxs[d] = rem % b[d];
        rem /= b[d];
      }
      uint64_t st = 1;
      for (int d = D - 1; d >= 0; d--) {
        const uint64_t i = o[d] + (d + 2 < D ? idxs[d] : 0);
        woff += i * st;
        st *= sl.shape[d];
      }
      char *w = (char *)whole + (size_t)woff * sl.es;
      char *a = (char *)slab + (size_t)p * rows * width;
      if (to_slab)
        X_CUDA(cudaMemcpy2DAsync(a, width, w, pitch, width, rows, cudaMemcpyDefault,
                                 stream));
      else
        X_CUDA(cudaMemcpy2DAsync(w, pitch, a, width, width, rows, cudaMemcpyDefault,
                                 stream));
    }
    return MGX_SUCCESS;
  }
  uint64_t s[5], off;
  sl.sub_shape(id, s, &off);
  const size_t width = (size_t)s[sl.dim] * sl.inner * sl.es;
  const size_t pitch = (size_t)sl.shape[sl.dim] * sl.inner * sl.es;
  char *w = (char *)whole + (size_t)off * sl.inner * sl.es;
  if (sl.outer == 1 || !sl.decomposed) {
    if (to_slab)
      X_CUDA(cudaMemcpyAsync(slab, w, width, cudaMemcpyDefault, stream));
    else
      X_CUDA(cudaMemcpyAsync(w, slab, width, cudaMemcpyDefault, stream));
    return MGX_SUCCESS;
  }
  if (pitch < ((size_t)1 << 31)) {
    if (to_slab)
      X_CUDA(cudaMemcpy2DAsync(slab, width, w, pitch, width, sl.outer,
                               cudaMemcpyDefault, stream));
    else
      X_CUDA(cudaMemcpy2DAsync(w, pitch, slab, width, width, sl.outer,
                               cudaMemcpyDefault, stream));
    return MGX_SUCCESS;
  }
  for (uint64_t o = 0; o < sl.outer; o++) { // rows too long for a pitched copy
    char *a = (char *)slab + o * width, *b = w + o * pitch;
    if (to_slab)
      X_CUDA(cudaMemcpyAsync(a, b, width, cudaMemcpyDefault, stream));
    else
      X_CUDA(cudaMemcpyAsync(b, a, width, cudaMemcpyDefault, stream));
  }
  return MGX_SUCCESS;
}

bool uses_correction(int D, double s) { return !(std::isinf(s) && D <= 3); }

uint64_t footprint(int D, const uint64_t *shape, int dtype, double s, int prefetch) {
  const uint64_t es = dtype == MGX_F64 ? 8 : 4;
  uint64_t n = 1;
  for (int d = 0; d < D; d++)
    n *= shape[d];
  // subdomain buffer + payload staging buffer (a payload larger than the raw
  // data is replaced by the raw data, so N*es + container overhead bounds it)
  uint64_t io = n * es + (n * es + (n / 1024 + 1) * 16 + (1u << 17));
  if (prefetch)
    io *= 2;
  // context: 16-bit symbols, chunk tables, outlier list, and for the
  // orthogonal-projection path the level workspace + value scratch +
  // correction ping/pong buffers
  uint64_t ws = n * 2 + (n / 256 + 1) * 24 + n / 8 + (1u << 18);
  if (uses_correction(D, s))
    ws += n * es * 2 + (n * es * 3) / 4;
  else if (D == 3) {
    // level-by-level inverse: compact nodal arrays of levels L-1 and L-2
    uint64_t c1 = es, c2 = es;
    for (int d = 0; d < D; d++) {
      const uint64_t h = shape[d] / 2 + 1;
      c1 *= h;
      c2 *= h / 2 + 1;
    }
    ws += c1 + c2 + (1u << 16);
  } else
    ws += (n * es) / 256 + (1u << 16); // coarse lattice of the fused inverse
  return io + ws;
}

int decide_slabs(int D, const uint64_t *shape, int dtype, double s,
                 uint64_t available, Slabs *sl, const mgx_config *cfg = nullptr) {
  sl->D = D;
  sl->es = dtype == MGX_F64 ? 8 : 4;
  for (int d = 0; d < D; d++)
    sl->shape[d] = shape[d];
  if (cfg && cfg->domain_decomposition == MGX_DD_BLOCK) {
    // generate_block_domain_decomposition_strategy (DomainDecomposer.hpp:226-251):
    // cubes of Config::block_size, halved (rounding up) until one fits the device
    uint64_t size = cfg->block_size ? cfg->block_size : 256;
    for (;;) {
      uint64_t chunk[5], n = 1;
      for (int d = 0; d < D; d++) {
        chunk[d] = size;
        n *= (shape[d] - 1) / size + 1;
      }
      if (footprint(D, chunk, dtype, s, n > 1) < available)
        break;
      if (size <= 1)
        return fail(MGX_FAILURE, "dataset cannot be decomposed to fit the device");
      size = (size - 1) / 2 + 1;
    }
    sl->decomposed = true;
    sl->type = MGX_DD_BLOCK;
    sl->dim = 0;
    sl->size = size;
    sl->finish();
    for (uint64_t i = 0; i < sl->num; i++) {
      uint64_t b[5], o[5];
      sl->sub_box(i, b, o);
      for (int d = 0; d < D; d++)
        if (b[d] < 3)
          return fail(MGX_FAILURE, "block decomposition leaves a block with fewer "
                                   "than 3 nodes along a dimension");
    }
    return MGX_SUCCESS;
  }
  int dec = 0, dim = 0;
  uint64_t size = 0, num = 1;
  int rc = mgx_max_dim_by_memory(D, shape, dtype, s, available, &dec, &dim, &size, &num);
  if (rc)
    return fail(rc, "dataset cannot be decomposed to fit the device");
  sl->decomposed = dec != 0;
  sl->dim = dim;
  sl->size = size;
  sl->num = num;
  sl->finish();
  return MGX_SUCCESS;
}

int check_config(const mgx_config &cfg) {
  if (cfg.dev_type != MGX_DEV_AUTO && cfg.dev_type != MGX_DEV_CUDA)
    return fail(MGX_BACKEND_UNAVAILABLE,
                "only the CUDA backend is built (no CPU fallback by design)");
  if (cfg.lossless != MGX_LL_HUFFMAN && cfg.lossless != MGX_LL_HUFFMAN_ZSTD &&
      cfg.lossless != MGX_LL_HUFFMAN_LZ4)
    return fail(MGX_FAILURE,
                "lossless type not supported: Huffman, Huffman_LZ4 and Huffman_Zstd only");
  if (cfg.lossless == MGX_LL_HUFFMAN_ZSTD && !zstd_api().ok)
    return fail(MGX_FAILURE, "Huffman_Zstd needs libzstd.so.1 on this machine");
  if (cfg.lossless == MGX_LL_HUFFMAN_LZ4 && !lz4_api().ok)
    return fail(MGX_FAILURE, "Huffman_LZ4 needs liblz4.so.1 on this machine");
  if (cfg.lossless == MGX_LL_HUFFMAN_LZ4 && cfg.lz4_block_size == 0)
    return fail(MGX_FAILURE, "lz4_block_size must be non-zero");
  if (cfg.decomposition != MGX_DC_MULTIDIM)
    return fail(MGX_FAILURE, "decomposition type not supported: MultiDim only");
  if (cfg.domain_decomposition != MGX_DD_MAXDIM &&
      cfg.domain_decomposition != MGX_DD_BLOCK)
    return fail(MGX_FAILURE, "domain decomposition not supported: MaxDim and Block only");
  if (cfg.huff_dict_size == 0 || cfg.huff_block_size == 0)
    return fail(MGX_FAILURE, "huff_dict_size / huff_block_size must be non-zero");
  return MGX_SUCCESS;
}

uint64_t available_bytes(const DeviceCache &c, const mgx_config &cfg) {
  size_t free_b = 0, total_b = 0;
  if (cudaMemGetInfo(&free_b, &total_b) != cudaSuccess) {
    cudaGetLastError();
    free_b = 0;
  }
  uint64_t avail = (uint64_t)free_b + c.held();
  return avail < cfg.max_memory_footprint ? avail : cfg.max_memory_footprint;
}

void put_u64_le(uint8_t *p, uint64_t v) {
  for (int i = 0; i < 8; i++)
    p[i] = (uint8_t)(v >> (8 * i));
}

// ---- compress ------------------------------------------------------------------
int x_compress(int D, int dtype, const uint64_t *shape, double tol, double s,
               int mode, const void *in, void **out, size_t *out_size,
               const mgx_config &cfg, int prealloc,
               const void *const *coords = nullptr) {
  if (D > 5)
    return fail(MGX_TOO_MANY_DIMENSIONS, "more than 5 dimensions");
  if (dtype != MGX_F32 && dtype != MGX_F64)
    return fail(MGX_BAD_DATA_TYPE, "data type not supported");
  if (D < 1 || !shape || !in || !out || !out_size)
    return fail(MGX_FAILURE, "bad argument");
  int rc = check_config(cfg);
  if (rc)
    return rc;
  const size_t es = dtype == MGX_F64 ? 8 : 4;
  // the templated reference entry takes tol and s as T (CompressionHighLevel.hpp:311)
  const double tolT = dtype == MGX_F32 ? (double)(float)tol : tol;
  const double sT = dtype == MGX_F32 ? (double)(float)s : s;
  const int rel = mode == MGX_EB_REL;
  const int linf = std::isinf(sT) ? 1 : 0;
  uint64_t N = 1;
  for (int d = 0; d < D; d++) {
    if (shape[d] == 0)
      return fail(MGX_FAILURE, "empty dimension");
    N *= shape[d];
  }

  const PtrLoc in_loc = locate(in);
  const int dev = in_loc.device ? in_loc.dev : cfg.dev_id;
  DeviceCache *cp = nullptr;
  rc = cache_get(dev, &cp);
  if (rc)
    return rc;
  DeviceCache &c = *cp;

  Slabs sl;
  rc = decide_slabs(D, shape, dtype, sT, available_bytes(c, cfg), &sl, &cfg);
  if (rc)
    return rc;

  // output buffer in the same memory space as the input (:141-154)
  size_t cap;
  bool allocated_here = false;
  if (!prealloc) {
    cap = N * es + 1000000;
    if (in_loc.device) {
      X_CUDA(cudaMalloc(out, cap));
    } else {
      *out = malloc(cap);
      if (!*out)
        return fail(MGX_FAILURE, "out of host memory");
    }
    allocated_here = true;
  } else {
    cap = *out_size;
    if (!*out)
      return fail(MGX_FAILURE, "output_pre_allocated without a buffer");
  }
  struct FreeOnError {
    void **out;
    bool dev, armed;
    ~FreeOnError() {
      if (armed) {
        if (dev)
          cudaFree(*out);
        else
          free(*out);
        *out = nullptr;
      }
    }
  } guard{out, in_loc.device, allocated_here};

  const PtrLoc out_loc = locate(*out);
  ScopedPin pin_in, pin_out;
  if (!in_loc.device)
    pin_in.pin(in, N * es, cfg.auto_pin_host_buffers);
  if (!out_loc.device)
    pin_out.pin(*out, cap, cfg.auto_pin_host_buffers);

  mgx_metadata md;
  memset(&md, 0, sizeof(md));
  md.version[0] = FILE_VERSION_MAJOR; // see mgard_b200_x.h: the reference stores
                                      // its file version in mgard_version
  md.D = D;
  for (int d = 0; d < D; d++)
    md.shape[d] = shape[d];
  md.dtype = dtype;
  md.mode = mode == MGX_EB_REL ? MGX_EB_REL : MGX_EB_ABS;
  md.linf = linf;
  md.s = sT;
  md.tol = tolT;
  md.norm = 1;
  md.decomposed = sl.decomposed;
  md.dd_type = sl.type;
  md.dd_dim = sl.dim;
  md.dd_size = sl.decomposed ? sl.size : shape[0];
  md.decomposition = MGX_DC_MULTIDIM;
  md.lossless = cfg.lossless;
  md.huff_dict_size = cfg.huff_dict_size;
  md.huff_block_size = cfg.huff_block_size;
  md.reorder = cfg.reorder;
  md.backend = 3; // X_CUDA
  // non-uniform grid: the node coordinates travel in the header as doubles
  // (Metadata.hpp:138-153)
  std::vector<double> hcoords;
  if (coords) {
    md.nonuniform = 1;
    for (int d = 0; d < D; d++) {
      if (!coords[d])
        return fail(MGX_FAILURE, "coordinate array missing");
      for (uint64_t i = 0; i < shape[d]; i++)
        hcoords.push_back(dtype == MGX_F32 ? (double)((const float *)coords[d])[i]
                                           : ((const double *)coords[d])[i]);
    }
  }
  std::vector<uint8_t> head = header_bytes(md, coords ? hcoords.data() : nullptr);
  const uint64_t M = PREAMBLE + head.size();
  if (cap < M)
    return fail(MGX_OUTPUT_TOO_LARGE, "output buffer smaller than the header");

  // input of subdomain i as a dense device array
  const bool in_place = in_loc.device && sl.contiguous();
  uint64_t max_elems = 0;
  for (uint64_t i = 0; i < sl.num; i++)
    max_elems = std::max(max_elems, sl.elems(i));
  const size_t pay_cap = max_elems * es + (max_elems / cfg.huff_block_size + 2) * 16 +
                         (size_t)(128 + cfg.huff_dict_size) * 8 + 4096;
  if (!in_place) {
    const int nb = sl.num > 1 ? 2 : 1;
    for (int b = 0; b < nb; b++)
      if ((rc = ensure_buffer(c.sub[b], c.sub_bytes[b], max_elems * es)))
        return rc;
  }
  for (int b = 0; b < (sl.num > 1 ? 2 : 1); b++)
    if ((rc = ensure_buffer(c.cbuf[b], c.cbuf_bytes[b], pay_cap)))
      return rc;

  auto slab_ptr = [&](uint64_t i, int b) -> const void * {
    if (!in_place)
      return c.sub[b];
    uint64_t sh[5], off;
    sl.sub_shape(i, sh, &off);
    return (const char *)in + (size_t)off * sl.inner * es;
  };
  auto fetch = [&](uint64_t i, int b) -> int { // OriginalToSubdomain
    if (in_place)
      return MGX_SUCCESS;
    // the buffer is free once the kernels that read it are done
    X_CUDA(cudaStreamWaitEvent(c.copy_in, c.ev_free[b], 0));
    int r = copy_slab(sl, i, c.sub[b], (void *)in, true, c.copy_in);
    if (r)
      return r;
    X_CUDA(cudaEventRecord(c.ev_in[b], c.copy_in));
    return MGX_SUCCESS;
  };

  // global norm + per-subdomain absolute tolerance (:100-131,
  // ErrorToleranceCalculator.hpp:61-147)
  double norm = 1, local_tol = tolT;
  int local_rel = rel;
  if (sl.decomposed) {
    if (rel) {
      double acc = 0;
      if ((rc = fetch(0, 0)))
        return rc;
      for (uint64_t i = 0; i < sl.num; i++) {
        const int b = (int)(i & 1);
        if (i + 1 < sl.num && (rc = fetch(i + 1, b ^ 1)))
          return rc;
        uint64_t sh[5], off;
        sl.sub_shape(i, sh, &off);
    uint64_t offs[5];
    sl.sub_box(i, sh, offs);
        mgx_ctx *ctx;
        SubCoords sc;
        if ((rc = cache_ctx(c, D, sh, dtype, cfg, &ctx,
                            sc.get(coords, D, offs, es))))
          return rc;
        if (!in_place)
          X_CUDA(cudaStreamWaitEvent(c.compute, c.ev_in[b], 0));
        double part = 0;
        rc = mgx_norm_partial(ctx, slab_ptr(i, b), linf, &part);
        if (rc)
          return fail(rc, std::string("norm: ") + mgx_last_error(ctx));
        X_CUDA(cudaEventRecord(c.ev_free[b], c.compute));
        acc = linf ? std::max(acc, part) : acc + part;
      }
      norm = mgx_finish_norm(dtype, linf, acc, N, cfg.normalize_coordinates ? 1 : 0);
    }
    local_tol = mgx_local_abs_tol(dtype, rel, norm, tolT, sT, sl.num);
    local_rel = 0; // ABS inside every subdomain (:126-129)
  }

  uint8_t *obase = (uint8_t *)*out;
  uint64_t offset = M; // bytes written so far
  uint8_t size_le[8];
  if ((rc = fetch(0, 0)))
    return rc;
  for (uint64_t i = 0; i < sl.num; i++) {
    const int b = (int)(i & 1);
    if (i + 1 < sl.num && (rc = fetch(i + 1, b ^ 1)))
      return rc;
    uint64_t sh[5], off;
    sl.sub_shape(i, sh, &off);
    uint64_t offs[5];
    sl.sub_box(i, sh, offs);
    const uint64_t n_sub = sl.elems(i);
    mgx_ctx *ctx;
    SubCoords sc;
    if ((rc = cache_ctx(c, D, sh, dtype, cfg, &ctx,
                        sc.get(coords, D, offs, es))))
      return rc;
    if (!in_place)
      X_CUDA(cudaStreamWaitEvent(c.compute, c.ev_in[b], 0));
    // the payload buffer is free once its previous D2H has finished
    X_CUDA(cudaStreamWaitEvent(c.compute, c.ev_out[b], 0));
    const void *d_in = slab_ptr(i, b);
    uint64_t bytes = 0;
    double nrm = local_rel ? 0.0 : norm;
    rc = mgx_compress(ctx, d_in, local_rel, local_tol, sT, &nrm,
                      (uint8_t *)c.cbuf[b], pay_cap, &bytes);
    if (rc != MGX_SUCCESS && rc != MGX_OUTPUT_TOO_LARGE)
      return fail(rc, std::string("compress: ") + mgx_last_error(ctx));
    if (!sl.decomposed && rel)
      norm = nrm;
    const void *payload = c.cbuf[b];
    const uint64_t huff_bytes = bytes; // the Huffman container in c.cbuf[b]
    bool raw = rc == MGX_OUTPUT_TOO_LARGE || bytes >= n_sub * es;
    if (!raw && cfg.lossless == MGX_LL_HUFFMAN_ZSTD) {
      // Lossless.hpp:188-192, Zstd.hpp:72-104: the record is [size_t n][ZSTD frame of
      // the n bytes of the serialised Huffman container]; libzstd runs on the host
      const ZstdApi &z = zstd_api();
      X_CUDA(cudaEventSynchronize(c.ev_out[b])); // the previous record left zout[b]
      c.zin.resize(bytes);
      X_CUDA(cudaMemcpy(c.zin.data(), c.cbuf[b], bytes, cudaMemcpyDeviceToHost));
      const size_t bound = z.compressBound(bytes);
      c.zout[b].resize(8 + bound);
      const size_t zs = z.compress(c.zout[b].data() + 8, bound, c.zin.data(), bytes,
                                   cfg.zstd_compress_level);
      if (z.isError(zs))
        return fail(MGX_FAILURE, std::string("ZSTD_compress: ") + z.getErrorName(zs));
      put_u64_le(c.zout[b].data(), bytes);
      payload = c.zout[b].data();
      bytes = 8 + zs;
      raw = bytes >= n_sub * es;
    } else if (!raw && cfg.lossless == MGX_LL_HUFFMAN_LZ4) {
      // Lossless.hpp:182-186: the serialised Huffman container goes through the
      // portable LZ4 stage (lz4_wrap above), on the host like the Zstd one
      X_CUDA(cudaEventSynchronize(c.ev_out[b]));
      c.zin.resize(bytes);
      X_CUDA(cudaMemcpy(c.zin.data(), c.cbuf[b], bytes, cudaMemcpyDeviceToHost));
      if ((rc = lz4_wrap(c.zin.data(), bytes, cfg.lz4_block_size, c.zout[b])))
        return rc;
      payload = c.zout[b].data();
      bytes = c.zout[b].size();
      raw = bytes >= n_sub * es;
    }
    // CR < 1: store the subdomain itself (GPUPipelines.hpp:139-156)
    // (>=: the decoder classifies a record of exactly n_sub * es bytes as raw)
    if (raw) {
      payload = d_in;
      bytes = n_sub * es;
    }
    if (offset + 8 > cap || bytes > cap - offset - 8)
      return fail(MGX_OUTPUT_TOO_LARGE, "output too large");
    // dry run (Compressor.hpp:239-242: Dequantize + Recompose after Serialize;
    // GPUPipelines.hpp:198-202: SubdomainToOriginal): the reconstruction replaces the
    // caller's INPUT array.  It is what mgx_x_decompress returns for this record -- a
    // subdomain stored raw (CR < 1) is its own reconstruction and stays as it is.
    if (cfg.compress_with_dryrun && !raw) {
      void *recon = in_place ? (void *)d_in : c.sub[b];
      int r2 = mgx_decompress(ctx, (const uint8_t *)c.cbuf[b], huff_bytes, local_rel,
                              local_tol, sT, nrm, recon);
      if (r2)
        return fail(r2, std::string("dry run: ") + mgx_last_error(ctx));
      if (!in_place && (r2 = copy_slab(sl, i, c.sub[b], (void *)in, false, c.compute)))
        return r2;
    }
    put_u64_le(size_le, bytes);
    if (out_loc.device)
      X_CUDA(cudaMemcpy(obase + offset, size_le, 8, cudaMemcpyHostToDevice));
    else
      memcpy(obase + offset, size_le, 8);
    offset += 8;
    // mgx_compress has synchronised the compute stream: the payload is final
    X_CUDA(cudaMemcpyAsync(obase + offset, payload, bytes, cudaMemcpyDefault,
                           c.copy_out));
    X_CUDA(cudaEventRecord(c.ev_out[b], c.copy_out));
    if (payload == d_in && !in_place) { // raw copy still reads the slab buffer
      X_CUDA(cudaStreamWaitEvent(c.compute, c.ev_out[b], 0));
    }
    X_CUDA(cudaEventRecord(c.ev_free[b], c.compute));
    offset += bytes;
  }
  X_CUDA(cudaStreamSynchronize(c.copy_out));
  X_CUDA(cudaStreamSynchronize(c.compute));

  // header with the final norm (:230-262)
  md.norm = dtype == MGX_F32 ? (double)(float)norm : norm;
  head = header_bytes(md, coords ? hcoords.data() : nullptr);
  std::vector<uint8_t> pre(PREAMBLE + head.size());
  memcpy(pre.data(), SIGNATURE, 5);
  put_u64_le(pre.data() + 5, head.size());
  const uint32_t crc = mgx_crc32(head.data(), head.size());
  for (int i = 0; i < 4; i++)
    pre[13 + i] = (uint8_t)(crc >> (8 * i));
  memcpy(pre.data() + PREAMBLE, head.data(), head.size());
  if (pre.size() != M)
    return fail(MGX_FAILURE, "header size changed");
  if (out_loc.device)
    X_CUDA(cudaMemcpy(obase, pre.data(), M, cudaMemcpyHostToDevice));
  else
    memcpy(obase, pre.data(), M);
  *out_size = offset;
  guard.armed = false;
  if (cfg.auto_cache_release)
    cache_release(c);
  return MGX_SUCCESS;
}

// ---- compress on several GPUs of one process ------------------------------------
// One host thread per device, MaxDim subdomains dealt to the devices in contiguous
// blocks.  The only coupling between the devices is the scalar of a REL bound: every
// thread leaves the partials of its subdomains in a shared array, a barrier, and
// every thread adds them up in subdomain order (the same fp64 result on all of them,
// and the same one x_compress gets).  The ABS bound per subdomain is a host-side
// argument of mgx_compress, so the scalar passes through the host anyway: an 8-byte
// read per subdomain and a host barrier replace the all-reduce of the N-process
// path.  Payloads stay on their device until every size is known, then rank-ordered
// records are copied behind the header: the stream is byte-identical to the one
// x_compress writes for the same decomposition.
struct HostBarrier {
  std::mutex m;
  std::condition_variable cv;
  int n, waiting = 0, gen = 0;
  explicit HostBarrier(int n_) : n(n_) {}
  void wait() {
    std::unique_lock<std::mutex> l(m);
    const int g = gen;
    if (++waiting == n) {
      waiting = 0;
      ++gen;
      cv.notify_all();
    } else {
      cv.wait(l, [&] { return gen != g; });
    }
  }
};

struct MultiShared {
  Slabs sl;
  int D, dtype, rel, linf, ndev;
  double tolT, sT;
  uint64_t N;
  const void *in;
  mgx_config cfg;
  std::vector<double> part;      // per subdomain: max |u| or sum u^2
  std::vector<void *> d_payload; // per subdomain: record payload on its device
  std::vector<uint64_t> bytes;
  double norm = 1, local_tol = 0;
  int local_rel = 0;
  std::vector<int> rc;
  std::vector<std::string> msg;
};

void multi_worker(MultiShared &S, HostBarrier &bar, int k, int dev, DeviceCache *cp) {
  DeviceCache &c = *cp;
  const size_t es = S.sl.es;
  const uint64_t per = (S.sl.num + S.ndev - 1) / S.ndev;
  const uint64_t lo = std::min<uint64_t>((uint64_t)k * per, S.sl.num),
                 hi = std::min<uint64_t>(lo + per, S.sl.num);
  auto bail = [&](int rc, const std::string &m) {
    if (!S.rc[k]) {
      S.rc[k] = rc;
      S.msg[k] = m;
    }
  };
  bool ok = cudaSetDevice(dev) == cudaSuccess;
  if (!ok)
    bail(MGX_FAILURE, "cudaSetDevice failed");
  uint64_t max_elems = 0;
  for (uint64_t i = lo; i < hi; i++)
    max_elems = std::max(max_elems, S.sl.elems(i));
  const size_t pay_cap = max_elems * es + (max_elems / S.cfg.huff_block_size + 2) * 16 +
                         (size_t)(128 + S.cfg.huff_dict_size) * 8 + 4096;
  if (ok && hi > lo) {
    if (ensure_buffer(c.sub[0], c.sub_bytes[0], max_elems * es) ||
        ensure_buffer(c.cbuf[0], c.cbuf_bytes[0], pay_cap)) {
      bail(MGX_FAILURE, "out of device memory");
      ok = false;
    }
  }
  auto load = [&](uint64_t i) -> bool { // OriginalToSubdomain, synchronous
    if (copy_slab(S.sl, i, c.sub[0], (void *)S.in, true, c.copy_in) ||
        cudaStreamSynchronize(c.copy_in) != cudaSuccess) {
      bail(MGX_FAILURE, "host -> device copy failed");
      return false;
    }
    return true;
  };
  auto ctx_of = [&](uint64_t i, mgx_ctx **ctx) -> bool {
    uint64_t sh[5], off;
    S.sl.sub_shape(i, sh, &off);
    const int r = cache_ctx(c, S.D, sh, S.dtype, S.cfg, ctx);
    if (r)
      bail(r, "plan failed");
    return r == 0;
  };
  // ---- global norm (calc_norm_decomposed, ErrorToleranceCalculator.hpp:61-123)
  if (S.rel && S.sl.decomposed) {
    for (uint64_t i = lo; ok && i < hi; i++) {
      mgx_ctx *ctx;
      if (!(ok = load(i) && ctx_of(i, &ctx)))
        break;
      const int r = mgx_norm_partial(ctx, c.sub[0], S.linf, &S.part[i]);
      if (r) {
        bail(r, std::string("norm: ") + mgx_last_error(ctx));
        ok = false;
      }
    }
    bar.wait();
    bool all_ok = true;
    for (int j = 0; j < S.ndev; j++)
      all_ok = all_ok && S.rc[j] == 0;
    if (!all_ok)
      ok = false;
    if (k == 0 && all_ok) {
      double acc = 0;
      for (uint64_t i = 0; i < S.sl.num; i++)
        acc = S.linf ? std::max(acc, S.part[i]) : acc + S.part[i];
      S.norm = mgx_finish_norm(S.dtype, S.linf, acc, S.N,
                               S.cfg.normalize_coordinates ? 1 : 0);
      S.local_tol = mgx_local_abs_tol(S.dtype, S.rel, S.norm, S.tolT, S.sT, S.sl.num);
    }
    bar.wait();
  }
  // ---- compress this device's subdomains, payloads stay on the device
  for (uint64_t i = lo; ok && i < hi; i++) {
    mgx_ctx *ctx;
    if (!(ok = load(i) && ctx_of(i, &ctx)))
      break;
    const uint64_t n_sub = S.sl.elems(i);
    uint64_t bytes = 0;
    double nrm = S.local_rel ? 0.0 : S.norm;
    int r = mgx_compress(ctx, c.sub[0], S.local_rel, S.local_tol, S.sT, &nrm,
                         (uint8_t *)c.cbuf[0], pay_cap, &bytes);
    if (r != MGX_SUCCESS && r != MGX_OUTPUT_TOO_LARGE) {
      bail(r, std::string("compress: ") + mgx_last_error(ctx));
      ok = false;
      break;
    }
    if (!S.sl.decomposed && S.rel)
      S.norm = nrm;
    const void *payload = c.cbuf[0];
    if (r == MGX_OUTPUT_TOO_LARGE || bytes >= n_sub * es) { // raw fallback
      payload = c.sub[0];
      bytes = n_sub * es;
    }
    void *keep = nullptr;
    if (cudaMalloc(&keep, bytes) != cudaSuccess ||
        cudaMemcpy(keep, payload, bytes, cudaMemcpyDeviceToDevice) != cudaSuccess) {
      cudaGetLastError();
      bail(MGX_FAILURE, "out of device memory for a payload");
      ok = false;
      break;
    }
    S.d_payload[i] = keep;
    S.bytes[i] = bytes;
  }
}

int x_compress_multi(int D, int dtype, const uint64_t *shape, double tol, double s,
                     int mode, const void *in, void **out, size_t *out_size,
                     const int *devices, int ndev, const mgx_config &cfg, int prealloc) {
  if (D > 5)
    return fail(MGX_TOO_MANY_DIMENSIONS, "more than 5 dimensions");
  if (dtype != MGX_F32 && dtype != MGX_F64)
    return fail(MGX_BAD_DATA_TYPE, "data type not supported");
  if (D < 1 || !shape || !in || !out || !out_size || !devices || ndev < 1 || ndev > 64)
    return fail(MGX_FAILURE, "bad argument");
  int rc = check_config(cfg);
  if (rc)
    return rc;
  // (the one-GPU pipeline does it, x_compress below "dry run")
  if (cfg.compress_with_dryrun)
    return fail(MGX_FAILURE, "compress_with_dryrun is not supported by the multi-GPU entry");
  if (cfg.lossless != MGX_LL_HUFFMAN)
    return fail(MGX_FAILURE, "the multi-GPU entry writes Huffman streams only");
  if (locate(in).device)
    return fail(MGX_FAILURE, "mgx_x_compress_multi takes host buffers");
  if (cfg.domain_decomposition != MGX_DD_MAXDIM)
    return fail(MGX_FAILURE, "mgx_x_compress_multi: MaxDim decomposition only");
  for (int a = 0; a < ndev; a++)
    for (int b = a + 1; b < ndev; b++)
      if (devices[a] == devices[b])
        return fail(MGX_FAILURE, "a device is listed twice");
  MultiShared S;
  S.D = D;
  S.dtype = dtype;
  S.ndev = ndev;
  S.cfg = cfg;
  S.in = in;
  const size_t es = dtype == MGX_F64 ? 8 : 4;
  S.tolT = dtype == MGX_F32 ? (double)(float)tol : tol;
  S.sT = dtype == MGX_F32 ? (double)(float)s : s;
  S.rel = mode == MGX_EB_REL;
  S.linf = std::isinf(S.sT) ? 1 : 0;
  S.N = 1;
  for (int d = 0; d < D; d++) {
    if (shape[d] == 0)
      return fail(MGX_FAILURE, "empty dimension");
    S.N *= shape[d];
  }
  // devices first (streams, caches), the smallest free memory bounds a subdomain
  std::vector<DeviceCache *> caches(ndev);
  uint64_t avail = ~0ull;
  for (int k = 0; k < ndev; k++) {
    if ((rc = cache_get(devices[k], &caches[k])))
      return rc;
    avail = std::min(avail, available_bytes(*caches[k], cfg));
  }
  // at least one subdomain per device (DomainDecomposer.hpp:192-223 halving rule),
  // further halved until one fits the memory
  Slabs &sl = S.sl;
  if ((rc = decide_slabs(D, shape, dtype, S.sT, avail, &sl)))
    return rc;
  if (sl.num < (uint64_t)ndev) {
    int dim = 0;
    uint64_t size = 0, num = 1;
    mgx_max_dim_decomposition(D, shape, (uint64_t)ndev, &dim, &size, &num);
    if (num > 1) {
      sl.decomposed = true;
      sl.dim = dim;
      sl.size = size;
      sl.num = num;
      sl.finish();
    }
  }
  S.part.assign(sl.num, 0.0);
  S.d_payload.assign(sl.num, nullptr);
  S.bytes.assign(sl.num, 0);
  S.rc.assign(ndev, 0);
  S.msg.assign(ndev, "");
  S.local_rel = S.rel;
  S.local_tol = S.tolT;
  if (sl.decomposed) {
    S.local_rel = 0; // ABS inside every subdomain (CompressionHighLevel.hpp:126-129)
    if (!S.rel)
      S.local_tol = mgx_local_abs_tol(dtype, 0, 1.0, S.tolT, S.sT, sl.num);
  }
  ScopedPin pin_in;
  pin_in.pin(in, S.N * es, cfg.auto_pin_host_buffers);
  {
    HostBarrier bar(ndev);
    std::vector<std::thread> th;
    for (int k = 0; k < ndev; k++)
      th.emplace_back(multi_worker, std::ref(S), std::ref(bar), k, devices[k], caches[k]);
    for (auto &t : th)
      t.join();
  }
  struct FreePayloads {
    MultiShared &S;
    const int *devices;
    ~FreePayloads() {
      const uint64_t per = (S.sl.num + S.ndev - 1) / S.ndev;
      for (uint64_t i = 0; i < S.sl.num; i++)
        if (S.d_payload[i]) {
          cudaSetDevice(devices[i / per]);
          cudaFree(S.d_payload[i]);
        }
    }
  } freer{S, devices};
  for (int k = 0; k < ndev; k++)
    if (S.rc[k])
      return fail(S.rc[k], "device " + std::to_string(devices[k]) + ": " + S.msg[k]);

  mgx_metadata md;
  memset(&md, 0, sizeof(md));
  md.version[0] = FILE_VERSION_MAJOR;
  md.D = D;
  for (int d = 0; d < D; d++)
    md.shape[d] = shape[d];
  md.dtype = dtype;
  md.mode = S.rel ? MGX_EB_REL : MGX_EB_ABS;
  md.linf = S.linf;
  md.s = S.sT;
  md.tol = S.tolT;
  md.norm = dtype == MGX_F32 ? (double)(float)S.norm : S.norm;
  md.decomposed = sl.decomposed;
  md.dd_type = MGX_DD_MAXDIM;
  md.dd_dim = sl.dim;
  md.dd_size = sl.decomposed ? sl.size : shape[0];
  md.decomposition = MGX_DC_MULTIDIM;
  md.lossless = MGX_LL_HUFFMAN;
  md.huff_dict_size = cfg.huff_dict_size;
  md.huff_block_size = cfg.huff_block_size;
  md.reorder = cfg.reorder;
  md.backend = 3; // X_CUDA
  std::vector<uint8_t> head = header_bytes(md);
  uint64_t total = PREAMBLE + head.size();
  for (uint64_t i = 0; i < sl.num; i++)
    total += 8 + S.bytes[i];
  if (!prealloc) {
    *out = malloc(total);
    if (!*out)
      return fail(MGX_FAILURE, "out of host memory");
  } else if (!*out || *out_size < total) {
    return fail(MGX_OUTPUT_TOO_LARGE, "output too large");
  }
  uint8_t *o = (uint8_t *)*out;
  memcpy(o, SIGNATURE, 5);
  put_u64_le(o + 5, head.size());
  const uint32_t crc = mgx_crc32(head.data(), head.size());
  for (int i = 0; i < 4; i++)
    o[13 + i] = (uint8_t)(crc >> (8 * i));
  memcpy(o + PREAMBLE, head.data(), head.size());
  uint64_t offset = PREAMBLE + head.size();
  const uint64_t per = (sl.num + ndev - 1) / ndev;
  for (uint64_t i = 0; i < sl.num; i++) {
    put_u64_le(o + offset, S.bytes[i]);
    offset += 8;
    X_CUDA(cudaSetDevice(devices[i / per]));
    X_CUDA(cudaMemcpy(o + offset, S.d_payload[i], S.bytes[i], cudaMemcpyDeviceToHost));
    offset += S.bytes[i];
  }
  *out_size = offset;
  return MGX_SUCCESS;
}

// ---- decompress ------------------------------------------------------------------
int fetch_header(const void *in, size_t in_size, const PtrLoc &loc,
                 std::vector<uint8_t> *buf) {
  if (in_size < PREAMBLE)
    return fail(MGX_FAILURE, "stream shorter than the MGARD preamble");
  uint8_t pre[PREAMBLE];
  if (loc.device)
    X_CUDA(cudaMemcpy(pre, in, PREAMBLE, cudaMemcpyDeviceToHost));
  else
    memcpy(pre, in, PREAMBLE);
  uint64_t total = 0;
  int rc = mgx_metadata_size(pre, PREAMBLE, &total);
  if (rc)
    return rc;
  if (total > in_size)
    return fail(MGX_FAILURE, "stream shorter than its header");
  buf->resize(total);
  if (loc.device)
    X_CUDA(cudaMemcpy(buf->data(), in, total, cudaMemcpyDeviceToHost));
  else
    memcpy(buf->data(), in, total);
  return MGX_SUCCESS;
}

// rec_lo / rec_hi: only the subdomain records [rec_lo, rec_hi) are decoded (the
// several-GPU entry hands every device its block of records)
int x_decompress(const void *in, size_t in_size, void **out, uint64_t *shape_out,
                 int *D_out, int *dtype_out, const mgx_config &cfg_in, int prealloc,
                 uint64_t rec_lo = 0, uint64_t rec_hi = ~0ull) {
  if (!in || !out)
    return fail(MGX_FAILURE, "bad argument");
  if (cfg_in.dev_type != MGX_DEV_AUTO && cfg_in.dev_type != MGX_DEV_CUDA)
    return fail(MGX_BACKEND_UNAVAILABLE,
                "only the CUDA backend is built (no CPU fallback by design)");
  int count = 0;
  if (cudaGetDeviceCount(&count) != cudaSuccess || count <= 0) {
    cudaGetLastError();
    return fail(MGX_BACKEND_UNAVAILABLE, "no CUDA device");
  }
  const PtrLoc in_loc = locate(in);
  std::vector<uint8_t> hb;
  int rc = fetch_header(in, in_size, in_loc, &hb);
  if (rc)
    return rc;
  mgx_metadata md;
  uint64_t M = 0;
  rc = mgx_metadata_parse(hb.data(), hb.size(), &md, &M);
  if (rc)
    return rc;
  if (md.D > 5)
    return fail(MGX_TOO_MANY_DIMENSIONS, "more than 5 dimensions");
  if (md.D < 1)
    return fail(MGX_FAILURE, "header without a shape");
  // non-uniform grid: node coordinates from the header.  The reference narrows them
  // to float whatever the data type ("coords[d][i] = (float)m.coords[d][i]",
  // CompressionHighLevel.hpp:443-450); reproduced so that both decoders rebuild the
  // same hierarchy from the same stream.
  std::vector<std::vector<unsigned char>> tcoords;
  const void *cptr[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};
  if (md.nonuniform) {
    std::vector<double> hc;
    mgx_metadata md2;
    rc = parse_header(hb.data() + PREAMBLE, M - PREAMBLE, md2, &hc);
    if (rc)
      return rc;
    uint64_t total = 0;
    for (uint32_t d = 0; d < md.D; d++)
      total += md.shape[d];
    if (hc.size() != total)
      return fail(MGX_FAILURE,
                  "mismatch between number of node coordinates and grid shape.");
    tcoords.resize(md.D);
    uint64_t at = 0;
    for (uint32_t d = 0; d < md.D; d++) {
      const uint64_t n = md.shape[d];
      tcoords[d].resize(n * (md.dtype == MGX_F64 ? 8 : 4));
      for (uint64_t i = 0; i < n; i++) {
        const float f = (float)hc[at + i];
        if (md.dtype == MGX_F64)
          ((double *)tcoords[d].data())[i] = (double)f;
        else
          ((float *)tcoords[d].data())[i] = f;
      }
      at += n;
      cptr[d] = tcoords[d].data();
    }
  }
  const void *const *coords = md.nonuniform ? cptr : nullptr;
  // MetadataBase::InitializeConfig (Metadata.cpp:129-149)
  mgx_config cfg = cfg_in;
  cfg.domain_decomposition = md.dd_type;
  cfg.decomposition = md.decomposition;
  cfg.lossless = md.lossless;
  if (md.huff_dict_size)
    cfg.huff_dict_size = md.huff_dict_size;
  if (md.huff_block_size)
    cfg.huff_block_size = md.huff_block_size;
  cfg.reorder = md.reorder;
  if ((rc = check_config(cfg)))
    return rc;

  const int D = (int)md.D, dtype = md.dtype;
  const size_t es = dtype == MGX_F64 ? 8 : 4;
  uint64_t N = 1;
  for (int d = 0; d < D; d++) {
    // (only the header's own bytes are covered by its CRC: a zero extent or a
    // product that does not fit 64 bits must be refused here)
    if (md.shape[d] == 0 || md.shape[d] > (~0ull / es) / N)
      return fail(MGX_FAILURE, "bad shape in the header");
    N *= md.shape[d];
  }
  if (shape_out)
    for (int d = 0; d < D; d++)
      shape_out[d] = md.shape[d];
  if (D_out)
    *D_out = D;
  if (dtype_out)
    *dtype_out = dtype;

  const int dev = in_loc.device ? in_loc.dev : cfg.dev_id;
  DeviceCache *cp = nullptr;
  if ((rc = cache_get(dev, &cp)))
    return rc;
  DeviceCache &c = *cp;

  bool allocated_here = false;
  if (!prealloc) {
    if (in_loc.device) {
      X_CUDA(cudaMalloc(out, N * es));
    } else {
      *out = malloc(N * es);
      if (!*out)
        return fail(MGX_FAILURE, "out of host memory");
    }
    allocated_here = true;
  } else if (!*out) {
    return fail(MGX_FAILURE, "output_pre_allocated without a buffer");
  }
  struct FreeOnError {
    void **out;
    bool dev, armed;
    ~FreeOnError() {
      if (armed) {
        if (dev)
          cudaFree(*out);
        else
          free(*out);
        *out = nullptr;
      }
    }
  } guard{out, in_loc.device, allocated_here};
  const PtrLoc out_loc = locate(*out);
  ScopedPin pin_in, pin_out;
  if (!in_loc.device)
    pin_in.pin(in, in_size, cfg.auto_pin_host_buffers);
  if (!out_loc.device)
    pin_out.pin(*out, N * es, cfg.auto_pin_host_buffers);

  Slabs sl;
  sl.D = D;
  sl.es = es;
  for (int d = 0; d < D; d++)
    sl.shape[d] = md.shape[d];
  sl.decomposed = md.decomposed != 0;
  if (sl.decomposed) {
    if ((md.dd_type != MGX_DD_MAXDIM && md.dd_type != MGX_DD_BLOCK) ||
        md.dd_dim >= (uint32_t)D || md.dd_size == 0)
      return fail(MGX_FAILURE, "bad domain decomposition in the header");
    sl.type = md.dd_type;
    sl.dim = (int)md.dd_dim;
    sl.size = md.dd_size;
    sl.num = (md.shape[sl.dim] - 1) / sl.size + 1;
  }
  sl.finish();
  if (sl.decomposed && sl.num > in_size / 8) // a record starts with an 8-byte size
    return fail(MGX_FAILURE, "more subdomains in the header than the stream can hold");
  if (sl.block())
    for (uint64_t i = 0; i < sl.num; i++) {
      uint64_t b[5], o[5];
      sl.sub_box(i, b, o);
      for (int d = 0; d < D; d++)
        if (b[d] < 3)
          return fail(MGX_FAILURE, "bad block decomposition in the header");
    }

  const int rel = md.mode == MGX_EB_REL;
  const double sT = dtype == MGX_F32 ? (double)(float)md.s : md.s;
  const double normT = dtype == MGX_F32 ? (double)(float)md.norm : md.norm;
  double local_tol = dtype == MGX_F32 ? (double)(float)md.tol : md.tol;
  int local_rel = rel;
  if (sl.decomposed) {
    local_tol = mgx_local_abs_tol(dtype, rel, md.norm, md.tol, md.s, sl.num);
    local_rel = 0;
  }

  const bool in_place = out_loc.device && sl.contiguous();
  uint64_t max_elems = 0;
  for (uint64_t i = 0; i < sl.num; i++)
    max_elems = std::max(max_elems, sl.elems(i));
  if (!in_place)
    for (int b = 0; b < (sl.num > 1 ? 2 : 1); b++)
      if ((rc = ensure_buffer(c.sub[b], c.sub_bytes[b], max_elems * es)))
        return rc;

  // walk the records once on the host side to know sizes and offsets
  // (a device-resident stream costs one 8-byte D2H per subdomain, like the
  // reference's Deserialize<SIZE>, GPUPipelines.hpp:338-340)
  std::vector<uint64_t> rec_off(sl.num), rec_bytes(sl.num);
  uint64_t pos = M, max_rec = 0;
  for (uint64_t i = 0; i < sl.num; i++) {
    if (pos + 8 > in_size)
      return fail(MGX_FAILURE, "stream ends inside a subdomain record");
    uint8_t le[8];
    if (in_loc.device)
      X_CUDA(cudaMemcpy(le, (const uint8_t *)in + pos, 8, cudaMemcpyDeviceToHost));
    else
      memcpy(le, (const uint8_t *)in + pos, 8);
    uint64_t n = 0;
    for (int k = 0; k < 8; k++)
      n |= (uint64_t)le[k] << (8 * k);
    pos += 8;
    if (n > in_size - pos)
      return fail(MGX_FAILURE, "stream ends inside a subdomain record");
    rec_off[i] = pos;
    rec_bytes[i] = n;
    max_rec = std::max(max_rec, n);
    pos += n;
  }
  for (int b = 0; b < (sl.num > 1 ? 2 : 1); b++)
    if ((rc = ensure_buffer(c.cbuf[b], c.cbuf_bytes[b], max_rec + 64)))
      return rc;

  auto fetch = [&](uint64_t i, int b) -> int {
    X_CUDA(cudaStreamWaitEvent(c.copy_in, c.ev_free[b], 0));
    X_CUDA(cudaMemcpyAsync(c.cbuf[b], (const uint8_t *)in + rec_off[i], rec_bytes[i],
                           cudaMemcpyDefault, c.copy_in));
    X_CUDA(cudaEventRecord(c.ev_in[b], c.copy_in));
    return MGX_SUCCESS;
  };

  const uint64_t first = std::min<uint64_t>(rec_lo, sl.num),
                 last = std::min<uint64_t>(rec_hi, sl.num);
  if (first < last && (rc = fetch(first, 0)))
    return rc;
  for (uint64_t i = first; i < last; i++) {
    const int b = (int)((i - first) & 1);
    if (i + 1 < last && (rc = fetch(i + 1, b ^ 1)))
      return rc;
    uint64_t sh[5], off;
    sl.sub_shape(i, sh, &off);
    uint64_t offs[5];
    sl.sub_box(i, sh, offs);
    const uint64_t n_sub = sl.elems(i);
    void *d_out = in_place ? (void *)((char *)*out + (size_t)off * sl.inner * es)
                           : c.sub[b];
    X_CUDA(cudaStreamWaitEvent(c.compute, c.ev_in[b], 0));
    if (!in_place) // previous copy-out of this buffer
      X_CUDA(cudaStreamWaitEvent(c.compute, c.ev_out[b], 0));
    // CR > 1 -> compressed; otherwise the record is the raw subdomain
    // (GPUPipelines.hpp:407-421)
    const double CR = (double)n_sub * es / (double)rec_bytes[i];
    if (CR > 1.0) {
      mgx_ctx *ctx;
      SubCoords sc;
      if ((rc = cache_ctx(c, D, sh, dtype, cfg, &ctx,
                          sc.get(coords, D, offs, es))))
        return rc;
      uint64_t huff_bytes = rec_bytes[i];
      if (md.lossless == MGX_LL_HUFFMAN_ZSTD) {
        // Zstd.hpp:106-141: [size_t n][ZSTD frame] -> the n bytes of the Huffman
        // container, decoded by libzstd on the host and sent back to the device
        const ZstdApi &z = zstd_api();
        if (rec_bytes[i] < 8)
          return fail(MGX_FAILURE, "Zstd record shorter than its size field");
        X_CUDA(cudaEventSynchronize(c.ev_in[b]));
        c.zin.resize(rec_bytes[i]);
        X_CUDA(cudaMemcpy(c.zin.data(), c.cbuf[b], rec_bytes[i], cudaMemcpyDeviceToHost));
        uint64_t n = 0;
        for (int k = 0; k < 8; k++)
          n |= (uint64_t)c.zin[k] << (8 * k);
        // a Huffman container is smaller than the field plus its fixed sections
        if (n > n_sub * es + 16 * (n_sub / cfg.huff_block_size + 2) +
                    8ull * (128 + cfg.huff_dict_size) + 4096)
          return fail(MGX_FAILURE, "Zstd record announces an impossible size");
        c.zout[b].resize(n);
        const size_t got = z.decompress(c.zout[b].data(), n, c.zin.data() + 8,
                                        rec_bytes[i] - 8);
        if (z.isError(got))
          return fail(MGX_FAILURE, std::string("ZSTD_decompress: ") + z.getErrorName(got));
        if (got != n)
          return fail(MGX_FAILURE, "Zstd frame shorter than announced");
        if ((rc = ensure_buffer(c.cbuf[b], c.cbuf_bytes[b], n)))
          return rc;
        // on the compute stream: a cudaMemcpy from pageable memory returns once the
        // bytes are STAGED, and the kernels' stream does not wait for the legacy one
        X_CUDA(cudaMemcpyAsync(c.cbuf[b], c.zout[b].data(), n, cudaMemcpyHostToDevice,
                               c.compute));
        huff_bytes = n;
      } else if (md.lossless == MGX_LL_HUFFMAN_LZ4) {
        // LZ4.hpp:183-262: container -> the bytes of the Huffman container
        X_CUDA(cudaEventSynchronize(c.ev_in[b]));
        c.zin.resize(rec_bytes[i]);
        X_CUDA(cudaMemcpy(c.zin.data(), c.cbuf[b], rec_bytes[i], cudaMemcpyDeviceToHost));
        const uint64_t limit = n_sub * es + 16 * (n_sub / cfg.huff_block_size + 2) +
                               8ull * (128 + cfg.huff_dict_size) + 4096;
        if ((rc = lz4_unwrap(c.zin.data(), rec_bytes[i], limit, c.zout[b])))
          return rc;
        const uint64_t n = c.zout[b].size();
        if ((rc = ensure_buffer(c.cbuf[b], c.cbuf_bytes[b], n)))
          return rc;
        X_CUDA(cudaMemcpyAsync(c.cbuf[b], c.zout[b].data(), n, cudaMemcpyHostToDevice,
                               c.compute));
        huff_bytes = n;
      }
      rc = mgx_decompress(ctx, (const uint8_t *)c.cbuf[b], huff_bytes, local_rel,
                          local_tol, sT, normT, d_out);
      if (rc)
        return fail(rc, std::string("decompress: ") + mgx_last_error(ctx));
    } else {
      if (rec_bytes[i] != n_sub * es)
        return fail(MGX_FAILURE, "raw subdomain record of the wrong size");
      X_CUDA(cudaMemcpyAsync(d_out, c.cbuf[b], rec_bytes[i], cudaMemcpyDeviceToDevice,
                             c.compute));
      X_CUDA(cudaStreamSynchronize(c.compute));
    }
    X_CUDA(cudaEventRecord(c.ev_free[b], c.compute));
    if (!in_place) { // SubdomainToOriginal, after the kernels that fill the slab
      X_CUDA(cudaStreamWaitEvent(c.copy_out, c.ev_free[b], 0));
      if ((rc = copy_slab(sl, i, c.sub[b], *out, false, c.copy_out)))
        return rc;
      X_CUDA(cudaEventRecord(c.ev_out[b], c.copy_out));
    }
  }
  X_CUDA(cudaStreamSynchronize(c.copy_out));
  X_CUDA(cudaStreamSynchronize(c.compute));
  guard.armed = false;
  if (cfg.auto_cache_release)
    cache_release(c);
  return MGX_SUCCESS;
}

// decompress on several GPUs of one process: devices[k] decodes the k-th contiguous
// block of subdomain records on its own host thread, straight into the caller's host
// array (the counterpart of x_compress_multi; any stream with a domain decomposition)
int x_decompress_multi(const void *in, size_t in_size, void **out, uint64_t *shape_out,
                       int *D_out, int *dtype_out, const int *devices, int ndev,
                       const mgx_config &cfg, int prealloc) {
  if (!in || !out || !devices || ndev < 1 || ndev > 64)
    return fail(MGX_FAILURE, "bad argument");
  if (locate(in).device)
    return fail(MGX_FAILURE, "mgx_x_decompress_multi takes host buffers");
  for (int a = 0; a < ndev; a++)
    for (int b = a + 1; b < ndev; b++)
      if (devices[a] == devices[b])
        return fail(MGX_FAILURE, "a device is listed twice");
  mgx_metadata md;
  uint64_t M = 0;
  int rc = mgx_metadata_parse((const uint8_t *)in, in_size, &md, &M);
  if (rc)
    return rc;
  if (md.D < 1 || md.D > 5)
    return fail(MGX_FAILURE, "bad shape in the header");
  if (md.lossless != MGX_LL_HUFFMAN)
    return fail(MGX_FAILURE, "the multi-GPU entry reads Huffman streams only");
  const size_t es = md.dtype == MGX_F64 ? 8 : 4;
  uint64_t N = 1, num = 1;
  for (uint32_t d = 0; d < md.D; d++) {
    if (md.shape[d] == 0 || md.shape[d] > (~0ull / es) / N)
      return fail(MGX_FAILURE, "bad shape in the header");
    N *= md.shape[d];
  }
  if (md.decomposed) {
    if (md.dd_size == 0 || md.dd_dim >= md.D)
      return fail(MGX_FAILURE, "bad domain decomposition in the header");
    if (md.dd_type == MGX_DD_BLOCK) {
      for (uint32_t d = 0; d < md.D; d++)
        num *= (md.shape[d] - 1) / md.dd_size + 1;
    } else {
      num = (md.shape[md.dd_dim] - 1) / md.dd_size + 1;
    }
  }
  bool allocated_here = false;
  if (!prealloc) {
    *out = malloc(N * es);
    if (!*out)
      return fail(MGX_FAILURE, "out of host memory");
    allocated_here = true;
  } else if (!*out) {
    return fail(MGX_FAILURE, "output_pre_allocated without a buffer");
  }
  std::vector<DeviceCache *> caches(ndev);
  for (int k = 0; k < ndev; k++)
    if ((rc = cache_get(devices[k], &caches[k]))) {
      if (allocated_here) {
        free(*out);
        *out = nullptr;
      }
      return rc;
    }
  ScopedPin pin_in, pin_out; // once, here: the workers must not unpin under each other
  pin_in.pin(in, in_size, cfg.auto_pin_host_buffers);
  pin_out.pin(*out, N * es, cfg.auto_pin_host_buffers);
  const uint64_t per = (num + ndev - 1) / ndev;
  std::vector<int> rcs(ndev, 0);
  {
    std::vector<std::thread> th;
    for (int k = 0; k < ndev; k++) {
      const uint64_t lo = std::min<uint64_t>((uint64_t)k * per, num),
                     hi = std::min<uint64_t>(lo + per, num);
      if (lo >= hi)
        continue;
      th.emplace_back([&, k, lo, hi]() {
        mgx_config c = cfg;
        c.dev_id = devices[k];
        c.auto_pin_host_buffers = 0;
        c.auto_cache_release = 0;
        void *o = *out;
        rcs[k] = x_decompress(in, in_size, &o, nullptr, nullptr, nullptr, c, 1, lo, hi);
      });
    }
    for (auto &t : th)
      t.join();
  }
  for (int k = 0; k < ndev; k++)
    if (rcs[k]) {
      if (allocated_here) {
        free(*out);
        *out = nullptr;
      }
      return rcs[k]; // the message of the (last) failing worker is in g_err
    }
  if (shape_out)
    for (uint32_t d = 0; d < md.D; d++)
      shape_out[d] = md.shape[d];
  if (D_out)
    *D_out = (int)md.D;
  if (dtype_out)
    *dtype_out = md.dtype;
  return MGX_SUCCESS;
}

} // namespace

// the caller's current device is the caller's business: put it back on return
struct DeviceRestore {
  int dev = -1;
  DeviceRestore() {
    if (cudaGetDevice(&dev) != cudaSuccess)
      dev = -1;
  }
  ~DeviceRestore() {
    if (dev >= 0)
      cudaSetDevice(dev);
  }
};

extern "C" {

void mgx_config_default(mgx_config *cfg) {
  if (!cfg)
    return;
  cfg->dev_type = MGX_DEV_AUTO;
  cfg->dev_id = 0;
  cfg->domain_decomposition = MGX_DD_MAXDIM;
  cfg->decomposition = MGX_DC_MULTIDIM;
  cfg->lossless = MGX_LL_HUFFMAN;
  cfg->zstd_compress_level = 3; // Config.cpp:44
  cfg->lz4_block_size = 1u << 15;
  cfg->huff_dict_size = 8192;
  cfg->huff_block_size = 1024;
  cfg->reorder = 0;
  cfg->normalize_coordinates = 1;
  cfg->auto_pin_host_buffers = 1;
  cfg->compress_with_dryrun = 0;
  cfg->block_size = 256;
  cfg->auto_cache_release = 0;
  cfg->log_level = 1;
  cfg->max_larget_level = std::numeric_limits<uint64_t>::max();
  cfg->max_memory_footprint = std::numeric_limits<uint64_t>::max();
}

uint32_t mgx_crc32(const void *data, uint64_t bytes) {
  if (!g_crc_ready)
    crc_init();
  const uint8_t *p = (const uint8_t *)data;
  uint32_t c = 0xFFFFFFFFu;
  for (uint64_t i = 0; i < bytes; i++)
    c = g_crc_table[(c ^ p[i]) & 0xFF] ^ (c >> 8);
  return c ^ 0xFFFFFFFFu;
}

int mgx_metadata_serialize(const mgx_metadata *m, uint8_t *out, uint64_t cap,
                           uint64_t *bytes) {
  return mgx_metadata_serialize_coords(m, nullptr, out, cap, bytes);
}

int mgx_metadata_coords(const uint8_t *in, uint64_t bytes, double *coords, uint64_t cap,
                        uint64_t *count) {
  if (!count)
    return MGX_FAILURE;
  uint64_t total = 0;
  int rc = mgx_metadata_size(in, bytes, &total);
  if (rc)
    return rc;
  if (bytes < total)
    return fail(MGX_FAILURE, "stream shorter than its header");
  mgx_metadata md;
  std::vector<double> c;
  rc = parse_header(in + PREAMBLE, total - PREAMBLE, md, &c);
  if (rc)
    return rc;
  *count = md.nonuniform ? c.size() : 0;
  if (*count && (!coords || cap < *count))
    return MGX_OUTPUT_TOO_LARGE;
  if (*count)
    memcpy(coords, c.data(), c.size() * 8);
  return MGX_SUCCESS;
}

int mgx_metadata_serialize_coords(const mgx_metadata *m, const double *coords,
                                  uint8_t *out, uint64_t cap, uint64_t *bytes) {
  if (!m || !bytes || m->D < 1 || m->D > 5)
    return MGX_FAILURE;
  if (m->nonuniform && !coords)
    return fail(MGX_FAILURE, "a non-uniform grid needs its node coordinates");
  std::vector<uint8_t> head = header_bytes(*m, coords);
  *bytes = PREAMBLE + head.size();
  if (!out || cap < *bytes)
    return MGX_OUTPUT_TOO_LARGE;
  memcpy(out, SIGNATURE, 5);
  put_u64_le(out + 5, head.size());
  const uint32_t crc = mgx_crc32(head.data(), head.size());
  for (int i = 0; i < 4; i++)
    out[13 + i] = (uint8_t)(crc >> (8 * i));
  memcpy(out + PREAMBLE, head.data(), head.size());
  return MGX_SUCCESS;
}

int mgx_metadata_size(const uint8_t *in, uint64_t bytes, uint64_t *header_bytes_out) {
  if (!in || bytes < 13 || !header_bytes_out)
    return fail(MGX_FAILURE, "stream shorter than the MGARD preamble");
  if (memcmp(in, SIGNATURE, 5) != 0)
    return fail(MGX_FAILURE, "signature mismatch.");
  uint64_t n = 0;
  for (int k = 0; k < 8; k++)
    n |= (uint64_t)in[5 + k] << (8 * k);
  if (n > ((uint64_t)1 << 31))
    return fail(MGX_FAILURE, "header is too large (size would overflow).");
  *header_bytes_out = n + PREAMBLE;
  return MGX_SUCCESS;
}

int mgx_metadata_parse(const uint8_t *in, uint64_t bytes, mgx_metadata *m,
                       uint64_t *header_bytes_out) {
  if (!m)
    return MGX_FAILURE;
  uint64_t total = 0;
  int rc = mgx_metadata_size(in, bytes, &total);
  if (rc)
    return rc;
  if (bytes < total)
    return fail(MGX_FAILURE, "stream shorter than its header");
  uint32_t crc = 0;
  for (int i = 0; i < 4; i++)
    crc |= (uint32_t)in[13 + i] << (8 * i);
  if (crc != mgx_crc32(in + PREAMBLE, total - PREAMBLE))
    return fail(MGX_FAILURE, "header CRC32 mismatch.");
  rc = parse_header(in + PREAMBLE, total - PREAMBLE, *m);
  if (rc)
    return rc;
  if (header_bytes_out)
    *header_bytes_out = total;
  return MGX_SUCCESS;
}

uint64_t mgx_estimate_footprint(int D, const uint64_t *shape, int dtype, double s,
                                int prefetch) {
  if (D < 1 || D > 5 || !shape)
    return 0;
  return footprint(D, shape, dtype, s, prefetch);
}

int mgx_max_dim_by_memory(int D, const uint64_t *shape, int dtype, double s,
                          uint64_t available, int *decomposed, int *dim,
                          uint64_t *size, uint64_t *num) {
  if (D > 5)
    return MGX_TOO_MANY_DIMENSIONS;
  if (D < 1 || !shape || !decomposed || !dim || !size || !num)
    return MGX_FAILURE;
  // DomainDecomposer(shape, config), DomainDecomposer.hpp:293-330
  if (footprint(D, shape, dtype, s, 0) < available) {
    *decomposed = 0;
    *dim = 0;
    *size = shape[0];
    *num = 1;
    return MGX_SUCCESS;
  }
  uint64_t chunk[5];
  uint64_t best = 0;
  int bd = 0;
  for (int d = 0; d < D; d++) {
    chunk[d] = shape[d];
    if (shape[d] > best) {
      best = shape[d];
      bd = d;
    }
  }
  int prefetch = 0;
  uint64_t n = 1;
  while (footprint(D, chunk, dtype, s, prefetch) >= available) {
    if (chunk[bd] <= 1)
      return MGX_FAILURE; // not even one row fits
    chunk[bd] = (chunk[bd] - 1) / 2 + 1; // divide by 2 and round up
    n = (best - 1) / chunk[bd] + 1;
    prefetch = n > 1;
  }
  *decomposed = 1;
  *dim = bd;
  *size = chunk[bd];
  *num = (best - 1) / chunk[bd] + 1;
  return MGX_SUCCESS;
}

int mgx_x_compress(int D, int dtype, const uint64_t *shape, double tol, double s,
                   int mode, const void *in, void **out, size_t *out_size,
                   const mgx_config *cfg, int output_pre_allocated) {
  std::lock_guard<std::mutex> lock(g_mu);
  mgx_config c;
  if (cfg)
    c = *cfg;
  else
    mgx_config_default(&c);
  g_err.clear();
  DeviceRestore restore;
  try {
    return x_compress(D, dtype, shape, tol, s, mode, in, out, out_size, c,
                      output_pre_allocated);
  } catch (const std::exception &e) { // no exception crosses the C ABI
    return fail(MGX_FAILURE, std::string("compress: ") + e.what());
  } catch (...) {
    return fail(MGX_FAILURE, "compress: unknown exception");
  }
}

int mgx_x_compress_coords(int D, int dtype, const uint64_t *shape, double tol, double s,
                          int mode, const void *in, void **out, size_t *out_size,
                          const void *const *coords, const mgx_config *cfg,
                          int output_pre_allocated) {
  std::lock_guard<std::mutex> lock(g_mu);
  mgx_config c;
  if (cfg)
    c = *cfg;
  else
    mgx_config_default(&c);
  g_err.clear();
  if (!coords)
    return fail(MGX_FAILURE, "coords missing");
  DeviceRestore restore;
  try {
    return x_compress(D, dtype, shape, tol, s, mode, in, out, out_size, c,
                      output_pre_allocated, coords);
  } catch (const std::exception &e) {
    return fail(MGX_FAILURE, std::string("compress: ") + e.what());
  } catch (...) {
    return fail(MGX_FAILURE, "compress: unknown exception");
  }
}

int mgx_x_compress_multi(int D, int dtype, const uint64_t *shape, double tol, double s,
                         int mode, const void *in, void **out, size_t *out_size,
                         const int *devices, int num_devices, const mgx_config *cfg,
                         int output_pre_allocated) {
  std::lock_guard<std::mutex> lock(g_mu);
  mgx_config c;
  if (cfg)
    c = *cfg;
  else
    mgx_config_default(&c);
  g_err.clear();
  DeviceRestore restore;
  try {
    return x_compress_multi(D, dtype, shape, tol, s, mode, in, out, out_size, devices,
                            num_devices, c, output_pre_allocated);
  } catch (const std::exception &e) {
    return fail(MGX_FAILURE, std::string("compress: ") + e.what());
  } catch (...) {
    return fail(MGX_FAILURE, "compress: unknown exception");
  }
}

int mgx_x_decompress_multi(const void *in, size_t in_size, void **out,
                           uint64_t *shape_out, int *D_out, int *dtype_out,
                           const int *devices, int num_devices, const mgx_config *cfg,
                           int output_pre_allocated) {
  std::lock_guard<std::mutex> lock(g_mu);
  mgx_config c;
  if (cfg)
    c = *cfg;
  else
    mgx_config_default(&c);
  g_err.clear();
  DeviceRestore restore;
  try {
    return x_decompress_multi(in, in_size, out, shape_out, D_out, dtype_out, devices,
                              num_devices, c, output_pre_allocated);
  } catch (const std::exception &e) {
    return fail(MGX_FAILURE, std::string("decompress: ") + e.what());
  } catch (...) {
    return fail(MGX_FAILURE, "decompress: unknown exception");
  }
}

int mgx_x_decompress(const void *in, size_t in_size, void **out,
                     uint64_t *shape_out, int *D_out, int *dtype_out,
                     const mgx_config *cfg, int output_pre_allocated) {
  std::lock_guard<std::mutex> lock(g_mu);
  mgx_config c;
  if (cfg)
    c = *cfg;
  else
    mgx_config_default(&c);
  g_err.clear();
  DeviceRestore restore;
  try {
    return x_decompress(in, in_size, out, shape_out, D_out, dtype_out, c,
                        output_pre_allocated);
  } catch (const std::exception &e) {
    return fail(MGX_FAILURE, std::string("decompress: ") + e.what());
  } catch (...) {
    return fail(MGX_FAILURE, "decompress: unknown exception");
  }
}

static int backend_ok(const mgx_config *cfg) {
  int t = cfg ? cfg->dev_type : MGX_DEV_AUTO;
  return t == MGX_DEV_AUTO || t == MGX_DEV_CUDA;
}

int mgx_x_release_cache(const mgx_config *cfg) {
  std::lock_guard<std::mutex> lock(g_mu);
  if (!backend_ok(cfg))
    return MGX_BACKEND_UNAVAILABLE;
  for (auto &kv : g_cache)
    cache_release(kv.second);
  return MGX_SUCCESS;
}

int mgx_x_pin_memory(void *ptr, uint64_t bytes, const mgx_config *cfg) {
  if (!backend_ok(cfg))
    return MGX_BACKEND_UNAVAILABLE;
  if (cudaHostRegister(ptr, bytes, cudaHostRegisterDefault) != cudaSuccess) {
    cudaGetLastError();
    return MGX_FAILURE;
  }
  return MGX_SUCCESS;
}

int mgx_x_check_memory_pinned(void *ptr, const mgx_config *cfg) {
  if (!backend_ok(cfg))
    return 0;
  return host_registered(ptr) ? 1 : 0;
}

int mgx_x_unpin_memory(void *ptr, const mgx_config *cfg) {
  if (!backend_ok(cfg))
    return MGX_BACKEND_UNAVAILABLE;
  if (cudaHostUnregister(ptr) != cudaSuccess) {
    cudaGetLastError();
    return MGX_FAILURE;
  }
  return MGX_SUCCESS;
}

const char *mgx_x_last_error(void) { return g_err.c_str(); }

int mgx_x_have_zstd(void) { return zstd_api().ok ? 1 : 0; }

int mgx_x_have_lz4(void) { return lz4_api().ok ? 1 : 0; }

int mgx_x_lz4_wrap(const uint8_t *in, uint64_t n, uint64_t chunk_size, uint8_t *out,
                   uint64_t cap, uint64_t *out_bytes) {
  if (!in || !out_bytes)
    return fail(MGX_FAILURE, "bad argument");
  std::vector<uint8_t> tmp;
  int rc = lz4_wrap(in, n, chunk_size, tmp);
  if (rc)
    return rc;
  *out_bytes = tmp.size();
  if (!out || cap < tmp.size())
    return MGX_OUTPUT_TOO_LARGE;
  memcpy(out, tmp.data(), tmp.size());
  return MGX_SUCCESS;
}

int mgx_x_lz4_unwrap(const uint8_t *in, uint64_t bytes, uint8_t *out, uint64_t cap,
                     uint64_t *out_bytes) {
  if (!in || !out_bytes)
    return fail(MGX_FAILURE, "bad argument");
  std::vector<uint8_t> tmp;
  int rc = lz4_unwrap(in, bytes, ~0ull >> 1, tmp);
  if (rc)
    return rc;
  *out_bytes = tmp.size();
  if (!out || cap < tmp.size())
    return MGX_OUTPUT_TOO_LARGE;
  memcpy(out, tmp.data(), tmp.size());
  return MGX_SUCCESS;
}

uint64_t mgx_x_launch_count(void) {
  std::lock_guard<std::mutex> lock(g_mu);
  uint64_t n = 0;
  for (auto &kv : g_cache) {
    n += kv.second.retired_launches;
    for (auto &p : kv.second.plans)
      n += mgx_launch_count(p.ctx);
  }
  return n;
}

} // extern "C"
