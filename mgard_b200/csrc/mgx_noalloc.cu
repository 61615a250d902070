// mgx_norm_partial_noalloc: the norm reduction (CalculateNorm's device part)
// for callers that may not allocate -- code under CUDA graph capture, tight
// serving loops, harnesses that forbid cudaMalloc.  No counterpart in the
// reference.
//
// The regular entry point (mgx_norm_partial) keeps its per-block partial sums
// in a workspace the context cudaMalloc'ed.  Here that scratch is a file-scope
// `static __device__` array of a fixed size: every GPU that loads the library
// holds MGX_NOALLOC_SCRATCH_BYTES (16 KB) until the process exits, and the
// entry point calls no cudaMalloc / cudaFree / allocating library routine.
//
// One block reduces one fixed segment of MGX_NOALLOC_SEG elements into
// scratch[block]; a second, single-block kernel combines the first `nblocks`
// entries in a fixed order and writes the result to the caller's device
// double.  Hence: at most MGX_NOALLOC_MAX_BLOCKS * MGX_NOALLOC_SEG = 2^31
// elements per call, results independent of the grid, bit-identical from run to
// run.  Nothing in the scratch has to start at zero (every entry that is read
// was written by the same call; stale entries of a longer earlier call are
// never read).  ALL calls on a device share the one array: calls that overlap on
// the GPU (different streams, or replays of two graphs at once) race.
#include <atomic>
#include <cstdio>
#include <cstring>
#include <type_traits>

#include "mgx_device.cuh"
#include "mgx_internal.h"

namespace {

constexpr int NA_MAX_BLOCKS = 2048;
constexpr uint64_t NA_SEG = 1ull << 20; // elements per block
constexpr int NA_THREADS = 512;
constexpr int NA_MAX_DEVICES = 64;

// the scratch: per-block partial results (max |u| or sum u^2, as doubles)
static __device__ __align__(256) double g_noalloc_scratch[NA_MAX_BLOCKS];

// device address of g_noalloc_scratch per device.  In host code the array's
// name is only a host-side stand-in for the symbol, NOT a device address: the
// address a kernel may be given comes from cudaGetSymbolAddress, asked once per
// device with that device current.
std::atomic<double *> g_scratch_addr[NA_MAX_DEVICES]; // device address of the scratch, per device

template <typename T, bool LINF>
__global__ void __launch_bounds__(NA_THREADS)
    k_norm_segment(const T *__restrict__ v, uint64_t n, double *__restrict__ part) {
  const uint64_t b0 = (uint64_t)blockIdx.x * NA_SEG;
  const uint64_t b1 = min(n, b0 + NA_SEG);
  double acc = 0.0;
  T tmax = 0;
  for (uint64_t i = b0 + threadIdx.x; i < b1; i += NA_THREADS) {
    const T x = v[i];
    if (LINF)
      tmax = fmax(tmax, fabs(x));
    else
      acc += (double)x * (double)x;
  }
  if (LINF)
    acc = (double)tmax;
  for (int o = 16; o > 0; o >>= 1) {
    const double other = __shfl_xor_sync(0xffffffffu, acc, o);
    acc = LINF ? fmax(acc, other) : acc + other;
  }
  __shared__ double s[NA_THREADS / 32];
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0)
    s[w] = acc;
  __syncthreads();
  if (w == 0) {
    acc = lane < NA_THREADS / 32 ? s[lane] : 0.0;
    for (int o = 8; o > 0; o >>= 1) {
      const double other = __shfl_xor_sync(0xffffffffu, acc, o);
      acc = LINF ? fmax(acc, other) : acc + other;
    }
    if (lane == 0)
      part[blockIdx.x] = acc;
  }
}

// fixed-order combination of part[0 .. nblocks)
template <bool LINF>
__global__ void __launch_bounds__(1024)
    k_norm_combine(const double *__restrict__ part, int nblocks,
                   double *__restrict__ out) {
  __shared__ double s[1024];
  double acc = 0.0;
  for (int i = threadIdx.x; i < nblocks; i += 1024)
    acc = LINF ? fmax(acc, part[i]) : acc + part[i];
  s[threadIdx.x] = acc;
  __syncthreads();
  for (int o = 512; o > 0; o >>= 1) {
    if ((int)threadIdx.x < o)
      s[threadIdx.x] = LINF ? fmax(s[threadIdx.x], s[threadIdx.x + o])
                            : s[threadIdx.x] + s[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0)
    *out = s[0];
}

int fail(char *err, uint64_t cap, int rc, const char *msg) {
  if (err && cap) {
    std::strncpy(err, msg, cap - 1);
    err[cap - 1] = 0;
  }
  return rc;
}

} // namespace

extern "C" int mgx_norm_partial_noalloc(int device, void *cuda_stream,
                                        const void *d_in, int dtype, uint64_t n,
                                        int linf, double *d_partial, char *err,
                                        uint64_t err_cap) {
  if (err && err_cap)
    err[0] = 0;
  if (dtype != MGX_F32 && dtype != MGX_F64)
    return fail(err, err_cap, MGX_BAD_DATA_TYPE, "dtype must be MGX_F32 or MGX_F64");
  if (n == 0 || !d_in || !d_partial)
    return fail(err, err_cap, MGX_FAILURE, "null pointer or empty input");
  const uint64_t max_n = (uint64_t)NA_MAX_BLOCKS * NA_SEG;
  if (n > max_n) {
    char msg[160];
    std::snprintf(msg, sizeof msg,
                  "mgx_norm_partial_noalloc: %llu elements exceed the fixed scratch "
                  "(maximum %llu = %d blocks x %llu)",
                  (unsigned long long)n, (unsigned long long)max_n, NA_MAX_BLOCKS,
                  (unsigned long long)NA_SEG);
    return fail(err, err_cap, MGX_OUTPUT_TOO_LARGE, msg);
  }
  if (device < 0 || device >= NA_MAX_DEVICES)
    return fail(err, err_cap, MGX_BACKEND_UNAVAILABLE, "device index out of range");
  int prev = -1;
  if (cudaGetDevice(&prev) != cudaSuccess || cudaSetDevice(device) != cudaSuccess)
    return fail(err, err_cap, MGX_BACKEND_UNAVAILABLE,
                "no such CUDA device (there is no CPU fallback)");
  int rc = MGX_SUCCESS;
  const char *what = nullptr;
  if (!g_scratch_addr[device].load(std::memory_order_acquire)) { // set-up, once per device, that device current
    double *p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress((void **)&p, g_noalloc_scratch);
    if (e != cudaSuccess || !p) {
      rc = MGX_FAILURE;
      what = cudaGetErrorString(e);
    } else {
      g_scratch_addr[device].store(p, std::memory_order_release);
    }
  }
  if (!rc) {
    double *scratch = g_scratch_addr[device].load(std::memory_order_acquire);
    cudaStream_t st = (cudaStream_t)cuda_stream;
    const int nblocks = (int)((n + NA_SEG - 1) / NA_SEG);
    if (dtype == MGX_F32) {
      if (linf)
        k_norm_segment<float, true><<<nblocks, NA_THREADS, 0, st>>>((const float *)d_in, n, scratch);
      else
        k_norm_segment<float, false><<<nblocks, NA_THREADS, 0, st>>>((const float *)d_in, n, scratch);
    } else {
      if (linf)
        k_norm_segment<double, true><<<nblocks, NA_THREADS, 0, st>>>((const double *)d_in, n, scratch);
      else
        k_norm_segment<double, false><<<nblocks, NA_THREADS, 0, st>>>((const double *)d_in, n, scratch);
    }
    if (linf)
      k_norm_combine<true><<<1, 1024, 0, st>>>(scratch, nblocks, d_partial);
    else
      k_norm_combine<false><<<1, 1024, 0, st>>>(scratch, nblocks, d_partial);
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
      rc = MGX_FAILURE;
      what = cudaGetErrorString(e);
    }
  }
  if (prev >= 0 && prev != device)
    cudaSetDevice(prev);
  if (rc)
    return fail(err, err_cap, rc, what ? what : "CUDA error");
  return MGX_SUCCESS;
}
