// cgb_op_dot_static: a dot product that allocates nothing.
//
// Who needs this: code that is not allowed to call cudaMalloc at the point where it needs a dot
// product. Typical cases are a solver step called from inside someone else's time-critical loop,
// where an allocation means an implicit device-wide synchronisation and an unbounded stall; a
// plug-in running under a host framework that owns all device memory and forbids its guests to
// allocate; and out-of-memory or teardown paths that still want a residual norm when the
// allocator has nothing left to give. cgb_op_dot() asks such callers for a scratch buffer they
// cannot provide; this variant brings its own.
//
// The scratch (one 4-byte arrival counter plus one partial sum per 32768-element tile) lives in a
// file-scope __device__ array of kStaticBytes = 65,792 bytes. Every GPU that loads libcgb200.so
// keeps that array until the process exits, whether or not this function is ever called. Its size
// fixes the largest input: kMaxTiles * kTile = 2^28 elements (2 GiB per operand). Larger n is
// rejected with CGB_ERR_INVALID and a message naming the limit; use cgb_op_dot() there.
//
// On the host, the array's name is only a handle (a shadow symbol), not an address the GPU can
// use. The first call on each device therefore asks the runtime for the device address with
// cudaGetSymbolAddress while that device is current, checks the result, and caches it per device.
// That first call is the set-up: make it before any stream capture begins.
//
// All calls on one device share the one array, and it still holds whatever the previous call left
// there. The arrival counter must start at zero, so every call clears it on the caller's stream
// before the kernel; partial sums are fully overwritten before they are read. Two calls that are
// in flight on the same GPU at the same time (different streams, or different host threads) race
// on the counter and the partials and return garbage: serialise them.
//
// The sum is taken tile by tile in a fixed order, so the result does not depend on the grid size
// and repeats bit for bit from call to call.
#include <mutex>

#include "../common/status.hpp"
#include "../kernels/dev_common.cuh"

using namespace cgb;

namespace {

constexpr int kTile = 32768;                 // elements per partial sum
constexpr int kMaxTiles = 8192;              // partial sums the static scratch can hold
constexpr long long kMaxN = (long long)kTile * kMaxTiles;
constexpr size_t kPartialsAt = 256;          // byte offset of the partial sums
constexpr size_t kStaticBytes = kPartialsAt + (size_t)kMaxTiles * sizeof(double);

static __device__ __align__(256) unsigned char g_static_scratch[kStaticBytes];

__global__ void __launch_bounds__(kThreads) k_dot_static(const double* __restrict__ a, const double* __restrict__ b,
                                                          long long n, int ntiles, unsigned char* scratch,
                                                          double* out) {
  __shared__ double red_smem[64];
  __shared__ int s_last;
  unsigned int* ticket = reinterpret_cast<unsigned int*>(scratch);
  double* partials = reinterpret_cast<double*>(scratch + kPartialsAt);
  for (int t = blockIdx.x; t < ntiles; t += gridDim.x) {
    const long long lo = (long long)t * kTile;
    const long long hi = lo + kTile < n ? lo + kTile : n;
    double acc = 0.0, z = 0.0;
    for (long long i = lo + threadIdx.x; i < hi; i += kThreads) acc += a[i] * b[i];
    block_sum2(acc, z, red_smem);
    if (threadIdx.x == 0) __stcg(&partials[t], acc);
  }
  if (threadIdx.x == 0) {
    __threadfence();
    s_last = (atomicAdd(ticket, 1u) == gridDim.x - 1);
  }
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  double sum = 0.0, z = 0.0;
  for (int t = threadIdx.x; t < ntiles; t += kThreads) sum += __ldcg(&partials[t]);
  block_sum2(sum, z, red_smem);
  if (threadIdx.x == 0) {
    *out = sum;
    *ticket = 0u;
  }
}

std::mutex g_mu;
unsigned char* g_scratch_addr[64] = {};  // device address of g_static_scratch, per device

// Once per device, with that device current: resolve the symbol to a device address.
int static_scratch(unsigned char** out) {
  int dev = 0;
  CGB_CUDA(cudaGetDevice(&dev));
  CGB_REQUIRE(dev >= 0 && dev < 64, "device index %d out of range", dev);
  std::lock_guard<std::mutex> lock(g_mu);
  if (!g_scratch_addr[dev]) {
    unsigned char* ptr = nullptr;
    cudaError_t e = cudaGetSymbolAddress((void**)&ptr, g_static_scratch);
    if (e != cudaSuccess || !ptr)
      return fail(CGB_ERR_CUDA, "cudaGetSymbolAddress(static dot scratch) on device %d: %s", dev,
                  cudaGetErrorString(e));
    g_scratch_addr[dev] = ptr;
  }
  *out = g_scratch_addr[dev];
  return CGB_OK;
}

}  // namespace

extern "C" long long cgb_op_dot_static_max_n(void) { return kMaxN; }

extern "C" int cgb_op_dot_static(const double* a, const double* b, int64_t n, double* out, void* stream) {
  CGB_REQUIRE(a && b && out, "cgb_op_dot_static: NULL argument");
  CGB_REQUIRE(n >= 0, "cgb_op_dot_static: negative n");
  if (n > kMaxN)
    return fail(CGB_ERR_INVALID,
                "cgb_op_dot_static: n = %lld exceeds the %lld elements the built-in scratch (%zu bytes) can cover; "
                "use cgb_op_dot with caller-provided scratch",
                (long long)n, kMaxN, kStaticBytes);
  unsigned char* scratch = nullptr;
  CGB_TRY(static_scratch(&scratch));
  cudaStream_t s = (cudaStream_t)stream;
  const int ntiles = (int)((n + kTile - 1) / kTile);
  CGB_CUDA(cudaMemsetAsync(scratch, 0, sizeof(unsigned int), s));  // arrival counter starts at zero
  int grid = ntiles < 1 ? 1 : (ntiles > 1184 ? 1184 : ntiles);
  k_dot_static<<<grid, kThreads, 0, s>>>(a, b, (long long)n, ntiles, scratch, out);
  CGB_CUDA(cudaGetLastError());
  return CGB_OK;
}
