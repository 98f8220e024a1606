// Error plumbing, device info and the device-wide scan shared by the pcc_b200 kernels.
#include <stdarg.h>

#include "pcc_common.cuh"

namespace pcc {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int cuda_fail(cudaError_t e, const char* what) {
  set_error("CUDA error %d (%s) at %s", (int)e, cudaGetErrorString(e), what);
  return e == cudaErrorMemoryAllocation ? PCC_ERR_OOM : PCC_ERR_CUDA;
}

int sm_count() {
  static int cached = 0;
  if (cached == 0) {
    int dev = 0, n = 0;
    if (cudaGetDevice(&dev) == cudaSuccess &&
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && n > 0)
      cached = n;
    else
      return 148;
  }
  return cached;
}

// ---------------------------------------------------------------- scan
__global__ void __launch_bounds__(SCAN_THREADS)
scan_reduce_kernel(const int32_t* __restrict__ in, int64_t n, int32_t* __restrict__ sums) {
  __shared__ int32_t warp_sums[SCAN_THREADS / 32];
  const int64_t base = (int64_t)blockIdx.x * SCAN_TILE;
  int32_t s = 0;
#pragma unroll
  for (int i = 0; i < SCAN_ITEMS; ++i) {
    const int64_t idx = base + (int64_t)i * SCAN_THREADS + threadIdx.x;
    if (idx < n) s += in[idx];
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if ((threadIdx.x & 31) == 0) warp_sums[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    int32_t t = 0;
    for (int w = 0; w < SCAN_THREADS / 32; ++w) t += warp_sums[w];
    sums[blockIdx.x] = t;
  }
}

// single block: exclusive scan of `m` block sums in place; total -> sums[m] (and total_out)
__global__ void __launch_bounds__(1024)
scan_sums_kernel(int32_t* __restrict__ sums, int64_t m, int32_t* __restrict__ total_out) {
  __shared__ int32_t warp_tot[32];
  __shared__ int32_t carry_s;
  if (threadIdx.x == 0) carry_s = 0;
  __syncthreads();
  for (int64_t base = 0; base < m; base += 1024) {
    const int64_t idx = base + threadIdx.x;
    const int32_t v = idx < m ? sums[idx] : 0;
    int32_t inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int32_t t = __shfl_up_sync(0xffffffffu, inc, o);
      if ((threadIdx.x & 31) >= o) inc += t;
    }
    if ((threadIdx.x & 31) == 31) warp_tot[threadIdx.x >> 5] = inc;
    __syncthreads();
    if (threadIdx.x < 32) {
      int32_t w = warp_tot[threadIdx.x];
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int32_t t = __shfl_up_sync(0xffffffffu, w, o);
        if (threadIdx.x >= o) w += t;
      }
      warp_tot[threadIdx.x] = w;  // inclusive over warps
    }
    __syncthreads();
    const int32_t carry = carry_s;
    const int32_t warp_prefix = (threadIdx.x >> 5) ? warp_tot[(threadIdx.x >> 5) - 1] : 0;
    if (idx < m) sums[idx] = carry + warp_prefix + inc - v;
    __syncthreads();
    if (threadIdx.x == 0) carry_s = carry + warp_tot[31];
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    sums[m] = carry_s;
    if (total_out) *total_out = carry_s;
  }
}

__global__ void __launch_bounds__(SCAN_THREADS)
scan_downsweep_kernel(const int32_t* __restrict__ in, int32_t* __restrict__ out, int64_t n,
                      const int32_t* __restrict__ sums) {
  // items are laid out blocked per thread so that the scan order equals the index order
  __shared__ int32_t warp_tot[SCAN_THREADS / 32];
  const int64_t base = (int64_t)blockIdx.x * SCAN_TILE + (int64_t)threadIdx.x * SCAN_ITEMS;
  int32_t v[SCAN_ITEMS];
  int32_t tsum = 0;
#pragma unroll
  for (int i = 0; i < SCAN_ITEMS; ++i) {
    v[i] = (base + i < n) ? in[base + i] : 0;
    tsum += v[i];
  }
  int32_t inc = tsum;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int32_t t = __shfl_up_sync(0xffffffffu, inc, o);
    if ((threadIdx.x & 31) >= o) inc += t;
  }
  if ((threadIdx.x & 31) == 31) warp_tot[threadIdx.x >> 5] = inc;
  __syncthreads();
  int32_t prefix = sums[blockIdx.x] + inc - tsum;
  for (int w = 0; w < (int)(threadIdx.x >> 5); ++w) prefix += warp_tot[w];
#pragma unroll
  for (int i = 0; i < SCAN_ITEMS; ++i) {
    if (base + i < n) out[base + i] = prefix;
    prefix += v[i];
  }
}

size_t scan_ws_bytes(int64_t n) { return align_up((size_t)(ceil_div(n, SCAN_TILE) + 2) * 4, 256); }

int exclusive_scan_i32(const int32_t* in, int32_t* out, int64_t n, int32_t* total_out, void* ws,
                       cudaStream_t stream) {
  int32_t* sums = reinterpret_cast<int32_t*>(ws);
  const int64_t tiles = ceil_div(n, SCAN_TILE);
  if (n > 0) {
    scan_reduce_kernel<<<(unsigned)tiles, SCAN_THREADS, 0, stream>>>(in, n, sums);
    PCC_CHECK_LAUNCH("scan_reduce");
  }
  scan_sums_kernel<<<1, 1024, 0, stream>>>(sums, tiles, total_out);
  PCC_CHECK_LAUNCH("scan_sums");
  if (n > 0) {
    scan_downsweep_kernel<<<(unsigned)tiles, SCAN_THREADS, 0, stream>>>(in, out, n, sums);
    PCC_CHECK_LAUNCH("scan_downsweep");
  }
  return PCC_OK;
}

}  // namespace pcc

extern "C" {

int pcc_version(void) { return 100; }

const char* pcc_last_error(void) { return pcc::g_err; }

int pcc_device_info(int* sm_count, int* cc) {
  int dev = 0, n = 0, major = 0, minor = 0;
  PCC_CHECK_CUDA(cudaGetDevice(&dev));
  PCC_CHECK_CUDA(cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev));
  PCC_CHECK_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
  PCC_CHECK_CUDA(cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, dev));
  if (sm_count) *sm_count = n;
  if (cc) *cc = major * 10 + minor;
  return PCC_OK;
}

}  // extern "C"
