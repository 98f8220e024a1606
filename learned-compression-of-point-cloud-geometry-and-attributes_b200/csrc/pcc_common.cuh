// Shared device/host helpers for the pcc_b200 kernels (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include "../../include/pcc_b200.h"

namespace pcc {

// ---------------------------------------------------------------- error plumbing
void set_error(const char* fmt, ...);
int cuda_fail(cudaError_t e, const char* what);

#define PCC_CHECK_CUDA(expr)                                   \
  do {                                                         \
    cudaError_t _e = (expr);                                   \
    if (_e != cudaSuccess) return ::pcc::cuda_fail(_e, #expr); \
  } while (0)

#define PCC_CHECK_LAUNCH(name)                                   \
  do {                                                           \
    cudaError_t _e = cudaGetLastError();                         \
    if (_e != cudaSuccess) return ::pcc::cuda_fail(_e, name);    \
  } while (0)

#define PCC_REQUIRE(cond, msg)               \
  do {                                       \
    if (!(cond)) {                           \
      ::pcc::set_error("%s: %s", __func__, msg); \
      return PCC_ERR_INVALID_ARG;            \
    }                                        \
  } while (0)

int sm_count();

static inline int64_t ceil_div(int64_t a, int64_t b) { return (a + b - 1) / b; }
static inline size_t align_up(size_t a, size_t b) { return (a + b - 1) / b * b; }

// ---------------------------------------------------------------- coordinate keys
// key order == lexicographic (b, x, y, z); see include/pcc_b200.h
__host__ __device__ __forceinline__ uint64_t pack_key(int b, int x, int y, int z) {
  return ((uint64_t)(uint16_t)b << 48) | ((uint64_t)(uint16_t)(x + 32768) << 32) |
         ((uint64_t)(uint16_t)(y + 32768) << 16) | (uint64_t)(uint16_t)(z + 32768);
}
__host__ __device__ __forceinline__ void unpack_key(uint64_t k, int& b, int& x, int& y, int& z) {
  b = (int)(k >> 48);
  x = (int)((k >> 32) & 0xFFFF) - 32768;
  y = (int)((k >> 16) & 0xFFFF) - 32768;
  z = (int)(k & 0xFFFF) - 32768;
}
__host__ __device__ __forceinline__ bool coord_in_range(int v) { return v >= -32768 && v < 32768; }

// ---------------------------------------------------------------- hash table
// Open addressing over 32-byte buckets of four 8-byte keys (one L2 sector per probe step);
// values live in a parallel int32 array.  Keys are unique at insert time (the maps are sorted
// and de-duplicated before hashing), so insert is a single atomicCAS on the first free slot.
constexpr uint64_t HASH_EMPTY = 0xFFFFFFFFFFFFFFFFull;
constexpr int HASH_BUCKET = 4;

__host__ __device__ __forceinline__ uint64_t hash_mix(uint64_t k) {
  k ^= k >> 33;
  k *= 0xff51afd7ed558ccdull;
  k ^= k >> 33;
  k *= 0xc4ceb9fe1a85ec53ull;
  k ^= k >> 33;
  return k;
}

// Returns the value stored for `key`, or -1.  cap is a power of two (multiple of HASH_BUCKET).
__device__ __forceinline__ int hash_find(const uint64_t* __restrict__ tkeys,
                                         const int32_t* __restrict__ tvals, uint64_t cap_mask,
                                         uint64_t key) {
  uint64_t b = (hash_mix(key) & cap_mask) & ~(uint64_t)(HASH_BUCKET - 1);
  for (;;) {
    const ulonglong2 lo = __ldg(reinterpret_cast<const ulonglong2*>(tkeys + b));
    const ulonglong2 hi = __ldg(reinterpret_cast<const ulonglong2*>(tkeys + b + 2));
    if (lo.x == key) return __ldg(tvals + b);
    if (lo.y == key) return __ldg(tvals + b + 1);
    if (hi.x == key) return __ldg(tvals + b + 2);
    if (hi.y == key) return __ldg(tvals + b + 3);
    // slots of a bucket fill front to back, so an empty last slot ends the probe sequence
    if (hi.y == HASH_EMPTY) return -1;
    b = (b + HASH_BUCKET) & cap_mask;
  }
}

// round an fp32 value to the nearest tf32 (10-bit mantissa), ties away from zero
__device__ __forceinline__ float round_tf32(float x) {
  uint32_t u;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(u) : "f"(x));
  return __uint_as_float(u);
}

// ---------------------------------------------------------------- device-wide exclusive scan
// Three-phase int32 exclusive scan (per-block reduce, single-block scan of the block sums,
// per-block down-sweep).  ws holds ceil(n / SCAN_TILE) + 1 int32; the grand total is written to
// ws_sums[num_tiles] for the caller (device side) and can be copied to the host.
constexpr int SCAN_THREADS = 256;
constexpr int SCAN_ITEMS = 8;
constexpr int SCAN_TILE = SCAN_THREADS * SCAN_ITEMS;
size_t scan_ws_bytes(int64_t n);
// in/out may alias. total_out [dev, 1 int32] may be NULL.
int exclusive_scan_i32(const int32_t* in, int32_t* out, int64_t n, int32_t* total_out, void* ws,
                       cudaStream_t stream);

}  // namespace pcc
