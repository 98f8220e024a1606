// Lossless side channel for the latent coordinates.
//
// The reference ships the stride-8 latent coordinates through the G-PCC reference codec
// (model/model.py:318-395: ASCII PLY temp file -> `tmc3` subprocess -> bytes and back).  tmc3 is an
// empty submodule here (.SUBMODULES.json) and the subprocess + temp files are the largest
// non-network share of the published encode/decode times, so the side channel is restated as an
// octree occupancy code built on the GPU: breadth-first, Morton order, one occupancy byte per
// internal node (bit c set <=> child c occupied, c = 4 x_bit + 2 y_bit + z_bit).  The stream is
// self-delimiting given the depth: level 0 has one node and the node count of level l + 1 is the
// popcount of level l.  Not G-PCC compatible (stated in DESIGN.md).  HBM/latency-bound integer work
// on 18 k (800 k-point frame) .. 112 k (5 M-point scene) voxels: one single-CTA kernel per
// direction, block-wide scans, no host round trips except the final byte count.
#include <cub/block/block_scan.cuh>
#include <cub/device/device_radix_sort.cuh>

#include "pcc_common.cuh"

namespace pcc {

constexpr int OT_THREADS = 1024;

__host__ __device__ __forceinline__ uint64_t spread3(uint32_t v) {  // 16 bits -> every third bit
  uint64_t x = v & 0xFFFFu;
  x = (x | (x << 16)) & 0x00000000FF0000FFull;
  x = (x | (x << 8)) & 0xF00F00F00F00F00Full;
  x = (x | (x << 4)) & 0x30C30C30C30C30C3ull;
  x = (x | (x << 2)) & 0x9249249249249249ull;
  return x;
}
__host__ __device__ __forceinline__ uint32_t compact3(uint64_t x) {
  x &= 0x9249249249249249ull;
  x = (x | (x >> 2)) & 0x30C30C30C30C30C3ull;
  x = (x | (x >> 4)) & 0xF00F00F00F00F00Full;
  x = (x | (x >> 8)) & 0x00FF0000FF0000FFull;
  x = (x | (x >> 16)) & 0x00FF00000000FFFFull;
  x = (x | (x >> 32)) & 0xFFFFull;
  return (uint32_t)x;
}

// keys (b, x, y, z packed) -> Morton codes of (x, y, z) / stride; status |= 1 on a coordinate that is
// negative, not a multiple of the stride, does not fit `depth` bits, or has batch != 0
__global__ void morton_from_keys_kernel(const uint64_t* __restrict__ keys, int64_t n, int stride, int depth,
                                        uint64_t* __restrict__ codes, int32_t* __restrict__ status) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  int b, x, y, z;
  unpack_key(keys[i], b, x, y, z);
  const bool ok = b == 0 && x >= 0 && y >= 0 && z >= 0 && x % stride == 0 && y % stride == 0 && z % stride == 0 &&
                  (x / stride) < (1 << depth) && (y / stride) < (1 << depth) && (z / stride) < (1 << depth);
  if (!ok) atomicOr(status, 1);
  const uint32_t qx = (uint32_t)(x / stride), qy = (uint32_t)(y / stride), qz = (uint32_t)(z / stride);
  codes[i] = (spread3(qx) << 2) | (spread3(qy) << 1) | spread3(qz);
}

__global__ void keys_from_morton_kernel(const uint64_t* __restrict__ codes, int64_t n, int stride,
                                        uint64_t* __restrict__ keys) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint64_t m = codes[i];
  const int x = (int)compact3(m >> 2) * stride, y = (int)compact3(m >> 1) * stride, z = (int)compact3(m) * stride;
  keys[i] = pack_key(0, x, y, z);
}

// One CTA.  codes: sorted unique Morton codes.  out32: zero-initialised, byte b of the stream lives in
// bits 8 (b & 3) .. of out32[b >> 2].  total[0] = stream length in bytes.
__global__ void __launch_bounds__(OT_THREADS)
octree_encode_kernel(const uint64_t* __restrict__ codes, int64_t n, int depth, uint32_t* __restrict__ out32,
                     int64_t* __restrict__ total) {
  using Scan = cub::BlockScan<int, OT_THREADS>;
  __shared__ typename Scan::TempStorage tmp;
  __shared__ int64_t s_carry;
  int64_t base = 0;  // bytes written by the previous levels (uniform across the block)
  for (int l = 0; l < depth; ++l) {
    const int s = 3 * (depth - l);  // node prefix = code >> s; child index = (code >> (s - 3)) & 7
    if (threadIdx.x == 0) s_carry = 0;
    __syncthreads();
    for (int64_t i0 = 0; i0 < n; i0 += OT_THREADS) {
      const int64_t i = i0 + threadIdx.x;
      int head = 0;
      uint64_t m = 0;
      if (i < n) {
        m = codes[i];
        head = (i == 0) || ((m >> s) != (codes[i - 1] >> s));
      }
      int incl, agg;
      Scan(tmp).InclusiveSum(head, incl, agg);
      const int64_t carry = s_carry;
      if (i < n) {
        const int64_t node = base + carry + incl - 1;
        atomicOr(out32 + (node >> 2), (1u << (int)((m >> (s - 3)) & 7)) << (8 * (int)(node & 3)));
      }
      __syncthreads();
      if (threadIdx.x == 0) s_carry = carry + agg;
      __syncthreads();
    }
    base += s_carry;
    __syncthreads();
  }
  if (threadIdx.x == 0) total[0] = base;
}

// One CTA.  bytes: the stream; a / b: ping-pong code buffers of capacity cap.  result[0] = number of
// leaves (or -1 when the stream is truncated / exceeds cap), result[1] = bytes consumed,
// result[2] = 0 when the leaves ended in buffer a, 1 for b.
__global__ void __launch_bounds__(OT_THREADS)
octree_decode_kernel(const uint8_t* __restrict__ bytes, int64_t nbytes, int depth, uint64_t* __restrict__ a,
                     uint64_t* __restrict__ b, int64_t cap, int64_t* __restrict__ result) {
  using Scan = cub::BlockScan<int, OT_THREADS>;
  __shared__ typename Scan::TempStorage tmp;
  __shared__ int64_t s_carry;
  uint64_t* cur = a;
  uint64_t* nxt = b;
  int64_t count = 1, pos = 0;
  if (threadIdx.x == 0) cur[0] = 0;
  __syncthreads();
  for (int l = 0; l < depth; ++l) {
    if (pos + count > nbytes) {  // uniform
      if (threadIdx.x == 0) result[0] = -1;
      return;
    }
    if (threadIdx.x == 0) s_carry = 0;
    __syncthreads();
    for (int64_t i0 = 0; i0 < count; i0 += OT_THREADS) {
      const int64_t i = i0 + threadIdx.x;
      const uint32_t occ = i < count ? bytes[pos + i] : 0u;
      const int c = __popc(occ);
      int excl, agg;
      Scan(tmp).ExclusiveSum(c, excl, agg);
      const int64_t carry = s_carry;
      if (carry + agg > cap) {  // uniform
        if (threadIdx.x == 0) result[0] = -1;
        return;
      }
      if (i < count) {
        const uint64_t node = cur[i];
        int64_t w = carry + excl;
        uint32_t o = occ;
        while (o) {
          const int ch = __ffs(o) - 1;
          o &= o - 1;
          nxt[w++] = (node << 3) | (uint64_t)ch;
        }
      }
      __syncthreads();
      if (threadIdx.x == 0) s_carry = carry + agg;
      __syncthreads();
    }
    pos += count;
    count = s_carry;
    uint64_t* t = cur;
    cur = nxt;
    nxt = t;
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    result[0] = count;
    result[1] = pos;
    result[2] = (cur == a) ? 0 : 1;
  }
}

static size_t cub_sort_codes_bytes(int64_t n) {
  size_t bytes = 0;
  cub::DeviceRadixSort::SortKeys(nullptr, bytes, (const uint64_t*)nullptr, (uint64_t*)nullptr, (int)n);
  return bytes;
}

static int64_t* pinned_result() {
  static thread_local int64_t* p = [] {
    int64_t* q = nullptr;
    if (cudaHostAlloc(reinterpret_cast<void**>(&q), 64, cudaHostAllocDefault) != cudaSuccess) q = nullptr;
    return q;
  }();
  return p;
}

}  // namespace pcc

using namespace pcc;

extern "C" {

size_t pcc_octree_ws_bytes(int64_t n) {
  if (n < 1) n = 1;
  return 2 * align_up((size_t)n * 8, 256) + align_up(cub_sort_codes_bytes(n), 256) + 512;
}

int pcc_octree_encode(const uint64_t* keys, int64_t n, int32_t stride, int32_t depth, uint8_t* out, int64_t out_cap,
                      int64_t* nbytes, void* ws, size_t ws_bytes, void* stream) {
  PCC_REQUIRE(nbytes, "null nbytes");
  *nbytes = 0;
  PCC_REQUIRE(n >= 0 && n < (1ll << 31) && stride > 0 && depth >= 1 && depth <= 16, "bad argument");
  if (n == 0) return PCC_OK;
  PCC_REQUIRE(keys && out && ws, "null pointer");
  PCC_REQUIRE(ws_bytes >= pcc_octree_ws_bytes(n), "workspace too small");
  PCC_REQUIRE(out_cap >= ((n * depth + 3) / 4) * 4 && (reinterpret_cast<uintptr_t>(out) & 3) == 0,
              "out must be 4-byte aligned with capacity >= n * depth rounded up to 4");
  cudaStream_t st = (cudaStream_t)stream;
  char* w = reinterpret_cast<char*>(ws);
  uint64_t* codes = reinterpret_cast<uint64_t*>(w);
  uint64_t* sorted = reinterpret_cast<uint64_t*>(w + align_up((size_t)n * 8, 256));
  char* cub_ws = w + 2 * align_up((size_t)n * 8, 256);
  size_t cub_bytes = align_up(cub_sort_codes_bytes(n), 256);
  int64_t* total = reinterpret_cast<int64_t*>(cub_ws + cub_bytes);
  int32_t* status = reinterpret_cast<int32_t*>(total + 1);
  PCC_CHECK_CUDA(cudaMemsetAsync(total, 0, 16, st));
  PCC_CHECK_CUDA(cudaMemsetAsync(out, 0, (size_t)((n * depth + 3) / 4) * 4, st));
  morton_from_keys_kernel<<<(unsigned)ceil_div(n, 256), 256, 0, st>>>(keys, n, stride, depth, codes, status);
  PCC_CHECK_LAUNCH("morton_from_keys_kernel");
  PCC_CHECK_CUDA(cub::DeviceRadixSort::SortKeys(cub_ws, cub_bytes, codes, sorted, (int)n, 0, 3 * depth, st));
  octree_encode_kernel<<<1, OT_THREADS, 0, st>>>(sorted, n, depth, reinterpret_cast<uint32_t*>(out), total);
  PCC_CHECK_LAUNCH("octree_encode_kernel");
  int64_t hbuf[2] = {0, 0};
  int64_t* h = pinned_result() ? pinned_result() : hbuf;
  PCC_CHECK_CUDA(cudaMemcpyAsync(h, total, 16, cudaMemcpyDeviceToHost, st));
  PCC_CHECK_CUDA(cudaStreamSynchronize(st));
  if ((int32_t)h[1] != 0) {
    set_error("pcc_octree_encode: coordinate negative, off the stride grid, beyond 2^depth, or batch != 0");
    return PCC_ERR_INVALID_ARG;
  }
  *nbytes = h[0];
  return PCC_OK;
}

int pcc_octree_decode(const uint8_t* bytes, int64_t nbytes, int32_t stride, int32_t depth, uint64_t* keys_out,
                      int64_t cap, int64_t* n_out, void* ws, size_t ws_bytes, void* stream) {
  PCC_REQUIRE(n_out, "null n_out");
  *n_out = 0;
  PCC_REQUIRE(nbytes >= 0 && stride > 0 && depth >= 1 && depth <= 16 && cap >= 1, "bad argument");
  if (nbytes == 0) return PCC_OK;
  PCC_REQUIRE(bytes && keys_out && ws, "null pointer");
  PCC_REQUIRE(ws_bytes >= pcc_octree_ws_bytes(cap), "workspace too small");
  cudaStream_t st = (cudaStream_t)stream;
  char* w = reinterpret_cast<char*>(ws);
  uint64_t* a = reinterpret_cast<uint64_t*>(w);
  uint64_t* b = reinterpret_cast<uint64_t*>(w + align_up((size_t)cap * 8, 256));
  int64_t* result = reinterpret_cast<int64_t*>(w + 2 * align_up((size_t)cap * 8, 256));
  octree_decode_kernel<<<1, OT_THREADS, 0, st>>>(bytes, nbytes, depth, a, b, cap, result);
  PCC_CHECK_LAUNCH("octree_decode_kernel");
  int64_t hbuf[3] = {0, 0, 0};
  int64_t* h = pinned_result() ? pinned_result() : hbuf;
  PCC_CHECK_CUDA(cudaMemcpyAsync(h, result, 24, cudaMemcpyDeviceToHost, st));
  PCC_CHECK_CUDA(cudaStreamSynchronize(st));
  if (h[0] < 0 || h[1] != nbytes) {
    set_error("pcc_octree_decode: malformed stream (truncated, trailing bytes, or more than cap leaves)");
    return PCC_ERR_INVALID_ARG;
  }
  const int64_t n = h[0];
  keys_from_morton_kernel<<<(unsigned)ceil_div(n, 256), 256, 0, st>>>(h[2] ? b : a, n, stride, keys_out);
  PCC_CHECK_LAUNCH("keys_from_morton_kernel");
  *n_out = n;
  return PCC_OK;
}

}  // extern "C"
