// Coordinate handling: key packing, canonical sort, hash build/probe, stride-2 and generative
// output maps, kernel maps.  Replaces MinkowskiEngine's coordinate manager on the reference
// path (see include/pcc_b200.h for the call sites).  All HBM-bound integer work.
#include <cub/device/device_radix_sort.cuh>

#include "pcc_common.cuh"

namespace pcc {

constexpr int THREADS = 256;

static inline unsigned grid_for(int64_t n, int threads = THREADS) {
  { const int64_t g = ceil_div(n, threads); return (unsigned)(g < 1 ? 1 : g); }
}

// ---------------------------------------------------------------- pack / unpack
__global__ void pack_kernel(const int4* __restrict__ coords, int64_t n, uint64_t* __restrict__ keys,
                            int32_t* __restrict__ status) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int4 c = coords[i];
  if (status && !(coord_in_range(c.y) && coord_in_range(c.z) && coord_in_range(c.w) && c.x >= 0 &&
                  c.x < 65536))
    atomicOr(status, 1);
  keys[i] = pack_key(c.x, c.y, c.z, c.w);
}

__global__ void unpack_kernel(const uint64_t* __restrict__ keys, int64_t n, int4* __restrict__ coords) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  int4 c;
  unpack_key(keys[i], c.x, c.y, c.z, c.w);
  coords[i] = c;
}

__global__ void iota_kernel(int32_t* __restrict__ v, int64_t n) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) v[i] = (int32_t)i;
}

__global__ void take_keys_kernel(const uint64_t* __restrict__ keys, const int32_t* __restrict__ rows,
                                 int64_t n, uint64_t* __restrict__ out) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = keys[rows[i]];
}

// ---------------------------------------------------------------- hash build / probe
__global__ void hash_fill_kernel(uint64_t* __restrict__ tkeys, int64_t cap) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < cap) tkeys[i] = HASH_EMPTY;
}

__global__ void hash_insert_kernel(const uint64_t* __restrict__ keys, int64_t n,
                                   uint64_t* __restrict__ tkeys, int32_t* __restrict__ tvals,
                                   uint64_t cap_mask) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint64_t key = keys[i];
  uint64_t b = (hash_mix(key) & cap_mask) & ~(uint64_t)(HASH_BUCKET - 1);
  for (;;) {
#pragma unroll
    for (int s = 0; s < HASH_BUCKET; ++s) {
      unsigned long long* slot = reinterpret_cast<unsigned long long*>(tkeys + b + s);
      unsigned long long cur = *reinterpret_cast<volatile unsigned long long*>(slot);
      if (cur == HASH_EMPTY) cur = atomicCAS(slot, (unsigned long long)HASH_EMPTY, (unsigned long long)key);
      if (cur == HASH_EMPTY || cur == key) {  // claimed (or duplicate key: last writer wins)
        tvals[b + s] = (int32_t)i;
        return;
      }
    }
    b = (b + HASH_BUCKET) & cap_mask;
  }
}

__global__ void hash_probe_kernel(const uint64_t* __restrict__ tkeys, const int32_t* __restrict__ tvals,
                                  uint64_t cap_mask, const uint64_t* __restrict__ q, int64_t nq,
                                  int32_t* __restrict__ rows) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < nq) rows[i] = hash_find(tkeys, tvals, cap_mask, q[i]);
}

// ---------------------------------------------------------------- stride map
// parent key = floor(c / ts) * ts per spatial axis; ts is a power of two on this path but any
// positive stride is handled (floor division, valid for negative coordinates).
__device__ __forceinline__ int floor_to(int v, int ts) {
  int q = v / ts;
  if ((v % ts) != 0 && v < 0) --q;
  return q * ts;
}

__global__ void stride_keys_kernel(const uint64_t* __restrict__ keys, int64_t n, int ts,
                                   uint64_t* __restrict__ out) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  int b, x, y, z;
  unpack_key(keys[i], b, x, y, z);
  out[i] = pack_key(b, floor_to(x, ts), floor_to(y, ts), floor_to(z, ts));
}

// flags[i] = 1 when sorted[i] starts a new run
__global__ void head_flags_kernel(const uint64_t* __restrict__ sorted, int64_t n, int32_t* __restrict__ flags) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) flags[i] = (i == 0 || sorted[i] != sorted[i - 1]) ? 1 : 0;
}

__global__ void compact_heads_kernel(const uint64_t* __restrict__ sorted, int64_t n,
                                     const int32_t* __restrict__ pos, uint64_t* __restrict__ out) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n && (i == 0 || sorted[i] != sorted[i - 1])) out[pos[i]] = sorted[i];
}

// ---------------------------------------------------------------- generative expand
// Candidate q = p + d * ts_out (ts_out = ts_in / 2).  Along an axis with d != 0, q lies between
// p and p + 2 d ts_out, so the parents that also generate q are p + (e .* d) * ts_in for
// e in {0,1}^3 (e_a = 0 where d_a = 0).  p owns q iff none of those that is lexicographically
// smaller than p exists.  The alternatives are among p's 13 lexicographically smaller coarse neighbours,
// so 13 probes per parent decide ownership of all 27 candidates.
__device__ __forceinline__ uint32_t coarse_neighbour_mask(const uint64_t* tkeys, const int32_t* tvals,
                                                          uint64_t cap_mask, int b, int x, int y, int z,
                                                          int ts) {
  uint32_t m = 1u << 13;
  // ownership only asks about the lexicographically SMALLER coarse neighbours (owned_candidates reads no other
  // bit): 13 of the 26, probed as independent, fully unrolled sequences so that they are in flight together
#pragma unroll
  for (int k = 0; k < 27; ++k) {
    const int dx = k % 3 - 1, dy = (k / 3) % 3 - 1, dz = k / 9 - 1;
    const int first = dx != 0 ? dx : (dy != 0 ? dy : dz);
    if (first >= 0) continue;  // compile-time
    const int qx = x + dx * ts, qy = y + dy * ts, qz = z + dz * ts;
    if (coord_in_range(qx) && coord_in_range(qy) && coord_in_range(qz) &&
        hash_find(tkeys, tvals, cap_mask, pack_key(b, qx, qy, qz)) >= 0)
      m |= 1u << k;
  }
  return m;
}

// bit k of the result: this parent owns candidate d_k
__device__ __forceinline__ uint32_t owned_candidates(uint32_t nmask) {
  uint32_t own = 0;
  for (int k = 0; k < 27; ++k) {
    const int d[3] = {k % 3 - 1, (k / 3) % 3 - 1, k / 9 - 1};
    bool owner = true;
    for (int e = 1; e < 8 && owner; ++e) {
      int a[3];
      bool valid = true;
      for (int ax = 0; ax < 3; ++ax) {
        const int on = (e >> ax) & 1;
        if (on && d[ax] == 0) valid = false;
        a[ax] = on ? d[ax] : 0;
      }
      if (!valid) continue;
      // lexicographic order is (x, y, z): first non-zero component decides
      const int first = a[0] != 0 ? a[0] : (a[1] != 0 ? a[1] : a[2]);
      if (first < 0) {
        const int kk = (a[0] + 1) + 3 * (a[1] + 1) + 9 * (a[2] + 1);
        if (nmask & (1u << kk)) owner = false;
      }
    }
    if (owner) own |= 1u << k;
  }
  return own;
}

__global__ void expand_count_kernel(const uint64_t* __restrict__ tkeys, const int32_t* __restrict__ tvals,
                                    uint64_t cap_mask, const uint64_t* __restrict__ keys, int64_t n,
                                    int kernel_size, int ts_in, uint32_t* __restrict__ own,
                                    int32_t* __restrict__ counts) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint32_t o;
  if (kernel_size == 3) {
    int b, x, y, z;
    unpack_key(keys[i], b, x, y, z);
    o = owned_candidates(coarse_neighbour_mask(tkeys, tvals, cap_mask, b, x, y, z, ts_in));
  } else if (kernel_size == 2) {
    o = 0xFFu;  // the 8 children of distinct parents are disjoint
  } else {
    o = 1u;
  }
  own[i] = o;
  counts[i] = __popc(o);
}

__global__ void expand_emit_kernel(const uint64_t* __restrict__ keys, int64_t n, int kernel_size,
                                   int ts_out, const uint32_t* __restrict__ own,
                                   const int32_t* __restrict__ pos, uint64_t* __restrict__ out,
                                   int32_t* __restrict__ status) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  int b, x, y, z;
  unpack_key(keys[i], b, x, y, z);
  uint32_t o = own[i];
  int64_t w = pos[i];
  const int lo = kernel_size == 3 ? -1 : 0;
  while (o) {
    const int k = __ffs(o) - 1;
    o &= o - 1;
    const int dx = k % kernel_size + lo, dy = (k / kernel_size) % kernel_size + lo,
              dz = k / (kernel_size * kernel_size) + lo;
    const int qx = x + dx * ts_out, qy = y + dy * ts_out, qz = z + dz * ts_out;
    if (!(coord_in_range(qx) && coord_in_range(qy) && coord_in_range(qz))) atomicOr(status, 1);
    out[w++] = pack_key(b, qx, qy, qz);
  }
}

// ---------------------------------------------------------------- kernel map
// One thread per tile row, looping over the K offsets: nbr writes are coalesced across the warp
// (nbr is offset-major), the warp votes the per-tile tap mask, and the K independent probes per
// thread give the memory system K-deep parallelism.
//
// Tile rows.  The convolution kernels skip, per 128-row tile, every kernel offset no row of the
// tile has a neighbour at.  In lexicographic row order a tile of a voxelised surface has nearly all
// 27 offsets active although a row has only 10-14 neighbours (and a generative transposed
// convolution has 1-8 of 27 by parity class), so most of the tile's multiply-adds would be spent on
// zero rows.  With `order` the tile rows are a permutation of the output rows: rows sorted by
// (row / seg_rows, 27-bit neighbour mask), i.e. rows with the same neighbour pattern share tiles
// while a tile's gathers stay inside one L2-sized segment of the map.  Pass 1 (MASK_ONLY) computes
// the sort keys, a radix sort yields `order`, pass 2 writes nbr / tapmask in tile-row order.
// MASK_ONLY: one thread per output row (natural order) -> its neighbour mask over all K offsets.
// Otherwise: one thread per TILE ROW j of offset group [k_lo, k_hi): output row o = order[j], writes
// nbr[k][j] for the group's offsets and the tile's tap mask (bits k_lo .. k_hi - 1 only).
// ints per row of the mask pass's probe results: a multiple of 4 (16-byte accesses); 32 = one 128-byte line for K = 27
__host__ __device__ constexpr int km_found_pitch(int K) { return K > 16 ? 32 : ((K + 3) & ~3); }

template <int KS, bool MASK_ONLY>
__global__ void __launch_bounds__(PCC_TILE_M)
kernel_map_kernel(const uint64_t* __restrict__ tkeys, const int32_t* __restrict__ tvals,
                  uint64_t cap_mask, const uint64_t* __restrict__ out_keys, int64_t n_out, int step,
                  int in_align, const int32_t* __restrict__ order, int32_t* __restrict__ nbr,
                  int64_t nbr_stride, uint32_t* __restrict__ tapmask, uint32_t* __restrict__ row_masks,
                  int k_lo, int k_hi, int32_t* __restrict__ found) {
  constexpr int K = KS * KS * KS;
  constexpr int LO = (KS == 3) ? -1 : 0;
  __shared__ uint32_t smask;
  if (!MASK_ONLY) {
    if (threadIdx.x == 0) smask = 0;
    __syncthreads();
  }
  const int64_t j = (int64_t)blockIdx.x * PCC_TILE_M + threadIdx.x;  // tile row
  const bool live = j < n_out;
  int b = 0, x = 0, y = 0, z = 0;
  int64_t o_row = 0;
  if (live) {
    o_row = (!MASK_ONLY && order) ? (int64_t)__ldg(order + j) : j;
    unpack_key(out_keys[o_row], b, x, y, z);
  }
  uint32_t mine = 0;
  const int amask = in_align - 1;  // in_align is a power of two
  // write pass after a mask pass: the row's mask says which neighbours exist, so the (longer) probe sequences
  // of the missing ones are not walked a second time
  const uint32_t known = (!MASK_ONLY && row_masks && live) ? __ldg(row_masks + o_row) : 0xffffffffu;
  auto probe = [&](int k) {
    const int dx = k % KS + LO, dy = (k / KS) % KS + LO, dz = k / (KS * KS) + LO;
    const int qx = x + dx * step, qy = y + dy * step, qz = z + dz * step;
    int r = -1;
    if (live && ((known >> k) & 1u) && coord_in_range(qx) && coord_in_range(qy) && coord_in_range(qz) &&
        (((qx | qy | qz) & amask) == 0))
      r = hash_find(tkeys, tvals, cap_mask, pack_key(b, qx, qy, qz));
    if (r >= 0) mine |= 1u << k;
    return r;
  };
  if (MASK_ONLY && found) {
    // probe results kept ROW-major, found[row][KM_FOUND_PITCH], written 16 bytes at a time: the write pass reads a
    // row by its tile position (a random row), and a row's 27 entries are then four whole sectors instead of
    // 27 sectors with 4 useful bytes each
    constexpr int FP = km_found_pitch(K);
#pragma unroll 1
    for (int k4 = 0; k4 < FP; k4 += 4) {
      int r[4];
#pragma unroll
      for (int e = 0; e < 4; ++e) r[e] = (k4 + e < K) ? probe(k4 + e) : -1;
      if (live) *reinterpret_cast<int4*>(found + j * FP + k4) = make_int4(r[0], r[1], r[2], r[3]);
    }
  } else {
#pragma unroll 3
    for (int k = 0; k < K; ++k) {
      if (!MASK_ONLY && (k < k_lo || k >= k_hi)) continue;
      const int r = probe(k);
      if (!MASK_ONLY && live) nbr[(int64_t)k * nbr_stride + j] = r;
    }
  }
  if (MASK_ONLY) {
    if (live) row_masks[j] = mine;
    return;
  }
#pragma unroll
  for (int s = 16; s > 0; s >>= 1) mine |= __shfl_xor_sync(0xffffffffu, mine, s);
  if ((threadIdx.x & 31) == 0 && mine) atomicOr(&smask, mine);
  __syncthreads();
  if (threadIdx.x == 0 && tapmask) tapmask[blockIdx.x] = smask;
}

// Write pass after a mask pass that kept its probe results (found: [n_out][km_found_pitch(K)], output-row order): tile row j of
// the group takes the entries of output row order[j] — a permuted copy, no second walk through the hash table —
// and the warp votes the tile's tap mask as kernel_map_kernel does.
template <int KS>
__global__ void __launch_bounds__(PCC_TILE_M)
kernel_map_permute_kernel(const int32_t* __restrict__ found, int64_t n_out, const int32_t* __restrict__ order,
                          int32_t* __restrict__ nbr, int64_t nbr_stride, uint32_t* __restrict__ tapmask, int k_lo,
                          int k_hi) {
  constexpr int K = KS * KS * KS;
  __shared__ uint32_t smask;
  if (threadIdx.x == 0) smask = 0;
  __syncthreads();
  const int64_t j = (int64_t)blockIdx.x * PCC_TILE_M + threadIdx.x;
  const bool live = j < n_out;
  const int64_t o = live ? (int64_t)__ldg(order + j) : 0;
  uint32_t mine = 0;
  constexpr int FP = km_found_pitch(K);
  int32_t vals[FP];
#pragma unroll
  for (int q = 0; q < FP / 4; ++q) {
    int4 v = make_int4(-1, -1, -1, -1);
    if (live && 4 * q < k_hi && 4 * q + 3 >= k_lo) v = __ldg(reinterpret_cast<const int4*>(found + o * FP) + q);
    vals[4 * q] = v.x; vals[4 * q + 1] = v.y; vals[4 * q + 2] = v.z; vals[4 * q + 3] = v.w;
  }
#pragma unroll
  for (int k = 0; k < K; ++k) {
    if (k < k_lo || k >= k_hi) continue;
    if (live) {
      const int32_t r = vals[k];
      nbr[(int64_t)k * nbr_stride + j] = r;
      if (r >= 0) mine |= 1u << k;
    }
  }
#pragma unroll
  for (int s = 16; s > 0; s >>= 1) mine |= __shfl_xor_sync(0xffffffffu, mine, s);
  if ((threadIdx.x & 31) == 0 && mine) atomicOr(&smask, mine);
  __syncthreads();
  if (threadIdx.x == 0 && tapmask) tapmask[blockIdx.x] = smask;
}

// sort key of row j for one offset group: (j / seg_rows, the group's bits of the row's neighbour mask).
// The sort is by mask VALUE on purpose.  A surface map has thousands of distinct masks, many with fewer rows than a
// tile, so most tiles hold several masks; in value order neighbouring masks share their high offsets and the union
// a tile computes stays small.  Replacing the mask by a 16-bit hash (two radix passes instead of four) was tried in
// round 2: tiles then mix unrelated masks, the useful-MAC share of the 799,458-voxel map fell below the 0.72
// threshold and the map was cut into three offset groups (tests/test_gpu_fullsize.py caught it).
template <typename KeyT>
__global__ void group_key_kernel(const uint32_t* __restrict__ row_masks, int64_t n, int k_lo, int k_hi, int seg_shift,
                                 KeyT* __restrict__ keys, int32_t* __restrict__ iota) {
  const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n) return;
  const uint32_t sub = (row_masks[j] >> k_lo) & (uint32_t)((1ull << (k_hi - k_lo)) - 1ull);
  keys[j] = (KeyT)(((uint64_t)(j >> seg_shift) << (k_hi - k_lo)) | sub);
  iota[j] = (int32_t)j;
}

// ---------------------------------------------------------------- sort helpers
static size_t cub_sort_pairs_bytes(int64_t n) {
  size_t bytes = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, bytes, (const uint64_t*)nullptr, (uint64_t*)nullptr,
                                  (const int32_t*)nullptr, (int32_t*)nullptr, (int)n);
  return bytes;
}
static size_t cub_sort_keys_bytes(int64_t n) {
  size_t bytes = 0;
  cub::DeviceRadixSort::SortKeys(nullptr, bytes, (const uint64_t*)nullptr, (uint64_t*)nullptr, (int)n);
  return bytes;
}

// order_out = rows sorted (stably) by the group key.  Keys of <= 32 bits (every whole-map sort: 27 mask bits; the
// segmented sorts up to 4 M rows) are sorted as 32-bit keys: a third less traffic per radix pass.
static int sort_rows_by_group_key(const uint32_t* row_masks, int64_t n, int k_lo, int k_hi, int seg_shift, int seg_bits,
                                  void* k_in, void* k_out, int32_t* iota, int32_t* order_out, void* cub_ws,
                                  size_t cub_bytes, cudaStream_t st) {
  const int bits = (k_hi - k_lo) + seg_bits;
  const unsigned grid = (unsigned)ceil_div(n, 256);
  if (bits <= 32) {
    size_t need = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, need, (const uint32_t*)nullptr, (uint32_t*)nullptr, (const int32_t*)nullptr,
                                    (int32_t*)nullptr, (int)n);
    if (need <= cub_bytes) {
      group_key_kernel<uint32_t><<<grid, 256, 0, st>>>(row_masks, n, k_lo, k_hi, seg_shift,
                                                       reinterpret_cast<uint32_t*>(k_in), iota);
      PCC_CHECK_LAUNCH("group_key_kernel");
      PCC_CHECK_CUDA(cub::DeviceRadixSort::SortPairs(cub_ws, cub_bytes, reinterpret_cast<const uint32_t*>(k_in),
                                                     reinterpret_cast<uint32_t*>(k_out), iota, order_out, (int)n, 0,
                                                     bits, st));
      return PCC_OK;
    }
  }
  group_key_kernel<uint64_t><<<grid, 256, 0, st>>>(row_masks, n, k_lo, k_hi, seg_shift,
                                                   reinterpret_cast<uint64_t*>(k_in), iota);
  PCC_CHECK_LAUNCH("group_key_kernel");
  PCC_CHECK_CUDA(cub::DeviceRadixSort::SortPairs(cub_ws, cub_bytes, reinterpret_cast<const uint64_t*>(k_in),
                                                 reinterpret_cast<uint64_t*>(k_out), iota, order_out, (int)n, 0, bits, st));
  return PCC_OK;
}

static int sort_keys_only(const uint64_t* in, uint64_t* out, int64_t n, void* ws, size_t ws_bytes,
                          cudaStream_t st) {
  if (n == 0) return PCC_OK;
  size_t need = ws_bytes;
  PCC_CHECK_CUDA(cub::DeviceRadixSort::SortKeys(ws, need, in, out, (int)n, 0, 64, st));
  return PCC_OK;
}

}  // namespace pcc

using namespace pcc;

extern "C" {

int pcc_coords_pack(const int32_t* coords, int64_t n, uint64_t* keys, int32_t* status, void* stream) {
  PCC_REQUIRE(n >= 0 && (n == 0 || (coords && keys)), "null pointer");
  if (n == 0) return PCC_OK;
  pack_kernel<<<grid_for(n), THREADS, 0, (cudaStream_t)stream>>>((const int4*)coords, n, keys, status);
  PCC_CHECK_LAUNCH("pack_kernel");
  return PCC_OK;
}

int pcc_coords_unpack(const uint64_t* keys, int64_t n, int32_t* coords, void* stream) {
  PCC_REQUIRE(n >= 0 && (n == 0 || (coords && keys)), "null pointer");
  if (n == 0) return PCC_OK;
  unpack_kernel<<<grid_for(n), THREADS, 0, (cudaStream_t)stream>>>(keys, n, (int4*)coords);
  PCC_CHECK_LAUNCH("unpack_kernel");
  return PCC_OK;
}

size_t pcc_sort_ws_bytes(int64_t n) {
  if (n <= 0) return 256;
  return align_up(cub_sort_pairs_bytes(n), 256) + align_up((size_t)n * 4, 256);
}

int pcc_sort_keys(const uint64_t* keys_in, int64_t n, uint64_t* keys_out, int32_t* perm_out, void* ws,
                  size_t ws_bytes, void* stream) {
  PCC_REQUIRE(n >= 0 && n < (1ll << 31), "n out of range");
  if (n == 0) return PCC_OK;
  PCC_REQUIRE(keys_in && keys_out && ws, "null pointer");
  PCC_REQUIRE(ws_bytes >= pcc_sort_ws_bytes(n), "workspace too small");
  cudaStream_t st = (cudaStream_t)stream;
  if (!perm_out) return sort_keys_only(keys_in, keys_out, n, ws, ws_bytes, st);
  size_t cub_bytes = align_up(cub_sort_pairs_bytes(n), 256);
  int32_t* iota = reinterpret_cast<int32_t*>(reinterpret_cast<char*>(ws) + cub_bytes);
  iota_kernel<<<grid_for(n), THREADS, 0, st>>>(iota, n);
  PCC_CHECK_LAUNCH("iota_kernel");
  PCC_CHECK_CUDA(cub::DeviceRadixSort::SortPairs(ws, cub_bytes, keys_in, keys_out, iota, perm_out,
                                                 (int)n, 0, 64, st));
  return PCC_OK;
}

int pcc_take_keys(const uint64_t* keys, const int32_t* rows, int64_t n, uint64_t* out, void* stream) {
  if (n == 0) return PCC_OK;
  PCC_REQUIRE(keys && rows && out, "null pointer");
  take_keys_kernel<<<grid_for(n), THREADS, 0, (cudaStream_t)stream>>>(keys, rows, n, out);
  PCC_CHECK_LAUNCH("take_keys_kernel");
  return PCC_OK;
}

int64_t pcc_hash_capacity(int64_t n) {
  int64_t cap = 1024;
  while (cap < 2 * n) cap <<= 1;
  return cap;
}

int pcc_hash_build(const uint64_t* keys, int64_t n, uint64_t* tkeys, int32_t* tvals, int64_t cap,
                   void* stream) {
  PCC_REQUIRE(tkeys && tvals, "null table");
  PCC_REQUIRE(cap >= 1024 && (cap & (cap - 1)) == 0 && cap >= 2 * n, "capacity must be a power of two >= 2n");
  cudaStream_t st = (cudaStream_t)stream;
  hash_fill_kernel<<<grid_for(cap), THREADS, 0, st>>>(tkeys, cap);
  PCC_CHECK_LAUNCH("hash_fill_kernel");
  if (n > 0) {
    PCC_REQUIRE(keys, "null keys");
    hash_insert_kernel<<<grid_for(n), THREADS, 0, st>>>(keys, n, tkeys, tvals, (uint64_t)cap - 1);
    PCC_CHECK_LAUNCH("hash_insert_kernel");
  }
  return PCC_OK;
}

int pcc_hash_probe(const uint64_t* tkeys, const int32_t* tvals, int64_t cap, const uint64_t* q_keys,
                   int64_t nq, int32_t* rows, void* stream) {
  if (nq == 0) return PCC_OK;
  PCC_REQUIRE(tkeys && tvals && q_keys && rows, "null pointer");
  PCC_REQUIRE((cap & (cap - 1)) == 0, "capacity must be a power of two");
  hash_probe_kernel<<<grid_for(nq), THREADS, 0, (cudaStream_t)stream>>>(tkeys, tvals, (uint64_t)cap - 1,
                                                                        q_keys, nq, rows);
  PCC_CHECK_LAUNCH("hash_probe_kernel");
  return PCC_OK;
}

// ws layout for stride / expand:  [A: raw keys][B: flags/pos int32][scan ws][cub ws]
static size_t unique_ws_bytes(int64_t m) {
  return align_up((size_t)m * 8, 256) * 2 + align_up((size_t)m * 4, 256) + scan_ws_bytes(m) +
         align_up(cub_sort_keys_bytes(m), 256) + 256;
}

// Page-locked per-thread staging for the counts a function hands back to the host: a D2H copy into
// pageable memory is staged by the driver, into pinned memory it is a plain DMA + one stream sync.
static int32_t* pinned_counts() {
  static thread_local int32_t* p = [] {
    int32_t* q = nullptr;
    if (cudaHostAlloc(reinterpret_cast<void**>(&q), 64, cudaHostAllocDefault) != cudaSuccess) q = nullptr;
    return q;
  }();
  return p;
}

// sorts raw[m] and writes the unique keys to out_keys; returns count through the host pointer
static int sort_unique(uint64_t* raw, int64_t m, uint64_t* out_keys, int64_t* n_out, char* ws,
                       size_t ws_bytes, cudaStream_t st) {
  // raw lives at ws offset 0 (caller filled it)
  size_t off = align_up((size_t)m * 8, 256);
  uint64_t* sorted = reinterpret_cast<uint64_t*>(ws + off);
  off += align_up((size_t)m * 8, 256);
  int32_t* flags = reinterpret_cast<int32_t*>(ws + off);
  off += align_up((size_t)m * 4, 256);
  void* scan_ws = ws + off;
  off += scan_ws_bytes(m);
  int32_t* total = reinterpret_cast<int32_t*>(ws + off);
  off += 256;
  void* cub_ws = ws + off;
  size_t cub_bytes = ws_bytes - off;
  int rc = sort_keys_only(raw, sorted, m, cub_ws, cub_bytes, st);
  if (rc) return rc;
  head_flags_kernel<<<grid_for(m), THREADS, 0, st>>>(sorted, m, flags);
  PCC_CHECK_LAUNCH("head_flags_kernel");
  rc = exclusive_scan_i32(flags, flags, m, total, scan_ws, st);
  if (rc) return rc;
  compact_heads_kernel<<<grid_for(m), THREADS, 0, st>>>(sorted, m, flags, out_keys);
  PCC_CHECK_LAUNCH("compact_heads_kernel");
  int32_t hbuf = 0;
  int32_t* h = pinned_counts() ? pinned_counts() : &hbuf;
  PCC_CHECK_CUDA(cudaMemcpyAsync(h, total, 4, cudaMemcpyDeviceToHost, st));
  PCC_CHECK_CUDA(cudaStreamSynchronize(st));
  *n_out = h[0];
  return PCC_OK;
}

size_t pcc_stride_ws_bytes(int64_t n) { return unique_ws_bytes(n > 0 ? n : 1); }

int pcc_coords_stride(const uint64_t* keys, int64_t n, int32_t ts_out, uint64_t* out_keys, int64_t* n_out,
                      void* ws, size_t ws_bytes, void* stream) {
  PCC_REQUIRE(n_out, "null n_out");
  *n_out = 0;
  if (n == 0) return PCC_OK;
  PCC_REQUIRE(keys && out_keys && ws && ts_out > 0, "bad argument");
  PCC_REQUIRE(ws_bytes >= pcc_stride_ws_bytes(n), "workspace too small");
  cudaStream_t st = (cudaStream_t)stream;
  uint64_t* raw = reinterpret_cast<uint64_t*>(ws);
  stride_keys_kernel<<<grid_for(n), THREADS, 0, st>>>(keys, n, ts_out, raw);
  PCC_CHECK_LAUNCH("stride_keys_kernel");
  return sort_unique(raw, n, out_keys, n_out, reinterpret_cast<char*>(ws), ws_bytes, st);
}

size_t pcc_expand_ws_bytes(int64_t n, int kernel_size) {
  if (n <= 0) n = 1;
  const int64_t m = n * kernel_size * kernel_size * kernel_size;
  // own[n] u32 | counts/pos[n] i32 | scan ws | total | raw[m] u64 | cub
  return align_up((size_t)n * 4, 256) * 2 + scan_ws_bytes(n) + 512 + align_up((size_t)m * 8, 256) +
         align_up(cub_sort_keys_bytes(m), 256);
}

int pcc_coords_expand(const uint64_t* tkeys, const int32_t* tvals, int64_t cap, const uint64_t* keys,
                      int64_t n, int kernel_size, int32_t ts_in, uint64_t* out_keys, int64_t* n_out,
                      void* ws, size_t ws_bytes, void* stream) {
  PCC_REQUIRE(n_out, "null n_out");
  *n_out = 0;
  if (n == 0) return PCC_OK;
  PCC_REQUIRE(kernel_size >= 1 && kernel_size <= 3, "kernel_size must be 1, 2 or 3");
  PCC_REQUIRE(keys && out_keys && ws && tkeys && tvals, "null pointer");
  PCC_REQUIRE(ts_in >= 2 && (ts_in % 2) == 0, "input tensor stride must be even");
  PCC_REQUIRE(ws_bytes >= pcc_expand_ws_bytes(n, kernel_size), "workspace too small");
  cudaStream_t st = (cudaStream_t)stream;
  char* w = reinterpret_cast<char*>(ws);
  size_t off = 0;
  uint32_t* own = reinterpret_cast<uint32_t*>(w + off);
  off += align_up((size_t)n * 4, 256);
  int32_t* pos = reinterpret_cast<int32_t*>(w + off);
  off += align_up((size_t)n * 4, 256);
  void* scan_ws = w + off;
  off += scan_ws_bytes(n);
  int32_t* total = reinterpret_cast<int32_t*>(w + off);
  int32_t* status = total + 1;
  off += 512;
  uint64_t* raw = reinterpret_cast<uint64_t*>(w + off);
  PCC_CHECK_CUDA(cudaMemsetAsync(total, 0, 8, st));
  expand_count_kernel<<<grid_for(n), THREADS, 0, st>>>(tkeys, tvals, (uint64_t)cap - 1, keys, n,
                                                       kernel_size, ts_in, own, pos);
  PCC_CHECK_LAUNCH("expand_count_kernel");
  int rc = exclusive_scan_i32(pos, pos, n, total, scan_ws, st);
  if (rc) return rc;
  expand_emit_kernel<<<grid_for(n), THREADS, 0, st>>>(keys, n, kernel_size, ts_in / 2, own, pos, raw, status);
  PCC_CHECK_LAUNCH("expand_emit_kernel");
  int32_t hbuf[2] = {0, 0};
  int32_t* h = pinned_counts() ? pinned_counts() : hbuf;
  PCC_CHECK_CUDA(cudaMemcpyAsync(h, total, 8, cudaMemcpyDeviceToHost, st));
  PCC_CHECK_CUDA(cudaStreamSynchronize(st));
  if (h[1]) {
    set_error("pcc_coords_expand: generated coordinate outside the packable range");
    return PCC_ERR_CAPACITY;
  }
  const int64_t m = h[0];
  off += align_up((size_t)(n * kernel_size * kernel_size * kernel_size) * 8, 256);
  rc = sort_keys_only(raw, out_keys, m, w + off, ws_bytes - off, st);
  if (rc) return rc;
  *n_out = m;
  return PCC_OK;
}

size_t pcc_kernel_map_ws_bytes(int64_t n_out) {
  if (n_out <= 0) return 256;
  // sort keys in | sort keys out | iota | row masks | cub | two counters | the mask pass's probe results [n_out][32]
  return align_up((size_t)n_out * 8, 256) * 2 + align_up((size_t)n_out * 4, 256) * 2 +
         align_up(cub_sort_pairs_bytes(n_out), 256) + 256 + align_up((size_t)n_out * 32 * 4, 256);
}

}  // extern "C"

// useful[0] += pairs of the tile, useful[1] += 128 * offsets the tile would compute (tile rows = order)
__global__ void __launch_bounds__(PCC_TILE_M)
tile_efficiency_kernel(const uint32_t* __restrict__ row_masks, const int32_t* __restrict__ order, int64_t n,
                       unsigned long long* __restrict__ sums) {
  __shared__ uint32_t s_or;
  __shared__ int s_pairs;
  if (threadIdx.x == 0) {
    s_or = 0;
    s_pairs = 0;
  }
  __syncthreads();
  const int64_t j = (int64_t)blockIdx.x * PCC_TILE_M + threadIdx.x;
  uint32_t m = j < n ? row_masks[order[j]] : 0u;
  int pc = __popc(m);
#pragma unroll
  for (int s = 16; s > 0; s >>= 1) {
    m |= __shfl_xor_sync(0xffffffffu, m, s);
    pc += __shfl_xor_sync(0xffffffffu, pc, s);
  }
  if ((threadIdx.x & 31) == 0) {
    atomicOr(&s_or, m);
    atomicAdd(&s_pairs, pc);
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    atomicAdd(sums, (unsigned long long)s_pairs);
    atomicAdd(sums + 1, (unsigned long long)(PCC_TILE_M * __popc(s_or)));
  }
}

// Launch order of the tiles of one offset group: most active offsets first (the block scheduler hands
// out CTAs in index order, so the long tiles start early and the short ones fill the tail).  One CTA,
// counting sort on popc(tapmask); the order inside a bin is arbitrary (it only affects scheduling).
__global__ void __launch_bounds__(1024)
tile_rank_kernel(const uint32_t* __restrict__ tapmask, int tiles, int32_t* __restrict__ tile_order) {
  __shared__ int hist[33], base[33], cursor[33];
  if (threadIdx.x < 33) hist[threadIdx.x] = cursor[threadIdx.x] = 0;
  __syncthreads();
  for (int t = threadIdx.x; t < tiles; t += blockDim.x) atomicAdd(&hist[__popc(tapmask[t])], 1);
  __syncthreads();
  if (threadIdx.x == 0) {
    int run = 0;
    for (int b = 32; b >= 0; --b) {
      base[b] = run;
      run += hist[b];
    }
  }
  __syncthreads();
  for (int t = threadIdx.x; t < tiles; t += blockDim.x) {
    const int b = __popc(tapmask[t]);
    tile_order[base[b] + atomicAdd(&cursor[b], 1)] = t;
  }
}

// n_groups == 0 (automatic).  Measured with the TMA kernel on the surface levels of the 799,458-voxel frame
// (tools/ragged_conv_bench.py, profiles/r02_ragged_groups.txt; 128 -> 128, useful-MAC share of the mask-sorted
// tiles with 1 / 2 / 3 groups in brackets):
//   249,574 rows (0.858 / 0.957 / 0.986): 2.376 / 2.221 / 2.250 ms      68,423 rows (0.756 / 0.889 / 0.959):
//   0.862 / 0.768 / 0.768 ms      17,957 rows (0.664 / 0.772 / 0.868): 0.347 / 0.361 / 0.372 ms
// i.e. every extra chained pass costs ~4 % of the padded-MAC rate, and a map that cannot fill the GPU twice gains
// nothing from shorter tiles.  Rule: one group when the single-group tiles are >= 90 % useful or the map has
// fewer than 2 x SM-count tiles; two groups down to 70 %; three below.
constexpr double AUTO_GROUP_ONE_MIN_EFFICIENCY = 0.90;
constexpr double AUTO_GROUP_TWO_MIN_EFFICIENCY = 0.70;
constexpr int AUTO_GROUPS = 3;

template <int KS>
static int kernel_map_launch(const uint64_t* tkeys, const int32_t* tvals, uint64_t cm, const uint64_t* out_keys,
                             int64_t n_out, int32_t step, int32_t in_align, int32_t* nbr, int64_t nbr_stride,
                             uint32_t* tapmask, int32_t* order, int32_t* tile_order, int64_t seg_rows,
                             int32_t n_groups, int32_t* n_groups_used, void* ws, size_t ws_bytes, cudaStream_t st) {
  constexpr int K = KS * KS * KS;
  if (n_groups_used) *n_groups_used = n_groups > 0 ? n_groups : 1;
  const unsigned grid = (unsigned)ceil_div(n_out, PCC_TILE_M);
  if (!order) {
    kernel_map_kernel<KS, false><<<grid, PCC_TILE_M, 0, st>>>(tkeys, tvals, cm, out_keys, n_out, step, in_align,
                                                             nullptr, nbr, nbr_stride, tapmask, nullptr, 0, K, nullptr);
    PCC_CHECK_LAUNCH("kernel_map_kernel");
    if (tile_order && tapmask) {
      tile_rank_kernel<<<1, 1024, 0, st>>>(tapmask, (int)grid, tile_order);
      PCC_CHECK_LAUNCH("tile_rank_kernel");
    }
    return PCC_OK;
  }
  int seg_shift = 0;
  while (((int64_t)1 << seg_shift) < seg_rows) ++seg_shift;
  int seg_bits = 0;
  while ((((n_out - 1) >> seg_shift) >> seg_bits) > 0) ++seg_bits;
  char* w = reinterpret_cast<char*>(ws);
  uint64_t* k_in = reinterpret_cast<uint64_t*>(w);
  uint64_t* k_out = reinterpret_cast<uint64_t*>(w + align_up((size_t)n_out * 8, 256));
  int32_t* iota = reinterpret_cast<int32_t*>(w + 2 * align_up((size_t)n_out * 8, 256));
  uint32_t* row_masks = reinterpret_cast<uint32_t*>(w + 2 * align_up((size_t)n_out * 8, 256) + align_up((size_t)n_out * 4, 256));
  void* cub_ws = w + 2 * align_up((size_t)n_out * 8, 256) + 2 * align_up((size_t)n_out * 4, 256);
  size_t cub_bytes = align_up(cub_sort_pairs_bytes(n_out), 256);
  // every neighbour is probed ONCE: the mask pass keeps what it found, the write passes permute it
  int32_t* found = reinterpret_cast<int32_t*>(reinterpret_cast<char*>(cub_ws) + cub_bytes + 256);
  kernel_map_kernel<KS, true><<<grid, PCC_TILE_M, 0, st>>>(tkeys, tvals, cm, out_keys, n_out, step, in_align,
                                                          nullptr, nullptr, 0, nullptr, row_masks, 0, K, found);
  PCC_CHECK_LAUNCH("kernel_map_kernel<mask>");
  bool have_order0 = false;
  if (n_groups == 0) {
    // sort by the whole mask, measure how well such tiles would be used, decide
    unsigned long long* sums = reinterpret_cast<unsigned long long*>(reinterpret_cast<char*>(cub_ws) + cub_bytes);
    PCC_CHECK_CUDA(cudaMemsetAsync(sums, 0, 16, st));
    {
      const int rc = sort_rows_by_group_key(row_masks, n_out, 0, K, seg_shift, seg_bits, k_in, k_out, iota, order,
                                            cub_ws, cub_bytes, st);
      if (rc) return rc;
    }
    tile_efficiency_kernel<<<grid, PCC_TILE_M, 0, st>>>(row_masks, order, n_out, sums);
    PCC_CHECK_LAUNCH("tile_efficiency_kernel");
    unsigned long long hbuf[2] = {0, 0};
    unsigned long long* h = pinned_counts() ? reinterpret_cast<unsigned long long*>(pinned_counts()) : hbuf;
    PCC_CHECK_CUDA(cudaMemcpyAsync(h, sums, 16, cudaMemcpyDeviceToHost, st));
    PCC_CHECK_CUDA(cudaStreamSynchronize(st));
    const double useful = (double)h[0], padded = (double)h[1];
    const bool single = K < AUTO_GROUPS || (int64_t)grid < 2 * (int64_t)sm_count() ||
                        useful >= AUTO_GROUP_ONE_MIN_EFFICIENCY * padded;
    n_groups = single ? 1 : (useful >= AUTO_GROUP_TWO_MIN_EFFICIENCY * padded ? 2 : AUTO_GROUPS);
    have_order0 = single;
    if (n_groups_used) *n_groups_used = n_groups;
  }
  const int per = (K + n_groups - 1) / n_groups;
  for (int g = 0; g < n_groups; ++g) {
    const int k_lo = g * per, k_hi = (g + 1) * per < K ? (g + 1) * per : K;
    int32_t* order_g = order + (int64_t)g * n_out;
    if (!have_order0) {
      const int rc = sort_rows_by_group_key(row_masks, n_out, k_lo, k_hi, seg_shift, seg_bits, k_in, k_out, iota,
                                            order_g, cub_ws, cub_bytes, st);
      if (rc) return rc;
    }
    kernel_map_permute_kernel<KS><<<grid, PCC_TILE_M, 0, st>>>(found, n_out, order_g, nbr, nbr_stride,
                                                              tapmask ? tapmask + (int64_t)g * grid : nullptr, k_lo, k_hi);
    PCC_CHECK_LAUNCH("kernel_map_permute_kernel");
    if (tile_order && tapmask) {
      tile_rank_kernel<<<1, 1024, 0, st>>>(tapmask + (int64_t)g * grid, (int)grid, tile_order + (int64_t)g * grid);
      PCC_CHECK_LAUNCH("tile_rank_kernel");
    }
  }
  return PCC_OK;
}

extern "C" {

int pcc_kernel_map(const uint64_t* tkeys, const int32_t* tvals, int64_t cap, const uint64_t* out_keys,
                   int64_t n_out, int kernel_size, int32_t step, int32_t in_align, int32_t* nbr,
                   int64_t nbr_stride, uint32_t* tapmask, int32_t* order, int32_t* tile_order, int64_t seg_rows,
                   int32_t n_groups, int32_t* n_groups_used, void* ws, size_t ws_bytes, void* stream) {
  if (n_groups_used) *n_groups_used = n_groups > 0 ? n_groups : 1;
  if (n_out == 0) return PCC_OK;
  PCC_REQUIRE(tkeys && tvals && out_keys && nbr, "null pointer");
  PCC_REQUIRE(n_out < (1ll << 31), "n_out out of range");
  PCC_REQUIRE(nbr_stride >= n_out, "nbr_stride < n_out");
  PCC_REQUIRE(in_align > 0 && (in_align & (in_align - 1)) == 0, "in_align must be a power of two");
  PCC_REQUIRE((cap & (cap - 1)) == 0, "capacity must be a power of two");
  PCC_REQUIRE(n_groups >= 0 && n_groups <= kernel_size * kernel_size * kernel_size, "n_groups out of range");
  PCC_REQUIRE(n_groups == 1 || order, "n_groups != 1 needs an order array");
  PCC_REQUIRE(n_groups != 0 || n_groups_used, "automatic n_groups needs n_groups_used");
  PCC_REQUIRE(!tile_order || tapmask, "tile_order needs tapmask");
  if (order) {
    PCC_REQUIRE(seg_rows >= PCC_TILE_M && (seg_rows & (seg_rows - 1)) == 0, "seg_rows must be a power of two >= 128");
    PCC_REQUIRE(ws && ws_bytes >= pcc_kernel_map_ws_bytes(n_out), "workspace too small");
  }
  cudaStream_t st = (cudaStream_t)stream;
  const uint64_t cm = (uint64_t)cap - 1;
  switch (kernel_size) {
    case 1:
      return kernel_map_launch<1>(tkeys, tvals, cm, out_keys, n_out, step, in_align, nbr, nbr_stride, tapmask, order, tile_order, seg_rows, n_groups, n_groups_used, ws, ws_bytes, st);
    case 2:
      return kernel_map_launch<2>(tkeys, tvals, cm, out_keys, n_out, step, in_align, nbr, nbr_stride, tapmask, order, tile_order, seg_rows, n_groups, n_groups_used, ws, ws_bytes, st);
    case 3:
      return kernel_map_launch<3>(tkeys, tvals, cm, out_keys, n_out, step, in_align, nbr, nbr_stride, tapmask, order, tile_order, seg_rows, n_groups, n_groups_used, ws, ws_bytes, st);
    default:
      PCC_REQUIRE(false, "kernel_size must be 1, 2 or 3");
  }
  return PCC_OK;
}

}  // extern "C"
