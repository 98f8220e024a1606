// Weight gradient of the sparse convolution on the tensor cores (PCC_MODE_TF32):
//   dW[k][ci][co] = sum_j in[nbr[k][j]][ci] * gout[order[j]][co]
// is, per kernel offset k, a (cin x rows) x (rows x cout) product whose CONTRACTION dimension is the
// row index.  Both operands are gathered rows, i.e. they arrive in shared memory with the M / N
// dimension (channels) contiguous: MN-major operands.  tcgen05.mma kind::tf32 takes them as they are
// (instruction-descriptor bits 15 / 16), so no transpose pass is needed:
//   * a stage = 32 gathered rows of `in` (A: 128 ci as four [32 rows x 128 B] swizzled slabs) and of
//     `gout` (B: cout / 32 slabs); 8 producer warps fill it with zero-filling cp.async and publish it
//     with cp.async.mbarrier.arrive.noinc (no thread waits for its copies);
//   * one elected lane issues 4 x tcgen05.mma (M = 128, N = cout tile, K = 8 rows) per stage; the fp32
//     accumulator [128 ci x cout] stays in TMEM across ALL row tiles of the CTA;
//   * layers with cin = 64 / 32 put 2 / 4 kernel offsets side by side in the M = 128 rows (each offset gathers its
//     own `in` rows into its slabs; the gout rows are shared), so narrow layers use the whole MMA and gather B once;
//   * the kernel map's tap masks skip row tiles without offset k; the row range is cut into S splits
//     whose partial tiles are added in a fixed order by wgrad_reduce (deterministic).
// MN-major operands of a 32-bit type have ONE legal shared-memory layout, SWIZZLE_128B_BASE32B (descriptor
// layout type 1; CUTLASS sm100_common.inl: "for mn-major tf32 operands, SW128_32B is the only available smem
// layout" — with plain SWIZZLE_128B the MMA silently produces zeros, measured): atoms of 4 rows x 128
// contiguous bytes of channels, rows 128 B apart, the 32-BYTE chunk index XORed with row % 4
// (Swizzle<2,5,2> on the byte address); LBO = distance between 32-channel slabs, SBO = 512 B between
// 4-row atoms.  One K = 8 MMA reads two atoms.
#include "pcc_common.cuh"

#include <algorithm>

namespace pcc {
namespace wgtc {

constexpr int ROWS = 32;                 // rows per stage = 4 MMAs of K = 8
constexpr int SLAB = ROWS * 128;         // bytes of one 32-channel slab of a stage
constexpr int PWARPS = 8;
constexpr int PRODUCERS = 32 * PWARPS;
constexpr int THREADS = PRODUCERS + 32;

struct Params {
  const float* in;
  const float* gout;
  const int32_t* nbr;
  const uint32_t* tapmask;
  const int32_t* order;
  float* part;  // [S][K][cin][cout]
  int64_t nbr_stride, n_out;
  int cin, cout, K, k_lo, k_hi, S;
};

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void cp16(uint32_t dst, const void* src, int bytes) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(dst), "l"(src), "r"(bytes) : "memory");
}
__device__ __forceinline__ void cp_arrive_noinc(uint32_t bar) {
  asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];\n" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void cp_commit_wait_all() {
  asm volatile("cp.async.commit_group;\n cp.async.wait_group 0;\n" ::: "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory"); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n"
      " .reg .pred p;\n"
      "WG_WAIT:\n"
      " mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      " @p bra WG_DONE;\n"
      " bra WG_WAIT;\n"
      "WG_DONE:\n"
      "}\n" ::"r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;\n" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory"); }
__device__ __forceinline__ void tmem_alloc(uint32_t dst_smem, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;\n" ::"r"(dst_smem), "r"(ncols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;\n" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;\n" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void umma_tf32(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n"
      " .reg .pred p;\n"
      " setp.ne.b32 p, %4, 0;\n"
      " tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n"
      "}\n" ::"r"(tmem_d), "l"(da), "l"(db), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];\n"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
        "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
        "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
        "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr)
      : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;\n" ::: "memory");
}

// MN-major descriptor: start >> 4 at [0,14), LBO >> 4 at [16,30) (between 32-channel slabs), SBO >> 4 at
// [32,46) (between 4-row atoms), version 1 at [46,48), layout type at [61,64)
__device__ __forceinline__ uint64_t mn_desc(uint32_t addr, uint32_t lbo = SLAB, uint32_t sbo = 512u, uint32_t layout = 1u) {
  uint64_t d = 0;
  d |= (uint64_t)((addr & 0x3FFFFu) >> 4);
  d |= (uint64_t)(lbo >> 4) << 16;
  d |= (uint64_t)(sbo >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)layout << 61;  // 1 = SWIZZLE_128B_BASE32B: the only MN-major layout of 32-bit operands
  return d;
}
// kind::tf32, D = f32, A = B = tf32, both MN-major (bits 15, 16), N >> 3 at [17,23), M >> 4 at [24,29)
__host__ __device__ constexpr uint32_t idesc_mn(int m, int n) {
  return (1u << 4) | (2u << 7) | (2u << 10) | (1u << 15) | (1u << 16) | ((uint32_t)(n >> 3) << 17) |
         ((uint32_t)(m >> 4) << 24);
}
// byte offset of (row, 16-byte unit) inside a [rows x 128 B] slab in the SWIZZLE_128B_BASE32B pattern
// (Swizzle<2,5,2> on the byte address: the 32-byte chunk index is XORed with row % 4)
__device__ __forceinline__ uint32_t sw128(int row, int unit) {
  return (uint32_t)(row * 128 + ((((unit >> 1) ^ (row & 3)) << 5) | ((unit & 1) << 4)));
}

// index loads that stay where they are written (issued a whole row tile ahead of their first use)
__device__ __forceinline__ int32_t ld_idx(const int32_t* ptr) {
  int32_t v;
  asm volatile("ld.global.nc.s32 %0, [%1];\n" : "=r"(v) : "l"(ptr));
  return v;
}

// The row tiles of this CTA (blockIdx.x, + S, ...) that have one of the CTA's kernel offsets (kmask).  A warp reads the tap masks of
// its next 32 candidate tiles at once and walks the ballot, so a skipped tile costs no memory latency and an
// active one 1/32 of one.  Every lane of the warp sees the same sequence.
struct TileWalk {
  const uint32_t* tapmask;
  int64_t tiles, next_base, cur_base;
  int S, lane;
  uint32_t kmask, m;
  __device__ TileWalk(const uint32_t* tm, int64_t tiles_, int first, int S_, uint32_t kmask_, int lane_)
      : tapmask(tm), tiles(tiles_), next_base(first), cur_base(first), S(S_), lane(lane_), kmask(kmask_), m(0u) {}
  __device__ __forceinline__ int64_t next() {
    while (m == 0u) {
      if (next_base >= tiles) return -1;
      cur_base = next_base;
      const int64_t t = cur_base + (int64_t)lane * S;
      const bool act = t < tiles && (!tapmask || (__ldg(tapmask + t) & kmask) != 0u);
      m = __ballot_sync(0xffffffffu, act);
      next_base = cur_base + (int64_t)32 * S;
    }
    const int j = __ffs(m) - 1;
    m &= m - 1u;
    return cur_base + (int64_t)j * S;
  }
};

template <int NT>
struct Smem {
  static constexpr int STAGES = NT <= 64 ? 4 : 3;
  static constexpr int A_BYTES = 4 * SLAB;
  static constexpr int B_BYTES = (NT / 32) * SLAB;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int BAR_OFFSET = STAGES * STAGE_BYTES;
  static constexpr int TOTAL = BAR_OFFSET + 256 + 1024;  // barriers + alignment slack
};

// grid (S, offsets of this launch / NS, ci tiles); NT = cout rounded up to 64 / 128 / 256.
// NS > 1 (cin = 128 / NS): the M = 128 rows of the MMA hold NS kernel offsets side by side — offset k0 + j owns
// A slabs [j * cin / 32, (j + 1) * cin / 32) and accumulator lanes [j * cin, (j + 1) * cin) — so a narrow layer
// fills the tensor core's M dimension, and the gathered gout rows (B) are shared by the NS offsets.
template <int NT, int NS>
__global__ void __launch_bounds__(THREADS, NT <= 128 ? 2 : 1)
wgrad_tf32_kernel(const Params p) {
  extern __shared__ unsigned char smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  unsigned char* sbase = smem_raw + (base - raw);
  using S = Smem<NT>;
  constexpr int NST = S::STAGES;
  const uint32_t bar_base = base + S::BAR_OFFSET;
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (NST + s); };
  const uint32_t accum_bar = bar_base + 8u * (2 * NST);
  const uint32_t tmem_slot = bar_base + 8u * (2 * NST + 1);
  volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(sbase + S::BAR_OFFSET + 8 * (2 * NST + 1));
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int k0 = p.k_lo + blockIdx.y * NS;               // first kernel offset of this CTA
  const int nk = min(NS, p.k_hi - k0);                    // live offsets
  const uint32_t kmask = ((1u << nk) - 1u) << k0;
  const int ci0 = NS > 1 ? 0 : blockIdx.z * 128;
  const int spo = NS > 1 ? p.cin / 32 : min(4, (p.cin - ci0) / 32);  // live 32-channel slabs of A per offset
  const int nb = p.cout / 32;                 // live slabs of B
  const int64_t tiles = (p.n_out + PCC_TILE_M - 1) / PCC_TILE_M;

  // slabs beyond the layer's channels are never written by the producers: they must read as zeros
  for (int e = tid; e < NST * S::STAGE_BYTES / 16; e += THREADS)
    reinterpret_cast<float4*>(sbase)[e] = make_float4(0.f, 0.f, 0.f, 0.f);
  if (tid == 0) {
    for (int s = 0; s < NST; ++s) {
      mbar_init(full_bar(s), PRODUCERS);
      mbar_init(empty_bar(s), 1);
    }
    mbar_init(accum_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
  }
  if (warp == PWARPS) tmem_alloc(tmem_slot, NT);
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_acc = *tmem_slot_ptr;

  int issued = 0;  // stages this CTA goes through (identical in every thread)
  if (warp < PWARPS) {
    // ============================ producers ============================
    const int r = tid >> 3, u0 = tid & 7;
    const uint32_t off = sw128(r, u0);
    const char* in_b = reinterpret_cast<const char*>(p.in) + (int64_t)(ci0 + u0 * 4) * 4;
    const char* go_b = reinterpret_cast<const char*>(p.gout) + (int64_t)(u0 * 4) * 4;
    const int64_t in_row = (int64_t)p.cin * 4, go_row = (int64_t)p.cout * 4;
    int it = 0;
    TileWalk walk(p.tapmask, tiles, (int)blockIdx.x, p.S, kmask, lane);
    // this thread's row of each of a tile's four stages; the indices of the NEXT tile are requested before the
    // copies of the current one are issued, so the index latency is off the tile-to-tile chain
    int32_t ri[4][NS], ro[4], ri_n[4][NS], ro_n[4];
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      ro[c] = ro_n[c] = 0;
#pragma unroll
      for (int j = 0; j < NS; ++j) ri[c][j] = ri_n[c][j] = -1;
    }
    auto request = [&](int64_t tile, int32_t (&a)[4][NS], int32_t (&b)[4]) {
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        const int64_t t = tile * PCC_TILE_M + c * ROWS + r;
        const bool in_range = t < p.n_out;
#pragma unroll
        for (int j = 0; j < NS; ++j)
          a[c][j] = (in_range && j < nk) ? ld_idx(p.nbr + (int64_t)(k0 + j) * p.nbr_stride + t) : -1;
        b[c] = in_range ? (p.order ? ld_idx(p.order + t) : (int32_t)t) : 0;
      }
    };
    int64_t tile = walk.next();
    if (tile >= 0) request(tile, ri, ro);
    while (tile >= 0) {
      const int64_t nxt = walk.next();
      if (nxt >= 0) request(nxt, ri_n, ro_n);
#pragma unroll
      for (int c = 0; c < 4; ++c, ++it) {
        const int stage = it % NST;
        if (it >= NST) mbar_wait(empty_bar(stage), (uint32_t)(((it / NST) - 1) & 1));
        const uint32_t a_stage = base + stage * S::STAGE_BYTES;
        const uint32_t b_stage = a_stage + S::A_BYTES;
        bool any = false;
#pragma unroll
        for (int j = 0; j < NS; ++j) {
          const bool ok = ri[c][j] >= 0;
          any |= ok;
          const char* src_a = ok ? in_b + ri[c][j] * in_row : reinterpret_cast<const char*>(p.in);
          for (int s = 0; s < spo; ++s)
            cp16(a_stage + (j * spo + s) * SLAB + off, ok ? src_a + s * 128 : src_a, ok ? 16 : 0);
        }
        const char* src_b = any ? go_b + ro[c] * go_row : reinterpret_cast<const char*>(p.gout);
        for (int s = 0; s < nb; ++s) cp16(b_stage + s * SLAB + off, any ? src_b + s * 128 : src_b, any ? 16 : 0);
        cp_arrive_noinc(full_bar(stage));
      }
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        ro[c] = ro_n[c];
#pragma unroll
        for (int j = 0; j < NS; ++j) ri[c][j] = ri_n[c][j];
      }
      tile = nxt;
    }
    issued = it;
    cp_commit_wait_all();
  } else {
    // ============================ MMA issuer ============================
    int it = 0;
    constexpr uint32_t idesc = idesc_mn(128, NT);
    TileWalk walk(p.tapmask, tiles, (int)blockIdx.x, p.S, kmask, lane);
    for (int64_t tile = walk.next(); tile >= 0; tile = walk.next()) {
      if (lane == 0) {
        for (int c = 0; c < 4; ++c) {
          const int stage = (it + c) % NST;
          mbar_wait(full_bar(stage), (uint32_t)(((it + c) / NST) & 1));
          fence_proxy_async();
          tc_fence_after();
          const uint32_t a_stage = base + stage * S::STAGE_BYTES;
          const uint32_t b_stage = a_stage + S::A_BYTES;
#pragma unroll
          for (int kk = 0; kk < ROWS / 8; ++kk)
            umma_tf32(tmem_acc, mn_desc(a_stage + kk * 1024), mn_desc(b_stage + kk * 1024), idesc,
                      (it + c > 0 || kk > 0) ? 1u : 0u);
          umma_commit(empty_bar(stage));
        }
      }
      it += 4;
      __syncwarp();
    }
    if (lane == 0 && it > 0) umma_commit(accum_bar);
    __syncwarp();
  }

  // ============================ epilogue: TMEM -> partial tile ============================
  if (warp < PWARPS) {
    if (issued > 0) {
      mbar_wait(accum_bar, 0);
      tc_fence_after();
    }
    const int quad = warp & 3;
    const int row = quad * 32 + lane;                       // accumulator lane
    const int j = NS > 1 ? row / p.cin : 0;                 // which of the CTA's offsets it belongs to
    const int ci = NS > 1 ? row - j * p.cin : ci0 + row;
    const bool row_live = NS > 1 ? j < nk : ci < p.cin;
    float* dst = p.part + (((int64_t)blockIdx.x * p.K + (k0 + (row_live ? j : 0))) * p.cin + (row_live ? ci : 0)) * p.cout;
    constexpr int HALF = NT / 2 >= 32 ? NT / 2 : 32;
    for (int c0 = (warp >> 2) * HALF; c0 < ((warp >> 2) + 1) * HALF && c0 < p.cout; c0 += 32) {
      uint32_t v[32];
      if (issued > 0) {
        tmem_ld32(tmem_acc + ((uint32_t)(quad * 32) << 16) + (uint32_t)c0, v);
      } else {
#pragma unroll
        for (int i = 0; i < 32; ++i) v[i] = 0u;
      }
      if (row_live) {
#pragma unroll
        for (int i = 0; i < 32; i += 4)
          *reinterpret_cast<float4*>(dst + c0 + i) = make_float4(__uint_as_float(v[i]), __uint_as_float(v[i + 1]),
                                                                 __uint_as_float(v[i + 2]), __uint_as_float(v[i + 3]));
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == PWARPS) tmem_dealloc(tmem_acc, NT);
}

template <int NT, int NS>
static int launch(const Params& p, int k_count, int ci_tiles, cudaStream_t st) {
  static bool attr_done = false;
  if (!attr_done) {
    PCC_CHECK_CUDA(cudaFuncSetAttribute(wgrad_tf32_kernel<NT, NS>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        Smem<NT>::TOTAL));
    attr_done = true;
  }
  dim3 grid((unsigned)p.S, (unsigned)ceil_div(k_count, NS), (unsigned)ci_tiles);
  wgrad_tf32_kernel<NT, NS><<<grid, THREADS, Smem<NT>::TOTAL, st>>>(p);
  PCC_CHECK_LAUNCH("wgrad_tf32_kernel");
  return PCC_OK;
}

template <int NT>
static int launch_ns(const Params& p, int ns, int k_count, int ci_tiles, cudaStream_t st) {
  if (ns == 4) return launch<NT, 4>(p, k_count, ci_tiles, st);
  if (ns == 2) return launch<NT, 2>(p, k_count, ci_tiles, st);
  return launch<NT, 1>(p, k_count, ci_tiles, st);
}

}  // namespace wgtc

bool wgrad_tf32_supported(const pcc_wgrad_args* a) {
  return a->cin % 32 == 0 && a->cout % 32 == 0 && a->cin >= 32 && a->cout >= 32 && a->cout <= 256 &&
         (reinterpret_cast<uintptr_t>(a->in) & 15) == 0 && (reinterpret_cast<uintptr_t>(a->gout) & 15) == 0 &&
         (reinterpret_cast<uintptr_t>(a->ws) & 15) == 0;
}

// Launches the partial-sum kernels for every offset group; *splits receives the number of partial tiles
// per (k, ci, co) the caller has to add (wgrad_reduce_kernel).
int wgrad_tf32(const pcc_wgrad_args* a, int* splits, cudaStream_t st) {
  using namespace wgtc;
  Params p;
  p.in = a->in; p.gout = a->gout; p.nbr = a->nbr; p.part = reinterpret_cast<float*>(a->ws);
  p.nbr_stride = a->nbr_stride; p.n_out = a->n_out; p.cin = a->cin; p.cout = a->cout; p.K = a->K;
  const int ng = a->n_groups > 1 ? a->n_groups : 1;
  const int per = (a->K + ng - 1) / ng;
  const int64_t tiles = ceil_div(a->n_out, PCC_TILE_M);
  const int ci_tiles = (int)ceil_div(a->cin, 128);
  // narrow layers put 128 / cin kernel offsets side by side in the MMA's M dimension
  const int ns = a->cin == 32 ? 4 : a->cin == 64 ? 2 : 1;
  // two CTAs per SM over (splits x offset stacks of a launch x ci tiles); never more splits than row tiles, at most 64
  int64_t s = (int64_t)2 * sm_count() / std::max(1, (int)ceil_div(per, ns) * ci_tiles);
  s = std::max<int64_t>(1, std::min<int64_t>(std::min<int64_t>(s, tiles), 64));
  p.S = (int)s;
  *splits = p.S;
  for (int g = 0; g < ng; ++g) {
    p.k_lo = g * per;
    const int k_hi = std::min((g + 1) * per, (int)a->K);
    if (p.k_lo >= k_hi) break;
    p.k_hi = k_hi;
    p.order = a->order ? a->order + (int64_t)g * a->n_out : nullptr;
    p.tapmask = a->tapmask ? a->tapmask + (int64_t)g * tiles : nullptr;
    int rc;
    if (a->cout <= 64) rc = launch_ns<64>(p, ns, k_hi - p.k_lo, ci_tiles, st);
    else if (a->cout <= 128) rc = launch_ns<128>(p, ns, k_hi - p.k_lo, ci_tiles, st);
    else rc = launch_ns<256>(p, ns, k_hi - p.k_lo, ci_tiles, st);
    if (rc != PCC_OK) return rc;
  }
  return PCC_OK;
}

}  // namespace pcc
