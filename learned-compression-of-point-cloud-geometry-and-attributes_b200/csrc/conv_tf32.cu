// Sparse convolution forward, tensor-core mode (PCC_MODE_TF32): tcgen05.mma kind::tf32 with the
// fp32 accumulator in TMEM.
//
// Same output-stationary implicit GEMM as conv_exact.cu (one CTA = PCC_TILE_M = 128 output rows x
// BN output channels; kernel offsets skipped through the tile's tap mask), but the contraction
// runs on the 5th-generation tensor cores:
//   * warps 0-3 (128 threads) gather: for every (offset k, 32-channel chunk) step they cp.async
//     the 128 neighbour rows (zero-fill when missing) and the BN x 32 slice of the pre-transposed
//     weights Wt[k][co][ci] into shared memory in the canonical K-major SWIZZLE_128B layout
//     (8-row x 128-byte atoms, 16-byte chunk index XOR row%8), 4 stages deep.  A stage is
//     published with  cp.async.wait_group -> fence.proxy.async -> mbarrier.arrive  (generic-proxy
//     writes must be fenced before the tensor core's async-proxy reads);
//   * warp 4, one elected lane, waits for the stage, issues 4 x tcgen05.mma (M=128, N=BN, K=8)
//     and tcgen05.commit's the stage's "empty" mbarrier; after the last step it commits the
//     accumulator barrier;
//   * warps 0-3 then read their 32 TMEM lanes (tcgen05.ld 32x32b.x32), apply the fused epilogue
//     (bias -> activation -> FiLM -> residual) and store.
// Operands are fp32 bit patterns read as tf32 (low 13 mantissa bits ignored); products are exact,
// accumulation is fp32 in TMEM.  Tolerance vs the exact mode is stated in tests/.
//
// Roofline: tensor pipe (tf32 dense) for cin, cout >= 64; the gather (128 B per row per step from
// L2) is the competing bound, see DESIGN.md.
#include <cuda.h>
#include <stdlib.h>

#include "pcc_common.cuh"

namespace pcc {

constexpr int TC_TM = PCC_TILE_M;      // 128 rows = UMMA M
constexpr int TC_TK = 32;              // fp32 channels per stage = 128 bytes = one swizzle row
constexpr int TC_STAGES = 3;          // = TC_LAG + 2: one stage being consumed, one published, one landing
constexpr int TC_PWARPS = 8;           // producer warps.  tools/gather_bench.cu (profiles/r01_gather_microbench.txt):
                                       // the shared-memory fill rate of gathered 128 B row chunks grows with the
                                       // number of issuing threads, not with the groups in flight: 2 CTAs x 128
                                       // threads x 8 copies reach 37-50 GB/s per SM, 2 x 256 x 4 reach 113
constexpr int TC_PRODUCERS = 32 * TC_PWARPS;
constexpr int TC_ROWS_PER_THREAD = TC_TM * 8 / TC_PRODUCERS;  // 16-byte copies per thread and stage
constexpr int TC_THREADS = TC_PRODUCERS + 32;  // + the MMA-issue warp (also the TMEM allocator)
constexpr int TC_LAG = 1;              // cp.async groups a producer thread keeps in flight
// Two CTAs are resident per SM (3 x 32 KB stages + the tile's kernel-map rows each): one CTA's
// epilogue and pipeline fill overlap the other's MMA stream.

struct TcParams {
  const float* in;
  const float* weight_t;  // [K][cout][cin]
  const float* bias;
  const int32_t* nbr;
  const uint32_t* tapmask;
  const float* film;
  const int32_t* film_rows;
  const float* residual;
  const int32_t* res_rows;
  const int32_t* order;
  const int32_t* tile_order;  // CTA b works on tile tile_order[b] (longest tiles first), NULL = identity
  float* out;
  int64_t nbr_stride, n_out, out_stride;
  int cin, cout, K, act, film_c, flags;
  float slope;
};

// ---------------------------------------------------------------- PTX wrappers
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void tc_cp_async16(uint32_t dst, const void* src, int src_bytes) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void tc_cp_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
template <int N>
__device__ __forceinline__ void tc_cp_wait() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N) : "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory"); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("{\n .reg .b64 st;\n mbarrier.arrive.shared::cta.b64 st, [%0];\n}\n" ::"r"(bar) : "memory");
}
// the barrier receives this thread's arrival when all of its earlier cp.async copies have landed
__device__ __forceinline__ void cp_async_arrive_noinc(uint32_t bar) {
  asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];\n" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n"
      " .reg .pred p;\n"
      "WAIT_LOOP:\n"
      " mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1, %2;\n"  // suspend until the phase flips (time cap)
      " @p bra DONE;\n"
      " bra WAIT_LOOP;\n"
      "DONE:\n"
      "}\n" ::"r"(bar), "r"(parity), "r"(0x989680u) : "memory");
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
// commit that arrives on the barrier at the same CTA-relative offset in every CTA of `mask`
__device__ __forceinline__ void umma_commit_mc(uint32_t bar, uint16_t mask) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;\n" ::"r"(bar),
               "h"(mask)
               : "memory");
}
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;\n" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n barrier.cluster.wait.acquire.aligned;\n" ::: "memory");
}
__device__ __forceinline__ void umma_tf32(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                          uint32_t accumulate) {
  asm volatile(
      "{\n"
      " .reg .pred p;\n"
      " setp.ne.b32 p, %4, 0;\n"
      " tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n"
      "}\n" ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate) : "memory");
}
// 32 lanes x 32 consecutive 32-bit columns -> 32 registers per thread (thread i <-> lane base+i)
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

__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("{\n .reg .b64 st;\n mbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1;\n}\n" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tma_gather4(uint32_t dst, const void* desc, uint32_t bar, int c0, int r0, int r1,
                                            int r2, int r3) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile::gather4.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4, %5, %6, %7}], [%2];\n" ::"r"(dst), "l"(desc), "r"(bar), "r"(c0), "r"(r0), "r"(r1), "r"(r2),
      "r"(r3)
      : "memory");
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const void* desc, uint32_t bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4}], [%2];\n" ::"r"(dst), "l"(desc), "r"(bar), "r"(c0), "r"(c1)
      : "memory");
}

// tiled 2-D load delivered to the same CTA-relative address (and signalled on the same CTA-relative
// barrier) in every CTA of `mask`
__device__ __forceinline__ void tma_load_2d_mc(uint32_t dst, const void* desc, uint32_t bar, int c0, int c1, uint16_t mask) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes.multicast::cluster"
      " [%0], [%1, {%3, %4}], [%2], %5;\n" ::"r"(dst), "l"(desc), "r"(bar), "r"(c0), "r"(c1), "h"(mask)
      : "memory");
}

// K-major, SWIZZLE_128B shared-memory matrix descriptor (sm_100 encoding):
//   [0,14) start address >> 4, [16,30) leading byte offset >> 4 (unused for swizzled K-major),
//   [32,46) stride byte offset >> 4 (1024 B between 8-row atoms), [46,48) version = 1,
//   [61,64) layout type = 2 (SWIZZLE_128B)
__device__ __forceinline__ uint64_t make_sw128_desc(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr & 0x3FFFFu) >> 4);
  d |= (uint64_t)(1024u >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}

// instruction descriptor, kind::tf32: D = f32 (bits 4-5 = 1), A = B = tf32 (format 2 at bits 7-9,
// 10-12), both K-major, N >> 3 at [17,23), M >> 4 at [24,29)
__host__ __device__ constexpr uint32_t make_idesc_tf32(int m, int n) {
  return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(m >> 4) << 24);
}

// byte offset of (row, 16-byte chunk) inside a [rows x 128 B] SWIZZLE_128B K-major tile
__device__ __forceinline__ uint32_t sw128_offset(int row, int chunk) {
  return (uint32_t)((row >> 3) * 1024 + (row & 7) * 128 + ((chunk ^ (row & 7)) << 4));
}

// stages of the hybrid kernel by tile width: the narrower tiles have smaller stages, so a fourth one
// still leaves room for 2 CTAs per SM — three stages in flight instead of two (the kernel is
// latency-bound on bytes in flight)
template <int BN>
struct TcSmem {
  static constexpr int STAGES = BN <= 64 ? 4 : TC_STAGES;
  static constexpr int LAG = STAGES - 2;  // cp.async groups a producer thread keeps in flight
  static constexpr int A_BYTES = TC_TM * 128;
  static constexpr int B_BYTES = BN * 128;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int BAR_OFFSET = STAGES * STAGE_BYTES;
  static constexpr int NBR_OFFSET = BAR_OFFSET + 128;      // int32 [active taps <= 27][128 rows]
  static constexpr int TOTAL = NBR_OFFSET + 27 * TC_TM * 4 + 1024;  // + alignment slack
};

// CL = 2: the CTAs 2c and 2c + 1 (row tiles 2c, 2c + 1) form a cluster and walk the union of their
// offsets in lockstep; each loads one half of every weight tile and multicasts it to both, so the
// weight traffic out of L2 halves.  A stage is refilled only after BOTH CTAs' MMAs released it.
template <int BN, int CL>
__global__ void __launch_bounds__(TC_THREADS, 2)
conv_tf32_kernel(const __grid_constant__ CUtensorMap tm_b, const TcParams p) {
  extern __shared__ unsigned char smem_raw[];
  // 1024-byte alignment for the swizzle atoms
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  unsigned char* sbase = smem_raw + (base - raw);
  using S = TcSmem<BN>;
  constexpr int NST = S::STAGES, LAG = S::LAG;
  const uint32_t bar_base = base + S::BAR_OFFSET;
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (NST + s); };
  const uint32_t accum_bar = bar_base + 8u * (2 * NST);
  const uint32_t tmem_slot = bar_base + 8u * (2 * NST + 1);
  volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(sbase + S::BAR_OFFSET + 8 * (2 * NST + 1));

  const int tid = threadIdx.x;
  const int warp = tid >> 5, lane = tid & 31;
  const unsigned tile = (CL == 1 && p.tile_order) ? (unsigned)__ldg(p.tile_order + blockIdx.x) : blockIdx.x;
  const int64_t tile_base = (int64_t)tile * TC_TM;
  const int co_base = blockIdx.y * BN;
  const int nchunks = p.cin / TC_TK;
  const uint32_t kfull = p.K >= 32 ? 0xffffffffu : ((1u << p.K) - 1u);
  const unsigned n_tiles = (unsigned)((p.n_out + TC_TM - 1) / TC_TM);
  uint32_t mask = kfull;
  if (p.tapmask) {
    mask = tile < n_tiles ? (p.tapmask[tile] & kfull) : 0u;
    if (CL == 2) {
      const unsigned other = tile ^ 1u;
      if (other < n_tiles) mask |= p.tapmask[other] & kfull;
    }
  }
  const uint32_t crank = CL == 2 ? cluster_ctarank() : 0u;
  const int nsteps = __popc(mask) * nchunks;
  constexpr uint32_t TMEM_COLS = BN <= 32 ? 32 : (BN <= 64 ? 64 : (BN <= 128 ? 128 : 256));

  if (tid == 0) {
    for (int s = 0; s < NST; ++s) {
      mbar_init(full_bar(s), TC_PRODUCERS + 1);  // the gather threads' arrivals + the weight tile's expect_tx
      mbar_init(empty_bar(s), CL);               // one commit per CTA of the cluster
    }
    mbar_init(accum_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
  }
  if (warp == TC_PWARPS) tmem_alloc(tmem_slot, TMEM_COLS);
  // the tile's kernel-map rows for its active offsets, staged once (coalesced) so that gather
  // addresses never wait on a dependent global load
  int32_t* s_nbr = reinterpret_cast<int32_t*>(sbase + S::NBR_OFFSET);
  {
    const int nact = __popc(mask);
    for (int a = warp; a < nact; a += TC_THREADS / 32) {
      const int k = __fns(mask, 0, a + 1);
#pragma unroll
      for (int r = lane; r < TC_TM; r += 32) {
        const int64_t t = tile_base + r;
        s_nbr[a * TC_TM + r] = (t < p.n_out) ? __ldg(p.nbr + (int64_t)k * p.nbr_stride + t) : -1;
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (CL == 2) cluster_sync_all();  // the partner's barriers exist before anything is sent to them
  tc_fence_after();
  const uint32_t tmem_acc = *tmem_slot_ptr;

  if (warp < TC_PWARPS) {
    // ============================ producers: gather A rows, load B slice ============================
    const int chunk = tid & 7;   // 16-byte chunk inside the 128-byte row
    constexpr int RSTEP = TC_PRODUCERS / 8;  // rows rbase + RSTEP j
    uint32_t pmask = mask;
    int pk = pmask ? (__ffs(pmask) - 1) : 0;
    int pa = 0;  // ordinal of pk among the active offsets
    int pchunk = 0;
    // per offset: source pointer of this thread's 16-byte chunk in each of its 8 rows (nullptr =
    // missing neighbour -> zero fill); per step only the channel-chunk displacement is added.
    // Destination: rows rbase + RSTEP j share (row & 7), so the swizzled offset is off0 + 128 RSTEP j.
    const int rbase = tid >> 3;
    const char* rowp[TC_ROWS_PER_THREAD];
    bool need_idx = true;
    const uint32_t off0 = sw128_offset(rbase, chunk);
    const char* in_b = reinterpret_cast<const char*>(p.in) + chunk * 16;
    const int64_t row_bytes = (int64_t)p.cin * 4;
    for (int s = 0; s < nsteps; ++s) {
      const int stage = s % NST;
      if (s >= NST) mbar_wait(empty_bar(stage), (uint32_t)(((s / NST) - 1) & 1));
      if (need_idx) {
#pragma unroll
        for (int j = 0; j < TC_ROWS_PER_THREAD; ++j) {
          const int ix = s_nbr[pa * TC_TM + rbase + RSTEP * j];
          rowp[j] = ix >= 0 ? in_b + ix * row_bytes : nullptr;
        }
        need_idx = false;
      }
      const uint32_t a_stage = base + stage * S::STAGE_BYTES;
      const uint32_t b_stage = a_stage + S::A_BYTES;
      const int cdisp = pchunk * (TC_TK * 4);
#pragma unroll
      for (int j = 0; j < TC_ROWS_PER_THREAD; ++j) {
        const bool ok = rowp[j] != nullptr;
        tc_cp_async16(a_stage + off0 + (uint32_t)(128 * RSTEP) * j, ok ? rowp[j] + cdisp : reinterpret_cast<const char*>(p.in), ok ? 16 : 0);
      }
      // weight slice Wt[pk][co_base .. +BN][ci chunk]: one tiled TMA load (SWIZZLE_128B box) by one thread
      if (tid == 0) {
        mbar_expect_tx(full_bar(stage), (uint32_t)S::B_BYTES);
        if (CL == 2)  // this CTA's half of the weight rows, to both CTAs
          tma_load_2d_mc(b_stage + crank * (S::B_BYTES / 2), &tm_b, full_bar(stage), pchunk * TC_TK,
                         pk * p.cout + co_base + (int)crank * (BN / 2), (uint16_t)3);
        else
          tma_load_2d(b_stage, &tm_b, full_bar(stage), pchunk * TC_TK, pk * p.cout + co_base);
      }
      // asynchronous publish: no thread waits for its copies; the stage's barrier completes when the
      // last copy of the last thread (and the weight tile) has landed
      cp_async_arrive_noinc(full_bar(stage));
      if (++pchunk == nchunks) {
        pchunk = 0;
        pmask &= pmask - 1;
        pk = pmask ? (__ffs(pmask) - 1) : 0;
        ++pa;
        need_idx = true;
      }
    }
    tc_cp_commit();
    tc_cp_wait<0>();  // nothing of this CTA's copies is in flight when the epilogue starts
  } else {
    // ============================ MMA issuer (one elected lane) ============================
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc_tf32(TC_TM, BN);
      for (int s = 0; s < nsteps; ++s) {
        const int stage = s % NST;
        mbar_wait(full_bar(stage), (uint32_t)((s / NST) & 1));
        fence_proxy_async();  // the gathered rows were written through the generic proxy (cp.async)
        tc_fence_after();
        const uint32_t a_stage = base + stage * S::STAGE_BYTES;
        const uint32_t b_stage = a_stage + S::A_BYTES;
#pragma unroll
        for (int kk = 0; kk < TC_TK / 8; ++kk) {
          const uint64_t da = make_sw128_desc(a_stage + kk * 32);
          const uint64_t db = make_sw128_desc(b_stage + kk * 32);
          umma_tf32(tmem_acc, da, db, idesc, (s > 0 || kk > 0) ? 1u : 0u);
        }
        if (CL == 2)
          umma_commit_mc(empty_bar(stage), (uint16_t)3);  // both CTAs' producers wait for both MMAs
        else
          umma_commit(empty_bar(stage));  // frees the smem stage when these MMAs retire
      }
      umma_commit(accum_bar);           // accumulator complete
    }
    __syncwarp();
  }

  // ============================ epilogue ============================
  // warp w reads TMEM lanes 32 (w % 4) .. + 31 (the hardware's lane quadrant of a warp); warps 0-3 take
  // the first EPI_COLS columns, warps 4-7 the next, ...
  constexpr int EPI_COLS = (BN / (TC_PWARPS / 4)) >= 32 ? BN / (TC_PWARPS / 4) : 32;  // columns per warp group
  if (warp < TC_PWARPS && (warp >> 2) * EPI_COLS < BN) {
    if (nsteps > 0) {
      mbar_wait(accum_bar, 0);
      tc_fence_after();
    }
    const int quad = warp & 3;
    const int64_t t = tile_base + quad * 32 + lane;  // tile row -> output row
    const bool live = t < p.n_out;
    const int64_t o = (live && p.order) ? (int64_t)__ldg(p.order + t) : t;
    int64_t frow = o, rrow = o;
    if (live && p.film_c && p.film_rows) frow = __ldg(p.film_rows + o);
    if (live && p.residual && p.res_rows) rrow = __ldg(p.res_rows + o);
    const bool vec4 = (p.out_stride % 4 == 0) && ((reinterpret_cast<uintptr_t>(p.out) & 15) == 0);
    const bool fvec = p.cout % 4 == 0 && (reinterpret_cast<uintptr_t>(p.bias) & 15) == 0 &&
                      (reinterpret_cast<uintptr_t>(p.residual) & 15) == 0 &&
                      (!p.film_c || ((reinterpret_cast<uintptr_t>(p.film) & 15) == 0 && (p.film_c == 1 || p.film_c % 4 == 0)));
#pragma unroll 1
    for (int c0 = (warp >> 2) * EPI_COLS; c0 < ((warp >> 2) + 1) * EPI_COLS && c0 < BN; c0 += 32) {
      if (co_base + c0 >= p.cout) break;  // warp-uniform
      uint32_t v[32];
      if (nsteps > 0) {
        tmem_ld32(tmem_acc + ((uint32_t)(quad * 32) << 16) + (uint32_t)c0, v);
      } else {
#pragma unroll
        for (int i = 0; i < 32; ++i) v[i] = 0u;
      }
      if (live && vec4 && co_base + c0 + 32 <= p.cout && fvec) {
        // 16-byte accesses for every per-row operand: this thread owns 32 consecutive columns of one output row
        float* dst = p.out + o * p.out_stride + co_base + c0;
#pragma unroll
        for (int i = 0; i < 32; i += 4) {
          const int col = co_base + c0 + i;
          float x[4] = {__uint_as_float(v[i]), __uint_as_float(v[i + 1]), __uint_as_float(v[i + 2]),
                        __uint_as_float(v[i + 3])};
          if (p.bias) {
            const float4 b = __ldg(reinterpret_cast<const float4*>(p.bias + col));
            x[0] = __fadd_rn(x[0], b.x); x[1] = __fadd_rn(x[1], b.y); x[2] = __fadd_rn(x[2], b.z); x[3] = __fadd_rn(x[3], b.w);
          }
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            if (p.act == PCC_ACT_RELU)
              x[j] = fmaxf(x[j], 0.0f);
            else if (p.act == PCC_ACT_LEAKY)
              x[j] = x[j] > 0.0f ? x[j] : __fmul_rn(x[j], p.slope);
          }
          if (p.film_c) {
            float4 beta = make_float4(0.f, 0.f, 0.f, 0.f), gamma = beta;
            if (frow >= 0) {
              if (p.film_c == 1) {
                const float b = __ldg(p.film + frow * 2), g = __ldg(p.film + frow * 2 + 1);
                beta = make_float4(b, b, b, b);
                gamma = make_float4(g, g, g, g);
              } else {
                beta = __ldg(reinterpret_cast<const float4*>(p.film + frow * 2 * p.film_c + col));
                gamma = __ldg(reinterpret_cast<const float4*>(p.film + frow * 2 * p.film_c + p.film_c + col));
              }
            }
            x[0] = __fadd_rn(__fmul_rn(x[0], beta.x), gamma.x); x[1] = __fadd_rn(__fmul_rn(x[1], beta.y), gamma.y);
            x[2] = __fadd_rn(__fmul_rn(x[2], beta.z), gamma.z); x[3] = __fadd_rn(__fmul_rn(x[3], beta.w), gamma.w);
          }
          if (p.residual) {
            float4 r4 = make_float4(0.f, 0.f, 0.f, 0.f);
            if (rrow >= 0) r4 = __ldg(reinterpret_cast<const float4*>(p.residual + rrow * p.cout + col));
            x[0] = __fadd_rn(x[0], r4.x); x[1] = __fadd_rn(x[1], r4.y); x[2] = __fadd_rn(x[2], r4.z); x[3] = __fadd_rn(x[3], r4.w);
          }
          if (p.flags & PCC_CONV_ROUND_OUT_TF32) {
#pragma unroll
            for (int j = 0; j < 4; ++j) x[j] = round_tf32(x[j]);
          }
          *reinterpret_cast<float4*>(dst + i) = make_float4(x[0], x[1], x[2], x[3]);
        }
      } else if (live) {
        float r[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          const int col = co_base + c0 + i;
          float x = __uint_as_float(v[i]);
          if (col < p.cout) {
            if (p.bias) x = __fadd_rn(x, __ldg(p.bias + col));
            if (p.act == PCC_ACT_RELU)
              x = fmaxf(x, 0.0f);
            else if (p.act == PCC_ACT_LEAKY)
              x = x > 0.0f ? x : __fmul_rn(x, p.slope);
            if (p.film_c) {
              float beta = 0.0f, gamma = 0.0f;
              if (frow >= 0) {
                const int fc = p.film_c == 1 ? 0 : col;
                beta = __ldg(p.film + frow * 2 * p.film_c + fc);
                gamma = __ldg(p.film + frow * 2 * p.film_c + p.film_c + fc);
              }
              x = __fadd_rn(__fmul_rn(x, beta), gamma);
            }
            if (p.residual) x = __fadd_rn(x, rrow >= 0 ? __ldg(p.residual + rrow * p.cout + col) : 0.0f);
            if (p.flags & PCC_CONV_ROUND_OUT_TF32) x = round_tf32(x);
          }
          r[i] = x;
        }
        float* dst = p.out + o * p.out_stride + co_base + c0;
        if (vec4 && co_base + c0 + 32 <= p.cout) {
#pragma unroll
          for (int i = 0; i < 32; i += 4)
            *reinterpret_cast<float4*>(dst + i) = make_float4(r[i], r[i + 1], r[i + 2], r[i + 3]);
        } else {
#pragma unroll
          for (int i = 0; i < 32; ++i)
            if (co_base + c0 + i < p.cout) dst[i] = r[i];
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (CL == 2) cluster_sync_all();  // no CTA leaves while its partner may still send to it
  if (warp == TC_PWARPS) tmem_dealloc(tmem_acc, TMEM_COLS);
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_tiled() {
  static EncodeTiledFn fn = [] {
    void* f = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
      f = nullptr;
    return reinterpret_cast<EncodeTiledFn>(f);
  }();
  return fn;
}

// 2-D fp32 tensor [rows x cols] (row-major), box = box_rows x 32 columns (128 B), SWIZZLE_128B
static int make_map_2d(CUtensorMap* m, const float* ptr, int64_t rows, int64_t cols, int box_rows) {
  EncodeTiledFn enc = encode_tiled();
  if (!enc) {
    set_error("cuTensorMapEncodeTiled is not available from the driver");
    return PCC_ERR_CUDA;
  }
  const cuuint64_t gdim[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  const cuuint64_t gstr[1] = {(cuuint64_t)cols * 4};
  const cuuint32_t box[2] = {32u, (cuuint32_t)box_rows};
  const cuuint32_t estr[2] = {1u, 1u};
  const CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(ptr), gdim, gstr, box, estr,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed (%d) for a [%lld x %lld] tensor", (int)r, (long long)rows, (long long)cols);
    return PCC_ERR_CUDA;
  }
  return PCC_OK;
}

static int tf32_cluster() {
  static int v = [] {
    const char* e = getenv("PCC_TF32_CLUSTER");
    return e ? atoi(e) : 0;
  }();
  return v;
}

template <int BN>
static int launch_tc(const TcParams& p, cudaStream_t st) {
  const size_t smem = TcSmem<BN>::TOTAL;
  const unsigned tiles = (unsigned)ceil_div(p.n_out, TC_TM);
  if (tf32_cluster() == 2 && BN >= 16) {
    CUtensorMap tm_b;
    const int rc = make_map_2d(&tm_b, p.weight_t, (int64_t)p.K * p.cout, p.cin, BN / 2);
    if (rc) return rc;
    static const cudaError_t attr =
        cudaFuncSetAttribute(conv_tf32_kernel<BN, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    PCC_CHECK_CUDA(attr);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((tiles + 1u) & ~1u, (unsigned)ceil_div(p.cout, BN));
    cfg.blockDim = dim3(TC_THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = 2;
    at[0].val.clusterDim.y = 1;
    at[0].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    PCC_CHECK_CUDA(cudaLaunchKernelEx(&cfg, conv_tf32_kernel<BN, 2>, tm_b, p));
    return PCC_OK;
  }
  CUtensorMap tm_b;
  const int rc = make_map_2d(&tm_b, p.weight_t, (int64_t)p.K * p.cout, p.cin, BN);
  if (rc) return rc;
  static const cudaError_t attr =
      cudaFuncSetAttribute(conv_tf32_kernel<BN, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  PCC_CHECK_CUDA(attr);
  dim3 grid(tiles, (unsigned)ceil_div(p.cout, BN));
  conv_tf32_kernel<BN, 1><<<grid, TC_THREADS, smem, st>>>(tm_b, p);
  PCC_CHECK_LAUNCH("conv_tf32_kernel");
  return PCC_OK;
}

// ======================================================================================
// TMA variant: the gathered operand rows are fetched by cp.async.bulk.tensor ... tile::gather4
// (one instruction = four rows of the feature map by row index, 128 B each, zero-filled when the
// index is out of bounds — which is how a missing neighbour, index -1, becomes a zero row), the
// weight slice by a plain 2-D tiled TMA load; both land in the SWIZZLE_128B layout the UMMA
// descriptors expect and complete on the stage's mbarrier (expect_tx), so there are no per-thread
// copies, address arithmetic, cp.async waits or proxy fences left:
//   warp 0   producer: per step one elected lane arms the barrier, then each of the 32 lanes issues
//            one gather4 (rows 4 lane .. 4 lane + 3 of the tile) and lane 0 the weight tile
//   warp 1   lane 0 issues tcgen05.mma and commits the stage / accumulator barriers
//   warps 2-5 epilogue (TMEM lanes 32 (warp % 4) ..), fused bias / activation / FiLM / residual
constexpr int TM_THREADS = 192;
constexpr int TM_STAGES = 3;

template <int BN>
struct TmSmem {
  static constexpr int A_BYTES = TC_TM * 128;
  static constexpr int B_BYTES = BN * 128;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int BAR_OFFSET = TM_STAGES * STAGE_BYTES;
  static constexpr int NBR_OFFSET = BAR_OFFSET + 128;
  static constexpr int TOTAL = NBR_OFFSET + 27 * TC_TM * 4 + 1024;
};

template <int BN>
__global__ void __launch_bounds__(TM_THREADS, 2)
conv_tma_kernel(const __grid_constant__ CUtensorMap tm_a, const __grid_constant__ CUtensorMap tm_b, const TcParams p) {
  extern __shared__ unsigned char smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  unsigned char* sbase = smem_raw + (base - raw);
  using S = TmSmem<BN>;
  const uint32_t bar_base = base + S::BAR_OFFSET;
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (TM_STAGES + s); };
  const uint32_t accum_bar = bar_base + 8u * (2 * TM_STAGES);
  const uint32_t tmem_slot = bar_base + 8u * (2 * TM_STAGES + 1);
  volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(sbase + S::BAR_OFFSET + 8 * (2 * TM_STAGES + 1));
  int32_t* s_nbr = reinterpret_cast<int32_t*>(sbase + S::NBR_OFFSET);

  const int tid = threadIdx.x;
  const int warp = tid >> 5, lane = tid & 31;
  const unsigned tile = p.tile_order ? (unsigned)__ldg(p.tile_order + blockIdx.x) : blockIdx.x;
  const int64_t tile_base = (int64_t)tile * TC_TM;
  const int co_base = blockIdx.y * BN;
  const int nchunks = p.cin / TC_TK;
  const uint32_t kfull = p.K >= 32 ? 0xffffffffu : ((1u << p.K) - 1u);
  const uint32_t mask = p.tapmask ? (p.tapmask[tile] & kfull) : kfull;
  const int nact = __popc(mask);
  const int nsteps = nact * nchunks;
  constexpr uint32_t TMEM_COLS = BN <= 32 ? 32 : (BN <= 64 ? 64 : (BN <= 128 ? 128 : 256));

  if (tid == 0) {
    for (int s = 0; s < TM_STAGES; ++s) {
      mbar_init(full_bar(s), 1);
      mbar_init(empty_bar(s), 1);
    }
    mbar_init(accum_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
  }
  if (warp == 1) tmem_alloc(tmem_slot, TMEM_COLS);
  for (int a = warp; a < nact; a += TM_THREADS / 32) {
    const int k = __fns(mask, 0, a + 1);
#pragma unroll
    for (int r = lane; r < TC_TM; r += 32) {
      const int64_t t = tile_base + r;
      s_nbr[a * TC_TM + r] = (t < p.n_out) ? __ldg(p.nbr + (int64_t)k * p.nbr_stride + t) : -1;
    }
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_acc = *tmem_slot_ptr;

  if (warp == 0) {
    // ============================ TMA producer ============================
    uint32_t pmask = mask;
    int pk = pmask ? (__ffs(pmask) - 1) : 0;
    int pa = 0, pchunk = 0;
    int4 rows = make_int4(-1, -1, -1, -1);
    bool need_idx = true;
    const uint32_t a_off = (uint32_t)((lane >> 1) * 1024 + (lane & 1) * 512);  // tile rows 4 lane .. 4 lane + 3
    for (int s = 0; s < nsteps; ++s) {
      const int stage = s % TM_STAGES;
      if (s >= TM_STAGES) mbar_wait(empty_bar(stage), (uint32_t)(((s / TM_STAGES) - 1) & 1));
      if (need_idx) {
        rows = *reinterpret_cast<const int4*>(s_nbr + pa * TC_TM + 4 * lane);
        need_idx = false;
      }
      const uint32_t a_stage = base + stage * S::STAGE_BYTES;
      if (lane == 0) mbar_expect_tx(full_bar(stage), (uint32_t)S::STAGE_BYTES);
      __syncwarp();
      tma_gather4(a_stage + a_off, &tm_a, full_bar(stage), pchunk * TC_TK, rows.x, rows.y, rows.z, rows.w);
      if (lane == 0)
        tma_load_2d(a_stage + S::A_BYTES, &tm_b, full_bar(stage), pchunk * TC_TK, pk * p.cout + co_base);
      if (++pchunk == nchunks) {
        pchunk = 0;
        pmask &= pmask - 1;
        pk = pmask ? (__ffs(pmask) - 1) : 0;
        ++pa;
        need_idx = true;
      }
    }
  } else if (warp == 1) {
    // ============================ MMA issuer (one elected lane) ============================
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc_tf32(TC_TM, BN);
      for (int s = 0; s < nsteps; ++s) {
        const int stage = s % TM_STAGES;
        mbar_wait(full_bar(stage), (uint32_t)((s / TM_STAGES) & 1));
        tc_fence_after();
        const uint32_t a_stage = base + stage * S::STAGE_BYTES;
        const uint32_t b_stage = a_stage + S::A_BYTES;
#pragma unroll
        for (int kk = 0; kk < TC_TK / 8; ++kk) {
          const uint64_t da = make_sw128_desc(a_stage + kk * 32);
          const uint64_t db = make_sw128_desc(b_stage + kk * 32);
          umma_tf32(tmem_acc, da, db, idesc, (s > 0 || kk > 0) ? 1u : 0u);
        }
        umma_commit(empty_bar(stage));
      }
      umma_commit(accum_bar);
    }
    __syncwarp();
  } else {
    // ============================ epilogue (warps 2-5 own TMEM lanes 32 (warp % 4) ..) ============================
    const int q = warp & 3;
    if (nsteps > 0) {
      mbar_wait(accum_bar, 0);
      tc_fence_after();
    }
    const int64_t t = tile_base + q * 32 + lane;  // tile row -> output row
    const bool live = t < p.n_out;
    const int64_t o = (live && p.order) ? (int64_t)__ldg(p.order + t) : t;
    int64_t frow = o, rrow = o;
    if (live && p.film_c && p.film_rows) frow = __ldg(p.film_rows + o);
    if (live && p.residual && p.res_rows) rrow = __ldg(p.res_rows + o);
    const bool vec4 = (p.out_stride % 4 == 0) && ((reinterpret_cast<uintptr_t>(p.out) & 15) == 0);
    const bool fvec = p.cout % 4 == 0 && (reinterpret_cast<uintptr_t>(p.bias) & 15) == 0 &&
                      (reinterpret_cast<uintptr_t>(p.residual) & 15) == 0 &&
                      (!p.film_c || ((reinterpret_cast<uintptr_t>(p.film) & 15) == 0 && (p.film_c == 1 || p.film_c % 4 == 0)));
#pragma unroll 1
    for (int c0 = 0; c0 < BN; c0 += 32) {
      if (co_base + c0 >= p.cout) break;  // warp-uniform
      uint32_t v[32];
      if (nsteps > 0) {
        tmem_ld32(tmem_acc + ((uint32_t)(q * 32) << 16) + (uint32_t)c0, v);
      } else {
#pragma unroll
        for (int i = 0; i < 32; ++i) v[i] = 0u;
      }
      if (live && vec4 && co_base + c0 + 32 <= p.cout && fvec) {
        // 16-byte accesses for every per-row operand: this thread owns 32 consecutive columns of one output row
        float* dst = p.out + o * p.out_stride + co_base + c0;
#pragma unroll
        for (int i = 0; i < 32; i += 4) {
          const int col = co_base + c0 + i;
          float x[4] = {__uint_as_float(v[i]), __uint_as_float(v[i + 1]), __uint_as_float(v[i + 2]),
                        __uint_as_float(v[i + 3])};
          if (p.bias) {
            const float4 b = __ldg(reinterpret_cast<const float4*>(p.bias + col));
            x[0] = __fadd_rn(x[0], b.x); x[1] = __fadd_rn(x[1], b.y); x[2] = __fadd_rn(x[2], b.z); x[3] = __fadd_rn(x[3], b.w);
          }
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            if (p.act == PCC_ACT_RELU)
              x[j] = fmaxf(x[j], 0.0f);
            else if (p.act == PCC_ACT_LEAKY)
              x[j] = x[j] > 0.0f ? x[j] : __fmul_rn(x[j], p.slope);
          }
          if (p.film_c) {
            float4 beta = make_float4(0.f, 0.f, 0.f, 0.f), gamma = beta;
            if (frow >= 0) {
              if (p.film_c == 1) {
                const float b = __ldg(p.film + frow * 2), g = __ldg(p.film + frow * 2 + 1);
                beta = make_float4(b, b, b, b);
                gamma = make_float4(g, g, g, g);
              } else {
                beta = __ldg(reinterpret_cast<const float4*>(p.film + frow * 2 * p.film_c + col));
                gamma = __ldg(reinterpret_cast<const float4*>(p.film + frow * 2 * p.film_c + p.film_c + col));
              }
            }
            x[0] = __fadd_rn(__fmul_rn(x[0], beta.x), gamma.x); x[1] = __fadd_rn(__fmul_rn(x[1], beta.y), gamma.y);
            x[2] = __fadd_rn(__fmul_rn(x[2], beta.z), gamma.z); x[3] = __fadd_rn(__fmul_rn(x[3], beta.w), gamma.w);
          }
          if (p.residual) {
            float4 r4 = make_float4(0.f, 0.f, 0.f, 0.f);
            if (rrow >= 0) r4 = __ldg(reinterpret_cast<const float4*>(p.residual + rrow * p.cout + col));
            x[0] = __fadd_rn(x[0], r4.x); x[1] = __fadd_rn(x[1], r4.y); x[2] = __fadd_rn(x[2], r4.z); x[3] = __fadd_rn(x[3], r4.w);
          }
          if (p.flags & PCC_CONV_ROUND_OUT_TF32) {
#pragma unroll
            for (int j = 0; j < 4; ++j) x[j] = round_tf32(x[j]);
          }
          *reinterpret_cast<float4*>(dst + i) = make_float4(x[0], x[1], x[2], x[3]);
        }
      } else if (live) {
        float r[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          const int col = co_base + c0 + i;
          float x = __uint_as_float(v[i]);
          if (col < p.cout) {
            if (p.bias) x = __fadd_rn(x, __ldg(p.bias + col));
            if (p.act == PCC_ACT_RELU)
              x = fmaxf(x, 0.0f);
            else if (p.act == PCC_ACT_LEAKY)
              x = x > 0.0f ? x : __fmul_rn(x, p.slope);
            if (p.film_c) {
              float beta = 0.0f, gamma = 0.0f;
              if (frow >= 0) {
                const int fc = p.film_c == 1 ? 0 : col;
                beta = __ldg(p.film + frow * 2 * p.film_c + fc);
                gamma = __ldg(p.film + frow * 2 * p.film_c + p.film_c + fc);
              }
              x = __fadd_rn(__fmul_rn(x, beta), gamma);
            }
            if (p.residual) x = __fadd_rn(x, rrow >= 0 ? __ldg(p.residual + rrow * p.cout + col) : 0.0f);
            if (p.flags & PCC_CONV_ROUND_OUT_TF32) x = round_tf32(x);
          }
          r[i] = x;
        }
        float* dst = p.out + o * p.out_stride + co_base + c0;
        if (vec4 && co_base + c0 + 32 <= p.cout) {
#pragma unroll
          for (int i = 0; i < 32; i += 4)
            *reinterpret_cast<float4*>(dst + i) = make_float4(r[i], r[i + 1], r[i + 2], r[i + 3]);
        } else {
#pragma unroll
          for (int i = 0; i < 32; ++i)
            if (co_base + c0 + i < p.cout) dst[i] = r[i];
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_acc, TMEM_COLS);
}

template <int BN>
static int launch_tma(const TcParams& p, int64_t n_in, cudaStream_t st) {
  CUtensorMap tm_a, tm_b;
  int rc = make_map_2d(&tm_a, p.in, n_in, p.cin, 1);
  if (rc) return rc;
  rc = make_map_2d(&tm_b, p.weight_t, (int64_t)p.K * p.cout, p.cin, BN);
  if (rc) return rc;
  const size_t smem = TmSmem<BN>::TOTAL;
  static const cudaError_t attr =
      cudaFuncSetAttribute(conv_tma_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  PCC_CHECK_CUDA(attr);
  dim3 grid((unsigned)ceil_div(p.n_out, TC_TM), (unsigned)ceil_div(p.cout, BN));
  conv_tma_kernel<BN><<<grid, TM_THREADS, smem, st>>>(tm_a, tm_b, p);
  PCC_CHECK_LAUNCH("conv_tma_kernel");
  return PCC_OK;
}

static bool use_tma() {
  static bool v = [] {
    const char* e = getenv("PCC_TF32_TMA");
    return e && atoi(e) == 1;  // measured 1.6x slower than the cp.async gather: off by default
  }();
  return v;
}

bool conv_tf32_supported(const pcc_conv_args* a) {
  return a->cin % 32 == 0 && a->cout % 32 == 0 && a->cout >= 32 && a->cout <= 256 && a->weight_t != nullptr &&
         (reinterpret_cast<uintptr_t>(a->in) & 15) == 0 && (reinterpret_cast<uintptr_t>(a->weight_t) & 15) == 0;
}

int conv_tf32(const pcc_conv_args* a, cudaStream_t st) {
  TcParams p;
  p.in = a->in; p.weight_t = a->weight_t; p.bias = a->bias; p.nbr = a->nbr; p.tapmask = a->tapmask;
  p.film = a->film; p.film_rows = a->film_rows; p.residual = a->residual; p.res_rows = a->res_rows;
  p.order = a->order; p.tile_order = a->tile_order; p.out = a->out; p.nbr_stride = a->nbr_stride; p.n_out = a->n_out; p.out_stride = a->out_stride;
  p.cin = a->cin; p.cout = a->cout; p.K = a->K; p.act = a->act; p.film_c = a->film_c; p.slope = a->slope; p.flags = a->flags;
  if (use_tma() && a->n_in > 0) {
    if (a->cout > 64) return launch_tma<128>(p, a->n_in, st);
    if (a->cout > 32) return launch_tma<64>(p, a->n_in, st);
    return launch_tma<32>(p, a->n_in, st);
  }
  if (a->cout % 128 == 0 || a->cout > 128) return launch_tc<128>(p, st);
  if (a->cout > 64) return launch_tc<128>(p, st);
  if (a->cout > 32) return launch_tc<64>(p, st);
  return launch_tc<32>(p, st);
}

// wt[k][co][ci] = tf32_round(w[k][ci][co])  (the tensor core would otherwise truncate)
__global__ void weight_transpose_kernel(const float* __restrict__ w, int K, int cin, int cout,
                                        float* __restrict__ wt) {
  __shared__ float tile[32][33];
  const int k = blockIdx.z;
  const int ci0 = blockIdx.y * 32, co0 = blockIdx.x * 32;
  const float* wk = w + (int64_t)k * cin * cout;
  float* wtk = wt + (int64_t)k * cin * cout;
  for (int r = threadIdx.y; r < 32; r += blockDim.y) {
    const int ci = ci0 + r, co = co0 + threadIdx.x;
    tile[r][threadIdx.x] = (ci < cin && co < cout) ? round_tf32(wk[(int64_t)ci * cout + co]) : 0.f;
  }
  __syncthreads();
  for (int r = threadIdx.y; r < 32; r += blockDim.y) {
    const int co = co0 + r, ci = ci0 + threadIdx.x;
    if (co < cout && ci < cin) wtk[(int64_t)co * cin + ci] = tile[threadIdx.x][r];
  }
}

}  // namespace pcc

extern "C" int pcc_weight_transpose(const float* w, int32_t K, int32_t cin, int32_t cout, float* wt, void* stream) {
  PCC_REQUIRE(w && wt && K > 0 && cin > 0 && cout > 0, "bad argument");
  dim3 grid((unsigned)pcc::ceil_div(cout, 32), (unsigned)pcc::ceil_div(cin, 32), (unsigned)K);
  pcc::weight_transpose_kernel<<<grid, dim3(32, 8), 0, (cudaStream_t)stream>>>(w, K, cin, cout, wt);
  PCC_CHECK_LAUNCH("weight_transpose_kernel");
  return PCC_OK;
}
