// Sparse convolution forward, exact fp32 mode (PCC_MODE_EXACT).
//
// Output-stationary implicit GEMM: one CTA owns PCC_TILE_M = 128 output rows x TN output
// channels and walks the kernel offsets k = 0..K-1 (skipping offsets no row of the tile has, per
// the kernel map's tap mask) and, inside an offset, the input channels in chunks of TK = 32.
// Gathered input rows and the weight slice are staged into shared memory with cp.async
// (zero-fill for missing neighbours), 3 stages deep; each thread keeps an 8 x (TN/16) register
// tile and issues fp32 FFMA in ascending (k, ci) order, which is the oracle's summation order,
// so the result is bit-identical to the CPU oracle's sparse-conv restatement.  No atomics.
//
// Roofline: FP32 FMA pipe for cin, cout >= 32 (148 SMs x 128 lanes x 2 flop x clock); the
// gathers are served by L2 because rows are in lexicographic order.
#include <cuda.h>
#include <stdlib.h>

#include "pcc_common.cuh"

namespace pcc {

constexpr int CE_TM = PCC_TILE_M;  // 128 output rows
constexpr int CE_TK = 32;          // input channels per stage
constexpr int CE_THREADS = 256;    // 16 x 16
constexpr int CE_STAGES = 3;
constexpr int CE_APITCH = CE_TK + 4;
#ifndef CE_KV_UNROLL
#define CE_KV_UNROLL 8  // 4-channel steps of a stage unrolled in the FMA loop (8 = the whole 32-channel stage)
#endif  // floats; keeps 16 B alignment, rows r and r+1 4 banks apart

__device__ __forceinline__ void cp_async16(void* smem, const void* gmem, int src_bytes) {
  const unsigned s = (unsigned)__cvta_generic_to_shared(smem);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(s), "l"(gmem), "r"(src_bytes));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;\n" ::"n"(N));
}

struct ConvParams {
  const float* in;
  const float* weight;
  const float* bias;
  const int32_t* nbr;
  const uint32_t* tapmask;
  const float* film;
  const int32_t* film_rows;
  const float* residual;
  const int32_t* res_rows;
  const int32_t* order;
  float* out;
  int64_t nbr_stride, n_out, out_stride, n_in;
  int cin, cout, K, act, film_c, flags;
  float slope;
  // offset group of this launch (kernel maps with n_groups > 1: one launch per group, chained through
  // `out`, which holds the raw fp32 partial sums between the launches)
  uint32_t kbits;   // offsets this launch may touch
  int first, last;  // first: accumulators start at 0; last: epilogue applied
  const int32_t* tile_order;  // CTA b works on tile tile_order[b] (longest tiles first), NULL = identity
  int vec4;  // out / bias / FiLM / residual rows can be accessed as aligned float4 (conv_epilogue4)
};

// bias -> activation -> FiLM -> residual -> (tf32 rounding) -> store, one output element
__device__ __forceinline__ void conv_epilogue(const ConvParams& p, int64_t o, int64_t frow, int64_t rrow, int col,
                                              float v) {
  if (p.bias) v = __fadd_rn(v, __ldg(p.bias + col));
  if (p.act == PCC_ACT_RELU)
    v = fmaxf(v, 0.0f);
  else if (p.act == PCC_ACT_LEAKY)
    v = v > 0.0f ? v : __fmul_rn(v, p.slope);
  if (p.film_c) {
    float beta = 0.0f, gamma = 0.0f;
    if (frow >= 0) {
      const int fc = p.film_c == 1 ? 0 : col;
      beta = __ldg(p.film + frow * 2 * p.film_c + fc);
      gamma = __ldg(p.film + frow * 2 * p.film_c + p.film_c + fc);
    }
    v = __fadd_rn(__fmul_rn(v, beta), gamma);
  }
  if (p.residual) {
    const float r = rrow >= 0 ? __ldg(p.residual + rrow * p.cout + col) : 0.0f;
    v = __fadd_rn(v, r);
  }
  if (p.flags & PCC_CONV_ROUND_OUT_TF32) v = round_tf32(v);
  p.out[o * p.out_stride + col] = v;
}

// The same element-wise chain for four consecutive columns col .. col + 3 (col % 4 == 0, p.vec4 set): one 16-byte
// access per operand and row instead of four 4-byte ones — the rows of a tile are scattered over the output
// (mask-sorted tile order), so a warp's scalar store touches 16 sectors for 128 useful bytes.
__device__ __forceinline__ void conv_epilogue4(const ConvParams& p, int64_t o, int64_t frow, int64_t rrow, int col,
                                               float (&v)[4]) {
  if (p.bias) {
    const float4 b = __ldg(reinterpret_cast<const float4*>(p.bias + col));
    v[0] = __fadd_rn(v[0], b.x); v[1] = __fadd_rn(v[1], b.y); v[2] = __fadd_rn(v[2], b.z); v[3] = __fadd_rn(v[3], b.w);
  }
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    if (p.act == PCC_ACT_RELU)
      v[j] = fmaxf(v[j], 0.0f);
    else if (p.act == PCC_ACT_LEAKY)
      v[j] = v[j] > 0.0f ? v[j] : __fmul_rn(v[j], p.slope);
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
    v[0] = __fadd_rn(__fmul_rn(v[0], beta.x), gamma.x); v[1] = __fadd_rn(__fmul_rn(v[1], beta.y), gamma.y);
    v[2] = __fadd_rn(__fmul_rn(v[2], beta.z), gamma.z); v[3] = __fadd_rn(__fmul_rn(v[3], beta.w), gamma.w);
  }
  if (p.residual) {
    float4 r = make_float4(0.f, 0.f, 0.f, 0.f);
    if (rrow >= 0) r = __ldg(reinterpret_cast<const float4*>(p.residual + rrow * p.cout + col));
    v[0] = __fadd_rn(v[0], r.x); v[1] = __fadd_rn(v[1], r.y); v[2] = __fadd_rn(v[2], r.z); v[3] = __fadd_rn(v[3], r.w);
  }
  if (p.flags & PCC_CONV_ROUND_OUT_TF32) {
#pragma unroll
    for (int j = 0; j < 4; ++j) v[j] = round_tf32(v[j]);
  }
  *reinterpret_cast<float4*>(p.out + o * p.out_stride + col) = make_float4(v[0], v[1], v[2], v[3]);
}

// TK: input channels per stage (32; 4 for layers with cin <= 4, which would otherwise multiply 28
// padded zeros per neighbour)
template <int TN, int TK>
struct ExactSmem {
  float a[CE_STAGES][CE_TM][TK + 4];
  float b[CE_STAGES][TK][TN];
};

// VEC: cin % 4 == 0 && cout % 4 == 0 (16-byte cp.async sources are aligned)
// F2: accumulate with the packed fma.rn.f32x2 (two independent IEEE fp32 FMAs per lane and
// instruction — the same roundings as two scalar FMAs, half the issue slots)
__device__ __forceinline__ unsigned long long pack2(float lo, float hi) {
  unsigned long long r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void unpack2(unsigned long long v, float& lo, float& hi) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ void ffma2(unsigned long long& d, unsigned long long a, unsigned long long b) {
  asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(d) : "l"(a), "l"(b));
}


// One pipeline stage of the register-tile loop: TK input channels of the staged rows (a: [CE_TM][TK + 4]) times the
// staged weight slice (b: [TK][TN]) into the thread's 8 x CN accumulators, ascending ci.  Shared by the cp.async
// kernel and the bulk-copy kernel below, so both issue the same FMA chain.
// SW128 = false: a is [CE_TM][TK + 4] (padded rows), rbase = the thread's first row; its rows are rbase + 4 i.
// SW128 = true (the TMA kernel): a is the dense [CE_TM][32] tile in the TMA unit's SWIZZLE_128B layout — the 16-byte
// chunk c of row r sits at chunk c ^ (r % 8) — which is also what keeps the four row-lanes of a warp (rows r, r+1,
// r+2, r+3) on different banks without padding.  rbase then is the float index 32 r0 + 4 (r0 % 8) of the thread's
// first row r0 (r0 % 8 < 4): chunk c of row r0 + 4 i is at  (rbase ^ 4 (c ^ 4 (i & 1))) + 128 i.
template <int TN, int TK, bool PK, bool SW128 = false, int CN = TN / 16>
__device__ __forceinline__ void exact_stage_fma(const float* __restrict__ sa, const float* __restrict__ sb, int rbase,
                                                int cbase, float (&acc)[8][CN],
                                                unsigned long long (&acc2)[8][PK ? CN / 2 : 1]) {
  constexpr int AV = 4;
  constexpr int CP = PK ? CN / 2 : 1;
  constexpr int AP = TK + 4;
  constexpr int KVU = CE_KV_UNROLL;
#pragma unroll KVU
  for (int kv = 0; kv < TK / AV; ++kv) {
    float a[8][AV];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const float* ap = SW128 ? sa + ((rbase ^ (4 * (kv ^ (4 * (i & 1))))) + 128 * i)
                              : sa + (rbase + 4 * i) * AP + kv * AV;
      const float4 v = *reinterpret_cast<const float4*>(ap);
      a[i][0] = v.x; a[i][1] = v.y; a[i][2] = v.z; a[i][3] = v.w;
    }
#pragma unroll
    for (int e = 0; e < AV; ++e) {
      float b[CN];
      const float* brow = sb + (kv * AV + e) * TN;
      if (CN == 8) {
        const float4 b0 = *reinterpret_cast<const float4*>(brow + cbase);
        const float4 b1 = *reinterpret_cast<const float4*>(brow + cbase + 32);
        b[0] = b0.x; b[1] = b0.y; b[2] = b0.z; b[3] = b0.w;
        b[4 % CN] = b1.x; b[5 % CN] = b1.y; b[6 % CN] = b1.z; b[7 % CN] = b1.w;
      } else if (CN == 4) {
        const float4 b0 = *reinterpret_cast<const float4*>(brow + cbase);
        b[0] = b0.x; b[1 % CN] = b0.y; b[2 % CN] = b0.z; b[3 % CN] = b0.w;
      } else if (CN == 2) {
        const float2 b0 = *reinterpret_cast<const float2*>(brow + cbase);
        b[0] = b0.x; b[1 % CN] = b0.y;
      } else {
        b[0] = brow[cbase];
      }
      if (PK) {
        unsigned long long bp[CP];
#pragma unroll
        for (int c = 0; c < CP; ++c) bp[c] = pack2(b[(2 * c) % CN], b[(2 * c + 1) % CN]);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const unsigned long long aa = pack2(a[i][e], a[i][e]);
#pragma unroll
          for (int c = 0; c < CP; ++c) ffma2(acc2[i][c], aa, bp[c]);
        }
      } else {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
#pragma unroll
          for (int c = 0; c < CN; ++c) acc[i][c] = __fmaf_rn(a[i][e], b[c], acc[i][c]);
        }
      }
    }
  }
}

template <int TN, bool VEC, int F2, int TK = CE_TK>
__global__ void __launch_bounds__(CE_THREADS, 2)
conv_exact_kernel(const ConvParams p) {
  constexpr int CN = TN / 16;  // output channels per thread
  constexpr bool PK = (F2 != 0) && (CN % 2 == 0);
  constexpr int CP = PK ? CN / 2 : 1;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  ExactSmem<TN, TK>& sm = *reinterpret_cast<ExactSmem<TN, TK>*>(smem_raw);

  const int tid = threadIdx.x;
  // compute mapping: warp w owns rows 32 (w >> 1) .. +31 and columns (TN / 2)(w & 1) .. ; inside the
  // warp, lane = ty * 8 + tx: thread rows rbase + 4 i, thread columns cbase (+ 32 for the second
  // quad when CN == 8).  For one k-slice the four ty of a warp read 4 rows x 16 B (one conflict-free
  // shared-memory wavefront, broadcast over tx) and the eight tx read 128 contiguous bytes of the
  // weight row (one wavefront, broadcast over ty).
  const int lane = tid & 31, wid = tid >> 5;
  const int tx = lane & 7, ty = lane >> 3;
  const int rbase = (wid >> 1) * 32 + ty;
  const int cbase = (wid & 1) * (TN / 2) + tx * (CN == 8 ? 4 : CN);
  const unsigned tile = p.tile_order ? (unsigned)__ldg(p.tile_order + blockIdx.x) : blockIdx.x;
  const int64_t tile_base = (int64_t)tile * CE_TM;
  const int co_base = blockIdx.y * TN;
  const int nchunks = (p.cin + TK - 1) / TK;
  const uint32_t mask = p.tapmask ? (p.tapmask[tile] & p.kbits) : p.kbits;
  const int nsteps = __popc(mask) * nchunks;
  if (nsteps == 0 && !p.first && !p.last) return;  // nothing to add to the partial sums of this tile

  float acc[8][CN];
  unsigned long long acc2[8][CP];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
#pragma unroll
    for (int c = 0; c < CN; ++c) acc[i][c] = 0.0f;
#pragma unroll
    for (int c = 0; c < CP; ++c) acc2[i][c] = 0ull;
  }
  if (!p.first) {  // continue the FMA chains from the previous group's partial sums
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int64_t t = tile_base + rbase + 4 * i;
      if (t >= p.n_out) continue;
      const int64_t o = p.order ? (int64_t)__ldg(p.order + t) : t;
#pragma unroll
      for (int c = 0; c < CN; ++c) {
        const int col = co_base + cbase + ((CN == 8) ? ((c >> 2) * 32 + (c & 3)) : c);
        if (col < p.cout) acc[i][c] = p.out[o * p.out_stride + col];
      }
    }
    if (PK) {
#pragma unroll
      for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int c = 0; c < CP; ++c) acc2[i][c] = pack2(acc[i][(2 * c) % CN], acc[i][(2 * c + 1) % CN]);
    }
  }

  // producer cursor.  The kernel-map rows of the current offset (idx) and of the next active
  // offset (nidx) are held in registers: the next offset's indices are requested when the current
  // one starts, so the gather addresses never wait on a dependent load.
  uint32_t pmask = mask;
  int pk = pmask ? (__ffs(pmask) - 1) : 0;
  int pchunk = 0;
  int idx[4], nidx[4];
  auto load_idx = [&](int k, int (&dst)[4]) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int64_t t = tile_base + (tid >> 3) + 32 * j;
      dst[j] = (t < p.n_out) ? __ldg(p.nbr + (int64_t)k * p.nbr_stride + t) : -1;
    }
  };
  {
    load_idx(pk, idx);
    const uint32_t rest = pmask & (pmask - 1);
    if (rest) load_idx(__ffs(rest) - 1, nidx);
  }

  auto issue = [&](int stage) {
    const int ci0 = pchunk * TK;
    // ---- A: 128 rows x TK / 4 quads; thread -> quad q = tid & 7, rows (tid >> 3) + 32 j
    if ((tid & 7) < TK / 4) {
      const int q = tid & 7;
      const int ci = ci0 + q * 4;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int r = (tid >> 3) + 32 * j;
        const int ix = idx[j];
        float* dst = &sm.a[stage][r][q * 4];
        if (VEC) {
          const bool ok = ix >= 0 && ci < p.cin;
          const float* src = ok ? p.in + (int64_t)ix * p.cin + ci : p.in;
          cp_async16(dst, src, ok ? 16 : 0);
        } else {
          float v[4] = {0.f, 0.f, 0.f, 0.f};
          if (ix >= 0) {
#pragma unroll
            for (int e = 0; e < 4; ++e)
              if (ci + e < p.cin) v[e] = __ldg(p.in + (int64_t)ix * p.cin + ci + e);
          }
          *reinterpret_cast<float4*>(dst) = make_float4(v[0], v[1], v[2], v[3]);
        }
      }
    }
    // ---- B: TK rows x TN/4 quads
    {
      constexpr int QPR = TN / 4;                 // quads per weight row
      constexpr int TOTAL = TK * QPR;             // quads per stage
      const float* wk = p.weight + (int64_t)pk * p.cin * p.cout;
#pragma unroll
      for (int t = tid; t < TOTAL; t += CE_THREADS) {
        const int kk = t / QPR, q = t % QPR;
        const int ci = ci0 + kk, co = co_base + q * 4;
        float* dst = &sm.b[stage][kk][q * 4];
        if (VEC) {
          const bool ok = ci < p.cin && co < p.cout;
          const float* src = ok ? wk + (int64_t)ci * p.cout + co : p.weight;
          cp_async16(dst, src, ok ? 16 : 0);
        } else {
          float v[4] = {0.f, 0.f, 0.f, 0.f};
          if (ci < p.cin) {
#pragma unroll
            for (int e = 0; e < 4; ++e)
              if (co + e < p.cout) v[e] = __ldg(wk + (int64_t)ci * p.cout + co + e);
          }
          *reinterpret_cast<float4*>(dst) = make_float4(v[0], v[1], v[2], v[3]);
        }
      }
    }
    // advance cursor
    if (++pchunk == nchunks) {
      pchunk = 0;
      pmask &= pmask - 1;
      pk = pmask ? (__ffs(pmask) - 1) : 0;
#pragma unroll
      for (int j = 0; j < 4; ++j) idx[j] = nidx[j];
      const uint32_t rest = pmask & (pmask - 1);
      if (rest) load_idx(__ffs(rest) - 1, nidx);
    }
  };

  // prologue
#pragma unroll
  for (int s = 0; s < CE_STAGES - 1; ++s) {
    if (s < nsteps) issue(s);
    cp_async_commit();
  }

  for (int s = 0; s < nsteps; ++s) {
    cp_async_wait<CE_STAGES - 2>();
    __syncthreads();
    if (s + CE_STAGES - 1 < nsteps) issue((s + CE_STAGES - 1) % CE_STAGES);
    cp_async_commit();

    const int st = s % CE_STAGES;
    exact_stage_fma<TN, TK, PK>(&sm.a[st][0][0], &sm.b[st][0][0], rbase, cbase, acc, acc2);
  }
  cp_async_wait<0>();
  if (PK) {
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
      for (int c = 0; c < CP; ++c) unpack2(acc2[i][c], acc[i][(2 * c) % CN], acc[i][(2 * c + 1) % CN]);
  }

  // ---- epilogue: bias -> activation -> FiLM -> residual -> store
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int64_t t = tile_base + rbase + 4 * i;  // tile row -> output row
    if (t >= p.n_out) continue;
    const int64_t o = p.order ? (int64_t)__ldg(p.order + t) : t;
    int64_t frow = o, rrow = o;
    if (p.film_c && p.film_rows) frow = __ldg(p.film_rows + o);
    if (p.residual && p.res_rows) rrow = __ldg(p.res_rows + o);
#pragma unroll
    for (int c = 0; c < CN; ++c) {
      const int col = co_base + cbase + ((CN == 8) ? ((c >> 2) * 32 + (c & 3)) : c);
      if (col >= p.cout) continue;
      if (p.last)
        conv_epilogue(p, o, frow, rrow, col, acc[i][c]);
      else
        p.out[o * p.out_stride + col] = acc[i][c];
    }
  }
}

// ---------------------------------------------------------------------------------------------
// Bulk-copy producer (cin % 32 == 0, 16-byte aligned operands, packed FMA): same tile, same register mapping and
// the same exact_stage_fma as conv_exact_kernel, but the producer side of a stage — 8 cp.async per thread with
// their address arithmetic, 1024 + 1024 lane-instructions per CTA — becomes ONE cp.async.bulk instruction per warp:
// lanes 0-15 of warp w copy the 128-byte chunks of tile rows 16 w .. 16 w + 15 (a missing neighbour copies the
// zero row), lanes 16-19 the weight rows 4 w .. 4 w + 3 of the stage.  The copies complete on the stage's
// mbarrier (complete_tx), which takes the place of cp.async.wait_group; the one __syncthreads per stage that
// frees the slot being refilled stays.  The FMA order is untouched, so results are bit-identical.
__device__ __align__(128) float g_zero_row[CE_TK];  // 128 zero bytes: the source of a missing neighbour's chunk

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("{\n .reg .b64 st;\n mbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1;\n}\n" ::"r"(smem_u32(bar)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n .reg .pred p;\n"
      "WAIT_%=:\n"
      " mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1, 0x989680;\n"
      " @!p bra WAIT_%=;\n}\n" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n" ::"r"(
                   smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}

template <int TN>
struct BulkSmem {
  float a[CE_STAGES][CE_TM][CE_TK + 4];
  float b[CE_STAGES][CE_TK][TN];
  uint64_t full[CE_STAGES];
};

template <int TN>
__global__ void __launch_bounds__(CE_THREADS, 2)
conv_exact_bulk_kernel(const ConvParams p) {
  constexpr int TK = CE_TK;
  constexpr int CN = TN / 16;
  constexpr int CP = CN / 2;
  static_assert(CN % 2 == 0, "packed accumulators");
  extern __shared__ __align__(16) unsigned char smem_raw[];
  BulkSmem<TN>& sm = *reinterpret_cast<BulkSmem<TN>*>(smem_raw);

  const int tid = threadIdx.x;
  const int lane = tid & 31, wid = tid >> 5;
  const int tx = lane & 7, ty = lane >> 3;
  const int rbase = (wid >> 1) * 32 + ty;
  const int cbase = (wid & 1) * (TN / 2) + tx * (CN == 8 ? 4 : CN);
  const unsigned tile = p.tile_order ? (unsigned)__ldg(p.tile_order + blockIdx.x) : blockIdx.x;
  const int64_t tile_base = (int64_t)tile * CE_TM;
  const int co_base = blockIdx.y * TN;
  const int nchunks = p.cin / TK;
  const uint32_t mask = p.tapmask ? (p.tapmask[tile] & p.kbits) : p.kbits;
  const int nsteps = __popc(mask) * nchunks;
  if (nsteps == 0 && !p.first && !p.last) return;  // CTA-uniform

  if (tid == 0) {
#pragma unroll
    for (int s = 0; s < CE_STAGES; ++s) mbar_init(&sm.full[s], CE_THREADS / 32);  // one expect_tx arrival per warp
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
  }
  __syncthreads();

  float acc[8][CN];
  unsigned long long acc2[8][CP];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
#pragma unroll
    for (int c = 0; c < CN; ++c) acc[i][c] = 0.0f;
#pragma unroll
    for (int c = 0; c < CP; ++c) acc2[i][c] = 0ull;
  }
  if (!p.first) {  // continue the FMA chains from the previous group's partial sums
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int64_t t = tile_base + rbase + 4 * i;
      if (t >= p.n_out) continue;
      const int64_t o = p.order ? (int64_t)__ldg(p.order + t) : t;
#pragma unroll
      for (int c = 0; c < CN; ++c) {
        const int col = co_base + cbase + ((CN == 8) ? ((c >> 2) * 32 + (c & 3)) : c);
        if (col < p.cout) acc[i][c] = p.out[o * p.out_stride + col];
      }
    }
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
      for (int c = 0; c < CP; ++c) acc2[i][c] = pack2(acc[i][2 * c], acc[i][2 * c + 1]);
  }

  // producer cursor: lanes 0-15 hold the kernel-map entry of tile row 16 wid + lane for the current offset (idx)
  // and for the next active one (nidx)
  uint32_t pmask = mask;
  int pk = pmask ? (__ffs(pmask) - 1) : 0;
  int pchunk = 0;
  const int arow = wid * 16 + (lane & 15);
  auto load_idx = [&](int k) -> int {
    const int64_t t = tile_base + arow;
    return (lane < 16 && t < p.n_out) ? __ldg(p.nbr + (int64_t)k * p.nbr_stride + t) : -1;
  };
  int idx = load_idx(pk), nidx = -1;
  {
    const uint32_t rest = pmask & (pmask - 1);
    if (rest) nidx = load_idx(__ffs(rest) - 1);
  }
  const int ncols = (p.cout - co_base) < TN ? (p.cout - co_base) : TN;  // valid columns of this CTA's weight slice
  const uint32_t b_row_bytes = (uint32_t)ncols * 4u;
  const uint32_t warp_tx = 16u * (TK * 4u) + 4u * b_row_bytes;

  auto issue = [&](int stage) {
    const int ci0 = pchunk * TK;
    if (lane == 0) mbar_expect_tx(&sm.full[stage], warp_tx);
    __syncwarp();
    if (lane < 16) {
      const float* src = idx >= 0 ? p.in + (int64_t)idx * p.cin + ci0 : g_zero_row;
      bulk_g2s(&sm.a[stage][arow][0], src, TK * 4u, &sm.full[stage]);
    } else if (lane < 20) {
      const int kk = wid * 4 + (lane - 16);
      const float* src = p.weight + ((int64_t)pk * p.cin + ci0 + kk) * p.cout + co_base;
      bulk_g2s(&sm.b[stage][kk][0], src, b_row_bytes, &sm.full[stage]);
    }
    if (++pchunk == nchunks) {
      pchunk = 0;
      pmask &= pmask - 1;
      pk = pmask ? (__ffs(pmask) - 1) : 0;
      idx = nidx;
      const uint32_t rest = pmask & (pmask - 1);
      if (rest) nidx = load_idx(__ffs(rest) - 1);
    }
  };

#pragma unroll
  for (int s = 0; s < CE_STAGES - 1; ++s)
    if (s < nsteps) issue(s);

  for (int s = 0; s < nsteps; ++s) {
    const int st = s % CE_STAGES;
    mbar_wait(&sm.full[st], (uint32_t)((s / CE_STAGES) & 1));  // this stage's copies have landed
    __syncthreads();  // every warp is done with step s - 1: its slot may be refilled
    if (s + CE_STAGES - 1 < nsteps) issue((s + CE_STAGES - 1) % CE_STAGES);
    exact_stage_fma<TN, TK, true>(&sm.a[st][0][0], &sm.b[st][0][0], rbase, cbase, acc, acc2);
  }
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int c = 0; c < CP; ++c) unpack2(acc2[i][c], acc[i][2 * c], acc[i][2 * c + 1]);

  // ---- epilogue: bias -> activation -> FiLM -> residual -> store
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int64_t t = tile_base + rbase + 4 * i;  // tile row -> output row
    if (t >= p.n_out) continue;
    const int64_t o = p.order ? (int64_t)__ldg(p.order + t) : t;
    int64_t frow = o, rrow = o;
    if (p.film_c && p.film_rows) frow = __ldg(p.film_rows + o);
    if (p.residual && p.res_rows) rrow = __ldg(p.res_rows + o);
#pragma unroll
    for (int c = 0; c < CN; ++c) {
      const int col = co_base + cbase + ((CN == 8) ? ((c >> 2) * 32 + (c & 3)) : c);
      if (col >= p.cout) continue;
      if (p.last)
        conv_epilogue(p, o, frow, rrow, col, acc[i][c]);
      else
        p.out[o * p.out_stride + col] = acc[i][c];
    }
  }
}

template <int TN>
static int launch_exact_bulk(const ConvParams& p, cudaStream_t st) {
  const size_t smem = sizeof(BulkSmem<TN>);
  dim3 grid((unsigned)ceil_div(p.n_out, CE_TM), (unsigned)ceil_div(p.cout, TN));
  static const cudaError_t attr = cudaFuncSetAttribute(conv_exact_bulk_kernel<TN>,
                                                       cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  PCC_CHECK_CUDA(attr);
  conv_exact_bulk_kernel<TN><<<grid, CE_THREADS, smem, st>>>(p);
  PCC_CHECK_LAUNCH("conv_exact_bulk_kernel");
  return PCC_OK;
}

// PCC_EXACT_BULK=0 keeps the cp.async producer everywhere (for A/B timing)
static int exact_bulk() {
  static int v = [] {
    const char* e = getenv("PCC_EXACT_BULK");
    return e ? atoi(e) : 1;
  }();
  return v;
}

// ---------------------------------------------------------------------------------------------
// TMA producer (cin % 32 == 0, cout % 4 == 0): the gathered rows of a stage arrive by 32
// cp.async.bulk.tensor ... tile::gather4 instructions (four rows of the feature matrix by row index, 128 bytes
// each; an index of -1 — a missing neighbour — is out of bounds and is zero-filled by the TMA unit), the weight
// slice by ONE tiled load; lanes 0-3 of each warp issue one gather4 each, thread 0 the weight tile and the
// barrier's expect_tx.  34 producer instructions per CTA and stage instead of 160 bulk copies (or 2048 cp.async lane
// operations), no per-lane address arithmetic.  Same tile, same columns per thread, same exact_stage_fma: the FMA
// chain of every output element is unchanged, so results are bit-identical.  The gathered tile is the dense
// [128][32] SWIZZLE_128B layout (TMA destinations must be 128-byte aligned, so rows cannot be padded); the swizzle
// doubles as the bank-conflict cure, at the price of one XOR per shared-memory address (exact_stage_fma).
__device__ __forceinline__ void tma_gather4(uint32_t dst, const void* desc, uint32_t bar, int c0, int r0, int r1, int r2,
                                            int r3) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile::gather4.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4, %5, %6, %7}], [%2];\n" ::"r"(dst), "l"(desc), "r"(bar), "r"(c0), "r"(r0), "r"(r1), "r"(r2),
      "r"(r3)
      : "memory");
}
__device__ __forceinline__ void tma_tile_2d(uint32_t dst, const void* desc, uint32_t bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4}], [%2];\n" ::"r"(dst), "l"(desc), "r"(bar), "r"(c0), "r"(c1)
      : "memory");
}

template <int TN>
struct TmaSmem {
  float a[CE_STAGES][CE_TM * CE_TK];  // 16 KB per stage, 1024-byte aligned: SWIZZLE_128B atoms of 8 rows
  float b[CE_STAGES][CE_TK][TN];
  uint64_t full[CE_STAGES];
};

// CN: output channels per thread.  16 row-threads x TN / CN column-threads: <128, 8> and <64, 4> are the 256-thread
// shapes of conv_exact_kernel; <64, 8> covers a 64-wide tile with 128 threads that keep the 8 x 8 register tile
// (a third fewer shared-memory operand bytes per FMA than the 8 x 4 tile), three CTAs per SM.
template <int TN, int CN>
__global__ void __launch_bounds__(16 * TN / CN, (16 * TN / CN) == 128 ? 3 : 2)
conv_exact_tma_kernel(const __grid_constant__ CUtensorMap tm_a, const __grid_constant__ CUtensorMap tm_b,
                      const ConvParams p) {
  constexpr int TK = CE_TK;
  constexpr int NT = 16 * TN / CN;     // threads
  constexpr int WCOLS = 8 * CN;        // columns a warp covers (CN == 8: two quads 32 apart)
  constexpr int NWX = TN / WCOLS;      // warps side by side
  constexpr int GPW = 32 / (NT / 32);  // gather groups issued per warp
  constexpr int CP = CN / 2;
  static_assert(CN % 2 == 0, "packed accumulators");
  extern __shared__ unsigned char smem_tma_raw[];
  // the swizzle pattern is a function of the shared-memory address: the tile must start on a 1024-byte boundary
  unsigned char* smem_tma = smem_tma_raw + ((1024u - (smem_u32(smem_tma_raw) & 1023u)) & 1023u);
  TmaSmem<TN>& sm = *reinterpret_cast<TmaSmem<TN>*>(smem_tma);

  const int tid = threadIdx.x;
  const int lane = tid & 31, wid = tid >> 5;
  const int tx = lane & 7, ty = lane >> 3;
  const int rbase = (wid / NWX) * 32 + ty;  // rows rbase + 4 i
  const int cbase = (wid % NWX) * WCOLS + tx * (CN == 8 ? 4 : CN);
  const unsigned tile = p.tile_order ? (unsigned)__ldg(p.tile_order + blockIdx.x) : blockIdx.x;
  const int64_t tile_base = (int64_t)tile * CE_TM;
  const int co_base = blockIdx.y * TN;
  const int nchunks = p.cin / TK;
  const uint32_t mask = p.tapmask ? (p.tapmask[tile] & p.kbits) : p.kbits;
  const int nsteps = __popc(mask) * nchunks;
  if (nsteps == 0 && !p.first && !p.last) return;  // CTA-uniform

  if (tid == 0) {
#pragma unroll
    for (int s = 0; s < CE_STAGES; ++s) mbar_init(&sm.full[s], 1);  // thread 0's expect_tx arrival
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
  }
  __syncthreads();

  float acc[8][CN];
  unsigned long long acc2[8][CP];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
#pragma unroll
    for (int c = 0; c < CN; ++c) acc[i][c] = 0.0f;
#pragma unroll
    for (int c = 0; c < CP; ++c) acc2[i][c] = 0ull;
  }
  if (!p.first) {  // continue the FMA chains from the previous group's partial sums
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int64_t t = tile_base + rbase + 4 * i;
      if (t >= p.n_out) continue;
      const int64_t o = p.order ? (int64_t)__ldg(p.order + t) : t;
      if constexpr (CN % 4 == 0) {
        if (p.vec4) {  // cout % 4 == 0: a quad is inside the row or entirely outside
#pragma unroll
          for (int q = 0; q < CN / 4; ++q) {
            const int col = co_base + cbase + 32 * q;
            if (col < p.cout) {
              const float4 v = *reinterpret_cast<const float4*>(p.out + o * p.out_stride + col);
              acc[i][4 * q] = v.x; acc[i][4 * q + 1] = v.y; acc[i][4 * q + 2] = v.z; acc[i][4 * q + 3] = v.w;
            }
          }
          continue;
        }
      }
#pragma unroll
      for (int c = 0; c < CN; ++c) {
        const int col = co_base + cbase + ((CN == 8) ? ((c >> 2) * 32 + (c & 3)) : c);
        if (col < p.cout) acc[i][c] = p.out[o * p.out_stride + col];
      }
    }
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
      for (int c = 0; c < CP; ++c) acc2[i][c] = pack2(acc[i][2 * c], acc[i][2 * c + 1]);
  }

  // producer cursor: lanes 0 .. GPW-1 of warp w hold the kernel-map entries of one gather group each (tile rows
  // 4 (GPW w + lane) .. + 3) for the offset being issued; they are re-read for the next active offset right after its last chunk
  // went out, a whole stage before they are needed
  uint32_t pmask = mask;
  int pk = pmask ? (__ffs(pmask) - 1) : 0;
  int pchunk = 0;
  const int grp = wid * GPW + (lane % GPW);  // gather group issued by this lane (lanes < GPW)
  int idx[4];
  auto load_idx = [&](int k) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int64_t t = tile_base + 4 * grp + j;
      idx[j] = (lane < GPW && t < p.n_out) ? __ldg(p.nbr + (int64_t)k * p.nbr_stride + t) : -1;
    }
  };
  load_idx(pk);
  constexpr uint32_t STAGE_TX = CE_TM * TK * 4u + TK * TN * 4u;  // the TMA unit counts whole boxes, zero fill included

  auto issue = [&](int stage) {
    const uint32_t bar = smem_u32(&sm.full[stage]);
    if (tid == 0) {
      mbar_expect_tx(&sm.full[stage], STAGE_TX);
      tma_tile_2d(smem_u32(&sm.b[stage][0][0]), &tm_b, bar, co_base, pk * p.cin + pchunk * TK);
    }
    if (lane < GPW)
      tma_gather4(smem_u32(&sm.a[stage][grp * 4 * TK]), &tm_a, bar, pchunk * TK, idx[0], idx[1], idx[2], idx[3]);
    if (++pchunk == nchunks) {
      pchunk = 0;
      pmask &= pmask - 1;
      pk = pmask ? (__ffs(pmask) - 1) : 0;
      if (pmask) load_idx(pk);
    }
  };

#pragma unroll
  for (int s = 0; s < CE_STAGES - 1; ++s)
    if (s < nsteps) issue(s);

  const int aidx = rbase * TK + 4 * ty;  // float index of the thread's first row in the swizzled tile (see exact_stage_fma)
  for (int s = 0; s < nsteps; ++s) {
    const int st = s % CE_STAGES;
    mbar_wait(&sm.full[st], (uint32_t)((s / CE_STAGES) & 1));  // this stage's copies have landed
    __syncthreads();  // every warp is done with step s - 1: its slot may be refilled
    if (s + CE_STAGES - 1 < nsteps) issue((s + CE_STAGES - 1) % CE_STAGES);
    exact_stage_fma<TN, TK, true, true, CN>(&sm.a[st][0], &sm.b[st][0][0], aidx, cbase, acc, acc2);
  }
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int c = 0; c < CP; ++c) unpack2(acc2[i][c], acc[i][2 * c], acc[i][2 * c + 1]);

  // ---- epilogue: bias -> activation -> FiLM -> residual -> store
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int64_t t = tile_base + rbase + 4 * i;  // tile row -> output row
    if (t >= p.n_out) continue;
    const int64_t o = p.order ? (int64_t)__ldg(p.order + t) : t;
    int64_t frow = o, rrow = o;
    if (p.film_c && p.film_rows) frow = __ldg(p.film_rows + o);
    if (p.residual && p.res_rows) rrow = __ldg(p.res_rows + o);
    if constexpr (CN % 4 == 0) {
      if (p.vec4) {
#pragma unroll
        for (int q = 0; q < CN / 4; ++q) {
          const int col = co_base + cbase + 32 * q;
          if (col >= p.cout) continue;
          float v[4] = {acc[i][4 * q], acc[i][4 * q + 1], acc[i][4 * q + 2], acc[i][4 * q + 3]};
          if (p.last)
            conv_epilogue4(p, o, frow, rrow, col, v);
          else
            *reinterpret_cast<float4*>(p.out + o * p.out_stride + col) = make_float4(v[0], v[1], v[2], v[3]);
        }
        continue;
      }
    }
#pragma unroll
    for (int c = 0; c < CN; ++c) {
      const int col = co_base + cbase + ((CN == 8) ? ((c >> 2) * 32 + (c & 3)) : c);
      if (col >= p.cout) continue;
      if (p.last)
        conv_epilogue(p, o, frow, rrow, col, acc[i][c]);
      else
        p.out[o * p.out_stride + col] = acc[i][c];
    }
  }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn exact_encode_tiled() {
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

// fp32 matrix [rows x cols] row-major, box = box_rows x box_cols, optional SWIZZLE_128B, out-of-bounds elements read as zero
static int exact_map_2d(CUtensorMap* m, const float* ptr, int64_t rows, int64_t cols, int box_rows, int box_cols,
                        bool sw128) {
  EncodeTiledFn enc = exact_encode_tiled();
  if (!enc) return PCC_ERR_CUDA;
  const cuuint64_t gdim[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  const cuuint64_t gstr[1] = {(cuuint64_t)cols * 4};
  const cuuint32_t box[2] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows};
  const cuuint32_t estr[2] = {1u, 1u};
  const CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(ptr), gdim, gstr, box, estr,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, sw128 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE,
                         CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? PCC_OK : PCC_ERR_CUDA;
}

// -> PCC_OK when launched; 1 when the tensor maps cannot describe this call (the caller then uses another producer)
template <int TN, int CN>
static int launch_exact_tma(const ConvParams& p, int64_t n_in, cudaStream_t st) {
  if (n_in <= 0 || n_in >= (1ll << 31) || (p.cout * 4) % 16 != 0 || (reinterpret_cast<uintptr_t>(p.in) & 15) != 0 ||
      (reinterpret_cast<uintptr_t>(p.weight) & 15) != 0)
    return 1;
  CUtensorMap tm_a, tm_b;
  if (exact_map_2d(&tm_a, p.in, n_in, p.cin, 1, CE_TK, true) != PCC_OK) return 1;
  if (exact_map_2d(&tm_b, p.weight, (int64_t)p.K * p.cin, p.cout, CE_TK, TN, false) != PCC_OK) return 1;
  const size_t smem = sizeof(TmaSmem<TN>) + 1024;  // + slack for the 1024-byte alignment of the tile
  dim3 grid((unsigned)ceil_div(p.n_out, CE_TM), (unsigned)ceil_div(p.cout, TN));
  static const cudaError_t attr = cudaFuncSetAttribute(conv_exact_tma_kernel<TN, CN>,
                                                       cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  PCC_CHECK_CUDA(attr);
  conv_exact_tma_kernel<TN, CN><<<grid, 16 * TN / CN, smem, st>>>(tm_a, tm_b, p);
  PCC_CHECK_LAUNCH("conv_exact_tma_kernel");
  return PCC_OK;
}

// PCC_EXACT_TMA: 0 = never, 1 = 128-wide tiles only, 2 = 128- and 64-wide tiles (default).  Measured on the
// 799,458-point frame, same box: dominant-kernel fraction 0.590 (off) / 0.610 (1) / 0.624 (2), 4.40 / 4.54 / 4.62 M
// points/s; every >= 64-channel layer gains 4-7 %.
static int exact_tma() {
  static int v = [] {
    const char* e = getenv("PCC_EXACT_TMA");
    return e ? atoi(e) : 2;
  }();
  return v;
}

// ---------------------------------------------------------------------------------------------
// Narrow-output variant (cout <= 4: occupancy / quality / colour heads).  With so few output
// channels the register-tile kernel above spends its time on padded columns; here one thread owns
// one output row and all (<= 4) channels, so the arithmetic is negligible and the kernel runs at
// the speed the gathered rows arrive (L2 -> shared memory, cp.async, CH_STAGES deep, 4 CTAs / SM).
// Same summation order (k ascending, ci ascending, one FMA chain per output element).
constexpr int CH_STAGES = 3;

struct HeadSmem {
  float a[CH_STAGES][CE_TM][CE_APITCH];
  float b[CH_STAGES][CE_TK][4];
};

__global__ void __launch_bounds__(CE_THREADS, 4)
conv_head_kernel(const ConvParams p) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  HeadSmem& sm = *reinterpret_cast<HeadSmem*>(smem_raw);
  const int tid = threadIdx.x;
  const unsigned tile = p.tile_order ? (unsigned)__ldg(p.tile_order + blockIdx.x) : blockIdx.x;
  const int64_t tile_base = (int64_t)tile * CE_TM;
  const int nchunks = (p.cin + CE_TK - 1) / CE_TK;
  const uint32_t mask = p.tapmask ? (p.tapmask[tile] & p.kbits) : p.kbits;
  const int nsteps = __popc(mask) * nchunks;
  if (nsteps == 0 && !p.first && !p.last) return;

  uint32_t pmask = mask;
  int pk = pmask ? (__ffs(pmask) - 1) : 0;
  int pchunk = 0;
  int idx[4], nidx[4];
  auto load_idx = [&](int k, int (&dst)[4]) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int64_t t = tile_base + (tid >> 3) + 32 * j;
      dst[j] = (t < p.n_out) ? __ldg(p.nbr + (int64_t)k * p.nbr_stride + t) : -1;
    }
  };
  {
    load_idx(pk, idx);
    const uint32_t rest = pmask & (pmask - 1);
    if (rest) load_idx(__ffs(rest) - 1, nidx);
  }

  auto issue = [&](int stage) {
    const int ci0 = pchunk * CE_TK;
    const int q = tid & 7;
    const int ci = ci0 + q * 4;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int r = (tid >> 3) + 32 * j;
      const int ix = idx[j];
      const bool ok = ix >= 0 && ci < p.cin;
      cp_async16(&sm.a[stage][r][q * 4], ok ? p.in + (int64_t)ix * p.cin + ci : p.in, ok ? 16 : 0);
    }
    if (tid < CE_TK) {  // weight rows ci0 + tid of offset pk, padded to 4 columns
      const int wc = ci0 + tid;
      float w[4] = {0.f, 0.f, 0.f, 0.f};
      if (wc < p.cin) {
        const float* wr = p.weight + ((int64_t)pk * p.cin + wc) * p.cout;
#pragma unroll
        for (int c = 0; c < 4; ++c)
          if (c < p.cout) w[c] = __ldg(wr + c);
      }
      *reinterpret_cast<float4*>(&sm.b[stage][tid][0]) = make_float4(w[0], w[1], w[2], w[3]);
    }
    if (++pchunk == nchunks) {
      pchunk = 0;
      pmask &= pmask - 1;
      pk = pmask ? (__ffs(pmask) - 1) : 0;
#pragma unroll
      for (int j = 0; j < 4; ++j) idx[j] = nidx[j];
      const uint32_t rest = pmask & (pmask - 1);
      if (rest) load_idx(__ffs(rest) - 1, nidx);
    }
  };

#pragma unroll
  for (int s = 0; s < CH_STAGES - 1; ++s) {
    if (s < nsteps) issue(s);
    cp_async_commit();
  }

  float acc[4] = {0.f, 0.f, 0.f, 0.f};
  if (!p.first && tid < CE_TM && tile_base + tid < p.n_out) {
    const int64_t t = tile_base + tid;
    const int64_t o = p.order ? (int64_t)__ldg(p.order + t) : t;
#pragma unroll
    for (int c = 0; c < 4; ++c)
      if (c < p.cout) acc[c] = p.out[o * p.out_stride + c];
  }
  for (int s = 0; s < nsteps; ++s) {
    cp_async_wait<CH_STAGES - 2>();
    __syncthreads();
    if (s + CH_STAGES - 1 < nsteps) issue((s + CH_STAGES - 1) % CH_STAGES);
    cp_async_commit();
    const int st = s % CH_STAGES;
    if (tid < CE_TM) {
#pragma unroll
      for (int kv = 0; kv < CE_TK / 4; ++kv) {
        const float4 a = *reinterpret_cast<const float4*>(&sm.a[st][tid][kv * 4]);
        const float av[4] = {a.x, a.y, a.z, a.w};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const float4 w = *reinterpret_cast<const float4*>(&sm.b[st][kv * 4 + e][0]);
          acc[0] = __fmaf_rn(av[e], w.x, acc[0]);
          acc[1] = __fmaf_rn(av[e], w.y, acc[1]);
          acc[2] = __fmaf_rn(av[e], w.z, acc[2]);
          acc[3] = __fmaf_rn(av[e], w.w, acc[3]);
        }
      }
    }
  }
  cp_async_wait<0>();

  const int64_t t = tile_base + tid;
  if (tid >= CE_TM || t >= p.n_out) return;
  const int64_t o = p.order ? (int64_t)__ldg(p.order + t) : t;
  int64_t frow = o, rrow = o;
  if (p.film_c && p.film_rows) frow = __ldg(p.film_rows + o);
  if (p.residual && p.res_rows) rrow = __ldg(p.res_rows + o);
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    if (c >= p.cout) continue;
    if (p.last)
      conv_epilogue(p, o, frow, rrow, c, acc[c]);
    else
      p.out[o * p.out_stride + c] = acc[c];
  }
}

static int launch_head(const ConvParams& p, cudaStream_t st) {
  const size_t smem = sizeof(HeadSmem);
  static const cudaError_t attr =
      cudaFuncSetAttribute(conv_head_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  PCC_CHECK_CUDA(attr);
  conv_head_kernel<<<(unsigned)ceil_div(p.n_out, CE_TM), CE_THREADS, smem, st>>>(p);
  PCC_CHECK_LAUNCH("conv_head_kernel");
  return PCC_OK;
}

// ---------------------------------------------------------------------------------------------
// Thin-input variant (cin <= 4, 27 offsets: the quality-map branch 2 -> 2 / 2 -> 16 / 2 -> 128, the condition
// convolutions 4 -> 2 and the xyz/rgb input layer 4 -> 64).  A neighbour's whole feature row is 8 or 16 bytes, so
// the tiled kernels spend their time staging 27 nearly empty shared-memory stages per tile.  Here one thread owns
// one output row: it reads its 27 kernel-map entries (coalesced) and the <= 27 neighbour rows (one 8 / 16-byte
// load each) into REGISTERS once, then walks the output channels CC at a time with the layer's weights broadcast
// from shared memory (loaded once per CTA; CTAs stride over the tiles).  Finished columns go through a shared
// staging tile so that the epilogue (bias, activation, FiLM, residual) and the stores run row-contiguous.
// Same summation order as everywhere (k ascending, ci ascending, one FMA chain per output element; a missing
// neighbour contributes fma(0, w, acc) like a zero-filled tile row), so results are bit-identical.
template <int CIN, int CC>
__global__ void __launch_bounds__(CE_TM, CIN * 27 + CC + 56 <= 128 ? 4 : (CIN * 27 + CC + 56 <= 184 ? 3 : 2))
conv_thin_kernel(const ConvParams p, int tiles) {
  constexpr int K = 27;
  constexpr int CP = CC / 2;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int pitch = (p.cout + CC - 1) / CC * CC;               // weight row pitch: whole column passes
  float* sw = reinterpret_cast<float*>(smem_raw);              // [K * CIN][pitch]
  float* stg = sw + K * CIN * pitch;                           // [CE_TM][CC + 1]
  int* srow = reinterpret_cast<int*>(stg + CE_TM * (CC + 1));  // [3][CE_TM]: output row, FiLM row, residual row
  const int tid = threadIdx.x;
  for (int i = tid; i < K * CIN * pitch; i += CE_TM) {
    const int r = i / pitch, c = i - r * pitch;
    sw[i] = c < p.cout ? __ldg(p.weight + (int64_t)r * p.cout + c) : 0.0f;
  }
  for (int b = blockIdx.x; b < tiles; b += gridDim.x) {
    const unsigned tile = p.tile_order ? (unsigned)__ldg(p.tile_order + b) : (unsigned)b;
    const int64_t tile_base = (int64_t)tile * CE_TM;
    const uint32_t mask = p.tapmask ? (p.tapmask[tile] & p.kbits) : p.kbits;
    if (mask == 0 && !p.first && !p.last) continue;  // CTA-uniform
    const int64_t t = tile_base + tid;
    const bool valid = t < p.n_out;
    const int64_t o = valid ? (p.order ? (int64_t)__ldg(p.order + t) : t) : -1;
    __syncthreads();  // weights visible (first tile) / the previous tile's staging reads are done
    {
      int frow = (int)o, rrow = (int)o;
      if (valid && p.film_c && p.film_rows) frow = __ldg(p.film_rows + o);
      if (valid && p.residual && p.res_rows) rrow = __ldg(p.res_rows + o);
      srow[tid] = (int)o;
      srow[CE_TM + tid] = frow;
      srow[2 * CE_TM + tid] = rrow;
    }
    // the tile rows are mask-sorted, so the 32 rows of a warp mostly share their neighbour set: offsets that no
    // row of the warp has are skipped by the whole warp (wmask), which leaves ~45 % of the 27 on a surface
    float x[K][CIN];
    uint32_t wmask = 0;
    {
      // three straight-line phases (no branch on a loaded value), so the 27 index loads and then the <= 27 row
      // loads are all in flight together instead of one round trip per offset
      int ixs[K];
#pragma unroll
      for (int k = 0; k < K; ++k)
        ixs[k] = (valid && ((mask >> k) & 1u)) ? __ldg(p.nbr + (int64_t)k * p.nbr_stride + t) : -1;
#pragma unroll
      for (int k = 0; k < K; ++k)
        if (__ballot_sync(0xffffffffu, ixs[k] >= 0)) wmask |= 1u << k;
#pragma unroll
      for (int k = 0; k < K; ++k) {
        const bool has = ixs[k] >= 0;
        const float* src = p.in + (int64_t)(has ? ixs[k] : 0) * CIN;
        if constexpr (CIN == 4) {
          float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
          if (has) v = __ldg(reinterpret_cast<const float4*>(src));
          x[k][0] = v.x; x[k][1] = v.y; x[k][2] = v.z; x[k][3] = v.w;
        } else if constexpr (CIN == 2) {
          float2 v = make_float2(0.f, 0.f);
          if (has) v = __ldg(reinterpret_cast<const float2*>(src));
          x[k][0] = v.x; x[k][1] = v.y;
        } else {
#pragma unroll
          for (int ci = 0; ci < CIN; ++ci) {
            float v = 0.f;
            if (has) v = __ldg(src + ci);
            x[k][ci] = v;
          }
        }
      }
    }
    for (int c0 = 0; c0 < p.cout; c0 += CC) {
      float acc[CC];
#pragma unroll
      for (int j = 0; j < CC; ++j) acc[j] = 0.0f;
      if (!p.first && valid) {  // continue the chains of the previous offset group
#pragma unroll
        for (int j = 0; j < CC; ++j)
          if (c0 + j < p.cout) acc[j] = p.out[o * p.out_stride + c0 + j];
      }
      unsigned long long acc2[CP];
#pragma unroll
      for (int j = 0; j < CP; ++j) acc2[j] = pack2(acc[2 * j], acc[2 * j + 1]);
#pragma unroll
      for (int k = 0; k < K; ++k) {
        if (!((wmask >> k) & 1u)) continue;  // warp-uniform: every lane would add fma(0, w, acc) = acc
#pragma unroll
        for (int ci = 0; ci < CIN; ++ci) {
          const float* wr = sw + (k * CIN + ci) * pitch + c0;
          const unsigned long long xx = pack2(x[k][ci], x[k][ci]);
#pragma unroll
          for (int j4 = 0; j4 < CC / 4; ++j4) {
            const float4 w = *reinterpret_cast<const float4*>(wr + 4 * j4);
            ffma2(acc2[2 * j4], xx, pack2(w.x, w.y));
            ffma2(acc2[2 * j4 + 1], xx, pack2(w.z, w.w));
          }
        }
      }
#pragma unroll
      for (int j = 0; j < CP; ++j) unpack2(acc2[j], acc[2 * j], acc[2 * j + 1]);
      if (c0) __syncthreads();  // the previous pass's staging reads are done
#pragma unroll
      for (int j = 0; j < CC; ++j) stg[tid * (CC + 1) + j] = acc[j];
      __syncthreads();
      const int ncol = p.cout - c0 < CC ? p.cout - c0 : CC;
      for (int e = tid; e < CE_TM * CC; e += CE_TM) {
        const int r = e / CC, c = e % CC;  // CC = 32: one warp stores 128 contiguous bytes of one row
        const int64_t orow = srow[r];
        if (c >= ncol || orow < 0) continue;
        const float v = stg[r * (CC + 1) + c];
        if (p.last)
          conv_epilogue(p, orow, srow[CE_TM + r], srow[2 * CE_TM + r], c0 + c, v);
        else
          p.out[orow * p.out_stride + c0 + c] = v;
      }
    }
  }
}

// Thin input, wide output (cin <= 4, cout > 16: the first q_predict layer 2 -> 128, the xyz/rgb input layer
// 4 -> 64).  One CTA = 128 output rows x 128 output channels with the 8 x 8 register tile of the main kernel, but
// WITHOUT its 27 nearly empty pipeline stages: the tile's whole gathered operand, 27 cin values per row, is
// staged once (transposed, [27 cin][128 rows]), the layer's weights once per CTA ([27 cin][128]), and the 27 cin
// rank-1 updates run back to back from shared memory (4 LDS.128 per 32 FFMA2, no barrier in between).  A warp
// (32 rows x 64 columns) skips the offsets none of its rows has.  Same FMA chain per output element.
template <int CIN>
__global__ void __launch_bounds__(CE_THREADS, 2)
conv_thin_wide_kernel(const ConvParams p, int tiles) {
  constexpr int K = 27, KC = K * CIN, KH = 14;  // KH: offsets gathered per thread (two threads per row)
  extern __shared__ __align__(16) unsigned char smem_raw[];
  float* sw = reinterpret_cast<float*>(smem_raw);              // [KC][128] columns co_base .. + 127
  float* xs = sw + KC * CE_TM;                                 // [KC][128 rows]
  uint32_t* bmask = reinterpret_cast<uint32_t*>(xs + KC * CE_TM);  // [4]: offsets present in each 32-row block
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const int co_base = blockIdx.y * 128;
  for (int i = tid; i < KC * 128; i += CE_THREADS) {
    const int r = i >> 7, c = i & 127;
    sw[i] = (co_base + c < p.cout) ? __ldg(p.weight + (int64_t)r * p.cout + co_base + c) : 0.0f;
  }
  const int tx = lane & 7, ty = lane >> 3;
  const int wr = wid >> 1, wc = wid & 1;  // warp: tile rows 32 wr .. + 31, columns 64 wc .. + 63
  const int r0 = wr * 32 + ty * 8;        // this thread's 8 consecutive tile rows
  const int c0 = wc * 64 + tx * 4;        // columns c0 .. c0 + 3 and c0 + 32 .. c0 + 35
  for (int b = blockIdx.x; b < tiles; b += gridDim.x) {
    const unsigned tile = p.tile_order ? (unsigned)__ldg(p.tile_order + b) : (unsigned)b;
    const int64_t tile_base = (int64_t)tile * CE_TM;
    const uint32_t mask = p.tapmask ? (p.tapmask[tile] & p.kbits) : p.kbits;
    if (mask == 0 && !p.first && !p.last) continue;  // CTA-uniform
    __syncthreads();  // the previous tile's reads of xs / bmask are done
    if (tid < 4) bmask[tid] = 0u;
    __syncthreads();
    {
      const int row = tid & (CE_TM - 1), half = tid >> 7;
      const int64_t t = tile_base + row;
      const bool valid = t < p.n_out;
      int ixs[KH];
#pragma unroll
      for (int j = 0; j < KH; ++j) {
        const int k = half * KH + j;
        ixs[j] = (k < K && valid && ((mask >> k) & 1u)) ? __ldg(p.nbr + (int64_t)k * p.nbr_stride + t) : -1;
      }
      uint32_t wm = 0;
#pragma unroll
      for (int j = 0; j < KH; ++j)
        if (__ballot_sync(0xffffffffu, ixs[j] >= 0)) wm |= 1u << ((half * KH + j) & 31);
      if (lane == 0 && wm) atomicOr(&bmask[row >> 5], wm);
#pragma unroll
      for (int j = 0; j < KH; ++j) {
        const int k = half * KH + j;
        if (k >= K) continue;
        const bool has = ixs[j] >= 0;
        const float* src = p.in + (int64_t)(has ? ixs[j] : 0) * CIN;
        float v[CIN];
        if constexpr (CIN == 4) {
          float4 q = make_float4(0.f, 0.f, 0.f, 0.f);
          if (has) q = __ldg(reinterpret_cast<const float4*>(src));
          v[0] = q.x; v[1] = q.y; v[2] = q.z; v[3] = q.w;
        } else if constexpr (CIN == 2) {
          float2 q = make_float2(0.f, 0.f);
          if (has) q = __ldg(reinterpret_cast<const float2*>(src));
          v[0] = q.x; v[1] = q.y;
        } else {
#pragma unroll
          for (int ci = 0; ci < CIN; ++ci) {
            v[ci] = 0.f;
            if (has) v[ci] = __ldg(src + ci);
          }
        }
#pragma unroll
        for (int ci = 0; ci < CIN; ++ci) xs[(k * CIN + ci) * CE_TM + row] = v[ci];
      }
    }
    __syncthreads();
    if (co_base + wc * 64 >= p.cout) continue;  // a warp whose 64 columns lie beyond cout (cout <= 64) only gathers
    const uint32_t wmask = bmask[wr];

    float acc[8][8];
    unsigned long long acc2[8][4];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
      for (int c = 0; c < 8; ++c) acc[i][c] = 0.0f;
    if (!p.first) {  // continue the chains of the previous offset group
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const int64_t t = tile_base + r0 + i;
        if (t >= p.n_out) continue;
        const int64_t o = p.order ? (int64_t)__ldg(p.order + t) : t;
        if (p.vec4) {
#pragma unroll
          for (int q = 0; q < 2; ++q) {
            const int col = co_base + c0 + 32 * q;
            if (col < p.cout) {
              const float4 v = *reinterpret_cast<const float4*>(p.out + o * p.out_stride + col);
              acc[i][4 * q] = v.x; acc[i][4 * q + 1] = v.y; acc[i][4 * q + 2] = v.z; acc[i][4 * q + 3] = v.w;
            }
          }
          continue;
        }
#pragma unroll
        for (int c = 0; c < 8; ++c) {
          const int col = co_base + c0 + (c >> 2) * 32 + (c & 3);
          if (col < p.cout) acc[i][c] = p.out[o * p.out_stride + col];
        }
      }
    }
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
      for (int c = 0; c < 4; ++c) acc2[i][c] = pack2(acc[i][2 * c], acc[i][2 * c + 1]);
#pragma unroll 1  // a rolled loop: 27 cin unrolled bodies of 32 FFMA2 would not fit the instruction cache
    for (int k = 0; k < K; ++k) {
      if (!((wmask >> k) & 1u)) continue;  // warp-uniform: every lane would add fma(0, w, acc) = acc
#pragma unroll
      for (int ci = 0; ci < CIN; ++ci) {
        const float* xr = xs + (k * CIN + ci) * CE_TM + r0;
        const float4 xa = *reinterpret_cast<const float4*>(xr);
        const float4 xb = *reinterpret_cast<const float4*>(xr + 4);
        const float* wrow = sw + (k * CIN + ci) * 128 + c0;
        const float4 w0 = *reinterpret_cast<const float4*>(wrow);
        const float4 w1 = *reinterpret_cast<const float4*>(wrow + 32);
        const unsigned long long bp[4] = {pack2(w0.x, w0.y), pack2(w0.z, w0.w), pack2(w1.x, w1.y), pack2(w1.z, w1.w)};
        const float xv[8] = {xa.x, xa.y, xa.z, xa.w, xb.x, xb.y, xb.z, xb.w};
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const unsigned long long aa = pack2(xv[i], xv[i]);
#pragma unroll
          for (int c = 0; c < 4; ++c) ffma2(acc2[i][c], aa, bp[c]);
        }
      }
    }
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
      for (int c = 0; c < 4; ++c) unpack2(acc2[i][c], acc[i][2 * c], acc[i][2 * c + 1]);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int64_t t = tile_base + r0 + i;
      if (t >= p.n_out) continue;
      const int64_t o = p.order ? (int64_t)__ldg(p.order + t) : t;
      int64_t frow = o, rrow = o;
      if (p.film_c && p.film_rows) frow = __ldg(p.film_rows + o);
      if (p.residual && p.res_rows) rrow = __ldg(p.res_rows + o);
      if (p.vec4) {
#pragma unroll
        for (int q = 0; q < 2; ++q) {
          const int col = co_base + c0 + 32 * q;
          if (col >= p.cout) continue;
          float v[4] = {acc[i][4 * q], acc[i][4 * q + 1], acc[i][4 * q + 2], acc[i][4 * q + 3]};
          if (p.last)
            conv_epilogue4(p, o, frow, rrow, col, v);
          else
            *reinterpret_cast<float4*>(p.out + o * p.out_stride + col) = make_float4(v[0], v[1], v[2], v[3]);
        }
        continue;
      }
#pragma unroll
      for (int c = 0; c < 8; ++c) {
        const int col = co_base + c0 + (c >> 2) * 32 + (c & 3);
        if (col >= p.cout) continue;
        if (p.last)
          conv_epilogue(p, o, frow, rrow, col, acc[i][c]);
        else
          p.out[o * p.out_stride + col] = acc[i][c];
      }
    }
  }
}

template <int CIN>
static int launch_thin_wide(const ConvParams& p, cudaStream_t st) {
  const size_t smem = (size_t)2 * 27 * CIN * CE_TM * 4 + 16;
  const int tiles = (int)ceil_div(p.n_out, CE_TM);
  const int gx = tiles < 2 * sm_count() ? tiles : 2 * sm_count();
  static const cudaError_t attr =
      cudaFuncSetAttribute(conv_thin_wide_kernel<CIN>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  PCC_CHECK_CUDA(attr);
  conv_thin_wide_kernel<CIN><<<dim3((unsigned)gx, (unsigned)ceil_div(p.cout, 128)), CE_THREADS, smem, st>>>(p, tiles);
  PCC_CHECK_LAUNCH("conv_thin_wide_kernel");
  return PCC_OK;
}

// -> PCC_OK when launched, 1 when this call is not a thin-input shape (the caller falls through)
template <int CIN>
static int launch_thin_cin(const ConvParams& p, cudaStream_t st) {
  const int cc = (p.cout > 16 && CIN <= 2) ? 32 : (p.cout > 4 ? 16 : 4);  // x[27][CIN] + CC accumulators in registers
  const int pitch = (p.cout + cc - 1) / cc * cc;
  const size_t smem = (size_t)27 * CIN * pitch * 4 + (size_t)CE_TM * (cc + 1) * 4 + 3 * CE_TM * 4;
  if (smem > 96 * 1024) return 1;
  const int tiles = (int)ceil_div(p.n_out, CE_TM);
  const int grid = tiles < 4 * sm_count() ? tiles : 4 * sm_count();
#define PCC_THIN_LAUNCH(CCV)                                                                                     \
  {                                                                                                              \
    static const cudaError_t attr = cudaFuncSetAttribute(conv_thin_kernel<CIN, CCV>,                             \
                                                         cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024); \
    PCC_CHECK_CUDA(attr);                                                                                        \
    conv_thin_kernel<CIN, CCV><<<grid, CE_TM, smem, st>>>(p, tiles);                                             \
  }
  if (cc == 32) PCC_THIN_LAUNCH(32) else if (cc == 16) PCC_THIN_LAUNCH(16) else PCC_THIN_LAUNCH(4)
#undef PCC_THIN_LAUNCH
  PCC_CHECK_LAUNCH("conv_thin_kernel");
  return PCC_OK;
}

// PCC_EXACT_THIN=0 keeps the tiled kernel with 4-channel stages for cin <= 4 (A/B timing)
static int launch_thin(const ConvParams& p, cudaStream_t st) {
  static const int on = [] {
    const char* e = getenv("PCC_EXACT_THIN");
    return e ? atoi(e) : 1;
  }();
  if (!on || p.K != 27 || p.cin < 1 || p.cin > 4 || p.n_out >= (1ll << 31) - CE_TM) return 1;
  if (p.cout > 16 && on < 2) {  // PCC_EXACT_THIN=2: the row-per-thread kernel for every width (A/B timing)
    const uintptr_t ai = reinterpret_cast<uintptr_t>(p.in);
    switch (p.cin) {
      case 1: return launch_thin_wide<1>(p, st);
      case 2: return (ai & 7) ? 1 : launch_thin_wide<2>(p, st);
      case 3: return launch_thin_wide<3>(p, st);
      default: return (ai & 15) ? 1 : launch_thin_wide<4>(p, st);
    }
  }
  const uintptr_t a = reinterpret_cast<uintptr_t>(p.in);
  switch (p.cin) {
    case 1: return launch_thin_cin<1>(p, st);
    case 2: return (a & 7) ? 1 : launch_thin_cin<2>(p, st);
    case 3: return launch_thin_cin<3>(p, st);
    default: return (a & 15) ? 1 : launch_thin_cin<4>(p, st);
  }
}

template <int TN, bool VEC, int F2, int TK = CE_TK>
static int launch_exact_v(const ConvParams& p, cudaStream_t st) {
  const size_t smem = sizeof(ExactSmem<TN, TK>);
  dim3 grid((unsigned)ceil_div(p.n_out, CE_TM), (unsigned)ceil_div(p.cout, TN));
  // once per instantiation and process (thread-safe static init): keeps driver calls off the launch path
  static const cudaError_t attr = cudaFuncSetAttribute(conv_exact_kernel<TN, VEC, F2, TK>,
                                                       cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  PCC_CHECK_CUDA(attr);
  conv_exact_kernel<TN, VEC, F2, TK><<<grid, CE_THREADS, smem, st>>>(p);
  PCC_CHECK_LAUNCH("conv_exact_kernel");
  return PCC_OK;
}

static int exact_f2() {
  static int f2 = [] {
    const char* e = getenv("PCC_EXACT_F2");
    return e ? atoi(e) : 1;
  }();
  return f2;
}

template <int TN>
static int launch_exact(const ConvParams& p, cudaStream_t st) {
  const bool vec = (p.cin % 4 == 0) && (p.cout % 4 == 0) &&
                   ((reinterpret_cast<uintptr_t>(p.in) & 15) == 0) &&
                   ((reinterpret_cast<uintptr_t>(p.weight) & 15) == 0);
  const int f2 = exact_f2();
  // measured on the 799,458-point frame (profiles/r02_*): the bulk-copy producer is 2 % faster on the 128-wide
  // tiles (1.19 M x 128 -> 128: 14.44 -> 14.18 ms) and 7 % slower on the 64-wide ones (twice the copies per FMA), so
  // only the 128-wide kernel uses it; PCC_EXACT_BULK=2 forces it for 64- and 32-wide tiles too
  if constexpr (TN >= 64) {
    const int tma = exact_tma();
    if (vec && f2 != 0 && p.cin % CE_TK == 0 && (tma >= 2 || (tma == 1 && TN == 128))) {
      // 64-wide tiles: 128 threads with 8 x 8 register tiles (3 CTAs per SM) for maps of >= 262,144 rows (with the
      // 16-byte epilogue: 3 x 64->64 on the 3.98 M candidates 36.7 -> 33.6 ms, 64->64 on 799,458 rows 1.80 -> 1.68,
      // the generative 128->64 2.45 -> 2.33); 256 threads with 8 x 4 tiles on the smaller levels, which cannot
      // fill 3 x 148 four-warp CTAs (17,957 rows: 3.63 vs 3.73 ms).  PCC_EXACT_TMA64 = 8 / 4 forces one shape.
      static const int cn64 = [] {
        const char* e = getenv("PCC_EXACT_TMA64");
        return e ? atoi(e) : 0;
      }();
      int rc;
      if constexpr (TN == 64)
        rc = (cn64 == 8 || (cn64 == 0 && p.n_out >= (1ll << 18)))
                 ? launch_exact_tma<64, 8>(p, p.n_in, st)
                 : launch_exact_tma<64, 4>(p, p.n_in, st);
      else
        rc = launch_exact_tma<TN, TN / 16>(p, p.n_in, st);
      if (rc <= 0) return rc;  // launched, or a CUDA error; 1 = tensor maps not applicable -> next producer
    }
  }
  if constexpr (TN >= 32) {
    const int bulk = exact_bulk();
    if (vec && f2 != 0 && p.cin % CE_TK == 0 && (bulk >= 2 || (bulk == 1 && TN == 128)))
      return launch_exact_bulk<TN>(p, st);
  }
  if (p.cin <= 4 && f2 != 0) {  // thin input (the Q branch, the xyz/rgb input layer)
    const int rc = launch_thin(p, st);
    if (rc <= 0) return rc;  // launched or failed; 1 = not a thin shape -> tiled kernel with 4-channel stages
  }
  if (p.cin <= 4 && f2 != 0)
    return vec ? launch_exact_v<TN, true, 1, 4>(p, st) : launch_exact_v<TN, false, 1, 4>(p, st);
  if (f2 != 0) return vec ? launch_exact_v<TN, true, 1>(p, st) : launch_exact_v<TN, false, 1>(p, st);
  return vec ? launch_exact_v<TN, true, 0>(p, st) : launch_exact_v<TN, false, 0>(p, st);
}

static int conv_exact_group(const pcc_conv_args* a, const ConvParams& p, cudaStream_t st) {
  if (a->cout <= 4 && a->cin % 4 == 0 && a->cin >= 16 && (reinterpret_cast<uintptr_t>(p.in) & 15) == 0)
    return launch_head(p, st);
  // a grid that cannot fill 2 CTAs on every SM is split further along the output channels
  const int64_t tiles = ceil_div(p.n_out, CE_TM);
  const bool small_grid = tiles * ceil_div(p.cout, 128) < 2 * (int64_t)sm_count();
  // PCC_EXACT_WIDE=0: cover wide layers with 64-wide tiles too (A/B of the 128-thread 8 x 8 kernel)
  static const int wide = [] {
    const char* e = getenv("PCC_EXACT_WIDE");
    return e ? atoi(e) : 1;
  }();
  if (a->cout > 64 && !small_grid && wide) return launch_exact<128>(p, st);
  if (a->cout > 32) return launch_exact<64>(p, st);
  if (a->cout > 16) return launch_exact<32>(p, st);
  return launch_exact<16>(p, st);
}

int conv_exact(const pcc_conv_args* a, cudaStream_t st) {
  ConvParams p;
  p.in = a->in; p.weight = a->weight; p.bias = a->bias; p.nbr = a->nbr; p.tapmask = a->tapmask;
  p.film = a->film; p.film_rows = a->film_rows; p.residual = a->residual; p.res_rows = a->res_rows;
  p.order = a->order; p.out = a->out; p.nbr_stride = a->nbr_stride; p.n_out = a->n_out; p.out_stride = a->out_stride;
  p.n_in = a->n_in;
  p.cin = a->cin; p.cout = a->cout; p.K = a->K; p.act = a->act; p.film_c = a->film_c; p.slope = a->slope; p.flags = a->flags;
  const uint32_t kfull = p.K >= 32 ? 0xffffffffu : ((1u << p.K) - 1u);
  p.tile_order = a->tile_order;
  {
    auto al16 = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15) == 0; };
    p.vec4 = p.cout % 4 == 0 && p.out_stride % 4 == 0 && al16(p.out) && al16(p.bias) && al16(p.residual) &&
             (!p.film_c || (al16(p.film) && (p.film_c == 1 || p.film_c % 4 == 0)));
  }
  const int ng = a->n_groups > 1 ? a->n_groups : 1;
  if (ng == 1) {
    p.kbits = kfull; p.first = 1; p.last = 1;
    return conv_exact_group(a, p, st);
  }
  // one launch per offset group: group g's tile rows follow order[g], its tap masks tapmask[g]
  const int64_t tiles = ceil_div(p.n_out, CE_TM);
  const int per = (p.K + ng - 1) / ng;
  for (int g = 0; g < ng; ++g) {
    const int k_lo = g * per, k_hi = (g + 1) * per < p.K ? (g + 1) * per : p.K;
    p.kbits = (((k_hi >= 32) ? 0xffffffffu : ((1u << k_hi) - 1u)) & ~((1u << k_lo) - 1u)) & kfull;
    p.first = g == 0; p.last = g == ng - 1;
    p.order = a->order + (int64_t)g * p.n_out;
    p.tapmask = a->tapmask ? a->tapmask + (int64_t)g * tiles : nullptr;
    p.tile_order = a->tile_order ? a->tile_order + (int64_t)g * tiles : nullptr;
    const int rc = conv_exact_group(a, p, st);
    if (rc) return rc;
  }
  return PCC_OK;
}

}  // namespace pcc
