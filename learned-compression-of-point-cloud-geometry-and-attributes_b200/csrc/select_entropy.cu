// Row gather, top-k radix select (pruning), entropy-stage element-wise kernels, colour
// finalisation.  All HBM-bound; see include/pcc_b200.h for the reference call sites.
#include <math_constants.h>
#include <algorithm>

#include "pcc_common.cuh"

namespace pcc {

constexpr int THREADS = 256;
static inline unsigned grid_for(int64_t n, int threads = THREADS) {
  { const int64_t g = ceil_div(n, threads); return (unsigned)(g < 1 ? 1 : g); }
}

// ---------------------------------------------------------------- gather rows (features_at)
// one thread per float4 (or scalar when c % 4 != 0)
template <bool VEC4>
__global__ void gather_rows_kernel(const float* __restrict__ in, int c, const int32_t* __restrict__ rows,
                                   int64_t n, const float* __restrict__ add, float* __restrict__ out) {
  const int per_row = VEC4 ? c / 4 : c;
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n * per_row) return;
  const int64_t i = t / per_row;
  const int j = (int)(t % per_row);
  const int32_t r = rows ? __ldg(rows + i) : (int32_t)i;
  if (VEC4) {
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (r >= 0) v = __ldg(reinterpret_cast<const float4*>(in + (int64_t)r * c) + j);
    if (add) {
      const float4 a = __ldg(reinterpret_cast<const float4*>(add + i * c) + j);
      v.x = __fadd_rn(a.x, v.x); v.y = __fadd_rn(a.y, v.y);
      v.z = __fadd_rn(a.z, v.z); v.w = __fadd_rn(a.w, v.w);
    }
    reinterpret_cast<float4*>(out + i * c)[j] = v;
  } else {
    float v = r >= 0 ? __ldg(in + (int64_t)r * c + j) : 0.f;
    if (add) v = __fadd_rn(__ldg(add + i * c + j), v);
    out[i * c + j] = v;
  }
}

// ---------------------------------------------------------------- occupancy head, factorised form
// Channel 0 of a convolution whose other output channels nobody reads (the occupancy head,
// model/blocks.py:142): out[o] = b0 + sum_k <in[nbr[k][o]], W[k][:, 0]>.  Gathering 64-128-channel rows
// for ONE output channel makes the direct kernel gather-bound (19 GB of rows per 4 M-candidate
// layer); projecting first, P[i][k] = <in[i], W[k][:, 0]> (one streaming pass, 32 floats per row),
// turns the gather into one 4-byte read per pair.  fp32 throughout, but the sum over ci is taken
// before the sum over k — a different rounding order than the exact path's single (k, ci) FMA
// chain, so this form serves the tolerance mode (PCC_MODE_TF32) only.
constexpr int HP_ROWS = 128;  // rows per block
constexpr int HP_K = 32;      // padded offsets per row of P
constexpr int HP_RPW = HP_ROWS / 8;  // rows per warp

// lane = offset k, warp w = rows 16 w .. 16 w + 15 of the block's 128: per four input channels a thread reads its
// four weights (conflict-free, 32 consecutive floats per channel) and one broadcast float4 per row, for 64 FMAs —
// 0.3 shared-memory loads per FMA (the first version, one thread per output with two loads per FMA, spent 4 ms per
// frame on the three heads; ncu launch list profiles/r02_launch_summary_tf32.txt).  Same ascending-ci FMA chain.
__global__ void __launch_bounds__(256)
head_project_kernel(const float* __restrict__ in, int64_t n, int cin, const float* __restrict__ w, int K, int cout,
                    float* __restrict__ P) {
  extern __shared__ __align__(16) float hp_smem[];
  float* w_s = hp_smem;                 // [cin][HP_K]
  float* x_s = hp_smem + cin * HP_K;    // [HP_ROWS][cin + 4]
  const int pitch = cin + 4;
  const int tid = threadIdx.x;
  for (int e = tid; e < cin * HP_K; e += 256) {
    const int ci = e / HP_K, k = e % HP_K;
    w_s[e] = k < K ? __ldg(w + ((int64_t)k * cin + ci) * cout) : 0.f;
  }
  const int64_t r0 = (int64_t)blockIdx.x * HP_ROWS;
  const int q = cin / 4;  // float4 per row (cin % 4 == 0 is required by the caller)
  for (int e = tid; e < HP_ROWS * q; e += 256) {
    const int r = e / q, c4 = e % q;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (r0 + r < n) v = __ldg(reinterpret_cast<const float4*>(in + (r0 + r) * cin) + c4);
    *reinterpret_cast<float4*>(x_s + r * pitch + 4 * c4) = v;
  }
  __syncthreads();
  const int k = tid & 31, wrp = tid >> 5;
  float acc[HP_RPW];
#pragma unroll
  for (int j = 0; j < HP_RPW; ++j) acc[j] = 0.f;
  const float* xb = x_s + (wrp * HP_RPW) * pitch;
  for (int ci = 0; ci < cin; ci += 4) {
    const float w0 = w_s[(ci + 0) * HP_K + k], w1 = w_s[(ci + 1) * HP_K + k];
    const float w2 = w_s[(ci + 2) * HP_K + k], w3 = w_s[(ci + 3) * HP_K + k];
#pragma unroll
    for (int j = 0; j < HP_RPW; ++j) {
      const float4 x = *reinterpret_cast<const float4*>(xb + j * pitch + ci);
      acc[j] = fmaf(x.x, w0, acc[j]);
      acc[j] = fmaf(x.y, w1, acc[j]);
      acc[j] = fmaf(x.z, w2, acc[j]);
      acc[j] = fmaf(x.w, w3, acc[j]);
    }
  }
#pragma unroll
  for (int j = 0; j < HP_RPW; ++j) {
    const int64_t r = r0 + wrp * HP_RPW + j;
    if (r < n) P[r * HP_K + k] = acc[j];
  }
}

__global__ void head_gather_kernel(const float* __restrict__ P, const int32_t* __restrict__ nbr, int64_t nbr_stride,
                                   const int32_t* __restrict__ order, int64_t n_out, int K,
                                   const float* __restrict__ bias, float* __restrict__ out) {
  const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n_out) return;
  float acc = bias ? __ldg(bias) : 0.f;
  for (int k = 0; k < K; ++k) {
    const int32_t i = __ldg(nbr + (int64_t)k * nbr_stride + j);
    if (i >= 0) acc += __ldg(P + (int64_t)i * HP_K + k);
  }
  out[order ? (int64_t)__ldg(order + j) : j] = acc;
}

// ---------------------------------------------------------------- top-k radix select
__device__ __forceinline__ uint32_t float_key(float f) {
  const uint32_t u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}

struct TopkState {
  uint32_t prefix;     // selected high bits of the threshold key so far
  uint32_t prefix_mask;
  int64_t remaining;   // how many still to take among keys matching the prefix
};

constexpr int TK_BINS = 2048;

// pass 0: bits 31..21, pass 1: bits 20..10, pass 2: bits 9..0
__device__ __forceinline__ int pass_shift(int pass) { return pass == 0 ? 21 : (pass == 1 ? 10 : 0); }
__device__ __forceinline__ int pass_bins(int pass) { return pass == 2 ? 1024 : 2048; }

__global__ void __launch_bounds__(THREADS)
topk_hist_kernel(const float* __restrict__ logits, int64_t stride, int64_t m, int pass,
                 const TopkState* __restrict__ state, uint32_t* __restrict__ hist) {
  __shared__ uint32_t sh[TK_BINS];
  for (int i = threadIdx.x; i < TK_BINS; i += blockDim.x) sh[i] = 0;
  __syncthreads();
  const uint32_t prefix = state->prefix, pmask = state->prefix_mask;
  const int shift = pass_shift(pass);
  const uint32_t bmask = (uint32_t)pass_bins(pass) - 1;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < m;
       i += (int64_t)gridDim.x * blockDim.x) {
    const uint32_t key = float_key(__ldg(logits + i * stride));
    if ((key & pmask) == prefix) atomicAdd(&sh[(key >> shift) & bmask], 1u);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < TK_BINS; i += blockDim.x)
    if (sh[i]) atomicAdd(&hist[i], sh[i]);
}

// One block: the first bin from the top whose cumulative count covers `remaining` becomes the next digit of the
// threshold key.  Thread t owns bins/256 consecutive bins (descending), a block scan gives the count above them;
// exactly one bin satisfies  count_above < remaining <= count_above + h  (the top bin when remaining == 0).
constexpr int TF_THREADS = 256;
__global__ void __launch_bounds__(TF_THREADS)
topk_find_kernel(int pass, TopkState* __restrict__ state, uint32_t* __restrict__ hist) {
  __shared__ long long part[TF_THREADS];
  __shared__ int s_sel;
  __shared__ long long s_cum;
  const int t = threadIdx.x;
  const int bins = pass_bins(pass);
  const int per = bins / TF_THREADS;  // 8 or 4
  const long long rem = state->remaining;
  uint32_t h[TK_BINS / TF_THREADS];
  long long tsum = 0;
#pragma unroll
  for (int j = 0; j < TK_BINS / TF_THREADS; ++j) {
    h[j] = j < per ? hist[bins - 1 - (t * per + j)] : 0u;
    tsum += h[j];
  }
  part[t] = tsum;
  if (t == 0) s_sel = -1;
  __syncthreads();
  for (int off = 1; off < TF_THREADS; off <<= 1) {  // inclusive scan of the per-thread counts
    const long long v = t >= off ? part[t - off] : 0;
    __syncthreads();
    part[t] += v;
    __syncthreads();
  }
  long long cum = part[t] - tsum;  // count in the bins above this thread's
#pragma unroll
  for (int j = 0; j < TK_BINS / TF_THREADS; ++j) {
    if (j < per) {
      const int d = t * per + j;
      const long long incl = cum + h[j];
      if (incl >= rem && (d == 0 || cum < rem)) {
        s_sel = bins - 1 - d;
        s_cum = cum;
      }
      cum = incl;
    }
  }
  __syncthreads();
  if (t == 0) {
    int sel = s_sel;
    long long above = s_cum;
    if (sel < 0) {  // fewer keys than `remaining` (k <= m rules it out): bottom bin, everything counted
      sel = 0;
      above = part[TF_THREADS - 1];
    }
    state->remaining = rem - above;
    const int shift = pass_shift(pass);
    state->prefix |= (uint32_t)sel << shift;
    state->prefix_mask |= ((uint32_t)bins - 1) << shift;
  }
  for (int b = t; b < TK_BINS; b += TF_THREADS) hist[b] = 0;
}

__global__ void topk_eq_flags_kernel(const float* __restrict__ logits, int64_t stride, int64_t m,
                                     const TopkState* __restrict__ state, int32_t* __restrict__ eq) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < m) eq[i] = float_key(__ldg(logits + i * stride)) == state->prefix ? 1 : 0;
}

__global__ void topk_sel_flags_kernel(const float* __restrict__ logits, int64_t stride, int64_t m,
                                      const TopkState* __restrict__ state,
                                      const int32_t* __restrict__ eq_rank, int32_t* __restrict__ sel) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= m) return;
  const uint32_t key = float_key(__ldg(logits + i * stride));
  const uint32_t T = state->prefix;
  sel[i] = (key > T || (key == T && (int64_t)eq_rank[i] < state->remaining)) ? 1 : 0;
}

__global__ void topk_emit_kernel(const float* __restrict__ logits, int64_t stride, int64_t m,
                                 const TopkState* __restrict__ state, const int32_t* __restrict__ eq_rank,
                                 const int32_t* __restrict__ pos, int64_t k, int32_t* __restrict__ sel_rows) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= m) return;
  const uint32_t key = float_key(__ldg(logits + i * stride));
  const uint32_t T = state->prefix;
  if (key > T || (key == T && (int64_t)eq_rank[i] < state->remaining)) {
    const int32_t w = pos[i];
    if (w < k) sel_rows[w] = (int32_t)i;
  }
}

__global__ void topk_init_kernel(TopkState* state, uint32_t* hist, int64_t k) {
  for (int i = threadIdx.x; i < TK_BINS; i += blockDim.x) hist[i] = 0;
  if (threadIdx.x == 0) {
    state->prefix = 0;
    state->prefix_mask = 0;
    state->remaining = k;
  }
}

// ---------------------------------------------------------------- entropy stage
// 32 x 32 transposing tiles: read row-major [N, C] coalesced along c, write channel-major [C, N]
// coalesced along n.
constexpr int ET = 32;

__device__ __forceinline__ int scale_index(float s, const float* __restrict__ table, int n_scales) {
  int idx = n_scales - 1;
  for (int j = 0; j < n_scales - 1; ++j) idx -= (s <= table[j]) ? 1 : 0;
  return idx;
}

// MODE 0: symbols + indexes from y; MODE 1: indexes only
template <int MODE>
__global__ void __launch_bounds__(ET * 8)
gc_quantize_kernel(const float* __restrict__ y, const float* __restrict__ params,
                   const int32_t* __restrict__ rows, int64_t n, int c,
                   const float* __restrict__ scale_table, int n_scales, float scale_bound,
                   int32_t* __restrict__ symbols, int32_t* __restrict__ indexes) {
  __shared__ int32_t t_sym[ET][ET + 1];
  __shared__ int32_t t_idx[ET][ET + 1];
  __shared__ float table[64];
  const int tid = threadIdx.y * ET + threadIdx.x;
  if (tid < n_scales) table[tid] = scale_table[tid];
  __syncthreads();
  const int64_t n0 = (int64_t)blockIdx.x * ET;
  const int c0 = blockIdx.y * ET;
  for (int r = threadIdx.y; r < ET; r += 8) {
    const int64_t i = n0 + r;
    const int ch = c0 + threadIdx.x;
    int32_t sym = 0, idx = 0;
    if (i < n && ch < c) {
      const int32_t pr = rows ? __ldg(rows + i) : (int32_t)i;
      float scale = 0.f, mean = 0.f;
      if (pr >= 0) {
        scale = __ldg(params + (int64_t)pr * 2 * c + ch);
        mean = __ldg(params + (int64_t)pr * 2 * c + c + ch);
      }
      idx = scale_index(fmaxf(scale, scale_bound), table, n_scales);
      if (MODE == 0) sym = (int32_t)rintf(__fsub_rn(__ldg(y + i * c + ch), mean));
    }
    t_sym[r][threadIdx.x] = sym;
    t_idx[r][threadIdx.x] = idx;
  }
  __syncthreads();
  for (int r = threadIdx.y; r < ET; r += 8) {
    const int ch = c0 + r;
    const int64_t i = n0 + threadIdx.x;
    if (i < n && ch < c) {
      if (MODE == 0) symbols[(int64_t)ch * n + i] = t_sym[threadIdx.x][r];
      indexes[(int64_t)ch * n + i] = t_idx[threadIdx.x][r];
    }
  }
}

// symbols [C, N] -> y_hat [N, C] = float(sym) + mean
__global__ void __launch_bounds__(ET * 8)
gc_dequantize_kernel(const int32_t* __restrict__ symbols, const float* __restrict__ params,
                     const int32_t* __restrict__ rows, int64_t n, int c, float* __restrict__ y_hat) {
  __shared__ int32_t t_sym[ET][ET + 1];
  const int64_t n0 = (int64_t)blockIdx.x * ET;
  const int c0 = blockIdx.y * ET;
  for (int r = threadIdx.y; r < ET; r += 8) {
    const int ch = c0 + r;
    const int64_t i = n0 + threadIdx.x;
    t_sym[r][threadIdx.x] = (i < n && ch < c) ? __ldg(symbols + (int64_t)ch * n + i) : 0;
  }
  __syncthreads();
  for (int r = threadIdx.y; r < ET; r += 8) {
    const int64_t i = n0 + r;
    const int ch = c0 + threadIdx.x;
    if (i < n && ch < c) {
      const int32_t pr = rows ? __ldg(rows + i) : (int32_t)i;
      const float mean = pr >= 0 ? __ldg(params + (int64_t)pr * 2 * c + c + ch) : 0.f;
      y_hat[i * c + ch] = __fadd_rn((float)t_sym[threadIdx.x][r], mean);
    }
  }
}

__device__ __forceinline__ float std_cumulative(float x) {
  // 0.5 * erfc(-(2^-0.5) * x)
  return 0.5f * erfcf(-0.70710678118654752440f * x);
}

__global__ void __launch_bounds__(ET * 8)
gc_likelihood_kernel(const float* __restrict__ y, const float* __restrict__ params,
                     const int32_t* __restrict__ rows, int64_t n, int c, float scale_bound,
                     float lik_bound, float* __restrict__ y_hat, float* __restrict__ lik) {
  __shared__ float t_lik[ET][ET + 1];
  const int64_t n0 = (int64_t)blockIdx.x * ET;
  const int c0 = blockIdx.y * ET;
  for (int r = threadIdx.y; r < ET; r += 8) {
    const int64_t i = n0 + r;
    const int ch = c0 + threadIdx.x;
    float l = 0.f;
    if (i < n && ch < c) {
      const int32_t pr = rows ? __ldg(rows + i) : (int32_t)i;
      float scale = 0.f, mean = 0.f;
      if (pr >= 0) {
        scale = __ldg(params + (int64_t)pr * 2 * c + ch);
        mean = __ldg(params + (int64_t)pr * 2 * c + c + ch);
      }
      const float yh = __fadd_rn(rintf(__fsub_rn(__ldg(y + i * c + ch), mean)), mean);
      if (y_hat) y_hat[i * c + ch] = yh;
      const float s = fmaxf(scale, scale_bound);
      const float v = fabsf(__fsub_rn(yh, mean));
      l = std_cumulative((0.5f - v) / s) - std_cumulative((-0.5f - v) / s);
      l = fmaxf(l, lik_bound);
    }
    t_lik[r][threadIdx.x] = l;
  }
  __syncthreads();
  for (int r = threadIdx.y; r < ET; r += 8) {
    const int ch = c0 + r;
    const int64_t i = n0 + threadIdx.x;
    if (i < n && ch < c) lik[(int64_t)ch * n + i] = t_lik[threadIdx.x][r];
  }
}

// MODE 0: quantise z -> symbols [C,N] (+ z_hat [N,C]);  MODE 1: symbols [C,N] -> z_hat [N,C]
template <int MODE>
__global__ void __launch_bounds__(ET * 8)
eb_quant_kernel(const float* __restrict__ z, const int32_t* __restrict__ sym_in,
                const float* __restrict__ medians, int64_t n, int c, int32_t* __restrict__ sym_out,
                float* __restrict__ z_hat) {
  __shared__ int32_t t_sym[ET][ET + 1];
  const int64_t n0 = (int64_t)blockIdx.x * ET;
  const int c0 = blockIdx.y * ET;
  if (MODE == 0) {
    for (int r = threadIdx.y; r < ET; r += 8) {
      const int64_t i = n0 + r;
      const int ch = c0 + threadIdx.x;
      int32_t s = 0;
      if (i < n && ch < c) {
        const float med = __ldg(medians + ch);
        const float q = rintf(__fsub_rn(__ldg(z + i * c + ch), med));
        s = (int32_t)q;
        if (z_hat) z_hat[i * c + ch] = __fadd_rn(q, med);
      }
      t_sym[r][threadIdx.x] = s;
    }
    __syncthreads();
    for (int r = threadIdx.y; r < ET; r += 8) {
      const int ch = c0 + r;
      const int64_t i = n0 + threadIdx.x;
      if (i < n && ch < c) sym_out[(int64_t)ch * n + i] = t_sym[threadIdx.x][r];
    }
  } else {
    for (int r = threadIdx.y; r < ET; r += 8) {
      const int ch = c0 + r;
      const int64_t i = n0 + threadIdx.x;
      t_sym[r][threadIdx.x] = (i < n && ch < c) ? __ldg(sym_in + (int64_t)ch * n + i) : 0;
    }
    __syncthreads();
    for (int r = threadIdx.y; r < ET; r += 8) {
      const int64_t i = n0 + r;
      const int ch = c0 + threadIdx.x;
      if (i < n && ch < c)
        z_hat[i * c + ch] = __fadd_rn((float)t_sym[threadIdx.x][r], __ldg(medians + ch));
    }
  }
}

// factorised prior, filters (3,3,3,3): p = 58 pre-transformed floats of one channel
__device__ __forceinline__ float eb_logits(const float* __restrict__ p, float v) {
  float h[3], g[3];
#pragma unroll
  for (int j = 0; j < 3; ++j) {
    const float t = p[j] * v + p[3 + j];
    h[j] = t + p[6 + j] * tanhf(t);
  }
  p += 9;
#pragma unroll 1
  for (int layer = 0; layer < 3; ++layer) {
#pragma unroll
    for (int j = 0; j < 3; ++j) {
      const float t = p[3 * j] * h[0] + p[3 * j + 1] * h[1] + p[3 * j + 2] * h[2] + p[9 + j];
      g[j] = t + p[12 + j] * tanhf(t);
    }
    h[0] = g[0]; h[1] = g[1]; h[2] = g[2];
    p += 15;
  }
  return p[0] * h[0] + p[1] * h[1] + p[2] * h[2] + p[3];
}

__device__ __forceinline__ float sigmoidf_(float x) { return 1.0f / (1.0f + expf(-x)); }

__global__ void __launch_bounds__(ET * 8)
eb_likelihood_kernel(const float* __restrict__ z, const float* __restrict__ medians,
                     const float* __restrict__ eb_params, int64_t n, int c, float lik_bound,
                     float* __restrict__ z_hat, float* __restrict__ lik) {
  __shared__ float t_lik[ET][ET + 1];
  const int64_t n0 = (int64_t)blockIdx.x * ET;
  const int c0 = blockIdx.y * ET;
  for (int r = threadIdx.y; r < ET; r += 8) {
    const int64_t i = n0 + r;
    const int ch = c0 + threadIdx.x;
    float l = 0.f;
    if (i < n && ch < c) {
      const float med = __ldg(medians + ch);
      const float zh = __fadd_rn(rintf(__fsub_rn(__ldg(z + i * c + ch), med)), med);
      if (z_hat) z_hat[i * c + ch] = zh;
      const float* p = eb_params + (int64_t)ch * PCC_EB_PARAMS_PER_CHANNEL;
      const float lower = eb_logits(p, zh - 0.5f), upper = eb_logits(p, zh + 0.5f);
      const float su = lower + upper;
      const float sign = su > 0.f ? -1.f : (su < 0.f ? 1.f : 0.f);
      l = fmaxf(fabsf(sigmoidf_(sign * upper) - sigmoidf_(sign * lower)), lik_bound);
    }
    t_lik[r][threadIdx.x] = l;
  }
  __syncthreads();
  for (int r = threadIdx.y; r < ET; r += 8) {
    const int ch = c0 + r;
    const int64_t i = n0 + threadIdx.x;
    if (i < n && ch < c) lik[(int64_t)ch * n + i] = t_lik[threadIdx.x][r];
  }
}

__global__ void color_finalize_kernel(const uint64_t* __restrict__ keys, const float* __restrict__ feats,
                                      int64_t n, int c, float* __restrict__ out) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  int b, x, y, z;
  unpack_key(keys[i], b, x, y, z);
  float* o = out + i * (3 + c);
  o[0] = (float)x; o[1] = (float)y; o[2] = (float)z;
  for (int j = 0; j < c; ++j) {
    // torch.clamp(torch.round(F * 255), 0.0, 255.0) / 255 (model/model.py:206): comparisons, not
    // fmaxf/fminf, so that round(-0.3) = -0.0 stays -0.0 exactly as torch.clamp leaves it
    const float r = rintf(__fmul_rn(feats[i * c + j], 255.0f));
    const float q = r < 0.0f ? 0.0f : (r > 255.0f ? 255.0f : r);
    o[3 + j] = __fdiv_rn(q, 255.0f);
  }
}

static dim3 tile_grid(int64_t n, int c) { return dim3((unsigned)ceil_div(n, ET), (unsigned)ceil_div(c, ET)); }

}  // namespace pcc

using namespace pcc;

extern "C" {

size_t pcc_conv_head0_ws_bytes(int64_t n_in) { return (size_t)(n_in > 0 ? n_in : 1) * HP_K * sizeof(float); }

int pcc_conv_head0(const float* in, int64_t n_in, int32_t cin, const float* weight, int32_t cout, const float* bias,
                   const int32_t* nbr, int64_t nbr_stride, const int32_t* order, int64_t n_out, int32_t K,
                   float* out, void* ws, size_t ws_bytes, void* stream) {
  if (n_out == 0) return PCC_OK;
  PCC_REQUIRE(in && weight && nbr && out && ws, "null pointer");
  PCC_REQUIRE(cin > 0 && cout > 0 && K > 0 && K <= 27 && nbr_stride >= n_out, "bad shape");
  PCC_REQUIRE(cin % 4 == 0 && (reinterpret_cast<uintptr_t>(in) & 15) == 0, "cin must be a multiple of 4 and `in` 16-byte aligned");
  PCC_REQUIRE(ws_bytes >= pcc_conv_head0_ws_bytes(n_in), "workspace too small");
  cudaStream_t st = (cudaStream_t)stream;
  float* P = reinterpret_cast<float*>(ws);
  const size_t smem = ((size_t)cin * HP_K + (size_t)HP_ROWS * (cin + 4)) * sizeof(float);
  PCC_REQUIRE(smem <= 96 * 1024, "cin too large for the projection tile");
  static bool attr_done = false;
  if (!attr_done) {
    PCC_CHECK_CUDA(cudaFuncSetAttribute(head_project_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024));
    attr_done = true;
  }
  head_project_kernel<<<(unsigned)ceil_div(n_in, HP_ROWS), 256, smem, st>>>(in, n_in, cin, weight, K, cout, P);
  PCC_CHECK_LAUNCH("head_project_kernel");
  head_gather_kernel<<<grid_for(n_out), THREADS, 0, st>>>(P, nbr, nbr_stride, order, n_out, K, bias, out);
  PCC_CHECK_LAUNCH("head_gather_kernel");
  return PCC_OK;
}

int pcc_gather_rows(const float* in, int32_t c, const int32_t* rows, int64_t n, const float* add,
                    float* out, void* stream) {
  if (n == 0) return PCC_OK;
  PCC_REQUIRE(in && out && c > 0, "bad argument");
  cudaStream_t st = (cudaStream_t)stream;
  const bool vec = (c % 4 == 0) && ((reinterpret_cast<uintptr_t>(in) & 15) == 0) &&
                   ((reinterpret_cast<uintptr_t>(out) & 15) == 0) &&
                   (!add || (reinterpret_cast<uintptr_t>(add) & 15) == 0);
  if (vec)
    gather_rows_kernel<true><<<grid_for(n * (c / 4)), THREADS, 0, st>>>(in, c, rows, n, add, out);
  else
    gather_rows_kernel<false><<<grid_for(n * c), THREADS, 0, st>>>(in, c, rows, n, add, out);
  PCC_CHECK_LAUNCH("gather_rows_kernel");
  return PCC_OK;
}

size_t pcc_topk_ws_bytes(int64_t m) {
  if (m <= 0) m = 1;
  return 256 /*state*/ + TK_BINS * 4 + align_up((size_t)m * 4, 256) * 2 + scan_ws_bytes(m);
}

int pcc_topk_select(const float* logits, int64_t stride, int64_t m, int64_t k, int32_t* sel_rows,
                    void* ws, size_t ws_bytes, void* stream) {
  PCC_REQUIRE(k >= 0 && k <= m, "k must be in [0, m] (torch.topk raises otherwise)");
  if (k == 0) return PCC_OK;
  PCC_REQUIRE(logits && sel_rows && ws && stride >= 1, "bad argument");
  PCC_REQUIRE(ws_bytes >= pcc_topk_ws_bytes(m), "workspace too small");
  cudaStream_t st = (cudaStream_t)stream;
  char* w = reinterpret_cast<char*>(ws);
  TopkState* state = reinterpret_cast<TopkState*>(w);
  uint32_t* hist = reinterpret_cast<uint32_t*>(w + 256);
  size_t off = 256 + TK_BINS * 4;
  int32_t* eq = reinterpret_cast<int32_t*>(w + off);
  off += align_up((size_t)m * 4, 256);
  int32_t* sel = reinterpret_cast<int32_t*>(w + off);
  off += align_up((size_t)m * 4, 256);
  void* scan_ws = w + off;
  topk_init_kernel<<<1, THREADS, 0, st>>>(state, hist, k);
  PCC_CHECK_LAUNCH("topk_init_kernel");
  const unsigned hgrid = (unsigned)std::min<int64_t>(ceil_div(m, THREADS), (int64_t)sm_count() * 8);
  for (int pass = 0; pass < 3; ++pass) {
    topk_hist_kernel<<<hgrid, THREADS, 0, st>>>(logits, stride, m, pass, state, hist);
    PCC_CHECK_LAUNCH("topk_hist_kernel");
    topk_find_kernel<<<1, TF_THREADS, 0, st>>>(pass, state, hist);
    PCC_CHECK_LAUNCH("topk_find_kernel");
  }
  topk_eq_flags_kernel<<<grid_for(m), THREADS, 0, st>>>(logits, stride, m, state, eq);
  PCC_CHECK_LAUNCH("topk_eq_flags_kernel");
  int rc = exclusive_scan_i32(eq, eq, m, nullptr, scan_ws, st);
  if (rc) return rc;
  topk_sel_flags_kernel<<<grid_for(m), THREADS, 0, st>>>(logits, stride, m, state, eq, sel);
  PCC_CHECK_LAUNCH("topk_sel_flags_kernel");
  rc = exclusive_scan_i32(sel, sel, m, nullptr, scan_ws, st);
  if (rc) return rc;
  topk_emit_kernel<<<grid_for(m), THREADS, 0, st>>>(logits, stride, m, state, eq, sel, k, sel_rows);
  PCC_CHECK_LAUNCH("topk_emit_kernel");
  return PCC_OK;
}

int pcc_gc_quantize_index(const float* y, const float* params, const int32_t* rows, int64_t n,
                          int32_t c, const float* scale_table, int32_t n_scales, float scale_bound,
                          int32_t* symbols, int32_t* indexes, void* stream) {
  if (n == 0) return PCC_OK;
  PCC_REQUIRE(y && params && scale_table && symbols && indexes, "null pointer");
  PCC_REQUIRE(n_scales >= 1 && n_scales <= 64, "n_scales must be in [1, 64]");
  gc_quantize_kernel<0><<<tile_grid(n, c), dim3(ET, 8), 0, (cudaStream_t)stream>>>(
      y, params, rows, n, c, scale_table, n_scales, scale_bound, symbols, indexes);
  PCC_CHECK_LAUNCH("gc_quantize_kernel");
  return PCC_OK;
}

int pcc_gc_build_indexes(const float* params, const int32_t* rows, int64_t n, int32_t c,
                         const float* scale_table, int32_t n_scales, float scale_bound,
                         int32_t* indexes, void* stream) {
  if (n == 0) return PCC_OK;
  PCC_REQUIRE(params && scale_table && indexes, "null pointer");
  PCC_REQUIRE(n_scales >= 1 && n_scales <= 64, "n_scales must be in [1, 64]");
  gc_quantize_kernel<1><<<tile_grid(n, c), dim3(ET, 8), 0, (cudaStream_t)stream>>>(
      nullptr, params, rows, n, c, scale_table, n_scales, scale_bound, nullptr, indexes);
  PCC_CHECK_LAUNCH("gc_quantize_kernel");
  return PCC_OK;
}

int pcc_gc_dequantize(const int32_t* symbols, const float* params, const int32_t* rows, int64_t n,
                      int32_t c, float* y_hat, void* stream) {
  if (n == 0) return PCC_OK;
  PCC_REQUIRE(symbols && params && y_hat, "null pointer");
  gc_dequantize_kernel<<<tile_grid(n, c), dim3(ET, 8), 0, (cudaStream_t)stream>>>(symbols, params, rows, n, c, y_hat);
  PCC_CHECK_LAUNCH("gc_dequantize_kernel");
  return PCC_OK;
}

int pcc_gc_likelihood(const float* y, const float* params, const int32_t* rows, int64_t n, int32_t c,
                      float scale_bound, float likelihood_bound, float* y_hat, float* likelihood,
                      void* stream) {
  if (n == 0) return PCC_OK;
  PCC_REQUIRE(y && params && likelihood, "null pointer");
  gc_likelihood_kernel<<<tile_grid(n, c), dim3(ET, 8), 0, (cudaStream_t)stream>>>(
      y, params, rows, n, c, scale_bound, likelihood_bound, y_hat, likelihood);
  PCC_CHECK_LAUNCH("gc_likelihood_kernel");
  return PCC_OK;
}

int pcc_eb_quantize(const float* z, const float* medians, int64_t n, int32_t c, int32_t* symbols,
                    float* z_hat, void* stream) {
  if (n == 0) return PCC_OK;
  PCC_REQUIRE(z && medians && symbols, "null pointer");
  eb_quant_kernel<0><<<tile_grid(n, c), dim3(ET, 8), 0, (cudaStream_t)stream>>>(z, nullptr, medians, n, c, symbols, z_hat);
  PCC_CHECK_LAUNCH("eb_quant_kernel");
  return PCC_OK;
}

int pcc_eb_dequantize(const int32_t* symbols, const float* medians, int64_t n, int32_t c, float* z_hat,
                      void* stream) {
  if (n == 0) return PCC_OK;
  PCC_REQUIRE(symbols && medians && z_hat, "null pointer");
  eb_quant_kernel<1><<<tile_grid(n, c), dim3(ET, 8), 0, (cudaStream_t)stream>>>(nullptr, symbols, medians, n, c, nullptr, z_hat);
  PCC_CHECK_LAUNCH("eb_quant_kernel");
  return PCC_OK;
}

int pcc_eb_likelihood(const float* z, const float* medians, const float* eb_params, int64_t n,
                      int32_t c, float likelihood_bound, float* z_hat, float* likelihood, void* stream) {
  if (n == 0) return PCC_OK;
  PCC_REQUIRE(z && medians && eb_params && likelihood, "null pointer");
  eb_likelihood_kernel<<<tile_grid(n, c), dim3(ET, 8), 0, (cudaStream_t)stream>>>(
      z, medians, eb_params, n, c, likelihood_bound, z_hat, likelihood);
  PCC_CHECK_LAUNCH("eb_likelihood_kernel");
  return PCC_OK;
}

int pcc_color_finalize(const uint64_t* keys, const float* feats, int64_t n, int32_t c, float* out,
                       void* stream) {
  if (n == 0) return PCC_OK;
  PCC_REQUIRE(keys && feats && out && c > 0, "bad argument");
  color_finalize_kernel<<<grid_for(n), THREADS, 0, (cudaStream_t)stream>>>(keys, feats, n, c, out);
  PCC_CHECK_LAUNCH("color_finalize_kernel");
  return PCC_OK;
}

}  // extern "C"
