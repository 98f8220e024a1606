// Training-path kernels (reference: ColorModel.forward model/model.py:51-93 with
// loss.backward() in train.py:189-203).  The reference gets these from MinkowskiEngine's
// ConvolutionBackward (data + weight gradients) and torch autograd over CompressAI's modules.
//
//  * data gradient: no kernel here — din = conv(gout, reversed kernel map, W^T) runs on the forward
//    kernels (pcc_conv_fwd) with the map built from the output side.
//  * weight gradient (this file): dW[k] = sum over pairs of in[nbr[k][j]]^T (x) gout[order[j]], an
//    (cin x rows) x (rows x cout) product per offset with both operands gathered.  Output-stationary in
//    registers, cp.async double-buffered row stages, tiles without the offset skipped by the kernel
//    map's tap masks.  The row range is cut into S splits whose partial sums are added in a fixed order
//    by a second kernel, so the gradient is deterministic (no float atomics).
//  * rows scatter-add (backward of pcc_gather_rows), Gaussian likelihood with noise-quantised inputs
//    and its derivatives (CompressAI GaussianConditional.forward in training mode, LowerBound's
//    gradient rule included).
#include "pcc_common.cuh"

#include <algorithm>
#include <mutex>

namespace pcc {

__device__ __forceinline__ void cpa16(void* smem, const void* gmem, bool valid) {
  const uint32_t s = (uint32_t)__cvta_generic_to_shared(smem);
  const int bytes = valid ? 16 : 0;  // src-size 0 -> 16 zero bytes
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(s), "l"(gmem), "r"(bytes));
}
__device__ __forceinline__ void cpa_commit() { asm volatile("cp.async.commit_group;\n" ::); }
template <int N>
__device__ __forceinline__ void cpa_wait() {
  asm volatile("cp.async.wait_group %0;\n" ::"n"(N));
}

constexpr int WG_ROWS = 32;  // rows per stage

// packed fma.rn.f32x2: two independent IEEE fp32 FMAs per issue slot (as conv_exact.cu)
__device__ __forceinline__ unsigned long long wg_pack2(float lo, float hi) {
  unsigned long long v;
  asm("mov.b64 %0, {%1, %2};" : "=l"(v) : "f"(lo), "f"(hi));
  return v;
}
__device__ __forceinline__ void wg_unpack2(unsigned long long v, float& lo, float& hi) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ void wg_ffma2(unsigned long long& d, unsigned long long a, unsigned long long b) {
  asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(d) : "l"(a), "l"(b));
}

struct WgradParams {
  const float* in;
  const float* gout;
  const int32_t* nbr;       // [K x nbr_stride], tile-row indexed
  const uint32_t* tapmask;  // this offset group's masks or NULL
  const int32_t* order;     // this offset group's tile row -> output row or NULL
  float* part;              // [S * NZ][K][cin][cout]
  int64_t nbr_stride, n_out;
  int cin, cout, K, k_lo, k_hi, S;
};

// CTA: one offset k, one (ci tile, co tile), row tiles s, s + S, ...  The NY x NX x NZ threads form NZ
// row-split groups of an NY x NX arrangement; thread (tz, ty, tx) owns TM x TN elements of the
// (NY TM) x (NX TN) channel tile and accumulates the rows r = tz (mod NZ) of every stage.  Wide layers
// use NZ = 1 (16 x 16 threads, 8 x 8 registers: 128 x 128 channels); narrow ones (the 4-channel heads,
// the 2-channel quality branch, the xyz/rgb input layer) keep all lanes busy by splitting rows
// instead of channels.  Every row split writes its own partial tile.
template <int TM, int TN, int NY, int NX, int NZ, bool VEC>
__global__ void __launch_bounds__(NY * NX * NZ)
conv_wgrad_kernel(const WgradParams p) {
  constexpr int THREADS = NY * NX * NZ;
  constexpr int CI_T = NY * TM, CO_T = NX * TN;
  constexpr int STAGE = WG_ROWS * (CI_T + CO_T);
  static_assert(WG_ROWS % NZ == 0, "row splits must divide the stage");
  extern __shared__ __align__(16) float smem[];
  __shared__ int32_t s_in[PCC_TILE_M], s_go[PCC_TILE_M];
  const int k = p.k_lo + blockIdx.y;
  const int co_tiles = (p.cout + CO_T - 1) / CO_T;
  const int ci0 = (blockIdx.z / co_tiles) * CI_T, co0 = (blockIdx.z % co_tiles) * CO_T;
  const int tid = threadIdx.x;
  const int tx = tid % NX, ty = (tid / NX) % NY, tz = tid / (NX * NY);
  unsigned long long acc2[TM][TN / 2];
#pragma unroll
  for (int i = 0; i < TM; ++i)
#pragma unroll
    for (int j = 0; j < TN / 2; ++j) acc2[i][j] = 0ull;

  const int64_t tiles = (p.n_out + PCC_TILE_M - 1) / PCC_TILE_M;
  for (int64_t tile = blockIdx.x; tile < tiles; tile += p.S) {
    if (p.tapmask && !((p.tapmask[tile] >> k) & 1u)) continue;
    __syncthreads();  // previous tile's readers of s_in / s_go and of the stages are done
    for (int e = tid; e < PCC_TILE_M; e += THREADS) {
      const int64_t t = tile * PCC_TILE_M + e;
      int32_t ri = -1, ro = -1;
      if (t < p.n_out) {
        ri = __ldg(p.nbr + (int64_t)k * p.nbr_stride + t);
        ro = p.order ? __ldg(p.order + t) : (int32_t)t;
      }
      s_in[e] = ri;
      s_go[e] = ri >= 0 ? ro : -1;
    }
    __syncthreads();
    auto load_stage = [&](int chunk, int buf) {
      float* a_s = smem + buf * STAGE;
      float* g_s = a_s + WG_ROWS * CI_T;
      if (VEC) {
        for (int e = tid; e < WG_ROWS * (CI_T / 4); e += THREADS) {
          const int r = e / (CI_T / 4), c4 = e % (CI_T / 4);
          const int32_t row = s_in[chunk * WG_ROWS + r];
          const bool ok = row >= 0 && ci0 + c4 * 4 < p.cin;
          cpa16(a_s + r * CI_T + c4 * 4, ok ? p.in + (int64_t)row * p.cin + ci0 + c4 * 4 : p.in, ok);
        }
        for (int e = tid; e < WG_ROWS * (CO_T / 4); e += THREADS) {
          const int r = e / (CO_T / 4), c4 = e % (CO_T / 4);
          const int32_t row = s_go[chunk * WG_ROWS + r];
          const bool ok = row >= 0 && co0 + c4 * 4 < p.cout;
          cpa16(g_s + r * CO_T + c4 * 4, ok ? p.gout + (int64_t)row * p.cout + co0 + c4 * 4 : p.gout, ok);
        }
      } else {
        for (int e = tid; e < WG_ROWS * CI_T; e += THREADS) {
          const int r = e / CI_T, c = e % CI_T;
          const int32_t row = s_in[chunk * WG_ROWS + r];
          a_s[e] = (row >= 0 && ci0 + c < p.cin) ? __ldg(p.in + (int64_t)row * p.cin + ci0 + c) : 0.f;
        }
        for (int e = tid; e < WG_ROWS * CO_T; e += THREADS) {
          const int r = e / CO_T, c = e % CO_T;
          const int32_t row = s_go[chunk * WG_ROWS + r];
          g_s[e] = (row >= 0 && co0 + c < p.cout) ? __ldg(p.gout + (int64_t)row * p.cout + co0 + c) : 0.f;
        }
      }
      cpa_commit();
    };
    constexpr int CHUNKS = PCC_TILE_M / WG_ROWS;
    load_stage(0, 0);
#pragma unroll 1
    for (int c = 0; c < CHUNKS; ++c) {
      if (c + 1 < CHUNKS) {
        load_stage(c + 1, (c + 1) & 1);
        cpa_wait<1>();
      } else {
        cpa_wait<0>();
      }
      __syncthreads();
      const float* a_s = smem + (c & 1) * STAGE;
      const float* g_s = a_s + WG_ROWS * CI_T;
#pragma unroll 4
      for (int r = tz; r < WG_ROWS; r += NZ) {
        float a[TM], b[TN];
#pragma unroll
        for (int i = 0; i < TM; i += 4)
          *reinterpret_cast<float4*>(a + i) = *reinterpret_cast<const float4*>(a_s + r * CI_T + ty * TM + i);
#pragma unroll
        for (int j = 0; j < TN; j += 4)
          *reinterpret_cast<float4*>(b + j) = *reinterpret_cast<const float4*>(g_s + r * CO_T + tx * TN + j);
        unsigned long long bp[TN / 2];
#pragma unroll
        for (int j = 0; j < TN / 2; ++j) bp[j] = wg_pack2(b[2 * j], b[2 * j + 1]);
#pragma unroll
        for (int i = 0; i < TM; ++i) {
          const unsigned long long aa = wg_pack2(a[i], a[i]);
#pragma unroll
          for (int j = 0; j < TN / 2; ++j) wg_ffma2(acc2[i][j], aa, bp[j]);
        }
      }
      __syncthreads();  // stage (c & 1) is overwritten by the load issued in the next iteration
    }
  }
  float acc[TM][TN];
#pragma unroll
  for (int i = 0; i < TM; ++i)
#pragma unroll
    for (int j = 0; j < TN / 2; ++j) wg_unpack2(acc2[i][j], acc[i][2 * j], acc[i][2 * j + 1]);
  float* dst = p.part + (((int64_t)blockIdx.x * NZ + tz) * p.K + k) * p.cin * p.cout;
#pragma unroll
  for (int i = 0; i < TM; ++i) {
    const int ci = ci0 + ty * TM + i;
    if (ci >= p.cin) continue;
#pragma unroll
    for (int j = 0; j < TN; ++j) {
      const int co = co0 + tx * TN + j;
      if (co < p.cout) dst[(int64_t)ci * p.cout + co] = acc[i][j];
    }
  }
}

__global__ void wgrad_reduce_kernel(const float* __restrict__ part, int S, int64_t n, float* __restrict__ dw) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  float s = 0.f;
  for (int j = 0; j < S; ++j) s = __fadd_rn(s, __ldg(part + (int64_t)j * n + i));
  dw[i] = s;
}

// channel-tile shape of one side: c <= 4: 4 x 1 lanes, <= 16: 4 x 4, <= 32: 4 x 8, <= 64: 8 x 8, else 8 x 16
struct Side { int t, n; };
static Side side_for(int c) {
  if (c <= 4) return {4, 1};
  if (c <= 16) return {4, 4};
  if (c <= 32) return {4, 8};
  if (c <= 64) return {8, 8};
  return {8, 16};
}
static int side_idx(int c) { return c <= 4 ? 0 : c <= 16 ? 1 : c <= 32 ? 2 : c <= 64 ? 3 : 4; }
static int splits_nz(Side a, Side b) { return std::min(32, 256 / (a.n * b.n)); }

constexpr int WG_MAX_SPLITS = 64;
static int wgrad_splits(int k_launch, int zt, int64_t tiles) {
  // about four CTAs per SM over (splits x offsets of the launch x channel tiles); never more splits than row tiles
  const int target = 4 * sm_count();
  int64_t s = target / std::max(1, k_launch * zt);
  s = std::min<int64_t>(s, tiles);
  return (int)std::max<int64_t>(1, std::min<int64_t>(s, WG_MAX_SPLITS));
}

template <int TM, int NY, int TN, int NX, bool VEC>
static int launch_wgrad(const WgradParams& p, int zt, cudaStream_t st) {
  constexpr int NZ = (256 / (NY * NX)) < 32 ? (256 / (NY * NX)) : 32;
  constexpr size_t smem = 2 * WG_ROWS * (NY * TM + NX * TN) * sizeof(float);
  static bool attr_done = false;
  if (!attr_done) {
    PCC_CHECK_CUDA(cudaFuncSetAttribute(conv_wgrad_kernel<TM, TN, NY, NX, NZ, VEC>,
                                        cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    attr_done = true;
  }
  dim3 grid((unsigned)p.S, (unsigned)(p.k_hi - p.k_lo), (unsigned)zt);
  conv_wgrad_kernel<TM, TN, NY, NX, NZ, VEC><<<grid, NY * NX * NZ, smem, st>>>(p);
  PCC_CHECK_LAUNCH("conv_wgrad_kernel");
  return PCC_OK;
}

template <int TM, int NY, bool VEC>
static int launch_wgrad_co(const WgradParams& p, int zt, int co_idx, cudaStream_t st) {
  switch (co_idx) {
    case 0: return launch_wgrad<TM, NY, 4, 1, VEC>(p, zt, st);
    case 1: return launch_wgrad<TM, NY, 4, 4, VEC>(p, zt, st);
    case 2: return launch_wgrad<TM, NY, 4, 8, VEC>(p, zt, st);
    case 3: return launch_wgrad<TM, NY, 8, 8, VEC>(p, zt, st);
    default: return launch_wgrad<TM, NY, 8, 16, VEC>(p, zt, st);
  }
}

template <bool VEC>
static int launch_wgrad_ci(const WgradParams& p, int zt, int ci_idx, int co_idx, cudaStream_t st) {
  switch (ci_idx) {
    case 0: return launch_wgrad_co<4, 1, VEC>(p, zt, co_idx, st);
    case 1: return launch_wgrad_co<4, 4, VEC>(p, zt, co_idx, st);
    case 2: return launch_wgrad_co<4, 8, VEC>(p, zt, co_idx, st);
    case 3: return launch_wgrad_co<8, 8, VEC>(p, zt, co_idx, st);
    default: return launch_wgrad_co<8, 16, VEC>(p, zt, co_idx, st);
  }
}

// ---------------------------------------------------------------- scatter-add of rows
// Column sums of a row-major [n x c] matrix (the bias gradient): block b adds rows b, b + B, ... per column in
// registers (consecutive threads on consecutive columns, so every row is one coalesced read), its row lanes are
// combined through shared memory in lane order, and col_sum_final_kernel adds the B partial rows in block order:
// the same bits on every run.  HBM-bound: 4 * n * c bytes read once.
constexpr int COLSUM_BLOCKS = 148 * 4;

template <int VEC>
__global__ void __launch_bounds__(256) col_sum_partial_kernel(const float* __restrict__ in, int64_t n, int c,
                                                              float* __restrict__ part) {
  __shared__ float sm[256 * VEC];
  const int cv = c / VEC;                       // column groups per row (c <= 1024, so cv <= 256 for VEC = 4)
  const int lanes = cv >= 256 ? 1 : 256 / cv;   // rows a block reads at a time
  float* dst = part + (int64_t)blockIdx.x * c;
  for (int col0 = 0; col0 < cv; col0 += 256) {  // more than one pass only for VEC = 1 and c > 256
    const int lane = threadIdx.x / cv, col = col0 + (cv >= 256 ? threadIdx.x : threadIdx.x % cv);
    float acc[VEC];
#pragma unroll
    for (int v = 0; v < VEC; ++v) acc[v] = 0.f;
    if (lane < lanes && col < cv) {
      for (int64_t r = (int64_t)blockIdx.x * lanes + lane; r < n; r += (int64_t)gridDim.x * lanes) {
        if (VEC == 4) {
          const float4 x = __ldg(reinterpret_cast<const float4*>(in + r * c) + col);
          acc[0] += x.x; acc[1 % VEC] += x.y; acc[2 % VEC] += x.z; acc[3 % VEC] += x.w;
        } else {
          acc[0] += __ldg(in + r * c + col);
        }
      }
    }
#pragma unroll
    for (int v = 0; v < VEC; ++v) sm[threadIdx.x * VEC + v] = acc[v];
    __syncthreads();
    if (lane == 0 && col < cv) {
#pragma unroll
      for (int v = 0; v < VEC; ++v) {
        float t = sm[threadIdx.x * VEC + v];
        for (int l = 1; l < lanes; ++l) t += sm[(threadIdx.x + l * cv) * VEC + v];
        dst[col * VEC + v] = t;
      }
    }
    __syncthreads();
  }
}

__global__ void col_sum_final_kernel(const float* __restrict__ part, int blocks, int c, float* __restrict__ out) {
  const int col = blockIdx.x * blockDim.x + threadIdx.x;
  if (col >= c) return;
  float t = 0.f;
  for (int b = 0; b < blocks; ++b) t += part[(int64_t)b * c + col];
  out[col] = t;
}

__global__ void scatter_rows_add_kernel(const float* __restrict__ g, int c, const int32_t* __restrict__ rows,
                                        int64_t n, float* __restrict__ out) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n * c) return;
  const int64_t i = t / c;
  const int j = (int)(t % c);
  const int32_t r = rows ? __ldg(rows + i) : (int32_t)i;
  if (r >= 0) atomicAdd(out + (int64_t)r * c + j, __ldg(g + t));
}

// ---------------------------------------------------------------- Gaussian likelihood, training mode
__device__ __forceinline__ float std_cdf(float x) { return 0.5f * erfcf(-0.70710678118654752440f * x); }
__device__ __forceinline__ float std_pdf(float x) { return 0.3989422804014327f * __expf(-0.5f * x * x); }

// lik[i][ch] = max(Phi((0.5 - v) / s) - Phi((-0.5 - v) / s), bound), v = |y~ - mean|, s = max(scale, sb)
__global__ void gc_train_fwd_kernel(const float* __restrict__ yt, const float* __restrict__ params,
                                    const int32_t* __restrict__ rows, int64_t n, int c, float sb, float lb,
                                    float* __restrict__ lik) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n * c) return;
  const int64_t i = t / c;
  const int ch = (int)(t % c);
  const int32_t pr = rows ? __ldg(rows + i) : (int32_t)i;
  float scale = 0.f, mean = 0.f;
  if (pr >= 0) {
    scale = __ldg(params + (int64_t)pr * 2 * c + ch);
    mean = __ldg(params + (int64_t)pr * 2 * c + c + ch);
  }
  const float s = fmaxf(scale, sb);
  const float v = fabsf(__ldg(yt + t) - mean);
  const float l = std_cdf((0.5f - v) / s) - std_cdf((-0.5f - v) / s);
  lik[t] = fmaxf(l, lb);
}

// g_lik -> g_y [n, c] and g_prow [n, 2c] (d/d scale | d/d mean of the gathered parameter row).
// LowerBound backward (compressai.ops.LowerBound): the gradient passes when x >= bound or it is negative.
__global__ void gc_train_bwd_kernel(const float* __restrict__ yt, const float* __restrict__ params,
                                    const int32_t* __restrict__ rows, int64_t n, int c, float sb, float lb,
                                    const float* __restrict__ g_lik, float* __restrict__ g_y,
                                    float* __restrict__ g_prow) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n * c) return;
  const int64_t i = t / c;
  const int ch = (int)(t % c);
  const int32_t pr = rows ? __ldg(rows + i) : (int32_t)i;
  float scale = 0.f, mean = 0.f;
  if (pr >= 0) {
    scale = __ldg(params + (int64_t)pr * 2 * c + ch);
    mean = __ldg(params + (int64_t)pr * 2 * c + c + ch);
  }
  const float s = fmaxf(scale, sb);
  const float d = __ldg(yt + t) - mean;
  const float v = fabsf(d);
  const float u = (0.5f - v) / s, w = (-0.5f - v) / s;
  const float l = std_cdf(u) - std_cdf(w);
  float g = __ldg(g_lik + t);
  if (!(l >= lb || g < 0.f)) g = 0.f;
  const float pu = std_pdf(u), pw = std_pdf(w);
  const float dl_dv = (pw - pu) / s;
  const float sgn = d > 0.f ? 1.f : (d < 0.f ? -1.f : 0.f);  // torch.abs' subgradient
  const float gy = g * dl_dv * sgn;
  float gs = g * (w * pw - u * pu) / s;
  if (!(scale >= sb || gs < 0.f)) gs = 0.f;
  g_y[t] = gy;
  g_prow[i * 2 * c + ch] = gs;
  g_prow[i * 2 * c + c + ch] = -gy;
}

}  // namespace pcc

namespace pcc {
bool wgrad_tf32_supported(const pcc_wgrad_args* a);
int wgrad_tf32(const pcc_wgrad_args* a, int* splits, cudaStream_t st);
}  // namespace pcc

namespace pcc {
// ---------------------------------------------------------------- pcc_col_sum without any allocation
// Same two kernels as pcc_col_sum, but the block partials live in a file-scope device array instead of a
// caller-supplied workspace, for callers that may not allocate (CUDA graph capture, serving loops).
// The array is part of the module image: every GPU that loads the library holds its 606,208 bytes until the
// process exits.  `g_colsum_static_scratch` in HOST code is only a handle for the runtime's symbol table, not
// a device address; the address is asked for once per device with cudaGetSymbolAddress and cached.
constexpr int COLSUM_STATIC_MAX_C = PCC_COL_SUM_STATIC_MAX_C;
static __device__ __align__(256) float g_colsum_static_scratch[COLSUM_BLOCKS * COLSUM_STATIC_MAX_C];

constexpr int COLSUM_MAX_DEVICES = 64;

// Device address of the scratch array on the CURRENT device (looked up on first use on that device).
int colsum_static_scratch(float** out) {
  static std::mutex mu;
  static float* addr[COLSUM_MAX_DEVICES] = {};
  int dev = -1;
  PCC_CHECK_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= COLSUM_MAX_DEVICES) {
    set_error("pcc_col_sum_static: device ordinal %d outside the supported range [0, %d)", dev, COLSUM_MAX_DEVICES);
    return PCC_ERR_INVALID_ARG;
  }
  std::lock_guard<std::mutex> lock(mu);
  if (!addr[dev]) {
    void* p = nullptr;
    PCC_CHECK_CUDA(cudaGetSymbolAddress(&p, g_colsum_static_scratch));
    if (!p) {
      set_error("pcc_col_sum_static: cudaGetSymbolAddress returned a null address on device %d", dev);
      return PCC_ERR_CUDA;
    }
    addr[dev] = static_cast<float*>(p);
  }
  *out = addr[dev];
  return PCC_OK;
}

}  // namespace pcc

using namespace pcc;

extern "C" {

size_t pcc_conv_wgrad_ws_bytes(int32_t K, int32_t cin, int32_t cout) {
  const Side a = side_for(cin), b = side_for(cout);
  // sized for the largest split count a launch may pick (it depends on the offset groups of the map)
  return (size_t)WG_MAX_SPLITS * splits_nz(a, b) * K * cin * cout * sizeof(float) + 256;
}

int pcc_conv_wgrad(const pcc_wgrad_args* a, void* stream) {
  PCC_REQUIRE(a != nullptr, "null args");
  PCC_REQUIRE(a->dweight && a->cin > 0 && a->cout > 0 && a->K > 0 && a->K <= 27, "bad shape");
  cudaStream_t st = (cudaStream_t)stream;
  const int64_t wn = (int64_t)a->K * a->cin * a->cout;
  if (a->n_out == 0) {
    PCC_CHECK_CUDA(cudaMemsetAsync(a->dweight, 0, wn * sizeof(float), st));
    return PCC_OK;
  }
  PCC_REQUIRE(a->in && a->gout && a->nbr && a->ws, "null pointer");
  PCC_REQUIRE(a->nbr_stride >= a->n_out, "nbr_stride < n_out");
  PCC_REQUIRE(a->ws_bytes >= pcc_conv_wgrad_ws_bytes(a->K, a->cin, a->cout), "workspace too small");
  PCC_REQUIRE(a->n_groups <= 1 || (a->order && a->n_groups <= a->K), "n_groups > 1 needs order and n_groups <= K");
  if (a->mode == PCC_MODE_TF32 && wgrad_tf32_supported(a)) {
    int splits = 1;
    const int rc = wgrad_tf32(a, &splits, st);
    if (rc != PCC_OK) return rc;
    wgrad_reduce_kernel<<<(unsigned)ceil_div(wn, 256), 256, 0, st>>>(reinterpret_cast<const float*>(a->ws), splits, wn,
                                                                   a->dweight);
    PCC_CHECK_LAUNCH("wgrad_reduce_kernel");
    return PCC_OK;
  }
  const Side sa = side_for(a->cin), sb = side_for(a->cout);
  const int zt = (int)(ceil_div(a->cin, sa.t * sa.n) * ceil_div(a->cout, sb.t * sb.n));
  WgradParams p;
  p.in = a->in; p.gout = a->gout; p.nbr = a->nbr; p.part = reinterpret_cast<float*>(a->ws);
  p.nbr_stride = a->nbr_stride; p.n_out = a->n_out; p.cin = a->cin; p.cout = a->cout; p.K = a->K;
  const int ng = a->n_groups > 1 ? a->n_groups : 1;
  const int per = (a->K + ng - 1) / ng;
  const int64_t tiles = ceil_div(a->n_out, PCC_TILE_M);
  p.S = wgrad_splits(per, zt, tiles);  // the same for every group: the partial buffer is [S * NZ][K]
  const bool vec = a->cin % 4 == 0 && a->cout % 4 == 0 &&
                   (reinterpret_cast<uintptr_t>(a->in) & 15) == 0 && (reinterpret_cast<uintptr_t>(a->gout) & 15) == 0;
  for (int g = 0; g < ng; ++g) {
    p.k_lo = g * per;
    p.k_hi = std::min((g + 1) * per, (int)a->K);
    if (p.k_lo >= p.k_hi) break;
    p.order = a->order ? a->order + (int64_t)g * a->n_out : nullptr;
    p.tapmask = a->tapmask ? a->tapmask + (int64_t)g * tiles : nullptr;
    const int rc = vec ? launch_wgrad_ci<true>(p, zt, side_idx(a->cin), side_idx(a->cout), st)
                       : launch_wgrad_ci<false>(p, zt, side_idx(a->cin), side_idx(a->cout), st);
    if (rc != PCC_OK) return rc;
  }
  wgrad_reduce_kernel<<<(unsigned)ceil_div(wn, 256), 256, 0, st>>>(p.part, p.S * splits_nz(sa, sb), wn, a->dweight);
  PCC_CHECK_LAUNCH("wgrad_reduce_kernel");
  return PCC_OK;
}

size_t pcc_col_sum_ws_bytes(int32_t c) { return c > 0 ? (size_t)COLSUM_BLOCKS * (size_t)c * sizeof(float) : 0; }

int pcc_col_sum(const float* in, int64_t n, int32_t c, float* out, void* ws, size_t ws_bytes, void* stream) {
  PCC_REQUIRE(out && c > 0 && c <= 1024 && n >= 0, "bad argument");
  cudaStream_t st = (cudaStream_t)stream;
  if (n == 0) {
    PCC_CHECK_CUDA(cudaMemsetAsync(out, 0, (size_t)c * sizeof(float), st));
    return PCC_OK;
  }
  PCC_REQUIRE(in && ws && ws_bytes >= pcc_col_sum_ws_bytes(c), "workspace too small (pcc_col_sum_ws_bytes)");
  const bool vec = (c % 4 == 0) && ((reinterpret_cast<uintptr_t>(in) & 15) == 0);
  const int blocks = (int)std::min<int64_t>(COLSUM_BLOCKS, ceil_div(n, 64));
  if (vec) col_sum_partial_kernel<4><<<blocks, 256, 0, st>>>(in, n, c, reinterpret_cast<float*>(ws));
  else col_sum_partial_kernel<1><<<blocks, 256, 0, st>>>(in, n, c, reinterpret_cast<float*>(ws));
  PCC_CHECK_LAUNCH("col_sum_partial_kernel");
  col_sum_final_kernel<<<(unsigned)ceil_div(c, 128), 128, 0, st>>>(reinterpret_cast<const float*>(ws), blocks, c, out);
  PCC_CHECK_LAUNCH("col_sum_final_kernel");
  return PCC_OK;
}

int pcc_col_sum_static(const float* in, int64_t n, int32_t c, float* out, void* stream) {
  PCC_REQUIRE(out && c > 0 && n >= 0, "bad argument");
  if (c > COLSUM_STATIC_MAX_C) {
    set_error("pcc_col_sum_static: c = %d exceeds the static scratch limit PCC_COL_SUM_STATIC_MAX_C = %d columns "
              "(use pcc_col_sum with a caller workspace)", (int)c, COLSUM_STATIC_MAX_C);
    return PCC_ERR_CAPACITY;
  }
  cudaStream_t st = (cudaStream_t)stream;
  if (n == 0) {
    PCC_CHECK_CUDA(cudaMemsetAsync(out, 0, (size_t)c * sizeof(float), st));
    return PCC_OK;
  }
  PCC_REQUIRE(in, "null input");
  float* scratch = nullptr;
  const int rc = colsum_static_scratch(&scratch);
  if (rc != PCC_OK) return rc;
  // Nothing has to start at zero: every block 0..blocks-1 overwrites all c entries of its partial row before
  // col_sum_final_kernel reads exactly those blocks x c entries; what an earlier call left beyond them is not read.
  const bool vec = (c % 4 == 0) && ((reinterpret_cast<uintptr_t>(in) & 15) == 0);
  const int blocks = (int)std::min<int64_t>(COLSUM_BLOCKS, ceil_div(n, 64));
  if (vec) col_sum_partial_kernel<4><<<blocks, 256, 0, st>>>(in, n, c, scratch);
  else col_sum_partial_kernel<1><<<blocks, 256, 0, st>>>(in, n, c, scratch);
  PCC_CHECK_LAUNCH("col_sum_partial_kernel (static scratch)");
  col_sum_final_kernel<<<(unsigned)ceil_div(c, 128), 128, 0, st>>>(scratch, blocks, c, out);
  PCC_CHECK_LAUNCH("col_sum_final_kernel (static scratch)");
  return PCC_OK;
}

int pcc_scatter_rows_add(const float* g, int32_t c, const int32_t* rows, int64_t n, float* out, void* stream) {
  if (n == 0) return PCC_OK;
  PCC_REQUIRE(g && out && c > 0, "bad argument");
  scatter_rows_add_kernel<<<(unsigned)ceil_div(n * c, 256), 256, 0, (cudaStream_t)stream>>>(g, c, rows, n, out);
  PCC_CHECK_LAUNCH("scatter_rows_add_kernel");
  return PCC_OK;
}

int pcc_gc_likelihood_train_fwd(const float* y_noisy, const float* params, const int32_t* rows, int64_t n,
                                int32_t c, float scale_bound, float likelihood_bound, float* likelihood,
                                void* stream) {
  if (n == 0) return PCC_OK;
  PCC_REQUIRE(y_noisy && params && likelihood && c > 0, "bad argument");
  gc_train_fwd_kernel<<<(unsigned)ceil_div(n * c, 256), 256, 0, (cudaStream_t)stream>>>(
      y_noisy, params, rows, n, c, scale_bound, likelihood_bound, likelihood);
  PCC_CHECK_LAUNCH("gc_train_fwd_kernel");
  return PCC_OK;
}

int pcc_gc_likelihood_train_bwd(const float* y_noisy, const float* params, const int32_t* rows, int64_t n,
                                int32_t c, float scale_bound, float likelihood_bound, const float* g_likelihood,
                                float* g_y, float* g_param_rows, void* stream) {
  if (n == 0) return PCC_OK;
  PCC_REQUIRE(y_noisy && params && g_likelihood && g_y && g_param_rows && c > 0, "bad argument");
  gc_train_bwd_kernel<<<(unsigned)ceil_div(n * c, 256), 256, 0, (cudaStream_t)stream>>>(
      y_noisy, params, rows, n, c, scale_bound, likelihood_bound, g_likelihood, g_y, g_param_rows);
  PCC_CHECK_LAUNCH("gc_train_bwd_kernel");
  return PCC_OK;
}

}  // extern "C"
