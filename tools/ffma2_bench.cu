// Microbenchmark: what does the FP32 pipe deliver for the exact conv kernel's inner loop?
//
// conv_exact_kernel<128> (csrc/conv_exact.cu) keeps an 8 x 8 register tile per thread (32 packed
// fma.rn.f32x2 accumulators) and, per 4 input channels, reads 8 LDS.128 of gathered rows and 8 LDS.128
// of weights for 128 FFMA2.  ncu shows the FMA pipe 75 % active.  This tool runs that loop alone:
//   mode 0  operands in registers only (pure FFMA2 issue)
//   mode 1  + the kernel's shared-memory operand loads (same addressing: 4 rows x 16 B broadcast,
//           8 x 16 B weight quads)
//   mode 2  + a __syncthreads() every 32 channels (one pipeline step)
//   mode 3  mode 2 + 8 cp.async of 16 B per thread and step into a 3-stage ring (the producer side)
//   mode 4  mode 3 without the 4 weight-tile copies (as if the weight tile came by one TMA load)
// for 1 and 2 CTAs of 256 threads per SM.  Output: TFLOP/s and share of the FP32 peak.
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ffma2_bench tools/ffma2_bench.cu && ./ffma2_bench
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s:%d %s\n", __FILE__, __LINE__, cudaGetErrorString(e)); exit(1); } } while (0)

__device__ __forceinline__ unsigned long long pack2(float lo, float hi) {
  unsigned long long r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void ffma2(unsigned long long& d, unsigned long long a, unsigned long long b) {
  asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(d) : "l"(a), "l"(b));
}
__device__ __forceinline__ void cp16(void* smem, const void* g) {
  const unsigned s = (unsigned)__cvta_generic_to_shared(smem);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(s), "l"(g));
}

constexpr int TM = 128, TK = 32, TN = 128, AP = TK + 4, ST = 3;
struct Smem {
  float a[ST][TM][AP];
  float b[ST][TK][TN];
};

template <int MODE>
__global__ void __launch_bounds__(256, 2) loop_kernel(const float* __restrict__ g, int steps, float* __restrict__ out) {
  extern __shared__ __align__(16) unsigned char raw[];
  Smem& sm = *reinterpret_cast<Smem*>(raw);
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const int tx = lane & 7, ty = lane >> 3;
  const int rbase = (wid >> 1) * 32 + ty, cbase = (wid & 1) * 64 + tx * 4;
  for (int i = tid; i < (int)(sizeof(Smem) / 4); i += 256) reinterpret_cast<float*>(raw)[i] = 1.0f + (i & 7) * 0.125f;
  __syncthreads();
  unsigned long long acc[8][4];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int c = 0; c < 4; ++c) acc[i][c] = 0ull;
  float areg[8] = {1.f, 1.5f, 2.f, 2.5f, 3.f, 3.5f, 4.f, 4.5f};
  unsigned long long breg[4] = {pack2(1.f, 2.f), pack2(3.f, 4.f), pack2(5.f, 6.f), pack2(7.f, 8.f)};
  for (int s = 0; s < steps; ++s) {
    const int st = s % ST;
    if (MODE >= 2) __syncthreads();
    if (MODE >= 3) {
      const int nst = (s + 2) % ST;
#pragma unroll
      for (int j = 0; j < 4; ++j) cp16(&sm.a[nst][(tid >> 3) + 32 * j][(tid & 7) * 4], g + ((size_t)(tid + 256 * j) * 4) % 4096);
      if (MODE == 3) {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int t = tid + 256 * j;
          cp16(&sm.b[nst][t / 32][(t % 32) * 4], g + ((size_t)t * 4) % 4096);
        }
      }
      asm volatile("cp.async.commit_group;\n cp.async.wait_group 1;\n" ::);
    }
#pragma unroll
    for (int kv = 0; kv < TK / 4; ++kv) {
      float a[8][4];
      if (MODE >= 1) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const float4 v = *reinterpret_cast<const float4*>(&sm.a[st][rbase + 4 * i][kv * 4]);
          a[i][0] = v.x; a[i][1] = v.y; a[i][2] = v.z; a[i][3] = v.w;
        }
      } else {
#pragma unroll
        for (int i = 0; i < 8; ++i)
#pragma unroll
          for (int e = 0; e < 4; ++e) a[i][e] = areg[i] + (float)e;
      }
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        unsigned long long bp[4];
        if (MODE >= 1) {
          const float* brow = &sm.b[st][kv * 4 + e][0];
          const float4 b0 = *reinterpret_cast<const float4*>(brow + cbase);
          const float4 b1 = *reinterpret_cast<const float4*>(brow + cbase + 32);
          bp[0] = pack2(b0.x, b0.y); bp[1] = pack2(b0.z, b0.w); bp[2] = pack2(b1.x, b1.y); bp[3] = pack2(b1.z, b1.w);
        } else {
#pragma unroll
          for (int c = 0; c < 4; ++c) bp[c] = breg[c];
        }
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const unsigned long long aa = pack2(a[i][e], a[i][e]);
#pragma unroll
          for (int c = 0; c < 4; ++c) ffma2(acc[i][c], aa, bp[c]);
        }
      }
    }
    if (MODE == 0) {  // keep the register operands live and changing
      areg[s & 7] += 1e-3f;
    }
  }
  asm volatile("cp.async.wait_group 0;\n" ::);
  unsigned long long x = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int c = 0; c < 4; ++c) x ^= acc[i][c];
  if (x == 0x1234567ull) out[blockIdx.x * 256 + tid] = 1.0f;
}

// Variant: 16 x 8 register tile, 128 threads per CTA (warp w: rows 32 w .. + 31 as 2 ty x 16, columns
// 16 tx x 8): 24 operand loads per 256 FFMA2 instead of 32.  Same modes.
template <int MODE>
__global__ void __launch_bounds__(128, 2) loop16_kernel(const float* __restrict__ g, int steps, float* __restrict__ out) {
  extern __shared__ __align__(16) unsigned char raw[];
  Smem& sm = *reinterpret_cast<Smem*>(raw);
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const int tx = lane & 15, ty = lane >> 4;
  const int rbase = wid * 32 + ty, cbase = tx * 4;  // rows rbase + 2 i (i < 16); columns cbase .. +3 and 64 + cbase .. +3
  for (int i = tid; i < (int)(sizeof(Smem) / 4); i += 128) reinterpret_cast<float*>(raw)[i] = 1.0f + (i & 7) * 0.125f;
  __syncthreads();
  unsigned long long acc[16][4];
#pragma unroll
  for (int i = 0; i < 16; ++i)
#pragma unroll
    for (int c = 0; c < 4; ++c) acc[i][c] = 0ull;
  for (int s = 0; s < steps; ++s) {
    const int st = s % ST;
    if (MODE >= 2) __syncthreads();
    if (MODE >= 3) {
      const int nst = (s + 2) % ST;
#pragma unroll
      for (int j = 0; j < 8; ++j) cp16(&sm.a[nst][(tid >> 3) + 16 * j][(tid & 7) * 4], g + ((size_t)(tid + 128 * j) * 4) % 4096);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int t = tid + 128 * j;
        cp16(&sm.b[nst][t / 32][(t % 32) * 4], g + ((size_t)t * 4) % 4096);
      }
      asm volatile("cp.async.commit_group;\n cp.async.wait_group 1;\n" ::);
    }
#pragma unroll
    for (int kv = 0; kv < TK / 4; ++kv) {
      float a[16][4];
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        const float4 v = *reinterpret_cast<const float4*>(&sm.a[st][rbase + 2 * i][kv * 4]);
        a[i][0] = v.x; a[i][1] = v.y; a[i][2] = v.z; a[i][3] = v.w;
      }
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        unsigned long long bp[4];
        const float* brow = &sm.b[st][kv * 4 + e][0];
        const float4 b0 = *reinterpret_cast<const float4*>(brow + cbase);
        const float4 b1 = *reinterpret_cast<const float4*>(brow + cbase + 64);
        bp[0] = pack2(b0.x, b0.y); bp[1] = pack2(b0.z, b0.w); bp[2] = pack2(b1.x, b1.y); bp[3] = pack2(b1.z, b1.w);
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          const unsigned long long aa = pack2(a[i][e], a[i][e]);
#pragma unroll
          for (int c = 0; c < 4; ++c) ffma2(acc[i][c], aa, bp[c]);
        }
      }
    }
  }
  asm volatile("cp.async.wait_group 0;\n" ::);
  unsigned long long x = 0;
#pragma unroll
  for (int i = 0; i < 16; ++i)
#pragma unroll
    for (int c = 0; c < 4; ++c) x ^= acc[i][c];
  if (x == 0x1234567ull) out[blockIdx.x * 128 + tid] = 1.0f;
}

template <int MODE>
static void run16(int ctas_per_sm, int sms, const float* g, float* out, double peak) {
  const size_t smem = sizeof(Smem);
  CK(cudaFuncSetAttribute(loop16_kernel<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const int steps = 4000, grid = sms * ctas_per_sm;
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  loop16_kernel<MODE><<<grid, 128, smem>>>(g, 100, out);
  CK(cudaDeviceSynchronize());
  CK(cudaEventRecord(e0));
  loop16_kernel<MODE><<<grid, 128, smem>>>(g, steps, out);
  CK(cudaEventRecord(e1));
  CK(cudaDeviceSynchronize());
  float ms = 0;
  CK(cudaEventElapsedTime(&ms, e0, e1));
  const double flop = (double)grid * steps * 2.0 * TM * TN * TK;
  const double tf = flop / (ms * 1e-3) / 1e12;
  printf("16x8 tile, 128 threads: mode %d  CTAs/SM %d : %6.2f TFLOP/s  = %.3f of %.1f\n", MODE, ctas_per_sm, tf, tf / peak, peak);
}

template <int MODE>
static void run(int ctas_per_sm, int sms, const float* g, float* out, double peak) {
  const size_t smem = sizeof(Smem);
  CK(cudaFuncSetAttribute(loop_kernel<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const int steps = 4000, grid = sms * ctas_per_sm;
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  loop_kernel<MODE><<<grid, 256, smem>>>(g, 100, out);
  CK(cudaDeviceSynchronize());
  CK(cudaEventRecord(e0));
  loop_kernel<MODE><<<grid, 256, smem>>>(g, steps, out);
  CK(cudaEventRecord(e1));
  CK(cudaDeviceSynchronize());
  float ms = 0;
  CK(cudaEventElapsedTime(&ms, e0, e1));
  const double flop = (double)grid * steps * 2.0 * TM * TN * TK;
  const double tf = flop / (ms * 1e-3) / 1e12;
  printf("mode %d  CTAs/SM %d : %6.2f TFLOP/s  = %.3f of %.1f\n", MODE, ctas_per_sm, tf, tf / peak, peak);
}

int main() {
  int dev = 0, sms = 0, khz = 0;
  CK(cudaGetDevice(&dev));
  CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  CK(cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, dev));
  const double peak = sms * 128.0 * 2.0 * khz * 1e3 / 1e12;
  float *g, *out;
  CK(cudaMalloc(&g, 1 << 20));
  CK(cudaMemset(g, 0, 1 << 20));
  CK(cudaMalloc(&out, (size_t)sms * 2 * 256 * 4));
  printf("%d SMs, %d MHz, FP32 peak %.2f TFLOP/s\n", sms, khz / 1000, peak);
  for (int c = 1; c <= 2; ++c) {
    run<0>(c, sms, g, out, peak);
    run<1>(c, sms, g, out, peak);
    run<2>(c, sms, g, out, peak);
    run<3>(c, sms, g, out, peak);
    run<4>(c, sms, g, out, peak);
  }
  for (int c = 1; c <= 2; ++c) {
    run16<1>(c, sms, g, out, peak);
    run16<2>(c, sms, g, out, peak);
    run16<3>(c, sms, g, out, peak);
  }
  return 0;
}
