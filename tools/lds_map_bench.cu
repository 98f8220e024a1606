// Microbenchmark: which thread arrangement of the exact conv kernel's 128 x 128 register-tile loop gets the most
// out of the FP32 pipe when the operands come from shared memory?  Every variant issues the same FFMA2 count
// (64 accumulators per thread, 256 threads, 2 CTAs per SM); they differ in how the 32 lanes of a warp are laid
// over rows (LY) and columns (LX) and in the thread tile RT x CT, i.e. in how many distinct addresses one
// LDS.128 of the gathered-row tile (A, [128][36]) and of the weight tile (B, [32][128]) touches:
//     RT CT LY LX
//      8  8  4  8   the shipped mapping: A 4 addresses / warp, B 8 quads x 4 copies
//     16  4  1 32   all lanes on one row set: A pure broadcast, B 32 distinct quads
//      8  8  2 16   A 2 addresses, B 16 quads x 2 copies
//      4 16 32  1   lane = row: A 32 distinct rows, B pure broadcast
//      8  8  8  4   A 8 addresses, B 4 quads x 8 copies
//      8  8 16  2
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o lds_map_bench tools/lds_map_bench.cu && ./lds_map_bench
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
constexpr int TM = 128, TK = 32, TN = 128, AP = TK + 4, ST = 3;
struct Smem {
  float a[ST][TM][AP];
  float b[ST][TK][TN];
};

template <int RT, int CT, int LY, int LX, int SYNC>
__global__ void __launch_bounds__(256, 2) loop_kernel(int steps, float* __restrict__ out) {
  static_assert(LY * LX == 32 && RT * CT == 64 && CT % 4 == 0, "tile");
  constexpr int WROWS = RT * LY, WCOLS = CT * LX, WX = TN / WCOLS;
  static_assert((TM / WROWS) * WX == 8, "8 warps must cover the tile");
  extern __shared__ __align__(16) unsigned char raw[];
  Smem& sm = *reinterpret_cast<Smem*>(raw);
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const int lx = lane % LX, ly = lane / LX;
  const int rbase = (wid / WX) * WROWS + ly, cbase = (wid % WX) * WCOLS + lx * 4;
  for (int i = tid; i < (int)(sizeof(Smem) / 4); i += 256) reinterpret_cast<float*>(raw)[i] = 1.0f + (i & 7) * 0.125f;
  __syncthreads();
  unsigned long long acc[RT][CT / 2];
#pragma unroll
  for (int i = 0; i < RT; ++i)
#pragma unroll
    for (int c = 0; c < CT / 2; ++c) acc[i][c] = 0ull;
  for (int s = 0; s < steps; ++s) {
    const int st = s % ST;
    if (SYNC) __syncthreads();
#pragma unroll
    for (int kv = 0; kv < TK / 4; ++kv) {
      float a[RT][4];
#pragma unroll
      for (int i = 0; i < RT; ++i) {
        const float4 v = *reinterpret_cast<const float4*>(&sm.a[st][rbase + LY * i][kv * 4]);
        a[i][0] = v.x; a[i][1] = v.y; a[i][2] = v.z; a[i][3] = v.w;
      }
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        unsigned long long bp[CT / 2];
#pragma unroll
        for (int q = 0; q < CT / 4; ++q) {
          const float4 v = *reinterpret_cast<const float4*>(&sm.b[st][kv * 4 + e][cbase + q * (LX * 4)]);
          bp[2 * q] = pack2(v.x, v.y);
          bp[2 * q + 1] = pack2(v.z, v.w);
        }
#pragma unroll
        for (int i = 0; i < RT; ++i) {
          const unsigned long long aa = pack2(a[i][e], a[i][e]);
#pragma unroll
          for (int c = 0; c < CT / 2; ++c) ffma2(acc[i][c], aa, bp[c]);
        }
      }
    }
  }
  unsigned long long s = 0;
#pragma unroll
  for (int i = 0; i < RT; ++i)
#pragma unroll
    for (int c = 0; c < CT / 2; ++c) s ^= acc[i][c];
  if (s == 0x1234567ull) out[blockIdx.x * 256 + tid] = 1.f;
}

template <int RT, int CT, int LY, int LX, int SYNC>
void run(float* out, int sms, double peak) {
  auto k = loop_kernel<RT, CT, LY, LX, SYNC>;
  CK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(Smem)));
  int occ = 0;
  CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k, 256, sizeof(Smem)));
  const int steps = 4000, grid = sms * 2;
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0));
  CK(cudaEventCreate(&e1));
  k<<<grid, 256, sizeof(Smem)>>>(200, out);
  CK(cudaDeviceSynchronize());
  CK(cudaEventRecord(e0));
  k<<<grid, 256, sizeof(Smem)>>>(steps, out);
  CK(cudaEventRecord(e1));
  CK(cudaDeviceSynchronize());
  float ms = 0;
  CK(cudaEventElapsedTime(&ms, e0, e1));
  const double fl = 2.0 * TM * TN * TK * (double)steps * grid;
  printf("RT %2d CT %2d LY %2d LX %2d sync %d  CTAs/SM %d : %6.2f TFLOP/s = %.3f of %.1f\n", RT, CT, LY, LX, SYNC, occ,
         fl / (ms * 1e-3) / 1e12, fl / (ms * 1e-3) / 1e12 / peak, peak);
}


// Software-pipelined loop, AV k-values per step: the operands of step s + 1 (8 rows x AV of A, AV x 8 of B) are
// requested before the FFMA2 block of step s is issued, into a second register set, so that no warp waits for
// shared-memory latency at a step boundary (all warps of an SM sub-partition run the same code at the same pace
// and reach the boundary together).
template <int AV, int SYNC>
__global__ void __launch_bounds__(256, 2) piped_kernel(int steps, float* __restrict__ out) {
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
  constexpr int NS = TK / AV;  // steps per stage
  float a[2][8][AV];
  unsigned long long b[2][AV][4];
  auto load = [&](int buf, int st, int kv) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const float* ap = &sm.a[st][rbase + 4 * i][kv * AV];
      if (AV == 4) {
        const float4 v = *reinterpret_cast<const float4*>(ap);
        a[buf][i][0] = v.x; a[buf][i][1 % AV] = v.y; a[buf][i][2 % AV] = v.z; a[buf][i][3 % AV] = v.w;
      } else if (AV == 2) {
        const float2 v = *reinterpret_cast<const float2*>(ap);
        a[buf][i][0] = v.x; a[buf][i][1 % AV] = v.y;
      } else {
        a[buf][i][0] = *ap;
      }
    }
#pragma unroll
    for (int e = 0; e < AV; ++e) {
      const float4 v0 = *reinterpret_cast<const float4*>(&sm.b[st][kv * AV + e][cbase]);
      const float4 v1 = *reinterpret_cast<const float4*>(&sm.b[st][kv * AV + e][cbase + 32]);
      b[buf][e][0] = pack2(v0.x, v0.y); b[buf][e][1] = pack2(v0.z, v0.w);
      b[buf][e][2] = pack2(v1.x, v1.y); b[buf][e][3] = pack2(v1.z, v1.w);
    }
  };
  load(0, 0, 0);
  for (int s = 0; s < steps; ++s) {
    const int st = s % ST;
    if (SYNC) __syncthreads();
#pragma unroll
    for (int kv = 0; kv < NS; ++kv) {
      const int cur = kv & 1;
      if (kv + 1 < NS) load(cur ^ 1, st, kv + 1);
      else load(cur ^ 1, (s + 1) % ST, 0);
#pragma unroll
      for (int e = 0; e < AV; ++e)
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const unsigned long long aa = pack2(a[cur][i][e], a[cur][i][e]);
#pragma unroll
          for (int c = 0; c < 4; ++c) ffma2(acc[i][c], aa, b[cur][e][c]);
        }
    }
  }
  unsigned long long x = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int c = 0; c < 4; ++c) x ^= acc[i][c];
  if (x == 0x1234567ull) out[blockIdx.x * 256 + tid] = 1.f;
}

template <int AV, int SYNC>
void run_piped(float* out, int sms, double peak) {
  auto k = piped_kernel<AV, SYNC>;
  CK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(Smem)));
  int occ = 0;
  CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k, 256, sizeof(Smem)));
  const int steps = 4000, grid = sms * 2;
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0));
  CK(cudaEventCreate(&e1));
  k<<<grid, 256, sizeof(Smem)>>>(200, out);
  CK(cudaDeviceSynchronize());
  CK(cudaEventRecord(e0));
  k<<<grid, 256, sizeof(Smem)>>>(steps, out);
  CK(cudaEventRecord(e1));
  CK(cudaDeviceSynchronize());
  float ms = 0;
  CK(cudaEventElapsedTime(&ms, e0, e1));
  const double fl = 2.0 * TM * TN * TK * (double)steps * grid;
  printf("piped AV %d sync %d  CTAs/SM %d : %6.2f TFLOP/s = %.3f of %.1f\n", AV, SYNC, occ, fl / (ms * 1e-3) / 1e12,
         fl / (ms * 1e-3) / 1e12 / peak, peak);
}

// Row-block structured loop: a thread owns 2 rows in each of the tile's four 32-row blocks; per 4-channel step the
// weights are loaded once (8 LDS.128) and each block that has the offset (runtime `bits`, CTA-uniform) loads its
// 2 rows and issues 32 FFMA2.  bits = 0xF does the full tile; fewer bits skip whole blocks.
template <int SYNC>
__global__ void __launch_bounds__(256, 2) rb_kernel(int steps, unsigned bits, float* __restrict__ out) {
  extern __shared__ __align__(16) unsigned char raw[];
  Smem& sm = *reinterpret_cast<Smem*>(raw);
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const int tx = lane & 7, ty = lane >> 3;
  const int r0 = (wid >> 1) * 8 + ty, cbase = (wid & 1) * 64 + tx * 4;
  for (int i = tid; i < (int)(sizeof(Smem) / 4); i += 256) reinterpret_cast<float*>(raw)[i] = 1.0f + (i & 7) * 0.125f;
  __syncthreads();
  unsigned long long acc[8][4];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int c = 0; c < 4; ++c) acc[i][c] = 0ull;
  for (int s = 0; s < steps; ++s) {
    const int st = s % ST;
    if (SYNC) __syncthreads();
#pragma unroll
    for (int kv = 0; kv < TK / 4; ++kv) {
      unsigned long long bp[4][4];
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const float4 v0 = *reinterpret_cast<const float4*>(&sm.b[st][kv * 4 + e][cbase]);
        const float4 v1 = *reinterpret_cast<const float4*>(&sm.b[st][kv * 4 + e][cbase + 32]);
        bp[e][0] = pack2(v0.x, v0.y); bp[e][1] = pack2(v0.z, v0.w);
        bp[e][2] = pack2(v1.x, v1.y); bp[e][3] = pack2(v1.z, v1.w);
      }
#pragma unroll
      for (int rb = 0; rb < 4; ++rb) {
        if (!((bits >> rb) & 1u)) continue;
        float a[2][4];
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          const float4 v = *reinterpret_cast<const float4*>(&sm.a[st][r0 + 32 * rb + 4 * j][kv * 4]);
          a[j][0] = v.x; a[j][1] = v.y; a[j][2] = v.z; a[j][3] = v.w;
        }
#pragma unroll
        for (int e = 0; e < 4; ++e)
#pragma unroll
          for (int j = 0; j < 2; ++j) {
            const unsigned long long aa = pack2(a[j][e], a[j][e]);
#pragma unroll
            for (int c = 0; c < 4; ++c) ffma2(acc[2 * rb + j][c], aa, bp[e][c]);
          }
      }
    }
  }
  unsigned long long x = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int c = 0; c < 4; ++c) x ^= acc[i][c];
  if (x == 0x1234567ull) out[blockIdx.x * 256 + tid] = 1.f;
}

template <int SYNC>
void run_rb(float* out, int sms, double peak, unsigned bits) {
  auto k = rb_kernel<SYNC>;
  CK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(Smem)));
  const int steps = 4000, grid = sms * 2;
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0));
  CK(cudaEventCreate(&e1));
  k<<<grid, 256, sizeof(Smem)>>>(200, bits, out);
  CK(cudaDeviceSynchronize());
  CK(cudaEventRecord(e0));
  k<<<grid, 256, sizeof(Smem)>>>(steps, bits, out);
  CK(cudaEventRecord(e1));
  CK(cudaDeviceSynchronize());
  float ms = 0;
  CK(cudaEventElapsedTime(&ms, e0, e1));
  const double fl = 2.0 * TM * TN * TK * (double)steps * grid * __builtin_popcount(bits) / 4.0;
  printf("row-block loop sync %d bits 0x%x : %6.2f useful TFLOP/s = %.3f of %.1f, %.3f ms (full-tile time ratio %.2f)\n", SYNC,
         bits, fl / (ms * 1e-3) / 1e12, fl / (ms * 1e-3) / 1e12 / peak, peak, ms, 0.0);
}

int main() {
  cudaDeviceProp p;
  CK(cudaGetDeviceProperties(&p, 0));
  int khz = 0;
  CK(cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0));
  const double peak = p.multiProcessorCount * 128.0 * 2.0 * khz * 1e3 / 1e12;
  printf("%d SMs, %d MHz, FP32 peak %.2f TFLOP/s\n", p.multiProcessorCount, khz / 1000, peak);
  float* out;
  CK(cudaMalloc(&out, 1 << 22));
  run<8, 8, 4, 8, 0>(out, p.multiProcessorCount, peak);
  run<8, 8, 4, 8, 1>(out, p.multiProcessorCount, peak);
  run<16, 4, 1, 32, 0>(out, p.multiProcessorCount, peak);
  run<16, 4, 1, 32, 1>(out, p.multiProcessorCount, peak);
  run<8, 8, 2, 16, 0>(out, p.multiProcessorCount, peak);
  run<8, 8, 2, 16, 1>(out, p.multiProcessorCount, peak);
  run<4, 16, 32, 1, 0>(out, p.multiProcessorCount, peak);
  run<4, 16, 32, 1, 1>(out, p.multiProcessorCount, peak);
  run<8, 8, 8, 4, 0>(out, p.multiProcessorCount, peak);
  run<8, 8, 16, 2, 0>(out, p.multiProcessorCount, peak);
  run<16, 4, 2, 16, 0>(out, p.multiProcessorCount, peak);
  run<4, 16, 16, 2, 0>(out, p.multiProcessorCount, peak);
  for (unsigned bits : {0xFu, 0x7u, 0x5u, 0x1u}) run_rb<1>(out, p.multiProcessorCount, peak, bits);
  run_rb<0>(out, p.multiProcessorCount, peak, 0xFu);
  run_piped<4, 0>(out, p.multiProcessorCount, peak);
  run_piped<2, 0>(out, p.multiProcessorCount, peak);
  run_piped<2, 1>(out, p.multiProcessorCount, peak);
  run_piped<1, 0>(out, p.multiProcessorCount, peak);
  run_piped<1, 1>(out, p.multiProcessorCount, peak);
  return 0;
}
