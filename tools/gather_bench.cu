// Microbenchmark: how fast can one SM fill shared memory with GATHERED 128-byte row chunks?
//
// The tensor-core sparse-conv kernel (csrc/conv_tf32.cu) stages, per k-step, 128 gathered rows x
// 128 B (one 32-channel fp32 chunk of each neighbour row) with cp.async from 128 producer threads
// and is limited by that fill.  This tool measures the fill alone, without
// MMA or barriers: persistent CTAs, each thread issues `PER` 16-byte cp.async per group (8 lanes
// cover one 128 B row chunk, rows chosen at random from a table of `table_rows` rows x `row_bytes`),
// keeps `G` groups in flight (cp.async.wait_group G-1) and overwrites a ring of shared-memory stages.
// Output: GB/s per SM for a sweep of (CTAs per SM, threads, groups in flight, table size).
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o gather_bench tools/gather_bench.cu && ./gather_bench
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s:%d %s\n", __FILE__, __LINE__, cudaGetErrorString(e)); exit(1); } } while (0)

__device__ __forceinline__ void cp16(uint32_t dst, const void* src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
template <int N> __device__ __forceinline__ void wait_group() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N) : "memory"); }

// G groups in flight, PER 16-byte copies per thread per group
template <int G, int PER>
__global__ void gather_kernel(const char* __restrict__ table, const int32_t* __restrict__ idx, int64_t n_idx,
                              int row_bytes, int iters, unsigned long long* __restrict__ sink) {
  extern __shared__ __align__(128) unsigned char smem[];
  const int tid = threadIdx.x, nthr = blockDim.x;
  const int chunk = tid & 7;               // 16-byte chunk of the 128-byte row piece
  const int rows_per_group = (nthr / 8) * PER;
  const uint32_t sbase = (uint32_t)__cvta_generic_to_shared(smem);
  const int stage_bytes = rows_per_group * 128;
  int64_t cursor = ((int64_t)blockIdx.x * 7919) % n_idx;
  for (int it = 0; it < iters; ++it) {
    const int stage = it % (G + 1);
#pragma unroll
    for (int j = 0; j < PER; ++j) {
      const int r = (tid >> 3) + (nthr / 8) * j;
      int64_t q = cursor + r;
      if (q >= n_idx) q -= n_idx;
      const int32_t row = __ldg(idx + q);
      cp16(sbase + stage * stage_bytes + r * 128 + chunk * 16, table + (int64_t)row * row_bytes + chunk * 16);
    }
    commit();
    wait_group<G - 1>();
    cursor += rows_per_group;
    if (cursor >= n_idx) cursor -= n_idx;
  }
  wait_group<0>();
  __syncthreads();
  if (tid == 0) sink[blockIdx.x] = smem[0];
}

template <int G, int PER>
static void run(const char* table, const int32_t* idx, int64_t n_idx, int row_bytes, int threads, int ctas_per_sm,
                int sms, unsigned long long* sink, const char* label) {
  const int rows_per_group = (threads / 8) * PER;
  const size_t smem = (size_t)(G + 1) * rows_per_group * 128;
  if (smem * ctas_per_sm > 220 * 1024) return;
  CK(cudaFuncSetAttribute(gather_kernel<G, PER>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int occ = 0;
  CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, gather_kernel<G, PER>, threads, smem));
  if (occ < ctas_per_sm) return;
  const int iters = 4000;
  const int grid = sms * ctas_per_sm;
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  gather_kernel<G, PER><<<grid, threads, smem>>>(table, idx, n_idx, row_bytes, 200, sink);
  CK(cudaDeviceSynchronize());
  CK(cudaEventRecord(e0));
  gather_kernel<G, PER><<<grid, threads, smem>>>(table, idx, n_idx, row_bytes, iters, sink);
  CK(cudaEventRecord(e1));
  CK(cudaDeviceSynchronize());
  float ms = 0;
  CK(cudaEventElapsedTime(&ms, e0, e1));
  const double bytes = (double)grid * iters * rows_per_group * 128.0;
  printf("%-10s ctas/SM %d threads %4d per-thread copies %2d groups in flight %d  stage %3d KB  in flight/SM %4d KB : "
         "%7.1f GB/s per SM  %6.2f TB/s total\n",
         label, ctas_per_sm, threads, PER, G, rows_per_group * 128 / 1024,
         (int)(ctas_per_sm * G * rows_per_group * 128 / 1024), bytes / (ms * 1e-3) / 1e9 / sms, bytes / (ms * 1e-3) / 1e12);
}

int main() {
  int dev = 0, sms = 0;
  CK(cudaGetDevice(&dev));
  CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  const int row_bytes = 512;  // 128 fp32 channels; one 128-byte chunk of a row is gathered per copy group
  unsigned long long* sink;
  CK(cudaMalloc(&sink, 8 * 4096));
  for (int64_t table_rows : {(int64_t)65536, (int64_t)1200000}) {  // 32 MB (L2-resident) / 614 MB (HBM)
    char* table;
    CK(cudaMalloc(&table, table_rows * row_bytes));
    CK(cudaMemset(table, 1, table_rows * row_bytes));
    // neighbour-like indices: runs of 8 consecutive rows at random positions (a surface tile's rows are
    // partly contiguous), plus a fully random set
    for (int contiguous : {8, 1}) {
      const int64_t n_idx = 1 << 22;
      std::vector<int32_t> h(n_idx);
      uint64_t s = 88172645463325252ull;
      for (int64_t i = 0; i < n_idx; i += contiguous) {
        s ^= s << 13; s ^= s >> 7; s ^= s << 17;
        const int64_t r0 = (int64_t)(s % (uint64_t)(table_rows - contiguous));
        for (int j = 0; j < contiguous && i + j < n_idx; ++j) h[i + j] = (int32_t)(r0 + j);
      }
      int32_t* idx;
      CK(cudaMalloc(&idx, n_idx * 4));
      CK(cudaMemcpy(idx, h.data(), n_idx * 4, cudaMemcpyHostToDevice));
      char label[64];
      snprintf(label, sizeof label, "%s/%s", table_rows < 100000 ? "L2" : "HBM", contiguous == 8 ? "run8" : "rand");
      printf("---- table %lld rows x %d B (%s), index runs of %d\n", (long long)table_rows, row_bytes,
             table_rows < 100000 ? "L2-resident" : "larger than L2", contiguous);
      // the kernel's shape: 128 threads x 8 copies = 128 rows (16 KB) per group, 2 groups in flight, 2 CTAs/SM
      run<2, 8>(table, idx, n_idx, row_bytes, 128, 2, sms, sink, label);
      run<2, 8>(table, idx, n_idx, row_bytes, 128, 1, sms, sink, label);
      run<3, 8>(table, idx, n_idx, row_bytes, 128, 2, sms, sink, label);
      run<4, 8>(table, idx, n_idx, row_bytes, 128, 2, sms, sink, label);
      run<2, 4>(table, idx, n_idx, row_bytes, 256, 2, sms, sink, label);
      run<4, 4>(table, idx, n_idx, row_bytes, 256, 2, sms, sink, label);
      run<2, 2>(table, idx, n_idx, row_bytes, 512, 2, sms, sink, label);
      run<4, 2>(table, idx, n_idx, row_bytes, 512, 2, sms, sink, label);
      run<8, 2>(table, idx, n_idx, row_bytes, 512, 2, sms, sink, label);
      run<2, 16>(table, idx, n_idx, row_bytes, 128, 1, sms, sink, label);
      run<3, 16>(table, idx, n_idx, row_bytes, 128, 1, sms, sink, label);
      run<4, 8>(table, idx, n_idx, row_bytes, 256, 1, sms, sink, label);
      run<8, 4>(table, idx, n_idx, row_bytes, 512, 1, sms, sink, label);
      run<8, 2>(table, idx, n_idx, row_bytes, 1024, 1, sms, sink, label);
      CK(cudaFree(idx));
    }
    CK(cudaFree(table));
  }
  return 0;
}
