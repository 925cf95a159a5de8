// Microbenchmark: throughput of shared-memory atomic adds on B200 as the histogram kernel issues them
// (256 threads, 3 CTAs/SM via 72 KB of dynamic shared memory, counters [12800] u32, one atomic per lane).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o smem_atomics smem_atomics.cu ; run: ./smem_atomics
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t lcg(uint32_t& s) { s = s * 1664525u + 1013904223u; return s >> 8; }

template <int MODE>
__global__ void __launch_bounds__(256) k(int iters, uint32_t* sink) {
  extern __shared__ uint32_t sh[];
  const int n = 12800;
  for (int i = threadIdx.x; i < n; i += 256) sh[i] = 0;
  __syncthreads();
  uint32_t s = blockIdx.x * 977u + threadIdx.x * 7919u + 1u;
  const int lane = threadIdx.x & 31;
  uint32_t acc = 0;
  for (int it = 0; it < iters; it++) {
#pragma unroll 8
    for (int u = 0; u < 8; u++) {
      const uint32_t r = lcg(s);
      if (MODE == 0) atomicAdd(&sh[r % n], 1u);                          // random address
      if (MODE == 1) atomicAdd(&sh[((r % 400) * 32 + lane)], 1u);        // conflict-free: bank = lane
      if (MODE == 2) { if (lane < 16) atomicAdd(&sh[r % n], 1u); }       // half the lanes
      if (MODE == 3) { const uint32_t a = r % n; sh[a] = sh[a] + 1u; }   // plain load + store (racy, timing only)
      if (MODE == 4) atomicAdd(&sh[(r % 200) * 50 + (lane % 50)], 1u);   // histogram-like: (bin*C + label)
      if (MODE == 5) acc += sh[r % n];                                   // random LDS only
      if (MODE == 6) { if (lane < 8) atomicAdd(&sh[r % n], 1u); }        // quarter of the lanes
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) sink[blockIdx.x] = sh[1] + acc;
}

template <int MODE>
void run(const char* name, int sms) {
  const int iters = 4000, grid = sms * 3;
  uint32_t* sink;
  cudaMalloc(&sink, grid * 4);
  cudaFuncSetAttribute(k<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, 72 * 1024);
  cudaEvent_t a, b;
  cudaEventCreate(&a); cudaEventCreate(&b);
  k<MODE><<<grid, 256, 72 * 1024>>>(100, sink);
  cudaEventRecord(a);
  k<MODE><<<grid, 256, 72 * 1024>>>(iters, sink);
  cudaEventRecord(b);
  cudaEventSynchronize(b);
  float ms;
  cudaEventElapsedTime(&ms, a, b);
  const double warp_instr_per_sm = (double)iters * 8 * 8 * 3;   // 8 warps x 3 CTAs
  int clk;
  cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
  const double cycles = ms * 1e-3 * clk * 1e3;
  printf("%-44s %.3f ms  %.2f cycles per warp instruction per SM\n", name, ms, cycles / warp_instr_per_sm);
  cudaFree(sink);
}

int main() {
  int sms;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  run<0>("atomicAdd random address, 32 lanes", sms);
  run<1>("atomicAdd conflict-free (bank = lane)", sms);
  run<2>("atomicAdd random address, 16 lanes", sms);
  run<6>("atomicAdd random address, 8 lanes", sms);
  run<4>("atomicAdd (bin*50 + label) pattern", sms);
  run<3>("plain LDS + STS random address", sms);
  run<5>("LDS random address", sms);
  return 0;
}
