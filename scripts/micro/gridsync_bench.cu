// Micro-benchmark (not part of the product): cost of a grid-wide barrier on B200, cooperative_groups vs hand-rolled variants.
#include <cooperative_groups.h>
#include <cstdio>
#include <cuda_runtime.h>
namespace cg = cooperative_groups;

__global__ void k_cg(int iters, float* out) {
  cg::grid_group grid = cg::this_grid();
  float acc = 0;
  for (int i = 0; i < iters; i++) { acc += 1.0f; grid.sync(); }
  if (threadIdx.x == 0 && blockIdx.x == 0) *out = acc;
}

// sense-reversing counter barrier: one atomic per CTA, everyone polls one word
__device__ __forceinline__ void bar_counter(unsigned* counter, unsigned* gen, unsigned nblocks, unsigned& localGen) {
  __syncthreads();
  if (threadIdx.x == 0) {
    localGen++;
    __threadfence();
    unsigned old = atomicAdd(counter, 1u);
    if (old == nblocks - 1) { *counter = 0; __threadfence(); atomicExch(gen, localGen); }
    else { while (*(volatile unsigned*)gen != localGen) {} }
    __threadfence();
  }
  __syncthreads();
}
__global__ void k_counter(int iters, unsigned* counter, unsigned* gen, float* out) {
  unsigned localGen = 0; float acc = 0;
  for (int i = 0; i < iters; i++) { acc += 1.0f; bar_counter(counter, gen, gridDim.x, localGen); }
  if (threadIdx.x == 0 && blockIdx.x == 0) *out = acc;
}

// release/acquire PTX version: red.release + ld.acquire polling, monotonically increasing target (no reset)
__device__ __forceinline__ void bar_relacq(unsigned* counter, unsigned nblocks, unsigned& target) {
  __syncthreads();
  if (threadIdx.x == 0) {
    target += nblocks;
    asm volatile("red.release.gpu.global.add.u32 [%0], 1;" :: "l"(counter) : "memory");
    unsigned v;
    do { asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(counter) : "memory"); } while ((int)(v - target) < 0);
  }
  __syncthreads();
}
__global__ void k_relacq(int iters, unsigned* counter, float* out) {
  unsigned target = 0; float acc = 0;
  for (int i = 0; i < iters; i++) { acc += 1.0f; bar_relacq(counter, gridDim.x, target); }
  if (threadIdx.x == 0 && blockIdx.x == 0) *out = acc;
}

// flags: each CTA writes its own flag (no atomic contention); CTA 0's warp(s) gather all flags then release one word
__device__ __forceinline__ void bar_flags(volatile unsigned* flags, volatile unsigned* go, unsigned nblocks, unsigned& ep) {
  __syncthreads();
  ep++;
  if (blockIdx.x == 0) {
    if (threadIdx.x < nblocks && threadIdx.x > 0) { while (flags[threadIdx.x * 32] != ep) {} }
    __syncthreads();
    if (threadIdx.x == 0) { __threadfence(); *go = ep; }
  } else {
    if (threadIdx.x == 0) { __threadfence(); flags[blockIdx.x * 32] = ep; while (*go != ep) {} __threadfence(); }
  }
  __syncthreads();
}
__global__ void k_flags(int iters, unsigned* flags, unsigned* go, float* out) {
  unsigned ep = 0; float acc = 0;
  for (int i = 0; i < iters; i++) { acc += 1.0f; bar_flags(flags, go, gridDim.x, ep); }
  if (threadIdx.x == 0 && blockIdx.x == 0) *out = acc;
}

int main() {
  int dev = 0; cudaSetDevice(dev);
  cudaDeviceProp p; cudaGetDeviceProperties(&p, dev);
  int nb = p.multiProcessorCount, iters = 2000;
  float* out; cudaMalloc(&out, 4);
  unsigned* buf; cudaMalloc(&buf, 1 << 20); cudaMemset(buf, 0, 1 << 20);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int threads : {256, 512}) {
    for (int variant = 0; variant < 4; variant++) {
      float best = 1e9;
      for (int rep = 0; rep < 3; rep++) {
        cudaMemset(buf, 0, 1 << 20);
        unsigned* counter = buf; unsigned* gen = buf + 64; unsigned* flags = buf + 1024; unsigned* go = buf + 128;
        void* a0[] = {&iters, &out}; void* a1[] = {&iters, &counter, &gen, &out}; void* a2[] = {&iters, &counter, &out}; void* a3[] = {&iters, &flags, &go, &out};
        cudaEventRecord(e0);
        if (variant == 0) cudaLaunchCooperativeKernel((void*)k_cg, dim3(nb), dim3(threads), a0, 0, 0);
        if (variant == 1) cudaLaunchCooperativeKernel((void*)k_counter, dim3(nb), dim3(threads), a1, 0, 0);
        if (variant == 2) cudaLaunchCooperativeKernel((void*)k_relacq, dim3(nb), dim3(threads), a2, 0, 0);
        if (variant == 3) cudaLaunchCooperativeKernel((void*)k_flags, dim3(nb), dim3(threads), a3, 0, 0);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
      }
      const char* names[] = {"cooperative_groups grid.sync", "counter + generation word", "red.release / ld.acquire", "per-CTA flags gathered by CTA 0"};
      printf("threads %d  %-34s %.3f us per barrier  (%s)\n", threads, names[variant], best * 1000.0f / iters, cudaGetErrorString(cudaGetLastError()));
    }
  }
  return 0;
}
