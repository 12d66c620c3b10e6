// Probe 6: cp.async.bulk global -> shared throughput per SM: D copies of `bytes` in flight on a ring of D slots, `grid` CTAs
// all reading the SAME 2 MB of global memory (as the eval kernel's weight stream does), or each its own region (own = 1).
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../oak_b200/csrc/umma.cuh"

__global__ void probe(long long* out, const uint8_t* g, int depth, int bytes, int iters, int own) {
  extern __shared__ __align__(1024) uint8_t dsm[];
  __shared__ __align__(8) uint64_t bars[8];
  const uint32_t smem = umma::smem_u32(dsm);
  if (threadIdx.x == 0) {
    for (int i = 0; i < 8; i++) umma::mbar_init(umma::smem_u32(&bars[i]), 1);
    umma::mbar_fence_init();
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    const uint8_t* base = g + (own ? (size_t)blockIdx.x * (2u << 20) : 0);
    const int per = (2 << 20) / bytes;
    const long long t0 = clock64();
    for (int i = 0; i < iters + depth; i++) {
      const int s = i % depth;
      if (i >= depth) umma::mbar_wait(umma::smem_u32(&bars[s]), ((i / depth) - 1) & 1);
      if (i < iters) {
        umma::mbar_expect_tx(umma::smem_u32(&bars[s]), bytes);
        umma::bulk_g2s(smem + s * bytes, base + (size_t)(i % per) * bytes, bytes, umma::smem_u32(&bars[s]));
      }
    }
    out[blockIdx.x] = clock64() - t0;
  }
}

int main(int argc, char** argv) {
  long long* out;
  uint8_t* g;
  cudaMallocManaged(&out, 256 * sizeof(long long));
  cudaMalloc(&g, (size_t)150 * (2u << 20));
  cudaMemset(g, 1, (size_t)150 * (2u << 20));
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 131072);
  const int iters = 512;
  for (int own = 0; own < 2; own++)
    for (int grid : {1, 148})
      for (int bytes : {32768, 16384, 8192})
        for (int depth : {1, 2, 4}) {
          if (depth * bytes > 131072) continue;
          for (int rep = 0; rep < 2; rep++) {
            probe<<<grid, 32, 131072>>>(out, g, depth, bytes, iters, own);
            if (cudaDeviceSynchronize() != cudaSuccess) { printf("error\n"); return 1; }
          }
          long long worst = 0;
          for (int i = 0; i < grid; i++) worst = out[i] > worst ? out[i] : worst;
          printf("own %d grid %3d bytes %5d depth %d: %.1f B/cycle per SM (%.0f cycles per copy)\n", own, grid, bytes, depth,
                 (double)iters * bytes / worst, (double)worst / iters);
        }
  return 0;
}
