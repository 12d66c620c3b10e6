// Probe 4: cycles per tcgen05.mma (M = 128, K = 16, bf16) in the eval kernel's operand layouts: R instructions issued back
// to back by one thread, then one commit; "issue" = clock64 around the issue loop, "total" = until the commit's barrier flips.
//   ./probe4 R        prints SS N=128, SS N=96, SS N=64, TS N=128, TS N=64 (A in shared memory un-swizzled K-major / in TMEM)
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../oak_b200/csrc/umma.cuh"

__global__ void probe(long long* out, int R) {
  extern __shared__ __align__(1024) uint8_t dsm[];
  __shared__ uint32_t tmem_slot;
  __shared__ __align__(8) uint64_t bar;
  const uint32_t smem = umma::smem_u32(dsm);
  const int warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < 65536 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(dsm)[i] = 0;
  if (warp == 0) umma::tmem_alloc(umma::smem_u32(&tmem_slot), 512);
  if (threadIdx.x == 0) { umma::mbar_init(umma::smem_u32(&bar), 1); umma::mbar_fence_init(); }
  umma::fence_proxy_async();
  __syncthreads();
  umma::tc_fence_after();
  const uint32_t tmem = *reinterpret_cast<volatile uint32_t*>(&tmem_slot);
  const uint32_t b = umma::smem_u32(&bar);
  if (warp == 1 && umma::elect_one()) {
    uint32_t phase = 0;
    for (int cfg = 0; cfg < 5; cfg++) {
      const uint32_t N = cfg == 0 || cfg == 3 ? 128u : (cfg == 1 ? 96u : 64u);
      const bool ts = cfg >= 3;
      const uint32_t idesc = umma::idesc_bf16_f32(N);
      const uint64_t da = umma::smem_desc(smem, 2048, 128), db = umma::smem_desc(smem + 32768, N * 16, 128);
      const long long t0 = clock64();
      for (int i = 0; i < R; i++) {
        const uint32_t ks = (uint32_t)(i & 7);  // walk 8 K steps like a real chunk
        if (ts) umma::mma_bf16_ts(tmem + 256, tmem + 8 * ks, db + ks * (N * 32 >> 4), idesc, 1u);
        else umma::mma_bf16(tmem + 256, da + ks * (4096 >> 4), db + ks * (N * 32 >> 4), idesc, 1u);
      }
      const long long t1 = clock64();
      umma::mma_commit(b);
      umma::mbar_wait(b, phase & 1);
      phase++;
      const long long t2 = clock64();
      out[2 * cfg] = t1 - t0;
      out[2 * cfg + 1] = t2 - t0;
    }
  }
  __syncthreads();
  if (warp == 0) umma::tmem_dealloc(tmem, 512);
}

int main(int argc, char** argv) {
  const int R = argc > 1 ? atoi(argv[1]) : 256;
  long long* out;
  cudaMallocManaged(&out, 10 * sizeof(long long));
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
  for (int rep = 0; rep < 2; rep++) {
    probe<<<1, 128, 65536>>>(out, R);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("%s\n", cudaGetErrorString(e)); return 1; }
  }
  const char* names[5] = {"SS N=128", "SS N=96", "SS N=64", "TS N=128", "TS N=64"};
  for (int c = 0; c < 5; c++)
    printf("%-9s R=%d: issue %.1f cycles/MMA, issue+drain %.1f cycles/MMA\n", names[c], R, (double)out[2 * c] / R, (double)out[2 * c + 1] / R);
  return 0;
}
