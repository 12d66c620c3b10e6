// Probe 5: cycles per SS tcgen05.mma (M = 128, N = 128, K = 16, bf16, un-swizzled K-major operands) while the rest of the
// CTA loads the shared-memory port: mode 0 nothing, 1 = 8 warps storing 16 B per lane in a loop, 2 = 8 warps loading,
// 3 = one thread streaming 32 KB bulk copies from global memory into shared memory, 4 = 1 + 3. `pair` = 1: the two MMAs
// of a K step share A through the collector (fill / lastuse).
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../oak_b200/csrc/umma.cuh"

__global__ void probe(long long* out, const uint8_t* g, int R, int mode, int pair) {
  extern __shared__ __align__(1024) uint8_t dsm[];
  __shared__ uint32_t tmem_slot;
  __shared__ __align__(8) uint64_t bar, tbar;
  __shared__ volatile int stop;
  const uint32_t smem = umma::smem_u32(dsm);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < 131072 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(dsm)[i] = 0;
  if (warp == 0) umma::tmem_alloc(umma::smem_u32(&tmem_slot), 512);
  if (threadIdx.x == 0) {
    umma::mbar_init(umma::smem_u32(&bar), 1);
    umma::mbar_init(umma::smem_u32(&tbar), 1);
    umma::mbar_fence_init();
    stop = 0;
  }
  umma::fence_proxy_async();
  __syncthreads();
  umma::tc_fence_after();
  const uint32_t tmem = *reinterpret_cast<volatile uint32_t*>(&tmem_slot);
  if (warp == 1) {
    if (umma::elect_one()) {
      const uint32_t N = 128, idesc = umma::idesc_bf16_f32(N);
      const uint64_t da = umma::smem_desc(smem, 2048, 128), db = umma::smem_desc(smem + 32768, N * 16, 128);
      const long long t0 = clock64();
      for (int i = 0; i < R; i++) {
        const uint32_t ks = (uint32_t)(i & 7);
        if (pair) umma::mma_bf16_pair(tmem + 256, da + ks * 256, db + ks * 256, db + ks * 256 + 16, idesc, 1u);
        else { umma::mma_bf16(tmem + 256, da + ks * 256, db + ks * 256, idesc, 1u); umma::mma_bf16(tmem + 256, da + ks * 256, db + ks * 256 + 16, idesc, 1u); }
      }
      const long long t1 = clock64();
      umma::mma_commit(umma::smem_u32(&bar));
      umma::mbar_wait(umma::smem_u32(&bar), 0);
      const long long t2 = clock64();
      out[0] = t1 - t0;
      out[1] = t2 - t0;
      stop = 1;
    }
  } else if (warp >= 2 && warp < 10) {
    if (mode == 1 || mode == 2 || mode == 4) {
      const uint32_t a = smem + 65536 + (uint32_t)((warp - 2) * 512 + lane * 16);
      uint32_t acc = 0;
      while (!stop) {
        for (int j = 0; j < 16; j++) {
          if (mode == 2) {
            uint32_t x, y, z, w;
            asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(x), "=r"(y), "=r"(z), "=r"(w) : "r"(a + (uint32_t)j * 4096));
            acc += x + y + z + w;
          } else {
            asm volatile("st.shared.v4.u32 [%0], {%1, %1, %1, %1};" ::"r"(a + (uint32_t)j * 4096), "r"(acc) : "memory");
          }
        }
      }
      if (acc == 12345) out[7] = acc;
    }
  } else if (warp == 10 && lane == 0 && (mode == 3 || mode == 4)) {
    uint32_t ph = 0;
    long long copies = 0;
    while (!stop) {
      umma::mbar_expect_tx(umma::smem_u32(&tbar), 32768);
      umma::bulk_g2s(smem + 98304, g + (copies & 31) * 32768, 32768, umma::smem_u32(&tbar));
      umma::mbar_wait(umma::smem_u32(&tbar), ph & 1);
      ph++;
      copies++;
    }
    out[2] = copies;
  }
  __syncthreads();
  if (warp == 0) umma::tmem_dealloc(tmem, 512);
}

int main(int argc, char** argv) {
  const int R = argc > 1 ? atoi(argv[1]) : 512;
  long long* out;
  uint8_t* g;
  cudaMallocManaged(&out, 8 * sizeof(long long));
  cudaMalloc(&g, 1 << 20);
  cudaMemset(g, 0, 1 << 20);
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 131072);
  for (int pair = 0; pair < 2; pair++)
    for (int mode = 0; mode < 5; mode++) {
      out[2] = 0;
      for (int rep = 0; rep < 2; rep++) {
        probe<<<1, 352, 131072>>>(out, g, R, mode, pair);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("%s\n", cudaGetErrorString(e)); return 1; }
      }
      printf("pair %d mode %d: %.1f cycles per MMA issued, %.1f with drain; bulk copies %lld (%.1f B/cycle)\n", pair, mode,
             (double)out[0] / (2 * R), (double)out[1] / (2 * R), out[2], out[2] * 32768.0 / (double)out[1]);
    }
  return 0;
}
