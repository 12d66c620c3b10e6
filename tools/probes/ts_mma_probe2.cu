// Probe 2: the eval kernel's second-layer MMA sequence with A in Tensor Memory: N = 64, K = 128 as 8 K-steps whose A columns
// follow tc::hidden_tmem_col, lo plane 32 columns on, B hi / lo planes in shared memory, D at column 128; issued by an
// elected lane of warp 9 of a 448-thread CTA. D[r][n] = sum_k A[r][k] B[n][k] with B[n][k] = (k == 2 n), A[r][k] = r % 8 + k / 256.
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../oak_b200/csrc/umma.cuh"

__device__ uint16_t f2bf(float f) { return (uint16_t)(__float_as_uint(f) >> 16); }
__host__ __device__ inline uint32_t hidden_tmem_col(uint32_t kstep) { return 64u * (kstep >> 2) + 8u * (kstep & 3u); }

__global__ void probe(float* out, int mode) {
  extern __shared__ __align__(1024) uint8_t dsm[];
  __shared__ uint32_t tmem_slot;
  __shared__ __align__(8) uint64_t bar;
  const uint32_t smem = umma::smem_u32(dsm);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int N = 64, KS = 8;
  uint16_t* B = reinterpret_cast<uint16_t*>(dsm);  // hi plane [K/8][N][8], then lo plane (zeros)
  for (int i = threadIdx.x; i < 2 * KS * 2 * N * 8; i += blockDim.x) {
    const int plane = i / (KS * 2 * N * 8), j = i % (KS * 2 * N * 8);
    const int kc = j / (N * 8), n = (j / 8) % N, k = kc * 8 + (j & 7);
    B[i] = plane == 0 ? f2bf(k == 2 * n ? 1.0f : 0.0f) : 0;
  }
  if (warp == 0) umma::tmem_alloc(umma::smem_u32(&tmem_slot), 512);
  if (threadIdx.x == 0) { umma::mbar_init(umma::smem_u32(&bar), 1); umma::mbar_fence_init(); }
  umma::fence_proxy_async();
  __syncthreads();
  umma::tc_fence_after();
  const uint32_t tmem = *reinterpret_cast<volatile uint32_t*>(&tmem_slot);
  if (threadIdx.x < 128) {  // sentinel in every column: what does the sequence below write?
    const uint32_t tlq = tmem + ((uint32_t)((warp & 3) * 32) << 16);
    uint32_t sv[16];
    for (int c = 0; c < 16; c++) sv[c] = __float_as_uint(777.0f);
    for (int c0 = 0; c0 < 512; c0 += 16) umma::tmem_st16(tlq + c0, sv);
    umma::tmem_st_wait();
  }
  umma::tc_fence_before();
  __syncthreads();
  umma::tc_fence_after();
  if (mode >= 2) {  // first an ordinary MMA whose accumulator is the columns the TS operand will live in (N = 128 at column 0)
    if (warp == 9 && umma::elect_one()) {
      const uint64_t da = umma::smem_desc(smem + 32768, 2048, 128), db = umma::smem_desc(smem, 128 * 16, 128);
      umma::mma_bf16(tmem + 0, da, db, umma::idesc_bf16_f32(128), 0u);
      umma::mma_commit(umma::smem_u32(&bar));
    }
    umma::mbar_wait(umma::smem_u32(&bar), 0);
    umma::tc_fence_after();
  }
  if (threadIdx.x < 128) {
    const uint32_t tl = tmem + ((uint32_t)((warp & 3) * 32) << 16);
    for (int ks = 0; ks < KS; ks += 2) {  // two K-steps = 16 columns per store
      uint32_t h[16], l[16];
      for (int c = 0; c < 16; c++) {
        const int k = 16 * ks + 2 * c;
        const float x0 = (float)(threadIdx.x % 8) + (float)k / 256.0f, x1 = (float)(threadIdx.x % 8) + (float)(k + 1) / 256.0f;
        h[c] = (uint32_t)f2bf(x0) | ((uint32_t)f2bf(x1) << 16);
        l[c] = 0;
      }
      umma::tmem_st16(tl + hidden_tmem_col(ks), h);
      umma::tmem_st16(tl + hidden_tmem_col(ks) + 32, l);
    }
    umma::tmem_st_wait();
  }
  umma::tc_fence_before();
  __syncthreads();
  umma::tc_fence_after();
  if (warp == 9 && umma::elect_one()) {
    const uint32_t idesc = umma::idesc_bf16_f32(N);
    const uint32_t kstep_b = N * 32;
    uint64_t dbh = umma::smem_desc(smem, N * 16, 128), dbl = umma::smem_desc(smem + KS * kstep_b, N * 16, 128);
    uint32_t acc = 0;
    for (int i = 0; i < KS; i++) {
      const uint32_t a_t = tmem + hidden_tmem_col(i);
      if (mode == 0 || mode == 2) {
        umma::mma_bf16_ts(tmem + 128, a_t, dbh, idesc, acc);
        umma::mma_bf16_ts(tmem + 128, a_t, dbl, idesc, 1u);
        umma::mma_bf16_ts(tmem + 128, a_t + 32u, dbh, idesc, 1u);
      } else {
        umma::mma_bf16_ts(tmem + 128, a_t, dbh, idesc, acc);
      }
      acc = 1u;
      dbh += kstep_b >> 4;
      dbl += kstep_b >> 4;
    }
    umma::mma_commit(umma::smem_u32(&bar));
  }
  umma::mbar_wait(umma::smem_u32(&bar), mode >= 2 ? 1 : 0);
  umma::tc_fence_after();
  if (threadIdx.x < 128) {
    const uint32_t tl = tmem + ((uint32_t)((warp & 3) * 32) << 16);
    uint32_t d[16];
    umma::tmem_ld16(tl + 128, d);
    umma::tmem_ld_wait();
    for (int c = 0; c < 16; c++) out[threadIdx.x * 16 + c] = __uint_as_float(d[c]);
    for (int c0 = 0; c0 < 512; c0 += 16) {
      umma::tmem_ld16(tl + c0, d);
      umma::tmem_ld_wait();
      if (threadIdx.x == 3)
        for (int c = 0; c < 16; c++) out[2048 + c0 + c] = __uint_as_float(d[c]);
    }
  }
  __syncthreads();
  if (warp == 0) umma::tmem_dealloc(tmem, 512);
}

int main(int argc, char** argv) {
  const int mode = argc > 1 ? atoi(argv[1]) : 0;
  float* out;
  cudaMallocManaged(&out, (128 * 16 + 512) * 4);
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
  probe<<<1, 448, 65536>>>(out, mode);
  cudaError_t e = cudaDeviceSynchronize();
  printf("mode %d: %s\n", mode, cudaGetErrorString(e));
  if (e == cudaSuccess)
    for (int r : {0, 3, 127}) {
      printf("  row %3d (want r%%8 + 2n/256):", r);
      for (int c = 0; c < 16; c++) printf(" %.4g", out[r * 16 + c]);
      printf("\n");
    }
  if (e == cudaSuccess) {
    printf("  row 3, columns that no longer hold the sentinel:");
    int run0 = -1;
    for (int c = 0; c <= 512; c++) {
      const bool changed = c < 512 && out[2048 + c] != 777.0f;
      if (changed && run0 < 0) run0 = c;
      if (!changed && run0 >= 0) { printf(" [%d, %d)", run0, c); run0 = -1; }
    }
    printf("\n");
  }
  return 0;
}
