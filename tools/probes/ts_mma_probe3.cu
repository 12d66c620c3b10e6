// Probe 3: the second-layer TS sequence at arbitrary Tensor Memory columns: ./probe3 H_HI H_LO D [warps_st] [extra_commits]
// A hi plane at H_HI + 8 ks, lo plane at H_LO + 8 ks, D (N = 64) at column D; an N = 128 SS MMA at column 0 runs first
// (the first layer), the planes are stored by `warps_st` (4 or 8) warps, 24 TS MMAs are issued by an elected lane of warp 9.
// D[r][n] = A[r][2n] + A_lo[r][2n] + 0.5 A[r][2n+1] with B_hi[n][k] = (k == 2n), B_lo[n][k] = 0.5 (k == 2n + 1).
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../oak_b200/csrc/umma.cuh"

__host__ __device__ inline uint16_t f2bf(float f) {
  union { float f; uint32_t u; } v; v.f = f; return (uint16_t)(v.u >> 16);
}
__host__ __device__ inline float a_hi(int r, int k) { return (float)(r % 8) + (float)k / 256.0f; }
__host__ __device__ inline float a_lo(int r, int k) { return (float)((r + k) % 4) / 1024.0f; }

__global__ void probe(float* out, int H_HI, int H_LO, int DC, int warps_st, int extra) {
  extern __shared__ __align__(1024) uint8_t dsm[];
  __shared__ uint32_t tmem_slot;
  __shared__ __align__(8) uint64_t bars[4];
  const uint32_t smem = umma::smem_u32(dsm);
  const int warp = threadIdx.x >> 5;
  const int N = 64, KS = 8;
  uint16_t* B = reinterpret_cast<uint16_t*>(dsm);  // hi plane [K/8][N][8], then lo plane
  for (int i = threadIdx.x; i < 2 * KS * 2 * N * 8; i += blockDim.x) {
    const int plane = i / (KS * 2 * N * 8), j = i % (KS * 2 * N * 8);
    const int kc = j / (N * 8), n = (j / 8) % N, k = kc * 8 + (j & 7);
    B[i] = plane == 0 ? f2bf(k == 2 * n ? 1.0f : 0.0f) : f2bf(k == 2 * n + 1 ? 0.5f : 0.0f);
  }
  for (int i = threadIdx.x; i < 16384; i += blockDim.x) reinterpret_cast<uint16_t*>(dsm + 32768)[i] = 0;
  if (warp == 0) umma::tmem_alloc(umma::smem_u32(&tmem_slot), 512);
  if (threadIdx.x == 0) { for (int i = 0; i < 4; i++) umma::mbar_init(umma::smem_u32(&bars[i]), 1); umma::mbar_fence_init(); }
  umma::fence_proxy_async();
  __syncthreads();
  umma::tc_fence_after();
  const uint32_t tmem = *reinterpret_cast<volatile uint32_t*>(&tmem_slot);
  const uint32_t bar = umma::smem_u32(&bars[0]);
  if (threadIdx.x < 128) {
    const uint32_t tlq = tmem + ((uint32_t)((warp & 3) * 32) << 16);
    uint32_t sv[16];
    for (int c = 0; c < 16; c++) sv[c] = __float_as_uint(777.0f);
    for (int c0 = 0; c0 < 512; c0 += 16) umma::tmem_st16(tlq + c0, sv);
    umma::tmem_st_wait();
  }
  umma::tc_fence_before();
  __syncthreads();
  umma::tc_fence_after();
  if (warp == 9 && umma::elect_one()) {  // "first layer": N = 128 at column 0
    const uint64_t da = umma::smem_desc(smem + 32768, 2048, 128), db = umma::smem_desc(smem, 128 * 16, 128);
    umma::mma_bf16(tmem + 0, da, db, umma::idesc_bf16_f32(128), 0u);
    umma::mma_commit(bar);
  }
  umma::mbar_wait(bar, 0);
  umma::tc_fence_after();
  if (threadIdx.x < 32 * warps_st) {  // warps 0..3 store K steps 0..3 (columns +0..31), warps 4..7 the rest; or 4 warps store all
    const uint32_t tl = tmem + ((uint32_t)((warp & 3) * 32) << 16);
    const int r = (warp & 3) * 32 + (threadIdx.x & 31);
    const int ks0 = warps_st == 8 ? 4 * (warp >> 2) : 0, ks1 = warps_st == 8 ? ks0 + 4 : 8;
    for (int ks = ks0; ks < ks1; ks += 2) {
      uint32_t h[16], l[16];
      for (int c = 0; c < 16; c++) {
        const int k = 16 * ks + 2 * c;
        h[c] = (uint32_t)f2bf(a_hi(r, k)) | ((uint32_t)f2bf(a_hi(r, k + 1)) << 16);
        l[c] = (uint32_t)f2bf(a_lo(r, k)) | ((uint32_t)f2bf(a_lo(r, k + 1)) << 16);
      }
      umma::tmem_st16(tl + H_HI + 8 * ks, h);
      umma::tmem_st16(tl + H_LO + 8 * ks, l);
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
      const uint32_t a_t = tmem + H_HI + 8 * i;
      umma::mma_bf16_ts(tmem + DC, a_t, dbh, idesc, acc);
      umma::mma_bf16_ts(tmem + DC, a_t, dbl, idesc, 1u);
      umma::mma_bf16_ts(tmem + DC, a_t - H_HI + H_LO, dbh, idesc, 1u);
      acc = 1u;
      dbh += kstep_b >> 4;
      dbl += kstep_b >> 4;
    }
    for (int i = 0; i < extra; i++) umma::mma_commit(umma::smem_u32(&bars[1 + i]));
    umma::mma_commit(bar);
  }
  umma::mbar_wait(bar, 1);
  umma::tc_fence_after();
  if (threadIdx.x < 128) {
    const uint32_t tl = tmem + ((uint32_t)((warp & 3) * 32) << 16);
    uint32_t d[16];
    for (int c0 = 0; c0 < 64; c0 += 16) {
      umma::tmem_ld16(tl + DC + c0, d);
      umma::tmem_ld_wait();
      for (int c = 0; c < 16; c++) out[threadIdx.x * 64 + c0 + c] = __uint_as_float(d[c]);
    }
    for (int c0 = 0; c0 < 512; c0 += 16) {
      umma::tmem_ld16(tl + c0, d);
      umma::tmem_ld_wait();
      if (threadIdx.x == 3)
        for (int c = 0; c < 16; c++) out[128 * 64 + c0 + c] = __uint_as_float(d[c]);
    }
  }
  __syncthreads();
  if (warp == 0) umma::tmem_dealloc(tmem, 512);
}

int main(int argc, char** argv) {
  const int H_HI = argc > 1 ? atoi(argv[1]) : 64, H_LO = argc > 2 ? atoi(argv[2]) : 416, DC = argc > 3 ? atoi(argv[3]) : 128;
  const int warps_st = argc > 4 ? atoi(argv[4]) : 4, extra = argc > 5 ? atoi(argv[5]) : 0;
  float* out;
  cudaMallocManaged(&out, (128 * 64 + 512) * 4);
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
  probe<<<1, 448, 65536>>>(out, H_HI, H_LO, DC, warps_st, extra);
  cudaError_t e = cudaDeviceSynchronize();
  printf("H_HI %d H_LO %d D %d warps %d extra commits %d: %s", H_HI, H_LO, DC, warps_st, extra, cudaGetErrorString(e));
  if (e != cudaSuccess) { printf("\n"); return 1; }
  double worst = 0;
  for (int r = 0; r < 128; r++)
    for (int n = 0; n < 64; n++) {
      auto bf = [](float x) { union { float f; uint32_t u; } v; v.u = (uint32_t)f2bf(x) << 16; return v.f; };
      const double want = (double)bf(a_hi(r, 2 * n)) + (double)bf(a_lo(r, 2 * n)) + 0.5 * (double)bf(a_hi(r, 2 * n + 1));
      worst = fmax(worst, fabs(want - (double)out[r * 64 + n]));
    }
  printf("  max |D - want| = %.3g; row 3 changed columns:", worst);
  int run0 = -1;
  for (int c = 0; c <= 512; c++) {
    const bool changed = c < 512 && out[128 * 64 + c] != 777.0f;
    if (changed && run0 < 0) run0 = c;
    if (!changed && run0 >= 0) { printf(" [%d, %d)", run0, c); run0 = -1; }
  }
  printf("\n");
  return 0;
}
