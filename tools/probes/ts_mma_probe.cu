// Probe: tcgen05.mma with the A operand in Tensor Memory (bf16, M = 128, K = 16): which column offsets are legal and how the
// K elements are packed. D[r][n] = sum_k A[r][k] B[n][k] with B = identity on the first 16 rows, so D[r][n] = A[r][n].
// nvcc -gencode arch=compute_100a,code=sm_100a -o ts_probe tools/probes/ts_mma_probe.cu && ./ts_probe <a_col> <d_col>
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../oak_b200/csrc/umma.cuh"

__device__ uint16_t f2bf(float f) { return (uint16_t)(__float_as_uint(f) >> 16); }

__global__ void probe(uint32_t a_col, uint32_t d_col, float* out) {
  extern __shared__ __align__(1024) uint8_t dsm[];
  __shared__ uint32_t tmem_slot;
  __shared__ __align__(8) uint64_t bar;
  const uint32_t smem = umma::smem_u32(dsm);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int N = 32;
  // B[n][k], K-major no swizzle: [K/8][N][8 elements] 16 bytes per row per core column
  uint16_t* B = reinterpret_cast<uint16_t*>(dsm);
  for (int i = threadIdx.x; i < 2 * N * 8; i += blockDim.x) {
    const int kc = i / (N * 8), n = (i / 8) % N, k = kc * 8 + (i & 7);
    B[i] = f2bf(n == k ? 1.0f : 0.0f);
  }
  if (warp == 0) umma::tmem_alloc(umma::smem_u32(&tmem_slot), 512);
  if (threadIdx.x == 0) { umma::mbar_init(umma::smem_u32(&bar), 1); umma::mbar_fence_init(); }
  umma::fence_proxy_async();
  __syncthreads();
  umma::tc_fence_after();
  const uint32_t tmem = *reinterpret_cast<volatile uint32_t*>(&tmem_slot);
  const uint32_t tl = tmem + ((uint32_t)((warp & 3) * 32) << 16);
  // A[r][k] = r % 8 + k / 16 (exact in bf16); column c of lane r holds (k = 2c) | (k = 2c + 1) << 16
  uint32_t v[16];
  for (int c = 0; c < 16; c++) {
    const float x0 = (float)(threadIdx.x % 8) + (float)(2 * c) / 16.0f, x1 = (float)(threadIdx.x % 8) + (float)(2 * c + 1) / 16.0f;
    v[c] = (uint32_t)f2bf(x0) | ((uint32_t)f2bf(x1) << 16);
  }
  umma::tmem_st16(tl + a_col, v);
  umma::tmem_st_wait();
  umma::tc_fence_before();
  __syncthreads();
  umma::tc_fence_after();
  if (threadIdx.x == 0) {
    const uint64_t db = umma::smem_desc(smem, N * 16, 128);
    umma::mma_bf16_ts(tmem + d_col, tmem + a_col, db, umma::idesc_bf16_f32(N), 0u);
    umma::mma_commit(umma::smem_u32(&bar));
  }
  umma::mbar_wait(umma::smem_u32(&bar), 0);
  umma::tc_fence_after();
  uint32_t d[16];
  umma::tmem_ld16(tl + d_col, d);
  umma::tmem_ld_wait();
  for (int c = 0; c < 16; c++) out[threadIdx.x * 16 + c] = __uint_as_float(d[c]);
  __syncthreads();
  if (warp == 0) umma::tmem_dealloc(tmem, 512);
}

int main(int argc, char** argv) {
  const uint32_t a_col = argc > 1 ? atoi(argv[1]) : 256, d_col = argc > 2 ? atoi(argv[2]) : 0;
  float* out;
  cudaMallocManaged(&out, 128 * 16 * 4);
  probe<<<1, 128, 4096>>>(a_col, d_col, out);
  cudaError_t e = cudaDeviceSynchronize();
  printf("a_col %u d_col %u: %s\n", a_col, d_col, cudaGetErrorString(e));
  if (e == cudaSuccess)
    for (int r : {0, 3, 37, 127}) {
      printf("  row %3d:", r);
      for (int c = 0; c < 16; c++) printf(" %.4g", out[r * 16 + c]);
      printf("\n");
    }
  return 0;
}
