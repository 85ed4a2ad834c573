// Where do the rows of an M = 64 tcgen05.mma accumulator (cta_group::1) land in tensor memory?
// A[r][0] = r (all other elements zero), B[n][0] = 1  ->  D[r][n] = r.  All 128 lanes x 32 columns are dumped.
// Second pass: the same M = 64 product with the accumulator address moved by 16 lanes.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -I sketch_rnn_b200/csrc -o tools/probe_m64 tools/probe_m64.cu -lcuda
#include <cuda.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include "tc_ptx.cuh"

__global__ void __launch_bounds__(128, 1) probe(float *out, int M, int lane_off) {
  extern __shared__ uint8_t raw[];
  const uint32_t base = (smem_u32(raw) + 1023u) & ~1023u;
  uint8_t *gen = raw + (base - smem_u32(raw));
  const uint32_t sA = base, sB = base + 16384, bar = base + 32768, holder = base + 32768 + 16;
  for (int i = threadIdx.x; i < 32768 / 4; i += blockDim.x) reinterpret_cast<uint32_t *>(gen)[i] = 0;
  __syncthreads();
  if (threadIdx.x < M) {
    const int r = threadIdx.x;
    *reinterpret_cast<__nv_bfloat16 *>(gen + r * 128 + ((0 ^ (r & 7)) << 4)) = __float2bfloat16_rn((float)r);
  }
  if (threadIdx.x < 32) {
    const int n = threadIdx.x;
    *reinterpret_cast<__nv_bfloat16 *>(gen + 16384 + n * 128 + ((0 ^ (n & 7)) << 4)) = __float2bfloat16_rn(1.0f);
  }
  if (threadIdx.x == 0) { mbar_init(bar, 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
  if (threadIdx.x < 32) tc_alloc(holder, 32);
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tm = *reinterpret_cast<volatile uint32_t *>(gen + 32768 + 16);
  // poison the accumulator so that lanes the product does not write are recognisable
  {
    uint32_t v[32];
    for (int j = 0; j < 32; ++j) v[j] = __float_as_uint(-1.0f);
    tc_st32(tm + ((uint32_t)(threadIdx.x & ~31) << 16), v);
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (threadIdx.x == 0) {
    const uint32_t idesc = make_idesc_bf16(M, 32, false, false);
    tc_mma_bf16(tm + ((uint32_t)lane_off << 16), make_desc(sA, 0, 1024), make_desc(sB, 0, 1024), idesc, 0u);
    tc_commit(bar);
  }
  mbar_wait(bar, 0);
  tc_fence_after();
  uint32_t v[32];
  tc_ld32(tm + ((uint32_t)(threadIdx.x & ~31) << 16), v);
  for (int j = 0; j < 32; ++j) out[threadIdx.x * 32 + j] = __uint_as_float(v[j]);
  tc_fence_before();
  __syncthreads();
  if (threadIdx.x < 32) tc_dealloc(tm, 32);
}

int main() {
  float *d, h[128 * 32];
  cudaMalloc(&d, sizeof(h));
  for (int pass = 0; pass < 3; ++pass) {
    const int M = pass == 2 ? 128 : 64, lane_off = pass == 1 ? 16 : 0;
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 40000);
    probe<<<1, 128, 40000>>>(d, M, lane_off);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("M=%d: %s\n", M, cudaGetErrorString(e)); return 1; }
    cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
    printf("M=%d, accumulator lane offset %d: lane -> D row (column 0 | column 5)\n", M, lane_off);
    for (int l = 0; l < 128; ++l) printf("%s%3d:%4.0f|%4.0f", (l % 8) ? "  " : "\n", l, h[l * 32], h[l * 32 + 5]);
    printf("\n");
  }
  return 0;
}
