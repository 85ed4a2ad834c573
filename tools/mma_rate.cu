// mma_rate.cu -- microbenchmark: cycles per tcgen05.mma (M=128, K=16, bf16) as a function of N, of whether consecutive
// instructions share an accumulator (dependent chain) or rotate over several, and of where the A operand lives.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/mma_rate tools/mma_rate.cu ; run on a B200.
#include <cstdio>
#include <cuda_runtime.h>
#include "../sketch_rnn_b200/csrc/tc_ptx.cuh"

template <int N, int NACC, bool ATMEM>
__global__ void __launch_bounds__(128, 1) rate_kernel(int iters, long long *out) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t *gen = smem_raw + (base - smem_u32(smem_raw));
  const uint32_t s_a = base, s_b = base + 4 * 16384, s_bar = s_b + 4 * N * 128, holder = s_bar + 16;
  for (int i = threadIdx.x; i < (4 * 16384 + 4 * N * 128) / 4; i += 128) reinterpret_cast<uint32_t *>(gen)[i] = 0;
  if (threadIdx.x == 0) { mbar_init(s_bar, 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
  if (threadIdx.x < 32) tc_alloc(holder, 512);
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *reinterpret_cast<volatile uint32_t *>(gen + (holder - base));
  if (threadIdx.x < 32) {
    if (elect_one()) {
      const uint32_t idesc = make_idesc_bf16(128, N, false, false);
      const uint64_t dA = make_desc(s_a, 0, 1024), dB = make_desc(s_b, 0, 1024);
      const uint32_t ta = tmem + 256;
      long long t0 = clock64();
      for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int kb = 0; kb < 4; ++kb) {
#pragma unroll
          for (int k = 0; k < 4; ++k) {
#pragma unroll
            for (int a = 0; a < NACC; ++a) {
              if (ATMEM) tc_mma_bf16_ts(tmem + a * N, ta + kb * 32 + k * 8, dB + (uint64_t)(kb * (N * 8) + 2 * k), idesc, 1u);
              else tc_mma_bf16(tmem + a * N, dA + (uint64_t)(kb * 1024 + 2 * k), dB + (uint64_t)(kb * (N * 8) + 2 * k), idesc, 1u);
            }
          }
        }
      }
      long long t1 = clock64();
      tc_commit(s_bar);
      mbar_wait(s_bar, 0);
      long long t2 = clock64();
      out[0] = t1 - t0;
      out[1] = t2 - t0;
    }
    __syncwarp();
  }
  tc_fence_before();
  __syncthreads();
  if (threadIdx.x < 32) tc_dealloc(tmem, 512);
}

template <int N, int NACC, bool ATMEM>
void run(const char *name) {
  long long *d, h[2];
  cudaMalloc(&d, 16);
  const int iters = 64;
  size_t smem = 4 * 16384 + 4 * N * 128 + 2048;
  cudaFuncSetAttribute(rate_kernel<N, NACC, ATMEM>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  for (int r = 0; r < 2; ++r) rate_kernel<N, NACC, ATMEM><<<1, 128, smem>>>(iters, d);
  cudaError_t e = cudaDeviceSynchronize();
  cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
  const double n = (double)iters * 16 * NACC;
  printf("%-34s N=%3d acc=%d A=%s: issue %.1f cyc/MMA, complete %.1f cyc/MMA (%s)\n", name, N, NACC, ATMEM ? "tmem" : "smem",
         h[0] / n, h[1] / n, cudaGetErrorString(e));
  cudaFree(d);
}

int main() {
  run<16, 1, true>("chain");   run<16, 2, true>("2 accumulators");  run<16, 4, true>("4 accumulators");
  run<16, 1, false>("chain");  run<16, 4, false>("4 accumulators");
  run<32, 1, true>("chain");   run<32, 4, true>("4 accumulators");
  run<64, 1, true>("chain");   run<64, 2, true>("2 accumulators");
  run<64, 1, false>("chain");
  run<128, 1, true>("chain");  run<128, 1, false>("chain");
  run<256, 1, false>("chain");
  return 0;
}
