// Issue-rate ceiling of tcgen05.mma.cta_group::2.kind::f16 M256 N128 K16 with both operands in shared memory:
// every CTA pair issues `n` MMAs on fixed (zero) operands into one accumulator, commits once, waits.
#include <cstdio>
#include <cuda_runtime.h>
#include "../../vae-extent-search_b200/csrc/tc_ptx.cuh"
using namespace vaex;
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(128, 1) k(int n, int N, long long* out, int ksteps) {
  extern __shared__ uint8_t smem[];
  const uint32_t sbase = ptx::smem_u32(smem);
  const uint32_t rank = ptx::cluster_ctarank();
  const int warp = threadIdx.x >> 5;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + 131072);
  uint32_t* slot = reinterpret_cast<uint32_t*>(smem + 131072 + 64);
  for (int i = threadIdx.x; i < 131072 / 4; i += 128) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  if (threadIdx.x == 0) { ptx::mbar_init(ptx::smem_u32(bars), 1); ptx::fence_mbar_init(); }
  if (warp == 1) ptx::tmem_alloc_2cta(ptx::smem_u32(slot), 512);
  ptx::fence_proxy_async(); ptx::tc_fence_before(); __syncthreads(); ptx::cluster_sync(); ptx::tc_fence_after();
  const uint32_t tmem = *slot;
  if (warp == 0 && rank == 0) {
    if (ptx::elect_one()) {
      const uint32_t idesc = ptx::umma_idesc_f16(256, N);
      const uint64_t dA = ptx::umma_desc_sw128(sbase), dB = ptx::umma_desc_sw128(sbase + 65536);
      const long long t0 = clock64();
      for (int i = 0; i < n; ++i) {
        const int kk = i % ksteps;     // walk K like a real layer: 4 k-steps per atom
        ptx::umma_f16_2cta(tmem + (uint32_t)((i / ksteps) & 1) * 128u, dA + (uint64_t)(((kk >> 2) * 16384 + (kk & 3) * 32) >> 4),
                           dB + (uint64_t)(((kk >> 2) * (N / 2) * 128 + (kk & 3) * 32) >> 4), idesc, kk > 0);
      }
      ptx::umma_commit_2cta(ptx::smem_u32(bars), 1);
      ptx::mbar_wait(ptx::smem_u32(bars), 0);
      out[blockIdx.x >> 1] = clock64() - t0;
    }
    __syncwarp();
  }
  ptx::tc_fence_before(); __syncthreads(); ptx::cluster_sync();
  if (warp == 1) ptx::tmem_dealloc_2cta(tmem, 512);
}
int main() {
  long long* d; cudaMalloc(&d, 8 * 128);
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 131072 + 128);
  for (int N : {128, 64, 256}) for (int pairs : {1, 74}) for (int ks : {16, 4}) {
    const int n = 8192;
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    k<<<2 * pairs, 128, 131072 + 128>>>(n, N, d, ks); cudaDeviceSynchronize();
    cudaEventRecord(a);
    for (int r = 0; r < 20; ++r) k<<<2 * pairs, 128, 131072 + 128>>>(n, N, d, ks);
    cudaEventRecord(b); cudaDeviceSynchronize();
    float ms; cudaEventElapsedTime(&ms, a, b); ms /= 20;
    long long h[128]; cudaMemcpy(h, d, 8 * pairs, cudaMemcpyDeviceToHost);
    const double flop = 2.0 * 256 * N * 16 * n * pairs;
    printf("N=%d pairs=%d ksteps=%d: %.1f cycles/MMA (pair 0), %.3f ms, %.0f TFLOP/s  err=%s\n", N, pairs, ks, (double)h[0] / n, ms, flop / ms / 1e9,
           cudaGetErrorString(cudaGetLastError()));
  }
  return 0;
}
