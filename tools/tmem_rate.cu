// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o /tmp/tmem_rate tools/tmem_rate.cu && /tmp/tmem_rate
// Read throughput of tensor memory: W warps of one CTA each issue `iters` tcgen05.ld of a given shape on their own lane
// quarter (back to back, one wait at the end of every 4), bytes per SM clock = W * iters * bytes_per_ld / cycles.
#include <cstdio>
#include <cuda_runtime.h>
#include "../vae-extent-search_b200/csrc/tc_ptx.cuh"
using namespace vaex;
template <int SHAPE>
__device__ __forceinline__ void ld(uint32_t t, uint32_t& sink) {
  if (SHAPE == 32) { uint32_t v[32]; ptx::tmem_ld32(t, v); sink ^= v[0] ^ v[31]; }
  if (SHAPE == 16) { uint32_t v[16]; ptx::tmem_ld16(t, v); sink ^= v[0] ^ v[15]; }
  if (SHAPE == 8) { uint32_t v[8]; ptx::tmem_ld8(t, v); sink ^= v[0] ^ v[7]; }
  if (SHAPE == 64) {
    uint32_t v[64];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x64.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, "
        "%32, %33, %34, %35, %36, %37, %38, %39, %40, %41, %42, %43, %44, %45, %46, %47, %48, %49, %50, %51, %52, %53, %54, %55, %56, %57, %58, %59, %60, %61, %62, %63}, [%64];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
          "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31]),
          "=r"(v[32]), "=r"(v[33]), "=r"(v[34]), "=r"(v[35]), "=r"(v[36]), "=r"(v[37]), "=r"(v[38]), "=r"(v[39]), "=r"(v[40]), "=r"(v[41]), "=r"(v[42]), "=r"(v[43]), "=r"(v[44]), "=r"(v[45]), "=r"(v[46]), "=r"(v[47]),
          "=r"(v[48]), "=r"(v[49]), "=r"(v[50]), "=r"(v[51]), "=r"(v[52]), "=r"(v[53]), "=r"(v[54]), "=r"(v[55]), "=r"(v[56]), "=r"(v[57]), "=r"(v[58]), "=r"(v[59]), "=r"(v[60]), "=r"(v[61]), "=r"(v[62]), "=r"(v[63])
        : "r"(t) : "memory");
    sink ^= v[0] ^ v[63];
  }
}
template <int SHAPE>
__global__ void __cluster_dims__(2, 1, 1) k(int iters, long long* out, uint32_t* sinkp) {
  __shared__ uint32_t slot;
  const int warp = threadIdx.x >> 5;
  if (warp == 0) ptx::tmem_alloc_2cta(ptx::smem_u32(&slot), 512);
  ptx::tc_fence_before(); __syncthreads(); ptx::cluster_sync(); ptx::tc_fence_after();
  const uint32_t t = slot + ((uint32_t)(32 * (warp & 3)) << 16) + (uint32_t)((warp >> 2) * SHAPE) % (512 - SHAPE);
  uint32_t sink = 0;
  __syncthreads();
  const long long t0 = clock64();
  for (int i = 0; i < iters; i += 4) {
    ld<SHAPE>(t, sink); ld<SHAPE>(t, sink); ld<SHAPE>(t, sink); ld<SHAPE>(t, sink);
    ptx::tmem_ld_wait();
  }
  __syncthreads();
  const long long t1 = clock64();
  if (threadIdx.x == 0) out[blockIdx.x] = t1 - t0;
  if (sink == 0xdeadbeef) *sinkp = sink;
  ptx::tc_fence_before(); __syncthreads(); ptx::cluster_sync();
  if (warp == 0) ptx::tmem_dealloc_2cta(slot, 512);
}
template <int SHAPE>
void run(int warps, long long* d, uint32_t* s) {
  const int iters = 4096;
  k<SHAPE><<<2, 32 * warps>>>(iters, d, s); cudaDeviceSynchronize();
  k<SHAPE><<<2, 32 * warps>>>(iters, d, s); cudaDeviceSynchronize();
  long long h[2]; cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
  const double bytes = (double)warps * iters * 32 * SHAPE * 4;
  printf("32x32b.x%-3d warps=%2d: %8lld clk, %.1f B/clk per SM, %.1f clk per ld  (%s)\n", SHAPE, warps, h[0], bytes / h[0], (double)h[0] / iters,
         cudaGetErrorString(cudaGetLastError()));
}
int main() {
  long long* d; uint32_t* s; cudaMalloc(&d, 64); cudaMalloc(&s, 4);
  for (int w : {1, 4, 8, 16}) { run<8>(w, d, s); run<16>(w, d, s); run<32>(w, d, s); run<64>(w, d, s); }
  return 0;
}
