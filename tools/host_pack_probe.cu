// Probe (not product): can the box's host cores narrow fp32 features to fp16 faster than PCIe moves fp32?
// Prints GB/s of fp32 input consumed by T threads (AVX2/F16C, streaming stores), alone and while an H2D copy
// of the packed halves runs.  Build: nvcc -O3 -Xcompiler "-mavx2 -mf16c -pthread" tools/host_pack_probe.cu vae-extent-search_b200/csrc/build/host_pack.o -o tools/host_pack_probe.bin
#include <cuda_runtime.h>
#include <immintrin.h>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>
namespace vaex { bool host_pack_f16(const float* in, uint16_t* out, size_t n); }
static void pack(const float* in, uint16_t* out, size_t n) { vaex::host_pack_f16(in, out, n); }
static void pack_avx2_probe(const float* in, uint16_t* out, size_t n) {
  const __m256 lim = _mm256_set1_ps(65504.f), nlim = _mm256_set1_ps(-65504.f);
  for (size_t i = 0; i < n; i += 16) {
    __m256 a = _mm256_min_ps(_mm256_max_ps(_mm256_loadu_ps(in + i), nlim), lim);
    __m256 b = _mm256_min_ps(_mm256_max_ps(_mm256_loadu_ps(in + i + 8), nlim), lim);
    __m128i ha = _mm256_cvtps_ph(a, _MM_FROUND_TO_NEAREST_INT), hb = _mm256_cvtps_ph(b, _MM_FROUND_TO_NEAREST_INT);
    _mm256_stream_si256((__m256i*)(out + i), _mm256_set_m128i(hb, ha));
  }
}
static double now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
int main() {
  const size_t n = (size_t)1 << 30;  // 4 GiB of fp32
  float* in; uint16_t* out; void* dev;
  cudaMallocHost(&in, n * 4); cudaMallocHost(&out, n * 2); cudaMalloc(&dev, n * 4);
  { std::vector<std::thread> th; for (int t = 0; t < 16; ++t) th.emplace_back([=] { for (size_t i = n / 16 * t; i < n / 16 * (t + 1); ++i) in[i] = (float)(i & 1023) * 0.37f; }); for (auto& x : th) x.join(); }
  cudaStream_t st; cudaStreamCreate(&st);
  for (int rep = 0; rep < 2; ++rep) {
    double t0 = now(); cudaMemcpyAsync(dev, in, n * 4, cudaMemcpyHostToDevice, st); cudaStreamSynchronize(st);
    printf("h2d fp32 alone: %.1f GB/s\n", n * 4 / (now() - t0) / 1e9);
  }
  for (int T : {4, 8, 12, 16}) for (int with_copy = 0; with_copy < 2; ++with_copy) {
    double t0 = now();
    if (with_copy) cudaMemcpyAsync(dev, out, n * 2, cudaMemcpyHostToDevice, st);
    std::vector<std::thread> th;
    for (int t = 0; t < T; ++t) th.emplace_back([=] { pack(in + n / T / 16 * 16 * t, out + n / T / 16 * 16 * t, n / T / 16 * 16); });
    for (auto& x : th) x.join();
    double t1 = now();
    cudaStreamSynchronize(st);
    double t2 = now();
    printf("pack T=%2d copy=%d: %.1f GB/s of fp32 in (pack %.3fs, h2d of 2 GiB done after %.3fs)\n", T, with_copy, n * 4 / (t1 - t0) / 1e9, t1 - t0, t2 - t0);
  }
  return 0;
}
