// PCIe probe for vmasb200_step_host's copies: flat vs pitched (2-D) copies of SoA rows, pinned host memory.
#include <cstdio>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e) { printf("%s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)
int main() {
  const size_t ld = 32768, rows = 72;
  float *h, *d;
  CK(cudaMallocHost(&h, rows * ld * 4)); CK(cudaMalloc(&d, rows * ld * 4));
  cudaStream_t s1, s2; CK(cudaStreamCreate(&s1)); CK(cudaStreamCreate(&s2));
  cudaEvent_t a, b; CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
  auto time = [&](const char* name, auto fn) {
    for (int i = 0; i < 3; ++i) fn();
    cudaDeviceSynchronize();
    cudaEventRecord(a, s1);
    for (int i = 0; i < 20; ++i) fn();
    cudaEventRecord(b, s1); cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b);
    printf("%-46s %8.1f us  %6.1f GB/s\n", name, ms * 1000 / 20, rows * ld * 4 / (ms / 20 * 1e-3) / 1e9);
    return 0;
  };
  time("D2H flat 9.4 MB", [&] { cudaMemcpyAsync(h, d, rows * ld * 4, cudaMemcpyDeviceToHost, s1); });
  time("H2D flat 9.4 MB", [&] { cudaMemcpyAsync(d, h, rows * ld * 4, cudaMemcpyHostToDevice, s1); });
  time("D2H 2D 72 rows x 128 KB (full width)", [&] { cudaMemcpy2DAsync(h, ld * 4, d, ld * 4, ld * 4, rows, cudaMemcpyDeviceToHost, s1); });
  time("D2H 2D 2 x (72 rows x 64 KB)", [&] { for (int c = 0; c < 2; ++c) cudaMemcpy2DAsync(h + c * ld / 2, ld * 4, d + c * ld / 2, ld * 4, ld * 2, rows, cudaMemcpyDeviceToHost, s1); });
  time("D2H 2D 4 x (72 rows x 32 KB)", [&] { for (int c = 0; c < 4; ++c) cudaMemcpy2DAsync(h + c * ld / 4, ld * 4, d + c * ld / 4, ld * 4, ld, rows, cudaMemcpyDeviceToHost, s1); });
  time("H2D 2D 2 x (72 rows x 64 KB)", [&] { for (int c = 0; c < 2; ++c) cudaMemcpy2DAsync(d + c * ld / 2, ld * 4, h + c * ld / 2, ld * 4, ld * 2, rows, cudaMemcpyHostToDevice, s1); });
  time("D2H 72 flat copies of 128 KB", [&] { for (size_t r = 0; r < rows; ++r) cudaMemcpyAsync(h + r * ld, d + r * ld, ld * 4, cudaMemcpyDeviceToHost, s1); });
  // small-copy latency
  time("D2H flat 64 B x 1 (reported GB/s meaningless)", [&] { cudaMemcpyAsync(h, d, 64, cudaMemcpyDeviceToHost, s1); });
  return 0;
}
