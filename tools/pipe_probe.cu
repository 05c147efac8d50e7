// Pipeline probe for vmasb200_step_host: per chunk H2D (24 + 12 rows) -> kernel (~10 us) -> D2H (24 + 48 rows), chunks on
// round-robin streams, one host sync at the end; navigation at 32768 envs (ld = 32768).
#include <cstdio>
#include <chrono>
#include <cuda_runtime.h>
__global__ void spin(long long cycles) { long long t0 = clock64(); while (clock64() - t0 < cycles) {} }
int main() {
  const size_t ld = 32768;
  float *h_in, *h_out, *d_in, *d_out;
  cudaMallocHost(&h_in, 36 * ld * 4); cudaMallocHost(&h_out, 72 * ld * 4);
  cudaMalloc(&d_in, 36 * ld * 4); cudaMalloc(&d_out, 72 * ld * 4);
  cudaStream_t st[8]; for (auto& s : st) cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking);
  for (int nchunk : {1, 2, 3, 4, 8}) {
    for (int split = 0; split < 2; ++split) {
      const size_t cs = ld / nchunk;
      auto step = [&] {
        for (int k = 0; k < nchunk; ++k) {
          cudaStream_t s = st[k % 8];
          const size_t c0 = k * cs;
          if (split) {  // 3 H2D + 2 D2H copies per chunk, as the library issues them
            cudaMemcpy2DAsync(d_in + c0, ld * 4, h_in + c0, ld * 4, cs * 4, 24, cudaMemcpyHostToDevice, s);
            cudaMemcpy2DAsync(d_in + 24 * ld + c0, ld * 4, h_in + 24 * ld + c0, ld * 4, cs * 4, 8, cudaMemcpyHostToDevice, s);
            cudaMemcpy2DAsync(d_in + 32 * ld + c0, ld * 4, h_in + 32 * ld + c0, ld * 4, cs * 4, 4, cudaMemcpyHostToDevice, s);
          } else {
            cudaMemcpy2DAsync(d_in + c0, ld * 4, h_in + c0, ld * 4, cs * 4, 36, cudaMemcpyHostToDevice, s);
          }
          spin<<<148, 128, 0, s>>>(16000);
          if (split) {
            cudaMemcpy2DAsync(h_out + c0, ld * 4, d_out + c0, ld * 4, cs * 4, 24, cudaMemcpyDeviceToHost, s);
            cudaMemcpy2DAsync(h_out + 24 * ld + c0, ld * 4, d_out + 24 * ld + c0, ld * 4, cs * 4, 48, cudaMemcpyDeviceToHost, s);
          } else {
            cudaMemcpy2DAsync(h_out + c0, ld * 4, d_out + c0, ld * 4, cs * 4, 72, cudaMemcpyDeviceToHost, s);
          }
        }
        for (int k = 0; k < nchunk && k < 8; ++k) cudaStreamSynchronize(st[k]);
      };
      for (int i = 0; i < 5; ++i) step();
      auto t0 = std::chrono::steady_clock::now();
      for (int i = 0; i < 50; ++i) step();
      double us = std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - t0).count() / 50;
      printf("chunks %d  %s copies: %7.1f us per step = %6.1f M env-steps/s\n", nchunk, split ? "split " : "merged", us, 32768 / us);
    }
  }
  return 0;
}
