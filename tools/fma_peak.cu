// fma_peak.cu - the FFMA micro-benchmark behind the non-tensor FP32 peak of bench.py's roofline ("SMs x 128
// lanes x 2 x clock ... to be confirmed with an FMA micro-benchmark committed next to the ncu captures").
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o /tmp/fma_peak tools/fma_peak.cu && /tmp/fma_peak > profiles/r2_fma_peak.json
// 8 independent FFMA chains per thread (enough ILP for the 4-cycle dependent-issue latency at 16 warps per scheduler),
// 1024 threads x 2 CTAs per SM, timed with CUDA events after a warm-up; the SM clock is sampled through NVML-free
// cudaDeviceGetAttribute (max) and from the cycle counter (achieved).
#include <cstdio>
#include <cuda_runtime.h>

constexpr int kIters = 4096, kChains = 8;

__global__ void __launch_bounds__(1024, 2) ffma_kernel(float* out, float a, float b, long long* cycles) {
  float x[kChains];
#pragma unroll
  for (int k = 0; k < kChains; ++k) x[k] = a + threadIdx.x * 1e-6f + k;
  const long long c0 = clock64();
#pragma unroll 1
  for (int i = 0; i < kIters; ++i) {
#pragma unroll
    for (int r = 0; r < 8; ++r)
#pragma unroll
      for (int k = 0; k < kChains; ++k) x[k] = fmaf(x[k], a, b);
  }
  const long long c1 = clock64();
  float s = 0.f;
#pragma unroll
  for (int k = 0; k < kChains; ++k) s += x[k];
  if (s == 12345.678f) out[0] = s;  // never true: keeps the chains alive
  if (threadIdx.x == 0 && blockIdx.x == 0) *cycles = c1 - c0;
}

int main() {
  cudaDeviceProp p;
  cudaGetDeviceProperties(&p, 0);
  int clock_khz = 0;
  cudaDeviceGetAttribute(&clock_khz, cudaDevAttrClockRate, 0);
  float* out;
  long long *d_cyc, h_cyc = 0;
  cudaMalloc(&out, 4);
  cudaMalloc(&d_cyc, 8);
  const int grid = p.multiProcessorCount * 2;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int w = 0; w < 5; ++w) ffma_kernel<<<grid, 1024>>>(out, 0.999f, 1e-3f, d_cyc);
  cudaDeviceSynchronize();
  float best = 1e30f;
  for (int rep = 0; rep < 10; ++rep) {
    cudaEventRecord(e0);
    ffma_kernel<<<grid, 1024>>>(out, 0.999f, 1e-3f, d_cyc);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    if (ms < best) best = ms;
  }
  cudaMemcpy(&h_cyc, d_cyc, 8, cudaMemcpyDeviceToHost);
  const double ffma = (double)grid * 1024 * kIters * 8 * kChains;
  const double tflops = 2.0 * ffma / (best * 1e-3) / 1e12;
  // the clock the SMs really ran at: FFMA warp instructions per SM sub-partition (one issue slot each) over the duration
  (void)h_cyc;
  const double sm_mhz = ffma / 32.0 / (p.multiProcessorCount * 4.0) / (best * 1e-3) / 1e6;
  const double nominal = p.multiProcessorCount * 128 * 2.0 * clock_khz * 1e3 / 1e12;
  printf("{\"gpu\": \"%s\", \"sms\": %d, \"ffma_tflops\": %.2f, \"ms\": %.4f, \"sm_mhz\": %.0f, \"max_sm_mhz\": %.0f, "
         "\"nominal_tflops_at_max_clock\": %.2f, \"frac_of_nominal\": %.3f, "
         "\"how\": \"tools/fma_peak.cu: %d CTAs x 1024 threads x %d iterations x 8 x %d independent FFMA chains, best of 10, CUDA events\"}\n",
         p.name, p.multiProcessorCount, tflops, best, sm_mhz, clock_khz / 1e3, nominal, tflops / nominal, grid, kIters, kChains);
  return cudaGetLastError() != cudaSuccess;
}
