// ffma_peak.cu — measured fp32 FFMA peak of the box (the denominator of the FP32_EXACT / exact re-score rooflines;
// BASELINE.md 2 promised a measured number instead of the nominal 2*128*SMs*clock).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/ffma_peak tools/ffma_peak.cu ; run: tools/ffma_peak
// Each thread runs 16 independent fmaf chains (enough ILP to hide the 4-cycle FMA latency at 8+ warps per scheduler).
#include <cuda_runtime.h>
#include <stdio.h>

__global__ void __launch_bounds__(512) ffma_kernel(float *out, int iters, float a, float b) {
  float x[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) x[i] = (float)(threadIdx.x + i) * 1e-3f;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int r = 0; r < 8; ++r) {
#pragma unroll
      for (int i = 0; i < 16; ++i) x[i] = __fmaf_rn(x[i], a, b);
    }
  }
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < 16; ++i) s += x[i];
  if (s == 123.456f) out[0] = s;  // never true: keeps the chains alive
}

int main() {
  int dev = 0, sms = 0, clk = 0;
  cudaSetDevice(dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, dev);
  float *out;
  cudaMalloc(&out, 4);
  const int iters = 20000, blocks = sms * 4, threads = 512;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  float best = 1e30f;
  for (int rep = 0; rep < 6; ++rep) {
    cudaEventRecord(e0);
    ffma_kernel<<<blocks, threads>>>(out, iters, 0.999f, 1e-4f);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    if (rep >= 1 && ms < best) best = ms;
  }
  const double flop = 2.0 * 16 * 8 * (double)iters * blocks * threads;
  const double tf = flop / (best * 1e-3) / 1e12;
  printf("{\"ffma_tflops\": %.2f, \"ms\": %.3f, \"sms\": %d, \"clock_rate_khz\": %d, \"nominal_tflops_at_clock_rate\": %.2f, "
         "\"blocks\": %d, \"threads\": %d, \"chains_per_thread\": 16, \"err\": \"%s\"}\n",
         tf, best, sms, clk, 2.0 * 128 * sms * (clk * 1e3) / 1e12, blocks, threads, cudaGetErrorString(cudaGetLastError()));
  return 0;
}
