// fma_peak.cu -- measured FP32 FMA throughput of the GPU (the denominator of bench.py's
// secondary roofline, SURVEY.md 8(d): "FP32 peak = 148 * 128 * 2 * f_clk; measure").
// Every thread runs 16 independent FFMA chains from registers; 8 CTAs of 256 threads per SM
// keep all four schedulers of an SM fed.  C ABI: fma_peak_run(iters, reps, &best_ms)
// returns the best-of-reps rate in TFLOP/s (2 flops per FMA), timed with CUDA events.
// Build: tools/fma_peak.py (nvcc -gencode arch=compute_100a,code=sm_100a).
#include <cuda_runtime.h>

__global__ void __launch_bounds__(256) fma_chain_kernel(float* out, int iters, float a, float b) {
  float v[16];
#pragma unroll
  for (int i = 0; i < 16; i++) v[i] = (float)(threadIdx.x + i) * 1e-3f;
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int i = 0; i < 16; i++) v[i] = __fmaf_rn(v[i], a, b);
  }
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < 16; i++) s += v[i];
  if (s == 123.456f) out[0] = s;  // never true: keeps the chains alive
}

extern "C" double fma_peak_run(int iters, int reps, float* best_ms) {
  int dev = 0, sms = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  float* out = nullptr;
  if (cudaMalloc(&out, 4) != cudaSuccess) return -1.0;
  const int blocks = sms * 8;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  fma_chain_kernel<<<blocks, 256>>>(out, iters, 0.999f, 1e-3f);  // warm-up
  float best = 1e30f;
  for (int r = 0; r < reps; r++) {
    cudaEventRecord(e0);
    fma_chain_kernel<<<blocks, 256>>>(out, iters, 0.999f, 1e-3f);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e0, e1);
    if (ms < best) best = ms;
  }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  cudaFree(out);
  if (cudaGetLastError() != cudaSuccess) return -1.0;
  if (best_ms) *best_ms = best;
  const double flops = 2.0 * 16.0 * (double)iters * 256.0 * (double)blocks;
  return flops / ((double)best * 1e-3) / 1e12;
}
