// Microbenchmarks that size the AR-chain kernel: FP64 FMA throughput / latency per SM, shuffle, block barrier.
#include <cstdio>
#include <cuda_runtime.h>
__global__ void k_dfma_tp(double* out, int iters, long long* cyc) {
  double a0 = threadIdx.x, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5, a6 = a0 + 6, a7 = a0 + 7;
  const double b = 1.0000001, c = 1e-9;
  __syncthreads();
  long long t0 = clock64();
  for (int i = 0; i < iters; i++) {
    a0 = fma(a0, b, c); a1 = fma(a1, b, c); a2 = fma(a2, b, c); a3 = fma(a3, b, c);
    a4 = fma(a4, b, c); a5 = fma(a5, b, c); a6 = fma(a6, b, c); a7 = fma(a7, b, c);
  }
  __syncthreads();
  long long t1 = clock64();
  out[blockIdx.x * blockDim.x + threadIdx.x] = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}
__global__ void k_dfma_lat(double* out, int iters, long long* cyc) {
  double a0 = threadIdx.x;
  const double b = 1.0000001, c = 1e-9;
  long long t0 = clock64();
  for (int i = 0; i < iters; i++) { a0 = fma(a0, b, c); a0 = fma(a0, b, c); a0 = fma(a0, b, c); a0 = fma(a0, b, c); }
  long long t1 = clock64();
  out[threadIdx.x] = a0;
  if (threadIdx.x == 0) cyc[0] = t1 - t0;
}
__global__ void k_ffma_lat(float* out, int iters, long long* cyc) {
  float a0 = threadIdx.x;
  const float b = 1.0000001f, c = 1e-9f;
  long long t0 = clock64();
  for (int i = 0; i < iters; i++) { a0 = fmaf(a0, b, c); a0 = fmaf(a0, b, c); a0 = fmaf(a0, b, c); a0 = fmaf(a0, b, c); }
  long long t1 = clock64();
  out[threadIdx.x] = a0;
  if (threadIdx.x == 0) cyc[0] = t1 - t0;
}
__global__ void k_bar(double* out, int iters, long long* cyc) {
  long long t0 = clock64();
  for (int i = 0; i < iters; i++) __syncthreads();
  long long t1 = clock64();
  if (threadIdx.x == 0) cyc[0] = t1 - t0;
  out[threadIdx.x] = 0;
}
__global__ void k_shfl(double* out, int iters, long long* cyc) {
  double a = threadIdx.x;
  long long t0 = clock64();
  for (int i = 0; i < iters; i++) { a += __shfl_xor_sync(0xffffffffu, a, 1); a += __shfl_xor_sync(0xffffffffu, a, 2); }
  long long t1 = clock64();
  out[threadIdx.x] = a;
  if (threadIdx.x == 0) cyc[0] = t1 - t0;
}
int main() {
  double* out; long long* cyc; float* fo;
  cudaMalloc(&out, 1 << 20); cudaMalloc(&cyc, 1024 * 8); cudaMalloc(&fo, 1 << 20);
  long long h[8];
  const int it = 4096;
  for (int threads : {32, 128, 256, 512, 1024}) {
    k_dfma_tp<<<1, threads>>>(out, it, cyc); cudaMemcpy(h, cyc, 8, cudaMemcpyDeviceToHost);
    printf("DFMA throughput: %4d threads/SM: %.2f cycles per warp-DFMA -> %.1f lanes/clk/SM\n", threads,
           (double)h[0] / (it * 8.0 * threads / 32), 32.0 * (it * 8.0 * threads / 32) / h[0]);
  }
  k_dfma_lat<<<1, 32>>>(out, it, cyc); cudaMemcpy(h, cyc, 8, cudaMemcpyDeviceToHost);
  printf("DFMA dependent latency: %.1f cycles\n", (double)h[0] / (it * 4.0));
  k_ffma_lat<<<1, 32>>>(fo, it, cyc); cudaMemcpy(h, cyc, 8, cudaMemcpyDeviceToHost);
  printf("FFMA dependent latency: %.1f cycles\n", (double)h[0] / (it * 4.0));
  for (int threads : {64, 512}) {
    k_bar<<<1, threads>>>(out, it, cyc); cudaMemcpy(h, cyc, 8, cudaMemcpyDeviceToHost);
    printf("__syncthreads (%d threads): %.1f cycles\n", threads, (double)h[0] / it);
  }
  k_shfl<<<1, 32>>>(out, it, cyc); cudaMemcpy(h, cyc, 8, cudaMemcpyDeviceToHost);
  printf("64-bit shfl_xor + DADD dependent: %.1f cycles per stage\n", (double)h[0] / (it * 2.0));
  printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
  return 0;
}
