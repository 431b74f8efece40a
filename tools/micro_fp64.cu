// micro_fp64.cu -- DFMA / DSETP issue rate per SM (experiment)
#include <cuda_runtime.h>
#include <cstdio>
__global__ void __launch_bounds__(512, 1) kf(double *out, long long *clk, int reps) {
  double a0 = threadIdx.x, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5, a6 = a0 + 6, a7 = a0 + 7;
  const double b = 1.0000001, c = 0.5;
  __syncthreads();
  long long t0 = clock64();
  for (int r = 0; r < reps; r++) {
    a0 = fma(a0, b, c); a1 = fma(a1, b, c); a2 = fma(a2, b, c); a3 = fma(a3, b, c);
    a4 = fma(a4, b, c); a5 = fma(a5, b, c); a6 = fma(a6, b, c); a7 = fma(a7, b, c);
  }
  __syncthreads();
  long long t1 = clock64();
  if (threadIdx.x == 0) clk[blockIdx.x] = t1 - t0;
  out[blockIdx.x * blockDim.x + threadIdx.x] = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7;
}
__global__ void __launch_bounds__(512, 1) kc(int *out, long long *clk, int reps, double x) {
  int c0 = 0, c1 = 0, c2 = 0, c3 = 0;
  double a = threadIdx.x * 0.001;
  __syncthreads();
  long long t0 = clock64();
  for (int r = 0; r < reps; r++) {
    c0 += (a < x + r); c1 += (a > x - r); c2 += (a + 1.0 < x * r) ; c3 += (a - 1.0 > x + 2.0 * r);
  }
  __syncthreads();
  long long t1 = clock64();
  if (threadIdx.x == 0) clk[blockIdx.x] = t1 - t0;
  out[blockIdx.x * blockDim.x + threadIdx.x] = c0 + c1 + c2 + c3;
}
__global__ void __launch_bounds__(512, 1) ki(int *out, long long *clk, int reps) {
  int a0 = threadIdx.x, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5, a6 = a0 + 6, a7 = a0 + 7;
  __syncthreads();
  long long t0 = clock64();
  for (int r = 0; r < reps; r++) {
    a0 = a0 * 3 + r; a1 = a1 * 3 + r; a2 = a2 * 3 + r; a3 = a3 * 3 + r; a4 = a4 * 3 + r; a5 = a5 * 3 + r; a6 = a6 * 3 + r; a7 = a7 * 3 + r;
  }
  __syncthreads();
  long long t1 = clock64();
  if (threadIdx.x == 0) clk[blockIdx.x] = t1 - t0;
  out[blockIdx.x * blockDim.x + threadIdx.x] = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7;
}
int main() {
  double *o; int *oi; long long *clk, h;
  cudaMalloc(&o, 148 * 512 * 8); cudaMalloc(&oi, 148 * 512 * 4); cudaMalloc(&clk, 148 * 8);
  const int reps = 4096;
  kf<<<148, 512>>>(o, clk, reps); cudaDeviceSynchronize(); cudaMemcpy(&h, clk, 8, cudaMemcpyDeviceToHost);
  printf("DFMA: %lld clk for %d x 8 per thread, 512 threads -> %.2f DFMA/clk/SM\n", h, reps, 512.0 * 8 * reps / h);
  kc<<<148, 512>>>(oi, clk, reps, 0.3); cudaDeviceSynchronize(); cudaMemcpy(&h, clk, 8, cudaMemcpyDeviceToHost);
  printf("fp64 compare loop: %lld clk for %d iterations (4 compares + few DADD/DMUL each)\n", h, reps);
  ki<<<148, 512>>>(oi, clk, reps); cudaDeviceSynchronize(); cudaMemcpy(&h, clk, 8, cudaMemcpyDeviceToHost);
  printf("IMAD: %lld clk -> %.2f IMAD/clk/SM\n", h, 512.0 * 8 * reps / h);
  return 0;
}
