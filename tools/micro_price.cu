// micro_price.cu -- why is the smem pricing loop slow?  (experiment, not product code)
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#define TOL 1e-9
struct Key { unsigned long long k; int var; int idx; };
__device__ __forceinline__ bool key_less(const Key &a, const Key &b) { return a.k < b.k || (a.k == b.k && a.var < b.var); }
__device__ __forceinline__ bool elig(int f, double dj) {
  if (f == 0) return dj < -TOL;
  if (f == 1) return dj > TOL;
  if (f == 2) return dj < -TOL || dj > TOL;
  return false;
}
extern __shared__ __align__(16) unsigned char sm[];
template <int V>
__global__ void __launch_bounds__(512, 1) k(int n, int reps, long long *out, int *sink, int nblk_code) {
  double *d = (double *)sm;
  int *at = (int *)(sm + n * 8), *nb = at + n;
  for (int j = threadIdx.x; j < n; j += blockDim.x) { d[j] = -1.0 / (1 + (j * 7919) % 1013); at[j] = (j % 5 == 0) ? 1 : 0; nb[j] = (j * 31) % n; }
  __syncthreads();
  long long t0 = clock64();
  int acc = 0;
  for (int r = 0; r < reps; r++) {
    Key bq{~0ULL, 0x7fffffff, -1};
    if (V == 0) {
      for (int j = threadIdx.x; j < n; j += blockDim.x) {
        const double dj = d[j];
        if (!elig(at[j], dj)) continue;
        const Key kk{~(unsigned long long)__double_as_longlong(fabs(dj)), nb[j], j};
        if (key_less(kk, bq)) bq = kk;
      }
    } else if (V == 1) {  // branch-free
      for (int j = threadIdx.x; j < n; j += blockDim.x) {
        const double dj = d[j];
        const int f = at[j];
        const bool e = (f == 0 && dj < -TOL) || (f == 1 && dj > TOL) || (f == 2 && fabs(dj) > TOL);
        const unsigned long long kb = e ? ~(unsigned long long)__double_as_longlong(fabs(dj)) : ~0ULL;
        const Key kk{kb, nb[j], j};
        if (key_less(kk, bq)) bq = kk;
      }
    } else {  // 4 consecutive per thread, vector loads
      for (int j0 = 4 * threadIdx.x; j0 < n; j0 += 4 * blockDim.x) {
        const double2 a = *(const double2 *)(d + j0), b = *(const double2 *)(d + j0 + 2);
        const int4 f = *(const int4 *)(at + j0);
        const int4 v = *(const int4 *)(nb + j0);
        const double dd[4] = {a.x, a.y, b.x, b.y};
        const int ff[4] = {f.x, f.y, f.z, f.w};
        const int vv[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int e4 = 0; e4 < 4; e4++) {
          const double dj = dd[e4];
          const bool e = (ff[e4] == 0 && dj < -TOL) || (ff[e4] == 1 && dj > TOL) || (ff[e4] == 2 && fabs(dj) > TOL);
          const unsigned long long kb = e ? ~(unsigned long long)__double_as_longlong(fabs(dj)) : ~0ULL;
          const Key kk{kb, vv[e4], j0 + e4};
          if (key_less(kk, bq)) bq = kk;
        }
      }
    }
    acc += bq.idx + (int)(bq.k >> 40) + bq.var;
    __syncthreads();
    if (threadIdx.x == r % 512) d[(r * 13) % n] *= 1.0000001;  // keep the loop honest
    __syncthreads();
  }
  long long t1 = clock64();
  if (threadIdx.x == 0) out[blockIdx.x] = t1 - t0;
  sink[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}
int main() {
  const int n = 2048, reps = 2000;
  long long *out; int *sink;
  cudaMalloc(&out, 148 * 8); cudaMalloc(&sink, 148 * 512 * 4);
  long long h[148];
  for (int v = 0; v < 3; v++) {
    for (int blocks : {1, 148}) {
      if (v == 0) k<0><<<blocks, 512, n * 16>>>(n, reps, out, sink, 0);
      if (v == 1) k<1><<<blocks, 512, n * 16>>>(n, reps, out, sink, 0);
      if (v == 2) k<2><<<blocks, 512, n * 16>>>(n, reps, out, sink, 0);
      cudaError_t e = cudaDeviceSynchronize();
      cudaMemcpy(h, out, blocks * 8, cudaMemcpyDeviceToHost);
      printf("variant %d blocks %3d: %8.1f clk per pricing pass (incl. 2 syncs) %s\n", v, blocks, (double)h[0] / reps, cudaGetErrorString(e));
    }
  }
  return 0;
}
