// micro_update.cu -- L2-resident rank-1 update bandwidth experiments (not product code).
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -fmad=false -o micro_update micro_update.cu
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <vector>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

__device__ __forceinline__ double2 ldcg2(const double *p) { return __ldcg(reinterpret_cast<const double2 *>(p)); }
__device__ __forceinline__ void stcg2(double *p, double2 v) { __stcg(reinterpret_cast<double2 *>(p), v); }

// Variant A: column-chunk x row-range units (the shape engine_ps.cuh uses), UNROLL rows in flight
template <int UNROLL>
__global__ void __launch_bounds__(512, 1) varA(double *T0, double *T1, int R, int ld, int iters, const double *col, const double *rho) {
  const int g = blockIdx.x, G = gridDim.x;
  const int tcols = 256, trows = 2, tx = threadIdx.x & 255, ty = threadIdx.x >> 8;
  const int CW = 512, nch = ld / CW;
  for (int it = 0; it < iters; it++) {
    const double *Ts = (it & 1) ? T1 : T0;
    double *Td = (it & 1) ? T0 : T1;
    const long long nunits = (long long)nch * R;
    long long u0 = nunits * g / G;
    const long long u1 = nunits * (g + 1) / G;
    while (u0 < u1) {
      const int ch = (int)(u0 / R), ia = (int)(u0 - (long long)ch * R);
      const long long segend = u1 < (long long)(ch + 1) * R ? u1 : (long long)(ch + 1) * R;
      const int ib = ia + (int)(segend - u0);
      const int c = ch * CW + 2 * tx;
      const double2 rh = *reinterpret_cast<const double2 *>(rho + c);
      for (int i = ia + ty; i < ib; i += trows * UNROLL) {
        double2 t[UNROLL];
#pragma unroll
        for (int u = 0; u < UNROLL; u++) { const int ii = i + u * trows; if (ii < ib) t[u] = ldcg2(Ts + (size_t)ii * ld + c); }
#pragma unroll
        for (int u = 0; u < UNROLL; u++) {
          const int ii = i + u * trows;
          if (ii < ib) { const double cc = col[ii]; double2 o; o.x = fma(cc, rh.x, t[u].x); o.y = fma(cc, rh.y, t[u].y); stcg2(Td + (size_t)ii * ld + c, o); }
        }
      }
      u0 = segend;
    }
    __syncthreads();
  }
}

// Variant B: flat split of the whole (R*ld) array in 16-byte elements; every thread
// issues ALL of its loads (NL of them) before the first use.
template <int NL, int NT>
__global__ void __launch_bounds__(NT, 1) varB(double *T0, double *T1, int R, int ld, int iters, const double *col, const double *rho) {
  const int g = blockIdx.x, G = gridDim.x;
  const long long nvec = (long long)R * ld / 2;
  const long long v0 = nvec * g / G, v1 = nvec * (g + 1) / G;
  for (int it = 0; it < iters; it++) {
    const double *Ts = (it & 1) ? T1 : T0;
    double *Td = (it & 1) ? T0 : T1;
    for (long long base = v0; base < v1; base += (long long)NL * NT) {
      double2 t[NL];
#pragma unroll
      for (int u = 0; u < NL; u++) { const long long v = base + u * NT + threadIdx.x; if (v < v1) t[u] = ldcg2(Ts + 2 * v); }
#pragma unroll
      for (int u = 0; u < NL; u++) {
        const long long v = base + u * NT + threadIdx.x;
        if (v < v1) {
          const int i = (int)(2 * v / ld), c = (int)(2 * v - (long long)i * ld);
          const double cc = col[i];
          const double2 rh = *reinterpret_cast<const double2 *>(rho + c);
          double2 o; o.x = fma(cc, rh.x, t[u].x); o.y = fma(cc, rh.y, t[u].y);
          stcg2(Td + 2 * v, o);
        }
      }
    }
    __syncthreads();
  }
}

// Variant R: read only (sum), Variant W: write only -- to see the two directions apart
template <int NL, int NT>
__global__ void __launch_bounds__(NT, 1) varR(double *T0, double *T1, int R, int ld, int iters, double *sink) {
  const int g = blockIdx.x, G = gridDim.x;
  const long long nvec = (long long)R * ld / 2;
  const long long v0 = nvec * g / G, v1 = nvec * (g + 1) / G;
  double acc = 0;
  for (int it = 0; it < iters; it++) {
    const double *Ts = (it & 1) ? T1 : T0;
    for (long long base = v0; base < v1; base += (long long)NL * NT) {
      double2 t[NL];
#pragma unroll
      for (int u = 0; u < NL; u++) { const long long v = base + u * NT + threadIdx.x; t[u] = v < v1 ? ldcg2(Ts + 2 * v) : make_double2(0, 0); }
#pragma unroll
      for (int u = 0; u < NL; u++) acc += t[u].x + t[u].y;
    }
  }
  if (acc == 12345.678) sink[0] = acc;
}
template <int NL, int NT>
__global__ void __launch_bounds__(NT, 1) varW(double *T0, double *T1, int R, int ld, int iters) {
  const int g = blockIdx.x, G = gridDim.x;
  const long long nvec = (long long)R * ld / 2;
  const long long v0 = nvec * g / G, v1 = nvec * (g + 1) / G;
  for (int it = 0; it < iters; it++) {
    double *Td = (it & 1) ? T0 : T1;
    for (long long base = v0; base < v1; base += (long long)NL * NT) {
#pragma unroll
      for (int u = 0; u < NL; u++) { const long long v = base + u * NT + threadIdx.x; if (v < v1) stcg2(Td + 2 * v, make_double2(it, u)); }
    }
  }
}

int main(int argc, char **argv) {
  const int R = 1025, ld = 2048, iters = 2000;
  double *T0, *T1, *col, *rho, *sink;
  const size_t bytes = (size_t)R * ld * 8;
  CK(cudaMalloc(&T0, bytes)); CK(cudaMalloc(&T1, bytes)); CK(cudaMalloc(&col, R * 8)); CK(cudaMalloc(&rho, ld * 8)); CK(cudaMalloc(&sink, 8));
  CK(cudaMemset(T0, 0, bytes)); CK(cudaMemset(T1, 0, bytes)); CK(cudaMemset(col, 0, R * 8)); CK(cudaMemset(rho, 0, ld * 8));
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  auto report = [&](const char *name, float ms, double bytes_per_iter) {
    printf("%-28s %7.3f us/update  %7.1f GB/s\n", name, 1e3 * ms / iters, bytes_per_iter * iters / (ms * 1e-3) / 1e9);
  };
#define RUN(name, bpi, ...) do { __VA_ARGS__; CK(cudaDeviceSynchronize()); cudaEventRecord(e0); __VA_ARGS__; cudaEventRecord(e1); CK(cudaDeviceSynchronize()); float ms; cudaEventElapsedTime(&ms, e0, e1); report(name, ms, bpi); } while (0)
  for (int G : {148, 296}) {
    printf("G=%d\n", G);
    RUN("A unroll8", 2.0 * bytes, (varA<8><<<G, 512>>>(T0, T1, R, ld, iters, col, rho)));
    RUN("A unroll14", 2.0 * bytes, (varA<14><<<G, 512>>>(T0, T1, R, ld, iters, col, rho)));
    RUN("B NL=7 NT=1024", 2.0 * bytes, (varB<7, 1024><<<G, 1024>>>(T0, T1, R, ld, iters, col, rho)));
    RUN("B NL=14 NT=512", 2.0 * bytes, (varB<14, 512><<<G, 512>>>(T0, T1, R, ld, iters, col, rho)));
    RUN("B NL=8 NT=512", 2.0 * bytes, (varB<8, 512><<<G, 512>>>(T0, T1, R, ld, iters, col, rho)));
    RUN("B NL=4 NT=512", 2.0 * bytes, (varB<4, 512><<<G, 512>>>(T0, T1, R, ld, iters, col, rho)));
    RUN("B NL=4 NT=1024", 2.0 * bytes, (varB<4, 1024><<<G, 1024>>>(T0, T1, R, ld, iters, col, rho)));
    RUN("R NL=14 NT=512", 1.0 * bytes, (varR<14, 512><<<G, 512>>>(T0, T1, R, ld, iters, sink)));
    RUN("R NL=7 NT=1024", 1.0 * bytes, (varR<7, 1024><<<G, 1024>>>(T0, T1, R, ld, iters, sink)));
    RUN("W NL=14 NT=512", 1.0 * bytes, (varW<14, 512><<<G, 512>>>(T0, T1, R, ld, iters)));
  }
  return 0;
}
