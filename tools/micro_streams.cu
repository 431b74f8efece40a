// micro_streams.cu -- what HBM delivers to MANY independent A->B streams (not product code).
// The streaming batch of the lazy loop (configs[3] shape) is 296 CTAs, each copying its own
// 1.05 MB tableau A -> B with K fma per element.  This measures the same traffic without the
// simplex around it: `nstreams` CTAs x `nt` threads, each CTA passes `passes` times over its own
// region of `kb` KB (ping-pong), UN 16-byte loads in flight per thread, KF fma per element.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -fmad=false -o micro_streams micro_streams.cu
//   ./micro_streams
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

__device__ __forceinline__ double2 ldcg2(const double *p) { return __ldcg(reinterpret_cast<const double2 *>(p)); }
__device__ __forceinline__ void stcg2(double *p, double2 v) { __stcg(reinterpret_cast<double2 *>(p), v); }

// private: CTA b owns region b (A and B halves); shared == 1: the CTAs split ONE big region evenly
template <int UN, int KF>
__global__ void __launch_bounds__(512, 1) streams(double *base, size_t region_doubles, int passes, int interleave) {
  const int NT = blockDim.x;
  double *A, *B;
  size_t n2;  // double2 elements this CTA moves per pass
  if (!interleave) {
    A = base + (size_t)blockIdx.x * 2 * region_doubles;
    B = A + region_doubles;
    n2 = region_doubles / 2;
  } else {
    // same total bytes, but the CTAs' regions are interleaved in 4 KB rows: row r of CTA b sits at
    // (r * gridDim.x + b) * 512 doubles -- at any time the grid touches a compact address range
    A = base;
    B = base + (size_t)gridDim.x * region_doubles;
    n2 = region_doubles / 2;
  }
  double2 rh[KF > 0 ? KF : 1];
#pragma unroll
  for (int p = 0; p < KF; p++) rh[p] = make_double2(1e-9 * (p + 1), 1e-9 * (p + 2));
  for (int it = 0; it < passes; it++) {
    const double *s = (it & 1) ? B : A;
    double *d = (it & 1) ? A : B;
    for (size_t v = threadIdx.x; v < n2; v += (size_t)NT * UN) {
      double2 t[UN];
      size_t off[UN];
#pragma unroll
      for (int u = 0; u < UN; u++) {
        const size_t e = v + (size_t)u * NT;  // double2 index inside the CTA's region
        if (!interleave) off[u] = 2 * e;
        else {
          const size_t row = e / 256, c = e % 256;  // 256 double2 = one 4 KB row
          off[u] = (row * gridDim.x + blockIdx.x) * 512 + 2 * c;
        }
        if (e < n2) t[u] = ldcg2(s + off[u]);
      }
#pragma unroll
      for (int u = 0; u < UN; u++) {
        const size_t e = v + (size_t)u * NT;
        if (e < n2) {
          double2 o = t[u];
#pragma unroll
          for (int p = 0; p < KF; p++) { o.x = fma(1.0 + 1e-12 * p, rh[p].x, o.x); o.y = fma(1.0 + 1e-12 * p, rh[p].y, o.y); }
          stcg2(d + off[u], o);
        }
      }
    }
    __syncthreads();
  }
}

template <int UN, int KF>
static void run(const char *name, int nstreams, int nt, size_t kb, int passes, int interleave, double *buf) {
  const size_t region = kb * 1024 / 8;
  cudaEvent_t a, b;
  CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
  streams<UN, KF><<<nstreams, nt>>>(buf, region, 2, interleave);
  CK(cudaDeviceSynchronize());
  CK(cudaEventRecord(a));
  streams<UN, KF><<<nstreams, nt>>>(buf, region, passes, interleave);
  CK(cudaEventRecord(b));
  CK(cudaDeviceSynchronize());
  float ms;
  CK(cudaEventElapsedTime(&ms, a, b));
  const double bytes = 2.0 * (double)nstreams * region * 8 * passes;
  printf("%-34s streams=%4d nt=%3d region=%5zu KB UN=%d KF=%d %s: %8.2f ms  %7.1f GB/s (read+write)\n", name, nstreams, nt, kb, UN, KF,
         interleave ? "interleaved" : "private    ", ms, bytes / ms / 1e6);
}

int main() {
  const size_t total = (size_t)6 << 30;  // 6 GB
  double *buf;
  CK(cudaMalloc(&buf, total));
  CK(cudaMemset(buf, 0, total));
  // the batch's shape: 296 streams x 256 threads, 1028 KB regions (257 rows x 4 KB); far beyond L2 in total
  run<4, 0>("copy", 296, 256, 1028, 40, 0, buf);
  run<4, 8>("copy + 8 fma", 296, 256, 1028, 40, 0, buf);
  run<8, 0>("copy UN=8", 296, 256, 1028, 40, 0, buf);
  run<8, 8>("copy + 8 fma UN=8", 296, 256, 1028, 40, 0, buf);
  run<4, 0>("copy, 1 CTA/SM x 512", 148, 512, 1028, 40, 0, buf);
  run<4, 8>("copy + 8 fma, 1 CTA/SM x 512", 148, 512, 1028, 40, 0, buf);
  run<4, 0>("copy interleaved rows", 296, 256, 1028, 40, 1, buf);
  run<4, 8>("copy + 8 fma interleaved rows", 296, 256, 1028, 40, 1, buf);
  // more streams than resident CTAs (second wave), and bigger regions
  run<4, 8>("copy + 8 fma, 592 streams", 592, 256, 1028, 20, 0, buf);
  run<4, 8>("copy + 8 fma, 4 MB regions", 296, 256, 4112, 10, 0, buf);
  // L2-resident for comparison (296 x 2 x 128 KB = 74 MB)
  run<4, 8>("copy + 8 fma, L2-resident", 296, 256, 128, 200, 0, buf);
  return 0;
}
