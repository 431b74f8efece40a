// engine_cuda.cu -- the LP-relaxation engine on B200 (sm_100a): dense fp64
// bounded-variable tableau simplex, device resident, behind the C ABI of
// include/mvolps_engine.h.  Replaces what the reference reaches through GLPK:
// glp_simplex (bs.cpp:117,279,287), glp_eval_tab_row (gmi.cpp:36) and the
// tableau part of glp_copy_prob (util.cpp:34, bs.cpp:116).
//
// One kernel (`simplex_kernel`) runs whole solves: an LP is owned by a GROUP
// of G co-resident CTAs (G = 1 for batched node LPs, G = all SMs for one large
// tableau).  Per iteration every CTA of the group takes the pivot decisions
// redundantly from L2-resident vectors (pricing / ratio test as warp-shuffle +
// block reductions on a total (value, variable-id) order), then updates its
// share of the tableau from buffer A into buffer B (ping-pong, so that no CTA
// overwrites what another still reads) and the group meets at ONE barrier.
// The arithmetic is the specification of oracle/simplex_oracle.c, operation
// for operation (explicit fma, IEEE divide, -fmad=false), so pivot sequences
// and tableaus are bit-identical to the CPU oracle.
//
// There is no CPU fallback in this file: without a CUDA device every entry
// point fails with an error.
#include <cuda_runtime.h>

#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/mvolps_engine.h"
#include "../../include/mvolps_spec.h"

// ------------------------------------------------------------------ device
struct DevResult {
  mvb_result r;
  int curT, curS;  // which ping-pong buffers hold the final state
};

struct LPDev {
  int m, n, ld, mcap;
  double *T[2];     // (mcap+1) x ld; row 0 = reduced costs
  double *beta[2];  // mcap+1, 1-based
  int *head[2];     // mcap+1, 1-based
  int *nb[2];       // n
  int *at[2];       // n
  double *lb, *ub;  // n+mcap by variable id
  double *cost;     // n
  double *p1part;   // phase-1 partial sums: ceil(mcap/32) x ld
  // outbox: one contiguous block copied back after the solve
  DevResult *res;
  double *xval;     // n+mcap
  int *ohead;       // mcap
  int *onb, *oat;   // n
  int *pivlog;      // 2*pivlog_cap or null
  int pivlog_cap;
  int curT, curS;   // parities at entry
  int max_iters;
};

namespace {

constexpr int STAGE_ROWS = 128;  // pivot-column values staged in smem per pass
constexpr int UNROLL = 8;        // independent 128-bit loads in flight per thread

struct Key {
  unsigned long long k;
  int var;
  int idx;
};

__device__ __forceinline__ Key key_none() { return Key{~0ULL, 0x7fffffff, -1}; }
__device__ __forceinline__ bool key_less(const Key &a, const Key &b) {
  return a.k < b.k || (a.k == b.k && a.var < b.var);
}
__device__ __forceinline__ unsigned long long dbits(double x) {
  return (unsigned long long)__double_as_longlong(x);
}

__device__ __forceinline__ double ldcg(const double *p) { return __ldcg(p); }
__device__ __forceinline__ int ldcg(const int *p) { return __ldcg(p); }
__device__ __forceinline__ double2 ldcg2(const double *p) {
  return __ldcg(reinterpret_cast<const double2 *>(p));
}
__device__ __forceinline__ void stcg2(double *p, double2 v) {
  __stcg(reinterpret_cast<double2 *>(p), v);
}

// warp-shuffle + block reduction to the minimum key; result on every thread
__device__ Key block_argmin(Key v, Key *red) {
#pragma unroll
  for (int off = 16; off >= 1; off >>= 1) {
    Key o;
    o.k = __shfl_xor_sync(0xffffffffu, v.k, off);
    o.var = __shfl_xor_sync(0xffffffffu, v.var, off);
    o.idx = __shfl_xor_sync(0xffffffffu, v.idx, off);
    if (key_less(o, v)) v = o;
  }
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  __syncthreads();  // red[] free again
  if (lane == 0) red[w] = v;
  __syncthreads();
  v = lane < nw ? red[lane] : key_none();
#pragma unroll
  for (int off = 16; off >= 1; off >>= 1) {
    Key o;
    o.k = __shfl_xor_sync(0xffffffffu, v.k, off);
    o.var = __shfl_xor_sync(0xffffffffu, v.var, off);
    o.idx = __shfl_xor_sync(0xffffffffu, v.idx, off);
    if (key_less(o, v)) v = o;
  }
  return v;
}

__device__ int block_sum(int v, int *red) {
#pragma unroll
  for (int off = 16; off >= 1; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  __syncthreads();
  if (lane == 0) red[w] = v;
  __syncthreads();
  v = lane < nw ? red[lane] : 0;
#pragma unroll
  for (int off = 16; off >= 1; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
  return v;
}

// barrier of the G CTAs that own one LP.  *ctr counts arrivals monotonically
// (zeroed by the host before the launch); generation `gen` is tracked in
// registers.  Data written before the barrier is read after it with ld.cg.
__device__ __forceinline__ void group_sync(unsigned *ctr, int G, unsigned &gen) {
  __syncthreads();
  gen++;
  if (G > 1) {
    if (threadIdx.x == 0) {
      __threadfence();
      atomicAdd(ctr, 1u);
      const unsigned target = gen * (unsigned)G;
      unsigned seen;
      do {
        asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(seen) : "l"(ctr) : "memory");
      } while (seen < target);
    }
    __syncthreads();
  }
}

__device__ __forceinline__ int row_infeas(double b, double l, double u) {
  if (l - b > MVB_TOL_BND * (1.0 + fabs(l))) return -1;
  if (b - u > MVB_TOL_BND * (1.0 + fabs(u))) return 1;
  return 0;
}

__device__ __forceinline__ bool dj_eligible(int f, double dj) {
  if (f == MVB_AT_LB) return dj < -MVB_TOL_DJ;
  if (f == MVB_AT_UB) return dj > MVB_TOL_DJ;
  if (f == MVB_AT_FR) return dj < -MVB_TOL_DJ || dj > MVB_TOL_DJ;
  return false;
}

__device__ __forceinline__ double nb_value(int f, double l, double u) {
  if (f == MVB_AT_LB || f == MVB_AT_FX) return l;
  if (f == MVB_AT_UB) return u;
  return 0.0;
}

}  // namespace

namespace {

struct Smem {
  Key redK[32];
  int redI[32];
  double colS[STAGE_ROWS];
};

// ---- K4: rank-1 tableau update, buffer Ts -> Td -------------------------
// Work unit = (row, column chunk of 2*tcols doubles); the group's units are
// split evenly over its G CTAs.  Each thread owns two adjacent columns
// (128-bit ld.cg/st.cg), keeps its rho pair in registers and streams UNROLL
// rows at a time; the pivot-column values of the rows in flight are staged
// in shared memory.   rho_q = 1/p, rho_j = -(T_rj * (1/p));
//   T'_ij = fma(T_iq, rho_j, j == q ? 0 : T_ij),  T'_rj = rho_j.
__device__ void tile_update(const LPDev &L, const double *__restrict__ Ts,
                            double *__restrict__ Td, int r, int q, double inv, int g, int G,
                            int tcl, Smem &sm) {
  const int R = L.m + 1, ld = L.ld;
  const int tcols = 1 << tcl, trows = blockDim.x >> tcl;
  const int tx = threadIdx.x & (tcols - 1), ty = threadIdx.x >> tcl;
  const int CW = 2 * tcols, nch = (ld + CW - 1) / CW;
  const long long nunits = (long long)nch * R;
  long long u0 = nunits * g / G;
  const long long u1 = nunits * (g + 1) / G;
  while (u0 < u1) {
    const int ch = (int)(u0 / R);
    const int ia = (int)(u0 - (long long)ch * R);
    const long long segend = u1 < (long long)(ch + 1) * R ? u1 : (long long)(ch + 1) * R;
    const int ib = ia + (int)(segend - u0);
    const int col = ch * CW + 2 * tx;
    const bool act = col < ld;
    const bool qx = col == q, qy = col + 1 == q;
    double2 rho = make_double2(0.0, 0.0);
    if (act) {
      const double2 t = ldcg2(Ts + (size_t)r * ld + col);
      rho.x = qx ? inv : -(t.x * inv);
      rho.y = qy ? inv : -(t.y * inv);
    }
    for (int s0 = ia; s0 < ib; s0 += STAGE_ROWS) {
      const int s1 = ib < s0 + STAGE_ROWS ? ib : s0 + STAGE_ROWS;
      __syncthreads();
      for (int k = threadIdx.x; k < s1 - s0; k += blockDim.x)
        sm.colS[k] = ldcg(Ts + (size_t)(s0 + k) * ld + q);
      __syncthreads();
      if (!act) continue;
      for (int i = s0 + ty; i < s1; i += trows * UNROLL) {
        double2 t[UNROLL];
#pragma unroll
        for (int u = 0; u < UNROLL; u++) {
          const int ii = i + u * trows;
          if (ii < s1) t[u] = ldcg2(Ts + (size_t)ii * ld + col);
        }
#pragma unroll
        for (int u = 0; u < UNROLL; u++) {
          const int ii = i + u * trows;
          if (ii < s1) {
            const double c = sm.colS[ii - s0];
            double2 o;
            o.x = fma(c, rho.x, qx ? 0.0 : t[u].x);
            o.y = fma(c, rho.y, qy ? 0.0 : t[u].y);
            if (ii == r) o = rho;
            stcg2(Td + (size_t)ii * ld + col, o);
          }
        }
      }
    }
    u0 = segend;
  }
}

// beta, basis heads and non-basic flags of the next state (ping-pong), split
// over all threads of the group.  r == 0: bound flip (no basis change).
__device__ void state_update(const LPDev &L, int cT, int cS, int r, int q, double delta,
                             double newbr, int kin, int kout, int newat, int g, int G) {
  const double *Ts = L.T[cT];
  const double *bs = L.beta[cS];
  double *bd = L.beta[cS ^ 1];
  const int *hs = L.head[cS], *ns = L.nb[cS], *as = L.at[cS];
  int *hd = L.head[cS ^ 1], *nd = L.nb[cS ^ 1], *ad = L.at[cS ^ 1];
  const int gt = g * blockDim.x + threadIdx.x, stride = G * blockDim.x;
  for (int i = 1 + gt; i <= L.m; i += stride) {
    const double a = ldcg(Ts + (size_t)i * L.ld + q);
    bd[i] = i == r ? newbr : fma(a, delta, ldcg(bs + i));
    hd[i] = i == r ? kin : ldcg(hs + i);
  }
  for (int j = gt; j < L.n; j += stride) {
    nd[j] = j == q ? kout : ldcg(ns + j);
    ad[j] = j == q ? newat : ldcg(as + j);
  }
}

__device__ int count_primal_infeas(const LPDev &L, int cS, Smem &sm) {
  const double *b = L.beta[cS];
  const int *h = L.head[cS];
  int cnt = 0;
  for (int i = 1 + threadIdx.x; i <= L.m; i += blockDim.x) {
    const int k = ldcg(h + i);
    cnt += row_infeas(ldcg(b + i), ldcg(L.lb + k), ldcg(L.ub + k)) != 0;
  }
  return block_sum(cnt, sm.redI);
}

__device__ int count_dual_infeas(const LPDev &L, int cT, int cS, Smem &sm) {
  const double *d = L.T[cT];
  const int *at = L.at[cS];
  int cnt = 0;
  for (int j = threadIdx.x; j < L.n; j += blockDim.x) cnt += dj_eligible(ldcg(at + j), ldcg(d + j));
  return block_sum(cnt, sm.redI);
}

// phase-1 partial sums: p1part[c][j] = sum over the 32 rows of chunk c of
// sign_i * T_ij, rows in order
__device__ void p1_partials(const LPDev &L, int cT, int cS, int g, int G) {
  const double *T = L.T[cT], *b = L.beta[cS];
  const int *h = L.head[cS];
  const int m = L.m, n = L.n, ld = L.ld;
  const int nchunks = (m + MVB_P1_CHUNK - 1) / MVB_P1_CHUNK;
  const long long items = (long long)nchunks * n;
  for (long long it = (long long)g * blockDim.x + threadIdx.x; it < items;
       it += (long long)G * blockDim.x) {
    const int c = (int)(it / n), j = (int)(it - (long long)c * n);
    const int i0 = 1 + c * MVB_P1_CHUNK;
    const int i1 = i0 + MVB_P1_CHUNK - 1 < m ? i0 + MVB_P1_CHUNK - 1 : m;
    double acc = 0.0;
    for (int i = i0; i <= i1; i++) {
      const int k = ldcg(h + i);
      const int s = row_infeas(ldcg(b + i), ldcg(L.lb + k), ldcg(L.ub + k));
      if (s > 0) acc = acc + ldcg(T + (size_t)i * ld + j);
      else if (s < 0) acc = acc - ldcg(T + (size_t)i * ld + j);
    }
    L.p1part[(size_t)c * ld + j] = acc;
  }
}

__device__ __forceinline__ double p1_cost(const LPDev &L, int j) {
  const int nchunks = (L.m + MVB_P1_CHUNK - 1) / MVB_P1_CHUNK;
  double d = 0.0;
  for (int c = 0; c < nchunks; c++) d = d + ldcg(L.p1part + (size_t)c * L.ld + j);
  return d;
}

}  // namespace
