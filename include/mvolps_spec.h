/* mvolps_spec.h -- numeric constants of the simplex specification shared by
 * the CUDA engine and the CPU oracle (constants only; each side implements
 * the arithmetic independently).  See DESIGN.md section 4 for the algorithm.
 */
#ifndef MVOLPS_SPEC_H
#define MVOLPS_SPEC_H

/* reduced-cost (dual feasibility) tolerance, absolute */
#define MVB_TOL_DJ 1e-9
/* bound (primal feasibility) tolerance: viol > TOL_BND*(1+|bound|) */
#define MVB_TOL_BND 1e-9
/* minimum |pivot candidate| admitted by the ratio tests, absolute */
#define MVB_TOL_PIV 1e-9
/* consecutive degenerate iterations before Bland's rule takes over */
#define MVB_BLAND_AFTER 40
/* rows per partial sum of the phase-1 reduced-cost rebuild */
#define MVB_P1_CHUNK 32
/* lanes of the deterministic dot product (beta rebuild) */
#define MVB_DOT_LANES 32
/* row stride alignment of the tableau in doubles (128 bytes) */
#define MVB_LD_ALIGN 16

/* FNV-1a 64 */
#define MVB_FNV_OFFSET 1469598103934665603ULL
#define MVB_FNV_PRIME 1099511628211ULL

#endif
