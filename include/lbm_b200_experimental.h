/*
 * lbm_b200_experimental.h -- knobs of the EXPERIMENT build of the library only (make -C latticeboltzmann_b200/csrc
 * experimental -> latticeboltzmann_b200/exp/liblbm_b200_{dp,sp}.so, compiled with -DLBM_EXPERIMENTAL).  The product
 * libraries do not export lbm_set_experimental and do not contain these kernels.
 *
 * What is behind it: the in-place step with threads walking several rows and the NEXT row's nine populations staged in
 * shared memory while the current row is relaxed -- by cp.async (LDGSTS), or, for whole-row launches of the local phase,
 * by TMA bulk copies both ways (cp.async.bulk + mbarrier; UBLKCP in SASS).  north_star names "TMA or shared-memory
 * staging of the neighbouring x-rows"; it is bit-identical (tests/test_gpu_parity.py::test_staged_kernel_is_bit_identical)
 * and measured SLOWER than the register path on every lattice (profiles/r02_run1: DP 8192x32768 52.3 vs 61.6 GLUPS, SP
 * 400x4000 86.7 vs 119.2), because the path is bound by L2 bandwidth, not by load latency.
 */
#ifndef LBM_B200_EXPERIMENTAL_H
#define LBM_B200_EXPERIMENTAL_H
#include "lbm_b200.h"
#ifdef __cplusplus
extern "C" {
#endif
typedef struct {
	int stage_smem;       /* 0 off; 1 = cp.async staging in both phases; 2 = phase 2 by TMA bulk copies (phase 1: cp.async) */
	int rows_per_thread;  /* >= 2: rows each thread walks (the staging pipelines over them); vec must be 1 */
	int block;            /* threads per CTA of the step kernels: 32..512, a multiple of 32; 0 = the product's 128 (the SP in-place kernels run
	                         at most 256) */
	int smem_pad;         /* bytes of (unused) dynamic shared memory requested by every in-place step launch: caps the CTAs
	                         resident per SM -- the residency probe (a CTA's lifetime is what one launch loses at its drain) */
} lbm_experimental;
int lbm_set_experimental(lbm_ctx *ctx, const lbm_experimental *x);
#ifdef __cplusplus
}
#endif
#endif
