/* Double precision on the B200 path.  Same idiom as the reference's
 * src/prec_double_{avx,sse,serial}.h:1-4 -- include exactly one prec_*_cuda.h
 * before lbm_b200.h; it fixes real_t for the whole translation unit and selects
 * which library (liblbm_b200_dp.so) the program must be linked against. */
#ifndef LBM_PREC_SELECTED
#define LBM_PREC_SELECTED 1
#define DOUBLEPREC 1
typedef double real_t;
#define REAL_T(x) (x)
#define LBM_PREC_TAG "dp"
/* SIMD group width of the reference build whose rounding we reproduce
 * (prec_double_avx.h:8-9).  On the GPU it only decides where the scalar
 * tail starts (NYVECMAX, LB.c:76); see lbm_config.vecwidth. */
#define VECWIDTH 4
#define ALIGNREQUIREMENT 4
#else
#error "include only one prec_*_cuda.h"
#endif
