/* Single precision on the B200 path.  Same idiom as the reference's
 * src/prec_float_{avx,sse,serial}.h:1-4 -- include exactly one prec_*_cuda.h
 * before lbm_b200.h; it fixes real_t for the whole translation unit and selects
 * which library (liblbm_b200_sp.so) the program must be linked against. */
#ifndef LBM_PREC_SELECTED
#define LBM_PREC_SELECTED 1
#define SINGLEPREC 1
typedef float real_t;
#define REAL_T(x) x##f
#define LBM_PREC_TAG "sp"
/* SIMD group width of the reference build whose rounding we reproduce
 * (prec_float_avx.h:8-9).  In SP it matters numerically: an aligned group of
 * VECWIDTH sites that contains a wall relaxes its fluid lanes with 1/tau held
 * in double (LB.c:452-465); see lbm_config.vecwidth. */
#define VECWIDTH 8
#define ALIGNREQUIREMENT 8
#else
#error "include only one prec_*_cuda.h"
#endif
