#!/bin/sh
# TEST INFRASTRUCTURE ONLY -- not part of the product path.
#
# Compiles the UNMODIFIED reference translation unit (src/latticeboltzmann.c,
# 662 lines + the prec_*.h headers) from where it lies under $LB_REFERENCE
# (default /root/reference) into oracle/_ref/.  No reference source is copied
# into this repository: the file is streamed through `sed` straight into gcc's
# stdin and only binaries land in oracle/_ref/ (git-ignored, NOT gpurun-ignored,
# so the prebuilt files travel to the GPU box where /root/reference is absent).
#
# Two kinds of artefact:
#
#   strict parity libraries  liblbref_{dp,sp}_{avx,sse,serial}.so
#       -O2 -ffp-contract=off (IEEE semantics, no reassociation, no FMA).
#       NX, NY, TAU, ACCEL, INITIALDENSITY become run-time globals
#       (lbref_nx, ...) so ONE library serves every parity lattice; main is
#       renamed lbref_main.  The functions the tests call are the reference's
#       own: DoTimeStep, ApplySource, StreamCollide[Vec], ComputeReynolds,
#       InitializeArrays, PrintLattice (LB.c:85-123).
#
#   fast timed executables   lbref_fast_{dp,sp}_<NX>x<NY>_<STEPS>
#       Makefile:2 flags (-Ofast -funroll-loops -march=core-avx-i -fopenmp)
#       with compile-time lattice constants, exactly like the shipped program;
#       timed by the program's own "Runtime:" line (LB.c:166,173).
#
# The edits applied on the fly (none of them touches arithmetic):
#   * the precision/vector header chosen at LB.c:59-65
#   * `default(none) shared(...)` dropped from the four OpenMP pragmas
#     (LB.c:226,328,559,580): gcc >= 9 rejects them because const int
#     parameters / file-scope variables are no longer predetermined shared
#   * the unconditional #defines of NX, NY, NTIMESTEPS (LB.c:46-51) rewritten
set -eu

REF="${LB_REFERENCE:-/root/reference}"
HERE="$(cd "$(dirname "$0")" && pwd)"
OUT="$HERE/_ref"
SRC="$REF/src/latticeboltzmann.c"

if [ ! -f "$SRC" ]; then
	echo "build_ref.sh: $SRC not found (reference absent) -- keeping prebuilt files in $OUT" >&2
	exit 0
fi
mkdir -p "$OUT"

# deliberately NOT $CC: this image exports CC=/opt/gcc/bin/gcc, which has no libgomp.spec
CC="${LB_HOST_CC:-/usr/bin/gcc}"

# $1 = header stem (e.g. double_avx); emits the patched TU with run-time lattice parameters
patch_runtime() {
	sed -e 's|^#include "prec_float_avx.h"|#include "prec_'"$1"'.h"\
int lbref_nx = 400, lbref_ny = 2000;\
real_t lbref_tau = REAL_T(0.7), lbref_accel = REAL_T(0.005), lbref_rho0 = REAL_T(0.1);|' \
	    -e 's|^#define NX 400|#define NX lbref_nx|' \
	    -e 's|^#define NY 2000|#define NY lbref_ny|' \
	    -e 's|^#define TAU REAL_T(0.7)|#define TAU lbref_tau|' \
	    -e 's|^#define ACCEL REAL_T(0.005)|#define ACCEL lbref_accel|' \
	    -e 's|^#define INITIALDENSITY REAL_T(0.1)|#define INITIALDENSITY lbref_rho0|' \
	    -e 's| default(none) shared([A-Za-z,]*)||' \
	    -e 's|^int main(void)|int lbref_main(void)|' \
	    "$SRC"
}

# $1 = header stem, $2 = NX, $3 = NY, $4 = NTIMESTEPS
patch_fixed() {
	sed -e 's|^#include "prec_float_avx.h"|#include "prec_'"$1"'.h"|' \
	    -e 's|^#define NX 400|#define NX '"$2"'|' \
	    -e 's|^#define NY 2000|#define NY '"$3"'|' \
	    -e 's|^#define NTIMESTEPS 10000|#define NTIMESTEPS '"$4"'|' \
	    -e 's| default(none) shared([A-Za-z,]*)||' \
	    "$SRC"
}

STRICT="-std=gnu99 -O2 -ffp-contract=off -march=core-avx-i -fopenmp -fPIC -shared"
for prec in double float; do
	case $prec in double) tag=dp ;; float) tag=sp ;; esac
	for isa in avx sse serial; do
		patch_runtime "${prec}_${isa}" | \
			$CC -x c - -I"$REF/src" $STRICT -o "$OUT/liblbref_${tag}_${isa}.so" -lm
	done
done

# Timed baseline: Makefile:2 flags.  core-avx-i (Ivy Bridge: AVX, no FMA) runs on
# any AVX host, so the prebuilt binary is safe on the GPU box.
FAST="-std=gnu99 -Ofast -funroll-loops -march=core-avx-i -fopenmp"
build_fast() { # prec tag NX NY STEPS
	patch_fixed "$1_avx" "$3" "$4" "$5" | \
		$CC -x c - -I"$REF/src" $FAST -o "$OUT/lbref_fast_$2_$3x$4_$5" -lrt -lm
}
build_fast double dp 400 2000 1000
build_fast double dp 400 2000 10000
build_fast double dp 400 4000 1000
build_fast double dp 400 4000 10000
build_fast float  sp 400 4000 1000
build_fast float  sp 400 4000 10000
build_fast double dp 800 4000 1000

# The weak-scaling unit (BASELINE configs[4], 8192x32768 per GPU) for `bench.py --impl reference`: 9*NX*NY = 2.4e9
# elements overflow the reference's int indexing (LB.c:72 `I()`, LB.c:136 `int allocSize`), so this one build also
# widens exactly those two expressions to size_t (SURVEY.md 8(c) patch (5)); nothing else changes.
build_fast_big() { # prec tag NX NY STEPS
	patch_fixed "$1_avx" "$3" "$4" "$5" | \
	sed -e 's|^#define I(i,j, speed) ((speed)\*NX\*NYPADDED + (i)\*NYPADDED + (j))|#define I(i,j, speed) ((size_t)(speed)*NX*NYPADDED + (size_t)(i)*NYPADDED + (j))|' \
	    -e 's|int allocSize = NX \* NYPADDED \* NSPEEDS;|size_t allocSize = (size_t)NX * NYPADDED * NSPEEDS;|' \
	    -e 's|walls = _mm_malloc(NX \* NYPADDED \* sizeof \*walls, 32);|walls = _mm_malloc((size_t)NX * NYPADDED * sizeof *walls, 32);|' | \
		$CC -x c - -I"$REF/src" $FAST -o "$OUT/lbref_fast_$2_$3x$4_$5" -lrt -lm
}
build_fast_big double dp 8192 32768 10
build_fast_big float  sp 8192 32768 10
# the same patch on a small lattice: tests/test_oracle_golden.py checks it prints the Reynolds number of the unpatched build
build_fast_big double dp 400 2000 1000
mv "$OUT/lbref_fast_dp_400x2000_1000" "$OUT/lbref_fastbig_dp_400x2000_1000"
build_fast double dp 400 2000 1000

echo "build_ref.sh: built into $OUT:" >&2
ls "$OUT" >&2
