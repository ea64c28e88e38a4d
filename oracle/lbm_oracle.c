/*
 * lbm_oracle.c -- TEST INFRASTRUCTURE ONLY.  CPU restatement of the D2Q9 BGK
 * timestep path of jodavies/latticeboltzmann (src/latticeboltzmann.c, "LB.c").
 *
 * Nothing in the product (latticeboltzmann_b200/, include/, driver/) may link,
 * import or execute this file.  It is the checker used by tests/, by
 * __graft_entry__.smoke() and by bench.py's cpu_baseline leg.
 *
 * Parity status: PINNED.  The reference ships no golden vectors (it has no
 * tests), so this restatement is pinned against the reference itself: the
 * unmodified LB.c compiled by oracle/build_ref.sh (oracle/_ref/liblbref_*.so,
 * -O2 -ffp-contract=off) must agree with it bit for bit, in SP and DP, for the
 * AVX, SSE and serial headers (tests/test_oracle_vs_ref.py, run where
 * /root/reference exists) and the resulting digests are committed under
 * tests/golden/ (tests/golden/make_golden.py).
 *
 * Compile once per precision:  -DLBO_DOUBLE  or  -DLBO_FLOAT,
 * always with  -O2 -ffp-contract=off  (no FMA contraction, no reassociation).
 *
 * Unlike the reference, lattice size, tau, accel and the emulated SIMD width
 * are run-time arguments, rows are indexed with size_t (LB.c:72,136 overflow
 * int above 2^31 elements) and the per-speed work is table driven.
 *
 * Layout (LB.c:71-72): f[s][i][j], element (s,i,j) at s*nx*ldy + i*ldy + j,
 * j contiguous; walls[i][j] chars with the same row stride ldy.
 */
#include <stddef.h>
#include <stdio.h>
#include <string.h>

#if defined(LBO_DOUBLE)
typedef double real;
#define R(x) (x)
#elif defined(LBO_FLOAT)
typedef float real;
#define R(x) x##f
#else
#error "define LBO_DOUBLE or LBO_FLOAT"
#endif

#define Q 9

/* where speed s is pulled from: (row shift, column shift); LB.c:235-243 */
static const int PULL_DI[Q] = { 0, 0, -1, 0, +1, -1, -1, +1, +1 };
static const int PULL_DJ[Q] = { 0, -1, 0, +1, 0, -1, +1, +1, -1 };
/* bounce-back partner: dst[s] = pulled[OPP[s]]; LB.c:247-254 */
static const int OPP[Q] = { 0, 3, 4, 1, 2, 7, 8, 5, 6 };
/* "reflect (slip)" walls -- named at LB.c:21, not implemented by the reference; defined in include/lbm_b200.h (wall
 * values 2 and 3): specular reflection about a wall whose normal is x (dx -> -dx) or y (dy -> -dy).  Parity unpinned. */
static const int MIRROR_X[Q] = { 0, 1, 4, 3, 2, 8, 7, 6, 5 };
static const int MIRROR_Y[Q] = { 0, 3, 2, 1, 4, 6, 5, 8, 7 };

typedef struct {
	int nx, ny;      /* lattice rows (x, slow) and columns (y, contiguous) */
	int ldy;         /* row stride in elements = NYPADDED (LB.c:71) */
	int vw;          /* emulated VECWIDTH: 1 serial, 2/4 SSE, 4/8 AVX */
	real tau, accel, rho0;
	int source_col;  /* column ApplySource acts on: 0 in the reference (LB.c:500-515).  Test windows cut out of a larger
	                    lattice (tests/test_gpu_fullsize.py) put the lattice's column 0 elsewhere, or nowhere (-1). */
	int outflow;     /* 0: y wraps (the reference, LB.c:233-234).  1: open y ends -- the option the reference's dead
	                    WRAPAROUND macro names (LB.c:42-43) and never implements; defined in include/lbm_b200.h
	                    (LBM_Y_OUTFLOW): what would be pulled from beyond an end column comes from that end column
	                    itself.  NOT pinned against the reference (it has no such code): parity unpinned. */
} lbo_params;

int lbo_sizeof_real(void) { return (int)sizeof(real); }

/* NYPADDED for a given ALIGNREQUIREMENT (LB.c:71) */
int lbo_padded_ny(int ny, int align) { return align * ((ny - 1) / align) + align; }

static inline size_t at(const lbo_params *p, int s, int i, int j)
{
	return ((size_t)s * (size_t)p->nx + (size_t)i) * (size_t)p->ldy + (size_t)j;
}

static void weights(real w[Q])
{
	/* OMEGA0/14/58 formed from REAL_T literals in real precision (LB.c:38-40) */
	w[0] = R(4.0) / R(9.0);
	for (int s = 1; s <= 4; s++) w[s] = R(1.0) / R(9.0);
	for (int s = 5; s <= 8; s++) w[s] = R(1.0) / R(36.0);
}

/* gather the nine incoming populations of site (i,j) with periodic wrap */
static inline void pull(const lbo_params *p, const real *src, int i, int j, real f[Q])
{
	const int up = (i + 1) % p->nx, dn = i ? i - 1 : p->nx - 1;   /* LB.c:231-232 */
	int rt = (j + 1) % p->ny, lf = j ? j - 1 : p->ny - 1;         /* LB.c:233-234 */
	if (p->outflow) {                                             /* open ends: the end column itself */
		if (j == p->ny - 1) rt = j;
		if (j == 0) lf = j;
	}
	for (int s = 0; s < Q; s++) {
		const int ii = PULL_DI[s] < 0 ? dn : PULL_DI[s] > 0 ? up : i;
		const int jj = PULL_DJ[s] < 0 ? lf : PULL_DJ[s] > 0 ? rt : j;
		f[s] = src[at(p, s, ii, jj)];
	}
}

/* wall site: kind 1 bounces back (LB.c:246-255); 2 / 3 reflect (see MIRROR_X / MIRROR_Y) */
static inline void bounce(const lbo_params *p, real *dst, int i, int j, const real f[Q], int kind)
{
	const int *perm = kind == 2 ? MIRROR_X : kind == 3 ? MIRROR_Y : OPP;
	for (int s = 1; s < Q; s++)            /* speed 0 is NOT written at walls */
		dst[at(p, s, i, j)] = f[perm[s]];
}

/*
 * Equilibrium in the operation order of the SIMD path, LB.c:360-433.
 * Every intermediate is a separate statement in `real`, so with
 * -ffp-contract=off each one is a single correctly rounded IEEE operation.
 */
static inline void feq_vector_order(const real f[Q], const real w[Q], real feq[Q])
{
	real rho = R(0.0);
	for (int s = 0; s < Q; s++) rho = rho + f[s];              /* LB.c:360,372-380 */

	real ux = f[6] + f[2];  ux = ux + f[5];  ux = ux - f[7];   /* LB.c:382-387 */
	ux = ux - f[4];  ux = ux - f[8];  ux = ux / rho;
	real uy = f[5] + f[1];  uy = uy + f[8];  uy = uy - f[6];   /* LB.c:389-394 */
	uy = uy - f[3];  uy = uy - f[7];  uy = uy / rho;

	const real icsq = R(1.0) / R(1.0);                         /* CSQ == 1, LB.c:49,318 */
	const real xx = ux * ux, yy = uy * uy;
	const real usq = R(1.5) * (icsq * (xx + yy));              /* LB.c:397 */

	const real k = R(3.0) * icsq;                              /* LB.c:400-401 */
	ux = ux * k;
	uy = uy * k;
	const real sum = ux + uy, dif = ux - uy;
	real e[Q];
	e[1] = uy;            e[2] = ux;                           /* LB.c:402-409 */
	e[3] = R(-1.0) * uy;  e[4] = R(-1.0) * ux;
	e[5] = sum;           e[6] = dif;
	e[7] = R(-1.0) * sum; e[8] = R(-1.0) * dif;

	feq[0] = R(1.0) - usq;                                     /* LB.c:413 */
	for (int s = 1; s < Q; s++) {                              /* LB.c:414-421 */
		const real sq = e[s] * e[s];
		const real hs = R(0.5) * sq;
		const real d = hs - usq;
		const real a = e[s] + d;
		feq[s] = R(1.0) + a;
	}
	for (int s = 0; s < Q; s++) {                              /* LB.c:425-433 */
		const real wr = w[s] * rho;
		feq[s] = feq[s] * wr;
	}
}

/* relaxation with 1/tau held in DOUBLE: the scalar statements LB.c:295, 456-464 */
static inline real relax_double(real f, real feq, real tau)
{
	const real d = feq - f;                     /* real - real, FLT_EVAL_METHOD 0 */
	const double itau = 1.0 / (double)tau;
	return (real)((double)f + itau * (double)d);
}

/* one aligned group of vw sites starting at column j: LB.c:331-479 */
static void collide_group(const lbo_params *p, const real *src, real *dst,
                          const char *walls, const real w[Q], int i, int j)
{
	real f[8][Q], feq[8][Q];
	int blocked = 0;
	for (int k = 0; k < p->vw; k++) {
		pull(p, src, i, j + k, f[k]);
		feq_vector_order(f[k], w, feq[k]);
		blocked += walls[(size_t)i * p->ldy + j + k];          /* LB.c:437-440 */
	}
	const real itau = R(1.0) / p->tau;                         /* LB.c:323 */
	for (int k = 0; k < p->vw; k++) {
		if (!blocked) {                                        /* LB.c:441-451 */
			for (int s = 0; s < Q; s++) {
				const real d = feq[k][s] - f[k][s];
				const real t = itau * d;
				dst[at(p, s, i, j + k)] = f[k][s] + t;
			}
		} else if (walls[(size_t)i * p->ldy + j + k] == 0) {   /* LB.c:455-465 */
			for (int s = 0; s < Q; s++)
				dst[at(p, s, i, j + k)] = relax_double(f[k][s], feq[k][s], p->tau);
		} else {
			bounce(p, dst, i, j + k, f[k], walls[(size_t)i * p->ldy + j + k]);   /* LB.c:466-475 */
		}
	}
}

/* scalar site update in the rounding order of StreamCollide, LB.c:227-300 */
static void collide_scalar(const lbo_params *p, const real *src, real *dst,
                           const char *walls, const real w[Q], int i, int j)
{
	real f[Q];
	pull(p, src, i, j, f);
	if (walls[(size_t)i * p->ldy + j] != 0) {                  /* LB.c:246 (== 1 there; 2, 3: the reflect walls) */
		bounce(p, dst, i, j, f, walls[(size_t)i * p->ldy + j]);
		return;
	}
	real rho = R(0.0);
	for (int s = 0; s < Q; s++) rho += f[s];                   /* LB.c:258-261 */

	const real ux = ((f[6] + f[2] + f[5]) - (f[7] + f[4] + f[8])) / rho;   /* LB.c:263-264 */
	const real uy = ((f[5] + f[1] + f[8]) - (f[6] + f[3] + f[7])) / rho;   /* LB.c:265-266 */
	const real usq = ux * ux + uy * uy;                        /* LB.c:268 */

	real e[Q];                                                 /* LB.c:271-279 */
	e[0] = R(0.0);
	e[1] = +uy;       e[2] = +ux;
	e[3] = -uy;       e[4] = -ux;
	e[5] = +ux + uy;  e[6] = +ux - uy;
	e[7] = -ux - uy;  e[8] = -ux + uy;

	const real csq = R(1.0);
	real feq[Q];
	/* LB.c:283: speed 0 has no e.u terms at all */
	feq[0] = w[0] * rho * (R(1.0) - R(3.0) / R(2.0) * usq / csq);
	for (int s = 1; s < Q; s++)                                /* LB.c:284-291, left to right */
		feq[s] = w[s] * rho * (R(1.0) + R(3.0) * e[s] / csq
		                       + R(9.0) / R(2.0) * e[s] * e[s] / csq / csq
		                       - R(3.0) / R(2.0) * usq / csq);

	for (int s = 0; s < Q; s++)                                /* LB.c:294-296 */
		dst[at(p, s, i, j)] = relax_double(f[s], feq[s], p->tau);
}

/* LB.c:216-302 (jmin..jmax scalar) */
void lbo_stream_collide(const lbo_params *p, int jmin, int jmax,
                        const real *src, real *dst, const char *walls)
{
	real w[Q];
	weights(w);
#pragma omp parallel for schedule(static)
	for (int i = 0; i < p->nx; i++)
		for (int j = jmin; j < jmax; j++)
			collide_scalar(p, src, dst, walls, w, i, j);
}

/* LB.c:307-484 (jmin..jmax in aligned groups of vw) */
void lbo_stream_collide_vec(const lbo_params *p, int jmin, int jmax,
                            const real *src, real *dst, const char *walls)
{
	real w[Q];
	weights(w);
#pragma omp parallel for schedule(static)
	for (int i = 0; i < p->nx; i++)
		for (int j = jmin; j < jmax; j += p->vw)
			collide_group(p, src, dst, walls, w, i, j);
}

/* body force on column 0 (p->source_col), in place; LB.c:489-518 */
void lbo_apply_source(const lbo_params *p, real *f, const char *walls)
{
	real w[Q];
	weights(w);
	const real a14 = p->accel * w[1], a58 = p->accel * w[5];
	const int c = p->source_col;
	if (c < 0 || c >= p->ny) return;
	for (int i = 0; i < p->nx; i++) {
		if (walls[(size_t)i * p->ldy + c] != 0) continue;
		real *f1 = &f[at(p, 1, i, c)], *f3 = &f[at(p, 3, i, c)], *f5 = &f[at(p, 5, i, c)];
		real *f6 = &f[at(p, 6, i, c)], *f7 = &f[at(p, 7, i, c)], *f8 = &f[at(p, 8, i, c)];
		if (*f6 - a58 > R(0.0) && *f3 - a14 > R(0.0) && *f7 - a58 > R(0.0)) {
			*f6 -= a58;  *f3 -= a14;  *f7 -= a58;
			*f5 += a58;  *f1 += a14;  *f8 += a58;
		}
	}
}

static void one_step(const lbo_params *p, real *src, real *dst, const char *walls)
{
	lbo_apply_source(p, src, walls);
	if (p->vw > 1) {
		const int nyvec = p->vw * (p->ny / p->vw);             /* NYVECMAX, LB.c:76 */
		lbo_stream_collide_vec(p, 0, nyvec, src, dst, walls);
		lbo_stream_collide(p, nyvec, p->ny, src, dst, walls);
	} else {
		lbo_stream_collide(p, 0, p->ny, src, dst, walls);
	}
}

/* two lattice steps A->B->A; LB.c:186-210 */
void lbo_do_timestep(const lbo_params *p, real *fA, real *fB, const char *walls)
{
	one_step(p, fA, fB, walls);
	one_step(p, fB, fA, walls);
}

/* nsteps must be even (the reference loop advances by 2, LB.c:148) */
void lbo_run(const lbo_params *p, real *fA, real *fB, const char *walls, int nsteps)
{
	for (int n = 0; n < nsteps; n += 2) lbo_do_timestep(p, fA, fB, walls);
}

/* LB.c:522-547 */
real lbo_compute_reynolds(const lbo_params *p, const real *f, const char *walls)
{
	const int j = (int)((real)p->ny / 2.0);
	real total = 0;
	int count = 0;
	for (int i = 0; i < p->nx; i++) {
		if (walls[(size_t)i * p->ldy + j] != 0) continue;
		real v[Q], rho = 0.0;
		for (int s = 0; s < Q; s++) { v[s] = f[at(p, s, i, j)]; rho += v[s]; }
		total += ((v[5] + v[1] + v[8]) - (v[6] + v[3] + v[7])) / rho;
		count++;
	}
	const real nu = R(1.0) / R(3.0) * (p->tau - R(1.0) / R(2.0));
	return total / (real)count * R(10.0) / nu;
}

/*
 * Built-in scenario, LB.c:551-606: channel walls on rows 0 and nx-1, a barrier
 * rows [20,220) x columns 100..104 (absolute indices, clipped here to the
 * lattice so small test lattices stay legal), fA at rest with density rho0,
 * fB zero.  Padding columns j >= ny are left untouched, as in the reference.
 */
void lbo_initialize_arrays(const lbo_params *p, real *fA, real *fB, char *walls)
{
	real w[Q];
	weights(w);
	for (int i = 0; i < p->nx; i++)
		for (int j = 0; j < p->ny; j++)
			walls[(size_t)i * p->ldy + j] = 0;
	for (int i = 20; i < 220 && i < p->nx; i++)
		for (int j = 100; j < 105 && j < p->ny; j++)
			walls[(size_t)i * p->ldy + j] = 1;
	for (int j = 0; j < p->ny; j++) {
		walls[j] = 1;
		walls[(size_t)(p->nx - 1) * p->ldy + j] = 1;
	}
	for (int s = 0; s < Q; s++) {
		const real v = p->rho0 * w[s];
		for (int i = 0; i < p->nx; i++)
			for (int j = 0; j < p->ny; j++) {
				fA[at(p, s, i, j)] = v;
				fB[at(p, s, i, j)] = R(0.0);
			}
	}
}

/* per-site rho, u_x, u_y in the order of PrintLattice, LB.c:621-629 */
void lbo_macroscopic(const lbo_params *p, const real *f, real *rho_out, real *ux_out, real *uy_out)
{
	for (int i = 0; i < p->nx; i++)
		for (int j = 0; j < p->ny; j++) {
			real v[Q], rho = 0;
			for (int s = 0; s < Q; s++) { v[s] = f[at(p, s, i, j)]; rho += v[s]; }
			const size_t o = (size_t)i * p->ny + j;
			rho_out[o] = rho;
			ux_out[o] = ((v[6] + v[2] + v[5]) - (v[7] + v[4] + v[8])) / rho;
			uy_out[o] = ((v[5] + v[1] + v[8]) - (v[6] + v[3] + v[7])) / rho;
		}
}

/* CSV of u_x^2+u_y^2, "%.10lf" joined by ", ", one row per line; LB.c:610-639 */
int lbo_print_lattice(const lbo_params *p, const real *f, const char *path)
{
	FILE *fp = fopen(path, "w");
	if (!fp) return -1;
	for (int i = 0; i < p->nx; i++) {
		for (int j = 0; j < p->ny; j++) {
			real v[Q], rho = 0;
			for (int s = 0; s < Q; s++) { v[s] = f[at(p, s, i, j)]; rho += v[s]; }
			const real ux = ((v[6] + v[2] + v[5]) - (v[7] + v[4] + v[8])) / rho;
			const real uy = ((v[5] + v[1] + v[8]) - (v[6] + v[3] + v[7])) / rho;
			const real u2 = ux * ux + uy * uy;
			fprintf(fp, "%.10lf", u2);
			if (j < p->ny - 1) fprintf(fp, ", ");
		}
		fprintf(fp, "\n");
	}
	fclose(fp);
	return 0;
}
