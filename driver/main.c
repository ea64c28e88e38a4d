/*
 * main.c -- the reference's driver (LB.c:127-182) over the B200 C ABI.  Plain C (gnu99),
 * no CUDA in this file: it only calls include/lbm_b200.h.
 *
 * Precision is chosen like LB.c:59-65, by header: compile with -DLBM_DRIVER_FLOAT for
 * prec_float_cuda.h, otherwise prec_double_cuda.h, and link the matching library.
 *
 * The reference's compile-time #defines (LB.c:46-56) are command-line arguments here:
 *   lbm_b200_{dp,sp} [NX NY NTIMESTEPS [NGPU [SCENE [DUMP_EVERY [INPLACE]]]]]
 *     SCENE: 0 = InitializeArrays (channel + barrier, LB.c:559-578)
 *            1 = channel + disc obstacle (BASELINE config 3: centre (NX/2, NY/5), radius NX/10)
 *     DUMP_EVERY > 0 writes data/<step>.csv like SAVELATTICE (LB.c:155-159; data/ must exist), one file for the
 *            whole lattice whatever NGPU
 *     INPLACE = 1 (default): one buffer per slab instead of fA/fB (lbm_config.inplace) -- this main never reads fB
 *            (LB.c:152,157,172-173 take fA), fA after every DoTimeStep is bit-identical, and the steps between two
 *            observation points run temporally blocked in L2; 0 = the reference's fA/fB pair
 * Defaults reproduce the shipped program: 400 x 2000, 10000 steps.
 */
#ifdef LBM_DRIVER_FLOAT
#include "prec_float_cuda.h"
#else
#include "prec_double_cuda.h"
#endif
#include "lbm_b200.h"

#include <stdio.h>
#include <stdlib.h>
#include <time.h>

#define NSPEEDS 9
#define TAU REAL_T(0.7)
#define ACCEL REAL_T(0.005)
#define PRINTSTATSEVERY 1000
#define FLOPPERLATTICEPOINT (124.0)   /* LB.c:80 */
#define MAXGPU 16

static int NX = 400, NY = 2000, NTIMESTEPS = 10000, NGPU = 1;

static double GetWallTime(void)       /* LB.c:643-648 */
{
	struct timespec tv;
	clock_gettime(CLOCK_REALTIME, &tv);
	return (double)tv.tv_sec + 1e-9 * (double)tv.tv_nsec;
}

/* PrintRunStats (LB.c:652-662); bandwidth counted as 2*9*sizeof(real_t) bytes per site update
 * (what the kernel moves), plus MLUPS */
static void PrintRunStats(int n, double startTime)
{
	double complete = (double)n / (double)NTIMESTEPS;
	double timeElap = GetWallTime() - startTime;
	double timeRem = timeElap / complete * (1.0 - complete);
	double updates = (double)NX * NY * n;
	double avgbw = 2.0 * NSPEEDS * sizeof(real_t) * updates / timeElap / 1e9;
	printf("%5.2lf%%--Elapsed: %3dm%02ds, Remaining: %3dm%02ds. [Updates/s: %.3le, MLUPS: %.1lf, Update BW: ~%.3lf GB/s, GFLOPs: ~%.3lf]\n",
	       complete * 100, (int)timeElap / 60, (int)timeElap % 60, (int)timeRem / 60, (int)timeRem % 60, n / timeElap,
	       updates / timeElap / 1e6, avgbw, FLOPPERLATTICEPOINT * updates / timeElap / 1e9);
}

#define CHECK(call)                                                            \
	do {                                                                       \
		int rc_ = (call);                                                      \
		if (rc_ != LBM_OK) {                                                   \
			fprintf(stderr, "%s failed (%d): %s\n", #call, rc_, lbm_last_error()); \
			return EXIT_FAILURE;                                               \
		}                                                                      \
	} while (0)

static lbm_ctx *slabs[MAXGPU];

static int sync_all(void)
{
	for (int g = 0; g < NGPU; g++)
		if (lbm_sync(slabs[g]) != LBM_OK) return 1;
	return 0;
}

/* The reference runs with FP traps on (feenableexcept, LB.c:129) and dies on the first NaN / overflow / division by
 * zero; here every observation point checks the lattice instead (one pass per PRINTSTATSEVERY steps). */
static int finite_all(void)
{
	for (int g = 0; g < NGPU; g++)
		if (lbm_check_finite(slabs[g], NULL) != LBM_OK) return 1;
	return 0;
}

/* ComputeReynolds over the slabs: concatenate the centre column in row order (LB.c:522-547) */
static int reynolds(real_t *re)
{
	real_t *uy = malloc((size_t)NX * sizeof *uy);
	char *fluid = malloc((size_t)NX);
	int at = 0, bad = !uy || !fluid;
	for (int g = 0; g < NGPU && !bad; g++) {
		lbm_slab_info info;
		bad = lbm_slab_info_get(slabs[g], &info) != LBM_OK || lbm_reynolds_column(slabs[g], uy + at, fluid + at) != LBM_OK;
		at += info.nx_local;
	}
	if (!bad) *re = lbm_reynolds_from_column(uy, fluid, NX, TAU);
	free(uy);
	free(fluid);
	return bad;
}

int main(int argc, char **argv)
{
	int scene_id = 0, dump_every = 0, inplace = 1;
	if (argc >= 4) { NX = atoi(argv[1]); NY = atoi(argv[2]); NTIMESTEPS = atoi(argv[3]); }
	if (argc >= 5) NGPU = atoi(argv[4]);
	if (argc >= 6) scene_id = atoi(argv[5]);
	if (argc >= 7) dump_every = atoi(argv[6]);
	if (argc >= 8) inplace = atoi(argv[7]) != 0;
	if (NGPU < 1 || NGPU > MAXGPU || NX < NGPU || (NTIMESTEPS & 1) || dump_every < 0) {
		fprintf(stderr, "usage: %s [NX NY NTIMESTEPS(even) [NGPU [SCENE [DUMP_EVERY [INPLACE]]]]]\n", argv[0]);
		return EXIT_FAILURE;
	}
	/* the loop advances two steps at a time (LB.c:148): an odd SAVELATTICEEVERY fires when n is a multiple of both */
	if (dump_every & 1) dump_every *= 2;

	/* LB.c:136-137 (the reference prints "%lf.2 MB") */
	printf("Lattice Size: %dx%d (%.2lf MB per buffer, %s, %d GPU%s%s)\n", NX, NY,
	       (double)NX * NY * NSPEEDS * sizeof(real_t) / 1024.0 / 1024.0, LBM_PREC_TAG, NGPU, NGPU > 1 ? "s" : "",
	       inplace ? ", one buffer in place" : "");

	lbm_scene scene = lbm_scene_reference();
	if (scene_id == 1) {
		scene.bar_i0 = scene.bar_i1 = 0;
		scene.disc_ci = NX / 2; scene.disc_cj = NY / 5; scene.disc_r = NX / 10;
	}
	for (int g = 0; g < NGPU; g++) {
		lbm_config cfg = {0};
		cfg.nx = NX; cfg.ny = NY; cfg.tau = TAU; cfg.accel = ACCEL;
		cfg.device = g;
		cfg.inplace = inplace;
		cfg.row_lo = (int)((long long)NX * g / NGPU);
		cfg.row_hi = (int)((long long)NX * (g + 1) / NGPU);
		CHECK(lbm_create(&slabs[g], &cfg));
		CHECK(lbm_initialize_arrays(slabs[g], &scene));          /* LB.c:142 */
	}
	if (NGPU > 1)
		for (int g = 0; g < NGPU; g++)
			CHECK(lbm_connect_local(slabs[g], slabs[(g + 1) % NGPU], slabs[(g + NGPU - 1) % NGPU]));
	if (sync_all()) { fprintf(stderr, "%s\n", lbm_last_error()); return EXIT_FAILURE; }

	/* Begin iterations (LB.c:146-164).  Steps between two observation points are enqueued in
	 * one call; the loop structure and the observation points are the reference's. */
	double startTime = GetWallTime();
	int chunk = PRINTSTATSEVERY;
	if (dump_every > 0)                                          /* both kinds of observation point fall on chunk boundaries */
		for (int a = chunk, b = dump_every; b; ) { int t = a % b; a = b; b = t; chunk = a; }
	for (int n = 0; n < NTIMESTEPS; n += chunk) {
		if (n % PRINTSTATSEVERY == 0 && n != 0) {
			if (sync_all() || finite_all()) { fprintf(stderr, "%s\n", lbm_last_error()); return EXIT_FAILURE; }
			PrintRunStats(n, startTime);
		}
		if (dump_every > 0 && n % dump_every == 0) {
			char filename[100];
			sprintf(filename, "data/%d.csv", n);                 /* LB.c:613 */
			CHECK(lbm_print_lattice_group(slabs, NGPU, filename));
		}
		int todo = NTIMESTEPS - n < chunk ? NTIMESTEPS - n : chunk;
		CHECK(lbm_group_run(slabs, NGPU, todo));                 /* DoTimeStep x todo/2, LB.c:162 */
	}
	if (sync_all() || finite_all()) { fprintf(stderr, "%s\n", lbm_last_error()); return EXIT_FAILURE; }
	double timeElapsed = GetWallTime() - startTime;

	PrintRunStats(NTIMESTEPS, startTime);                        /* LB.c:172 */
	real_t re = 0;
	if (reynolds(&re)) { fprintf(stderr, "%s\n", lbm_last_error()); return EXIT_FAILURE; }
	printf("Runtime: %lf Re %.10le\n", timeElapsed, (double)re); /* LB.c:173 */

	for (int g = 0; g < NGPU; g++) lbm_destroy(slabs[g]);
	return EXIT_SUCCESS;
}
