/*
 * lbm_b200.h -- C ABI of the B200-native D2Q9 BGK timestep path.
 *
 * Drop-in boundary for the hot path of jodavies/latticeboltzmann
 * (src/latticeboltzmann.c, "LB.c").  The reference has no plugin layer; its
 * boundary is the prototype block LB.c:85-123.  Every entry point below names
 * the reference function (file:line) it replaces.  Plain pointers and sizes
 * only; no CUDA or torch types cross this boundary.
 *
 * Precision is chosen at compile time exactly as in the reference
 * (LB.c:59-65): include prec_double_cuda.h or prec_float_cuda.h first, link
 * against liblbm_b200_dp.so or liblbm_b200_sp.so.  Both libraries export the
 * same symbol names; lbm_sizeof_real() lets a caller check the pairing.
 *
 * Host array layout accepted by upload/download is the reference's
 * (LB.c:71-72): f[s][i][j] at s*nx*ldy + i*ldy + j with j contiguous, walls
 * as char 0/1 with the same row stride ldy (= NYPADDED).  Packing the mask to
 * one bit per site and padding rows to 128 B happens inside.
 *
 * There is no CPU fallback: every call fails with LBM_ERR_CUDA when no
 * sm_100 device is usable.
 */
#ifndef LBM_B200_H
#define LBM_B200_H

#include <stddef.h>

#ifndef LBM_PREC_SELECTED
#error "include prec_double_cuda.h or prec_float_cuda.h before lbm_b200.h (cf. LB.c:59-65)"
#endif

#ifdef __cplusplus
extern "C" {
#endif

typedef struct lbm_ctx lbm_ctx; /* one slab of the lattice on one GPU */

enum {
	LBM_OK = 0,
	LBM_ERR_ARG = 1,    /* bad argument (odd step count, mask value not 0/1, ...) */
	LBM_ERR_CUDA = 2,   /* CUDA runtime/driver error, text in lbm_last_error() */
	LBM_ERR_STATE = 3,  /* call order (e.g. run before upload/initialize) */
	LBM_ERR_IO = 4,     /* PrintLattice / lbm_load_walls could not open or parse its file */
	LBM_ERR_FPE = 5     /* lbm_check_finite found NaN/Inf (stands in for the reference's FP traps, LB.c:129) */
};

enum { LBM_BUF_A = 0, LBM_BUF_B = 1 };

/* y (column) ends of the lattice -- the reference's dead WRAPAROUND macro (LB.c:42-43): "Wrap-around (periodic) or not
 * (fluid flows out of the domain)".  LBM_Y_PERIODIC is what the reference computes (LB.c:233-234, 337-338).
 * LBM_Y_OUTFLOW opens both ends: a population that would be pulled from beyond column 0 / NY-1 is taken from the end
 * column itself (same speed, same row shift: a zero-gradient end), what streams past an end leaves the lattice, and
 * ApplySource keeps forcing column 0.  The reference has no such code path: parity for it can only be checked against a CPU
 * restatement of THIS definition ("parity unpinned"). */
enum { LBM_Y_PERIODIC = 0, LBM_Y_OUTFLOW = 1 };

/* wall mask values (LB.c:140: char per site).  0 and 1 are the reference's; 2 and 3 are the "reflect (slip)" walls its
 * header comment names (LB.c:21) and never implements: specular reflection instead of reversal, about a wall whose normal
 * is the x (row) axis -- e.g. the channel walls -- or the y axis.  f0 of a wall site is left untouched either way
 * (LB.c:246-255).  Parity unpinned like LBM_Y_OUTFLOW. */
enum { LBM_FLUID = 0, LBM_WALL_BOUNCE = 1, LBM_WALL_SLIP_X = 2, LBM_WALL_SLIP_Y = 3 };

/* The compile-time constants of LB.c:46-56, as run-time fields. */
typedef struct {
	int nx, ny;        /* NX, NY (LB.c:46-47): global rows (x, slow) / columns (y, contiguous) */
	real_t tau;        /* TAU (LB.c:48) */
	real_t accel;      /* ACCEL (LB.c:55) */
	int vecwidth;      /* VECWIDTH of the reference build being matched (prec_*.h):
	                      4/2/1 DP, 8/4/1 SP; 0 = the prec_*_cuda.h default (AVX).
	                      1 selects the rounding order of StreamCollide (LB.c:216-302)
	                      everywhere, >1 that of StreamCollideVec (LB.c:307-484) for
	                      j < NYVECMAX and of StreamCollide for the tail columns. */
	int device;        /* CUDA device ordinal */
	int row_lo, row_hi;/* this slab owns global rows [row_lo,row_hi); 0,0 = whole lattice.
	                      The reference's MPI row-slab decomposition (README.md:44-53). */
	int inplace;       /* 0: A/B double buffer as LB.c:186-210 (fA and fB both observable).
	                      1: ONE buffer, in-place "AA pattern" streaming: half the memory, the
	                      working set of a step is the lattice once instead of twice.  Buffer A
	                      after every DoTimeStep / even step count is bit-identical to the
	                      A/B result; buffer B does not exist (fB arguments must be NULL,
	                      lbm_stream_collide and the staged halo transport are refused). */
	int boundary_y;    /* LBM_Y_PERIODIC (0, the reference) or LBM_Y_OUTFLOW.  With inplace = 1 open ends need the
	                      vector kernels on every column (vecwidth > 1 and ny a multiple of it). */
} lbm_config;

/* Closed-form scenes.  lbm_scene_reference() is InitializeArrays (LB.c:551-606). */
typedef struct {
	real_t rho0;                 /* INITIALDENSITY (LB.c:56): fA = rho0*omega_s at rest, fB = 0 */
	int channel_walls;           /* 1: rows 0 and nx-1 solid, bounce-back (LB.c:575-578); 2: the same rows as reflect (slip) walls; 0: none */
	int bar_i0, bar_i1, bar_j0, bar_j1; /* solid box rows [i0,i1) x cols [j0,j1) (LB.c:567-573); empty if i0>=i1 */
	long long disc_ci, disc_cj, disc_r; /* solid disc (i-ci)^2+(j-cj)^2 <= r^2 in integers; r<0: none */
	unsigned perturb_seed;       /* 0: uniform density rho0 (LB.c:583-591).  != 0: a closed-form rough initial state for
	                                full-size parity checks, rho(i,j) = rho0*(1 + amp*(h/2^32 - 1/2)) with h the 32-bit
	                                integer hash of (global row, column, seed); fA = rho*omega_s, every operation rounded
	                                once in real_t so that a host can reproduce any window of it bit for bit */
	real_t perturb_amp;
} lbm_scene;

int lbm_sizeof_real(void);            /* 8 (dp library) or 4 (sp library) */
const char *lbm_last_error(void);     /* text of the last failure on this thread */
const char *lbm_version(void);

/* ---- lifetime (replaces the _mm_malloc/_mm_free block LB.c:136-140,177-179) ---- */
int lbm_create(lbm_ctx **out, const lbm_config *cfg);
int lbm_destroy(lbm_ctx *ctx);

/* ---- inputs ---- */
/* InitializeArrays (LB.c:551-606), evaluated on the device for the rows this slab owns. */
lbm_scene lbm_scene_reference(void);
int lbm_initialize_arrays(lbm_ctx *ctx, const lbm_scene *scene);
/* Host arrays in the reference layout, covering the GLOBAL lattice (nx rows); the slab
 * takes its own rows.  fB may be NULL (zeros, LB.c:593-601).  walls values: LBM_FLUID .. LBM_WALL_SLIP_Y. */
int lbm_upload(lbm_ctx *ctx, const real_t *fA, const real_t *fB, const char *walls, int ldy);
/* The wall mask alone (global layout, as for lbm_upload); f is left as it is. */
int lbm_set_walls(lbm_ctx *ctx, const char *walls, int ldy);
/* Geometry from a file instead of the hard-wired barrier of LB.c:567-578: text, one lattice row per line, NY entries
 * per line, each one of the digits 0..3 (the mask values above), optionally separated by blanks and/or commas;
 * lines starting with '#' are comments; exactly NX rows.  Every slab reads the whole file and keeps its rows. */
int lbm_load_walls(lbm_ctx *ctx, const char *path);
/* Copies this slab's rows into global-layout host arrays; either pointer may be NULL.
 * walls_out (optional) receives the unpacked device bitmask as chars. */
int lbm_download(lbm_ctx *ctx, real_t *fA, real_t *fB, char *walls_out, int ldy);
/* Same, for host arrays that hold ONLY this slab's rows (what one MPI rank of the reference
 * would own): planes of nx_local rows, the slab's first row at row 0. */
int lbm_upload_slab(lbm_ctx *ctx, const real_t *fA, const real_t *fB, const char *walls, int ldy);
int lbm_download_slab(lbm_ctx *ctx, real_t *fA, real_t *fB, char *walls_out, int ldy);

/* A rectangular window of buffer `buf`: local rows [r0,r1) x columns [j0,j1) of this slab into a dense
 * out[9][r1-r0][j1-j0] (the reference has only whole-array observers; lattices of tens of GB want windows). */
int lbm_download_window(lbm_ctx *ctx, int buf, int r0, int r1, int j0, int j1, real_t *out);

/* ---- the timestep path ---- */
/* DoTimeStep (LB.c:186-210): Source(A); A->B; Source(B); B->A.  Asynchronous. */
int lbm_do_timestep(lbm_ctx *ctx);
/* The loop LB.c:148-164 without the prints: nsteps (even) lattice steps.  Asynchronous. */
int lbm_run(lbm_ctx *ctx, int nsteps);
/* Same, bracketed by CUDA events on the context's stream; *ms = device time. Synchronous. */
int lbm_run_timed(lbm_ctx *ctx, int nsteps, float *ms);
/* ApplySource (LB.c:489-518) on device buffer `buf`, in place. */
int lbm_apply_source(lbm_ctx *ctx, int buf);
/* StreamCollideVec + StreamCollide tail (LB.c:195-196 / 204-205): one pull step
 * src_buf -> the other buffer, no source term. */
int lbm_stream_collide(lbm_ctx *ctx, int src_buf);
int lbm_sync(lbm_ctx *ctx);

/* ---- observers ---- */
/* ComputeReynolds (LB.c:522-547) on buffer A: the centre column is gathered on the
 * device, summed on the host in the reference's serial order.  `column`/`valid`
 * variant below is for slab runs: each slab contributes its rows. */
int lbm_compute_reynolds(lbm_ctx *ctx, real_t *re);
/* u_y of the centre column for this slab's rows (uy[nx_local]) and the fluid flags
 * (LB.c:533-540); a multi-slab caller concatenates them in row order and finishes with
 * lbm_reynolds_from_column. */
int lbm_reynolds_column(lbm_ctx *ctx, real_t *uy, char *is_fluid);
/* The same column, for callers that must not allocate (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc).  NO COUNTERPART IN THE REFERENCE.  Calls no cudaMalloc / cudaFree /
 * cudaMallocAsync and no allocating library routine; the device scratch that lbm_reynolds_column allocates is a static device
 * array inside the library here (LBM_STATIC_MAX_ROWS entries of u_y + fluid flags: 576 KB DP / 320 KB SP, held by every GPU that
 * loads the library until exit).  Contexts with more than LBM_STATIC_MAX_ROWS local rows are refused with LBM_ERR_ARG.  Output
 * identical to lbm_reynolds_column's, bit for bit.  All contexts on one GPU SHARE the array, and it keeps the previous call's
 * contents (every entry that is read back is written first, nothing needs clearing): calls that overlap on the same GPU -- two
 * host threads, two contexts -- race; serialise them.  Synchronises like lbm_reynolds_column. */
#define LBM_STATIC_MAX_ROWS 65536
int lbm_reynolds_column_static(lbm_ctx *ctx, real_t *uy, char *is_fluid);
real_t lbm_reynolds_from_column(const real_t *uy, const char *is_fluid, int nx, real_t tau);
/* rho, u_x, u_y of every site of buffer A in the order of PrintLattice (LB.c:621-629);
 * outputs are dense [nx_local][ny]; any may be NULL. */
int lbm_macroscopic(lbm_ctx *ctx, real_t *rho, real_t *ux, real_t *uy);
/* PrintLattice (LB.c:610-639): CSV of u_x^2+u_y^2, "%.10lf", ", " separated, one lattice
 * row per line.  `path` replaces the hard-wired data/<timeStep>.csv. */
int lbm_print_lattice(lbm_ctx *ctx, const char *path);
/* The same file for a lattice spread over slabs of ONE process (slabs[k] owns the k-th row range, as for
 * lbm_group_run): SAVELATTICE (LB.c:155-159) of a multi-GPU run. */
int lbm_print_lattice_group(lbm_ctx **slabs, int nslabs, const char *path);

/* 64-bit checksum of buffer `buf` over this slab's elements: the sum (mod 2^64) of
 * mix64(bits(f[s][i][j]) ^ mix64((s*NX + i)*NY + j)) with GLOBAL row i -- the checksums of the slabs of a ring add up to
 * the single-GPU value, whatever the partition.  (No reference counterpart: a size-independent parity property for the
 * lattices too large to download and compare.) */
int lbm_checksum(lbm_ctx *ctx, int buf, unsigned long long *sum);
/* Stand-in for the reference's floating-point traps (feenableexcept(FE_ALL_EXCEPT & ~FE_INEXACT), LB.c:129: the
 * reference aborts on NaN, overflow and division by zero): counts the non-finite populations of buffer A;
 * LBM_ERR_FPE if there are any.  *nonfinite may be NULL.  Synchronous; one pass over the lattice. */
int lbm_check_finite(lbm_ctx *ctx, unsigned long long *nonfinite);

/* ---- slab decomposition / halo exchange (README.md:44-53; no reference source) ---- */
typedef struct {
	int nx_local, ny, row_lo, row_hi;
	size_t ldy;            /* device row stride in elements */
	size_t plane;          /* device plane stride in elements = (nx_local+2)*ldy */
	size_t halo_bytes;     /* bytes of one packed halo message: 3*ny*sizeof(real_t) */
	void *dev_f[2];        /* device base of buffers A, B (row 0 = lower halo row) */
	void *dev_flags;       /* device int[4]: neighbour-barrier slots */
} lbm_slab_info;
int lbm_slab_info_get(lbm_ctx *ctx, lbm_slab_info *info);

/* 64-byte CUDA IPC handles for fA, fB and the flag words, to be shipped to the ring
 * neighbours by the caller (torch.distributed / MPI / a pipe). */
#define LBM_IPC_BLOB_BYTES 256
int lbm_ipc_export(lbm_ctx *ctx, unsigned char blob[LBM_IPC_BLOB_BYTES]);
/* Connect the ring: `up` owns rows starting at row_hi, `dn` owns rows ending at row_lo
 * (both wrap periodically, LB.c:231-232).  After this call the boundary-row kernels store
 * their halo speeds (2,5,6 upwards; 4,7,8 downwards; LB.c:237-243) straight into the
 * neighbour's halo rows over NVLink, and steps are ordered by device-side flags. */
int lbm_ipc_connect(lbm_ctx *ctx, const unsigned char up[LBM_IPC_BLOB_BYTES],
                    const unsigned char dn[LBM_IPC_BLOB_BYTES]);
/* Same wiring for slabs living in ONE process (several GPUs with peer access, or several
 * logical slabs on one GPU for decomposition-invariance tests).  Ordering between slabs is
 * then the caller's: use lbm_group_run. */
int lbm_connect_local(lbm_ctx *ctx, lbm_ctx *up, lbm_ctx *dn);
/* Steps a ring of in-process slabs together (slabs[k] owns the k-th row range). */
int lbm_group_run(lbm_ctx **slabs, int nslabs, int nsteps);

/* Staged exchange for a transport the caller owns (NCCL send/recv through
 * torch.distributed): pack this slab's boundary rows of `buf` into two contiguous device
 * messages, and unpack two received messages into the halo rows of `buf`. */
int lbm_halo_pack(lbm_ctx *ctx, int buf, void *dev_to_up, void *dev_to_dn);
int lbm_halo_unpack(lbm_ctx *ctx, int buf, const void *dev_from_dn, const void *dev_from_up);
/* One lattice step src_buf -> other restricted to local rows [r0,r1), used by a caller that
 * overlaps its own exchange with inner-row compute.  stream_id 0 = main, 1 = boundary. */
int lbm_step_rows(lbm_ctx *ctx, int src_buf, int r0, int r1, int tail_source, int stream_id);
int lbm_stream_handle(lbm_ctx *ctx, int stream_id, void **cuda_stream);
int lbm_stream_join(lbm_ctx *ctx);   /* main stream waits for the boundary stream and vice versa */

/* ---- tuning / accounting ---- */
/* What round 1 measured and dropped (persistent kernel, dataflow kernel, fused step pairs, warp-shuffle neighbours, software
 * pipelining, capped grids: all slower than the defaults, numbers in DESIGN.md / profiles/) is no longer here; the
 * shared-memory staged kernels (cp.async / TMA) and the CTA size knob (128 threads measured best at every size, profiles/r01_run2)
 * live in the experiment build only (include/lbm_b200_experimental.h). */
typedef struct {
	int vec;          /* sites per thread: 1, 2 or 4; 0 = by lattice size (DP: 1 up to 16 Mi sites and in place, else 4; SP: 4) */
	int hints;        /* bit 0: A/B kernel: L1::no_allocate loads + streaming stores; in-place kernels under temporal blocking: the
	                     LAST level of a macro-step stores with the L2 evict-first policy (its results are not read again
	                     while the tile is L2-resident).  bit 1 (in-place contexts): the other levels load and store with the
	                     evict-last policy.  Never changes results.  Default 3 for DP in-place lattices above 16 Mi sites (half
	                     the DRAM traffic, +1.8 % in long runs), 0 elsewhere (SP and the small lattices lose 2-6 % with it) */
	int pdl;          /* 1: step kernels are launched with programmatic dependent launch, so the next
	                     step's CTAs are resident (waiting in griddepcontrol.wait) while this one drains */
	int graph;        /* >0: steps are replayed from a captured CUDA graph of `graph` step pairs (single-slab contexts) */
	int tb_depth;     /* in-place contexts only.  >= 2 (even): temporal blocking in L2 -- the lattice
	                     is swept in bands of tb_rows rows, and each band is advanced tb_depth timesteps (one
	                     kernel launch per band and timestep, band shifted down one row per timestep) while it
	                     is L2-resident, so HBM sees each site once per tb_depth steps instead of every step.
	                     Possible because the in-place kernels have no write-after-read hazards: a row may take
	                     step t as soon as rows i-1, i, i+1 have taken step t-1.  0 = off (one launch per step). */
	int tb_rows;      /* band height H in rows; 0 = chosen so that a tile of (H + tb_depth) rows takes its share of L2: one chain
	                     45 MB (DP) / 60 MB (SP) in full-width bands, 60 MB in 2-D tiles; tb_chains = C: 60 MB (DP) / 54 MB (SP) / C */
	int tb_cols;      /* column-strip width W of the 2-D schedule (multiple of 8): wide lattices are cut into strips as well
	                     as bands, the tile (H + k) x (W + 8k) sites is what stays in L2.  0 = automatic (strips of 16 KB
	                     row segments, 32 KB on slabs, only where full-width bands would be lower than 64 rows), >= ny = full-width rows */
	int tb_chains;    /* concurrent tile chains of the temporal-blocking schedule (1..4; 0 = the default, 2): the slab's rows are cut into this many
	                     lanes that run their tiles on separate streams, each with its share of the L2 tile budget, the boundaries
	                     between lanes handled like slab seams.  A second chain fills the drain every tile level pays. */
} lbm_tuning;
int lbm_set_tuning(lbm_ctx *ctx, const lbm_tuning *t);
int lbm_get_tuning(lbm_ctx *ctx, lbm_tuning *t);
/* Kernels launched by this context since creation (bench.py's gpu_launches). */
long long lbm_kernel_launches(lbm_ctx *ctx);
/* The launch list of ONE temporal-blocking macro-step for a slab of nx rows x ny columns (host logic only, no GPU
 * needed): launches[6n..6n+5] = (level, r0, r1, j0, j1, seam): rows [r0,r1) and columns [j0,j1), either range may
 * run past nx / ny and then wraps (rows: into the ring neighbours when multi != 0); j0 == j1 == 0 means whole rows.
 * seam (slabs of a multi-slab ring, multi != 0): bit 0 = wait for both neighbours' previous event before this
 * launch, bit 1 = announce this slab's event after it.  *depth_used = the depth used (single slab: clamped to the
 * ring, 0 = lattice too small, plain stepping).  Returns the number of launches (which may exceed cap; only cap
 * entries are written), -1 on bad arguments. */
int lbm_tb_schedule(int nx, int ny, int multi, int tb_depth, int tb_rows, int tb_cols, int *depth_used, int *launches, int cap);
/* The same for lbm_tuning.tb_chains > 1 (concurrent tile chains): 7 ints per launch, (level, r0, r1, j0, j1, stream, outer), in
 * enqueue order, which is also a valid sequential order.  stream 0..C-1 = the lanes' own tiles, C..2C-1 = the seam chain of the
 * boundary after lane (stream - C), whose launches of one column piece start after lanes b and b+1 have finished their tiles of
 * that piece; outer = 1 marks the boundary between the last lane and the first (the periodic wrap / the slab seam).
 * *chains_used < 2: the slab is too low for lanes, lbm_tb_schedule describes what runs.  Host logic only. */
int lbm_tb_schedule_chains(int nx, int ny, int multi, int chains, int tb_depth, int tb_rows, int tb_cols, int *depth_used, int *chains_used,
                           int *launches, int cap);

#ifdef __cplusplus
}
#endif
#endif /* LBM_B200_H */
