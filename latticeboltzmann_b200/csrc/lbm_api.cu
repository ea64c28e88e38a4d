// lbm_api.cu -- the C ABI declared in include/lbm_b200.h (host side, thin).
//
// Compiled twice: -DLBM_DOUBLE -> liblbm_b200_dp.so, -DLBM_FLOAT -> liblbm_b200_sp.so
// (the reference chooses precision the same way, by header, LB.c:59-65).
// No CPU fallback anywhere: without a usable sm_100 device every call fails.
#if defined(LBM_DOUBLE)
#include "prec_double_cuda.h"
#elif defined(LBM_FLOAT)
#include "prec_float_cuda.h"
#else
#error "compile with -DLBM_DOUBLE or -DLBM_FLOAT"
#endif
#include "lbm_b200.h"
#include "lbm_kernels.cuh"
#ifdef LBM_EXPERIMENTAL     /* the experiment build (make experimental): shared-memory staged step kernels behind lbm_set_experimental */
#include "lbm_b200_experimental.h"
#include "lbm_kernels_experimental.cuh"
#endif

#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

using namespace lbm;

namespace {

thread_local char g_err[512] = "";

int fail(int code, const char *fmt, ...)
{
	va_list ap;
	va_start(ap, fmt);
	vsnprintf(g_err, sizeof g_err, fmt, ap);
	va_end(ap);
	return code;
}

#define CK(call)                                                                                   \
	do {                                                                                           \
		cudaError_t e_ = (call);                                                                   \
		if (e_ != cudaSuccess)                                                                     \
			return fail(LBM_ERR_CUDA, "%s:%d %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_)); \
	} while (0)

enum { SYNC_SELF = 0, SYNC_LOCAL = 1, SYNC_FLAGS = 2 };


struct Peer {
	real_t *f[2];      // neighbour's buffers A, B (storage row 0)
	int *flags;        // neighbour's flag words
	int nxl;           // neighbour's owned rows
};

struct IpcBlob {           // what lbm_ipc_export ships (<= LBM_IPC_BLOB_BYTES)
	char magic[8];
	cudaIpcMemHandle_t f[2];
	cudaIpcMemHandle_t flags;
	int nxl, ny, device, pid;
	unsigned long long uid;
	int inplace;
};
static_assert(sizeof(IpcBlob) <= LBM_IPC_BLOB_BYTES, "blob too large");

} // namespace

enum { LBM_MAX_CHAINS = 4 };

struct lbm_ctx {
	lbm_config cfg;
	int nxl, ny, vw, nyvec, ldy, ldw;
	int inplace;                 // one buffer, AA-pattern steps (lbm_config.inplace)
	size_t plane;
	real_t *f[2];                // f[1] == nullptr when inplace
	uint32_t *mask;
	int *flags;                  // device int[4]: [0] from dn, [1] from up, [3] timeout
	cudaStream_t stream[2];      // 0 main (inner rows), 1 boundary rows
	cudaEvent_t ev_main, ev_bnd[2], ev_t0, ev_t1;
	Peer up, dn;
	int sync_mode;
	int epoch;
	unsigned long long wait_timeout_ns;   /* neighbour barrier: give up (and raise the ring's abort word) after this long */
	void *ipc_opened[6];
	int n_ipc_opened;
	lbm_tuning tune;
	long long launches;
	int has_state;
	real_t w0, w14, w58, itau, a14, a58;
	double itau_d;
	unsigned *seam_done;         // device counter of the fused seam barrier (CTAs of the running seam launch that have finished)
	uint32_t *slip;              // device [2][nxl][ldw]: "reflect" walls with normal x / normal y (mask values 2 / 3)
	int has_slip;                // any bit set in `slip` (else the kernels are told there is none)
	int outflow;                 // lbm_config.boundary_y
	cudaGraphExec_t graph_exec;  // captured `graph_pairs` x (A->B, B->A) steps with the current tuning
	int graph_pairs;
	cudaGraphExec_t wf_exec;     // captured wavefront macro-step (wf_k timesteps, bands of wf_h rows)
	long long wf_launches;       // kernels in that graph
	int in_wavefront;            // set while run_wavefront enqueues (step kernels may then use PDL on slabs too)
	/* concurrent tile chains (lbm_tuning.tb_chains): lane l > 0 runs on lane_stream[l] (lane 0 on stream[0]); the seam chain of
	 * the boundary between lanes b and b+1 on chain_stream[b] (the outer boundary -- the periodic wrap / the slab seam -- on stream[1]) */
	cudaStream_t lane_stream[LBM_MAX_CHAINS], chain_stream[LBM_MAX_CHAINS];
	cudaEvent_t ev_lane[LBM_MAX_CHAINS], ev_chain[LBM_MAX_CHAINS], ev_fork;
	int lanes_ready;
#ifdef LBM_EXPERIMENTAL
	lbm_experimental exp;
#endif
};

namespace {

int default_vec(const lbm_ctx *c)
{
	if (sizeof(real_t) == 4) return 4;
	return (unsigned long long)c->nxl * c->ny <= (16ull << 20) ? 1 : 4;
}

int default_small_dp(const lbm_ctx *c)
{
	return sizeof(real_t) == 8 && (unsigned long long)c->nxl * c->ny <= (16ull << 20);
}

void drop_graph(lbm_ctx *c)
{
	if (c->graph_exec) cudaGraphExecDestroy(c->graph_exec);
	c->graph_exec = nullptr;
	c->graph_pairs = 0;
	if (c->wf_exec) cudaGraphExecDestroy(c->wf_exec);
	c->wf_exec = nullptr;
}

int set_device(const lbm_ctx *c) { CK(cudaSetDevice(c->cfg.device)); return LBM_OK; }

StepParams<real_t> make_params(const lbm_ctx *c, int src, int r0, int r1, int tail_source)
{
	StepParams<real_t> p;
	/* in place: `src` only selects the phase (0 -> phase 1, 1 -> phase 2); one buffer everywhere */
	const int dst = c->inplace ? 0 : src ^ 1;
	if (c->inplace) src = 0;
	p.src = c->f[src];
	p.dst = c->f[dst];
	p.mask = c->mask;
	p.plane = c->plane;
	p.up_plane = (size_t)(c->up.nxl + 2) * c->ldy;
	p.dn_plane = (size_t)(c->dn.nxl + 2) * c->ldy;
	p.up_halo = c->up.f[dst];                                          // neighbour's storage row 0
	p.dn_halo = c->dn.f[dst] + (size_t)(c->dn.nxl + 1) * c->ldy;       // neighbour's storage row nxl'+1
	p.ldy = c->ldy; p.ldw = c->ldw;
	p.nxl = c->nxl; p.ny = c->ny;
	p.jmin = 0; p.jmax = c->nyvec;
	p.r0 = r0; p.r1 = r1;
	p.vw = c->vw;
	p.tail_source = tail_source;
	p.outflow = c->outflow;
	p.slip = c->has_slip ? c->slip : nullptr;
	p.seam_flags = nullptr; p.seam_up = p.seam_dn = nullptr; p.seam_done = nullptr;
	p.seam_wait = p.seam_signal = 0; p.seam_timeout_ns = 0;
	p.w0 = c->w0; p.w14 = c->w14; p.w58 = c->w58;
	p.itau = c->itau; p.itau_d = c->itau_d;
	p.a14 = c->a14; p.a58 = c->a58;
	return p;
}

} // namespace

extern "C" {

int lbm_sizeof_real(void) { return (int)sizeof(real_t); }
const char *lbm_last_error(void) { return g_err; }
const char *lbm_version(void) { return "latticeboltzmann_b200 0.1 (" LBM_PREC_TAG ", sm_100a)"; }

lbm_scene lbm_scene_reference(void)
{
	lbm_scene s;
	s.rho0 = REAL_T(0.1);                 /* LB.c:56 */
	s.channel_walls = 1;                  /* LB.c:575-578 */
	s.bar_i0 = 20; s.bar_i1 = 220;        /* LB.c:567 */
	s.bar_j0 = 100; s.bar_j1 = 105;       /* LB.c:568-572 */
	s.disc_ci = s.disc_cj = 0; s.disc_r = -1;
	s.perturb_seed = 0; s.perturb_amp = REAL_T(0.0);
	return s;
}

int lbm_create(lbm_ctx **out, const lbm_config *cfg)
{
	if (!out || !cfg) return fail(LBM_ERR_ARG, "lbm_create: null argument");
	*out = nullptr;
	if (cfg->nx < 1 || cfg->ny < 2) return fail(LBM_ERR_ARG, "lbm_create: lattice %dx%d too small", cfg->nx, cfg->ny);
	int vw = cfg->vecwidth ? cfg->vecwidth : VECWIDTH;
	if (vw != 1 && vw != 2 && vw != 4 && vw != 8) return fail(LBM_ERR_ARG, "lbm_create: vecwidth %d not in {1,2,4,8}", vw);
	if (cfg->inplace != 0 && cfg->inplace != 1) return fail(LBM_ERR_ARG, "lbm_create: inplace %d not in {0,1}", cfg->inplace);
	if (cfg->boundary_y != LBM_Y_PERIODIC && cfg->boundary_y != LBM_Y_OUTFLOW) return fail(LBM_ERR_ARG, "lbm_create: boundary_y %d not in {0,1}", cfg->boundary_y);
	if (cfg->boundary_y == LBM_Y_OUTFLOW && cfg->inplace && (vw == 1 || cfg->ny % vw != 0))
		return fail(LBM_ERR_ARG, "lbm_create: open y ends with in-place streaming need the vector kernels everywhere (vecwidth > 1 and ny a "
		                         "multiple of it): the end column's site shares one slot with its in-row neighbour; use inplace = 0");
	int lo = cfg->row_lo, hi = cfg->row_hi;
	if (lo == 0 && hi == 0) hi = cfg->nx;
	if (lo < 0 || hi > cfg->nx || lo >= hi) return fail(LBM_ERR_ARG, "lbm_create: bad slab rows [%d,%d) of %d", lo, hi, cfg->nx);

	int ndev = 0;
	CK(cudaGetDeviceCount(&ndev));
	if (cfg->device < 0 || cfg->device >= ndev) return fail(LBM_ERR_CUDA, "lbm_create: device %d of %d", cfg->device, ndev);
	CK(cudaSetDevice(cfg->device));
	cudaDeviceProp prop;
	CK(cudaGetDeviceProperties(&prop, cfg->device));
	if (prop.major != 10) return fail(LBM_ERR_CUDA, "lbm_create: device is sm_%d%d, this library carries sm_100a code only", prop.major, prop.minor);

	lbm_ctx *c = new lbm_ctx();
	memset(c, 0, sizeof *c);
	c->cfg = *cfg;
	c->cfg.row_lo = lo; c->cfg.row_hi = hi; c->cfg.vecwidth = vw;
	c->nxl = hi - lo;
	c->ny = cfg->ny;
	c->vw = vw;
	c->nyvec = vw * (cfg->ny / vw);                           /* NYVECMAX, LB.c:76 */
	const int per128 = 128 / (int)sizeof(real_t);
	c->ldy = ((cfg->ny + per128 - 1) / per128) * per128;      /* rows padded to 128 B */
	c->ldw = (((cfg->ny + 31) / 32) + 3) / 4 * 4;
	c->plane = (size_t)(c->nxl + 2) * c->ldy;
	if ((unsigned long long)c->nxl * (unsigned long long)c->ny >= 0xffffffffull) {
		delete c;
		return fail(LBM_ERR_ARG, "lbm_create: slab has >= 2^32 sites");
	}

	/* constants formed in real_t exactly as the reference's literals are (LB.c:38-40,323,502) */
	c->w0 = REAL_T(4.0) / REAL_T(9.0);
	c->w14 = REAL_T(1.0) / REAL_T(9.0);
	c->w58 = REAL_T(1.0) / REAL_T(36.0);
	c->itau = REAL_T(1.0) / cfg->tau;
	c->itau_d = 1.0 / (double)cfg->tau;
	c->a14 = cfg->accel * c->w14;
	c->a58 = cfg->accel * c->w58;

	const size_t fbytes = 9 * c->plane * sizeof(real_t);
	c->inplace = cfg->inplace ? 1 : 0;
	c->outflow = cfg->boundary_y == LBM_Y_OUTFLOW;
	cudaError_t e = cudaMalloc(&c->f[0], fbytes);
	if (e == cudaSuccess && !c->inplace) e = cudaMalloc(&c->f[1], fbytes);
	if (e == cudaSuccess) e = cudaMalloc(&c->mask, (size_t)c->nxl * c->ldw * sizeof(uint32_t));
	if (e == cudaSuccess) e = cudaMalloc(&c->flags, 4 * sizeof(int));
	if (e == cudaSuccess) e = cudaMalloc(&c->seam_done, sizeof(unsigned));
	if (e == cudaSuccess) e = cudaMalloc(&c->slip, (size_t)2 * c->nxl * c->ldw * sizeof(uint32_t));
	for (int k = 0; k < 2 && e == cudaSuccess; k++) e = cudaStreamCreateWithFlags(&c->stream[k], cudaStreamNonBlocking);
	/* Zero everything ON THE CONTEXT'S OWN STREAM and wait: the streams are non-blocking, so a cudaMemset on the legacy
	 * default stream is not ordered with them -- on a 19 GB lattice it was still running when lbm_initialize_arrays'
	 * kernels had finished, and wiped the wall mask they wrote (found by the cross-engine checksum at 8192x32768). */
	if (e == cudaSuccess) e = cudaMemsetAsync(c->f[0], 0, fbytes, c->stream[0]);
	if (e == cudaSuccess && c->f[1]) e = cudaMemsetAsync(c->f[1], 0, fbytes, c->stream[0]);
	if (e == cudaSuccess) e = cudaMemsetAsync(c->mask, 0, (size_t)c->nxl * c->ldw * sizeof(uint32_t), c->stream[0]);
	if (e == cudaSuccess) e = cudaMemsetAsync(c->slip, 0, (size_t)2 * c->nxl * c->ldw * sizeof(uint32_t), c->stream[0]);
	if (e == cudaSuccess) e = cudaMemsetAsync(c->flags, 0, 4 * sizeof(int), c->stream[0]);
	if (e == cudaSuccess) e = cudaMemsetAsync(c->seam_done, 0, sizeof(unsigned), c->stream[0]);
	if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream[0]);
	if (e == cudaSuccess) e = cudaEventCreateWithFlags(&c->ev_main, cudaEventDisableTiming);
	for (int k = 0; k < 2 && e == cudaSuccess; k++) e = cudaEventCreateWithFlags(&c->ev_bnd[k], cudaEventDisableTiming);
	if (e == cudaSuccess) e = cudaEventCreate(&c->ev_t0);
	if (e == cudaSuccess) e = cudaEventCreate(&c->ev_t1);
	if (e != cudaSuccess) {
		fail(LBM_ERR_CUDA, "lbm_create: %s (%.2f GB requested)", cudaGetErrorString(e), (c->inplace ? 1.0 : 2.0) * fbytes / 1e9);
		lbm_destroy(c);
		return LBM_ERR_CUDA;
	}
	/* P == 1: the slab is its own ring neighbour (x wraps periodically, LB.c:231-232) */
	c->up.f[0] = c->dn.f[0] = c->f[0];
	c->up.f[1] = c->dn.f[1] = c->f[1];
	c->up.flags = c->dn.flags = c->flags;
	c->up.nxl = c->dn.nxl = c->nxl;
	c->sync_mode = SYNC_SELF;
	c->wait_timeout_ns = 20ull * 1000 * 1000 * 1000;
	if (const char *ms = getenv("LBM_WAIT_TIMEOUT_MS"))         /* tests of the ring's abort path */
		if (atoll(ms) > 0) c->wait_timeout_ns = (unsigned long long)atoll(ms) * 1000ull * 1000ull;

	/* measured best on B200 at every size (profiles/r01_run2): direct neighbour loads, 128-thread CTAs */
	c->tune.vec = default_vec(c);
	c->tune.hints = 0;
	/* small DP lattices are launch-bound (a step is ~35 us): PDL + graph replay of 10 step pairs measured
	 * 43.7 vs 41.1 GLUPS at 400x4000 (profiles/r01_run3); no gain for SP or for large lattices */
	c->tune.pdl = default_small_dp(c);
	c->tune.graph = default_small_dp(c) ? 10 : 0;
	c->tune.tb_chains = 1;
	if (c->inplace && (unsigned long long)c->nxl * c->ny > (16ull << 20)) {
		/* wide / large in-place lattices (profiles/r01_run14): 2-D tiles of ~60 MB, plain PDL launches (replaying
		 * thousands of kernel nodes from a graph measured 5-10 % slower).  DP 8192x32768: 62.1 vs 47.8 GLUPS. */
		c->tune.pdl = 1;
		c->tune.graph = 0;
		c->tune.tb_depth = 16;                                   /* SP since its kernels run at 64 registers: 104.4 (k 16) vs 102.1 (k 8), r01_run28 */
		c->tune.vec = sizeof(real_t) == 8 ? 1 : 4;
		c->tune.tb_chains = 2;                                   /* two concurrent tile chains: DP 57.4 -> 65.6 GLUPS, SP 102.2 -> 119.5 (profiles/r02_run6) */
		/* L2 eviction priorities (last level evict-first, the others evict-last): DP +1.8 % in 1000-timestep runs under the power
		 * cap (profiles/r02_run11 long_hints.log), +0.1-0.8 % in short ones; SP -1.6 to -2.6 %, and DP -6 % on the small
		 * lattices below (profiles/r02_run7): on for DP here only */
		if (sizeof(real_t) == 8) c->tune.hints = 3;
	}
	if (c->inplace && (unsigned long long)c->nxl * c->ny <= (16ull << 20)) {
		/* in-place small lattices (profiles/r01_run8, r01_run9): PDL + graph replay for both precisions; lattices
		 * beyond what the L2 keeps (~60 MB) whose rows are narrow enough for >= 64-row bands are swept with
		 * temporal blocking (DP 400x4000: 59.3 vs 43.8 GLUPS, DP 800x4000 57.0 vs 44.7, SP 800x4000 90.0 vs 83.5) */
		c->tune.pdl = 1;
		c->tune.graph = 10;
		const double row_bytes = 9.0 * c->ldy * sizeof(real_t);
		const int k = 16;
		const double band = sizeof(real_t) == 8 ? 45e6 : 60e6;
		/* SP: vec 4 and k = 16 since the SP kernels run at 64 registers -- 800x4000 105.7 GLUPS against 98.7 for the
		 * earlier vec 2 / k = 8, 1600x4000 109.9 against 101.9 (profiles/r01_run28) */
		if (row_bytes * c->nxl > 64e6 && band / row_bytes - k >= 64 && c->nxl >= 2 * k + 2) {
			c->tune.tb_depth = k;
			c->tune.tb_chains = 2;                               /* DP 400x4000 54.1 -> 65.1 GLUPS, 800x4000 55.6 -> 67.1 (profiles/r02_run7) */
		}
	}
	*out = c;
	return LBM_OK;
}

int lbm_destroy(lbm_ctx *c)
{
	if (!c) return LBM_OK;
	cudaSetDevice(c->cfg.device);
	cudaDeviceSynchronize();
	for (int k = 0; k < c->n_ipc_opened; k++) cudaIpcCloseMemHandle(c->ipc_opened[k]);
	for (int k = 0; k < 2; k++) {
		if (c->f[k]) cudaFree(c->f[k]);
		if (c->stream[k]) cudaStreamDestroy(c->stream[k]);
		if (c->ev_bnd[k]) cudaEventDestroy(c->ev_bnd[k]);
	}
	if (c->mask) cudaFree(c->mask);
	drop_graph(c);
	if (c->flags) cudaFree(c->flags);
	if (c->slip) cudaFree(c->slip);
	if (c->seam_done) cudaFree(c->seam_done);
	for (int k = 0; k < LBM_MAX_CHAINS; k++) {
		if (c->lane_stream[k]) cudaStreamDestroy(c->lane_stream[k]);
		if (c->chain_stream[k]) cudaStreamDestroy(c->chain_stream[k]);
		if (c->ev_lane[k]) cudaEventDestroy(c->ev_lane[k]);
		if (c->ev_chain[k]) cudaEventDestroy(c->ev_chain[k]);
	}
	if (c->ev_fork) cudaEventDestroy(c->ev_fork);
	if (c->ev_main) cudaEventDestroy(c->ev_main);
	if (c->ev_t0) cudaEventDestroy(c->ev_t0);
	if (c->ev_t1) cudaEventDestroy(c->ev_t1);
	delete c;
	return LBM_OK;
}

} // extern "C"

// ---------------------------------------------------------------------------
// step launch
// ---------------------------------------------------------------------------
namespace {

typedef void (*step_kernel_t)(const StepParams<real_t>);

cudaError_t launch_kernel(step_kernel_t kern, const StepParams<real_t> &p, dim3 grid, unsigned block, cudaStream_t st, int pdl, size_t smem = 0)
{
	cudaLaunchConfig_t cfg = {};
	cfg.gridDim = grid; cfg.blockDim = dim3(block); cfg.dynamicSmemBytes = smem; cfg.stream = st;
	cudaLaunchAttribute at[1];
	at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
	at[0].val.programmaticStreamSerializationAllowed = 1;
	cfg.attrs = at; cfg.numAttrs = pdl ? 1 : 0;
	return cudaLaunchKernelEx(&cfg, kern, p);
}

/* in-place (AA) step kernels: phase 1 (from the reference layout) / phase 2 (back to it); hint 1 = streaming stores */
template <int V>
cudaError_t launch_aa_v(int phase, int hint, const StepParams<real_t> &p, dim3 grid, unsigned block, cudaStream_t st, int pdl, size_t smem = 0)
{
	if (p.seam_flags)                                            /* a slab's seam launch: carries the neighbour barrier, default eviction priority */
		return phase == 1 ? launch_kernel(stream_collide_aa_kernel<real_t, V, 1, 0, true>, p, grid, block, st, pdl, smem)
		                  : launch_kernel(stream_collide_aa_kernel<real_t, V, 2, 0, true>, p, grid, block, st, pdl, smem);
	if (hint == 1)
		return phase == 1 ? launch_kernel(stream_collide_aa_kernel<real_t, V, 1, 1>, p, grid, block, st, pdl, smem)
		                  : launch_kernel(stream_collide_aa_kernel<real_t, V, 2, 1>, p, grid, block, st, pdl, smem);
	if (hint == 2)
		return phase == 1 ? launch_kernel(stream_collide_aa_kernel<real_t, V, 1, 2>, p, grid, block, st, pdl, smem)
		                  : launch_kernel(stream_collide_aa_kernel<real_t, V, 2, 2>, p, grid, block, st, pdl, smem);
	return phase == 1 ? launch_kernel(stream_collide_aa_kernel<real_t, V, 1, 0>, p, grid, block, st, pdl, smem)
	                  : launch_kernel(stream_collide_aa_kernel<real_t, V, 2, 0>, p, grid, block, st, pdl, smem);
}

template <int V>
cudaError_t launch_vec_v(int hint, const StepParams<real_t> &p, unsigned grid, unsigned block, cudaStream_t st, int pdl)
{
	return hint ? launch_kernel(stream_collide_vec_kernel<real_t, V, 1>, p, dim3(grid), block, st, pdl)
	            : launch_kernel(stream_collide_vec_kernel<real_t, V, 0>, p, dim3(grid), block, st, pdl);
}

int sm_count(int device)
{
	static int cached[64] = {0};
	if (device < 64 && cached[device]) return cached[device];
	int n = 148;
	cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, device);
	if (device < 64) cached[device] = n;
	return n;
}

#ifdef LBM_EXPERIMENTAL
/* experiment build only: the in-place step with the next row's populations staged in shared memory (cp.async, or TMA bulk
 * copies for whole-row launches of phase 2); returns cudaErrorNotYetImplemented where it does not apply */
cudaError_t launch_staged(lbm_ctx *c, int phase, const StepParams<real_t> &p, unsigned rows, unsigned ncols, bool whole_rows, cudaStream_t st, int pdl)
{
	if (!c->exp.stage_smem || c->exp.rows_per_thread < 2 || c->nyvec != c->ny || c->outflow) return cudaErrorNotYetImplemented;
	unsigned blk = c->exp.block ? (unsigned)c->exp.block : 128u;
	if (blk > 256u) blk = 256u;
	unsigned gy = (rows + (unsigned)c->exp.rows_per_thread - 1) / (unsigned)c->exp.rows_per_thread;
	if (gy > 65535u) gy = 65535u;
	if (phase == 2 && c->exp.stage_smem == 2 && whole_rows && ((size_t)c->ny * sizeof(real_t)) % 16 == 0) {   /* TMA both ways; whole rows of 16-byte units */
		const unsigned b2 = blk > 128 ? 128 : blk;
		return launch_kernel(stream_collide_aa_bulk2_kernel<real_t>, p, dim3(((unsigned)c->ny + b2 - 1) / b2, gy), b2, st, pdl, (size_t)4 * 9 * b2 * sizeof(real_t) + 16);
	}
	const dim3 g3((ncols + blk - 1) / blk, gy);
	const size_t smem = (size_t)2 * 9 * blk * sizeof(real_t);
	return phase == 1 ? launch_kernel(stream_collide_aa_staged_kernel<real_t, 1>, p, g3, blk, st, pdl, smem)
	                  : launch_kernel(stream_collide_aa_staged_kernel<real_t, 2>, p, g3, blk, st, pdl, smem);
}
#endif

// One pull step src -> src^1 over owned rows [r0,r1): vector kernel for j < NYVECMAX,
// scalar-order kernel for the ragged tail (LB.c:195-196) or for everything (LB.c:198).
// jc0 < jc1: restrict the launch to columns [jc0, jc1) (multiples of 8, jc1 may exceed ny and wraps) -- only the
// in-place vector kernel takes a column range (2-D temporal blocking; needs nyvec == ny)
// stream_out: in-place kernels store with the evict-first policy (lbm_tuning.hints on the last level of a macro-step)
/* L2 eviction priority of level s of a macro-step of kk levels (lbm_tuning.hints: bit 0 = the last level stores evict-first,
 * bit 1 = the other levels load and store evict-last) */
int tb_hint(const lbm_ctx *c, int s, int kk)
{
	if (s == kk - 1) return (c->tune.hints & 1) ? 1 : 0;
	return (c->tune.hints & 2) ? 2 : 0;
}

/* the in-place vector kernel can carry a slab's neighbour barrier itself (wait at its start, announce at its end) when it is
 * the only kernel of the step */
bool seam_fusable(const lbm_ctx *c)
{
	bool ok = c->sync_mode == SYNC_FLAGS && c->inplace && c->vw > 1 && c->nyvec == c->ny;
#ifdef LBM_EXPERIMENTAL
	ok = ok && !c->exp.stage_smem;
#endif
	return ok;
}

int launch_step(lbm_ctx *c, int src, int r0, int r1, int tail_source, cudaStream_t st, int jc0 = 0, int jc1 = 0, int stream_out = 0, int seam = 0)
{
	if (r1 <= r0) return LBM_OK;
	StepParams<real_t> p = make_params(c, src, r0, r1, tail_source);
	if (seam) {                                                  /* wait for epoch, then announce epoch + 1 (launch_wait / launch_signal in one) */
		if (!seam_fusable(c)) return fail(LBM_ERR_STATE, "launch_step: this step cannot carry the neighbour barrier");
		p.seam_flags = c->flags; p.seam_wait = c->epoch;
		p.seam_up = c->up.flags + 0; p.seam_dn = c->dn.flags + 1; p.seam_signal = ++c->epoch;
		p.seam_done = c->seam_done; p.seam_timeout_ns = c->wait_timeout_ns;
	}
	const unsigned rows = (unsigned)(r1 - r0);
	int jtail = 0;
	unsigned ncols = (unsigned)c->nyvec;
	if (jc1 > jc0) {
		if (!c->inplace || c->nyvec != c->ny || c->vw <= 1) return fail(LBM_ERR_STATE, "launch_step: column ranges need the in-place vector kernel");
		p.jmin = jc0; p.jmax = jc1;
		ncols = (unsigned)(jc1 - jc0);
	}
	if (c->vw > 1 && c->nyvec > 0) {
		int V = c->tune.vec;
		if (V > c->vw) V = c->vw;
		unsigned block = 128;                                    /* measured best at every size (profiles/r01_run2) */
#ifdef LBM_EXPERIMENTAL
		if (c->exp.block) block = (unsigned)c->exp.block;
#endif
		if (c->inplace && sizeof(real_t) == 4 && block > 256) block = 256;      /* launch bounds of the SP in-place kernels */
		/* PDL only where consecutive step kernels follow each other on one stream with nothing between */
		const int pdl = c->tune.pdl && (c->sync_mode == SYNC_SELF || c->in_wavefront) && c->nyvec == c->ny;
		cudaError_t e = cudaErrorNotYetImplemented;
		if (c->inplace) {
			const int phase = src == 0 ? 1 : 2;
#ifdef LBM_EXPERIMENTAL
			e = launch_staged(c, phase, p, rows, ncols, jc1 <= jc0, st, pdl);
#endif
			if (e == cudaErrorNotYetImplemented) {
				/* x: the vectors of one row; y: rows (grid-stride above the 65535 limit) */
				const dim3 g2((ncols / (unsigned)V + block - 1) / block, rows > 65535u ? 65535u : rows);
				size_t pad = 0;
#ifdef LBM_EXPERIMENTAL
				pad = (size_t)c->exp.smem_pad;
#endif
				switch (V) {
				case 1: e = launch_aa_v<1>(phase, stream_out, p, g2, block, st, pdl, pad); break;
				case 2: e = launch_aa_v<2>(phase, stream_out, p, g2, block, st, pdl, pad); break;
				case 4: e = launch_aa_v<4>(phase, stream_out, p, g2, block, st, pdl, pad); break;
				default: return fail(LBM_ERR_ARG, "launch_step: vec %d", V);
				}
			}
		} else {
			const unsigned long long total = (unsigned long long)rows * (unsigned)(c->nyvec / V);
			const unsigned grid = (unsigned)((total + block - 1) / block);
			switch (V) {
			case 1: e = launch_vec_v<1>(c->tune.hints & 1, p, grid, block, st, pdl); break;
			case 2: e = launch_vec_v<2>(c->tune.hints & 1, p, grid, block, st, pdl); break;
			case 4: e = launch_vec_v<4>(c->tune.hints & 1, p, grid, block, st, pdl); break;
			default: return fail(LBM_ERR_ARG, "launch_step: vec %d", V);
			}
		}
		if (e != cudaSuccess) return fail(LBM_ERR_CUDA, "stream_collide kernel: %s", cudaGetErrorString(e));
		c->launches++;
		jtail = c->nyvec;
	}
	if (jtail < c->ny) {
		p.jmin = jtail; p.jmax = c->ny;
		const unsigned long long total = (unsigned long long)rows * (unsigned)(c->ny - jtail);
		const unsigned block = 256;
		unsigned long long grid = (total + block - 1) / block;
		const unsigned long long cap = (unsigned long long)sm_count(c->cfg.device) * 16;
		if (grid > cap) grid = cap;
		if (!c->inplace) stream_collide_scalar_kernel<real_t, 0><<<(unsigned)grid, block, 0, st>>>(p);
		else if (src == 0) stream_collide_scalar_kernel<real_t, 1><<<(unsigned)grid, block, 0, st>>>(p);
		else stream_collide_scalar_kernel<real_t, 2><<<(unsigned)grid, block, 0, st>>>(p);
		cudaError_t e = cudaGetLastError();
		if (e != cudaSuccess) return fail(LBM_ERR_CUDA, "stream_collide_scalar_kernel: %s", cudaGetErrorString(e));
		c->launches++;
	}
	return LBM_OK;
}

int launch_source(lbm_ctx *c, int buf, cudaStream_t st)
{
	StepParams<real_t> p = make_params(c, buf, 0, c->nxl, 0);
	apply_source_kernel<real_t><<<(c->nxl + 127) / 128, 128, 0, st>>>(c->f[buf], p);
	CK(cudaGetLastError());
	c->launches++;
	return LBM_OK;
}

// refresh the neighbours' (or, P == 1, our own) halo rows of buffer `buf` from our boundary rows
int launch_halo_push(lbm_ctx *c, int buf, cudaStream_t st)
{
	StepParams<real_t> p = make_params(c, buf ^ 1, 0, c->nxl, 0);   // "dst" == buf
	halo_push_kernel<real_t><<<(c->ny + 255) / 256, 256, 0, st>>>(c->f[buf], p);
	CK(cudaGetLastError());
	c->launches++;
	return LBM_OK;
}

// neighbour barrier through device flags (ranks in other processes): announce, then wait
int launch_signal(lbm_ctx *c, cudaStream_t st)
{
	c->epoch++;
	flag_signal_kernel<<<1, 1, 0, st>>>(c->up.flags + 0, c->dn.flags + 1, c->epoch);
	CK(cudaGetLastError());
	c->launches++;
	return LBM_OK;
}
int launch_wait(lbm_ctx *c, cudaStream_t st)
{
	flag_wait_kernel<<<1, 1, 0, st>>>(c->flags, c->up.flags + 3, c->dn.flags + 3, c->epoch, c->wait_timeout_ns);
	CK(cudaGetLastError());
	c->launches++;
	return LBM_OK;
}

// One lattice step of this slab, with the neighbour ordering its sync mode needs.
//   SYNC_SELF : single launch over all rows (halo push lands in our own halo rows).
//   SYNC_FLAGS: boundary rows first on stream 1 (wait for the neighbours' previous step,
//               compute rows 0 and nxl-1 storing their halos into the neighbours over NVLink,
//               announce), inner rows meanwhile on stream 0, then join -- the reference's
//               "hide the exchange behind the inner rows" (README.md:45-51).
//   SYNC_LOCAL: like SYNC_FLAGS but ordered by the caller (lbm_group_run) with events.
int enqueue_step(lbm_ctx *c, int src, int tail_source)
{
	int rc;
	if (c->sync_mode == SYNC_SELF) return launch_step(c, src, 0, c->nxl, tail_source, c->stream[0]);
	if (c->sync_mode == SYNC_FLAGS) {
		CK(cudaEventRecord(c->ev_main, c->stream[0]));
		CK(cudaStreamWaitEvent(c->stream[1], c->ev_main, 0));
		if ((rc = launch_wait(c, c->stream[1]))) return rc;
		if ((rc = launch_step(c, src, 0, 1, tail_source, c->stream[1]))) return rc;
		if (c->nxl > 1 && (rc = launch_step(c, src, c->nxl - 1, c->nxl, tail_source, c->stream[1]))) return rc;
		if ((rc = launch_signal(c, c->stream[1]))) return rc;
		CK(cudaEventRecord(c->ev_bnd[0], c->stream[1]));
		if ((rc = launch_step(c, src, 1, c->nxl - 1, tail_source, c->stream[0]))) return rc;
		CK(cudaStreamWaitEvent(c->stream[0], c->ev_bnd[0], 0));
		return LBM_OK;
	}
	return fail(LBM_ERR_STATE, "slab is wired with lbm_connect_local: step it through lbm_group_run");
}

// ---------------------------------------------------------------------------
// Temporal blocking in L2 (in-place, single slab).  The in-place kernels only have read-after-write
// dependencies: row i may take timestep t as soon as rows i-1, i, i+1 (periodic) have taken t-1.
// One "macro-step" advances the whole ring k timesteps (levels s = 0..k-1) as
//   pyramid          rows [s, H0-s)                       H0 >= 2k
//   parallelograms   rows [a-s, b-s) for bands [a,b) of H rows tiling [H0, nx)     band after band
//   inverted pyramid rows [nx-s, nx+s) mod nx             closes the periodic ring
// Each (band, level) is one launch of the ordinary in-place kernel on a row range; stream order
// satisfies every dependency.  A band of H+k rows stays L2-resident while it takes its k steps, so
// HBM traffic per site update drops from 18 reals to ~18/k.  Results are bit-identical: the same
// kernels do the same arithmetic on the same values, only the order of independent work changes.
// ---------------------------------------------------------------------------
// One dimension of the schedule: a periodic axis of n cells cut into pieces whose extent at level s is
//   pyramid [a + g s, b - g s) | parallelogram [a - g s, b - g s) | inverted pyramid [n - g s, n + g s) | whole axis
// g = cells the piece boundaries move per level (>= the stencil reach 1; 8 along j keeps vectors aligned).
struct TbPiece { int type, a, b; };
enum { TB_PYRAMID = 0, TB_PARALLELOGRAM = 1, TB_INVERTED = 2, TB_WHOLE = 3 };
struct TbAxis { int n, g, e, npieces; TbPiece piece[1026]; };   /* e: cells at either end of the axis kept for the inverted pyramid at level 0 */
struct TbPlan { int k; TbAxis rows, cols; };

inline void tb_range(const TbAxis &ax, const TbPiece &pc, int s, int *lo, int *hi)
{
	switch (pc.type) {
	case TB_PYRAMID: *lo = pc.a + ax.g * s; *hi = pc.b - ax.g * s; break;
	case TB_PARALLELOGRAM: *lo = pc.a - ax.g * s; *hi = pc.b - ax.g * s; break;
	case TB_INVERTED: *lo = ax.n - ax.e - ax.g * s; *hi = ax.n + ax.e + ax.g * s; break;
	default: *lo = 0; *hi = ax.n; break;
	}
}

/* pieces of one axis: pyramid [e, P0) with P0 - e >= 2 g k, then equal parallelograms of ~P cells up to n - e, then
 * the inverted pyramid.  e = 1 on the row axis of a slab: its first and last row (whose neighbours live in other
 * slabs) then belong to the inverted pyramid at EVERY level, so that piece alone talks to the neighbours. */
void tb_axis(TbAxis *ax, int n, int g, int e, int k, int P)
{
	ax->n = n; ax->g = g; ax->e = e; ax->npieces = 0;
	if (P < 2 * g) P = 2 * g;
	int P0 = P < 2 * g * k + e ? 2 * g * k + e : P;
	if (P0 > n - e - g) P0 = n - e - g;
	P0 = P0 / g * g;
	ax->piece[ax->npieces++] = TbPiece{TB_PYRAMID, e, P0};
	const int end = n - e;
	if (end - P0 > 0) {
		int nb = (end - P0 + P - 1) / P;
		if (nb > 1023) nb = 1023;
		int Pe = ((end - P0 + nb - 1) / nb + g - 1) / g * g;     /* equal pieces instead of a short last one */
		for (int a = P0; a < end; a += Pe) ax->piece[ax->npieces++] = TbPiece{TB_PARALLELOGRAM, a, a + Pe < end ? a + Pe : end};
	}
	ax->piece[ax->npieces++] = TbPiece{TB_INVERTED, 0, 0};
}

/* depth, row bands and column strips actually used for a ring of nx rows x ny columns (k == 0: plain stepping).
 * tb_cols: 0 = automatic (strips only where full-width bands would be lower than 64 rows), >= ny = full width */
/* multi != 0: this slab is one of several.  Every slab of a ring must then walk the SAME sequence of (column piece,
 * level) events, so the depth is never clamped to the slab: a slab too low for the row pyramid (2k + 2 > nx)
 * takes whole-slab launches per event instead of row pieces. */
void tb_plan(TbPlan *pl, int nx, int ny, int ldy, int strips_ok, int multi, int tb_depth, int tb_rows, int tb_cols, int chains = 1, int chains_asked = 1)
{
	const int edge = multi != 0 || chains > 1;                  /* first and last row belong to the seam piece (a slab of a ring, a lane) */
	const int G = 8;                                             /* column skew per level: keeps every vector width aligned */
	int k = tb_depth & ~1;
	if (!edge)
		while (k >= 2 && 2 * k + 2 > nx) k -= 2;                 /* pyramid base >= 2k inside the ring */
	pl->k = k < 2 ? 0 : k;
	if (!pl->k) return;
	const double cell = 9.0 * sizeof(real_t);
	/* bytes of a tile incl. its skew margins, measured best: bands 45 MB (DP) / 60 MB (SP) (profiles/r01_run9,10),
	 * 2-D tiles 60 MB with 16 KB row segments (profiles/r01_run14) */
	double budget = sizeof(real_t) == 8 ? 45e6 : 60e6;
	int W = tb_cols;
	if (!strips_ok || ny % G || 2 * G * k + G > ny) W = ny;      /* strips need aligned, wide enough rows */
	else if (W == 0) W = budget / (cell * ldy) - k >= 64 ? ny : (int)(((multi && chains_asked < 2) ? 32768 : 16384) / sizeof(real_t));   /* single-chain
	                                                                  slabs: fewer, wider strips -- every (strip, level) costs a handshake with both
	                                                                  neighbours; with lanes 16 KB segments win on slabs too (profiles/r02_run8_2gpu).
	                                                                  chains_asked, not the lanes this slab has room for: every slab of a ring must
	                                                                  cut the same strips */
	if (W >= ny) {
		pl->cols.n = ny; pl->cols.g = G; pl->cols.e = 0; pl->cols.npieces = 1;
		pl->cols.piece[0] = TbPiece{TB_WHOLE, 0, ny};
		W = ldy;
	} else {
		W = (W + G - 1) / G * G;
		tb_axis(&pl->cols, ny, G, 0, k, W);
		W += G * k;
		budget = 60e6;
	}
	/* several concurrent chains share the L2: measured best per chain at 8192x32768 with two chains: 30 MB DP, 24-30 MB SP
	 * (profiles/r02_run7 and r02_run10 chain_sweep.log, r02_run8 chain_sweep_sp.log) */
	if (chains > 1) budget = (sizeof(real_t) == 8 ? 60e6 : 54e6) / chains;
	if (2 * k + 2 + 2 * edge > nx) {                             /* (slabs, lanes) no room for the row pyramid */
		pl->rows.n = nx; pl->rows.g = 1; pl->rows.e = 0; pl->rows.npieces = 1;
		pl->rows.piece[0] = TbPiece{TB_WHOLE, 0, nx};
		return;
	}
	int H = tb_rows;
	if (H <= 0) H = (int)(budget / (cell * W)) - k;
	tb_axis(&pl->rows, nx, 1, edge, k, H < 2 ? 2 : H);
}

/* the launches of one macro-step of kk <= pl.k levels, in order: visit(level, r0, r1, j0, j1); r1 / j1 may exceed
 * nx / ny (the range wraps); j0 == j1 == 0 means whole rows.  Column pieces outermost, row pieces next, levels
 * innermost: valid because each axis' piece order is (1-D argument in DESIGN.md) and the tile is their product. */
/* `seam` (slabs of a multi-slab ring): 3 = this launch contains the slab's first and last row, whose neighbours
 * live in other slabs -- wait for both neighbours' previous event before it (bit 0) and announce this slab's event
 * after it (bit 1); 0 = the launch depends on this slab's own rows only.  With e = 1 on the row axis those are
 * exactly the inverted-pyramid launches (one per column piece and level); whole-slab pieces likewise. */
template <typename F> int tb_for_each_launch(const TbPlan &pl, int kk, F visit)
{
	int rc;
	for (int ci = 0; ci < pl.cols.npieces; ci++)
		for (int ri = 0; ri < pl.rows.npieces; ri++)
			for (int s = 0; s < kk; s++) {
				int r0, r1, j0, j1;
				const int rt = pl.rows.piece[ri].type;
				tb_range(pl.rows, pl.rows.piece[ri], s, &r0, &r1);
				tb_range(pl.cols, pl.cols.piece[ci], s, &j0, &j1);
				if (r1 <= r0 || j1 <= j0) continue;                  /* inverted pyramids are empty at level 0 */
				if (pl.cols.piece[ci].type == TB_WHOLE) j0 = j1 = 0;
				const int seam = (rt == TB_WHOLE || rt == TB_INVERTED) ? 3 : 0;
				if ((rc = visit(s, r0, r1, j0, j1, seam))) return rc;
			}
	return 0;
}

/* Single slab: everything on stream 0.  Slab of a ring: the seam launches (inverted row pyramid: the only ones that
 * read or write another slab) form a chain on stream 1 -- wait for the neighbours' previous event, launch, announce --
 * that starts once this column piece's own tiles are done and runs beside the NEXT column piece's tiles on stream 0
 * (those never depend on seam rows of the same macro-step); the streams join at the end of the macro-step. */
int enqueue_macro(lbm_ctx *c, int t0, int kk, const TbPlan &pl, int nsteps)
{
	const bool flags = c->sync_mode == SYNC_FLAGS;
	bool chain_open = false;
	cudaError_t ce = cudaSuccess;
	int rc = tb_for_each_launch(pl, kk, [&](int s, int r0, int r1, int j0, int j1, int seam) {
		int r;
		const int tail = t0 + s != nsteps - 1;
		if (!(flags && seam)) {
			chain_open = false;
			return launch_step(c, (t0 + s) & 1, r0, r1, tail, c->stream[0], j0, j1, tb_hint(c, s, kk));
		}
		if (!chain_open) {
			if ((ce = cudaEventRecord(c->ev_main, c->stream[0])) != cudaSuccess) return (int)LBM_ERR_CUDA;
			if ((ce = cudaStreamWaitEvent(c->stream[1], c->ev_main, 0)) != cudaSuccess) return (int)LBM_ERR_CUDA;
			chain_open = true;
		}
		if (seam_fusable(c))                                         /* wait + step + announce in ONE launch */
			return launch_step(c, (t0 + s) & 1, r0, r1, tail, c->stream[1], j0, j1, tb_hint(c, s, kk), 1);
		if ((r = launch_wait(c, c->stream[1]))) return r;
		if ((r = launch_step(c, (t0 + s) & 1, r0, r1, tail, c->stream[1], j0, j1, tb_hint(c, s, kk)))) return r;
		return launch_signal(c, c->stream[1]);
	});
	if (ce != cudaSuccess) return fail(LBM_ERR_CUDA, "enqueue_macro: %s", cudaGetErrorString(ce));
	if (rc) return rc;
	if (flags) {
		CK(cudaEventRecord(c->ev_bnd[0], c->stream[1]));
		CK(cudaStreamWaitEvent(c->stream[0], c->ev_bnd[0], 0));
	}
	return LBM_OK;
}

/* ---- concurrent tile chains ("lanes") ----
 * Every level of a tile is one launch, and a launch loses about one CTA lifetime (~2 us) at its drain: the next level's CTAs are
 * resident (PDL) but must wait until the last CTA of this level is done.  A second, independent chain of launches fills that
 * gap (tools/two_chain_probe.py: DP 8192x32768 57.4 -> 65.6 GLUPS, SP 102.2 -> 119.5).  The rows of the slab are
 * cut into C lanes; each lane walks the schedule of a slab of a ring on its own stream (row axis with e = 1: pyramid from its
 * second row, parallelograms up to its last but one), and the boundary between lanes b and b+1 gets what a slab seam gets: an
 * inverted pyramid (rows [B-1-s, B+1+s) at level s) whose chain of launches runs on a stream of its own after both lanes'
 * tiles of the same column piece, beside their tiles of the next one.  Same buffer, same kernels, no halo in between: the
 * dependency argument is the one of the slab schedule (a lane's tiles never need seam rows of the same macro-step; a seam
 * launch needs the two lanes' tiles of its column piece and its own previous level).  The boundary between the last lane and
 * the first is the periodic wrap of a single slab or the seam of a ring (then it carries the neighbour barrier as before). */
struct LanePlan { int k, C; int base[LBM_MAX_CHAINS + 1]; TbAxis rows[LBM_MAX_CHAINS]; TbAxis cols; };

/* returns the number of lanes used (< 2: take the single-chain schedule) */
int tb_plan_lanes(LanePlan *lp, int nx, int ny, int ldy, int strips_ok, int multi, int tb_depth, int tb_rows, int tb_cols, int chains)
{
	const int asked = chains;
	static thread_local TbPlan pl;
	int k = tb_depth & ~1;
	if (k < 2 || chains < 2) return 1;
	int C = chains > LBM_MAX_CHAINS ? LBM_MAX_CHAINS : chains;
	if (C > nx / (2 * k + 4)) C = nx / (2 * k + 4);              /* every lane needs room for its row pyramid */
	if (C < 2) return 1;
	lp->k = k; lp->C = C;
	for (int l = 0; l <= C; l++) lp->base[l] = (int)((long long)nx * l / C);
	for (int l = 0; l < C; l++) {
		tb_plan(&pl, lp->base[l + 1] - lp->base[l], ny, ldy, strips_ok, multi, k, tb_rows, tb_cols, C, asked);
		if (pl.k != k || pl.rows.npieces < 2) return 1;
		lp->rows[l] = pl.rows;
		if (l == 0) lp->cols = pl.cols;
	}
	return C;
}

/* launch(stream, s, r0, r1, j0, j1, outer): stream 0..C-1 = lane, C..2C-1 = the chain of boundary b = stream - C; rows are
 * slab rows (r1 may exceed nx: the outer boundary wraps); outer = 1 for the launches of the last boundary.
 * event(kind, b_or_lane): 0 = lane finished its tiles of this column piece, 1 = chain b is about to start this column piece
 * (needs lanes b and b+1), 2 = chain b finished it. */
template <typename L, typename E> int tb_lanes_walk(const LanePlan &lp, int kk, L launch, E event)
{
	int rc;
	const int C = lp.C;
	for (int ci = 0; ci < lp.cols.npieces; ci++) {
		for (int l = 0; l < C; l++) {
			const TbAxis &rows = lp.rows[l];
			for (int ri = 0; ri < rows.npieces; ri++) {
				if (rows.piece[ri].type == TB_INVERTED) continue;
				for (int s = 0; s < kk; s++) {
					int r0, r1, j0, j1;
					tb_range(rows, rows.piece[ri], s, &r0, &r1);
					tb_range(lp.cols, lp.cols.piece[ci], s, &j0, &j1);
					if (r1 <= r0 || j1 <= j0) continue;
					if (lp.cols.piece[ci].type == TB_WHOLE) j0 = j1 = 0;
					if ((rc = launch(l, s, lp.base[l] + r0, lp.base[l] + r1, j0, j1, 0))) return rc;
				}
			}
			if ((rc = event(0, l))) return rc;
		}
		for (int b = 0; b < C; b++) {
			const TbAxis &rows = lp.rows[b];
			if ((rc = event(1, b))) return rc;
			for (int s = 0; s < kk; s++) {
				int r0, r1, j0, j1;
				tb_range(rows, rows.piece[rows.npieces - 1], s, &r0, &r1);
				tb_range(lp.cols, lp.cols.piece[ci], s, &j0, &j1);
				if (r1 <= r0 || j1 <= j0) continue;
				if (lp.cols.piece[ci].type == TB_WHOLE) j0 = j1 = 0;
				if ((rc = launch(C + b, s, lp.base[b] + r0, lp.base[b] + r1, j0, j1, b == C - 1))) return rc;
			}
			if ((rc = event(2, b))) return rc;
		}
	}
	return 0;
}

int ensure_lanes(lbm_ctx *c)
{
	if (c->lanes_ready) return LBM_OK;
	for (int k = 0; k < LBM_MAX_CHAINS; k++) {
		if (k > 0) CK(cudaStreamCreateWithFlags(&c->lane_stream[k], cudaStreamNonBlocking));
		CK(cudaStreamCreateWithFlags(&c->chain_stream[k], cudaStreamNonBlocking));
		CK(cudaEventCreateWithFlags(&c->ev_lane[k], cudaEventDisableTiming));
		CK(cudaEventCreateWithFlags(&c->ev_chain[k], cudaEventDisableTiming));
	}
	CK(cudaEventCreateWithFlags(&c->ev_fork, cudaEventDisableTiming));
	c->lanes_ready = 1;
	return LBM_OK;
}

inline cudaStream_t lane_st(lbm_ctx *c, int l) { return l == 0 ? c->stream[0] : c->lane_stream[l]; }
inline cudaStream_t chain_st(lbm_ctx *c, int b, int C) { return b == C - 1 ? c->stream[1] : c->chain_stream[b]; }

/* fork: every stream first waits for what stream 0 holds so far; join: stream 0 ends up behind everything enqueued here */
int enqueue_macro_lanes(lbm_ctx *c, int t0, int kk, const LanePlan &lp, int nsteps, bool fork, bool join)
{
	const bool flags = c->sync_mode == SYNC_FLAGS;
	const int C = lp.C;
	cudaError_t ce = cudaSuccess;
	if (fork) {
		CK(cudaEventRecord(c->ev_fork, c->stream[0]));
		for (int l = 1; l < C; l++) CK(cudaStreamWaitEvent(lane_st(c, l), c->ev_fork, 0));
		for (int b = 0; b < C; b++) CK(cudaStreamWaitEvent(chain_st(c, b, C), c->ev_fork, 0));
	}
	int rc = tb_lanes_walk(lp, kk,
		[&](int st, int s, int r0, int r1, int j0, int j1, int outer) {
			int r;
			const int tail = t0 + s != nsteps - 1;
			cudaStream_t q = st < C ? lane_st(c, st) : chain_st(c, st - C, C);
			if (!(outer && flags)) return launch_step(c, (t0 + s) & 1, r0, r1, tail, q, j0, j1, tb_hint(c, s, kk));
			if (seam_fusable(c)) return launch_step(c, (t0 + s) & 1, r0, r1, tail, q, j0, j1, tb_hint(c, s, kk), 1);
			if ((r = launch_wait(c, q))) return r;
			if ((r = launch_step(c, (t0 + s) & 1, r0, r1, tail, q, j0, j1, tb_hint(c, s, kk)))) return r;
			return launch_signal(c, q);
		},
		[&](int kind, int i) {
			if (kind == 0) ce = cudaEventRecord(c->ev_lane[i], lane_st(c, i));
			else if (kind == 1) {
				ce = cudaStreamWaitEvent(chain_st(c, i, C), c->ev_lane[i], 0);
				if (ce == cudaSuccess) ce = cudaStreamWaitEvent(chain_st(c, i, C), c->ev_lane[(i + 1) % C], 0);
			} else ce = cudaEventRecord(c->ev_chain[i], chain_st(c, i, C));
			return ce == cudaSuccess ? 0 : (int)LBM_ERR_CUDA;
		});
	if (ce != cudaSuccess) return fail(LBM_ERR_CUDA, "enqueue_macro_lanes: %s", cudaGetErrorString(ce));
	if (rc) return rc;
	for (int l = 0; l < C; l++) {                                /* the next macro-step's tiles of a lane need its two seam chains */
		CK(cudaStreamWaitEvent(lane_st(c, l), c->ev_chain[l], 0));
		CK(cudaStreamWaitEvent(lane_st(c, l), c->ev_chain[(l + C - 1) % C], 0));
	}
	if (join)
		for (int l = 1; l < C; l++) {
			CK(cudaEventRecord(c->ev_lane[l], lane_st(c, l)));
			CK(cudaStreamWaitEvent(c->stream[0], c->ev_lane[l], 0));
		}
	return LBM_OK;
}

int run_wavefront(lbm_ctx *c, int nsteps)
{
	int rc;
	static thread_local TbPlan pl;                               /* ~25 KB: not on the stack of every call */
	static thread_local LanePlan lp;
	const int strips_ok = c->inplace && c->vw > 1 && c->nyvec == c->ny;
	tb_plan(&pl, c->nxl, c->ny, c->ldy, strips_ok, c->sync_mode == SYNC_FLAGS, c->tune.tb_depth, c->tune.tb_rows, c->tune.tb_cols, 1, strips_ok ? c->tune.tb_chains : 1);
	const int k = pl.k;
	if (k < 2) {
		for (int t = 0; t < nsteps; t++)
			if ((rc = enqueue_step(c, t & 1, t != nsteps - 1))) return rc;
		return LBM_OK;
	}
	/* concurrent chains need the in-place vector kernels on every column (a launch is then one kernel) */
	const int C = strips_ok ? tb_plan_lanes(&lp, c->nxl, c->ny, c->ldy, strips_ok, c->sync_mode == SYNC_FLAGS, k, c->tune.tb_rows, c->tune.tb_cols, c->tune.tb_chains) : 1;
	if (C >= 2 && (rc = ensure_lanes(c))) return rc;
	auto macro = [&](int t0, int kk, int ns, bool fork, bool join) {
		return C >= 2 ? enqueue_macro_lanes(c, t0, kk, lp, ns, fork, join) : enqueue_macro(c, t0, kk, pl, ns);
	};
	int t = 0;
	struct Scope { lbm_ctx *c; Scope(lbm_ctx *c_) : c(c_) { c->in_wavefront = 1; } ~Scope() { c->in_wavefront = 0; } } scope(c);
	if (c->tune.graph > 0 && nsteps - k >= k && c->sync_mode == SYNC_SELF) {      /* slabs: the epochs of wait/signal change per launch */
		if (!c->wf_exec) {                                       /* any change of tuning drops the graph */
			cudaGraph_t g = nullptr;
			CK(cudaStreamBeginCapture(c->stream[0], cudaStreamCaptureModeThreadLocal));
			const long long l0 = c->launches;
			rc = macro(0, k, 1 << 30, true, true);               /* every level with its tail source */
			c->wf_launches = c->launches - l0;
			c->launches = l0;
			cudaError_t e = cudaStreamEndCapture(c->stream[0], &g);
			if (rc != LBM_OK) { if (g) cudaGraphDestroy(g); return rc; }
			CK(e);
			e = cudaGraphInstantiate(&c->wf_exec, g, 0);
			cudaGraphDestroy(g);
			CK(e);
		}
		for (; nsteps - k - t >= k; t += k) {                    /* keep the last macro-step out: its last level leaves A un-sourced */
			CK(cudaGraphLaunch(c->wf_exec, c->stream[0]));
			c->launches += c->wf_launches;
		}
	}
	bool first = true;
	while (t < nsteps) {
		const int kk = nsteps - t < k ? nsteps - t : k;          /* even; pyramid bases >= 2k >= 2kk */
		if ((rc = macro(t, kk, nsteps, first, t + kk >= nsteps))) return rc;
		first = false;
		t += kk;
	}
	return LBM_OK;
}

// ApplySource + halo refresh + neighbour barrier that open every run (the first step's source
// cannot be fused into a previous kernel's tail)
int enqueue_prologue(lbm_ctx *c, int buf)
{
	int rc;
	if ((rc = launch_source(c, buf, c->stream[0]))) return rc;
	if ((rc = launch_halo_push(c, buf, c->stream[0]))) return rc;
	if (c->sync_mode == SYNC_FLAGS && (rc = launch_signal(c, c->stream[0]))) return rc;
	return LBM_OK;
}

} // namespace

extern "C" {

// ---------------------------------------------------------------------------
// inputs
// ---------------------------------------------------------------------------
int lbm_initialize_arrays(lbm_ctx *c, const lbm_scene *scene)
{
	if (!c || !scene) return fail(LBM_ERR_ARG, "lbm_initialize_arrays: null argument");
	int rc;
	if ((rc = set_device(c))) return rc;
	SceneDev sc;
	sc.nx = c->cfg.nx;
	sc.channel_walls = scene->channel_walls;
	sc.bar_i0 = scene->bar_i0; sc.bar_i1 = scene->bar_i1; sc.bar_j0 = scene->bar_j0; sc.bar_j1 = scene->bar_j1;
	sc.disc_ci = scene->disc_ci; sc.disc_cj = scene->disc_cj;
	sc.disc_r2 = scene->disc_r < 0 ? -1 : scene->disc_r * scene->disc_r;
	cudaStream_t st = c->stream[0];
	const size_t nwords = (size_t)c->nxl * c->ldw, nsites = (size_t)c->nxl * c->ny;
	if (scene->channel_walls < 0 || scene->channel_walls > 2) return fail(LBM_ERR_ARG, "lbm_initialize_arrays: channel_walls %d not in {0,1,2}", scene->channel_walls);
	scene_mask_kernel<<<(unsigned)((nwords + 255) / 256), 256, 0, st>>>(c->mask, c->slip, sc, c->cfg.row_lo, c->nxl, c->ny, c->ldw);
	CK(cudaGetLastError());
	c->has_slip = scene->channel_walls == 2;
	drop_graph(c);                                               /* captured launches carry the slip pointer */
	/* INITIALDENSITY * OMEGA_s in real_t (LB.c:583-591) */
	scene_fill_kernel<real_t><<<(unsigned)((nsites + 255) / 256), 256, 0, st>>>(
		c->f[0], c->f[1] /* nullptr when in place */, c->plane, c->ldy, c->nxl, c->ny, c->cfg.row_lo, scene->rho0, c->w0, c->w14, c->w58,
		scene->perturb_seed, scene->perturb_amp);
	CK(cudaGetLastError());
	c->launches += 2;
	c->has_state = 1;
	return LBM_OK;
}

/* walls: chars (0 fluid, 1 bounce-back, 2 / 3 reflect with normal x / y) -> device scratch -> packed bit planes.
 * The slab's first row is host row `first`; row stride ldy.  Synchronous (the values are checked). */
static int set_walls_impl(lbm_ctx *c, const char *walls, int ldy, int first)
{
	cudaStream_t st = c->stream[0];
	char *dwalls = nullptr;
	int *dflags = nullptr;
	const size_t ldc = (size_t)c->ny;
	CK(cudaMallocAsync(&dwalls, (size_t)c->nxl * ldc, st));
	CK(cudaMallocAsync(&dflags, 2 * sizeof(int), st));
	CK(cudaMemsetAsync(dflags, 0, 2 * sizeof(int), st));
	CK(cudaMemcpy2DAsync(dwalls, ldc, walls + (size_t)first * ldy, (size_t)ldy, (size_t)c->ny, (size_t)c->nxl, cudaMemcpyHostToDevice, st));
	const size_t nwords = (size_t)c->nxl * c->ldw;
	mask_pack_kernel<<<(unsigned)((nwords + 255) / 256), 256, 0, st>>>(dwalls, ldc, c->mask, c->slip, c->nxl, c->ny, c->ldw, dflags);
	CK(cudaGetLastError());
	c->launches++;
	int fl[2] = {0, 0};
	CK(cudaMemcpyAsync(fl, dflags, sizeof fl, cudaMemcpyDeviceToHost, st));
	CK(cudaFreeAsync(dwalls, st));
	CK(cudaFreeAsync(dflags, st));
	CK(cudaStreamSynchronize(st));
	if (fl[0]) return fail(LBM_ERR_ARG, "walls must hold 0 (fluid), 1 (bounce-back: LB.c:246 tests == 1, LB.c:455 tests == 0), 2 or 3 (reflect, normal x / y)");
	c->has_slip = fl[1] != 0;
	drop_graph(c);
	return LBM_OK;
}

/* host arrays hold `host_rows` lattice rows per plane; the slab's first row is host row `first` */
static int upload_impl(lbm_ctx *c, const real_t *fA, const real_t *fB, const char *walls, int ldy, int host_rows, int first)
{
	if (!c || !fA || !walls) return fail(LBM_ERR_ARG, "lbm_upload: null argument");
	if (ldy < c->ny) return fail(LBM_ERR_ARG, "lbm_upload: ldy %d < ny %d", ldy, c->ny);
	if (fB && c->inplace) return fail(LBM_ERR_STATE, "lbm_upload: an in-place context has no buffer B (pass fB = NULL)");
	int rc;
	if ((rc = set_device(c))) return rc;
	cudaStream_t st = c->stream[0];
	const size_t hplane = (size_t)host_rows * ldy;              /* host plane stride, I() of LB.c:72 */
	const size_t row0 = (size_t)first * ldy;
	for (int b = 0; b < 2; b++) {
		const real_t *src = b == 0 ? fA : fB;
		if (!src) {                                              /* fB = 0 (LB.c:593-601) */
			if (c->f[1]) CK(cudaMemsetAsync(c->f[1], 0, 9 * c->plane * sizeof(real_t), st));
			continue;
		}
		for (int s = 0; s < 9; s++)
			CK(cudaMemcpy2DAsync(c->f[b] + s * c->plane + c->ldy, (size_t)c->ldy * sizeof(real_t),
			                     src + s * hplane + row0, (size_t)ldy * sizeof(real_t),
			                     (size_t)c->ny * sizeof(real_t), (size_t)c->nxl, cudaMemcpyHostToDevice, st));
	}
	if ((rc = set_walls_impl(c, walls, ldy, first))) return rc;
	c->has_state = 1;
	return LBM_OK;
}

int lbm_upload(lbm_ctx *c, const real_t *fA, const real_t *fB, const char *walls, int ldy)
{
	if (!c) return fail(LBM_ERR_ARG, "lbm_upload: null ctx");
	return upload_impl(c, fA, fB, walls, ldy, c->cfg.nx, c->cfg.row_lo);
}

int lbm_upload_slab(lbm_ctx *c, const real_t *fA, const real_t *fB, const char *walls, int ldy)
{
	if (!c) return fail(LBM_ERR_ARG, "lbm_upload_slab: null ctx");
	return upload_impl(c, fA, fB, walls, ldy, c->nxl, 0);
}

int lbm_set_walls(lbm_ctx *c, const char *walls, int ldy)
{
	if (!c || !walls) return fail(LBM_ERR_ARG, "lbm_set_walls: null argument");
	if (ldy < c->ny) return fail(LBM_ERR_ARG, "lbm_set_walls: ldy %d < ny %d", ldy, c->ny);
	int rc;
	if ((rc = lbm_sync(c))) return rc;
	return set_walls_impl(c, walls, ldy, c->cfg.row_lo);
}

/* Geometry from a file instead of the hard-wired barrier of LB.c:567-578: text, one lattice row per line, NY entries
 * per line, each one of the digits 0 1 2 3, optionally separated by blanks and/or commas (so both "0100" and the
 * "0, 1, 0, 0" style of the reference's own data/<step>.csv dumps parse); lines starting with '#' are comments. */
int lbm_load_walls(lbm_ctx *c, const char *path)
{
	if (!c || !path) return fail(LBM_ERR_ARG, "lbm_load_walls: null argument");
	FILE *fp = fopen(path, "r");
	if (!fp) return fail(LBM_ERR_IO, "Error opening file %s", path);
	const int nx = c->cfg.nx, ny = c->ny;
	std::vector<char> walls((size_t)nx * ny);
	int row = 0, col = 0, ch, rc = LBM_OK;
	bool comment = false, line_start = true;
	while (rc == LBM_OK && (ch = fgetc(fp)) != EOF) {
		if (comment) { if (ch == '\n') { comment = false; line_start = true; } continue; }
		if (ch == '#' && line_start) { comment = true; continue; }
		if (ch == '\n') {
			if (col == 0) { line_start = true; continue; }                         /* blank line */
			if (col != ny) rc = fail(LBM_ERR_IO, "%s: row %d has %d entries, the lattice has %d columns", path, row, col, ny);
			row++; col = 0; line_start = true;
			continue;
		}
		line_start = false;
		if (ch == ' ' || ch == '\t' || ch == ',' || ch == '\r') continue;
		if (ch < '0' || ch > '3') { rc = fail(LBM_ERR_IO, "%s: row %d: unexpected character '%c' (entries are 0, 1, 2 or 3)", path, row, ch); break; }
		if (row >= nx || col >= ny) { rc = fail(LBM_ERR_IO, "%s: more than %d rows of %d entries", path, nx, ny); break; }
		walls[(size_t)row * ny + col++] = (char)(ch - '0');
	}
	fclose(fp);
	if (rc != LBM_OK) return rc;
	if (col == ny) { row++; col = 0; }                                              /* last line without a newline */
	if (col != 0 || row != nx) return fail(LBM_ERR_IO, "%s: %d complete rows, the lattice has %d", path, row, nx);
	return lbm_set_walls(c, walls.data(), ny);
}

static int download_impl(lbm_ctx *c, real_t *fA, real_t *fB, char *walls_out, int ldy, int host_rows, int first)
{
	if (!c) return fail(LBM_ERR_ARG, "lbm_download: null ctx");
	if (ldy < c->ny) return fail(LBM_ERR_ARG, "lbm_download: ldy %d < ny %d", ldy, c->ny);
	if (fB && c->inplace) return fail(LBM_ERR_STATE, "lbm_download: an in-place context has no buffer B");
	int rc;
	if ((rc = lbm_sync(c))) return rc;
	cudaStream_t st = c->stream[0];
	const size_t hplane = (size_t)host_rows * ldy, row0 = (size_t)first * ldy;
	for (int b = 0; b < 2; b++) {
		real_t *dst = b == 0 ? fA : fB;
		if (!dst) continue;
		for (int s = 0; s < 9; s++)
			CK(cudaMemcpy2DAsync(dst + s * hplane + row0, (size_t)ldy * sizeof(real_t),
			                     c->f[b] + s * c->plane + c->ldy, (size_t)c->ldy * sizeof(real_t),
			                     (size_t)c->ny * sizeof(real_t), (size_t)c->nxl, cudaMemcpyDeviceToHost, st));
	}
	if (walls_out) {
		char *dwalls = nullptr;
		const size_t ldc = (size_t)c->ny, nsites = (size_t)c->nxl * c->ny;
		CK(cudaMallocAsync(&dwalls, nsites, st));
		mask_unpack_kernel<<<(unsigned)((nsites + 255) / 256), 256, 0, st>>>(c->mask, c->has_slip ? c->slip : nullptr, dwalls, ldc, c->nxl, c->ny, c->ldw);
		CK(cudaGetLastError());
		c->launches++;
		CK(cudaMemcpy2DAsync(walls_out + row0, (size_t)ldy, dwalls, ldc, (size_t)c->ny, (size_t)c->nxl, cudaMemcpyDeviceToHost, st));
		CK(cudaFreeAsync(dwalls, st));
	}
	CK(cudaStreamSynchronize(st));
	return LBM_OK;
}

int lbm_download_window(lbm_ctx *c, int buf, int r0, int r1, int j0, int j1, real_t *out)
{
	if (!c || !out || (buf != 0 && buf != 1)) return fail(LBM_ERR_ARG, "lbm_download_window: bad argument");
	if (buf == 1 && c->inplace) return fail(LBM_ERR_STATE, "lbm_download_window: an in-place context has no buffer B");
	if (r0 < 0 || r1 > c->nxl || r0 >= r1 || j0 < 0 || j1 > c->ny || j0 >= j1)
		return fail(LBM_ERR_ARG, "lbm_download_window: rows [%d,%d) of %d, columns [%d,%d) of %d", r0, r1, c->nxl, j0, j1, c->ny);
	int rc;
	if ((rc = lbm_sync(c))) return rc;
	cudaStream_t st = c->stream[0];
	const size_t w = (size_t)(j1 - j0), h = (size_t)(r1 - r0);
	for (int s = 0; s < 9; s++)
		CK(cudaMemcpy2DAsync(out + (size_t)s * h * w, w * sizeof(real_t),
		                     c->f[buf] + s * c->plane + (size_t)(r0 + 1) * c->ldy + j0, (size_t)c->ldy * sizeof(real_t),
		                     w * sizeof(real_t), h, cudaMemcpyDeviceToHost, st));
	CK(cudaStreamSynchronize(st));
	return LBM_OK;
}

int lbm_download(lbm_ctx *c, real_t *fA, real_t *fB, char *walls_out, int ldy)
{
	if (!c) return fail(LBM_ERR_ARG, "lbm_download: null ctx");
	return download_impl(c, fA, fB, walls_out, ldy, c->cfg.nx, c->cfg.row_lo);
}

int lbm_download_slab(lbm_ctx *c, real_t *fA, real_t *fB, char *walls_out, int ldy)
{
	if (!c) return fail(LBM_ERR_ARG, "lbm_download_slab: null ctx");
	return download_impl(c, fA, fB, walls_out, ldy, c->nxl, 0);
}

// ---------------------------------------------------------------------------
// the timestep path
// ---------------------------------------------------------------------------
int lbm_sync(lbm_ctx *c)
{
	if (!c) return fail(LBM_ERR_ARG, "lbm_sync: null ctx");
	int rc;
	if ((rc = set_device(c))) return rc;
	CK(cudaStreamSynchronize(c->stream[1]));
	CK(cudaStreamSynchronize(c->stream[0]));
	if (c->sync_mode == SYNC_FLAGS) {
		int fl[4];
		CK(cudaMemcpy(fl, c->flags, sizeof fl, cudaMemcpyDeviceToHost));
		if (fl[3]) return fail(LBM_ERR_STATE, "neighbour barrier timed out (a ring neighbour stopped stepping)");
	}
	return LBM_OK;
}

int lbm_run(lbm_ctx *c, int nsteps)
{
	if (!c) return fail(LBM_ERR_ARG, "lbm_run: null ctx");
	if (nsteps < 0 || (nsteps & 1)) return fail(LBM_ERR_ARG, "lbm_run: nsteps %d must be even (LB.c:148 advances by 2)", nsteps);
	if (!c->has_state) return fail(LBM_ERR_STATE, "lbm_run: no lattice loaded (lbm_upload / lbm_initialize_arrays first)");
	if (nsteps == 0) return LBM_OK;
	int rc;
	if ((rc = set_device(c))) return rc;
	/* Source(A) of the first step; every later Source is fused into the tail of the step
	 * before it, the last step leaves A un-sourced as the reference does (LB.c:192-205). */
	if ((rc = enqueue_prologue(c, LBM_BUF_A))) return rc;
	if (c->inplace && (c->sync_mode == SYNC_SELF || c->sync_mode == SYNC_FLAGS) && c->tune.tb_depth >= 2) return run_wavefront(c, nsteps);
	int t = 0;
	if (c->tune.graph > 0 && c->sync_mode == SYNC_SELF) {
		/* replay whole (A->B, B->A) pairs, all with the tail source, from a captured graph; the last
		 * pair (whose final step must leave A un-sourced) is launched directly below */
		const int pairs = c->tune.graph;
		if (nsteps - 2 >= 2 * pairs) {
			if (!c->graph_exec || c->graph_pairs != pairs) {
				drop_graph(c);
				cudaGraph_t g = nullptr;
				CK(cudaStreamBeginCapture(c->stream[0], cudaStreamCaptureModeThreadLocal));
				rc = LBM_OK;
				const long long l0 = c->launches;
				for (int k = 0; k < 2 * pairs && rc == LBM_OK; k++) rc = enqueue_step(c, k & 1, 1);
				c->launches = l0;                               /* captured, not launched */
				cudaError_t e = cudaStreamEndCapture(c->stream[0], &g);
				if (rc != LBM_OK) { if (g) cudaGraphDestroy(g); return rc; }
				CK(e);
				e = cudaGraphInstantiate(&c->graph_exec, g, 0);
				cudaGraphDestroy(g);
				CK(e);
				c->graph_pairs = pairs;
			}
			const int per_graph = (c->nyvec == c->ny || c->vw == 1) ? 1 : 2;   /* kernels per step */
			for (; nsteps - 2 - t >= 2 * pairs; t += 2 * pairs) {
				CK(cudaGraphLaunch(c->graph_exec, c->stream[0]));
				c->launches += (long long)2 * pairs * per_graph;
			}
		}
	}
	for (; t < nsteps; t++)
		if ((rc = enqueue_step(c, t & 1, t != nsteps - 1))) return rc;
	return LBM_OK;
}

int lbm_do_timestep(lbm_ctx *c) { return lbm_run(c, 2); }

int lbm_run_timed(lbm_ctx *c, int nsteps, float *ms)
{
	if (!c || !ms) return fail(LBM_ERR_ARG, "lbm_run_timed: null argument");
	int rc;
	if ((rc = lbm_sync(c))) return rc;
	CK(cudaEventRecord(c->ev_t0, c->stream[0]));
	if ((rc = lbm_run(c, nsteps))) return rc;
	CK(cudaEventRecord(c->ev_t1, c->stream[0]));
	if ((rc = lbm_sync(c))) return rc;
	CK(cudaEventElapsedTime(ms, c->ev_t0, c->ev_t1));
	return LBM_OK;
}

int lbm_apply_source(lbm_ctx *c, int buf)
{
	if (!c || (buf != 0 && buf != 1)) return fail(LBM_ERR_ARG, "lbm_apply_source: bad argument");
	if (c->sync_mode != SYNC_SELF) return fail(LBM_ERR_STATE, "lbm_apply_source: single-slab contexts only (use lbm_run)");
	if (c->inplace && buf != 0) return fail(LBM_ERR_STATE, "lbm_apply_source: an in-place context has no buffer B");
	int rc;
	if ((rc = set_device(c))) return rc;
	if ((rc = launch_source(c, buf, c->stream[0]))) return rc;
	return launch_halo_push(c, buf, c->stream[0]);
}

int lbm_stream_collide(lbm_ctx *c, int src_buf)
{
	if (!c || (src_buf != 0 && src_buf != 1)) return fail(LBM_ERR_ARG, "lbm_stream_collide: bad argument");
	if (c->sync_mode != SYNC_SELF) return fail(LBM_ERR_STATE, "lbm_stream_collide: single-slab contexts only (use lbm_run)");
	if (!c->has_state) return fail(LBM_ERR_STATE, "lbm_stream_collide: no lattice loaded");
	if (c->inplace) return fail(LBM_ERR_STATE, "lbm_stream_collide: in-place contexts step in pairs (lbm_do_timestep / lbm_run); "
	                                            "between the two steps of a pair the buffer is not in the reference layout");
	int rc;
	if ((rc = set_device(c))) return rc;
	/* the source buffer may have been written by upload/apply_source: make its halo rows current */
	if ((rc = launch_halo_push(c, src_buf, c->stream[0]))) return rc;
	return launch_step(c, src_buf, 0, c->nxl, 0, c->stream[0]);
}

} // extern "C"

extern "C" {

// ---------------------------------------------------------------------------
// observers
// ---------------------------------------------------------------------------
int lbm_reynolds_column(lbm_ctx *c, real_t *uy, char *is_fluid)
{
	if (!c || !uy || !is_fluid) return fail(LBM_ERR_ARG, "lbm_reynolds_column: null argument");
	int rc;
	if ((rc = lbm_sync(c))) return rc;
	cudaStream_t st = c->stream[0];
	const int jc = (int)((real_t)c->ny / 2.0);                  /* LB.c:527 */
	real_t *duy = nullptr;
	CK(cudaMallocAsync(&duy, (size_t)c->nxl * sizeof(real_t), st));
	macroscopic_kernel<real_t><<<(c->nxl + 127) / 128, 128, 0, st>>>(c->f[0], c->plane, c->ldy, c->nxl, jc, 1,
	                                                                 (real_t *)nullptr, (real_t *)nullptr, duy);
	CK(cudaGetLastError());
	c->launches++;
	CK(cudaMemcpyAsync(uy, duy, (size_t)c->nxl * sizeof(real_t), cudaMemcpyDeviceToHost, st));
	std::vector<uint32_t> words((size_t)c->nxl);
	CK(cudaMemcpy2DAsync(words.data(), sizeof(uint32_t), c->mask + (jc >> 5), (size_t)c->ldw * sizeof(uint32_t),
	                     sizeof(uint32_t), (size_t)c->nxl, cudaMemcpyDeviceToHost, st));
	CK(cudaFreeAsync(duy, st));
	CK(cudaStreamSynchronize(st));
	for (int r = 0; r < c->nxl; r++) is_fluid[r] = (char)(((words[r] >> (jc & 31)) & 1u) == 0u);
	return LBM_OK;
}

/* ---- lbm_reynolds_column without any allocation (include/lbm_b200.h) ----
 * The scratch lbm_reynolds_column takes from cudaMallocAsync lives here in a file-scope device array instead: LBM_STATIC_MAX_ROWS
 * values of u_y followed by as many fluid flags (576 KB in the DP library, 320 KB in the SP one), held by every GPU that loads
 * the library until the process exits.  In host code `g_static_scratch` is only a stand-in for the symbol, NOT a device address:
 * the address is asked for once per device, with that device current (cudaGetSymbolAddress), and that pointer goes to the kernel. */
static __device__ __align__(256) unsigned char g_static_scratch[(size_t)LBM_STATIC_MAX_ROWS * (sizeof(real_t) + 1)];

static int static_scratch_address(const lbm_ctx *c, unsigned char **out)
{
	static unsigned char *cached[64] = {nullptr};                /* per device; racing first calls store the same value */
	const int dev = c->cfg.device;
	if (dev < 0 || dev >= 64) return fail(LBM_ERR_ARG, "lbm_reynolds_column_static: device %d", dev);
	if (!cached[dev]) {
		void *ptr = nullptr;
		cudaError_t e = cudaGetSymbolAddress(&ptr, g_static_scratch);   /* the caller made c's device current (set_device) */
		if (e != cudaSuccess || !ptr) return fail(LBM_ERR_CUDA, "lbm_reynolds_column_static: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
		cached[dev] = (unsigned char *)ptr;
	}
	*out = cached[dev];
	return LBM_OK;
}

int lbm_reynolds_column_static(lbm_ctx *c, real_t *uy, char *is_fluid)
{
	if (!c || !uy || !is_fluid) return fail(LBM_ERR_ARG, "lbm_reynolds_column_static: null argument");
	if (c->nxl > LBM_STATIC_MAX_ROWS)
		return fail(LBM_ERR_ARG, "lbm_reynolds_column_static: %d local rows exceed the static scratch of LBM_STATIC_MAX_ROWS = %d rows "
		                         "(use lbm_reynolds_column)", c->nxl, LBM_STATIC_MAX_ROWS);
	if (!c->has_state) return fail(LBM_ERR_STATE, "lbm_reynolds_column_static: no lattice loaded");
	int rc;
	if ((rc = lbm_sync(c))) return rc;                           /* sets the device; no allocation */
	unsigned char *scratch = nullptr;
	if ((rc = static_scratch_address(c, &scratch))) return rc;
	real_t *duy = (real_t *)scratch;
	char *dfl = (char *)(scratch + (size_t)LBM_STATIC_MAX_ROWS * sizeof(real_t));
	cudaStream_t st = c->stream[0];
	const int jc = (int)((real_t)c->ny / 2.0);                  /* LB.c:527 */
	/* nothing in the scratch has to start at zero: the kernel writes every entry [0, nxl) that is copied out below */
	reynolds_column_static_kernel<real_t><<<(c->nxl + 127) / 128, 128, 0, st>>>(c->f[0], c->plane, c->ldy, c->nxl, jc, c->mask, c->ldw, duy, dfl);
	CK(cudaGetLastError());
	c->launches++;
	CK(cudaMemcpyAsync(uy, duy, (size_t)c->nxl * sizeof(real_t), cudaMemcpyDeviceToHost, st));
	CK(cudaMemcpyAsync(is_fluid, dfl, (size_t)c->nxl, cudaMemcpyDeviceToHost, st));
	CK(cudaStreamSynchronize(st));
	return LBM_OK;
}

/* the serial accumulation and the final scaling of ComputeReynolds, LB.c:529-546 */
real_t lbm_reynolds_from_column(const real_t *uy, const char *is_fluid, int nx, real_t tau)
{
	real_t total = 0;
	int points = 0;
	for (int i = 0; i < nx; i++)
		if (is_fluid[i]) { total += uy[i]; points++; }
	const real_t viscosity = REAL_T(1.0) / REAL_T(3.0) * (tau - REAL_T(1.0) / REAL_T(2.0));
	return total / (real_t)points * REAL_T(10.0) / viscosity;
}

int lbm_compute_reynolds(lbm_ctx *c, real_t *re)
{
	if (!c || !re) return fail(LBM_ERR_ARG, "lbm_compute_reynolds: null argument");
	if (c->nxl != c->cfg.nx) return fail(LBM_ERR_STATE, "lbm_compute_reynolds: slab context; gather lbm_reynolds_column over the ring");
	std::vector<real_t> uy((size_t)c->nxl);
	std::vector<char> fl((size_t)c->nxl);
	int rc;
	if ((rc = lbm_reynolds_column(c, uy.data(), fl.data()))) return rc;
	*re = lbm_reynolds_from_column(uy.data(), fl.data(), c->nxl, c->cfg.tau);
	return LBM_OK;
}

int lbm_macroscopic(lbm_ctx *c, real_t *rho, real_t *ux, real_t *uy)
{
	if (!c) return fail(LBM_ERR_ARG, "lbm_macroscopic: null ctx");
	int rc;
	if ((rc = lbm_sync(c))) return rc;
	cudaStream_t st = c->stream[0];
	const size_t n = (size_t)c->nxl * c->ny;
	real_t *d[3] = {nullptr, nullptr, nullptr};
	real_t *h[3] = {rho, ux, uy};
	for (int k = 0; k < 3; k++)
		if (h[k]) CK(cudaMallocAsync(&d[k], n * sizeof(real_t), st));
	macroscopic_kernel<real_t><<<(unsigned)((n + 255) / 256), 256, 0, st>>>(c->f[0], c->plane, c->ldy, c->nxl, 0, c->ny, d[0], d[1], d[2]);
	CK(cudaGetLastError());
	c->launches++;
	for (int k = 0; k < 3; k++)
		if (h[k]) {
			CK(cudaMemcpyAsync(h[k], d[k], n * sizeof(real_t), cudaMemcpyDeviceToHost, st));
			CK(cudaFreeAsync(d[k], st));
		}
	CK(cudaStreamSynchronize(st));
	return LBM_OK;
}

/* the rows of one slab, appended to an open file: PrintLattice's loop body (LB.c:618-637) */
static int print_rows(lbm_ctx *c, FILE *fp)
{
	const size_t n = (size_t)c->nxl * c->ny;
	std::vector<real_t> ux(n), uy(n);
	int rc;
	if ((rc = lbm_macroscopic(c, nullptr, ux.data(), uy.data()))) return rc;
	std::vector<char> line((size_t)c->ny * 400 + 8);                       /* "%.10lf" of DBL_MAX is 320 characters (a diverged run) */
	for (int i = 0; i < c->nxl; i++) {
		size_t pos = 0;
		for (int j = 0; j < c->ny; j++) {
			const real_t a = ux[(size_t)i * c->ny + j], b = uy[(size_t)i * c->ny + j];
			const real_t u2 = a * a + b * b;                                 /* LB.c:631, real_t arithmetic */
			pos += (size_t)snprintf(&line[pos], 398, "%.10lf", (double)u2);  /* LB.c:633 */
			if (j < c->ny - 1) { line[pos++] = ','; line[pos++] = ' '; }     /* LB.c:634 */
		}
		line[pos++] = '\n';
		if (fwrite(line.data(), 1, pos, fp) != pos) return fail(LBM_ERR_IO, "PrintLattice: short write");
	}
	return LBM_OK;
}

int lbm_print_lattice_group(lbm_ctx **slabs, int n, const char *path)
{
	if (!slabs || n < 1 || !path) return fail(LBM_ERR_ARG, "lbm_print_lattice_group: bad argument");
	int next = 0;
	for (int k = 0; k < n; k++) {                                          /* one file = the whole lattice, rows in order */
		if (!slabs[k] || slabs[k]->cfg.row_lo != next) return fail(LBM_ERR_ARG, "lbm_print_lattice_group: slab %d does not continue at row %d", k, next);
		next = slabs[k]->cfg.row_hi;
	}
	if (next != slabs[0]->cfg.nx) return fail(LBM_ERR_ARG, "lbm_print_lattice_group: the slabs cover rows [0,%d) of %d", next, slabs[0]->cfg.nx);
	FILE *fp = fopen(path, "w");
	if (!fp) return fail(LBM_ERR_IO, "Error opening file %s", path);       /* LB.c:616 */
	int rc = LBM_OK;
	for (int k = 0; k < n && rc == LBM_OK; k++) rc = print_rows(slabs[k], fp);
	fclose(fp);
	return rc;
}

int lbm_print_lattice(lbm_ctx *c, const char *path)
{
	if (!c || !path) return fail(LBM_ERR_ARG, "lbm_print_lattice: null argument");
	FILE *fp = fopen(path, "w");
	if (!fp) return fail(LBM_ERR_IO, "Error opening file %s", path);       /* LB.c:616 */
	const int rc = print_rows(c, fp);
	fclose(fp);
	return rc;
}

static int reduce_u64(lbm_ctx *c, int buf, int what, unsigned long long *out)
{
	int rc;
	if ((rc = lbm_sync(c))) return rc;
	cudaStream_t st = c->stream[0];
	unsigned long long *d = nullptr;
	CK(cudaMallocAsync(&d, sizeof *d, st));
	CK(cudaMemsetAsync(d, 0, sizeof *d, st));
	const size_t total = (size_t)c->nxl * c->ny;
	size_t grid = (total + 255) / 256;
	const size_t cap = (size_t)sm_count(c->cfg.device) * 8;              /* one resident wave, grid-stride */
	if (grid > cap) grid = cap;
	if (what == 0) checksum_kernel<real_t><<<(unsigned)grid, 256, 0, st>>>(c->f[buf], c->plane, c->ldy, c->nxl, c->ny, c->cfg.row_lo, c->cfg.nx, d);
	else nonfinite_kernel<real_t><<<(unsigned)grid, 256, 0, st>>>(c->f[buf], c->plane, c->ldy, c->nxl, c->ny, d);
	CK(cudaGetLastError());
	c->launches++;
	CK(cudaMemcpyAsync(out, d, sizeof *d, cudaMemcpyDeviceToHost, st));
	CK(cudaFreeAsync(d, st));
	CK(cudaStreamSynchronize(st));
	return LBM_OK;
}

int lbm_checksum(lbm_ctx *c, int buf, unsigned long long *sum)
{
	if (!c || !sum || (buf != 0 && buf != 1)) return fail(LBM_ERR_ARG, "lbm_checksum: bad argument");
	if (buf == 1 && c->inplace) return fail(LBM_ERR_STATE, "lbm_checksum: an in-place context has no buffer B");
	return reduce_u64(c, buf, 0, sum);
}

int lbm_check_finite(lbm_ctx *c, unsigned long long *nonfinite)
{
	if (!c) return fail(LBM_ERR_ARG, "lbm_check_finite: null ctx");
	unsigned long long n = 0;
	int rc;
	if ((rc = reduce_u64(c, LBM_BUF_A, 1, &n))) return rc;
	if (nonfinite) *nonfinite = n;
	if (n) return fail(LBM_ERR_FPE, "lbm_check_finite: %llu non-finite populations in buffer A (the reference would have trapped, LB.c:129)", n);
	return LBM_OK;
}

// ---------------------------------------------------------------------------
// tuning / accounting
// ---------------------------------------------------------------------------
int lbm_set_tuning(lbm_ctx *c, const lbm_tuning *t)
{
	if (!c || !t) return fail(LBM_ERR_ARG, "lbm_set_tuning: null argument");
	lbm_tuning n = *t;
	if (n.vec == 0) n.vec = default_vec(c);
	if (n.vec != 1 && n.vec != 2 && n.vec != 4) return fail(LBM_ERR_ARG, "lbm_set_tuning: vec %d", n.vec);
	if (n.hints < 0 || n.hints > 3 || (n.hints > 1 && !c->inplace)) return fail(LBM_ERR_ARG, "lbm_set_tuning: hints %d", n.hints);
	if (n.pdl < 0 || n.pdl > 1) return fail(LBM_ERR_ARG, "lbm_set_tuning: pdl %d", n.pdl);
	if (n.graph < 0 || n.graph > 1000) return fail(LBM_ERR_ARG, "lbm_set_tuning: graph %d", n.graph);
	if (n.tb_chains < 0 || n.tb_chains > LBM_MAX_CHAINS) return fail(LBM_ERR_ARG, "lbm_set_tuning: tb_chains %d", n.tb_chains);
	if (n.tb_chains == 0) n.tb_chains = 2;
	if (n.tb_depth < 0 || n.tb_depth > 1024 || (n.tb_depth & 1)) return fail(LBM_ERR_ARG, "lbm_set_tuning: tb_depth %d (even, 0..1024)", n.tb_depth);
	if (n.tb_rows < 0 || n.tb_cols < 0) return fail(LBM_ERR_ARG, "lbm_set_tuning: tb_rows %d, tb_cols %d", n.tb_rows, n.tb_cols);
	if (n.tb_depth && !c->inplace) return fail(LBM_ERR_STATE, "lbm_set_tuning: tb_depth needs an in-place context (lbm_config.inplace): "
	                                                           "the A/B buffers have write-after-read hazards between bands");
	drop_graph(c);
	c->tune = n;
	return LBM_OK;
}
#ifdef LBM_EXPERIMENTAL
int lbm_set_experimental(lbm_ctx *c, const lbm_experimental *x)
{
	if (!c || !x) return fail(LBM_ERR_ARG, "lbm_set_experimental: null argument");
	if (x->stage_smem < 0 || x->stage_smem > 2 || x->rows_per_thread < 0 || x->rows_per_thread > 1024 || x->smem_pad < 0 || x->smem_pad > 48 * 1024 ||
	    (x->block != 0 && (x->block < 32 || x->block > 512 || (x->block & 31)))) return fail(LBM_ERR_ARG, "lbm_set_experimental: stage_smem %d, rows_per_thread %d", x->stage_smem, x->rows_per_thread);
	drop_graph(c);
	c->exp = *x;
	return LBM_OK;
}
#endif
int lbm_get_tuning(lbm_ctx *c, lbm_tuning *t)
{
	if (!c || !t) return fail(LBM_ERR_ARG, "lbm_get_tuning: null argument");
	*t = c->tune;
	return LBM_OK;
}
long long lbm_kernel_launches(lbm_ctx *c) { return c ? c->launches : 0; }

int lbm_tb_schedule(int nx, int ny, int multi, int tb_depth, int tb_rows, int tb_cols, int *depth_used, int *launches, int cap)
{
	if (nx < 1 || ny < 2 || tb_depth < 0 || tb_rows < 0 || tb_cols < 0 || (cap > 0 && !launches)) { fail(LBM_ERR_ARG, "lbm_tb_schedule: bad argument"); return -1; }
	const int per128 = 128 / (int)sizeof(real_t);
	const int ldy = ((ny + per128 - 1) / per128) * per128;
	static thread_local TbPlan pl;
	if (multi < 0 || multi > 1) { fail(LBM_ERR_ARG, "lbm_tb_schedule: multi %d", multi); return -1; }
	tb_plan(&pl, nx, ny, ldy, ny % VECWIDTH == 0, multi == 1, tb_depth, tb_rows, tb_cols);
	if (depth_used) *depth_used = pl.k;
	if (pl.k < 2) return 0;
	int n = 0;
	tb_for_each_launch(pl, pl.k, [&](int s, int r0, int r1, int j0, int j1, int seam) {
		if (n < cap) { int *o = launches + 6 * n; o[0] = s; o[1] = r0; o[2] = r1; o[3] = j0; o[4] = j1; o[5] = seam; }
		n++;
		return 0;
	});
	return n;
}

/* The launch list of one macro-step with `chains` concurrent tile chains, in enqueue order (which is also a valid sequential
 * order): 7 ints per launch -- level, r0, r1, j0, j1, stream, outer.  stream 0..C-1 = the lanes, C..2C-1 = the seam chains of the
 * boundaries after lane (stream - C); a chain's launches of one column piece start after lanes b and b+1 have finished their
 * tiles of that piece; outer = 1: the boundary between the last lane and the first (periodic wrap / slab seam).
 * Returns the number of launches, *chains_used < 2 when the lattice is too low for lanes (use lbm_tb_schedule then). */
int lbm_tb_schedule_chains(int nx, int ny, int multi, int chains, int tb_depth, int tb_rows, int tb_cols, int *depth_used, int *chains_used, int *launches, int cap)
{
	if (nx < 1 || ny < 2 || tb_depth < 0 || tb_rows < 0 || tb_cols < 0 || chains < 1 || multi < 0 || multi > 1 || (cap > 0 && !launches)) { fail(LBM_ERR_ARG, "lbm_tb_schedule_chains: bad argument"); return -1; }
	const int per128 = 128 / (int)sizeof(real_t);
	const int ldy = ((ny + per128 - 1) / per128) * per128;
	static thread_local TbPlan pl;
	static thread_local LanePlan lp;
	tb_plan(&pl, nx, ny, ldy, ny % VECWIDTH == 0, multi == 1, tb_depth, tb_rows, tb_cols, 1, chains);
	if (depth_used) *depth_used = pl.k;
	const int C = (pl.k >= 2 && ny % VECWIDTH == 0) ? tb_plan_lanes(&lp, nx, ny, ldy, 1, multi == 1, pl.k, tb_rows, tb_cols, chains) : 1;
	if (chains_used) *chains_used = C;
	if (C < 2) return 0;
	int n = 0;
	tb_lanes_walk(lp, lp.k,
		[&](int st, int s, int r0, int r1, int j0, int j1, int outer) {
			if (n < cap) { int *o = launches + 7 * n; o[0] = s; o[1] = r0; o[2] = r1; o[3] = j0; o[4] = j1; o[5] = st; o[6] = outer; }
			n++;
			return 0;
		},
		[&](int, int) { return 0; });
	return n;
}

} // extern "C"

extern "C" {

// ---------------------------------------------------------------------------
// slab decomposition / halo exchange
// ---------------------------------------------------------------------------
int lbm_slab_info_get(lbm_ctx *c, lbm_slab_info *info)
{
	if (!c || !info) return fail(LBM_ERR_ARG, "lbm_slab_info_get: null argument");
	info->nx_local = c->nxl; info->ny = c->ny;
	info->row_lo = c->cfg.row_lo; info->row_hi = c->cfg.row_hi;
	info->ldy = (size_t)c->ldy; info->plane = c->plane;
	info->halo_bytes = (size_t)3 * c->ny * sizeof(real_t);
	info->dev_f[0] = c->f[0]; info->dev_f[1] = c->f[1];
	info->dev_flags = c->flags;
	return LBM_OK;
}

int lbm_ipc_export(lbm_ctx *c, unsigned char blob[LBM_IPC_BLOB_BYTES])
{
	if (!c || !blob) return fail(LBM_ERR_ARG, "lbm_ipc_export: null argument");
	int rc;
	if ((rc = set_device(c))) return rc;
	IpcBlob b;
	memset(&b, 0, sizeof b);
	memcpy(b.magic, "LBMIPC1", 8);
	CK(cudaIpcGetMemHandle(&b.f[0], c->f[0]));
	if (!c->inplace) CK(cudaIpcGetMemHandle(&b.f[1], c->f[1]));
	b.inplace = c->inplace;
	CK(cudaIpcGetMemHandle(&b.flags, c->flags));
	b.nxl = c->nxl; b.ny = c->ny; b.device = c->cfg.device;
	b.uid = (unsigned long long)(uintptr_t)c->f[0] ^ ((unsigned long long)c->cfg.row_lo << 48);
	memset(blob, 0, LBM_IPC_BLOB_BYTES);
	memcpy(blob, &b, sizeof b);
	return LBM_OK;
}

int lbm_ipc_connect(lbm_ctx *c, const unsigned char up[LBM_IPC_BLOB_BYTES], const unsigned char dn[LBM_IPC_BLOB_BYTES])
{
	if (!c || !up || !dn) return fail(LBM_ERR_ARG, "lbm_ipc_connect: null argument");
	int rc;
	if ((rc = set_device(c))) return rc;
	IpcBlob b[2];
	memcpy(&b[0], up, sizeof(IpcBlob));
	memcpy(&b[1], dn, sizeof(IpcBlob));
	Peer peer[2];
	for (int k = 0; k < 2; k++) {
		if (memcmp(b[k].magic, "LBMIPC1", 8)) return fail(LBM_ERR_ARG, "lbm_ipc_connect: bad blob");
		if (b[k].ny != c->ny) return fail(LBM_ERR_ARG, "lbm_ipc_connect: neighbour ny %d != %d", b[k].ny, c->ny);
		if (b[k].inplace != c->inplace) return fail(LBM_ERR_ARG, "lbm_ipc_connect: neighbour and this slab differ in lbm_config.inplace");
		if (k == 1 && !memcmp(&b[0], &b[1], sizeof(IpcBlob))) { peer[1] = peer[0]; break; }   /* P == 2: one neighbour, both sides */
		void *ptr[3] = {nullptr, nullptr, nullptr};
		const cudaIpcMemHandle_t *h[3] = {&b[k].f[0], &b[k].f[1], &b[k].flags};
		for (int m = 0; m < 3; m++) {
			if (m == 1 && c->inplace) continue;
			CK(cudaIpcOpenMemHandle(&ptr[m], *h[m], cudaIpcMemLazyEnablePeerAccess));
			c->ipc_opened[c->n_ipc_opened++] = ptr[m];
		}
		peer[k].f[0] = (real_t *)ptr[0]; peer[k].f[1] = (real_t *)ptr[1];
		peer[k].flags = (int *)ptr[2];
		peer[k].nxl = b[k].nxl;
	}
	c->up = peer[0];
	c->dn = peer[1];
	c->sync_mode = SYNC_FLAGS;
	c->epoch = 0;
	CK(cudaMemsetAsync(c->flags, 0, 4 * sizeof(int), c->stream[0]));
	CK(cudaStreamSynchronize(c->stream[0]));
	return LBM_OK;
}

int lbm_connect_local(lbm_ctx *c, lbm_ctx *up, lbm_ctx *dn)
{
	if (!c || !up || !dn) return fail(LBM_ERR_ARG, "lbm_connect_local: null argument");
	if (up->ny != c->ny || dn->ny != c->ny) return fail(LBM_ERR_ARG, "lbm_connect_local: ny mismatch");
	if (up->inplace != c->inplace || dn->inplace != c->inplace) return fail(LBM_ERR_ARG, "lbm_connect_local: slabs differ in lbm_config.inplace");
	lbm_ctx *nb[2] = {up, dn};
	for (int k = 0; k < 2; k++) {
		if (nb[k]->cfg.device != c->cfg.device) {
			int can = 0;
			CK(cudaDeviceCanAccessPeer(&can, c->cfg.device, nb[k]->cfg.device));
			if (!can) return fail(LBM_ERR_CUDA, "lbm_connect_local: no peer access %d -> %d", c->cfg.device, nb[k]->cfg.device);
			CK(cudaSetDevice(c->cfg.device));
			cudaError_t e = cudaDeviceEnablePeerAccess(nb[k]->cfg.device, 0);
			if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) CK(e);
			cudaGetLastError();
		}
	}
	c->up.f[0] = up->f[0]; c->up.f[1] = up->f[1]; c->up.flags = up->flags; c->up.nxl = up->nxl;
	c->dn.f[0] = dn->f[0]; c->dn.f[1] = dn->f[1]; c->dn.flags = dn->flags; c->dn.nxl = dn->nxl;
	c->sync_mode = SYNC_LOCAL;
	return LBM_OK;
}

/* lbm_group_run for in-place slabs with temporal blocking.  Every slab walks its own launch list (tb_plan, multi):
 * tiles that touch only the slab's own rows on stream 0, the seam chain (inverted row pyramid, or whole-slab
 * launches where the slab is too low) on stream 1.  Chain event (column piece, level) of a slab waits for the
 * previous chain event of both neighbours; events are recorded only after every slab has enqueued its waits on
 * the previous ones (two event objects per slab, alternating), as in the plain loop below. */
static int group_run_wavefront(lbm_ctx **slabs, int n, int nsteps)
{
	int rc;
	std::vector<TbPlan> plans((size_t)n);
	for (int p = 0; p < n; p++) {
		lbm_ctx *c = slabs[p];
		tb_plan(&plans[p], c->nxl, c->ny, c->ldy, c->vw > 1 && c->nyvec == c->ny, 1, c->tune.tb_depth, c->tune.tb_rows, c->tune.tb_cols);
		c->in_wavefront = 1;
	}
	struct Done { lbm_ctx **s; int n; ~Done() { for (int p = 0; p < n; p++) s[p]->in_wavefront = 0; } } done{slabs, n};
	const int k = plans[0].k;
	const TbAxis &cols = plans[0].cols;                          /* identical on every slab (same ny, depth, tb_cols) */
	int ev = 0;                                                  /* chain events so far: ev_bnd[ev & 1] is recorded next */
	for (int t0 = 0; t0 < nsteps; t0 += k) {
		const int kk = nsteps - t0 < k ? nsteps - t0 : k;
		for (int ci = 0; ci < cols.npieces; ci++) {
			for (int p = 0; p < n; p++) {                        /* own tiles of this column piece, all levels */
				lbm_ctx *c = slabs[p];
				const TbAxis &rows = plans[p].rows;
				if ((rc = set_device(c))) return rc;
				for (int ri = 0; ri < rows.npieces; ri++) {
					if (rows.piece[ri].type == TB_INVERTED || rows.piece[ri].type == TB_WHOLE) continue;
					for (int s = 0; s < kk; s++) {
						int r0, r1, j0, j1;
						tb_range(rows, rows.piece[ri], s, &r0, &r1);
						tb_range(cols, cols.piece[ci], s, &j0, &j1);
						if (r1 <= r0 || j1 <= j0) continue;
						if (cols.piece[ci].type == TB_WHOLE) j0 = j1 = 0;
						if ((rc = launch_step(c, (t0 + s) & 1, r0, r1, t0 + s != nsteps - 1, c->stream[0], j0, j1, tb_hint(c, s, kk)))) return rc;
					}
				}
				CK(cudaEventRecord(c->ev_main, c->stream[0]));
				CK(cudaStreamWaitEvent(c->stream[1], c->ev_main, 0));
			}
			for (int s = 0; s < kk; s++) {                       /* the seam chain, level by level, all slabs */
				int j0, j1;
				tb_range(cols, cols.piece[ci], s, &j0, &j1);
				if (j1 <= j0) continue;
				if (cols.piece[ci].type == TB_WHOLE) j0 = j1 = 0;
				const int cur = ev & 1, prev = cur ^ 1;
				for (int p = 0; p < n; p++) {
					lbm_ctx *c = slabs[p], *up = slabs[(p + 1) % n], *dn = slabs[(p + n - 1) % n];
					const TbAxis &rows = plans[p].rows;
					int r0, r1;
					tb_range(rows, rows.piece[rows.npieces - 1], s, &r0, &r1);   /* inverted pyramid or the whole slab */
					if ((rc = set_device(c))) return rc;
					CK(cudaStreamWaitEvent(c->stream[1], up->ev_bnd[prev], 0));
					CK(cudaStreamWaitEvent(c->stream[1], dn->ev_bnd[prev], 0));
					if ((rc = launch_step(c, (t0 + s) & 1, r0, r1, t0 + s != nsteps - 1, c->stream[1], j0, j1, tb_hint(c, s, kk)))) return rc;
				}
				for (int p = 0; p < n; p++) {
					lbm_ctx *c = slabs[p];
					if ((rc = set_device(c))) return rc;
					CK(cudaEventRecord(c->ev_bnd[cur], c->stream[1]));
				}
				ev++;
			}
		}
		for (int p = 0; p < n; p++) {                            /* the next macro-step's tiles need the chains */
			lbm_ctx *c = slabs[p];
			if ((rc = set_device(c))) return rc;
			CK(cudaStreamWaitEvent(c->stream[0], c->ev_bnd[(ev - 1) & 1], 0));
		}
	}
	return LBM_OK;
}

/* All slabs of a ring that live in this process step together.  Per step and slab:
 * boundary rows on stream 1 once the neighbours' previous boundary kernels are done
 * (cross-device event waits), inner rows on stream 0 meanwhile, then join. */
int lbm_group_run(lbm_ctx **slabs, int n, int nsteps)
{
	if (!slabs || n < 1) return fail(LBM_ERR_ARG, "lbm_group_run: bad argument");
	if (nsteps < 0 || (nsteps & 1)) return fail(LBM_ERR_ARG, "lbm_group_run: nsteps %d must be even", nsteps);
	if (nsteps == 0) return LBM_OK;
	int rc;
	for (int k = 0; k < n; k++) {
		lbm_ctx *c = slabs[k];
		if (!c->has_state) return fail(LBM_ERR_STATE, "lbm_group_run: slab %d has no lattice", k);
		if (n > 1 && c->sync_mode != SYNC_LOCAL) return fail(LBM_ERR_STATE, "lbm_group_run: slab %d not wired (lbm_connect_local)", k);
	}
	if (n == 1) return lbm_run(slabs[0], nsteps);
	/* prologue: Source(A) and halo refresh everywhere, recorded as "boundary of step -1" */
	for (int k = 0; k < n; k++) {
		lbm_ctx *c = slabs[k];
		if ((rc = set_device(c))) return rc;
		if ((rc = launch_source(c, LBM_BUF_A, c->stream[0]))) return rc;
		if ((rc = launch_halo_push(c, LBM_BUF_A, c->stream[0]))) return rc;
		CK(cudaEventRecord(c->ev_bnd[1], c->stream[0]));
	}
	/* in-place slabs that all ask for the same temporal blocking: the slab schedule of run_wavefront, with events
	 * between the seam chains instead of device flags */
	bool tb = slabs[0]->inplace && slabs[0]->tune.tb_depth >= 2;
	for (int k = 0; k < n && tb; k++)
		tb = slabs[k]->inplace && slabs[k]->tune.tb_depth == slabs[0]->tune.tb_depth && slabs[k]->tune.tb_cols == slabs[0]->tune.tb_cols &&
		     slabs[k]->ny == slabs[0]->ny && slabs[k]->nyvec == slabs[0]->nyvec && slabs[k]->vw == slabs[0]->vw;
	if (tb) return group_run_wavefront(slabs, n, nsteps);
	for (int t = 0; t < nsteps; t++) {
		const int src = t & 1, tail = t != nsteps - 1, cur = t & 1, prev = cur ^ 1;
		for (int k = 0; k < n; k++) {
			lbm_ctx *c = slabs[k], *up = slabs[(k + 1) % n], *dn = slabs[(k + n - 1) % n];
			if ((rc = set_device(c))) return rc;
			CK(cudaEventRecord(c->ev_main, c->stream[0]));
			CK(cudaStreamWaitEvent(c->stream[1], c->ev_main, 0));
			CK(cudaStreamWaitEvent(c->stream[1], up->ev_bnd[prev], 0));
			CK(cudaStreamWaitEvent(c->stream[1], dn->ev_bnd[prev], 0));
			if ((rc = launch_step(c, src, 0, 1, tail, c->stream[1]))) return rc;
			if (c->nxl > 1 && (rc = launch_step(c, src, c->nxl - 1, c->nxl, tail, c->stream[1]))) return rc;
		}
		/* record after ALL slabs queued their waits on the previous events (an event re-recorded
		 * before a neighbour's wait was enqueued would make that wait target the newer record) */
		for (int k = 0; k < n; k++) {
			lbm_ctx *c = slabs[k];
			if ((rc = set_device(c))) return rc;
			CK(cudaEventRecord(c->ev_bnd[cur], c->stream[1]));
			if ((rc = launch_step(c, src, 1, c->nxl - 1, tail, c->stream[0]))) return rc;
			CK(cudaStreamWaitEvent(c->stream[0], c->ev_bnd[cur], 0));
		}
	}
	return LBM_OK;
}

int lbm_halo_pack(lbm_ctx *c, int buf, void *to_up, void *to_dn)
{
	if (!c || !to_up || !to_dn || (buf != 0 && buf != 1)) return fail(LBM_ERR_ARG, "lbm_halo_pack: bad argument");
	if (c->inplace) return fail(LBM_ERR_STATE, "lbm_halo_pack: the staged transport needs the A/B layout (in-place slabs exchange by peer stores)");
	int rc;
	if ((rc = set_device(c))) return rc;
	halo_pack_kernel<real_t><<<(c->ny + 255) / 256, 256, 0, c->stream[1]>>>(c->f[buf], (real_t *)to_up, (real_t *)to_dn, c->plane, c->ldy, c->nxl, c->ny);
	CK(cudaGetLastError());
	c->launches++;
	return LBM_OK;
}

int lbm_halo_unpack(lbm_ctx *c, int buf, const void *from_dn, const void *from_up)
{
	if (!c || !from_dn || !from_up || (buf != 0 && buf != 1)) return fail(LBM_ERR_ARG, "lbm_halo_unpack: bad argument");
	if (c->inplace) return fail(LBM_ERR_STATE, "lbm_halo_unpack: the staged transport needs the A/B layout (in-place slabs exchange by peer stores)");
	int rc;
	if ((rc = set_device(c))) return rc;
	halo_unpack_kernel<real_t><<<(c->ny + 255) / 256, 256, 0, c->stream[1]>>>(c->f[buf], (const real_t *)from_dn, (const real_t *)from_up, c->plane, c->ldy, c->nxl, c->ny);
	CK(cudaGetLastError());
	c->launches++;
	return LBM_OK;
}

int lbm_step_rows(lbm_ctx *c, int src_buf, int r0, int r1, int tail_source, int stream_id)
{
	if (!c || (src_buf != 0 && src_buf != 1) || (stream_id != 0 && stream_id != 1)) return fail(LBM_ERR_ARG, "lbm_step_rows: bad argument");
	if (r0 < 0 || r1 > c->nxl || r0 > r1) return fail(LBM_ERR_ARG, "lbm_step_rows: rows [%d,%d) of %d", r0, r1, c->nxl);
	int rc;
	if ((rc = set_device(c))) return rc;
	return launch_step(c, src_buf, r0, r1, tail_source, c->stream[stream_id]);
}

int lbm_stream_handle(lbm_ctx *c, int stream_id, void **out)
{
	if (!c || !out || (stream_id != 0 && stream_id != 1)) return fail(LBM_ERR_ARG, "lbm_stream_handle: bad argument");
	*out = (void *)c->stream[stream_id];
	return LBM_OK;
}

int lbm_stream_join(lbm_ctx *c)
{
	if (!c) return fail(LBM_ERR_ARG, "lbm_stream_join: null ctx");
	int rc;
	if ((rc = set_device(c))) return rc;
	CK(cudaEventRecord(c->ev_bnd[0], c->stream[1]));
	CK(cudaStreamWaitEvent(c->stream[0], c->ev_bnd[0], 0));
	CK(cudaEventRecord(c->ev_main, c->stream[0]));
	CK(cudaStreamWaitEvent(c->stream[1], c->ev_main, 0));
	return LBM_OK;
}

} // extern "C"
