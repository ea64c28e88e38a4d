// lbm_kernels_experimental.cuh -- EXPERIMENT build only (-DLBM_EXPERIMENTAL, include/lbm_b200_experimental.h): the in-place
// step with the next row staged in shared memory by cp.async or by TMA bulk copies.  Bit-identical, measured slower than
// the register path (profiles/r02_run1); kept out of the product libraries.
#pragma once
#include "lbm_kernels.cuh"

namespace lbm {

// ---------------------------------------------------------------------------
// In-place step, one site per thread, threads walking rows blockIdx.y, +gridDim.y, ... like PIPE == 1 -- but the nine
// populations of the NEXT row travel global -> shared memory with cp.async (LDGSTS, 4 or 8 bytes per copy, each thread
// its own nine) while the current row is relaxed, instead of into a second register set: the L2 latency is covered
// without the registers that cost PIPE == 1 its occupancy, and without the serial load -> arithmetic -> store of a
// looping thread (capped grids, DESIGN.md "What did NOT help" (2), (8)).  No barrier: a thread reads only what it copied
// itself (cp.async.wait_group).  Safe in place for the reason PIPE == 1 is: a site's nine slots are its own.
// Dynamic shared memory: 2 x 9 x blockDim.x elements.
// STATUS (round 2, profiles/r02_run1): runs, bit-identical, slower than one row per thread at every rows_per_thread.
// ---------------------------------------------------------------------------
template <typename T> __device__ __forceinline__ void cp_async_elem(T *smem_dst, const T *gsrc)
{
	const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
	asm volatile("cp.async.ca.shared.global [%0], [%1], %2;" ::"r"(d), "l"(gsrc), "n"(sizeof(T)) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// same addresses as aa_load<T, 1, PHASE>; dst = this thread's element of stage slot s, slots `stride` elements apart
template <typename T, int PHASE>
__device__ __forceinline__ void aa_stage_issue(const StepParams<T> &p, int r, int j0, T *dst, uint32_t stride)
{
	const T *const own = p.dst + (size_t)(r + 1) * p.ldy;
	if constexpr (PHASE == 2) {
#pragma unroll
		for (int s = 0; s < 9; s++) cp_async_elem<T>(dst + s * stride, own + (size_t)opposite(s) * p.plane + j0);
	} else {
		const int jl = j0 == 0 ? p.ny - 1 : j0 - 1;
		const int jr = j0 + 1 == p.ny ? 0 : j0 + 1;
		const T *lrow[3] = {own - p.ldy, own, own + p.ldy};
#pragma unroll
		for (int s = 0; s < 9; s++)
			cp_async_elem<T>(dst + s * stride, lrow[pull_di(s) + 1] + (size_t)s * p.plane + (pull_dj(s) < 0 ? jl : pull_dj(s) > 0 ? jr : j0));
	}
}

template <typename T, int PHASE>
__global__ void __launch_bounds__(256)
stream_collide_aa_staged_kernel(const __grid_constant__ StepParams<T> p)
{
	extern __shared__ __align__(16) unsigned char aa_stage_raw[];
	T *const stage = reinterpret_cast<T *>(aa_stage_raw);        // [2][9][blockDim.x]
	asm volatile("griddepcontrol.launch_dependents;");
	const uint32_t ncol = (uint32_t)(p.jmax - p.jmin);
	const uint32_t rows = (uint32_t)(p.r1 - p.r0);
	const uint32_t jv = blockIdx.x * blockDim.x + threadIdx.x;
	int j0 = p.jmin + (int)jv;
	if (j0 >= p.ny) j0 -= p.ny;
	asm volatile("griddepcontrol.wait;" ::: "memory");
	uint32_t rr = blockIdx.y;
	if (jv >= ncol || rr >= rows) return;
	const uint32_t st = gridDim.y, bd = blockDim.x;
	T *const mine = stage + threadIdx.x;
	aa_stage_issue<T, PHASE>(p, aa_row(p, rr), j0, mine, bd);
	cp_async_commit();
	uint32_t b = 0;
#pragma unroll 1
	for (; rr < rows; rr += st) {
		const int r = aa_row(p, rr);
		if (rr + st < rows) aa_stage_issue<T, PHASE>(p, aa_row(p, rr + st), j0, mine + (b ^ 1u) * 9u * bd, bd);
		cp_async_commit();                                       // (possibly empty) so that "all but the newest group" is this row's
		cp_async_wait<1>();
		T f[9][1], o[9][1];
#pragma unroll
		for (int s = 0; s < 9; s++) f[s][0] = mine[(b * 9u + (uint32_t)s) * bd];
		relax_sites<T, 1>(p, f, o, r, j0, p.tail_source);
		aa_store<T, 1, PHASE, 0>(p, r, j0, o, 0);
		b ^= 1u;
	}
}

// ---------------------------------------------------------------------------
// Phase 2 with TMA (1-D bulk copies): the phase is local and aligned, so what a CTA needs for a row segment of
// blockDim.x sites is nine contiguous pieces of memory in and nine out.  Thread 0 moves them with cp.async.bulk
// (global -> shared completing on an mbarrier; shared -> global as a bulk group), two stages each way; every thread
// is left with 9 LDS, relax_sites and 9 STS at 32-bit shared addresses -- no global address arithmetic, no load
// latency in any thread.  CTAs walk rows blockIdx.y, +gridDim.y, ...; whole rows only (segments must not wrap).
//   iteration i:  T0: loads of row i+1 -> in[b^1] (read last in iteration i-1, before its barrier)
//                 all: wait full[b]; LDS; relax; STS -> out[b]; fence.proxy.async
//                 T0: wait until the bulk stores of iterations <= i-1 have read their shared source (that of i-1
//                     reads out[b^1], which iteration i+1 overwrites after this iteration's barrier)
//                 barrier;  T0: bulk store out[b] -> global
// Dynamic shared memory: (2 x 9 + 2 x 9) x blockDim.x elements + two mbarriers.
// STATUS (round 2, profiles/r02_run1): runs, bit-identical; phase 2 alone 57.9 GLUPS (DP 400x2000, L2-resident) against
// 75.1 for the register kernel.  lbm_experimental.stage_smem = 2 selects it for phase 2 (phase 1: the cp.async kernel).
// ---------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count)
{
	asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes)
{
	asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity)
{
	uint32_t ok;
	do {
		asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
		             : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
	} while (!ok);
}
__device__ __forceinline__ void bulk_g2s(void *smem_dst, const void *gsrc, uint32_t bytes, uint64_t *bar)
{
	asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
	             ::"r"(smem_u32(smem_dst)), "l"(gsrc), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void bulk_s2g(void *gdst, const void *smem_src, uint32_t bytes)
{
	asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(smem_u32(smem_src)), "r"(bytes) : "memory");
}

template <typename T>
__global__ void __launch_bounds__(128)
stream_collide_aa_bulk2_kernel(const __grid_constant__ StepParams<T> p)
{
	extern __shared__ __align__(128) unsigned char aa_bulk_raw[];
	const uint32_t B = blockDim.x;
	T *const in = reinterpret_cast<T *>(aa_bulk_raw);            // [2][9][B]
	T *const out = in + 2u * 9u * B;                             // [2][9][B]
	uint64_t *const full = reinterpret_cast<uint64_t *>(out + 2u * 9u * B);   // [2]
	asm volatile("griddepcontrol.launch_dependents;");
	const int c0 = (int)(blockIdx.x * B);
	const uint32_t ncols = (uint32_t)min((int)B, p.ny - c0);
	const uint32_t bytes = ncols * (uint32_t)sizeof(T);          // multiple of 16: ny is a multiple of the vector width
	const bool live = threadIdx.x < ncols;
	const int j0 = c0 + (int)threadIdx.x;
	const uint32_t rows = (uint32_t)(p.r1 - p.r0), st = gridDim.y;
	if (threadIdx.x == 0) {
		mbar_init(&full[0], 1);
		mbar_init(&full[1], 1);
		asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
		asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
	}
	__syncthreads();
	asm volatile("griddepcontrol.wait;" ::: "memory");
	uint32_t rr = blockIdx.y;
	if (rr >= rows) return;
	auto issue_loads = [&](int r, uint32_t stage) {              // thread 0 only
		const T *const own = p.dst + (size_t)(r + 1) * p.ldy + c0;
		mbar_expect_tx(&full[stage], 9u * bytes);
#pragma unroll
		for (int s = 0; s < 9; s++) bulk_g2s(in + (stage * 9u + (uint32_t)s) * B, own + (size_t)opposite(s) * p.plane, bytes, &full[stage]);
	};
	if (threadIdx.x == 0) issue_loads(aa_row(p, rr), 0);
	uint32_t b = 0, parity = 0;                                  // bit b of `parity`: phase of full[b] to wait for
#pragma unroll 1
	for (; rr < rows; rr += st) {
		const int r = aa_row(p, rr);
		if (threadIdx.x == 0 && rr + st < rows) issue_loads(aa_row(p, rr + st), b ^ 1u);
		mbar_wait(&full[b], (parity >> b) & 1u);
		parity ^= 1u << b;
		if (live) {
			T f[9][1], o[9][1];
#pragma unroll
			for (int s = 0; s < 9; s++) f[s][0] = in[(b * 9u + (uint32_t)s) * B + threadIdx.x];
			relax_sites<T, 1>(p, f, o, r, j0, p.tail_source);
#pragma unroll
			for (int s = 0; s < 9; s++) out[(b * 9u + (uint32_t)s) * B + threadIdx.x] = o[s][0];
			if (r == 0) {                                        // halo pushes as in aa_store<T, 1, 2>
				p.dn_halo[4 * p.dn_plane + j0] = o[4][0]; p.dn_halo[7 * p.dn_plane + j0] = o[7][0]; p.dn_halo[8 * p.dn_plane + j0] = o[8][0];
			}
			if (r == p.nxl - 1) {
				p.up_halo[2 * p.up_plane + j0] = o[2][0]; p.up_halo[5 * p.up_plane + j0] = o[5][0]; p.up_halo[6 * p.up_plane + j0] = o[6][0];
			}
		}
		asm volatile("fence.proxy.async.shared::cta;" ::: "memory");    // my STS before the bulk store reads them
		if (threadIdx.x == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
		__syncthreads();
		if (threadIdx.x == 0) {
			T *const own = p.dst + (size_t)(r + 1) * p.ldy + c0;
#pragma unroll
			for (int s = 0; s < 9; s++) bulk_s2g(own + (size_t)s * p.plane, out + (b * 9u + (uint32_t)s) * B, bytes);
			asm volatile("cp.async.bulk.commit_group;" ::: "memory");
		}
		b ^= 1u;
	}
	if (threadIdx.x == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}


} // namespace lbm
