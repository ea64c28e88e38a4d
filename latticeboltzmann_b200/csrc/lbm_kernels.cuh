// lbm_kernels.cuh -- sm_100a device code of the D2Q9 BGK timestep path.
//
// What it computes is StreamCollideVec / StreamCollide / ApplySource of the
// reference (src/latticeboltzmann.c = "LB.c", lines cited at each function); how
// it computes it is B200-first:
//   * one fused pull kernel: stream + collide + bounce-back + body force of the
//     NEXT step + halo push, A/B double buffer, SoA f[s][row][j] planes;
//   * each thread owns V consecutive sites of one row (128-bit aligned vector
//     loads/stores along j); the +-1 column neighbours of the six shifted
//     streams come from the adjacent lane by warp shuffle, so every element of
//     f is fetched from memory exactly once per step (144 B DP / 72 B SP per
//     site update);
//   * walls are a packed bitmask (1 bit/site); x-periodicity and the multi-GPU
//     slab decomposition are the same mechanism: every slab carries two halo
//     rows, and the threads of the first/last owned row store the three speeds
//     the neighbour will pull (LB.c:237-243) straight into the neighbour's
//     halo row -- a peer pointer over NVLink, or the slab itself when P == 1;
//   * every floating-point operation is an explicit round-to-nearest intrinsic
//     (__dadd_rn, __fmul_rn, ...): never contracted to FMA, evaluated in the
//     reference's operation order, so DP and SP results are bit-identical to
//     the reference's -O2 AVX build.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace lbm {

// ---------------------------------------------------------------------------
// correctly rounded, never-fused arithmetic in T
// ---------------------------------------------------------------------------
template <typename T> struct fp;
template <> struct fp<double> {
	static __device__ __forceinline__ double add(double a, double b) { return __dadd_rn(a, b); }
	static __device__ __forceinline__ double sub(double a, double b) { return __dsub_rn(a, b); }
	static __device__ __forceinline__ double mul(double a, double b) { return __dmul_rn(a, b); }
	static __device__ __forceinline__ double div(double a, double b) { return __ddiv_rn(a, b); }
};
template <> struct fp<float> {
	static __device__ __forceinline__ float add(float a, float b) { return __fadd_rn(a, b); }
	static __device__ __forceinline__ float sub(float a, float b) { return __fsub_rn(a, b); }
	static __device__ __forceinline__ float mul(float a, float b) { return __fmul_rn(a, b); }
	static __device__ __forceinline__ float div(float a, float b) { return __fdiv_rn(a, b); }
};

// ---------------------------------------------------------------------------
// parameters of one launch
// ---------------------------------------------------------------------------
template <typename T> struct StepParams {
	const T *src;          // source buffer, element (s=0, storage row 0, j=0)
	T *dst;                // destination buffer
	const uint32_t *mask;  // wall bits, [nxl][ldw] words, bit (j&31) of word j>>5
	T *up_halo;            // neighbour above: (s=0, its storage row 0, j=0) of ITS dst buffer
	T *dn_halo;            // neighbour below: (s=0, its storage row nxl'+1, j=0) of ITS dst buffer
	size_t up_plane, dn_plane, plane;   // plane strides (elements)
	int ldy, ldw;          // row strides: elements of T, mask words
	int nxl, ny;           // owned rows, columns
	int jmin, jmax;        // column range of this launch (vector kernel: [0,NYVECMAX))
	int r0, r1;            // owned-row range of this launch
	int vw;                // VECWIDTH of the reference build being matched
	int tail_source;       // apply the NEXT step's ApplySource to column 0 before storing
	T w0, w14, w58;        // OMEGA0/14/58 (LB.c:38-40), formed on the host in T
	T itau;                // REAL_T(1.0)/TAU (LB.c:323)
	double itau_d;         // (1.0/TAU) in double (LB.c:295,456)
	T a14, a58;            // ACCEL*OMEGA14, ACCEL*OMEGA58 (LB.c:502-512)
};

// pull offsets (LB.c:235-243): speed -> (row shift, column shift)
//   s : 0  1  2  3  4  5  6  7  8
//   di: 0  0 -1  0 +1 -1 -1 +1 +1
//   dj: 0 -1  0 +1  0 -1 +1 +1 -1
__device__ __forceinline__ constexpr int pull_di(int s) { return (s == 2 || s == 5 || s == 6) ? -1 : (s == 4 || s == 7 || s == 8) ? 1 : 0; }
__device__ __forceinline__ constexpr int pull_dj(int s) { return (s == 1 || s == 5 || s == 8) ? -1 : (s == 3 || s == 6 || s == 7) ? 1 : 0; }
// bounce-back partner (LB.c:247-254)
__device__ __forceinline__ constexpr int opposite(int s) { return s == 0 ? 0 : (s <= 4 ? ((s + 1) & 3) + 1 : ((s - 3) & 3) + 5); }

// ---------------------------------------------------------------------------
// vector memory access: V elements of T, V*sizeof(T) in {4,8,16,32} bytes
// ---------------------------------------------------------------------------
template <int HINT> __device__ __forceinline__ double ld1(const double *p)
{
	if (HINT == 0) return __ldg(p);
	if (HINT == 2) return __ldcg(p);
	double v;
	asm volatile("ld.global.nc.L1::no_allocate.f64 %0, [%1];" : "=d"(v) : "l"(p));
	return v;
}
template <int HINT> __device__ __forceinline__ float ld1(const float *p)
{
	if (HINT == 0) return __ldg(p);
	if (HINT == 2) return __ldcg(p);
	float v;
	asm volatile("ld.global.nc.L1::no_allocate.f32 %0, [%1];" : "=f"(v) : "l"(p));
	return v;
}
template <int HINT> __device__ __forceinline__ void ld16(const double *p, double &a, double &b)
{
	if (HINT == 0) { const double2 t = __ldg(reinterpret_cast<const double2 *>(p)); a = t.x; b = t.y; }
	else if (HINT == 2) { const double2 t = __ldcg(reinterpret_cast<const double2 *>(p)); a = t.x; b = t.y; }
	else asm volatile("ld.global.nc.L1::no_allocate.v2.f64 {%0,%1}, [%2];" : "=d"(a), "=d"(b) : "l"(p));
}
template <int HINT> __device__ __forceinline__ void ld16(const float *p, float &a, float &b, float &c, float &d)
{
	if (HINT == 0) { const float4 t = __ldg(reinterpret_cast<const float4 *>(p)); a = t.x; b = t.y; c = t.z; d = t.w; }
	else if (HINT == 2) { const float4 t = __ldcg(reinterpret_cast<const float4 *>(p)); a = t.x; b = t.y; c = t.z; d = t.w; }
	else asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(a), "=f"(b), "=f"(c), "=f"(d) : "l"(p));
}
template <int HINT> __device__ __forceinline__ void ld8(const float *p, float &a, float &b)
{
	if (HINT == 0) { const float2 t = __ldg(reinterpret_cast<const float2 *>(p)); a = t.x; b = t.y; }
	else if (HINT == 2) { const float2 t = __ldcg(reinterpret_cast<const float2 *>(p)); a = t.x; b = t.y; }
	else asm volatile("ld.global.nc.L1::no_allocate.v2.f32 {%0,%1}, [%2];" : "=f"(a), "=f"(b) : "l"(p));
}

template <int HINT, int V> __device__ __forceinline__ void ldv(const double *p, double (&v)[V])
{
	if constexpr (V == 1) v[0] = ld1<HINT>(p);
	else {
#pragma unroll
		for (int k = 0; k < V; k += 2) ld16<HINT>(p + k, v[k], v[k + 1]);
	}
}
template <int HINT, int V> __device__ __forceinline__ void ldv(const float *p, float (&v)[V])
{
	if constexpr (V == 1) v[0] = ld1<HINT>(p);
	else if constexpr (V == 2) ld8<HINT>(p, v[0], v[1]);
	else {
#pragma unroll
		for (int k = 0; k < V; k += 4) ld16<HINT>(p + k, v[k], v[k + 1], v[k + 2], v[k + 3]);
	}
}

template <int HINT> __device__ __forceinline__ void st1(double *p, double v)
{
	if (HINT != 1) *p = v; else __stcs(p, v);
}
template <int HINT> __device__ __forceinline__ void st1(float *p, float v)
{
	if (HINT != 1) *p = v; else __stcs(p, v);
}
template <int HINT, int V> __device__ __forceinline__ void stv(double *p, const double (&v)[V])
{
	if constexpr (V == 1) st1<HINT>(p, v[0]);
	else {
#pragma unroll
		for (int k = 0; k < V; k += 2) {
			const double2 t = make_double2(v[k], v[k + 1]);
			if (HINT != 1) *reinterpret_cast<double2 *>(p + k) = t; else __stcs(reinterpret_cast<double2 *>(p + k), t);
		}
	}
}
template <int HINT, int V> __device__ __forceinline__ void stv(float *p, const float (&v)[V])
{
	if constexpr (V == 1) st1<HINT>(p, v[0]);
	else if constexpr (V == 2) {
		const float2 t = make_float2(v[0], v[1]);
		if (HINT != 1) *reinterpret_cast<float2 *>(p) = t; else __stcs(reinterpret_cast<float2 *>(p), t);
	} else {
#pragma unroll
		for (int k = 0; k < V; k += 4) {
			const float4 t = make_float4(v[k], v[k + 1], v[k + 2], v[k + 3]);
			if (HINT != 1) *reinterpret_cast<float4 *>(p + k) = t; else __stcs(reinterpret_cast<float4 *>(p + k), t);
		}
	}
}

// ---------------------------------------------------------------------------
// per-site arithmetic
// ---------------------------------------------------------------------------

// relaxation of the scalar statements LB.c:295 and 456-464: (feq-f) in T, the
// product and the sum in double, rounded to T on store.
template <typename T> __device__ __forceinline__ T relax_double(T f, T feq, double itau_d)
{
	const T d = fp<T>::sub(feq, f);
	return (T)__dadd_rn((double)f, __dmul_rn(itau_d, (double)d));
}

// Fluid site, operation order of the SIMD path LB.c:360-451.
// dbl != 0 selects the mixed-group fallback LB.c:455-465 (identical bits in DP).
template <typename T>
__device__ __forceinline__ void collide_vec_order(const T (&f)[9], T (&o)[9], const StepParams<T> &p, bool dbl)
{
	typedef fp<T> A;
	T rho = A::add(T(0), f[0]);                                  // LB.c:360,372
#pragma unroll
	for (int s = 1; s < 9; s++) rho = A::add(rho, f[s]);         // LB.c:373-380
	T ux = A::add(f[6], f[2]);                                   // LB.c:382-387
	ux = A::add(ux, f[5]); ux = A::sub(ux, f[7]); ux = A::sub(ux, f[4]); ux = A::sub(ux, f[8]);
	ux = A::div(ux, rho);
	T uy = A::add(f[5], f[1]);                                   // LB.c:389-394
	uy = A::add(uy, f[8]); uy = A::sub(uy, f[6]); uy = A::sub(uy, f[3]); uy = A::sub(uy, f[7]);
	uy = A::div(uy, rho);
	// CSQ == 1 (LB.c:49): _ICSQ multiplications are exact identities and are dropped
	const T usq = A::mul(T(1.5), A::add(A::mul(ux, ux), A::mul(uy, uy)));   // LB.c:397
	ux = A::mul(ux, T(3));                                       // LB.c:400-401
	uy = A::mul(uy, T(3));
	const T sum = A::add(ux, uy), dif = A::sub(ux, uy);
	T e[9];
	e[0] = T(0);
	e[1] = uy;      e[2] = ux;                                   // LB.c:402-409; (-1)*x == -x exactly
	e[3] = -uy;     e[4] = -ux;
	e[5] = sum;     e[6] = dif;
	e[7] = -sum;    e[8] = -dif;
	T q[9];
	q[0] = A::sub(T(1), usq);                                    // LB.c:413
#pragma unroll
	for (int s = 1; s < 9; s++)                                  // LB.c:414-421
		q[s] = A::add(T(1), A::add(e[s], A::sub(A::mul(T(0.5), A::mul(e[s], e[s])), usq)));
	const T r0 = A::mul(p.w0, rho), r14 = A::mul(p.w14, rho), r58 = A::mul(p.w58, rho);   // LB.c:425-433
#pragma unroll
	for (int s = 0; s < 9; s++) q[s] = A::mul(q[s], s == 0 ? r0 : (s <= 4 ? r14 : r58));
	// Mixed group (LB.c:455-465): fluid lanes relax with 1/tau in double; only SP can tell the
	// difference.  ONE branch around all nine fp64 relaxations: groups with a wall are rare, and
	// nine predicated-off cvt/dmul/dadd/cvt sequences per site cost issue slots the L2-resident
	// (in-place, temporally blocked) runs are bound by.
	if (sizeof(T) == 4 && dbl) {
#pragma unroll
		for (int s = 0; s < 9; s++) o[s] = relax_double<T>(f[s], q[s], p.itau_d);             // LB.c:456-464
	} else {
		const T itau = p.itau;
#pragma unroll
		for (int s = 0; s < 9; s++) o[s] = A::add(f[s], A::mul(itau, A::sub(q[s], f[s])));    // LB.c:442-450
	}
}

// Fluid site, rounding order of the scalar function StreamCollide LB.c:257-297.
template <typename T>
__device__ __forceinline__ void collide_scalar_order(const T (&f)[9], T (&o)[9], const StepParams<T> &p)
{
	typedef fp<T> A;
	T rho = A::add(T(0), f[0]);
#pragma unroll
	for (int s = 1; s < 9; s++) rho = A::add(rho, f[s]);         // LB.c:258-261
	const T ux = A::div(A::sub(A::add(A::add(f[6], f[2]), f[5]), A::add(A::add(f[7], f[4]), f[8])), rho);   // LB.c:263-264
	const T uy = A::div(A::sub(A::add(A::add(f[5], f[1]), f[8]), A::add(A::add(f[6], f[3]), f[7])), rho);   // LB.c:265-266
	const T usq = A::add(A::mul(ux, ux), A::mul(uy, uy));        // LB.c:268
	T e[9];                                                      // LB.c:271-279
	e[0] = T(0);
	e[1] = uy;                e[2] = ux;
	e[3] = -uy;               e[4] = -ux;
	e[5] = A::add(ux, uy);    e[6] = A::sub(ux, uy);
	e[7] = A::sub(-ux, uy);   e[8] = A::add(-ux, uy);
	const T t3 = A::mul(T(1.5), usq);                            // REAL_T(3.0)/REAL_T(2.0)*uDotu (/CSQ exact)
	const T r0 = A::mul(p.w0, rho), r14 = A::mul(p.w14, rho), r58 = A::mul(p.w58, rho);
#pragma unroll
	for (int s = 0; s < 9; s++) {
		T g;
		if (s == 0) g = A::sub(T(1), t3);                        // LB.c:283
		else g = A::sub(A::add(A::add(T(1), A::mul(T(3), e[s])), A::mul(A::mul(T(4.5), e[s]), e[s])), t3);   // LB.c:284-291
		const T feq = A::mul(s == 0 ? r0 : (s <= 4 ? r14 : r58), g);
		o[s] = relax_double<T>(f[s], feq, p.itau_d);             // LB.c:294-296
	}
}

// wall site: reflect (LB.c:246-255, 466-475); o[0] is never stored
template <typename T> __device__ __forceinline__ void bounce(const T (&f)[9], T (&o)[9])
{
	o[0] = f[0];
#pragma unroll
	for (int s = 1; s < 9; s++) o[s] = f[opposite(s)];
}

// ApplySource (LB.c:489-518) on the nine values of one fluid site of column 0
template <typename T> __device__ __forceinline__ void source_site(T (&o)[9], const StepParams<T> &p)
{
	typedef fp<T> A;
	const T n6 = A::sub(o[6], p.a58), n3 = A::sub(o[3], p.a14), n7 = A::sub(o[7], p.a58);
	if (n6 > T(0) && n3 > T(0) && n7 > T(0)) {
		o[6] = n6; o[3] = n3; o[7] = n7;
		o[5] = A::add(o[5], p.a58); o[1] = A::add(o[1], p.a14); o[8] = A::add(o[8], p.a58);
	}
}

// ---------------------------------------------------------------------------
// wall bits + collide/bounce + tail source for the V sites (r, j0..j0+V-1) whose pulled
// populations are f[s][k]; returns the wall bits of the V sites (LB.c:437-475)
// ---------------------------------------------------------------------------
template <typename T, int V>
__device__ __forceinline__ uint32_t relax_sites(const StepParams<T> &p, const T (&f)[9][V], T (&o)[9][V], int r, int j0, int tail_source)
{
	// wall bits of my V sites and of my aligned VECWIDTH group (LB.c:437-440)
	const uint32_t word = __ldg(p.mask + (size_t)r * p.ldw + (j0 >> 5));
	const uint32_t sh = (uint32_t)j0 & 31u;
	const uint32_t bits = (word >> sh) & ((1u << V) - 1u);
	bool dbl = false;
	if (sizeof(T) == 4) dbl = ((word >> (sh & ~(uint32_t)(p.vw - 1))) & ((1u << p.vw) - 1u)) != 0u;
#pragma unroll
	for (int k = 0; k < V; k++) {
		T fk[9], ok[9];
#pragma unroll
		for (int s = 0; s < 9; s++) fk[s] = f[s][k];
		if ((bits >> k) & 1u) {
			bounce<T>(fk, ok);
		} else {
			collide_vec_order<T>(fk, ok, p, dbl);
			if (k == 0 && j0 == 0 && tail_source) source_site<T>(ok, p);
		}
#pragma unroll
		for (int s = 0; s < 9; s++) o[s][k] = ok[s];
	}
	return bits;
}

// ---------------------------------------------------------------------------
// the fused pull update of one V-wide vector of sites (StreamCollideVec, LB.c:307-484,
// + tail ApplySource + halo push)
//   V     sites per thread (contiguous in j, V*sizeof(T) <= 32 bytes)
//   NB    0: +-1 column element from the adjacent lane (shuffle), 1: direct load
//   HINT  0: default caching, 1: L1::no_allocate loads + .cs stores, 2: L2-coherent (.cg) loads
//   WRAP  true: rows wrap inside this slab (no halo rows read or written) -- the persistent
//         multi-step kernel; false: rows i-1/i+1 come from the halo rows, boundary rows push
// All 32 lanes of a warp must call this together (shuffles); `active` predicates the stores.
// ---------------------------------------------------------------------------
template <typename T, int V, int NB, int HINT, bool WRAP>
__device__ __forceinline__ void update_vector(const StepParams<T> &p, const T *__restrict__ src, T *__restrict__ dst,
                                              uint32_t nvec, uint32_t g, bool active, uint32_t lane, int tail_source)
{
	const uint32_t rr = g / nvec;
	const uint32_t jv = g - rr * nvec;
	const int r = p.r0 + (int)rr;                            // owned row; storage row r+1
	const int j0 = p.jmin + (int)jv * V;

	// wrapped columns of the two out-of-vector neighbours (LB.c:337-338)
	const int jl = j0 == 0 ? p.ny - 1 : j0 - 1;
	const int jr = j0 + V == p.ny ? 0 : j0 + V;
	const bool edge_l = (NB == 1) || V == 1 || lane == 0 || jv == 0;
	const bool edge_r = (NB == 1) || V == 1 || lane == 31 || jv == nvec - 1;

	const T *row[3];                                         // rows i-1, i, i+1 of the source
	row[1] = src + (size_t)(r + 1) * p.ldy;
	if (WRAP) {                                              // x_d, x_u of LB.c:335-336
		row[0] = src + (size_t)(r == 0 ? p.nxl : r) * p.ldy;
		row[2] = src + (size_t)(r == p.nxl - 1 ? 1 : r + 2) * p.ldy;
	} else {                                                 // halo rows make the wrap / the slab seam
		row[0] = row[1] - p.ldy;
		row[2] = row[1] + p.ldy;
	}

	T f[9][V];
#pragma unroll
	for (int s = 0; s < 9; s++) {
		const T *base_s = row[pull_di(s) + 1] + (size_t)s * p.plane;
		if constexpr (V == 1) {
			const int jj = pull_dj(s) < 0 ? jl : pull_dj(s) > 0 ? jr : j0;
			f[s][0] = ld1<HINT>(base_s + jj);
		} else {
			T own[V];
			ldv<HINT, V>(base_s + j0, own);
			if (pull_dj(s) == 0) {
#pragma unroll
				for (int k = 0; k < V; k++) f[s][k] = own[k];
			} else if (pull_dj(s) < 0) {                     // element k comes from column j0+k-1
				T nb = T(0);
				if (NB == 0) nb = __shfl_up_sync(0xffffffffu, own[V - 1], 1);
				if (edge_l) nb = ld1<HINT>(base_s + jl);
				f[s][0] = nb;
#pragma unroll
				for (int k = 1; k < V; k++) f[s][k] = own[k - 1];
			} else {                                         // element k comes from column j0+k+1
				T nb = T(0);
				if (NB == 0) nb = __shfl_down_sync(0xffffffffu, own[0], 1);
				if (edge_r) nb = ld1<HINT>(base_s + jr);
				f[s][V - 1] = nb;
#pragma unroll
				for (int k = 0; k < V - 1; k++) f[s][k] = own[k + 1];
			}
		}
	}

	T o[9][V];
	const uint32_t bits = relax_sites<T, V>(p, f, o, r, j0, tail_source);

	if (active) {
		T *d = dst + (size_t)(r + 1) * p.ldy + j0;
		if (bits == 0u) stv<HINT, V>(d, o[0]);
		else {
#pragma unroll
			for (int k = 0; k < V; k++) if (!((bits >> k) & 1u)) d[k] = o[0][k];     // f0 untouched at walls
		}
#pragma unroll
		for (int s = 1; s < 9; s++) stv<HINT, V>(d + (size_t)s * p.plane, o[s]);
		if (!WRAP) {
			if (r == 0) {                                    // what the slab below pulls from its row "hi" (LB.c:239,242,243)
				stv<HINT, V>(p.dn_halo + 4 * p.dn_plane + j0, o[4]);
				stv<HINT, V>(p.dn_halo + 7 * p.dn_plane + j0, o[7]);
				stv<HINT, V>(p.dn_halo + 8 * p.dn_plane + j0, o[8]);
			}
			if (r == p.nxl - 1) {                            // what the slab above pulls from its row "lo-1" (LB.c:237,240,241)
				stv<HINT, V>(p.up_halo + 2 * p.up_plane + j0, o[2]);
				stv<HINT, V>(p.up_halo + 5 * p.up_plane + j0, o[5]);
				stv<HINT, V>(p.up_halo + 6 * p.up_plane + j0, o[6]);
			}
		}
	}
}

// One lattice step per launch.  1-D grid, blockDim multiple of 32; grid-stride over (row, vector) pairs.
template <typename T, int V, int NB, int HINT>
__global__ void __launch_bounds__(512)
stream_collide_vec_kernel(const __grid_constant__ StepParams<T> p)
{
	// Programmatic dependent launch (no-ops unless launched with the PDL attribute): let the NEXT step's
	// CTAs become resident while this step drains, and do not touch f before the PREVIOUS step is complete.
	asm volatile("griddepcontrol.launch_dependents;");
	const uint32_t nvec = (uint32_t)(p.jmax - p.jmin) / V;
	const uint32_t total = (uint32_t)(p.r1 - p.r0) * nvec;
	const uint32_t lane = threadIdx.x & 31u;
	const uint32_t stride = gridDim.x * blockDim.x;
	asm volatile("griddepcontrol.wait;" ::: "memory");
	for (uint32_t base = blockIdx.x * blockDim.x; base < total; base += stride) {   // warp-uniform trip count
		const uint32_t g0 = base + threadIdx.x;
		const bool active = g0 < total;
		const uint32_t g = active ? g0 : total - 1;          // idle lanes shadow the last vector (loads stay legal)
		update_vector<T, V, NB, HINT, false>(p, p.src, p.dst, nvec, g, active, lane, p.tail_source);
	}
}

// ---------------------------------------------------------------------------
// In-place streaming ("AA pattern": ONE buffer instead of the A/B pair of LB.c:186-210).
// Every site reads nine slots and stores its nine results into the SAME nine slots, so no
// other site's data is ever overwritten and no second buffer is needed.  Steps alternate:
//   PHASE 1 (steps 1,3,5,... of a run; memory holds the reference layout, slot s of site x =
//            post-collide f_s of x): site x pulls f_s from slot s of (x - e_s) exactly like the
//            A/B kernel (LB.c:235-243) and stores its post-collide o_s into the slot it read
//            f_opp(s) from, i.e. slot opp(s) of site (x + e_s).
//   PHASE 2 (steps 2,4,6,...): site x finds what the pull of the next step would fetch, f_s, in
//            slot opp(s) of x itself; it stores o_s into slot s of x -- the reference layout again,
//            so after every DoTimeStep (two steps, LB.c:186-210) buffer A is bit-identical to
//            the reference's fA.  Purely local: no neighbour rows, perfectly aligned accesses.
// Wall sites write back exactly what they read (bounce-back, LB.c:246-255, maps slot to itself),
// which also keeps the stale f0 of wall sites.  The arithmetic is relax_sites(), shared with the
// A/B kernel.  Loads are coherent (never ld.global.nc): the buffer is written by this kernel.
//
// Slab seams: phase 1 reads rows i-1 / i+1 of boundary rows from the halo rows (filled by the
// neighbours' phase-2 push) and stores straight into the NEIGHBOUR'S OWNED boundary row (peer
// memory, or this slab itself when P == 1); phase 2 pushes the three halo speeds like the A/B
// kernel does.  Same bytes over NVLink per two steps as A/B.
// ---------------------------------------------------------------------------
// CG: ld.global.cg (L2 only) -- for the dataflow kernel, whose CTAs read what other SMs wrote during the same launch
template <bool CG, typename X> __device__ __forceinline__ X ld_coh(const X *p)
{
	if constexpr (CG) return __ldcg(p);
	else return *p;
}
template <int V, bool CG = false> __device__ __forceinline__ void ldc(const double *p, double (&v)[V])
{
	if constexpr (V == 1) v[0] = ld_coh<CG>(p);
	else {
#pragma unroll
		for (int k = 0; k < V; k += 2) { const double2 t = ld_coh<CG>(reinterpret_cast<const double2 *>(p + k)); v[k] = t.x; v[k + 1] = t.y; }
	}
}
template <int V, bool CG = false> __device__ __forceinline__ void ldc(const float *p, float (&v)[V])
{
	if constexpr (V == 1) v[0] = ld_coh<CG>(p);
	else if constexpr (V == 2) { const float2 t = ld_coh<CG>(reinterpret_cast<const float2 *>(p)); v[0] = t.x; v[1] = t.y; }
	else {
#pragma unroll
		for (int k = 0; k < V; k += 4) { const float4 t = ld_coh<CG>(reinterpret_cast<const float4 *>(p + k)); v[k] = t.x; v[k + 1] = t.y; v[k + 2] = t.z; v[k + 3] = t.w; }
	}
}
// V values to columns (j0-1 .. j0+V-2) [DJ < 0, first one wraps to jl] or (j0+1 .. j0+V) [DJ > 0, last one wraps to jr]
template <typename T, int V, int DJ> __device__ __forceinline__ void st_shifted(T *row, int j0, int jl, int jr, const T (&v)[V])
{
	if constexpr (DJ < 0) {
		row[jl] = v[0];
		if constexpr (V == 2) row[j0] = v[1];
		if constexpr (V >= 4) {
#pragma unroll
			for (int k = 1; k + 1 < V; k += 2) {            // pairs (1,2),(3,4),... start at the even column j0+k-1
				T pr[2] = {v[k], v[k + 1]};
				stv<0, 2>(row + j0 + k - 1, pr);
			}
			row[j0 + V - 2] = v[V - 1];
		}
	} else {
		row[jr] = v[V - 1];
		if constexpr (V == 2) row[j0 + 1] = v[0];
		if constexpr (V >= 4) {
			row[j0 + 1] = v[0];
#pragma unroll
			for (int k = 1; k + 1 < V; k += 2) {            // pairs (1,2),(3,4),... land on the even column j0+k+1
				T pr[2] = {v[k], v[k + 1]};
				stv<0, 2>(row + j0 + k + 1, pr);
			}
		}
	}
}

// the nine populations the V sites (r, j0..j0+V-1) relax this step
template <typename T, int V, int PHASE, bool CG = false>
__device__ __forceinline__ void aa_load(const StepParams<T> &p, int r, int j0, T (&f)[9][V])
{
	const T *const own = p.dst + (size_t)(r + 1) * p.ldy;    // p.src == p.dst
	if constexpr (PHASE == 2) {
#pragma unroll
		for (int s = 0; s < 9; s++) ldc<V, CG>(own + (size_t)opposite(s) * p.plane + j0, f[s]);
	} else {
		const int jl = j0 == 0 ? p.ny - 1 : j0 - 1;          // LB.c:337-338
		const int jr = j0 + V == p.ny ? 0 : j0 + V;
		const T *lrow[3] = {own - p.ldy, own, own + p.ldy};  // rows i-1 (halo when r == 0), i, i+1 (halo when r == nxl-1)
#pragma unroll
		for (int s = 0; s < 9; s++) {
			const T *b = lrow[pull_di(s) + 1] + (size_t)s * p.plane;
			if constexpr (V == 1) {
				f[s][0] = ld_coh<CG>(b + (pull_dj(s) < 0 ? jl : pull_dj(s) > 0 ? jr : j0));
			} else {
				T al[V];
				ldc<V, CG>(b + j0, al);
				if (pull_dj(s) == 0) {
#pragma unroll
					for (int k = 0; k < V; k++) f[s][k] = al[k];
				} else if (pull_dj(s) < 0) {
					f[s][0] = ld_coh<CG>(b + jl);
#pragma unroll
					for (int k = 1; k < V; k++) f[s][k] = al[k - 1];
				} else {
					f[s][V - 1] = ld_coh<CG>(b + jr);
#pragma unroll
					for (int k = 0; k < V - 1; k++) f[s][k] = al[k + 1];
				}
			}
		}
	}
}

template <typename T, int V, int PHASE>
__device__ __forceinline__ void aa_store(const StepParams<T> &p, int r, int j0, const T (&o)[9][V])
{
	T *const own = p.dst + (size_t)(r + 1) * p.ldy;
	if constexpr (PHASE == 2) {
#pragma unroll
		for (int s = 0; s < 9; s++) stv<0, V>(own + (size_t)s * p.plane + j0, o[s]);
		if (r == 0) {                                        // as in the A/B kernel (LB.c:239,242,243)
			stv<0, V>(p.dn_halo + 4 * p.dn_plane + j0, o[4]);
			stv<0, V>(p.dn_halo + 7 * p.dn_plane + j0, o[7]);
			stv<0, V>(p.dn_halo + 8 * p.dn_plane + j0, o[8]);
		}
		if (r == p.nxl - 1) {                                // (LB.c:237,240,241)
			stv<0, V>(p.up_halo + 2 * p.up_plane + j0, o[2]);
			stv<0, V>(p.up_halo + 5 * p.up_plane + j0, o[5]);
			stv<0, V>(p.up_halo + 6 * p.up_plane + j0, o[6]);
		}
	} else {
		const int jl = j0 == 0 ? p.ny - 1 : j0 - 1;
		const int jr = j0 + V == p.ny ? 0 : j0 + V;
		// slot k of site (x - e_k) receives o[opp(k)]; rows beyond the slab are the neighbour's owned boundary rows
		T *srow[3];
		size_t spl[3];
		srow[1] = own;              spl[1] = p.plane;
		if (r == 0) { srow[0] = p.dn_halo - p.ldy; spl[0] = p.dn_plane; } else { srow[0] = own - p.ldy; spl[0] = p.plane; }
		if (r == p.nxl - 1) { srow[2] = p.up_halo + p.ldy; spl[2] = p.up_plane; } else { srow[2] = own + p.ldy; spl[2] = p.plane; }
#pragma unroll
		for (int k = 0; k < 9; k++) {
			T *b = srow[pull_di(k) + 1] + (size_t)k * spl[pull_di(k) + 1];
			if (pull_dj(k) == 0) stv<0, V>(b + j0, o[opposite(k)]);
			else if (pull_dj(k) < 0) st_shifted<T, V, -1>(b, j0, jl, jr, o[opposite(k)]);
			else st_shifted<T, V, 1>(b, j0, jl, jr, o[opposite(k)]);
		}
	}
}

// ---------------------------------------------------------------------------
// Phase 1 with every global access an ALIGNED vector (V > 1): the element a shifted stream needs from column
// j0-1 / j0+V is the adjacent lane's aligned load (warp shuffle), and the element a thread owes to the
// adjacent lane's aligned vector travels the same way before the store; only the lanes at a warp edge or at
// an end of the row's column range fall back to one scalar access.  Safe in place: the shuffled element
// belongs to the receiving thread's own site set, and a warp issues its loads before its stores.
// All 32 lanes must call these (idle lanes shadow the last vector of the range and do not store).
// ---------------------------------------------------------------------------
template <typename T, int V>
__device__ __forceinline__ void aa_load_shf(const StepParams<T> &p, int r, int j0, bool edge_l, bool edge_r, T (&f)[9][V])
{
	const T *const own = p.dst + (size_t)(r + 1) * p.ldy;
	const int jl = j0 == 0 ? p.ny - 1 : j0 - 1;
	const int jr = j0 + V == p.ny ? 0 : j0 + V;
	const T *lrow[3] = {own - p.ldy, own, own + p.ldy};
#pragma unroll
	for (int s = 0; s < 9; s++) {
		const T *b = lrow[pull_di(s) + 1] + (size_t)s * p.plane;
		T al[V];
		ldc<V>(b + j0, al);
		if (pull_dj(s) == 0) {
#pragma unroll
			for (int k = 0; k < V; k++) f[s][k] = al[k];
		} else if (pull_dj(s) < 0) {
			T nb = __shfl_up_sync(0xffffffffu, al[V - 1], 1);
			if (edge_l) nb = b[jl];
			f[s][0] = nb;
#pragma unroll
			for (int k = 1; k < V; k++) f[s][k] = al[k - 1];
		} else {
			T nb = __shfl_down_sync(0xffffffffu, al[0], 1);
			if (edge_r) nb = b[jr];
			f[s][V - 1] = nb;
#pragma unroll
			for (int k = 0; k < V - 1; k++) f[s][k] = al[k + 1];
		}
	}
}

template <typename T, int V>
__device__ __forceinline__ void aa_store_shf(const StepParams<T> &p, int r, int j0, bool edge_l, bool edge_r, bool active, const T (&o)[9][V])
{
	T *const own = p.dst + (size_t)(r + 1) * p.ldy;
	const int jl = j0 == 0 ? p.ny - 1 : j0 - 1;
	const int jr = j0 + V == p.ny ? 0 : j0 + V;
	T *srow[3];
	size_t spl[3];
	srow[1] = own;              spl[1] = p.plane;
	if (r == 0) { srow[0] = p.dn_halo - p.ldy; spl[0] = p.dn_plane; } else { srow[0] = own - p.ldy; spl[0] = p.plane; }
	if (r == p.nxl - 1) { srow[2] = p.up_halo + p.ldy; spl[2] = p.up_plane; } else { srow[2] = own + p.ldy; spl[2] = p.plane; }
#pragma unroll
	for (int k = 0; k < 9; k++) {
		T *b = srow[pull_di(k) + 1] + (size_t)k * spl[pull_di(k) + 1];
		const T (&v)[V] = o[opposite(k)];
		if (pull_dj(k) == 0) {
			if (active) stv<0, V>(b + j0, v);
		} else if (pull_dj(k) < 0) {                         // my values belong at columns j0-1 .. j0+V-2
			const T recv = __shfl_down_sync(0xffffffffu, v[0], 1);   // the right lane's first value: my vector's last column
			if (active) {
				if (!edge_r) {
					T w[V];
#pragma unroll
					for (int m = 0; m < V - 1; m++) w[m] = v[m + 1];
					w[V - 1] = recv;
					stv<0, V>(b + j0, w);
				} else {
#pragma unroll
					for (int m = 1; m < V; m++) b[j0 + m - 1] = v[m];
				}
				if (edge_l) b[jl] = v[0];
			}
		} else {                                             // my values belong at columns j0+1 .. j0+V
			const T recv = __shfl_up_sync(0xffffffffu, v[V - 1], 1);  // the left lane's last value: my vector's first column
			if (active) {
				if (!edge_l) {
					T w[V];
					w[0] = recv;
#pragma unroll
					for (int m = 1; m < V; m++) w[m] = v[m - 1];
					stv<0, V>(b + j0, w);
				} else {
#pragma unroll
					for (int m = 0; m < V - 1; m++) b[j0 + m + 1] = v[m];
				}
				if (edge_r) b[jr] = v[V - 1];
			}
		}
	}
}

// row ranges of the wavefront schedule may wrap (r1 > nxl)
template <typename T> __device__ __forceinline__ int aa_row(const StepParams<T> &p, uint32_t rr)
{
	const int r = p.r0 + (int)rr;
	return r >= p.nxl ? r - p.nxl : r;
}

// 2-D grid: x over the vectors of a row, y over rows -- no integer division per site.
// PIPE == 0: one row per thread and pass (grid-stride over y).
// PIPE == 1: each thread walks rows blockIdx.y, +gridDim.y, ... and loads the populations of its NEXT
//            row before it relaxes the current one (two register sets, ping-pong), so the L2 latency
//            of the loads is covered by ~300 instructions of arithmetic.  For the L2-resident runs
//            (in-place small lattices, temporal blocking) that are bound by load latency at the
//            occupancy 60-120 registers allow.  Safe in place: a site's nine slots are its own.
#ifdef LBM_AA_MINB      /* experiment builds (tools/): 128-thread CTAs, at least LBM_AA_MINB of them per SM */
#define LBM_AA_BOUNDS __launch_bounds__(128, LBM_AA_MINB)
#else
// SP: CTAs of at most 256 threads, four of them per SM -> 64 registers.  Measured on the L2-resident 400x4000 lattice
// (profiles/r01_run24/occ_probe_sp.log): phase 1 at vec 4 drops from 104 registers / 22.4 us to 64 / 17.7 us, the
// run from 113.2 to 119.7 GLUPS.  DP (54-58 registers at vec 1) gains nothing from a cap (profiles/r01_run12).
#define LBM_AA_BOUNDS __launch_bounds__(sizeof(T) == 4 ? 256 : 512, sizeof(T) == 4 ? 4 : 0)
#endif
template <typename T, int V, int PHASE, int PIPE>
__global__ void LBM_AA_BOUNDS
stream_collide_aa_kernel(const __grid_constant__ StepParams<T> p)
{
	asm volatile("griddepcontrol.launch_dependents;");
	const uint32_t nvec = (uint32_t)(p.jmax - p.jmin) / V;
	const uint32_t rows = (uint32_t)(p.r1 - p.r0);
	const uint32_t jv = blockIdx.x * blockDim.x + threadIdx.x;
	int j0 = p.jmin + (int)jv * V;
	if (j0 >= p.ny) j0 -= p.ny;                              // column ranges of the 2-D wavefront schedule may wrap (jmax > ny)
	asm volatile("griddepcontrol.wait;" ::: "memory");
	if constexpr (PIPE == 2 && PHASE == 1 && V > 1) {            // PIPE == 2: the aligned-access (warp shuffle) phase 1
		const bool active = jv < nvec;
		const uint32_t jc = active ? jv : nvec - 1;              // idle lanes shadow the last vector (loads stay legal)
		int jj = p.jmin + (int)jc * V;
		if (jj >= p.ny) jj -= p.ny;
		const uint32_t lane = threadIdx.x & 31u;
		const bool edge_l = lane == 0 || jc == 0, edge_r = lane == 31 || jc >= nvec - 1;
		for (uint32_t rr = blockIdx.y; rr < rows; rr += gridDim.y) {
			const int r = aa_row(p, rr);
			T f[9][V], o[9][V];
			aa_load_shf<T, V>(p, r, jj, edge_l, edge_r, f);
			relax_sites<T, V>(p, f, o, r, jj, p.tail_source);
			aa_store_shf<T, V>(p, r, jj, edge_l, edge_r, active, o);
		}
		return;
	}
	if (jv >= nvec) return;
	if constexpr (PIPE != 1) {
		for (uint32_t rr = blockIdx.y; rr < rows; rr += gridDim.y) {
			const int r = aa_row(p, rr);
			T f[9][V], o[9][V];
			aa_load<T, V, PHASE>(p, r, j0, f);
			relax_sites<T, V>(p, f, o, r, j0, p.tail_source);
			aa_store<T, V, PHASE>(p, r, j0, o);
		}
	} else {
		uint32_t rr = blockIdx.y;
		if (rr >= rows) return;
		const uint32_t st = gridDim.y;
		T fa[9][V], fb[9][V], o[9][V];
		aa_load<T, V, PHASE>(p, aa_row(p, rr), j0, fa);
#pragma unroll 1
		for (; rr < rows; rr += st) {
			const int r = aa_row(p, rr);
			const uint32_t nx = rr + st < rows ? rr + st : rr;       // last row: a harmless reload of its own (already consumed) slots
			aa_load<T, V, PHASE>(p, aa_row(p, nx), j0, fb);
			relax_sites<T, V>(p, fa, o, r, j0, p.tail_source);
			aa_store<T, V, PHASE>(p, r, j0, o);
#pragma unroll
			for (int s = 0; s < 9; s++)
#pragma unroll
				for (int k = 0; k < V; k++) fa[s][k] = fb[s][k];
		}
	}
}

// ---------------------------------------------------------------------------
// In-place step, one site per thread, threads walking rows blockIdx.y, +gridDim.y, ... like PIPE == 1 -- but the nine
// populations of the NEXT row travel global -> shared memory with cp.async (LDGSTS, 4 or 8 bytes per copy, each thread
// its own nine) while the current row is relaxed, instead of into a second register set: the L2 latency is covered
// without the registers that cost PIPE == 1 its occupancy, and without the serial load -> arithmetic -> store of a
// looping thread (capped grids, DESIGN.md "What did NOT help" (2), (8)).  No barrier: a thread reads only what it copied
// itself (cp.async.wait_group).  Safe in place for the reason PIPE == 1 is: a site's nine slots are its own.
// Dynamic shared memory: 2 x 9 x blockDim.x elements.
// STATUS: written after the round-1 GPU budget was spent; compiles (SASS: LDGSTS), not yet run on a GPU.  Off unless
// lbm_tuning.stage_smem is set; its GPU test is gated by LBM_TEST_EXPERIMENTAL=1.
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
		aa_store<T, 1, PHASE>(p, r, j0, o);
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
// STATUS: like stream_collide_aa_staged_kernel -- compiled (SASS: UBLKCP, SYNCS), not yet run on a GPU;
// lbm_tuning.stage_smem = 2 selects it for phase 2 (phase 1 then runs the cp.async kernel above).
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

// ---------------------------------------------------------------------------
// Persistent multi-step kernel for lattices whose step is short (tens of microseconds): ONE
// cooperative launch runs `nsteps` lattice steps.  Each CTA owns a fixed contiguous range of the
// flattened (row, vector) space; instead of a grid-wide barrier per step, a CTA starts step t as
// soon as the CTAs owning the rows above/below its range have finished step t-1 (their progress
// counters, polled with acquire loads) -- the D2Q9 pull only reaches one row and one column away
// (LB.c:235-243), and the same condition covers the write-after-read hazard of the A/B buffers.
// No launch gaps, no per-step ramp-up/drain.  Source data is read with .cg (L2-coherent) loads:
// L1 is not coherent across CTAs and the same addresses are rewritten every second step.
// ---------------------------------------------------------------------------
struct PersistParams {
	int nsteps;
	int last_tail_source;       // tail_source of the last step (0: leave A un-sourced, LB.c:192-205)
	uint32_t chunk;             // vectors per CTA (multiple of 32)
	int *progress;              // [gridDim.x] completed steps per CTA, zeroed before launch
	int *abort;                 // set when a wait timed out
	unsigned long long timeout_ns;
};

__device__ __forceinline__ int ld_acquire_gpu(const int *p)
{
	int v;
	asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
	return v;
}
__device__ __forceinline__ void st_release_gpu(int *p, int v)
{
	asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

template <typename T, int V, int NB>
__global__ void __launch_bounds__(512)
stream_collide_persistent_kernel(const __grid_constant__ StepParams<T> p, const __grid_constant__ PersistParams q)
{
	const uint32_t nvec = (uint32_t)(p.jmax - p.jmin) / V;
	const uint32_t total = (uint32_t)p.nxl * nvec;           // r0 == 0, r1 == nxl
	const uint32_t lane = threadIdx.x & 31u;
	const uint32_t lo = blockIdx.x * q.chunk;
	const uint32_t hi = min(lo + q.chunk, total);
	__shared__ int s_abort, s_ndeps;
	__shared__ uint32_t s_dep0[3], s_depn[3];

	// CTAs whose ranges touch [lo - 2*nvec, hi + 2*nvec] (periodic in the flattened index): one row up
	// and down, plus the y-wrap of the diagonal pulls at j = 0 and j = ny-1 (e.g. f8 of (r, 0) comes
	// from (r+1, ny-1) = g + 2*nvec - 1).  Up to three runs of consecutive CTA ids.
	if (threadIdx.x == 0) {
		s_abort = 0;
		const long long a0 = (long long)lo - 2ll * nvec, b0 = (long long)hi + 2ll * nvec;
		int n = 0;
		if ((unsigned long long)(b0 - a0 + 1) >= total) {    // tiny lattice: depend on everybody
			s_dep0[0] = 0; s_depn[0] = gridDim.x; n = 1;
		} else {
			const uint32_t ia = (uint32_t)max(a0, 0ll) / q.chunk, ib = (uint32_t)min(b0, (long long)total - 1) / q.chunk;
			s_dep0[n] = ia; s_depn[n] = ib - ia + 1; n++;
			if (a0 < 0) { const uint32_t ja = (uint32_t)(a0 + total) / q.chunk, jb = (total - 1) / q.chunk; s_dep0[n] = ja; s_depn[n] = jb - ja + 1; n++; }
			if (b0 >= (long long)total) { const uint32_t jb = (uint32_t)(b0 - total) / q.chunk; s_dep0[n] = 0; s_depn[n] = jb + 1; n++; }
		}
		s_ndeps = n;
	}
	__syncthreads();

	const T *src = p.src;
	T *dst = p.dst;
	for (int t = 0; t < q.nsteps; t++) {
		if (t > 0) {
			if (threadIdx.x < 32) {                          // warp 0: one dependency per lane, polled concurrently
				unsigned long long t0 = 0;
				uint32_t spins = 0;
				for (int k = 0; k < s_ndeps; k++)
					for (uint32_t c = s_dep0[k] + lane; c < s_dep0[k] + s_depn[k]; c += 32)
						while (ld_acquire_gpu(q.progress + c) < t) {
							if ((++spins & 255u) != 0) continue;         // look at the clock only now and then
							unsigned long long now;
							asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
							if (t0 == 0) t0 = now;
							if (now - t0 > q.timeout_ns || ld_acquire_gpu(q.abort)) { *q.abort = 1; s_abort = 1; break; }
						}
			}
			__syncthreads();
			if (s_abort) return;
		}
		const int tail = (t == q.nsteps - 1) ? q.last_tail_source : 1;
		for (uint32_t base = lo; base < hi; base += blockDim.x) {                    // warp-uniform trip count
			const uint32_t g0 = base + threadIdx.x;
			const bool active = g0 < hi;
			const uint32_t g = active ? g0 : hi - 1;
			update_vector<T, V, NB, 2, true>(p, src, dst, nvec, g, active, lane, tail);
		}
		__syncthreads();                                     // all of this CTA's stores issued ...
		if (threadIdx.x == 0) {
			__threadfence();                                 // ... and visible device-wide before the counter moves
			st_release_gpu(q.progress + blockIdx.x, t + 1);
		}
		const T *tmp = src; src = dst; dst = const_cast<T *>(tmp);
	}
}

// ---------------------------------------------------------------------------
// Dataflow form of the temporally blocked in-place engine (single slab, full-width row bands): ONE launch runs a
// whole macro-step, i.e. every (row piece, level) of the wavefront schedule.  The list of those "launches" is cut
// into items of one row x one column chunk; CTA b takes items b, b + gridDim.x, ... in list order.  What the
// kernel boundary between two levels used to guarantee is checked per item instead: (level s, row r) starts when
// rows r-1, r, r+1 (periodic) have finished level s-1 -- the only dependency the in-place kernels have (read
// after write; see the temporal blocking notes in lbm_api.cu) -- read from per-(level,row) counters of finished
// chunks.  The list order is a topological order (tests/test_tb_schedule.py) and all CTAs are co-resident
// (cooperative launch), so the lowest unfinished item can always run: no deadlock.  No drain / refill between
// the ~60 short launches of a macro-step; the small inverted-pyramid levels hide between the big ones.
// Level 0 depends on the previous macro-step only, i.e. on the previous launch.
// ---------------------------------------------------------------------------
struct DfLaunch {
	int level, r0, rows, tail_source;
	unsigned first;             // index of this launch's first item
};
struct DfParams {
	const DfLaunch *list;
	int nlaunch;
	unsigned nitems;
	int chunks;                 // column chunks per row
	int iters;                  // vectors per thread and item; a chunk is blockDim.x * iters vectors
	int *cnt;                   // [levels][nxl] finished chunks, zeroed before the launch
	int *abort;                 // set when a wait timed out
	int levels;
	unsigned long long timeout_ns;
};

// Both phases in one kernel: left alone ptxas spends 88 registers on the vec-1 DP instance (the phase kernels take 56-59,
// and the L2-resident DP kernel loses 15 % at 104, profiles/r01_run25/NOTE.md) -- capped at 64 like the SP kernels.
#ifndef LBM_DF_MINB
#define LBM_DF_MINB 8
#endif
struct DfItem { int level, r, tail_source; unsigned ch; };

__device__ __forceinline__ DfItem df_locate(const DfParams &q, unsigned n, int &cur, int nxl)
{
	while (cur + 1 < q.nlaunch && n >= q.list[cur + 1].first) cur++;
	const DfLaunch L = q.list[cur];
	const unsigned local = n - L.first;
	const unsigned rr = local / (unsigned)q.chunks;
	DfItem it;
	it.level = L.level; it.tail_source = L.tail_source;
	it.ch = local - rr * (unsigned)q.chunks;
	it.r = L.r0 + (int)rr;
	if (it.r >= nxl) it.r -= nxl;
	return it;
}
// the counter thread `t` (0..2) of a CTA watches for an item: row r-1+t of the level below
__device__ __forceinline__ const int *df_dep(const DfParams &q, const DfItem &it, int t, int nxl)
{
	int rd = it.r - 1 + t;
	if (rd < 0) rd += nxl;
	if (rd >= nxl) rd -= nxl;
	return q.cnt + (size_t)(it.level - 1) * nxl + rd;
}
__device__ __forceinline__ void red_release_gpu_add(int *p, int v)
{
	asm volatile("red.release.gpu.global.add.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

template <typename T, int V, bool CG>
__global__ void __launch_bounds__(sizeof(T) == 4 ? 256 : 128, sizeof(T) == 4 ? 4 : LBM_DF_MINB)
stream_collide_aa_dataflow_kernel(const __grid_constant__ StepParams<T> p, const __grid_constant__ DfParams q)
{
	const uint32_t nvec = (uint32_t)p.ny / V;
	__shared__ int s_abort;
	if (threadIdx.x == 0) s_abort = 0;
	const bool watcher = threadIdx.x < 3;
	int cur = 0;
	unsigned n = blockIdx.x;
	if (n >= q.nitems) return;
	DfItem it = df_locate(q, n, cur, p.nxl);
	int seen = (watcher && it.level > 0) ? ld_acquire_gpu(df_dep(q, it, (int)threadIdx.x, p.nxl)) : q.chunks;
	for (;;) {
		if (watcher && it.level > 0 && seen < q.chunks) {        // usually already seen full one item ago (below)
			const int *c = df_dep(q, it, (int)threadIdx.x, p.nxl);
			unsigned long long t0 = 0;
			uint32_t spins = 0;
			while (ld_acquire_gpu(c) < q.chunks) {
				if ((++spins & 255u) != 0) continue;
				unsigned long long now;
				asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
				if (t0 == 0) t0 = now;
				if (now - t0 > q.timeout_ns || ld_acquire_gpu(q.abort)) { *q.abort = 1; s_abort = 1; break; }
			}
		}
		__syncthreads();
		if (s_abort) return;
		// next item of this CTA: look at its counters now, the answer arrives while this item is computed
		const unsigned nn = n + gridDim.x;
		DfItem nit = it;
		if (nn < q.nitems) {
			nit = df_locate(q, nn, cur, p.nxl);
			seen = (watcher && nit.level > 0) ? ld_acquire_gpu(df_dep(q, nit, (int)threadIdx.x, p.nxl)) : q.chunks;
		}
		const uint32_t v0 = it.ch * blockDim.x * (uint32_t)q.iters + threadIdx.x;
		for (int u = 0; u < q.iters; u++) {
			const uint32_t jv = v0 + (uint32_t)u * blockDim.x;
			if (jv < nvec) {
				const int j0 = (int)jv * V;
				T f[9][V], o[9][V];
				if (it.level & 1) {
					aa_load<T, V, 2, CG>(p, it.r, j0, f);
					relax_sites<T, V>(p, f, o, it.r, j0, it.tail_source);
					aa_store<T, V, 2>(p, it.r, j0, o);
				} else {
					aa_load<T, V, 1, CG>(p, it.r, j0, f);
					relax_sites<T, V>(p, f, o, it.r, j0, it.tail_source);
					aa_store<T, V, 1>(p, it.r, j0, o);
				}
			}
		}
		__syncthreads();                                         // all of this item's stores issued ...
		if (threadIdx.x == 0)                                    // ... and visible device-wide before the row's counter moves
			red_release_gpu_add(q.cnt + (size_t)it.level * p.nxl + it.r, 1);
		if (nn >= q.nitems) return;
		n = nn; it = nit;
	}
}

// ---------------------------------------------------------------------------
// Fused pair: phase 2 of one step and phase 1 of the next in ONE pass over memory.
//   Phase 2 is local (site x relaxes the nine slots of x), phase 1 of the following step pulls f_s from slot s of
//   x + pull(s) -- i.e. the phase-2 result o_s of that neighbour.  A CTA owns a tile of `tile_rows` rows x TC
//   columns, one thread per column, and walks the rows: at row r it runs phase 2 of (r, c), leaves the nine
//   results in a ring of rows in shared memory, and then runs phase 1 of (r-1, c) from the ring.  For a site
//   whose 3x3 neighbourhood lies inside the tile ("interior") the intermediate populations never travel: 9 loads
//   and 9 stores for TWO timesteps, half the traffic of the two separate kernels (to L2 and, where the lattice
//   does not fit it, to HBM).
//   Sites on the rim of a tile ("edge": first/last row, first/last column) cannot take phase 1 here (a neighbour's
//   phase-2 result belongs to another CTA); they store their phase-2 result the ordinary way and
//   stream_collide_aa_edges_kernel gives them phase 1 afterwards.  That kernel reads slot k of interior sites next to
//   the rim as well, so an interior site stores o_k into its slot k wherever the slot's consumer x - pull(k) is an
//   edge site; every other slot of an interior site receives the phase-1 result of its interior consumer, exactly
//   where the ordinary phase 1 puts it.  Each slot is written by one thread per sub-step and tiles read and write
//   their own sites only: no ordering between CTAs is needed.  Same relax_sites() as everywhere: bit-identical.
//   Rows 0 and nxl-1 of a slab must be tile edges (their halo traffic is the ordinary kernels' business).
// ---------------------------------------------------------------------------
struct PairParams {
	int tile_rows;              // rows per tile (>= 3 for any interior)
	int tail_source2;           // tail_source of the phase-1 half (StepParams.tail_source is the phase-2 half's)
};

#ifndef LBM_PAIR_MINB
#define LBM_PAIR_MINB 0
#endif
template <typename T, int TC>
__global__ void __launch_bounds__(TC, LBM_PAIR_MINB)
stream_collide_aa_pair_kernel(const __grid_constant__ StepParams<T> p, const __grid_constant__ PairParams q)
{
	asm volatile("griddepcontrol.launch_dependents;");
	// four-row ring: row r is written while rows r-3..r-1 may still be read, so one barrier per row is enough (a thread
	// reaches the write of row r+2, which reuses the slot of r-2, only past the barrier of row r+1, i.e. after every
	// thread has finished reading for row r)
	__shared__ T sh[4][9][TC];
	const int j = (int)threadIdx.x;
	const int c0 = (int)blockIdx.x * TC;
	const int c1 = min(c0 + TC, p.ny);
	const int c = c0 + j;
	const bool live = c < c1;
	const bool col_edge = c == c0 || c == c1 - 1;
	const int tr0 = p.r0 + (int)blockIdx.y * q.tile_rows;
	const int tr1 = min(tr0 + q.tile_rows, p.r1);
	asm volatile("griddepcontrol.wait;" ::: "memory");
	// the nine populations of row r are loaded one row ahead (nothing writes a row of this tile before its own phase 2
	// has read it: phase 1 of row r-1 reaches rows r-2..r only, and other CTAs never write this tile's sites)
	T fn[9][1];
	if (live) aa_load<T, 1, 2>(p, tr0, c, fn);
	int cur = 0;                                             // ring slot of row r; row r-1 in (cur+3)&3, row r-2 in (cur+2)&3
	for (int r = tr0; r < tr1; r++) {
		if (live) {
			T f[9][1], o[9][1];
#pragma unroll
			for (int s = 0; s < 9; s++) f[s][0] = fn[s][0];
			if (r + 1 < tr1) aa_load<T, 1, 2>(p, r + 1, c, fn);
			relax_sites<T, 1>(p, f, o, r, c, p.tail_source);
#pragma unroll
			for (int s = 0; s < 9; s++) sh[cur][s][j] = o[s][0];
			if (col_edge || r == tr0 || r == tr1 - 1) {
				aa_store<T, 1, 2>(p, r, c, o);
			} else {
				T *const own = p.dst + (size_t)(r + 1) * p.ldy + c;
				const bool dn = r - 1 == tr0, up = r + 1 == tr1 - 1, lf = c - 1 == c0, rt = c + 1 == c1 - 1;
#pragma unroll
				for (int k = 1; k < 9; k++) {                    // consumer of slot k: site (r - pull_di(k), c - pull_dj(k))
					const bool row_hit = pull_di(k) > 0 ? dn : pull_di(k) < 0 ? up : false;
					const bool col_hit = pull_dj(k) > 0 ? lf : pull_dj(k) < 0 ? rt : false;
					if (row_hit || col_hit) own[(size_t)k * p.plane] = o[k][0];
				}
			}
		}
		__syncthreads();
		if (live && !col_edge && r - 1 > tr0) {                  // phase 1 of (r-1, c): rows r-2, r-1, r are in the ring
			const int slot_of[3] = {(cur + 2) & 3, (cur + 3) & 3, cur};
			T f[9][1], o[9][1];
#pragma unroll
			for (int s = 0; s < 9; s++) f[s][0] = sh[slot_of[pull_di(s) + 1]][s][j + pull_dj(s)];
			relax_sites<T, 1>(p, f, o, r - 1, c, q.tail_source2);
			aa_store<T, 1, 1>(p, r - 1, c, o);
		}
		cur = (cur + 1) & 3;
	}
}

// phase 1 for the sites the pair kernel left out: the first and last row of every row tile (whole rows), and the first
// and last column of every column tile in the other rows.  blockIdx.y < edge_rows: row ordinal; beyond: rows of the
// launch's range, threads over the column-tile rims.
struct EdgeParams {
	int tile_rows, tile_cols;
	int edge_rows;              // 2 per row tile
	int col_tiles;
};

template <typename T>
__global__ void __launch_bounds__(128)
stream_collide_aa_edges_kernel(const __grid_constant__ StepParams<T> p, const __grid_constant__ EdgeParams q)
{
	asm volatile("griddepcontrol.launch_dependents;");
	int r, c;
	if ((int)blockIdx.y < q.edge_rows) {
		const int t = (int)blockIdx.y >> 1;
		const int tr0 = p.r0 + t * q.tile_rows, tr1 = min(tr0 + q.tile_rows, p.r1);
		r = (blockIdx.y & 1) ? tr1 - 1 : tr0;
		if ((blockIdx.y & 1) && tr1 - 1 == tr0) return;          // one-row tile: once
		c = (int)(blockIdx.x * blockDim.x + threadIdx.x);
		if (c >= p.ny) return;
	} else {
		r = p.r0 + (int)blockIdx.y - q.edge_rows;
		const int rel = (r - p.r0) % q.tile_rows;
		if (rel == 0 || rel == q.tile_rows - 1 || r == p.r1 - 1) return;     // a tile's first / last row: done above
		const int e = (int)(blockIdx.x * blockDim.x + threadIdx.x);
		if (e >= 2 * q.col_tiles) return;
		const int c0 = (e >> 1) * q.tile_cols, c1 = min(c0 + q.tile_cols, p.ny);
		c = (e & 1) ? c1 - 1 : c0;
		if ((e & 1) && c1 - 1 == c0) return;
	}
	asm volatile("griddepcontrol.wait;" ::: "memory");
	T f[9][1], o[9][1];
	aa_load<T, 1, 1>(p, r, c, f);
	relax_sites<T, 1>(p, f, o, r, c, p.tail_source);
	aa_store<T, 1, 1>(p, r, c, o);
}

// ---------------------------------------------------------------------------
// scalar-order kernel: StreamCollide (LB.c:216-302) for columns [jmin,jmax) --
// the ragged tail j >= NYVECMAX of a vector build, or everything when the
// matched reference build is a prec_*_serial.h one.  One site per thread.
// ---------------------------------------------------------------------------
// PHASE 0: A/B (src -> dst); 1, 2: the in-place phases described at update_vector_aa.
template <typename T, int PHASE>
__global__ void __launch_bounds__(256)
stream_collide_scalar_kernel(const __grid_constant__ StepParams<T> p)
{
	const uint32_t ncol = (uint32_t)(p.jmax - p.jmin);
	const uint32_t total = (uint32_t)(p.r1 - p.r0) * ncol;
	for (uint32_t g = blockIdx.x * blockDim.x + threadIdx.x; g < total; g += gridDim.x * blockDim.x) {
		const uint32_t rr = g / ncol;
		int r = p.r0 + (int)rr;
		if (PHASE != 0 && r >= p.nxl) r -= p.nxl;            // in-place wavefront schedule: wrapping row range
		const int j = p.jmin + (int)(g - rr * ncol);
		const int jl = j == 0 ? p.ny - 1 : j - 1;
		const int jr = j + 1 == p.ny ? 0 : j + 1;
		T f[9], o[9];
		if (PHASE == 0) {
			const T *row0 = p.src + (size_t)r * p.ldy;
#pragma unroll
			for (int s = 0; s < 9; s++) {
				const int jj = pull_dj(s) < 0 ? jl : pull_dj(s) > 0 ? jr : j;
				f[s] = __ldg(row0 + (size_t)(pull_di(s) + 1) * p.ldy + (size_t)s * p.plane + jj);
			}
		} else if (PHASE == 1) {
			const T *row0 = p.dst + (size_t)r * p.ldy;
#pragma unroll
			for (int s = 0; s < 9; s++) {
				const int jj = pull_dj(s) < 0 ? jl : pull_dj(s) > 0 ? jr : j;
				f[s] = row0[(size_t)(pull_di(s) + 1) * p.ldy + (size_t)s * p.plane + jj];
			}
		} else {
			const T *site = p.dst + (size_t)(r + 1) * p.ldy + j;
#pragma unroll
			for (int s = 0; s < 9; s++) f[s] = site[(size_t)opposite(s) * p.plane];
		}
		const bool wall = (__ldg(p.mask + (size_t)r * p.ldw + (j >> 5)) >> (j & 31)) & 1u;
		if (wall) bounce<T>(f, o);
		else {
			collide_scalar_order<T>(f, o, p);
			if (j == 0 && p.tail_source) source_site<T>(o, p);
		}
		if (PHASE == 1) {
			T *own = p.dst + (size_t)(r + 1) * p.ldy;
#pragma unroll
			for (int k = 0; k < 9; k++) {
				const int jj = pull_dj(k) < 0 ? jl : pull_dj(k) > 0 ? jr : j;
				T *b;
				if (pull_di(k) < 0) b = r == 0 ? p.dn_halo - p.ldy + (size_t)k * p.dn_plane : own - p.ldy + (size_t)k * p.plane;
				else if (pull_di(k) > 0) b = r == p.nxl - 1 ? p.up_halo + p.ldy + (size_t)k * p.up_plane : own + p.ldy + (size_t)k * p.plane;
				else b = own + (size_t)k * p.plane;
				b[jj] = o[opposite(k)];
			}
			continue;
		}
		T *d = p.dst + (size_t)(r + 1) * p.ldy + j;
		if (!wall) d[0] = o[0];
#pragma unroll
		for (int s = 1; s < 9; s++) d[(size_t)s * p.plane] = o[s];
		if (r == 0) {
			p.dn_halo[4 * p.dn_plane + j] = o[4]; p.dn_halo[7 * p.dn_plane + j] = o[7]; p.dn_halo[8 * p.dn_plane + j] = o[8];
		}
		if (r == p.nxl - 1) {
			p.up_halo[2 * p.up_plane + j] = o[2]; p.up_halo[5 * p.up_plane + j] = o[5]; p.up_halo[6 * p.up_plane + j] = o[6];
		}
	}
}

// ---------------------------------------------------------------------------
// small kernels around the step
// ---------------------------------------------------------------------------

// ApplySource (LB.c:489-518), in place on column 0 of owned rows of one buffer.
template <typename T>
__global__ void apply_source_kernel(T *f, const StepParams<T> p)
{
	const int r = blockIdx.x * blockDim.x + threadIdx.x;
	if (r >= p.nxl) return;
	if (__ldg(p.mask + (size_t)r * p.ldw) & 1u) return;          // walls[I(i,0,0)] != 0
	T *site = f + (size_t)(r + 1) * p.ldy;
	T v[9];
	v[0] = T(0);
#pragma unroll
	for (int s = 1; s < 9; s++) v[s] = site[(size_t)s * p.plane];
	source_site<T>(v, p);
	site[1 * p.plane] = v[1]; site[3 * p.plane] = v[3]; site[5 * p.plane] = v[5];
	site[6 * p.plane] = v[6]; site[7 * p.plane] = v[7]; site[8 * p.plane] = v[8];
}

// Copy the halo speeds of the first/last owned row of `f` into the neighbours' halo rows
// (same destinations as the fused push in the step kernel).
template <typename T>
__global__ void halo_push_kernel(const T *f, const StepParams<T> p)
{
	const int j = blockIdx.x * blockDim.x + threadIdx.x;
	if (j >= p.ny) return;
	const T *lo = f + (size_t)1 * p.ldy + j, *hi = f + (size_t)p.nxl * p.ldy + j;
	p.dn_halo[4 * p.dn_plane + j] = lo[4 * p.plane];
	p.dn_halo[7 * p.dn_plane + j] = lo[7 * p.plane];
	p.dn_halo[8 * p.dn_plane + j] = lo[8 * p.plane];
	p.up_halo[2 * p.up_plane + j] = hi[2 * p.plane];
	p.up_halo[5 * p.up_plane + j] = hi[5 * p.plane];
	p.up_halo[6 * p.up_plane + j] = hi[6 * p.plane];
}

// Staged exchange: messages are [3][ny] contiguous.  to_up = speeds 2,5,6 of the last owned
// row, to_dn = speeds 4,7,8 of the first owned row.
template <typename T>
__global__ void halo_pack_kernel(const T *f, T *to_up, T *to_dn, size_t plane, int ldy, int nxl, int ny)
{
	const int j = blockIdx.x * blockDim.x + threadIdx.x;
	if (j >= ny) return;
	const T *lo = f + (size_t)1 * ldy + j, *hi = f + (size_t)nxl * ldy + j;
	to_up[j] = hi[2 * plane]; to_up[ny + j] = hi[5 * plane]; to_up[2 * (size_t)ny + j] = hi[6 * plane];
	to_dn[j] = lo[4 * plane]; to_dn[ny + j] = lo[7 * plane]; to_dn[2 * (size_t)ny + j] = lo[8 * plane];
}
template <typename T>
__global__ void halo_unpack_kernel(T *f, const T *from_dn, const T *from_up, size_t plane, int ldy, int nxl, int ny)
{
	const int j = blockIdx.x * blockDim.x + threadIdx.x;
	if (j >= ny) return;
	T *lo = f + j, *hi = f + (size_t)(nxl + 1) * ldy + j;            // halo rows: storage 0 and nxl+1
	lo[2 * plane] = from_dn[j]; lo[5 * plane] = from_dn[ny + j]; lo[6 * plane] = from_dn[2 * (size_t)ny + j];
	hi[4 * plane] = from_up[j]; hi[7 * plane] = from_up[ny + j]; hi[8 * plane] = from_up[2 * (size_t)ny + j];
}

// Scene description, evaluated per site (InitializeArrays, LB.c:559-578, plus a disc)
struct SceneDev {
	int nx, channel_walls;
	int bar_i0, bar_i1, bar_j0, bar_j1;
	long long disc_ci, disc_cj, disc_r2;   // disc_r2 < 0: none
};
__device__ __forceinline__ bool scene_is_wall(const SceneDev &sc, int gi, int j)
{
	if (sc.channel_walls && (gi == 0 || gi == sc.nx - 1)) return true;
	if (gi >= sc.bar_i0 && gi < sc.bar_i1 && j >= sc.bar_j0 && j < sc.bar_j1) return true;
	if (sc.disc_r2 >= 0) {
		const long long di = gi - sc.disc_ci, dj = j - sc.disc_cj;
		if (di * di + dj * dj <= sc.disc_r2) return true;
	}
	return false;
}
// one thread per mask word
__global__ void scene_mask_kernel(uint32_t *mask, SceneDev sc, int row_lo, int nxl, int ny, int ldw)
{
	const size_t g = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (g >= (size_t)nxl * ldw) return;
	const int r = (int)(g / ldw), w = (int)(g % ldw);
	uint32_t bits = 0;
	for (int b = 0; b < 32; b++) {
		const int j = w * 32 + b;
		if (j < ny && scene_is_wall(sc, row_lo + r, j)) bits |= 1u << b;
	}
	mask[g] = bits;
}
// 32-bit integer hash (lowbias32) -- the closed-form density perturbation of lbm_scene.perturb_seed
__host__ __device__ __forceinline__ uint32_t hash32(uint32_t x)
{
	x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
	return x;
}
// fA = rho*omega_s at every owned site (walls included), fB = 0 (LB.c:580-603); j < ny only.
// rho = rho0 (LB.c:56), or, with a perturbation seed, rho0 * (1 + amp * (u - 1/2)) with u = h / 2^32 in [0,1) from the
// integer hash h of (global row, column, seed): a closed-form rough initial state, every operation one rounding in T
// (geometry.perturbed_rest evaluates the same expression with numpy for any window of the lattice).
template <typename T>
__global__ void scene_fill_kernel(T *fA, T *fB, size_t plane, int ldy, int nxl, int ny, int row_lo, T rho0, T w0, T w14, T w58, uint32_t seed, T amp)
{
	typedef fp<T> A;
	const size_t g = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (g >= (size_t)nxl * ny) return;
	const int r = (int)(g / ny), j = (int)(g % ny);
	const size_t o = (size_t)(r + 1) * ldy + j;
	T rho = rho0;
	if (seed) {
		const uint32_t h = hash32((uint32_t)(row_lo + r) * 0x9E3779B1u ^ hash32((uint32_t)j + seed));
		const T u = A::mul((T)h, T(1.0 / 4294967296.0));             /* exact scaling; (T)h rounds to nearest in SP */
		rho = A::mul(rho0, A::add(T(1), A::mul(amp, A::sub(u, T(0.5)))));
	}
#pragma unroll
	for (int s = 0; s < 9; s++) {
		fA[o + s * plane] = A::mul(rho, s == 0 ? w0 : (s <= 4 ? w14 : w58));
		if (fB) fB[o + s * plane] = T(0);
	}
}

// Decomposition-invariant 64-bit checksum of one buffer: the sum (mod 2^64) over every owned element (s, global row i,
// column j < ny) of mix64(bit pattern ^ mix64(global element index)), global index = (s*NX + i)*NY + j.  Addition is
// commutative, so the checksums of the slabs of a ring add up to the single-GPU value whatever the partition and the
// reduction order; position enters through the inner mix, so permuted or shifted data does not cancel.
__host__ __device__ __forceinline__ unsigned long long mix64(unsigned long long z)
{
	z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
	z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
	return z ^ (z >> 31);
}
__device__ __forceinline__ unsigned long long bits_of(double v) { return (unsigned long long)__double_as_longlong(v); }
__device__ __forceinline__ unsigned long long bits_of(float v) { return (unsigned long long)__float_as_uint(v); }
template <typename T>
__global__ void __launch_bounds__(256)
checksum_kernel(const T *f, size_t plane, int ldy, int nxl, int ny, int row_lo, int nx_global, unsigned long long *out)
{
	unsigned long long acc = 0;
	const size_t total = (size_t)nxl * ny;
	for (size_t g = (size_t)blockIdx.x * blockDim.x + threadIdx.x; g < total; g += (size_t)gridDim.x * blockDim.x) {
		const int r = (int)(g / ny), j = (int)(g % ny);
		const T *site = f + (size_t)(r + 1) * ldy + j;
#pragma unroll
		for (int s = 0; s < 9; s++) {
			const unsigned long long idx = ((unsigned long long)s * nx_global + (unsigned long long)(row_lo + r)) * ny + j;
			acc += mix64(bits_of(site[(size_t)s * plane]) ^ mix64(idx));
		}
	}
#pragma unroll
	for (int d = 16; d > 0; d >>= 1) acc += __shfl_down_sync(0xffffffffu, acc, d);
	__shared__ unsigned long long part[8];
	if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = acc;
	__syncthreads();
	if (threadIdx.x == 0) {
		unsigned long long t = 0;
		for (int w = 0; w < (int)(blockDim.x >> 5); w++) t += part[w];
		atomicAdd(out, t);
	}
}

// isfinite guard (the reference traps NaN / overflow / division by zero with feenableexcept, LB.c:129): counts the
// owned elements of one buffer that are not finite
template <typename T>
__global__ void __launch_bounds__(256)
nonfinite_kernel(const T *f, size_t plane, int ldy, int nxl, int ny, unsigned long long *out)
{
	unsigned int bad = 0;
	const size_t total = (size_t)nxl * ny;
	for (size_t g = (size_t)blockIdx.x * blockDim.x + threadIdx.x; g < total; g += (size_t)gridDim.x * blockDim.x) {
		const int r = (int)(g / ny), j = (int)(g % ny);
		const T *site = f + (size_t)(r + 1) * ldy + j;
#pragma unroll
		for (int s = 0; s < 9; s++) bad += !isfinite(site[(size_t)s * plane]);
	}
	bad = __reduce_add_sync(0xffffffffu, bad);
	if ((threadIdx.x & 31) == 0 && bad) atomicAdd(out, (unsigned long long)bad);
}

// walls chars [nxl][ldc] -> bit words; *bad is set if a value other than 0/1 is seen
__global__ void mask_pack_kernel(const char *walls, size_t ldc, uint32_t *mask, int nxl, int ny, int ldw, int *bad)
{
	const size_t g = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (g >= (size_t)nxl * ldw) return;
	const int r = (int)(g / ldw), w = (int)(g % ldw);
	uint32_t bits = 0;
	for (int b = 0; b < 32; b++) {
		const int j = w * 32 + b;
		if (j < ny) {
			const char c = walls[(size_t)r * ldc + j];
			if (c == 1) bits |= 1u << b; else if (c != 0) *bad = 1;
		}
	}
	mask[g] = bits;
}
__global__ void mask_unpack_kernel(const uint32_t *mask, char *walls, size_t ldc, int nxl, int ny, int ldw)
{
	const size_t g = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (g >= (size_t)nxl * ny) return;
	const int r = (int)(g / ny), j = (int)(g % ny);
	walls[(size_t)r * ldc + j] = (char)((mask[(size_t)r * ldw + (j >> 5)] >> (j & 31)) & 1u);
}

// rho, u_x, u_y in the order of PrintLattice / ComputeReynolds (LB.c:621-629, 535-540).
// Dense outputs [nxl][ncol]; column range [j0, j0+ncol).
template <typename T>
__global__ void macroscopic_kernel(const T *f, size_t plane, int ldy, int nxl, int j0, int ncol, T *rho_o, T *ux_o, T *uy_o)
{
	typedef fp<T> A;
	const size_t g = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (g >= (size_t)nxl * ncol) return;
	const int r = (int)(g / ncol), j = j0 + (int)(g % ncol);
	const T *site = f + (size_t)(r + 1) * ldy + j;
	T v[9], rho = T(0);
#pragma unroll
	for (int s = 0; s < 9; s++) { v[s] = site[(size_t)s * plane]; rho = A::add(rho, v[s]); }
	if (rho_o) rho_o[g] = rho;
	if (ux_o) ux_o[g] = A::div(A::sub(A::add(A::add(v[6], v[2]), v[5]), A::add(A::add(v[7], v[4]), v[8])), rho);
	if (uy_o) uy_o[g] = A::div(A::sub(A::add(A::add(v[5], v[1]), v[8]), A::add(A::add(v[6], v[3]), v[7])), rho);
}

// ---- neighbour barrier between ranks living in different processes / GPUs ----
// signal: my epoch into the slot each neighbour reserves for me.  Runs after the boundary
// kernel in stream order, so the halo stores it announces have been performed.
__global__ void flag_signal_kernel(volatile int *up_slot, volatile int *dn_slot, int epoch)
{
	__threadfence_system();
	*up_slot = epoch;
	*dn_slot = epoch;
	__threadfence_system();
}
// wait: spin until both neighbours have announced `epoch`.  The spinners live on DIFFERENT
// GPUs than the signallers.  flags[3] is the ring's abort word: a wait that gives up after
// ~timeout_ns raises it here AND in both neighbours' flag blocks, every wait returns at once
// when it finds it raised (at entry or while spinning), so one stalled rank stops the whole
// ring within one time-out instead of one time-out per remaining wait; lbm_sync reports it.
__global__ void flag_wait_kernel(volatile int *flags, volatile int *up_abort, volatile int *dn_abort, int epoch, unsigned long long timeout_ns)
{
	if (flags[3]) return;
	unsigned long long t0;
	asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
	while (flags[0] < epoch || flags[1] < epoch) {
		unsigned long long t;
		asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
		if (flags[3] || t - t0 > timeout_ns) {
			flags[3] = 1; *up_abort = 1; *dn_abort = 1;
			break;
		}
		__nanosleep(200);
	}
	__threadfence_system();
}

} // namespace lbm
