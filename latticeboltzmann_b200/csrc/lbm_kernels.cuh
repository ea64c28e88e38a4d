// lbm_kernels.cuh -- sm_100a device code of the D2Q9 BGK timestep path.
//
// What it computes is StreamCollideVec / StreamCollide / ApplySource of the
// reference (src/latticeboltzmann.c = "LB.c", lines cited at each function); how
// it computes it is B200-first:
//   * one fused pull kernel: stream + collide + bounce-back + body force of the
//     NEXT step + halo push, A/B double buffer, SoA f[s][row][j] planes;
//   * each thread owns V consecutive sites of one row (aligned vector loads /
//     stores along j); the +-1 column element of the six shifted streams is one
//     extra element load that hits the line a neighbouring lane fetches anyway
//     (coalesced in L1), so HBM sees every element of f once per step (144 B DP /
//     72 B SP per site update) -- measured equal or faster than exchanging it by
//     warp shuffle (profiles/r01_run2, r01_run17);
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
	int outflow;           // 0: y wraps periodically (WRAPAROUND 1, LB.c:42-43,233-234); 1: open ends, see pull_columns()
	const uint32_t *slip;  // nullptr: every wall site bounces back (LB.c:246-255); else two bit planes [2][nxl][ldw]:
	                       // plane 0 = "reflect" (slip, LB.c:21) walls whose normal is x, plane 1 = normal y
	// Neighbour barrier folded into a seam launch of the in-place kernel (slabs of a ring living in other processes):
	int *seam_flags;       // nullptr: none.  Else this slab's flag block: every CTA waits until [0] and [1] (the epochs the
	                       // two neighbours have announced) reach seam_wait, or gives up (abort word [3], see flag_wait_kernel)
	int *seam_up, *seam_dn;// the slots this slab announces in (neighbours' flag blocks): the LAST CTA to finish stores seam_signal
	unsigned *seam_done;   // CTAs finished so far (zero between launches)
	int seam_wait, seam_signal;
	unsigned long long seam_timeout_ns;
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

// "reflect" (slip) walls, named at LB.c:21 but not implemented by the reference: specular reflection instead of
// reversal.  NORMAL_X: the wall's normal is the x (row) axis -- the channel walls -- so (dx,dy) leaves as (-dx,dy);
// otherwise the normal is y and (dx,dy) leaves as (dx,-dy).  Like bounce-back a permutation of the site's own nine
// values (in-place safe); o[0] is never stored.
__device__ __forceinline__ constexpr int mirror_x(int s) { return s == 2 ? 4 : s == 4 ? 2 : s == 5 ? 8 : s == 8 ? 5 : s == 6 ? 7 : s == 7 ? 6 : s; }
__device__ __forceinline__ constexpr int mirror_y(int s) { return s == 1 ? 3 : s == 3 ? 1 : s == 5 ? 6 : s == 6 ? 5 : s == 7 ? 8 : s == 8 ? 7 : s; }
template <typename T, bool NORMAL_X> __device__ __forceinline__ void reflect(const T (&f)[9], T (&o)[9])
{
#pragma unroll
	for (int s = 0; s < 9; s++) o[s] = f[NORMAL_X ? mirror_x(s) : mirror_y(s)];
}

// columns the two out-of-vector neighbours are pulled from: wrapped (LB.c:337-338) or, with open ends ("fluid flows
// out of the domain", LB.c:42: the reference names the option, its WRAPAROUND macro is dead), the end column itself --
// a zero-gradient end: what would be pulled from beyond the lattice is the same speed of the end column's own site in
// that row, and what streams past the end is dropped.
template <typename T> __device__ __forceinline__ void pull_columns(const StepParams<T> &p, int j0, int V, int &jl, int &jr)
{
	jl = j0 == 0 ? (p.outflow ? 0 : p.ny - 1) : j0 - 1;
	jr = j0 + V == p.ny ? (p.outflow ? p.ny - 1 : 0) : j0 + V;
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
	uint32_t sx = 0u, sy = 0u;                               // slip walls among my V sites (subsets of `bits`)
	if (bits != 0u && p.slip != nullptr) {
		const size_t w = (size_t)r * p.ldw + (j0 >> 5);
		sx = (__ldg(p.slip + w) >> sh) & bits;
		sy = (__ldg(p.slip + (size_t)p.nxl * p.ldw + w) >> sh) & bits;
	}
#pragma unroll
	for (int k = 0; k < V; k++) {
		T fk[9], ok[9];
#pragma unroll
		for (int s = 0; s < 9; s++) fk[s] = f[s][k];
		if ((bits >> k) & 1u) {
			if ((sx >> k) & 1u) reflect<T, true>(fk, ok);
			else if ((sy >> k) & 1u) reflect<T, false>(fk, ok);
			else bounce<T>(fk, ok);
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
//   HINT  0: default caching, 1: L1::no_allocate loads + .cs stores
// Rows i-1 / i+1 come from the halo rows (the periodic wrap of a single slab, the seam of a ring); threads of the
// first / last owned row push the three speeds the neighbour will pull into its halo row.
// ---------------------------------------------------------------------------
template <typename T, int V, int HINT>
__device__ __forceinline__ void update_vector(const StepParams<T> &p, uint32_t nvec, uint32_t g)
{
	const uint32_t rr = g / nvec;
	const uint32_t jv = g - rr * nvec;
	const int r = p.r0 + (int)rr;                            // owned row; storage row r+1
	const int j0 = p.jmin + (int)jv * V;
	int jl, jr;
	pull_columns(p, j0, V, jl, jr);

	const T *row[3];                                         // rows i-1, i, i+1 of the source
	row[1] = p.src + (size_t)(r + 1) * p.ldy;
	row[0] = row[1] - p.ldy;
	row[2] = row[1] + p.ldy;

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
				f[s][0] = ld1<HINT>(base_s + jl);
#pragma unroll
				for (int k = 1; k < V; k++) f[s][k] = own[k - 1];
			} else {                                         // element k comes from column j0+k+1
				f[s][V - 1] = ld1<HINT>(base_s + jr);
#pragma unroll
				for (int k = 0; k < V - 1; k++) f[s][k] = own[k + 1];
			}
		}
	}

	T o[9][V];
	const uint32_t bits = relax_sites<T, V>(p, f, o, r, j0, p.tail_source);

	T *d = p.dst + (size_t)(r + 1) * p.ldy + j0;
	if (bits == 0u) stv<HINT, V>(d, o[0]);
	else {
#pragma unroll
		for (int k = 0; k < V; k++) if (!((bits >> k) & 1u)) d[k] = o[0][k];     // f0 untouched at walls
	}
#pragma unroll
	for (int s = 1; s < 9; s++) stv<HINT, V>(d + (size_t)s * p.plane, o[s]);
	if (r == 0) {                                            // what the slab below pulls from its row "hi" (LB.c:239,242,243)
		stv<HINT, V>(p.dn_halo + 4 * p.dn_plane + j0, o[4]);
		stv<HINT, V>(p.dn_halo + 7 * p.dn_plane + j0, o[7]);
		stv<HINT, V>(p.dn_halo + 8 * p.dn_plane + j0, o[8]);
	}
	if (r == p.nxl - 1) {                                    // what the slab above pulls from its row "lo-1" (LB.c:237,240,241)
		stv<HINT, V>(p.up_halo + 2 * p.up_plane + j0, o[2]);
		stv<HINT, V>(p.up_halo + 5 * p.up_plane + j0, o[5]);
		stv<HINT, V>(p.up_halo + 6 * p.up_plane + j0, o[6]);
	}
}

// One lattice step per launch.  1-D grid; grid-stride over (row, vector) pairs.
template <typename T, int V, int HINT>
__global__ void __launch_bounds__(512)
stream_collide_vec_kernel(const __grid_constant__ StepParams<T> p)
{
	// Programmatic dependent launch (no-ops unless launched with the PDL attribute): let the NEXT step's
	// CTAs become resident while this step drains, and do not touch f before the PREVIOUS step is complete.
	asm volatile("griddepcontrol.launch_dependents;");
	const uint32_t nvec = (uint32_t)(p.jmax - p.jmin) / V;
	const uint32_t total = (uint32_t)(p.r1 - p.r0) * nvec;
	const uint32_t stride = gridDim.x * blockDim.x;
	asm volatile("griddepcontrol.wait;" ::: "memory");
	for (uint32_t g = blockIdx.x * blockDim.x + threadIdx.x; g < total; g += stride)
		update_vector<T, V, HINT>(p, nvec, g);
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
// L2 eviction priority of the in-place kernels' accesses (temporal blocking keeps a tile L2-resident for k levels):
//   EP 0  default
//   EP 1  evict_first: the stores of a tile's LAST level -- nothing reads them again before the tile has left L2
//   EP 2  evict_last:  the loads and stores of the other levels -- the next level reads exactly these lines
// (createpolicy + .L2::cache_hint accesses; always coherent loads, never ld.global.nc: the buffer is written by the same kernel)
template <int EP> __device__ __forceinline__ uint64_t l2_policy()
{
	uint64_t pol = 0;
	if constexpr (EP == 1) asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
	if constexpr (EP == 2) asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
	return pol;
}
template <int EP> __device__ __forceinline__ double ld_ep(const double *p, uint64_t pol)
{
	if constexpr (EP == 0) return *p;
	double v;
	asm volatile("ld.global.L2::cache_hint.f64 %0, [%1], %2;" : "=d"(v) : "l"(p), "l"(pol));
	return v;
}
template <int EP> __device__ __forceinline__ float ld_ep(const float *p, uint64_t pol)
{
	if constexpr (EP == 0) return *p;
	float v;
	asm volatile("ld.global.L2::cache_hint.f32 %0, [%1], %2;" : "=f"(v) : "l"(p), "l"(pol));
	return v;
}
template <int EP> __device__ __forceinline__ void st_ep(double *p, double v, uint64_t pol)
{
	if constexpr (EP == 0) *p = v;
	else asm volatile("st.global.L2::cache_hint.f64 [%0], %1, %2;" ::"l"(p), "d"(v), "l"(pol));
}
template <int EP> __device__ __forceinline__ void st_ep(float *p, float v, uint64_t pol)
{
	if constexpr (EP == 0) *p = v;
	else asm volatile("st.global.L2::cache_hint.f32 [%0], %1, %2;" ::"l"(p), "f"(v), "l"(pol));
}
template <int V, int EP> __device__ __forceinline__ void ldc(const double *p, double (&v)[V], uint64_t pol)
{
	if constexpr (V == 1) v[0] = ld_ep<EP>(p, pol);
	else {
#pragma unroll
		for (int k = 0; k < V; k += 2) {
			if constexpr (EP == 0) { const double2 t = *reinterpret_cast<const double2 *>(p + k); v[k] = t.x; v[k + 1] = t.y; }
			else asm volatile("ld.global.L2::cache_hint.v2.f64 {%0,%1}, [%2], %3;" : "=d"(v[k]), "=d"(v[k + 1]) : "l"(p + k), "l"(pol));
		}
	}
}
template <int V, int EP> __device__ __forceinline__ void ldc(const float *p, float (&v)[V], uint64_t pol)
{
	if constexpr (V == 1) v[0] = ld_ep<EP>(p, pol);
	else if constexpr (V == 2) {
		if constexpr (EP == 0) { const float2 t = *reinterpret_cast<const float2 *>(p); v[0] = t.x; v[1] = t.y; }
		else asm volatile("ld.global.L2::cache_hint.v2.f32 {%0,%1}, [%2], %3;" : "=f"(v[0]), "=f"(v[1]) : "l"(p), "l"(pol));
	} else {
#pragma unroll
		for (int k = 0; k < V; k += 4) {
			if constexpr (EP == 0) { const float4 t = *reinterpret_cast<const float4 *>(p + k); v[k] = t.x; v[k + 1] = t.y; v[k + 2] = t.z; v[k + 3] = t.w; }
			else asm volatile("ld.global.L2::cache_hint.v4.f32 {%0,%1,%2,%3}, [%4], %5;" : "=f"(v[k]), "=f"(v[k + 1]), "=f"(v[k + 2]), "=f"(v[k + 3]) : "l"(p + k), "l"(pol));
		}
	}
}
template <int V, int EP> __device__ __forceinline__ void ste(double *p, const double (&v)[V], uint64_t pol)
{
	if constexpr (EP == 0) stv<0, V>(p, v);
	else if constexpr (V == 1) st_ep<EP>(p, v[0], pol);
	else {
#pragma unroll
		for (int k = 0; k < V; k += 2)
			asm volatile("st.global.L2::cache_hint.v2.f64 [%0], {%1,%2}, %3;" ::"l"(p + k), "d"(v[k]), "d"(v[k + 1]), "l"(pol));
	}
}
template <int V, int EP> __device__ __forceinline__ void ste(float *p, const float (&v)[V], uint64_t pol)
{
	if constexpr (EP == 0) stv<0, V>(p, v);
	else if constexpr (V == 1) st_ep<EP>(p, v[0], pol);
	else if constexpr (V == 2) asm volatile("st.global.L2::cache_hint.v2.f32 [%0], {%1,%2}, %3;" ::"l"(p), "f"(v[0]), "f"(v[1]), "l"(pol));
	else {
#pragma unroll
		for (int k = 0; k < V; k += 4)
			asm volatile("st.global.L2::cache_hint.v4.f32 [%0], {%1,%2,%3,%4}, %5;" ::"l"(p + k), "f"(v[k]), "f"(v[k + 1]), "f"(v[k + 2]), "f"(v[k + 3]), "l"(pol));
	}
}
// V values to columns (j0-1 .. j0+V-2) [DJ < 0, the first one goes to jl] or (j0+1 .. j0+V) [DJ > 0, the last one goes to jr].
// Periodic ends: jl / jr are the wrapped columns.  Open ends (outflow): what would leave the lattice is dropped and the
// end column's site stores a second copy of the value that stays at the end (the zero-gradient end of pull_columns()):
// the slot the wrap would have filled.
template <typename T, int V, int DJ, int EP> __device__ __forceinline__ void st_shifted(T *row, int j0, int jl, int jr, int ny, int outflow, const T (&v)[V], uint64_t pol)
{
	if constexpr (DJ < 0) {
		if (!(outflow && j0 == 0)) st_ep<EP>(row + jl, v[0], pol);
		if (outflow && j0 + V == ny) st_ep<EP>(row + ny - 1, v[V - 1], pol);
		if constexpr (V == 2) st_ep<EP>(row + j0, v[1], pol);
		if constexpr (V >= 4) {
#pragma unroll
			for (int k = 1; k + 1 < V; k += 2) {            // pairs (1,2),(3,4),... start at the even column j0+k-1
				T pr[2] = {v[k], v[k + 1]};
				ste<2, EP>(row + j0 + k - 1, pr, pol);
			}
			st_ep<EP>(row + j0 + V - 2, v[V - 1], pol);
		}
	} else {
		if (!(outflow && j0 + V == ny)) st_ep<EP>(row + jr, v[V - 1], pol);
		if (outflow && j0 == 0) st_ep<EP>(row, v[0], pol);
		if constexpr (V == 2) st_ep<EP>(row + j0 + 1, v[0], pol);
		if constexpr (V >= 4) {
			st_ep<EP>(row + j0 + 1, v[0], pol);
#pragma unroll
			for (int k = 1; k + 1 < V; k += 2) {            // pairs (1,2),(3,4),... land on the even column j0+k+1
				T pr[2] = {v[k], v[k + 1]};
				ste<2, EP>(row + j0 + k + 1, pr, pol);
			}
		}
	}
}

// the nine populations the V sites (r, j0..j0+V-1) relax this step
template <typename T, int V, int PHASE, int EP = 0>
__device__ __forceinline__ void aa_load(const StepParams<T> &p, int r, int j0, T (&f)[9][V], unsigned lanes, uint64_t pol = 0)
{
	const T *const own = p.dst + (size_t)(r + 1) * p.ldy;    // p.src == p.dst
	if constexpr (PHASE == 2) {
#pragma unroll
		for (int s = 0; s < 9; s++) ldc<V, EP>(own + (size_t)opposite(s) * p.plane + j0, f[s], pol);
	} else {
		int jl, jr;
		pull_columns(p, j0, V, jl, jr);
		const T *lrow[3] = {own - p.ldy, own, own + p.ldy};  // rows i-1 (halo when r == 0), i, i+1 (halo when r == nxl-1)
#pragma unroll
		for (int s = 0; s < 9; s++) {
			const T *b = lrow[pull_di(s) + 1] + (size_t)s * p.plane;
			if constexpr (V == 1) {
				f[s][0] = ld_ep<EP>(b + (pull_dj(s) < 0 ? jl : pull_dj(s) > 0 ? jr : j0), pol);
			} else {
				T al[V];
				ldc<V, EP>(b + j0, al, pol);
				if (pull_dj(s) == 0) {
#pragma unroll
					for (int k = 0; k < V; k++) f[s][k] = al[k];
				} else if (pull_dj(s) < 0) {
					f[s][0] = ld_ep<EP>(b + jl, pol);
#pragma unroll
					for (int k = 1; k < V; k++) f[s][k] = al[k - 1];
				} else {
					f[s][V - 1] = ld_ep<EP>(b + jr, pol);
#pragma unroll
					for (int k = 0; k < V - 1; k++) f[s][k] = al[k + 1];
				}
			}
		}
		// Open ends only: the end column's site reads a slot its in-row neighbour also reads and then overwrites (the one
		// slot with two readers).  Both are lanes of one warp (V == 1: adjacent lanes, column ranges start at multiples
		// of 8; V > 1: one thread); order every lane's loads before any lane's stores (`lanes`: the warp's lanes that work).
		if (p.outflow) __syncwarp(lanes);
	}
}

template <typename T, int V, int PHASE, int EP>
__device__ __forceinline__ void aa_store(const StepParams<T> &p, int r, int j0, const T (&o)[9][V], uint64_t pol = 0)
{
	T *const own = p.dst + (size_t)(r + 1) * p.ldy;
	if constexpr (PHASE == 2) {
#pragma unroll
		for (int s = 0; s < 9; s++) ste<V, EP>(own + (size_t)s * p.plane + j0, o[s], pol);
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
		int jl, jr;
		pull_columns(p, j0, V, jl, jr);
		// slot k of site (x - e_k) receives o[opp(k)]; rows beyond the slab are the neighbour's owned boundary rows
		T *srow[3];
		size_t spl[3];
		srow[1] = own;              spl[1] = p.plane;
		if (r == 0) { srow[0] = p.dn_halo - p.ldy; spl[0] = p.dn_plane; } else { srow[0] = own - p.ldy; spl[0] = p.plane; }
		if (r == p.nxl - 1) { srow[2] = p.up_halo + p.ldy; spl[2] = p.up_plane; } else { srow[2] = own + p.ldy; spl[2] = p.plane; }
#pragma unroll
		for (int k = 0; k < 9; k++) {
			T *b = srow[pull_di(k) + 1] + (size_t)k * spl[pull_di(k) + 1];
			if (pull_dj(k) == 0) ste<V, EP>(b + j0, o[opposite(k)], pol);
			else if (pull_dj(k) < 0) st_shifted<T, V, -1, EP>(b, j0, jl, jr, p.ny, p.outflow, o[opposite(k)], pol);
			else st_shifted<T, V, 1, EP>(b, j0, jl, jr, p.ny, p.outflow, o[opposite(k)], pol);
		}
	}
}

// row ranges of the wavefront schedule may wrap (r1 > nxl)
template <typename T> __device__ __forceinline__ int aa_row(const StepParams<T> &p, uint32_t rr)
{
	const int r = p.r0 + (int)rr;
	return r >= p.nxl ? r - p.nxl : r;
}

// One in-place step over rows [r0, r1) x columns [jmin, jmax) (either range may wrap).  2-D grid: x over the vectors
// of a row, y over rows (grid-stride above 65535 rows) -- no integer division per site.
// SP: CTAs of at most 256 threads, four of them per SM -> 64 registers.  Measured on the L2-resident 400x4000 lattice
// (profiles/r01_run24/occ_probe_sp.log): phase 1 at vec 4 drops from 104 registers / 22.4 us to 64 / 17.7 us, the
// run from 113.2 to 119.7 GLUPS.  DP (54-58 registers at vec 1) gains nothing from a cap (profiles/r01_run12).
#ifdef LBM_AA_MINB      /* experiment builds (tools/): 128-thread CTAs, at least LBM_AA_MINB of them per SM */
#define LBM_AA_BOUNDS __launch_bounds__(128, LBM_AA_MINB)
#else
#define LBM_AA_BOUNDS __launch_bounds__(sizeof(T) == 4 ? 256 : 512, sizeof(T) == 4 ? 4 : 0)
#endif
// ---- neighbour barrier between ranks living in different processes / GPUs ----
// The spinners live on DIFFERENT GPUs than the signallers.  flags[3] is the ring's abort word: a wait that gives up after
// ~timeout_ns raises it here AND in both neighbours' flag blocks, every wait returns at once when it finds it raised, so
// one stalled rank stops the whole ring within one time-out instead of one per remaining wait; lbm_sync reports it.
// Returns false if the ring is aborting (the caller then skips its work: the halos it would read are stale).
__device__ __forceinline__ bool seam_wait_flags(volatile int *flags, volatile int *up_abort, volatile int *dn_abort, int epoch, unsigned long long timeout_ns)
{
	if (flags[3]) return false;
	if (flags[0] >= epoch && flags[1] >= epoch) { __threadfence_system(); return true; }
	unsigned long long t0;
	asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
	while (flags[0] < epoch || flags[1] < epoch) {
		unsigned long long t;
		asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
		if (flags[3] || t - t0 > timeout_ns) {
			flags[3] = 1; *up_abort = 1; *dn_abort = 1;
			__threadfence_system();
			return false;
		}
		__nanosleep(100);
	}
	__threadfence_system();
	return true;
}

template <typename T, int V, int PHASE, int HINT, bool SEAM = false>
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
	bool live = jv < nvec;
	unsigned lanes = 0xffffffffu;                            // the lanes of this warp that work (exited ones need not be named)
	if constexpr (SEAM) {                                    // seam launch of a slab: both neighbours must have announced seam_wait
		__shared__ int go;
		if (threadIdx.x == 0) go = seam_wait_flags(p.seam_flags, p.seam_up + 3, p.seam_dn + 2, p.seam_wait, p.seam_timeout_ns);
		__syncthreads();
		live = live && go;
		lanes = __ballot_sync(0xffffffffu, live);            // idle lanes stay for the closing __syncthreads()
	} else if (!live) return;
	const uint64_t pol = l2_policy<HINT>();                  // HINT 1: evict-first stores; HINT 2: evict-last loads and stores
	if (live)
		for (uint32_t rr = blockIdx.y; rr < rows; rr += gridDim.y) {
			const int r = aa_row(p, rr);
			T f[9][V], o[9][V];
			aa_load<T, V, PHASE, HINT == 2 ? 2 : 0>(p, r, j0, f, lanes, pol);
			relax_sites<T, V>(p, f, o, r, j0, p.tail_source);
			aa_store<T, V, PHASE, HINT>(p, r, j0, o, pol);
		}
	if constexpr (SEAM) {                                    // announce: the last CTA to finish publishes this slab's epoch
		__syncthreads();                                         // every store of this CTA is issued ...
		if (threadIdx.x == 0) {
			__threadfence_system();                              // ... and performed (peer stores included) before the count moves
			const unsigned n = gridDim.x * gridDim.y;
			if (atomicAdd(p.seam_done, 1u) == n - 1u) {
				*p.seam_done = 0u;
				__threadfence_system();
				*(volatile int *)p.seam_up = p.seam_signal;
				*(volatile int *)p.seam_dn = p.seam_signal;
				__threadfence_system();
			}
		}
	}
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
		int jl, jr;                                          // open ends (p.outflow): the A/B phase only, lbm_create refuses the rest
		pull_columns(p, j, 1, jl, jr);
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
		if (wall) {
			uint32_t kind = 0u;
			if (p.slip != nullptr) {
				const size_t w = (size_t)r * p.ldw + (j >> 5);
				kind = ((__ldg(p.slip + w) >> (j & 31)) & 1u) | (((__ldg(p.slip + (size_t)p.nxl * p.ldw + w) >> (j & 31)) & 1u) << 1);
			}
			if (kind & 1u) reflect<T, true>(f, o);
			else if (kind & 2u) reflect<T, false>(f, o);
			else bounce<T>(f, o);
		} else {
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
	int nx, channel_walls;                 // channel_walls 2: rows 0 and nx-1 are "reflect" (slip) walls, normal x
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
// one thread per mask word; slip = [2][nxl][ldw] bit planes of the reflect walls (only the channel rows can be one)
__global__ void scene_mask_kernel(uint32_t *mask, uint32_t *slip, SceneDev sc, int row_lo, int nxl, int ny, int ldw)
{
	const size_t g = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (g >= (size_t)nxl * ldw) return;
	const int r = (int)(g / ldw), w = (int)(g % ldw);
	const int gi = row_lo + r;
	uint32_t bits = 0, sx = 0;
	for (int b = 0; b < 32; b++) {
		const int j = w * 32 + b;
		if (j < ny && scene_is_wall(sc, gi, j)) bits |= 1u << b;
	}
	if (sc.channel_walls == 2 && (gi == 0 || gi == sc.nx - 1)) {
		SceneDev inner = sc;                                     // sites of the channel rows that are NOT also barrier / disc sites
		inner.channel_walls = 0;
		for (int b = 0; b < 32; b++) {
			const int j = w * 32 + b;
			if (j < ny && !scene_is_wall(inner, gi, j)) sx |= 1u << b;
		}
	}
	mask[g] = bits;
	slip[g] = sx;
	slip[(size_t)nxl * ldw + g] = 0u;
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

// walls chars [nxl][ldc] -> bit planes: mask = any wall (1, 2, 3); slip[0] = value 2 (reflect, normal x), slip[1] = value 3
// (reflect, normal y).  flags[0] is set if another value is seen, flags[1] if any reflect wall is.
__global__ void mask_pack_kernel(const char *walls, size_t ldc, uint32_t *mask, uint32_t *slip, int nxl, int ny, int ldw, int *flags)
{
	const size_t g = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (g >= (size_t)nxl * ldw) return;
	const int r = (int)(g / ldw), w = (int)(g % ldw);
	uint32_t bits = 0, sx = 0, sy = 0;
	for (int b = 0; b < 32; b++) {
		const int j = w * 32 + b;
		if (j < ny) {
			const char c = walls[(size_t)r * ldc + j];
			if (c == 1 || c == 2 || c == 3) bits |= 1u << b; else if (c != 0) flags[0] = 1;
			if (c == 2) sx |= 1u << b;
			if (c == 3) sy |= 1u << b;
		}
	}
	mask[g] = bits;
	slip[g] = sx;
	slip[(size_t)nxl * ldw + g] = sy;
	if (sx | sy) flags[1] = 1;
}
__global__ void mask_unpack_kernel(const uint32_t *mask, const uint32_t *slip, char *walls, size_t ldc, int nxl, int ny, int ldw)
{
	const size_t g = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (g >= (size_t)nxl * ny) return;
	const int r = (int)(g / ny), j = (int)(g % ny);
	const size_t w = (size_t)r * ldw + (j >> 5);
	char v = (char)((mask[w] >> (j & 31)) & 1u);
	if (v && slip) {
		if ((slip[w] >> (j & 31)) & 1u) v = 2;
		else if ((slip[(size_t)nxl * ldw + w] >> (j & 31)) & 1u) v = 3;
	}
	walls[(size_t)r * ldc + j] = v;
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

// The centre column of macroscopic_kernel for lbm_reynolds_column_static: u_y of column jc in the same operation order, plus
// the fluid flag of each site from the packed wall mask, both into caller-provided device memory (there: the static scratch).
template <typename T>
__global__ void reynolds_column_static_kernel(const T *f, size_t plane, int ldy, int nxl, int jc, const uint32_t *mask, int ldw,
                                              T *uy_o, char *fluid_o)
{
	typedef fp<T> A;
	const int r = (int)(blockIdx.x * blockDim.x + threadIdx.x);
	if (r >= nxl) return;
	const T *site = f + (size_t)(r + 1) * ldy + jc;
	T v[9], rho = T(0);
#pragma unroll
	for (int s = 0; s < 9; s++) { v[s] = site[(size_t)s * plane]; rho = A::add(rho, v[s]); }
	uy_o[r] = A::div(A::sub(A::add(A::add(v[5], v[1]), v[8]), A::add(A::add(v[6], v[3]), v[7])), rho);
	fluid_o[r] = (char)(((mask[(size_t)r * ldw + (jc >> 5)] >> (jc & 31)) & 1u) == 0u);
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
// wait (the unfused form, for the launches that cannot carry the barrier themselves): see seam_wait_flags
__global__ void flag_wait_kernel(int *flags, int *up_abort, int *dn_abort, int epoch, unsigned long long timeout_ns)
{
	seam_wait_flags(flags, up_abort, dn_abort, epoch, timeout_ns);
}

} // namespace lbm
