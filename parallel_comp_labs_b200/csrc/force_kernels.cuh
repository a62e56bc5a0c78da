// force_kernels.cuh -- K1: the all-pairs clamped-softening acceleration for sm_100a.
//
// Replaces the reference's hot loop
//   CPU  lab_4_N_body_MPI/src/include/N_body/struct_ode_system.hpp:60-75
//   GPU  lab_5_N_body_CUDA/src/N_body.cu:134-195 (calc_acceleration)
// a_i = -G * sum_j m_j (r_i - r_j) / max(|r_i - r_j|^3, eps^3), self term included.
// This kernel produces the raw sum over one j-slice; -G and the slice reduction are
// applied (in fixed order) by the fused stage kernel (stage_kernels.cuh).
//
// Design (not a port of calc_acceleration):
//  * bodies are packed {x,y,z,m} (double4 / float4): one array is the j-tile source,
//    the i-body source and the all-gather payload;
//  * persistent CTAs draw work units (i-tile, j-slice) from a ticket counter; a j-slice is a
//    run of TJ-body tiles fetched by ONE elected thread with 1-D bulk-async copies
//    (cp.async.bulk -> SASS UBLKCP) into a STAGES-deep shared-memory ring, completion
//    signalled through mbarrier transaction counts;
//  * every thread keeps I i-bodies and their accumulators in registers; each j body is
//    read from shared memory once per thread (warp-broadcast LDS.128) and reused I times;
//  * the clamp moves onto r^2 (max(d^3,eps^3) == max(d^2,eps^2)^{3/2}) and OFF the hot loop:
//    the hot loop only tracks the smallest key(r^2) it saw (VIMNMX3, one per two pairs); a
//    warp that may have met a pair inside the clamp radius (always: the self pair) restores
//    its accumulators and redoes that tile with the exact 64-bit integer max;
//  * FP64: 1/d^3 from the MUFU.RSQ64H seed y0 with a folded correction
//      e = 1 - x y0^2,  x^{-3/2} = y0^3 (1 + e(3/2 + 15/8 e)) + O(e^3)
//    -> 16 FP64-pipe instructions per pair (3 sub, 3 r^2, 7 inverse-cube*m, 3 FMA), 15 when
//    all masses are equal and the mass is factored out (UM);
//  * FP32: MUFU.RSQ directly (tolerance 1e-4), 13-14 FP32 issue slots per pair.
#pragma once
// NBODY_HOST_EMU: tests/tools/cuda_emu.h runs these kernels lane by lane on the CPU (test
// infrastructure; it supplies the CUDA vocabulary and the PTX helpers below as plain C++).
#ifndef NBODY_HOST_EMU
#include <cuda_runtime.h>
#define NB_DYN_SMEM(name) extern __shared__ __align__(128) unsigned char name[]
#endif
#include <stdint.h>

// Device-side bounds asserts (compute-sanitizer is closed on the GPU pool): with
// -DNBODY_BOUNDS_CHECK every global store / RED / bulk-copy source of K1, K1s and K2 checks its
// index against the allocation and traps (the host then sees a launch failure).  The product
// build compiles them out (SASS unchanged); `make -C parallel_comp_labs_b200/csrc check` builds
// libnbody_b200_check.so, which the -m gpu suite can be pointed at with NBODY_B200_LIB.
#if defined(NBODY_BOUNDS_CHECK) && !defined(NBODY_HOST_EMU)
#define NB_CHECK(cond) do { if (!(cond)) __trap(); } while (0)
#else
#define NB_CHECK(cond) do { } while (0)
#endif

namespace nbody {

template <typename T> struct V4;
template <> struct V4<double> { typedef double4 type; };
template <> struct V4<float> { typedef float4 type; };

// ---------------------------------------------------------------- PTX helpers
#ifndef NBODY_HOST_EMU
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
	return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
	asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_fence_init() {
	asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
	asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
				 "r"(bytes)
				 : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
	uint32_t done;
	do {
		asm volatile(
			"{\n"
			".reg .pred p;\n"
			"mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
			"selp.u32 %0, 1, 0, p;\n"
			"}\n"
			: "=r"(done)
			: "r"(smem_u32(bar)), "r"(parity)
			: "memory");
	} while (!done);
}
// 1-D bulk async copy global -> shared (TMA engine), bytes multiple of 16.
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes,
										 uint64_t* bar) {
	asm volatile(
		"cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::
			"r"(smem_u32(dst)),
		"l"(src), "r"(bytes), "r"(smem_u32(bar))
		: "memory");
}

// ---------------------------------------------------------------- per-pair math
__device__ __forceinline__ double rsqrt_seed(double x) {
	double y;
	asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
	return y;
}
__device__ __forceinline__ float rsqrt_seed(float x) {
	float y;
	asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
	return y;
}
#endif	// !NBODY_HOST_EMU

// Clamp constants prepared on the host.  key(r2) is an int that orders like r2 for r2 >= +0
// (FP64: the high word; FP32: the bit pattern); "key(r2) <= thresh" is a conservative test
// for "r2 may be below eps^2".
template <typename T> struct Clamp;
template <> struct Clamp<double> {
	long long eps2_bits;
	int thresh;	 // high word of eps^2
	__device__ __forceinline__ double apply(double r2) const {
		// r2 is a sum of squares (>= +0): integer max on the bit patterns == fmax.
		return __longlong_as_double(max(__double_as_longlong(r2), eps2_bits));
	}
	static __device__ __forceinline__ int key(double r2) { return __double2hiint(r2); }
};
template <> struct Clamp<float> {
	float eps2;
	int thresh;	 // bit pattern of eps^2
	__device__ __forceinline__ float apply(float r2) const { return fmaxf(r2, eps2); }
	static __device__ __forceinline__ int key(float r2) { return __float_as_int(r2); }
};

// w = m * x^{-3/2}, x = max(r2,eps2).  ORD = 2: second-order correction of the seed,
// truncation error (35/16)|e|^3 < 2^-55 (full double accuracy, ~2-3 ulp with rounding);
// ORD = 1: first-order only, error 1.875 e^2 (|e| measured < 2^-19.04: <= 6.5e-12 per pair,
// 2.5e-13 on a whole RHS); selectable, not the default.
// UM (uniform mass): every body has the same mass, which is then factored out of the pair
// sum (a_i = -G m sum_j (r_i-r_j)/den, applied once in K2): one FP64 instruction less.
template <int ORD, bool UM>
__device__ __forceinline__ double inv_cube_m(double x, double m) {
	double y0 = rsqrt_seed(x);				 // MUFU.RSQ64H
	double t = __dmul_rn(y0, y0);			 // y0^2
	double e = __fma_rn(-x, t, 1.0);		 // 1 - x y0^2
	double u = __dmul_rn(y0, t);			 // y0^3
	double um = UM ? u : __dmul_rn(u, m);	 // m y0^3
	// (1 + e p) first, then one multiply: no DFMA with three distinct register operands
	// (those need a third register-file cycle: two banks, 64-bit operands straddle both)
	if (ORD == 1) return __dmul_rn(um, __fma_rn(e, 1.5, 1.0));	// m y0^3 (1 + 3/2 e)
	double p = __fma_rn(e, 1.875, 1.5);		 // 3/2 + 15/8 e
	double q = __fma_rn(e, p, 1.0);			 // 1 + e p
	return __dmul_rn(um, q);				 // m y0^3 (1 + e p)
}
template <int ORD, bool UM>
__device__ __forceinline__ float inv_cube_m(float x, float m) {
	float y = rsqrt_seed(x);  // MUFU.RSQ
	float y2 = y * y;
	float my = UM ? y : m * y;
	return my * y2;
}

template <typename T, int I>
struct Acc {
	T x[I], y[I], z[I];
};

// One j-body against the I i-bodies of this thread.
//  EXACT = false: the hot path.  No clamp at all; instead the smallest key(r2) seen is
//    tracked (one VIMNMX3 per two pairs).  Measured on B200: a warp-instruction on the
//    FP64 pipe costs 2 issue cycles and every other instruction 1 more, additively, so the
//    exact 64-bit clamp (2 ISETP + 2 SEL per pair) is worth taking off the common path.
//  EXACT = true: the reference's clamp applied to every pair (cold path; the tile is redone
//    with it whenever the tracked minimum says some pair -- always including the self pair,
//    r2 == 0 -- may lie inside the clamp radius).
template <typename T, int I, int ORD, bool UM, bool EXACT>
__device__ __forceinline__ void interact(const typename V4<T>::type pj, const T (&xi)[I],
										  const T (&yi)[I], const T (&zi)[I], Acc<T, I>& a,
										  const Clamp<T>& cl, int& mn) {
	int key[I];
#pragma unroll
	for (int k = 0; k < I; ++k) {
		T dx = xi[k] - pj.x;  // r_i - r_j, the reference's sign (struct_ode_system.hpp:65)
		T dy = yi[k] - pj.y;
		T dz = zi[k] - pj.z;
		T r2 = fma(dz, dz, fma(dy, dy, dx * dx));
		if (EXACT) r2 = cl.apply(r2);
		else key[k] = Clamp<T>::key(r2);
		T w = inv_cube_m<ORD, UM>(r2, pj.w);
		a.x[k] = fma(w, dx, a.x[k]);
		a.y[k] = fma(w, dy, a.y[k]);
		a.z[k] = fma(w, dz, a.z[k]);
	}
	if (!EXACT) {
#pragma unroll
		for (int k = 0; k + 1 < I; k += 2) mn = min(mn, min(key[k], key[k + 1]));
		if (I & 1) mn = min(mn, key[I - 1]);
	}
}

// ---------------------------------------------------------------- per-thread state ("lane")
// Scalar lane: I i-bodies and their accumulators in registers, one FP instruction per pair.
template <typename T, int I, int ORD, bool UM>
struct LaneScalar {
	T xi[I], yi[I], zi[I];
	Acc<T, I> a;
	__device__ __forceinline__ void init(const T (&x)[I], const T (&y)[I], const T (&z)[I]) {
#pragma unroll
		for (int k = 0; k < I; ++k) {
			xi[k] = x[k]; yi[k] = y[k]; zi[k] = z[k];
			a.x[k] = a.y[k] = a.z[k] = T(0);
		}
	}
	template <bool EXACT>
	__device__ __forceinline__ void pair(const typename V4<T>::type pj, const Clamp<T>& cl, int& mn) {
		interact<T, I, ORD, UM, EXACT>(pj, xi, yi, zi, a, cl, mn);
	}
	__device__ __forceinline__ void result(T (&x)[I], T (&y)[I], T (&z)[I]) const {
#pragma unroll
		for (int k = 0; k < I; ++k) { x[k] = a.x[k]; y[k] = a.y[k]; z[k] = a.z[k]; }
	}
};

// (A packed FP32 lane built on Blackwell's f32x2 instructions -- SASS FADD2/FMUL2/FFMA2, two
// i-bodies per 64-bit register pair -- was measured in round 1: 2150-2280 G int/s against
// 2340-2390 for the scalar lane at N = 1M; the packed instructions do not relieve the issue
// port and the extra pack MOVs cost more than they save.  Not kept.)

// ---------------------------------------------------------------- launch parameters
struct ForceParams {
	// Positions {x,y,z,m} as one block of `jt_per_rank` j-tiles per rank.  With NCCL all-gathers
	// all blocks live in this GPU's own array (blk[r] = pos + r*block); with the peer-pull
	// exchange (nbody_p2p_connect) blk[r] for r != rank points INTO RANK r's HBM over NVLink and
	// the bulk-async copies fetch the remote tiles directly: no separate collective.
	const void* blk[8];
	const void* own;	   // this rank's block (i-body source)
	unsigned contiguous;   // 1: all blocks are consecutive in ONE array starting at blk[0] (NCCL
						   //    exchange, any number of ranks); 0: per-rank table (peer-pull, <= 8)
	unsigned jt_per_rank;  // j-tiles per block
	unsigned rank;
	void* apart;		   // V4[n_slices_total][own_pad] raw partial sums (w unused)
	unsigned own0;		   // first owned global row
	unsigned own_pad;	   // padded owned rows (row stride of apart)
	unsigned n_itiles;	   // i-tiles of THREADS*I rows covering the real owned rows
	unsigned jt_begin;	   // first j-tile (absolute), wraps modulo jt_total
	unsigned jt_count;	   // j-tiles in this launch
	unsigned jt_total;	   // n_pad / TJ
	unsigned n_slices;	   // j-slices in this launch
	unsigned slice_base;   // first slice row of apart written by this launch
	unsigned n_real;	   // bodies that exist; rows >= n_real of `pos` are padding
	unsigned* counter;	   // work-unit ticket counter, zeroed before the launch
	unsigned apart_rows;   // rows allocated in apart (bounds asserts only)
};

template <typename T> struct Tile {
	static constexpr int TJ = 256;
};

// global address of j-tile `jt` (absolute tile index): block of its owner + offset inside it
template <typename vec4, int TJ>
__device__ __forceinline__ const vec4* tile_src(const ForceParams& p, unsigned jt) {
	NB_CHECK(jt < p.jt_total);
	if (p.contiguous) return reinterpret_cast<const vec4*>(p.blk[0]) + (size_t)jt * TJ;
	const unsigned r = jt / p.jt_per_rank;
	return reinterpret_cast<const vec4*>(p.blk[r]) + (size_t)(jt - r * p.jt_per_rank) * TJ;
}

// ---------------------------------------------------------------- the kernel
template <typename T, int THREADS, int I, int STAGES, int MINB, int UM, int UNROLL, int ORD>
__global__ void __launch_bounds__(THREADS, MINB)
	force_kernel(const ForceParams p, const Clamp<T> cl) {
	typedef typename V4<T>::type vec4;
	constexpr int TJ = Tile<T>::TJ;
	constexpr uint32_t TILE_BYTES = TJ * sizeof(vec4);

	NB_DYN_SMEM(smem_raw);
	vec4* tiles = reinterpret_cast<vec4*>(smem_raw);
	uint64_t* full = reinterpret_cast<uint64_t*>(smem_raw + (size_t)STAGES * TILE_BYTES);
	volatile unsigned* next_unit_s = reinterpret_cast<volatile unsigned*>(full + STAGES);

	const int tid = threadIdx.x;
	const vec4* __restrict__ own = reinterpret_cast<const vec4*>(p.own);
	vec4* __restrict__ apart = reinterpret_cast<vec4*>(p.apart);

	if (tid == 0) {
#pragma unroll
		for (int s = 0; s < STAGES; ++s) mbar_init(&full[s], 1);
		mbar_fence_init();
	}
	__syncthreads();

	uint32_t ring = 0;	// tiles consumed so far by this CTA (== tiles issued - in flight)
	const unsigned n_units = p.n_itiles * p.n_slices;

	// Dynamic unit scheduling: the warp schedulers favour some resident CTAs over others
	// (ncu: with a static round-robin walk the average residency fell to 2.1 of 3 warps per
	// SMSP because favoured CTAs finished early), so every CTA draws tickets until the pool
	// is empty; the tail is at most one unit long.
	unsigned unit = blockIdx.x;
	while (unit < n_units) {
		unsigned next_ticket = 0;
		if (tid == 0) next_ticket = gridDim.x + atomicAdd(p.counter, 1u);  // consumed at the end
		const unsigned itile = unit / p.n_slices;
		const unsigned slice = unit - itile * p.n_slices;
		const unsigned t0 = (unsigned)(((unsigned long long)slice * p.jt_count) / p.n_slices);
		const unsigned t1 = (unsigned)(((unsigned long long)(slice + 1) * p.jt_count) / p.n_slices);
		const unsigned ntiles = t1 - t0;

		// prologue: the ring is empty here (previous unit ended with __syncthreads)
		if (tid == 0) {
			const unsigned pre = ntiles < (unsigned)STAGES ? ntiles : (unsigned)STAGES;
			for (unsigned k = 0; k < pre; ++k) {
				const unsigned st = (ring + k) % STAGES;
				unsigned jt = p.jt_begin + t0 + k;
				if (jt >= p.jt_total) jt -= p.jt_total;
				mbar_expect_tx(&full[st], TILE_BYTES);
				bulk_g2s(tiles + (size_t)st * TJ, tile_src<vec4, TJ>(p, jt), TILE_BYTES, &full[st]);
			}
		}

		// i-bodies of this thread: rows own0 + itile*THREADS*I + k*THREADS + tid
		LaneScalar<T, I, ORD, UM != 0> lane;
		const unsigned row0 = itile * (THREADS * I) + tid;
		{
			T bx[I], by[I], bz[I];
#pragma unroll
			for (int k = 0; k < I; ++k) {
				// the last i-tile may reach past the owned block: clamp the load, drop the store
				const unsigned row = min(row0 + k * THREADS, p.own_pad - 1u);
				const vec4 b = own[row];
				bx[k] = b.x;
				by[k] = b.y;
				bz[k] = b.z;
			}
			lane.init(bx, by, bz);
		}

		for (unsigned t = 0; t < ntiles; ++t) {
			const unsigned st = ring % STAGES;
			mbar_wait(&full[st], (ring / STAGES) & 1u);
			const vec4* tile = tiles + (size_t)st * TJ;
			// bodies of this tile that exist (only the last tile of the array is ragged; for the
			// reference's 4-body problem that is 4 of 256, which is what keeps a step at
			// launch latency instead of 250 wasted iterations)
			unsigned jt_here = p.jt_begin + t0 + t;
			if (jt_here >= p.jt_total) jt_here -= p.jt_total;
			const unsigned j_first = jt_here * TJ;
			const int cnt = j_first >= p.n_real ? 0 : (int)min((unsigned)TJ, p.n_real - j_first);
			const auto saved = lane;
			int mn = 0x7fffffff;
			if (cnt == TJ) {  // full tile: compile-time trip count for the hot loop
#pragma unroll UNROLL
				for (int jj = 0; jj < TJ; ++jj) lane.template pair<false>(tile[jj], cl, mn);
			} else {
#pragma unroll 2
				for (int jj = 0; jj < cnt; ++jj) lane.template pair<false>(tile[jj], cl, mn);
			}
			if (__any_sync(0xffffffffu, mn <= cl.thresh)) {	 // rare: redo this tile exactly
				lane = saved;
#pragma unroll 2
				for (int jj = 0; jj < cnt; ++jj) lane.template pair<true>(tile[jj], cl, mn);
			}
			++ring;
			__syncthreads();  // every warp is done with stage st -> it may be refilled
			if (tid == 0 && t + STAGES < ntiles) {
				unsigned jt = p.jt_begin + t0 + t + STAGES;
				if (jt >= p.jt_total) jt -= p.jt_total;
				mbar_expect_tx(&full[st], TILE_BYTES);
				bulk_g2s(tiles + (size_t)st * TJ, tile_src<vec4, TJ>(p, jt), TILE_BYTES, &full[st]);
			}
		}

		vec4* out = apart + (size_t)(p.slice_base + slice) * p.own_pad;
		T rx[I], ry[I], rz[I];
		lane.result(rx, ry, rz);
#pragma unroll
		for (int k = 0; k < I; ++k) {
			vec4 o;
			o.x = rx[k];
			o.y = ry[k];
			o.z = rz[k];
			o.w = T(0);
			if (row0 + k * THREADS < p.own_pad) {
				NB_CHECK((size_t)(p.slice_base + slice) * p.own_pad + row0 + k * THREADS < p.apart_rows);
				out[row0 + k * THREADS] = o;
			}
		}
		if (tid == 0) *next_unit_s = next_ticket;
		__syncthreads();
		unit = *next_unit_s;
		__syncthreads();  // everyone has read it before thread 0 can overwrite it
	}
}

template <typename T, int STAGES>
constexpr size_t force_smem_bytes() {
	return (size_t)STAGES * Tile<T>::TJ * sizeof(typename V4<T>::type) + STAGES * sizeof(uint64_t) + 16;
}

}  // namespace nbody
