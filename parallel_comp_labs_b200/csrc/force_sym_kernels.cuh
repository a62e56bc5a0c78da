// force_sym_kernels.cuh -- K1s: the pair-symmetric kernel of the all-pairs acceleration (the default
// for n >= 2048, option "pair_symmetry"; on several GPUs the ranks split the blocks of one schedule).
//
// Same operator as force_kernels.cuh (reference: lab_4 .../struct_ode_system.hpp:60-75,
// lab_5 N_body.cu:134-195), same per-pair arithmetic, but every pair of DISTINCT bodies is
// evaluated once and credited to both (a_i += m_j w d, a_j -= m_i w d, d = r_i - r_j): the
// differences, r^2 and the inverse cube -- 12 of the 15 FP64 instructions of a pair -- are shared
// by the two interactions, 18 FP64 instructions per pair evaluation = 9 per delivered
// interaction (20 = 10 with general masses).  The reference evaluates both ordered pairs
// separately; (r_j - r_i) == -(r_i - r_j) exactly in floating point, so each pair contributes
// the same bits as in the one-sided kernel and only the summation order differs.
//
// How the reaction force is collected without shared-memory traffic in the hot loop:
//  * a CTA owns one block of THREADS*I bodies as i-bodies (I per thread, in registers) and walks
//    the j-tiles of its partner blocks (sym_schedule.h), staged by the same bulk-async ring as K1;
//  * a tile is consumed in strips of 32 j-bodies.  Lane l of a warp starts with j-body l of the
//    strip AND its accumulator in registers; after every step (I pair evaluations per lane) the
//    j-body and its accumulator ROTATE to the next lane (warp shuffles), so that after 32 steps
//    every i-body of the warp has met every j-body of the strip and each j accumulator has
//    collected the reaction of all 32*I i-bodies and is back in its home lane.  It is then added
//    to the j-side sum array with one RED per component (fire-and-forget atomics: the summation
//    ORDER of the j-side contributions depends on scheduling, so results are reproducible only
//    to rounding, unlike the one-sided kernel -- or, ACC = 2, as exact fixed-point integers);
//  * the block's own tiles are met one-sidedly with the exact clamp (this is where the self
//    pairs are), exactly as K1 does;
//  * clamp: as in K1 the hot loop is unclamped and tracks the smallest key(r^2); a strip that may
//    contain a pair inside the clamp radius is redone with the exact clamp.  Strip results live
//    in strip-local accumulators (zeroed per strip) and are committed to the running i-side sums
//    in shared memory afterwards, so a redo needs no backup registers.
#pragma once
#include "fixed_sums.cuh"
#include "force_kernels.cuh"
#include "sym_schedule.h"

namespace nbody {

// K1s waits on the tile ring through a watchdog: a tile that has not landed 10 s after the wait
// began is a protocol error (it cannot happen with the schedule in sym_schedule.h; the CPU
// execution of this file checks the ring for deadlocks), and the kernel then traps so that the
// host sees a launch failure instead of a hung device.  The fast path is the same single
// try_wait as K1's mbar_wait; the timer is read only after 4096 failed polls.
#ifndef NBODY_HOST_EMU
__device__ __forceinline__ void mbar_wait_guarded(uint64_t* bar, uint32_t parity) {
	uint32_t done, polls = 0;
	unsigned long long t_begin = 0;
	for (;;) {
		asm volatile(
			"{\n"
			".reg .pred p;\n"
			"mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
			"selp.u32 %0, 1, 0, p;\n"
			"}\n"
			: "=r"(done)
			: "r"(smem_u32(bar)), "r"(parity)
			: "memory");
		if (done) return;
		if ((++polls & 4095u) == 0) {
			unsigned long long now;
			asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
			if (t_begin == 0) t_begin = now;
			else if (now - t_begin > 10000000000ull) __trap();
		}
	}
}
#else
inline void mbar_wait_guarded(uint64_t* bar, uint32_t parity) { mbar_wait(bar, parity); }
#endif

struct SymParams {
	const void* pos;	 // V4[(jt_total + tpb) * TJ] positions {x,y,z,m}; rows >= n_pad unspecified
	void* aj;			 // T[3][rows_pad] acceleration sums (global row index), zeroed before the launch:
						 //   BOTH sides of every pair are added here with RED.ADD -- the i-side sum of a
						 //   unit once at its end, the j-side sum of a strip once per warp and strip
	unsigned rows_pad;	 // nb * BLOCK, BLOCK = THREADS * I
	unsigned nb;		 // blocks of the whole problem
	unsigned block0;	 // first block this GPU owns (multi-GPU: the ranks split the blocks, every
	unsigned nb_own;	 //   rank walks the same global schedule, j-side sums are reduce-scattered)
	unsigned jt_total;	 // tiles that exist (n_pad / TJ)
	unsigned n_pad;		 // rows of pos that are initialised (real bodies + pad bodies)
	unsigned n_real;	 // bodies that exist
	unsigned n_slices;	 // slices of a block's walk
	unsigned* counter;	 // work-unit ticket counter, zeroed before the launch
	double pad_pos;		 // coordinates given to i-rows >= n_pad (see pack_aos_kernel)
	unsigned pos_rows;	 // rows allocated in pos (bounds asserts only)
	int fx_e0;			 // ACC = 2 only: exponent of the lowest bit of the fixed-point row sums; `aj` is
						 //   then uint64[FX_PLANES][rows_pad] (fixed_sums.cuh) instead of T[3][rows_pad]
};

// One j-body (travelling with its accumulator) against the I i-bodies of this thread.
// UM: all masses equal and factored out (applied once in K2); else m_j weights the action on i
// and m_i the reaction on j.
__device__ __forceinline__ double sym_mul(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ double sym_add(double a, double b) { return __dadd_rn(a, b); }
__device__ __forceinline__ float sym_mul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float sym_add(float a, float b) { return __fadd_rn(a, b); }

// ACC = 0 (every shipped build): both sides accumulate with FMA (a += w d).
// ACC = 1: f = w d once (3 multiplies), then a_i += f, a_j -= f (6 adds): 9 two-operand instructions
// instead of 6 three-operand FMAs.  Built and measured in round 2 to test whether the FMAs' third
// register operand is what costs the FP64 pipe its last 17 %: it is not -- 21 instructions per
// evaluation ran at 2.26 cycles each (1565 vs 1718 G int/s, profiles/r02_experiment_acc_muladd.jsonl),
// the 18-instruction build at 2.40: the pipe delivers ~84-88 % for this instruction stream either
// way.  The template parameter stays for the record; no variant table entry uses it.
// ACC = 2 (option "deterministic"): the pair arithmetic of ACC = 0; what changes is how a unit's
// sums reach the global row sums -- as exact fixed-point integers (fixed_sums.cuh) instead of
// floating-point RED, so the result does not depend on the order in which units finish.
template <typename T, int I, int ORD, bool UM, bool EXACT, int ACC = 0>
__device__ __forceinline__ void interact_sym(const T jx, const T jy, const T jz, const T jm,
											  const T (&xi)[I], const T (&yi)[I], const T (&zi)[I],
											  const T (&mi)[I], Acc<T, I>& a, T& ajx, T& ajy, T& ajz,
											  const Clamp<T>& cl, int& mn) {
	int key[I];
#pragma unroll
	for (int k = 0; k < I; ++k) {
		const T dx = xi[k] - jx;  // r_i - r_j, the reference's sign (struct_ode_system.hpp:65)
		const T dy = yi[k] - jy;
		const T dz = zi[k] - jz;
		T r2 = fma(dz, dz, fma(dy, dy, dx * dx));
		if (EXACT) r2 = cl.apply(r2);
		else key[k] = Clamp<T>::key(r2);
		const T w = inv_cube_m<ORD, true>(r2, T(1));  // max(r2,eps2)^{-3/2}
		const T wi = UM ? w : w * jm;
		const T wj = UM ? w : w * mi[k];
		if (ACC == 1 && UM) {
			// explicit roundings: the compiler must not contract these back into FMAs
			const T fx = sym_mul(w, dx), fy = sym_mul(w, dy), fz = sym_mul(w, dz);
			a.x[k] = sym_add(a.x[k], fx); a.y[k] = sym_add(a.y[k], fy); a.z[k] = sym_add(a.z[k], fz);
			ajx = sym_add(ajx, -fx); ajy = sym_add(ajy, -fy); ajz = sym_add(ajz, -fz);
		} else {
			a.x[k] = fma(wi, dx, a.x[k]);
			a.y[k] = fma(wi, dy, a.y[k]);
			a.z[k] = fma(wi, dz, a.z[k]);
			ajx = fma(-wj, dx, ajx);  // force on j = -(force on i) per unit of the other's mass
			ajy = fma(-wj, dy, ajy);
			ajz = fma(-wj, dz, ajz);
		}
	}
	if (!EXACT) {
#pragma unroll
		for (int k = 0; k + 1 < I; k += 2) mn = min(mn, min(key[k], key[k + 1]));
		if (I & 1) mn = min(mn, key[I - 1]);
	}
}

// One strip: 32 rotation steps.  On return `loc` holds the strip's contribution to this thread's
// i-bodies and (ajx,ajy,ajz) the strip's contribution to the j-body this lane STARTED with.
// ROT = 0: the j-body travels with its accumulator (6 shuffled doubles per step, no shared-memory
// read in the loop).  ROT = 1: only the accumulator travels; lane l re-reads the j-body that is
// visiting it, strip[(l + r + 1) & 31], from the staged tile (distinct consecutive addresses:
// conflict-free; 2 LDS instead of 6 SHFL per step).  Same arithmetic, same bits.
template <typename T, int I, int ORD, bool UM, bool EXACT, bool ROT, int ACC = 0>
__device__ __forceinline__ void sym_strip(const typename V4<T>::type* strip, const int lane,
										   const T (&xi)[I],
										   const T (&yi)[I], const T (&zi)[I], const T (&mi)[I],
										   Acc<T, I>& loc, T& ajx, T& ajy, T& ajz,
										   const Clamp<T>& cl, int& mn, const int next_lane) {
	const typename V4<T>::type pj = strip[lane];
	T jx = pj.x, jy = pj.y, jz = pj.z, jm = pj.w;
	ajx = ajy = ajz = T(0);
#pragma unroll
	for (int k = 0; k < I; ++k) loc.x[k] = loc.y[k] = loc.z[k] = T(0);
	mn = 0x7fffffff;
#pragma unroll 2
	for (int r = 0; r < 32; ++r) {
		interact_sym<T, I, ORD, UM, EXACT, ACC>(jx, jy, jz, jm, xi, yi, zi, mi, loc, ajx, ajy, ajz, cl, mn);
		// pass the j-body and its accumulator on; after 32 passes they are home again
		if (ROT) {
			const typename V4<T>::type q = strip[(lane + r + 1) & 31];
			jx = q.x; jy = q.y; jz = q.z;
			if (!UM) jm = q.w;
		} else {
			jx = __shfl_sync(0xffffffffu, jx, next_lane);
			jy = __shfl_sync(0xffffffffu, jy, next_lane);
			jz = __shfl_sync(0xffffffffu, jz, next_lane);
			if (!UM) jm = __shfl_sync(0xffffffffu, jm, next_lane);
		}
		ajx = __shfl_sync(0xffffffffu, ajx, next_lane);
		ajy = __shfl_sync(0xffffffffu, ajy, next_lane);
		ajz = __shfl_sync(0xffffffffu, ajz, next_lane);
	}
}

// ACC = 2: one row's share of a unit (or of a warp's strip) goes to the fixed-point row sums.
template <typename T>
__device__ __forceinline__ void sym_commit_fx(const SymParams& p, const unsigned row, const T ax,
											   const T ay, const T az) {
	unsigned long long* planes = reinterpret_cast<unsigned long long*>(p.aj);
	const size_t stride = p.rows_pad;
	unsigned long long* poison = planes + (size_t)(3 * FX_LIMBS) * stride + row;
	fx_add(planes + row, stride, poison, (double)ax, p.fx_e0);
	fx_add(planes + (size_t)FX_LIMBS * stride + row, stride, poison, (double)ay, p.fx_e0);
	fx_add(planes + (size_t)(2 * FX_LIMBS) * stride + row, stride, poison, (double)az, p.fx_e0);
}

template <typename T, int THREADS, int I, int STAGES>
constexpr size_t force_sym_smem_bytes() {
	return (size_t)STAGES * Tile<T>::TJ * sizeof(typename V4<T>::type) +
		   (size_t)3 * I * THREADS * sizeof(T) + STAGES * sizeof(uint64_t) + 16;
}

template <typename T, int THREADS, int I, int STAGES, int MINB, int UM, int ORD, int ROT = 0, int ACC = 0>
__global__ void __launch_bounds__(THREADS, MINB)
	force_sym_kernel(const SymParams p, const Clamp<T> cl) {
	typedef typename V4<T>::type vec4;
	constexpr int TJ = Tile<T>::TJ;
	constexpr int BLOCK = THREADS * I;
	static_assert(BLOCK % TJ == 0, "a block must be a whole number of j-tiles");
	constexpr unsigned TPB = BLOCK / TJ;
	constexpr unsigned SPT = TJ / 32;  // strips per tile
	constexpr uint32_t TILE_BYTES = TJ * sizeof(vec4);
	constexpr size_t ACC_BYTES = (size_t)3 * I * THREADS * sizeof(T);

	NB_DYN_SMEM(smem_raw);
	vec4* tiles = reinterpret_cast<vec4*>(smem_raw);
	T* accm = reinterpret_cast<T*>(smem_raw + (size_t)STAGES * TILE_BYTES);	 // [3*I][THREADS]
	uint64_t* full = reinterpret_cast<uint64_t*>(smem_raw + (size_t)STAGES * TILE_BYTES + ACC_BYTES);
	volatile unsigned* next_unit_s = reinterpret_cast<volatile unsigned*>(full + STAGES);

	const int tid = threadIdx.x;
	const int next_lane = (tid + 1) & 31;
	const vec4* __restrict__ pos = reinterpret_cast<const vec4*>(p.pos);
	T* __restrict__ gjx = reinterpret_cast<T*>(p.aj);
	T* __restrict__ gjy = gjx + p.rows_pad;
	T* __restrict__ gjz = gjy + p.rows_pad;

	if (tid == 0) {
#pragma unroll
		for (int s = 0; s < STAGES; ++s) mbar_init(&full[s], 1);
		mbar_fence_init();
	}
	__syncthreads();

	uint32_t ring = 0;	// tiles consumed so far by this CTA
	const unsigned n_units = p.nb_own * p.n_slices;

	unsigned unit = blockIdx.x;
	while (unit < n_units) {
		unsigned next_ticket = 0;
		if (tid == 0) next_ticket = gridDim.x + atomicAdd(p.counter, 1u);  // consumed at the end
		const unsigned a_loc = unit / p.n_slices;
		const unsigned slice = unit - a_loc * p.n_slices;
		const unsigned a = p.block0 + a_loc;
		unsigned q0, q1, t0, t1;  // strips of the walk this unit takes, and the tiles they lie in
		sym_slice_range(sym_ntiles(p.nb, TPB, a), SPT, p.n_slices, slice, &q0, &q1);
		sym_slice_tiles(SPT, q0, q1, &t0, &t1);
		const unsigned ntiles = t1 - t0;

		// prologue: the ring is empty here (previous unit ended with __syncthreads)
		if (tid == 0) {
			const unsigned pre = ntiles < (unsigned)STAGES ? ntiles : (unsigned)STAGES;
			for (unsigned k = 0; k < pre; ++k) {
				const unsigned st = (ring + k) % STAGES;
				bool d;
				const unsigned jt = sym_tile(p.nb, TPB, a, t0 + k, &d);
				NB_CHECK((size_t)(jt + 1) * TJ <= p.pos_rows);
				mbar_expect_tx(&full[st], TILE_BYTES);
				bulk_g2s(tiles + (size_t)st * TJ, pos + (size_t)jt * TJ, TILE_BYTES, &full[st]);
			}
		}

		// i-bodies of this thread: rows a*BLOCK + k*THREADS + tid; rows past the initialised
		// array are ghosts that act on nothing (m = 0 far away / at the pad position)
		T xi[I], yi[I], zi[I], mi[I];
		const unsigned row0 = a * BLOCK + tid;
#pragma unroll
		for (int k = 0; k < I; ++k) {
			const unsigned row = row0 + k * THREADS;
			if (row < p.n_pad) {
				const vec4 b = pos[row];
				xi[k] = b.x; yi[k] = b.y; zi[k] = b.z; mi[k] = b.w;
			} else {
				xi[k] = yi[k] = zi[k] = (T)p.pad_pos;
				mi[k] = T(0);
			}
			accm[(0 * I + k) * THREADS + tid] = T(0);
			accm[(1 * I + k) * THREADS + tid] = T(0);
			accm[(2 * I + k) * THREADS + tid] = T(0);
		}

		for (unsigned t = 0; t < ntiles; ++t) {
			const unsigned st = ring % STAGES;
			mbar_wait_guarded(&full[st], (ring / STAGES) & 1u);
			const vec4* tile = tiles + (size_t)st * TJ;
			bool diag;
			const unsigned jt = sym_tile(p.nb, TPB, a, t0 + t, &diag);
			const unsigned j_first = jt * TJ;
			// bodies of this tile that exist (0 for pad-only and phantom tiles)
			int cnt = j_first >= p.n_real ? 0 : (int)min((unsigned)TJ, p.n_real - j_first);
			unsigned part_lo, part_hi;	// the bodies of this tile that belong to the unit's strips
			sym_tile_part(t0 + t, SPT, q0, q1, &part_lo, &part_hi);
			const int lo = (int)part_lo;
			cnt = min(cnt, (int)part_hi);
			Acc<T, I> loc;
			int mn;
			if (diag) {
				// the block against itself: one-sided, exact clamp (self pairs live here)
#pragma unroll
				for (int k = 0; k < I; ++k) loc.x[k] = loc.y[k] = loc.z[k] = T(0);
#pragma unroll 2
				for (int jj = lo; jj < cnt; ++jj)
					interact<T, I, ORD, UM != 0, true>(tile[jj], xi, yi, zi, loc, cl, mn);
#pragma unroll
				for (int k = 0; k < I; ++k) {
					accm[(0 * I + k) * THREADS + tid] += loc.x[k];
					accm[(1 * I + k) * THREADS + tid] += loc.y[k];
					accm[(2 * I + k) * THREADS + tid] += loc.z[k];
				}
			} else {
				for (int s0 = lo; s0 < cnt; s0 += 32) {	// strips that hold a real body
					T ajx, ajy, ajz;
					sym_strip<T, I, ORD, UM != 0, false, ROT != 0, ACC>(tile + s0, tid & 31, xi, yi, zi, mi,
																		loc, ajx, ajy, ajz, cl, mn, next_lane);
					if (__any_sync(0xffffffffu, mn <= cl.thresh))  // rare: redo the strip exactly
						sym_strip<T, I, ORD, UM != 0, true, ROT != 0, ACC>(tile + s0, tid & 31, xi, yi, zi,
																		   mi, loc, ajx, ajy, ajz, cl, mn,
																		   next_lane);
					const unsigned j = j_first + (unsigned)s0 + (unsigned)(tid & 31);
					NB_CHECK(j < p.rows_pad);
					if (ACC == 2) {
						sym_commit_fx<T>(p, j, ajx, ajy, ajz);
					} else {
						atomicAdd(gjx + j, ajx);  // result unused -> RED.ADD
						atomicAdd(gjy + j, ajy);
						atomicAdd(gjz + j, ajz);
					}
#pragma unroll
					for (int k = 0; k < I; ++k) {
						accm[(0 * I + k) * THREADS + tid] += loc.x[k];
						accm[(1 * I + k) * THREADS + tid] += loc.y[k];
						accm[(2 * I + k) * THREADS + tid] += loc.z[k];
					}
				}
			}
			++ring;
			__syncthreads();  // every warp is done with stage st -> it may be refilled
			if (tid == 0 && t + STAGES < ntiles) {
				bool d;
				const unsigned nt = sym_tile(p.nb, TPB, a, t0 + t + STAGES, &d);
				NB_CHECK((size_t)(nt + 1) * TJ <= p.pos_rows);
				mbar_expect_tx(&full[st], TILE_BYTES);
				bulk_g2s(tiles + (size_t)st * TJ, pos + (size_t)nt * TJ, TILE_BYTES, &full[st]);
			}
		}

		// i-side sums of this unit: into the same global rows the j-side reactions go to (the
		// summation order over units and strips is scheduling-dependent either way, so the slices
		// of a block's walk need no partial-sum rows of their own and can be as fine as the tail
		// of the launch wants them)
#pragma unroll
		for (int k = 0; k < I; ++k) {
			const unsigned row = row0 + k * THREADS;	 // < rows_pad = nb * BLOCK
			NB_CHECK(row < p.rows_pad);
			if (ACC == 2) {
				sym_commit_fx<T>(p, row, accm[(0 * I + k) * THREADS + tid],
								 accm[(1 * I + k) * THREADS + tid], accm[(2 * I + k) * THREADS + tid]);
			} else {
				atomicAdd(gjx + row, accm[(0 * I + k) * THREADS + tid]);  // result unused -> RED.ADD
				atomicAdd(gjy + row, accm[(1 * I + k) * THREADS + tid]);
				atomicAdd(gjz + row, accm[(2 * I + k) * THREADS + tid]);
			}
		}
		if (tid == 0) *next_unit_s = next_ticket;
		__syncthreads();
		unit = *next_unit_s;
		__syncthreads();  // everyone has read it before thread 0 can overwrite it
	}
}

}  // namespace nbody
