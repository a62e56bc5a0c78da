// fixed_sums.cuh -- order-independent row sums for the pair-symmetric kernel (option
// "deterministic", force_sym_kernel<..., ACC = 2>).
//
// K1s adds contributions to a row from many work units (the i-side sum of every slice of the
// row's own block, the j-side sum of every warp of every partner block); with floating-point RED
// the ORDER of those additions depends on which CTA finishes first, so results are reproducible
// only to rounding.  Here every contribution is added as an INTEGER instead: a row sum is a
// fixed-point number of FX_LIMBS limbs, limb k holding the bits [32k, 32k+32) above 2^e0 in a
// 64-bit word (32 payload bits + 31 bits of headroom for deferred carries + sign).  A double
// x = +-M 2^E (M < 2^53) is split at the limb boundaries into three 32-bit chunks which are added
// to (subtracted from) three consecutive limbs with 64-bit integer RED.ADD.  Integer addition is
// associative and commutative, so the limbs -- and the double that fx_read() rounds them to, ONCE,
// correctly -- do not depend on the order: bit-reproducible runs whatever the scheduling, and the
// sum across units is exact instead of accumulating one rounding per contribution.
//
// The window [2^e0, 2^(e0 + 32 FX_LIMBS)) is chosen on the host (nbody_capi.cu: fx_window) from a
// rigorous bound on any single contribution, n m_max / eps^2; bits of a contribution below 2^e0
// are dropped (truncation towards zero: still order-independent), a non-finite contribution or one
// that would not fit sets the row's poison word and the row reads as NaN.
//
// The reference has no counterpart (it sums sequentially on one thread per row,
// lab_4_N_body_MPI/src/include/N_body/struct_ode_system.hpp:60-75); the result stays within the
// same tolerances of it as the floating-point sums.
#pragma once
#include <stdint.h>

namespace nbody {

constexpr int FX_LIMBS = 6;					// 192-bit window
constexpr int FX_PLANES = 3 * FX_LIMBS + 1;	// x, y, z limbs + one poison word per row
constexpr int FX_WINDOW_BITS = 32 * FX_LIMBS;

// The host keeps |e0| <= FX_E0_LIMIT (nbody_capi.cu: fx_window refuses anything else), so that in
// fx_add a zero / subnormal (biased exponent 0) lands below the window and inf / NaN (0x7ff) above
// the fast path's range without separate tests.
constexpr int FX_E0_LIMIT = 800;

// Anything the fast path of fx_add does not take: zero, bits below the window (dropped towards
// zero), inf / NaN and values at the top of the window (poison).
__device__ __forceinline__ void fx_add_slow(unsigned long long* __restrict__ limb0,
											const size_t stride,
											unsigned long long* __restrict__ poison, const double x,
											const int e0) {
	const unsigned long long b = (unsigned long long)__double_as_longlong(x);
	const int ex = (int)((b >> 52) & 0x7ffu);
	if (ex == 0) return;  // zero (and subnormals: far below any window in use)
	if (ex == 0x7ff) {	  // inf / NaN: the row is not a number any more
		atomicAdd(poison, 1ull);
		return;
	}
	unsigned long long m = (b & 0xfffffffffffffull) | 0x10000000000000ull;
	int p = ex - 1075 - e0;	 // position of m's lowest bit in the window
	if (p < 0) {
		if (p <= -53) return;
		m >>= -p;  // drop the bits below the window (towards zero)
		p = 0;
	}
	const int k = p >> 5, r = p & 31;
	const unsigned long long lo = m << r;
	unsigned long long c0 = lo & 0xffffffffull, c1 = lo >> 32, c2 = r ? m >> (64 - r) : 0ull;
	if (k + 2 >= FX_LIMBS &&
		((k >= FX_LIMBS && c0) || (k + 1 >= FX_LIMBS && c1) || c2)) {
		atomicAdd(poison, 1ull);  // above the window: cannot happen with the host's bound
		return;
	}
	if ((long long)b < 0) {	 // two's complement per limb
		c0 = 0ull - c0;
		c1 = 0ull - c1;
		c2 = 0ull - c2;
	}
	if (c0) atomicAdd(limb0 + (size_t)k * stride, c0);
	if (c1) atomicAdd(limb0 + (size_t)(k + 1) * stride, c1);
	if (c2) atomicAdd(limb0 + (size_t)(k + 2) * stride, c2);
}

// planes[(c * FX_LIMBS + k) * stride + row], poison = planes[3 * FX_LIMBS * stride + row]
// Fast path (a normal number whose 53 bits lie inside the window with two limbs to spare): the
// 53-bit mantissa shifted by r = p mod 32 IS three 32-bit words -- three funnel shifts -- each
// added to (subtracted from) its limb with one 64-bit integer RED; zero words are skipped.
__device__ __forceinline__ void fx_add(unsigned long long* __restrict__ limb0, const size_t stride,
									   unsigned long long* __restrict__ poison, const double x,
									   const int e0) {
	const unsigned hi = (unsigned)__double2hiint(x), lo = (unsigned)__double2loint(x);
	const int p = (int)((hi >> 20) & 0x7ffu) - 1075 - e0;  // position of the lowest mantissa bit
	if ((unsigned)p >= (unsigned)(32 * (FX_LIMBS - 2))) {
		fx_add_slow(limb0, stride, poison, x, e0);
		return;
	}
	const unsigned mh = (hi & 0xfffffu) | 0x100000u;
	const int r = p & 31;
	const unsigned c0 = lo << r;
	const unsigned c1 = __funnelshift_l(lo, mh, r);
	const unsigned c2 = __funnelshift_l(mh, 0u, r);
	const bool neg = (int)hi < 0;
	unsigned long long* q = limb0 + (size_t)(p >> 5) * stride;
	if (c0) atomicAdd(q, neg ? 0ull - (unsigned long long)c0 : (unsigned long long)c0);
	if (c1) atomicAdd(q + stride, neg ? 0ull - (unsigned long long)c1 : (unsigned long long)c1);
	if (c2) atomicAdd(q + 2 * stride, neg ? 0ull - (unsigned long long)c2 : (unsigned long long)c2);
}

// The double nearest to the limbs' value (round to nearest even, one rounding).
__device__ __forceinline__ double fx_read(const unsigned long long* __restrict__ limb0,
										  const size_t stride, const int e0) {
	uint32_t d[FX_LIMBS + 2];
	long long carry = 0;
#pragma unroll
	for (int k = 0; k < FX_LIMBS; ++k) {
		const long long t = (long long)limb0[(size_t)k * stride] + carry;
		d[k] = (uint32_t)t;
		carry = t >> 32;  // arithmetic: floor division, the remainder d[k] is non-negative
	}
	d[FX_LIMBS] = (uint32_t)carry;
	d[FX_LIMBS + 1] = (uint32_t)(carry >> 32);
	const bool neg = carry < 0;
	if (neg) {	// magnitude of the (FX_LIMBS + 2)-digit two's complement number
		unsigned long long c = 1;
#pragma unroll
		for (int k = 0; k < FX_LIMBS + 2; ++k) {
			const unsigned long long v = (unsigned long long)(uint32_t)~d[k] + c;
			d[k] = (uint32_t)v;
			c = v >> 32;
		}
	}
	// leading digit h, the two digits below it, and "anything non-zero further down"
	int h = -1;
	uint32_t a = 0, b1 = 0, b2 = 0, rest = 0;
#pragma unroll
	for (int k = FX_LIMBS + 1; k >= 0; --k) {
		if (h < 0) {
			if (d[k]) { h = k; a = d[k]; }
		} else if (h - k == 1) b1 = d[k];
		else if (h - k == 2) b2 = d[k];
		else rest |= d[k];
	}
	if (h < 0) return 0.;
	const int lz = __clz((int)a);
	const unsigned long long hi = ((unsigned long long)a << 32) | b1;
	unsigned long long m64 = lz ? (hi << lz) | ((unsigned long long)b2 >> (32 - lz)) : hi;
	const uint32_t tail = lz ? (b2 << lz) : b2;
	if (tail | rest) m64 |= 1ull;  // sticky bit, 10 places below the rounding position
	const double v = __ull2double_rn(m64);	// m64 has its leading one at bit 63
	const double s = ldexp(v, 32 * (h - 1) - lz + e0);
	return neg ? -s : s;
}

}  // namespace nbody
