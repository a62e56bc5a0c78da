// cuda_emu.h -- TEST INFRASTRUCTURE: just enough of the CUDA execution model to run this
// repository's kernels (parallel_comp_labs_b200/csrc/*.cuh, compiled with NBODY_HOST_EMU) on the
// CPU, thread by thread, so that kernel SOURCE is exercised by the `-m "not gpu"` suite:
//   * one CTA at a time, every CUDA thread a ucontext fiber, round-robin scheduling;
//   * __syncthreads, __shfl_sync, __any_sync as real rendezvous between fibers (a lane that
//     skips a collective, or a barrier not reached by every thread, is reported as a deadlock);
//   * mbarrier + bulk-async copy: the copy happens at issue and completes the barrier's phase;
//     a wait on a phase that was never armed is a deadlock; a refill of a stage is checked
//     against "every thread has passed a __syncthreads since the stage was last waited on";
//   * the MUFU seeds are modelled with their measured accuracy (rsqrt to ~2^-20, low word zero);
//   * atomics are plain read-modify-write (fibers never run concurrently).
// It does not model memory ordering, bank conflicts, registers or timing.  Never linked into the
// product; the GPU parity tests remain the proof of the bits.
#pragma once
#define NBODY_HOST_EMU 1

#include <ucontext.h>

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <map>
#include <stdexcept>
#include <string>
#include <vector>

// ---------------------------------------------------------------- CUDA vocabulary
#define __global__
#define __device__
#define __host__
#define __forceinline__ inline
#define __launch_bounds__(...)
#define __align__(n) __attribute__((aligned(n)))
#define __shared__ static thread_local

struct double4 { double x, y, z, w; };
struct float4 { float x, y, z, w; };
struct double2 { double x, y; };
struct dim3 { unsigned x = 1, y = 1, z = 1; dim3(unsigned a = 1, unsigned b = 1, unsigned c = 1) : x(a), y(b), z(c) {} };
inline double4 make_double4(double x, double y, double z, double w) { return {x, y, z, w}; }

namespace emu {

struct Fiber {
	ucontext_t ctx;
	unsigned char* stack = nullptr;	 // from the per-thread pool below (never zero-filled)
	dim3 tid;
	bool done = false;
	unsigned long long sync_epoch = 0;	// __syncthreads passed so far
};

struct Cta {
	std::vector<Fiber> fibers;
	std::vector<unsigned char> smem;
	dim3 block_idx, block_dim, grid_dim;
	unsigned arrived = 0;
	unsigned long long generation = 0;
	// warp collectives: [warp][parity][lane]
	std::vector<uint64_t> slot;
	std::vector<unsigned> w_arrived;
	std::vector<unsigned long long> w_generation;
	unsigned long long progress = 0;  // bumped whenever any rendezvous completes or a fiber ends
	// per mbarrier: __syncthreads epoch at which each thread last waited on it (-1: never)
	std::map<const void*, std::vector<long long>> waited;
};

// all emulation state is per OS thread: rank threads of one process may emulate concurrently
constexpr size_t STACK_BYTES = 128 * 1024;
inline std::vector<unsigned char*>& stack_pool() {
	static thread_local std::vector<unsigned char*> pool;	// kept for the life of the thread
	return pool;
}
inline uint64_t& sched_seed() {
	static thread_local uint64_t seed = [] {
		const char* e = getenv("EMU_SCHED_SEED");
		return e ? (uint64_t)strtoull(e, nullptr, 0) * 0x9E3779B97F4A7C15ull + 1 : (uint64_t)0;
	}();
	return seed;
}
inline Cta*& cta() { static thread_local Cta* c = nullptr; return c; }
inline Fiber*& cur() { static thread_local Fiber* f = nullptr; return f; }
inline ucontext_t& sched_ctx() { static thread_local ucontext_t c; return c; }
inline std::string& failure() { static thread_local std::string s; return s; }

inline void yield() { swapcontext(&cur()->ctx, &sched_ctx()); }

[[noreturn]] inline void fail(const std::string& msg) {
	failure() = msg;
	cur()->done = true;
	cta()->progress++;
	swapcontext(&cur()->ctx, &sched_ctx());
	abort();
}

inline unsigned linear_tid() { return cur()->tid.x; }

inline void syncthreads() {
	Cta* c = cta();
	const unsigned long long gen = c->generation;
	cur()->sync_epoch++;
	if (++c->arrived == c->fibers.size()) {
		c->arrived = 0;
		c->generation++;
		c->progress++;
		return;
	}
	while (c->generation == gen) yield();
}

// one rendezvous of the 32 lanes of the calling warp; returns the parity of the slot set used
inline unsigned warp_rendezvous(uint64_t mine) {
	Cta* c = cta();
	const unsigned w = linear_tid() / 32, lane = linear_tid() % 32;
	const unsigned long long gen = c->w_generation[w];
	const unsigned par = (unsigned)(gen & 1u);
	c->slot[(w * 2 + par) * 32 + lane] = mine;
	if (++c->w_arrived[w] == 32) {
		c->w_arrived[w] = 0;
		c->w_generation[w]++;
		c->progress++;
	} else {
		while (c->w_generation[w] == gen) yield();
	}
	return par;
}
inline uint64_t warp_slot(unsigned par, unsigned lane) {
	return cta()->slot[((linear_tid() / 32) * 2 + par) * 32 + lane];
}

struct Entry { std::function<void()> fn; };
inline Entry*& entry() { static thread_local Entry* e = nullptr; return e; }
inline void trampoline() {
	entry()->fn();
	cur()->done = true;
	cta()->progress++;
	swapcontext(&cur()->ctx, &sched_ctx());
}

// Run `kernel()` for every thread of every CTA of the grid.  CTAs run one after another (a
// persistent kernel's first CTA then drains the whole ticket pool, later CTAs find it empty:
// both paths get exercised).  Throws std::runtime_error on deadlock or a reported failure.
inline void launch(dim3 grid, dim3 block, size_t smem_bytes, const std::function<void()>& kernel) {
	if (block.x % 32 || block.y != 1 || block.z != 1) throw std::runtime_error("emu: block must be (32k,1,1)");
	Entry e{kernel};
	entry() = &e;
	// EMU_CTA_SHUFFLE=<seed> (read at every launch): run the CTAs in a pseudo-random order instead
	// of 0,1,2,...  With persistent kernels that draw work from a ticket counter this changes which
	// CTA takes which unit and the ORDER in which units reach global memory -- the emulated
	// counterpart of the GPU's scheduling freedom (used to show that the fixed-point row sums of
	// option "deterministic" do not depend on it while the floating-point ones do).
	std::vector<unsigned> cta_order(grid.x);
	for (unsigned b = 0; b < grid.x; ++b) cta_order[b] = b;
	if (const char* e = getenv("EMU_CTA_SHUFFLE")) {
		uint64_t x = (uint64_t)strtoull(e, nullptr, 0) * 0x9E3779B97F4A7C15ull + 1;
		for (size_t t = cta_order.size(); t > 1; --t) {
			x ^= x << 13; x ^= x >> 7; x ^= x << 17;
			std::swap(cta_order[t - 1], cta_order[(size_t)(x % t)]);
		}
	}
	for (unsigned b : cta_order) {
		Cta c;
		c.block_idx = dim3(b);
		c.block_dim = block;
		c.grid_dim = grid;
		c.smem.assign(smem_bytes + 128, 0xCD);
		c.fibers.resize(block.x);
		c.slot.assign((size_t)(block.x / 32) * 2 * 32, 0);
		c.w_arrived.assign(block.x / 32, 0);
		c.w_generation.assign(block.x / 32, 0);
		cta() = &c;
		for (unsigned t = 0; t < block.x; ++t) {
			Fiber& f = c.fibers[t];
			f.tid = dim3(t);
			if (stack_pool().size() <= t) stack_pool().push_back(static_cast<unsigned char*>(malloc(STACK_BYTES)));
			f.stack = stack_pool()[t];
			getcontext(&f.ctx);
			f.ctx.uc_stack.ss_sp = f.stack;
			f.ctx.uc_stack.ss_size = STACK_BYTES;
			f.ctx.uc_link = &sched_ctx();
			makecontext(&f.ctx, (void (*)())trampoline, 0);
		}
		failure().clear();
		for (;;) {
			const unsigned long long before = c.progress;
			unsigned alive = 0;
			// EMU_SCHED_SEED=<n>: visit the threads in a fresh pseudo-random order on every sweep
			// instead of 0,1,2,...: which thread reaches a rendezvous first, and how far thread 0
			// (the producer of the bulk copies) runs ahead of its consumers, then varies from run
			// to run -- a missing barrier shows up as a wrong result or a failed ring check
			std::vector<unsigned> order(c.fibers.size());
			for (unsigned t = 0; t < order.size(); ++t) order[t] = t;
			if (sched_seed()) {
				uint64_t& x = sched_seed();
				for (size_t t = order.size(); t > 1; --t) {
					x ^= x << 13; x ^= x >> 7; x ^= x << 17;
					std::swap(order[t - 1], order[(size_t)(x % t)]);
				}
			}
			for (unsigned t : order) {
				Fiber& f = c.fibers[t];
				if (f.done) continue;
				++alive;
				cur() = &f;
				swapcontext(&sched_ctx(), &f.ctx);
				if (!failure().empty()) throw std::runtime_error("emu: " + failure());
			}
			if (!alive) break;
			if (c.progress == before) {
				unsigned waiting = 0;
				for (Fiber& f : c.fibers) waiting += !f.done;
				throw std::runtime_error("emu: deadlock in CTA " + std::to_string(b) + " (" +
										 std::to_string(waiting) + " threads blocked)");
			}
		}
		cta() = nullptr;
	}
	entry() = nullptr;
}

}  // namespace emu

#define threadIdx (emu::cur()->tid)
#define blockIdx (emu::cta()->block_idx)
#define blockDim (emu::cta()->block_dim)
#define gridDim (emu::cta()->grid_dim)
#define NB_DYN_SMEM(name)                                                                       \
	unsigned char* name = reinterpret_cast<unsigned char*>(                                     \
		(reinterpret_cast<uintptr_t>(emu::cta()->smem.data()) + 127) & ~(uintptr_t)127)

inline void __syncthreads() { emu::syncthreads(); }

template <typename T>
inline T __shfl_sync(unsigned mask, T v, int src_lane) {
	if (mask != 0xffffffffu) emu::fail("__shfl_sync with a partial mask is not modelled");
	static_assert(sizeof(T) <= 8, "shuffle of at most 64 bits");
	uint64_t bits = 0;
	memcpy(&bits, &v, sizeof(T));
	const unsigned par = emu::warp_rendezvous(bits);
	const uint64_t got = emu::warp_slot(par, (unsigned)src_lane & 31u);
	T out;
	memcpy(&out, &got, sizeof(T));
	return out;
}
template <typename T>
inline T __shfl_down_sync(unsigned mask, T v, unsigned delta) {
	const unsigned lane = emu::linear_tid() % 32;
	const unsigned src = lane + delta < 32 ? lane + delta : lane;
	return __shfl_sync(mask, v, (int)src);
}
inline int __any_sync(unsigned mask, int pred) {
	if (mask != 0xffffffffu) emu::fail("__any_sync with a partial mask is not modelled");
	const unsigned par = emu::warp_rendezvous(pred ? 1 : 0);
	int any = 0;
	for (unsigned l = 0; l < 32; ++l) any |= (int)emu::warp_slot(par, l);
	return any;
}

// ---------------------------------------------------------------- intrinsics
inline double __dmul_rn(double a, double b) { return a * b; }
inline double __dadd_rn(double a, double b) { return a + b; }
inline double __fma_rn(double a, double b, double c) { return std::fma(a, b, c); }
inline float __fmul_rn(float a, float b) { return a * b; }
inline float __fadd_rn(float a, float b) { return a + b; }
inline long long __double_as_longlong(double x) { long long v; memcpy(&v, &x, 8); return v; }
inline double __longlong_as_double(long long v) { double x; memcpy(&x, &v, 8); return x; }
inline int __double2hiint(double x) { return (int)(__double_as_longlong(x) >> 32); }
inline int __float_as_int(float x) { int v; memcpy(&v, &x, 4); return v; }
inline int __double2loint(double x) { return (int)(unsigned)(__double_as_longlong(x) & 0xffffffffLL); }
// upper 32 bits of (hi:lo) << (shift & 31)
inline unsigned __funnelshift_l(unsigned lo, unsigned hi, unsigned shift) {
	const unsigned s = shift & 31u;
	return s ? (hi << s) | (lo >> (32u - s)) : hi;
}
inline int __clz(int x) { return x == 0 ? 32 : __builtin_clz((unsigned)x); }
inline double __ull2double_rn(unsigned long long v) { return (double)v; }  // round to nearest even
using std::ldexp;
inline double rsqrt(double x) { return 1.0 / std::sqrt(x); }
using std::fma;
using std::max;
using std::min;
inline void __nanosleep(unsigned) { emu::yield(); }
inline void __threadfence_system() {}

template <typename T>
inline T atomicMax(T* p, T v) {
	const T old = *p;
	if (v > old) *p = v;
	return old;
}
template <typename T>
inline T atomicAdd(T* p, T v) {
	const T old = *p;
	*p = old + v;
	return old;
}

namespace nbody {
// MUFU.RSQ64H: works on the high word only, result has a zero low word, relative error measured
// below 2^-19.04 on B200 (profiles/r01_issue_probe.log).  Modelled as: truncate the operand to
// its high word, correctly rounded 1/sqrt, keep 20 mantissa bits of the result.
inline double rsqrt_seed(double x) {
	long long b = __double_as_longlong(x) & ~0xffffffffLL;
	const double xh = __longlong_as_double(b);
	double y = 1.0 / std::sqrt(xh);
	long long yb = __double_as_longlong(y) & ~((1LL << 32) - 1);
	return __longlong_as_double(yb);
}
// MUFU.RSQ (FP32): ~2 ulp; flush-to-zero is irrelevant for the inputs used here
inline float rsqrt_seed(float x) {
	float y = (float)(1.0 / std::sqrt((double)x));
	int b = __float_as_int(y) ^ 1;	// 1 ulp off, deliberately
	float o;
	memcpy(&o, &b, 4);
	return std::isfinite(y) && y != 0.f ? o : y;
}

// mbarrier, modelled as a count of completed phases; bulk copies complete at issue.
inline void mbar_init(uint64_t* bar, uint32_t) {
	*bar = 0;
	emu::cta()->waited[bar].assign(emu::cta()->fibers.size(), -1);
}
inline void mbar_fence_init() {}
inline void mbar_expect_tx(uint64_t*, uint32_t) {}
inline void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
	if (bytes % 16 || (reinterpret_cast<uintptr_t>(dst) & 15) || (reinterpret_cast<uintptr_t>(src) & 15))
		emu::fail("bulk copy: size and addresses must be multiples of 16");
	const unsigned char* lo = emu::cta()->smem.data();
	if ((const unsigned char*)dst < lo || (const unsigned char*)dst + bytes > lo + emu::cta()->smem.size())
		emu::fail("bulk copy: destination outside the CTA's shared memory");
	// write-after-read check: whoever consumed this stage must have passed a __syncthreads since
	auto it = emu::cta()->waited.find(bar);
	if (it == emu::cta()->waited.end()) emu::fail("bulk copy: mbarrier was never initialised");
	for (size_t t = 0; t < it->second.size(); ++t) {
		const emu::Fiber& f = emu::cta()->fibers[t];
		if (!f.done && it->second[t] >= 0 && (long long)f.sync_epoch == it->second[t])
			emu::fail("bulk copy refills a stage that thread " + std::to_string(t) + " may still be reading");
	}
	memcpy(dst, src, bytes);
	(*bar)++;  // this phase is complete
	emu::cta()->progress++;
}
inline void mbar_wait(uint64_t* bar, uint32_t parity) {
	// try_wait.parity(P) succeeds once the phase of parity P has completed, i.e. when the
	// barrier's CURRENT phase has the other parity
	while (((*bar) & 1u) == parity) emu::yield();
	emu::cta()->waited[bar][emu::linear_tid()] = (long long)emu::cur()->sync_epoch;
}
}  // namespace nbody
