// nbody_capi.cu -- the C ABI declared in include/nbody_b200.h (libnbody_b200.so).
//
// One context == one GPU == one rank.  Streams: `sc` compute, `sm` comm (NCCL
// all-gather of the stage positions, overlapped with the local-tile force launch),
// `sx` snapshot copies.  NCCL is bound at run time with dlopen so that the single-GPU
// path has no NCCL dependency and a host process that already carries an NCCL (torch)
// shares it.  There is NO CPU path in this file: every entry point that computes
// launches sm_100a kernels or fails.
// Under NBODY_HOST_EMU (tests/tools/nbody_capi_emu.cpp, TEST INFRASTRUCTURE) this file is compiled
// by g++ against a stand-in CUDA runtime and every NB_LAUNCH runs the kernel SOURCE on the CPU, so
// that the host logic below is exercised by the CPU test suite.  The product build never defines it.
#ifndef NBODY_HOST_EMU
#include <cuda_runtime.h>
#define NB_LAUNCH(kern, grid, block, smem, stream, ...) kern<<<grid, block, smem, stream>>>(__VA_ARGS__)
#endif
#include <dlfcn.h>

#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/nbody_b200.h"
#include "force_kernels.cuh"
#include "force_sym_kernels.cuh"
#include "stage_kernels.cuh"

using namespace nbody;

// ------------------------------------------------------------------ NCCL (dlopen)
namespace {
typedef struct ncclComm* ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId;
typedef int ncclResult_t;
enum { NCCL_INT8 = 0, NCCL_UINT64 = 5, NCCL_FLOAT = 7, NCCL_DOUBLE = 8, NCCL_SUM = 0, NCCL_MAX = 2,
	   NCCL_MIN = 3 };
struct NcclApi {
	void* lib = nullptr;
	ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
	ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
	ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
	ncclResult_t (*AllGather)(const void*, void*, size_t, int, ncclComm_t, cudaStream_t) = nullptr;
	ncclResult_t (*AllReduce)(const void*, void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
	ncclResult_t (*ReduceScatter)(const void*, void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
	ncclResult_t (*GroupStart)() = nullptr;
	ncclResult_t (*GroupEnd)() = nullptr;
	const char* (*GetErrorString)(ncclResult_t) = nullptr;
	std::string err;
	std::mutex mu;
	bool load() {
		std::lock_guard<std::mutex> lk(mu);	 // rank threads of one process may race here
		if (lib) return true;
		const char* env = getenv("NBODY_NCCL_LIB");
		const char* names[] = {env, "libnccl.so.2", "libnccl.so"};
		for (const char* nm : names) {
			if (!nm) continue;
			lib = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
			if (lib) break;
		}
		if (!lib) {
			err = std::string("cannot dlopen NCCL: ") + dlerror();
			return false;
		}
#define NB_SYM(f) *(void**)(&f) = dlsym(lib, "nccl" #f); if (!f) { err = "missing nccl" #f; lib = nullptr; return false; }
		NB_SYM(GetUniqueId) NB_SYM(CommInitRank) NB_SYM(CommDestroy) NB_SYM(AllGather)
		NB_SYM(AllReduce) NB_SYM(ReduceScatter) NB_SYM(GroupStart) NB_SYM(GroupEnd) NB_SYM(GetErrorString)
#undef NB_SYM
		return true;
	}
};
NcclApi g_nccl;
thread_local std::string g_err;

struct Variant {
	int threads, ipt;  // i-bodies per thread
	size_t smem;
	const void* fn[2];	// [0] general masses, [1] uniform mass factored out
	void (*launch[2])(dim3, cudaStream_t, size_t, const ForceParams&, double eps2);
	const char* name;
};

template <typename T> Clamp<T> make_clamp(double eps2);
template <> Clamp<double> make_clamp<double>(double eps2) {
	Clamp<double> c;
	memcpy(&c.eps2_bits, &eps2, 8);
	c.thresh = (int)(c.eps2_bits >> 32);
	return c;
}
template <> Clamp<float> make_clamp<float>(double eps2) {
	Clamp<float> c;
	c.eps2 = (float)eps2;
	memcpy(&c.thresh, &c.eps2, 4);
	return c;
}

template <typename T, int TH, int I, int ST, int MB, int UM, int UN, int ORD>
void launch_force(dim3 grid, cudaStream_t s, size_t smem, const ForceParams& p, double eps2) {
	auto kern = force_kernel<T, TH, I, ST, MB, UM, UN, ORD>;
	NB_LAUNCH(kern, grid, dim3(TH), smem, s, p, make_clamp<T>(eps2));
}
#define NB_VARIANT(T, TH, I, ST, MB, UN, ORD)                                                  \
	Variant { TH, I, force_smem_bytes<T, ST>(),                                                \
			  {(const void*)force_kernel<T, TH, I, ST, MB, 0, UN, ORD>,                         \
			   (const void*)force_kernel<T, TH, I, ST, MB, 1, UN, ORD>},                        \
			  {launch_force<T, TH, I, ST, MB, 0, UN, ORD>, launch_force<T, TH, I, ST, MB, 1, UN, ORD>}, \
			  #T "/t" #TH "/i" #I "/s" #ST "/b" #MB "/u" #UN "/o" #ORD }

// {threads, i-bodies per thread, ring stages, min CTAs/SM, j unroll,
//  correction order of the inverse cube (2 = full double accuracy)}
const Variant g_var64[] = {
	NB_VARIANT(double, 128, 2, 3, 8, 8, 2),	 // default: best at every N measured (profiles/)
	NB_VARIANT(double, 128, 4, 3, 4, 8, 2),
	NB_VARIANT(double, 128, 2, 3, 6, 8, 2),
	NB_VARIANT(double, 256, 2, 3, 3, 8, 2),
	NB_VARIANT(double, 128, 4, 3, 3, 4, 2),
	NB_VARIANT(double, 128, 2, 3, 8, 8, 1),	 // first-order correction only
	NB_VARIANT(double, 128, 4, 3, 4, 8, 1),
	NB_VARIANT(double, 128, 3, 3, 6, 4, 2),
};
const Variant g_var32[] = {
	NB_VARIANT(float, 128, 2, 3, 8, 8, 2),
	NB_VARIANT(float, 128, 4, 3, 4, 4, 2),
	NB_VARIANT(float, 128, 8, 3, 4, 4, 2),
	NB_VARIANT(float, 256, 4, 3, 2, 4, 2),
};

// Pair-symmetric kernels (force_sym_kernels.cuh), opt-in with option "pair_symmetry".
struct SymVariant {
	int threads, ipt;
	size_t smem;
	const void* fn[2];	// [0] general masses, [1] uniform mass factored out
	void (*launch[2])(dim3, cudaStream_t, size_t, const SymParams&, double eps2);
	const char* name;
	bool fx;  // ACC = 2: fixed-point row sums (option "deterministic", fixed_sums.cuh)
};
template <typename T, int TH, int I, int ST, int MB, int UM, int ORD, int ROT, int ACC>
void launch_force_sym(dim3 grid, cudaStream_t s, size_t smem, const SymParams& p, double eps2) {
	auto kern = force_sym_kernel<T, TH, I, ST, MB, UM, ORD, ROT, ACC>;
	NB_LAUNCH(kern, grid, dim3(TH), smem, s, p, make_clamp<T>(eps2));
}
#define NB_SYM_VARIANT_A(T, TH, I, ST, MB, ORD, ROT, ACC)                                      \
	SymVariant { TH, I, force_sym_smem_bytes<T, TH, I, ST>(),                                  \
				 {(const void*)force_sym_kernel<T, TH, I, ST, MB, 0, ORD, ROT, ACC>,            \
				  (const void*)force_sym_kernel<T, TH, I, ST, MB, 1, ORD, ROT, ACC>},           \
				 {launch_force_sym<T, TH, I, ST, MB, 0, ORD, ROT, ACC>,                         \
				  launch_force_sym<T, TH, I, ST, MB, 1, ORD, ROT, ACC>},                        \
				 #T "/sym/t" #TH "/i" #I "/s" #ST "/b" #MB "/o" #ORD "/rot" #ROT "/acc" #ACC, ACC == 2 }
#define NB_SYM_VARIANT(T, TH, I, ST, MB, ORD, ROT) NB_SYM_VARIANT_A(T, TH, I, ST, MB, ORD, ROT, 0)
#define NB_SYM_VARIANT_FX(T, TH, I, ST, MB, ORD, ROT) NB_SYM_VARIANT_A(T, TH, I, ST, MB, ORD, ROT, 2)
// {threads, i-bodies per thread, ring stages, min CTAs/SM, correction order, rot}
// rot = 0: the j-body is shuffled round the warp with its accumulator; rot = 1: it is re-read
// from the staged tile and only the accumulator is shuffled (force_sym_kernels.cuh, sym_strip).
// New variants go at the END of a table: indices are part of the option interface.
const SymVariant g_sym64[] = {
	NB_SYM_VARIANT(double, 128, 4, 2, 5, 2, 0),	 // 96 registers, 20 warps/SM
	NB_SYM_VARIANT(double, 128, 4, 2, 4, 2, 0),	 // 128 registers, 16 warps/SM
	NB_SYM_VARIANT(double, 128, 6, 2, 4, 2, 0),
	NB_SYM_VARIANT(double, 128, 2, 2, 8, 2, 0),
	NB_SYM_VARIANT(double, 128, 8, 2, 3, 2, 0),
	NB_SYM_VARIANT(double, 128, 4, 2, 5, 2, 1),
	NB_SYM_VARIANT(double, 128, 4, 2, 4, 2, 1),
	NB_SYM_VARIANT(double, 128, 6, 2, 4, 2, 1),
	NB_SYM_VARIANT(double, 128, 8, 2, 3, 2, 1),
	// 64-thread CTAs: the same 3 warps per scheduler as entry 8 in blocks half the size, which
	// halves the partition and tail granularity (profiles/r02_sym_sweep2.md)
	NB_SYM_VARIANT(double, 64, 8, 2, 6, 2, 1),	 // 9: 512-row blocks
	NB_SYM_VARIANT(double, 64, 4, 2, 10, 2, 1),	 // 10: 256-row blocks
	// 11-15: the builds the automatic choice uses (8, 9, 7, 5, 3) with fixed-point row sums
	// (option "deterministic": the same pair arithmetic, order-independent sums across units)
	NB_SYM_VARIANT_FX(double, 128, 8, 2, 3, 2, 1),
	NB_SYM_VARIANT_FX(double, 64, 8, 2, 6, 2, 1),
	NB_SYM_VARIANT_FX(double, 128, 6, 2, 4, 2, 1),
	NB_SYM_VARIANT_FX(double, 128, 4, 2, 5, 2, 1),
	NB_SYM_VARIANT_FX(double, 128, 2, 2, 8, 2, 0),
};
// deterministic twin of each entry: the fixed-point build with the same block size (an fx entry
// is its own twin)
const int g_sym64_fx_twin[] = {14, 14, 13, 15, 11, 14, 14, 13, 11, 12, 15, 11, 12, 13, 14, 15};
const SymVariant g_sym32[] = {
	NB_SYM_VARIANT(float, 128, 4, 2, 8, 2, 0),
	NB_SYM_VARIANT(float, 128, 8, 2, 5, 2, 0),
	NB_SYM_VARIANT(float, 128, 2, 2, 8, 2, 0),
	NB_SYM_VARIANT(float, 128, 4, 2, 8, 2, 1),
	NB_SYM_VARIANT(float, 128, 8, 2, 5, 2, 1),
	NB_SYM_VARIANT(float, 64, 8, 2, 10, 2, 1),	// 5: 64-thread CTAs, 512-row blocks
	// 6-9: entries 1, 5, 3, 2 with fixed-point row sums (option "deterministic")
	NB_SYM_VARIANT_FX(float, 128, 8, 2, 5, 2, 0),
	NB_SYM_VARIANT_FX(float, 64, 8, 2, 10, 2, 1),
	NB_SYM_VARIANT_FX(float, 128, 4, 2, 8, 2, 1),
	NB_SYM_VARIANT_FX(float, 128, 2, 2, 8, 2, 0),
};
const int g_sym32_fx_twin[] = {8, 6, 9, 8, 6, 7, 6, 7, 8, 9};
constexpr int N_SYM64 = sizeof(g_sym64) / sizeof(SymVariant);
constexpr int N_SYM32 = sizeof(g_sym32) / sizeof(SymVariant);
static_assert(sizeof(g_sym64_fx_twin) / sizeof(int) == N_SYM64, "one twin entry per FP64 build");
static_assert(sizeof(g_sym32_fx_twin) / sizeof(int) == N_SYM32, "one twin entry per FP32 build");
constexpr unsigned SYM_EXTRA_TILES = 4;	 // rows allocated past n_pad: phantom tiles of the last block
constexpr int N_VAR64 = sizeof(g_var64) / sizeof(Variant);
constexpr int N_VAR32 = sizeof(g_var32) / sizeof(Variant);
constexpr unsigned PAD_UNIT = 256;	 // == TJ: per-rank blocks are whole j-tiles
constexpr unsigned MAX_SLICES = 64;		  // default kernel: one partial-sum row per slice
constexpr unsigned MAX_SYM_SLICES = 4096;  // pair-symmetric kernel: slices cost no memory
}  // namespace

struct nbody_ctx {
	size_t n = 0;
	int prec = 64, device = 0, rank = 0, world = 1;
	unsigned n_pad = 0, own_pad = 0, own0 = 0, own_real = 0;
	double G = 6.67e-11, eps = 1e-6;
	size_t esz = 8;	 // sizeof(T)
	void *Y = nullptr, *P = nullptr, *vel_n = nullptr, *vel_s = nullptr, *acc_r = nullptr,
		 *acc_v = nullptr, *apart = nullptr, *rhs_out = nullptr, *vel_all = nullptr,
		 *d_aos = nullptr;
	unsigned apart_slices = 0;
	size_t apart_bytes = 0;
	void* aj = nullptr;		 // T[3][rows] row sums of the pair-symmetric kernel (fixed-point builds:
	size_t aj_bytes = 0;	 // uint64[FX_PLANES][rows])
	long opt_pair_symmetry = -1, opt_sym_variant = -1;
	// option "deterministic": fixed-point row sums.  What the window needs to know about the input,
	// scanned at upload: largest |mass|, smallest non-zero |mass|, largest |coordinate|.
	long opt_deterministic = 0;
	bool stats_valid = false;
	double m_absmax = 0., m_absmin = 0., coord_absmax = 0.;
	int fx_e0 = 0;			 // window in use by the cached graph / the last plan
	bool fx_e0_valid = false;
	// general tableaux (nbody_step_rk / nbody_step_embedded): resident stage derivatives, the
	// norms of the last embedded step, and the state before it (nbody_step_undo)
	void *kr = nullptr, *kv = nullptr;
	int k_stages = 0;
	double* d_norms = nullptr;
	void *undo_Y = nullptr, *undo_vel = nullptr;
	bool undo_valid = false;
	int k0_slot = -1;  // slot of kr/kv that holds f(current state), or -1 (see enqueue_step_general)
	uint64_t sym_pairs = 0;	 // pair evaluations of one symmetric RHS (cached with its key)
	int sym_pairs_key = -1;
	double* d_energy = nullptr;
	unsigned* d_counters = nullptr;	 // work-unit ticket counters (one per force launch of a stage)
	void* snap_dev[NBODY_SNAPSHOT_SLOTS] = {};
	void* snap_host[NBODY_SNAPSHOT_SLOTS] = {};
	cudaEvent_t snap_ev[NBODY_SNAPSHOT_SLOTS] = {};
	// nbody_snapshot_wait / _release may run on ONE other thread (the trajectory writer) while the
	// integrator thread is inside any other call: the slot table is the only state they share.
	// A slot is published by the release-store of snap_state (after its ticket, event and buffers
	// are in place) and given back by the release-store of 0.
	std::atomic<int> snap_state[NBODY_SNAPSHOT_SLOTS] = {};	 // 0 free, 1 in flight / held
	std::atomic<uint64_t> snap_ticket[NBODY_SNAPSHOT_SLOTS] = {};
	uint64_t next_ticket = 1;
	cudaStream_t sc = nullptr, sm = nullptr, sx = nullptr;
	cudaEvent_t e_gy = nullptr;	 // completion of the eager end-of-step gather of Y
	cudaEvent_t e_k2 = nullptr, e_g = nullptr, e_snap = nullptr, ev_start = nullptr,
				ev_stop = nullptr;
	std::vector<cudaEvent_t> kev;  // pairs around force launches (time_kernels)
	size_t kev_used = 0;
	// option "trace": event pairs around every operation of a call (nbody_get_trace)
	struct TraceEv { uint32_t kind, stream; cudaEvent_t e0, e1; };
	std::vector<TraceEv> tev;
	size_t tev_used = 0;
	long opt_trace = 0;
	ncclComm_t comm = nullptr;
	// peer-pull exchange (nbody_p2p_connect): remote j-tiles are fetched from the peers' HBM by
	// the force kernel's bulk-async copies; no all-gather on the step path
	bool p2p = false;
	void* P2 = nullptr;				  // second stage buffer (stages alternate P, P2)
	unsigned* d_flags = nullptr;	  // [8] production epochs published by the peers
	unsigned** d_peer_flags = nullptr;  // device copy of peer_flags[]
	void* peerY[8] = {};
	void* peerP[2][8] = {};			  // [0] = P2 (even stages >= 2), [1] = P (odd stages)
	unsigned* peer_flags[8] = {};
	unsigned epoch = 0;				  // productions so far (same sequence on every rank)
	unsigned epoch_of[3] = {0, 0, 0};  // epoch at which Y, P2, P were last produced
	bool y_stale = false;			  // p2p: remote blocks of the local Y are out of date
	bool y_pending = false;	   // a gather of Y is in flight on sm; sc has not waited for e_gy yet
	bool uploaded = false;
	bool unimass = false;	   // all bodies share one mass: factored out of the pair sum
	double uniform_mass = 0.;
	long opt_unimass = -1;	   // -1 auto-detect at upload, 0 never
	long opt_time_kernels = 0, opt_slices = 0, opt_variant = -1;
	long opt_use_graph = -1;  // -1 auto: replay a captured step when the run is launch-bound
	int n_sm = 148;
	nbody_timing tm = {};
	bool timing_pending = false;
	std::string err;
	// cached CUDA graph of one step
	cudaGraphExec_t graph = nullptr;
	uint64_t graph_captures = 0;
	int graph_scheme = 0;
	double graph_tau = 0;
};

namespace {
int fail(nbody_ctx* c, int code, const std::string& msg) {
	if (c) c->err = msg;
	g_err = msg;
	return code;
}
#define CK(call)                                                                              \
	do {                                                                                      \
		cudaError_t e__ = (call);                                                             \
		if (e__ != cudaSuccess)                                                               \
			return fail(ctx, NBODY_ERR_CUDA,                                                  \
						std::string(#call) + ": " + cudaGetErrorString(e__));                 \
	} while (0)
#define CKN(call)                                                                             \
	do {                                                                                      \
		ncclResult_t r__ = (call);                                                            \
		if (r__ != 0)                                                                         \
			return fail(ctx, NBODY_ERR_NCCL, std::string(#call) + ": " +                      \
												 g_nccl.GetErrorString(r__));                 \
	} while (0)

// option "trace": bracket one enqueued operation with an event pair on its stream
int trace_begin(nbody_ctx* ctx, cudaStream_t st, uint32_t kind) {
	if (!ctx->opt_trace) return -1;
	if (ctx->tev_used == ctx->tev.size()) {
		nbody_ctx::TraceEv t{kind, 0, nullptr, nullptr};
		if (cudaEventCreate(&t.e0) != cudaSuccess || cudaEventCreate(&t.e1) != cudaSuccess) return -1;
		ctx->tev.push_back(t);
	}
	nbody_ctx::TraceEv& t = ctx->tev[ctx->tev_used];
	t.kind = kind;
	t.stream = st == ctx->sc ? 0u : st == ctx->sm ? 1u : 2u;
	cudaEventRecord(t.e0, st);
	return (int)ctx->tev_used++;
}
void trace_end(nbody_ctx* ctx, cudaStream_t st, int idx) {
	if (idx >= 0) cudaEventRecord(ctx->tev[idx].e1, st);
}

// variant -1 (default) = entry 0 of the table.
const Variant& variant_of(const nbody_ctx* c) {
	const int nv = c->prec == 64 ? N_VAR64 : N_VAR32;
	const Variant* tab = c->prec == 64 ? g_var64 : g_var32;
	long v = c->opt_variant;
	if (v < 0 || v >= nv) v = 0;
	return tab[v];
}

// Pair-symmetric variant in use.  An explicit "sym_variant" wins; otherwise the first entry of a
// preference list whose block (threads * i-bodies) divides the per-rank block of the partition --
// on several GPUs the ranks must own whole blocks; the 256-row builds always fit.  The lists are
// measured (B200, G interactions/s; profiles/r02_sym_sweep.md, r02_sym_sweep2.md at N = 1M,
// r02_sym_crossover.md and r02_sym_sweep3.md per size):
//   FP64 equal masses   N = 1M: 9 (64 threads, I=8) 1722 | 8 (128 threads, I=8) 1720 | 7 (I=6) 1720 | 5 (I=4) 1676
//                       N = 32k-256k: 9 ahead by 0.4-2 %;  N = 16k-20k: 5;  N <= 8k: 3 (256-row blocks)
//   FP64 general masses N >= 512k: 7 1529 | 9 1525 | 5 1518 | 8 1472;  N = 64k-256k: 9 >= 5 >= 7;  below: 5
//   FP32 equal masses   N >= 16k: 1 (I=8, shuffled j);  general masses: 5 (64 threads, I=8, re-read j);  N <= 12k: 3
// On several GPUs the size that matters is the per-rank share of the pairs: n / sqrt(world).
// Option "deterministic" = 1 then swaps the choice for its fixed-point twin (same block size).
// (pure host arithmetic on the numbers that decide: nbody_sym_variant_for exposes it to the tests)
// Several GPUs: every list is followed by a 256-row build (FP64 entry 10, FP32 entry 2; their
// fixed-point twins 15 and 9) -- nbody_partition pads the per-rank block to 256 rows only, so when
// that block is an odd multiple of 256 (n = 1 100 000 on 2 GPUs) no 512- or 1024-row build fits and
// the 256-row one is what runs, instead of the call failing.
const SymVariant& sym_variant_pick(int prec, size_t n, int world, unsigned own_pad, bool unimass,
								   bool det, long forced) {
	const int nv = prec == 64 ? N_SYM64 : N_SYM32;
	const SymVariant* tab = prec == 64 ? g_sym64 : g_sym32;
	const int* twin = prec == 64 ? g_sym64_fx_twin : g_sym32_fx_twin;
	if (forced >= 0 && forced < nv) return tab[det ? twin[forced] : forced];
	const double n_eff = (double)n / std::sqrt((double)world);
	// per-size winners of profiles/r02_sym_crossover.md and r02_sym_sweep3.md
	static const int p64_small[] = {3, 3, 3}, p64_mid[] = {5, 3, 3}, p64_fine[] = {9, 5, 3},
					 p64_um[] = {8, 9, 5}, p64_gen[] = {7, 9, 5}, p32_small[] = {3, 2, 2},
					 p32_um[] = {1, 3, 2}, p32_gen[] = {5, 3, 2};
	const int* pref;
	if (prec == 32) pref = n_eff < 14000 ? p32_small : unimass ? p32_um : p32_gen;
	else if (n_eff < 10000) pref = p64_small;
	else if (unimass) pref = n_eff < 26000 ? p64_mid : n_eff < 400000 ? p64_fine : p64_um;
	else pref = n_eff < 50000 ? p64_mid : (n_eff < 500000 || world > 1) ? p64_fine : p64_gen;
	for (int k = 0; k < 3; ++k) {
		const SymVariant& sv = tab[det ? twin[pref[k]] : pref[k]];
		const unsigned rows = (unsigned)(sv.threads * sv.ipt);
		// FP64, large n, only a 256-row build fits: entry 10 (64 threads, I = 4, j re-read) rather
		// than the small-n entry 3 (I = 2) the lists end on -- measured on 2 GPUs at n = 1 100 000
		// (per-rank block 2149 x 256): entry 10 3336 G int/s with equal masses, entry 3 2417 with
		// general masses (profiles/r03_multi_2gpu.md)
		if (world > 1 && rows == 256 && prec == 64 && n_eff >= 26000) break;
		if (world == 1 || own_pad % rows == 0) return sv;
	}
	const int fits_all = prec == 64 ? 10 : 2;  // 256-row builds: divide every per-rank block
	return tab[det ? twin[fits_all] : fits_all];
}
const SymVariant& sym_variant_of(const nbody_ctx* c) {
	return sym_variant_pick(c->prec, c->n, c->world, c->own_pad, c->unimass,
							c->opt_deterministic != 0, c->opt_sym_variant);
}

// Window of the fixed-point row sums (fixed_sums.cuh): the top comes from a rigorous bound on any
// single contribution -- a unit's share of a row is at most n pair terms, each at most
// m_max |d| / max(|d|, eps)^3 <= m_max / eps^2 (m = 1 with the mass factored out) -- and the lowest
// bit lies FX_WINDOW_BITS below it.  Refused (the caller reports it) when the window would not
// leave 60 bits under the SMALLEST pair term the input can produce, m_min / D^2 with D the
// diagonal of the bounding box at upload: the sums would silently lose precision.
int fx_window(const nbody_ctx* c, int* e0, std::string* why) {
	if (!c->stats_valid) {
		*why = "option \"deterministic\" must be set before the upload (the window of the "
			   "fixed-point sums is derived from the uploaded masses and positions)";
		return NBODY_ERR_STATE;
	}
	const double eps2 = c->eps * c->eps;
	const double mmax = c->unimass ? 1. : c->m_absmax, mmin = c->unimass ? 1. : c->m_absmin;
	if (!(eps2 > 0.) || !std::isfinite(1. / eps2) || !std::isfinite(mmax)) {
		*why = "option \"deterministic\" needs eps > 0 (eps^2 representable) and finite masses: the "
			   "fixed-point window is bounded by n m_max / eps^2";
		return NBODY_ERR_STATE;
	}
	int ex = 0;
	// log2 of the bound, rounded up, + 2 bits for the rounding of the pair arithmetic
	std::frexp(mmax > 0. ? mmax : 1., &ex);
	int top = ex;
	std::frexp(1. / eps2, &ex);
	top += ex;
	std::frexp((double)c->n, &ex);
	top += ex + 2;
	*e0 = top - FX_WINDOW_BITS;
	if (*e0 < -FX_E0_LIMIT || *e0 > FX_E0_LIMIT) {	// fixed_sums.cuh: fx_add relies on it
		*why = "option \"deterministic\": n m_max / eps^2 = 2^" + std::to_string(top) +
			   " is outside the range the fixed-point sums are built for";
		return NBODY_ERR_STATE;
	}
	if (c->coord_absmax > 0. && mmin > 0.) {
		int em = 0, ed = 0;
		std::frexp(mmin, &em);					 // mmin >= 2^(em-1)
		std::frexp(4. * c->coord_absmax, &ed);	 // D <= 2 sqrt(3) max|coordinate| < 2^ed
		const int smallest = (em - 1) - 2 * ed;	 // log2 of a lower bound of m_min / D^2
		if (*e0 + 60 > smallest) {
			*why = "option \"deterministic\": the dynamic range of this input (n m_max / eps^2 against "
				   "m_min / D^2) does not fit the " + std::to_string(FX_WINDOW_BITS) +
				   "-bit fixed-point window; use pair_symmetry = 0 for bit-reproducible sums";
			return NBODY_ERR_STATE;
		}
	}
	return NBODY_OK;
}

// Which force kernel runs.  Option "pair_symmetry": 1 = the pair-symmetric kernel, 0 = the
// one-sided kernel (every ordered pair, bit-reproducible), -1 (default) = automatic: the
// pair-symmetric kernel from SYM_AUTO_MIN_N bodies on (with slices cut at strip granularity it
// wins from ~2k bodies: profiles/r02_sym_crossover.md; below that a step is launch latency and
// the one-sided kernel's ragged-tile loop is the shorter path) unless the peer-pull exchange is
// connected.
constexpr size_t SYM_AUTO_MIN_N = 2048;
// CUDA-graph replay of a step (one GPU): while a step is shorter than ~10 ms the gaps between its
// 3 operations per stage are worth removing (4 % of a step at N = 20k, r02_bench_uniform20k.json)
constexpr size_t GRAPH_AUTO_MAX_N = 131072;
bool sym_in_use(const nbody_ctx* c) {
	if (c->opt_pair_symmetry == 0) return false;
	if (c->opt_pair_symmetry > 0) return true;
	if (c->world > 1 && c->p2p) return false;
	return c->n >= SYM_AUTO_MIN_N;
}

// j-slices per launch.  Units (i-tile, j-slice) are handed out dynamically, so the tail is
// about half a unit long: aim at >= UNITS_PER_SLOT units per resident CTA, keep a slice at
// least MIN_TILES tiles long (pipeline prologue + partial-sum traffic), cap the partial buffer.
// Measured at N = 1M (tools/slices_probe.py): 6 / 12 / 20 / 32 / 64 slices = 1093 / 1103 / 1109 /
// 1111 / 1113 G int/s, i.e. the half-unit tail model; 100 units per CTA (29 slices there) keeps
// the tail at 0.5 % for 0.9 GB of partial sums per launch (0.1 % of the launch's HBM time).
unsigned choose_slices(unsigned n_itiles, unsigned jt_count, unsigned slots, long forced,
					   unsigned max_slices = MAX_SLICES, unsigned units_per_slot = 100) {
	const unsigned MIN_TILES = 2;
	unsigned smax = std::min<unsigned>(max_slices, std::max(1u, jt_count / MIN_TILES));
	if (forced > 0) return std::min<unsigned>((unsigned)forced, std::max(1u, jt_count));
	if (n_itiles == 0) return 1;
	unsigned want = (units_per_slot * slots + n_itiles - 1) / n_itiles;
	return std::max(1u, std::min(want, smax));
}

// Units per resident CTA for the pair-symmetric kernel.  A unit costs a fixed o ~ 2 us (i-bodies
// loaded, shared sums zeroed, 3 * BLOCK REDs at its end) and the launch ends with a tail of about
// half a unit, so with T = the launch's time per CTA slot the loss o * U / T + 1 / (2 U) is
// smallest at U = sqrt(T / (2 o)).  (Measured: 256k bodies on 8 GPUs, T = 5 ms: U = 100 units of
// 52 us lost 4.3 % against the one-GPU rate; the formula gives U = 35.)
unsigned sym_units_per_slot(double pair_evaluations, unsigned prec) {
	const double rate = prec == 64 ? 0.85e12 : 1.8e12;	// pair evaluations / s of one B200
	const double t_us = pair_evaluations / rate * 1e6;
	const double u = std::sqrt(t_us / 4.0);
	return (unsigned)std::max(4.0, std::min(400.0, u));
}
}  // namespace

// ------------------------------------------------------------------ helpers on a context
namespace {

struct LaunchPlan {
	unsigned n_itiles, slots, grid;
	unsigned s_local, s_remote;	 // slices
	unsigned jt_total, jt_per_rank;
	unsigned apart_stride;	// rows between slices of apart
	// pair-symmetric kernel (one GPU): blocks of tpb j-tiles, s_local slices of a block's walk
	bool sym;
	unsigned nb, tpb, rows_pad;	 // blocks of the whole problem, tiles per block, nb * block rows
	unsigned block0, nb_own;	 // blocks this rank owns
	bool fx;					 // fixed-point row sums (SymVariant::fx), window exponent fx_e0
	int fx_e0;
};

// bytes of the row-sum array of the pair-symmetric kernel
size_t sym_sum_bytes(const nbody_ctx* ctx, const LaunchPlan& lp) {
	return lp.fx ? (size_t)FX_PLANES * lp.rows_pad * sizeof(unsigned long long)
				 : (size_t)3 * lp.rows_pad * ctx->esz;
}

int ensure_apart(nbody_ctx* ctx, const LaunchPlan& lp) {
	// the pair-symmetric kernel adds both sides of every pair into `aj`: no partial-sum rows
	const size_t bytes = lp.sym ? 0 : (size_t)(lp.s_local + lp.s_remote) * lp.apart_stride * 4 * ctx->esz;
	if (bytes > ctx->apart_bytes) {
		CK(cudaStreamSynchronize(ctx->sc));
		if (ctx->apart) CK(cudaFree(ctx->apart));
		ctx->apart = nullptr;
		ctx->apart_bytes = 0;
		CK(cudaMalloc(&ctx->apart, bytes));
		ctx->apart_bytes = bytes;
	}
	ctx->apart_slices = lp.sym ? 0 : lp.s_local + lp.s_remote;
	const size_t aj_bytes = lp.sym ? sym_sum_bytes(ctx, lp) : 0;
	if (aj_bytes > ctx->aj_bytes) {
		CK(cudaStreamSynchronize(ctx->sc));
		if (ctx->aj) CK(cudaFree(ctx->aj));
		ctx->aj = nullptr;
		ctx->aj_bytes = 0;
		CK(cudaMalloc(&ctx->aj, aj_bytes));
		ctx->aj_bytes = aj_bytes;
	}
	return NBODY_OK;
}

// pair evaluations of one pair-symmetric RHS: block rows x the j-bodies their walk touches
// (whole strips of 32 in partner tiles, existing bodies in the block's own tiles)
uint64_t sym_pair_evaluations(unsigned n_real, unsigned nb, unsigned tpb, unsigned block0,
							  unsigned nb_own) {
	const unsigned TJ = 256, block = tpb * TJ;
	std::vector<uint64_t> own(nb), strips(nb + 1, 0);  // strips: prefix sums over blocks
	for (unsigned b = 0; b < nb; ++b) {
		uint64_t o = 0, st = 0;
		for (unsigned q = 0; q < tpb; ++q) {
			const uint64_t first = (uint64_t)(b * tpb + q) * TJ;
			const uint64_t cnt = first >= n_real ? 0 : std::min<uint64_t>(TJ, n_real - first);
			o += cnt;
			st += (cnt + 31) / 32 * 32;
		}
		own[b] = o;
		strips[b + 1] = strips[b] + st;
	}
	uint64_t total = 0;
	for (unsigned a = block0; a < block0 + nb_own; ++a) {
		const unsigned h = sym_half(nb, a);
		const unsigned lo = a + 1, hi = a + 1 + h;	// partner blocks [lo,hi) cyclic
		uint64_t part = hi <= nb ? strips[hi] - strips[lo]
								 : (strips[nb] - strips[std::min(lo, nb)]) + strips[hi - nb];
		total += (uint64_t)block * (own[a] + part);
	}
	return total;
}

int make_plan(nbody_ctx* ctx, LaunchPlan* lp) {
	lp->sym = sym_in_use(ctx);
	lp->fx = false;
	lp->fx_e0 = 0;
	if (lp->sym) {
		const SymVariant& sv = sym_variant_of(ctx);
		const unsigned TJ = 256, block = sv.threads * sv.ipt;
		if (sv.fx) {
			std::string why;
			const int rc = fx_window(ctx, &lp->fx_e0, &why);
			if (rc) return fail(ctx, rc, why);
			lp->fx = true;
			// a captured step holds the window by value
			if (ctx->graph && (!ctx->fx_e0_valid || ctx->fx_e0 != lp->fx_e0)) {
				cudaGraphExecDestroy(ctx->graph);
				ctx->graph = nullptr;
			}
			ctx->fx_e0 = lp->fx_e0;
			ctx->fx_e0_valid = true;
		}
		if (ctx->world > 1 && ctx->p2p)
			return fail(ctx, NBODY_ERR_STATE,
						"option \"pair_symmetry\" is not available with the peer-pull exchange");
		if (ctx->world > 1 && ctx->own_pad % block)
			return fail(ctx, NBODY_ERR_STATE,
						"option \"pair_symmetry\" on several GPUs needs the per-rank block (" +
							std::to_string(ctx->own_pad) + " rows) to be a multiple of " +
							std::to_string(block) + " rows");
		lp->tpb = block / TJ;
		lp->jt_total = ctx->n_pad / TJ;
		lp->jt_per_rank = ctx->own_pad / TJ;
		lp->nb = (lp->jt_total + lp->tpb - 1) / lp->tpb;
		lp->rows_pad = lp->nb * block;
		lp->nb_own = ctx->world > 1 ? ctx->own_pad / block : lp->nb;
		lp->block0 = (unsigned)ctx->rank * (ctx->world > 1 ? lp->nb_own : 0u);
		lp->apart_stride = lp->nb_own * block;
		lp->n_itiles = lp->nb_own;
		int per_sm = 0;
		CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, sv.fn[ctx->unimass ? 1 : 0],
														sv.threads, sv.smem));
		if (per_sm < 1) return fail(ctx, NBODY_ERR_CUDA, "force kernel does not fit on an SM");
		lp->slots = (unsigned)per_sm * (unsigned)ctx->n_sm;
		// shortest walk of a block: tpb * (1 + h) tiles with h >= (nb-1)/2 - (nb even ? 0 : 0)
		const unsigned h_min = (lp->nb & 1u) ? (lp->nb - 1u) / 2u : lp->nb / 2u - 1u;
		const unsigned min_walk = lp->tpb * (1u + h_min);
		// slices are cut at strip granularity (32 j-bodies; sym_schedule.h), at least 2 strips each
		const int key = (int)(sv.threads * sv.ipt);
		if (ctx->sym_pairs_key != key) {
			ctx->sym_pairs = sym_pair_evaluations((unsigned)ctx->n, lp->nb, lp->tpb, lp->block0, lp->nb_own);
			ctx->sym_pairs_key = key;
		}
		lp->s_local = choose_slices(lp->nb_own, min_walk * (TJ / 32), lp->slots, ctx->opt_slices,
									MAX_SYM_SLICES, sym_units_per_slot((double)ctx->sym_pairs, ctx->prec));
		lp->s_remote = 0;
		return NBODY_OK;
	}
	lp->nb = lp->tpb = lp->rows_pad = lp->block0 = lp->nb_own = 0;
	lp->apart_stride = ctx->own_pad;
	const Variant& v = variant_of(ctx);
	const unsigned itile = v.threads * v.ipt;
	const unsigned TJ = 256;
	lp->n_itiles = (ctx->own_real + itile - 1) / itile;
	int per_sm = 0;
	CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, v.fn[ctx->unimass ? 1 : 0], v.threads,
													v.smem));
	if (per_sm < 1) return fail(ctx, NBODY_ERR_CUDA, "force kernel does not fit on an SM");
	lp->slots = (unsigned)per_sm * (unsigned)ctx->n_sm;
	lp->jt_total = ctx->n_pad / TJ;
	lp->jt_per_rank = ctx->own_pad / TJ;
	// only tiles that hold real bodies matter for the tail (pad tiles are all-zero-mass),
	// but keeping them makes the per-rank tile ranges uniform; n_pad - n < world*PAD_UNIT.
	lp->s_local = choose_slices(lp->n_itiles, lp->jt_per_rank, lp->slots, ctx->opt_slices);
	lp->s_remote = ctx->world > 1 ? choose_slices(lp->n_itiles, lp->jt_total - lp->jt_per_rank,
												  lp->slots, ctx->opt_slices)
								  : 0;
	return NBODY_OK;
}

// Stage position buffers.  X_0 = Y; with NCCL every later stage lives in P; with the
// peer-pull exchange stages alternate P (odd) / P2 (even) so that a rank never overwrites a
// block its peers may still be reading.
enum StageBuf { BUF_Y = 0, BUF_P2 = 1, BUF_P = 2 };

void* buf_ptr(nbody_ctx* ctx, StageBuf b) { return b == BUF_Y ? ctx->Y : b == BUF_P ? ctx->P : ctx->P2; }
StageBuf stage_buf(const nbody_ctx* ctx, int s) {
	if (s == 0) return BUF_Y;
	return (ctx->p2p && (s % 2 == 0)) ? BUF_P2 : BUF_P;
}

// one force launch over j-tiles [jt_begin, jt_begin + jt_count) (wrapping); `remote` selects
// the peers' copies of the blocks this rank does not own (peer-pull exchange only)
int launch_force_range(nbody_ctx* ctx, const LaunchPlan& lp, StageBuf buf, bool remote,
					   unsigned jt_begin, unsigned jt_count, unsigned n_slices,
					   unsigned slice_base) {
	if (lp.n_itiles == 0 || jt_count == 0) return NBODY_OK;
	const Variant& v = variant_of(ctx);
	ForceParams p;
	const size_t block_bytes = (size_t)ctx->own_pad * 4 * ctx->esz;
	// NCCL exchange: every block lives in this GPU's own array -> one base pointer, any number of
	// ranks.  Peer-pull exchange (<= 8 ranks): blk[r] points into rank r's HBM.
	p.contiguous = !(remote && ctx->p2p);
	for (int r = 0; r < 8; ++r) {
		const char* base = (const char*)buf_ptr(ctx, buf);
		if (remote && ctx->p2p && r < ctx->world && r != ctx->rank)
			base = (const char*)(buf == BUF_Y ? ctx->peerY[r] : ctx->peerP[buf == BUF_P ? 1 : 0][r]);
		p.blk[r] = r < ctx->world ? base + (size_t)r * block_bytes : nullptr;
	}
	p.own = (const char*)buf_ptr(ctx, buf) + (size_t)ctx->rank * block_bytes;
	p.jt_per_rank = lp.jt_per_rank;
	p.rank = (unsigned)ctx->rank;
	p.apart = ctx->apart;
	p.own0 = ctx->own0;
	p.own_pad = ctx->own_pad;
	p.n_itiles = lp.n_itiles;
	p.jt_begin = jt_begin % lp.jt_total;
	p.jt_count = jt_count;
	p.jt_total = lp.jt_total;
	p.n_slices = n_slices;
	p.slice_base = slice_base;
	p.n_real = (unsigned)ctx->n;
	p.counter = ctx->d_counters + (slice_base ? 1 : 0);
	p.apart_rows = (unsigned)std::min<size_t>(ctx->apart_bytes / (4 * ctx->esz), 0xffffffffu);
	{
		const int tr = trace_begin(ctx, ctx->sc, NBODY_TRACE_MEMSET);
		CK(cudaMemsetAsync(p.counter, 0, sizeof(unsigned), ctx->sc));
		trace_end(ctx, ctx->sc, tr);
	}
	const unsigned long long units = (unsigned long long)lp.n_itiles * n_slices;
	const unsigned grid = (unsigned)std::min<unsigned long long>(units, lp.slots);
	cudaEvent_t e0 = nullptr, e1 = nullptr;
	if (ctx->opt_time_kernels) {
		if (ctx->kev_used + 2 > ctx->kev.size()) {
			for (int k = 0; k < 2; ++k) {
				cudaEvent_t e;
				CK(cudaEventCreate(&e));
				ctx->kev.push_back(e);
			}
		}
		e0 = ctx->kev[ctx->kev_used++];
		e1 = ctx->kev[ctx->kev_used++];
		CK(cudaEventRecord(e0, ctx->sc));
	}
	const int tr = trace_begin(ctx, ctx->sc, NBODY_TRACE_FORCE);
	v.launch[ctx->unimass ? 1 : 0](dim3(grid), ctx->sc, v.smem, p, ctx->eps * ctx->eps);
	trace_end(ctx, ctx->sc, tr);
	CK(cudaGetLastError());
	if (e1) CK(cudaEventRecord(e1, ctx->sc));
	ctx->tm.force_launches++;
	ctx->tm.total_launches++;
	return NBODY_OK;
}

// zero the ticket counter and the row-sum array of the pair-symmetric kernel
int sym_prepare(nbody_ctx* ctx, const LaunchPlan& lp) {
	const int tr = trace_begin(ctx, ctx->sc, NBODY_TRACE_MEMSET);
	CK(cudaMemsetAsync(ctx->d_counters, 0, sizeof(unsigned), ctx->sc));
	CK(cudaMemsetAsync(ctx->aj, 0, sym_sum_bytes(ctx, lp), ctx->sc));
	trace_end(ctx, ctx->sc, tr);
	return NBODY_OK;
}

// one pair-symmetric force launch over all bodies of `buf`
int launch_force_sym_all(nbody_ctx* ctx, const LaunchPlan& lp, StageBuf buf) {
	const SymVariant& v = sym_variant_of(ctx);
	SymParams p;
	p.pos = buf_ptr(ctx, buf);
	p.aj = ctx->aj;
	p.rows_pad = lp.rows_pad;
	p.nb = lp.nb;
	p.block0 = lp.block0;
	p.nb_own = lp.nb_own;
	p.jt_total = lp.jt_total;
	p.n_pad = ctx->n_pad;
	p.n_real = (unsigned)ctx->n;
	p.n_slices = lp.s_local;
	p.counter = ctx->d_counters;
	p.pad_pos = ctx->unimass ? (ctx->prec == 64 ? 1e150 : 1e18) : 0.;
	p.pos_rows = ctx->n_pad + SYM_EXTRA_TILES * PAD_UNIT;
	p.fx_e0 = lp.fx_e0;
	// (the ticket counter and the row sums were zeroed by sym_prepare, before the wait for the
	// exchange, so that the memsets overlap the all-gather)
	const unsigned long long units = (unsigned long long)lp.nb_own * lp.s_local;
	const unsigned grid = (unsigned)std::min<unsigned long long>(units, lp.slots);
	cudaEvent_t e0 = nullptr, e1 = nullptr;
	if (ctx->opt_time_kernels) {
		if (ctx->kev_used + 2 > ctx->kev.size()) {
			for (int k = 0; k < 2; ++k) {
				cudaEvent_t e;
				CK(cudaEventCreate(&e));
				ctx->kev.push_back(e);
			}
		}
		e0 = ctx->kev[ctx->kev_used++];
		e1 = ctx->kev[ctx->kev_used++];
		CK(cudaEventRecord(e0, ctx->sc));
	}
	const int tr = trace_begin(ctx, ctx->sc, NBODY_TRACE_FORCE);
	v.launch[ctx->unimass ? 1 : 0](dim3(grid), ctx->sc, v.smem, p, ctx->eps * ctx->eps);
	trace_end(ctx, ctx->sc, tr);
	CK(cudaGetLastError());
	if (e1) CK(cudaEventRecord(e1, ctx->sc));
	ctx->tm.force_launches++;
	ctx->tm.total_launches++;
	if (ctx->world > 1) {
		// every rank has collected reactions for ALL bodies its walk touched: sum them over the
		// ranks, each rank keeping the rows it owns (in place: rows_pad == world * own_pad here)
		// (fixed-point sums: integer limbs, so the order in which NCCL adds the ranks' shares does not
		// matter either)
		const int dt = lp.fx ? NCCL_UINT64 : ctx->prec == 64 ? NCCL_DOUBLE : NCCL_FLOAT;
		const size_t esz = lp.fx ? sizeof(unsigned long long) : ctx->esz;
		const int planes = lp.fx ? FX_PLANES : 3;
		const int tr2 = trace_begin(ctx, ctx->sc, NBODY_TRACE_REDUCE);
		CKN(g_nccl.GroupStart());
		for (int c = 0; c < planes; ++c) {
			char* base = (char*)ctx->aj + (size_t)c * lp.rows_pad * esz;
			CKN(g_nccl.ReduceScatter(base, base + (size_t)ctx->rank * ctx->own_pad * esz,
									 ctx->own_pad, dt, NCCL_SUM, ctx->comm, ctx->sc));
		}
		CKN(g_nccl.GroupEnd());
		trace_end(ctx, ctx->sc, tr2);
	}
	return NBODY_OK;
}

// all-gather the owned block of `buf` (V4[n_pad]) across ranks on the comm stream,
// after everything enqueued so far on sc; leaves `done` recorded on sm.
int enqueue_gather(nbody_ctx* ctx, void* buf, cudaEvent_t done) {
	CK(cudaEventRecord(ctx->e_k2, ctx->sc));
	CK(cudaStreamWaitEvent(ctx->sm, ctx->e_k2, 0));
	const size_t bytes = (size_t)ctx->own_pad * 4 * ctx->esz;
	char* base = (char*)buf;
	const int tr = trace_begin(ctx, ctx->sm, NBODY_TRACE_GATHER);
	CKN(g_nccl.AllGather(base + (size_t)ctx->rank * bytes, base, bytes, NCCL_INT8, ctx->comm,
						 ctx->sm));
	trace_end(ctx, ctx->sm, tr);
	CK(cudaEventRecord(done, ctx->sm));
	return NBODY_OK;
}

// peer-pull exchange: this rank has just produced its block of `buf` on sc
int p2p_publish(nbody_ctx* ctx, StageBuf buf) {
	ctx->epoch++;
	ctx->epoch_of[buf] = ctx->epoch;
	NB_LAUNCH(p2p_signal_kernel, dim3(1), dim3(32), 0, ctx->sc,
			  ctx->d_peer_flags, ctx->world, ctx->rank, ctx->epoch);
	CK(cudaGetLastError());
	ctx->tm.total_launches++;
	return NBODY_OK;
}
int p2p_wait(nbody_ctx* ctx, StageBuf buf) {
	NB_LAUNCH(p2p_wait_kernel, dim3(1), dim3(32), 0, ctx->sc,
			  ctx->d_flags, ctx->world, ctx->rank, ctx->epoch_of[buf]);
	CK(cudaGetLastError());
	ctx->tm.total_launches++;
	return NBODY_OK;
}

enum GatherMode { GATHER_NONE, GATHER_NOW, GATHER_PENDING_Y };

// acceleration of the owned rows from the positions in `buf`; the owned block is current.
// NCCL exchange -- GATHER_NOW: the other ranks' blocks are stale -> all-gather them now,
// overlapped with the local-tile launch; GATHER_PENDING_Y: that gather was already enqueued at
// the end of the previous step (buf == Y), the remote launch waits for its event.
// Peer-pull exchange -- the remote launch waits for the peers' epoch flags and then fetches
// their tiles straight from their HBM.
int enqueue_forces(nbody_ctx* ctx, const LaunchPlan& lp, StageBuf buf, GatherMode mode) {
	int rc;
	ctx->tm.interactions += (uint64_t)ctx->own_real * (uint64_t)ctx->n;	 // self pairs counted,
	// as the reference's loop bounds do (struct_ode_system.hpp:64); pad bodies are not.
	if (lp.sym) {
		ctx->tm.pair_evaluations += ctx->sym_pairs;
		// The walk of a block reaches half-way round the ring of blocks, so the whole stage array
		// must be present before the launch: the exchange is waited for (tens of microseconds
		// against a launch of milliseconds) instead of splitting the launch into local and remote
		// tiles; the memsets of the launch go between "gather enqueued" and "gather awaited", so
		// they overlap it.
		if (ctx->world > 1 && mode == GATHER_NOW &&
			(rc = enqueue_gather(ctx, buf_ptr(ctx, buf), ctx->e_g)))
			return rc;
		if ((rc = sym_prepare(ctx, lp))) return rc;
		if (ctx->world > 1) {
			if (mode == GATHER_NOW) {
				CK(cudaStreamWaitEvent(ctx->sc, ctx->e_g, 0));
			} else if (mode == GATHER_PENDING_Y) {
				CK(cudaStreamWaitEvent(ctx->sc, ctx->e_gy, 0));
				ctx->y_pending = false;
			}
		}
		return launch_force_sym_all(ctx, lp, buf);
	}
	ctx->tm.pair_evaluations += (uint64_t)ctx->own_real * (uint64_t)ctx->n;
	if (ctx->world == 1)
		return launch_force_range(ctx, lp, buf, false, 0, lp.jt_total, lp.s_local, 0);
	if (!ctx->p2p && mode == GATHER_NOW && (rc = enqueue_gather(ctx, buf_ptr(ctx, buf), ctx->e_g)))
		return rc;
	// local tiles first: overlaps the exchange
	if ((rc = launch_force_range(ctx, lp, buf, false, ctx->rank * lp.jt_per_rank, lp.jt_per_rank,
								 lp.s_local, 0)))
		return rc;
	if (ctx->p2p) {
		if ((rc = p2p_wait(ctx, buf))) return rc;
	} else {
		if (mode == GATHER_NOW) CK(cudaStreamWaitEvent(ctx->sc, ctx->e_g, 0));
		if (mode == GATHER_PENDING_Y) {
			CK(cudaStreamWaitEvent(ctx->sc, ctx->e_gy, 0));
			ctx->y_pending = false;
		}
	}
	return launch_force_range(ctx, lp, buf, true, (ctx->rank + 1) * lp.jt_per_rank,
							  lp.jt_total - lp.jt_per_rank, lp.s_remote, lp.s_local);
}

template <typename T>
int launch_stage(nbody_ctx* ctx, const LaunchPlan& lp, void* rhs_out, void* next_buf,
				 double c_next, double b, double final_scale, int first, int last) {
	if (ctx->own_real == 0) return NBODY_OK;
	StageParams<T> p;
	p.apart = ctx->apart;
	p.Y = ctx->Y;
	p.P = next_buf;
	p.vel_n = ctx->vel_n;
	p.vel_s = ctx->vel_s;
	p.acc_r = ctx->acc_r;
	p.acc_v = ctx->acc_v;
	p.rhs_out = rhs_out;
	p.own0 = ctx->own0;
	p.own_pad = ctx->own_pad;
	p.own_real = ctx->own_real;
	p.n_slices = lp.sym ? 0 : lp.s_local + lp.s_remote;
	p.apart_stride = lp.apart_stride;
	p.aj = lp.sym ? ctx->aj : nullptr;
	p.aj_stride = lp.rows_pad;
	p.aj_fx = lp.fx ? 1 : 0;
	p.aj_e0 = lp.fx_e0;
	p.mass_scale = ctx->unimass ? (T)ctx->uniform_mass : T(1);
	p.minusG = (T)(-(T)ctx->G);
	p.c_next = (T)c_next;
	p.b = (T)b;
	p.final_scale = (T)final_scale;
	p.first = first;
	p.last = last;
	auto kern = stage_kernel<T>;
	const int tr = trace_begin(ctx, ctx->sc, NBODY_TRACE_STAGE);
	NB_LAUNCH(kern, dim3((ctx->own_real + 255) / 256), dim3(256), 0, ctx->sc, p);
	trace_end(ctx, ctx->sc, tr);
	CK(cudaGetLastError());
	ctx->tm.total_launches++;
	return NBODY_OK;
}

// Coefficients of one step of a general explicit RK scheme, already multiplied by tau in T.
struct GenCoef {
	int ns = 0;
	bool want_err = false;
	bool fsal = false;	// "first same as last": the last stage is f(y_{n+1}) = the next step's k_0
	double a[NBODY_MAX_STAGES][NBODY_MAX_STAGES] = {};	// tau * a[s][j], j < s
	double w[NBODY_MAX_STAGES] = {};					// tau * b[j]
	double e[NBODY_MAX_STAGES] = {};					// tau * (b[j] - b_err[j])
};

template <typename T>
int launch_stage_general(nbody_ctx* ctx, const LaunchPlan& lp, void* next_buf, const GenCoef& k,
						 int s, int k_from) {
	if (ctx->own_real == 0) return NBODY_OK;
	StageGenParams<T> g;
	StageParams<T>& p = g.base;
	p.apart = ctx->apart;
	p.Y = ctx->Y;
	p.P = next_buf;
	p.vel_n = ctx->vel_n;
	p.vel_s = ctx->vel_s;
	p.acc_r = ctx->acc_r;
	p.acc_v = ctx->acc_v;
	p.rhs_out = nullptr;
	p.own0 = ctx->own0;
	p.own_pad = ctx->own_pad;
	p.own_real = ctx->own_real;
	p.n_slices = lp.sym ? 0 : lp.s_local + lp.s_remote;
	p.apart_stride = lp.apart_stride;
	p.aj = lp.sym ? ctx->aj : nullptr;
	p.aj_stride = lp.rows_pad;
	p.aj_fx = lp.fx ? 1 : 0;
	p.aj_e0 = lp.fx_e0;
	p.mass_scale = ctx->unimass ? (T)ctx->uniform_mass : T(1);
	p.minusG = (T)(-(T)ctx->G);
	p.c_next = p.b = p.final_scale = T(0);
	p.first = s == 0;
	p.last = s == k.ns - 1;
	g.kr = ctx->kr;
	g.kv = ctx->kv;
	g.s = s;
	g.n_stages = k.ns;
	g.last = s == k.ns - 1;
	g.want_err = g.last && k.want_err;
	g.k_from = k_from;
	g.store_last = k.fsal ? 1 : 0;
	g.norms = g.want_err ? ctx->d_norms : nullptr;
	for (int j = 0; j < NBODY_MAX_STAGES; ++j) {
		g.c[j] = j <= s ? (T)(g.last ? k.w[j] : k.a[s + 1][j]) : T(0);
		g.e[j] = (g.want_err && j <= s) ? (T)k.e[j] : T(0);
	}
	auto kern = stage_general_kernel<T>;
	const int tr = trace_begin(ctx, ctx->sc, NBODY_TRACE_STAGE);
	NB_LAUNCH(kern, dim3((ctx->own_real + 255) / 256), dim3(256), 0, ctx->sc, g);
	trace_end(ctx, ctx->sc, tr);
	CK(cudaGetLastError());
	ctx->tm.total_launches++;
	return NBODY_OK;
}

int stage_dispatch(nbody_ctx* ctx, const LaunchPlan& lp, void* rhs_out, void* next_buf,
				   double c_next, double b, double final_scale, int first, int last) {
	return ctx->prec == 64
			   ? launch_stage<double>(ctx, lp, rhs_out, next_buf, c_next, b, final_scale, first, last)
			   : launch_stage<float>(ctx, lp, rhs_out, next_buf, c_next, b, final_scale, first, last);
}

// Coefficients of one step of an explicit RK scheme whose stage s+1 depends on stage s only:
// stage state = y_n + c[s] k_s, y_{n+1} = y_n + fs * sum_s b[s] k_s (c and fs include tau).
struct StepCoef {
	int ns = 0;
	double c[NBODY_MAX_STAGES] = {}, b[NBODY_MAX_STAGES] = {}, fs = 0;
};

// tau/2, tau, tau/6 formed in T on the host exactly as the oracle forms them.
StepCoef builtin_coef(const nbody_ctx* ctx, int scheme, double tau) {
	StepCoef k;
	if (ctx->prec == 64) {
		if (scheme == NBODY_RK2) { k.ns = 2; k.c[0] = tau / 2; k.b[0] = 0.; k.b[1] = 1.; k.fs = tau; }
		else { k.ns = 4; k.c[0] = tau / 2; k.c[1] = tau / 2; k.c[2] = tau; k.b[0] = 1.; k.b[1] = 2.; k.b[2] = 2.; k.b[3] = 1.; k.fs = tau / 6; }
	} else {
		const float t = (float)tau;
		if (scheme == NBODY_RK2) { k.ns = 2; k.c[0] = t / 2; k.b[0] = 0.; k.b[1] = 1.; k.fs = t; }
		else { k.ns = 4; k.c[0] = t / 2; k.c[1] = t / 2; k.c[2] = t; k.b[0] = 1.; k.b[1] = 2.; k.b[2] = 2.; k.b[3] = 1.; k.fs = t / 6; }
	}
	return k;
}

// one integrator step, enqueued on sc (+ sm)
int enqueue_step(nbody_ctx* ctx, const LaunchPlan& lp, const StepCoef& k) {
	int rc;
	const int ns = k.ns;
	// peer-pull exchange: a 1-stage scheme overwrites this rank's block of Y right after its own
	// force launch, while peers may still be pulling it (no later epoch covers that read)
	if (ctx->p2p && ns < 2)
		return fail(ctx, NBODY_ERR_STATE, "1-stage schemes need the NCCL exchange (peer-pull has no "
										  "acknowledgement before the state is overwritten)");
	const double* c = k.c;
	const double* b = k.b;
	const double fs = k.fs;
	for (int s = 0; s < ns; ++s) {
		const StageBuf buf = stage_buf(ctx, s);
		const GatherMode mode = s > 0 ? GATHER_NOW : (ctx->y_pending ? GATHER_PENDING_Y : GATHER_NONE);
		if ((rc = enqueue_forces(ctx, lp, buf, mode))) return rc;
		const int last = s == ns - 1;
		const StageBuf nxt = last ? BUF_Y : stage_buf(ctx, s + 1);
		if ((rc = stage_dispatch(ctx, lp, nullptr, buf_ptr(ctx, nxt), last ? 0. : c[s], b[s], fs,
								 s == 0, last)))
			return rc;
		if (ctx->p2p && (rc = p2p_publish(ctx, nxt))) return rc;
	}
	if (ctx->p2p) {
		ctx->y_stale = true;  // the peers' blocks of the local Y are out of date; pulled on demand
	} else if (ctx->world > 1) {
		// Eager gather of y_{n+1}: every rank enqueues it here, so Y is whole on all ranks after
		// every step without any later collective (snapshots, downloads and the energy of ONE
		// rank only wait for e_gy); the next step's local-tile launch still overlaps it.
		if ((rc = enqueue_gather(ctx, ctx->Y, ctx->e_gy))) return rc;
		ctx->y_pending = true;
	}
	return NBODY_OK;
}

// one step of a general explicit RK scheme (all k_j resident), enqueued on sc (+ sm)
int enqueue_step_general(nbody_ctx* ctx, const LaunchPlan& lp, const GenCoef& k) {
	int rc;
	if (ctx->p2p && k.ns < 2)
		return fail(ctx, NBODY_ERR_STATE, "1-stage schemes need the NCCL exchange (peer-pull has no "
										  "acknowledgement before the state is overwritten)");
	for (int s = 0; s < k.ns; ++s) {
		const StageBuf buf = stage_buf(ctx, s);
		const GatherMode mode = s > 0 ? GATHER_NOW : (ctx->y_pending ? GATHER_PENDING_Y : GATHER_NONE);
		// k_0 = f(y_n) may already be resident: the last stage of a "first same as last" scheme
		// (Dormand-Prince, Bogacki-Shampine) or stage 0 of a step that was just undone
		const int k_from = s == 0 ? ctx->k0_slot : -1;
		if (k_from < 0) {
			if ((rc = enqueue_forces(ctx, lp, buf, mode))) return rc;
		} else if (mode == GATHER_PENDING_Y) {	// nobody else waits for the gather of y_n before
			CK(cudaStreamWaitEvent(ctx->sc, ctx->e_gy, 0));	 // this step overwrites its block of Y
			ctx->y_pending = false;
		}
		const bool last = s == k.ns - 1;
		const StageBuf nxt = last ? BUF_Y : stage_buf(ctx, s + 1);
		rc = ctx->prec == 64 ? launch_stage_general<double>(ctx, lp, buf_ptr(ctx, nxt), k, s, k_from)
							 : launch_stage_general<float>(ctx, lp, buf_ptr(ctx, nxt), k, s, k_from);
		if (rc) return rc;
		if (ctx->p2p && (rc = p2p_publish(ctx, nxt))) return rc;
	}
	if (ctx->p2p) {
		ctx->y_stale = true;
	} else if (ctx->world > 1) {
		if ((rc = enqueue_gather(ctx, ctx->Y, ctx->e_gy))) return rc;
		ctx->y_pending = true;
	}
	ctx->k0_slot = (k.fsal && k.ns > 1) ? k.ns - 1 : -1;
	return NBODY_OK;
}

// make the local copy of Y whole (snapshots, downloads, diagnostics).  Never a collective.
int ensure_y_gathered(nbody_ctx* ctx) {
	if (ctx->p2p) {
		if (!ctx->y_stale) return NBODY_OK;
		int rc;
		if ((rc = p2p_wait(ctx, BUF_Y))) return rc;
		const size_t block_bytes = (size_t)ctx->own_pad * 4 * ctx->esz;
		for (int r = 0; r < ctx->world; ++r)
			if (r != ctx->rank)
				CK(cudaMemcpyAsync((char*)ctx->Y + (size_t)r * block_bytes,
								   (const char*)ctx->peerY[r] + (size_t)r * block_bytes, block_bytes,
								   cudaMemcpyDefault, ctx->sc));
		ctx->y_stale = false;
		return NBODY_OK;
	}
	if (!ctx->y_pending) return NBODY_OK;
	CK(cudaStreamWaitEvent(ctx->sc, ctx->e_gy, 0));
	ctx->y_pending = false;
	return NBODY_OK;
}
}  // namespace

// ------------------------------------------------------------------ the C ABI
extern "C" {

const char* nbody_version(void) { return "nbody_b200 0.1 (sm_100a)"; }

int nbody_device_count(void) {
	int ndev = 0, ok = 0;
	if (cudaGetDeviceCount(&ndev) != cudaSuccess) return 0;
	for (int d = 0; d < ndev; ++d) {
		cudaDeviceProp prop;
		if (cudaGetDeviceProperties(&prop, d) == cudaSuccess && prop.major == 10) ++ok;
	}
	return ok;
}

const char* nbody_last_error(const nbody_ctx* ctx) {
	return ctx ? ctx->err.c_str() : g_err.c_str();
}

int nbody_nccl_unique_id(void* out) {
	nbody_ctx* ctx = nullptr;
	if (!out) return fail(ctx, NBODY_ERR_ARG, "nbody_nccl_unique_id: null output");
	if (!g_nccl.load()) return fail(ctx, NBODY_ERR_NCCL, g_nccl.err);
	ncclUniqueId id;
	CKN(g_nccl.GetUniqueId(&id));
	memcpy(out, &id, NBODY_NCCL_UID_BYTES);
	return NBODY_OK;
}

void nbody_destroy(nbody_ctx* ctx) {
	if (!ctx) return;
	cudaSetDevice(ctx->device);
	if (ctx->sc) cudaStreamSynchronize(ctx->sc);
	if (ctx->sm) cudaStreamSynchronize(ctx->sm);
	if (ctx->sx) cudaStreamSynchronize(ctx->sx);
	if (ctx->graph) cudaGraphExecDestroy(ctx->graph);
	if (ctx->p2p)
		for (int r = 0; r < ctx->world; ++r) {
			if (r == ctx->rank) continue;
			void* mapped[] = {ctx->peerY[r], ctx->peerP[0][r], ctx->peerP[1][r], ctx->peer_flags[r]};
			for (void* m : mapped)
				if (m) cudaIpcCloseMemHandle(m);
		}
	if (ctx->comm && g_nccl.lib) g_nccl.CommDestroy(ctx->comm);
	void* bufs[] = {ctx->Y, ctx->P, ctx->vel_n, ctx->vel_s, ctx->acc_r, ctx->acc_v, ctx->apart,
					ctx->rhs_out, ctx->vel_all, ctx->d_aos, ctx->d_energy, ctx->d_counters, ctx->aj,
					ctx->kr, ctx->kv, ctx->d_norms, ctx->undo_Y, ctx->undo_vel,
					ctx->P2, ctx->d_flags, ctx->d_peer_flags};
	for (void* b : bufs)
		if (b) cudaFree(b);
	for (int s = 0; s < NBODY_SNAPSHOT_SLOTS; ++s) {
		if (ctx->snap_dev[s]) cudaFree(ctx->snap_dev[s]);
		if (ctx->snap_host[s]) cudaFreeHost(ctx->snap_host[s]);
		if (ctx->snap_ev[s]) cudaEventDestroy(ctx->snap_ev[s]);
	}
	for (cudaEvent_t e : ctx->kev) cudaEventDestroy(e);
	for (auto& t : ctx->tev) { cudaEventDestroy(t.e0); cudaEventDestroy(t.e1); }
	cudaEvent_t evs[] = {ctx->e_k2, ctx->e_g, ctx->e_gy, ctx->e_snap, ctx->ev_start, ctx->ev_stop};
	for (cudaEvent_t e : evs)
		if (e) cudaEventDestroy(e);
	cudaStream_t ss[] = {ctx->sc, ctx->sm, ctx->sx};
	for (cudaStream_t s : ss)
		if (s) cudaStreamDestroy(s);
	delete ctx;
}

static int create_impl(nbody_ctx* ctx, const void* nccl_uid) {
	int ndev = 0;
	if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
		return fail(ctx, NBODY_ERR_NO_DEVICE,
					"no CUDA device: this library has no CPU fallback (sm_100a required)");
	if (ctx->device < 0 || ctx->device >= ndev)
		return fail(ctx, NBODY_ERR_ARG, "device index out of range");
	cudaDeviceProp prop;
	CK(cudaGetDeviceProperties(&prop, ctx->device));
	if (prop.major != 10)
		return fail(ctx, NBODY_ERR_NO_DEVICE,
					std::string("device '") + prop.name +
						"' is not sm_100 (Blackwell B200); the kernels are built for sm_100a only");
	ctx->n_sm = prop.multiProcessorCount;
	CK(cudaSetDevice(ctx->device));
	CK(cudaStreamCreateWithFlags(&ctx->sc, cudaStreamNonBlocking));
	CK(cudaStreamCreateWithFlags(&ctx->sm, cudaStreamNonBlocking));
	CK(cudaStreamCreateWithFlags(&ctx->sx, cudaStreamNonBlocking));
	CK(cudaEventCreateWithFlags(&ctx->e_k2, cudaEventDisableTiming));
	CK(cudaEventCreateWithFlags(&ctx->e_g, cudaEventDisableTiming));
	CK(cudaEventCreateWithFlags(&ctx->e_gy, cudaEventDisableTiming));
	CK(cudaEventCreateWithFlags(&ctx->e_snap, cudaEventDisableTiming));
	CK(cudaEventCreate(&ctx->ev_start));
	CK(cudaEventCreate(&ctx->ev_stop));
	const size_t v4 = 4 * ctx->esz;
	// SYM_EXTRA_TILES tiles past n_pad: the last block of the pair-symmetric walk may name tiles
	// that hold no bodies; they are fetched (and ignored), so they must be mapped
	const size_t pos_rows = (size_t)ctx->n_pad + SYM_EXTRA_TILES * PAD_UNIT;
	CK(cudaMalloc(&ctx->Y, pos_rows * v4));
	CK(cudaMalloc(&ctx->P, pos_rows * v4));
	CK(cudaMemset((char*)ctx->Y + (size_t)ctx->n_pad * v4, 0, SYM_EXTRA_TILES * PAD_UNIT * v4));
	CK(cudaMemset((char*)ctx->P + (size_t)ctx->n_pad * v4, 0, SYM_EXTRA_TILES * PAD_UNIT * v4));
	CK(cudaMalloc(&ctx->vel_n, (size_t)ctx->own_pad * v4));
	CK(cudaMalloc(&ctx->vel_s, (size_t)ctx->own_pad * v4));
	CK(cudaMalloc(&ctx->acc_r, (size_t)ctx->own_pad * v4));
	CK(cudaMalloc(&ctx->acc_v, (size_t)ctx->own_pad * v4));
	CK(cudaMalloc(&ctx->d_energy, 9 * sizeof(double)));
	CK(cudaMalloc(&ctx->d_counters, 4 * sizeof(unsigned)));
	CK(cudaMemset(ctx->vel_s, 0, (size_t)ctx->own_pad * v4));
	for (int u = 0; u < 2; ++u) {
		for (int i = 0; i < N_VAR64; ++i)
			CK(cudaFuncSetAttribute(g_var64[i].fn[u], cudaFuncAttributeMaxDynamicSharedMemorySize,
									(int)g_var64[i].smem));
		for (int i = 0; i < N_VAR32; ++i)
			CK(cudaFuncSetAttribute(g_var32[i].fn[u], cudaFuncAttributeMaxDynamicSharedMemorySize,
									(int)g_var32[i].smem));
		for (int i = 0; i < N_SYM64; ++i)
			CK(cudaFuncSetAttribute(g_sym64[i].fn[u], cudaFuncAttributeMaxDynamicSharedMemorySize,
									(int)g_sym64[i].smem));
		for (int i = 0; i < N_SYM32; ++i)
			CK(cudaFuncSetAttribute(g_sym32[i].fn[u], cudaFuncAttributeMaxDynamicSharedMemorySize,
									(int)g_sym32[i].smem));
	}
	if (ctx->world > 1) {
		if (ctx->world <= 8) {	// buffers the peer-pull exchange exports (nbody_p2p_export)
			CK(cudaMalloc(&ctx->P2, (size_t)ctx->n_pad * v4));
			CK(cudaMalloc(&ctx->d_flags, 8 * sizeof(unsigned)));
			CK(cudaMemset(ctx->d_flags, 0, 8 * sizeof(unsigned)));
			CK(cudaMalloc(&ctx->d_peer_flags, 8 * sizeof(unsigned*)));
		}
		if (!nccl_uid) return fail(ctx, NBODY_ERR_ARG, "world > 1 needs an NCCL unique id");
		if (!g_nccl.load()) return fail(ctx, NBODY_ERR_NCCL, g_nccl.err);
		ncclUniqueId id;
		memcpy(&id, nccl_uid, NBODY_NCCL_UID_BYTES);
		CKN(g_nccl.CommInitRank(&ctx->comm, ctx->world, id, ctx->rank));
	}
	return NBODY_OK;
}

int nbody_partition(size_t n, int world, int rank, size_t* begin, size_t* end,
					size_t* block_rows) {
	if (n == 0 || world < 1 || rank < 0 || rank >= world || !begin || !end || !block_rows)
		return fail(nullptr, NBODY_ERR_ARG, "nbody_partition: bad argument");
	if ((size_t)world > n)	// class_Solver_N_body.hpp:115-117
		return fail(nullptr, NBODY_ERR_ARG,
					"The number of processes is greater than the problem dimension");
	const size_t per = (n + world - 1) / world;
	const size_t blk = (per + PAD_UNIT - 1) / PAD_UNIT * PAD_UNIT;
	const size_t b = std::min(n, blk * (size_t)rank);
	*begin = b;
	*end = std::min(n, blk * (size_t)(rank + 1));
	*block_rows = blk;
	return NBODY_OK;
}

int nbody_sym_variant_for(size_t n, int precision, int world, int uniform_mass, int deterministic,
						  int* variant, size_t* kernel_block_rows, size_t* rank_block_rows) {
	nbody_ctx* ctx = nullptr;
	if ((precision != 64 && precision != 32) || !variant || !kernel_block_rows || !rank_block_rows)
		return fail(ctx, NBODY_ERR_ARG, "nbody_sym_variant_for: bad argument");
	size_t b = 0, e = 0, blk = 0;
	const int rc = nbody_partition(n, world, 0, &b, &e, &blk);
	if (rc) return rc;
	const SymVariant& sv = sym_variant_pick(precision, n, world, (unsigned)blk, uniform_mass != 0,
											deterministic != 0, -1);
	*variant = (int)(&sv - (precision == 64 ? g_sym64 : g_sym32));
	*kernel_block_rows = (size_t)(sv.threads * sv.ipt);
	*rank_block_rows = blk;
	return NBODY_OK;
}

int nbody_create(nbody_ctx** out, size_t n, int precision, int device, int rank, int world,
				 const void* nccl_uid) {
	if (!out) return fail(nullptr, NBODY_ERR_ARG, "nbody_create: null output");
	*out = nullptr;
	if (n == 0) return fail(nullptr, NBODY_ERR_ARG, "nbody_create: n == 0");
	if (precision != 64 && precision != 32)
		return fail(nullptr, NBODY_ERR_ARG, "nbody_create: precision must be 64 or 32");
	if (world < 1 || rank < 0 || rank >= world)
		return fail(nullptr, NBODY_ERR_ARG, "nbody_create: bad rank/world");
	if ((size_t)world > n)	// class_Solver_N_body.hpp:115-117
		return fail(nullptr, NBODY_ERR_ARG,
					"The number of processes is greater than the problem dimension");
	if (n > 0x7fffffffu - 2u * PAD_UNIT * (unsigned)world)
		return fail(nullptr, NBODY_ERR_ARG, "nbody_create: n too large (32-bit row indices)");
	nbody_ctx* ctx = new nbody_ctx();
	ctx->n = n;
	ctx->prec = precision;
	ctx->esz = precision == 64 ? 8 : 4;
	ctx->device = device;
	ctx->rank = rank;
	ctx->world = world;
	size_t pb = 0, pe = 0, blk = 0;
	nbody_partition(n, world, rank, &pb, &pe, &blk);
	ctx->own_pad = (unsigned)blk;
	ctx->n_pad = ctx->own_pad * (unsigned)world;
	ctx->own0 = ctx->own_pad * (unsigned)rank;
	ctx->own_real = (unsigned)(pe - pb);
	int rc = create_impl(ctx, nccl_uid);
	if (rc != NBODY_OK) {
		std::string e = ctx->err;
		nbody_destroy(ctx);
		g_err = e;
		return rc;
	}
	*out = ctx;
	return NBODY_OK;
}

// ---- peer-pull exchange setup ------------------------------------------------------------
struct P2PHandles {
	cudaIpcMemHandle_t Y, P2, P, flags;
};
static_assert(sizeof(P2PHandles) <= NBODY_P2P_HANDLE_BYTES, "handle blob too small");

int nbody_p2p_export(nbody_ctx* ctx, void* handle_out) {
	if (!ctx || !handle_out) return fail(ctx, NBODY_ERR_ARG, "null argument");
	if (ctx->world < 2 || !ctx->P2)
		return fail(ctx, NBODY_ERR_STATE, "peer-pull exchange needs 2..8 ranks");
	CK(cudaSetDevice(ctx->device));
	P2PHandles h;
	memset(&h, 0, sizeof(h));
	CK(cudaIpcGetMemHandle(&h.Y, ctx->Y));
	CK(cudaIpcGetMemHandle(&h.P2, ctx->P2));
	CK(cudaIpcGetMemHandle(&h.P, ctx->P));
	CK(cudaIpcGetMemHandle(&h.flags, ctx->d_flags));
	memset(handle_out, 0, NBODY_P2P_HANDLE_BYTES);
	memcpy(handle_out, &h, sizeof(h));
	return NBODY_OK;
}

int nbody_p2p_connect(nbody_ctx* ctx, const void* all_handles) {
	if (!ctx || !all_handles) return fail(ctx, NBODY_ERR_ARG, "null argument");
	if (ctx->world < 2 || !ctx->P2)
		return fail(ctx, NBODY_ERR_STATE, "peer-pull exchange needs 2..8 ranks");
	if (ctx->uploaded) return fail(ctx, NBODY_ERR_STATE, "connect before the first upload");
	CK(cudaSetDevice(ctx->device));
	for (int r = 0; r < ctx->world; ++r) {
		if (r == ctx->rank) {
			ctx->peer_flags[r] = ctx->d_flags;
			continue;
		}
		P2PHandles h;
		memcpy(&h, (const char*)all_handles + (size_t)r * NBODY_P2P_HANDLE_BYTES, sizeof(h));
		CK(cudaIpcOpenMemHandle(&ctx->peerY[r], h.Y, cudaIpcMemLazyEnablePeerAccess));
		CK(cudaIpcOpenMemHandle(&ctx->peerP[0][r], h.P2, cudaIpcMemLazyEnablePeerAccess));
		CK(cudaIpcOpenMemHandle(&ctx->peerP[1][r], h.P, cudaIpcMemLazyEnablePeerAccess));
		void* f = nullptr;
		CK(cudaIpcOpenMemHandle(&f, h.flags, cudaIpcMemLazyEnablePeerAccess));
		ctx->peer_flags[r] = (unsigned*)f;
	}
	CK(cudaMemcpy(ctx->d_peer_flags, ctx->peer_flags, 8 * sizeof(unsigned*), cudaMemcpyHostToDevice));
	ctx->p2p = true;
	return NBODY_OK;
}

int nbody_set_params(nbody_ctx* ctx, double G, double eps) {
	if (!ctx) return fail(ctx, NBODY_ERR_ARG, "null context");
	if (!(eps >= 0.)) return fail(ctx, NBODY_ERR_ARG, "eps must be >= 0");
	ctx->G = G;
	ctx->eps = eps;
	ctx->k0_slot = -1;
	if (ctx->graph) { cudaGraphExecDestroy(ctx->graph); ctx->graph = nullptr; }
	return NBODY_OK;
}

int nbody_set_option(nbody_ctx* ctx, const char* name, long value) {
	if (!ctx || !name) return fail(ctx, NBODY_ERR_ARG, "null argument");
	std::string k(name);
	if (k == "time_kernels") ctx->opt_time_kernels = value;
	else if (k == "trace") ctx->opt_trace = value != 0;
	else if (k == "slices") ctx->opt_slices = value;
	else if (k == "variant") {
		const int nv = ctx->prec == 64 ? N_VAR64 : N_VAR32;
		if (value < -1 || value >= nv) return fail(ctx, NBODY_ERR_ARG, "variant out of range");
		ctx->opt_variant = value;
	} else if (k == "use_graph") ctx->opt_use_graph = value;
	else if (k == "pair_symmetry") ctx->opt_pair_symmetry = value < 0 ? -1 : (value != 0);
	else if (k == "sym_variant") {
		const int nv = ctx->prec == 64 ? N_SYM64 : N_SYM32;
		if (value < -1 || value >= nv) return fail(ctx, NBODY_ERR_ARG, "sym_variant out of range");
		ctx->opt_sym_variant = value;
	}
	else if (k == "deterministic") {
		// fixed-point row sums in the pair-symmetric kernel (the one-sided kernel is bit-reproducible
		// as it is); the window needs the input statistics gathered by the NEXT upload
		if (ctx->uploaded && value != 0 && !ctx->stats_valid)
			return fail(ctx, NBODY_ERR_STATE, "set \"deterministic\" before the upload");
		ctx->opt_deterministic = value != 0;
	}
	else if (k == "unimass") {
		if (ctx->uploaded) return fail(ctx, NBODY_ERR_STATE, "set \"unimass\" before the upload");
		ctx->opt_unimass = value;
	}
	else return fail(ctx, NBODY_ERR_ARG, "unknown option " + k);
	ctx->k0_slot = -1;
	if (ctx->graph) { cudaGraphExecDestroy(ctx->graph); ctx->graph = nullptr; }
	return NBODY_OK;
}

int nbody_owned_range(const nbody_ctx* ctx, size_t* begin, size_t* end) {
	if (!ctx || !begin || !end) return NBODY_ERR_ARG;
	*begin = std::min<size_t>(ctx->own0, ctx->n);
	*end = *begin + ctx->own_real;
	return NBODY_OK;
}

// The one thing an upload can change that a captured step has baked into its kernel parameters by
// value: which kernels run (equal masses or not) and the common mass (K2's mass_scale).  The cached
// graph survives an upload that leaves both alone -- the end-to-end loop that re-uploads a state
// every step would otherwise re-capture and re-instantiate it each time -- and is dropped otherwise.
static void set_uniform_mass(nbody_ctx* ctx, bool uni, double m0) {
	const double um = uni ? m0 : 0.;
	if (ctx->graph && (uni != ctx->unimass || um != ctx->uniform_mass)) {
		cudaGraphExecDestroy(ctx->graph);
		ctx->graph = nullptr;
	}
	ctx->unimass = uni;
	ctx->uniform_mass = um;
}

// Option "deterministic": what fx_window needs to know about the input -- largest |mass|, smallest
// non-zero |mass|, largest |coordinate| -- over `count` bodies whose masses are `mstride` elements
// apart and whose x,y,z start at `coords` and are `cstride` elements apart.  acc = {max|m|,
// min|m| != 0 (HUGE_VAL if none), max|x|}; non-finite values are left to the kernels (poison word).
static void scan_stats(nbody_ctx* ctx, const void* masses, size_t mstride, const void* coords,
					   size_t cstride, size_t count, double acc[3]) {
	const bool f64 = ctx->prec == 64;
	auto at = [f64](const void* base, size_t idx) {
		return f64 ? ((const double*)base)[idx] : (double)((const float*)base)[idx];
	};
	double mx = acc[0], mn = acc[1], cx = acc[2];
	for (size_t i = 0; i < count; ++i) {
		const double m = std::fabs(at(masses, i * mstride));
		if (m > mx) mx = m;
		if (m > 0. && m < mn) mn = m;
		const double a = std::max(std::fabs(at(coords, i * cstride)),
								  std::max(std::fabs(at(coords, i * cstride + 1)),
										   std::fabs(at(coords, i * cstride + 2))));
		if (std::isfinite(a) && a > cx) cx = a;
	}
	acc[0] = mx; acc[1] = mn; acc[2] = cx;
}
// the fixed-point builds are asked for: option "deterministic", or an fx entry named explicitly
static bool wants_fx(const nbody_ctx* ctx) {
	if (ctx->opt_deterministic) return true;
	const int nv = ctx->prec == 64 ? N_SYM64 : N_SYM32;
	const long v = ctx->opt_sym_variant;
	return v >= 0 && v < nv && (ctx->prec == 64 ? g_sym64 : g_sym32)[v].fx;
}
static void set_stats(nbody_ctx* ctx, const double acc[3]) {
	ctx->m_absmax = acc[0];
	ctx->m_absmin = acc[1] == HUGE_VAL ? 0. : acc[1];
	ctx->coord_absmax = acc[2];
	ctx->stats_valid = true;
}

// Every body has the same mass (e.g. the Plummer configs, the 4-body test)?  Then the mass is
// factored out of the pair sum.  `stride` = elements between consecutive masses.
static void detect_uniform_mass(nbody_ctx* ctx, const void* masses, size_t stride) {
	bool uni = ctx->opt_unimass != 0;
	double m0 = 0.;
	if (ctx->prec == 64) {
		const double* p = (const double*)masses;
		m0 = p[0];
		for (size_t i = 1; uni && i < ctx->n; ++i) uni = p[i * stride] == p[0];
	} else {
		const float* p = (const float*)masses;
		m0 = p[0];
		for (size_t i = 1; uni && i < ctx->n; ++i) uni = p[i * stride] == p[0];
	}
	uni = uni && std::isfinite(m0);
	set_uniform_mass(ctx, uni, m0);
}

static int upload_common(nbody_ctx* ctx) {
	ctx->undo_valid = false;
	ctx->k0_slot = -1;
	// (the cached graph: see set_uniform_mass, which every upload path goes through)
	if (ctx->P2)  // pad rows and masses of the second stage buffer
		CK(cudaMemcpy(ctx->P2, ctx->P, (size_t)ctx->n_pad * 4 * ctx->esz, cudaMemcpyDeviceToDevice));
	if (ctx->p2p) {	 // peers read this rank's block of Y from here on: publish it
		int rc = p2p_publish(ctx, BUF_Y);
		if (rc) return rc;
		CK(cudaStreamSynchronize(ctx->sc));
		ctx->y_stale = false;  // every rank uploaded the whole array
	}
	ctx->y_pending = false;	 // every rank uploaded the whole array
	ctx->uploaded = true;
	ctx->tm.steps_done = 0;
	return NBODY_OK;
}

int nbody_upload_aos(nbody_ctx* ctx, const void* bodies) {
	if (!ctx || !bodies) return fail(ctx, NBODY_ERR_ARG, "null argument");
	CK(cudaSetDevice(ctx->device));
	const size_t bytes = ctx->n * 7 * ctx->esz;
	if (!ctx->d_aos) CK(cudaMalloc(&ctx->d_aos, bytes + 64));
	CK(cudaMemcpyAsync(ctx->d_aos, bodies, bytes, cudaMemcpyHostToDevice, ctx->sc));
	detect_uniform_mass(ctx, bodies, 7);
	ctx->stats_valid = false;
	if (wants_fx(ctx)) {
		double acc[3] = {0., HUGE_VAL, 0.};
		scan_stats(ctx, bodies, 7, (const char*)bodies + ctx->esz, 7, ctx->n, acc);
		set_stats(ctx, acc);
	}
	const unsigned g = (ctx->n_pad + 255) / 256;
	if (ctx->prec == 64)
		NB_LAUNCH(pack_aos_kernel<double>, dim3(g), dim3(256), 0, ctx->sc,
				  (const double*)ctx->d_aos, ctx->Y, ctx->P, ctx->vel_n, (unsigned)ctx->n, ctx->n_pad,
				  ctx->own0, ctx->own_pad, ctx->unimass ? 1e150 : 0.,
				  ctx->unimass ? ctx->uniform_mass : 0.);
	else
		NB_LAUNCH(pack_aos_kernel<float>, dim3(g), dim3(256), 0, ctx->sc,
				  (const float*)ctx->d_aos, ctx->Y, ctx->P, ctx->vel_n, (unsigned)ctx->n, ctx->n_pad,
				  ctx->own0, ctx->own_pad, ctx->unimass ? 1e18f : 0.f,
				  ctx->unimass ? (float)ctx->uniform_mass : 0.f);
	CK(cudaGetLastError());
	ctx->tm.total_launches++;
	CK(cudaStreamSynchronize(ctx->sc));	 // the host buffer is borrowed for the call only
	return upload_common(ctx);
}

// Every rank hands over ONLY the rows it owns (own_real x Body<T>, nbody_owned_range); the full
// position array is assembled on the devices by one all-gather over NVLink instead of every rank
// pushing all n bodies through PCIe.  Collective: all ranks of the world call it together.
int nbody_upload_aos_owned(nbody_ctx* ctx, const void* owned_bodies) {
	if (!ctx || (!owned_bodies && ctx->own_real)) return fail(ctx, NBODY_ERR_ARG, "null argument");
	if (ctx->world == 1) return nbody_upload_aos(ctx, owned_bodies);
	if (ctx->p2p)
		return fail(ctx, NBODY_ERR_STATE, "nbody_upload_aos_owned needs the NCCL exchange");
	CK(cudaSetDevice(ctx->device));
	const size_t bytes = (size_t)ctx->own_real * 7 * ctx->esz;
	if (!ctx->d_aos) CK(cudaMalloc(&ctx->d_aos, ctx->n * 7 * ctx->esz + 64));
	if (bytes) CK(cudaMemcpyAsync(ctx->d_aos, owned_bodies, bytes, cudaMemcpyHostToDevice, ctx->sc));
	// equal masses everywhere?  min / max of the masses, all-reduced (MIN over {min, -max})
	// (+ with option "deterministic" the input statistics of fx_window: {-max|m|, min|m| != 0,
	// -max|coordinate|} ride along)
	double mm[5] = {HUGE_VAL, HUGE_VAL, 0., HUGE_VAL, 0.};
	ctx->stats_valid = false;
	if (wants_fx(ctx) && ctx->own_real) {
		double acc[3] = {0., HUGE_VAL, 0.};
		scan_stats(ctx, owned_bodies, 7, (const char*)owned_bodies + ctx->esz, 7, ctx->own_real, acc);
		mm[2] = -acc[0]; mm[3] = acc[1]; mm[4] = -acc[2];
	}
	for (unsigned i = 0; i < ctx->own_real; ++i) {
		const double m = ctx->prec == 64 ? ((const double*)owned_bodies)[7 * (size_t)i]
										  : (double)((const float*)owned_bodies)[7 * (size_t)i];
		if (!(m == m)) { mm[0] = -HUGE_VAL; mm[1] = -HUGE_VAL; break; }	 // NaN mass: never uniform
		mm[0] = std::min(mm[0], m);
		mm[1] = std::min(mm[1], -m);
	}
	if (!ctx->d_norms) CK(cudaMalloc(&ctx->d_norms, 8 * sizeof(double)));
	CK(cudaMemcpyAsync(ctx->d_norms, mm, sizeof mm, cudaMemcpyHostToDevice, ctx->sc));
	CKN(g_nccl.AllReduce(ctx->d_norms, ctx->d_norms, 5, NCCL_DOUBLE, NCCL_MIN, ctx->comm, ctx->sc));
	CK(cudaMemcpyAsync(mm, ctx->d_norms, sizeof mm, cudaMemcpyDeviceToHost, ctx->sc));
	CK(cudaStreamSynchronize(ctx->sc));
	if (wants_fx(ctx)) {
		const double acc[3] = {-mm[2], mm[3], -mm[4]};
		set_stats(ctx, acc);
	}
	const bool uni = ctx->opt_unimass != 0 && mm[0] == -mm[1] && std::isfinite(mm[0]);
	set_uniform_mass(ctx, uni, mm[0]);
	const unsigned g = (ctx->own_pad + 255) / 256;
	if (ctx->prec == 64)
		NB_LAUNCH(pack_aos_owned_kernel<double>, dim3(g), dim3(256), 0, ctx->sc,
				  (const double*)ctx->d_aos, ctx->Y, ctx->P, ctx->P2, ctx->vel_n, ctx->own0, ctx->own_real,
				  ctx->own_pad, uni ? 1e150 : 0., uni ? ctx->uniform_mass : 0.);
	else
		NB_LAUNCH(pack_aos_owned_kernel<float>, dim3(g), dim3(256), 0, ctx->sc,
				  (const float*)ctx->d_aos, ctx->Y, ctx->P, ctx->P2, ctx->vel_n, ctx->own0, ctx->own_real,
				  ctx->own_pad, uni ? 1e18f : 0.f, uni ? (float)ctx->uniform_mass : 0.f);
	CK(cudaGetLastError());
	ctx->tm.total_launches++;
	CK(cudaStreamSynchronize(ctx->sc));	 // the host buffer is borrowed for the call only
	ctx->undo_valid = false;
	ctx->k0_slot = -1;
	if (ctx->graph) { cudaGraphExecDestroy(ctx->graph); ctx->graph = nullptr; }
	int rc;
	if ((rc = enqueue_gather(ctx, ctx->Y, ctx->e_gy))) return rc;  // the first force launch waits for it
	ctx->y_pending = true;
	ctx->uploaded = true;
	ctx->tm.steps_done = 0;
	return NBODY_OK;
}

int nbody_upload_soa(nbody_ctx* ctx, const void* m, const void* r4, const void* v4) {
	if (!ctx || !m || !r4 || !v4) return fail(ctx, NBODY_ERR_ARG, "null argument");
	CK(cudaSetDevice(ctx->device));
	const size_t e = ctx->esz, n = ctx->n;
	const size_t bytes = n * 9 * e;	 // m | r4 | v4
	if (!ctx->d_aos) CK(cudaMalloc(&ctx->d_aos, std::max(bytes, n * 7 * e) + 64));
	else {
		CK(cudaFree(ctx->d_aos));
		ctx->d_aos = nullptr;
		CK(cudaMalloc(&ctx->d_aos, std::max(bytes, n * 7 * e) + 64));
	}
	char* d = (char*)ctx->d_aos;
	char* dr = d + (n * e + 31) / 32 * 32;
	char* dv = dr + n * 4 * e;
	CK(cudaMemcpyAsync(d, m, n * e, cudaMemcpyHostToDevice, ctx->sc));
	CK(cudaMemcpyAsync(dr, r4, n * 4 * e, cudaMemcpyHostToDevice, ctx->sc));
	CK(cudaMemcpyAsync(dv, v4, n * 4 * e, cudaMemcpyHostToDevice, ctx->sc));
	detect_uniform_mass(ctx, m, 1);
	ctx->stats_valid = false;
	if (wants_fx(ctx)) {
		double acc[3] = {0., HUGE_VAL, 0.};
		scan_stats(ctx, m, 1, r4, 4, ctx->n, acc);
		set_stats(ctx, acc);
	}
	const unsigned g = (ctx->n_pad + 255) / 256;
	if (ctx->prec == 64)
		NB_LAUNCH(pack_soa_kernel<double>, dim3(g), dim3(256), 0, ctx->sc,
				  (const double*)d, dr, dv, ctx->Y, ctx->P, ctx->vel_n, (unsigned)n, ctx->n_pad,
				  ctx->own0, ctx->own_pad, ctx->unimass ? 1e150 : 0.,
				  ctx->unimass ? ctx->uniform_mass : 0.);
	else
		NB_LAUNCH(pack_soa_kernel<float>, dim3(g), dim3(256), 0, ctx->sc,
				  (const float*)d, dr, dv, ctx->Y, ctx->P, ctx->vel_n, (unsigned)n, ctx->n_pad,
				  ctx->own0, ctx->own_pad, ctx->unimass ? 1e18f : 0.f,
				  ctx->unimass ? (float)ctx->uniform_mass : 0.f);
	CK(cudaGetLastError());
	ctx->tm.total_launches++;
	CK(cudaStreamSynchronize(ctx->sc));
	return upload_common(ctx);
}

static int begin_timed_call(nbody_ctx* ctx) {
	ctx->tm.force_launches = 0;
	ctx->tm.interactions = 0;
	ctx->tm.pair_evaluations = 0;
	ctx->tm.force_ms = 0;
	ctx->kev_used = 0;
	ctx->tev_used = 0;
	CK(cudaEventRecord(ctx->ev_start, ctx->sc));
	return NBODY_OK;
}
static int end_timed_call(nbody_ctx* ctx) {
	CK(cudaEventRecord(ctx->ev_stop, ctx->sc));
	ctx->timing_pending = true;
	return NBODY_OK;
}

int nbody_rhs(nbody_ctx* ctx, void* acc_host) {
	if (!ctx || !acc_host) return fail(ctx, NBODY_ERR_ARG, "null argument");
	if (!ctx->uploaded) return fail(ctx, NBODY_ERR_STATE, "nbody_rhs before upload");
	CK(cudaSetDevice(ctx->device));
	LaunchPlan lp;
	int rc;
	if ((rc = make_plan(ctx, &lp))) return rc;
	if ((rc = ensure_apart(ctx, lp))) return rc;
	if (!ctx->rhs_out) CK(cudaMalloc(&ctx->rhs_out, (size_t)ctx->own_pad * 3 * ctx->esz));
	if ((rc = begin_timed_call(ctx))) return rc;
	if ((rc = enqueue_forces(ctx, lp, BUF_Y, ctx->y_pending ? GATHER_PENDING_Y : GATHER_NONE)))
		return rc;
	if ((rc = stage_dispatch(ctx, lp, ctx->rhs_out, ctx->P, 0., 0., 0., 1, 0))) return rc;
	if ((rc = end_timed_call(ctx))) return rc;
	if (ctx->own_real)
		CK(cudaMemcpyAsync((char*)acc_host + (size_t)ctx->own0 * 3 * ctx->esz, ctx->rhs_out,
						   (size_t)ctx->own_real * 3 * ctx->esz, cudaMemcpyDeviceToHost, ctx->sc));
	CK(cudaStreamSynchronize(ctx->sc));
	return NBODY_OK;
}

static int step_impl(nbody_ctx* ctx, const StepCoef& k, int graph_key, double tau, long n_steps) {
	ctx->undo_valid = false;
	ctx->k0_slot = -1;
	if (n_steps < 0) return fail(ctx, NBODY_ERR_ARG, "n_steps < 0");
	if (!ctx->uploaded) return fail(ctx, NBODY_ERR_STATE, "nbody_step before upload");
	CK(cudaSetDevice(ctx->device));
	LaunchPlan lp;
	int rc;
	if ((rc = make_plan(ctx, &lp))) return rc;
	if ((rc = ensure_apart(ctx, lp))) return rc;
	if ((rc = begin_timed_call(ctx))) return rc;
	long done = 0;
	// Launch-bound regime (small n, single rank): replay one captured step as a CUDA graph
	// (3 nodes per stage: counter memset, K1, K2).  The reference's own 4-body run is pure
	// launch latency (165 us/step in lab_5_N_body_CUDA/output/res_4body_float_64_.csv:2).
	const bool graph_on = ctx->opt_use_graph > 0 || (ctx->opt_use_graph < 0 && ctx->n <= GRAPH_AUTO_MAX_N);
	if (graph_on && graph_key > 0 && ctx->world == 1 && !ctx->opt_time_kernels && !ctx->opt_trace &&
		n_steps >= 1) {
		if (!ctx->graph || ctx->graph_scheme != graph_key || ctx->graph_tau != tau) {
			if (ctx->graph) { cudaGraphExecDestroy(ctx->graph); ctx->graph = nullptr; }
			cudaGraph_t g = nullptr;
			const nbody_timing keep = ctx->tm;
			CK(cudaStreamBeginCapture(ctx->sc, cudaStreamCaptureModeThreadLocal));
			rc = enqueue_step(ctx, lp, k);
			cudaError_t ce = cudaStreamEndCapture(ctx->sc, &g);
			ctx->tm = keep;
			if (rc) return rc;
			CK(ce);
			CK(cudaGraphInstantiate(&ctx->graph, g, 0));
			cudaGraphDestroy(g);
			++ctx->graph_captures;
			ctx->graph_scheme = graph_key;
			ctx->graph_tau = tau;
		}
		for (; done < n_steps; ++done) {
			CK(cudaGraphLaunch(ctx->graph, ctx->sc));
			ctx->tm.force_launches += k.ns;
			ctx->tm.total_launches += 2 * k.ns;
			ctx->tm.interactions += (uint64_t)k.ns * ctx->own_real * ctx->n;
			ctx->tm.pair_evaluations +=
				(uint64_t)k.ns * (lp.sym ? ctx->sym_pairs : (uint64_t)ctx->own_real * ctx->n);
		}
	}
	for (; done < n_steps; ++done)
		if ((rc = enqueue_step(ctx, lp, k))) return rc;
	ctx->tm.steps_done += (uint64_t)n_steps;
	return end_timed_call(ctx);
}

int nbody_step(nbody_ctx* ctx, int scheme, double tau, long n_steps) {
	if (!ctx) return fail(ctx, NBODY_ERR_ARG, "null context");
	if (scheme != NBODY_RK2 && scheme != NBODY_RK4)
		return fail(ctx, NBODY_ERR_ARG, "scheme must be 2 (RK2) or 4 (RK4)");
	return step_impl(ctx, builtin_coef(ctx, scheme, tau), scheme, tau, n_steps);
}

int nbody_step_tableau(nbody_ctx* ctx, int n_stages, const double* a_sub, const double* b,
					   double scale, double tau, long n_steps) {
	if (!ctx || !b || (n_stages > 1 && !a_sub)) return fail(ctx, NBODY_ERR_ARG, "null argument");
	if (n_stages < 1 || n_stages > NBODY_MAX_STAGES)
		return fail(ctx, NBODY_ERR_ARG, "n_stages out of range");
	StepCoef k;
	k.ns = n_stages;
	if (ctx->prec == 64) {
		for (int s = 0; s + 1 < n_stages; ++s) k.c[s] = tau * a_sub[s];
		k.fs = tau * scale;
	} else {
		for (int s = 0; s + 1 < n_stages; ++s) k.c[s] = (float)tau * (float)a_sub[s];
		k.fs = (float)tau * (float)scale;
	}
	for (int s = 0; s < n_stages; ++s) k.b[s] = b[s];
	return step_impl(ctx, k, /*graph_key: tableaux are not cached*/ 0, tau, n_steps);
}

// ---- general tableaux ------------------------------------------------------------------------
static int make_gen_coef(nbody_ctx* ctx, int n_stages, const double* a, const double* b,
						 const double* b_err, double tau, GenCoef* k) {
	if (!a && n_stages > 1) return fail(ctx, NBODY_ERR_ARG, "null tableau");
	if (!b) return fail(ctx, NBODY_ERR_ARG, "null weights");
	if (n_stages < 1 || n_stages > NBODY_MAX_STAGES)
		return fail(ctx, NBODY_ERR_ARG, "n_stages out of range");
	k->ns = n_stages;
	k->want_err = b_err != nullptr;
	const bool f64 = ctx->prec == 64;
	auto mul = [&](double x, double y) { return f64 ? x * y : (double)((float)x * (float)y); };
	for (int s = 0; s < n_stages; ++s) {
		for (int j = 0; j < s; ++j) k->a[s][j] = mul(tau, a[s * n_stages + j]);
		k->w[s] = mul(tau, b[s]);
		k->e[s] = b_err ? mul(tau, f64 ? b[s] - b_err[s] : (double)((float)b[s] - (float)b_err[s])) : 0.;
	}
	// first same as last: the last stage is evaluated at y_n + tau sum_j b[j] k_j = y_{n+1}
	k->fsal = n_stages > 1 && b[n_stages - 1] == 0.0;
	for (int j = 0; k->fsal && j + 1 < n_stages; ++j)
		k->fsal = a[(n_stages - 1) * n_stages + j] == b[j];
	return NBODY_OK;
}

static int ensure_stage_store(nbody_ctx* ctx, int n_stages) {
	if (n_stages <= ctx->k_stages) return NBODY_OK;
	CK(cudaStreamSynchronize(ctx->sc));
	if (ctx->kr) CK(cudaFree(ctx->kr));
	if (ctx->kv) CK(cudaFree(ctx->kv));
	ctx->kr = ctx->kv = nullptr;
	ctx->k_stages = 0;
	ctx->k0_slot = -1;
	const size_t bytes = (size_t)n_stages * ctx->own_pad * 4 * ctx->esz;
	CK(cudaMalloc(&ctx->kr, bytes));
	CK(cudaMalloc(&ctx->kv, bytes));
	ctx->k_stages = n_stages;
	return NBODY_OK;
}

int nbody_step_rk(nbody_ctx* ctx, int n_stages, const double* a, const double* b, double tau,
				  long n_steps) {
	if (!ctx) return fail(ctx, NBODY_ERR_ARG, "null context");
	if (n_steps < 0) return fail(ctx, NBODY_ERR_ARG, "n_steps < 0");
	if (!ctx->uploaded) return fail(ctx, NBODY_ERR_STATE, "nbody_step_rk before upload");
	GenCoef k;
	int rc;
	if ((rc = make_gen_coef(ctx, n_stages, a, b, nullptr, tau, &k))) return rc;
	CK(cudaSetDevice(ctx->device));
	LaunchPlan lp;
	if ((rc = make_plan(ctx, &lp))) return rc;
	if ((rc = ensure_apart(ctx, lp))) return rc;
	if ((rc = ensure_stage_store(ctx, n_stages))) return rc;
	if ((rc = begin_timed_call(ctx))) return rc;
	for (long done = 0; done < n_steps; ++done)
		if ((rc = enqueue_step_general(ctx, lp, k))) return rc;
	ctx->tm.steps_done += (uint64_t)n_steps;
	ctx->undo_valid = false;
	return end_timed_call(ctx);
}

int nbody_step_embedded(nbody_ctx* ctx, int n_stages, const double* a, const double* b,
						const double* b_err, double tau, double norms[4]) {
	if (!ctx || !b_err || !norms) return fail(ctx, NBODY_ERR_ARG, "null argument");
	if (!ctx->uploaded) return fail(ctx, NBODY_ERR_STATE, "nbody_step_embedded before upload");
	if (ctx->p2p)
		return fail(ctx, NBODY_ERR_STATE, "nbody_step_embedded needs the NCCL exchange (undo copies Y)");
	GenCoef k;
	int rc;
	if ((rc = make_gen_coef(ctx, n_stages, a, b, b_err, tau, &k))) return rc;
	CK(cudaSetDevice(ctx->device));
	LaunchPlan lp;
	if ((rc = make_plan(ctx, &lp))) return rc;
	if ((rc = ensure_apart(ctx, lp))) return rc;
	if ((rc = ensure_stage_store(ctx, n_stages))) return rc;
	const size_t v4 = 4 * ctx->esz;
	if (!ctx->undo_Y) {
		CK(cudaMalloc(&ctx->undo_Y, (size_t)ctx->n_pad * v4));
		CK(cudaMalloc(&ctx->undo_vel, (size_t)ctx->own_pad * v4));
	}
	if (!ctx->d_norms) CK(cudaMalloc(&ctx->d_norms, 8 * sizeof(double)));
	// the state this step starts from, whole (every rank holds all positions)
	if ((rc = ensure_y_gathered(ctx))) return rc;
	CK(cudaMemcpyAsync(ctx->undo_Y, ctx->Y, (size_t)ctx->n_pad * v4, cudaMemcpyDeviceToDevice, ctx->sc));
	CK(cudaMemcpyAsync(ctx->undo_vel, ctx->vel_n, (size_t)ctx->own_pad * v4, cudaMemcpyDeviceToDevice,
					   ctx->sc));
	CK(cudaMemsetAsync(ctx->d_norms, 0, 4 * sizeof(double), ctx->sc));
	if ((rc = begin_timed_call(ctx))) return rc;
	if ((rc = enqueue_step_general(ctx, lp, k))) return rc;
	if ((rc = end_timed_call(ctx))) return rc;
	if (ctx->world > 1)
		CKN(g_nccl.AllReduce(ctx->d_norms, ctx->d_norms, 4, NCCL_DOUBLE, NCCL_MAX, ctx->comm, ctx->sc));
	CK(cudaMemcpyAsync(norms, ctx->d_norms, 4 * sizeof(double), cudaMemcpyDeviceToHost, ctx->sc));
	CK(cudaStreamSynchronize(ctx->sc));
	ctx->tm.steps_done += 1;
	ctx->undo_valid = true;
	return NBODY_OK;
}

int nbody_step_undo(nbody_ctx* ctx) {
	if (!ctx) return fail(ctx, NBODY_ERR_ARG, "null context");
	if (!ctx->undo_valid)
		return fail(ctx, NBODY_ERR_STATE, "nbody_step_undo: the last call was not nbody_step_embedded");
	CK(cudaSetDevice(ctx->device));
	const size_t v4 = 4 * ctx->esz;
	if (ctx->y_pending) {  // the eager gather of the rejected y_{n+1} must land before Y is rewritten
		CK(cudaStreamWaitEvent(ctx->sc, ctx->e_gy, 0));
		ctx->y_pending = false;
	}
	CK(cudaMemcpyAsync(ctx->Y, ctx->undo_Y, (size_t)ctx->n_pad * v4, cudaMemcpyDeviceToDevice, ctx->sc));
	CK(cudaMemcpyAsync(ctx->vel_n, ctx->undo_vel, (size_t)ctx->own_pad * v4, cudaMemcpyDeviceToDevice,
					   ctx->sc));
	ctx->tm.steps_done -= 1;
	ctx->undo_valid = false;
	ctx->k0_slot = ctx->k_stages > 1 ? 0 : -1;	 // stage 0 of the rejected step was f(this state)
	return NBODY_OK;
}

int nbody_sync(nbody_ctx* ctx) {
	if (!ctx) return fail(ctx, NBODY_ERR_ARG, "null context");
	CK(cudaSetDevice(ctx->device));
	CK(cudaStreamSynchronize(ctx->sc));
	CK(cudaStreamSynchronize(ctx->sm));
	return NBODY_OK;
}

int nbody_get_timing(nbody_ctx* ctx, nbody_timing* out) {
	if (!ctx || !out) return fail(ctx, NBODY_ERR_ARG, "null argument");
	CK(cudaSetDevice(ctx->device));
	if (ctx->timing_pending) {
		CK(cudaEventSynchronize(ctx->ev_stop));
		float ms = 0;
		CK(cudaEventElapsedTime(&ms, ctx->ev_start, ctx->ev_stop));
		ctx->tm.last_call_ms = ms;
		double f = 0;
		for (size_t k = 0; k + 1 < ctx->kev_used; k += 2) {
			float t = 0;
			CK(cudaEventElapsedTime(&t, ctx->kev[k], ctx->kev[k + 1]));
			f += t;
		}
		ctx->tm.force_ms = f;
		ctx->timing_pending = false;
	}
	ctx->tm.uniform_mass = ctx->unimass ? 1u : 0u;
	ctx->tm.pair_symmetry = sym_in_use(ctx) ? 1u : 0u;
	ctx->tm.graph_captures = ctx->graph_captures;
	ctx->tm.sym_variant = 0;
	ctx->tm.deterministic = sym_in_use(ctx) ? 0u : 1u;	// the one-sided kernel sums in a fixed order
	ctx->tm.reserved0 = 0;
	if (sym_in_use(ctx)) {
		const SymVariant* tab = ctx->prec == 64 ? g_sym64 : g_sym32;
		const SymVariant& sv = sym_variant_of(ctx);
		ctx->tm.sym_variant = (uint32_t)(&sv - tab);
		ctx->tm.deterministic = sv.fx ? 1u : 0u;
	}
	{
		const int nv = ctx->prec == 64 ? N_VAR64 : N_VAR32;
		ctx->tm.variant = (ctx->opt_variant < 0 || ctx->opt_variant >= nv) ? 0u : (uint32_t)ctx->opt_variant;
	}
	*out = ctx->tm;
	return NBODY_OK;
}

int nbody_get_trace(nbody_ctx* ctx, nbody_trace_entry* out, size_t cap, size_t* n_out) {
	if (!ctx || !n_out || (!out && cap)) return fail(ctx, NBODY_ERR_ARG, "null argument");
	CK(cudaSetDevice(ctx->device));
	CK(cudaStreamSynchronize(ctx->sc));
	CK(cudaStreamSynchronize(ctx->sm));
	size_t k = 0;
	for (; k < ctx->tev_used && k < cap; ++k) {
		const nbody_ctx::TraceEv& t = ctx->tev[k];
		float a = 0, d = 0;
		CK(cudaEventElapsedTime(&a, ctx->ev_start, t.e0));
		CK(cudaEventElapsedTime(&d, t.e0, t.e1));
		out[k].kind = t.kind;
		out[k].stream = t.stream;
		out[k].start_ms = a;
		out[k].dur_ms = d;
	}
	*n_out = ctx->tev_used;
	return NBODY_OK;
}

// ------------------------------------------------------------------ K4: snapshots
int nbody_snapshot_async(nbody_ctx* ctx, uint64_t* ticket) {
	if (!ctx || !ticket) return fail(ctx, NBODY_ERR_ARG, "null argument");
	if (!ctx->uploaded) return fail(ctx, NBODY_ERR_STATE, "snapshot before upload");
	CK(cudaSetDevice(ctx->device));
	int slot = -1;
	for (int s = 0; s < NBODY_SNAPSHOT_SLOTS; ++s)
		if (!ctx->snap_state[s].load(std::memory_order_acquire)) { slot = s; break; }
	if (slot < 0) return fail(ctx, NBODY_ERR_STATE, "all snapshot slots are held");
	const size_t bytes = ctx->n * 3 * ctx->esz;
	if (!ctx->snap_dev[0]) {  // first snapshot: set the whole ring up at once
		for (int s = 0; s < NBODY_SNAPSHOT_SLOTS; ++s) {
			CK(cudaMalloc(&ctx->snap_dev[s], bytes));
			CK(cudaHostAlloc(&ctx->snap_host[s], bytes, cudaHostAllocDefault));
			CK(cudaEventCreateWithFlags(&ctx->snap_ev[s], cudaEventDisableTiming));
		}
	}
	int rc;
	if ((rc = ensure_y_gathered(ctx))) return rc;
	const unsigned g = ((unsigned)ctx->n + 255) / 256;
	if (ctx->prec == 64)
		NB_LAUNCH(snapshot_pack_kernel<double>, dim3(g), dim3(256), 0, ctx->sc,
				  ctx->Y, (double*)ctx->snap_dev[slot], (unsigned)ctx->n);
	else
		NB_LAUNCH(snapshot_pack_kernel<float>, dim3(g), dim3(256), 0, ctx->sc,
				  ctx->Y, (float*)ctx->snap_dev[slot], (unsigned)ctx->n);
	CK(cudaGetLastError());
	ctx->tm.total_launches++;
	CK(cudaEventRecord(ctx->e_snap, ctx->sc));
	CK(cudaStreamWaitEvent(ctx->sx, ctx->e_snap, 0));
	CK(cudaMemcpyAsync(ctx->snap_host[slot], ctx->snap_dev[slot], bytes, cudaMemcpyDeviceToHost,
					   ctx->sx));
	CK(cudaEventRecord(ctx->snap_ev[slot], ctx->sx));
	const uint64_t tk = ctx->next_ticket++;
	ctx->snap_ticket[slot].store(tk, std::memory_order_relaxed);
	ctx->snap_state[slot].store(1, std::memory_order_release);
	*ticket = tk;
	return NBODY_OK;
}

static int find_slot(nbody_ctx* ctx, uint64_t ticket) {
	for (int s = 0; s < NBODY_SNAPSHOT_SLOTS; ++s)
		if (ctx->snap_state[s].load(std::memory_order_acquire) &&
			ctx->snap_ticket[s].load(std::memory_order_relaxed) == ticket)
			return s;
	return -1;
}

// These two may be called from a second thread (see nbody_ctx::snap_state).  They touch the slot
// table only, and report errors through the calling thread's own message (nbody_last_error(NULL)):
// the context's message string belongs to the thread that drives the context.
int nbody_snapshot_wait(nbody_ctx* owner, uint64_t ticket, const void** r_host) {
	nbody_ctx* ctx = nullptr;  // what fail / CK write to: the thread-local message only
	if (!owner || !r_host) return fail(ctx, NBODY_ERR_ARG, "null argument");
	const int s = find_slot(owner, ticket);
	if (s < 0) return fail(ctx, NBODY_ERR_ARG, "unknown snapshot ticket");
	CK(cudaEventSynchronize(owner->snap_ev[s]));
	*r_host = owner->snap_host[s];
	return NBODY_OK;
}

int nbody_snapshot_release(nbody_ctx* owner, uint64_t ticket) {
	nbody_ctx* ctx = nullptr;
	if (!owner) return fail(ctx, NBODY_ERR_ARG, "null context");
	const int s = find_slot(owner, ticket);
	if (s < 0) return fail(ctx, NBODY_ERR_ARG, "unknown snapshot ticket");
	CK(cudaEventSynchronize(owner->snap_ev[s]));
	owner->snap_state[s].store(0, std::memory_order_release);
	return NBODY_OK;
}

// ------------------------------------------------------------------ download / energy
int nbody_download_aos(nbody_ctx* ctx, void* bodies) {
	if (!ctx || !bodies) return fail(ctx, NBODY_ERR_ARG, "null argument");
	if (!ctx->uploaded) return fail(ctx, NBODY_ERR_STATE, "download before upload");
	CK(cudaSetDevice(ctx->device));
	int rc;
	if ((rc = ensure_y_gathered(ctx))) return rc;
	const size_t v4 = 4 * ctx->esz;
	const void* vel = ctx->vel_n;
	if (ctx->world > 1) {
		if (!ctx->vel_all) CK(cudaMalloc(&ctx->vel_all, (size_t)ctx->n_pad * v4));
		CK(cudaMemcpyAsync((char*)ctx->vel_all + (size_t)ctx->own0 * v4, ctx->vel_n,
						   (size_t)ctx->own_pad * v4, cudaMemcpyDeviceToDevice, ctx->sc));
		if ((rc = enqueue_gather(ctx, ctx->vel_all, ctx->e_g))) return rc;
		CK(cudaStreamWaitEvent(ctx->sc, ctx->e_g, 0));
		vel = ctx->vel_all;
	}
	const size_t bytes = ctx->n * 7 * ctx->esz;
	if (!ctx->d_aos) CK(cudaMalloc(&ctx->d_aos, bytes + 64));
	const unsigned g = ((unsigned)ctx->n + 255) / 256;
	if (ctx->prec == 64)
		NB_LAUNCH(unpack_aos_kernel<double>, dim3(g), dim3(256), 0, ctx->sc,
				  ctx->Y, vel, (double*)ctx->d_aos, (unsigned)ctx->n);
	else
		NB_LAUNCH(unpack_aos_kernel<float>, dim3(g), dim3(256), 0, ctx->sc,
				  ctx->Y, vel, (float*)ctx->d_aos, (unsigned)ctx->n);
	CK(cudaGetLastError());
	ctx->tm.total_launches++;
	CK(cudaMemcpyAsync(bodies, ctx->d_aos, bytes, cudaMemcpyDeviceToHost, ctx->sc));
	CK(cudaStreamSynchronize(ctx->sc));
	return NBODY_OK;
}

// The owned rows only (own_real x Body<T>): no collective, no copy of other ranks' state.
int nbody_download_aos_owned(nbody_ctx* ctx, void* owned_bodies) {
	if (!ctx || (!owned_bodies && ctx->own_real)) return fail(ctx, NBODY_ERR_ARG, "null argument");
	if (!ctx->uploaded) return fail(ctx, NBODY_ERR_STATE, "download before upload");
	CK(cudaSetDevice(ctx->device));
	if (ctx->own_real) {
		const size_t bytes = (size_t)ctx->own_real * 7 * ctx->esz;
		if (!ctx->d_aos) CK(cudaMalloc(&ctx->d_aos, ctx->n * 7 * ctx->esz + 64));
		const unsigned g = (ctx->own_real + 255) / 256;
		if (ctx->prec == 64)
			NB_LAUNCH(unpack_aos_owned_kernel<double>, dim3(g), dim3(256), 0, ctx->sc,
					  ctx->Y, ctx->vel_n, (double*)ctx->d_aos, ctx->own0, ctx->own_real);
		else
			NB_LAUNCH(unpack_aos_owned_kernel<float>, dim3(g), dim3(256), 0, ctx->sc,
					  ctx->Y, ctx->vel_n, (float*)ctx->d_aos, ctx->own0, ctx->own_real);
		CK(cudaGetLastError());
		ctx->tm.total_launches++;
		CK(cudaMemcpyAsync(owned_bodies, ctx->d_aos, bytes, cudaMemcpyDeviceToHost, ctx->sc));
	}
	CK(cudaStreamSynchronize(ctx->sc));
	return NBODY_OK;
}

int nbody_diagnostics(nbody_ctx* ctx, double out[9]) {
	if (!ctx || !out) return fail(ctx, NBODY_ERR_ARG, "null argument");
	if (!ctx->uploaded) return fail(ctx, NBODY_ERR_STATE, "diagnostics before upload");
	CK(cudaSetDevice(ctx->device));
	int rc;
	if ((rc = ensure_y_gathered(ctx))) return rc;
	CK(cudaMemsetAsync(ctx->d_energy, 0, 9 * sizeof(double), ctx->sc));
	if (ctx->own_real) {
		const unsigned g = (ctx->own_real + 255) / 256;
		const unsigned gm = std::min(g, 4u * (unsigned)ctx->n_sm);
		if (ctx->prec == 64) {
			NB_LAUNCH(energy_kernel<double>, dim3(g), dim3(256), 0, ctx->sc,
					  ctx->Y, ctx->vel_n, (unsigned)ctx->n, ctx->own0, ctx->own_real, ctx->eps, ctx->G,
					  ctx->d_energy);
			NB_LAUNCH(moments_kernel<double>, dim3(gm), dim3(256), 0, ctx->sc,
					  ctx->Y, ctx->vel_n, ctx->own0, ctx->own_real, ctx->d_energy + 2);
		} else {
			NB_LAUNCH(energy_kernel<float>, dim3(g), dim3(256), 0, ctx->sc,
					  ctx->Y, ctx->vel_n, (unsigned)ctx->n, ctx->own0, ctx->own_real, ctx->eps, ctx->G,
					  ctx->d_energy);
			NB_LAUNCH(moments_kernel<float>, dim3(gm), dim3(256), 0, ctx->sc,
					  ctx->Y, ctx->vel_n, ctx->own0, ctx->own_real, ctx->d_energy + 2);
		}
		CK(cudaGetLastError());
		ctx->tm.total_launches += 2;
	}
	if (ctx->world > 1)
		CKN(g_nccl.AllReduce(ctx->d_energy, ctx->d_energy, 9, NCCL_DOUBLE, NCCL_SUM, ctx->comm,
							 ctx->sc));
	CK(cudaMemcpyAsync(out, ctx->d_energy, 9 * sizeof(double), cudaMemcpyDeviceToHost, ctx->sc));
	CK(cudaStreamSynchronize(ctx->sc));
	return NBODY_OK;
}

int nbody_energy(nbody_ctx* ctx, double* kinetic, double* potential) {
	if (!kinetic || !potential) return fail(ctx, NBODY_ERR_ARG, "null argument");
	double d[9];
	const int rc = nbody_diagnostics(ctx, d);
	if (rc) return rc;
	*kinetic = d[0];
	*potential = d[1];
	return NBODY_OK;
}

// ------------------------------------------------------------------ misc
void* nbody_host_alloc(size_t bytes) {
	void* p = nullptr;
	if (cudaHostAlloc(&p, bytes, cudaHostAllocDefault) != cudaSuccess) return nullptr;
	return p;
}
void nbody_host_free(void* p) {
	if (p) cudaFreeHost(p);
}

int nbody_measure_fma_peak(int device, int precision, double ms_target, double* tflops) {
	nbody_ctx* ctx = nullptr;
	if (!tflops || (precision != 64 && precision != 32))
		return fail(ctx, NBODY_ERR_ARG, "nbody_measure_fma_peak: bad argument");
	int ndev = 0;
	if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev)
		return fail(ctx, NBODY_ERR_NO_DEVICE, "no such CUDA device");
	CK(cudaSetDevice(device));
	cudaDeviceProp prop;
	CK(cudaGetDeviceProperties(&prop, device));
	const int blocks = prop.multiProcessorCount * 8, threads = 256;
	void* d = nullptr;
	CK(cudaMalloc(&d, 64));
	cudaEvent_t e0, e1;
	CK(cudaEventCreate(&e0));
	CK(cudaEventCreate(&e1));
	int iters = 256;
	float ms = 0;
	for (int pass = 0; pass < 8; ++pass) {
		CK(cudaEventRecord(e0, 0));
		if (precision == 64)
			NB_LAUNCH(fma_peak_kernel<double>, dim3(blocks), dim3(threads), 0, 0,
					  (double*)d, 0.999999, 1e-9, iters);
		else
			NB_LAUNCH(fma_peak_kernel<float>, dim3(blocks), dim3(threads), 0, 0,
					  (float*)d, 0.9999f, 1e-5f, iters);
		CK(cudaEventRecord(e1, 0));
		CK(cudaEventSynchronize(e1));
		CK(cudaGetLastError());
		CK(cudaEventElapsedTime(&ms, e0, e1));
		if (ms >= ms_target * 0.7) break;
		const double scale = ms > 0.05 ? ms_target / ms : 16.0;
		iters = (int)std::min(2.0e9 / 128, std::max(1.0, iters * std::min(scale, 64.0)));
	}
	const double flops = 2.0 * 8 * 16 * (double)iters * (double)blocks * threads;
	*tflops = flops / (ms * 1e-3) / 1e12;
	cudaEventDestroy(e0);
	cudaEventDestroy(e1);
	cudaFree(d);
	return NBODY_OK;
}

}  // extern "C"

// ------------------------------------------------------------------ nbody_rhs_noalloc
// The acceleration of n bodies for hosts that cannot let this library reserve device memory (an
// application pool that already holds the whole GPU, a fixed per-process budget, a loop that must
// not run into the device-wide synchronisation of cudaMalloc / cudaFree): see include/nbody_b200.h.
// Nothing below calls cudaMalloc, cudaFree or any routine that allocates on the device.  The
// working memory of the pair-symmetric kernel sits in g_noalloc_scratch, an array that is part of
// the library image: every GPU on which this library's code gets loaded carries these 4 MiB until
// the process ends, which is why the entry point stops at NBODY_NOALLOC_MAX_BODIES.
//
// In HOST code the name g_noalloc_scratch is only a handle for the runtime, not a device address
// (passing it to a kernel would make the kernel touch host memory or fault), so the host asks the
// runtime once per device -- with that device current -- where the array lives there
// (cudaGetSymbolAddress) and hands the kernels that pointer.
//
// Layout (bytes):  [0, 256)  work-ticket counter of the force kernel
//                  [256, 256 + MAX*32)                 V4[n_pad] padded {x,y,z,m}
//                  [256 + MAX*32, 256 + MAX*32 + MAX*24)  T[3][n_pad] row sums
// with MAX = NBODY_NOALLOC_MAX_BODIES and the element sizes of FP64 (FP32 uses half of each part).
namespace nbody {
constexpr size_t NOALLOC_POS_OFF = 256;
constexpr size_t NOALLOC_SUM_OFF = NOALLOC_POS_OFF + (size_t)NBODY_NOALLOC_MAX_BODIES * 32;
constexpr size_t NOALLOC_END = NOALLOC_SUM_OFF + (size_t)NBODY_NOALLOC_MAX_BODIES * 24;
static_assert(NOALLOC_END <= NBODY_NOALLOC_SCRATCH_BYTES, "scratch layout exceeds the stated size");
static_assert(NBODY_NOALLOC_MAX_BODIES % 256 == 0, "the limit is a whole number of j-tiles");

static __device__ __align__(256) unsigned char g_noalloc_scratch[NBODY_NOALLOC_SCRATCH_BYTES];

// Everything a call needs from the shared array in the state it must start from: the bodies in the
// kernel's layout (pad rows: zero mass at the origin, they add exactly 0 to every sum), the row
// sums and the ticket counter at zero -- whatever the previous call left there.
template <typename T>
__global__ void __launch_bounds__(256)
	noalloc_prepare_kernel(const T* __restrict__ m, const typename V4<T>::type* __restrict__ r,
						   unsigned char* scratch, unsigned n, unsigned n_pad) {
	typedef typename V4<T>::type vec4;
	vec4* pos = reinterpret_cast<vec4*>(scratch + NOALLOC_POS_OFF);
	T* sums = reinterpret_cast<T*>(scratch + NOALLOC_SUM_OFF);
	const unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i == 0) *reinterpret_cast<unsigned*>(scratch) = 0u;
	if (i >= n_pad) return;
	vec4 q;
	q.x = q.y = q.z = q.w = T(0);
	if (i < n) {
		const vec4 p = r[i];
		q.x = p.x; q.y = p.y; q.z = p.z;
		q.w = m[i];
	}
	pos[i] = q;
	sums[i] = T(0);
	sums[(size_t)n_pad + i] = T(0);
	sums[2 * (size_t)n_pad + i] = T(0);
}

// row sums -> accelerations, the same two roundings as stage_kernel (mass scale 1, then -G)
template <typename T>
__global__ void __launch_bounds__(256)
	noalloc_finish_kernel(const unsigned char* scratch, T* __restrict__ acc, unsigned n,
						  unsigned n_pad, T minusG) {
	const T* sums = reinterpret_cast<const T*>(scratch + NOALLOC_SUM_OFF);
	const unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
#pragma unroll
	for (int c = 0; c < 3; ++c) {
		const T s = add_rn(T(0), sums[(size_t)c * n_pad + i]);
		acc[3 * (size_t)i + c] = mul_rn(mul_rn(s, T(1)), minusG);
	}
}
}  // namespace nbody

namespace {
// device address of g_noalloc_scratch per device (nullptr = not looked up yet) and what the set-up
// learnt about the device on the way
struct NoallocDev {
	unsigned char* scratch = nullptr;
	int n_sm = 0, per_sm[2] = {0, 0};  // resident CTAs per SM of the FP64 / FP32 build
};
constexpr int NOALLOC_MAX_DEVICES = 64;
NoallocDev g_noalloc_dev[NOALLOC_MAX_DEVICES];
std::mutex g_noalloc_mu;

// the 256-row builds of the pair-symmetric kernel (128 threads x 2 i-bodies), general masses
const SymVariant& noalloc_variant(int precision) { return precision == 64 ? g_sym64[3] : g_sym32[2]; }

// once per device, with that device current: where does the array live on THIS GPU?
int noalloc_setup(int device, NoallocDev** out) {
	nbody_ctx* ctx = nullptr;  // CK / fail report through nbody_last_error(NULL)
	std::lock_guard<std::mutex> lock(g_noalloc_mu);
	NoallocDev& d = g_noalloc_dev[device];
	*out = &d;
	if (d.scratch) return NBODY_OK;
	cudaDeviceProp prop;
	CK(cudaGetDeviceProperties(&prop, device));
	if (prop.major != 10)
		return fail(ctx, NBODY_ERR_NO_DEVICE,
					std::string("device '") + prop.name +
						"' is not sm_100 (Blackwell B200); the kernels are built for sm_100a only");
	void* p = nullptr;
	const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
	if (e != cudaSuccess || !p)
		return fail(ctx, NBODY_ERR_CUDA,
					std::string("nbody_rhs_noalloc: cudaGetSymbolAddress(g_noalloc_scratch): ") +
						cudaGetErrorString(e));
	for (int k = 0; k < 2; ++k) {
		const SymVariant& v = noalloc_variant(k == 0 ? 64 : 32);
		CK(cudaFuncSetAttribute(v.fn[0], cudaFuncAttributeMaxDynamicSharedMemorySize, (int)v.smem));
		CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&d.per_sm[k], v.fn[0], v.threads, v.smem));
		if (d.per_sm[k] < 1) return fail(ctx, NBODY_ERR_CUDA, "force kernel does not fit on an SM");
	}
	d.n_sm = prop.multiProcessorCount;
	d.scratch = static_cast<unsigned char*>(p);	 // last: marks the set-up as complete
	return NBODY_OK;
}

template <typename T>
int noalloc_enqueue(const NoallocDev& d, size_t n, const void* m_dev, const void* r4_dev,
					void* acc_dev, double G, double eps, cudaStream_t st) {
	nbody_ctx* ctx = nullptr;
	typedef typename V4<T>::type vec4;
	const int prec = sizeof(T) == 8 ? 64 : 32;
	const SymVariant& v = noalloc_variant(prec);
	const unsigned TJ = 256, block = (unsigned)(v.threads * v.ipt);	 // == TJ: one tile per block
	const unsigned n_pad = (unsigned)((n + TJ - 1) / TJ * TJ);
	const unsigned nb = n_pad / block;
	NB_LAUNCH(noalloc_prepare_kernel<T>, dim3(n_pad / 256), dim3(256), 0, st,
			  (const T*)m_dev, (const vec4*)r4_dev, d.scratch, (unsigned)n, n_pad);
	CK(cudaGetLastError());

	SymParams p;
	p.pos = d.scratch + NOALLOC_POS_OFF;
	p.aj = d.scratch + NOALLOC_SUM_OFF;
	p.rows_pad = n_pad;
	p.nb = nb;
	p.block0 = 0;
	p.nb_own = nb;
	p.jt_total = n_pad / TJ;
	p.n_pad = n_pad;
	p.n_real = (unsigned)n;
	p.counter = reinterpret_cast<unsigned*>(d.scratch);
	p.pad_pos = 0.;
	p.pos_rows = n_pad;	 // with one tile per block the walk names no tile past the last block
	p.fx_e0 = 0;
	// the same slicing rule as make_plan: enough units per resident CTA for a short tail
	const unsigned slots = (unsigned)d.per_sm[prec == 64 ? 0 : 1] * (unsigned)d.n_sm;
	const unsigned h_min = (nb & 1u) ? (nb - 1u) / 2u : nb / 2u - 1u;
	const uint64_t pairs = sym_pair_evaluations((unsigned)n, nb, 1u, 0u, nb);
	p.n_slices = choose_slices(nb, (1u + h_min) * (TJ / 32), slots, 0, MAX_SYM_SLICES,
							   sym_units_per_slot((double)pairs, (unsigned)prec));
	const unsigned long long units = (unsigned long long)nb * p.n_slices;
	const unsigned grid = (unsigned)std::min<unsigned long long>(units, slots);
	v.launch[0](dim3(grid), st, v.smem, p, eps * eps);
	CK(cudaGetLastError());

	NB_LAUNCH(noalloc_finish_kernel<T>, dim3((unsigned)((n + 255) / 256)), dim3(256), 0, st,
			  d.scratch, (T*)acc_dev, (unsigned)n, n_pad, (T)(-(T)G));
	CK(cudaGetLastError());
	return NBODY_OK;
}
}  // namespace

extern "C" int nbody_rhs_noalloc(int device, int precision, size_t n, const void* m_dev,
								 const void* r4_dev, void* acc_dev, double grav_G, double eps,
								 void* stream) {
	nbody_ctx* ctx = nullptr;
	if (precision != 64 && precision != 32)
		return fail(ctx, NBODY_ERR_ARG, "nbody_rhs_noalloc: precision must be 64 or 32");
	if (n > (size_t)NBODY_NOALLOC_MAX_BODIES)
		return fail(ctx, NBODY_ERR_ARG,
					"nbody_rhs_noalloc: n = " + std::to_string(n) + " exceeds NBODY_NOALLOC_MAX_BODIES = " +
						std::to_string(NBODY_NOALLOC_MAX_BODIES) +
						" (its working memory is a fixed array inside the library; use a context, "
						"nbody_create + nbody_rhs, for larger systems)");
	if (n == 0) return NBODY_OK;
	if (!m_dev || !r4_dev || !acc_dev) return fail(ctx, NBODY_ERR_ARG, "nbody_rhs_noalloc: null array");
	int ndev = 0;
	if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
		return fail(ctx, NBODY_ERR_NO_DEVICE,
					"no CUDA device: this library has no CPU fallback (sm_100a required)");
	if (device < 0 || device >= ndev || device >= NOALLOC_MAX_DEVICES)
		return fail(ctx, NBODY_ERR_ARG, "device index out of range");
	CK(cudaSetDevice(device));
	NoallocDev* d = nullptr;
	const int rc = noalloc_setup(device, &d);
	if (rc) return rc;
	cudaStream_t st = static_cast<cudaStream_t>(stream);
	return precision == 64
			   ? noalloc_enqueue<double>(*d, n, m_dev, r4_dev, acc_dev, grav_G, eps, st)
			   : noalloc_enqueue<float>(*d, n, m_dev, r4_dev, acc_dev, grav_G, eps, st);
}
