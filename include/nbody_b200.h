/*
 * nbody_b200.h -- C ABI of the B200-native N-body hot path (libnbody_b200.so).
 *
 * Drop-in boundary for the ONE data-parallel path of sanonesan/parallel_comp_labs:
 * the all-pairs clamped-softening acceleration evaluated at every stage of the
 * explicit Runge-Kutta N-body integrator (lab_4_N_body_MPI / lab_5_N_body_CUDA).
 * Plain pointers and sizes only; no C++/torch types; never throws; no CPU fallback
 * (nbody_create fails without an sm_100 device).
 *
 * Each entry point cites the reference interface it replaces (file:line relative to
 * the reference repository root).  One context == one GPU == one "rank"
 * (the MPI rank of lab_4).  A context is not thread-safe; distinct contexts may be
 * driven from distinct threads or processes.  ONE exception, made for the trajectory writer:
 * nbody_snapshot_wait and nbody_snapshot_release may be called from one other thread while the
 * thread that drives the context is inside any other call (they share only the snapshot slot
 * table, which is atomic); on that thread their error text is nbody_last_error(NULL).
 *
 * Body layouts accepted at the boundary
 *   AoS  Body<T> {T m; T r[3]; T v[3];}            lab_4 .../struct_Body.hpp:5-10
 *   SoA  T m[n]; T3 r[n]; T3 v[n] (T3 = float4/double4, .w ignored)
 *                                                   lab_5 N_body.cu:19-27,40-62
 * T is double (precision 64) or float (precision 32).
 */
#ifndef NBODY_B200_H
#define NBODY_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct nbody_ctx nbody_ctx;

enum { NBODY_OK = 0, NBODY_ERR_ARG = 1, NBODY_ERR_CUDA = 2, NBODY_ERR_NCCL = 3,
	   NBODY_ERR_NO_DEVICE = 4, NBODY_ERR_STATE = 5 };

/* Integrator schemes.  2 = explicit midpoint, the reference's only scheme
 * (lab_4 .../schemes/ode_RK2_MPI.hpp:72-100, lab_5 N_body.cu:299-316);
 * 4 = classical RK4 over the same RHS (BASELINE.json's metric; not in the reference). */
enum { NBODY_RK2 = 2, NBODY_RK4 = 4 };

#define NBODY_NCCL_UID_BYTES 128

typedef struct nbody_timing {
	double last_call_ms;		 /* CUDA-event time of the last nbody_step / nbody_rhs call */
	double force_ms;			 /* summed CUDA-event time of force-kernel launches in that call
									(only when option "time_kernels" is 1; else 0) */
	uint64_t force_launches;	 /* force-kernel launches in that call */
	uint64_t total_launches;	 /* all kernel launches of this library since create */
	uint64_t interactions;		 /* body-body interactions evaluated in that call, this rank */
	uint64_t steps_done;		 /* integrator steps since the last upload */
	uint32_t uniform_mass;		 /* 1 if the uniform-mass kernels are in use (see "unimass") */
	uint32_t variant;			 /* force-kernel variant index in use */
	uint64_t pair_evaluations;	 /* pair evaluations the kernels actually performed in that call:
									== interactions for the default kernels, about half of it with
									option "pair_symmetry" (each evaluation serves both bodies) */
	uint32_t pair_symmetry;		 /* 1 if the pair-symmetric kernels are in use */
	uint32_t sym_variant;		 /* pair-symmetric kernel variant index in use (0 if not in use) */
	uint64_t graph_captures;	 /* steps captured into a CUDA graph since create: a replayed step
									(option "use_graph") is captured once per scheme, tau and mass
									state, and again only after an option, parameter or upload
									changed what the capture holds by value */
	uint32_t deterministic;		 /* 1 if the row sums of the force kernel in use do not depend on
									scheduling: the one-sided kernel (fixed order) or the
									pair-symmetric kernel with option "deterministic" */
	uint32_t reserved0;
} nbody_timing;

/* Option "trace" = 1: every operation a call enqueues (force launch, fused stage kernel, memsets,
 * all-gather, reduce-scatter) is bracketed by a CUDA-event pair on its stream; nbody_get_trace
 * returns them for the LAST nbody_step* / nbody_rhs call, times relative to the start of the call.
 * A per-rank launch list without a profiler (ncu cannot be run under a multi-rank job). */
enum { NBODY_TRACE_FORCE = 0, NBODY_TRACE_STAGE = 1, NBODY_TRACE_MEMSET = 2, NBODY_TRACE_GATHER = 3,
	   NBODY_TRACE_REDUCE = 4 };
typedef struct nbody_trace_entry {
	uint32_t kind;	 /* NBODY_TRACE_* */
	uint32_t stream; /* 0 compute, 1 comm */
	double start_ms; /* after the call began on the compute stream */
	double dur_ms;
} nbody_trace_entry;

/* Fills out[NBODY_NCCL_UID_BYTES] with a fresh NCCL unique id (rank 0 calls this and
 * distributes the bytes; replaces MPI_Init's rendezvous, lab_4 N_body.cpp:67).  One id
 * creates ONE communicator: every nbody_create of a world needs a fresh id. */
int nbody_nccl_unique_id(void* out);

/* Create the context of rank `rank` of `world` on CUDA device `device`.
 * Replaces Solver_N_body<T>::solve_parallel_MPI's setup (rank/size query, block
 * partition, datatype creation: lab_4 .../class_Solver_N_body.hpp:93-179) and
 * ode_RK2_CUDA's allocation block (lab_5 N_body.cu:257-288).
 * n = number of bodies (every rank holds all n positions and masses, as every MPI
 * rank of lab_4 holds the whole array); precision = 64 or 32; world > n is an error
 * (the reference throws, class_Solver_N_body.hpp:115-117).  nccl_uid may be NULL
 * when world == 1.  On failure *out is NULL and the message is available through
 * nbody_last_error(NULL). */
int nbody_create(nbody_ctx** out, size_t n, int precision, int device, int rank, int world,
				 const void* nccl_uid);
void nbody_destroy(nbody_ctx* ctx);

/* Peer-pull exchange (optional; one process per GPU, 2..8 ranks on one NVLink domain).
 * By default the stage positions travel by one NCCL all-gather per RK stage (the counterpart
 * of lab_4's MPI_Allgatherv, .../schemes/ode_RK2_MPI.hpp:83,91).  After every rank has called
 * nbody_p2p_export, the blobs have been exchanged (rank order) and every rank has called
 * nbody_p2p_connect -- all before the first upload -- there is no collective on the step path
 * any more: the force kernel's bulk-async copies fetch the other ranks' j-tiles straight from
 * their HBM over NVLink, gated by per-rank epoch flags.  nbody_download_aos and
 * nbody_energy / nbody_diagnostics remain collectives. */
#define NBODY_P2P_HANDLE_BYTES 256
int nbody_p2p_export(nbody_ctx* ctx, void* handle_out /* NBODY_P2P_HANDLE_BYTES */);
int nbody_p2p_connect(nbody_ctx* ctx, const void* all_handles /* world x NBODY_P2P_HANDLE_BYTES */);

/* The i-row partition nbody_create uses, as pure host arithmetic (no GPU needed):
 * rank r owns global rows [begin,end); all ranks exchange equal blocks of block_rows
 * rows (padded with zero-mass bodies that contribute exactly 0), block_rows * world >= n.
 * Counterpart of counts[]/offsets[] in lab_4 .../class_Solver_N_body.hpp:119-130 (which
 * gives the remainder to the last rank; NCCL all-gather wants equal counts instead).
 * Returns NBODY_ERR_ARG where the reference throws (world > n). */
int nbody_partition(size_t n, int world, int rank, size_t* begin, size_t* end,
					size_t* block_rows);

/* Which build of the pair-symmetric kernel the automatic choice ("sym_variant" -1) runs for n
 * bodies on `world` GPUs, as pure host arithmetic (no GPU needed): *variant = its index,
 * *kernel_block_rows = the rows one thread block owns, *rank_block_rows = block_rows of
 * nbody_partition.  On several GPUs the ranks own whole kernel blocks, so the choice always
 * satisfies rank_block_rows % kernel_block_rows == 0 (the 256-row builds divide every per-rank
 * block).  No counterpart in the reference (one kernel, BLOCK_SIZE fixed at 64, lab_5
 * N_body.cu:29); lets a host -- and the CPU tests -- check a configuration before any rank
 * enters a collective. */
int nbody_sym_variant_for(size_t n, int precision, int world, int uniform_mass, int deterministic,
						  int* variant, size_t* kernel_block_rows, size_t* rank_block_rows);

/* G and eps captured by the RHS lambda (lab_4 .../struct_ode_system.hpp:34-36,53;
 * lab_5 N_body.cu:22,30).  Defaults 6.67e-11 and 1e-6. */
int nbody_set_params(nbody_ctx* ctx, double grav_G, double eps); /* not "G": lab_5 #defines G (N_body.cu:30) */

/* Named integer options: "time_kernels" (0/1), "slices" (0 = auto j-slices per launch),
 * "variant" (force-kernel variant id, -1 = auto, the default), "use_graph" (-1 = auto: CUDA-graph replay of a step when n <= 131072 on one GPU; 0/1),
 * "unimass" (-1 = at upload detect "all masses equal" and factor the mass out of the pair
 * sum, the default; 0 = never; set before the upload),
 * "pair_symmetry" (which force kernel runs.  0 = every ordered pair is evaluated separately, as
 * the reference does: the one-sided kernel, BIT-REPRODUCIBLE run to run (fixed-order sums);
 * 1 = every pair of distinct bodies is evaluated once and credited to both bodies (Newton's third
 * law) -- same per-pair bits, about half the arithmetic, 1.5x the throughput; both sides are
 * summed with floating-point atomics, so results are reproducible to ROUNDING only (same
 * tolerances against the reference); -1 = automatic, the default: the pair-symmetric kernel for
 * n >= 2048 unless the peer-pull exchange is connected, the one-sided kernel below that.
 * On several GPUs the ranks split the blocks of one global schedule and the sums are
 * reduce-scattered (NCCL exchange only; with "sym_variant" -1 a kernel variant is chosen whose
 * block divides the per-rank block of nbody_partition -- one always exists, see
 * nbody_sym_variant_for -- a forced variant that does not is NBODY_ERR_STATE).  nbody_timing.pair_symmetry reports which kernel ran),
 * "sym_variant" (kernel variant id of that mode, -1 = default),
 * "deterministic" (0 default / 1; set before the upload): the pair-symmetric kernel adds every
 * unit's share of a row as an exact FIXED-POINT integer (192-bit window derived from n, the
 * largest mass and eps; 64-bit integer atomics) instead of a floating-point atomic, and the stage
 * kernel rounds each row sum to T once: BIT-REPRODUCIBLE run to run whatever the scheduling, on
 * any number of GPUs (the reduce-scatter sums integers), same tolerances against the reference.
 * NBODY_ERR_STATE from the step / RHS call if eps == 0 or the input's dynamic range does not fit
 * the window.  The window is FIXED from the state at upload (masses, bounding box): every
 * contribution is truncated at an absolute 2^e0 (e0 = log2(n m_max / eps^2) - 190), and the check
 * that 60 bits remain under the smallest pair term m_min / D^2 is made against the bounding box
 * D of the UPLOADED positions only.  Bodies that leave that box during a long run (Plummer
 * escapers) keep bit-reproducible sums, but the precision floor is then the absolute 2^e0, not the
 * 60 relative bits; it is not re-derived after nbody_step_undo or an adaptive step either --
 * upload again to re-derive it.  No effect where the one-sided kernel runs (it is bit-reproducible
 * as it is).
 * nbody_timing.deterministic reports what the kernel in use guarantees). */
int nbody_set_option(nbody_ctx* ctx, const char* name, long value);

/* Rows [begin,end) this rank owns (lab_4's offsets[id], counts[id],
 * class_Solver_N_body.hpp:119-130; here equal-sized padded blocks). */
int nbody_owned_range(const nbody_ctx* ctx, size_t* begin, size_t* end);

/* Upload the whole state.  AoS: n x Body<T> = sys.traj_init
 * (lab_4 .../schemes/ode_RK2_MPI.hpp:55-56).  SoA: lab_5's m, r, v host vectors
 * (N_body.cu:279-288).  Host pointers are borrowed for the call only. */
int nbody_upload_aos(nbody_ctx* ctx, const void* bodies);
int nbody_upload_soa(nbody_ctx* ctx, const void* m, const void* r4, const void* v4);
/* Scalable variant for several GPUs: every rank passes ONLY the rows it owns (own_real x Body<T>,
 * rows [begin,end) of nbody_owned_range -- &sys.traj_init[offsets[id]] in lab_4's terms,
 * class_Solver_N_body.hpp:119-130); the full position array is assembled on the devices by one
 * all-gather over NVLink (what MPI_Allgatherv of `sol` does at ode_RK2_MPI.hpp:91) instead of
 * every rank pushing all n bodies through PCIe.  Collective over the world; NCCL exchange only. */
int nbody_upload_aos_owned(nbody_ctx* ctx, const void* owned_bodies);

/* Test hook == sys.N_body_func(traj, diff, begin, end) with [begin,end) = owned rows
 * (lab_4 .../struct_ode_system.hpp:27-30,53-76; lab_5 calc_acceleration N_body.cu:134-195).
 * Writes diff[i].v (acceleration, 3 x T per row) for owned rows into acc_host[3*i..],
 * i global; rows not owned are left untouched.  Synchronous. */
int nbody_rhs(nbody_ctx* ctx, void* acc_host);

/* Advance n_steps steps of size tau with the given scheme; everything stays on the
 * device.  Replaces the body of the time loop (lab_4 ode_RK2_MPI.hpp:72-92, lab_5
 * N_body.cu:299-316).  Loop control stays with the caller.
 * Asynchronous: returns once the work is enqueued; nbody_sync waits.  Collective:
 * all ranks of the world must call it with the same arguments. */
int nbody_step(nbody_ctx* ctx, int scheme, double tau, long n_steps);
int nbody_sync(nbody_ctx* ctx);

/* The same fused stage kernel with the tableau as data (not in the
 * reference): explicit RK schemes whose stage s+1 depends on stage s only,
 *   k_1 = f(y), k_{s+1} = f(y + tau a_sub[s] k_s),  y += tau scale sum_s b[s] k_s,
 * e.g. Heun {a={1}, b={1,1}, scale=1/2}, Ralston {a={2/3}, b={1,3}, scale=1/4}, Heun's RK3
 * {a={1/3,2/3}, b={1,0,3}, scale=1/4}; midpoint and classical RK4 are nbody_step. */
#define NBODY_MAX_STAGES 8
int nbody_step_tableau(nbody_ctx* ctx, int n_stages, const double* a_sub /* n_stages-1 */,
					   const double* b /* n_stages */, double scale, double tau, long n_steps);

/* General explicit Runge-Kutta schemes, the whole Butcher tableau as data (SURVEY.md 8f row 4:
 * "adaptive/other tableaux (RK3/RK45)"; not in the reference, whose only scheme is the midpoint
 * rule of ode_RK2_MPI.hpp:72-92).  All stage derivatives of the owned rows stay resident;
 *   k_s = f(y + tau sum_{j<s} a[s][j] k_j),   y += tau sum_s b[s] k_s.
 * a: n_stages x n_stages, row-major, only the strictly lower triangle is read; b: n_stages.
 * Same RHS kernels and exchange as nbody_step; one fused stage kernel per stage.  A force
 * evaluation whose result is still resident is not repeated: "first same as last" tableaux
 * (a[s-1][j] == b[j], b[s-1] == 0: Dormand-Prince, Bogacki-Shampine) reuse the last stage as the
 * next step's first, and the step after nbody_step_undo reuses the rejected step's first stage. */
int nbody_step_rk(nbody_ctx* ctx, int n_stages, const double* a, const double* b, double tau,
				  long n_steps);

/* ONE step of an embedded pair (b: weights of the solution that is propagated, b_err: weights of
 * the companion): additionally returns, synchronously, the max-norms over all bodies and
 * components  norms = {|err_r|, |err_v|, |r_new|, |v_new|},  err = tau sum_s (b[s]-b_err[s]) k_s,
 * and keeps the state the step started from so that a step-size controller on the host can
 * reject it with nbody_step_undo (valid until the next call that changes the state).
 * Collective when world > 1 (an all-reduce of the norms); NCCL exchange only. */
int nbody_step_embedded(nbody_ctx* ctx, int n_stages, const double* a, const double* b,
						const double* b_err, double tau, double norms[4]);
int nbody_step_undo(nbody_ctx* ctx);

/* Asynchronous position snapshot (replaces the per-step cudaMemcpyAsync D2H +
 * write_bodies_coords of lab_5 N_body.cu:352-367 and rank 0's row write of lab_4
 * ode_RK2_MPI.hpp:94-96).  Enqueues a device-side pack of all n positions (3 x T per
 * body) and a copy into a pinned ring slot; returns a ticket.  NOT collective: after every
 * step each rank already holds all positions (the end-of-step all-gather), so rank 0 alone
 * may take snapshots, as rank 0 alone writes rows in the reference.  nbody_snapshot_wait
 * blocks until that copy has landed and returns the pinned pointer, valid until
 * nbody_snapshot_release.  The ring has NBODY_SNAPSHOT_SLOTS slots; asking for a
 * snapshot while all are held is NBODY_ERR_STATE.  _wait / _release: callable from a second
 * thread, see the threading note at the top; their errors go to nbody_last_error(NULL). */
#define NBODY_SNAPSHOT_SLOTS 4
int nbody_snapshot_async(nbody_ctx* ctx, uint64_t* ticket);
int nbody_snapshot_wait(nbody_ctx* ctx, uint64_t ticket, const void** r_host);
int nbody_snapshot_release(nbody_ctx* ctx, uint64_t ticket);

/* Whole state back as n x Body<T> (superset of the reference, whose final state is
 * only visible through the CSV; needed for parity on velocities).  Collective when
 * world > 1 (velocities are owner-local and are gathered here): all ranks must call it. */
int nbody_download_aos(nbody_ctx* ctx, void* bodies);
/* The owned rows only (own_real x Body<T>): not a collective, nothing but this rank's block moves. */
int nbody_download_aos_owned(nbody_ctx* ctx, void* owned_bodies);

/* Kinetic and potential energy of the current state, potential consistent with the
 * clamp force.  Not in the reference.  Collective when world > 1. */
int nbody_energy(nbody_ctx* ctx, double* kinetic, double* potential);

/* Energy plus first moments, out[9] = {kinetic, potential, px, py, pz (sum m v),
 * mx, my, mz (sum m r; centre of mass = that / M), M (sum m)}, accumulated in FP64.
 * Not in the reference.  Collective when world > 1. */
int nbody_diagnostics(nbody_ctx* ctx, double out[9]);

/* Acceleration of n bodies WITHOUT a context and WITHOUT any device allocation.
 * No counterpart in the reference (its programs cudaMalloc whatever they need,
 * lab_5 N_body.cu:257-288); the argument list is the one lab_5's calc_acceleration takes.
 *
 * Who needs it: every other compute entry point works on an nbody_ctx, and nbody_create (and a few
 * calls after it, lazily) reserve device memory with cudaMalloc.  Some hosts cannot afford that:
 * an application whose own pool already holds all of the GPU's memory (a cudaMalloc from this
 * library would fail or evict nothing), a service with a fixed per-process memory budget, or a
 * latency-sensitive loop that must not hit the device-wide synchronisation and driver lock that
 * cudaMalloc / cudaFree imply.  Such a host already OWNS device arrays of masses and positions; it
 * passes them in, together with the array the accelerations go to, and this call does the rest.
 *
 * What it computes: the operator of nbody_rhs (struct_ode_system.hpp:60-75: clamped softening,
 * times -G) for all n rows, by the pair-symmetric force kernel in its 256-row build with general
 * masses.  The kernel's working memory -- a padded {x,y,z,m} copy of the bodies, the row sums both
 * sides of a pair are added to, the work-ticket counter -- lives in ONE array that is part of the
 * library image (about NBODY_NOALLOC_SCRATCH_BYTES on every GPU the library is used on, from the
 * moment the library's code is loaded there until the process ends).  Hence the limit:
 * n <= NBODY_NOALLOC_MAX_BODIES, larger n is NBODY_ERR_ARG and nothing is launched.
 *
 *   device      CUDA device the three arrays live on (made current for the call)
 *   precision   64: T = double, T4 = double4;   32: T = float, T4 = float4
 *   m_dev       device, T[n]   masses
 *   r4_dev      device, T4[n]  positions, .w ignored           (16-/32-byte aligned)
 *   acc_dev     device, T[3n]  accelerations, 3 x T per body as nbody_rhs writes them
 *   stream      a cudaStream_t passed as void*; NULL = the default stream
 *
 * The work (three kernels) is enqueued on `stream` and the call returns; the caller synchronises
 * the stream before it reads acc_dev.  Sums are floating-point atomics as in the default kernel:
 * equal to nbody_rhs up to rounding, not bit for bit.
 * The first call on a device looks the array's device address up and records it (a few
 * microseconds, no allocation); make that first call outside any stream capture.
 * The array is shared by all calls and keeps what the last call left in it; each call clears the
 * parts that must start at zero itself.  Two calls whose work can be on the same GPU at the same
 * time (different streams, or different host threads with different streams) would overwrite each
 * other's working memory: order them on one stream or with events. */
#define NBODY_NOALLOC_MAX_BODIES 65536
#define NBODY_NOALLOC_SCRATCH_BYTES (4u << 20)
int nbody_rhs_noalloc(int device, int precision, size_t n, const void* m_dev, const void* r4_dev,
					  void* acc_dev, double grav_G, double eps, void* stream);

int nbody_get_timing(nbody_ctx* ctx, nbody_timing* out);
/* Up to `cap` entries of the last call's trace (option "trace"); *n_out = entries recorded. Syncs. */
int nbody_get_trace(nbody_ctx* ctx, nbody_trace_entry* out, size_t cap, size_t* n_out);

/* Pinned host memory for callers that want zero-staging transfers. */
void* nbody_host_alloc(size_t bytes);
void nbody_host_free(void* p);

/* Sustained FMA-pipe peak of this device in TFLOP/s (2 flops per FMA), measured by a
 * register-resident dependent-chain microbenchmark of `ms_target` milliseconds:
 * the roofline denominator for the force kernel (MEASURED_PEAKS.json has no FP64/FP32
 * vector figure).  precision = 64 or 32. */
int nbody_measure_fma_peak(int device, int precision, double ms_target, double* tflops);

const char* nbody_last_error(const nbody_ctx* ctx);
const char* nbody_version(void);
/* Number of sm_100 devices this process can use (0 if none or no driver): lets a multi-GPU host
 * validate `--gpus` BEFORE it spawns ranks that would otherwise wait for each other forever in
 * the communicator set-up (the reference has no counterpart: mpirun sizes the world). */
int nbody_device_count(void);

#ifdef __cplusplus
}
#endif
#endif /* NBODY_B200_H */
