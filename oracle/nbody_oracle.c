/*
 * nbody_oracle.c -- CPU restatement of the reference N-body hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing in the product (parallel_comp_labs_b200/,
 * the C-ABI library, the C++ host driver) may include, link or call this file.
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs use it, and only as the checker / reported baseline.
 *
 * Parity pinning: every function here is checked bit-for-bit against the
 * reference's own code compiled from /root/reference (oracle/_ref, see
 * oracle/ref_lab4.cpp and tests/test_oracle_vs_reference.py), against the
 * golden fixtures generated from that build (tests/golden/, script
 * tests/golden/make_golden.py) and against the reference's own known-answer
 * files traj{1..4}.csv / res4body_async.csv.
 *
 * All file:line citations are relative to the reference repository root
 * (sanonesan/parallel_comp_labs).
 *
 * Build: gcc -O3 -std=c11 -fPIC -shared -fopenmp (NO -ffast-math, NO -march=native:
 * FMA contraction would change the rounding sequence the reference's
 * `mpicxx -O3 -std=c++17` build (lab_4_N_body_MPI/src/build.sh:1) produces).
 */
#include <math.h>
#include <stddef.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#ifdef _OPENMP
#include <omp.h>
#endif

/* Body<T>: lab_4_N_body_MPI/src/include/N_body/struct_Body.hpp:5-10
 * AoS record {m; r[3]; v[3]} -- 7*T, no padding. */
typedef struct {
	double m;
	double r[3];
	double v[3];
} body_f64;

typedef struct {
	float m;
	float r[3];
	float v[3];
} body_f32;

/* ------------------------------------------------------------------------- */
/* RHS operator, FP64.
 * lab_4_N_body_MPI/src/include/N_body/struct_ode_system.hpp:53-76 (and the
 * identical copy :105-128); arithmetic order from
 * lab_4_N_body_MPI/src/include/N_body/utils/array_utils.hpp:10,37,45,61:
 *   ri_rj = traj[i].r - traj[j].r            (operator-      :45-50)
 *   den   = sqrt(((0 + x*x) + y*y) + z*z)    (arr_norm       :61-65)
 *   den   = max(den*den*den, eps*eps*eps)    (std::max: (a<b)?b:a)
 *   ri_rj *= traj[j].m / den                 (one divide, operator*= :10-13)
 *   a_tmp += ri_rj                           (operator+=     :37-41)
 *   a_tmp *= -G ; diff[i].v = a_tmp ; diff[i].r = traj[i].v
 * diff[i].m is never written (struct_ode_system.hpp:60-75).
 * Rows are independent, so threading over i is bit-neutral. */
void oracle_rhs_f64(const body_f64* traj, body_f64* diff, size_t n, size_t begin,
					size_t end, double eps, double G, int n_threads) {
	const double eps3 = eps * eps * eps;
	const double mG = -G;
	if (n_threads < 1) n_threads = 1;
#pragma omp parallel for schedule(static) num_threads(n_threads) if (n_threads > 1)
	for (ptrdiff_t ii = (ptrdiff_t)begin; ii < (ptrdiff_t)end; ++ii) {
		const size_t i = (size_t)ii;
		double a0 = 0., a1 = 0., a2 = 0.;
		diff[i].r[0] = traj[i].v[0];
		diff[i].r[1] = traj[i].v[1];
		diff[i].r[2] = traj[i].v[2];
		for (size_t j = 0; j < n; ++j) {
			double d0 = traj[i].r[0] - traj[j].r[0];
			double d1 = traj[i].r[1] - traj[j].r[1];
			double d2 = traj[i].r[2] - traj[j].r[2];
			double sum = 0.0;
			sum += d0 * d0;
			sum += d1 * d1;
			sum += d2 * d2;
			double den = sqrt(sum);
			double den3 = den * den * den;
			den = (den3 < eps3) ? eps3 : den3;
			double c = traj[j].m / den;
			d0 = d0 * c;
			d1 = d1 * c;
			d2 = d2 * c;
			a0 += d0;
			a1 += d1;
			a2 += d2;
		}
		diff[i].v[0] = a0 * mG;
		diff[i].v[1] = a1 * mG;
		diff[i].v[2] = a2 * mG;
	}
}

/* RHS operator, FP32, following the GPU flavour's arithmetic:
 * lab_5_N_body_CUDA/src/N_body.cu:163-184 (float typedefs :19-22):
 *   den = x*x + y*y + z*z ; den = sqrtf(den) ; den = den*den*den ;
 *   den = __fdividef(m_j, fmax(den, EPS*EPS*EPS)) ; a += ri_rj * den ;
 *   out = (-G) * a                               (:190-193)
 * __fdividef and device FMA contraction are not bit-reproducible on a CPU, so
 * this restatement pins the FP32 path to FP32 round-off only (tolerance 1e-4
 * relative, BASELINE.json north_star), not bit-for-bit. */
void oracle_rhs_f32(const body_f32* traj, body_f32* diff, size_t n, size_t begin,
					size_t end, float eps, float G, int n_threads) {
	const float eps3 = eps * eps * eps;
	if (n_threads < 1) n_threads = 1;
#pragma omp parallel for schedule(static) num_threads(n_threads) if (n_threads > 1)
	for (ptrdiff_t ii = (ptrdiff_t)begin; ii < (ptrdiff_t)end; ++ii) {
		const size_t i = (size_t)ii;
		float a0 = 0.f, a1 = 0.f, a2 = 0.f;
		diff[i].r[0] = traj[i].v[0];
		diff[i].r[1] = traj[i].v[1];
		diff[i].r[2] = traj[i].v[2];
		for (size_t j = 0; j < n; ++j) {
			float d0 = traj[i].r[0] - traj[j].r[0];
			float d1 = traj[i].r[1] - traj[j].r[1];
			float d2 = traj[i].r[2] - traj[j].r[2];
			float den = d0 * d0 + d1 * d1 + d2 * d2;
			den = sqrtf(den);
			den = den * den * den;
			den = traj[j].m / fmaxf(den, eps3);
			a0 += d0 * den;
			a1 += d1 * den;
			a2 += d2 * den;
		}
		diff[i].v[0] = (-G) * a0;
		diff[i].v[1] = (-G) * a1;
		diff[i].v[2] = (-G) * a2;
	}
}

/* ------------------------------------------------------------------------- */
/* Loop control (finding F4 of SURVEY.md).
 * lab_4 .../schemes/ode_RK2_MPI.hpp:65,72,98-99 and lab_5 N_body.cu:292,299,369-370:
 *   T t = t_start + tau; while (true) { step; if (t > t_final) break; t += tau; }
 * with t accumulated in T, so the step count depends on T's rounding. */
long oracle_step_count_f64(double t_start, double t_final, double tau) {
	long steps = 0;
	double t = t_start + tau;
	while (1) {
		++steps;
		if (t > t_final) break;
		t += tau;
		if (steps > (1L << 40)) return -1;
	}
	return steps;
}

long oracle_step_count_f32(float t_start, float t_final, float tau) {
	long steps = 0;
	float t = t_start + tau;
	while (1) {
		++steps;
		if (t > t_final) break;
		t += tau;
		if (steps > (1L << 40)) return -1;
	}
	return steps;
}

/* Block partition: lab_4 .../class_Solver_N_body.hpp:114-130.
 * count_all = N / P; last rank takes the remainder; offsets[i] = i*count_all.
 * Returns -1 where the reference throws (P > N, :115-117). */
int oracle_partition(size_t n, int n_procs, int* counts, int* offsets) {
	if ((size_t)n_procs > n) return -1;
	size_t count_all = n / (size_t)n_procs;
	size_t count_last = n - count_all * (size_t)(n_procs - 1);
	for (int i = 0; i < n_procs; ++i) counts[i] = (int)count_all;
	counts[n_procs - 1] = (int)count_last;
	for (int i = 0; i < n_procs; ++i) offsets[i] = (int)((size_t)i * count_all);
	return 0;
}

/* ------------------------------------------------------------------------- */
/* Stage arithmetic: lab_4 .../schemes/ode_RK2_MPI.hpp:26-43
 *   mult: vec[i].r *= alpha; vec[i].v *= alpha      (A[i] = A[i]*alpha)
 *   add : vec1[i].r += vec2[i].r; vec1[i].v += vec2[i].v */
static void st_mult(body_f64* k, double alpha, size_t n) {
	for (size_t i = 0; i < n; ++i)
		for (int c = 0; c < 3; ++c) {
			k[i].r[c] = k[i].r[c] * alpha;
			k[i].v[c] = k[i].v[c] * alpha;
		}
}
static void st_add(body_f64* a, const body_f64* b, size_t n) {
	for (size_t i = 0; i < n; ++i)
		for (int c = 0; c < 3; ++c) {
			a[i].r[c] += b[i].r[c];
			a[i].v[c] += b[i].v[c];
		}
}

/* Explicit midpoint (RK2) integrator, single rank (begin=0,end=N).
 * lab_4 .../schemes/ode_RK2_MPI.hpp:55-100:
 *   k = f(sol); k *= tau/2; tmp = sol; tmp += k; k = f(tmp); k *= tau; sol += k;
 *   [row]; if (t > t_final) break; t += tau;
 * The three MPI_Allgatherv calls (:76,83,91) are identities on one rank and
 * do not change values on P ranks (each row is computed by exactly one rank).
 * rows_out (optional): (max_rows) x (1 + 3n) doubles -- the time and positions
 * the reference passes to write_bodies_coords each step (:94-96); the t_start
 * row (:63) is NOT included here. Returns the number of steps executed. */
long oracle_rk2_f64(body_f64* sol, size_t n, double tau, double t_start,
					double t_final, double eps, double G, double* rows_out,
					long max_rows, long max_steps, int n_threads) {
	body_f64* k = (body_f64*)calloc(n, sizeof(body_f64));
	body_f64* tmp = (body_f64*)calloc(n, sizeof(body_f64));
	for (size_t i = 0; i < n; ++i) tmp[i].m = sol[i].m; /* :69-70 */
	long steps = 0;
	double t = t_start + tau;
	while (1) {
		oracle_rhs_f64(sol, k, n, 0, n, eps, G, n_threads);
		st_mult(k, tau / 2, n);
		for (size_t i = 0; i < n; ++i) tmp[i] = sol[i]; /* copy_n :79 */
		st_add(tmp, k, n);
		oracle_rhs_f64(tmp, k, n, 0, n, eps, G, n_threads);
		st_mult(k, tau, n);
		st_add(sol, k, n);
		if (rows_out && steps < max_rows) {
			double* row = rows_out + (size_t)steps * (1 + 3 * n);
			row[0] = t;
			for (size_t i = 0; i < n; ++i)
				for (int c = 0; c < 3; ++c) row[1 + 3 * i + c] = sol[i].r[c];
		}
		++steps;
		if (max_steps > 0 && steps >= max_steps) break;
		if (t > t_final) break;
		t += tau;
	}
	free(k);
	free(tmp);
	return steps;
}

/* Classical RK4 over the same RHS.  NOT IN THE REFERENCE (the
 * reference only implements RK2); BASELINE.json's metric names RK4, so this is
 * "the reference's RHS + the classical tableau", written in the reference's
 * style (mult/add on whole stage vectors, same loop control):
 *   k1 = f(y); tmp = y + (tau/2) k1; k2 = f(tmp); tmp = y + (tau/2) k2;
 *   k3 = f(tmp); tmp = y + tau k3; k4 = f(tmp);
 *   y += (tau/6) * (((k1 + 2 k2) + 2 k3) + k4)
 * The product's fused stage kernel follows this exact rounding sequence. */
long oracle_rk4_f64(body_f64* sol, size_t n, double tau, double t_start,
					double t_final, double eps, double G, double* rows_out,
					long max_rows, long max_steps, int n_threads) {
	body_f64* k = (body_f64*)calloc(n, sizeof(body_f64));
	body_f64* acc = (body_f64*)calloc(n, sizeof(body_f64));
	body_f64* tmp = (body_f64*)calloc(n, sizeof(body_f64));
	body_f64* kk = (body_f64*)calloc(n, sizeof(body_f64));
	for (size_t i = 0; i < n; ++i) tmp[i].m = sol[i].m;
	long steps = 0;
	double t = t_start + tau;
	const double cs[3] = {tau / 2, tau / 2, tau};
	const double bs[4] = {1., 2., 2., 1.};
	while (1) {
		oracle_rhs_f64(sol, k, n, 0, n, eps, G, n_threads); /* k1 */
		memcpy(acc, k, n * sizeof(body_f64));				/* acc = 1*k1 */
		for (int s = 0; s < 3; ++s) {
			memcpy(kk, k, n * sizeof(body_f64));
			st_mult(kk, cs[s], n);
			for (size_t i = 0; i < n; ++i) {
				double m = tmp[i].m;
				tmp[i] = sol[i];
				tmp[i].m = m;
			}
			st_add(tmp, kk, n);
			oracle_rhs_f64(tmp, k, n, 0, n, eps, G, n_threads); /* k_{s+2} */
			memcpy(kk, k, n * sizeof(body_f64));
			st_mult(kk, bs[s + 1], n);
			st_add(acc, kk, n);
		}
		st_mult(acc, tau / 6, n);
		st_add(sol, acc, n);
		if (rows_out && steps < max_rows) {
			double* row = rows_out + (size_t)steps * (1 + 3 * n);
			row[0] = t;
			for (size_t i = 0; i < n; ++i)
				for (int c = 0; c < 3; ++c) row[1 + 3 * i + c] = sol[i].r[c];
		}
		++steps;
		if (max_steps > 0 && steps >= max_steps) break;
		if (t > t_final) break;
		t += tau;
	}
	free(k);
	free(acc);
	free(tmp);
	free(kk);
	return steps;
}

/* Explicit RK with a SUB-DIAGONAL tableau (stage s+1 depends on stage s only):
 *   k_1 = f(y); k_{s+1} = f(y + (tau*a[s]) k_s); y += (tau*scale) * sum_s b[s] k_s
 * NOT IN THE REFERENCE (generalisation of the reference-style RK2/RK4 loops above; Heun,
 * Ralston, Heun's RK3 are of this form).  Rounding sequence: c_s = tau*a[s] and
 * fs = tau*scale formed once; stage vectors multiplied, then added, as in st_mult/st_add. */
long oracle_rk_subdiag_f64(body_f64* sol, size_t n, int n_stages, const double* a,
						   const double* b, double scale, double tau, double t_start,
						   double t_final, double eps, double G, long max_steps, int n_threads) {
	body_f64* k = (body_f64*)calloc(n, sizeof(body_f64));
	body_f64* acc = (body_f64*)calloc(n, sizeof(body_f64));
	body_f64* tmp = (body_f64*)calloc(n, sizeof(body_f64));
	body_f64* kk = (body_f64*)calloc(n, sizeof(body_f64));
	for (size_t i = 0; i < n; ++i) tmp[i].m = sol[i].m;
	long steps = 0;
	double t = t_start + tau;
	const double fs = tau * scale;
	while (1) {
		oracle_rhs_f64(sol, k, n, 0, n, eps, G, n_threads);
		memcpy(acc, k, n * sizeof(body_f64));
		st_mult(acc, b[0], n);
		for (int s = 0; s + 1 < n_stages; ++s) {
			memcpy(kk, k, n * sizeof(body_f64));
			st_mult(kk, tau * a[s], n);
			for (size_t i = 0; i < n; ++i) {
				double m = tmp[i].m;
				tmp[i] = sol[i];
				tmp[i].m = m;
			}
			st_add(tmp, kk, n);
			oracle_rhs_f64(tmp, k, n, 0, n, eps, G, n_threads);
			memcpy(kk, k, n * sizeof(body_f64));
			st_mult(kk, b[s + 1], n);
			st_add(acc, kk, n);
		}
		st_mult(acc, fs, n);
		st_add(sol, acc, n);
		++steps;
		if (max_steps > 0 && steps >= max_steps) break;
		if (t > t_final) break;
		t += tau;
	}
	free(k);
	free(acc);
	free(tmp);
	free(kk);
	return steps;
}

/* General explicit Runge-Kutta step(s) with the whole tableau as data:
 *   k_s = f(y + tau sum_{j<s} a[s][j] k_j),  y_new = y + tau sum_s b[s] k_s,
 *   err = tau sum_s (b[s] - b_err[s]) k_s   (embedded pair; b_err may be NULL).
 * NOT IN THE REFERENCE ("adaptive/other tableaux (RK3/RK45)"); the RHS is
 * the reference operator above.  a is n_stages x n_stages row-major, strictly lower triangle
 * read.  Rounding sequence: every coefficient is first multiplied by tau (tau*a, tau*b,
 * tau*(b-b_err)), stage combinations are accumulated left to right as x = x + (c_j * k_j), zero
 * coefficients skipped.  err_out[4] (last step) = {max|err_r|, max|err_v|, max|r_new|, max|v_new|}. */
long oracle_rk_general_f64(body_f64* sol, size_t n, int n_stages, const double* a, const double* b,
						   const double* b_err, double tau, long n_steps, double eps, double G,
						   int n_threads, double* err_out) {
	body_f64** k = (body_f64**)calloc((size_t)n_stages, sizeof(body_f64*));
	for (int s = 0; s < n_stages; ++s) k[s] = (body_f64*)calloc(n, sizeof(body_f64));
	body_f64* tmp = (body_f64*)calloc(n, sizeof(body_f64));
	long steps = 0;
	for (; steps < n_steps; ++steps) {
		for (int s = 0; s < n_stages; ++s) {
			for (size_t i = 0; i < n; ++i) {
				tmp[i] = sol[i];
				for (int j = 0; j < s; ++j) {
					const double c = tau * a[s * n_stages + j];
					if (c == 0.0) continue;
					for (int d = 0; d < 3; ++d) {
						tmp[i].r[d] = tmp[i].r[d] + c * k[j][i].r[d];
						tmp[i].v[d] = tmp[i].v[d] + c * k[j][i].v[d];
					}
				}
			}
			oracle_rhs_f64(tmp, k[s], n, 0, n, eps, G, n_threads);
		}
		double e[4] = {0, 0, 0, 0};
		for (size_t i = 0; i < n; ++i) {
			double er[3] = {0, 0, 0}, ev[3] = {0, 0, 0};
			for (int j = 0; j < n_stages; ++j) {
				const double c = tau * b[j];
				if (c != 0.0)
					for (int d = 0; d < 3; ++d) {
						sol[i].r[d] = sol[i].r[d] + c * k[j][i].r[d];
						sol[i].v[d] = sol[i].v[d] + c * k[j][i].v[d];
					}
				if (b_err) {
					const double ce = tau * (b[j] - b_err[j]);
					if (ce != 0.0)
						for (int d = 0; d < 3; ++d) {
							er[d] = er[d] + ce * k[j][i].r[d];
							ev[d] = ev[d] + ce * k[j][i].v[d];
						}
				}
			}
			for (int d = 0; d < 3; ++d) {
				if (fabs(er[d]) > e[0]) e[0] = fabs(er[d]);
				if (fabs(ev[d]) > e[1]) e[1] = fabs(ev[d]);
				if (fabs(sol[i].r[d]) > e[2]) e[2] = fabs(sol[i].r[d]);
				if (fabs(sol[i].v[d]) > e[3]) e[3] = fabs(sol[i].v[d]);
			}
		}
		if (err_out) memcpy(err_out, e, sizeof(e));
	}
	for (int s = 0; s < n_stages; ++s) free(k[s]);
	free(k);
	free(tmp);
	return steps;
}

/* FP32 explicit midpoint following the GPU flavour's update form.
 * lab_5_N_body_CUDA/src/N_body.cu:299-316 with mult_n_add (:120-127):
 *   k = f(r,v); r_tmp = r + (tau/2)*k_r; v_tmp = v + (tau/2)*k_v;
 *   k = f(r_tmp,v_tmp); r = r + tau*k_r; v = v + tau*k_v
 * rows_out as above but float (time is printed as T, :366,67-76). */
long oracle_rk2_f32(body_f32* sol, size_t n, float tau, float t_start,
					float t_final, float eps, float G, float* rows_out,
					long max_rows, long max_steps, int n_threads) {
	body_f32* k = (body_f32*)calloc(n, sizeof(body_f32));
	body_f32* tmp = (body_f32*)calloc(n, sizeof(body_f32));
	for (size_t i = 0; i < n; ++i) tmp[i].m = sol[i].m;
	long steps = 0;
	float t = t_start + tau;
	while (1) {
		oracle_rhs_f32(sol, k, n, 0, n, eps, G, n_threads);
		for (size_t i = 0; i < n; ++i)
			for (int c = 0; c < 3; ++c) {
				tmp[i].r[c] = sol[i].r[c] + (tau / 2) * k[i].r[c];
				tmp[i].v[c] = sol[i].v[c] + (tau / 2) * k[i].v[c];
			}
		oracle_rhs_f32(tmp, k, n, 0, n, eps, G, n_threads);
		for (size_t i = 0; i < n; ++i)
			for (int c = 0; c < 3; ++c) {
				sol[i].r[c] = sol[i].r[c] + tau * k[i].r[c];
				sol[i].v[c] = sol[i].v[c] + tau * k[i].v[c];
			}
		if (rows_out && steps < max_rows) {
			float* row = rows_out + (size_t)steps * (1 + 3 * n);
			row[0] = t;
			for (size_t i = 0; i < n; ++i)
				for (int c = 0; c < 3; ++c) row[1 + 3 * i + c] = sol[i].r[c];
		}
		++steps;
		if (max_steps > 0 && steps >= max_steps) break;
		if (t > t_final) break;
		t += tau;
	}
	free(k);
	free(tmp);
	return steps;
}

/* ------------------------------------------------------------------------- */
/* Diagnostics (NOT IN THE REFERENCE; BASELINE.json asks for an energy-drift
 * report).  Potential consistent with the clamp force of
 * struct_ode_system.hpp:69-70: pair term -G m_i m_j / d for d >= eps,
 * -G m_i m_j (3 eps^2 - d^2) / (2 eps^3) for d < eps; i<j only. */
void oracle_energy_f64(const body_f64* b, size_t n, double eps, double G,
					   double* kin, double* pot, int n_threads) {
	double K = 0., U = 0.;
	if (n_threads < 1) n_threads = 1;
#pragma omp parallel for schedule(dynamic, 64) reduction(+ : K, U) num_threads(n_threads) if (n_threads > 1)
	for (ptrdiff_t ii = 0; ii < (ptrdiff_t)n; ++ii) {
		const size_t i = (size_t)ii;
		K += 0.5 * b[i].m *
			 (b[i].v[0] * b[i].v[0] + b[i].v[1] * b[i].v[1] + b[i].v[2] * b[i].v[2]);
		double u = 0.;
		for (size_t j = i + 1; j < n; ++j) {
			double d0 = b[i].r[0] - b[j].r[0];
			double d1 = b[i].r[1] - b[j].r[1];
			double d2 = b[i].r[2] - b[j].r[2];
			double r2 = d0 * d0 + d1 * d1 + d2 * d2;
			double d = sqrt(r2);
			if (d >= eps)
				u += b[j].m / d;
			else
				u += b[j].m * (3. * eps * eps - r2) / (2. * eps * eps * eps);
		}
		U += -G * b[i].m * u;
	}
	*kin = K;
	*pot = U;
}

int oracle_max_threads(void) {
#ifdef _OPENMP
	return omp_get_max_threads();
#else
	return 1;
#endif
}
