/*
 * ref_lab4.cpp -- thin extern "C" driver around the REFERENCE'S OWN lab_4 code,
 * compiled from the sources where they lie under /root/reference (nothing is
 * copied into this repository).  TEST INFRASTRUCTURE ONLY; output goes to
 * oracle/_ref/libref_lab4.so (git-ignored, travels to the GPU box).
 *
 * What comes from the reference (unmodified, via #include):
 *   Body<T>                         include/N_body/struct_Body.hpp:5-10
 *   ode_system<T> + N_body_func     include/N_body/struct_ode_system.hpp:20-134
 *   ode_RK2_MPI                     include/N_body/schemes/ode_RK2_MPI.hpp:19-107
 *   Solver_N_body::solve_parallel_MPI  include/N_body/class_Solver_N_body.hpp:87-208
 *   read_data / write_bodies_coords include/N_body/utils/read_write_utils.hpp:12-52
 * What is supplied here: mpi.h (threads-as-ranks shim), config.hpp, and the
 * RK4 driver (the reference has no RK4 -- SURVEY.md F2 -- so ref_rk4 is the
 * classical tableau written over the reference's own N_body_func lambda).
 */
#include <mpi.h>
#include <sys/stat.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <string>
#include <vector>

#include "include/N_body/class_Solver_N_body.hpp"
#include "include/N_body/struct_Body.hpp"
#include "include/N_body/struct_ode_system.hpp"

static_assert(sizeof(Body<double>) == 56, "Body<double> must be 7 packed doubles");

namespace {

std::string scratch_dir() {
	const char* e = std::getenv("NBODY_REF_SCRATCH");
	std::string d = e ? e : "/tmp/nbody_ref";
	mkdir(d.c_str(), 0777);
	mkdir((d + "/output").c_str(), 0777);
	return d;
}

/* The factory reads <PATH_output_folder>/rand_bodies.txt (struct_ode_system.hpp:101-103)
 * and throws int 1 if it cannot; give it a one-body placeholder and then
 * overwrite traj_init with the caller's exact binary state. */
void make_system(ode_system<double>& sys, const double* aos, std::size_t n, double tau,
				 double t_start, double t_final, double eps, double G,
				 const char* name) {
	std::string d = scratch_dir();
	config_lab_4_MPI::PATH_main_folder = d;
	config_lab_4_MPI::PATH_src_folder = d + "/src";
	config_lab_4_MPI::PATH_output_folder = d + "/output";
	{
		std::ifstream probe(d + "/output/rand_bodies.txt");
		if (!probe.is_open()) {
			std::ofstream f(d + "/output/rand_bodies.txt");
			f << "1\n1 0 0 0 0 0 0\n";
		}
	}
	sys.ode_system_random_data(n, tau, eps, G);	 // installs the reference RHS lambda
	sys.traj_init.resize(n);
	std::memcpy(sys.traj_init.data(), aos, n * sizeof(Body<double>));
	sys.t_start = t_start;
	sys.t_final = t_final;
	sys.tau = tau;
	sys.n_time_steps = (int)((t_final - t_start) / tau);
	sys.N_bodies = (int)n;
	sys.name = name;
}

}  // namespace

extern "C" {

/* Reference RHS on rows [begin,end): struct_ode_system.hpp:105-128. */
int ref_rhs(const double* aos, double* diff_aos, std::size_t n, std::size_t begin,
			std::size_t end, double eps, double G) {
	try {
		ode_system<double> sys;
		make_system(sys, aos, n, 0.1, 0., 1., eps, G, "4Body_random");
		std::vector<Body<double>> diff(n);
		std::memcpy(diff.data(), diff_aos, n * sizeof(Body<double>));
		sys.N_body_func(sys.traj_init, diff, begin, end);
		std::memcpy(diff_aos, diff.data(), n * sizeof(Body<double>));
		return 0;
	} catch (...) {
		return 1;
	}
}

/* One RHS evaluation the way P MPI ranks do it inside an RK stage: rank r calls the
 * reference's N_body_func(traj, diff, begin_r, end_r) on its own copy of the bodies, with
 * begin_r = r * rows_per_rank (a BOUNDED sample of the rows when rows_per_rank * n_ranks < n;
 * the timed CPU baseline of bench.py).  diff_aos (n x 7) receives the computed rows.
 * Returns wall seconds of the slowest rank (MPI_Wtime pair around the call), <0 on error. */
double ref_rhs_ranks(const double* aos, double* diff_aos, std::size_t n,
					 std::size_t rows_per_rank, int n_ranks, double eps, double G) {
	if (rows_per_rank * (std::size_t)n_ranks > n) return -2.;
	try {
		std::vector<ode_system<double>> syss(n_ranks);
		std::vector<std::vector<Body<double>>> diffs(n_ranks);
		for (int r = 0; r < n_ranks; ++r) {
			make_system(syss[r], aos, n, 0.1, 0., 1., eps, G, "4Body_random");
			diffs[r].resize(n);
		}
		std::vector<double> t(n_ranks, 0.);
		mpi_shim::run(n_ranks, [&](int rank) {
			const std::size_t b = rows_per_rank * (std::size_t)rank;
			MPI_Barrier(MPI_COMM_WORLD);
			const double t0 = MPI_Wtime();
			syss[rank].N_body_func(syss[rank].traj_init, diffs[rank], b, b + rows_per_rank);
			t[rank] = MPI_Wtime() - t0;
		});
		double wall = 0.;
		for (int r = 0; r < n_ranks; ++r) {
			const std::size_t b = rows_per_rank * (std::size_t)r;
			if (diff_aos)
				std::memcpy(diff_aos + 7 * b, diffs[r].data() + b,
							rows_per_rank * sizeof(Body<double>));
			if (t[r] > wall) wall = t[r];
		}
		return wall;
	} catch (...) {
		return -1.;
	}
}

/* Reference RK2 integrator run on n_ranks thread-ranks through the reference's
 * ode_RK2_MPI, with the reference's block partition (class_Solver_N_body.hpp:119-130)
 * restated for the call.  use_bodyr != 0 gathers positions only (MPI_BODYR, the
 * 4-body flavour :150-170,188); else the whole Body (MPI_BODY :141-148,192).
 * csv_path (may be "") receives the reference's trajectory CSV (rank 0).
 * Returns wall seconds measured by the reference's own MPI_Wtime pair, <0 on error. */
double ref_rk2_run(const double* aos, std::size_t n, double tau, double t_start,
				   double t_final, double eps, double G, int n_ranks, int use_bodyr,
				   const char* csv_path) {
	if ((std::size_t)n_ranks > n) return -2.;
	double wall = -1.;
	try {
		std::vector<ode_system<double>> syss(n_ranks);
		for (int r = 0; r < n_ranks; ++r)
			make_system(syss[r], aos, n, tau, t_start, t_final, eps, G, "4Body_random");
		std::vector<int> counts(n_ranks), offsets(n_ranks);
		std::size_t count_all = n / n_ranks;
		for (int r = 0; r < n_ranks; ++r) {
			counts[r] = (int)count_all;
			offsets[r] = (int)(r * count_all);
		}
		counts[n_ranks - 1] = (int)(n - count_all * (n_ranks - 1));

		MPI_Datatype body, bodyr_tmp, bodyr;
		{
			MPI_Aint dis[3] = {offsetof(Body<double>, m), offsetof(Body<double>, r),
							   offsetof(Body<double>, v)};
			int len[3] = {1, 3, 3};
			MPI_Datatype ty[3] = {MPI_DOUBLE, MPI_DOUBLE, MPI_DOUBLE};
			MPI_Type_create_struct(3, len, dis, ty, &body);
			MPI_Aint dis1[1] = {offsetof(Body<double>, r)};
			int len1[1] = {3};
			MPI_Type_create_struct(1, len1, dis1, ty, &bodyr_tmp);
			MPI_Type_create_resized(bodyr_tmp, 0, sizeof(Body<double>), &bodyr);
		}
		std::string path = csv_path ? csv_path : "";
		std::vector<double> timers(n_ranks, 0.);
		mpi_shim::run(n_ranks, [&](int rank) {
			std::size_t iter = 0;
			ode_RK2_MPI<double>(syss[rank], iter, timers[rank], use_bodyr ? bodyr : body,
								counts.data(), offsets.data(), (std::size_t)rank, path);
		});
		wall = timers[0];
	} catch (...) {
		return -1.;
	}
	return wall;
}

/* The reference's whole harness (Solver_N_body::solve_parallel_MPI, "Server"
 * flavour, name "4Body_random": timing CSV only, no trajectory) on n_ranks
 * thread-ranks -- what `mpirun -np P ./N_body_server` runs after main()'s setup
 * (lab_4_N_body_MPI/src/N_body.cpp:61-84).  Returns the reference's own timer. */
double ref_solve_parallel(const double* aos, std::size_t n, double tau, double t_start,
						  double t_final, double eps, double G, int n_ranks) {
	double wall = -1.;
	try {
		std::vector<ode_system<double>> syss(n_ranks);
		for (int r = 0; r < n_ranks; ++r)
			make_system(syss[r], aos, n, tau, t_start, t_final, eps, G, "4Body_random");
		std::vector<Solver_N_body<double>> solvers(n_ranks);
		mpi_shim::run(n_ranks, [&](int rank) {
			solvers[rank].notifications = false;
			solvers[rank].solve_parallel_MPI(syss[rank], "Server");
		});
		wall = solvers[0].timer;
	} catch (...) {
		return -1.;
	}
	return wall;
}

/* Classical RK4 over the reference's own N_body_func (NOT a reference
 * integrator; see header).  In/out state, optional position rows as in
 * oracle_rk4_f64.  Rounding sequence identical to oracle/nbody_oracle.c. */
long ref_rk4(double* aos, std::size_t n, double tau, double t_start, double t_final,
			 double eps, double G, double* rows_out, long max_rows, long max_steps) {
	try {
		ode_system<double> sys;
		make_system(sys, aos, n, tau, t_start, t_final, eps, G, "4Body_random");
		std::vector<Body<double>> sol(sys.traj_init), k(n), kk(n), acc(n), tmp(n);
		for (std::size_t i = 0; i < n; ++i) tmp[i].m = sol[i].m;
		auto mult = [n](std::vector<Body<double>>& v, double a) {
			for (std::size_t i = 0; i < n; ++i) {
				v[i].r *= a;
				v[i].v *= a;
			}
		};
		auto add = [n](std::vector<Body<double>>& a, const std::vector<Body<double>>& b) {
			for (std::size_t i = 0; i < n; ++i) {
				a[i].r += b[i].r;
				a[i].v += b[i].v;
			}
		};
		const double cs[3] = {tau / 2, tau / 2, tau};
		const double bs[4] = {1., 2., 2., 1.};
		long steps = 0;
		double t = t_start + tau;
		while (true) {
			sys.N_body_func(sol, k, 0, n);
			acc = k;
			for (int s = 0; s < 3; ++s) {
				kk = k;
				mult(kk, cs[s]);
				for (std::size_t i = 0; i < n; ++i) {
					tmp[i].r = sol[i].r;
					tmp[i].v = sol[i].v;
				}
				add(tmp, kk);
				sys.N_body_func(tmp, k, 0, n);
				kk = k;
				mult(kk, bs[s + 1]);
				add(acc, kk);
			}
			mult(acc, tau / 6);
			add(sol, acc);
			if (rows_out && steps < max_rows) {
				double* row = rows_out + (std::size_t)steps * (1 + 3 * n);
				row[0] = t;
				for (std::size_t i = 0; i < n; ++i)
					for (int c = 0; c < 3; ++c) row[1 + 3 * i + c] = sol[i].r[c];
			}
			++steps;
			if (max_steps > 0 && steps >= max_steps) break;
			if (t > t_final) break;
			t += tau;
		}
		std::memcpy(aos, sol.data(), n * sizeof(Body<double>));
		return steps;
	} catch (...) {
		return -1;
	}
}

/* read_data<double> (read_write_utils.hpp:12-39): returns N, fills up to cap bodies;
 * -1 where the reference throws. */
long ref_read_data(const char* path, double* aos, std::size_t cap) {
	try {
		std::vector<Body<double>> b = read_data<double>(path);
		std::size_t n = b.size() < cap ? b.size() : cap;
		if (aos) std::memcpy(aos, b.data(), n * sizeof(Body<double>));
		return (long)b.size();
	} catch (...) {
		return -1;
	}
}

/* write_bodies_coords (read_write_utils.hpp:42-52) appended to path. */
int ref_write_bodies_coords(const char* path, double time, const double* aos, std::size_t n) {
	std::vector<Body<double>> b(n);
	std::memcpy(b.data(), aos, n * sizeof(Body<double>));
	std::ofstream f(path, std::ios::app);
	if (!f.is_open()) return 1;
	write_bodies_coords(f, time, b);
	return 0;
}

}  // extern "C"
