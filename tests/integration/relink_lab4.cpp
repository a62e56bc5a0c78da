// relink_lab4.cpp -- the reference's OWN lab_4 program relinked against libnbody_b200.so.
// TEST INFRASTRUCTURE (built by oracle/Makefile into oracle/_ref/relink_lab4 when /root/reference
// is present; the binary travels to the GPU box, the sources are never copied).
//
// From the reference, unmodified, via #include where they lie:
//   N_body.cpp (main :38-87), class_Solver_N_body.hpp (solve_parallel_MPI :87-208),
//   struct_ode_system.hpp (ode_system_default_test / ode_system_random_data), struct_Body.hpp,
//   utils/read_write_utils.hpp (read_data, write_bodies_coords).
// Replaced: schemes/ode_RK2_MPI.hpp's function body -> tests/integration/lab4_ode_RK2_MPI_b200.hpp
// (the patch INTEGRATION.md shows).  Mechanism: the reference's header is included FIRST with its
// function renamed by the preprocessor, so it still compiles but is never called; being
// `#pragma once` it is skipped when class_Solver_N_body.hpp includes it again, and the call sites
// there (:188,192) bind to the patched template.
// Supplied: mpi.h (oracle/shim, threads-as-ranks; ranks = GPUs), config.hpp.
//
// usage (cwd = <scratch>/src, with ./include/N_body/N_body_config/4body.txt and ../output/):
//   relink_lab4 default [n_ranks]   what N_body.cpp:69-79 (commented) does: tau = 0.1, 0.05, 0.025
//   relink_lab4 main [n_ranks]      the reference's main as it is (random bodies, "Server")
#include <mpi.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <string>

#define ode_RK2_MPI ode_RK2_MPI_reference_cpu_body
#include "include/N_body/schemes/ode_RK2_MPI.hpp"
#undef ode_RK2_MPI
#include "lab4_ode_RK2_MPI_b200.hpp"

#define main reference_lab4_main
#include "N_body.cpp"
#undef main

int main(int argc, char** argv) {
	const std::string mode = argc > 1 ? argv[1] : "default";
	const int n_ranks = argc > 2 ? std::atoi(argv[2]) : 1;
	if (const char* d = std::getenv("NBODY_REF_SCRATCH")) {
		config_lab_4_MPI::PATH_main_folder = d;
		config_lab_4_MPI::PATH_src_folder = std::string(d) + "/src";
		config_lab_4_MPI::PATH_output_folder = std::string(d) + "/output";
	}
	int rc = 0;
	try {
		mpi_shim::run(n_ranks, [&](int) {
			if (mode == "main") {
				reference_lab4_main(argc, argv);
				return;
			}
			for (double tau : {0.1, 0.05, 0.025}) {	 // N_body.cpp:69-79
				ode_system<double> sys;
				Solver_N_body<double> solver;
				solver.notifications = false;
				MPI_Barrier(MPI_COMM_WORLD);
				sys.ode_system_default_test(tau);
				solver.solve_parallel_MPI(sys);
			}
		});
	} catch (const std::exception& e) {
		std::fprintf(stderr, "relink_lab4: %s\n", e.what());
		rc = 2;
	} catch (int e) {
		std::fprintf(stderr, "relink_lab4: reference threw int %d\n", e);
		rc = 3;
	} catch (const char* e) {
		std::fprintf(stderr, "relink_lab4: reference threw \"%s\"\n", e);
		rc = 4;
	}
	return rc;
}
