// lab4_ode_RK2_MPI_b200.hpp -- THE PATCH of INTEGRATION.md section 1, as a compilable file.
//
// A maintainer of the reference replaces the body of
//     template <typename T> void ode_RK2_MPI(ode_system<T>&, std::size_t& iter, T& timer,
//         MPI_Datatype, int* counts, int* offsets, std::size_t id, const std::string path, const T& eps)
// (lab_4_N_body_MPI/src/include/N_body/schemes/ode_RK2_MPI.hpp:19-107) with this one.  Signature,
// loop control (:65,72,98-99: it fixes the step count) and the rank-0 trajectory
// output (:45-53,63,94-96, through the reference's own write_bodies_coords) stay; the two
// sys.N_body_func calls, the mult/add lambdas and the three MPI_Allgatherv calls (:73-92) become
// nbody_step; the MPI rank becomes the GPU rank.  tests/integration/relink_lab4.cpp compiles the
// reference's UNMODIFIED N_body.cpp, class_Solver_N_body.hpp, struct_ode_system.hpp and
// read_write_utils.hpp against it; tests/test_gpu_reference_relink.py runs the result on a B200.
#pragma once
#include <mpi.h>

#include <cstdint>
#include <fstream>
#include <stdexcept>
#include <string>
#include <vector>

#include "nbody_b200.h"	 // include/nbody_b200.h of this repository

template <typename T>
void ode_RK2_MPI(ode_system<T>& sys, std::size_t& iter, T& timer, MPI_Datatype /*type*/,
				 int* /*counts*/, int* /*offsets*/, std::size_t id, const std::string path = "",
				 const T& eps = 1e-6) {
	(void)iter;
	(void)eps;	// unused by the reference too
	int world = 1;
	MPI_Comm_size(MPI_COMM_WORLD, &world);
	char uid[NBODY_NCCL_UID_BYTES] = {};
	if (world > 1) {  // was: MPI_Init's rendezvous
		if (id == 0 && nbody_nccl_unique_id(uid) != NBODY_OK)
			throw std::runtime_error(nbody_last_error(nullptr));
		MPI_Bcast(uid, (int)sizeof uid, MPI_BYTE, 0, MPI_COMM_WORLD);
	}
	nbody_ctx* ctx = nullptr;  // was: counts/offsets + MPI datatypes
	if (nbody_create(&ctx, sys.traj_init.size(), 8 * (int)sizeof(T), /*device*/ (int)id, (int)id, world,
					 world > 1 ? uid : nullptr) != NBODY_OK)
		throw std::runtime_error(nbody_last_error(nullptr));
	auto ck = [&](int rc) {
		if (rc != NBODY_OK) {
			const std::string msg = nbody_last_error(ctx);
			nbody_destroy(ctx);
			throw std::runtime_error(msg);
		}
	};
	ck(nbody_set_params(ctx, 6.67e-11, 1e-6));		   // the lambda's captures G, eps (:53)
	ck(nbody_upload_aos(ctx, sys.traj_init.data()));   // Body<T> is 7 packed T: accepted as is

	std::ofstream fout;
	if (!path.empty() && id == 0) fout = std::ofstream(path);
	if (!path.empty() && id == 0 && !fout.is_open()) {
		std::cout << "Could not open file for writing" << std::endl;
	} else {
		fout << "time;r11;r12;r13;r21;r22;r23;r31;r32;r33;r41;r42;r43\n";
	}
	std::vector<Body<T>> sol(sys.traj_init);
	if (!path.empty() && id == 0) write_bodies_coords(fout, sys.t_start, sol);

	T t = sys.t_start + sys.tau;
	timer = -MPI_Wtime();
	while (true) {
		ck(nbody_step(ctx, NBODY_RK2, (double)sys.tau, 1));	 // was :73-92
		if (!path.empty() && id == 0) {						 // was :94-96
			uint64_t ticket = 0;
			const void* r = nullptr;
			ck(nbody_snapshot_async(ctx, &ticket));
			ck(nbody_snapshot_wait(ctx, ticket, &r));  // n x 3 T in pinned host memory
			for (std::size_t i = 0; i < sol.size(); ++i)
				for (int c = 0; c < 3; ++c) sol[i].r[c] = ((const T*)r)[3 * i + c];
			ck(nbody_snapshot_release(ctx, ticket));
			write_bodies_coords(fout, t, sol);
		}
		if (t > sys.t_final) break;
		t += sys.tau;
	}
	ck(nbody_sync(ctx));
	timer += MPI_Wtime();
	fout.close();
	nbody_destroy(ctx);
}
