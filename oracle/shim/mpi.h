/*
 * mpi.h -- threads-as-ranks MPI shim.  TEST INFRASTRUCTURE ONLY (oracle/).
 *
 * Neither this container nor the GPU box has an MPI.  The
 * reference's lab_4 sources need exactly twelve entry points
 * (lab_4_N_body_MPI/src/N_body.cpp:67,81,86;
 *  include/N_body/class_Solver_N_body.hpp:95-96,141-179,196-197;
 *  include/N_body/schemes/ode_RK2_MPI.hpp:66,76,83,91,102).
 * Each "rank" is a std::thread; rank identity is thread_local; every rank owns
 * its own buffers (as separate address spaces would), and the only data-moving
 * call, the in-place MPI_Allgatherv, copies the other ranks' [displ,count)
 * segments honouring the committed datatype layout.
 */
#pragma once
#include <chrono>
#include <condition_variable>
#include <cstddef>
#include <cstring>
#include <functional>
#include <mutex>
#include <thread>
#include <vector>

typedef int MPI_Comm;
typedef int MPI_Datatype;
typedef std::ptrdiff_t MPI_Aint;

#define MPI_COMM_WORLD 0
#define MPI_DATATYPE_NULL 0
#define MPI_DOUBLE 1
#define MPI_BYTE (-1)
#define MPI_SUCCESS 0
#define MPI_IN_PLACE ((void*)-1)

namespace mpi_shim {

struct TypeLayout {
	std::vector<int> lens;		  // block lengths, in doubles
	std::vector<MPI_Aint> disps;  // block byte displacements
	MPI_Aint lb = 0, extent = 0;
};

struct World {
	int size = 1;
	std::mutex mu;
	std::condition_variable cv;
	int waiting = 0;
	long generation = 0;
	std::vector<void*> bufs;
	std::vector<TypeLayout> types;	// index = datatype id - 2
	World() : bufs(1, nullptr) {}
};

inline World& world() {
	static World w;
	return w;
}
inline int& my_rank() {
	static thread_local int r = 0;
	return r;
}

inline void barrier() {
	World& w = world();
	if (w.size == 1) return;
	std::unique_lock<std::mutex> lk(w.mu);
	long gen = w.generation;
	if (++w.waiting == w.size) {
		w.waiting = 0;
		++w.generation;
		w.cv.notify_all();
	} else {
		w.cv.wait(lk, [&] { return w.generation != gen; });
	}
}

/* Run fn(rank) on n_ranks threads ("mpirun -np n_ranks"). */
inline void run(int n_ranks, const std::function<void(int)>& fn) {
	World& w = world();
	w.size = n_ranks;
	w.bufs.assign(n_ranks, nullptr);
	w.waiting = 0;
	std::vector<std::thread> th;
	for (int r = 1; r < n_ranks; ++r)
		th.emplace_back([r, &fn] {
			my_rank() = r;
			fn(r);
		});
	my_rank() = 0;
	fn(0);
	for (auto& t : th) t.join();
	w.size = 1;
	w.bufs.assign(1, nullptr);
}

}  // namespace mpi_shim

inline int MPI_Init(int*, char***) { return MPI_SUCCESS; }
inline int MPI_Finalize() { return MPI_SUCCESS; }
inline int MPI_Barrier(MPI_Comm) {
	mpi_shim::barrier();
	return MPI_SUCCESS;
}
inline int MPI_Comm_rank(MPI_Comm, int* r) {
	*r = mpi_shim::my_rank();
	return MPI_SUCCESS;
}
inline int MPI_Comm_size(MPI_Comm, int* s) {
	*s = mpi_shim::world().size;
	return MPI_SUCCESS;
}
inline double MPI_Wtime() {
	using clk = std::chrono::steady_clock;
	return std::chrono::duration<double>(clk::now().time_since_epoch()).count();
}

inline int MPI_Type_create_struct(int count, const int* lens, const MPI_Aint* disps,
								  const MPI_Datatype* /*all MPI_DOUBLE*/,
								  MPI_Datatype* newtype) {
	mpi_shim::World& w = mpi_shim::world();
	std::lock_guard<std::mutex> lk(w.mu);
	mpi_shim::TypeLayout t;
	MPI_Aint lo = disps[0], hi = disps[0];
	for (int i = 0; i < count; ++i) {
		t.lens.push_back(lens[i]);
		t.disps.push_back(disps[i]);
		if (disps[i] < lo) lo = disps[i];
		MPI_Aint e = disps[i] + (MPI_Aint)lens[i] * (MPI_Aint)sizeof(double);
		if (e > hi) hi = e;
	}
	t.lb = lo;
	t.extent = hi - lo;
	w.types.push_back(t);
	*newtype = (int)w.types.size() + 1;
	return MPI_SUCCESS;
}
inline int MPI_Type_create_resized(MPI_Datatype old, MPI_Aint lb, MPI_Aint extent,
								   MPI_Datatype* newtype) {
	mpi_shim::World& w = mpi_shim::world();
	std::lock_guard<std::mutex> lk(w.mu);
	mpi_shim::TypeLayout t = w.types[old - 2];
	t.lb = lb;
	t.extent = extent;
	w.types.push_back(t);
	*newtype = (int)w.types.size() + 1;
	return MPI_SUCCESS;
}
inline int MPI_Type_get_extent(MPI_Datatype t, MPI_Aint* lb, MPI_Aint* extent) {
	mpi_shim::World& w = mpi_shim::world();
	std::lock_guard<std::mutex> lk(w.mu);
	*lb = w.types[t - 2].lb;
	*extent = w.types[t - 2].extent;
	return MPI_SUCCESS;
}
inline int MPI_Type_commit(MPI_Datatype*) { return MPI_SUCCESS; }
inline int MPI_Type_free(MPI_Datatype*) { return MPI_SUCCESS; }

/* Not used by the reference's current sources (only by its stale binary); used by
 * the INTEGRATION.md patch to hand the NCCL unique id to every rank.  MPI_BYTE only. */
inline int MPI_Bcast(void* buf, int count, MPI_Datatype /*MPI_BYTE*/, int root, MPI_Comm) {
	mpi_shim::World& w = mpi_shim::world();
	if (w.size == 1) return MPI_SUCCESS;
	const int me = mpi_shim::my_rank();
	{
		std::lock_guard<std::mutex> lk(w.mu);
		w.bufs[me] = buf;
	}
	mpi_shim::barrier();
	if (me != root) std::memcpy(buf, w.bufs[root], (std::size_t)count);
	mpi_shim::barrier();
	return MPI_SUCCESS;
}

/* Only the form the reference uses: sendbuf == MPI_IN_PLACE. */
inline int MPI_Allgatherv(const void* /*MPI_IN_PLACE*/, int, MPI_Datatype, void* recvbuf,
						  const int* counts, const int* displs, MPI_Datatype type,
						  MPI_Comm) {
	mpi_shim::World& w = mpi_shim::world();
	if (w.size == 1) return MPI_SUCCESS;
	const int me = mpi_shim::my_rank();
	mpi_shim::TypeLayout t;
	{
		std::lock_guard<std::mutex> lk(w.mu);
		t = w.types[type - 2];
		w.bufs[me] = recvbuf;
	}
	mpi_shim::barrier();
	for (int r = 0; r < w.size; ++r) {
		if (r == me) continue;
		const char* src = (const char*)w.bufs[r];
		char* dst = (char*)recvbuf;
		for (int e = displs[r]; e < displs[r] + counts[r]; ++e)
			for (std::size_t b = 0; b < t.lens.size(); ++b)
				std::memcpy(dst + (MPI_Aint)e * t.extent + t.disps[b],
							src + (MPI_Aint)e * t.extent + t.disps[b],
							(std::size_t)t.lens[b] * sizeof(double));
	}
	mpi_shim::barrier();
	return MPI_SUCCESS;
}
