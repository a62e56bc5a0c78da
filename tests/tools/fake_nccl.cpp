// fake_nccl.cpp -- TEST INFRASTRUCTURE: the eight NCCL entry points nbody_capi.cu binds with dlsym,
// implemented for ranks that are THREADS of one process sharing host memory (the emulated
// library's "device" memory).  Loaded through NBODY_NCCL_LIB by tests/test_capi_emulation.py so
// that the multi-rank host logic (exchange sequencing, in-place all-gather layout, reduce-scatter
// of the pair-symmetric kernel's j-side sums, all-reduce of the diagnostics) runs in the CPU
// suite.  Collectives block until every rank has called them, in call order; the stream argument
// is ignored because the emulated streams execute at issue.  Never used by the product.
#include <condition_variable>
#include <cstdint>
#include <cstring>
#include <map>
#include <mutex>
#include <vector>

namespace {
struct Group {
	int nranks = 0, joined = 0;
	std::mutex mu;
	std::condition_variable cv;
	int arrived = 0;
	unsigned long long generation = 0;
	std::vector<const void*> send;
	std::vector<void*> recv;
	void barrier() {
		std::unique_lock<std::mutex> lk(mu);
		const unsigned long long gen = generation;
		if (++arrived == nranks) {
			arrived = 0;
			++generation;
			cv.notify_all();
		} else {
			cv.wait(lk, [&] { return generation != gen; });
		}
	}
};
struct Comm { Group* g; int rank; };
std::mutex g_mu;
std::map<uint64_t, Group*> g_groups;
uint64_t g_next_id = 1;

size_t elem_size(int dt) { return (dt == 8 || dt == 5) ? 8 : dt == 7 ? 4 : 1; }  // ncclUint64 = 5

template <typename T>
void sum_into(T* out, const std::vector<const void*>& send, size_t offset, size_t count, int op = 0) {
	for (size_t k = 0; k < count; ++k) {
		T s = static_cast<const T*>(send[0])[offset + k];
		for (size_t r = 1; r < send.size(); ++r) {	// rank order: deterministic
			const T v = static_cast<const T*>(send[r])[offset + k];
			if (op == 0) s += v;
			else if (op == 3) { if (v < s || v != v) s = v; }  // ncclMin
			else if (v > s || v != v) s = v;  // ncclMax (NaN wins, as it must stay visible)
		}
		out[k] = s;
	}
}
}  // namespace

extern "C" {
typedef struct { char internal[128]; } ncclUniqueId;
typedef Comm* ncclComm_t;

int ncclGetUniqueId(ncclUniqueId* id) {
	std::lock_guard<std::mutex> lk(g_mu);
	memset(id, 0, sizeof(*id));
	const uint64_t v = g_next_id++;
	memcpy(id->internal, &v, 8);
	return 0;
}

int ncclCommInitRank(ncclComm_t* comm, int nranks, ncclUniqueId id, int rank) {
	uint64_t v;
	memcpy(&v, id.internal, 8);
	Group* g;
	{
		std::lock_guard<std::mutex> lk(g_mu);
		Group*& slot = g_groups[v];
		if (!slot) {
			slot = new Group();
			slot->nranks = nranks;
			slot->send.assign(nranks, nullptr);
			slot->recv.assign(nranks, nullptr);
		}
		g = slot;
		if (g->nranks != nranks || rank < 0 || rank >= nranks) return 4;  // ncclInvalidArgument
	}
	*comm = new Comm{g, rank};
	g->barrier();
	return 0;
}

int ncclCommDestroy(ncclComm_t comm) {
	delete comm;  // groups are leaked on purpose: a few hundred bytes per test context
	return 0;
}

int ncclGroupStart() { return 0; }
int ncclGroupEnd() { return 0; }
const char* ncclGetErrorString(int) { return "fake NCCL error"; }

int ncclAllGather(const void* send, void* recv, size_t count, int dt, ncclComm_t c, void*) {
	Group* g = c->g;
	const size_t bytes = count * elem_size(dt);
	g->send[c->rank] = send;
	g->barrier();
	for (int r = 0; r < g->nranks; ++r) {
		char* dst = static_cast<char*>(recv) + (size_t)r * bytes;
		if (dst != g->send[r]) memcpy(dst, g->send[r], bytes);
	}
	g->barrier();
	return 0;
}

int ncclAllReduce(const void* send, void* recv, size_t count, int dt, int op, ncclComm_t c, void*) {
	if ((op != 0 && op != 2 && op != 3) || (dt != 8 && dt != 7)) return 4;	 // ncclSum, ncclMax, ncclMin
	Group* g = c->g;
	g->send[c->rank] = send;
	g->barrier();
	std::vector<char> tmp(count * elem_size(dt));
	if (dt == 8) sum_into(reinterpret_cast<double*>(tmp.data()), g->send, 0, count, op);
	else sum_into(reinterpret_cast<float*>(tmp.data()), g->send, 0, count, op);
	g->barrier();  // everyone has read every send buffer: in-place results may be written
	memcpy(recv, tmp.data(), tmp.size());
	g->barrier();
	return 0;
}

int ncclReduceScatter(const void* send, void* recv, size_t recvcount, int dt, int op, ncclComm_t c, void*) {
	if (op != 0 || (dt != 8 && dt != 7 && dt != 5)) return 4;
	Group* g = c->g;
	g->send[c->rank] = send;
	g->barrier();
	std::vector<char> tmp(recvcount * elem_size(dt));
	const size_t off = (size_t)c->rank * recvcount;
	if (dt == 8) sum_into(reinterpret_cast<double*>(tmp.data()), g->send, off, recvcount);
	else if (dt == 5) sum_into(reinterpret_cast<uint64_t*>(tmp.data()), g->send, off, recvcount);
	else sum_into(reinterpret_cast<float*>(tmp.data()), g->send, off, recvcount);
	g->barrier();
	memcpy(recv, tmp.data(), tmp.size());
	g->barrier();
	return 0;
}
}  // extern "C"
