// emu_selftest.cpp -- TEST INFRASTRUCTURE: does the kernel emulation (cuda_emu.h) actually catch
// the mistakes it claims to catch?  Four tiny kernels built on the product's own ring helpers, each
// with one deliberate defect, plus a correct twin; selftest(case) returns the error text ("" = ran
// clean).  tests/test_kernel_emulation.py asserts on the texts.
#include "cuda_emu.h"

#include "../../parallel_comp_labs_b200/csrc/force_kernels.cuh"

using namespace nbody;

namespace {
// two-stage ring over `ntiles` tiles of 32 doubles; every thread sums its element of every tile
template <int DEFECT>
void ring_kernel(const double* src, double* out, int ntiles) {
	NB_DYN_SMEM(smem);
	double* tiles = reinterpret_cast<double*>(smem);			   // [2][32]
	uint64_t* full = reinterpret_cast<uint64_t*>(smem + 2 * 256);  // [2]
	const int tid = threadIdx.x;
	if (tid == 0) {
		mbar_init(&full[0], 1);
		mbar_init(&full[1], 1);
		mbar_fence_init();
	}
	__syncthreads();
	if (tid == 0)
		for (int k = 0; k < 2 && k < ntiles; ++k) {
			if (DEFECT == 3 && k == 1) break;  // defect 3: second stage never armed
			mbar_expect_tx(&full[k], 256);
			bulk_g2s(tiles + 32 * k, src + 32 * k, 256, &full[k]);
		}
	double acc = 0;
	for (int t = 0; t < ntiles; ++t) {
		const int st = t & 1;
		mbar_wait(&full[st], (t >> 1) & 1);
		acc += tiles[32 * st + (tid & 31)];
		if (DEFECT != 1) __syncthreads();  // defect 1: stage refilled while others may still read it
		if (tid == 0 && t + 2 < ntiles) {
			mbar_expect_tx(&full[st], 256);
			bulk_g2s(tiles + 32 * st, src + 32 * (t + 2), 256, &full[st]);
		}
		if (DEFECT == 1) __syncthreads();
	}
	if (DEFECT == 2 && tid == 5) {	// defect 2: one lane skips a warp collective
		out[tid] = acc;
		return;
	}
	acc += __shfl_sync(0xffffffffu, acc, (tid + 1) & 31);
	out[tid] = acc;
}
}  // namespace

extern "C" const char* selftest(int defect) {
	static thread_local std::string msg;
	msg.clear();
	std::vector<double> src(32 * 6), out(64, 0.0);
	for (size_t i = 0; i < src.size(); ++i) src[i] = (double)i;
	const double* s = src.data();
	double* o = out.data();
	try {
		emu::launch(dim3(1), dim3(64), 2 * 256 + 16, [&] {
			switch (defect) {
				case 0: ring_kernel<0>(s, o, 6); break;
				case 1: ring_kernel<1>(s, o, 6); break;
				case 2: ring_kernel<2>(s, o, 6); break;
				default: ring_kernel<3>(s, o, 6); break;
			}
		});
		if (defect == 0) {	// the correct twin must also compute the right thing
			for (int t = 0; t < 64; ++t) {
				double a = 0, b = 0;
				for (int k = 0; k < 6; ++k) { a += src[32 * k + (t & 31)]; b += src[32 * k + ((t + 1) & 31)]; }
				if (out[t] != a + b) msg = "wrong result in the correct kernel";
			}
		}
	} catch (const std::exception& e) {
		msg = e.what();
	}
	return msg.c_str();
}
