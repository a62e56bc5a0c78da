// fx_sums_host.cpp -- TEST INFRASTRUCTURE: the fixed-point row-sum arithmetic of option
// "deterministic" (parallel_comp_labs_b200/csrc/fixed_sums.cuh, the SOURCE the GPU runs) compiled
// for the host through tests/tools/cuda_emu.h, so that the CPU suite can compare it with exact
// rational arithmetic.  Never linked into the product.
#include "cuda_emu.h"

#include "../../parallel_comp_labs_b200/csrc/fixed_sums.cuh"

extern "C" {
int fx_limbs(void) { return nbody::FX_LIMBS; }

// adds x[0..n) into the (zeroed by the caller) limbs[FX_LIMBS] / poison word, in the order given
void fx_accumulate(const double* x, size_t n, int e0, unsigned long long* limbs,
				   unsigned long long* poison) {
	for (size_t i = 0; i < n; ++i) nbody::fx_add(limbs, 1, poison, x[i], e0);
}

double fx_value(const unsigned long long* limbs, int e0) { return nbody::fx_read(limbs, 1, e0); }
}
