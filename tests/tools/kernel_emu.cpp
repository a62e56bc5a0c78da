// kernel_emu.cpp -- TEST INFRASTRUCTURE: runs the product's kernel SOURCE (pack, K1 / K1s, K2)
// on the CPU through tests/tools/cuda_emu.h and returns the accelerations, so that the
// `-m "not gpu"` suite executes the code the GPU will run (index arithmetic, shared-memory
// layout, ring protocol, clamp redo, rotation, slice reduction).  The launch parameters are
// filled the way nbody_capi.cu fills them for a single-GPU context.  Never linked into the
// product.
#include "cuda_emu.h"

#include "../../parallel_comp_labs_b200/csrc/force_kernels.cuh"
#include "../../parallel_comp_labs_b200/csrc/force_sym_kernels.cuh"
#include "../../parallel_comp_labs_b200/csrc/stage_kernels.cuh"

using namespace nbody;

namespace {
constexpr unsigned TJ = 256, EXTRA_TILES = 4;

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

template <typename T, int I, int UM>
void run_default(const ForceParams& p, double eps2, unsigned grid) {
	const Clamp<T> cl = make_clamp<T>(eps2);
	emu::launch(dim3(grid), dim3(128), force_smem_bytes<T, 3>(),
				[&] { force_kernel<T, 128, I, 3, 1, UM, 8, 2>(p, cl); });
}
template <typename T, int I, int UM>
void run_sym(const SymParams& p, double eps2, unsigned grid) {
	const Clamp<T> cl = make_clamp<T>(eps2);
	emu::launch(dim3(grid), dim3(128), force_sym_smem_bytes<T, 128, I, 2>(),
				[&] { force_sym_kernel<T, 128, I, 2, 1, UM, 2>(p, cl); });
}

template <typename T>
int rhs_impl(unsigned n, const double* bodies, double G, double eps, int sym, int ipt,
			 unsigned n_slices, int unimass, unsigned grid, double* acc_out) {
	typedef typename V4<T>::type vec4;
	const unsigned n_pad = (n + TJ - 1) / TJ * TJ, own_pad = n_pad;
	const unsigned jt_total = n_pad / TJ;
	std::vector<T> aos((size_t)n * 7);
	for (size_t k = 0; k < aos.size(); ++k) aos[k] = (T)bodies[k];
	bool um = unimass != 0;
	for (unsigned i = 1; um && i < n; ++i) um = aos[7 * (size_t)i] == aos[0];
	const T mass = um ? aos[0] : T(0);
	const T pad_pos = um ? (sizeof(T) == 8 ? (T)1e150 : (T)1e18) : T(0);

	// device buffers; rows past n_pad are mapped but hold rubbish, as on the GPU
	std::vector<vec4> Y((size_t)n_pad + EXTRA_TILES * TJ), P(Y.size()), vel(own_pad);
	memset((void*)Y.data(), 0xFF, Y.size() * sizeof(vec4));	 // NaN pattern
	{
		const T* a = aos.data();
		void *y = Y.data(), *pp = P.data(), *v = vel.data();
		emu::launch(dim3((n_pad + 255) / 256), dim3(256), 0, [&] {
			pack_aos_kernel<T>(a, y, pp, v, n, n_pad, 0u, own_pad, pad_pos, um ? mass : T(0));
		});
	}
	unsigned counter = 0;
	std::vector<vec4> apart;
	std::vector<T> aj;
	unsigned apart_stride = own_pad, aj_stride = 0, slices = n_slices;
	if (!sym) {
		const unsigned itile = 128 * ipt;
		ForceParams p;
		memset(&p, 0, sizeof(p));
		p.blk[0] = Y.data();
		p.own = Y.data();
		p.contiguous = 1;
		p.jt_per_rank = jt_total;
		p.rank = 0;
		p.own0 = 0;
		p.own_pad = own_pad;
		p.n_itiles = (n + itile - 1) / itile;
		p.jt_begin = 0;
		p.jt_count = jt_total;
		p.jt_total = jt_total;
		slices = std::min(n_slices, std::max(1u, jt_total));
		p.n_slices = slices;
		p.slice_base = 0;
		p.n_real = n;
		p.counter = &counter;
		apart.assign((size_t)slices * own_pad, vec4{(T)NAN, (T)NAN, (T)NAN, (T)NAN});
		p.apart = apart.data();
		const unsigned units = p.n_itiles * slices;
		const unsigned g = std::min(units, grid);
		if (ipt == 2) um ? run_default<T, 2, 1>(p, eps * eps, g) : run_default<T, 2, 0>(p, eps * eps, g);
		else if (ipt == 4) um ? run_default<T, 4, 1>(p, eps * eps, g) : run_default<T, 4, 0>(p, eps * eps, g);
		else return -10;
	} else {
		const unsigned block = 128 * ipt;
		if (block % TJ) return -11;
		const unsigned tpb = block / TJ;
		if (tpb - 1 > EXTRA_TILES) return -12;
		SymParams p;
		memset(&p, 0, sizeof(p));
		p.pos = Y.data();
		p.nb = (jt_total + tpb - 1) / tpb;
		p.rows_pad = p.nb * block;
		p.block0 = 0;
		p.nb_own = p.nb;
		p.jt_total = jt_total;
		p.n_pad = n_pad;
		p.n_real = n;
		p.n_slices = slices;
		p.counter = &counter;
		p.pad_pos = (double)pad_pos;
		apart_stride = p.rows_pad;
		aj_stride = p.rows_pad;
		aj.assign((size_t)3 * p.rows_pad, T(0));  // cudaMemsetAsync before the launch
		p.aj = aj.data();
		const unsigned g = std::min(p.nb * slices, grid);
		switch (ipt * 2 + (um ? 1 : 0)) {
			case 4: run_sym<T, 2, 0>(p, eps * eps, g); break;
			case 5: run_sym<T, 2, 1>(p, eps * eps, g); break;
			case 8: run_sym<T, 4, 0>(p, eps * eps, g); break;
			case 9: run_sym<T, 4, 1>(p, eps * eps, g); break;
			case 12: run_sym<T, 6, 0>(p, eps * eps, g); break;
			case 13: run_sym<T, 6, 1>(p, eps * eps, g); break;
			case 16: run_sym<T, 8, 0>(p, eps * eps, g); break;
			case 17: run_sym<T, 8, 1>(p, eps * eps, g); break;
			default: return -13;
		}
	}
	// K2 in its nbody_rhs form
	std::vector<T> rhs((size_t)3 * own_pad, (T)NAN);
	StageParams<T> sp;
	memset((void*)&sp, 0, sizeof(sp));
	sp.apart = apart.data();
	sp.Y = Y.data();
	sp.P = P.data();
	sp.rhs_out = rhs.data();
	sp.own0 = 0;
	sp.own_pad = own_pad;
	sp.own_real = n;
	sp.n_slices = sym ? 0 : slices;	 // the pair-symmetric kernel adds both sides into aj
	sp.apart_stride = apart_stride;
	sp.aj = sym ? aj.data() : nullptr;
	sp.aj_stride = aj_stride;
	sp.mass_scale = um ? mass : T(1);
	sp.minusG = (T)(-(T)G);
	sp.first = 1;
	emu::launch(dim3((n + 255) / 256), dim3(256), 0, [&] { stage_kernel<T>(sp); });
	for (size_t k = 0; k < (size_t)3 * n; ++k) acc_out[k] = (double)rhs[k];
	return 0;
}

// The pair-symmetric kernel on `world` ranks (FP64, I = 4): every rank packs the whole array,
// walks its own blocks of the global schedule and collects reactions for all bodies; the j-side
// arrays are then summed over the ranks (what ncclReduceScatter does) and every rank's K2 reduces
// its owned rows.  Parameters as launch_force_sym_all / make_plan fill them for world > 1.
int rhs_sym_ranks(unsigned n, const double* bodies, double G, double eps, unsigned world,
				  unsigned n_slices, int unimass, double* acc_out) {
	typedef double T;
	typedef double4 vec4;
	constexpr unsigned BLOCK = 512, TPB = 2;
	const unsigned per = (n + world - 1) / world;
	const unsigned own_pad = (per + TJ - 1) / TJ * TJ, n_pad = own_pad * world;
	if (own_pad % BLOCK) return -20;
	const unsigned jt_total = n_pad / TJ, nb = jt_total / TPB, nb_own = own_pad / BLOCK;
	bool um = unimass != 0;
	for (unsigned i = 1; um && i < n; ++i) um = bodies[7 * (size_t)i] == bodies[0];
	const T mass = um ? bodies[0] : 0., pad_pos = um ? 1e150 : 0.;
	std::vector<T> aj_sum((size_t)3 * n_pad, 0.);
	std::vector<std::vector<vec4>> apart(world);
	std::vector<vec4> Y((size_t)n_pad + EXTRA_TILES * TJ), P(Y.size());
	for (unsigned r = 0; r < world; ++r) {
		std::vector<vec4> vel(own_pad);
		void *y = Y.data(), *pp = P.data(), *v = vel.data();
		emu::launch(dim3((n_pad + 255) / 256), dim3(256), 0, [&] {
			pack_aos_kernel<T>(bodies, y, pp, v, n, n_pad, r * own_pad, own_pad, pad_pos, um ? mass : 0.);
		});
		unsigned counter = 0;
		std::vector<T> aj((size_t)3 * n_pad, 0.);
		SymParams p;
		memset(&p, 0, sizeof(p));
		p.pos = Y.data();
		p.aj = aj.data();
		p.rows_pad = nb * BLOCK;
		p.nb = nb;
		p.block0 = r * nb_own;
		p.nb_own = nb_own;
		p.jt_total = jt_total;
		p.n_pad = n_pad;
		p.n_real = n;
		p.n_slices = n_slices;
		p.counter = &counter;
		p.pad_pos = pad_pos;
		const unsigned g = std::min(nb_own * n_slices, 2u);
		um ? run_sym<T, 4, 1>(p, eps * eps, g) : run_sym<T, 4, 0>(p, eps * eps, g);
		for (size_t k = 0; k < aj.size(); ++k) aj_sum[k] += aj[k];
	}
	for (unsigned r = 0; r < world; ++r) {
		const unsigned own0 = r * own_pad;
		const unsigned own_real = own0 >= n ? 0 : std::min(own_pad, n - own0);
		if (!own_real) continue;
		std::vector<T> rhs((size_t)3 * own_pad, NAN);
		StageParams<T> sp;
		memset((void*)&sp, 0, sizeof(sp));
		sp.apart = apart[r].data();
		sp.Y = Y.data();
		sp.P = P.data();
		sp.rhs_out = rhs.data();
		sp.own0 = own0;
		sp.own_pad = own_pad;
		sp.own_real = own_real;
		sp.n_slices = 0;
		sp.apart_stride = nb_own * BLOCK;
		sp.aj = aj_sum.data();
		sp.aj_stride = nb * BLOCK;
		sp.mass_scale = um ? mass : 1.;
		sp.minusG = -G;
		sp.first = 1;
		emu::launch(dim3((own_real + 255) / 256), dim3(256), 0, [&] { stage_kernel<T>(sp); });
		for (size_t k = 0; k < (size_t)3 * own_real; ++k) acc_out[3 * (size_t)own0 + k] = rhs[k];
	}
	return 0;
}
}  // namespace

extern "C" int emu_rhs_sym_ranks(unsigned n, const double* bodies, double G, double eps,
								 unsigned world, unsigned n_slices, int unimass, double* acc_out,
								 char* err, int err_len) {
	try {
		return rhs_sym_ranks(n, bodies, G, eps, world, n_slices, unimass, acc_out);
	} catch (const std::exception& e) {
		if (err && err_len > 0) snprintf(err, (size_t)err_len, "%s", e.what());
		return -1;
	}
}

// bodies: n x {m,x,y,z,vx,vy,vz} doubles; acc_out: n x 3 doubles.  precision 64 | 32.
// sym = 0: force_kernel<T,128,ipt,...> (ipt 2 | 4); sym = 1: force_sym_kernel (ipt 2|4|6|8).
// grid = CTAs launched (the first one drains the ticket pool; the others must cope with that).
extern "C" int emu_rhs(unsigned n, const double* bodies, int precision, double G, double eps,
					   int sym, int ipt, unsigned n_slices, int unimass, unsigned grid,
					   double* acc_out, char* err, int err_len) {
	try {
		return precision == 64
				   ? rhs_impl<double>(n, bodies, G, eps, sym, ipt, n_slices, unimass, grid, acc_out)
				   : rhs_impl<float>(n, bodies, G, eps, sym, ipt, n_slices, unimass, grid, acc_out);
	} catch (const std::exception& e) {
		if (err && err_len > 0) snprintf(err, (size_t)err_len, "%s", e.what());
		return -1;
	}
}
