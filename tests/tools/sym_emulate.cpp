// sym_emulate.cpp -- TEST INFRASTRUCTURE: a lane-by-lane CPU walk of the pair-symmetric force
// kernel's decomposition (parallel_comp_labs_b200/csrc/force_sym_kernels.cuh), driven by the
// SAME schedule header the kernel includes (sym_schedule.h).  It proves on the CPU, before any
// GPU time is spent, that
//   * every ordered pair (i, j) of existing bodies is credited exactly once (coverage matrix),
//   * the rotating j-bodies and their accumulators are back in their home lane after 32 steps,
//   * the sums the kernel would hand to the stage kernel (slice rows + j-side array) equal the
//     reference operator (compared with the oracle by tests/test_host_logic.py).
// The arithmetic here is plain double (1/(x*sqrt(x))), not the kernel's seed+correction chain:
// this file checks the bookkeeping, the GPU parity tests check the bits.
// Never linked into the product.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>

#include "../../parallel_comp_labs_b200/csrc/sym_schedule.h"

using namespace nbody;

namespace {
constexpr unsigned TJ = 256, THREADS = 128;
struct B4 { double x, y, z, m; };
}  // namespace

// bodies: n x {m, x, y, z, vx, vy, vz}.  acc_out: n x 3.  cover: n x n uint8 (may be null).
// uniform_mass != 0 emulates the UM kernels (mass factored out, pads parked far away).
// Returns 0, or a negative code for a bookkeeping violation.
extern "C" int sym_emulate_rhs(unsigned n, const double* bodies, double G, double eps, unsigned ipt,
							   unsigned n_slices, int uniform_mass, double* acc_out,
							   unsigned char* cover) {
	const unsigned block = THREADS * ipt;
	if (block % TJ) return -1;
	const unsigned tpb = block / TJ;
	const unsigned n_pad = (n + TJ - 1) / TJ * TJ;	// world == 1: own_pad == n_pad
	const unsigned jt_total = n_pad / TJ;
	const unsigned nb = (jt_total + tpb - 1) / tpb;
	const unsigned rows_pad = nb * block;
	const double pad_pos = uniform_mass ? 1e150 : 0.;
	const double um = uniform_mass ? bodies[0] : 1.0;
	const double eps2 = eps * eps;

	// pos as the pack kernel lays it out, plus phantom rows (poisoned: must never be used)
	std::vector<B4> pos((size_t)(jt_total + tpb) * TJ);
	for (size_t i = 0; i < pos.size(); ++i) {
		if (i < n) pos[i] = {bodies[7 * i + 1], bodies[7 * i + 2], bodies[7 * i + 3], bodies[7 * i]};
		else if (i < n_pad) pos[i] = {pad_pos, pad_pos, pad_pos, uniform_mass ? um : 0.};
		else pos[i] = {NAN, NAN, NAN, NAN};
	}
	std::vector<double> apart((size_t)n_slices * rows_pad * 3, 0.), aj((size_t)3 * rows_pad, 0.);
	if (cover) memset(cover, 0, (size_t)n * n);
	auto credit = [&](unsigned i, unsigned j) {
		if (cover && i < n && j < n) cover[(size_t)i * n + j]++;
	};
	auto inv_cube = [&](double r2) {
		const double x = std::max(r2, eps2);
		return 1.0 / (x * std::sqrt(x));
	};

	for (unsigned a = 0; a < nb; ++a)
		for (unsigned slice = 0; slice < n_slices; ++slice) {
			unsigned q0, q1, t0, t1;
			const unsigned spt = TJ / 32;
			sym_slice_range(sym_ntiles(nb, tpb, a), spt, n_slices, slice, &q0, &q1);
			sym_slice_tiles(spt, q0, q1, &t0, &t1);
			// per-thread state of the CTA
			std::vector<double> xi(block), yi(block), zi(block), mi(block), accm(3 * (size_t)block, 0.);
			std::vector<unsigned> row(block);
			for (unsigned tid = 0; tid < THREADS; ++tid)
				for (unsigned k = 0; k < ipt; ++k) {
					const unsigned r = a * block + k * THREADS + tid, s = k * THREADS + tid;
					row[s] = r;
					if (r < n_pad) { xi[s] = pos[r].x; yi[s] = pos[r].y; zi[s] = pos[r].z; mi[s] = pos[r].m; }
					else { xi[s] = yi[s] = zi[s] = pad_pos; mi[s] = 0.; }
				}
			for (unsigned t = t0; t < t1; ++t) {
				bool diag;
				const unsigned jt = sym_tile(nb, tpb, a, t, &diag);
				if (jt >= jt_total + tpb) return -2;  // outside the allocation
				const unsigned j_first = jt * TJ;
				unsigned cnt = j_first >= n ? 0 : std::min(TJ, n - j_first);
				unsigned lo, hi;  // the bodies of this tile that belong to the unit's strips
				sym_tile_part(t, spt, q0, q1, &lo, &hi);
				cnt = std::min(cnt, hi);
				const B4* tile = pos.data() + (size_t)jt * TJ;
				if (diag) {
					for (unsigned s = 0; s < block; ++s)
						for (unsigned jj = lo; jj < cnt; ++jj) {
							const double dx = xi[s] - tile[jj].x, dy = yi[s] - tile[jj].y, dz = zi[s] - tile[jj].z;
							const double w = inv_cube(dx * dx + dy * dy + dz * dz) * (uniform_mass ? 1.0 : tile[jj].m);
							accm[0 * block + s] += w * dx;
							accm[1 * block + s] += w * dy;
							accm[2 * block + s] += w * dz;
							credit(row[s], j_first + jj);
						}
					continue;
				}
				for (unsigned s0 = lo; s0 < cnt; s0 += 32)
					for (unsigned warp = 0; warp < THREADS / 32; ++warp) {
						// lane state: the travelling j-body, its accumulator and (for the check) its index
						double jx[32], jy[32], jz[32], jm[32], ax[32], ay[32], az[32];
						unsigned jidx[32];
						for (unsigned l = 0; l < 32; ++l) {
							const B4& b = tile[s0 + l];
							jx[l] = b.x; jy[l] = b.y; jz[l] = b.z; jm[l] = b.m;
							ax[l] = ay[l] = az[l] = 0.;
							jidx[l] = j_first + s0 + l;
						}
						for (unsigned r = 0; r < 32; ++r) {
							for (unsigned l = 0; l < 32; ++l)
								for (unsigned k = 0; k < ipt; ++k) {
									const unsigned s = k * THREADS + warp * 32 + l;
									const double dx = xi[s] - jx[l], dy = yi[s] - jy[l], dz = zi[s] - jz[l];
									const double w = inv_cube(dx * dx + dy * dy + dz * dz);
									const double wi = uniform_mass ? w : w * jm[l];
									const double wj = uniform_mass ? w : w * mi[s];
									accm[0 * block + s] += wi * dx;
									accm[1 * block + s] += wi * dy;
									accm[2 * block + s] += wi * dz;
									ax[l] -= wj * dx;
									ay[l] -= wj * dy;
									az[l] -= wj * dz;
									credit(row[s], jidx[l]);
									if (row[s] < n_pad) credit(jidx[l], row[s]);  // ghosts have no mass
								}
							// __shfl_sync(v, (lane + 1) & 31): lane l takes lane l+1's value
							auto rot = [](auto* v) {
								auto first = v[0];
								for (unsigned l = 0; l < 31; ++l) v[l] = v[l + 1];
								v[31] = first;
							};
							rot(jx); rot(jy); rot(jz); rot(jm); rot(ax); rot(ay); rot(az); rot(jidx);
						}
						for (unsigned l = 0; l < 32; ++l) {
							if (jidx[l] != j_first + s0 + l) return -3;	 // not home
							if (jidx[l] >= rows_pad) return -4;
							aj[0 * (size_t)rows_pad + jidx[l]] += ax[l];
							aj[1 * (size_t)rows_pad + jidx[l]] += ay[l];
							aj[2 * (size_t)rows_pad + jidx[l]] += az[l];
						}
					}
			}
			for (unsigned s = 0; s < block; ++s) {
				if (row[s] >= rows_pad) return -5;
				for (int c = 0; c < 3; ++c)
					apart[((size_t)slice * rows_pad + row[s]) * 3 + c] = accm[c * (size_t)block + s];
			}
		}
	// what the stage kernel does: slice rows in order, then the j-side sum, mass scale, -G
	for (unsigned i = 0; i < n; ++i)
		for (int c = 0; c < 3; ++c) {
			double v = 0.;
			for (unsigned s = 0; s < n_slices; ++s) v += apart[((size_t)s * rows_pad + i) * 3 + c];
			v += aj[(size_t)c * rows_pad + i];
			acc_out[3 * (size_t)i + c] = v * um * (-G);
		}
	return 0;
}
