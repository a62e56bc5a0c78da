// sym_schedule.h -- work decomposition of the pair-symmetric force kernel (force_sym_kernels.cuh).
//
// Plain integer arithmetic shared by the device kernel, the host planner (nbody_capi.cu) and the
// CPU coverage test (tests/tools/sym_coverage.cpp), so that what the test proves is what the
// kernel walks.
//
// The bodies are cut into `nb` blocks of `tpb` j-tiles (one block == the i-rows of one CTA).
// Block a meets
//   * itself, one-sided (every ordered pair inside the block, self pairs included), and
//   * the next h(a) blocks in cyclic order, symmetrically: every pair (i in a, j in b) is
//     evaluated once and credited to both bodies (Newton's third law).
// With h(a) = (nb-1)/2 for odd nb every unordered pair of distinct blocks is met exactly once.  For
// even nb the blocks half-way round the ring, (a, a + nb/2), would be met from both sides with
// h = nb/2 and from neither with nb/2 - 1: exactly one of the two takes the pair -- the lower block
// when its index is even, the upper one when it is odd -- so that the blocks with the extra
// partner ALTERNATE along the ring and every contiguous run of blocks (= every rank of a
// multi-GPU run) carries the same load up to one partner block.  (Giving the extra partner to the
// whole lower half cost the lower ranks 1/(nb/2) more work: 0.8 % at N = 256k, measured on 2 GPUs.)
#pragma once

#if defined(__CUDACC__)
#define NB_HD __host__ __device__ __forceinline__
#else
#define NB_HD inline
#endif

namespace nbody {

// number of partner blocks of block a
NB_HD unsigned sym_half(unsigned nb, unsigned a) {
	if (nb & 1u) return (nb - 1u) / 2u;
	const unsigned half = nb / 2u;
	const unsigned lower = a < half ? a : a - half;	 // index of the lower block of the pair (a, a +- nb/2)
	const bool takes = a < half ? (lower & 1u) == 0u : (lower & 1u) != 0u;
	return takes ? half : half - 1u;
}

// j-tiles block a walks: tpb diagonal tiles first, then tpb per partner block
NB_HD unsigned sym_ntiles(unsigned nb, unsigned tpb, unsigned a) {
	return tpb * (1u + sym_half(nb, a));
}

// q-th tile of block a's walk (absolute tile index; may be a phantom tile >= jt_total when the
// last block is incomplete -- such tiles hold no bodies).  *diag = tile of the block itself.
NB_HD unsigned sym_tile(unsigned nb, unsigned tpb, unsigned a, unsigned q, bool* diag) {
	*diag = q < tpb;
	if (q < tpb) return a * tpb + q;
	const unsigned step = (q - tpb) / tpb;	// 0 .. h-1
	unsigned b = a + 1u + step;
	if (b >= nb) b -= nb;
	return b * tpb + (q - tpb) % tpb;
}

// A walk is cut into S slices at STRIP granularity (a strip = 32 consecutive j-bodies of a tile, the
// unit the kernel's warps consume): slice s takes strips [q0,q1) of the ntiles * spt strips of the
// walk (spt = strips per tile).  Fine slices are what lets small problems fill the GPU.
NB_HD void sym_slice_range(unsigned ntiles, unsigned spt, unsigned S, unsigned s, unsigned* q0,
						   unsigned* q1) {
	const unsigned long long nq = (unsigned long long)ntiles * spt;
	*q0 = (unsigned)(((unsigned long long)s * nq) / S);
	*q1 = (unsigned)(((unsigned long long)(s + 1u) * nq) / S);
}
// tiles [t0,t1) of the walk that the strip range [q0,q1) touches (t0 == t1 for an empty range)
NB_HD void sym_slice_tiles(unsigned spt, unsigned q0, unsigned q1, unsigned* t0, unsigned* t1) {
	*t0 = q0 / spt;
	*t1 = q1 > q0 ? (q1 + spt - 1u) / spt : q0 / spt;
}
// bodies [lo,hi) of tile t (index in the walk) that belong to the strip range [q0,q1)
NB_HD void sym_tile_part(unsigned t, unsigned spt, unsigned q0, unsigned q1, unsigned* lo,
						 unsigned* hi) {
	const unsigned first = t * spt, last = first + spt;
	*lo = ((q0 > first ? q0 : first) - first) * 32u;
	*hi = ((q1 < last ? q1 : last) - first) * 32u;
}

}  // namespace nbody
