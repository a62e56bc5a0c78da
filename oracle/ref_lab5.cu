/*
 * ref_lab5.cu -- harness around the REFERENCE'S OWN CUDA program
 * (lab_5_N_body_CUDA/src/N_body.cu, included unmodified from /root/reference; its main() is
 * renamed by the preprocessor).  Built for sm_100 exactly as the reference builds
 * (`nvcc -O3`, build.sh:1) -> oracle/_ref/lab5_ref.  TEST / REPORTING INFRASTRUCTURE ONLY:
 * it provides the "reference GPU kernel recompiled on the same B200" row and
 * an FP32 cross-check; nothing in the product links it.
 *
 * usage: lab5_ref N STEPS [dump.bin]
 *   runs the reference's ode_RK2_CUDA (N_body.cu:208-406; FP32 typedefs :19-22, block 64) on N
 *   uniform bodies (splitmix64, same distributions as lab_4's generate_data) for STEPS RK2
 *   steps and prints {"n":..,"steps":..,"seconds":..,"G_interactions_per_s":..}.
 *   N must be a multiple of 64: the reference kernel reads past the arrays otherwise
 *   (N_body.cu:157-158).
 */
#define main lab5_reference_main
#include "N_body.cu"
#undef main

#include <cstdint>
#include <cstdio>
#include <sstream>

static double u01(uint64_t seed, uint64_t c) {
	uint64_t z = seed + (c + 1) * 0x9E3779B97F4A7C15ull;
	z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
	z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
	z = z ^ (z >> 31);
	return (double)(z >> 11) * (1.0 / 9007199254740992.0);
}

int main(int argc, char** argv) {
	if (argc < 3) {
		std::fprintf(stderr, "usage: lab5_ref N STEPS\n");
		return 64;
	}
	const size_t n = std::strtoull(argv[1], nullptr, 10);
	const int steps = std::atoi(argv[2]);
	if (n % 64 != 0) {
		std::fprintf(stderr, "N must be a multiple of 64 (reference kernel reads out of bounds)\n");
		return 64;
	}
	std::vector<T> m(n);
	std::vector<T3> r(n), v(n);
	const uint64_t seed = 0x5EED0001;
	for (size_t i = 0; i < n; ++i) {
		m[i] = (T)(1e6 + (1e8 - 1e6) * u01(seed, 7 * i));
		r[i].x = (T)(-1e2 + 2e2 * u01(seed, 7 * i + 1));
		r[i].y = (T)(-1e2 + 2e2 * u01(seed, 7 * i + 2));
		r[i].z = (T)(-1e2 + 2e2 * u01(seed, 7 * i + 3));
		v[i].x = (T)(-1e1 + 2e1 * u01(seed, 7 * i + 4));
		v[i].y = (T)(-1e1 + 2e1 * u01(seed, 7 * i + 5));
		v[i].z = (T)(-1e1 + 2e1 * u01(seed, 7 * i + 6));
		r[i].w = v[i].w = 0;
	}
	// the reference loop runs while t <= t_final with t += tau in float: tau = 1, t_final =
	// steps - 0.5 gives exactly `steps` steps
	const std::string csv = "/tmp/lab5_ref_timing.csv";
	ode_RK2_CUDA(m, r, v, (T)0., (T)(steps - 0.5), (T)1., csv);
	std::ifstream in(csv);
	std::string header, row;
	std::getline(in, header);
	std::getline(in, row);
	// type;n_bodies;n_steps;block_size;time_taken;avg_exec_time
	std::vector<std::string> f;
	std::stringstream ss(row);
	for (std::string tok; std::getline(ss, tok, ';');) f.push_back(tok);
	const double seconds = f.size() > 4 ? std::atof(f[4].c_str()) : -1.;
	std::printf("{\"impl\": \"reference lab_5 N_body.cu recompiled\", \"type\": \"%s\", \"n\": %zu, "
				"\"steps\": %d, \"block_size\": %d, \"seconds\": %.6f, \"G_interactions_per_s\": %.3f}\n",
				T_TYPE_NAME.c_str(), n, steps, BLOCK_SIZE, seconds,
				2.0 * (double)n * (double)n * steps / seconds / 1e9);
	return 0;
}
