// nbody_main.cpp -- C++ host driver (the lab's `N_body_server`, B200 edition).
//
// Mirrors what the reference's two mains do by editing source
// (lab_4_N_body_MPI/src/N_body.cpp:38-87, lab_5_N_body_CUDA/src/N_body.cu:408-497),
// as run-time flags with the reference's parameter names and defaults:
//   --system default_test|random   ode_system_default_test / ode_system_random_data
//   --input PATH                    body file "N" + "m x y z vx vy vz" per line
//   --tau --t-start --t-final --eps --G
//   --precision 64|32   --scheme rk2|rk4   --gpus P   --flavour lab4|lab5
//   --traj-out PATH  --timing-out PATH  --final-state PATH  --out-every K  --max-steps K
//   --traj-bin PATH (binary sidecar)  --per-body-dir DIR [--per-body-first K --per-body-count M]
//   --bin2csv BIN --bin2csv-out CSV   --bin2perbody BIN --per-body-dir DIR   (converters, no GPU)
//   --output-folder DIR (where the reference-named CSVs go)  --report (energy, rates)
// All computation happens behind the C ABI (include/nbody_b200.h); no CPU fallback.
#include <array>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iostream>
#include <map>
#include <stdexcept>
#include <string>

#include "nbody_host.hpp"

using namespace nbody_host;

namespace {
struct Args {
	std::map<std::string, std::string> kv;
	bool has(const std::string& k) const { return kv.count(k) != 0; }
	std::string str(const std::string& k, const std::string& d) const {
		auto it = kv.find(k);
		return it == kv.end() ? d : it->second;
	}
	double num(const std::string& k, double d) const {
		auto it = kv.find(k);
		return it == kv.end() ? d : std::strtod(it->second.c_str(), nullptr);
	}
};

void usage() {
	std::puts(
		"nbody_b200 --input FILE [--system default_test|random] [--tau 0.1] [--t-start 0]\n"
		"           [--t-final 20|1] [--eps 1e-6] [--G 6.67e-11] [--precision 64|32]\n"
		"           [--scheme rk2|rk4|euler|midpoint|heun|ralston|kutta3|rk38|bs23|rkf45|cashkarp|dopri5]\n"
		"           [--adaptive-rtol R [--adaptive-atol A]]   (step-size control, embedded pairs; --tau = first step)\n"
		"           [--gpus 1] [--flavour lab4|lab5] [--traj-out FILE]\n"
		"           [--timing-out FILE] [--final-state FILE] [--out-every 1] [--max-steps 0]\n"
		"           [--traj-bin FILE] [--per-body-dir DIR [--per-body-first 1] [--per-body-count 0]]\n"
		"           [--output-folder ../output] [--variant -1] [--pair-symmetry [0|1]] [--deterministic] [--report]\n"
		"nbody_b200 --generate uniform|plummer --n N [--seed S] --generate-out FILE   (input generator)\n"
		"nbody_b200 --lab5-sweep DIR [--precision 32] [--output-folder ../output]   (lab_5 test_random_bodies)\n"
		"nbody_b200 --reformat IN --reformat-out OUT   (parse a body file, rewrite it losslessly; no GPU)\n"
		"nbody_b200 --convergence FIXTURE_DIR --input 4body.txt [--scheme rk2] [--taus 0.1,0.05,0.025]\n"
		"           (the reference's commented convergence runs, N_body.cpp:69-79: error of each run against\n"
		"            FIXTURE_DIR/traj<k>.csv and the observed order)\n"
		"nbody_b200 --bin2csv BIN --bin2csv-out CSV   (binary trajectory -> the reference's CSV; no GPU)\n"
		"nbody_b200 --bin2perbody BIN --per-body-dir DIR [--per-body-first 1] [--per-body-count 0]   (-> traj<k>.csv; no GPU)\n"
		"nbody_b200 --list-tableaux   (the schemes --scheme accepts, with their coefficients; no GPU)");
}

// The reference's commented convergence runs (lab_4 N_body.cpp:69-79: ode_system_default_test at
// tau = 0.1, 0.05, 0.025) as one command: each run's rows at t = 0.1 k against the reference's
// accurate per-body solution FIXTURE_DIR/traj<k>.csv, max |error| over bodies, samples and
// coordinates, and the observed order log2(err(tau) / err(tau/2)).  One JSON line on stdout.
template <typename T>
int run_convergence(const Args& a) {
	const std::string dir = a.str("convergence", ".");
	std::vector<double> taus;
	{
		std::string list = a.str("taus", "0.1,0.05,0.025");
		std::size_t pos = 0;
		while (pos < list.size()) {
			std::size_t end = list.find(',', pos);
			if (end == std::string::npos) end = list.size();
			taus.push_back(std::strtod(list.substr(pos, end - pos).c_str(), nullptr));
			pos = end + 1;
		}
	}
	RunOptions opt;
	const std::string scheme = a.str("scheme", "rk2");
	opt.scheme = scheme == "rk4" ? NBODY_RK4 : NBODY_RK2;
	if (scheme != "rk2" && scheme != "rk4") {
		if (!find_tableau(scheme)) {
			std::cerr << "unknown --scheme " << scheme << std::endl;
			return 2;
		}
		opt.tableau = scheme;
	}
	opt.n_gpus = (int)a.num("gpus", 1);
	opt.keep_rows = true;
	opt.timing_path = a.str("timing-out", "/dev/null");
	opt.pair_symmetry = a.has("pair-symmetry") ? (a.str("pair-symmetry", "1") == "0" ? 0 : 1) : -1;
	std::vector<std::vector<std::array<double, 4>>> truth;
	std::vector<double> errs;
	try {
		for (const double tau : taus) {
			ode_system<T> sys;
			sys.ode_system_default_test((T)tau, (T)a.num("eps", 1e-6), (T)a.num("G", 6.67e-11),
										a.str("input", "./include/N_body/N_body_config/4body.txt"));
			sys.n_time_steps = (int)((sys.t_final - sys.t_start) / sys.tau);
			sys.N_bodies = (int)sys.traj_init.size();
			if (truth.empty())
				for (std::size_t k = 1; k <= sys.traj_init.size(); ++k)
					truth.push_back(read_per_body_file(dir + "/traj" + std::to_string(k) + ".csv"));
			Solver_N_body<T> solver;
			solver.output_folder = a.str("output-folder", "../output");
			solver.solve_parallel_CUDA(sys, opt);
			// rows: t_start, then one per step; the fixtures sample t = 0, 0.1, 0.2, ...
			const double dt_fix = truth[0].size() > 1 ? truth[0][1][0] - truth[0][0][0] : 0.1;
			const long stride = std::lround(dt_fix / tau);
			if (stride < 1 || std::fabs((double)stride * tau - dt_fix) > 1e-9 * dt_fix) {
				std::cerr << "--taus: " << tau << " does not divide the fixtures' sampling interval " << dt_fix
						  << std::endl;
				return 2;
			}
			double err = 0;
			for (std::size_t k = 0; k < truth.size(); ++k)
				for (std::size_t s = 0; s < truth[k].size(); ++s) {
					const std::size_t row = s * (std::size_t)stride;
					if (row >= solver.kept_rows.size()) break;
					for (int c = 0; c < 3; ++c)
						err = std::max(err, std::fabs(solver.kept_rows[row][1 + 3 * k + c] - truth[k][s][1 + c]));
				}
			errs.push_back(err);
		}
	} catch (int code) {
		return code;
	} catch (const char* msg) {
		std::cerr << msg << std::endl;
		return 2;
	} catch (const std::exception& e) {
		std::cerr << "error: " << e.what() << std::endl;
		return 3;
	}
	std::printf("{\"scheme\": \"%s\", \"taus\": [", scheme.c_str());
	for (std::size_t i = 0; i < taus.size(); ++i) std::printf("%s%.17g", i ? ", " : "", taus[i]);
	std::printf("], \"max_abs_err\": [");
	for (std::size_t i = 0; i < errs.size(); ++i) std::printf("%s%.10e", i ? ", " : "", errs[i]);
	std::printf("], \"observed_order\": [");
	for (std::size_t i = 0; i + 1 < errs.size(); ++i)
		std::printf("%s%.4f", i ? ", " : "", std::log(errs[i] / errs[i + 1]) / std::log(taus[i] / taus[i + 1]));
	std::printf("]}\n");
	return 0;
}

template <typename T>
int run(const Args& a) {
	ode_system<T> sys;
	const std::string system = a.str("system", "random");
	const T tau = (T)a.num("tau", 0.1);
	const T eps = (T)a.num("eps", 1e-6);
	const T G = (T)a.num("G", 6.67e-11);
	try {
		if (system == "default_test")
			sys.ode_system_default_test(tau, eps, G,
										a.str("input", "./include/N_body/N_body_config/4body.txt"));
		else
			sys.ode_system_random_data(0, tau, eps, G, a.str("input", "../output/rand_bodies.txt"));
	} catch (int code) {
		return code;  // the reference lets `throw 1` terminate the program
	}
	if (a.has("t-start")) sys.t_start = (T)a.num("t-start", 0);
	if (a.has("t-final")) sys.t_final = (T)a.num("t-final", 1);
	sys.n_time_steps = (int)((sys.t_final - sys.t_start) / sys.tau);
	sys.N_bodies = (int)sys.traj_init.size();

	RunOptions opt;
	const std::string scheme = a.str("scheme", "rk2");
	opt.scheme = scheme == "rk4" ? NBODY_RK4 : NBODY_RK2;
	if (scheme != "rk2" && scheme != "rk4") {  // a named tableau (nbody_tableaux.hpp)
		if (!find_tableau(scheme)) {
			std::cerr << "unknown --scheme " << scheme << std::endl;
			return 2;
		}
		opt.tableau = scheme;
	}
	opt.adaptive_rtol = a.num("adaptive-rtol", 0);
	opt.adaptive_atol = a.num("adaptive-atol", 0);
	if (opt.adaptive_rtol > 0 && opt.tableau.empty()) opt.tableau = "dopri5";
	opt.n_gpus = (int)a.num("gpus", 1);
	opt.flavour = a.str("flavour", "lab4");
	opt.traj_path = a.str("traj-out", "");
	opt.traj_bin_path = a.str("traj-bin", "");
	opt.per_body_dir = a.str("per-body-dir", "");
	{  // bodies are numbered from 1 on the command line, as the reference's traj1..traj4 are
		const long f = (long)a.num("per-body-first", 1);
		opt.per_body_first = f > 1 ? (std::size_t)(f - 1) : 0;
		const long c = (long)a.num("per-body-count", 0);
		opt.per_body_count = c > 0 ? (std::size_t)c : 0;
	}
	opt.timing_path = a.str("timing-out", "");
	opt.final_state_path = a.str("final-state", "");
	opt.out_every = (long)a.num("out-every", 1);
	if (opt.out_every < 1) opt.out_every = 1;
	opt.max_steps = (long)a.num("max-steps", 0);
	opt.variant = (long)a.num("variant", -1);
	// --pair-symmetry [0|1]: 1 (or bare flag) forces the pair-symmetric kernel, 0 the one-sided
	// bit-reproducible kernel; absent = the library's automatic choice
	opt.pair_symmetry = a.has("pair-symmetry") ? (a.str("pair-symmetry", "1") == "0" ? 0 : 1) : -1;
	// --deterministic: fixed-point row sums in the pair-symmetric kernel (same bits every run)
	opt.deterministic = a.has("deterministic") ? (a.str("deterministic", "1") == "0" ? 0 : 1) : 0;
	opt.report = a.has("report");

	Solver_N_body<T> solver;
	solver.output_folder = a.str("output-folder", "../output");
	solver.notifications = a.has("notifications");
	// the reference derives the trajectory name for the 4-body test
	// (class_Solver_N_body.hpp:185-187)
	if (opt.traj_path.empty() && sys.name == "4Body_test" && !a.has("no-traj"))
		opt.traj_path = solver.output_folder + "/results_" + std::to_string(sys.tau) + ".csv";
	try {
		solver.solve_parallel_CUDA(sys, opt);
	} catch (const char* msg) {
		std::cerr << msg << std::endl;
		return 2;
	} catch (const std::exception& e) {
		std::cerr << "error: " << e.what() << std::endl;
		return 3;
	}
	if (opt.report) {
		const RunResult& r = solver.result;
		const double e0 = r.e_kin0 + r.e_pot0, e1 = r.e_kin1 + r.e_pot1;
		std::printf(
			"{\"n\": %zu, \"precision\": %d, \"scheme\": \"%s\", \"gpus\": %d, \"steps\": %ld, \"rejected\": %ld, "
			"\"wall_s\": %.6f, \"loop_s\": %.6f, \"G_interactions_per_s\": %.3f, \"writer_busy_s\": %.6f, "
			"\"energy0\": %.17g, \"energy1\": %.17g, \"energy_drift\": %.3e}\n",
			sys.traj_init.size(), (int)(8 * sizeof(T)),
			!opt.tableau.empty() ? opt.tableau.c_str() : opt.scheme == NBODY_RK4 ? "rk4" : "rk2",
			opt.n_gpus, r.steps, r.rejected, r.wall_s, r.loop_s, r.interactions / r.loop_s / 1e9, r.writer_busy_s, e0,
			e1,
			e0 != 0 ? std::fabs(e1 - e0) / std::fabs(e0) : 0.0);
	}
	return 0;
}
// one entry of lab_5's sweep: read_data into SoA vectors, ode_RK2_CUDA(m, r, v, 0, 2, 0.1, path)
template <typename T>
void lab5_run(const std::string& in, const std::string& timing_csv) {
	const std::vector<Body<T>> b = read_data<T>(in);
	std::vector<T> m(b.size());
	std::vector<Vec4<T>> r(b.size()), v(b.size());
	for (std::size_t i = 0; i < b.size(); ++i) {
		m[i] = b[i].m;
		r[i] = {b[i].r[0], b[i].r[1], b[i].r[2], T(0)};
		v[i] = {b[i].v[0], b[i].v[1], b[i].v[2], T(0)};
	}
	ode_RK2_CUDA<T>(m, r, v, T(0.), T(2.), T(0.1), timing_csv);
}
}  // namespace

int main(int argc, char** argv) {
	Args a;
	for (int i = 1; i < argc; ++i) {
		std::string k = argv[i];
		if (k == "-h" || k == "--help") {
			usage();
			return 0;
		}
		if (k.rfind("--", 0) != 0) {
			std::cerr << "unexpected argument " << k << std::endl;
			usage();
			return 64;
		}
		k = k.substr(2);
		if (i + 1 < argc && std::strncmp(argv[i + 1], "--", 2) != 0) a.kv[k] = argv[++i];
		else a.kv[k] = "1";
	}
	if (a.has("list-tableaux")) {  // name stages order err_order | a (row-major) | b | b_err; no GPU
		for (const Tableau& t : tableaux()) {
			std::printf("%s %d %d %d", t.name, t.stages, t.order, t.embedded ? t.err_order : 0);
			for (double v : t.flat_a()) std::printf(" %.17g", v);
			for (int j = 0; j < t.stages; ++j) std::printf(" %.17g", t.b[j]);
			for (int j = 0; j < t.stages; ++j) std::printf(" %.17g", t.embedded ? t.b_err[j] : 0.0);
			std::printf("\n");
		}
		return 0;
	}
	if (a.has("generate")) {  // --generate uniform|plummer --n N [--seed S] --generate-out FILE
		const std::size_t n = (std::size_t)a.num("n", 0);
		const std::string kind = a.str("generate", "uniform"), out = a.str("generate-out", "");
		if (n == 0 || out.empty()) {
			std::cerr << "--generate needs --n N and --generate-out FILE" << std::endl;
			return 64;
		}
		const bool plummer = kind == "plummer";
		const std::uint64_t seed =
			a.has("seed") ? std::strtoull(a.str("seed", "0").c_str(), nullptr, 0)
						  : (plummer ? 0x5EED0004ull : 0x5EED0001ull);
		write_input_format(out, plummer ? generate_plummer(n, seed, a.num("G", 6.67e-11))
										: generate_uniform(n, seed));
		if (!a.has("input")) return 0;
	}
	if (a.has("reformat")) {  // --reformat IN --reformat-out OUT: read_data -> input format (no GPU)
		try {
			write_input_format(a.str("reformat-out", "/dev/stdout"),
							   read_data<double>(a.str("reformat", "")));
		} catch (int code) {
			return code;
		}
		return 0;
	}
	if (a.has("bin2csv") || a.has("bin2perbody")) {	 // converters of the --traj-bin sidecar (no GPU)
		std::string why;
		long rows;
		if (a.has("bin2csv"))
			rows = traj_bin_to_csv(a.str("bin2csv", ""), a.str("bin2csv-out", "/dev/stdout"), &why);
		else {
			const long f = (long)a.num("per-body-first", 1), c = (long)a.num("per-body-count", 0);
			rows = traj_bin_to_per_body(a.str("bin2perbody", ""), a.str("per-body-dir", "."),
										f > 1 ? (std::size_t)(f - 1) : 0, c > 0 ? (std::size_t)c : 0, &why);
		}
		if (rows < 0) {
			std::cerr << why << std::endl;
			return 1;
		}
		std::cerr << rows << " rows converted" << std::endl;
		return 0;
	}
	if (a.has("lab5-sweep")) {	// lab_5's test_random_bodies() (N_body.cu:438-491): RK2, T=[0,2], tau=0.1
		const std::string dir = a.str("lab5-sweep", "./include");
		const std::string out = a.str("output-folder", "../output");
		const int prec5 = (int)a.num("precision", 32);	// the reference's committed typedef is float
		int ran = 0;
		for (const char* name : {"10k", "50k", "100k", "500k"}) {
			const std::string in = dir + "/" + name + "_bodies.txt";
			if (!std::ifstream(in).good()) continue;  // the reference commits only 10k_bodies.txt
			try {
				if (prec5 == 32) lab5_run<float>(in, out + "/res_" + name + "_float_128_.csv");
				else lab5_run<double>(in, out + "/res_" + name + "_double_128_.csv");
			} catch (int code) {
				return code;
			} catch (const std::exception& e) {
				std::cerr << "error: " << e.what() << std::endl;
				return 3;
			}
			++ran;
		}
		std::cout << "lab5 sweep: " << ran << " input file(s) run" << std::endl;
		return ran ? 0 : 1;
	}
	const int prec = (int)a.num("precision", 64);
	if (a.has("convergence")) return prec == 32 ? run_convergence<float>(a) : run_convergence<double>(a);
	if (prec == 64) return run<double>(a);
	if (prec == 32) return run<float>(a);
	std::cerr << "--precision must be 64 or 32" << std::endl;
	return 64;
}
