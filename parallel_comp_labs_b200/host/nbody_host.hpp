// nbody_host.hpp -- C++ host side of the B200 N-body path: the lab's own interface
// (same type and function names, argument meaning, file formats and error behaviour)
// re-implemented over the C ABI of include/nbody_b200.h.  Nothing here computes forces
// or integrates: that is the CUDA library's job, and there is no CPU fallback.
//
// Reference interfaces mirrored (file:line relative to the reference root):
//   Body<T>                               lab_4 .../struct_Body.hpp:5-10
//   read_data<T>, write_bodies_coords     lab_4 .../utils/read_write_utils.hpp:12-52
//   ode_system<T> + two factories         lab_4 .../struct_ode_system.hpp:20-134
//   Solver_N_body<T>::solve_parallel_MPI  lab_4 .../class_Solver_N_body.hpp:87-208
//   time loop / loop control              lab_4 .../schemes/ode_RK2_MPI.hpp:55-104
//   ode_RK2_CUDA(m,r,v,t0,t1,tau,p,p1)    lab_5 N_body.cu:208-406
#pragma once
#include <array>
#include <atomic>
#include <algorithm>
#include <charconv>
#include <cmath>
#include <cstdint>
#include <chrono>
#include <condition_variable>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <fstream>
#include <functional>
#include <iostream>
#include <memory>
#include <mutex>
#include <sstream>
#include <stdexcept>
#include <string>
#include <thread>
#include <type_traits>
#include <vector>

#include "nbody_b200.h"
#include "nbody_tableaux.hpp"

namespace nbody_host {

template <typename T>
struct Body {  // memory image identical to the reference's: 7 x T, no padding
	T m;
	std::array<T, 3> r;
	std::array<T, 3> v;
};
static_assert(sizeof(Body<double>) == 56 && sizeof(Body<float>) == 28, "Body<T> must be packed");

// lab_5's T3 (float4 / double4 used as a 3-vector, .w ignored)
template <typename T>
struct alignas(4 * sizeof(T) > 16 ? 16 : 4 * sizeof(T)) Vec4 {
	T x, y, z, w;
};

// ---- input ---------------------------------------------------------------------------
// "N" then N records "m x y z vx vy vz", any whitespace (tabs, CRLF).  Unopenable file:
// prints the reference's message and throws int 1 (read_write_utils.hpp:15-19).
template <typename T>
std::vector<Body<T>> read_data(const std::string& path) {
	std::ifstream in(path, std::ios::binary);
	if (!in.is_open()) {
		std::cout << "file is't open" << std::endl;
		throw 1;
	}
	std::string text((std::istreambuf_iterator<char>(in)), std::istreambuf_iterator<char>());
	const char* p = text.data();
	const char* end = p + text.size();
	auto skip = [&] { while (p < end && (*p == ' ' || *p == '\t' || *p == '\n' || *p == '\r' || *p == '\f' || *p == '\v')) ++p; };
	auto next = [&](double& out) {
		skip();
		if (p < end && *p == '+') ++p;
		auto res = std::from_chars(p, end, out);
		if (res.ec != std::errc()) return false;
		p = res.ptr;
		return true;
	};
	double nn = 0;
	std::vector<Body<T>> bodies;
	if (!next(nn) || nn < 0) return bodies;
	const std::size_t n = (std::size_t)nn;
	bodies.reserve(n);
	// operator>> semantics on a short file (read_write_utils.hpp:27-33): fields are assigned one
	// by one; at end of input the extraction fails without touching its field and sets failbit,
	// after which nothing is assigned any more and every later record repeats `body`
	// (checked against the reference build: tests/test_host_logic.py).
	Body<T> b{};
	bool failed = false;
	for (std::size_t i = 0; i < n; ++i) {
		T* field[7] = {&b.m, &b.r[0], &b.r[1], &b.r[2], &b.v[0], &b.v[1], &b.v[2]};
		for (int k = 0; k < 7 && !failed; ++k) {
			double f = 0;
			if (next(f)) *field[k] = (T)f;
			else failed = true;
		}
		bodies.push_back(b);
	}
	return bodies;
}

// ---- output --------------------------------------------------------------------------
// Appends "%.16e" of x (17 significant digits, at least 2 exponent digits) + ';'.
inline void append_sci16(std::string& s, double x) {
	char buf[40];
	auto res = std::to_chars(buf, buf + sizeof(buf), x, std::chars_format::scientific, 16);
	s.append(buf, res.ptr);
	s.push_back(';');
}

// One row "t;x1;y1;z1;...;zN;\n" (write_bodies_coords, read_write_utils.hpp:42-52);
// `pos` is n x 3.  The row is flushed like std::endl does.
template <typename T>
void write_positions_row(std::ostream& out, double time, const T* pos, std::size_t n) {
	std::string row;
	row.reserve(24 * (3 * n + 1) + 2);
	append_sci16(row, time);
	for (std::size_t i = 0; i < 3 * n; ++i) append_sci16(row, (double)pos[i]);
	row.push_back('\n');
	out.write(row.data(), (std::streamsize)row.size());
	out.flush();
}

template <typename T>
void write_bodies_coords(std::ostream& out, double time, const std::vector<Body<T>>& bodies) {
	std::vector<T> pos(3 * bodies.size());
	for (std::size_t i = 0; i < bodies.size(); ++i)
		for (int k = 0; k < 3; ++k) pos[3 * i + k] = bodies[i].r[k];
	write_positions_row(out, time, pos.data(), bodies.size());
}

inline const char* lab4_header() {	// ode_RK2_MPI.hpp:52 (hard-coded for 4 bodies)
	return "time;r11;r12;r13;r21;r22;r23;r31;r32;r33;r41;r42;r43\n";
}
inline std::string lab5_header(std::size_t n) {	 // N_body.cu:236-246
	std::string h = "time;";
	for (std::size_t i = 1; i <= n; ++i)
		for (int j = 1; j <= 3; ++j) {
			h += "r_" + std::to_string(i) + "_" + std::to_string(j);
			h += (i == n && j == 3) ? "\n" : ";";
		}
	return h;
}

// Final state in the INPUT format (restartable runs; SURVEY.md 5 "checkpoint").
template <typename T>
void write_input_format(const std::string& path, const std::vector<Body<T>>& bodies) {
	std::ofstream out(path);
	out << bodies.size() << "\n";
	char buf[40];
	std::string line;
	for (const auto& b : bodies) {
		line.clear();
		const double f[7] = {(double)b.m, (double)b.r[0], (double)b.r[1], (double)b.r[2],
							 (double)b.v[0], (double)b.v[1], (double)b.v[2]};
		for (int k = 0; k < 7; ++k) {
			auto res = std::to_chars(buf, buf + sizeof(buf), f[k], std::chars_format::scientific, 16);
			line.append(buf, res.ptr);
			line.push_back(k == 6 ? '\n' : '\t');
		}
		out << line;
	}
}

// ---- synthetic inputs ---------------------------------------------------
// Same counter-based splitmix64 streams as parallel_comp_labs_b200/generators.py.  `uniform` is
// bit-identical to the Python generator; `plummer` uses the same draws but libm's pow/cos/sin,
// so it may differ from numpy's in the last ulp.
inline double splitmix_u01(std::uint64_t seed, std::uint64_t counter) {
	std::uint64_t z = seed + (counter + 1) * 0x9E3779B97F4A7C15ull;
	z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
	z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
	z = z ^ (z >> 31);
	return (double)(z >> 11) * (1.0 / 9007199254740992.0);
}

// the distributions of the reference's commented-out generate_data (lab_4 N_body.cpp:15-35)
inline std::vector<Body<double>> generate_uniform(std::size_t n, std::uint64_t seed = 0x5EED0001) {
	std::vector<Body<double>> b(n);
	for (std::size_t i = 0; i < n; ++i) {
		b[i].m = 1e6 + (1e8 - 1e6) * splitmix_u01(seed, 7 * i);
		for (int k = 0; k < 3; ++k) {
			b[i].r[k] = -1e2 + 2e2 * splitmix_u01(seed, 7 * i + 1 + k);
			b[i].v[k] = -1e1 + 2e1 * splitmix_u01(seed, 7 * i + 4 + k);
		}
	}
	return b;
}

// Plummer sphere, a = 1, G M = 1, equal masses, radius cut 22.8 a (Aarseth-Henon-Wielen)
inline std::vector<Body<double>> generate_plummer(std::size_t n, std::uint64_t seed = 0x5EED0004,
												  double G = 6.67e-11, double r_cut = 22.8) {
	const double pi = 3.14159265358979323846;
	std::vector<Body<double>> b(n);
	const std::uint64_t s2 = seed ^ 0xA5A5A5A5ull, s3 = seed ^ 0x3C3C3C3Cull;
	double com[3] = {0, 0, 0}, vcm[3] = {0, 0, 0};
	for (std::size_t i = 0; i < n; ++i) {
		b[i].m = (1.0 / G) / (double)n;
		double r = 0;
		for (std::uint64_t a = 0;; ++a) {
			double x1 = splitmix_u01(seed, 64 * i + a);
			x1 = std::min(std::max(x1, 1e-12), 1.0 - 1e-12);
			r = 1.0 / std::sqrt(std::pow(x1, -2.0 / 3.0) - 1.0);
			if (r <= r_cut) break;
		}
		const double x2 = splitmix_u01(s2, 4 * i), x3 = splitmix_u01(s2, 4 * i + 1);
		double z = (1.0 - 2.0 * x2) * r, rxy = std::sqrt(std::max(r * r - z * z, 0.0));
		b[i].r = {rxy * std::cos(2 * pi * x3), rxy * std::sin(2 * pi * x3), z};
		double q = 0;
		for (std::uint64_t a = 0;; ++a) {
			const double x4 = splitmix_u01(s3, 512 * i + 2 * a), x5 = splitmix_u01(s3, 512 * i + 2 * a + 1);
			if (0.1 * x5 < x4 * x4 * std::pow(1.0 - x4 * x4, 3.5)) { q = x4; break; }
		}
		const double v = q * std::sqrt(2.0) * std::pow(1.0 + r * r, -0.25);
		const double x6 = splitmix_u01(s2, 4 * i + 2), x7 = splitmix_u01(s2, 4 * i + 3);
		z = (1.0 - 2.0 * x6) * v;
		rxy = std::sqrt(std::max(v * v - z * z, 0.0));
		b[i].v = {rxy * std::cos(2 * pi * x7), rxy * std::sin(2 * pi * x7), z};
		for (int k = 0; k < 3; ++k) { com[k] += b[i].r[k]; vcm[k] += b[i].v[k]; }
	}
	for (auto& q : b)
		for (int k = 0; k < 3; ++k) { q.r[k] -= com[k] / (double)n; q.v[k] -= vcm[k] / (double)n; }
	return b;
}

// ---- problem descriptor ------------------------------------------------------------------
template <typename T>
struct ode_system {
	T tau = T(0.1);
	T t_start = 0, t_final = 1;
	int n_time_steps = 10, N_bodies = 0;
	std::vector<Body<T>> traj_init;
	std::string name;
	// the RHS lambda's captures (struct_ode_system.hpp:53): the operator itself lives on
	// the GPU behind nbody_rhs / nbody_step
	T eps = T(1e-6), G = T(6.67e-11);

	// struct_ode_system.hpp:34-81
	ode_system<T>& ode_system_default_test(const T time_step = 0.1, const T eps_ = 1e-6,
										   const T G_ = 6.67e-11,
										   const std::string& path =
											   "./include/N_body/N_body_config/4body.txt") {
		name = "4Body_test";
		t_start = 0.;
		t_final = 20.;
		tau = time_step;
		n_time_steps = (int)((t_final - t_start) / tau);
		traj_init = read_data<T>(path);
		N_bodies = (int)traj_init.size();
		eps = eps_;
		G = G_;
		return *this;
	}
	// struct_ode_system.hpp:84-133; N_of_bodies is ignored as in the reference
	ode_system<T>& ode_system_random_data(const std::size_t N_of_bodies, const T time_step,
										  const T eps_, const T G_, const std::string& path) {
		name = "4Body_random";
		t_start = 0.;
		t_final = 1.;
		tau = time_step;
		n_time_steps = (int)((t_final - t_start) / tau);
		N_bodies = (int)N_of_bodies;
		traj_init = read_data<T>(path);
		eps = eps_;
		G = G_;
		return *this;
	}
};

// ---- where a trajectory row goes ---------------------------------------------------------------
// One row = (time, n x 3 positions).  Three independent outputs:
//  * the whole-system CSV of the reference, "t;x1;y1;z1;...;zN;" (write_bodies_coords,
//    read_write_utils.hpp:42-52) -- byte-compatible, 72 MB per row at N = 1M;
//  * a binary sidecar: 32-byte header {"NBTRJ01", n, sizeof(T), flavour} + per row {double t,
//    T pos[3n]} -- what a long run at large N should write (3n sizeof(T) bytes per row instead of
//    72 n characters); `nbody_b200 --bin2csv` / `--bin2perbody` turn it into the text formats
//    afterwards, byte-identical to what a direct run writes;
//  * per-body files traj<k>.csv, k = 1..: "time;r1;r2;r3" then "t;x;y;z" per row -- the layout of the
//    reference's own fixtures (N_body_config/traj1.csv:1), for diffing body by body.
constexpr std::size_t MAX_PER_BODY_FILES = 256;	 // open at once (fd limit); select a range above that

struct TrajBinHeader {
	char magic[8];		 // "NBTRJ01\0"
	std::uint64_t n;	 // bodies per row
	std::uint32_t elem;	 // sizeof(T): 8 or 4
	std::uint32_t flavour;	// 4 = lab_4 (4-body header, t_start row), 5 = lab_5 (N-body header)
	std::uint64_t reserved;
};
static_assert(sizeof(TrajBinHeader) == 32, "binary trajectory header is 32 bytes");

inline void append_sci16_bare(std::string& s, double x) {
	char buf[40];
	auto res = std::to_chars(buf, buf + sizeof(buf), x, std::chars_format::scientific, 16);
	s.append(buf, res.ptr);
}

struct PerBodyFiles {
	std::vector<std::ofstream> files;
	std::size_t first = 0;	// 0-based index of the body traj<first+1>.csv follows
	// opens DIR/traj<k>.csv for bodies [first, first + count); false (with `why`) on failure
	bool open(const std::string& dir, std::size_t n, std::size_t first_body, std::size_t count,
			  std::string* why) {
		if (first_body >= n) { *why = "--per-body-first is past the last body"; return false; }
		if (count == 0 || first_body + count > n) count = n - first_body;
		if (count > MAX_PER_BODY_FILES) {
			*why = "per-body trajectory files: " + std::to_string(count) + " bodies selected, at most " +
				   std::to_string(MAX_PER_BODY_FILES) + " files are kept open at once; narrow the range "
				   "with --per-body-first / --per-body-count (or write --traj-bin and convert ranges "
				   "with --bin2perbody afterwards)";
			return false;
		}
		first = first_body;
		files.resize(count);
		for (std::size_t k = 0; k < count; ++k) {
			const std::string path = dir + "/traj" + std::to_string(first + k + 1) + ".csv";
			files[k].open(path);
			if (!files[k].is_open()) { *why = "cannot open " + path; return false; }
			files[k] << "time;r1;r2;r3\n";	 // N_body_config/traj1.csv:1
		}
		return true;
	}
	template <typename T>
	void row(double time, const T* pos) {
		std::string line;
		for (std::size_t k = 0; k < files.size(); ++k) {
			line.clear();
			append_sci16(line, time);
			const T* p = pos + 3 * (first + k);
			append_sci16(line, (double)p[0]);
			append_sci16(line, (double)p[1]);
			append_sci16_bare(line, (double)p[2]);
			line.push_back('\n');
			files[k].write(line.data(), (std::streamsize)line.size());
		}
	}
	void flush() { for (auto& f : files) f.flush(); }
};

template <typename T>
struct TrajectorySinks {
	std::ostream* csv = nullptr;   // whole-system CSV (header already written by the caller)
	std::ofstream bin;			   // binary sidecar
	PerBodyFiles per_body;
	std::vector<std::vector<double>>* keep = nullptr;  // rows {t, 3n positions} kept in memory
	std::size_t n = 0;
	bool any() const { return csv || keep || bin.is_open() || !per_body.files.empty(); }
	bool open_bin(const std::string& path, std::size_t n_bodies, const std::string& flavour) {
		bin.open(path, std::ios::binary);
		if (!bin.is_open()) return false;
		TrajBinHeader h;
		std::memset(&h, 0, sizeof h);
		std::memcpy(h.magic, "NBTRJ01", 8);
		h.n = n_bodies;
		h.elem = (std::uint32_t)sizeof(T);
		h.flavour = flavour == "lab5" ? 5u : 4u;
		bin.write(reinterpret_cast<const char*>(&h), sizeof h);
		return true;
	}
	void row(double time, const T* pos) {
		if (csv) write_positions_row(*csv, time, pos, n);
		if (bin.is_open()) {
			bin.write(reinterpret_cast<const char*>(&time), sizeof time);
			bin.write(reinterpret_cast<const char*>(pos), (std::streamsize)(3 * n * sizeof(T)));
		}
		if (!per_body.files.empty()) per_body.row(time, pos);
		if (keep) {
			std::vector<double> r(1 + 3 * n);
			r[0] = time;
			for (std::size_t i = 0; i < 3 * n; ++i) r[1 + i] = (double)pos[i];
			keep->push_back(std::move(r));
		}
	}
	void finish() {
		if (bin.is_open()) bin.flush();
		per_body.flush();
	}
};

// ---- converters of the binary sidecar (no GPU) -------------------------------------------------
// Reads the header; returns false with `why` if the file is not a trajectory sidecar.
inline bool read_traj_bin_header(std::ifstream& in, TrajBinHeader* h, std::string* why) {
	in.read(reinterpret_cast<char*>(h), sizeof *h);
	if (!in || std::memcmp(h->magic, "NBTRJ01", 8) != 0 || (h->elem != 8 && h->elem != 4) || h->n == 0 ||
		(h->flavour != 4 && h->flavour != 5)) {
		*why = "not a trajectory sidecar written by --traj-bin";
		return false;
	}
	return true;
}

// Calls row(time, pos) for every complete record; returns the number of rows.
template <typename T, typename F>
long for_each_traj_row(std::ifstream& in, const TrajBinHeader& h, F&& row) {
	std::vector<T> pos(3 * (std::size_t)h.n);
	long rows = 0;
	for (;;) {
		double t;
		in.read(reinterpret_cast<char*>(&t), sizeof t);
		in.read(reinterpret_cast<char*>(pos.data()), (std::streamsize)(pos.size() * sizeof(T)));
		if (!in) break;	 // a truncated last record (a killed run) is dropped
		row(t, pos.data());
		++rows;
	}
	return rows;
}

// sidecar -> the whole-system CSV a direct --traj-out run writes, byte for byte
inline long traj_bin_to_csv(const std::string& bin_path, const std::string& csv_path, std::string* why) {
	std::ifstream in(bin_path, std::ios::binary);
	if (!in.is_open()) { *why = "cannot open " + bin_path; return -1; }
	TrajBinHeader h;
	if (!read_traj_bin_header(in, &h, why)) return -1;
	std::ofstream out(csv_path);
	if (!out.is_open()) { *why = "cannot open " + csv_path; return -1; }
	out << (h.flavour == 5 ? lab5_header((std::size_t)h.n) : std::string(lab4_header()));
	const std::size_t n = (std::size_t)h.n;
	if (h.elem == 8)
		return for_each_traj_row<double>(in, h, [&](double t, const double* p) { write_positions_row(out, t, p, n); });
	return for_each_traj_row<float>(in, h, [&](double t, const float* p) { write_positions_row(out, t, p, n); });
}

// sidecar -> per-body files for bodies [first, first + count)
inline long traj_bin_to_per_body(const std::string& bin_path, const std::string& dir, std::size_t first,
								 std::size_t count, std::string* why) {
	std::ifstream in(bin_path, std::ios::binary);
	if (!in.is_open()) { *why = "cannot open " + bin_path; return -1; }
	TrajBinHeader h;
	if (!read_traj_bin_header(in, &h, why)) return -1;
	PerBodyFiles files;
	if (!files.open(dir, (std::size_t)h.n, first, count, why)) return -1;
	long rows;
	if (h.elem == 8) rows = for_each_traj_row<double>(in, h, [&](double t, const double* p) { files.row(t, p); });
	else rows = for_each_traj_row<float>(in, h, [&](double t, const float* p) { files.row(t, p); });
	files.flush();
	return rows;
}

// A per-body file "time;r1;r2;r3" + "t;x;y;z" rows -- ours or the reference's fixtures
// (N_body_config/traj1.csv, which write "0." and "1." for whole numbers): rows of {t, x, y, z}.
inline std::vector<std::array<double, 4>> read_per_body_file(const std::string& path) {
	std::ifstream in(path);
	if (!in.is_open()) {
		std::cout << "file is't open" << std::endl;	 // as read_data does (read_write_utils.hpp:15-19)
		throw 1;
	}
	std::vector<std::array<double, 4>> rows;
	std::string line;
	std::getline(in, line);	 // header
	while (std::getline(in, line)) {
		std::array<double, 4> r{};
		const char* p = line.c_str();
		int k = 0;
		for (; k < 4; ++k) {
			char* end = nullptr;
			r[k] = std::strtod(p, &end);
			if (end == p) break;
			p = *end == ';' ? end + 1 : end;
		}
		if (k == 4) rows.push_back(r);
	}
	return rows;
}

// ---- asynchronous trajectory sink (K4, host half) -------------------------------------------
// The integrator thread hands (time, snapshot ticket) pairs over; this thread waits for
// the pinned copy, formats the row and releases the ring slot, so text formatting and
// file I/O stay off the GPU's critical path.
template <typename T>
class TrajectoryWriter {
   public:
	TrajectoryWriter(nbody_ctx* ctx, TrajectorySinks<T>* out, std::size_t n)
		: ctx_(ctx), out_(out), n_(n), th_([this] { run(); }) {}
	~TrajectoryWriter() { finish(); }
	// blocks while all ring slots are in flight (back-pressure)
	void push(double time) {
		{
			std::unique_lock<std::mutex> lk(mu_);
			cv_.wait(lk, [&] { return q_.size() < (std::size_t)NBODY_SNAPSHOT_SLOTS; });
		}
		uint64_t ticket = 0;
		if (nbody_snapshot_async(ctx_, &ticket) != NBODY_OK) {
			std::cout << "snapshot failed: " << nbody_last_error(ctx_) << std::endl;
			return;
		}
		std::lock_guard<std::mutex> lk(mu_);
		q_.push_back({time, ticket});
		cv_.notify_all();
	}
	void finish() {
		{
			std::lock_guard<std::mutex> lk(mu_);
			if (done_) return;
			done_ = true;
			cv_.notify_all();
		}
		th_.join();
	}
	double busy_seconds() const { return busy_; }

   private:
	struct Item { double time; uint64_t ticket; };
	void run() {
		for (;;) {
			Item it;
			{
				std::unique_lock<std::mutex> lk(mu_);
				cv_.wait(lk, [&] { return done_ || !q_.empty(); });
				if (q_.empty()) return;
				it = q_.front();
			}
			const void* p = nullptr;
			if (nbody_snapshot_wait(ctx_, it.ticket, &p) == NBODY_OK) {
				auto t0 = std::chrono::steady_clock::now();
				out_->row(it.time, (const T*)p);
				busy_ += std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
			}
			nbody_snapshot_release(ctx_, it.ticket);
			{
				std::lock_guard<std::mutex> lk(mu_);
				q_.pop_front();
				cv_.notify_all();
			}
		}
	}
	nbody_ctx* ctx_;
	TrajectorySinks<T>* out_;
	std::size_t n_;
	std::mutex mu_;
	std::condition_variable cv_;
	std::deque<Item> q_;
	bool done_ = false;
	double busy_ = 0;
	std::thread th_;
};

struct RunOptions {
	int scheme = NBODY_RK2;	 // the reference's scheme; NBODY_RK4 is the extension
	int n_gpus = 1;
	std::string flavour = "lab4";  // lab4: t_start row + 4-body header; lab5: N-body header, no t_start row
	std::string traj_path;		   // "" = no trajectory (as for "4Body_random")
	std::string traj_bin_path;	   // "" = no binary sidecar
	std::string per_body_dir;	   // "" = no per-body traj<k>.csv files
	std::size_t per_body_first = 0, per_body_count = 0;	 // bodies [first, first + count); 0 = to the end
	bool keep_rows = false;		   // keep every trajectory row in Solver_N_body::kept_rows
	std::string timing_path;	   // "" = derive as the reference does
	std::string final_state_path;
	long out_every = 1;
	long max_steps = 0;
	long variant = -1;
	std::string tableau;	 // "" = `scheme`; else a name from nbody_tableaux.hpp (nbody_step_rk)
	double adaptive_rtol = 0;  // > 0: step-size control over the embedded pair `tableau`, tau = first step
	double adaptive_atol = 0;
	long pair_symmetry = -1;  // -1 auto (library default), 0 one-sided bit-reproducible kernel, 1 pair-symmetric
	long deterministic = 0;	  // 1: fixed-point row sums in the pair-symmetric kernel (bit-reproducible)
	bool report = false;
};

struct RunResult {
	long steps = 0, rejected = 0;
	double wall_s = 0, loop_s = 0, gpu_ms = 0, writer_busy_s = 0;  // loop_s: until the GPU is idle;
	// wall_s additionally includes draining the trajectory writer (the last row's formatting)
	double interactions = 0;
	double e_kin0 = 0, e_pot0 = 0, e_kin1 = 0, e_pot1 = 0;
};

template <typename T>
class Solver_N_body {
   public:
	T tol = T(1e-6);
	bool notifications = false;
	T timer = 0;
	std::size_t iter = 0;
	std::string file_name;
	std::string output_folder = "../output/";
	std::vector<Body<T>> final_state;
	RunResult result;
	std::vector<std::vector<double>> kept_rows;	 // RunOptions::keep_rows: {t, x1, y1, z1, ...} per row

	Solver_N_body() = default;
	Solver_N_body(T tol_, std::string output_folder_, std::string file_name_)
		: tol(tol_), file_name(std::move(file_name_)), output_folder(std::move(output_folder_)) {}

	// Counterpart of solve_parallel_MPI(sys, OS): one thread per GPU plays the MPI rank.
	void solve_parallel_CUDA(ode_system<T>& sys, const RunOptions& opt, std::string OS = "PC") {
		(void)OS;
		if (notifications) std::cout << file_name << ": solving... \t";
		const std::size_t n = sys.traj_init.size();
		const int world = opt.n_gpus;
		if ((std::size_t)world > n)	 // class_Solver_N_body.hpp:115-117
			throw "The number of processes is greater than the problem dimension";
		// Everything that can fail without a GPU call is checked BEFORE ranks are spawned: a rank
		// that returns early would leave its peers waiting in the communicator set-up forever.
		if (world < 1) throw std::runtime_error("--gpus must be >= 1");
		if (world > 1 && world > nbody_device_count())	// (one rank: nbody_create reports it)
			throw std::runtime_error("--gpus " + std::to_string(world) + " but only " +
									 std::to_string(nbody_device_count()) +
									 " sm_100 device(s) are visible (no CPU fallback)");
		if (!opt.tableau.empty()) {
			const nbody_host::Tableau* tab = nbody_host::find_tableau(opt.tableau);
			if (!tab) throw std::runtime_error("unknown tableau " + opt.tableau);
			if (opt.adaptive_rtol > 0 && !tab->embedded)
				throw std::runtime_error("tableau " + opt.tableau + " has no embedded companion");
		}
		char uid[NBODY_NCCL_UID_BYTES] = {};
		if (world > 1 && nbody_nccl_unique_id(uid) != NBODY_OK)
			throw std::runtime_error(std::string("NCCL: ") + nbody_last_error(nullptr));
		std::vector<std::string> errors(world);
		std::vector<std::thread> ranks;
		for (int r = 1; r < world; ++r)
			ranks.emplace_back([&, r] { rank_main(sys, opt, r, world, uid, errors[r]); });
		rank_main(sys, opt, 0, world, uid, errors[0]);
		for (auto& t : ranks) t.join();
		for (auto& e : errors)
			if (!e.empty()) throw std::runtime_error(e);
		if (notifications) std::cout << "  Done!\n";
	}

   private:
	// Several ranks, one of them fails: its peers may already be blocked in a collective that it
	// will never join, so the process has to end (a join() would hang).  What rank 0 has written so
	// far must not be lost with it: rank 0 registers how to flush its open sinks, and whichever rank
	// ends the process runs that first (std::_Exit itself flushes nothing and runs no destructors).
	std::mutex abort_mu_;
	std::function<void()> abort_flush_;
	struct AbortFlushScope {  // registration that ends with rank 0's sinks
		Solver_N_body* s;
		AbortFlushScope(Solver_N_body* s_, std::function<void()> f) : s(s_) {
			std::lock_guard<std::mutex> g(s->abort_mu_);
			s->abort_flush_ = std::move(f);
		}
		~AbortFlushScope() {
			std::lock_guard<std::mutex> g(s->abort_mu_);
			s->abort_flush_ = nullptr;
		}
	};
	[[noreturn]] void abort_world(int rank, const std::string& error) {
		std::cerr << "rank " << rank << ": " << error << std::endl;
		{
			std::lock_guard<std::mutex> g(abort_mu_);
			if (abort_flush_) abort_flush_();
		}
		std::cout.flush();
		std::fflush(nullptr);
		std::_Exit(3);
	}

	void rank_main(ode_system<T>& sys, const RunOptions& opt, int rank, int world,
				   const char* uid, std::string& error) {
		const std::size_t n = sys.traj_init.size();
		nbody_ctx* ctx = nullptr;
		const int prec = std::is_same<T, double>::value ? 64 : 32;
		if (nbody_create(&ctx, n, prec, rank, rank, world, world > 1 ? uid : nullptr) != NBODY_OK) {
			error = std::string("nbody_create: ") + nbody_last_error(nullptr);
			if (world > 1) abort_world(rank, error);  // peers are waiting in ncclCommInitRank
			return;
		}
		auto fail = [&](const char* what) {
			error = std::string(what) + ": " + nbody_last_error(ctx);
			if (world > 1) abort_world(rank, error);  // never leave the peers waiting (see abort_world)
			nbody_destroy(ctx);
		};
		if (nbody_set_params(ctx, (double)sys.G, (double)sys.eps)) return fail("set_params");
		if (opt.variant >= 0 && nbody_set_option(ctx, "variant", opt.variant)) return fail("variant");
		if (nbody_set_option(ctx, "pair_symmetry", opt.pair_symmetry)) return fail("pair_symmetry");
		if (nbody_set_option(ctx, "deterministic", opt.deterministic)) return fail("deterministic");
		if (nbody_upload_aos(ctx, sys.traj_init.data())) return fail("upload");

		std::ofstream traj;
		TrajectorySinks<T> sinks;
		sinks.n = n;
		if (rank == 0 && !opt.traj_path.empty()) {
			traj.open(opt.traj_path);
			if (!traj.is_open()) std::cout << "Could not open file for writing" << std::endl;
			if (opt.flavour == "lab5") traj << lab5_header(n);
			else traj << lab4_header();
			sinks.csv = &traj;
		}
		auto fail_io = [&](const std::string& why) {	// as `fail`: never leave the peers waiting
			error = why;
			if (world > 1) abort_world(rank, error);
			nbody_destroy(ctx);
		};
		std::unique_ptr<AbortFlushScope> abort_scope;
		if (rank == 0 && world > 1)
			abort_scope.reset(new AbortFlushScope(this, [&traj, &sinks] {
				if (traj.is_open()) traj.flush();
				sinks.finish();
			}));
		if (rank == 0 && !opt.traj_bin_path.empty() && !sinks.open_bin(opt.traj_bin_path, n, opt.flavour))
			return fail_io("cannot open " + opt.traj_bin_path);
		std::string why;
		if (rank == 0 && !opt.per_body_dir.empty() &&
			!sinks.per_body.open(opt.per_body_dir, n, opt.per_body_first, opt.per_body_count, &why))
			return fail_io(why);
		if (rank == 0 && opt.keep_rows) {
			kept_rows.clear();
			sinks.keep = &kept_rows;
		}
		const bool writes = rank == 0 && sinks.any();
		if (writes && opt.flavour != "lab5") {	// the t_start row (ode_RK2_MPI.hpp:63)
			std::vector<T> pos0(3 * n);
			for (std::size_t i = 0; i < n; ++i)
				for (int k = 0; k < 3; ++k) pos0[3 * i + k] = sys.traj_init[i].r[k];
			sinks.row((double)sys.t_start, pos0.data());
		}
		if (opt.report && nbody_energy(ctx, &result_slot(rank).e_kin0, &result_slot(rank).e_pot0))
			return fail("energy");

		RunResult& res = result_slot(rank);
		{
			std::unique_ptr<TrajectoryWriter<T>> writer;
			if (writes) writer.reset(new TrajectoryWriter<T>(ctx, &sinks, n));
			const auto w0 = std::chrono::steady_clock::now();
			// The reference's loop, in T: ode_RK2_MPI.hpp:65,72,98-99
			T t = sys.t_start + sys.tau;
			long steps = 0;
			const nbody_host::Tableau* tab = nullptr;
			std::vector<double> tab_a;
			if (!opt.tableau.empty()) {
				tab = nbody_host::find_tableau(opt.tableau);
				if (!tab) { error = "unknown tableau " + opt.tableau; nbody_destroy(ctx); return; }
				if (opt.adaptive_rtol > 0 && !tab->embedded) {
					error = "tableau " + opt.tableau + " has no embedded companion"; nbody_destroy(ctx); return;
				}
				tab_a = tab->flat_a();
			}
			if (tab && opt.adaptive_rtol > 0) {
				// Step-size control (not in the reference): rows are written at the accepted times.
				// Every rank sees the same all-reduced norms, so all take the same decisions.
				double time = (double)sys.t_start, tau = (double)sys.tau;
				const int q = tab->order < tab->err_order ? tab->order : tab->err_order;
				const long cap = opt.max_steps > 0 ? opt.max_steps : 100000000L;
				while (time < (double)sys.t_final && steps < cap) {
					if (tau > (double)sys.t_final - time) tau = (double)sys.t_final - time;
					double nrm[4];
					if (nbody_step_embedded(ctx, tab->stages, tab_a.data(), tab->b, tab->b_err, tau, nrm))
						return fail("step_embedded");
					const double er = nrm[0] / (opt.adaptive_atol + opt.adaptive_rtol * nrm[2]);
					const double ev = nrm[1] / (opt.adaptive_atol + opt.adaptive_rtol * nrm[3]);
					const double err = er > ev ? er : ev;
					if (!(err == err)) { error = "non-finite error estimate"; nbody_destroy(ctx); return; }
					if (err <= 1.0) {
						time += tau;
						++steps;
						if (writer && (steps % opt.out_every == 0)) writer->push(time);
					} else {
						if (nbody_step_undo(ctx)) return fail("step_undo");
						++res.rejected;
					}
					tau *= nbody_host::controller_factor(err, q);
				}
			} else
			while (true) {
				if (tab ? nbody_step_rk(ctx, tab->stages, tab_a.data(), tab->b, (double)sys.tau, 1)
						: nbody_step(ctx, opt.scheme, (double)sys.tau, 1))
					return fail("step");
				++steps;
				if (writer && (steps % opt.out_every == 0)) writer->push((double)t);
				if (opt.max_steps > 0 && steps >= opt.max_steps) break;
				if (t > sys.t_final) break;
				t += sys.tau;
			}
			if (nbody_sync(ctx)) return fail("sync");
			res.loop_s = std::chrono::duration<double>(std::chrono::steady_clock::now() - w0).count();
			if (writer) {
				writer->finish();
				sinks.finish();
				res.writer_busy_s = writer->busy_seconds();
			}
			res.wall_s = std::chrono::duration<double>(std::chrono::steady_clock::now() - w0).count();
			res.steps = steps;
			const int stages = tab ? tab->stages : (opt.scheme == NBODY_RK2 ? 2 : 4);
			res.interactions = (double)stages * (double)n * (double)n * (double)(steps + res.rejected);
		}
		if (opt.report && nbody_energy(ctx, &res.e_kin1, &res.e_pot1)) return fail("energy");
		std::vector<Body<T>> fin(n);
		if (nbody_download_aos(ctx, fin.data())) return fail("download");
		if (rank == 0) {
			final_state.swap(fin);
			timer = (T)res.wall_s;
			iter = (std::size_t)res.steps;
			result = res;
			if (!opt.final_state_path.empty()) write_input_format(opt.final_state_path, final_state);
			write_timing_csv(sys, opt, world);
		}
		nbody_destroy(ctx);
	}

	RunResult& result_slot(int rank) {
		static thread_local RunResult other;
		return rank == 0 ? result : other;
	}

	void write_timing_csv(const ode_system<T>& sys, const RunOptions& opt, int world) {
		std::string path = opt.timing_path;
		if (path.empty())  // class_Solver_N_body.hpp:98-108
			path = output_folder + "/output_" + sys.name + "_" + std::to_string(sys.tau) + "_np_" +
				   std::to_string(world) + ".csv";
		std::ofstream out(path);
		if (!out.is_open()) return;
		if (opt.flavour == "lab5") {  // N_body.cu:226,397-399
			const int n_steps = (int)((sys.t_final - sys.t_start) / sys.tau);
			out << "type;n_bodies;n_steps;block_size;time_taken;avg_exec_time\n";
			out << (std::is_same<T, double>::value ? "double" : "float") << ";"
				<< sys.traj_init.size() << ";" << n_steps << ";" << 128 << ";" << result.wall_s
				<< ";" << result.wall_s / (n_steps > 0 ? n_steps : 1) << "\n";
		} else {  // class_Solver_N_body.hpp:110-111,199-202
			out << "size;t_start;t_finish;tau;time_steps;num_procs;time_taken;avg_time_for_step\n";
			out << sys.traj_init.size() << ";" << sys.t_start << ";" << sys.t_final << ";" << sys.tau
				<< ";" << sys.n_time_steps << ";" << world << ";" << timer << ";"
				<< timer / sys.n_time_steps << "\n";
		}
	}
};

// lab_5's entry point with its own signature (N_body.cu:208-210): SoA host vectors in,
// timing CSV at `path`, trajectory CSV at `path_1` (lab_5 flavour).
template <typename T>
void ode_RK2_CUDA(const std::vector<T>& m, const std::vector<Vec4<T>>& r,
				  const std::vector<Vec4<T>>& v, T t_start, T t_final, T tau,
				  std::string path = "", std::string path_1 = "", T eps = T(1e-6),
				  T G = T(6.67e-11)) {
	ode_system<T> sys;
	sys.name = "lab5";
	sys.t_start = t_start;
	sys.t_final = t_final;
	sys.tau = tau;
	sys.n_time_steps = (int)((t_final - t_start) / tau);
	sys.eps = eps;
	sys.G = G;
	sys.traj_init.resize(m.size());
	for (std::size_t i = 0; i < m.size(); ++i) {
		sys.traj_init[i].m = m[i];
		sys.traj_init[i].r = {r[i].x, r[i].y, r[i].z};
		sys.traj_init[i].v = {v[i].x, v[i].y, v[i].z};
	}
	RunOptions opt;
	opt.flavour = "lab5";
	opt.scheme = NBODY_RK2;
	opt.timing_path = path.empty() ? "/dev/null" : path;
	opt.traj_path = path_1;
	Solver_N_body<T> solver;
	solver.solve_parallel_CUDA(sys, opt);
}

}  // namespace nbody_host
