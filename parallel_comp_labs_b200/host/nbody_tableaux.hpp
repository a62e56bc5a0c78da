// nbody_tableaux.hpp -- named Butcher tableaux for nbody_step_rk / nbody_step_embedded and the
// step-size controller of the host driver ("adaptive/other tableaux
// (RK3/RK45)").  Not in the reference, whose only scheme is the explicit midpoint rule
// (lab_4 .../schemes/ode_RK2_MPI.hpp:72-92) -- "midpoint" below.  Same coefficients as
// parallel_comp_labs_b200/tableaux.py (the CPU tests compare the two tables).
#pragma once
#include <cmath>
#include <cstring>
#include <string>
#include <vector>

#include "nbody_b200.h"

namespace nbody_host {

struct Tableau {
	const char* name;
	int stages, order, err_order;
	bool embedded;
	double a[NBODY_MAX_STAGES][NBODY_MAX_STAGES];  // strictly lower triangle
	double b[NBODY_MAX_STAGES];
	double b_err[NBODY_MAX_STAGES];
	// a as the stages x stages row-major array the C ABI takes
	std::vector<double> flat_a() const {
		std::vector<double> f((size_t)stages * stages, 0.0);
		for (int i = 0; i < stages; ++i)
			for (int j = 0; j < i; ++j) f[(size_t)i * stages + j] = a[i][j];
		return f;
	}
};

inline const std::vector<Tableau>& tableaux() {
	static const std::vector<Tableau> t = {
		{"euler", 1, 1, 0, false,
			{{0.0}},
			{1.0},
			{}},
		{"midpoint", 2, 2, 0, false,
			{{0.0, 0.0},
			 {1.0 / 2.0, 0.0}},
			{0.0, 1.0},
			{}},
		{"heun", 2, 2, 0, false,
			{{0.0, 0.0},
			 {1.0, 0.0}},
			{1.0 / 2.0, 1.0 / 2.0},
			{}},
		{"ralston", 2, 2, 0, false,
			{{0.0, 0.0},
			 {2.0 / 3.0, 0.0}},
			{1.0 / 4.0, 3.0 / 4.0},
			{}},
		{"kutta3", 3, 3, 0, false,
			{{0.0, 0.0, 0.0},
			 {1.0 / 2.0, 0.0, 0.0},
			 {-1.0, 2.0, 0.0}},
			{1.0 / 6.0, 2.0 / 3.0, 1.0 / 6.0},
			{}},
		{"rk4", 4, 4, 0, false,
			{{0.0, 0.0, 0.0, 0.0},
			 {1.0 / 2.0, 0.0, 0.0, 0.0},
			 {0.0, 1.0 / 2.0, 0.0, 0.0},
			 {0.0, 0.0, 1.0, 0.0}},
			{1.0 / 6.0, 1.0 / 3.0, 1.0 / 3.0, 1.0 / 6.0},
			{}},
		{"rk38", 4, 4, 0, false,
			{{0.0, 0.0, 0.0, 0.0},
			 {1.0 / 3.0, 0.0, 0.0, 0.0},
			 {-1.0 / 3.0, 1.0, 0.0, 0.0},
			 {1.0, -1.0, 1.0, 0.0}},
			{1.0 / 8.0, 3.0 / 8.0, 3.0 / 8.0, 1.0 / 8.0},
			{}},
		{"bs23", 4, 3, 2, true,
			{{0.0, 0.0, 0.0, 0.0},
			 {1.0 / 2.0, 0.0, 0.0, 0.0},
			 {0.0, 3.0 / 4.0, 0.0, 0.0},
			 {2.0 / 9.0, 1.0 / 3.0, 4.0 / 9.0, 0.0}},
			{2.0 / 9.0, 1.0 / 3.0, 4.0 / 9.0, 0.0},
			{7.0 / 24.0, 1.0 / 4.0, 1.0 / 3.0, 1.0 / 8.0}},
		{"rkf45", 6, 4, 5, true,
			{{0.0, 0.0, 0.0, 0.0, 0.0, 0.0},
			 {1.0 / 4.0, 0.0, 0.0, 0.0, 0.0, 0.0},
			 {3.0 / 32.0, 9.0 / 32.0, 0.0, 0.0, 0.0, 0.0},
			 {1932.0 / 2197.0, -7200.0 / 2197.0, 7296.0 / 2197.0, 0.0, 0.0, 0.0},
			 {439.0 / 216.0, -8.0, 3680.0 / 513.0, -845.0 / 4104.0, 0.0, 0.0},
			 {-8.0 / 27.0, 2.0, -3544.0 / 2565.0, 1859.0 / 4104.0, -11.0 / 40.0, 0.0}},
			{25.0 / 216.0, 0.0, 1408.0 / 2565.0, 2197.0 / 4104.0, -1.0 / 5.0, 0.0},
			{16.0 / 135.0, 0.0, 6656.0 / 12825.0, 28561.0 / 56430.0, -9.0 / 50.0, 2.0 / 55.0}},
		{"cashkarp", 6, 5, 4, true,
			{{0.0, 0.0, 0.0, 0.0, 0.0, 0.0},
			 {1.0 / 5.0, 0.0, 0.0, 0.0, 0.0, 0.0},
			 {3.0 / 40.0, 9.0 / 40.0, 0.0, 0.0, 0.0, 0.0},
			 {3.0 / 10.0, -9.0 / 10.0, 6.0 / 5.0, 0.0, 0.0, 0.0},
			 {-11.0 / 54.0, 5.0 / 2.0, -70.0 / 27.0, 35.0 / 27.0, 0.0, 0.0},
			 {1631.0 / 55296.0, 175.0 / 512.0, 575.0 / 13824.0, 44275.0 / 110592.0, 253.0 / 4096.0, 0.0}},
			{37.0 / 378.0, 0.0, 250.0 / 621.0, 125.0 / 594.0, 0.0, 512.0 / 1771.0},
			{2825.0 / 27648.0, 0.0, 18575.0 / 48384.0, 13525.0 / 55296.0, 277.0 / 14336.0, 1.0 / 4.0}},
		{"dopri5", 7, 5, 4, true,
			{{0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0},
			 {1.0 / 5.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0},
			 {3.0 / 40.0, 9.0 / 40.0, 0.0, 0.0, 0.0, 0.0, 0.0},
			 {44.0 / 45.0, -56.0 / 15.0, 32.0 / 9.0, 0.0, 0.0, 0.0, 0.0},
			 {19372.0 / 6561.0, -25360.0 / 2187.0, 64448.0 / 6561.0, -212.0 / 729.0, 0.0, 0.0, 0.0},
			 {9017.0 / 3168.0, -355.0 / 33.0, 46732.0 / 5247.0, 49.0 / 176.0, -5103.0 / 18656.0, 0.0, 0.0},
			 {35.0 / 384.0, 0.0, 500.0 / 1113.0, 125.0 / 192.0, -2187.0 / 6784.0, 11.0 / 84.0, 0.0}},
			{35.0 / 384.0, 0.0, 500.0 / 1113.0, 125.0 / 192.0, -2187.0 / 6784.0, 11.0 / 84.0, 0.0},
			{5179.0 / 57600.0, 0.0, 7571.0 / 16695.0, 393.0 / 640.0, -92097.0 / 339200.0, 187.0 / 2100.0, 1.0 / 40.0}},
	};
	return t;
}

inline const Tableau* find_tableau(const std::string& name) {
	for (const Tableau& t : tableaux())
		if (name == t.name) return &t;
	return nullptr;
}

// tau_new / tau of the standard controller for an embedded pair: err is the scaled error norm
// (<= 1 accepts), q = min(order, err_order).
inline double controller_factor(double err, int q, double safety = 0.9) {
	const double e = err > 0 ? err : 1e-16;
	return std::fmin(5.0, std::fmax(0.2, safety * std::pow(e, -1.0 / (q + 1))));
}

}  // namespace nbody_host
