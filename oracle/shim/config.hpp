/*
 * config.hpp -- the reference's git-ignored path configuration,
 * supplied by the oracle build.  TEST INFRASTRUCTURE ONLY.
 * Shape: lab_3_Helmholtz_MPI/src/config.hpp:7-18; names used at
 * lab_4_N_body_MPI/src/N_body.cpp:64 and
 * lab_4_N_body_MPI/src/include/N_body/struct_ode_system.hpp:101-102,
 * class_Solver_N_body.hpp:104.
 * The strings are run-time settable (inline variables) so that the oracle
 * driver can point them at a scratch directory.
 */
#pragma once
#include <string>
namespace config_lab_4_MPI {
inline std::string PATH_main_folder = "/tmp/nbody_ref";
inline std::string PATH_src_folder = "/tmp/nbody_ref/src";
inline std::string PATH_output_folder = "/tmp/nbody_ref/output";
}  // namespace config_lab_4_MPI
