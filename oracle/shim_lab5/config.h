/* config.h -- lab_5's git-ignored path configuration, supplied by the oracle
 * build.  TEST INFRASTRUCTURE ONLY.  Names used at lab_5_N_body_CUDA/src/N_body.cu:428-435.
 * The root is taken from $NBODY_REF5_SCRATCH at start-up (default /tmp/nbody_ref5) so that the
 * reference's unmodified main() can be pointed at a test's scratch directory. */
#pragma once
#include <cstdlib>
#include <string>
namespace config_lab_5_CUDA {
inline std::string root_() {
	const char* e = std::getenv("NBODY_REF5_SCRATCH");
	return e ? e : "/tmp/nbody_ref5";
}
inline std::string PATH_main_folder = root_();
inline std::string PATH_src_folder = root_() + "/src";
inline std::string PATH_output_folder = root_() + "/output";
}  // namespace config_lab_5_CUDA
