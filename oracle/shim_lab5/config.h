/* config.h -- lab_5's git-ignored path configuration, supplied by the oracle
 * build.  TEST INFRASTRUCTURE ONLY.  Names used at lab_5_N_body_CUDA/src/N_body.cu:428-435. */
#pragma once
#include <string>
namespace config_lab_5_CUDA {
inline std::string PATH_main_folder = "/tmp/nbody_ref5";
inline std::string PATH_src_folder = "/tmp/nbody_ref5/src";
inline std::string PATH_output_folder = "/tmp/nbody_ref5/output";
}  // namespace config_lab_5_CUDA
