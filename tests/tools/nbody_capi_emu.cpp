// nbody_capi_emu.cpp -- TEST INFRASTRUCTURE: the whole C ABI (parallel_comp_labs_b200/csrc/
// nbody_capi.cu, unmodified) built by g++ against the stand-in CUDA runtime (cuda_rt_emu.h) and
// the kernel emulation (cuda_emu.h).  tests/test_capi_emulation.py loads the result with ctypes
// and runs single-rank parity cases through it, so the library's HOST logic and kernel SOURCE are
// executed by the `-m "not gpu"` suite.  The product (libnbody_b200.so) is never built this way
// and has no CPU path; nothing outside tests/ refers to this file.
#include "cuda_rt_emu.h"

#include "../../parallel_comp_labs_b200/csrc/nbody_capi.cu"
