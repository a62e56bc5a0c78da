// cuda_rt_emu.h -- TEST INFRASTRUCTURE: the slice of the CUDA runtime API that
// parallel_comp_labs_b200/csrc/nbody_capi.cu uses, as plain host code, so that the whole C ABI
// can be built by g++ (tests/tools/nbody_capi_emu.cpp) and its host logic -- launch plans,
// buffer sizing, stage sequencing, snapshot ring, CUDA-graph replay -- runs in the CPU suite with
// the kernels executed by cuda_emu.h.  Eight identical "devices", sm_100-like, 2 SMs each; device memory is host
// memory filled with 0xFF (NaN) so reads of never-written rows show up; streams execute in issue
// order at once; a capturing stream records closures that cudaGraphLaunch replays.
// Never linked into the product, which has no CPU path.
#pragma once
#include <chrono>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <vector>

#include "cuda_emu.h"

typedef int cudaError_t;
enum { cudaSuccess = 0, cudaErrorMemoryAllocation = 2, cudaErrorInvalidValue = 1, cudaErrorNotSupported = 801,
	   cudaErrorLaunchFailure = 719 };
namespace emu_rt {
inline std::string& launch_error() { static thread_local std::string s; return s; }
}
inline const char* cudaGetErrorString(cudaError_t e) {
	return e == cudaSuccess ? "no error" : e == cudaErrorNotSupported ? "not supported by the emulation"
		   : e == cudaErrorMemoryAllocation ? "out of memory"
		   : e == cudaErrorLaunchFailure ? emu_rt::launch_error().c_str() : "invalid value";
}

namespace emu_rt {
typedef std::vector<std::function<void()>> Graph;
struct Stream { Graph* capturing = nullptr; };
struct Event { std::chrono::steady_clock::time_point t; };
// run now, or record into the graph being captured on this stream
inline void enqueue(Stream* s, std::function<void()> fn) {
	if (s && s->capturing) { s->capturing->push_back(std::move(fn)); return; }
	try {
		fn();
	} catch (const std::exception& e) {	 // a deadlock or a violated check inside an emulated kernel
		launch_error() = e.what();
	}
}
}  // namespace emu_rt

typedef emu_rt::Stream* cudaStream_t;
typedef emu_rt::Event* cudaEvent_t;
typedef emu_rt::Graph* cudaGraph_t;
typedef emu_rt::Graph* cudaGraphExec_t;
struct cudaIpcMemHandle_t { char reserved[64]; };
struct cudaDeviceProp { char name[256]; int major, minor, multiProcessorCount; };
enum { cudaStreamNonBlocking = 1, cudaEventDisableTiming = 2, cudaHostAllocDefault = 0,
	   cudaIpcMemLazyEnablePeerAccess = 1, cudaStreamCaptureModeThreadLocal = 1,
	   cudaFuncAttributeMaxDynamicSharedMemorySize = 8 };
enum cudaMemcpyKind { cudaMemcpyHostToDevice, cudaMemcpyDeviceToHost, cudaMemcpyDeviceToDevice, cudaMemcpyDefault };

#define NB_LAUNCH(kern, grid, block, smem, stream, ...) \
	emu_rt::enqueue(stream, [=] { emu::launch(grid, block, smem, [&] { kern(__VA_ARGS__); }); })

inline cudaError_t cudaGetDeviceCount(int* n) { *n = 8; return cudaSuccess; }  // rank threads name devices 0..7
inline cudaError_t cudaGetDeviceProperties(cudaDeviceProp* p, int) {
	memset(p, 0, sizeof(*p));
	strcpy(p->name, "emulated sm_100 (CPU test build)");
	p->major = 10;
	p->multiProcessorCount = 2;
	return cudaSuccess;
}
inline cudaError_t cudaSetDevice(int) { return cudaSuccess; }
inline cudaError_t cudaGetLastError() {
	return emu_rt::launch_error().empty() ? cudaSuccess : cudaErrorLaunchFailure;
}
inline cudaError_t cudaMalloc(void** p, size_t bytes) {
	*p = aligned_alloc(256, (bytes + 255) / 256 * 256 + 256);
	if (!*p) return cudaErrorMemoryAllocation;
	memset(*p, 0xFF, bytes);
	return cudaSuccess;
}
template <typename T> inline cudaError_t cudaMalloc(T** p, size_t bytes) { return cudaMalloc((void**)p, bytes); }
inline cudaError_t cudaFree(void* p) { free(p); return cudaSuccess; }
inline cudaError_t cudaHostAlloc(void** p, size_t bytes, unsigned) {
	*p = aligned_alloc(256, (bytes + 255) / 256 * 256 + 256);
	return *p ? cudaSuccess : cudaErrorMemoryAllocation;
}
inline cudaError_t cudaFreeHost(void* p) { free(p); return cudaSuccess; }
inline cudaError_t cudaMemset(void* p, int v, size_t n) { memset(p, v, n); return cudaSuccess; }
inline cudaError_t cudaMemsetAsync(void* p, int v, size_t n, cudaStream_t s) {
	emu_rt::enqueue(s, [=] { memset(p, v, n); });
	return cudaSuccess;
}
inline cudaError_t cudaMemcpy(void* d, const void* s, size_t n, cudaMemcpyKind) { memmove(d, s, n); return cudaSuccess; }
inline cudaError_t cudaMemcpyAsync(void* d, const void* s, size_t n, cudaMemcpyKind, cudaStream_t st) {
	emu_rt::enqueue(st, [=] { memmove(d, s, n); });
	return cudaSuccess;
}
inline cudaError_t cudaStreamCreateWithFlags(cudaStream_t* s, unsigned) { *s = new emu_rt::Stream(); return cudaSuccess; }
inline cudaError_t cudaStreamDestroy(cudaStream_t s) { delete s; return cudaSuccess; }
inline cudaError_t cudaStreamSynchronize(cudaStream_t) { return cudaSuccess; }
inline cudaError_t cudaStreamWaitEvent(cudaStream_t, cudaEvent_t, unsigned) { return cudaSuccess; }
inline cudaError_t cudaEventCreate(cudaEvent_t* e) { *e = new emu_rt::Event(); return cudaSuccess; }
inline cudaError_t cudaEventCreateWithFlags(cudaEvent_t* e, unsigned) { return cudaEventCreate(e); }
inline cudaError_t cudaEventDestroy(cudaEvent_t e) { delete e; return cudaSuccess; }
inline cudaError_t cudaEventRecord(cudaEvent_t e, cudaStream_t s) {
	if (!(s && s->capturing)) e->t = std::chrono::steady_clock::now();
	return cudaSuccess;
}
inline cudaError_t cudaEventSynchronize(cudaEvent_t) { return cudaSuccess; }
inline cudaError_t cudaEventElapsedTime(float* ms, cudaEvent_t a, cudaEvent_t b) {
	*ms = std::chrono::duration<float, std::milli>(b->t - a->t).count();
	return cudaSuccess;
}
inline cudaError_t cudaFuncSetAttribute(const void*, int, int) { return cudaSuccess; }
// a `static __device__` array is an ordinary host array here: its "device address" is its address
template <typename S> inline cudaError_t cudaGetSymbolAddress(void** p, const S& sym) { *p = (void*)&sym; return cudaSuccess; }
inline cudaError_t cudaOccupancyMaxActiveBlocksPerMultiprocessor(int* n, const void*, int, size_t) {
	*n = 2;
	return cudaSuccess;
}
inline cudaError_t cudaIpcGetMemHandle(cudaIpcMemHandle_t*, void*) { return cudaErrorNotSupported; }
inline cudaError_t cudaIpcOpenMemHandle(void**, cudaIpcMemHandle_t, unsigned) { return cudaErrorNotSupported; }
inline cudaError_t cudaIpcCloseMemHandle(void*) { return cudaErrorNotSupported; }
inline cudaError_t cudaStreamBeginCapture(cudaStream_t s, int) {
	if (s->capturing) return cudaErrorInvalidValue;
	s->capturing = new emu_rt::Graph();
	return cudaSuccess;
}
inline cudaError_t cudaStreamEndCapture(cudaStream_t s, cudaGraph_t* g) {
	if (!s->capturing) return cudaErrorInvalidValue;
	*g = s->capturing;
	s->capturing = nullptr;
	return cudaSuccess;
}
inline cudaError_t cudaGraphInstantiate(cudaGraphExec_t* x, cudaGraph_t g, unsigned) {
	*x = new emu_rt::Graph(*g);
	return cudaSuccess;
}
inline cudaError_t cudaGraphDestroy(cudaGraph_t g) { delete g; return cudaSuccess; }
inline cudaError_t cudaGraphExecDestroy(cudaGraphExec_t g) { delete g; return cudaSuccess; }
inline cudaError_t cudaGraphLaunch(cudaGraphExec_t g, cudaStream_t) {
	for (auto& fn : *g) emu_rt::enqueue(nullptr, fn);
	return cudaGetLastError();
}

// the peer-pull exchange needs several devices: its flag kernels are system-scope PTX and are
// never reached by a single-rank emulated context
namespace nbody {
inline void p2p_signal_kernel(unsigned* const*, int, int, unsigned) {}
inline void p2p_wait_kernel(const unsigned*, int, int, unsigned) {}
}  // namespace nbody
