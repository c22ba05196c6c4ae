// np_rt.h — thin runtime layer: device memory, streams, events and the kernel-launch macro.
//
// Product build (nvcc): everything maps 1:1 onto the CUDA runtime; there is no other backend.
// NP_EMU build (g++, tests only, see tests/emu/cuda_emu.h): the same kernels and the same host driver
// run on a host emulation of blocks/threads so that kernel logic can be checked without a GPU.
#pragma once
#include <cstddef>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#ifdef NP_EMU
#include "cuda_emu.h"
#include <chrono>
typedef void* np_stream_t;
struct np_event_s { std::chrono::steady_clock::time_point t; };
typedef np_event_s* np_event_t;
#define NP_LAUNCH(sync, kernel, grid, block, smem, stream, ...) \
    np_emu::launch((grid), (block), (smem), (sync), [=]() { kernel(__VA_ARGS__); })
#define NP_DYN_SMEM(type, name) type* name = reinterpret_cast<type*>((reinterpret_cast<uintptr_t>(np_emu::st().dyn.data()) + 15) & ~uintptr_t(15))
namespace np_rt {
inline int init() { return 0; }
inline int dev_malloc(void** p, size_t n) { *p = std::calloc(n ? n : 1, 1); return *p ? 0 : 1; }
inline int dev_free(void* p) { std::free(p); return 0; }
inline int host_malloc(void** p, size_t n) { *p = std::malloc(n ? n : 1); return *p ? 0 : 1; }
inline int host_free(void* p) { std::free(p); return 0; }
inline int memset_async(void* p, int v, size_t n, np_stream_t) { std::memset(p, v, n); return 0; }
inline int copy_async(void* d, const void* s, size_t n, np_stream_t) { std::memmove(d, s, n); return 0; }
inline int stream_sync(np_stream_t) { return 0; }
inline int stream_create(np_stream_t* s) { *s = nullptr; return 0; }
inline int stream_create_high_priority(np_stream_t* s) { *s = nullptr; return 0; }
inline int stream_destroy(np_stream_t) { return 0; }
inline int stream_wait_event(np_stream_t, np_event_t) { return 0; }     // the emulation runs everything in program order
inline int ipc_export(void*, void*) { return 1; }                         // no peer memory in the emulation: NCCL-style callbacks only
inline int ipc_open(const void*, void**) { return 1; }
inline int ipc_close(void*) { return 0; }
inline int is_device_ptr(const void*) { return 1; }
inline int event_create(np_event_t* e) { *e = new np_event_s(); return 0; }
inline int event_destroy(np_event_t e) { delete e; return 0; }
inline int event_record(np_event_t e, np_stream_t) { e->t = std::chrono::steady_clock::now(); return 0; }
inline float event_ms(np_event_t a, np_event_t b) { return std::chrono::duration<float, std::milli>(b->t - a->t).count(); }
inline int last_error() { return 0; }
inline const char* last_error_string() { return "emu"; }
inline int sm_count() { return 4; }
// CUDA graphs do not exist in the emulation: every call fails and the solver keeps launching kernel by kernel
typedef void* np_graph_t;
inline int capture_begin(np_stream_t) { return 1; }
inline int capture_end(np_stream_t, np_graph_t*) { return 1; }
inline int graph_launch(np_graph_t, np_stream_t) { return 1; }
inline int graph_destroy(np_graph_t) { return 0; }
inline int event_record_in_graph(np_event_t e, np_stream_t s) { return event_record(e, s); }
struct Range { explicit Range(const char*) {} };          // NVTX range (no-op here)
} // namespace np_rt
#else
#include <cuda_runtime.h>
#include <nvtx3/nvToolsExt.h>
typedef cudaStream_t np_stream_t;
typedef cudaEvent_t np_event_t;
#define NP_LAUNCH(sync, kernel, grid, block, smem, stream, ...) \
    kernel<<<(grid), (block), (smem), (stream)>>>(__VA_ARGS__)
#define NP_DYN_SMEM(type, name) extern __shared__ __align__(16) unsigned char name##_raw_[]; type* name = reinterpret_cast<type*>(name##_raw_)
namespace np_rt {
inline int init() {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) { cudaGetLastError(); return 1; }
    return 0;
}
inline int dev_malloc(void** p, size_t n) { return cudaMalloc(p, n ? n : 1) == cudaSuccess ? 0 : 1; }
inline int dev_free(void* p) { return cudaFree(p) == cudaSuccess ? 0 : 1; }
inline int host_malloc(void** p, size_t n) { return cudaMallocHost(p, n ? n : 1) == cudaSuccess ? 0 : 1; }
inline int host_free(void* p) { return cudaFreeHost(p) == cudaSuccess ? 0 : 1; }
inline int memset_async(void* p, int v, size_t n, np_stream_t s) { return cudaMemsetAsync(p, v, n, s) == cudaSuccess ? 0 : 1; }
inline int copy_async(void* d, const void* s, size_t n, np_stream_t st) { return cudaMemcpyAsync(d, s, n, cudaMemcpyDefault, st) == cudaSuccess ? 0 : 1; }
inline int stream_sync(np_stream_t s) { return cudaStreamSynchronize(s) == cudaSuccess ? 0 : 1; }
inline int stream_create(np_stream_t* s) { return cudaStreamCreateWithFlags(s, cudaStreamNonBlocking) == cudaSuccess ? 0 : 1; }
inline int stream_create_high_priority(np_stream_t* s) {     // the halo kernels must get SM slots while a big kernel of the main stream runs
    int lo = 0, hi = 0; cudaDeviceGetStreamPriorityRange(&lo, &hi);
    return cudaStreamCreateWithPriority(s, cudaStreamNonBlocking, hi) == cudaSuccess ? 0 : 1;
}
inline int stream_destroy(np_stream_t s) { return cudaStreamDestroy(s) == cudaSuccess ? 0 : 1; }
inline int stream_wait_event(np_stream_t s, np_event_t e) { return cudaStreamWaitEvent(s, e, 0) == cudaSuccess ? 0 : 1; }
// CUDA IPC: one process per GPU, so a neighbour's arrays are reached through exported allocation handles (64 bytes)
inline int ipc_export(void* dev_ptr, void* handle64) {
    cudaIpcMemHandle_t h;
    if (cudaIpcGetMemHandle(&h, dev_ptr) != cudaSuccess) { cudaGetLastError(); return 1; }
    static_assert(sizeof(h) == 64, "cudaIpcMemHandle_t is 64 bytes");
    memcpy(handle64, &h, 64); return 0;
}
inline int ipc_open(const void* handle64, void** dev_ptr) {
    cudaIpcMemHandle_t h; memcpy(&h, handle64, 64);
    if (cudaIpcOpenMemHandle(dev_ptr, h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { cudaGetLastError(); return 1; }
    return 0;
}
inline int ipc_close(void* dev_ptr) { return cudaIpcCloseMemHandle(dev_ptr) == cudaSuccess ? 0 : 1; }
inline int is_device_ptr(const void* p) {
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return 0; }
    return a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged;
}
inline int event_create(np_event_t* e) { return cudaEventCreate(e) == cudaSuccess ? 0 : 1; }
inline int event_destroy(np_event_t e) { return cudaEventDestroy(e) == cudaSuccess ? 0 : 1; }
inline int event_record(np_event_t e, np_stream_t s) { return cudaEventRecord(e, s) == cudaSuccess ? 0 : 1; }
inline float event_ms(np_event_t a, np_event_t b) { float ms = 0; cudaEventSynchronize(b); cudaEventElapsedTime(&ms, a, b); return ms; }
inline int last_error() { return cudaGetLastError() == cudaSuccess ? 0 : 1; }
inline const char* last_error_string() { return cudaGetErrorString(cudaPeekAtLastError()); }
inline int sm_count() { int dev = 0, n = 148; cudaGetDevice(&dev); cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev); return n; }
// CUDA graph of a launch sequence: captured from a private stream (the caller's may be the legacy default stream, which
// cannot be captured), replayed on the caller's
typedef cudaGraphExec_t np_graph_t;
inline int capture_begin(np_stream_t s) { return cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal) == cudaSuccess ? 0 : 1; }
inline int capture_end(np_stream_t s, np_graph_t* out) {
    cudaGraph_t g = nullptr; *out = nullptr;
    if (cudaStreamEndCapture(s, &g) != cudaSuccess || !g) { cudaGetLastError(); return 1; }
    const cudaError_t e = cudaGraphInstantiate(out, g, 0);
    cudaGraphDestroy(g);
    if (e != cudaSuccess) { cudaGetLastError(); *out = nullptr; return 1; }
    return 0;
}
inline int graph_launch(np_graph_t g, np_stream_t s) { return cudaGraphLaunch(g, s) == cudaSuccess ? 0 : 1; }
inline int graph_destroy(np_graph_t g) { return cudaGraphExecDestroy(g) == cudaSuccess ? 0 : 1; }
// inside a capture a plain cudaEventRecord only marks a dependency; this one becomes a node that records the event at every replay
inline int event_record_in_graph(np_event_t e, np_stream_t s) { return cudaEventRecordWithFlags(e, s, cudaEventRecordExternal) == cudaSuccess ? 0 : 1; }
// NVTX range with the reference's BL_PROFILE region names ("MLMG::solve", ...): shows up in a timeline when a profiler is attached,
// costs a few nanoseconds otherwise (header-only NVTX3, no link dependency)
struct Range { explicit Range(const char* name) { nvtxRangePushA(name); } ~Range() { nvtxRangePop(); } Range(const Range&) = delete; };
} // namespace np_rt
#endif
