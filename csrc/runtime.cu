// Runtime services of libb200ddl: error reporting, device info, symmetric-memory
// allocation with CUDA IPC handle export/import (target layer T1).
// No counterpart in the reference (it has no native code).
#include <stdarg.h>
#include <string.h>

#include "../include/b200ddl.h"
#include "common.cuh"

static thread_local char g_err[1024] = "";

void ddl_set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

DDL_API const char* ddl_last_error(void) { return g_err; }

DDL_API int ddl_version(void) { return 100; }

DDL_API int ddl_device_info(int device, int* sm_count, int* cc_major, int* cc_minor, size_t* total_mem) {
    cudaDeviceProp p;
    DDL_CHECK(cudaGetDeviceProperties(&p, device));
    if (sm_count) *sm_count = p.multiProcessorCount;
    if (cc_major) *cc_major = p.major;
    if (cc_minor) *cc_minor = p.minor;
    if (total_mem) *total_mem = p.totalGlobalMem;
    return 0;
}

// ---- symmetric heap --------------------------------------------------------------------
// Every rank allocates one slab with cudaMalloc, exports a 64-byte IPC handle, and opens
// every peer's handle.  The Python side (parallel/symm.py) moves the handles between
// processes over the bootstrap store; nothing here talks to NCCL.

DDL_API int ddl_symm_alloc(size_t bytes, void** ptr) {
    DDL_REQUIRE(ptr != nullptr && bytes > 0, "ddl_symm_alloc: bad arguments");
    DDL_CHECK(cudaMalloc(ptr, bytes));
    DDL_CHECK(cudaMemset(*ptr, 0, bytes));
    DDL_CHECK(cudaDeviceSynchronize());
    return 0;
}

DDL_API int ddl_symm_free(void* ptr) {
    DDL_CHECK(cudaFree(ptr));
    return 0;
}

DDL_API int ddl_ipc_export(void* ptr, unsigned char* handle64) {
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    cudaIpcMemHandle_t h;
    DDL_CHECK(cudaIpcGetMemHandle(&h, ptr));
    memcpy(handle64, &h, 64);
    return 0;
}

DDL_API int ddl_ipc_open(const unsigned char* handle64, void** ptr) {
    cudaIpcMemHandle_t h;
    memcpy(&h, handle64, 64);
    DDL_CHECK(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess));
    return 0;
}

DDL_API int ddl_ipc_close(void* ptr) {
    DDL_CHECK(cudaIpcCloseMemHandle(ptr));
    return 0;
}

DDL_API int ddl_can_access_peer(int device, int peer, int* ok) {
    DDL_CHECK(cudaDeviceCanAccessPeer(ok, device, peer));
    return 0;
}

DDL_API int ddl_stream_sync(void* stream) {
    DDL_CHECK(cudaStreamSynchronize((cudaStream_t)stream));
    return 0;
}
