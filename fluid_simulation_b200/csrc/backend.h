// backend.h -- the narrow seam between the host logic (api.cpp: validation, planning, the
// solver handle) and the device. The product implements it with CUDA (kernels.cu). The
// CPU-only test build in tests/emu implements it with host threads so the same host logic and
// the same warp task bodies can be exercised without a GPU; that build is never shipped.
#pragma once
#include <stddef.h>
#include <stdint.h>

#include "warp_args.h"

namespace fbk {

enum CopyKind { H2D = 0, D2H = 1, D2D = 2 };

const char* backend_name();
int be_set_device(int device);
int be_sm_count();  // SMs of the current device (148 on B200)
int be_malloc(void** p, size_t bytes);
int be_free(void* p);
int be_memset_async(void* p, int value, size_t bytes, void* stream);
int be_memcpy_async(void* dst, const void* src, size_t bytes, CopyKind kind, void* stream);
int be_memcpy2d_async(void* dst, size_t dpitch, const void* src, size_t spitch, size_t width_bytes,
                      size_t height, CopyKind kind, void* stream);
int be_stream_create(void** stream);
int be_stream_destroy(void* stream);
int be_stream_sync(void* stream);
// Drop the helper stream a split Jacobi launch forked from `stream` (if any); be_stream_destroy does it too.
void be_side_stream_release(void* stream);
// events, for ordering work across streams (copy / compute pipelines)
int be_event_create(void** ev);
int be_event_destroy(void* ev);
int be_event_record(void* ev, void* stream);
int be_stream_wait_event(void* stream, void* ev);
const char* be_last_error_string();
// Kernels launched (or recorded into a graph capture) by the calling host thread so far: a split
// Jacobi launch counts its three kernels. The solver handle keeps its launch count with it.
long long be_kernel_count();

// CUDA-graph capture of a launch sequence (returns non-zero when unsupported).
int be_graph_begin(void* stream);
int be_graph_end(void* stream, void** graph_exec);
int be_graph_launch(void* graph_exec, void* stream);
int be_graph_destroy(void* graph_exec);

// Kernel launches. `nchunks` x ceil(nstrips / warps-per-CTA) CTAs; all asynchronous.
template <typename T> int launch_jacobi(const fb::JacobiArgs<T>& a, int sweeps, bool mask, int nchunks, void* stream);
template <typename T> int launch_predictor(const fb::PredictorArgs<T>& a, int nchunks, void* stream);
template <typename T> int launch_corrector(const fb::CorrectorArgs<T>& a, bool mask, int nchunks, void* stream);
template <typename T>
int launch_init_cavity(T* const fields[7], long long pitch, int X, int Y, int row0, int rows, T lid, void* stream);


// Resident solver for small grids (resident_kernels.cuh). be_resident_max_threads: threads per CTA the
// kernel with patch height R may be launched with (0: no such kernel, or no resident kernels at all in
// this backend); resident_capacity: CTAs of that plan that can be co-resident on the current device.
int be_resident_max_threads(int R);
template <typename T> int resident_capacity(const fb::ResidentPlan& plan);
template <typename T> int launch_resident(const fb::ResidentArgs<T>& a, const fb::ResidentPlan& plan, void* stream);

}  // namespace fbk
