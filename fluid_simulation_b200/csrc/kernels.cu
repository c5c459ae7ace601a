// kernels.cu -- CUDA (sm_100a) implementation of backend.h: __global__ wrappers around the warp
// tasks of warp_kernels.cuh, their launch dispatch, and the device-memory plumbing.
//
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -fmad=false
//        (no -use_fast_math: IEEE division, no flush-to-zero; -fmad=false as a second guard
//        next to the *_rn intrinsics, see the arithmetic contract in warp_kernels.cuh)
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <map>
#include <mutex>
#include <utility>

#include "backend.h"
#include "warp_kernels.cuh"
#include "resident_kernels.cuh"

namespace {

thread_local cudaError_t g_last = cudaSuccess;
thread_local long long g_kernels = 0;   // kernels launched by this host thread (see be_kernel_count)
inline int rc_of(cudaError_t e) {
    if (e != cudaSuccess) {
        g_last = e;
        return -1;
    }
    return 0;
}

constexpr int kThreads = fb::kWarpsPerCta * 32;

// Resident CTAs per SM the register allocator must leave room for: the pipeline is latency bound
// (ncu: ~50 % issue, ~50 % DRAM at 12 warps/SM), so shallow depths trade registers for warps.
// Depth 8 is capped at 168 registers (56 bytes spilled) although the kernel wants 210: in a split launch it
// runs the wall chunks underneath the lean kernel, and a 210-register CTA takes the place of TWO lean CTAs
// (3 x 162 registers fill the file) where a 168-register one takes one: 568 -> 564 us per launch at 16384^2,
// 321 -> 316 us on a 4096 x 32768 slab (depth 7: no gain, left at 2).
#ifndef FLUIDB200_JACOBI_MIN_BLOCKS
#define FLUIDB200_JACOBI_MIN_BLOCKS(K) ((K) <= 2 ? 5 : (K) <= 4 ? 4 : (K) <= 6 ? 3 : (K) == 8 ? 3 : 2)
#endif
// Programmatic dependent launch (sm_90+): a kernel launched with the attribute below may be scheduled
// while its predecessor in the stream is still running; it announces at once that its own successor
// may do the same and then waits, before touching any memory, until the predecessor has completed
// and its writes are visible. What overlaps is launch latency and CTA scheduling, which is most of
// a ~10 us launch on a small grid. Without the attribute both instructions are no-ops.
__device__ __forceinline__ void pdl_prologue() {
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");
}
bool pdl_enabled() {
    static const bool on = [] { const char* e = getenv("FLUIDB200_PDL"); return !e || atoi(e) != 0; }();
    return on;
}
template <typename A>
cudaError_t launch_k(void (*kern)(A), dim3 grid, size_t smem, cudaStream_t st, const A& arg) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = dim3(kThreads, 1, 1);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = pdl_enabled() ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kern, arg);
}

template <typename T, int V, int K, bool MASK>
__global__ void __launch_bounds__(kThreads, FLUIDB200_JACOBI_MIN_BLOCKS(K)) jacobi_kernel(const __grid_constant__ fb::JacobiArgs<T> A) {
    pdl_prologue();
    int strip, chunk;  // whole warps only; the tasks use no block barrier
    if (!fb::jacobi_decode_task(A, (int)blockIdx.x, (int)blockIdx.y, (int)(threadIdx.x >> 5), fb::kWarpsPerCta, &strip, &chunk)) return;
    fb::jacobi_general_task<T, V, K, MASK>(A, strip, chunk, threadIdx.x & 31, threadIdx.x, kThreads);
}

// The lean twin (see jacobi_fast_task): without the general path it needs fewer registers, so it
// is compiled for one more resident CTA per SM than jacobi_kernel.
#ifndef FLUIDB200_JACOBI_FAST_MIN_BLOCKS
#define FLUIDB200_JACOBI_FAST_MIN_BLOCKS(K) ((K) <= 2 ? 5 : (K) <= 6 ? 4 : 3)
#endif
template <typename T, int V, int K>
__global__ void __launch_bounds__(kThreads, FLUIDB200_JACOBI_FAST_MIN_BLOCKS(K)) jacobi_fast_kernel(const __grid_constant__ fb::JacobiArgs<T> A) {
    pdl_prologue();
    int strip, chunk;
    if (!fb::jacobi_decode_task(A, (int)blockIdx.x, (int)blockIdx.y, (int)(threadIdx.x >> 5), fb::kWarpsPerCta, &strip, &chunk)) return;
    fb::jacobi_fast_task<T, V, K>(A, strip, chunk, threadIdx.x & 31, threadIdx.x, kThreads);
}

// The lean pipeline plus the y-edge cells, for the strips that hold column 0 or Y-1.
#ifndef FLUIDB200_JACOBI_EDGE_MIN_BLOCKS
#define FLUIDB200_JACOBI_EDGE_MIN_BLOCKS(K) ((K) <= 2 ? 5 : (K) <= 4 ? 4 : (K) <= 6 ? 3 : 2)
#endif
template <typename T, int V, int K>
__global__ void __launch_bounds__(kThreads, FLUIDB200_JACOBI_EDGE_MIN_BLOCKS(K)) jacobi_edge_kernel(const __grid_constant__ fb::JacobiArgs<T> A) {
    pdl_prologue();
    int strip, chunk;
    if (!fb::jacobi_decode_task(A, (int)blockIdx.x, (int)blockIdx.y, (int)(threadIdx.x >> 5), fb::kWarpsPerCta, &strip, &chunk)) return;
    fb::jacobi_edge_task<T, V, K>(A, strip, chunk, threadIdx.x & 31, threadIdx.x, kThreads);
}

#ifndef FLUIDB200_PRED_MIN_BLOCKS
#define FLUIDB200_PRED_MIN_BLOCKS 4
#endif
template <typename T, int V, int MODE>
__global__ void __launch_bounds__(kThreads, FLUIDB200_PRED_MIN_BLOCKS) predictor_kernel(const __grid_constant__ fb::PredictorArgs<T> A) {
    pdl_prologue();
    const int strip = blockIdx.x * fb::kWarpsPerCta + (threadIdx.x >> 5);
    if (strip >= A.g.nstrips) return;
    fb::predictor_task<T, V, MODE>(A, strip, blockIdx.y, threadIdx.x & 31, threadIdx.x, kThreads);
}

template <typename T, int V, bool MASK, int MODE>
__global__ void __launch_bounds__(kThreads, 4) corrector_kernel(const __grid_constant__ fb::CorrectorArgs<T> A) {
    pdl_prologue();
    const int strip = blockIdx.x * fb::kWarpsPerCta + (threadIdx.x >> 5);
    if (strip >= A.g.nstrips) return;
    fb::corrector_task<T, V, MASK, MODE>(A, strip, blockIdx.y, threadIdx.x & 31);
}

// Whole timesteps of a small grid in one cooperative launch (resident_kernels.cuh). One CTA per SM;
// the patch height R fixes how many threads the registers allow.
constexpr int resident_max_threads(int R) { return R <= 4 ? 384 : 256; }
template <typename T, int V, int R>
__global__ void __launch_bounds__(resident_max_threads(R), 1) resident_kernel(const __grid_constant__ fb::ResidentArgs<T> A) {
    fb::resident_body<T, V, R>(A);
}
// ... and grids of at most NC * 1024 cells in ONE CTA, a cell per thread and slot (plan.R == 0, plan.wy == NC)
template <typename T, int NC>
__global__ void __launch_bounds__(fb::kCellThreads, 1) resident_cell_kernel(const __grid_constant__ fb::ResidentArgs<T> A) {
    fb::resident_cell_body<T, NC>(A);
}
template <typename T> void (*resident_cell_entry(int NC))(fb::ResidentArgs<T>) {
    switch (NC) {
        case 2: return resident_cell_kernel<T, 2>;
        case 4: return resident_cell_kernel<T, 4>;
        default: return nullptr;
    }
}
template <typename T, int R> void (*resident_entry())(fb::ResidentArgs<T>) {
    return resident_kernel<T, fb::VecWidth<T>::V, R>;
}
template <typename T> void (*resident_entry_r(int R))(fb::ResidentArgs<T>) {
    switch (R) {
        case 4: return resident_entry<T, 4>();
        case 8: return resident_entry<T, 8>();
        default: return nullptr;
    }
}

template <typename T> struct FieldPtrs {
    T* f[7];
};

// zero state + cavity boundary values (fluid_cpp/main.cpp:43-55): only u on the lid is non-zero
template <typename T>
__global__ void init_cavity_kernel(FieldPtrs<T> F, long long pitch, int Y, long long n, T lid) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const int col = (int)(i % pitch);
#pragma unroll
        for (int k = 0; k < 7; ++k)
            if (F.f[k]) F.f[k][i] = (k == 0 && col == Y - 1) ? lid : T(0);
    }
}

inline dim3 grid_for(int nstrips, int nchunks) {
    return dim3((unsigned)((nstrips + fb::kWarpsPerCta - 1) / fb::kWarpsPerCta), (unsigned)nchunks, 1);
}

// Side stream for the GENERAL and EDGE parts of a split Jacobi launch: forked from and joined back
// into the caller's stream with events, high priority so that its few, long tasks start first and
// run underneath the FAST kernel instead of after it. One per (device, caller stream): two solver
// handles (each on its own stream) never share one, whichever host threads drive them. Created on
// the first split launch from a stream (one cudaStreamCreateWithPriority + two cudaEventCreate);
// be_side_stream_release() gives it back when the caller's stream goes away.
struct SideStream {
    cudaStream_t stream = nullptr;
    cudaEvent_t fork = nullptr, join = nullptr;
};
std::mutex g_side_mutex;
std::map<std::pair<int, cudaStream_t>, SideStream> g_side_streams;
SideStream* side_stream(cudaStream_t caller) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return nullptr;
    std::lock_guard<std::mutex> lock(g_side_mutex);
    SideStream& s = g_side_streams[std::make_pair(dev, caller)];  // (references into a std::map stay valid)
    if (!s.stream) {
        int lo = 0, hi = 0;
        cudaDeviceGetStreamPriorityRange(&lo, &hi);  // hi = numerically smallest = highest priority
        if (cudaStreamCreateWithPriority(&s.stream, cudaStreamNonBlocking, hi) != cudaSuccess ||
            cudaEventCreateWithFlags(&s.fork, cudaEventDisableTiming) != cudaSuccess ||
            cudaEventCreateWithFlags(&s.join, cudaEventDisableTiming) != cudaSuccess) {
            (void)cudaGetLastError();
            if (s.stream) cudaStreamDestroy(s.stream);
            if (s.fork) cudaEventDestroy(s.fork);
            g_side_streams.erase(std::make_pair(dev, caller));
            return nullptr;
        }
    }
    return &s;
}

template <typename T, int K> int launch_jacobi_k(const fb::JacobiArgs<T>& a, bool mask, int nchunks, cudaStream_t st) {
    constexpr int V = fb::VecWidth<T>::V;
    // dynamic shared memory: the per-lane cp.async prefetch ring (D slots x 2 arrays x 16 B per thread)
    size_t smem = (size_t)fb::JacobiRing<K>::D * 2 * 16 * kThreads;
#ifdef FLUIDB200_TMA_RING
    smem += (size_t)fb::JacobiRing<K>::D * 8 * fb::kWarpsPerCta;   // one mbarrier per (warp, slot)
#endif
    SideStream* side = a.part != 0 ? side_stream(st) : nullptr;
    if (!side) {  // one kernel does every task
        fb::JacobiArgs<T> all = a;
        all.part = 0;
        const dim3 grid = grid_for(a.g.nstrips, nchunks);
        if (mask) launch_k(jacobi_kernel<T, V, K, true>, grid, smem, st, all);
        else launch_k(jacobi_kernel<T, V, K, false>, grid, smem, st, all);
        ++g_kernels;
        return rc_of(cudaGetLastError());
    }
    // split launch: GENERAL (wall chunks) and EDGE (y-edge strips) tasks on the side stream, FAST tasks
    // on the caller's stream, then join
    fb::JacobiArgs<T> gen = a, edge = a, fast = a;
    gen.part = 2;
    edge.part = 3;
    fast.part = 1;
    if (rc_of(cudaEventRecord(side->fork, st)) || rc_of(cudaStreamWaitEvent(side->stream, side->fork, 0))) return -1;
    const int ngen = fb::jacobi_part2_tasks(gen), nedge = fb::jacobi_part3_tasks(edge);
    if (nedge > 0) {
        const dim3 egrid((unsigned)((nedge + fb::kWarpsPerCta - 1) / fb::kWarpsPerCta), 1, 1);
        jacobi_edge_kernel<T, V, K><<<egrid, kThreads, smem, side->stream>>>(edge);
        ++g_kernels;
        if (rc_of(cudaGetLastError())) return -1;
    }
    if (ngen > 0) {
        const dim3 ggrid((unsigned)((ngen + fb::kWarpsPerCta - 1) / fb::kWarpsPerCta), 1, 1);
        if (mask) jacobi_kernel<T, V, K, true><<<ggrid, kThreads, smem, side->stream>>>(gen);
        else jacobi_kernel<T, V, K, false><<<ggrid, kThreads, smem, side->stream>>>(gen);
        ++g_kernels;
        if (rc_of(cudaGetLastError())) return -1;
    }
    if (rc_of(cudaEventRecord(side->join, side->stream))) return -1;
    launch_k(jacobi_fast_kernel<T, V, K>, grid_for(a.sf_hi - a.sf_lo, a.cf_hi - a.cf_lo), smem, st, fast);
    ++g_kernels;
    if (rc_of(cudaGetLastError())) return -1;
    return rc_of(cudaStreamWaitEvent(st, side->join, 0));
}

}  // namespace

namespace fbk {

const char* backend_name() { return "cuda-sm_100a"; }
const char* be_last_error_string() { return cudaGetErrorString(g_last); }
long long be_kernel_count() { return g_kernels; }
int be_set_device(int device) { return rc_of(cudaSetDevice(device)); }
int be_sm_count() {
    static thread_local int cached_dev = -1, cached = 148;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 148;
    if (dev != cached_dev) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && n > 0) cached = n;
        cached_dev = dev;
    }
    return cached;
}

int be_malloc(void** p, size_t bytes) { return rc_of(cudaMalloc(p, bytes)); }
int be_free(void* p) { return rc_of(cudaFree(p)); }
int be_memset_async(void* p, int value, size_t bytes, void* stream) {
    return rc_of(cudaMemsetAsync(p, value, bytes, (cudaStream_t)stream));
}
static cudaMemcpyKind kind_of(CopyKind k) {
    return k == H2D ? cudaMemcpyHostToDevice : (k == D2H ? cudaMemcpyDeviceToHost : cudaMemcpyDeviceToDevice);
}
int be_memcpy_async(void* dst, const void* src, size_t bytes, CopyKind kind, void* stream) {
    return rc_of(cudaMemcpyAsync(dst, src, bytes, kind_of(kind), (cudaStream_t)stream));
}
int be_memcpy2d_async(void* dst, size_t dpitch, const void* src, size_t spitch, size_t width_bytes, size_t height,
                      CopyKind kind, void* stream) {
    return rc_of(cudaMemcpy2DAsync(dst, dpitch, src, spitch, width_bytes, height, kind_of(kind), (cudaStream_t)stream));
}
int be_stream_create(void** stream) {
    cudaStream_t s;
    const int rc = rc_of(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
    *stream = rc ? nullptr : (void*)s;
    return rc;
}
int be_stream_destroy(void* stream) {
    be_side_stream_release(stream);
    return rc_of(cudaStreamDestroy((cudaStream_t)stream));
}
void be_side_stream_release(void* stream) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return;
    std::lock_guard<std::mutex> lock(g_side_mutex);
    auto it = g_side_streams.find(std::make_pair(dev, (cudaStream_t)stream));
    if (it == g_side_streams.end()) return;
    cudaStreamSynchronize(it->second.stream);
    cudaEventDestroy(it->second.fork);
    cudaEventDestroy(it->second.join);
    cudaStreamDestroy(it->second.stream);
    g_side_streams.erase(it);
}
int be_stream_sync(void* stream) { return rc_of(cudaStreamSynchronize((cudaStream_t)stream)); }

// A failed capture call leaves its (non-sticky) error in the runtime's per-thread slot, where the
// next launch's cudaGetLastError() would find it: the caller falls back to plain launches, so the
// slot is cleared here.
int be_event_create(void** ev) {
    cudaEvent_t e;
    const int rc = rc_of(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    *ev = rc ? nullptr : (void*)e;
    return rc;
}
int be_event_destroy(void* ev) { return rc_of(cudaEventDestroy((cudaEvent_t)ev)); }
int be_event_record(void* ev, void* stream) { return rc_of(cudaEventRecord((cudaEvent_t)ev, (cudaStream_t)stream)); }
int be_stream_wait_event(void* stream, void* ev) { return rc_of(cudaStreamWaitEvent((cudaStream_t)stream, (cudaEvent_t)ev, 0)); }

int be_graph_begin(void* stream) {
    const int rc = rc_of(cudaStreamBeginCapture((cudaStream_t)stream, cudaStreamCaptureModeThreadLocal));
    if (rc) (void)cudaGetLastError();
    return rc;
}
int be_graph_end(void* stream, void** graph_exec) {
    cudaGraph_t g = nullptr;
    *graph_exec = nullptr;
    if (rc_of(cudaStreamEndCapture((cudaStream_t)stream, &g))) {
        (void)cudaGetLastError();
        return -1;
    }
    cudaGraphExec_t ex = nullptr;
    const int rc = rc_of(cudaGraphInstantiate(&ex, g, 0));
    cudaGraphDestroy(g);
    if (rc) {
        (void)cudaGetLastError();
        return rc;
    }
    *graph_exec = (void*)ex;
    return 0;
}
int be_graph_launch(void* graph_exec, void* stream) {
    return rc_of(cudaGraphLaunch((cudaGraphExec_t)graph_exec, (cudaStream_t)stream));
}
int be_graph_destroy(void* graph_exec) { return rc_of(cudaGraphExecDestroy((cudaGraphExec_t)graph_exec)); }

template <typename T>
int launch_jacobi(const fb::JacobiArgs<T>& a, int sweeps, bool mask, int nchunks, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    switch (sweeps) {
        case 1: return launch_jacobi_k<T, 1>(a, mask, nchunks, st);
        case 2: return launch_jacobi_k<T, 2>(a, mask, nchunks, st);
        case 3: return launch_jacobi_k<T, 3>(a, mask, nchunks, st);
        case 4: return launch_jacobi_k<T, 4>(a, mask, nchunks, st);
        case 5: return launch_jacobi_k<T, 5>(a, mask, nchunks, st);
        case 6: return launch_jacobi_k<T, 6>(a, mask, nchunks, st);
        case 7: return launch_jacobi_k<T, 7>(a, mask, nchunks, st);
        case 8: return launch_jacobi_k<T, 8>(a, mask, nchunks, st);
        default: g_last = cudaErrorInvalidValue; return -1;
    }
}

template <typename T> int launch_predictor(const fb::PredictorArgs<T>& a, int nchunks, void* stream) {
    constexpr int V = fb::VecWidth<T>::V;
    const dim3 grid = grid_for(a.g.nstrips, nchunks);
    cudaStream_t st = (cudaStream_t)stream;
    // the arithmetic mode is a compile-time property of the kernel (see Metric::mode)
    const size_t smem = (size_t)FLUIDB200_PRED_D * 2 * 16 * kThreads;  // per-lane cp.async ring: D slots x {u, v} x 16 B
    if (a.m.mode == 2) launch_k(predictor_kernel<T, V, 2>, grid, smem, st, a);
    else if (a.m.mode == 1) launch_k(predictor_kernel<T, V, 1>, grid, smem, st, a);
    else launch_k(predictor_kernel<T, V, 0>, grid, smem, st, a);
    ++g_kernels;
    return rc_of(cudaGetLastError());
}

template <typename T> int launch_corrector(const fb::CorrectorArgs<T>& a, bool mask, int nchunks, void* stream) {
    constexpr int V = fb::VecWidth<T>::V;
    const dim3 grid = grid_for(a.g.nstrips, nchunks);
    cudaStream_t st = (cudaStream_t)stream;
    const int mode = a.m.mode;
    if (mask) {  // the general-BC path is not specialised further
        if (mode != 0) launch_k(corrector_kernel<T, V, true, 1>, grid, 0, st, a);
        else launch_k(corrector_kernel<T, V, true, 0>, grid, 0, st, a);
    } else if (mode == 2) launch_k(corrector_kernel<T, V, false, 2>, grid, 0, st, a);
    else if (mode == 1) launch_k(corrector_kernel<T, V, false, 1>, grid, 0, st, a);
    else launch_k(corrector_kernel<T, V, false, 0>, grid, 0, st, a);
    ++g_kernels;
    return rc_of(cudaGetLastError());
}

template <typename T>
int launch_init_cavity(T* const fields[7], long long pitch, int X, int Y, int row0, int rows, T lid, void* stream) {
    (void)X; (void)row0;
    FieldPtrs<T> F;
    for (int k = 0; k < 7; ++k) F.f[k] = fields[k];
    const long long n = (long long)rows * pitch;
    long long blocks = (n + 255) / 256;
    if (blocks > 148LL * 16) blocks = 148LL * 16;
    init_cavity_kernel<T><<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(F, pitch, Y, n, lid);
    ++g_kernels;
    return rc_of(cudaGetLastError());
}

int be_resident_max_threads(int R) { return R == 0 ? fb::kCellThreads : (R == 4 || R == 8) ? resident_max_threads(R) : 0; }

// CTAs of the resident kernel that can be co-resident on the current device (0: not launchable)
template <typename T> int resident_capacity(const fb::ResidentPlan& pl) {
    auto kern = pl.R == 0 ? resident_cell_entry<T>(pl.wy) : resident_entry_r<T>(pl.R);
    if (!kern || pl.threads > (pl.R == 0 ? fb::kCellThreads : resident_max_threads(pl.R))) return 0;
    int dev = 0, coop = 0, max_optin = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 0;
    cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev);
    cudaDeviceGetAttribute(&max_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
    if (!coop || pl.smem > (size_t)max_optin) return 0;
    if (cudaFuncSetAttribute((const void*)kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl.smem) != cudaSuccess) {
        (void)cudaGetLastError();
        return 0;
    }
    int per_sm = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, (const void*)kern, pl.threads, pl.smem) != cudaSuccess) {
        (void)cudaGetLastError();
        return 0;
    }
    return per_sm * be_sm_count();
}
template <typename T> int launch_resident(const fb::ResidentArgs<T>& a, const fb::ResidentPlan& pl, void* stream) {
    auto kern = pl.R == 0 ? resident_cell_entry<T>(pl.wy) : resident_entry_r<T>(pl.R);
    if (!kern) { g_last = cudaErrorInvalidValue; return -1; }
    cudaStream_t st = (cudaStream_t)stream;
    if (pl.R == 0) {   // one CTA: an ordinary launch, no scratch
        if (rc_of(cudaFuncSetAttribute((const void*)kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl.smem))) return -1;
        kern<<<1, pl.threads, pl.smem, st>>>(a);
        ++g_kernels;
        return rc_of(cudaGetLastError());
    }
    const int ntiles = pl.tiles_x * pl.tiles_y;
    if (rc_of(cudaMemsetAsync(a.sync, 0, fb::resident_scratch_bytes<T>(pl), st))) return -1;   // counters, packets
    if (rc_of(cudaFuncSetAttribute((const void*)kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl.smem))) return -1;
    fb::ResidentArgs<T> args = a;
    void* params[1] = {(void*)&args};
    if (rc_of(cudaLaunchCooperativeKernel((const void*)kern, dim3((unsigned)ntiles), dim3((unsigned)pl.threads), params, pl.smem, st)))
        return -1;
    ++g_kernels;
    return 0;
}

#define INSTANTIATE(T)                                                                                   \
    template int launch_jacobi<T>(const fb::JacobiArgs<T>&, int, bool, int, void*);                     \
    template int launch_predictor<T>(const fb::PredictorArgs<T>&, int, void*);                          \
    template int launch_corrector<T>(const fb::CorrectorArgs<T>&, bool, int, void*);                    \
    template int launch_init_cavity<T>(T* const[7], long long, int, int, int, int, T, void*);       \
    template int resident_capacity<T>(const fb::ResidentPlan&);                                          \
    template int launch_resident<T>(const fb::ResidentArgs<T>&, const fb::ResidentPlan&, void*);
INSTANTIATE(float)
INSTANTIATE(double)

}  // namespace fbk

// =========================================================================================
// Allocation-free Poisson entry point (see include/fluidb200.h for the user-facing contract)
// =========================================================================================
// The Jacobi sweeps ping-pong between two pressure buffers. Every other entry point gets the
// second buffer from the caller or from cudaMalloc. Some callers can do neither: code running
// while a memory pool is frozen or nearly exhausted, real-time loops that must not enter the
// driver's allocator (cudaMalloc serialises against the whole device and may block for
// milliseconds), and plug-in hosts that hand a library its input and output buffers and forbid
// any allocation of its own. For them the partner buffer lives in this module's own image:
// a file-scope __device__ array, reserved on every GPU that loads the library, for the life of
// the process.
#include "internal.h"

static __device__ __align__(256) unsigned char g_noalloc_scratch[FLUIDB200_NOALLOC_MAX_BYTES];

namespace {

constexpr int kMaxDevices = 64;
void* g_noalloc_ptr[kMaxDevices] = {};   // device address of g_noalloc_scratch, per device
std::mutex g_noalloc_mutex;

// In host code `g_noalloc_scratch` is only a handle for the symbol, not a device address; the
// address a kernel may dereference has to be asked for, with the device in question current.
int noalloc_scratch_for_current_device(void** out) {
    int dev = -1;
    if (rc_of(cudaGetDevice(&dev))) return -1;
    if (dev < 0 || dev >= kMaxDevices) {
        g_last = cudaErrorInvalidDevice;
        return -1;
    }
    std::lock_guard<std::mutex> lock(g_noalloc_mutex);
    if (!g_noalloc_ptr[dev]) {
        void* p = nullptr;
        if (rc_of(cudaGetSymbolAddress(&p, g_noalloc_scratch))) return -1;
        g_noalloc_ptr[dev] = p;
    }
    *out = g_noalloc_ptr[dev];
    return 0;
}

thread_local char g_noalloc_err[256];

}  // namespace

extern "C" int fluidb200_poisson_solve_noalloc(const fluidb200_slab* g, void* p, const void* rhs,
                                               const fluidb200_edges* p_edges, int sweeps, void* stream) {
    if (!g || !p || !rhs || !p_edges) return fluidb200_report_error(FLUIDB200_EINVAL, "poisson_solve_noalloc: NULL argument");
    if (sweeps < 0) return fluidb200_report_error(FLUIDB200_EINVAL, "poisson_solve_noalloc: sweeps must be >= 0");
    if (g->row0 != 0 || g->rows != g->X)
        return fluidb200_report_error(FLUIDB200_EINVAL, "poisson_solve_noalloc: works on a whole grid (row0 = 0, rows = X)");
    const size_t esize = g->dtype == FLUIDB200_F64 ? 8 : 4;
    if (g->pitch <= 0 || g->rows <= 0) return fluidb200_report_error(FLUIDB200_EINVAL, "poisson_solve_noalloc: bad geometry");
    const unsigned long long need = (unsigned long long)g->rows * (unsigned long long)g->pitch * esize;
    if (need > (unsigned long long)FLUIDB200_NOALLOC_MAX_BYTES) {
        snprintf(g_noalloc_err, sizeof(g_noalloc_err),
                 "poisson_solve_noalloc: the grid needs %llu bytes of scratch (rows x pitch x element size) but the "
                 "built-in scratch holds %llu; use fluidb200_poisson_sweeps with your own second buffer",
                 need, (unsigned long long)FLUIDB200_NOALLOC_MAX_BYTES);
        return fluidb200_report_error(FLUIDB200_EINVAL, g_noalloc_err);
    }
    void* scratch = nullptr;
    if (noalloc_scratch_for_current_device(&scratch)) {
        snprintf(g_noalloc_err, sizeof(g_noalloc_err), "poisson_solve_noalloc: cannot resolve the scratch symbol: %s",
                 cudaGetErrorString(g_last));
        return fluidb200_report_error(FLUIDB200_ECUDA, g_noalloc_err);
    }
    {   // the full argument check of a launch (dtype, pitch, alignment, edges), also when sweeps == 0
        const int rc = fbapi::poisson_sweeps(g, p, scratch, rhs, p_edges, nullptr, 1, 0, g->X, stream,
                                             fbapi::POISSON_VALIDATE_ONLY);
        if (rc) return rc;
    }
    if (sweeps == 0) return 0;
    // Split the sweeps into launches of at most 4, and into an EVEN number of launches when
    // possible, so the ping-pong ends in the caller's buffer. A single sweep ends in the scratch
    // and is copied back.
    const int kmax = 4;
    int nl = (sweeps + kmax - 1) / kmax;
    if (sweeps >= 2 && (nl & 1)) ++nl;
    void* src = p;
    void* dst = scratch;
    int left = sweeps;
    for (int i = 0; i < nl; ++i) {
        const int k = (left + (nl - i) - 1) / (nl - i);   // spread the sweeps evenly over the launches
        // always the single-kernel form of a launch: the split form would fork onto a helper stream,
        // which the library creates on first use -- a driver-side allocation this entry point promises not to make
        const int rc = fbapi::poisson_sweeps(g, src, dst, rhs, p_edges, nullptr, k, 0, g->X, stream, fbapi::POISSON_NO_SPLIT);
        if (rc) return rc;
        left -= k;
        void* t = src; src = dst; dst = t;
    }
    if (src != p) {  // odd number of launches (sweeps == 1): result sits in the scratch
        if (rc_of(cudaMemcpyAsync(p, scratch, (size_t)need, cudaMemcpyDeviceToDevice, (cudaStream_t)stream))) {
            snprintf(g_noalloc_err, sizeof(g_noalloc_err), "poisson_solve_noalloc: copy back failed: %s", cudaGetErrorString(g_last));
            return fluidb200_report_error(FLUIDB200_ECUDA, g_noalloc_err);
        }
    }
    return 0;
}

// =========================================================================================
// Halo exchange between GPUs without a collective: one-sided pushes over NVLink
// (see include/fluidb200.h, "x-slab halo links"). Plain peer copies and a flag word; no TMA.
// =========================================================================================
#include <cuda.h>

namespace {

__global__ void halo_flag_kernel(volatile unsigned* flag, unsigned seq) {
    *flag = seq;
    __threadfence_system();
}
__global__ void halo_spin_kernel(const volatile unsigned* flag, unsigned seq) {
    while ((int)(*flag - seq) < 0) {}
}
typedef CUresult (*wait_value32_fn)(CUstream, CUdeviceptr, cuuint32_t, unsigned int);
wait_value32_fn get_wait_value32() {
    static const wait_value32_fn fn = [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (getenv("FLUIDB200_HALO_SPIN")) return (wait_value32_fn) nullptr;   // (measurement knob: wait in a kernel)
        if (cudaGetDriverEntryPoint("cuStreamWaitValue32", &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) {
            (void)cudaGetLastError();
            return (wait_value32_fn) nullptr;
        }
        return (wait_value32_fn)p;
    }();
    return fn;
}
int halo_fail(const char* what) {
    static thread_local char msg[256];
    snprintf(msg, sizeof(msg), "%s: %s", what, cudaGetErrorString(g_last));
    return fluidb200_report_error(FLUIDB200_ECUDA, msg);
}

}  // namespace

extern "C" int fluidb200_device_alloc(void** ptr, size_t bytes) {
    if (!ptr || bytes == 0) return fluidb200_report_error(FLUIDB200_EINVAL, "device_alloc: bad argument");
    if (rc_of(cudaMalloc(ptr, bytes))) return halo_fail("device_alloc");
    if (rc_of(cudaMemset(*ptr, 0, bytes))) return halo_fail("device_alloc: clear");
    return 0;
}
extern "C" int fluidb200_device_free(void* ptr) {
    if (ptr && rc_of(cudaFree(ptr))) return halo_fail("device_free");
    return 0;
}
extern "C" int fluidb200_ipc_export(const void* ptr, unsigned char handle[64]) {
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "cudaIpcMemHandle_t is 64 bytes");
    if (!ptr || !handle) return fluidb200_report_error(FLUIDB200_EINVAL, "ipc_export: NULL argument");
    cudaIpcMemHandle_t h;
    if (rc_of(cudaIpcGetMemHandle(&h, const_cast<void*>(ptr)))) return halo_fail("ipc_export");
    memcpy(handle, &h, 64);
    return 0;
}
extern "C" int fluidb200_ipc_open(const unsigned char handle[64], void** ptr) {
    if (!ptr || !handle) return fluidb200_report_error(FLUIDB200_EINVAL, "ipc_open: NULL argument");
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, 64);
    if (rc_of(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess))) return halo_fail("ipc_open");
    return 0;
}
extern "C" int fluidb200_ipc_close(void* ptr) {
    if (ptr && rc_of(cudaIpcCloseMemHandle(ptr))) return halo_fail("ipc_close");
    return 0;
}
extern "C" int fluidb200_peer_enable(int peer_device) {
    int dev = 0, can = 0;
    if (rc_of(cudaGetDevice(&dev))) return halo_fail("peer_enable");
    if (dev == peer_device) return 0;
    if (rc_of(cudaDeviceCanAccessPeer(&can, dev, peer_device))) return halo_fail("peer_enable");
    if (!can) return fluidb200_report_error(FLUIDB200_ESTATE, "peer_enable: the two devices cannot reach each other's memory");
    const cudaError_t e = cudaDeviceEnablePeerAccess(peer_device, 0);
    if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) {
        g_last = e;
        return halo_fail("peer_enable");
    }
    (void)cudaGetLastError();
    return 0;
}
extern "C" int fluidb200_halo_push(void* dst, const void* src, size_t bytes, void* flag, uint32_t seq, void* stream) {
    if ((bytes && (!dst || !src)) || !flag) return fluidb200_report_error(FLUIDB200_EINVAL, "halo_push: NULL argument");
    cudaStream_t st = (cudaStream_t)stream;
    if (bytes && rc_of(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDefault, st))) return halo_fail("halo_push: copy");
    halo_flag_kernel<<<1, 1, 0, st>>>((volatile unsigned*)flag, seq);
    ++g_kernels;
    if (rc_of(cudaGetLastError())) return halo_fail("halo_push: flag");
    return 0;
}
extern "C" int fluidb200_halo_copy(void* dst, const void* src, size_t bytes, void* stream) {
    if (!dst || !src) return fluidb200_report_error(FLUIDB200_EINVAL, "halo_copy: NULL argument");
    if (rc_of(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDefault, (cudaStream_t)stream))) return halo_fail("halo_copy");
    return 0;
}
extern "C" int fluidb200_halo_wait(void* flag, uint32_t seq, void* stream) {
    if (!flag) return fluidb200_report_error(FLUIDB200_EINVAL, "halo_wait: NULL argument");
    if (wait_value32_fn fn = get_wait_value32()) {
        // (flag - seq) as a signed number >= 0: sequence numbers stay far below 2^31 apart
        const CUresult r = fn((CUstream)stream, (CUdeviceptr)flag, seq, CU_STREAM_WAIT_VALUE_GEQ);
        if (r == CUDA_SUCCESS) return 0;
    }
    halo_spin_kernel<<<1, 1, 0, (cudaStream_t)stream>>>((const volatile unsigned*)flag, seq);
    ++g_kernels;
    if (rc_of(cudaGetLastError())) return halo_fail("halo_wait");
    return 0;
}
