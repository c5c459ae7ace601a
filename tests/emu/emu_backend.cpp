// emu_backend.cpp -- HOST implementation of fluid_simulation_b200/csrc/backend.h.
// TEST INFRASTRUCTURE ONLY: lets `pytest -m "not gpu"` run the product's host logic (api.cpp)
// and the very same warp task bodies (warp_kernels.cuh) on the CPU, 32 threads per warp with
// barrier-emulated shuffles, and compare them bit-for-bit with the oracle. It is built into
// tests/emu/libfluidb200_emu.so, which nothing under fluid_simulation_b200/ ever loads.
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <atomic>
#include <functional>
#include <thread>
#include <vector>

#define FLUIDB200_HOST_EMU 1
#include "backend.h"
#include "warp_kernels.cuh"

namespace {

constexpr size_t kFiberStack = 256 * 1024;

struct WarpRunner {
    EmuWarp w;
    std::vector<char> stacks;
    std::function<void(int)>* body = nullptr;
    WarpRunner() : stacks(32 * kFiberStack) {}

    static void trampoline(unsigned lo, unsigned hi, int lane) {
        WarpRunner* self = (WarpRunner*)(((uintptr_t)hi << 32) | (uintptr_t)lo);
        (*self->body)(lane);
        self->w.done[lane] = true;
        // returning switches to uc_link = the scheduler
    }

    // run one warp task: body(lane) for 32 lanes in lock step
    void run(std::function<void(int)>& fn) {
        body = &fn;
        emu_warp = &w;
        w.arrived = 0;
        for (int l = 0; l < 32; ++l) {
            w.par[l] = 0;
            w.waiting[l] = false;
            w.done[l] = false;
            getcontext(&w.ctx[l]);
            w.ctx[l].uc_stack.ss_sp = stacks.data() + (size_t)l * kFiberStack;
            w.ctx[l].uc_stack.ss_size = kFiberStack;
            w.ctx[l].uc_link = &w.sched;
            const uintptr_t p = (uintptr_t)this;
            makecontext(&w.ctx[l], (void (*)())trampoline, 3, (unsigned)(p & 0xffffffffu), (unsigned)(p >> 32), l);
        }
        for (;;) {
            int live = 0;
            for (int l = 0; l < 32; ++l) {
                if (w.done[l] || w.waiting[l]) continue;
                emu_lane = l;
                swapcontext(&w.sched, &w.ctx[l]);  // until the lane reaches a shuffle or returns
            }
            for (int l = 0; l < 32; ++l) live += !w.done[l];
            if (live == 0) break;
            if (w.arrived != live || live != 32) {
                // on the GPU this is a shuffle executed by only part of the warp: undefined behaviour
                std::fprintf(stderr, "EMU: divergent warp shuffle (%d of 32 lanes arrived, %d live)\n", w.arrived, live);
                std::abort();
            }
            w.arrived = 0;
            for (int l = 0; l < 32; ++l) w.waiting[l] = false;
        }
    }
};

static thread_local long long g_kernels = 0;  // one per emulated kernel launch (be_kernel_count)

// run fn(strip, chunk, lane) for every warp task. Tasks are independent (disjoint outputs), so they
// are spread over the host cores; each OS thread emulates one warp at a time.
template <typename F> void run_grid(int nstrips, int nchunks, F fn) {
    ++g_kernels;
    const int njobs = nstrips * nchunks;
    std::atomic<int> next{0};
    auto worker = [&] {
        WarpRunner runner;
        for (;;) {
            const int j = next.fetch_add(1);
            if (j >= njobs) break;
            const int strip = j % nstrips, chunk = j / nstrips;
            std::function<void(int)> body = [&, strip, chunk](int lane) { fn(strip, chunk, lane); };
            runner.run(body);
        }
    };
    int nthreads = (int)std::thread::hardware_concurrency();
    nthreads = std::max(1, std::min(nthreads, std::min(njobs, 16)));
    if (nthreads == 1) {
        worker();
    } else {
        std::vector<std::thread> pool;
        for (int t = 0; t < nthreads; ++t) pool.emplace_back(worker);
        for (auto& t : pool) t.join();
    }
}
}  // namespace

namespace fbk {
const char* backend_name() { return "host-emulation (tests only)"; }
const char* be_last_error_string() { return "host emulation error"; }
int be_set_device(int) { return 0; }
int be_sm_count() { return 148; }
int be_malloc(void** p, size_t bytes) {
    *p = aligned_alloc(256, (bytes + 255) / 256 * 256);
    return *p ? 0 : -1;
}
int be_free(void* p) { free(p); return 0; }
int be_memset_async(void* p, int v, size_t n, void*) { memset(p, v, n); return 0; }
int be_memcpy_async(void* d, const void* s, size_t n, CopyKind, void*) { memcpy(d, s, n); return 0; }
int be_memcpy2d_async(void* d, size_t dp, const void* s, size_t sp, size_t w, size_t h, CopyKind, void*) {
    for (size_t i = 0; i < h; ++i) memcpy((char*)d + i * dp, (const char*)s + i * sp, w);
    return 0;
}
int be_stream_create(void** s) { *s = nullptr; return 0; }
int be_stream_destroy(void*) { return 0; }
int be_stream_sync(void*) { return 0; }
void be_side_stream_release(void*) {}
int be_event_create(void** ev) { *ev = (void*)1; return 0; }   // (the emulation runs everything in order)
int be_event_destroy(void*) { return 0; }
int be_event_record(void*, void*) { return 0; }
int be_stream_wait_event(void*, void*) { return 0; }
int be_graph_begin(void*) { return -1; }
int be_graph_end(void*, void**) { return -1; }
int be_graph_launch(void*, void*) { return -1; }
long long be_kernel_count() { return g_kernels; }
int be_graph_destroy(void*) { return 0; }

static void set_ranges(std::initializer_list<const void*> ptrs, size_t bytes) {
    emu_nranges = 0;
    for (const void* p : ptrs)
        if (p) emu_ranges[emu_nranges++] = {(const char*)p, (const char*)p + bytes};
}
// Emulates the CUDA wrappers literally: (blockIdx.x, blockIdx.y, warp) -> task through the same
// jacobi_decode_task, so the split launch (FAST, EDGE and GENERAL kernel) is what gets tested.
template <typename T, int K>
static void jac_k(const fb::JacobiArgs<T>& a, bool mask, int nchunks) {
    set_ranges({a.p_in, a.p_out, a.rhs}, (size_t)a.g.rows * a.g.pitch * sizeof(T));
    constexpr int V = fb::VecWidth<T>::V;
    constexpr int W = fb::kWarpsPerCta;
    auto general = [&](const fb::JacobiArgs<T>& g, int nwarps_x, int ny) {
        run_grid(nwarps_x, ny, [&](int wx, int by, int l) {
            int strip, chunk;
            if (!fb::jacobi_decode_task(g, wx / W, by, wx % W, W, &strip, &chunk)) return;
            if (mask) fb::jacobi_general_task<T, V, K, true>(g, strip, chunk, l, l, 32);
            else fb::jacobi_general_task<T, V, K, false>(g, strip, chunk, l, l, 32);
        });
    };
    auto round_up = [](int n) { return (n + W - 1) / W * W; };
    if (a.part == 0) {
        fb::JacobiArgs<T> all = a;
        all.part = 0;
        general(all, round_up(a.g.nstrips), nchunks);
        return;
    }
    fb::JacobiArgs<T> gen = a, edge = a, fast = a;
    gen.part = 2;
    edge.part = 3;
    fast.part = 1;
    if (getenv("FLUIDB200_EMU_VERBOSE"))
        std::fprintf(stderr, "EMU split K=%d: strips [%d,%d) of %d, chunks [%d,%d) of %d -> %d fast, %d edge, %d general tasks (head %d, tail %d)\n", K,
                     a.sf_lo, a.sf_hi, a.g.nstrips, a.cf_lo, a.cf_hi, nchunks, (a.sf_hi - a.sf_lo) * (a.cf_hi - a.cf_lo),
                     fb::jacobi_part3_tasks(edge), fb::jacobi_part2_tasks(gen), a.head, a.tail);
    general(gen, round_up(fb::jacobi_part2_tasks(gen)), 1);
    run_grid(round_up(fb::jacobi_part3_tasks(edge)), 1, [&](int wx, int by, int l) {
        int strip, chunk;
        if (!fb::jacobi_decode_task(edge, wx / W, by, wx % W, W, &strip, &chunk)) return;
        fb::jacobi_edge_task<T, V, K>(edge, strip, chunk, l, l, 32);
    });
    run_grid(round_up(a.sf_hi - a.sf_lo), a.cf_hi - a.cf_lo, [&](int wx, int by, int l) {
        int strip, chunk;
        if (!fb::jacobi_decode_task(fast, wx / W, by, wx % W, W, &strip, &chunk)) return;
        fb::jacobi_fast_task<T, V, K>(fast, strip, chunk, l, l, 32);
    });
}
template <typename T> int launch_jacobi(const fb::JacobiArgs<T>& a, int sweeps, bool mask, int nchunks, void*) {
    switch (sweeps) {
        case 1: jac_k<T, 1>(a, mask, nchunks); break;
        case 2: jac_k<T, 2>(a, mask, nchunks); break;
        case 3: jac_k<T, 3>(a, mask, nchunks); break;
        case 4: jac_k<T, 4>(a, mask, nchunks); break;
        case 5: jac_k<T, 5>(a, mask, nchunks); break;
        case 6: jac_k<T, 6>(a, mask, nchunks); break;
        case 7: jac_k<T, 7>(a, mask, nchunks); break;
        case 8: jac_k<T, 8>(a, mask, nchunks); break;
        default: return -1;
    }
    return 0;
}
template <typename T> int launch_predictor(const fb::PredictorArgs<T>& a, int nchunks, void*) {
    constexpr int V = fb::VecWidth<T>::V;
    set_ranges({a.u, a.v, a.us, a.vs, a.rhs}, (size_t)a.g.rows * a.g.pitch * sizeof(T));
    const int mode = a.m.mode;
    if (mode == 2) run_grid(a.g.nstrips, nchunks, [&](int s, int c, int l) { fb::predictor_task<T, V, 2>(a, s, c, l, l, 32); });
    else if (mode == 1) run_grid(a.g.nstrips, nchunks, [&](int s, int c, int l) { fb::predictor_task<T, V, 1>(a, s, c, l, l, 32); });
    else run_grid(a.g.nstrips, nchunks, [&](int s, int c, int l) { fb::predictor_task<T, V, 0>(a, s, c, l, l, 32); });
    return 0;
}
template <typename T> int launch_corrector(const fb::CorrectorArgs<T>& a, bool mask, int nchunks, void*) {
    constexpr int V = fb::VecWidth<T>::V;
    set_ranges({a.p, a.us, a.vs, a.u, a.v}, (size_t)a.g.rows * a.g.pitch * sizeof(T));
    const int mode = a.m.mode;
    if (mask) {
        if (mode != 0) run_grid(a.g.nstrips, nchunks, [&](int s, int c, int l) { fb::corrector_task<T, V, true, 1>(a, s, c, l); });
        else run_grid(a.g.nstrips, nchunks, [&](int s, int c, int l) { fb::corrector_task<T, V, true, 0>(a, s, c, l); });
    } else if (mode == 2) run_grid(a.g.nstrips, nchunks, [&](int s, int c, int l) { fb::corrector_task<T, V, false, 2>(a, s, c, l); });
    else if (mode == 1) run_grid(a.g.nstrips, nchunks, [&](int s, int c, int l) { fb::corrector_task<T, V, false, 1>(a, s, c, l); });
    else run_grid(a.g.nstrips, nchunks, [&](int s, int c, int l) { fb::corrector_task<T, V, false, 0>(a, s, c, l); });
    return 0;
}
template <typename T>
int launch_init_cavity(T* const f[7], long long pitch, int, int Y, int, int rows, T lid, void*) {
    for (int k = 0; k < 7; ++k) {
        if (!f[k]) continue;
        for (long long i = 0; i < (long long)rows * pitch; ++i)
            f[k][i] = (k == 0 && (int)(i % pitch) == Y - 1) ? lid : T(0);
    }
    return 0;
}
// the resident (cooperative, shared-memory) kernels are not emulated: the host logic never selects them here
int be_resident_max_threads(int) { return 0; }
template <typename T> int resident_capacity(const fb::ResidentPlan&) { return 0; }
template <typename T> int launch_resident(const fb::ResidentArgs<T>&, const fb::ResidentPlan&, void*) { return -1; }
#define INSTANTIATE(T)                                                                \
    template int launch_jacobi<T>(const fb::JacobiArgs<T>&, int, bool, int, void*);  \
    template int launch_predictor<T>(const fb::PredictorArgs<T>&, int, void*);       \
    template int launch_corrector<T>(const fb::CorrectorArgs<T>&, bool, int, void*); \
    template int launch_init_cavity<T>(T* const[7], long long, int, int, int, int, T, void*); \
    template int resident_capacity<T>(const fb::ResidentPlan&);                      \
    template int launch_resident<T>(const fb::ResidentArgs<T>&, const fb::ResidentPlan&, void*);
INSTANTIATE(float)
INSTANTIATE(double)
}  // namespace fbk

// The allocation-free entry point needs a __device__ symbol; the host emulation has none.
#include "../../include/fluidb200.h"
extern "C" int fluidb200_poisson_solve_noalloc(const fluidb200_slab*, void*, const void*, const fluidb200_edges*, int, void*) {
    return fluidb200_report_error(FLUIDB200_ESTATE, "poisson_solve_noalloc is not available in the host emulation");
}

// GPU-to-GPU halo links do not exist in the host emulation (its multi-rank tests exchange through gloo).
#define EMU_NO_LINK(name, ...) extern "C" int name(__VA_ARGS__) { return fluidb200_report_error(FLUIDB200_ESTATE, #name " is not available in the host emulation"); }
EMU_NO_LINK(fluidb200_device_alloc, void**, size_t)
EMU_NO_LINK(fluidb200_device_free, void*)
EMU_NO_LINK(fluidb200_ipc_export, const void*, unsigned char*)
EMU_NO_LINK(fluidb200_ipc_open, const unsigned char*, void**)
EMU_NO_LINK(fluidb200_ipc_close, void*)
EMU_NO_LINK(fluidb200_peer_enable, int)
EMU_NO_LINK(fluidb200_halo_push, void*, const void*, size_t, void*, uint32_t, void*)
EMU_NO_LINK(fluidb200_halo_copy, void*, const void*, size_t, void*)
EMU_NO_LINK(fluidb200_halo_wait, void*, uint32_t, void*)

// Test hooks: the product's exact-rounding arithmetic (Ar<float>), element-wise, so that its
// division-free modes can be hammered against the reference formulas in NumPy. A flagged
// element is recomputed in the mode's fallback, like the kernels do per row.
static long long g_second_fallbacks = 0;   // elements that failed the checked products and took the divisions
extern "C" long long emu_second_fallbacks(void) { const long long n = g_second_fallbacks; g_second_fallbacks = 0; return n; }
template <int MODE> static void cdiff_hook(const float* a, const float* b, float h, long long n, float* out) {
    const fb::Metric<float> m = fb::make_metric<float>(h, h);
    for (long long i = 0; i < n; ++i) {
        fb::Flags fl;
        out[i] = fb::Ar<float>::cdiff<MODE>(a[i], b[i], m.hx, m.hx2, m.inv_hx2, m.inv_hx2f, fl);
        if (fl.raised(m.pow2_lo_m1)) {
            constexpr int FB = fb::FallbackMode<float, MODE>::value;
            fb::Flags f2;
            out[i] = fb::Ar<float>::cdiff<FB>(a[i], b[i], m.hx, m.hx2, m.inv_hx2, m.inv_hx2f, f2);
            if (FB == fb::DIV_FAST_CHECKED && f2.bad) { out[i] = fb::Ar<float>::cdiff<fb::DIV_EXACT>(a[i], b[i], m.hx, m.hx2, m.inv_hx2, m.inv_hx2f, f2); ++g_second_fallbacks; }
        }
    }
}
template <int MODE> static void lap_hook(const float* s, const float* c, float hx, float hy, long long n, float* out) {
    const fb::Metric<float> m = fb::make_metric<float>(hx, hy);
    for (long long i = 0; i < n; ++i) {
        fb::Flags fl;
        out[i] = fb::Ar<float>::lap_tail<MODE>(s[i], c[i], m, fl);
        if (fl.raised(m.pow2_lo_m1)) {
            constexpr int FB = fb::FallbackMode<float, MODE>::value;
            fb::Flags f2;
            out[i] = fb::Ar<float>::lap_tail<FB>(s[i], c[i], m, f2);
            if (FB == fb::DIV_FAST_CHECKED && f2.bad) { out[i] = fb::Ar<float>::lap_tail<fb::DIV_EXACT>(s[i], c[i], m, f2); ++g_second_fallbacks; }
        }
    }
}
extern "C" int emu_metric_mode_f32(float hx, float hy) { return fb::make_metric<float>(hx, hy).mode; }
extern "C" int emu_metric_mode_f64(double hx, double hy) { return fb::make_metric<double>(hx, hy).mode; }
// fp64: mode 2 (power-of-two spacing) against mode 0, element-wise
// fp64: modes 2 (power-of-two spacing) and 1 (quotients by FMA corrections, flagged elements redone) against mode 0, element-wise
template <int MODE> static void cdiff64_hook(const double* a, const double* b, double h, long long n, double* out) {
    const fb::Metric<double> m = fb::make_metric<double>(h, h);
    for (long long i = 0; i < n; ++i) {
        fb::Flags fl;
        out[i] = fb::Ar<double>::cdiff<MODE>(a[i], b[i], m.hx, m.hx2, m.inv_hx2, 0.0, fl);
        if (fl.raised(m.pow2_lo_m1)) { out[i] = fb::Ar<double>::cdiff<fb::DIV_EXACT>(a[i], b[i], m.hx, m.hx2, m.inv_hx2, 0.0, fl); ++g_second_fallbacks; }
    }
}
template <int MODE> static void lap64_hook(const double* s, const double* c, double hx, double hy, long long n, double* out) {
    const fb::Metric<double> m = fb::make_metric<double>(hx, hy);
    for (long long i = 0; i < n; ++i) {
        fb::Flags fl;
        out[i] = fb::Ar<double>::lap_tail<MODE>(s[i], c[i], m, fl);
        if (fl.raised(m.pow2_lo_m1)) { out[i] = fb::Ar<double>::lap_tail<fb::DIV_EXACT>(s[i], c[i], m, fl); ++g_second_fallbacks; }
    }
}
extern "C" void emu_cdiff_f64(int mode, const double* a, const double* b, double h, long long n, double* out) {
    if (mode == 2) cdiff64_hook<2>(a, b, h, n, out); else if (mode == 1) cdiff64_hook<1>(a, b, h, n, out); else cdiff64_hook<0>(a, b, h, n, out);
}
extern "C" void emu_lap_tail_f64(int mode, const double* s, const double* c, double hx, double hy, long long n, double* out) {
    if (mode == 2) lap64_hook<2>(s, c, hx, hy, n, out); else if (mode == 1) lap64_hook<1>(s, c, hx, hy, n, out); else lap64_hook<0>(s, c, hx, hy, n, out);
}
extern "C" void emu_cdiff_f32(int mode, const float* a, const float* b, float h, long long n, float* out) {
    if (mode == 2) cdiff_hook<2>(a, b, h, n, out); else if (mode == 1) cdiff_hook<1>(a, b, h, n, out); else cdiff_hook<0>(a, b, h, n, out);
}
extern "C" void emu_lap_tail_f32(int mode, const float* s, const float* c, float hx, float hy, long long n, float* out) {
    if (mode == 2) lap_hook<2>(s, c, hx, hy, n, out); else if (mode == 1) lap_hook<1>(s, c, hx, hy, n, out); else lap_hook<0>(s, c, hx, hy, n, out);
}
