// emu_shim.h -- HOST emulation shims so fluid_simulation_b200/csrc/warp_kernels.cuh can be
// compiled with g++ and its warp tasks run on the CPU (TEST INFRASTRUCTURE ONLY).
// A warp is 32 std::threads, one per lane; a shuffle is "publish, barrier, read, barrier".
// Warp-uniform control flow around shuffles is therefore REQUIRED, exactly as on the GPU:
// a divergent shuffle deadlocks here instead of being undefined behaviour there.
// Build with -ffp-contract=off so the *_rn shims below are never fused.
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>

#define FB_DEV inline
#define FB_DEV_NOINLINE inline __attribute__((noinline))
using std::max;
using std::min;

// Bounds checking: the emulation backend registers the arrays of the launch being emulated and every
// vector load/store/prefetch of the kernels is checked against them (compute-sanitizer is not
// available on the GPU pool, so this is the memcheck of this repository).
#include <cstdio>
#include <cstdlib>
struct EmuRange { const char* lo; const char* hi; };
inline EmuRange emu_ranges[16];
inline int emu_nranges = 0;
inline void emu_check(const void* p, size_t n, const char* what) {
    if (emu_nranges == 0) return;
    const char* c = (const char*)p;
    for (int i = 0; i < emu_nranges; ++i)
        if (c >= emu_ranges[i].lo && c + n <= emu_ranges[i].hi) return;
    std::fprintf(stderr, "EMU: out-of-bounds %s of %zu bytes at %p\n", what, n, p);
    std::abort();
}
#define FB_EMU_CHECK(p, n, what) emu_check((p), (n), (what))

// A warp is 32 user-level fibers (ucontext) on ONE OS thread, run in lock step by a small
// scheduler: a lane runs until it reaches a shuffle (or returns); when every live lane has
// published its value the scheduler resumes them one by one to read their source lane's value.
// Values are double buffered by shuffle parity, so a resumed lane may already publish for the
// NEXT shuffle while later lanes still read this one. No kernel calls are involved (the earlier
// std::thread + std::barrier version spent minutes in futexes).
#include <ucontext.h>
struct EmuWarp {
    ucontext_t sched;
    ucontext_t ctx[32];
    alignas(8) unsigned char slot[2][32][8];
    unsigned char par[32];
    bool waiting[32];
    bool done[32];
    int arrived;
};
inline thread_local EmuWarp* emu_warp = nullptr;
inline thread_local int emu_lane = 0;

template <typename T> inline T emu_exchange(T x, int src) {
    EmuWarp* w = emu_warp;
    const int lane = emu_lane;
    const int par = w->par[lane];
    w->par[lane] ^= 1;
    std::memcpy(w->slot[par][lane], &x, sizeof(T));
    w->arrived++;
    w->waiting[lane] = true;
    swapcontext(&w->ctx[lane], &w->sched);  // resumed once every live lane has published
    T r;
    std::memcpy(&r, w->slot[par][src], sizeof(T));
    return r;
}
template <typename T> inline T __shfl_up_sync(unsigned, T x, int d) {
    return emu_exchange(x, emu_lane - d >= 0 ? emu_lane - d : emu_lane);
}
template <typename T> inline T __shfl_down_sync(unsigned, T x, int d) {
    return emu_exchange(x, emu_lane + d < 32 ? emu_lane + d : emu_lane);
}
template <typename T> inline T __shfl_sync(unsigned, T x, int src) { return emu_exchange(x, src & 31); }

inline float __fadd_rn(float a, float b) { return a + b; }
inline float __fsub_rn(float a, float b) { return a - b; }
inline float __fmul_rn(float a, float b) { return a * b; }
inline float __fdiv_rn(float a, float b) { return a / b; }
inline double __dadd_rn(double a, double b) { return a + b; }
inline double __dsub_rn(double a, double b) { return a - b; }
inline double __dmul_rn(double a, double b) { return a * b; }
inline double __ddiv_rn(double a, double b) { return a / b; }
inline double __fma_rn(double a, double b, double c) { return std::fma(a, b, c); }
inline float __fmaf_rn(float a, float b, float c) { return std::fma(a, b, c); }
inline float __double2float_rn(double a) { return (float)a; }
inline unsigned __float_as_uint(float a) { unsigned b; std::memcpy(&b, &a, 4); return b; }
inline double __hiloint2double(int hi, int lo) { unsigned long long b = ((unsigned long long)(unsigned)hi << 32) | (unsigned)lo; double d; std::memcpy(&d, &b, 8); return d; }
inline int __double2hiint(double a) { long long b; std::memcpy(&b, &a, 8); return (int)((unsigned long long)b >> 32); }
inline int __double2loint(double a) { long long b; std::memcpy(&b, &a, 8); return (int)(b & 0xffffffffLL); }
