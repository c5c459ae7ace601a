// api.cpp -- host side of the C ABI declared in include/fluidb200.h: argument validation,
// strip/chunk planning, the stateless launch entry points and the stateful solver handle.
// Device work goes through backend.h only (kernels.cu in the product).
#include <math.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <new>
#include <chrono>
#include <vector>

#include "../../include/fluidb200.h"
#include "backend.h"
#include "internal.h"

namespace {

thread_local char g_err[512] = "";

int fail(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}
int fail_backend(const char* what) {
    return fail(FLUIDB200_ECUDA, "%s: %s", what, fbk::be_last_error_string());
}
#define BE_CHECK(expr, what)                      \
    do {                                          \
        if ((expr) != 0) return fail_backend(what); \
    } while (0)

int env_int(const char* name, int dflt) {
    const char* s = getenv(name);
    if (!s || !*s) return dflt;
    return atoi(s);
}

template <typename T> constexpr int vec_width() { return 16 / (int)sizeof(T); }
inline int halo_cols(int sweeps) { return (sweeps + 3) & ~3; }

template <typename T> int num_strips(int Y, int H) {
    const int W = 32 * vec_width<T>();
    const int S = W - 2 * H;
    const int n = (Y - H + S - 1) / S;
    return n < 1 ? 1 : n;
}

// Warp tasks aimed for per launch (measured on B200 at 16384^2): the Jacobi pipeline pays K halo rows
// per chunk end and is best around 9 x 2368 tasks up to depth 6 and 5 x 2368 at depth 7-8 (longer
// chunks: fewer recomputed rows); the two one-pass kernels have a 1-2 row halo: the corrector is best
// around 40 x 2368 (0.89 ms), the predictor keeps improving down to 16-row chunks (10 x 2368: 1.76 ms,
// 40 x: 1.14 ms, 64 x: 1.08 ms; its ring depth makes no difference).
constexpr int kJacobiTasks = 148 * 16 * 9, kJacobiTasksDeep = 148 * 16 * 5, kLightTasks = 148 * 16 * 40,
              kPredictorTasks = 148 * 16 * 64;

// Split `n` output rows into chunks. More chunks = more warps in flight and a smaller tail, but
// every chunk recomputes `halo` rows at each end. The chunk count is chosen so that the grid is
// (just under) a whole number of waves over the SMs: `ctas_per_sm` is the occupancy the kernel is
// compiled for, and a partial last wave showed up as a 10-15 % loss at 16384^2. A chunk that
// touches a wall must hold >= 2 rows because the wall row is derived while its neighbour row is
// produced.
void plan_chunks(int n, int halo, int nstrips, int ctas_per_sm, int default_tasks, int* rpc_out, int* nchunks_out, int rpc_hint = 0) {
    const long long target_tasks = env_int("FLUIDB200_TARGET_TASKS", default_tasks);
    int min_rpc = env_int("FLUIDB200_MIN_RPC", 0);
    const long long ctas_x = (nstrips + fb::kWarpsPerCta - 1) / fb::kWarpsPerCta;
    const long long slots = (long long)fbk::be_sm_count() * std::max(1, ctas_per_sm);
    if (min_rpc <= 0) {
        // Large grids: keep the recomputed halo below ~25 % of a chunk. Small grids (not even one wave
        // of CTAs at that chunk length) are latency bound, so trade redundant rows for parallelism.
        min_rpc = std::max(8, 8 * halo);
        if ((long long)((n + min_rpc - 1) / min_rpc) * ctas_x < slots) min_rpc = std::max(4, 2 * halo);
    }
    long long want_chunks = std::max(1LL, (target_tasks + nstrips - 1) / nstrips);
    long long waves = std::max(1LL, (want_chunks * ctas_x + slots / 2) / slots);
    long long aligned = (waves * slots) / ctas_x;  // chunks that fill `waves` waves exactly (rounded down)
    if (aligned >= 1 && !env_int("FLUIDB200_NO_WAVE_ALIGN", 0)) want_chunks = aligned;
    int rpc = (int)((n + want_chunks - 1) / want_chunks);
    rpc = std::max(rpc, min_rpc);
    if (rpc_hint > 0 && target_tasks == default_tasks) rpc = rpc_hint;   // (the caller's measured choice; the task knob overrides it)
    const int fixed = env_int("FLUIDB200_RPC", 0);
    if (fixed > 0) rpc = fixed;
    rpc = std::max(2, std::min(rpc, n));
    while (rpc < n && n % rpc == 1) ++rpc;  // never leave a 1-row tail chunk
    int nchunks = (n + rpc - 1) / rpc;
    if (nchunks > 65535) {
        rpc = (n + 65534) / 65535;
        while (rpc < n && n % rpc == 1) ++rpc;
        nchunks = (n + rpc - 1) / rpc;
    }
    *rpc_out = rpc;
    *nchunks_out = nchunks;
}

int check_slab(const fluidb200_slab* g) {
    if (!g) return fail(FLUIDB200_EINVAL, "slab is NULL");
    if (g->dtype != FLUIDB200_F32 && g->dtype != FLUIDB200_F64)
        return fail(FLUIDB200_EINVAL, "dtype must be FLUIDB200_F32 or FLUIDB200_F64, got %d", g->dtype);
    if (g->X < 3 || g->Y < 3)
        return fail(FLUIDB200_EINVAL, "box must be at least 3x3 (got %dx%d): no interior cells", g->X, g->Y);
    if (g->pitch < g->Y || g->pitch % 32 != 0)
        return fail(FLUIDB200_EINVAL, "pitch %lld must be >= Y=%d and a multiple of 32 elements",
                    (long long)g->pitch, g->Y);
    if (g->row0 < 0 || g->rows < 1 || (long long)g->row0 + g->rows > g->X)
        return fail(FLUIDB200_EINVAL, "local rows [%d,%d) outside the box [0,%d)", g->row0,
                    g->row0 + g->rows, g->X);
    if (!(g->hx > 0) || !(g->hy > 0)) return fail(FLUIDB200_EINVAL, "element lengths must be positive");
    return 0;
}
int check_ptr(const void* p, const char* name) {
    if (!p) return fail(FLUIDB200_EINVAL, "%s is NULL", name);
    if ((uintptr_t)p % 16 != 0) return fail(FLUIDB200_EINVAL, "%s must be 16-byte aligned", name);
    return 0;
}
// rows [lo,hi) to write, needing `halo` extra input rows on each non-wall side
int check_range(const fluidb200_slab* g, int lo, int hi, int halo, const char* what) {
    if (lo < 0 || hi > g->X || lo >= hi)
        return fail(FLUIDB200_EINVAL, "%s: row range [%d,%d) is empty or outside [0,%d)", what, lo, hi, g->X);
    const int need_lo = std::max(lo - halo, 0), need_hi = std::min(hi + halo, g->X);
    if (need_lo < g->row0 || need_hi > g->row0 + g->rows)
        return fail(FLUIDB200_EINVAL, "%s: needs rows [%d,%d) but the slab holds [%d,%d)", what, need_lo,
                    need_hi, g->row0, g->row0 + g->rows);
    if ((lo == 0 && hi < 2) || (hi == g->X && lo > g->X - 2))
        return fail(FLUIDB200_EINVAL, "%s: a range touching a wall must include the row next to it", what);
    return 0;
}

template <typename T> fb::Edges<T> to_dev_edges(const fluidb200_edges* e) {
    fb::Edges<T> d;
    for (int i = 0; i < 4; ++i) {
        d.mode[i] = e ? e->mode[i] : FLUIDB200_EDGE_UNSET;
        d.cval[i] = e ? (T)e->cval[i] : T(0);
        d.val[i] = e ? (const T*)e->val[i] : nullptr;
        d.set[i] = e ? e->set[i] : nullptr;
    }
    return d;
}
int check_edges(const fluidb200_edges* e, const char* name) {
    if (!e) return fail(FLUIDB200_EINVAL, "%s is NULL", name);
    for (int i = 0; i < 4; ++i) {
        if (e->mode[i] < 0 || e->mode[i] > 2)
            return fail(FLUIDB200_EINVAL, "%s: edge %d has bad mode %d", name, i, e->mode[i]);
        if (e->mode[i] == FLUIDB200_EDGE_ARRAY && (!e->val[i] || !e->set[i]))
            return fail(FLUIDB200_EINVAL, "%s: edge %d is ARRAY but val/set is NULL", name, i);
    }
    return 0;
}
fb::Geom make_geom(const fluidb200_slab* g, int nstrips, int rpc) {
    fb::Geom q;
    q.pitch = g->pitch;
    q.X = g->X;
    q.Y = g->Y;
    q.row0 = g->row0;
    q.rows = g->rows;
    q.nstrips = nstrips;
    q.rows_per_chunk = rpc;
    return q;
}

template <typename T>
int predictor_t(const fluidb200_slab* g, const void* u, const void* v, void* us, void* vs, void* rhs,
                double dt, double nu, double rho, int star_lo, int star_hi, int rhs_lo, int rhs_hi,
                void* stream) {
    fb::PredictorArgs<T> a;
    a.u = (const T*)u; a.v = (const T*)v;
    a.us = (T*)us; a.vs = (T*)vs; a.rhs = (T*)rhs;
    a.m = fb::make_metric<T>((T)g->hx, (T)g->hy, env_int("FLUIDB200_DIVMODE", -1));
    a.nu = (T)nu; a.dt = (T)dt;
    a.rho_over_dt = (T)rho / (T)dt;  // (density / time_step) evaluated in T, SerialNavierStokes.cpp:332
    a.star_lo = star_lo; a.star_hi = star_hi; a.rhs_lo = rhs_lo; a.rhs_hi = rhs_hi;
    a.lo = std::min(star_lo, rhs_lo);
    a.hi = std::max(star_hi, rhs_hi);
    const int nstrips = num_strips<T>(g->Y, 4);
    int rpc, nchunks;
    plan_chunks(a.hi - a.lo, 2, nstrips, 4, env_int("FLUIDB200_LIGHT_TASKS", kPredictorTasks), &rpc, &nchunks);
    a.g = make_geom(g, nstrips, rpc);
    BE_CHECK(fbk::launch_predictor<T>(a, nchunks, stream), "predictor launch");
    return 0;
}

// Decide which tasks of a Jacobi launch may go to the lean kernels (jacobi_fast_task / jacobi_edge_task):
// every step of such a task must be a plain interior step in x, i.e.
//   chunks [cf_lo, cf_hi): pipeline starts at c0-K >= max(row0, K+2) and ends at c1-1+K <= X-2,
//                          and the chunk spans fewer than 2^31 elements (32-bit running offsets);
//   strips [sf_lo, sf_hi): no y-edge column inside the strip's 32*V loaded columns (FAST kernel);
//                          the other strips of those chunks go to the EDGE kernel.
// The remaining (wall) chunks stay with the general kernel. part = 1 asks the backend for the split
// launch, part = 0 for the single general kernel.
template <typename T>
void split_jacobi(fb::JacobiArgs<T>* a, const fluidb200_slab* g, int K, int nchunks, bool use_mask) {
    a->part = 0;
    a->nchunks = nchunks;
    a->sf_lo = a->sf_hi = a->cf_lo = a->cf_hi = 0;
    a->edge_lean = 0;
    a->gen_all = 0;
    if ((use_mask && !a->mtiles) || env_int("FLUIDB200_NO_SPLIT", 0)) return;
    const int W = 32 * vec_width<T>(), S = W - 2 * halo_cols(K);
    const int nstrips = a->g.nstrips;
    int sf_lo = 1, sf_hi = 1;
    while (sf_hi < nstrips && sf_hi * S + W <= g->Y - 1) ++sf_hi;  // strip sf_hi is the first that holds column Y-1
    int cf_lo = nchunks, cf_hi = 0;
    for (int c = 0; c < nchunks; ++c) {
        int c0, c1;
        fb::jacobi_chunk_rows(*a, c, &c0, &c1);
        const bool ok = c0 < c1 && c0 - K >= std::max(g->row0, K + 2) && c1 - 1 + K <= g->X - 2 &&
                        c1 + K <= g->row0 + g->rows && (long long)(c1 - c0 + 2 * K + 1) * g->pitch < (1LL << 31);
        if (ok) {
            cf_lo = std::min(cf_lo, c);
            cf_hi = std::max(cf_hi, c + 1);
        } else if (cf_hi > 0) {
            break;  // fast chunks are one contiguous run
        }
    }
    if (sf_hi - sf_lo < 1 || cf_hi - cf_lo < 1) return;
    // a launch that cannot fill the GPU a few times over is latency bound: the extra kernels and the
    // fork/join events cost more than the extra occupancy brings (measured, us per depth-3 launch,
    // split vs single kernel: 1000^2 18.0 vs 14.8, 1536^2 19.1 vs 19.3, 2048^2 21.7 vs 27.3)
    // Exception: a range inside a SLAB of a wide grid (the halo strips of the multi-GPU drivers, swept beside the
    // slab's big interior launch). There the general kernel's 218 registers are the cost: one of its CTAs on an SM
    // pushes two of the interior launch's lean CTAs (3 x 162 registers fill the file) off it, whereas lean CTAs of the
    // same kind pack with them. Measured on a 4096 x 32768 slab: 313 us per launch with ~40 us of that displacement.
    const bool slab_strip = nstrips >= 32 && (g->row0 > 0 || g->row0 + g->rows < g->X);
    if ((long long)(sf_hi - sf_lo) * (cf_hi - cf_lo) < (long long)env_int("FLUIDB200_SPLIT_MIN_TASKS", 2048) && !slab_strip) return;
    a->part = 1;
    a->sf_lo = sf_lo; a->sf_hi = sf_hi; a->cf_lo = cf_lo; a->cf_hi = cf_hi;
    // the lean EDGE kernel knows constant and copy boundary cells; per-cell arrays stay with the general kernel
    a->edge_lean = a->e.mode[fb::BOTTOM] != fb::EDGE_ARRAY && a->e.mode[fb::TOP] != fb::EDGE_ARRAY;
    a->gen_all = use_mask ? 1 : 0;  // masked: the general kernel is offered every task (see jacobi_general_task)
}

template <typename T>
int poisson_t(const fluidb200_slab* g, const void* p_in, void* p_out, const void* rhs,
              const fluidb200_edges* e, const fluidb200_mask* mask, int sweeps, int out_lo, int out_hi,
              void* stream, bool allow_split) {
    fb::JacobiArgs<T> a;
    a.p_in = (const T*)p_in; a.p_out = (T*)p_out; a.rhs = (const T*)rhs;
    const bool use_mask = mask && mask->set;
    a.mset = use_mask ? mask->set : nullptr;
    a.mval = use_mask ? (const T*)mask->val : nullptr;
    a.mtiles = use_mask ? mask->tiles : nullptr;
    a.mtiles_nc = (g->Y + fb::kMaskTileCols - 1) / fb::kMaskTileCols;
    a.hx = (T)g->hx; a.hy = (T)g->hy;
    a.out_lo = out_lo; a.out_hi = out_hi;
    a.e = to_dev_edges<T>(e);
    const int nstrips = num_strips<T>(g->Y, halo_cols(sweeps));
    int rpc, nchunks;
    // resident CTAs per SM of the kernel that does the bulk of the work (the lean FAST kernel)
    const int ctas_per_sm = sweeps <= 2 ? 5 : sweeps <= 6 ? 4 : 3;
    // Waves of tasks: few on a short range (every chunk recomputes 2K halo rows, so short chunks cost more than a
    // ragged last wave does), the measured optimum of the tall grid from about 13000 rows up. Depth 8 at 32768 columns,
    // us per launch with 2 / 3 / 5 / 7 waves: 4096 rows 315 / 315 / 331 / 340; 16384 rows (16384 columns) - / 590 / 584 / 591.
    const int nrows = out_hi - out_lo;
    const int tasks = sweeps >= 7 ? 148 * 16 * std::max(2, std::min(5, nrows / 2700)) : 148 * 16 * std::max(3, std::min(9, nrows / 1800));
    // Deep launches on big ranges, from scans of the rows per chunk (profiles/README_r2.md section 5; tools/scan_rpc.sh): a grid
    // that is an exact number of waves is the WORST choice there, because the wall and y-edge kernels underneath take slots
    // from the first wave and push the last CTAs into one more. (1) Tall ranges: chunks of ~270 rows (6 % of recomputed halo
    // rows) once that makes at least 4.4 waves: 16384^2 544-550 us against 563 at exactly 7 waves, 32768 x 8192 556 against
    // 580. (2) Short, wide ranges (the 4096 x 32768 slab of 8 GPUs): 4.4 waves if the chunks stay >= 150 rows: 305 against 315.
    int rpc_hint = 0;
    if (sweeps >= 7) {
        const double slots = (double)fbk::be_sm_count() * ctas_per_sm, ctas_x = (nstrips + fb::kWarpsPerCta - 1) / fb::kWarpsPerCta;
        const int inner = nrows - (out_lo == 0 ? 2 * sweeps + 2 : 0) - (out_hi == g->X ? sweeps + 1 : 0);
        const int by_len = (inner + 269) / 270;
        const int by_waves = (int)(4.4 * slots / ctas_x);
        if (by_len * ctas_x >= 4.4 * slots) rpc_hint = (inner + by_len - 1) / by_len;
        else if (by_waves >= 1 && inner / by_waves >= 150) rpc_hint = (inner + by_waves - 1) / by_waves;
    }
    // First try: the chunks at an x wall as short as the pipeline allows (the next chunk must start
    // at row >= 2K+2, the previous one end at row <= X-K-2), because only the register-heavy general
    // kernel can run them; the rows between are chunked uniformly and go to the lean kernels.
    const int n = out_hi - out_lo;
    a.head = out_lo == 0 ? 2 * sweeps + 2 : 0;
    a.tail = out_hi == g->X ? sweeps + 1 : 0;
    a.part = 0;
    if (allow_split && (!use_mask || a.mtiles) && (a.head || a.tail) && n - a.head - a.tail >= 4 * sweeps + 8) {
        plan_chunks(n - a.head - a.tail, sweeps, nstrips, ctas_per_sm, tasks, &rpc, &nchunks, rpc_hint);
        nchunks += (a.head > 0) + (a.tail > 0);
        a.g = make_geom(g, nstrips, rpc);
        split_jacobi<T>(&a, g, sweeps, nchunks, use_mask);
    }
    if (a.part == 0) {  // no split (small grid, masked sweep, no wall in range): plain uniform chunks
        a.head = a.tail = 0;
        plan_chunks(n, sweeps, nstrips, ctas_per_sm, tasks, &rpc, &nchunks, rpc_hint);
        a.g = make_geom(g, nstrips, rpc);
        split_jacobi<T>(&a, g, sweeps, nchunks, use_mask);
        if (!allow_split) a.part = 0;
    }
    BE_CHECK(fbk::launch_jacobi<T>(a, sweeps, use_mask, nchunks, stream), "poisson launch");
    return 0;
}

template <typename T>
int corrector_t(const fluidb200_slab* g, const void* p, const void* us, const void* vs, void* u, void* v,
                const fluidb200_edges* eu, const fluidb200_edges* ev, const fluidb200_mask* mu,
                const fluidb200_mask* mv, double dt, double rho, int out_lo, int out_hi, void* stream) {
    fb::CorrectorArgs<T> a;
    a.p = (const T*)p; a.us = (const T*)us; a.vs = (const T*)vs;
    a.u = (T*)u; a.v = (T*)v;
    const bool um = mu && mu->set, vm = mv && mv->set;
    a.uset = um ? mu->set : nullptr; a.uval = um ? (const T*)mu->val : nullptr;
    a.vset = vm ? mv->set : nullptr; a.vval = vm ? (const T*)mv->val : nullptr;
    a.m = fb::make_metric<T>((T)g->hx, (T)g->hy, env_int("FLUIDB200_DIVMODE", -1));
    a.dt_over_rho = (T)dt / (T)rho;  // (time_step / density) in T, SerialNavierStokes.cpp:346
    a.out_lo = out_lo; a.out_hi = out_hi;
    a.eu = to_dev_edges<T>(eu);
    a.ev = to_dev_edges<T>(ev);
    const int nstrips = num_strips<T>(g->Y, 4);
    int rpc, nchunks;
    plan_chunks(out_hi - out_lo, 1, nstrips, 4, env_int("FLUIDB200_LIGHT_TASKS", kLightTasks), &rpc, &nchunks);
    a.g = make_geom(g, nstrips, rpc);
    BE_CHECK(fbk::launch_corrector<T>(a, um || vm, nchunks, stream), "corrector launch");
    return 0;
}

}  // namespace

// =========================================================================================
// stateless entry points
// =========================================================================================
extern "C" int fluidb200_predictor(const fluidb200_slab* g, const void* u, const void* v, void* u_next,
                                   void* v_next, void* rhs, double dt, double nu, double rho, int star_lo,
                                   int star_hi, int rhs_lo, int rhs_hi, void* stream) {
    int rc;
    if ((rc = check_slab(g))) return rc;
    if ((rc = check_ptr(u, "u")) || (rc = check_ptr(v, "v")) || (rc = check_ptr(u_next, "u_next")) ||
        (rc = check_ptr(v_next, "v_next")) || (rc = check_ptr(rhs, "rhs")))
        return rc;
    if (!(dt != 0)) return fail(FLUIDB200_EINVAL, "time_step must be non-zero");
    if ((rc = check_range(g, rhs_lo, rhs_hi, 2, "predictor rhs"))) return rc;
    if ((rc = check_range(g, star_lo, star_hi, 1, "predictor u_next"))) return rc;
    return g->dtype == FLUIDB200_F32
               ? predictor_t<float>(g, u, v, u_next, v_next, rhs, dt, nu, rho, star_lo, star_hi, rhs_lo, rhs_hi, stream)
               : predictor_t<double>(g, u, v, u_next, v_next, rhs, dt, nu, rho, star_lo, star_hi, rhs_lo, rhs_hi, stream);
}

extern "C" int fluidb200_poisson_sweeps(const fluidb200_slab* g, const void* p_in, void* p_out,
                                        const void* rhs, const fluidb200_edges* p_edges,
                                        const fluidb200_mask* mask, int sweeps, int out_lo, int out_hi,
                                        void* stream) {
    return fbapi::poisson_sweeps(g, p_in, p_out, rhs, p_edges, mask, sweeps, out_lo, out_hi, stream, 0);
}

int fbapi::poisson_sweeps(const fluidb200_slab* g, const void* p_in, void* p_out, const void* rhs,
                          const fluidb200_edges* p_edges, const fluidb200_mask* mask, int sweeps, int out_lo,
                          int out_hi, void* stream, int flags) {
    int rc;
    if ((rc = check_slab(g))) return rc;
    if ((rc = check_ptr(p_in, "p_in")) || (rc = check_ptr(p_out, "p_out")) || (rc = check_ptr(rhs, "rhs")))
        return rc;
    if (p_in == p_out) return fail(FLUIDB200_EINVAL, "p_out must not alias p_in (Jacobi ping-pong)");
    if ((rc = check_edges(p_edges, "p_edges"))) return rc;
    if (sweeps < 1 || sweeps > FLUIDB200_MAX_SWEEPS_PER_LAUNCH)
        return fail(FLUIDB200_EINVAL, "sweeps per launch must be in [1,%d], got %d",
                    FLUIDB200_MAX_SWEEPS_PER_LAUNCH, sweeps);
    if (mask && mask->set) {
        if (!mask->val) return fail(FLUIDB200_EINVAL, "mask.val is NULL");
        if ((uintptr_t)mask->set % 4 != 0 || (uintptr_t)mask->val % 16 != 0)
            return fail(FLUIDB200_EINVAL, "mask.set must be 4-byte and mask.val 16-byte aligned");
    }
    if ((rc = check_range(g, out_lo, out_hi, sweeps, "poisson"))) return rc;
    if (flags & POISSON_VALIDATE_ONLY) return 0;
    const bool allow_split = !(flags & POISSON_NO_SPLIT);
    return g->dtype == FLUIDB200_F32
               ? poisson_t<float>(g, p_in, p_out, rhs, p_edges, mask, sweeps, out_lo, out_hi, stream, allow_split)
               : poisson_t<double>(g, p_in, p_out, rhs, p_edges, mask, sweeps, out_lo, out_hi, stream, allow_split);
}

extern "C" int fluidb200_corrector(const fluidb200_slab* g, const void* p, const void* u_next,
                                   const void* v_next, void* u, void* v, const fluidb200_edges* u_edges,
                                   const fluidb200_edges* v_edges, const fluidb200_mask* u_mask,
                                   const fluidb200_mask* v_mask, double dt, double rho, int out_lo,
                                   int out_hi, void* stream) {
    int rc;
    if ((rc = check_slab(g))) return rc;
    if ((rc = check_ptr(p, "p")) || (rc = check_ptr(u_next, "u_next")) || (rc = check_ptr(v_next, "v_next")) ||
        (rc = check_ptr(u, "u")) || (rc = check_ptr(v, "v")))
        return rc;
    if ((rc = check_edges(u_edges, "u_edges")) || (rc = check_edges(v_edges, "v_edges"))) return rc;
    if (!(rho != 0)) return fail(FLUIDB200_EINVAL, "density must be non-zero");
    if ((u_mask && u_mask->set && !u_mask->val) || (v_mask && v_mask->set && !v_mask->val))
        return fail(FLUIDB200_EINVAL, "mask.val is NULL");
    if ((rc = check_range(g, out_lo, out_hi, 1, "corrector"))) return rc;
    return g->dtype == FLUIDB200_F32
               ? corrector_t<float>(g, p, u_next, v_next, u, v, u_edges, v_edges, u_mask, v_mask, dt, rho, out_lo, out_hi, stream)
               : corrector_t<double>(g, p, u_next, v_next, u, v, u_edges, v_edges, u_mask, v_mask, dt, rho, out_lo, out_hi, stream);
}

extern "C" int fluidb200_init_cavity(const fluidb200_slab* g, void* u, void* v, void* p, void* u_next,
                                     void* v_next, void* p_next, void* rhs, double lid_velocity,
                                     void* stream) {
    int rc;
    if ((rc = check_slab(g))) return rc;
    void* f[7] = {u, v, p, u_next, v_next, p_next, rhs};
    for (int i = 0; i < 7; ++i)
        if (f[i] && (uintptr_t)f[i] % 16 != 0) return fail(FLUIDB200_EINVAL, "field %d must be 16-byte aligned", i);
    if (g->dtype == FLUIDB200_F32) {
        float* ff[7];
        for (int i = 0; i < 7; ++i) ff[i] = (float*)f[i];
        BE_CHECK(fbk::launch_init_cavity<float>(ff, g->pitch, g->X, g->Y, g->row0, g->rows, (float)lid_velocity, stream),
                 "init_cavity launch");
    } else {
        double* ff[7];
        for (int i = 0; i < 7; ++i) ff[i] = (double*)f[i];
        BE_CHECK(fbk::launch_init_cavity<double>(ff, g->pitch, g->X, g->Y, g->row0, g->rows, lid_velocity, stream),
                 "init_cavity launch");
    }
    return 0;
}

extern "C" int fluidb200_cavity_edges(char field, double lid_velocity, fluidb200_edges* out) {
    if (!out) return fail(FLUIDB200_EINVAL, "out is NULL");
    memset(out, 0, sizeof(*out));
    if (field == 'u' || field == 'v') {
        for (int i = 0; i < 4; ++i) out->mode[i] = FLUIDB200_EDGE_CONST;
        out->cval[FLUIDB200_EDGE_TOP] = (field == 'u') ? lid_velocity : 0.0;
    } else if (field == 'p') {
        out->mode[FLUIDB200_EDGE_TOP] = FLUIDB200_EDGE_CONST;
    } else {
        return fail(FLUIDB200_EINVAL, "field must be 'u', 'v' or 'p'");
    }
    return 0;
}

// =========================================================================================
// stateful solver handle
// =========================================================================================
struct fluidb200_solver {
    int dtype = 0, X = 0, Y = 0;
    double wx = 0, wy = 0;
    double hx = 0, hy = 0;           // exact T values
    double dt = 0.001, nu = 0.1, rho = 1.0;
    int K = 50, S = 1000;            // NavierStokesSolver.h:22-26 defaults
    int64_t pitch = 0;
    size_t esize = 4;
    void* f[FLUIDB200_NUM_FIELDS] = {};
    bool p_next_is_p = false;        // after any sweep the reference has p_next == p
    fluidb200_edges edges[3] = {};   // u, v, p
    void* edge_val[3][4] = {};
    uint8_t* edge_set[3][4] = {};
    fluidb200_mask mask[3] = {};     // interior flags (device), NULL when none
    bool mask_sparse = false;        // pressure flags touch few tiles: masked launches keep the full blocking depth
    void* stream = nullptr;
    bool own_stream = false;
    int kmax = 4;
    bool use_graph = false;
    void* graph[2] = {nullptr, nullptr};  // one timestep, for each ping-pong parity
    int64_t graph_kernels[2] = {0, 0};    // kernels recorded in each
    double graph_key[6] = {};
    void* graph_home_p = nullptr;     // where p lived when graph[0] was captured
    int64_t launches = 0;
    // resident solver for small grids (resident_kernels.cuh)
    int resident = -1;                // -1: by FLUIDB200_RESIDENT (default on), 0 / 1: fluidb200_set_resident
    bool res_planned = false, res_ok = false;
    fb::ResidentPlan res_plan = {};
    void* res_scratch = nullptr;      // device: tile counters + 4 values for the corner rule
    // chunk pipeline of fluidb200_migrate_solve: copy streams, events, output arrays, a chunk's pressure ping-pong pair
    void* pipe_up = nullptr;
    void* pipe_down = nullptr;
    void* pipe_up2 = nullptr;     // second copy stream per direction (chunks alternate: two DMA engines per direction)
    void* pipe_down2 = nullptr;
    void* pipe_out[2] = {nullptr, nullptr};   // new u, v (whole grid): the inputs must stay intact for the next chunk's halo
    void* pipe_p[2] = {nullptr, nullptr};
    size_t pipe_p_rows = 0;
    std::vector<void*> pipe_ev;
    double pipe_issue_ms = 0;   // host time the last pipelined migrate_solve spent issuing its work (before waiting for it)
    int64_t resident_launches = 0;
};

namespace {

size_t field_bytes(const fluidb200_solver* s) { return (size_t)s->X * (size_t)s->pitch * s->esize; }

void fill_slab(const fluidb200_solver* s, fluidb200_slab* g) {
    g->dtype = s->dtype; g->X = s->X; g->Y = s->Y; g->pitch = s->pitch;
    g->row0 = 0; g->rows = s->X; g->hx = s->hx; g->hy = s->hy;
}

void drop_graphs(fluidb200_solver* s) {
    for (int i = 0; i < 2; ++i)
        if (s->graph[i]) { fbk::be_graph_destroy(s->graph[i]); s->graph[i] = nullptr; }
}

void free_bc(fluidb200_solver* s, int f) {
    for (int e = 0; e < 4; ++e) {
        if (s->edge_val[f][e]) fbk::be_free(s->edge_val[f][e]);
        if (s->edge_set[f][e]) fbk::be_free(s->edge_set[f][e]);
        s->edge_val[f][e] = nullptr; s->edge_set[f][e] = nullptr;
        s->edges[f].mode[e] = FLUIDB200_EDGE_UNSET; s->edges[f].cval[e] = 0;
        s->edges[f].val[e] = nullptr; s->edges[f].set[e] = nullptr;
    }
    if (s->mask[f].set) fbk::be_free((void*)s->mask[f].set);
    if (s->mask[f].val) fbk::be_free((void*)s->mask[f].val);
    if (s->mask[f].tiles) fbk::be_free((void*)s->mask[f].tiles);
    s->mask[f].set = nullptr; s->mask[f].val = nullptr; s->mask[f].tiles = nullptr;
}

// Turn dense host flag/value arrays of one field into the edge description (+ interior mask).
template <typename T>
int upload_bc_t(fluidb200_solver* s, int f, const T* val, const uint8_t* set) {
    const int X = s->X, Y = s->Y;
    free_bc(s, f);
    drop_graphs(s);
    for (int e = 0; e < 4; ++e) {
        const bool along_x = (e == FLUIDB200_EDGE_BOTTOM || e == FLUIDB200_EDGE_TOP);
        const int len = along_x ? X : Y;
        const int first = along_x ? 0 : 1, last = along_x ? X - 1 : Y - 2;  // entries the kernels use
        std::vector<T> ev(len, T(0));
        std::vector<uint8_t> es(len, 0);
        for (int i = 0; i < len; ++i) {
            size_t idx;
            if (e == FLUIDB200_EDGE_BOTTOM) idx = (size_t)i * Y;
            else if (e == FLUIDB200_EDGE_TOP) idx = (size_t)i * Y + (Y - 1);
            else if (e == FLUIDB200_EDGE_LEFT) idx = (size_t)i;
            else idx = (size_t)(X - 1) * Y + i;
            es[i] = set[idx] ? 1 : 0;
            ev[i] = val[idx];
        }
        bool any = false, all = true, same = true;
        for (int i = first; i <= last; ++i) {
            any = any || es[i];
            all = all && es[i];
            if (memcmp(&ev[i], &ev[first], sizeof(T)) != 0) same = false;
        }
        if (!any) {
            s->edges[f].mode[e] = FLUIDB200_EDGE_UNSET;
        } else if (all && same) {
            s->edges[f].mode[e] = FLUIDB200_EDGE_CONST;
            s->edges[f].cval[e] = (double)ev[first];
        } else {
            void *dv = nullptr, *ds = nullptr;
            BE_CHECK(fbk::be_malloc(&dv, len * sizeof(T)), "edge alloc");
            s->edge_val[f][e] = dv;
            BE_CHECK(fbk::be_malloc(&ds, len), "edge alloc");
            s->edge_set[f][e] = (uint8_t*)ds;
            BE_CHECK(fbk::be_memcpy_async(dv, ev.data(), len * sizeof(T), fbk::H2D, s->stream), "edge upload");
            BE_CHECK(fbk::be_memcpy_async(ds, es.data(), len, fbk::H2D, s->stream), "edge upload");
            BE_CHECK(fbk::be_stream_sync(s->stream), "edge upload");  // host vectors go out of scope
            s->edges[f].mode[e] = FLUIDB200_EDGE_ARRAY;
            s->edges[f].val[e] = dv;
            s->edges[f].set[e] = (const uint8_t*)ds;
        }
    }
    // interior flags -> dense pitched mask (the general-BC path)
    bool interior = false;
    for (int x = 1; x < X - 1 && !interior; ++x)
        for (int y = 1; y < Y - 1; ++y)
            if (set[(size_t)x * Y + y]) { interior = true; break; }
    if (interior) {
        void *ds = nullptr, *dv = nullptr;
        const size_t n = (size_t)X * s->pitch;
        BE_CHECK(fbk::be_malloc(&ds, n), "mask alloc");
        s->mask[f].set = (const uint8_t*)ds;
        BE_CHECK(fbk::be_malloc(&dv, n * sizeof(T)), "mask alloc");
        s->mask[f].val = dv;
        BE_CHECK(fbk::be_memset_async(ds, 0, n, s->stream), "mask clear");
        BE_CHECK(fbk::be_memset_async(dv, 0, n * sizeof(T), s->stream), "mask clear");
        BE_CHECK(fbk::be_memcpy2d_async(ds, s->pitch, set, Y, Y, X, fbk::H2D, s->stream), "mask upload");
        BE_CHECK(fbk::be_memcpy2d_async(dv, s->pitch * sizeof(T), val, Y * sizeof(T), Y * sizeof(T), X, fbk::H2D,
                                        s->stream), "mask upload");
        if (f == 2 && !env_int("FLUIDB200_NO_MASK_TILES", 0)) {
            // tile summary of the interior pressure flags: lets a masked Jacobi launch keep its lean kernels
            // wherever there is no flag (see fluidb200_mask::tiles)
            const int TR = FLUIDB200_MASK_TILE_ROWS, TC = FLUIDB200_MASK_TILE_COLS;
            const int ntr = (X + TR - 1) / TR, ntc = (Y + TC - 1) / TC;
            std::vector<uint8_t> tiles((size_t)ntr * ntc, 0);
            size_t flagged = 0;
            for (int x = 1; x < X - 1; ++x) {
                const uint8_t* row = set + (size_t)x * Y;
                uint8_t* trow = tiles.data() + (size_t)(x / TR) * ntc;
                for (int y = 1; y < Y - 1; ++y)
                    if (row[y]) trow[y / TC] = 1;
            }
            for (uint8_t t : tiles) flagged += t;
            s->mask_sparse = flagged * 20 <= tiles.size();  // flags in at most 5 % of the tiles
            void* dt = nullptr;
            BE_CHECK(fbk::be_malloc(&dt, tiles.size()), "mask tiles alloc");
            s->mask[f].tiles = (const uint8_t*)dt;
            BE_CHECK(fbk::be_memcpy_async(dt, tiles.data(), tiles.size(), fbk::H2D, s->stream), "mask tiles upload");
        }
        BE_CHECK(fbk::be_stream_sync(s->stream), "mask upload");
    }
    return 0;
}

int step_predictor(fluidb200_solver* s) {
    fluidb200_slab g;
    fill_slab(s, &g);
    const long long k0 = fbk::be_kernel_count();
    int rc = fluidb200_predictor(&g, s->f[FLUIDB200_FIELD_U], s->f[FLUIDB200_FIELD_V], s->f[FLUIDB200_FIELD_U_NEXT],
                                 s->f[FLUIDB200_FIELD_V_NEXT], s->f[FLUIDB200_FIELD_RHS], s->dt, s->nu, s->rho, 0,
                                 s->X, 0, s->X, s->stream);
    s->launches += fbk::be_kernel_count() - k0;
    return rc;
}

// Blocking depth when interior pressure flags are present. Tasks with a flagged cell in their footprint
// run the general path, which is instruction and latency bound (no row prefetch), so a deep pipeline
// only makes those few tasks the critical path of the launch. Measured at 8192^2, us per sweep:
//   no tile summary (every task general): depth 1 255, depth 2 192, depth 4 203, depth 8 366 -> 2;
//   tile summary, one 64 x 64 obstacle  : depth 2 76, depth 4 41, depth 6 40, depth 8 41 -> 4
//   (the same grid without flags: 23 at depth 8).
constexpr int kMaskedKmaxDefault = 2, kSparseMaskedKmax = 4;
int masked_kmax() { return std::max(1, env_int("FLUIDB200_MASK_KMAX", kMaskedKmaxDefault)); }
// flags in few tiles (and FLUIDB200_MASK_KMAX not set: tests and measurements force a depth with it)
bool sparse_mask(const fluidb200_solver* s) { return s->mask_sparse && s->mask[2].tiles && getenv("FLUIDB200_MASK_KMAX") == nullptr; }
int poisson_kmax(const fluidb200_solver* s) {
    int k = s->kmax;
    if (s->mask[2].set) k = std::min(k, sparse_mask(s) ? kSparseMaskedKmax : masked_kmax());
    return std::max(1, std::min(k, FLUIDB200_MAX_SWEEPS_PER_LAUNCH));
}

int step_poisson(fluidb200_solver* s, int sweeps) {
    fluidb200_slab g;
    fill_slab(s, &g);
    const bool masked = s->mask[2].set != nullptr;
    const int kmax = poisson_kmax(s);
    // n = ceil(sweeps / kmax) launches of nearly equal depth (100 sweeps at depth 7: 10 x 7 + 5 x 6)
    // rather than full launches and a shallow remainder: a launch costs about the same at any depth.
    const int nlaunch = (sweeps + kmax - 1) / kmax;
    int left = sweeps;
    const long long k0 = fbk::be_kernel_count();
    struct Count {   // kernels of this call, however it returns
        fluidb200_solver* s; long long k0;
        ~Count() { s->launches += fbk::be_kernel_count() - k0; }
    } count{s, k0};
    for (int i = 0; i < nlaunch; ++i) {
        const int k = (left + (nlaunch - i) - 1) / (nlaunch - i);
        int rc = fluidb200_poisson_sweeps(&g, s->f[FLUIDB200_FIELD_P], s->f[FLUIDB200_FIELD_P_NEXT],
                                          s->f[FLUIDB200_FIELD_RHS], &s->edges[2], masked ? &s->mask[2] : nullptr, k, 0,
                                          s->X, s->stream);
        if (rc) return rc;
        std::swap(s->f[FLUIDB200_FIELD_P], s->f[FLUIDB200_FIELD_P_NEXT]);  // ping-pong instead of p = p_next
        s->p_next_is_p = true;
        left -= k;
    }
    return 0;
}

int step_corrector(fluidb200_solver* s) {
    fluidb200_slab g;
    fill_slab(s, &g);
    const long long k0 = fbk::be_kernel_count();
    int rc = fluidb200_corrector(&g, s->f[FLUIDB200_FIELD_P], s->f[FLUIDB200_FIELD_U_NEXT], s->f[FLUIDB200_FIELD_V_NEXT],
                                 s->f[FLUIDB200_FIELD_U], s->f[FLUIDB200_FIELD_V], &s->edges[0], &s->edges[1],
                                 s->mask[0].set ? &s->mask[0] : nullptr, s->mask[1].set ? &s->mask[1] : nullptr, s->dt,
                                 s->rho, 0, s->X, s->stream);
    s->launches += fbk::be_kernel_count() - k0;
    return rc;
}

int one_step(fluidb200_solver* s) {
    int rc;
    if ((rc = step_predictor(s))) return rc;
    if ((rc = step_poisson(s, s->K))) return rc;
    return step_corrector(s);
}

// ---- resident solver: whole timesteps of a small grid in one cooperative launch -----------------
// Tiling: a thread owns R rows x 16 bytes, a warp 32 such patches side by side, a CTA wx x wy warps,
// the grid tiles_x x tiles_y CTAs, at most one per SM (they wait for each other). A sweep costs about
// R rows of arithmetic per warp, times the warps sharing a scheduler, plus -- once tiles exist -- a
// round trip through L2 to the neighbouring tile, which is worth roughly 1000 cycles: so a grid that
// fits one CTA stays in one CTA, and a larger one spreads as widely as it can.
template <typename T> bool plan_resident(int X, int Y, fb::ResidentPlan* out) {
    const int V = 16 / (int)sizeof(T), W = 32 * V;
    const int nsm = fbk::be_sm_count();
    const int ncw = (Y + W - 1) / W;
    const int fR = env_int("FLUIDB200_RESIDENT_R", 0), fwy = env_int("FLUIDB200_RESIDENT_WY", 0),
              fwx = env_int("FLUIDB200_RESIDENT_WX", 0);   // (measurement knobs)
    // a grid of at most 4096 cells: one CTA, a cell per thread and slot, everything in shared memory
    // (FLUIDB200_RESIDENT_R = 4 or 8 forces the patch layout instead)
    if (!fR && (long long)X * Y <= 4 * 1024 && fbk::be_resident_max_threads(0) > 0) {
        const int n = X * Y;
        const int nc = env_int("FLUIDB200_CELL_NC", n <= 2048 ? 2 : 4);   // cells per thread (2 or 4; a measurement knob)
        // just the warps the cells need (41^2: 27 of them, 17.6 ms per 1000 x 50 against 18.3 with all 32; 4 cells per
        // thread on 14 warps: 19.9)
        const int threads = std::min(fbk::be_resident_max_threads(0), ((n + nc - 1) / nc + 31) / 32 * 32);
        if ((nc == 2 || nc == 4) && (long long)threads * nc >= n) {
            out->R = 0; out->wy = nc; out->wx = 1; out->tiles_y = out->tiles_x = 1;
            out->threads = threads;
            out->smem = (size_t)(2 * (n + 4) + 4 * n) * sizeof(T);
            return true;
        }
    }
    double best = 1e300;
    bool found = false;
    const int Rs[2] = {8, 4};
    for (int R : Rs) {
        if (fR && R != fR) continue;
        const int maxw = fbk::be_resident_max_threads(R) / 32;
        if (maxw <= 0) continue;
        const int nrw = (X + R - 1) / R;
        for (int wy = 1; wy <= std::min(ncw, maxw); ++wy) {
            if (fwy && wy != fwy) continue;
            const int ty = (ncw + wy - 1) / wy;
            for (int wx = 1; wx * wy <= maxw && wx <= nrw; ++wx) {
                if (fwx && wx != fwx) continue;
                const int tx = (nrw + wx - 1) / wx;
                if ((long long)ty * tx > nsm) continue;
                const size_t smem = fb::resident_smem_bytes<T>(R, wy, wx);
                if (smem > 220u * 1024u) continue;
                const int warps = wx * wy;
                // Cycles per sweep, roughly: the arithmetic of the warps sharing a scheduler, plus -- once there are several
                // tiles -- what crosses them through L2: ~720 cycles per polling round a warp makes one after the other (the
                // two rows are polled together; a column crossing tiles is one more round, two when a warp has both), and
                // ~0.62 cycles per value of the CTA's borders. Fitted to forced tilings of 1000^2 (8 x 1024-cell tiles 99.5 ms
                // per 1000 x 50, 16 x 512 102.3, 32 x 256 94.5, 64 x 128 102.6) and checked on 1000 x 500 and 800 x 1200
                // (profiles/README_r2.md section 2).
                const int TC = wy * W, TR = wx * R;
                const double rounds = (tx > 1 ? 1.0 : 0.0) + (ty > 1 ? (wy == 1 ? 2.0 : 1.0) : 0.0);
                const double values = (tx > 1 ? 2.0 * TC : 0.0) + (ty > 1 ? 2.0 * TR : 0.0);
                const double cost = (R * 20.0 + 40.0) * ((warps + 3) / 4) + 720.0 * rounds + 0.62 * values;
                if (cost < best) {   // (ties keep the taller patch: fewer rows cross shared memory per sweep)
                    best = cost;
                    found = true;
                    out->R = R; out->wy = wy; out->wx = wx; out->tiles_y = ty; out->tiles_x = tx;
                    out->threads = warps * 32; out->smem = smem;
                }
            }
        }
    }
    return found;
}

bool resident_wanted(const fluidb200_solver* s) {
    const int on = s->resident >= 0 ? s->resident : env_int("FLUIDB200_RESIDENT", 1);
    if (!on) return false;
    if ((long long)s->S * s->K < 2) return false;   // the corner rule needs a sweep before the last one
    if ((long long)s->S * s->K > (1LL << 30)) return false;   // sweeps are numbered in 32 bits (packet sequence numbers)
    for (int f = 0; f < 3; ++f)
        if (s->mask[f].set) return false;
    for (int e = 0; e < 4; ++e) {
        if (s->edges[0].mode[e] != FLUIDB200_EDGE_CONST || s->edges[1].mode[e] != FLUIDB200_EDGE_CONST) return false;
        if (s->edges[2].mode[e] == FLUIDB200_EDGE_ARRAY) return false;
    }
    return true;
}

template <typename T> int solve_resident_t(fluidb200_solver* s, bool* done) {
    *done = false;
    if (!s->res_planned) {
        s->res_planned = true;
        s->res_ok = fbk::be_resident_max_threads(4) > 0 && plan_resident<T>(s->X, s->Y, &s->res_plan) &&
                    fbk::resident_capacity<T>(s->res_plan) >= s->res_plan.tiles_x * s->res_plan.tiles_y;
        if (s->res_ok) {
            if (fbk::be_malloc(&s->res_scratch, fb::resident_scratch_bytes<T>(s->res_plan)) != 0) s->res_ok = false;
        }
    }
    if (!s->res_ok) return 0;
    const fb::ResidentPlan& pl = s->res_plan;
    fb::ResidentArgs<T> a = {};
    for (int i = 0; i < FLUIDB200_NUM_FIELDS; ++i) a.f[i] = (T*)s->f[i];
    a.pitch = s->pitch; a.X = s->X; a.Y = s->Y;
    a.pa.m = fb::make_metric<T>((T)s->hx, (T)s->hy, env_int("FLUIDB200_DIVMODE", -1));
    a.pa.nu = (T)s->nu; a.pa.dt = (T)s->dt;
    a.pa.rho_over_dt = (T)s->rho / (T)s->dt;   // SerialNavierStokes.cpp:332
    a.ca.m = a.pa.m;
    a.ca.dt_over_rho = (T)s->dt / (T)s->rho;   // :346
    a.hx = (T)s->hx; a.hy = (T)s->hy;
    a.K = s->K; a.S = s->S;
    a.wy = pl.wy; a.wx = pl.wx; a.tiles_y = pl.tiles_y; a.tiles_x = pl.tiles_x;
    for (int e = 0; e < 4; ++e) {
        a.pmode[e] = s->edges[2].mode[e];
        a.pval[e] = (T)s->edges[2].cval[e];
        a.uval[e] = (T)s->edges[0].cval[e];
        a.vval[e] = (T)s->edges[1].cval[e];
    }
    const size_t words = fb::resident_sync_words(pl.tiles_x * pl.tiles_y);
    a.sync = (unsigned*)s->res_scratch;
    a.stash = (T*)((char*)s->res_scratch + words * sizeof(unsigned));
    a.band = (char*)s->res_scratch + words * sizeof(unsigned) + 256;
    a.band_len = fb::resident_band_len<T>(pl);
    a.prof = env_int("FLUIDB200_RESIDENT_PROF", 0) ? (unsigned long long*)((char*)s->res_scratch + words * sizeof(unsigned) + 64) : nullptr;
    BE_CHECK(fbk::launch_resident<T>(a, pl, s->stream), "resident launch");
    s->launches += 1;
    s->resident_launches += 1;
    s->p_next_is_p = true;
    *done = true;
    return 0;
}

int launches_per_step(const fluidb200_solver* s) {
    const int kmax = poisson_kmax(s);
    return 2 + (s->K + kmax - 1) / kmax;
}

}  // namespace

extern "C" int fluidb200_create(fluidb200_solver** out, int dtype, int X, int Y, double domain_x, double domain_y) {
    if (!out) return fail(FLUIDB200_EINVAL, "out is NULL");
    *out = nullptr;
    if (dtype != FLUIDB200_F32 && dtype != FLUIDB200_F64) return fail(FLUIDB200_EINVAL, "bad dtype %d", dtype);
    if (X < 3 || Y < 3) return fail(FLUIDB200_EINVAL, "box must be at least 3x3 (got %dx%d)", X, Y);
    fluidb200_solver* s = new (std::nothrow) fluidb200_solver();
    if (!s) return fail(FLUIDB200_ENOMEM, "host allocation failed");
    s->dtype = dtype; s->X = X; s->Y = Y;
    s->esize = dtype == FLUIDB200_F32 ? 4 : 8;
    s->pitch = ((int64_t)Y + 31) / 32 * 32;
    // Temporal-blocking depth, from depth sweeps on B200 (fp32, Gcell-sweeps/s at the chosen depth):
    // 8 from 8192^2 up (2880; 16384^2: 3640), 7 around 6144^2 (2290), 5 around 3072^2-4096^2 (940-1460),
    // 3 below (2048^2: 590, 1000^2: 170): short pipelines fill faster and the recomputed halo weighs
    // more on small chunks. FLUIDB200_KMAX / set_sweeps_per_launch override.
    {
        const long long cells = (long long)X * Y;
        int k = cells >= 60000000LL ? 8 : cells >= 30000000LL ? 7 : cells >= 8000000LL ? 5 : 3;
        // fp64: half the FP rate and twice the bytes; depth 6 is DRAM bound already (16384^2: 1570
        // Gcell-sweeps/s at depth 6, 1520 at 8; 8192^2: 1300 at 6, 1340 at 7)
        if (dtype == FLUIDB200_F64) k = std::min(k, 6);
        s->kmax = env_int("FLUIDB200_KMAX", k);
        // Small grids are launch bound (a timestep is ~20 launches of ~10 us): solve() replays each
        // timestep as a CUDA graph there (41^2: 231 -> 190 ms per 1000 steps, 1000^2: 287 -> 242 ms).
        // fluidb200_set_use_graph / FLUIDB200_GRAPH override; a stream that cannot be captured falls
        // back to plain launches.
        s->use_graph = env_int("FLUIDB200_GRAPH", cells <= 4000000LL ? 1 : 0) != 0;
    }
    fluidb200_set_domain(s, domain_x, domain_y);
    if (fbk::be_stream_create(&s->stream) != 0) {
        delete s;
        return fail_backend("stream create");
    }
    s->own_stream = true;
    for (int i = 0; i < FLUIDB200_NUM_FIELDS; ++i) {
        if (fbk::be_malloc(&s->f[i], field_bytes(s)) != 0 ||
            fbk::be_memset_async(s->f[i], 0, field_bytes(s), s->stream) != 0) {
            int rc = fail(FLUIDB200_ENOMEM, "device allocation of %zu bytes failed: %s", field_bytes(s),
                          fbk::be_last_error_string());
            fluidb200_destroy(s);
            return rc;
        }
    }
    *out = s;
    return 0;
}

extern "C" int fluidb200_destroy(fluidb200_solver* s) {
    if (!s) return 0;
    if (s->stream) fbk::be_stream_sync(s->stream);
    drop_graphs(s);
    for (int f = 0; f < 3; ++f) free_bc(s, f);
    for (int i = 0; i < FLUIDB200_NUM_FIELDS; ++i)
        if (s->f[i]) fbk::be_free(s->f[i]);
    if (s->res_scratch) fbk::be_free(s->res_scratch);
    for (void* p : {s->pipe_out[0], s->pipe_out[1], s->pipe_p[0], s->pipe_p[1]})
        if (p) fbk::be_free(p);
    for (void* e : s->pipe_ev) fbk::be_event_destroy(e);
    if (s->pipe_up) fbk::be_stream_destroy(s->pipe_up);
    if (s->pipe_down) fbk::be_stream_destroy(s->pipe_down);
    if (s->pipe_up2) fbk::be_stream_destroy(s->pipe_up2);
    if (s->pipe_down2) fbk::be_stream_destroy(s->pipe_down2);
    if (s->own_stream && s->stream) fbk::be_stream_destroy(s->stream);
    delete s;
    return 0;
}

extern "C" int fluidb200_set_params(fluidb200_solver* s, double time_step, double kinematic_viscosity,
                                    double density, int num_poisson_iterations, int num_iterations) {
    if (!s) return fail(FLUIDB200_EINVAL, "solver is NULL");
    if (num_poisson_iterations < 0 || num_iterations < 0)
        return fail(FLUIDB200_EINVAL, "iteration counts must be >= 0");
    s->dt = time_step; s->nu = kinematic_viscosity; s->rho = density;
    s->K = num_poisson_iterations; s->S = num_iterations;
    return 0;
}

extern "C" double fluidb200_element_length(int dtype, double domain, int n) {
    if (dtype == FLUIDB200_F32) return (double)((float)domain / (float)(n - 1));
    return domain / (double)(n - 1);
}

extern "C" int fluidb200_set_domain(fluidb200_solver* s, double domain_x, double domain_y) {
    if (!s) return fail(FLUIDB200_EINVAL, "solver is NULL");
    s->wx = domain_x; s->wy = domain_y;
    s->hx = fluidb200_element_length(s->dtype, domain_x, s->X);
    s->hy = fluidb200_element_length(s->dtype, domain_y, s->Y);
    return 0;
}

extern "C" int fluidb200_upload(fluidb200_solver* s, const void* const fields[FLUIDB200_NUM_FIELDS],
                                const void* const bc_val[3], const uint8_t* const bc_set[3]) {
    if (!s) return fail(FLUIDB200_EINVAL, "solver is NULL");
    const size_t row = (size_t)s->Y * s->esize;
    if (fields) {
        for (int i = 0; i < FLUIDB200_NUM_FIELDS; ++i) {
            if (!fields[i]) continue;
            BE_CHECK(fbk::be_memcpy2d_async(s->f[i], (size_t)s->pitch * s->esize, fields[i], row, row, s->X, fbk::H2D,
                                            s->stream), "field upload");
            if (i == FLUIDB200_FIELD_P || i == FLUIDB200_FIELD_P_NEXT) s->p_next_is_p = false;
        }
    }
    for (int f = 0; f < 3; ++f) {
        const void* v = bc_val ? bc_val[f] : nullptr;
        const uint8_t* b = bc_set ? bc_set[f] : nullptr;
        if (!v && !b) continue;
        if (!v || !b) return fail(FLUIDB200_EINVAL, "bc_val[%d] and bc_set[%d] must both be given", f, f);
        int rc = s->dtype == FLUIDB200_F32 ? upload_bc_t<float>(s, f, (const float*)v, b)
                                           : upload_bc_t<double>(s, f, (const double*)v, b);
        if (rc) return rc;
    }
    BE_CHECK(fbk::be_stream_sync(s->stream), "upload");  // caller may free its host buffers on return
    return 0;
}

extern "C" int fluidb200_upload_cavity(fluidb200_solver* s, double lid_velocity) {
    if (!s) return fail(FLUIDB200_EINVAL, "solver is NULL");
    fluidb200_slab g;
    fill_slab(s, &g);
    for (int f = 0; f < 3; ++f) free_bc(s, f);
    drop_graphs(s);
    fluidb200_cavity_edges('u', lid_velocity, &s->edges[0]);
    fluidb200_cavity_edges('v', lid_velocity, &s->edges[1]);
    fluidb200_cavity_edges('p', lid_velocity, &s->edges[2]);
    s->p_next_is_p = false;
    int rc = fluidb200_init_cavity(&g, s->f[0], s->f[1], s->f[2], s->f[3], s->f[4], s->f[5], s->f[6], lid_velocity,
                                   s->stream);
    if (rc == 0) s->launches++;
    return rc;
}

extern "C" int fluidb200_step_predictor(fluidb200_solver* s) {
    if (!s) return fail(FLUIDB200_EINVAL, "solver is NULL");
    return step_predictor(s);
}
extern "C" int fluidb200_step_poisson(fluidb200_solver* s, int sweeps) {
    if (!s) return fail(FLUIDB200_EINVAL, "solver is NULL");
    if (sweeps < 0) return fail(FLUIDB200_EINVAL, "sweeps must be >= 0");
    return step_poisson(s, sweeps);
}
extern "C" int fluidb200_step_corrector(fluidb200_solver* s) {
    if (!s) return fail(FLUIDB200_EINVAL, "solver is NULL");
    return step_corrector(s);
}

extern "C" int fluidb200_solve(fluidb200_solver* s) {
    if (!s) return fail(FLUIDB200_EINVAL, "solver is NULL");
    if (s->S == 0) return 0;
    if (!(s->dt != 0) || !(s->rho != 0)) return fail(FLUIDB200_EINVAL, "time_step and density must be non-zero");
    if (resident_wanted(s)) {   // small grid, cavity-style boundaries: every timestep in ONE launch
        bool done = false;
        const int rc = s->dtype == FLUIDB200_F32 ? solve_resident_t<float>(s, &done) : solve_resident_t<double>(s, &done);
        if (rc || done) return rc;
    }
    if (!s->use_graph) {
        for (int i = 0; i < s->S; ++i) {
            int rc = one_step(s);
            if (rc) return rc;
        }
        return 0;
    }
    // CUDA-graph replay: one captured timestep per ping-pong parity of p (a step with an odd
    // number of sweep launches leaves p in the other buffer). The graphs bake the scalar
    // parameters in, so they are rebuilt when any of them changes.
    const double key[6] = {s->dt, s->nu, s->rho, (double)s->K + 4096.0 * s->kmax, s->hx, s->hy};
    if (memcmp(key, s->graph_key, sizeof(key)) != 0) {
        drop_graphs(s);
        memcpy(s->graph_key, key, sizeof(key));
    }
    const bool odd = (launches_per_step(s) - 2) % 2 != 0;
    for (int i = 0; i < s->S; ++i) {
        const int parity = (s->graph[0] && s->f[FLUIDB200_FIELD_P] != s->graph_home_p) ? 1 : 0;
        if (!s->graph[parity]) {
            if (parity == 0) s->graph_home_p = s->f[FLUIDB200_FIELD_P];
            const int64_t before = s->launches;
            if (fbk::be_graph_begin(s->stream) != 0) {
                s->use_graph = false;  // backend cannot capture: plain launches from here on
                for (; i < s->S; ++i) {
                    int rc = one_step(s);
                    if (rc) return rc;
                }
                return 0;
            }
            const int rc = one_step(s);  // recorded, not executed
            void* exec = nullptr;
            const int rc2 = fbk::be_graph_end(s->stream, &exec);
            s->graph_kernels[parity] = s->launches - before;
            s->launches = before;
            if (odd) std::swap(s->f[FLUIDB200_FIELD_P], s->f[FLUIDB200_FIELD_P_NEXT]);  // undo recorded swaps
            if (rc) return rc;
            if (rc2 != 0) {  // the capture did not yield a graph (nothing ran): plain launches from this step on
                s->use_graph = false;
                for (; i < s->S; ++i) {
                    int rc3 = one_step(s);
                    if (rc3) return rc3;
                }
                return 0;
            }
            s->graph[parity] = exec;
        }
        BE_CHECK(fbk::be_graph_launch(s->graph[parity], s->stream), "graph launch");
        s->launches += s->graph_kernels[parity];
        if (s->K > 0) s->p_next_is_p = true;
        if (odd) std::swap(s->f[FLUIDB200_FIELD_P], s->f[FLUIDB200_FIELD_P_NEXT]);
    }
    return 0;
}

extern "C" int fluidb200_synchronize(fluidb200_solver* s) {
    if (!s) return fail(FLUIDB200_EINVAL, "solver is NULL");
    BE_CHECK(fbk::be_stream_sync(s->stream), "synchronize");
    return 0;
}

extern "C" int fluidb200_download(fluidb200_solver* s, int field, void* out) {
    if (!s) return fail(FLUIDB200_EINVAL, "solver is NULL");
    if (field < 0 || field >= FLUIDB200_NUM_FIELDS) return fail(FLUIDB200_EINVAL, "bad field %d", field);
    if (!out) return fail(FLUIDB200_EINVAL, "out is NULL");
    // updatePressure leaves p_next == p after every sweep (SerialNavierStokes.cpp:192-198); with
    // ping-pong buffers the "other" buffer is one launch stale, so p_next reads p.
    int src = field;
    if (field == FLUIDB200_FIELD_P_NEXT && s->p_next_is_p) src = FLUIDB200_FIELD_P;
    const size_t row = (size_t)s->Y * s->esize;
    BE_CHECK(fbk::be_memcpy2d_async(out, row, s->f[src], (size_t)s->pitch * s->esize, row, s->X, fbk::D2H, s->stream),
             "download");
    BE_CHECK(fbk::be_stream_sync(s->stream), "download");
    return 0;
}

// ---- migrateSolve with host buffers: upload, timesteps, download -----------------------------------
namespace {

// One timestep of a LARGE grid, pipelined over row chunks so that the three legs -- host -> device,
// compute, device -> host -- of different chunks run at the same time and the link works in both
// directions at once. Rows [a, b) after one timestep depend on rows [a - H, b + H) before it with
// H = K + 2 (K sweeps reach K rows, the corrector one more, the predictor's u_next another), so a chunk
// is computed on its rows plus H halo rows from launches whose valid range shrinks sweep by sweep; the
// halo is recomputed (2H / chunk rows of extra arithmetic, hidden behind the transfers), never sent twice:
// every input row crosses the link once, into the solver's own u, v, p, which stay intact as inputs;
// results go to other arrays (new u, v: two extra arrays; new p: p_next) and the roles are swapped at the end.
int migrate_solve_pipelined(fluidb200_solver* s, const void* const host_in[3], void* const host_out[3], int chunk_rows) {
    const int X = s->X, Y = s->Y, K = s->K, H = K + 2;
    const size_t es = s->esize, pitchB = (size_t)s->pitch * es, rowB = (size_t)Y * es;
    // Chunk boundaries: uniform, except that the last chunk is cut in three (1/2, 1/4, 1/4): what follows the last upload
    // -- that chunk's timestep and its download, with the link idle in the other direction -- shrinks with it (about 1.5 ms
    // of a 16384^2 step), while three short chunks add little to the rows computed twice.
    std::vector<int> cut;   // chunk c covers rows [cut[c], cut[c + 1])
    for (int a = 0; a < X; a += chunk_rows) cut.push_back(a);
    if (env_int("FLUIDB200_E2E_TAPER", 1) && cut.size() >= 4) {
        const int a = cut.back(), n = X - a;
        if (n >= 64) {
            cut.push_back(a + n / 2);
            cut.push_back(a + n / 2 + n / 4);
        }
    }
    cut.push_back(X);
    const int nchunks = (int)cut.size() - 1;
    // resources (kept with the solver)
    if (!s->pipe_up && (fbk::be_stream_create(&s->pipe_up) || fbk::be_stream_create(&s->pipe_down) || fbk::be_stream_create(&s->pipe_up2) ||
                        fbk::be_stream_create(&s->pipe_down2)))
        return fail_backend("pipeline streams");
    for (int i = 0; i < 2; ++i)
        if (!s->pipe_out[i]) {
            BE_CHECK(fbk::be_malloc(&s->pipe_out[i], field_bytes(s)), "pipeline output arrays");
            BE_CHECK(fbk::be_memset_async(s->pipe_out[i], 0, field_bytes(s), s->stream), "pipeline output arrays");
        }
    const size_t prow = (size_t)chunk_rows + 2 * H;
    if (s->pipe_p_rows < prow) {
        for (int i = 0; i < 2; ++i) {
            if (s->pipe_p[i]) fbk::be_free(s->pipe_p[i]);
            s->pipe_p[i] = nullptr;
            BE_CHECK(fbk::be_malloc(&s->pipe_p[i], prow * pitchB), "pipeline pressure buffers");
        }
        s->pipe_p_rows = prow;
    }
    while ((int)s->pipe_ev.size() < 2 * nchunks + 3) {
        void* e = nullptr;
        BE_CHECK(fbk::be_event_create(&e), "pipeline events");
        s->pipe_ev.push_back(e);
    }
    // (measurement knob) FLUIDB200_E2E_STREAMS=2: chunks alternate between two copy streams per direction. Measured: 92.8-94.9 ms
    // per 16384^2 step (34 GB/s each way) against 71.3-71.5 (45 GB/s) with one stream per direction: copies of one direction
    // in flight together get in each other's way. Default 1.
    const bool two = env_int("FLUIDB200_E2E_STREAMS", 1) >= 2;
    void* const ups[2] = {s->pipe_up, two ? s->pipe_up2 : s->pipe_up};
    void* const downs[2] = {s->pipe_down, two ? s->pipe_down2 : s->pipe_down};
    void* const comp = s->stream;
    // the copy streams start after whatever the solver's stream was doing (and the output arrays' clearing)
    void* ev0 = s->pipe_ev[2 * nchunks];
    BE_CHECK(fbk::be_event_record(ev0, comp), "pipeline");
    for (int i = 0; i < (two ? 2 : 1); ++i) {
        BE_CHECK(fbk::be_stream_wait_event(ups[i], ev0), "pipeline");
        BE_CHECK(fbk::be_stream_wait_event(downs[i], ev0), "pipeline");
    }

    char* const fU = (char*)s->f[FLUIDB200_FIELD_U];
    char* const fV = (char*)s->f[FLUIDB200_FIELD_V];
    char* const fP = (char*)s->f[FLUIDB200_FIELD_P];
    char* const fUS = (char*)s->f[FLUIDB200_FIELD_U_NEXT];
    char* const fVS = (char*)s->f[FLUIDB200_FIELD_V_NEXT];
    char* const fR = (char*)s->f[FLUIDB200_FIELD_RHS];
    char* const oP = (char*)s->f[FLUIDB200_FIELD_P_NEXT];
    char* const oU = (char*)s->pipe_out[0];
    char* const oV = (char*)s->pipe_out[1];
    char* const dev_in[3] = {fU, fV, fP};
    char* const dev_out[3] = {oU, oV, oP};
    const int kmax = poisson_kmax(s);
    const int nlaunch = (K + kmax - 1) / kmax;
    const long long k0 = fbk::be_kernel_count();
    const auto t_issue0 = std::chrono::steady_clock::now();
    int uploaded = 0;   // rows [0, uploaded) are on their way to the device
    for (int c = 0; c < nchunks; ++c) {
        const int a = cut[c], b = cut[c + 1];
        const int lo = std::max(0, a - H), hi = std::min(X, b + H);
        void* const up = ups[c & 1];
        void* const down = downs[c & 1];
        // 1. host -> device: the rows this chunk needs and no earlier chunk brought
        if (hi > uploaded) {
            for (int f = 0; f < 3; ++f)
                BE_CHECK(fbk::be_memcpy2d_async(dev_in[f] + (size_t)uploaded * pitchB, pitchB, (const char*)host_in[f] + (size_t)uploaded * rowB,
                                                rowB, rowB, (size_t)(hi - uploaded), fbk::H2D, up), "pipeline upload");
            uploaded = hi;
        }
        BE_CHECK(fbk::be_event_record(s->pipe_ev[2 * c], up), "pipeline");
        BE_CHECK(fbk::be_stream_wait_event(comp, s->pipe_ev[2 * c]), "pipeline");
        // 2. one timestep on rows [a, b), computed on the slab [lo, hi)
        fluidb200_slab g;
        fill_slab(s, &g);
        g.row0 = lo;
        g.rows = hi - lo;
        const size_t off = (size_t)lo * pitchB;
        const int star_lo = lo == 0 ? 0 : lo + 1, star_hi = hi == X ? X : hi - 1;
        const int rhs_lo = lo == 0 ? 0 : lo + 2, rhs_hi = hi == X ? X : hi - 2;
        int rc = fluidb200_predictor(&g, fU + off, fV + off, fUS + off, fVS + off, fR + off, s->dt, s->nu, s->rho, star_lo, star_hi,
                                     rhs_lo, rhs_hi, comp);
        if (rc) return rc;
        const char* src = fP + off;   // the first launch reads the pristine p; then the chunk's own pair
        int done = 0, left = K, which = 0;
        for (int i = 0; i < nlaunch; ++i) {
            const int k = (left + (nlaunch - i) - 1) / (nlaunch - i);
            done += k;
            left -= k;
            const int out_lo = lo == 0 ? 0 : lo + 1 + done, out_hi = hi == X ? X : hi - 1 - done;
            rc = fluidb200_poisson_sweeps(&g, src, s->pipe_p[which], fR + off, &s->edges[2], nullptr, k, out_lo, out_hi, comp);
            if (rc) return rc;
            src = (const char*)s->pipe_p[which];
            which ^= 1;
        }
        rc = fluidb200_corrector(&g, src, fUS + off, fVS + off, oU + off, oV + off, &s->edges[0], &s->edges[1], nullptr, nullptr, s->dt, s->rho,
                                 a, b, comp);
        if (rc) return rc;
        // the chunk's final p joins the whole-grid array (rows [a, b); device to device)
        BE_CHECK(fbk::be_memcpy_async(oP + (size_t)a * pitchB, src + (size_t)(a - lo) * pitchB, (size_t)(b - a) * pitchB, fbk::D2D, comp),
                 "pipeline p");
        BE_CHECK(fbk::be_event_record(s->pipe_ev[2 * c + 1], comp), "pipeline");
        // 3. device -> host
        BE_CHECK(fbk::be_stream_wait_event(down, s->pipe_ev[2 * c + 1]), "pipeline");
        for (int f = 0; f < 3; ++f)
            if (host_out[f])
                BE_CHECK(fbk::be_memcpy2d_async((char*)host_out[f] + (size_t)a * rowB, rowB, dev_out[f] + (size_t)a * pitchB, pitchB, rowB,
                                                (size_t)(b - a), fbk::D2H, down), "pipeline download");
    }
    s->launches += fbk::be_kernel_count() - k0;
    s->pipe_issue_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_issue0).count();
    // the new state takes the place of the old one
    std::swap(s->f[FLUIDB200_FIELD_U], s->pipe_out[0]);
    std::swap(s->f[FLUIDB200_FIELD_V], s->pipe_out[1]);
    std::swap(s->f[FLUIDB200_FIELD_P], s->f[FLUIDB200_FIELD_P_NEXT]);
    s->p_next_is_p = true;
    drop_graphs(s);   // (captured graphs hold the old pointers)
    for (int i = 0; i < (two ? 2 : 1); ++i) {   // later work on the solver's stream comes after the downloads
        void* evd = s->pipe_ev[2 * nchunks + 1 + i];
        BE_CHECK(fbk::be_event_record(evd, downs[i]), "pipeline");
        BE_CHECK(fbk::be_stream_wait_event(comp, evd), "pipeline");
    }
    for (int i = 0; i < (two ? 2 : 1); ++i) BE_CHECK(fbk::be_stream_sync(downs[i]), "pipeline download");
    BE_CHECK(fbk::be_stream_sync(comp), "pipeline");
    if (env_int("FLUIDB200_E2E_DEBUG", 0))
        fprintf(stderr, "[fluidb200] migrate_solve: %d chunks issued in %.2f ms of host time, done after %.2f ms\n", nchunks, s->pipe_issue_ms,
                std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_issue0).count());
    return 0;
}

}  // namespace

extern "C" int fluidb200_migrate_solve(fluidb200_solver* s, const void* const host_in[3], void* const host_out[3]) {
    if (!s) return fail(FLUIDB200_EINVAL, "solver is NULL");
    if (!host_in || !host_out) return fail(FLUIDB200_EINVAL, "host_in / host_out is NULL");
    for (int f = 0; f < 3; ++f)
        if (!host_in[f]) return fail(FLUIDB200_EINVAL, "host_in[%d] is NULL (u, v and p are all needed)", f);
    if (!(s->dt != 0) || !(s->rho != 0)) return fail(FLUIDB200_EINVAL, "time_step and density must be non-zero");
    // chunk pipeline: one timestep, no interior flags, a grid tall enough that the halo is a fraction of a chunk
    const int H = s->K + 2;
    int chunk = env_int("FLUIDB200_E2E_CHUNK", std::max(4 * H, s->X / 32));
    chunk = std::max(chunk, 2);
    const bool masks = s->mask[0].set || s->mask[1].set || s->mask[2].set;
    if (s->S == 1 && s->K >= 1 && !masks && chunk + 2 * H <= s->X / 2 && env_int("FLUIDB200_E2E_PIPELINE", 1))
        return migrate_solve_pipelined(s, host_in, host_out, chunk);
    // otherwise: the plain sequence, with the three downloads queued together
    const void* fields[FLUIDB200_NUM_FIELDS] = {host_in[0], host_in[1], host_in[2], nullptr, nullptr, nullptr, nullptr};
    int rc = fluidb200_upload(s, fields, nullptr, nullptr);
    if (rc) return rc;
    if ((rc = fluidb200_solve(s))) return rc;
    const size_t row = (size_t)s->Y * s->esize;
    for (int f = 0; f < 3; ++f)
        if (host_out[f])
            BE_CHECK(fbk::be_memcpy2d_async(host_out[f], row, s->f[f], (size_t)s->pitch * s->esize, row, s->X, fbk::D2H, s->stream),
                     "download");
    BE_CHECK(fbk::be_stream_sync(s->stream), "download");
    return 0;
}

extern "C" int fluidb200_get_slab(const fluidb200_solver* s, fluidb200_slab* out) {
    if (!s || !out) return fail(FLUIDB200_EINVAL, "NULL argument");
    fill_slab(s, out);
    return 0;
}
extern "C" void* fluidb200_field_ptr(fluidb200_solver* s, int field) {
    if (!s || field < 0 || field >= FLUIDB200_NUM_FIELDS) return nullptr;
    return s->f[field];
}
extern "C" int fluidb200_set_stream(fluidb200_solver* s, void* stream) {
    if (!s) return fail(FLUIDB200_EINVAL, "solver is NULL");
    fbk::be_stream_sync(s->stream);
    drop_graphs(s);
    if (s->own_stream && s->stream) fbk::be_stream_destroy(s->stream);
    s->stream = stream;
    s->own_stream = false;
    return 0;
}
extern "C" int fluidb200_set_sweeps_per_launch(fluidb200_solver* s, int k) {
    if (!s) return fail(FLUIDB200_EINVAL, "solver is NULL");
    if (k < 1 || k > FLUIDB200_MAX_SWEEPS_PER_LAUNCH)
        return fail(FLUIDB200_EINVAL, "sweeps per launch must be in [1,%d]", FLUIDB200_MAX_SWEEPS_PER_LAUNCH);
    s->kmax = k;
    drop_graphs(s);
    return 0;
}
extern "C" int fluidb200_set_use_graph(fluidb200_solver* s, int enable) {
    if (!s) return fail(FLUIDB200_EINVAL, "solver is NULL");
    s->use_graph = enable != 0;
    if (!enable) drop_graphs(s);
    return 0;
}
extern "C" int fluidb200_set_resident(fluidb200_solver* s, int enable) {
    if (!s) return fail(FLUIDB200_EINVAL, "solver is NULL");
    s->resident = enable < 0 ? -1 : (enable != 0);
    return 0;
}
// (measurement aid) cycles per phase of the last resident launch made with FLUIDB200_RESIDENT_PROF=1:
// out[0..5] = ring load + grid barrier, tentative velocity, right-hand side, sweeps (the rest after the split below), corrector, write-back;
// out[8..12] = inside the sweeps: waiting for the borders, first + last row and their publication, rows between, fix-ups + columns, CTA barrier;
// out[13] = polling rounds
extern "C" int fluidb200_resident_profile(fluidb200_solver* s, unsigned long long out[24]) {
    if (!s || !out || !s->res_scratch) return fail(FLUIDB200_ESTATE, "no resident launch to report on");
    BE_CHECK(fbk::be_stream_sync(s->stream), "resident profile");
    const size_t words = fb::resident_sync_words(s->res_plan.tiles_x * s->res_plan.tiles_y);
    BE_CHECK(fbk::be_memcpy_async(out, (char*)s->res_scratch + words * sizeof(unsigned) + 64, 192, fbk::D2H, s->stream), "resident profile");
    BE_CHECK(fbk::be_stream_sync(s->stream), "resident profile");
    return 0;
}
extern "C" int64_t fluidb200_resident_launch_count(const fluidb200_solver* s) { return s ? s->resident_launches : 0; }
extern "C" int64_t fluidb200_launch_count(const fluidb200_solver* s) { return s ? s->launches : 0; }

extern "C" int fluidb200_set_device(int device) {
    BE_CHECK(fbk::be_set_device(device), "set_device");
    return 0;
}
extern "C" int fluidb200_report_error(int code, const char* message) { return fail(code, "%s", message ? message : ""); }
extern "C" const char* fluidb200_last_error(void) { return g_err; }
extern "C" const char* fluidb200_version(void) { return "fluidb200 0.1 (sm_100a)"; }
