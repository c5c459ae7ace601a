// warp_args.h -- plain argument structs of the three kernels (no CUDA types), shared by the
// device code (warp_kernels.cuh), the host planner (api.cpp) and the backend seam (backend.h).
#pragma once
#include <stdint.h>

namespace fb {

constexpr int kWarpsPerCta = 4;

// ---------------------------------------------------------------------------------------
// boundary description on the device (see fluidb200_edges in include/fluidb200.h)
// ---------------------------------------------------------------------------------------
enum { EDGE_UNSET = 0, EDGE_CONST = 1, EDGE_ARRAY = 2 };
enum { BOTTOM = 0, TOP = 1, LEFT = 2, RIGHT = 3 };

template <typename T> struct Edges {
    int mode[4];
    T cval[4];
    const T* val[4];
    const uint8_t* set[4];
};

// Geometry shared by the kernels. All row indices are GLOBAL x; the arrays hold rows
// [row0, row0+rows).
struct Geom {
    long long pitch;
    int X, Y;
    int row0, rows;
    int nstrips;         // strips along y
    int rows_per_chunk;  // x rows written by one warp task
};

// Grid spacing as the kernels consume it. hx2 = 2*hx (exact). The reciprocals feed the fp32
// build's division-free arithmetic (see Ar<float> in warp_kernels.cuh); hx*hy is exact in double
// (two 24-bit significands). `mode` selects that arithmetic: 0 = IEEE divisions (always valid),
// 1 = division-free with flagged exact fallback (fp32 only), 2 = both spacings are powers of two,
// divisions are exact scalings (fp32 and fp64).
template <typename T> struct Metric {
    T hx, hy, hx2, hy2;
    double inv_hx2, inv_hy2, inv_hxhy;
    float inv_hx2f, inv_hy2f, inv_hxhyf;
    int mode;
    // mode 2: the Laplacian numerator t is safe to scale when t == 0 or pow2_lo <= |t| < inf, with
    // pow2_lo = 2^-120 * max(1, hx*hy) so that t / (hx*hy) is a normal float as well; stored as the
    // bit pattern minus one (see Flags in warp_kernels.cuh)
    unsigned pow2_lo_m1;
};
inline bool metric_is_pow2(double h) {
    if (!(h > 0)) return false;
    union { double d; unsigned long long u; } b;
    b.d = h;
    return (b.u & 0x000fffffffffffffULL) == 0;
}
template <typename T> inline Metric<T> make_metric(T hx, T hy, int forced_mode = -1) {
    Metric<T> m;
    m.hx = hx; m.hy = hy;
    m.hx2 = (T)2 * hx; m.hy2 = (T)2 * hy;
    m.inv_hx2 = 1.0 / (double)m.hx2;
    m.inv_hy2 = 1.0 / (double)m.hy2;
    m.inv_hxhy = 1.0 / ((double)hx * (double)hy);
    m.inv_hx2f = (float)m.inv_hx2; m.inv_hy2f = (float)m.inv_hy2; m.inv_hxhyf = (float)m.inv_hxhy;
    m.mode = 0;
    {
        const double hxhy = (double)hx * (double)hy;
        union { float f; unsigned u; } lo;
        lo.f = (float)(7.52316384526264e-37 /* 2^-120 */ * (hxhy > 1.0 ? hxhy : 1.0));
        m.pow2_lo_m1 = lo.u - 1u;
    }
    {
        const double lo = 9.094947017729282e-13 /* 2^-40 */, hi = 1099511627776.0 /* 2^40 */;
        const bool sane = (double)hx >= lo && (double)hx <= hi && (double)hy >= lo && (double)hy <= hi;
        const bool pow2 = metric_is_pow2((double)hx) && metric_is_pow2((double)hy);
        if (sane) m.mode = pow2 ? 2 : 1;
        if (forced_mode >= 0 && forced_mode < m.mode) m.mode = forced_mode;  // only ever towards "more exact"
        if (m.mode == 1 && sizeof(T) == 8) {
            // fp64 mode 1 corrects quotients by the constants 2hx, 2hy, hx, hy with fused multiply-adds; a divisor whose
            // significand is all ones is the one case that correction is not proven for: IEEE divisions there
            union { double d; unsigned long long u; } a, b;
            a.d = (double)hx; b.d = (double)hy;
            const unsigned long long ones = 0x000fffffffffffffULL;
            if ((a.u & ones) == ones || (b.u & ones) == ones) m.mode = 0;
        }
    }
    return m;
}

// Jacobi sweeps (see jacobi_task in warp_kernels.cuh)
template <typename T> struct JacobiArgs {
    const T* p_in;
    T* p_out;
    const T* rhs;
    const uint8_t* mset;  // interior pressure flags (MASK build only)
    const T* mval;
    // Optional summary of mset: one byte per tile of kMaskTileRows x kMaskTileCols cells (rows counted
    // from g.row0), non-zero when the tile holds a flagged interior cell; mtiles_nc tiles per tile row.
    // With it a masked launch can be split: tasks whose footprint holds no flag run the lean kernels.
    const uint8_t* mtiles;
    int mtiles_nc;
    Geom g;
    T hx, hy;
    int out_lo, out_hi;   // global rows to write
    Edges<T> e;
    // Split launch (split_jacobi in api.cpp): a launch is cut into a lean FAST kernel over the tasks
    // whose whole pipeline runs on plain interior rows and columns, a lean EDGE kernel over the
    // strips that hold a y-edge column (same rows), and the GENERAL kernel over the chunks at an x wall.
    int part;             // 0 = one kernel does every task, 1 = fast, 2 = general (wall chunks), 3 = edge strips
    int sf_lo, sf_hi;     // strips [sf_lo, sf_hi) hold no y-edge column
    int cf_lo, cf_hi;     // chunks [cf_lo, cf_hi) neither touch an x wall nor lack halo rows
    int nchunks;
    // Chunking of the output rows [out_lo, out_hi): an optional short first chunk of `head` rows and
    // last chunk of `tail` rows (the wall chunks of a split launch, kept as short as the pipeline
    // allows because only the general kernel can run them), uniform chunks of g.rows_per_chunk between.
    int head, tail;
    // 1: the strips holding a y-edge column have their own lean kernel (part 3); 0 (a y edge with
    // per-cell boundary arrays): they are general tasks, appended to the part 2 list
    int edge_lean;
    // 1 (masked split launch): the part 2 list is EVERY task; a task skips itself unless it is a wall
    // chunk, a strip no lean kernel takes, or flagged by mtiles -- and the lean kernels skip flagged tasks
    int gen_all;
};

constexpr int kMaskTileRows = 16, kMaskTileCols = 64;

#ifdef __CUDACC__
#define FB_HD __host__ __device__
#else
#define FB_HD
#endif

// rows [c0, c1) written by chunk `chunk` (c0 >= c1: nothing)
template <typename T> inline FB_HD void jacobi_chunk_rows(const JacobiArgs<T>& a, int chunk, int* c0, int* c1) {
    const int lo = a.out_lo + a.head, hi = a.out_hi - a.tail;  // the uniformly chunked middle
    if (a.head > 0 && chunk == 0) {
        *c0 = a.out_lo; *c1 = lo;
    } else if (a.tail > 0 && chunk == a.nchunks - 1) {
        *c0 = hi; *c1 = a.out_hi;
    } else {
        const int i = chunk - (a.head > 0 ? 1 : 0);
        *c0 = lo + i * a.g.rows_per_chunk;
        *c1 = *c0 + a.g.rows_per_chunk < hi ? *c0 + a.g.rows_per_chunk : hi;
    }
}

// Task decoding shared by the CUDA wrappers and the emulator. Returns false for padding threads.
// part 1: grid (ceil((sf_hi-sf_lo)/W), cf_hi-cf_lo), warps = adjacent fast strips of one fast chunk.
// part 3: one-dimensional list: edge strips x fast chunks.
// part 2: one-dimensional list: every strip x the remaining (wall) chunks.
template <typename T>
inline FB_HD bool jacobi_decode_task(const JacobiArgs<T>& a, int bx, int by, int warp, int warps_per_cta, int* strip, int* chunk) {
    if (a.part == 0) {
        *strip = bx * warps_per_cta + warp;
        *chunk = by;
        return *strip < a.g.nstrips;
    }
    if (a.part == 1) {
        *strip = a.sf_lo + bx * warps_per_cta + warp;
        *chunk = a.cf_lo + by;
        return *strip < a.sf_hi;
    }
    const int id = bx * warps_per_cta + warp;
    if (a.part == 3) {
        const int E = a.sf_lo + (a.g.nstrips - a.sf_hi);
        if (id >= E * (a.cf_hi - a.cf_lo)) return false;
        const int e = id % E;
        *strip = e < a.sf_lo ? e : a.sf_hi + (e - a.sf_lo);
        *chunk = a.cf_lo + id / E;
        return true;
    }
    if (a.gen_all) {
        *strip = id % a.g.nstrips;
        *chunk = id / a.g.nstrips;
        return *chunk < a.nchunks;
    }
    const int E = a.sf_lo + (a.g.nstrips - a.sf_hi), nwall = a.nchunks - (a.cf_hi - a.cf_lo);
    const int ci = id / a.g.nstrips;
    if (ci < nwall) {
        *strip = id % a.g.nstrips;
        *chunk = ci < a.cf_lo ? ci : ci - a.cf_lo + a.cf_hi;
        return true;
    }
    const int id2 = id - nwall * a.g.nstrips;  // edge strips x fast chunks, when no lean kernel takes them
    if (a.edge_lean || id2 >= E * (a.cf_hi - a.cf_lo)) return false;
    const int e = id2 % E;
    *strip = e < a.sf_lo ? e : a.sf_hi + (e - a.sf_lo);
    *chunk = a.cf_lo + id2 / E;
    return true;
}
template <typename T> inline int jacobi_part3_tasks(const JacobiArgs<T>& a) {
    return a.edge_lean ? (a.sf_lo + (a.g.nstrips - a.sf_hi)) * (a.cf_hi - a.cf_lo) : 0;
}
template <typename T> inline int jacobi_part2_tasks(const JacobiArgs<T>& a) {
    if (a.gen_all) return a.g.nstrips * a.nchunks;
    const int E = a.sf_lo + (a.g.nstrips - a.sf_hi);
    return a.g.nstrips * (a.nchunks - (a.cf_hi - a.cf_lo)) + (a.edge_lean ? 0 : E * (a.cf_hi - a.cf_lo));
}

// Fused tentative velocity + right-hand side (see predictor_task)
template <typename T> struct PredictorArgs {
    const T *u, *v;
    T *us, *vs, *rhs;     // u_next, v_next, right_hand_size
    Geom g;
    Metric<T> m;          // element lengths (+ exact doubles, reciprocals)
    T nu, dt, rho_over_dt;
    int lo, hi;           // chunked global row range = union of the two below
    int star_lo, star_hi; // rows of u_next/v_next to write
    int rhs_lo, rhs_hi;   // rows of rhs to write
};

// Fused correction + velocity boundary rows (see corrector_task)
template <typename T> struct CorrectorArgs {
    const T* p;
    const T *us, *vs;   // u_next, v_next
    T *u, *v;
    const uint8_t *uset, *vset;  // interior flags (MASK build only)
    const T *uval, *vval;
    Geom g;
    Metric<T> m;
    T dt_over_rho;
    int out_lo, out_hi;
    Edges<T> eu, ev;
};

}  // namespace fb

// ---------------------------------------------------------------------------------------
// Resident solver (resident_kernels.cuh): whole timesteps of a SMALL grid in one cooperative
// launch, the pressure field living in registers, one patch of R rows x 16 bytes per thread.
// ---------------------------------------------------------------------------------------
namespace fb {

template <typename T> struct ResidentArgs {
    T* f[7];                  // u, v, p, u_next, v_next, p_next, rhs (device, pitched, whole grid)
    long long pitch;
    int X, Y;
    PredictorArgs<T> pa;      // m, nu, dt, rho_over_dt are used
    CorrectorArgs<T> ca;      // m, dt_over_rho are used
    T hx, hy;
    int K, S;                 // sweeps per step, steps in this launch (K * S >= 2)
    int wy, wx;               // warps of a CTA along y and along x
    int tiles_y, tiles_x;     // CTAs along y and along x; tile = blockIdx.x = ty + tiles_y * tx
    int pmode[4];             // pressure edges: EDGE_UNSET or EDGE_CONST
    T pval[4];
    T uval[4], vval[4];       // velocity edges: all EDGE_CONST
    unsigned* sync;           // zeroed before the launch: [0] grid barrier, [32 * (1 + tile)] rows published by a tile
    T* stash;                 // 4 values for the stale-corner rule
    // Rows and columns a tile publishes for its neighbours: [parity][tile][top row, bottom row, left column,
    // right column][band_len elements, as 8-byte packets of 32 data bits + sequence number], zeroed before
    // the launch; band_len >= max(columns, rows) of a tile
    void* band;
    int band_len;
    unsigned long long* prof;  // optional (may be NULL): cycles per phase of one tile, for measurements
};

struct ResidentPlan {   // R == 0: the one-CTA cell layout, wy = cells per thread
    int R, wy, wx, tiles_y, tiles_x, threads;
    size_t smem;
};
constexpr int kResidentRowHalo = 2;  // rows of u, v kept around a tile: u_next at x +- 1 of the rhs stencil
constexpr int kResidentColHalo = 4;  // columns: 2 are needed, 4 keep a lane's 16 bytes aligned inside the frame
template <typename T> inline size_t resident_smem_bytes(int R, int wy, int wx) {
    const int W = 32 * (16 / (int)sizeof(T));
    const size_t FR = (size_t)wx * R + 2 * kResidentRowHalo, FP = (size_t)wy * W + 2 * kResidentColHalo;
    const size_t nwarps = (size_t)wy * wx;
    return (4 * FR * FP + 2 * nwarps * 2 * W + 2 * nwarps * 2 * R) * sizeof(T);
}
inline size_t resident_sync_words(int ntiles) { return 32 * (size_t)(1 + ntiles); }
// device scratch of a launch: [counters][256 bytes: corner values][bands]
template <typename T> inline int resident_band_len(const ResidentPlan& pl) {
    const int W = 32 * (16 / (int)sizeof(T));
    const int n = pl.wy * W > pl.wx * pl.R ? pl.wy * W : pl.wx * pl.R;
    return (n + 31) / 32 * 32;
}
template <typename T> inline size_t resident_scratch_bytes(const ResidentPlan& pl) {
    const size_t nt = (size_t)pl.tiles_x * pl.tiles_y;
    return resident_sync_words((int)nt) * sizeof(unsigned) + 256 + 2 * nt * 4 * (size_t)resident_band_len<T>(pl) * sizeof(T) * 2;
}

}  // namespace fb
