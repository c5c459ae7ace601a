// warp_kernels.cuh -- the three hot-path kernels of the cavity projection timestep, written as
// warp-synchronous *streaming* stencils for sm_100a.
//
// Common shape (all three kernels):
//   * fields are SoA, row x = `pitch` contiguous elements along y (the reference's device
//     order, fluid_cpp/cuda/ParallelNavierStokes.cu:22, minus the 92-byte AoS cell);
//   * one WARP owns a strip of 32*V contiguous y-columns (V elements = 16 bytes per lane,
//     so every global access is a coalesced 128-bit load/store) and marches along x;
//   * x-neighbours live in the lane's own registers (a sliding window of rows),
//     y-neighbours of the lane's outer elements come from the adjacent lane by one
//     warp shuffle -- no block barrier, no atomics; shared memory only as a per-lane
//     prefetch ring (cp.async), never to exchange data between lanes;
//   * strips overlap by a 2*H column halo and x-chunks by the stencil radius, the overlap is
//     recomputed (a few percent) so that warps never communicate.
//
// The Jacobi kernel additionally blocks K sweeps in TIME: level t of the pipeline lags level
// t-1 by one row, all K levels of a row strip stay in registers, and HBM sees one read of p,
// one read of rhs and one write of p per K sweeps. Jacobi (p -> p_next, then copy; reference
// SerialNavierStokes.cpp:140-198) has no intra-sweep dependence, so this is bit-identical to
// K separate sweeps. The boundary rows of every intermediate sweep are applied inside the
// pipeline (see jacobi_task).
//
// Arithmetic contract: operand order and rounding points of the
// reference's serial fp32 path are reproduced exactly with *_rn intrinsics (never contracted
// into FMAs), IEEE division and no flush-to-zero; the fp64 Laplacian tail of the fp32 path is
// kept in fp64.
//
// This header is also compiled for the HOST by tests/emu (FLUIDB200_HOST_EMU) where a warp
// is 32 threads and a shuffle is a barrier exchange; that build is test infrastructure only.
#pragma once
#include <stdint.h>

#include "warp_args.h"

#ifdef FLUIDB200_HOST_EMU
#include "emu_shim.h"
#else
#include <cuda_runtime.h>
#define FB_DEV __device__ __forceinline__
#define FB_DEV_NOINLINE __device__ __noinline__
#define FB_EMU_CHECK(p, n, what) ((void)0)
#endif

namespace fb {

// ---------------------------------------------------------------------------------------
// exact-rounding arithmetic
// ---------------------------------------------------------------------------------------
template <typename T> struct Ar;

// The two divisions of the path come in three flavours, selected by MODE (Metric::mode):
//   DIV_EXACT  the reference's operations literally (IEEE divisions);
//   DIV_FAST   a division-free evaluation that is provably equal except for inputs it flags in
//              `bad`; the caller evaluates a flagged row a second time (FallbackMode below). Flags are
//              OR-ed, never branched on, so the ~40 independent chains of a row can overlap. fp32: one
//              product in double per division; fp64: reciprocal multiply + two FMA remainder corrections;
//   DIV_POW2   hx and hy are powers of two (unit spacing, 2^k+1 points on a unit box, ...): every
//              division is an exact scaling, i.e. one correctly rounded multiplication.
enum { DIV_EXACT = 0, DIV_FAST = 1, DIV_POW2 = 2, DIV_POW2_WIDE = 3, DIV_FAST_CHECKED = 4 };
// What the cells a mode flags are recomputed in (each of these costs the same for every input, unlike an
// IEEE division, whose slow path the subnormal values at the front of an impulsively started flow take):
//   DIV_POW2 -> DIV_POW2_WIDE. With power-of-two spacings every division of the reference is an exact
//     scaling, so the flagged cells (results in the subnormal range, where the float product would round
//     twice) need no division either: the Laplacian numerator is formed in double, scaled and rounded
//     once, which is the reference's evaluation to the letter.
//   DIV_FAST -> DIV_FAST_CHECKED -> DIV_EXACT (fp32). The same products as DIV_FAST, each one verified by
//     rounding the two ends of an interval that holds the reference's double (Ar<float>::rounds_alike);
//     only a cell that fails this (a tie, or nearly one) takes the divisions.
//   fp64: DIV_FAST -> DIV_EXACT (numerators whose corrections would underflow or overflow).
template <typename T, int MODE> struct FallbackMode {
    static constexpr int value = sizeof(T) != 4 ? DIV_EXACT : MODE == DIV_POW2 ? DIV_POW2_WIDE : MODE == DIV_FAST ? DIV_FAST_CHECKED : DIV_EXACT;
};
#ifndef FLUIDB200_EXACT_WIDE
#define FLUIDB200_EXACT_WIDE 1   // fp32 DIV_EXACT central differences go through one double division (0: __fdiv_rn)
#endif

// What the division-free modes collect over the cells of one row strip. The row is recomputed with
// DIV_EXACT when raised() says so; nothing branches per cell.
struct Flags {
    bool bad = false;        // DIV_FAST: some cell was flagged
    // DIV_POW2: over the Laplacian numerators t of the row, the minimum of (bits(|t|) - 1), where a zero
    // wraps to the top and so never counts as small, and the maximum of bits(|t|): two integer
    // min/max per value instead of a range test per cell
    unsigned lo = 0xffffffffu, hi = 0u;
    FB_DEV bool raised(unsigned pow2_lo_m1) const { return bad || lo < pow2_lo_m1 || hi >= 0x7f800000u; }
};

template <> struct Ar<float> {
    static FB_DEV float add(float a, float b) { return __fadd_rn(a, b); }
    static FB_DEV float sub(float a, float b) { return __fsub_rn(a, b); }
    static FB_DEV float mul(float a, float b) { return __fmul_rn(a, b); }
    // (a - b) / 2. / h of the reference: float subtract, exact halving in double, one double
    // division, one rounding to float. 53 >= 2*24+2, so the double rounding is innocuous and
    // the result equals the correctly rounded float quotient d/(2h), d = a-b; h2 = 2h is exact.
    //
    // DIV_FAST: q~ = (double)d * RN64(1/h2) is within 2^-52 (relative) of d/h2, while the quotient
    // of two 24-bit floats is never closer than ~2^-49 (relative) to a rounding boundary of a
    // NORMAL float result, so RN32(q~) is the correctly rounded quotient. Results in or near the
    // subnormal range (coarser, absolute rounding grid; exact ties are possible) are
    // flagged. The cost is data independent, unlike __fdiv_rn whose slow path the
    // mostly-zero fields of an impulsively started cavity hit all the time.
    // DIV_POW2: d * (1/h2) is the same real number as d / h2 and is rounded once, like the division.
    template <int MODE>
    static FB_DEV float cdiff(float a, float b, float h, float h2, double inv_h2, float inv_h2f, Flags& fl) {
        (void)h; (void)h2;
        const float d = __fsub_rn(a, b);
#if FLUIDB200_EXACT_WIDE
        // the reference's own evaluation, literally: d / 2. (exact) and one double division, rounded to float.
        // Every float, subnormals included, is a NORMAL double, so the cost does not depend on the data.
        if (MODE == DIV_EXACT) return __double2float_rn(__ddiv_rn(__dmul_rn((double)d, 0.5), (double)h));
#else
        if (MODE == DIV_EXACT) return __fdiv_rn(d, h2);
#endif
        if (MODE == DIV_POW2 || MODE == DIV_POW2_WIDE) return __fmul_rn(d, inv_h2f);
        // (float -> double is exact for every input, subnormals included; inf/nan propagate like the division)
        const double q = __dmul_rn((double)d, inv_h2);
        const float r = __double2float_rn(q);
        if (MODE == DIV_FAST_CHECKED) fl.bad |= !rounds_alike(q);
        else fl.bad |= (d != 0.0f) && (fabsf(r) < 7.52316385e-37f /* 2^-120 */);
        return r;
    }
    // q~ and the reference's double T are both within 2^-52 (relative) of the exact quotient, so T lies
    // inside q~ (1 +- 2^-50) even after the two products below are rounded. Rounding to float is monotone:
    // when both ends of that interval round to the same float, so do q~ and T. (DIV_FAST_CHECKED: the
    // second evaluation of a cell DIV_FAST flagged, typically for a result in the subnormal range, whose
    // coarse, absolute float grid allows ties -- if very rarely.)
    static FB_DEV bool rounds_alike(double q) {
        const float lo = __double2float_rn(__dmul_rn(q, 1.0 - 0x1p-50));
        const float hi = __double2float_rn(__dmul_rn(q, 1.0 + 0x1p-50));
        return lo == hi;
    }
    // (s - 4.*c) / hx / hy with s already rounded to float: the one place the fp32 path is
    // observably mixed precision (SerialNavierStokes.cpp:308-311). 4.*c is exact in double, so
    // s - 4.*c is a single rounding == fma(-4, c, s). Reference value: RN32(T), T = RN64(RN64(d/hx)/hy).
    //
    // DIV_FAST (Ziv's strategy): q~ = d * RN64(1/(hx*hy)) and T are both within 2^-52 (relative) of
    // the exact quotient, hence within 4 ulp64 of each other; RN32(q~) = RN32(T) unless a float
    // rounding boundary lies between them, i.e. unless the 29 bits of q~ below float precision
    // are within a few units of the midpoint pattern. Those (about 1 in 2^24), special inputs and
    // results near the subnormal range are flagged.
    // DIV_POW2: T = D * 2^k exactly with D = RN64(s - 4c), and for a normal result
    // RN32(D * 2^k) = RN32(D) * 2^k = fmaf(-4, c, s) * 2^k (double rounding of a float sum is
    // innocuous); values at the edge of the subnormal range are flagged.
    template <int MODE>
    static FB_DEV float lap_tail(float s, float c, const Metric<float>& m, Flags& fl) {
        const float hx = m.hx, hy = m.hy, inv_hxhyf = m.inv_hxhyf;
        const double inv_hxhy = m.inv_hxhy;
        if (MODE == DIV_EXACT) {
            double t = __fma_rn(-4.0, (double)c, (double)s);
            t = __ddiv_rn(t, (double)hx);
            t = __ddiv_rn(t, (double)hy);
            return __double2float_rn(t);
        }
        if (MODE == DIV_POW2) {
            const float t = __fmaf_rn(-4.0f, c, s);
            const float r = __fmul_rn(t, inv_hxhyf);
            // t == 0 or pow2_lo <= |t| < inf (then r is normal too), checked per row by Flags::raised
            const unsigned ta = __float_as_uint(t) & 0x7fffffffu;
            fl.lo = min(fl.lo, ta - 1u);
            fl.hi = max(fl.hi, ta);
            return r;
        }
        const double d = __fma_rn(-4.0, (double)c, (double)s);
        const double q = __dmul_rn(d, inv_hxhy);
        const float r = __double2float_rn(q);
        if (MODE == DIV_POW2_WIDE) return r;   // q == T: the two divisions are exact scalings
        // distance of the 29 sub-float bits of q from the midpoint pattern, as one unsigned compare
        const unsigned low = (unsigned)__double2loint(q) & 0x1fffffffu;
        if (MODE == DIV_FAST_CHECKED) fl.bad |= !rounds_alike(q);   // (|T - q~| <= 2^-51 |q~| here too)
        else fl.bad |= (low - 0x0ffffff0u) <= 32u || (fabsf(r) < 7.52316385e-37f && d != 0.0);
        return r;
    }
    // t *= -0.25 : the double multiply is exact, so this is one float rounding (subnormals).
    static FB_DEV float neg_quarter(float t) { return __fmul_rn(t, -0.25f); }
};

template <> struct Ar<double> {
    static FB_DEV double add(double a, double b) { return __dadd_rn(a, b); }
    static FB_DEV double sub(double a, double b) { return __dsub_rn(a, b); }
    static FB_DEV double mul(double a, double b) { return __dmul_rn(a, b); }
    // DIV_POW2 (hx, hy powers of two): x / h and x * (1/h) are the same real number rounded once, for
    // every x including subnormal results, so each division becomes one multiplication by an exact
    // reciprocal and nothing needs a flag.
    // DIV_FAST (any other spacing): the quotient by a constant h in five fp64 instructions instead of
    // __ddiv_rn (a reciprocal seed, its refinement and a range branch: the predictor spent 4.1 ms of a
    // 8192^2 step in them). y = RN(1/h), correctly rounded (the host's division); q0 = RN(d y) is within
    // 2 ulp of d / h; the remainder d - q h of such a q is a double, so each fused multiply-add computes it
    // exactly; one correction makes q faithful, and for a faithful q and a correctly rounded reciprocal
    // the next correction RN(q + r y) is the correctly rounded quotient (Markstein, IBM J. Res. Dev. 34,
    // 1990; make_metric keeps divisors whose significand is all ones, the one doubtful case, out of
    // this mode) -- as long as nothing underflows or overflows on the way: d = 0 or 2^-800 <= |d| < 2^800
    // with the spacings bounded by make_metric; anything else is flagged and redone with IEEE divisions.
    static FB_DEV double quot(double d, double h, double y, Flags& fl) {
        const double q0 = __dmul_rn(d, y);
        double r = __fma_rn(-q0, h, d);
        double q = __fma_rn(r, y, q0);
        r = __fma_rn(-q, h, d);
        q = __fma_rn(r, y, q);
        const unsigned hi = (unsigned)__double2hiint(d) & 0x7fffffffu;
        const bool zero = (hi | (unsigned)__double2loint(d)) == 0u;
        fl.bad |= !zero && (hi - 0x0df00000u /* 2^-800 */) >= (0x71f00000u /* 2^800 */ - 0x0df00000u);
        return zero ? q0 : q;   // (a zero keeps its sign: the corrections would add +0 to -0)
    }
    template <int MODE> static FB_DEV double cdiff(double a, double b, double h, double h2, double inv_h2, double, Flags& fl) {
        const double d = __dsub_rn(a, b);
        if (MODE == DIV_FAST) return quot(d, h2, inv_h2, fl);   // d / (2h) is the same real number as (d / 2.) / h, rounded once either way
        const double half = __dmul_rn(d, 0.5);  // "/ 2." is an exact halving
        if (MODE == DIV_POW2) return __dmul_rn(half, inv_h2 * 2.0);
        return __ddiv_rn(half, h);
    }
    template <int MODE> static FB_DEV double lap_tail(double s, double c, const Metric<double>& m, Flags& fl) {
        const double hx = m.hx, hy = m.hy, inv_hxhy = m.inv_hxhy;
        double d = __fma_rn(-4.0, c, s);  // 4.*c exact -> single rounding
        if (MODE == DIV_POW2) {
            d = __dmul_rn(d, inv_hxhy * hy);       // 1/hx, exact
            return __dmul_rn(d, inv_hxhy * hx);    // 1/hy, exact
        }
        if (MODE == DIV_FAST) {   // two divisions, each one rounded like the reference's; 2 RN(1/(2h)) = RN(1/h) exactly
            d = quot(d, hx, __dmul_rn(2.0, m.inv_hx2), fl);
            return quot(d, hy, __dmul_rn(2.0, m.inv_hy2), fl);
        }
        d = __ddiv_rn(d, hx);
        return __ddiv_rn(d, hy);
    }
    static FB_DEV double neg_quarter(double t) { return __dmul_rn(t, -0.25); }
};

// ---------------------------------------------------------------------------------------
// 16-byte lane vectors
// ---------------------------------------------------------------------------------------
template <typename T, int V> struct alignas(sizeof(T) * V) Vec {
    T a[V];
};
template <typename T> struct VecWidth;
template <> struct VecWidth<float> { static constexpr int V = 4; };
template <> struct VecWidth<double> { static constexpr int V = 2; };

template <typename T, int V> FB_DEV Vec<T, V> vzero() {
    Vec<T, V> r;
#pragma unroll
    for (int e = 0; e < V; ++e) r.a[e] = T(0);
    return r;
}
// read-only streaming load (inputs never alias outputs in the kernels that use it)
template <typename T, int V> FB_DEV Vec<T, V> ld_ro(const T* p) {
#ifdef FLUIDB200_HOST_EMU
    FB_EMU_CHECK(p, sizeof(Vec<T, V>), "load");
    return *reinterpret_cast<const Vec<T, V>*>(p);
#else
    const int4 r = __ldg(reinterpret_cast<const int4*>(p));
    Vec<T, V> v;
    *reinterpret_cast<int4*>(&v) = r;
    return v;
#endif
}
template <typename T, int V> FB_DEV Vec<T, V> ld_rw(const T* p) {
    FB_EMU_CHECK(p, sizeof(Vec<T, V>), "load");
    return *reinterpret_cast<const Vec<T, V>*>(p);
}
template <typename T, int V> FB_DEV void st_vec(T* p, const Vec<T, V>& v) {
    FB_EMU_CHECK(p, sizeof(Vec<T, V>), "store");
    *reinterpret_cast<Vec<T, V>*>(p) = v;
}


// V consecutive flag bytes as one integer (byte e = flag of element e); the address is V-aligned
template <int V> FB_DEV unsigned load_flags(const uint8_t* p) {
    static_assert(V == 2 || V == 4, "16 bytes per lane: 4 floats or 2 doubles");
    if (V == 4) return *reinterpret_cast<const uint32_t*>(p);
    return *reinterpret_cast<const uint16_t*>(p);
}

// Per-lane view of a strip.
template <int V, int H> struct Strip {
    static constexpr int W = 32 * V;      // columns loaded by the warp
    static constexpr int S = W - 2 * H;   // columns owned (stride between strips)
    int lane, col0;                       // first column of this lane
    int lcol;                             // column used for LOADS: col0 clamped into the row, so
                                          // loads never need a predicate (such lanes never store)
    bool has_bottom, has_top, ld_ok, st_ok;
    unsigned top_bits;                    // bit e set <=> element e of this lane is column Y-1
    int ym2_lane, ym2_e;                  // where column Y-2 lives (valid when has_top)
    FB_DEV void init(const Geom& g, int strip, int lane_) {
        lane = lane_;
        const int ys = strip * S;
        col0 = ys + lane * V;
        has_bottom = (strip == 0);
        has_top = (ys <= g.Y - 1) && (g.Y - 1 < ys + W);
        const int top_lane = (g.Y - 1 - ys) / V, top_e = (g.Y - 1 - ys) % V;
        top_bits = (has_top && lane == top_lane) ? (1u << top_e) : 0u;
        ym2_lane = (g.Y - 2 - ys) / V;
        ym2_e = (g.Y - 2 - ys) % V;
        ld_ok = (long long)col0 + V <= g.pitch;
        lcol = ld_ok ? col0 : (int)(g.pitch - V);
        const int vlo = has_bottom ? 0 : ys + H;
        const int vhi = (strip == g.nstrips - 1) ? g.Y : ys + W - H;
        st_ok = ld_ok && col0 >= vlo && col0 < vhi;
    }
};

template <typename T> FB_DEV T shfl_up1(T x) { return __shfl_up_sync(0xffffffffu, x, 1); }
template <typename T> FB_DEV T shfl_down1(T x) { return __shfl_down_sync(0xffffffffu, x, 1); }
template <typename T> FB_DEV T shfl_idx(T x, int src) { return __shfl_sync(0xffffffffu, x, src); }

// y-neighbours of every element of a lane vector: ym[e] = value at column-1, yp[e] = column+1
template <typename T, int V>
FB_DEV void y_neighbours(const Vec<T, V>& c, Vec<T, V>& ym, Vec<T, V>& yp) {
    const T from_prev = shfl_up1(c.a[V - 1]);
    const T from_next = shfl_down1(c.a[0]);
#pragma unroll
    for (int e = 0; e < V; ++e) {
        ym.a[e] = (e > 0) ? c.a[e - 1] : from_prev;
        yp.a[e] = (e < V - 1) ? c.a[e + 1] : from_next;
    }
}

// The value held at lane `lane_`, element `e_` of a row spread over the warp (warp-uniform call).
template <typename T, int V> FB_DEV T pick(const Vec<T, V>& row, int lane_, int e_) {
    T r = T(0);
#pragma unroll
    for (int e = 0; e < V; ++e) {
        const T s = shfl_idx(row.a[e], lane_ & 31);
        r = (e == e_) ? s : r;
    }
    return r;
}

// resolved flag/value of one edge cell
template <typename T> struct EdgeCell {
    bool use;
    T v;
};
template <typename T> FB_DEV EdgeCell<T> edge_cell(const Edges<T>& e, int edge, int idx) {
    EdgeCell<T> r;
    const int m = e.mode[edge];
    if (m == EDGE_UNSET) {
        r.use = false; r.v = T(0);
    } else if (m == EDGE_CONST) {
        r.use = true; r.v = e.cval[edge];
    } else {
        r.use = e.set[edge][idx] != 0; r.v = e.val[edge][idx];
    }
    return r;
}

// The two y-edge cells of interior row `row`: the boundary value when the flag is set, else a copy
// of column 1 of `bottom_src` / column Y-2 of `top_src` (SerialNavierStokes.cpp:163-175; the
// velocity version copies u_next at the top, :246). Element selection uses compile-time indices
// and a per-lane bit mask so the vectors stay in registers.
template <typename T, int V, int H>
FB_DEV void apply_y_edges(const Edges<T>& e, const Strip<V, H>& st, int row, Vec<T, V>& out,
                          const Vec<T, V>& bottom_src, const Vec<T, V>& top_src) {
    if (st.has_bottom) {
        const EdgeCell<T> b = edge_cell(e, BOTTOM, row);
        const T nv = b.use ? b.v : bottom_src.a[1];
        out.a[0] = (st.lane == 0) ? nv : out.a[0];
    }
    if (st.has_top) {
        const EdgeCell<T> t = edge_cell(e, TOP, row);
        const T below = shfl_up1(top_src.a[V - 1]);
        Vec<T, V> cand;
#pragma unroll
        for (int k = 0; k < V; ++k) cand.a[k] = (k > 0) ? top_src.a[k > 0 ? k - 1 : 0] : below;
#pragma unroll
        for (int k = 0; k < V; ++k) out.a[k] = ((st.top_bits >> k) & 1u) ? (t.use ? t.v : cand.a[k]) : out.a[k];
    }
}

// The same two cells for y edges whose boundary value does not depend on the row (UNSET = copy the
// neighbour, CONST): everything that is uniform over the rows is decided once per task, and a row
// costs a shuffle and a few selects, without branches. Used by the lean Jacobi pipeline.
template <typename T, int V> struct YEdgeUniform {
    bool lane0_bottom, bot_use, top_use;
    T bot_v, top_v;
    unsigned top_bits;
    template <int H> FB_DEV void init(const Edges<T>& e, const Strip<V, H>& st) {
        lane0_bottom = st.has_bottom && st.lane == 0;
        bot_use = e.mode[BOTTOM] == EDGE_CONST;
        top_use = e.mode[TOP] == EDGE_CONST;
        bot_v = e.cval[BOTTOM];
        top_v = e.cval[TOP];
        top_bits = st.top_bits;  // 0 in every lane of a strip without column Y-1
    }
    FB_DEV void apply(Vec<T, V>& out) const {
        const T below = shfl_up1(out.a[V - 1]);
        const T nb = bot_use ? bot_v : out.a[1];
        Vec<T, V> nt;
#pragma unroll
        for (int k = 0; k < V; ++k) nt.a[k] = top_use ? top_v : (k > 0 ? out.a[k > 0 ? k - 1 : 0] : below);
#pragma unroll
        for (int k = V - 1; k >= 0; --k) out.a[k] = ((top_bits >> k) & 1u) ? nt.a[k] : out.a[k];
        out.a[0] = lane0_bottom ? nb : out.a[0];
    }
};

// store only interior columns [1, Y-2] of a row (edge cells of u_next/v_next/rhs are never
// written by the reference, SerialNavierStokes.cpp:299-300,326-327)
template <typename T, int V>
FB_DEV void st_interior(T* rowp, int col0, int Y, const Vec<T, V>& v) {
    if (col0 >= 1 && col0 + V - 1 <= Y - 2) {
        st_vec<T, V>(rowp + col0, v);
    } else {
#pragma unroll
        for (int e = 0; e < V; ++e)
            if (col0 + e >= 1 && col0 + e <= Y - 2) rowp[col0 + e] = v.a[e];
    }
}

// ---------------------------------------------------------------------------------------
// Per-lane asynchronous prefetch ring (cp.async / LDGSTS into shared memory).
// Every lane copies its OWN 16 bytes of a row into its own slot and later reads the same 16 bytes
// back, so completion is tracked per thread (cp.async.wait_group) and no barrier of any kind is
// needed. Register prefetching did not survive the compiler: under the register cap ptxas sank
// the LDGs to just before their first use (ncu: 43 % of all stall samples sat on that first use),
// whereas an asynchronous copy is issued where it is written and holds no registers in flight.
// D slots x 2 arrays x 16 B per thread.
// ---------------------------------------------------------------------------------------
template <typename T, int V, int D> struct LaneRing {
#ifdef FLUIDB200_HOST_EMU
    Vec<T, V> buf[D][2];
    FB_DEV void init(int, int) {}
    FB_DEV void fetch(int slot, int arr, const T* src) {
        FB_EMU_CHECK(src, sizeof(Vec<T, V>), "prefetch");
        buf[slot][arr] = *reinterpret_cast<const Vec<T, V>*>(src);
    }
    FB_DEV Vec<T, V> read(int slot, int arr) const { return buf[slot][arr]; }
    static FB_DEV void commit() {}
    template <int N> static FB_DEV void wait() {}
#else
    unsigned char* mine;   // generic address of this thread's 16 bytes in slot 0, array 0
    unsigned stride;       // bytes between consecutive (slot, array) planes
    FB_DEV void init(int tid, int nthreads) {
        extern __shared__ __align__(16) unsigned char fb_ring_smem[];
        mine = fb_ring_smem + tid * 16;
        stride = (unsigned)nthreads * 16u;
    }
    FB_DEV void fetch(int slot, int arr, const T* src) {
        const unsigned dst = (unsigned)__cvta_generic_to_shared(mine + (slot * 2 + arr) * stride);
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
    }
    FB_DEV Vec<T, V> read(int slot, int arr) const {
        return *reinterpret_cast<const Vec<T, V>*>(mine + (slot * 2 + arr) * stride);
    }
    static FB_DEV void commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
    template <int N> static FB_DEV void wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }
#endif
};
// ring depth used by the Jacobi fast path for blocking depth K (U = steps per unrolled iteration)
// Unroll factor U of the fast loop. U = K (rounded up to even) lets the row windows and the
// right-hand-side shift register rotate by register renaming alone. Round 1 ran K >= 7 with 4 steps per
// iteration because the 8-step body (27 KB then) fell out of the instruction caches; after that round's
// instruction diet the 8-step body is 22 KB (1405 instructions, 34 moves instead of 180 per 8 steps) and is
// the faster one again: 7.36 -> 7.18 ms for the 100 sweeps of a 16384^2 step (profiles/README_r2.md).
// Build-time knobs (0 = the default rule), kept for such measurements:
#ifndef FLUIDB200_U_DEEP
#define FLUIDB200_U_DEEP 8   // K = 7, 8
#endif
#ifndef FLUIDB200_U_MID
#define FLUIDB200_U_MID 0    // K = 5, 6
#endif
template <int K> struct JacobiRing {
    static constexpr int U0 = (K + 1) & ~1;
    static constexpr int U = (K >= 7 && FLUIDB200_U_DEEP) ? FLUIDB200_U_DEEP : ((K == 5 || K == 6) && FLUIDB200_U_MID) ? FLUIDB200_U_MID : U0;
    static constexpr int D0 = U0 < 4 ? 4 : U0;
    static constexpr int D = (D0 % U == 0) ? D0 : ((D0 + U - 1) / U) * U;
};

// =========================================================================================
// Jacobi pressure-Poisson sweeps, K sweeps per launch, boundary rows fused in.
// Replaces poisson_step_kernel + enforce_pressure_BC_kernel + update_pressure_kernel
// (fluid_cpp/cuda/ParallelNavierStokes.cu:172-192, :14-42, :44-56).
// =========================================================================================

// The x = 0 (LEFT) or x = X-1 (RIGHT) boundary row of one sweep level, derived from that
// level's adjacent interior row `inner` (SerialNavierStokes.cpp:177-188). Corner elements are
// placeholders here (corners feed no 5-point stencil); wall_row_out() fixes them for output.
template <typename T, int V>
FB_DEV Vec<T, V> wall_row(const Edges<T>& e, int edge, const Vec<T, V>& inner, int col0, int Y) {
    Vec<T, V> r = inner;
    const int m = e.mode[edge];
    if (m != EDGE_UNSET) {
#pragma unroll
        for (int k = 0; k < V; ++k) {
            const int col = col0 + k;
            if (col >= 1 && col <= Y - 2) {
                if (m == EDGE_CONST) r.a[k] = e.cval[edge];
                else if (e.set[edge][col]) r.a[k] = e.val[edge][col];
            }
        }
    }
    return r;
}

// Final-level boundary row including the corners. `x` is 0 or X-1, `prev_wall` is the same
// boundary row one sweep earlier: an unset corner copies its y-neighbour on the wall BEFORE
// that neighbour is refreshed (the y-edge loop runs first, SerialNavierStokes.cpp:163-175 vs
// :177-188), i.e. it is one sweep stale.
template <typename T, int V, int H>
FB_DEV Vec<T, V> wall_row_out(const Edges<T>& e, int edge, int x, const Vec<T, V>& inner,
                              const Vec<T, V>& prev_wall, const Strip<V, H>& st, int Y) {
    Vec<T, V> r = wall_row<T, V>(e, edge, inner, st.col0, Y);
    if (st.has_bottom) {
        const EdgeCell<T> b = edge_cell(e, BOTTOM, x);
        const T nv = b.use ? b.v : prev_wall.a[1];  // column 1 (V >= 2)
        r.a[0] = (st.lane == 0) ? nv : r.a[0];
    }
    if (st.has_top) {
        const EdgeCell<T> t = edge_cell(e, TOP, x);
        const T stale = pick<T, V>(prev_wall, st.ym2_lane, st.ym2_e);
        const T nv = t.use ? t.v : stale;
#pragma unroll
        for (int k = 0; k < V; ++k) r.a[k] = ((st.top_bits >> k) & 1u) ? nv : r.a[k];
    }
    return r;
}

#if !defined(FLUIDB200_NO_PACKED_F32X2) && !defined(FLUIDB200_HOST_EMU)
FB_DEV float2 fsub2(float2 a, float2 b) {  // two correctly rounded differences in one instruction
    unsigned long long xa, xb, xr;
    float2 r;
    asm("mov.b64 %0, {%1,%2};" : "=l"(xa) : "f"(a.x), "f"(a.y));
    asm("mov.b64 %0, {%1,%2};" : "=l"(xb) : "f"(b.x), "f"(b.y));
    asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(xr) : "l"(xa), "l"(xb));
    asm("mov.b64 {%0,%1}, %2;" : "=f"(r.x), "=f"(r.y) : "l"(xr));
    return r;
}
#endif

// One Jacobi update of a row strip: out = ((b - (((xm + xp) + yp) + ym)) * -0.25
// (SerialNavierStokes.cpp:144-146, operand order kept). ym / yp: the y-neighbours of every element.
template <typename T, int V>
FB_DEV Vec<T, V> jacobi_row_yn(const Vec<T, V>& xm, const Vec<T, V>& xp, const Vec<T, V>& ym, const Vec<T, V>& yp,
                               const Vec<T, V>& b) {
    using A_ = Ar<T>;
    Vec<T, V> out;
#if !defined(FLUIDB200_NO_PACKED_F32X2) && !defined(FLUIDB200_HOST_EMU)
    if constexpr (sizeof(T) == 4 && V == 4) {
        // sm_100 packed fp32: one FADD2 rounds two independent sums exactly like two FADDs. Only the
        // first sum and the subtraction pair up without register moves; the multiplication stays
        // scalar because ptxas (12.9) contracts mul.rn.f32x2 + add.rn.f32x2 into FFMA2.
        float2 s0 = __fadd2_rn(make_float2(xm.a[0], xm.a[1]), make_float2(xp.a[0], xp.a[1]));
        float2 s1 = __fadd2_rn(make_float2(xm.a[2], xm.a[3]), make_float2(xp.a[2], xp.a[3]));
        s0.x = __fadd_rn(__fadd_rn(s0.x, yp.a[0]), ym.a[0]);
        s0.y = __fadd_rn(__fadd_rn(s0.y, yp.a[1]), ym.a[1]);
        s1.x = __fadd_rn(__fadd_rn(s1.x, yp.a[2]), ym.a[2]);
        s1.y = __fadd_rn(__fadd_rn(s1.y, yp.a[3]), ym.a[3]);
        const float2 d0 = fsub2(make_float2(b.a[0], b.a[1]), s0), d1 = fsub2(make_float2(b.a[2], b.a[3]), s1);
        out.a[0] = __fmul_rn(d0.x, -0.25f);
        out.a[1] = __fmul_rn(d0.y, -0.25f);
        out.a[2] = __fmul_rn(d1.x, -0.25f);
        out.a[3] = __fmul_rn(d1.y, -0.25f);
        return out;
    }
#endif
#pragma unroll
    for (int e = 0; e < V; ++e) {
        const T s = A_::add(A_::add(A_::add(xm.a[e], xp.a[e]), yp.a[e]), ym.a[e]);
        out.a[e] = A_::neg_quarter(A_::sub(b.a[e], s));
    }
    return out;
}
template <typename T, int V>
FB_DEV Vec<T, V> jacobi_row(const Vec<T, V>& xm, const Vec<T, V>& c, const Vec<T, V>& xp, const Vec<T, V>& b) {
    Vec<T, V> ym, yp;
    y_neighbours<T, V>(c, ym, yp);
    return jacobi_row_yn<T, V>(xm, xp, ym, yp, b);
}

// Pipeline: at step r the warp loads row r of p (level 0) and level t = 1..K produces row r-t
// from rows r-t-1, r-t, r-t+1 of level t-1. w[t] holds the two newest rows of level t, bw[j]
// the scaled right-hand side of row r-1-j. Steps whose K rows are all plain interior rows take
// the FAST path (no per-level range checks, no wall logic); pipeline fill/drain at the box walls
// and masked sweeps take the GENERAL path. Both act on the same register state.
template <typename T, int V, int K, bool MASK>
FB_DEV void jacobi_task(const JacobiArgs<T>& A, int strip, int chunk, int lane, int tid, int nthreads) {
    constexpr int H = (K + 3) & ~3;
    using A_ = Ar<T>;
    using VT = Vec<T, V>;
    const Geom& g = A.g;
    const int X = g.X, Y = g.Y;
    Strip<V, H> st;
    st.init(g, strip, lane);

    int c0, c1;
    jacobi_chunk_rows(A, chunk, &c0, &c1);
    if (c0 >= c1) return;
    const int slab_hi = g.row0 + g.rows;
    const int r_begin = max(c0 - K, g.row0);
    const int load_hi = min(c1 + K, slab_hi);
    const int r_end = c1 - 1 + K;

    const long long base = (long long)st.lcol - (long long)g.row0 * g.pitch;
    const T* pin = A.p_in + base;
    const T* rin = A.rhs + base;
    T* pout = A.p_out + (long long)st.col0 - (long long)g.row0 * g.pitch;
    const bool y_edges = st.has_bottom || st.has_top;
    const bool ye_uniform = A.e.mode[BOTTOM] != EDGE_ARRAY && A.e.mode[TOP] != EDGE_ARRAY;
    YEdgeUniform<T, V> ye;
    ye.init(A.e, st);

    VT w[K][2];
    VT bw[K];
#pragma unroll
    for (int t = 0; t < K; ++t) {
        w[t][0] = vzero<T, V>();
        w[t][1] = vzero<T, V>();
        bw[t] = vzero<T, V>();
    }

    // (rhs * hx) * hy, SerialNavierStokes.cpp:144 -- once per row, reused by all K levels; with unit
    // spacing both multiplications are by 1 and exact, so they are skipped (warp-uniform branch)
    const bool scale_b = !(A.hx == T(1) && A.hy == T(1));
    auto push_b = [&](VT bnew) {
        if (scale_b) {
#pragma unroll
            for (int e = 0; e < V; ++e) bnew.a[e] = A_::mul(A_::mul(bnew.a[e], A.hx), A.hy);
        }
#pragma unroll
        for (int t = K - 1; t > 0; --t) bw[t] = bw[t - 1];
        bw[0] = bnew;
    };

    // Rows reach the pipeline through the per-lane cp.async ring, D steps ahead, on BOTH paths: one
    // commit group per step r = {row r of p, row r-1 of rhs}, in slot (r - r_begin) % D. Row indices
    // are clamped to rows that exist; a row the reference would not read (beyond the loaded range,
    // rhs of a wall row) is replaced by zeros when it is taken out of the ring.
    constexpr int U = (K + 1) & ~1;
    constexpr int D = JacobiRing<K>::D;
    static_assert(D % U == 0, "ring depth must be a multiple of the unroll factor");
    const int pmax = load_hi - 1, bmax = min(load_hi - 1, X - 2);
    LaneRing<T, V, D> ring;
    ring.init(tid, nthreads);
    auto fetch_step = [&](int slot, int rr) {
        ring.fetch(slot, 0, pin + (long long)min(rr, pmax) * g.pitch);
        ring.fetch(slot, 1, rin + (long long)max(min(rr - 1, bmax), r_begin) * g.pitch);
        ring.commit();
    };
#pragma unroll
    for (int j = 0; j < D; ++j) fetch_step(j, r_begin + j);

    // steps r in [fast_lo, fast_hi]: every level's row r-t lies in [2, X-3]. Strips that hold a
    // y-edge column and masked sweeps always take the general path. The fast loop is unrolled U
    // steps deep (U even and >= K) so that the row windows rotate by register renaming instead of
    // moves; it starts on a step whose rows sit in slot 0 so that its slot numbers are constants.
    int fast_lo = max(r_begin, K + 2);
    fast_lo += (D - (fast_lo - r_begin) % D) % D;
    int nfast = min(r_end, X - 2) - fast_lo + 1;
    if (MASK || y_edges || nfast < 0) nfast = 0;
    if ((long long)(load_hi - r_begin + 1) * g.pitch >= (1LL << 31)) nfast = 0;  // 32-bit offsets in the fast loop
    nfast -= nfast % U;
    const int fast_hi = fast_lo + nfast - 1;

    unsigned fw[K], fnext = 0u;  // MASK: flags of rows r-1 .. r-K, and of row r (nothing precedes r_begin)
#pragma unroll
    for (int t = 0; t < K; ++t) fw[t] = 0u;
    int r = r_begin;
#pragma unroll 1
    for (int phase = 0; phase < 2; ++phase) {
        // ---------------- GENERAL path: [r, fast_lo) first, then (fast_hi, r_end] ----------------
        const int gen_hi = (phase == 0 && nfast > 0) ? fast_lo - 1 : r_end;
        int gslot = (r - r_begin) % D;
#pragma unroll 1
        for (; r <= gen_hi; ++r) {
            ring.template wait<D - 1>();          // the oldest group = this step's rows has landed
            VT cur = ring.read(gslot, 0);
            VT bnew = ring.read(gslot, 1);
            fetch_step(gslot, r + D);
            gslot = (gslot + 1 == D) ? 0 : gslot + 1;
            if (r >= load_hi) cur = vzero<T, V>();
            if (!(r - 1 >= r_begin && r - 1 < load_hi && r - 1 >= 1 && r - 1 <= X - 2)) bnew = vzero<T, V>();
            push_b(bnew);
            if (MASK) {
                // The flags of a row travel down the levels like its right-hand side: read once, one step
                // before level 1 needs them (V flag bytes in one load: col0 and the pitch are multiples
                // of V), instead of once per level with the load latency inside every level.
#pragma unroll
                for (int t = K - 1; t > 0; --t) fw[t] = fw[t - 1];
                fw[0] = fnext;
                fnext = (st.ld_ok && r < load_hi)
                            ? load_flags<V>(A.mset + (long long)(r - g.row0) * g.pitch + st.col0) : 0u;
            }
#pragma unroll
            for (int t = 1; t <= K; ++t) {
                const int rr = r - t;  // row this level produces now
                VT out = cur;
                if (rr >= 1 && rr <= X - 2) {
                    VT xm = w[t - 1][0];
                    const VT c = w[t - 1][1];
                    VT xp = cur;
                    if (t > 1) {
                        // levels >= 1 never materialise their boundary rows: rebuild them from the
                        // adjacent interior row of the same level
                        if (rr == 1) xm = wall_row<T, V>(A.e, LEFT, c, st.col0, Y);
                        if (rr == X - 2) xp = wall_row<T, V>(A.e, RIGHT, c, st.col0, Y);
                    }
                    out = jacobi_row<T, V>(xm, c, xp, bw[t - 1]);
                    if (MASK) {  // interior flags (:154-160); fw[t-1] = this lane's flags of row rr
                        const unsigned f4 = fw[t - 1];
                        if (f4) {  // rare: only then are the values fetched
                            const long long off = (long long)(rr - g.row0) * g.pitch + st.col0;
                            const VT mv = *reinterpret_cast<const VT*>(A.mval + off);
#pragma unroll
                            for (int e = 0; e < V; ++e) {
                                const int col = st.col0 + e;
                                if (col >= 1 && col <= Y - 2 && ((f4 >> (8 * e)) & 0xffu)) out.a[e] = mv.a[e];
                            }
                        }
                    }
                    if (y_edges) {  // (:163-175)
                        if (ye_uniform) ye.apply(out);  // constant / copy edges: a shuffle and a few selects
                        else apply_y_edges<T, V, H>(A.e, st, rr, out, out, out);
                    }
                    if (t == K) {
                        if (rr >= c0 && rr < c1 && st.st_ok) st_vec<T, V>(pout + (long long)rr * g.pitch, out);
                        if (rr == 1 && c0 == 0) {
                            const VT wr = wall_row_out<T, V, H>(A.e, LEFT, 0, out, xm, st, Y);
                            if (st.st_ok) st_vec<T, V>(pout, wr);
                        }
                        if (rr == X - 2 && c1 == X) {
                            const VT wr = wall_row_out<T, V, H>(A.e, RIGHT, X - 1, out, xp, st, Y);
                            if (st.st_ok) st_vec<T, V>(pout + (long long)(X - 1) * g.pitch, wr);
                        }
                    }
                }
                w[t - 1][0] = w[t - 1][1];
                w[t - 1][1] = cur;
                cur = out;
            }
        }
        if (phase == 1 || nfast == 0) break;
        // ---------------- FAST path: [fast_lo, fast_hi], U steps per iteration ----------------
        // Addresses inside the loop are 32-bit element offsets from row r_begin of this chunk (the
        // host keeps a chunk below 2^31 elements), advanced by one pitch per step and clamped.
        const T* pbase = pin + (long long)r_begin * g.pitch;
        const T* bbase = rin + (long long)r_begin * g.pitch;
        T* obase = pout + (long long)r_begin * g.pitch;
        const unsigned pitch32 = (unsigned)g.pitch;
        const unsigned pmax_off = (unsigned)(pmax - r_begin) * pitch32;
        const unsigned bmax_off = (unsigned)(bmax - r_begin) * pitch32;
        unsigned poff = (unsigned)(min(r + D, pmax) - r_begin) * pitch32;          // next p row to fetch
        unsigned boff = (unsigned)(min(r + D - 1, bmax) - r_begin) * pitch32;      // next rhs row to fetch
        int soff = (r - K - r_begin) * (int)pitch32;                               // row stored by this step
        const int first_store = c0 + K;  // step at which level K reaches row c0
        int sbase = 0;                   // ring slot of step r (advances by U per iteration, mod D)
#pragma unroll 1
        for (; r <= fast_hi; r += U) {
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int slot = (D == U) ? u : sbase + u;
                ring.template wait<D - 1>();
                VT cur = ring.read(slot, 0);
                push_b(ring.read(slot, 1));
                ring.fetch(slot, 0, pbase + poff);
                ring.fetch(slot, 1, bbase + boff);
                ring.commit();
                poff = min(poff + pitch32, pmax_off);
                boff = min(boff + pitch32, bmax_off);
#pragma unroll
                for (int t = 1; t <= K; ++t) {
                    const VT out = jacobi_row<T, V>(w[t - 1][0], w[t - 1][1], cur, bw[t - 1]);
                    w[t - 1][0] = w[t - 1][1];
                    w[t - 1][1] = cur;
                    cur = out;
                }
                if (st.st_ok && r + u >= first_store) st_vec<T, V>(obase + soff, cur);
                soff += (int)pitch32;
            }
            if (D != U) sbase = (sbase + U) % D;
        }
        // (fast_hi + 1 - r_begin is a multiple of U, so the general path resumes in the slot the loop left)
    }
    ring.template wait<0>();  // nothing may still be landing in shared memory when the CTA retires
}

// The FAST-ONLY twin of jacobi_task: same pipeline, same arithmetic, but compiled without any of
// the general path (walls, y-edges, masks, range checks), which is what sets the register count
// of jacobi_task (K=6: 166 -> 128 registers, K=8: 214 -> 162), i.e. one more resident CTA per SM.
// The host (split_jacobi in api.cpp) only sends it tasks for which EVERY step is a plain step in x:
// r_begin = c0-K >= max(row0, K+2), r_end = c1-1+K <= X-2. YEDGE = false: the strip holds no y-edge
// column (FAST kernel); YEDGE = true: it does and the two y-edge cells of every level are added,
// branch-free, for constant / copy edges (EDGE kernel). SCALE_B: the grid is not unit spaced.
// The step count is rounded UP to a multiple of U; the extra steps read clamped rows and are
// never stored.
template <typename T, int V, int K, bool SCALE_B, bool YEDGE>
FB_DEV void jacobi_fast_body(const JacobiArgs<T>& A, int strip, int chunk, int lane, int tid, int nthreads) {
    constexpr int H = (K + 3) & ~3;
    constexpr int U = JacobiRing<K>::U, D = JacobiRing<K>::D;
    static_assert(D % U == 0, "ring depth must be a multiple of the unroll factor");
    using A_ = Ar<T>;
    using VT = Vec<T, V>;
    const Geom& g = A.g;
    Strip<V, H> st;
    st.init(g, strip, lane);
    int c0, c1;
    jacobi_chunk_rows(A, chunk, &c0, &c1);
    if (c0 >= c1) return;
    const int r_begin = c0 - K;
    const int load_hi = min(c1 + K, g.row0 + g.rows);
    const int r_end = c1 - 1 + K;
    const int pmax = load_hi - 1, bmax = min(load_hi - 1, g.X - 2);

    const long long base = (long long)st.lcol + (long long)(r_begin - g.row0) * g.pitch;
    const T* pbase = A.p_in + base;
    const T* bbase = A.rhs + base;
    // row stored by step r is r-K; the pointer runs along with the steps and is only dereferenced
    // for rows of the chunk
    T* optr = A.p_out + (long long)st.col0 + (long long)(r_begin - K - g.row0) * g.pitch;
    const unsigned pitch32 = (unsigned)g.pitch;
    const unsigned pmax_off = (unsigned)(pmax - r_begin) * pitch32;
    const unsigned bmax_off = (unsigned)(bmax - r_begin) * pitch32;

    VT w[K][2];
    VT bw[K];
#pragma unroll
    for (int t = 0; t < K; ++t) {
        w[t][0] = vzero<T, V>();
        w[t][1] = vzero<T, V>();
        bw[t] = vzero<T, V>();
    }

    YEdgeUniform<T, V> ye;
    if (YEDGE) ye.init(A.e, st);

    LaneRing<T, V, D> ring;
    ring.init(tid, nthreads);
#pragma unroll
    for (int j = 0; j < D; ++j) {
        ring.fetch(j, 0, pbase + min((unsigned)j * pitch32, pmax_off));
        // rhs of row r_begin+j-1; for j = 0 that row feeds only pipeline-fill values: any row will do
        ring.fetch(j, 1, bbase + min((unsigned)(j > 0 ? j - 1 : 0) * pitch32, bmax_off));
        ring.commit();
    }
    unsigned poff = min((unsigned)D * pitch32, pmax_off);
    unsigned boff = min((unsigned)(D - 1) * pitch32, bmax_off);
    int sbase = 0;
#pragma unroll 1
    for (int r = r_begin; r <= r_end; r += U) {
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int slot = (D == U) ? u : sbase + u;
            ring.template wait<D - 1>();
            VT cur = ring.read(slot, 0);
            VT bnew = ring.read(slot, 1);
            ring.fetch(slot, 0, pbase + poff);
            ring.fetch(slot, 1, bbase + boff);
            ring.commit();
            poff = min(poff + pitch32, pmax_off);
            boff = min(boff + pitch32, bmax_off);
            if (SCALE_B) {
#pragma unroll
                for (int e = 0; e < V; ++e) bnew.a[e] = A_::mul(A_::mul(bnew.a[e], A.hx), A.hy);
            }
#pragma unroll
            for (int t = K - 1; t > 0; --t) bw[t] = bw[t - 1];
            bw[0] = bnew;
#pragma unroll
            for (int t = 1; t <= K; ++t) {
                VT out = jacobi_row<T, V>(w[t - 1][0], w[t - 1][1], cur, bw[t - 1]);
                // strips holding column 0 or Y-1: the y-edge cells of every level (:163-175)
                if (YEDGE) ye.apply(out);
                w[t - 1][0] = w[t - 1][1];
                w[t - 1][1] = cur;
                cur = out;
            }
            const int row = r + u - K;  // row level K produced in this step
            if (st.st_ok && row >= c0 && row < c1) st_vec<T, V>(optr, cur);
            optr += pitch32;
        }
        if (D != U) sbase = (sbase + U) % D;
    }
    ring.template wait<0>();  // nothing may still be landing in shared memory when the CTA retires
}

#if defined(FLUIDB200_TMA_RING) && !defined(FLUIDB200_HOST_EMU)
// EXPERIMENT (build with -DFLUIDB200_TMA_RING; not the default): the lean pipeline
// fed by 1-D bulk copies (cp.async.bulk, the TMA engine's linear mode) instead of the per-lane cp.async ring.
// One elected lane asks for the warp's 512 contiguous bytes of p and of rhs; an mbarrier per (warp, slot)
// counts the bytes in; every lane then reads its own 16 bytes. This GPU's memory only, no multicast.
FB_DEV void mbar_init(unsigned bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
FB_DEV void mbar_expect_tx(unsigned bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
FB_DEV void mbar_wait(unsigned bar, unsigned phase) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t}" ::"r"(bar), "r"(phase) : "memory");
}
FB_DEV void bulk_g2s(unsigned dst, const void* src, unsigned bytes, unsigned bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src), "r"(bytes),
                 "r"(bar)
                 : "memory");
}
template <typename T, int V, int K, bool SCALE_B>
FB_DEV void jacobi_fast_body_bulk(const JacobiArgs<T>& A, int strip, int chunk, int lane, int tid, int nthreads) {
    constexpr int H = (K + 3) & ~3;
    constexpr int U = JacobiRing<K>::U, D = JacobiRing<K>::D;
    static_assert(D == U, "the bulk-copy experiment keeps one ring slot per unrolled step");
    using A_ = Ar<T>;
    using VT = Vec<T, V>;
    constexpr unsigned ROW = 32 * 16;   // bytes of a warp's row piece
    const Geom& g = A.g;
    Strip<V, H> st;
    st.init(g, strip, lane);
    int c0, c1;
    jacobi_chunk_rows(A, chunk, &c0, &c1);
    if (c0 >= c1) return;
    const int r_begin = c0 - K;
    const int load_hi = min(c1 + K, g.row0 + g.rows);
    const int r_end = c1 - 1 + K;
    const int pmax = load_hi - 1, bmax = min(load_hi - 1, g.X - 2);
    // lane 0's addresses: the start of the warp's 512 bytes
    const long long base0 = (long long)(st.col0 - lane * V) + (long long)(r_begin - g.row0) * g.pitch;
    const T* pbase = A.p_in + base0;
    const T* bbase = A.rhs + base0;
    T* optr = A.p_out + (long long)st.col0 + (long long)(r_begin - K - g.row0) * g.pitch;
    const unsigned pitch32 = (unsigned)g.pitch;
    const unsigned pmax_off = (unsigned)(pmax - r_begin) * pitch32;
    const unsigned bmax_off = (unsigned)(bmax - r_begin) * pitch32;

    extern __shared__ __align__(16) unsigned char fb_ring_smem[];
    const int warp = tid >> 5, nwarps = nthreads >> 5;
    // [warp][slot][p, rhs][512 B], then the mbarriers [warp][slot]
    const unsigned smem0 = (unsigned)__cvta_generic_to_shared(fb_ring_smem);
    const unsigned wbase = smem0 + (unsigned)warp * D * 2 * ROW;
    const unsigned bar0 = smem0 + (unsigned)nwarps * D * 2 * ROW + (unsigned)warp * D * 8;
    if (lane == 0) {
#pragma unroll
        for (int j = 0; j < D; ++j) mbar_init(bar0 + j * 8, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();

    VT w[K][2];
    VT bw[K];
#pragma unroll
    for (int t = 0; t < K; ++t) {
        w[t][0] = vzero<T, V>();
        w[t][1] = vzero<T, V>();
        bw[t] = vzero<T, V>();
    }
    if (lane == 0) {
#pragma unroll
        for (int j = 0; j < D; ++j) {
            mbar_expect_tx(bar0 + j * 8, 2 * ROW);
            bulk_g2s(wbase + j * 2 * ROW, pbase + min((unsigned)j * pitch32, pmax_off), ROW, bar0 + j * 8);
            bulk_g2s(wbase + j * 2 * ROW + ROW, bbase + min((unsigned)(j > 0 ? j - 1 : 0) * pitch32, bmax_off), ROW, bar0 + j * 8);
        }
    }
    unsigned poff = min((unsigned)D * pitch32, pmax_off);
    unsigned boff = min((unsigned)(D - 1) * pitch32, bmax_off);
    unsigned phase = 0;
    const unsigned char* mine = fb_ring_smem + (size_t)warp * D * 2 * ROW + lane * 16;
#pragma unroll 1
    for (int r = r_begin; r <= r_end; r += U) {
#pragma unroll
        for (int u = 0; u < U; ++u) {
            mbar_wait(bar0 + u * 8, phase);
            VT cur = *reinterpret_cast<const VT*>(mine + u * 2 * ROW);
            VT bnew = *reinterpret_cast<const VT*>(mine + u * 2 * ROW + ROW);
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // my reads come before the next bulk write of this slot
            __syncwarp();
            if (lane == 0) {
                mbar_expect_tx(bar0 + u * 8, 2 * ROW);
                bulk_g2s(wbase + u * 2 * ROW, pbase + poff, ROW, bar0 + u * 8);
                bulk_g2s(wbase + u * 2 * ROW + ROW, bbase + boff, ROW, bar0 + u * 8);
            }
            poff = min(poff + pitch32, pmax_off);
            boff = min(boff + pitch32, bmax_off);
            if (SCALE_B) {
#pragma unroll
                for (int e = 0; e < V; ++e) bnew.a[e] = A_::mul(A_::mul(bnew.a[e], A.hx), A.hy);
            }
#pragma unroll
            for (int t = K - 1; t > 0; --t) bw[t] = bw[t - 1];
            bw[0] = bnew;
#pragma unroll
            for (int t = 1; t <= K; ++t) {
                VT out = jacobi_row<T, V>(w[t - 1][0], w[t - 1][1], cur, bw[t - 1]);
                w[t - 1][0] = w[t - 1][1];
                w[t - 1][1] = cur;
                cur = out;
            }
            const int row = r + u - K;
            if (st.st_ok && row >= c0 && row < c1) st_vec<T, V>(optr, cur);
            optr += pitch32;
        }
        phase ^= 1u;
    }
    // every copy that was asked for has to land before the CTA may retire
#pragma unroll
    for (int u = 0; u < D; ++u) mbar_wait(bar0 + u * 8, phase);
}
#endif

// Masked split launch: does the footprint of task (strip, chunk) -- every row and column its pipeline
// loads -- hold a flagged interior cell? Tile granularity (JacobiArgs::mtiles); all lanes read the
// same few bytes.
template <typename T, int V, int K>
FB_DEV bool jacobi_task_flagged(const JacobiArgs<T>& A, int strip, int chunk) {
    constexpr int H = (K + 3) & ~3, W = 32 * V, S = W - 2 * H;
    const Geom& g = A.g;
    int c0, c1;
    jacobi_chunk_rows(A, chunk, &c0, &c1);
    if (c0 >= c1) return false;
    const int r0 = max(c0 - K, g.row0), r1 = min(c1 + K, g.row0 + g.rows);
    const int y0 = strip * S, y1 = min(strip * S + W, g.Y);
    if (r0 >= r1 || y0 >= y1) return false;
    const int tr1 = (r1 - 1 - g.row0) / kMaskTileRows, tc0 = y0 / kMaskTileCols, tc1 = (y1 - 1) / kMaskTileCols;
    unsigned any = 0;
    for (int tr = (r0 - g.row0) / kMaskTileRows; tr <= tr1; ++tr)
        for (int tc = tc0; tc <= tc1; ++tc) any |= A.mtiles[(long long)tr * A.mtiles_nc + tc];
    return any != 0;
}

// The scaling of the right-hand side by hx*hy (two exact multiplications by 1 on a unit-spaced grid)
// is decided once per task, not once per row.
template <typename T, int V, int K>
FB_DEV void jacobi_fast_task(const JacobiArgs<T>& A, int strip, int chunk, int lane, int tid, int nthreads) {
    if (A.mtiles && jacobi_task_flagged<T, V, K>(A, strip, chunk)) return;  // the general kernel runs it
#if defined(FLUIDB200_TMA_RING) && !defined(FLUIDB200_HOST_EMU)
    if constexpr (JacobiRing<K>::D == JacobiRing<K>::U) {
        if (A.hx == T(1) && A.hy == T(1)) jacobi_fast_body_bulk<T, V, K, false>(A, strip, chunk, lane, tid, nthreads);
        else jacobi_fast_body_bulk<T, V, K, true>(A, strip, chunk, lane, tid, nthreads);
        return;
    }
#endif
    if (A.hx == T(1) && A.hy == T(1)) jacobi_fast_body<T, V, K, false, false>(A, strip, chunk, lane, tid, nthreads);
    else jacobi_fast_body<T, V, K, true, false>(A, strip, chunk, lane, tid, nthreads);
}

// One task of the EDGE kernel of a split launch: a strip that holds a y-edge column, in a chunk
// with no x wall within reach of the pipeline. The lean pipeline with the y-edge cells added.
template <typename T, int V, int K>
FB_DEV void jacobi_edge_task(const JacobiArgs<T>& A, int strip, int chunk, int lane, int tid, int nthreads) {
    if (A.mtiles && jacobi_task_flagged<T, V, K>(A, strip, chunk)) return;  // the general kernel runs it
    if (A.hx == T(1) && A.hy == T(1)) jacobi_fast_body<T, V, K, false, true>(A, strip, chunk, lane, tid, nthreads);
    else jacobi_fast_body<T, V, K, true, true>(A, strip, chunk, lane, tid, nthreads);
}

// One task of the general kernel. In a masked split launch it is offered every task and keeps the
// ones no lean kernel runs: wall chunks, strips without a lean kernel, tasks with a flagged cell.
template <typename T, int V, int K, bool MASK>
FB_DEV void jacobi_general_task(const JacobiArgs<T>& A, int strip, int chunk, int lane, int tid, int nthreads) {
    if (MASK && A.part == 2 && A.gen_all) {
        const bool lean_chunk = chunk >= A.cf_lo && chunk < A.cf_hi;
        const bool lean_strip = (strip >= A.sf_lo && strip < A.sf_hi) || A.edge_lean;
        if (lean_chunk && lean_strip && !jacobi_task_flagged<T, V, K>(A, strip, chunk)) return;
    }
    jacobi_task<T, V, K, MASK>(A, strip, chunk, lane, tid, nthreads);
}

// =========================================================================================
// Fused tentative velocity + divergence right-hand side ("predictor").
// Replaces unified_timestep_kernel + compute_righthand_kernel
// (fluid_cpp/cuda/ParallelNavierStokes.cu:113-148, :150-170).
// =========================================================================================

// A pair of results per cell (u and v component).
template <typename T> struct UV {
    T u, v;
};

// u_next, v_next of ONE interior cell from its own values (uc, vc) and its four neighbours in u
// and v (SerialNavierStokes.cpp:298-322, operand order kept). All arguments are scalars so that
// the out-of-line exact variant below gets them in registers: passing the row vectors by reference
// would pin them to local memory in the hot loop.
template <typename T, int MODE>
FB_DEV UV<T> tentative_cell(const PredictorArgs<T>& A, T uc, T vc, T uxm, T uxp, T uym, T uyp, T vxm, T vxp, T vym,
                            T vyp, Flags& fl) {
    using A_ = Ar<T>;
    const T du_dx = A_::template cdiff<MODE>(uxp, uxm, A.m.hx, A.m.hx2, A.m.inv_hx2, A.m.inv_hx2f, fl);
    const T dv_dx = A_::template cdiff<MODE>(vxp, vxm, A.m.hx, A.m.hx2, A.m.inv_hx2, A.m.inv_hx2f, fl);
    const T du_dy = A_::template cdiff<MODE>(uyp, uym, A.m.hy, A.m.hy2, A.m.inv_hy2, A.m.inv_hy2f, fl);
    const T dv_dy = A_::template cdiff<MODE>(vyp, vym, A.m.hy, A.m.hy2, A.m.inv_hy2, A.m.inv_hy2f, fl);
    const T su = A_::add(A_::add(A_::add(uxm, uxp), uyp), uym);
    const T sv = A_::add(A_::add(A_::add(vxm, vxp), vyp), vym);
    const T u_lap = A_::template lap_tail<MODE>(su, uc, A.m, fl);
    const T v_lap = A_::template lap_tail<MODE>(sv, vc, A.m, fl);
    const T du_dt = A_::sub(A_::sub(A_::mul(A.nu, u_lap), A_::mul(uc, du_dx)), A_::mul(vc, du_dy));
    const T dv_dt = A_::sub(A_::sub(A_::mul(A.nu, v_lap), A_::mul(uc, dv_dx)), A_::mul(vc, dv_dy));
    UV<T> r;
    r.u = A_::add(uc, A_::mul(A.dt, du_dt));
    r.v = A_::add(vc, A_::mul(A.dt, dv_dt));
    return r;
}
template <typename T, int MODE>
FB_DEV_NOINLINE UV<T> tentative_cell_exact(const PredictorArgs<T>& A, T uc, T vc, T uxm, T uxp, T uym, T uyp, T vxm,
                                           T vxp, T vym, T vyp) {
    constexpr int FB = FallbackMode<T, MODE>::value;
    Flags fl;
    UV<T> r = tentative_cell<T, FB>(A, uc, vc, uxm, uxp, uym, uyp, vxm, vxp, vym, vyp, fl);
    if (FB == DIV_FAST_CHECKED && fl.bad) r = tentative_cell<T, DIV_EXACT>(A, uc, vc, uxm, uxp, uym, uyp, vxm, vxp, vym, vyp, fl);
    return r;
}

// u_next, v_next of one interior row strip from rows x-1 (um,vm), x (u0,v0), x+1 (up,vp) of u, v.
// All cells are first evaluated in MODE without branching; if the division-free arithmetic raised a
// flag on any of this lane's cells they are recomputed with IEEE divisions (per lane; the shuffles
// stay outside).
template <typename T, int V, int MODE>
FB_DEV void tentative_row(const PredictorArgs<T>& A, const Vec<T, V>& um, const Vec<T, V>& u0, const Vec<T, V>& up_,
                          const Vec<T, V>& vm, const Vec<T, V>& v0, const Vec<T, V>& vp_, Vec<T, V>& us, Vec<T, V>& vs) {
    Vec<T, V> uym, uyp, vym, vyp;
    y_neighbours<T, V>(u0, uym, uyp);
    y_neighbours<T, V>(v0, vym, vyp);
    Flags fl;
#pragma unroll
    for (int e = 0; e < V; ++e) {
        const UV<T> r = tentative_cell<T, MODE>(A, u0.a[e], v0.a[e], um.a[e], up_.a[e], uym.a[e], uyp.a[e], vm.a[e],
                                                vp_.a[e], vym.a[e], vyp.a[e], fl);
        us.a[e] = r.u;
        vs.a[e] = r.v;
    }
    if (MODE != DIV_EXACT && fl.raised(A.m.pow2_lo_m1)) {
#pragma unroll
        for (int e = 0; e < V; ++e) {
            const UV<T> r = tentative_cell_exact<T, MODE>(A, u0.a[e], v0.a[e], um.a[e], up_.a[e], uym.a[e], uyp.a[e],
                                                    vm.a[e], vp_.a[e], vym.a[e], vyp.a[e]);
            us.a[e] = r.u;
            vs.a[e] = r.v;
        }
    }
}

// right-hand side of one interior cell from u_next at x-1, x+1 and v_next at y-1, y+1
// (SerialNavierStokes.cpp:325-335)
template <typename T, int MODE>
FB_DEV T rhs_cell(const PredictorArgs<T>& A, T us_m, T us_p, T vs_ym, T vs_yp, Flags& fl) {
    using A_ = Ar<T>;
    const T du_next_dx = A_::template cdiff<MODE>(us_p, us_m, A.m.hx, A.m.hx2, A.m.inv_hx2, A.m.inv_hx2f, fl);
    const T dv_next_dy = A_::template cdiff<MODE>(vs_yp, vs_ym, A.m.hy, A.m.hy2, A.m.inv_hy2, A.m.inv_hy2f, fl);
    return A_::mul(A.rho_over_dt, A_::add(du_next_dx, dv_next_dy));
}
template <typename T, int MODE>
FB_DEV_NOINLINE T rhs_cell_exact(const PredictorArgs<T>& A, T us_m, T us_p, T vs_ym, T vs_yp) {
    constexpr int FB = FallbackMode<T, MODE>::value;
    Flags fl;
    T r = rhs_cell<T, FB>(A, us_m, us_p, vs_ym, vs_yp, fl);
    if (FB == DIV_FAST_CHECKED && fl.bad) r = rhs_cell<T, DIV_EXACT>(A, us_m, us_p, vs_ym, vs_yp, fl);
    return r;
}
template <typename T, int V, int MODE>
FB_DEV Vec<T, V> rhs_row(const PredictorArgs<T>& A, const Vec<T, V>& us_m, const Vec<T, V>& us_p, const Vec<T, V>& vs0) {
    Vec<T, V> vym, vyp, q;
    y_neighbours<T, V>(vs0, vym, vyp);
    Flags fl;
#pragma unroll
    for (int e = 0; e < V; ++e) q.a[e] = rhs_cell<T, MODE>(A, us_m.a[e], us_p.a[e], vym.a[e], vyp.a[e], fl);
    if (MODE != DIV_EXACT && fl.raised(A.m.pow2_lo_m1)) {
#pragma unroll
        for (int e = 0; e < V; ++e) q.a[e] = rhs_cell_exact<T, MODE>(A, us_m.a[e], us_p.a[e], vym.a[e], vyp.a[e]);
    }
    return q;
}


#ifndef FLUIDB200_PRED_D
#define FLUIDB200_PRED_D 4   // rows of u and v in flight per lane in the predictor's fast loop
#endif
// Pipeline: step r loads row r of u, v; level A produces u_next, v_next of row r-1, level B the
// right-hand side of row r-2 (it needs u_next of rows r-3 and r-1). Interior strips away from the
// box walls take the FAST path (2 steps per iteration, no range checks).
template <typename T, int V, int MODE>
FB_DEV void predictor_task(const PredictorArgs<T>& A, int strip, int chunk, int lane, int tid, int nthreads) {
    constexpr int H = 4;
    using VT = Vec<T, V>;
    const Geom& g = A.g;
    const int X = g.X, Y = g.Y;
    Strip<V, H> st;
    st.init(g, strip, lane);

    const int c0 = A.lo + chunk * g.rows_per_chunk;
    const int c1 = min(c0 + g.rows_per_chunk, A.hi);
    if (c0 >= c1) return;
    const int slab_hi = g.row0 + g.rows;
    // rows this task stores
    const int s0 = max(c0, A.star_lo), s1 = min(c1, A.star_hi);
    const int b0 = max(c0, A.rhs_lo), b1 = min(c1, A.rhs_hi);
    const int r_begin = max(c0 - 2, g.row0);
    const int load_hi = min(c1 + 2, slab_hi);
    const int r_end = c1 + 1;

    const long long lbase = (long long)st.lcol - (long long)g.row0 * g.pitch;
    const long long sbase = (long long)st.col0 - (long long)g.row0 * g.pitch;
    const T* up = A.u + lbase;
    const T* vp = A.v + lbase;
    const bool y_edges = st.has_bottom || st.has_top;

    VT wu[2] = {vzero<T, V>(), vzero<T, V>()};   // u rows r-2, r-1
    VT wv[2] = {vzero<T, V>(), vzero<T, V>()};
    VT wus[2] = {vzero<T, V>(), vzero<T, V>()};  // u_next rows r-3, r-2
    VT wvs = vzero<T, V>();                      // v_next row r-2

    auto load = [&](const T* f, int r) -> VT {
        if (r >= r_begin && r < load_hi) return ld_ro<T, V>(f + (long long)r * g.pitch);
        return vzero<T, V>();
    };

    // fast steps: level A row r-1 in [1, X-2] and level B rows r-3, r-1 both computed interior rows
    constexpr int U = 2;
    const int fast_lo = max(r_begin, 4);
    int nfast = min(r_end, X - 1) - fast_lo + 1;
    if (y_edges || nfast < 0) nfast = 0;
    if ((long long)(load_hi - r_begin + 1) * g.pitch >= (1LL << 31)) nfast = 0;  // 32-bit offsets in the fast loop
    nfast -= nfast % U;
    const int fast_hi = fast_lo + nfast - 1;

    VT nu_ = load(up, r_begin), nv_ = load(vp, r_begin);
    int r = r_begin;
#pragma unroll 1
    for (int phase = 0; phase < 2; ++phase) {
        const int gen_hi = (phase == 0 && nfast > 0) ? fast_lo - 1 : r_end;
#pragma unroll 1
        for (; r <= gen_hi; ++r) {
            const VT cu = nu_, cv = nv_;
            nu_ = load(up, r + 1);
            nv_ = load(vp, r + 1);
            // ---- level A: u_next, v_next of row ra = r-1
            const int ra = r - 1;
            VT us = vzero<T, V>(), vs = vzero<T, V>();
            if (ra >= 1 && ra <= X - 2) {
                tentative_row<T, V, MODE>(A, wu[0], wu[1], cu, wv[0], wv[1], cv, us, vs);
                // edge columns of u_next/v_next keep whatever is stored (never written by the
                // reference); the right-hand side below reads them as y-neighbours
                // (rows outside the loaded range only feed discarded pipeline-fill values: skip the loads)
                const bool in_slab = ra >= r_begin && ra < load_hi;
                const long long ro = in_slab ? (long long)(ra - g.row0) * g.pitch : 0;
                if (st.has_bottom && in_slab) {
                    FB_EMU_CHECK(A.us + ro, sizeof(T), "edge load");
                    const T ue = A.us[ro], ve = A.vs[ro];
                    us.a[0] = (lane == 0) ? ue : us.a[0];
                    vs.a[0] = (lane == 0) ? ve : vs.a[0];
                }
                if (st.has_top && in_slab) {
                    FB_EMU_CHECK(A.us + ro + (Y - 1), sizeof(T), "edge load");
                    const T ue = A.us[ro + (Y - 1)], ve = A.vs[ro + (Y - 1)];
#pragma unroll
                    for (int e = 0; e < V; ++e) {
                        us.a[e] = ((st.top_bits >> e) & 1u) ? ue : us.a[e];
                        vs.a[e] = ((st.top_bits >> e) & 1u) ? ve : vs.a[e];
                    }
                }
                if (ra >= s0 && ra < s1 && st.st_ok) {
                    st_interior<T, V>(A.us + ro, st.col0, Y, us);
                    st_interior<T, V>(A.vs + ro, st.col0, Y, vs);
                }
            } else if ((ra == 0 || ra == X - 1) && ra >= r_begin && ra < load_hi) {
                // wall rows of u_next / v_next: stored values (x-neighbours of rows 1 and X-2)
                us = ld_rw<T, V>(A.us + lbase + (long long)ra * g.pitch);
                vs = ld_rw<T, V>(A.vs + lbase + (long long)ra * g.pitch);
            }
            // ---- level B: rhs of row rb = r-2
            const int rb = r - 2;
            if (rb >= 1 && rb <= X - 2) {
                const VT q = rhs_row<T, V, MODE>(A, wus[0], us, wvs);
                if (rb >= b0 && rb < b1 && st.st_ok)
                    st_interior<T, V>(A.rhs + (long long)(rb - g.row0) * g.pitch, st.col0, Y, q);
            }
            wu[0] = wu[1]; wu[1] = cu;
            wv[0] = wv[1]; wv[1] = cv;
            wus[0] = wus[1]; wus[1] = us;
            wvs = vs;
        }
        if (phase == 1 || nfast == 0) break;
        // ---------------- FAST path ----------------
        const int pmax = load_hi - 1;
        // rows of u and v are prefetched D steps ahead through the per-lane cp.async ring; inside the
        // loop the row to fetch is a 32-bit element offset from row r (clamped at the last loaded row)
        // and the three store pointers run along with the steps
        constexpr int D = FLUIDB200_PRED_D;
        static_assert(D % U == 0, "ring depth must be a multiple of the unroll factor");
        LaneRing<T, V, D> ring;
        ring.init(tid, nthreads);
        const T* ubase = up + (long long)r * g.pitch;
        const T* vbase = vp + (long long)r * g.pitch;
        const unsigned pitch32 = (unsigned)g.pitch;
        const unsigned pmax_off = (unsigned)(pmax - r) * pitch32;
#pragma unroll
        for (int j = 0; j < D; ++j) {
            const unsigned off = min((unsigned)j * pitch32, pmax_off);
            ring.fetch(j, 0, ubase + off);
            ring.fetch(j, 1, vbase + off);
            ring.commit();
        }
        unsigned poff = min((unsigned)D * pitch32, pmax_off);
        T* uso = A.us + sbase + (long long)(r - 1) * g.pitch;   // row ra = r-1 of u_next, v_next
        T* vso = A.vs + sbase + (long long)(r - 1) * g.pitch;
        T* rho = A.rhs + sbase + (long long)(r - 2) * g.pitch;  // row rb = r-2 of the right-hand side
        int ring_pos = 0;
#pragma unroll 1
        for (; r <= fast_hi; r += U) {
#pragma unroll
            for (int k = 0; k < U; ++k) {
                const int slot = ring_pos + k;
                ring.template wait<D - 1>();
                const VT cu = ring.read(slot, 0), cv = ring.read(slot, 1);
                ring.fetch(slot, 0, ubase + poff);
                ring.fetch(slot, 1, vbase + poff);
                ring.commit();
                poff = min(poff + pitch32, pmax_off);
                VT us, vs;
                tentative_row<T, V, MODE>(A, wu[0], wu[1], cu, wv[0], wv[1], cv, us, vs);
                const int ra = r + k - 1, rb = r + k - 2;
                if (st.st_ok && ra >= s0 && ra < s1) {
                    st_vec<T, V>(uso, us);
                    st_vec<T, V>(vso, vs);
                }
                const VT q = rhs_row<T, V, MODE>(A, wus[0], us, wvs);
                if (st.st_ok && rb >= b0 && rb < b1) st_vec<T, V>(rho, q);
                uso += pitch32;
                vso += pitch32;
                rho += pitch32;
                wu[0] = wu[1]; wu[1] = cu;
                wv[0] = wv[1]; wv[1] = cv;
                wus[0] = wus[1]; wus[1] = us;
                wvs = vs;
            }
            ring_pos = (ring_pos + U) % D;
        }
        ring.template wait<0>();
        nu_ = (r < load_hi) ? ring.read(ring_pos, 0) : vzero<T, V>();
        nv_ = (r < load_hi) ? ring.read(ring_pos, 1) : vzero<T, V>();
    }
}

// =========================================================================================
// Fused pressure-gradient correction + velocity boundary rows ("corrector").
// Replaces velocity_correction_kerenl + enforce_velocity_BC_kernel
// (fluid_cpp/cuda/ParallelNavierStokes.cu:194-217, :58-111).
// =========================================================================================

// One velocity component of a boundary row x = 0 / X-1 (SerialNavierStokes.cpp:236-294).
//   inner_copy : what an UNSET wall cell copies (u_next[1][y] on the left, corrected u[X-2][y]
//                on the right -- the reference's asymmetry, :269 vs :276)
//   old_wall   : this field's wall row BEFORE the step (unset bottom corner copies old [x][1])
//   star_wall  : u_next's wall row (unset top corner copies u_next[x][Y-2], :246)
template <typename T, int V, int H>
FB_DEV Vec<T, V> velocity_wall_row(const Edges<T>& e, int edge, int x, const Vec<T, V>& inner_copy,
                                   const Vec<T, V>& old_wall, const Vec<T, V>& star_wall,
                                   const Strip<V, H>& st, int Y) {
    Vec<T, V> r = wall_row<T, V>(e, edge, inner_copy, st.col0, Y);
    if (st.has_bottom) {
        const EdgeCell<T> b = edge_cell(e, BOTTOM, x);
        const T nv = b.use ? b.v : old_wall.a[1];
        r.a[0] = (st.lane == 0) ? nv : r.a[0];
    }
    if (st.has_top) {
        const EdgeCell<T> t = edge_cell(e, TOP, x);
        const T sv = pick<T, V>(star_wall, st.ym2_lane, st.ym2_e);
        const T nv = t.use ? t.v : sv;
#pragma unroll
        for (int k = 0; k < V; ++k) r.a[k] = ((st.top_bits >> k) & 1u) ? nv : r.a[k];
    }
    return r;
}

// corrected u, v of one interior cell from p at x-1, x+1, y-1, y+1 and u_next, v_next
// (SerialNavierStokes.cpp:338-350)
template <typename T, int MODE>
FB_DEV UV<T> correct_cell(const CorrectorArgs<T>& A, T pxm, T pxp, T pym, T pyp, T us, T vs, Flags& fl) {
    using A_ = Ar<T>;
    const T dp_dx = A_::template cdiff<MODE>(pxp, pxm, A.m.hx, A.m.hx2, A.m.inv_hx2, A.m.inv_hx2f, fl);
    const T dp_dy = A_::template cdiff<MODE>(pyp, pym, A.m.hy, A.m.hy2, A.m.inv_hy2, A.m.inv_hy2f, fl);
    UV<T> r;
    r.u = A_::sub(us, A_::mul(A.dt_over_rho, dp_dx));
    r.v = A_::sub(vs, A_::mul(A.dt_over_rho, dp_dy));
    return r;
}
template <typename T, int MODE>
FB_DEV_NOINLINE UV<T> correct_cell_exact(const CorrectorArgs<T>& A, T pxm, T pxp, T pym, T pyp, T us, T vs) {
    constexpr int FB = FallbackMode<T, MODE>::value;
    Flags fl;
    UV<T> r = correct_cell<T, FB>(A, pxm, pxp, pym, pyp, us, vs, fl);
    if (FB == DIV_FAST_CHECKED && fl.bad) r = correct_cell<T, DIV_EXACT>(A, pxm, pxp, pym, pyp, us, vs, fl);
    return r;
}
template <typename T, int V, int MODE>
FB_DEV void correct_row(const CorrectorArgs<T>& A, const Vec<T, V>& pm, const Vec<T, V>& p0, const Vec<T, V>& pp_,
                        const Vec<T, V>& us, const Vec<T, V>& vs, Vec<T, V>& un, Vec<T, V>& vn) {
    Vec<T, V> pym, pyp;
    y_neighbours<T, V>(p0, pym, pyp);
    Flags fl;
#pragma unroll
    for (int e = 0; e < V; ++e) {
        const UV<T> r = correct_cell<T, MODE>(A, pm.a[e], pp_.a[e], pym.a[e], pyp.a[e], us.a[e], vs.a[e], fl);
        un.a[e] = r.u;
        vn.a[e] = r.v;
    }
    if (MODE != DIV_EXACT && fl.raised(A.m.pow2_lo_m1)) {
#pragma unroll
        for (int e = 0; e < V; ++e) {
            const UV<T> r = correct_cell_exact<T, MODE>(A, pm.a[e], pp_.a[e], pym.a[e], pyp.a[e], us.a[e], vs.a[e]);
            un.a[e] = r.u;
            vn.a[e] = r.v;
        }
    }
}


template <typename T, int V, bool MASK, int MODE>
FB_DEV void corrector_task(const CorrectorArgs<T>& A, int strip, int chunk, int lane) {
    constexpr int H = 4;
    using VT = Vec<T, V>;
    const Geom& g = A.g;
    const int X = g.X, Y = g.Y;
    Strip<V, H> st;
    st.init(g, strip, lane);

    const int c0 = A.out_lo + chunk * g.rows_per_chunk;
    const int c1 = min(c0 + g.rows_per_chunk, A.out_hi);
    if (c0 >= c1) return;
    const int slab_hi = g.row0 + g.rows;
    const int r_begin = max(c0 - 1, g.row0);
    const int load_hi = min(c1 + 1, slab_hi);
    const int r_end = c1;

    const long long lbase = (long long)st.lcol - (long long)g.row0 * g.pitch;
    const long long sbase = (long long)st.col0 - (long long)g.row0 * g.pitch;
    const T* pp = A.p + lbase;
    const T* usp = A.us + lbase;
    const T* vsp = A.vs + lbase;
    T* up = A.u + sbase;
    T* vp = A.v + sbase;
    const bool y_edges = st.has_bottom || st.has_top;
    const int i_lo = max(c0, 1), i_hi = min(c1, X - 1);  // interior rows this task corrects: [i_lo, i_hi)

    VT wp[2] = {vzero<T, V>(), vzero<T, V>()};  // p rows r-2, r-1

    auto load = [&](const T* f, int r) -> VT {
        if (r >= r_begin && r < load_hi) return ld_ro<T, V>(f + (long long)r * g.pitch);
        return vzero<T, V>();
    };
    auto load_star = [&](const T* f, int r) -> VT {  // u_next / v_next of an interior row
        if (r >= i_lo && r < i_hi) return ld_ro<T, V>(f + (long long)r * g.pitch);
        return vzero<T, V>();
    };

    // fast steps: row rr = r-1 in [i_lo, i_hi) and not next to a wall (rows 1 and X-2 also emit the
    // boundary rows), interior strips, no interior flags
    constexpr int U = 2;
    const int fast_lo = max(max(i_lo, 2) + 1, r_begin + 2);
    int nfast = min(i_hi - 1, X - 3) + 1 - fast_lo + 1;
    if (MASK || y_edges || nfast < 0) nfast = 0;
    nfast -= nfast % U;
    const int fast_hi = fast_lo + nfast - 1;

    VT np_ = load(pp, r_begin);
    VT nus = load_star(usp, r_begin - 1), nvs = load_star(vsp, r_begin - 1);
    int r = r_begin;
#pragma unroll 1
    for (int phase = 0; phase < 2; ++phase) {
        const int gen_hi = (phase == 0 && nfast > 0) ? fast_lo - 1 : r_end;
#pragma unroll 1
        for (; r <= gen_hi; ++r) {
            const VT cp = np_;
            const VT us = nus, vs = nvs;
            np_ = load(pp, r + 1);
            nus = load_star(usp, r);
            nvs = load_star(vsp, r);
            const int rr = r - 1;
            if (rr >= i_lo && rr < i_hi) {
                VT un, vn;
                correct_row<T, V, MODE>(A, wp[0], wp[1], cp, us, vs, un, vn);
                if (MASK) {  // interior flags (SerialNavierStokes.cpp:224-233)
                    const long long off = (long long)(rr - g.row0) * g.pitch + st.col0;
                    if (st.ld_ok) {
#pragma unroll
                        for (int e = 0; e < V; ++e) {
                            const int col = st.col0 + e;
                            if (col >= 1 && col <= Y - 2) {
                                if (A.uset && A.uset[off + e]) un.a[e] = A.uval[off + e];
                                if (A.vset && A.vset[off + e]) vn.a[e] = A.vval[off + e];
                            }
                        }
                    }
                }
                // y-edges: bottom copies the corrected value (:239,:253), top copies u_next (:246,:260)
                if (y_edges) {
                    apply_y_edges<T, V, H>(A.eu, st, rr, un, un, us);
                    apply_y_edges<T, V, H>(A.ev, st, rr, vn, vn, vs);
                }
                if (st.st_ok) {
                    st_vec<T, V>(up + (long long)rr * g.pitch, un);
                    st_vec<T, V>(vp + (long long)rr * g.pitch, vn);
                }
                if (rr == 1 && c0 == 0) {  // boundary row x = 0: unset cells copy u_next[1][y] (:269,:283)
                    const VT ou = ld_rw<T, V>(A.u + lbase), ov = ld_rw<T, V>(A.v + lbase);
                    const VT su = ld_rw<T, V>(usp), sv = ld_rw<T, V>(vsp);
                    const VT wu_ = velocity_wall_row<T, V, H>(A.eu, LEFT, 0, us, ou, su, st, Y);
                    const VT wv_ = velocity_wall_row<T, V, H>(A.ev, LEFT, 0, vs, ov, sv, st, Y);
                    if (st.st_ok) {
                        st_vec<T, V>(up, wu_);
                        st_vec<T, V>(vp, wv_);
                    }
                }
                if (rr == X - 2 && c1 == X) {  // boundary row x = X-1: unset cells copy corrected u[X-2][y]
                    const long long o = (long long)(X - 1) * g.pitch;
                    const VT ou = ld_rw<T, V>(A.u + lbase + o), ov = ld_rw<T, V>(A.v + lbase + o);
                    const VT su = ld_rw<T, V>(usp + o), sv = ld_rw<T, V>(vsp + o);
                    const VT wu_ = velocity_wall_row<T, V, H>(A.eu, RIGHT, X - 1, un, ou, su, st, Y);
                    const VT wv_ = velocity_wall_row<T, V, H>(A.ev, RIGHT, X - 1, vn, ov, sv, st, Y);
                    if (st.st_ok) {
                        st_vec<T, V>(up + o, wu_);
                        st_vec<T, V>(vp + o, wv_);
                    }
                }
            }
            wp[0] = wp[1];
            wp[1] = cp;
        }
        if (phase == 1 || nfast == 0) break;
        // ---------------- FAST path: rows r-1 all plain interior rows of this task ----------------
        const int pmax = load_hi - 1, smax = i_hi - 1;
        VT pfp[U], pfu[U], pfv[U];
        pfp[0] = np_; pfu[0] = nus; pfv[0] = nvs;
#pragma unroll
        for (int k = 1; k < U; ++k) {
            pfp[k] = ld_ro<T, V>(pp + (long long)min(r + k, pmax) * g.pitch);
            pfu[k] = ld_ro<T, V>(usp + (long long)min(r + k - 1, smax) * g.pitch);
            pfv[k] = ld_ro<T, V>(vsp + (long long)min(r + k - 1, smax) * g.pitch);
        }
#pragma unroll 1
        for (; r <= fast_hi; r += U) {
#pragma unroll
            for (int k = 0; k < U; ++k) {
                const VT cp = pfp[k], us = pfu[k], vs = pfv[k];
                pfp[k] = ld_ro<T, V>(pp + (long long)min(r + k + U, pmax) * g.pitch);
                pfu[k] = ld_ro<T, V>(usp + (long long)min(r + k + U - 1, smax) * g.pitch);
                pfv[k] = ld_ro<T, V>(vsp + (long long)min(r + k + U - 1, smax) * g.pitch);
                VT un, vn;
                correct_row<T, V, MODE>(A, wp[0], wp[1], cp, us, vs, un, vn);
                if (st.st_ok) {
                    st_vec<T, V>(up + (long long)(r + k - 1) * g.pitch, un);
                    st_vec<T, V>(vp + (long long)(r + k - 1) * g.pitch, vn);
                }
                wp[0] = wp[1];
                wp[1] = cp;
            }
        }
        np_ = pfp[0]; nus = pfu[0]; nvs = pfv[0];
    }
}

}  // namespace fb
