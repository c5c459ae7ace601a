// resident_kernels.cuh -- the projection timestep for SMALL grids (41^2 ... about 1000^2), where
// the streaming kernels of warp_kernels.cuh are launch-latency bound (19 launches of 9-15 us per
// timestep). One cooperative launch runs num_iterations WHOLE timesteps -- predictor, K Jacobi
// sweeps with their boundary rows, corrector -- replacing the 4 + 3K launches per step of
// fluid_cpp/cuda/ParallelNavierStokes.cu:269-288.
//
// Layout: the grid is cut into tiles, one CTA (= one SM) per tile; a CTA is wx x wy warps; a warp
// covers 32 * V contiguous columns (V elements = 16 bytes per lane); a thread owns a PATCH of R rows
// x V columns whose pressure p and scaled right-hand side b stay in REGISTERS for the whole launch.
// One Jacobi sweep is then: fetch the four sides of the patch (rows from the warps above / below
// through shared memory, columns from the adjacent lanes by shuffle; across tiles through the
// pressure arrays in L2 with a release / acquire counter per tile), R rows of arithmetic, publish
// the patch's outer rows and columns, one __syncthreads. Nothing else touches memory.
//
// Same arithmetic as the streaming kernels (Ar<T>, tentative_cell, rhs_cell, correct_cell,
// jacobi_row_yn: operand order of SerialNavierStokes.cpp:140-146, :298-350), so the result is
// bit-identical to them and to the serial reference. Boundary rules (SerialNavierStokes.cpp:152-189):
//   * y edges (columns 0 and Y-1) are patch cells like any other and are refreshed after every sweep
//     from their freshly computed inward neighbour (or the constant), YEdgeUniform;
//   * x walls (rows 0 and X-1) are not stored during the sweeps: the row next to a wall substitutes
//     the wall value (its own previous value, or the constant) for the missing neighbour -- except in
//     the very first sweep of a launch, which reads the wall row as it was uploaded;
//   * wall rows and corners (one sweep stale) are written once, at the end of the launch.
// Supported: no interior flags; pressure edges UNSET or CONST; velocity edges all CONST (the cavity).
// Everything else takes the streaming path (fluidb200_solve decides).
#pragma once
#include "warp_kernels.cuh"

#ifndef FLUIDB200_HOST_EMU
namespace fb {

FB_DEV unsigned ld_acquire_gpu(const unsigned* p) {
    unsigned v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
FB_DEV void red_release_gpu(unsigned* p, unsigned v) {
    asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
// loads that must see what OTHER SMs wrote during this launch: L2, never this SM's L1
template <typename T, int V> FB_DEV Vec<T, V> ld_cg(const T* p) {
    const int4 r = __ldcg(reinterpret_cast<const int4*>(p));
    Vec<T, V> v;
    *reinterpret_cast<int4*>(&v) = r;
    return v;
}
template <typename T> FB_DEV T ld_cg1(const T* p) { return __ldcg(p); }

// All CTAs of the (cooperative, hence co-resident) launch meet here. `epoch` counts the barriers
// this thread has passed; the counter only ever grows.
FB_DEV void resident_grid_barrier(unsigned* bar, unsigned& epoch, unsigned nctas) {
    __syncthreads();
    ++epoch;
    if (nctas > 1 && threadIdx.x == 0) {
        __threadfence();
        red_release_gpu(bar, 1u);
        while (ld_acquire_gpu(bar) < epoch * nctas) {}
        __threadfence();
    }
    __syncthreads();
}

template <int M> struct ModeTag { static constexpr int value = M; };

// N consecutive elements (N * sizeof(T) is 8 or a multiple of 16 bytes, the address aligned likewise) from
// shared memory, or -- CG -- from L2 past this SM's L1.
template <typename T, int N, bool CG> FB_DEV void load_run(const T* src, T (&out)[N]) {
    if constexpr ((N * sizeof(T)) % 16 == 0) {
#pragma unroll
        for (int k = 0; k < (int)(N * sizeof(T) / 16); ++k) {
            const int4 q = CG ? __ldcg(reinterpret_cast<const int4*>(src) + k) : reinterpret_cast<const int4*>(src)[k];
            *reinterpret_cast<int4*>(&out[k * (16 / sizeof(T))]) = q;
        }
    } else {
        static_assert(N * sizeof(T) == 8, "runs are 8 bytes or whole 16-byte vectors");
        const int2 q = CG ? __ldcg(reinterpret_cast<const int2*>(src)) : *reinterpret_cast<const int2*>(src);
        *reinterpret_cast<int2*>(&out[0]) = q;
    }
}
template <typename T, int N> FB_DEV void store_run(T* dst, const T (&v)[N]) {
    if constexpr ((N * sizeof(T)) % 16 == 0) {
#pragma unroll
        for (int k = 0; k < (int)(N * sizeof(T) / 16); ++k)
            reinterpret_cast<int4*>(dst)[k] = *reinterpret_cast<const int4*>(&v[k * (16 / sizeof(T))]);
    } else {
        *reinterpret_cast<int2*>(dst) = *reinterpret_cast<const int2*>(&v[0]);
    }
}

template <typename T, int V, int R>
FB_DEV void resident_body(const ResidentArgs<T>& A) {
    extern __shared__ __align__(16) unsigned char fb_res_smem[];
    using A_ = Ar<T>;
    using VT = Vec<T, V>;
    constexpr int W = 32 * V;
    constexpr int RH = kResidentRowHalo, CH = kResidentColHalo;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int wy = A.wy, wx = A.wx;
    const int nwarps = wx * wy, nthreads = nwarps * 32;
    const int wyi = warp % wy, wxi = warp / wy;
    const int tile = blockIdx.x, tiles_y = A.tiles_y, tiles_x = A.tiles_x, ntiles = tiles_x * tiles_y;
    const int tyi = tile % tiles_y, txi = tile / tiles_y;
    const int X = A.X, Y = A.Y;
    const long long pitch = A.pitch;
    const int TR = wx * R, TC = wy * W;
    const int tr0 = txi * TR, tc0 = tyi * TC;   // first row / column of the tile
    const int x0 = tr0 + wxi * R;               // first row of this thread's patch
    const int wc0 = tc0 + wyi * W;              // first column of this warp
    const int c0 = wc0 + lane * V;              // first column of this thread's patch
    const bool col_ok = (long long)c0 + V <= pitch;   // the lane's 16 bytes exist in memory
    const bool l0 = lane == 0, l31 = lane == 31;

    // shared memory: four frames (tile + halo) for the predictor / corrector, then the exchange buffers of the sweeps
    const int FR = TR + 2 * RH, FP = TC + 2 * CH;
    T* su = reinterpret_cast<T*>(fb_res_smem);
    T* sv = su + FR * FP;
    T* sus = sv + FR * FP;      // u_next
    T* svs = sus + FR * FP;     // v_next
    T* rowbuf = svs + FR * FP;                  // [parity][warp][first / last row of the patch][W]
    T* colbuf = rowbuf + 2 * nwarps * 2 * W;    // [parity][warp][first / last column of the warp][R]
    const int rowpar = nwarps * 2 * W, colpar = nwarps * 2 * R;   // elements between the two parities
    const int fbase = (x0 - tr0 + RH) * FP + (c0 - tc0 + CH);   // frame index of patch cell (row 0, element 0)

    T* const U = A.f[0];
    T* const Vf = A.f[1];

    // ---- who this thread exchanges with (all of it fixed for the launch) -------------------------------
    // kinds: 0 nothing there (box wall), 1 another warp of this CTA (shared memory), 2 another tile (L2 band)
    const int up_kind = wxi > 0 ? 1 : (txi > 0 ? 2 : 0);
    const int dn_kind = wxi < wx - 1 ? 1 : (txi < tiles_x - 1 ? 2 : 0);
    const int lf_kind = wyi > 0 ? 1 : (tyi > 0 ? 2 : 0);
    const int rt_kind = wyi < wy - 1 ? 1 : (tyi < tiles_y - 1 ? 2 : 0);
    // (kept as 32-bit offsets: the pointers would cost two registers each)
    const int pub_row = (warp * 2) * W + lane * V;             // into rowbuf
    const int pub_col = (warp * 2) * R;                         // into colbuf
    const int up_s = ((warp - wy) * 2 + 1) * W + lane * V;
    const int dn_s = ((warp + wy) * 2) * W + lane * V;
    const int lf_s = ((warp - 1) * 2 + 1) * R;
    const int rt_s = ((warp + 1) * 2) * R;
    // Between tiles a border travels as 8-byte PACKETS {32 data bits, sequence number}: an aligned 8-byte
    // store is seen whole or not at all, so the reader spins on the data itself until every packet carries
    // the number of the publication it waits for -- no flag, no fence, one trip through L2. A border of the
    // tile is `band_words` packets: [parity][tile][top row, bottom row, left column, right column].
    constexpr int WPE = sizeof(T) / 4;                          // 32-bit words per element
    constexpr int ROWW = V * WPE, COLW = R * WPE;               // packets of a lane's row piece / of a warp's column piece
    const int LW = A.band_len * WPE;
    const int bandpar = ntiles * 4 * LW;                        // packets between the two parities
    const int roffw = (wyi * W + lane * V) * WPE, coffw = wxi * R * WPE;
    auto band_off = [&](int t, int side, int o) { return (t * 4 + side) * LW + o; };   // (packets; < 2^31)
    const int up_g = up_kind == 2 ? band_off(tile - tiles_y, 1, roffw) : 0;
    const int dn_g = dn_kind == 2 ? band_off(tile + tiles_y, 0, roffw) : 0;
    const int lf_g = lf_kind == 2 ? band_off(tile - 1, 3, coffw) : 0;
    const int rt_g = rt_kind == 2 ? band_off(tile + 1, 2, coffw) : 0;
    const int gb_top = (wxi == 0 && txi > 0) ? band_off(tile, 0, roffw) : -1;
    const int gb_bot = (wxi == wx - 1 && txi < tiles_x - 1) ? band_off(tile, 1, roffw) : -1;
    const int gb_lf = (wyi == 0 && tyi > 0) ? band_off(tile, 2, coffw) : -1;
    const int gb_rt = (wyi == wy - 1 && tyi < tiles_y - 1) ? band_off(tile, 3, coffw) : -1;
    unsigned long long* const band = reinterpret_cast<unsigned long long*>(A.band);

    // ---- boundary rules of this thread (SerialNavierStokes.cpp:152-189) -----------------------------
    const bool lconst = A.pmode[LEFT] == EDGE_CONST, rconst = A.pmode[RIGHT] == EDGE_CONST;
    const bool bot_use = A.pmode[BOTTOM] == EDGE_CONST, top_use = A.pmode[TOP] == EDGE_CONST;
    const T bot_v = A.pval[BOTTOM], top_v = A.pval[TOP];
    // Column Y-1 can be the FIRST column of a warp (Y = 1 + m * 32V: 129, 257, 1025 ...): its inward
    // neighbour Y-2 then belongs to another warp, so the column cannot be refreshed in the same sweep.
    // It is not kept during the sweeps: column Y-2 substitutes the value it would hold, and it is written
    // at the end. Likewise row X-1 when it is the first row of a patch (`wall_below`).
    const bool top_detached = ((Y - 1) % W) == 0;
    const bool warp_top_det = top_detached && (wc0 + W - 1 == Y - 2);   // this warp's last column is Y-2
    const bool has_bottom = wc0 == 0;
    const bool has_top = !top_detached && wc0 <= Y - 1 && Y - 1 < wc0 + W;
    const int eT = (Y - 1) % V;                           // element that holds column Y-1 (in the lane that has it)
    const bool top_lane = has_top && (c0 + eT == Y - 1);
    const bool bconst = has_bottom && l0 && bot_use, bcopy = has_bottom && l0 && !bot_use;
    const bool tconst = top_lane && top_use, tcopy = top_lane && !top_use;
    const int rX1 = X - 1 - x0;                            // index of row X-1 in this patch (if in [1, R-1]: kept here)
    const bool wall_below = x0 + R - 1 == X - 2;           // row X-1 is the first row of the patch below
    // warps that have anything to fix up around the plain sweep
    const bool general = has_bottom || has_top || warp_top_det || x0 == 0 || (rX1 >= 1 && rX1 <= R - 1) || wall_below;

    // Values the sweep loop uses over and over. Made opaque so that the compiler keeps them in registers
    // instead of recomputing each from the kernel arguments in every sweep (which it did: ~150
    // instructions per sweep).
    const int rb0 = (int)(rowbuf - su), cb0 = (int)(colbuf - su);   // (offsets below count from the start of shared memory)
    int o_general = general, o_pub_row = rb0 + pub_row, o_pub_col = cb0 + pub_col, o_up_s = rb0 + up_s, o_dn_s = rb0 + dn_s,
        o_lf_s = cb0 + lf_s, o_rt_s = cb0 + rt_s, o_up_k = up_kind, o_dn_k = dn_kind, o_lf_k = lf_kind, o_rt_k = rt_kind, o_up_g = up_g, o_dn_g = dn_g,
        o_lf_g = lf_g, o_rt_g = rt_g, o_gb_top = gb_top, o_gb_bot = gb_bot, o_gb_lf = gb_lf, o_gb_rt = gb_rt,
        o_rowpar = rowpar, o_colpar = colpar, o_bandpar = bandpar;
    asm volatile("" : "+r"(o_general), "+r"(o_pub_row), "+r"(o_pub_col), "+r"(o_up_s), "+r"(o_dn_s), "+r"(o_lf_s), "+r"(o_rt_s));
    asm volatile("" : "+r"(o_up_k), "+r"(o_dn_k), "+r"(o_lf_k), "+r"(o_rt_k), "+r"(o_up_g), "+r"(o_dn_g), "+r"(o_lf_g), "+r"(o_rt_g));
    asm volatile("" : "+r"(o_gb_top), "+r"(o_gb_bot), "+r"(o_gb_lf), "+r"(o_gb_rt), "+r"(o_rowpar), "+r"(o_colpar), "+r"(o_bandpar));

    VT p[R], b[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const int x = x0 + r;
        p[r] = (x < X && col_ok) ? ld_cg<T, V>(A.f[2] + (long long)x * pitch + c0) : vzero<T, V>();
        b[r] = vzero<T, V>();
    }

    // FLUIDB200_RESIDENT_PROF=1 at run time: cycles per phase of a timestep. The split of a SWEEP (five more ticks per
    // sweep, ~4 % of a 1000^2 step even when switched off) is compiled in only with -DFLUIDB200_RESIDENT_SUBPROF=1.
#ifndef FLUIDB200_RESIDENT_SUBPROF
#define FLUIDB200_RESIDENT_SUBPROF 0
#endif
    long long t_prev = 0;
    const bool profiled = A.prof && tid == 0 && tile == ntiles / 2;
    __shared__ unsigned long long prof_s[16];   // (added up in shared memory with fire-and-forget reductions: a tick must not stall the warp)
    if (profiled) {
        for (int k = 0; k < 16; ++k) prof_s[k] = 0;
    }
    auto tick = [&](int k) {   // (only when asked for: one thread of a middle tile adds up its cycles per phase)
        if (k >= 8 && !FLUIDB200_RESIDENT_SUBPROF) return;
        if (profiled) {
            const long long t = clock64();
            atomicAdd(&prof_s[k], (unsigned long long)(t - t_prev));
            t_prev = t;
        }
    };
    unsigned polls = 0;

    // ---- exchange of patch borders ----------------------------------------------------------
    // NW 32-bit words -> NW packets carrying sequence number `seq`, stored 16 bytes (two packets) at a time
    auto put_packets = [&](unsigned long long* dst, const unsigned* w, auto nw_tag, unsigned seq) {
        constexpr int NW = decltype(nw_tag)::value;
        static_assert(NW % 2 == 0, "packets are stored in pairs");
#pragma unroll
        for (int k = 0; k < NW; k += 2) {
            const unsigned long long a = (unsigned long long)w[k] | ((unsigned long long)seq << 32);
            const unsigned long long c = (unsigned long long)w[k + 1] | ((unsigned long long)seq << 32);
            asm volatile("st.volatile.global.v2.u64 [%0], {%1, %2};" ::"l"(dst + k), "l"(a), "l"(c) : "memory");
        }
    };
    auto get_packets = [&](const unsigned long long* src, unsigned* w, auto nw_tag, unsigned seq) {
        constexpr int NW = decltype(nw_tag)::value;
        bool ok;
        do {
            ok = true;
            if (FLUIDB200_RESIDENT_SUBPROF) ++polls;
#pragma unroll
            for (int k = 0; k < NW; k += 2) {
                unsigned long long a, c;
                asm volatile("ld.volatile.global.v2.u64 {%0, %1}, [%2];" : "=l"(a), "=l"(c) : "l"(src + k) : "memory");
                w[k] = (unsigned)a;
                w[k + 1] = (unsigned)c;
                ok = ok && (unsigned)(a >> 32) == seq && (unsigned)(c >> 32) == seq;
            }
        } while (!ok);
    };
    // publication number `seq` (>= 1) of this patch's borders, into parity `par`: first and last row ...
    auto publish_rows = [&](int par, unsigned seq, const VT& first, const VT& last) {
        T* rb = su + (o_pub_row + par * o_rowpar);
        st_vec<T, V>(rb, first);
        st_vec<T, V>(rb + W, last);
        if (o_gb_top >= 0) put_packets(band + (o_gb_top + par * o_bandpar), reinterpret_cast<const unsigned*>(&first), ModeTag<ROWW>{}, seq);
        if (o_gb_bot >= 0) put_packets(band + (o_gb_bot + par * o_bandpar), reinterpret_cast<const unsigned*>(&last), ModeTag<ROWW>{}, seq);
    };
    // ... and first and last column of the warp
    auto publish_cols = [&](int par, unsigned seq) {
        if (l0 || l31) {
            T edge[R];
#pragma unroll
            for (int r = 0; r < R; ++r) edge[r] = l0 ? p[r].a[0] : p[r].a[V - 1];
            store_run<T, R>(su + (o_pub_col + par * o_colpar + (l0 ? 0 : R)), edge);
            const int g_off = l0 ? o_gb_lf : o_gb_rt;
            if (g_off >= 0) put_packets(band + (g_off + par * o_bandpar), reinterpret_cast<const unsigned*>(edge), ModeTag<COLW>{}, seq);
        }
    };
    // the four sides of the patch as published (number `seq`) with parity `par`
    auto fetch = [&](int par, unsigned seq, VT& up, VT& dn, T (&lf)[R], T (&rt)[R]) {
        up = vzero<T, V>();
        dn = vzero<T, V>();
#pragma unroll
        for (int r = 0; r < R; ++r) {
            lf[r] = T(0);
            rt[r] = T(0);
        }
        const bool need_up = o_up_k == 2, need_dn = o_dn_k == 2;
        if (need_up && need_dn) {
            // both rows come from other tiles: ask for both at once and spin on both (one trip through L2, not two)
            const unsigned long long* sa = band + o_up_g + par * o_bandpar;
            const unsigned long long* sb = band + o_dn_g + par * o_bandpar;
            unsigned* wa = reinterpret_cast<unsigned*>(&up);
            unsigned* wb = reinterpret_cast<unsigned*>(&dn);
            bool ok;
            do {
                ok = true;
                if (FLUIDB200_RESIDENT_SUBPROF) ++polls;
#pragma unroll
                for (int k = 0; k < ROWW; k += 2) {
                    unsigned long long a, c, d, e;
                    asm volatile("ld.volatile.global.v2.u64 {%0, %1}, [%2];" : "=l"(a), "=l"(c) : "l"(sa + k) : "memory");
                    asm volatile("ld.volatile.global.v2.u64 {%0, %1}, [%2];" : "=l"(d), "=l"(e) : "l"(sb + k) : "memory");
                    wa[k] = (unsigned)a; wa[k + 1] = (unsigned)c;
                    wb[k] = (unsigned)d; wb[k + 1] = (unsigned)e;
                    ok = ok && (unsigned)(a >> 32) == seq && (unsigned)(c >> 32) == seq && (unsigned)(d >> 32) == seq && (unsigned)(e >> 32) == seq;
                }
            } while (!ok);
        } else {
            if (o_up_k == 1) up = *reinterpret_cast<const VT*>(su + o_up_s + par * o_rowpar);
            else if (need_up) get_packets(band + o_up_g + par * o_bandpar, reinterpret_cast<unsigned*>(&up), ModeTag<ROWW>{}, seq);
            if (o_dn_k == 1) dn = *reinterpret_cast<const VT*>(su + o_dn_s + par * o_rowpar);
            else if (need_dn) get_packets(band + o_dn_g + par * o_bandpar, reinterpret_cast<unsigned*>(&dn), ModeTag<ROWW>{}, seq);
        }
        // (every lane reads the same run: one broadcast access; only lane 0 / 31 use it)
        if (o_lf_k == 1) load_run<T, R, false>(su + o_lf_s + par * o_colpar, lf);
        else if (o_lf_k == 2) get_packets(band + o_lf_g + par * o_bandpar, reinterpret_cast<unsigned*>(lf), ModeTag<COLW>{}, seq);
        if (o_rt_k == 1) load_run<T, R, false>(su + o_rt_s + par * o_colpar, rt);
        else if (o_rt_k == 2) get_packets(band + o_rt_g + par * o_bandpar, reinterpret_cast<unsigned*>(rt), ModeTag<COLW>{}, seq);
    };
    auto splat = [](T v) {
        VT r;
#pragma unroll
        for (int e = 0; e < V; ++e) r.a[e] = v;
        return r;
    };
    // What a `general` warp adds around the plain arithmetic. BEFORE: neighbours that are not kept
    // (row X-1 below this patch, column Y-1 beside this warp) take the value they would hold -- the
    // constant, or a copy of this patch's own border as of the previous sweep; the first sweep of a
    // launch (`subst` false) reads them as they were uploaded instead.
    auto fix_neighbours = [&](bool subst, VT& dn, T (&rt)[R]) {
        if (!subst) return;
        if (wall_below) dn = rconst ? splat(A.pval[RIGHT]) : p[R - 1];                      // :183-187
        if (warp_top_det) {
#pragma unroll
            for (int r = 0; r < R; ++r) rt[r] = top_use ? top_v : p[r].a[V - 1];            // :170-174
        }
    };
    // AFTER: the boundary cells this patch does keep are refreshed from their freshly computed inward
    // neighbours (or the constant): wall rows 0 and X-1 (:177-188), columns 0 and Y-1 (:163-175). Corners
    // end up with placeholders; they feed no stencil and are written at the end of the launch.
    auto fix_boundary_cells = [&]() {
        if (x0 == 0) p[0] = lconst ? splat(A.pval[LEFT]) : p[1];
        if (rX1 >= 1 && rX1 <= R - 1) {
#pragma unroll
            for (int r = 1; r < R; ++r)
                if (r == rX1) p[r] = rconst ? splat(A.pval[RIGHT]) : p[r - 1];
        }
        if (has_bottom) {
#pragma unroll
            for (int r = 0; r < R; ++r) p[r].a[0] = bcopy ? p[r].a[1] : (bconst ? bot_v : p[r].a[0]);
        }
        if (has_top) {   // (warp-uniform, like eT)
            if (eT == 0) {
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    const T below = shfl_up1(p[r].a[V - 1]);
                    p[r].a[0] = tcopy ? below : (tconst ? top_v : p[r].a[0]);
                }
            }
#pragma unroll
            for (int e = 1; e < V; ++e) {
                if (eT == e) {
#pragma unroll
                    for (int r = 0; r < R; ++r) p[r].a[e] = tcopy ? p[r].a[e - 1] : (tconst ? top_v : p[r].a[e]);
                }
            }
        }
    };

    // ---- one Jacobi sweep: p_g from p_{g-1} (SerialNavierStokes.cpp:140-198) ---------------------
    // Every row of the patch is computed, wall rows and rows beyond the box included: what must not
    // change is put right by fix_boundary_cells, what lies outside is never stored and feeds nothing.
    // The first and the last row are computed FIRST and published at once: they are what the patches
    // above and below (other warps, other tiles) wait for, and the rows between are computed while
    // those travel.
    auto one_row = [&](const VT& xm, const VT& cur, const VT& xp, T lfv, T rtv, const VT& bb) {
        VT ym, yp;
        y_neighbours<T, V>(cur, ym, yp);
        ym.a[0] = l0 ? lfv : ym.a[0];
        yp.a[V - 1] = l31 ? rtv : yp.a[V - 1];
        return jacobi_row_yn<T, V>(xm, xp, ym, yp, bb);
    };
    auto sweep = [&](int g) {
        static_assert(R >= 2, "a patch holds its wall row and the row next to it");
        VT up, dn;
        T lf[R], rt[R];
        fetch((g + 1) & 1, (unsigned)(g + 1), up, dn, lf, rt);
        tick(8);
        if (o_general) fix_neighbours(g > 0, dn, rt);
        const VT old0 = p[0], oldL = p[R - 1];
        const VT new0 = one_row(up, old0, p[1], lf[0], rt[0], b[0]);
        const VT newL = one_row(p[R - 2], oldL, dn, lf[R - 1], rt[R - 1], b[R - 1]);
        publish_rows(g & 1, (unsigned)(g + 2), new0, newL);
        tick(9);
        VT prev = old0;
#pragma unroll
        for (int r = 1; r < R - 1; ++r) {
            const VT cur = p[r];
            const VT nxt = (r < R - 2) ? p[r < R - 2 ? r + 1 : r] : oldL;
            p[r] = one_row(prev, cur, nxt, lf[r], rt[r], b[r]);
            prev = cur;
        }
        p[0] = new0;
        p[R - 1] = newL;
        tick(10);
        if (o_general) fix_boundary_cells();
        publish_cols(g & 1, (unsigned)(g + 2));
        tick(11);
    };

    // ---- predictor: u_next, v_next on the tile + 1 ring, rhs on the patch (SerialNavierStokes.cpp:298-335) ----
    // u and v of the tile live in the frames su / sv for the whole launch (the corrector writes them there);
    // only the two-cell ring around the tile comes from the other tiles, through L2, once per step.
    auto load_ring = [&]() {
        const int nj = TC + 4, n1 = 2 * RH * nj, n = n1 + TR * 4;
        for (int idx = tid; idx < n; idx += nthreads) {
            int i, j;
            if (idx < n1) {
                const int q = idx / nj;
                i = q < RH ? q : RH + TR + (q - RH);
                j = CH - 2 + (idx - q * nj);
            } else {
                const int k = idx - n1, q = k >> 2, m = k & 3;
                i = RH + q;
                j = m < 2 ? CH - 2 + m : CH + TC + (m - 2);
            }
            const int gx = tr0 - RH + i, gy = tc0 - CH + j;
            if (gx >= 0 && gx < X && gy >= 0 && gy < Y) {
                const long long o = (long long)gx * pitch + gy;
                su[i * FP + j] = ld_cg1<T>(U + o);
                sv[i * FP + j] = ld_cg1<T>(Vf + o);
            }
        }
    };
    auto tentative_scalar = [&](auto mode_tag, int c) {   // one frame cell (an interior cell of the box)
        constexpr int MODE = decltype(mode_tag)::value;
        Flags fl;
        UV<T> r = tentative_cell<T, MODE>(A.pa, su[c], sv[c], su[c - FP], su[c + FP], su[c - 1], su[c + 1], sv[c - FP],
                                          sv[c + FP], sv[c - 1], sv[c + 1], fl);
        if (MODE != DIV_EXACT && fl.raised(A.pa.m.pow2_lo_m1))
            r = tentative_cell_exact<T, MODE>(A.pa, su[c], sv[c], su[c - FP], su[c + FP], su[c - 1], su[c + 1], sv[c - FP],
                                        sv[c + FP], sv[c - 1], sv[c + 1]);
        sus[c] = r.u;
        svs[c] = r.v;
    };
    auto tentative = [&](auto mode_tag) {
        // the patch
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const int x = x0 + r;
            if (x < 1 || x > X - 2) continue;
#pragma unroll
            for (int e = 0; e < V; ++e) {
                const int col = c0 + e;
                if (col >= 1 && col <= Y - 2) tentative_scalar(mode_tag, fbase + r * FP + e);
            }
        }
        // the one-cell ring around the tile (its cells belong to other tiles, which compute them too)
        const int nj = TC + 2, n1 = 2 * nj, n = n1 + 2 * TR;
        for (int idx = tid; idx < n; idx += nthreads) {
            int i, j;
            if (idx < n1) {
                i = idx < nj ? RH - 1 : RH + TR;
                j = CH - 1 + (idx < nj ? idx : idx - nj);
            } else {
                const int k = idx - n1;
                i = RH + (k >> 1);
                j = (k & 1) ? CH + TC : CH - 1;
            }
            const int gx = tr0 - RH + i, gy = tc0 - CH + j;
            if (gx >= 1 && gx <= X - 2 && gy >= 1 && gy <= Y - 2) tentative_scalar(mode_tag, i * FP + j);
        }
    };
    auto rhs_patch = [&](auto mode_tag, bool last_step) {
        constexpr int MODE = decltype(mode_tag)::value;
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const int x = x0 + r;
            if (x < 1 || x > X - 2) continue;
            const int c = fbase + r * FP;
            VT q = vzero<T, V>();
#pragma unroll
            for (int e = 0; e < V; ++e) {
                const int col = c0 + e;
                if (col >= 1 && col <= Y - 2) {
                    Flags fl;
                    T v = rhs_cell<T, MODE>(A.pa, sus[c + e - FP], sus[c + e + FP], svs[c + e - 1], svs[c + e + 1], fl);
                    if (MODE != DIV_EXACT && fl.raised(A.pa.m.pow2_lo_m1))
                        v = rhs_cell_exact<T, MODE>(A.pa, sus[c + e - FP], sus[c + e + FP], svs[c + e - 1], svs[c + e + 1]);
                    q.a[e] = v;
                }
                b[r].a[e] = A_::mul(A_::mul(q.a[e], A.hx), A.hy);   // (rhs * hx) * hy, :144, once per step
            }
            if (last_step && col_ok) st_interior<T, V>(A.f[6] + (long long)x * pitch, c0, Y, q);
        }
    };

    // ---- corrector: u, v from the final p of the step, velocity edges (SerialNavierStokes.cpp:338-350, :222-295) ----
    auto corrector = [&](auto mode_tag, int g_done) {
        constexpr int MODE = decltype(mode_tag)::value;
        VT up, dn;
        T lf[R], rt[R];
        fetch((g_done + 1) & 1, (unsigned)(g_done + 1), up, dn, lf, rt);
        if (o_general) fix_neighbours(g_done > 0, dn, rt);
        VT prev = up;
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const int x = x0 + r;
            const VT cur = p[r];
            const VT nxt = (r < R - 1) ? p[r < R - 1 ? r + 1 : r] : dn;
            if (x < X) {   // (warp-uniform)
                VT un = vzero<T, V>(), vn = vzero<T, V>();
                if (x >= 1 && x <= X - 2) {
                    const VT pm = prev, pp = nxt;   // (wall rows are kept up to date in the patch, or were fixed above)
                    VT pym, pyp;
                    y_neighbours<T, V>(cur, pym, pyp);
                    pym.a[0] = l0 ? lf[r] : pym.a[0];
                    pyp.a[V - 1] = l31 ? rt[r] : pyp.a[V - 1];
                    const int c = fbase + r * FP;
#pragma unroll
                    for (int e = 0; e < V; ++e) {
                        const int col = c0 + e;
                        if (col >= 1 && col <= Y - 2) {
                            Flags fl;
                            UV<T> q = correct_cell<T, MODE>(A.ca, pm.a[e], pp.a[e], pym.a[e], pyp.a[e], sus[c + e], svs[c + e], fl);
                            if (MODE != DIV_EXACT && fl.raised(A.ca.m.pow2_lo_m1))
                                q = correct_cell_exact<T, MODE>(A.ca, pm.a[e], pp.a[e], pym.a[e], pyp.a[e], sus[c + e], svs[c + e]);
                            un.a[e] = q.u;
                            vn.a[e] = q.v;
                        } else if (col == 0) {          // the y edges own their cells for ALL x (:236-264) ...
                            un.a[e] = A.uval[BOTTOM]; vn.a[e] = A.vval[BOTTOM];
                        } else if (col == Y - 1) {
                            un.a[e] = A.uval[TOP]; vn.a[e] = A.vval[TOP];
                        }
                    }
                } else {                                // ... the x walls the cells between (:266-294)
                    const int edge = (x == 0) ? LEFT : RIGHT;
#pragma unroll
                    for (int e = 0; e < V; ++e) {
                        const int col = c0 + e;
                        if (col == 0) { un.a[e] = A.uval[BOTTOM]; vn.a[e] = A.vval[BOTTOM]; }
                        else if (col == Y - 1) { un.a[e] = A.uval[TOP]; vn.a[e] = A.vval[TOP]; }
                        else if (col < Y - 1) { un.a[e] = A.uval[edge]; vn.a[e] = A.vval[edge]; }
                    }
                }
                st_vec<T, V>(su + fbase + r * FP, un);   // next step's predictor reads them here
                st_vec<T, V>(sv + fbase + r * FP, vn);
                if (col_ok) {
                    st_vec<T, V>(U + (long long)x * pitch + c0, un);
                    st_vec<T, V>(Vf + (long long)x * pitch + c0, vn);
                }
            }
            prev = cur;
        }
    };
    using M0 = ModeTag<DIV_EXACT>;
    using M1 = ModeTag<DIV_FAST>;
    using M2 = ModeTag<DIV_POW2>;
    const int mode = A.pa.m.mode;   // (grid-uniform)

    // ---- the launch -------------------------------------------------------------------------------
    // frames: this patch's u, v; the stored u_next / v_next of the box edges (never written, read by the rhs)
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const int x = x0 + r;
        const bool in = x < X && col_ok;
        st_vec<T, V>(su + fbase + r * FP, in ? ld_cg<T, V>(U + (long long)x * pitch + c0) : vzero<T, V>());
        st_vec<T, V>(sv + fbase + r * FP, in ? ld_cg<T, V>(Vf + (long long)x * pitch + c0) : vzero<T, V>());
    }
    for (int idx = tid; idx < FR * FP; idx += nthreads) {
        const int i = idx / FP, j = idx - i * FP;
        const int gx = tr0 - RH + i, gy = tc0 - CH + j;
        T us = T(0), vs = T(0);
        if (gx >= 0 && gx < X && gy >= 0 && gy < Y && (gx == 0 || gx == X - 1 || gy == 0 || gy == Y - 1)) {
            us = A.f[3][(long long)gx * pitch + gy];
            vs = A.f[4][(long long)gx * pitch + gy];
        }
        sus[idx] = us;
        svs[idx] = vs;
    }
    publish_rows(1, 1u, p[0], p[R - 1]);   // the uploaded p, for the first sweep's fetch
    publish_cols(1, 1u);
    __syncthreads();
    unsigned epoch = 0;
    t_prev = clock64();
    const int G = A.S * A.K;
    int g = 0;
    for (int step = 0; step < A.S; ++step) {
        if (step > 0) resident_grid_barrier(A.sync, epoch, (unsigned)ntiles);   // every tile's new u, v are in L2
        load_ring();
        __syncthreads();
        tick(0);
        if (mode == 2) tentative(M2{}); else if (mode == 1) tentative(M1{}); else tentative(M0{});
        __syncthreads();
        tick(1);
        const bool last = step == A.S - 1;
        if (mode == 2) rhs_patch(M2{}, last); else if (mode == 1) rhs_patch(M1{}, last); else rhs_patch(M0{}, last);
        tick(2);
        for (int s = 0; s < A.K; ++s, ++g) {
            if (g == G - 1) {   // the values an unset corner copies are one sweep old
#pragma unroll
                for (int r = 0; r < R; ++r) {
#pragma unroll
                    for (int e = 0; e < V; ++e) {
                        const int x = x0 + r, col = c0 + e;
                        if (x == 1 && col == 1) A.stash[0] = p[r].a[e];
                        if (x == 1 && col == Y - 2) A.stash[1] = p[r].a[e];
                        if (x == X - 2 && col == 1) A.stash[2] = p[r].a[e];
                        if (x == X - 2 && col == Y - 2) A.stash[3] = p[r].a[e];
                    }
                }
            }
            sweep(g);
            __syncthreads();
            tick(12);
        }
        // a tile may overwrite u, v only when no other tile still reads them for ITS predictor: three
        // sweeps of neighbour hand-shakes order that even for the diagonal tiles; fewer do not
        tick(3);
        if (A.K < 3) resident_grid_barrier(A.sync, epoch, (unsigned)ntiles);
        if (mode == 2) corrector(M2{}, g); else if (mode == 1) corrector(M1{}, g); else corrector(M0{}, g);
        tick(4);
    }

    // ---- write the state back: p with its wall rows and corners, u_next, v_next ----------------------
    resident_grid_barrier(A.sync, epoch, (unsigned)ntiles);   // nobody reads a published border any more
    T* const P = A.f[2];
    // (the wall rows and y-edge columns this patch keeps are final already; corners are fixed below)
    const bool deferred = x0 == X - 1;   // row X-1 is the first row of this patch: its inward neighbour is another thread's
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const int x = x0 + r;
        if (x >= X) continue;
        if (col_ok && !(deferred && r == 0)) st_vec<T, V>(P + (long long)x * pitch + c0, p[r]);
        if (x >= 1 && x <= X - 2 && col_ok) {
            const int c = fbase + r * FP;
            st_interior<T, V>(A.f[3] + (long long)x * pitch, c0, Y, *reinterpret_cast<const VT*>(sus + c));
            st_interior<T, V>(A.f[4] + (long long)x * pitch, c0, Y, *reinterpret_cast<const VT*>(svs + c));
        }
    }
    resident_grid_barrier(A.sync, epoch, (unsigned)ntiles);
    if (deferred && col_ok) {
        const VT inner = ld_cg<T, V>(P + (long long)(X - 2) * pitch + c0);
        VT out;
#pragma unroll
        for (int e = 0; e < V; ++e) out.a[e] = rconst ? A.pval[RIGHT] : inner.a[e];
        st_interior<T, V>(P + (long long)(X - 1) * pitch, c0, Y, out);
    }
    if (top_detached && c0 == Y - 1) {   // the detached column Y-1, rows 1 .. X-2 (:163-175)
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const int x = x0 + r;
            if (x >= 1 && x <= X - 2)
                P[(long long)x * pitch + Y - 1] = top_use ? top_v : ld_cg1<T>(P + (long long)x * pitch + Y - 2);
        }
    }
    if (tile == 0 && tid == 0) {   // corners: the y-edge loop runs first (:163-175), so an unset corner copies the
                                   // wall cell next to it as it was BEFORE this sweep's x-edge loop refreshed it
        const T c00 = lconst ? A.pval[LEFT] : ld_cg1<T>(A.stash + 0), c01 = lconst ? A.pval[LEFT] : ld_cg1<T>(A.stash + 1);
        const T c10 = rconst ? A.pval[RIGHT] : ld_cg1<T>(A.stash + 2), c11 = rconst ? A.pval[RIGHT] : ld_cg1<T>(A.stash + 3);
        P[0] = bot_use ? bot_v : c00;
        P[Y - 1] = top_use ? top_v : c01;
        P[(long long)(X - 1) * pitch] = bot_use ? bot_v : c10;
        P[(long long)(X - 1) * pitch + Y - 1] = top_use ? top_v : c11;
    }
    tick(5);
    if (profiled) {
        for (int k = 0; k < 13; ++k) A.prof[k] += prof_s[k];
        A.prof[13] += polls;
    }
}


// =========================================================================================
// Grids that fit ONE CTA outright (up to 4096 cells: 41^2, 64^2): one cell per thread and slot, every
// field in shared memory, one __syncthreads per sweep. With so few columns a 16-byte-per-lane patch
// layout leaves two lanes in three idle; here every lane has a cell.
//   * p ping-pongs between two shared arrays (Jacobi, SerialNavierStokes.cpp:140-198);
//   * boundary cells are not updated during the sweeps: a cell next to a boundary reads, instead of
//     the boundary cell, what that cell would hold -- ITSELF (the copy rule, :163-188: the boundary
//     took this cell's value at the end of the previous sweep) or a slot holding the constant. That
//     is a redirected neighbour ADDRESS, decided once per cell: no test in the loop. The very first
//     sweep of a launch reads the true neighbours (the uploaded boundary values);
//   * boundary cells and corners (one sweep stale, SURVEY A5: the other array still holds the state
//     before the last sweep) are written once, at the end.
// =========================================================================================
constexpr int kCellThreads = 1024;

template <typename T, int NC>
FB_DEV void resident_cell_body(const ResidentArgs<T>& A) {
    extern __shared__ __align__(16) unsigned char fb_res_smem[];
    using A_ = Ar<T>;
    const int tid = threadIdx.x, NT = blockDim.x;   // (NC * NT >= cells: the planner's choice)
    const int X = A.X, Y = A.Y, n = X * Y;
    const long long pitch = A.pitch;
    const int NP = n + 4;                      // + one slot per edge for its constant
    T* pbuf0 = reinterpret_cast<T*>(fb_res_smem);
    T* pbuf1 = pbuf0 + NP;
    T* su = pbuf1 + NP;
    T* sv = su + n;
    T* sus = sv + n;
    T* svs = sus + n;

    // per slot: the cell, its four (redirected) neighbours, its scaled right-hand side
    int cell[NC], axm[NC], axp[NC], aym[NC], ayp[NC];
    bool interior[NC];
    T b[NC];
    const bool lconst = A.pmode[LEFT] == EDGE_CONST, rconst = A.pmode[RIGHT] == EDGE_CONST;
    const bool bconst = A.pmode[BOTTOM] == EDGE_CONST, tconst = A.pmode[TOP] == EDGE_CONST;
#pragma unroll
    for (int k = 0; k < NC; ++k) {
        const int c = tid + k * NT;
        const int x = c / Y, y = c - x * Y;
        cell[k] = c;
        interior[k] = c < n && x >= 1 && x <= X - 2 && y >= 1 && y <= Y - 2;
        axm[k] = (x == 1) ? (lconst ? n + LEFT : c) : c - Y;
        axp[k] = (x == X - 2) ? (rconst ? n + RIGHT : c) : c + Y;
        aym[k] = (y == 1) ? (bconst ? n + BOTTOM : c) : c - 1;
        ayp[k] = (y == Y - 2) ? (tconst ? n + TOP : c) : c + 1;
        b[k] = T(0);
    }
    // load the state
    for (int c = tid; c < n; c += NT) {
        const int x = c / Y, y = c - x * Y;
        const long long o = (long long)x * pitch + y;
        pbuf0[c] = A.f[2][o];
        pbuf1[c] = T(0);
        su[c] = A.f[0][o];
        sv[c] = A.f[1][o];
        const bool edge = x == 0 || x == X - 1 || y == 0 || y == Y - 1;
        sus[c] = edge ? A.f[3][o] : T(0);   // edge u_next / v_next: stored values, never written, read by the rhs
        svs[c] = edge ? A.f[4][o] : T(0);
    }
    if (tid < 4) {
        pbuf0[n + tid] = A.pval[tid];
        pbuf1[n + tid] = A.pval[tid];
    }
    __syncthreads();

    const int mode = A.pa.m.mode;
    auto tentative = [&](auto mode_tag) {
        constexpr int MODE = decltype(mode_tag)::value;
#pragma unroll
        for (int k = 0; k < NC; ++k) {
            if (!interior[k]) continue;
            const int c = cell[k];
            Flags fl;
            UV<T> r = tentative_cell<T, MODE>(A.pa, su[c], sv[c], su[c - Y], su[c + Y], su[c - 1], su[c + 1], sv[c - Y],
                                              sv[c + Y], sv[c - 1], sv[c + 1], fl);
            if (MODE != DIV_EXACT && fl.raised(A.pa.m.pow2_lo_m1))
                r = tentative_cell_exact<T, MODE>(A.pa, su[c], sv[c], su[c - Y], su[c + Y], su[c - 1], su[c + 1], sv[c - Y], sv[c + Y],
                                            sv[c - 1], sv[c + 1]);
            sus[c] = r.u;
            svs[c] = r.v;
        }
    };
    auto rhs = [&](auto mode_tag, bool last_step) {
        constexpr int MODE = decltype(mode_tag)::value;
#pragma unroll
        for (int k = 0; k < NC; ++k) {
            if (!interior[k]) continue;
            const int c = cell[k];
            Flags fl;
            T q = rhs_cell<T, MODE>(A.pa, sus[c - Y], sus[c + Y], svs[c - 1], svs[c + 1], fl);
            if (MODE != DIV_EXACT && fl.raised(A.pa.m.pow2_lo_m1)) q = rhs_cell_exact<T, MODE>(A.pa, sus[c - Y], sus[c + Y], svs[c - 1], svs[c + 1]);
            b[k] = A_::mul(A_::mul(q, A.hx), A.hy);   // (rhs * hx) * hy, :144, once per step
            if (last_step) {
                const int x = c / Y;
                A.f[6][(long long)x * pitch + (c - x * Y)] = q;
            }
        }
    };
    // p neighbours of slot k in array `src`: redirected, or -- before any sweep of this launch -- the true ones
    auto corrector = [&](auto mode_tag, const T* src, bool redirected) {
        constexpr int MODE = decltype(mode_tag)::value;
#pragma unroll
        for (int k = 0; k < NC; ++k) {
            const int c = cell[k];
            if (c >= n) continue;
            const int x = c / Y, y = c - x * Y;
            T un, vn;
            if (interior[k]) {
                const T pxm = src[redirected ? axm[k] : c - Y], pxp = src[redirected ? axp[k] : c + Y];
                const T pym = src[redirected ? aym[k] : c - 1], pyp = src[redirected ? ayp[k] : c + 1];
                Flags fl;
                UV<T> q = correct_cell<T, MODE>(A.ca, pxm, pxp, pym, pyp, sus[c], svs[c], fl);
                if (MODE != DIV_EXACT && fl.raised(A.ca.m.pow2_lo_m1)) q = correct_cell_exact<T, MODE>(A.ca, pxm, pxp, pym, pyp, sus[c], svs[c]);
                un = q.u;
                vn = q.v;
            } else {   // the y edges own their cells for all x (:236-264), the x walls the cells between (:266-294)
                const int e = y == 0 ? BOTTOM : y == Y - 1 ? TOP : x == 0 ? LEFT : RIGHT;
                un = A.uval[e];
                vn = A.vval[e];
            }
            su[c] = un;
            sv[c] = vn;
        }
    };
    using M0 = ModeTag<DIV_EXACT>;
    using M1 = ModeTag<DIV_FAST>;
    using M2 = ModeTag<DIV_POW2>;

    T* src = pbuf0;
    T* dst = pbuf1;
    int g = 0;
    for (int step = 0; step < A.S; ++step) {
        if (mode == 2) tentative(M2{}); else if (mode == 1) tentative(M1{}); else tentative(M0{});
        __syncthreads();
        const bool last = step == A.S - 1;
        if (mode == 2) rhs(M2{}, last); else if (mode == 1) rhs(M1{}, last); else rhs(M0{}, last);
        for (int s = 0; s < A.K; ++s, ++g) {
            if (g == 0) {   // the uploaded boundary values
#pragma unroll
                for (int k = 0; k < NC; ++k) {
                    if (!interior[k]) continue;
                    const int c = cell[k];
                    const T sum = A_::add(A_::add(A_::add(src[c - Y], src[c + Y]), src[c + 1]), src[c - 1]);
                    dst[c] = A_::neg_quarter(A_::sub(b[k], sum));
                }
            } else {
#pragma unroll
                for (int k = 0; k < NC; ++k) {
                    if (!interior[k]) continue;
                    const T sum = A_::add(A_::add(A_::add(src[axm[k]], src[axp[k]]), src[ayp[k]]), src[aym[k]]);
                    dst[cell[k]] = A_::neg_quarter(A_::sub(b[k], sum));
                }
            }
            __syncthreads();
            T* t = src; src = dst; dst = t;
        }
        if (mode == 2) corrector(M2{}, src, g > 0); else if (mode == 1) corrector(M1{}, src, g > 0); else corrector(M0{}, src, g > 0);
        __syncthreads();
    }

    // write the state back. `src` holds the final p on the interior, `dst` the state before the last sweep.
    for (int c = tid; c < n; c += NT) {
        const int x = c / Y, y = c - x * Y;
        const long long o = (long long)x * pitch + y;
        const bool xin = x >= 1 && x <= X - 2, yin = y >= 1 && y <= Y - 2;
        T pv;
        if (xin && yin) {
            pv = src[c];
            A.f[3][o] = sus[c];
            A.f[4][o] = svs[c];
        } else if (xin) {            // y edges, rows 1 .. X-2 (:163-175)
            pv = y == 0 ? (bconst ? A.pval[BOTTOM] : src[c + 1]) : (tconst ? A.pval[TOP] : src[c - 1]);
        } else if (yin) {            // x walls, columns 1 .. Y-2 (:177-188)
            pv = x == 0 ? (lconst ? A.pval[LEFT] : src[c + Y]) : (rconst ? A.pval[RIGHT] : src[c - Y]);
        } else {                     // corners: copy the wall cell beside them as it was one sweep earlier
            const int inner = (x == 0 ? c + Y : c - Y) + (y == 0 ? 1 : -1);   // the interior cell diagonal to the corner
            const T wall_prev = x == 0 ? (lconst ? A.pval[LEFT] : dst[inner]) : (rconst ? A.pval[RIGHT] : dst[inner]);
            pv = y == 0 ? (bconst ? A.pval[BOTTOM] : wall_prev) : (tconst ? A.pval[TOP] : wall_prev);
        }
        A.f[2][o] = pv;
        A.f[0][o] = su[c];
        A.f[1][o] = sv[c];
    }
}

}  // namespace fb
#endif  // !FLUIDB200_HOST_EMU
