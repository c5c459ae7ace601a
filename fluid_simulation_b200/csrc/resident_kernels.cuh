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
    const int c0 = tc0 + wyi * W + lane * V;    // first column of this thread's patch
    const bool col_ok = (long long)c0 + V <= pitch;   // the lane's 16 bytes exist in memory

    // shared memory: four frames (tile + halo) for the predictor, then the exchange buffers of the sweeps
    const int FR = TR + 2 * RH, FP = TC + 2 * CH;
    T* su = reinterpret_cast<T*>(fb_res_smem);
    T* sv = su + FR * FP;
    T* sus = sv + FR * FP;      // u_next
    T* svs = sus + FR * FP;     // v_next
    T* rowbuf = svs + FR * FP;                  // [parity][warp][first / last row of the patch][W]
    T* colbuf = rowbuf + 2 * nwarps * 2 * W;    // [parity][warp][first / last column of the warp][R]
    const int fbase = (x0 - tr0 + RH) * FP + (c0 - tc0 + CH);   // frame index of patch cell (row 0, element 0)

    T* const U = A.f[0];
    T* const Vf = A.f[1];
    T* const gp[2] = {A.f[5], A.f[2]};   // pressure as published for the neighbours: p_g lives in gp[g & 1]; the
                                         // state before the first sweep (g = -1) is the solver's p
    unsigned* const my_cnt = A.sync + 32 * (1 + tile);
    const bool bwarp = ntiles > 1 && (wxi == 0 || wxi == wx - 1 || wyi == 0 || wyi == wy - 1);
    const unsigned nbw = (unsigned)(nwarps - (wx > 2 ? wx - 2 : 0) * (wy > 2 ? wy - 2 : 0));  // boundary warps of a tile

    // y edges of this warp (SerialNavierStokes.cpp:163-175)
    const int wc0 = tc0 + wyi * W;
    const bool has_yedge = (wc0 == 0) || (wc0 <= Y - 1 && Y - 1 < wc0 + W);
    YEdgeUniform<T, V> ye;
    ye.lane0_bottom = (wc0 == 0) && lane == 0;
    ye.bot_use = A.pmode[BOTTOM] == EDGE_CONST;
    ye.top_use = A.pmode[TOP] == EDGE_CONST;
    ye.bot_v = A.pval[BOTTOM];
    ye.top_v = A.pval[TOP];
    ye.top_bits = 0u;
#pragma unroll
    for (int e = 0; e < V; ++e) ye.top_bits |= (c0 + e == Y - 1) ? (1u << e) : 0u;
    const bool lconst = A.pmode[LEFT] == EDGE_CONST, rconst = A.pmode[RIGHT] == EDGE_CONST;
    // Column Y-1 can be the FIRST column of a warp (Y = 1 + m * 32V: 129, 257, 1025 ...): its inward
    // neighbour Y-2 then belongs to another warp, so the column cannot be refreshed in the same sweep.
    // It is treated like an x wall instead: not stored, substituted where column Y-2 needs it, written
    // at the end.
    const bool top_detached = ((Y - 1) % W) == 0;
    if (top_detached) ye.top_bits = 0u;
    const bool warp_top_det = top_detached && (wc0 + W - 1 == Y - 2);   // this warp's last column is Y-2

    VT p[R], b[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const int x = x0 + r;
        p[r] = (x < X && col_ok) ? ld_cg<T, V>(A.f[2] + (long long)x * pitch + c0) : vzero<T, V>();
        b[r] = vzero<T, V>();
    }

    // ---- exchange of patch borders ----------------------------------------------------------
    auto publish = [&](int par, T* gdst) {
        T* rb = rowbuf + (size_t)((par * nwarps + warp) * 2) * W + lane * V;
        st_vec<T, V>(rb, p[0]);
        st_vec<T, V>(rb + W, p[R - 1]);
        T* cb = colbuf + (size_t)((par * nwarps + warp) * 2) * R;
        if (lane == 0) {
#pragma unroll
            for (int r = 0; r < R; ++r) cb[r] = p[r].a[0];
        }
        if (lane == 31) {
#pragma unroll
            for (int r = 0; r < R; ++r) cb[R + r] = p[r].a[V - 1];
        }
        if (bwarp && gdst) {   // (warp-uniform) the tile's outer rows and columns, for the adjacent tiles
            if (wxi == 0 && txi > 0 && col_ok && x0 < X) st_vec<T, V>(gdst + (long long)x0 * pitch + c0, p[0]);
            if (wxi == wx - 1 && txi < tiles_x - 1 && col_ok && x0 + R - 1 < X)
                st_vec<T, V>(gdst + (long long)(x0 + R - 1) * pitch + c0, p[R - 1]);
            if (wyi == 0 && tyi > 0 && lane == 0) {
#pragma unroll
                for (int r = 0; r < R; ++r)
                    if (x0 + r < X) gdst[(long long)(x0 + r) * pitch + c0] = p[r].a[0];
            }
            if (wyi == wy - 1 && tyi < tiles_y - 1 && lane == 31) {
#pragma unroll
                for (int r = 0; r < R; ++r)
                    if (x0 + r < X) gdst[(long long)(x0 + r) * pitch + c0 + V - 1] = p[r].a[V - 1];
            }
            __syncwarp();
            if (lane == 0) {
                __threadfence();
                red_release_gpu(my_cnt, 1u);
            }
        }
    };
    auto wait_tile = [&](int other, unsigned need) {
        const unsigned* c = A.sync + 32 * (1 + other);
        while (ld_acquire_gpu(c) < need) {}
    };
    // the four sides of the patch as published with parity `par` (`need` publications of a neighbouring tile)
    auto fetch = [&](int par, const T* gsrc, unsigned need, VT& up, VT& dn, T (&lf)[R], T (&rt)[R]) {
        const T* rb = rowbuf + (size_t)(par * nwarps * 2) * W + lane * V;
        const T* cb = colbuf + (size_t)(par * nwarps * 2) * R;
        up = vzero<T, V>();
        dn = vzero<T, V>();
        if (wxi > 0) {
            up = *reinterpret_cast<const VT*>(rb + (size_t)((warp - wy) * 2 + 1) * W);
        } else if (txi > 0) {
            wait_tile(tile - tiles_y, need);
            if (col_ok) up = ld_cg<T, V>(gsrc + (long long)(x0 - 1) * pitch + c0);
        }
        if (wxi < wx - 1) {
            dn = *reinterpret_cast<const VT*>(rb + (size_t)((warp + wy) * 2) * W);
        } else if (txi < tiles_x - 1) {
            wait_tile(tile + tiles_y, need);
            if (col_ok && x0 + R < X) dn = ld_cg<T, V>(gsrc + (long long)(x0 + R) * pitch + c0);
        }
#pragma unroll
        for (int r = 0; r < R; ++r) {
            lf[r] = T(0);
            rt[r] = T(0);
        }
        if (wyi > 0) {
            if (lane == 0) {
#pragma unroll
                for (int r = 0; r < R; ++r) lf[r] = cb[(size_t)((warp - 1) * 2 + 1) * R + r];
            }
        } else if (tyi > 0) {
            wait_tile(tile - 1, need);
            if (lane == 0) {
#pragma unroll
                for (int r = 0; r < R; ++r)
                    if (x0 + r < X) lf[r] = ld_cg1<T>(gsrc + (long long)(x0 + r) * pitch + c0 - 1);
            }
        }
        if (wyi < wy - 1) {
            if (lane == 31) {
#pragma unroll
                for (int r = 0; r < R; ++r) rt[r] = cb[(size_t)((warp + 1) * 2) * R + r];
            }
        } else if (tyi < tiles_y - 1) {
            wait_tile(tile + 1, need);
            if (lane == 31) {
#pragma unroll
                for (int r = 0; r < R; ++r)
                    if (x0 + r < X) rt[r] = ld_cg1<T>(gsrc + (long long)(x0 + r) * pitch + c0 + V);
            }
        }
    };
    auto splat = [](T v) {
        VT r;
#pragma unroll
        for (int e = 0; e < V; ++e) r.a[e] = v;
        return r;
    };

    // ---- one Jacobi sweep: p_g from p_{g-1} (SerialNavierStokes.cpp:140-198) ---------------------
    auto sweep = [&](int g) {
        const int sp = (g + 1) & 1;   // parity of g - 1
        VT up, dn;
        T lf[R], rt[R];
        fetch(sp, gp[sp], nbw * (unsigned)g, up, dn, lf, rt);
        const bool subst = g > 0;     // the first sweep of a launch reads the wall rows as they were uploaded
        VT prev = up;
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const int x = x0 + r;
            const VT cur = p[r];
            const VT nxt = (r < R - 1) ? p[r < R - 1 ? r + 1 : r] : dn;
            if (x >= 1 && x <= X - 2) {   // (warp-uniform)
                VT xm = prev, xp = nxt;
                if (subst && x == 1) xm = lconst ? splat(A.pval[LEFT]) : cur;         // :177-181 as of the previous sweep
                if (subst && x == X - 2) xp = rconst ? splat(A.pval[RIGHT]) : cur;    // :183-187
                VT ym, yp;
                y_neighbours<T, V>(cur, ym, yp);
                ym.a[0] = (lane == 0) ? lf[r] : ym.a[0];
                yp.a[V - 1] = (lane == 31) ? rt[r] : yp.a[V - 1];
                if (warp_top_det && subst) yp.a[V - 1] = (lane == 31) ? (ye.top_use ? ye.top_v : cur.a[V - 1]) : yp.a[V - 1];
                VT out = jacobi_row_yn<T, V>(xm, xp, ym, yp, b[r]);
                if (has_yedge) ye.apply(out);   // columns 0 and Y-1 of this row (:163-175)
                p[r] = out;
            }
            prev = cur;
        }
        publish(g & 1, gp[g & 1]);
    };

    // ---- predictor: u_next, v_next on the tile + 1 ring, rhs on the patch (SerialNavierStokes.cpp:298-335) ----
    auto load_frames = [&]() {
        for (int idx = tid; idx < FR * FP; idx += nthreads) {
            const int i = idx / FP, j = idx - i * FP;
            const int gx = tr0 - RH + i, gy = tc0 - CH + j;
            T uu = T(0), vv = T(0), us = T(0), vs = T(0);
            if (gx >= 0 && gx < X && gy >= 0 && gy < Y) {
                const long long o = (long long)gx * pitch + gy;
                uu = ld_cg1<T>(U + o);
                vv = ld_cg1<T>(Vf + o);
                if (gx == 0 || gx == X - 1 || gy == 0 || gy == Y - 1) {  // edge u_next / v_next: stored values, never written
                    us = A.f[3][o];
                    vs = A.f[4][o];
                }
            }
            su[idx] = uu;
            sv[idx] = vv;
            sus[idx] = us;
            svs[idx] = vs;
        }
    };
    auto tentative_frame = [&](auto mode_tag) {
        constexpr int MODE = decltype(mode_tag)::value;
        const int nj = TC + 2, n = (TR + 2) * nj;
        for (int idx = tid; idx < n; idx += nthreads) {
            const int ii = idx / nj, jj = idx - ii * nj;
            const int i = RH - 1 + ii, j = CH - 1 + jj;
            const int gx = tr0 - RH + i, gy = tc0 - CH + j;
            if (gx < 1 || gx > X - 2 || gy < 1 || gy > Y - 2) continue;
            const int c = i * FP + j;
            Flags fl;
            UV<T> r = tentative_cell<T, MODE>(A.pa, su[c], sv[c], su[c - FP], su[c + FP], su[c - 1], su[c + 1], sv[c - FP],
                                              sv[c + FP], sv[c - 1], sv[c + 1], fl);
            if (MODE != DIV_EXACT && fl.raised(A.pa.m.pow2_lo_m1))
                r = tentative_cell_exact<T>(A.pa, su[c], sv[c], su[c - FP], su[c + FP], su[c - 1], su[c + 1], sv[c - FP],
                                            sv[c + FP], sv[c - 1], sv[c + 1]);
            sus[c] = r.u;
            svs[c] = r.v;
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
                        v = rhs_cell_exact<T>(A.pa, sus[c + e - FP], sus[c + e + FP], svs[c + e - 1], svs[c + e + 1]);
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
        const int sp = (g_done + 1) & 1;
        VT up, dn;
        T lf[R], rt[R];
        fetch(sp, gp[sp], nbw * (unsigned)g_done, up, dn, lf, rt);
        const bool subst = g_done > 0;
        VT prev = up;
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const int x = x0 + r;
            const VT cur = p[r];
            const VT nxt = (r < R - 1) ? p[r < R - 1 ? r + 1 : r] : dn;
            if (x < X) {   // (warp-uniform)
                VT un = vzero<T, V>(), vn = vzero<T, V>();
                if (x >= 1 && x <= X - 2) {
                    VT pm = prev, pp = nxt;
                    if (subst && x == 1) pm = lconst ? splat(A.pval[LEFT]) : cur;
                    if (subst && x == X - 2) pp = rconst ? splat(A.pval[RIGHT]) : cur;
                    VT pym, pyp;
                    y_neighbours<T, V>(cur, pym, pyp);
                    pym.a[0] = (lane == 0) ? lf[r] : pym.a[0];
                    pyp.a[V - 1] = (lane == 31) ? rt[r] : pyp.a[V - 1];
                    if (warp_top_det && subst) pyp.a[V - 1] = (lane == 31) ? (ye.top_use ? ye.top_v : cur.a[V - 1]) : pyp.a[V - 1];
                    const int c = fbase + r * FP;
#pragma unroll
                    for (int e = 0; e < V; ++e) {
                        const int col = c0 + e;
                        if (col >= 1 && col <= Y - 2) {
                            Flags fl;
                            UV<T> q = correct_cell<T, MODE>(A.ca, pm.a[e], pp.a[e], pym.a[e], pyp.a[e], sus[c + e], svs[c + e], fl);
                            if (MODE != DIV_EXACT && fl.raised(A.ca.m.pow2_lo_m1))
                                q = correct_cell_exact<T>(A.ca, pm.a[e], pp.a[e], pym.a[e], pyp.a[e], sus[c + e], svs[c + e]);
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
                if (col_ok) {
                    st_vec<T, V>(U + (long long)x * pitch + c0, un);
                    st_vec<T, V>(Vf + (long long)x * pitch + c0, vn);
                }
            }
            prev = cur;
        }
    };
    using M0 = ModeTag<DIV_EXACT>;
    using M1 = ModeTag<(sizeof(T) == 4) ? DIV_FAST : DIV_EXACT>;
    using M2 = ModeTag<DIV_POW2>;
    const int mode = A.pa.m.mode;   // (grid-uniform)

    // ---- the launch -------------------------------------------------------------------------------
    publish(1, nullptr);   // the uploaded p, for the first sweep's fetch (its global copy is the solver's p itself)
    __syncthreads();
    unsigned epoch = 0;
    const int G = A.S * A.K;
    int g = 0;
    for (int step = 0; step < A.S; ++step) {
        if (step > 0) resident_grid_barrier(A.sync, epoch, (unsigned)ntiles);   // every tile's new u, v are in L2
        load_frames();
        __syncthreads();
        if (mode == 2) tentative_frame(M2{}); else if (mode == 1) tentative_frame(M1{}); else tentative_frame(M0{});
        __syncthreads();
        const bool last = step == A.S - 1;
        if (mode == 2) rhs_patch(M2{}, last); else if (mode == 1) rhs_patch(M1{}, last); else rhs_patch(M0{}, last);
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
        }
        // a tile may overwrite u, v only when no other tile still reads them for ITS predictor: three
        // sweeps of neighbour hand-shakes order that even for the diagonal tiles; fewer do not
        if (A.K < 3) resident_grid_barrier(A.sync, epoch, (unsigned)ntiles);
        if (mode == 2) corrector(M2{}, g); else if (mode == 1) corrector(M1{}, g); else corrector(M0{}, g);
    }

    // ---- write the state back: p with its wall rows and corners, u_next, v_next ----------------------
    resident_grid_barrier(A.sync, epoch, (unsigned)ntiles);   // nobody reads a published border any more
    T* const P = A.f[2];
    bool deferred = false;   // row X-1 is the first row of this patch: its inward neighbour is another thread's
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const int x = x0 + r;
        if (x >= X) continue;
        VT out = p[r];
        if (x == 0 || x == X - 1) {   // x walls, columns 1 .. Y-2 (:177-188); corners are fixed below
            const bool cst = (x == 0) ? lconst : rconst;
            const T cv = (x == 0) ? A.pval[LEFT] : A.pval[RIGHT];
            if (x == X - 1 && r == 0) {
                deferred = true;
            } else {
                const VT inner = (x == 0) ? p[r < R - 1 ? r + 1 : r] : p[r > 0 ? r - 1 : 0];
#pragma unroll
                for (int e = 0; e < V; ++e) {
                    const int col = c0 + e;
                    if (col >= 1 && col <= Y - 2) out.a[e] = cst ? cv : inner.a[e];
                }
            }
        }
        if (col_ok && !(x == X - 1 && r == 0)) st_vec<T, V>(P + (long long)x * pitch + c0, out);
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
                P[(long long)x * pitch + Y - 1] = ye.top_use ? ye.top_v : ld_cg1<T>(P + (long long)x * pitch + Y - 2);
        }
    }
    if (tile == 0 && tid == 0) {   // corners: the y-edge loop runs first (:163-175), so an unset corner copies the
                                   // wall cell next to it as it was BEFORE this sweep's x-edge loop refreshed it
        const T l0 = lconst ? A.pval[LEFT] : ld_cg1<T>(A.stash + 0), l1 = lconst ? A.pval[LEFT] : ld_cg1<T>(A.stash + 1);
        const T r0 = rconst ? A.pval[RIGHT] : ld_cg1<T>(A.stash + 2), r1 = rconst ? A.pval[RIGHT] : ld_cg1<T>(A.stash + 3);
        P[0] = ye.bot_use ? ye.bot_v : l0;
        P[Y - 1] = ye.top_use ? ye.top_v : l1;
        P[(long long)(X - 1) * pitch] = ye.bot_use ? ye.bot_v : r0;
        P[(long long)(X - 1) * pitch + Y - 1] = ye.top_use ? ye.top_v : r1;
    }
}

}  // namespace fb
#endif  // !FLUIDB200_HOST_EMU
