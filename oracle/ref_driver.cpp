// ref_driver.cpp -- thin C-ABI driver around the UNMODIFIED reference CPU solvers.
//
// TEST INFRASTRUCTURE ONLY (see oracle/fluid_oracle.c). This file contains no
// reference code: it #includes the reference's headers from /root/reference and is
// linked with the reference's own .cpp files, compiled where they lie (oracle/Makefile,
// outputs only under oracle/_ref/). It exists so that
//   * tests can pin oracle/fluid_oracle.c against the real SerialNavierStokes<T>, and
//   * bench.py can time the reference's serial / threaded CPU path beside the GPU.
//
// The solvers are driven only through their public API (fields, set*BoundaryCondition,
// solve / safeSolve -- the call sequence of fluid_cpp/main.cpp:31-59 and
// fluid_cpp/tests/test_main.cpp:19-65). Results are read through a subclass that can
// see the protected `cells` member (NavierStokesSolver.h:19) because the public getters
// use stride box_dimension_x (NavierStokesSolver.cpp:91) and are only right for square
// grids. Output layout is i = x*Y + y.
#include <chrono>
#include <cstddef>

#include "cpp/SerialNavierStokes.h"
#include "cpp/ThreadedNavierStokes.h"

namespace {

template <class Solver, class T>
struct Probe : public Solver {
    Probe(int X, int Y, T wx, T wy) : Solver(X, Y, wx, wy) {}
    void dump(T* u, T* v, T* p, T* un, T* vn, T* pn, T* rhs) {
        const int X = this->box_dimension_x, Y = this->box_dimension_y;
        for (int x = 0; x < X; x++) {
            for (int y = 0; y < Y; y++) {
                const size_t i = (size_t)x * Y + y;
                const NavierStokesCell<T>& c = this->cells[x][y];
                if (u) u[i] = c.u;
                if (v) v[i] = c.v;
                if (p) p[i] = c.p;
                if (un) un[i] = c.u_next;
                if (vn) vn[i] = c.v_next;
                if (pn) pn[i] = c.p_next;
                if (rhs) rhs[i] = c.right_hand_size;
            }
        }
    }
    T hx() const { return this->element_length_x; }
    T hy() const { return this->element_length_y; }
};

template <class Solver, class T>
int run(int kind, int X, int Y, double wx, double wy, double dt, double nu, double rho, int K,
        int S, const unsigned char* u_set, const void* u_bc, const unsigned char* v_set,
        const void* v_bc, const unsigned char* p_set, const void* p_bc, void* u, void* v, void* p,
        void* un, void* vn, void* pn, void* rhs, double* elapsed_ms) {
    Probe<Solver, T> s(X, Y, (T)wx, (T)wy);
    s.time_step = (T)dt;
    s.kinematic_viscosity = (T)nu;
    s.density = (T)rho;
    s.num_poisson_iterations = K;
    s.num_iterations = S;
    const T* ub = (const T*)u_bc;
    const T* vb = (const T*)v_bc;
    const T* pb = (const T*)p_bc;
    for (int x = 0; x < X; x++) {
        for (int y = 0; y < Y; y++) {
            const size_t i = (size_t)x * Y + y;
            if (u_set && u_set[i]) s.setUBoundaryCondition(x, y, ub[i]);
            if (v_set && v_set[i]) s.setVBoundaryCondition(x, y, vb[i]);
            if (p_set && p_set[i]) s.setPBoundaryCondition(x, y, pb[i]);
        }
    }
    auto t0 = std::chrono::steady_clock::now();
    if constexpr (requires { s.safeSolve(); }) {
        if (kind == 1) s.safeSolve(); else s.solve();
    } else {
        s.solve();
    }
    auto t1 = std::chrono::steady_clock::now();
    if (elapsed_ms) *elapsed_ms = std::chrono::duration<double, std::milli>(t1 - t0).count();
    s.dump((T*)u, (T*)v, (T*)p, (T*)un, (T*)vn, (T*)pn, (T*)rhs);
    return 0;
}

}  // namespace

// kind: 0 = SerialNavierStokes::solve, 1 = SerialNavierStokes::safeSolve,
//       2 = ThreadedNavierStokes::solve (fixed 8 threads, ThreadedNavierStokes.h:13).
// *_set / *_bc: X*Y arrays (i = x*Y + y); a non-zero flag means "call the setter with bc[i]".
// Output pointers may be NULL. Returns 0, or -1 for a bad `kind`.
extern "C" int ref_solve(int kind, int is_double, int X, int Y, double wx, double wy, double dt,
                         double nu, double rho, int K, int S, const unsigned char* u_set,
                         const void* u_bc, const unsigned char* v_set, const void* v_bc,
                         const unsigned char* p_set, const void* p_bc, void* u, void* v, void* p,
                         void* un, void* vn, void* pn, void* rhs, double* elapsed_ms) {
#define ARGS kind, X, Y, wx, wy, dt, nu, rho, K, S, u_set, u_bc, v_set, v_bc, p_set, p_bc, u, v, p, \
             un, vn, pn, rhs, elapsed_ms
    if (kind == 0 || kind == 1) {
        return is_double ? run<SerialNavierStokes<double>, double>(ARGS)
                         : run<SerialNavierStokes<float>, float>(ARGS);
    }
    if (kind == 2) {
        return is_double ? run<ThreadedNavierStokes<double>, double>(ARGS)
                         : run<ThreadedNavierStokes<float>, float>(ARGS);
    }
#undef ARGS
    return -1;
}

extern "C" int ref_threaded_grid_size(void) { return THREADED_GRID_SIZE; }
