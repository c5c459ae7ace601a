// test_main_b200.cpp -- the C++ face of the parity test, written to read like the reference's
// fluid_cpp/tests/test_main.cpp: same set-up (41 x 41, 1000 steps, 50 Poisson iterations), same
// boundary loops, same `==` comparison at index x*num_y_bins + y -- but it checks U, V and P (the
// reference checks V only), tests TiledNavierStokes as well, and EXITS NON-ZERO on a mismatch (the
// reference always returns 0). The "unit solver" is the oracle (oracle/liboracle.so, the pinned C
// restatement of SerialNavierStokes), loaded with dlopen so that no reference code is linked.
//
// build: g++ -std=c++17 -Iinclude -Iinclude/fluid_cpp tests/cpp/test_main_b200.cpp \
//            -Lfluid_simulation_b200 -lfluidb200 -ldl -Wl,-rpath,<repo>/fluid_simulation_b200
#include <dlfcn.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <vector>

#include "cuda/ParallelNavierStokes.cuh"
#include "cuda/TiledNavierStokes.cuh"
#include "cuda/MultiGpuNavierStokes.cuh"

struct OracleStateF32 {  // mirrors oracle_state_f32 in oracle/fluid_oracle_impl.inc
    int X, Y;
    float hx, hy, dt, nu, rho;
    int K, S;
    float *u, *v, *p, *u_next, *v_next, *p_next, *rhs, *u_bc, *v_bc, *p_bc;
    unsigned char *u_set, *v_set, *p_set;
};

template <class Solver>
static void cavity(Solver& solver, int num_x_bins, int num_y_bins) {
    for (int x = 0; x < num_x_bins; x++) {
        solver.setUBoundaryCondition(x, 0, 0.);               // no flow inside of the floor
        solver.setVBoundaryCondition(x, 0, 0.);               // no flow into the floor
        solver.setUBoundaryCondition(x, num_y_bins - 1, 1.);  // lid moves to the right
        solver.setVBoundaryCondition(x, num_y_bins - 1, 0.);  // no flow into the top
        solver.setPBoundaryCondition(x, num_y_bins - 1, 0.);  // no pressure at the top
    }
    for (int y = 1; y < num_y_bins - 1; y++) {
        solver.setUBoundaryCondition(0, y, 0.);
        solver.setVBoundaryCondition(0, y, 0.);
        solver.setUBoundaryCondition(num_x_bins - 1, y, 0.);
        solver.setVBoundaryCondition(num_x_bins - 1, y, 0.);
    }
}

struct OracleRun {
    int X, Y;
    std::vector<float> f[10];
    std::vector<unsigned char> s[3];
    OracleStateF32 st;
    OracleRun(int X_, int Y_) : X(X_), Y(Y_) {
        for (auto& a : f) a.assign((size_t)X * Y, 0.f);
        for (auto& a : s) a.assign((size_t)X * Y, 0);
    }
    int setBc(int k, int x, int y, float bc) {
        const size_t i = (size_t)x * Y + y;
        f[7 + k][i] = bc; f[k][i] = bc; s[k][i] = 1;
        return 0;
    }
    int setUBoundaryCondition(int x, int y, float bc) { return setBc(0, x, y, bc); }
    int setVBoundaryCondition(int x, int y, float bc) { return setBc(1, x, y, bc); }
    int setPBoundaryCondition(int x, int y, float bc) { return setBc(2, x, y, bc); }
};

template <class Solver>
static int compare(const char* name, Solver& tested, OracleRun& unit, int nx, int ny, bool verbose) {
    int total_errors = 0;
    std::vector<float> tested_values((size_t)nx * ny);
    const char* field_names[3] = {"U", "V", "P"};
    for (int field = 0; field < 3; field++) {
        if (field == 0) tested.getUValues(tested_values.data());
        if (field == 1) tested.getVValues(tested_values.data());
        if (field == 2) tested.getPValues(tested_values.data());
        const float* unit_values = unit.f[field].data();
        int num_errors = 0;
        for (int x = 0; x < nx; x++) {
            for (int y = 0; y < ny; y++) {
                int ind = x * ny + y;
                if (!(unit_values[ind] == tested_values[ind])) {
                    num_errors++;
                    if (verbose && num_errors <= 5)
                        std::cout << "FAILURE " << unit_values[ind] << " :: " << tested_values[ind] << " x=" << x << " y=" << y << std::endl;
                }
            }
        }
        std::cout << name << " " << field_names[field] << ": Passed "
                  << 100.f * float(nx * ny - num_errors) / float(nx * ny) << "% cells" << std::endl;
        total_errors += num_errors;
    }
    return total_errors;
}

int main(int argc, char** argv) {
    const char* oracle_path = argc > 1 ? argv[1] : "oracle/liboracle.so";
    void* lib = dlopen(oracle_path, RTLD_NOW);
    if (!lib) { std::cerr << "cannot load the oracle: " << dlerror() << std::endl; return 2; }
    auto oracle_solve = (void (*)(OracleStateF32*))dlsym(lib, "oracle_solve_f32");
    auto element_length = (float (*)(float, int))dlsym(lib, "oracle_element_length_f32");
    if (!oracle_solve || !element_length) { std::cerr << "oracle symbols missing" << std::endl; return 2; }

    int num_x_bins = 41, num_y_bins = 41;
    float width = 1.0, height = 1.0;
    OracleRun unit_solver(num_x_bins, num_y_bins);
    ParallelNavierStokes<float> tested_solver(num_x_bins, num_y_bins, width, height);
    TiledNavierStokes<float> tiled_solver(num_x_bins, num_y_bins, width, height);
    // constructed with another shape, then resized through the base-class method
    // (NavierStokesSolver.h:31): the device state must follow
    ParallelNavierStokes<float> resized_solver(20, 30, width, height);
    if (resized_solver.setBoxDimenension(num_x_bins, num_y_bins) != 0) { std::cerr << "resize failed" << std::endl; return 4; }

    tested_solver.density = 1.0;
    tested_solver.kinematic_viscosity = 0.1;
    tested_solver.num_iterations = 1000;
    tested_solver.num_poisson_iterations = 50;
    tested_solver.time_step = 0.001;
    tested_solver.stability_safety_factor = 0.5;
    tiled_solver.density = tested_solver.density;
    tiled_solver.kinematic_viscosity = tested_solver.kinematic_viscosity;
    tiled_solver.num_iterations = tested_solver.num_iterations;
    tiled_solver.num_poisson_iterations = tested_solver.num_poisson_iterations;
    tiled_solver.time_step = tested_solver.time_step;
    resized_solver.density = tested_solver.density;
    resized_solver.kinematic_viscosity = tested_solver.kinematic_viscosity;
    resized_solver.num_iterations = tested_solver.num_iterations;
    resized_solver.num_poisson_iterations = tested_solver.num_poisson_iterations;
    resized_solver.time_step = tested_solver.time_step;

    cavity(unit_solver, num_x_bins, num_y_bins);
    cavity(tested_solver, num_x_bins, num_y_bins);
    cavity(tiled_solver, num_x_bins, num_y_bins);
    cavity(resized_solver, num_x_bins, num_y_bins);

    OracleStateF32& st = unit_solver.st;
    st.X = num_x_bins; st.Y = num_y_bins;
    st.hx = element_length(width, num_x_bins); st.hy = element_length(height, num_y_bins);
    st.dt = tested_solver.time_step; st.nu = tested_solver.kinematic_viscosity; st.rho = tested_solver.density;
    st.K = tested_solver.num_poisson_iterations; st.S = tested_solver.num_iterations;
    float** fp[] = {&st.u, &st.v, &st.p, &st.u_next, &st.v_next, &st.p_next, &st.rhs, &st.u_bc, &st.v_bc, &st.p_bc};
    for (int k = 0; k < 10; k++) *fp[k] = unit_solver.f[k].data();
    st.u_set = unit_solver.s[0].data(); st.v_set = unit_solver.s[1].data(); st.p_set = unit_solver.s[2].data();

    // Solve and retrieve the solution
    oracle_solve(&st);
    tested_solver.migrateSolve();
    tiled_solver.migrateSolve();
    resized_solver.migrateSolve();
    if (tested_solver.lastError() || tiled_solver.lastError() || resized_solver.lastError()) {
        std::cerr << "fluidb200 error: " << tested_solver.lastErrorMessage() << std::endl;
        return 3;
    }
    int errors = compare("ParallelNavierStokes<float>", tested_solver, unit_solver, num_x_bins, num_y_bins, true);
    errors += compare("TiledNavierStokes<float>", tiled_solver, unit_solver, num_x_bins, num_y_bins, true);
    errors += compare("ParallelNavierStokes<float> after setBoxDimenension", resized_solver, unit_solver, num_x_bins, num_y_bins, true);
    {   // the multi-GPU class on the devices named in FLUIDB200_TEST_DEVICES (default: device 0 alone)
        std::vector<int> devices;
        if (const char* e = getenv("FLUIDB200_TEST_DEVICES"))
            for (const char* c = e; *c; ++c)
                if (*c >= '0' && *c <= '9') devices.push_back(*c - '0');
        MultiGpuNavierStokes<float> multi_solver(num_x_bins, num_y_bins, width, height, devices);
        multi_solver.density = tested_solver.density;
        multi_solver.kinematic_viscosity = tested_solver.kinematic_viscosity;
        multi_solver.num_iterations = tested_solver.num_iterations;
        multi_solver.num_poisson_iterations = tested_solver.num_poisson_iterations;
        multi_solver.time_step = tested_solver.time_step;
        multi_solver.uploadCavity(1.0f);
        multi_solver.solve();
        multi_solver.migrateDeviceToHost();
        if (multi_solver.lastError()) {
            std::cerr << "fluidb200 multi error: " << multi_solver.lastErrorMessage() << std::endl;
            return 3;
        }
        errors += compare("MultiGpuNavierStokes<float>", multi_solver, unit_solver, num_x_bins, num_y_bins, true);
    }
    return errors == 0 ? 0 : 1;
}
