// ref_cuda_driver.cu -- thin C-ABI driver around the UNMODIFIED reference CUDA solver, for a same-box
// comparison in bench.py: ParallelNavierStokes<float> compiled for sm_100a.
//
// MEASUREMENT INFRASTRUCTURE ONLY (see oracle/fluid_oracle.c). This file contains no reference code: it
// #includes the reference's header from /root/reference and is linked with the reference's own
// cuda/ParallelNavierStokes.cu and cpp/NavierStokesSolver.cpp, compiled where they lie (oracle/Makefile
// target `ref_cuda`, output only under oracle/_ref/). Nothing in fluid_simulation_b200/ loads it.
//
// Timing follows fluid_cpp/cuda_main.cu:96-99 (steady clock around solve(); cudaDeviceSynchronize())
// with one difference the reference's own driver forgets: migrateHostToDevice() is
// called after the boundary conditions are set, so the GPU works on the cavity and not on an all-zero grid.
#include <cuda_runtime.h>

#include <chrono>
#include <cstddef>

#include "cuda/ParallelNavierStokes.cuh"

// v_out (optional): X*Y floats, i = x*Y + y, the layout fluid_cpp/tests/test_main.cpp:76 reads.
extern "C" int ref_cuda_cavity(int X, int Y, double wx, double wy, double dt, double nu, double rho, int K, int S,
                               int warmup_steps, float* v_out, double* elapsed_ms) {
    ParallelNavierStokes<float> s(X, Y, (float)wx, (float)wy);
    s.time_step = (float)dt;
    s.kinematic_viscosity = (float)nu;
    s.density = (float)rho;
    s.num_poisson_iterations = K;
    for (int x = 0; x < X; x++) {   // boundary set-up of fluid_cpp/main.cpp:43-55
        s.setUBoundaryCondition(x, 0, 0.);
        s.setVBoundaryCondition(x, 0, 0.);
        s.setUBoundaryCondition(x, Y - 1, 1.);
        s.setVBoundaryCondition(x, Y - 1, 0.);
        s.setPBoundaryCondition(x, Y - 1, 0.);
    }
    for (int y = 1; y < Y - 1; y++) {
        s.setUBoundaryCondition(0, y, 0.);
        s.setVBoundaryCondition(0, y, 0.);
        s.setUBoundaryCondition(X - 1, y, 0.);
        s.setVBoundaryCondition(X - 1, y, 0.);
    }
    s.migrateHostToDevice();
    if (warmup_steps > 0) {
        s.num_iterations = warmup_steps;
        s.solve();
        cudaDeviceSynchronize();
        s.migrateHostToDevice();   // start the timed run from the initial state again
    }
    s.num_iterations = S;
    const auto t0 = std::chrono::steady_clock::now();
    s.solve();
    const cudaError_t rc = cudaDeviceSynchronize();
    const auto t1 = std::chrono::steady_clock::now();
    *elapsed_ms = std::chrono::duration<double, std::milli>(t1 - t0).count();
    if (v_out && X == Y) {   // (the reference's getters use stride box_dimension_x: square grids only)
        s.migrateDeviceToHost();
        s.getVValues(v_out);
    }
    return rc == cudaSuccess ? 0 : (int)rc;
}
