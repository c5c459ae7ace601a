// ParallelNavierStokes.cuh -- drop-in for fluid_cpp/cuda/ParallelNavierStokes.{cuh,cu}: the same
// class template, constructor and methods (ParallelNavierStokes.cuh:16-47), implemented as a
// header-only shim over the C ABI of libfluidb200.so (include/fluidb200.h). Plain C++: it needs
// no nvcc, only the library at link time.
//
// Two ways to use it:
//  (a) stand-alone: put include/fluid_cpp on the include path; the base class is this repository's
//      cpp/NavierStokesSolver.h (structure-of-arrays host state);
//  (b) inside the reference's tree: replace its cuda/ParallelNavierStokes.{cuh,cu} (and
//      TiledNavierStokes.{cuh,cu}) with these files. "../cpp/NavierStokesSolver.h" then resolves to
//      the REFERENCE's base class and this shim packs/unpacks its array-of-cells host grid
//      (`cells[x][y]`, NavierStokesSolver.h:19) -- the reference's own tests/test_main.cpp compiles
//      and runs unmodified (oracle/Makefile target `ref_test_main` does exactly that).
// The mode is picked by the reference's include guard NAVIERSTOKESCELL_H.
//
// Semantics kept from the reference: setters touch host state only; migrateHostToDevice uploads
// the whole host grid (:244-249); solve() is asynchronous and continues from the device state
// (:269-288); getters read host state, so results need migrateDeviceToHost() first (:252-257).
// The reference ignores every CUDA error; this shim keeps the void signatures and records the
// last C-ABI error, readable with lastError().
#ifndef PARALLELNAVIERSTOKES_CUH
#define PARALLELNAVIERSTOKES_CUH

#include <cstddef>
#include <cstdint>
#include <vector>

#include "../cpp/NavierStokesSolver.h"
extern "C" {
#include "fluidb200.h"
}

template <class T>
class ParallelNavierStokes : public NavierStokesSolver<T> {
protected:
    fluidb200_solver* handle = nullptr;
    int last_error = 0;

    static int dtypeCode() { return sizeof(T) == 4 ? FLUIDB200_F32 : FLUIDB200_F64; }
    void note(int rc) { if (rc != 0) last_error = rc; }
    // setBoxDimenension (NavierStokesSolver.h:31) is a base-class method that knows nothing of the
    // device: the reference's CUDA class keeps its old d_cells after a resize. Here every device
    // call first checks that the handle still has the host grid's dimensions and re-creates it
    // (pristine, zero-filled, no flags -- what a newly constructed solver holds) when it does not.
    void ensureHandle() {
        fluidb200_slab g;
        if (handle && fluidb200_get_slab(handle, &g) == 0 && g.X == this->box_dimension_x && g.Y == this->box_dimension_y)
            return;
        if (handle) fluidb200_destroy(handle);
        handle = nullptr;
        note(fluidb200_create(&handle, dtypeCode(), this->box_dimension_x, this->box_dimension_y,
                              (double)this->domain_size_x, (double)this->domain_size_y));
        if (handle && sweeps_per_launch > 0) note(fluidb200_set_sweeps_per_launch(handle, sweeps_per_launch));
    }
    int sweeps_per_launch = 0;  // 0 = the library's choice by grid size
    void pushParams() {
        note(fluidb200_set_domain(handle, (double)this->domain_size_x, (double)this->domain_size_y));
        note(fluidb200_set_params(handle, (double)this->time_step, (double)this->kinematic_viscosity,
                                  (double)this->density, this->num_poisson_iterations, this->num_iterations));
    }

public:
    ParallelNavierStokes(int box_dim_x, int box_dim_y, T dom_x, T dom_y)
        : NavierStokesSolver<T>(box_dim_x, box_dim_y, dom_x, dom_y) {
        // the reference allocates the device grid and uploads the pristine (all-zero, flag-less) host
        // grid here (:226-235); fluidb200_create leaves the device state in exactly that condition
        note(fluidb200_create(&handle, dtypeCode(), box_dim_x, box_dim_y, (double)dom_x, (double)dom_y));
    }
    ~ParallelNavierStokes() { fluidb200_destroy(handle); }
    ParallelNavierStokes(const ParallelNavierStokes&) = delete;
    ParallelNavierStokes& operator=(const ParallelNavierStokes&) = delete;

    void migrateHostToDevice() {
        ensureHandle();
        const int X = this->box_dimension_x, Y = this->box_dimension_y;
        const void* fields[FLUIDB200_NUM_FIELDS];
        const void* bcv[3];
        const std::uint8_t* bcs[3];
#ifdef NAVIERSTOKESCELL_H  // (b) the reference's array-of-cells host grid
        const std::size_t n = (std::size_t)X * Y;
        std::vector<T> f[FLUIDB200_NUM_FIELDS], v[3];
        std::vector<std::uint8_t> s[3];
        for (auto& a : f) a.resize(n);
        for (int k = 0; k < 3; ++k) { v[k].resize(n); s[k].resize(n); }
        for (int x = 0; x < X; x++) {
            for (int y = 0; y < Y; y++) {
                const NavierStokesCell<T>& c = this->cells[x][y];
                const std::size_t i = (std::size_t)x * Y + y;
                f[FLUIDB200_FIELD_U][i] = c.u; f[FLUIDB200_FIELD_V][i] = c.v; f[FLUIDB200_FIELD_P][i] = c.p;
                f[FLUIDB200_FIELD_U_NEXT][i] = c.u_next; f[FLUIDB200_FIELD_V_NEXT][i] = c.v_next;
                f[FLUIDB200_FIELD_P_NEXT][i] = c.p_next; f[FLUIDB200_FIELD_RHS][i] = c.right_hand_size;
                v[0][i] = c.u_boundary; v[1][i] = c.v_boundary; v[2][i] = c.p_boundary;
                s[0][i] = c.u_boundary_set; s[1][i] = c.v_boundary_set; s[2][i] = c.p_boundary_set;
            }
        }
        for (int k = 0; k < FLUIDB200_NUM_FIELDS; ++k) fields[k] = f[k].data();
        for (int k = 0; k < 3; ++k) { bcv[k] = v[k].data(); bcs[k] = s[k].data(); }
        note(fluidb200_upload(handle, fields, bcv, bcs));
#else                      // (a) this repository's structure-of-arrays base
        (void)X; (void)Y;
        for (int k = 0; k < FLUIDB200_NUM_FIELDS; ++k) fields[k] = this->hostField(k).data();
        for (int k = 0; k < 3; ++k) {
            this->touchBc(k);
            bcv[k] = this->bc_val_[k].data();
            bcs[k] = this->bc_set_[k].data();
        }
        note(fluidb200_upload(handle, fields, bcv, bcs));
#endif
    }

    void migrateDeviceToHost() {
        ensureHandle();
#ifdef NAVIERSTOKESCELL_H
        const int X = this->box_dimension_x, Y = this->box_dimension_y;
        const std::size_t n = (std::size_t)X * Y;
        std::vector<T> f[FLUIDB200_NUM_FIELDS];
        for (int k = 0; k < FLUIDB200_NUM_FIELDS; ++k) {
            f[k].resize(n);
            note(fluidb200_download(handle, k, f[k].data()));
        }
        for (int x = 0; x < X; x++) {
            for (int y = 0; y < Y; y++) {
                NavierStokesCell<T>& c = this->cells[x][y];
                const std::size_t i = (std::size_t)x * Y + y;
                c.u = f[FLUIDB200_FIELD_U][i]; c.v = f[FLUIDB200_FIELD_V][i]; c.p = f[FLUIDB200_FIELD_P][i];
                c.u_next = f[FLUIDB200_FIELD_U_NEXT][i]; c.v_next = f[FLUIDB200_FIELD_V_NEXT][i];
                c.p_next = f[FLUIDB200_FIELD_P_NEXT][i]; c.right_hand_size = f[FLUIDB200_FIELD_RHS][i];
            }
        }
#else
        for (int k = 0; k < FLUIDB200_NUM_FIELDS; ++k) note(fluidb200_download(handle, k, this->hostField(k).data()));
#endif
    }

    // asynchronous, like the reference (:269-288); callers synchronise (cuda_main.cu:98) or migrate
    void solve() {
        ensureHandle();
        pushParams();
        note(fluidb200_solve(handle));
    }
    void migrateSolve() {
        this->migrateHostToDevice();
        this->solve();
        this->migrateDeviceToHost();
    }

    // additions (not in the reference)
    void synchronize() { note(fluidb200_synchronize(handle)); }
    void uploadCavity(T lid_velocity) { ensureHandle(); note(fluidb200_upload_cavity(handle, (double)lid_velocity)); }
    int lastError() const { return last_error; }
    const char* lastErrorMessage() const { return fluidb200_last_error(); }
    fluidb200_solver* nativeHandle() { ensureHandle(); return handle; }
};

// explicit instantiation allows float and double precision types (ParallelNavierStokes.cuh:50-51)
template class ParallelNavierStokes<float>;
template class ParallelNavierStokes<double>;

#endif  // PARALLELNAVIERSTOKES_CUH
