// MultiGpuNavierStokes.cuh -- the reference's solver interface on SEVERAL GPUs of one box. The reference has no
// such class (no cudaSetDevice anywhere in its tree); this one keeps ParallelNavierStokes<T>'s shape
// (fluid_cpp/cuda/ParallelNavierStokes.cuh:16-47: constructor, the public parameter fields, solve(),
// migrateDeviceToHost(), the get*Values getters of NavierStokesSolver.h:37-42) and adds the device list. It is a
// header-only shim over fluidb200_multi_* (include/fluidb200.h); the grid is cut into x-slabs, one per device,
// halos travel over NVLink. Boundaries are the lid-driven cavity's (fluid_cpp/main.cpp:43-55), set up on the
// devices by uploadCavity(): per-cell setters are a single-device feature. Results equal ParallelNavierStokes<T>'s
// bit for bit.
#ifndef MULTIGPUNAVIERSTOKES_CUH
#define MULTIGPUNAVIERSTOKES_CUH

#include <vector>

#include "../cpp/NavierStokesSolver.h"
extern "C" {
#include "fluidb200.h"
}

template <class T>
class MultiGpuNavierStokes : public NavierStokesSolver<T> {
protected:
    fluidb200_multi* handle = nullptr;
    int last_error = 0;
    void note(int rc) { if (rc != 0) last_error = rc; }

public:
    // devices: one slab per entry (e.g. {0, 1, 2, 3}); empty = device 0 only
    MultiGpuNavierStokes(int box_dim_x, int box_dim_y, T dom_x, T dom_y, const std::vector<int>& devices = {})
        : NavierStokesSolver<T>(box_dim_x, box_dim_y, dom_x, dom_y) {
        const std::vector<int> d = devices.empty() ? std::vector<int>{0} : devices;
        note(fluidb200_multi_create(&handle, sizeof(T) == 4 ? FLUIDB200_F32 : FLUIDB200_F64, box_dim_x, box_dim_y, (double)dom_x,
                                    (double)dom_y, (int)d.size(), d.data()));
    }
    ~MultiGpuNavierStokes() { fluidb200_multi_destroy(handle); }
    MultiGpuNavierStokes(const MultiGpuNavierStokes&) = delete;
    MultiGpuNavierStokes& operator=(const MultiGpuNavierStokes&) = delete;

    void uploadCavity(T lid_velocity) { note(fluidb200_multi_upload_cavity(handle, (double)lid_velocity)); }
    // asynchronous, like the reference's solve() (ParallelNavierStokes.cu:269-288)
    void solve() {
        note(fluidb200_multi_set_params(handle, (double)this->time_step, (double)this->kinematic_viscosity, (double)this->density,
                                        this->num_poisson_iterations, this->num_iterations));
        note(fluidb200_multi_solve(handle));
    }
    void synchronize() { note(fluidb200_multi_synchronize(handle)); }
#ifndef NAVIERSTOKESCELL_H   // (stand-alone mode: this repository's structure-of-arrays base class)
    void migrateDeviceToHost() {
        for (int k = 0; k < FLUIDB200_NUM_FIELDS; ++k) note(fluidb200_multi_download(handle, k, this->hostField(k).data()));
    }
#endif
    int lastError() const { return last_error; }
    const char* lastErrorMessage() const { return fluidb200_last_error(); }
    fluidb200_multi* nativeHandle() { return handle; }
};

template class MultiGpuNavierStokes<float>;
template class MultiGpuNavierStokes<double>;

#endif  // MULTIGPUNAVIERSTOKES_CUH
