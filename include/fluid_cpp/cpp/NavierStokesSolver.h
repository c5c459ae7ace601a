// NavierStokesSolver.h -- stand-alone base class with the reference's public surface
// (fluid_cpp/cpp/NavierStokesSolver.h:9-44): constructor (int, int, T, T), the six public
// parameter fields, set{U,V,P}BoundaryCondition, setDomainSize, setBoxDimenension [sic] and the six
// get*Values getters. Use this header when the reference's tree is NOT on the include path; when
// it is, cuda/ParallelNavierStokes.cuh derives from the reference's own base instead (see there).
//
// Differences, all deliberate:
//  * host state is structure-of-arrays and allocated lazily, not an array of 92-byte cells
//    (fluid_cpp/cpp/NavierStokesCell.h:10-43, NavierStokesSolver.cpp:15-18), so a 16384^2 solver
//    driven through upload_cavity() costs no host memory;
//  * getters write output[x*Y + y] -- the order of the reference's device array
//    (cuda/ParallelNavierStokes.cu:22) and of its test (tests/test_main.cpp:76); the reference's
//    x*box_dimension_x + y (NavierStokesSolver.cpp:91) is the same for square grids only;
//  * setBoxDimenension really resizes (the reference frees with the wrong count and mixes
//    free/new[], NavierStokesSolver.cpp:33-50).
#ifndef FLUIDB200_NAVIERSTOKESSOLVER_H
#define FLUIDB200_NAVIERSTOKESSOLVER_H

#include <cstddef>
#include <cstdint>
#include <vector>

template <class T = float>
class NavierStokesSolver {
protected:
    int box_dimension_x;
    int box_dimension_y;
    T domain_size_x;
    T domain_size_y;
    T element_length_x;
    T element_length_y;

    // lazily allocated dense host arrays, index x*Y + y
    enum { F_U = 0, F_V, F_P, F_U_NEXT, F_V_NEXT, F_P_NEXT, F_RHS, F_COUNT };
    std::vector<T> host_[F_COUNT];
    std::vector<T> bc_val_[3];
    std::vector<std::uint8_t> bc_set_[3];

    std::size_t cellCount() const { return (std::size_t)box_dimension_x * (std::size_t)box_dimension_y; }
    std::vector<T>& hostField(int f) {
        if (host_[f].empty()) host_[f].assign(cellCount(), T(0));
        return host_[f];
    }
    void touchBc(int f) {
        if (bc_val_[f].empty()) {
            bc_val_[f].assign(cellCount(), T(0));
            bc_set_[f].assign(cellCount(), 0);
        }
    }
    int setBc(int f, int x, int y, T BC) {
        touchBc(f);
        const std::size_t i = (std::size_t)x * box_dimension_y + y;
        bc_val_[f][i] = BC;
        bc_set_[f][i] = 1;
        hostField(f)[i] = BC;  // the setter also seeds the live field, NavierStokesSolver.cpp:64-85
        return 0;
    }
    int getField(int f, T* output) {
        const std::vector<T>& a = hostField(f);
        for (std::size_t i = 0; i < a.size(); ++i) output[i] = a[i];
        return 0;
    }

public:
    T time_step = 0.001;
    T kinematic_viscosity = 0.1;
    T density = 1.0;
    int num_poisson_iterations = 50;
    int num_iterations = 1000;
    T stability_safety_factor = 0.5;  // read by no C++ solver of the reference either

    NavierStokesSolver(int box_dim_x, int box_dim_y, T dom_x, T dom_y)
        : box_dimension_x(box_dim_x), box_dimension_y(box_dim_y) {
        setDomainSize(dom_x, dom_y);
    }
    ~NavierStokesSolver() {}

    int setBoxDimenension(int x_dim, int y_dim) {
        if (x_dim == box_dimension_x && y_dim == box_dimension_y) return -1;
        box_dimension_x = x_dim;
        box_dimension_y = y_dim;
        for (auto& a : host_) a.clear();
        for (auto& a : bc_val_) a.clear();
        for (auto& a : bc_set_) a.clear();
        setDomainSize(domain_size_x, domain_size_y);
        return 0;
    }
    int setDomainSize(T dom_x, T dom_y) {
        domain_size_x = dom_x;
        domain_size_y = dom_y;
        element_length_x = dom_x / (box_dimension_x - 1);
        element_length_y = dom_y / (box_dimension_y - 1);
        return 0;
    }
    int setUBoundaryCondition(int x, int y, T BC) { return setBc(0, x, y, BC); }
    int setVBoundaryCondition(int x, int y, T BC) { return setBc(1, x, y, BC); }
    int setPBoundaryCondition(int x, int y, T BC) { return setBc(2, x, y, BC); }

    int getUValues(T* output) { return getField(F_U, output); }
    int getVValues(T* output) { return getField(F_V, output); }
    int getPValues(T* output) { return getField(F_P, output); }
    int getUNextValues(T* output) { return getField(F_U_NEXT, output); }
    int getRightHand(T* output) { return getField(F_RHS, output); }
    int getPNext(T* output) { return getField(F_P_NEXT, output); }
};

#endif  // FLUIDB200_NAVIERSTOKESSOLVER_H
