// CPU restatement of the fused diffusion step: correctness oracle and timed CPU baseline.
//
// The reference has no CPU code path (every array is a ROCArray); this is the same arithmetic as
// its fused kernel (scripts/diffusion_2D_perf.jl:8-10) written as a plain C++ loop nest, row-major
// T[ix][iy] with iy contiguous.  Compiled with -ffp-contract=off so that it rounds exactly like the
// PyTorch statement in ops/reference.py.
#include <stddef.h>
#include <utility>

#include "../include/igg_b200.h"

extern "C" int igg_cpu_diffusion2d_step(const double* T, double* T2, const double* Cp, int nx, int ny, double lam,
                                        double dt, double _dx, double _dy) {
    if (!T || !T2 || !Cp || nx < 3 || ny < 3) return IGG_ERR_INVALID;
    // rows are independent: OpenMP over ix when the library was built with it (IGG_CPU_THREADS / OMP_NUM_THREADS)
#pragma omp parallel for schedule(static)
    for (int ix = 1; ix < nx - 1; ++ix) {
        const double* c = T + (size_t)ix * ny;
        const double* u = c - ny;
        const double* d = c + ny;
        const double* cp = Cp + (size_t)ix * ny;
        double* o = T2 + (size_t)ix * ny;
        for (int iy = 1; iy < ny - 1; ++iy) {
            const double qxr = -lam * (d[iy] - c[iy]) * _dx;
            const double qxl = -lam * (c[iy] - u[iy]) * _dx;
            const double qyt = -lam * (c[iy + 1] - c[iy]) * _dy;
            const double qyb = -lam * (c[iy] - c[iy - 1]) * _dy;
            o[iy] = c[iy] + dt * (cp[iy] * (-(qxr - qxl) * _dx - (qyt - qyb) * _dy));
        }
    }
    return 0;
}

// nt steps with pointer swapping; the final field ends up in T (copied back if nt is odd).
extern "C" int igg_cpu_diffusion2d_run(double* T, double* T2, const double* Cp, int nx, int ny, double lam, double dt,
                                       double _dx, double _dy, int nt) {
    if (nt < 0) return IGG_ERR_INVALID;
    double* a = T;
    double* b = T2;
    for (int it = 0; it < nt; ++it) {
        int rc = igg_cpu_diffusion2d_step(a, b, Cp, nx, ny, lam, dt, _dx, _dy);
        if (rc) return rc;
        std::swap(a, b);
    }
    if (a != T) {
        for (size_t k = 0; k < (size_t)nx * ny; ++k) T[k] = a[k];
    }
    return 0;
}
