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

// ---- three-kernel form (reference scripts/diffusion_2D_kp.jl:16-54), same loop nests on the CPU ----------
// qx is (nx-1) x (ny-2), qy is (nx-2) x (ny-1), dTdt is (nx-2) x (ny-2); all row-major.
extern "C" int igg_cpu_flux2d(double* qx, double* qy, const double* T, int nx, int ny, double lam, double _dx,
                              double _dy) {
    if (!qx || !qy || !T || nx < 3 || ny < 3) return IGG_ERR_INVALID;
#pragma omp parallel for schedule(static)
    for (int ix = 0; ix < nx - 1; ++ix) {
        for (int iy = 0; iy < ny - 2; ++iy)
            qx[(size_t)ix * (ny - 2) + iy] = -lam * (T[(size_t)(ix + 1) * ny + iy + 1] - T[(size_t)ix * ny + iy + 1]) * _dx;
        if (ix < nx - 2)
            for (int iy = 0; iy < ny - 1; ++iy)
                qy[(size_t)ix * (ny - 1) + iy] =
                    -lam * (T[(size_t)(ix + 1) * ny + iy + 1] - T[(size_t)(ix + 1) * ny + iy]) * _dy;
    }
    return 0;
}

extern "C" int igg_cpu_residual2d(double* dTdt, const double* qx, const double* qy, const double* Cp, int nx, int ny,
                                  double _dx, double _dy) {
    if (!dTdt || !qx || !qy || !Cp || nx < 3 || ny < 3) return IGG_ERR_INVALID;
#pragma omp parallel for schedule(static)
    for (int ix = 0; ix < nx - 2; ++ix)
        for (int iy = 0; iy < ny - 2; ++iy) {
            const double dqx = (qx[(size_t)(ix + 1) * (ny - 2) + iy] - qx[(size_t)ix * (ny - 2) + iy]) * _dx;
            const double dqy = (qy[(size_t)ix * (ny - 1) + iy + 1] - qy[(size_t)ix * (ny - 1) + iy]) * _dy;
            dTdt[(size_t)ix * (ny - 2) + iy] = 1.0 / Cp[(size_t)(ix + 1) * ny + iy + 1] * -(dqx + dqy);
        }
    return 0;
}

extern "C" int igg_cpu_update2d(double* T, const double* dTdt, int nx, int ny, double dt) {
    if (!T || !dTdt || nx < 3 || ny < 3) return IGG_ERR_INVALID;
#pragma omp parallel for schedule(static)
    for (int ix = 0; ix < nx - 2; ++ix)
        for (int iy = 0; iy < ny - 2; ++iy)
            T[(size_t)(ix + 1) * ny + iy + 1] += dt * dTdt[(size_t)ix * (ny - 2) + iy];
    return 0;
}

// ---- 3-D fused step (7-point), the CPU restatement of igg_diffusion3d_step mode 0 --------------------------
extern "C" int igg_cpu_diffusion3d_step(const double* T, double* T2, const double* Cp, int nx, int ny, int nz,
                                        double lam, double dt, double _dx, double _dy, double _dz) {
    if (!T || !T2 || !Cp || nx < 3 || ny < 3 || nz < 3) return IGG_ERR_INVALID;
    const size_t sx = (size_t)ny * nz, sy = (size_t)nz;
#pragma omp parallel for collapse(2) schedule(static)
    for (int ix = 1; ix < nx - 1; ++ix)
        for (int iy = 1; iy < ny - 1; ++iy) {
            const size_t row = (size_t)ix * sx + (size_t)iy * sy;
            for (int iz = 1; iz < nz - 1; ++iz) {
                const size_t k = row + iz;
                const double c = T[k];
                const double qxr = -lam * (T[k + sx] - c) * _dx;
                const double qxl = -lam * (c - T[k - sx]) * _dx;
                const double qyr = -lam * (T[k + sy] - c) * _dy;
                const double qyl = -lam * (c - T[k - sy]) * _dy;
                const double qzr = -lam * (T[k + 1] - c) * _dz;
                const double qzl = -lam * (c - T[k - 1]) * _dz;
                T2[k] = c + dt * (Cp[k] * (-(qxr - qxl) * _dx - (qyr - qyl) * _dy - (qzr - qzl) * _dz));
            }
        }
    return 0;
}
