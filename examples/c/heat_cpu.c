/* Calling the native library from plain C (no Python, no GPU needed for this one).
 *
 *   gcc -std=c99 -O2 -I include examples/c/heat_cpu.c -o heat_cpu \
 *       -L rocm_mpi_b200/_lib -ligg_b200 -Wl,-rpath,$PWD/rocm_mpi_b200/_lib -lm
 *
 * Runs the diffusion of reference scripts/diffusion_2D_perf.jl (Gaussian initial condition :35, fused step
 * :8-10, nt steps) on one CPU domain with the OpenMP restatement, prints max(T) and a checksum.  The GPU entry
 * points (igg_diffusion2d_step, ...) are called the same way with device pointers and a stream.
 */
#include <math.h>
#include <stdio.h>
#include <stdlib.h>

#include "igg_b200.h"

int main(int argc, char** argv) {
    const int nx = argc > 1 ? atoi(argv[1]) : 254, ny = argc > 2 ? atoi(argv[2]) : 254;
    const int nt = argc > 3 ? atoi(argv[3]) : 1000;
    const double lx = 10.0, ly = 10.0, lam = 1.0, Cp0 = 1.0;
    const double dx = lx / nx, dy = ly / ny;
    const double dt = (dx < dy ? dx : dy) * (dx < dy ? dx : dy) * Cp0 / lam / 4.1;
    double* T = malloc(sizeof(double) * nx * ny);
    double* T2 = malloc(sizeof(double) * nx * ny);
    double* Cp = malloc(sizeof(double) * nx * ny);
    if (!T || !T2 || !Cp) return 2;
    for (int ix = 0; ix < nx; ++ix)
        for (int iy = 0; iy < ny; ++iy) {
            const double x = ix * dx + dx / 2 - lx / 2, y = iy * dy + dy / 2 - ly / 2;
            T[ix * ny + iy] = T2[ix * ny + iy] = exp(-x * x - y * y);
            Cp[ix * ny + iy] = Cp0;
        }
    const int rc = igg_cpu_diffusion2d_run(T, T2, Cp, nx, ny, lam, dt, 1.0 / dx, 1.0 / dy, nt);
    if (rc) {
        fprintf(stderr, "igg_cpu_diffusion2d_run failed (%d): %s\n", rc, igg_last_error());
        return 1;
    }
    double mx = -1.0, sum = 0.0;
    for (int ix = 1; ix < nx - 1; ++ix)
        for (int iy = 1; iy < ny - 1; ++iy) {
            const double v = T[ix * ny + iy];
            mx = v > mx ? v : mx;
            sum += v;
        }
    printf("libigg_b200 version %d: %d x %d, %d steps, max(T_v) = %.16f, sum = %.10f\n", igg_version(), nx, ny, nt, mx,
           sum);
    free(T);
    free(T2);
    free(Cp);
    return 0;
}
