/* The fused GPU time loop from plain C: no Python, no PyTorch -- the CUDA runtime and libigg_b200.so.
 *
 *   gcc -std=c99 -O2 -I include -I /usr/local/cuda/include examples/c/heat_gpu.c -o heat_gpu \
 *       -L rocm_mpi_b200/_lib -ligg_b200 -Wl,-rpath,$PWD/rocm_mpi_b200/_lib -L /usr/local/cuda/lib64 -lcudart
 *
 * One B200, nx x ny cells: Gaussian initial condition on the device (reference scripts/diffusion_2D_perf.jl:35),
 * nt fused steps from the library's native loop (:47-52), CUDA-event timing after 10 warm-up steps and the
 * T_eff line (:55-58), then max / sum of the field with the entry point that needs no scratch allocation.
 * (Multi-GPU from C: fill igg_push2d with peer pointers obtained through igg_ipc_export / igg_ipc_open.)
 */
#include <cuda_runtime_api.h>
#include <stdio.h>
#include <stdlib.h>

#include "igg_b200.h"

#define CUDA_OK(x)                                                                          \
    do {                                                                                    \
        cudaError_t e_ = (x);                                                               \
        if (e_ != cudaSuccess) {                                                            \
            fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e_));                        \
            return 1;                                                                       \
        }                                                                                   \
    } while (0)
#define IGG_OK(x)                                                                           \
    do {                                                                                    \
        int r_ = (x);                                                                       \
        if (r_) {                                                                           \
            fprintf(stderr, "%s failed (%d): %s\n", #x, r_, igg_last_error());              \
            return 1;                                                                       \
        }                                                                                   \
    } while (0)

int main(int argc, char** argv) {
    const int nx = argc > 1 ? atoi(argv[1]) : 16384, ny = argc > 2 ? atoi(argv[2]) : 16384;
    const int nt = argc > 3 ? atoi(argv[3]) : 110, warm = 10;
    const double lx = 10.0, ly = 10.0, lam = 1.0;
    const double dx = lx / nx, dy = ly / ny;
    const double dt = (dx < dy ? dx : dy) * (dx < dy ? dx : dy) / lam / 4.1;
    const size_t bytes = sizeof(double) * (size_t)nx * ny;
    double *T, *T2, *Cp, *out;
    CUDA_OK(cudaMalloc((void**)&T, bytes));
    CUDA_OK(cudaMalloc((void**)&T2, bytes));
    CUDA_OK(cudaMalloc((void**)&Cp, bytes));
    CUDA_OK(cudaMalloc((void**)&out, 2 * sizeof(double)));
    cudaStream_t st;
    CUDA_OK(cudaStreamCreate(&st));
    IGG_OK(igg_gaussian_ic2d(T, nx, ny, 0.0, 0.0, dx, dy, lx, ly, st));
    IGG_OK(igg_gaussian_ic2d(Cp, nx, ny, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, st));   /* all-zero geometry: exp(0) = 1 = Cp0 */
    CUDA_OK(cudaMemcpyAsync(T2, T, bytes, cudaMemcpyDeviceToDevice, st));
    double one[2];
    IGG_OK(igg_reduce_max_sum_static(Cp, (size_t)nx * ny, out, st));          /* first call: before the loop */
    CUDA_OK(cudaMemcpyAsync(one, out, sizeof(one), cudaMemcpyDeviceToHost, st));
    CUDA_OK(cudaStreamSynchronize(st));
    if (one[0] != 1.0) {
        fprintf(stderr, "Cp initialisation failed (max = %g)\n", one[0]);
        return 1;
    }
    /* even step counts leave the result in T (see igg_diffusion2d_run) */
    IGG_OK(igg_diffusion2d_run(T, T2, Cp, nx, ny, lam, dt, 1.0 / dx, 1.0 / dy, 16, 4, 1, warm, NULL, NULL, st));
    cudaEvent_t e0, e1;
    CUDA_OK(cudaEventCreate(&e0));
    CUDA_OK(cudaEventCreate(&e1));
    CUDA_OK(cudaEventRecord(e0, st));
    IGG_OK(igg_diffusion2d_run(T, T2, Cp, nx, ny, lam, dt, 1.0 / dx, 1.0 / dy, 16, 4, 1, nt - warm, NULL, NULL, st));
    CUDA_OK(cudaEventRecord(e1, st));
    CUDA_OK(cudaEventSynchronize(e1));
    float ms = 0.f;
    CUDA_OK(cudaEventElapsedTime(&ms, e0, e1));
    const double* res = ((nt - warm) % 2) ? T2 : T;
    double ms_it = ms / (nt - warm), a_eff = 3.0 * nx * (double)ny * 8.0 / 1e9;
    IGG_OK(igg_reduce_max_sum_static(res, (size_t)nx * ny, out, st));
    CUDA_OK(cudaMemcpyAsync(one, out, sizeof(one), cudaMemcpyDeviceToHost, st));
    CUDA_OK(cudaStreamSynchronize(st));
    printf("Executed %d steps in = %1.3e sec (@ T_eff = %1.0f GB/s); max(T) = %.12f, launches = %llu\n", nt,
           ms * 1e-3, a_eff / (ms_it * 1e-3), one[0], igg_launch_count());
    cudaFree(T);
    cudaFree(T2);
    cudaFree(Cp);
    cudaFree(out);
    return 0;
}
